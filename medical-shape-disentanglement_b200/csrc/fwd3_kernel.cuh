// Fused 9-layer DeepSDF forward chain, v3: persistent CTAs, 128 rows per CTA tile, lean
// single-thread issue loops, optional CTA pair (CG = 2: tcgen05 cta_group::2, M = 256 over two SMs).
//
//   warp 0      : weight producer (one elected lane): K-steps ([256 n x 16 k] bf16, layout.cuh)
//                 global -> smem ring with the TMA engine.  CG = 2: each CTA of the pair loads ITS HALF
//                 of every step (128 n rows) into its own ring with a tensor-map load
//                 (cp.async.bulk.tensor.2d.cta_group::2, SASS UTMALDG.2D.2CTA) whose complete_tx lands on the
//                 LEADER's `full` barrier; the leader's producer posts one arrive.expect_tx for both halves.
//   warp 1      : MMA issuer (one elected lane of the leader CTA; the peer's warp 1 idles): tcgen05.mma
//                 kind::f16, M = 128*CG, N = 256, K = 16; a try_wait + UTCHMMA + UTCBAR per step keeps the
//                 tensor pipe at its 128-cycle-per-step floor (tools/gpu_mma_rate.py).  After the last
//                 half's K-steps 0..15 it commits `lo_free`.
//   warp 2      : TMEM allocator (512 columns); warp 3: activation-stash warp (TRAIN)
//   warps 4..11 : epilogue warps: each layer is two 256-column halves; h0's ReLU'd bf16 activations are
//                 held in registers until `lo_free` (the layer's last read of A columns 0..255 has
//                 retired) and then become the next layer's A columns 0..255 UNDER the K-steps that still
//                 read columns 256..511, so the next layer's first MMAs follow without a bubble; h1 follows
//                 straight from TMEM once the layer has retired.
//
// E steps (xyz hi/lo + bias through the ones columns; in FOLD mode the per-segment bias
// b + W[:, latent] z of lin0 / lin4) travel through their own 2-slot ring so that the W ring always
// advances in multiples of 16 steps.
//
// TRAIN mode also writes, per layer, the activation stash (bf16 images = the smem A tile byte for byte, copied
// out by warp 3 with coalesced stores; consumed by the wgrad kernel) and a 1-bit/element ReLU mask (consumed by
// the backward chain instead of re-reading the 1 KB/row activations).
//
// Replaces, per 128-row tile, reference networks/deep_sdf_decoder.py:74-109 (9x addmm, 2x cat,
// 8x relu, tanh), deep_sdf/utils.py:86-97 (expand + cat) and, in TRAIN mode,
// train_deep_sdf.py:598-606 (gather, cat, clamp, L1/N).
#pragma once
#include "common.cuh"
#include "layout.cuh"
#include "ptx.cuh"

namespace dsdf {

template <int CG>
struct F3 {
  static constexpr int kRing = CG == 2 ? DSDF_FWD_RING : 8;   // W ring stages (per CTA), any depth <= 32
  static constexpr uint32_t kSlice = kStepBytes / CG;    // bytes of one K-step held by one CTA
  static constexpr uint32_t kSmemA = 0;                  // 8 SUBs (SWIZZLE_128B)
  static constexpr uint32_t kSmemEA = kTileBytes;        // [128 x 16] no-swizzle, 4 KB
  static constexpr uint32_t kSmemRing = kSmemEA + 4096;
  static constexpr uint32_t kSmemERing = kSmemRing + kRing * kSlice;   // 64 KB W ring
  static constexpr uint32_t kSmemAux = kSmemERing + 2 * kSlice;
  static constexpr uint32_t kAuxTmem = 1024;             // after 128 barrier slots
  static constexpr uint32_t kAuxPart = 1040;             // 128 floats
  static constexpr uint32_t kAuxRed = 1552;              // 16 floats
  static constexpr uint32_t kAuxW8 = 1616;               // 512 floats
  static constexpr uint32_t kAuxG = kAuxW8 + 2048;       // 128 floats: dL/ds of the tile's rows
  static constexpr uint32_t kSmemBytes = kSmemAux + kAuxG + 512;
};
constexpr int kF3Threads = 384;

// bit (31 - i) of the result = (v[i] > 0) for fp32 accumulators (sign-bit test; +0 counts as
// positive, which only ever happens on zero-padded columns whose gradients are discarded)
__device__ __forceinline__ uint32_t pos_mask32(const uint32_t (&v)[32]) {
  uint32_t m = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) m = __funnelshift_l(v[i], m, 1);
  return ~m;
}

// 16 packed words = columns col..col+31 of row r -> activation-stash image in HBM (same
// SWIZZLE_128B SUB layout as the smem A tile).  Straight from the epilogue registers: a 128 KB
// bulk store per layer would sit in the SM's TMA queue in front of the 4 KB weight-ring loads and
// starve the MMA (measured: 46 % of the issuer's time waiting on `full`).
__device__ __forceinline__ void stash_chunk32(uint8_t* slot_base, uint32_t r, uint32_t col,
                                              const uint32_t* p) {
  uint8_t* sub = slot_base + (size_t)(col >> 6) * kSubBytes;
  const uint32_t ch0 = (col & 63) >> 3;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<uint4*>(sub + sub_chunk_off(r, ch0 + i)) =
        make_uint4(p[4 * i], p[4 * i + 1], p[4 * i + 2], p[4 * i + 3]);
}

// 32 fp32 columns -> zero where mask bit (31 - i) is clear, else x * scale -> 16 packed bf16x2
// (ReLU and dropout in one select: the mask is (x > 0) & keep)
__device__ __forceinline__ void mask_scale_pack32(const uint32_t (&v)[32], uint32_t m, float scale,
                                                  uint32_t* p) {
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float lo = (m & (0x80000000u >> (2 * i))) ? __uint_as_float(v[2 * i]) * scale : 0.f;
    const float hi = (m & (0x40000000u >> (2 * i))) ? __uint_as_float(v[2 * i + 1]) * scale : 0.f;
    p[i] = pack_bf16x2(lo, hi);
  }
}

// 32 fp32 columns -> zero where the mask bit (31 - i) is clear -> 16 packed bf16x2
__device__ __forceinline__ void mask_pack32(const uint32_t (&v)[32], uint32_t m, uint32_t* p) {
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float lo = (m & (0x80000000u >> (2 * i))) ? __uint_as_float(v[2 * i]) : 0.f;
    const float hi = (m & (0x40000000u >> (2 * i))) ? __uint_as_float(v[2 * i + 1]) : 0.f;
    p[i] = pack_bf16x2(lo, hi);
  }
}

#ifdef DSDF_ARCH_SIREN
// sin / cos of y = 30 x with |y| possibly tens of radians (trained SIRENs): the hardware
// approximation's absolute error grows with |y|, so reduce first -- Cody-Waite with 2 pi = C1 + C2, C1 exact in
// 9 bits (k C1 is exact for |k| < 2^15, i.e. |y| < 2e5) -- and evaluate on r in [-pi, pi] where it is ~2^-21.
__device__ __forceinline__ void sincos_reduced(float y, float& s, float& c) {
  // round-to-nearest of y / 2 pi by the 1.5 * 2^23 add trick: two FMA-pipe ops instead of F2F/rint on the
  // quarter-rate conversion pipe the SFU-bound epilogue already saturates (valid for |y / 2 pi| < 2^22)
  const float k = fmaf(y, 0.15915494309189535f, 12582912.f) - 12582912.f;
  float r = fmaf(-k, 6.28125f, y);
  r = fmaf(-k, 1.9353071795864769e-3f, r);
  __sincosf(r, &s, &c);
}
// networks/modules.py:4-14 Sine: 32 fp32 pre-activations -> sin(30 x) (16 packed bf16x2 in p) and the derivative's
// cos(30 x) as 32 signed bytes round(127 cos) in d[0..7] (cos is bounded, so a linear 8-bit code has the absolute
// error of bf16 at |cos| ~ 1 and a smaller one elsewhere: half the stash bytes of a bf16 derivative; the backward
// chain applies the factor 30 / 127); returns sum_i sin(30 x_i) w[i] when w != nullptr.
// byte = low 8 bits of float_as_int(127 c + 1.5 * 2^23): rounding on the FMA pipe, no conversion instruction.
__device__ __forceinline__ float sine_pack32(const uint32_t (&v)[32], uint32_t* p, uint32_t* d, const float* w) {
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float s0, c0, s1, c1;
    sincos_reduced(30.f * __uint_as_float(v[2 * i]), s0, c0);
    sincos_reduced(30.f * __uint_as_float(v[2 * i + 1]), s1, c1);
    p[i] = pack_bf16x2(s0, s1);
    const uint32_t q0 = __float_as_uint(fmaf(c0, 127.f, 12582912.f)), q1 = __float_as_uint(fmaf(c1, 127.f, 12582912.f));
    const uint32_t two = __byte_perm(q0, q1, 0x0040);            // bytes [q0.b0, q1.b0, q0.b0, q0.b0] -> low 16 bits used
    if (i & 1) d[i >> 1] = __byte_perm(d[i >> 1], two, 0x5410);  // [old.b0, old.b1, two.b0, two.b1]
    else d[i >> 1] = two;
    if (w) acc = fmaf(s1, w[2 * i + 1], fmaf(s0, w[2 * i], acc));
  }
  return acc;
}
__device__ __forceinline__ void deriv_store(uint8_t* aux_tile, int l, int chunk, uint32_t r, const uint32_t* d) {
  uint8_t* b = aux_tile + (size_t)(l * 16 + chunk) * (2 * kTileM * 16) + (size_t)r * 16;
#pragma unroll
  for (int i = 0; i < 2; ++i)
    *reinterpret_cast<uint4*>(b + (size_t)i * kTileM * 16) = make_uint4(d[4 * i], d[4 * i + 1], d[4 * i + 2], d[4 * i + 3]);
}
__device__ __forceinline__ void deriv_load(const uint8_t* aux_tile, int l, int chunk, uint32_t r, uint32_t* d) {
  const uint8_t* b = aux_tile + (size_t)(l * 16 + chunk) * (2 * kTileM * 16) + (size_t)r * 16;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(b + (size_t)i * kTileM * 16));
    d[4 * i] = q.x; d[4 * i + 1] = q.y; d[4 * i + 2] = q.z; d[4 * i + 3] = q.w;
  }
}
#endif

// TAN: the Eikonal recompute pass (train_deep_sdf.py:724-729 without double backward, tests/eikonal_algebra.py):
// the same chain over the perturbed per-row input (x', y', z', eps) of FwdArgs::eik_row, with the ReLU decisions
// READ from the first pass's bit-masks instead of taken from the sign (the map is then linear in the input), no
// output / loss; stashes Xt_l = (1 + eps) h_l + tangent and accumulates dW8 = sum_r g_r Xt8[r, :].
template <int MODE, bool FOLD, bool TRAIN, int CG, bool DROP = false, bool TAN = false>
__global__ void __launch_bounds__(kF3Threads, 1) k_forward3(const __grid_constant__ FwdArgs a) {
  static_assert(!DROP || TRAIN, "dropout exists in training mode only");
  static_assert(!TAN || (TRAIN && FOLD && !DROP && MODE == MODE_TABLE), "the recompute pass is a fold-mode training pass");
  using C = F3<CG>;
  constexpr int kRing = C::kRing;
  constexpr uint32_t kSlice = C::kSlice;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kSmemAux);
  static_assert(kRing <= 32, "barrier layout holds 32 ring stages");
  uint64_t* full = bars;                 // [kRing]  producer -> MMA (leader's copy is the live one)
  uint64_t* empty = bars + 32;           // [kRing]  MMA -> producer (committed to both CTAs)
  uint64_t* e_full = bars + 64;          // [2]
  uint64_t* e_empty = bars + 66;         // [2]
  uint64_t* a_lo = bars + 68;            // epilogue -> MMA: A SUB 0..3 (+E_A) ready, TMEM h0 drained
  uint64_t* a_hi = bars + 69;            // epilogue -> MMA: A SUB 4..7 ready, TMEM h1 drained
  uint64_t* d_done = bars + 70;          // [2] MMA -> epilogue: half h accumulated
  uint64_t* s_lo = bars + 72;            // TRAIN: epilogue -> stash warp: A SUB 0..3 written
  uint64_t* s_hi = bars + 73;            //        epilogue -> stash warp: A SUB 4..7 written
  uint64_t* s_done = bars + 74;          //        stash warp -> epilogue: A SUB 0..3 copied to HBM
  uint64_t* s_done_hi = bars + 75;       //        stash warp -> epilogue: A SUB 4..7 copied to HBM
  uint64_t* lo_free = bars + 80;         // MMA -> epilogue: the layer's last read of A SUB 0..3 has retired (both CTAs)
  uint64_t* a_sub = bars + 76;           // [3] epilogue -> MMA at a layer boundary: A SUB 1 / 2 / 3 written
                                         // (a_lo then only means SUB 0: the MMAs of the next layer start
                                         //  after 16 KB of A stores instead of 64 KB)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + C::kSmemAux + C::kAuxTmem);
  float* part = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxPart);
  float* red = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxRed);
  float* w8s = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxW8);
  float* gsm = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxG);

  if ((s_base & 1023u) != 0) __trap();
  const uint32_t crank = CG > 1 ? cluster_ctarank() : 0u;
  const bool leader = crank == 0;

  if (threadIdx.x == 0) {
    // full barriers: ONE arrive.expect_tx (the leader's producer, for the bytes of both CTAs); in a pair every
    // CTA's tensor-map copy reports its bytes to the leader's barrier (the peer's own copies are unused)
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&e_full[i], 1);
      mbar_init(&e_empty[i], 1);
      mbar_init(&d_done[i], 1);
    }
    mbar_init(a_lo, 8 * CG);             // one elected arrive per epilogue warp of every CTA
    mbar_init(a_hi, 8 * CG);
    for (int i = 0; i < 3; ++i) mbar_init(&a_sub[i], 4 * CG);   // the 4 warps that own the SUB's columns
    mbar_init(s_lo, 8);
    mbar_init(s_hi, 8);
    mbar_init(s_done, 1);
    mbar_init(s_done_hi, 1);
    mbar_init(lo_free, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    if (CG > 1) tmem_alloc2<512>(tmem_slot); else tmem_alloc<512>(tmem_slot);
  }
  for (int i = threadIdx.x; i < kHidden; i += blockDim.x) w8s[i] = a.w8[i];
  tc_fence_before();
  __syncthreads();
  if (CG > 1) cluster_sync_all();        // the peer's barriers exist before anything is signalled
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // programmatic dependent launch (dsdf_reconstruct's latency-bound loop; a no-op for ordinary launches): the next
  // grid may start its own prologue now, and nothing the previous grid wrote is read before it has completed
  grid_dep_launch();
  grid_dep_wait();

  const long long n_tiles = (a.n_rows + kTileM - 1) / kTileM;
  // every CTA runs the same number of iterations (a pair is lock-stepped through the MMA);
  // iterations past the last tile recompute the last tile and write nothing.
  const long long n_iter = (n_tiles + gridDim.x - 1) / gridDim.x;
  // barriers the MMA issuer waits on live in the leader CTA
  const uint32_t full_c = CG > 1 ? mapa_u32(smem_u32(full), 0) : smem_u32(full);
  const uint32_t e_full_c = CG > 1 ? mapa_u32(smem_u32(e_full), 0) : smem_u32(e_full);
  const uint32_t a_lo_c = CG > 1 ? mapa_u32(smem_u32(a_lo), 0) : smem_u32(a_lo);
  const uint32_t a_hi_c = CG > 1 ? mapa_u32(smem_u32(a_hi), 0) : smem_u32(a_hi);
  const uint32_t a_sub_c = CG > 1 ? mapa_u32(smem_u32(a_sub), 0) : smem_u32(a_sub);

  if (warp == 0) {
    // ===================== weight producer =====================
    if (elect_one()) {
      uint32_t ph = 0, st = 0, es = 0, eph = 0;   // st/ph: ring stage and its phase, running through layers and tiles
      for (long long it = 0; it < n_iter; ++it) {
        long long tile = it * gridDim.x + blockIdx.x;
        tile = tile < n_tiles ? tile : n_tiles - 1;
        long long se_off = 0;                 // byte offset of this tile's scene E steps in scene_e
        if (FOLD) {
          const long long seg = (MODE == MODE_TABLE && a.seg_ids) ? (tile * kTileM) / a.rows_per_seg : 0;
          se_off = seg * (long long)(kSceneESteps * kStepBytes);
        }
#pragma unroll 1
        for (int l = 0; l < 8; ++l) {
          const int ksteps = f2_used_ksteps<FOLD>(l);
#pragma unroll 1
          for (int h = 0; h < w2_halves(l); ++h) {
            const long long w_off = (long long)w2_step(l, h, 0) * kStepBytes + crank * kSlice;
            const bool scene_step = FOLD && (l == 0 || l == kSkipLayer);
            const long long e_off = scene_step ? se_off + (long long)((l == 0 ? 0 : 2) + h) * kStepBytes + crank * kSlice
                                               : w_off;
            mbar_wait(&e_empty[es], eph ^ 1);
            if (CG > 1) {
              if (leader) mbar_arrive_expect_tx(&e_full[es], CG * kSlice);
              tma2_load_2d(s_base + C::kSmemERing + es * kSlice, scene_step ? &a.tmap_e : &a.tmap_w, 0,
                           (int)(e_off / kMapRowBytes), e_full_c + es * 8);
            } else {
              mbar_arrive_expect_tx(&e_full[es], kSlice);
              bulk_g2s(smem + C::kSmemERing + es * kSlice, (scene_step ? a.scene_e : a.w2img) + e_off, kSlice, &e_full[es]);
            }
            es ^= 1;
            eph ^= (es == 0);
            if (CG > 1) {
              int y = (int)((w_off + kStepBytes) / kMapRowBytes);
#pragma unroll 4
              for (int k = 0; k < ksteps; ++k) {
                mbar_wait(&empty[st], ph ^ 1);
                if (leader) mbar_arrive_expect_tx(&full[st], CG * kSlice);
                tma2_load_2d(s_base + C::kSmemRing + st * kSlice, &a.tmap_w, 0, y, full_c + st * 8);
                y += (int)(kStepBytes / kMapRowBytes);
                if (++st == (uint32_t)kRing) { st = 0; ph ^= 1; }
              }
            } else {
              const uint8_t* w = a.w2img + w_off + kStepBytes;
#pragma unroll 4
              for (int k = 0; k < ksteps; ++k) {
                mbar_wait(&empty[st], ph ^ 1);
                mbar_arrive_expect_tx(&full[st], kSlice);
                bulk_g2s(smem + C::kSmemRing + st * kSlice, w, kSlice, &full[st]);
                w += kStepBytes;
                if (++st == (uint32_t)kRing) { st = 0; ph ^= 1; }
              }
            }
          }
        }
      }
    }
  } else if (warp == 1 && leader) {
    // ===================== MMA issuer (leader CTA only; the peer's warp 1 idles) =====================
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128 * CG, 256, false, false);
      const uint64_t ea_desc = umma_smem_desc_ns(s_base + C::kSmemEA, 128, 256);
      const uint64_t a_desc0 = umma_smem_desc(s_base + C::kSmemA, 16, 1024);
      const uint64_t b_desc0 = umma_smem_desc_ns(s_base + C::kSmemRing, 128, 256);
      const uint64_t eb_desc0 = umma_smem_desc_ns(s_base + C::kSmemERing, 128, 256);
      uint32_t ph = 0, st = 0, es = 0, eph = 0, par_lo = 0, par_hi = 0, par_sub = 0;
      long long dbg_wf = 0, dbg_wl = 0, dbg_wh = 0, dbg_we = 0, dbg_first = 0;
      const long long dbg_t0 = DSDF_DBG_ON(a) ? clock64() : 0;
      auto timed_wait = [&](uint64_t* bar, uint32_t par, long long& acc) {
        if (DSDF_DBG_ON(a)) {
          const long long c0 = clock64();
          mbar_wait(bar, par);
          acc += clock64() - c0;
        } else {
          mbar_wait(bar, par);
        }
      };
      auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t acc) {
        if (CG > 1) umma2_bf16(d, ad, bd, idesc, acc); else umma_bf16(d, ad, bd, idesc, acc);
      };
      auto commit = [&](uint64_t* bar) {
        if (CG > 1) umma2_commit_mc(bar, (uint16_t)3); else umma_commit(bar);
      };
      for (long long it = 0; it < n_iter; ++it) {
#pragma unroll 1
        for (int l = 0; l < 8; ++l) {
          const int ksteps = f2_used_ksteps<FOLD>(l);
          timed_wait(a_lo, par_lo, dbg_wl);
          par_lo ^= 1;
          tc_fence_after();
#pragma unroll 1
          for (int h = 0; h < w2_halves(l); ++h) {
            const uint32_t d_addr = tmem + h * 256;
            // E step: xyz (hi/lo) + bias through the ones columns; overwrites the accumulator
            timed_wait(&e_full[es], eph, dbg_we);
            tc_fence_after();
            mma(d_addr, ea_desc, eb_desc0 + (uint64_t)(es * (kSlice >> 4)), 0u);
            if (DSDF_DBG_ON(a) && h == 0 && l >= 1) dbg_first += clock64();   // first MMA of layer l issued
            commit(&e_empty[es]);
            es ^= 1;
            eph ^= (es == 0);
#pragma unroll 1
            for (int sub = 0; sub < (ksteps >> 2); ++sub) {   // one A SUB = 64 K = 4 K-steps
              if (h == 0 && sub == 4) {          // first K-step that reads A SUB 4..7
                timed_wait(a_hi, par_hi, dbg_wh);
                par_hi ^= 1;
                tc_fence_after();
              }
              if (h == 0 && l >= 1 && sub >= 1 && sub <= 3) {
                // an A SUB the epilogue wrote at the layer boundary (SUB 0 = a_lo)
                timed_wait(&a_sub[sub - 1], par_sub, dbg_wl);
                tc_fence_after();
              }
              const uint64_t a_desc = a_desc0 + (uint64_t)((sub * kSubBytes) >> 4);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                timed_wait(&full[st], ph, dbg_wf);
                tc_fence_after();
                mma(d_addr, a_desc + (uint64_t)((k * 32) >> 4), b_desc0 + (uint64_t)((st * kSlice) >> 4), 1u);
                commit(&empty[st]);
                if (++st == (uint32_t)kRing) { st = 0; ph ^= 1; }
              }
              // the layer's last half has read A SUB 0..3 for the last time: the epilogue may overwrite them with
              // the next layer's input while K-steps 16..31 (SUB 4..7) are still running
              if (h == w2_halves(l) - 1 && sub == 3 && ksteps > 16) commit(lo_free);
            }
            if (h == w2_halves(l) - 1 && ksteps <= 16) commit(lo_free);
            commit(&d_done[h]);
          }
          if (l >= 1) par_sub ^= 1;
        }
      }
      if (DSDF_DBG_ON(a)) {
        unsigned long long* d = a.dbg + (size_t)blockIdx.x * kDbgSlots;
        d[0] = clock64() - dbg_t0; d[1] = dbg_wf; d[2] = dbg_wl; d[3] = dbg_wh; d[8] = dbg_we; d[11] = dbg_first;
      }
    }
  } else if (warp == 3) {
    // ===================== TRAIN: activation-stash warp =====================
    // The stash image in HBM is byte-for-byte the smem A tile, so this is a linear copy with fully
    // coalesced 512 B warp stores.  (Not a TMA bulk store: 128 KB per layer in the SM's TMA queue
    // starves the 4 KB weight-ring loads; not per-row stores from the epilogue registers either:
    // 16 B per 128 B line.)
    if (TRAIN) {
      const uint64_t pol = stash_policy();
      uint32_t par_lo = 0, par_hi = 0;
      for (long long it = 0; it < n_iter; ++it) {
        long long tile = it * gridDim.x + blockIdx.x;
        const bool tile_valid = tile < n_tiles;
        tile = tile_valid ? tile : n_tiles - 1;
        uint8_t* xs_tile = a.xstash + (size_t)tile * kStashTileBytes;
#pragma unroll 1
        for (int l = 0; l < 8; ++l) {
          uint4* dst = reinterpret_cast<uint4*>(xs_tile + (size_t)l * kTileBytes);
          const uint32_t src = s_base + C::kSmemA;
#pragma unroll 1
          for (int half = 0; half < 2; ++half) {
            if (half == 1 && l == 3 && FOLD) break;          // lin3 has 256 outputs (x3 part): nothing to copy, and no
                                                             // s_done_hi phase (an ungated arrive could lap a slow warp)
            if (half == 0) { mbar_wait(s_lo, par_lo); par_lo ^= 1; }
            else { mbar_wait(s_hi, par_hi); par_hi ^= 1; }
            // frozen-decoder callers pass no stash (no wgrad); X8 (slot 7) is consumed on-chip when the
            // epilogue accumulates dW8 itself
            if (tile_valid && a.xstash && !(l == 7 && a.dw8_partials)) {
#pragma unroll 1
              for (int i = 0; i < 128; i += 8) {              // 4 SUBs = 4096 x 16 B = 128 x (32 lanes)
                uint4 v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                  const uint32_t off = (uint32_t)(half * 4 * kSubBytes) + (uint32_t)((i + k) * 32 + lane) * 16u;
                  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                               : "=r"(v[k].x), "=r"(v[k].y), "=r"(v[k].z), "=r"(v[k].w)
                               : "r"(src + off));
                }
#pragma unroll
                for (int k = 0; k < 8; ++k)
                  st_stash(dst + (size_t)(half * 4 * kSubBytes) / 16 + (size_t)(i + k) * 32 + lane, v[k], pol);
              }
            }
            __syncwarp();                                     // every lane's ld.shared of this half has returned
            if (lane == 0) mbar_arrive(half == 0 ? s_done : s_done_hi);
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps =====================
    const int t = threadIdx.x - 128;      // 0..255
    const int q = warp & 3;               // TMEM lane quarter this warp may access
    const int j = (warp - 4) >> 2;        // which 128 columns of a half
    const uint32_t r = q * 32 + lane;     // row inside the tile
    const uint32_t tlane = (uint32_t)(q * 32) << 16;
    const uint32_t a_base = s_base + C::kSmemA;
    uint32_t par_d0 = 0, par_d1 = 0, par_lf = 0, par_s = 1, par_sh = 1;   // par_s / par_sh: first wait passes (nothing copied yet)
    auto stash_ready = [&](uint64_t* bar) {       // after publish(): this warp's A stores are done
      if (TRAIN && lane == 0) mbar_arrive(bar);
    };
    long long dbg_d0 = 0, dbg_d1 = 0, dbg_rd = 0, dbg_free = 0, dbg_st = 0, dbg_h1 = 0;
    float dw8a = 0.f, dw8b = 0.f;           // dW8 columns 2t, 2t+1 (fused-loss TRAIN)
    const long long dbg_e0 = DSDF_DBG_ON(a) ? clock64() : 0;
    // generic-proxy writes of this warp -> visible to the (leader's) MMA, one arrive per warp
    // (each SM's tensor core reads its OWN A tile, so cta-scope release + the proxy fence order
    // the data; only the arrival itself travels to the leader)
    auto publish = [&](uint32_t bar_c) {
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(reinterpret_cast<uint64_t*>(smem) + (bar_c - s_base) / 8);
        else { __threadfence_block(); mbar_arrive_cluster_relaxed(bar_c); }
      }
    };
    // this thread's input row of a tile: (x, y, z, gt) -- or the recompute pass's (x', y', z', eps) -- fetched one
    // tile ahead so the global-load latency is not part of the tile boundary
    auto fetch_row = [&](long long it_) -> float4 {
      float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
      if (t < kTileM && it_ < n_iter) {
        long long tl = it_ * gridDim.x + blockIdx.x;
        tl = tl < n_tiles ? tl : n_tiles - 1;
        long long row = tl * kTileM + t;
        row = row < a.n_rows ? row : a.n_rows - 1;
        if (TAN) e = __ldg(a.eik_row + row);
        else load_xyz<MODE>(a, row, e.x, e.y, e.z, e.w);
      }
      return e;
    };
    // dW8[c] += sum_r g[r] X8[r, c] straight from the A tile (train_deep_sdf.py:731 for lin8.weight): thread t owns
    // columns 2t, 2t+1; a warp reads one 128-byte swizzled row per step (conflict-free).  Runs AFTER the next tile's
    // prologue has been published, i.e. under the next tile's layer-0 MMAs.
    int dw8_pending = 0;                    // 1: accumulate the finished tile, 2: only the barrier (padding tile)
    auto dw8_flush = [&]() {
      if (!dw8_pending) return;
      if (dw8_pending == 1) {
        const uint32_t c0 = (uint32_t)(2 * t) & 63u;
        const uint32_t sub = a_base + (uint32_t)(t >> 5) * kSubBytes;
#pragma unroll 8
        for (uint32_t rr = 0; rr < (uint32_t)kTileM; ++rr) {
          uint32_t wv;
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(wv) : "r"(sub + sub_off(rr, c0)));
          const float gr = gsm[rr];
          dw8a = fmaf(gr, __uint_as_float(wv << 16), dw8a);
          dw8b = fmaf(gr, __uint_as_float(wv & 0xFFFF0000u), dw8b);
        }
      }
      named_bar_sync(1, kActThreads);       // the tile in flight rewrites gsm and the A tile
      dw8_pending = 0;
    };
    float4 row_cur = fetch_row(0);
    for (long long it = 0; it < n_iter; ++it) {
      long long tile = it * gridDim.x + blockIdx.x;
      const bool tile_valid = tile < n_tiles;
      tile = tile_valid ? tile : n_tiles - 1;
      const long long row0 = tile * kTileM;
#ifdef DSDF_ARCH_SIREN
      uint8_t* aux_tile = TRAIN ? reinterpret_cast<uint8_t*>(a.relu_mask) + (size_t)tile * kActAuxTileBytes : nullptr;
      // sin(30 x) of 32 accumulator columns -> packed bf16 (+ derivative stash for the backward chain)
      auto act_pack32 = [&](const uint32_t (&v)[32], int l, uint32_t col, uint32_t* p) -> uint32_t {
        uint32_t d[16];
        sine_pack32(v, p, d, nullptr);
        if (TRAIN && tile_valid) deriv_store(aux_tile, l, (int)(col >> 5), r, d);
        return 0u;
      };
      auto load_masks = [](int) {};        // (the recompute pass exists for the ReLU family only)
#else
      uint32_t* mask_tile = TRAIN ? a.relu_mask + (size_t)tile * (8 * 16 * kTileM) + r : nullptr;
      [[maybe_unused]] uint32_t mk[8];     // TAN: this thread's 8 mask words of the current layer (h0: 0..3, h1: 4..7)
      auto load_masks = [&](int l) {
        if (TAN) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            mk[c] = __ldg(mask_tile + (size_t)(l * 16 + j * 4 + c) * kTileM);
            mk[4 + c] = (l != 3) ? __ldg(mask_tile + (size_t)(l * 16 + 8 + j * 4 + c) * kTileM) : 0u;
          }
        }
      };
      // ReLU (+ dropout) of 32 accumulator columns -> packed bf16; stores and returns the mask word
      auto act_pack32 = [&](const uint32_t (&v)[32], int l, uint32_t col, uint32_t* p) -> uint32_t {
        uint32_t m;
        if (TAN) {                          // the first pass's decision (preloaded per layer), not this pass's sign
          m = mk[((col >> 8) << 2) | ((col >> 5) & 3u)];
          mask_pack32(v, m, p);
          return m;
        }
        if (DROP && ((a.drop.layers >> l) & 1u)) {
          m = pos_mask32(v) & dropout_keep32(a.drop, a.drop.row_base + row0 + r, l, (int)(col >> 5));
          mask_scale_pack32(v, m, a.drop.scale, p);
        } else {
          relu_pack32(v, p);
          m = TRAIN ? pos_mask32(v) : 0u;
        }
        if (TRAIN && tile_valid) mask_tile[(size_t)(l * 16 + (col >> 5)) * kTileM] = m;
        return m;
      };
#endif
      // ---- tile prologue: E_A rows (+ layer-0 latent gather when not folded)
      if (!FOLD) {
        // dense input: the latent goes into A SUB 0..3 -- after the stash warp has copied the previous tile's X8 out of
        // them (layer 0's epilogue below then has nothing to wait for: nothing is stashed in between)
        if (TRAIN) { mbar_wait(s_done, par_s); par_s ^= 1; }
        gather_latent<MODE>(a, row0, a_base, t);
      }
      const float gt = row_cur.w;
      if (t < kTileM) {
        if (TAN) write_e_row_tan(s_base + C::kSmemEA, t, row_cur.x, row_cur.y, row_cur.z, row_cur.w);
        else write_e_row_ns(s_base + C::kSmemEA, t, row_cur.x, row_cur.y, row_cur.z);
      }
      publish(a_lo_c);
      dw8_flush();                          // the previous tile's dW8 share, under this tile's layer-0 MMAs

#pragma unroll 1
      for (int l = 0; l < 7; ++l) {
        const bool has_h1 = (l != 3);
        load_masks(l);
        // ---- half 0: TMEM -> ReLU -> bf16, held in registers while the MMAs still read A
        uint32_t held[64];
        {
          const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
          mbar_wait(&d_done[0], par_d0);
          if (DSDF_DBG_ON(a)) dbg_d0 += clock64() - c0;
        }
        par_d0 ^= 1;
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(tmem + tlane + (uint32_t)(j * 128 + c * 32), v);
          tmem_ld_wait();
          act_pack32(v, l, (uint32_t)(j * 128 + c * 32), held + 16 * c);
        }
        {
          // the layer's last read of A SUB 0..3 has retired (mid-way through the last half's K-steps): the next
          // layer's SUB 0..3 go in NOW, under the K-steps that still read SUB 4..7, so the next layer's first 16
          // K-steps are issued back to back with this layer's last ones (no layer-boundary bubble)
          const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
          mbar_wait(lo_free, par_lf);
          if (DSDF_DBG_ON(a)) dbg_d1 += clock64() - c0;
          par_lf ^= 1;
          tc_fence_after();
        }
        if (TRAIN && (FOLD || l > 0)) {     // ... and the stash warp has copied the old SUB 0..3
          const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
          mbar_wait(s_done, par_s);
          par_s ^= 1;
          if (DSDF_DBG_ON(a)) dbg_rd += clock64() - c0;
        }
        // SUB by SUB; warps 4..7 own SUB 0,1 (columns 0..127), warps 8..11 own SUB 2,3
        tc_fence_before();
        if (j == 0) {
          store_chunk32(a_base, r, 0u, held);
          store_chunk32(a_base, r, 32u, held + 16);
          publish(a_lo_c);
          store_chunk32(a_base, r, 64u, held + 32);
          store_chunk32(a_base, r, 96u, held + 48);
          publish(a_sub_c);
        } else {
          publish(a_lo_c);                  // this warp's h0 columns left TMEM long ago; SUB 0 is not its business
          store_chunk32(a_base, r, 128u, held);
          store_chunk32(a_base, r, 160u, held + 16);
          publish(a_sub_c + 8);
          store_chunk32(a_base, r, 192u, held + 32);
          store_chunk32(a_base, r, 224u, held + 48);
          publish(a_sub_c + 16);
        }
        stash_ready(s_lo);
        // ---- half 1: straight from TMEM into SUB 4..7 once every MMA of layer l has retired
        if (has_h1) {
          const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
          mbar_wait(&d_done[1], par_d1);
          par_d1 ^= 1;
          tc_fence_after();
          const long long c_free = DSDF_DBG_ON(a) ? clock64() : 0;   // layer l retired
          if (DSDF_DBG_ON(a)) { dbg_st += c_free - c0; dbg_free += c_free; }
          if (TRAIN && !(FOLD && l == 4)) { mbar_wait(s_done_hi, par_sh); par_sh ^= 1; }   // (slot 3 has no hi half)
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t v[32], p[16];
            tmem_ld_32x32(tmem + tlane + (uint32_t)(256 + j * 128 + c * 32), v);
            tmem_ld_wait();
            act_pack32(v, l, (uint32_t)(256 + j * 128 + c * 32), p);
            store_chunk32(a_base, r, (uint32_t)(256 + j * 128 + c * 32), p);
          }
          tc_fence_before();
          if (f2_in_hi<FOLD>(l + 1)) publish(a_hi_c); else __syncwarp();
          stash_ready(s_hi);
          if (DSDF_DBG_ON(a)) dbg_h1 += clock64() - c_free;
        } else {
          if (DSDF_DBG_ON(a)) dbg_free += clock64();                 // (no h1: layer l retired at d_done[0])
          if (TRAIN) { mbar_wait(s_done_hi, par_sh); par_sh ^= 1; }
          if (!FOLD) {
            // skip connection: latent -> SUB 4..7 (reference cat order [x3 | latent | xyz])
            gather_latent<MODE>(a, row0, a_base + 4 * kSubBytes, t);
            publish(a_hi_c);
            stash_ready(s_hi);
          }
        }
      }

      // ---- lin7 epilogue fused with lin8 (512 -> 1), tanh, loss
      const float4 row_next = fetch_row(it + 1);
      float acc = 0.f;
      if (TAN) {
        // recompute pass: Xt8 = mask7 . (lin7 output) goes through the A tile only for dW8 = g^T Xt8 below
        uint32_t held[64];
        load_masks(7);
        mbar_wait(&d_done[0], par_d0);
        par_d0 ^= 1;
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          const uint32_t col = (uint32_t)(j * 128 + c * 32);
          tmem_ld_32x32(tmem + tlane + col, v);
          tmem_ld_wait();
          act_pack32(v, 7, col, held + 16 * c);
        }
        mbar_wait(lo_free, par_lf);
        par_lf ^= 1;
        tc_fence_after();
        mbar_wait(s_done, par_s);
        par_s ^= 1;
#pragma unroll
        for (int c = 0; c < 4; ++c) store_chunk32(a_base, r, (uint32_t)(j * 128 + c * 32), held + 16 * c);
        __syncwarp();
        stash_ready(s_lo);
        mbar_wait(&d_done[1], par_d1);
        par_d1 ^= 1;
        tc_fence_after();
        mbar_wait(s_done_hi, par_sh);
        par_sh ^= 1;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32], p[16];
          const uint32_t col = (uint32_t)(256 + j * 128 + c * 32);
          tmem_ld_32x32(tmem + tlane + col, v);
          tmem_ld_wait();
          act_pack32(v, 7, col, p);
          store_chunk32(a_base, r, col, p);
        }
        __syncwarp();
        stash_ready(s_hi);
      } else {
        uint32_t held[TRAIN ? 64 : 1];
        mbar_wait(&d_done[0], par_d0);
        par_d0 ^= 1;
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          const uint32_t col = (uint32_t)(j * 128 + c * 32);
          tmem_ld_32x32(tmem + tlane + col, v);
          tmem_ld_wait();
#ifdef DSDF_ARCH_SIREN
          {
            uint32_t pp[16], d[16];
            acc += sine_pack32(v, pp, d, w8s + col);
            if (TRAIN) {
#pragma unroll
              for (int i = 0; i < 16; ++i) held[16 * c + i] = pp[i];
              if (tile_valid) deriv_store(aux_tile, 7, (int)(col >> 5), r, d);
            }
          }
#else
          if (DROP && ((a.drop.layers >> 7) & 1u)) {
            const uint32_t mw = act_pack32(v, 7, col, held + 16 * c);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              acc = fmaf((mw & (0x80000000u >> i)) ? __uint_as_float(v[i]) * a.drop.scale : 0.f, w8s[col + i], acc);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) acc = fmaf(fmaxf(__uint_as_float(v[i]), 0.f), w8s[col + i], acc);
            if (TRAIN) act_pack32(v, 7, col, held + 16 * c);
          }
#endif
        }
        mbar_wait(lo_free, par_lf);
        par_lf ^= 1;
        tc_fence_after();
        if (TRAIN) {   // X8 goes through the A tile only to be stashed (wgrad of lin8, ReLU of lin7)
          mbar_wait(s_done, par_s);
          par_s ^= 1;
#pragma unroll
          for (int c = 0; c < 4; ++c) store_chunk32(a_base, r, (uint32_t)(j * 128 + c * 32), held + 16 * c);
          __syncwarp();
          stash_ready(s_lo);
        }
        mbar_wait(&d_done[1], par_d1);
        par_d1 ^= 1;
        tc_fence_after();
        if (TRAIN) { mbar_wait(s_done_hi, par_sh); par_sh ^= 1; }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          const uint32_t col = (uint32_t)(256 + j * 128 + c * 32);
          tmem_ld_32x32(tmem + tlane + col, v);
          tmem_ld_wait();
#ifdef DSDF_ARCH_SIREN
          {
            uint32_t pp[16], d[16];
            acc += sine_pack32(v, pp, d, w8s + col);
            if (TRAIN) {
              store_chunk32(a_base, r, col, pp);
              if (tile_valid) deriv_store(aux_tile, 7, (int)(col >> 5), r, d);
            }
          }
#else
          if (DROP && ((a.drop.layers >> 7) & 1u)) {
            uint32_t p[16];
            const uint32_t mw = act_pack32(v, 7, col, p);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              acc = fmaf((mw & (0x80000000u >> i)) ? __uint_as_float(v[i]) * a.drop.scale : 0.f, w8s[col + i], acc);
            store_chunk32(a_base, r, col, p);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) acc = fmaf(fmaxf(__uint_as_float(v[i]), 0.f), w8s[col + i], acc);
            if (TRAIN) {
              uint32_t p[16];
              act_pack32(v, 7, col, p);
              store_chunk32(a_base, r, col, p);
            }
          }
#endif
        }
        if (TRAIN) {
          __syncwarp();
          stash_ready(s_hi);
        }
      }
      tc_fence_before();
      if (TAN) {
        if (j == 0) {
          const long long row = row0 + r;
          gsm[r] = (tile_valid && row < a.n_rows) ? __ldg(a.g_tan + row) : 0.f;
        }
        named_bar_sync(1, kActThreads);     // gsm and the A tile (Xt8) are complete
      } else {
      if (j == 1) part[r] = acc;
      named_bar_sync(1, kActThreads);
      if (j == 0) {
        const long long row = row0 + r;
        const bool in_range = tile_valid && row < a.n_rows;
        const bool valid = in_range && (a.valid_per_seg == 0 || (row % a.rows_per_seg) < a.valid_per_seg);
        const float s = acc + part[r] + __ldg(a.b8);
#ifdef DSDF_ARCH_SIREN
        const float th = s;                 // use_tanh=False: no output nonlinearity (siren_decoder.py:233-237)
        const float dth = 1.f;
#else
        const float th = tanhf(s);          // self.th, always (deep_sdf_decoder.py:71,106-107)
        const float dth = 1.f - th * th;
#endif
        if (in_range) a.out[row] = th;
        if (TRAIN && a.fused_loss) {
          // train_deep_sdf.py:565-567,604-606: clamp both, L1(sum)/N_total.
          const float d = a.clamp_dist;
          const float pc = fminf(fmaxf(th, -d), d);
          const float gc = fminf(fmaxf(gt, -d), d);
          const float diff = pc - gc;
          float lossr = valid ? fabsf(diff) : 0.f;
          const float sgn = (diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f);
          const float inside = (th >= -d && th <= d) ? 1.f : 0.f;
          float g = valid ? sgn * inside * dth * a.inv_n_total : 0.f;
          if (in_range) a.gout[row] = g;
          gsm[r] = g;
          if (tile_valid && a.g_img) {
            const __nv_bfloat16 gh = __float2bfloat16_rn(g);
            const __nv_bfloat16 gl = __float2bfloat16_rn(g - __bfloat162float(gh));
            uint8_t* gi = a.g_img + (size_t)tile * kSubBytes;
            *reinterpret_cast<uint4*>(gi + sub_chunk_off(r, 0)) = make_uint4(
                (uint32_t)__bfloat16_as_ushort(gh) | ((uint32_t)__bfloat16_as_ushort(gl) << 16), 0u, 0u, 0u);
#pragma unroll
            for (int ch = 1; ch < 8; ++ch)
              *reinterpret_cast<uint4*>(gi + sub_chunk_off(r, ch)) = make_uint4(0u, 0u, 0u, 0u);
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            lossr += __shfl_xor_sync(0xffffffffu, lossr, o);
            g += __shfl_xor_sync(0xffffffffu, g, o);
          }
          if (lane == 0) { red[q * 2] = lossr; red[q * 2 + 1] = g; }
        }
      }
      }   // !TAN
      if (TAN || (TRAIN && a.fused_loss)) {
        if (!TAN) {
          named_bar_sync(1, kActThreads);
          if (t == 0 && tile_valid) {
            a.tile_partials[tile * 2] = (red[0] + red[2]) + (red[4] + red[6]);
            a.tile_partials[tile * 2 + 1] = (red[1] + red[3]) + (red[5] + red[7]);
          }
        }
        if (a.dw8_partials) dw8_pending = tile_valid ? 1 : 2;
      } else {
        named_bar_sync(1, kActThreads);   // part[] is rewritten by the next tile
      }
      row_cur = row_next;
    }
    dw8_flush();
    if ((TAN || (TRAIN && a.fused_loss)) && a.dw8_partials) {
      float* o = a.dw8_partials + (size_t)blockIdx.x * a.dw8_stride;
      o[2 * t] = dw8a;
      o[2 * t + 1] = dw8b;
    }
    if (DSDF_DBG_ON(a) && t == 0) {
      unsigned long long* d = a.dbg + (size_t)blockIdx.x * kDbgSlots;
      d[5] = dbg_d0; d[6] = dbg_d1; d[7] = clock64() - dbg_e0; d[9] = dbg_rd;
      d[10] = dbg_free; d[12] = dbg_st; d[13] = dbg_h1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CG > 1) cluster_sync_all();   // no CTA exits (or frees TMEM) while the pair is still running
  if (warp == 2) {
    if (CG > 1) tmem_dealloc2<512>(tmem); else tmem_dealloc<512>(tmem);
  }
}

}  // namespace dsdf

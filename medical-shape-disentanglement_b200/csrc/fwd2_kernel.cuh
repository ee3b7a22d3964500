// Fused 9-layer DeepSDF forward chain, v2: one persistent CTA per SM, 128 rows per tile,
// optional thread-block cluster (CL CTAs) sharing the weight stream by TMA-engine multicast.
//
//   warp 0      : weight producer: 8 KB K-steps ([256 n x 16 k], layout.cuh) global -> smem ring;
//                 with CL > 1 every CTA loads 1/CL of each step and multicasts it to the cluster
//   warp 1      : MMA issuer (one thread): tcgen05.mma cta_group::1 kind::f16, M=128 N=256 K=16.
//                 Each layer is two output halves h0/h1 = two 256-column TMEM accumulators, so
//                 the epilogue of one half runs under the MMAs of the other
//   warp 2      : TMEM allocator (512 columns), warp 3 idle
//   warps 4..11 : epilogue warps.  h0's ReLU'd bf16 activations are HELD IN REGISTERS until the
//                 layer's last MMA has retired (the MMAs still read the old A tile), then written
//                 to smem as the next layer's A tile; h1 follows straight from TMEM.  The next
//                 layer's first 16 K-steps only need the h0 part, so they start while the h1
//                 epilogue is still running.
//
// Fold mode (FOLD = true: every row of a tile shares one latent -- scene-contiguous training
// batches, decode_sdf, the create_mesh grid): W[:, latent] z + b is a per-segment bias, computed
// once per step in fp32 (k_scene_fold) and fed through the E step, so lin0 is a single K=16 MMA
// per half and lin4 only contracts the 256 x3 columns.  Executed MACs/row: 1 573 376 (nominal
// 1 835 520; roofline numerators always use the nominal count).
//
// Replaces, per 128-row tile, reference networks/deep_sdf_decoder.py:74-109 (9x addmm, 2x cat,
// 8x relu, tanh), deep_sdf/utils.py:86-97 (expand + cat) and, in TRAIN mode,
// train_deep_sdf.py:598-606 (gather, cat, clamp, L1/N).
#pragma once
#include "fwd_kernel.cuh"
#include "layout.cuh"
#include "ptx.cuh"

// per-CTA cycle counters are compiled in only with -DDSDF_DBG_BUILD (tools/gpu_dbg.py builds a
// separate library); the product build carries no instrumentation.
#ifdef DSDF_DBG_BUILD
#define DSDF_DBG_ON(a) ((a).dbg != nullptr)
#else
#define DSDF_DBG_ON(a) false
#endif
// timing ablations for diagnostics builds only (results are wrong): bit0 = producer skips the
// weight copies, bit1 = MMA issuer skips the MMAs
#ifndef DSDF_ABLATE
#define DSDF_ABLATE 0
#endif

namespace dsdf {

constexpr int kF2Stages = 11;
constexpr int kF2Threads = 384;
constexpr uint32_t kF2SmemA = 0;                                         // 8 SUBs (SWIZZLE_128B)
constexpr uint32_t kF2SmemEA = kTileBytes;                               // [128 x 16] no-swizzle, 4 KB
constexpr uint32_t kF2SmemRing = kF2SmemEA + 4096;
constexpr uint32_t kF2SmemAux = kF2SmemRing + kF2Stages * kStepBytes;    // 225280
constexpr uint32_t kF2AuxBars = 0;      // 32 x 8 B
constexpr uint32_t kF2AuxTmem = 256;
constexpr uint32_t kF2AuxPart = 272;    // 128 floats
constexpr uint32_t kF2AuxRed = 784;     // 16 floats
constexpr uint32_t kF2AuxW8 = 848;      // 512 floats
constexpr uint32_t kF2SmemBytes = kF2SmemAux + kF2AuxW8 + 2048;          // 228176

template <bool FOLD>
__host__ __device__ constexpr int f2_used_ksteps(int l) {
  return FOLD ? (l == 0 ? 0 : (l == kSkipLayer ? 16 : 32)) : w2_ksteps(l);
}
// does layer l contract the "hi" half (SUB 4..7) of the A tile?
template <bool FOLD>
__host__ __device__ constexpr bool f2_in_hi(int l) { return f2_used_ksteps<FOLD>(l) > 16; }

// A-side E row, no-swizzle [128 x 16]: [xh yh zh xl yl zl xh yh | zh 1 1 0 0 0 0 0]
__device__ __forceinline__ void write_e_row_ns(uint32_t e_base, uint32_t r, float x, float y,
                                               float z) {
  const __nv_bfloat16 xh = __float2bfloat16_rn(x), yh = __float2bfloat16_rn(y),
                      zh = __float2bfloat16_rn(z);
  const __nv_bfloat16 xl = __float2bfloat16_rn(x - __bfloat162float(xh));
  const __nv_bfloat16 yl = __float2bfloat16_rn(y - __bfloat162float(yh));
  const __nv_bfloat16 zl = __float2bfloat16_rn(z - __bfloat162float(zh));
  auto u = [](__nv_bfloat16 b) -> uint32_t { return (uint32_t)__bfloat16_as_ushort(b); };
  const uint32_t one = 0x3F80u;
  const uint32_t base = e_base + (r >> 3) * 256u + (r & 7u) * 16u;
  st_shared_v4(base, u(xh) | (u(yh) << 16), u(zh) | (u(xl) << 16), u(yl) | (u(zl) << 16),
               u(xh) | (u(yh) << 16));
  st_shared_v4(base + 128u, u(zh) | (one << 16), one, 0u, 0u);
}

// 32 fp32 accumulator columns -> ReLU -> 16 packed bf16x2
__device__ __forceinline__ void relu_pack32(const uint32_t (&v)[32], uint32_t* p) {
#pragma unroll
  for (int i = 0; i < 16; ++i)
    p[i] = pack_relu_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
}
// 16 packed words = columns col..col+31 of row r -> A tile (SWIZZLE_128B SUBs)
__device__ __forceinline__ void store_chunk32(uint32_t a_base, uint32_t r, uint32_t col,
                                              const uint32_t* p) {
  const uint32_t sub = a_base + (col >> 6) * kSubBytes;
  const uint32_t ch0 = (col & 63) >> 3;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    st_shared_v4(sub + sub_chunk_off(r, ch0 + i), p[4 * i], p[4 * i + 1], p[4 * i + 2],
                 p[4 * i + 3]);
}

template <int MODE, bool FOLD, bool TRAIN, int CL>
__global__ void __launch_bounds__(kF2Threads, 1) k_forward2(const FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kF2SmemAux + kF2AuxBars);
  uint64_t* full = bars;                    // [kF2Stages]
  uint64_t* empty = bars + kF2Stages;       // [kF2Stages]
  uint64_t* a_lo = bars + 2 * kF2Stages;    // epilogue -> MMA: A SUB 0..3 (+E_A) ready, TMEM h0 drained
  uint64_t* a_hi = a_lo + 1;                // epilogue -> MMA: A SUB 4..7 ready, TMEM h1 drained
  uint64_t* d_done = a_lo + 2;              // [2] MMA -> epilogue: half h accumulated
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kF2SmemAux + kF2AuxTmem);
  float* part = reinterpret_cast<float*>(smem + kF2SmemAux + kF2AuxPart);
  float* red = reinterpret_cast<float*>(smem + kF2SmemAux + kF2AuxRed);
  float* w8s = reinterpret_cast<float*>(smem + kF2SmemAux + kF2AuxW8);

  if ((s_base & 1023u) != 0) __trap();
  const uint32_t crank = CL > 1 ? cluster_ctarank() : 0u;
  constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1u);

  if (threadIdx.x == 0) {
    for (int i = 0; i < kF2Stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], CL);
    }
    mbar_init(a_lo, kActThreads);
    mbar_init(a_hi, kActThreads);
    mbar_init(&d_done[0], 1);
    mbar_init(&d_done[1], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  for (int i = threadIdx.x; i < kHidden; i += blockDim.x) w8s[i] = a.w8[i];
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();   // peers' barriers are initialised before any multicast lands
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const long long n_tiles = (a.n_rows + kTileM - 1) / kTileM;
  // every CTA of a cluster runs the same number of iterations (the ring is cluster-synchronous);
  // iterations past the last tile recompute the last tile and write nothing.
  const long long n_iter = (n_tiles + gridDim.x - 1) / gridDim.x;

  if (warp == 0) {
    // ===================== weight producer =====================
    // the whole warp runs the loop (warp-uniform control flow keeps the barrier / descriptor
    // arithmetic on the uniform datapath); one elected lane issues the copies
    {
      uint32_t stage = 0, phase = 0;
      long long dbg_we = 0;
      auto load = [&](const uint8_t* src, bool mc) {
        const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
        mbar_wait(&empty[stage], phase ^ 1);
        if (DSDF_DBG_ON(a)) dbg_we += clock64() - c0;
        if (elect_one()) {
          mbar_arrive_expect_tx(&full[stage], kStepBytes);
          uint8_t* dst = smem + kF2SmemRing + stage * kStepBytes;
          if (DSDF_ABLATE & 1) {
            asm volatile("mbarrier.complete_tx.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[stage])), "r"((uint32_t)kStepBytes) : "memory");
          } else if (CL > 1 && mc) {
            constexpr uint32_t slice = kStepBytes / CL;
            bulk_g2s_mc(dst + crank * slice, src + crank * slice, slice, &full[stage], kMask);
          } else {
            bulk_g2s(dst, src, kStepBytes, &full[stage]);
          }
        }
        __syncwarp();
        if (++stage == kF2Stages) { stage = 0; phase ^= 1; }
      };
      for (long long it = 0; it < n_iter; ++it) {
        long long tile = it * gridDim.x + blockIdx.x;
        tile = tile < n_tiles ? tile : n_tiles - 1;
        const uint8_t* se = nullptr;
        if (FOLD) {
          const long long seg = (MODE == MODE_TABLE && a.seg_ids) ? (tile * kTileM) / a.rows_per_seg : 0;
          se = a.scene_e + (size_t)seg * kSceneESteps * kStepBytes;
        }
#pragma unroll 1
        for (int l = 0; l < 8; ++l) {
          const int used = f2_used_ksteps<FOLD>(l);
          for (int h = 0; h < w2_halves(l); ++h) {
            const uint8_t* w = a.w2img + (size_t)w2_step(l, h, 0) * kStepBytes;
            if (FOLD && (l == 0 || l == kSkipLayer))
              load(se + (size_t)((l == 0 ? 0 : 2) + h) * kStepBytes, false);
            else
              load(w, true);
            for (int s = 0; s < used; ++s) load(w + (size_t)(1 + s) * kStepBytes, true);
          }
        }
      }
      if (DSDF_DBG_ON(a) && lane == 0) a.dbg[(size_t)blockIdx.x * kDbgSlots + 4] = dbg_we;
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // warp-uniform loop; the same elected lane issues every tcgen05.mma / tcgen05.commit
    {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 256, false, false);
      uint32_t stage = 0, phase = 0, par_lo = 0, par_hi = 0;
      const uint64_t ea_desc = umma_smem_desc_ns(s_base + kF2SmemEA, 128, 256);
      long long dbg_wf = 0, dbg_wl = 0, dbg_wh = 0;
      const long long dbg_t0 = DSDF_DBG_ON(a) ? clock64() : 0;
      for (long long it = 0; it < n_iter; ++it) {
#pragma unroll 1
        for (int l = 0; l < 8; ++l) {
          const int used = f2_used_ksteps<FOLD>(l);
          {
            const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
            mbar_wait(a_lo, par_lo);
            if (DSDF_DBG_ON(a)) dbg_wl += clock64() - c0;
          }
          par_lo ^= 1;
          tc_fence_after();
          for (int h = 0; h < w2_halves(l); ++h) {
            const uint32_t d_addr = tmem + h * 256;
            // E step: xyz (hi/lo) + bias through the ones columns; overwrites the accumulator
            {
              const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
              mbar_wait(&full[stage], phase);
              if (DSDF_DBG_ON(a)) dbg_wf += clock64() - c0;
            }
            tc_fence_after();
            if (elect_one()) {
              if (!(DSDF_ABLATE & 2))
              umma_bf16(d_addr, ea_desc,
                        umma_smem_desc_ns(s_base + kF2SmemRing + stage * kStepBytes, 128, 256), idesc, 0u);
              if (CL > 1) umma_commit_mc(&empty[stage], kMask); else umma_commit(&empty[stage]);
            }
            __syncwarp();
            if (++stage == kF2Stages) { stage = 0; phase ^= 1; }
            for (int s = 0; s < used; ++s) {
              if (h == 0 && s == 16) {   // first K-step that reads A SUB 4..7
                const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
                mbar_wait(a_hi, par_hi);
                if (DSDF_DBG_ON(a)) dbg_wh += clock64() - c0;
                par_hi ^= 1;
                tc_fence_after();
              }
              {
                const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
                mbar_wait(&full[stage], phase);
                if (DSDF_DBG_ON(a)) dbg_wf += clock64() - c0;
              }
              tc_fence_after();
              const uint32_t a_addr = s_base + kF2SmemA + (s >> 2) * kSubBytes + (s & 3) * 32;
              if (elect_one()) {
                if (!(DSDF_ABLATE & 2))
                umma_bf16(d_addr, umma_smem_desc(a_addr, 16, 1024),
                          umma_smem_desc_ns(s_base + kF2SmemRing + stage * kStepBytes, 128, 256), idesc, 1u);
                if (CL > 1) umma_commit_mc(&empty[stage], kMask); else umma_commit(&empty[stage]);
              }
              __syncwarp();
              if (++stage == kF2Stages) { stage = 0; phase ^= 1; }
            }
            if (elect_one()) umma_commit(&d_done[h]);
            __syncwarp();
          }
        }
      }
      if (DSDF_DBG_ON(a) && lane == 0) {
        unsigned long long* d = a.dbg + (size_t)blockIdx.x * kDbgSlots;
        d[0] = clock64() - dbg_t0; d[1] = dbg_wf; d[2] = dbg_wl; d[3] = dbg_wh;
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps =====================
    const int t = threadIdx.x - 128;      // 0..255
    const int q = warp & 3;               // TMEM lane quarter this warp may access
    const int j = (warp - 4) >> 2;        // which 128 columns of a half
    const uint32_t r = q * 32 + lane;     // row inside the tile
    const uint32_t tlane = (uint32_t)(q * 32) << 16;
    const uint32_t a_base = s_base + kF2SmemA;
    uint32_t par_d0 = 0, par_d1 = 0;
    long long dbg_d0 = 0, dbg_d1 = 0;
    const long long dbg_e0 = DSDF_DBG_ON(a) ? clock64() : 0;
    for (long long it = 0; it < n_iter; ++it) {
      long long tile = it * gridDim.x + blockIdx.x;
      const bool tile_valid = tile < n_tiles;
      tile = tile_valid ? tile : n_tiles - 1;
      const long long row0 = tile * kTileM;
      // ---- tile prologue: E_A rows (+ layer-0 latent gather when not folded)
      if (TRAIN && !FOLD) {   // the previous tile's slot-7 stash may still be reading A
        if (t == 0) bulk_wait_read_all();
        named_bar_sync(1, kActThreads);
      }
      if (!FOLD) gather_latent<MODE>(a, row0, a_base, t);
      float gt = 0.f;
      if (t < kTileM) {
        long long row = row0 + t;
        row = row < a.n_rows ? row : a.n_rows - 1;
        float x, y, z;
        load_xyz<MODE>(a, row, x, y, z, gt);
        write_e_row_ns(s_base + kF2SmemEA, t, x, y, z);
      }
      fence_proxy_async_smem();
      mbar_arrive(a_lo);

#pragma unroll 1
      for (int l = 0; l < 7; ++l) {
        const bool has_h1 = (l != 3);
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
          relu_pack32(v, held + 16 * c);
        }
        if (has_h1) {
          const long long c0 = DSDF_DBG_ON(a) ? clock64() : 0;
          mbar_wait(&d_done[1], par_d1);    // every MMA of layer l has retired: A is free
          if (DSDF_DBG_ON(a)) dbg_d1 += clock64() - c0;
          par_d1 ^= 1;
          tc_fence_after();
        }
        if (TRAIN) {                        // ... and the stash copy of A has been read out
          if (t == 0) bulk_wait_read_all();
          named_bar_sync(1, kActThreads);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) store_chunk32(a_base, r, (uint32_t)(j * 128 + c * 32), held + 16 * c);
        tc_fence_before();
        fence_proxy_async_smem();
        mbar_arrive(a_lo);
        // ---- half 1: straight from TMEM into SUB 4..7
        if (has_h1) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t v[32], p[16];
            tmem_ld_32x32(tmem + tlane + (uint32_t)(256 + j * 128 + c * 32), v);
            tmem_ld_wait();
            relu_pack32(v, p);
            store_chunk32(a_base, r, (uint32_t)(256 + j * 128 + c * 32), p);
          }
          tc_fence_before();
          fence_proxy_async_smem();
          if (f2_in_hi<FOLD>(l + 1)) mbar_arrive(a_hi);
        } else if (!FOLD) {
          // skip connection: latent -> SUB 4..7 (reference cat order [x3 | latent | xyz])
          gather_latent<MODE>(a, row0, a_base + 4 * kSubBytes, t);
          fence_proxy_async_smem();
          mbar_arrive(a_hi);
        }
        if (TRAIN) {
          // stash X_{l+1} (input of lin(l+1)); for l == 3 only the x3 part (SUB 0..3)
          named_bar_sync(1, kActThreads);
          if (t == 0 && tile_valid) {
            const uint32_t bytes = (l == 3) ? 4 * kSubBytes : kTileBytes;
            bulk_s2g(a.xstash + (size_t)tile * kStashTileBytes + (size_t)l * kTileBytes,
                     smem + kF2SmemA, bytes);
            bulk_commit();
          }
        }
      }

      // ---- lin7 epilogue fused with lin8 (512 -> 1), tanh, loss
      float acc = 0.f;
      {
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
#pragma unroll
          for (int i = 0; i < 32; ++i) acc = fmaf(fmaxf(__uint_as_float(v[i]), 0.f), w8s[col + i], acc);
          if (TRAIN) relu_pack32(v, held + 16 * c);
        }
        mbar_wait(&d_done[1], par_d1);
        par_d1 ^= 1;
        tc_fence_after();
        if (TRAIN) {
          if (t == 0) bulk_wait_read_all();
          named_bar_sync(1, kActThreads);
#pragma unroll
          for (int c = 0; c < 4; ++c) store_chunk32(a_base, r, (uint32_t)(j * 128 + c * 32), held + 16 * c);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          const uint32_t col = (uint32_t)(256 + j * 128 + c * 32);
          tmem_ld_32x32(tmem + tlane + col, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) acc = fmaf(fmaxf(__uint_as_float(v[i]), 0.f), w8s[col + i], acc);
          if (TRAIN) {
            uint32_t p[16];
            relu_pack32(v, p);
            store_chunk32(a_base, r, col, p);
          }
        }
      }
      tc_fence_before();
      if (j == 1) part[r] = acc;
      if (TRAIN) fence_proxy_async_smem();
      named_bar_sync(1, kActThreads);
      if (TRAIN && t == 0 && tile_valid) {
        bulk_s2g(a.xstash + (size_t)tile * kStashTileBytes + (size_t)7 * kTileBytes,
                 smem + kF2SmemA, kTileBytes);
        bulk_commit();
      }
      if (j == 0) {
        const long long row = row0 + r;
        const bool valid = tile_valid && row < a.n_rows;
        const float s = acc + part[r] + __ldg(a.b8);
        const float th = tanhf(s);
        if (valid) a.out[row] = th;
        if (TRAIN && a.fused_loss) {
          // train_deep_sdf.py:565-567,604-606: clamp both, L1(sum)/N_total.
          const float d = a.clamp_dist;
          const float pc = fminf(fmaxf(th, -d), d);
          const float gc = fminf(fmaxf(gt, -d), d);
          const float diff = pc - gc;
          float lossr = valid ? fabsf(diff) : 0.f;
          const float sgn = (diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f);
          const float inside = (th >= -d && th <= d) ? 1.f : 0.f;
          float g = valid ? sgn * inside * (1.f - th * th) * a.inv_n_total : 0.f;
          if (valid) a.gout[row] = g;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            lossr += __shfl_xor_sync(0xffffffffu, lossr, o);
            g += __shfl_xor_sync(0xffffffffu, g, o);
          }
          if (lane == 0) { red[q * 2] = lossr; red[q * 2 + 1] = g; }
        }
      }
      if (TRAIN && a.fused_loss) {
        named_bar_sync(1, kActThreads);
        if (t == 0 && tile_valid) {
          a.tile_partials[tile * 2] = (red[0] + red[2]) + (red[4] + red[6]);
          a.tile_partials[tile * 2 + 1] = (red[1] + red[3]) + (red[5] + red[7]);
        }
      } else {
        named_bar_sync(1, kActThreads);   // part[] is rewritten by the next tile
      }
    }
    if (TRAIN && t == 0) bulk_wait_all();
    if (DSDF_DBG_ON(a) && t == 0) {
      unsigned long long* d = a.dbg + (size_t)blockIdx.x * kDbgSlots;
      d[5] = dbg_d0; d[6] = dbg_d1; d[7] = clock64() - dbg_e0;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();   // no CTA exits while a peer can still multicast into it
  if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace dsdf

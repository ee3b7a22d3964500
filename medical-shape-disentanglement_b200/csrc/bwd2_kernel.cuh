// Fused backward dgrad chain, v2 (scene mode): persistent CTAs, 128 rows per CTA tile, the same
// lean issue loops / CTA pairing (tcgen05 cta_group::2) as the v3 forward chain.
//
//   warp 0       : weight producer: K-steps of W^T ([256 i x 16 o], layout.cuh dg2_*) -> smem ring
//   warp 1       : MMA issuer (leader; the peer's warp 1 idles):  dX_l = dY_l W_l as two
//                  256-column halves, M = 128*CG, N = 256, K = 16
//   warp 2       : TMEM allocator; warp 3 idle
//   warps 4..11  : epilogue warps: dY7 = g w8 (X8 > 0) from the forward's dL/ds and ReLU bit-mask;
//                  per layer TMEM -> (X_l > 0) mask -> bf16; half 0 is held in registers until the
//                  layer's last MMA has retired, then becomes the next A tile; dY stash for wgrad
//   warps 12..15 : reducer warps (in the shadow of the MMAs): column sums of every dY tile
//                  -> bias grads, per-tile colsums of dY0/dY4 (-> latent grads and the latent
//                  columns of dW0/dW4 by the segmented kernels), xyz columns of dW0/dW4.
//                  (dW8 is a wgrad item, wgrad2_kernel.cuh)
//
// The ReLU masks come from the forward kernel's 1-bit/element image (8 KB per layer per tile)
// instead of re-reading the 128 KB bf16 activation tiles.
//
// Replaces the autograd backward of reference train_deep_sdf.py:731 for the decoder
// (9 dgrad GEMMs, ReLU masks, tanh', bias grads).  In scene mode the latent/xyz columns of
// dX0 and dX4 are never materialised: sum_rows dX[:, latent] = colsum(dY) * W[:, latent].
#pragma once
#include "common.cuh"
#include "fwd3_kernel.cuh"
#include "layout.cuh"
#include "ptx.cuh"

namespace dsdf {

struct alignas(64) Bwd2Args {
  CUtensorMap tmap_dg;        // CTA-pair launches: tensor map over dg2img (common.cuh kMapRowBytes)
  long long n_rows;
  const uint8_t* dg2img;      // dgrad K-step stream
  const float* w8;            // [512]
  const float* g;             // [n] dL/ds
  const uint32_t* relu_mask;  // forward ReLU bit-masks [tiles][8][16][128]
  uint8_t* dystash;           // out: dY_l images, slot l (1..7)
  const float* xyz;           // xyz of row r at xyz[r * xyz_stride + 0..2]
  long long xyz_stride;
  float* tile_colsum;         // [tiles][2][512]: colsum(dY0), colsum(dY4)
  float* cta_partials;        // [grid][kCtaPartialFloats]
  DropoutArgs drop;           // DROP: layers + scale (the keep bits already sit in relu_mask)
  int stash_dy0;              // dense-input backward: also stash dY0 (slot 0) for the dInput / dW0 kernels
  int skip_dy7_stash;         // the wgrad kernel rebuilds dY7 itself (wgrad2_kernel.cuh REBUILD): slot 7 is not written
  // EIK: the Eikonal term's per-row quantities (train_deep_sdf.py:724-729; algebra in tests/eikonal_algebra.py)
  const float* pred;          // [n] tanh output of the forward pass
  const float4* wx;           // [2][512] fp32 (x, y, z, 0) xyz columns of lin0 / lin4 (k_fold_images)
  float clamp_dist;
  float eik_c2;               // 2 * lambda / N_total
  float4* eik_row;            // out [n]: perturbed input of the recompute pass (x', y', z', eps)
  float* eik_tile;            // out [tiles][2]: sum_rows (1 - ||G||)^2, sum_rows eps * g
};

template <int CG>
struct B2 {
  static constexpr int kRing = 8 * CG;
  static constexpr uint32_t kSlice = kStepBytes / CG;
  static constexpr uint32_t kSmemA = 0;
  static constexpr uint32_t kSmemRing = kTileBytes;
  static constexpr uint32_t kSmemAux = kSmemRing + kRing * kSlice;   // + 64 KB
  static constexpr uint32_t kAuxTmem = 512;     // after 64 barriers
  static constexpr uint32_t kAuxW8 = 528;       // 512 floats
  static constexpr uint32_t kAuxG = kAuxW8 + 2048;      // 128 floats
  static constexpr uint32_t kAuxXyz = kAuxG + 512;      // 128 x 3 floats
  static constexpr uint32_t kAuxWx = kAuxXyz + 1536;    // EIK: 2 x 512 float4
  static constexpr uint32_t kAuxJ = kAuxWx + 16384;     // EIK: 128 float4 (partial d s / d xyz of the other column half)
  static constexpr uint32_t kAuxRho = kAuxJ + 2048;     // EIK: 128 floats (1 + eps)
  static constexpr uint32_t kAuxRed = kAuxRho + 512;    // EIK: 16 floats
  static constexpr uint32_t kSmemBytes = kSmemAux + kAuxRed + 64;
};
constexpr int kB2Threads = 512;

__host__ __device__ constexpr bool b2_has_h1(int l) { return l != kSkipLayer; }   // scene mode: lin4 -> x3 part only
__host__ __device__ constexpr bool b2_in_hi(int l) { return dg2_ksteps(l) > 16; }

template <int N>
__device__ __forceinline__ void setmaxnreg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N));
}
template <int N>
__device__ __forceinline__ void setmaxnreg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N));
}

#ifdef DSDF_ARCH_SIREN
// 32 fp32 dX columns * stashed activation derivative 30 cos(30 x) -> 16 packed bf16x2.  The stash holds
// round(127 cos) as signed bytes (sine_pack32, fwd3_kernel.cuh): byte b -> float by the exponent trick
// float(0x4B400000 | (b ^ 0x80)) - (1.5 * 2^23 + 128), no conversion instruction; the factor 30 / 127 rides on dX.
__device__ __forceinline__ void deriv_mul_pack32(const uint32_t (&v)[32], const uint32_t* d, uint32_t* p) {
  constexpr float kScale = 30.f / 127.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const uint32_t w = d[i >> 1] ^ 0x80808080u;
    const uint32_t b0 = __byte_perm(w, 0x4B400000u, (i & 1) ? 0x7652 : 0x7650);   // [w.b(2 or 0), 00, 40, 4B]
    const uint32_t b1 = __byte_perm(w, 0x4B400000u, (i & 1) ? 0x7653 : 0x7651);
    const float c0 = __uint_as_float(b0) - 12583040.f, c1 = __uint_as_float(b1) - 12583040.f;
    p[i] = pack_bf16x2(__uint_as_float(v[2 * i]) * kScale * c0, __uint_as_float(v[2 * i + 1]) * kScale * c1);
  }
}
#endif

// partial d s / d xyz of one row over 32 columns: j += sum_i (mask_i ? v_i : 0) * w[i].xyz  (w in shared memory,
// the same address for the whole warp: broadcast loads)
__device__ __forceinline__ void j_acc32(const uint32_t (&v)[32], uint32_t m, const float4* w, float& jx, float& jy,
                                        float& jz) {
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const float d = (m & (0x80000000u >> i)) ? __uint_as_float(v[i]) : 0.f;
    const float4 ww = w[i];
    jx = fmaf(d, ww.x, jx);
    jy = fmaf(d, ww.y, jy);
    jz = fmaf(d, ww.z, jz);
  }
}

// EIK: additionally accumulates, per row, j = g (d s / d xyz) from dY4 and dY0, turns it into the Eikonal term's
// scalars at the end of the tile (||G||, eps, xyz'), writes them for the recompute pass and lets the reducers
// weight the layer-0 sums with them (the sums of layers 1..7 are corrected by k_eik_corr).
template <int CG, bool DROP = false, bool EIK = false>
__global__ void __launch_bounds__(kB2Threads, 1) k_backward2(const __grid_constant__ Bwd2Args a) {
  static_assert(!(EIK && DROP), "the Eikonal pass is not built together with dropout");
  using C = B2<CG>;
  constexpr int kRing = C::kRing;
  constexpr uint32_t kSlice = C::kSlice;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kSmemAux);
  uint64_t* full = bars;                 // [kRing]
  uint64_t* empty = bars + 16;           // [kRing]
  uint64_t* a_lo = bars + 32;            // epilogue -> MMA: dY columns 0..255 ready
  uint64_t* a_hi = bars + 33;            // epilogue -> MMA: dY columns 256..511 ready
  uint64_t* d_done = bars + 34;          // [2] MMA -> epilogue
  uint64_t* dy_ready = bars + 36;        // epilogue -> reducers: dY tile complete in A
  uint64_t* red_done = bars + 37;        // reducers -> epilogue: A may be overwritten
  uint64_t* s_lo = bars + 38;            // epilogue -> stash warp: dY columns 0..255 written
  uint64_t* s_hi = bars + 39;            // epilogue -> stash warp: dY columns 256..511 written
  uint64_t* s_done = bars + 40;          // stash warp -> epilogue: dY columns 0..255 copied to HBM
  uint64_t* s_done_hi = bars + 41;       // stash warp -> epilogue: dY columns 256..511 copied to HBM
  uint64_t* red_done_hi = bars + 42;     // reducer warps 14,15 (columns 256..511) -> epilogue (red_done: warps 12,13)
  uint64_t* lo_free = bars + 43;         // MMA -> epilogue: the layer's last read of A columns 0..255 has retired
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + C::kSmemAux + C::kAuxTmem);
  float* w8s = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxW8);
  float* gs = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxG);
  float* xyzs = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxXyz);
  [[maybe_unused]] float4* wxs = reinterpret_cast<float4*>(smem + C::kSmemAux + C::kAuxWx);
  [[maybe_unused]] float4* jpart = reinterpret_cast<float4*>(smem + C::kSmemAux + C::kAuxJ);
  [[maybe_unused]] float* rhos = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxRho);
  [[maybe_unused]] float* ered = reinterpret_cast<float*>(smem + C::kSmemAux + C::kAuxRed);

  if ((s_base & 1023u) != 0) __trap();
  const uint32_t crank = CG > 1 ? cluster_ctarank() : 0u;
  const bool leader = crank == 0;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&full[i], 1);              // one arrive.expect_tx by the leader's producer for both CTAs' bytes
      mbar_init(&empty[i], 1);
    }
    mbar_init(a_lo, 8 * CG);
    mbar_init(a_hi, 8 * CG);
    mbar_init(&d_done[0], 1);
    mbar_init(&d_done[1], 1);
    mbar_init(dy_ready, 8);
    mbar_init(red_done, 2);
    mbar_init(red_done_hi, 2);
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
  if (EIK)
    for (int i = threadIdx.x; i < 2 * kHidden; i += blockDim.x) wxs[i] = a.wx[i];
  tc_fence_before();
  __syncthreads();
  if (CG > 1) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  grid_dep_launch();                     // programmatic dependent launch, see k_forward3
  grid_dep_wait();

  const long long n_tiles = (a.n_rows + kTileM - 1) / kTileM;
  const long long n_iter = (n_tiles + gridDim.x - 1) / gridDim.x;
  const uint32_t full_c = CG > 1 ? mapa_u32(smem_u32(full), 0) : smem_u32(full);
  const uint32_t a_lo_c = CG > 1 ? mapa_u32(smem_u32(a_lo), 0) : smem_u32(a_lo);
  const uint32_t a_hi_c = CG > 1 ? mapa_u32(smem_u32(a_hi), 0) : smem_u32(a_hi);

  if (warp < 4) {
    setmaxnreg_dec<56>();
    if (warp == 0) {
      // ===================== weight producer =====================
      if (elect_one()) {
        uint32_t ph = 0;
        for (long long it = 0; it < n_iter; ++it) {
#pragma unroll 1
          for (int l = 7; l >= 1; --l) {
            const int groups = dg2_ksteps(l) / kRing;
            const int halves = b2_has_h1(l) ? 2 : 1;
#pragma unroll 1
            for (int h = 0; h < halves; ++h) {
              const long long w_off = (long long)dg2_step(l, h, 0) * kStepBytes + crank * kSlice;
              const uint8_t* w = a.dg2img + w_off;
              int y = (int)(w_off / kMapRowBytes);
#pragma unroll 1
              for (int g = 0; g < groups; ++g) {
#pragma unroll
                for (int u = 0; u < kRing; ++u) {
                  mbar_wait(&empty[u], ph ^ 1);
                  if (CG > 1) {
                    // each CTA's half of the K-step reports to the LEADER's barrier (UTMALDG.2CTA): no relay hop
                    if (leader) mbar_arrive_expect_tx(&full[u], CG * kSlice);
                    tma2_load_2d(s_base + C::kSmemRing + u * kSlice, &a.tmap_dg, 0, y, full_c + u * 8);
                    y += (int)(kStepBytes / kMapRowBytes);
                  } else {
                    mbar_arrive_expect_tx(&full[u], kSlice);
                    bulk_g2s(smem + C::kSmemRing + u * kSlice, w, kSlice, &full[u]);
                    w += kStepBytes;
                  }
                }
                ph ^= 1;
              }
            }
          }
        }
      }
    } else if (warp == 1 && leader) {
      // ===================== MMA issuer (leader CTA only; the peer's warp 1 idles) =====================
      if (elect_one()) {
        constexpr uint32_t idesc = umma_idesc_bf16(128 * CG, 256, false, false);
        const uint64_t a_desc0 = umma_smem_desc(s_base + C::kSmemA, 16, 1024);
        const uint64_t b_desc0 = umma_smem_desc_ns(s_base + C::kSmemRing, 128, 256);
        uint32_t ph = 0, par_lo = 0, par_hi = 0;
        auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t acc) {
          if (CG > 1) umma2_bf16(d, ad, bd, idesc, acc); else umma_bf16(d, ad, bd, idesc, acc);
        };
        auto commit = [&](uint64_t* bar) {
          if (CG > 1) umma2_commit_mc(bar, (uint16_t)3); else umma_commit(bar);
        };
        for (long long it = 0; it < n_iter; ++it) {
#pragma unroll 1
          for (int l = 7; l >= 1; --l) {
            const int groups = dg2_ksteps(l) / kRing;
            const int halves = b2_has_h1(l) ? 2 : 1;
            DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 1u);
            mbar_wait(a_lo, par_lo);
            par_lo ^= 1;
            tc_fence_after();
#pragma unroll 1
            for (int h = 0; h < halves; ++h) {
              const uint32_t d_addr = tmem + h * 256;
#pragma unroll 1
              for (int g = 0; g < groups; ++g) {
                DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | (unsigned)(16 + h * 4 + g));
                if (h == 0 && g * kRing == 16) {   // first K-step that reads dY columns 256..511
                  mbar_wait(a_hi, par_hi);
                  par_hi ^= 1;
                  tc_fence_after();
                }
                const uint64_t a_desc = a_desc0 + (uint64_t)((g * (kRing / 4) * kSubBytes) >> 4);
#pragma unroll
                for (int u = 0; u < kRing; ++u) {
                  mbar_wait(&full[u], ph);
                  tc_fence_after();
                  mma(d_addr, a_desc + (uint64_t)(((u >> 2) * kSubBytes + (u & 3) * 32) >> 4),
                      b_desc0 + (uint64_t)((u * kSlice) >> 4), (g > 0 || u > 0) ? 1u : 0u);
                  commit(&empty[u]);
                }
                ph ^= 1;
                // K-steps 0..15 (A columns 0..255) of the layer's last half are issued: once they retire the epilogue
                // may put the next layer's columns 0..255 in place, under K-steps 16..31
                if (h == halves - 1 && (g + 1) * kRing == 16) commit(lo_free);
              }
              commit(&d_done[h]);
            }
          }
        }
      }
    }
    if (warp == 3) {
      // ===================== dY-stash warp =====================
      // dY_7..dY_1 images for the wgrad kernel: byte-for-byte the smem A tile -> linear copy with
      // coalesced 512 B warp stores (see the forward kernel's stash warp for why not TMA).
      const uint64_t pol = stash_policy();
      uint32_t par_lo = 0, par_hi = 0;
      for (long long it = 0; it < n_iter; ++it) {
        long long tile = it * gridDim.x + blockIdx.x;
        const bool tile_valid = tile < n_tiles;
        tile = tile_valid ? tile : n_tiles - 1;
        uint8_t* dy_tile = a.dystash + (size_t)tile * kStashTileBytes;
#pragma unroll 1
        for (int k = 7; k >= (a.stash_dy0 ? 0 : 1); --k) {
          uint4* dst = reinterpret_cast<uint4*>(dy_tile + (size_t)k * kTileBytes);
          const uint32_t src = s_base + C::kSmemA;
#pragma unroll 1
          for (int half = 0; half < 2; ++half) {
            if (half == 1 && k == 3) break;                   // dY3 has 256 columns (and no s_done_hi phase: an arrive
                                                              // not gated by s_hi could lap a slow epilogue warp)
            DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)k << 8) | (unsigned)(1 + half));
            if (half == 0) { mbar_wait(s_lo, par_lo); par_lo ^= 1; }
            else { mbar_wait(s_hi, par_hi); par_hi ^= 1; }
            if (tile_valid && a.dystash && !(k == 7 && a.skip_dy7_stash)) {    // frozen-decoder callers pass no stash (no wgrad)
#pragma unroll 1
              for (int i = 0; i < 128; i += 8) {
                uint4 v[8];
#pragma unroll
                for (int q8 = 0; q8 < 8; ++q8) {
                  const uint32_t off = (uint32_t)(half * 4 * kSubBytes) + (uint32_t)((i + q8) * 32 + lane) * 16u;
                  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                               : "=r"(v[q8].x), "=r"(v[q8].y), "=r"(v[q8].z), "=r"(v[q8].w)
                               : "r"(src + off));
                }
#pragma unroll
                for (int q8 = 0; q8 < 8; ++q8)
                  st_stash(dst + (size_t)(half * 4 * kSubBytes) / 16 + (size_t)(i + q8) * 32 + lane, v[q8], pol);
              }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(half == 0 ? s_done : s_done_hi);
          }
        }
      }
    }
  } else if (warp < 12) {
    // ===================== epilogue warps =====================
    setmaxnreg_inc<176>();
    [[maybe_unused]] const int t = threadIdx.x - 128;      // 0..255
    const int q = warp & 3;               // TMEM lane quarter this warp may access
    const int j = (warp - 4) >> 2;        // which 128 columns of a half
    const uint32_t r = q * 32 + lane;     // row inside the tile
    const uint32_t tlane = (uint32_t)(q * 32) << 16;
    const uint32_t a_base = s_base + C::kSmemA;
    // phase parities of the seven barriers this role waits on, packed into one register (the epilogue is at its
    // register cap): bit 0 d_done[0], 1 d_done[1], 2 red_done, 3 red_done_hi, 4 s_done, 5 s_done_hi, 6 lo_free
    uint32_t pars = 0;
    auto wait_flip = [&](uint64_t* bar, uint32_t bit) {
      mbar_wait(bar, (pars >> bit) & 1u);
      pars ^= 1u << bit;
    };
    auto stash_ready = [&](uint64_t* bar) {   // after this warp's A stores (+ __syncwarp)
      if (lane == 0) mbar_arrive(bar);
    };
    auto publish = [&](uint64_t* bar, uint32_t bar_c) {
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(bar);
        else { __threadfence_block(); mbar_arrive_cluster_relaxed(bar_c); }
      }
    };
    auto tell_reducers = [&]() {
      __syncwarp();
      if (lane == 0) mbar_arrive(dy_ready);
    };
    bool first = true;
    for (long long it = 0; it < n_iter; ++it) {
      long long tile = it * gridDim.x + blockIdx.x;
      const bool tile_valid = tile < n_tiles;
      tile = tile_valid ? tile : n_tiles - 1;
      const long long row0 = tile * kTileM;
      const long long row = row0 + r;
      const bool valid = tile_valid && row < a.n_rows;
#ifdef DSDF_ARCH_SIREN
      const uint8_t* aux_tile = reinterpret_cast<const uint8_t*>(a.relu_mask) + (size_t)tile * kActAuxTileBytes;
#else
      const uint32_t* mask_tile = a.relu_mask + (size_t)tile * (8 * 16 * kTileM) + r;
#endif
      // previous tile: reducers done with dY0, stash stores done reading A
      DSDF_PROG(0, ((unsigned)it << 16) | (9u << 8) | 1u);
      if (!first) {
        wait_flip(red_done, 2u);
        wait_flip(red_done_hi, 3u);
        if (a.stash_dy0) {                  // ... and the stash warp has copied the previous tile's dY0
          wait_flip(s_done, 4u);
          wait_flip(s_done_hi, 5u);
        }
      }
      first = false;
      float gr = valid ? __ldg(a.g + row) : 0.f;
      [[maybe_unused]] const float th = (EIK && valid) ? __ldg(a.pred + row) : 0.f;
      [[maybe_unused]] float jx = 0.f, jy = 0.f, jz = 0.f;         // EIK: this thread's share of g (d s / d xyz)
      if (DROP && ((a.drop.layers >> 7) & 1u)) gr *= a.drop.scale;   // dropout after lin7's ReLU
      if (j == 0) {
        gs[r] = gr;
        const long long rr = row < a.n_rows ? row : a.n_rows - 1;
        xyzs[r * 3 + 0] = __ldg(a.xyz + rr * a.xyz_stride + 0);
        xyzs[r * 3 + 1] = __ldg(a.xyz + rr * a.xyz_stride + 1);
        xyzs[r * 3 + 2] = __ldg(a.xyz + rr * a.xyz_stride + 2);
      }
      // ---- dY7[r,c] = g[r] * w8[c] * (X8[r,c] > 0)
#pragma unroll 1
      for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t col = (uint32_t)(hh * 256 + j * 128 + c * 32);
          uint32_t v[32], p[16];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(gr * w8s[col + i]);
#ifdef DSDF_ARCH_SIREN
          uint32_t d[16];
          deriv_load(aux_tile, 7, hh * 8 + j * 4 + c, r, d);
          deriv_mul_pack32(v, d, p);
#else
          const uint32_t m = __ldg(mask_tile + (size_t)(7 * 16 + hh * 8 + j * 4 + c) * kTileM);
          mask_pack32(v, m, p);
#endif
          store_chunk32(a_base, r, col, p);
        }
      }
      publish(a_lo, a_lo_c);
      publish(a_hi, a_hi_c);
      stash_ready(s_lo);
      stash_ready(s_hi);
      tell_reducers();

#pragma unroll 1
      for (int l = 7; l >= 1; --l) {
        const bool has_h1 = b2_has_h1(l);
        const int lm = l - 1;   // X_l = ReLU output of lin(l-1): forward mask layer l-1
#ifndef DSDF_ARCH_SIREN
        uint32_t mk[8];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          mk[c] = __ldg(mask_tile + (size_t)(lm * 16 + j * 4 + c) * kTileM);
          mk[4 + c] = has_h1 ? __ldg(mask_tile + (size_t)(lm * 16 + 8 + j * 4 + c) * kTileM) : 0u;
        }
#endif
        // ---- half 0: TMEM -> mask -> bf16, held in registers while the MMAs still read A
        uint32_t held[64];
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 2u);
        wait_flip(&d_done[0], 0u);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(tmem + tlane + (uint32_t)(j * 128 + c * 32), v);
          tmem_ld_wait();
#ifdef DSDF_ARCH_SIREN
          uint32_t d[16];
          deriv_load(aux_tile, lm, j * 4 + c, r, d);
          deriv_mul_pack32(v, d, held + 16 * c);
#else
          if (DROP && ((a.drop.layers >> lm) & 1u)) mask_scale_pack32(v, mk[c], a.drop.scale, held + 16 * c);
          else mask_pack32(v, mk[c], held + 16 * c);
          if (EIK && (lm == kSkipLayer || lm == 0))
            j_acc32(v, mk[c], wxs + (lm == 0 ? 0 : kHidden) + j * 128 + c * 32, jx, jy, jz);
#endif
        }
        // columns 0..255 of A are free once the layer's last half has retired its K-steps 0..15 (lo_free), the
        // reducer warps 12,13 have summed them and the stash warp has copied them: the next layer's columns 0..255
        // go in under K-steps 16..31, so the next layer's first MMAs follow this layer's last ones without a bubble
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 3u);
        wait_flip(lo_free, 6u);
        tc_fence_after();
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 4u);
        wait_flip(red_done, 2u);
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 5u);
        wait_flip(s_done, 4u);
#pragma unroll
        for (int c = 0; c < 4; ++c) store_chunk32(a_base, r, (uint32_t)(j * 128 + c * 32), held + 16 * c);
        tc_fence_before();
        if (l - 1 >= 1) { publish(a_lo, a_lo_c); stash_ready(s_lo); }
        else if (a.stash_dy0) { __syncwarp(); stash_ready(s_lo); }
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 6u);
        if (has_h1) {
          wait_flip(&d_done[1], 1u);    // every MMA of layer l has retired: columns 256..511 are free ...
          tc_fence_after();
        }
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 7u);
        wait_flip(red_done_hi, 3u);     // ... the reducers have read them
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 8u);
        if (l != 3) {                       // ... and the stash warp has copied them (dY3 has no such half)
          wait_flip(s_done_hi, 5u);
        }
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 9u);
        if (has_h1) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t v[32], p[16];
            tmem_ld_32x32(tmem + tlane + (uint32_t)(256 + j * 128 + c * 32), v);
            tmem_ld_wait();
#ifdef DSDF_ARCH_SIREN
            uint32_t d[16];
            deriv_load(aux_tile, lm, 8 + j * 4 + c, r, d);
            deriv_mul_pack32(v, d, p);
#else
            if (DROP && ((a.drop.layers >> lm) & 1u)) mask_scale_pack32(v, mk[4 + c], a.drop.scale, p);
            else mask_pack32(v, mk[4 + c], p);
            if (EIK && (lm == kSkipLayer || lm == 0))
              j_acc32(v, mk[4 + c], wxs + (lm == 0 ? 0 : kHidden) + 256 + j * 128 + c * 32, jx, jy, jz);
#endif
            store_chunk32(a_base, r, (uint32_t)(256 + j * 128 + c * 32), p);
          }
          tc_fence_before();
          if (l - 1 >= 1 && b2_in_hi(l - 1)) publish(a_hi, a_hi_c); else __syncwarp();
          if (l - 1 >= 1 || a.stash_dy0) stash_ready(s_hi);
        }
        if (!(EIK && l == 1)) tell_reducers();
      }
      if (EIK) {
        // ---- end of the tile: dY0 sits in A.  Per row (tests/eikonal_algebra.py):
        //   G = sigma J, J = j / g;  coef = c2 (1 - 1/||G||);  eps = -2 t coef ||G||^2 / g;  xyz' = (1+eps) xyz + (sigma/g) coef G
        // The reducers then weight the layer-0 sums: bias / latent column sums with 1 + eps, xyz columns with xyz'.
        if (j == 1) jpart[r] = make_float4(jx, jy, jz, 0.f);
        named_bar_sync(1, 256);
        if (j == 0) {
          const float4 o = jpart[r];
          jx += o.x; jy += o.y; jz += o.z;
          const float d = a.clamp_dist;
          const float sigma = (th >= -d && th <= d) ? 1.f - th * th : 0.f;
          float x = xyzs[r * 3], y = xyzs[r * 3 + 1], z = xyzs[r * 3 + 2];
          float eps = 0.f, gn = 0.f;
          if (gr != 0.f) {
            const float ratio = sigma / gr;               // = N sign(pred - gt)
            const float gx = jx * ratio, gy = jy * ratio, gz = jz * ratio;
            gn = sqrtf(fmaf(gx, gx, fmaf(gy, gy, gz * gz)));
            if (gn > 0.f) {
              const float coef = a.eik_c2 * (1.f - 1.f / gn);
              eps = -2.f * th * coef * gn * gn / gr;
              const float k = ratio * coef;
              x = fmaf(eps, x, x) + k * gx;
              y = fmaf(eps, y, y) + k * gy;
              z = fmaf(eps, z, z) + k * gz;
            }
          }
          float e_loss = valid ? (1.f - gn) * (1.f - gn) : 0.f;
          float e_g = eps * gr;
          if (valid) a.eik_row[row] = make_float4(x, y, z, eps);
          rhos[r] = 1.f + eps;
          xyzs[r * 3] = x; xyzs[r * 3 + 1] = y; xyzs[r * 3 + 2] = z;
#pragma unroll
          for (int o2 = 16; o2 > 0; o2 >>= 1) {
            e_loss += __shfl_xor_sync(0xffffffffu, e_loss, o2);
            e_g += __shfl_xor_sync(0xffffffffu, e_g, o2);
          }
          if (lane == 0) { ered[q * 2] = e_loss; ered[q * 2 + 1] = e_g; }
        }
        named_bar_sync(1, 256);
        if (t == 0 && tile_valid) {
          a.eik_tile[tile * 2] = (ered[0] + ered[2]) + (ered[4] + ered[6]);
          a.eik_tile[tile * 2 + 1] = (ered[1] + ered[3]) + (ered[5] + ered[7]);
        }
        tell_reducers();
      }
    }

  } else {
    // ===================== reducer warps =====================
    setmaxnreg_dec<96>();
    const int rt = threadIdx.x - 384;      // 0..127, columns 4rt..4rt+3
    const uint32_t sub_o = (rt >> 4) * kSubBytes;
    const uint32_t chk = (rt & 15) >> 1;
    const uint32_t hb = (rt & 1) * 8;
    float db[6][4];
    float dw8[4];
    float dwx[2][4][3];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k) db[i][k] = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) dw8[k] = 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 3; ++e) dwx[i][k][e] = 0.f;
    uint32_t par_dy = 0;
    for (long long it = 0; it < n_iter; ++it) {
      long long tile = it * gridDim.x + blockIdx.x;
      const bool tile_valid = tile < n_tiles;
      tile = tile_valid ? tile : n_tiles - 1;
#pragma unroll
      for (int l = 7; l >= 0; --l) {
        DSDF_PROG(0, ((unsigned)it << 16) | ((unsigned)l << 8) | 1u);
        mbar_wait(dy_ready, par_dy);
        par_dy ^= 1;
        const bool active = (l != 3) || (rt < 64);   // dY3 has 256 columns
        float s[4] = {0.f, 0.f, 0.f, 0.f};
        if (active) {
          if (false) {
          } else if (l == 4 || l == 0) {
            const int xi = (l == 0) ? 0 : 1;
#pragma unroll 4
            for (int r = 0; r < kTileM; ++r) {
              const uint32_t off = sub_o + sub_chunk_off(r, chk) + hb;
              uint32_t w0, w1;
              asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(s_base + C::kSmemA + off));
              const float d0 = bf16_lo(w0), d1 = bf16_hi(w0), d2 = bf16_lo(w1), d3 = bf16_hi(w1);
              if (EIK && l == 0) {            // bias / latent column sums see the row's total upstream (1 + eps) g
                const float rho = rhos[r];
                s[0] = fmaf(rho, d0, s[0]); s[1] = fmaf(rho, d1, s[1]); s[2] = fmaf(rho, d2, s[2]); s[3] = fmaf(rho, d3, s[3]);
              } else {
                s[0] += d0; s[1] += d1; s[2] += d2; s[3] += d3;
              }
              const float x = xyzs[r * 3], y = xyzs[r * 3 + 1], z = xyzs[r * 3 + 2];
              dwx[xi][0][0] = fmaf(d0, x, dwx[xi][0][0]); dwx[xi][0][1] = fmaf(d0, y, dwx[xi][0][1]); dwx[xi][0][2] = fmaf(d0, z, dwx[xi][0][2]);
              dwx[xi][1][0] = fmaf(d1, x, dwx[xi][1][0]); dwx[xi][1][1] = fmaf(d1, y, dwx[xi][1][1]); dwx[xi][1][2] = fmaf(d1, z, dwx[xi][1][2]);
              dwx[xi][2][0] = fmaf(d2, x, dwx[xi][2][0]); dwx[xi][2][1] = fmaf(d2, y, dwx[xi][2][1]); dwx[xi][2][2] = fmaf(d2, z, dwx[xi][2][2]);
              dwx[xi][3][0] = fmaf(d3, x, dwx[xi][3][0]); dwx[xi][3][1] = fmaf(d3, y, dwx[xi][3][1]); dwx[xi][3][2] = fmaf(d3, z, dwx[xi][3][2]);
            }
            if (tile_valid)
              *reinterpret_cast<float4*>(a.tile_colsum + ((size_t)tile * 2 + xi) * kHidden + 4 * rt) =
                  make_float4(s[0], s[1], s[2], s[3]);
          } else {
#pragma unroll 8
            for (int r = 0; r < kTileM; ++r) {
              const uint32_t off = sub_o + sub_chunk_off(r, chk) + hb;
              uint32_t w0, w1;
              asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(s_base + C::kSmemA + off));
              s[0] += bf16_lo(w0); s[1] += bf16_hi(w0); s[2] += bf16_lo(w1); s[3] += bf16_hi(w1);
            }
          }
          if (l != 0 && l != 4) {
            const int sl = db_slot(l);
#pragma unroll
            for (int k = 0; k < 4; ++k) db[sl][k] += s[k];
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(rt < 64 ? red_done : red_done_hi);
      }
    }
    float* out = a.cta_partials + (size_t)blockIdx.x * kCtaPartialFloats;
#pragma unroll
    for (int i = 0; i < 6; ++i)
      *reinterpret_cast<float4*>(out + kPartDb + i * kHidden + 4 * rt) =
          make_float4(db[i][0], db[i][1], db[i][2], db[i][3]);
    // (the kPartW8 slot is written by the forward kernel: dW8 is accumulated there)
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 3; ++e)
          out[kPartXyz + (i * kHidden + 4 * rt + k) * 3 + e] = dwx[i][k][e];
  }

  tc_fence_before();
  __syncthreads();
  if (CG > 1) cluster_sync_all();
  if (warp == 2) {
    if (CG > 1) tmem_dealloc2<512>(tmem); else tmem_dealloc<512>(tmem);
  }
}

}  // namespace dsdf

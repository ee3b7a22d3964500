// Weight-gradient GEMMs dW_l = dY_l^T * X_l over the stashed activation images (split-K over
// row tiles).  Both operands are consumed MN-major straight from the SWIZZLE_128B images the
// forward / backward kernels stashed (rows = K), so no transpose pass exists anywhere.
//
//   one CTA = one work item (layer l, 128-wide out chunk oc, K split): D[128 o x N i] in TMEM
//   warp 0 producer (4 KB bulk copies: 32-row slices of 2 dY SUBs + N/64 X SUBs per stage)
//   warp 1 MMA issuer (M=128, N=256, K=16; A and B MN-major)
//   warp 2 TMEM alloc; warps 4..7 epilogue (TMEM -> fp32 split-K partial in HBM)
//
// Replaces the 9 wgrad `mm` calls of the reference's autograd backward
// (train_deep_sdf.py:731) for lin1..lin7; lin0/lin8 and all bias grads come from the
// backward kernel's reducer warps and the segmented kernels.
#pragma once
#include "layout.cuh"
#include "ptx.cuh"

namespace dsdf {

struct WgItem {
  int layer;       // 1..7
  int oc;          // out chunk (128 rows of dW)
  int n_sub;       // N / 64 (8, or 4 for lin4's x3 part)
  int split;       // K-split index
  long long tile_lo, tile_hi;
};

struct WgArgs {
  const WgItem* items;
  const uint8_t* xstash;
  const uint8_t* dystash;
  float* partials;  // [n_items][128][512]
};

constexpr int kWgStages = 5;
constexpr int kWgThreads = 256;
constexpr uint32_t kWgSlice = 4096;                    // 32 rows of one SUB
constexpr uint32_t kWgStageBytes = 10 * kWgSlice;      // 2 A slices + 8 B slices
constexpr uint32_t kWgSmemAux = kWgStages * kWgStageBytes;  // 204800
constexpr uint32_t kWgSmemBytes = kWgSmemAux + 256;

__global__ void __launch_bounds__(kWgThreads, 1) k_wgrad(const WgArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWgSmemAux);
  uint64_t* full = bars;
  uint64_t* empty = bars + kWgStages;
  uint64_t* done = bars + 2 * kWgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kWgSmemAux + 128);

  if ((s_base & 1023u) != 0) __trap();
  if (threadIdx.x == 0) {
    for (int i = 0; i < kWgStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const WgItem it = a.items[blockIdx.x];
  const long long n_tiles = it.tile_hi - it.tile_lo;
  const uint32_t stage_tx = (2 + it.n_sub) * kWgSlice;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (long long tile = it.tile_lo; tile < it.tile_hi; ++tile) {
        const uint8_t* dy = a.dystash + (size_t)tile * kStashTileBytes +
                            (size_t)it.layer * kTileBytes + (size_t)(2 * it.oc) * kSubBytes;
        const uint8_t* x = a.xstash + (size_t)tile * kStashTileBytes +
                           (size_t)(it.layer - 1) * kTileBytes;
        for (int s = 0; s < 4; ++s) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], stage_tx);
          uint8_t* dst = smem + stage * kWgStageBytes;
          bulk_g2s(dst, dy + s * kWgSlice, kWgSlice, &full[stage]);
          bulk_g2s(dst + kWgSlice, dy + kSubBytes + s * kWgSlice, kWgSlice, &full[stage]);
          for (int b = 0; b < it.n_sub; ++b)
            bulk_g2s(dst + (2 + b) * kWgSlice, x + (size_t)b * kSubBytes + s * kWgSlice, kWgSlice,
                     &full[stage]);
          if (++stage == kWgStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 256, true, true);
      uint32_t stage = 0, phase = 0;
      const int n_half = it.n_sub / 4;
      for (long long i = 0; i < n_tiles * 4; ++i) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t sa = s_base + stage * kWgStageBytes;
        const uint32_t sb = sa + 2 * kWgSlice;
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
          for (int hf = 0; hf < n_half; ++hf) {
            umma_bf16(tmem + hf * 256, umma_smem_desc(sa + ks * 2048, kWgSlice, 1024),
                      umma_smem_desc(sb + hf * 4 * kWgSlice + ks * 2048, kWgSlice, 1024), idesc,
                      (i > 0 || ks > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty[stage]);
        if (++stage == kWgStages) { stage = 0; phase ^= 1; }
      }
      umma_commit(done);
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const uint32_t r = q * 32 + lane;
    float* out = a.partials + ((size_t)blockIdx.x * kTileM + r) * kHidden;
    const int ncols = it.n_sub * 64;
    mbar_wait(done, 0);
    tc_fence_after();
    for (int c = 0; c < ncols; c += 32) {
      uint32_t v[32];
      if (n_tiles > 0) {
        tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + c, v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0u;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint4*>(out + c + 4 * i) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace dsdf

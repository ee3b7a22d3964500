// Single-tile GEMM self-tests for the two UMMA operand conventions used by the hot path,
// built from the same descriptor helpers (ptx.cuh) and the same SUB layout (layout.cuh):
//   kind 0: K-major  A[128x64] * B[128x64]^T           -> D[128x128]   (forward / dgrad)
//   kind 1: MN-major A[k<32][m<128] , B[k<32][n<256]   -> D[128x256]   (wgrad: D = A^T B)
// Exposed through dsdf_selftest_gemm() so the GPU tests can localise a descriptor bug.
#pragma once
#include "layout.cuh"
#include "ptx.cuh"

namespace dsdf {

// kind 0: a_img, b_img = one SUB each.  kind 1: a_img = 2 SUBs (M atoms), b_img = 4 SUBs (N atoms);
// only rows 0..31 (the first 4 KB) of every SUB are used as the K extent.
__global__ void __launch_bounds__(128, 1)
k_selftest_gemm(int kind, const uint8_t* a_img, const uint8_t* b_img, float* d_out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t& bar_full = *reinterpret_cast<uint64_t*>(smem + 65536 - 64);
  uint64_t& bar_done = *reinterpret_cast<uint64_t*>(smem + 65536 - 56);
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(smem + 65536 - 48);
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if ((s_base & 1023u) != 0) __trap();
  if (threadIdx.x == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<256>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t sA = 0, sB = 32768;
  if (threadIdx.x == 0) {
    if (kind == 0) {
      mbar_arrive_expect_tx(&bar_full, 2 * kSubBytes);
      bulk_g2s(smem + sA, a_img, kSubBytes, &bar_full);
      bulk_g2s(smem + sB, b_img, kSubBytes, &bar_full);
    } else {
      mbar_arrive_expect_tx(&bar_full, 6 * 4096);
      for (int i = 0; i < 2; ++i) bulk_g2s(smem + sA + i * 4096, a_img + (size_t)i * kSubBytes, 4096, &bar_full);
      for (int i = 0; i < 4; ++i) bulk_g2s(smem + sB + i * 4096, b_img + (size_t)i * kSubBytes, 4096, &bar_full);
    }
    mbar_wait(&bar_full, 0);
    tc_fence_after();
    if (kind == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 128, false, false);
      for (int ks = 0; ks < 4; ++ks)
        umma_bf16(tmem, umma_smem_desc(s_base + sA + ks * 32, 16, 1024),
                  umma_smem_desc(s_base + sB + ks * 32, 16, 1024), idesc, ks > 0);
    } else {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 256, true, true);
      for (int ks = 0; ks < 2; ++ks)
        umma_bf16(tmem, umma_smem_desc(s_base + sA + ks * 2048, 4096, 1024),
                  umma_smem_desc(s_base + sB + ks * 2048, 4096, 1024), idesc, ks > 0);
    }
    umma_commit(&bar_done);
  }
  __syncwarp();
  mbar_wait(&bar_done, 0);
  tc_fence_after();
  const int ncols = kind == 0 ? 128 : 256;
  const uint32_t r = warp * 32 + lane;
  for (int c = 0; c < ncols; c += 32) {
    uint32_t v[32];
    tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) d_out[(size_t)r * ncols + c + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(tmem);
}

}  // namespace dsdf

namespace dsdf {
// MMA issue-rate micro-benchmark (diagnostics): `iters` x 8 back-to-back tcgen05.mma M=128 N=256 K=16
// on fixed (zeroed) smem operands with precomputed descriptors; out[blockIdx.x] = total cycles.
//   kind & 7  : B layout 0 = no-swizzle K-step, 1 = SWIZZLE_128B, 4 = SWIZZLE_64B, 6 = SWIZZLE_32B
//   kind & 8  : tcgen05.commit to a scratch barrier after every MMA
//   kind & 16 : try_wait on an already-completed barrier before every MMA
//   kind & 32 : commit after every 2nd MMA only
template <int LT, bool CM, bool TW, bool CM2>
__device__ __forceinline__ void mma_rate_loop(uint32_t s_base, uint32_t tmem, int iters, uint64_t* bar_scr,
                                              uint64_t* bar_rdy) {
  constexpr uint32_t idesc = umma_idesc_bf16(128, 256, false, false);
  const uint64_t a0 = umma_smem_desc(s_base, 16, 1024);
  uint64_t b0;
  if (LT == 1) b0 = umma_smem_desc(s_base + 131072, 16, 1024);
  else if (LT == 4) b0 = umma_smem_desc_ns(s_base + 131072, 16, 512) | (4ull << 61);
  else if (LT == 6) b0 = umma_smem_desc_ns(s_base + 131072, 16, 256) | (6ull << 61);
  else b0 = umma_smem_desc_ns(s_base + 131072, 128, 256);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (TW) mbar_wait(bar_rdy, 0);
      const uint64_t ad = a0 + (uint64_t)(((u >> 2) * kSubBytes + (u & 3) * 32) >> 4);
      uint64_t bd;
      if (LT == 1) bd = b0 + (uint64_t)(((u & 3) * 32) >> 4);
      else if (LT == 4) bd = b0 + (uint64_t)((((u >> 1) & 3) * 16384 + (u & 1) * 32) >> 4);
      else bd = b0 + (uint64_t)((u * kStepBytes) >> 4);
      umma_bf16(tmem + (i & 1) * 256, ad, bd, idesc, 1u);
      if (CM || (CM2 && (u & 1))) umma_commit(bar_scr);
    }
  }
}
__global__ void __launch_bounds__(128, 1) k_mma_rate(int kind, int iters, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr uint32_t kAux = 196608;
  uint64_t* bar_done = reinterpret_cast<uint64_t*>(smem + kAux);
  uint64_t* bar_scr = bar_done + 1;
  uint64_t* bar_rdy = bar_done + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kAux + 32);
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  for (uint32_t i = threadIdx.x; i < kAux / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(bar_done, 1);
    mbar_init(bar_scr, 1);
    mbar_init(bar_rdy, 1);
    fence_mbar_init();
    mbar_arrive(bar_rdy);   // phase 0 complete
  }
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (warp == 1) {
    const long long t0 = clock64();
    if (elect_one()) {
#define RATE_CASE(LT)                                                                           \
      if ((kind & 7) == LT) {                                                                   \
        if (kind & 8) mma_rate_loop<LT, true, false, false>(s_base, tmem, iters, bar_scr, bar_rdy);        \
        else if (kind & 16) mma_rate_loop<LT, true, true, false>(s_base, tmem, iters, bar_scr, bar_rdy);   \
        else if (kind & 32) mma_rate_loop<LT, false, false, true>(s_base, tmem, iters, bar_scr, bar_rdy);  \
        else mma_rate_loop<LT, false, false, false>(s_base, tmem, iters, bar_scr, bar_rdy);      \
      }
      RATE_CASE(0) RATE_CASE(1) RATE_CASE(4) RATE_CASE(6)
#undef RATE_CASE
      umma_commit(bar_done);
    }
    __syncwarp();
    mbar_wait(bar_done, 0);
    if (threadIdx.x == 32) out[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}
}  // namespace dsdf

namespace dsdf {
// cta_group::2 variant of the issue-rate micro-benchmark (cluster of 2; the leader issues
// M=256 N=256 K=16 MMAs).  kind & 7: B layout (0 no-swizzle, 1 SWIZZLE_128B), kind & 8: multicast
// commit after every MMA, kind & 16: N = 128.
__global__ void __launch_bounds__(128, 1) k_mma_rate2(int kind, int iters, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr uint32_t kAux = 196608;
  uint64_t* bar_done = reinterpret_cast<uint64_t*>(smem + kAux);
  uint64_t* bar_scr = bar_done + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kAux + 32);
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const bool leader = cluster_ctarank() == 0;
  for (uint32_t i = threadIdx.x; i < kAux / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(bar_done, 1);
    mbar_init(bar_scr, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc2<512>(tmem_slot);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (warp == 1) {
    const long long t0 = clock64();
    if (leader && elect_one()) {
      const bool sw = kind & 1, cm = kind & 8, n128 = kind & 16;
      const uint32_t idesc = n128 ? umma_idesc_bf16(256, 128, false, false) : umma_idesc_bf16(256, 256, false, false);
      const uint64_t a0 = umma_smem_desc(s_base, 16, 1024);
      const uint64_t b0 = sw ? umma_smem_desc(s_base + 131072, 16, 1024) : umma_smem_desc_ns(s_base + 131072, 128, 256);
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const uint64_t ad = a0 + (uint64_t)(((u >> 2) * kSubBytes + (u & 3) * 32) >> 4);
          const uint64_t bd = sw ? b0 + (uint64_t)(((u & 3) * 32) >> 4) : b0 + (uint64_t)((u * 4096) >> 4);
          umma2_bf16(tmem + (i & 1) * 256, ad, bd, idesc, 1u);
          if (cm) umma2_commit_mc(bar_scr, (uint16_t)3);
        }
      }
      umma2_commit_mc(bar_done, (uint16_t)3);
    }
    __syncwarp();
    mbar_wait(bar_done, 0);
    if ((threadIdx.x & 31) == 0) out[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc2<512>(tmem);
}
}  // namespace dsdf

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

// Dense-input backward pieces of the general autograd path (Decoder.forward(input[N,259]) with
// arbitrary per-row latents, differentiable w.r.t. parameters AND input: reference
// networks/deep_sdf_decoder.py:74-109 under torch autograd, used by reconstruct.py:90-94,130 and any
// caller that feeds its own latents).  The tensor-core chain (forward, dgrad lin7..lin1, wgrad
// lin1..lin8) is shared with the fused step; what only exists for dense input is
//   dInput[r, j]   = sum_o dY0[r,o] W0[o,j] + sum_o dY4[r,o] W4[o,253+j]       (j < 259)
//   dW0[o, j]      = sum_r dY0[r,o] latent[r,j],  dW4[o, 253+j] = sum_r dY4[r,o] latent[r,j]  (j < 256)
// (in scene mode both collapse to per-scene column sums).  These two products are 14 % of the
// backward FLOPs and run here as plain fp32 CUDA-core tile GEMMs straight from the stashed bf16 dY
// images: this path is the compatibility path, not the measured hot path.
#pragma once
#include "fold_kernels.cuh"
#include "layout.cuh"

namespace dsdf {

constexpr int kDsLd = 128;   // leading dimension of the [k][m] smem operand (32 KB + 16 KB = the 48 KB static limit)

// C[8 x 4] per thread += As[k][tm*8 .. +7] * Bs[k][tn*4 .. +3] over 64 k
__device__ __forceinline__ void dense_mac64(const float (*As)[kDsLd], const float (*Bs)[64], int tm, int tn,
                                            float (&acc)[8][4]) {
#pragma unroll 8
  for (int k = 0; k < 64; ++k) {
    const float4 a0 = *reinterpret_cast<const float4*>(&As[k][tm * 8]);
    const float4 a1 = *reinterpret_cast<const float4*>(&As[k][tm * 8 + 4]);
    const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tn * 4]);
    const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) acc[i][jj] = fmaf(av[i], bv[jj], acc[i][jj]);
  }
}

__device__ __forceinline__ void unpack8(const uint4 v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}

// ---- dInput: grid (tiles, 5), 256 threads; block = 128 rows x columns [64 y, 64 y + 64) ---------
__global__ void __launch_bounds__(256) k_dense_dinput(const uint8_t* __restrict__ dystash,
                                                      const float* __restrict__ weff, long long n_rows,
                                                      float* __restrict__ dinput) {
  __shared__ __align__(16) float As[64][kDsLd];
  __shared__ __align__(16) float Bs[64][64];
  const long long tile = blockIdx.x;
  const int n0 = blockIdx.y * 64;
  const int t = threadIdx.x, tm = t >> 4, tn = t & 15;
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) acc[i][jj] = 0.f;
  for (int which = 0; which < 2; ++which) {
    const int l = which == 0 ? 0 : kSkipLayer;
    const int ld = ref_in(l), coff = which == 0 ? 0 : kX3;
    const float* w = weff + weff_off(l);
    const uint8_t* dy = dystash + (size_t)tile * kStashTileBytes + (size_t)l * kTileBytes;
    for (int sub = 0; sub < 8; ++sub) {       // o = 64 sub .. 64 sub + 63
      __syncthreads();
#pragma unroll
      for (int q = 0; q < 4; ++q) {           // 128 rows x 8 chunks
        const int id = t + 256 * q, r = id >> 3, ch = id & 7;
        float f[8];
        unpack8(*reinterpret_cast<const uint4*>(dy + (size_t)sub * kSubBytes + sub_chunk_off(r, ch)), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) As[ch * 8 + e][r] = f[e];
      }
#pragma unroll
      for (int q = 0; q < 16; ++q) {          // 64 k x 64 n
        const int id = t + 256 * q, k = id >> 6, n = id & 63;
        const int j = n0 + n;
        const bool has = j < kIn0 && (which == 0 || kXyzAtSkip || j < kLatent);   // SIREN: no xyz into lin4
        Bs[k][n] = has ? __ldg(w + (size_t)(sub * 64 + k) * ld + coff + j) : 0.f;
      }
      __syncthreads();
      dense_mac64(As, Bs, tm, tn, acc);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long row = tile * kTileM + tm * 8 + i;
    if (row >= n_rows) continue;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = n0 + tn * 4 + jj;
      if (j < kIn0) dinput[row * kIn0 + j] = acc[i][jj];
    }
  }
}

// ---- latent columns of dW0 / dW4: grid (16, 2, n_split), 256 threads ------------------------------
// block = out rows [128 (x/4), +128) x latent columns [64 (x%4), +64) of layer `which`, rows of
// tiles [tile_lo, tile_hi) of its split; fp32 partial per split, summed in order by the reducer.
__global__ void __launch_bounds__(256) k_dense_dw_latent(const uint8_t* __restrict__ dystash,
                                                         const float* __restrict__ input, long long n_rows,
                                                         long long n_tiles, float* __restrict__ partials) {
  __shared__ __align__(16) float As[64][kDsLd];
  __shared__ __align__(16) float Bs[64][64];
  const int o0 = (blockIdx.x >> 2) * 128, j0 = (blockIdx.x & 3) * 64;
  const int which = blockIdx.y, l = which == 0 ? 0 : kSkipLayer;
  const long long tile_lo = n_tiles * blockIdx.z / gridDim.z, tile_hi = n_tiles * (blockIdx.z + 1) / gridDim.z;
  const int t = threadIdx.x, tm = t >> 4, tn = t & 15;
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) acc[i][jj] = 0.f;
  for (long long tile = tile_lo; tile < tile_hi; ++tile) {
    const uint8_t* dy = dystash + (size_t)tile * kStashTileBytes + (size_t)l * kTileBytes +
                        (size_t)(o0 >> 6) * kSubBytes;
    for (int half = 0; half < 2; ++half) {    // 64 rows (= K) at a time
      __syncthreads();
#pragma unroll
      for (int q = 0; q < 4; ++q) {           // 64 rows x 2 SUBs x 8 chunks
        const int id = t + 256 * q, k = id >> 4, s = (id >> 3) & 1, ch = id & 7;
        float f[8];
        unpack8(*reinterpret_cast<const uint4*>(dy + (size_t)s * kSubBytes + sub_chunk_off(half * 64 + k, ch)), f);
        *reinterpret_cast<float4*>(&As[k][s * 64 + ch * 8]) = make_float4(f[0], f[1], f[2], f[3]);
        *reinterpret_cast<float4*>(&As[k][s * 64 + ch * 8 + 4]) = make_float4(f[4], f[5], f[6], f[7]);
      }
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        const int id = t + 256 * q, k = id >> 6, n = id & 63;
        const long long row = tile * kTileM + half * 64 + k;
        Bs[k][n] = row < n_rows ? __ldg(input + row * kIn0 + j0 + n) : 0.f;
      }
      __syncthreads();
      dense_mac64(As, Bs, tm, tn, acc);
    }
  }
  float* p = partials + ((size_t)(blockIdx.z * 2 + which) * kHidden) * kLatent;
#pragma unroll
  for (int i = 0; i < 8; ++i)
    *reinterpret_cast<float4*>(p + (size_t)(o0 + tm * 8 + i) * kLatent + j0 + tn * 4) =
        make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
}
// grid (512, 2), 256 threads
__global__ void k_dense_dw_reduce(const float* __restrict__ partials, int n_split, float* __restrict__ dweff) {
  const int o = blockIdx.x, which = blockIdx.y, j = threadIdx.x;
  float s = 0.f;
  for (int k = 0; k < n_split; ++k) s += partials[((size_t)(k * 2 + which) * kHidden + o) * kLatent + j];
  if (which == 0) dweff[weff_off(0) + (size_t)o * kIn0 + j] = s;
  else dweff[weff_off(4) + (size_t)o * kHidden + kX3 + j] = s;
}

// ---- g = dOut (1 - out^2) -> fp32 g, the wgrad kernel's lin8 "dY" image, per-tile sum of g ---------
// grid tiles, 128 threads
__global__ void k_dense_g(const float* __restrict__ dout, const float* __restrict__ out, long long n_rows,
                          float* __restrict__ g, uint8_t* __restrict__ g_img, float* __restrict__ tile_partials) {
  __shared__ float sh[4];
  const long long tile = blockIdx.x;
  const uint32_t r = threadIdx.x;
  const long long row = tile * kTileM + r;
  float gv = 0.f;
  if (row < n_rows) {
    const float o = out[row];
#ifdef DSDF_ARCH_SIREN
    gv = dout[row];                       // no output nonlinearity
    (void)o;
#else
    gv = dout[row] * (1.f - o * o);       // through self.th
#endif
    g[row] = gv;
  }
  const __nv_bfloat16 gh = __float2bfloat16_rn(gv);
  const __nv_bfloat16 gl = __float2bfloat16_rn(gv - __bfloat162float(gh));
  uint8_t* gi = g_img + (size_t)tile * kSubBytes;
  *reinterpret_cast<uint4*>(gi + sub_chunk_off(r, 0)) = make_uint4(
      (uint32_t)__bfloat16_as_ushort(gh) | ((uint32_t)__bfloat16_as_ushort(gl) << 16), 0u, 0u, 0u);
#pragma unroll
  for (int ch = 1; ch < 8; ++ch) *reinterpret_cast<uint4*>(gi + sub_chunk_off(r, ch)) = make_uint4(0u, 0u, 0u, 0u);
  float s = gv;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((r & 31) == 0) sh[r >> 5] = s;
  __syncthreads();
  if (r == 0) {
    tile_partials[tile * 2] = 0.f;
    tile_partials[tile * 2 + 1] = (sh[0] + sh[1]) + (sh[2] + sh[3]);
  }
}

// ---- column sums of dY0 / dY4 over chunks of tiles: grid (n_chunks, 2), 512 threads ---------------
__global__ void k_colsum_chunks(const float* __restrict__ tile_colsum, long long n_tiles,
                                float* __restrict__ chunk_sum) {
  const int c = blockIdx.x, which = blockIdx.y, col = threadIdx.x;
  const long long lo = n_tiles * c / gridDim.x, hi = n_tiles * (c + 1) / gridDim.x;
  float s = 0.f;
  for (long long t = lo; t < hi; ++t) s += tile_colsum[((size_t)t * 2 + which) * kHidden + col];
  chunk_sum[((size_t)c * 2 + which) * kHidden + col] = s;
}

}  // namespace dsdf

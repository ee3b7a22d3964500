// Weight preparation, once per optimiser step (north_star (3): "the weight-norm scale g/||v||
// is folded into the weights once per step"):
//   k_rowscale   : scale[l][o] = g/||v||_row for weight-normalised layers, 1 otherwise
//                  (nn.utils.weight_norm, reference networks/deep_sdf_decoder.py:49-53)
//   k_weff       : fp32 effective weights in reference layout (used by the small fp32 kernels)
//   k_fold_fwd   : bf16 forward image  (B tiles [128 n x 64 k] + E tiles, layout.cuh)
//   k_fold_dgrad : bf16 dgrad image    (B tiles [128 i x 64 o] of W^T)
#pragma once
#include "layout.cuh"
#include "ptx.cuh"

namespace dsdf {

// Parameter addressing: one flat fp32 buffer + per-layer element offsets, reference shapes:
//   weight (or weight_v) [out,in], weight_g [out,1] (or -1), bias [out].
struct ParamLayout {
  long long w[kNumLin];
  long long g[kNumLin];
  long long b[kNumLin];
};
// offsets of layer l inside the fp32 effective-weight buffer (dense, reference [out,in] layout)
__host__ __device__ constexpr long long weff_off(int l) {
  long long o = 0;
  for (int i = 0; i < l; ++i) o += (long long)ref_out(i) * ref_in(i);
  return o;
}
constexpr long long kWeffElems = weff_off(kNumLin);  // 1 836 032 weights
__host__ __device__ constexpr int scale_off(int l) {
  int o = 0;
  for (int i = 0; i < l; ++i) o += ref_out(i);
  return o;
}
constexpr int kScaleElems = scale_off(kNumLin);  // 3838 rows

// one warp per output row
__global__ void k_rowscale(const float* __restrict__ params, ParamLayout pl,
                           float* __restrict__ scale, float* __restrict__ inv_norm) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw >= kScaleElems) return;
  int l = 0;
  while (l + 1 < kNumLin && gw >= scale_off(l + 1)) ++l;
  const int o = gw - scale_off(l);
  if (pl.g[l] < 0) {
    if (lane == 0) { scale[gw] = 1.f; inv_norm[gw] = 1.f; }
    return;
  }
  const int in = ref_in(l);
  const float* v = params + pl.w[l] + (long long)o * in;
  float ss = 0.f;
  for (int i = lane; i < in; i += 32) ss = fmaf(v[i], v[i], ss);
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, s);
  if (lane == 0) {
    const float nrm = sqrtf(ss);
    inv_norm[gw] = 1.f / nrm;
    scale[gw] = params[pl.g[l] + o] / nrm;
  }
}

__global__ void k_weff(const float* __restrict__ params, ParamLayout pl,
                       const float* __restrict__ scale, float* __restrict__ weff) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) weff[kWeffElems] = params[pl.b[8]];   // lin8 bias rides behind the weights
  if (i >= kWeffElems) return;
  int l = 0;
  while (l + 1 < kNumLin && i >= weff_off(l + 1)) ++l;
  const long long e = i - weff_off(l);
  const int o = (int)(e / ref_in(l));
  weff[i] = params[pl.w[l] + e] * scale[scale_off(l) + o];
}

__device__ __forceinline__ float weff_at(const float* weff, int l, int o, int i) {
  return weff[weff_off(l) + (long long)o * ref_in(l) + i];
}

// one thread per 16-byte chunk of the forward image
__global__ void k_fold_fwd(const float* __restrict__ weff, const float* __restrict__ params,
                           ParamLayout pl, uint8_t* __restrict__ img) {
  const long long cidx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)kFwdTilesTotal * (kSubBytes / 16);
  if (cidx >= total) return;
  const int tile = (int)(cidx / (kSubBytes / 16));
  const int within = (int)(cidx % (kSubBytes / 16));
  const int r = within >> 3;   // n local
  const int ch = within & 7;   // logical k chunk
  int l = 0;
  while (l + 1 < 8 && tile >= fwd_tile_base(l + 1)) ++l;
  const int tl = tile - fwd_tile_base(l);
  const int KB = fwd_kb(l), NC = fwd_nc(l);
  float vals[8];
  if (tl < KB * NC) {
    const int kb = tl / NC, nc = tl % NC;
    const int o = nc * 128 + r;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int kk = kb * 64 + ch * 8 + e;
      const int ic = pad_to_ref_col(l, kk);
      vals[e] = (o < ref_out(l) && ic >= 0) ? weff_at(weff, l, o, ic) : 0.f;
    }
  } else {
    // E tile: columns c = 16*j + e for n-chunk j
#pragma unroll
    for (int e8 = 0; e8 < 8; ++e8) {
      const int c = ch * 8 + e8;
      const int j = c >> 4, e = c & 15;
      const int o = j * 128 + r;
      float v = 0.f;
      if (j < NC && o < ref_out(l)) {
        if (e < 9) {
          const int xc = xyz_col(l, e % 3);
          if (xc >= 0) {
            const float w = weff_at(weff, l, o, xc);
            const float wh = __bfloat162float(__float2bfloat16_rn(w));
            v = (e < 6) ? wh : (w - wh);
          }
        } else if (e < 11) {
          const float b = params[pl.b[l] + o];
          const float bh = __bfloat162float(__float2bfloat16_rn(b));
          v = (e == 9) ? bh : (b - bh);
        }
      }
      vals[e8] = v;
    }
  }
  uint4 w;
  w.x = pack_bf16x2(vals[0], vals[1]);
  w.y = pack_bf16x2(vals[2], vals[3]);
  w.z = pack_bf16x2(vals[4], vals[5]);
  w.w = pack_bf16x2(vals[6], vals[7]);
  *reinterpret_cast<uint4*>(img + (size_t)tile * kSubBytes + sub_chunk_off(r, ch)) = w;
}

// one thread per 16-byte chunk of the dgrad image: tile (l, kb = o block, nc = i chunk),
// element (r = i local, c = o local) = W_l[o, pad_to_ref_col(l, i)]
__global__ void k_fold_dgrad(const float* __restrict__ weff, uint8_t* __restrict__ img) {
  const long long cidx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)kDgTilesTotal * (kSubBytes / 16);
  if (cidx >= total) return;
  const int tile = (int)(cidx / (kSubBytes / 16));
  const int within = (int)(cidx % (kSubBytes / 16));
  const int r = within >> 3;
  const int ch = within & 7;
  int l = 1;
  while (l + 1 < 8 && tile >= dg_tile_base(l + 1)) ++l;
  const int tl = tile - dg_tile_base(l);
  const int NC = dg_nc(l);
  const int kb = tl / NC, nc = tl % NC;
  const int ic = pad_to_ref_col(l, nc * 128 + r);
  float vals[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int o = kb * 64 + ch * 8 + e;
    vals[e] = (ic >= 0 && o < ref_out(l)) ? weff_at(weff, l, o, ic) : 0.f;
  }
  uint4 w;
  w.x = pack_bf16x2(vals[0], vals[1]);
  w.y = pack_bf16x2(vals[2], vals[3]);
  w.z = pack_bf16x2(vals[4], vals[5]);
  w.w = pack_bf16x2(vals[6], vals[7]);
  *reinterpret_cast<uint4*>(img + (size_t)tile * kSubBytes + sub_chunk_off(r, ch)) = w;
}

}  // namespace dsdf

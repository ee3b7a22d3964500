// Weight preparation, once per optimiser step (north_star (3): "the weight-norm scale g/||v||
// is folded into the weights once per step"):
//   k_rowscale_weff : scale[l][o] = g/||v||_row for weight-normalised layers, 1 otherwise
//                  (nn.utils.weight_norm, reference networks/deep_sdf_decoder.py:49-53) and the fp32
//                  effective weights in reference layout (read by the image builders and the small fp32 kernels)
//   k_fold_images: bf16 forward K-step stream (layout.cuh w2_*, E steps included), bf16 dgrad K-step stream of W^T
//                  (layout.cuh dg2_*) and the fp32 xyz-column table, one launch
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

// Weight-norm fold, one warp per output row: scale = g / ||v|| (1 for a plain Linear), inv_norm for the
// weight-norm backward, and the effective fp32 row weff = v * scale in the reference [out, in] layout.
__global__ void k_rowscale_weff(const float* __restrict__ params, ParamLayout pl,
                                float* __restrict__ scale, float* __restrict__ inv_norm,
                                float* __restrict__ weff) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw == 0 && lane == 0) weff[kWeffElems] = params[pl.b[8]];   // lin8 bias rides behind the weights
  if (gw >= kScaleElems) return;
  int l = 0;
  while (l + 1 < kNumLin && gw >= scale_off(l + 1)) ++l;
  const int o = gw - scale_off(l);
  const int in = ref_in(l);
  const float* v = params + pl.w[l] + (long long)o * in;
  float sc = 1.f;
  if (pl.g[l] >= 0) {
    float ss = 0.f;
    for (int i = lane; i < in; i += 32) ss = fmaf(v[i], v[i], ss);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, s);
    const float nrm = sqrtf(ss);
    sc = params[pl.g[l] + o] / nrm;
    if (lane == 0) { inv_norm[gw] = 1.f / nrm; scale[gw] = sc; }
  } else if (lane == 0) {
    scale[gw] = 1.f;
    inv_norm[gw] = 1.f;
  }
  float* w = weff + weff_off(l) + (long long)o * in;
  for (int i = lane; i < in; i += 32) w[i] = v[i] * sc;
}

__device__ __forceinline__ float weff_at(const float* weff, int l, int o, int i) {
  return weff[weff_off(l) + (long long)o * ref_in(l) + i];
}

}  // namespace dsdf

namespace dsdf {

// E-step row of output feature o of layer l with bias value `bias`:
// [wxh wyh wzh | wxh wyh wzh | wxl wyl wzl | bh | bl | bh | wxh wyh wzh | 0]  (two 16 B chunks)
// Columns 11..14 only meet non-zero A columns in the Eikonal recompute pass (k_forward3<TAN>: A = [.. | eps |
// third-order xyz residual]); every other caller writes zeros there, so they change nothing.
__device__ __forceinline__ void e_step_row(const float* weff, int l, int o, bool o_valid,
                                           float bias, uint4& c0, uint4& c1) {
  float v[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) v[e] = 0.f;
  if (o_valid) {
#pragma unroll
    for (int e = 0; e < 3; ++e) {
      const int xc = xyz_col(l, e);
      if (xc >= 0) {
        const float w = weff_at(weff, l, o, xc);
        const float wh = __bfloat162float(__float2bfloat16_rn(w));
        v[e] = wh; v[3 + e] = wh; v[6 + e] = w - wh; v[12 + e] = wh;
      }
    }
    const float bh = __bfloat162float(__float2bfloat16_rn(bias));
    v[9] = bh; v[10] = bias - bh; v[11] = bh;
  }
  c0 = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  c1 = make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]), pack_bf16x2(v[12], v[13]), pack_bf16x2(v[14], v[15]));
}

// v2 forward K-step stream (layout.cuh): one thread per 16-byte chunk
__device__ __forceinline__ void fold_fwd2_chunk(long long cidx, const float* __restrict__ weff,
                                                const float* __restrict__ params, const ParamLayout& pl,
                                                uint8_t* __restrict__ img) {
  const long long total = (long long)kW2Steps * (kStepBytes / 16);
  if (cidx >= total) return;
  const int step = (int)(cidx / (kStepBytes / 16));
  const int within = (int)(cidx % (kStepBytes / 16));
  const int n = ((within >> 4) << 3) | (within & 7);   // [n/8][k/8][n%8]
  const int kc = (within >> 3) & 1;
  int l = 0;
  while (l + 1 < 8 && step >= w2_layer_base(l + 1)) ++l;
  const int sl = step - w2_layer_base(l);
  const int h = sl / w2_half_steps(l), s = sl % w2_half_steps(l);
  const int o = h * 256 + n;
  const bool ov = o < ref_out(l);
  uint4 w;
  if (s == 0) {
    uint4 c0, c1;
    e_step_row(weff, l, o, ov, ov ? params[pl.b[l] + o] : 0.f, c0, c1);
    w = kc == 0 ? c0 : c1;
  } else {
    float vals[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int ic = pad_to_ref_col(l, (s - 1) * 16 + kc * 8 + e);
      vals[e] = (ov && ic >= 0) ? weff_at(weff, l, o, ic) : 0.f;
    }
    w.x = pack_bf16x2(vals[0], vals[1]); w.y = pack_bf16x2(vals[2], vals[3]);
    w.z = pack_bf16x2(vals[4], vals[5]); w.w = pack_bf16x2(vals[6], vals[7]);
  }
  *reinterpret_cast<uint4*>(img + (size_t)step * kStepBytes + (size_t)within * 16) = w;
}

// v2 dgrad K-step stream (layout.cuh): one thread per 16-byte chunk
__device__ __forceinline__ void fold_dgrad2_chunk(long long cidx, const float* __restrict__ weff,
                                                  uint8_t* __restrict__ img) {
  const long long total = (long long)kDg2Steps * (kStepBytes / 16);
  if (cidx >= total) return;
  const int step = (int)(cidx / (kStepBytes / 16));
  const int within = (int)(cidx % (kStepBytes / 16));
  const int n = ((within >> 4) << 3) | (within & 7);   // [n/8][k/8][n%8]
  const int kc = (within >> 3) & 1;
  int l = 7;
  while (l > 0 && step >= dg2_layer_base(l - 1)) --l;
  const int sl = step - dg2_layer_base(l);
  const int h = sl / dg2_ksteps(l), s = sl % dg2_ksteps(l);
  const int ic = pad_to_ref_col(l, h * 256 + n);
  float vals[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int o = s * 16 + kc * 8 + e;
    vals[e] = (ic >= 0 && o < ref_out(l)) ? weff_at(weff, l, o, ic) : 0.f;
  }
  uint4 w;
  w.x = pack_bf16x2(vals[0], vals[1]); w.y = pack_bf16x2(vals[2], vals[3]);
  w.z = pack_bf16x2(vals[4], vals[5]); w.w = pack_bf16x2(vals[6], vals[7]);
  *reinterpret_cast<uint4*>(img + (size_t)step * kStepBytes + (size_t)within * 16) = w;
}

// fp32 xyz columns of lin0 / lin4 as float4 rows (x, y, z, 0): wx[c] = W0[c, 256..258], wx[512 + c] = W4[c, 509..511]
// (the Eikonal backward's per-row d s / d xyz = sum_c dY0[c] W0x[c] + dY4[c] W4x[c]).  1024 threads.
__device__ __forceinline__ void build_wx_elem(int i, const float* __restrict__ weff, float4* __restrict__ wx) {
  if (i >= 2 * kHidden) return;
  const int l = i < kHidden ? 0 : kSkipLayer, c = i & (kHidden - 1);
  float v[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int e = 0; e < 3; ++e)
    if (xyz_col(l, e) >= 0) v[e] = weff_at(weff, l, c, xyz_col(l, e));
  wx[i] = make_float4(v[0], v[1], v[2], 0.f);
}

// The three images are independent given weff: ONE launch (256 threads), blocks [0, kFoldFwdBlocks) build the forward
// stream, the next kFoldDgBlocks the dgrad stream, the last 4 the xyz-column table (ReLU family only).
constexpr unsigned kFoldFwdBlocks = (unsigned)(((long long)kW2Steps * (kStepBytes / 16) + 255) / 256);
constexpr unsigned kFoldDgBlocks = (unsigned)(((long long)kDg2Steps * (kStepBytes / 16) + 255) / 256);
constexpr unsigned kFoldWxBlocks = 4;
__global__ void __launch_bounds__(256) k_fold_images(const float* __restrict__ weff, const float* __restrict__ params,
                                                     ParamLayout pl, uint8_t* __restrict__ w2_img,
                                                     uint8_t* __restrict__ dg2_img, float4* __restrict__ wx) {
  const unsigned b = blockIdx.x;
  if (b < kFoldFwdBlocks) {
    fold_fwd2_chunk((long long)b * 256 + threadIdx.x, weff, params, pl, w2_img);
  } else if (b < kFoldFwdBlocks + kFoldDgBlocks) {
    fold_dgrad2_chunk((long long)(b - kFoldFwdBlocks) * 256 + threadIdx.x, weff, dg2_img);
  } else {
    build_wx_elem((int)((b - kFoldFwdBlocks - kFoldDgBlocks) * 256 + threadIdx.x), weff, wx);
  }
}

// Fold mode: per-segment E steps of lin0 / lin4 with bias' = b + W[:, latent] z_seg (fp32).
// grid (n_seg, 2, kFoldSplit), 512 threads: block z owns 512 / kFoldSplit output rows (so a single segment --
// reconstruct, grid decode -- still spreads over 8 SMs); warp w computes rows w, w+16, ... of its slice
// (coalesced over the latent).  seg_ids == nullptr: one segment whose latent is `table` itself.
constexpr int kFoldSplit = 4;
constexpr int kFoldRows = kHidden / kFoldSplit;
__global__ void __launch_bounds__(512) k_scene_fold(const float* __restrict__ weff,
                                                    const float* __restrict__ params, ParamLayout pl,
                                                    const float* __restrict__ table,
                                                    const long long* __restrict__ seg_ids,
                                                    uint8_t* __restrict__ scene_e) {
  __shared__ float z[kLatent];
  __shared__ float bias[kFoldRows];
  const int b = blockIdx.x, which = blockIdx.y, o0 = blockIdx.z * kFoldRows;
  const int l = which == 0 ? 0 : kSkipLayer;
  const int lat0 = which == 0 ? 0 : kX3;   // first latent column in the reference weight
  const float* zr = table + (seg_ids ? seg_ids[b] * kLatent : 0);
  if (threadIdx.x < kLatent) z[threadIdx.x] = zr[threadIdx.x];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp; r < kFoldRows; r += 16) {
    const int o = o0 + r;
    const float* wr = weff + weff_off(l) + (size_t)o * ref_in(l) + lat0;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < kLatent / 32; ++k) acc = fmaf(wr[lane + 32 * k], z[lane + 32 * k], acc);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    if (lane == 0) bias[r] = acc + params[pl.b[l] + o];
  }
  __syncthreads();
  if (threadIdx.x >= kFoldRows) return;
  const int o = o0 + threadIdx.x, h = o >> 8, n = o & 255;
  uint4 c0, c1;
  e_step_row(weff, l, o, true, bias[threadIdx.x], c0, c1);
  uint8_t* dst = scene_e + ((size_t)b * kSceneESteps + which * 2 + h) * kStepBytes +
                 (size_t)(n >> 3) * 256 + (size_t)(n & 7) * 16;
  *reinterpret_cast<uint4*>(dst) = c0;
  *reinterpret_cast<uint4*>(dst + 128) = c1;
}

}  // namespace dsdf

// Small kernels of the latent-only optimisation (frozen decoder) of reference
// reconstruct.py:16-151, batched over independent shapes:
//   k_recon_gather  deep_sdf/data.py:112-136 (unpack_sdf_samples_from_ram): the per-iteration sample
//                   window cat(pos[ps:ps+half], neg[ns:ns+half]) of every shape, written into the
//                   tile-padded [shape][rows_pad][4] buffer the forward chain reads
//   k_recon_update  reconstruct.py:102-140: loss value, regulariser gradients, torch.optim.Adam
//                   step on the latent and the optional hard norm bound, one block per shape
#pragma once
#include "fold_kernels.cuh"
#include "layout.cuh"
#include "small_kernels.cuh"

namespace dsdf {

// grid (ceil(rows_pad / 256), n_shapes), 256 threads.  Padding rows repeat the shape's first
// window row (finite coordinates; the forward kernel gives them zero loss and zero gradient).
__global__ void k_recon_gather(const float4* __restrict__ samples,
                               const long long* __restrict__ shape_base,   // [n_shapes][2] pos, neg row offsets
                               const long long* __restrict__ win,          // [n_shapes][2] window starts
                               long long half, long long rows_pad, float4* __restrict__ out,
                               long long* __restrict__ ids) {
  const int s = blockIdx.y;
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r == 0) ids[s] = s;
  if (r >= rows_pad) return;
  const long long pos0 = shape_base[2 * s] + win[2 * s];
  const long long neg0 = shape_base[2 * s + 1] + win[2 * s + 1];
  long long src;
  if (r < half) src = pos0 + r;
  else if (r < 2 * half) src = neg0 + (r - half);
  else src = pos0;
  out[(long long)s * rows_pad + r] = __ldg(samples + src);
}

struct ReconUpdateArgs {
  const float* seg_sum;        // [n_shapes][2][512] column sums of dY0 / dY4
  const float* weff;           // effective fp32 weights
  const float* tile_partials;  // [tiles][2]
  int tiles_per_seg;
  float inv_n;                 // 1 / num_samples (the L1 partial sums are unscaled)
  int reg_kind;                // 0 none, 1 lambda*mean(z^2), 2 lambda*||z||
  float reg_lambda;
  int dist_kind;               // 0 none, 1 mean(diff^2), 2 mean(|diff|)
  float dist_weight;
  const float* dist_mean;      // [256] or null
  const float* dist_std;       // [256] or null (already clamped >= 1e-8 by the host side)
  float step_size, bc2_sqrt, b1, b2, eps;   // Adam, torch.optim.Adam single-tensor arithmetic
  float code_bound;            // <= 0: none
  float* latents;              // [n_shapes][256] in/out
  float* exp_avg;
  float* exp_avg_sq;
  float* loss_out;             // [n_shapes] loss of this iteration (before the update)
  float* grad_out;             // [n_shapes][256] (nullable) dL/dz of this iteration
};

// grid n_shapes, 1024 threads: thread (part = t / 256, j = t % 256)
__global__ void __launch_bounds__(1024) k_recon_update(const ReconUpdateArgs a) {
  __shared__ float s0[kHidden], s4[kHidden], accs[4][kLatent], sh[8];
  const int b = blockIdx.x, t = threadIdx.x, j = t & 255, part = t >> 8;
  for (int c = t; c < 2 * kHidden; c += 1024) {
    const float v = a.seg_sum[(size_t)b * 2 * kHidden + c];
    if (c < kHidden) s0[c] = v; else s4[c - kHidden] = v;
  }
  __syncthreads();
  {
    const float* w0 = a.weff + weff_off(0) + j;                 // [512][259], latent columns 0..255
    const float* w4 = a.weff + weff_off(4) + kX3 + j;           // [512][512], latent columns 253..508
    float acc0 = 0.f, acc4 = 0.f;
    const int c0 = part * (kHidden / 4);
#pragma unroll 8
    for (int c = c0; c < c0 + kHidden / 4; ++c) {
      acc0 = fmaf(s0[c], __ldg(w0 + (size_t)c * kIn0), acc0);
      acc4 = fmaf(s4[c], __ldg(w4 + (size_t)c * kHidden), acc4);
    }
    accs[part][j] = acc0 + acc4;
  }
  __syncthreads();
  if (t >= 256) return;   // the rest is one thread per latent element (no block-wide barrier below
                          // other than block_sum_256-style reductions over these 256 threads)
  float g = (accs[0][j] + accs[1][j]) + (accs[2][j] + accs[3][j]);
  float* zp = a.latents + (size_t)b * kLatent + j;
  const float z = *zp;
  // ---- loss value + regulariser gradients (reconstruct.py:102-129)
  auto sum256 = [&](float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    named_bar_sync(2, 256);
    if ((t & 31) == 0) sh[t >> 5] = v;
    named_bar_sync(2, 256);
    float r = 0.f;
    for (int i = 0; i < 8; ++i) r += sh[i];
    return r;
  };
  float l1 = 0.f;
  for (int i = t; i < a.tiles_per_seg; i += 256)
    l1 += a.tile_partials[((size_t)b * a.tiles_per_seg + i) * 2];
  float loss = sum256(l1) * a.inv_n;
  const float zz = sum256(z * z);
  if (a.reg_kind == 1) {
    loss += a.reg_lambda * zz * (1.f / kLatent);
    g += a.reg_lambda * 2.f * z * (1.f / kLatent);
  } else if (a.reg_kind == 2) {
    const float nrm = sqrtf(zz);
    loss += a.reg_lambda * nrm;
    if (nrm > 0.f) g += a.reg_lambda * z / nrm;
  }
  if (a.dist_kind != 0 && a.dist_mean != nullptr) {
    const float inv_s = a.dist_std ? 1.f / a.dist_std[j] : 1.f;
    const float diff = (z - a.dist_mean[j]) * inv_s;
    if (a.dist_kind == 1) {
      loss += a.dist_weight * sum256(diff * diff) * (1.f / kLatent);
      g += a.dist_weight * 2.f * diff * inv_s * (1.f / kLatent);
    } else {
      loss += a.dist_weight * sum256(fabsf(diff)) * (1.f / kLatent);
      const float sg = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);
      g += a.dist_weight * sg * inv_s * (1.f / kLatent);
    }
  }
  if (t == 0) a.loss_out[b] = loss;
  if (a.grad_out) a.grad_out[(size_t)b * kLatent + j] = g;
  // ---- Adam (reconstruct.py:56,131), same arithmetic as k_adam
  float* mp = a.exp_avg + (size_t)b * kLatent + j;
  float* vp = a.exp_avg_sq + (size_t)b * kLatent + j;
  const float mi = fmaf(1.f - a.b1, g - *mp, *mp);
  const float vi = fmaf(1.f - a.b2, g * g, a.b2 * *vp);
  *mp = mi;
  *vp = vi;
  float zn = z - a.step_size * (mi / (sqrtf(vi) / a.bc2_sqrt + a.eps));
  // ---- hard norm bound (reconstruct.py:134-140)
  if (a.code_bound > 0.f) {
    const float n = sqrtf(sum256(zn * zn));
    zn *= fminf(a.code_bound / (n + 1e-12f), 1.f);
  }
  *zp = zn;
}

}  // namespace dsdf

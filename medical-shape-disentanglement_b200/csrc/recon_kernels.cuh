// Small kernels of the latent-only optimisation (frozen decoder) of reference
// reconstruct.py:16-151, batched over independent shapes:
//   k_recon_gather  deep_sdf/data.py:112-136 (unpack_sdf_samples_from_ram): the per-iteration sample
//                   window cat(pos[ps:ps+half], neg[ns:ns+half]) of every shape, written into the
//                   tile-padded [shape][rows_pad][4] buffer the forward chain reads
//   k_recon_dlat_partial  the shape's latent gradient sum_rows dX[:, latent] = colsum(dY0) W0[:, lat] +
//                   colsum(dY4) W4[:, lat], cut into kReconParts slices of the 512 hidden channels so that one
//                   shape keeps 8 SMs busy instead of one (the loop is latency-bound: a single shape is the
//                   reference's use case)
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

constexpr int kReconParts = 8;                      // slices of the hidden dimension
constexpr int kReconPartC = kHidden / kReconParts;  // 64 channels each

// grid (n_shapes, kReconParts), 256 threads.  Phase 1: column sums of dY0 / dY4 over the shape's tiles (in tile
// order) for this slice's channels; phase 2: their product with the latent columns of W0 / W4 (thread = latent j).
__global__ void __launch_bounds__(256) k_recon_dlat_partial(const float* __restrict__ tile_colsum, int tiles_per_seg,
                                                            const float* __restrict__ weff,
                                                            float* __restrict__ dlat_part) {
  __shared__ float sc[2][kReconPartC];
  const int b = blockIdx.x, part = blockIdx.y, t = threadIdx.x;
  const int cbase = part * kReconPartC;
  if (t < 2 * kReconPartC) {
    const int which = t / kReconPartC, c = cbase + t % kReconPartC;
    const float* src = tile_colsum + ((size_t)b * tiles_per_seg * 2 + which) * kHidden + c;
    float s = 0.f;
#pragma unroll 8
    for (int tt = 0; tt < tiles_per_seg; ++tt) s += src[(size_t)tt * 2 * kHidden];
    sc[which][t % kReconPartC] = s;
  }
  __syncthreads();
  const float* w0 = weff + weff_off(0) + (size_t)cbase * kIn0 + t;                  // [512][259], latent cols 0..255
  const float* w4 = weff + weff_off(4) + (size_t)cbase * kHidden + kX3 + t;         // [512][512], latent cols 253..508
  float acc0 = 0.f, acc4 = 0.f;
#pragma unroll 16
  for (int c = 0; c < kReconPartC; ++c) {
    acc0 = fmaf(sc[0][c], __ldg(w0 + (size_t)c * kIn0), acc0);
    acc4 = fmaf(sc[1][c], __ldg(w4 + (size_t)c * kHidden), acc4);
  }
  dlat_part[((size_t)b * kReconParts + part) * kLatent + t] = acc0 + acc4;
}

struct ReconUpdateArgs {
  const float* dlat_part;      // [n_shapes][kReconParts][256] from k_recon_dlat_partial
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

// grid n_shapes, 256 threads: thread = latent element j
__global__ void __launch_bounds__(256) k_recon_update(const ReconUpdateArgs a) {
  __shared__ float sh[8];
  const int b = blockIdx.x, t = threadIdx.x, j = t;
  float g = 0.f;
#pragma unroll
  for (int p = 0; p < kReconParts; ++p) g += a.dlat_part[((size_t)b * kReconParts + p) * kLatent + j];   // fixed order
  float* zp = a.latents + (size_t)b * kLatent + j;
  const float z = *zp;
  // ---- loss value + regulariser gradients (reconstruct.py:102-129)
  auto sum256 = [&](float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((t & 31) == 0) sh[t >> 5] = v;
    __syncthreads();
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

// The small-kernel side of the latent-only optimisation (frozen decoder) of reference reconstruct.py:16-151,
// batched over independent shapes.  The loop is strictly sequential and latency-bound (the reference's use case is
// ONE shape: 63 tiles on 64 SMs), so everything between one iteration's dgrad chain and the next iteration's
// forward chain is ONE launch, k_recon_tail, run by a cluster of 8 CTAs per shape:
//   A  the shape's latent gradient sum_rows dX[:, latent] = colsum(dY0) W0[:, lat] + colsum(dY4) W4[:, lat]: CTA p
//      takes 64 of the 512 hidden channels (column sums over the shape's tiles, then the product with the latent
//      columns of W0 / W4) and leaves its partial [256] in shared memory;
//   B  every CTA sums the 8 partials in rank order over distributed shared memory and runs reconstruct.py:102-140
//      (loss value, regulariser gradients, torch.optim.Adam step, optional hard norm bound) redundantly -- identical
//      arithmetic, so identical new latents in all 8 CTAs; rank 0 writes the state;
//   C  the NEXT iteration's per-shape E step (what k_scene_fold does for training segments): CTA p folds
//      128 of the 2 x 512 latent products with the new latent;
//   D  the NEXT iteration's sample window cat(pos[ps:ps+half], neg[ns:ns+half]) (deep_sdf/data.py:112-136,
//      unpack_sdf_samples_from_ram) into the tile-padded [shape][rows_pad][4] buffer the forward chain reads.
// The first launch of a call runs C + D only (do_update = 0), the last A + B only (do_next = 0).
#pragma once
#include "fold_kernels.cuh"
#include "layout.cuh"
#include "small_kernels.cuh"

namespace dsdf {

constexpr int kReconParts = 8;                      // CTAs per shape = cluster size = slices of the hidden dimension
constexpr int kReconPartC = kHidden / kReconParts;  // 64 channels each
constexpr int kReconThreads = 512;
static_assert(kReconParts == 2 * kFoldSplit, "phase C maps CTA p to (layer p / kFoldSplit, row slice p % kFoldSplit)");

struct ReconTailArgs {
  int do_update, do_next;
  // ---- A + B
  const float* tile_colsum;    // [tiles][2][512] column sums of dY0 / dY4 per tile (k_backward2)
  const float* tile_partials;  // [tiles][2] L1 partial sums per tile (k_forward3)
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
  // ---- C
  const float* weff;
  const float* params;
  ParamLayout pl;
  uint8_t* scene_e;
  // ---- D
  const float4* samples;
  const long long* shape_base; // [n_shapes][2] pos, neg row offsets
  const long long* win;        // [n_shapes][2] window starts of the NEXT iteration
  long long half, rows_pad;
  float4* sdf;                 // [n_shapes][rows_pad] out
  long long* ids;              // [n_shapes] out: segment s reads latent row s
};

// grid (kReconParts, n_shapes), cluster (kReconParts, 1, 1), 512 threads
__global__ void __launch_bounds__(kReconThreads) k_recon_tail(const ReconTailArgs a) {
  __shared__ float sc4[4][2 * kReconPartC];
  __shared__ float sc[2][kReconPartC];
  __shared__ float part[kLatent];      // read by the whole cluster in phase B
  __shared__ float hi_half[kLatent];
  __shared__ float z_s[kLatent];
  __shared__ float bias[kFoldRows];
  __shared__ float sh[kReconThreads / 32];
  const int p = blockIdx.x, b = blockIdx.y, t = threadIdx.x, j = t & (kLatent - 1);
  const bool lo = t < kLatent;
  grid_dep_launch();
  grid_dep_wait();
  float z = 0.f, m = 0.f, v = 0.f;
  if (a.do_update) {
    if (lo) {   // every CTA reads the old state BEFORE the cluster barrier, rank 0 writes it after
      z = a.latents[(size_t)b * kLatent + j];
      m = a.exp_avg[(size_t)b * kLatent + j];
      v = a.exp_avg_sq[(size_t)b * kLatent + j];
    }
    // ---- A1: column sums of this CTA's 64 channels of dY0 / dY4; 4 thread groups take a quarter of the tiles each
    {
      const int q = t >> 7, i = t & 127, which = i / kReconPartC, c = p * kReconPartC + i % kReconPartC;
      const int t0 = q * a.tiles_per_seg / 4, t1 = (q + 1) * a.tiles_per_seg / 4;
      const float* src = a.tile_colsum + ((size_t)b * a.tiles_per_seg * 2 + which) * kHidden + c;
      float s = 0.f;
#pragma unroll 8
      for (int tt = t0; tt < t1; ++tt) s += src[(size_t)tt * 2 * kHidden];
      sc4[q][i] = s;
    }
    __syncthreads();
    if (t < 2 * kReconPartC) sc[t / kReconPartC][t % kReconPartC] = ((sc4[0][t] + sc4[1][t]) + sc4[2][t]) + sc4[3][t];
    __syncthreads();
    // ---- A2: product with the latent columns of W0 / W4; thread = (channel half, latent j)
    {
      const int ch = (t >> 8) * (kReconPartC / 2);
      const int cbase = p * kReconPartC + ch;
      const float* w0 = a.weff + weff_off(0) + (size_t)cbase * kIn0 + j;                // [512][259], latent cols 0..255
      const float* w4 = a.weff + weff_off(4) + (size_t)cbase * kHidden + kX3 + j;       // [512][512], latent cols 253..508
      float acc0 = 0.f, acc4 = 0.f;
#pragma unroll 16
      for (int c = 0; c < kReconPartC / 2; ++c) {
        acc0 = fmaf(sc[0][ch + c], __ldg(w0 + (size_t)c * kIn0), acc0);
        acc4 = fmaf(sc[1][ch + c], __ldg(w4 + (size_t)c * kHidden), acc4);
      }
      if (!lo) hi_half[j] = acc0 + acc4;
      __syncthreads();
      if (lo) part[j] = (acc0 + acc4) + hi_half[j];
    }
  }
  cluster_sync_all();
  if (a.do_update) {
    float g = 0.f;
    if (lo) {
      const uint32_t mine = smem_u32(&part[j]);
#pragma unroll
      for (int r = 0; r < kReconParts; ++r) g += ld_shared_cluster_f32(mapa_u32(mine, r));   // fixed order
    }
    cluster_sync_all();          // no CTA may leave (or reuse `part`) while a peer still reads it
    // ---- B: loss value + regulariser gradients (reconstruct.py:102-129); threads >= 256 only feed zeros
    auto sum_block = [&](float x) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      __syncthreads();
      if ((t & 31) == 0) sh[t >> 5] = x;
      __syncthreads();
      float r = 0.f;
      for (int i = 0; i < kReconThreads / 32; ++i) r += sh[i];
      return r;
    };
    float l1 = 0.f;
    for (int i = t; i < a.tiles_per_seg; i += kReconThreads)
      l1 += a.tile_partials[((size_t)b * a.tiles_per_seg + i) * 2];
    float loss = sum_block(l1) * a.inv_n;
    const float zz = sum_block(z * z);
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
      const float diff = lo ? (z - a.dist_mean[j]) * inv_s : 0.f;
      if (a.dist_kind == 1) {
        loss += a.dist_weight * sum_block(diff * diff) * (1.f / kLatent);
        g += a.dist_weight * 2.f * diff * inv_s * (1.f / kLatent);
      } else {
        loss += a.dist_weight * sum_block(fabsf(diff)) * (1.f / kLatent);
        const float sg = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);
        g += a.dist_weight * sg * inv_s * (1.f / kLatent);
      }
    }
    // ---- Adam (reconstruct.py:56,131), same arithmetic as k_adam
    const float mi = fmaf(1.f - a.b1, g - m, m);
    const float vi = fmaf(1.f - a.b2, g * g, a.b2 * v);
    float zn = z - a.step_size * (mi / (sqrtf(vi) / a.bc2_sqrt + a.eps));
    if (!lo) zn = 0.f;
    // ---- hard norm bound (reconstruct.py:134-140)
    if (a.code_bound > 0.f) {
      const float n = sqrtf(sum_block(zn * zn));
      zn *= fminf(a.code_bound / (n + 1e-12f), 1.f);
    }
    if (lo) z_s[j] = zn;
    if (p == 0) {
      if (t == 0) a.loss_out[b] = loss;
      if (lo) {
        a.latents[(size_t)b * kLatent + j] = zn;
        a.exp_avg[(size_t)b * kLatent + j] = mi;
        a.exp_avg_sq[(size_t)b * kLatent + j] = vi;
      }
    }
  } else if (lo) {
    z_s[j] = a.latents[(size_t)b * kLatent + j];
  }
  if (!a.do_next) return;
  __syncthreads();
  // ---- C: this CTA's 128 rows of the next iteration's E step (same arithmetic and order as k_scene_fold)
  {
    const int which = p / kFoldSplit, o0 = (p % kFoldSplit) * kFoldRows;
    const int l = which == 0 ? 0 : kSkipLayer;
    const int lat0 = which == 0 ? 0 : kX3;   // first latent column in the reference weight
    const int warp = t >> 5, lane = t & 31;
    for (int r = warp; r < kFoldRows; r += kReconThreads / 32) {
      const int o = o0 + r;
      const float* wr = a.weff + weff_off(l) + (size_t)o * ref_in(l) + lat0;
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < kLatent / 32; ++k) acc = fmaf(wr[lane + 32 * k], z_s[lane + 32 * k], acc);
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
      if (lane == 0) bias[r] = acc + a.params[a.pl.b[l] + o];
    }
    __syncthreads();
    if (t < kFoldRows) {
      const int o = o0 + t, h = o >> 8, n = o & 255;
      uint4 c0, c1;
      e_step_row(a.weff, l, o, true, bias[t], c0, c1);
      uint8_t* dst = a.scene_e + ((size_t)b * kSceneESteps + which * 2 + h) * kStepBytes +
                     (size_t)(n >> 3) * 256 + (size_t)(n & 7) * 16;
      *reinterpret_cast<uint4*>(dst) = c0;
      *reinterpret_cast<uint4*>(dst + 128) = c1;
    }
  }
  // ---- D: the next iteration's window.  Padding rows repeat the shape's first window row (finite coordinates;
  // the forward kernel gives them zero loss and zero gradient).
  {
    if (p == 0 && t == 0) a.ids[b] = b;
    const long long pos0 = a.shape_base[2 * b] + a.win[2 * b];
    const long long neg0 = a.shape_base[2 * b + 1] + a.win[2 * b + 1];
    for (long long r = (long long)p * kReconThreads + t; r < a.rows_pad; r += (long long)kReconParts * kReconThreads) {
      long long src;
      if (r < a.half) src = pos0 + r;
      else if (r < 2 * a.half) src = neg0 + (r - a.half);
      else src = pos0;
      a.sdf[(long long)b * a.rows_pad + r] = __ldg(a.samples + src);
    }
  }
}

}  // namespace dsdf

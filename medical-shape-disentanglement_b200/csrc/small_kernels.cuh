// HBM-bound / tiny kernels around the tensor-core chain: embedding max_norm renorm, segmented
// (scene-contiguous) latent-gradient reduction, split-K wgrad reduction, weight-norm backward,
// deterministic loss / grad-norm reductions, fused clip + Adam.
//
// Reference semantics restated (train_deep_sdf.py):
//   :429,598  nn.Embedding(max_norm) in-place renorm of looked-up rows: row *= max/(norm+1e-7)
//   :609-613  code regulariser lambda*min(1,epoch/100)*sum_rows||z_row|| / N_total
//   :731      embedding_dense_backward (here: deterministic, no float atomics)
//   :780-782  clip_grad_norm_: coef = min(1, c/(||g||_2 + 1e-6)) over decoder params only
//   :784      Adam(betas=(.9,.999), eps=1e-8) over decoder params and ALL latent rows
#pragma once
#include "common.cuh"
#include "fold_kernels.cuh"
#include "layout.cuh"
#include "ptx.cuh"
#include "wgrad2_kernel.cuh"

namespace dsdf {

__device__ __forceinline__ float block_sum_256(float v, float* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  float t = 0.f;
  const int nw = blockDim.x >> 5;
  for (int i = 0; i < nw; ++i) t += sh[i];   // fixed order -> deterministic
  return t;
}

// ---- embedding max_norm: grid n_seg, 256 threads.  A scene that appears more than once in the batch is
// handled by the block of its FIRST occurrence only (renorm is idempotent -- torch renorms the unique ids --
// and two blocks rescaling one row concurrently would mix old and new elements in the norm).
__device__ __forceinline__ bool first_occurrence(const long long* __restrict__ seg_ids, int b) {
  const long long id = seg_ids[b];
  for (int k = 0; k < b; ++k)
    if (seg_ids[k] == id) return false;
  return true;
}

__global__ void k_renorm_rows(float* __restrict__ table, const long long* __restrict__ seg_ids,
                              int n_seg, long long num_scenes, float max_norm) {
  __shared__ float sh[8];
  const int b = blockIdx.x;
  if (b >= n_seg || !first_occurrence(seg_ids, b)) return;      // block-uniform
  if (seg_ids[b] < 0 || seg_ids[b] >= num_scenes) return;       // reported by k_sanitize_ids in the step itself
  float* row = table + seg_ids[b] * kLatent;
  const float x = row[threadIdx.x];
  const float nrm = sqrtf(block_sum_256(x * x, sh));
  if (nrm > max_norm) row[threadIdx.x] = x * (max_norm / (nrm + 1e-7f));
}

// ---- scene-id guard: nn.Embedding raises on an index outside [0, num_embeddings) (a device-side assert on CUDA).
// Here: ids are copied into the workspace with bad ones replaced by 0 (so no kernel of the step can index the
// latent table out of bounds) and status[0] = 1 + position of the first bad id (0 = all good); the host wrapper
// turns a non-zero status into an exception when it reads the step's scalars.  One block.
__global__ void k_sanitize_ids(const long long* __restrict__ ids, int n_seg, long long num_scenes,
                               long long* __restrict__ out, float* __restrict__ status) {
  __shared__ int first_bad;
  if (threadIdx.x == 0) first_bad = 0x7fffffff;
  __syncthreads();
  for (int b = threadIdx.x; b < n_seg; b += blockDim.x) {
    const long long id = ids[b];
    const bool ok = id >= 0 && id < num_scenes;
    out[b] = ok ? id : 0;
    if (!ok) atomicMin(&first_bad, b);
  }
  __syncthreads();
  if (threadIdx.x == 0) status[0] = first_bad == 0x7fffffff ? 0.f : (float)(first_bad + 1);
}

// ---- per-segment column sums of dY0 / dY4: seg_sum[b][which][c] -----------------------------
// grid (n_seg, 2), 512 threads.  tiles of segment b: [b*tps, (b+1)*tps), summed in order.
__global__ void k_seg_colsum(const float* __restrict__ tile_colsum, int tiles_per_seg,
                             float* __restrict__ seg_sum) {
  const int b = blockIdx.x, which = blockIdx.y, c = threadIdx.x;
  const float* src = tile_colsum + ((size_t)b * tiles_per_seg * 2 + which) * kHidden + c;
  float s = 0.f;
#pragma unroll 16
  for (int t = 0; t < tiles_per_seg; ++t) s += src[(size_t)t * 2 * kHidden];
  seg_sum[((size_t)b * 2 + which) * kHidden + c] = s;
}

// ---- latent gradient per segment ---------------------------------------------------------------
// dlat[b][j] = sum_c S0[b][c] W0[c][j] + sum_c S4[b][c] W4[c][kX3+j] + reg_coef * z[j]/||z||
// grid n_seg, 1024 threads: thread (part = t / 256, j = t % 256) sums a quarter of c; fixed-order combine
__global__ void __launch_bounds__(1024) k_lat_grad_seg(const float* __restrict__ seg_sum,
                                                       const float* __restrict__ weff,
                                                       const float* __restrict__ table,
                                                       const long long* __restrict__ seg_ids, float reg_coef,
                                                       float* __restrict__ dlat_seg, float* __restrict__ seg_norm) {
  __shared__ float s0[kHidden], s4[kHidden], accs[4][kLatent], sh[8];
  const int b = blockIdx.x, t = threadIdx.x, j = t & 255, part = t >> 8;
  for (int c = t; c < 2 * kHidden; c += 1024) {
    const float v = seg_sum[(size_t)b * 2 * kHidden + c];
    if (c < kHidden) s0[c] = v; else s4[c - kHidden] = v;
  }
  __syncthreads();
  const float* w0 = weff + weff_off(0) + j;             // [512][259]
  const float* w4 = weff + weff_off(4) + kX3 + j;       // [512][512]
  float a0 = 0.f, a4 = 0.f;
  const int c0 = part * (kHidden / 4);
#pragma unroll 16
  for (int c = c0; c < c0 + kHidden / 4; ++c) {
    a0 = fmaf(s0[c], __ldg(w0 + (size_t)c * kIn0), a0);
    a4 = fmaf(s4[c], __ldg(w4 + (size_t)c * kHidden), a4);
  }
  accs[part][j] = a0 + a4;
  __syncthreads();
  if (t >= 256) return;
  float acc = (accs[0][j] + accs[1][j]) + (accs[2][j] + accs[3][j]);
  const float z = table[seg_ids[b] * kLatent + j];
  float v = z * z;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((t & 31) == 0) sh[t >> 5] = v;
  named_bar_sync(2, 256);
  float zz = 0.f;
  for (int i = 0; i < 8; ++i) zz += sh[i];
  const float nrm = sqrtf(zz);
  if (reg_coef != 0.f && nrm > 0.f) acc += reg_coef * z / nrm;
  dlat_seg[(size_t)b * kLatent + j] = acc;
  if (j == 0) seg_norm[b] = nrm;
}

// dense latent grad rows = per-segment grads; duplicate ids are summed in segment order by the block of the
// id's first occurrence (deterministic).  grid n_seg, 256 threads.  lat_grad must be zeroed by the caller.
__global__ void k_lat_grad_scatter(const float* __restrict__ dlat_seg,
                                   const long long* __restrict__ seg_ids, int n_seg,
                                   float* __restrict__ lat_grad) {
  const int b = blockIdx.x;
  if (b >= n_seg || !first_occurrence(seg_ids, b)) return;      // block-uniform
  const long long id = seg_ids[b];
  float acc = dlat_seg[(size_t)b * kLatent + threadIdx.x];
  for (int k = b + 1; k < n_seg; ++k)
    if (seg_ids[k] == id) acc += dlat_seg[(size_t)k * kLatent + threadIdx.x];
  lat_grad[id * kLatent + threadIdx.x] = acc;
}

// ---- latent columns of dW0 / dW4: outer products of segment colsums and latents ---------------
// grid (512, 2), 256 threads: dW[which][c][lat col j] = sum_b S[b][which][c] * z_b[j]
__global__ void k_dw_latent(const float* __restrict__ seg_sum, const float* __restrict__ table,
                            const long long* __restrict__ seg_ids, int n_seg,
                            float* __restrict__ dweff) {
  const int c = blockIdx.x, which = blockIdx.y, j = threadIdx.x;
  float acc = 0.f;
  for (int b = 0; b < n_seg; ++b)
    acc = fmaf(seg_sum[((size_t)b * 2 + which) * kHidden + c], table[seg_ids[b] * kLatent + j], acc);
  if (which == 0) dweff[weff_off(0) + (size_t)c * kIn0 + j] = acc;
  else dweff[weff_off(4) + (size_t)c * kHidden + kX3 + j] = acc;
}

// ---- split-K wgrad partial reduction into the reference-layout fp32 gradient ----------------
// v2 schedule: item (layer l, out-chunk pair) -> contiguous partial slots [slot_lo, slot_hi)
struct Wg2Layout {
  int item_base[9];   // first item of layer l (1..8)
  int slot_lo[16];
  int slot_hi[16];
};
// grid (512 rows o, 8 layers; 7 when dW8 comes from the forward kernel), 512 threads (i pad).
// Fixed summation order -> deterministic.
__global__ void k_reduce_wgrad2(const float* __restrict__ partials, Wg2Layout wl,
                                float* __restrict__ dweff) {
  const int l = blockIdx.y + 1, o = blockIdx.x, ip = threadIdx.x;
  if (l == 8) {
    if (o != 0) return;
    const int item = wl.item_base[8];
    float s = 0.f;
    for (int k = wl.slot_lo[item]; k < wl.slot_hi[item]; ++k) {
      const float* p = partials + (size_t)k * 256 * kHidden + ip;
      s += p[0] + p[kHidden];   // D row 0 = bf16-hi part of g, row 1 = lo part
    }
    dweff[weff_off(8) + ip] = s;
    return;
  }
  if (o >= ref_out(l)) return;
  const int npad = (l == 4) ? 256 : 512;
  if (ip >= npad) return;
  const int ic = pad_to_ref_col(l, ip);
  if (ic < 0) return;
  const int item = wl.item_base[l] + (o >> 8);
  float s = 0.f;
  for (int k = wl.slot_lo[item]; k < wl.slot_hi[item]; ++k)
    s += partials[((size_t)k * 256 + (o & 255)) * kHidden + ip];
  dweff[weff_off(l) + (size_t)o * ref_in(l) + ic] = s;
}

// ---- Eikonal term: row-weighted column sums of the stashed dY_l, l = 1..7 ---------------------------------
// The dgrad chain runs with the L1 upstream g; the Eikonal term adds eps_r g_r on row r (known only at the end of
// the chain, tests/eikonal_algebra.py), so every sum over rows that is not a wgrad GEMM needs a correction:
//   db_l      += sum_r eps_r dY_l[r, :]                       l in {1,2,3,5,6,7}  -> corr[block][slot][512]
//   colsum(dY4) per tile += sum_r eps_r dY4[r, :]             (-> db4, latent grads, latent columns of dW4)
//   dW4[:, xyz] += sum_r dY4[r, :] (xyz'_r - xyz_r)                                -> corr[block][6*512 + c*3 + e]
// (layer 0 is weighted inside k_backward2<EIK>; the weight matrices get theirs through the recompute pass.)
// HBM-bound: reads the dY stash once (6.5 x 128 KB per tile).  grid = any, 256 threads; block b takes tiles
// b, b + grid, ...; thread = (16-byte chunk ch = t & 7, rows (t >> 3) + 32 i): a warp reads 4 full 128-byte rows.
constexpr int kCorrFloats = 6 * kHidden + kHidden * 3;
__global__ void __launch_bounds__(256, 4) k_eik_corr(const uint8_t* __restrict__ dystash,
                                                     const float4* __restrict__ eik_row,
                                                     const float* __restrict__ sdf_data, long long n_tiles,
                                                     float* __restrict__ tile_colsum, float* __restrict__ corr,
                                                     const float* __restrict__ g = nullptr,
                                                     const uint32_t* __restrict__ relu_mask = nullptr,
                                                     const float* __restrict__ w8 = nullptr) {
  __shared__ float w_eps[kTileM], w_dx[kTileM], w_dy[kTileM], w_dz[kTileM];
  __shared__ float red[8][4][64];
  __shared__ float db_acc[6][kHidden];
  __shared__ float dwx_acc[kHidden * 3];
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const uint32_t ch = t & 7, rg = t >> 3;
  for (int i = t; i < 6 * kHidden; i += 256) (&db_acc[0][0])[i] = 0.f;
  for (int i = t; i < 3 * kHidden; i += 256) dwx_acc[i] = 0.f;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    __syncthreads();
    if (t < kTileM) {
      const long long row = tile * kTileM + t;
      const float4 e = __ldg(eik_row + row);
      const float4 x = __ldg(reinterpret_cast<const float4*>(sdf_data) + row);
      w_eps[t] = e.w; w_dx[t] = e.x - x.x; w_dy[t] = e.y - x.y; w_dz[t] = e.z - x.z;
    }
    __syncthreads();
    const uint8_t* dy_tile = dystash + (size_t)tile * kStashTileBytes;
#pragma unroll 1
    for (int l = 1; l <= (g ? 6 : 7); ++l) {
      const int nsub = l == 3 ? 4 : 8;
      const bool skip = l == kSkipLayer;
#pragma unroll 1
      for (int sub = 0; sub < nsub; ++sub) {
        const uint8_t* base = dy_tile + (size_t)l * kTileBytes + (size_t)sub * kSubBytes;
        uint4 v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
          v[i] = __ldg(reinterpret_cast<const uint4*>(base + sub_chunk_off(rg + 32 * i, ch)));
        float acc[4][8];
#pragma unroll
        for (int w = 0; w < 4; ++w)
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[w][k] = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t r = rg + 32 * i;
          const uint32_t wd[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
          const float we = w_eps[r];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float lo = bf16_lo(wd[k]), hi = bf16_hi(wd[k]);
            acc[0][2 * k] = fmaf(we, lo, acc[0][2 * k]);
            acc[0][2 * k + 1] = fmaf(we, hi, acc[0][2 * k + 1]);
          }
          if (skip && kXyzAtSkip) {
            const float wx = w_dx[r], wy = w_dy[r], wz = w_dz[r];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float lo = bf16_lo(wd[k]), hi = bf16_hi(wd[k]);
              acc[1][2 * k] = fmaf(wx, lo, acc[1][2 * k]); acc[1][2 * k + 1] = fmaf(wx, hi, acc[1][2 * k + 1]);
              acc[2][2 * k] = fmaf(wy, lo, acc[2][2 * k]); acc[2][2 * k + 1] = fmaf(wy, hi, acc[2][2 * k + 1]);
              acc[3][2 * k] = fmaf(wz, lo, acc[3][2 * k]); acc[3][2 * k + 1] = fmaf(wz, hi, acc[3][2 * k + 1]);
            }
          }
        }
        const int nw = (skip && kXyzAtSkip) ? 4 : 1;
        // the 4 row groups of a warp (lane bits 3, 4), fixed order
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          if (w < nw) {                   // block-uniform
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              float a0 = acc[w][k];
              a0 += __shfl_xor_sync(0xffffffffu, a0, 8);
              a0 += __shfl_xor_sync(0xffffffffu, a0, 16);
              if (lane < 8) red[warp][w][ch * 8 + k] = a0;
            }
          }
        }
        __syncthreads();
        if (t < 64 * nw) {
          const int w = t >> 6, c = t & 63;
          float sum = 0.f;
#pragma unroll
          for (int k = 0; k < 8; ++k) sum += red[k][w][c];
          const int col = sub * 64 + c;
          if (w == 0) {
            if (skip) tile_colsum[((size_t)tile * 2 + 1) * kHidden + col] += sum;   // single writer per (tile, col)
            else db_acc[db_slot(l)][col] += sum;
          } else {
            dwx_acc[col * 3 + (w - 1)] += sum;
          }
        }
        __syncthreads();
      }
    }
    if (g) {
      // lin7 when the dgrad chain does not stash dY7 (the wgrad kernel rebuilds it): dY7[r, c] = g_r w8[c] where the
      // ReLU bit of X8[r, c] is set, so sum_r eps_r dY7[r, c] = w8[c] sum_{r: bit} eps_r g_r -- 2 KB of mask words per
      // tile instead of 128 KB of dY7.  thread = (32-column block cb, 8 rows); bit 31 - j of a word = column 32 cb + j
      const int cb = t >> 4, rq = t & 15;
      const uint32_t* mt = relu_mask + (size_t)tile * (8 * 16 * kTileM) + (size_t)(7 * 16 + cb) * kTileM;
      float acc7[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) acc7[j] = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = rq + 16 * i;
        const uint32_t m = __ldg(mt + r);
        const float v = w_eps[r] * __ldg(g + tile * kTileM + r);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc7[j] += (m & (0x80000000u >> j)) ? v : 0.f;
      }
      // the 16 row groups of a column block are 16 consecutive lanes: fixed-order butterfly, lane rq == 0 holds the sum
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float a0 = acc7[j];
        a0 += __shfl_xor_sync(0xffffffffu, a0, 1);
        a0 += __shfl_xor_sync(0xffffffffu, a0, 2);
        a0 += __shfl_xor_sync(0xffffffffu, a0, 4);
        a0 += __shfl_xor_sync(0xffffffffu, a0, 8);
        if (rq == 0) db_acc[db_slot(7)][cb * 32 + j] += a0 * __ldg(w8 + cb * 32 + j);
      }
    }
  }
  __syncthreads();
  float* out = corr + (size_t)blockIdx.x * kCorrFloats;
  for (int i = t; i < 6 * kHidden; i += 256) out[i] = (&db_acc[0][0])[i];
  for (int i = t; i < 3 * kHidden; i += 256) out[6 * kHidden + i] = dwx_acc[i];
}

// ---- reducer-warp partials (per CTA) -> bias grads, dW8, xyz columns of dW0/dW4 -----------------
// grid 16 jobs x 512 threads (column c); every job sums its partials in a fixed order (deterministic):
//   0..5   db of lin {1,2,3,5,6,7}        6..11  xyz column e of dW{0,4}   (job - 6 = 3 i + e)
//   12,13  db0 / db4 from the segment column sums      14  db8 = sum_rows g      15  dW8 (forward partials)
__global__ void k_reduce_cta(const float* __restrict__ cta_partials, int n_cta,
                             const float* __restrict__ seg_sum, int n_seg,
                             const float* __restrict__ tile_partials, long long n_tiles,
                             ParamLayout pl, float* __restrict__ grads,
                             float* __restrict__ dweff, int n_cta_dw8,
                             const float* __restrict__ corr = nullptr, int n_corr = 0,
                             const float* __restrict__ eik_tile = nullptr) {
  __shared__ float sh[16];
  const int c = threadIdx.x, job = blockIdx.x;
  if (job < 6) {
    const int ls[6] = {1, 2, 3, 5, 6, 7};
    float s = 0.f;
#pragma unroll 16
    for (int k = 0; k < n_cta; ++k) s += cta_partials[(size_t)k * kCtaPartialFloats + kPartDb + job * kHidden + c];
#pragma unroll 16
    for (int k = 0; k < n_corr; ++k) s += corr[(size_t)k * kCorrFloats + job * kHidden + c];   // Eikonal: sum eps dY
    if (c < ref_out(ls[job])) grads[pl.b[ls[job]] + c] = s;
  } else if (job < 12) {
    const int i = (job - 6) / 3, e = (job - 6) % 3, l = i == 0 ? 0 : 4;
    if (xyz_col(l, e) < 0) return;
    float s = 0.f;
#pragma unroll 16
    for (int k = 0; k < n_cta; ++k)
      s += cta_partials[(size_t)k * kCtaPartialFloats + kPartXyz + (i * kHidden + c) * 3 + e];
    if (i == 1) {
#pragma unroll 16
      for (int k = 0; k < n_corr; ++k) s += corr[(size_t)k * kCorrFloats + 6 * kHidden + c * 3 + e];
    }
    dweff[weff_off(l) + (size_t)c * ref_in(l) + xyz_col(l, e)] = s;
  } else if (job < 14) {
    const int i = job - 12;
    float s = 0.f;
#pragma unroll 16
    for (int b = 0; b < n_seg; ++b) s += seg_sum[((size_t)b * 2 + i) * kHidden + c];
    grads[pl.b[i == 0 ? 0 : 4] + c] = s;
  } else if (job == 14) {
    float s = 0.f;
    for (long long t = c; t < n_tiles; t += blockDim.x)
      s += tile_partials[t * 2 + 1] + (eik_tile ? eik_tile[t * 2 + 1] : 0.f);   // db8 = sum (1 + eps) g
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((c & 31) == 0) sh[c >> 5] = s;
    __syncthreads();
    if (c == 0) {
      float t = 0.f;
      for (int i = 0; i < 16; ++i) t += sh[i];
      grads[pl.b[8]] = t;
    }
  } else if (n_cta_dw8 > 0) {   // dW8 partials of the forward kernel's epilogue (or the v1 backward kernel);
    float s = 0.f;              // 0: the wgrad kernel produced dW8 (autograd entries: g unknown in the forward)
#pragma unroll 16
    for (int k = 0; k < n_cta_dw8; ++k) s += cta_partials[(size_t)k * kCtaPartialFloats + kPartW8 + c];
    dweff[weff_off(8) + c] = s;
  }
}

// ---- effective-weight grads -> parameter grads (weight-norm backward) -------------------------
// one warp per output row (all 9 layers).  plain layers: copy.
//   dg = sum_i dW v / ||v|| ;  dv = (g/||v||) (dW - (sum_i dW v)/||v||^2 v)
__global__ void k_param_grads(const float* __restrict__ dweff, const float* __restrict__ params,
                              ParamLayout pl, const float* __restrict__ scale,
                              const float* __restrict__ inv_norm, float* __restrict__ grads) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw >= kScaleElems) return;
  int l = 0;
  while (l + 1 < kNumLin && gw >= scale_off(l + 1)) ++l;
  const int o = gw - scale_off(l);
  const int in = ref_in(l);
  const float* dw = dweff + weff_off(l) + (size_t)o * in;
  float* gout = grads + pl.w[l] + (size_t)o * in;
  if (pl.g[l] < 0) {
    for (int i = lane; i < in; i += 32) gout[i] = dw[i];
    return;
  }
  const float* v = params + pl.w[l] + (size_t)o * in;
  float dot = 0.f;
  for (int i = lane; i < in; i += 32) dot = fmaf(dw[i], v[i], dot);
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, s);
  const float inv = inv_norm[gw], sc = scale[gw];
  const float k = dot * inv * inv;
  for (int i = lane; i < in; i += 32) gout[i] = sc * (dw[i] - k * v[i]);
  if (lane == 0) grads[pl.g[l] + o] = dot * inv;
}

// ---- loss: loss_out[0] = sum_tiles |.|/N (already scaled?) ; see host for the scaling ---------
// one block, 1024 threads.  loss_out[0] = sdf loss, loss_out[1] = reg loss
__global__ void k_loss(const float* __restrict__ tile_partials, long long n_tiles, float inv_n,
                       const float* __restrict__ seg_norm, int n_seg, float reg_coef,
                       float* __restrict__ loss_out, const float* __restrict__ eik_tile = nullptr,
                       float eik_lambda = 0.f) {
  __shared__ float sh[32], she[32];
  float s = 0.f, se = 0.f;
  for (long long t = threadIdx.x; t < n_tiles; t += blockDim.x) {
    s += tile_partials[t * 2];
    if (eik_tile) se += eik_tile[t * 2];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    se += __shfl_xor_sync(0xffffffffu, se, o);
  }
  if ((threadIdx.x & 31) == 0) { sh[threadIdx.x >> 5] = s; she[threadIdx.x >> 5] = se; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f, te = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { t += sh[i]; te += she[i]; }
    loss_out[0] = t * inv_n;
    loss_out[2] = eik_lambda * te * inv_n;     // train_deep_sdf.py:727 (this rank's rows / N_total)
    float r = 0.f;
    for (int b = 0; b < n_seg; ++b) r += seg_norm[b];
    loss_out[1] = reg_coef * r;   // reg_coef = lambda*ramp*P/N
  }
}

// ---- grad norm (two stages, deterministic) + fused clip/Adam --------------------------------
__global__ void k_sqnorm_partial(const float* __restrict__ g, long long n, float* __restrict__ part) {
  __shared__ float sh[8];
  float s = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    s = fmaf(g[i], g[i], s);
  const float t = block_sum_256(s, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}
// torch.optim.Adam (amsgrad=False, wd=0): algorithmic traffic 28 B/param (r: p,g,m,v  w: p,m,v).
// part != nullptr: clip_grad_norm_ first -- every block sums the n_part squared-norm partials in the same fixed order
// (bit-identical in all blocks), coef = min(1, max_norm / (||g|| + 1e-6)), the gradient is scaled in place as torch does;
// block 0 publishes scal[0] = ||g||_2, scal[1] = coef.
__global__ void k_adam(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                       float* __restrict__ v, long long n, const float* __restrict__ part, int n_part,
                       float max_norm, float* __restrict__ scal, float step_size, float bc2_sqrt, float b1,
                       float b2, float eps) {
  float coef = 1.f;
  if (part) {
    __shared__ float sh[8];
    float s = 0.f;
    for (int i = threadIdx.x; i < n_part; i += blockDim.x) s += part[i];
    const float nrm = sqrtf(block_sum_256(s, sh));          // the same value in every thread of every block
    coef = fminf(1.f, max_norm / (nrm + 1e-6f));
    if (blockIdx.x == 0 && threadIdx.x == 0) { scal[0] = nrm; scal[1] = coef; }
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * coef;
    const float mi = fmaf(1.f - b1, gi - m[i], m[i]);          // lerp, as torch
    const float vi = fmaf(1.f - b2, gi * gi, b2 * v[i]);
    m[i] = mi;
    v[i] = vi;
    p[i] -= step_size * (mi / (sqrtf(vi) / bc2_sqrt + eps));
    if (part) g[i] = gi;   // clip_grad_norm_ scales .grad in place
  }
}

// dOut -> dL/ds through the final tanh (general autograd path): g = dout * (1 - out^2)
__global__ void k_tanh_backward(const float* __restrict__ dout, const float* __restrict__ out,
                                long long n, float* __restrict__ g) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) g[i] = dout[i] * (1.f - out[i] * out[i]);
}

}  // namespace dsdf

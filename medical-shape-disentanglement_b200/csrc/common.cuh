// Shared pieces of the persistent chain kernels (fwd3_kernel.cuh, bwd2_kernel.cuh): launch argument
// structs, tile prologue helpers (xyz / grid coordinates, latent gather, E rows), the epilogue's
// ReLU/pack/store helpers and the per-CTA partial layout of the backward reducers.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

#include "layout.cuh"
#include "philox.cuh"
#include "ptx.cuh"

// per-CTA cycle counters are compiled in only with -DDSDF_DBG_BUILD (tools/gpu_dbg.py builds a
// separate library); the product build carries no instrumentation.
#ifdef DSDF_DBG_BUILD
#define DSDF_DBG_ON(a) ((a).dbg != nullptr)
#else
#define DSDF_DBG_ON(a) false
#endif

namespace dsdf {

enum : int { MODE_DENSE = 0, MODE_TABLE = 1, MODE_GRID = 2 };

// A linear byte stream (weight K-steps, per-scene E steps) seen as a 2-D tensor of 1 KB rows (256 x u32) so that a
// {256, 4} box is one 4 KB CTA slice of a K-step: y = byte offset / 1024 (host encoder: dsdf_api.cu make_stream_map)
constexpr uint32_t kMapRowBytes = 1024;

struct alignas(64) FwdArgs {
  // CTA-pair launches: tensor maps over w2img and scene_e (first members: 64-byte aligned in the parameter space)
  CUtensorMap tmap_w;
  CUtensorMap tmap_e;
  long long n_rows;
  // MODE_DENSE: input [n,259] fp32 (latent | xyz)
  const float* input;
  // MODE_TABLE: latent table [S,256], segment ids [B] (row r -> seg_ids[r / rows_per_seg]),
  //             sdf_data [n,4] (x,y,z,sdf).  MODE_GRID: table = one latent [256].
  const float* table;
  const long long* seg_ids;
  long long rows_per_seg;
  const float* xyzs;       // xyz of row r at xyzs[r * xyz_stride + 0..2] (+3: sdf when stride 4)
  long long xyz_stride;
  // MODE_GRID: create_mesh grid (deep_sdf/mesh.py:38-53), row r <-> linear index grid_lo + r
  int grid_n;
  long long grid_lo;
  float voxel_size;
  // weights
  const uint8_t* w2img;    // forward K-step stream (layout.cuh)
  const uint8_t* scene_e;  // v2 fold mode: per-segment E steps [seg][kSceneESteps][kStepBytes]
  const float* w8;      // [512] effective lin8 weight (fp32)
  const float* b8;      // [1]
  // outputs
  float* out;           // [n] tanh output (un-clamped)
  // TRAIN only
  uint8_t* xstash;      // [tiles][8][128 KB]
  uint32_t* relu_mask;  // [tiles][8 layers][16 chunks][128 rows]: bit (31-i) = column 32*chunk+i > 0
  uint8_t* g_img;       // [tiles][16 KB SUB]: column 0 = bf16 hi part of dL/ds, column 1 = lo part
                        // (the wgrad kernel's "dY" operand for lin8), fused-loss mode only
  float* gout;          // [n] dL/ds (s = pre-tanh lin8 output), fused-loss mode only
  float* tile_partials; // [tiles][2] {sum |clamp(pred)-clamp(gt)|, sum dL/ds}
  int fused_loss;       // 1: compute clamped-L1 loss + gout from sdf_data[:,3]
  float clamp_dist;
  float inv_n_total;
  // v3 fused-loss mode: rows r with (r % rows_per_seg) >= valid_per_seg are padding (zero loss,
  // zero dL/ds); 0 = every row is a sample.  Used by the latent-only optimisation whose
  // num_samples (reconstruct.py:22, 8000 on the CLI) is not a multiple of the 128-row tile.
  long long valid_per_seg;
  // v3 TRAIN + DROP: training-mode dropout after the ReLUs (deep_sdf_decoder.py:103-104), philox.cuh
  DropoutArgs drop;
  // v3 TRAIN + fused loss: dW8 = sum_r g_r X8[r,:] accumulated by the forward epilogue itself (X8 is in the
  // smem A tile, g is known there), one 512-float partial per CTA at dw8_partials[blockIdx.x * dw8_stride];
  // non-null => X8 is not stashed and the wgrad kernel has no lin8 item.
  float* dw8_partials;
  long long dw8_stride;
  // v3 TAN (Eikonal recompute pass, k_forward3<..., TAN>): per-row perturbed input (x', y', z', eps) written by
  // k_backward2<EIK>, and dL/ds of the L1 term (the forward's gout) for dW8 = sum_r g_r Xt8[r,:]
  const float4* eik_row;
  const float* g_tan;
  // diagnostics (nullable): per-CTA cycle counters, see dsdf_debug_read
  unsigned long long* dbg;
};
constexpr int kDbgSlots = 16;
constexpr int kActThreads = 256;   // the 8 epilogue warps

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c,
                                             uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c),
               "r"(d)
               : "memory");
}

// create_mesh grid coordinates, bit-exact with the reference's fp32 torch expressions
// (deep_sdf/mesh.py:45-53; closed form verified in oracle/mesh_grid.py).
__device__ __forceinline__ void grid_xyz(long long idx, int n, float vs, float& x, float& y,
                                         float& z) {
  const float nf = static_cast<float>(n);
  const float f = __ll2float_rn(idx);
  const float a = __fdiv_rn(f, nf);
  const float zi = static_cast<float>(static_cast<int>(idx % n));
  const float yi = fmodf(a, nf);
  const float xi = fmodf(__fdiv_rn(a, nf), nf);
  x = __fadd_rn(__fmul_rn(xi, vs), -1.0f);
  y = __fadd_rn(__fmul_rn(yi, vs), -1.0f);
  z = __fadd_rn(__fmul_rn(zi, vs), -1.0f);
}

template <int MODE>
__device__ __forceinline__ void load_xyz(const FwdArgs& a, long long row, float& x, float& y,
                                         float& z, float& gt) {
  gt = 0.f;
  if (MODE == MODE_DENSE) {
    const float* p = a.input + row * kIn0 + kLatent;
    x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2);
  } else if (MODE == MODE_TABLE) {
    const float* p = a.xyzs + row * a.xyz_stride;
    x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2);
    if (a.xyz_stride >= 4) gt = __ldg(p + 3);
  } else {
    grid_xyz(a.grid_lo + row, a.grid_n, a.voxel_size, x, y, z);
  }
}

// latent -> bf16 into 4 SUBs starting at `dst_base` (256 columns), all 128 rows.
template <int MODE>
__device__ __forceinline__ void gather_latent(const FwdArgs& a, long long row0, uint32_t dst_base,
                                              int t /*0..255*/) {
  const long long last = a.n_rows - 1;
  if (MODE == MODE_DENSE) {
    const int w = t >> 5, lane = t & 31;
    for (int j = 0; j < 16; ++j) {
      const uint32_t r = w * 16 + j;
      long long row = row0 + r;
      row = row < last ? row : last;
      const float* src = a.input + row * kIn0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t c = lane + 32 * k;
        const __nv_bfloat16 b = __float2bfloat16_rn(__ldg(src + c));
        const uint32_t addr = dst_base + (c >> 6) * kSubBytes + sub_off(r, c & 63);
        asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(__bfloat16_as_ushort(b))
                     : "memory");
      }
    }
  } else {
    const uint32_t chunk = t & 31;  // 8 floats
    for (int j = 0; j < 16; ++j) {
      const uint32_t r = (t >> 5) + 8 * j;
      long long row = row0 + r;
      row = row < last ? row : last;
      const float* src;
      if (MODE == MODE_TABLE) {
        const long long sid = a.seg_ids ? __ldg(a.seg_ids + row / a.rows_per_seg) : 0;
        src = a.table + sid * kLatent + chunk * 8;
      } else {
        src = a.table + chunk * 8;
      }
      const float4 v0 = __ldg(reinterpret_cast<const float4*>(src));
      const float4 v1 = __ldg(reinterpret_cast<const float4*>(src) + 1);
      st_shared_v4(dst_base + (chunk >> 3) * kSubBytes + sub_chunk_off(r, chunk & 7),
                   pack_bf16x2(v0.x, v0.y), pack_bf16x2(v0.z, v0.w), pack_bf16x2(v1.x, v1.y),
                   pack_bf16x2(v1.z, v1.w));
    }
  }
}

template <bool FOLD>
__host__ __device__ constexpr int f2_used_ksteps(int l) {
  return FOLD ? (l == 0 ? 0 : (l == kSkipLayer ? 16 : 32)) : w2_ksteps(l);
}
// does layer l contract the "hi" half (SUB 4..7) of the A tile?
template <bool FOLD>
__host__ __device__ constexpr bool f2_in_hi(int l) { return f2_used_ksteps<FOLD>(l) > 16; }

// A-side E row, no-swizzle [128 x 16]: [xh yh zh xl yl zl xh yh | zh 1 1 0 0 0 0 0]
__device__ __forceinline__ void write_e_row_ns(uint32_t e_base, uint32_t r, float x, float y,
                                               float z) {
  const __nv_bfloat16 xh = __float2bfloat16_rn(x), yh = __float2bfloat16_rn(y),
                      zh = __float2bfloat16_rn(z);
  const __nv_bfloat16 xl = __float2bfloat16_rn(x - __bfloat162float(xh));
  const __nv_bfloat16 yl = __float2bfloat16_rn(y - __bfloat162float(yh));
  const __nv_bfloat16 zl = __float2bfloat16_rn(z - __bfloat162float(zh));
  auto u = [](__nv_bfloat16 b) -> uint32_t { return (uint32_t)__bfloat16_as_ushort(b); };
  const uint32_t one = 0x3F80u;
  const uint32_t base = e_base + (r >> 3) * 256u + (r & 7u) * 16u;
  st_shared_v4(base, u(xh) | (u(yh) << 16), u(zh) | (u(xl) << 16), u(yl) | (u(zl) << 16),
               u(xh) | (u(yh) << 16));
  st_shared_v4(base + 128u, u(zh) | (one << 16), one, 0u, 0u);
}

// A-side E row of the Eikonal recompute pass: the perturbed coordinates xyz' split in THREE bf16 terms (the
// perturbation is ~1e-3 of xyz: two terms would resolve it to ~1 %) and the bias multiplier 1 + eps as 1 + bf16(eps):
// [xh yh zh xl yl zl xh yh | zh 1 1 eps xr yr zr 0]  against  [wh wh wh wh wh wh wl wl | wl bh bl bh wh wh wh 0]
__device__ __forceinline__ void write_e_row_tan(uint32_t e_base, uint32_t r, float x, float y, float z, float eps) {
  auto u = [](__nv_bfloat16 b) -> uint32_t { return (uint32_t)__bfloat16_as_ushort(b); };
  auto split3 = [](float v, __nv_bfloat16& h, __nv_bfloat16& l, __nv_bfloat16& rr) {
    h = __float2bfloat16_rn(v);
    const float d1 = v - __bfloat162float(h);
    l = __float2bfloat16_rn(d1);
    rr = __float2bfloat16_rn(d1 - __bfloat162float(l));
  };
  __nv_bfloat16 xh, xl, xr, yh, yl, yr, zh, zl, zr;
  split3(x, xh, xl, xr); split3(y, yh, yl, yr); split3(z, zh, zl, zr);
  const uint32_t one = 0x3F80u;
  const uint32_t base = e_base + (r >> 3) * 256u + (r & 7u) * 16u;
  st_shared_v4(base, u(xh) | (u(yh) << 16), u(zh) | (u(xl) << 16), u(yl) | (u(zl) << 16), u(xh) | (u(yh) << 16));
  st_shared_v4(base + 128u, u(zh) | (one << 16), one | (u(__float2bfloat16_rn(eps)) << 16), u(xr) | (u(yr) << 16), u(zr));
}

// 32 fp32 accumulator columns -> ReLU -> 16 packed bf16x2
__device__ __forceinline__ void relu_pack32(const uint32_t (&v)[32], uint32_t* p) {
#pragma unroll
  for (int i = 0; i < 16; ++i)
    p[i] = pack_relu_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
}
// 16 packed words = columns col..col+31 of row r -> A tile (SWIZZLE_128B SUBs)
__device__ __forceinline__ void store_chunk32(uint32_t a_base, uint32_t r, uint32_t col,
                                              const uint32_t* p) {
  const uint32_t sub = a_base + (col >> 6) * kSubBytes;
  const uint32_t ch0 = (col & 63) >> 3;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    st_shared_v4(sub + sub_chunk_off(r, ch0 + i), p[4 * i], p[4 * i + 1], p[4 * i + 2],
                 p[4 * i + 3]);
}

// per-CTA reducer partials: db for l in {1,2,3,5,6,7} | dW8 | dWxyz for l in {0,4} ([512][3])
constexpr int kPartDb = 0;
constexpr int kPartW8 = 6 * kHidden;
constexpr int kPartXyz = 7 * kHidden;
constexpr int kCtaPartialFloats = 7 * kHidden + 2 * kHidden * 3;  // 6656
__host__ __device__ constexpr int db_slot(int l) { return l < 4 ? l - 1 : l - 2; }  // 1,2,3,5,6,7 -> 0..5

}  // namespace dsdf

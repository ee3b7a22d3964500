// Fused 9-layer DeepSDF forward chain: one persistent CTA per SM, 128 rows per tile.
//
//   warp 0      : weight producer  (bulk async copies global -> smem ring, TMA engine)
//   warp 1      : MMA issuer       (one thread, tcgen05.mma cta_group::1 kind::f16, M=128 N=128 K=16)
//   warp 2      : TMEM allocator (512 columns = the whole 128x512 fp32 accumulator)
//   warp 3      : idle
//   warps 4..11 : activation warps (A-tile gather/convert, TMEM -> ReLU -> bf16 -> smem epilogue,
//                 lin8 dot + tanh (+ clamped-L1 loss and dL/ds) epilogue, activation stash)
//
// Replaces, per 128-row tile, reference networks/deep_sdf_decoder.py:74-109 (9x addmm, 2x cat,
// 8x relu, tanh), deep_sdf/utils.py:86-97 (expand + cat) and, in TRAIN mode,
// train_deep_sdf.py:598-606 (gather, cat, clamp, L1/N).
#pragma once
#include "layout.cuh"
#include "philox.cuh"
#include "ptx.cuh"

namespace dsdf {

enum : int { MODE_DENSE = 0, MODE_TABLE = 1, MODE_GRID = 2 };

struct FwdArgs {
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
  const uint8_t* wimg;  // forward weight image (layout.cuh)
  const uint8_t* w2img;    // v2 K-step stream (layout.cuh)
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
  // diagnostics (nullable): per-CTA cycle counters, see dsdf_debug_read
  unsigned long long* dbg;
};
constexpr int kDbgSlots = 16;

constexpr int kFwdStages = 5;
constexpr int kFwdThreads = 384;
constexpr int kActThreads = 256;
constexpr uint32_t kSmemA = 0;
constexpr uint32_t kSmemE = kTileBytes;                       // 131072
constexpr uint32_t kSmemRing = kSmemE + kSubBytes;            // 147456
constexpr uint32_t kSmemAux = kSmemRing + kFwdStages * kSubBytes;  // 229376
// aux: barriers (16 x 8 B) | tmem slot (16 B) | w8 fp32 (2048 B) | partial dot (512 B) | red (64 B)
constexpr uint32_t kAuxBars = 0;
constexpr uint32_t kAuxTmem = 128;
constexpr uint32_t kAuxW8 = 144;
constexpr uint32_t kAuxPart = kAuxW8 + 2048;
constexpr uint32_t kAuxRed = kAuxPart + 512;
constexpr uint32_t kFwdSmemBytes = kSmemAux + kAuxRed + 64;   // 232144 <= 232448

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

// E tile (A side): [xh yh zh xl yl zl xh yh | zh 1 1 0 0 0 0 0], one thread per row.
__device__ __forceinline__ void write_e_row(uint32_t e_base, uint32_t r, float x, float y,
                                            float z) {
  const __nv_bfloat16 xh = __float2bfloat16_rn(x), yh = __float2bfloat16_rn(y),
                      zh = __float2bfloat16_rn(z);
  const __nv_bfloat16 xl = __float2bfloat16_rn(x - __bfloat162float(xh));
  const __nv_bfloat16 yl = __float2bfloat16_rn(y - __bfloat162float(yh));
  const __nv_bfloat16 zl = __float2bfloat16_rn(z - __bfloat162float(zh));
  auto u = [](__nv_bfloat16 b) -> uint32_t { return (uint32_t)__bfloat16_as_ushort(b); };
  const uint32_t one = 0x3F80u;
  st_shared_v4(e_base + sub_chunk_off(r, 0), u(xh) | (u(yh) << 16), u(zh) | (u(xl) << 16),
               u(yl) | (u(zl) << 16), u(xh) | (u(yh) << 16));
  st_shared_v4(e_base + sub_chunk_off(r, 1), u(zh) | (one << 16), one, 0u, 0u);
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

template <int MODE, bool TRAIN>
__global__ void __launch_bounds__(kFwdThreads, 1) k_forward(const FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemAux + kAuxBars);
  uint64_t* full = bars;                  // [kFwdStages]
  uint64_t* empty = bars + kFwdStages;    // [kFwdStages]
  uint64_t* a_ready = bars + 2 * kFwdStages;
  uint64_t* mma_done = a_ready + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kSmemAux + kAuxTmem);
  float* w8s = reinterpret_cast<float*>(smem + kSmemAux + kAuxW8);
  float* part = reinterpret_cast<float*>(smem + kSmemAux + kAuxPart);
  float* red = reinterpret_cast<float*>(smem + kSmemAux + kAuxRed);

  if ((s_base & 1023u) != 0) __trap();

  if (threadIdx.x == 0) {
    for (int i = 0; i < kFwdStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(a_ready, kActThreads);
    mbar_init(mma_done, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  for (int i = threadIdx.x; i < kHidden; i += blockDim.x) w8s[i] = a.w8[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const long long n_tiles = (a.n_rows + kTileM - 1) / kTileM;

  if (warp == 0) {
    // ===================== weight producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int t = 0; t < kFwdTilesTotal; ++t) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], kSubBytes);
          bulk_g2s(smem + kSmemRing + stage * kSubBytes, a.wimg + (size_t)t * kSubBytes,
                   kSubBytes, &full[stage]);
          if (++stage == kFwdStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 128, false, false);
      uint32_t stage = 0, phase = 0, par_a = 0;
      for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int l = 0; l < 8; ++l) {
          mbar_wait(a_ready, par_a);
          par_a ^= 1;
          tc_fence_after();
          const int KB = fwd_kb(l), NC = fwd_nc(l);
          for (int kb = 0; kb < KB; ++kb) {
            for (int nc = 0; nc < NC; ++nc) {
              mbar_wait(&full[stage], phase);
              tc_fence_after();
              const uint32_t a_addr = s_base + kSmemA + kb * kSubBytes;
              const uint32_t b_addr = s_base + kSmemRing + stage * kSubBytes;
#pragma unroll
              for (int ks = 0; ks < 4; ++ks) {
                umma_bf16(tmem + nc * 128, umma_smem_desc(a_addr + ks * 32, 16, 1024),
                          umma_smem_desc(b_addr + ks * 32, 16, 1024), idesc,
                          (kb > 0 || ks > 0) ? 1u : 0u);
              }
              umma_commit(&empty[stage]);
              if (++stage == kFwdStages) { stage = 0; phase ^= 1; }
            }
          }
          // E step: xyz (hi/lo) + bias through the ones columns
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          for (int nc = 0; nc < NC; ++nc) {
            umma_bf16(tmem + nc * 128, umma_smem_desc(s_base + kSmemE, 16, 1024),
                      umma_smem_desc(s_base + kSmemRing + stage * kSubBytes + nc * 32, 16, 1024),
                      idesc, 1u);
          }
          umma_commit(&empty[stage]);
          if (++stage == kFwdStages) { stage = 0; phase ^= 1; }
          umma_commit(mma_done);
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== activation warps =====================
    const int t = threadIdx.x - 128;      // 0..255
    const int q = warp & 3;               // TMEM lane quarter this warp may access
    const int h = (warp - 4) >> 2;        // column half
    const uint32_t r = q * 32 + lane;     // row inside the tile
    const uint32_t tlane = (uint32_t)(q * 32) << 16;
    uint32_t par_md = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const long long row0 = tile * kTileM;
      if (TRAIN) {
        if (t == 0) bulk_wait_read_all();
        named_bar_sync(1, kActThreads);
      }
      // ---- layer-0 A tile: latent -> SUB 0..3, xyz -> E
      gather_latent<MODE>(a, row0, s_base + kSmemA, t);
      float gt = 0.f;
      if (t < kTileM) {
        long long row = row0 + t;
        row = row < a.n_rows ? row : a.n_rows - 1;
        float x, y, z;
        load_xyz<MODE>(a, row, x, y, z, gt);
        write_e_row(s_base + kSmemE, t, x, y, z);
      }
      fence_proxy_async_smem();
      mbar_arrive(a_ready);

      for (int l = 0; l < 7; ++l) {
        mbar_wait(mma_done, par_md);
        par_md ^= 1;
        tc_fence_after();
        if (TRAIN) {
          if (t == 0) bulk_wait_read_all();
          named_bar_sync(1, kActThreads);
        }
        const int ncols = fwd_nc(l) * 128;
        if (h * 256 < ncols) {
#pragma unroll 1
          for (int j = 0; j < 8; ++j) {
            const uint32_t col = h * 256 + j * 32;
            uint32_t v[32];
            tmem_ld_32x32(tmem + tlane + col, v);
            tmem_ld_wait();
            const uint32_t sub = s_base + kSmemA + (col >> 6) * kSubBytes;
            const uint32_t ch0 = (col & 63) >> 3;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              st_shared_v4(sub + sub_chunk_off(r, ch0 + i),
                           pack_relu_bf16x2(__uint_as_float(v[8 * i + 0]), __uint_as_float(v[8 * i + 1])),
                           pack_relu_bf16x2(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])),
                           pack_relu_bf16x2(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])),
                           pack_relu_bf16x2(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7])));
            }
          }
        }
        tc_fence_before();
        fence_proxy_async_smem();
        if (TRAIN) {
          // stash X_{l+1} (input of lin(l+1)); for l == 3 only the x3 part (SUB 0..3)
          named_bar_sync(1, kActThreads);
          if (t == 0) {
            const uint32_t bytes = (l == 3) ? 4 * kSubBytes : kTileBytes;
            bulk_s2g(a.xstash + (size_t)tile * kStashTileBytes + (size_t)l * kTileBytes,
                     smem + kSmemA, bytes);
            bulk_commit();
          }
        }
        if (l == 3) {
          // skip connection: latent -> SUB 4..7 (reference cat order [x3 | latent | xyz])
          gather_latent<MODE>(a, row0, s_base + kSmemA + 4 * kSubBytes, t);
          fence_proxy_async_smem();
        }
        mbar_arrive(a_ready);
      }

      // ---- lin7 epilogue fused with lin8 (512 -> 1), tanh, loss
      mbar_wait(mma_done, par_md);
      par_md ^= 1;
      tc_fence_after();
      if (TRAIN) {
        if (t == 0) bulk_wait_read_all();
        named_bar_sync(1, kActThreads);
      }
      float acc = 0.f;
#pragma unroll 1
      for (int j = 0; j < 8; ++j) {
        const uint32_t col = h * 256 + j * 32;
        uint32_t v[32];
        tmem_ld_32x32(tmem + tlane + col, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float x = fmaxf(__uint_as_float(v[i]), 0.f);
          acc = fmaf(x, w8s[col + i], acc);
        }
        if (TRAIN) {
          const uint32_t sub = s_base + kSmemA + (col >> 6) * kSubBytes;
          const uint32_t ch0 = (col & 63) >> 3;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            st_shared_v4(sub + sub_chunk_off(r, ch0 + i),
                         pack_relu_bf16x2(__uint_as_float(v[8 * i + 0]), __uint_as_float(v[8 * i + 1])),
                         pack_relu_bf16x2(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])),
                         pack_relu_bf16x2(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])),
                         pack_relu_bf16x2(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7])));
          }
        }
      }
      tc_fence_before();
      if (h == 1) part[r] = acc;
      if (TRAIN) fence_proxy_async_smem();
      named_bar_sync(1, kActThreads);
      if (TRAIN && t == 0) {
        bulk_s2g(a.xstash + (size_t)tile * kStashTileBytes + (size_t)7 * kTileBytes,
                 smem + kSmemA, kTileBytes);
        bulk_commit();
      }
      if (h == 0) {
        const long long row = row0 + r;
        const bool valid = row < a.n_rows;
        const float s = acc + part[r] + __ldg(a.b8);
        const float th = tanhf(s);
        if (valid) a.out[row] = th;
        if (TRAIN && a.fused_loss) {
          // train_deep_sdf.py:565-567,604-606: clamp both, L1(sum)/N_total.
          const float d = a.clamp_dist;
          const float pc = fminf(fmaxf(th, -d), d);
          const float gc = fminf(fmaxf(gt, -d), d);
          const float diff = pc - gc;
          float lossr = valid ? fabsf(diff) : 0.f;
          const float sgn = (diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f);
          const float inside = (th >= -d && th <= d) ? 1.f : 0.f;
          float g = valid ? sgn * inside * (1.f - th * th) * a.inv_n_total : 0.f;
          if (valid) a.gout[row] = g;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            lossr += __shfl_xor_sync(0xffffffffu, lossr, o);
            g += __shfl_xor_sync(0xffffffffu, g, o);
          }
          if (lane == 0) { red[q * 2] = lossr; red[q * 2 + 1] = g; }
        }
      }
      if (TRAIN && a.fused_loss) {
        named_bar_sync(1, kActThreads);
        if (t == 0) {
          a.tile_partials[tile * 2] = (red[0] + red[2]) + (red[4] + red[6]);
          a.tile_partials[tile * 2 + 1] = (red[1] + red[3]) + (red[5] + red[7]);
        }
      }
    }
    if (TRAIN && t == 0) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace dsdf

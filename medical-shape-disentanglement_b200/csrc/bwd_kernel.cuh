// Fused backward dgrad chain (scene mode): one persistent CTA per SM, 128 rows per tile.
//
//   warp 0       : weight producer (W^T image tiles, bulk async copies -> smem ring)
//   warp 1       : MMA issuer      (dX_l = dY_l * W_l, tcgen05.mma M=128 N=128 K=16)
//   warp 2       : TMEM allocator; warp 3 idle
//   warps 4..11  : activation warps: dY7 from g (lin8 + tanh' + clamp/L1 folded into g by the
//                  forward), TMEM -> ReLU mask (from the stashed activations) -> bf16 -> smem,
//                  dY stash for the wgrad kernel
//   warps 12..15 : reducer warps (run in the shadow of the MMAs): column sums of every dY tile
//                  -> bias grads, per-tile colsums of dY0/dY4 (-> latent grads and the latent
//                  columns of dW0/dW4 by the segmented kernels), xyz columns of dW0/dW4, dW8.
//
// Replaces the autograd backward of reference train_deep_sdf.py:731 for the decoder
// (9 dgrad GEMMs, ReLU masks, tanh', bias grads).  In scene mode the latent/xyz columns of
// dX0 and dX4 are never materialised: sum_rows dX[:, latent] = colsum(dY) * W[:, latent].
#pragma once
#include "layout.cuh"
#include "ptx.cuh"

namespace dsdf {

struct BwdArgs {
  long long n_rows;
  const uint8_t* dgimg;     // dgrad weight image
  const float* w8;          // [512]
  const float* g;           // [n] dL/ds
  const uint8_t* xstash;    // forward activation stash
  uint8_t* dystash;         // out: dY_l images, slot l (1..7)
  const float* xyz;         // xyz of row r at xyz[r * xyz_stride + 0..2]
  long long xyz_stride;
  float* tile_colsum;       // [tiles][2][512]: colsum(dY0), colsum(dY4)
  float* cta_partials;      // [grid][kCtaPartialFloats]
};

// per-CTA reducer partials: db for l in {1,2,3,5,6,7} | dW8 | dWxyz for l in {0,4} ([512][3])
constexpr int kPartDb = 0;
constexpr int kPartW8 = 6 * kHidden;
constexpr int kPartXyz = 7 * kHidden;
constexpr int kCtaPartialFloats = 7 * kHidden + 2 * kHidden * 3;  // 6656
__host__ __device__ constexpr int db_slot(int l) { return l < 4 ? l - 1 : l - 2; }  // 1,2,3,5,6,7 -> 0..5

constexpr int kBwdStages = 5;
constexpr int kBwdThreads = 512;
constexpr uint32_t kBSmemA = 0;
constexpr uint32_t kBSmemRing = kTileBytes;
constexpr uint32_t kBSmemAux = kBSmemRing + kBwdStages * kSubBytes;  // 212992
constexpr uint32_t kBAuxBars = 0;      // 16 x 8
constexpr uint32_t kBAuxTmem = 128;
constexpr uint32_t kBAuxW8 = 144;      // 2048
constexpr uint32_t kBAuxG = kBAuxW8 + 2048;        // 512
constexpr uint32_t kBAuxXyz = kBAuxG + 512;        // 128*3*4 = 1536
constexpr uint32_t kBwdSmemBytes = kBSmemAux + kBAuxXyz + 1536;

__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// zero the bf16 halves of d whose counterpart in x (post-ReLU activation, >= 0) is zero
__device__ __forceinline__ uint32_t relu_mask2(uint32_t d, uint32_t x) {
  __nv_bfloat162 xv;
  *reinterpret_cast<uint32_t*>(&xv) = x;
  const unsigned m = __hgt2_mask(xv, __float2bfloat162_rn(0.f));   // 0xFFFF per half where x > 0
  return d & m;
}
__device__ __forceinline__ void st_shared_v4b(uint32_t addr, uint32_t a, uint32_t b, uint32_t c,
                                              uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c),
               "r"(d)
               : "memory");
}

__global__ void __launch_bounds__(kBwdThreads, 1) k_backward(const BwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kBSmemAux + kBAuxBars);
  uint64_t* full = bars;
  uint64_t* empty = bars + kBwdStages;
  uint64_t* a_ready = bars + 2 * kBwdStages;
  uint64_t* mma_done = a_ready + 1;
  uint64_t* dy_ready = a_ready + 2;
  uint64_t* red_done = a_ready + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kBSmemAux + kBAuxTmem);
  float* w8s = reinterpret_cast<float*>(smem + kBSmemAux + kBAuxW8);
  float* gs = reinterpret_cast<float*>(smem + kBSmemAux + kBAuxG);
  float* xyzs = reinterpret_cast<float*>(smem + kBSmemAux + kBAuxXyz);

  if ((s_base & 1023u) != 0) __trap();
  if (threadIdx.x == 0) {
    for (int i = 0; i < kBwdStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(a_ready, 256);
    mbar_init(mma_done, 1);
    mbar_init(dy_ready, 256);
    mbar_init(red_done, 128);
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
        for (int l = 7; l >= 1; --l) {
          const int nt = dg_tiles(l);
          const uint8_t* src = a.dgimg + (size_t)dg_tile_base(l) * kSubBytes;
          for (int t = 0; t < nt; ++t) {
            mbar_wait(&empty[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full[stage], kSubBytes);
            bulk_g2s(smem + kBSmemRing + stage * kSubBytes, src + (size_t)t * kSubBytes,
                     kSubBytes, &full[stage]);
            if (++stage == kBwdStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 128, false, false);
      uint32_t stage = 0, phase = 0, par_a = 0;
      for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int l = 7; l >= 1; --l) {
          mbar_wait(a_ready, par_a);
          par_a ^= 1;
          tc_fence_after();
          const int KB = dg_kb(l), NC = dg_nc(l);
          for (int kb = 0; kb < KB; ++kb) {
            for (int nc = 0; nc < NC; ++nc) {
              mbar_wait(&full[stage], phase);
              tc_fence_after();
              const uint32_t a_addr = s_base + kBSmemA + kb * kSubBytes;
              const uint32_t b_addr = s_base + kBSmemRing + stage * kSubBytes;
#pragma unroll
              for (int ks = 0; ks < 4; ++ks) {
                umma_bf16(tmem + nc * 128, umma_smem_desc(a_addr + ks * 32, 16, 1024),
                          umma_smem_desc(b_addr + ks * 32, 16, 1024), idesc,
                          (kb > 0 || ks > 0) ? 1u : 0u);
              }
              umma_commit(&empty[stage]);
              if (++stage == kBwdStages) { stage = 0; phase ^= 1; }
            }
          }
          umma_commit(mma_done);
        }
      }
    }
  } else if (warp >= 4 && warp < 12) {
    // ===================== activation warps =====================
    const int t = threadIdx.x - 128;
    const int q = warp & 3;
    const int h = (warp - 4) >> 2;
    const uint32_t r = q * 32 + lane;
    const uint32_t tlane = (uint32_t)(q * 32) << 16;
    uint32_t par_md = 0, par_rd = 0;
    bool first = true;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const long long row0 = tile * kTileM;
      const long long row = row0 + r;
      const bool valid = row < a.n_rows;
      const uint8_t* xs_tile = a.xstash + (size_t)tile * kStashTileBytes;
      uint8_t* dy_tile = a.dystash + (size_t)tile * kStashTileBytes;
      // previous tile: reducers done with dY0, stash stores done reading A
      if (!first) {
        mbar_wait(red_done, par_rd);
        par_rd ^= 1;
      }
      first = false;
      if (t == 0) bulk_wait_read_all();
      named_bar_sync(1, 256);
      const float gr = valid ? __ldg(a.g + row) : 0.f;
      if (h == 0) {
        gs[r] = gr;
        const long long rr = valid ? row : a.n_rows - 1;
        xyzs[r * 3 + 0] = __ldg(a.xyz + rr * a.xyz_stride + 0);
        xyzs[r * 3 + 1] = __ldg(a.xyz + rr * a.xyz_stride + 1);
        xyzs[r * 3 + 2] = __ldg(a.xyz + rr * a.xyz_stride + 2);
      }
      // ---- dY7[r,c] = g[r] * w8[c] * (X8[r,c] > 0)
      {
        const uint8_t* x8 = xs_tile + (size_t)7 * kTileBytes;
#pragma unroll 1
        for (int j = 0; j < 8; ++j) {
          const uint32_t col = h * 256 + j * 32;
          const uint32_t sub_o = (col >> 6) * kSubBytes;
          const uint32_t ch0 = (col & 63) >> 3;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t off = sub_o + sub_chunk_off(r, ch0 + i);
            const uint4 x = ldg_nc_v4(x8 + off);
            const float* w = w8s + col + 8 * i;
            const uint32_t d0 = relu_mask2(pack_bf16x2(gr * w[0], gr * w[1]), x.x);
            const uint32_t d1 = relu_mask2(pack_bf16x2(gr * w[2], gr * w[3]), x.y);
            const uint32_t d2 = relu_mask2(pack_bf16x2(gr * w[4], gr * w[5]), x.z);
            const uint32_t d3 = relu_mask2(pack_bf16x2(gr * w[6], gr * w[7]), x.w);
            st_shared_v4b(s_base + kBSmemA + off, d0, d1, d2, d3);
          }
        }
      }
      fence_proxy_async_smem();
      named_bar_sync(1, 256);
      if (t == 0) {
        bulk_s2g(dy_tile + (size_t)7 * kTileBytes, smem + kBSmemA, kTileBytes);
        bulk_commit();
      }
      mbar_arrive(a_ready);
      mbar_arrive(dy_ready);

      for (int l = 7; l >= 1; --l) {
        mbar_wait(mma_done, par_md);
        par_md ^= 1;
        tc_fence_after();
        mbar_wait(red_done, par_rd);   // reducers finished reading dY_l from A
        par_rd ^= 1;
        if (t == 0) bulk_wait_read_all();
        named_bar_sync(1, 256);
        const int ncols = dg_nc(l) * 128;
        const uint8_t* xl = xs_tile + (size_t)(l - 1) * kTileBytes;   // X_l = input of lin l
        if (h * 256 < ncols) {
#pragma unroll 1
          for (int j = 0; j < 8; ++j) {
            const uint32_t col = h * 256 + j * 32;
            uint32_t v[32];
            tmem_ld_32x32(tmem + tlane + col, v);
            const uint32_t sub_o = (col >> 6) * kSubBytes;
            const uint32_t ch0 = (col & 63) >> 3;
            uint4 x[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) x[i] = ldg_nc_v4(xl + sub_o + sub_chunk_off(r, ch0 + i));
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t d0 = relu_mask2(pack_bf16x2(__uint_as_float(v[8 * i + 0]), __uint_as_float(v[8 * i + 1])), x[i].x);
              const uint32_t d1 = relu_mask2(pack_bf16x2(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])), x[i].y);
              const uint32_t d2 = relu_mask2(pack_bf16x2(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])), x[i].z);
              const uint32_t d3 = relu_mask2(pack_bf16x2(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7])), x[i].w);
              st_shared_v4b(s_base + kBSmemA + sub_o + sub_chunk_off(r, ch0 + i), d0, d1, d2, d3);
            }
          }
        }
        tc_fence_before();
        fence_proxy_async_smem();
        named_bar_sync(1, 256);
        if (l - 1 >= 1) {
          if (t == 0) {
            bulk_s2g(dy_tile + (size_t)(l - 1) * kTileBytes, smem + kBSmemA,
                     (uint32_t)(ncols / 64) * kSubBytes);
            bulk_commit();
          }
          mbar_arrive(a_ready);
        }
        mbar_arrive(dy_ready);
      }
    }
    if (t == 0) bulk_wait_all();
  } else if (warp >= 12) {
    // ===================== reducer warps =====================
    const int rt = threadIdx.x - 384;      // 0..127, columns 4rt..4rt+3
    const uint32_t sub_o = (rt >> 4) * kSubBytes;
    const uint32_t chk = (rt & 15) >> 1;
    const uint32_t hb = (rt & 1) * 8;
    float db[6][4];
    float dw8[4];
    float dwx[2][4][3];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k) db[i][k] = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) dw8[k] = 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 3; ++e) dwx[i][k][e] = 0.f;
    uint32_t par_dy = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const uint8_t* x8 = a.xstash + (size_t)tile * kStashTileBytes + (size_t)7 * kTileBytes;
#pragma unroll
      for (int l = 7; l >= 0; --l) {
        mbar_wait(dy_ready, par_dy);
        par_dy ^= 1;
        const bool active = (l != 3) || (rt < 64);   // dY3 has 256 columns
        float s[4] = {0.f, 0.f, 0.f, 0.f};
        if (active) {
          if (l == 7) {
#pragma unroll 4
            for (int r = 0; r < kTileM; ++r) {
              const uint32_t off = sub_o + sub_chunk_off(r, chk) + hb;
              uint32_t w0, w1;
              asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(s_base + kBSmemA + off));
              s[0] += bf16_lo(w0); s[1] += bf16_hi(w0); s[2] += bf16_lo(w1); s[3] += bf16_hi(w1);
              const uint2 x = __ldg(reinterpret_cast<const uint2*>(x8 + off));
              const float gr = gs[r];
              dw8[0] = fmaf(gr, bf16_lo(x.x), dw8[0]);
              dw8[1] = fmaf(gr, bf16_hi(x.x), dw8[1]);
              dw8[2] = fmaf(gr, bf16_lo(x.y), dw8[2]);
              dw8[3] = fmaf(gr, bf16_hi(x.y), dw8[3]);
            }
          } else if (l == 4 || l == 0) {
            const int xi = (l == 0) ? 0 : 1;
#pragma unroll 4
            for (int r = 0; r < kTileM; ++r) {
              const uint32_t off = sub_o + sub_chunk_off(r, chk) + hb;
              uint32_t w0, w1;
              asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(s_base + kBSmemA + off));
              const float d0 = bf16_lo(w0), d1 = bf16_hi(w0), d2 = bf16_lo(w1), d3 = bf16_hi(w1);
              s[0] += d0; s[1] += d1; s[2] += d2; s[3] += d3;
              const float x = xyzs[r * 3], y = xyzs[r * 3 + 1], z = xyzs[r * 3 + 2];
              dwx[xi][0][0] = fmaf(d0, x, dwx[xi][0][0]); dwx[xi][0][1] = fmaf(d0, y, dwx[xi][0][1]); dwx[xi][0][2] = fmaf(d0, z, dwx[xi][0][2]);
              dwx[xi][1][0] = fmaf(d1, x, dwx[xi][1][0]); dwx[xi][1][1] = fmaf(d1, y, dwx[xi][1][1]); dwx[xi][1][2] = fmaf(d1, z, dwx[xi][1][2]);
              dwx[xi][2][0] = fmaf(d2, x, dwx[xi][2][0]); dwx[xi][2][1] = fmaf(d2, y, dwx[xi][2][1]); dwx[xi][2][2] = fmaf(d2, z, dwx[xi][2][2]);
              dwx[xi][3][0] = fmaf(d3, x, dwx[xi][3][0]); dwx[xi][3][1] = fmaf(d3, y, dwx[xi][3][1]); dwx[xi][3][2] = fmaf(d3, z, dwx[xi][3][2]);
            }
            float4 o = make_float4(s[0], s[1], s[2], s[3]);
            *reinterpret_cast<float4*>(a.tile_colsum + ((size_t)tile * 2 + xi) * kHidden + 4 * rt) = o;
          } else {
#pragma unroll 8
            for (int r = 0; r < kTileM; ++r) {
              const uint32_t off = sub_o + sub_chunk_off(r, chk) + hb;
              uint32_t w0, w1;
              asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(s_base + kBSmemA + off));
              s[0] += bf16_lo(w0); s[1] += bf16_hi(w0); s[2] += bf16_lo(w1); s[3] += bf16_hi(w1);
            }
          }
          if (l != 0 && l != 4) {
            const int sl = db_slot(l);
#pragma unroll
            for (int k = 0; k < 4; ++k) db[sl][k] += s[k];
          }
        }
        mbar_arrive(red_done);
      }
    }
    float* out = a.cta_partials + (size_t)blockIdx.x * kCtaPartialFloats;
#pragma unroll
    for (int i = 0; i < 6; ++i)
      *reinterpret_cast<float4*>(out + kPartDb + i * kHidden + 4 * rt) =
          make_float4(db[i][0], db[i][1], db[i][2], db[i][3]);
    *reinterpret_cast<float4*>(out + kPartW8 + 4 * rt) = make_float4(dw8[0], dw8[1], dw8[2], dw8[3]);
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 3; ++e)
          out[kPartXyz + (i * kHidden + 4 * rt + k) * 3 + e] = dwx[i][k][e];
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace dsdf

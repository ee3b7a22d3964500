// Weight-gradient GEMMs dW_l = dY_l^T X_l, v2: CTA pairs (tcgen05 cta_group::2, M = 256 output
// rows per pair), lean issue loop, cost-balanced static schedule.
//
// Both operands are consumed MN-major straight from the SWIZZLE_128B activation images the forward /
// backward kernels stashed (rows = K), so no transpose pass exists anywhere.
//
//   work unit ("segment") = (layer l, 256-wide out chunk ocp, tile range): D[256 o x N i] over the
//   pair's two TMEMs (N = 512, or 256 for lin4's x3 part), accumulated over the tile range and
//   written as one fp32 partial; a deterministic kernel sums an item's partials in order.
//   The host cuts the (item, tile) space into equal-cost contiguous ranges, one per CTA pair.
//
//   warp 0 producer: per 32-row K-chunk each CTA copies ITS 128 o columns of dY (8 KB) and ITS half
//          of X's columns (16 KB; 8 KB for lin4) -> 24 KB stage, ring of 8
//   warp 1 MMA issuer (leader; M=256, N=256, K=16, A and B MN-major) / full-barrier relay (peer)
//   warp 2 TMEM alloc; warps 4..7 epilogue (TMEM -> fp32 partial in HBM)
//
// Layer 8 (dW8 = sum_r g_r X8[r,:]) rides along as an item whose "dY" is the forward kernel's
// g image (column 0 = bf16 hi part of g, column 1 = lo part).
//
// Replaces the 9 wgrad `mm` calls of the reference's autograd backward
// (train_deep_sdf.py:731) for lin1..lin8; lin0, the latent/xyz columns of lin4 and all bias grads
// come from the backward kernel's reducer warps and the segmented kernels.
#pragma once
#include "common.cuh"
#include "fwd3_kernel.cuh"   // mask_pack32
#include "layout.cuh"
#include "ptx.cuh"

namespace dsdf {

struct WgSeg {
  int layer;       // 1..8
  int ocp;         // out-chunk pair (256 rows of dW)
  int n_sub;       // N / 64 (8, or 4 for lin4's x3 part)
  int slot;        // partial slot
  long long tile_lo, tile_hi;
};

struct Wg2Args {
  const WgSeg* segs;
  const int* seg_begin;      // [n_pairs + 1]
  const uint8_t* xstash;
  const uint8_t* dystash;
  const uint8_t* g_img;      // [tiles][16 KB]: layer-8 "dY"
  float* partials;           // [slots][256][512]
  // REBUILD: dY7 = g w8^T . mask7 is rebuilt in shared memory by the (otherwise idle) epilogue warps from 512 B of g,
  // 512 B of w8 and 2 KB of ReLU bits per tile and CTA, instead of being stashed by the dgrad chain and read back
  // (1 KB/row written + 1 KB/row read less); byte-identical to what k_backward2 would have stashed
  const float* g;            // [n] dL/ds
  const float* w8;           // [512]
  const uint32_t* relu_mask; // forward ReLU bit-masks [tiles][8][16][128]
  int pair_base;             // schedule index of this launch's first CTA pair
};

constexpr int kW2Ring = 8;
constexpr int kW2Threads = 256;
constexpr uint32_t kW2Slice = 4096;                        // 32 rows of one SUB
constexpr uint32_t kW2StageBytes = 6 * kW2Slice;           // 2 A slices + 4 B slices
constexpr uint32_t kW2SmemAux = kW2Ring * kW2StageBytes;   // 196608
constexpr uint32_t kW2SmemBytes = kW2SmemAux + 512;

// QUAD: clusters of 4 CTAs = the two pairs of a quad (ranks 0,1: out chunk 0; ranks 2,3: out chunk 1) walking the same
// tile range.  The X slices both pairs need (ranks r and r ^ 2 hold the same half of X's columns) are loaded ONCE:
// each of the two CTAs fetches half of them and multicasts to both, so an X tile leaves HBM and the L2 exactly once;
// a stage may only be overwritten when BOTH pairs have consumed it, so both issuers commit every stage's `empty`
// barrier to all four CTAs (count 2).  lin3 (one out chunk) is split by X columns instead: both pairs read the same
// dY3 and their own half of X3 (no multicast) and write the two column halves of one partial slot.
template <bool REBUILD, bool QUAD = false>
__global__ void __launch_bounds__(kW2Threads, 1) k_wgrad2(const Wg2Args a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kW2SmemAux);
  uint64_t* full = bars;                  // [8]
  uint64_t* empty = bars + 8;             // [8]
  uint64_t* d_done = bars + 16;           // MMA -> epilogue: segment accumulated
  uint64_t* d_free = bars + 17;           // epilogue -> MMA: TMEM drained (both CTAs)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kW2SmemAux + 256);

  if ((s_base & 1023u) != 0) __trap();
  const uint32_t rank_all = cluster_ctarank();
  const uint32_t crank = QUAD ? (rank_all & 1u) : rank_all;      // rank inside the CTA pair
  const uint32_t pr = QUAD ? (rank_all >> 1) : 0u;               // which pair of the quad
  const uint32_t lead_rank = 2u * pr;                            // cluster rank of this pair's leader
  const bool leader = crank == 0;
  if (threadIdx.x == 0) {
    // full: the producer's arrive.expect_tx (+ the relayed arrive of the peer in the leader) and, REBUILD, one more
    // arrival per stage: the dY7 generator warps for a lin7 stage, the producer itself otherwise
    for (int i = 0; i < kW2Ring; ++i) {
      mbar_init(&full[i], (leader ? 2 : 1) + (REBUILD ? 1 : 0));
      mbar_init(&empty[i], QUAD ? 2 : 1);
    }
    mbar_init(d_done, 1);
    mbar_init(d_free, 8);                 // 4 epilogue warps of each CTA
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc2<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t full_c = mapa_u32(smem_u32(full), lead_rank);
  const uint32_t d_free_c = mapa_u32(smem_u32(d_free), lead_rank);
  const uint16_t pair_mask = (uint16_t)(3u << (2u * pr));        // this pair's two CTAs
  const uint16_t stage_mask = QUAD ? (uint16_t)0xF : (uint16_t)3;

  const int pair = a.pair_base + (int)(blockIdx.x >> 1);
  grid_dep_launch();   // every CTA of this launch is resident: a programmatically dependent launch may start beside it
  const int seg_lo = a.seg_begin[pair], seg_hi = a.seg_begin[pair + 1];

  if (warp == 0) {
    // ===================== producer =====================
    if (elect_one()) {
      uint32_t ph = 0, par = 0;           // par: which half of the ring this tile uses
      for (int si = seg_lo; si < seg_hi; ++si) {
        const WgSeg sg = a.segs[si];
        const int nb = sg.n_sub / 4;      // B slices per CTA per N half... (2 per half)
        const bool gen = REBUILD && sg.layer == 7;          // the A slices come from the generator warps
        const uint32_t tx = ((gen ? 0 : 2) + 2 * nb) * kW2Slice;
        for (long long tile = sg.tile_lo; tile < sg.tile_hi; ++tile) {
          const uint8_t* dy;
          uint32_t dy_sub2;               // offset of the second M atom
          if (sg.layer == 8) {
            dy = a.g_img + (size_t)tile * kSubBytes;
            dy_sub2 = 0;
          } else {
            dy = a.dystash + (size_t)tile * kStashTileBytes + (size_t)sg.layer * kTileBytes +
                 (size_t)(4 * sg.ocp + 2 * crank) * kSubBytes;
            dy_sub2 = kSubBytes;
          }
          // QUAD lin3: pair pr contracts X columns 256 pr .. (SUBs 4 pr ..)
          const uint8_t* x = a.xstash + (size_t)tile * kStashTileBytes + (size_t)(sg.layer - 1) * kTileBytes +
                             (size_t)(2 * crank + ((QUAD && sg.layer == 3) ? 4 * pr : 0)) * kSubBytes;
          const bool share_x = QUAD && sg.layer != 3;         // the partner pair needs the same X slices
          const uint16_t x_mask = (uint16_t)((1u << crank) | (1u << (crank + 2)));
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const uint32_t st = par * 4 + u;
            mbar_wait(&empty[st], ph ^ 1);
            mbar_arrive_expect_tx(&full[st], tx);
            uint8_t* dst = smem + st * kW2StageBytes;
            if (!gen) {
              if (REBUILD) mbar_arrive(&full[st]);
              bulk_g2s(dst, dy + u * kW2Slice, kW2Slice, &full[st]);
              bulk_g2s(dst + kW2Slice, dy + dy_sub2 + u * kW2Slice, kW2Slice, &full[st]);
            }
            // N half hf covers X SUBs 4hf..4hf+3; this CTA provides SUBs 4hf+2c, 4hf+2c+1
            if (share_x) {
              // 2 nb slices per CTA and stage; this CTA fetches every other one (slice parity = pr) for both holders
              for (int sl = (int)pr; sl < 2 * nb; sl += 2) {
                const int hf = sl >> 1, w = sl & 1;
                bulk_g2s_mc(dst + (2 + sl) * kW2Slice, x + (size_t)(4 * hf + w) * kSubBytes + u * kW2Slice, kW2Slice,
                            &full[st], x_mask);
              }
            } else {
              for (int hf = 0; hf < nb; ++hf) {
                bulk_g2s(dst + (2 + 2 * hf) * kW2Slice, x + (size_t)(4 * hf) * kSubBytes + u * kW2Slice,
                         kW2Slice, &full[st]);
                bulk_g2s(dst + (3 + 2 * hf) * kW2Slice, x + (size_t)(4 * hf + 1) * kSubBytes + u * kW2Slice,
                         kW2Slice, &full[st]);
              }
            }
          }
          par ^= 1;
          ph ^= (par == 0);
        }
      }
    }
  } else if (warp == 1 && !leader) {
    // ===================== peer: relay "my part of the stage has landed" =====================
    if (elect_one()) {
      uint32_t ph = 0, par = 0;
      for (int si = seg_lo; si < seg_hi; ++si) {
        const long long nt = a.segs[si].tile_hi - a.segs[si].tile_lo;
        for (long long i = 0; i < nt; ++i) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const uint32_t st = par * 4 + u;
            mbar_wait(&full[st], ph);
            mbar_arrive_cluster_relaxed(full_c + st * 8);
          }
          par ^= 1;
          ph ^= (par == 0);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader) =====================
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(256, 256, true, true);
      // MN-major, SWIZZLE_128B: LBO = distance between 64-wide MN atoms (one slice), SBO = 8 K rows
      const uint64_t a_desc0 = umma_smem_desc(s_base, kW2Slice, 1024);
      const uint64_t b_desc0 = umma_smem_desc(s_base + 2 * kW2Slice, kW2Slice, 1024);
      uint32_t ph = 0, par = 0, par_free = 0;
      bool first_seg = true;
      for (int si = seg_lo; si < seg_hi; ++si) {
        const WgSeg sg = a.segs[si];
        const int nb = sg.n_sub / 4;
        if (!first_seg) {                 // previous segment's accumulators have been read out
          mbar_wait(d_free, par_free);
          par_free ^= 1;
          tc_fence_after();
        }
        first_seg = false;
        uint32_t acc = 0;
        for (long long tile = sg.tile_lo; tile < sg.tile_hi; ++tile) {
          const uint64_t soff = (uint64_t)((par * 4 * kW2StageBytes) >> 4);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            mbar_wait(&full[par * 4 + u], ph);
            tc_fence_after();
            const uint64_t ad = a_desc0 + soff + (uint64_t)((u * kW2StageBytes) >> 4);
            const uint64_t bd = b_desc0 + soff + (uint64_t)((u * kW2StageBytes) >> 4);
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
              umma2_bf16(tmem, ad + (uint64_t)((ks * 2048) >> 4), bd + (uint64_t)((ks * 2048) >> 4), idesc,
                         acc | (uint32_t)ks);
              if (nb == 2)
                umma2_bf16(tmem + 256, ad + (uint64_t)((ks * 2048) >> 4),
                           bd + (uint64_t)((2 * kW2Slice + ks * 2048) >> 4), idesc, acc | (uint32_t)ks);
            }
            acc = 1;
            umma2_commit_mc(&empty[par * 4 + u], stage_mask);
          }
          par ^= 1;
          ph ^= (par == 0);
        }
        umma2_commit_mc(d_done, pair_mask);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> fp32 partial =====================
    const int q = warp & 3;
    const uint32_t r = q * 32 + lane;
    uint32_t par_done = 0;
    [[maybe_unused]] uint32_t g_ph = 0, g_par = 0;   // REBUILD: the producer's stage sequence, followed in lock step
    for (int si = seg_lo; si < seg_hi; ++si) {
      const WgSeg sg = a.segs[si];
      if (REBUILD) {
        const long long nt = sg.tile_hi - sg.tile_lo;
        if (sg.layer == 7) {
          // dY7[row, col] = bf16(g[row] * w8[col]) where ReLU bit of X8[row, col] is set: warp q builds the 32 columns
          // 32 q .. of this CTA's 128 (thread = one row of the 32-row K chunk), exactly k_backward2's expression
          const uint32_t col = (uint32_t)(256 * sg.ocp + 128 * crank + 32 * q);
          float w8r[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) w8r[i] = __ldg(a.w8 + col + i);
          for (long long tile = sg.tile_lo; tile < sg.tile_hi; ++tile) {
            const uint32_t* mt = a.relu_mask + (size_t)tile * (8 * 16 * kTileM) + (size_t)(7 * 16 + (col >> 5)) * kTileM;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint32_t st = g_par * 4 + u;
              const uint32_t row = 32 * u + lane;
              const float gr = __ldg(a.g + tile * kTileM + row);
              const uint32_t m = __ldg(mt + row);
              uint32_t v[32], p[16];
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(gr * w8r[i]);
              mask_pack32(v, m, p);
              mbar_wait(&empty[st], g_ph ^ 1);
              // slice (q >> 1) of the stage = SUB 2 crank + (q >> 1) of the out chunk, rows 32 u ..; columns 32 (q & 1) ..
              const uint32_t base = s_base + st * kW2StageBytes + (uint32_t)(q >> 1) * kW2Slice;
#pragma unroll
              for (int i = 0; i < 4; ++i)
                st_shared_v4(base + sub_chunk_off(lane, (uint32_t)((q & 1) * 4 + i)), p[4 * i], p[4 * i + 1], p[4 * i + 2],
                             p[4 * i + 3]);
              fence_proxy_async_smem();
              named_bar_sync(2, 128);
              if (threadIdx.x == 128) mbar_arrive(&full[st]);
            }
            g_par ^= 1;
            g_ph ^= (g_par == 0);
          }
        } else {
          g_ph ^= (uint32_t)((g_par + nt) >> 1) & 1u;
          g_par ^= (uint32_t)nt & 1u;
        }
      }
      float* out = a.partials + ((size_t)sg.slot * 256 + crank * 128 + r) * kHidden +
                   ((QUAD && sg.layer == 3) ? 256 * pr : 0);
      const int ncols = sg.n_sub * 64;
      mbar_wait(d_done, par_done);
      par_done ^= 1;
      tc_fence_after();
      for (int c = 0; c < ncols; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + c, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<uint4*>(out + c + 4 * i) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(d_free); else mbar_arrive_cluster_relaxed(d_free_c);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc2<512>(tmem);
}

}  // namespace dsdf

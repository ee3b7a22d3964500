// Data layout shared by every kernel of the DeepSDF hot path.
//
// Network (reference networks/deep_sdf_decoder.py:25-46 with dims=[512]*8, latent 256,
// latent_in=[4]): lin0 259->512, lin1,2 512->512, lin3 512->253, lin4 512->512 with input
// [x3(253) | latent(256) | xyz(3)], lin5..7 512->512, lin8 512->1.
//
// On-chip / in-HBM activation tile ("image"): 128 rows x 64 bf16 columns = 16 KB "SUB",
// in the canonical UMMA SWIZZLE_128B layout:
//     byte(r, c) = (r/8)*1024 + (r%8)*128 + (((c/8) ^ (r%8)) * 16) + (c%8)*2
// The same bytes are a K-major operand (rows = M or N, columns = K; forward / dgrad) and an
// MN-major operand (rows = K, columns = M or N; wgrad), so activations stashed by the forward
// and backward kernels are consumed by the wgrad kernel without any re-layout.
// A 128-row activation tile with 512 features = 8 consecutive SUBs (128 KB).
//
// Padded layer geometry used by the tensor-core kernels ("pad" order):
//   lin0 : K = latent(256)                         N = 512   + E
//   lin3 :                                          N = 256   (253 real + 3 zero rows)
//   lin4 : K = [x3(256: 253 real + 3 zero) | latent(256)]     + E
//   E    : one extra K-step of 16 columns
//          [xh xh xh | xl xl xl | xh xh xh | 1 | 1 | 0 x5]  (A side: xyz split hi/lo bf16, ones)
//          [wh wh wh | wh wh wh | wl wl wl | bh| bl| 0 x5]  (B side: xyz weights hi/lo, bias hi/lo)
//          => x*w to ~2^-17 relative and the bias added inside the MMA for every layer.
#pragma once
#include <stdint.h>

// forward weight-ring depth (4 KB stages per CTA of a pair).  Measured on B200: 8 stages cost 24 % of the chain
// kernels, 16 -> 20 gains 1-3 % (the knee of the latency curve); 20 is what shared memory allows.  The backward
// kernel keeps its 16-stage ring with compile-time stage descriptors (a running stage index measured 1.5 % slower
// there: its MMA issuer lives under a 56-register setmaxnreg budget).
#ifndef DSDF_FWD_RING
#define DSDF_FWD_RING 20
#endif

namespace dsdf {

constexpr int kLatent = 256;
constexpr int kHidden = 512;
constexpr int kNumLin = 9;       // lin0..lin8
constexpr int kSkipLayer = 4;    // latent_in = [4]
#ifdef DSDF_ARCH_SIREN
// networks/siren_decoder.py (8x512, latent_in=[4], no xyz_in): lin3 512->256, lin4 [x3(256) | latent(256)],
// sin(30 x) activations, no final tanh (SirenDecoder.__init__/forward :98-237).  Same padded geometry.
constexpr int kX3 = 256;         // out features of lin3
constexpr bool kXyzAtSkip = false;
#else
constexpr int kX3 = 253;         // out features of lin3
constexpr bool kXyzAtSkip = true;
#endif
constexpr int kIn0 = 259;

constexpr int kTileM = 128;      // rows per tile (= TMEM lanes)
constexpr int kSubCols = 64;     // bf16 columns per SUB (128 B rows)
constexpr int kSubBytes = kTileM * kSubCols * 2;  // 16384
constexpr int kTileBytes = 8 * kSubBytes;         // 131072: 128 rows x 512 bf16

// byte offset of element (r, c) inside one SUB (r < 128, c < 64)
__host__ __device__ __forceinline__ uint32_t sub_off(uint32_t r, uint32_t c) {
  return (r >> 3) * 1024u + (r & 7u) * 128u + ((((c >> 3) ^ r) & 7u) << 4) + ((c & 7u) << 1);
}
// byte offset of 16-byte chunk `ch` (8 bf16) of row r inside one SUB
__host__ __device__ __forceinline__ uint32_t sub_chunk_off(uint32_t r, uint32_t ch) {
  return (r >> 3) * 1024u + (r & 7u) * 128u + (((ch ^ r) & 7u) << 4);
}

// ---- v2 forward weight stream ("K-step image") ----------------------------------------------
// The v2 forward chain consumes B as a stream of K-steps: one step = [256 n x 16 k] bf16 = 8 KB in
// the canonical no-swizzle K-major UMMA layout (8 n-rows x 16 B core matrices, 128 B each):
//     byte(n, k) = (n/8)*256 + (k/8)*128 + (n%8)*16 + (k%8)*2          (LBO = 128, SBO = 256)
// Per layer l (0..7) and output half h (n in [256h, 256h+256); lin3 has only h = 0):
//     step 0      : E step  [wxh wyh wzh | wxh wyh wzh | wxl wyl wzl | bh | bl | 0 x5]
//     step 1 + s  : padded-K columns 16s .. 16s+15 of W_l   (pad order: pad_to_ref_col)
// lin4's first 16 W steps are the x3 part, the last 16 the latent part; lin0 has 16 (latent).
// In "fold" mode (all rows of a tile share one latent) the latent steps are skipped and the E
// steps of lin0/lin4 come from a per-segment image with bias' = b + W[:, latent] z (fp32, hi/lo).
constexpr int kStepBytes = 8192;
__host__ __device__ __forceinline__ uint32_t step_off(uint32_t n, uint32_t k) {
  return (n >> 3) * 256u + (k >> 3) * 128u + (n & 7u) * 16u + (k & 7u) * 2u;
}
__host__ __device__ constexpr int w2_ksteps(int l) { return l == 0 ? 16 : 32; }
__host__ __device__ constexpr int w2_halves(int l) { return l == 3 ? 1 : 2; }
__host__ __device__ constexpr int w2_half_steps(int l) { return 1 + w2_ksteps(l); }
__host__ __device__ constexpr int w2_layer_base(int l) {
  int b = 0;
  for (int i = 0; i < l; ++i) b += w2_halves(i) * w2_half_steps(i);
  return b;
}
__host__ __device__ constexpr int w2_step(int l, int h, int s) {
  return w2_layer_base(l) + h * w2_half_steps(l) + s;
}
constexpr int kW2Steps = w2_layer_base(8);  // 34 + 66*2 + 33 + 66*4 = 463 steps = 3.79 MB
// per-segment E image (fold mode): [seg][which: lin0, lin4][h] steps
constexpr int kSceneESteps = 4;

// ---- v2 dgrad weight stream ("K-step image" of W^T) -------------------------------------------
// The v2 backward chain computes dX_l[128 x in] = dY_l[128 x out] W_l as, per output half h
// (i in [256h, 256h+256) of the padded input order), a stream of K-steps over o:
//     step s: [256 i x 16 o] bf16, same no-swizzle layout as the forward stream (step_off),
//             element (n, k) = W_l[o = 16 s + k, pad_to_ref_col(l, 256 h + n)]   (0 when padded)
// Layers in stream order 7, 6, 5, 4, 3, 2, 1, then 0 (dense-input backward only: latent part).
// lin4's second half (latent columns) is only consumed by the dense-input backward.
__host__ __device__ constexpr int dg2_ksteps(int l) { return l == 3 ? 16 : 32; }          // out / 16
__host__ __device__ constexpr int dg2_halves(int l) { return l == 0 ? 1 : 2; }            // in(pad) / 256
__host__ __device__ constexpr int dg2_layer_base(int l) {   // stream order 7..0
  int b = 0;
  for (int i = 7; i > l; --i) b += dg2_halves(i) * dg2_ksteps(i);
  return b;
}
__host__ __device__ constexpr int dg2_step(int l, int h, int s) {
  return dg2_layer_base(l) + h * dg2_ksteps(l) + s;
}
constexpr int kDg2Steps = dg2_layer_base(0) + dg2_halves(0) * dg2_ksteps(0);  // 64*6 + 32 + 32 = 448

// padded K index of layer l -> reference input column (or -1 for a zero pad column)
__host__ __device__ __forceinline__ int pad_to_ref_col(int l, int kk) {
  if (l == 0) return kk;  // kk < 256: latent
  if (l == kSkipLayer) {
    if (kk < 256) return kk < kX3 ? kk : -1;
    return kX3 + (kk - 256);  // latent part: ref cols 253..508
  }
  return kk;
}
__host__ __device__ constexpr int ref_in(int l) { return l == 0 ? kIn0 : kHidden; }
__host__ __device__ constexpr int ref_out(int l) { return l == 3 ? kX3 : (l == 8 ? 1 : kHidden); }
// reference column of xyz component e (0..2) for layers with an xyz input, else -1
__host__ __device__ constexpr int xyz_col(int l, int e) {
  return l == 0 ? 256 + e : ((l == kSkipLayer && kXyzAtSkip) ? kX3 + 256 + e : -1);
}

// ---- activation stash in HBM ---------------------------------------------------------------
// X stash : [tile][slot 0..7][8 SUB]  slot s = input of lin(s+1) (= relu output of lin s);
//           slot 3 (input of lin4) holds only the x3 part in SUB 0..3.
// dY stash: [tile][slot 0..7][8 SUB]  slot l = gradient w.r.t. the pre-activation of lin l
//           (l = 1..7 used in scene mode; lin3 has 4 SUBs).
constexpr size_t kStashTileBytes = 8 * (size_t)kTileBytes;  // 1 MiB per 128-row tile
// ---- activation-derivative side channel forward -> backward, per 128-row tile ----------------------
// ReLU nets : 1 bit / element: [8 layers][16 chunks of 32 columns][128 rows] words, bit (31-i) = column
//             32*chunk+i active (and kept by dropout)
// SIREN     : round(127 cos(30 x)) as a signed byte / element: [8 layers][16 chunks][2][128 rows][16 B] (a warp's
//             32 rows write 512 contiguous bytes per store instruction)
#ifdef DSDF_ARCH_SIREN
constexpr size_t kActAuxTileBytes = 8 * 16 * (size_t)kTileM * 32;   // 512 KiB: one signed byte per activation
#else
constexpr size_t kActAuxTileBytes = 8 * 16 * (size_t)kTileM * 4;    // 64 KB
#endif

}  // namespace dsdf

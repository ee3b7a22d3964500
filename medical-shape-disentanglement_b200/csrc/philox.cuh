// Counter-based dropout stream: Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as
// 1, 2, 3", SC'11; the generator behind torch.cuda / curand Philox), keyed by (seed) and counted by
// (row, layer, column block, step offset), so the keep/drop decision of every activation element is
// a pure function of its coordinates: no RNG state, no ordering dependence, identical under any
// tiling or sharding, and the backward pass needs no regeneration (the decision is folded into the
// forward kernel's 1-bit/element ReLU mask).
//
// Replaces F.dropout(x, p=dropout_prob, training=True) of reference
// networks/deep_sdf_decoder.py:103-104: y = keep ? x / (1 - p) : 0 with P(keep) = 1 - p.
// Each Philox call yields 4 x 32 bits = eight 16-bit uniforms; element e of an 8-column block is
// dropped iff its 16-bit uniform < round(p * 65536).
#pragma once
#include <stdint.h>

namespace dsdf {

struct DropoutArgs {
  uint32_t layers;      // bit l set: dropout after the ReLU of lin l (l = 0..7)
  uint32_t thresh16;    // round(p * 65536): drop iff u16 < thresh16
  float scale;          // 1 / (1 - p)
  uint32_t seed_lo, seed_hi;
  uint32_t offset;      // step counter (a new mask every optimisation step)
  long long row_base;   // global index of row 0 of this call (chunked / sharded callers)
};

__host__ __device__ __forceinline__ void philox_mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
  const uint64_t p = (uint64_t)a * (uint64_t)b;
  hi = (uint32_t)(p >> 32);
  lo = (uint32_t)p;
}

__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0, lo0, hi1, lo1;
    philox_mulhilo(0xD2511F53u, c0, hi0, lo0);
    philox_mulhilo(0xCD9E8D57u, c2, hi1, lo1);
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// keep bits of columns 32*chunk32 .. +31 of global row `row` after lin `layer`:
// bit (31 - i) set <=> column 32*chunk32 + i is kept (same bit order as the ReLU mask words)
__host__ __device__ __forceinline__ uint32_t dropout_keep32(const DropoutArgs& d, long long row, int layer,
                                                            int chunk32) {
  uint32_t keep = 0;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    uint32_t o[4];
    philox4x32_10((uint32_t)row, (uint32_t)((unsigned long long)row >> 32),
                  ((uint32_t)layer << 16) | (uint32_t)(chunk32 * 4 + b), d.offset, d.seed_lo, d.seed_hi, o);
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      keep = (keep << 1) | ((o[w] & 0xFFFFu) >= d.thresh16 ? 1u : 0u);
      keep = (keep << 1) | ((o[w] >> 16) >= d.thresh16 ? 1u : 0u);
    }
  }
  return keep;
}

}  // namespace dsdf

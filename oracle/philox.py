"""Oracle restatement (TEST INFRASTRUCTURE ONLY) of the counter-based dropout stream of
medical-shape-disentanglement_b200/csrc/philox.cuh: Philox4x32-10 (Salmon, Moraes, Dror, Shaw,
"Parallel random numbers: as easy as 1, 2, 3", SC'11 -- the published algorithm, pinned by the
Random123 known-answer vectors in tests/test_oracle_cpu.py) and the keep/drop rule that stands in
for the RNG draw of F.dropout (reference networks/deep_sdf_decoder.py:103-104).  numpy, vectorised."""
from __future__ import annotations

import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr: np.ndarray, key: np.ndarray) -> np.ndarray:
    """ctr [..., 4] uint32, key [2] uint32 -> [..., 4] uint32."""
    c = [ctr[..., i].astype(np.uint64) for i in range(4)]
    k0, k1 = int(key[0]), int(key[1])
    for _ in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return np.stack(c, -1).astype(np.uint32)


def dropout_keep(seed: int, offset: int, layer: int, row_lo: int, n_rows: int, prob: float,
                 n_cols: int = 512) -> np.ndarray:
    """keep[n_rows, n_cols] uint8 (1 = kept): element (row, col) uses call block col // 8 with counter
    (row_lo32, row_hi32, layer << 16 | block, offset) and its 16-bit uniform #(col % 8)
    (word (col % 8) // 2, low half first); dropped iff u16 < round(prob * 65536)."""
    thresh = int(round(prob * 65536.0))
    rows = np.arange(row_lo, row_lo + n_rows, dtype=np.uint64)
    blocks = np.arange(n_cols // 8, dtype=np.uint32)
    ctr = np.zeros((n_rows, n_cols // 8, 4), dtype=np.uint32)
    ctr[..., 0] = (rows & MASK).astype(np.uint32)[:, None]
    ctr[..., 1] = (rows >> np.uint64(32)).astype(np.uint32)[:, None]
    ctr[..., 2] = (np.uint32(layer) << np.uint32(16)) | blocks[None, :]
    ctr[..., 3] = np.uint32(offset & 0xFFFFFFFF)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32)
    o = philox4x32_10(ctr, key)                                    # [rows, blocks, 4]
    halves = np.stack([o & np.uint32(0xFFFF), o >> np.uint32(16)], -1)   # [rows, blocks, 4, 2]
    u16 = halves.reshape(n_rows, n_cols)
    return (u16 >= thresh).astype(np.uint8)

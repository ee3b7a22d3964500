"""Oracle restatement of create_mesh's grid + chunked decode (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/deep_sdf/mesh.py:35-71 and deep_sdf/utils.py:86-97.

The reference builds the grid with *true* division on a LongTensor
(`(overall_index.long() / N) % N`, mesh.py:46-47), so the y/x "indices" are fractional:
   z_i = float(idx % N)
   y_i = fmod(float(idx) / N, N)
   x_i = fmod((float(idx) / N) / N, N)
and coordinate = v * float32(2/(N-1)) + (-1)  (mul, then add; mesh.py:51-53; column 0 is
the slowest index).  `grid_coords_reference_ops` replays the reference's torch expressions
literally; `grid_coords_closed_form` is the numpy closed form the CUDA kernel implements.
tests/test_oracle_cpu.py asserts they are bit-equal.
"""
from __future__ import annotations

import numpy as np
import torch

from . import deepsdf


def grid_coords_reference_ops(N: int, lo: int = 0, hi: int | None = None) -> torch.Tensor:
    """Literal torch replay of mesh.py:38-53 restricted to linear indices [lo, hi)."""
    hi = N ** 3 if hi is None else hi
    voxel_origin = [-1, -1, -1]
    voxel_size = 2.0 / (N - 1)
    overall_index = torch.arange(lo, hi, 1, dtype=torch.long)
    samples = torch.zeros(hi - lo, 3)
    samples[:, 2] = overall_index % N
    samples[:, 1] = (overall_index.long() / N) % N
    samples[:, 0] = ((overall_index.long() / N) / N) % N
    samples[:, 0] = (samples[:, 0] * voxel_size) + voxel_origin[2]
    samples[:, 1] = (samples[:, 1] * voxel_size) + voxel_origin[1]
    samples[:, 2] = (samples[:, 2] * voxel_size) + voxel_origin[0]
    return samples


def grid_coords_closed_form(N: int, lo: int = 0, hi: int | None = None) -> np.ndarray:
    """fp32 closed form of the same coordinates (what the CUDA grid generator computes)."""
    hi = N ** 3 if hi is None else hi
    idx = np.arange(lo, hi, dtype=np.int64)
    f = idx.astype(np.float32)              # int64 -> fp32 (round to nearest even)
    nf = np.float32(N)
    a = f / nf                              # fp32 true division
    zi = (idx % N).astype(np.float32)
    yi = np.fmod(a, nf)
    xi = np.fmod(a / nf, nf)
    vs = np.float32(2.0 / (N - 1))
    one = np.float32(1.0)
    out = np.empty((hi - lo, 3), dtype=np.float32)
    out[:, 0] = xi * vs - one
    out[:, 1] = yi * vs - one
    out[:, 2] = zi * vs - one
    return out


def decode_sdf(params, spec: deepsdf.DecoderSpec, latent: torch.Tensor | None,
               queries: torch.Tensor) -> torch.Tensor:
    """deep_sdf/utils.py:86-97."""
    if latent is None:
        inputs = queries
    else:
        inputs = torch.cat([latent.expand(queries.shape[0], -1), queries], 1)
    return deepsdf.forward(params, spec, inputs, training=False)


def decode_grid(params, spec: deepsdf.DecoderSpec, latent: torch.Tensor, N: int,
                max_batch: int = 2 ** 18, lo: int = 0, hi: int | None = None,
                device: str = "cpu") -> torch.Tensor:
    """mesh.py:55-71 restricted to [lo, hi): returns sdf values fp32 [hi-lo] (CPU)."""
    hi = N ** 3 if hi is None else hi
    samples = grid_coords_reference_ops(N, lo, hi)
    out = torch.empty(hi - lo)
    params = {k: v.to(device) for k, v in params.items()}
    latent = latent.to(device)
    head = 0
    with torch.no_grad():
        while head < hi - lo:
            sub = samples[head:head + max_batch].to(device)
            out[head:head + max_batch] = decode_sdf(params, spec, latent, sub).squeeze(1).cpu()
            head += max_batch
    return out

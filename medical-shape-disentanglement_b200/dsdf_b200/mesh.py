"""Grid decode of deep_sdf/mesh.create_mesh (reference deep_sdf/mesh.py:21-71) on the fused
forward kernel: coordinates are generated in-kernel (bit-exact with the reference's fp32 torch
expressions, including the true-division shear of mesh.py:46-47), the latent is read once per
tile instead of expand+cat per chunk, and there is one D2H of N^3 floats at the end instead
of one H2D + one D2H per 2^18-point chunk.

Marching cubes / PLY writing (mesh.py:96-165) stay on the host in the reference's own code
(skimage / plyfile are not available in this image): `create_mesh_sdf` returns exactly the
`sdf_values.reshape(N, N, N)` tensor that `convert_sdf_samples_to_ply` consumes (mesh.py:70-71).
"""
from __future__ import annotations

from typing import Optional

import torch

from .decoder import Decoder


@torch.no_grad()
def create_mesh_sdf(decoder: Decoder, latent_vec: torch.Tensor, N: int = 256,
                    max_batch: int = 32 ** 3, to_cpu: bool = True) -> torch.Tensor:
    """Drop-in for the sampling half of create_mesh: returns sdf_values [N,N,N] fp32.

    `max_batch` is accepted for signature compatibility; the persistent kernel streams all
    N^3 points in one launch (the chunking existed to bound the reference's [m,259] cat).
    """
    decoder.eval()
    dev = next(decoder.parameters()).device
    lat = latent_vec.to(dev)
    out = decoder.decode_grid(lat, N)
    out = out.view(N, N, N)
    return out.cpu() if to_cpu else out


@torch.no_grad()
def decode_grid_sharded(decoder: Decoder, latent_vec: torch.Tensor, N: int, rank: int,
                        world_size: int) -> tuple:
    """Voxel-slab shard of the grid: rank r decodes x-slabs [r*N/W, (r+1)*N/W) = a contiguous
    linear-index range.  Returns (lo, hi, sdf[hi-lo]) on this rank's device; no collective."""
    lo = (N * rank // world_size) * N * N
    hi = (N * (rank + 1) // world_size) * N * N
    dev = next(decoder.parameters()).device
    return lo, hi, decoder.decode_grid(latent_vec.to(dev), N, lo, hi)

"""Synthetic SDF samples of the reference's on-disk shape (deep_sdf/data.py:99-136,
src/PreprocessMesh.cpp:291-313,530-546): per scene a pool of points, 6 % uniform in the cube and
94 % near the surface of a sphere of radius 0.5 (two noise variances), sdf = |xyz| - 0.5,
each scene's P rows = P/2 positive-sdf rows followed by P/2 negative-sdf rows."""
from __future__ import annotations

import math

import torch


def scene_samples(num_scenes: int, samples_per_scene: int, seed: int = 0,
                  device: str = "cpu") -> torch.Tensor:
    """[num_scenes, P, 4] fp32 (x, y, z, sdf)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    P = samples_per_scene
    half = P // 2
    pool = 2 * P
    out = torch.empty(num_scenes, P, 4)
    for s in range(num_scenes):
        n_uni = int(0.06 * pool)
        n_surf = pool - n_uni
        uni = torch.rand(n_uni, 3, generator=g) * 2 - 1
        d = torch.randn(n_surf, 3, generator=g)
        d = d / d.norm(dim=1, keepdim=True) * (0.5 + 0.02 * math.sin(s))
        sig = torch.where(torch.arange(n_surf) % 2 == 0, math.sqrt(0.005), math.sqrt(0.0005))
        surf = d + torch.randn(n_surf, 3, generator=g) * sig[:, None]
        xyz = torch.cat([uni, surf], 0)
        sdf = xyz.norm(dim=1) - 0.5
        pos = torch.nonzero(sdf > 0).flatten()
        neg = torch.nonzero(sdf <= 0).flatten()
        ip = pos[torch.randint(0, pos.numel(), (half,), generator=g)]
        ineg = neg[torch.randint(0, neg.numel(), (P - half,), generator=g)]
        idx = torch.cat([ip, ineg])
        out[s, :, :3] = xyz[idx]
        out[s, :, 3] = sdf[idx]
    return out.to(device)


def fast_scene_samples(num_scenes: int, samples_per_scene: int, seed: int = 0,
                       device: str = "cuda") -> torch.Tensor:
    """Same distribution family generated directly on `device` (bench-size inputs)."""
    g = torch.Generator(device=device).manual_seed(seed)
    S, P = num_scenes, samples_per_scene
    d = torch.randn(S, P, 3, generator=g, device=device)
    d = d / d.norm(dim=2, keepdim=True) * 0.5
    sig = torch.where(torch.arange(P, device=device) % 2 == 0, math.sqrt(0.005), math.sqrt(0.0005))
    xyz = d + torch.randn(S, P, 3, generator=g, device=device) * sig[None, :, None]
    uni = torch.rand(S, P, 3, generator=g, device=device) * 2 - 1
    pick = (torch.rand(S, P, 1, generator=g, device=device) < 0.06)
    xyz = torch.where(pick, uni, xyz)
    sdf = xyz.norm(dim=2, keepdim=True) - 0.5
    # first half positive, second half negative (mirror through the surface where needed)
    want_pos = (torch.arange(P, device=device) < P // 2)[None, :, None]
    flip = (sdf > 0) != want_pos
    r = xyz.norm(dim=2, keepdim=True).clamp_min(1e-6)
    xyz = torch.where(flip, xyz / r * (1.0 - r).abs().clamp(0.0, 1.7), xyz)
    sdf = xyz.norm(dim=2, keepdim=True) - 0.5
    return torch.cat([xyz, sdf], 2).float().contiguous()

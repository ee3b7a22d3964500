"""Oracle restatement of reconstruct.reconstruct() (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/reconstruct.py:16-151 (frozen decoder, Adam on one latent) and the
sample window draw of /root/reference/deep_sdf/data.py:112-136.  The optional
`code_reg_lambda` (`l2_sq` / `l2_norm`) and `dist_*` regularisers of the reference default to off.

Window offsets come from Python's `random.randint` in the reference; here they are drawn
up front (`draw_windows`) so the CUDA path can replay the identical sequence.
"""
from __future__ import annotations

import random
from typing import List, Optional, Tuple

import torch

from . import deepsdf


def draw_windows(pos_size: int, neg_size: int, num_samples: int, iters: int,
                 rng: random.Random) -> List[Tuple[int, int]]:
    """data.py:118-134 for the (usual) neg_size > half case."""
    half = int(num_samples / 2)
    assert neg_size > half and pos_size >= half
    out = []
    for _ in range(iters):
        ps = rng.randint(0, pos_size - half)
        ns = rng.randint(0, neg_size - half)
        out.append((ps, ns))
    return out


def draw_plan(pos: torch.Tensor, neg: torch.Tensor, num_samples: int, iters: int, rng=random):
    """data.py:112-134 including the small-shape branch (:123-125): when neg_size <= half the reference draws
    `(torch.rand(half) * neg_size).long()` from torch's global CPU generator and index_selects.  Returns
    (neg_pool, windows) such that window(pos, neg_pool, num_samples, windows[e]) is iteration e's sample: in the
    small-shape branch neg_pool is the concatenation of every iteration's selected rows."""
    half = int(num_samples / 2)
    if neg.shape[0] > half:
        return neg, draw_windows(pos.shape[0], neg.shape[0], num_samples, iters, rng)
    wins, picked = [], []
    for e in range(iters):
        ps = rng.randint(0, pos.shape[0] - half)                         # :120
        random_neg = (torch.rand(half) * neg.shape[0]).long()            # :124
        picked.append(torch.index_select(neg, 0, random_neg))            # :125
        wins.append((ps, e * half))
    return torch.cat(picked, 0), wins


def window(pos: torch.Tensor, neg: torch.Tensor, num_samples: int, w: Tuple[int, int]):
    half = int(num_samples / 2)
    return torch.cat([pos[w[0]:w[0] + half], neg[w[1]:w[1] + half]], 0)


def reconstruct(params, spec: deepsdf.DecoderSpec, num_iterations: int, pos: torch.Tensor,
                neg: torch.Tensor, windows, latent_init: torch.Tensor, clamp_dist: float,
                num_samples: int = 8000, lr: float = 5e-3, l2reg: bool = True,
                code_reg_lambda: Optional[float] = None, code_reg_type: str = "l2_sq",
                code_bound: Optional[float] = None, device: str = "cpu",
                dist_mean: Optional[torch.Tensor] = None, dist_std: Optional[torch.Tensor] = None,
                dist_weight: float = 0.0, dist_type: str = "zscore_l2", return_first_grad: bool = False):
    decreased_by = 10
    adjust_lr_every = max(1, int(num_iterations / 2))                    # :43-44
    params = {k: v.to(device) for k, v in params.items()}
    latent = latent_init.clone().to(device).requires_grad_(True)         # :49-54
    opt = torch.optim.Adam([latent], lr=lr)                              # :56
    losses = []
    first_grad = None
    if dist_mean is not None:                                            # :58-70
        dist_mean = dist_mean.to(device)
    if dist_std is not None:
        dist_std = torch.clamp(dist_std.to(device), min=1e-8)
    for e in range(num_iterations):
        sdf_data = window(pos, neg, num_samples, windows[e]).to(device)  # :80
        xyz = sdf_data[:, 0:3]
        gt = torch.clamp(sdf_data[:, 3:4], -clamp_dist, clamp_dist)      # :84
        cur_lr = lr * ((1 / decreased_by) ** (e // adjust_lr_every))     # :35-44
        for g in opt.param_groups:
            g["lr"] = cur_lr
        opt.zero_grad()
        inputs = torch.cat([latent.expand(num_samples, -1), xyz], 1)     # :90-92
        pred = deepsdf.forward(params, spec, inputs, training=False)     # :94
        pred = torch.clamp(pred, -clamp_dist, clamp_dist)                # :100
        loss = torch.nn.functional.l1_loss(pred, gt)                     # :102 (mean)
        if code_reg_lambda is not None and float(code_reg_lambda) > 0.0: # :106-113
            if code_reg_type in ("l2_norm", "l2norm", "norm"):
                loss = loss + float(code_reg_lambda) * latent.norm(dim=1).mean()
            else:
                loss = loss + float(code_reg_lambda) * torch.mean(latent.pow(2))
        elif l2reg:
            loss = loss + 1e-4 * torch.mean(latent.pow(2))               # :114-116
        if dist_weight > 0.0 and dist_mean is not None:                  # :117-129
            diff = (latent - dist_mean) / dist_std if dist_std is not None else latent - dist_mean
            if dist_type in ("zscore_l2", "diag_gaussian", "gaussian", "l2", "mse"):
                loss = loss + dist_weight * torch.mean(diff.pow(2))
            else:
                loss = loss + dist_weight * torch.mean(diff.abs())
        loss.backward()                                                  # :130
        if e == 0:
            first_grad = latent.grad.detach().clone().cpu()
        opt.step()                                                       # :131
        if code_bound is not None and float(code_bound) > 0:             # :134-140
            with torch.no_grad():
                n = latent.norm(dim=1, keepdim=True)
                latent.mul_(torch.clamp(float(code_bound) / (n + 1e-12), max=1.0))
        losses.append(float(loss.detach().cpu()))
    if return_first_grad:
        return losses, latent.detach().cpu(), first_grad
    return losses, latent.detach().cpu()

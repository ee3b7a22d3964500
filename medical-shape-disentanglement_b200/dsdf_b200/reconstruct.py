"""Latent-only optimisation against a frozen decoder: drop-in for reference
reconstruct.reconstruct() (reconstruct.py:16-151), plus the batched form (many independent
shapes in one launch sequence) that the reference loops over one shape at a time
(reconstruct.py:312-420, and the in-training evaluation of train_deep_sdf.py).

The whole `num_iterations` loop is ONE call into libdsdf_b200.so (dsdf_reconstruct, include/dsdf.h):
per iteration a window gather, the fused tcgen05 forward + clamped-L1 + dgrad chain, the per-shape
latent gradient (segmented column sums; sum_rows dX[:, latent] = colsum(dY) W[:, latent]), Adam and
the optional norm bound -- no wgrad, no activation stash, no per-iteration host sync.

Randomness: the reference draws the two window starts of every iteration with Python's
`random.randint` (deep_sdf/data.py:112-131) and the initial latent with torch's RNG
(reconstruct.py:49-52).  `reconstruct()` here consumes both generators in the same order, so with
the same `random.seed` / `torch.manual_seed` it walks the reference's exact sample sequence.
"""
from __future__ import annotations

import ctypes as C
import random
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from .decoder import LATENT, Decoder, _stream_ptr

_REG_KINDS = {"l2_sq": 1, "l2_squared": 1, "l2": 1, "sq": 1, "squared": 1,
              "l2_norm": 2, "l2norm": 2, "norm": 2}
_DIST_KINDS = {"zscore_l2": 1, "diag_gaussian": 1, "gaussian": 1, "l2": 1, "mse": 1, "l1": 2, "abs": 2}


class ReconstructCfg(C.Structure):
    _fields_ = [("num_iterations", C.c_int64), ("num_samples", C.c_int64), ("clamp_dist", C.c_float),
                ("lr", C.c_float), ("reg_kind", C.c_int32), ("reg_lambda", C.c_float),
                ("code_bound", C.c_float), ("dist_kind", C.c_int32), ("dist_weight", C.c_float),
                ("dist_mean", C.c_void_p), ("dist_std", C.c_void_p)]


def draw_windows(pos_size: int, neg_size: int, num_samples: int, num_iterations: int,
                 rng=random) -> List[Tuple[int, int]]:
    """Window starts of unpack_sdf_samples_from_ram (deep_sdf/data.py:112-131), one pair per
    iteration, consuming `rng` exactly as the reference loop does."""
    half = int(num_samples / 2)
    if neg_size <= half:
        raise NotImplementedError(
            f"shape has {neg_size} negative samples <= num_samples/2 = {half}: the reference falls back to "
            "random index_select (data.py:123-125); use plan_samples(), which lays the selected rows out as windows")
    if pos_size < half:
        raise ValueError(f"shape has {pos_size} positive samples < num_samples/2 = {half}")
    out = []
    for _ in range(num_iterations):
        ps = rng.randint(0, pos_size - half)
        ns = rng.randint(0, neg_size - half)
        out.append((ps, ns))
    return out


def plan_samples(pos: torch.Tensor, neg: torch.Tensor, num_samples: int, num_iterations: int, rng=random):
    """Every iteration's sample of unpack_sdf_samples_from_ram (deep_sdf/data.py:112-134) as (neg_pool, windows)
    for the window gather, consuming `rng` and torch's global CPU generator exactly as the reference loop does.
    Usual case (neg_size > half): neg_pool is `neg` and both starts are random.randint draws.  Small-shape case
    (:123-125, neg_size <= half): the reference draws `(torch.rand(half) * neg_size).long()` and index_selects;
    the selected rows of all iterations are laid end to end (num_iterations * half * 16 B) so iteration e's
    negatives are again a window, starting at e * half."""
    half = int(num_samples / 2)
    if neg.shape[0] > half:
        return neg, draw_windows(pos.shape[0], neg.shape[0], num_samples, num_iterations, rng)
    if pos.shape[0] < half:
        raise ValueError(f"shape has {pos.shape[0]} positive samples < num_samples/2 = {half}")
    if neg.shape[0] == 0:
        raise ValueError("shape has no negative samples")
    src = neg.detach().cpu()
    wins, picked = [], []
    for e in range(num_iterations):
        ps = rng.randint(0, pos.shape[0] - half)
        picked.append(torch.index_select(src, 0, (torch.rand(half) * src.shape[0]).long()))
        wins.append((ps, e * half))
    pool = torch.cat(picked, 0) if picked else src[:0]
    return pool, wins


def _reg_args(l2reg, code_reg_lambda, code_reg_type):
    if code_reg_lambda is not None and float(code_reg_lambda) > 0.0:      # reconstruct.py:106-113
        t = str(code_reg_type).lower()
        if t not in _REG_KINDS:
            raise ValueError(f"Unknown code_reg_type: {code_reg_type}")
        return _REG_KINDS[t], float(code_reg_lambda)
    if l2reg:                                                             # :114-116
        return 1, 1e-4
    return 0, 0.0


@torch.no_grad()
def reconstruct_batch(decoder: Decoder, num_iterations: int, shapes: Sequence[Tuple[torch.Tensor, torch.Tensor]],
                      latents: torch.Tensor, windows: torch.Tensor, clamp_dist: float,
                      num_samples: int = 30000, lr: float = 5e-4, l2reg: bool = False,
                      code_reg_lambda=None, code_reg_type: str = "l2_sq", code_bound=None,
                      dist_mean=None, dist_std=None, dist_weight=0.0, dist_type: str = "zscore_l2",
                      return_state: bool = False, first_step: int = 1, n_steps: Optional[int] = None,
                      state: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
    """Optimise `latents[n_shapes, 256]` (initial values, not modified) against `shapes[i] = (pos, neg)`
    with per-iteration windows `windows[num_iterations, n_shapes, 2]` (int64).  Returns
    (loss_hist [num_iterations, n_shapes] on the device, optimised latents [n_shapes, 256]) and,
    with return_state, the Adam moments (exp_avg, exp_avg_sq) as well.

    The schedule can be run in pieces (to look at the losses, stop early, checkpoint): `first_step` (1-based) and
    `n_steps` select iterations [first_step, first_step + n_steps) of the `num_iterations` schedule -- the learning
    rate drop at num_iterations/2 and Adam's bias correction follow the absolute iteration -- and `state` carries
    the Adam moments returned by the previous piece; `latents` are then the previous piece's result.  Pieces
    chained this way equal the single call bit for bit."""
    dev = next(decoder.parameters()).device
    if dev.type != "cuda":
        raise _lib.DsdfError("reconstruct needs the decoder on a CUDA device (no CPU path)")
    L = decoder.L
    n_shapes = len(shapes)
    if latents.shape != (n_shapes, LATENT):
        raise ValueError(f"latents must be [{n_shapes}, {LATENT}], got {tuple(latents.shape)}")
    if num_samples % 2:
        raise ValueError("num_samples must be even (the reference concatenates latent.expand(num_samples) "
                         "with 2*int(num_samples/2) rows, reconstruct.py:90-92)")
    base, blocks, off = [], [], 0
    for pos, neg in shapes:
        base.append((off, off + pos.shape[0]))
        off += pos.shape[0] + neg.shape[0]
        blocks += [pos.reshape(-1, 4), neg.reshape(-1, 4)]
    samples = torch.cat(blocks, 0).to(dev, torch.float32).contiguous()
    shape_base = torch.tensor(base, dtype=torch.int64).to(dev)
    win = torch.as_tensor(windows, dtype=torch.int64).reshape(num_iterations, n_shapes, 2).contiguous()
    # every window must lie inside its shape's pos / neg block (the gather kernel reads start .. start + half)
    half = num_samples // 2
    if win.numel():
        sizes = torch.tensor([[pos.shape[0], neg.shape[0]] for pos, neg in shapes], dtype=torch.int64)
        wc = win.cpu()
        if int(wc.min()) < 0 or bool((wc + half > sizes[None]).any()):
            bad = ((wc < 0) | (wc + half > sizes[None])).nonzero()[0].tolist()
            raise IndexError(f"window {bad} (iteration, shape, pos/neg) = {int(wc[tuple(bad)])} + {half} rows leaves the "
                             f"shape's block of {int(sizes[bad[1], bad[2]])} rows")
    win = win.to(dev)
    lat = latents.detach().to(dev, torch.float32).clone().contiguous()
    if n_steps is None:
        n_steps = num_iterations - (first_step - 1)
    if first_step < 1 or n_steps < 0 or first_step - 1 + n_steps > num_iterations:
        raise ValueError(f"iterations [{first_step}, {first_step + n_steps}) are outside the schedule 1..{num_iterations}")
    if state is None:
        if first_step != 1:
            raise ValueError("continuing a schedule (first_step > 1) needs the Adam moments of the previous piece")
        m = torch.zeros_like(lat)
        v = torch.zeros_like(lat)
    else:
        m = state[0].detach().to(dev, torch.float32).clone().contiguous()
        v = state[1].detach().to(dev, torch.float32).clone().contiguous()
        if m.shape != lat.shape or v.shape != lat.shape:
            raise ValueError("state must be (exp_avg, exp_avg_sq), each shaped like latents")
    hist = torch.zeros(num_iterations, n_shapes, dtype=torch.float32, device=dev)
    reg_kind, reg_lambda = _reg_args(l2reg, code_reg_lambda, code_reg_type)
    dist_w = float(dist_weight) if dist_weight is not None else 0.0
    dm = ds = None
    dist_kind = 0
    if dist_w > 0.0 and dist_mean is not None:                           # reconstruct.py:58-70,117-129
        t = str(dist_type).lower()
        if t not in _DIST_KINDS:
            raise ValueError(f"Unknown dist_type: {dist_type}")
        dist_kind = _DIST_KINDS[t]
        # the kernel indexes [j] for j < 256: broadcast scalars / [1,256] the way torch would (reconstruct.py:58-70)
        dm = torch.as_tensor(dist_mean, dtype=torch.float32).reshape(-1).to(dev)
        dm = dm.expand(LATENT).contiguous() if dm.numel() == 1 else dm.contiguous()
        if dm.numel() != LATENT:
            raise ValueError(f"dist_mean must have 1 or {LATENT} elements, got {dm.numel()}")
        if dist_std is not None:
            ds = torch.clamp(torch.as_tensor(dist_std, dtype=torch.float32).reshape(-1).to(dev), min=1e-8)
            ds = ds.expand(LATENT).contiguous() if ds.numel() == 1 else ds.contiguous()
            if ds.numel() != LATENT:
                raise ValueError(f"dist_std must have 1 or {LATENT} elements, got {ds.numel()}")
    cfg = ReconstructCfg(num_iterations, num_samples, float(clamp_dist), float(lr), reg_kind, reg_lambda,
                         float(code_bound) if code_bound is not None else 0.0, dist_kind, dist_w,
                         dm.data_ptr() if dm is not None else None, ds.data_ptr() if ds is not None else None)
    decoder.eval()                                                       # reconstruct.py:79
    h = decoder.sync_weights()
    flat = decoder.flat_params()
    nbytes = L.dsdf_reconstruct_workspace_bytes(h, n_shapes, num_samples)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.dsdf_reconstruct(h, flat.data_ptr(), C.byref(cfg), n_shapes, samples.data_ptr(),
                                      shape_base.data_ptr(), win.data_ptr(), int(first_step), int(n_steps),
                                      lat.data_ptr(), m.data_ptr(), v.data_ptr(), ws.data_ptr(), ws.numel(),
                                      hist.data_ptr(), _stream_ptr(dev)), h, L)
    # keep every buffer the enqueued kernels use alive until they have run
    torch.cuda.current_stream(dev).synchronize()
    if return_state:
        return hist, lat, m, v
    return hist, lat


def reconstruct_sharded(decoder: Decoder, num_iterations: int, shapes: Sequence[Tuple[torch.Tensor, torch.Tensor]],
                        latents: torch.Tensor, windows: torch.Tensor, clamp_dist: float, rank: int, world_size: int,
                        **kw):
    """Shape shard of `reconstruct_batch` (the reference loops over the test split one shape at a time,
    reconstruct.py:312-420; shapes are independent): rank r optimises the contiguous block
    [S*r/W, S*(r+1)/W) of the global shape list on its own GPU -- `shapes`, `latents[S,256]` and
    `windows[num_iterations, S, 2]` are the GLOBAL arguments, identical on every rank.  Returns
    (lo, hi, loss_hist[num_iterations, hi-lo], latents[hi-lo, 256]); no collective (gather the blocks with
    all_gather_object / a file per rank if one process needs them all).  A shape's result does not depend on which
    rank, or which batch mates, it ran with: the blocks concatenated equal the single-GPU call bit for bit."""
    from .dist import owned_scenes
    own = owned_scenes(len(shapes), rank, world_size)
    lo, hi = own.start, own.stop
    if hi == lo:
        dev = next(decoder.parameters()).device
        return lo, hi, torch.zeros(num_iterations, 0, device=dev), torch.zeros(0, LATENT, device=dev)
    win = torch.as_tensor(windows, dtype=torch.int64).reshape(num_iterations, len(shapes), 2)[:, lo:hi]
    out = reconstruct_batch(decoder, num_iterations, list(shapes[lo:hi]), latents[lo:hi], win, clamp_dist, **kw)
    return (lo, hi) + tuple(out)


def reconstruct(decoder, num_iterations, latent_size, test_sdf, stat, clamp_dist, num_samples=30000,
                lr=5e-4, l2reg=False, code_reg_lambda=None, code_reg_type="l2_sq", code_bound=None,
                return_loss_hist=False, dist_mean=None, dist_std=None, dist_weight=0.0,
                dist_type="zscore_l2"):
    """Same signature and return values as reference reconstruct.reconstruct (reconstruct.py:16-151).
    `test_sdf` = [pos, neg] as returned by read_sdf_samples_into_ram (deep_sdf/data.py:83-88)."""
    if latent_size != LATENT:
        raise NotImplementedError(f"latent_size={latent_size}: fused path supports 256")
    device = next(decoder.parameters()).device
    if type(stat) == type(0.1):                                          # reconstruct.py:49-52
        latent = torch.ones(1, latent_size, device=device).normal_(mean=0, std=stat)
    else:
        latent = torch.normal(stat[0].detach(), stat[1].detach()).to(device)
    pos, neg = test_sdf[0], test_sdf[1]
    neg, wins = plan_samples(pos, neg, num_samples, num_iterations, random)
    hist, lat = reconstruct_batch(
        decoder, num_iterations, [(pos, neg)], latent.reshape(1, latent_size),
        torch.tensor(wins, dtype=torch.int64).reshape(num_iterations, 1, 2), clamp_dist,
        num_samples=num_samples, lr=lr, l2reg=l2reg, code_reg_lambda=code_reg_lambda,
        code_reg_type=code_reg_type, code_bound=code_bound, dist_mean=dist_mean, dist_std=dist_std,
        dist_weight=dist_weight, dist_type=dist_type)
    losses = hist[:, 0].cpu().numpy()                                    # one D2H at the end
    latent = lat.reshape(1, latent_size).requires_grad_(True)            # the reference returns a leaf
    if return_loss_hist:
        return [losses[i] for i in range(num_iterations)], latent
    return (losses[-1] if num_iterations > 0 else 0), latent

"""Scene-owner sharding for one-process-per-GPU training.

The reference's only parallelism is torch.nn.DataParallel (train_deep_sdf.py:352): GPU0 scatters
[M,259] inputs, broadcasts all parameters, gathers outputs and reduce-adds replica gradients every
step.  Here scenes are partitioned over ranks; a rank owns its scenes' latent codes, their Adam
state and their samples, so latent gradients never leave the rank and the only exchange per step is
one all-reduce(SUM) of the flat decoder gradient (~1.84 M fp32 = 7.4 MB) over NCCL/NVLink.
The loss is normalised by the GLOBAL row count (train_deep_sdf.py:558,606), hence SUM not MEAN.
"""
from __future__ import annotations

from typing import List, Tuple

import torch


def owned_scenes(num_scenes: int, rank: int, world_size: int) -> range:
    """Contiguous block partition: rank r owns scenes [r*S/W, (r+1)*S/W)."""
    lo = num_scenes * rank // world_size
    hi = num_scenes * (rank + 1) // world_size
    return range(lo, hi)


def owner_of(scene: int, num_scenes: int, world_size: int) -> int:
    # inverse of owned_scenes
    r = (scene * world_size) // num_scenes
    while scene < num_scenes * r // world_size:
        r -= 1
    while scene >= num_scenes * (r + 1) // world_size:
        r += 1
    return r


def local_batch(global_scenes_per_batch: int, rank: int, world_size: int) -> int:
    """ScenesPerBatch is a global quantity; rank r draws B/W of its own scenes per step."""
    lo = global_scenes_per_batch * rank // world_size
    hi = global_scenes_per_batch * (rank + 1) // world_size
    return hi - lo


def global_rows(local_rows: int, group=None) -> int:
    """Sum of per-rank row counts (the N_total every rank divides its loss by)."""
    if not torch.distributed.is_available() or not torch.distributed.is_initialized():
        return local_rows
    t = torch.tensor([local_rows], dtype=torch.int64)
    if torch.distributed.get_backend(group) == "nccl":
        t = t.cuda()
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.SUM, group=group)
    return int(t.item())


def allreduce_decoder_grads(flat_grad: torch.Tensor, group=None) -> None:
    """The one exchange step of training: SUM of the flat decoder gradient."""
    torch.distributed.all_reduce(flat_grad, op=torch.distributed.ReduceOp.SUM, group=group)


def gather_latents_for_checkpoint(local_rows: torch.Tensor, num_scenes: int, rank: int,
                                  world_size: int, group=None) -> torch.Tensor | None:
    """Reassemble the [num_scenes, L] latent table on rank 0 in scene order
    (for the reference's save_latent_vectors, train_deep_sdf.py:70-79)."""
    sizes = [len(owned_scenes(num_scenes, r, world_size)) for r in range(world_size)]
    mx = max(sizes)
    pad = torch.zeros(mx, local_rows.shape[1], dtype=local_rows.dtype, device=local_rows.device)
    pad[: local_rows.shape[0]] = local_rows
    bufs = [torch.empty_like(pad) for _ in range(world_size)]
    torch.distributed.all_gather(bufs, pad, group=group)
    if rank != 0:
        return None
    return torch.cat([b[:s] for b, s in zip(bufs, sizes)], 0)

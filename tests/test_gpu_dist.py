"""Multi-rank correctness of the scene-sharded fused step ON THE GPU.

Two processes, both on cuda:0 (the driver's GPU box for `pytest -m gpu` has one GPU), gloo process group over CUDA
tensors: every rank owns half of the scenes, runs FusedTrainer(process_group=...) on its own scenes' rows with the
GLOBAL row count as the loss normaliser, all-reduces (SUM) only the flat decoder gradient and applies the same
clip + Adam.  Checked against ONE process stepping on the union batch:
  * the all-reduced decoder gradient equals the union-batch gradient (cosine >= 0.999; in fact ~1e-6 apart: the same
    kernels, only the summation is partitioned differently),
  * each rank's latent gradients are bit-equal to the union run's rows of its scenes (a scene's latent gradient
    never depends on its batch mates and never leaves its owner),
  * after 3 optimiser steps the two replicas' decoder parameters are BIT-identical, and equal to the union run's
    to Adam-step accuracy.
Replaces the reference's nn.DataParallel (train_deep_sdf.py:352).
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist_
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
KW = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False,
          use_tanh=False, latent_dropout=False, weight_norm=True)
S, B, P, W, STEPS = 8, 4, 512, 2, 3


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _batches():
    g = torch.Generator().manual_seed(11)
    out = []
    for _ in range(STEPS):
        # B/W scenes from every rank's shard, as dist.local_batch prescribes
        per_rank = []
        for r in range(W):
            lo, hi = S * r // W, S * (r + 1) // W
            per_rank.append(lo + torch.randperm(hi - lo, generator=g)[:B // W])
        out.append(per_rank)
    return out


def _worker(rank, world, port, eik, out_dir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.join(here, ".."), os.path.join(here, "..", "medical-shape-disentanglement_b200")):
        sys.path.insert(0, p)
    import dsdf_b200
    from dsdf_b200 import dist, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist_.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, norm_layers=list(range(8)), **KW).cuda()       # replicas start identical
    mine = dist.owned_scenes(S, rank, world)
    lat_all = torch.randn(S, 256, generator=torch.Generator().manual_seed(4)) / 16.0
    cfg = dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, use_eikonal=eik)
    tr = dsdf_b200.FusedTrainer(dec, len(mine), cfg, process_group=dist_.group.WORLD, latents=lat_all[list(mine)])
    data = synth.scene_samples(S, P, seed=1).cuda()
    first = None
    for it, per_rank in enumerate(_batches()):
        gids = per_rank[rank]
        local = gids - mine[0]
        assert dist.global_rows(gids.numel() * P) == B * P
        if it == 0:
            tr.forward_backward(data[gids], local.cuda(), 60)              # n_total defaults to rows * world
            g = tr.grads.clone()
            dist.allreduce_decoder_grads(g)
            first = {"grads": g.cpu(), "lat_grad": tr.lat_grad.cpu().clone(), "loss": tr.loss_buf.cpu().clone()}
        st = tr.step(data[gids], local.cuda(), 60)
    torch.cuda.synchronize()
    tr.check_status()
    torch.save({"first": first, "params": dec.flat_params().cpu().clone(), "lat": tr.lat_vecs.weight.data.cpu().clone(),
                "loss": tr.loss_buf.cpu().clone()}, os.path.join(out_dir, f"r{rank}.pt"))
    dist_.destroy_process_group()


@pytest.mark.parametrize("eik", [False, True])
def test_two_rank_sharded_trainer_equals_the_union_batch_step(eik, tmp_path):
    import dsdf_b200
    from common import cos
    from dsdf_b200 import synth
    mp.spawn(_worker, args=(W, _free_port(), eik, str(tmp_path)), nprocs=W, join=True)
    r = [torch.load(os.path.join(tmp_path, f"r{i}.pt")) for i in range(W)]
    # ---- replicas: identical summed gradient, bit-identical parameters after 3 steps
    assert torch.equal(r[0]["first"]["grads"], r[1]["first"]["grads"])
    assert torch.equal(r[0]["params"], r[1]["params"])
    # ---- one process, union batch
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, norm_layers=list(range(8)), **KW).cuda()
    lat_all = torch.randn(S, 256, generator=torch.Generator().manual_seed(4)) / 16.0
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, use_eikonal=eik), latents=lat_all)
    data = synth.scene_samples(S, P, seed=1).cuda()
    for it, per_rank in enumerate(_batches()):
        ids = torch.cat(per_rank)
        if it == 0:
            tr.forward_backward(data[ids], ids.cuda(), 60)
            torch.cuda.synchronize()
            g = tr.grads.cpu()
            assert cos(g, r[0]["first"]["grads"]) >= 0.999
            assert float((g - r[0]["first"]["grads"]).norm() / g.norm()) <= 1e-4
            lg = tr.lat_grad.cpu()
            half = S // W
            assert torch.equal(lg[:half], r[0]["first"]["lat_grad"]) and torch.equal(lg[half:], r[1]["first"]["lat_grad"])
            tot = r[0]["first"]["loss"] + r[1]["first"]["loss"]              # every rank reports its share of the global loss
            assert tot[0].item() == pytest.approx(tr.loss_buf[0].item(), rel=1e-5)
            assert tot[2].item() == pytest.approx(tr.loss_buf[2].item(), rel=1e-5, abs=1e-12)
        tr.step(data[ids], ids.cuda(), 60)
    torch.cuda.synchronize()
    # Adam moves every weight by <= lr per step; a sign flip of a ~zero gradient bounds the difference by 2 lr per step
    assert (dec.flat_params().cpu() - r[0]["params"]).abs().max().item() <= 2.2 * 5e-4 * STEPS
    assert float((dec.flat_params().cpu() - r[0]["params"]).norm() / (5e-4 * STEPS * dec.flat_params().numel() ** 0.5)) <= 0.05
    lat = torch.cat([r[0]["lat"], r[1]["lat"]])
    assert (tr.lat_vecs.weight.data.cpu() - lat).abs().max().item() <= 2.2 * 1e-3 * STEPS

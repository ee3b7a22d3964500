"""Parity at BASELINE.json's FULL sizes (configs[1]: 64 scenes x 16384 samples = 1 048 576 rows per step, 800-scene
latent table; create_mesh N = 256; reconstruct 800 iterations x 8000 samples).  Two kinds of checks:

* size-independent properties of the CUDA path alone: run-to-run bit determinism (no float atomics anywhere),
  additivity of the step over scene sub-batches (loss and decoder gradient of the 64-scene batch = sum over its two
  32-scene halves at the same global normaliser; latent gradients of a scene do not depend on its batch mates),
  voxel-slab shards reproducing the single launch bit for bit;
* the oracle itself at the full size, run in fp32 torch on the GPU (it needs tens of GB of autograd state, which a
  180 GB board has): same tolerances as the small cases -- SDF <= 2e-3, gradient cosine >= 0.999, loss <= 1 %.
"""
import random

import numpy as np
import pytest
import torch

import dsdf_b200
from common import cos, spec_for
from dsdf_b200 import synth
from oracle import mesh_grid
from oracle import reconstruct as orec
from oracle import train as otrain

pytestmark = pytest.mark.gpu
KW = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False,
          use_tanh=False, latent_dropout=False, weight_norm=True)
S, B, P = 800, 64, 16384


def make(seed=0, scale=1.5):
    torch.manual_seed(seed)
    dec = dsdf_b200.Decoder(256, norm_layers=list(range(8)), **KW)
    params = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    for k in params:
        if k.endswith("weight_g"):
            params[k] *= scale
    dec.load_state_dict(params)
    return dec.cuda(), spec_for(True), params


def batch():
    data = synth.fast_scene_samples(B, P, seed=3, device="cuda")          # [64, 16384, 4]
    ids = torch.randperm(S, generator=torch.Generator().manual_seed(1))[:B].sort().values
    return data, ids


def test_full_size_step_matches_the_oracle_run_on_the_gpu():
    dec, spec, params = make()
    lat0 = torch.randn(S, 256, generator=torch.Generator().manual_seed(4)) / 16.0
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0)
    orc = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params, latents=lat0)
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0), latents=lat0)
    data, ids = batch()
    orc.set_epoch_lr(60); orc.opt.zero_grad()
    sl, rl, pred_ref = orc.forward_loss(data, ids, 60)
    pred = tr.forward_backward(data, ids.cuda(), 60)
    torch.cuda.synchronize()
    assert pred.numel() == B * P == 1048576
    assert abs(tr.loss_buf[0].item() - sl) <= 1e-2 * abs(sl)
    assert abs(tr.loss_buf[1].item() - rl) <= 1e-4 * abs(rl) + 1e-12
    assert (pred.clamp(-0.1, 0.1)[:, None] - pred_ref).abs().max().item() <= 2e-3
    ref_flat = torch.cat([orc.params[n].grad.flatten() for n, _ in dec.named_parameters()])
    assert cos(tr.grads, ref_flat) >= 0.999
    lg = orc.lat_vecs.weight.grad
    assert cos(tr.lat_grad, lg) >= 0.999
    assert torch.equal(tr.lat_grad != 0, lg != 0)                 # exactly the 64 gathered rows carry gradient
    assert int((tr.lat_grad != 0).any(dim=1).sum()) == B


def test_full_size_step_is_deterministic_and_additive_over_scene_halves():
    dec, _, _ = make()
    lat0 = torch.randn(S, 256, generator=torch.Generator().manual_seed(4)) / 16.0
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0), latents=lat0)
    data, ids = batch()
    idc = ids.cuda()
    n_total = B * P

    def run(lo, hi):
        pred = tr.forward_backward(data[lo:hi], idc[lo:hi], 60, n_total=n_total)
        torch.cuda.synchronize()
        return pred.clone(), tr.grads.clone(), tr.lat_grad.clone(), tr.loss_buf[:2].clone()

    p1, g1, l1, s1 = run(0, B)
    p2, g2, l2, s2 = run(0, B)
    assert torch.equal(p1, p2) and torch.equal(g1, g2) and torch.equal(l1, l2) and torch.equal(s1, s2)
    pa, ga, la, sa = run(0, B // 2)
    pb, gb, lb, sb = run(B // 2, B)
    assert torch.equal(torch.cat([pa, pb]), p1)                   # a row's SDF does not depend on its batch mates
    torch.testing.assert_close(sa + sb, s1, rtol=1e-5, atol=0)
    assert torch.equal(la + lb, l1)                               # per-scene segmented reduction: fixed order
    assert cos(ga + gb, g1) >= 0.999999
    assert ((ga + gb) - g1).norm().item() <= 1e-4 * g1.norm().item()


def test_grid_decode_256_properties_and_sampled_oracle():
    dec, spec, params = make()
    lat = torch.randn(1, 256, generator=torch.Generator().manual_seed(5)) * 0.05
    N = 256
    out = dsdf_b200.create_mesh_sdf(dec, lat, N)
    assert out.shape == (N, N, N) and torch.isfinite(out).all()
    again = dsdf_b200.create_mesh_sdf(dec, lat, N)
    assert torch.equal(out, again)
    parts = [dsdf_b200.decode_grid_sharded(dec, lat, N, r, 8) for r in range(8)]
    assert parts[0][0] == 0 and parts[-1][1] == N ** 3
    assert torch.equal(torch.cat([p[2] for p in parts]).cpu(), out.reshape(-1))
    ref = mesh_grid.decode_grid(params, spec, lat, N, device="cuda")      # all 16.7 M voxels through the oracle
    assert (out.reshape(-1) - ref.cpu().reshape(-1)).abs().max().item() <= 2e-3


def test_reconstruct_full_schedule_follows_the_oracle():
    """reconstruct.py's shipped schedule (800 iterations, 8000 samples, lr 5e-3, l2reg) end to end."""
    dec, spec, params = make()
    d = synth.scene_samples(1, 32768, seed=11)[0]
    pos, neg = d[:16384].contiguous(), d[16384:].contiguous()
    iters, num = 800, 8000
    wins = orec.draw_windows(pos.shape[0], neg.shape[0], num, iters, random.Random(2))
    lat0 = torch.ones(1, 256).normal_(mean=0, std=0.01, generator=torch.Generator().manual_seed(6))
    losses, lat_ref = orec.reconstruct(params, spec, iters, pos, neg, wins, lat0, 0.1, num_samples=num, lr=5e-3,
                                       device="cuda", l2reg=True)
    hist, lat = dsdf_b200.reconstruct_batch(dec, iters, [(pos, neg)], lat0, torch.tensor(wins).view(iters, 1, 2), 0.1,
                                            num_samples=num, lr=5e-3, l2reg=True)
    got = hist[:, 0].cpu().numpy()
    ref = np.array(losses)
    assert got.shape == (iters,)
    np.testing.assert_allclose(got, ref, rtol=1e-2)
    assert got[-20:].mean() < got[:20].mean()
    # 800 Adam steps: the per-step loss stays within 1 % the whole way, but Adam normalises every coordinate's
    # step to ~lr, so bf16-level gradient differences in the flat directions of the code space accumulate and the
    # two final codes are only loosely aligned (measured cosine 0.80).  What the caller consumes is the shape the
    # code decodes to: score both final codes with the ORACLE decoder on all 32768 samples.
    assert cos(lat.cpu() - lat0, lat_ref - lat0) >= 0.6
    pc = {k: v.cuda() for k, v in params.items()}
    gt = d[:, 3:4].cuda().clamp(-0.1, 0.1)

    def score(code):
        with torch.no_grad():
            return (mesh_grid.decode_sdf(pc, spec, code.cuda(), d[:, :3].cuda()).clamp(-0.1, 0.1) - gt).abs().mean().item()
    assert score(lat) == pytest.approx(score(lat_ref), rel=2e-2)

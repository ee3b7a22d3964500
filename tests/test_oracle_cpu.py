"""CPU tests of the oracle: against the committed golden fixtures (generated from the live
reference by tests/golden/make_golden.py) and, where /root/reference is mounted, against the
live reference modules themselves."""
import math
import os
import sys
import warnings

import numpy as np
import pytest
import torch

from common import REF_DIR, check_checksums, cos, golden, spec_for
from oracle import deepsdf, mesh_grid
from oracle import reconstruct as orec
from oracle import train as otrain

warnings.simplefilter("ignore")
have_ref = os.path.isdir(os.path.join(REF_DIR, "networks"))


@pytest.mark.parametrize("wn", [False, True])
def test_layer_shapes_and_param_count(wn):
    spec = spec_for(wn)
    assert deepsdf.layer_shapes(spec) == [(259, 512), (512, 512), (512, 512), (512, 253),
                                          (512, 512), (512, 512), (512, 512), (512, 512), (512, 1)]
    assert deepsdf.param_count(spec) == (1843195 if wn else 1839358)
    assert deepsdf.macs_per_row(spec) == 1835520


@pytest.mark.parametrize("wn,name", [(False, "decoder_plain.npz"), (True, "decoder_wn.npz")])
def test_forward_backward_matches_golden(wn, name):
    g = golden(name)
    spec = spec_for(wn)
    p = deepsdf.init_params(spec, seed=0)
    check_checksums(p, g)
    p = {k: v.requires_grad_(True) for k, v in p.items()}
    x = torch.from_numpy(g["input"]).requires_grad_(True)
    y = deepsdf.forward(p, spec, x)
    np.testing.assert_allclose(y.detach().numpy(), g["output"], atol=2e-6)
    (y * torch.from_numpy(g["w"])).sum().backward()
    np.testing.assert_allclose(x.grad.numpy(), g["dinput"], atol=1e-6, rtol=1e-4)
    for k, v in p.items():
        np.testing.assert_allclose(float(v.grad.norm()), float(g["gradnorm/" + k]), rtol=1e-4)
    for k in ("lin8.bias", "lin7.bias", "lin0.bias"):
        np.testing.assert_allclose(p[k].grad.numpy(), g["grad/" + k], rtol=1e-3, atol=1e-6)


@pytest.mark.parametrize("wn,name", [(True, "train_wn.npz"), (False, "train_plain.npz")])
def test_train_trajectory_matches_golden(wn, name):
    from dsdf_b200 import synth
    g = golden(name)
    S, B, P = int(g["S"]), int(g["B"]), int(g["P"])
    spec = spec_for(wn)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0)
    tr = otrain.OracleTrainer(spec, cfg, S, seed=0, latents=torch.from_numpy(g["lat0"]))
    data = synth.scene_samples(S, P, seed=1)
    steps = g["ids"].shape[0]
    for it in range(steps):
        ids = torch.from_numpy(g["ids"][it])
        st = tr.step(data[ids], ids, 50 + it)
        assert st["sdf_loss"] == pytest.approx(float(g["loss"][it]), rel=2e-4)
        assert st["reg_loss"] == pytest.approx(float(g["reg"][it]), rel=2e-4)
        assert st["grad_norm"] == pytest.approx(float(g["gnorm"][it]), rel=2e-3)
        if it == 0:
            np.testing.assert_allclose(st["pred"].numpy(), g["pred0"], atol=2e-6)
            np.testing.assert_allclose(tr.latent_grad().numpy(), g["latgrad0"], atol=1e-8, rtol=1e-3)
            # bookkeeping: exactly the batch rows have gradient, exactly row 1 was renormed
            assert np.array_equal(tr.latent_grad().numpy() != 0, g["latgrad0"] != 0)
    np.testing.assert_allclose(tr.lat_vecs.weight.data.numpy(), g["lat_final"], atol=2e-4)


def test_grid_closed_form_matches_golden_and_reference_ops():
    g = golden("grid.npz")
    np.testing.assert_array_equal(mesh_grid.grid_coords_closed_form(8), g["coords8"])
    for N in (8, 64, 256):
        xor = np.zeros(3, dtype=np.uint64)
        tot = np.zeros(3, dtype=np.uint64)
        step = 1 << 22
        for lo in range(0, N ** 3, step):
            hi = min(N ** 3, lo + step)
            c = mesh_grid.grid_coords_closed_form(N, lo, hi)
            bits = c.view(np.uint32).astype(np.uint64)
            xor ^= np.bitwise_xor.reduce(bits, axis=0)
            tot += bits.sum(axis=0)
        np.testing.assert_array_equal(xor, g[f"xor{N}"])
        np.testing.assert_array_equal(tot, g[f"sum{N}"])
        stride = max(1, N ** 3 // 4096)
        np.testing.assert_array_equal(mesh_grid.grid_coords_closed_form(N)[::stride], g[f"sample{N}"])
    a = mesh_grid.grid_coords_reference_ops(100, 1000, 50000).numpy()
    np.testing.assert_array_equal(a, mesh_grid.grid_coords_closed_form(100, 1000, 50000))


def test_restated_optimizer_math_matches_torch():
    torch.manual_seed(0)
    p = torch.randn(1000); g = torch.randn(1000) * 0.1
    q = torch.nn.Parameter(p.clone()); opt = torch.optim.Adam([q], lr=1e-3)
    m = torch.zeros(1000); v = torch.zeros(1000); pp = p.clone()
    for step in range(1, 6):
        q.grad = g.clone() * step
        opt.step()
        otrain.adam_update(pp, g * step, m, v, step, 1e-3)
    np.testing.assert_allclose(pp.numpy(), q.data.numpy(), atol=1e-6)
    # clip coefficient
    w = torch.nn.Parameter(torch.ones(10)); w.grad = torch.full((10,), 3.0)
    tn = w.grad.norm().clone()
    torch.nn.utils.clip_grad_norm_([w], 1.0)
    np.testing.assert_allclose(w.grad.numpy(), (3.0 * otrain.clip_coef(tn, 1.0)).item(), rtol=1e-6)
    # embedding renorm side effect
    emb = torch.nn.Embedding(5, 8, max_norm=1.0)
    emb.weight.data.normal_(0, 1.0)
    t = emb.weight.data.clone()
    ids = torch.tensor([3, 1, 3])
    emb(ids)
    otrain.renorm_rows(t, ids, 1.0)
    np.testing.assert_allclose(t.numpy(), emb.weight.data.numpy(), atol=1e-7)
    assert torch.equal(t[[0, 2, 4]], emb.weight.data[[0, 2, 4]])


def test_reconstruct_oracle_runs_and_descends():
    import random
    from dsdf_b200 import synth
    spec = spec_for(True)
    p = deepsdf.init_params(spec, seed=0)
    d = synth.scene_samples(1, 2048, seed=3)[0]
    pos, neg = d[:1024], d[1024:]
    wins = orec.draw_windows(1024, 1024, 256, 12, random.Random(0))
    lat0 = torch.randn(1, 256, generator=torch.Generator().manual_seed(1)) * 0.01
    losses, lat = orec.reconstruct(p, spec, 12, pos, neg, wins, lat0, 0.1, num_samples=256, lr=5e-3)
    assert len(losses) == 12 and lat.shape == (1, 256) and all(math.isfinite(x) for x in losses)


@pytest.mark.skipif(not have_ref, reason="reference tree not mounted")
@pytest.mark.parametrize("wn", [False, True])
def test_oracle_equals_live_reference(wn):
    sys.path.insert(0, REF_DIR)
    from networks.deep_sdf_decoder import Decoder as RefDecoder
    spec = spec_for(wn)
    torch.manual_seed(0)
    ref = RefDecoder(256, **spec.network_specs())
    p = deepsdf.init_params(spec, seed=0)
    sd = ref.state_dict()
    assert list(sd.keys()) == list(p.keys())
    for k in sd:
        assert torch.equal(sd[k], p[k]), k
    x = torch.randn(300, 259, generator=torch.Generator().manual_seed(2)) * 0.3
    assert torch.equal(ref(x), deepsdf.forward(p, spec, x))
    ref.train()
    assert torch.equal(ref(x), deepsdf.forward(p, spec, x, training=True))

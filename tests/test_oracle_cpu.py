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


def test_philox4x32_10_known_answer_vectors():
    """Random123's published KATs for philox4x32-10 pin the oracle of the dropout stream."""
    from oracle import philox
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for c, k, want in kat:
        got = philox.philox4x32_10(np.array(c, dtype=np.uint32), np.array(k, dtype=np.uint32))
        assert got.tolist() == list(want)
    keep = philox.dropout_keep(seed=7, offset=3, layer=2, row_lo=1000, n_rows=512, prob=0.2)
    assert keep.shape == (512, 512) and abs(keep.mean() - 0.8) < 5e-3
    assert np.array_equal(keep[10:20], philox.dropout_keep(7, 3, 2, 1010, 10, 0.2))   # pure function of the row
    assert not np.array_equal(keep, philox.dropout_keep(7, 4, 2, 1000, 512, 0.2))


def test_oracle_dropout_with_explicit_mask_equals_f_dropout_arithmetic():
    spec = deepsdf.DecoderSpec(norm_layers=tuple(range(8)), weight_norm=True, dropout=(0, 3, 7), dropout_prob=0.2)
    p = deepsdf.init_params(spec, seed=0)
    x = torch.randn(64, 259, generator=torch.Generator().manual_seed(1)) * 0.3
    ones = {l: torch.ones(64, 253 if l == 3 else 512) for l in (0, 3, 7)}
    a = deepsdf.forward(p, spec, x, training=True, dropout_keep=ones)
    spec_scaled = deepsdf.DecoderSpec(norm_layers=tuple(range(8)), weight_norm=True)
    # all-kept mask = plain network with activations of layers 0,3,7 scaled by 1/(1-p)
    h = x
    for l in range(9):
        if l == 4:
            h = torch.cat([h, x], 1)
        h = torch.nn.functional.linear(h, deepsdf.effective_weight(p, spec_scaled, l), p[f"lin{l}.bias"])
        if l < 8:
            h = torch.relu(h)
            if l in (0, 3, 7):
                h = h * (1.0 / (1.0 - 0.2))
    np.testing.assert_allclose(a.numpy(), torch.tanh(h).numpy(), atol=1e-6)


def test_siren_oracle_matches_golden_from_the_live_reference():
    from oracle import siren
    g = golden("decoder_siren.npz")
    p = siren.init_params(seed=0)
    check_checksums(p, g)
    assert sum(v.numel() for v in p.values()) == 1_840_897 and siren.macs_per_row() == 1_837_056
    x = torch.from_numpy(g["input"]).requires_grad_(True)
    pr = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    y = siren.forward(pr, x)
    (y * torch.from_numpy(g["w"])).sum().backward()
    np.testing.assert_allclose(y.detach().numpy(), g["output"], atol=1e-6)
    np.testing.assert_allclose(x.grad.numpy(), g["dinput"], atol=1e-6, rtol=1e-4)
    for k in ("decoder.lin8.weight", "decoder.lin4.bias", "decoder.lin0.bias"):
        np.testing.assert_allclose(pr[k].grad.numpy(), g["grad/" + k], atol=1e-6, rtol=1e-4)
    for k, v in pr.items():
        assert float(v.grad.norm()) == pytest.approx(float(g["gradnorm/" + k]), rel=1e-4)


@pytest.mark.skipif(not have_ref, reason="reference tree not mounted")
def test_siren_oracle_and_host_decoder_equal_live_reference():
    sys.path.insert(0, REF_DIR)
    from networks.siren_decoder import Decoder as RefSiren
    from oracle import siren
    import dsdf_b200.siren as S
    torch.manual_seed(0)
    ref = RefSiren(256, **siren.network_specs())
    p = siren.init_params(seed=0)
    sd = ref.state_dict()
    assert list(sd.keys()) == list(p.keys()) and all(torch.equal(sd[k], p[k]) for k in sd)
    x = torch.randn(200, 259, generator=torch.Generator().manual_seed(2)) * 0.3
    assert torch.equal(ref(x), siren.forward(p, x))
    torch.manual_seed(0)
    mine = S.Decoder(256, **siren.network_specs())
    sd2 = mine.state_dict()
    assert list(sd2.keys()) == list(sd.keys()) and all(torch.equal(sd2[k], sd[k]) for k in sd)
    mine.load_state_dict(sd)            # reference checkpoints load unchanged
    for kw in (dict(nonlinearity="relu"), dict(xyz_in=[2]), dict(encoding_features=8), dict(use_tanh=True)):
        spec = dict(siren.network_specs()); spec.update(kw)
        with pytest.raises(NotImplementedError):
            S.Decoder(256, **spec)


def _callers_setup():
    """Decoder parameters, shape and RNG replay of tests/golden/make_golden_callers.py."""
    import random
    from dsdf_b200 import synth
    g = golden("callers_ref.npz")
    spec = spec_for(True)
    p = deepsdf.init_params(spec, seed=0)
    for l in range(8):
        p[f"lin{l}.weight_g"] = p[f"lin{l}.weight_g"] * 1.5
    check_checksums(p, g)
    d = synth.scene_samples(1, 4096, seed=9)[0]
    return g, spec, p, d[:2048].contiguous(), d[2048:].contiguous(), random


def test_oracle_callers_match_the_executed_reference():
    """callers_ref.npz was produced by RUNNING the reference's reconstruct.reconstruct, data.unpack_sdf_samples_from_ram
    and mesh.create_mesh (sampling loop) on CPU: the oracle restatements must reproduce them."""
    g, spec, p, pos, neg, random = _callers_setup()
    # sampler (deep_sdf/data.py:112-136): same random.randint stream -> identical rows
    random.seed(3)
    w = orec.draw_windows(2048, 2048, 800, 1, random)[0]
    np.testing.assert_array_equal(orec.window(pos, neg, 800, w).numpy(), g["window_sample"])
    # reconstruct (reconstruct.py:16-151): initial latent from torch's RNG, then two randints per iteration
    for name, kw in (("a", dict(l2reg=True)), ("b", dict(l2reg=False, code_reg_lambda=1e-2, code_reg_type="l2_norm", code_bound=0.2))):
        random.seed(0); torch.manual_seed(0)
        lat0 = torch.ones(1, 256).normal_(mean=0, std=0.01)
        wins = orec.draw_windows(2048, 2048, 800, 30, random)
        losses, lat = orec.reconstruct(p, spec, 30, pos, neg, wins, lat0, 0.1, num_samples=800, lr=5e-3, **kw)
        np.testing.assert_allclose(np.array(losses), g[f"recon_{name}/losses"], rtol=2e-5, atol=1e-7)
        np.testing.assert_allclose(lat.numpy(), g[f"recon_{name}/latent"], atol=2e-5)
    # create_mesh sampling (mesh.py:32-71): closed-form grid + decode_sdf
    lat = torch.from_numpy(g["recon_a/latent"])
    coords = torch.from_numpy(mesh_grid.grid_coords_closed_form(24))
    sdf = deepsdf.forward(p, spec, torch.cat([lat.expand(24 ** 3, -1), coords], 1)).reshape(24, 24, 24)
    np.testing.assert_allclose(sdf.numpy(), g["mesh/sdf24"], atol=2e-6)


def test_oracle_small_shape_sampler_branch_matches_the_executed_reference():
    """callers_fallback_ref.npz: the reference's unpack_sdf_samples_from_ram / reconstruct.reconstruct RUN on a shape
    with 300 negatives < num_samples/2 = 400 (data.py:123-125: torch.rand + index_select every iteration)."""
    _, spec, p, pos, neg, random = _callers_setup()
    g = golden("callers_fallback_ref.npz")
    neg = neg[:300].contiguous()
    random.seed(3); torch.manual_seed(3)
    pool, wins = orec.draw_plan(pos, neg, 800, 3, random)
    assert pool.shape == (1200, 4)
    for k in range(3):
        np.testing.assert_array_equal(orec.window(pos, pool, 800, wins[k]).numpy(), g[f"sample_{k}"])
    random.seed(0); torch.manual_seed(0)
    lat0 = torch.ones(1, 256).normal_(mean=0, std=0.01)
    pool, wins = orec.draw_plan(pos, neg, 800, 20, random)
    losses, lat = orec.reconstruct(p, spec, 20, pos, pool, wins, lat0, 0.1, num_samples=800, lr=5e-3, l2reg=True)
    np.testing.assert_allclose(np.array(losses), g["recon_c/losses"], rtol=2e-5, atol=1e-7)
    np.testing.assert_allclose(lat.numpy(), g["recon_c/latent"], atol=2e-5)


@pytest.mark.parametrize("wn,name", [(False, "train_eikonal_plain.npz"), (True, "train_eikonal_wn.npz")])
def test_oracle_eikonal_step_matches_golden_from_the_live_reference(wn, name):
    """The oracle's UseEikonal branch against the fixture made by driving the live reference Decoder with
    train_deep_sdf.py:724-729 verbatim (tests/golden/make_golden_eikonal.py), with and without the term."""
    from oracle import train as otrain
    from dsdf_b200 import synth
    from common import eikonal_golden_check
    g = golden(name)
    S, B, P = int(g["S"]), int(g["B"]), int(g["P"])
    data = synth.scene_samples(S, P, seed=1)
    ids = torch.from_numpy(g["ids"])
    for use, prefix in ((True, "eik/"), (False, "base/")):
        cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0, use_eikonal=use)
        orc = otrain.OracleTrainer(spec_for(wn), cfg, S, seed=0, latents=torch.from_numpy(g["lat0"]))
        orc.opt.zero_grad()
        sl, rl, pred = orc.forward_loss(data[ids], ids, int(g["epoch"]))
        assert sl == pytest.approx(float(g[prefix + "sdf_loss"]), rel=1e-6)
        assert rl == pytest.approx(float(g[prefix + "reg_loss"]), rel=1e-6)
        assert orc.last_eikonal_loss == pytest.approx(float(g[prefix + "eikonal_loss"]), rel=1e-6, abs=1e-12)
        np.testing.assert_allclose(pred.numpy(), g[prefix + "pred"], atol=1e-7)
        eikonal_golden_check(g, orc.decoder_grads(), orc.latent_grad(), prefix, rtol_norm=1e-4, min_cos=0.999999)
    # the term moves the gradient: the two arms differ measurably
    assert float(g["eik/gnorm/lin7.weight_v" if wn else "eik/gnorm/lin7.weight"]) != \
        float(g["base/gnorm/lin7.weight_v" if wn else "base/gnorm/lin7.weight"])


def test_fused_eikonal_algebra_equals_double_backward():
    """The single-backward algebra the CUDA kernels implement (tests/eikonal_algebra.py) against torch's
    create_graph=True double backward of the oracle step, in fp32 on CPU: same losses, same gradients."""
    from oracle import train as otrain
    from dsdf_b200 import synth
    from eikonal_algebra import fused_eikonal_grads
    S, B, P = 4, 2, 256
    spec = spec_for(False)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=None, grad_clip=None,
                             code_reg=False, use_eikonal=True)
    params = deepsdf.init_params(spec, seed=3)
    # a few "training-like" perturbations so that ||G|| is not ~0 everywhere and the clamp is active on some rows
    g = torch.Generator().manual_seed(5)
    for k in params:
        if k.endswith("weight"):
            params[k] = params[k] * 2.4
    orc = otrain.OracleTrainer(spec, cfg, S, seed=0, params=params)
    data = synth.scene_samples(S, P, seed=2)
    ids = torch.tensor([3, 1])
    orc.opt.zero_grad()
    sl, rl, pred = orc.forward_loss(data[ids], ids, 50)
    W = [orc.params[f"lin{l}.weight"].detach() for l in range(9)]
    b = [orc.params[f"lin{l}.bias"].detach() for l in range(9)]
    sdf = data[ids].reshape(-1, 4)
    idx = ids.unsqueeze(-1).repeat(1, P).view(-1)
    lat_rows = orc.lat_vecs.weight.data[idx]
    out = fused_eikonal_grads(W, b, lat_rows, sdf[:, :3], sdf[:, 3:4].clamp(-0.1, 0.1), 0.1, 0.002, sdf.shape[0])
    assert float(out["sdf_loss"]) == pytest.approx(sl, rel=1e-5)
    assert float(out["eik_loss"]) == pytest.approx(orc.last_eikonal_loss, rel=1e-4)
    inside = (pred.abs() < 0.1).float().mean().item()
    assert 0.05 < inside < 1.0                      # both clamp branches are exercised
    for l in range(9):
        gw = orc.params[f"lin{l}.weight"].grad
        gb = orc.params[f"lin{l}.bias"].grad
        assert cos(out["dW"][l], gw) > 0.99999, l
        assert float((out["dW"][l] - gw).norm() / gw.norm()) < 2e-3, l
        assert cos(out["db"][l], gb) > 0.99999, l
    lat_ref = orc.latent_grad()
    dlat = torch.zeros_like(lat_ref).index_add_(0, idx, out["dlat_rows"])
    assert cos(dlat, lat_ref) > 0.99999
    # and the Eikonal part on its own (difference to the plain step) is reproduced, not just swamped by the L1 part
    cfg0 = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=None, grad_clip=None,
                              code_reg=False, use_eikonal=False)
    orc0 = otrain.OracleTrainer(spec, cfg0, S, seed=0, params=params)
    orc0.opt.zero_grad()
    orc0.forward_loss(data[ids], ids, 50)
    for l in (0, 2, 4, 7, 8):
        d_ref = orc.params[f"lin{l}.weight"].grad - orc0.params[f"lin{l}.weight"].grad
        d_got = out["dW"][l] - orc0.params[f"lin{l}.weight"].grad
        assert float(d_ref.norm()) > 0
        assert cos(d_got, d_ref) > 0.999, (l, cos(d_got, d_ref))


@pytest.mark.parametrize("eik", [False, True])
def test_oracle_port_equals_the_step_loop_around_the_reference_module(eik):
    """oracle/ref_step.py runs train_deep_sdf.py's step statements around the reference's own Decoder class
    (oracle/_ref, copied unmodified by oracle/build_ref.py); the oracle port (oracle/train.py, functional restatement)
    must walk the same trajectory: losses and gradient norm of 3 optimiser steps, then the parameters."""
    from oracle import ref_step
    from oracle import train as otrain
    from dsdf_b200 import synth
    if not ref_step.available():
        pytest.skip("oracle/_ref not built (python oracle/build_ref.py needs /root/reference)")
    S, B, P = 4, 2, 256
    specs = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, norm_layers=[], latent_in=[4], xyz_in_all=False,
                 use_tanh=False, latent_dropout=False, weight_norm=True)
    ref = ref_step.ReferenceStepLoop(S, specs, use_eikonal=eik, seed=0)
    lat0 = ref.lat_vecs.weight.data.clone()
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0, use_eikonal=eik)
    orc = otrain.OracleTrainer(spec_for(False), cfg, S, seed=0, latents=lat0)
    data = synth.scene_samples(S, P, seed=1)
    g = torch.Generator().manual_seed(5)
    for it in range(3):
        ids = torch.randperm(S, generator=g)[:B]
        a = ref.step(data[ids].clone(), ids, 50 + it)
        b = orc.step(data[ids], ids, 50 + it)
        assert a["sdf_loss"] == pytest.approx(b["sdf_loss"], rel=1e-6)
        assert a["reg_loss"] == pytest.approx(b["reg_loss"], rel=1e-6)
        assert a["grad_norm"] == pytest.approx(b["grad_norm"], rel=1e-5)
        if eik:
            assert a["eikonal_loss"] == pytest.approx(b["eikonal_loss"], rel=1e-6)
    for k, v in ref.decoder.state_dict().items():
        torch.testing.assert_close(v, orc.params[k].data, rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(ref.lat_vecs.weight.data, orc.lat_vecs.weight.data, rtol=1e-5, atol=1e-7)


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference checkout exists in the build container only")
@pytest.mark.parametrize("eik", [False, True])
def test_oracle_step_equals_the_executed_reference_training_script(eik):
    """SURVEY.md 8(c) / VERDICT r1 weak #6: the oracle's train step is no longer pinned to torch primitives only -- the
    reference's OWN train_deep_sdf.main_function (train_deep_sdf.py:211-1123) is executed here, unmodified, on CPU for three
    epochs of a tiny experiment (tests/golden/make_golden_trainloop.py: import stubs, .cuda() -> identity, recording
    DataLoader / Adam / SummaryWriter; none touches the arithmetic).  The oracle, started from the script's initial
    decoder and latent table and fed the very batches the script's DataLoader yielded, must reproduce the script's
    per-batch loss, its per-epoch loss terms and its final checkpoint (decoder + latent codes)."""
    import importlib.util
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "make_golden_trainloop.py")
    spec_ = importlib.util.spec_from_file_location("make_golden_trainloop", path)
    mod = importlib.util.module_from_spec(spec_)
    spec_.loader.exec_module(mod)
    r = mod.run_reference(eik)
    sp = r["specs"]
    sch = sp["LearningRateSchedule"]
    cfg = otrain.TrainConfig(samples_per_scene=sp["SamplesPerScene"], scenes_per_batch=sp["ScenesPerBatch"],
                             clamp_dist=sp["ClampingDistance"], code_reg=sp["CodeRegularization"],
                             code_reg_lambda=sp["CodeRegularizationLambda"], code_bound=sp["CodeBound"],
                             grad_clip=sp["GradientClipNorm"], lr_decoder=sch[0]["Initial"], lr_latent=sch[1]["Initial"],
                             lr_interval=sch[0]["Interval"], lr_factor=sch[0]["Factor"], use_eikonal=eik)
    names = [k[len("module."):] if k.startswith("module.") else k for k in r["final_model"].keys()]
    # Adam's first parameter group = decoder.parameters() in registration order = state_dict order (no buffers)
    assert len(names) == len(r["init_decoder"])
    params = {n: p.clone() for n, p in zip(names, r["init_decoder"])}
    S = r["init_latents"].shape[0]
    orc = otrain.OracleTrainer(deepsdf.DecoderSpec(norm_layers=tuple(range(8)), weight_norm=True), cfg, S, seed=0, device="cpu",
                               params=params, latents=r["init_latents"].clone())
    per_batch_ref = [v for t, v, _ in r["scalars"] if t == "Loss/train_vanilla"]
    assert len(per_batch_ref) == len(r["batches"]) == r["epochs"] * r["batches_per_epoch"]
    ours, eik_terms, reg_terms = [], [], []
    for i, (sdf_data, indices) in enumerate(r["batches"]):
        epoch = 1 + i // r["batches_per_epoch"]
        st = orc.step(sdf_data, indices, epoch)
        ours.append(st["sdf_loss"] + st["reg_loss"])      # "Loss/train_vanilla" is logged before the Eikonal term is added (:723)
        eik_terms.append(st["eikonal_loss"])
        reg_terms.append(st["reg_loss"])
    assert np.allclose(ours, per_batch_ref, rtol=2e-6, atol=0), (ours, per_batch_ref)
    per_epoch = lambda xs: [float(np.mean(xs[e * r["batches_per_epoch"]:(e + 1) * r["batches_per_epoch"]])) for e in range(r["epochs"])]
    ref_reg = [v for t, v, _ in r["scalars"] if t == "Loss/train_reg"]
    assert np.allclose(per_epoch(reg_terms), ref_reg, rtol=1e-5)
    if eik:
        ref_eik = [v for t, v, _ in r["scalars"] if t == "Loss/train_eikonal"]
        assert np.allclose(per_epoch(eik_terms), ref_eik, rtol=1e-5) and min(ref_eik) > 0
    # the script's final checkpoint
    for k, v in r["final_model"].items():
        n = k[len("module."):] if k.startswith("module.") else k
        assert torch.allclose(orc.params[n].detach(), v, rtol=0, atol=2e-7), n
    assert torch.allclose(orc.lat_vecs.weight.data, r["final_latents"], rtol=0, atol=2e-7)

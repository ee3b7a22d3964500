"""CPU tests of the host side: reference-compatible module surface, configuration errors,
parameter addressing, the C-ABI library's exports (no compute calls without a GPU)."""
import ctypes as C
import os
import re
import sys

import pytest
import torch

import dsdf_b200
from common import spec_for
from dsdf_b200 import _lib, dist
from dsdf_b200.trainer import StepConfig
from oracle import deepsdf

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KW = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False,
          use_tanh=False, latent_dropout=False, weight_norm=True)


@pytest.mark.parametrize("wn", [False, True])
def test_decoder_has_reference_parameters(wn):
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, norm_layers=list(range(8)) if wn else [], **KW)
    p = deepsdf.init_params(spec_for(wn), seed=0)
    sd = dec.state_dict()
    assert list(sd.keys()) == list(p.keys())
    for k in sd:
        assert torch.equal(sd[k], p[k]), k          # same RNG stream as the reference ctor
    assert [n for n, _ in dec.named_parameters()] == list(p.keys())


def test_flat_parameter_views_and_layout():
    dec = dsdf_b200.Decoder(256, norm_layers=list(range(8)), **KW)
    before = {k: v.clone() for k, v in dec.state_dict().items()}
    flat = dec.flat_params()
    assert flat.numel() == 1843195 == dec._layout.total
    off = 0
    for name, p in dec.named_parameters():
        assert p.data_ptr() == flat.data_ptr() + 4 * off
        assert torch.equal(p.data, before[name])
        off += p.numel()
    lay = dec._layout
    assert lay.g_off[8] == -1 and all(lay.g_off[l] >= 0 for l in range(8))
    # load_state_dict keeps the views (in-place copy)
    dec.load_state_dict({k: v * 2 for k, v in before.items()})
    assert dec.flat_params().data_ptr() == flat.data_ptr()
    assert torch.equal(dec.lin3.weight_v.data, before["lin3.weight_v"] * 2)


@pytest.mark.parametrize("kw,key", [
    (dict(dims=[256] * 8), "dims"), (dict(latent_in=[2]), "latent_in"), (dict(xyz_in_all=True), "xyz_in_all"),
    (dict(use_tanh=True), "use_tanh"), (dict(latent_dropout=True), "latent_dropout"),
    (dict(weight_norm=False, norm_layers=[0, 1]), "norm_layers"),
])
def test_unsupported_specs_raise_naming_the_key(kw, key):
    args = dict(KW); args.setdefault("norm_layers", []); args.update(kw)
    with pytest.raises(NotImplementedError, match=key):
        dsdf_b200.Decoder(256, **args)
    with pytest.raises(NotImplementedError, match="CodeLength"):
        dsdf_b200.Decoder(128, norm_layers=[], **KW)


def test_no_cpu_fallback():
    dec = dsdf_b200.Decoder(256, norm_layers=[], **KW)
    with pytest.raises(dsdf_b200.DsdfError):
        dec(torch.zeros(4, 259))


def test_library_loads_and_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "dsdf.h")).read()
    declared = sorted(set(re.findall(r"\b(dsdf_[a-z_0-9]+)\s*\(", hdr)))
    assert len(declared) >= 16
    for arch in ("deep_sdf_decoder", "siren_decoder"):     # one source tree, one build per family
        L = _lib.lib(arch)
        for name in declared:
            assert hasattr(L, name), (arch, name)
        assert L.dsdf_arch().decode() == arch
    assert sorted(_lib.EXPORTS) == declared
    assert C.sizeof(_lib.DsdfConfig) == 24 and C.sizeof(_lib.DsdfParamLayout) == 8 * 28


def test_create_fails_loudly_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    cfg = _lib.DsdfConfig(256, 512, 8, 4, 0, 0)
    h = C.c_void_p()
    rc = _lib.lib().dsdf_create(C.byref(cfg), C.byref(h))
    assert rc == -2 and b"no CPU fallback" in _lib.lib().dsdf_last_error(None)
    # unsupported network -> DSDF_EUNSUPPORTED before touching the device
    cfg = _lib.DsdfConfig(128, 512, 8, 4, 0, 0)
    assert _lib.lib().dsdf_create(C.byref(cfg), C.byref(h)) == -4


def test_step_config_from_reference_specs():
    import json
    specs = json.loads("""{"ClampingDistance":0.1,"CodeRegularization":true,"CodeRegularizationLambda":1e-4,
      "CodeBound":1.0,"GradientClipNorm":1.0,"LearningRateSchedule":[
      {"Type":"Step","Initial":0.0005,"Interval":500,"Factor":0.5},{"Type":"Step","Initial":0.001,"Interval":500,"Factor":0.5}]}""")
    c = StepConfig.from_specs(specs)
    assert (c.clamp_dist, c.code_bound, c.grad_clip, c.lr_decoder, c.lr_latent) == (0.1, 1.0, 1.0, 5e-4, 1e-3)


def test_scene_partition_is_a_partition():
    for S, W in [(800, 8), (16, 3), (7, 8), (4096, 4)]:
        seen = []
        for r in range(W):
            o = dist.owned_scenes(S, r, W)
            seen += list(o)
            for s in o:
                assert dist.owner_of(s, S, W) == r
        assert seen == list(range(S))
    assert sum(dist.local_batch(64, r, 8) for r in range(8)) == 64
    assert sum(dist.local_batch(24, r, 8) for r in range(8)) == 24


def test_sample_file_contract_matches_reference_data_module(tmp_path):
    """dsdf_b200.data restates deep_sdf/data.py:76-136,196-206: same rows for the same RNG streams as the golden
    drawn by the executed reference (tests/golden/callers_ref.npz) and, when the reference tree is mounted, as the
    live functions."""
    import random
    import numpy as np
    from common import golden
    from dsdf_b200 import data as D, synth
    d = synth.scene_samples(1, 4096, seed=9)[0]
    pos, neg = d[:2048].contiguous(), d[2048:].contiguous()
    random.seed(3)
    got = D.unpack_sdf_samples_from_ram([pos, neg], 800)
    np.testing.assert_array_equal(got.numpy(), golden("callers_ref.npz")["window_sample"])
    # file round trip incl. NaN rows (dropped, data.py:76-78) and the load-time shuffle
    pos_n = pos.clone(); pos_n[5, 3] = float("nan")
    f = str(tmp_path / "shape.npz")
    D.write_sdf_samples(f, pos_n, neg)
    p2, n2 = D.read_sdf_samples_into_ram(f)
    assert torch.equal(n2, neg) and torch.isnan(p2[5, 3])
    torch.manual_seed(1); random.seed(1)
    ds = D.SDFSamples([f], 512)
    item, idx = ds[0]
    assert item.shape == (512, 4) and idx == 0 and not torch.isnan(item).any()
    assert ds.loaded_data[0][0].shape[0] == 2047 and (item[:256, 3] > 0).all() and (item[256:, 3] <= 0).all()
    torch.manual_seed(2)
    a = D.unpack_sdf_samples(f, 100)
    assert a.shape == (100, 4) and not torch.isnan(a).any()
    if os.path.isdir("/root/reference/deep_sdf"):
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
        import make_golden_callers as mk
        mk.install_stubs()
        import deep_sdf.data as R
        random.seed(7); ref = R.unpack_sdf_samples_from_ram([pos, neg], 600)
        random.seed(7); mine = D.unpack_sdf_samples_from_ram([pos, neg], 600)
        assert torch.equal(ref, mine)
        torch.manual_seed(5); ref = R.unpack_sdf_samples(f, 64)
        torch.manual_seed(5); mine = D.unpack_sdf_samples(f, 64)
        assert torch.equal(ref, mine)
        assert torch.equal(R.remove_nans(pos_n), D.remove_nans(pos_n))


def test_small_shape_sampler_branch_is_laid_out_as_windows():
    """plan_samples (host side of dsdf_reconstruct) and data.unpack_sdf_samples_from_ram on a shape with fewer
    negatives than num_samples/2: bit-equal rows to the executed reference (callers_fallback_ref.npz)."""
    import random
    import numpy as np
    import torch
    from common import golden
    import dsdf_b200
    from dsdf_b200 import data as D, synth
    g = golden("callers_fallback_ref.npz")
    d = synth.scene_samples(1, 4096, seed=9)[0]
    pos, neg = d[:2048].contiguous(), d[2048:2348].contiguous()
    random.seed(3); torch.manual_seed(3)
    for k in range(3):
        np.testing.assert_array_equal(D.unpack_sdf_samples_from_ram([pos, neg], 800).numpy(), g[f"sample_{k}"])
    random.seed(3); torch.manual_seed(3)
    pool, wins = dsdf_b200.plan_samples(pos, neg, 800, 3, random)
    assert pool.shape == (1200, 4) and [w[1] for w in wins] == [0, 400, 800]
    for k in range(3):
        got = torch.cat([pos[wins[k][0]:wins[k][0] + 400], pool[wins[k][1]:wins[k][1] + 400]], 0)
        np.testing.assert_array_equal(got.numpy(), g[f"sample_{k}"])
    # usual case: the pool is the shape's own negatives and the draws are draw_windows'
    random.seed(5)
    pool2, wins2 = dsdf_b200.plan_samples(pos, d[2048:], 800, 4, random)
    random.seed(5)
    assert pool2 is not neg and pool2.shape[0] == 2048 and wins2 == dsdf_b200.draw_windows(2048, 2048, 800, 4, random)

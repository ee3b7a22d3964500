"""Shared helpers for the test-suite (not a test module)."""
import os

import numpy as np
import torch

from oracle import deepsdf

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REF_DIR = "/root/reference"


def spec_for(wn: bool) -> deepsdf.DecoderSpec:
    return deepsdf.DecoderSpec(norm_layers=tuple(range(8)) if wn else (), weight_norm=True)


def golden(name: str):
    return np.load(os.path.join(GOLDEN, name))


def check_checksums(params, g, prefix="cks/", rtol=1e-6):
    for k, v in params.items():
        want = g[prefix + k]
        got = np.array([float(v.double().sum()), float(v.double().abs().sum())])
        np.testing.assert_allclose(got, want, rtol=rtol, atol=1e-6, err_msg=k)


def cos(a: torch.Tensor, b: torch.Tensor) -> float:
    a = a.double().flatten().cpu()
    b = b.double().flatten().cpu()
    return float(a @ b / (a.norm() * b.norm() + 1e-300))


def eikonal_golden_check(g, named_grads, latgrad, prefix, rtol_norm, min_cos):
    """Compare {param name: grad tensor} + latent grad with a tests/golden/train_eikonal_*.npz arm
    (`prefix` = "eik/" or "base/"): per-parameter norms, full small gradients, strided samples of the rest."""
    stride = int(g["stride"])
    for k, v in named_grads.items():
        v = v.detach().float().cpu()
        want_n = float(g[prefix + "gnorm/" + k])
        assert abs(float(v.norm()) - want_n) <= rtol_norm * want_n + 1e-12, (k, float(v.norm()), want_n)
        if prefix + "g/" + k in g.files:
            assert cos(v, torch.from_numpy(g[prefix + "g/" + k])) >= min_cos, k
        else:
            assert cos(v.reshape(-1)[::stride], torch.from_numpy(g[prefix + "gs/" + k])) >= min_cos, k
    assert cos(latgrad, torch.from_numpy(g[prefix + "latgrad"])) >= min_cos

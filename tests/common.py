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

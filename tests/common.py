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


def icosphere(subdiv: int = 2, radius: float = 0.5, center=(0.0, 0.0, 0.0)):
    """(verts [nv,3] f32, faces [nf,3] i64) of a subdivided icosahedron, outward winding."""
    import numpy as np
    t = (1.0 + 5 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
         (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7),
         (9, 8, 1)]
    v = [np.array(p, dtype=np.float64) / np.linalg.norm(p) for p in v]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[k] = len(v) - 1
            return cache[k]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    verts = (np.array(v) * radius + np.array(center)).astype(np.float32)
    return verts, np.array(f, dtype=np.int64)

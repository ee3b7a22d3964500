"""Generate tests/golden/callers_ref.npz by EXECUTING the reference's own caller code (build container only):

    python tests/golden/make_golden_callers.py

Runs, unmodified from /root/reference: reconstruct.reconstruct (reconstruct.py:16-151) with
deep_sdf.data.unpack_sdf_samples_from_ram (data.py:112-136), deep_sdf.mesh.create_mesh's sampling loop
(mesh.py:21-71) with deep_sdf.utils.decode_sdf (utils.py:86-97), around networks.deep_sdf_decoder.Decoder,
all on CPU.  Three shims make that possible here, none touches the arithmetic:
  * empty stub modules for imports the container lacks and the executed lines never call
    (trimesh, plyfile, skimage.measure, pytorch3d.*, robust_laplacian);
  * torch.Tensor.cuda -> identity (create_mesh hard-codes .cuda() on each chunk, mesh.py:61; no GPU here);
  * deep_sdf.mesh.convert_sdf_samples_to_ply -> a recorder (marching cubes needs skimage): it receives exactly
    the sdf_values tensor create_mesh computed (mesh.py:68-71,80-87).
"""
import os
import random
import sys
import types
import warnings

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, "..", "..", "medical-shape-disentanglement_b200"))
warnings.simplefilter("ignore")

STUBS = ["trimesh", "plyfile", "skimage", "skimage.measure", "pytorch3d", "pytorch3d.structures",
         "pytorch3d.loss", "robust_laplacian"]


def _stub(name):
    m = types.ModuleType(name)
    m.__path__ = []

    def _attr(k):                                          # attribute access at import time only
        if k.startswith("__"):
            raise AttributeError(k)
        return types.SimpleNamespace()
    m.__getattr__ = _attr
    sys.modules[name] = m


def install_stubs():
    """Stub the known-missing imports, then any further one the import of reconstruct.py trips over."""
    for name in STUBS:
        if name not in sys.modules:
            _stub(name)
    for _ in range(40):
        try:
            import reconstruct  # noqa: F401
            return
        except ModuleNotFoundError as e:
            STUBS.append(e.name)
            _stub(e.name)
    raise RuntimeError("could not import the reference's reconstruct.py")


def make_decoder():
    from networks.deep_sdf_decoder import Decoder as RefDecoder
    torch.manual_seed(0)
    dec = RefDecoder(256, dims=[512] * 8, dropout=[], dropout_prob=0.2, norm_layers=list(range(8)), latent_in=[4],
                     xyz_in_all=False, use_tanh=False, latent_dropout=False, weight_norm=True)
    with torch.no_grad():                      # a livelier SDF than the default init (documented rule)
        for l in range(8):
            getattr(dec, f"lin{l}").weight_g.mul_(1.5)
    return dec


def main():
    install_stubs()
    import reconstruct as ref_reconstruct          # /root/reference/reconstruct.py
    import deep_sdf.data as ref_data
    import deep_sdf.mesh as ref_mesh
    from dsdf_b200 import synth                     # synthetic inputs only
    dec = make_decoder()
    out = {}
    for k, v in dec.state_dict().items():
        out["cks/" + k] = np.array([float(v.double().sum()), float(v.double().abs().sum())])
    d = synth.scene_samples(1, 4096, seed=9)[0]
    pos, neg = d[:2048].contiguous(), d[2048:].contiguous()
    # ---- data.unpack_sdf_samples_from_ram
    random.seed(3)
    out["window_sample"] = ref_data.unpack_sdf_samples_from_ram([pos, neg], 800).numpy()
    # ---- reconstruct.reconstruct
    for name, kw in (("a", dict(l2reg=True)), ("b", dict(code_reg_lambda=1e-2, code_reg_type="l2_norm", code_bound=0.2))):
        random.seed(0)
        torch.manual_seed(0)
        losses, latent = ref_reconstruct.reconstruct(dec, 30, 256, [pos, neg], 0.01, 0.1, num_samples=800, lr=5e-3,
                                                     return_loss_hist=True, **kw)
        out[f"recon_{name}/losses"] = np.array([float(x) for x in losses], dtype=np.float64)
        out[f"recon_{name}/latent"] = latent.detach().numpy()
    # ---- create_mesh sampling loop
    captured = {}
    ref_mesh.convert_sdf_samples_to_ply = lambda sdf, *a, **k: captured.setdefault("sdf", sdf.clone()) is None
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        lat = torch.from_numpy(out["recon_a/latent"])
        ref_mesh.create_mesh(dec, lat, filename="/tmp/_unused", N=24, max_batch=5000)
    finally:
        torch.Tensor.cuda = orig_cuda
    out["mesh/sdf24"] = captured["sdf"].numpy()
    np.savez_compressed(os.path.join(HERE, "callers_ref.npz"), **out)
    print({k: v.shape for k, v in out.items() if not k.startswith("cks")})
    print("losses a:", out["recon_a/losses"][:3], "...", out["recon_a/losses"][-1])


def main_fallback():
    """The sampler's small-shape branch (data.py:123-125: neg_size <= half -> torch.rand + index_select) and a
    reconstruct() run that takes it every iteration; written to callers_fallback_ref.npz."""
    install_stubs()
    import reconstruct as ref_reconstruct
    import deep_sdf.data as ref_data
    from dsdf_b200 import synth
    dec = make_decoder()
    out = {}
    d = synth.scene_samples(1, 4096, seed=9)[0]
    pos, neg = d[:2048].contiguous(), d[2048:2348].contiguous()          # 300 negatives < half = 400
    random.seed(3)
    torch.manual_seed(3)
    for k in range(3):
        out[f"sample_{k}"] = ref_data.unpack_sdf_samples_from_ram([pos, neg], 800).numpy()
    random.seed(0)
    torch.manual_seed(0)
    losses, latent = ref_reconstruct.reconstruct(dec, 20, 256, [pos, neg], 0.01, 0.1, num_samples=800, lr=5e-3,
                                                 return_loss_hist=True, l2reg=True)
    out["recon_c/losses"] = np.array([float(x) for x in losses], dtype=np.float64)
    out["recon_c/latent"] = latent.detach().numpy()
    np.savez_compressed(os.path.join(HERE, "callers_fallback_ref.npz"), **out)
    print({k: v.shape for k, v in out.items()})
    print("losses c:", out["recon_c/losses"][:3], "...", out["recon_c/losses"][-1])


if __name__ == "__main__":
    if "--fallback" in sys.argv:
        main_fallback()
    else:
        main()

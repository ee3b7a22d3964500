"""Diagnostic (not a test): SIREN parity as a function of the bias phase shift 2 pi k / 30."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-shape-disentanglement_b200"), os.path.join(ROOT, "tests")]
import torch
from common import cos
from oracle import siren
import test_gpu_siren as T

n = 2048
gen = torch.Generator().manual_seed(6)
xyz = (torch.rand(n, 3, generator=gen) * 2 - 1).cuda()
lat = (torch.randn(1, 256, generator=gen) * 0.5).cuda()
x = torch.cat([lat.expand(n, -1), xyz], 1)
w = torch.linspace(-1, 1, n, device="cuda")[:, None]
for k in (0, 1, 2, 5):
    for only in (None, 0, 3, 7):
        dec, params = T.make()
        big = T._phase_shifted_params(params, k=k)
        if only is not None:
            big = {kk: (big[kk] if kk == f"decoder.lin{only}.bias" else params[kk].clone()) for kk in params}
        dec.load_state_dict(big)
        p_ref = {kk: v.clone().cuda().double().requires_grad_(True) for kk, v in big.items()}
        out_ref = siren.forward(p_ref, x.double())
        (out_ref * w.double()).sum().backward()
        x1 = x.clone().requires_grad_(True)
        out = dec(x1)
        (out * w).sum().backward()
        with torch.no_grad():
            fold = dec.decode_sdf(lat, xyz)
        per = []
        for name, p in dec.named_parameters():
            a, b = p.grad.flatten().double(), p_ref[name].grad.flatten()
            per.append(f"{name.split('.')[1][3:]}{name.split('.')[-1][0]}:{float(a.norm()/b.norm()):.3f}/{cos(a, b):.4f}")
        e = (out.double() - out_ref).flatten()
        ef = (fold.double().flatten() - out_ref.flatten())
        print(f"k={k} only={only} out err max {float(e.abs().max()):.2e} mean {float(e.mean()):+.2e} rms {float(e.pow(2).mean().sqrt()):.2e} | fold rms {float(ef.pow(2).mean().sqrt()):.2e} | |out| rms {float(out_ref.pow(2).mean().sqrt()):.2e} | " + " ".join(per))

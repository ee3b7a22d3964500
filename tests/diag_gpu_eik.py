"""Diagnostic (not a test): per-iteration gradient norm / cosine of the fused step against the oracle, with and
without the Eikonal term, on the 'trained' weights of test_gpu_eikonal.  Run on the GPU box:
python tests/diag_gpu_eik.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-shape-disentanglement_b200"), os.path.join(ROOT, "tests")]
import torch
from dsdf_b200 import synth
from common import cos
import test_gpu_eikonal as T

for eik in (False, True):
    for P in (512, 4096):
        S, B = 8, 4
        dec, orc, tr = T.pair(False, S, B, P, "trained", eik)
        data = synth.scene_samples(S, P, seed=1).cuda()
        g = torch.Generator().manual_seed(3)
        for it in range(4):
            ids = torch.randperm(S, generator=g)[:B]
            orc.opt.zero_grad()
            orc.forward_loss(data[ids], ids, 50 + it)
            ref = T.flat_ref(dec, orc).clone()
            tr.forward_backward(data[ids], ids.cuda(), 50 + it)
            torch.cuda.synchronize()
            got = tr.grads.clone()
            off = 0
            per = []
            for n_, p_ in dec.named_parameters():
                a, b = got[off:off + p_.numel()], ref[off:off + p_.numel()]
                off += p_.numel()
                per.append(f"{n_.split('.')[0][3:]}{n_.split('.')[-1][0]}:{float(a.norm()/b.norm()):.3f}/{cos(a, b):.4f}")
            print(f"eik={eik} P={P} it={it} |got|={float(got.norm()):.5f} |ref|={float(ref.norm()):.5f} ratio={float(got.norm()/ref.norm()):.4f} "
                  f"cos={cos(got, ref):.6f}  " + " ".join(per))

"""Loss-trajectory diagnostic (test infrastructure, not collected by pytest): fused path vs oracle vs a 1e-6-perturbed
oracle (inherent chaos).  python tests/diag_gpu_traj.py  -> profiles/r01_trajectory_chaos*.log"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import dsdf_b200
from dsdf_b200 import synth
from oracle import deepsdf, train as otrain
from common import spec_for

def run(S, B, P, steps, lr_scale=1.0, wn=True):
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, dims=[512]*8, dropout=[], latent_in=[4], norm_layers=list(range(8)) if wn else [], weight_norm=True)
    params = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    dec = dec.cuda(); spec = spec_for(wn)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0, lr_decoder=5e-4*lr_scale, lr_latent=1e-3*lr_scale)
    orc = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params)
    lat0 = orc.lat_vecs.weight.data.clone()
    orc2 = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params, latents=lat0 * (1 + 1e-6 * torch.randn_like(lat0)))
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, lr_decoder=5e-4*lr_scale, lr_latent=1e-3*lr_scale), latents=lat0.clone())
    data = synth.scene_samples(S, P, seed=2).cuda()
    g = torch.Generator().manual_seed(11)
    a, b, c = [], [], []
    for it in range(steps):
        ids = torch.randperm(S, generator=g)[:B]
        epoch = 1 + it // 4
        a.append(orc.step(data[ids], ids, epoch)["loss"])
        c.append(orc2.step(data[ids], ids, epoch)["loss"])
        st = tr.step(data[ids], ids.cuda(), epoch)
        b.append((st["sdf_loss"] + st["reg_loss"]).item())
    a, b, c = map(np.array, (a, b, c))
    print(f"S={S} B={B} P={P} lr_scale={lr_scale} wn={wn}")
    for s in range(0, steps, 50):
        e = min(steps, s + 50)
        print(f"  steps {s:3d}-{e:3d}: ref mean {a[s:e].mean():.5f} ours rel {np.abs(b[s:e]-a[s:e]).max()/a[s:e].mean():.4f} (mean {np.abs(b[s:e].mean()-a[s:e].mean())/a[s:e].mean():.4f}) | perturbed-oracle rel {np.abs(c[s:e]-a[s:e]).max()/a[s:e].mean():.4f} (mean {np.abs(c[s:e].mean()-a[s:e].mean())/a[s:e].mean():.4f})")

if __name__ == "__main__":
    lr = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    run(16, 8, 4096, 500, lr_scale=lr)
    run(16, 8, 4096, 500, lr_scale=lr, wn=False)

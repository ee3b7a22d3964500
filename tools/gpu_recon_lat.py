"""Latency of the latent-only optimisation loop (reconstruct.py defaults: 800 iterations x 8000 samples, one shape)
through dsdf_reconstruct.   python tools/gpu_recon_lat.py [iterations] [n_shapes]
Under `ncu --metrics gpu__time_duration.sum` with a few iterations it gives the per-kernel durations."""
import os, random, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch
import dsdf_b200
from dsdf_b200 import synth

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 800
n_shapes = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda:0")
torch.manual_seed(0)
dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
dec.eval()
shapes = []
for s_ in range(n_shapes):
    d = synth.fast_scene_samples(1, 500000, seed=100 + s_, device="cuda")[0]
    shapes.append((d[:250000].contiguous(), d[250000:].contiguous()))
rng = random.Random(0)
wins = torch.tensor([[dsdf_b200.draw_windows(250000, 250000, 8000, 1, rng)[0] for _ in range(n_shapes)]
                     for _ in range(iters)], dtype=torch.int64)
lat0 = torch.randn(n_shapes, 256) * 0.01
best = None
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hist, lat = dsdf_b200.reconstruct_batch(dec, iters, shapes, lat0, wins, 0.1, num_samples=8000, lr=5e-3, l2reg=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    best = dt if best is None else min(best, dt)
print(f"{os.path.basename(os.environ.get('DSDF_LIB', 'product')):24s} reconstruct {iters} it x 8000 samples x {n_shapes} shape(s): "
      f"{best * 1e3:.2f} ms wall = {best / iters * 1e6:.1f} us/iteration; loss {float(hist[0].mean()):.5f} -> "
      f"{float(hist[-1].mean()):.5f}; |z| {float(lat.norm(dim=1).mean()):.4f}")

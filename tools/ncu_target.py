"""Small fixed workload for ncu captures: 2 training steps (16 scenes x 16384 rows) + one 128^3 grid decode."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch
import dsdf_b200
from dsdf_b200 import synth
dev = torch.device("cuda:0")
torch.manual_seed(0)
B, P = 16, 16384
dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
tr = dsdf_b200.FusedTrainer(dec, 32, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0))
data = synth.fast_scene_samples(B, P, seed=0, device="cuda")
ids = torch.arange(B, device=dev)
for _ in range(2):
    tr.step(data, ids, 10)
lat = torch.randn(1, 256, device=dev) * 0.05
dec.eval()
dec.decode_grid(lat, 128)
torch.cuda.synchronize()
print("ok", tr.loss_buf[:2].tolist())

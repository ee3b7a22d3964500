"""Fixed workload for ncu captures: 2 training steps (NCU_B scenes x 16384 rows; 64 = the bench's
1 048 576 rows/step) + one 128^3 grid decode + (NCU_SIREN=1) one SIREN step + a short reconstruct."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch
import dsdf_b200
from dsdf_b200 import synth
dev = torch.device("cuda:0")
torch.manual_seed(0)
B, P = int(os.environ.get("NCU_B", "16")), 16384
dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
tr = dsdf_b200.FusedTrainer(dec, max(32, B), dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0))
data = synth.fast_scene_samples(B, P, seed=0, device="cuda")
ids = torch.arange(B, device=dev)
for _ in range(2):
    tr.step(data, ids, 10)
lat = torch.randn(1, 256, device=dev) * 0.05
dec.eval()
dec.decode_grid(lat, 128)
torch.cuda.synchronize()
print("ok", tr.loss_buf[:2].tolist())
if os.environ.get("NCU_SIREN"):
    import random
    import dsdf_b200.siren as S
    sd = S.Decoder(256, dims=[512] * 8, latent_in=[4], nonlinearity="sine").to(dev)
    st = dsdf_b200.FusedTrainer(sd, max(32, B), dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, lr_decoder=1e-5))
    st.step(data, ids, 10)
    d = synth.fast_scene_samples(1, 65536, seed=3, device="cuda")[0]
    wins = torch.tensor(dsdf_b200.draw_windows(32768, 32768, 8000, 3, random.Random(0))).view(3, 1, 2)
    dsdf_b200.reconstruct_batch(dec, 3, [(d[:32768].contiguous(), d[32768:].contiguous())], torch.zeros(1, 256), wins, 0.1,
                                num_samples=8000, lr=5e-3, l2reg=True)
    torch.cuda.synchronize()
    print("siren ok", st.loss_buf[:2].tolist())

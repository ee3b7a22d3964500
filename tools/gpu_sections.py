"""Per-section CUDA-event timings of the fused step (dsdf_profile_*) for the library selected by DSDF_LIB.
python tools/gpu_sections.py [steps] [eik]   -- one line, for same-box A/B runs (tools/gpu_ab.sh)"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch
import dsdf_b200
from dsdf_b200 import _lib, synth

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
eik = len(sys.argv) > 2 and sys.argv[2] == "eik"
dev = torch.device("cuda:0")
torch.manual_seed(0)
dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
B, P = 64, 16384
tr = dsdf_b200.FusedTrainer(dec, 2 * B, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, use_eikonal=eik))
data = synth.fast_scene_samples(B, P, seed=0, device="cuda")
ids = torch.arange(B, device=dev)
for _ in range(3):
    tr.step(data, ids, 10)
torch.cuda.synchronize()
L, h = _lib.lib(), dec.handle()
L.dsdf_profile_enable(h, 1)
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    tr.step(data, ids, 10)
e1.record()
torch.cuda.synchronize()
ms = (C.c_float * 8)(); n = C.c_int32()
_lib.check(L.dsdf_profile_read(h, ms, C.byref(n)), h)
per = [ms[i] / max(1, n.value) for i in range(8)]
lat = torch.randn(1, 256, device=dev) * 0.05
dec.eval()
for _ in range(2):
    dec.decode_grid(lat, 256)
d0 = torch.cuda.Event(enable_timing=True); d1 = torch.cuda.Event(enable_timing=True)
d0.record(); dec.decode_grid(lat, 256); d1.record(); torch.cuda.synchronize()
print(f"{os.path.basename(os.environ.get('DSDF_LIB', 'product')):28s} step {e0.elapsed_time(e1) / steps:6.3f} ms | fwd {per[0]:.3f} dgrad {per[1]:.3f} tan {per[2]:.3f} "
      f"corr {per[3]:.3f} wgrad {per[4]:.3f} rest {sum(per[5:]):.3f} | decode256 {d0.elapsed_time(d1):.2f} ms")

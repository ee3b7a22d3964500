"""Per-section CUDA-event timings of the fused SIREN step (BASELINE config 5: 64 scenes x 16384 rows) through
libdsdf_b200_siren.so.   python tools/gpu_sections_siren.py [steps]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch
import dsdf_b200
import dsdf_b200.siren as S
from dsdf_b200 import _lib, synth

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda:0")
torch.manual_seed(0)
dec = S.Decoder(256, [512] * 8, latent_in=[4], nonlinearity="sine").to(dev)
B, P = 64, 16384
tr = dsdf_b200.FusedTrainer(dec, 2 * B, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, lr_decoder=2e-6, lr_latent=1e-4))
data = synth.fast_scene_samples(B, P, seed=0, device="cuda")
ids = torch.arange(B, device=dev)
for _ in range(3):
    tr.step(data, ids, 10)
torch.cuda.synchronize()
L, h = dec.L, dec.handle()
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
print(f"siren step {e0.elapsed_time(e1) / steps:6.3f} ms | fwd {per[0]:.3f} dgrad {per[1]:.3f} wgrad {per[4]:.3f} "
      f"rest {sum(per[5:]) + per[2] + per[3]:.3f}")

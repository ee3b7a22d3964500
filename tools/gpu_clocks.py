"""Run the train step / grid decode for a few seconds and sample SM clock + power (nvidia-smi)."""
import os, sys, subprocess, time, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch
import dsdf_b200
from dsdf_b200 import synth
dev = torch.device("cuda:0")
torch.manual_seed(0)
dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
B, P = 64, 16384
tr = dsdf_b200.FusedTrainer(dec, 2 * B, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0))
data = synth.fast_scene_samples(B, P, seed=0, device="cuda")
ids = torch.arange(B, device=dev)
lat = torch.randn(1, 256, device=dev) * 0.05
rows = []
stop = False
def sample():
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,temperature.gpu,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown",
                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
    for line in p.stdout:
        rows.append((time.time(), line.strip()))
        if stop:
            break
    p.terminate()
th = threading.Thread(target=sample, daemon=True); th.start()
def run(name, fn, secs):
    fn(); torch.cuda.synchronize()
    t0 = time.time(); n = 0
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < secs:
        for _ in range(10):
            fn()
        n += 10
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    t1 = time.time()
    r = [x for (t, x) in rows if t0 + 0.5 <= t <= t1]
    print(f"== {name}: {e0.elapsed_time(e1) / n:.3f} ms/iter over {n} iters")
    for x in r[:: max(1, len(r) // 8)]:
        print("   ", x)
time.sleep(0.5)
run("train step 1M rows", lambda: tr.step(data, ids, 10), 4.0)
run("decode 256^3", lambda: dec.decode_grid(lat, 256), 4.0)
run("decode 128^3", lambda: dec.decode_grid(lat, 128), 3.0)
stop = True

"""Quick kernel timings (CUDA events) for the fused path at benchmark shapes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch
import dsdf_b200
from dsdf_b200 import synth

dev = torch.device("cuda:0")
FWD_FLOP = 3671040.0


def timeit(fn, n=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), sum(ts) / len(ts)


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
    dec.eval()
    lat = torch.randn(1, 256, device=dev) * 0.05
    for N in (128, 256):
        t, ta = timeit(lambda: dec.decode_grid(lat, N))
        pts = N ** 3
        print(f"decode_grid N={N}: {t:.3f} ms (avg {ta:.3f}) -> {pts/t/1e3:.1f} M pts/s, {pts*FWD_FLOP/t/1e9:.1f} TFLOP/s")
    S = max(B * 2, 16)
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0))
    data = synth.fast_scene_samples(B, P, seed=0, device="cuda")
    ids = torch.arange(B, device=dev)
    M = B * P
    t, ta = timeit(lambda: tr.forward_backward(data, ids, 10))
    print(f"fwd+bwd M={M}: {t:.3f} ms (avg {ta:.3f}) -> {M/t/1e3:.2f} M samples/s, {M*3*FWD_FLOP/t/1e9:.1f} TFLOP/s")
    t, ta = timeit(lambda: tr.step(data, ids, 10))
    print(f"full step M={M}: {t:.3f} ms (avg {ta:.3f}) -> {M/t/1e3:.2f} M samples/s, {M*3*FWD_FLOP/t/1e9:.1f} TFLOP/s")
    # per-kernel breakdown via the profiler
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        tr.step(data, ids, 10)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=20, max_name_column_width=60))


if __name__ == "__main__":
    main()

"""8-GPU probe (torchrun): (a) the 7.4 MB fp32 decoder-gradient all-reduce alone, back to back, (b) per-rank time of the
strong-scaled step WITHOUT the collective (rank skew), (c) the step with it.  Prints one line per item on rank 0."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import torch, torch.distributed as dist
import dsdf_b200
from dsdf_b200 import synth
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
def timed(fn, n, warm=5):
    for _ in range(warm): fn()
    dist.barrier(); torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for nbytes in (7372780, 1 << 20, 64 << 20):
    x = torch.zeros(nbytes // 4, device=dev)
    ms = timed(lambda: dist.all_reduce(x), 50)
    if rank == 0: print(f"all_reduce alone {nbytes/1e6:.1f} MB fp32: {ms*1e3:.1f} us per call (back to back, {world} ranks)")
torch.manual_seed(0)
dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
B, P = 64 // world, 16384
data = synth.fast_scene_samples(B, P, seed=rank, device=str(dev)); ids = torch.arange(B, device=dev)
cfg = dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0)
tr0 = dsdf_b200.FusedTrainer(dec, 2 * B, cfg)                       # no process group: no collective
ms0 = timed(lambda: tr0.step(data, ids, 10, n_total=64 * P), 100)
all0 = [None] * world; dist.all_gather_object(all0, ms0)
tr1 = dsdf_b200.FusedTrainer(dec, 2 * B, cfg, process_group=dist.group.WORLD)
ms1 = timed(lambda: tr1.step(data, ids, 10, n_total=64 * P), 100)
all1 = [None] * world; dist.all_gather_object(all1, ms1)
if rank == 0:
    print("step without collective, per rank ms:", " ".join(f"{v:.3f}" for v in all0))
    print("step with collective,    per rank ms:", " ".join(f"{v:.3f}" for v in all1))
dist.destroy_process_group()

"""Per-CTA cycle counters of the persistent kernels (DSDF_DBG=1): where do the MMA issuer, the
weight producer and the epilogue warps wait?  python tools/gpu_dbg.py [decode|train]"""
import os, sys
os.environ["DSDF_DBG"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import ctypes as C
import numpy as np
import torch
import dsdf_b200
from dsdf_b200 import _lib, synth

dev = torch.device("cuda:0")
NAMES = ["mma_total", "mma_wait_full", "mma_wait_a_lo", "mma_wait_a_hi", "prod_wait_empty",
         "epi_wait_d0", "epi_wait_d1", "epi_total", "mma_wait_e_full", "epi_wait_stash_read", "epi_sum_t_free", "mma_sum_t_first", "epi_free_to_publish_lo", "epi_half1", "s14", "s15"]


def read(dec, label, tiles_per_cta):
    L = _lib.lib()
    h = dec.handle()
    n = L.dsdf_num_sms(h) * 16
    buf = (C.c_uint64 * n)()
    got = L.dsdf_debug_read(h, buf, n)
    a = np.frombuffer(buf, dtype=np.uint64).reshape(-1, 16).astype(np.float64)
    a = a[a[:, 7] > 0]
    if a.shape[0] == 0:
        print(f"== {label}: no counters (build with DSDF_BUILD_VARIANT=dbg DSDF_BUILD_DEFS=-DDSDF_DBG_BUILD and set DSDF_LIB)")
        return
    print(f"== {label}: {a.shape[0]} CTAs, {tiles_per_cta:.1f} tiles/CTA")
    lead = a[a[:, 11] > 0]
    if lead.shape[0]:
        bub = (lead[:, 11] - lead[:, 10]) / (7 * tiles_per_cta)
        print(f"  layer-boundary bubble (layer l retired -> first MMA of l+1 issued): mean {bub.mean():8.0f} cyc/layer  min {bub.min():8.0f} max {bub.max():8.0f}")
    for i, nm in enumerate(NAMES):
        if i in (10, 11):
            continue
        if a[:, i].max() > 0:
            print(f"  {nm:18s} mean {a[:, i].mean():12.0f} cyc  ({a[:, i].mean() / tiles_per_cta:9.0f} /tile)  min {a[:, i].min():12.0f} max {a[:, i].max():12.0f}")


def main():
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=list(range(8)), weight_norm=True, dropout=[]).to(dev)
    dec.eval()
    lat = torch.randn(1, 256, device=dev) * 0.05
    N = 128
    dec.decode_grid(lat, N); torch.cuda.synchronize()
    read(dec, "warm-up (discard)", 1)
    dec.decode_grid(lat, N); torch.cuda.synchronize()
    read(dec, f"decode_grid N={N}", N ** 3 / 128 / 148)
    B, P = 64, 16384
    tr = dsdf_b200.FusedTrainer(dec, 2 * B, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0))
    data = synth.fast_scene_samples(B, P, seed=0, device="cuda")
    ids = torch.arange(B, device=dev)
    tr.forward_backward(data, ids, 10); torch.cuda.synchronize()
    read(dec, "warm-up (discard)", 1)
    tr.forward_backward(data, ids, 10); torch.cuda.synchronize()
    read(dec, f"train fwd+bwd M={B * P}", B * P / 128 / 148)


if __name__ == "__main__":
    main()

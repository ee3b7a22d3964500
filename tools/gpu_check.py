"""Diagnostic driver for the first GPU trips: prints per-stage error figures instead of
pass/fail.  Usage: python tools/gpu_check.py {selftest|forward|grid|train|all}"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))

import ctypes as C

import numpy as np
import torch

import dsdf_b200
from dsdf_b200 import _lib, synth
from oracle import deepsdf, mesh_grid
from oracle import train as otrain

dev = torch.device("cuda:0")


def sub_image(mat):
    """[128, 64] float -> 16 KB SUB image (SWIZZLE_128B) as uint8 tensor."""
    b = mat.to(torch.bfloat16).view(torch.int16).numpy().astype(np.uint16)
    img = np.zeros(8192, dtype=np.uint16)
    r = np.arange(128)[:, None]
    c = np.arange(64)[None, :]
    off = (r // 8) * 1024 + (r % 8) * 128 + ((((c // 8) ^ r) & 7) << 4) + ((c & 7) << 1)
    img[(off // 2).reshape(-1)] = b.reshape(-1)
    return torch.from_numpy(img.view(np.uint8).copy())


def selftest():
    L = _lib.lib()
    torch.manual_seed(1)
    # kind 0: K-major
    A = torch.randn(128, 64).bfloat16().float()
    B = torch.randn(128, 64).bfloat16().float()
    a = sub_image(A).to(dev)
    b = sub_image(B).to(dev)
    d = torch.zeros(128, 128, device=dev)
    rc = L.dsdf_selftest_gemm(0, a.data_ptr(), b.data_ptr(), d.data_ptr(), None)
    torch.cuda.synchronize()
    ref = A @ B.t()
    print("selftest K-major rc", rc, "max err", (d.cpu() - ref).abs().max().item(), "ref max", ref.abs().max().item())
    # kind 1: MN-major. A^T: [k<32][m<128] -> 2 SUBs (rows = k, cols = 64 m each)
    At = torch.randn(32, 128).bfloat16().float()
    Bt = torch.randn(32, 256).bfloat16().float()
    subs = []
    for j in range(2):
        m = torch.zeros(128, 64)
        m[:32] = At[:, 64 * j:64 * j + 64]
        subs.append(sub_image(m))
    a = torch.cat(subs).to(dev)
    subs = []
    for j in range(4):
        m = torch.zeros(128, 64)
        m[:32] = Bt[:, 64 * j:64 * j + 64]
        subs.append(sub_image(m))
    b = torch.cat(subs).to(dev)
    d = torch.zeros(128, 256, device=dev)
    rc = L.dsdf_selftest_gemm(1, a.data_ptr(), b.data_ptr(), d.data_ptr(), None)
    torch.cuda.synchronize()
    ref = At.t() @ Bt
    print("selftest MN-major rc", rc, "max err", (d.cpu() - ref).abs().max().item(), "ref max", ref.abs().max().item())


def make(wn, seed=0, scale=1.0):
    nl = list(range(8)) if wn else []
    torch.manual_seed(seed)
    dec = dsdf_b200.Decoder(256, [512] * 8, latent_in=[4], norm_layers=nl, weight_norm=True, dropout=[])
    spec = deepsdf.DecoderSpec(norm_layers=tuple(nl), weight_norm=True)
    params = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    if scale != 1.0:
        for k in params:
            if k.endswith("weight") or k.endswith("weight_g"):
                params[k] *= scale
        dec.load_state_dict(params)
    return dec.to(dev), spec, params


def forward():
    for wn in (False, True):
        for scale in (1.0, 1.6):
            dec, spec, params = make(wn, scale=scale)
            dec.eval()
            g = torch.Generator().manual_seed(3)
            for n in (1, 100, 128, 1000, 40000):
                x = torch.cat([torch.randn(n, 256, generator=g) * 0.06, torch.rand(n, 3, generator=g) * 2 - 1], 1)
                ref = deepsdf.forward({k: v.to(dev) for k, v in params.items()}, spec, x.to(dev))
                t0 = time.time()
                with torch.no_grad():
                    out = dec(x.to(dev))
                torch.cuda.synchronize()
                err = (out - ref).abs().max().item()
                print(f"forward wn={wn} scale={scale} n={n}: max err {err:.3e}  ref range [{ref.min().item():.3f},{ref.max().item():.3f}]  ({time.time()-t0:.3f}s)")
            lat = torch.randn(256, generator=g) * 0.06
            q = torch.rand(5000, 3, generator=g) * 2 - 1
            ref = mesh_grid.decode_sdf({k: v.to(dev) for k, v in params.items()}, spec, lat.to(dev)[None], q.to(dev))
            out = dec.decode_sdf(lat.to(dev), q.to(dev))
            print(f"decode_sdf wn={wn}: max err {(out-ref).abs().max().item():.3e}")


def grid():
    L = _lib.lib()
    dec, spec, params = make(True)
    h = dec.sync_weights()
    for N in (16, 64, 100):
        xyz = torch.empty(N ** 3, 3, device=dev)
        _lib.check(L.dsdf_grid_coords(h, N, 0, N ** 3, xyz.data_ptr(), None), h)
        ref = mesh_grid.grid_coords_reference_ops(N)
        print(f"grid coords N={N} bit-exact:", torch.equal(xyz.cpu(), ref))
    N = 64
    g = torch.Generator().manual_seed(5)
    lat = torch.randn(1, 256, generator=g) * 0.05
    ref = mesh_grid.decode_grid(params, spec, lat, N, device="cuda")
    out = dsdf_b200.create_mesh_sdf(dec, lat, N)
    print("grid decode N=64 max err", (out.reshape(-1) - ref).abs().max().item())
    lo, hi, part = dsdf_b200.decode_grid_sharded(dec, lat, N, 1, 4)
    print("slab shard equal:", torch.equal(part.cpu(), out.reshape(-1)[lo:hi]))


def cos(a, b):
    a = a.double().flatten(); b = b.double().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def train(wn=True, S=8, B=4, P=256, steps=3):
    dec, spec, params = make(wn)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0)
    orc = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params)
    lat0 = orc.lat_vecs.weight.data.clone()
    lat0[1] *= 30.0   # exceeds CodeBound -> exercises the renorm
    orc.lat_vecs.weight.data.copy_(lat0)
    scfg = dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0)
    tr = dsdf_b200.FusedTrainer(dec, S, scfg, latents=lat0)
    data = synth.scene_samples(S, P, seed=1).to(dev)
    g = torch.Generator().manual_seed(7)
    for it in range(steps):
        ids = torch.randperm(S, generator=g)[:B]
        epoch = 50 + it
        batch = data[ids]
        # oracle grads (no optimiser yet)
        orc.set_epoch_lr(epoch)
        orc.opt.zero_grad()
        sl, rl, pred_ref = orc.forward_loss(batch, ids, epoch)
        pred = tr.forward_backward(batch, ids.to(dev), epoch)
        torch.cuda.synchronize()
        print(f"[it {it}] loss ref {sl:.6e}+{rl:.6e}  ours {tr.loss_buf[0].item():.6e}+{tr.loss_buf[1].item():.6e}")
        print("   pred max err (clamped)", (pred.clamp(-0.1, 0.1)[:, None] - pred_ref).abs().max().item())
        flat = dec.flat_params()
        off = 0
        worst = 1.0
        for (name, p) in dec.named_parameters():
            n = p.numel()
            gr = tr.grads[off:off + n].view(p.shape)
            c = cos(gr, orc.params[name].grad)
            rn = (gr.norm() / (orc.params[name].grad.norm() + 1e-30)).item()
            worst = min(worst, c)
            if c < 0.9995 or abs(rn - 1) > 0.02:
                print(f"   {name}: cos {c:.6f} norm ratio {rn:.4f}")
            off += n
        print("   worst decoder-grad cos", worst, " total cos", cos(tr.grads, torch.cat([orc.params[n_].grad.flatten() for n_, _ in dec.named_parameters()])))
        lg = orc.lat_vecs.weight.grad
        print("   latent grad cos", cos(tr.lat_grad, lg), " zero rows equal:", torch.equal(tr.lat_grad == 0, lg == 0),
              " table equal after renorm:", torch.equal(tr.lat_vecs.weight.data, orc.lat_vecs.weight.data))
        # now apply the optimiser on both
        gn = torch.nn.utils.clip_grad_norm_(list(orc.params.values()), cfg.grad_clip)
        orc.opt.step()
        st = tr.step(batch, ids.to(dev), epoch)
        torch.cuda.synchronize()
        print("   grad norm ref", float(gn), "ours", st["grad_norm"].item())
        pf = torch.cat([orc.params[n_].data.flatten() for n_, _ in dec.named_parameters()])
        d = (dec.flat_params() - pf).abs()
        print("   param diff after Adam: max", d.max().item(), "mean", d.mean().item(), "(lr 5e-4)",
              " latent max diff", (tr.lat_vecs.weight.data - orc.lat_vecs.weight.data).abs().max().item())
        # re-synchronise both sides so the next iteration compares gradients at identical weights
        with torch.no_grad():
            for n_, p_ in dec.named_parameters():
                p_.copy_(orc.params[n_].data)
            tr.lat_vecs.weight.data.copy_(orc.lat_vecs.weight.data)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    print("device:", torch.cuda.get_device_name(0))
    if what in ("selftest", "all"):
        selftest()
    if what in ("forward", "all"):
        forward()
    if what in ("grid", "all"):
        grid()
    if what in ("train", "all"):
        train(True)
        train(False, S=6, B=3, P=384, steps=2)

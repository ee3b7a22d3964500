"""Generate the golden fixtures from the LIVE reference (run in the build container only):

    python tests/golden/make_golden.py

Imports networks.deep_sdf_decoder.Decoder from /root/reference (imports unchanged: torch only)
and drives it with the reference's own torch calls (nn.Embedding(max_norm), L1Loss(sum),
clip_grad_norm_, Adam) exactly as train_deep_sdf.py:429-454,555-616,731,780-784 does.  The
reference ships no tests or known-answer vectors, so these files are the
pin.  Weights are regenerated from the seed at test time (default nn.Linear init is
deterministic for a given torch CPU generator); per-tensor checksums are stored to prove it.
"""
import math
import os
import sys
import warnings

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, "..", "..", "medical-shape-disentanglement_b200"))
warnings.simplefilter("ignore")

from networks.deep_sdf_decoder import Decoder as RefDecoder  # noqa: E402

from dsdf_b200 import synth  # noqa: E402  (synthetic inputs only; no product compute)

SPECS = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False,
             use_tanh=False, latent_dropout=False, weight_norm=True)


def make_decoder(wn: bool, seed: int = 0):
    torch.manual_seed(seed)
    return RefDecoder(256, norm_layers=list(range(8)) if wn else [], **SPECS)


def checksums(dec):
    return {k: np.array([float(v.double().sum()), float(v.double().abs().sum())])
            for k, v in dec.state_dict().items()}


def golden_forward(wn: bool, name: str):
    dec = make_decoder(wn)
    g = torch.Generator().manual_seed(11)
    n = 192
    x = torch.cat([torch.randn(n, 256, generator=g) * 0.06, torch.rand(n, 3, generator=g) * 2 - 1], 1)
    x.requires_grad_(True)
    y = dec(x)
    w = torch.linspace(-1, 1, n)[:, None]
    (y * w).sum().backward()
    out = {"input": x.detach().numpy(), "output": y.detach().numpy(), "dinput": x.grad.numpy(),
           "w": w.numpy()}
    for k, v in dec.named_parameters():
        out["gradnorm/" + k] = np.array(float(v.grad.norm()))
    for k, v in checksums(dec).items():
        out["cks/" + k] = v
    # a few gradient slices for direct comparison
    out["grad/lin8.bias"] = dec.lin8.bias.grad.numpy()
    out["grad/lin7.bias"] = dec.lin7.bias.grad.numpy()
    out["grad/lin0.bias"] = dec.lin0.bias.grad.numpy()
    np.savez_compressed(os.path.join(HERE, name), **out)


def golden_train(wn: bool, name: str, S=8, B=4, P=256, steps=40):
    dec = make_decoder(wn)
    torch.manual_seed(123)
    lat = torch.nn.Embedding(S, 256, max_norm=1.0)
    torch.nn.init.normal_(lat.weight.data, 0.0, 1.0 / math.sqrt(256))
    lat.weight.data[1] *= 30.0        # exceeds CodeBound: exercises the renorm side effect
    lat0 = lat.weight.data.clone()
    opt = torch.optim.Adam([{"params": dec.parameters(), "lr": 5e-4},
                            {"params": lat.parameters(), "lr": 1e-3}])
    loss_l1 = torch.nn.L1Loss(reduction="sum")
    data = synth.scene_samples(S, P, seed=1)
    g = torch.Generator().manual_seed(7)
    losses, regs, gnorms, ids_all = [], [], [], []
    first = {}
    for it in range(steps):
        ids = torch.randperm(S, generator=g)[:B]
        epoch = 50 + it
        sdf_data = data[ids].reshape(-1, 4)
        n = sdf_data.shape[0]
        xyz = sdf_data[:, 0:3]
        gt = torch.clamp(sdf_data[:, 3:4], -0.1, 0.1)
        idx = ids.unsqueeze(-1).repeat(1, P).view(-1)
        opt.zero_grad()
        bv = lat(idx)
        pred = torch.clamp(dec(torch.cat([bv, xyz], 1)), -0.1, 0.1)
        loss = loss_l1(pred, gt) / n
        reg = (1e-4 * min(1, epoch / 100) * torch.sum(torch.norm(bv, dim=1))) / n
        (loss + reg).backward()
        if it == 0:
            first["pred0"] = pred.detach().numpy()
            first["latgrad0"] = lat.weight.grad.numpy().copy()
            first["table_after_renorm0"] = lat.weight.data.numpy().copy()
            for k, v in dec.named_parameters():
                first["gradnorm0/" + k] = np.array(float(v.grad.norm()))
            first["grad0/lin4.bias"] = dec.lin4.bias.grad.numpy().copy()
            first["grad0/lin8.weight"] = dec.lin8.weight.grad.numpy().copy()
        gn = torch.nn.utils.clip_grad_norm_(dec.parameters(), 1.0)
        opt.step()
        losses.append(float(loss)); regs.append(float(reg)); gnorms.append(float(gn)); ids_all.append(ids.numpy())
    out = {"lat0": lat0.numpy(), "ids": np.stack(ids_all), "loss": np.array(losses),
           "reg": np.array(regs), "gnorm": np.array(gnorms), "lat_final": lat.weight.data.numpy(),
           "S": np.array(S), "B": np.array(B), "P": np.array(P)}
    out.update(first)
    for k, v in checksums(dec).items():
        out["cks_final/" + k] = v
    np.savez_compressed(os.path.join(HERE, name), **out)


def golden_grid():
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    out = {}
    for N in (8, 64, 256, 512):
        vs = 2.0 / (N - 1)
        idx = torch.arange(0, N ** 3, 1, out=torch.LongTensor())
        s = torch.zeros(N ** 3, 3)
        s[:, 2] = idx % N
        s[:, 1] = (idx.long() / N) % N
        s[:, 0] = ((idx.long() / N) / N) % N
        s[:, 0] = (s[:, 0] * vs) + -1
        s[:, 1] = (s[:, 1] * vs) + -1
        s[:, 2] = (s[:, 2] * vs) + -1
        if N == 8:
            out["coords8"] = s.numpy()
        # order-independent bit checksum + a strided sample
        bits = s.numpy().view(np.uint32).astype(np.uint64)
        out[f"xor{N}"] = np.bitwise_xor.reduce(bits, axis=0)
        out[f"sum{N}"] = bits.sum(axis=0)
        out[f"sample{N}"] = s[:: max(1, N ** 3 // 4096)].numpy()
    np.savez_compressed(os.path.join(HERE, "grid.npz"), **out)


if __name__ == "__main__":
    golden_forward(False, "decoder_plain.npz")
    golden_forward(True, "decoder_wn.npz")
    golden_train(True, "train_wn.npz")
    golden_train(False, "train_plain.npz")
    golden_grid()
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))

"""Golden fixture for the Eikonal term of the Stage-1 step, from the LIVE reference decoder (build container only):

    python tests/golden/make_golden_eikonal.py

Drives networks.deep_sdf_decoder.Decoder (imported from /root/reference) with the reference's own statements
for the step body -- train_deep_sdf.py:555-616 and, for the term under test, :724-729 verbatim:

    grad_outputs = torch.ones_like(pred_sdf, requires_grad=True)
    gradients = torch.autograd.grad(pred_sdf, [xyz[i]], grad_outputs=grad_outputs, create_graph=True,
                                    allow_unused=True, retain_graph=True)[0]
    eikonal_loss = 0.002 * ((1. - torch.linalg.vector_norm(gradients, dim=1))**2).mean()

Every shipped Stage-1 specs.json sets UseEikonal: true with norm_layers [] (plain Linear layers), so the plain
variant is the primary fixture; the weight-normalised one is stored as well.  Stored: the three loss terms,
per-row ||d pred / d xyz||, per-parameter gradient norms with and without the Eikonal term, all small gradients
in full (biases, lin8.weight, the xyz columns of lin0/lin4, the latent gradient) and a strided sample of every
weight matrix gradient.
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
STRIDE = 97


def one_step(wn: bool, use_eikonal: bool, S=4, B=2, P=256, epoch=50):
    torch.manual_seed(0)
    decoder = RefDecoder(256, norm_layers=list(range(8)) if wn else [], **SPECS)
    torch.manual_seed(123)
    lat_vecs = torch.nn.Embedding(S, 256, max_norm=1.0)
    torch.nn.init.normal_(lat_vecs.weight.data, 0.0, 1.0 / math.sqrt(256))
    loss_l1 = torch.nn.L1Loss(reduction="sum")
    data = synth.scene_samples(S, P, seed=1)
    ids = torch.tensor([2, 0])[:B]
    minT, maxT = -0.1, 0.1
    code_reg_lambda = 1e-4
    # ---- train_deep_sdf.py:555-577
    sdf_data = data[ids].reshape(-1, 4)
    num_sdf_samples = sdf_data.shape[0]
    sdf_data.requires_grad = False
    xyz = sdf_data[:, 0:3]
    xyz.requires_grad = True
    sdf_gt = sdf_data[:, 3].unsqueeze(1)
    sdf_gt = torch.clamp(sdf_gt, minT, maxT)
    indices = ids.unsqueeze(-1).repeat(1, P).view(-1)
    # ---- :598-616
    batch_vecs = lat_vecs(indices)
    input = torch.cat([batch_vecs, xyz], dim=1)
    pred_sdf = decoder(input)
    pred_sdf = torch.clamp(pred_sdf, minT, maxT)
    chunk_loss = loss_l1(pred_sdf, sdf_gt) / num_sdf_samples
    sdf_loss = float(chunk_loss.item())
    l2_size_loss = torch.sum(torch.norm(batch_vecs, dim=1))
    reg_loss = (code_reg_lambda * min(1, epoch / 100) * l2_size_loss) / num_sdf_samples
    chunk_loss = chunk_loss + reg_loss
    out = {}
    eik = 0.0
    if use_eikonal:
        # ---- :724-729
        grad_outputs = torch.ones_like(pred_sdf, requires_grad=True)
        gradients = torch.autograd.grad(pred_sdf, [xyz], grad_outputs=grad_outputs, create_graph=True,
                                        allow_unused=True, retain_graph=True)[0]
        eikonal_loss = 0.002 * ((1. - torch.linalg.vector_norm(gradients, dim=1)) ** 2).mean()
        chunk_loss += eikonal_loss
        eik = float(eikonal_loss.item())
        out["gradnorm_xyz"] = torch.linalg.vector_norm(gradients, dim=1).detach().numpy()
    chunk_loss.backward()          # :731
    out["sdf_loss"] = np.array(sdf_loss)
    out["reg_loss"] = np.array(float(reg_loss))
    out["eikonal_loss"] = np.array(eik)
    out["pred"] = pred_sdf.detach().numpy()
    out["latgrad"] = lat_vecs.weight.grad.numpy().copy()
    for k, v in decoder.named_parameters():
        g = v.grad
        out["gnorm/" + k] = np.array(float(g.norm()))
        if g.numel() <= 4096:
            out["g/" + k] = g.numpy().copy()
        else:
            out["gs/" + k] = g.reshape(-1)[::STRIDE].numpy().copy()
    wname = "weight_v" if wn else "weight"
    out["gxyz/lin0"] = getattr(decoder.lin0, wname).grad[:, 256:259].numpy().copy()
    out["gxyz/lin4"] = getattr(decoder.lin4, wname).grad[:, 509:512].numpy().copy()
    return out, lat_vecs.weight.data.numpy().copy(), ids.numpy()


def main():
    for wn, name in ((False, "train_eikonal_plain.npz"), (True, "train_eikonal_wn.npz")):
        with_e, lat0, ids = one_step(wn, True)
        without, _, _ = one_step(wn, False)
        out = {"lat0": lat0, "ids": ids, "S": np.array(4), "B": np.array(2), "P": np.array(256),
               "epoch": np.array(50), "stride": np.array(STRIDE)}
        for k, v in with_e.items():
            out["eik/" + k] = v
        for k, v in without.items():
            out["base/" + k] = v
        np.savez_compressed(os.path.join(HERE, name), **out)
        print(name, os.path.getsize(os.path.join(HERE, name)), "eikonal", float(with_e["eikonal_loss"]),
              "sdf", float(with_e["sdf_loss"]))


if __name__ == "__main__":
    main()

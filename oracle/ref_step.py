"""The reference's Stage-1 step loop around the reference's OWN Decoder class (TEST / BASELINE INFRASTRUCTURE ONLY).

`oracle/_ref/networks/deep_sdf_decoder.py` is the unmodified reference module (oracle/build_ref.py).  The statements
below are train_deep_sdf.py's, in order, with `.cuda()` replaced by a device argument and file I/O / logging removed:
    :429-435  lat_vecs = nn.Embedding(num_scenes, latent, max_norm=CodeBound); normal_(0, std / sqrt(latent))
    :443-454  loss_l1 = L1Loss(reduction="sum"); optimizer_all = Adam([{decoder}, {lat_vecs}])
    :555-577  reshape, xyz / sdf_gt split, clamp, index expansion
    :598-616  gather, cat, decoder, clamp, L1 / N, code regulariser
    :724-729  Eikonal term (create_graph=True)
    :731      backward
    :780-784  clip_grad_norm_, optimizer_all.step()
Used by bench.py as the timed CPU baseline (kind "reference") and by tests/test_oracle_cpu.py to check that the
oracle port (oracle/train.py) and this loop produce the same numbers.
"""
from __future__ import annotations

import math
import os
import sys
import warnings

import torch

_REF = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available() -> bool:
    return os.path.isfile(os.path.join(_REF, "networks", "deep_sdf_decoder.py"))


def reference_decoder_class():
    """networks.deep_sdf_decoder.Decoder of the reference, imported from oracle/_ref under a private module name."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_dsdf_ref_deep_sdf_decoder",
                                                  os.path.join(_REF, "networks", "deep_sdf_decoder.py"))
    mod = importlib.util.module_from_spec(spec)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        spec.loader.exec_module(mod)
    return mod.Decoder


class ReferenceStepLoop:
    def __init__(self, num_scenes: int, network_specs: dict, latent_size: int = 256, code_bound=1.0, grad_clip=1.0,
                 clamp_dist: float = 0.1, code_reg_lambda: float = 1e-4, use_eikonal: bool = False,
                 lr=(5e-4, 1e-3), code_init_std: float = 1.0, seed: int = 0, device: str = "cpu"):
        Decoder = reference_decoder_class()
        torch.manual_seed(seed)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            self.decoder = Decoder(latent_size, **network_specs).to(device)
        self.lat_vecs = torch.nn.Embedding(num_scenes, latent_size, max_norm=code_bound)
        torch.nn.init.normal_(self.lat_vecs.weight.data, 0.0, code_init_std / math.sqrt(latent_size))
        self.lat_vecs = self.lat_vecs.to(device)
        self.loss_l1 = torch.nn.L1Loss(reduction="sum")
        self.optimizer_all = torch.optim.Adam([
            {"params": self.decoder.parameters(), "lr": lr[0]},
            {"params": self.lat_vecs.parameters(), "lr": lr[1]},
        ])
        self.grad_clip, self.clamp_dist, self.code_reg_lambda = grad_clip, clamp_dist, code_reg_lambda
        self.use_eikonal, self.device = use_eikonal, device

    def step(self, sdf_data: torch.Tensor, indices: torch.Tensor, epoch: int):
        minT, maxT = -self.clamp_dist, self.clamp_dist
        num_samp_per_scene = sdf_data.shape[1]
        sdf_data = sdf_data.reshape(-1, 4).to(self.device)
        indices = indices.to(self.device)
        num_sdf_samples = sdf_data.shape[0]
        sdf_data.requires_grad = False
        xyz = sdf_data[:, 0:3]
        xyz.requires_grad = True
        sdf_gt = sdf_data[:, 3].unsqueeze(1)
        sdf_gt = torch.clamp(sdf_gt, minT, maxT)
        indices = indices.unsqueeze(-1).repeat(1, num_samp_per_scene).view(-1)
        self.optimizer_all.zero_grad()
        batch_vecs = self.lat_vecs(indices)
        input = torch.cat([batch_vecs, xyz], dim=1)
        pred_sdf = self.decoder(input)
        pred_sdf = torch.clamp(pred_sdf, minT, maxT)
        chunk_loss = self.loss_l1(pred_sdf, sdf_gt) / num_sdf_samples
        out = {"sdf_loss": chunk_loss.item()}
        l2_size_loss = torch.sum(torch.norm(batch_vecs, dim=1))
        reg_loss = (self.code_reg_lambda * min(1, epoch / 100) * l2_size_loss) / num_sdf_samples
        chunk_loss = chunk_loss + reg_loss
        out["reg_loss"] = reg_loss.item()
        if self.use_eikonal:
            grad_outputs = torch.ones_like(pred_sdf, requires_grad=True)
            gradients = torch.autograd.grad(pred_sdf, [xyz], grad_outputs=grad_outputs, create_graph=True,
                                            allow_unused=True, retain_graph=True)[0]
            eikonal_loss = 0.002 * ((1. - torch.linalg.vector_norm(gradients, dim=1)) ** 2).mean()
            chunk_loss += eikonal_loss
            out["eikonal_loss"] = eikonal_loss.item()
        chunk_loss.backward()
        if self.grad_clip is not None:
            out["grad_norm"] = float(torch.nn.utils.clip_grad_norm_(self.decoder.parameters(), self.grad_clip, norm_type=2))
        self.optimizer_all.step()
        return out

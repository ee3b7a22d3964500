"""Oracle restatement of one Stage-1 auto-decoder optimisation step (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/train_deep_sdf.py:
  * latent table + init       :429-435  nn.Embedding(S, L, max_norm=CodeBound), N(0,(std/sqrt(L))^2)
  * optimizer                 :445-454  one Adam over {decoder params, lr0} and {latents, lr1}
  * LR schedule               :263-265  (deep_sdf/lr_scheduling.py Step schedule restated below)
  * step prologue             :555-577  reshape, gt clamp, scene-id expansion, batch_split chunks
  * chunk body                :594-616  gather (renorm side effect) -> cat -> decoder -> clamp ->
                                        L1(sum)/N_total + lambda*min(1,epoch/100)*sum||z||/N_total
  * Eikonal term              :286,563,724-729  0.002*mean((1-||d clamp(pred)/d xyz||)^2), create_graph=True
  * backward                  :731
  * clip + step               :780-784  clip_grad_norm_(decoder.parameters(), c) then Adam step

Pinned against the EXECUTED script: tests/golden/make_golden_trainloop.py runs train_deep_sdf.main_function itself on CPU
(import stubs, .cuda() -> identity) and tests/test_oracle_cpu.py requires this restatement to reproduce its per-batch
losses and final checkpoint from the same initial state and batches, with and without UseEikonal.

Everything is torch fp32 and device agnostic (the script itself hard-codes .cuda()).
torch's own Embedding/Adam/clip_grad_norm_ are used where the reference uses them
(third-party arithmetic: the reference pins pytorch 1.13, this container has 2.11; the
semantics relied on were re-verified), plus an explicit restatement
(`adam_update`, `clip_coef`, `renorm_rows`) that the tests cross-check against torch.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

from . import deepsdf, siren


@dataclass
class TrainConfig:
    samples_per_scene: int = 16384
    scenes_per_batch: int = 64
    clamp_dist: float = 0.1
    code_reg: bool = True
    code_reg_lambda: float = 1e-4
    code_bound: Optional[float] = 1.0
    grad_clip: Optional[float] = 1.0
    code_init_std: float = 1.0
    lr_decoder: float = 5e-4
    lr_latent: float = 1e-3
    lr_interval: int = 500
    lr_factor: float = 0.5
    batch_split: int = 1
    use_eikonal: bool = False        # specs["UseEikonal"] (train_deep_sdf.py:286)


def step_lr(initial: float, interval: int, factor: float, epoch: int) -> float:
    """deep_sdf/lr_scheduling.py:16-24 (StepLearningRateSchedule)."""
    return initial * (factor ** (epoch // interval))


def renorm_rows(table: torch.Tensor, ids: torch.Tensor, max_norm: float) -> None:
    """Embedding(max_norm) side effect: rows that are looked up and exceed the bound are
    scaled in place by max_norm/(norm+1e-7); other rows untouched (SURVEY.md App. A probe)."""
    with torch.no_grad():
        u = torch.unique(ids)
        rows = table[u]
        n = rows.norm(dim=1)
        scale = torch.where(n > max_norm, max_norm / (n + 1e-7), torch.ones_like(n))
        table[u] = rows * scale[:, None]


def clip_coef(total_norm: torch.Tensor, max_norm: float) -> torch.Tensor:
    """clip_grad_norm_: coef = max_norm/(total+1e-6), clamped to 1."""
    return torch.clamp(max_norm / (total_norm + 1e-6), max=1.0)


def adam_update(p, g, m, v, step: int, lr: float, b1=0.9, b2=0.999, eps=1e-8):
    """torch.optim.Adam (amsgrad=False, wd=0) single-tensor math, in place."""
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-(lr / bc1))


class OracleTrainer:
    """The reference's training state + step, single process, any device."""

    def __init__(self, spec: deepsdf.DecoderSpec, cfg: TrainConfig, num_scenes: int,
                 seed: int = 0, device: str = "cpu", params=None, latents=None,
                 arch: str = "deep_sdf_decoder"):
        # arch selects the decoder the step loop wraps (specs["NetworkArch"], train_deep_sdf.py:225,348)
        self.spec, self.cfg, self.device, self.arch = spec, cfg, torch.device(device), arch
        if params is None:
            params = deepsdf.init_params(spec, seed) if arch == "deep_sdf_decoder" else siren.init_params(seed)
        self.params = {k: torch.nn.Parameter(v.clone().to(self.device)) for k, v in params.items()}
        # train_deep_sdf.py:429-435
        self.lat_vecs = torch.nn.Embedding(num_scenes, spec.latent_size, max_norm=cfg.code_bound)
        if latents is None:
            torch.nn.init.normal_(self.lat_vecs.weight.data, 0.0,
                                  cfg.code_init_std / math.sqrt(spec.latent_size))
        else:
            self.lat_vecs.weight.data.copy_(latents)
        self.lat_vecs = self.lat_vecs.to(self.device)
        # train_deep_sdf.py:445-454
        self.opt = torch.optim.Adam([
            {"params": list(self.params.values()), "lr": cfg.lr_decoder},
            {"params": list(self.lat_vecs.parameters()), "lr": cfg.lr_latent},
        ])

    def set_epoch_lr(self, epoch: int) -> None:
        c = self.cfg
        self.opt.param_groups[0]["lr"] = step_lr(c.lr_decoder, c.lr_interval, c.lr_factor, epoch)
        self.opt.param_groups[1]["lr"] = step_lr(c.lr_latent, c.lr_interval, c.lr_factor, epoch)

    def forward_loss(self, sdf_data: torch.Tensor, scene_ids: torch.Tensor, epoch: int,
                     n_total: Optional[int] = None, pred_noise: float = 0.0, dropout_keep=None):
        """Gather + forward + loss for all chunks, calling backward per chunk like the
        reference (:594-731).  sdf_data [B,P,4] (or [M,4] with P=cfg.samples_per_scene)."""
        c = self.cfg
        sdf_data = sdf_data.reshape(-1, 4).to(self.device)
        scene_ids = scene_ids.to(self.device)
        m = sdf_data.shape[0]
        n_total = m if n_total is None else n_total      # F9: global row count
        p = m // scene_ids.numel()
        xyz = sdf_data[:, 0:3]
        if c.use_eikonal:
            xyz = xyz.detach().clone().requires_grad_(True)      # :562-563 (xyz.requires_grad = True)
        gt = torch.clamp(sdf_data[:, 3:4], -c.clamp_dist, c.clamp_dist)
        idx = scene_ids.unsqueeze(-1).repeat(1, p).view(-1)
        xyz_c = torch.chunk(xyz, c.batch_split)
        idx_c = torch.chunk(idx, c.batch_split)
        gt_c = torch.chunk(gt, c.batch_split)
        sdf_loss = 0.0
        reg_loss = 0.0
        self.last_eikonal_loss = 0.0
        preds = []
        for i in range(c.batch_split):
            bv = self.lat_vecs(idx_c[i])
            inp = torch.cat([bv, xyz_c[i]], dim=1)
            assert dropout_keep is None or c.batch_split == 1
            if self.arch == "siren_decoder":
                pred = siren.forward(self.params, inp)
            else:
                pred = deepsdf.forward(self.params, self.spec, inp, training=True, dropout_keep=dropout_keep)
            if pred_noise:
                # test diagnostic only (not reference behaviour): uniform +-pred_noise on the SDF
                # output, to measure how far the clamped-L1 gradient moves under an output
                # perturbation far inside the 2e-3 parity tolerance (sign(pred-gt) is discontinuous)
                gen = torch.Generator(device=pred.device).manual_seed(1234 + i)
                pred = pred + pred_noise * (2 * torch.rand(pred.shape, generator=gen, device=pred.device) - 1)
            pred = torch.clamp(pred, -c.clamp_dist, c.clamp_dist)
            chunk_loss = torch.nn.functional.l1_loss(pred, gt_c[i], reduction="sum") / n_total
            sdf_loss += float(chunk_loss.item())
            if c.code_reg:
                l2 = torch.sum(torch.norm(bv, dim=1))
                reg = (c.code_reg_lambda * min(1, epoch / 100) * l2) / n_total
                chunk_loss = chunk_loss + reg
                reg_loss += float(reg.item())
            if c.use_eikonal:
                # train_deep_sdf.py:724-729, verbatim arithmetic (pred is the CLAMPED prediction)
                grad_outputs = torch.ones_like(pred, requires_grad=True)
                gradients = torch.autograd.grad(pred, [xyz_c[i]], grad_outputs=grad_outputs, create_graph=True,
                                                allow_unused=True, retain_graph=True)[0]
                eikonal_loss = 0.002 * ((1. - torch.linalg.vector_norm(gradients, dim=1)) ** 2).mean()
                chunk_loss = chunk_loss + eikonal_loss
                self.last_eikonal_loss += float(eikonal_loss.item())
            chunk_loss.backward()
            preds.append(pred.detach())
        return sdf_loss, reg_loss, torch.cat(preds, 0)

    def step(self, sdf_data: torch.Tensor, scene_ids: torch.Tensor, epoch: int,
             n_total: Optional[int] = None, apply: bool = True) -> Dict[str, float]:
        self.set_epoch_lr(epoch)
        self.opt.zero_grad()
        sdf_loss, reg_loss, pred = self.forward_loss(sdf_data, scene_ids, epoch, n_total)
        gnorm = None
        if self.cfg.grad_clip is not None:
            gnorm = float(torch.nn.utils.clip_grad_norm_(list(self.params.values()),
                                                         self.cfg.grad_clip, norm_type=2))
        if apply:
            self.opt.step()
        return {"loss": sdf_loss + reg_loss + self.last_eikonal_loss, "sdf_loss": sdf_loss, "reg_loss": reg_loss,
                "eikonal_loss": self.last_eikonal_loss, "grad_norm": gnorm, "pred": pred}

    # convenience for tests
    def decoder_grads(self) -> Dict[str, torch.Tensor]:
        return {k: v.grad.detach().clone() for k, v in self.params.items()}

    def latent_grad(self) -> torch.Tensor:
        return self.lat_vecs.weight.grad.detach().clone()

"""Fused auto-decoder training step: the body of the reference's hot loop
(train_deep_sdf.py:551-616, 731, 780-784) as a handful of CUDA launches through the C ABI.

    trainer = FusedTrainer(decoder, num_scenes, StepConfig(...))
    for epoch ...:
        for sdf_data[B,P,4], indices[B] in loader:        # same tensors the reference loop sees
            stats = trainer.step(sdf_data, indices, epoch)

State that the reference checkpoints (train_deep_sdf.py:32-79) stays in torch tensors with the
reference's shapes: `decoder.state_dict()`, `trainer.lat_vecs` (an nn.Embedding with
max_norm=CodeBound) and Adam moments addressable in torch.optim.Adam's state_dict format
(`optimizer_state_dict()`).

Multi-GPU (one process per GPU): scenes are owned by ranks; every rank steps on its own scenes'
samples and latent codes (latent grads never leave the rank) and only the flat decoder gradient
is all-reduced (SUM, F9: the loss is normalised by the *global* row count).
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Dict, Optional

import torch

from . import _lib
from .decoder import LATENT, Decoder, _stream_ptr


# loss switches of train_deep_sdf.py:286-327 that the fused step does not compute.  `UseEikonal` is built
# (fused second-order path, StepConfig.use_eikonal).  The two latent-only priors act on the [B, 256] batch
# latents, not on the decoder path: they can be supplied as `aux_latent_loss` to FusedTrainer.step.
_UNSUPPORTED_LOSS_KEYS = ("UseIsometryLoss", "UseGradMetricIsotropyLoss")
_LATENT_ONLY_LOSS_KEYS = ("UseCovarianceLoss", "UseGMMPriorLoss")


@dataclass
class StepConfig:
    """Subset of specs.json read by the step (train_deep_sdf.py:243-346)."""
    clamp_dist: float = 0.1                 # ClampingDistance
    code_reg: bool = True                   # CodeRegularization
    code_reg_lambda: float = 1e-4           # CodeRegularizationLambda
    code_bound: Optional[float] = None      # CodeBound
    grad_clip: Optional[float] = None       # GradientClipNorm
    code_init_std: float = 1.0              # CodeInitStdDev
    lr_decoder: float = 5e-4                # LearningRateSchedule[0] (Step)
    lr_latent: float = 1e-3                 # LearningRateSchedule[1] (Step)
    lr_interval: int = 500                  # LearningRateSchedule[0].Interval
    lr_factor: float = 0.5                  # LearningRateSchedule[0].Factor
    lr_interval_latent: Optional[int] = None    # LearningRateSchedule[1].Interval (None: same as [0])
    lr_factor_latent: Optional[float] = None    # LearningRateSchedule[1].Factor
    betas: tuple = (0.9, 0.999)
    eps: float = 1e-8
    dropout_seed: int = 0                   # key of the Philox dropout stream (NetworkSpecs.dropout)
    use_eikonal: bool = False               # UseEikonal (train_deep_sdf.py:286,724-729)
    eikonal_lambda: float = 0.002           # hard-coded in the reference (:727)

    @staticmethod
    def from_specs(specs: dict, latent_only_losses_handled: bool = False) -> "StepConfig":
        """Read the step's keys the way train_deep_sdf.py does (get_spec_with_default, :191-195).  A switch
        that changes the objective and is not computed by the fused step raises, naming the key (SURVEY.md
        H13): training a different loss silently is worse than not training.  `latent_only_losses_handled`:
        the caller passes the covariance / GMM prior term as `aux_latent_loss` to FusedTrainer.step."""
        for k in _UNSUPPORTED_LOSS_KEYS:
            if specs.get(k, False):
                raise NotImplementedError(
                    f"specs[{k!r}] = true: this loss back-propagates through decoder Jacobians "
                    "(deep_sdf/loss.py) and is not part of the fused step; use the reference NetworkArch for it")
        for k in _LATENT_ONLY_LOSS_KEYS:
            if specs.get(k, False) and not latent_only_losses_handled:
                raise NotImplementedError(
                    f"specs[{k!r}] = true: a prior on the batch latents (train_deep_sdf.py:737-764), outside the "
                    "decoder path; compute it in torch on trainer.lat_vecs rows and pass it to "
                    "FusedTrainer.step(aux_latent_loss=...), then call from_specs(..., latent_only_losses_handled=True)")
        sch = specs["LearningRateSchedule"]
        for i, s in enumerate(sch[:2]):
            if s["Type"] != "Step":
                raise NotImplementedError(f"LearningRateSchedule[{i}].Type = {s['Type']!r}: only 'Step' is built in; "
                                          "pass lr_decoder / lr_latent explicitly to step()")
        return StepConfig(
            clamp_dist=specs["ClampingDistance"],
            code_reg=specs.get("CodeRegularization", True),
            code_reg_lambda=specs.get("CodeRegularizationLambda", 1e-4),
            code_bound=specs.get("CodeBound", None),
            grad_clip=specs.get("GradientClipNorm", None),
            code_init_std=specs.get("CodeInitStdDev", 1.0),
            lr_decoder=sch[0]["Initial"], lr_latent=sch[1]["Initial"],
            lr_interval=sch[0]["Interval"], lr_factor=sch[0]["Factor"],
            lr_interval_latent=sch[1]["Interval"], lr_factor_latent=sch[1]["Factor"],
            use_eikonal=bool(specs.get("UseEikonal", False)),
        )


class FusedTrainer:
    def __init__(self, decoder: Decoder, num_scenes: int, cfg: StepConfig,
                 process_group=None, latents: Optional[torch.Tensor] = None):
        self.decoder = decoder
        self.cfg = cfg
        self.num_scenes = num_scenes
        self.pg = process_group
        dev = next(decoder.parameters()).device
        if dev.type != "cuda":
            raise _lib.DsdfError("FusedTrainer needs the decoder on a CUDA device (no CPU path)")
        self.device = dev
        # train_deep_sdf.py:429-435
        self.lat_vecs = torch.nn.Embedding(num_scenes, LATENT, max_norm=cfg.code_bound)
        if latents is None:
            torch.nn.init.normal_(self.lat_vecs.weight.data, 0.0, cfg.code_init_std / math.sqrt(LATENT))
        else:
            self.lat_vecs.weight.data.copy_(latents)
        self.lat_vecs = self.lat_vecs.to(dev)
        flat = decoder.flat_params()
        self.grads = torch.zeros_like(flat)
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.lat_grad = torch.zeros_like(self.lat_vecs.weight.data)
        self.lat_exp_avg = torch.zeros_like(self.lat_grad)
        self.lat_exp_avg_sq = torch.zeros_like(self.lat_grad)
        self.step_count = 0
        self.scratch = torch.zeros(4096, dtype=torch.float32, device=dev)
        self.scratch_lat = torch.zeros(4096, dtype=torch.float32, device=dev)
        self.loss_buf = torch.zeros(4, dtype=torch.float32, device=dev)
        self._ws = None
        self._ws_key = None
        self._pred = None
        self.dropout_row_base = 0               # global index of this rank's first row (sharded runs)
        self._last_ids = None
        # scene-sharded runs: the decoder-gradient all-reduce goes to a side stream as soon as the gradient is final
        # (event recorded inside dsdf_train_fwd_bwd), under the latent-gradient kernels and the latent Adam
        self._comm_stream = None
        self._grads_ready = None
        self.time_comm = False                  # bench.py: record CUDA events around the collective
        self.comm_events = []
        if process_group is not None and torch.distributed.get_world_size(process_group) > 1:
            self._comm_stream = torch.cuda.Stream(device=dev)
            self._grads_ready = torch.cuda.Event()
            self._grads_ready.record(torch.cuda.current_stream(dev))     # instantiates the cudaEvent_t handle
        if cfg.use_eikonal and decoder._arch != "deep_sdf_decoder":
            raise NotImplementedError("specs['UseEikonal'] = true is built for NetworkArch deep_sdf_decoder only "
                                      "(sine layers have a non-zero second derivative)")
        if cfg.use_eikonal and decoder.dropout and decoder.dropout_prob > 0:
            raise NotImplementedError("specs['UseEikonal'] together with a non-empty NetworkSpecs.dropout is not built")

    # ---------------------------------------------------------------- helpers
    def lr(self, which: int, epoch: int) -> float:
        # one StepLearningRateSchedule per parameter group (train_deep_sdf.py:247-265, lr_scheduling.py:16-24)
        c = self.cfg
        if which == 0:
            init, interval, factor = c.lr_decoder, c.lr_interval, c.lr_factor
        else:
            init = c.lr_latent
            interval = c.lr_interval if c.lr_interval_latent is None else c.lr_interval_latent
            factor = c.lr_factor if c.lr_factor_latent is None else c.lr_factor_latent
        return init * (factor ** (epoch // interval))

    def _workspace(self, h, rows: int, segs: int):
        key = (rows, segs)
        if self._ws_key != key:
            nbytes = self.decoder.L.dsdf_train_workspace_bytes(h, rows, segs)
            self._ws = None
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            self._pred = torch.empty(rows, dtype=torch.float32, device=self.device)
            self._ws_key = key
        return self._ws

    # ---------------------------------------------------------------- the step
    def forward_backward(self, sdf_data: torch.Tensor, indices: torch.Tensor, epoch: int,
                         n_total: Optional[int] = None):
        """gather + forward + loss + backward; fills self.grads / self.lat_grad (no optimiser)."""
        c = self.cfg
        dec = self.decoder
        L = dec.L
        h = dec.sync_weights()
        flat = dec.flat_params()
        if self.grads.data_ptr() == 0 or self.grads.numel() != flat.numel():
            raise RuntimeError("decoder parameters were re-laid out after the trainer was built")
        if not sdf_data.is_cuda:
            raise _lib.DsdfError("sdf_data must already be on the GPU (there is no CPU path)")
        sdf = sdf_data.reshape(-1, 4).contiguous().float()
        ids = indices.reshape(-1).contiguous().to(torch.int64)
        n_seg = ids.numel()
        rows = sdf.shape[0]
        if n_seg == 0 or rows % n_seg != 0:
            raise ValueError(f"{rows} sample rows do not split evenly over {n_seg} scenes "
                             "(train_deep_sdf.py:571-577 repeats every scene id SamplesPerScene times)")
        p = rows // n_seg
        # nn.Embedding raises on an index outside the table (train_deep_sdf.py:598).  Host-resident ids (what the
        # reference's DataLoader yields) are checked here for free; device-resident ids are checked by the step
        # itself (loss_buf[3], see check_status) so that no host sync is forced on the hot path.
        if not ids.is_cuda:
            if n_seg and (int(ids.min()) < 0 or int(ids.max()) >= self.num_scenes):
                raise IndexError(f"scene index out of range [0, {self.num_scenes}): "
                                 f"min {int(ids.min())}, max {int(ids.max())}")
            ids = ids.to(self.device, non_blocking=True)
        if n_total is None:
            n_total = rows
            if self.pg is not None:
                n_total = rows * torch.distributed.get_world_size(self.pg)
        ws = self._workspace(h, rows, n_seg)
        st = _stream_ptr(self.device)
        table = self.lat_vecs.weight.data
        with torch.cuda.device(self.device):
            # decoder.train() + NetworkSpecs.dropout: F.dropout after the listed ReLUs
            # (deep_sdf_decoder.py:103-104), fused into the forward epilogues; a new mask per step
            drop_layers = 0
            if dec.training and dec.dropout and dec.dropout_prob > 0:
                for l in dec.dropout:
                    drop_layers |= 1 << int(l)
            _lib.check(L.dsdf_set_dropout(h, drop_layers, float(dec.dropout_prob), int(c.dropout_seed),
                                          self.step_count, int(self.dropout_row_base)), h, L)
            if c.code_bound is not None:
                _lib.check(L.dsdf_renorm_rows(h, table.data_ptr(), self.num_scenes, ids.data_ptr(), n_seg,
                                              float(c.code_bound), st), h, L)
            ramp = c.code_reg_lambda * min(1, epoch / 100) if c.code_reg else 0.0
            _lib.check(L.dsdf_train_fwd_bwd(
                h, flat.data_ptr(), table.data_ptr(), self.num_scenes, ids.data_ptr(), n_seg, p,
                sdf.data_ptr(), int(n_total), float(c.clamp_dist), float(ramp),
                float(c.eikonal_lambda) if c.use_eikonal else 0.0, ws.data_ptr(),
                ws.numel(), self.grads.data_ptr(), self.lat_grad.data_ptr(),
                self._pred.data_ptr(), self.loss_buf.data_ptr(),
                self._grads_ready.cuda_event if self._grads_ready is not None else None, st), h, L)
        self._last_ids = ids
        return self._pred

    def check_status(self) -> None:
        """Raise if the last step saw a scene index outside [0, num_scenes) (one D2H read; nn.Embedding's
        IndexError / device-side assert).  The step itself ran with that id replaced by 0."""
        bad = int(self.loss_buf[3].item())
        if bad:
            raise IndexError(f"scene index out of range [0, {self.num_scenes}) at batch position {bad - 1}")

    def step(self, sdf_data: torch.Tensor, indices: torch.Tensor, epoch: int,
             n_total: Optional[int] = None, lr_decoder: Optional[float] = None,
             lr_latent: Optional[float] = None, aux_latent_loss=None) -> Dict[str, torch.Tensor]:
        """One optimisation step.  `aux_latent_loss(latents[U, 256]) -> scalar`: an extra torch loss on the rows of
        the UNIQUE batch scenes (the reference's covariance / GMM prior terms, train_deep_sdf.py:737-764); its
        gradient is added to the latent gradient before Adam (its own parameters, if any, are the caller's)."""
        L = self.decoder.L
        c = self.cfg
        self.forward_backward(sdf_data, indices, epoch, n_total)
        if aux_latent_loss is not None:
            uniq = torch.unique(self._last_ids)
            z = self.lat_vecs.weight.data.index_select(0, uniq).requires_grad_(True)
            with torch.enable_grad():
                aux = aux_latent_loss(z)
            (gz,) = torch.autograd.grad(aux, [z], allow_unused=True)
            if gz is not None:
                self.lat_grad.index_add_(0, uniq, gz.to(self.lat_grad.dtype))
        self.step_count += 1
        h = self.decoder.handle()
        flat = self.decoder.flat_params()
        st = _stream_ptr(self.device)
        cur = torch.cuda.current_stream(self.device)
        lr0 = self.lr(0, epoch) if lr_decoder is None else lr_decoder
        lr1 = self.lr(1, epoch) if lr_latent is None else lr_latent
        b1, b2 = c.betas
        work = None
        if self._comm_stream is not None:
            # only the ~1.84 M-element decoder gradient crosses NVLink; SUM, not mean (F9: the loss is normalised
            # by the GLOBAL row count).  Issued on the side stream behind the grads-ready event, so it runs under
            # the rest of the backward (latent-gradient reduction) and the latent Adam below.
            with torch.cuda.stream(self._comm_stream):
                self._comm_stream.wait_event(self._grads_ready)
                if self.time_comm:                             # measurement only (bench.py): collective start / end
                    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                    e0.record(self._comm_stream)
                work = torch.distributed.all_reduce(self.grads, op=torch.distributed.ReduceOp.SUM, group=self.pg,
                                                    async_op=True)
        with torch.cuda.device(self.device):
            table = self.lat_vecs.weight.data
            _lib.check(L.dsdf_clip_adam(
                h, table.data_ptr(), self.lat_grad.data_ptr(), self.lat_exp_avg.data_ptr(),
                self.lat_exp_avg_sq.data_ptr(), table.numel(), 0.0, float(lr1), b1, b2, c.eps,
                self.step_count, self.scratch_lat.data_ptr(), st), h, L)
            if work is not None:
                with torch.cuda.stream(self._comm_stream):
                    work.wait()                               # comm stream waits for the collective ...
                    if self.time_comm:
                        e1.record(self._comm_stream)
                        self.comm_events.append((e0, e1))
                cur.wait_stream(self._comm_stream)            # ... and the step's stream for the comm stream
            _lib.check(L.dsdf_clip_adam(
                h, flat.data_ptr(), self.grads.data_ptr(), self.exp_avg.data_ptr(),
                self.exp_avg_sq.data_ptr(), flat.numel(),
                float(c.grad_clip) if c.grad_clip is not None else 0.0, float(lr0), b1, b2, c.eps,
                self.step_count, self.scratch.data_ptr(), st), h, L)
        self.decoder.sync_weights(force=True)
        # device scalars, no host sync: [0] L1 term, [1] regulariser term, [2] Eikonal term (this rank's share),
        # [3] scene-id status (check_status).  The gradient norm is only computed when the step clips
        # (train_deep_sdf.py:780-782), otherwise None.
        return {"sdf_loss": self.loss_buf[0], "reg_loss": self.loss_buf[1], "eikonal_loss": self.loss_buf[2],
                "status": self.loss_buf[3],
                "grad_norm": self.scratch[0] if c.grad_clip is not None else None}

    # ---------------------------------------------------------------- checkpoint interchange
    def optimizer_state_dict(self, epoch: int = 0) -> dict:
        """torch.optim.Adam-format state (param order: decoder.parameters(), then lat_vecs), what the reference's
        save_optimizer writes (train_deep_sdf.py:42-49); loads into a torch.optim.Adam built the reference's way
        (:454-466) and back through load_optimizer_state_dict."""
        c = self.cfg
        state = {}
        off = 0
        plist = list(self.decoder.parameters())
        for i, p in enumerate(plist):
            n = p.numel()
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self.exp_avg[off:off + n].view(p.shape).clone(),
                        "exp_avg_sq": self.exp_avg_sq[off:off + n].view(p.shape).clone()}
            off += n
        state[len(plist)] = {"step": torch.tensor(float(self.step_count)),
                             "exp_avg": self.lat_exp_avg.clone(),
                             "exp_avg_sq": self.lat_exp_avg_sq.clone()}
        common = {"betas": tuple(c.betas), "eps": c.eps, "weight_decay": 0, "amsgrad": False, "maximize": False,
                  "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                  "decoupled_weight_decay": False}
        groups = [dict(common, lr=self.lr(0, epoch), params=list(range(len(plist)))),
                  dict(common, lr=self.lr(1, epoch), params=[len(plist)])]
        return {"state": state, "param_groups": groups}

    def load_optimizer_state_dict(self, sd: dict) -> None:
        """Resume (`--continue`, train_deep_sdf.py:469-493 -> load_optimizer :52-61): take the Adam moments and
        step count from a torch.optim.Adam state_dict (ours or one the reference saved)."""
        plist = list(self.decoder.parameters())
        st = sd["state"]
        if len(st) == 0:
            self.exp_avg.zero_(); self.exp_avg_sq.zero_()
            self.lat_exp_avg.zero_(); self.lat_exp_avg_sq.zero_()
            self.step_count = 0
            return
        if len(st) != len(plist) + 1:
            raise ValueError(f"optimizer state has {len(st)} entries, this decoder + lat_vecs need {len(plist) + 1}")
        off = 0
        for i, p in enumerate(plist):
            n = p.numel()
            if tuple(st[i]["exp_avg"].shape) != tuple(p.shape):
                raise ValueError(f"optimizer state entry {i} has shape {tuple(st[i]['exp_avg'].shape)}, "
                                 f"parameter has {tuple(p.shape)}")
            self.exp_avg[off:off + n].copy_(st[i]["exp_avg"].reshape(-1))
            self.exp_avg_sq[off:off + n].copy_(st[i]["exp_avg_sq"].reshape(-1))
            off += n
        lat = st[len(plist)]
        if tuple(lat["exp_avg"].shape) != tuple(self.lat_exp_avg.shape):
            raise ValueError("latent-code optimizer state does not match lat_vecs")
        self.lat_exp_avg.copy_(lat["exp_avg"]); self.lat_exp_avg_sq.copy_(lat["exp_avg_sq"])
        steps = {int(float(v["step"])) for v in st.values()}
        if len(steps) != 1:
            raise ValueError(f"per-parameter step counts differ ({sorted(steps)}): one shared count is supported")
        self.step_count = steps.pop()

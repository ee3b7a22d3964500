"""`Decoder`: drop-in for the reference's networks/deep_sdf_decoder.Decoder.

Same constructor keywords (reference networks/deep_sdf_decoder.py:10-22), same parameter
names / shapes / registration order (`linK.weight`, `linK.bias` or `linK.bias`,
`linK.weight_g`, `linK.weight_v`), same RNG consumption at construction, so reference
checkpoints load unchanged and `torch.manual_seed(s); Decoder(...)` gives the reference's
initial weights.  `forward(input[N, L+3]) -> [N, 1]` (deep_sdf_decoder.py:74-109) runs on the
hand-written sm_100a kernels through the C ABI (include/dsdf.h).  Unsupported NetworkSpecs
raise at construction; there is no PyTorch/CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import warnings
from typing import List, Optional

import torch
import torch.nn as nn

from . import _lib
from ._lib import DsdfConfig, DsdfError, DsdfParamLayout

LATENT = 256
HIDDEN = 512
NUM_LIN = 9


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class _FusedDecoderBase(nn.Module):
    """Runtime machinery shared by the decoder families (deep_sdf_decoder, siren_decoder): flat fp32
    parameter buffer, C handle, weight folding, forward / decode entry points.  A subclass builds the
    nn.Linear layers with the reference's names in its constructor, sets `_wn_mask`, `dropout`,
    `dropout_prob`, calls `_init_runtime()`, and names its library with `_arch`."""
    _arch = "deep_sdf_decoder"

    @property
    def L(self):
        return _lib.lib(self._arch)

    def _init_runtime(self):
        # runtime state (not part of state_dict)
        self._handle = None
        self._handle_device = None
        self._flat: Optional[torch.Tensor] = None
        self._layout: Optional[DsdfParamLayout] = None
        self._folded_key = None

    def _lin(self, l: int):
        return getattr(self, "lin" + str(l))

    # ------------------------------------------------------------------ parameter plumbing
    def _lin_params(self, l: int):
        lin = self._lin(l)
        if (self._wn_mask >> l) & 1:
            return lin.weight_v, lin.weight_g, lin.bias
        return lin.weight, None, lin.bias

    def flat_params(self) -> torch.Tensor:
        """All parameters as views of one flat fp32 buffer (parameters() order)."""
        plist = list(self.parameters())
        dev = plist[0].device
        ok = self._flat is not None and self._flat.device == dev
        if ok:
            off = 0
            base = self._flat.data_ptr()
            for p in plist:
                if p.data_ptr() != base + 4 * off or not p.is_contiguous():
                    ok = False
                    break
                off += p.numel()
        if not ok:
            total = sum(p.numel() for p in plist)
            flat = torch.empty(total, dtype=torch.float32, device=dev)
            off = 0
            for p in plist:
                n = p.numel()
                flat[off:off + n].copy_(p.data.reshape(-1).float())
                p.data = flat[off:off + n].view(p.shape)
                off += n
            self._flat = flat
            self._layout = None
            self._folded_key = None
        if self._layout is None:
            offs = {}
            off = 0
            for p in plist:
                offs[id(p)] = off
                off += p.numel()
            lay = DsdfParamLayout()
            for l in range(NUM_LIN):
                w, g, b = self._lin_params(l)
                lay.w_off[l] = offs[id(w)]
                lay.g_off[l] = offs[id(g)] if g is not None else -1
                lay.b_off[l] = offs[id(b)]
            lay.total = off
            self._layout = lay
        return self._flat

    def handle(self):
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise DsdfError("dsdf_b200.Decoder computes on CUDA (sm_100a) only; "
                            "there is no CPU fallback -- move the module with .cuda()")
        if self._handle is not None and self._handle_device == dev:
            return self._handle
        self._release()
        cfg = DsdfConfig(LATENT, HIDDEN, 8, 4, self._wn_mask, dev.index or 0)
        h = C.c_void_p()
        _lib.check(self.L.dsdf_create(C.byref(cfg), C.byref(h)))
        self._handle, self._handle_device = h, dev
        self._folded_key = None
        return h

    def _release(self):
        h = self.__dict__.get("_handle", None)
        if h is not None:
            self.__dict__["_handle"] = None      # plain dict write: safe during interpreter teardown
            try:
                self.L.dsdf_destroy(h)
            except Exception:
                pass

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    def sync_weights(self, force: bool = False):
        """Fold fp32 masters -> bf16 tensor-core images if any parameter changed.

        "Changed" is detected through the parameters' autograd version counters (optimiser steps, `p.copy_()`,
        `load_state_dict` all bump them).  Writes that go through `p.data` (`p.data.copy_`, `p.data.mul_` -- common in
        hand-written init / EMA code) do NOT bump that counter: after such a write call `sync_weights(force=True)`
        (`FusedTrainer.step` always re-folds after its own Adam update)."""
        h = self.handle()
        flat = self.flat_params()
        key = (flat.data_ptr(), tuple(p._version for p in self.parameters()))
        if force or key != self._folded_key:
            dev = flat.device
            with torch.cuda.device(dev):
                _lib.check(self.L.dsdf_set_weights(h, flat.data_ptr(), C.byref(self._layout),
                                                       _stream_ptr(dev)), h, self.L)
            self._folded_key = key
        return h

    # ------------------------------------------------------------------ forward
    def forward(self, input):
        if input.dim() != 2 or input.shape[1] != LATENT + 3:
            raise ValueError(f"expected input [N, {LATENT + 3}], got {tuple(input.shape)}")
        if self.training and self.dropout and self.dropout_prob > 0:
            raise NotImplementedError(
                "training-mode dropout runs inside the fused step (dsdf_b200.FusedTrainer, Philox stream "
                "fused into the forward epilogues); Decoder.forward serves eval mode -- call .eval() first")
        if not input.is_cuda:
            raise DsdfError("dsdf_b200.Decoder has no CPU path; pass CUDA tensors")
        need_grad = torch.is_grad_enabled() and (
            input.requires_grad or any(p.requires_grad for p in self.parameters()))
        if need_grad:
            from .autograd_fn import decoder_apply
            return decoder_apply(self, input)
        return self.forward_inference(input)

    @torch.no_grad()
    def forward_inference(self, input: torch.Tensor) -> torch.Tensor:
        h = self.sync_weights()
        x = input.detach().contiguous().float()
        n = x.shape[0]
        out = torch.empty(n, dtype=torch.float32, device=x.device)
        if n == 0:
            return out.view(0, 1)
        with torch.cuda.device(x.device):
            _lib.check(self.L.dsdf_forward_dense(h, x.data_ptr(), n, out.data_ptr(),
                                                     _stream_ptr(x.device)), h, self.L)
        return out.view(n, 1)

    @torch.no_grad()
    def decode_sdf(self, latent_vector: torch.Tensor, queries: torch.Tensor) -> torch.Tensor:
        """deep_sdf/utils.py:86-97 without materialising expand+cat: one latent, queries [n,3]."""
        h = self.sync_weights()
        lat = latent_vector.detach().reshape(-1).contiguous().float()
        if lat.numel() != LATENT:
            raise ValueError("latent_vector must have 256 elements")
        q = queries.detach().contiguous().float()
        n = q.shape[0]
        out = torch.empty(n, dtype=torch.float32, device=q.device)
        if n == 0:
            return out.view(0, 1)
        with torch.cuda.device(q.device):
            _lib.check(self.L.dsdf_forward_segments(h, lat.data_ptr(), None, 1, n, q.data_ptr(),
                                                        q.shape[1], n, out.data_ptr(),
                                                        _stream_ptr(q.device)), h, self.L)
        return out.view(n, 1)

    @torch.no_grad()
    def decode_grid(self, latent_vector: torch.Tensor, N: int, lo: int = 0,
                    hi: Optional[int] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """SDF on linear indices [lo,hi) of create_mesh's N^3 grid (deep_sdf/mesh.py:38-71)."""
        h = self.sync_weights()
        hi = N ** 3 if hi is None else hi
        lat = latent_vector.detach().reshape(-1).contiguous().float()
        dev = lat.device
        if out is None:
            out = torch.empty(hi - lo, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(self.L.dsdf_decode_grid(h, lat.data_ptr(), N, lo, hi, out.data_ptr(),
                                                   _stream_ptr(dev)), h, self.L)
        return out


class Decoder(_FusedDecoderBase):
    """networks/deep_sdf_decoder.Decoder (reference networks/deep_sdf_decoder.py:10-109)."""
    _arch = "deep_sdf_decoder"

    def __init__(self, latent_size, dims, dropout=None, dropout_prob=0.0, norm_layers=(),
                 latent_in=(), weight_norm=False, xyz_in_all=None, use_tanh=False,
                 latent_dropout=False):
        super().__init__()
        dims = list(dims)
        # ---- supported matrix; anything else raises, naming the key
        if latent_size != LATENT:
            raise NotImplementedError(f"CodeLength={latent_size}: fused path supports 256")
        if dims != [HIDDEN] * 8:
            raise NotImplementedError(f"NetworkSpecs.dims={dims}: fused path supports [512]*8")
        if list(latent_in) != [4]:
            raise NotImplementedError(f"NetworkSpecs.latent_in={list(latent_in)}: fused path supports [4]")
        if xyz_in_all:
            raise NotImplementedError("NetworkSpecs.xyz_in_all=true is not supported by the fused path")
        if use_tanh:
            raise NotImplementedError("NetworkSpecs.use_tanh=true (double tanh) is not supported")
        if latent_dropout:
            raise NotImplementedError("NetworkSpecs.latent_dropout=true is not supported")
        norm_layers = list(norm_layers or [])
        if norm_layers and not weight_norm:
            raise NotImplementedError("NetworkSpecs.norm_layers without weight_norm (LayerNorm) is not supported")
        if any(l not in range(8) for l in norm_layers):
            raise NotImplementedError(f"NetworkSpecs.norm_layers={norm_layers}: only hidden layers 0..7")
        if dropout is not None and any(int(l) not in range(8) for l in dropout):
            raise NotImplementedError(f"NetworkSpecs.dropout={list(dropout)}: only hidden layers 0..7")

        widths = [latent_size + 3] + dims + [1]
        self.num_layers = len(widths)
        self.norm_layers = norm_layers
        self.latent_in = list(latent_in)
        self.latent_dropout = latent_dropout
        self.xyz_in_all = xyz_in_all
        self.weight_norm = weight_norm
        self.use_tanh = use_tanh
        self.dropout_prob = dropout_prob
        self.dropout = list(dropout) if dropout is not None else None
        self._wn_mask = 0
        # identical construction order / RNG stream to the reference ctor (:39-55)
        for layer in range(self.num_layers - 1):
            out_dim = widths[layer + 1] - widths[0] if (layer + 1) in self.latent_in else widths[layer + 1]
            lin = nn.Linear(widths[layer], out_dim)
            if weight_norm and layer in norm_layers:
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    lin = nn.utils.weight_norm(lin)
                self._wn_mask |= 1 << layer
            setattr(self, "lin" + str(layer), lin)
        self.relu = nn.ReLU()
        self.th = nn.Tanh()
        self._init_runtime()

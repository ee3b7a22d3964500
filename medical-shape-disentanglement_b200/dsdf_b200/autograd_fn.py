"""torch.autograd bridge of `Decoder.forward` (reference networks/deep_sdf_decoder.py:74-109 when
called with grad enabled): forward = dsdf_dense_forward_train (the fused tcgen05 chain, keeping the
activation stash + ReLU masks in a workspace), backward = dsdf_dense_backward (dgrad chain, wgrad,
weight-norm backward, dInput).  Gradients flow to every decoder parameter and to `input`
(latent and xyz columns), which is what reconstruct.py:90-94,130 and Stage-2 callers rely on.

Python only owns the tensors; all arithmetic is in libdsdf_b200.so.
"""
from __future__ import annotations

import torch
from torch.autograd.function import once_differentiable

from . import _lib
from .decoder import _stream_ptr


class _DecoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dec, inp, *params):
        L = dec.L
        h = dec.sync_weights()
        x = inp.detach().contiguous().float()
        n = x.shape[0]
        dev = x.device
        out = torch.empty(n, dtype=torch.float32, device=dev)
        ws = torch.empty(L.dsdf_dense_workspace_bytes(h, max(n, 1)), dtype=torch.uint8, device=dev)
        if n > 0:
            with torch.cuda.device(dev):
                _lib.check(L.dsdf_dense_forward_train(h, x.data_ptr(), n, out.data_ptr(), ws.data_ptr(),
                                                      ws.numel(), _stream_ptr(dev)), h, L)
        ctx.dec, ctx.ws, ctx.x, ctx.out = dec, ws, x, out
        ctx.param_shapes = [p.shape for p in params]
        ctx.weights_key = dec._folded_key
        return out.view(n, 1)

    @staticmethod
    @once_differentiable   # opaque C kernels: double backward / create_graph=True must raise, not return zeros
    def backward(ctx, dout):
        dec, x, out = ctx.dec, ctx.x, ctx.out
        L = dec.L
        n = x.shape[0]
        dev = x.device
        if dec._folded_key != ctx.weights_key:
            raise RuntimeError("decoder parameters were modified between forward and backward")
        h = dec.handle()
        flat = dec.flat_params()
        grads = torch.zeros_like(flat)
        need_dx = ctx.needs_input_grad[1]
        dx = torch.empty(n, x.shape[1], dtype=torch.float32, device=dev) if need_dx else None
        if n > 0:
            d = dout.detach().reshape(-1).contiguous().float()
            with torch.cuda.device(dev):
                _lib.check(L.dsdf_dense_backward(h, flat.data_ptr(), x.data_ptr(), n, out.data_ptr(), d.data_ptr(),
                                                 ctx.ws.data_ptr(), ctx.ws.numel(), grads.data_ptr(),
                                                 dx.data_ptr() if need_dx else None, _stream_ptr(dev)), h, L)
        ctx.ws = None
        pg, off = [], 0
        for i, shp in enumerate(ctx.param_shapes):
            k = 1
            for s in shp:
                k *= s
            pg.append(grads[off:off + k].view(shp) if ctx.needs_input_grad[2 + i] else None)
            off += k
        return (None, dx) + tuple(pg)


def decoder_apply(dec, inp: torch.Tensor) -> torch.Tensor:
    return _DecoderFn.apply(dec, inp, *list(dec.parameters()))


class _SegmentsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dec, latents, xyz, *params):
        L = dec.L
        h = dec.sync_weights()
        flat = dec.flat_params()
        lat = latents.detach().contiguous().float()
        q = xyz.detach().contiguous().float()
        n_seg, p = q.shape[0], q.shape[1]
        dev = q.device
        out = torch.empty(n_seg * p, dtype=torch.float32, device=dev)
        ws = torch.empty(L.dsdf_train_workspace_bytes(h, n_seg * p, n_seg), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.check(L.dsdf_segments_forward_train(h, flat.data_ptr(), lat.data_ptr(), n_seg, p, q.data_ptr(),
                                                     q.shape[2], out.data_ptr(), ws.data_ptr(), ws.numel(),
                                                     _stream_ptr(dev)), h, L)
        ctx.dec, ctx.ws, ctx.lat, ctx.q, ctx.out = dec, ws, lat, q, out
        ctx.param_shapes = [t.shape for t in params]
        ctx.weights_key = dec._folded_key
        return out.view(n_seg * p, 1)

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        dec, lat, q, out = ctx.dec, ctx.lat, ctx.q, ctx.out
        L = dec.L
        if dec._folded_key != ctx.weights_key:
            raise RuntimeError("decoder parameters were modified between forward and backward")
        h = dec.handle()
        flat = dec.flat_params()
        dev = q.device
        grads = torch.zeros_like(flat)
        dlat = torch.empty_like(lat)
        d = dout.detach().reshape(-1).contiguous().float()
        with torch.cuda.device(dev):
            _lib.check(L.dsdf_segments_backward(h, flat.data_ptr(), lat.data_ptr(), q.shape[0], q.shape[1],
                                                q.data_ptr(), q.shape[2], out.data_ptr(), d.data_ptr(),
                                                ctx.ws.data_ptr(), ctx.ws.numel(), grads.data_ptr(), dlat.data_ptr(),
                                                _stream_ptr(dev)), h, L)
        ctx.ws = None
        pg, off = [], 0
        for i, shp in enumerate(ctx.param_shapes):
            k = 1
            for s in shp:
                k *= s
            pg.append(grads[off:off + k].view(shp) if ctx.needs_input_grad[3 + i] else None)
            off += k
        return (None, dlat if ctx.needs_input_grad[1] else None, None) + tuple(pg)


def decode_segments(dec, latents: torch.Tensor, xyz: torch.Tensor) -> torch.Tensor:
    """decoder(cat([latents[b].expand(P, -1), xyz[b]], 1)) for every shape b, as one differentiable call:
    latents [B, 256] (may require grad: e.g. a VAE's z_hat, train_MLP_VAE_deep_sdf.py:3097-3133),
    xyz [B, P, 3] (or [B, P, 4] sample rows; the 4th column is ignored), P % 128 == 0.  Returns [B*P, 1].
    Gradients reach `latents` (deterministic segmented reduction) and the decoder parameters; no gradient
    is produced for xyz.  Uses the fold-mode kernels of the fused step (no [B*P, 259] input is built)."""
    if xyz.dim() != 3 or xyz.shape[2] < 3 or xyz.shape[1] % 128 != 0:
        raise ValueError(f"xyz must be [B, P, >=3] with P % 128 == 0, got {tuple(xyz.shape)}")
    if latents.shape != (xyz.shape[0], 256):
        raise ValueError(f"latents must be [{xyz.shape[0]}, 256], got {tuple(latents.shape)}")
    if not xyz.is_cuda:
        raise _lib.DsdfError("decode_segments has no CPU path; pass CUDA tensors")
    if dec.training and dec.dropout and dec.dropout_prob > 0:
        raise NotImplementedError("training-mode dropout runs inside the fused step only; call decoder.eval()")
    return _SegmentsFn.apply(dec, latents, xyz, *list(dec.parameters()))

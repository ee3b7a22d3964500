"""`siren.Decoder`: drop-in for the reference's networks/siren_decoder.Decoder (SIREN variant of the
DeepSDF auto-decoder: sin(30 x) activations, separate latent / xyz in-streaming, no output tanh).

Same constructor keywords (reference networks/siren_decoder.py:35-44 + the SirenDecoder kwargs of
:100-113), same parameter names (`decoder.linK.weight`, `decoder.linK.bias`), shapes, registration
order and RNG consumption (nn.Linear default init, then sine_init on every layer, then
first_layer_sine_init on lin0, :181-186), so reference checkpoints load unchanged and a seeded
construction gives the reference's initial weights.  The arithmetic runs in libdsdf_b200_siren.so:
the same fused tcgen05 chain as the ReLU decoder built with -DDSDF_ARCH_SIREN (sine epilogues,
bf16 derivative stash for the backward chain, lin3 512->256, lin4 = [x3 | latent]).

Supported: dims=[512]*8, latent 256, latent_in=[4], xyz_in=() and xyz_in_all=False,
encoding_features=1, nonlinearity="sine", use_tanh=False, no dropout / norm layers / latent dropout.
Anything else raises at construction naming the key (no fallback).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .decoder import HIDDEN, LATENT, _FusedDecoderBase


class _SirenCore(nn.Module):
    """Holds lin0..lin8 under the attribute name `decoder`, as the reference's SirenDecoder does."""


class Decoder(_FusedDecoderBase):
    _arch = "siren_decoder"

    def __init__(self, latent_size, dims, encoding_features=1, encoding_sigma=0, xyz_in=(), xyz_in_all=False,
                 dropout=None, dropout_prob=0.0, norm_layers=(), latent_in=(), weight_norm=False,
                 latent_dropout=False, nonlinearity="relu", use_tanh=False):
        super().__init__()
        dims = list(dims)
        if latent_size != LATENT:
            raise NotImplementedError(f"CodeLength={latent_size}: fused path supports 256")
        if dims != [HIDDEN] * 8:
            raise NotImplementedError(f"NetworkSpecs.dims={dims}: fused path supports [512]*8")
        # the reference appends 0 to the caller's list (siren_decoder.py:129); accept both spellings
        if sorted(set(latent_in) - {0}) != [4]:
            raise NotImplementedError(f"NetworkSpecs.latent_in={list(latent_in)}: fused path supports [4]")
        if encoding_features != 1:
            raise NotImplementedError("NetworkSpecs.encoding_features > 1 (Fourier features) is not supported")
        if xyz_in_all or sorted(set(xyz_in) - {0}):
            raise NotImplementedError("NetworkSpecs.xyz_in / xyz_in_all in-streaming is not supported")
        if nonlinearity != "sine":
            raise NotImplementedError(f"NetworkSpecs.nonlinearity={nonlinearity!r}: the SIREN build supports 'sine' "
                                      "(use NetworkArch deep_sdf_decoder for ReLU)")
        if use_tanh:
            raise NotImplementedError("NetworkSpecs.use_tanh=true is not supported")
        if latent_dropout:
            raise NotImplementedError("NetworkSpecs.latent_dropout=true is not supported")
        if norm_layers:
            raise NotImplementedError("NetworkSpecs.norm_layers (weight norm / BatchNorm1d) is not supported for SIREN")
        if dropout and dropout_prob > 0:
            raise NotImplementedError("NetworkSpecs.dropout is not supported for SIREN")
        self.encoding_features = encoding_features
        self.encoding_sigma = encoding_sigma
        self.encoding = None
        self.dropout = None
        self.dropout_prob = 0.0
        self._wn_mask = 0
        core = _SirenCore()
        # fc widths of siren_decoder.py:131-136: the skip layer's INPUT carries the latent, so lin3 has
        # dims[3] - latent_size outputs
        ins = [latent_size + 3, 512, 512, 512, 512, 512, 512, 512, 512]
        outs = [512, 512, 512, 512 - latent_size, 512, 512, 512, 512, 1]
        for i in range(9):
            lin = nn.Linear(ins[i], outs[i])
            with torch.no_grad():                                  # sine_init (:8-13)
                b = math.sqrt(6 / ins[i]) / 30
                lin.weight.uniform_(-b, b)
            setattr(core, "lin" + str(i), lin)
        with torch.no_grad():                                      # first_layer_sine_init (:16-21)
            core.lin0.weight.uniform_(-1 / ins[0], 1 / ins[0])
        self.decoder = core
        self._init_runtime()

    def _lin(self, l: int):
        return getattr(self.decoder, "lin" + str(l))

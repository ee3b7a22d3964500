"""Oracle restatement (TEST INFRASTRUCTURE ONLY) of the reference's SIREN decoder for the supported
configuration (8x512, latent 256, latent_in=[4], no xyz in-streaming, nonlinearity="sine",
use_tanh=False): networks/siren_decoder.py:8-21 (inits), :83-93 (Decoder.forward), :98-237
(SirenDecoder) and networks/modules.py:4-14 (Sine).  Plain torch fp32; pinned bit-for-bit against the
live reference in tests/test_oracle_cpu.py and through tests/golden/decoder_siren.npz."""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict

import torch
import torch.nn.functional as F

LATENT = 256
INS = [LATENT + 3, 512, 512, 512, 512, 512, 512, 512, 512]
OUTS = [512, 512, 512, 512 - LATENT, 512, 512, 512, 512, 1]      # siren_decoder.py:131-136


def network_specs() -> dict:
    """The NetworkSpecs block that selects this configuration in the reference."""
    return dict(dims=[512] * 8, latent_in=[4], nonlinearity="sine", encoding_features=1, xyz_in=[],
                xyz_in_all=False, dropout=None, dropout_prob=0.0, norm_layers=[], weight_norm=False,
                latent_dropout=False, use_tanh=False)


def init_params(seed: int | None = None) -> "OrderedDict[str, torch.Tensor]":
    """Same RNG consumption as the reference constructor (:148-186): per layer nn.Linear's default init,
    then sine_init on the weight; after the loop first_layer_sine_init on lin0."""
    if seed is not None:
        torch.manual_seed(seed)
    p: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    lins = []
    for i in range(9):
        lin = torch.nn.Linear(INS[i], OUTS[i])
        with torch.no_grad():
            b = math.sqrt(6 / INS[i]) / 30
            lin.weight.uniform_(-b, b)
        lins.append(lin)
    with torch.no_grad():
        lins[0].weight.uniform_(-1 / INS[0], 1 / INS[0])
    for i, lin in enumerate(lins):
        p[f"decoder.lin{i}.weight"] = lin.weight.detach().clone()
        p[f"decoder.lin{i}.bias"] = lin.bias.detach().clone()
    return p


def forward(p: Dict[str, torch.Tensor], inp: torch.Tensor) -> torch.Tensor:
    """input [N, L+3] -> [N, 1] (siren_decoder.py:83-93, 188-237)."""
    xyz = inp[:, -3:]
    lat = inp[:, :-3]
    x = torch.cat([lat, xyz], 1)
    for i in range(9):
        if i == 4:
            x = torch.cat([x, lat], 1)
        x = F.linear(x, p[f"decoder.lin{i}.weight"], p[f"decoder.lin{i}.bias"])
        if i < 8:
            x = torch.sin(30 * x)
    return x


def macs_per_row() -> int:
    return sum(a * b for a, b in zip(INS, OUTS))

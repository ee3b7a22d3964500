"""Oracle restatement of the reference DeepSDF decoder (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/networks/deep_sdf_decoder.py:
  * layer widths               :25-46   (`dims = [L+3] + dims + [1]`, skip layer shrinks the
                                         previous layer's out by `dims[0]`, xyz_in_all by 3)
  * weight-norm / LayerNorm    :48-62   (weight_norm only `if weight_norm and layer in norm_layers`)
  * forward                    :74-109  (cat at latent_in, ReLU on hidden layers, optional tanh,
                                         and the unconditional final `self.th` tanh, :106-107)

Written functionally over an ordered dict of tensors that carries the reference's
state_dict key names (`linK.weight`/`linK.bias` or `linK.weight_g`/`linK.weight_v`/`linK.bias`,
`bnK.weight`/`bnK.bias`), so oracle parameters and reference checkpoints interchange.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F


@dataclass
class DecoderSpec:
    latent_size: int = 256
    dims: Sequence[int] = tuple([512] * 8)
    dropout: Sequence[int] = ()
    dropout_prob: float = 0.0
    norm_layers: Sequence[int] = ()
    latent_in: Sequence[int] = (4,)
    weight_norm: bool = False
    xyz_in_all: bool = False
    use_tanh: bool = False
    latent_dropout: bool = False

    @staticmethod
    def from_network_specs(latent_size: int, ns: dict) -> "DecoderSpec":
        return DecoderSpec(
            latent_size=latent_size,
            dims=tuple(ns["dims"]),
            dropout=tuple(ns.get("dropout") or ()),
            dropout_prob=float(ns.get("dropout_prob", 0.0)),
            norm_layers=tuple(ns.get("norm_layers") or ()),
            latent_in=tuple(ns.get("latent_in") or ()),
            weight_norm=bool(ns.get("weight_norm", False)),
            xyz_in_all=bool(ns.get("xyz_in_all") or False),
            use_tanh=bool(ns.get("use_tanh", False)),
            latent_dropout=bool(ns.get("latent_dropout", False)),
        )

    def network_specs(self) -> dict:
        return dict(
            dims=list(self.dims), dropout=list(self.dropout), dropout_prob=self.dropout_prob,
            norm_layers=list(self.norm_layers), latent_in=list(self.latent_in),
            weight_norm=self.weight_norm, xyz_in_all=self.xyz_in_all, use_tanh=self.use_tanh,
            latent_dropout=self.latent_dropout,
        )


def layer_shapes(spec: DecoderSpec) -> List[Tuple[int, int]]:
    """(in_features, out_features) per Linear, deep_sdf_decoder.py:27,39-46."""
    widths = [spec.latent_size + 3] + list(spec.dims) + [1]
    n_lin = len(widths) - 1
    shapes = []
    for l in range(n_lin):
        if (l + 1) in spec.latent_in:
            out = widths[l + 1] - widths[0]
        else:
            out = widths[l + 1]
            if spec.xyz_in_all and l != n_lin - 1:
                out -= 3
        shapes.append((widths[l], out))
    return shapes


def is_wn(spec: DecoderSpec, l: int) -> bool:
    return bool(spec.weight_norm) and l in spec.norm_layers


def is_ln(spec: DecoderSpec, l: int) -> bool:
    return (not spec.weight_norm) and spec.norm_layers is not None and l in spec.norm_layers


def init_params(spec: DecoderSpec, seed: int | None = None) -> "OrderedDict[str, torch.Tensor]":
    """Create parameters with the same RNG consumption as the reference constructor.

    The reference builds `nn.Linear(in, out)` for lin0..linN in order
    (deep_sdf_decoder.py:39-55); each draws kaiming-uniform weight then uniform bias.
    `weight_norm` consumes no RNG (g = ||v|| per output row).  Re-using nn.Linear for
    the draws keeps the streams identical, so `torch.manual_seed(s)` gives the same
    numbers as `Decoder(...)` in the reference.
    """
    if seed is not None:
        torch.manual_seed(seed)
    p: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for l, (fin, fout) in enumerate(layer_shapes(spec)):
        lin = torch.nn.Linear(fin, fout)
        w = lin.weight.detach().clone()
        b = lin.bias.detach().clone()
        if is_wn(spec, l):
            # torch parametrisation registers bias first, then weight_g, weight_v
            p[f"lin{l}.bias"] = b
            p[f"lin{l}.weight_g"] = w.norm(dim=1, keepdim=True)
            p[f"lin{l}.weight_v"] = w
        else:
            p[f"lin{l}.weight"] = w
            p[f"lin{l}.bias"] = b
        if is_ln(spec, l):
            p[f"bn{l}.weight"] = torch.ones(fout)
            p[f"bn{l}.bias"] = torch.zeros(fout)
    return p


def effective_weight(p: Dict[str, torch.Tensor], spec: DecoderSpec, l: int) -> torch.Tensor:
    """w = g * v / ||v||_row for weight-normalised layers (deep_sdf_decoder.py:49-53)."""
    if is_wn(spec, l):
        v = p[f"lin{l}.weight_v"]
        g = p[f"lin{l}.weight_g"]
        # same torch primitive the reference's WeightNorm pre-forward hook calls
        # (== v * (g / ||v||_row) up to 1 ulp); keeps the oracle bit-equal to the reference
        return torch._weight_norm(v, g, 0)
    return p[f"lin{l}.weight"]


def forward(p: Dict[str, torch.Tensor], spec: DecoderSpec, inp: torch.Tensor,
            training: bool = False, dropout_keep: Dict[int, torch.Tensor] | None = None) -> torch.Tensor:
    """input [N, L+3] -> [N, 1]; deep_sdf_decoder.py:74-109.

    `dropout_keep[l]` ([N, out] bool/0-1), when given, replaces the RNG draw of F.dropout for layer l
    by an explicit keep mask (y = x * keep / (1 - p), the same arithmetic F.dropout applies), so a
    counter-based dropout stream can be checked against this oracle element by element."""
    n_lin = len(spec.dims) + 1
    xyz = inp[:, -3:]
    if inp.shape[1] > 3 and spec.latent_dropout:
        lat = F.dropout(inp[:, :-3], p=0.2, training=training)
        x = torch.cat([lat, xyz], 1)
    else:
        x = inp
    for l in range(n_lin):
        if l in spec.latent_in:
            x = torch.cat([x, inp], 1)
        elif l != 0 and spec.xyz_in_all:
            x = torch.cat([x, xyz], 1)
        x = F.linear(x, effective_weight(p, spec, l), p[f"lin{l}.bias"])
        if l == n_lin - 1 and spec.use_tanh:
            x = torch.tanh(x)
        if l < n_lin - 1:
            if is_ln(spec, l):
                x = F.layer_norm(x, (x.shape[1],), p[f"bn{l}.weight"], p[f"bn{l}.bias"])
            x = torch.relu(x)
            if spec.dropout is not None and l in spec.dropout:
                if dropout_keep is not None and training:
                    x = x * dropout_keep[l].to(x.dtype) * (1.0 / (1.0 - spec.dropout_prob))
                else:
                    x = F.dropout(x, p=spec.dropout_prob, training=training)
    return torch.tanh(x)  # `self.th` is always present (:71,106-107)


def param_count(spec: DecoderSpec) -> int:
    return sum(int(t.numel()) for t in init_params(spec, seed=0).values())


# FLOP accounting used by bench.py / DESIGN.md (nominal, unpadded; 1 MAC = 2 FLOP)
def macs_per_row(spec: DecoderSpec) -> int:
    return sum(fin * fout for fin, fout in layer_shapes(spec))

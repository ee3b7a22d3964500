"""Generate tests/golden/decoder_siren.npz from the LIVE reference SIREN decoder (build container only):

    python tests/golden/make_golden_siren.py

networks.siren_decoder.Decoder from /root/reference, seed 0, the supported NetworkSpecs
(oracle/siren.py:network_specs).  The reference's only executable check for this class is the
unseeded __main__ smoke (siren_decoder.py:240-268), so this file is the pin: parameter checksums
(weights are regenerated from the seed at test time), an input, its output, dInput and gradient
slices under a fixed upstream gradient."""
import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
warnings.simplefilter("ignore")
from networks.siren_decoder import Decoder as RefSiren  # noqa: E402

from oracle import siren  # noqa: E402  (only for the NetworkSpecs dict)

torch.manual_seed(0)
dec = RefSiren(256, **siren.network_specs())
g = torch.Generator().manual_seed(21)
n = 192
x = torch.cat([torch.randn(n, 256, generator=g) * 0.06, torch.rand(n, 3, generator=g) * 2 - 1], 1)
x.requires_grad_(True)
y = dec(x)
w = torch.linspace(-1, 1, n)[:, None]
(y * w).sum().backward()
out = {"input": x.detach().numpy(), "output": y.detach().numpy(), "dinput": x.grad.numpy(), "w": w.numpy()}
for k, v in dec.named_parameters():
    out["gradnorm/" + k] = np.array(float(v.grad.norm()))
for k, v in dec.state_dict().items():
    out["cks/" + k] = np.array([float(v.double().sum()), float(v.double().abs().sum())])
out["grad/decoder.lin8.weight"] = dec.decoder.lin8.weight.grad.numpy()
out["grad/decoder.lin4.bias"] = dec.decoder.lin4.bias.grad.numpy()
out["grad/decoder.lin0.bias"] = dec.decoder.lin0.bias.grad.numpy()
np.savez_compressed(os.path.join(HERE, "decoder_siren.npz"), **out)
print("wrote decoder_siren.npz", {k: v.shape for k, v in out.items() if not k.startswith(("cks", "gradnorm"))})

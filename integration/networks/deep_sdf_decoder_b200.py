"""networks/deep_sdf_decoder_b200.py -- the file a maintainer adds to the reference's `networks/` package.

Select it with  "NetworkArch": "deep_sdf_decoder_b200"  in specs.json; nothing else changes:
`arch.Decoder(latent_size, **specs["NetworkSpecs"])` (train_deep_sdf.py:348, reconstruct.py:229,
deep_sdf/workspace.py:66, generate_training_meshes.py:29) builds the B200-native module with the reference's
constructor keywords, parameter names and checkpoint keys (networks/deep_sdf_decoder.py:10-71).
DSDF_B200_HOME points at this repository's package directory when it is not already importable.
"""
import os
import sys

_home = os.environ.get("DSDF_B200_HOME") or os.path.join(
    os.path.dirname(os.path.abspath(__file__)), "..", "..", "medical-shape-disentanglement_b200")
if os.path.isdir(_home) and _home not in sys.path:
    sys.path.insert(0, _home)

from dsdf_b200 import Decoder  # noqa: E402,F401  (same ctor kwargs / parameter names as networks.deep_sdf_decoder.Decoder)

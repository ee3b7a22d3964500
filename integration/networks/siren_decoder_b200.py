"""networks/siren_decoder_b200.py -- drop-in for networks/siren_decoder.Decoder (networks/siren_decoder.py:64-93),
selected with  "NetworkArch": "siren_decoder_b200".  Same constructor keywords and `decoder.linK.*` checkpoint keys;
runs on libdsdf_b200_siren.so (sine epilogues, no output tanh)."""
import os
import sys

_home = os.environ.get("DSDF_B200_HOME") or os.path.join(
    os.path.dirname(os.path.abspath(__file__)), "..", "..", "medical-shape-disentanglement_b200")
if os.path.isdir(_home) and _home not in sys.path:
    sys.path.insert(0, _home)

from dsdf_b200.siren import Decoder  # noqa: E402,F401

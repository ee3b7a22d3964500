"""Reference-side package stub.  In the reference tree the two `*_b200.py` files of this directory are dropped
into its own `networks/` package; here the directory stands in for that package so the selection line every
reference caller uses --  __import__("networks." + specs["NetworkArch"], fromlist=["Decoder"])
(train_deep_sdf.py:225, reconstruct.py:225, deep_sdf/workspace.py:56) -- can be exercised by the tests.  With the
reference checkout on sys.path as well, both packages merge (extend_path) and the reference's own architectures
stay importable next to the drop-ins."""
from pkgutil import extend_path

__path__ = extend_path(__path__, __name__)

"""tcgen05.mma issue-rate micro-benchmark (diagnostics)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200"))
import ctypes as C
import numpy as np
import torch
from dsdf_b200 import _lib
torch.cuda.init(); torch.zeros(1, device="cuda")
L = _lib.lib()
for n_cta in (148,):
    for kind in (0, 16, 64, 65, 64 + 8, 65 + 8, 64 + 16, 65 + 16):
        buf = (C.c_uint64 * n_cta)()
        for rep in range(2):
            rc = L.dsdf_selftest_mma_rate(kind, 512, n_cta, buf)
        a = np.frombuffer(buf, dtype=np.uint64).astype(np.float64)
        print(("PAIR " if kind & 64 else "") + f"n_cta {n_cta:4d} kind {kind} (layout={kind&7} commit_each={(kind>>3)&1} trywait={(kind>>4)&1} commit2={(kind>>5)&1}): rc {rc} {a.mean()/4096:8.1f} cyc/MMA (max {a.max()/4096:8.1f})")

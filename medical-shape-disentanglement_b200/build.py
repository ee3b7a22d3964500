"""Build libdsdf_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python medical-shape-disentanglement_b200/build.py [--force]

nvcc cross-compiles without a GPU; the .so lands next to the Python package
(dsdf_b200/libdsdf_b200.so) so it travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "dsdf_b200", "libdsdf_b200.so")
# diagnostics variants: DSDF_BUILD_VARIANT=name DSDF_BUILD_DEFS="-DDSDF_DBG_BUILD -DDSDF_ABLATE=1"
# builds dsdf_b200/libdsdf_b200.<name>.so (loaded with DSDF_LIB=<path>); never the product library
VARIANT = os.environ.get("DSDF_BUILD_VARIANT", "")
EXTRA = os.environ.get("DSDF_BUILD_DEFS", "").split()
if VARIANT:
    OUT = os.path.join(HERE, "dsdf_b200", f"libdsdf_b200.{VARIANT}.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC))


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = sources() + [os.path.join(HERE, "..", "include", "dsdf.h")]
    return any(os.path.getmtime(s) > t for s in deps)


def build_lib(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    cmd = [NVCC] + FLAGS + EXTRA + ["-o", OUT, os.path.join(CSRC, "dsdf_api.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libdsdf_b200.so")
    if not VARIANT:
        with open(os.path.join(HERE, "dsdf_b200", "ptxas_report.txt"), "w") as f:
            f.write(r.stdout + r.stderr)
    return OUT


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose=True))

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


# the same source tree builds one library per network family (layout.cuh: geometry + activation are
# compile-time constants): libdsdf_b200.so = networks/deep_sdf_decoder, libdsdf_b200_siren.so =
# networks/siren_decoder (-DDSDF_ARCH_SIREN).  Same C ABI (include/dsdf.h); dsdf_arch() tells which.
OUT_SIREN = OUT[:-3] + "_siren.so"


def _compile(out: str, defs, report: str | None, verbose: bool) -> None:
    cmd = [NVCC] + FLAGS + EXTRA + list(defs) + ["-o", out, os.path.join(CSRC, "dsdf_api.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building " + os.path.basename(out))
    if report and not VARIANT:
        # ptxas' own compile-time lines differ from run to run: leave them out so that a rebuild of unchanged
        # sources leaves the committed report untouched
        text = "".join(l for l in (r.stdout + r.stderr).splitlines(True) if "Compile time" not in l)
        with open(os.path.join(HERE, "dsdf_b200", report), "w") as f:
            f.write(text)


def _stale(out: str) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    deps = sources() + [os.path.join(HERE, "..", "include", "dsdf.h")]
    return any(os.path.getmtime(s) > t for s in deps)


def build_lib(force: bool = False, verbose: bool = False) -> str:
    if force or _stale(OUT):
        _compile(OUT, [], "ptxas_report.txt", verbose)
    if force or _stale(OUT_SIREN):
        _compile(OUT_SIREN, ["-DDSDF_ARCH_SIREN"], "ptxas_report_siren.txt", verbose)
    return OUT


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose=True))

"""Recipe for oracle/_ref: the reference's OWN decoder modules, taken unmodified from where they lie under
/root/reference (TEST / BASELINE INFRASTRUCTURE ONLY; run by __graft_entry__.build() in the build container).

    python oracle/build_ref.py

The reference is a Python program, not an installable package (no setup.py / pyproject), and its entry scripts do
not import here (trimesh / skimage / pytorch3d are absent, train_deep_sdf.py hard-codes .cuda()); the modules that
define the arithmetic of this path import with torch alone:
    networks/__init__.py, networks/deep_sdf_decoder.py, networks/siren_decoder.py, networks/modules.py,
    deep_sdf/lr_scheduling.py
They are copied byte for byte into oracle/_ref/ (git-ignored, so no reference source enters the history; not
gpurun-ignored, so it travels to the GPU box like a built .so).  bench.py's CPU legs (`cpu_baseline`,
`--impl reference`) then time the reference's own Decoder class inside the step statements of
train_deep_sdf.py:555-616,731,780-784 (oracle/ref_step.py) -- kind "reference"; without oracle/_ref they fall back
to the oracle port and say kind "port".  The product never imports anything from here.
"""
import os
import shutil
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
FILES = ["networks/__init__.py", "networks/deep_sdf_decoder.py", "networks/siren_decoder.py", "networks/modules.py",
         "deep_sdf/lr_scheduling.py"]


def build_ref(verbose: bool = False) -> bool:
    if not os.path.isdir(REF):
        return os.path.isdir(OUT)          # GPU box: use what travelled with the snapshot
    for rel in FILES:
        dst = os.path.join(OUT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF, rel), dst)
        if verbose:
            print("oracle/_ref/" + rel)
    return True


if __name__ == "__main__":
    sys.exit(0 if build_ref(verbose=True) else 1)

"""Commit a copy of the four shipped Stage-1 experiment specifications (build container only):

    python tests/golden/make_golden_specs.py

Reads /root/reference/examples/{ADNI,OAI-ZIB}/minimal_eikonal{,_gmm}/specs.json and stores every key the Stage-1
trainer reads (train_deep_sdf.py:219-434) with its value -- free-text and path-valued keys are dropped -- as
tests/golden/stage1_specs.json, so the spec-reader tests run where /root/reference does not exist.
"""
import json
import os

REF = "/root/reference/examples"
HERE = os.path.dirname(os.path.abspath(__file__))
DROP = {"Description", "DataSource", "DataSourceMesh", "TrainSplit", "TestSplit", "TorusPath", "PretrainedSDFDecoderDir"}

out = {}
for ds in ("ADNI", "OAI-ZIB"):
    for exp in ("minimal_eikonal", "minimal_eikonal_gmm"):
        with open(os.path.join(REF, ds, exp, "specs.json")) as f:
            d = json.load(f)
        out[f"{ds}/{exp}"] = {k: v for k, v in d.items() if k not in DROP}
with open(os.path.join(HERE, "stage1_specs.json"), "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
print({k: len(v) for k, v in out.items()})

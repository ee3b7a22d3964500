"""Generate tests/golden/trainloop_ref.npz by EXECUTING the reference's own training script (build container only):

    python tests/golden/make_golden_trainloop.py

Runs train_deep_sdf.main_function (train_deep_sdf.py:211-1123) unmodified from /root/reference for a few epochs on a tiny
synthetic experiment directory, on CPU, with the Stage-1 objective the shipped specs use (Step schedules, CodeBound,
GradientClipNorm, code regularisation) both without and with UseEikonal.  Shims, none of which touches the arithmetic:
  * stub modules for imports this container lacks and the executed lines never call (trimesh, skimage, plyfile,
    pytorch3d, matplotlib, pyrender, ... -- whatever `import train_deep_sdf` trips over);
  * `.cuda()` on tensors / modules -> identity, torch.cuda.device_count() -> 1 (the script hard-codes .cuda());
  * torch.utils.tensorboard.SummaryWriter -> a no-op recorder; evaluation / plotting frequencies set beyond the run;
  * torch.utils.data.DataLoader -> a subclass that records every batch it yields (sdf_data, indices), so the oracle can
    be driven with exactly the batches the script saw (the loader's shuffle / window RNG is not part of the path).
Recorded: the initial decoder state_dict and latent table (captured at the first batch), every batch, the script's own
per-batch loss log and its final decoder / latent checkpoints.
"""
import json
import os
import random
import shutil
import sys
import tempfile
import types
import warnings

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, "..", "..", "medical-shape-disentanglement_b200"))
warnings.simplefilter("ignore")


class _Anything:
    """absorbs module-level decoration the executed path never depends on (plt.rc(...), matplotlib.use(...), base classes)"""
    def __call__(self, *a, **k): return self
    def __getattr__(self, k):
        if k.startswith("__"):
            raise AttributeError(k)
        return self
    def __mro_entries__(self, bases): return (object,)
    def __iter__(self): return iter(())
    def __getitem__(self, k): return self
    def __setitem__(self, k, v): pass


def _stub(name):
    m = types.ModuleType(name)
    m.__path__ = []

    def _attr(k):
        if k.startswith("__"):
            raise AttributeError(k)
        return _Anything()
    m.__getattr__ = _attr
    sys.modules[name] = m


def import_reference_script():
    for _ in range(80):
        try:
            import train_deep_sdf
            return train_deep_sdf
        except ModuleNotFoundError as e:
            _stub(e.name)
            # a failed package import leaves its already-imported submodules in sys.modules; the retry would then not
            # re-bind them as attributes of the fresh package object (main_function uses `deep_sdf.data.SDFSamples`)
            for k in [k for k in sys.modules if k == "deep_sdf" or k.startswith("deep_sdf.") or k in ("reconstruct", "train_deep_sdf")]:
                del sys.modules[k]
    raise RuntimeError("could not import the reference's train_deep_sdf.py")


_STATE = {}


def _install():
    """one-time shims (idempotent): .cuda() -> identity, recording SummaryWriter / DataLoader / Adam"""
    if _STATE:
        return _STATE
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    torch.cuda.device_count = lambda: 1
    scalars, batches, adam_init = [], [], []

    class _Writer:
        def __init__(self, *a, **k): pass
        def add_scalar(self, tag, value, global_step=None, *a, **k):
            scalars.append((tag, float(value), global_step))
        def __getattr__(self, k): return lambda *a, **kw: None
    import torch.utils.tensorboard as tb
    tb.SummaryWriter = _Writer
    ref = import_reference_script()
    ref.SummaryWriter = _Writer
    import deep_sdf.data  # noqa: F401  (main_function reaches it as an attribute of the package)
    import torch.utils.data as data_utils
    BaseLoader = data_utils.DataLoader

    class RecordingLoader(BaseLoader):
        def __iter__(self):
            for batch in super().__iter__():
                batches.append((batch[0].clone(), batch[1].clone()))
                yield batch
    ref.data_utils.DataLoader = RecordingLoader
    BaseAdam = torch.optim.Adam

    class RecordingAdam(BaseAdam):
        def __init__(self, params, *a, **k):
            params = [dict(g, params=list(g["params"])) for g in params]     # decoder.parameters() is a generator: keep it intact
            adam_init.append([[p.detach().clone() for p in g["params"]] for g in params])
            super().__init__(params, *a, **k)
    ref.torch.optim.Adam = RecordingAdam
    _STATE.update(ref=ref, scalars=scalars, batches=batches, adam_init=adam_init, BaseAdam=BaseAdam)
    return _STATE


def run_reference(eik: bool, S=6, P=1024, B=2, EPOCHS=3):
    """Execute train_deep_sdf.main_function on a temporary experiment; returns what the script did."""
    st = _install()
    ref = st["ref"]
    from dsdf_b200 import synth, data as ddata
    for k in ("scalars", "batches", "adam_init"):
        st[k].clear()
    root = tempfile.mkdtemp(prefix="dsdf_exp_")
    try:
        data_dir = os.path.join(root, "data")
        os.makedirs(data_dir)
        names = []
        for s in range(S):       # this fork's layout: <DataSource>/<instance>.npz, split = flat list of "<instance>.obj"
            d = synth.scene_samples(1, 8192, seed=10 + s)[0]
            ddata.write_sdf_samples(os.path.join(data_dir, f"s{s}.npz"), d[d[:, 3] > 0], d[d[:, 3] <= 0])
            names.append(f"s{s}.obj")
        for f in ("train.json", "test.json"):
            json.dump(names, open(os.path.join(root, f), "w"))
        specs = {
            "Description": "trainloop golden", "DataSource": data_dir, "TrainSplit": os.path.join(root, "train.json"),
            "TestSplit": os.path.join(root, "test.json"), "NetworkArch": "deep_sdf_decoder",
            "NetworkSpecs": {"dims": [512] * 8, "dropout": [], "dropout_prob": 0.2, "norm_layers": list(range(8)),
                             "latent_in": [4], "xyz_in_all": False, "use_tanh": False, "latent_dropout": False,
                             "weight_norm": True},
            "CodeLength": 256, "NumEpochs": EPOCHS, "SnapshotFrequency": EPOCHS, "AdditionalSnapshots": [],
            "LearningRateSchedule": [{"Type": "Step", "Initial": 0.0005, "Interval": 2, "Factor": 0.5},
                                     {"Type": "Step", "Initial": 0.001, "Interval": 2, "Factor": 0.5}],
            "SamplesPerScene": P, "ScenesPerBatch": B, "DataLoaderThreads": 0, "ClampingDistance": 0.1,
            "CodeRegularization": True, "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0, "GradientClipNorm": 1.0,
            "UseEikonal": eik, "LoadDatasetIntoRAM": True, "LogFrequency": 10 ** 6, "EvalTrainFrequency": 10 ** 6,
            "EvalTestFrequency": 10 ** 6, "EvalTrainSceneNumber": 0, "EvalTestSceneNumber": 0,
        }
        json.dump(specs, open(os.path.join(root, "specs.json"), "w"))
        torch.manual_seed(0)
        random.seed(0)
        np.random.seed(0)
        ref.main_function(root, None, 1)
        model = torch.load(os.path.join(root, "ModelParameters", f"{EPOCHS}.pth"), weights_only=False)
        lat = torch.load(os.path.join(root, "LatentCodes", f"{EPOCHS}.pth"), weights_only=False)
        return {"specs": specs, "batches": list(st["batches"]), "scalars": list(st["scalars"]),
                "init_decoder": st["adam_init"][0][0], "init_latents": st["adam_init"][0][1][0],
                "final_model": model["model_state_dict"], "final_latents": lat["latent_codes"]["weight"],
                "batches_per_epoch": S // B, "epochs": EPOCHS}
    finally:
        shutil.rmtree(root, ignore_errors=True)


def main():
    for eik in (False, True):
        r = run_reference(eik)
        per_batch = [v for t, v, _ in r["scalars"] if t == "Loss/train_vanilla"]
        print("UseEikonal", eik, "batches", len(r["batches"]), "per-batch loss", [round(v, 6) for v in per_batch])


if __name__ == "__main__":
    main()

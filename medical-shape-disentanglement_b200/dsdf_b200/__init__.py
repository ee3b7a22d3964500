"""dsdf_b200: B200-native (sm_100a) implementation of the DeepSDF auto-decoder hot path of
anonymous-submission01/medical-shape-disentanglement, behind the reference's `Decoder` API.

Python here is host-side plumbing only (device memory, streams, torch.distributed); the
arithmetic is in libdsdf_b200.so (hand-written tcgen05/TMEM/TMA-engine CUDA, csrc/), reached
through the C ABI declared in include/dsdf.h.
"""
from ._lib import DsdfError, LIB_PATH  # noqa: F401
from .decoder import Decoder  # noqa: F401
from .trainer import FusedTrainer, StepConfig  # noqa: F401
from .mesh import (convert_sdf_samples_to_ply, create_mesh, create_mesh_sdf, decode_grid_sharded,  # noqa: F401
                   marching_cubes, read_ply, write_ply)
from .reconstruct import reconstruct, reconstruct_batch, reconstruct_sharded, draw_windows, plan_samples  # noqa: F401
from . import data, siren  # noqa: F401,E402  (sample-file contract; SIREN decoder family)
from .autograd_fn import decode_segments  # noqa: F401,E402
from .preprocess import GpuSDFSamples, nearest_visible_sdf, preprocess_mesh, sample_sdf_near_surface, visible_surface_points  # noqa: F401,E402

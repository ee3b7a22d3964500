"""SDF sample-file contract of the reference (deep_sdf/data.py:76-136, 268-304; writer
src/PreprocessMesh.cpp:196-226): one `.npz` per shape with `pos` and `neg`, each [n, 4] fp32 rows
(x, y, z, sdf), sdf > 0 in `pos`.  Host-side I/O only (numpy + torch tensors); the per-iteration
window gather itself runs on the GPU (k_recon_tail / the fused step's A-tile load).

Same function names, arguments and RNG consumption as the reference module, minus its mesh helpers
(trimesh / marching-cubes utilities are outside the path).
"""
from __future__ import annotations

import random
from typing import List, Sequence

import numpy as np
import torch


def remove_nans(tensor: torch.Tensor) -> torch.Tensor:
    """data.py:76-78: drop rows whose sdf is NaN."""
    tensor_nan = torch.isnan(tensor[:, 3])
    return tensor[~tensor_nan, :]


def read_sdf_samples_into_ram(filename) -> List[torch.Tensor]:
    """data.py:83-88: [pos, neg] tensors of one shape."""
    npz = np.load(filename)
    return [torch.from_numpy(npz["pos"]), torch.from_numpy(npz["neg"])]


def unpack_sdf_samples(filename, subsample=None):
    """data.py:91-109: load + NaN filter + `subsample/2` random rows of each sign (torch.rand stream)."""
    npz = np.load(filename)
    if subsample is None:
        return npz
    pos_tensor = remove_nans(torch.from_numpy(npz["pos"]))
    neg_tensor = remove_nans(torch.from_numpy(npz["neg"]))
    half = int(subsample / 2)
    random_pos = (torch.rand(half) * pos_tensor.shape[0]).long()
    random_neg = (torch.rand(half) * neg_tensor.shape[0]).long()
    return torch.cat([torch.index_select(pos_tensor, 0, random_pos),
                      torch.index_select(neg_tensor, 0, random_neg)], 0)


def unpack_sdf_samples_from_ram(data: Sequence[torch.Tensor], subsample=None):
    """data.py:112-136: contiguous windows of `subsample/2` rows from pos and from neg, window starts from
    Python's `random.randint` (index_select fallback with torch.rand when neg is too short)."""
    if subsample is None:
        return data
    pos_tensor, neg_tensor = data[0], data[1]
    half = int(subsample / 2)
    pos_size, neg_size = pos_tensor.shape[0], neg_tensor.shape[0]
    pos_start_ind = random.randint(0, pos_size - half)
    sample_pos = pos_tensor[pos_start_ind:(pos_start_ind + half)]
    if neg_size <= half:
        random_neg = (torch.rand(half) * neg_tensor.shape[0]).long()
        sample_neg = torch.index_select(neg_tensor, 0, random_neg)
    else:
        neg_start_ind = random.randint(0, neg_size - half)
        sample_neg = neg_tensor[neg_start_ind:(neg_start_ind + half)]
    return torch.cat([sample_pos, sample_neg], 0)


def write_sdf_samples(filename, pos: torch.Tensor, neg: torch.Tensor) -> None:
    """The writer's layout (PreprocessMesh.cpp:196-226 via cnpy): `pos`, `neg` [n, 4] fp32."""
    np.savez(filename, pos=pos.detach().cpu().numpy().astype(np.float32),
             neg=neg.detach().cpu().numpy().astype(np.float32))


class SDFSamples(torch.utils.data.Dataset):
    """data.py:268-304 with load_ram=True semantics: item = (cat(half pos, half neg) [subsample, 4], idx).
    `npz_files` replaces the reference's split-file walk (get_instance_filenames), which is off the path."""

    def __init__(self, npz_files: Sequence[str], subsample: int, load_ram: bool = True):
        self.subsample = subsample
        self.npzfiles = list(npz_files)
        self.load_ram = load_ram
        self.loaded_data = []
        if load_ram:
            for f in self.npzfiles:
                pos, neg = read_sdf_samples_into_ram(f)
                pos, neg = remove_nans(pos), remove_nans(neg)
                # the reference shuffles each tensor once at load time (data.py:289-296)
                self.loaded_data.append([pos[torch.randperm(pos.shape[0])], neg[torch.randperm(neg.shape[0])]])

    def __len__(self):
        return len(self.npzfiles)

    def __getitem__(self, idx):
        if self.load_ram:
            return unpack_sdf_samples_from_ram(self.loaded_data[idx], self.subsample), idx
        return unpack_sdf_samples(self.npzfiles[idx], self.subsample), idx

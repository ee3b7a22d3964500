"""SDF-sample generation and a GPU-resident sample store.

`preprocess_mesh` restates the reference's `bin/PreprocessMesh` (src/PreprocessMesh.cpp:289-565) on the device through
the C ABI (dsdf_rasterize_views / dsdf_sample_points / dsdf_knn_sdf): centre the mesh on its bounding box, render it from
100 equidistant views with an orthographic camera into 400 x 400 position / normal images, keep the visible surface
points, draw the sample cloud (surface points perturbed with variance `variance` and `variance / 10`, plus uniform points
in the cube of side 2) and give every sample the distance to its nearest visible point, signed by a unanimous vote of the
11 nearest normals.  Returns the (pos, neg) tensors the reference writes into the .npz (writeSDFToNPZ :196-226).
The reference's OpenGL pipeline is unavailable on this box; nothing here needs a graphics API.

`GpuSDFSamples` keeps many scenes' samples resident in HBM and serves whole batches with the reference's sampling rule
(deep_sdf/data.py:112-136: one contiguous window of pos rows and one of neg rows per scene, starts from Python's
`random.randint`) in one gather kernel -- the step consumes ~1.9 GB/s of fresh rows per GPU, which a host DataLoader
cannot feed.
"""
from __future__ import annotations

import ctypes as C
import math
import random
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib

RAST_W = RAST_H = 400          # PreprocessMesh.cpp:423-424
NUM_VIEWS = 100                # :431
NUM_VOTES = 11                 # :537


def equidist_points_on_sphere(num_samples: int, radius: float) -> np.ndarray:
    """src/Utils.cpp:7-26, in fp32 like the C++ (golden-angle spiral)."""
    f = np.float32
    pts = np.zeros((num_samples, 3), dtype=np.float32)
    offset = f(2.0) / f(num_samples)
    increment = f(math.pi) * (f(3.0) - np.sqrt(f(5.0)))
    for i in range(num_samples):
        y = ((f(i) * offset) - f(1)) + (offset / f(2))
        r = np.sqrt(f(1) - np.power(y, f(2.0)))
        phi = (f(i) + f(1.0)) * increment
        pts[i] = f(radius) * np.array([np.cos(phi) * r, y, np.sin(phi) * r], dtype=np.float32)
    return pts


def look_at_views(eyes: np.ndarray) -> np.ndarray:
    """pangolin::ModelViewLookAt(eye, 0, AxisY) (right / up / back camera frame, as gluLookAt) for every eye:
    [n, 12] rows of [R | t] with camera = R p + t."""
    out = np.zeros((len(eyes), 12), dtype=np.float32)
    up = np.array([0, 1, 0], dtype=np.float32)
    for i, e in enumerate(eyes.astype(np.float32)):
        z = e / np.linalg.norm(e)
        x = np.cross(up, z)
        x = x / np.linalg.norm(x)
        y = np.cross(z, x)
        y = y / np.linalg.norm(y)
        R = np.stack([x, y, z]).astype(np.float32)
        t = -(R @ e)
        out[i] = np.concatenate([R, t[:, None]], 1).reshape(-1)
    return out


def bounding_cube_center(verts: np.ndarray, faces: np.ndarray) -> np.ndarray:
    """src/Utils.cpp:171-244 BoundingCubeNormalization as this fork ships it: subtract the bounding-box centre of the
    vertices that some face uses; NO rescaling (max_dist is returned as 1)."""
    used = np.unique(faces.reshape(-1))
    v = verts[used]
    return ((v.max(0) + v.min(0)) / np.float32(2.0)).astype(np.float32)


def triangle_area_cdf(verts: np.ndarray, faces: np.ndarray) -> np.ndarray:
    """src/Utils.cpp:70-91 TriangleArea per face (fp32, NaN -> 0) and the running sum of PreprocessMesh.cpp:46-65."""
    a, b, c = verts[faces[:, 0]], verts[faces[:, 1]], verts[faces[:, 2]]
    ab, ac = (b - a).astype(np.float32), (c - a).astype(np.float32)
    nab, nac = np.linalg.norm(ab, axis=1).astype(np.float32), np.linalg.norm(ac, axis=1).astype(np.float32)
    with np.errstate(invalid="ignore", divide="ignore"):
        cos = (np.einsum("ij,ij->i", ab, ac).astype(np.float32) / (nab * nac)).astype(np.float32)
        cos = np.where(cos < -1, np.float32(math.cos(math.pi * 359.0 / 360)), cos)
        cos = np.where(cos > 1, np.float32(math.cos(math.pi * 1.0 / 360)), cos).astype(np.float32)
        area = (np.float32(0.5) * nab * nac * np.sqrt(np.float32(1) - cos * cos)).astype(np.float32)
    area = np.where(np.isnan(area), np.float32(0), area).astype(np.float32)
    return np.cumsum(area, dtype=np.float32)           # sequential fp32 accumulation, like the C++ loop


def _st(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


@torch.no_grad()
def visible_surface_points(verts: torch.Tensor, faces: torch.Tensor, views: Optional[np.ndarray] = None,
                           return_images: bool = False):
    """Render the (centred) mesh from the 100 views and return (points [n,4], normals [n,4]) of the valid pixels in the
    reference's order (view by view, x-major inside a view); w = triangle id + 1 (PreprocessMesh.cpp:431-477)."""
    if not verts.is_cuda:
        raise _lib.DsdfError("visible_surface_points needs the mesh on a CUDA device (there is no CPU path)")
    L = _lib.lib()
    dev = verts.device
    v = verts.contiguous().float()
    f = faces.contiguous().to(torch.int32)
    if views is None:
        views = look_at_views(equidist_points_on_sphere(NUM_VIEWS, 1.0 * 1.1))
    vw = torch.from_numpy(np.ascontiguousarray(views, dtype=np.float32)).to(dev)
    nv = vw.shape[0]
    with torch.cuda.device(dev):
        zbuf = torch.empty(nv * RAST_W * RAST_H, dtype=torch.int64, device=dev)
        ov = torch.empty(nv, RAST_W, RAST_H, 4, dtype=torch.float32, device=dev)
        on = torch.empty_like(ov)
        ok = torch.empty(nv, RAST_W, RAST_H, dtype=torch.uint8, device=dev)
        _lib.check(L.dsdf_rasterize_views(v.data_ptr(), v.shape[0], f.data_ptr(), f.shape[0], vw.data_ptr(), nv,
                                          zbuf.data_ptr(), ov.data_ptr(), on.data_ptr(), ok.data_ptr(), _st(dev)), None, L)
    m = ok.view(-1).bool()
    pts, nrm = ov.view(-1, 4)[m], on.view(-1, 4)[m]
    if return_images:
        return pts, nrm, ov, on, ok
    return pts, nrm


@torch.no_grad()
def nearest_visible_sdf(xyz: torch.Tensor, pts: torch.Tensor, nrm: torch.Tensor, stdv: float, method: str = "grid",
                        return_neighbours: bool = False):
    """SampleSDFNearSurface :137-172 for every row of xyz [n,3]: the signed distance from the 11 nearest visible points
    (pts / nrm float4 rows), NaN where the sign vote is split.  method "grid" (uniform grid, dsdf_knn_sdf_grid) and
    "brute" (all pairs, dsdf_knn_sdf) give bit-identical results; the reference uses a nanoflann kd-tree."""
    L = _lib.lib()
    dev = xyz.device
    if dev.type != "cuda":
        raise _lib.DsdfError("nearest_visible_sdf needs CUDA tensors (there is no CPU path)")
    q = xyz.contiguous().float()
    p4, n4 = pts.contiguous().float(), nrm.contiguous().float()
    n = q.shape[0]
    with torch.cuda.device(dev):
        sdf = torch.empty(n, dtype=torch.float32, device=dev)
        nn = torch.empty(n, NUM_VOTES, dtype=torch.int32, device=dev) if return_neighbours else None
        nn_ptr = nn.data_ptr() if nn is not None else None
        if method == "brute":
            _lib.check(L.dsdf_knn_sdf(q.data_ptr(), n, p4.data_ptr(), n4.data_ptr(), p4.shape[0], float(stdv), sdf.data_ptr(),
                                      nn_ptr, _st(dev)), None, L)
        elif method == "grid":
            # one thread walks one query's cells: hand the kernel the queries in cell order (a 64^3 key over the
            # points' bounding box) so that the threads of a warp walk the same neighbourhood; every query's result is
            # independent of its position in the batch, so this changes nothing but the memory behaviour
            lo = p4[:, :3].amin(0)
            span = (p4[:, :3].amax(0) - lo).clamp_min(1e-12)
            cell = ((q - lo) / span * 64.0).clamp_(0.0, 63.0).long()
            perm = torch.argsort((cell[:, 2] * 64 + cell[:, 1]) * 64 + cell[:, 0])
            qs = q[perm].contiguous()
            sdf_s = torch.empty_like(sdf)
            nn_s = torch.empty_like(nn) if nn is not None else None
            ws = torch.empty(L.dsdf_knn_grid_workspace_bytes(p4.shape[0]), dtype=torch.uint8, device=dev)
            _lib.check(L.dsdf_knn_sdf_grid(qs.data_ptr(), n, p4.data_ptr(), n4.data_ptr(), p4.shape[0], float(stdv),
                                           ws.data_ptr(), ws.numel(), sdf_s.data_ptr(),
                                           nn_s.data_ptr() if nn_s is not None else None, _st(dev)), None, L)
            sdf[perm] = sdf_s
            if nn is not None:
                nn[perm] = nn_s
            torch.cuda.current_stream(dev).synchronize()      # the workspace must outlive the enqueued kernels
        else:
            raise ValueError(f"unknown method {method!r} (grid | brute)")
    return (sdf, nn) if return_neighbours else sdf


@torch.no_grad()
def sample_sdf_near_surface(verts: torch.Tensor, faces: torch.Tensor, pts: torch.Tensor, nrm: torch.Tensor,
                            num_sample: int = 500000, variance: float = 0.005, second_variance: Optional[float] = None,
                            seed: int = 0, near_ratio: Tuple[int, int] = (47, 50), return_all: bool = False,
                            knn: str = "grid"):
    """SampleFromSurface + SampleSDFNearSurface (PreprocessMesh.cpp:22-178, call site :516-537).  Returns (xyz [n,3],
    sdf [n]) of the samples whose 11-neighbour vote was unanimous, in generation order."""
    L = _lib.lib()
    dev = verts.device
    v = verts.contiguous().float()
    f = faces.contiguous().to(torch.int32)
    if second_variance is None:
        second_variance = variance / 10
    num_near = int(near_ratio[0] * num_sample / near_ratio[1])       # (int)(47 * num_sample / 50)
    n_surf = num_near // 2
    n_rand = num_sample - num_near
    cdf = torch.from_numpy(triangle_area_cdf(v.cpu().numpy(), f.cpu().numpy().astype(np.int64))).to(dev)
    n = 2 * n_surf + n_rand
    stdv = math.sqrt(variance)
    with torch.cuda.device(dev):
        xyz = torch.empty(n, 3, dtype=torch.float32, device=dev)
        _lib.check(L.dsdf_sample_points(v.data_ptr(), f.data_ptr(), cdf.data_ptr(), f.shape[0], n_surf, n_rand,
                                        float(np.float32(stdv)), float(np.float32(math.sqrt(second_variance))), 2.0,
                                        int(seed) & 0xFFFFFFFFFFFFFFFF, xyz.data_ptr(), None, _st(dev)), None, L)
    sdf = nearest_visible_sdf(xyz, pts, nrm, float(np.float32(stdv)), method=knn)
    if return_all:
        return xyz, sdf
    keep = ~torch.isnan(sdf)
    return xyz[keep], sdf[keep]


@torch.no_grad()
def preprocess_mesh(verts, faces, num_sample: int = 500000, variance: float = 0.005, seed: int = 0,
                    test_flag: bool = False, device="cuda") -> Tuple[torch.Tensor, torch.Tensor]:
    """`PreprocessMesh -m mesh -o out.npz [-t]` (PreprocessMesh.cpp:289-565): (pos [n,4], neg [m,4]) CUDA tensors of
    (x, y, z, sdf) rows, sdf > 0 in pos (writeSDFToNPZ :196-226).  verts [nv,3] / faces [nf,3] as arrays or tensors."""
    v = torch.as_tensor(np.asarray(verts, dtype=np.float32) if not torch.is_tensor(verts) else verts).float()
    f = torch.as_tensor(np.asarray(faces) if not torch.is_tensor(faces) else faces).to(torch.int64)
    near = (47, 50)
    second = variance / 10
    if test_flag:                                                     # :318-327
        variance = 0.05
        second = variance / 100
        near = (45, 50)
        num_sample = 250000
    centre = torch.from_numpy(bounding_cube_center(v.cpu().numpy(), f.cpu().numpy()))
    v = (v.cpu() - centre).to(device)
    f = f.to(device)
    pts, nrm = visible_surface_points(v, f)
    if pts.shape[0] < NUM_VOTES:
        raise ValueError("the mesh has fewer than 11 visible surface points")
    xyz, sdf = sample_sdf_near_surface(v, f, pts, nrm, num_sample, variance, second, seed, near)
    rows = torch.cat([xyz, sdf[:, None]], 1)
    return rows[sdf > 0], rows[~(sdf > 0)]


class GpuSDFSamples:
    """Scenes' (pos, neg) sample tensors resident on one GPU + the reference's per-item sampling rule for whole batches.

    `batch(scene_ids, subsample)` = torch.stack([unpack_sdf_samples_from_ram(data[i], subsample) for i in scene_ids])
    (deep_sdf/data.py:112-136, 298-304) with the same `random.randint` consumption, as ONE gather kernel; the scene
    tensors are shuffled once at load like the reference's loader does (data.py:289-296)."""

    def __init__(self, scenes: Sequence[Tuple[torch.Tensor, torch.Tensor]], device="cuda", shuffle: bool = True):
        self.device = torch.device(device)
        chunks, base = [], []
        off = 0
        self.sizes: List[Tuple[int, int]] = []
        for pos, neg in scenes:
            pos = pos[~torch.isnan(pos[:, 3])]                       # remove_nans (data.py:76-80)
            neg = neg[~torch.isnan(neg[:, 3])]
            if shuffle:
                pos = pos[torch.randperm(pos.shape[0])]
                neg = neg[torch.randperm(neg.shape[0])]
            base.append((off, off + pos.shape[0]))
            self.sizes.append((pos.shape[0], neg.shape[0]))
            chunks += [pos, neg]
            off += pos.shape[0] + neg.shape[0]
        self.samples = torch.cat(chunks, 0).float().contiguous().to(self.device)
        self.base = torch.tensor(base, dtype=torch.int64)
        self._base_dev = self.base.to(self.device)

    def __len__(self):
        return len(self.sizes)

    def plan(self, scene_ids: Sequence[int], subsample: int) -> torch.Tensor:
        """window starts [B, 2], consuming `random.randint` exactly as the reference's per-item calls would"""
        half = int(subsample / 2)
        win = []
        for s in scene_ids:
            ps, ns = self.sizes[int(s)]
            if ns <= half or ps < half:
                raise NotImplementedError("a scene with fewer than subsample/2 rows of one sign takes the reference's "
                                          "index_select fallback (data.py:123-126), which is not a contiguous window")
            a = random.randint(0, ps - half)
            b = random.randint(0, ns - half)
            win.append((a, b))
        return torch.tensor(win, dtype=torch.int64)

    @torch.no_grad()
    def batch(self, scene_ids: Sequence[int], subsample: int, windows: Optional[torch.Tensor] = None) -> torch.Tensor:
        half = int(subsample / 2)
        ids = torch.as_tensor(list(scene_ids), dtype=torch.int64)
        win = self.plan(ids.tolist(), subsample) if windows is None else windows
        L = _lib.lib()
        with torch.cuda.device(self.device):
            out = torch.empty(len(ids), 2 * half, 4, dtype=torch.float32, device=self.device)
            base = self._base_dev.index_select(0, ids.to(self.device)).contiguous()
            wd = win.to(self.device)
            _lib.check(L.dsdf_gather_windows(self.samples.data_ptr(), base.data_ptr(), wd.data_ptr(), len(ids), half,
                                             out.data_ptr(), _st(self.device)), None, L)
        return out

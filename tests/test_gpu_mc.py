"""GPU iso-surface extraction (reference deep_sdf/mesh.py:96-165) through the C ABI
(dsdf_mc_classify / dsdf_mc_emit) against the CPU restatement oracle/marching_cubes.py: vertices bit-exact and in the same
order, faces identical; at create_mesh's full size (256^3) through the table-free vertex enumeration and the mesh
invariants; and `create_mesh` end to end (decode + extraction on the device, PLY on disk)."""
import numpy as np
import pytest
import torch

import dsdf_b200
from oracle import marching_cubes as omc

pytestmark = pytest.mark.gpu


def _sphere(n, r=0.6, c=(0.03, -0.02, 0.01)):
    ax = np.linspace(-1, 1, n, dtype=np.float32)
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    return np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2).astype(np.float32) - np.float32(r)


@pytest.mark.parametrize("shape,kind", [((14, 14, 14), "sphere"), ((9, 7, 6), "noise"), ((2, 2, 2), "noise"),
                                        ((5, 17, 3), "noise"), ((12, 12, 12), "two_blobs")])
def test_kernel_equals_oracle(shape, kind):
    rng = np.random.default_rng(3)
    if kind == "sphere":
        vol = _sphere(shape[0])
    elif kind == "noise":        # white noise: every one of the 256 cube configurations, ambiguous faces included
        vol = rng.standard_normal(shape).astype(np.float32)
        vol[0, 0, 0], vol[-1, -1, -1] = -1.0, 1.0
    else:
        vol = np.minimum(_sphere(12, 0.35, (-0.4, 0, 0)), _sphere(12, 0.3, (0.45, 0.1, 0)))
    sp = (0.5, 0.25, 2.0)
    v_ref, f_ref = omc.marching_cubes(vol, 0.0, sp)
    v, f = dsdf_b200.marching_cubes(torch.from_numpy(vol).cuda(), 0.0, sp)
    assert np.array_equal(v.cpu().numpy(), v_ref)                 # bit-exact, same order
    assert np.array_equal(f.cpu().numpy(), f_ref)
    # a level other than 0
    lvl = 0.17
    v_ref, f_ref = omc.marching_cubes(vol, lvl, sp)
    v, f = dsdf_b200.marching_cubes(torch.from_numpy(vol).cuda(), lvl, sp)
    assert np.array_equal(v.cpu().numpy(), v_ref) and np.array_equal(f.cpu().numpy(), f_ref)


def test_empty_surface_raises_like_skimage_and_the_caller_returns_false(tmp_path):
    vol = torch.ones(8, 8, 8, device="cuda")
    with pytest.raises(ValueError):
        dsdf_b200.marching_cubes(vol)
    assert dsdf_b200.convert_sdf_samples_to_ply(vol.cpu(), [-1, -1, -1], 2 / 7, str(tmp_path / "x.ply")) is False
    with pytest.raises(dsdf_b200._lib.DsdfError):
        dsdf_b200.marching_cubes(torch.ones(8, 8, 8))            # host tensor: no CPU path


def test_full_size_256_grid_invariants_and_vertex_set():
    n = 256
    vol = torch.from_numpy(_sphere(n)).cuda()
    vs = 2.0 / (n - 1)
    v, f = dsdf_b200.marching_cubes(vol, 0.0, (vs,) * 3)
    v, f = v.cpu().numpy(), f.cpu().numpy()
    want = omc.crossing_vertices(vol.cpu().numpy(), 0.0, (vs,) * 3)     # table-free numpy enumeration
    assert np.array_equal(v, want)
    inv = omc.mesh_invariants(v, f)
    assert inv["closed"] and inv["oriented"] and inv["all_half_edges_paired"] and inv["euler"] == 2, inv
    nrm = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
    ctr = v[f].mean(1) - 1.0 - np.array([0.03, -0.02, 0.01], dtype=np.float32)
    assert np.all(np.einsum("ij,ij->i", nrm, ctr) > 0)                  # normals towards increasing SDF
    assert np.abs(np.linalg.norm(v - 1.0 - np.array([0.03, -0.02, 0.01], dtype=np.float32), axis=1) - 0.6).max() < 1e-4


def test_create_mesh_end_to_end_on_the_device(tmp_path):
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False,
                            use_tanh=False, latent_dropout=False, weight_norm=True, norm_layers=list(range(8))).cuda()
    # an untrained decoder has no guaranteed zero crossing: shift lin8's bias to the median output on a coarse grid
    lat = torch.randn(1, 256, generator=torch.Generator().manual_seed(1)).cuda() * 0.05
    N = 64
    sdf0 = dsdf_b200.create_mesh_sdf(dec, lat, N, to_cpu=False)
    with torch.no_grad():
        pre = torch.atanh(sdf0.clamp(-0.999, 0.999))
        dec.lin8.bias -= pre.median()
    dec.sync_weights(force=True)
    name = str(tmp_path / "mesh")
    out = dsdf_b200.create_mesh(dec, lat, name, N=N, max_batch=2 ** 18, offset=np.array([0.1, 0.0, -0.2], dtype=np.float32),
                                scale=2.0, return_trimesh=True)
    v, f = out
    v_file, f_file = dsdf_b200.read_ply(name + ".ply")
    assert np.array_equal(v, v_file) and np.array_equal(f, f_file)
    # the same mesh from the oracle on the decoded grid (the grid itself is parity-tested in test_gpu_parity)
    sdf = dsdf_b200.create_mesh_sdf(dec, lat, N, to_cpu=True).numpy()
    vs = 2.0 / (N - 1)
    want_v = omc.crossing_vertices(sdf, 0.0, (vs,) * 3)
    want = omc.transform_vertices(want_v, [-1, -1, -1], np.array([0.1, 0.0, -0.2], dtype=np.float32), 2.0)
    assert v.shape == want.shape and np.abs(v - want).max() <= 1e-6
    inv = omc.mesh_invariants(v, f)
    assert inv["oriented"] and inv["degenerate_faces"] == 0 and f.max() < len(v) and f.min() >= 0

"""GPU SDF-sample generation (reference src/PreprocessMesh.cpp, src/Utils.cpp, src/ShaderProgram.cpp,
deep_sdf/data.py:112-136) through the C ABI against the CPU restatement oracle/preprocess.py: rendered position / normal
images bit-exact, the sample cloud from the same Philox words within libm rounding, the 11-NN sign vote bit-exact on the
same inputs, the whole pipeline against a sphere's analytic distance, and the HBM-resident batch sampler bit-exact
against the reference's per-item rule under the same `random` stream."""
import ctypes as C
import math
import random

import numpy as np
import pytest
import torch

import dsdf_b200
from common import icosphere
from dsdf_b200 import _lib, data as ddata, preprocess as P
from oracle import preprocess as opp

pytestmark = pytest.mark.gpu


def test_rendered_images_equal_the_oracle_bit_for_bit():
    v, f = icosphere(1, 0.5, (0.03, -0.02, 0.05))
    # an inner, inward-facing sphere: occluded everywhere, and two triangles sharing depth along an edge
    v2, f2 = icosphere(0, 0.2)
    verts = np.concatenate([v, v2]).astype(np.float32)
    faces = np.concatenate([f, f2[:, ::-1] + len(v)])
    eyes = P.equidist_points_on_sphere(100, 1.1)[[3, 50, 97]]
    views = P.look_at_views(eyes)
    pts, nrm, ov, on, ok = P.visible_surface_points(torch.from_numpy(verts).cuda(), torch.from_numpy(faces).cuda(), views,
                                                    return_images=True)
    for i, view in enumerate(views):
        vimg, nimg, valid = opp.render_view(verts, faces, view)
        assert np.array_equal(ok[i].cpu().numpy(), valid)
        assert np.array_equal(ov[i].cpu().numpy(), vimg)
        assert np.array_equal(on[i].cpu().numpy(), nimg)
    # compaction order = ValidPointsFromIm's walk: view, then x, then y
    want = np.concatenate([opp.render_view(verts, faces, vw)[0][opp.render_view(verts, faces, vw)[2] == 1] for vw in views[:1]])
    assert np.array_equal(pts[:len(want)].cpu().numpy(), want)
    assert pts.shape == nrm.shape and torch.equal(pts[:, 3], nrm[:, 3])
    assert int(pts[:, 3].max()) <= len(f)                       # the inner sphere is never visible


def test_sample_cloud_matches_the_oracle_from_the_same_philox_words():
    v, f = icosphere(2, 0.5)
    cdf = P.triangle_area_cdf(v, f)
    n_surf, n_rand, seed = 5000, 700, 0x1234ABCD9876
    s1, s2 = np.float32(math.sqrt(0.005)), np.float32(math.sqrt(0.0005))
    L = _lib.lib()
    vd, fd, cd = torch.from_numpy(v).cuda(), torch.from_numpy(f).to(torch.int32).cuda(), torch.from_numpy(cdf).cuda()
    xyz = torch.empty(2 * n_surf + n_rand, 3, device="cuda")
    surf = torch.empty(n_surf, 3, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(L.dsdf_sample_points(vd.data_ptr(), fd.data_ptr(), cd.data_ptr(), len(f), n_surf, n_rand, float(s1), float(s2),
                                    2.0, seed, xyz.data_ptr(), surf.data_ptr(), st), None, L)
    want, p, tri = opp.sample_points(v, f, cdf, n_surf, n_rand, s1, s2, 2.0, seed)
    assert np.abs(surf.cpu().numpy() - p).max() <= 2e-6        # same triangle picked, same barycentric draw
    assert np.abs(xyz.cpu().numpy() - want).max() <= 5e-6      # Box-Muller through two libms
    got = xyz.cpu().numpy()
    assert np.array_equal(got[2 * n_surf:], want[2 * n_surf:])  # the uniform part has no libm call: bit-exact
    assert np.abs(got[2 * n_surf:]).max() <= 1.0
    d1 = got[0:2 * n_surf:2] - p
    d2 = got[1:2 * n_surf:2] - p
    assert abs(d1.std() / s1 - 1) < 0.03 and abs(d2.std() / s2 - 1) < 0.03 and abs(d1.mean()) < 3e-3
    # surface points are area-uniform: they lie on the polyhedron
    assert np.abs(np.linalg.norm(p, axis=1) - 0.5).max() < 0.5 * 0.03


def test_knn_sign_vote_equals_the_oracle_on_the_same_inputs():
    v, f = icosphere(2, 0.5)
    vd, fd = torch.from_numpy(v).cuda(), torch.from_numpy(f).cuda()
    views = P.look_at_views(P.equidist_points_on_sphere(100, 1.1)[::9])
    pts, nrm = P.visible_surface_points(vd, fd, views)
    sub = torch.arange(0, pts.shape[0], 23, device="cuda")       # ~12 k points: the oracle is O(n q) numpy
    pts, nrm = pts[sub].contiguous(), nrm[sub].contiguous()
    g = torch.Generator().manual_seed(5)
    q = torch.cat([(torch.rand(600, 3, generator=g) * 2 - 1),
                   torch.nn.functional.normalize(torch.randn(900, 3, generator=g), dim=1) * 0.5 +
                   torch.randn(900, 3, generator=g) * 0.03]).cuda()
    stdv = np.float32(math.sqrt(0.005))
    L = _lib.lib()
    sdf = torch.empty(len(q), device="cuda")
    nn = torch.empty(len(q), 11, dtype=torch.int32, device="cuda")
    _lib.check(L.dsdf_knn_sdf(q.data_ptr(), len(q), pts.data_ptr(), nrm.data_ptr(), pts.shape[0], float(stdv), sdf.data_ptr(),
                              nn.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream)), None, L)
    want, want_nn = opp.knn_sdf(q.cpu().numpy(), pts.cpu().numpy(), nrm.cpu().numpy(), stdv)
    assert np.array_equal(nn.cpu().numpy(), want_nn)
    got = sdf.cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])
    assert 0.005 < np.isnan(got).mean() < 0.6                   # both outcomes of the vote are exercised


def _same_knn(q, pts, nrm, stdv):
    a, ann = P.nearest_visible_sdf(q, pts, nrm, stdv, method="brute", return_neighbours=True)
    b, bnn = P.nearest_visible_sdf(q, pts, nrm, stdv, method="grid", return_neighbours=True)
    assert torch.equal(ann, bnn)
    assert torch.equal(a.view(torch.int32), b.view(torch.int32))             # NaNs included
    return a, ann


def test_grid_search_equals_the_all_pairs_search_bit_for_bit():
    """dsdf_knn_sdf_grid (uniform grid, growing shells) against dsdf_knn_sdf (all pairs), itself pinned to the oracle
    above: neighbour lists and distances identical -- near and far queries, queries outside the points' bounding box,
    duplicated points (distance ties go to the lower index), fewer than 11 points, degenerate point sets."""
    v, f = icosphere(3, 0.5, (0.05, 0.0, -0.02))
    vd, fd = torch.from_numpy(v).cuda(), torch.from_numpy(f).cuda()
    pts, nrm = P.visible_surface_points(vd, fd, P.look_at_views(P.equidist_points_on_sphere(100, 1.1)[::5]))
    assert pts.shape[0] > 200000
    g = torch.Generator().manual_seed(7)
    q = torch.cat([torch.rand(3000, 3, generator=g) * 2 - 1,                              # uniform in the cube
                   torch.nn.functional.normalize(torch.randn(6000, 3, generator=g), dim=1) * 0.5 +
                   torch.randn(6000, 3, generator=g) * 0.02,                              # hugging the surface
                   torch.randn(500, 3, generator=g) * 3.0,                                # far outside the bounding box
                   pts[::997, :3].cpu()]).cuda()                                          # exactly on visible points
    stdv = float(np.float32(math.sqrt(0.005)))
    sdf, nn = _same_knn(q, pts, nrm, stdv)
    assert 0.0 < float(torch.isnan(sdf).float().mean()) < 0.6
    # ties: every point three times (and shuffled), so each distance occurs at three indices
    perm = torch.randperm(3 * 20000, generator=g).cuda()
    p3 = pts[:20000].repeat(3, 1)[perm].contiguous()
    n3 = nrm[:20000].repeat(3, 1)[perm].contiguous()
    _, nn3 = _same_knn(q[:4000], p3, n3, stdv)
    d = (q[:4000, None, :] - p3[nn3.long()][:, :, :3]).norm(dim=2)
    assert bool((d[:, 1:] >= d[:, :-1] - 1e-6).all())
    same_pt = (p3[nn3[:, 1:].long()][:, :, :3] == p3[nn3[:, :-1].long()][:, :, :3]).all(2)   # copies of one point: exact ties
    assert bool(same_pt.any()) and bool((nn3[:, 1:][same_pt] > nn3[:, :-1][same_pt]).all())  # ... lower index first
    # the same tie rule as the oracle's stable argsort, on a set small enough for it (each point three times)
    perm_s = torch.randperm(3 * 1500, generator=g).cuda()
    ps = pts[::400][:1500].repeat(3, 1)[perm_s].contiguous()
    ns = nrm[::400][:1500].repeat(3, 1)[perm_s].contiguous()
    qs = torch.cat([q[:150], q[3000:3150]])
    s_s, nn_s = _same_knn(qs, ps, ns, stdv)
    want_s, want_nn = opp.knn_sdf(qs.cpu().numpy(), ps.cpu().numpy(), ns.cpu().numpy(), np.float32(stdv))
    assert np.array_equal(nn_s.cpu().numpy(), want_nn)
    assert np.array_equal(s_s.cpu().numpy().view(np.int32), want_s.view(np.int32))
    # fewer than 11 points: every vote is incomplete -> NaN, the list is padded with -1
    s7, nn7 = _same_knn(q[:100], pts[:7].contiguous(), nrm[:7].contiguous(), stdv)
    assert bool(torch.isnan(s7).all()) and bool((nn7[:, 7:] == -1).all()) and bool((nn7[:, :7] >= 0).all())
    # degenerate sets: all points identical; all points in one plane
    one = pts[:1].repeat(50, 1).contiguous()
    _same_knn(q[:200], one, nrm[:50].contiguous(), stdv)
    flat = pts[:5000].clone(); flat[:, 2] = 0.125
    _same_knn(q[:2000], flat.contiguous(), nrm[:5000].contiguous(), stdv)


def test_preprocess_mesh_end_to_end_against_the_analytic_sphere():
    v, f = icosphere(4, 0.5, (0.2, -0.1, 0.05))                  # 5120 faces; the pipeline re-centres it
    pos, neg = dsdf_b200.preprocess_mesh(v, f, num_sample=60000, seed=3)
    assert pos.shape[1] == 4 and neg.shape[1] == 4 and pos.is_cuda
    assert bool((pos[:, 3] > 0).all()) and bool((neg[:, 3] <= 0).all())
    kept = pos.shape[0] + neg.shape[0]
    assert kept > 0.85 * 60000                                    # few split votes on a convex shape
    rows = torch.cat([pos, neg]).cpu().numpy()
    true = np.linalg.norm(rows[:, :3], axis=1) - 0.5
    far = np.abs(true) > 0.01
    assert np.all(np.sign(rows[far, 3]) == np.sign(true[far]))
    assert np.abs(rows[:, 3] - true).max() < 0.012                # 0.005 pixel pitch + facet sagitta
    near = np.abs(true) < 0.07                                    # inside stdv: point-plane distance
    assert np.abs(rows[near, 3] - true[near]).mean() < 2e-3
    # 47/50 of the cloud hugs the surface (two variances), the rest is uniform in the cube
    assert 0.80 < (np.abs(true) < 0.25).mean() < 0.99
    # deterministic in the seed
    pos2, neg2 = dsdf_b200.preprocess_mesh(v, f, num_sample=60000, seed=3)
    assert torch.equal(pos, pos2) and torch.equal(neg, neg2)
    pos3, _ = dsdf_b200.preprocess_mesh(v, f, num_sample=60000, seed=4)
    assert pos3.shape != pos.shape or not torch.equal(pos3, pos)


def test_gpu_resident_batches_equal_the_reference_sampling_rule():
    g = torch.Generator().manual_seed(1)
    scenes = []
    for s in range(5):
        npos, nneg = 3000 + 137 * s, 2500 + 91 * s
        pos = torch.rand(npos, 4, generator=g); pos[:, 3] = pos[:, 3].abs() + 1e-3
        neg = torch.rand(nneg, 4, generator=g); neg[:, 3] = -neg[:, 3].abs()
        scenes.append((pos, neg))
    store = dsdf_b200.GpuSDFSamples(scenes, shuffle=False)
    ids = [3, 0, 4, 3]
    random.seed(77)
    got = store.batch(ids, 2048)
    random.seed(77)
    want = torch.stack([ddata.unpack_sdf_samples_from_ram(list(scenes[i]), 2048) for i in ids])
    assert got.shape == (4, 2048, 4) and torch.equal(got.cpu(), want)
    # straight into the fused step (the [B, P, 4] layout train_deep_sdf.py:555-556 gets from the DataLoader)
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False, use_tanh=False,
                            latent_dropout=False, weight_norm=True, norm_layers=list(range(8))).cuda()
    tr = dsdf_b200.FusedTrainer(dec, 5, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0))
    st = tr.step(got, torch.tensor(ids).cuda(), 1)
    assert math.isfinite(float(st["sdf_loss"]))


@pytest.mark.parametrize("n_faces_subdiv,num_sample", [(5, 500000)])
def test_full_size_cli_defaults(n_faces_subdiv, num_sample):
    """PreprocessMesh's defaults (-s 500000, 100 views, 11 votes) on a 20 480-face mesh."""
    import time
    v, f = icosphere(n_faces_subdiv, 0.6)
    torch.cuda.synchronize()
    t0 = time.time()
    pos, neg = dsdf_b200.preprocess_mesh(v, f, num_sample=num_sample, seed=1)
    torch.cuda.synchronize()
    dt = time.time() - t0
    n = pos.shape[0] + neg.shape[0]
    print(f"preprocess_mesh 20480 faces, {num_sample} samples: {dt:.2f} s, kept {n} ({pos.shape[0]} pos / {neg.shape[0]} neg)")
    assert n > 0.9 * num_sample
    rows = torch.cat([pos, neg])
    true = rows[:, :3].norm(dim=1) - 0.6
    assert float((rows[:, 3] - true).abs().max()) < 0.01
    assert 0.35 < pos.shape[0] / n < 0.65

"""CPU sanity of the SDF-sample-generation oracle: the software rasteriser restating
src/ShaderProgram.cpp and the 11-NN sign vote of src/PreprocessMesh.cpp:137-172 behave as the reference describes on a
shape with a known answer; host-side pieces of dsdf_b200.preprocess (views, CDF)."""
import numpy as np

from common import icosphere
from dsdf_b200 import preprocess as P
from oracle import preprocess as opp


def test_views_are_the_reference_spiral_and_orthonormal_look_at_frames():
    eyes = P.equidist_points_on_sphere(100, 1.1)
    assert eyes.shape == (100, 3) and np.allclose(np.linalg.norm(eyes, axis=1), 1.1, atol=1e-5)
    assert np.allclose(eyes[:, 1], np.linspace(-1 + 0.01, 1 - 0.01, 100) * 1.1, atol=1e-5)       # y = i * 2/n - 1 + 1/n
    V = P.look_at_views(eyes)
    for e, v in zip(eyes, V):
        R, t = v.reshape(3, 4)[:, :3], v.reshape(3, 4)[:, 3]
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-5)
        assert np.allclose(R @ e + t, 0, atol=1e-5)                      # the eye is the camera origin
        assert np.allclose(R @ (0 * e) + t, [0, 0, -1.1], atol=1e-5)     # the target sits 1.1 in front (camera looks down -z)


def test_area_cdf_matches_heron_and_is_monotone():
    v, f = icosphere(1, 0.5)
    cdf = P.triangle_area_cdf(v, f)
    a, b, c = v[f[:, 0]].astype(np.float64), v[f[:, 1]].astype(np.float64), v[f[:, 2]].astype(np.float64)
    area = 0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1)
    assert np.all(np.diff(cdf) > 0) and abs(cdf[-1] - area.sum()) < 1e-4 * area.sum()


def test_oracle_render_of_a_sphere_one_view():
    v, f = icosphere(1, 0.5)
    eye = P.equidist_points_on_sphere(100, 1.1)[37]
    view = P.look_at_views(eye[None])[0]
    vimg, nimg, valid = opp.render_view(v, f, view)
    n = int(valid.sum())
    assert 0.90 * np.pi * 100 ** 2 < n <= np.pi * 100 ** 2               # disc of radius 0.5 = 100 pixels, minus the 80-face silhouette
    pts, nrm = vimg[valid == 1], nimg[valid == 1]
    r = np.linalg.norm(pts[:, :3], axis=1)
    assert r.min() > 0.5 * 0.93 and r.max() <= 0.5 + 1e-6                # on the inscribed polyhedron
    side = pts[:, :3] @ eye                                               # the near hemisphere only (depth test); the limb facets
    assert side.min() > -0.15 and (side > 0).mean() > 0.90               # of an 80-face sphere reach a little behind the equator
    assert np.all(np.einsum("ij,ij->i", nrm[:, :3], pts[:, :3]) > 0)     # normals turned towards the camera = outward here
    assert np.array_equal(pts[:, 3], nrm[:, 3]) and pts[:, 3].min() >= 1 and pts[:, 3].max() <= len(f)
    assert np.allclose(np.linalg.norm(nrm[:, :3], axis=1), 1, atol=1e-5)


def test_oracle_knn_sign_vote_on_a_sphere_point_cloud():
    rng = np.random.default_rng(0)
    p = rng.standard_normal((4000, 3))
    p = (p / np.linalg.norm(p, axis=1, keepdims=True) * 0.5).astype(np.float32)
    nrm = np.concatenate([p / 0.5, np.ones((len(p), 1), np.float32)], 1).astype(np.float32)
    pts = np.concatenate([p, np.ones((len(p), 1), np.float32)], 1).astype(np.float32)
    q = (rng.random((300, 3)) * 2 - 1).astype(np.float32)
    sdf, nn = opp.knn_sdf(q, pts, nrm, np.sqrt(0.005))
    true = np.linalg.norm(q, axis=1) - 0.5
    ok = ~np.isnan(sdf)
    assert ok.mean() > 0.9
    assert np.all(np.sign(sdf[ok]) == np.sign(true[ok]))
    assert np.abs(sdf[ok] - true[ok]).max() < 0.03                       # point spacing of a 4000-point cloud
    assert np.all(nn[:, 0] == np.argmin(((q[:, None] - p[None]) ** 2).sum(-1), 1))


def test_grid_search_order_and_stopping_rule_give_the_exact_neighbours():
    """The algorithm of k_knn_sdf_grid, restated in numpy (oracle.preprocess.knn_grid_neighbours), against the stable
    argsort of all distances: points on a sphere with every point duplicated (distance ties), queries near the surface,
    far inside, outside the points' bounding box, and exactly on points; coarse and fine grids."""
    rng = np.random.default_rng(3)
    p = rng.normal(size=(1500, 3)); p = (0.5 * p / np.linalg.norm(p, axis=1, keepdims=True)).astype(np.float32)
    p = np.concatenate([p, p[::3]])[rng.permutation(2000)]
    n = (p / np.linalg.norm(p, axis=1, keepdims=True)).astype(np.float32)
    q = np.concatenate([rng.uniform(-1, 1, (40, 3)), p[:40] + rng.normal(size=(40, 3)) * 0.03, rng.normal(size=(15, 3)) * 3,
                        p[100:110], np.zeros((1, 3))]).astype(np.float32)
    _, want = opp.knn_sdf(q, np.c_[p, np.zeros(len(p))], np.c_[n, np.zeros(len(p))], np.float32(0.07))
    for grid, block in ((16, 4), (32, 8), (8, 2)):
        got = opp.knn_grid_neighbours(q, p, grid=grid, block=block)
        assert np.array_equal(got, want), (grid, block)
    # fewer points than neighbours; all points in one spot
    assert np.array_equal(opp.knn_grid_neighbours(q[:5], p[:7], grid=16, block=4)[:, 7:], np.full((5, 4), -1))
    one = np.repeat(p[:1], 30, 0)
    assert np.array_equal(opp.knn_grid_neighbours(q[:5], one, grid=16, block=4), np.tile(np.arange(11), (5, 1)))

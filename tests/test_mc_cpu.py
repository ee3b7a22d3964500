"""CPU checks of the iso-surface extraction pieces: the oracle's invariants, the kernel's generated
case table against the oracle's table-free tracing for all 256 cases, and the PLY writer."""
import importlib.util
import os

import numpy as np

from oracle import marching_cubes as omc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gen():
    spec = importlib.util.spec_from_file_location("gen_mc_tables", os.path.join(ROOT, "tools", "gen_mc_tables.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _sphere(n, r=0.6, c=(0.03, -0.02, 0.01)):
    ax = np.linspace(-1, 1, n, dtype=np.float32)
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    return np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2).astype(np.float32) - np.float32(r)


def test_generated_table_equals_table_free_tracing_for_all_256_cases():
    gen = _gen()
    for case in range(256):
        inside = {p: bool((case >> c) & 1) for c, p in enumerate(omc._CORNERS)}
        want = omc.cube_triangles(inside)
        got = gen.case_triangles(case)
        # edge id (generator) -> corner-offset pair (oracle)
        def edge_pair(e):
            axis, a, b = e >> 2, e & 1, (e >> 1) & 1
            lo = [0, 0, 0]
            others = [i for i in range(3) if i != axis]
            lo[others[0]], lo[others[1]] = a, b
            hi = list(lo)
            hi[axis] = 1
            return (tuple(lo), tuple(hi))
        got_p = [tuple(edge_pair(e) for e in t) for t in got]
        norm = lambda tris: sorted(min(t[i:] + t[:i] for i in range(3)) for t in tris)   # rotation-invariant, keeps winding
        assert norm(got_p) == norm(want), case
        assert len(got) <= 5


def test_committed_table_header_is_what_the_generator_emits():
    import io
    import contextlib
    gen = _gen()
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        gen.main()
    path = os.path.join(ROOT, "medical-shape-disentanglement_b200", "csrc", "mc_tables.cuh")
    assert open(path).read() == buf.getvalue()


def test_oracle_sphere_is_a_closed_oriented_genus_0_surface_with_outward_normals():
    vol = _sphere(14)
    v, f = omc.marching_cubes(vol, 0.0, (2 / 13,) * 3)
    inv = omc.mesh_invariants(v, f)
    assert inv["closed"] and inv["oriented"] and inv["euler"] == 2 and inv["degenerate_faces"] == 0, inv
    assert np.array_equal(v, omc.crossing_vertices(vol, 0.0, (2 / 13,) * 3))       # bit-equal, same order
    p = v - 1.0 + np.array([-0.03, 0.02, -0.01], dtype=np.float32)                   # relative to the centre
    n = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
    assert np.all(np.einsum("ij,ij->i", n, p[f].mean(1)) > 0)                       # towards increasing SDF
    r = np.linalg.norm(p, axis=1)
    assert np.abs(r - 0.6).max() < 0.01                                              # on the sphere (linear interp)


def test_oracle_ambiguous_faces_stay_crack_free():
    rng = np.random.default_rng(0)
    vol = rng.standard_normal((7, 6, 5)).astype(np.float32)      # noise: every ambiguous configuration shows up
    v, f = omc.marching_cubes(vol, 0.0)
    inv = omc.mesh_invariants(v, f)
    # closed except where the surface leaves the grid: boundary edges lie on the grid's outer faces only
    f64 = f.astype(np.int64)
    he = np.concatenate([f64[:, [0, 1]], f64[:, [1, 2]], f64[:, [2, 0]]])
    und, cnt = np.unique(np.sort(he, axis=1), axis=0, return_counts=True)
    assert cnt.max() == 2 and inv["oriented"]
    hi = np.array(vol.shape, dtype=np.float32) - 1
    for a, b in und[cnt == 1]:
        on_face = ((v[a] == 0) & (v[b] == 0)) | ((v[a] == hi) & (v[b] == hi))
        assert on_face.any(), (v[a], v[b])


def test_ply_round_trip_and_header(tmp_path):
    import dsdf_b200
    v = np.random.default_rng(1).standard_normal((11, 3)).astype(np.float32)
    f = np.random.default_rng(2).integers(0, 11, (7, 3)).astype(np.int32)
    path = str(tmp_path / "m.ply")
    dsdf_b200.write_ply(path, v, f)
    head = open(path, "rb").read(200).decode("ascii", "ignore")
    assert head.startswith("ply\nformat binary_little_endian 1.0\nelement vertex 11\nproperty float x\nproperty float y\n"
                           "property float z\nelement face 7\nproperty list uchar int vertex_indices\nend_header\n")
    v2, f2 = dsdf_b200.read_ply(path)
    assert np.array_equal(v, v2) and np.array_equal(f, f2)
    assert os.path.getsize(path) == head.index("end_header\n") + 11 + 11 * 12 + 7 * 13


def test_vertex_transform_matches_mesh_py():
    v = np.array([[0.5, 1.0, 1.5]], dtype=np.float32)
    out = omc.transform_vertices(v, [-1, -1, -1], offset=np.array([0.1, 0.2, 0.3], dtype=np.float32), scale=2.0)
    assert np.allclose(out, (v - 1) / 2 - np.array([0.1, 0.2, 0.3]))

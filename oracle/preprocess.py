"""CPU restatement (numpy) of the reference's SDF-sample generation, src/PreprocessMesh.cpp + src/Utils.cpp +
src/ShaderProgram.cpp, for the pieces csrc/sampler_kernels.cuh computes.  TEST INFRASTRUCTURE: imported only by tests/.

Parity status: the reference renders through OpenGL (pangolin), which is unavailable on this box, and links nanoflann /
Eigen / cnpy that are not vendored; the binary cannot be built or run here, and its RNG is seeded from
std::random_device, so no golden vectors can exist -- "parity unpinned" for this row.  What this file restates is the
shader MATH (ShaderProgram.cpp:5-141: world position varying, flat normal turned towards the camera, the |d| < 0.001
grazing cut) behind a plain software rasteriser with OpenGL's conventions (pixel centres, top-left rule, GL_LESS in
submission order), and SampleSDFNearSurface's arithmetic (:137-172) around a brute-force exact 11-NN.
All fp32, one rounding per operation, in the same operation order as the kernels -> bit-comparable.
"""
from __future__ import annotations

import numpy as np

F = np.float32
W = H = 400


def _mad3(a0, b0, a1, b1, a2, b2, c):
    return F(F(F(F(a0 * b0) + F(a1 * b1)) + F(a2 * b2)) + c)


def _edge(ax, ay, bx, by, px, py):
    return F(F(F(bx - ax) * F(py - ay)) - F(F(by - ay) * F(px - ax)))


def _owns(ax, ay, bx, by):
    return (ay == by and bx < ax) or (by > ay)


def _setup(verts, tri, view):
    wx, wy, z = [], [], []
    for vi in tri:
        px, py, pz = (F(c) for c in verts[vi])
        xc = _mad3(view[0], px, view[1], py, view[2], pz, view[3])
        yc = _mad3(view[4], px, view[5], py, view[6], pz, view[7])
        zc = _mad3(view[8], px, view[9], py, view[10], pz, view[11])
        wx.append(F(F(xc + F(1)) * F(0.5 * W)))
        wy.append(F(F(-yc + F(1)) * F(0.5 * H)))
        z.append(F(-zc))
    return wx, wy, z


def _cover(wx, wy, px, py):
    cx, cy = F(px + 0.5), F(py + 0.5)
    area = _edge(wx[0], wy[0], wx[1], wy[1], wx[2], wy[2])
    if area == 0 or np.isnan(area):
        return None
    e0 = _edge(wx[1], wy[1], wx[2], wy[2], cx, cy)
    e1 = _edge(wx[2], wy[2], wx[0], wy[0], cx, cy)
    e2 = _edge(wx[0], wy[0], wx[1], wy[1], cx, cy)
    if area < 0:
        area, e0, e1, e2 = F(-area), F(-e0), F(-e1), F(-e2)
        o0, o1, o2 = _owns(wx[2], wy[2], wx[1], wy[1]), _owns(wx[0], wy[0], wx[2], wy[2]), _owns(wx[1], wy[1], wx[0], wy[0])
    else:
        o0, o1, o2 = _owns(wx[1], wy[1], wx[2], wy[2]), _owns(wx[2], wy[2], wx[0], wy[0]), _owns(wx[0], wy[0], wx[1], wy[1])
    if e0 < 0 or e1 < 0 or e2 < 0:
        return None
    if (e0 == 0 and not o0) or (e1 == 0 and not o1) or (e2 == 0 and not o2):
        return None
    return F(e0 / area), F(e1 / area), F(e2 / area)


def _bary(b, v0, v1, v2):
    return F(F(F(b[0] * v0) + F(b[1] * v1)) + F(b[2] * v2))


def render_view(verts, faces, view):
    """One view: (vert_img [W,H,4], norm_img [W,H,4], valid [W,H]) indexed [x][y] as Utils.cpp:28-43 walks them."""
    verts = np.asarray(verts, dtype=np.float32)
    view = np.asarray(view, dtype=np.float32)
    depth = np.full((H, W), np.inf, dtype=np.float32)
    winner = np.full((H, W), -1, dtype=np.int64)
    with np.errstate(all="ignore"):
        for t, tri in enumerate(faces):
            wx, wy, z = _setup(verts, tri, view)
            if any(np.isnan(v) for v in wx + wy):
                continue
            x0, x1 = max(0, int(np.floor(min(wx) - F(0.5)))), min(W - 1, int(np.ceil(max(wx) - F(0.5))))
            y0, y1 = max(0, int(np.floor(min(wy) - F(0.5)))), min(H - 1, int(np.ceil(max(wy) - F(0.5))))
            for py in range(y0, y1 + 1):
                for px in range(x0, x1 + 1):
                    b = _cover(wx, wy, px, py)
                    if b is None:
                        continue
                    d = _bary(b, z[0], z[1], z[2])
                    if not (d >= 0 and d <= F(2.5)):
                        continue
                    if d < depth[py, px]:                       # GL_LESS, triangles submitted in index order
                        depth[py, px] = d
                        winner[py, px] = t
        vimg = np.zeros((W, H, 4), dtype=np.float32)
        nimg = np.zeros((W, H, 4), dtype=np.float32)
        valid = np.zeros((W, H), dtype=np.uint8)
        for py in range(H):
            for px in range(W):
                t = winner[py, px]
                if t < 0:
                    continue
                tri = faces[t]
                wx, wy, z = _setup(verts, tri, view)
                b = _cover(wx, wy, px, py)
                A, B, Cc = (verts[i].astype(np.float32) for i in tri)
                u = (B - A).astype(np.float32)
                w = (Cc - A).astype(np.float32)
                n = np.array([F(F(u[1] * w[2]) - F(u[2] * w[1])), F(F(u[2] * w[0]) - F(u[0] * w[2])),
                              F(F(u[0] * w[1]) - F(u[1] * w[0]))], dtype=np.float32)
                nl = np.sqrt(F(F(F(n[0] * n[0]) + F(n[1] * n[1])) + F(n[2] * n[2])))
                n = (n / nl).astype(np.float32)
                cn = [_mad3(view[4 * r], n[0], view[4 * r + 1], n[1], view[4 * r + 2], n[2], F(0)) for r in range(3)]
                cl = np.sqrt(F(F(F(cn[0] * cn[0]) + F(cn[1] * cn[1])) + F(cn[2] * cn[2])))
                d = F(cn[2] / cl)
                if abs(d) < F(0.001):
                    continue
                s = F(-1) if d < 0 else F(1)
                vimg[px, py] = [_bary(b, A[0], B[0], Cc[0]), _bary(b, A[1], B[1], Cc[1]), _bary(b, A[2], B[2], Cc[2]), F(t) + F(1)]
                nimg[px, py] = [s * n[0], s * n[1], s * n[2], F(t) + F(1)]
                valid[px, py] = 1
    return vimg, nimg, valid


def knn_sdf(xyz, pts, nrm, stdv, k=11):
    """SampleSDFNearSurface :137-172 for every row of xyz: (sdf [n] with NaN where the vote is split, nn [n,k])."""
    xyz = np.asarray(xyz, dtype=np.float32)
    p = np.asarray(pts, dtype=np.float32)[:, :3]
    m = np.asarray(nrm, dtype=np.float32)[:, :3]
    out = np.full(len(xyz), np.nan, dtype=np.float32)
    nn = np.full((len(xyz), k), -1, dtype=np.int32)
    stdv = F(stdv)
    with np.errstate(all="ignore"):
        for q, s in enumerate(xyz):
            dx, dy, dz = (s[0] - p[:, 0]).astype(np.float32), (s[1] - p[:, 1]).astype(np.float32), (s[2] - p[:, 2]).astype(np.float32)
            d2 = ((dx * dx).astype(np.float32) + (dy * dy).astype(np.float32)).astype(np.float32)
            d2 = (d2 + (dz * dz).astype(np.float32)).astype(np.float32)
            order = np.argsort(d2, kind="stable")[:k]             # ties -> lower index first
            nn[q, :len(order)] = order
            if len(order) < k:
                continue
            num_pos = 0
            sdf = F(0)
            for j, ci in enumerate(order):
                r = np.array([F(s[0] - p[ci, 0]), F(s[1] - p[ci, 1]), F(s[2] - p[ci, 2])], dtype=np.float32)
                ln = np.sqrt(F(F(F(r[0] * r[0]) + F(r[1] * r[1])) + F(r[2] * r[2])))
                if j == 0:
                    pl = F(F(F(m[ci, 0] * r[0]) + F(m[ci, 1] * r[1])) + F(m[ci, 2] * r[2]))
                    sdf = abs(pl) if ln < stdv else ln
                d = F(F(F(m[ci, 0] * F(r[0] / ln)) + F(m[ci, 1] * F(r[1] / ln))) + F(m[ci, 2] * F(r[2] / ln)))
                if d > 0:
                    num_pos += 1
            if num_pos == 0 or num_pos == k:
                out[q] = -sdf if num_pos <= k // 2 else sdf
    return out, nn


def knn_grid_neighbours(xyz, pts, k=11, grid=128, block=8, fine_r=1):
    """The visiting order and the stopping rule of k_knn_sdf_grid (csrc/sampler_kernels.cuh), restated for small inputs:
    cubic cells over the points' bounding box (block^3 cells = one coarse block), the query's own cell and fine_r shells
    around it, then shells of blocks with box-distance pruning of blocks and cells; stop when the k-th distance is
    strictly below the distance to the nearest face that still has cells behind it (minus the rounding margin).  Returns
    nn [n, k] (-1 padded).  Its claim -- tested in tests/test_preprocess_cpu.py -- is that this equals the stable
    argsort of all distances (knn_sdf above) whatever the geometry."""
    xyz = np.asarray(xyz, dtype=np.float32)
    p = np.asarray(pts, dtype=np.float32)[:, :3]
    G, B = grid, block
    C = G // B
    lo = p.min(0)
    ext = F(max(float((p.max(0) - lo).max()), 1e-6))
    h = F(F(ext * F(1 + 1e-3)) / F(G))
    inv_h = F(F(1) / h)
    org = (lo - F(0.5e-3) * ext).astype(np.float32)
    margin = F(F(1e-3) * h)

    def axis_cell(v, a):
        return int(min(max(np.floor(F(F(v - org[a]) * inv_h)), 0), G - 1))

    cells = {}
    for i, q in enumerate(p):
        cells.setdefault(tuple(axis_cell(q[a], a) for a in range(3)), []).append(i)
    blocks = {}
    for c in cells:
        blocks.setdefault(tuple(v // B for v in c), []).append(c)

    def shell(ctr, R, n):
        l, u = [c - R for c in ctr], [c + R for c in ctr]
        for z in range(max(l[2], 0), min(u[2], n - 1) + 1):
            for y in range(max(l[1], 0), min(u[1], n - 1) + 1):
                if z in (l[2], u[2]) or y in (l[1], u[1]):
                    xs = range(max(l[0], 0), min(u[0], n - 1) + 1)
                else:
                    xs = ([l[0]] if l[0] >= 0 else []) + ([u[0]] if R > 0 and u[0] <= n - 1 else [])
                for x in xs:
                    yield x, y, z

    def box_d2(q, corner, w):
        d2 = F(0)
        for a in range(3):
            t = max(max(F(corner[a] - q[a]), F(q[a] - F(corner[a] + w))) - margin, F(0))
            d2 = F(d2 + F(t * t))
        return d2

    nn = np.full((len(xyz), k), -1, dtype=np.int32)
    for qi, q in enumerate(xyz):
        d = (q - p).astype(np.float32)
        d2 = (((d[:, 0] * d[:, 0]).astype(np.float32) + (d[:, 1] * d[:, 1]).astype(np.float32)).astype(np.float32)
              + (d[:, 2] * d[:, 2]).astype(np.float32)).astype(np.float32)
        best = []                                               # sorted (distance, index) pairs, at most k

        def scan(idx):
            nonlocal best
            best = sorted(best + [(d2[i], i) for i in idx])[:k]

        def finished(ctr, R, n, w):
            df, op = F(3e38), False
            for a in range(3):
                if ctr[a] - R > 0:
                    df, op = min(df, F(q[a] - F(org[a] + F(ctr[a] - R) * w))), True
                if ctr[a] + R < n - 1:
                    df, op = min(df, F(F(org[a] + F(ctr[a] + R + 1) * w) - q[a])), True
            if not op:
                return True
            df = F(df - margin)
            return len(best) == k and df > 0 and best[k - 1][0] < F(df * df)

        c = [axis_cell(q[a], a) for a in range(3)]
        done = False
        for R in range(fine_r + 1):
            for cell in shell(c, R, G):
                scan(cells.get(cell, []))
            done = finished(c, R, G, h)
            if done:
                break
        if not done:
            cc, wc = [v // B for v in c], F(F(B) * h)
            for R in range(C):
                for blk in shell(cc, R, C):
                    if blk not in blocks:
                        continue
                    corner = [F(org[a] + F(blk[a] * B) * h) for a in range(3)]
                    if len(best) == k and box_d2(q, corner, wc) > best[k - 1][0]:
                        continue
                    for cell in sorted(blocks[blk], key=lambda t: (t[2] % B, t[1] % B, t[0] % B)):
                        if max(abs(cell[a] - c[a]) for a in range(3)) <= fine_r:
                            continue
                        if len(best) == k and box_d2(q, [F(org[a] + F(cell[a]) * h) for a in range(3)], h) > best[k - 1][0]:
                            continue
                        scan(cells[cell])
                if finished(cc, R, C, wc):
                    break
        for j, (_, i) in enumerate(best):
            nn[qi, j] = i
    return nn


def sample_points(verts, faces, cdf, n_surf, n_rand, stdv1, stdv2, cube_dim, seed):
    """The sample cloud of k_sample_points from the same Philox words (oracle/philox.py); libm calls (sqrt / log / sin /
    cos) are numpy's, so values agree with the device's to ~1e-6, not bit for bit."""
    from oracle import philox
    verts = np.asarray(verts, dtype=np.float32)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32)

    def words(idx, stream):
        ctr = np.zeros((len(idx), 4), dtype=np.uint32)
        ctr[:, 0] = (idx & 0xFFFFFFFF).astype(np.uint32)
        ctr[:, 1] = (idx >> 32).astype(np.uint32)
        ctr[:, 2] = stream
        return philox.philox4x32_10(ctr, key)

    u01 = lambda x: ((x >> np.uint32(8)).astype(np.float32) * F(1.0 / 16777216.0)).astype(np.float32)

    def normal2(a, b):
        u1 = (((a >> np.uint32(8)).astype(np.float32) + F(1)) * F(1.0 / 16777216.0)).astype(np.float32)
        u2 = u01(b)
        rr = np.sqrt(F(-2) * np.log(u1)).astype(np.float32)
        ang = (F(6.283185307179586) * u2).astype(np.float32)
        return (rr * np.cos(ang)).astype(np.float32), (rr * np.sin(ang)).astype(np.float32)

    i = np.arange(n_surf, dtype=np.uint64)
    r0, r1, r2 = words(i, 0), words(i, 1), words(i, 2)
    pick = (u01(r0[:, 0]) * cdf[-1]).astype(np.float32)
    tri = np.searchsorted(cdf, pick, side="left")               # lower_bound
    tri = np.minimum(tri, len(cdf) - 1)
    a, b, c = verts[faces[tri, 0]], verts[faces[tri, 1]], verts[faces[tri, 2]]
    sq, u2 = np.sqrt(u01(r0[:, 1])).astype(np.float32), u01(r0[:, 2])
    p = ((F(1) - sq)[:, None] * a + (sq * (F(1) - u2))[:, None] * b + (u2 * sq)[:, None] * c).astype(np.float32)
    g = np.zeros((n_surf, 6), dtype=np.float32)
    g[:, 0], g[:, 1] = normal2(r1[:, 0], r1[:, 1])
    g[:, 2], g[:, 3] = normal2(r1[:, 2], r1[:, 3])
    g[:, 4], g[:, 5] = normal2(r2[:, 0], r2[:, 1])
    xyz = np.zeros((2 * n_surf + n_rand, 3), dtype=np.float32)
    xyz[0:2 * n_surf:2] = p + F(stdv1) * g[:, 0::2]
    xyz[1:2 * n_surf:2] = p + F(stdv2) * g[:, 1::2]
    j = np.arange(n_surf, n_surf + n_rand, dtype=np.uint64)
    r = words(j, 3)
    xyz[2 * n_surf:] = (u01(r[:, :3]) * F(cube_dim) - F(cube_dim / 2)).astype(np.float32)
    return xyz, p, tri

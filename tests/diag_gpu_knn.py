"""diagnostic: where do the grid and the all-pairs nearest-neighbour searches differ?  python tests/diag_gpu_knn.py"""
import math, os, sys
import numpy as np, torch
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.join(HERE, "..")); sys.path.insert(0, os.path.join(HERE, "..", "medical-shape-disentanglement_b200"))
from common import icosphere
from dsdf_b200 import preprocess as P
v, f = icosphere(3, 0.5, (0.05, 0.0, -0.02))
vd, fd = torch.from_numpy(v).cuda(), torch.from_numpy(f).cuda()
pts, nrm = P.visible_surface_points(vd, fd, P.look_at_views(P.equidist_points_on_sphere(100, 1.1)[::5]))
g = torch.Generator().manual_seed(7)
q = torch.cat([torch.rand(3000, 3, generator=g) * 2 - 1,
               torch.nn.functional.normalize(torch.randn(6000, 3, generator=g), dim=1) * 0.5 + torch.randn(6000, 3, generator=g) * 0.02,
               torch.randn(500, 3, generator=g) * 3.0, pts[::997, :3].cpu()]).cuda()
stdv = float(np.float32(math.sqrt(0.005)))
a, ann = P.nearest_visible_sdf(q, pts, nrm, stdv, method="brute", return_neighbours=True)
b, bnn = P.nearest_visible_sdf(q, pts, nrm, stdv, method="grid", return_neighbours=True)
bad = (ann != bnn).any(1).nonzero().flatten()
print("points", pts.shape[0], "queries", q.shape[0], "mismatching queries", bad.numel(), "bbox", pts[:, :3].min(0).values.tolist(), pts[:, :3].max(0).values.tolist())
for i in bad[:12].tolist():
    da = ((q[i] - pts[ann[i].long(), :3]) ** 2).sum(1)
    db = ((q[i] - pts[bnn[i].long(), :3]) ** 2).sum(1)
    print(i, q[i].tolist())
    print("  brute", ann[i].tolist(), [f"{x:.9g}" for x in da.tolist()])
    print("  grid ", bnn[i].tolist(), [f"{x:.9g}" for x in db.tolist()])
print("region histogram of bad:", [(int(((bad >= lo) & (bad < hi)).sum())) for lo, hi in ((0, 3000), (3000, 9000), (9000, 9500), (9500, 10**9))])

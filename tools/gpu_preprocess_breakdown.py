"""Where preprocess_mesh (CLI defaults: 500 000 samples, 20 480-face mesh) spends its wall time.  python tools/gpu_preprocess_breakdown.py"""
import math, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "medical-shape-disentanglement_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import dsdf_b200
from dsdf_b200 import preprocess as P
from common import icosphere

def T():
    torch.cuda.synchronize(); return time.perf_counter()
v, f = icosphere(5, 0.6)
dsdf_b200.preprocess_mesh(v, f, num_sample=500000, seed=1)           # warm-up (library load, allocator)
for rep in range(2):
    t0 = T()
    vt = torch.as_tensor(v).float(); ft = torch.as_tensor(f).to(torch.int64)
    centre = torch.from_numpy(P.bounding_cube_center(vt.numpy(), ft.numpy()))
    vd = (vt - centre).cuda(); fd = ft.cuda()
    t1 = T()
    pts, nrm = P.visible_surface_points(vd, fd)
    t2 = T()
    cdf = torch.from_numpy(P.triangle_area_cdf(vd.cpu().numpy(), fd.cpu().numpy().astype(np.int64))).cuda()
    t3 = T()
    xyz, sdf = P.sample_sdf_near_surface(vd, fd, pts, nrm, 500000, 0.005, None, 1)
    t4 = T()
    pos, neg = dsdf_b200.preprocess_mesh(v, f, num_sample=500000, seed=1)
    t5 = T()
    print(f"centre+upload {1e3*(t1-t0):.1f} ms | rasterise 100 views + compaction {1e3*(t2-t1):.1f} ms ({pts.shape[0]} visible points) | "
          f"area cdf on the host {1e3*(t3-t2):.1f} ms | sample cloud + cdf again + 11-NN {1e3*(t4-t3):.1f} ms | whole call {1e3*(t5-t4):.1f} ms")

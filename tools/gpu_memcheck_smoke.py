"""Small end-to-end pass over every kernel family for `compute-sanitizer --tool memcheck` (tiny sizes: the tool slows
kernels by 10-100x):  compute-sanitizer --tool memcheck python tools/gpu_memcheck_smoke.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-shape-disentanglement_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import torch
import dsdf_b200
from dsdf_b200 import synth, preprocess as P
from common import icosphere

torch.manual_seed(0)
kw = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False, use_tanh=False,
          latent_dropout=False, weight_norm=True, norm_layers=list(range(8)))
dec = dsdf_b200.Decoder(256, **kw).cuda()
S, B, Pn = 6, 4, 384                                       # 12 tiles: CTA pairs, two tiles per CTA on a few
data = synth.scene_samples(S, Pn, seed=1).cuda()
ids = torch.tensor([2, 0, 5, 3]).cuda()
for eik in (False, True):
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=1.0, use_eikonal=eik))
    for _ in range(2):
        st = tr.step(data[ids.cpu()], ids, 3)
    torch.cuda.synchronize()
    print("step eik", eik, float(st["sdf_loss"]))
# dense autograd path + inference + grid decode
x = torch.randn(300, 259, device="cuda", requires_grad=True)
dec(x * 0.1).sum().backward()
with torch.no_grad():
    dec.eval()
    y = dec(torch.randn(200, 259, device="cuda") * 0.1)
lat = torch.randn(1, 256, device="cuda") * 0.05
sdf = dsdf_b200.create_mesh_sdf(dec, lat, 24, to_cpu=False)
# reconstruct
pos = synth.scene_samples(1, 4096, seed=3)[0]
err, latent = dsdf_b200.reconstruct(dec, 4, 256, [pos[:2048], pos[2048:]], 0.01, 0.1, num_samples=512, lr=5e-3, l2reg=True)
print("reconstruct", float(err))
# marching cubes
ax = torch.linspace(-1, 1, 20, device="cuda")
vol = (ax[:, None, None] ** 2 + ax[None, :, None] ** 2 + ax[None, None, :] ** 2).sqrt() - 0.6
v, f = dsdf_b200.marching_cubes(vol, 0.0, (0.1, 0.1, 0.1))
print("mc", v.shape, f.shape)
# sample generation + resident sampler
vv, ff = icosphere(2, 0.5)
views = P.look_at_views(P.equidist_points_on_sphere(100, 1.1)[:2])
pts, nrm = P.visible_surface_points(torch.from_numpy(vv).cuda(), torch.from_numpy(ff).cuda(), views)
xyz, sd = P.sample_sdf_near_surface(torch.from_numpy(vv).cuda(), torch.from_numpy(ff).cuda(), pts[::7].contiguous(),
                                    nrm[::7].contiguous(), num_sample=2000, seed=1)
print("samples", xyz.shape, int(torch.isnan(sd).sum()))
store = dsdf_b200.GpuSDFSamples([(torch.rand(700, 4), -torch.rand(600, 4)) for _ in range(3)], shuffle=False)
print("batch", store.batch([2, 0], 256).shape)
torch.cuda.synchronize()
print("memcheck smoke done")

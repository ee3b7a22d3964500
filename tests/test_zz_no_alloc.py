"""dsdf_rasterize_views_static: the allocation-free variant of dsdf_rasterize_views (depth buffer in a file-scope
static __device__ array of the library, include/dsdf.h).  It has no counterpart in the reference; it was added
for callers that cannot call cudaMalloc.  Called through the C ABI and compared with the
existing entry point (caller-provided depth buffer) bit for bit -- the same tolerance test_gpu_preprocess.py uses for
the rendered images -- on two calls in a row whose meshes and view counts differ, so stale depths from the first call
would show in the second.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from common import icosphere
from dsdf_b200 import _lib, preprocess as P

pytestmark = pytest.mark.gpu

W = H = 400
MAX_VIEWS = 8                   # DSDF_RAST_STATIC_MAX_VIEWS in include/dsdf.h


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _outputs(nv):
    ov = torch.full((nv, W, H, 4), float("nan"), device="cuda")
    on = torch.full((nv, W, H, 4), float("nan"), device="cuda")
    ok = torch.full((nv, W, H), 77, dtype=torch.uint8, device="cuda")
    return ov, on, ok


def _mesh(level, radius, centre, view_ids):
    v, f = icosphere(level, radius, centre)
    views = P.look_at_views(P.equidist_points_on_sphere(100, 1.1)[view_ids])
    return (torch.from_numpy(v.astype(np.float32)).cuda(), torch.from_numpy(f).to(torch.int32).cuda(),
            torch.from_numpy(np.ascontiguousarray(views, dtype=np.float32)).cuda())


def _with_caller_buffer(L, v, f, vw):
    nv = vw.shape[0]
    zbuf = torch.empty(nv * W * H, dtype=torch.int64, device="cuda")
    ov, on, ok = _outputs(nv)
    _lib.check(L.dsdf_rasterize_views(v.data_ptr(), v.shape[0], f.data_ptr(), f.shape[0], vw.data_ptr(), nv,
                                      zbuf.data_ptr(), ov.data_ptr(), on.data_ptr(), ok.data_ptr(), _stream()), None, L)
    return ov, on, ok


def _static(L, v, f, vw, out):
    ov, on, ok = out
    return L.dsdf_rasterize_views_static(v.data_ptr(), v.shape[0], f.data_ptr(), f.shape[0], vw.data_ptr(), vw.shape[0],
                                         ov.data_ptr(), on.data_ptr(), ok.data_ptr(), _stream())


def _same(a, b):
    # bit-for-bit, NaN-safe (background pixels hold whatever the kernel writes there; both paths run the same kernels)
    return torch.equal(a.view(torch.int32) if a.dtype == torch.float32 else a,
                       b.view(torch.int32) if b.dtype == torch.float32 else b)


def test_static_scratch_rasteriser_equals_the_caller_buffer_one_on_two_calls_in_a_row():
    L = _lib.lib()
    # call 1: a big sphere seen from 5 views; call 2: a small off-centre sphere from 3 other views -- every pixel the
    # first call covered and the second does not must come back as background
    first = _mesh(2, 0.8, (0.0, 0.0, 0.0), [0, 17, 42, 73, 99])
    second = _mesh(1, 0.3, (0.2, -0.1, 0.15), [5, 60, 88])
    want1 = _with_caller_buffer(L, *first)
    want2 = _with_caller_buffer(L, *second)
    out1, out2 = _outputs(5), _outputs(3)
    torch.cuda.synchronize()
    _lib.check(_static(L, *first, out1), None, L)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    _lib.check(_static(L, *second, out2), None, L)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    for got, want in ((out1, want1), (out2, want2)):
        assert torch.equal(got[2], want[2])
        assert _same(got[0], want[0])
        assert _same(got[1], want[1])
    assert int(want1[2].sum()) > int(want2[2].sum()) > 0              # the two calls really differ, both see the mesh
    assert free_before - free_after < (1 << 20)                       # no device allocation inside the call
    # the same call again gives the same answer (the array keeps the last call's depths; they must be cleared)
    out3 = _outputs(5)
    _lib.check(_static(L, *first, out3), None, L)
    torch.cuda.synchronize()
    assert torch.equal(out3[2], want1[2]) and _same(out3[0], want1[0]) and _same(out3[1], want1[1])


def test_static_scratch_rasteriser_rejects_more_views_than_the_array_holds():
    L = _lib.lib()
    v, f, vw = _mesh(0, 0.5, (0.0, 0.0, 0.0), list(range(MAX_VIEWS + 1)))
    out = _outputs(MAX_VIEWS + 1)
    rc = _static(L, v, f, vw, out)
    torch.cuda.synchronize()
    assert rc != 0
    msg = L.dsdf_last_error(None).decode()
    assert "exceeds the static depth buffer" in msg and str(MAX_VIEWS) in msg
    assert int(out[2].min()) == 77 and int(out[2].max()) == 77        # nothing was launched
    # exactly the maximum is accepted
    out = _outputs(MAX_VIEWS)
    _lib.check(_static(L, v, f, vw[:MAX_VIEWS].contiguous(), out), None, L)
    want = _with_caller_buffer(L, v, f, vw[:MAX_VIEWS].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(out[2], want[2]) and _same(out[0], want[0]) and _same(out[1], want[1])

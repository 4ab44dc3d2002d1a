"""GPU test of gp_raster_backward_static, the backward entry point for callers that cannot allocate device memory
(real-time render loops where cudaMalloc is forbidden, plugins inside an engine that owns the allocator, bindings
without a device allocator).  Its per-Gaussian gradient scratch is a fixed 12 MiB array inside the library instead of a
caller workspace.

Called through the C ABI (ctypes) and compared with gp_raster_backward on the same frame, with the gradient
tolerance used everywhere else in this suite (1e-3 relative per tensor; the accumulation order of the atomics is
not deterministic, so the two are not bit-identical).  It runs twice in a row with different upstream gradients:
the second call must not see what the first left in the shared array.  The file name sorts last so that a run
stopped at the first failure has already executed every other test.
"""
import ctypes

import pytest
import torch

from helpers import raster_case, rel_err, settings_of

pytestmark = pytest.mark.gpu
GRAD_TOL = 1e-3


def _forward(case):
    from gaussianprediction_b200 import rasterizer as R
    c = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in case.items()}
    rs = settings_of(c)
    color, radii, depth, tidx, st = R.raster_forward(c["means3D"], c["shs"], None, c["opacities"], c["scales"],
                                                     c["rotations"], None, rs)
    return c, rs, st


def _backward(c, st, grad, static):
    """Both entry points through ctypes; the caller owns every input and output buffer."""
    from gaussianprediction_b200 import _lib
    from gaussianprediction_b200._lib import check, ptr
    L = _lib.lib()
    P, M = c["means3D"].shape[0], c["shs"].shape[1]
    f = dict(dtype=torch.float32, device="cuda")
    out = dict(means2D=torch.full((P, 3), float("nan"), **f), opacity=torch.full((P, 1), float("nan"), **f),
               means3D=torch.full((P, 3), float("nan"), **f), sh=torch.full((P, M, 3), float("nan"), **f),
               scales=torch.full((P, 3), float("nan"), **f), rotations=torch.full((P, 4), float("nan"), **f))
    head = (ctypes.byref(st.cfg), ptr(c["bg"]), ptr(c["means3D"]), ptr(c["shs"]), None, ptr(c["scales"]),
            ptr(c["rotations"]), None, ptr(c["view"]), ptr(c["proj"]), ptr(c["campos"]), ptr(st.radii), ptr(st.geom),
            ptr(st.binning), st.capacity, ptr(st.image), ptr(grad))
    tail = (ptr(out["means2D"]), None, ptr(out["opacity"]), ptr(out["means3D"]), None, ptr(out["sh"]),
            ptr(out["scales"]), ptr(out["rotations"]), _lib.stream_ptr())
    if static:
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        check(L.gp_raster_backward_static(*head, *tail), "gp_raster_backward_static")
        torch.cuda.synchronize()
        # the property the entry point exists for: the device's free memory is the same before and after the call
        # (cudaMemGetInfo; the first call on a device has already resolved the scratch address in the warm-up below)
        if getattr(_backward, "warmed", False):
            assert torch.cuda.mem_get_info()[0] == free0, "gp_raster_backward_static changed the device's free memory"
        _backward.warmed = True
    else:
        ws = torch.empty((L.gp_raster_backward_ws_bytes(P),), dtype=torch.uint8, device="cuda")
        check(L.gp_raster_backward(*head, ptr(ws), *tail), "gp_raster_backward")
    torch.cuda.synchronize()
    return out


def test_static_backward_matches_workspace_backward_twice(gplib):
    case = raster_case(N=3000, W=160, H=112, seed=3)
    c, rs, st = _forward(case)
    gen = torch.Generator().manual_seed(31)
    for call in range(2):      # two calls in a row share the array: stale sums from call 0 must not leak into call 1
        grad = (torch.randn(3, c["H"], c["W"], generator=gen) * (1e-2 if call == 0 else 3e-2)).cuda().contiguous()
        ref = _backward(c, st, grad, static=False)
        got = _backward(c, st, grad, static=True)
        for k in ref:
            assert torch.isfinite(got[k]).all(), (call, k)
            assert rel_err(got[k], ref[k]) <= GRAD_TOL, (call, k, rel_err(got[k], ref[k]))


def test_static_backward_at_its_limit(gplib):
    """Exactly gp_raster_backward_static_max_gaussians() Gaussians: the whole static array is in use."""
    from gaussianprediction_b200 import _lib
    limit = _lib.lib().gp_raster_backward_static_max_gaussians()
    case = raster_case(N=limit, W=192, H=128, seed=5)
    c, rs, st = _forward(case)
    grad = (torch.randn(3, c["H"], c["W"], generator=torch.Generator().manual_seed(33)) * 1e-2).cuda().contiguous()
    ref = _backward(c, st, grad, static=False)
    got = _backward(c, st, grad, static=True)
    for k in ref:
        assert torch.isfinite(got[k]).all(), k
        assert rel_err(got[k], ref[k]) <= GRAD_TOL, (k, rel_err(got[k], ref[k]))


def test_static_backward_rejects_sizes_beyond_its_limit(gplib):
    from gaussianprediction_b200 import _lib
    L = _lib.lib()
    limit = L.gp_raster_backward_static_max_gaussians()
    assert limit == 262144
    cfg = _lib.RasterCfg(limit + 1, 64, 64, 3, 16, 0.5, 0.5, 1.0, 0)
    rc = L.gp_raster_backward_static(ctypes.byref(cfg), *([None] * 13), 0, *([None] * 11))
    assert rc == -4 and b"limit" in L.gp_error_string(rc)       # GP_ERR_LIMIT, reported before anything is launched

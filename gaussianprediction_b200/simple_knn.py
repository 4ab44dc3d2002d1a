"""Drop-in for ``simple_knn._C.distCUDA2`` (submodules/simple-knn/ext.cpp:15-17, spatial.cu:15-26)."""
import torch

from . import _lib
from ._lib import check, ptr


@torch.no_grad()
def distCUDA2(points):
    """Mean of the 3 smallest squared distances from each point to the others; float32 [P]."""
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 needs a CUDA tensor (there is no CPU path)")
    pts = points.float().contiguous()
    P = int(pts.shape[0])
    out = torch.zeros((P,), dtype=torch.float32, device=pts.device)
    if P == 0:
        return out
    L = _lib.lib()
    ws = torch.empty((L.gp_dist2_ws_bytes(P),), dtype=torch.uint8, device=pts.device)
    check(L.gp_dist2_knn3(P, ptr(pts), ptr(out), ptr(ws), _lib.stream_ptr()), "gp_dist2_knn3")
    return out

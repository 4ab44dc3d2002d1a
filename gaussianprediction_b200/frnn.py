"""Drop-in for ``frnn.frnn_grid_points`` (submodules/FRNN/frnn/frnn.py:210-347) on the B200 kNN kernel.

Same signature, return tuple ``(dists, idxs, nn, grid)``, padding (-1) and errors (``ValueError`` on batch / dimension
mismatch :283-289, ``TypeError`` for non-CUDA input :295-296).  ``grid`` is accepted and returned as ``None``: the
search keeps the reference points in shared memory, there is no grid to cache.  Ties between equal distances are
broken by the smaller index (the reference's order there is nondeterministic).
"""
from typing import Union

import torch

from . import _lib
from ._lib import check, ptr


def frnn_gather(x, idxs, lengths=None):
    """x[n, idxs[n,l,k]] with zeros where idxs < 0 (frnn.py:350-398)."""
    N, P2, D = x.shape
    K = idxs.shape[2]
    tmp = idxs.clone().detach()
    tmp[idxs < 0] = 0
    out = x[:, :, None].expand(-1, -1, K, -1).gather(1, tmp[:, :, :, None].expand(-1, -1, -1, D))
    out = out.clone()
    out[(idxs < 0)[:, :, :, None].expand(-1, -1, -1, D)] = 0.0
    return out


@torch.no_grad()
def frnn_grid_points(points1, points2, lengths1=None, lengths2=None, K: int = -1,
                     r: Union[float, torch.Tensor] = -1, grid=None, return_nn: bool = False,
                     return_sorted: bool = True, radius_cell_ratio: float = 2.0):
    if points1.shape[0] != points2.shape[0]:
        raise ValueError("points1 and points2 must have the same batch  dimension")
    if points1.shape[2] != points2.shape[2]:
        raise ValueError(f"dimension mismatch: points1 of dimension {points1.shape[2]} while points2 of dimension "
                         f"{points2.shape[2]}")
    if not points1.is_cuda or not points2.is_cuda:
        raise TypeError("for now only cuda version is supported")
    points1, points2 = points1.float().contiguous(), points2.float().contiguous()
    N, P1, D = points1.shape
    P2 = points2.shape[1]
    if isinstance(r, torch.Tensor):
        rs = [float(v) for v in r.reshape(-1).tolist()]
        rs = rs * N if len(rs) == 1 else rs
    else:
        rs = [float(r)] * N
    idxs = torch.full((N, P1, K), -1, dtype=torch.int64, device=points1.device)
    dists = torch.full((N, P1, K), -1.0, dtype=torch.float32, device=points1.device)
    L = _lib.lib()
    for n in range(N):
        n1 = P1 if lengths1 is None else int(lengths1[n])
        n2 = P2 if lengths2 is None else int(lengths2[n])
        if n1 == 0:
            continue
        check(L.gp_knn_points(n1, n2, D, int(K), rs[n], ptr(points1[n]), ptr(points2[n]), ptr(idxs[n]), ptr(dists[n]),
                              _lib.stream_ptr()), "gp_knn_points")
    nn = frnn_gather(points2, idxs, lengths2) if return_nn else None
    return dists, idxs, nn, None

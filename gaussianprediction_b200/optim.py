"""The step after the backward: fused Adam and the densification statistics.

``FusedAdam`` mirrors the part of ``torch.optim.Adam`` the reference uses (scene/gaussian_model.py:439-452:
``Adam(l, lr=0.0, eps=1e-15)`` over named parameter groups, no weight decay, no amsgrad): ``param_groups`` are plain
dicts whose ``'lr'`` the reference's ``update_learning_rate`` rewrites every iteration (:474-496), ``state[p]`` holds
``'step'``, ``'exp_avg'``, ``'exp_avg_sq'`` under the names its densification code edits (:530-600), ``step()`` and
``zero_grad(set_to_none=True)`` are what train.py:192-194 calls.  One kernel launch updates every tensor.
"""
import ctypes

import torch

from . import _lib
from ._lib import check, ptr


class FusedAdam:
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        groups = list(params)
        if groups and not isinstance(groups[0], dict):
            groups = [{"params": groups}]
        self.defaults = dict(lr=lr, betas=betas, eps=eps)
        self.param_groups = []
        for g in groups:
            g = dict(g)
            g["params"] = [g["params"]] if torch.is_tensor(g["params"]) else list(g["params"])
            for k, v in self.defaults.items():
                g.setdefault(k, v)
            self.param_groups.append(g)
        self.state = {}

    def zero_grad(self, set_to_none=True):
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()

    @torch.no_grad()
    def step(self):
        by_hyper = {}
        keep = []                                 # contiguous gradient copies must outlive the launch
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue                      # torch.optim.Adam skips parameters without a gradient
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise TypeError("FusedAdam: parameters must be contiguous float32 CUDA tensors (no CPU path)")
                st = self.state.setdefault(p, {})
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] = int(st["step"]) + 1
                grad = p.grad if (p.grad.is_contiguous() and p.grad.dtype == torch.float32) else p.grad.float().contiguous()
                keep.append(grad)
                by_hyper.setdefault((tuple(g["betas"]), float(g["eps"])), []).append(
                    _lib.AdamTensor(ptr(p), ptr(grad), ptr(st["exp_avg"]), ptr(st["exp_avg_sq"]), p.numel(),
                                    float(g["lr"]), st["step"]))
        for ((b1, b2), eps), items in by_hyper.items():
            arr = (_lib.AdamTensor * len(items))(*items)
            check(_lib.lib().gp_adam_step(len(items), arr, b1, b2, eps, _lib.stream_ptr()), "gp_adam_step")
        del keep


@torch.no_grad()
def add_densification_stats(viewspace_grad, radii, xyz_gradient_accum, denom, xyz_gradient_accum_max=None,
                            max_radii2D=None):
    """train.py:166-167 + ``GaussianModel.add_densification_stats`` (scene/gaussian_model.py:756-760) for the frame
    whose rasterizer returned ``radii`` (visibility_filter = radii > 0), in place, one launch:
    ``max_radii2D[vis] = max(max_radii2D[vis], radii[vis])``, ``g = |viewspace_grad[vis, :2]|``,
    ``xyz_gradient_accum[vis] += g``, ``denom[vis] += 1``, ``xyz_gradient_accum_max[vis] = max(., g)``."""
    P = int(radii.shape[0])
    for t in (viewspace_grad, radii, xyz_gradient_accum, denom, xyz_gradient_accum_max, max_radii2D):
        if t is not None and (not t.is_cuda or not t.is_contiguous()):
            raise TypeError("add_densification_stats: contiguous CUDA tensors required (no CPU path)")
    assert viewspace_grad.shape == (P, 3) and radii.dtype == torch.int32 and viewspace_grad.dtype == torch.float32
    check(_lib.lib().gp_densification_stats(P, ptr(viewspace_grad), ptr(radii), ptr(max_radii2D), ptr(xyz_gradient_accum),
                                            ptr(denom), ptr(xyz_gradient_accum_max), _lib.stream_ptr()),
          "gp_densification_stats")

"""Shared helpers for the parity tests (scene construction, comparisons)."""
import numpy as np
import torch

from gaussianprediction_b200 import synthetic as syn
from gaussianprediction_b200.rasterizer import GaussianRasterizationSettings


def raster_case(N=3000, W=160, H=112, seed=3, shape="dnerf", bg=(0.0, 0.0, 0.0), scale_mul=1.0, cam_kwargs=None,
                device="cpu"):
    """Inputs of one rasterizer call (activated parameters) + camera; all float32 CPU tensors unless device given."""
    gs = syn.make_gaussians(N, seed=seed, shape=shape, scale_mul=scale_mul)
    a = syn.activated(gs)
    if shape == "hyper":
        cam = syn.hyper_camera(W, H, focal=700.0 * W / 536.0)
    else:
        cam = syn.make_camera(W, H, **(cam_kwargs or {}))
    t = lambda x: x.to(device)
    case = dict(means3D=t(a.means3D), opacities=t(a.opacities), scales=t(a.scales), rotations=t(a.rotations),
                shs=t(a.shs), bg=t(torch.tensor(bg, dtype=torch.float32)), view=t(cam.world_view_transform),
                proj=t(cam.full_proj_transform), campos=t(cam.camera_center), W=W, H=H, tanfovx=cam.tanfovx,
                tanfovy=cam.tanfovy, sh_degree=3)
    return case


def settings_of(case, sh_degree=None, scale_modifier=1.0):
    return GaussianRasterizationSettings(
        image_height=case["H"], image_width=case["W"], tanfovx=case["tanfovx"], tanfovy=case["tanfovy"], bg=case["bg"],
        scale_modifier=scale_modifier, viewmatrix=case["view"], projmatrix=case["proj"],
        sh_degree=case["sh_degree"] if sh_degree is None else sh_degree, campos=case["campos"], prefiltered=False)


def to_np(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def rel_err(a, b, eps=1e-12):
    """||a-b||_inf / max(||b||_inf, eps): the per-tensor gradient metric of SURVEY.md §8c."""
    a, b = to_np(a).astype(np.float64), to_np(b).astype(np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), eps)) if a.size else 0.0


def bits(x):
    return to_np(x).view(np.uint32) if to_np(x).dtype == np.float32 else to_np(x)


def count_diff(a, b):
    a, b = to_np(a), to_np(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.dtype == np.float32:
        return int((a.view(np.uint32) != b.view(np.uint32)).sum())
    return int((a != b).sum())

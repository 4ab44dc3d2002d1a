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


# ---- the reference's own GaussianModel.forward chain (tests/golden/blend_chain.npz, see make_golden.py:make_chain) ----

CHAIN_CASES = {   # tag -> (iteration, flags) exactly as make_chain ran the reference source
    "A": (50000, dict(norm_rotation=True)),
    "A2": (50000, dict(norm_rotation=False, knn_type="3D")),
    "B": (50000, dict(norm_rotation=False, step_opacity=True, opacity_type="implicit")),
    "C": (50000, dict(norm_rotation=True, step_opacity=True, opacity_type="explicit")),
    "D": (20000, dict(norm_rotation=True)),
    "D2": (20000, dict(norm_rotation=False, step_opacity=True)),
    "E": (500, dict()),
    "F": (32000, dict(norm_rotation=True)),
    "W": (50000, dict(norm_rotation=True)),
}
CHAIN_LEAVES = ("xyz", "rotation", "scaling", "opacity", "motion_feature", "kp", "kf", "logits", "thres")


def chain_args(flags):
    from gaussianprediction_b200.deform import default_args
    a = dict(feature_amplify=0.5, w=64, d=4, feature_dim=32)
    a.update(flags)
    return default_args(**a)


def chain_state_dict(gold, out_dim):
    pre = f"sd{out_dim}_"
    return {k[len(pre):]: torch.tensor(gold[k]) for k in gold.files if k.startswith(pre)}

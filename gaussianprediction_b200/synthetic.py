"""Seeded synthetic scenes, cameras, keypoints and trajectories (SURVEY.md §8d recipes).

Everything is generated on the CPU in float32 so that every implementation (this library, the
reference CUDA build, the CPU oracle) sees identical bits.  Camera matrices follow the reference's
conventions: ``getWorld2View2`` / ``getProjectionMatrix`` (utils/graphics_utils.py:38-71) transposed as
in ``scene/cameras.py:59-62`` (row-vector convention), znear=0.01, zfar=100.
"""
import math
from types import SimpleNamespace

import numpy as np
import torch


def _gen(seed):
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def make_gaussians(N, seed=2024, shape="dnerf", sh_coeffs=16, scale_mul=1.0):
    """Canonical Gaussian parameters in the reference's storage form (scene/gaussian_model.py:348-353):
    ``_xyz[N,3]``, ``_scaling[N,3]`` (log), ``_rotation[N,4]`` (raw), ``_opacity[N,1]`` (logit),
    ``_features_dc[N,1,3]``, ``_features_rest[N,M-1,3]``, plus ``motion_feature[N,32]``."""
    g = _gen(seed)
    N = int(N)
    if shape == "dnerf":
        n_surf = int(0.7 * N)
        n_obj = 8
        centers = (torch.rand(n_obj, 3, generator=g) * 2 - 1) * 0.8
        radii = 0.2 + 0.4 * torch.rand(n_obj, 3, generator=g)
        which = torch.randint(0, n_obj, (n_surf,), generator=g)
        d = torch.randn(n_surf, 3, generator=g)
        d = d / d.norm(dim=1, keepdim=True).clamp_min(1e-8)
        surf = centers[which] + d * radii[which] * (1 + 0.02 * torch.randn(n_surf, 1, generator=g))
        vol = (torch.rand(N - n_surf, 3, generator=g) * 2 - 1) * 1.3  # reference random-init box, dataset_readers.py:259
        xyz = torch.cat([surf, vol], 0)
        extent = 2.6
    elif shape == "hyper":
        xy = (torch.rand(N, 2, generator=g) * 2 - 1) * 2.0
        z = 1.0 + 5.0 * torch.rand(N, 1, generator=g)
        xyz = torch.cat([xy, z], 1)
        extent = 4.0
    else:
        raise ValueError(shape)
    xyz = xyz[torch.randperm(N, generator=g)].contiguous()
    # isotropic spacing estimate standing in for sqrt(distCUDA2) (scene/gaussian_model.py:341-342), with anisotropy
    base = 0.12 * extent / (N ** (1.0 / 3.0)) * scale_mul
    # wide log-normal size spread (a few large splats, many small) + per-axis anisotropy: at 100k / 800x800 this gives
    # a median screen radius of ~7 px and ~7-8 tile instances per Gaussian
    log_s = math.log(base) + 0.9 * torch.randn(N, 1, generator=g) + 0.3 * torch.randn(N, 3, generator=g)
    rot = torch.randn(N, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True) * (0.5 + 1.5 * torch.rand(N, 1, generator=g))
    u = 0.02 + 0.96 * torch.rand(N, 1, generator=g)
    opacity = torch.log(u / (1 - u))
    f_dc = 0.5 * torch.randn(N, 1, 3, generator=g)
    f_rest = 0.05 * torch.randn(N, sh_coeffs - 1, 3, generator=g)
    motion_feature = (torch.rand(N, 32, generator=g) * 2 - 1) * 0.1
    return SimpleNamespace(xyz=xyz.float(), scaling=log_s.float().contiguous(), rotation=rot.float().contiguous(),
                           opacity=opacity.float(), features_dc=f_dc.float(), features_rest=f_rest.float(),
                           motion_feature=motion_feature.float(), N=N)


def _projection(znear, zfar, fovX, fovY):
    tanY, tanX = math.tan(fovY / 2), math.tan(fovX / 2)
    top, right = tanY * znear, tanX * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W=800, H=800, fovx=0.6911, azimuth=0.3, elevation=0.4, radius=4.0, look_at=(0.0, 0.0, 0.0),
                fovy=None, position=None):
    """Camera on a sphere looking at ``look_at``; returns the tensors ``render()`` puts in the raster settings
    (gaussian_renderer/__init__.py:38-50)."""
    if position is None:
        pos = np.array([radius * math.cos(elevation) * math.cos(azimuth), radius * math.cos(elevation) * math.sin(azimuth),
                        radius * math.sin(elevation)], dtype=np.float64)
    else:
        pos = np.asarray(position, dtype=np.float64)
    tgt = np.asarray(look_at, dtype=np.float64)
    fwd = tgt - pos
    fwd /= np.linalg.norm(fwd)
    up = np.array([0.0, 0.0, 1.0])
    if abs(np.dot(up, fwd)) > 0.99:
        up = np.array([0.0, 1.0, 0.0])
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    # camera-to-world with +x right, +y down, +z forward (the convention after dataset_readers.py:201)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, pos
    w2c = np.linalg.inv(c2w)
    R = np.transpose(w2c[:3, :3])
    T = w2c[:3, 3]
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    world_view = torch.tensor(np.float32(Rt)).transpose(0, 1).contiguous()
    if fovy is None:
        focal = W / (2 * math.tan(fovx / 2))
        fovy = 2 * math.atan(H / (2 * focal))
    proj = _projection(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = (world_view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = world_view.inverse()[3, :3].contiguous()
    return SimpleNamespace(image_width=W, image_height=H, FoVx=fovx, FoVy=fovy, tanfovx=math.tan(fovx * 0.5),
                           tanfovy=math.tan(fovy * 0.5), world_view_transform=world_view.float(),
                           full_proj_transform=full.float(), camera_center=center.float())


def camera_ring(n, W=800, H=800, fovx=0.6911, seed=7, radius=4.0):
    g = _gen(seed)
    az = torch.rand(n, generator=g) * 2 * math.pi
    el = (torch.rand(n, generator=g) - 0.3) * 1.0
    return [make_camera(W, H, fovx, float(a), float(e), radius) for a, e in zip(az, el)]


def hyper_camera(W=536, H=960, focal=700.0, jitter=(0.0, 0.0)):
    """Forward-facing camera for the HyperNeRF-shaped slab (looks down +z from the origin)."""
    fovx = 2 * math.atan(W / (2 * focal))
    fovy = 2 * math.atan(H / (2 * focal))
    tgt = (math.tan(jitter[0]) * 3.5, math.tan(jitter[1]) * 3.5, 3.5)
    cam = make_camera(W, H, fovx, position=(0.0, 0.0, 0.0), look_at=tgt, fovy=fovy)
    return cam


def make_keypoints(gs, Kp=512, seed=11, feature_dim=32):
    """``super_gaussians[Kp,3]`` = random subset of xyz + N(0,0.02); ``super_gaussians_feature[Kp,F]`` = U(-1,1)*0.1."""
    g = _gen(seed)
    idx = torch.randperm(gs.N, generator=g)[:Kp]
    kp = gs.xyz[idx] + 0.02 * torch.randn(Kp, 3, generator=g)
    feat = (torch.rand(Kp, feature_dim, generator=g) * 2 - 1) * 0.1
    return kp.float().contiguous(), feat.float().contiguous()


def make_logits(N, K=6, seed=13):
    """Stand-in for the tinycudann blend-weight network output [N, 2K] (parity unpinned)."""
    return torch.randn(N, 2 * K, generator=_gen(seed)).float()


def keypoint_trajectory(kp, frames=200, seed=17, amp=0.3):
    """C4 extrapolation inputs: kpts[f] = kp + amp*sin(2*pi*f/frames + phi_k); small normalised rotations."""
    g = _gen(seed)
    Kp = kp.shape[0]
    phi = torch.rand(Kp, 3, generator=g) * 2 * math.pi
    f = torch.arange(frames).float().view(-1, 1, 1)
    kpts = kp.unsqueeze(0) + amp * torch.sin(2 * math.pi * f / frames + phi.unsqueeze(0))
    axis = torch.randn(Kp, 3, generator=g)
    axis = axis / axis.norm(dim=1, keepdim=True)
    ang = 0.3 * torch.sin(2 * math.pi * f / frames + phi[:, :1].unsqueeze(0))  # [F,Kp,1]
    rot = torch.cat([torch.cos(ang / 2), torch.sin(ang / 2) * axis.unsqueeze(0)], dim=-1)
    return kpts.float().contiguous(), rot.float().contiguous()


def activated(gs, device="cpu"):
    """What ``render()`` feeds the rasterizer for the canonical scene: exp(scale), normalised rotation, sigmoid(opacity),
    SH features concatenated (scene/gaussian_model.py:138-162)."""
    shs = torch.cat([gs.features_dc, gs.features_rest], dim=1).contiguous()
    return SimpleNamespace(
        means3D=gs.xyz.to(device), scales=torch.exp(gs.scaling).to(device),
        rotations=torch.nn.functional.normalize(gs.rotation).to(device), opacities=torch.sigmoid(gs.opacity).to(device),
        shs=shs.to(device))

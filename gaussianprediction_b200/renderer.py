"""``render()`` / ``render_motion()`` with the reference's signatures (gaussian_renderer/__init__.py:18-191).

``pc`` is any object with the reference model's surface: ``get_xyz``, ``get_opacity``, ``get_scaling``,
``get_rotation``, ``get_features``, ``active_sh_degree`` and ``pc(time, it)`` -- e.g. ``deform.KeypointDeformation``
wrapped by ``SceneModel`` below, or the reference's own ``GaussianModel`` once the drop-in modules are installed.
"""
import math

import torch

from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer


def _settings(viewpoint_camera, pc, bg_color, scaling_modifier):
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    return GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False)


def _sh_inputs(pc, override_color):
    """(shs, shs_rest, colors_precomp) for the rasterizer.  The reference passes ``pc.get_features`` = cat(_features_dc,
    _features_rest) (gaussian_renderer/__init__.py:74-78); when the model exposes the two parameters in the usual
    [N,1,3] / [N,15,3] layout they are handed over separately and read in place (no 192 B/Gaussian copy per frame, and
    no split of the gradient on the way back)."""
    if override_color is not None:
        return None, None, override_color
    dc, rest = getattr(pc, "_features_dc", None), getattr(pc, "_features_rest", None)
    if torch.is_tensor(dc) and torch.is_tensor(rest) and dc.is_cuda and dc.dim() == 3 and dc.shape[1] == 1 \
            and rest.dim() == 3 and rest.shape[1] == 15 and dc.dtype == torch.float32 and rest.dtype == torch.float32 \
            and dc.is_contiguous() and rest.is_contiguous():
        return dc, rest, None
    return pc.get_features, None, None


def render(viewpoint_camera, pc, pipe=None, bg_color=None, scaling_modifier=1.0, override_color=None, delta=None,
           time=None, it=1, capacity=None, deformed=None):
    """``deformed`` (extra, keyword only): the result of an earlier ``pc(time, it)`` call to use instead of calling it
    here -- lets a captured step put the deformation forward and the rasterizer into different CUDA graphs."""
    # (the reference adds 0 to get a non-leaf tensor and retains its grad, gaussian_renderer/__init__.py; a leaf receives
    # .grad just the same and saves a 24 MB read-modify-write per frame at 2 M Gaussians)
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda")
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    rasterizer = GaussianRasterizer(_settings(viewpoint_camera, pc, bg_color, scaling_modifier), capacity=capacity)
    if time is None:
        means3D = pc.get_xyz + delta if delta is not None else pc.get_xyz
        scales, rotations, opacity = pc.get_scaling, pc.get_rotation, pc.get_opacity
    elif deformed is not None:
        means3D, rotations, scales, opacity = deformed
    else:
        means3D, rotations, scales, opacity = pc(time, it)
    shs, shs_rest, colors_precomp = _sh_inputs(pc, override_color)
    rendered_image, radii, depth, tidx = rasterizer(means3D=means3D, means2D=screenspace_points, shs=shs,
                                                    colors_precomp=colors_precomp, opacities=opacity, scales=scales,
                                                    rotations=rotations, cov3D_precomp=None, shs_rest=shs_rest)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii, "depth": depth, "tidx": tidx,
            # extra key (not in the reference's dict): the deformed parameters the rasterizer consumed, so that a
            # caller can run the backward in two stages (pipeline.TrainStep overlaps the gradient exchange with it)
            "deformed": (means3D, rotations, scales, opacity)}


def render_motion(viewpoint_camera, pc, pipe=None, bg_color=None, scaling_modifier=1.0, override_color=None, xyz_t=None,
                  r_t=None, opacity=None, capacity=None):
    # (the reference adds 0 to get a non-leaf tensor and retains its grad, gaussian_renderer/__init__.py; a leaf receives
    # .grad just the same and saves a 24 MB read-modify-write per frame at 2 M Gaussians)
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda")
    rasterizer = GaussianRasterizer(_settings(viewpoint_camera, pc, bg_color, scaling_modifier), capacity=capacity)
    opacity = pc.get_opacity if opacity is None else opacity
    shs, shs_rest, colors_precomp = _sh_inputs(pc, override_color)
    rendered_image, radii, depth, tidx = rasterizer(means3D=xyz_t, means2D=screenspace_points, shs=shs,
                                                    colors_precomp=colors_precomp, opacities=opacity,
                                                    scales=pc.get_scaling, rotations=r_t, cov3D_precomp=None,
                                                    shs_rest=shs_rest)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii}


class SceneModel(torch.nn.Module):
    """KeypointDeformation + SH features: the subset of ``GaussianModel`` that ``render()`` touches."""

    def __init__(self, deformation, features_dc, features_rest, active_sh_degree=3):
        super().__init__()
        self.deformation = deformation
        self._features_dc = torch.nn.Parameter(features_dc)
        self._features_rest = torch.nn.Parameter(features_rest)
        self.active_sh_degree = active_sh_degree
        self.max_sh_degree = active_sh_degree

    def forward(self, t, iteration, return_weights=False):
        return self.deformation(t, iteration, return_weights)

    get_xyz = property(lambda self: self.deformation.get_xyz)
    get_opacity = property(lambda self: self.deformation.get_opacity)
    get_scaling = property(lambda self: self.deformation.get_scaling)
    get_rotation = property(lambda self: self.deformation.get_rotation)

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)  # scene/gaussian_model.py:155-158

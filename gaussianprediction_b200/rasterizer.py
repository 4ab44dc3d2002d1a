"""Drop-in for the reference's ``diff_gaussian_rasterization`` package, backed by libgp_b200.so.

Mirrors submodules/diff-gaussian-rasterization-w-depth/diff_gaussian_rasterization/__init__.py:
``GaussianRasterizationSettings`` (:134-145), ``GaussianRasterizer`` (:147-196), the autograd
function ``_RasterizeGaussians`` (:40-132) and ``rasterize_gaussians`` (:16-38): same names, argument
order, outputs ``(color[3,H,W], radii[P] int32, depth[1,H,W], tidx[1,H,W] int32)``, gradient order and
the same ``Exception`` when not exactly one of SHs/colours or scale+rotation/covariance is given.

Two ways to size the tile-instance buffers:
  * default ("exact"): one host read of the instance count R per frame, like the reference's own
    ``cudaMemcpy`` of ``num_rendered`` (rasterizer_impl.cu:283-284);
  * ``capacity=...`` on ``rasterize_gaussians``/``GaussianRasterizer``: no host round trip at all; if R
    exceeds the capacity the frame is not rendered and ``last_overflow()`` reports it (used by the
    frame pipelines in ``pipeline.py``, which own the retry policy).
"""
import ctypes
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _lib
from ._sinks import acc_sink_for, out_like, sink_for
from ._lib import RasterCfg, check, ptr


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool


def _f32c(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _opt(t):
    """Empty tensors stand for 'not given' in the reference API (:173-183)."""
    return None if t is None or t.numel() == 0 else _f32c(t)


def workspace_sizes(P, W, H, capacity):
    g, b, i = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
    check(_lib.lib().gp_raster_sizes(P, W, H, int(capacity), ctypes.byref(g), ctypes.byref(b), ctypes.byref(i)),
          "gp_raster_sizes")
    return g.value, b.value, i.value


def _make_cfg(P, raster_settings, sh, sh_rest=None):
    M = 0 if sh is None else int(sh.shape[1]) + (0 if sh_rest is None else int(sh_rest.shape[1]))
    return RasterCfg(P, int(raster_settings.image_width), int(raster_settings.image_height),
                     int(raster_settings.sh_degree), M, float(raster_settings.tanfovx), float(raster_settings.tanfovy),
                     float(raster_settings.scale_modifier), 1 if raster_settings.prefiltered else 0,
                     ptr(sh_rest), None)


class RasterState:
    """What the backward needs from a forward (the reference keeps geom/binning/img byte tensors, :84)."""
    __slots__ = ("cfg", "geom", "binning", "image", "radii", "capacity", "num_rendered")


LAST_STATE = None   # state of the most recent forward (diagnostics: header words, overflow flag)


def last_overflow():
    """(num_rendered, overflowed) of the most recent forward, read from the device header (synchronises)."""
    st = LAST_STATE
    if st is None or st.cfg is None:
        return 0, False
    hdr = st.geom[:16].view(torch.int32).cpu()
    return int(hdr[_lib.HDR_NUM_RENDERED]), bool(int(hdr[_lib.HDR_OVERFLOW]))


def raster_forward(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings,
                   capacity: Optional[int] = None, sh_rest=None):
    """Runs the two forward stages; returns (color, radii, depth, tidx, RasterState).  With ``sh_rest`` ([P,15,3]),
    ``sh`` is only coefficient 0 ([P,1,3]): the two parameters of the reference model, read in place."""
    L = _lib.lib()
    dev = means3D.device
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # reference rasterize_points.cu:56-58
    if not means3D.is_cuda:
        raise RuntimeError("gaussianprediction_b200 rasterizer needs CUDA tensors (there is no CPU path)")
    P = int(means3D.shape[0])
    H, W = int(raster_settings.image_height), int(raster_settings.image_width)
    color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    tidx = torch.empty((1, H, W), dtype=torch.int32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    st = RasterState()
    if P == 0:  # reference returns its zero-filled outputs untouched (rasterize_points.cu:67-82)
        color.zero_(); depth.zero_(); tidx.zero_()
        st.cfg = None
        st.num_rendered = 0
        return color, radii, depth, tidx, st

    means3D = _f32c(means3D)
    sh, colors_precomp, sh_rest = _opt(sh), _opt(colors_precomp), _opt(sh_rest)
    scales, rotations, cov3Ds_precomp = _opt(scales), _opt(rotations), _opt(cov3Ds_precomp)
    opacities = _f32c(opacities)
    bg = _f32c(raster_settings.bg)
    view, proj, campos = _f32c(raster_settings.viewmatrix), _f32c(raster_settings.projmatrix), _f32c(raster_settings.campos)
    cfg = _make_cfg(P, raster_settings, sh, sh_rest)
    gbytes, _, ibytes = workspace_sizes(P, W, H, 0)
    geom = torch.empty((gbytes,), dtype=torch.uint8, device=dev)
    image = torch.empty((ibytes,), dtype=torch.uint8, device=dev)
    stream = _lib.stream_ptr()
    check(L.gp_raster_preprocess(ctypes.byref(cfg), ptr(means3D), ptr(sh), ptr(colors_precomp), ptr(opacities),
                                 ptr(scales), ptr(rotations), ptr(cov3Ds_precomp), ptr(view), ptr(proj), ptr(campos),
                                 ptr(geom), ptr(radii), stream), "gp_raster_preprocess")
    if capacity is None:
        n = ctypes.c_int(0)
        check(L.gp_raster_num_rendered(ptr(geom), stream, ctypes.byref(n)), "gp_raster_num_rendered")
        st.num_rendered = int(n.value)
        capacity = max(int(n.value), 1)
    else:
        st.num_rendered = None  # still on the device: geom header word 0
    _, bbytes, _ = workspace_sizes(P, W, H, capacity)
    binning = torch.empty((bbytes,), dtype=torch.uint8, device=dev)
    check(L.gp_raster_render(ctypes.byref(cfg), ptr(bg), ptr(colors_precomp), ptr(geom), ptr(binning), int(capacity),
                             ptr(image), ptr(radii), ptr(color), ptr(depth), ptr(tidx), stream), "gp_raster_render")
    st.cfg, st.geom, st.binning, st.image, st.radii, st.capacity = cfg, geom, binning, image, radii, int(capacity)
    global LAST_STATE
    LAST_STATE = st
    return color, radii, depth, tidx, st


def raster_backward(st, grad_out_color, means3D, sh, colors_precomp, scales, rotations, cov3Ds_precomp,
                    raster_settings, sh_rest=None, sinks=(None, None), sh_accumulate=False):
    """Returns (grad_means2D, grad_colors, grad_opacity, grad_means3D, grad_cov3D, grad_sh, grad_scales, grad_rots);
    with split SH storage grad_sh is the pair (grad_sh_dc [P,1,3], grad_sh_rest [P,15,3])."""
    L = _lib.lib()
    dev = means3D.device
    P = int(means3D.shape[0])
    sh, colors_precomp, sh_rest = _opt(sh), _opt(colors_precomp), _opt(sh_rest)
    scales, rotations, cov3Ds_precomp = _opt(scales), _opt(rotations), _opt(cov3Ds_precomp)
    M = 0 if sh is None else int(sh.shape[1])
    f = dict(dtype=torch.float32, device=dev)
    if P == 0:
        z = lambda *s: torch.zeros(s, **f)
        g_sh0 = z(0, M, 3) if sh_rest is None else (z(0, M, 3), z(0, int(sh_rest.shape[1]), 3))
        return z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 6), g_sh0, z(0, 3), z(0, 4)
    g_means2D = torch.empty((P, 3), **f)
    g_opacity = torch.empty((P, 1), **f)
    g_means3D = torch.empty((P, 3), **f)
    g_colors = torch.empty((P, 3), **f) if colors_precomp is not None else None
    g_sh = out_like(sinks[0], (P, M, 3), **f) if sh is not None else None
    g_scales = torch.empty((P, 3), **f) if scales is not None else None
    g_rots = torch.empty((P, 4), **f) if scales is not None else None
    g_cov3D = torch.empty((P, 6), **f) if cov3Ds_precomp is not None else None
    ws = torch.empty((L.gp_raster_backward_ws_bytes(P),), dtype=torch.uint8, device=dev)
    cfg = st.cfg
    g_sh_rest = None
    if sh_rest is not None:
        g_sh_rest = out_like(sinks[1], (P, int(sh_rest.shape[1]), 3), **f)
        cfg = RasterCfg.from_buffer_copy(st.cfg)
        cfg.sh_rest, cfg.dL_dsh_rest = ptr(sh_rest), ptr(g_sh_rest)
    if sh_accumulate:      # ``sinks`` hold a running sum: the kernel adds the SH gradients to it (gp_b200.h)
        cfg = RasterCfg.from_buffer_copy(cfg)
        cfg.dL_dsh_accumulate = 1
    bg = _f32c(raster_settings.bg)
    view, proj, campos = _f32c(raster_settings.viewmatrix), _f32c(raster_settings.projmatrix), _f32c(raster_settings.campos)
    means3D, grad_out_color = _f32c(means3D), _f32c(grad_out_color)   # named: copies must outlive the launch
    check(L.gp_raster_backward(ctypes.byref(cfg), ptr(bg), ptr(means3D), ptr(sh), ptr(colors_precomp),
                               ptr(scales), ptr(rotations), ptr(cov3Ds_precomp), ptr(view), ptr(proj), ptr(campos),
                               ptr(st.radii), ptr(st.geom), ptr(st.binning), st.capacity, ptr(st.image),
                               ptr(grad_out_color), ptr(ws), ptr(g_means2D), ptr(g_colors), ptr(g_opacity),
                               ptr(g_means3D), ptr(g_cov3D), ptr(g_sh), ptr(g_scales), ptr(g_rots),
                               _lib.stream_ptr()), "gp_raster_backward")
    if sh_rest is not None:
        g_sh = (g_sh, g_sh_rest)
    return g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rots


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, capacity=None, sh_rest=None):
        color, radii, depth, tidx, st = raster_forward(means3D, sh, colors_precomp, opacities, scales, rotations,
                                                       cov3Ds_precomp, raster_settings, capacity, sh_rest)
        ctx.raster_settings = raster_settings
        ctx.state = st
        ctx.num_rendered = st.num_rendered
        ctx.split_sh = sh_rest is not None
        ctx.sinks = (sink_for(sh), sink_for(sh_rest))      # where the SH gradients go, if the caller registered a place
        # running sums the SH gradients are added to (gradient accumulation): needs the staged kernel path, i.e. 16
        # coefficients in 16-byte aligned arrays, and a sum for every SH array in use
        acc = (acc_sink_for(sh), acc_sink_for(sh_rest))
        arrays = [(t, a) for t, a in zip((sh, sh_rest), acc) if t is not None and t.numel()]
        M = sum(int(t.shape[1]) for t, _ in arrays)
        ctx.sh_accumulate = bool(arrays) and M == 16 and all(
            a is not None and t.data_ptr() % 16 == 0 and a.data_ptr() % 16 == 0 for t, a in arrays)
        if ctx.sh_accumulate:
            ctx.sinks = acc
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, sh, sh_rest)
        ctx.mark_non_differentiable(radii, depth, tidx)  # the reference has no depth gradient (:88)
        ctx.set_materialize_grads(False)                 # no zero tensors built for the outputs that take no gradient
        return color, radii, depth, tidx

    @staticmethod
    def backward(ctx, grad_out_color, _radii, _depth, _tidx):
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, sh, sh_rest = ctx.saved_tensors
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, int(ctx.raster_settings.image_height),
                                          int(ctx.raster_settings.image_width)), dtype=torch.float32, device=means3D.device)
        (g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rots) = raster_backward(
            ctx.state, grad_out_color, means3D, sh, colors_precomp, scales, rotations, cov3Ds_precomp,
            ctx.raster_settings, sh_rest if ctx.split_sh else None, ctx.sinks, ctx.sh_accumulate)
        g_sh, g_sh_rest = g_sh if ctx.split_sh else (g_sh, None)
        # order of the reference's backward (:120-130)
        return (g_means3D, g_means2D, g_sh, g_colors, g_opacity, g_scales, g_rots, g_cov3D, None, None, g_sh_rest)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, capacity=None, sh_rest=None):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings, capacity, sh_rest)


def mark_visible(positions, viewmatrix, projmatrix):
    """Stand-in for ``_C.mark_visible`` (rasterize_points.cu:199-218): bool[P], view-space z > 0.2."""
    P = int(positions.shape[0])
    out = torch.zeros((P,), dtype=torch.uint8, device=positions.device)
    if P:
        positions, viewmatrix, projmatrix = _f32c(positions), _f32c(viewmatrix), _f32c(projmatrix)
        check(_lib.lib().gp_mark_visible(P, ptr(positions), ptr(viewmatrix), ptr(projmatrix),
                                         ptr(out), _lib.stream_ptr()), "gp_mark_visible")
    return out.bool()


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings, capacity=None):
        super().__init__()
        self.raster_settings = raster_settings
        self.capacity = capacity

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, shs_rest=None):
        """The reference's call (diff_gaussian_rasterization/__init__.py:186-221).  Extension: with ``shs_rest``
        ([P,15,3], CUDA float32) ``shs`` is coefficient 0 only ([P,1,3]) -- the model's ``_features_dc`` /
        ``_features_rest`` read in place, without the per-frame ``torch.cat`` of ``get_features``."""
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        if shs_rest is not None and (shs.dim() != 3 or shs.shape[1] != 1 or shs_rest.shape[1] != 15):
            shs, shs_rest = torch.cat((shs, shs_rest), dim=1), None      # only the 1 + 15 layout is read in place
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   raster_settings, self.capacity, shs_rest)


def export_state(st):
    """Test/introspection: the internal forward state in the reference's array shapes (dict of torch tensors)."""
    L = _lib.lib()
    cfg = st.cfg
    P, W, H = cfg.P, cfg.width, cfg.height
    dev = st.geom.device
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    f = dict(dtype=torch.float32, device=dev)
    o = dict(depths=torch.zeros(P, **f), means2D=torch.zeros(P, 2, **f), cov3D=torch.zeros(P, 6, **f),
             conic_opacity=torch.zeros(P, 4, **f), rgb=torch.zeros(P, 3, **f),
             tiles_touched=torch.zeros(P, dtype=torch.int32, device=dev),
             clamped=torch.zeros(P, 3, dtype=torch.uint8, device=dev),
             ranges=torch.zeros(tiles, 2, dtype=torch.int32, device=dev), final_T=torch.zeros(H * W, **f),
             n_contrib=torch.zeros(H * W, dtype=torch.int32, device=dev))
    stream = _lib.stream_ptr()
    check(L.gp_raster_export_state(ctypes.byref(cfg), ptr(st.geom), ptr(st.image), ptr(o["depths"]), ptr(o["means2D"]),
                                   ptr(o["cov3D"]), ptr(o["conic_opacity"]), ptr(o["rgb"]), ptr(o["tiles_touched"]),
                                   ptr(o["clamped"]), ptr(o["ranges"]), ptr(o["final_T"]), ptr(o["n_contrib"]), stream),
          "gp_raster_export_state")
    hdr = st.geom[:16].view(torch.int32).cpu()
    R = int(hdr[_lib.HDR_NUM_SORTED])
    o["num_rendered"] = int(hdr[_lib.HDR_NUM_RENDERED])
    o["overflow"] = int(hdr[_lib.HDR_OVERFLOW])
    o["point_list"] = torch.zeros(max(R, 1), dtype=torch.int32, device=dev)
    o["point_list_keys"] = torch.zeros(max(R, 1), dtype=torch.int64, device=dev)
    check(L.gp_raster_export_binning(ctypes.byref(cfg), ptr(st.geom), ptr(st.binning), st.capacity, ptr(o["point_list"]),
                                     ptr(o["point_list_keys"]), stream), "gp_raster_export_binning")
    o["point_list"], o["point_list_keys"] = o["point_list"][:R], o["point_list_keys"][:R]
    return o

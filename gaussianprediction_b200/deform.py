"""Keypoint-driven deformation: the ``GaussianModel.forward`` call of the reference, on B200 kernels.

Mirrors scene/gaussian_model.py:231-304 (control flow), :110-125 (``get_nearest_mask``), :214-229
(``fill_nearest``), :180-189 (``get_motion_delta``), :314-315 (``get_rotation_``) and
scene/deformable_field.py:63-127 (``positional_encoding``, ``Deformable_Field``).  The K-nearest-keypoint
search and the blend (softmax, gather, weighted sums, quaternion compose, activations) run in
libgp_b200.so, and so does the small keypoint MLP: north_star leaves it a cuBLAS/torch call "unless ncu shows it
matters", and as ~70 tiny launches it was 0.26 ms of a 1.18 ms iteration (profiles/r1_keypoint_mlp.md).

The hash-grid blend-weight network (tinycudann, scene/gaussian_model.py:370-392) is not available
(un-vendored third-party code): ``weights_model`` is any callable [N,3] -> [N,2K];
the default is a learnable logits table.
"""
import ctypes
from types import SimpleNamespace

import torch
import torch.nn as nn

from . import _lib
from ._lib import check, ptr
from ._sinks import acc_sink_for, out_like, sink_for


# ------------------------------------------------------------------------------------------------ small torch pieces

_FREQ_BANDS = {}


def _freq_bands(freqs, device):
    # cached per device: building it on the CPU and copying it over is a pageable H2D copy, which is slow in the
    # frame loop and illegal while a CUDA graph is being captured
    key = (int(freqs), str(device))
    fb = _FREQ_BANDS.get(key)
    if fb is None:
        fb = (2 ** torch.arange(freqs).float()).to(device)
        _FREQ_BANDS[key] = fb
    return fb


def positional_encoding(positions, freqs):
    """[sin(2^k x), cos(2^k x)] interleaved per input dim (scene/deformable_field.py:63-72, ori=False)."""
    freq_bands = _freq_bands(freqs, positions.device)
    pts = (positions[..., None] * freq_bands).reshape(positions.shape[:-1] + (freqs * positions.shape[-1],))
    pts = torch.stack([torch.sin(pts), torch.cos(pts)], dim=-1).reshape(pts.shape[:-1] + (pts.shape[-1] * 2,))
    return pts


class DeformableField(nn.Module):
    """d x (Linear + ReLU) then Linear: the non-split ``Deformable_Field`` (scene/deformable_field.py:74-127).
    Sub-module names match the reference so a reference ``state_dict`` loads unchanged."""

    def __init__(self, input_dim, output_dim=7, d=4, w=256):
        super().__init__()
        layers = []
        for i in range(d):
            layers.append(nn.Linear(input_dim if i == 0 else w, w))
            layers.append(nn.ReLU())
        self.mlp = nn.Sequential(*layers)
        self.feature_to_deformation = nn.Sequential(nn.Linear(w, output_dim))

    def forward(self, x):
        return self.feature_to_deformation(self.mlp(x))

    def linears(self):
        """The Linear layers in order, or None if the module is not the plain Linear/ReLU stack."""
        mods = list(self.mlp) + list(self.feature_to_deformation)
        lin = [m for m in mods if isinstance(m, nn.Linear)]
        if any(not isinstance(m, (nn.Linear, nn.ReLU)) for m in mods) or len(self.feature_to_deformation) != 1:
            return None
        return lin


def quat_mul(q1, q2):
    """(w,x,y,z) product with the operand order of utils/camera_utils.py:158-170 (torch; used off the hot path)."""
    w = q2[..., 0] * q1[..., 0] - q2[..., 1] * q1[..., 1] - q2[..., 2] * q1[..., 2] - q2[..., 3] * q1[..., 3]
    x = q2[..., 1] * q1[..., 0] + q2[..., 0] * q1[..., 1] + q2[..., 3] * q1[..., 2] - q2[..., 2] * q1[..., 3]
    y = q2[..., 2] * q1[..., 0] - q2[..., 3] * q1[..., 1] + q2[..., 0] * q1[..., 2] + q2[..., 1] * q1[..., 3]
    z = q2[..., 3] * q1[..., 0] + q2[..., 2] * q1[..., 1] - q2[..., 1] * q1[..., 2] + q2[..., 0] * q1[..., 3]
    return torch.stack([w, x, y, z], dim=-1)


# ------------------------------------------------------------------------------------------------ kernels

def _c(t):
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise TypeError("for now only cuda version is supported")  # same error as frnn.py:295-296


@torch.no_grad()
def spatial_order(xyz):
    """Morton-order permutation of the points (int32 [N]): an optional speed hint for ``knn_keypoints``."""
    _need_cuda(xyz)
    N = int(xyz.shape[0])
    order = torch.empty((N,), dtype=torch.int32, device=xyz.device)
    if N:
        L = _lib.lib()
        ws = torch.empty((L.gp_spatial_order_ws_bytes(N),), dtype=torch.uint8, device=xyz.device)
        xyz_c = _c(xyz)   # named: a contiguous copy must outlive the launch
        check(L.gp_spatial_order(N, ptr(xyz_c), ptr(order), ptr(ws), _lib.stream_ptr()), "gp_spatial_order")
    return order


@torch.no_grad()
def knn_keypoints(xyz, motion_feature, kp_xyz, kp_feature, K, knn_type="hybird", feature_amplify=5.0,
                  return_dists=False, idx32_only=False, order=None, hint=None):
    """``get_nearest_mask`` (scene/gaussian_model.py:110-125) without the torch.cat and the FRNN grid set-up.
    Returns (idx64 [N,K] or None, idx32 [N,K], dists or None).  ``order``: optional ``spatial_order(xyz)``;
    ``hint``: optional int32 [N,K] neighbours of a previous call (speed only; it is overwritten with the new result
    when given, so the same buffer serves as hint and output across frames)."""
    _need_cuda(xyz, kp_xyz)
    N, Kp = int(xyz.shape[0]), int(kp_xyz.shape[0])
    dev = xyz.device
    if knn_type == "3D":
        F, r, mf, kf = 0, 1e3, None, None
    elif knn_type == "hybird":
        F, r, mf, kf = int(motion_feature.shape[1]), 1e8, _c(motion_feature), _c(kp_feature)
    else:
        raise ValueError('knn_type should be "3D" or "hybird"')
    if hint is not None and (hint.shape != (N, K) or hint.dtype != torch.int32 or not hint.is_contiguous()):
        hint = None
    idx32 = hint if hint is not None else torch.empty((N, K), dtype=torch.int32, device=dev)
    idx64 = None if idx32_only else torch.empty((N, K), dtype=torch.int64, device=dev)
    dists = torch.empty((N, K), dtype=torch.float32, device=dev) if return_dists else None
    xyz_c, kp_c = _c(xyz), _c(kp_xyz)   # named: contiguous copies must outlive the launch
    check(_lib.lib().gp_knn_keypoints(N, Kp, F, int(K), float(r), float(feature_amplify), ptr(xyz_c), ptr(mf),
                                      ptr(kp_c), ptr(kf), ptr(order), ptr(hint), ptr(idx64), ptr(idx32), ptr(dists),
                                      _lib.stream_ptr()),
          "gp_knn_keypoints")
    return idx64, idx32, dists


class _KeypointBlend(torch.autograd.Function):
    """xyz_t, q_t, s_t, o_t = blend(logits, idx, kp deltas, canonical params) with the sparse kernels."""

    @staticmethod
    def forward(ctx, logits, idx32, kp_dxyz, kp_dq, xyz, rot, scaling, opacity):
        L = _lib.lib()
        N, K, Kp = int(xyz.shape[0]), int(idx32.shape[1]), int(kp_dxyz.shape[0])
        dev = xyz.device
        f = dict(dtype=torch.float32, device=dev)
        logits_c, kpx, kpq, xyz_c, rot_c = _c(logits), _c(kp_dxyz), _c(kp_dq), _c(xyz), _c(rot)
        sc_c, op_c = _c(scaling), _c(opacity)
        xyz_t, q_t = torch.empty((N, 3), **f), torch.empty((N, 4), **f)
        s_t, o_t = torch.empty((N, 3), **f), torch.empty((N, 1), **f)
        check(L.gp_blend_forward(N, Kp, K, ptr(logits_c), ptr(idx32), None, ptr(kpx), ptr(kpq), ptr(xyz_c), ptr(rot_c),
                                 ptr(sc_c), ptr(op_c), ptr(xyz_t), ptr(q_t), ptr(s_t), ptr(o_t), _lib.stream_ptr()),
              "gp_blend_forward")
        ctx.save_for_backward(logits_c, idx32, kpx, kpq, rot_c, s_t, o_t)
        ctx.dims = (N, Kp, K)
        ctx.sinks = [sink_for(t) for t in (logits, xyz, rot, scaling, opacity)]
        ctx.acc_sinks = [acc_sink_for(t) for t in (logits, xyz, rot, scaling, opacity)]
        return xyz_t, q_t, s_t, o_t

    @staticmethod
    def backward(ctx, g_xyz_t, g_q_t, g_s_t, g_o_t):
        L = _lib.lib()
        logits, idx32, kpx, kpq, rot, s_t, o_t = ctx.saved_tensors
        N, Kp, K = ctx.dims
        f = dict(dtype=torch.float32, device=logits.device)
        need = ctx.needs_input_grad
        sk = ctx.sinks
        # every per-Gaussian output has a running sum to add to: the kernel adds in place (gp_blend_backward_accumulate)
        accumulate = all(need[i] and a is not None for i, a in zip((0, 4, 5, 6, 7), ctx.acc_sinks))
        if accumulate:
            sk = ctx.acc_sinks
        g_logits = out_like(sk[0], (N, 2 * K), **f) if need[0] else None
        g_kpx, g_kpq = torch.empty((Kp, 3), **f), torch.empty((Kp, 4), **f)
        g_xyz = out_like(sk[1], (N, 3), **f) if need[4] else None
        g_rot = out_like(sk[2], (N, 4), **f) if need[5] else None
        g_sc = out_like(sk[3], (N, 3), **f) if need[6] else None
        g_op = out_like(sk[4], (N, 1), **f) if need[7] else None
        g_in = [None if g is None else _c(g) for g in (g_xyz_t, g_q_t, g_s_t, g_o_t)]   # kept alive until the launch
        entry = L.gp_blend_backward_accumulate if accumulate else L.gp_blend_backward
        check(entry(N, Kp, K, ptr(logits), ptr(idx32), None, ptr(kpx), ptr(kpq), ptr(rot), ptr(s_t), ptr(o_t),
                    ptr(g_in[0]), ptr(g_in[1]), ptr(g_in[2]), ptr(g_in[3]), ptr(g_logits),
                    ptr(g_xyz), ptr(g_rot), ptr(g_sc), ptr(g_op), ptr(g_kpx), ptr(g_kpq),
                    _lib.stream_ptr()), "gp_blend_backward")
        return g_logits, None, (g_kpx if need[2] else None), (g_kpq if need[3] else None), g_xyz, g_rot, g_sc, g_op


def keypoint_blend(logits, idx32, kp_delta_xyz, kp_delta_q, xyz, rot, scaling, opacity):
    return _KeypointBlend.apply(logits, idx32, kp_delta_xyz, kp_delta_q, xyz, rot, scaling, opacity)


class _KeypointMLP(torch.autograd.Function):
    """(delta_xyz, delta_q[, delta_o]) = get_motion_delta at the keypoints (cat([feature, PE(xyz), PE(t)]) -> df_model,
    optional quaternion normalisation), forward and backward in libgp_b200.so."""

    @staticmethod
    def forward(ctx, dims, kp_xyz, kp_feat, t, *params):
        L = _lib.lib()
        xyz_freq, time_freq, depth, width, out_dim, normalize_q = dims
        n = depth + 1
        weights, biases = [_c(w) for w in params[:n]], [_c(b) for b in params[n:]]
        xyz_c, feat_c = _c(kp_xyz), _c(kp_feat)
        t_c = _c(t).reshape(-1)[:1]
        cfg = _lib.MlpCfg(int(xyz_c.shape[0]), int(feat_c.shape[1]), xyz_freq, time_freq, depth, width, out_dim,
                          int(bool(normalize_q)))
        sv, bs = ctypes.c_size_t(), ctypes.c_size_t()
        check(L.gp_kpmlp_sizes(ctypes.byref(cfg), ctypes.byref(sv), ctypes.byref(bs)), "gp_kpmlp_sizes")
        dev = xyz_c.device
        f = dict(dtype=torch.float32, device=dev)
        saved = torch.empty(sv.value, dtype=torch.uint8, device=dev)
        dxyz, dq = torch.empty((cfg.rows, 3), **f), torch.empty((cfg.rows, 4), **f)
        do = torch.empty((cfg.rows, 1), **f) if out_dim == 8 else None
        wp = (ctypes.c_void_p * n)(*[w.data_ptr() for w in weights])
        bp = (ctypes.c_void_p * n)(*[b.data_ptr() for b in biases])
        check(L.gp_kpmlp_forward(ctypes.byref(cfg), ptr(xyz_c), ptr(feat_c), ptr(t_c), wp, bp, ptr(saved), ptr(dxyz),
                                 ptr(dq), ptr(do), _lib.stream_ptr()), "gp_kpmlp_forward")
        ctx.save_for_backward(xyz_c, t_c, saved, *weights)
        ctx.cfg, ctx.bwd_bytes = cfg, bs.value
        ctx.sinks = [sink_for(t) for t in (kp_xyz, kp_feat) + tuple(params)]
        ctx.set_materialize_grads(False)     # an unused output arrives as None (= zeros for the kernel), not as a zero tensor
        return (dxyz, dq) if do is None else (dxyz, dq, do)

    @staticmethod
    def backward(ctx, g_dxyz, g_dq, g_do=None):
        L = _lib.lib()
        xyz_c, t_c, saved, *weights = ctx.saved_tensors
        cfg = ctx.cfg
        n = cfg.depth + 1
        dev = xyz_c.device
        need = ctx.needs_input_grad
        sk = ctx.sinks
        f = dict(dtype=torch.float32, device=dev)
        g_xyz = out_like(sk[0], tuple(xyz_c.shape), **f) if need[1] else None
        g_feat = out_like(sk[1], (cfg.rows, cfg.feature_dim), **f) if need[2] else None
        gw = [out_like(sk[2 + i], tuple(w.shape), **f) for i, w in enumerate(weights)]
        gb = [out_like(sk[2 + n + i], (w.shape[0],), **f) for i, w in enumerate(weights)]
        scratch = torch.empty(ctx.bwd_bytes, dtype=torch.uint8, device=dev)
        wp = (ctypes.c_void_p * n)(*[w.data_ptr() for w in weights])
        gwp = (ctypes.c_void_p * n)(*[g.data_ptr() for g in gw])
        gbp = (ctypes.c_void_p * n)(*[g.data_ptr() for g in gb])
        # contiguous copies (if any) must stay referenced until the launch: a temporary freed right after taking its
        # address could be handed to the next allocation and overwritten before the kernel reads it
        g_in = [None if g is None else _c(g) for g in (g_dxyz, g_dq, g_do)]
        check(L.gp_kpmlp_backward(ctypes.byref(cfg), ptr(xyz_c), ptr(t_c), wp, ptr(saved), ptr(g_in[0]), ptr(g_in[1]),
                                  ptr(g_in[2]), ptr(scratch), ptr(g_xyz), ptr(g_feat), gwp, gbp, _lib.stream_ptr()),
              "gp_kpmlp_backward")
        del g_in
        return (None, g_xyz, g_feat, None) + tuple(gw) + tuple(gb)


_WARNED = set()


def _warn_torch_mlp(df_model, x):
    """There is no silent torch fallback: when the MLP cannot run on this library's kernels it says so (once per cause)."""
    import warnings
    why = "fused_mlp disabled / not a CUDA float32 [rows, dim] input" if (hasattr(df_model, "linears") and df_model.linears()) \
        else "df_model is not a plain Linear/ReLU stack"
    if why not in _WARNED:
        _WARNED.add(why)
        warnings.warn(f"gaussianprediction_b200: Deformable_Field runs through torch modules ({why}); "
                      "this is off the accelerated path", RuntimeWarning, stacklevel=3)


def keypoint_mlp_supported(df_model, feature_dim, xyz_freq, time_freq):
    """Shapes the fused kernels take (include/gp_b200.h, gp_kpmlp_*); anything else goes through the torch modules."""
    lin = df_model.linears() if hasattr(df_model, "linears") else None
    if lin is None:
        return False
    din = feature_dim + 6 * xyz_freq + 2 * time_freq
    w = lin[0].out_features
    ok = din % 4 == 0 and 4 <= din <= 256 and w % 4 == 0 and 4 <= w <= 256 and 2 <= len(lin) <= 9
    ok = ok and lin[0].in_features == din and lin[-1].out_features in (7, 8) and lin[-1].in_features == w
    ok = ok and all(l.in_features == w and l.out_features == w for l in lin[1:-1])
    return ok and all(l.bias is not None and l.weight.is_cuda and l.weight.dtype == torch.float32 for l in lin)


def keypoint_mlp(df_model, kp_xyz, kp_feat, t, xyz_freq, time_freq, normalize_q=False):
    """``get_motion_delta(PE(kp_xyz), kp_feat, PE(t))`` (scene/gaussian_model.py:180-189 at the keypoint call site)
    through the fused kernels: returns (delta_xyz [Kp,3], delta_q [Kp,4], delta_o [Kp,1] or None); with
    ``normalize_q`` delta_q is already ``rotation_activation(delta_q)`` (:266-267).  ``t``: tensor with one element."""
    lin = df_model.linears()
    dims = (int(xyz_freq), int(time_freq), len(lin) - 1, lin[0].out_features, lin[-1].out_features, bool(normalize_q))
    out = _KeypointMLP.apply(dims, kp_xyz, kp_feat, t, *[l.weight for l in lin], *[l.bias for l in lin])
    return out if len(out) == 3 else (out[0], out[1], None)


def _linear_tc(a, w, bias, relu, mask=None):
    """act(a @ w.T + bias) [* (mask > 0)] through gp_linear_forward (tcgen05, error-compensated TF32)."""
    M, K = int(a.shape[0]), int(a.shape[1])
    n_out = int(w.shape[0])
    out = torch.empty((M, n_out), dtype=torch.float32, device=a.device)
    check(_lib.lib().gp_linear_forward(M, K, n_out, ptr(a), K, ptr(mask), n_out if mask is not None else 0, ptr(w),
                                       ptr(bias), int(bool(relu)), ptr(out), n_out, _lib.stream_ptr()), "gp_linear_forward")
    return out


class _GaussianMLP(torch.autograd.Function):
    """``Deformable_Field`` (d x (Linear + ReLU), Linear) over ALL rows of x -- the per-Gaussian MLP of the first stage
    and of the step-opacity pass (scene/gaussian_model.py:240-244,265,291-298).  Forward and the data gradients run
    layer by layer on the tensor cores (csrc/linear_tc.cu); the weight gradients dZ^T H are plain library GEMMs
    (torch.mm) and the bias gradients column sums."""

    @staticmethod
    def forward(ctx, x, n_layers, *params):
        weights = [_c(w) for w in params[:n_layers]]
        biases = [None if b is None else _c(b) for b in params[n_layers:]]      # the hash-grid MLP has no biases
        hs = [_c(x)]
        for i, (w, b) in enumerate(zip(weights, biases)):
            hs.append(_linear_tc(hs[-1], w, b, relu=i < n_layers - 1))
        ctx.save_for_backward(*hs[:-1], *weights)
        ctx.n_layers = n_layers
        return hs[-1]

    @staticmethod
    def backward(ctx, g_out):
        n = ctx.n_layers
        saved = ctx.saved_tensors
        hs, weights = saved[:n], saved[n:]
        need = ctx.needs_input_grad
        dz = _c(g_out)
        gw, gb = [None] * n, [None] * n
        gx = None
        for i in range(n - 1, -1, -1):
            if need[2 + i]:
                gw[i] = dz.t().mm(hs[i])
            if len(need) > 2 + n + i and need[2 + n + i]:
                gb[i] = dz.sum(0)
            if i > 0:        # through the ReLU that produced hs[i]
                dz = _linear_tc(dz, weights[i].t().contiguous(), None, relu=False, mask=hs[i])
            elif need[0]:
                gx = _linear_tc(dz, weights[0].t().contiguous(), None, relu=False)
        return (gx, None) + tuple(gw) + tuple(gb)


def gaussian_mlp_supported(df_model, x):
    lin = df_model.linears() if hasattr(df_model, "linears") else None
    return lin is not None and x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and \
        all(l.bias is not None and l.out_features <= 256 and l.weight.dtype == torch.float32 and l.weight.is_cuda for l in lin)


def gaussian_mlp(df_model, x):
    lin = df_model.linears()
    return _GaussianMLP.apply(x, len(lin), *[l.weight for l in lin], *[l.bias for l in lin])


@torch.no_grad()
def dense_weights(logits, idx32, Kp, want_sum=False):
    """Dense ``weights_xyz, weights_r`` [N,Kp] as ``fill_nearest`` builds them (scene/gaussian_model.py:214-229)."""
    N, K = int(idx32.shape[0]), int(idx32.shape[1])
    f = dict(dtype=torch.float32, device=logits.device)
    wx, wr = torch.empty((N, Kp), **f), torch.empty((N, Kp), **f)
    ws = torch.empty((N, Kp), **f) if want_sum else None
    logits_c = _c(logits)
    check(_lib.lib().gp_blend_dense_weights(N, Kp, K, ptr(logits_c), ptr(idx32), None, ptr(wx), ptr(wr), ptr(ws),
                                            _lib.stream_ptr()), "gp_blend_dense_weights")
    return (wx, wr, ws) if want_sum else (wx, wr)


# ------------------------------------------------------------------------------------------------ the model call

def default_args(**over):
    """The hot-path sizing flags with the reference's defaults (options/gaussian_option.py:41-90,
    arguments/__init__.py:97-98)."""
    a = dict(nearest_num=6, knn_type="hybird", feature_amplify=5.0, norm_rotation=False, step_opacity=False,
             step_opacity_iteration=5000, opacity_type="implicit", beta=0.1, jointly_iteration=1000,
             second_stage_iteration=30000, xyz_noise_iteration=10000, feature_dim=32, d=4, w=256, max_points=100,
             xyz_freq=10, time_freq=6)
    a.update(over)
    return SimpleNamespace(**a)


class LogitsTable(nn.Module):
    """Stand-in for the absent tinycudann network: one learnable row of 2K logits per Gaussian."""

    def __init__(self, logits):
        super().__init__()
        self.logits = nn.Parameter(logits.clone())

    def forward(self, xyz):
        return self.logits


class KeypointDeformation(nn.Module):
    """Holds the tensors ``GaussianModel.forward`` reads and reproduces its call:
    ``xyz, q, s, opacity = model(t, iteration)`` (+ dense weights with ``return_weights=True``).
    Side-effect attributes other reference code reads are kept: ``kpts_xyz_motion``, ``kpts_rotation_motion``,
    ``nearest_mask``, ``lifecycle_opacity``, ``weights_sum`` (lazy: built on first access)."""

    def __init__(self, args, xyz, rotation, scaling, opacity, motion_feature, super_gaussians, super_gaussians_feature,
                 weights_model=None, logits=None):
        super().__init__()
        self.args = args
        P = nn.Parameter
        self._xyz, self._rotation, self._scaling, self._opacity = P(xyz), P(rotation), P(scaling), P(opacity)
        self.motion_feature = P(motion_feature)
        self.super_gaussians, self.super_gaussians_feature = P(super_gaussians), P(super_gaussians_feature)
        self.time_input_dim = 2 * args.time_freq
        self.xyz_input_dim = 6 * args.xyz_freq
        delta_dim = 8 if args.step_opacity else 7
        if args.step_opacity:  # step-opacity threshold (scene/gaussian_model.py:358-361)
            self.opacity_thres = P(-2 * torch.ones_like(opacity))
        self.df_model = DeformableField(self.time_input_dim + self.xyz_input_dim + args.feature_dim, delta_dim, args.d, args.w)
        if weights_model is None:
            if logits is None:
                logits = torch.zeros(xyz.shape[0], 2 * args.nearest_num)
            weights_model = LogitsTable(logits)
        self.weights_model = weights_model
        self.second_stage_iter = args.second_stage_iteration
        self.beta = args.beta
        self._nearest_mask = None
        self._nearest_flat = False
        self.lifecycle_opacity = None
        self._idx32 = None
        self._last_logits = None
        self.keep_kp_deltas, self._kp_deltas = False, None
        self._pre_blend_hook = None
        self.fused_mlp = True   # keypoint MLP through gp_kpmlp_* (False: the torch modules, e.g. to compare)
        self.knn_cache = False  # eval: the search inputs are parameters only, so it can be hoisted out of the frame loop
        # Morton order of the Gaussians, a speed hint for the search (results do not depend on it).  Built on first use
        # and whenever N changes; call refresh_knn_order() after the cloud has moved a lot (e.g. after densification).
        self._knn_order = None

    # getters with the reference's names (scene/gaussian_model.py:138-162)
    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_scaling(self):
        return torch.exp(self._scaling)

    @property
    def get_rotation(self):
        return torch.nn.functional.normalize(self._rotation)

    @property
    def get_opacity(self):
        return torch.sigmoid(self._opacity)

    @property
    def weights_sum(self):
        if self._idx32 is None:
            return None
        return dense_weights(self._last_logits, self._idx32, self.super_gaussians.shape[0], want_sum=True)[2]

    def refresh_knn_order(self):
        self._knn_order = spatial_order(self._xyz)

    def get_nearest_mask(self, keepshape=False):
        if self._knn_order is None or self._knn_order.shape[0] != self._xyz.shape[0]:
            self.refresh_knn_order()
        _, idx32, _ = knn_keypoints(self._xyz, self.motion_feature, self.super_gaussians, self.super_gaussians_feature,
                                    self.args.nearest_num, self.args.knn_type, self.args.feature_amplify,
                                    idx32_only=True, order=self._knn_order, hint=self._idx32)
        self._idx32 = idx32
        self._nearest_mask, self._nearest_flat = None, not keepshape

    @property
    def nearest_mask(self):
        """The reference's int64 attribute (scene/gaussian_model.py:123-125), widened from the int32 search result on
        access: the blend kernels read the int32 form, so a training step never writes the 48 B per Gaussian.  Not
        cached: a replayed CUDA graph refreshes the int32 buffer in place."""
        if self._nearest_mask is None and self._idx32 is not None:
            m = self._idx32.long()
            return m.view([-1]) if self._nearest_flat else m
        return self._nearest_mask

    @nearest_mask.setter
    def nearest_mask(self, value):
        self._nearest_mask = value

    def get_motion_delta(self, xyz_embed, motion_feature, t):
        t = t.repeat((motion_feature.shape[0], 1))
        x = torch.cat([motion_feature, xyz_embed, t], dim=-1)
        if self.fused_mlp and gaussian_mlp_supported(self.df_model, x):
            delta = gaussian_mlp(self.df_model, x)           # tensor-core path (csrc/linear_tc.cu)
        else:
            _warn_torch_mlp(self.df_model, x)
            delta = self.df_model(x)
        delta_o = delta[..., 7:8] if self.args.step_opacity else None
        return delta[..., 0:3], delta[..., 3:7], delta_o

    def forward(self, t, iteration, return_weights=False):
        args = self.args
        t_input = t
        t_embed = lambda: positional_encoding(t_input, self.time_input_dim // 2).unsqueeze(0)  # (:233), where it is used
        if torch.is_tensor(iteration):
            iteration = iteration.item()
        if iteration < args.jointly_iteration:  # warm-up: canonical parameters (:237-238)
            return self.get_xyz, self.get_rotation, self.get_scaling, self.get_opacity
        normalize = torch.nn.functional.normalize
        if iteration <= self.second_stage_iter:
            # first stage (:240-244): per-Gaussian MLP, torch ops (not the keypoint hot path)
            noise = torch.randn_like(self._xyz) * 0.1 * (1 - min(1, iteration / args.xyz_noise_iteration))
            xyz_embed = positional_encoding(self._xyz.detach() + noise, int(self.xyz_input_dim / 6))
            delta_xyz, delta_q, _ = self.get_motion_delta(xyz_embed, self.motion_feature, t_embed())
            if args.norm_rotation:
                delta_q = normalize(delta_q)
            self.kpts_xyz_motion, self.kpts_rotation_motion = delta_xyz.detach(), delta_q.detach()
            xyz = self._xyz + delta_xyz
            q = normalize(quat_mul(normalize(delta_q), self._rotation))
            s, opacity = self.get_scaling, self.get_opacity
        else:
            # second stage (:251-273, :285-287): keypoints
            factor = 1 - min(1, (iteration - self.second_stage_iter) / args.xyz_noise_iteration)
            noise = torch.randn_like(self.super_gaussians)  # RNG consumed even when the factor is 0 (:254)
            # same product order as the reference (randn * 0.1 * factor); + 0 * noise == the keypoints themselves
            kp_in = self.super_gaussians + noise * 0.1 * factor if factor != 0 else self.super_gaussians
            logits = self.weights_model(self._xyz.detach())
            if not (self.knn_cache and self._idx32 is not None):
                self.get_nearest_mask(keepshape=True)
            self._last_logits = logits.detach()
            xf, tf = int(self.xyz_input_dim / 6), self.time_input_dim // 2
            if self.fused_mlp and self.super_gaussians.is_cuda and \
                    keypoint_mlp_supported(self.df_model, args.feature_dim, xf, tf):
                delta_xyz, delta_q, _ = keypoint_mlp(self.df_model, kp_in,
                                                     self.super_gaussians_feature, t_input, xf, tf,
                                                     normalize_q=args.norm_rotation)
            else:
                # shapes outside the fused keypoint kernel's limits (include/gp_b200.h, gp_kpmlp_*): layer by layer
                # through gp_linear_forward; only a model that is not a plain Linear/ReLU stack reaches torch (warned)
                xyz_embed = positional_encoding(kp_in, xf)
                delta_xyz, delta_q, _ = self.get_motion_delta(xyz_embed, self.super_gaussians_feature, t_embed())
                if args.norm_rotation:
                    delta_q = normalize(delta_q)
            # detached: the readers (motion_model/dataset.py:120-123) only take values, and a live reference would keep
            # this frame's autograd graph (and its stream-bound AccumulateGrad nodes) alive across iterations
            self.kpts_xyz_motion, self.kpts_rotation_motion = delta_xyz.detach(), delta_q.detach()
            if self.keep_kp_deltas:          # pipeline.TrainStep cuts the backward here (and drops the reference again)
                self._kp_deltas = (delta_xyz, delta_q)
            if self._pre_blend_hook is not None:      # pipeline.TrainStep ends one captured graph here and starts the next:
                self._pre_blend_hook()                # what precedes reads only xyz / keypoints / MLP, what follows the rest
            xyz, q, s, opacity = keypoint_blend(logits, self._idx32, delta_xyz, delta_q, self._xyz, self._rotation,
                                                self._scaling, self._opacity)
        self.lifecycle_opacity = None
        base_opacity = opacity
        if args.step_opacity and iteration > args.step_opacity_iteration:  # (:291-301) per-Gaussian MLP, torch ops
            xyz_embed = positional_encoding(self._xyz, int(self.xyz_input_dim / 6))
            _, _, delta_o = self.get_motion_delta(xyz_embed, self.motion_feature, t_embed())
            if args.opacity_type == "explicit":
                opacity = base_opacity * (1 / (1 + torch.exp(-(t_input - self.opacity_thres) / self.beta)))
            else:
                opacity = base_opacity * (1 / (1 + torch.exp(-delta_o / self.beta)))
            self.lifecycle_opacity = opacity
        if return_weights and iteration > self.second_stage_iter:
            wx, wr = dense_weights(self._last_logits, self._idx32, self.super_gaussians.shape[0])
            return xyz, q, s, base_opacity, wx, wr
        return xyz, q, s, opacity

    @torch.no_grad()
    def extrapolate(self, kpts_i, kpts_rotation_i):
        """One frame of ``render_kpts`` (eval.py:140-141): xyz + W_x (kpts_i - kpts), quaternion compose with
        W_r kpts_rotation_i, using the cached indices / logits of the last forward (no dense matrices)."""
        xyz_t, q_t, _, _ = keypoint_blend(self._last_logits, self._idx32, kpts_i - self.super_gaussians, kpts_rotation_i,
                                          self._xyz, self._rotation, self._scaling, self._opacity)
        return xyz_t, q_t

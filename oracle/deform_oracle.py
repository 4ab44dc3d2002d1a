"""CPU restatement (torch / numpy, float32) of the reference's keypoint-stage deformation.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference legs; never by gaussianprediction_b200/.

Parity status: PINNED for ``positional_encoding`` and ``Deformable_Field`` against the reference's own
modules (loaded by file path from /root/reference/scene/deformable_field.py by tests/golden/make_golden.py,
vectors in tests/golden/deform_*.npz) and for the kNN against the reference's CUDA kernels
(oracle/_ref/ref_frnn.so, same script, run on a B200) and its CPU brute force ``frnn_bf_cpu``.
The blend / quaternion chain (``fill_nearest``, ``quat_mul``, ``keypoint_forward``, ``model_forward``,
``extrapolate_frame``) is PINNED too: ``scene.gaussian_model`` cannot be imported (tinycudann, pytorch3d, ... are
absent), so tests/golden/make_golden.py cuts ``GaussianModel.forward / fill_nearest / get_motion_delta / get_rotation_ /
get_nearest_mask``, ``quat_mul`` and the two extrapolation statements of ``render_kpts`` out of the reference files as
text, executes them against a stub ``self`` and stores outputs + autograd gradients in tests/golden/blend_chain.npz
(all three stages, both step-opacity types, keypoint noise, ``return_weights``); tests/test_oracle_golden.py checks
this file against it.  Only the tinycudann hash-grid network stays unpinned (its source is not in the tree).
"""
import numpy as np
import torch


def positional_encoding(positions, freqs):
    """scene/deformable_field.py:63-72 (ori=False)."""
    freq_bands = (2 ** torch.arange(freqs).float()).to(positions.device)
    pts = (positions[..., None] * freq_bands).reshape(positions.shape[:-1] + (freqs * positions.shape[-1],))
    return torch.stack([torch.sin(pts), torch.cos(pts)], dim=-1).reshape(pts.shape[:-1] + (pts.shape[-1] * 2,))


def mlp_forward(state_dict, x, d=4):
    """``Deformable_Field.forward`` (scene/deformable_field.py:112-127, split_xyz=False): d x (Linear, ReLU), Linear."""
    h = x
    for i in range(d):
        h = torch.relu(torch.nn.functional.linear(h, state_dict[f"mlp.{2 * i}.weight"], state_dict[f"mlp.{2 * i}.bias"]))
    return torch.nn.functional.linear(h, state_dict["feature_to_deformation.0.weight"],
                                      state_dict["feature_to_deformation.0.bias"])


def quat_mul(q1, q2):
    """utils/camera_utils.py:158-170."""
    q1q2 = torch.zeros_like(q1)
    q1q2[..., 0] = q2[..., 0] * q1[..., 0] - q2[..., 1] * q1[..., 1] - q2[..., 2] * q1[..., 2] - q2[..., 3] * q1[..., 3]
    q1q2[..., 1] = q2[..., 1] * q1[..., 0] + q2[..., 0] * q1[..., 1] + q2[..., 3] * q1[..., 2] - q2[..., 2] * q1[..., 3]
    q1q2[..., 2] = q2[..., 2] * q1[..., 0] - q2[..., 3] * q1[..., 1] + q2[..., 0] * q1[..., 2] + q2[..., 1] * q1[..., 3]
    q1q2[..., 3] = q2[..., 3] * q1[..., 0] + q2[..., 2] * q1[..., 1] - q2[..., 1] * q1[..., 2] + q2[..., 0] * q1[..., 3]
    return q1q2


def knn(points1, points2, K, r=1e8, fma=True):
    """Exhaustive K nearest points2 rows for every points1 row; returns (dists [N,K] f32, idxs [N,K] int64).

    Distance = sequential accumulation over d of (p2 - p1)^2 exactly as FindNbrsNDKernelV0
    (frnn/csrc/grid/find_nbrs.cu:308-314): with ``fma=True`` each step is a fused multiply-add (what nvcc emits for
    the GPU kernel), with ``fma=False`` product and sum round separately (the reference's CPU code,
    bruteforce_cpu.cpp:36-41).  Selection = MinK (utils/mink.cuh:55-97) on tie-free data: the K smallest, ascending;
    exact ties resolve to the smaller index."""
    p1 = np.ascontiguousarray(points1, dtype=np.float32)
    p2 = np.ascontiguousarray(points2, dtype=np.float32)
    N, D = p1.shape
    Kp = p2.shape[0]
    acc = np.zeros((N, Kp), dtype=np.float32)
    for d in range(D):
        diff = (p2[None, :, d] - p1[:, None, d]).astype(np.float32)
        if fma:
            acc = (diff.astype(np.float64) * diff.astype(np.float64) + acc.astype(np.float64)).astype(np.float32)
        else:
            acc = (acc + (diff * diff).astype(np.float32)).astype(np.float32)
    r2 = np.float32(r) * np.float32(r)
    key = np.where(acc <= r2, acc, np.float32(np.inf))
    order = np.argsort(key, axis=1, kind="stable")[:, :K]   # stable: ties -> smaller index
    d_sorted = np.take_along_axis(key, order, axis=1)
    idx = order.astype(np.int64)
    if Kp < K:
        pad = K - Kp
        idx = np.concatenate([idx, -np.ones((N, pad), np.int64)], 1)
        d_sorted = np.concatenate([d_sorted, -np.ones((N, pad), np.float32)], 1)
    miss = ~np.isfinite(d_sorted)
    idx[miss] = -1
    d_sorted = np.where(miss, np.float32(-1), d_sorted).astype(np.float32)
    return d_sorted, idx


def knn_rel_gaps(points1, points2, K):
    """Per query: the smallest relative gap between consecutive entries among its K+1 nearest distances.  The tie
    policy of SURVEY.md §8c excludes queries where this is < 1e-6 (exact ties are broken nondeterministically by
    the reference, and an FMA-vs-no-FMA ulp can swap near-ties)."""
    d, _ = knn(points1, points2, min(K + 1, points2.shape[0]), fma=False)
    gap = (d[:, 1:] - d[:, :-1]) / np.maximum(d[:, 1:], 1e-30)
    return gap.min(axis=1)


def knn_min_rel_gap(points1, points2, K):
    return float(knn_rel_gaps(points1, points2, K).min())


def hybrid_points(xyz, motion_feature, kp_xyz, kp_feature, feature_amplify):
    """The two torch.cat of get_nearest_mask (scene/gaussian_model.py:115-116)."""
    return (torch.cat([xyz, motion_feature * feature_amplify], dim=-1),
            torch.cat([kp_xyz, kp_feature * feature_amplify], dim=-1))


def fill_nearest(weights, nearest_mask, Kp, K):
    """scene/gaussian_model.py:214-229: two softmaxes scattered into dense zeros [N,Kp]."""
    weights_xyz = torch.softmax(weights[..., 0:K], dim=-1)
    weights_r = torch.softmax(weights[..., K:2 * K], dim=-1)
    zeros = torch.zeros(weights.shape[0], Kp, dtype=weights.dtype, device=weights.device)
    return torch.scatter(zeros, -1, nearest_mask, weights_xyz), torch.scatter(zeros, -1, nearest_mask, weights_r)


def keypoint_forward(t, xyz, rotation, scaling, opacity, super_gaussians, super_gaussians_feature, logits, nearest_mask,
                     df_state, time_freq=6, xyz_freq=10, d=4, norm_rotation=False, noise=None):
    """Second-stage ``GaussianModel.forward`` (scene/gaussian_model.py:231-304) with the dense formulation
    (scatter + [N,Kp] matmuls), differentiable in torch.  Returns (xyz_t, q_t, s_t, o_t, kp_dxyz, kp_dq)."""
    normalize = torch.nn.functional.normalize
    K = nearest_mask.shape[1]
    Kp = super_gaussians.shape[0]
    t_enc = positional_encoding(t, time_freq).unsqueeze(0)                       # :233
    kp = super_gaussians if noise is None else super_gaussians + noise           # :254-255
    xyz_embed = positional_encoding(kp, xyz_freq)
    weights_xyz, weights_r = fill_nearest(logits, nearest_mask, Kp, K)           # :261
    tt = t_enc.repeat((super_gaussians_feature.shape[0], 1))                     # :181
    delta = mlp_forward(df_state, torch.cat([super_gaussians_feature, xyz_embed, tt], dim=-1), d)   # :183
    delta_xyz, delta_q = delta[..., 0:3], delta[..., 3:7]
    if norm_rotation:
        delta_q = normalize(delta_q)                                             # :266-267
    dq = weights_r @ delta_q                                                     # :272
    dx = weights_xyz @ delta_xyz                                                 # :273
    xyz_t = xyz + dx                                                             # :285
    q_t = normalize(quat_mul(normalize(dq), rotation))                           # :286, :314-315
    return xyz_t, q_t, torch.exp(scaling), torch.sigmoid(opacity), delta_xyz, delta_q


def model_forward(t, iteration, p, df_state, args, nearest_mask=None, noise=None, return_weights=False):
    """All branches of ``GaussianModel.forward`` (scene/gaussian_model.py:231-304).  ``p``: dict with xyz, rotation,
    scaling, opacity, motion_feature, kp, kf, logits and (step opacity) thres; ``args``: the reference's flag object;
    ``noise``: the ``randn_like`` draw of :240 / :254 (None = zeros).  Returns what the reference returns, plus a dict
    of the side-effect attributes (kpts_xyz_motion, kpts_rotation_motion, weights_sum, lifecycle_opacity)."""
    normalize = torch.nn.functional.normalize
    side = {}
    t_input = t
    t_enc = positional_encoding(t, args.time_freq).unsqueeze(0)                              # :233
    get_opacity, get_scaling = torch.sigmoid(p["opacity"]), torch.exp(p["scaling"])
    if iteration < args.jointly_iteration:                                                   # :237-238
        return (p["xyz"], normalize(p["rotation"]), get_scaling, get_opacity), side
    second = iteration > args.second_stage_iteration
    d = getattr(args, "d", 4)

    def motion_delta(xyz_embed, feat):                                                       # :180-189
        tt = t_enc.repeat((feat.shape[0], 1))
        delta = mlp_forward(df_state, torch.cat([feat, xyz_embed, tt], dim=-1), d)
        return delta[..., 0:3], delta[..., 3:7], (delta[..., 7:8] if args.step_opacity else None)

    if not second:                                                                           # :240-244
        f = 1 - min(1, iteration / args.xyz_noise_iteration)
        x = p["xyz"].detach() + (0 if noise is None else noise * 0.1 * f)    # same product order as :240
        xyz_embed, feat = positional_encoding(x, args.xyz_freq), p["motion_feature"]
    else:                                                                                    # :251-263
        f = 1 - min(1, (iteration - args.second_stage_iteration) / args.xyz_noise_iteration)
        xyz_embed = positional_encoding(p["kp"] + (0 if noise is None else noise * 0.1 * f), args.xyz_freq)   # :254-255
        feat = p["kf"]
        K, Kp = args.nearest_num, p["kp"].shape[0]
        weights_xyz, weights_r = fill_nearest(p["logits"], nearest_mask, Kp, K)
        side["weights_sum"] = torch.abs(weights_xyz) + torch.abs(weights_r)
    delta_xyz, delta_q, _ = motion_delta(xyz_embed, feat)                                    # :265
    if args.norm_rotation:
        delta_q = normalize(delta_q)                                                         # :266-267
    side["kpts_xyz_motion"], side["kpts_rotation_motion"] = delta_xyz, delta_q               # :269-270
    if second:
        delta_q = weights_r @ delta_q                                                        # :272
        delta_xyz = weights_xyz @ delta_xyz                                                  # :273
    xyz = p["xyz"] + delta_xyz                                                               # :285
    q = normalize(quat_mul(normalize(delta_q), p["rotation"]))                               # :286, :314-315
    opacity = get_opacity
    if args.step_opacity and iteration > args.step_opacity_iteration:                        # :291-301
        _, _, delta_o = motion_delta(positional_encoding(p["xyz"], args.xyz_freq), p["motion_feature"])
        if args.opacity_type == "explicit":
            opacity = get_opacity * (1 / (1 + torch.exp(-(t_input - p["thres"]) / args.beta)))   # :50
        else:
            opacity = get_opacity * (1 / (1 + torch.exp(-delta_o / args.beta)))                  # :51
        side["lifecycle_opacity"] = opacity
    if return_weights and second:
        return (xyz, q, get_scaling, get_opacity, weights_xyz, weights_r), side              # :299-300, :302-303
    return (xyz, q, get_scaling, opacity), side


def extrapolate_frame(xyz, rotation, weights_xyz, weights_r, kpts_i, kpts, kpts_rotation_i):
    """One frame of render_kpts (eval.py:140-141)."""
    normalize = torch.nn.functional.normalize
    xyz_final = xyz + (weights_xyz @ (kpts_i - kpts))
    delta_r = normalize(quat_mul(normalize(weights_r @ kpts_rotation_i), rotation))
    return xyz_final, delta_r


def dist2_knn3(points):
    """distCUDA2 (submodules/simple-knn/simple_knn.cu:185-221): mean of the three smallest squared distances to
    OTHER points, distance = fma(dz,dz, fma(dx,dx, dy*dy)) as the CUDA build contracts :141-142, summed ascending,
    divided by 3.0f (:182).  O(P^2): small P only."""
    p = np.ascontiguousarray(points, dtype=np.float32)
    P = p.shape[0]
    d = (p[None, :, :] - p[:, None, :]).astype(np.float32)
    f64 = np.float64
    yy = (d[..., 1] * d[..., 1]).astype(np.float32)
    a = (d[..., 0].astype(f64) * d[..., 0].astype(f64) + yy.astype(f64)).astype(np.float32)
    dist = (d[..., 2].astype(f64) * d[..., 2].astype(f64) + a.astype(f64)).astype(np.float32)
    dist[np.arange(P), np.arange(P)] = np.float32(np.finfo(np.float32).max)
    best = np.sort(dist, axis=1)[:, :3]
    s = (best[:, 0] + best[:, 1]).astype(np.float32)
    s = (s + best[:, 2]).astype(np.float32)
    return (s / np.float32(3.0)).astype(np.float32)

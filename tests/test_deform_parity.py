"""GPU parity of the deformation stage (kNN, blend fwd/bwd, distCUDA2, the model call) through the C ABI.

  kNN      : bit-exact indices AND squared distances vs the reference FRNN CUDA path (oracle/_ref/ref_frnn.so) on
             tie-free inputs (SURVEY.md §8c tie policy: min relative gap among the K+1 nearest >= 1e-6), and vs the
             FMA-ordered CPU oracle; golden vectors from tests/golden/knn_gpu.npz.
  blend    : forward <= 1e-5 max-abs, gradients <= 1e-3 relative vs the dense torch formulation of the oracle
             (scatter + [N,Kp] matmul, autograd).
  distCUDA2: bit-exact vs the reference simple-knn build and the O(P^2) oracle.
"""
import os

import numpy as np
import pytest
import torch

from helpers import count_diff, rel_err, to_np

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def scene(N=20000, Kp=512, seed=2024, amplify=0.5):
    from gaussianprediction_b200 import synthetic as syn
    gs = syn.make_gaussians(N, seed=seed)
    kp, kf = syn.make_keypoints(gs, Kp=Kp, seed=seed + 1)
    logits = syn.make_logits(N, 6, seed=seed + 2)
    return gs, kp, kf, logits


@pytest.mark.parametrize("N,Kp", [(700, 64), (20000, 512), (5000, 37)])
@pytest.mark.parametrize("mode", ["hybird", "3D"])
def test_knn_bit_exact_vs_reference(gplib, N, Kp, mode):
    from gaussianprediction_b200.deform import knn_keypoints
    from oracle import deform_oracle as do
    from oracle import ref_cuda
    gs, kp, kf, _ = scene(N, Kp, seed=5)
    amplify = 0.5
    p1, p2 = do.hybrid_points(gs.xyz, gs.motion_feature, kp, kf, amplify)
    if mode == "3D":
        p1, p2 = gs.xyz, kp
    ok = do.knn_rel_gaps(p1.numpy(), p2.numpy(), 6) >= 1e-6      # tie policy: near-tied queries are excluded
    assert ok.mean() > 0.999, "too many near-ties in the input; regenerate"
    idx64, idx32, dists = knn_keypoints(gs.xyz.cuda(), gs.motion_feature.cuda(), kp.cuda(), kf.cuda(), 6, mode, amplify,
                                        return_dists=True)
    od, oi = do.knn(p1.numpy(), p2.numpy(), 6, r=1e8 if mode == "hybird" else 1e3, fma=True)
    assert count_diff(to_np(idx64)[ok], oi[ok]) == 0 and count_diff(to_np(idx32)[ok], oi[ok].astype(np.int32)) == 0
    assert np.abs(to_np(dists) - od)[ok].max() <= 1e-6 * max(od.max(), 1.0)
    if ref_cuda.available("ref_frnn") and ref_cuda.available("ref_prefix_sum"):
        rd, ri = ref_cuda.frnn_grid_points(p1[None].cuda().contiguous(), p2[None].cuda().contiguous(), 6,
                                           1e8 if mode == "hybird" else 1e3)
        assert count_diff(to_np(idx64)[ok], to_np(ri[0])[ok]) == 0, "indices differ from the reference CUDA kNN"
        assert count_diff(to_np(dists)[ok], to_np(rd[0])[ok]) == 0, "squared distances are not bit-identical to the reference CUDA kNN"
        # on near-tied queries the SET of distances must still agree bit for bit
        assert count_diff(np.sort(to_np(dists), 1), np.sort(to_np(rd[0]), 1)) == 0


def test_knn_golden_and_frnn_api(gplib):
    from gaussianprediction_b200 import frnn
    g = np.load(os.path.join(GOLD, "knn_gpu.npz")) if os.path.exists(os.path.join(GOLD, "knn_gpu.npz")) else None
    c = np.load(os.path.join(GOLD, "knn_cpu.npz"))
    p1, p2 = torch.tensor(c["p1"]).cuda(), torch.tensor(c["p2"]).cuda()
    dists, idxs, nn, grid = frnn.frnn_grid_points(p1[None], p2[None], K=6, r=1e8, return_nn=True)
    assert idxs.dtype == torch.int64 and idxs.shape == (1, 700, 6) and nn.shape == (1, 700, 6, 35) and grid is None
    assert count_diff(idxs[0], c["i35"]) == 0                       # the reference's CPU brute force
    assert np.abs(to_np(dists[0]) - c["d35"]).max() <= 1e-5
    if g is not None:
        assert count_diff(idxs[0], g["i35"]) == 0 and count_diff(dists[0], g["d35"]) == 0   # reference CUDA, bit-exact
        d3, i3, _, _ = frnn.frnn_grid_points(p1[None, :, :3].contiguous(), p2[None, :, :3].contiguous(), K=6, r=1e3)
        assert count_diff(i3[0], g["i3"]) == 0 and count_diff(d3[0], g["d3"]) == 0
    # generic-D path (D = 7), radius that excludes most points, K > number of references
    q, r = torch.randn(1, 300, 7, device="cuda"), torch.randn(1, 5, 7, device="cuda")
    d, i, _, _ = frnn.frnn_grid_points(q, r, K=8, r=2.0)
    from oracle import deform_oracle as do
    od, oi = do.knn(to_np(q[0]), to_np(r[0]), 8, r=2.0, fma=True)
    assert count_diff(i[0], oi) == 0 and (to_np(i[0])[:, 5:] == -1).all()
    with pytest.raises(ValueError):
        frnn.frnn_grid_points(q, r[:, :, :3], K=2, r=1.0)
    with pytest.raises(TypeError):
        frnn.frnn_grid_points(q.cpu(), r.cpu(), K=2, r=1.0)


def test_knn_ties_break_by_index(gplib):
    from gaussianprediction_b200 import frnn
    r = torch.zeros(1, 8, 3, device="cuda")          # eight identical reference points: all distances tie
    q = torch.ones(1, 4, 3, device="cuda")
    _, i, _, _ = frnn.frnn_grid_points(q, r, K=3, r=10.0)
    assert (to_np(i[0]) == np.array([0, 1, 2])).all()


def _dense_reference(gs, kp, kf, logits, idx64, kp_dxyz, kp_dq, dev):
    from oracle import deform_oracle as do
    leaf = dict(logits=logits, kp_dxyz=kp_dxyz, kp_dq=kp_dq, xyz=gs.xyz, rot=gs.rotation, sc=gs.scaling, op=gs.opacity)
    leaf = {k: v.clone().to(dev).requires_grad_(True) for k, v in leaf.items()}
    K, Kp = idx64.shape[1], kp.shape[0]
    wx, wr = do.fill_nearest(leaf["logits"], idx64.to(dev), Kp, K)
    normalize = torch.nn.functional.normalize
    dq = wr @ leaf["kp_dq"]
    dx = wx @ leaf["kp_dxyz"]
    xyz_t = leaf["xyz"] + dx
    q_t = normalize(do.quat_mul(normalize(dq), leaf["rot"]))
    return leaf, (xyz_t, q_t, torch.exp(leaf["sc"]), torch.sigmoid(leaf["op"]))


@pytest.mark.parametrize("N,Kp", [(3000, 100), (20000, 512)])
def test_blend_forward_backward_vs_dense_torch(gplib, N, Kp):
    from gaussianprediction_b200.deform import dense_weights, keypoint_blend, knn_keypoints
    from oracle import deform_oracle as do
    gs, kp, kf, logits = scene(N, Kp, seed=8)
    g = torch.Generator().manual_seed(3)
    kp_dxyz = 0.2 * torch.randn(Kp, 3, generator=g)
    kp_dq = torch.nn.functional.normalize(torch.randn(Kp, 4, generator=g))
    idx64, idx32, _ = knn_keypoints(gs.xyz.cuda(), gs.motion_feature.cuda(), kp.cuda(), kf.cuda(), 6, "hybird", 0.5)
    mine_leaf = dict(logits=logits, kp_dxyz=kp_dxyz, kp_dq=kp_dq, xyz=gs.xyz, rot=gs.rotation, sc=gs.scaling, op=gs.opacity)
    mine_leaf = {k: v.clone().cuda().requires_grad_(True) for k, v in mine_leaf.items()}
    outs = keypoint_blend(mine_leaf["logits"], idx32, mine_leaf["kp_dxyz"], mine_leaf["kp_dq"], mine_leaf["xyz"],
                          mine_leaf["rot"], mine_leaf["sc"], mine_leaf["op"])
    ref_leaf, ref_outs = _dense_reference(gs, kp, kf, logits, idx64.cpu(), kp_dxyz, kp_dq, "cpu")
    for a, b, name in zip(outs, ref_outs, ("xyz_t", "q_t", "s_t", "o_t")):
        assert float((a.cpu() - b).abs().max()) <= 1e-5, name
    gens = [torch.randn(o.shape, generator=g) for o in ref_outs]
    sum((o * w.cuda()).sum() for o, w in zip(outs, gens)).backward()
    sum((o * w).sum() for o, w in zip(ref_outs, gens)).backward()
    for k in mine_leaf:
        assert rel_err(mine_leaf[k].grad, ref_leaf[k].grad) <= 1e-3, k
    # dense weights for API parity (return_weights=True)
    wx, wr = dense_weights(mine_leaf["logits"].detach(), idx32, Kp)
    ox, orr = do.fill_nearest(logits, idx64.cpu(), Kp, 6)
    assert float((wx.cpu() - ox).abs().max()) <= 1e-6 and float((wr.cpu() - orr).abs().max()) <= 1e-6


def test_dist2_bit_exact(gplib):
    from gaussianprediction_b200.simple_knn import distCUDA2
    from oracle import deform_oracle as do
    from oracle import ref_cuda
    gs, *_ = scene(3000, 8, seed=11)
    pts = gs.xyz[:3000].contiguous()
    mine = distCUDA2(pts.cuda())
    assert count_diff(mine, do.dist2_knn3(pts.numpy())) == 0
    gpath = os.path.join(GOLD, "dist2.npz")
    if os.path.exists(gpath):
        g = np.load(gpath)
        assert count_diff(distCUDA2(torch.tensor(g["points"]).cuda()), g["dist2"]) == 0
    if ref_cuda.available("ref_sknn"):
        for n, seed, shape in [(50000, 12, "dnerf"), (200000, 13, "hyper"), (5, 14, "dnerf"), (1000, 15, "dnerf")]:
            from gaussianprediction_b200 import synthetic as syn
            p = syn.make_gaussians(n, seed=seed, shape=shape).xyz.cuda()
            if n == 1000:
                p = torch.cat([p, p[:100]], 0).contiguous()      # duplicated points -> zero distances
            assert count_diff(distCUDA2(p), ref_cuda.dist_cuda2(p)) == 0, (n, shape)


def test_model_call_matches_oracle_and_render_runs(gplib):
    from gaussianprediction_b200 import synthetic as syn
    from gaussianprediction_b200.deform import KeypointDeformation, default_args
    from gaussianprediction_b200.renderer import SceneModel, render
    from oracle import deform_oracle as do
    N, Kp = 6000, 128
    gs, kp, kf, logits = scene(N, Kp, seed=21)
    args = default_args(feature_amplify=0.5, norm_rotation=True, second_stage_iteration=100, jointly_iteration=10,
                        xyz_noise_iteration=50)
    torch.manual_seed(0)
    model = KeypointDeformation(args, gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf, logits=logits).cuda()
    t = torch.tensor([0.37], device="cuda")
    it = 1000    # past second_stage + noise window: deterministic
    xyz_t, q_t, s_t, o_t, wx, wr = model(t, it, return_weights=True)
    assert wx.shape == (N, Kp) and model.nearest_mask.shape == (N, 6) and model.nearest_mask.dtype == torch.int64
    assert model.kpts_xyz_motion.shape == (Kp, 3) and model.kpts_rotation_motion.shape == (Kp, 4)
    sd = {k: v.detach().cpu() for k, v in model.df_model.state_dict().items()}
    o = do.keypoint_forward(t.cpu(), gs.xyz, gs.rotation, gs.scaling, gs.opacity, kp, kf, logits, model.nearest_mask.cpu(),
                            sd, norm_rotation=True)
    for a, b, name in zip((xyz_t, q_t, s_t, o_t), o[:4], ("xyz", "q", "s", "o")):
        assert float((a.detach().cpu() - b).abs().max()) <= 2e-5, name
    # warm-up stage returns canonical parameters
    c = model(t, 0)
    assert float((c[0] - model.get_xyz).abs().max()) == 0 and float((c[1] - model.get_rotation).abs().max()) == 0
    # extrapolation frame (eval.py:140-141)
    kpts, krot = syn.keypoint_trajectory(kp, frames=4)
    xe, qe = model.extrapolate(kpts[2].cuda(), krot[2].cuda())
    xo, qo = do.extrapolate_frame(gs.xyz, gs.rotation, wx.cpu(), wr.cpu(), kpts[2], kp, krot[2])
    assert float((xe.cpu() - xo).abs().max()) <= 2e-5 and float((qe.cpu() - qo).abs().max()) <= 2e-5
    # full render(): deformation -> rasterizer, forward + backward through both
    cam = syn.make_camera(160, 112)
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).cuda())
    pc = SceneModel(model, gs.features_dc, gs.features_rest).cuda()
    pkg = render(cam, pc, None, torch.zeros(3, device="cuda"), time=t, it=it)
    assert pkg["render"].shape == (3, 112, 160) and pkg["depth"].shape == (1, 112, 160)
    pkg["render"].square().mean().backward()
    for name, p in pc.named_parameters():
        if name.endswith("motion_feature"):
            continue            # only feeds the (non-differentiable) search
        assert p.grad is not None and torch.isfinite(p.grad).all(), name
    assert float(model.super_gaussians.grad.abs().sum()) > 0 and float(model.weights_model.logits.grad.abs().sum()) > 0


def test_graph_step_matches_eager(gplib):
    """pipeline.TrainStep (whole iteration as one CUDA graph, sync-free capacity mode) gives the eager results."""
    from gaussianprediction_b200 import synthetic as syn
    from gaussianprediction_b200.deform import KeypointDeformation, default_args
    from gaussianprediction_b200.pipeline import RenderStep, TrainStep
    from gaussianprediction_b200.renderer import SceneModel, render
    N, Kp, W, H = 8000, 128, 160, 112
    gs, kp, kf, logits = scene(N, Kp, seed=31)
    args = default_args(feature_amplify=0.5, norm_rotation=True)
    torch.manual_seed(0)
    model = KeypointDeformation(args, gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf, logits=logits)
    pc = SceneModel(model, gs.features_dc, gs.features_rest).cuda()
    cams = syn.camera_ring(3, W, H, seed=5)
    for c in cams:
        for k in ("world_view_transform", "full_proj_transform", "camera_center"):
            setattr(c, k, getattr(c, k).cuda())
    bg = torch.zeros(3, device="cuda")
    g = torch.Generator().manual_seed(1)
    targets = [(torch.randn(3, H, W, generator=g) * 1e-2).cuda() for _ in cams]
    times = [torch.tensor([0.2 + 0.3 * i], device="cuda") for i in range(len(cams))]
    step = TrainStep(pc, cams[0], capacity=400_000, bg=bg)
    rstep = RenderStep(pc, cams[0], capacity=400_000, bg=bg)
    for cam, t, tgt in zip(cams, times, targets):
        loss_g = float(step(cam, t, tgt))
        R, over = step.overflowed()
        assert not over and R > 0
        grads_g = {n: p.grad.clone() for n, p in pc.named_parameters() if p.grad is not None}
        img_g = step.image.clone()
        for p in pc.parameters():
            p.grad = None
        pkg = render(cam, pc, None, bg, time=t, it=50_000)
        loss_e = (pkg["render"] * tgt).sum()
        loss_e.backward()
        assert count_diff(img_g, pkg["render"]) == 0                 # same kernels, same inputs: bit-identical image
        assert abs(loss_g - float(loss_e)) <= 1e-6 * max(1.0, abs(float(loss_e)))
        for n, p in pc.named_parameters():
            if p.grad is not None:
                assert rel_err(grads_g[n], p.grad) <= 1e-3, n
        assert count_diff(rstep(cam, t)["render"], pkg["render"]) == 0
    # a capacity that is too small is reported, not silently rendered
    small = TrainStep(pc, cams[0], capacity=1000, bg=bg)
    small(cams[0], times[0], targets[0])
    assert small.overflowed()[1]


def _torch_motion_delta(net, kp, kf, t, xf, tf):
    """get_motion_delta at the keypoint call site with the torch modules (scene/gaussian_model.py:180-189)."""
    from gaussianprediction_b200.deform import positional_encoding
    cols = [kf]
    if xf:
        cols.append(positional_encoding(kp, xf))
    if tf:
        cols.append(positional_encoding(t, tf).unsqueeze(0).repeat((kp.shape[0], 1)))
    return net(torch.cat(cols, dim=-1))


@pytest.mark.parametrize("Kp,F,xf,tf,d,w,out,norm", [(512, 32, 10, 6, 4, 256, 7, True), (515, 32, 10, 6, 4, 256, 8, False),
                                                     (1, 32, 10, 6, 4, 256, 7, True), (37, 8, 4, 2, 2, 64, 7, True),
                                                     (130, 16, 0, 0, 1, 32, 8, True), (64, 0, 10, 6, 8, 128, 7, False)])
def test_keypoint_mlp_vs_torch_modules(gplib, Kp, F, xf, tf, d, w, out, norm):
    """gp_kpmlp_forward/backward against the torch fp32 modules they replace (PE + cat + Linear/ReLU stack +
    F.normalize + autograd).  Tolerances: values 1e-5 of the output scale, gradients 1e-4 relative (fp32 sums in a
    different order than cuBLAS)."""
    from gaussianprediction_b200.deform import DeformableField, keypoint_mlp, keypoint_mlp_supported
    dev = "cuda"
    torch.manual_seed(Kp + w)
    net = DeformableField(F + 6 * xf + 2 * tf, out, d, w).to(dev)
    assert keypoint_mlp_supported(net, F, xf, tf)
    g = torch.Generator().manual_seed(3)
    kp = (torch.rand(Kp, 3, generator=g) * 2 - 1).to(dev).requires_grad_(True)
    kf = torch.randn(Kp, F, generator=g).to(dev).requires_grad_(True)
    t = torch.tensor([0.37], device=dev)
    gout = torch.randn(Kp, out, generator=g).to(dev)
    y_ref = _torch_motion_delta(net, kp, kf, t, xf, tf)
    if norm:
        y_ref = torch.cat([y_ref[:, :3], torch.nn.functional.normalize(y_ref[:, 3:7]), y_ref[:, 7:]], dim=1)
    y_ref.backward(gout)
    ref = {n: p.grad.clone() for n, p in net.named_parameters()}
    ref["kp"], ref["kf"] = (kp.grad.clone() if xf else None), (kf.grad.clone() if F else None)

    def run():
        for p in list(net.parameters()) + [kp, kf]:
            p.grad = None
        parts = keypoint_mlp(net, kp, kf, t, xf, tf, normalize_q=norm)
        assert parts[0].shape == (Kp, 3) and parts[1].shape == (Kp, 4) and (parts[2] is None) == (out == 7)
        assert all(p.is_contiguous() for p in parts if p is not None)
        y = torch.cat([p for p in parts if p is not None], dim=1)
        y.backward(gout)
        return y.detach(), [p.grad.clone() for p in net.parameters()]
    y, grads = run()
    assert float((y - y_ref).abs().max()) <= 1e-5 * max(1.0, float(y_ref.abs().max()))
    for (n, _), gr in zip(net.named_parameters(), grads):
        assert rel_err(gr, ref[n]) <= 1e-4, n
    if xf:
        assert rel_err(kp.grad, ref["kp"]) <= 1e-4
    else:
        assert float(kp.grad.abs().max()) == 0
    if F:
        assert rel_err(kf.grad, ref["kf"]) <= 1e-4
    # an unused output (no incoming gradient for delta_xyz) counts as zeros
    for p in list(net.parameters()) + [kp, kf]:
        p.grad = None
    keypoint_mlp(net, kp, kf, t, xf, tf, normalize_q=norm)[1].backward(gout[:, 3:7])
    part = [p.grad.clone() for p in net.parameters()]
    for p in list(net.parameters()) + [kp, kf]:
        p.grad = None
    y_ref2 = _torch_motion_delta(net, kp, kf, t, xf, tf)[:, 3:7]
    (torch.nn.functional.normalize(y_ref2) if norm else y_ref2).backward(gout[:, 3:7])
    for a, p in zip(part, net.parameters()):
        assert rel_err(a, p.grad) <= 1e-4
    # bit-reproducible: fixed summation order, no atomics
    y2, grads2 = run()
    assert count_diff(y, y2) == 0 and all(count_diff(a, b) == 0 for a, b in zip(grads, grads2))


def test_keypoint_mlp_golden_and_limits(gplib):
    """The reference's own Deformable_Field output (tests/golden/deform_mlp_pe.npz, made by importing
    scene/deformable_field.py) through the fused kernel: feature_dim = the whole 104-column input, no encodings."""
    from gaussianprediction_b200.deform import DeformableField, keypoint_mlp, keypoint_mlp_supported
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "deform_mlp_pe.npz"))
    net = DeformableField(104, 7, 4, 32)
    net.load_state_dict({k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd_")})
    net = net.cuda()
    x = torch.from_numpy(z["x"]).cuda()
    parts = keypoint_mlp(net, torch.zeros(x.shape[0], 3, device="cuda"), x, torch.zeros(1, device="cuda"), 0, 0)
    y = torch.cat(parts[:2], dim=1)
    assert float((y.cpu() - torch.from_numpy(z["y"])).abs().max()) <= 1e-5
    # the encoding columns: PE(pos) / PE(t) of the golden file pass through a 1-layer identity-like probe
    pos, tt = torch.from_numpy(z["pos"]).cuda(), torch.from_numpy(z["t"]).cuda()
    probe = DeformableField(60 + 12, 8, 1, 72).cuda()
    with torch.no_grad():
        probe.mlp[0].weight.copy_(torch.eye(72)); probe.mlp[0].bias.fill_(2.0)       # +2: keeps sin/cos above the ReLU
        probe.feature_to_deformation[0].weight.zero_(); probe.feature_to_deformation[0].bias.zero_()
        sel = [0, 1, 19, 20, 59, 60, 61, 71]
        for j, c in enumerate(sel):
            probe.feature_to_deformation[0].weight[j, c] = 1.0
    yp = torch.cat(keypoint_mlp(probe, pos, torch.zeros(pos.shape[0], 0, device="cuda"), tt, 10, 6), dim=1).cpu() - 2.0
    want = torch.cat([torch.from_numpy(z["pe_xyz"]), torch.from_numpy(z["pe_t"]).reshape(1, -1).repeat(pos.shape[0], 1)], 1)[:, sel]
    assert float((yp - want).abs().max()) <= 1e-6
    # shapes outside the kernel's limits are refused by the C ABI and routed to the torch modules by the model
    assert not keypoint_mlp_supported(DeformableField(104, 7, 4, 260).cuda(), 32, 10, 6)
    assert not keypoint_mlp_supported(DeformableField(102, 7, 4, 64).cuda(), 30, 10, 6)
    assert not keypoint_mlp_supported(DeformableField(104, 3, 4, 64).cuda(), 32, 10, 6)


def test_model_fused_mlp_equals_torch_mlp(gplib):
    """KeypointDeformation with the fused keypoint MLP vs the same model on the torch modules."""
    from gaussianprediction_b200.deform import KeypointDeformation, default_args
    N, Kp = 5000, 200
    gs, kp, kf, logits = scene(N, Kp, seed=41)
    args = default_args(feature_amplify=0.5, norm_rotation=True)
    torch.manual_seed(0)
    model = KeypointDeformation(args, gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf, logits=logits).cuda()
    t = torch.tensor([0.61], device="cuda")
    outs, grads = [], []
    for fused in (True, False):
        model.fused_mlp = fused
        for p in model.parameters():
            p.grad = None
        o = model(t, 50_000)
        sum((a * (i + 1)).sum() for i, a in enumerate(o)).backward()
        outs.append([a.detach().clone() for a in o])
        grads.append({n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None})
    for a, b in zip(*outs):
        assert float((a - b).abs().max()) <= 2e-5
    assert grads[0].keys() == grads[1].keys()
    for n in grads[0]:
        assert rel_err(grads[0][n], grads[1][n]) <= 1e-3, n


@pytest.mark.parametrize("N", [6000, 6001])      # 6001: not a multiple of 4 (the reducer pads every view to 16 bytes)
def test_two_stage_step_with_grad_sinks_matches_eager(gplib, N):
    """pipeline.TrainStep with a reducer: two graphs with the gradient exchange issued between them, gradients written
    by the kernels straight into the reducer's flat buffer.  One-rank NCCL group: the sums are the local gradients."""
    import torch.distributed as dist
    from gaussianprediction_b200 import synthetic as syn
    from gaussianprediction_b200.deform import KeypointDeformation, default_args
    from gaussianprediction_b200.parallel import FlatGradAllReduce
    from gaussianprediction_b200.pipeline import TrainStep
    from gaussianprediction_b200.renderer import SceneModel, render
    own_group = not dist.is_initialized()
    if own_group:
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % (32500 + os.getpid() % 2000), rank=0, world_size=1)
    try:
        Kp, W, H = 96, 128, 96
        gs, kp, kf, logits = scene(N, Kp, seed=51)
        args = default_args(feature_amplify=0.5, norm_rotation=True)
        torch.manual_seed(0)
        model = KeypointDeformation(args, gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf, logits=logits)
        pc = SceneModel(model, gs.features_dc, gs.features_rest).cuda()
        cams = syn.camera_ring(2, W, H, seed=5)
        for c in cams:
            for k in ("world_view_transform", "full_proj_transform", "camera_center"):
                setattr(c, k, getattr(c, k).cuda())
        bg = torch.zeros(3, device="cuda")
        g = torch.Generator().manual_seed(1)
        targets = [(torch.randn(3, H, W, generator=g) * 1e-2).cuda() for _ in cams]
        times = [torch.tensor([0.2 + 0.3 * i], device="cuda") for i in range(len(cams))]
        with_grad = [p for n, p in pc.named_parameters() if not n.endswith("motion_feature")]
        red = FlatGradAllReduce(with_grad, first=[pc._features_dc, pc._features_rest],
                                second=[model._xyz, model._rotation, model._scaling, model._opacity, model.weights_model.logits])
        step = TrainStep(pc, cams[0], capacity=400_000, bg=bg, reducer=red)
        assert step.two_stage and len(step.stage_graphs) == 3 and len(red.buckets()) == 3
        assert step.packed == 0            # every gradient reached the flat buffer through a sink: nothing filled or copied
        for cam, t, tgt in zip(cams, times, targets):
            loss_g = float(step(cam, t, tgt))
            assert not step.overflowed()[1]
            grads_g = {n: p.grad.clone() for n, p in pc.named_parameters() if p.grad is not None}
            assert all(p.grad.data_ptr() == v.data_ptr() for p, v in zip(red.params, red.views))
            vs_g = step.viewspace_grad.clone()
            for p in pc.parameters():
                p.grad = None
            pkg = render(cam, pc, None, bg, time=t, it=50_000)
            loss_e = (pkg["render"] * tgt).sum()
            loss_e.backward()
            assert abs(loss_g - float(loss_e)) <= 1e-6 * max(1.0, abs(float(loss_e)))
            for n, p in pc.named_parameters():
                if p.grad is not None:
                    assert rel_err(grads_g[n], p.grad) <= 1e-3, n
            assert rel_err(vs_g, pkg["viewspace_points"].grad) <= 1e-3
    finally:
        if own_group:
            dist.destroy_process_group()


def test_batch_step_accumulates_samples(gplib):
    """pipeline.BatchStep (the reference's --batch loop, train.py:113-133): after a step over three samples ``p.grad`` is
    the SUM of the three per-sample gradients, the loss the sum of the losses, ``radii`` the element-wise maximum --
    from device tensors and from pinned host tensors (prefetch path)."""
    from gaussianprediction_b200 import synthetic as syn
    from gaussianprediction_b200.deform import KeypointDeformation, default_args
    from gaussianprediction_b200.parallel import FlatGradAllReduce
    from gaussianprediction_b200.pipeline import BatchStep
    from gaussianprediction_b200.renderer import SceneModel, render
    N, Kp, W, H = 6001, 96, 128, 96
    gs, kp, kf, logits = scene(N, Kp, seed=52)
    args = default_args(feature_amplify=0.5, norm_rotation=True)
    torch.manual_seed(0)
    model = KeypointDeformation(args, gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf, logits=logits)
    pc = SceneModel(model, gs.features_dc, gs.features_rest).cuda()
    cams = syn.camera_ring(3, W, H, seed=6)
    host = []
    g = torch.Generator().manual_seed(2)
    for i, c in enumerate(cams):
        hc = type(c)(**vars(c))
        for k in ("world_view_transform", "full_proj_transform", "camera_center"):
            setattr(hc, k, getattr(c, k).pin_memory())
            setattr(c, k, getattr(c, k).cuda())
        host.append((hc, torch.tensor([0.2 + 0.3 * i]).pin_memory(), (torch.randn(3, H, W, generator=g) * 1e-2).pin_memory()))
    dev_samples = [(c, h[1].cuda(), h[2].cuda()) for c, h in zip(cams, host)]
    bg = torch.zeros(3, device="cuda")
    with_grad = [p for n, p in pc.named_parameters() if not n.endswith("motion_feature")]
    # eager truth: autograd accumulation over the three samples
    want_loss, want_radii = 0.0, None
    for cam, t, tgt in dev_samples:
        pkg = render(cam, pc, None, bg, time=t, it=50_000)
        loss = (pkg["render"] * tgt).sum()
        loss.backward()
        want_loss += float(loss)
        want_radii = pkg["radii"] if want_radii is None else torch.maximum(want_radii, pkg["radii"])
    want = {n: p.grad.clone() for n, p in pc.named_parameters() if p.grad is not None}
    for p in pc.parameters():
        p.grad = None
    del pkg, loss
    batch = BatchStep(pc, cams[0], 400_000, lambda: FlatGradAllReduce(with_grad), 3, bg=bg)
    # samples 2.. add their SH and per-Gaussian gradients to the running sum inside the kernels (no scratch copy)
    assert batch.rest.in_place == 7, batch.rest.in_place
    for mode in ("device", "host", "host"):
        if mode == "device":
            loss = batch.step(dev_samples)
        else:
            loss = batch.step_host(host, next_first=host[0])
        torch.cuda.synchronize()
        assert not batch.overflowed()[1]
        assert abs(float(loss) - want_loss) <= 1e-5 * max(1.0, abs(want_loss)), mode
        assert torch.equal(batch.radii, want_radii), mode
        for n, p in pc.named_parameters():
            if n in want:
                assert rel_err(p.grad, want[n]) <= 1e-3, (mode, n, rel_err(p.grad, want[n]))


def test_overflow_policy_nan_loss_then_grow(gplib):
    """A frame with more tile instances than the captured capacity: defined outputs (empty image, zero gradients), NaN
    loss on the device (no extra sync), and ``step_checked`` re-captures with more room and reruns the step."""
    from gaussianprediction_b200 import synthetic as syn
    from gaussianprediction_b200.deform import KeypointDeformation, default_args
    from gaussianprediction_b200.parallel import FlatGradAllReduce
    from gaussianprediction_b200.pipeline import BatchStep
    from gaussianprediction_b200.renderer import SceneModel, render
    N, Kp, W, H = 6000, 96, 128, 96
    gs, kp, kf, logits = scene(N, Kp, seed=53)
    args = default_args(feature_amplify=0.5, norm_rotation=True)
    torch.manual_seed(0)
    model = KeypointDeformation(args, gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf, logits=logits)
    pc = SceneModel(model, gs.features_dc, gs.features_rest).cuda()
    cam = syn.camera_ring(1, W, H, seed=7)[0]
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).cuda())
    bg = torch.tensor([0.25, 0.5, 0.75], device="cuda")
    t = torch.tensor([0.4], device="cuda")
    tgt = (torch.randn(3, H, W, generator=torch.Generator().manual_seed(3)) * 1e-2).cuda()
    pkg = render(cam, pc, None, bg, time=t, it=50_000)
    want = float((pkg["render"] * tgt).sum())
    del pkg
    with_grad = [p for n, p in pc.named_parameters() if not n.endswith("motion_feature")]
    batch = BatchStep(pc, cam, 1000, lambda: FlatGradAllReduce(with_grad), 1, bg=bg)      # far too small on purpose
    loss = batch.step([(cam, t, tgt)])
    torch.cuda.synchronize()
    R, over = batch.overflowed()
    assert over and R > 1000 and bool(torch.isnan(loss))
    img = batch.first.pkg["render"]
    assert float((img - bg.view(3, 1, 1)).abs().max()) == 0.0                # the empty image, not stale memory
    assert int((batch.first.pkg["tidx"] != -1).sum()) == 0 and float((batch.first.pkg["depth"] - 15.0).abs().max()) == 0.0
    assert all(float(p.grad.abs().max()) == 0.0 for p in with_grad if p.grad is not None)
    loss = batch.step_checked([(cam, t, tgt)])
    torch.cuda.synchronize()
    assert not batch.overflowed()[1] and batch.first.capacity >= R
    assert abs(float(loss) - want) <= 1e-5 * max(1.0, abs(want))


def test_full_size_deformation_properties_1m(gplib):
    """BASELINE.json's large configs (C4/C5: 1-2 M Gaussians) are too big for the brute-force oracle, so the
    deformation stage is checked there through size-independent properties: the neighbour lists are sorted, unique and
    in range; a random sample of queries agrees with the oracle bit for bit; the order / hint arguments (speed hints)
    do not change the result; the position blend is linear in the keypoint translations; the quaternions come out
    normalised."""
    from gaussianprediction_b200.deform import keypoint_blend, knn_keypoints, spatial_order
    from oracle import deform_oracle as do
    N, Kp, K = 1_000_000, 512, 6
    gs, kp, kf, logits = scene(N, Kp, seed=77)
    dev = "cuda"
    xyz, mf, kpd, kfd = gs.xyz.to(dev), gs.motion_feature.to(dev), kp.to(dev), kf.to(dev)
    idx64, idx32, dists = knn_keypoints(xyz, mf, kpd, kfd, K, "hybird", 0.5, return_dists=True)
    d, i = to_np(dists), to_np(idx32)
    assert (np.diff(d, axis=1) >= 0).all() and i.min() >= 0 and i.max() < Kp
    assert (np.sort(i, axis=1)[:, 1:] != np.sort(i, axis=1)[:, :-1]).all()            # no keypoint twice in a row
    assert count_diff(to_np(idx64), i.astype(np.int64)) == 0
    sel = np.random.default_rng(0).choice(N, 3000, replace=False)
    p1, p2 = do.hybrid_points(gs.xyz[sel], gs.motion_feature[sel], kp, kf, 0.5)
    ok = do.knn_rel_gaps(p1.numpy(), p2.numpy(), K) >= 1e-6
    od, oi = do.knn(p1.numpy(), p2.numpy(), K, fma=True)
    assert ok.mean() > 0.995 and count_diff(i[sel][ok], oi[ok].astype(np.int32)) == 0
    assert np.abs(d[sel] - od)[ok].max() <= 1e-6 * max(float(od.max()), 1.0)
    # speed hints change nothing
    order = spatial_order(xyz)
    _, idx_b, dist_b = knn_keypoints(xyz, mf, kpd, kfd, K, "hybird", 0.5, return_dists=True, order=order, hint=idx32.clone())
    assert count_diff(idx_b, idx32) == 0 and count_diff(dist_b, dists) == 0
    # blend: linear in the keypoint translations, unit quaternions
    g = torch.Generator().manual_seed(3)
    kdx = (torch.randn(Kp, 3, generator=g) * 0.05).to(dev)
    kdq = torch.nn.functional.normalize(torch.randn(Kp, 4, generator=g)).to(dev)
    args = (logits.to(dev), idx32)
    rest = (xyz, gs.rotation.to(dev), gs.scaling.to(dev), gs.opacity.to(dev))
    with torch.no_grad():
        x1, q1, s1, o1 = keypoint_blend(*args, kdx, kdq, *rest)
        x2, q2, _, _ = keypoint_blend(*args, 2.0 * kdx, kdq, *rest)
    assert float(((x2 - xyz) - 2.0 * (x1 - xyz)).abs().max()) <= 2e-6
    assert float((q1.norm(dim=1) - 1).abs().max()) <= 1e-6 and count_diff(q1, q2) == 0
    assert float((s1 - gs.scaling.to(dev).exp()).abs().max()) <= 1e-6 * float(s1.max())
    assert float((o1 - torch.sigmoid(gs.opacity.to(dev))).abs().max()) <= 1e-6


@pytest.mark.parametrize("tag", ["A", "A2", "B", "C", "D", "D2", "E", "F", "W"])
def test_model_forward_vs_reference_source(gplib, tag, monkeypatch):
    """KeypointDeformation on the GPU == the reference's own ``GaussianModel.forward`` source, executed on the CPU by
    tests/golden/make_golden.py:make_chain (blend_chain.npz): all three stages, both step-opacity types, keypoint noise,
    ``return_weights``, the side-effect attributes, the search result, and every autograd gradient.
    Tolerances (north_star): outputs 1e-5 abs (fp32, different summation order), gradients 1e-3 relative."""
    from helpers import CHAIN_CASES, CHAIN_LEAVES, chain_args, chain_state_dict
    from gaussianprediction_b200.deform import KeypointDeformation
    g = np.load(os.path.join(GOLD, "blend_chain.npz"))
    it, flags = CHAIN_CASES[tag]
    args = chain_args(flags)
    leaf = {k: torch.tensor(g[k]) for k in CHAIN_LEAVES}
    model = KeypointDeformation(args, leaf["xyz"], leaf["rotation"], leaf["scaling"], leaf["opacity"], leaf["motion_feature"],
                                leaf["kp"], leaf["kf"], logits=leaf["logits"])
    if args.step_opacity:
        model.opacity_thres.data.copy_(leaf["thres"])
    model.df_model.load_state_dict(chain_state_dict(g, 8 if args.step_opacity else 7))
    model = model.cuda()
    if (tag + "_noise") in g.files:      # the reference's randn draw (CPU generator) injected in place of the CUDA one
        noise = torch.tensor(g[tag + "_noise"]).cuda()
        monkeypatch.setattr(torch, "randn_like", lambda x, *a, **k: noise)
    t = torch.tensor([float(g["t"])], device="cuda")
    res = model(t, it, return_weights=(tag == "W"))
    for n, v in zip(("xyz_t", "q_t", "s_t", "o_t", "weights_xyz", "weights_r"), res):
        assert float(np.abs(to_np(v) - g[f"{tag}_{n}"]).max()) <= 1e-5, n
    if it > 30000:
        assert np.array_equal(to_np(model.nearest_mask), g[tag + "_nearest_mask"])
        assert float(np.abs(to_np(model.weights_sum) - g[tag + "_weights_sum"]).max()) <= 1e-5
    if it >= 1000:
        assert float(np.abs(to_np(model.kpts_xyz_motion) - g[tag + "_kpts_xyz_motion"]).max()) <= 1e-5
        assert float(np.abs(to_np(model.kpts_rotation_motion) - g[tag + "_kpts_rotation_motion"]).max()) <= 1e-5
    if (tag + "_lifecycle_opacity") in g.files:
        assert float(np.abs(to_np(model.lifecycle_opacity) - g[tag + "_lifecycle_opacity"]).max()) <= 1e-5
    else:
        assert model.lifecycle_opacity is None
    cot = [torch.tensor(g[k]).cuda() for k in ("cot_xyz", "cot_q", "cot_s", "cot_o")]
    sum((a * b).sum() for a, b in zip(res[:4], cot)).backward()
    names = dict(xyz=model._xyz, rotation=model._rotation, scaling=model._scaling, opacity=model._opacity,
                 motion_feature=model.motion_feature, kp=model.super_gaussians, kf=model.super_gaussians_feature,
                 logits=model.weights_model.logits)
    if args.step_opacity:
        names["thres"] = model.opacity_thres
    names.update({"net_" + k: p for k, p in model.df_model.named_parameters()})
    for n, p in names.items():
        key = f"{tag}_g_{n}"
        if key in g.files:
            assert p.grad is not None, n
            assert rel_err(p.grad, g[key]) <= 1e-3, (n, rel_err(p.grad, g[key]))
        else:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, n


def test_knn_c2_size_with_order_and_hint_vs_reference(gplib):
    """The benchmarked configuration of the search -- 100k x 512, D = 35, Morton order + previous-frame hint, i.e. the
    pruned path bench.py times -- against the reference's CUDA FRNN (bit-exact indices and squared distances)."""
    from gaussianprediction_b200.deform import knn_keypoints, spatial_order
    from oracle import deform_oracle as do
    from oracle import ref_cuda
    if not (ref_cuda.available("ref_frnn") and ref_cuda.available("ref_prefix_sum")):
        pytest.skip("oracle/_ref/ref_frnn.so not built")
    gs, kp, kf, _ = scene(100_000, 512, seed=2024)
    amplify = 0.5
    p1, p2 = do.hybrid_points(gs.xyz, gs.motion_feature, kp, kf, amplify)
    rd, ri = ref_cuda.frnn_grid_points(p1[None].cuda().contiguous(), p2[None].cuda().contiguous(), 6, 1e8)
    rd, ri = to_np(rd[0]), to_np(ri[0])
    xyz, mf, kpc, kfc = gs.xyz.cuda(), gs.motion_feature.cuda(), kp.cuda(), kf.cuda()
    order = spatial_order(xyz)
    # first call: no hint; second call: the first result as hint (the steady state of the frame loop); third: a hint
    # that is valid but useless (neighbours of a shifted cloud), which must not change the result either
    _, idx_a, d_a = knn_keypoints(xyz, mf, kpc, kfc, 6, "hybird", amplify, return_dists=True, order=order)
    hint = idx_a.clone()
    _, idx_b, d_b = knn_keypoints(xyz, mf, kpc, kfc, 6, "hybird", amplify, return_dists=True, order=order, hint=hint)
    bad = torch.roll(idx_a, 7919, 0).contiguous()
    _, idx_c, d_c = knn_keypoints(xyz, mf, kpc, kfc, 6, "hybird", amplify, return_dists=True, order=order, hint=bad)
    # tie policy (SURVEY.md par. 8c): rows whose K+1 nearest are closer than 1e-6 relative are compared as sets
    for idx, d in ((idx_a, d_a), (idx_b, d_b), (idx_c, d_c)):
        idx, d = to_np(idx), to_np(d)
        assert count_diff(np.sort(d, 1), np.sort(rd, 1)) == 0            # distances: bit-identical multiset per row
        rows = (idx.astype(np.int64) != ri).any(1)
        assert rows.mean() < 1e-4                                         # index order may differ only on near-ties
        for r in np.nonzero(rows)[0]:
            gap = np.diff(rd[r]) / np.maximum(rd[r][1:], 1e-30)
            assert gap.min() < 1e-6, "index mismatch on a tie-free row"


def test_knn_pruning_adversarial_lattice(gplib):
    """Queries and keypoints on lattices: keypoints sit exactly ON the faces of the warps' query bounding boxes and
    exact distance ties are everywhere -- the worst case for the box-bound pruning (csrc/knn.cu step (b): a keypoint is
    dropped only if its shrunk box distance is strictly greater than the K-th bound).  Must equal the exhaustive
    oracle, ties resolved to the smaller index, with and without order / hint."""
    from gaussianprediction_b200.deform import knn_keypoints, spatial_order
    from oracle import deform_oracle as do
    ax = torch.arange(17, dtype=torch.float32) / 16.0
    q = torch.stack(torch.meshgrid(ax, ax, ax, indexing="ij"), -1).reshape(-1, 3)            # 4913 lattice queries
    bx = torch.arange(5, dtype=torch.float32) / 4.0
    k = torch.stack(torch.meshgrid(bx, bx, bx, indexing="ij"), -1).reshape(-1, 3)            # 125 lattice keypoints
    g = torch.Generator().manual_seed(3)
    q = q[torch.randperm(q.shape[0], generator=g)].contiguous()
    for D_feat in (0, 32):
        mf = torch.zeros(q.shape[0], 32)
        kf = torch.zeros(k.shape[0], 32)
        mode = "3D" if D_feat == 0 else "hybird"
        p1 = q if D_feat == 0 else torch.cat([q, mf], 1)
        p2 = k if D_feat == 0 else torch.cat([k, kf], 1)
        od, oi = do.knn(p1.numpy(), p2.numpy(), 6, r=1e3 if D_feat == 0 else 1e8, fma=True)
        qc, kc = q.cuda(), k.cuda()
        order = spatial_order(qc)
        for o, h in ((None, None), (order, None), (order, "self")):
            hint = None
            if h == "self":
                hint = knn_keypoints(qc, mf.cuda(), kc, kf.cuda(), 6, mode, 1.0, order=order)[1].clone()
            i64, _, d = knn_keypoints(qc, mf.cuda(), kc, kf.cuda(), 6, mode, 1.0, return_dists=True, order=o, hint=hint)
            assert count_diff(to_np(d), od) == 0, (mode, o is not None, h)
            assert count_diff(to_np(i64), oi) == 0, (mode, o is not None, h)


def test_blend_backward_adds_to_a_running_sum(gplib):
    """gp_blend_backward_accumulate (accumulating gradient sinks): the per-Gaussian gradients of a second backward are
    added to the destination; keypoint gradients are returned as usual."""
    from gaussianprediction_b200._sinks import grad_sinks
    from gaussianprediction_b200.deform import keypoint_blend, knn_keypoints
    N, Kp = 5003, 64
    gs, kp, kf, logits = scene(N, Kp, seed=9)
    g = torch.Generator().manual_seed(4)
    leaf = dict(logits=logits, kp_dxyz=0.2 * torch.randn(Kp, 3, generator=g),
                kp_dq=torch.nn.functional.normalize(torch.randn(Kp, 4, generator=g)), xyz=gs.xyz, rot=gs.rotation,
                sc=gs.scaling, op=gs.opacity)
    leaf = {k: v.clone().cuda().requires_grad_(True) for k, v in leaf.items()}
    _, idx32, _ = knn_keypoints(gs.xyz.cuda(), gs.motion_feature.cuda(), kp.cuda(), kf.cuda(), 6, "hybird", 0.5)
    per_gaussian = ("logits", "xyz", "rot", "sc", "op")
    gens = None

    def run(scale, sums=None):
        nonlocal gens
        pairs = [(leaf[k], sums[k]) for k in per_gaussian] if sums is not None else ()
        with grad_sinks((), accumulate=pairs):
            outs = keypoint_blend(leaf["logits"], idx32, leaf["kp_dxyz"], leaf["kp_dq"], leaf["xyz"], leaf["rot"],
                                  leaf["sc"], leaf["op"])
        if gens is None:
            gens = [torch.randn(o.shape, generator=g).cuda() for o in outs]
        for v in leaf.values():
            v.grad = None
        sum((o * w * scale).sum() for o, w in zip(outs, gens)).backward()
        return {k: v.grad for k, v in leaf.items()}
    g1 = {k: v.clone() for k, v in run(1.0).items()}
    sums = {k: g1[k].clone() for k in per_gaussian}
    g2 = run(0.5, sums)
    for k in per_gaussian:
        assert g2[k].data_ptr() == sums[k].data_ptr()
        assert float(g1[k].abs().sum()) > 0 and rel_err(sums[k], 1.5 * g1[k]) <= 1e-5, k
    for k in ("kp_dxyz", "kp_dq"):
        assert rel_err(g2[k], 0.5 * g1[k]) <= 1e-4, k


def test_blend_entries_refuse_misaligned_quaternion_rows(gplib):
    """rot / q_t / g_rot / g_q_t move as 128-bit rows: a base pointer off a 16-byte boundary is an error, not a fault."""
    from gaussianprediction_b200 import _lib
    from gaussianprediction_b200._lib import ptr
    L = gplib
    N, Kp, K = 64, 8, 6
    dev = "cuda"
    f = dict(dtype=torch.float32, device=dev)
    logits, idx = torch.zeros(N, 2 * K, **f), torch.zeros(N, K, dtype=torch.int32, device=dev)
    kpx, kpq = torch.zeros(Kp, 3, **f), torch.zeros(Kp, 4, **f)
    xyz, sc, op = torch.zeros(N, 3, **f), torch.zeros(N, 3, **f), torch.zeros(N, 1, **f)
    pool = torch.ones(4 * N + 4, **f)
    rot_ok, rot_bad = pool[:4 * N].view(N, 4), pool[1:4 * N + 1].view(N, 4)          # the second starts 4 bytes in
    outs = [torch.empty(N, 3, **f), torch.empty(N, 4, **f), torch.empty(N, 3, **f), torch.empty(N, 1, **f)]
    s = _lib.stream_ptr()
    call = lambda rot: L.gp_blend_forward(N, Kp, K, ptr(logits), ptr(idx), None, ptr(kpx), ptr(kpq), ptr(xyz), ptr(rot),
                                          ptr(sc), ptr(op), ptr(outs[0]), ptr(outs[1]), ptr(outs[2]), ptr(outs[3]), s)
    assert call(rot_ok) == 0
    assert call(rot_bad) != 0
    torch.cuda.synchronize()

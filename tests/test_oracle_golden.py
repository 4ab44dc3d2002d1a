"""CPU tests: the oracle (oracle/raster_oracle.c, oracle/deform_oracle.py) against golden vectors produced by the
REFERENCE itself (tests/golden/make_golden.py: reference CUDA kernels on a B200, reference CPU kNN, reference torch
modules).  This is what pins the oracle; the GPU tests then compare the CUDA library with both."""
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12))


@pytest.fixture(scope="module")
def raster_gold():
    return np.load(os.path.join(GOLD, "raster_small.npz"))


@pytest.fixture(scope="module")
def oracle_fwd(raster_gold):
    from oracle import raster_oracle as ro
    g = raster_gold
    return ro.forward(g["in_means3D"], g["in_opacities"], g["in_bg"], g["in_view"], g["in_proj"], g["in_campos"],
                      int(g["in_W"]), int(g["in_H"]), float(g["in_tanfovx"]), float(g["in_tanfovy"]), shs=g["in_shs"],
                      scales=g["in_scales"], rotations=g["in_rotations"], sh_degree=int(g["in_sh_degree"]))


def test_raster_oracle_forward_vs_reference_cuda(raster_gold, oracle_fwd):
    g, o = raster_gold, oracle_fwd
    # integer / index outputs: the CPU evaluation rounds differently from the GPU in the last ulp, so an isolated
    # flip is possible in principle; on this fixture there is none.
    assert (o["radii"] != g["radii"]).sum() == 0
    assert (o["tiles_touched"] != g["tiles_touched"].astype(np.uint32)).sum() == 0
    assert o["num_rendered"] == int(g["num_rendered"])
    assert (o["point_list"] != g["point_list"].astype(np.uint32)).sum() == 0
    assert (o["ranges"] != g["ranges"].astype(np.uint32)).sum() == 0
    assert ((o["point_list_keys"] >> 32) != (g["point_list_keys"].astype(np.uint64) >> 32)).sum() == 0   # tile ids
    vis = g["radii"] > 0
    for k, tol in (("depths", 1e-6), ("means2D", 1e-4), ("conic_opacity", 1e-5), ("cov3D", 1e-6), ("rgb", 1e-6)):
        assert rel(o[k][vis], g[k][vis]) <= tol, k
    assert (o["clamped"][vis] != g["clamped"][vis]).sum() == 0
    assert (o["n_contrib"] != g["n_contrib"].astype(np.uint32)).mean() < 2e-3
    assert (o["tidx"] != g["tidx"]).mean() < 2e-3
    same = o["tidx"] == g["tidx"]
    dcol = np.abs(o["color"] - g["color"])
    assert (dcol > 1e-4).mean() < 5e-3 and dcol.max() <= 1e-2        # north_star tolerance 1e-4, bar isolated alpha flips
    assert np.abs(o["depth"] - g["depth"])[same].max() <= 1e-4
    assert np.abs(o["final_T"] - g["final_T"]).max() <= 1e-2 and np.median(np.abs(o["final_T"] - g["final_T"])) <= 1e-6


def test_raster_oracle_backward_vs_reference_cuda(raster_gold, oracle_fwd):
    from oracle import raster_oracle as ro
    g = raster_gold
    og = ro.backward(oracle_fwd, g["grad_in"])
    for k in ("means2D", "opacity", "means3D", "sh", "scales", "rotations"):
        assert rel(og[k], g["g_" + k]) <= 1e-3, (k, rel(og[k], g["g_" + k]))       # north_star: 1e-3 relative
    inv = ~(g["radii"] > 0)
    assert np.abs(og["means3D"][inv]).sum() == 0


def test_raster_oracle_edge_cases():
    from oracle import raster_oracle as ro
    from helpers import raster_case, to_np
    # every Gaussian behind the camera -> background only, R = 0
    c = raster_case(N=50, W=40, H=24, seed=1, bg=(0.25, 0.5, 0.75))
    n = {k: (to_np(v) if torch.is_tensor(v) else v) for k, v in c.items()}
    means = n["means3D"] + 100.0 * (n["campos"] - 0)       # push far behind/away along the camera direction
    o = ro.forward(means, n["opacities"], n["bg"], n["view"], n["proj"], n["campos"], n["W"], n["H"], n["tanfovx"],
                   n["tanfovy"], shs=n["shs"], scales=n["scales"], rotations=n["rotations"])
    assert o["num_rendered"] == 0 and (o["radii"] == 0).all() and (o["tidx"] == -1).all()
    assert np.allclose(o["color"][:, 3, 5], [0.25, 0.5, 0.75]) and (o["depth"] == 15.0).all()   # D = 15 sentinel
    # ragged image (not a multiple of 16) and precomputed colours
    c = raster_case(N=300, W=37, H=21, seed=2)
    n = {k: (to_np(v) if torch.is_tensor(v) else v) for k, v in c.items()}
    o1 = ro.forward(n["means3D"], n["opacities"], n["bg"], n["view"], n["proj"], n["campos"], n["W"], n["H"], n["tanfovx"],
                    n["tanfovy"], shs=n["shs"], scales=n["scales"], rotations=n["rotations"])
    o2 = ro.forward(n["means3D"], n["opacities"], n["bg"], n["view"], n["proj"], n["campos"], n["W"], n["H"], n["tanfovx"],
                    n["tanfovy"], colors_precomp=o1["rgb"], cov3D_precomp=o1["cov3D"])
    assert np.array_equal(o1["color"], o2["color"]) and np.array_equal(o1["radii"], o2["radii"])
    # sortedness of the instance list: keys ascending, ties by index
    k = o1["point_list_keys"].astype(np.uint64)
    assert (k[1:] >= k[:-1]).all()
    eq = k[1:] == k[:-1]
    assert (o1["point_list"][1:][eq] > o1["point_list"][:-1][eq]).all()


def test_deform_oracle_vs_reference_modules():
    from oracle import deform_oracle as do
    m = np.load(os.path.join(GOLD, "deform_mlp_pe.npz"))
    sd = {k[3:]: torch.tensor(m[k]) for k in m.files if k.startswith("sd_")}
    assert float((do.mlp_forward(sd, torch.tensor(m["x"])) - torch.tensor(m["y"])).abs().max()) <= 1e-6
    assert float((do.positional_encoding(torch.tensor(m["pos"]), 10) - torch.tensor(m["pe_xyz"])).abs().max()) == 0
    assert float((do.positional_encoding(torch.tensor(m["t"]), 6) - torch.tensor(m["pe_t"])).abs().max()) == 0
    # the product's torch-side restatements of the same two pieces
    from gaussianprediction_b200.deform import DeformableField, positional_encoding
    net = DeformableField(104, 7, d=4, w=32)
    net.load_state_dict(sd)                                  # reference state_dict keys load unchanged
    assert float((net(torch.tensor(m["x"])) - torch.tensor(m["y"])).abs().max()) <= 1e-6
    assert float((positional_encoding(torch.tensor(m["pos"]), 10) - torch.tensor(m["pe_xyz"])).abs().max()) == 0


def test_knn_oracle_vs_reference_cpu_and_cuda():
    from oracle import deform_oracle as do
    c = np.load(os.path.join(GOLD, "knn_cpu.npz"))
    d, i = do.knn(c["p1"], c["p2"], 6, fma=False)           # reference CPU brute force: bit-exact
    assert (d.view(np.uint32) != c["d35"].view(np.uint32)).sum() == 0 and (i != c["i35"]).sum() == 0
    d3, i3 = do.knn(c["p1"][:, :3], c["p2"][:, :3], 6, r=1e3, fma=False)
    assert (d3.view(np.uint32) != c["d3"].view(np.uint32)).sum() == 0 and (i3 != c["i3"]).sum() == 0
    g = np.load(os.path.join(GOLD, "knn_gpu.npz"))           # reference CUDA grid path (FMA contraction)
    d, i = do.knn(g["p1"], g["p2"], 6, fma=True)
    assert (i != g["i35"]).sum() == 0
    assert (d.view(np.uint32) != g["d35"].view(np.uint32)).mean() < 1e-3      # float64-emulated FMA: double rounding is rare
    assert do.knn_min_rel_gap(g["p1"], g["p2"], 6) >= 1e-6                     # the fixture is tie-free


def test_dist2_oracle_vs_reference_cuda():
    from oracle import deform_oracle as do
    g = np.load(os.path.join(GOLD, "dist2.npz"))
    mine = do.dist2_knn3(g["points"])
    assert (mine.view(np.uint32) != g["dist2"].view(np.uint32)).mean() < 1e-3
    assert rel(mine, g["dist2"]) <= 1e-6


def test_blend_oracle_dense_vs_sparse_identity():
    """Size-independent property: the dense scatter+matmul form equals the explicit K-term sum."""
    from oracle import deform_oracle as do
    torch.manual_seed(0)
    N, Kp, K = 200, 40, 6
    logits = torch.randn(N, 2 * K)
    idx = torch.stack([torch.randperm(Kp)[:K] for _ in range(N)])
    dq, dx = torch.randn(Kp, 4), torch.randn(Kp, 3)
    wx, wr = do.fill_nearest(logits, idx, Kp, K)
    sx = (torch.softmax(logits[:, :K], -1)[..., None] * dx[idx]).sum(1)
    sr = (torch.softmax(logits[:, K:], -1)[..., None] * dq[idx]).sum(1)
    assert float((wx @ dx - sx).abs().max()) <= 1e-5 and float((wr @ dq - sr).abs().max()) <= 1e-5
    assert float((wx.sum(1) - 1).abs().max()) <= 1e-5


def test_loss_oracle_vs_reference_functions():
    """oracle/loss_oracle.py against the reference's own l1_loss / ssim and their autograd gradient (loss.npz)."""
    from oracle import loss_oracle as lo
    g = np.load(os.path.join(GOLD, "loss.npz"))
    img = torch.tensor(g["img"]).requires_grad_(True)
    gt = torch.tensor(g["gt"])
    assert abs(float(lo.l1_loss(img, gt)) - float(g["l1"])) <= 1e-7
    assert abs(float(lo.ssim(img, gt)) - float(g["ssim"])) <= 1e-6
    loss = lo.image_loss(img, gt, 0.2)
    loss.backward()
    assert abs(float(loss) - float(g["loss"])) <= 1e-6
    assert rel(img.grad.numpy(), g["grad"]) <= 1e-5


def test_adam_oracle_vs_torch_optim_adam():
    """oracle/optim_oracle.adam_step against torch.optim.Adam (what the reference's optimizer.step() is), CPU."""
    import torch
    from oracle import optim_oracle as oo
    torch.manual_seed(3)
    ps = [torch.nn.Parameter(torch.randn(37, 3)), torch.nn.Parameter(torch.randn(11))]
    lrs = [1.6e-4, 2.5e-3]
    opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(ps, lrs)], lr=0.0, eps=1e-15)
    mine = [dict(p=p.detach().numpy().copy(), m=np.zeros(p.shape, np.float32), v=np.zeros(p.shape, np.float32)) for p in ps]
    for step in range(1, 6):
        for p, st, lr in zip(ps, mine, lrs):
            g = torch.randn(p.shape) * (10.0 ** (-step))
            p.grad = g.clone()
            oo.adam_step(st["p"], g.numpy(), st["m"], st["v"], step, lr)
        opt.step()
    for p, st in zip(ps, mine):
        assert np.abs(st["p"] - p.detach().numpy()).max() <= 2e-6 * max(1.0, float(p.detach().abs().max()))
        assert np.allclose(st["m"], opt.state[p]["exp_avg"].numpy(), rtol=1e-5, atol=1e-12)
        assert np.allclose(st["v"], opt.state[p]["exp_avg_sq"].numpy(), rtol=1e-5, atol=1e-20)


def test_densification_stats_oracle_vs_reference_expressions():
    """oracle/optim_oracle.densification_stats against train.py:166-167 + scene/gaussian_model.py:756-760 restated."""
    import torch
    from oracle import optim_oracle as oo
    g = torch.Generator().manual_seed(5)
    P = 500
    radii = (torch.rand(P, generator=g) * 30).int() * (torch.rand(P, generator=g) > 0.3).int()
    vgrad = torch.randn(P, 3, generator=g) * 1e-3
    max_r = torch.rand(P, generator=g) * 20
    accum, denom, amax = torch.rand(P, 1, generator=g), torch.rand(P, 1, generator=g).round(), torch.rand(P, 1, generator=g) * 1e-3
    o = [t.numpy().copy() for t in (max_r, accum, denom, amax)]
    oo.densification_stats(vgrad.numpy(), radii.numpy(), *o)
    vis = radii > 0
    max_r[vis] = torch.max(max_r[vis], radii[vis])                        # train.py:166
    grad = torch.norm(vgrad[vis, :2], dim=-1, keepdim=True)                # gaussian_model.py:757
    accum[vis] += grad
    denom[vis] += 1
    amax[vis] = torch.where(grad > amax[vis], grad, amax[vis])
    for a, b in zip(o, (max_r, accum, denom, amax)):
        assert np.allclose(a, b.numpy(), rtol=1e-6, atol=0)


# ---- blend / quaternion chain: the oracle against the reference's OWN source (blend_chain.npz) ----

@pytest.fixture(scope="module")
def chain_gold():
    return np.load(os.path.join(GOLD, "blend_chain.npz"))


def test_quat_mul_and_fill_nearest_vs_reference_source(chain_gold):
    from oracle import deform_oracle as do
    g = chain_gold
    assert np.array_equal(do.quat_mul(torch.tensor(g["Q_a"]), torch.tensor(g["Q_b"])).numpy(), g["Q_ab"])
    wx, wr = do.fill_nearest(torch.tensor(g["logits"]), torch.tensor(g["W_nearest_mask"]), g["kp"].shape[0], 6)
    assert np.array_equal(wx.numpy(), g["W_weights_xyz"]) and np.array_equal(wr.numpy(), g["W_weights_r"])


@pytest.mark.parametrize("tag", ["A", "A2", "B", "C", "D", "D2", "E", "F", "W"])
def test_model_forward_oracle_vs_reference_source(chain_gold, tag):
    """oracle.model_forward == the reference's GaussianModel.forward executed from its own source text: outputs,
    side-effect attributes and every autograd gradient."""
    from helpers import CHAIN_CASES, CHAIN_LEAVES, chain_args, chain_state_dict
    from oracle import deform_oracle as do
    g = chain_gold
    it, flags = CHAIN_CASES[tag]
    args = chain_args(flags)
    leaf = {k: torch.tensor(g[k]).requires_grad_(True) for k in CHAIN_LEAVES}
    sd = {k: v.requires_grad_(True) for k, v in chain_state_dict(g, 8 if args.step_opacity else 7).items()}
    mask = torch.tensor(g[tag + "_nearest_mask"]) if it > 30000 else None
    if mask is not None:      # the search itself: oracle kNN == the reference's get_nearest_mask (on its CPU brute force)
        p1, p2 = (leaf["xyz"].detach(), leaf["kp"].detach()) if args.knn_type == "3D" else \
            do.hybrid_points(leaf["xyz"].detach(), leaf["motion_feature"].detach(), leaf["kp"].detach(), leaf["kf"].detach(), 0.5)
        _, idx = do.knn(p1.numpy(), p2.numpy(), 6, fma=False)
        assert np.array_equal(idx, g[tag + "_nearest_mask"])
    noise = torch.tensor(g[tag + "_noise"]) if (tag + "_noise") in g.files else None
    res, side = do.model_forward(torch.tensor([float(g["t"])]), it, leaf, sd, args, nearest_mask=mask, noise=noise,
                                 return_weights=(tag == "W"))
    for n, v in zip(("xyz_t", "q_t", "s_t", "o_t", "weights_xyz", "weights_r"), res):
        assert np.abs(v.detach().numpy() - g[f"{tag}_{n}"]).max() <= 1e-6, n
    for n, v in side.items():
        assert np.abs(v.detach().numpy() - g[f"{tag}_{n}"]).max() <= 1e-6, n
    cot = [torch.tensor(g[k]) for k in ("cot_xyz", "cot_q", "cot_s", "cot_o")]
    loss = sum((a * b).sum() for a, b in zip(res[:4], cot))
    names = list(leaf) + ["net_" + k for k in sd]
    grads = torch.autograd.grad(loss, list(leaf.values()) + list(sd.values()), allow_unused=True)
    for n, gr in zip(names, grads):
        key = f"{tag}_g_{n}"
        assert (gr is not None) == (key in g.files), n
        if gr is not None:
            assert rel(gr.numpy(), g[key]) <= 1e-5, (n, rel(gr.numpy(), g[key]))


def test_extrapolate_frame_oracle_vs_reference_source(chain_gold):
    from oracle import deform_oracle as do
    g = chain_gold
    wx, wr = torch.tensor(g["W_weights_xyz"]), torch.tensor(g["W_weights_r"])
    for i in range(3):
        xf, dr = do.extrapolate_frame(torch.tensor(g["xyz"]), torch.tensor(g["rotation"]), wx, wr, torch.tensor(g["X_kpts"][i]),
                                      torch.tensor(g["kp"]), torch.tensor(g["X_kpts_rotation"][i]))
        assert np.abs(xf.numpy() - g[f"X_xyz_{i}"]).max() <= 1e-6 and np.abs(dr.numpy() - g[f"X_r_{i}"]).max() <= 1e-6

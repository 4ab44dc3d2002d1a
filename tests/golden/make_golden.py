"""Generates the golden vectors under tests/golden/ from the REFERENCE itself.

  python tests/golden/make_golden.py --cpu     # here (needs /root/reference and oracle/_ref/ref_frnn.so)
      deform_mlp_pe.npz : reference ``positional_encoding`` / ``Deformable_Field`` (scene/deformable_field.py, loaded
                          by file path) on seeded inputs, incl. the module's state_dict
      knn_cpu.npz       : the reference's own CPU brute force ``frnn._C.frnn_bf_cpu``
      blend_chain.npz   : the reference's own ``GaussianModel.forward`` / ``fill_nearest`` / ``get_motion_delta`` /
                          ``get_rotation_`` / ``get_nearest_mask`` (scene/gaussian_model.py), ``quat_mul``
                          (utils/camera_utils.py) and the two extrapolation statements of ``render_kpts``
                          (eval.py:140-141), cut out of the reference files as text and executed against a stub
                          ``self`` (the module itself cannot be imported: tinycudann, pytorch3d, ... are absent)
  python tests/golden/make_golden.py --gpu     # on the B200 box (gpurun); needs only oracle/_ref/*.so
      raster_small.npz  : inputs + every output of the reference CUDA rasterizer, forward and backward
      knn_gpu.npz       : reference FRNN CUDA path (insert/prefix-sum/counting-sort/find_nbrs), D=35 and D=3
      dist2.npz         : reference simple-knn distCUDA2
    written to gpurun_out/golden/ (copied into tests/golden/ afterwards).

The vectors pin oracle/ (CPU tests) and are also checked against the CUDA library (GPU tests).
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def knn_inputs(seed, N, Kp, F=32, amplify=0.5):
    from gaussianprediction_b200 import synthetic as syn
    gs = syn.make_gaussians(N, seed=seed)
    kp, kf = syn.make_keypoints(gs, Kp=Kp, seed=seed + 1, feature_dim=F)
    p1 = torch.cat([gs.xyz, gs.motion_feature[:, :F] * amplify], dim=-1).contiguous()
    p2 = torch.cat([kp, kf * amplify], dim=-1).contiguous()
    return gs, kp, kf, p1, p2


def make_cpu(out_dir):
    ref_root = os.environ.get("GP_REFERENCE_ROOT", "/root/reference")
    spec = importlib.util.spec_from_file_location("ref_deformable_field", os.path.join(ref_root, "scene", "deformable_field.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    torch.manual_seed(2024)
    net = m.Deformable_Field(12 + 60 + 32, d=4, w=32, output_dim=7)
    x = torch.randn(24, 104)
    pos = torch.randn(9, 3)
    tt = torch.tensor([0.37])
    with torch.no_grad():
        y = net(x)
        pe_xyz = m.positional_encoding(pos, 10)
        pe_t = m.positional_encoding(tt, 6)
    sd = {k: v.numpy() for k, v in net.state_dict().items()}
    np.savez_compressed(os.path.join(out_dir, "deform_mlp_pe.npz"), x=x.numpy(), y=y.numpy(), pos=pos.numpy(),
                        pe_xyz=pe_xyz.numpy(), t=tt.numpy(), pe_t=pe_t.numpy(), **{"sd_" + k: v for k, v in sd.items()})
    from oracle import ref_cuda
    _, _, _, p1, p2 = knn_inputs(5, 700, 64)
    d35, i35 = ref_cuda.frnn_bf_cpu(p1[None], p2[None], 6, 1e8)
    d3, i3 = ref_cuda.frnn_bf_cpu(p1[None, :, :3].contiguous(), p2[None, :, :3].contiguous(), 6, 1e3)
    np.savez_compressed(os.path.join(out_dir, "knn_cpu.npz"), p1=p1.numpy(), p2=p2.numpy(), d35=d35[0].numpy(),
                        i35=i35[0].numpy(), d3=d3[0].numpy(), i3=i3[0].numpy())
    # image loss: the reference's own l1_loss / ssim (utils/loss_utils.py) and their autograd gradient
    spec = importlib.util.spec_from_file_location("ref_loss_utils", os.path.join(ref_root, "utils", "loss_utils.py"))
    lu = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(lu)
    g = torch.Generator().manual_seed(7)
    img = torch.rand(3, 45, 70, generator=g).requires_grad_(True)
    gt = (img.detach() + 0.1 * torch.randn(3, 45, 70, generator=g)).clamp(0, 1)
    l1 = lu.l1_loss(img, gt)
    ss = lu.ssim(img, gt)
    loss = (1.0 - 0.2) * l1 + 0.2 * (1.0 - ss)
    loss.backward()
    np.savez_compressed(os.path.join(out_dir, "loss.npz"), img=img.detach().numpy(), gt=gt.numpy(), l1=l1.item(),
                        ssim=ss.item(), loss=loss.item(), grad=img.grad.numpy())
    print("wrote deform_mlp_pe.npz, knn_cpu.npz, loss.npz to", out_dir)


# ------------------------------------------------------------------------------------------------ blend chain

CHAIN_METHODS = ("setup_functions", "get_nearest_mask", "get_scaling", "get_rotation", "get_xyz", "get_superGaussians",
                 "get_opacity", "get_motion_delta", "fill_nearest", "forward", "get_rotation_")


def _slice_defs(path, names, cls=None):
    """Source text of the named top-level functions (or methods of ``cls``) of a reference file, decorators included."""
    import ast
    src = open(path).read()
    body = ast.parse(src).body
    if cls is not None:
        body = [n for n in body if isinstance(n, ast.ClassDef) and n.name == cls][0].body
    lines = src.splitlines()
    out = []
    for n in body:
        if isinstance(n, ast.FunctionDef) and n.name in names:
            first = min([d.lineno for d in n.decorator_list] + [n.lineno])
            out.append("\n".join(lines[first - 1:n.end_lineno]))
    assert len(out) == len(names), f"{path}: expected {names}"
    return out


def _load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def reference_chain(ref_root):
    """(RefChain class, quat_mul, extrapolate(xyz, kpts, weights_xyz, weights_r, kpts_i, rot_i, gaussians)) built from
    the reference's own source text."""
    import ast
    import textwrap
    import torch.nn as nn
    from oracle import ref_cuda
    df = _load_by_path("ref_deformable_field", os.path.join(ref_root, "scene", "deformable_field.py"))
    gu = _load_by_path("ref_general_utils", os.path.join(ref_root, "utils", "general_utils.py"))

    def frnn_grid_points(p1, p2, K=-1, r=-1):
        # the reference's own CPU brute force stands in for its CUDA grid search (frnn raises on CPU tensors,
        # frnn/frnn.py:295-296); the CUDA path is pinned separately (knn_gpu.npz)
        d, i = ref_cuda.frnn_bf_cpu(p1.contiguous(), p2.contiguous(), K, r)
        return d, i, None, None

    from types import SimpleNamespace
    glob = dict(torch=torch, nn=nn, positional_encoding=df.positional_encoding, inverse_sigmoid=gu.inverse_sigmoid,
                build_scaling_rotation=gu.build_scaling_rotation, strip_symmetric=gu.strip_symmetric,
                frnn=SimpleNamespace(frnn_grid_points=frnn_grid_points))
    exec(_slice_defs(os.path.join(ref_root, "utils", "camera_utils.py"), ("quat_mul",))[0], glob)
    methods = _slice_defs(os.path.join(ref_root, "scene", "gaussian_model.py"), CHAIN_METHODS, cls="GaussianModel")
    exec("class RefChain:\n" + "\n\n".join(methods), glob)
    # eval.py:140-141: the two assignments inside render_kpts' frame loop
    src = open(os.path.join(ref_root, "eval.py")).read()
    fn = [n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "render_kpts"][0]
    loop = [n for n in fn.body if isinstance(n, ast.For)][0]
    wanted = {"gaussian_xyz", "gaussian_kpts", "xyz_final", "delta_r"}
    keep = [n for n in loop.body if isinstance(n, ast.Assign) and
            {x.id for t in n.targets for x in ast.walk(t) if isinstance(x, ast.Name)} & wanted]
    assert len(keep) == 3
    lines = src.splitlines()
    body = textwrap.dedent("\n".join("\n".join(lines[n.lineno - 1:n.end_lineno]) for n in keep))
    exec("def extrapolate(gaussians, weights_xyz, weights_r, kpts, kpts_rotation, i):\n" + textwrap.indent(body, "    ") +
         "\n    return xyz_final, delta_r\n", glob)
    return glob["RefChain"], glob["quat_mul"], glob["extrapolate"], df


def chain_inputs(seed=11, N=257, Kp=37, K=6, F=32):
    from gaussianprediction_b200 import synthetic as syn
    gs = syn.make_gaussians(N, seed=seed)
    kp, kf = syn.make_keypoints(gs, Kp=Kp, seed=seed + 1, feature_dim=F)
    g = torch.Generator().manual_seed(seed + 2)
    logits = torch.randn(N, 2 * K, generator=g)
    thres = -2 + 2.5 * torch.rand(N, 1, generator=g)           # opacity_thres (scene/gaussian_model.py:358-361 inits it to -2)
    return gs, kp, kf, logits, thres


def make_chain(out_dir, ref_root):
    from types import SimpleNamespace
    RefChain, quat_mul, extrapolate, df = reference_chain(ref_root)
    N, Kp, K, F, W = 257, 37, 6, 32, 64
    gs, kp, kf, logits, thres = chain_inputs(N=N, Kp=Kp, K=K, F=F)
    out = dict(xyz=gs.xyz.numpy(), rotation=gs.rotation.numpy(), scaling=gs.scaling.numpy(), opacity=gs.opacity.numpy(),
               motion_feature=gs.motion_feature[:, :F].numpy(), kp=kp.numpy(), kf=kf.numpy(), logits=logits.numpy(),
               thres=thres.numpy(), t=np.float32(0.37))
    nets = {}
    for od in (7, 8):
        torch.manual_seed(100 + od)
        nets[od] = df.Deformable_Field(12 + 60 + F, d=4, w=W, output_dim=od)
        out.update({f"sd{od}_" + k: v.numpy() for k, v in nets[od].state_dict().items()})
    g = torch.Generator().manual_seed(77)
    cot = [torch.randn(N, 3, generator=g), torch.randn(N, 4, generator=g), torch.randn(N, 3, generator=g),
           torch.randn(N, 1, generator=g)]
    out.update(cot_xyz=cot[0].numpy(), cot_q=cot[1].numpy(), cot_s=cot[2].numpy(), cot_o=cot[3].numpy())

    def run(tag, iteration, norm_rotation=True, step_opacity=False, opacity_type="implicit", feature_amplify=0.5,
            knn_type="hybird", return_weights=False, seed=None):
        leaf = {k: torch.tensor(out[k]).requires_grad_(True) for k in
                ("xyz", "rotation", "scaling", "opacity", "motion_feature", "kp", "kf", "logits", "thres")}
        net = df.Deformable_Field(12 + 60 + F, d=4, w=W, output_dim=8 if step_opacity else 7)
        net.load_state_dict(nets[8 if step_opacity else 7].state_dict())
        m = RefChain.__new__(RefChain)
        m.args = SimpleNamespace(nearest_num=K, knn_type=knn_type, feature_amplify=feature_amplify, norm_rotation=norm_rotation,
                                 step_opacity=step_opacity, step_opacity_iteration=5000, opacity_type=opacity_type,
                                 jointly_iteration=1000, xyz_noise_iteration=10000, densify_from_teaching=False)
        m.beta = 0.1
        m.second_stage_iter, m.third_stage_iter = 30000, 10 ** 9
        m.time_input_dim, m.xyz_input_dim = 12, 60
        m._xyz, m._rotation, m._scaling, m._opacity = leaf["xyz"], leaf["rotation"], leaf["scaling"], leaf["opacity"]
        m.motion_feature, m.super_gaussians, m.super_gaussians_feature = leaf["motion_feature"], leaf["kp"], leaf["kf"]
        m.opacity_thres = leaf["thres"]
        m.df_model = net
        m.weights_model = lambda x: leaf["logits"]            # stand-in for the absent tinycudann network (:257)
        m.weights_fill = torch.zeros(N, Kp + 11)              # dense zeros [N, Kp_max] (:226-227 slice it)
        m.setup_functions()
        if seed is not None:
            torch.manual_seed(seed)
            out[tag + "_noise"] = (torch.randn_like(leaf["kp"] if iteration > 30000 else leaf["xyz"])).numpy()
            torch.manual_seed(seed)
        res = m.forward(torch.tensor([out["t"]]), iteration, return_weights=return_weights)
        names = ("xyz_t", "q_t", "s_t", "o_t", "weights_xyz", "weights_r")
        for n, v in zip(names, res):
            out[f"{tag}_{n}"] = v.detach().numpy()
        loss = sum((a * b).sum() for a, b in zip(res[:4], cot))
        params = list(leaf.values()) + list(net.parameters())
        grads = torch.autograd.grad(loss, params, allow_unused=True)
        for n, gr in zip(list(leaf) + ["net_" + k for k, _ in net.named_parameters()], grads):
            if gr is not None:
                out[f"{tag}_g_{n}"] = gr.numpy()
        if iteration >= 1000:
            out[tag + "_kpts_xyz_motion"] = m.kpts_xyz_motion.detach().numpy()
            out[tag + "_kpts_rotation_motion"] = m.kpts_rotation_motion.detach().numpy()
        if iteration > 30000:
            out[tag + "_nearest_mask"] = m.nearest_mask.numpy()
            out[tag + "_weights_sum"] = m.weights_sum.detach().numpy()
        if getattr(m, "lifecycle_opacity", None) is not None:
            out[tag + "_lifecycle_opacity"] = m.lifecycle_opacity.detach().numpy()
        return m, res

    run("A", 50000, norm_rotation=True)
    run("A2", 50000, norm_rotation=False, knn_type="3D")
    run("B", 50000, norm_rotation=False, step_opacity=True, opacity_type="implicit")
    run("C", 50000, norm_rotation=True, step_opacity=True, opacity_type="explicit")
    run("D", 20000, norm_rotation=True)                                  # first stage, noise factor 0
    run("D2", 20000, norm_rotation=False, step_opacity=True)             # first stage + implicit step opacity
    run("E", 500)                                                        # warm-up
    run("F", 32000, norm_rotation=True, seed=5)                          # second stage with keypoint noise (factor 0.08)
    m, res = run("W", 50000, norm_rotation=True, return_weights=True)
    g = torch.Generator().manual_seed(78)
    kpts = (kp[None] + 0.3 * torch.randn(3, Kp, 3, generator=g))
    krot = torch.nn.functional.normalize(torch.tensor([1.0, 0, 0, 0]) + 0.2 * torch.randn(3, Kp, 4, generator=g), dim=-1)
    gaussians = SimpleNamespace(get_xyz=m.get_xyz.detach(), super_gaussians=m.super_gaussians.detach(),
                                get_rotation_=lambda d: m.get_rotation_(d).detach(), rotation_activation=m.rotation_activation)
    out["X_kpts"], out["X_kpts_rotation"] = kpts.numpy(), krot.numpy()
    for i in range(3):
        xf, dr = extrapolate(gaussians, res[4].detach(), res[5].detach(), kpts, krot, i)
        out[f"X_xyz_{i}"], out[f"X_r_{i}"] = xf.numpy(), dr.numpy()
    g = torch.Generator().manual_seed(79)
    a, b = torch.randn(50, 4, generator=g), torch.randn(50, 4, generator=g)
    out["Q_a"], out["Q_b"], out["Q_ab"] = a.numpy(), b.numpy(), quat_mul(a, b).numpy()
    np.savez_compressed(os.path.join(out_dir, "blend_chain.npz"), **out)
    print("wrote blend_chain.npz to", out_dir, f"({len(out)} arrays)")


def make_gpu(out_dir):
    from helpers import raster_case
    from oracle import ref_cuda
    os.makedirs(out_dir, exist_ok=True)
    dev = "cuda"
    # rasterizer
    case = raster_case(N=600, W=64, H=48, seed=21)
    c = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in case.items()}
    fwd = ref_cuda.rasterize(c["means3D"], c["opacities"], c["bg"], c["view"], c["proj"], c["campos"], c["W"], c["H"],
                             c["tanfovx"], c["tanfovy"], shs=c["shs"], scales=c["scales"], rotations=c["rotations"],
                             sh_degree=3)
    st = ref_cuda.decode_state(fwd)
    grad = torch.randn(3, c["H"], c["W"], generator=torch.Generator().manual_seed(9)) * 1e-2
    g = ref_cuda.rasterize_backward(fwd, grad.to(dev))
    vis = (fwd["radii"] > 0)

    def clean(x, mask_rows=True):
        x = x.clone()
        if mask_rows:
            x[~vis] = 0          # rows of culled Gaussians are uninitialised memory in the reference buffers
        return x.cpu().numpy()
    out = {("in_" + k): (v.cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in case.items()}
    out.update(color=fwd["color"].cpu().numpy(), depth=fwd["depth"].cpu().numpy(), tidx=fwd["tidx"].cpu().numpy(),
               radii=fwd["radii"].cpu().numpy(), num_rendered=np.int64(fwd["num_rendered"]),
               depths=clean(st["depths"]), means2D=clean(st["means2D"]), cov3D=clean(st["cov3D"]),
               conic_opacity=clean(st["conic_opacity"]), rgb=clean(st["rgb"]), clamped=clean(st["clamped"]),
               tiles_touched=st["tiles_touched"].cpu().numpy(), point_list=st["point_list"].cpu().numpy(),
               point_list_keys=st["point_list_keys"].cpu().numpy(), ranges=st["ranges"].cpu().numpy(),
               final_T=st["final_T"].cpu().numpy(), n_contrib=st["n_contrib"].cpu().numpy(), grad_in=grad.numpy())
    out.update({("g_" + k): v.cpu().numpy() for k, v in g.items()})
    np.savez_compressed(os.path.join(out_dir, "raster_small.npz"), **out)
    # kNN through the reference's CUDA grid path
    _, _, _, p1, p2 = knn_inputs(5, 700, 64)
    d35, i35 = ref_cuda.frnn_grid_points(p1[None].to(dev), p2[None].to(dev), 6, 1e8)
    d3, i3 = ref_cuda.frnn_grid_points(p1[None, :, :3].contiguous().to(dev), p2[None, :, :3].contiguous().to(dev), 6, 1e3)
    np.savez_compressed(os.path.join(out_dir, "knn_gpu.npz"), p1=p1.numpy(), p2=p2.numpy(), d35=d35[0].cpu().numpy(),
                        i35=i35[0].cpu().numpy(), d3=d3[0].cpu().numpy(), i3=i3[0].cpu().numpy())
    # distCUDA2
    pts = p1[:, :3].contiguous()
    d2 = ref_cuda.dist_cuda2(pts.to(dev))
    np.savez_compressed(os.path.join(out_dir, "dist2.npz"), points=pts.numpy(), dist2=d2.cpu().numpy())
    print("wrote raster_small.npz, knn_gpu.npz, dist2.npz to", out_dir)


if __name__ == "__main__":
    if "--cpu" in sys.argv:
        make_cpu(HERE)
    if "--chain" in sys.argv or "--cpu" in sys.argv:
        make_chain(HERE, os.environ.get("GP_REFERENCE_ROOT", "/root/reference"))
    if "--gpu" in sys.argv:
        make_gpu(os.path.join(ROOT, "gpurun_out", "golden"))

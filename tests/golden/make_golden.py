"""Generates the golden vectors under tests/golden/ from the REFERENCE itself.

  python tests/golden/make_golden.py --cpu     # here (needs /root/reference and oracle/_ref/ref_frnn.so)
      deform_mlp_pe.npz : reference ``positional_encoding`` / ``Deformable_Field`` (scene/deformable_field.py, loaded
                          by file path) on seeded inputs, incl. the module's state_dict
      knn_cpu.npz       : the reference's own CPU brute force ``frnn._C.frnn_bf_cpu``
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
    if "--gpu" in sys.argv:
        make_gpu(os.path.join(ROOT, "gpurun_out", "golden"))

"""Loader and thin callers for the UNMODIFIED reference CUDA extensions built into oracle/_ref/.

TEST INFRASTRUCTURE ONLY (parity checker and the ``--impl reference`` bench arm).  The .so files are
built from /root/reference in place by oracle/build_ref.py; this module contains no reference source,
only the argument plumbing needed to call the pybind entry points and to read back their buffers.

The Python glue of the reference (diff_gaussian_rasterization/__init__.py, frnn/frnn.py) cannot travel
to the GPU box, so the call sequences are restated here with citations.
"""
import importlib.util
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
_mods = {}


def available(name="ref_dgr"):
    return os.path.exists(os.path.join(REF_DIR, name + ".so"))


def load(name):
    if name not in _mods:
        path = os.path.join(REF_DIR, name + ".so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `python oracle/build_ref.py` where /root/reference is mounted")
        spec = importlib.util.spec_from_file_location(name, path)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        _mods[name] = m
    return _mods[name]


def _align(x, a=128):
    return (x + a - 1) // a * a


def _e(t, dev):
    """The reference passes empty tensors for absent inputs (diff_gaussian_rasterization/__init__.py:173-183)."""
    return torch.empty(0, device=dev) if t is None else t


def rasterize(means3D, opacities, bg, view, proj, campos, W, H, tanfovx, tanfovy, shs=None, colors_precomp=None,
              scales=None, rotations=None, cov3D_precomp=None, sh_degree=3, scale_modifier=1.0, prefiltered=False):
    """``_C.rasterize_gaussians(*args)`` with the argument order of
    diff_gaussian_rasterization/__init__.py:56-79.  Returns a dict incl. the three opaque buffers."""
    C = load("ref_dgr")
    dev = means3D.device
    args = (bg, means3D, _e(colors_precomp, dev), opacities, _e(scales, dev), _e(rotations, dev), float(scale_modifier),
            _e(cov3D_precomp, dev), view, proj, float(tanfovx), float(tanfovy), int(H), int(W), _e(shs, dev),
            int(sh_degree), campos, bool(prefiltered))
    num_rendered, color, radii, geom, binning, img, depth, tidx = C.rasterize_gaussians(*args)
    return dict(num_rendered=int(num_rendered), color=color, radii=radii, depth=depth, tidx=tidx, geom=geom,
                binning=binning, img=img, _args=dict(
                    means3D=means3D, opacities=opacities, bg=bg, view=view, proj=proj, campos=campos, W=W, H=H,
                    tanfovx=tanfovx, tanfovy=tanfovy, shs=shs, colors_precomp=colors_precomp, scales=scales,
                    rotations=rotations, cov3D_precomp=cov3D_precomp, sh_degree=sh_degree,
                    scale_modifier=scale_modifier))


def rasterize_backward(fwd, grad_out_color):
    """``_C.rasterize_gaussians_backward(*args)``, argument order of diff_gaussian_rasterization/__init__.py:97-118.
    Returns dict(means2D, colors, opacity, means3D, cov3D, sh, scales, rotations)."""
    C = load("ref_dgr")
    a = fwd["_args"]
    dev = a["means3D"].device
    args = (a["bg"], a["means3D"], fwd["radii"], _e(a["colors_precomp"], dev), _e(a["scales"], dev),
            _e(a["rotations"], dev), float(a["scale_modifier"]), _e(a["cov3D_precomp"], dev), a["view"], a["proj"],
            float(a["tanfovx"]), float(a["tanfovy"]), grad_out_color, _e(a["shs"], dev), int(a["sh_degree"]),
            a["campos"], fwd["geom"], int(fwd["num_rendered"]), fwd["binning"], fwd["img"])
    g2d, gcol, gop, g3d, gcov, gsh, gsc, grot = C.rasterize_gaussians_backward(*args)
    return dict(means2D=g2d, colors=gcol, opacity=gop, means3D=g3d, cov3D=gcov, sh=gsh, scales=gsc, rotations=grot)


def decode_state(fwd):
    """Unpack the reference's geom/binning/img byte buffers.  Layout = the ``obtain`` sequence of
    GeometryState/BinningState/ImageState::fromChunk (rasterizer_impl.cu:155-196) with 128-byte alignment
    (rasterizer_impl.h:22-27).  Only arrays in front of the CUB scratch space are decoded."""
    a = fwd["_args"]
    P, W, H, R = a["means3D"].shape[0], a["W"], a["H"], fwd["num_rendered"]
    out = {}

    def carve(buf, specs):
        base = buf.data_ptr()
        off = 0
        for name, dtype, count, shape in specs:
            start = _align(base + off) - base
            nbytes = count * torch.tensor([], dtype=dtype).element_size()
            out[name] = buf[start:start + nbytes].view(dtype).view(shape).clone()
            off = start + nbytes

    carve(fwd["geom"], [
        ("depths", torch.float32, P, (P,)), ("clamped", torch.uint8, 3 * P, (P, 3)),
        ("internal_radii", torch.int32, P, (P,)), ("means2D", torch.float32, 2 * P, (P, 2)),
        ("cov3D", torch.float32, 6 * P, (P, 6)), ("conic_opacity", torch.float32, 4 * P, (P, 4)),
        ("rgb", torch.float32, 3 * P, (P, 3)), ("tiles_touched", torch.int32, P, (P,))])
    if R > 0:
        carve(fwd["binning"], [
            ("point_list", torch.int32, R, (R,)), ("point_list_unsorted", torch.int32, R, (R,)),
            ("point_list_keys", torch.int64, R, (R,))])
    else:
        out["point_list"] = torch.zeros(0, dtype=torch.int32, device=fwd["geom"].device)
        out["point_list_keys"] = torch.zeros(0, dtype=torch.int64, device=fwd["geom"].device)
    N = W * H
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    carve(fwd["img"], [("final_T", torch.float32, N, (N,)), ("n_contrib", torch.int32, N, (N,)),
                       ("ranges_full", torch.int32, 2 * N, (N, 2))])
    out["ranges"] = out.pop("ranges_full")[:tiles].clone()
    return out


def mark_visible(positions, view, proj):
    return load("ref_dgr").mark_visible(positions, view, proj)


def dist_cuda2(points):
    """simple_knn._C.distCUDA2 (submodules/simple-knn/spatial.cu:15-26)."""
    return load("ref_sknn").distCUDA2(points)


def frnn_grid_points(points1, points2, K, r, radius_cell_ratio=2.0):
    """The call sequence of ``_frnn_grid_points.forward`` (submodules/FRNN/frnn/frnn.py:22-170) for one
    un-cached query with D >= 3, as ``get_nearest_mask`` issues it (scene/gaussian_model.py:110-125):
    grid set-up from points2's extent, insert + prefix-sum + counting-sort for both clouds, then
    ``find_nbrs_cuda``.  Returns (dists [N,P1,K], idxs int64 [N,P1,K])."""
    F = load("ref_frnn")
    PS = load("ref_prefix_sum")
    dev = points1.device
    N, P1, D = points1.shape
    P2 = points2.shape[1]
    assert D >= 3
    lengths1 = torch.full((N,), P1, dtype=torch.long, device=dev)
    lengths2 = torch.full((N,), P2, dtype=torch.long, device=dev)
    rt = torch.ones((N,), dtype=torch.float32, device=dev) * float(r)
    grid_params = torch.zeros((N, 8), dtype=torch.float, device=dev)
    G = -1
    for i in range(N):                                              # frnn.py:72-87
        grid_min = points2[i, :lengths2[i], :3].min(dim=0)[0]
        grid_max = points2[i, :lengths2[i], :3].max(dim=0)[0]
        grid_params[i, :3] = grid_min
        grid_size = grid_max - grid_min
        cell_size = rt[i].item() / radius_cell_ratio
        if cell_size < grid_size.max() / 64:
            cell_size = grid_size.max() / 64
        grid_params[i, 3] = 1 / cell_size
        grid_params[i, 4:7] = torch.floor(grid_size / cell_size) + 1
        grid_params[i, 7] = torch.prod(grid_params[i, 4:7])
        if G < grid_params[i, 7]:
            G = int(grid_params[i, 7].item())

    def sort_cloud(points, lengths, Pn):                            # frnn.py:90-130 / 137-170
        cnt = torch.zeros((N, G), dtype=torch.int, device=dev)
        cell = torch.full((N, Pn), -1, dtype=torch.int, device=dev)
        idx = torch.full((N, Pn), -1, dtype=torch.int, device=dev)
        F.insert_points_cuda(points, lengths, grid_params, cnt, cell, idx, G)
        gp_cpu = grid_params.cpu()
        off = torch.full((N, G), 0, dtype=torch.int, device=dev)
        for i in range(N):
            PS.prefix_sum_cuda(cnt[i], int(gp_cpu[i, 7].item()), off[i])
        sorted_points = torch.zeros_like(points)
        sorted_idxs = torch.full((N, Pn), -1, dtype=torch.int, device=dev)
        F.counting_sort_cuda(points, lengths, cell, idx, off, sorted_points, sorted_idxs)
        return sorted_points, sorted_idxs, off

    sp2, si2, off2 = sort_cloud(points2.contiguous(), lengths2, P2)
    sp1, si1, _ = sort_cloud(points1.contiguous(), lengths1, P1)
    idxs, dists = F.find_nbrs_cuda(sp1, sp2, lengths1, lengths2, off2, si1, si2, grid_params, int(K), rt, rt * rt)
    return dists, idxs


def frnn_bf_cpu(points1, points2, K, r):
    """The reference's own single-threaded CPU brute force (frnn/csrc/bruteforce/bruteforce_cpu.cpp:5-58)."""
    F = load("ref_frnn")
    N, P1, _ = points1.shape
    P2 = points2.shape[1]
    l1 = torch.full((N,), P1, dtype=torch.long)
    l2 = torch.full((N,), P2, dtype=torch.long)
    idxs, dists = F.frnn_bf_cpu(points1.contiguous(), points2.contiguous(), l1, l2, int(K), float(r))
    return dists, idxs

"""numpy front end of oracle/raster_oracle.c (CPU restatement of the reference rasterizer).

TEST INFRASTRUCTURE ONLY -- see the header of raster_oracle.c.  Nothing in gaussianprediction_b200/
imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "raster_oracle.c")
LIB = os.path.join(HERE, "_build", "libraster_oracle.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", SRC, "-o", LIB, "-lm"])
    return LIB


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.oracle_raster_forward.restype = ctypes.c_longlong
        _lib.oracle_raster_backward.restype = None
        _lib.oracle_free.restype = None
        _lib.oracle_free.argtypes = [ctypes.c_void_p]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def forward(means3D, opacities, bg, view, proj, campos, W, H, tanfovx, tanfovy, shs=None, colors_precomp=None,
            scales=None, rotations=None, cov3D_precomp=None, sh_degree=3, scale_modifier=1.0):
    """All inputs numpy (or array-likes).  Returns a dict with every array the reference produces."""
    L = _load()
    means3D, opacities, bg, view, proj, campos = map(_f, (means3D, opacities, bg, view, proj, campos))
    shs, colors_precomp, scales, rotations, cov3D_precomp = map(_f, (shs, colors_precomp, scales, rotations, cov3D_precomp))
    P = means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    o = dict(
        radii=np.zeros(P, np.int32), depths=np.zeros(P, np.float32), means2D=np.zeros((P, 2), np.float32),
        cov3D=np.zeros((P, 6), np.float32), conic_opacity=np.zeros((P, 4), np.float32), rgb=np.zeros((P, 3), np.float32),
        tiles_touched=np.zeros(P, np.uint32), clamped=np.zeros((P, 3), np.uint8), ranges=np.zeros((tiles, 2), np.uint32),
        color=np.zeros((3, H, W), np.float32), depth=np.zeros((1, H, W), np.float32), tidx=np.zeros((1, H, W), np.int32),
        final_T=np.zeros(H * W, np.float32), n_contrib=np.zeros(H * W, np.uint32))
    plist, keys = ctypes.c_void_p(), ctypes.c_void_p()
    R = L.oracle_raster_forward(
        ctypes.c_int(P), ctypes.c_int(sh_degree), ctypes.c_int(M), _p(bg), ctypes.c_int(W), ctypes.c_int(H), _p(means3D),
        _p(shs), _p(colors_precomp), _p(opacities), _p(scales), ctypes.c_float(scale_modifier), _p(rotations),
        _p(cov3D_precomp), _p(view), _p(proj), _p(campos), ctypes.c_float(tanfovx), ctypes.c_float(tanfovy),
        _p(o["radii"]), _p(o["depths"]), _p(o["means2D"]), _p(o["cov3D"]), _p(o["conic_opacity"]), _p(o["rgb"]),
        _p(o["tiles_touched"]), _p(o["clamped"]), ctypes.byref(plist), ctypes.byref(keys), _p(o["ranges"]),
        _p(o["color"]), _p(o["depth"]), _p(o["tidx"]), _p(o["final_T"]), _p(o["n_contrib"]))
    R = int(R)
    o["num_rendered"] = R
    o["point_list"] = np.ctypeslib.as_array(ctypes.cast(plist, ctypes.POINTER(ctypes.c_uint32)), shape=(max(R, 1),))[:R].copy()
    o["point_list_keys"] = np.ctypeslib.as_array(ctypes.cast(keys, ctypes.POINTER(ctypes.c_uint64)), shape=(max(R, 1),))[:R].copy()
    L.oracle_free(plist)
    L.oracle_free(keys)
    o["_inputs"] = dict(means3D=means3D, opacities=opacities, bg=bg, view=view, proj=proj, campos=campos, W=W, H=H,
                        tanfovx=tanfovx, tanfovy=tanfovy, shs=shs, colors_precomp=colors_precomp, scales=scales,
                        rotations=rotations, cov3D_precomp=cov3D_precomp, sh_degree=sh_degree,
                        scale_modifier=scale_modifier)
    return o


def backward(fwd, dL_dpix):
    """Gradients for the forward result ``fwd`` given dL/dcolor [3,H,W].  Returns a dict of numpy arrays."""
    L = _load()
    i = fwd["_inputs"]
    P = i["means3D"].shape[0]
    M = 0 if i["shs"] is None else i["shs"].shape[1]
    dL_dpix = _f(dL_dpix)
    g = dict(means2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 4), np.float32), opacity=np.zeros((P, 1), np.float32),
             colors=np.zeros((P, 3), np.float32), means3D=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32),
             sh=np.zeros((P, M, 3), np.float32), scales=np.zeros((P, 3), np.float32), rotations=np.zeros((P, 4), np.float32))
    L.oracle_raster_backward(
        ctypes.c_int(P), ctypes.c_int(i["sh_degree"]), ctypes.c_int(M), _p(i["bg"]), ctypes.c_int(i["W"]), ctypes.c_int(i["H"]),
        _p(i["means3D"]), _p(i["shs"]), _p(i["colors_precomp"]), _p(i["scales"]), ctypes.c_float(i["scale_modifier"]),
        _p(i["rotations"]), _p(i["cov3D_precomp"]), _p(i["view"]), _p(i["proj"]), _p(i["campos"]),
        ctypes.c_float(i["tanfovx"]), ctypes.c_float(i["tanfovy"]), _p(fwd["radii"]), _p(fwd["means2D"]), _p(fwd["cov3D"]),
        _p(fwd["conic_opacity"]), _p(fwd["rgb"]), _p(fwd["clamped"]), _p(fwd["point_list"]), _p(fwd["ranges"]),
        _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dL_dpix), _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]),
        _p(g["colors"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]))
    return g

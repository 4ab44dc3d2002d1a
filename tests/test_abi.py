"""CPU-side checks of the drop-in boundary: the shared library loads and exports exactly what include/gp_b200.h
declares, the ctypes table matches the header, and argument errors are reported without touching a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "gp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(gp_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_symbols_exported(gplib):
    from gaussianprediction_b200 import _lib
    names = header_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(gplib, n), f"{n} declared in include/gp_b200.h but not exported by libgp_b200.so"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in _lib.SIGNATURES"
    for n in _lib.SIGNATURES:
        assert n in names, f"{n} bound in _lib.py but not declared in the header"


def test_version_and_error_strings(gplib):
    assert gplib.gp_abi_version() == 1
    assert b"bad argument" in gplib.gp_error_string(-1)
    assert b"limit" in gplib.gp_error_string(-4)
    assert gplib.gp_error_string(0) == b"ok"


def test_sizes_and_argument_errors_without_gpu(gplib):
    g, b, i = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
    assert gplib.gp_raster_sizes(1000, 640, 480, 50000, ctypes.byref(g), ctypes.byref(b), ctypes.byref(i)) == 0
    assert g.value > 1000 * 100 and b.value >= 50000 * 16 and i.value >= 640 * 480 * 8
    assert gplib.gp_raster_sizes(-1, 640, 480, 0, None, None, None) == -1
    assert gplib.gp_knn_points(10, 5, 3, 64, 1.0, None, None, None, None, None) != 0      # K > 32 / NULL pointers
    assert gplib.gp_raster_preprocess(None, *([None] * 13)) == -1
    assert gplib.gp_dist2_ws_bytes(10) > 0 and gplib.gp_raster_backward_ws_bytes(10) >= 480


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from gaussianprediction_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.GpError):
        _lib.lib()


def test_cpu_tensors_are_rejected(gplib):
    import torch
    from gaussianprediction_b200 import GaussianRasterizer, frnn
    from gaussianprediction_b200.simple_knn import distCUDA2
    from helpers import raster_case, settings_of
    case = raster_case(N=50, W=32, H=32)
    with pytest.raises(RuntimeError):
        GaussianRasterizer(settings_of(case))(case["means3D"], torch.zeros(50, 3), case["opacities"], shs=case["shs"],
                                              scales=case["scales"], rotations=case["rotations"])
    with pytest.raises(TypeError):
        frnn.frnn_grid_points(torch.zeros(1, 4, 3), torch.zeros(1, 4, 3), K=2, r=1.0)
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(4, 3))


def test_dropin_import_names():
    import gaussianprediction_b200 as g
    g.install_dropin()
    import diff_gaussian_rasterization as dgr
    import frnn
    from simple_knn._C import distCUDA2
    assert dgr.GaussianRasterizer is g.GaussianRasterizer and callable(frnn.frnn_grid_points) and callable(distCUDA2)
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered")


def test_keypoint_mlp_limits_without_gpu(gplib):
    from gaussianprediction_b200._lib import MlpCfg
    sv, bs = ctypes.c_size_t(), ctypes.c_size_t()
    ok = MlpCfg(512, 32, 10, 6, 4, 256, 7, 1)
    assert gplib.gp_kpmlp_sizes(ctypes.byref(ok), ctypes.byref(sv), ctypes.byref(bs)) == 0
    assert sv.value >= 512 * (104 + 4 * 256 + 8) * 4 and bs.value >= 512 * (4 * 256 + 8) * 4
    for bad in (MlpCfg(512, 32, 10, 6, 4, 260, 7, 0), MlpCfg(512, 30, 10, 6, 4, 256, 7, 0), MlpCfg(512, 32, 10, 6, 9, 256, 7, 0),
                MlpCfg(512, 32, 10, 6, 4, 256, 9, 0), MlpCfg(512, 32, 10, 6, 4, 256, 3, 0)):
        assert gplib.gp_kpmlp_sizes(ctypes.byref(bad), None, None) == -4
    assert gplib.gp_kpmlp_sizes(None, None, None) == -1
    assert gplib.gp_kpmlp_forward(ctypes.byref(ok), *([None] * 10)) == -1


def test_header_is_plain_c_and_struct_layouts_match_ctypes(tmp_path):
    """include/gp_b200.h compiles as C99 (what a cgo / JNI / ctypes-free binding would include) and the structs that
    cross the ABI by pointer have the size and field offsets the ctypes mirrors in _lib.py assume."""
    import shutil
    import subprocess
    from gaussianprediction_b200 import _lib
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    src = tmp_path / "abi_probe.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "gp_b200.h"\n'
        'int main(void) {\n'
        '  printf("%zu %zu %zu %zu\\n", sizeof(gp_raster_cfg), offsetof(gp_raster_cfg, tanfovx), '
        'offsetof(gp_raster_cfg, sh_rest), offsetof(gp_raster_cfg, dL_dsh_rest));\n'
        '  printf("%zu %zu\\n", sizeof(gp_mlp_cfg), offsetof(gp_mlp_cfg, normalize_q));\n'
        '  printf("%zu %zu %zu %zu\\n", sizeof(gp_adam_tensor), offsetof(gp_adam_tensor, n), '
        'offsetof(gp_adam_tensor, lr), offsetof(gp_adam_tensor, step));\n'
        '  return 0;\n}\n')
    exe = tmp_path / "abi_probe"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)],
                   check=True)
    rows = [[int(x) for x in line.split()] for line in subprocess.run([str(exe)], check=True, capture_output=True,
                                                                        text=True).stdout.strip().splitlines()]
    R, M, A = _lib.RasterCfg, _lib.MlpCfg, _lib.AdamTensor
    assert rows[0] == [ctypes.sizeof(R), R.tanfovx.offset, R.sh_rest.offset, R.dL_dsh_rest.offset]
    assert rows[1] == [ctypes.sizeof(M), M.normalize_q.offset]
    assert rows[2] == [ctypes.sizeof(A), A.n.offset, A.lr.offset, A.step.offset]


def test_hashgrid_level_table_matches_oracle(gplib):
    """Host-side level geometry of the hash grid (entries per level: dense below 2^log2_T, aligned to 8) -- no GPU needed."""
    import numpy as np
    from oracle import hashgrid_oracle as ho
    for log2_t in (12, 19):
        _, total = ho.levels(log2_hashmap_size=log2_t)
        assert gplib.gp_hashgrid_entries(16, 16, float(np.exp(np.log(2048 / 16) / 15)), log2_t) == total
    assert gplib.gp_hashgrid_entries(0, 16, 1.5, 19) == -1

"""Compile csrc/*.cu for sm_100a into gaussianprediction_b200/libgp_b200.so (in-tree, so it ships to the GPU box).

    python -m gaussianprediction_b200.build [--force] [-v]

nvcc cross-compiles without a GPU.  The library has no torch dependency: it is a plain C-ABI
shared object (include/gp_b200.h) linked only against the CUDA runtime.
"""
import glob
import hashlib
import os
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libgp_b200.so")
STAMP = os.path.join(PKG_DIR, ".libgp_b200.stamp")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc():
    cand = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")
    return cand if os.path.exists(cand) else "nvcc"


def _source_hash():
    h = hashlib.sha256()
    root = os.path.dirname(PKG_DIR)
    files = sorted(glob.glob(os.path.join(CSRC, "*"))) + [os.path.join(root, "include", "gp_b200.h")]
    for f in files:
        h.update(os.path.basename(f).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    want = _source_hash()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == want:
                return LIB_PATH
    sources = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    objdir = os.path.join(PKG_DIR, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for src in sources:
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"[gp_b200.build] nvcc failed on {src}:\n{out}\n")
        elif verbose:
            print(out)
    if failed:
        raise RuntimeError("nvcc compilation failed")
    cmd = [_nvcc(), "-shared", "-Wno-deprecated-gpu-targets", "-o", LIB_PATH] + objs  # static cudart: no runtime library path needed
    subprocess.check_call(cmd)
    with open(STAMP, "w") as fh:
        fh.write(want)
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)

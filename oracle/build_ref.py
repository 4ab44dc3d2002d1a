"""Build the UNMODIFIED reference CUDA extensions for sm_100 into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is imported by the product
package; only tests/, __graft_entry__.smoke() and bench.py's baseline legs use it.

The sources are compiled where they lie under /root/reference (read-only); only
object files and the final pybind modules are written, all below oracle/_ref/
(git-ignored, but shipped to the GPU box with the snapshot).  No reference
source is copied into this repository.

What gets built:
  ref_dgr        <- submodules/diff-gaussian-rasterization-w-depth  (needs -include cstdint under gcc 13)
  ref_sknn       <- submodules/simple-knn                           (needs -include cfloat)
  ref_prefix_sum <- submodules/FRNN/external/prefix_sum/prefix_sum.cu
  ref_frnn       <- submodules/FRNN/frnn/csrc                       (slow: ~1000 template instantiations per file)

Usage:  python oracle/build_ref.py [dgr sknn prefix_sum frnn]     (default: all)
Idempotent: a module whose .so already exists is skipped unless --force.
"""
import glob
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("GP_REFERENCE_ROOT", "/root/reference")
SUB = os.path.join(REF, "submodules")


def _specs():
    dgr = os.path.join(SUB, "diff-gaussian-rasterization-w-depth")
    sknn = os.path.join(SUB, "simple-knn")
    frnn = os.path.join(SUB, "FRNN", "frnn", "csrc")
    psum = os.path.join(SUB, "FRNN", "external", "prefix_sum")
    return {
        "dgr": dict(
            name="ref_dgr",
            sources=[os.path.join(dgr, p) for p in (
                "cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
                "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp")],
            include=[os.path.join(dgr, "third_party/glm"), dgr],
            nvcc=["-include", "cstdint"], cxx=[]),
        "sknn": dict(
            name="ref_sknn",
            sources=[os.path.join(sknn, p) for p in ("spatial.cu", "simple_knn.cu", "ext.cpp")],
            include=[sknn], nvcc=["-include", "cfloat"], cxx=[]),
        "prefix_sum": dict(
            name="ref_prefix_sum",
            sources=[os.path.join(psum, "prefix_sum.cu")],
            include=[psum], nvcc=[], cxx=[]),
        "frnn": dict(
            name="ref_frnn",
            sources=[os.path.join(frnn, "ext.cpp")]
            + sorted(glob.glob(os.path.join(frnn, "*", "*.cpp")))
            + sorted(glob.glob(os.path.join(frnn, "*", "*.cu"))),
            include=[frnn],
            nvcc=["-DCUDA_HAS_FP16=1", "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
                  "-D__CUDA_NO_HALF2_OPERATORS__", "--ftemplate-depth=2048"],
            cxx=["-std=c++17"]),
    }


def so_path(key):
    return os.path.join(OUT, _specs()[key]["name"] + ".so")


def build(keys=None, force=False, verbose=False):
    if not os.path.isdir(REF):
        raise RuntimeError(f"reference tree not found at {REF}; oracle/_ref can only be built where it is mounted")
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0"   # -gencode arch=compute_100,code=sm_100: the reference's stock flags
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
    from torch.utils import cpp_extension
    specs = _specs()
    for key in (keys or list(specs)):
        s = specs[key]
        target = os.path.join(OUT, s["name"] + ".so")
        if os.path.exists(target) and not force:
            print(f"[build_ref] {s['name']}: up to date")
            continue
        bdir = os.path.join(OUT, "build_" + key)
        os.makedirs(bdir, exist_ok=True)
        print(f"[build_ref] building {s['name']} from {len(s['sources'])} sources ...", flush=True)
        cpp_extension.load(
            name=s["name"], sources=s["sources"], extra_include_paths=s["include"],
            extra_cuda_cflags=s["nvcc"], extra_cflags=s["cxx"], build_directory=bdir,
            verbose=verbose, is_python_module=False)
        os.replace(os.path.join(bdir, s["name"] + ".so"), target)
        print(f"[build_ref] wrote {target}")


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("-")]
    build(args or None, force="--force" in sys.argv, verbose="-v" in sys.argv)

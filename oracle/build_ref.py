"""Build the reference's own CUDA extensions, UNMODIFIED, for sm_100a -> oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is imported by the product
package (nerfsafetyvalidation_b200/); only tests/, __graft_entry__.smoke() and
bench.py's baseline legs may load what this script produces.

The four extensions on the hot path are compiled from the sources where they lie
under /root/reference (nothing is copied into this repo):

    _raymarching   raymarching/src/{raymarching.cu,bindings.cpp}
    _gridencoder   gridencoder/src/{gridencoder.cu,bindings.cpp}
    _shencoder     shencoder/src/{shencoder.cu,bindings.cpp}
    _ffmlp         ffmlp/src/{ffmlp.cu,bindings.cpp} (+ vendored CUTLASS 2.8 includes)

Flags are the reference's (``<pkg>/backend.py:6-9`` and ``ffmlp/backend.py:6-13``) with the
single change -std=c++14 -> -std=c++17 (torch 2.11 headers need it) and an explicit
``-gencode arch=compute_100a,code=sm_100a`` (the reference names no arch at all).

The resulting ``oracle/_ref/_<name>.so`` files are git-ignored but travel to the GPU box
with the gpurun snapshot; there they are the bit-exactness oracle for march/composite/
near-far/packbits/morton/grid-index and the "reference rebuilt on the same B200" speed bar.

Usage:  python oracle/build_ref.py [name ...]        (default: all four)
"""
import os
import shutil
import sys

REF = os.environ.get("NGP_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

COMMON = ["-O3", "-std=c++17",
          "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
          "-U__CUDA_NO_HALF2_OPERATORS__",
          "-gencode", "arch=compute_100a,code=sm_100a"]

EXTS = {
    "_raymarching": dict(dir="raymarching", cu="raymarching.cu", extra=[], inc=[]),
    "_gridencoder": dict(dir="gridencoder", cu="gridencoder.cu", extra=[], inc=[]),
    "_shencoder": dict(dir="shencoder", cu="shencoder.cu", extra=[], inc=[]),
    "_ffmlp": dict(dir="ffmlp", cu="ffmlp.cu",
                   extra=["--expt-extended-lambda", "--expt-relaxed-constexpr",
                          "-Xcompiler=-mf16c", "-Xcompiler=-Wno-float-conversion",
                          "-Xcompiler=-fno-strict-aliasing"],
                   inc=["dependencies/cutlass/include", "dependencies/cutlass/tools/util/include"]),
}


def build(name: str) -> str:
    from torch.utils.cpp_extension import load
    spec = EXTS[name]
    src = os.path.join(REF, spec["dir"], "src")
    bdir = os.path.join(OUT, "build", name)
    os.makedirs(bdir, exist_ok=True)
    # torch refuses to pick an arch without a visible GPU unless told; our -gencode wins anyway
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    load(name=name,
         sources=[os.path.join(src, spec["cu"]), os.path.join(src, "bindings.cpp")],
         extra_cflags=["-O3", "-std=c++17"],
         extra_cuda_cflags=COMMON + spec["extra"],
         extra_include_paths=[os.path.join(REF, spec["dir"], p) for p in spec["inc"]],
         build_directory=bdir, verbose=True, is_python_module=False)
    so = os.path.join(bdir, name + ".so")
    dst = os.path.join(OUT, name + ".so")
    shutil.copyfile(so, dst)
    return dst


# The reference's own Python on the path, unmodified, so that the GPU box (which has no /root/reference) can run its methods on the
# rebuilt extensions: tests/refext.reference_renderer_class() imports nerf/renderer.py from here behind three stub modules.  Like the
# .so files these are build OUTPUTS: oracle/_ref/ is git-ignored, nothing of it enters the repository's history.
PYREF = ["nerf/renderer.py", "nerf/network_ff.py", "validate.py",
         # the reference's autograd Functions and module glue (tests/refext.reference_training_stack(): the gradient oracle of
         # the differentiable path runs the reference's OWN backward kernels through its OWN Python)
         "encoding.py", "activation.py", "gridencoder/__init__.py", "gridencoder/grid.py", "shencoder/__init__.py",
         "shencoder/sphere_harmonics.py", "ffmlp/__init__.py", "ffmlp/ffmlp.py", "raymarching/__init__.py",
         "raymarching/raymarching.py"]


def copy_python():
    dst_root = os.path.join(OUT, "pyref")
    for rel in PYREF:
        dst = os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF, rel), dst)
    print("[build_ref] ->", dst_root, flush=True)


def main(argv):
    if not os.path.isdir(REF):
        print(f"[build_ref] {REF} not present (GPU box?) - using prebuilt oracle/_ref if any")
        return 0
    copy_python()
    names = argv or list(EXTS)
    for n in names:
        if not n.startswith("_"):
            n = "_" + n
        dst = os.path.join(OUT, n + ".so")
        if os.path.exists(dst) and not os.environ.get("NGP_REF_REBUILD"):
            print(f"[build_ref] {dst} exists, skip")
            continue
        print(f"[build_ref] building {n} ...", flush=True)
        print("[build_ref] ->", build(n), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))

"""Builds libngp_b200.so (hand-written sm_100a CUDA behind the C ABI of include/ngp_b200.h) in-tree.

    python -m nerfsafetyvalidation_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU; the resulting .so is git-ignored but travels with the gpurun snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(PKG, "_lib", "obj")
LIB = os.path.join(PKG, "_lib", "libngp_b200.so")

SOURCES = ["common.cu", "ops_raymarch.cu", "ops_encode.cu", "ops_mlp.cu", "ops_occupancy.cu", "fused.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr", "--expt-extended-lambda", "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def _newer(target, deps):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = os.environ.get("NVCC", "nvcc")
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "ngp_b200.h"))
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]

    def compile_one(src):
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        if not force and _newer(obj, [src] + headers):
            return obj
        # NGP_EXTRA_NVCC_FLAGS: development builds only (e.g. -DNGP_SPIN_TRAP: a barrier that never completes traps instead of hanging)
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("NGP_EXTRA_NVCC_FLAGS", "").split(), "-Xptxas", "-v", "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = os.path.join(OBJ, os.path.basename(src)[:-3] + ".ptxas.log")
        with open(log, "w") as f:
            f.write("".join(ln for ln in r.stderr.splitlines(True) if "Compile time" not in ln))   # keep the tracked log stable
        if r.returncode != 0:
            sys.stderr.write(r.stderr)
            raise RuntimeError(f"nvcc failed on {src}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
        objs = list(ex.map(compile_one, srcs))
    if force or not _newer(LIB, objs):
        subprocess.check_call([nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                               "-Xcompiler", "-fPIC", "-lcudart"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))

"""Build libmfv2d_cuda.so (sm_100a) in-tree with nvcc.

``python -m mfv2d_b200.build`` or ``mfv2d_b200.build.build()``.  nvcc cross-compiles without a GPU; the
resulting ``mfv2d_b200/lib/libmfv2d_cuda.so`` is git-ignored but travels with the repository snapshot to
the GPU box.  No PyTorch is involved anywhere in the build or in the library.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(HERE, "build")
LIB = os.path.join(LIBDIR, "libmfv2d_cuda.so")
SOURCES = ["context.cu", "tables.cu", "plan.cu", "assembly.cu", "apply.cu", "element_api.cu", "dense.cu", "system.cu",
           "comm.cu", "projector.cu", "schur_cg.cu", "peer.cu", "precond.cu", "micro.cu", "bicgstab.cu", "pcg.cu", "noalloc.cu", "refine.cu"]
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr",
              "-Xptxas", "-v"]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libmfv2d_cuda.so cannot be built (there is no CPU fallback)")


def _digest() -> str:
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(os.path.dirname(HERE), "include")):
        for name in sorted(os.listdir(root)):
            with open(os.path.join(root, name), "rb") as f:
                h.update(name.encode())
                h.update(f.read())
    h.update(" ".join(ARCH_FLAGS + NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    stamp = os.path.join(LIBDIR, "libmfv2d_cuda.stamp")
    digest = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == digest:
        return LIB
    nvcc = find_nvcc()
    host_cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else None

    def compile_one(src: str) -> tuple[str, str]:
        obj = os.path.join(OBJDIR, src.replace(".cu", ".o"))
        cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if host_cxx:
            cmd[1:1] = ["-ccbin", host_cxx]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{res.stdout}\n{res.stderr}")
        return obj, res.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as pool:
        results = list(pool.map(compile_one, SOURCES))
    log = os.path.join(OBJDIR, "ptxas.log")
    with open(log, "w") as f:
        for obj, err in results:
            f.write(f"==== {os.path.basename(obj)}\n{err}\n")
    if verbose:
        print(open(log).read())
    cmd = [nvcc, *ARCH_FLAGS, "-shared", "-o", LIB, *[o for o, _ in results], "-lcudart", "-ldl"]
    if host_cxx:
        cmd[1:1] = ["-ccbin", host_cxx]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    with open(stamp, "w") as f:
        f.write(digest)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))

"""TEST INFRASTRUCTURE ONLY.  Build tests/emu/_build/libmfv2d_emu.so: the sources of libmfv2d_cuda.so
compiled with plain g++ and -DMFV2D_EMU (see emu_cuda.h) so that the host logic and the kernels' index
arithmetic can be exercised in the GPU-less build container.  Never imported by the mfv2d_b200 package."""
from __future__ import annotations

import hashlib
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "mfv2d_b200", "csrc")
OUT = os.path.join(HERE, "_build")
LIB = os.path.join(OUT, "libmfv2d_emu.so")


def build(force: bool = False) -> str:
    import sys

    sys.path.insert(0, ROOT)
    from mfv2d_b200.build import SOURCES

    os.makedirs(OUT, exist_ok=True)
    h = hashlib.sha256()
    for d in (CSRC, HERE, os.path.join(ROOT, "include")):
        for n in sorted(os.listdir(d)):
            p = os.path.join(d, n)
            if os.path.isfile(p) and not n.endswith(".py"):
                h.update(open(p, "rb").read())
    stamp = os.path.join(OUT, "stamp")
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == h.hexdigest():
        return LIB
    cxx = "/usr/bin/g++"

    def one(src):
        path = os.path.join(CSRC, src) if src.endswith(".cu") else os.path.join(HERE, src)
        obj = os.path.join(OUT, src.replace(".cu", ".o").replace(".cpp", ".o"))
        cmd = [cxx, "-x", "c++", "-std=c++17", "-O1", "-g", "-fPIC", "-fvisibility=hidden", "-DMFV2D_EMU", "-w", "-c", path, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"emu build failed on {src}:\n{r.stderr[-4000:]}")
        return obj

    with ThreadPoolExecutor(8) as pool:
        objs = list(pool.map(one, list(SOURCES) + ["emu_globals.cpp"]))
    r = subprocess.run([cxx, "-shared", "-o", LIB, *objs], capture_output=True, text=True)
    if r.returncode:
        raise RuntimeError(r.stderr)
    open(stamp, "w").write(h.hexdigest())
    return LIB


if __name__ == "__main__":
    print(build(True))

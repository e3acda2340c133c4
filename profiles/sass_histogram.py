"""Opcode histogram per kernel of libmfv2d_cuda.so (cuobjdump -sass), written to profiles/sass_r02.txt.

    python profiles/sass_histogram.py [out]

Per kernel: instruction count and the counts of the opcodes that say which hardware path the kernel uses (fp64 FMA /
tensor-core DMMA, shared and global loads and stores by width, asynchronous copies LDGSTS / bulk copies UBLKCP / TMA
UTMALDG-UTMASTG, tcgen05 UTC*MMA / LDTM / STTM, warp shuffles, barriers), then the library-wide totals.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mfv2d_b200", "lib", "libmfv2d_cuda.so")
KEYS = ["DFMA", "DMUL", "DADD", "DMMA", "LDS", "LDS.64", "LDS.128", "STS", "STS.64", "STS.128", "LDG", "LDG.64", "LDG.128", "STG", "STG.64", "STG.128",
        "LDGSTS", "UBLKCP", "UTMALDG", "UTMASTG", "UTCMMA", "LDTM", "STTM", "SHFL", "BAR", "ATOM", "RED", "LDC", "ULDC", "MUFU", "MEMBAR", "LDL", "STL"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "sass_r02.txt")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur is not None:
            op, mods = m.group(1), m.group(2)
            cur["_total"] += 1
            cur[op] += 1
            if op in ("LDS", "STS", "LDG", "STG"):
                width = "128" if ".128" in mods else "64" if ".64" in mods else None
                if width:
                    cur[f"{op}.{width}"] += 1
            if op.startswith("UTC") and "MMA" in op:
                cur["UTCMMA"] += 1
    names = demangle(list(kernels))
    total = collections.Counter()
    with open(out_path, "w") as f:
        f.write(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} : opcode histogram per kernel (profiles/sass_histogram.py)\n")
        f.write("# LDS/STS/LDG/STG count every width; the .64 / .128 columns are the subsets with that width\n\n")
        for k, c in kernels.items():
            total.update(c)
            name = re.sub(r"\(.*", "", names.get(k, k))
            f.write(f"{name}  [{c['_total']} instructions]\n   ")
            f.write("  ".join(f"{key}={c[key]}" for key in KEYS if c[key]) + "\n")
        f.write(f"\nTOTAL over {len(kernels)} kernels [{total['_total']} instructions]\n   ")
        f.write("  ".join(f"{key}={total[key]}" for key in KEYS) + "\n")
    print(out_path, len(kernels), "kernels;", "  ".join(f"{key}={total[key]}" for key in ("DFMA", "DMMA", "LDGSTS", "UBLKCP", "UTMALDG", "UTMASTG", "UTCMMA", "STG.64", "STG.128")))


if __name__ == "__main__":
    main()

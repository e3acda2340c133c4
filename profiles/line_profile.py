#!/usr/bin/env python
"""Join an `ncu --page source --csv` SASS dump with `nvdisasm --print-line-info` output of the same cubin
and print executed warp instructions and stall samples per CUDA source line.

    python profiles/line_profile.py asm_src.csv asm.dis element_matrix_kernel
"""
import csv, re, sys, collections
src_csv, dis, kname = sys.argv[1:4]
# parse the disassembly: sequence of (line) per instruction within the kernel's text section
lines = open(dis).read().split("\n")
inst_lines = []
cur = None
inside = False
for ln in lines:
    if ln.startswith("\t.section") or ln.startswith(".section"):
        inside = (".text." in ln and kname in ln)
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.search(r"/\*[0-9a-f]{4,6}\*/", ln):
        inst_lines.append(cur)
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = rows[2:]
print(f"sass instructions: ncu {len(body)}, nvdisasm {len(inst_lines)}", file=sys.stderr)
agg = collections.defaultdict(lambda: [0, 0])
tot = [0, 0]
for r, l in zip(body, inst_lines):
    n, s = int(r[ii]), int(r[isamp])
    agg[l][0] += n; agg[l][1] += s
    tot[0] += n; tot[1] += s
for l, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[4]) if len(sys.argv)>4 else 60]:
    print(f"{str(l):40s} inst {n:12d} {100*n/tot[0]:5.1f}%   samples {s:8d} {100*s/tot[1]:5.1f}%")

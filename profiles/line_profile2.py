#!/usr/bin/env python
"""Like line_profile.py, plus shared-memory wavefronts (actual / ideal) and the main stall reasons per CUDA line.

    python profiles/line_profile2.py src.csv kernel.dis <kernel name fragment> [rows] [sort: samples|inst|wave]
"""
import csv, re, sys, collections
src_csv, dis, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
key = sys.argv[5] if len(sys.argv) > 5 else "samples"
inst_lines, cur, inside = [], None, False
for ln in open(dis).read().split("\n"):
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
col = {h: i for i, h in enumerate(hdr)}
body = rows[2:]
assert len(body) == len(inst_lines), (len(body), len(inst_lines))
stalls = ["stall_barrier", "stall_long_sb", "stall_short_sb", "stall_mio", "stall_math", "stall_wait", "stall_lg", "stall_not_selected", "stall_branch_resolving"]
agg = collections.defaultdict(lambda: collections.Counter())
tot = collections.Counter()
def num(r, name):
    try:
        return int(float(r[col[name]] or 0))
    except ValueError:
        return 0
for r, l in zip(body, inst_lines):
    d = {"inst": num(r, "Instructions Executed"), "samples": num(r, "# Samples"), "wave": num(r, "L1 Wavefronts Shared"),
         "ideal": num(r, "L1 Wavefronts Shared Ideal")}
    for s in stalls:
        d[s] = num(r, s)
    agg[l].update(d)
    tot.update(d)
print(f"total: inst {tot['inst']}  samples {tot['samples']}  shared wavefronts {tot['wave']} (ideal {tot['ideal']})")
print("stalls: " + "  ".join(f"{s[6:]} {100 * tot[s] / max(tot['samples'], 1):.1f}%" for s in stalls))
for l, d in sorted(agg.items(), key=lambda kv: -kv[1][key])[:top]:
    big = sorted(stalls, key=lambda s: -d[s])[:2]
    print(f"{str(l):24s} inst {100 * d['inst'] / tot['inst']:5.1f}%  samples {100 * d['samples'] / tot['samples']:5.1f}%  "
          f"wave {100 * d['wave'] / max(tot['wave'], 1):5.1f}% (x{d['wave'] / max(d['ideal'], 1):.1f} ideal)  " +
          " ".join(f"{s[6:]}:{d[s]}" for s in big))

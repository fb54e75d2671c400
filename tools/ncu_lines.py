"""Aggregate an ncu SASS source page (csv) by CUDA source line using nvdisasm -g line markers.
usage: python tools/ncu_lines.py <ncu_sass.csv> <nvdisasm_g.txt> <kernel mangled substring> [topN]"""
import csv
import re
import sys
from collections import defaultdict

sass_csv, dis_txt, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
# 1. instruction index -> source line from nvdisasm
lines = open(dis_txt).read().split("\n")
in_k = False
cur = None
idx2line = []
for ln in lines:
    if ln.startswith(".text.") and ln.endswith(":"):
        in_k = kname in ln
        continue
    if not in_k:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        idx2line.append(cur)
# 2. ncu rows for the first kernel instance
rows = list(csv.reader(open(sass_csv)))
hdr = None
agg = defaultdict(lambda: defaultdict(float))
n = 0
started = 0
for r in rows:
    if r and r[0] == "Kernel Name":
        started += 1
        if started > 1:
            break
        continue
    if r and r[0] == "Address":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    d = dict(zip(hdr, r))
    line = idx2line[n] if n < len(idx2line) else None
    n += 1
    a = agg[line]
    a["samples"] += float(d["# Samples"] or 0)
    a["inst"] += float(d["Instructions Executed"] or 0)
    for k in hdr:
        if k.startswith("stall_") and "Not Issued" not in k:
            a[k] += float(d[k] or 0)
print("instructions in ncu: %d, in nvdisasm: %d" % (n, len(idx2line)))
tot = sum(a["samples"] for a in agg.values())
toti = sum(a["inst"] for a in agg.values())
print("total samples %d, total warp-instructions %d" % (tot, toti))
for line, a in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:top]:
    st = sorted(((k, v) for k, v in a.items() if k.startswith("stall_") and v > 0), key=lambda kv: -kv[1])[:3]
    print("%-28s samples %5.1f%%  inst %5.1f%%  %s" % (line, 100 * a["samples"] / tot, 100 * a["inst"] / toti,
                                                      ", ".join("%s=%d" % (k[6:], v) for k, v in st)))

"""Aggregate an `ncu --page source --csv --print-source sass,cuda` dump by (file, CUDA source line).
usage: python profiles/srcagg.py dump.csv [kernel-substring] [top]
Only the per-source-line summary rows are used (rows whose first column is a line number)."""
import collections
import csv
import os
import sys

path = sys.argv[1]
want = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
def _rows(path):
    """ncu does not escape quotes inside the source column (inline asm): split on '","' and keep the fixed-width tail"""
    width = None
    for line in open(path, errors="replace"):
        line = line.rstrip("\n")
        if not (line.startswith('"') and line.endswith('"')):
            continue
        f = line[1:-1].split('","')
        if f[0] == "Line No":
            width = len(f)
        elif width and len(f) > width and f[0].isdigit():
            f = [f[0], '","'.join(f[1:len(f) - width + 2])] + f[len(f) - width + 2:]
        yield f


rows = list(_rows(path))
fn = None
hdr = None
cur_file = ""
agg = {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = os.path.basename(r[1])
        continue
    if len(r) == 2 and r[0] == "Function Name":
        fn = r[1]
        hdr = None
        continue
    if r and r[0] == "Line No":
        hdr = r
        col = {}
        for i, h in enumerate(hdr):
            col.setdefault(h, i)
        continue
    if hdr is None or fn is None or want not in fn or len(r) < 10 or not r[0].isdigit():
        continue

    def g(name):
        try:
            return int(r[col[name]] or 0)
        except (ValueError, KeyError):
            return 0
    stalls = {h[6:]: g(h) for h in col if h.startswith("stall_") and "Not Issued" not in h}
    agg[(fn.split("(")[0], cur_file, int(r[0]))] = dict(
        samples=g("# Samples"), inst=g("Instructions Executed"), tinst=g("Thread Instructions Executed"),
        wf=g("L1 Wavefronts Shared"), wfx=g("L1 Wavefronts Shared Excessive"), src=r[1][:90], stalls=stalls)
by_fn = collections.defaultdict(list)
for k, a in agg.items():
    by_fn[k[0]].append((k, a))
for f, items in by_fn.items():
    ts = sum(a["samples"] for _, a in items) or 1
    ti = sum(a["inst"] for _, a in items) or 1
    tt = sum(a["tinst"] for _, a in items)
    tw = sum(a["wf"] for _, a in items)
    print(f"== {f}: samples {ts} warp-inst {ti} lanes/inst {tt / ti:.1f} smem wavefronts {tw}")
    items.sort(key=lambda x: -x[1]["samples"])
    for (ff, file, l), a in items[:top]:
        st = sorted(a["stalls"].items(), key=lambda x: -x[1])[:2]
        sts = " ".join(f"{n}:{v}" for n, v in st if v)
        lanes = a["tinst"] / a["inst"] if a["inst"] else 0
        print(f"  {file[:14]:14s}:{l:<4d} samp {a['samples'] / ts * 100:5.1f}% inst {a['inst'] / ti * 100:5.1f}% lanes {lanes:4.1f} "
              f"wf {a['wf'] / max(tw, 1) * 100:4.1f}% (x{a['wfx']})  [{sts}]  {a['src']}")

"""Aggregate an `ncu --page source --csv --print-source sass,cuda` dump by PHASE (source-line ranges of one file).
usage: python profiles/phaseagg.py dump.csv kernel-substring file-prefix name:lo:hi [name:lo:hi ...]  [--n particles]"""
import collections
import csv
import os
import sys

path, want, fpre = sys.argv[1:4]
phases = []
npart = 8388608
for a in sys.argv[4:]:
    if a.startswith("--n="):
        npart = int(a[4:])
    else:
        n, lo, hi = a.split(":")
        phases.append((n, int(lo), int(hi)))
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
fn = hdr = None
cur = ""
agg = collections.defaultdict(lambda: [0, 0, 0])
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = os.path.basename(r[1])
        continue
    if len(r) == 2 and r[0] == "Function Name":
        fn, hdr = r[1], None
        continue
    if r and r[0] == "Line No":
        hdr, col = r, {}
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
    line = int(r[0])
    ph = "other:" + cur[:16]
    if cur.startswith(fpre):
        for n, lo, hi in phases:
            if lo <= line <= hi:
                ph = n
                break
    a = agg[ph]
    a[0] += g("# Samples")
    a[1] += g("Instructions Executed")
    a[2] += g("Thread Instructions Executed")
ts = sum(a[0] for a in agg.values()) or 1
ti = sum(a[1] for a in agg.values()) or 1
for k, a in sorted(agg.items(), key=lambda x: -x[1][0]):
    print(f"{k:28s} samples {a[0] / ts * 100:5.1f}%  inst {a[1] / ti * 100:5.1f}% ({a[1] / npart:6.2f} warp-inst/particle) lanes {a[2] / max(a[1], 1):4.1f}")
print(f"total warp-inst/particle {ti / npart:.2f}")

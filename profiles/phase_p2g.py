"""Phase breakdown (samples, instructions, top stalls) of k_mpm_p2g from an `ncu --page source --csv --print-source sass,cuda`
dump; phases are found from the marker comments of b2_mpm_p2gw.cuh.  usage: phase_p2g.py dump.csv [particles]"""
import collections
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
src = open(os.path.join(HERE, "..", "taichi_b200", "csrc", "b2_mpm_p2gw.cuh")).read().split("\n")


def find(pat, start=0):
    for i in range(start, len(src)):
        if re.search(pat, src[i]):
            return i + 1
    raise KeyError(pat)


marks = [("helpers", 1), ("stencil", find(r"void p2w_stencil")), ("kernel_head", find(r"__global__")),
         ("count", find(r"---- count:")), ("scan+units", find(r"---- scan:")), ("order", find(r"---- order:")), ("roundsetup", find(r"---- rounds:")),
         ("main", find(r"pragma unroll 1", find(r"---- rounds:"))), ("padflush", find(r"---- pad:")), ("globalflush", find(r"---- flush:"))]


def phase(line):
    name = marks[0][0]
    for n, l in marks:
        if line >= l:
            name = n
    return {"stencil": "main", "helpers": "main"}.get(name, name)


def rows(path):
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


npart = int(sys.argv[2]) if len(sys.argv) > 2 else 8388608
hdr = fn = None
cur = ""
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
for r in rows(sys.argv[1]):
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
    if hdr is None or fn is None or "k_mpm_p2g" not in fn or len(r) < 10 or not r[0].isdigit():
        continue
    ph = phase(int(r[0])) if cur.startswith("b2_mpm_p2gw") else "main(f32x2)" if cur.startswith("b2_f32x2") else "other:" + cur[:14]
    a = agg[ph]
    try:
        a[0] += int(r[col["# Samples"]] or 0)
        a[1] += int(r[col["Instructions Executed"]] or 0)
    except ValueError:
        pass
    for h, i in col.items():
        if h.startswith("stall_") and "Not Issued" not in h:
            try:
                a[2][h[6:]] += int(r[i] or 0)
            except ValueError:
                pass
ts = sum(a[0] for a in agg.values()) or 1
ti = sum(a[1] for a in agg.values()) or 1
for k, a in sorted(agg.items(), key=lambda x: -x[1][0]):
    t = sum(a[2].values()) or 1
    print(f"{k:20s} samples {a[0] / ts * 100:5.1f}%  inst {a[1] / ti * 100:5.1f}% ({a[1] / npart:5.2f} warp-inst/particle)  "
          + " ".join(f"{kk}:{v / t * 100:.0f}%" for kk, v in a[2].most_common(4)))
print(f"total {ti / npart:.2f} warp-inst/particle")

"""Aggregate an `ncu --page source --csv --print-source sass,cuda` dump by CUDA source line.
usage: python profiles/srcagg.py dump.csv [kernel-substring] [top]"""
import csv, sys, collections
path = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
rows = list(csv.reader(open(path, errors="replace")))
fn = None; hdr = None; agg = collections.defaultdict(lambda: [0, 0, 0, 0.0]); cur_file = ""
srcline = {}
for r in rows:
    if len(r) == 2 and r[0] == "Function Name": fn = r[1]; hdr = None; continue
    if len(r) == 2 and r[0] == "File Path": cur_file = r[1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or fn is None or want not in fn or len(r) < 10 or r[0] == '': continue
    d = dict(zip(hdr, r))
    # sass rows carry "Source" twice: first col = cuda line text? use positional
    try:
        samples = int(r[hdr.index("# Samples")] or 0); inst = int(r[hdr.index("Instructions Executed")] or 0)
    except ValueError:
        continue
    key = (fn.split("(")[0], r[0])
    a = agg[key]; a[0] += samples; a[1] += inst
    try: a[2] += int(r[hdr.index("L1 Wavefronts Shared")] or 0)
    except ValueError: pass
    srcline.setdefault(key, r[1][:110])
tot = collections.Counter(); toti = collections.Counter()
for (f, l), a in agg.items(): tot[f] += a[0]; toti[f] += a[1]
for f in tot:
    print(f"== {f}: samples {tot[f]} warp-inst {toti[f]}")
    items = sorted(((k, a) for k, a in agg.items() if k[0] == f), key=lambda x: -x[1][0])[:top]
    for (ff, l), a in items:
        print(f"  L{l:>5} samp {a[0]/max(tot[f],1)*100:5.1f}% inst {a[1]/max(toti[f],1)*100:5.1f}% smemwf {a[2]:>9}  {srcline[(ff,l)]}")

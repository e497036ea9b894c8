#!/usr/bin/env python
"""Group executed warp-instructions / stall samples of one kernel by source-line ranges.
Usage: tools/ncu_regions.py X.ncu-rep <function-substring> name:lo-hi [name:lo-hi ...]   (lines of the kernel's main .cu file)"""
import csv, io, subprocess, sys
rep, want = sys.argv[1], sys.argv[2]
regions = [(a.split(":")[0], int(a.split(":")[1].split("-")[0]), int(a.split(":")[1].split("-")[1])) for a in sys.argv[3:]]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur_file = cur_fn = hdr = None
acc = {}
tot = [0.0, 0.0, 0.0]
for row in csv.reader(io.StringIO(txt)):
    if not row: continue
    if row[0] == "File Path": cur_file = row[1]; continue
    if row[0] == "Function Name": cur_fn = row[1]; hdr = None; continue
    if row[0] == "Line No": hdr = row; ix = {}; [ix.setdefault(n, i) for i, n in enumerate(hdr)]; continue
    if hdr is None or row[0] in ("", "...") or want not in cur_fn: continue
    try: line = int(row[0])
    except ValueError: continue
    def fl(n):
        v = row[ix[n] - len(hdr)]  # index from the end: unescaped quotes in the source column can split it
        try: return float(v)
        except ValueError: return 0.0
    i, t, s = fl("Instructions Executed"), fl("Thread Instructions Executed"), fl("# Samples")
    name = "other:" + cur_file.split("/")[-1]
    if cur_file.endswith(".cu"):
        for n, lo, hi in regions:
            if lo <= line <= hi: name = n; break
    a = acc.setdefault(name, [0.0, 0.0, 0.0]); a[0] += i; a[1] += t; a[2] += s
    tot[0] += i; tot[1] += t; tot[2] += s
print(f"total warp-inst {tot[0]:.4g}  thread-inst {tot[1]:.4g}  samples {tot[2]:.0f}")
for n, a in sorted(acc.items(), key=lambda x: -x[1][0]):
    print(f"  {n:28s} inst {a[0]/tot[0]*100:5.1f}%  lanes {a[1]/max(a[0],1):5.1f}  samples {a[2]/tot[2]*100:5.1f}%")

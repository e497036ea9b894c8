#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv` per kernel function: source lines
ranked by stall samples / executed instructions and the stall-reason totals.
Usage: tools/ncu_source_summary.py X.ncu-rep [function-substring] [top-n]"""
import csv, io, subprocess, sys

def main():
    rep = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
    funcs = {}  # function name -> list of (file, line, src, row, hdr)
    cur_file = cur_fn = hdr = None
    for row in csv.reader(io.StringIO(txt)):
        if not row: continue
        if row[0] == "File Path": cur_file = row[1]; continue
        if row[0] == "Function Name": cur_fn = row[1]; hdr = None; continue
        if row[0] == "Line No": hdr = row; continue
        if hdr is None or row[0] in ("", "..."): continue
        lst = funcs.setdefault(cur_fn, {})
        key = (cur_file, row[0])
        if key in lst:
            old = lst[key][3]
            for i in range(4, len(row)):
                try: old[i] = str(float(old[i]) + float(row[i]))
                except ValueError: pass
        else: lst[key] = (cur_file, row[0], row[1], list(row), hdr)
    for fn, rows in funcs.items():
        if want and want not in fn: continue
        rows = list(rows.values())
        hdr = rows[0][4]
        ix = {}
        for i, n in enumerate(hdr): ix.setdefault(n, i)
        def f(r, n):
            v = r[ix[n]] if n in ix else ""
            try: return float(v)
            except ValueError: return 0.0
        tot_s = sum(f(r[3], "# Samples") for r in rows); tot_i = sum(f(r[3], "Instructions Executed") for r in rows)
        tot_t = sum(f(r[3], "Thread Instructions Executed") for r in rows)
        print(f"=== {fn}\n    samples {tot_s:.0f}, warp-inst {tot_i:.4g}, thread-inst {tot_t:.4g}, avg active threads/inst {tot_t/max(tot_i,1):.2f}")
        stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
        st = sorted(((sum(f(r[3], n) for r in rows), n) for n in stalls), reverse=True)
        print("    stall samples:", ", ".join(f"{n[6:]} {v/max(tot_s,1)*100:.1f}%" for v, n in st[:9]))
        print("    %samples %inst  thr/inst  top-stall   file:line | source")
        import os
        sortkey = os.environ.get("SORT", "# Samples")
        for file, line, src, r, _ in sorted(rows, key=lambda x: -f(x[3], sortkey))[:topn]:
            top = sorted(((f(r, n), n[6:]) for n in stalls), reverse=True)[0]
            ni = f(r, "Instructions Executed")
            print(f"    {f(r,'# Samples')/max(tot_s,1)*100:6.1f} {ni/max(tot_i,1)*100:6.1f} {ni:10.3g}   {f(r,'Thread Instructions Executed')/max(ni,1):5.1f}  {top[1]:>14}  {file.split('/')[-1]}:{line} | {src.strip()[:100]}")

main()

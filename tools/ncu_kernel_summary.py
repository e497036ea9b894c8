#!/usr/bin/env python
"""Per-launch summary of an `ncu --set full` report: the metrics DESIGN.md / BASELINE.json's north_star ask for (achieved L2 / HBM
GB/s, L1 hit rate, warp execution efficiency, issue-slot utilisation, occupancy, registers).  Usage: ncu_kernel_summary.py X.ncu-rep [--json out.json]"""
import csv, io, json, subprocess, sys
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {n: i for i, n in enumerate(hdr)}
M = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "duration"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
     ("launch__registers_per_thread", "regs"), ("launch__shared_mem_per_block_static", "smem_static_B"),
     ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved_occupancy_pct"),
     ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct_peak"),
     ("lts__t_bytes.sum", "l2_bytes"), ("lts__t_sector_hit_rate.pct", "l2_hit_pct"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_pct_peak"),
     ("l1tex__t_sector_hit_rate.pct", "l1_hit_pct"), ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1_pct_peak"),
     ("smsp__thread_inst_executed_per_inst_executed.ratio", "active_threads_per_inst(of 32)"),
     ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_slot_util_pct"), ("sm__inst_executed.sum", "warp_inst"),
     ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct_peak"), ("smsp__inst_executed.avg.per_cycle_active", "ipc_per_smsp"),
     ("sm__cycles_elapsed.avg.per_second", "sm_clock")]
out = []
for r in data:
    d = {}
    for name, short in M:
        if name in ix:
            d[short] = r[ix[name]] + ("" if short == "kernel" else " " + units[ix[name]])
    out.append(d)
    dur = float(r[ix["gpu__time_duration.sum"]]); du = units[ix["gpu__time_duration.sum"]]
    sec = dur * {"us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1}.get(du, 1e-6)
    def bytes_of(n):
        v = float(r[ix[n]]); u = units[ix[n]]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    d["dram_bytes_total"] = bytes_of("dram__bytes_read.sum") + bytes_of("dram__bytes_write.sum")
    d["dram_GBps"] = d["dram_bytes_total"] / sec / 1e9
    if "lts__t_bytes.sum" in ix:
        d["l2_GBps"] = bytes_of("lts__t_bytes.sum") / sec / 1e9
for d in out:
    print("; ".join(f"{k}={v if not isinstance(v, float) else round(v, 2)}" for k, v in d.items()))
if "--json" in sys.argv:
    json.dump(out, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)

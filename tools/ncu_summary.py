#!/usr/bin/env python
"""Summary of an `ncu --set full` capture of tools/ncu_target.py (the step's six MLP launches + three inference forwards):
usage: ncu -i X.ncu-rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv out_summary.csv out_traffic.json"""
import csv, json, sys

raw, out_csv, out_json = sys.argv[1:4]
rows = list(csv.reader(open(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
cols = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
        "FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size"]
idx = [hdr.index(c) for c in cols]
ki = hdr.index("Kernel Name")
labels = ["fwd<full,train> coarse 4096x64 rows", "fwd<full,train> fine 4096x128 rows", "dgrad<full> fine", "wgrad fine",
          "dgrad<full> coarse", "wgrad coarse", "fwd<full,inference> 4096x128 rows", "fwd<full,inference> 4096x128 rows",
          "fwd<full,inference> 4096x128 rows"]
with open(out_csv, "w", newline="") as f:
    f.write("# ncu --set full --clock-control none -k regex:mlp_ -s 12 -c 9 python tools/ncu_target.py (B200, final round-2 library); "
            "one training step of 4096 rays 64+64 + three inference forwards; ncu serialises and replays launches: durations are "
            "cold-cache at profiler clocks, compare SHARES and byte counts, not absolutes\n")
    w = csv.writer(f)
    w.writerow(["launch", "kernel"] + [f"{c} [{units[i]}]" for c, i in zip(cols, idx)])
    for lab, r in zip(labels, data):
        w.writerow([lab, r[ki]] + [r[i] for i in idx])


def gb(r, name):
    i = hdr.index(name)
    v = float(r[i].replace(",", ""))
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[units[i]]


def per_launch(sel):
    return sum(gb(data[i], "dram__bytes_read.sum") + gb(data[i], "dram__bytes_write.sum") for i in sel) / len(sel)


json.dump({"dram_bytes_per_launch": {"3": per_launch([0, 1]), "5": per_launch([2, 4]), "7": per_launch([3, 5])},
           "note": "average over the coarse (4096x64 rows) and fine (4096x128 rows) launch of one training step; ncu --set full "
                   "of the final round-2 library, profiles/r02_ncu_mlp_summary.csv"}, open(out_json, "w"), indent=1)

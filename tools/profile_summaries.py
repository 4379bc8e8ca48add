#!/usr/bin/env python
"""Turns the ncu outputs of a gpurun call into the tracked summaries under profiles/.

    python tools/profile_summaries.py <launches.csv> <full.ncu-rep> <tag>

  <tag>_launches_bench_r4.csv       the launch list as ncu wrote it (gpu__time_duration.sum per launch)
  <tag>_launch_shares_bench_r4.csv  per kernel: launches, total us, share of the build kernels, last launch us
  <tag>_ncu_full_summary.csv        selected metrics of the `--set full` capture, one row per captured launch
  ncu_summary.json                  what bench.py quotes beside its live numbers: per kernel the DRAM bytes per launch and the
                                    FP64-pipe busy share of that capture (last captured launch of each kernel), with the tag
"""
import collections, csv, io, json, re, shutil, subprocess, sys

launches, rep, tag = sys.argv[1:4]
shutil.copy(launches, "profiles/%s_launches_bench_r4.csv" % tag)
rows = [r for r in csv.reader(l for l in open(launches) if l.startswith('"'))]
hdr, rows = rows[0], rows[1:]
ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace(",", ";")
    a = agg.setdefault(name, [0, 0.0, 0.0])
    a[0] += 1; a[1] += float(r[iv]) / 1e3; a[2] = float(r[iv]) / 1e3
build = {k: v for k, v in agg.items() if "k_fp64_rate" not in k and "k_unary" not in k}
tot = sum(v[1] for v in build.values())
with open("profiles/%s_launch_shares_bench_r4.csv" % tag, "w") as f:
    f.write("kernel,launches,total_us,share_of_build_kernels,last_launch_us\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("%s,%d,%.1f,%s,%.1f\n" % (k, v[0], v[1], ("%.4f" % (v[1] / tot)) if k in build else "", v[2]))
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active", "smsp__inst_executed.sum",
        "sm__cycles_elapsed.max", "lts__t_sector_hit_rate.pct"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-units", "base"], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
h = rr[0]
stalls = [c for c in h if c.startswith("smsp__average_warps_issue_stalled") and c.endswith("per_issue_active.ratio")]
cols = ["ID", "Kernel Name", "Grid Size", "Block Size"] + [c for c in want if c in h] + stalls
with open("profiles/%s_ncu_full_summary.csv" % tag, "w", newline="") as f:
    w = csv.writer(f)
    for r in rr:
        w.writerow([r[h.index(c)] for c in cols])
print(open("profiles/%s_launch_shares_bench_r4.csv" % tag).read()[:1800])

# bench.py's copy of the hardware counters (keys = the library's timing names)
NAMES = {"k_dp_hsamples": "dp_hsamples", "k_bp_quad_fused": "bipolar_quad_fused", "k_band_fill": "band_fill", "k_quad_cells": "quad_cells"}
def num(r, c):
    return float(r[h.index(c)].replace(",", "")) if c in h and r[h.index(c)] not in ("", "n/a") else None
summary = {"capture": tag, "command": "ncu --set full --clock-control none --import-source on (one launch per kernel of `python bench.py --steps 2 --warmup 1`, -r 4 --r_dp 0.2, one GPU)", "kernels": {}}
for r in rr[2:]:
    kn = r[h.index("Kernel Name")]
    for pat, key in NAMES.items():
        if pat in kn:
            rd, wr = num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum")
            summary["kernels"][key] = {
                "kernel": re.sub(r"\(.*", "", kn).replace("void ", ""), "duration_us_under_ncu": (num(r, "gpu__time_duration.sum") or 0.0) / 1e3,  # base unit: ns,
                "dram_bytes_per_launch": (rd or 0.0) + (wr or 0.0), "dram_bytes_read": rd, "dram_bytes_write": wr,
                "pipe_fp64_pct": num(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
                "issue_active_pct": num(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                "registers_per_thread": num(r, "launch__registers_per_thread"),
                "warps_active_pct": num(r, "sm__warps_active.avg.pct_of_peak_sustained_active")}
json.dump(summary, open("profiles/ncu_summary.json", "w"), indent=1, sort_keys=True)
print(json.dumps(summary["kernels"], indent=1)[:1500])

#!/usr/bin/env python
"""Turns the ncu outputs of a gpurun call into the tracked summaries under profiles/.

    python tools/profile_summaries.py <launches.csv> <full.ncu-rep> <tag>

  <tag>_launches_bench_r4.csv       the launch list as ncu wrote it (gpu__time_duration.sum per launch)
  <tag>_launch_shares_bench_r4.csv  per kernel: launches, total us, share of the build kernels, last launch us
  <tag>_ncu_full_summary.csv        selected metrics of the `--set full` capture, one row per captured launch
"""
import collections, csv, io, re, shutil, subprocess, sys

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
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
h = rr[0]
stalls = [c for c in h if c.startswith("smsp__average_warps_issue_stalled") and c.endswith("per_issue_active.ratio")]
cols = ["ID", "Kernel Name", "Grid Size", "Block Size"] + [c for c in want if c in h] + stalls
with open("profiles/%s_ncu_full_summary.csv" % tag, "w", newline="") as f:
    w = csv.writer(f)
    for r in rr:
        w.writerow([r[h.index(c)] for c in cols])
print(open("profiles/%s_launch_shares_bench_r4.csv" % tag).read()[:1800])

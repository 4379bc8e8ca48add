#!/usr/bin/env python
"""Benchmark of the supergrid hot path (metric: supergrid points/s, coordinates + metrics, FP64).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the CPU restatement of the reference (oracle)

One step = one build of the whole workload supergrid (every rank builds its latitude row-slab of every
sub-grid, then the global-area check: device reduction + one all-reduce when N > 1).
N=1 workload: BASELINE.json configs[1], `-r 4 --r_dp 0.2` (2243 x 2881 vertices).  For N > 1 the resolution grows
with N so that the vertices per GPU stay ~6.5-7.3e6 (weak scaling): N=2: r=6, N=4: r=8, N=8: r=12.
Rank 0 prints ONE JSON line.  See DESIGN.md "Measurement" for the roofline arithmetic.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "supergrid points/s (coords+metrics, FP64)"
UNIT = "points/s"
WEAK_RESOLUTION = {1: 4, 2: 6, 4: 8, 8: 12}

# ---- algorithmic FP64 work per unit (DESIGN.md "Work per unit"): +,-,* count 1; / counts W_DIV and sqrt W_SQRT,
# the FP64-pipe instructions of the IEEE-correct expansions on sm_100a (fast paths seen in SASS)
W_DIV, W_SQRT = 10, 9
OPS_BP_SAMPLE = 165 + 7 * W_DIV + 3 * W_SQRT                 # bipolar metrics-only sample (k_bp_lattice<1>)
OPS_DP_VERTEX = 87 + 5 * W_DIV + 1 * W_SQRT
OPS_HAVERSINE = 192 + 1 * W_DIV + 1 * W_SQRT
OPS_DP_SAMPLE = 8 * OPS_DP_VERTEX + 4 * OPS_HAVERSINE + 8 + 24   # order-4 differences, 8 projections (k_dp_hsamples<2>)
BYTES_BAND_VERTEX = 48
# DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the default workload's kernels, from the committed
# `ncu --set full` capture profiles/r01e_ncu_full_summary.csv -- not measured in this run; valid for `-r 4 --r_dp 0.2` only.
# The cap kernels are FP64-bound and move almost nothing (O(Ni+Nj) tables in, samples parked in the 126 MB L2).
NCU_TRAFFIC_R4 = {"dp_hsamples": 774144 + 6492672, "bipolar_quad_fused": 352256 + 24576, "band_fill": 326400 + 140988672}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--resolution", type=float, default=0.0, help="override inverse resolution r (default: per --gpus)")
    p.add_argument("--r_dp", type=float, default=0.2)
    p.add_argument("--slabs", type=int, default=0,
                   help="strong-scaling mode for grids larger than one GPU's HBM: cut the grid into V row-slabs; rank g builds "
                        "slabs g, g+N, ... one after the other in one slab's worth of buffers (V %% N == 0)")
    p.add_argument("--no_cpu_baseline", action="store_true")
    p.add_argument("--no_e2e", action="store_true")
    return p.parse_args()


def workload(args):
    r = args.resolution if args.resolution > 0 else WEAK_RESOLUTION.get(args.gpus, 4)
    r = int(r) if float(r).is_integer() else r
    return r, "-r %s --r_dp %s" % (r, args.r_dp)


# ------------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------------
class ClockSampler(object):
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self._stop, self._t = index, [], threading.Event(), None

    def _init_nvml(self):
        import pynvml as N
        N.nvmlInit()
        self._nvml = (N, N.nvmlDeviceGetHandleByIndex(self.index))

    def _run_nvml(self):
        """Fast path: NVML samples every ~2 ms, so that even a 30 ms timed region gets several clock samples."""
        N, h = self._nvml
        mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
        bits = (("hw_slowdown", N.nvmlClocksThrottleReasonHwSlowdown),
                ("hw_thermal_slowdown", N.nvmlClocksThrottleReasonHwThermalSlowdown),
                ("sw_thermal_slowdown", N.nvmlClocksThrottleReasonSwThermalSlowdown),
                ("sw_power_cap", N.nvmlClocksThrottleReasonSwPowerCap))
        while not self._stop.is_set():
            sm = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)
            rs = N.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            self.samples.append([str(sm), str(mx)] + ["Active" if rs & b else "Not Active" for _, b in bits])
            self._stop.wait(0.002)

    def _run(self):
        try:
            if self._nvml is not None:
                self._run_nvml()
                return
        except Exception:
            pass
        while not self._stop.is_set():
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                               "--format=csv,noheader,nounits"], timeout=5).decode().strip()
                self.samples.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        try:
            self._init_nvml()      # before the timed region: library load + handle lookup take tens of ms
        except Exception:
            self._nvml = None
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------------
# CPU arm (oracle): the reference's algorithm restated in C, all host threads
# ------------------------------------------------------------------------------------------------------
def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 for every rank; the CPU arm runs on rank 0 alone and is to use every host
    thread, so the limit is lifted before the OpenMP runtime of the oracle library starts."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    os.environ["OMP_NUM_THREADS"] = str(n)


def cpu_pass(r, r_dp):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import supergrid_cpu
    t0 = time.time()
    out = supergrid_cpu.build_cpu(r, r_dp=r_dp)
    dt = time.time() - t0
    return out["x"].size / dt, dt, out["x"].shape


def cpu_baseline(r, r_dp):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ogg_oracle
    cores = ogg_oracle.num_threads()
    rr = r if r <= 4 else 4   # bounded sample: the 1/4-degree grid costs ~10-30 s of CPU work
    pps, dt, shape = cpu_pass(rr, r_dp)
    return {"value": pps, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "one pass of '-r %s --r_dp %s' (%dx%d vertices, whole grid, %.1f s) through oracle/ (C restatement "
                      "of the reference, OpenMP)" % (rr, r_dp, shape[0], shape[1], dt)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r, wl = workload(args)
    rr = r if r <= 4 else 4
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    use_all_host_threads()
    import ogg_oracle
    cores = ogg_oracle.num_threads()
    for _ in range(min(args.warmup, 1)):
        cpu_pass(rr, args.r_dp)
    steps = max(1, min(args.steps, 5))
    t0 = time.time()
    shape = None
    for _ in range(steps):
        _, _, shape = cpu_pass(rr, args.r_dp)
    dt = (time.time() - t0) / steps
    pps = shape[0] * shape[1] / dt
    line = {"impl": "reference", "metric": METRIC, "value": pps, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": min(args.warmup, 1), "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "-r %s --r_dp %s (%dx%d vertices); CPU arm bounded to r<=4" % (rr, args.r_dp, shape[0], shape[1])},
            "cpu_baseline": {"value": pps, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "whole '-r %s --r_dp %s' grid per step, oracle/ C restatement with OpenMP; the reference "
                                       "itself is pure Python + the absent numpypi and cannot travel to this box" % (rr, args.r_dp)},
            "e2e": {"value": pps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------
def run_ours(args):
    import ctypes
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from ocean_model_grid_generator_b200 import _cabi, device as D, supergrid as SG
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    D.ready()
    lib = _cabi.load()
    r, wl = workload(args)
    flags = vars(G._build_parser().parse_args(wl.split()))
    plan_args = {k: v for k, v in flags.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta",
                                                              "write_subgrid_files")}
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        plan = SG.plan_supergrid(**plan_args)
    V = args.slabs if args.slabs else world
    if V % world:
        raise SystemExit("--slabs must be a multiple of --gpus")
    mine = list(range(rank, V, world))
    sg = SG.Supergrid(plan, mine[0], V, capacity_world=V if args.slabs else None)
    npoints = plan.nyp * plan.nxp
    area_parts = torch.zeros(len(mine), dtype=torch.float64, device="cuda")
    my_slabs = []
    for v in mine:
        my_slabs.append(list(sg.set_slab(v, V).slabs))
    sg.set_slab(mine[0], V)

    def step(concurrent=True):
        for n, v in enumerate(mine):
            if len(mine) > 1:
                sg.set_slab(v, V)
            sg.build(concurrent=concurrent)
            _cabi.call("ogg_sum_strided", D.ptr(sg.area), sg.area.numel(), 1, D.ptr_off(area_parts, n), D.stream())
        if dist is not None:
            dist.all_reduce(area_parts)    # the single global-area all-reduce (NCCL, 8 bytes per slab held)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    total_area = float(area_parts.sum())
    # ---- timed region: K steps, CUDA events on the launching stream, per-kernel events inside --------------
    sampler = ClockSampler(local)
    lib.ogg_timing_reset()
    lib.ogg_timing_enable(2)     # event brackets around the three heavy kernel classes only (4 launches of 21 per step)
    l0 = lib.ogg_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    lib.ogg_timing_enable(0)
    launches = (lib.ogg_launch_count() - l0) // args.steps + (1 if dist is not None else 0)
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    value = npoints * args.steps / (ms * 1e-3)

    def read_kernels():
        out = {}
        for k in range(lib.ogg_timing_kernels()):
            tot, n = ctypes.c_double(), ctypes.c_longlong()
            lib.ogg_timing_get(k, ctypes.byref(tot), ctypes.byref(n))
            if n.value:
                out[lib.ogg_timing_name(k).decode()] = (tot.value, n.value)
        lib.ogg_timing_reset()
        return out

    kid = {lib.ogg_timing_name(k).decode(): k for k in range(lib.ogg_timing_kernels())}
    span_ms, span_n = ctypes.c_double(), ctypes.c_longlong()
    lib.ogg_timing_pair_span(kid["dp_hsamples"], kid["bipolar_quad_fused"], ctypes.byref(span_ms), ctypes.byref(span_n))
    kern = read_kernels()
    # second pass, not part of the metric: sub-grids one after the other on ONE stream with a bracket around EVERY
    # launch -- each kernel's duration when it has the GPU to itself (the per-kernel table and the band-fill roofline)
    lib.ogg_timing_enable(1)
    for _ in range(args.steps):
        step(concurrent=False)
    barrier()
    lib.ogg_timing_enable(0)
    kern_all = read_kernels()

    # ---- end to end through the public API: plan + H2D of the 1-D laws + build + D2H of the six arrays ------
    e2e = None
    if not args.no_e2e and not args.slabs:
        pinned = None
        area_e2e = D.empty(1)
        times = []
        for it in range(4):
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()):
                p2 = SG.plan_supergrid(**plan_args)
                s2 = SG.Supergrid(p2, rank, world)
            if pinned is None:
                pinned = s2.pinned_like()
            s2.build(host=pinned)        # every piece's rows are copied out on its own stream while the caps compute
            # the step's results read on the host: the six arrays (above) and the global-area check (device tree sum,
            # 8 bytes).  No host-side parallel reduction here: an OpenMP pool left spinning after torch's CPU sum
            # slowed the NEXT iteration's launches and PCIe copies by 60% on the 16-core box.
            _cabi.call("ogg_sum_strided", D.ptr(s2.area), s2.area.numel(), 1, D.ptr(area_e2e), D.stream())
            checksum = float(area_e2e[0])                # synchronises the stream: all copies have landed
            torch.cuda.synchronize()
            corner = float(pinned["area"][-1, -1]) if pinned["area"].numel() else 0.0
            if dist is not None:
                dist.barrier()
            times.append(time.perf_counter() - t0)
        te = torch.tensor([min(times[1:])], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        d2h = sum(v.numel() * 8 for v in pinned.values()) + 8
        h2d = sum(8 * (len(pc.p["lam"]) + len(pc.p["phi"])) for pc in plan.pieces if pc.kind == "band") + 8 * (plan.Ni + 1) + 4096
        e2e = {"value": npoints / float(te[0]), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms": float(te[0]) * 1e3, "checksum_area_local": checksum, "host_area_last_cell": corner}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ---------------------------------------------------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, hbm_src = (peaks["hbm_gbs"], "MEASURED_PEAKS.json") if "hbm_gbs" in peaks else (6650.0, "fallback")
    rate = ctypes.c_double()
    lib.ogg_fp64_issue_rate(20000, ctypes.byref(rate))
    fp64_peak_tf = rate.value / 1e3
    units = {}
    for bounds in my_slabs:
        for piece, (c0, c1) in zip(plan.pieces, bounds):
            if c1 <= c0:
                continue
            if piece.kind == "dp":
                units["dp_hsamples"] = units.get("dp_hsamples", 0) + (3 * (c1 - c0) + 1) * (3 * plan.Ni + 1)
            elif piece.kind == "bipolar":
                units["bipolar_quad_fused"] = units.get("bipolar_quad_fused", 0) + \
                    (4 * (c1 - c0) + 1 + (1 if piece.row0 + c1 == piece.p["Nj"] else 0)) * (4 * plan.Ni + 1)
            else:
                units["band_fill"] = units.get("band_fill", 0) + (c1 - c0) * plan.nxp
    # Dominant kernel class: the two FP64-bound cap quadratures.  They run concurrently on two streams and share the
    # SMs, so a bracket around one of them also spans the other's CTAs; what the timed region does give cleanly is the
    # span during which at least one of the pair was in flight (ogg_timing_pair_span) -- `achieved` = the algorithmic
    # ops of both / that span.  Each kernel alone (single-stream pass) is reported beside it.
    roof = None
    kernels_ms = {k: round(v[0] / args.steps, 4) for k, v in sorted(kern_all.items(), key=lambda kv: -kv[1][0])}
    quad_ops = OPS_DP_SAMPLE * units.get("dp_hsamples", 0) + OPS_BP_SAMPLE * units.get("bipolar_quad_fused", 0)

    def alone(name, ops):
        if name not in kern_all or name not in units:
            return None
        per = kern_all[name][0] / args.steps
        ach = ops * units[name] / (per * 1e-3) / 1e12
        return {"kernel": name, "bound": "fp64", "achieved": ach, "peak": fp64_peak_tf, "unit": "TFLOP/s", "frac": ach / fp64_peak_tf,
                "ops_per_unit": ops, "units_per_step": units[name], "kernel_ms_per_step": per,
                "measured_in": "single-stream pass: the kernel has the GPU to itself"}
    roof_dp, roof_bp = alone("dp_hsamples", OPS_DP_SAMPLE), alone("bipolar_quad_fused", OPS_BP_SAMPLE)
    if quad_ops and span_n.value:
        per_step_ms = span_ms.value / args.steps
        ach = quad_ops / (per_step_ms * 1e-3) / 1e12
        names = [n for n in ("dp_hsamples", "bipolar_quad_fused") if n in units]
        roof = {"kernel": " + ".join(names), "bound": "fp64", "achieved": ach, "peak": fp64_peak_tf, "unit": "TFLOP/s",
                "frac": ach / fp64_peak_tf,
                "traffic": sum(NCU_TRAFFIC_R4[n] for n in names) if (r == 4 and world == 1 and not args.slabs) else None,
                "traffic_source": "profiles/r01e_ncu_full_summary.csv (ncu --set full, bytes per launch, both kernels)",
                "ops_per_step": quad_ops, "ops_per_unit": {"dp_hsamples": OPS_DP_SAMPLE, "bipolar_quad_fused": OPS_BP_SAMPLE},
                "units_per_step": {n: units[n] for n in names}, "kernel_ms_per_step": per_step_ms,
                "share_of_step": per_step_ms / (ms / args.steps),
                "measured_in": "timed region: the two cap quadratures co-run on two streams; span during which at least one "
                               "of them is in flight (CUDA events on their streams)",
                "peak_source": "measured in this run: DADD+DMUL issue rate with FMA contraction off (ogg_fp64_issue_rate); "
                               "MEASURED_PEAKS.json has no FP64 entry"}
    elif "band_fill" in kern and "band_fill" in units:
        per_step_ms = kern["band_fill"][0] / args.steps
        ach = BYTES_BAND_VERTEX * units["band_fill"] / (per_step_ms * 1e-3) / 1e9
        roof = {"kernel": "band_fill", "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                "traffic": None, "peak_source": hbm_src, "kernel_ms_per_step": per_step_ms,
                "share_of_step": per_step_ms / (ms / args.steps)}
    roof_hbm = None
    if "band_fill" in kern_all and "band_fill" in units:
        per_step_ms = kern_all["band_fill"][0] / args.steps
        ach = BYTES_BAND_VERTEX * units["band_fill"] / (per_step_ms * 1e-3) / 1e9
        roof_hbm = {"kernel": "band_fill", "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                    "frac": ach / hbm_peak, "peak_source": hbm_src, "kernel_ms_per_step": per_step_ms,
                    "measured_in": "single-stream pass (in the timed region the band fills overlap the cap kernels)"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if args.slabs else "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": "%s: full tripolar supergrid %dx%d vertices (displaced-pole cap + Southern Ocean + Mercator + "
                                   "bipolar cap), x,y,dx,dy,area,angle_dx" % (wl, plan.nyp, plan.nxp),
                       "sharding": ("latitude row-slabs of every sub-grid, %d rank(s); weak scaling: r=4/6/8/12 at N=1/2/4/8" % world)
                       if not args.slabs else ("%d latitude row-slabs of every sub-grid, %d per rank, built one after the other in "
                                               "one slab's buffers (strong scaling: the grid is fixed)" % (V, len(mine))),
                       "cache": "each step rewrites %.0f MB of outputs + scratch, larger than the 126 MB L2; no flush needed"
                                % (npoints * 48 / 1e6),
                       "global_area_m2": total_area},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "roofline_dp_hsamples": roof_dp, "roofline_bipolar_quad": roof_bp, "roofline_band_hbm": roof_hbm,
            "kernels_ms_per_step_single_stream_pass": kernels_ms,
            "kernels_ms_per_step_timed_region": {k: round(v[0] / args.steps, 4) for k, v in kern.items()}, "fp64_issue_peak_tflops_measured": fp64_peak_tf}
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(r, args.r_dp)
    elif world > 1:
        line["cpu_baseline"] = None
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    # exactly ONE line on stdout: libraries that print there (NCCL's version banner) are sent to stderr
    _real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    _print = print

    def print(*args, **kw):  # noqa: A001  (only the JSON lines below call it)
        kw.setdefault("file", _real_stdout)
        kw.setdefault("flush", True)
        _print(*args, **kw)

    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)

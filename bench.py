#!/usr/bin/env python
"""Benchmark of the supergrid hot path (metric: supergrid points/s, coordinates + metrics, FP64).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the CPU restatement of the reference (oracle)

One step = one build of the whole workload supergrid (every rank builds its latitude row-slab of every
sub-grid, then the global-area check: device reduction + one all-reduce when N > 1).
N=1 workload: BASELINE.json configs[1], `-r 4 --r_dp 0.2` (2243 x 2881 vertices).  For N > 1 the resolution grows
with N so that the vertices per GPU stay ~6.5-7.3e6 (weak scaling): N=2: r=6, N=4: r=8, N=8: r=12.
The same line also carries, outside the timed region of the metric: `strong` -- BASELINE configs[2..4] (1/8, 1/30, 1/120
degree) as fixed-size runs in 8 row-slabs on the N GPUs of this launch (area and device checksums must not depend on N);
`halo_midas` (N > 1) -- the supplied-mesh MIDAS entry point with its one-row halo over NCCL send/recv, checked bit for
bit; `cli` (N = 1) -- wall time of the real entry point main() -> write_nc.
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
BYTES_BAND_VERTEX = 48
# BASELINE.json configs[2..4] as fixed-size (strong-scaling) runs: (flags, row-slabs V, timed steps).  V = 8 on every N so
# that the 1/120-degree grid (279 GB) fits one GPU slab by slab and the same slabs are built on 1, 2, 4 and 8 GPUs.
STRONG_CONFIGS = [("r8", "-r 8 --r_dp 0.2", 8, 10), ("r30", "-r 30 --r_dp 0.2 --ensure_nj_even", 8, 5),
                  ("r120", "-r 120 --r_dp 0.2", 8, 3)]


def load_opcount():
    """Per-unit FP64 operation counts DERIVED from the oracle by tools/opcount.py (profiles/opcount.json, committed; a CPU
    test re-derives them).  ops_8d: SURVEY 8(d) weights (every + - * / sqrt compare = 1) on what an implementation must
    execute for the same bits; ops_issue: FP64-pipe instructions of the IEEE expansions (division 8, cached divisor 3,
    square root 8); ops_reference: 8(d) weights on everything the reference evaluates per unique sample (orientation)."""
    d = json.load(open(os.path.join(ROOT, "profiles", "opcount.json")))
    t = d["totals"]
    return {"dp_hsamples": t["dp_hsample"], "bipolar_quad_fused": t["bipolar_sample"]}


def load_ncu_summary():
    """Hardware counters of the committed `ncu --set full` capture (profiles/ncu_summary.json, written by
    tools/profile_summaries.py): per kernel the DRAM bytes per launch and the FP64-pipe busy fraction, with the commit the
    capture was taken at.  Not measured in this run -- reported beside the live numbers and labelled as such."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))
    except Exception:
        return {}


def config_of(wl, nyp, nxp, world, slabs, per_rank):
    """The `config` object: identical for the GPU arm and the CPU reference arm of the same launch."""
    sharding = ("latitude row-slabs of every sub-grid, %d rank(s); weak scaling: r=4/6/8/12 at N=1/2/4/8" % world) if not slabs else (
        "%d latitude row-slabs of every sub-grid, %d per rank, built one after the other in one slab's buffers "
        "(strong scaling: the grid is fixed)" % (slabs, per_rank))
    return {"workload": "%s: full tripolar supergrid %dx%d vertices (displaced-pole cap + Southern Ocean + Mercator + "
                        "bipolar cap), x,y,dx,dy,area,angle_dx" % (wl, nyp, nxp), "sharding": sharding}


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
    p.add_argument("--no_strong", action="store_true", help="skip the fixed-size runs of BASELINE configs[2..4] (1/8, 1/30, 1/120 degree)")
    p.add_argument("--no_cli", action="store_true", help="skip the timing of the CLI entry point (main -> write_nc)")
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


def grid_shape_cpu(r, r_dp):
    """(nyp, nxp) of `-r R --r_dp X` from the sizing rules alone (OGG:969-1195, default flags): no array is built."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import ogg_oracle
    Ni = int(r * 2 * 360)
    phi_s, phi_n = (-68.0, 65.0) if r == 2 else (-66.85954725, 64.05895973)
    nM = ogg_oracle.mercator_phi_1d(Ni, phi_s, phi_n, True, False).shape[0]
    nBP = (119 * 2 if r == 2 else int(60 * r * 2)) + 1
    nSO = int(r * 55) + 1
    Nj_scap = int(r * 40) * 7 // 4
    if r_dp == 0.0:
        nSC = int((-78.0 + 90.0) / (1.0 / r / 2)) + 1
    else:
        jmin = np.ceil(0.49 * Nj_scap)
        nSC = Nj_scap + 1 - int(jmin + np.mod(jmin, 2))
    return (nSC - 1) + (nSO - 1) + (nM - 1) + nBP, Ni + 1


def cpu_pass(r, r_dp, world=1):
    """One timed pass of the CPU arm.  world == 1: the whole grid.  world > 1: rank 0's row-slab share of every sub-grid
    (1/world of the rows of each -- the work one GPU does per step), a bounded representative sample of the larger grid;
    points/s = vertices evaluated / time either way."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import supergrid_cpu
    t0 = time.time()
    out = supergrid_cpu.build_cpu(r, r_dp=r_dp, slab=(0, world) if world > 1 else None)
    dt = time.time() - t0
    _LAST_CPU_PASS[0] = out
    return out["x"].size / dt, dt, out["x"].shape


_LAST_CPU_PASS = [None]


FIELDS = ("x", "y", "dx", "dy", "area", "angle_dx")


def host_checksums(arrays):
    """The six per-array checksums ogg_checksum_accumulate computes on the device (sum over elements of the splitmix64
    finaliser of the element's bit pattern, mod 2**64), computed with numpy integers on host arrays."""
    import numpy as np
    out = []
    for k in FIELDS:
        z = np.ascontiguousarray(arrays[k], dtype=np.float64).reshape(-1).view(np.uint64).copy()
        z += np.uint64(0x9e3779b97f4a7c15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xbf58476d1ce4e5b9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94d049bb133111eb)
        z ^= z >> np.uint64(31)
        out.append("%016x" % int(z.sum(dtype=np.uint64)))
    return out


def own_rows(arrays, k, V):
    """A slab as supergrid_cpu evaluates it carries its top piece's closing vertex row; that row belongs to the next slab
    (to this one only when it is the last): trim x, y, dx, angle_dx to the rows the slab owns, as Supergrid.arrays() does."""
    n = arrays["area"].shape[0] + (1 if k == V - 1 else 0)
    return {f: (a[:n] if f in ("x", "y", "dx", "angle_dx") else a) for f, a in arrays.items()}


def cpu_slab_checksums(r, r_dp, k, V):
    """Checker, not a measurement: the CPU oracle's evaluation of row-slab k of V of the timed grid, as six checksums."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import supergrid_cpu
    out = own_rows(supergrid_cpu.build_cpu(r, r_dp=r_dp, slab=(k, V) if V > 1 else None), k, V)
    return host_checksums(out), list(out["x"].shape)


def cpu_baseline(r, r_dp):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ogg_oracle
    cores = ogg_oracle.num_threads()
    pps, dt, shape = cpu_pass(r, r_dp)
    o1 = {}
    try:
        g = json.load(open(os.path.join(ROOT, "tests", "golden", "configs.json")))["res0.25_full"]
        o1 = {"o1_unmodified_reference_runtime_secs": g["o1_runtime_secs"][0],
              "o1_note": "the unmodified Python reference (numpy + this repo's portable-trig twin, strict loops, 1 core) for "
                         "'-r 4 --r_dp 0.2', its own runtime(secs) line (OGG:1449), measured in the build container by "
                         "oracle/make_golden.py; the reference itself cannot travel to the GPU box"}
    except Exception:
        pass
    return dict({"value": pps, "unit": UNIT, "cores": cores, "kind": "port",
                 "sample": "one pass of '-r %s --r_dp %s' (%dx%d vertices, whole grid, %.1f s) through oracle/ (C restatement "
                           "of the reference, OpenMP)" % (r, r_dp, shape[0], shape[1], dt)}, **o1)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    r, wl = workload(args)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    use_all_host_threads()
    import ogg_oracle
    cores = ogg_oracle.num_threads()
    nyp, nxp = grid_shape_cpu(r, args.r_dp)
    for _ in range(args.warmup):
        cpu_pass(r, args.r_dp, world)
    t0 = time.time()
    pts = 0
    for _ in range(args.steps):
        _, _, shape = cpu_pass(r, args.r_dp, world)
        pts += shape[0] * shape[1]
    dt = (time.time() - t0) / args.steps
    pps = pts / (dt * args.steps)
    sample = ("whole '%s' grid per step" % wl) if world == 1 else (
        "per step: rank 0's row-slab share (1/%d of every sub-grid's rows, %dx%d vertices -- what one of the %d GPUs builds "
        "per step) of '%s'" % (world, shape[0], shape[1], world, wl))
    line = {"impl": "reference", "metric": METRIC, "value": pps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(wl, nyp, nxp, world, 0, 1),
            "cpu_baseline": {"value": pps, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": sample + "; oracle/ C restatement of the reference with OpenMP on all host threads (the "
                                       "reference itself is pure Python + the absent numpypi and cannot travel to this box)"},
            "e2e": {"value": pps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            # what this arm evaluated in its last timed step, as the same six checksums the GPU arm prints for rank 0's slab
            "rank0_slab_checksums_x_y_dx_dy_area_angle": host_checksums(own_rows(_LAST_CPU_PASS[0], 0, world)),
            "rank0_slab": {"slab": [0, world], "shape": list(own_rows(_LAST_CPU_PASS[0], 0, world)["x"].shape),
                           "evaluated_shape": list(shape)}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------
def _plan_for(flags):
    import contextlib
    import io
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G, supergrid as SG
    a = vars(G._build_parser().parse_args(flags.split()))
    plan_args = {k: v for k, v in a.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}
    with contextlib.redirect_stdout(io.StringIO()):
        return SG.plan_supergrid(**plan_args), plan_args


def _u64(t):
    return ["%016x" % (int(v) & 0xffffffffffffffff) for v in t.cpu().tolist()]


def run_fixed_grid(flags, V, steps, rank, world, dist, whole_check=False):
    """Strong scaling on a fixed grid cut into V row-slabs (V % world == 0): rank g builds slabs g, g+N, ... one after the
    other in one slab's buffers, then the single global-area all-reduce.  Returns ms per build of the whole grid (device
    time, max over ranks), the all-reduced area and the per-array device checksums summed over all slabs (additive, so
    they must not depend on N)."""
    import torch
    from ocean_model_grid_generator_b200 import _cabi, device as D, supergrid as SG
    plan, _ = _plan_for(flags)
    mine = list(range(rank, V, world))
    sg = SG.Supergrid(plan, mine[0], V, capacity_world=V)
    area_parts = torch.zeros(len(mine), dtype=torch.float64, device="cuda")
    acc = torch.zeros(6, dtype=torch.int64, device="cuda")

    def step(checks=False):
        for n, v in enumerate(mine):
            sg.set_slab(v, V).build()
            _cabi.call("ogg_sum_strided", D.ptr(sg.area), sg.area.numel(), 1, D.ptr_off(area_parts, n), D.stream())
            if checks:
                sg.checksums(acc)
        if dist is not None:
            dist.all_reduce(area_parts)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    step(checks=True)                       # warm-up pass; also the checksum pass (not timed)
    area_mine = area_parts.clone()          # after the all-reduce: every rank's per-position sums
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(acc)                # int64 sums wrap mod 2**64, which is the checksum's arithmetic
    ms = float(t[0]) / steps
    npoints = plan.nyp * plan.nxp
    out = {"workload": flags, "grid": [plan.nyp, plan.nxp], "slabs": V, "steps": steps, "ms_per_step": ms,
           "value": npoints / (ms * 1e-3), "unit": UNIT, "area_m2": float(area_mine.sum()),
           "checksums_x_y_dx_dy_area_angle": _u64(acc)}
    del sg
    if whole_check and rank == 0:
        whole = SG.Supergrid(plan).build()
        out["checksums_equal_whole_grid"] = _u64(whole.checksums()) == out["checksums_x_y_dx_dy_area_angle"]
        del whole
    torch.cuda.empty_cache()
    return out


def halo_midas_check(plan, rank, world, dist):
    """The general supplied-mesh entry point on a slab-distributed mesh: generate_grid_metrics_MIDAS (OGG:687-716) with the
    one-row halo received from the next rank over NCCL send/recv (slabs.metrics_midas_slab).  Mesh = this workload's bipolar
    cap (non-separable).  Every rank compares its dx, dy, area bit for bit with the rows of the whole-mesh result."""
    import contextlib
    import io
    import torch
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G, slabs
    p = [pc for pc in plan.pieces if pc.kind == "bipolar"][0].p
    with contextlib.redirect_stdout(io.StringIO()):
        xw, yw, _, _ = G.generate_bipolar_cap_mesh(plan.Ni, p["Nj"], p["lat0_bp"], p["lon_bp"], ensure_nj_even=False, device=True)
    nyp = xw.shape[0]
    v0, v1 = slabs.vertex_rows(nyp - 1, rank, world)
    c0, c1 = slabs.slab_bounds(nyp - 1, rank, world)
    xs, ys = xw[v0:v1].contiguous(), yw[v0:v1].contiguous()
    got = slabs.metrics_midas_slab(xs, ys)
    ref = G.generate_grid_metrics_MIDAS(xw, yw)
    ok = all(g.shape == r_.shape and bool((g.view(torch.int64) == r_.view(torch.int64)).all())
             for g, r_ in zip(got, (ref[0][v0:v1], ref[1][c0:c1], ref[2][c0:c1])))
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        slabs.metrics_midas_slab(xs, ys)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps, 0.0 if ok else 1.0], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"mesh": "bipolar cap %dx%d vertices in %d row-slabs" % (nyp, xw.shape[1], world), "ms": float(t[0]),
            "halo_bytes_per_boundary": 16 * int(xw.shape[1]), "transport": "NCCL send/recv (torch.distributed batch_isend_irecv)",
            "bit_equal_to_whole": float(t[1]) == 0.0}


def cli_timing(wl):
    """The reference's real entry point, end to end: main() -> plan -> build -> CHECK_metrics lines -> write_nc (OGG:855-1449),
    wall time from its own `runtime(secs)` line (OGG:1449); the file goes to a temporary directory."""
    import contextlib
    import io
    import shutil
    import tempfile
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    tmp = tempfile.mkdtemp(prefix="ogg_cli_")
    path = os.path.join(tmp, "ocean_hgrid.nc")
    runs, size = [], 0
    try:
        for _ in range(3):
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                G.cli(["-f", path] + wl.split())
            runs += [float(ln.split()[-1]) for ln in buf.getvalue().splitlines() if ln.startswith("runtime(secs)")]
            size = os.path.getsize(path)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    o1 = None
    try:
        o1 = json.load(open(os.path.join(ROOT, "tests", "golden", "configs.json")))["res0.25_full"]["o1_runtime_secs"][0]
    except Exception:
        pass
    return {"command": "ocean_grid_generator.py -f ocean_hgrid.nc " + wl, "runtime_secs": sorted(runs)[len(runs) // 2], "runs": runs,
            "file_bytes": size, "o1_unmodified_reference_runtime_secs_build_container": o1 if wl.strip() == "-r 4 --r_dp 0.2" else None}


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
    D.ready()
    lib = _cabi.load()
    r, wl = workload(args)
    import contextlib
    import io
    plan, plan_args = _plan_for(wl)
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
    # event brackets around the three heavy kernel classes, on every 4th step of the timed region only: an event
    # record between two kernels of a stream costs that stream a few microseconds, and six of them per step made the
    # measured step 2 % longer than the same step without them
    bracketed = [i for i in range(args.steps) if i % 4 == 0]
    l0 = lib.ogg_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    barrier()
    e0.record()
    for i in range(args.steps):
        lib.ogg_timing_enable(2 if i % 4 == 0 else 0)
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
        times, plans = [], []
        for it in range(6):
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()):
                p2 = SG.plan_supergrid(**plan_args)
                s2 = SG.Supergrid(p2, rank, world)
            t1 = time.perf_counter()
            if pinned is None:
                with D.host_near_device() as near:      # host pages on the GPU's own NUMA node (matters on 2-socket boxes)
                    pinned = s2.pinned_like()
                numa_bound = near.bound
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
            plans.append(t1 - t0)
        timed = sorted(times[1:])                         # the first iteration allocates the pinned buffers
        te = torch.tensor([timed[len(timed) // 2], sorted(plans[1:])[len(plans[1:]) // 2]], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        d2h = sum(v.numel() * 8 for v in pinned.values()) + 8
        h2d = sum(8 * (len(pc.p["lam"]) + len(pc.p["phi"])) for pc in plan.pieces if pc.kind == "band") + 8 * (plan.Ni + 1) + 4096
        e2e = {"value": npoints / float(te[0]), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms": float(te[0]) * 1e3, "statistic": "median of 5 timed iterations (max over ranks), after one untimed",
               "plan_ms": float(te[1]) * 1e3, "d2h_gbs_this_rank": d2h / max(1e-9, float(te[0]) - float(te[1])) / 1e9,
               "checksum_area_local": checksum, "host_area_last_cell": corner, "pinned_on_gpu_numa_node": numa_bound}
        del pinned

    # rank 0's first row-slab of the timed grid as the device holds it after one more build: six checksums, compared with
    # the CPU oracle's evaluation of the same slab when the line is assembled (outside every timed region)
    ours_chk = ours_shape = None
    if rank == 0:
        sg.set_slab(mine[0], V).build()
        ours_chk, ours_shape = _u64(sg.checksums()), [sg.vertex_rows, plan.nxp]
        torch.cuda.synchronize()

    # ---- N > 1: the halo path on the wire; every N: the fixed-size (strong-scaling) runs of configs[2..4] -------------
    halo = halo_midas_check(plan, rank, world, dist) if dist is not None else None
    strong = None
    if not args.no_strong and not args.slabs and args.resolution <= 0:
        del sg
        torch.cuda.empty_cache()
        strong = {}
        for tag, flags, Vs, nsteps in STRONG_CONFIGS:
            strong[tag] = run_fixed_grid(flags, Vs, nsteps, rank, world, dist, whole_check=(tag == "r8"))
    cli = None
    if rank == 0 and world == 1 and not args.no_cli and not args.slabs:
        cli = cli_timing(wl)

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
    ops = load_opcount()
    ncu = load_ncu_summary()
    kernels_ms = {k: round(v[0] / args.steps, 4) for k, v in sorted(kern_all.items(), key=lambda kv: -kv[1][0])}
    names = [n for n in ("dp_hsamples", "bipolar_quad_fused") if n in units]

    def rates(per_ms, which):
        tot = {w: sum(ops[n][w] * units[n] for n in which) for w in ("ops_8d", "ops_issue", "ops_reference")}
        return {w: tot[w] / (per_ms * 1e-3) / 1e12 for w in tot}, tot

    def roof_of(which, per_ms, where):
        rt, tot = rates(per_ms, which)
        hw = {n: ncu.get("kernels", {}).get(n) for n in which}
        traffic = None
        if all(hw.get(n) for n in which) and r == 4 and world == 1 and not args.slabs:
            traffic = sum(hw[n]["dram_bytes_per_launch"] for n in which)
        return {"kernel": " + ".join(which), "bound": "fp64", "unit": "TFLOP/s", "peak": fp64_peak_tf,
                "achieved": rt["ops_8d"], "frac": rt["ops_8d"] / fp64_peak_tf,
                "achieved_issue": rt["ops_issue"], "frac_issue": rt["ops_issue"] / fp64_peak_tf,
                "frac_reference_ops": rt["ops_reference"] / fp64_peak_tf,
                "pipe_fp64_busy_ncu": {n: (hw[n] or {}).get("pipe_fp64_pct") for n in which},
                "ncu_capture": ncu.get("capture"),
                "traffic": traffic,
                "ops_per_unit": {n: {w: round(ops[n][w], 2) for w in ("ops_8d", "ops_issue", "ops_reference")} for n in which},
                "units_per_step": {n: units[n] for n in which}, "ops_per_step": {w: tot[w] for w in tot},
                "kernel_ms_per_step": per_ms, "measured_in": where,
                "definitions": "frac: SURVEY 8(d) weights (+ - * / sqrt compare = 1 op) on the minimal per-unit count derived from "
                               "the oracle (profiles/opcount.json); frac_issue: the same count weighted by the FP64-pipe instructions "
                               "of the IEEE expansions (div 8, cached divisor 3, sqrt 8); frac_reference_ops: 8(d) weights on what the "
                               "reference evaluates per unique sample (un-hoisted, both trig polynomials) -- above 1 means the kernel "
                               "does less arithmetic than the reference for the same bits, not that it beats the pipe; "
                               "pipe_fp64_busy_ncu: sm__inst_executed_pipe_fp64 of the committed capture (not this run)",
                "peak_source": "measured in this run: DADD+DMUL issue rate with FMA contraction off (ogg_fp64_issue_rate); "
                               "MEASURED_PEAKS.json has no FP64 entry"}
    roof = roof_dp = roof_bp = None
    if names and span_n.value:
        per_step_ms = span_ms.value / len(bracketed)
        roof = roof_of(names, per_step_ms, "timed region: the two cap quadratures co-run on two streams; span during which at least "
                                           "one of them is in flight (CUDA events on their streams), averaged over the "
                                           "bracketed steps (every 4th of the timed steps)")
        roof["bracketed_steps"] = len(bracketed)
        roof["share_of_step"] = per_step_ms / (ms / args.steps)
    alone = "single-stream pass: the kernel has the GPU to itself"
    if "dp_hsamples" in units and "dp_hsamples" in kern_all:
        roof_dp = roof_of(["dp_hsamples"], kern_all["dp_hsamples"][0] / args.steps, alone)
    if "bipolar_quad_fused" in units and "bipolar_quad_fused" in kern_all:
        roof_bp = roof_of(["bipolar_quad_fused"], kern_all["bipolar_quad_fused"][0] / args.steps, alone)
    roof_hbm = None
    if "band_fill" in kern_all and "band_fill" in units:
        per_step_ms = kern_all["band_fill"][0] / args.steps
        ach = BYTES_BAND_VERTEX * units["band_fill"] / (per_step_ms * 1e-3) / 1e9
        roof_hbm = {"kernel": "band_fill", "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                    "frac": ach / hbm_peak, "peak_source": hbm_src, "kernel_ms_per_step": per_step_ms,
                    "traffic": ((ncu.get("kernels", {}).get("band_fill") or {}).get("dram_bytes_per_launch")
                                if r == 4 and world == 1 and not args.slabs else None),
                    "traffic_note": "DRAM bytes of the one launch that fills both bands, from the committed ncu capture; below the "
                                    "algorithmic bytes because part of the last rows is still in the 126 MB L2 when the kernel ends",
                    "algorithmic_bytes_per_step": BYTES_BAND_VERTEX * units["band_fill"],
                    "measured_in": "single-stream pass (in the timed region the band fills overlap the cap kernels)"}
    if roof is None:
        roof = roof_hbm

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if args.slabs else "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": config_of(wl, plan.nyp, plan.nxp, world, args.slabs, len(mine)),
            "cache": "each step rewrites %.0f MB of outputs + scratch, larger than the 126 MB L2; no flush needed" % (npoints * 48 / 1e6),
            "global_area_m2": total_area,
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "roofline_dp_hsamples": roof_dp, "roofline_bipolar_quad": roof_bp, "roofline_band_hbm": roof_hbm,
            "kernels_ms_per_step_single_stream_pass": kernels_ms,
            "kernels_ms_per_step_timed_region": {k: round(v[0] / len(bracketed), 4) for k, v in kern.items()}, "fp64_issue_peak_tflops_measured": fp64_peak_tf,
            "halo_midas": halo, "strong": strong, "cli": cli}
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(r, args.r_dp)
    elif world > 1:
        line["cpu_baseline"] = None
    # Parity of the grid that was timed, visible in the line itself: rank 0's first row-slab as the device holds it after
    # one more build, against the CPU oracle's evaluation of the same slab (outside the timed region; the oracle is the
    # checker here, never the thing measured).  Six splitmix64 checksums, x y dx dy area angle_dx.
    line["rank0_slab_checksums_x_y_dx_dy_area_angle"] = line["parity_rank0_slab_equals_cpu"] = None
    if rank == 0:
        line["rank0_slab_checksums_x_y_dx_dy_area_angle"] = ours_chk
        line["rank0_slab"] = {"slab": [mine[0], V], "shape": ours_shape}
        if not args.no_cpu_baseline:
            use_all_host_threads()
            theirs, shape = cpu_slab_checksums(r, args.r_dp, mine[0], V)
            line["parity_rank0_slab_equals_cpu"] = bool(ours_chk == theirs)
            line["rank0_slab"]["cpu_shape"] = shape
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

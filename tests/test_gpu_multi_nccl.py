"""Multi-GPU tests proper: one process per GPU over NCCL (skipped on a box with fewer than 2 GPUs).

Each rank builds ITS row-slab of every sub-grid through the C ABI, runs the general supplied-mesh MIDAS entry point with
the one-row halo received from the next rank over NCCL send/recv, and all-reduces the metrics_error sums.  Rank 0
also builds the whole grid alone; slabs gathered in rank order must equal it bit for bit (index work and FP64 alike),
the all-reduced area must agree within 1e-12 relative (north_star)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, flags, q):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import contextlib
    import io
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    ok, why = True, []
    try:
        from ocean_model_grid_generator_b200 import ocean_grid_generator as G, slabs, supergrid as SG
        args = vars(G._build_parser().parse_args(flags.split()))
        pa = {k: v for k, v in args.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta",
                                                          "write_subgrid_files")}
        with contextlib.redirect_stdout(io.StringIO()):
            plan = SG.plan_supergrid(**pa)
        mine = SG.build_supergrid(plan, rank, world)
        sums = slabs.global_sums(mine)                      # one NCCL all-reduce (4 doubles)
        torch.cuda.synchronize()
        # ---- gather every rank's slabs on rank 0 (test plumbing only) and compare with the whole grid -------------
        names = ("x", "y", "dx", "dy", "area", "angle_dx")
        arrs = mine.arrays()
        gathered = {}
        for k in names:
            t = arrs[k].contiguous()
            shapes = [None] * world
            dist.all_gather_object(shapes, tuple(t.shape))
            bufs = [torch.empty(s, dtype=torch.float64, device="cuda") for s in shapes]
            dist.all_gather(bufs, t) if len({tuple(s) for s in shapes}) == 1 else _gather_ragged(dist, bufs, t, rank, world)
            gathered[k] = [b.cpu().numpy() for b in bufs]
        slab_list = [None] * world
        dist.all_gather_object(slab_list, mine.slabs)
        if rank == 0:
            whole = SG.build_supergrid(plan).to_host()
            for k in names:
                rows = []
                for pi in range(len(plan.pieces)):
                    for r in range(world):
                        c0, c1 = slab_list[r][pi]
                        off = sum(b - a for a, b in slab_list[r][:pi])
                        rows.append(gathered[k][r][off:off + (c1 - c0)])
                if k in ("x", "y", "dx", "angle_dx"):
                    rows.append(gathered[k][world - 1][-1:])
                got = np.concatenate(rows, axis=0)
                if got.shape != whole[k].shape or not (got.view(np.int64) == whole[k].view(np.int64)).all():
                    ok = False
                    why.append("%s reassembled over NCCL ranks differs" % k)
            tot = float(whole["area"].sum())
            if abs(float(sums[0]) - tot) > 1e-12 * tot:
                ok = False
                why.append("all-reduced area %r vs %r" % (float(sums[0]), tot))
            col = float(whole["dy"][:, plan.Ni // 4].sum())
            if abs(float(sums[1]) - col) > 1e-12 * col:
                ok = False
                why.append("all-reduced dy column sum")
            if abs(float(sums[2]) - whole["dx"][0].sum()) > 1e-12 * whole["dx"][0].sum():
                ok = False
                why.append("dx first row sum")
            if abs(float(sums[3]) - whole["dx"][-1].sum()) > 1e-12 * whole["dx"][-1].sum():
                ok = False
                why.append("dx last row sum")
        # ---- supplied-mesh MIDAS with the NCCL halo row: mesh = the bipolar cap (non-separable) ------------------
        p = [pc for pc in plan.pieces if pc.kind == "bipolar"][0].p
        xw, yw, _, _ = G.generate_bipolar_cap_mesh(plan.Ni, p["Nj"], p["lat0_bp"], p["lon_bp"], device=True)
        nyp = xw.shape[0]
        v0, v1 = slabs.vertex_rows(nyp - 1, rank, world)
        c0, c1 = slabs.slab_bounds(nyp - 1, rank, world)
        dxs, dys, areas = slabs.metrics_midas_slab(xw[v0:v1].contiguous(), yw[v0:v1].contiguous())
        dxw, dyw, areaw = G.generate_grid_metrics_MIDAS(xw, yw)
        for got, ref, nm in ((dxs, dxw[v0:v1], "dx"), (dys, dyw[c0:c1], "dy"), (areas, areaw[c0:c1], "area")):
            if got.shape != ref.shape or not bool((got.view(torch.int64) == ref.view(torch.int64)).all()):
                ok = False
                why.append("rank %d: halo MIDAS %s differs from the whole-mesh result" % (rank, nm))
        q.put((rank, ok, why))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, False, [traceback.format_exc()]))
    finally:
        dist.destroy_process_group()


def _gather_ragged(dist, bufs, t, rank, world):
    for r in range(world):
        if r == rank:
            bufs[r].copy_(t)
        dist.broadcast(bufs[r], src=r)


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("flags", ["-r 1.0 --r_dp 0.2", "-r 2"])
def test_nccl_slabs_halo_allreduce(flags):
    import torch.multiprocessing as mp
    world = min(_ngpu(), 8)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, flags, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    bad = [(r, w) for r, ok, w in got if not ok]
    assert not bad, bad


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_one_process_drives_two_devices():
    """SURVEY 8(b) allows one process driving several devices: the library's per-device state (SM count, the opt-in to
    53 KB of dynamic shared memory of the fused bipolar kernel, the scratch pool, timing events) must follow
    cudaSetDevice.  The same calls on device 1 after device 0 give the same bits, and the timing brackets of each device
    are reported on that device."""
    import ctypes
    import torch
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from ocean_model_grid_generator_b200 import _cabi, device as D, ocean_grid_generator as G, supergrid as SG
    D.ready()
    lib = _cabi.load()
    lib.ogg_timing_reset()
    lib.ogg_timing_enable(1)
    results = []
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            rp = G.bipolar_rp(64.05895973)
            bp = G.bipolar_cap_metrics_quad_fast(5, 96, 16, 64.05895973, -300.0, rp)
            dp = G.displacedPoleCap_metrics_quad(4, 48, 10, -300.0, -78.0, 80.0, 0.2)
            a = vars(G._build_parser().parse_args("-r 0.5 --r_dp 0.2".split()))
            pa = {k: v for k, v in a.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}
            sg = SG.build_supergrid(SG.plan_supergrid(**pa)).to_host()
            torch.cuda.synchronize()
            tot, n = ctypes.c_double(), ctypes.c_longlong()
            kid = {lib.ogg_timing_name(k).decode(): k for k in range(lib.ogg_timing_kernels())}
            assert lib.ogg_timing_get(kid["bipolar_quad_fused"], ctypes.byref(tot), ctypes.byref(n)) == 0
            assert n.value >= 2 and tot.value > 0        # this device's brackets only; elapsed times are valid
            results.append([np.asarray(v) for v in bp + dp] + [sg[k] for k in ("x", "y", "dx", "dy", "area", "angle_dx")])
    lib.ogg_timing_enable(0)
    lib.ogg_timing_reset()
    for other in results[1:]:
        for u, v in zip(results[0], other):
            assert u.shape == v.shape and (u.view(np.int64) == v.view(np.int64)).all()

"""world_size-2/3 CPU tests (gloo) of the multi-GPU host logic: slab ownership, the one-row halo exchange and the
single all-reduce (ocean_model_grid_generator_b200/slabs.py).  Arithmetic on the CPU ranks is the ORACLE's
generate_grid_metrics_MIDAS (the checker), so the test proves the exchange protocol, not the kernels."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _mesh(nyp, nxp):
    """A non-separable mesh (skewed longitudes, curved latitudes) so that every halo value matters."""
    j, i = np.meshgrid(np.arange(nyp, dtype=np.float64), np.arange(nxp, dtype=np.float64), indexing="ij")
    x = -300.0 + 360.0 * i / (nxp - 1) + 0.37 * np.sin(0.11 * j)
    y = -70.0 + 120.0 * j / (nyp - 1) + 0.23 * np.cos(0.05 * i + 0.02 * j)
    return x, y


def _worker(rank, world, port, nyp, nxp, q):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import ogg_oracle as O
        from ocean_model_grid_generator_b200 import slabs
        from util import bits_equal
        x, y = _mesh(nyp, nxp)
        dx_w, dy_w, area_w = O.generate_grid_metrics_MIDAS(x, y)
        v0, v1 = slabs.vertex_rows(nyp - 1, rank, world)
        c0, c1 = slabs.slab_bounds(nyp - 1, rank, world)

        def midas(xs, ys, halo=None):
            xs, ys = xs.numpy(), ys.numpy()
            if halo is not None:      # the oracle has no halo argument: append the received row, drop the extra dx row
                xs = np.vstack((xs, halo[0].numpy()[None]))
                ys = np.vstack((ys, halo[1].numpy()[None]))
                dx, dy, area = O.generate_grid_metrics_MIDAS(xs, ys)
                return dx[:-1], dy, area
            return O.generate_grid_metrics_MIDAS(xs, ys)

        xs, ys = torch.from_numpy(x[v0:v1].copy()), torch.from_numpy(y[v0:v1].copy())
        halo = slabs.exchange_halo(xs, ys)
        ok = True
        if rank < world - 1:
            ok &= bits_equal(halo[0].numpy(), x[v1]) and bits_equal(halo[1].numpy(), y[v1])
        else:
            ok &= halo is None
        dx, dy, area = slabs.metrics_midas_slab(xs, ys, midas=midas)
        ok &= bits_equal(dx, dx_w[v0:v1]) and bits_equal(dy, dy_w[c0:c1]) and bits_equal(area, area_w[c0:c1])
        sums = slabs.allreduce_sums(torch.tensor([float(area.sum()), float(rank + 1)], dtype=torch.float64))
        ok &= abs(float(sums[0]) - float(area_w.sum())) <= 1e-12 * float(area_w.sum())
        ok &= float(sums[1]) == world * (world + 1) / 2
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,nyp,nxp", [(2, 17, 33), (3, 11, 20), (2, 3, 5)])
def test_halo_exchange_and_allreduce_gloo(world, nyp, nxp):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nyp, nxp, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    got = dict(q.get(timeout=5) for _ in range(world))
    assert got == {r: True for r in range(world)}


def _worker_validate(rank, world, port, q):
    """exchange_halo's cross-rank check: done on the first call for a (group, nxp) only, on every call when asked, and
    a mesh without a row to send makes every rank raise together when it is done."""
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ocean_model_grid_generator_b200 import slabs
        from util import bits_equal
        calls = []
        real = dist.all_reduce

        def counting(*a, **k):
            calls.append(1)
            return real(*a, **k)
        dist.all_reduce = counting
        ok = True
        x, y = _mesh(9, 12)
        v0, v1 = slabs.vertex_rows(8, rank, world)
        xs, ys = torch.from_numpy(x[v0:v1].copy()), torch.from_numpy(y[v0:v1].copy())
        for rep in range(3):
            halo = slabs.exchange_halo(xs, ys)
            if rank < world - 1:
                ok &= bits_equal(halo[0].numpy(), x[v1]) and bits_equal(halo[1].numpy(), y[v1])
        ok &= len(calls) == 1                      # "once": only the first of the three calls validated
        slabs.exchange_halo(xs, ys, validate="always")
        slabs.exchange_halo(xs, ys, validate="never")
        ok &= len(calls) == 2
        # a new number of columns is a new combination: checked again, and the last rank's empty slab is caught by all
        x2, y2 = _mesh(9, 7)
        xe = torch.from_numpy(x2[v0:v1].copy()) if rank < world - 1 else torch.empty((0, 7), dtype=torch.float64)
        ye = torch.from_numpy(y2[v0:v1].copy()) if rank < world - 1 else torch.empty((0, 7), dtype=torch.float64)
        try:
            slabs.exchange_halo(xe, ye)
            ok = False
        except Exception as e:
            ok &= "no mesh row" in str(e)
        ok &= len(calls) == 3
        # what a rank can see by itself is checked on every call, validated combination or not
        try:
            slabs.exchange_halo(xs, ys[:, :-1], validate="never")
            ok = False
        except Exception as e:
            ok &= "same shape" in str(e)
        dist.all_reduce = real
        dist.barrier()
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_halo_validation_modes_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_validate, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    got = dict(q.get(timeout=5) for _ in range(world))
    assert got == {r: True for r in range(world)}


def test_slab_bounds_cover_and_balance():
    from ocean_model_grid_generator_b200 import slabs
    for n in (0, 1, 7, 8, 143, 1400):
        for world in (1, 2, 3, 8):
            b = [slabs.slab_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[k][1] == b[k + 1][0] for k in range(world - 1))
            sizes = [c1 - c0 for c0, c1 in b]
            assert max(sizes) - min(sizes) <= 1
            v = [slabs.vertex_rows(n, r, world) for r in range(world)]
            assert v[-1][1] == n + 1

"""Latitude row-slabs across GPUs: ownership, the one-row halo exchange and the single all-reduce.

One process per GPU (torch.distributed; NCCL over NVLink on GPUs, gloo in the CPU tests).  The reference is a
single process (no communication backend); its only cross-row couplings are
  * generate_grid_metrics_MIDAS: dy and area of cell row j read mesh row j+1 (OGG:699-702, 713);
  * metrics_error: sums over whole arrays / whole columns (OGG:747-756).
Generated sub-grids need no exchange (the row above a slab is recomputed from the 1-D laws, supergrid.py); a
*supplied* slab-distributed mesh needs the next rank's first row -- `exchange_halo` -- and the sums need one
all-reduce -- `allreduce_sums`.  This module moves bytes only; all arithmetic is in the kernels.
"""
import torch
import torch.distributed as dist


def slab_bounds(ncells, rank, world):
    """Contiguous split of `ncells` cell rows over `world` ranks: [c0, c1) of rank `rank` (sizes differ by <= 1)."""
    base, rem = divmod(int(ncells), int(world))
    c0 = rank * base + min(rank, rem)
    return c0, c0 + base + (1 if rank < rem else 0)


def vertex_rows(ncells, rank, world):
    """Mesh (vertex) rows a rank owns: its cell rows' lower edges, plus the closing row on the last rank."""
    c0, c1 = slab_bounds(ncells, rank, world)
    return c0, c1 + (1 if rank == world - 1 else 0)


def _world(group):
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


# (group, number of columns, device type) combinations whose cross-rank agreement has been checked already
_validated = set()


def _validate_collectively(x, y, rank, group):
    """One tiny all-reduce (min of "my arguments are fine", and the row length, which must agree across ranks), read
    back with ONE device-to-host copy.  Every rank raises together, before any point-to-point operation is posted."""
    good = x.dim() == 2 and tuple(x.shape) == tuple(y.shape) and (rank == 0 or x.shape[0] >= 1)
    nxp = int(x.shape[1]) if x.dim() == 2 else -1
    check = torch.tensor([1 if good else 0, nxp, -nxp], dtype=torch.int64, device=x.device)
    dist.all_reduce(check, op=dist.ReduceOp.MIN, group=group)
    all_good, lo, neg_hi = check.tolist()
    if all_good == 0:
        if x.dim() != 2 or tuple(x.shape) != tuple(y.shape):
            raise Exception("Input arrays do not have the same shape!")
        raise Exception("exchange_halo: a rank owns no mesh row to send or passed mismatched arrays (this rank: %d)" % rank)
    if lo != -neg_hi:
        raise Exception("exchange_halo: the ranks' meshes differ in the number of columns")


def exchange_halo(x, y, group=None, validate=None):
    """Send this rank's first mesh row of (x, y) to the rank below, receive the first row of the rank above.

    x, y: [rows_local][nxp] float64 tensors (CUDA with NCCL, CPU with gloo).  Returns (halo_x, halo_y), each [nxp],
    or None on the last rank (its closing row is its own).  2*nxp doubles per rank boundary, sent straight from the
    rows (no staging copy), received into one [2][nxp] buffer.

    validate: "once" (default; OGG_HALO_VALIDATE overrides) checks across ranks that every rank has a row to send and
    that all meshes have the same number of columns the FIRST time a (group, nxp) combination is seen -- one all-reduce
    and one host read -- and never again, so that the steady-state exchange posts its two messages without the host
    waiting on the device.  What each rank can see by itself (x and y two-dimensional, same shape, a row to send) is
    still checked on every call; a rank that fails that check alone raises alone, and its neighbours find out through
    the communicator's timeout.  "always" repeats the collective check on every call (all ranks raise together, as the
    reference-facing tests expect); "never" skips it.  All ranks must pass the same mode.
    """
    rank, world = _world(group)
    if world == 1:
        return None
    if validate is None:
        import os
        validate = os.environ.get("OGG_HALO_VALIDATE", "once")
    if validate not in ("once", "always", "never"):
        raise ValueError("validate must be 'once', 'always' or 'never'")
    key = (id(group) if group is not None else None, int(x.shape[1]) if x.dim() == 2 else -1, x.device.type)
    if validate == "always" or (validate == "once" and key not in _validated):
        _validate_collectively(x, y, rank, group)
        _validated.add(key)
    else:
        if x.dim() != 2 or tuple(x.shape) != tuple(y.shape):
            raise Exception("Input arrays do not have the same shape!")
        if rank > 0 and x.shape[0] < 1:
            raise Exception("exchange_halo: this rank (%d) owns no mesh row to send" % rank)
    nxp = int(x.shape[1])
    ops, recv = [], None
    if rank > 0:
        below = dist.get_global_rank(group, rank - 1) if group else rank - 1
        for row in (x[0], y[0]):                      # a row of a contiguous mesh is contiguous: sent in place
            ops.append(dist.P2POp(dist.isend, row if row.is_contiguous() else row.contiguous(), below, group))
    if rank < world - 1:
        above = dist.get_global_rank(group, rank + 1) if group else rank + 1
        recv = torch.empty((2, nxp), dtype=x.dtype, device=x.device)
        for k in (0, 1):
            ops.append(dist.P2POp(dist.irecv, recv[k], above, group))
    # NCCL: wait() orders the caller's stream after the transfer, it does not block the host
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    if recv is None:
        return None
    return recv[0], recv[1]


def metrics_midas_slab(x, y, midas=None, group=None, **kwargs):
    """generate_grid_metrics_MIDAS (OGG:687-716) on a slab-distributed supplied mesh.

    Every rank passes the mesh rows it owns (see `vertex_rows`); returns this rank's dx [rows][nx], dy [cells][nxp],
    area [cells][nx].  `midas` defaults to the CUDA path; the CPU tests pass the oracle to check the protocol.
    """
    if midas is None:
        from . import ocean_grid_generator as ogg
        midas = ogg.generate_grid_metrics_MIDAS
    halo = exchange_halo(x, y, group)
    return midas(x, y, halo=halo, **kwargs)


def allreduce_sums(values, group=None):
    """The single all-reduce of the path: element-wise sum over ranks of a small float64 tensor (in place)."""
    rank, world = _world(group)
    if world > 1:
        dist.all_reduce(values, op=dist.ReduceOp.SUM, group=group)
    return values


def global_sums(sg, group=None):
    """Sums behind metrics_error / the global-area check for a slab-distributed supergrid (OGG:747-756):
    [sum(area), sum(dy[:, Ni//4]), sum(dx[0, :]) (first rank), sum(dx[-1, :]) (last rank)] -> all ranks."""
    from . import _cabi
    from . import device as _dev
    out = torch.zeros(4, dtype=torch.float64, device=sg.area.device)
    st = _dev.stream()
    nxp = sg.plan.nxp
    if sg.area.numel():
        _cabi.call("ogg_sum_strided", _dev.ptr(sg.area), sg.area.numel(), 1, _dev.ptr_off(out, 0), st)
        _cabi.call("ogg_sum_strided", _dev.ptr_off(sg.dy, sg.plan.Ni // 4), sg.dy.shape[0], nxp, _dev.ptr_off(out, 1), st)
    if sg.rank == 0 and sg.rows_local:
        _cabi.call("ogg_sum_strided", _dev.ptr(sg.dx), nxp - 1, 1, _dev.ptr_off(out, 2), st)
    if sg.rank == sg.world - 1:
        _cabi.call("ogg_sum_strided", _dev.ptr_off(sg.dx, sg.rows_local * (nxp - 1)), nxp - 1, 1, _dev.ptr_off(out, 3), st)
    return allreduce_sums(out, group)

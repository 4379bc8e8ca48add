"""ogg_dp_cap_mesh_noalloc: the displaced-pole cap mesh for callers that are not allowed to allocate device memory.

Some users of this library cannot let it call the CUDA allocator: an ocean-model component whose device-memory budget
is fixed at start-up, a thread inside a region where the allocator's driver-wide lock (and the device synchronisation a
first stream-ordered allocation can cause) is not acceptable, or a process whose device memory is managed by another
runtime.  For them the library carries one entry point whose work tables sit in a fixed 2 MiB array inside the library's
device image instead of coming from its memory pool.  What is checked here, through the C ABI only:

  * its lams / phis equal those of the allocating entry point ogg_dp_cap_mesh bit for bit (the tolerance every other
    parity test of this repo uses for the mesh arrays), on two calls in a row with DIFFERENT geometries, so that the
    second call would show anything the first one left in the shared array;
  * a request beyond the array's size is refused with OGG_ERR_TOO_LARGE and writes nothing;
  * the library's own memory pool does not grow across the calls.

This file sorts last on purpose (`pytest -x` stops at the first failure: a failure here must not hide other tests), and it
must pass both alone in a fresh process -- where its first call is the first use of the library on the device -- and after
every other test.
"""

import pytest

pytestmark = pytest.mark.gpu


def _dev():
    import torch
    from ocean_model_grid_generator_b200 import _cabi, device as dev
    dev.ready()
    return torch, _cabi, dev


def _mesh(entry, Ni, Nj, j0, j1, lon0, lat0, lon_dp, r_dp, fill):
    """One call of `entry` into freshly poisoned output arrays (so unwritten elements cannot pass as results)."""
    torch, _cabi, dev = _dev()
    lams = dev.full((j1 - j0, Ni + 1), fill)
    phis = dev.full((j1 - j0, Ni + 1), fill)
    _cabi.call(entry, Ni, Nj, j0, j1, lon0, lat0, lon_dp, r_dp, dev.ptr(lams), dev.ptr(phis), dev.stream())
    torch.cuda.synchronize()
    return dev.to_host(lams), dev.to_host(phis)


# two caps that differ in every size and parameter: a second call that reused any table of the first would be wrong
CAPS = [
    dict(Ni=1440, Nj=140, j0=0, j1=141, lon0=-300.0, lat0=-70.0, lon_dp=60.0, r_dp=0.2),     # whole cap, README-like
    dict(Ni=360, Nj=47, j0=5, j1=33, lon0=-280.0, lat0=-66.5, lon_dp=110.0, r_dp=0.35),      # a row range of another
]


def test_noalloc_mesh_equals_allocating_mesh_twice_in_a_row():
    from util import assert_bits
    for rep, cap in enumerate(CAPS + CAPS[:1]):          # third call: back to the first geometry after the second
        # the no-allocation call goes first: run alone, this is the library's very first call on the device
        got_l, got_p = _mesh("ogg_dp_cap_mesh_noalloc", fill=-9999.0, **cap)
        want_l, want_p = _mesh("ogg_dp_cap_mesh", fill=-7777.0, **cap)
        assert_bits(got_l, want_l, "call %d lams" % rep)
        assert_bits(got_p, want_p, "call %d phis" % rep)
        assert not (got_l == -9999.0).any() and not (got_p == -9999.0).any()
    # the same call twice with nothing in between gives the same bits both times
    a = _mesh("ogg_dp_cap_mesh_noalloc", fill=1.0, **CAPS[1])
    b = _mesh("ogg_dp_cap_mesh_noalloc", fill=2.0, **CAPS[1])
    assert_bits(a[0], b[0], "repeat lams")
    assert_bits(a[1], b[1], "repeat phis")


def test_noalloc_mesh_back_to_back_on_one_stream_without_host_sync():
    """Two calls queued on one stream with no synchronisation between them: stream order alone must keep the second
    call's tables from overwriting the first call's before its kernels have read them."""
    from util import assert_bits
    torch, _cabi, dev = _dev()
    outs = []
    for cap in CAPS:
        shape = (cap["j1"] - cap["j0"], cap["Ni"] + 1)
        outs.append((dev.full(shape, -9999.0), dev.full(shape, -9999.0)))
    for cap, (lams, phis) in zip(CAPS, outs):
        _cabi.call("ogg_dp_cap_mesh_noalloc", cap["Ni"], cap["Nj"], cap["j0"], cap["j1"], cap["lon0"], cap["lat0"],
                   cap["lon_dp"], cap["r_dp"], dev.ptr(lams), dev.ptr(phis), dev.stream())
    torch.cuda.synchronize()
    for k, (cap, (lams, phis)) in enumerate(zip(CAPS, outs)):
        want_l, want_p = _mesh("ogg_dp_cap_mesh", fill=-7777.0, **cap)
        assert_bits(dev.to_host(lams), want_l, "queued call %d lams" % k)
        assert_bits(dev.to_host(phis), want_p, "queued call %d phis" % k)


def test_noalloc_mesh_refuses_requests_beyond_its_fixed_array():
    torch, _cabi, dev = _dev()
    lib = _cabi.load()
    limit = _cabi.OGG_NOALLOC_SCRATCH_DOUBLES
    out = dev.full((4,), -9999.0)            # never written: the refusal happens before any launch
    launches = lib.ogg_launch_count()
    # too many columns for the array, then columns that fit but too many rows on top of them
    for Ni, Nj, j0, j1 in ((limit // 2, 10, 0, 1), (1000, limit, 0, limit - 2000)):
        rc = lib.ogg_dp_cap_mesh_noalloc(Ni, Nj, j0, j1, -300.0, -70.0, 60.0, 0.2, dev.ptr(out), dev.ptr(out),
                                         dev.stream())
        assert rc == _cabi.OGG_ERR_TOO_LARGE
        with pytest.raises(_cabi.OggError, match="fixed .* work array"):
            _cabi.check(rc, "ogg_dp_cap_mesh_noalloc")
    torch.cuda.synchronize()
    assert lib.ogg_launch_count() == launches
    assert (dev.to_host(out) == -9999.0).all()
    # the largest request that does fit is accepted: 32 + 2 * (Ni + 1) + rows == limit
    Ni = 131000                              # 110 rows of 131001 columns: 115 MB per array
    rows = limit - 32 - 2 * (Ni + 1)
    lams, phis = _mesh("ogg_dp_cap_mesh_noalloc", Ni, rows + 10, 0, rows, -300.0, -70.0, 60.0, 0.2, fill=-9999.0)
    want_l, want_p = _mesh("ogg_dp_cap_mesh", Ni, rows + 10, 0, rows, -300.0, -70.0, 60.0, 0.2, fill=-7777.0)
    from util import assert_bits
    assert_bits(lams, want_l, "largest request lams")
    assert_bits(phis, want_p, "largest request phis")


def test_noalloc_mesh_leaves_the_device_allocator_alone():
    """Free device memory as the driver reports it does not move across no-allocation calls (the outputs are allocated
    before the first reading).  cudaMemGetInfo is coarse -- the driver hands memory out in 2 MiB steps -- but the tables of
    this cap are far smaller than that, so any pool growth of the allocating kind would come with the first call."""
    torch, _cabi, dev = _dev()
    cap = CAPS[0]
    shape = (cap["j1"] - cap["j0"], cap["Ni"] + 1)
    lams, phis = dev.full(shape, 0.0), dev.full(shape, 0.0)
    args = (cap["Ni"], cap["Nj"], cap["j0"], cap["j1"], cap["lon0"], cap["lat0"], cap["lon_dp"], cap["r_dp"],
            dev.ptr(lams), dev.ptr(phis), dev.stream())
    _cabi.call("ogg_dp_cap_mesh_noalloc", *args)      # set-up call: address look-up, module load if this is the first use
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        _cabi.call("ogg_dp_cap_mesh_noalloc", *args)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0

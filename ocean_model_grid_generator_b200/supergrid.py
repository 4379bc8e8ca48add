"""Plan + in-place device build of the stitched supergrid (the body of the reference's main(), OGG:855-1449).

Host side (this file): the reference's sizing rules, cuts and joins reduced to a *plan* -- an ordered list of
sub-grid pieces (south to north) with their generator parameters, the rows kept, and the row offset each piece
is written at.  Device side: every piece is generated straight into the six final arrays at its offset
(no sub-grid temporaries, no concatenate copies -- OGG:1316-1365 becomes pointer arithmetic), optionally only a
latitude row-slab of each piece (one slab per GPU).

Joins follow OGG:1316-1365: x, y, dx, angle_dx = lower[:-1] + upper;  dy, area = lower + upper.
"""
import ctypes
import sys
import time

import numpy

from . import _cabi
from . import device as _dev
from . import devnp as np
from . import ocean_grid_generator as ogg

PI_180 = ogg.PI_180
Re = ogg._default_Re
FIELDS = ("x", "y", "dx", "dy", "area", "angle_dx")


class Piece(object):
    """One sub-grid as it enters the stitch.  Vertex rows kept: [row0, rows_full)."""

    def __init__(self, name, kind, rows_full, **params):
        self.name, self.kind, self.rows_full, self.row0 = name, kind, int(rows_full), 0
        self.p = params

    @property
    def rows(self):
        return self.rows_full - self.row0

    def cut(self, n):
        self.row0 += int(n)

    def copy(self):
        q = Piece(self.name, self.kind, self.rows_full, **self.p)
        q.row0 = self.row0
        return q


class Plan(object):
    def __init__(self, Ni, pieces, skip_metrics, info):
        self.Ni, self.pieces, self.skip_metrics, self.info = int(Ni), pieces, bool(skip_metrics), info

    @property
    def nxp(self):
        return self.Ni + 1

    @property
    def nyp(self):
        return sum(p.rows - 1 for p in self.pieces) + 1

    def offsets(self):
        off, out = 0, []
        for p in self.pieces:
            out.append(off)
            off += p.rows - 1
        return out


def _has(grids, name):
    return name in grids or "all" in grids


def plan_supergrid(inverse_resolution, r_dp=0.0, lon_dp=80.0, lat_dp=-99.0, exfracdp=None, south_cutoff_row=0,
                   south_cutoff_ang=-90.0, enhanced_equatorial=0, grids="all", match_dy=(), skip_metrics=False,
                   ensure_nj_even=False, shift_equator_to_u_point=True, bipolar_lower_lat=-99.0,
                   mercator_lower_lat=-99.0, mercator_upper_lat=-99.0, south_ocean_lower_lat=-99.0,
                   no_south_cap=False, **_unused):
    """Sizing rules of OGG:969-1195 and the south cut of OGG:1270-1313, as data."""
    doughnut = 0.28 * 7 / 4 if exfracdp is None else exfracdp                      # OGG:891-892
    refineS, refineR, lenlon, lon0 = 2, inverse_resolution, 360, -300.0            # OGG:969-972
    Ni = int(refineR * refineS * lenlon)                                           # OGG:974
    phi_s_Merc, phi_n_Merc = -66.85954725, 64.05895973                             # OGG:987
    if mercator_upper_lat > -90:
        phi_n_Merc = mercator_upper_lat
    if mercator_lower_lat > -90:
        phi_s_Merc = mercator_lower_lat
    if refineR == 2:
        phi_s_Merc, phi_n_Merc = -68.0, 65.0                                       # OGG:996-998
    if refineR == 1 and enhanced_equatorial:
        phi_s_Merc, phi_n_Merc = -77.8, 60.0                                       # OGG:999-1001
    info = {"lon0": lon0, "doughnut": doughnut}
    # every sub-grid as main() first generates it (before the doughnut / parity / south cuts): what its CHECK_metrics line
    # is evaluated on (OGG:1016, 1069, 1111, 1156, 1172) and, for Merc / BP / SO, what --write_subgrid_files writes
    as_generated = info["as_generated"] = {}
    lam1d = lon0 + numpy.arange(Ni + 1) * lenlon / float(Ni)                       # OGG:431, :834 (same law)

    merc = bp = so = sc = None
    if _has(grids, "mercator"):
        phiM = ogg.mercator_latitudes(Ni, phi_s_Merc, phi_n_Merc, refineR, shift_equator_to_u_point, ensure_nj_even,
                                      enhanced_equatorial)
        if phiM.shape[0] % 2 == 0 and ensure_nj_even:                               # OGG:434-437
            phiM = phiM[1:]
        merc = Piece("Merc", "band", phiM.shape[0], lam=lam1d, phi=phiM)
        as_generated["Merc"] = merc.copy()
        DeltaPhiMerc_so = phiM[1] - phiM[0]                                         # OGG:1024-1025
        DeltaPhiMerc_no = phiM[-1] - phiM[-2]
        lat0_bp = phiM[-1]                                                          # OGG:1030
        info["phiMerc"] = (phiM[0], phiM[-1])
    if _has(grids, "bipolar"):
        if bipolar_lower_lat > -90:
            lat0_bp = bipolar_lower_lat
        Nj_ncap = int(60 * refineR * refineS)                                       # OGG:1042
        if refineR == 2:
            Nj_ncap = 119 * refineS
        if refineR == 1 and enhanced_equatorial:
            Nj_ncap = 154
        if "bp" in match_dy:
            Nj_ncap = int(0.5 + (90.0 - lat0_bp) / DeltaPhiMerc_no)                 # OGG:1057
        if Nj_ncap % 2 != 0 and ensure_nj_even:                                     # OGG:107-111
            Nj_ncap -= 1
        bp = Piece("BP", "bipolar", Nj_ncap + 1, Nj=Nj_ncap, lat0_bp=float(lat0_bp), lon_bp=lon0,
                   rp=ogg.bipolar_rp(lat0_bp))
        as_generated["BP"] = bp.copy()
    # Southern Ocean band (OGG:1080-1116)
    lat0_SO = -78.0
    if south_ocean_lower_lat > -90:
        lat0_SO = south_ocean_lower_lat
    if merc is None:
        raise Exception("the Mercator sub-grid is required: the reference reads phiMerc at OGG:1083")
    latUp_SO = merc.p["phi"][0]
    lenlat_SO = latUp_SO - lat0_SO
    deltaPhiSO = 1.0 / refineR / refineS
    Nj_SO = int(refineR * 55)
    if refineR == 2 and enhanced_equatorial:
        Nj_SO = 109
    if refineR == 1 and enhanced_equatorial:
        Nj_SO = 0
    if "so" in match_dy:
        Nj_SO = int(0.5 + lenlat_SO / DeltaPhiMerc_so)                              # OGG:1097
    info.update(latUp_SO=latUp_SO, lat0_SO=lat0_SO)
    if Nj_SO != 0 and _has(grids, "so"):
        lamSO, phiSO = ogg.latlon_vectors(Ni, Nj_SO, lon0, lenlon, lat0_SO, lenlat_SO, ensure_nj_even)
        so = Piece("SO", "band", phiSO.shape[0], lam=lamSO, phi=phiSO)
        as_generated["SO"] = so.copy()
    if so is None:
        raise Exception("the Southern Ocean sub-grid is required: the reference reads phiSO at OGG:1122")
    # South cap (OGG:1122-1195)
    lat0_SC = so.p["phi"][0]
    if "p125sc" in match_dy:
        lat0_SC = lat0_SO
    Nj_scap = int(refineR * 40) * 7 // 4                                            # OGG:1127-1131
    if no_south_cap or (refineR == 2 and enhanced_equatorial) or (refineR == 1 and enhanced_equatorial):
        Nj_scap = 0
    info["r_dp"] = r_dp
    if Nj_scap != 0 and _has(grids, "sc"):
        if r_dp == 0.0 and lat_dp <= -90.0:                                         # regular pole, OGG:1145-1157
            fullArc = lat0_SC + 90.0
            Nj_scap = int(fullArc / deltaPhiSO)
            lamSC, phiSC = ogg.latlon_vectors(Ni, Nj_scap, lon0, lenlon, -90.0, 90 + lat0_SO, ensure_nj_even)
            sc = Piece("SC", "band", phiSC.shape[0], lam=lamSC, phi=phiSC, regular=True)
            as_generated["SC"] = sc.copy()
        else:                                                                       # displaced pole, OGG:1158-1195
            if lat_dp > -90:
                r_dp = float(np.tan((90 + lat_dp) * PI_180) / np.tan((90 + lat0_SC) * PI_180))   # OGG:1160
                info["r_dp"] = r_dp
            sc = Piece("SC", "dp", Nj_scap + 1, Ni=Ni, Nj=Nj_scap, lon0=lon0, lat0=float(lat0_SC), lon_dp=float(lon_dp),
                       r_dp=float(r_dp))
            as_generated["SC"] = sc.copy()
            if doughnut != 0.0:
                jmin = numpy.ceil(doughnut * Nj_scap)                               # OGG:1178-1180
                sc.cut(int(jmin + numpy.mod(jmin, 2)))
            if sc.rows % 2 == 0 and ensure_nj_even:                                 # OGG:1188-1195
                sc.cut(1)
    info["lat0_SC"] = lat0_SC
    if sc is not None:
        info["SC_file_piece"] = sc.copy()       # what --write_subgrid_files writes for SC (OGG:1227-1229): before the south cut

    # South cut (OGG:1270-1313)
    cut, jcut = False, 0
    if south_cutoff_row > 0:
        cut, jcut = True, south_cutoff_row - 1
    elif south_cutoff_ang > -90:
        cut = True
        jcut = 1 + numpy.nonzero(first_column_latitudes(sc) < south_cutoff_ang)[0][-1]
    if cut:
        if sc is not None and jcut < sc.rows:
            if (sc.rows - jcut) % 2 == 0 and ensure_nj_even:
                jcut += 1
            print("   Cutting SC grid rows 0 to ", jcut)
            sc.cut(jcut)
        elif so is not None:
            print("   Whole SC and some of SO need to be cut!")
            jcut_SO = jcut - (sc.rows if sc is not None else 0)
            sc = None
            if ((so.rows - 1) - jcut_SO - 1) % 2 == 0 and ensure_nj_even:
                jcut_SO += 1
            print("   No SC grid remained. Cutting SO grid rows 0 to ", jcut_SO)
            so.cut(jcut_SO)
    pieces = [p for p in (sc, so, merc, bp) if p is not None]
    info.update(hasSC=sc is not None, hasSO=so is not None, hasMerc=merc is not None, hasBP=bp is not None)
    return Plan(Ni, pieces, skip_metrics, info)


def first_column_latitudes(piece):
    """phi[:, 0] of a piece's kept rows (needed by --south_cutoff_ang, OGG:1278)."""
    if piece is None:
        raise Exception("south_cutoff_ang needs the southern cap sub-grid (OGG:1278)")
    if piece.kind == "band":
        return piece.p["phi"][piece.row0:]
    p = piece.p
    _, phis, _, _ = ogg.displacedPoleCap_mesh(numpy.arange(1), numpy.arange(piece.row0, piece.rows_full), p["Ni"],
                                              p["Nj"], p["lon0"], p["lat0"], p["lon_dp"], p["r_dp"])
    return phis[:, 0]


# --------------------------------------------------------------------------------------------------
# device build
# --------------------------------------------------------------------------------------------------
from .slabs import slab_bounds  # noqa: E402  (contiguous split of a piece's cell rows; every rank gets a share of every piece)


_POOL = {}


def _stream_pool():
    """Side streams shared by every Supergrid of this process and device (creating a stream costs ~0.1 ms): two for the
    caps' quadratures, four of higher priority for everything short; eight reusable events."""
    import torch
    dev = torch.cuda.current_device()
    if dev not in _POOL:
        # measured at r=4 (tools/stepbench.py): all equal 0.977 ms, short kernels above the quadratures 0.972, and the
        # second cap's quadrature (the cheaper bipolar one, which then finishes first and frees the SMs for the
        # displaced-pole tail) above the first 0.966
        import os
        p0, p1 = {"first": (-1, 0), "equal": (0, 0)}.get(os.environ.get("OGG_QUAD_PRIO", ""), (0, -1))   # A/B knob
        _POOL[dev] = ([torch.cuda.Stream(priority=p0), torch.cuda.Stream(priority=p1)] +
                      [torch.cuda.Stream(priority=-2) for _ in range(4)], [torch.cuda.Event() for _ in range(8)])
    return _POOL[dev]


class Supergrid(object):
    """The six arrays (device tensors) of one rank's share of the supergrid."""

    def __init__(self, plan, rank=0, world=1, capacity_world=None):
        """capacity_world=V: allocate once for the largest of V row-slabs; `set_slab(v, V)` then re-targets the same
        buffers at slab v (a 1/120-degree grid is 279 GB: one GPU streams it through one slab's worth of HBM)."""
        self.plan = plan
        self._iv = None
        self._lam_dev = {}
        cap = None
        if capacity_world is not None:
            cap = max(sum(slab_bounds(p.rows - 1, v, capacity_world)[1] - slab_bounds(p.rows - 1, v, capacity_world)[0]
                          for p in plan.pieces) for v in range(capacity_world))
        self._set_bounds(rank, world)
        nxp, nyl = plan.nxp, (self.rows_local if cap is None else cap)
        self._cap = nyl
        self._x, self._y, self._angle = _dev.empty(nyl + 1, nxp), _dev.empty(nyl + 1, nxp), _dev.empty(nyl + 1, nxp)
        self._dx = _dev.empty(nyl + 1, nxp - 1)
        self._dy, self._area = _dev.empty(nyl, nxp), _dev.empty(nyl, nxp - 1)

    def _set_bounds(self, rank, world):
        self.rank, self.world = rank, world
        self.slabs = [slab_bounds(p.rows - 1, rank, world) for p in self.plan.pieces]
        self.rows_local = sum(c1 - c0 for c0, c1 in self.slabs)    # cell rows held here

    def set_slab(self, rank, world):
        """Re-target the (already allocated) buffers at slab `rank` of `world`; the previous contents are overwritten
        by the next build()."""
        self._set_bounds(rank, world)
        if self.rows_local > self._cap:
            raise Exception("slab %d of %d needs %d cell rows, buffers hold %d" % (rank, world, self.rows_local, self._cap))
        return self

    # the six arrays, trimmed to the rows of the current slab (views of the capacity buffers)
    x = property(lambda self: self._x[:self.rows_local + 1])
    y = property(lambda self: self._y[:self.rows_local + 1])
    angle_dx = property(lambda self: self._angle[:self.rows_local + 1])
    dx = property(lambda self: self._dx[:self.rows_local + 1])
    dy = property(lambda self: self._dy[:self.rows_local])
    area = property(lambda self: self._area[:self.rows_local])

    # number of vertex rows that are really this rank's (the closing row belongs to the last rank only)
    @property
    def vertex_rows(self):
        return self.rows_local + (1 if self.rank == self.world - 1 else 0)

    def arrays(self):
        n = self.vertex_rows
        return dict(x=self.x[:n], y=self.y[:n], dx=self.dx[:n], dy=self.dy, area=self.area, angle_dx=self.angle_dx[:n])

    def checksums(self, acc=None):
        """Adds this slab's per-array 64-bit checksums (ogg_checksum_accumulate) to `acc` (uint64[6] device tensor, order of
        FIELDS); the sum over all slabs equals the whole grid's checksums."""
        import torch
        if acc is None:
            acc = torch.zeros(6, dtype=torch.int64, device="cuda")
        st = _dev.stream()
        for k, name in enumerate(FIELDS):
            t = self.arrays()[name]
            _cabi.call("ogg_checksum_accumulate", _dev.ptr(t), t.numel(), ctypes.c_void_p(acc.data_ptr() + 8 * k), st)
        return acc

    def build(self, concurrent=True, host=None):
        """Launch every piece's kernels (asynchronous).  The pieces write disjoint rows, so each gets its own stream
        forked from / joined to the current one: the caps' many small table kernels and the HBM-bound band fills
        overlap the FP64-bound cap quadratures instead of queueing behind them.

        host: dict of six pinned host tensors shaped like arrays(); each piece's rows are then copied out on the piece's
        own stream as soon as its kernels are queued, bands first, so that the device->host traffic of the bands (most
        of the bytes) runs under the caps' quadrature.  Synchronise the current stream before reading `host`."""
        import torch
        plan = self.plan
        if plan.skip_metrics:                                       # OGG:1009-1011 etc.: metrics stay -1
            self.dx.fill_(-1.0); self.dy.fill_(-1.0); self.area.fill_(-1.0)
        jobs, off = [], 0
        last = len(plan.pieces) - 1
        for k, (piece, (c0, c1)) in enumerate(zip(plan.pieces, self.slabs)):
            if c1 > c0:
                jobs.append((piece, c0, c1, off, k == last and self.rank == self.world - 1))
            off += c1 - c0
        if not concurrent or len(jobs) < 2:
            band_jobs = [j for j in jobs if j[0].kind == "band"]
            if band_jobs:
                self._build_bands(band_jobs)
            for job in jobs:
                if job[0].kind != "band":
                    self._build_piece(*job)
                if host is not None:
                    self._copy_piece_out(host, *job)
            return self
        # Two FP64-bound quadratures (one per cap) and many short kernels: table set-up, the HBM-bound band fills, the
        # caps' own coordinate meshes and angles.  The quadratures go on two normal-priority streams, everything else on
        # high-priority streams, so that the short kernels' CTAs are placed as soon as a quadrature CTA retires instead
        # of queueing behind 10^4 of them.  With `host`, the bands are queued first and their rows (most of the bytes)
        # start crossing PCIe at once; a cap's rows follow when both of its halves are done.
        main = torch.cuda.current_stream()
        streams, events = _stream_pool()
        fork = events[6]
        fork.record(main)
        caps = [j for j in jobs if j[0].kind != "band"]
        bands = [j for j in jobs if j[0].kind == "band"]
        quads = [] if plan.skip_metrics else [(n % 2, job, "quad") for n, job in enumerate(caps)]
        light_caps = [(2 + n % 2, job, "mesh") for n, job in enumerate(caps)]
        light_bands = [(4, bands, "bands")] if bands else []          # every band of the step in ONE persistent launch
        work = (light_bands + quads + light_caps) if host is not None else (quads + light_caps + light_bands)
        used = set()
        for sidx, job, part in work:
            side = streams[sidx]
            if sidx not in used:
                side.wait_event(fork)
                used.add(sidx)
            with torch.cuda.stream(side):
                if part == "bands":
                    self._build_bands(job)
                    if host is not None:
                        for bjob in job:
                            self._copy_piece_out(host, *bjob)
                    continue
                self._build_piece(*job, part=part)
                if host is not None and part == "mesh":
                    for qidx, qjob, _ in quads:
                        if qjob is job:
                            events[7].record(streams[qidx])
                            side.wait_event(events[7])
                    self._copy_piece_out(host, *job)
        for sidx in sorted(used):
            events[sidx].record(streams[sidx])
            main.wait_event(events[sidx])
        return self

    def _copy_piece_out(self, host, piece, c0, c1, off, closing):
        nv, nc = c1 - c0 + (1 if closing else 0), c1 - c0
        for name, dev, n in (("x", self._x, nv), ("y", self._y, nv), ("dx", self._dx, nv), ("angle_dx", self._angle, nv),
                             ("dy", self._dy, nc), ("area", self._area, nc)):
            host[name][off:off + n].copy_(dev[off:off + n], non_blocking=True)

    def pinned_like(self):
        """Six pinned host tensors shaped like arrays() (allocate once, reuse across builds)."""
        import torch
        return {k: torch.empty(tuple(v.shape), dtype=torch.float64, pin_memory=True) for k, v in self.arrays().items()}

    def _views(self, off, nrows_v):
        nxp = self.plan.nxp
        st = _dev.stream()
        return (_dev.ptr_off(self.x, off * nxp), _dev.ptr_off(self.y, off * nxp), _dev.ptr_off(self.dx, off * (nxp - 1)),
                _dev.ptr_off(self.dy, off * nxp), _dev.ptr_off(self.area, off * (nxp - 1)),
                _dev.ptr_off(self.angle_dx, off * nxp), st)

    def _band_desc(self, piece, c0, c1, off, closing):
        plan, nxp = self.plan, self.plan.nxp
        px, py, pdx, pdy, parea, pangle, _ = self._views(off, c1 - c0 + 1)
        key = id(piece)
        if key not in self._lam_dev:
            self._lam_dev[key] = (_dev.to_dev(piece.p["lam"]), _dev.to_dev(piece.p["phi"]))
        lam, phi = self._lam_dev[key]
        flags = 0 if closing else _cabi.OGG_BAND_SKIP_TOP_VERTEX_ROW
        if plan.skip_metrics:
            flags |= _cabi.OGG_BAND_NO_METRICS
            pdx = pdy = parea = None
        v = lambda q: q.value if q is not None else None            # noqa: E731
        return _cabi.OggBand(v(_dev.ptr(lam)), v(_dev.ptr_off(phi, piece.row0 + c0)), c1 - c0 + 1, nxp, flags,
                             v(px), v(py), v(pdx), v(pdy), v(parea), v(pangle))

    def _build_bands(self, jobs):
        """All band pieces of this rank's share in one launch of the fill kernel (ogg_band_fill_metrics_batch)."""
        for k in range(0, len(jobs), 4):
            group = jobs[k:k + 4]
            arr = (_cabi.OggBand * len(group))(*[self._band_desc(*job) for job in group])
            _cabi.call("ogg_band_fill_metrics_batch", arr, len(group), Re, _dev.stream())

    def _build_piece(self, piece, c0, c1, off, closing, part="all"):
        """part: "all", or for a cap "mesh" (x, y, angle_dx) / "quad" (dx, dy, area) -- independent outputs."""
        plan, nxp = self.plan, self.plan.nxp
        px, py, pdx, pdy, parea, pangle, st = self._views(off, c1 - c0 + 1)
        if plan.skip_metrics:
            pdx = pdy = parea = None
        if piece.kind == "band":
            key = id(piece)
            if key not in self._lam_dev:
                self._lam_dev[key] = (_dev.to_dev(piece.p["lam"]), _dev.to_dev(piece.p["phi"]))
            lam, phi = self._lam_dev[key]
            j0 = piece.row0 + c0
            flags = 0 if closing else _cabi.OGG_BAND_SKIP_TOP_VERTEX_ROW
            if plan.skip_metrics:
                flags |= _cabi.OGG_BAND_NO_METRICS
            _cabi.call("ogg_band_fill_metrics", _dev.ptr(lam), _dev.ptr_off(phi, j0), c1 - c0 + 1, nxp, Re, flags,
                       px, py, pdx, pdy, parea, pangle, st)
        elif piece.kind == "bipolar":
            # vertex rows written: this slab's lower edges, plus its top edge only when it closes the grid -- any other
            # top row belongs to the slab / sub-grid above, which writes it (possibly concurrently, on another stream)
            p = piece.p
            j0, j1 = piece.row0 + c0, piece.row0 + c1
            nv = c1 - c0 + (1 if closing else 0)
            if part != "quad":
                _cabi.call("ogg_bipolar_cap_mesh", plan.Ni, p["Nj"], j0, j0 + nv, p["lat0_bp"], p["lon_bp"], p["rp"],
                           px, py, None, None, st)
                _cabi.call("ogg_angle_x", px, py, nv, nxp, pangle, st)                            # OGG:1071
            if not plan.skip_metrics and part != "mesh":
                _cabi.call("ogg_bipolar_cap_metrics_quad", 5, plan.Ni, p["Nj"], j0, j1, p["lat0_bp"], p["lon_bp"],
                           p["rp"], Re, 0 if closing else _cabi.OGG_QUAD_NO_TOP_DX_ROW, pdx, pdy, parea, st)   # OGG:1068
        else:  # displaced pole
            p = piece.p
            j0, j1 = piece.row0 + c0, piece.row0 + c1
            nv = c1 - c0 + (1 if closing else 0)
            if part != "quad":
                _cabi.call("ogg_dp_cap_mesh", plan.Ni, p["Nj"], j0, j0 + nv, p["lon0"], p["lat0"], p["lon_dp"], p["r_dp"],
                           px, py, st)                                                             # OGG:1162
                _cabi.call("ogg_angle_x", px, py, nv, nxp, pangle, st)                            # OGG:1163
            if not plan.skip_metrics and part != "mesh":
                _cabi.call("ogg_dp_metrics_quad", 4, plan.Ni, p["Nj"], j0, j1, p["lon0"], p["lat0"], p["lon_dp"],
                           p["r_dp"], Re, 0 if closing else _cabi.OGG_QUAD_NO_TOP_DX_ROW, pdx, pdy, parea, st)  # OGG:1170

    # ---- reductions / checks -------------------------------------------------------------------------
    def area_sum(self):
        return ogg.device_sum(self.area) if self.area.numel() else 0.0

    def to_host(self):
        return {k: _dev.to_host(v) for k, v in self.arrays().items()}


def build_supergrid(plan, rank=0, world=1):
    return Supergrid(plan, rank, world).build()


def generate(host=None, **kwargs):
    """Public one-call API: plan + build on this GPU; returns (plan, Supergrid) with arrays resident in HBM.
    host: pinned tensors from Supergrid.pinned_like() to receive the six arrays (copies overlap the build)."""
    _dev.ready()
    plan = plan_supergrid(**kwargs)
    return plan, Supergrid(plan).build(host=host)


# --------------------------------------------------------------------------------------------------
# main(): the CHECK_metrics self-check lines, --write_subgrid_files, the stitched file
# --------------------------------------------------------------------------------------------------
def build_piece_standalone(plan, piece):
    sub = Plan(plan.Ni, [piece], plan.skip_metrics, plan.info)
    return build_supergrid(sub)


def _stitched_views(plan, sg, name):
    """(dx, dy, area) of sub-grid `name` as row ranges of the stitched arrays, or None when the stitched grid does not
    hold what metrics_error reads from that sub-grid as main() generated it (rows cut away, or its own top dx row)."""
    for piece, off in zip(plan.pieces, plan.offsets()):
        if piece.name != name:
            continue
        full = plan.info["as_generated"][name]
        if piece.row0 != 0 or piece.rows_full != full.rows_full or piece.kind == "dp":
            return None
        if piece.kind == "band" and piece.p.get("regular"):
            return None                      # OGG:1157 reads the regular cap's OWN top dx row, which the join drops
        closes = piece is plan.pieces[-1]
        if piece.kind == "bipolar" and not closes:
            return None
        n = piece.rows - 1
        return sg.dx[off:off + n + 1], sg.dy[off:off + n], sg.area[off:off + n]
    return None


def check_line(plan, name, sg=None, standalone=None):
    """Prints the reference's self-check line of one sub-grid (OGG:1016-1017, 1069-1070, 1111-1112, 1156-1157,
    1172-1174), evaluated on the sub-grid as main() first generates it.  The sums come from the stitched arrays where
    they hold those rows, from a stand-alone build of the sub-grid (`standalone`, or made here) otherwise, and for the
    displaced-pole cap -- whose rows inside the doughnut are never stored -- from a chunked reduction over the whole cap."""
    full = plan.info["as_generated"].get(name)
    if full is None or plan.skip_metrics:
        return
    if full.kind == "dp":
        p = full.p
        poles_i = int(plan.Ni * numpy.mod(p["lon_dp"] - p["lon0"], 360) / 360.0)                     # OGG:1171
        sums = ogg.displacedPoleCap_metrics_sums(4, plan.Ni, p["Nj"], p["lon0"], p["lat0"], p["lon_dp"], p["r_dp"], poles_i)
        err = ogg.metrics_error_from_sums(sums, p["lat0"], -90.0, displaced_pole=poles_i,
                                          excluded_fraction=plan.info["doughnut"])
        print("   CHECK_metrics_hquad: % errors in (area, lat arc, lon arc)", err)
        return
    views = _stitched_views(plan, sg, name) if sg is not None else None
    if views is None:
        if standalone is None:
            standalone = build_piece_standalone(plan, full)
        a = standalone.arrays()
        views = a["dx"], a["dy"], a["area"]
    dx_, dy_, area_ = views
    if full.kind == "bipolar":
        print("   CHECK_metrics_hquad: % errors in (area, lat arc, lon arc1, lon arc2)",
              ogg.metrics_error(dx_, dy_, area_, plan.Ni, full.p["lat0_bp"], 90.0, bipolar=True))
        return
    phi = full.p["phi"]
    lat1, lat2 = phi[0], phi[-1]
    if full.p.get("regular"):
        lat1, lat2 = phi[-1], phi[0]                                                # OGG:1157
    label = "CHECK_metrics" if name == "Merc" else "CHECK_metrics_MIDAS"
    print("   %s: %% errors in (area, lat arc, lon arc)" % label, ogg.metrics_error(dx_, dy_, area_, plan.Ni, lat1, lat2))


def run_main(args):
    """Body of main(**args) (OGG:855-1449)."""
    _dev.ready()
    match_dy, grids = args["match_dy"], args["grids"]
    known_options = ["bp", "so", "p125sc", ""]
    unknown = list(set(match_dy).difference(known_options))
    if len(unknown) != 0:                                                           # OGG:883-888
        print("Unknown options in match_dy: ", unknown)
        print("Known options are one or more of ", known_options)
        sys.exit(2)
    if args["r_dp"] != 0.0 and args["lat_dp"] > -90.0:                              # OGG:904-907
        print("Cannot specify both --rdp and --latdp for the displaced pole!")
        ogg.usage()
        sys.exit(2)
    gridfilename, no_changing_meta, debug = args["gridfilename"], args["no_changing_meta"], args["debug"]
    hist, desc, source = _metadata(args)
    start_time = time.time()
    plan_args = {k: v for k, v in args.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta",
                                                             "write_subgrid_files", "reproduce_MIDAS_grids")}
    plan = plan_supergrid(**plan_args)
    info = plan.info
    if len(plan.pieces) < 2:
        raise Exception("nothing to stitch: the reference needs at least two adjoining sub-grids (OGG:1237-1247)")
    sg = build_supergrid(plan)               # one in-place build of the stitched grid; everything below reads it
    # OGG_CHECK_LINES=0: leave out the CHECK_metrics lines.  They are the reference's self-check prints, not part of the
    # file; for the displaced-pole cap they cost a second pass over the whole cap's quadrature (the rows inside the
    # doughnut are never stored), and for sub-grids the stitched arrays cannot serve (a regular southern cap, a cut
    # Southern Ocean) a stand-alone build next to the resident grid -- at 1/30 degree and finer that is most of the
    # CLI's cap work and extra HBM a user may not have.
    import os
    check_lines = os.environ.get("OGG_CHECK_LINES", "1") != "0"
    if not check_lines:
        print("   self-check lines left out (OGG_CHECK_LINES=0)")
    # Per sub-grid, in main()'s order (Mercator, bipolar cap, Southern Ocean, southern cap): the CHECK_metrics line and,
    # with --write_subgrid_files, its own file.  A sub-grid's file holds its OWN closing vertex row (which the join
    # drops) and the rows main() had at that point -- Mercator / bipolar / Southern Ocean as generated (OGG:1020, 1074,
    # 1115: before any south cut), the southern cap after the doughnut cut (OGG:1228) -- so the files come from
    # stand-alone builds, which then also serve the check lines.
    for name in ("Merc", "BP", "SO", "SC"):
        if name not in info["as_generated"]:
            continue
        file_piece = info.get("SC_file_piece") if name == "SC" else info["as_generated"][name]
        standalone = None
        if args["write_subgrid_files"]:
            standalone = build_piece_standalone(plan, file_piece)
        full_build = standalone if (standalone is not None and file_piece.row0 == 0) else None
        if check_lines:
            check_line(plan, name, sg, full_build)
        if standalone is not None:
            a = standalone.arrays()
            ogg.write_nc(a["x"], a["y"], a["dx"], a["dy"], a["area"], a["angle_dx"], axis_units="degrees",
                         fnam=gridfilename + name + ".nc", description=desc, history=hist, source=source, debug=debug)
            del standalone, a
    print("Stitching the grids together...")
    a = sg.arrays()
    ycol = _dev.to_host(a["y"][:, plan.Ni // 4].contiguous())
    if info["hasBP"]:
        dy3_ = numpy.roll(ycol, shift=-1, axis=0) - ycol                            # OGG:1366-1370
        if numpy.any(dy3_ == 0):
            raise Exception("lattitude array has repeated values along symmetry meridian!")
    # description (OGG:1403-1423)
    desc = desc + "It consists of; "
    if info["hasMerc"]:
        desc += "a Mercator grid spanning " + str(info["phiMerc"][0]) + " to " + str(info["phiMerc"][1]) + " degrees; "
    if info["hasBP"]:
        desc += "a bipolar northern cap north of " + str(info["phiMerc"][1]) + " degrees; "
    if info["hasSO"]:
        desc += "a regular lat-lon grid spanning " + str(info["latUp_SO"]) + " to " + str(info["lat0_SO"]) + " degrees; "
    if info["hasSC"]:
        desc += "a " + ("displaced pole " if info["r_dp"] != 0.0 else "regular ")
        desc += "southern cap south of " + str(info["lat0_SO"]) + " degrees."
    if args["south_cutoff_ang"] > -90:
        desc += " It is cut south of " + str(args["south_cutoff_ang"]) + " degrees."
    if args["south_cutoff_row"] > 0:
        desc += " The first " + str(args["south_cutoff_row"]) + " rows at south are deleted."
    equator_index = numpy.searchsorted(ycol, 0.0)                                   # OGG:1425-1436
    if equator_index == 0:
        raise Exception("   Ooops: Equator is not in the grid")
    print("   Equator is at j=", equator_index)
    if equator_index % 2 == 0:
        raise Exception("Ooops: Equator is not going to be a u-point. Use option --south_cutoff_row to one more or on less row from south.")
    if a["y"].shape[0] % 2 == 0:
        raise Exception("Ooops: The number of j's in the supergrid is not even. Use option --south_cutoff_row to one more or on less row from south.")
    print("shapes: ", tuple(a["x"].shape), tuple(a["y"].shape), tuple(a["dx"].shape), tuple(a["dy"].shape),
          tuple(a["area"].shape), tuple(a["angle_dx"].shape))
    ogg.write_nc(a["x"], a["y"], a["dx"], a["dy"], a["area"], a["angle_dx"], axis_units="degrees", fnam=gridfilename,
                 description=desc, history=hist, source=source, no_changing_meta=no_changing_meta, debug=debug)
    print("Wrote the whole grid to file ", gridfilename)
    print("runtime(secs)  %s" % (time.time() - start_time))
    return sg


def _metadata(args):
    """history / description / source strings (OGG:909-963)."""
    import datetime
    import os
    import socket
    import subprocess
    hist = "This grid file was generated via command " + " ".join(sys.argv)
    desc = ("This is an orthogonal coordinate grid for the Earth with a nominal resoution of "
            + str(1 / args["inverse_resolution"]) + " degrees along the equator. ")
    source = ""
    if not args["no_changing_meta"]:
        hist += " on " + str(datetime.date.today()) + " on platform " + str(socket.gethostname())
        here = os.path.dirname(os.path.abspath(__file__))
        try:
            githash = subprocess.check_output("cd " + here + ";git rev-parse HEAD; exit 0", stderr=subprocess.STDOUT,
                                              shell=True).decode("ascii").rstrip("\n")
        except Exception:
            githash = "unknown"
        source = (sys.argv[0] + " had git hash " + githash
                  + ". B200-native build of the ocean_model_grid_generator supergrid path")
    return hist, desc, source

"""CPU composition of a whole default-flag tripolar supergrid from the C oracle -- TEST INFRASTRUCTURE.

Follows the reference's main() for the flags used by the benchmark workload and the golden configs
(`-r R [--r_dp X] [--south_cutoff_row K]`, everything else default): sizing OGG:969-1195, doughnut cut
OGG:1177-1186, south cut OGG:1270-1295, joins OGG:1316-1365.  Used (a) by tests to check the oracle against
the reference-made per-array hashes, (b) by bench.py as the timed CPU baseline.  Never imported by the product.

slab=(rank, world): only rank's contiguous share of every sub-grid's cell rows is evaluated (the work one of `world`
GPUs does per step), through the reference's own decomposition of the cap metrics (index vectors -> h samples ->
cell sums): a bounded, representative sample of a large grid for the timed CPU arm.  The slabs of all ranks, stacked
per sub-grid, equal the whole-grid result bit for bit (tests/test_oracle_cpu.py).
"""
import numpy as np

import ogg_oracle as O


def _bounds(ncells, rank, world):
    base, rem = divmod(int(ncells), int(world))
    c0 = rank * base + min(rank, rem)
    return c0, c0 + base + (1 if rank < rem else 0)


def build_cpu(r, r_dp=0.0, south_cutoff_row=0, exfracdp=0.49, timings=None, slab=None, ensure_nj_even=False):
    """ensure_nj_even (the CLI flag, OGG:873): coded for what it changes on a default-flag integer-r grid -- the Mercator
    row count (OGG:329-333) and the doughnut-cut displaced-pole cap (OGG:1188-1194).  The bipolar cap's 120*r rows are
    even already (OGG:107); a Southern Ocean band or regular cap that the flag would shorten raises (not coded here)."""
    import time
    t0 = time.time()
    refineS, lon0, lenlon = 2, -300.0, 360
    Ni = int(r * refineS * lenlon)
    phi_s, phi_n = (-68.0, 65.0) if r == 2 else (-66.85954725, 64.05895973)
    phiM = O.mercator_phi_1d(Ni, phi_s, phi_n, True, bool(ensure_nj_even))
    lam1 = lon0 + np.arange(Ni + 1) * lenlon / float(Ni)
    rank, world = slab if slab is not None else (0, 1)

    def rows_of(nrows_vertex, row0=0):
        """vertex rows [v0, v1] (inclusive) of this slab for a piece whose kept vertex rows are [row0, nrows_vertex)"""
        c0, c1 = _bounds(nrows_vertex - 1 - row0, rank, world)
        return row0 + c0, row0 + c1

    def band(phi, row0=0):
        v0, v1 = rows_of(phi.shape[0], row0)
        ph = phi[v0:v1 + 1]
        x = np.tile(lam1, (ph.shape[0], 1))
        y = np.tile(ph.reshape(-1, 1), (1, Ni + 1))
        dx, dy, area = O.generate_grid_metrics_MIDAS(x, y)
        return [x, y, dx, dy, area, O.angle_x(x, y)]

    lat0_bp = phiM[-1]
    Nj_ncap = 119 * refineS if r == 2 else int(60 * r * refineS)
    lat0_SO = -78.0
    lenlat_SO = phiM[0] - lat0_SO
    Nj_SO = int(r * 55)
    phiSO = lat0_SO + np.arange(Nj_SO + 1) * lenlat_SO / float(Nj_SO)
    if ensure_nj_even and (Nj_ncap % 2 != 0 or (Nj_SO + 1) % 2 == 0):
        raise ValueError("ensure_nj_even would cut a bipolar / Southern Ocean row at r=%r: not coded in this test helper" % r)
    lat0_SC = phiSO[0]
    Nj_scap = int(r * 40) * 7 // 4
    # south cut (OGG:1270-1313) as row offsets into SC / SO
    sc_row0, so_row0, has_sc = 0, 0, True
    if r_dp == 0.0:
        Nj_scap = int((lat0_SC + 90.0) / (1.0 / r / refineS))
        phiSC = -90.0 + np.arange(Nj_scap + 1) * (90 + lat0_SO) / float(Nj_scap)
        if ensure_nj_even and (Nj_scap + 1) % 2 == 0:
            raise ValueError("ensure_nj_even would cut a row of the regular southern cap at r=%r: not coded here" % r)
    else:
        jmin = np.ceil(exfracdp * Nj_scap)
        sc_row0 = int(jmin + np.mod(jmin, 2))
        if ensure_nj_even and (Nj_scap + 1 - sc_row0) % 2 == 0:                                        # OGG:1188-1194
            sc_row0 += 1
    sc_rows = Nj_scap + 1 - sc_row0
    if south_cutoff_row > 0:
        jcut = south_cutoff_row - 1
        if jcut < sc_rows:
            sc_row0 += jcut
        else:
            has_sc, so_row0 = False, jcut - sc_rows

    merc = band(phiM)
    # bipolar cap (OGG:103-188)
    rp = O.bipolar_rp(lat0_bp)
    if slab is None:
        xb, yb, _, _ = O.generate_bipolar_cap_mesh(Ni, Nj_ncap, lat0_bp, lon0, ensure_nj_even=False)
        dxb, dyb, ab = O.bipolar_cap_metrics_quad_fast(5, Ni, Nj_ncap, lat0_bp, lon0, rp)
    else:
        v0, v1 = rows_of(Nj_ncap + 1)
        lamg = np.tile(lon0 + np.arange(Ni + 1) * 360.0 / float(Ni), (v1 - v0 + 1, 1))                  # OGG:113-116
        phig = np.tile((lat0_bp + np.arange(v0, v1 + 1) * (90 - lat0_bp) / float(Nj_ncap)).reshape(-1, 1), (1, Ni + 1))
        xb, yb, _, _ = O.bipolar_projection(lamg, phig, lon0, rp)
        j1d = O.sample_index(5, v0, v1, Nj_ncap, fudge=True)                                            # OGG:143-147
        i1d = O.sample_index(5, 0, Ni, Ni)
        hi, hj = O.bipolar_cap_ij_array(i1d, j1d, Ni, Nj_ncap, lat0_bp, lon0, rp)                       # OGG:158
        dxb, dyb, ab = O.quad_cells(5, Ni, v1 - v0, hi, hj)
    bp = [xb, yb, dxb, dyb, ab, O.angle_x(xb, yb)]
    so = band(phiSO, so_row0)
    sc = None
    if has_sc and r_dp == 0.0:
        sc = band(phiSC, sc_row0)
    elif has_sc:
        if slab is None:
            xs, ys, _, _ = O.generate_displaced_pole_grid(Ni, Nj_scap, lon0, lat0_SC, 80.0, r_dp)
            dxs, dys, asx = O.displacedPoleCap_metrics_quad(4, Ni, Nj_scap, lon0, lat0_SC, 80.0, r_dp)
            sc = [a[sc_row0:] for a in (xs, ys, dxs, dys, asx, O.angle_x(xs, ys))]
        else:
            v0, v1 = rows_of(Nj_scap + 1, sc_row0)
            xs, ys, _, _ = O.displacedPoleCap_mesh(np.arange(Ni + 1), np.arange(v0, v1 + 1), Ni, Nj_scap, lon0, lat0_SC, 80.0, r_dp)
            j1d = O.sample_index(4, v0, v1, Nj_scap)                                                    # OGG:573-581
            i1d = O.sample_index(4, 0, Ni, Ni)
            hx = O.numerical_hi(j1d, i1d, Ni, Nj_scap, lon0, lat0_SC, 80.0, r_dp, 1e-3, 4)             # OGG:583
            hy = O.numerical_hj(j1d, i1d, Ni, Nj_scap, lon0, lat0_SC, 80.0, r_dp, 1e-3, 4)             # OGG:584
            dxs, dys, asx = O.quad_cells(4, Ni, v1 - v0, hx, hy)
            sc = [xs, ys, dxs, dys, asx, O.angle_x(xs, ys)]
    out = None
    for piece in (sc, so, merc, bp):
        if piece is None:
            continue
        if out is None:
            out = piece
            continue
        lo, up = out, piece
        out = [np.concatenate((lo[k][:-1], up[k]), axis=0) if k in (0, 1, 2, 5) else np.concatenate((lo[k], up[k]), axis=0)
               for k in range(6)]
    if timings is not None:
        timings.append(time.time() - t0)
    return dict(zip(("x", "y", "dx", "dy", "area", "angle_dx"), out))

"""CPU composition of a whole default-flag tripolar supergrid from the C oracle -- TEST INFRASTRUCTURE.

Follows the reference's main() for the flags used by the benchmark workload and the golden configs
(`-r R [--r_dp X] [--south_cutoff_row K]`, everything else default): sizing OGG:969-1195, doughnut cut
OGG:1177-1186, south cut OGG:1270-1295, joins OGG:1316-1365.  Used (a) by tests to check the oracle against
the reference-made per-array hashes, (b) by bench.py as the timed CPU baseline.  Never imported by the product.
"""
import numpy as np

import ogg_oracle as O


def build_cpu(r, r_dp=0.0, south_cutoff_row=0, exfracdp=0.49, timings=None):
    import time
    t0 = time.time()
    refineS, lon0, lenlon = 2, -300.0, 360
    Ni = int(r * refineS * lenlon)
    phi_s, phi_n = (-68.0, 65.0) if r == 2 else (-66.85954725, 64.05895973)
    phiM = O.mercator_phi_1d(Ni, phi_s, phi_n, True, False)
    lam1 = lon0 + np.arange(Ni + 1) * lenlon / float(Ni)

    def band(phi):
        x = np.tile(lam1, (phi.shape[0], 1))
        y = np.tile(phi.reshape(-1, 1), (1, Ni + 1))
        dx, dy, area = O.generate_grid_metrics_MIDAS(x, y)
        return [x, y, dx, dy, area, O.angle_x(x, y)]

    merc = band(phiM)
    lat0_bp = phiM[-1]
    Nj_ncap = 119 * refineS if r == 2 else int(60 * r * refineS)
    xb, yb, _, _ = O.generate_bipolar_cap_mesh(Ni, Nj_ncap, lat0_bp, lon0, ensure_nj_even=False)
    rp = O.bipolar_rp(lat0_bp)
    dxb, dyb, ab = O.bipolar_cap_metrics_quad_fast(5, Ni, Nj_ncap, lat0_bp, lon0, rp)
    bp = [xb, yb, dxb, dyb, ab, O.angle_x(xb, yb)]
    lat0_SO = -78.0
    lenlat_SO = phiM[0] - lat0_SO
    Nj_SO = int(r * 55)
    phiSO = lat0_SO + np.arange(Nj_SO + 1) * lenlat_SO / float(Nj_SO)
    so = band(phiSO)
    lat0_SC = phiSO[0]
    Nj_scap = int(r * 40) * 7 // 4
    if r_dp == 0.0:
        Nj_scap = int((lat0_SC + 90.0) / (1.0 / r / refineS))
        phiSC = -90.0 + np.arange(Nj_scap + 1) * (90 + lat0_SO) / float(Nj_scap)
        sc = band(phiSC)
    else:
        xs, ys, _, _ = O.generate_displaced_pole_grid(Ni, Nj_scap, lon0, lat0_SC, 80.0, r_dp)
        dxs, dys, asx = O.displacedPoleCap_metrics_quad(4, Ni, Nj_scap, lon0, lat0_SC, 80.0, r_dp)
        sc = [xs, ys, dxs, dys, asx, O.angle_x(xs, ys)]
        jmin = np.ceil(exfracdp * Nj_scap)
        jmint = int(jmin + np.mod(jmin, 2))
        sc = [a[jmint:] for a in sc]
    if south_cutoff_row > 0:
        jcut = south_cutoff_row - 1
        if jcut < sc[0].shape[0]:
            sc = [a[jcut:] for a in sc]
        else:
            jcut_so = jcut - sc[0].shape[0]
            sc = None
            so = [a[jcut_so:] for a in so]
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

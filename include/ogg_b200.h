/*
 * ogg_b200.h -- C ABI of libogg_b200.so: the B200 (sm_100a) supergrid hot path of
 * ocean_model_grid_generator.
 *
 * Each entry point stands where one Python function of the reference stands
 * (ocean_grid_generator.py, cited as OGG:first-last).  Conventions:
 *   - every `double*` / `const double*` is a DEVICE pointer owned by the caller, row-major, j (latitude
 *     row) outer, i (longitude) inner; the library never frees caller memory;
 *   - sizes are int64_t, reals are double (FP64 throughout, FMA contraction disabled in the kernels);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); calls are asynchronous
 *     with respect to the host unless stated; scratch memory is stream-ordered (cudaMallocAsync);
 *   - return value: 0 on success, OGG_ERR_* below, or OGG_ERR_CUDA - (int)cudaError_t.  Nothing throws.
 * There is no CPU fallback: without a CUDA device every compute entry point returns an error.
 *
 * Devices and threads.  Every call works on the CURRENT device of the calling thread (cudaSetDevice), and `stream`
 * must belong to it; one process may drive several devices (per-device state -- SM count, the shared-memory opt-in
 * of the fused bipolar kernel, the scratch pool, timing events -- is keyed by device id).  Scratch comes from a
 * private cudaMemPool per device whose release threshold is "never"; the application's default pool is not touched.
 * Compute entry points may be called from several host threads at once; the launch counter and the timing brackets
 * (ogg_launch_count, ogg_timing_*) are process-wide, mutex-guarded, and ogg_timing_get / _pair_span report the
 * brackets of the current device only.  A call that fails half-way may have queued some of its launches already.
 */
#ifndef OGG_B200_H_
#define OGG_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OGG_OK 0
#define OGG_ERR_UNCODED_ORDER (-1)   /* reference: Exception("Uncoded order")      OGG:204,222,255 */
#define OGG_ERR_ORDER_NOT_CODED (-2) /* reference: Exception("order not coded")    OGG:547,562     */
#define OGG_ERR_SHAPE (-3)           /* reference: shape-mismatch exceptions        OGG:229,722     */
#define OGG_ERR_ARG (-4)
#define OGG_ERR_TOO_LARGE (-5)       /* request exceeds the fixed work array of a *_noalloc entry point (see below) */
#define OGG_ERR_CUDA (-1000)

/* flags of ogg_band_fill_metrics */
#define OGG_BAND_SKIP_TOP_VERTEX_ROW 1 /* do not write row nyp-1 of x, y, dx, angle_dx (the "[:-1]" of OGG:1316-1362) */
#define OGG_BAND_NO_METRICS 2          /* --skip_metrics: dx, dy, area are left untouched                               */

const char* ogg_version(void);
/* number of CUDA devices visible (0 when there is no GPU; never fails) */
int ogg_device_count(void);

/* ---- portable elementary functions, element-wise (numpypi.numpypi_series stand-ins; OGG:5) ------------- */
int ogg_sin(const double* in, double* out, int64_t n, void* stream);
int ogg_cos(const double* in, double* out, int64_t n, void* stream);
int ogg_tan(const double* in, double* out, int64_t n, void* stream);
int ogg_arcsin(const double* in, double* out, int64_t n, void* stream);
int ogg_arccos(const double* in, double* out, int64_t n, void* stream);
int ogg_arctan(const double* in, double* out, int64_t n, void* stream);
int ogg_exp(const double* in, double* out, int64_t n, void* stream);
int ogg_log(const double* in, double* out, int64_t n, void* stream);
int ogg_sinh(const double* in, double* out, int64_t n, void* stream);
int ogg_cosh(const double* in, double* out, int64_t n, void* stream);
int ogg_mod360(const double* in, double* out, int64_t n, void* stream); /* numpy.mod(x, 360.0), OGG:684 */
int ogg_arctan2(const double* y, const double* x, double* out, int64_t n, void* stream);
/* y_mercator (OGG:292-295): out = R * log((1 + sin(phi)) / cos(phi)), phi in radians, R = Ni / (2 pi) from the caller.
 * phi_mercator (OGG:298-301): out = arctan(sinh(y / R)) * (180 / pi); y == NULL means y[t] = y_first + t, the
 * numpy.arange(y_star[0], y_star[1] + 1) of OGG:336.  One launch each instead of one per elementary function. */
int ogg_y_mercator(const double* phi, int64_t n, double R, double* out, void* stream);
int ogg_phi_mercator(const double* y, int64_t n, double y_first, double R, double* out, void* stream);

/* ---- spherical bands (Mercator / Southern Ocean / regular cap) ------------------------------------------ */
/* Fused tile + MIDAS metrics + angle for a band whose mesh is the outer product of lam[nxp] and phi[nyp]:
 * replaces np.tile (OGG:430-432, 840-841) + generate_grid_metrics_MIDAS (OGG:687-716, latlon_areafix=True)
 * + angle_x (OGG:719-729).  Outputs (any may be NULL): x, y, angle [nyp][nxp]; dx [nyp][nxp-1];
 * dy [nyp-1][nxp]; area [nyp-1][nxp-1]. */
int ogg_band_fill_metrics(const double* lam, const double* phi, int64_t nyp, int64_t nxp, double Re, int flags,
                          double* x, double* y, double* dx, double* dy, double* area, double* angle, void* stream);
/* Several bands in ONE persistent launch (the Southern Ocean and Mercator bands of a supergrid step, plus a regular
 * southern cap): same outputs per band as ogg_band_fill_metrics.  `bands` is a HOST array of nbands <= 4 descriptors
 * whose pointers are device pointers. */
typedef struct {
  const double *lam, *phi;
  int64_t nyp, nxp;
  int flags;
  double *x, *y, *dx, *dy, *area, *angle;
} ogg_band_t;
int ogg_band_fill_metrics_batch(const ogg_band_t* bands, int nbands, double Re, void* stream);
/* generate_grid_metrics_MIDAS (OGG:687-716) on a supplied mesh.  halo_x/halo_y (both or neither; [nxp]) are
 * the mesh row that follows row nyp-1 on the next row-slab; when given, dy and area get nyp rows instead of
 * nyp-1. */
int ogg_metrics_midas(const double* x, const double* y, int64_t nyp, int64_t nxp, double Re, int latlon_areafix,
                      const double* halo_x, const double* halo_y, double* dx, double* dy, double* area,
                      void* stream);
/* angle_x (OGG:719-729) */
int ogg_angle_x(const double* x, const double* y, int64_t nyp, int64_t nxp, double* angle, void* stream);

/* ---- bipolar (Murray tripolar) Arctic cap ----------------------------------------------------------------- */
/* bipolar_projection (OGG:33-100), element-wise on n points.  metrics_only!=0: lams, phis may be NULL. */
int ogg_bipolar_projection(const double* lamg, const double* phig, int64_t n, double lon_bp, double rp,
                           int metrics_only, double* lams, double* phis, double* h_i_inv, double* h_j_inv,
                           void* stream);
/* generate_bipolar_cap_mesh (OGG:103-122), mesh rows [j0, j1) of the (Nj+1) x (Ni+1) cap (row drop for
 * ensure_nj_even is the caller's).  lams, phis [j1-j0][Ni+1]; h_i_inv [j1-j0][Ni] and
 * h_j_inv [min(j1,Nj)-j0][Ni+1] may be NULL. */
int ogg_bipolar_cap_mesh(int64_t Ni, int64_t Nj, int64_t j0, int64_t j1, double lat0_bp, double lon_bp, double rp,
                         double* lams, double* phis, double* h_i_inv, double* h_j_inv, void* stream);
/* bipolar_cap_ij_array (OGG:125-133): fractional index vectors i[ni], j[nj] -> h_i_inv, h_j_inv [nj][ni] */
int ogg_bipolar_cap_ij_array(const double* i, int64_t ni, const double* j, int64_t nj, int64_t Ni, int64_t Nj_ncap,
                             double lat0_bp, double lon_bp, double rp, double* h_i_inv, double* h_j_inv,
                             void* stream);
/* bipolar_cap_metrics_quad_fast (OGG:136-188) for cell rows [j0, j1) of an nx x ny cap.
 * dxq [(j1-j0) + (j1==ny)][nx]  (the extra row is the pole row dxq[ny,:]);  dyq [j1-j0][nx+1];  daq [j1-j0][nx].
 * flags: OGG_QUAD_NO_TOP_DX_ROW leaves dxq[ny,:] unwritten (in a stitched grid that row belongs to the sub-grid above,
 * the [:-1] of OGG:1316-1362, which may be writing it concurrently on another stream). */
#define OGG_QUAD_NO_TOP_DX_ROW 1
int ogg_bipolar_cap_metrics_quad(int order, int64_t nx, int64_t ny, int64_t j0, int64_t j1, double lat0_bp,
                                 double lon_bp, double rp, double Re, int flags, double* dxq, double* dyq,
                                 double* daq, void* stream);

/* ---- quadrature helpers (OGG:191-255), batched --------------------------------------------------------------- */
/* a[n], b[n] are HOST pointers */
int ogg_quad_positions(int n, double* a_host, double* b_host);
/* quad_average on m vectors y[m][n] -> out[m] */
int ogg_quad_average(const double* y, int64_t m, int n, double* out, void* stream);
/* quad_average_2d on m squares y[m][n][n] -> out[m] */
int ogg_quad_average_2d(const double* y, int64_t m, int n, double* out, void* stream);

/* ---- displaced-pole southern cap ---------------------------------------------------------------------------------- */
/* r_joint and z_0 of OGG:494-495 -> out3 = {r_joint, Re z_0, Im z_0} (device) */
int ogg_dp_constants(double lat0, double lam_pole, double r_pole, double* out3, void* stream);
/* displacedPoleCap_projection (OGG:447-467) incl. monotonic_bounding with x_0 = lon_grid[0,0] */
int ogg_dp_projection(const double* lon_grid, const double* lat_grid, int64_t nj, int64_t ni, double z0_re,
                      double z0_im, double r_joint, double* lam, double* phi, void* stream);
/* monotonic_bounding (OGG:470-475), in place on x[nj][ni] */
int ogg_monotonic_bounding(double* x, int64_t nj, int64_t ni, double x_0, void* stream);
/* displacedPoleCap_mesh (OGG:478-506, excluded_fraction=None) on index vectors i[nil], j[njl] */
int ogg_dp_mesh(const double* i, int64_t nil, const double* j, int64_t njl, int64_t ni, int64_t nj, double lon0,
                double lat0, double lam_pole, double r_pole, double* lams, double* phis, void* stream);
/* generate_displaced_pole_grid (OGG:509-518) restricted to mesh rows [j0, j1) of the (Nj+1) x (Ni+1) cap: the integer
 * index vectors np.arange(Ni+1), np.arange(j0, j1) are formed on the device.  lams, phis [j1-j0][Ni+1]. */
int ogg_dp_cap_mesh(int64_t Ni, int64_t Nj, int64_t j0, int64_t j1, double lon0, double lat0, double lam_pole,
                    double r_pole, double* lams, double* phis, void* stream);
/* ogg_dp_cap_mesh for callers that must not allocate device memory.  No counterpart in the reference (numpy allocates
 * as it pleases); same arguments, same outputs, bit for bit, as ogg_dp_cap_mesh above.
 *
 * Who needs it.  ogg_dp_cap_mesh draws its work tables from a stream-ordered memory pool the library owns.  Three kinds of
 * caller cannot live with that: a model component with a fixed device-memory budget planned at start-up (a pool that
 * grows behind its back is memory it has not accounted for); a thread in a region where entering the CUDA allocator is
 * off limits because the allocator takes a driver-wide lock and its first use may synchronise the whole device; and a
 * process whose device memory belongs to another runtime (an MPI, OpenACC or Fortran host) that forbids foreign
 * allocations.  This entry point calls no cudaMalloc / cudaMallocAsync / cudaFree and no library routine that does; the
 * caller still supplies lams and phis.
 *
 * Where the work tables live.  In one fixed array of OGG_NOALLOC_SCRATCH_DOUBLES doubles (2 MiB) that is part of the
 * library's device image: every GPU on which the library is used holds it until the process exits.  The limit this
 * puts on a request is   32 + 2 * (Ni + 1) + (j1 - j0)  <=  OGG_NOALLOC_SCRATCH_DOUBLES   (e.g. Ni = 86400, the
 * 1/120-degree supergrid, with up to 89310 mesh rows per call; taller caps are done in row ranges, which is what
 * j0, j1 are for).  A larger request returns OGG_ERR_TOO_LARGE and launches nothing.
 *
 * First call on a device.  It looks up the device address of the array (cudaGetSymbolAddress, which may load the
 * library's device code onto that GPU); later calls on that device reuse the address.  Make that first call, on every
 * device you use, during set-up and before you begin capturing any stream.
 *
 * Sharing.  All calls on one device use the same array.  Each call rewrites every slot it reads before reading it, so
 * nothing has to be cleared between calls and one call never sees another's leftovers -- provided they do not run at
 * the same time: two calls whose kernels can overlap on the GPU (different streams without an event between them,
 * or two host threads) overwrite each other's tables and both results are garbage.  Order them on one stream, or with
 * events. */
#define OGG_NOALLOC_SCRATCH_DOUBLES 262144
int ogg_dp_cap_mesh_noalloc(int64_t Ni, int64_t Nj, int64_t j0, int64_t j1, double lon0, double lat0, double lam_pole,
                            double r_pole, double* lams, double* phis, void* stream);
/* great_arc_distance (OGG:522-532) -> out[njl][nil] */
int ogg_great_arc_distance(const double* j0, const double* i0, const double* j1, const double* i1, int64_t njl,
                           int64_t nil, int64_t nx, int64_t ny, double lon0, double lat0, double lon_dp,
                           double r_dp, double* out, void* stream);
/* numerical_hi (direction=0, OGG:535-547) / numerical_hj (direction=1, OGG:550-562) -> out[njl][nil] */
int ogg_numerical_h(int direction, const double* j, int64_t njl, const double* i, int64_t nil, int64_t nx, int64_t ny,
                    double lon0, double lat0, double lon_dp, double r_dp, double eps, int order, double* out,
                    void* stream);
/* displacedPoleCap_metrics_quad (OGG:565-601) for cell rows [j0, j1); output shapes as the bipolar twin */
int ogg_dp_metrics_quad(int order, int64_t nx, int64_t ny, int64_t j0, int64_t j1, double lon0, double lat0,
                        double lon_dp, double r_dp, double Re, int flags, double* dxq, double* dyq, double* daq,
                        void* stream);

/* ---- reductions for metrics_error (OGG:747-756) and the global-area check ------------------------------------------- */
/* out[0] = sum_{k<n} a[k*stride]   (fixed-shape tree: deterministic for given n) */
int ogg_sum_strided(const double* a, int64_t n, int64_t stride, double* out, void* stream);
/* acc[0] += sum_{k<n} splitmix64(bits of a[k]) mod 2^64 (device scalar, caller zeroes it): an order-independent
 * checksum, additive over row-slabs -- array-level parity where the arrays are too large to copy out; stands where the
 * reference prints chksum(x, label) (OGG:19-30, sha256 of the host array) */
int ogg_checksum_accumulate(const double* a, int64_t n, uint64_t* acc, void* stream);

/* ---- file output (write_nc, OGG:773-829) ---------------------------------------------------------------------
 * out[e] = byte-reversed in[e] for n 8-byte elements: the big-endian doubles the netCDF classic format stores.  `in` is
 * a device pointer; `out` is a device pointer OR pinned host memory (cudaHostAlloc: mapped under unified addressing), in
 * which case the kernel's stores are the device->host transfer. */
int ogg_bswap64(const void* in, void* out, int64_t n, void* stream);

/* ---- launch accounting (bench.py: gpu_launches and per-kernel CUDA-event durations) ---------------------------------- */
/* kernels launched by this library since it was loaded */
long long ogg_launch_count(void);
/* on = 1: every launch is bracketed by CUDA events on its own stream; on = 2: only the heavy kernel classes
   (dp_hsamples, bipolar_quad_fused, band_fill), so that the brackets do not perturb a timed step; 0: off */
int ogg_timing_enable(int on);
int ogg_timing_reset(void);
int ogg_timing_kernels(void);          /* number of kernel classes */
const char* ogg_timing_name(int id);   /* class name, e.g. "dp_hsamples" */
/* total milliseconds and launch count of class `id` since the last reset (synchronises on the recorded events) */
int ogg_timing_get(int id, double* ms_total, long long* launches);
/* two classes that run concurrently on different streams: length of the union of the [start, end] intervals of all
   their bracketed launches -- the time during which at least one of them was in flight; *launches = intervals used */
int ogg_timing_pair_span(int id_a, int id_b, double* ms_total, long long* launches);

/* ---- measurement aid: FP64 add/mul issue rate (no FMA), the cap kernels' roofline denominator ------------------------ */
/* runs iters dependent DADD/DMUL pairs x 8 independent chains per thread on every SM; returns via *gops_host the
 * achieved 1e9 instructions/s.  Synchronous. */
int ogg_fp64_issue_rate(int64_t iters, double* ginstr_per_s_host);

/* ---- self-test: the branch-free division / square-root sequences of the cap kernels (MathFast, csrc/ogg_trig.cuh) ------- */
/* n pseudo-random operand sets (seed, mode 0 random / 1 raw bit patterns / 2 near-boundary significands / 3 small integers):
 * counts3_host[0] = results accepted by the operand guards, [1] = accepted results that differ in any bit from the
 * compiler's IEEE-754 div.rn / sqrt.rn (must be 0), [2] = in-range operand sets a guard rejected (must be 0; mode 1
 * has no in-range promise).  mode 4: tan, arctan, arctan2, arcsin, arccos evaluated with those sequences against the
 * same functions with the compiler's operations ([0] accepted, [1] differing).  No reference counterpart: the reference's `/` and numpy.sqrt are IEEE operations, which
 * is what this verifies.  Synchronous. */
int ogg_selftest_divsqrt(int64_t n, uint64_t seed, int mode, uint64_t* counts3_host);

/* ---- audit of the elided operand guards (csrc/ogg_trig.cuh, "ELIDE") -------------------------------------------------- */
/* A library built with -DOGG_GUARD_AUDIT evaluates every operand-range claim the shipped build leaves out as proven and
 * counts, over all cap-kernel launches on the current device since the last reset: counts2_host[0] = units (one projected
 * vertex, one haversine, one quadrature sample) whose guard flag survived, [1] = those of them for which a claim was
 * false (must be 0).  The shipped build returns {0, 0}.  reset != 0 clears the counters after reading.  Synchronous. */
int ogg_guard_audit(uint64_t* counts2_host, int reset);

#ifdef __cplusplus
}
#endif
#endif /* OGG_B200_H_ */

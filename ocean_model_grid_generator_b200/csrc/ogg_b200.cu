// ogg_b200.cu -- sm_100a kernels + C ABI (include/ogg_b200.h) for the supergrid hot path.
//
// Compile with -fmad=false: the reference's arithmetic is numpy element-wise math, one rounding per
// operation, and bit parity with it (and with the CPU oracle) depends on never contracting a*b+c.
// "OGG:n" = /root/reference/ocean_grid_generator.py line n.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <utility>
#include <vector>

#include "../../include/ogg_b200.h"
#include "ogg_trig.cuh"

namespace ogg {

using i64 = long long;

// ------------------------------------------------------------------------------------------------
// host helpers
// ------------------------------------------------------------------------------------------------
#define OGG_CUDA(call)                                        \
  do {                                                        \
    cudaError_t e__ = (call);                                 \
    if (e__ != cudaSuccess) return OGG_ERR_CUDA - (int)e__;   \
  } while (0)
#define OGG_LAUNCH_CHECK() OGG_CUDA(cudaGetLastError())

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }
// Final outputs that no kernel of the build reads again (the cell metrics, the angles) are written with streaming
// (evict-first) stores: they should not push the quadratures' samples, which ARE read again, out of the L2.
#ifdef OGG_PLAIN_OUT_ST
#define OGG_OUT_ST(p, v) (p) = (v)
#else
#define OGG_OUT_ST(p, v) __stcs(&(p), (v))
#endif
static inline unsigned cdiv_u(i64 a, i64 b) { return static_cast<unsigned>((a + b - 1) / b); }

// Per-device state (one process may drive several devices): everything cached about "the device" is keyed
// by the current device id.
constexpr int kMaxDevices = 64;
static inline int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 0;
  return dev;
}
static int num_sms() {
  static int sms[kMaxDevices] = {0};
  const int dev = current_device();
  if (sms[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    sms[dev] = n;
  }
  return sms[dev];
}

// stream-ordered scratch from a PRIVATE pool per device (the host application's default pool is left alone) whose
// release threshold is "never": steady-state steps do not reach the physical allocator.
static std::mutex g_pool_mutex;
static cudaMemPool_t scratch_pool() {
  static cudaMemPool_t pools[kMaxDevices] = {nullptr};
  const int dev = current_device();
  std::lock_guard<std::mutex> lock(g_pool_mutex);
  if (!pools[dev]) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    cudaMemPool_t pool = nullptr;
    if (cudaMemPoolCreate(&pool, &props) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    unsigned long long thr = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    pools[dev] = pool;
  }
  return pools[dev];
}
struct Scratch {
  cudaStream_t st;
  cudaMemPool_t pool;
  void* ptrs[16];
  int n = 0;
  explicit Scratch(cudaStream_t s) : st(s), pool(scratch_pool()) {}
  template <typename T>
  cudaError_t get(T** p, size_t count) {
    void* q = nullptr;
    cudaError_t e = pool ? cudaMallocFromPoolAsync(&q, count * sizeof(T) + 16, pool, st)
                         : cudaMallocAsync(&q, count * sizeof(T) + 16, st);
    if (e != cudaSuccess) return e;
    ptrs[n++] = q;
    *p = static_cast<T*>(q);
    return cudaSuccess;
  }
  ~Scratch() {
    for (int k = 0; k < n; ++k) cudaFreeAsync(ptrs[k], st);
  }
};


// ------------------------------------------------------------------------------------------------
// launch accounting: every kernel launch goes through OGG_LAUNCH -- counted, and (when enabled by
// ogg_timing_enable) bracketed by CUDA events on the launching stream so that bench.py can report each
// kernel's own duration inside the timed region.
// ------------------------------------------------------------------------------------------------
enum KernelId { KID_ELEMENTWISE, KID_BAND_TABLES, KID_BAND_FILL, KID_MIDAS, KID_ANGLE, KID_BP_PROJECTION, KID_BP_TABLES,
                KID_BP_MESH, KID_BP_SAMPLES, KID_BP_QUAD, KID_SAMPLE_INDEX, KID_QUAD, KID_DP_SETUP, KID_DP_TABLES, KID_DP_RAW,
                KID_MONOTONIC, KID_DP_HSAMPLES, KID_REDUCE, KID_COUNT };
static const char* const kKernelNames[KID_COUNT] = {
    "elementwise", "band_tables", "band_fill", "midas", "angle_x", "bipolar_projection", "bipolar_tables", "bipolar_mesh",
    "bipolar_samples", "bipolar_quad_fused", "sample_index", "quad_cells", "dp_setup", "dp_tables", "dp_raw", "monotonic_bounding", "dp_hsamples",
    "reduce"};
// Thread safety: the counters and the list of bracketed launches are shared by all host threads and devices of the
// process and guarded by one mutex; whether the CURRENT launch is bracketed is per thread.
static std::mutex g_timing_mutex;
static long long g_launches = 0;
static int g_timing = 0;         // 0 off, 1 every launch, 2 only the heavy kernel classes (caps + band fill)
struct TimedLaunch { int id, dev; cudaEvent_t e0, e1; bool ended; };   // events belong to the device they were created on
static std::vector<TimedLaunch> g_timed;
static std::vector<cudaEvent_t> g_event_pool[kMaxDevices];
static thread_local long long t_timed_index = -1;   // index into g_timed of this thread's open bracket
static cudaEvent_t take_event(int dev) {   // g_timing_mutex held
  auto& pool = g_event_pool[dev];
  if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}
static inline void launch_begin(int id, cudaStream_t st) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  ++g_launches;
  t_timed_index = -1;
  if (g_timing == 1 || (g_timing == 2 && (id == KID_DP_HSAMPLES || id == KID_BP_QUAD || id == KID_BAND_FILL))) {
    const int dev = current_device();
    TimedLaunch t{id, dev, take_event(dev), take_event(dev), false};
    cudaEventRecord(t.e0, st);
    g_timed.push_back(t);
    t_timed_index = (long long)g_timed.size() - 1;
  }
}
// an entry whose launch_end never ran (an error return in between) stays !ended and is skipped by the readers
static inline void launch_end(cudaStream_t st) {
  if (t_timed_index < 0) return;
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  if (t_timed_index < (long long)g_timed.size()) {
    cudaEventRecord(g_timed[t_timed_index].e1, st);
    g_timed[t_timed_index].ended = true;
  }
  t_timed_index = -1;
}
template <typename K, typename... A>
static inline void launch_cfg(cudaStream_t st, dim3 g, dim3 b, K k, A... a) {
  k<<<g, b, 0, st>>>(a...);
}
#define OGG_UNPAREN(...) __VA_ARGS__
#define OGG_LAUNCH(ID, KERNEL, GRID_BLOCK, STREAM, ...)                              \
  do {                                                                               \
    cudaStream_t st__ = OGG_UNPAREN STREAM;                                          \
    launch_begin(ID, st__);                                                          \
    launch_cfg(st__, OGG_UNPAREN GRID_BLOCK, OGG_UNPAREN KERNEL, __VA_ARGS__);       \
    launch_end(st__);                                                                \
  } while (0)

// ------------------------------------------------------------------------------------------------
// element-wise portable functions
// ------------------------------------------------------------------------------------------------
enum Fn { F_SIN, F_COS, F_TAN, F_ASIN, F_ACOS, F_ATAN, F_EXP, F_LOG, F_SINH, F_COSH, F_MOD360 };

template <int F>
OGG_DI double apply_fn(double x) {
  if (F == F_SIN) return p_sin(x);
  if (F == F_COS) return p_cos(x);
  if (F == F_TAN) return p_tan(x);
  if (F == F_ASIN) return p_asin(x);
  if (F == F_ACOS) return p_acos(x);
  if (F == F_ATAN) return p_atan(x);
  if (F == F_EXP) return p_exp(x);
  if (F == F_LOG) return p_log(x);
  if (F == F_SINH) return p_sinh(x);
  if (F == F_COSH) return p_cosh(x);
  return mod360(x);
}

template <int F>
__global__ void __launch_bounds__(256) k_unary(const double* __restrict__ in, double* __restrict__ out, i64 n) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x)
    out[e] = apply_fn<F>(in[e]);
}
__global__ void __launch_bounds__(256) k_atan2(const double* __restrict__ y, const double* __restrict__ x,
                                               double* __restrict__ out, i64 n) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x)
    out[e] = p_atan2(y[e], x[e]);
}

// the two 1-D Mercator laws, each the reference's chain of element-wise operations in one launch
__global__ void __launch_bounds__(256) k_y_mercator(const double* __restrict__ phi, i64 n, double R, double* __restrict__ out) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x)
    out[e] = R * p_log((1.0 + p_sin(phi[e])) / p_cos(phi[e]));                              // :295
}
__global__ void __launch_bounds__(256) k_phi_mercator(const double* __restrict__ y, i64 n, double y_first, double R,
                                                      double* __restrict__ out) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) {
    const double yv = y ? y[e] : y_first + (double)e;                                       // numpy.arange(y0, y1 + 1)
    out[e] = p_atan(p_sinh(yv / R)) * kC[k180OverPi];                                       // :301
  }
}

static unsigned flat_grid(i64 n, int per_sm = 8) {
  i64 want = (n + 255) / 256;
  i64 cap = (i64)num_sms() * per_sm;
  return static_cast<unsigned>(want < 1 ? 1 : (want < cap ? want : cap));
}

template <int F>
static int launch_unary(const double* in, double* out, i64 n, void* stream) {
  if (n < 0 || (n > 0 && (!in || !out))) return OGG_ERR_ARG;
  if (n == 0) return OGG_OK;
  OGG_LAUNCH(KID_ELEMENTWISE, (k_unary<F>), (flat_grid(n), 256), (S(stream)), in, out, n);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

// ------------------------------------------------------------------------------------------------
// spherical bands: fused tile + MIDAS + angle  (x = lam[i], y = phi[j])
// ------------------------------------------------------------------------------------------------
// column table: dxi[i] (OGG:696), hx[i] = 0.5*(dxi+dxi) (OGG:713), xd[i] = x[ia]-x[ib] (OGG:725-727)
// row table:    cl[j] = cos(lv) (OGG:695,698), cy[j] = cos(y*PI_180) (OGG:725), dyr[j] (OGG:699-702),
//               dsl[j] = sl[j+1]-sl[j] (OGG:711,713)
__global__ void __launch_bounds__(256) k_band_tables(const double* __restrict__ lam, const double* __restrict__ phi,
                                                     i64 nyp, i64 nxp, double Re, double* __restrict__ dxi,
                                                     double* __restrict__ hx, double* __restrict__ xd,
                                                     double* __restrict__ cl, double* __restrict__ cy,
                                                     double* __restrict__ dyr, double* __restrict__ dsl) {
  const i64 t = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (t < nxp) {
    const i64 i = t;
    if (i < nxp - 1) {
      const double d = mdist(lam[i + 1], lam[i]) * kPi180;
      dxi[i] = d;
      hx[i] = 0.5 * (d + d);
    }
    const i64 ia = i == nxp - 1 ? i : i + 1, ib = i == 0 ? i : i - 1;
    xd[i] = lam[ia] - lam[ib];
  }
  if (t < nyp) {
    const i64 j = t;
    const double yj = phi[j];
    const double lv = (0.5 * (yj + yj)) * kPi180;
    const SinCos q = sincos_parts(lv);
    cl[j] = cos_of(q);
    cy[j] = p_cos(yj * kPi180);
    if (j < nyp - 1) {
      const double yu = phi[j + 1];
      const double lu = (0.5 * (yu + yj)) * kPi180;
      const double dx_j = mdist(lam[0], lam[0]) * kPi180;  // same-column longitudes: exactly what OGG:700 sees
      const double dy_j = (yu - yj) * kPi180;
      const double tt = dx_j * p_cos(lu);
      dyr[j] = Re * sqrt(dy_j * dy_j + tt * tt);
      const double lvu = (0.5 * (yu + yu)) * kPi180;
      dsl[j] = p_sin(lvu) - sin_of(q);
    }
  }
}

// One launch fills every band of a step: blockIdx.y runs over the row tiles of all bands, one thread per column,
// kBandRows (a compile-time 8) rows per CTA.  Choosing the rows per CTA on the host so that the grid comes close to a
// whole number of waves was tried with a run-time loop bound and with 4 / 12 / 16 compiled rows and measured slower
// than the fixed 8 rows x 5 CTAs/SM (DESIGN.md "tried and rejected"); it is not what ships.  A thread keeps the column-only factors in registers; the
// row-only factors are uniform loads.  Stores are plain 8-byte stores, a warp writing 256 contiguous bytes per array and
// row: 16-byte pair stores (which need a per-row alignment shift because the row length nxp is odd) and column strips
// with a one-wave persistent grid were both measured slower (DESIGN.md "tried and rejected").
struct BandDesc {
  const double *lam, *phi, *dxi, *hx, *xd, *cl, *cy, *dyr, *dsl;
  double *x, *y, *dx, *dy, *area, *angle;
  i64 nyp, nxp, nyp_v;
  i64 t_begin;                    // first row tile (blockIdx.y) of this band
  int metrics;
};
constexpr int kMaxBands = 4;
struct BandBatch { BandDesc b[kMaxBands]; int n; double Re, Re2; };

OGG_DI double band_angle(double xdv, double cyj, double dyc, double ang_pos, double ang_neg) {
  // angle_x on a band (:725-728): y[ia]-y[ib] = phi-phi = +0, so p_atan2 takes its ay == 0 branch: +0 for a
  // non-negative second argument, pi - (0 - pi_lo) for a negative one -- two constants.  What is left is a NaN first
  // argument (phi not finite) or a NaN second one: the spec's arctan2 returns NaN for both (its quotient is NaN and no
  // range test of arctan takes a NaN), which is what the sum below is -- without the general code's 90 registers.
  const double arg = xdv * cyj;
  if (dyc == 0.0 && arg >= 0.0) return ang_pos;
  if (dyc == 0.0 && arg < 0.0) return ang_neg;
  return dyc + arg;
}
OGG_DI double band_dx(double dxiv, double clj, double dyc, double Re) {
  // dx (:698): dy_i = +0 and sqrt(0 + tt*tt) = |tt|: sqrt(fl(t*t)) == |t| in binary floating point, round-to-nearest,
  // when t*t neither under- nor overflows (Boldo 2015); checked on 1e7 values in
  // tests/test_oracle_cpu.py::test_sqrt_of_square_identity
  const double dy_i = dyc * kPi180;
  const double tt = dxiv * clj;
  const double t2 = tt * tt;
  const double root = (dy_i == 0.0 && t2 >= 0x1p-1000 && t2 <= 0x1p1000) ? fabs(tt) : sqrt(dy_i * dy_i + t2);
  return Re * root;
}
#ifndef OGG_BAND_THREADS
#define OGG_BAND_THREADS 256
#endif
constexpr int kBandThreads = OGG_BAND_THREADS;
#ifndef OGG_BAND_ROWS
#define OGG_BAND_ROWS 8
#endif
#ifndef OGG_BAND_MINB
#define OGG_BAND_MINB 5
#endif
constexpr int kBandRows = OGG_BAND_ROWS;
// kBandRows rows of one band from row j0 on, one column per thread.  D is a statically indexed member of the kernel
// parameter, so its pointers and sizes are constant-bank operands (a dynamically indexed descriptor costs ~50 registers).
OGG_DI void band_rows(const BandDesc& D, i64 j0, double Re, double Re2) {
  const i64 nxp = D.nxp, nx = nxp - 1, nyp = D.nyp;
  const i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (i >= nxp) return;
  const bool cell = i < nx;
  const double lamv = D.lam[i], xdv = D.xd[i];
  const double dxiv = cell ? D.dxi[i] : 0.0, hxv = cell ? D.hx[i] : 0.0;
  const double ang_pos = 0.0 / kPi180, ang_neg = (kPi - (0.0 - kC[kPiLo])) / kPi180;
  // dy (:699-700) has x[j+1,i]-x[j,i] = lam-lam under its root: +0 for a finite longitude (then dy is the row table's
  // value), NaN for a NaN or infinite one (then dy is NaN)
  const double dxc = lamv - lamv;
#pragma unroll
  for (int r = 0; r < kBandRows; ++r) {
    const i64 j = j0 + r;
    if (j >= nyp) break;
    const double phij = D.phi[j];
    const double dyc = phij - phij;  // y[ia]-y[ib] on a band: +0 unless phi is not finite
    // streaming (evict-first) stores: nothing here is read again by this kernel, and the bytes should not push the cap
    // quadratures' samples out of L2.  Measured on the Mercator band alone: 4.96 -> 5.17 TB/s at r=4, 6.61 -> 6.91 TB/s
    // at r=12 (write-through / cache-global stores: 4.93 / 6.53).  -DOGG_BAND_PLAIN_ST restores plain stores.
#ifdef OGG_BAND_PLAIN_ST
#define OGG_BST(p, v) (p) = (v)
#else
#define OGG_BST(p, v) __stcs(&(p), (v))
#endif
    if (j < D.nyp_v) {
      if (D.x) OGG_BST(D.x[j * nxp + i], lamv);
      if (D.y) OGG_BST(D.y[j * nxp + i], phij);
      if (D.angle) OGG_BST(D.angle[j * nxp + i], band_angle(xdv, D.cy[j], dyc, ang_pos, ang_neg));
      if (D.metrics && cell && D.dx) OGG_BST(D.dx[j * nx + i], band_dx(dxiv, D.cl[j], dyc, Re));
    }
    if (D.metrics && j < nyp - 1) {
      if (D.dy) OGG_BST(D.dy[j * nxp + i], dxc == 0.0 ? D.dyr[j] : dxc);
      if (cell && D.area) OGG_BST(D.area[j * nx + i], Re2 * (hxv * D.dsl[j]));
    }
  }
}
__global__ void __launch_bounds__(kBandThreads, OGG_BAND_MINB) k_band_fill(const __grid_constant__ BandBatch B) {
  const i64 t = blockIdx.y;
  if (B.n > 3 && t >= B.b[3].t_begin) band_rows(B.b[3], (t - B.b[3].t_begin) * kBandRows, B.Re, B.Re2);
  else if (B.n > 2 && t >= B.b[2].t_begin) band_rows(B.b[2], (t - B.b[2].t_begin) * kBandRows, B.Re, B.Re2);
  else if (B.n > 1 && t >= B.b[1].t_begin) band_rows(B.b[1], (t - B.b[1].t_begin) * kBandRows, B.Re, B.Re2);
  else band_rows(B.b[0], t * kBandRows, B.Re, B.Re2);
}

// ------------------------------------------------------------------------------------------------
// general MIDAS metrics on a supplied mesh (OGG:687-716)
// ------------------------------------------------------------------------------------------------
// One thread per column, a strip of kMidasRows mesh rows per block row; the j/j+1 stencil is carried in
// registers down the strip (row j of one step = row j-1 of the next), the i+1 neighbour comes from the next
// lane by warp shuffle (lane 31 reads it from memory).
constexpr int kMidasRows = 32;
__global__ void __launch_bounds__(128) k_midas(const double* __restrict__ x, const double* __restrict__ y, i64 nyp,
                                               i64 nxp, double Re, double Re2, int areafix,
                                               const double* __restrict__ halo_x, const double* __restrict__ halo_y,
                                               double* __restrict__ dx, double* __restrict__ dy,
                                               double* __restrict__ area) {
  const i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  const i64 nx = nxp - 1;
  const i64 nrows = nyp + (halo_x ? 1 : 0);  // mesh rows available incl. halo
  const i64 jb = (i64)blockIdx.y * kMidasRows;
  const bool in = i < nxp, cell = i < nx;
  const unsigned lane = threadIdx.x & 31u;
  double xp = 0, yp = 0, dxip = 0, slp = 0;  // previous row
  for (int r = 0; r <= kMidasRows; ++r) {
    const i64 j = jb + r;
    if (j >= nrows) break;
    const double* xr = j < nyp ? x + j * nxp : halo_x;
    const double* yr = j < nyp ? y + j * nxp : halo_y;
    const double xc = in ? xr[i] : 0.0, yc = in ? yr[i] : 0.0;
    double xn = __shfl_down_sync(0xffffffffu, xc, 1);
    double yn = __shfl_down_sync(0xffffffffu, yc, 1);
    if (lane == 31u && cell) { xn = xr[i + 1]; yn = yr[i + 1]; }
    double dxi = 0, sl = 0;
    if (cell) {
      const double lv = (0.5 * (yn + yc)) * kPi180;       // :695
      dxi = mdist(xn, xc) * kPi180;                       // :696
      const double dyi = (yn - yc) * kPi180;              // :697
      const SinCos q = sincos_parts(lv);
      const double tt = dxi * cos_of(q);
      if (r < kMidasRows && j < nyp && dx) dx[j * nx + i] = Re * sqrt(dyi * dyi + tt * tt);  // :698
      sl = sin_of(q);                                     // :711
    }
    if (r > 0 && in) {
      const i64 jm = j - 1;
      if (dy) {
        const double lu = (0.5 * (yc + yp)) * kPi180;     // :699
        const double dxj = mdist(xc, xp) * kPi180;        // :700
        const double dyj = (yc - yp) * kPi180;            // :701
        const double tt = dxj * p_cos(lu);
        dy[jm * nxp + i] = Re * sqrt(dyj * dyj + tt * tt);  // :702
      }
      if (cell && area && areafix) area[jm * nx + i] = Re2 * ((0.5 * (dxi + dxip)) * (sl - slp));  // :713
    }
    xp = xc; yp = yc; dxip = dxi; slp = sl;
  }
}
// latlon_areafix=False branch (OGG:715); dx has nyp_a+1 rows available (halo-aware callers pass the extra row)
__global__ void __launch_bounds__(256) k_area_nofix(const double* __restrict__ dx, const double* __restrict__ dxtop,
                                                    const double* __restrict__ dy, i64 ny, i64 nx,
                                                    double* __restrict__ area) {
  const i64 n = ny * nx;
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) {
    const i64 j = e / nx, i = e - j * nx;
    const double up = (dxtop && j == ny - 1) ? dxtop[i] : dx[(j + 1) * nx + i];
    area[e] = 0.25 * ((up + dx[j * nx + i]) * (dy[j * (nx + 1) + i + 1] + dy[j * (nx + 1) + i]));
  }
}
// dx of a single (halo) row: used only by the areafix=False + halo combination
__global__ void __launch_bounds__(256) k_dx_row(const double* __restrict__ xr, const double* __restrict__ yr, i64 nx,
                                                double Re, double* __restrict__ out) {
  const i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (i >= nx) return;
  const double lv = (0.5 * (yr[i + 1] + yr[i])) * kPi180;
  const double dxi = mdist(xr[i + 1], xr[i]) * kPi180;
  const double dyi = (yr[i + 1] - yr[i]) * kPi180;
  const double tt = dxi * p_cos(lv);
  out[i] = Re * sqrt(dyi * dyi + tt * tt);
}

// angle_x (OGG:719-729)
__device__ __noinline__ double angle_ieee(double dy, double dx, double yc) {
  return p_atan2(dy, dx * p_cos(yc * kPi180)) / kPi180;
}
// pi180 != NULL: the two divisions of arctan2 and the one by pi/180 under MathFast, the compiler's code as redo (a zero
// difference -- a grid line along a parallel or a meridian -- fails the unit)
__global__ void __launch_bounds__(256) k_angle(const double* __restrict__ x, const double* __restrict__ y, i64 nyp,
                                               i64 nxp, double* __restrict__ angle, const double* __restrict__ pi180) {
  const i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (i >= nxp) return;
  const i64 ia = i == nxp - 1 ? i : i + 1, ib = i == 0 ? i : i - 1;
  if (pi180) {
    MathFast mf(pi180);
    for (i64 j = blockIdx.y; j < nyp; j += gridDim.y) {
      const double* xr = x + j * nxp;
      const double* yr = y + j * nxp;
      const double dy = yr[ia] - yr[ib], dx = xr[ia] - xr[ib], yc = yr[i];
      double a;
      if (dy == 0.0 || dx == 0.0) {
        // a grid line along a parallel or a meridian (every vertex of a regular or Mercator mesh): the fast path's
        // guard is known to fail, so do not evaluate it first
        a = angle_ieee(dy, dx, yc);
      } else {
        mf.reset();
        a = mf.div_pi180(p_atan2(mf, dy, dx * p_cos(mf, yc * kPi180)));
        if (!mf.ok()) a = angle_ieee(dy, dx, yc);
      }
      OGG_OUT_ST(angle[j * nxp + i], a);
    }
    return;
  }
  for (i64 j = blockIdx.y; j < nyp; j += gridDim.y) {
    const double* xr = x + j * nxp;
    const double* yr = y + j * nxp;
    const double a = p_atan2(yr[ia] - yr[ib], (xr[ia] - xr[ib]) * p_cos(yr[i] * kPi180));
    angle[j * nxp + i] = a / kPi180;
  }
}

// ------------------------------------------------------------------------------------------------
// bipolar cap (OGG:33-133)
// ------------------------------------------------------------------------------------------------
struct BpRow { double sphig, beta2_inv, N_inv; };   // depends on the latitude index only
struct BpCol { double sinla, alpha2, dl; };         // depends on the longitude index only

OGG_DI BpRow bp_row(double phig_in, double rp) {
  const double phig = 90 - ((2 * p_atan(p_tan((0.5 * (90 - phig_in)) * kPi180) / rp)) / kPi180);  // :41
  const SinCos q = sincos_parts(phig * kPi180);
  BpRow o;
  o.sphig = sin_of(q);                                   // :44
  const double tb = tan_of(q);
  o.beta2_inv = tb * tb;                                 // :46
  const double chig = (90 - phig) * kPi180;              // :75
  const double tg = p_tan(chig / 2);
  const double qg = rp * tg;
  const double rden2 = 1.0 / (1 + qg * qg);              // :76
  const double N = (rp * (1 + tg * tg)) * rden2;         // :77
  o.N_inv = 1 / N;                                       // :78
  return o;
}
OGG_DI BpCol bp_col(double lamg, double lon_bp) {
  const double tmp = mdist(lamg, lon_bp) * kPi180;       // :42
  const SinCos q = sincos_parts(tmp);
  BpCol o;
  o.sinla = sin_of(q);                                   // :43
  const double ca = cos_of(q);
  o.alpha2 = ca * ca;                                    // :45
  o.dl = lamg - lon_bp;
  return o;
}
// M: division / square-root policy (ogg_trig.cuh); with MathFast the caller must redo the sample if !m.ok(), and its
// base flag promises rp in [2**-63, 2**65) (bp_rp_ok).  The "proven" claims, in order:
//  (1) alpha2 = ca*ca and beta2_inv = tb*tb are >= 0 or NaN, so 1 + alpha2*beta2_inv is >= 1, +inf or NaN: "not huge"
//      is the whole guard.  (A NaN beta2_inv makes `huge` differ between the policies -- inside a failed unit.)
//  (2) tan(chic/2) = s/c with s, c in "mid" and |s|, |c| <= 2: |tc| in [2**-130, 2**128], so |q| = |rp*tc| is in
//      [2**-193, 2**193]: arctan's argument is not huge, 2*arctan(q) is in [2**-193, pi] (moderate), and 1 + q*q is in
//      [1, 2**387] (moderate).
template <bool FULL, class M, bool METRICS = true>
OGG_DI void bp_sample(M& m, const BpRow& R, const BpCol& C, double lon_bp, double rp,
                      double& lams, double& phis, double& h_i_inv, double& h_j_inv) {
  const double alpha2 = C.alpha2, beta2_inv = R.beta2_inv;
  const double den1 = 1.0 + alpha2 * beta2_inv;
  m.require_not_huge(den1);
  m.proven(rng_moderate(den1));                           // (1)
  const double rden = m.div_ok(1.0, den1);               // :47
  // |beta2_inv| > 1e30 happens on the exact pole row only: under MathFast it fails the unit (the redo applies :52, :86,
  // :94) instead of costing every sample a compare and five selects
  const bool huge = M::kFast ? false : M::abs_gt_huge(beta2_inv);
  if (M::kFast) m.require(fabsf(hi_as_float(beta2_inv)) < 0x1p13f);   // high word of 2**97 < 1e30; a NaN fails too
  if (FULL) {
    double B = C.sinla * m.sqrt_(rden);                  // :50
    if (huge) B = 0.0;                                   // :52
    double lamc = m.div_pi180(p_asin(m, B));             // :53
    const double dl = C.dl;
    if (dl > 90 && dl <= 180) lamc = 180 - lamc;         // :58
    if (dl > 180 && dl <= 270) lamc = 180 + lamc;        // :59
    if (dl > 270) lamc = 360 - lamc;                     // :60
    if (dl == 90) lamc = 90;                             // :62
    if (dl == 270) lamc = 270;                           // :63
    lams = lamc + lon_bp;                                // :64
  }
  const double A = C.sinla * R.sphig;                    // :68
  const double chic = p_acos(m, A);                      // :69
  const double tc = p_tan_q0<true>(m, chic / 2);        // chic in [0, pi]: quadrant 0 unless A < 0 (never on a cap)
  const double q = rp * tc;
  const double at2 = 2 * p_atan<false>(m, q);            // (2)
  m.proven(rng_moderate(at2));
  phis = 90 - m.div_pi180_seq(at2);                      // :70
  if (!METRICS) { h_i_inv = 0.0; h_j_inv = 0.0; return; } // coordinates only (the supergrid takes the metrics from the quadrature)
  const double den2 = 1 + q * q;
  m.proven(rng_moderate(den2));                          // (2)
  const double rden2 = m.div_ok(1.0, den2);              // :72
  const double M_inv = (rp * (1 + tc * tc)) * rden2;     // :73
  const double cp = p_cos_q1_unsigned(m, phis * kPi180);  // +-cos(phis): squared on the next line
  const double cos2phis = cp * cp;                       // :79
  const double MM = M_inv * M_inv, r2 = rden * rden;
  double hj = ((((cos2phis * alpha2) * (1 - alpha2)) * beta2_inv) * (1 + beta2_inv)) * r2 +
              (MM * (1 - alpha2)) * rden;                // :81-84
  if (huge) hj = MM;                                     // :86
  h_j_inv = m.sqrt_(hj) * R.N_inv;                       // :87
  double hi = (cos2phis * (1 + beta2_inv)) * r2 + ((MM * alpha2) * beta2_inv) * rden;  // :89-92
  if (huge) hi = MM;                                     // :94
  h_i_inv = m.sqrt_(hi);                                 // :95
}
OGG_DI bool bp_rp_ok(double rp) { const float f = hi_as_float(rp); return (f >= 0x1p-7f) && (f < 0x1p9f); }
template <bool FULL>
OGG_DI void bp_sample(const BpRow& R, const BpCol& C, double lon_bp, double rp, double& lams, double& phis,
                      double& h_i_inv, double& h_j_inv) {
  MathIeee m;
  bp_sample<FULL>(m, R, C, lon_bp, rp, lams, phis, h_i_inv, h_j_inv);
}

__global__ void __launch_bounds__(256) k_bp_projection(const double* __restrict__ lamg, const double* __restrict__ phig,
                                                       i64 n, double lon_bp, double rp, int metrics_only,
                                                       double* __restrict__ lams, double* __restrict__ phis,
                                                       double* __restrict__ hi, double* __restrict__ hj) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) {
    const BpRow R = bp_row(phig[e], rp);
    const BpCol C = bp_col(lamg[e], lon_bp);
    double l = 0, p, a, b;
    if (metrics_only) bp_sample<false>(R, C, lon_bp, rp, l, p, a, b);
    else bp_sample<true>(R, C, lon_bp, rp, l, p, a, b);
    if (!metrics_only) { lams[e] = l; phis[e] = p; }
    if (hi) hi[e] = a;
    if (hj) hj[e] = b;
  }
}

// index -> coordinate tables.  iv/jv NULL: integer indices i0+t / j0+t (generate_bipolar_cap_mesh :113,:115);
// otherwise fractional indices (bipolar_cap_ij_array :126-127).
__global__ void __launch_bounds__(256) k_bp_tables(const double* __restrict__ iv, i64 i0, i64 ncol,
                                                   const double* __restrict__ jv, i64 j0, i64 nrow, i64 Ni, i64 Nj,
                                                   double lat0_bp, double lon_bp, double rp, BpCol* __restrict__ ct,
                                                   BpRow* __restrict__ rt) {
  const i64 t = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (t < ncol) {
    const double idx = iv ? iv[t] : (double)(i0 + t);
    ct[t] = bp_col(lon_bp + (idx * 360.0) / (double)Ni, lon_bp);
  }
  if (t < nrow) {
    const double idx = jv ? jv[t] : (double)(j0 + t);
    rt[t] = bp_row(lat0_bp + (idx * (90 - lat0_bp)) / (double)Nj, rp);
  }
}

struct HiHj { double a, b; };
__device__ __noinline__ HiHj bp_vertex_ieee(BpRow R, BpCol C, double lon_bp, double rp) {   // (lams, phis), compiler's / and sqrt
  double l, p, a, b;
  MathIeee m;
  bp_sample<true, MathIeee, false>(m, R, C, lon_bp, rp, l, p, a, b);
  return HiHj{l, p};
}
// MODE 0: mesh (:118-120): lams, phis + optionally scaled h's with the mesh's trimming
// MODE 2: lams, phis only (pi180: the cached divisor pair of pi180_pair())
// MODE 1: metrics-only samples (:130-132): scaled h_i_inv, h_j_inv on the full lattice
template <int MODE>
__global__ void __launch_bounds__(256) k_bp_lattice(const BpCol* __restrict__ ct, const BpRow* __restrict__ rt,
                                                    i64 nrow, i64 ncol, i64 Ni, i64 Nj, i64 jrow0, double lat0_bp,
                                                    double lon_bp, double rp, double* __restrict__ o0,
                                                    double* __restrict__ o1, double* __restrict__ o2,
                                                    double* __restrict__ o3, const double* __restrict__ pi180) {
  const i64 c = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  const BpCol C = ct[c];
  for (i64 r = blockIdx.y; r < nrow; r += gridDim.y) {
    const BpRow R = rt[r];
    double l = 0, p, a, b;
    if (MODE == 0) {
      bp_sample<true>(R, C, lon_bp, rp, l, p, a, b);
      o0[r * ncol + c] = l;
      o1[r * ncol + c] = p;
      if (o2 && c < ncol - 1) o2[r * (ncol - 1) + c] = ((a * 2) * kPi) / (double)Ni;                        // :119
      if (o3 && jrow0 + r < Nj) o3[r * ncol + c] = ((b * kPi180) * (90 - lat0_bp)) / (double)Nj;            // :120
    } else if (MODE == 2) {
      // coordinates only, under MathFast (the supergrid's cap mesh: its metrics come from the quadrature)
      MathFast mf(pi180);
      mf.base = bp_rp_ok(rp);
      mf.reset();
      bp_sample<true, MathFast, false>(mf, R, C, lon_bp, rp, l, p, a, b);
      if (!mf.ok()) { const HiHj o = bp_vertex_ieee(R, C, lon_bp, rp); l = o.a; p = o.b; }
      o0[r * ncol + c] = l;
      o1[r * ncol + c] = p;
    } else {
      bp_sample<false>(R, C, lon_bp, rp, l, p, a, b);
      o0[r * ncol + c] = ((a * 2) * kPi) / (double)Ni;                                                      // :131
      o1[r * ncol + c] = ((b * (90 - lat0_bp)) * kPi180) / (double)Nj;                                      // :132
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Gauss-Lobatto quadrature (OGG:191-255) on a lattice of UNIQUE samples
// ------------------------------------------------------------------------------------------------
struct QuadNodes { double a[5], b[5]; };

static int quad_nodes_host(int n, QuadNodes* q) {  // :191-204 (plain IEEE ops, no trig)
  if (n == 2) { q->a[0] = 0.0; q->a[1] = 1.0; q->b[0] = 1.0; q->b[1] = 0.0; return 0; }
  if (n == 3) { q->a[0] = 0.0; q->a[1] = 0.5; q->a[2] = 1.0; q->b[0] = 1.0; q->b[1] = 0.5; q->b[2] = 0.0; return 0; }
  if (n == 4) {
    const double r5 = 0.5 / std::sqrt(5.0);
    q->a[0] = 0.0; q->a[1] = 0.5 - r5; q->a[2] = 0.5 + r5; q->a[3] = 1.0;
    q->b[0] = 1.0; q->b[1] = 0.5 + r5; q->b[2] = 0.5 - r5; q->b[3] = 0.0;
    return 0;
  }
  if (n == 5) {
    const double r37 = 0.5 * std::sqrt(3.0 / 7.0);
    q->a[0] = 0.0; q->a[1] = 0.5 - r37; q->a[2] = 0.5; q->a[3] = 0.5 + r37; q->a[4] = 1.0;
    q->b[0] = 1.0; q->b[1] = 0.5 + r37; q->b[2] = 0.5; q->b[3] = 0.5 - r37; q->b[4] = 0.0;
    return 0;
  }
  return -1;
}

// Unique sample indices for cells [c0, c1): position u = (n-1)*(c-c0)+k holds b[k]*c + a[k]*(c+1)  (:145,:155),
// k < n-1; the closing entry u = (n-1)*(c1-c0) is cell c1-1's last node b[n-1]*(c1-1) + a[n-1]*c1.
// fudge: if that closing value equals `total` it becomes total-0.001 (:146-147) and one more entry
// holding cell `total`'s first node (the exact pole row used by dxq[ny,:]) is appended.
__global__ void __launch_bounds__(256) k_sample_index(QuadNodes q, int n, i64 c0, i64 c1, i64 total, int fudge,
                                                      double* __restrict__ out) {
  const i64 nu = (n - 1) * (c1 - c0) + 1;
  const i64 u = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (u < nu) {
    i64 c = c0 + u / (n - 1);
    int k = (int)(u % (n - 1));
    if (u == nu - 1) { c = c1 - 1; k = n - 1; }
    double v = q.b[k] * (double)c + q.a[k] * (double)(c + 1);
    if (fudge && u == nu - 1 && v == (double)total) v = (double)total - 0.001;
    out[u] = v;
  }
  if (fudge && c1 == total && u == nu) out[u] = q.b[0] * (double)total + q.a[0] * (double)(total + 1);
}

template <int N>
OGG_DI double quad_avg_1d(const double* y) {  // :207-222
  if constexpr (N == 2) return (1.0 / 2.0) * (y[0] + y[1]);
  else if constexpr (N == 3) return (1.0 / 6.0) * (4.0 * y[1] + (y[0] + y[2]));
  else if constexpr (N == 4) return (1.0 / 12.0) * (5.0 * (y[1] + y[2]) + (y[0] + y[3]));
  else return (1.0 / 180.0) * ((64.0 * y[2] + (49.0 * (y[1] + y[3]))) + 9.0 * (y[0] + y[4]));
}
template <int N>
OGG_DI double quad_avg_2d(const double (&y)[N][N]) {  // :225-255
  if constexpr (N == 2) {
    const double d = 1.0 / 2.0;
    return (d * d) * (((y[0][0] + y[0][1]) + y[1][0]) + y[1][1]);
  } else if constexpr (N == 3) {
    const double d = 1.0 / 6.0;
    const double inner = (((y[0][1] + y[1][0]) + y[1][2]) + y[2][1]) + 4.0 * y[1][1];
    const double outer = (((y[0][0] + y[0][2]) + y[2][0]) + y[2][2]) + 4.0 * inner;
    return (d * d) * outer;
  } else {
    constexpr double w4[5] = {1.0, 5.0, 5.0, 1.0, 0.0};
    constexpr double w5[5] = {9.0, 49.0, 64.0, 49.0, 9.0};
    const double d = N == 4 ? 1.0 / 12.0 : 1.0 / 180.0;
    double ysum = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) ysum = ysum + ((N == 4 ? w4[i] : w5[i]) * (N == 4 ? w4[j] : w5[j])) * y[j][i];
    return (d * d) * ysum;
  }
}

// Cell sums (OGG:180-187 / :591-599) for cell rows [0, ncr) of a chunk whose unique sample rows start at the
// chunk's first cell.  sx, sy: [nur][nuc] unique samples (dx, dy).  top_row >= 0: unique row feeding the extra
// dxq row (pole row) written at dxq row ncr.
template <int N>
__global__ void __launch_bounds__(128) k_quad_cells(const double* __restrict__ sx, const double* __restrict__ sy,
                                                    i64 nuc, i64 ncr, i64 nx, i64 top_row, double Re,
                                                    double* __restrict__ dxq, double* __restrict__ dyq,
                                                    double* __restrict__ daq) {
  const i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (i > nx) return;
  for (i64 j = blockIdx.y; j <= ncr; j += gridDim.y) {
    const i64 c0 = (N - 1) * i;
    if (j == ncr) {
      if (top_row >= 0 && i < nx) OGG_OUT_ST(dxq[j * nx + i], quad_avg_1d<N>(sx + top_row * nuc + c0) * Re);  // :183,:186
      continue;
    }
    const i64 r0 = (N - 1) * j;
    if (i < nx) {
      double prod[N][N];
#pragma unroll
      for (int jj = 0; jj < N; ++jj)
#pragma unroll
        for (int ii = 0; ii < N; ++ii) {
          const i64 o = (r0 + jj) * nuc + c0 + ii;
          prod[jj][ii] = sx[o] * sy[o];                                                     // :178
        }
      OGG_OUT_ST(daq[j * nx + i], (quad_avg_2d<N>(prod) * Re) * Re);                        // :182,:185
      OGG_OUT_ST(dxq[j * nx + i], quad_avg_1d<N>(sx + r0 * nuc + c0) * Re);                 // :183,:186
    }
    double col[N];
#pragma unroll
    for (int jj = 0; jj < N; ++jj) col[jj] = sy[(r0 + jj) * nuc + c0];
    OGG_OUT_ST(dyq[j * (nx + 1) + i], quad_avg_1d<N>(col) * Re);                            // :184,:187
  }
}

// --- fused bipolar quadrature (OGG:136-188): samples never leave the SM ------------------------------------------
// Sample index of unique position u for cells [c0, c1) (same law as k_sample_index).
OGG_DI double sample_index_at(const QuadNodes& q, int n, i64 c0, i64 c1, i64 total, int fudge, i64 u) {
  const i64 nu = (n - 1) * (c1 - c0) + 1;
  if (u >= nu) return q.b[0] * (double)total + q.a[0] * (double)(total + 1);  // the exact pole row (:183 at j=ny)
  i64 c = c0 + u / (n - 1);
  int k = (int)(u % (n - 1));
  if (u == nu - 1) { c = c1 - 1; k = n - 1; }
  double v = q.b[k] * (double)c + q.a[k] * (double)(c + 1);                    // :145,:155
  if (fudge && u == nu - 1 && v == (double)total) v = (double)total - 0.001;   // :146-147
  return v;
}
// column table over all unique sample columns, row table over the unique sample rows of cell rows [j0, j1)
// (+ the exact pole row when j1 == ny): one launch replaces k_sample_index x2 + k_bp_tables x2.
__global__ void __launch_bounds__(256) k_bp_quad_tables(QuadNodes q, int n, i64 nx, i64 ny, i64 j0, i64 j1, i64 nuc,
                                                        i64 nur, double lat0_bp, double lon_bp, double rp,
                                                        BpCol* __restrict__ ct, BpRow* __restrict__ rt,
                                                        double* __restrict__ divs) {
  const i64 t = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (t < 3) MathFast::store_divisor(divs + 2 * t, t == 0 ? kPi180 : (t == 1 ? (double)nx : (double)ny));   // see MathFast::Divisor
  if (t < nuc) {
    const double idx = sample_index_at(q, n, 0, nx, nx, 0, t);
    ct[t] = bp_col(lon_bp + (idx * 360.0) / (double)nx, lon_bp);                 // :126
  }
  if (t < nur) {
    const double idx = sample_index_at(q, n, j0, j1, ny, 1, t);
    rt[t] = bp_row(lat0_bp + (idx * (90 - lat0_bp)) / (double)ny, rp);           // :127
  }
}

// One block = a tile of TR x TC cells.  Phase 1: the tile's unique samples ((N-1)TR+1) x ((N-1)TC+1) of the scaled
// h_i_inv, h_j_inv (:130-132) go to shared memory, columns stored residue-major (column c at
// (c mod (N-1))*(TC+1) + c/(N-1)) so that phase 2 -- one thread per cell, sums in the reference's order
// (:180-187) -- reads consecutive words across a warp.  Samples on tile edges are recomputed by the neighbour
// (5.0 % more samples at TR=6, TC=32) instead of travelling through HBM (16 B written + 16 B read per sample).
// the same scaled sample with the compiler's division and square root: taken when an operand leaves MathFast's range
#ifndef OGG_BP_MINB
#define OGG_BP_MINB 4
#endif
__device__ __noinline__ HiHj bp_qsample_ieee(BpRow R, BpCol C, double lon_bp, double rp, double lat0_bp, double dnx,
                                             double dny) {
  double l, p, a, b;
  bp_sample<false>(R, C, lon_bp, rp, l, p, a, b);
  return HiHj{((a * 2) * kPi) / dnx,                                              // :131
              ((b * (90 - lat0_bp)) * kPi180) / dny};                             // :132
}
#ifndef OGG_BP_THREADS
#define OGG_BP_THREADS 256
#endif
constexpr int kBpThreads = OGG_BP_THREADS;
template <int N, int TR, int TC>
__global__ void __launch_bounds__(kBpThreads, OGG_BP_MINB) k_bp_quad_fused(const BpCol* __restrict__ ct, const BpRow* __restrict__ rt,
                                                       const double* __restrict__ divs, i64 nuc,
                                                       i64 ncr, i64 nx, i64 ny, double lat0_bp, double lon_bp, double rp,
                                                       double Re, double* __restrict__ dxq, double* __restrict__ dyq,
                                                       double* __restrict__ daq) {
  constexpr int SR = (N - 1) * TR + 1, SC = (N - 1) * TC + 1, SCP = (N - 1) * (TC + 1);
  extern __shared__ double smem[];
  double* sx = smem;
  double* sy = smem + SR * SCP;
  // sample row / column numbers fit 31 bits (the host checks): 32-bit arithmetic in the per-sample bounds and addresses
  const int gr0 = (int)blockIdx.y * TR * (N - 1), gc0 = (int)blockIdx.x * TC * (N - 1);
  const int nreg = (N - 1) * (int)ncr + 1;  // regular unique rows of this slab
  // cached divisors pi/180, nx, ny (all moderate: nx, ny >= 1) with their refined reciprocals, stored by k_bp_quad_tables
  MathFast mf(divs);
  mf.base = bp_rp_ok(rp);
  const MathFast::Divisor dnx = MathFast::load_divisor(divs + 2), dny = MathFast::load_divisor(divs + 4);
  // one sample: scaled h_i_inv, h_j_inv of unique row r / column c of this tile -> shared memory
  auto sample = [&](int r, int cpos, const BpCol& C, bool col_ok) {   // cpos: the column's part of the smem position
    const int gr = gr0 + r;
    double a = 0.0, b = 0.0;
    if (col_ok && gr < nreg) {
      const BpRow R = rt[gr];
      double l, p;
      mf.reset();
      bp_sample<false>(mf, R, C, lon_bp, rp, l, p, a, b);
      // a = sqrt_(hi) passed the square root's guard: it is in [2**-224, 2**225), and so is 2*pi times it
      const double a2pi = (a * 2) * kPi;
      mf.proven(rng_moderate(a2pi));
      a = mf.div_by_seq(a2pi, dnx);                                              // :131
      b = mf.div_by((b * (90 - lat0_bp)) * kPi180, dny);                         // :132
      if (!mf.ok()) {   // a zero / non-finite / extreme operand somewhere (the pole row, the lon_bp meridian)
        const HiHj o = bp_qsample_ieee(R, C, lon_bp, rp, lat0_bp, (double)nx, (double)ny);
        a = o.a; b = o.b;
      }
    }
    const int pos = r * SCP + cpos;
    sx[pos] = a;
    sy[pos] = b;
  };
  auto col_pos = [](int c) { return (c % (N - 1)) * (TC + 1) + c / (N - 1); };
  // a thread keeps ONE sample column (its column-table entry stays in registers, a warp covers 32 consecutive
  // columns of one row, whose row-table entry is a broadcast load) and walks down the rows, G row-interleaved groups
  // of SC-1 threads; the tile's last column goes to the first SR threads of the last group, which has the fewest rows.
  constexpr int CM = SC - 1, G = kBpThreads / CM;
  static_assert(G >= 1 && SR <= CM, "a row of sample columns per thread group");
  const int grp = threadIdx.x / CM, cm = threadIdx.x - grp * CM;
  if (grp < G) {
    const bool col_ok = gc0 + cm < (int)nuc;
    BpCol C{0.0, 0.0, 0.0};
    if (col_ok) C = ct[gc0 + cm];
    const bool extra = grp == G - 1 && cm < SR;
    int cpos = col_pos(cm);
    for (int r = grp;; r += G) {        // one call site: the sample code is large
      int rr = r;
      bool ok = col_ok;
      const bool tail = r >= SR;
      if (tail) {
        if (!extra) break;
        rr = cm; cpos = col_pos(CM); ok = gc0 + CM < (int)nuc;
        if (ok) C = ct[gc0 + CM];
      }
      sample(rr, cpos, C, ok);
      if (tail) break;
    }
  }
  __syncthreads();
  const int lr = threadIdx.x / TC, li = threadIdx.x - lr * TC;
  const i64 j = (i64)blockIdx.y * TR + lr, i = (i64)blockIdx.x * TC + li;
  if (lr >= TR || j >= ncr || i >= nx) return;
  const int r0 = (N - 1) * lr;
  // sample column (N-1)*li + k sits at residue k, slot li (k < N-1) or residue 0, slot li+1 (k = N-1)
  auto at = [&](const double* a, int jj, int k) -> double {
    return a[(r0 + jj) * SCP + (k == N - 1 ? li + 1 : k * (TC + 1) + li)];
  };
  double prod[N][N], row[N], col[N];
#pragma unroll
  for (int jj = 0; jj < N; ++jj)
#pragma unroll
    for (int k = 0; k < N; ++k) prod[jj][k] = at(sx, jj, k) * at(sy, jj, k);      // :178
#pragma unroll
  for (int k = 0; k < N; ++k) { row[k] = at(sx, 0, k); col[k] = at(sy, k, 0); }
  OGG_OUT_ST(daq[j * nx + i], (quad_avg_2d<N>(prod) * Re) * Re);                  // :182,:185
  OGG_OUT_ST(dxq[j * nx + i], quad_avg_1d<N>(row) * Re);                          // :183,:186
  OGG_OUT_ST(dyq[j * (nx + 1) + i], quad_avg_1d<N>(col) * Re);                    // :184,:187
  if (i == nx - 1) {                                                              // dyq has nx+1 columns
#pragma unroll
    for (int k = 0; k < N; ++k) col[k] = at(sy, k, N - 1);
    OGG_OUT_ST(dyq[j * (nx + 1) + nx], quad_avg_1d<N>(col) * Re);
  }
}
// dxq[ny, :] (:183 at the ghost row j = ny, kept by :186): quadrature of the scaled h_i_inv along the exact pole row
template <int N>
__global__ void __launch_bounds__(128) k_bp_pole_row(const BpCol* __restrict__ ct, const BpRow* __restrict__ rt,
                                                     i64 top_row, i64 nx, double lon_bp, double rp, double Re,
                                                     double* __restrict__ dxq_row) {
  const i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (i >= nx) return;
  const BpRow R = rt[top_row];
  double v[N];
#pragma unroll
  for (int k = 0; k < N; ++k) {
    double l, p, a, b;
    bp_sample<false>(R, ct[(N - 1) * i + k], lon_bp, rp, l, p, a, b);
    v[k] = ((a * 2) * kPi) / (double)nx;
  }
  dxq_row[i] = quad_avg_1d<N>(v) * Re;
}

template <int N>
__global__ void __launch_bounds__(128) k_quad_avg(const double* __restrict__ y, i64 m, double* __restrict__ out) {
  const i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (e < m) out[e] = quad_avg_1d<N>(y + e * N);
}
template <int N>
__global__ void __launch_bounds__(128) k_quad_avg2d(const double* __restrict__ y, i64 m, double* __restrict__ out) {
  const i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (e >= m) return;
  double v[N][N];
#pragma unroll
  for (int j = 0; j < N; ++j)
#pragma unroll
    for (int i = 0; i < N; ++i) v[j][i] = y[e * N * N + j * N + i];
  out[e] = quad_avg_2d<N>(v);
}

// ------------------------------------------------------------------------------------------------
// displaced-pole cap (OGG:447-601)
// ------------------------------------------------------------------------------------------------
struct Cx { double re, im; };
OGG_DI Cx cmul(Cx a, Cx b) { return Cx{a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }  // numpy, strict
// RE_DOMINANT (MathFast only): the caller expects |b.re| >= |b.im| -- anything else fails the unit -- and the other
// branch of Smith's algorithm is not compiled in.
template <bool RE_DOMINANT = false, class M>
OGG_DI Cx cdivn(M& m, Cx a, Cx b) {  // numpy CDOUBLE_divide (Smith)
  // MathFast: both parts of the divisor are held to the class "mid" (a zero part -- an axis crossing -- fails the unit).
  // Then |rat| <= 1, the product b.im*rat = b.im**2/b.re has the sign of b.re and a magnitude in [2**-385, |b.re|], so
  // the second divisor b.re + b.im*rat is in [|b.re|, 2|b.re|]: moderate, no guard of its own.
  Cx r;
  if (RE_DOMINANT) m.require(M::abs_ge(b.re, b.im));
  if (RE_DOMINANT || M::abs_ge(b.re, b.im)) {
    // (0,0) divisor: numpy's inf/nan rule; under MathFast a zero b.re fails the guard of the first division instead
    if (!M::kFast && b.re == 0 && b.im == 0) { const double abr = fabs(b.re); r.re = a.re / abr; r.im = a.im / abr; return r; }
    const double rat = m.div_mid(b.im, b.re);
    const double den = b.re + b.im * rat;
    m.proven(rng_moderate(den));
    const double scl = m.div_ok(1.0, den);
    r.re = (a.re + a.im * rat) * scl;
    r.im = (a.im - a.re * rat) * scl;
  } else {
    const double rat = m.div_mid(b.re, b.im);
    const double den = b.im + b.re * rat;
    m.proven(rng_moderate(den));
    const double scl = m.div_ok(1.0, den);
    r.re = (a.re * rat + a.im) * scl;
    r.im = (a.im * rat - a.re) * scl;
  }
  return r;
}
OGG_DI Cx cdivn(Cx a, Cx b) { MathIeee m; return cdivn(m, a, b); }
OGG_DI Cx cis_np(double t) {  // cos(t) + 1j*sin(t) with numpy's promotions (:451,:495)
  const SinCos q = sincos_parts(t);
  const Cx js = cmul(Cx{0.0, 1.0}, Cx{sin_of(q), 0.0});
  return Cx{cos_of(q) + js.re, 0.0 + js.im};
}
struct DpConst { double r_joint; Cx z0; };
OGG_DI DpConst dp_load(const double* __restrict__ c) { return DpConst{c[0], Cx{c[1], c[2]}}; }

OGG_DI Cx dp_col(double lon, Cx z0) {  // e2ithetaprime (:451-452), a function of longitude only
  const Cx e = cis_np(lon * kPi180);
  const Cx p = cmul(Cx{z0.re, -z0.im}, e);
  return cdivn(Cx{e.re - z0.re, e.im - z0.im}, Cx{1.0 - p.re, 0.0 - p.im});
}
OGG_DI double dp_row(double lat, double r_joint) { return p_tan((90 + lat) * kPi180) / r_joint; }  // :448
template <class M>
OGG_DI void dp_vertex(M& m, double r, Cx ep, const DpConst& K, double& lam_raw, double& phi) {
  const Cx z = cmul(Cx{r, 0.0}, ep);                                                         // :454
  const Cx q = cmul(Cx{K.z0.re, -K.z0.im}, z);
  const Cx w = cdivn(m, Cx{z.re + K.z0.re, z.im + K.z0.im}, Cx{1.0 + q.re, 0.0 + q.im});     // :455
  lam_raw = p_atan2(m, w.im, w.re) * kC[k180OverPi];                                          // :457
  const double rw = m.sqrt_(w.re * w.re + w.im * w.im);                                      // :465
  phi = -90 + m.div_pi180(p_atan(m, rw * K.r_joint));                                        // :466
}
OGG_DI void dp_vertex(double r, Cx ep, const DpConst& K, double& lam_raw, double& phi) {
  MathIeee m;
  dp_vertex(m, r, ep, K, lam_raw, phi);
}
// The same vertex under MathFast for k_dp_hsamples.  Preconditions, promised by the caller through m.base (k_dp_quad_setup
// checks them once per table row / column): r, ep.re, ep.im and K.r_joint are in the class "mid" = [2**-127, 2**129).
//  * ZEROOPS: z = (r + 0i) * e' in numpy's strict form is r*e'.re - 0*e'.im, r*e'.im + 0*e'.re.  With finite e' the
//    products by 0 are +-0 and with r*e'.re, r*e'.im in [2**-254, 2**258] (non-zero) adding or subtracting them changes
//    nothing; likewise 0.0 + q.im differs from q.im only for q.im = -0, where the divisor's imaginary part is a zero
//    either way and the first division of cdivn fails the unit.  5 of a vertex's 146 FP64 operations.
//  * p_atan2 holds w.re, w.im to "mid", so w.re**2 + w.im**2 is in [2**-254, 2**259] (the square root needs no guard
//    of its own), rw * r_joint is in [2**-255, 2**260] (not huge) and its arctan in [2**-256, pi/2] (moderate).
OGG_DI bool dp_mid(double v) { return rng_mid(v); }
//  * SMALLAT (the host sets it for caps that end south of 67 S with a pole displaced by at most half the cap's radius:
//    every southern cap the reference's recipes make): tan(phi + 90) < 0.4375 at every vertex, so the latitude's arctan
//    is taken in its first range without the range selection (r_joint > 0 is then part of the caller's promise: the
//    argument rw * r_joint is positive -- no |x|, no copysign), and Smith's division takes its first branch.  A vertex
//    that is not like that fails its unit and is redone: the flag is about speed, never about results.
template <bool SMALLAT = false>
OGG_DI void dp_vertex_fast(MathFast& m, double r, Cx ep, const DpConst& K, double& lam_raw, double& phi) {
  Cx z = Cx{r * ep.re, r * ep.im};
  if (!kZeroOps) z = cmul(Cx{r, 0.0}, ep);                                                   // :454
#ifdef OGG_GUARD_AUDIT
  { const Cx zz = cmul(Cx{r, 0.0}, ep);
    m.proven(__double_as_longlong(zz.re) == __double_as_longlong(z.re) && __double_as_longlong(zz.im) == __double_as_longlong(z.im)); }
#endif
  const Cx q = cmul(Cx{K.z0.re, -K.z0.im}, z);
  // SMALLAT caps have |z0| <= 0.5 and |z| <= 1 (+ eps): the divisor 1 + conj(z0) z has the dominant real part
  const Cx w = cdivn<SMALLAT>(m, Cx{z.re + K.z0.re, z.im + K.z0.im}, Cx{1.0 + q.re, kZeroOps ? q.im : 0.0 + q.im});  // :455
  lam_raw = p_atan2(m, w.im, w.re) * kC[k180OverPi];                                          // :457
  const double ss = w.re * w.re + w.im * w.im;
  m.proven(rng_pos_moderate(ss));
  const double rw = m.sqrt_seq(ss);                                                          // :465
  const double at = SMALLAT ? atan_first_range(m, rw * K.r_joint) : p_atan<false>(m, rw * K.r_joint);
  m.proven(rng_moderate(at));
  phi = -90 + m.div_pi180_seq(at);                                                           // :466
}

struct LamPhi { double l, p; };
__device__ __noinline__ LamPhi dp_vertex_ieee(double r, double epre, double epim, double rj, double z0re, double z0im) {
  LamPhi o;
  dp_vertex(r, Cx{epre, epim}, DpConst{rj, Cx{z0re, z0im}}, o.l, o.p);
  return o;
}
__global__ void k_dp_constants(double lat0, double lam_pole, double r_pole, double* __restrict__ out3) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    out3[0] = p_tan((90 + lat0) * kPi180);                                                   // :494
    const Cx z0 = cmul(Cx{r_pole, 0.0}, cis_np(lam_pole * kPi180));                          // :495
    out3[1] = z0.re;
    out3[2] = z0.im;
  }
}
__global__ void k_set3(double a, double b, double c, double* __restrict__ out3) {
  if (threadIdx.x == 0 && blockIdx.x == 0) { out3[0] = a; out3[1] = b; out3[2] = c; }
}

// raw (not yet unwrapped) projection of arbitrary 2-D input
__global__ void __launch_bounds__(256) k_dp_projection_raw(const double* __restrict__ lon, const double* __restrict__ lat,
                                                           i64 n, const double* __restrict__ cst,
                                                           double* __restrict__ lam, double* __restrict__ phi) {
  const DpConst K = dp_load(cst);
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x)
    dp_vertex(dp_row(lat[e], K.r_joint), dp_col(lon[e], K.z0), K, lam[e], phi[e]);
}

// column / row tables for lattice evaluation: u = lon0 + (i*360)/ni (:479), v = -90 + (j*(lat0+90))/nj (:482)
// `shift` is added to the index first (numerical_hi/hj pass i+eps etc., :538-558)
__global__ void __launch_bounds__(256) k_dp_tables(const double* __restrict__ iv, i64 i0, i64 ncol, double ishift,
                                                   const double* __restrict__ jv, i64 j0, i64 nrow, double jshift, i64 ni,
                                                   i64 nj, double lon0, double lat0, const double* __restrict__ cst,
                                                   Cx* __restrict__ ct, double* __restrict__ rt) {
  const DpConst K = dp_load(cst);
  const i64 t = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (ct && t < ncol) {
    const double base = iv ? iv[t] : (double)(i0 + t);
    const double idx = ishift != 0.0 ? base + ishift : base;
    ct[t] = dp_col(lon0 + (idx * 360.0) / (double)ni, K.z0);
  }
  if (rt && t < nrow) {
    const double base = jv ? jv[t] : (double)(j0 + t);
    const double idx = jshift != 0.0 ? base + jshift : base;
    const double a = -90.0, b = lat0;
    rt[t] = dp_row(a + (idx * (b - a)) / (double)nj, K.r_joint);
  }
}
// pi180 != NULL: MathFast (dp_vertex_fast; its preconditions are checked here per vertex) with the compiler's code as redo
__global__ void __launch_bounds__(256) k_dp_lattice_raw(const Cx* __restrict__ ct, const double* __restrict__ rt,
                                                        i64 nrow, i64 ncol, const double* __restrict__ cst,
                                                        double* __restrict__ lam, double* __restrict__ phi,
                                                        const double* __restrict__ pi180) {
  const i64 c = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  const DpConst K = dp_load(cst);
  const Cx ep = ct[c];
  if (pi180) {
    MathFast mf(pi180);
    const bool col_ok = dp_mid(ep.re) && dp_mid(ep.im) && dp_mid(K.r_joint);
    for (i64 r = blockIdx.y; r < nrow; r += gridDim.y) {
      const double rv = rt[r];
      double l, p;
      mf.base = col_ok && dp_mid(rv);
      mf.reset();
      dp_vertex_fast(mf, rv, ep, K, l, p);
      if (!mf.ok()) { const LamPhi o = dp_vertex_ieee(rv, ep.re, ep.im, K.r_joint, K.z0.re, K.z0.im); l = o.l; p = o.p; }
      lam[r * ncol + c] = l;
      phi[r * ncol + c] = p;
    }
    return;
  }
  for (i64 r = blockIdx.y; r < nrow; r += gridDim.y) dp_vertex(rt[r], ep, K, lam[r * ncol + c], phi[r * ncol + c]);
}

// --- monotonic_bounding (OGG:470-475) as a scan over 2-state maps ---------------------------------------
// Element t with raw value x_t and predecessor raw value p.  If the predecessor ended up shifted its value is
// fl(p - 360), else p.  m0 = [fl(x_t - p) > 100], m1 = [fl(x_t - fl(p - 360)) > 100] say whether t is shifted
// given the predecessor's state; composition of such maps is associative, so one row is a prefix scan.
// Up to 16 independent scans ("variants") are packed bitwise into one word.
struct Map2 { unsigned f0, f1; };                      // image of state 0 / state 1, one bit per variant
OGG_DI Map2 compose(Map2 first, Map2 then) {           // then o first
  return Map2{(first.f0 & then.f1) | (~first.f0 & then.f0), (first.f1 & then.f1) | (~first.f1 & then.f0)};
}
OGG_DI unsigned apply_map(Map2 m, unsigned s) { return (s & m.f1) | (~s & m.f0); }

// Block-wide inclusive scan; returns the state AFTER this thread's element given the tile's carry-in state.
// sh: shared scratch of 2*kWarps+2 unsigned.  All threads of the block must call.
template <int THREADS>
OGG_DI unsigned tile_scan_states(Map2 m, unsigned s_in, unsigned* sh) {
  constexpr int kWarps = THREADS / 32;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    Map2 e;
    e.f0 = __shfl_up_sync(0xffffffffu, m.f0, off);
    e.f1 = __shfl_up_sync(0xffffffffu, m.f1, off);
    if (lane >= (unsigned)off) m = compose(e, m);
  }
  if (lane == 31u) { sh[2 * warp] = m.f0; sh[2 * warp + 1] = m.f1; }
  __syncthreads();
  unsigned s = s_in;
  for (unsigned w = 0; w < warp; ++w) s = apply_map(Map2{sh[2 * w], sh[2 * w + 1]}, s);
  const unsigned st = apply_map(m, s);
  __syncthreads();  // sh may be reused by the caller / next tile
  (void)kWarps;
  return st;
}

constexpr int kScanThreads = 256;
// in place on x[nj][ni]; x0 from x0_dev[0] if non-NULL (device scalar) else x0_host
__global__ void __launch_bounds__(kScanThreads) k_monotonic(double* __restrict__ x, i64 nj, i64 ni, double x0_host,
                                                            const double* __restrict__ x0_dev) {
  __shared__ double raw[kScanThreads + 1];
  __shared__ unsigned sh[2 * (kScanThreads / 32)];
  __shared__ unsigned s_carry;
  const double x0 = x0_dev ? x0_dev[0] : x0_host;
  for (i64 j = blockIdx.x; j < nj; j += gridDim.x) {
    double* row = x + j * ni;
    if (threadIdx.x == 0) { raw[0] = row[0] * 0 + x0; s_carry = 0u; }  // x_im1 = x[:,0]*0 + x_0 (:471)
    __syncthreads();
    for (i64 t0 = 0; t0 < ni; t0 += kScanThreads) {
      const i64 c = t0 + threadIdx.x;
      const bool valid = c < ni;
      const double v = valid ? row[c] : 0.0;
      raw[threadIdx.x + 1] = v;
      __syncthreads();
      const double p = raw[threadIdx.x];
      Map2 m{0u, 1u};  // identity for padding lanes
      if (valid) { m.f0 = (v - p) > 100 ? 1u : 0u; m.f1 = (v - (p - 360)) > 100 ? 1u : 0u; }
      const unsigned s_in = s_carry;
      const unsigned st = tile_scan_states<kScanThreads>(m, s_in, sh);
      if (valid) row[c] = (st & 1u) ? v - 360 : v;                       // :473
      const i64 last = (ni - t0 < kScanThreads ? ni - t0 : kScanThreads) - 1;
      if ((i64)threadIdx.x == last) { raw[0] = v; s_carry = st; }
      __syncthreads();
    }
  }
}

// great_arc_distance tail (OGG:527-532) on two projected meshes
template <class M>
OGG_DI double haversine(M& m, double lam0, double phi0, double lam1, double phi1) {
  const double l0 = lam0 * kPi180, p0 = phi0 * kPi180, l1 = lam1 * kPi180, p1 = phi1 * kPi180;
  const double dphi = p1 - p0, dlam = l1 - l0;
  const double s1 = p_sin(m, 0.5 * dphi), s2 = p_sin(m, 0.5 * dlam);
  const double d = s1 * s1 + ((s2 * s2) * p_cos(m, p0)) * p_cos(m, p1);  // :531
  // MathFast: d in [2**-127, 2**129) (coincident points -- d = 0 -- fail the unit), so its root is at least 2**-64
  return 2.0 * p_asin<true>(m, m.sqrt_mid(d));                     // :532
}
OGG_DI double haversine(double lam0, double phi0, double lam1, double phi1) {
  MathIeee m;
  return haversine(m, lam0, phi0, lam1, phi1);
}
__global__ void __launch_bounds__(256) k_haversine(const double* __restrict__ lam0, const double* __restrict__ phi0,
                                                   const double* __restrict__ lam1, const double* __restrict__ phi1,
                                                   i64 n, double* __restrict__ out) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x)
    out[e] = haversine(lam0[e], phi0[e], lam1[e], phi1[e]);
}
template <int ORDER>
OGG_DI double fd_combine(double ds2, double ds4, double ds6, double reps) {  // :539-546
  if (ORDER == 2) return (0.5 * ds2) * reps;
  if (ORDER == 4) return ((8.0 * ds2 - ds4) * (1.0 / 12.0)) * reps;
  return (((45.0 * ds2 - 9.0 * ds4) + ds6) * (1.0 / 60.0)) * reps;
}
__global__ void __launch_bounds__(256) k_fd_combine(const double* __restrict__ ds2, const double* __restrict__ ds4,
                                                    const double* __restrict__ ds6, int order, double reps, i64 n,
                                                    double* __restrict__ out) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) {
    if (order == 2) out[e] = fd_combine<2>(ds2[e], 0, 0, reps);
    else if (order == 4) out[e] = fd_combine<4>(ds2[e], ds4[e], 0, reps);
    else out[e] = fd_combine<6>(ds2[e], ds4[e], ds6[e], reps);
  }
}

// out-of-line copy for k_dp_hsamples: with the haversine inlined four times next to four inlined projections the
// kernel's loop body overflowed the instruction cache (stalled_no_instruction 2.1 per issue); one shared copy is 4.5 %
// faster.  Same instructions, same results.
__device__ __noinline__ double haversine_ieee(double lam0, double phi0, double lam1, double phi1) {
  return haversine(lam0, phi0, lam1, phi1);
}
// The haversine of k_dp_hsamples.  Its four trigonometric calls share one straight-line block in the case every sample
// of a real grid is in: the two half-differences are tiny (quadrant k = 0, where the Cody-Waite reduction subtracts three
// exact zeros and returns its argument: skipped) and the two latitudes sit in an odd quadrant, so that all four values
// come from the SINE polynomial -- evaluated four times on one set of coefficients in uniform registers, without the
// per-call quadrant branches.  sin(dphi/2) and sin(dlam/2) enter :531 squared: their sign (and the sign of a zero
// argument, the one thing the skipped reduction can change) does not reach the result.  Anything else takes the general
// code; both give the bits of haversine<MathIeee> (tests: goldens, oracle parity, test_dp_haversine_quadrants).
__device__ __noinline__ double haversine_ol(double lam0, double phi0, double lam1, double phi1) {
  MathFast m;
#ifndef OGG_NO_HAV_FAST
  const double2 ra = m.pair(T_RED0), rb = m.pair(T_RED1);
  const double l0 = lam0 * kPi180, p0 = phi0 * kPi180, l1 = lam1 * kPi180, p1 = phi1 * kPi180;
  const double x1 = 0.5 * (p1 - p0), x2 = 0.5 * (l1 - l0);
  const double k3 = rint(p0 * ra.x), k4 = rint(p1 * ra.x);
  // the half-differences' quadrant numbers are only tested for zero: converted with round-to-nearest-even (= rint) in one
  // instruction, as 64-bit integers so that a huge quotient cannot alias to 0
  const int i1 = (__double2ll_rn(x1 * ra.x) == 0LL) ? 0 : 1, i2 = (__double2ll_rn(x2 * ra.x) == 0LL) ? 0 : 1;
  const int i3 = static_cast<int>(static_cast<long long>(k3)), i4 = static_cast<int>(static_cast<long long>(k4));
  double d;
  if ((i1 | i2) == 0 && (i3 & i4 & 1) != 0) {
    const double r3 = ((p0 - k3 * ra.y) - k3 * rb.x) - k3 * rb.y;
    const double r4 = ((p1 - k4 * ra.y) - k4 * rb.x) - k4 * rb.y;
    const double s1 = sin_kernel(m, x1, x1 * x1), s2 = sin_kernel(m, x2, x2 * x2);
    double c0 = sin_kernel(m, r3, r3 * r3), c1 = sin_kernel(m, r4, r4 * r4);
    if ((i3 + 1) & 2) c0 = -c0;                                     // p_cos: quadrant n -> c, -s, -c, s
    if ((i4 + 1) & 2) c1 = -c1;
    d = s1 * s1 + ((s2 * s2) * c0) * c1;                            // :531
  } else {
    const double s1 = p_sin(m, x1), s2 = p_sin(m, x2);
    d = s1 * s1 + ((s2 * s2) * p_cos(m, p0)) * p_cos(m, p1);
  }
  // :532; d in [2**-127, 2**129) or the unit fails, so its root is at least 2**-64; a root of 0.5 or more (points more
  // than 60 degrees apart) fails the unit too
  const double v = 2.0 * p_asin_small_pos(m, m.sqrt_mid(d));
#else
  const double v = haversine(m, lam0, phi0, lam1, phi1);
#endif
  return m.ok() ? v : haversine_ieee(lam0, phi0, lam1, phi1);   // e.g. coincident points: sqrt(0), asin(0)
}
// --- fused numerical_hi + numerical_hj on a lattice of unique samples (OGG:583-584) ------------------------------
// For finite-difference order 2*NV there are 4*NV shifted meshes (i+-m*eps at this row; j+-m*eps at this column),
// each needing its own longitude unwrap along the row: their 2-state maps are packed into one word and scanned
// together.  Column table ct: [ncol][1+2NV] = e' at {i, i+eps, i-eps, i+2eps, i-2eps, ...}; row table rt likewise.
//
// Work item = (sample row, tile of THREADS-1 new columns), one item per CTA.  A CTA draws its item number from a
// ticket counter when it starts running, so item k is always taken by a CTA that is already resident -- no
// co-residency requirement on the grid (a plain launch: the block scheduler interleaves the CTAs of the other
// sub-grids' kernels as these retire) and an even load for ANY number of sample rows (a row-slab on one of 8 GPUs
// has far fewer rows than there are CTA slots).  Thread 0 of a tile re-evaluates the previous tile's last column (the
// predecessor raw longitudes of thread 1), so a tile needs from its left neighbours only the shift STATE (one bit
// per mesh) after that column: a decoupled look-back.  Each tile publishes a 64-bit descriptor: first its aggregate
// 2-state map, then -- once the state at its left edge is known -- the state after its last column; a tile resolves
// its left-edge state by walking back over aggregates until it meets a published state (tile 0 of a row knows its
// state: not shifted).  A tile waits only on lower-numbered items, which were drawn earlier by running CTAs, so the
// lowest unfinished item can always proceed.
constexpr unsigned long long kDescAgg = 1ull << 62, kDescState = 2ull << 62;
OGG_DI unsigned long long desc_load(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
OGG_DI void desc_store(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// The 4*NV projected meshes of a thread's column are parked in shared memory and the per-mesh work runs in partly
// rolled loops: fewer live registers, and less code than the fully unrolled form (which misses the instruction cache).
// Items are numbered tile-major (all rows' tile t before any tile t+1), so that a tile's left neighbour was drawn
// nrow tickets earlier and its descriptor is normally there when looked for.
template <int NV, int THREADS, int MINB, int UNR, bool SMALLAT>
__global__ void __launch_bounds__(THREADS, MINB) k_dp_hsamples(const Cx* __restrict__ ct, const double* __restrict__ rt,
                                                                 i64 nrow, i64 ncol, const double* __restrict__ x0v,
                                                                 const double* __restrict__ cst, double reps,
                                                                 double* __restrict__ hx, double* __restrict__ hy,
                                                                 unsigned long long* __restrict__ desc,
                                                                 const int* __restrict__ tok) {
  constexpr int NM = 4 * NV, NT = 1 + 2 * NV, TW = THREADS - 1, kWarps = THREADS / 32;
  constexpr unsigned kMask = (1u << NM) - 1u;
#ifndef OGG_NO_IDX32
  // one array for both parks (one base address + immediate offsets) and 32-bit item / column numbers (the host refuses
  // launches of 2**31 items or more): at 48 registers the compiler re-derives the table and shared-memory addresses
  // for every vertex instead of keeping them, and a 64-bit derivation is twice the instructions
  __shared__ double parkS[2 * NM][THREADS];
  double (*const lamS)[THREADS] = parkS;
  double (*const phiS)[THREADS] = parkS + NM;
#else
  __shared__ double lamS[NM][THREADS];
  __shared__ double phiS[NM][THREADS];
#endif
  __shared__ unsigned shw[2 * kWarps];
  __shared__ unsigned s_state;
  const DpConst K = dp_load(cst);
  MathFast mf(cst + 3);                                  // pi/180 and its refined reciprocal, stored by k_dp_quad_setup
  const i64 ntile = (ncol + TW - 1) / TW;
  const i64 nitem = nrow * ntile;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  __shared__ i64 s_item;
  {
    if (threadIdx.x == 0) s_item = (i64)atomicAdd(desc + nitem, 1ull);   // the ticket counter follows the descriptors
    __syncthreads();
#ifndef OGG_NO_IDX32
    const unsigned item = (unsigned)s_item;
    const unsigned t = item / (unsigned)nrow, r = item - t * (unsigned)nrow;
    const int c = (int)(t * (unsigned)TW) + (int)threadIdx.x - 1;
    const bool virt = c < 0;
    const bool valid = !virt && c < (int)ncol;
#else
    const i64 item = s_item;
    const i64 t = item / nrow, r = item - t * nrow;
    const i64 c = t * TW + (i64)threadIdx.x - 1;
    const bool virt = c < 0;
    const bool valid = !virt && c < ncol;
#endif
    {
      // every thread projects (the few past the last column redo column 0 and drop the result): no per-thread
      // predicate around the loop, so its constants stay in uniform registers (-3 %)
      const unsigned cidx = valid ? (unsigned)c : 0u;
      const Cx* cc = ct + (size_t)cidx * NT;
      const double* rr = rt + (size_t)r * NT;
      mf.base = (tok[cidx] & tok[ncol + r]) != 0 && (!SMALLAT || K.r_joint > 0.0);   // dp_vertex_fast's preconditions
      // A vertex whose guard flag fails (a zero / non-finite / extreme operand: the pole row, an axis crossing) is
      // redone with the compiler's operators AFTER the loop: a call inside it, even one never taken, makes the compiler
      // re-load the ~20 coefficients it keeps in uniform registers after every vertex (-80 LDCU per sample).
      unsigned bad = 0u;
#pragma unroll UNR
      for (int m = 0; m < NM; ++m) {
        const int lvl = m >> 2, kind = m & 3;
        const double rv = rr[kind < 2 ? 0 : 2 * lvl + kind - 1];        // r at j (kinds 0,1) or j +- (lvl+1) eps
        const Cx ep = cc[kind < 2 ? 2 * lvl + kind + 1 : 0];            // e' at i +- (lvl+1) eps (kinds 0,1) or i
        double l, p;
        mf.reset();
        dp_vertex_fast<SMALLAT>(mf, rv, ep, K, l, p);
#ifdef OGG_DP_REDO_INLOOP
        if (!mf.ok()) {
          const LamPhi o = dp_vertex_ieee(rv, ep.re, ep.im, K.r_joint, K.z0.re, K.z0.im);
          l = o.l; p = o.p;
        }
#else
        if (!mf.ok()) bad |= 1u << m;
#endif
        lamS[m][threadIdx.x] = l;
        phiS[m][threadIdx.x] = p;
      }
      if (bad) {
#pragma unroll 1
        for (int m = 0; m < NM; ++m) {
          if (!((bad >> m) & 1u)) continue;
          const int lvl = m >> 2, kind = m & 3;
          const double rv = rr[kind < 2 ? 0 : 2 * lvl + kind - 1];
          const Cx ep = cc[kind < 2 ? 2 * lvl + kind + 1 : 0];
          const LamPhi o = dp_vertex_ieee(rv, ep.re, ep.im, K.r_joint, K.z0.re, K.z0.im);
          lamS[m][threadIdx.x] = o.l;
          phiS[m][threadIdx.x] = o.p;
        }
      }
    }
    if (!valid) {
#pragma unroll
      for (int m = 0; m < NM; ++m) { lamS[m][threadIdx.x] = virt ? x0v[m] : 0.0; }
    }
    __syncthreads();
    Map2 mp{0u, 0xffffffffu};
    if (valid && threadIdx.x > 0) {
#ifndef OGG_NO_MAP_FAST
      // |v - p| < 100 decides both tests without evaluating them: v - p > 100 is false, and v - fl(p - 360) lies within
      // 1e-13 of (v - p) + 360 > 260, so the second is true -- the identity map.  Only a thread that sees a jump (or a
      // NaN: its high word is above every finite one's) evaluates the reference's two differences.
      int far = 0;
#pragma unroll
      for (int m = 0; m < NM; ++m)
        far = max(far, __double2hiint(lamS[m][threadIdx.x] - lamS[m][threadIdx.x - 1]) & 0x7fffffff);
      mp.f0 = 0u; mp.f1 = kMask;
      if (far >= 0x40590000)   // |v - p| >= 100 (high word of 100.0; low word 0)
#endif
      {
        mp.f0 = 0u; mp.f1 = 0u;
#pragma unroll
        for (int m = 0; m < NM; ++m) {
          const double v = lamS[m][threadIdx.x], p = lamS[m][threadIdx.x - 1];
          mp.f0 |= ((v - p) > 100 ? 1u : 0u) << m;
          mp.f1 |= ((v - (p - 360)) > 100 ? 1u : 0u) << m;
        }
      }
    }
#ifndef OGG_NO_SCAN_FAST
    // a tile without a single +-360 jump (almost every tile: a row wraps once) has only identity maps: every prefix is
    // the identity and the warp / block scans are skipped (the vote is the barrier the scan would need anyway)
    const bool all_ident = __syncthreads_and(((mp.f0 & kMask) == 0u && (mp.f1 & kMask) == kMask) ? 1 : 0) != 0;
    if (!all_ident)
#endif
    {
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        Map2 e;
        e.f0 = __shfl_up_sync(0xffffffffu, mp.f0, off);
        e.f1 = __shfl_up_sync(0xffffffffu, mp.f1, off);
        if (lane >= (unsigned)off) mp = compose(e, mp);
      }
      if (lane == 31u) { shw[2 * warp] = mp.f0; shw[2 * warp + 1] = mp.f1; }
      __syncthreads();
      if (warp > 0) {
        Map2 pre{shw[0], shw[1]};
        for (unsigned w = 1; w < warp; ++w) pre = compose(pre, Map2{shw[2 * w], shw[2 * w + 1]});
        mp = compose(pre, mp);
      }
    }
    if (threadIdx.x == THREADS - 1) {
      unsigned s_in = 0u;
      if (t > 0) {
        desc_store(desc + item, kDescAgg | ((unsigned long long)(mp.f0 & kMask) << 16) | (mp.f1 & kMask));
        Map2 acc{0u, 0xffffffffu};
        for (i64 k = item - nrow;; k -= nrow) {          // tiles t-1, t-2, ... of this row: tile 0 always has a state
          unsigned long long d;
          do { d = desc_load(desc + k); } while ((d >> 62) == 0ull);
          if ((d >> 62) == 2ull) { s_in = apply_map(acc, (unsigned)(d & 0xffffull)); break; }
          acc = compose(Map2{(unsigned)((d >> 16) & 0xffffull), (unsigned)(d & 0xffffull)}, acc);
        }
      }
      desc_store(desc + item, kDescState | (apply_map(mp, s_in) & kMask));
      s_state = s_in;
    }
    __syncthreads();
    if (valid && threadIdx.x > 0) {
      const unsigned st = apply_map(mp, s_state);
#pragma unroll UNR
      for (int q = 0; q < 2 * NV; ++q) {                                  // pair q: meshes 2q (plus) and 2q+1 (minus)
        double l0 = lamS[2 * q][threadIdx.x], l1 = lamS[2 * q + 1][threadIdx.x];
        if ((st >> (2 * q)) & 1u) l0 = l0 - 360;                           // :473
        if ((st >> (2 * q + 1)) & 1u) l1 = l1 - 360;
        phiS[2 * q][threadIdx.x] = haversine_ol(l0, phiS[2 * q][threadIdx.x], l1, phiS[2 * q + 1][threadIdx.x]);
      }
      // pair 2m = i-direction at level m, pair 2m+1 = j-direction (:538-546, :553-561)
      hx[r * ncol + c] = fd_combine<2 * NV>(phiS[0][threadIdx.x], NV > 1 ? phiS[4][threadIdx.x] : 0.0,
                                            NV > 2 ? phiS[8][threadIdx.x] : 0.0, reps);
      hy[r * ncol + c] = fd_combine<2 * NV>(phiS[2][threadIdx.x], NV > 1 ? phiS[6][threadIdx.x] : 0.0,
                                            NV > 2 ? phiS[10][threadIdx.x] : 0.0, reps);
    }
  }
}
// x_0 of every shifted mesh: lon0 + ((i[0] (+/-) s)*360)/ni  (:463 with :479)
template <int NV>
__global__ void k_dp_x0(const double* __restrict__ iv, double i0, double eps, i64 ni, double lon0,
                        double* __restrict__ x0v) {
  const int m = threadIdx.x;
  if (m >= 4 * NV) return;
  const int lvl = m / 4, kind = m % 4;
  const double s = lvl == 0 ? eps : (lvl == 1 ? 2.0 * eps : 3.0 * eps);
  const double first = iv ? iv[0] : i0;
  double idx = first;
  if (kind == 0) idx = first + s;
  if (kind == 1) idx = first - s;
  x0v[m] = lon0 + (idx * 360.0) / (double)ni;
}


// ------------------------------------------------------------------------------------------------
// self-test of the MathFast policy (ogg_trig.cuh) against the compiler's IEEE division / square root
// ------------------------------------------------------------------------------------------------
OGG_DI unsigned long long splitmix64(unsigned long long x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}
// operand generators.  mode 0: random significands, exponents spread over MathFast's whole guarded range;
// 1: raw random bit patterns (zeros, subnormals, infinities, NaNs, extremes: the guards must reject what the sequence
// cannot do); 2: significands within a few ulps of 1, 2 and of each other (hard rounding cases: quotients next to a
// rounding boundary); 3: small-integer significands (exact and half-way quotients).
OGG_DI double gen_operand(unsigned long long h, int mode) {
  const unsigned long long mant = h & 0xfffffffffffffull, sign = (h >> 63) << 63;
  if (mode == 1) {
    const unsigned sel = (unsigned)(h >> 52) & 7u;
    if (sel == 0) return __longlong_as_double((long long)(sign | (h & 0xffull)));                // zero / tiny subnormals
    if (sel == 1) return __longlong_as_double((long long)(sign | 0x7ff0000000000000ull | ((h >> 8) & 1 ? mant : 0)));
    return __longlong_as_double((long long)h);
  }
  const int span = 0x5c0 - 0x240;
  const unsigned long long e = 0x240 + ((h >> 52) & 0x7ff) % span;
  if (mode == 2) {
    const unsigned long long near = (h >> 8) & 1 ? (mant & 0x3f) : (0xfffffffffffffull - (mant & 0x3f));
    return __longlong_as_double((long long)(sign | (e << 52) | near));
  }
  if (mode == 3) return scale2((double)(1 + (mant & 0xfff)), (int)(e % 200) - 100) * (sign ? -1.0 : 1.0);
  return __longlong_as_double((long long)(sign | (e << 52) | mant));
}
// counts[0]: operand sets accepted by the guards; counts[1]: accepted sets whose MathFast result differs in any bit
// from the IEEE result; counts[2]: sets in MathFast's documented range that a guard rejected (must be 0 for modes != 1)
__global__ void __launch_bounds__(256) k_selftest_divsqrt(i64 n, unsigned long long seed, int mode,
                                                          unsigned long long* __restrict__ counts) {
  unsigned long long acc = 0, bad = 0, rej = 0;
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) {
    const unsigned long long h0 = splitmix64(seed + 2ull * (unsigned long long)e), h1 = splitmix64(h0 ^ 0x5851f42d4c957f2dull);
    const double a = gen_operand(h0, mode), b = gen_operand(h1, mode);
    MathFast m;
    const double q = m.div(a, b);
    if (m.ok()) { ++acc; if (__double_as_longlong(q) != __double_as_longlong(a / b)) ++bad; } else if (mode != 1) ++rej;
    m.reset();
    const double q1 = m.div_num_ok(1.0, b);
    if (m.ok()) { ++acc; if (__double_as_longlong(q1) != __double_as_longlong(1.0 / b)) ++bad; } else if (mode != 1) ++rej;
    m.reset();
    double by[2];
    MathFast::store_divisor(by, b);                       // what a set-up kernel stores ...
    const MathFast::Divisor d = MathFast::load_divisor(by);   // ... and a quadrature kernel loads
    const double q2 = m.div_by(a, d);
    if (m.ok() && d.y == d.y) { ++acc; if (__double_as_longlong(q2) != __double_as_longlong(a / b)) ++bad; } else if (mode != 1) ++rej;
    m.reset();
    const double x = MathFast::abs_(a);
    const double r = m.sqrt_(x);
    if (m.ok()) { ++acc; if (__double_as_longlong(r) != __double_as_longlong(sqrt(x))) ++bad; } else if (mode != 1) ++rej;
    m.reset();
    const double z = m.div_ok(0.0, MathFast::abs_(b));   // +0 numerators (atan_pos, asin_r) over a positive denominator
    m.require_moderate(b);
    if (m.ok()) { ++acc; if (__double_as_longlong(z) != __double_as_longlong(0.0 / MathFast::abs_(b))) ++bad; }
  }
  if (acc) atomicAdd(counts, acc);
  if (bad) atomicAdd(counts + 1, bad);
  if (rej) atomicAdd(counts + 2, rej);
}
// mode 4: the portable functions that divide (tan, arctan, arctan2, arcsin, arccos) under MathFast against the same
// functions under MathIeee, on arguments spread over [-8, 8] (tan, arctan*), [-1, 1] (arcsin, arccos) and tiny values
__global__ void __launch_bounds__(256) k_selftest_trig(i64 n, unsigned long long seed, unsigned long long* __restrict__ counts) {
  unsigned long long acc = 0, bad = 0;
  auto same = [](double u, double v) { return __double_as_longlong(u) == __double_as_longlong(v); };
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) {
    const unsigned long long h0 = splitmix64(seed + 2ull * (unsigned long long)e), h1 = splitmix64(h0 ^ 0x5851f42d4c957f2dull);
    const double u0 = (double)(h0 >> 11) * 0x1p-53, u1 = (double)(h1 >> 11) * 0x1p-53;   // [0, 1)
    const double scale = (h0 & 7) == 0 ? 1e-9 : 1.0;
    double x = (16.0 * u0 - 8.0) * scale, y = (16.0 * u1 - 8.0) * scale, c = (2.0 * u0 - 1.0) * scale;
    if (((h0 >> 3) & 15) == 1) {
      // every 16th set sits on or within 2 ulps of a range boundary of arctan / arcsin / arccos (the integer-pipe
      // comparisons of MathFast decide on the high word there), with either sign
      const double edges[8] = {0.4375, 0.6875, 1.1875, 2.4375, 0.5, 1.0, 0.25, 2.0};
      const double edge = edges[(h0 >> 7) & 7];
      const long long off = (long long)((h0 >> 10) % 5) - 2;
      const double v = __longlong_as_double(__double_as_longlong(edge) + off);
      x = (h0 >> 20) & 1 ? -v : v;
      c = x;
      if ((h0 >> 21) & 1) y = x * ((h1 >> 5) & 1 ? 1.0 : -1.0);   // arctan2 with |y/x| = 1
    }
    MathFast m;
    double v = p_tan(m, x);
    if (m.ok()) { ++acc; if (!same(v, p_tan(x))) ++bad; }
    m.reset(); v = p_atan(m, x);
    if (m.ok()) { ++acc; if (!same(v, p_atan(x))) ++bad; }
    m.reset(); v = p_atan2(m, y, x);
    if (m.ok()) { ++acc; if (!same(v, p_atan2(y, x))) ++bad; }
    m.reset(); v = p_asin(m, c);
    if (m.ok()) { ++acc; if (!same(v, p_asin(c))) ++bad; }
    m.reset(); v = p_acos(m, c);
    if (m.ok()) { ++acc; if (!same(v, p_acos(c))) ++bad; }
    // sin / cos do not divide: the two policies differ only in where the coefficient pairs come from
    ++acc; if (!same(p_sin(m, x), p_sin(x))) ++bad;
    ++acc; if (!same(p_cos(m, y), p_cos(y))) ++bad;
    const SinCos qa = sincos_parts(m, x), qb = sincos_parts(x);
    ++acc; if (!same(qa.s, qb.s) || !same(qa.c, qb.c) || qa.n != qb.n) ++bad;
  }
  if (acc) atomicAdd(counts, acc);
  if (bad) atomicAdd(counts + 1, bad);
}

// ------------------------------------------------------------------------------------------------
// reductions
// ------------------------------------------------------------------------------------------------
constexpr int kRedThreads = 256;
constexpr int kRedBlocks = 592;  // 4 x 148
OGG_DI double block_sum(double v, double* sh) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (warp == 0) {
    t = lane < kRedThreads / 32 ? sh[lane] : 0.0;
#pragma unroll
    for (int off = 4; off > 0; off >>= 1) t += __shfl_down_sync(0xffffffffu, t, off);
  }
  return t;  // valid in thread 0
}
__global__ void __launch_bounds__(kRedThreads) k_sum_partial(const double* __restrict__ a, i64 n, i64 stride,
                                                             double* __restrict__ partial) {
  __shared__ double sh[kRedThreads / 32];
  double v = 0.0;
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) v += a[e * stride];
  const double t = block_sum(v, sh);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}
__global__ void __launch_bounds__(kRedThreads) k_sum_final(const double* __restrict__ partial, int np,
                                                           double* __restrict__ out) {
  __shared__ double sh[kRedThreads / 32];
  double v = 0.0;
  for (int e = threadIdx.x; e < np; e += blockDim.x) v += partial[e];
  const double t = block_sum(v, sh);
  if (threadIdx.x == 0) out[0] = t;
}

// Order-independent 64-bit checksum of an array's bit patterns: sum over elements of mix(bits) mod 2^64 (mix = the
// splitmix64 finaliser).  The checksum of a grid is the sum of its row-slabs' checksums -- the array-level parity tool
// at resolutions whose arrays are too large to copy out (the role chksum, OGG:19-30, plays in the reference).
OGG_DI unsigned long long mix64(unsigned long long z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
__global__ void __launch_bounds__(kRedThreads) k_checksum(const double* __restrict__ a, i64 n,
                                                          unsigned long long* __restrict__ acc) {
  unsigned long long v = 0ull;
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x)
    v += mix64((unsigned long long)__double_as_longlong(a[e]));
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
  if ((threadIdx.x & 31u) == 0u) atomicAdd(acc, v);   // integer addition: the result does not depend on the order
}

// Byte-order reversal of 8-byte elements (the netCDF classic format is big-endian): `out` may be device memory or
// pinned host memory mapped into the device's address space -- the stores then cross PCIe directly, which replaces the
// host-side astype(">f8") pass AND the separate device->host copy of write_nc (OGG:773-829).
__global__ void __launch_bounds__(256) k_bswap64(const unsigned long long* __restrict__ in, unsigned long long* __restrict__ out,
                                                 i64 n) {
  for (i64 e = blockIdx.x * (i64)blockDim.x + threadIdx.x; e < n; e += (i64)gridDim.x * blockDim.x) {
    const unsigned long long v = in[e];
    const unsigned lo = (unsigned)v, hi = (unsigned)(v >> 32);
    out[e] = ((unsigned long long)__byte_perm(lo, 0u, 0x0123) << 32) | (unsigned long long)__byte_perm(hi, 0u, 0x0123);
  }
}

// FP64 issue-rate probe: 8 independent chains of dependent (mul, add) pairs
__global__ void __launch_bounds__(256) k_fp64_rate(i64 iters, double seed, double* __restrict__ sink) {
  double a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = seed + (double)(threadIdx.x + k);
  const double m = 1.0000001, c = 1e-9;
  for (i64 it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = a[k] * m + c;  // -fmad=false: DMUL + DADD
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k];
  if (s == 12345.678) sink[0] = s;
}

// ------------------------------------------------------------------------------------------------
// host-side composition helpers
// ------------------------------------------------------------------------------------------------
// pi/180 with its refined reciprocal (MathFast::Divisor) in device memory, computed once per device: the short mesh /
// angle kernels divide by it under MathFast without a set-up launch of their own.
__global__ void k_store_pi180(double* __restrict__ by) {
  if (threadIdx.x == 0 && blockIdx.x == 0) MathFast::store_divisor(by, kPi180);
}
static const double* pi180_pair() {
  static double* pairs[kMaxDevices] = {nullptr};
  static std::mutex mu;
  const int dev = current_device();
  std::lock_guard<std::mutex> lock(mu);
  if (!pairs[dev]) {
    double* d = nullptr;
    cudaStream_t st = nullptr;
    if (cudaMalloc(&d, 2 * sizeof(double)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); cudaFree(d); return nullptr; }
    k_store_pi180<<<1, 32, 0, st>>>(d);
    const cudaError_t e = cudaStreamSynchronize(st);   // visible to every stream before the pointer is handed out
    cudaStreamDestroy(st);
    if (e != cudaSuccess) { cudaGetLastError(); cudaFree(d); return nullptr; }
    pairs[dev] = d;
  }
  return pairs[dev];
}
static dim3 lattice_grid(i64 ncol, i64 nrow, int threads) {
  const unsigned gx = cdiv_u(ncol, threads);
  i64 gy = ((i64)num_sms() * 16 + gx - 1) / gx;
  if (gy > nrow) gy = nrow;
  if (gy < 1) gy = 1;
  if (gy > 65535) gy = 65535;
  return dim3(gx, (unsigned)gy, 1);
}

template <typename KERN, typename... Args>
static int launch_quad(int order, KERN k2, KERN k3, KERN k4, KERN k5, dim3 g, dim3 b, cudaStream_t st, Args... args) {
  if (order == 2) OGG_LAUNCH(KID_QUAD, (k2), (g, b), (st), args...);
  else if (order == 3) OGG_LAUNCH(KID_QUAD, (k3), (g, b), (st), args...);
  else if (order == 4) OGG_LAUNCH(KID_QUAD, (k4), (g, b), (st), args...);
  else OGG_LAUNCH(KID_QUAD, (k5), (g, b), (st), args...);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

static int run_quad_cells(int order, const double* sx, const double* sy, i64 nuc, i64 ncr, i64 nx, i64 top_row,
                          double Re, double* dxq, double* dyq, double* daq, cudaStream_t st) {
  dim3 g = lattice_grid(nx + 1, ncr + 1, 128);
  return launch_quad(order, k_quad_cells<2>, k_quad_cells<3>, k_quad_cells<4>, k_quad_cells<5>, g, dim3(128), st, sx, sy,
                     nuc, ncr, nx, top_row, Re, dxq, dyq, daq);
}

#ifndef OGG_BP_TR
#define OGG_BP_TR 6
#endif
#ifndef OGG_BP_TC
#define OGG_BP_TC 32
#endif
constexpr int kBpTileRows = OGG_BP_TR, kBpTileCols = OGG_BP_TC;  // 53 KB of samples per CTA: 4 CTAs/SM (8x32: 3 CTAs/SM, 3.6 % slower)
template <int N>
static int launch_bp_quad_fused(const BpCol* ct, const BpRow* rt, const double* divs, i64 nuc, i64 ncr, i64 nx, i64 ny, double lat0_bp,
                                double lon_bp, double rp, double Re, double* dxq, double* dyq, double* daq,
                                cudaStream_t st) {
  constexpr int TR = kBpTileRows, TC = kBpTileCols;
  static_assert(TR * TC <= kBpThreads, "one thread per cell");
  constexpr size_t smem = sizeof(double) * 2 * ((N - 1) * TR + 1) * ((N - 1) * (TC + 1));
  static bool configured[kMaxDevices] = {false};   // the opt-in to > 48 KB of dynamic shared memory is per device
  const int dev = current_device();
  if (!configured[dev]) {
    OGG_CUDA(cudaFuncSetAttribute(k_bp_quad_fused<N, TR, TC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev] = true;
  }
  if (nuc > 0x7fff0000LL || (N - 1) * ncr > 0x7fff0000LL) return OGG_ERR_ARG;   // the kernel's 32-bit sample numbers
  const dim3 g(cdiv_u(nx, TC), cdiv_u(ncr, TR));
  launch_begin(KID_BP_QUAD, st);
  k_bp_quad_fused<N, TR, TC><<<g, kBpThreads, smem, st>>>(ct, rt, divs, nuc, ncr, nx, ny, lat0_bp, lon_bp, rp, Re, dxq, dyq, daq);
  launch_end(st);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}
static int run_bp_quad_fused(int order, const BpCol* ct, const BpRow* rt, const double* divs, i64 nuc, i64 ncr, i64 nx, i64 ny, double lat0_bp,
                             double lon_bp, double rp, double Re, double* dxq, double* dyq, double* daq, cudaStream_t st) {
  if (order == 2) return launch_bp_quad_fused<2>(ct, rt, divs, nuc, ncr, nx, ny, lat0_bp, lon_bp, rp, Re, dxq, dyq, daq, st);
  if (order == 3) return launch_bp_quad_fused<3>(ct, rt, divs, nuc, ncr, nx, ny, lat0_bp, lon_bp, rp, Re, dxq, dyq, daq, st);
  if (order == 4) return launch_bp_quad_fused<4>(ct, rt, divs, nuc, ncr, nx, ny, lat0_bp, lon_bp, rp, Re, dxq, dyq, daq, st);
  return launch_bp_quad_fused<5>(ct, rt, divs, nuc, ncr, nx, ny, lat0_bp, lon_bp, rp, Re, dxq, dyq, daq, st);
}

// cap on the per-chunk sample scratch of the quadrature drivers (bytes of dx+dy samples)
static const i64 kQuadChunkBytes = (i64)1 << 30;

}  // namespace ogg

// ================================================================================================
// C ABI
// ================================================================================================
using namespace ogg;

extern "C" {

const char* ogg_version(void) { return "ogg_b200 0.1 (sm_100a, fp64, fmad=false)"; }

int ogg_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

long long ogg_launch_count(void) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  return g_launches;
}
int ogg_timing_enable(int on) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  g_timing = on < 0 ? 0 : (on > 2 ? 1 : on);
  return OGG_OK;
}
// Two kernel classes that run concurrently on different streams (the two cap quadratures): total time during which at
// least one bracketed launch of either class was in flight = length of the union of their [start, end] intervals
// (event times taken relative to the first bracketed launch).  Any number of launches per class (a cap's quadrature is
// chunked on large grids, a rank may hold several slabs per step).
int ogg_timing_pair_span(int id_a, int id_b, double* ms_total, long long* launches) {
  if (id_a < 0 || id_a >= KID_COUNT || id_b < 0 || id_b >= KID_COUNT || !ms_total || !launches) return OGG_ERR_ARG;
  const int dev = current_device();   // brackets of the current device only (event times of two devices do not compare)
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  std::vector<std::pair<double, double>> iv;
  const TimedLaunch* base = nullptr;
  for (auto& t : g_timed) {
    if (!t.ended || t.dev != dev || (t.id != id_a && t.id != id_b)) continue;
    if (!base) base = &t;
    OGG_CUDA(cudaEventSynchronize(t.e1));
    float s = 0.f, e = 0.f;
    if (&t != base) OGG_CUDA(cudaEventElapsedTime(&s, base->e0, t.e0));   // negative if it started before the base
    OGG_CUDA(cudaEventElapsedTime(&e, base->e0, t.e1));
    iv.emplace_back((double)s, (double)e);
  }
  std::sort(iv.begin(), iv.end());
  double tot = 0.0, lo = 0.0, hi = 0.0;
  bool open = false;
  for (auto& x : iv) {
    if (open && x.first <= hi) { if (x.second > hi) hi = x.second; continue; }
    if (open) tot += hi - lo;
    lo = x.first; hi = x.second; open = true;
  }
  if (open) tot += hi - lo;
  *ms_total = tot;
  *launches = (long long)iv.size();
  return OGG_OK;
}
int ogg_timing_reset(void) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  for (auto& t : g_timed) { g_event_pool[t.dev].push_back(t.e0); g_event_pool[t.dev].push_back(t.e1); }
  g_timed.clear();
  return OGG_OK;
}
int ogg_timing_kernels(void) { return KID_COUNT; }
const char* ogg_timing_name(int id) { return (id >= 0 && id < KID_COUNT) ? kKernelNames[id] : ""; }
int ogg_timing_get(int id, double* ms_total, long long* launches) {
  if (id < 0 || id >= KID_COUNT || !ms_total || !launches) return OGG_ERR_ARG;
  const int dev = current_device();
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  double tot = 0.0;
  long long n = 0;
  for (auto& t : g_timed) {
    if (!t.ended || t.dev != dev || t.id != id) continue;
    OGG_CUDA(cudaEventSynchronize(t.e1));
    float ms = 0.f;
    OGG_CUDA(cudaEventElapsedTime(&ms, t.e0, t.e1));
    tot += ms;
    ++n;
  }
  *ms_total = tot;
  *launches = n;
  return OGG_OK;
}

int ogg_sin(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_SIN>(in, out, n, s); }
int ogg_cos(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_COS>(in, out, n, s); }
int ogg_tan(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_TAN>(in, out, n, s); }
int ogg_arcsin(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_ASIN>(in, out, n, s); }
int ogg_arccos(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_ACOS>(in, out, n, s); }
int ogg_arctan(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_ATAN>(in, out, n, s); }
int ogg_exp(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_EXP>(in, out, n, s); }
int ogg_log(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_LOG>(in, out, n, s); }
int ogg_sinh(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_SINH>(in, out, n, s); }
int ogg_cosh(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_COSH>(in, out, n, s); }
int ogg_mod360(const double* in, double* out, int64_t n, void* s) { return launch_unary<F_MOD360>(in, out, n, s); }
int ogg_y_mercator(const double* phi, int64_t n, double R, double* out, void* stream) {
  if (n < 0 || (n > 0 && (!phi || !out))) return OGG_ERR_ARG;
  if (n == 0) return OGG_OK;
  OGG_LAUNCH(KID_ELEMENTWISE, (k_y_mercator), (flat_grid(n), 256), (S(stream)), phi, n, R, out);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}
int ogg_phi_mercator(const double* y, int64_t n, double y_first, double R, double* out, void* stream) {
  if (n < 0 || (n > 0 && !out)) return OGG_ERR_ARG;
  if (n == 0) return OGG_OK;
  OGG_LAUNCH(KID_ELEMENTWISE, (k_phi_mercator), (flat_grid(n), 256), (S(stream)), y, n, y_first, R, out);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}
int ogg_arctan2(const double* y, const double* x, double* out, int64_t n, void* stream) {
  if (n < 0 || (n > 0 && (!y || !x || !out))) return OGG_ERR_ARG;
  if (n == 0) return OGG_OK;
  OGG_LAUNCH(KID_ELEMENTWISE, (k_atan2), (flat_grid(n), 256), (S(stream)), y, x, out, n);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_band_fill_metrics_batch(const ogg_band_t* bands, int nbands, double Re, void* stream) {
  if (!bands || nbands < 1 || nbands > kMaxBands) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  BandBatch B;
  B.n = nbands; B.Re = Re; B.Re2 = Re * Re;
  i64 nxp_max = 0, words = 0;
  for (int k = 0; k < nbands; ++k) {
    const ogg_band_t& a = bands[k];
    if (!a.lam || !a.phi || a.nyp < 1 || a.nxp < 2) return OGG_ERR_ARG;
    words += 3 * (a.nxp + 2) + 4 * (a.nyp + 2);
  }
  double* tables;                                  // one block for every band's seven tables
  OGG_CUDA(ws.get(&tables, (size_t)words));
  for (int k = 0; k < nbands; ++k) {
    const ogg_band_t& a = bands[k];
    const i64 nyp = a.nyp, nxp = a.nxp;
    double *dxi = tables, *hx = dxi + nxp + 2, *xd = hx + nxp + 2, *cl = xd + nxp + 2, *cy = cl + nyp + 2, *dyr = cy + nyp + 2,
           *dsl = dyr + nyp + 2;
    tables = dsl + nyp + 2;
    const i64 m = nxp > nyp ? nxp : nyp;
    OGG_LAUNCH(KID_BAND_TABLES, (k_band_tables), (cdiv_u(m, 256), 256), (st), a.lam, a.phi, nyp, nxp, Re, dxi, hx, xd, cl, cy, dyr, dsl);
    OGG_LAUNCH_CHECK();
    BandDesc& D = B.b[k];
    D.lam = a.lam; D.phi = a.phi; D.dxi = dxi; D.hx = hx; D.xd = xd; D.cl = cl; D.cy = cy; D.dyr = dyr; D.dsl = dsl;
    D.x = a.x; D.y = a.y; D.dx = a.dx; D.dy = a.dy; D.area = a.area; D.angle = a.angle;
    D.nyp = nyp; D.nxp = nxp;
    D.nyp_v = (a.flags & OGG_BAND_SKIP_TOP_VERTEX_ROW) ? nyp - 1 : nyp;
    D.metrics = (a.flags & OGG_BAND_NO_METRICS) ? 0 : 1;
    if (nxp > nxp_max) nxp_max = nxp;
  }
  const i64 ncb = (nxp_max + kBandThreads - 1) / kBandThreads;
  const i64 R = kBandRows;
  i64 tiles = 0;
  for (int k = 0; k < nbands; ++k) { B.b[k].t_begin = tiles; tiles += (B.b[k].nyp + R - 1) / R; }
  if (tiles > 65535) return OGG_ERR_ARG;
  OGG_LAUNCH(KID_BAND_FILL, (k_band_fill), (dim3((unsigned)ncb, (unsigned)tiles), kBandThreads), (st), B);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_band_fill_metrics(const double* lam, const double* phi, int64_t nyp, int64_t nxp, double Re, int flags,
                          double* x, double* y, double* dx, double* dy, double* area, double* angle, void* stream) {
  ogg_band_t b;
  b.lam = lam; b.phi = phi; b.nyp = nyp; b.nxp = nxp; b.flags = flags;
  b.x = x; b.y = y; b.dx = dx; b.dy = dy; b.area = area; b.angle = angle;
  return ogg_band_fill_metrics_batch(&b, 1, Re, stream);
}

int ogg_metrics_midas(const double* x, const double* y, int64_t nyp, int64_t nxp, double Re, int areafix,
                      const double* halo_x, const double* halo_y, double* dx, double* dy, double* area, void* stream) {
  if (!x || !y || nyp < 1 || nxp < 2) return OGG_ERR_ARG;
  if ((halo_x == nullptr) != (halo_y == nullptr)) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  const i64 nrows = nyp + (halo_x ? 1 : 0);
  dim3 g(cdiv_u(nxp, 128), cdiv_u(nrows, kMidasRows));
  OGG_LAUNCH(KID_MIDAS, (k_midas), (g, 128), (st), x, y, nyp, nxp, Re, Re * Re, areafix, halo_x, halo_y, dx, dy, area);
  OGG_LAUNCH_CHECK();
  if (!areafix && area) {
    if (!dx || !dy) return OGG_ERR_ARG;
    const i64 ny = nrows - 1, nx = nxp - 1;
    if (ny > 0) {
      Scratch ws(st);
      double* top = nullptr;
      if (halo_x) {
        OGG_CUDA(ws.get(&top, nx));
        OGG_LAUNCH(KID_MIDAS, (k_dx_row), (cdiv_u(nx, 256), 256), (st), halo_x, halo_y, nx, Re, top);
        OGG_LAUNCH_CHECK();
      }
      OGG_LAUNCH(KID_MIDAS, (k_area_nofix), (flat_grid(ny * nx), 256), (st), dx, top, dy, ny, nx, area);
      OGG_LAUNCH_CHECK();
    }
  }
  return OGG_OK;
}

int ogg_angle_x(const double* x, const double* y, int64_t nyp, int64_t nxp, double* angle, void* stream) {
  if (!x || !y || !angle || nyp < 1 || nxp < 2) return OGG_ERR_ARG;
  OGG_LAUNCH(KID_ANGLE, (k_angle), (lattice_grid(nxp, nyp, 256), 256), (S(stream)), x, y, nyp, nxp, angle, pi180_pair());
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_bipolar_projection(const double* lamg, const double* phig, int64_t n, double lon_bp, double rp, int metrics_only,
                           double* lams, double* phis, double* hi, double* hj, void* stream) {
  if (n < 0 || (n > 0 && (!lamg || !phig))) return OGG_ERR_ARG;
  if (!metrics_only && (!lams || !phis)) return OGG_ERR_ARG;
  if (n == 0) return OGG_OK;
  OGG_LAUNCH(KID_BP_PROJECTION, (k_bp_projection), (flat_grid(n), 256), (S(stream)), lamg, phig, n, lon_bp, rp, metrics_only, lams, phis, hi, hj);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_bipolar_cap_mesh(int64_t Ni, int64_t Nj, int64_t j0, int64_t j1, double lat0_bp, double lon_bp, double rp,
                         double* lams, double* phis, double* h_i_inv, double* h_j_inv, void* stream) {
  if (Ni < 1 || Nj < 1 || j0 < 0 || j1 > Nj + 1 || j0 >= j1 || !lams || !phis) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  const i64 ncol = Ni + 1, nrow = j1 - j0;
  BpCol* ct; BpRow* rt;
  OGG_CUDA(ws.get(&ct, ncol)); OGG_CUDA(ws.get(&rt, nrow));
  const i64 m = ncol > nrow ? ncol : nrow;
  OGG_LAUNCH(KID_BP_TABLES, (k_bp_tables), (cdiv_u(m, 256), 256), (st), nullptr, 0, ncol, nullptr, j0, nrow, Ni, Nj, lat0_bp, lon_bp, rp, ct, rt);
  OGG_LAUNCH_CHECK();
  const double* pi180 = (!h_i_inv && !h_j_inv) ? pi180_pair() : nullptr;
  if (pi180)   // coordinates only: no metrics evaluated, division / square root without branches
    OGG_LAUNCH(KID_BP_MESH, (k_bp_lattice<2>), (lattice_grid(ncol, nrow, 256), 256), (st), ct, rt, nrow, ncol, Ni, Nj, j0, lat0_bp, lon_bp, rp, lams,
                                                                   phis, h_i_inv, h_j_inv, pi180);
  else
    OGG_LAUNCH(KID_BP_MESH, (k_bp_lattice<0>), (lattice_grid(ncol, nrow, 256), 256), (st), ct, rt, nrow, ncol, Ni, Nj, j0, lat0_bp, lon_bp, rp, lams,
                                                                   phis, h_i_inv, h_j_inv, pi180);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_bipolar_cap_ij_array(const double* i, int64_t ni, const double* j, int64_t nj, int64_t Ni, int64_t Nj_ncap,
                             double lat0_bp, double lon_bp, double rp, double* h_i_inv, double* h_j_inv, void* stream) {
  if (!i || !j || ni < 1 || nj < 1 || !h_i_inv || !h_j_inv) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  BpCol* ct; BpRow* rt;
  OGG_CUDA(ws.get(&ct, ni)); OGG_CUDA(ws.get(&rt, nj));
  const i64 m = ni > nj ? ni : nj;
  OGG_LAUNCH(KID_BP_TABLES, (k_bp_tables), (cdiv_u(m, 256), 256), (st), i, 0, ni, j, 0, nj, Ni, Nj_ncap, lat0_bp, lon_bp, rp, ct, rt);
  OGG_LAUNCH_CHECK();
  OGG_LAUNCH(KID_BP_SAMPLES, (k_bp_lattice<1>), (lattice_grid(ni, nj, 256), 256), (st), ct, rt, nj, ni, Ni, Nj_ncap, 0, lat0_bp, lon_bp, rp, h_i_inv,
                                                             h_j_inv, nullptr, nullptr, nullptr);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_bipolar_cap_metrics_quad(int order, int64_t nx, int64_t ny, int64_t j0, int64_t j1, double lat0_bp, double lon_bp,
                                 double rp, double Re, int flags, double* dxq, double* dyq, double* daq, void* stream) {
  QuadNodes q;
  if (quad_nodes_host(order, &q)) return OGG_ERR_UNCODED_ORDER;
  if (nx < 1 || ny < 1 || j0 < 0 || j1 > ny || j0 >= j1 || !dxq || !dyq || !daq) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  const bool top = j1 == ny && !(flags & OGG_QUAD_NO_TOP_DX_ROW);
  const i64 ncr = j1 - j0;
  const i64 nuc = (order - 1) * nx + 1;
  const i64 nur = (order - 1) * ncr + 1 + (top ? 1 : 0);  // + the exact pole row
  Scratch ws(st);
  BpCol* ct; BpRow* rt; double* divs;
  OGG_CUDA(ws.get(&ct, nuc)); OGG_CUDA(ws.get(&rt, nur)); OGG_CUDA(ws.get(&divs, 6));
  const i64 m = nuc > nur ? nuc : nur;
  OGG_LAUNCH(KID_BP_TABLES, (k_bp_quad_tables), (cdiv_u(m, 256), 256), (st), q, order, nx, ny, j0, j1, nuc, nur, lat0_bp, lon_bp, rp, ct, rt, divs);
  OGG_LAUNCH_CHECK();
  int rc = run_bp_quad_fused(order, ct, rt, divs, nuc, ncr, nx, ny, lat0_bp, lon_bp, rp, Re, dxq, dyq, daq, st);
  if (rc) return rc;
  if (top) {
    const dim3 g(cdiv_u(nx, 128)), b(128);
    rc = launch_quad(order, k_bp_pole_row<2>, k_bp_pole_row<3>, k_bp_pole_row<4>, k_bp_pole_row<5>, g, b, st,
                     (const BpCol*)ct, (const BpRow*)rt, nur - 1, (i64)nx, lon_bp, rp, Re, dxq + ncr * nx);
    if (rc) return rc;
  }
  return OGG_OK;
}

int ogg_quad_positions(int n, double* a, double* b) {
  QuadNodes q;
  if (quad_nodes_host(n, &q)) return OGG_ERR_UNCODED_ORDER;
  for (int k = 0; k < n; ++k) { a[k] = q.a[k]; b[k] = q.b[k]; }
  return OGG_OK;
}
int ogg_quad_average(const double* y, int64_t m, int n, double* out, void* stream) {
  if (n < 2 || n > 5) return OGG_ERR_UNCODED_ORDER;
  if (m < 1 || !y || !out) return OGG_ERR_ARG;
  return launch_quad(n, k_quad_avg<2>, k_quad_avg<3>, k_quad_avg<4>, k_quad_avg<5>, dim3(cdiv_u(m, 128)), dim3(128),
                     S(stream), y, (i64)m, out);
}
int ogg_quad_average_2d(const double* y, int64_t m, int n, double* out, void* stream) {
  if (n < 2 || n > 5) return OGG_ERR_UNCODED_ORDER;
  if (m < 1 || !y || !out) return OGG_ERR_ARG;
  return launch_quad(n, k_quad_avg2d<2>, k_quad_avg2d<3>, k_quad_avg2d<4>, k_quad_avg2d<5>, dim3(cdiv_u(m, 128)),
                     dim3(128), S(stream), y, (i64)m, out);
}

int ogg_dp_constants(double lat0, double lam_pole, double r_pole, double* out3, void* stream) {
  if (!out3) return OGG_ERR_ARG;
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_constants), (1, 32), (S(stream)), lat0, lam_pole, r_pole, out3);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_monotonic_bounding(double* x, int64_t nj, int64_t ni, double x_0, void* stream) {
  if (!x || nj < 1 || ni < 1) return OGG_ERR_ARG;
  const i64 cap = (i64)num_sms() * 8;
  OGG_LAUNCH(KID_MONOTONIC, (k_monotonic), ((unsigned)(nj < cap ? nj : cap), kScanThreads), (S(stream)), x, nj, ni, x_0, nullptr);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_dp_projection(const double* lon, const double* lat, int64_t nj, int64_t ni, double z0_re, double z0_im,
                      double r_joint, double* lam, double* phi, void* stream) {
  if (!lon || !lat || !lam || !phi || nj < 1 || ni < 1) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  double* cst;
  OGG_CUDA(ws.get(&cst, 3));
  OGG_LAUNCH(KID_DP_SETUP, (k_set3), (1, 32), (st), r_joint, z0_re, z0_im, cst);
  OGG_LAUNCH_CHECK();
  OGG_LAUNCH(KID_DP_RAW, (k_dp_projection_raw), (flat_grid(nj * ni), 256), (st), lon, lat, nj * ni, cst, lam, phi);
  OGG_LAUNCH_CHECK();
  const i64 cap = (i64)num_sms() * 8;
  OGG_LAUNCH(KID_MONOTONIC, (k_monotonic), ((unsigned)(nj < cap ? nj : cap), kScanThreads), (st), lam, nj, ni, 0.0, lon);  // x_0 = lon_grid[0,0] (:463)
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

}  // extern "C"

namespace ogg {
// single shifted mesh on index vectors: tables -> raw lattice -> unwrap.  cst: device {r_joint, z0}
static int dp_mesh_dev(const double* iv, i64 i0, i64 nil, double ishift, const double* jv, i64 j0, i64 njl, double jshift,
                       i64 ni, i64 nj,
                       double lon0, double lat0, const double* cst, double* lams, double* phis, cudaStream_t st) {
  Scratch ws(st);
  Cx* ct; double* rt; double* x0;
  OGG_CUDA(ws.get(&ct, nil)); OGG_CUDA(ws.get(&rt, njl)); OGG_CUDA(ws.get(&x0, 12));
  const i64 m = nil > njl ? nil : njl;
  OGG_LAUNCH(KID_DP_TABLES, (k_dp_tables), (cdiv_u(m, 256), 256), (st), iv, i0, nil, ishift, jv, j0, njl, jshift, ni, nj, lon0, lat0, cst, ct, rt);
  OGG_LAUNCH_CHECK();
  OGG_LAUNCH(KID_DP_RAW, (k_dp_lattice_raw), (lattice_grid(nil, njl, 256), 256), (st), ct, rt, njl, nil, cst, lams, phis, pi180_pair());
  OGG_LAUNCH_CHECK();
  // x_0 = lon_grid[0,0] = lon0 + ((i[0]+shift)*360)/ni : reuse k_dp_x0<1> slots (kind 0 = plus shift, kind 2 = none)
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_x0<1>), (1, 32), (st), iv, (double)i0, ishift, ni, lon0, x0);
  OGG_LAUNCH_CHECK();
  const i64 cap = (i64)num_sms() * 8;
  OGG_LAUNCH(KID_MONOTONIC, (k_monotonic), ((unsigned)(njl < cap ? njl : cap), kScanThreads), (st), lams, njl, nil, 0.0, ishift != 0.0 ? x0 + 0 : x0 + 2);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}
}  // namespace ogg

extern "C" {

int ogg_dp_mesh(const double* i, int64_t nil, const double* j, int64_t njl, int64_t ni, int64_t nj, double lon0, double lat0,
                double lam_pole, double r_pole, double* lams, double* phis, void* stream) {
  if (!i || !j || nil < 1 || njl < 1 || !lams || !phis) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  double* cst;
  OGG_CUDA(ws.get(&cst, 3));
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_constants), (1, 32), (st), lat0, lam_pole, r_pole, cst);
  OGG_LAUNCH_CHECK();
  return dp_mesh_dev(i, 0, nil, 0.0, j, 0, njl, 0.0, ni, nj, lon0, lat0, cst, lams, phis, st);
}

int ogg_dp_cap_mesh(int64_t Ni, int64_t Nj, int64_t j0, int64_t j1, double lon0, double lat0, double lam_pole, double r_pole,
                    double* lams, double* phis, void* stream) {
  if (Ni < 1 || Nj < 1 || j0 < 0 || j1 > Nj + 1 || j0 >= j1 || !lams || !phis) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  double* cst;
  OGG_CUDA(ws.get(&cst, 3));
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_constants), (1, 32), (st), lat0, lam_pole, r_pole, cst);
  OGG_LAUNCH_CHECK();
  return dp_mesh_dev(nullptr, 0, Ni + 1, 0.0, nullptr, j0, j1 - j0, 0.0, Ni, Nj, lon0, lat0, cst, lams, phis, st);
}

}  // extern "C"

namespace ogg {
// ------------------------------------------------------------------------------------------------
// ogg_dp_cap_mesh_noalloc: the displaced-pole cap mesh for hosts that must not allocate.
//
// Who that is: an ocean model's I/O or coupler layer that hands this library two output arrays from its own arena and
// has a fixed device-memory budget (every byte of HBM is accounted for at start-up, so a library that grows a memory
// pool behind its back breaks the budget); a thread that is inside a region where the CUDA allocator may not be entered
// (it takes a driver-wide lock, and a first stream-ordered allocation may synchronise the device); a process whose
// device memory is handed out by another runtime (an MPI / OpenACC / Fortran host) that forbids foreign cudaMalloc calls.
// ogg_dp_cap_mesh takes its three work tables (the column table, the row table, a handful of constants) from the
// library's cudaMallocFromPoolAsync pool; this variant keeps them in the array below instead, which is part of the
// loaded module: every device that loads libogg_b200.so holds these 2 MiB from the first use of the library on that
// device until the process exits, whether or not this entry point is ever called.
//
// The name g_noalloc_scratch is only meaningful in device code.  Host code gets the array's device address with
// cudaGetSymbolAddress, once per device and with that device current (noalloc_scratch below), and hands that pointer to
// the kernels; passing the bare name from the host would hand them a host-side shadow address.
// ------------------------------------------------------------------------------------------------
constexpr i64 kNoAllocDoubles = OGG_NOALLOC_SCRATCH_DOUBLES;   // include/ogg_b200.h: 262144 doubles = 2 MiB
constexpr i64 kNoAllocHeader = 32;                             // constants: [0,3) r_joint, z0; [4,6) pi/180 pair; [8,20) x_0 slots
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocDoubles];

// Device address of g_noalloc_scratch on the CURRENT device (nullptr + *err on failure).  The first call on a device
// asks the runtime for the address (which also makes sure the module is loaded there); later calls read the cached value.
static double* noalloc_scratch(cudaError_t* err) {
  static double* ptrs[kMaxDevices] = {nullptr};
  static std::mutex mu;
  *err = cudaSuccess;
  int dev = -1;
  *err = cudaGetDevice(&dev);
  if (*err != cudaSuccess) return nullptr;
  if (dev < 0 || dev >= kMaxDevices) { *err = cudaErrorInvalidDevice; return nullptr; }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    *err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (*err != cudaSuccess) return nullptr;
    ptrs[dev] = static_cast<double*>(p);
  }
  return ptrs[dev];
}
}  // namespace ogg

extern "C" {

int ogg_dp_cap_mesh_noalloc(int64_t Ni, int64_t Nj, int64_t j0, int64_t j1, double lon0, double lat0, double lam_pole,
                            double r_pole, double* lams, double* phis, void* stream) {
  if (Ni < 1 || Nj < 1 || j0 < 0 || j1 > Nj + 1 || j0 >= j1 || !lams || !phis) return OGG_ERR_ARG;
  const i64 nil = Ni + 1, njl = j1 - j0;
  // the whole request must fit the fixed array: header + one complex number per mesh column + one real per mesh row
  if (nil > kNoAllocDoubles || njl > kNoAllocDoubles || kNoAllocHeader + 2 * nil + njl > kNoAllocDoubles)
    return OGG_ERR_TOO_LARGE;
  cudaError_t err;
  double* ws = noalloc_scratch(&err);
  if (!ws) return OGG_ERR_CUDA - (int)err;
  cudaStream_t st = S(stream);
  // Every slot read below is written earlier in this same call, in stream order: nothing has to start at zero, and
  // whatever the previous call left behind is never looked at.
  double* cst = ws;            // r_joint, Re z_0, Im z_0
  double* pi180 = ws + 4;      // pi/180 and its refined reciprocal
  double* x0 = ws + 8;         // k_dp_x0<1> writes 4 slots; slot 2 is the unshifted x_0 = lon_grid[0,0]
  Cx* ct = reinterpret_cast<Cx*>(ws + kNoAllocHeader);
  double* rt = ws + kNoAllocHeader + 2 * nil;
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_constants), (1, 32), (st), lat0, lam_pole, r_pole, cst);
  OGG_LAUNCH_CHECK();
  OGG_LAUNCH(KID_DP_SETUP, (k_store_pi180), (1, 32), (st), pi180);
  OGG_LAUNCH_CHECK();
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_x0<1>), (1, 32), (st), (const double*)nullptr, 0.0, 0.0, Ni, lon0, x0);
  OGG_LAUNCH_CHECK();
  const i64 m = nil > njl ? nil : njl;
  OGG_LAUNCH(KID_DP_TABLES, (k_dp_tables), (cdiv_u(m, 256), 256), (st), (const double*)nullptr, (i64)0, nil, 0.0,
             (const double*)nullptr, (i64)j0, njl, 0.0, (i64)Ni, (i64)Nj, lon0, lat0, (const double*)cst, ct, rt);
  OGG_LAUNCH_CHECK();
  OGG_LAUNCH(KID_DP_RAW, (k_dp_lattice_raw), (lattice_grid(nil, njl, 256), 256), (st), (const Cx*)ct, (const double*)rt,
             njl, nil, (const double*)cst, lams, phis, (const double*)pi180);
  OGG_LAUNCH_CHECK();
  const i64 cap = (i64)num_sms() * 8;
  OGG_LAUNCH(KID_MONOTONIC, (k_monotonic), ((unsigned)(njl < cap ? njl : cap), kScanThreads), (st), lams, njl, nil, 0.0,
             (const double*)(x0 + 2));
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_great_arc_distance(const double* j0, const double* i0, const double* j1, const double* i1, int64_t njl, int64_t nil,
                           int64_t nx, int64_t ny, double lon0, double lat0, double lon_dp, double r_dp, double* out,
                           void* stream) {
  if (!j0 || !i0 || !j1 || !i1 || !out || njl < 1 || nil < 1) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  const i64 n = njl * nil;
  double *cst, *l0, *p0, *l1, *p1;
  OGG_CUDA(ws.get(&cst, 3));
  OGG_CUDA(ws.get(&l0, n)); OGG_CUDA(ws.get(&p0, n)); OGG_CUDA(ws.get(&l1, n)); OGG_CUDA(ws.get(&p1, n));
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_constants), (1, 32), (st), lat0, lon_dp, r_dp, cst);
  OGG_LAUNCH_CHECK();
  int rc = dp_mesh_dev(i0, 0, nil, 0.0, j0, 0, njl, 0.0, nx, ny, lon0, lat0, cst, l0, p0, st);
  if (rc) return rc;
  rc = dp_mesh_dev(i1, 0, nil, 0.0, j1, 0, njl, 0.0, nx, ny, lon0, lat0, cst, l1, p1, st);
  if (rc) return rc;
  OGG_LAUNCH(KID_ELEMENTWISE, (k_haversine), (flat_grid(n), 256), (st), l0, p0, l1, p1, n, out);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

}  // extern "C"

namespace ogg {
// tables for the fused h-sample kernel: ct [ncol][1+2nv], rt [nrow][1+2nv] (centre, +eps, -eps, +2eps, -2eps, ...)
// Everything k_dp_hsamples needs, in one launch (it used to be six short launches in front of the critical path):
// the projection constants (:494-495), the unique sample indices (:573-581), the column / row tables of e' and r at
// the centre and at +-eps, +-2eps.. (:538-559), the x_0 of every shifted mesh (:463, :479) and the cleared look-back
// descriptors + ticket counter.  Every thread derives the constants itself (a tan and a sincos) instead of waiting
// for another kernel to publish them; thread 0 stores them for the kernels that follow.
__global__ void __launch_bounds__(256) k_dp_quad_setup(QuadNodes q, int order, i64 nx, i64 ny, i64 c0, i64 c1, int do_cols,
                                                       double lon0, double lat0, double lam_pole, double r_pole, int nv,
                                                       double eps, double* __restrict__ cst, Cx* __restrict__ ct,
                                                       double* __restrict__ rt, double* __restrict__ x0v,
                                                       unsigned long long* __restrict__ desc, i64 ndesc,
                                                       int* __restrict__ tok) {
  const i64 t = blockIdx.x * (i64)blockDim.x + threadIdx.x;
  DpConst K;
  K.r_joint = p_tan((90 + lat0) * kPi180);                                                 // :494
  K.z0 = cmul(Cx{r_pole, 0.0}, cis_np(lam_pole * kPi180));                                 // :495
  if (t == 0) {
    cst[0] = K.r_joint; cst[1] = K.z0.re; cst[2] = K.z0.im;
    MathFast::store_divisor(cst + 3, kPi180);                                              // see MathFast::Divisor
  }
  const int nt = 1 + 2 * nv;
  const i64 nuc = (order - 1) * nx + 1, nur = (order - 1) * (c1 - c0) + 1;
  const bool col = do_cols && t < nuc, row = t < nur;
  const double ic = col ? sample_index_at(q, order, 0, nx, nx, 0, t) : 0.0;                // :578-581
  const double jc = row ? sample_index_at(q, order, c0, c1, ny, 0, t) : 0.0;               // :573-576
  // tok: dp_vertex_fast's preconditions, per table column [0, nuc) and per table row [nuc, nuc + nur)
  bool col_ok = true, row_ok = dp_mid(K.r_joint);
  for (int k = 0; k < nt; ++k) {
    const int lvl = (k + 1) / 2;  // 0 centre, 1: eps, 2: 2eps, 3: 3eps
    const double s = lvl == 1 ? eps : (lvl == 2 ? 2.0 * eps : 3.0 * eps);
    const bool plus = (k & 1) == 1;
    if (col) {
      const double idx = k == 0 ? ic : (plus ? ic + s : ic - s);                           // :538,:541,:544
      const Cx e = dp_col(lon0 + (idx * 360.0) / (double)nx, K.z0);
      ct[t * nt + k] = e;
      col_ok = col_ok && dp_mid(e.re) && dp_mid(e.im);
    }
    if (row) {
      const double idx = k == 0 ? jc : (plus ? jc + s : jc - s);                           // :553,:556,:559
      const double a = -90.0, b = lat0;
      const double rv = dp_row(a + (idx * (b - a)) / (double)ny, K.r_joint);
      rt[t * nt + k] = rv;
      row_ok = row_ok && dp_mid(rv);
    }
  }
  if (col) tok[t] = col_ok ? 1 : 0;
  if (row) tok[nuc + t] = row_ok ? 1 : 0;
  if (do_cols && t < 4 * nv) {                                                             // x_0 of mesh t (k_dp_x0)
    const int lvl = (int)t / 4, kind = (int)t % 4;
    const double s = lvl == 0 ? eps : (lvl == 1 ? 2.0 * eps : 3.0 * eps);
    const double first = sample_index_at(q, order, 0, nx, nx, 0, 0);
    double idx = first;
    if (kind == 0) idx = first + s;
    if (kind == 1) idx = first - s;
    x0v[t] = lon0 + (idx * 360.0) / (double)nx;
  }
  for (i64 e = t; e < ndesc; e += (i64)gridDim.x * blockDim.x) desc[e] = 0ull;
}
#ifndef OGG_DP_MINB   // tuning knobs of the order-4 instantiation (the one main() uses).  4 CTAs/SM = 64 registers: the
#define OGG_DP_MINB 4 // kernel is issue-bound (2F + O), so the registers buy more (fewer re-derived addresses
#endif                // and re-loaded coefficients) than a fifth resident CTA: 0.4060 -> 0.3938 ms

#ifndef OGG_DP_UNR
#define OGG_DP_UNR 4
#endif
#ifndef OGG_DP_UNR_SA  // the SMALLAT instantiation's loop body is shorter: all 8 meshes unrolled fit the instruction cache
#define OGG_DP_UNR_SA 8
#endif
#ifndef OGG_DP_THREADS
#define OGG_DP_THREADS 256
#endif
// nitem descriptors + the ticket counter, all zero before the launch (k_dp_quad_setup clears them)
static i64 dp_hsamples_items(int nv, i64 nrow, i64 ncol) {
  const int threads = nv == 3 ? 128 : OGG_DP_THREADS;
  return nrow * ((ncol + threads - 2) / (threads - 1));
}
template <int NV, int THREADS, int MINB, int UNR, bool SMALLAT = false>
static int dp_hsamples_t(const Cx* ct, const double* rt, i64 nrow, i64 ncol, const double* x0v, const double* cst,
                         double reps, double* hx, double* hy, unsigned long long* desc, const int* tok, cudaStream_t st) {
  const i64 nitem = nrow * ((ncol + THREADS - 2) / (THREADS - 1));
  // the kernel numbers items, columns and tiles in 32 bits (t * TW, (int)ncol, (unsigned)nrow): bound each of them,
  // not only their product (few rows of very many columns would pass the item count and wrap the column number)
  if (nitem > 0x7fffffffLL || ncol > 0x7fff0000LL || nrow > 0x7fff0000LL) return OGG_ERR_ARG;
  launch_begin(KID_DP_HSAMPLES, st);
  k_dp_hsamples<NV, THREADS, MINB, UNR, SMALLAT><<<(unsigned)nitem, THREADS, 0, st>>>(ct, rt, nrow, ncol, x0v, cst, reps, hx, hy, desc, tok);
  launch_end(st);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}
static int dp_hsamples(int nv, const Cx* ct, const double* rt, i64 nrow, i64 ncol, const double* x0v, const double* cst,
                       double reps, double* hx, double* hy, unsigned long long* desc, const int* tok, bool small_at, cudaStream_t st) {
#ifndef OGG_NO_SMALLAT
  if (nv == 2 && small_at)
    return dp_hsamples_t<2, OGG_DP_THREADS, OGG_DP_MINB, OGG_DP_UNR_SA, true>(ct, rt, nrow, ncol, x0v, cst, reps, hx, hy, desc, tok, st);
#endif
  if (nv == 1) return dp_hsamples_t<1, OGG_DP_THREADS, 4, 4>(ct, rt, nrow, ncol, x0v, cst, reps, hx, hy, desc, tok, st);
  if (nv == 2) return dp_hsamples_t<2, OGG_DP_THREADS, OGG_DP_MINB, OGG_DP_UNR>(ct, rt, nrow, ncol, x0v, cst, reps, hx, hy, desc, tok, st);
  return dp_hsamples_t<3, 128, 4, 4>(ct, rt, nrow, ncol, x0v, cst, reps, hx, hy, desc, tok, st);
}
}  // namespace ogg

extern "C" {

int ogg_numerical_h(int direction, const double* j, int64_t njl, const double* i, int64_t nil, int64_t nx, int64_t ny,
                    double lon0, double lat0, double lon_dp, double r_dp, double eps, int order, double* out, void* stream) {
  if (!j || !i || !out || njl < 1 || nil < 1 || (direction != 0 && direction != 1)) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  // the reference evaluates ds2 (and ds4, ds6) before it finds the order uncoded (:538-547); only the status matters here
  const int nds = order == 2 ? 1 : (order == 4 ? 2 : (order == 6 ? 3 : 0));
  if (nds == 0) return OGG_ERR_ORDER_NOT_CODED;
  Scratch ws(st);
  const i64 n = njl * nil;
  double *cst, *l0, *p0, *l1, *p1, *ds[3] = {nullptr, nullptr, nullptr};
  OGG_CUDA(ws.get(&cst, 3));
  OGG_CUDA(ws.get(&l0, n)); OGG_CUDA(ws.get(&p0, n)); OGG_CUDA(ws.get(&l1, n)); OGG_CUDA(ws.get(&p1, n));
  for (int m = 0; m < nds; ++m) OGG_CUDA(ws.get(&ds[m], n));
  OGG_LAUNCH(KID_DP_SETUP, (k_dp_constants), (1, 32), (st), lat0, lon_dp, r_dp, cst);
  OGG_LAUNCH_CHECK();
  for (int m = 0; m < nds; ++m) {
    const double s = m == 0 ? eps : (m == 1 ? 2.0 * eps : 3.0 * eps);
    int rc;
    if (direction == 0) {
      rc = dp_mesh_dev(i, 0, nil, s, j, 0, njl, 0.0, nx, ny, lon0, lat0, cst, l0, p0, st);
      if (!rc) rc = dp_mesh_dev(i, 0, nil, -s, j, 0, njl, 0.0, nx, ny, lon0, lat0, cst, l1, p1, st);
    } else {
      rc = dp_mesh_dev(i, 0, nil, 0.0, j, 0, njl, s, nx, ny, lon0, lat0, cst, l0, p0, st);
      if (!rc) rc = dp_mesh_dev(i, 0, nil, 0.0, j, 0, njl, -s, nx, ny, lon0, lat0, cst, l1, p1, st);
    }
    if (rc) return rc;
    OGG_LAUNCH(KID_ELEMENTWISE, (k_haversine), (flat_grid(n), 256), (st), l0, p0, l1, p1, n, ds[m]);
    OGG_LAUNCH_CHECK();
  }
  OGG_LAUNCH(KID_ELEMENTWISE, (k_fd_combine), (flat_grid(n), 256), (st), ds[0], ds[1], ds[2], order, 1.0 / eps, n, out);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_dp_metrics_quad(int order, int64_t nx, int64_t ny, int64_t j0, int64_t j1, double lon0, double lat0, double lon_dp,
                        double r_dp, double Re, int flags, double* dxq, double* dyq, double* daq, void* stream) {
  QuadNodes q;
  if (quad_nodes_host(order, &q)) return OGG_ERR_UNCODED_ORDER;      // quad_positions raises first (:567)
  if (order != 2 && order != 4) return OGG_ERR_ORDER_NOT_CODED;      // numerical_hi(order=order) (:583)
  if (nx < 1 || ny < 1 || j0 < 0 || j1 > ny || j0 >= j1 || !dxq || !dyq || !daq) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  const int nv = order / 2, nt = 1 + 2 * nv;
  const double eps = 1e-3;                                           // :583-584
  const i64 nuc = (order - 1) * nx + 1;
  i64 rows_per_chunk = kQuadChunkBytes / (16 * nuc * (order - 1));
  if (rows_per_chunk < 1) rows_per_chunk = 1;
  const i64 max_cr = rows_per_chunk < (j1 - j0) ? rows_per_chunk : (j1 - j0);
  const i64 max_nur = (order - 1) * max_cr + 2;
  Scratch ws(st);
  double *cst, *rt, *x0v, *sx, *sy;
  Cx* ct;
  unsigned long long* desc;
  const i64 max_items = dp_hsamples_items(nv, max_nur, nuc);
  OGG_CUDA(ws.get(&cst, 5));
  OGG_CUDA(ws.get(&ct, nuc * nt)); OGG_CUDA(ws.get(&rt, max_nur * nt)); OGG_CUDA(ws.get(&x0v, 12));
  OGG_CUDA(ws.get(&sx, max_nur * nuc)); OGG_CUDA(ws.get(&sy, max_nur * nuc));
  OGG_CUDA(ws.get(&desc, (size_t)max_items + 1));
  int* tok;
  OGG_CUDA(ws.get(&tok, nuc + max_nur));
  int rc;
  for (i64 c0 = j0; c0 < j1; c0 += rows_per_chunk) {
    const i64 c1 = c0 + rows_per_chunk < j1 ? c0 + rows_per_chunk : j1;
    const bool top = c1 == ny && !(flags & OGG_QUAD_NO_TOP_DX_ROW);
    const i64 nur = (order - 1) * (c1 - c0) + 1;                                           // closing row = row ny when top
    const i64 ndesc = dp_hsamples_items(nv, nur, nuc) + 1;
    const i64 nthr = c0 == j0 ? (nuc > nur ? nuc : nur) : nur;
    OGG_LAUNCH(KID_DP_SETUP, (k_dp_quad_setup), (cdiv_u(nthr, 256), 256), (st), q, order, nx, ny, c0, c1, c0 == j0 ? 1 : 0,
               lon0, lat0, lon_dp, r_dp, nv, eps, cst, ct, rt, x0v, desc, ndesc, tok);
    OGG_LAUNCH_CHECK();
    rc = dp_hsamples(nv, ct, rt, nur, nuc, x0v, cst, 1.0 / eps, sx, sy, desc, tok, lat0 < -67.0 && fabs(r_dp) <= 0.5, st);
    if (rc) return rc;
    rc = run_quad_cells(order, sx, sy, nuc, c1 - c0, nx, top ? nur - 1 : -1, Re, dxq + (c0 - j0) * nx,
                        dyq + (c0 - j0) * (nx + 1), daq + (c0 - j0) * nx, st);
    if (rc) return rc;
  }
  return OGG_OK;
}

int ogg_sum_strided(const double* a, int64_t n, int64_t stride, double* out, void* stream) {
  if (!a || !out || n < 0 || stride < 1) return OGG_ERR_ARG;
  cudaStream_t st = S(stream);
  Scratch ws(st);
  double* partial;
  OGG_CUDA(ws.get(&partial, kRedBlocks));
  OGG_LAUNCH(KID_REDUCE, (k_sum_partial), (kRedBlocks, kRedThreads), (st), a, n, stride, partial);
  OGG_LAUNCH_CHECK();
  OGG_LAUNCH(KID_REDUCE, (k_sum_final), (1, kRedThreads), (st), partial, kRedBlocks, out);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_checksum_accumulate(const double* a, int64_t n, uint64_t* acc, void* stream) {
  if (!acc || n < 0 || (n > 0 && !a)) return OGG_ERR_ARG;
  if (n == 0) return OGG_OK;
  cudaStream_t st = S(stream);
  const i64 want = (n + kRedThreads - 1) / kRedThreads;
  const unsigned g = (unsigned)(want < (i64)num_sms() * 8 ? want : (i64)num_sms() * 8);
  OGG_LAUNCH(KID_REDUCE, (k_checksum), (g, kRedThreads), (st), a, (i64)n, reinterpret_cast<unsigned long long*>(acc));
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_bswap64(const void* in, void* out, int64_t n, void* stream) {
  if (n < 0 || (n > 0 && (!in || !out))) return OGG_ERR_ARG;
  if (n == 0) return OGG_OK;
  OGG_LAUNCH(KID_ELEMENTWISE, (k_bswap64), (flat_grid(n, 16), 256), (S(stream)), static_cast<const unsigned long long*>(in),
             static_cast<unsigned long long*>(out), (i64)n);
  OGG_LAUNCH_CHECK();
  return OGG_OK;
}

int ogg_selftest_divsqrt(int64_t n, uint64_t seed, int mode, uint64_t* counts3_host) {
  if (!counts3_host || n < 1 || mode < 0 || mode > 4) return OGG_ERR_ARG;
  unsigned long long* dev;
  OGG_CUDA(cudaMalloc(&dev, 24));
  OGG_CUDA(cudaMemset(dev, 0, 24));
  if (mode == 4) k_selftest_trig<<<num_sms() * 8, 256>>>(n, (unsigned long long)seed, dev);
  else k_selftest_divsqrt<<<num_sms() * 8, 256>>>(n, (unsigned long long)seed, mode, dev);
  cudaError_t e = cudaMemcpy(counts3_host, dev, 24, cudaMemcpyDeviceToHost);
  cudaFree(dev);
  OGG_CUDA(e);
  return OGG_OK;
}

int ogg_guard_audit(uint64_t* counts2_host, int reset) {
  if (!counts2_host) return OGG_ERR_ARG;
  counts2_host[0] = counts2_host[1] = 0;
#ifdef OGG_GUARD_AUDIT
  OGG_CUDA(cudaDeviceSynchronize());
  unsigned long long h[2] = {0ull, 0ull};
  OGG_CUDA(cudaMemcpyFromSymbol(h, g_guard_audit, sizeof(h)));
  counts2_host[0] = h[0]; counts2_host[1] = h[1];
  if (reset) { h[0] = h[1] = 0ull; OGG_CUDA(cudaMemcpyToSymbol(g_guard_audit, h, sizeof(h))); }
#else
  (void)reset;
#endif
  return OGG_OK;
}

int ogg_fp64_issue_rate(int64_t iters, double* ginstr_per_s_host) {
  if (!ginstr_per_s_host || iters < 1) return OGG_ERR_ARG;
  double* sink;
  OGG_CUDA(cudaMalloc(&sink, 8));
  cudaEvent_t e0, e1;
  OGG_CUDA(cudaEventCreate(&e0)); OGG_CUDA(cudaEventCreate(&e1));
  const int blocks = num_sms() * 8;
  k_fp64_rate<<<blocks, 256>>>(iters / 8 + 1, 1.0, sink);  // warm-up
  OGG_CUDA(cudaEventRecord(e0));
  k_fp64_rate<<<blocks, 256>>>(iters, 1.0, sink);
  OGG_CUDA(cudaEventRecord(e1));
  OGG_CUDA(cudaEventSynchronize(e1));
  float ms = 0;
  OGG_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  const double instr = (double)blocks * 256.0 * (double)iters * 16.0;
  *ginstr_per_s_host = instr / (ms * 1e-3) / 1e9;
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
  return OGG_OK;
}

}  // extern "C"

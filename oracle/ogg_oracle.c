/*
 * ogg_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * Plain-C restatement of the supergrid hot path of ocean_model_grid_generator, function by
 * function, in the reference's operation order.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this library; the product path
 * (ocean_model_grid_generator_b200/) never does.
 *
 * Citations "OGG:n" are /root/reference/ocean_grid_generator.py line numbers.
 * Trigonometry is the repo's portable-trig spec, restated here independently of the
 * CUDA header; the numpy twin is oracle/numpypi/numpypi_series.py.
 * PARITY STATUS: pinned against the UNMODIFIED reference run in the build container with the
 * numpy twin standing in for the un-vendored `numpypi` (tests/golden/, made by
 * oracle/make_golden.py).  Unpinned against upstream numpypi's own polynomials (absent, unpinned
 * upstream), hence against the reference's whole-file md5/sha256 values.
 *
 * Build: gcc -O2 -fopenmp -ffp-contract=off -fno-fast-math -fPIC -shared (oracle/Makefile).
 * Every arithmetic operation below is one correctly rounded IEEE-754 binary64 operation.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef int64_t i64;

static const double PI_180 = 3.141592653589793 / 180.0; /* OGG:13 */
static const double PI_D = 3.141592653589793;
static const double HUGE_ = 1.0e30;                      /* OGG:16 */

/* ======================================================================================= */
/* Portable elementary functions (spec: DESIGN.md; twin: oracle/numpypi/numpypi_series.py)  */
/* ======================================================================================= */
static const double TWO_OVER_PI = 0x1.45f306dc9c883p-1;
static const double RP1 = 0x1.921fb54400000p+0, RP2 = 0x1.0b4611a600000p-34, RP3 = 0x1.3198a2e037073p-69;
static const double PIO2_HI = 0x1.921fb54442d18p+0, PIO2_LO = 6.12323399573676603587e-17;
static const double PI_LO = 1.2246467991473532e-16;
static const double SC[6] = {-1.66666666666666324348e-01, 8.33333333332248946124e-03,
                             -1.98412698298579493134e-04, 2.75573137070700676789e-06,
                             -2.50507602534068634195e-08, 1.58969099521155010221e-10};
static const double CC[6] = {4.16666666666666019037e-02, -1.38888888888741095749e-03,
                             2.48015872894767294178e-05, -2.75573143513906633035e-07,
                             2.08757232129817482790e-09, -1.13596475577881948265e-11};
static const double ATHI[4] = {4.63647609000806093515e-01, 7.85398163397448278999e-01,
                               9.82793723247329054082e-01, 1.57079632679489655800e+00};
static const double ATLO[4] = {2.26987774529616870924e-17, 3.06161699786838301793e-17,
                               1.39033110312309984516e-17, 6.12323399573676603587e-17};
static const double AT[11] = {3.33333333333329318027e-01, -1.99999999998764832476e-01,
                              1.42857142725034663711e-01, -1.11111104054623557880e-01,
                              9.09088713343650656196e-02, -7.69187620504482999495e-02,
                              6.66107313738753120669e-02, -5.83357013379057348645e-02,
                              4.97687799461593236017e-02, -3.65315727442169155270e-02,
                              1.62858201153657823623e-02};
static const double PS[6] = {1.66666666666666657415e-01, -3.25565818622400915405e-01,
                             2.01212532134862925881e-01, -4.00555345006794114027e-02,
                             7.91534994289814532176e-04, 3.47933107596021167570e-05};
static const double QS[4] = {-2.40339491173441421878e+00, 2.02094576023350569471e+00,
                             -6.88283971605453293030e-01, 7.70381505559019352791e-02};
static const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
static const double INV_LN2 = 1.44269504088896338700e+00;
static const double EP[5] = {1.66666666666666019037e-01, -2.77777777770155933842e-03,
                             6.61375632143793436117e-05, -1.65339022054652515390e-06,
                             4.13813679705723846039e-08};
static const double LG[7] = {6.666666666666735130e-01, 3.999999999940941908e-01,
                             2.857142874366239149e-01, 2.222219843214978396e-01,
                             1.818357216161805012e-01, 1.531383769920937332e-01,
                             1.479819860511658591e-01};

/* Operation-count hooks: no-ops here.  oracle/opcount.cpp compiles this file with an operator-counting scalar in
 * place of double and uses them to tell the two kernel polynomials of sincos_parts apart, so that it can also report
 * the count of an implementation that evaluates only the polynomial a sin / cos call selects (same result bits). */
#ifndef OGG_OPCOUNT_BEGIN
#define OGG_OPCOUNT_BEGIN(bucket)
#define OGG_OPCOUNT_END(bucket)
#define OGG_OPCOUNT_UNUSED(bucket)
#endif
static void sincos_parts(double x, int *n, double *s, double *c) {
  double k = rint(x * TWO_OVER_PI);
  double r = ((x - k * RP1) - k * RP2) - k * RP3;
  double z = r * r;
  OGG_OPCOUNT_BEGIN(sin_poly)
  double p = SC[1] + z * (SC[2] + z * (SC[3] + z * (SC[4] + z * SC[5])));
  *s = r + (z * r) * (SC[0] + z * p);
  OGG_OPCOUNT_END(sin_poly)
  OGG_OPCOUNT_BEGIN(cos_poly)
  double rr = z * (CC[0] + z * (CC[1] + z * (CC[2] + z * (CC[3] + z * (CC[4] + z * CC[5])))));
  double hz = 0.5 * z;
  double w = 1.0 - hz;
  *c = w + (((1.0 - w) - hz) + z * rr);
  OGG_OPCOUNT_END(cos_poly)
  *n = (int)(((i64)k) & 3);
}
static double p_sin(double x) {
  int n; double s, c; sincos_parts(x, &n, &s, &c);
  if (n & 1) { OGG_OPCOUNT_UNUSED(sin_poly) } else { OGG_OPCOUNT_UNUSED(cos_poly) }
  return n == 0 ? s : (n == 1 ? c : (n == 2 ? -s : -c));
}
static double p_cos(double x) {
  int n; double s, c; sincos_parts(x, &n, &s, &c);
  if (n & 1) { OGG_OPCOUNT_UNUSED(cos_poly) } else { OGG_OPCOUNT_UNUSED(sin_poly) }
  return n == 0 ? c : (n == 1 ? -s : (n == 2 ? -c : s));
}
static double p_tan(double x) {
  int n; double s, c; sincos_parts(x, &n, &s, &c);
  return (n & 1) ? (-c) / s : s / c;
}
static double atan_pos(double ax) {
  double num, den, hi = 0.0, lo = 0.0;
  int id;
  if (ax < 0.4375) { id = -1; num = ax; den = 1.0; }
  else if (ax < 0.6875) { id = 0; num = 2.0 * ax - 1.0; den = 2.0 + ax; }
  else if (ax < 1.1875) { id = 1; num = ax - 1.0; den = ax + 1.0; }
  else if (ax < 2.4375) { id = 2; num = ax - 1.5; den = 1.0 + 1.5 * ax; }
  else { id = 3; num = -1.0; den = ax; }
  if (id >= 0) { hi = ATHI[id]; lo = ATLO[id]; }
  double t = num / den;
  double z = t * t, w = z * z;
  double s1 = z * (AT[0] + w * (AT[2] + w * (AT[4] + w * (AT[6] + w * (AT[8] + w * AT[10])))));
  double s2 = w * (AT[1] + w * (AT[3] + w * (AT[5] + w * (AT[7] + w * AT[9]))));
  double ts = t * (s1 + s2);
  return id < 0 ? t - ts : hi - ((ts - lo) - t);
}
static double p_atan(double x) { return copysign(atan_pos(fabs(x)), x); }
static double p_atan2(double y, double x) {
  double ay = fabs(y), ax = fabs(x);
  double q = (ay == 0.0 && ax == 0.0) ? 0.0 : ay / ax;
  double a = atan_pos(q);
  if (x < 0.0) a = PI_D - (a - PI_LO);
  return copysign(a, y);
}
static double asin_r(double t) {
  double p = t * (PS[0] + t * (PS[1] + t * (PS[2] + t * (PS[3] + t * (PS[4] + t * PS[5])))));
  double q = 1.0 + t * (QS[0] + t * (QS[1] + t * (QS[2] + t * QS[3])));
  return p / q;
}
static double p_asin(double x) {
  double ax = fabs(x), y;
  if (ax > 1.0) return NAN;
  if (ax < 0.5) {
    double w = asin_r(ax * ax);
    y = ax + ax * w;
  } else {
    double t = (1.0 - ax) * 0.5;
    double w = asin_r(t);
    double s = sqrt(t);
    y = PIO2_HI - (2.0 * (s + s * w) - PIO2_LO);
  }
  return copysign(y, x);
}
static double p_acos(double x) {
  double ax = fabs(x);
  if (ax > 1.0) return NAN;
  if (ax < 0.5) {
    double w = asin_r(x * x);
    return PIO2_HI - (x - (PIO2_LO - x * w));
  }
  double t = (1.0 - ax) * 0.5;
  double w = asin_r(t);
  double s = sqrt(t);
  if (x > 0.0) return 2.0 * (s + s * w);
  return PI_D - 2.0 * (s + (s * w - PIO2_LO));
}
static double p_exp(double x) {
  double k = rint(x * INV_LN2);
  double hi = x - k * LN2_HI;
  double lo = k * LN2_LO;
  double r = hi - lo;
  double t = r * r;
  double c = r - t * (EP[0] + t * (EP[1] + t * (EP[2] + t * (EP[3] + t * EP[4]))));
  double y = 1.0 - ((lo - (r * c) / (2.0 - c)) - hi);
  return ldexp(y, (int)k);
}
static double p_log(double x) {
  if (x == 0.0) return -INFINITY;
  if (x < 0.0 || x != x) return NAN;
  int e;
  double m = frexp(x, &e);
  if (m < 0.70710678118654752440) { m = 2.0 * m; e -= 1; }
  double k = (double)e;
  double f = m - 1.0;
  double s = f / (2.0 + f);
  double z = s * s, w = z * z;
  double t1 = w * (LG[1] + w * (LG[3] + w * LG[5]));
  double t2 = z * (LG[0] + w * (LG[2] + w * (LG[4] + w * LG[6])));
  double R = t2 + t1;
  double hfsq = 0.5 * (f * f);
  return k * LN2_HI - ((hfsq - (s * (hfsq + R) + k * LN2_LO)) - f);
}
static double p_sinh(double x) {
  static const double SH[6] = {1.0 / 6.0, 1.0 / 120.0, 1.0 / 5040.0, 1.0 / 362880.0,
                               1.0 / 39916800.0, 1.0 / 6227020800.0};
  double ax = fabs(x), y;
  if (ax < 0.5) {
    double z = ax * ax;
    double p = SH[0] + z * (SH[1] + z * (SH[2] + z * (SH[3] + z * (SH[4] + z * SH[5]))));
    y = ax + (ax * z) * p;
  } else {
    double t = p_exp(ax);
    y = 0.5 * (t - 1.0 / t);
  }
  return copysign(y, x);
}
static double p_cosh(double x) {
  double t = p_exp(fabs(x));
  return 0.5 * (t + 1.0 / t);
}
/* numpy.mod(a, 360.0): fmod, then +360 if negative, +0.0 if zero */
static double np_mod360(double a) {
  double m = fmod(a, 360.0);
  if (m != 0.0) { if (m < 0.0) m += 360.0; }
  else m = 0.0;
  return m;
}
static double mdist(double x1, double x2) { /* OGG:682-684 */
  double a = np_mod360(x1 - x2), b = np_mod360(x2 - x1);
  return a < b ? a : b;
}

#define UNARY(name, fn)                                                  \
  void name(const double *in, double *out, i64 n) {                      \
    _Pragma("omp parallel for") for (i64 i = 0; i < n; ++i) out[i] = fn(in[i]); \
  }
UNARY(ogo_sin, p_sin)
UNARY(ogo_cos, p_cos)
UNARY(ogo_tan, p_tan)
UNARY(ogo_arcsin, p_asin)
UNARY(ogo_arccos, p_acos)
UNARY(ogo_arctan, p_atan)
UNARY(ogo_exp, p_exp)
UNARY(ogo_log, p_log)
UNARY(ogo_sinh, p_sinh)
UNARY(ogo_cosh, p_cosh)
void ogo_arctan2(const double *y, const double *x, double *out, i64 n) {
#pragma omp parallel for
  for (i64 i = 0; i < n; ++i) out[i] = p_atan2(y[i], x[i]);
}
void ogo_mod360(const double *in, double *out, i64 n) {
  for (i64 i = 0; i < n; ++i) out[i] = np_mod360(in[i]);
}

/* ======================================================================================= */
/* Bipolar cap                                                                              */
/* ======================================================================================= */
/* OGG:33-100, one element.  Any output pointer may be NULL. */
static void bipolar_point(double lamg, double phig, double lon_bp, double rp, int metrics_only,
                          double *lams, double *phis_o, double *hi_o, double *hj_o) {
  phig = 90 - ((2 * p_atan(p_tan((0.5 * (90 - phig)) * PI_180) / rp)) / PI_180);     /* :41 */
  double tmp = mdist(lamg, lon_bp) * PI_180;                                           /* :42 */
  double sinla = p_sin(tmp);                                                           /* :43 */
  double sphig = p_sin(phig * PI_180);                                                 /* :44 */
  double ca = p_cos(tmp);
  double alpha2 = ca * ca;                                                             /* :45 */
  double tb = p_tan(phig * PI_180);
  double beta2_inv = tb * tb;                                                          /* :46 */
  double rden = 1.0 / (1.0 + alpha2 * beta2_inv);                                      /* :47 */
  if (!metrics_only) {
    double B = sinla * sqrt(rden);                                                     /* :50 */
    if (fabs(beta2_inv) > HUGE_) B = 0.0;                                              /* :52 */
    double lamc = p_asin(B) / PI_180;                                                  /* :53 */
    double dl = lamg - lon_bp;
    if (dl > 90 && dl <= 180) lamc = 180 - lamc;                                       /* :58 */
    if (dl > 180 && dl <= 270) lamc = 180 + lamc;                                      /* :59 */
    if (dl > 270) lamc = 360 - lamc;                                                   /* :60 */
    if (dl == 90) lamc = 90;                                                           /* :62 */
    if (dl == 270) lamc = 270;                                                         /* :63 */
    if (lams) *lams = lamc + lon_bp;                                                   /* :64 */
  }
  double A = sinla * sphig;                                                            /* :68 */
  double chic = p_acos(A);                                                             /* :69 */
  double tc = p_tan(chic / 2);
  double phis = 90 - ((2 * p_atan(rp * tc)) / PI_180);                                 /* :70 */
  double q = rp * tc;
  double rden2 = 1.0 / (1 + q * q);                                                    /* :72 */
  double M_inv = (rp * (1 + tc * tc)) * rden2;                                         /* :73 */
  double chig = (90 - phig) * PI_180;                                                  /* :75 */
  double tg = p_tan(chig / 2);
  double qg = rp * tg;
  rden2 = 1.0 / (1 + qg * qg);                                                         /* :76 */
  double N = (rp * (1 + tg * tg)) * rden2;                                             /* :77 */
  double N_inv = 1 / N;                                                                /* :78 */
  double cp = p_cos(phis * PI_180);
  double cos2phis = cp * cp;                                                           /* :79 */
  double h_j = ((((cos2phis * alpha2) * (1 - alpha2)) * beta2_inv) * (1 + beta2_inv)) * (rden * rden)
               + ((M_inv * M_inv) * (1 - alpha2)) * rden;                              /* :81-84 */
  if (fabs(beta2_inv) > HUGE_) h_j = M_inv * M_inv;                                    /* :86 */
  h_j = sqrt(h_j) * N_inv;                                                             /* :87 */
  double h_i = (cos2phis * (1 + beta2_inv)) * (rden * rden)
               + (((M_inv * M_inv) * alpha2) * beta2_inv) * rden;                      /* :89-92 */
  if (fabs(beta2_inv) > HUGE_) h_i = M_inv * M_inv;                                    /* :94 */
  h_i = sqrt(h_i);                                                                     /* :95 */
  if (phis_o) *phis_o = phis;
  if (hi_o) *hi_o = h_i;
  if (hj_o) *hj_o = h_j;
}

/* OGG:33-100 on flat arrays of n points */
void ogo_bipolar_projection(const double *lamg, const double *phig, i64 n, double lon_bp, double rp,
                            int metrics_only, double *lams, double *phis, double *hi, double *hj) {
#pragma omp parallel for
  for (i64 e = 0; e < n; ++e)
    bipolar_point(lamg[e], phig[e], lon_bp, rp, metrics_only, lams ? lams + e : NULL,
                  phis ? phis + e : NULL, hi ? hi + e : NULL, hj ? hj + e : NULL);
}

/* OGG:103-122 (the ensure_nj_even row drop :107-111 is the caller's).  rp is returned. */
double ogo_generate_bipolar_cap_mesh(i64 Ni, i64 Nj, double lat0_bp, double lon_bp, double *lams,
                                     double *phis, double *h_i_inv /*[Nj+1][Ni]*/,
                                     double *h_j_inv /*[Nj][Ni+1]*/) {
  double rp = p_tan((0.5 * (90 - lat0_bp)) * PI_180);                                  /* :117 */
#pragma omp parallel for
  for (i64 j = 0; j <= Nj; ++j) {
    double phig = lat0_bp + ((double)j * (90 - lat0_bp)) / (double)Nj;                 /* :115 */
    for (i64 i = 0; i <= Ni; ++i) {
      double lamg = lon_bp + ((double)i * 360.0) / (double)Ni;                         /* :113 */
      double hi, hj;
      bipolar_point(lamg, phig, lon_bp, rp, 0, lams + j * (Ni + 1) + i, phis + j * (Ni + 1) + i, &hi, &hj);
      if (h_i_inv && i < Ni) h_i_inv[j * Ni + i] = ((hi * 2) * PI_D) / (double)Ni;     /* :119 */
      if (h_j_inv && j < Nj) h_j_inv[j * (Ni + 1) + i] = ((hj * PI_180) * (90 - lat0_bp)) / (double)Nj; /* :120 */
    }
  }
  return rp;
}

/* OGG:125-133 */
void ogo_bipolar_cap_ij_array(const double *i, i64 ni, const double *j, i64 nj, i64 Ni, i64 Nj_ncap,
                              double lat0_bp, double lon_bp, double rp, double *h_i_inv, double *h_j_inv) {
#pragma omp parallel for
  for (i64 jj = 0; jj < nj; ++jj) {
    double latg = lat0_bp + (j[jj] * (90 - lat0_bp)) / (double)Nj_ncap;                /* :127 */
    for (i64 ii = 0; ii < ni; ++ii) {
      double lng = lon_bp + (i[ii] * 360.0) / (double)Ni;                              /* :126 */
      double hi, hj;
      bipolar_point(lng, latg, lon_bp, rp, 1, NULL, NULL, &hi, &hj);
      h_i_inv[jj * ni + ii] = ((hi * 2) * PI_D) / (double)Ni;                          /* :131 */
      h_j_inv[jj * ni + ii] = ((hj * (90 - lat0_bp)) * PI_180) / (double)Nj_ncap;      /* :132 */
    }
  }
}

/* ---- quadrature helpers OGG:191-255 ---- */
int ogo_quad_positions(int n, double *a, double *b) {
  if (n == 2) { a[0] = 0.0; a[1] = 1.0; b[0] = 1.0; b[1] = 0.0; return 0; }
  if (n == 3) { a[0] = 0.0; a[1] = 0.5; a[2] = 1.0; b[0] = 1.0; b[1] = 0.5; b[2] = 0.0; return 0; }
  if (n == 4) {
    double r5 = 0.5 / sqrt(5.0);
    a[0] = 0.0; a[1] = 0.5 - r5; a[2] = 0.5 + r5; a[3] = 1.0;
    b[0] = 1.0; b[1] = 0.5 + r5; b[2] = 0.5 - r5; b[3] = 0.0;
    return 0;
  }
  if (n == 5) {
    double r37 = 0.5 * sqrt(3.0 / 7.0);
    a[0] = 0.0; a[1] = 0.5 - r37; a[2] = 0.5; a[3] = 0.5 + r37; a[4] = 1.0;
    b[0] = 1.0; b[1] = 0.5 + r37; b[2] = 0.5; b[3] = 0.5 - r37; b[4] = 0.0;
    return 0;
  }
  return -1; /* "Uncoded order" :204 */
}
/* y has stride `st` */
static double quad_average(const double *y, i64 st, int n) {                           /* :207-222 */
  if (n == 2) { double d = 1.0 / 2.0; return d * (y[0] + y[st]); }
  if (n == 3) { double d = 1.0 / 6.0; return d * (4.0 * y[st] + (y[0] + y[2 * st])); }
  if (n == 4) { double d = 1.0 / 12.0; return d * (5.0 * (y[st] + y[2 * st]) + (y[0] + y[3 * st])); }
  double d = 1.0 / 180.0;
  return d * ((64.0 * y[2 * st] + (49.0 * (y[st] + y[3 * st]))) + 9.0 * (y[0] + y[4 * st]));
}
/* y[j*sj + i*si] */
static double quad_average_2d(const double *y, i64 sj, i64 si, int n) {                /* :225-255 */
#define Y(j, i) y[(j) * sj + (i) * si]
  if (n == 2) { double d = 1.0 / 2.0; return (d * d) * (((Y(0, 0) + Y(0, 1)) + Y(1, 0)) + Y(1, 1)); }
  if (n == 3) {
    double d = 1.0 / 6.0;
    double inner = (((Y(0, 1) + Y(1, 0)) + Y(1, 2)) + Y(2, 1)) + 4.0 * Y(1, 1);
    double outer = (((Y(0, 0) + Y(0, 2)) + Y(2, 0)) + Y(2, 2)) + 4.0 * inner;
    return (d * d) * outer;
  }
  static const double w4[4] = {1.0, 5.0, 5.0, 1.0};
  static const double w5[5] = {9.0, 49.0, 64.0, 49.0, 9.0};
  const double *w = n == 4 ? w4 : w5;
  double d = n == 4 ? 1.0 / 12.0 : 1.0 / 180.0;
  double ysum = 0.0;
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < n; ++i) ysum = ysum + (w[i] * w[j]) * Y(j, i);
  return (d * d) * ysum;
#undef Y
}
void ogo_quad_average(const double *y, int n, double *out) { *out = quad_average(y, 1, n); }
void ogo_quad_average_2d(const double *y, int n, double *out) { *out = quad_average_2d(y, n, 1, n); }

/* sample index vector j1d of OGG:143-156 / 573-581; fudge => the ny-0.001 replacement :146-147 */
static double *sample_index(int order, i64 n, const double *a, const double *b, int fudge) {
  double *v = (double *)malloc(sizeof(double) * (size_t)(order * (n + 1)));
  for (i64 j = 0; j <= n; ++j) {
    for (int k = 0; k < order; ++k) v[j * order + k] = b[k] * (double)j + a[k] * (double)(j + 1);
    if (fudge && v[j * order + order - 1] == (double)n) v[j * order + order - 1] = (double)n - 0.001;
  }
  return v;
}

/* shared tail of OGG:174-187 and :585-599: dx, dy are [order*(ny+1)][order*(nx+1)] */
static void quad_cells(int order, i64 nx, i64 ny, const double *dx, const double *dy, double Re,
                       double *dxq, double *dyq, double *daq) {
  i64 ni = order * (nx + 1);
#pragma omp parallel for
  for (i64 j = 0; j <= ny; ++j) {
    double prod[25];
    for (i64 i = 0; i <= nx; ++i) {
      const double *px = dx + (j * order) * ni + i * order;
      const double *py = dy + (j * order) * ni + i * order;
      if (j < ny && i < nx) {
        for (int jj = 0; jj < order; ++jj)
          for (int ii = 0; ii < order; ++ii) prod[jj * order + ii] = px[jj * ni + ii] * py[jj * ni + ii];
        daq[j * nx + i] = (quad_average_2d(prod, order, 1, order) * Re) * Re;          /* :185 */
      }
      if (i < nx) dxq[j * nx + i] = quad_average(px, 1, order) * Re;                   /* :183,:186 */
      if (j < ny) dyq[j * (nx + 1) + i] = quad_average(py, ni, order) * Re;            /* :184,:187 */
    }
  }
}

/* exported face of quad_cells for row-slab compositions (oracle/supergrid_cpu.py): dx, dy are the [order*(ny+1)][order*(nx+1)]
 * sample arrays of ANY ny consecutive cell rows (+ the ghost row), exactly as OGG:174-187 / :585-599 consume them */
void ogo_quad_cells(int order, i64 nx, i64 ny, const double *dx, const double *dy, double Re, double *dxq, double *dyq,
                    double *daq) {
  quad_cells(order, nx, ny, dx, dy, Re, dxq, dyq, daq);
}

/* OGG:136-188. returns 0, or -1 for an uncoded order */
int ogo_bipolar_cap_metrics_quad_fast(int order, i64 nx, i64 ny, double lat0_bp, double lon_bp, double rp,
                                      double Re, double *dxq, double *dyq, double *daq) {
  double a[5], b[5];
  if (ogo_quad_positions(order, a, b)) return -1;
  double *j1d = sample_index(order, ny, a, b, 1);
  double *i1d = sample_index(order, nx, a, b, 0);
  i64 nj = order * (ny + 1), ni = order * (nx + 1);
  double *dx = (double *)malloc(sizeof(double) * (size_t)(nj * ni));
  double *dy = (double *)malloc(sizeof(double) * (size_t)(nj * ni));
  ogo_bipolar_cap_ij_array(i1d, ni, j1d, nj, nx, ny, lat0_bp, lon_bp, rp, dx, dy);
  quad_cells(order, nx, ny, dx, dy, Re, dxq, dyq, daq);
  free(dx); free(dy); free(j1d); free(i1d);
  return 0;
}

/* ======================================================================================= */
/* Displaced-pole cap                                                                       */
/* ======================================================================================= */
typedef struct { double re, im; } cplx;
static cplx cmul(cplx a, cplx b) { /* numpy CDOUBLE multiply, strict */
  cplx r = {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re};
  return r;
}
static cplx cdiv(cplx a, cplx b) { /* numpy CDOUBLE divide (Smith) */
  cplx r;
  double abr = fabs(b.re), abi = fabs(b.im);
  if (abr >= abi) {
    if (abr == 0 && abi == 0) { r.re = a.re / abr; r.im = a.im / abr; return r; }
    double rat = b.im / b.re;
    double scl = 1.0 / (b.re + b.im * rat);
    r.re = (a.re + a.im * rat) * scl;
    r.im = (a.im - a.re * rat) * scl;
  } else {
    double rat = b.re / b.im;
    double scl = 1.0 / (b.im + b.re * rat);
    r.re = (a.re * rat + a.im) * scl;
    r.im = (a.im * rat - a.re) * scl;
  }
  return r;
}
static cplx cfromreal(double x) { cplx r = {x, 0.0}; return r; }
static cplx cadd(cplx a, cplx b) { cplx r = {a.re + b.re, a.im + b.im}; return r; }
static cplx csub(cplx a, cplx b) { cplx r = {a.re - b.re, a.im - b.im}; return r; }
static cplx cconj(cplx a) { cplx r = {a.re, -a.im}; return r; }

/* cos(t) + 1j*sin(t) with numpy's promotions (OGG:451, :495) */
static cplx cis_np(double t) {
  cplx j1 = {0.0, 1.0};
  cplx js = cmul(j1, cfromreal(p_sin(t)));
  return cadd(cfromreal(p_cos(t)), js);
}

/* OGG:447-466 without the unwrap; raw lam in (-180,180] */
static void dp_point(double lon, double lat, cplx z0, double r_joint, double *lam, double *phi) {
  double r = p_tan((90 + lat) * PI_180) / r_joint;                                     /* :448 */
  cplx e = cis_np(lon * PI_180);                                                       /* :451 */
  cplx ep = cdiv(csub(e, z0), csub(cfromreal(1.0), cmul(cconj(z0), e)));               /* :452 */
  cplx z = cmul(cfromreal(r), ep);                                                     /* :454 */
  cplx w = cdiv(cadd(z, z0), cadd(cfromreal(1.0), cmul(cconj(z0), z)));                /* :455 */
  *lam = p_atan2(w.im, w.re) * (180.0 / PI_D);                                         /* :457 */
  double rw = sqrt(w.re * w.re + w.im * w.im);                                         /* :465 */
  *phi = -90 + p_atan(rw * r_joint) / PI_180;                                          /* :466 */
}

/* OGG:470-475, in place */
void ogo_monotonic_bounding(double *x, i64 nj, i64 ni, double x_0) {
#pragma omp parallel for
  for (i64 j = 0; j < nj; ++j) {
    double x_im1 = x[j * ni] * 0 + x_0;
    for (i64 i = 0; i < ni; ++i) {
      double v = x[j * ni + i];
      if (v - x_im1 > 100) v = v - 360;
      x[j * ni + i] = v;
      x_im1 = v;
    }
  }
}

/* OGG:447-467 on 2-D arrays */
void ogo_dp_projection(const double *lon, const double *lat, i64 nj, i64 ni, double z0_re, double z0_im,
                       double r_joint, double *lam, double *phi) {
  cplx z0 = {z0_re, z0_im};
#pragma omp parallel for
  for (i64 e = 0; e < nj * ni; ++e) dp_point(lon[e], lat[e], z0, r_joint, lam + e, phi + e);
  ogo_monotonic_bounding(lam, nj, ni, lon[0]);                                         /* :463 */
}

/* r_joint and z_0 of OGG:494-495 */
void ogo_dp_constants(double lat0, double lam_pole, double r_pole, double *r_joint, double *z0_re, double *z0_im) {
  *r_joint = p_tan((90 + lat0) * PI_180);
  cplx z0 = cmul(cfromreal(r_pole), cis_np(lam_pole * PI_180));
  *z0_re = z0.re; *z0_im = z0.im;
}

/* OGG:478-506 with excluded_fraction=None: i[nil], j[njl] (possibly fractional) -> lams, phis [njl][nil] */
void ogo_dp_mesh(const double *i, i64 nil, const double *j, i64 njl, i64 ni, i64 nj, double lon0, double lat0,
                 double lam_pole, double r_pole, double *lams, double *phis) {
  double r_joint, zr, zi;
  ogo_dp_constants(lat0, lam_pole, r_pole, &r_joint, &zr, &zi);
  cplx z0 = {zr, zi};
  double a = -90.0, b = lat0;
#pragma omp parallel for
  for (i64 jj = 0; jj < njl; ++jj) {
    double v = a + (j[jj] * (b - a)) / (double)nj;                                     /* :482 */
    for (i64 ii = 0; ii < nil; ++ii) {
      double u = lon0 + (i[ii] * 360.0) / (double)ni;                                  /* :479 */
      dp_point(u, v, z0, r_joint, lams + jj * nil + ii, phis + jj * nil + ii);
    }
  }
  double u0 = lon0 + (i[0] * 360.0) / (double)ni;
  ogo_monotonic_bounding(lams, njl, nil, u0);
}

/* OGG:522-532 */
void ogo_great_arc_distance(const double *j0, const double *i0, const double *j1, const double *i1, i64 njl,
                            i64 nil, i64 nx, i64 ny, double lon0, double lat0, double lon_dp, double r_dp,
                            double *out) {
  size_t n = (size_t)(njl * nil);
  double *lam0 = (double *)malloc(8 * n), *phi0 = (double *)malloc(8 * n);
  double *lam1 = (double *)malloc(8 * n), *phi1 = (double *)malloc(8 * n);
  ogo_dp_mesh(i0, nil, j0, njl, nx, ny, lon0, lat0, lon_dp, r_dp, lam0, phi0);
  ogo_dp_mesh(i1, nil, j1, njl, nx, ny, lon0, lat0, lon_dp, r_dp, lam1, phi1);
#pragma omp parallel for
  for (i64 e = 0; e < (i64)n; ++e) {
    double l0 = lam0[e] * PI_180, p0 = phi0[e] * PI_180;
    double l1 = lam1[e] * PI_180, p1 = phi1[e] * PI_180;
    double dphi = p1 - p0, dlam = l1 - l0;
    double s1 = p_sin(0.5 * dphi), s2 = p_sin(0.5 * dlam);
    double d = s1 * s1 + ((s2 * s2) * p_cos(p0)) * p_cos(p1);                          /* :531 */
    out[e] = 2.0 * p_asin(sqrt(d));                                                    /* :532 */
  }
  free(lam0); free(phi0); free(lam1); free(phi1);
}

static double *shifted(const double *v, i64 n, double s) {
  double *o = (double *)malloc(8 * (size_t)n);
  for (i64 k = 0; k < n; ++k) o[k] = v[k] + s;
  return o;
}
static double *shifted_m(const double *v, i64 n, double s) {
  double *o = (double *)malloc(8 * (size_t)n);
  for (i64 k = 0; k < n; ++k) o[k] = v[k] - s;
  return o;
}
/* OGG:535-562; dir 0 = hi (shift i), 1 = hj (shift j). returns -1 for "order not coded" */
int ogo_numerical_h(int dir, const double *j, i64 njl, const double *i, i64 nil, i64 nx, i64 ny, double lon0,
                    double lat0, double lon_dp, double r_dp, double eps, int order, double *out) {
  size_t n = (size_t)(njl * nil);
  double reps = 1.0 / eps;
  const double *v = dir == 0 ? i : j;
  i64 nv = dir == 0 ? nil : njl;
  int nds = order == 2 ? 1 : (order == 4 ? 2 : 3);
  double *ds[3] = {NULL, NULL, NULL};
  for (int m = 0; m < nds; ++m) {
    double sh = m == 0 ? eps : (m == 1 ? 2.0 * eps : 3.0 * eps);
    double *vp = shifted(v, nv, sh), *vm = shifted_m(v, nv, sh);
    ds[m] = (double *)malloc(8 * n);
    if (dir == 0) ogo_great_arc_distance(j, vp, j, vm, njl, nil, nx, ny, lon0, lat0, lon_dp, r_dp, ds[m]);
    else ogo_great_arc_distance(vp, i, vm, i, njl, nil, nx, ny, lon0, lat0, lon_dp, r_dp, ds[m]);
    free(vp); free(vm);
  }
  int rc = 0;
  if (order == 2) for (size_t e = 0; e < n; ++e) out[e] = (0.5 * ds[0][e]) * reps;
  else if (order == 4) for (size_t e = 0; e < n; ++e) out[e] = ((8.0 * ds[0][e] - ds[1][e]) * (1.0 / 12.0)) * reps;
  else if (order == 6)
    for (size_t e = 0; e < n; ++e) out[e] = (((45.0 * ds[0][e] - 9.0 * ds[1][e]) + ds[2][e]) * (1.0 / 60.0)) * reps;
  else rc = -1;
  for (int m = 0; m < 3; ++m) free(ds[m]);
  return rc;
}

/* OGG:565-601. returns -1 "Uncoded order" (quad_positions), -2 "order not coded" (numerical_h) */
int ogo_dp_metrics_quad(int order, i64 nx, i64 ny, double lon0, double lat0, double lon_dp, double r_dp, double Re,
                        double *dxq, double *dyq, double *daq) {
  double a[5], b[5];
  if (ogo_quad_positions(order, a, b)) return -1;
  if (order != 2 && order != 4) return -2;
  double *j1d = sample_index(order, ny, a, b, 0);
  double *i1d = sample_index(order, nx, a, b, 0);
  i64 nj = order * (ny + 1), ni = order * (nx + 1);
  double *dx = (double *)malloc(8 * (size_t)(nj * ni)), *dy = (double *)malloc(8 * (size_t)(nj * ni));
  ogo_numerical_h(0, j1d, nj, i1d, ni, nx, ny, lon0, lat0, lon_dp, r_dp, 1e-3, order, dx);
  ogo_numerical_h(1, j1d, nj, i1d, ni, nx, ny, lon0, lat0, lon_dp, r_dp, 1e-3, order, dy);
  quad_cells(order, nx, ny, dx, dy, Re, dxq, dyq, daq);
  free(dx); free(dy); free(j1d); free(i1d);
  return 0;
}

/* ======================================================================================= */
/* Spherical bands: MIDAS metrics, angle, lat-lon tile                                       */
/* ======================================================================================= */
/* OGG:687-716.  x,y [nyp][nxp] -> dx [nyp][nxp-1], dy [nyp-1][nxp], area [nyp-1][nxp-1] */
void ogo_metrics_midas(const double *x, const double *y, i64 nyp, i64 nxp, double Re, int areafix, double *dx,
                       double *dy, double *area) {
  i64 nx = nxp - 1, ny = nyp - 1;
  double *dxi = (double *)malloc(8 * (size_t)(nyp * nx)), *sl = (double *)malloc(8 * (size_t)(nyp * nx));
#pragma omp parallel for
  for (i64 j = 0; j < nyp; ++j)
    for (i64 i = 0; i < nx; ++i) {
      double yr = y[j * nxp + i + 1], yl = y[j * nxp + i], xr = x[j * nxp + i + 1], xl = x[j * nxp + i];
      double lv = (0.5 * (yr + yl)) * PI_180;                                          /* :695 */
      double dx_i = mdist(xr, xl) * PI_180;                                            /* :696 */
      double dy_i = (yr - yl) * PI_180;                                                /* :697 */
      double t = dx_i * p_cos(lv);
      dx[j * nx + i] = Re * sqrt(dy_i * dy_i + t * t);                                 /* :698 */
      dxi[j * nx + i] = dx_i;
      sl[j * nx + i] = p_sin(lv);                                                      /* :711 */
    }
#pragma omp parallel for
  for (i64 j = 0; j < ny; ++j)
    for (i64 i = 0; i < nxp; ++i) {
      double yu = y[(j + 1) * nxp + i], yd = y[j * nxp + i], xu = x[(j + 1) * nxp + i], xd = x[j * nxp + i];
      double lu = (0.5 * (yu + yd)) * PI_180;                                          /* :699 */
      double dx_j = mdist(xu, xd) * PI_180;                                            /* :700 */
      double dy_j = (yu - yd) * PI_180;                                                /* :701 */
      double t = dx_j * p_cos(lu);
      dy[j * nxp + i] = Re * sqrt(dy_j * dy_j + t * t);                                /* :702 */
    }
#pragma omp parallel for
  for (i64 j = 0; j < ny; ++j)
    for (i64 i = 0; i < nx; ++i) {
      if (areafix)
        area[j * nx + i] = (Re * Re) * ((0.5 * (dxi[(j + 1) * nx + i] + dxi[j * nx + i])) *
                                        (sl[(j + 1) * nx + i] - sl[j * nx + i]));       /* :713 */
      else
        area[j * nx + i] = 0.25 * ((dx[(j + 1) * nx + i] + dx[j * nx + i]) *
                                   (dy[j * nxp + i + 1] + dy[j * nxp + i]));            /* :715 */
    }
  free(dxi); free(sl);
}

/* OGG:719-729 */
void ogo_angle_x(const double *x, const double *y, i64 nyp, i64 nxp, double *angle) {
#pragma omp parallel for
  for (i64 j = 0; j < nyp; ++j) {
    const double *xr = x + j * nxp, *yr = y + j * nxp;
    for (i64 i = 0; i < nxp; ++i) {
      i64 ia = i == nxp - 1 ? i : i + 1, ib = i == 0 ? i : i - 1;                      /* :725-727 */
      double a = p_atan2(yr[ia] - yr[ib], (xr[ia] - xr[ib]) * p_cos(yr[i] * PI_180));
      angle[j * nxp + i] = a / PI_180;                                                 /* :728 */
    }
  }
}

/* OGG:832-846 1-D laws + tile (row drop for ensure_nj_even is the caller's) */
void ogo_latlon_grid(i64 lni, i64 lnj, double llon0, double llen_lon, double llat0, double llen_lat, double *lam,
                     double *phi) {
  for (i64 j = 0; j <= lnj; ++j) {
    double p = llat0 + ((double)j * llen_lat) / (double)lnj;                           /* :835 */
    for (i64 i = 0; i <= lni; ++i) {
      lam[j * (lni + 1) + i] = llon0 + ((double)i * llen_lon) / (double)lni;           /* :834 */
      phi[j * (lni + 1) + i] = p;
    }
  }
}

int ogo_num_threads(void) {
  int n = 1;
#ifdef _OPENMP
#pragma omp parallel
  {
#pragma omp master
    n = omp_get_num_threads();
  }
#endif
  return n;
}

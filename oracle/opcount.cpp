// opcount.cpp -- derives the per-unit FP64 operation counts of the supergrid hot path from the CPU oracle itself.
// TEST / MEASUREMENT INFRASTRUCTURE (tools/opcount.py builds and runs it); never part of the product.
//
// oracle/ogg_oracle.c (the plain-C restatement of the reference, OGG:n = ocean_grid_generator.py line n) is compiled
// here a second time with `double` replaced by an operator-counting scalar, so every + - * / sqrt fmod compare and rint
// the restatement performs is tallied -- nothing is typed in by hand.  The units are the ones
// bench.py's roofline uses:
//   bipolar metrics-only sample  OGG:33-100 with metrics_only + the scalings :131-132   (k_bp_quad_fused)
//   displaced-pole vertex        OGG:447-467 (projection) + :470-475 (unwrap step)      (8 per h-sample)
//   haversine                    OGG:527-532                                             (4 per h-sample)
//   displaced-pole h-sample      OGG:535-562 twice (h_i, h_j), order 4                   (k_dp_hsamples)
// For each unit three tallies are printed:
//   "reference"  everything the straight-line restatement evaluates per sample (both kernel polynomials of every
//                sin/cos, the row-only and column-only sub-expressions re-evaluated at every sample);
//   "hoisted"    the same minus the sub-expressions that depend on the row index only or on the column index only
//                (they are O(Ni+Nj) table work in any sensible implementation);
//   "minimal"    "hoisted" minus the kernel polynomial a sin/cos call does not select and minus divisions by 1.0
//                (arctan's first range) -- what an implementation must execute to produce the same bits.
// Counts are averages over the actual sample lattices of a cap (they depend on the data through the range branches).
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <type_traits>
#include <vector>

namespace cnt {
enum { ADD, MUL, DIV, DIV_CONST, DIV_ONE, SQRT, FMOD, CMP, ROUND, SIGN, NKIND };
static const char* const kNames[NKIND] = {"add_sub", "mul", "div", "div_by_constant", "div_by_one", "sqrt", "fmod", "compare",
                                          "round", "sign_ops"};
struct Tally {
  double n[NKIND];
  Tally() { memset(n, 0, sizeof n); }
  Tally operator-(const Tally& o) const { Tally r; for (int k = 0; k < NKIND; ++k) r.n[k] = n[k] - o.n[k]; return r; }
  Tally operator+(const Tally& o) const { Tally r; for (int k = 0; k < NKIND; ++k) r.n[k] = n[k] + o.n[k]; return r; }
  Tally operator*(double s) const { Tally r; for (int k = 0; k < NKIND; ++k) r.n[k] = n[k] * s; return r; }
};
static Tally g;                     // everything executed
static Tally g_unused;              // ops inside a kernel polynomial whose value the caller then discards
static Tally g_scope_start, g_last_sin_poly, g_last_cos_poly;
static double g_consts[8];
static int g_nconst = 0;
static inline void hit(int k) { g.n[k] += 1.0; }
}  // namespace cnt

struct Counted {
  double v;
  Counted() = default;
  template <class T, class = typename std::enable_if<std::is_arithmetic<T>::value>::type>
  Counted(T x) : v((double)x) {}   // int -> double conversions are literal promotions and index casts: not tallied
  explicit operator int() const { return (int)v; }
  explicit operator long() const { return (long)v; }
  explicit operator long long() const { return (long long)v; }
  explicit operator bool() const { return v != 0.0; }
};
static inline Counted raw(double x) { Counted c; c.v = x; return c; }
static inline Counted operator+(Counted a, Counted b) { cnt::hit(cnt::ADD); return raw(a.v + b.v); }
static inline Counted operator-(Counted a, Counted b) { cnt::hit(cnt::ADD); return raw(a.v - b.v); }
static inline Counted operator*(Counted a, Counted b) { cnt::hit(cnt::MUL); return raw(a.v * b.v); }
static inline Counted operator/(Counted a, Counted b) {
  int kind = cnt::DIV;
  if (b.v == 1.0) kind = cnt::DIV_ONE;
  for (int k = 0; k < cnt::g_nconst; ++k) if (b.v == cnt::g_consts[k]) kind = cnt::DIV_CONST;
  cnt::hit(kind);
  return raw(a.v / b.v);
}
static inline Counted operator-(Counted a) { cnt::hit(cnt::SIGN); return raw(-a.v); }
static inline Counted& operator+=(Counted& a, Counted b) { a = a + b; return a; }
static inline Counted& operator-=(Counted& a, Counted b) { a = a - b; return a; }
#define OGG_CMP(op) static inline bool operator op(Counted a, Counted b) { cnt::hit(cnt::CMP); return a.v op b.v; }
OGG_CMP(<) OGG_CMP(>) OGG_CMP(<=) OGG_CMP(>=) OGG_CMP(==) OGG_CMP(!=)
static inline Counted fabs(Counted a) { cnt::hit(cnt::SIGN); return raw(::fabs(a.v)); }
static inline Counted copysign(Counted a, Counted b) { cnt::hit(cnt::SIGN); return raw(::copysign(a.v, b.v)); }
static inline Counted sqrt(Counted a) { cnt::hit(cnt::SQRT); return raw(::sqrt(a.v)); }
static inline Counted rint(Counted a) { cnt::hit(cnt::ROUND); return raw(::rint(a.v)); }
static inline Counted fmod(Counted a, Counted b) { cnt::hit(cnt::FMOD); return raw(::fmod(a.v, b.v)); }
static inline Counted ldexp(Counted a, int e) { return raw(::ldexp(a.v, e)); }
static inline Counted frexp(Counted a, int* e) { return raw(::frexp(a.v, e)); }

#define OGG_OPCOUNT_BEGIN(bucket) cnt::g_scope_start = cnt::g;
#define OGG_OPCOUNT_END(bucket) cnt::g_last_##bucket = cnt::g - cnt::g_scope_start;
#define OGG_OPCOUNT_UNUSED(bucket) cnt::g_unused = cnt::g_unused + cnt::g_last_##bucket;

// ---- the oracle, re-typed ----------------------------------------------------------------------------------------
#define double Counted
#include "ogg_oracle.c"
#undef double

// ---- measurement ----------------------------------------------------------------------------------------------------
using cnt::Tally;
struct Snap { Tally all, unused; };
static Snap snap() { return Snap{cnt::g, cnt::g_unused}; }
struct Unit {
  Tally all, unused;   // totals over `count` evaluations
  double count = 0;
  void add(const Snap& a, const Snap& b, double n) { all = all + (b.all - a.all); unused = unused + (b.unused - a.unused); count += n; }
  Tally mean_all() const { return all * (1.0 / count); }
  Tally mean_unused() const { return unused * (1.0 / count); }
};
static void print_tally(const char* key, const Tally& t, bool last) {
  printf("    \"%s\": {", key);
  for (int k = 0; k < cnt::NKIND; ++k) printf("\"%s\": %.4f%s", cnt::kNames[k], t.n[k], k + 1 < cnt::NKIND ? ", " : "");
  printf("}%s\n", last ? "" : ",");
}
// minimal = hoisted - unused polynomial - divisions by one
static Tally minimal_of(Tally hoisted, const Tally& unused) {
  Tally m = hoisted - unused;
  m.n[cnt::DIV_ONE] = 0.0;
  return m;
}
static void print_unit(const char* name, const char* where, const Tally& reference, const Tally& hoisted, const Tally& unused_h,
                       bool last) {
  printf("  \"%s\": {\n    \"where\": \"%s\",\n", name, where);
  print_tally("reference", reference, false);
  print_tally("hoisted", hoisted, false);
  print_tally("minimal", minimal_of(hoisted, unused_h), true);
  printf("  }%s\n", last ? "" : ",");
}

int main(int argc, char** argv) {
  // lattice: the caps of `-r R --r_dp 0.2` (defaults: R = 1; the per-sample averages move in the third digit with R)
  const double R = argc > 1 ? atof(argv[1]) : 1.0;
  const i64 Ni = (i64)(R * 2 * 360), Nj_bp = (i64)(60 * R * 2), Nj_dp = ((i64)(R * 40) * 7) / 4;
  const double lat0_bp = 64.05895973, lon_bp = -300.0, lon0 = -300.0, lat0_dp = -78.0, lon_dp = 80.0, r_dp = 0.2;
  const Counted PI180 = PI_180;
  cnt::g_consts[cnt::g_nconst++] = PI180.v;
  cnt::g_consts[cnt::g_nconst++] = (double)Ni;
  cnt::g_consts[cnt::g_nconst++] = (double)Nj_bp;
  cnt::g_consts[cnt::g_nconst++] = (double)Nj_dp;

  Counted qa[5], qb[5];
  // ---------------- bipolar cap: unique Gauss-Lobatto samples of bipolar_cap_metrics_quad_fast(order 5) ----------------
  ogo_quad_positions(5, qa, qb);
  Counted* j1d = sample_index(5, Nj_bp, qa, qb, 1);
  Counted* i1d = sample_index(5, Ni, qa, qb, 0);
  const Counted rp = p_tan((0.5 * (90 - Counted(lat0_bp))) * PI180);
  Unit bp_full, bp_row, bp_col;
  // unique samples: nodes 0..3 of every cell + the closing node (node 4 of the last cell)
  std::vector<i64> ju, iu;
  for (i64 j = 0; j < 5 * (Nj_bp + 1); ++j) if (j % 5 != 4 || j / 5 == Nj_bp - 1) if (j / 5 < Nj_bp) ju.push_back(j);
  for (i64 i = 0; i < 5 * (Ni + 1); ++i) if (i % 5 != 4 || i / 5 == Ni - 1) if (i / 5 < Ni) iu.push_back(i);
  for (i64 jj : ju) {
    const Counted latg = Counted(lat0_bp) + (j1d[jj] * (90 - Counted(lat0_bp))) / Counted((double)Nj_bp);          // :127 (1-D)
    {  // row-only sub-expressions of OGG:41, 44, 46, 75-78
      Snap a = snap();
      Counted phig = 90 - ((2 * p_atan(p_tan((0.5 * (90 - latg)) * PI180) / rp)) / PI180);
      Counted sphig = p_sin(phig * PI180);
      Counted tb = p_tan(phig * PI180);
      Counted beta2_inv = tb * tb;
      Counted chig = (90 - phig) * PI180;
      Counted tg = p_tan(chig / 2);
      Counted qg = rp * tg;
      Counted rden2 = Counted(1.0) / (1 + qg * qg);
      Counted N = (rp * (1 + tg * tg)) * rden2;
      Counted N_inv = 1 / N;
      (void)sphig; (void)beta2_inv; (void)N_inv;
      bp_row.add(a, snap(), 1);
    }
    for (i64 ii : iu) {
      const Counted lng = Counted(lon_bp) + (i1d[ii] * 360.0) / Counted((double)Ni);                                // :126 (1-D)
      if (jj == ju[0]) {  // column-only sub-expressions of OGG:42, 43, 45
        Snap a = snap();
        Counted tmp = mdist(lng, lon_bp) * PI180;
        Counted sinla = p_sin(tmp);
        Counted ca = p_cos(tmp);
        Counted alpha2 = ca * ca;
        (void)sinla; (void)alpha2;
        bp_col.add(a, snap(), 1);
      }
      Snap a = snap();
      Counted hi, hj;
      bipolar_point(lng, latg, lon_bp, rp, 1, NULL, NULL, &hi, &hj);
      Counted s1 = ((hi * 2) * PI_D) / Counted((double)Ni);                                                          // :131
      Counted s2 = ((hj * (90 - Counted(lat0_bp))) * PI180) / Counted((double)Nj_bp);                                // :132
      (void)s1; (void)s2;
      bp_full.add(a, snap(), 1);
    }
  }
  // ---------------- displaced-pole cap: unique samples of displacedPoleCap_metrics_quad(order 4) ----------------------
  ogo_quad_positions(4, qa, qb);
  Counted* j1q = sample_index(4, Nj_dp, qa, qb, 0);
  Counted* i1q = sample_index(4, Ni, qa, qb, 0);
  std::vector<Counted> jv, iv;
  // cell rows kept after the doughnut cut (OGG:1177-1186 with the CLI default exfracdp = 0.49): the rows the product builds
  i64 jmint = (i64)ceil(0.49 * (double)Nj_dp);
  jmint += jmint % 2;
  for (i64 j = 4 * jmint; j < 4 * Nj_dp; ++j) if (j % 4 != 3 || j / 4 == Nj_dp - 1) jv.push_back(j1q[j]);
  for (i64 i = 0; i < 4 * Ni; ++i) if (i % 4 != 3 || i / 4 == Ni - 1) iv.push_back(i1q[i]);
  const i64 njl = (i64)jv.size(), nil = (i64)iv.size();
  const double nsamp = (double)njl * (double)nil;
  Counted r_joint, z0re, z0im;
  ogo_dp_constants(lat0_dp, lon_dp, r_dp, &r_joint, &z0re, &z0im);
  cplx z0 = {z0re, z0im};
  Unit dp_full, dp_row, dp_col, unwrap, hav, hsample, fdc;
  for (i64 jj = 0; jj < njl; ++jj) {
    const Counted v = Counted(-90.0) + (jv[jj] * (Counted(lat0_dp) - Counted(-90.0))) / Counted((double)Nj_dp);     // :482 (1-D)
    {
      Snap a = snap();
      Counted r = p_tan((90 + v) * PI180) / r_joint;                                                                 // :448
      (void)r;
      dp_row.add(a, snap(), 1);
    }
    for (i64 ii = 0; ii < nil; ++ii) {
      const Counted u = Counted(lon0) + (iv[ii] * 360.0) / Counted((double)Ni);                                      // :479 (1-D)
      if (jj == 0) {
        Snap a = snap();
        cplx e = cis_np(u * PI180);                                                                                  // :451
        cplx ep = cdiv(csub(e, z0), csub(cfromreal(1.0), cmul(cconj(z0), e)));                                       // :452
        (void)ep;
        dp_col.add(a, snap(), 1);
      }
      Snap a = snap();
      Counted lam, phi;
      dp_point(u, v, z0, r_joint, &lam, &phi);
      dp_full.add(a, snap(), 1);
    }
  }
  {  // unwrap step (:470-475) on a projected mesh
    std::vector<Counted> lams(nsamp), phis(nsamp);
    ogo_dp_mesh(iv.data(), nil, jv.data(), njl, Ni, Nj_dp, lon0, lat0_dp, lon_dp, r_dp, lams.data(), phis.data());
    for (auto& x : lams) if (x.v < -120.0) x.v += 360.0;   // undo, so that the measured pass shifts again
    Snap a = snap();
    ogo_monotonic_bounding(lams.data(), njl, nil, Counted(lon0) + (iv[0] * 360.0) / Counted((double)Ni));
    unwrap.add(a, snap(), nsamp);
  }
  {  // haversine = great_arc_distance - its two meshes, over the four (direction, shift) passes of order 4;
     // the two difference formulas = numerical_hi + numerical_hj - those four great_arc_distance calls
    std::vector<Counted> out(nsamp), l(nsamp), p(nsamp);
    const Counted eps = 1e-3;
    Tally gad_all, gad_unused;
    for (int dir = 0; dir < 2; ++dir)
      for (int m = 1; m <= 2; ++m) {
        const Counted sh = m == 1 ? eps : 2.0 * eps;
        Counted* vp = shifted(dir == 0 ? iv.data() : jv.data(), dir == 0 ? nil : njl, sh);
        Counted* vm = shifted_m(dir == 0 ? iv.data() : jv.data(), dir == 0 ? nil : njl, sh);
        const Counted *i0 = dir == 0 ? vp : iv.data(), *i1 = dir == 0 ? vm : iv.data();
        const Counted *j0 = dir == 0 ? jv.data() : vp, *j1 = dir == 0 ? jv.data() : vm;
        Snap a0 = snap();
        ogo_dp_mesh(i0, nil, j0, njl, Ni, Nj_dp, lon0, lat0_dp, lon_dp, r_dp, l.data(), p.data());
        ogo_dp_mesh(i1, nil, j1, njl, Ni, Nj_dp, lon0, lat0_dp, lon_dp, r_dp, l.data(), p.data());
        Snap a1 = snap();
        ogo_great_arc_distance(j0, i0, j1, i1, njl, nil, Ni, Nj_dp, lon0, lat0_dp, lon_dp, r_dp, out.data());
        Snap a2 = snap();
        hav.all = hav.all + ((a2.all - a1.all) - (a1.all - a0.all));
        hav.unused = hav.unused + ((a2.unused - a1.unused) - (a1.unused - a0.unused));
        hav.count += nsamp;
        gad_all = gad_all + (a2.all - a1.all);
        gad_unused = gad_unused + (a2.unused - a1.unused);
        free(vp); free(vm);
      }
    Snap b0 = snap();
    ogo_numerical_h(0, jv.data(), njl, iv.data(), nil, Ni, Nj_dp, lon0, lat0_dp, lon_dp, r_dp, eps, 4, out.data());
    ogo_numerical_h(1, jv.data(), njl, iv.data(), nil, Ni, Nj_dp, lon0, lat0_dp, lon_dp, r_dp, eps, 4, out.data());
    hsample.add(b0, snap(), nsamp);
    fdc.all = hsample.all - gad_all; fdc.unused = hsample.unused - gad_unused; fdc.count = nsamp;
  }

  // ---------------- portable functions: C_f on the argument ranges of SURVEY Appendix C ------------------------------------
  struct Fn { const char* name; Counted (*f)(Counted); double lo, hi; };
  const Fn fns[] = {{"sin", p_sin, -6.3, 6.3}, {"cos", p_cos, -6.3, 6.3}, {"tan", p_tan, 0.0, 1.57}, {"arctan", p_atan, 0.0, 30.0},
                    {"arcsin", p_asin, -1.0, 1.0}, {"arccos", p_acos, -1.0, 1.0}};
  std::map<std::string, Unit> cf;
  for (const Fn& f : fns) {
    Unit u;
    const int n = 20000;
    for (int k = 0; k < n; ++k) {
      const Counted x = raw(f.lo + (f.hi - f.lo) * (k + 0.5) / n);
      Snap a = snap();
      Counted y = f.f(x);
      (void)y;
      u.add(a, snap(), 1);
    }
    cf[f.name] = u;
  }
  {
    Unit u;
    const int n = 20000;
    for (int k = 0; k < n; ++k) {
      const double t = 6.283185307179586 * (k + 0.5) / n;
      Snap a = snap();
      Counted y = p_atan2(raw(::sin(t)), raw(::cos(t)));
      (void)y;
      u.add(a, snap(), 1);
    }
    cf["arctan2"] = u;
  }

  // ---------------- report ----------------------------------------------------------------------------------------------
  printf("{\n");
  printf("  \"lattice\": {\"inverse_resolution\": %g, \"Ni\": %lld, \"bipolar_unique_samples\": %.0f, \"displaced_pole_unique_samples\": %.0f},\n",
         R, (long long)Ni, bp_full.count, nsamp);
  // the row / column parts are evaluated once per row / column, the sample-level mean of the reference includes them at every sample
  const Tally bp_ref = bp_full.mean_all(), bp_h = bp_ref - bp_row.mean_all() - bp_col.mean_all();
  const Tally bp_unused_h = bp_full.mean_unused() - bp_row.mean_unused() - bp_col.mean_unused();
  print_unit("bipolar_sample", "OGG:33-100 (metrics_only) + :131-132; row-only parts :41,44,46,75-78, column-only parts :42,43,45",
             bp_ref, bp_h, bp_unused_h, false);
  const Tally dv_ref = dp_full.mean_all() + unwrap.mean_all();
  const Tally dv_h = dv_ref - dp_row.mean_all() - dp_col.mean_all();
  const Tally dv_unused_h = dp_full.mean_unused() - dp_row.mean_unused() - dp_col.mean_unused();
  print_unit("dp_vertex", "OGG:447-467 + unwrap step :470-475; row-only part :448, column-only parts :451-452", dv_ref, dv_h,
             dv_unused_h, false);
  print_unit("haversine", "OGG:527-532", hav.mean_all(), hav.mean_all(), hav.mean_unused(), false);
  print_unit("fd_combine_x2", "OGG:539-546 and :553-561 (order 4), both directions", fdc.mean_all(), fdc.mean_all(), fdc.mean_unused(), false);
  // h-sample: reference = the two numerical_h calls as restated (8 meshes incl. their 1-D index laws, 4 haversines);
  // hoisted = 8 hoisted vertices + 4 haversines + the two difference formulas
  const Tally hs_ref = hsample.mean_all();
  const Tally hs_h = dv_h * 8.0 + hav.mean_all() * 4.0 + fdc.mean_all();
  const Tally hs_unused_h = dv_unused_h * 8.0 + hav.mean_unused() * 4.0 + fdc.mean_unused();
  print_unit("dp_hsample", "OGG:535-562 x2 (numerical_hi + numerical_hj, order 4): 8 vertices + 4 haversines + 2 difference formulas",
             hs_ref, hs_h, hs_unused_h, false);
  printf("  \"C_f\": {\n");
  size_t k = 0;
  for (auto& kv : cf) {
    printf("  \"%s\": {\n", kv.first.c_str());
    print_tally("reference", kv.second.mean_all(), false);
    print_tally("minimal", minimal_of(kv.second.mean_all(), kv.second.mean_unused()), true);
    printf("  }%s\n", ++k < cf.size() ? "," : "");
  }
  printf("  }\n}\n");
  free(j1d); free(i1d); free(j1q); free(i1q);
  return 0;
}

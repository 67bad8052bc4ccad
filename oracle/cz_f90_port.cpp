// =====================================================================================
// oracle/cz_f90_port.cpp  --  TEST INFRASTRUCTURE ONLY (parity oracle + timed CPU baseline)
//
// CPU restatement, in C++, of the Fortran 90 kernels of kenoogl/CubeZ that sit behind the
// `extern "C"` boundary of src/cz_cpp/cz_Ffunc.h:16-578.  The reference kernels are Fortran
// and this image has no Fortran front-end, so they cannot be compiled here; each function
// below follows the cited reference routine statement by statement (same operand order,
// true division, REAL arithmetic, no FMA contraction: build with -ffp-contract=off).
//
// NOT part of the product.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this.  The product path (cubez_b200/) never
// links or calls it.
//
// PARITY PINNING: the reference ships no golden vectors for this path.  What pins this oracle:
// (1) OUTPUTS OF THE REFERENCE'S OWN FORTRAN, executed through a mechanical Fortran -> C translation of the unmodified
// src/cz_f90/*.f90 (oracle/f90_to_c.py -> oracle/_ref/libczf90c_*.so; no Fortran compiler exists here): every routine
// below, float and double, equals it bit for bit in every array argument -- and in res / flop / the dot products
// with ORACLE_SUM=real -- and the reference's unmodified driver prints identical history files and console lines over
// either library (tests/test_f90c_pin.py; committed vectors tests/golden/f90c_digests.json, f90c_histories.json);
// (2) OUTPUTS OF THE REFERENCE'S OWN C++ PCR (src/cz_cpp/cz_pcr.cpp, compiled unmodified behind
// oracle/ref_pcr_harness.cpp): the stage loop below (line_pcr_stages) equals them bit for bit in double,
// tests/golden/ref_pcr_stages.json; (3) the reference's own UNMODIFIED C++ driver linked against this file
// (oracle/Makefile -> oracle/_ref/cz_ref_*), reproducing the analytic-solution check `Error max` of
// cz_Evaluate.cpp:550-563; (4) the known answers of example/tri/test1.cpp, test3.cpp, test5.cpp for the tridiagonal
// line solves and doc/Memo.md:131-134 (pcr_eda history == pcr history); (5) an independent numpy restatement
// (oracle/np_restatement.py).  Not covered by any of these: the reference's OpenMP reduction ORDER (it is order
// dependent) -- hence the two reduction modes below.
//
// Reduction mode: env ORACLE_SUM=double (default) -> REAL products, double
// accumulation, j-plane partials summed in j order (deterministic for any thread count);
// ORACLE_SUM=real -> one REAL accumulator, serial j,i,k order (what the 1-thread reference
// does; used only to bound the reference's own drift).
//
// Racy reference kernels (psor, pcr, pcr_esa, pcr_eda: in-place Gauss-Seidel inside an
// OpenMP loop, cz_solver.f90:243-264, 711-871) are restated in their 1-thread order.
// =====================================================================================
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef _REAL_IS_DOUBLE_
typedef double REAL;
#else
typedef float REAL;
#endif

namespace {

// Fortran array p(1-g:sz(3)+g, 1-g:sz(1)+g, 1-g:sz(2)+g): index order (k,i,j), k contiguous
// (cz_solver.f90:218).
struct Grid {
  int ix, jx, kx, g;
  long nk, ni;
  Grid(const int* sz, int gg) : ix(sz[0]), jx(sz[1]), kx(sz[2]), g(gg),
                                nk(sz[2] + 2L * gg), ni(sz[0] + 2L * gg) {}
  inline long at(int k, int i, int j) const {
    return (long)(k + g - 1) + nk * ((long)(i + g - 1) + ni * (long)(j + g - 1));
  }
  inline long total() const { return nk * ni * (long)(jx + 2L * g); }
};

struct Range {
  int ist, ied, jst, jed, kst, ked;
  explicit Range(const int* idx)
      : ist(idx[0]), ied(idx[1]), jst(idx[2]), jed(idx[3]), kst(idx[4]), ked(idx[5]) {}
  double cells() const {
    return (double)(ied - ist + 1) * (double)(jed - jst + 1) * (double)(ked - kst + 1);
  }
};

int g_sum_mode = -1;     // -1: read ORACLE_SUM on first use; 0 double, 1 real (oracle_set_sum_mode)
bool real_mode() {
  if (g_sum_mode < 0) {
    const char* e = std::getenv("ORACLE_SUM");
    g_sum_mode = (e && !strcmp(e, "real")) ? 1 : 0;
  }
  return g_sum_mode == 1;
}

struct Acc {
  double d = 0.0;
  REAL r = 0;
  bool rm;
  Acc() : rm(real_mode()) {}
  inline void add(REAL v) { if (rm) r = r + v; else d += (double)v; }
  double value() const { return rm ? (double)r : d; }
};

// Parallel over j-planes with a deterministic reduction.
template <class F>
double reduce_planes(int jst, int jed, bool allow_parallel, F f) {
  if (jed < jst) return 0.0;
  if (real_mode() || !allow_parallel) {
    Acc a;
    for (int j = jst; j <= jed; ++j) f(j, a);
    return a.value();
  }
  std::vector<Acc> part(jed - jst + 1);
#pragma omp parallel for schedule(static)
  for (int j = jst; j <= jed; ++j) f(j, part[j - jst]);
  double s = 0.0;
  for (auto& a : part) s += a.d;
  return s;
}

template <class F>
void for_planes(int jst, int jed, F f) {
#pragma omp parallel for schedule(static)
  for (int j = jst; j <= jed; ++j) f(j);
}

inline REAL rsin(REAL x) { return std::sin(x); }
inline REAL rsinh(REAL x) { return std::sinh(x); }
inline REAL rasin(REAL x) { return std::asin(x); }
inline REAL rsqrt_(REAL x) { return std::sqrt(x); }

// ---------------------------------------------------------------------------------
// PCR line solver work space.  Index k in [kst-pad, ked+pad] maps to [0, n+2*pad).
// Every variant of the reference treats out-of-range entries as zero: pcr_rb/pcr clamp the
// index to kst-1 / ked+1 whose entries are 0 (cz_solver.f90:590-591, 724-736), the
// _esa/_eda variants zero-pad (cz_solver.f90:919-929, 1327-1345).  One padded layout with
// pad >= 2^pn therefore reproduces all of them.
// ---------------------------------------------------------------------------------
struct LineWork {
  int kst, ked, pad;
  std::vector<REAL> a, c, d, a1, c1, d1;
  LineWork(int kst_, int ked_, int pn) : kst(kst_), ked(ked_), pad(1 << pn) {
    size_t n = (size_t)(ked - kst + 1) + 2 * (size_t)pad;
    a.assign(n, 0); c.assign(n, 0); d.assign(n, 0);
    a1.assign(n, 0); c1.assign(n, 0); d1.assign(n, 0);
  }
  inline int o(int k) const { return k - kst + pad; }
};

// cz_solver.f90:556-564 (pcr_rb), 724-736 (pcr): a(kst)=0, a(kst+1..ked)=-r, c(kst..ked-1)=-r, c(ked)=0
void line_coef(LineWork& w, REAL r) {
  std::fill(w.a.begin(), w.a.end(), (REAL)0);
  std::fill(w.c.begin(), w.c.end(), (REAL)0);
  for (int k = w.kst + 1; k <= w.ked; ++k) w.a[w.o(k)] = -r;
  for (int k = w.kst; k <= w.ked - 1; ++k) w.c[w.o(k)] = -r;
}

// cz_solver.f90:584-608: `stages` PCR stages, stride 2^(p-1), Jacobi-style (a1,c1,d1 then copy).
void line_pcr_stages(LineWork& w, int stages) {
  for (int p = 1; p <= stages; ++p) {
    const int s = 1 << (p - 1);
    for (int k = w.kst; k <= w.ked; ++k) {
      const int o = w.o(k);
      const REAL ap = w.a[o];
      const REAL cp = w.c[o];
      const REAL e = (REAL)1.0 / ((REAL)1.0 - ap * w.c[o - s] - cp * w.a[o + s]);
      w.a1[o] = -e * ap * w.a[o - s];
      w.c1[o] = -e * cp * w.c[o + s];
      w.d1[o] = e * (w.d[o] - ap * w.d[o - s] - cp * w.d[o + s]);
    }
    for (int k = w.kst; k <= w.ked; ++k) {
      const int o = w.o(k);
      w.a[o] = w.a1[o]; w.c[o] = w.c1[o]; w.d[o] = w.d1[o];
    }
  }
}

// cz_solver.f90:612-630: last stage as a direct 2x2 solve, s = 2^(pn-1)
void line_last_2x2(LineWork& w, int pn) {
  const int s = 1 << (pn - 1);
  for (int k = w.kst; k <= w.kst + s - 1; ++k) {
    const int o = w.o(k), orr = o + s;
    const REAL cc1 = w.c[o];
    const REAL aa2 = w.a[orr];
    const REAL f1 = w.d[o];
    const REAL f2 = w.d[orr];
    const REAL jj = (REAL)1.0 / ((REAL)1.0 - aa2 * cc1);
    const REAL dd1 = (f1 - cc1 * f2) * jj;
    const REAL dd2 = (f2 - aa2 * f1) * jj;
    w.d1[o] = dd1;
    w.d1[orr] = dd2;
  }
}

// cz_solver.f90:788-844: last two stages as a direct 4x4 solve (Cramer), s = 2^(pn-2)
void line_last_4x4(LineWork& w, int pn) {
  const int s = 1 << (pn - 2);
  for (int k = w.kst; k <= w.kst + s - 1; ++k) {
    const int o = w.o(k), ol = o + s, om = o + 2 * s, orr = o + 3 * s;
    const REAL cc1 = w.c[o], cc2 = w.c[ol], cc3 = w.c[om];
    const REAL aa2 = w.a[ol], aa3 = w.a[om], aa4 = w.a[orr];
    const REAL dd1 = w.d[o], dd2 = w.d[ol], dd3 = w.d[om], dd4 = w.d[orr];
    const REAL one = (REAL)1.0;
    const REAL inv_detA = one / (one - aa4 * cc3 - aa3 * cc2 - aa2 * cc1 * (one - cc3 * aa4));
    const REAL detA1 = -cc3 * (aa4 * dd1 + cc1 * cc2 * dd4 - aa4 * cc1 * dd2)
                     + dd1 + cc1 * cc2 * dd3 - aa3 * cc2 * dd1 - cc1 * dd2;
    const REAL detA2 = dd2 + cc2 * cc3 * dd4 - aa4 * cc3 * dd2 - cc2 * dd3
                     - aa2 * (dd1 - aa4 * cc3 * dd1);
    const REAL detA3 = dd3 - cc3 * dd4 - aa3 * dd2
                     - aa2 * (cc1 * dd3 - cc1 * cc3 * dd4 - aa3 * dd1);
    const REAL detA4 = dd4 + aa3 * aa4 * dd2 - aa4 * dd3 - aa3 * cc2 * dd4
                     - aa2 * (cc1 * dd4 + aa3 * aa4 * dd1 - aa4 * cc1 * dd3);
    // written in this order; with zero padding the out-of-range stores are never read back
    w.d1[o] = detA1 * inv_detA;
    w.d1[ol] = detA2 * inv_detA;
    w.d1[om] = detA3 * inv_detA;
    w.d1[orr] = detA4 * inv_detA;
  }
}

// cz_solver.f90:569-579: source term from the four in-plane neighbours + Dirichlet ends
void line_source(LineWork& w, const Grid& G, const REAL* x, const REAL* rhs, const REAL* msk,
                 int i, int j, REAL r) {
  for (int k = w.kst; k <= w.ked; ++k) {
    w.d[w.o(k)] = ((x[G.at(k, i, j - 1)] + x[G.at(k, i, j + 1)] + x[G.at(k, i - 1, j)] +
                    x[G.at(k, i + 1, j)] - rhs[G.at(k, i, j)]) * r) * msk[G.at(k, i, j)];
  }
  const int kst = w.kst, ked = w.ked;
  w.d[w.o(kst)] = (w.d[w.o(kst)] + x[G.at(kst - 1, i, j)] * r) * msk[G.at(kst, i, j)];
  w.d[w.o(ked)] = (w.d[w.o(ked)] + x[G.at(ked + 1, i, j)] * r) * msk[G.at(ked, i, j)];
}

// cz_solver.f90:642-648: relaxation, dst may alias x
void line_relax(LineWork& w, const Grid& G, const REAL* x, REAL* dst, const REAL* msk,
                int i, int j, REAL omg, Acc& acc) {
  for (int k = w.kst; k <= w.ked; ++k) {
    const long q = G.at(k, i, j);
    const REAL pp = x[q];
    const REAL dp = (w.d1[w.o(k)] - pp) * omg * msk[q];
    dst[q] = pp + dp;
    acc.add(dp * dp);
  }
}

enum LastStage { LAST_2X2, LAST_4X4 };

void solve_line(LineWork& w, const Grid& G, REAL* x, REAL* dst, const REAL* rhs,
                const REAL* msk, int i, int j, int pn, REAL omg, LastStage ls, Acc& acc) {
  const REAL r = (REAL)1.0 / (REAL)6.0;
  line_coef(w, r);
  std::fill(w.d.begin(), w.d.end(), (REAL)0);
  line_source(w, G, x, rhs, msk, i, j, r);
  if (ls == LAST_2X2) { line_pcr_stages(w, pn - 1); line_last_2x2(w, pn); }
  else                { line_pcr_stages(w, pn - 2); line_last_4x4(w, pn); }
  line_relax(w, G, x, dst, msk, i, j, omg, acc);
}

double ipow2(int e) { return (double)(1L << e); }

}  // namespace

extern "C" {

// ------------------------------------------------------------------ cz_solver.f90:22-191
void bc_k_(int* sz, int* g, REAL* p, REAL* dh, REAL* org, int* nID) {
  const Grid G(sz, *g);
  const int ix = G.ix, jx = G.jx, kx = G.kx;
  const REAL pi = (REAL)2.0 * rasin((REAL)1.0);
  const REAL h = *dh;
  enum { I_MINUS = 0, I_PLUS, J_MINUS, J_PLUS, K_MINUS, K_PLUS };  // cz_fparam.fi:10-16
  if (nID[K_MINUS] < 0)
    for_planes(1, jx, [&](int j) {
      for (int i = 1; i <= ix; ++i) {
        const REAL x = org[0] + h * (REAL)(i - 1);
        const REAL y = org[1] + h * (REAL)(j - 1);
        p[G.at(1, i, j)] = rsin(pi * x) * rsin(pi * y);
      }
    });
  if (nID[K_PLUS] < 0)
    for_planes(1, jx, [&](int j) {
      for (int i = 1; i <= ix; ++i) {
        const REAL x = org[0] + h * (REAL)(i - 1);
        const REAL y = org[1] + h * (REAL)(j - 1);
        p[G.at(kx, i, j)] = rsin(pi * x) * rsin(pi * y);
      }
    });
  if (nID[I_MINUS] < 0)
    for_planes(1, jx, [&](int j) { for (int k = 1; k <= kx; ++k) p[G.at(k, 1, j)] = 0; });
  if (nID[I_PLUS] < 0)
    for_planes(1, jx, [&](int j) { for (int k = 1; k <= kx; ++k) p[G.at(k, ix, j)] = 0; });
  if (nID[J_MINUS] < 0)
    for_planes(1, ix, [&](int i) { for (int k = 1; k <= kx; ++k) p[G.at(k, i, 1)] = 0; });
  if (nID[J_PLUS] < 0)
    for_planes(1, ix, [&](int i) { for (int k = 1; k <= kx; ++k) p[G.at(k, i, jx)] = 0; });
}

// ------------------------------------------------------------------ cz_solver.f90:207-269
// Lexicographic point SOR; racy under OpenMP in the reference -> 1-thread order here.
void psor_(REAL* p, int* sz, int* idx, int* g, REAL* cf, REAL* omg, REAL* b, double* res,
           double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL c1 = cf[0], c2 = cf[1], c3 = cf[2], c4 = cf[3], c5 = cf[4], c6 = cf[5], dd = cf[6];
  const REAL w = *omg;
  *flop += 18.0 * R.cells();
  Acc acc;
  for (int j = R.jst; j <= R.jed; ++j)
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) {
        const REAL pp = p[G.at(k, i, j)];
        const REAL bb = b[G.at(k, i, j)];
        const REAL ss = c1 * p[G.at(k, i + 1, j)] + c2 * p[G.at(k, i - 1, j)] +
                        c3 * p[G.at(k, i, j + 1)] + c4 * p[G.at(k, i, j - 1)] +
                        c5 * p[G.at(k + 1, i, j)] + c6 * p[G.at(k - 1, i, j)];
        const REAL dp = ((ss - bb) / dd - pp) * w;
        p[G.at(k, i, j)] = pp + dp;
        acc.add(dp * dp);
      }
  *res += acc.value();
}

// ------------------------------------------------------------------ cz_solver.f90:284-387
void jacobi_(REAL* p, int* sz, int* idx, int* g, REAL* cf, REAL* omg, REAL* b, double* res,
             REAL* wk2, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL c1 = cf[0], c2 = cf[1], c3 = cf[2], c4 = cf[3], c5 = cf[4], c6 = cf[5], dd = cf[6];
  const REAL w = *omg;
  *flop += 18.0 * R.cells();
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const REAL* pc = p + G.at(0, i, j);
      const REAL* pe = p + G.at(0, i + 1, j);
      const REAL* pw = p + G.at(0, i - 1, j);
      const REAL* pn = p + G.at(0, i, j + 1);
      const REAL* ps = p + G.at(0, i, j - 1);
      const REAL* bc = b + G.at(0, i, j);
      REAL* wc = wk2 + G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) {
        const REAL pp = pc[k];
        const REAL bb = bc[k];
        const REAL ss = c1 * pe[k] + c2 * pw[k] + c3 * pn[k] + c4 * ps[k] + c5 * pc[k + 1] +
                        c6 * pc[k - 1];
        const REAL dp = ((ss - bb) / dd - pp) * w;
        wc[k] = pp + dp;
        acc.add(dp * dp);
      }
    }
  });
  for_planes(R.jst, R.jed, [&](int j) {  // cz_solver.f90:369-375 copy-back
    for (int i = R.ist; i <= R.ied; ++i) {
      const long q = G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) p[q + k] = wk2[q + k];
    }
  });
  *res += s;
}

// ------------------------------------------------------------------ cz_solver.f90:404-493
// One colour of red-black SOR: k = kst+mod(i+j+kp,2), ked, 2 with kp = ofst+color (:419,466)
void psor2sma_core_(REAL* p, int* sz, int* idx, int* g, REAL* cf, int* ofst, int* color,
                    REAL* omg, REAL* b, double* res, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL c1 = cf[0], c2 = cf[1], c3 = cf[2], c4 = cf[3], c5 = cf[4], c6 = cf[5], dd = cf[6];
  const REAL w = *omg;
  const int kp = *ofst + *color;
  *flop += 18.0 * 0.5 * R.cells();
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    for (int i = R.ist; i <= R.ied; ++i) {
      REAL* pc = p + G.at(0, i, j);
      const REAL* pe = p + G.at(0, i + 1, j);
      const REAL* pw = p + G.at(0, i - 1, j);
      const REAL* pn = p + G.at(0, i, j + 1);
      const REAL* ps = p + G.at(0, i, j - 1);
      const REAL* bc = b + G.at(0, i, j);
      for (int k = R.kst + ((i + j + kp) % 2 + 2) % 2; k <= R.ked; k += 2) {
        const REAL pp = pc[k];
        const REAL bb = bc[k];
        const REAL ss = c1 * pe[k] + c2 * pw[k] + c3 * pn[k] + c4 * ps[k] + c5 * pc[k + 1] +
                        c6 * pc[k - 1];
        const REAL dp = ((ss - bb) / dd - pp) * w;
        pc[k] = pp + dp;
        acc.add(dp * dp);
      }
    }
  });
  *res += s;
}

// ------------------------------------------------------------------ cz_solver.f90:497-662
// Red-black line SOR, PCR pn-1 stages + 2x2.  Colour test mod(i+j,2)==color (:549); `ofst`
// is ignored by the reference (:533, App. A).  res reduced directly in double (:541,647).
void pcr_rb_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, REAL* x, REAL* msk,
             REAL* rhs, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res,
             double* flop) {
  (void)ofst;
  const Grid G(sz, *g);
  const Range R(idx);
  const double nk = R.ked - R.kst + 1;
  *flop += (double)(R.jed - R.jst + 1) * (R.ied - R.ist + 1) *
           (nk * 6.0 + nk * (*pn - 1) * 14.0 + ipow2(*pn - 1) * 9.0 + nk * 6.0 + 6.0) * 0.5;
  if (real_mode()) {
    // 1-thread reference order: `res = res + dp*dp` (:647) adds every REAL product straight to the double
    // argument, line after line -- no REAL partial sum in this routine (res1 is commented out, :646,659)
    Acc acc;
    acc.rm = false;
    acc.d = *res;
    LineWork w(R.kst, R.ked, *pn);
    for (int j = R.jst; j <= R.jed; ++j)
      for (int i = R.ist; i <= R.ied; ++i) {
        if (((i + j) % 2 + 2) % 2 != *color) continue;
        solve_line(w, G, x, x, rhs, msk, i, j, *pn, *omg, LAST_2X2, acc);
      }
    *res = acc.d;
    return;
  }
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    LineWork w(R.kst, R.ked, *pn);
    for (int i = R.ist; i <= R.ied; ++i) {
      if (((i + j) % 2 + 2) % 2 != *color) continue;
      solve_line(w, G, x, x, rhs, msk, i, j, *pn, *omg, LAST_2X2, acc);
    }
  });
  *res += s;
}

// ------------------------------------------------------------------ cz_solver.f90:1261-1469
void pcr_rb_esa_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, int* s_, REAL* x,
                 REAL* msk, REAL* rhs, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg,
                 double* res, double* flop) {
  (void)ofst; (void)s_;
  const Grid G(sz, *g);
  const Range R(idx);
  const double nk = R.ked - R.kst + 1;
  *flop += (double)(R.jed - R.jst + 1) * (R.ied - R.ist + 1) *
           (nk * 6.0 + nk * (*pn - 2) * 14.0 + ipow2(*pn - 2) * 78.0 + nk * 6.0 + 6.0) * 0.5;
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    LineWork w(R.kst, R.ked, *pn);
    for (int i = R.ist; i <= R.ied; ++i) {
      if (((i + j) % 2 + 2) % 2 != *color) continue;
      solve_line(w, G, x, x, rhs, msk, i, j, *pn, *omg, LAST_4X4, acc);
    }
  });
  *res += s;
}

// ------------------------------------------------------------------ cz_solver.f90:666-878
// All lines in (j,i) order, in place (Gauss-Seidel between lines): 1-thread order.
void pcr_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL*, REAL*,
          REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const double nk = R.ked - R.kst + 1;
  *flop += (double)(R.jed - R.jst + 1) * (R.ied - R.ist + 1) *
           (nk * 6.0 + nk * (*pn - 2) * 14.0 + ipow2(*pn - 2) * 74.0 + nk * 6.0 + 6.0);
  Acc acc;
  LineWork w(R.kst, R.ked, *pn);
  for (int j = R.jst; j <= R.jed; ++j)
    for (int i = R.ist; i <= R.ied; ++i)
      solve_line(w, G, x, x, rhs, msk, i, j, *pn, *omg, LAST_4X4, acc);
  *res += acc.value();
}

// ------------------------------------------------------------------ cz_solver.f90:883-1045
void pcr_eda_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL*, REAL*,
              REAL*, REAL* omg, double* res, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const double nk = R.ked - R.kst + 1;
  *flop += (double)(R.jed - R.jst + 1) * (R.ied - R.ist + 1) *
           (nk * 6.0 + nk * (*pn - 1) * 14.0 + ipow2(*pn - 1) * 9.0 + nk * 6.0 + 6.0);
  Acc acc;
  LineWork w(R.kst, R.ked, *pn);
  for (int j = R.jst; j <= R.jed; ++j)
    for (int i = R.ist; i <= R.ied; ++i)
      solve_line(w, G, x, x, rhs, msk, i, j, *pn, *omg, LAST_2X2, acc);
  *res += acc.value();
}

// ------------------------------------------------------------------ cz_solver.f90:1050-1257
void pcr_esa_(int* sz, int* idx, int* g, int* pn, int* s_, REAL* x, REAL* msk, REAL* rhs,
              REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, double* flop) {
  (void)s_;
  const Grid G(sz, *g);
  const Range R(idx);
  const double nk = R.ked - R.kst + 1;
  *flop += (double)(R.jed - R.jst + 1) * (R.ied - R.ist + 1) *
           (nk * 6.0 + nk * (*pn - 2) * 14.0 + ipow2(*pn - 2) * 78.0 + nk * 6.0 + 6.0);
  Acc acc;
  LineWork w(R.kst, R.ked, *pn);
  for (int j = R.jst; j <= R.jed; ++j)
    for (int i = R.ist; i <= R.ied; ++i)
      solve_line(w, G, x, x, rhs, msk, i, j, *pn, *omg, LAST_4X4, acc);
  *res += acc.value();
}

// ------------------------------------------------------------------ cz_solver.f90:1473-1676
// Line Jacobi: source to `src` (:1521-1531), line solve reading old x, result to `wrk`
// (:1632-1637), copy back (:1659-1665).
void pcr_j_esa_(int* sz, int* idx, int* g, int* pn, int* s_, REAL* x, REAL* msk, REAL* rhs,
                REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* src, REAL* wrk, REAL* omg,
                double* res, double* flop) {
  (void)s_;
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL r = (REAL)1.0 / (REAL)6.0;
  const double nk = R.ked - R.kst + 1;
  *flop += (double)(R.jed - R.jst + 1) * (R.ied - R.ist + 1) *
           (nk * 6.0 + nk * (*pn - 1) * 14.0 + ipow2(*pn - 1) * 9.0 + nk * 6.0 + 6.0);
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k)
        src[G.at(k, i, j)] = ((x[G.at(k, i, j - 1)] + x[G.at(k, i, j + 1)] +
                               x[G.at(k, i - 1, j)] + x[G.at(k, i + 1, j)] -
                               rhs[G.at(k, i, j)]) * r) * msk[G.at(k, i, j)];
  });
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    LineWork w(R.kst, R.ked, *pn);
    for (int i = R.ist; i <= R.ied; ++i) {
      line_coef(w, r);
      std::fill(w.d.begin(), w.d.end(), (REAL)0);
      for (int k = R.kst; k <= R.ked; ++k) w.d[w.o(k)] = src[G.at(k, i, j)];
      w.d[w.o(R.kst)] = (w.d[w.o(R.kst)] + x[G.at(R.kst - 1, i, j)] * r) * msk[G.at(R.kst, i, j)];
      w.d[w.o(R.ked)] = (w.d[w.o(R.ked)] + x[G.at(R.ked + 1, i, j)] * r) * msk[G.at(R.ked, i, j)];
      line_pcr_stages(w, *pn - 1);
      line_last_2x2(w, *pn);
      line_relax(w, G, x, wrk, msk, i, j, *omg, acc);
    }
  });
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) x[G.at(k, i, j)] = wrk[G.at(k, i, j)];
  });
  *res += s;
}

// ------------------------------------------------------------------ cz_blas.f90:24-103
void imask_k_(REAL* x, int* sz, int* idx, int* g) {
  const Grid G(sz, *g);
  const Range R(idx);
  const long n = G.total();
#pragma omp parallel for schedule(static)
  for (long q = 0; q < n; ++q) x[q] = 0;
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) x[G.at(k, i, j)] = (REAL)1.0;
  });
}

// ------------------------------------------------------------------ cz_blas.f90:112-149
void blas_clear_(REAL* x, int* sz, int* g) {
  const Grid G(sz, *g);
  const long n = G.total();
#pragma omp parallel for schedule(static)
  for (long q = 0; q < n; ++q) x[q] = 0;
}

// ------------------------------------------------------------------ cz_blas.f90:159-195
void blas_copy_(REAL* y, REAL* x, int* sz, int* g) {
  const Grid G(sz, *g);
  const long n = G.total();
#pragma omp parallel for schedule(static)
  for (long q = 0; q < n; ++q) y[q] = x[q];
}

// ------------------------------------------------------------------ cz_blas.f90:205-242
void blas_copy_in_(REAL* y, REAL* x, int* sz, int* g) {
  const Grid G(sz, *g);
  for_planes(1, G.jx, [&](int j) {
    for (int i = 1; i <= G.ix; ++i)
      for (int k = 1; k <= G.kx; ++k) y[G.at(k, i, j)] = x[G.at(k, i, j)];
  });
}

// ------------------------------------------------------------------ cz_blas.f90:255-308
void blas_triad_(REAL* z, REAL* x, REAL* y, REAL* a, int* sz, int* idx, int* g, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL aa = *a;
  *flop += 2.0 * R.cells();
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const long q = G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) z[q + k] = aa * x[q + k] + y[q + k];
    }
  });
}

// ------------------------------------------------------------------ cz_blas.f90:320-373
void blas_dot1_(REAL* r, REAL* p, int* sz, int* idx, int* g, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  *flop += 2.0 * R.cells();
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const long q0 = G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) { const REAL q = p[q0 + k]; acc.add(q * q); }
    }
  });
  *r = (REAL)s;
}

// ------------------------------------------------------------------ cz_blas.f90:386-437
void blas_dot2_(REAL* r, REAL* p, REAL* q, int* sz, int* idx, int* g, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  *flop += 2.0 * R.cells();
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const long q0 = G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) acc.add(p[q0 + k] * q[q0 + k]);
    }
  });
  *r = (REAL)s;
}

// ------------------------------------------------------------------ cz_blas.f90:452-502
void blas_bicg_1_(REAL* p, REAL* r, REAL* q, REAL* beta, REAL* omg, int* sz, int* idx, int* g,
                  double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL be = *beta, om = *omg;
  *flop += 4.0 * R.cells();
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const long q0 = G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k)
        p[q0 + k] = r[q0 + k] + be * (p[q0 + k] - om * q[q0 + k]);
    }
  });
}

// ------------------------------------------------------------------ cz_blas.f90:517-566
void blas_bicg_2_(REAL* z, REAL* x, REAL* y, REAL* a, REAL* b, int* sz, int* idx, int* g,
                  double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL aa = *a, bb = *b;
  *flop += 4.0 * R.cells();
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const long q0 = G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k)
        z[q0 + k] = aa * x[q0 + k] + bb * y[q0 + k] + z[q0 + k];
    }
  });
}

// ------------------------------------------------------------------ cz_blas.f90:579-644
void blas_calc_ax_(REAL* ap, REAL* p, int* sz, int* idx, int* g, REAL* cf, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL c1 = cf[0], c2 = cf[1], c3 = cf[2], c4 = cf[3], c5 = cf[4], c6 = cf[5], dd = cf[6];
  *flop += 13.0 * R.cells();
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const REAL* pc = p + G.at(0, i, j);
      const REAL* pe = p + G.at(0, i + 1, j);
      const REAL* pw = p + G.at(0, i - 1, j);
      const REAL* pn = p + G.at(0, i, j + 1);
      const REAL* ps = p + G.at(0, i, j - 1);
      REAL* out = ap + G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) {
        const REAL ss = c1 * pe[k] + c2 * pw[k] + c3 * pn[k] + c4 * ps[k] + c5 * pc[k + 1] +
                        c6 * pc[k - 1];
        out[k] = (ss - dd * pc[k]);
      }
    }
  });
}

// ------------------------------------------------------------------ cz_blas.f90:658-723
void blas_calc_rk_(REAL* r, REAL* p, REAL* b, int* sz, int* idx, int* g, REAL* cf,
                   double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL c1 = cf[0], c2 = cf[1], c3 = cf[2], c4 = cf[3], c5 = cf[4], c6 = cf[5], dd = cf[6];
  *flop += 14.0 * R.cells();
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i) {
      const REAL* pc = p + G.at(0, i, j);
      const REAL* pe = p + G.at(0, i + 1, j);
      const REAL* pw = p + G.at(0, i - 1, j);
      const REAL* pn = p + G.at(0, i, j + 1);
      const REAL* ps = p + G.at(0, i, j - 1);
      const REAL* bc = b + G.at(0, i, j);
      REAL* out = r + G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) {
        const REAL ss = c1 * pe[k] + c2 * pw[k] + c3 * pn[k] + c4 * ps[k] + c5 * pc[k + 1] +
                        c6 * pc[k - 1];
        out[k] = (bc[k] - (ss - dd * pc[k]));
      }
    }
  });
}

// ------------------------------------------------------------------ cz_blas.f90:947-1039
// Called unconditionally by the reference driver (cz_Evaluate.cpp:369); result used only by
// the MAF solvers.
void search_pivot_(REAL* pvt, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z) {
  const Grid G(sz, *g);
  const Range R(idx);
  // X(-1:sz(1)+2): Fortran index i -> C offset i+1
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) {
        const REAL XG = (REAL)0.5 * (X[i + 2] - X[i]);
        const REAL YE = (REAL)0.5 * (Y[j + 2] - Y[j]);
        const REAL ZT = (REAL)0.5 * (Z[k + 2] - Z[k]);
        const REAL XGG = X[i + 2] - (REAL)2.0 * X[i + 1] + X[i];
        const REAL YEE = Y[j + 2] - (REAL)2.0 * Y[j + 1] + Y[j];
        const REAL ZTT = Z[k + 2] - (REAL)2.0 * Z[k + 1] + Z[k];
        const REAL YJA = XG * YE * ZT;
        const REAL YJAI = (REAL)1.0 / YJA;
        const REAL GX = YE * ZT * YJAI;
        const REAL EY = XG * ZT * YJAI;
        const REAL TZ = XG * YE * YJAI;
        const REAL C1 = GX * GX, C2 = EY * EY, C3 = TZ * TZ;
        const REAL C7 = -XGG * C1 * GX, C8 = -YEE * C2 * EY, C9 = -ZTT * C3 * TZ;
        const REAL s1 = std::fabs(C1 + (REAL)0.5 * C7), s2 = std::fabs(C1 - (REAL)0.5 * C7);
        const REAL s3 = std::fabs(C2 + (REAL)0.5 * C8), s4 = std::fabs(C2 - (REAL)0.5 * C8);
        const REAL s5 = std::fabs(C3 + (REAL)0.5 * C9), s6 = std::fabs(C3 - (REAL)0.5 * C9);
        const REAL s7 = std::fabs((REAL)2.0 * (C1 + C2 + C3));
        const REAL ss = std::max(std::max(std::max(s1, s2), std::max(s3, s4)),
                                 std::max(std::max(s5, s6), s7));
        pvt[G.at(k, i, j)] = (REAL)1.0 / ss;
      }
  });
}

// ------------------------------------------------------------------ cz_utility.f90:17-44
// Compiled out in the reference unless _aurora_==1: no-op.
void fileout_t_(int*, int*, REAL*, REAL*, REAL*, char*) {}

// ------------------------------------------------------------------ cz_utility.f90:52-82
void exact_t_(int* sz, int* g, REAL* e, REAL* dh, REAL* org) {
  const Grid G(sz, *g);
  const REAL r2 = rsqrt_((REAL)2.0);
  const REAL pi = (REAL)2.0 * rasin((REAL)1.0);
  const REAL h = *dh;
  for_planes(1, G.jx, [&](int j) {
    for (int i = 1; i <= G.ix; ++i)
      for (int k = 1; k <= G.kx; ++k) {
        const REAL x = org[0] + h * (REAL)(i - 1);
        const REAL y = org[1] + h * (REAL)(j - 1);
        const REAL z = org[2] + h * (REAL)(k - 1);
        e[G.at(k, i, j)] = rsin(pi * x) * rsin(pi * y) / rsinh(r2 * pi) *
                           (rsinh(r2 * pi * z) - rsinh(r2 * pi * (z - (REAL)1.0)));
      }
  });
}

// ------------------------------------------------------------------ cz_utility.f90:86-129
void err_t_(int* sz, int* idx, int* g, double* d, REAL* p, REAL* e, int* loc) {
  const Grid G(sz, *g);
  const Range R(idx);
  double dm = 0.0;
  loc[0] = loc[1] = loc[2] = -1;
  for (int j = R.jst; j <= R.jed; ++j)
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) {
        const REAL r = p[G.at(k, i, j)] - e[G.at(k, i, j)];
        e[G.at(k, i, j)] = r;
        const double q = std::fabs((double)r);
        if (dm < q) { dm = q; loc[0] = i; loc[1] = j; loc[2] = k; }
      }
  *d = dm;
}

// ------------------------------------------------------------------ src/cz_f90/obsolete.f90:65-84
// tdma_p: Thomas algorithm, unit diagonal; used by tests as an independent check of the PCR
// line solves.  d(nx) in/out, a lower, c upper, w work.
void tdma_p_(int* nx, REAL* d, REAL* a, REAL* c, REAL* w) {
  const int n = *nx;
  REAL e;
  w[0] = c[0];
  for (int i = 1; i < n; ++i) {
    e = (REAL)1.0 / ((REAL)1.0 - a[i] * w[i - 1]);
    w[i] = e * c[i];
    d[i] = (d[i] - a[i] * d[i - 1]) * e;
  }
  for (int i = n - 2; i >= 0; --i) d[i] = d[i] - w[i] * d[i + 1];
}

// ------------------------------------------------------------------ src/cz_f90/obsolete.f90:40-61
// tdma_0: Thomas algorithm with a general diagonal b (known answer example/tri/test1.cpp:25-31).
void tdma_0_(int* nx, REAL* d, REAL* a, REAL* b, REAL* c, REAL* w) {
  const int n = *nx;
  d[0] = d[0] / b[0];
  w[0] = c[0] / b[0];
  for (int i = 1; i < n; ++i) {
    const REAL e = (REAL)1.0 / (b[i] - a[i] * w[i - 1]);
    w[i] = e * c[i];
    d[i] = (d[i] - a[i] * d[i - 1]) * e;
  }
  for (int i = n - 2; i >= 0; --i) d[i] = d[i] - w[i] * d[i + 1];
}

// Stand-alone PCR line solve on a normalised tridiagonal system (unit diagonal) of length n,
// a(0)=0, c(n-1)=0: exposes the line solver of pcr_rb (2x2 last stage, mode 0) / pcr (4x4,
// mode 1) for the known-answer tests of example/tri/test3.cpp, test5.cpp.
void oracle_pcr_line_(int* n_, int* mode, REAL* a, REAL* c, REAL* d, REAL* xout) {
  const int n = *n_;
  int pn = -1, b = 1;
  for (int i = 1; i < 20; ++i) { b *= 2; if (n < b) { pn = i; break; } }  // cz.h:293-300
  LineWork w(1, n, pn);
  for (int k = 1; k <= n; ++k) { w.a[w.o(k)] = a[k - 1]; w.c[w.o(k)] = c[k - 1]; w.d[w.o(k)] = d[k - 1]; }
  if (*mode == 0) { line_pcr_stages(w, pn - 1); line_last_2x2(w, pn); }
  else            { line_pcr_stages(w, pn - 2); line_last_4x4(w, pn); }
  for (int k = 1; k <= n; ++k) xout[k - 1] = w.d1[w.o(k)];
}

// The PCR reduction stages alone, for pinning against the reference's own C++ PCR (src/cz_cpp/cz_pcr.cpp:42-96,
// CZ::pcr / CZ::pcr_kernel -- the same recurrences as cz_solver.f90:584-608): a, c, d hold cells 1..n on entry and the
// coefficients / right-hand side after `stages` stages on return.  See oracle/ref_pcr_harness.cpp.
void oracle_pcr_stages_(int* n_, int* stages, REAL* a, REAL* c, REAL* d) {
  const int n = *n_;
  int pn = -1, b = 1;
  for (int i = 1; i < 20; ++i) { b *= 2; if (n < b) { pn = i; break; } }  // cz.h:293-300
  LineWork w(1, n, pn);
  for (int k = 1; k <= n; ++k) { w.a[w.o(k)] = a[k - 1]; w.c[w.o(k)] = c[k - 1]; w.d[w.o(k)] = d[k - 1]; }
  line_pcr_stages(w, *stages);
  for (int k = 1; k <= n; ++k) { a[k - 1] = w.a[w.o(k)]; c[k - 1] = w.c[w.o(k)]; d[k - 1] = w.d[w.o(k)]; }
}

int oracle_real_bytes_() { return (int)sizeof(REAL); }

// ------------------------------------------------------------------ cz_maf.f90 / cz_blas.f90:738,845
// Stretched-grid ("MAF") twins.  Coordinates X(-1:sz(1)+2) etc.: Fortran index
// m -> C offset m+1.  The metric coefficients are recomputed per cell exactly as the reference
// does (they round differently per (i,j,k), so they cannot be tabulated per axis).
}  // extern "C"

namespace {

// cz_maf.f90:70-99 (psor_maf; identical block in jacobi_maf :205-234, psor2sma_core_maf
// :383-412, calc_rk_maf cz_blas.f90:785-812, calc_ax_maf :888-915, search_pivot :987-1014)
struct MafCell { REAL C1, C2, C3, C7, C8, C9; };
inline MafCell maf_cell(const REAL* X, const REAL* Y, const REAL* Z, int i, int j, int k) {
  const REAL XG = (REAL)0.5 * (X[i + 2] - X[i]);
  const REAL YE = (REAL)0.5 * (Y[j + 2] - Y[j]);
  const REAL ZT = (REAL)0.5 * (Z[k + 2] - Z[k]);
  const REAL XGG = X[i + 2] - (REAL)2.0 * X[i + 1] + X[i];
  const REAL YEE = Y[j + 2] - (REAL)2.0 * Y[j + 1] + Y[j];
  const REAL ZTT = Z[k + 2] - (REAL)2.0 * Z[k + 1] + Z[k];
  const REAL YJA = XG * YE * ZT;
  const REAL YJAI = (REAL)1.0 / YJA;
  const REAL GX = YE * ZT * YJAI;
  const REAL EY = XG * ZT * YJAI;
  const REAL TZ = XG * YE * YJAI;
  MafCell m;
  m.C1 = GX * GX; m.C2 = EY * EY; m.C3 = TZ * TZ;
  m.C7 = -XGG * m.C1 * GX; m.C8 = -YEE * m.C2 * EY; m.C9 = -ZTT * m.C3 * TZ;
  return m;
}

// cz_maf.f90:102-112: dd, rp (note "+ bb": the MAF point sweeps ADD the source term), dp
inline REAL maf_point_dp(const MafCell& m, const Grid& G, const REAL* p, long q, REAL bb, REAL pp,
                         REAL omg) {
  const REAL half = (REAL)0.5;
  const REAL dd = (REAL)2.0 * (m.C1 + m.C2 + m.C3);
  const REAL rp = (m.C1 + half * m.C7) * p[q + G.nk] + (m.C1 - half * m.C7) * p[q - G.nk] +
                  (m.C2 + half * m.C8) * p[q + G.nk * G.ni] + (m.C2 - half * m.C8) * p[q - G.nk * G.ni] +
                  (m.C3 + half * m.C9) * p[q + 1] + (m.C3 - half * m.C9) * p[q - 1] + bb;
  return (rp / dd - pp) * omg;
}

// One line of the MAF line-SOR routines: cz_maf.f90:522-647 (pcr_rb_maf); pcr_maf :744-870,
// pcr_eda_maf :981-1093, pcr_esa_maf :1194-1314, pcr_rb_esa_maf :1419-1539 run the same
// statements (pn-1 PCR stages + 2x2 last stage in all five; they differ in the line selection
// and in clamped vs zero-padded neighbour indices, which the padded LineWork covers: the
// clamped routines read a/c/d(kst-1), (ked+1), which they never initialise, only multiplied by
// coefficients that are exactly zero).
void solve_line_maf(LineWork& w, const Grid& G, REAL* x, const REAL* rhs, const REAL* msk,
                    const REAL* XX, const REAL* YY, const REAL* ZZ, int i, int j, int pn, REAL omg,
                    Acc& acc) {
  const REAL half = (REAL)0.5, two = (REAL)2.0;
  const int kst = w.kst, ked = w.ked;
  const REAL GX = two / (XX[i + 2] - XX[i]);
  const REAL EY = two / (YY[j + 2] - YY[j]);
  const REAL C1 = GX * GX;
  const REAL C2 = EY * EY;
  const REAL C7 = -(XX[i + 2] - two * XX[i + 1] + XX[i]) * C1 * GX;
  const REAL C8 = -(YY[j + 2] - two * YY[j + 1] + YY[j]) * C2 * EY;
  const REAL dd1 = C1 + half * C7, dd2 = C1 - half * C7;
  const REAL cc1 = C2 + half * C8, cc2 = C2 - half * C8;
  std::fill(w.a.begin(), w.a.end(), (REAL)0);
  std::fill(w.c.begin(), w.c.end(), (REAL)0);
  std::fill(w.d.begin(), w.d.end(), (REAL)0);
  for (int k = kst; k <= ked; ++k) {   // :534-543 -> a1 = C3, c1 = C9, d1 = 1/R7
    const REAL f1 = ZZ[k + 2], f2 = ZZ[k];
    const REAL TZ = two / (f1 - f2);
    const REAL ZTT = f1 - two * ZZ[k + 1] + f2;
    const REAL f3 = TZ * TZ;
    w.a1[w.o(k)] = f3;
    w.c1[w.o(k)] = -ZTT * f3 * TZ;
    w.d1[w.o(k)] = half / (C1 + C2 + f3);
  }
  const int ok = w.o(kst), oe = w.o(ked);
  w.a[ok] = (REAL)0;                                                   // :545
  w.c[ok] = -(w.a1[ok] + half * w.c1[ok]) * w.d1[ok];                  // :546
  for (int k = kst + 1; k <= ked - 1; ++k) {                           // :548-554
    const int o = w.o(k);
    const REAL f1 = w.a1[o], f2 = w.c1[o], aa3 = w.d1[o];
    w.a[o] = -(f1 - half * f2) * aa3;
    w.c[o] = -(f1 + half * f2) * aa3;
  }
  w.a[oe] = -(w.a1[oe] - half * w.c1[oe]) * w.d1[oe];                  // :556
  w.c[oe] = (REAL)0;                                                   // :557
  const long pl = G.nk * G.ni;
  for (int k = kst; k <= ked; ++k) {                                   // :560-568 (note "- rhs")
    const long q = G.at(k, i, j);
    w.d[w.o(k)] = (dd1 * x[q + G.nk] + dd2 * x[q - G.nk] + cc1 * x[q + pl] + cc2 * x[q - pl] - rhs[q]) *
                  w.d1[w.o(k)] * msk[q];
  }
  w.d[ok] = (w.d[ok] + (w.a1[ok] - half * w.c1[ok]) * w.d1[ok] * x[G.at(kst - 1, i, j)]) * msk[G.at(kst, i, j)];
  w.d[oe] = (w.d[oe] + (w.a1[oe] + half * w.c1[oe]) * w.d1[oe] * x[G.at(ked + 1, i, j)]) * msk[G.at(ked, i, j)];
  line_pcr_stages(w, pn - 1);                                          // :580-601
  line_last_2x2(w, pn);                                                // :605-620
  for (int k = kst; k <= ked; ++k) {                                   // :630-637
    const long q = G.at(k, i, j);
    const REAL pp = x[q];
    const REAL dp = (w.d1[w.o(k)] - pp) * omg * msk[q];
    x[q] = pp + dp;
    acc.add(dp * dp);                                                  // :645 (the #else branch: _SVR undefined)
  }
}

// flop formulas cz_maf.f90:493-505 (all exact small integers)
double pcr_maf_flop(const Range& R, int pn, double per2) {
  const double nk = R.ked - R.kst + 1;
  const double lines = (double)(R.jed - R.jst + 1) * (double)(R.ied - R.ist + 1);
  return lines * ((24.0 + 3.0 * 2.0 + 12.0) + nk * (11.0 + 10.0) + (nk - 2.0) * 6.0 + nk * (pn - 1) * 16.0 +
                  ipow2(pn - 1) * per2 + nk * 6.0);
}

// Residual of the five MAF line routines: `tmp(k) = tmp(k) + dp*dp` (+ `res1 = res1 + tmp(k)` after the loops) and
// `res1 = res1 + dp*dp` are the two branches of `#ifdef _SVR` (cz_maf.f90:642-646, 659-663; -D_SVR only with
// enable_VectorReduction=ON, CMakeLists.txt:227-230): every square is counted once, `tmp` is not touched in the
// default build.  (Until round 2 this port read both statements as unconditional and doubled the sum; executing the
// reference's own Fortran through oracle/f90_to_c.py showed the error.)
void pcr_maf_sweep(int* sz, int* idx, int* g, int pn, int color /* <0: all lines, (j,i) order */, REAL* x,
                   REAL* msk, REAL* rhs, REAL* XX, REAL* YY, REAL* ZZ, REAL omg, double* res, REAL* tmp) {
  (void)tmp;
  const Grid G(sz, *g);
  const Range R(idx);
  if (real_mode() || color < 0) {
    Acc acc;
    LineWork w(R.kst, R.ked, pn);
    for (int j = R.jst; j <= R.jed; ++j)
      for (int i = R.ist; i <= R.ied; ++i) {
        if (color >= 0 && ((i + j) % 2 + 2) % 2 != color) continue;
        solve_line_maf(w, G, x, rhs, msk, XX, YY, ZZ, i, j, pn, omg, acc);
      }
    *res += acc.value();
    return;
  }
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    LineWork w(R.kst, R.ked, pn);
    for (int i = R.ist; i <= R.ied; ++i) {
      if (((i + j) % 2 + 2) % 2 != color) continue;
      solve_line_maf(w, G, x, rhs, msk, XX, YY, ZZ, i, j, pn, omg, acc);
    }
  });
  *res += s;
}

}  // namespace

extern "C" {

// ------------------------------------------------------------------ cz_maf.f90:131-282
void jacobi_maf_(REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, REAL* omg, REAL* b,
                 double* res, REAL* wk2, REAL* tmp, double* flop) {
  (void)tmp;   // used only under _SVR
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL w = *omg;
  *flop += 66.0 * R.cells();
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) {
        const long q = G.at(k, i, j);
        const REAL bb = b[q], pp = p[q];
        const MafCell m = maf_cell(X, Y, Z, i, j, k);
        const REAL dp = maf_point_dp(m, G, p, q, bb, pp, w);
        wk2[q] = pp + dp;
        acc.add(dp * dp);
      }
  });
  for_planes(R.jst, R.jed, [&](int j) {   // :258-264 copy-back
    for (int i = R.ist; i <= R.ied; ++i) {
      const long q = G.at(0, i, j);
      for (int k = R.kst; k <= R.ked; ++k) p[q + k] = wk2[q + k];
    }
  });
  *res += s;
}

// ------------------------------------------------------------------ cz_maf.f90:23-114
// Lexicographic, racy under OpenMP in the reference -> 1-thread order.
void psor_maf_(REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, REAL* omg, REAL* b,
               double* res, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL w = *omg;
  *flop += 66.0 * R.cells();
  Acc acc;
  for (int j = R.jst; j <= R.jed; ++j)
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) {
        const long q = G.at(k, i, j);
        const REAL bb = b[q], pp = p[q];
        const MafCell m = maf_cell(X, Y, Z, i, j, k);
        const REAL dp = maf_point_dp(m, G, p, q, bb, pp, w);
        p[q] = pp + dp;
        acc.add(dp * dp);
      }
  *res += acc.value();
}

// ------------------------------------------------------------------ cz_maf.f90:301-438
void psor2sma_core_maf_(REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, int* ofst,
                        int* color, REAL* omg, REAL* b, double* res, REAL* tmp, double* flop) {
  (void)tmp;   // used only under _SVR
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL w = *omg;
  const int kp = *ofst + *color;
  *flop += 66.0 * 0.5 * R.cells();
  const double s = reduce_planes(R.jst, R.jed, true, [&](int j, Acc& acc) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst + ((i + j + kp) % 2 + 2) % 2; k <= R.ked; k += 2) {
        const long q = G.at(k, i, j);
        const REAL pp = p[q], bb = b[q];
        const MafCell m = maf_cell(X, Y, Z, i, j, k);
        const REAL dp = maf_point_dp(m, G, p, q, bb, pp, w);
        p[q] = pp + dp;
        acc.add(dp * dp);
      }
  });
  *res += s;
}

// ------------------------------------------------------------------ cz_maf.f90:442-668
void pcr_rb_maf_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, REAL* x, REAL* msk,
                 REAL* rhs, REAL* XX, REAL* YY, REAL* ZZ, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*,
                 REAL* omg, double* res, REAL* tmp, double* flop) {
  (void)ofst;   // ip = ofst + color is computed but not used (:507, 526)
  *flop += pcr_maf_flop(Range(idx), *pn, 11.0) * 0.5;
  pcr_maf_sweep(sz, idx, g, *pn, *color, x, msk, rhs, XX, YY, ZZ, *omg, res, tmp);
}

// ------------------------------------------------------------------ cz_maf.f90:1339-1560
void pcr_rb_esa_maf_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, int* s, REAL* x,
                     REAL* msk, REAL* rhs, REAL* XX, REAL* YY, REAL* ZZ, REAL*, REAL*, REAL*, REAL*,
                     REAL*, REAL*, REAL* omg, double* res, REAL* tmp, double* flop) {
  (void)ofst; (void)s;
  *flop += pcr_maf_flop(Range(idx), *pn, 11.0) * 0.5;
  pcr_maf_sweep(sz, idx, g, *pn, *color, x, msk, rhs, XX, YY, ZZ, *omg, res, tmp);
}

// ------------------------------------------------------------------ cz_maf.f90:672-892
// All lines in (j,i) order, in place: racy under OpenMP in the reference -> 1-thread order.
void pcr_maf_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL* XX, REAL* YY,
              REAL* ZZ, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, REAL* tmp,
              double* flop) {
  *flop += pcr_maf_flop(Range(idx), *pn, 11.0);
  pcr_maf_sweep(sz, idx, g, *pn, -1, x, msk, rhs, XX, YY, ZZ, *omg, res, tmp);
}

// ------------------------------------------------------------------ cz_maf.f90:896-1113
// Allocates a,c,d(kst-2^(pn-2) : ked+2^(pn-2)) itself; the 2x2 stage reads a(k+2^(pn-1)), which
// lies inside that allocation only when n >= 3*2^(pn-2) (true for the driver's 2^m-2 sizes).
void pcr_eda_maf_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL* XX,
                  REAL* YY, REAL* ZZ, REAL*, REAL*, REAL*, REAL* omg, double* res, REAL* tmp,
                  double* flop) {
  *flop += pcr_maf_flop(Range(idx), *pn, 9.0);
  pcr_maf_sweep(sz, idx, g, *pn, -1, x, msk, rhs, XX, YY, ZZ, *omg, res, tmp);
}

// ------------------------------------------------------------------ cz_maf.f90:1117-1335
void pcr_esa_maf_(int* sz, int* idx, int* g, int* pn, int* s, REAL* x, REAL* msk, REAL* rhs, REAL* XX,
                  REAL* YY, REAL* ZZ, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res,
                  REAL* tmp, double* flop) {
  (void)s;
  *flop += pcr_maf_flop(Range(idx), *pn, 9.0);
  pcr_maf_sweep(sz, idx, g, *pn, -1, x, msk, rhs, XX, YY, ZZ, *omg, res, tmp);
}

// ------------------------------------------------------------------ cz_blas.f90:738-829
void calc_rk_maf_(REAL* r, REAL* p, REAL* b, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z,
                  REAL* pvt, double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL half = (REAL)0.5;
  const long pl = G.nk * G.ni;
  *flop += 63.0 * R.cells();
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) {
        const long q = G.at(k, i, j);
        const MafCell m = maf_cell(X, Y, Z, i, j, k);
        r[q] = (b[q] + (REAL)2.0 * (m.C1 + m.C2 + m.C3) * p[q] - (m.C1 + half * m.C7) * p[q + G.nk] -
                (m.C1 - half * m.C7) * p[q - G.nk] - (m.C2 + half * m.C8) * p[q + pl] -
                (m.C2 - half * m.C8) * p[q - pl] - (m.C3 + half * m.C9) * p[q + 1] -
                (m.C3 - half * m.C9) * p[q - 1]) * pvt[q];
      }
  });
}

// ------------------------------------------------------------------ cz_blas.f90:845-935
void calc_ax_maf_(REAL* ap, REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, REAL* pvt,
                  double* flop) {
  const Grid G(sz, *g);
  const Range R(idx);
  const REAL half = (REAL)0.5;
  const long pl = G.nk * G.ni;
  *flop += 63.0 * R.cells();
  for_planes(R.jst, R.jed, [&](int j) {
    for (int i = R.ist; i <= R.ied; ++i)
      for (int k = R.kst; k <= R.ked; ++k) {
        const long q = G.at(k, i, j);
        const MafCell m = maf_cell(X, Y, Z, i, j, k);
        ap[q] = ((m.C1 + half * m.C7) * p[q + G.nk] + (m.C1 - half * m.C7) * p[q - G.nk] +
                 (m.C2 + half * m.C8) * p[q + pl] + (m.C2 - half * m.C8) * p[q - pl] +
                 (m.C3 + half * m.C9) * p[q + 1] + (m.C3 - half * m.C9) * p[q - 1] -
                 (REAL)2.0 * (m.C1 + m.C2 + m.C3) * p[q]) * pvt[q];
      }
  });
}

// ------------------------------------------------------------------ harness helpers (not reference symbols)
// bench.py's CPU arm: the OpenMP runtime reads OMP_NUM_THREADS once, when it is loaded; a launcher
// (torchrun) may have exported OMP_NUM_THREADS=1 before that.  These two calls set / report the team
// size the kernels above actually run with (cz_Evaluate.cpp:34-39 reads the same value for its banner).
// ORACLE_SUM from the harness: 1 = "real" (the reference's 1-thread order, one REAL accumulator), 0 = "double"
void oracle_set_sum_mode(int real_order) { g_sum_mode = real_order ? 1 : 0; }
void oracle_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
int oracle_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"

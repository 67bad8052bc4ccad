// cubez_b200/csrc/abi.cu -- PART 1 of include/czb200.h: the reference's Fortran kernel symbols
// (src/cz_cpp/cz_Ffunc.h:16-578) implemented over the CUDA kernels.  Per-call semantics are the
// reference's: res/flop accumulated into, dot results overwritten, the call returns when the
// result is visible to the host.  Array arguments are device pointers (zero copy) or host
// pointers (staged, runtime.h).  The flop formulas are the reference's analytic counts.
#include "runtime.h"
#include <vector>
#include "../../include/czb200.h"

using namespace czb;

namespace {

inline double ipow2(int e) { return (double)(1LL << e); }

// run one PCR-type sweep for the per-call ABI
void pcr_call(int* sz, int* idx, int* g, int* pn, int mode_kind /*0 rb,1 lex,2 jacobi*/, int color, bool last4,
              REAL* x, REAL* msk, REAL* rhs, REAL* wrk, REAL* omg, double* res) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  PcrTables& T = last4 ? R.pcr4 : R.pcr2;
  pcr_build_tables(R.stream, T, G.ked - G.kst + 1, *pn, last4);
  Arg<REAL> ax(x, n, true, true), am(msk, n, true, false), ar(rhs, n, true, false);
  Launch L = R.launch();
  if (mode_kind == 0) {
    launch_pcr_lines(L, G, T, color, 0, 0, ax.dev(), ax.dev(), am.dev(), ar.dev(), *omg, 0, false);
  } else if (mode_kind == 1) {
    bool first = true;
    for (int d = G.ist + G.jst; d <= G.ied + G.jed; ++d) {
      launch_pcr_lines(L, G, T, 3, 0, d, ax.dev(), ax.dev(), am.dev(), ar.dev(), *omg, 0, !first);
      first = false;
    }
  } else {
    Arg<REAL> aw(wrk, n, true, true);
    launch_pcr_lines(L, G, T, 2, 0, 0, ax.dev(), aw.dev(), am.dev(), ar.dev(), *omg, 0, false);
    launch_copy_inner(L, G, ax.dev(), aw.dev());
  }
  *res += R.read_acc(0);
}

inline double pcr_flop(const int* idx, int pn, bool last4, double per4, bool half) {
  const double nk = idx[5] - idx[4] + 1;
  const double lines = (double)(idx[3] - idx[2] + 1) * (idx[1] - idx[0] + 1);
  const double f = last4 ? (nk * 6.0 + nk * (pn - 2) * 14.0 + ipow2(pn - 2) * per4 + nk * 6.0 + 6.0)
                         : (nk * 6.0 + nk * (pn - 1) * 14.0 + ipow2(pn - 1) * 9.0 + nk * 6.0 + 6.0);
  return lines * f * (half ? 0.5 : 1.0);
}


// flop formulas of the MAF line routines (cz_maf.f90:493-505, 713-725, 924-936, 1150-1162, 1373-1385)
inline double pcr_maf_flop(const int* idx, int pn, double per2, bool half) {
  const double nk = idx[5] - idx[4] + 1;
  const double lines = (double)(idx[3] - idx[2] + 1) * (idx[1] - idx[0] + 1);
  const double f = (24.0 + 3.0 * 2.0 + 12.0) + nk * (11.0 + 10.0) + (nk - 2.0) * 6.0 + nk * (pn - 1) * 16.0 +
                   ipow2(pn - 1) * per2 + nk * 6.0;
  return lines * f * (half ? 0.5 : 1.0);
}

// coordinate vectors of the MAF ABI: X(-1:sz(1)+2), Y(-1:sz(2)+2), Z(-1:sz(3)+2), host or device
struct MafArgs {
  Arg<REAL> x, y, z;
  MafArgs(REAL* X, REAL* Y, REAL* Z, const int* sz, int g)
      : x(X, (size_t)sz[0] + 2 * g, true, false), y(Y, (size_t)sz[1] + 2 * g, true, false),
        z(Z, (size_t)sz[2] + 2 * g, true, false) {}
  MafAxes axes() const { return MafAxes{x.dev(), y.dev(), z.dev()}; }
};

// one MAF line-SOR sweep for the per-call ABI: color >= 0 red-black, < 0 lexicographic (wavefronts i+j = const)
void pcr_maf_call(int* sz, int* idx, int* g, int* pn, int color, REAL* x, REAL* msk, REAL* rhs, REAL* XX, REAL* YY,
                  REAL* ZZ, REAL* omg, double* res) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  {
    Arg<REAL> ax(x, n, true, true), am(msk, n, true, false), ar(rhs, n, true, false);
    MafArgs M(XX, YY, ZZ, sz, *g);
    Launch L = R.launch();
    if (color >= 0) {
      launch_pcr_lines_maf(L, G, *pn, color, 0, M.axes(), ax.dev(), am.dev(), ar.dev(), *omg, 0, false);
    } else {
      bool first = true;
      for (int d = G.ist + G.jst; d <= G.ied + G.jed; ++d) {
        launch_pcr_lines_maf(L, G, *pn, 3, d, M.axes(), ax.dev(), am.dev(), ar.dev(), *omg, 0, !first);
        first = false;
      }
    }
  }
  *res += R.read_acc(0);
}

}  // namespace

extern "C" {

// ------------------------------------------------------------------ cz_solver.f90:22-191
void bc_k_(int* sz, int* g, REAL* p, REAL* dh, REAL* org, int* nID) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, nullptr, *g);
  Arg<REAL> ap(p, (size_t)G.total, true, true);
  launch_bc_k(R.launch(), G, ap.dev(), *dh, org, nID);
}

// ------------------------------------------------------------------ cz_solver.f90:284-387
void jacobi_(REAL* p, int* sz, int* idx, int* g, REAL* cf, REAL* omg, REAL* b, double* res, REAL* wk2,
             double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 18.0 * G.cells();
  {
    Arg<REAL> ap(p, n, true, true), ab(b, n, true, false), aw(wk2, n, true, true);
    Launch L = R.launch();
    launch_jacobi(L, G, ap.dev(), aw.dev(), ab.dev(), cf, *omg, 0, false);
    launch_copy_inner(L, G, ap.dev(), aw.dev());   // cz_solver.f90:369-375
  }
  *res += R.read_acc(0);
}

// ------------------------------------------------------------------ cz_solver.f90:207-269
void psor_(REAL* p, int* sz, int* idx, int* g, REAL* cf, REAL* omg, REAL* b, double* res, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 18.0 * G.cells();
  {
    Arg<REAL> ap(p, n, true, true), ab(b, n, true, false);
    launch_psor_sweep(R.launch(), G, ap.dev(), ab.dev(), cf, *omg, 0, false);
  }
  *res += R.read_acc(0);
}

// ------------------------------------------------------------------ cz_solver.f90:404-493
void psor2sma_core_(REAL* p, int* sz, int* idx, int* g, REAL* cf, int* ip, int* color, REAL* omg, REAL* b,
                    double* res, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 18.0 * 0.5 * G.cells();
  {
    Arg<REAL> ap(p, n, true, true), ab(b, n, true, false);
    launch_rb_color(R.launch(), G, ap.dev(), ab.dev(), cf, *omg, *ip + *color, 0, false);
  }
  *res += R.read_acc(0);
}

// ------------------------------------------------------------------ cz_solver.f90:497-662
void pcr_rb_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, REAL* x, REAL* msk, REAL* rhs, REAL*,
             REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, double* flop) {
  (void)ofst;   // ignored by the reference too (cz_solver.f90:533,549)
  *flop += pcr_flop(idx, *pn, false, 0.0, true);
  pcr_call(sz, idx, g, pn, 0, *color, false, x, msk, rhs, nullptr, omg, res);
}

// ------------------------------------------------------------------ cz_solver.f90:1261-1469
void pcr_rb_esa_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, int* s, REAL* x, REAL* msk,
                 REAL* rhs, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, double* flop) {
  (void)ofst; (void)s;
  *flop += pcr_flop(idx, *pn, true, 78.0, true);
  pcr_call(sz, idx, g, pn, 0, *color, true, x, msk, rhs, nullptr, omg, res);
}

// ------------------------------------------------------------------ cz_solver.f90:666-878
void pcr_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL*, REAL*, REAL*, REAL*, REAL*,
          REAL*, REAL* omg, double* res, double* flop) {
  *flop += pcr_flop(idx, *pn, true, 74.0, false);
  pcr_call(sz, idx, g, pn, 1, 0, true, x, msk, rhs, nullptr, omg, res);
}

// ------------------------------------------------------------------ cz_solver.f90:883-1045
void pcr_eda_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL*, REAL*, REAL*, REAL* omg,
              double* res, double* flop) {
  *flop += pcr_flop(idx, *pn, false, 0.0, false);
  pcr_call(sz, idx, g, pn, 1, 0, false, x, msk, rhs, nullptr, omg, res);
}

// ------------------------------------------------------------------ cz_solver.f90:1050-1257
void pcr_esa_(int* sz, int* idx, int* g, int* pn, int* s, REAL* x, REAL* msk, REAL* rhs, REAL*, REAL*, REAL*,
              REAL*, REAL*, REAL*, REAL* omg, double* res, double* flop) {
  (void)s;
  *flop += pcr_flop(idx, *pn, true, 78.0, false);
  pcr_call(sz, idx, g, pn, 1, 0, true, x, msk, rhs, nullptr, omg, res);
}

// ------------------------------------------------------------------ cz_solver.f90:1473-1676
void pcr_j_esa_(int* sz, int* idx, int* g, int* pn, int* s, REAL* x, REAL* msk, REAL* rhs, REAL*, REAL*, REAL*,
                REAL*, REAL*, REAL*, REAL* src, REAL* wrk, REAL* omg, double* res, double* flop) {
  (void)s; (void)src;   // src is scratch in the reference; the source term is fused here
  *flop += pcr_flop(idx, *pn, false, 0.0, false);
  pcr_call(sz, idx, g, pn, 2, 0, false, x, msk, rhs, wrk, omg, res);
}

// ------------------------------------------------------------------ cz_blas.f90
void imask_k_(REAL* x, int* sz, int* idx, int* g) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  Arg<REAL> ax(x, (size_t)G.total, false, true);
  launch_imask(R.launch(), G, ax.dev());
}

void blas_clear_(REAL* x, int* sz, int* g) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, nullptr, *g);
  Arg<REAL> ax(x, (size_t)G.total, false, true);
  CZB_CHECK(cudaMemsetAsync(ax.dev(), 0, sizeof(REAL) * (size_t)G.total, R.stream));
}

void blas_copy_(REAL* dst, REAL* src, int* sz, int* g) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, nullptr, *g);
  const size_t n = (size_t)G.total;
  Arg<REAL> ad(dst, n, false, true), as(src, n, true, false);
  CZB_CHECK(cudaMemcpyAsync(ad.dev(), as.dev(), sizeof(REAL) * n, cudaMemcpyDeviceToDevice, R.stream));
}

void blas_copy_in_(REAL* dst, REAL* src, int* sz, int* g) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, nullptr, *g);
  const size_t n = (size_t)G.total;
  Arg<REAL> ad(dst, n, true, true), as(src, n, true, false);
  launch_copy_in(R.launch(), G, ad.dev(), as.dev());
}

void blas_triad_(REAL* z, REAL* x, REAL* y, REAL* a, int* sz, int* idx, int* g, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 2.0 * G.cells();
  Arg<REAL> az(z, n, true, true), ax(x, n, true, false), ay(y, n, true, false);
  launch_triad(R.launch(), G, az.dev(), ax.dev(), ay.dev(), *a, nullptr);
}

void blas_dot1_(REAL* r, REAL* p, int* sz, int* idx, int* g, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  *flop += 2.0 * G.cells();
  {
    Arg<REAL> ap(p, (size_t)G.total, true, false);
    launch_dot1(R.launch(), G, ap.dev(), 0);
  }
  *r = (REAL)R.read_acc(0);
}

void blas_dot2_(REAL* r, REAL* p, REAL* q, int* sz, int* idx, int* g, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 2.0 * G.cells();
  {
    Arg<REAL> ap(p, n, true, false), aq(q, n, true, false);
    launch_dot2(R.launch(), G, ap.dev(), aq.dev(), 0);
  }
  *r = (REAL)R.read_acc(0);
}

void blas_bicg_1_(REAL* p, REAL* r, REAL* q, REAL* beta, REAL* omg, int* sz, int* idx, int* g, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 4.0 * G.cells();
  Arg<REAL> ap(p, n, true, true), ar(r, n, true, false), aq(q, n, true, false);
  launch_bicg1(R.launch(), G, ap.dev(), ar.dev(), aq.dev(), *beta, *omg, nullptr, nullptr);
}

void blas_bicg_2_(REAL* z, REAL* x, REAL* y, REAL* a, REAL* b, int* sz, int* idx, int* g, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 4.0 * G.cells();
  Arg<REAL> az(z, n, true, true), ax(x, n, true, false), ay(y, n, true, false);
  launch_bicg2(R.launch(), G, az.dev(), ax.dev(), ay.dev(), *a, *b, nullptr, nullptr);
}

void blas_calc_ax_(REAL* ap, REAL* p, int* sz, int* idx, int* g, REAL* cf, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 13.0 * G.cells();
  Arg<REAL> aa(ap, n, true, true), a_p(p, n, true, false);
  launch_calc_ax(R.launch(), G, aa.dev(), a_p.dev(), cf, nullptr, 0);
}

void blas_calc_rk_(REAL* r, REAL* p, REAL* b, int* sz, int* idx, int* g, REAL* cf, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 14.0 * G.cells();
  Arg<REAL> ar(r, n, true, true), ap(p, n, true, false), ab(b, n, true, false);
  launch_calc_rk(R.launch(), G, ar.dev(), ap.dev(), ab.dev(), cf);
}

// cz_blas.f90:947-1039: pvt = 1 / max |row coefficient| on the inner cells (called unconditionally by the
// reference driver, cz_Evaluate.cpp:369; used by pbicgstab_maf only)
void search_pivot_(REAL* pvt, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  Arg<REAL> ap(pvt, (size_t)G.total, true, true);
  MafArgs M(X, Y, Z, sz, *g);
  launch_search_pivot(R.launch(), G, ap.dev(), M.axes());
}

// ------------------------------------------------------------------ cz_utility.f90
// cz_utility.f90:17-44.  The reference compiles the body out unless _aurora_==1, so by default this is a no-op as
// there; with CZB_FILEOUT=1 it writes the same .sph record sequence (Fortran unformatted sequential, 4-byte record
// markers as gfortran writes them): (1,1) / (ix,jx,kx) / org(1:3) / (dh,dh,dh) / (nn=0, rtime=0) /
// (((s(k,i,j), i=1,ix), j=1,jx), k=1,kx).  s may be a host or a device pointer.
void fileout_t_(int* sz, int* g, REAL* s, REAL* dh, REAL* org, char* fname) {
  static const bool on = getenv("CZB_FILEOUT") != nullptr;
  if (!on) return;
  const Geo G = make_geo(sz, nullptr, *g);
  const size_t n = (size_t)G.total;
  std::vector<REAL> host;
  const REAL* h = s;
  if (is_device_ptr(s)) {
    Runtime& R = rt();
    host.resize(n);
    CZB_CHECK(cudaMemcpyAsync(host.data(), s, sizeof(REAL) * n, cudaMemcpyDeviceToHost, R.stream));
    R.sync();
    h = host.data();
  }
  char name[21];                                   // character*20, blank padded on the Fortran side
  size_t len = 0;
  while (len < 20 && fname[len] != '\0' && fname[len] != ' ') { name[len] = fname[len]; ++len; }
  name[len] = '\0';
  FILE* f = fopen(name, "wb");
  if (!f) { fprintf(stderr, "czb200: fileout_t_: cannot open '%s'\n", name); return; }
  auto record = [&](const void* data, size_t bytes) {
    const int32_t m = (int32_t)bytes;
    fwrite(&m, 4, 1, f); fwrite(data, 1, bytes, f); fwrite(&m, 4, 1, f);
  };
  const int one[2] = {1, 1};
  record(one, sizeof(one));
  const int dims[3] = {G.ix, G.jx, G.kx};
  record(dims, sizeof(dims));
  const REAL o3[3] = {org[0], org[1], org[2]};
  record(o3, sizeof(o3));
  const REAL d3[3] = {*dh, *dh, *dh};
  record(d3, sizeof(d3));
  unsigned char step[sizeof(int) + sizeof(REAL)] = {0};   // nn = 0 (integer), rtime = 0.0 (REAL), packed
  record(step, sizeof(step));
  std::vector<REAL> out((size_t)G.ix * G.jx * G.kx);
  size_t q = 0;
  for (int k = 1; k <= G.kx; ++k)
    for (int j = 1; j <= G.jx; ++j)
      for (int i = 1; i <= G.ix; ++i) out[q++] = h[G.at(k, i, j)];
  if (out.size() * sizeof(REAL) > 0x7fffffffu) fprintf(stderr, "czb200: fileout_t_: record longer than 2 GiB, markers wrap\n");
  record(out.data(), out.size() * sizeof(REAL));
  fclose(f);
}

void exact_t_(int* sz, int* g, REAL* e, REAL* dh, REAL* org) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, nullptr, *g);
  Arg<REAL> ae(e, (size_t)G.total, true, true);
  launch_exact(R.launch(), G, ae.dev(), *dh, org);
}

void err_t_(int* sz, int* idx, int* g, double* d, REAL* p, REAL* e, int* loc) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  unsigned long long out[3];
  {
    Arg<REAL> ap(p, n, true, false), ae(e, n, true, true);
    launch_err(R.launch(), G, ap.dev(), ae.dev(), R.d_err);
    CZB_CHECK(cudaMemcpyAsync(out, R.d_err, sizeof(out), cudaMemcpyDeviceToHost, R.stream));
    R.sync();
  }
  double v; memcpy(&v, &out[0], sizeof(double));
  *d = v;
  if (out[1] == ~0ull) { loc[0] = loc[1] = loc[2] = -1; }
  else {
    const unsigned long long nkk = G.ked - G.kst + 1, nii = G.ied - G.ist + 1;
    const unsigned long long k = out[1] % nkk, i = (out[1] / nkk) % nii, j = out[1] / (nkk * nii);
    loc[0] = G.ist + (int)i; loc[1] = G.jst + (int)j; loc[2] = G.kst + (int)k;
  }
}

// ------------------------------------------------------------------ cz_maf.f90:131-282
void jacobi_maf_(REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, REAL* omg, REAL* b, double* res,
                 REAL* wk2, REAL* tmp, double* flop) {
  (void)tmp;   // used by the reference only under _SVR
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 66.0 * G.cells();
  {
    Arg<REAL> ap(p, n, true, true), ab(b, n, true, false), aw(wk2, n, true, true);
    MafArgs M(X, Y, Z, sz, *g);
    Launch L = R.launch();
    launch_jacobi_maf(L, G, ap.dev(), aw.dev(), ab.dev(), M.axes(), *omg, 0, false);
    launch_copy_inner(L, G, ap.dev(), aw.dev());   // cz_maf.f90:258-264
  }
  *res += R.read_acc(0);
}

// ------------------------------------------------------------------ cz_maf.f90:23-114
void psor_maf_(REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, REAL* omg, REAL* b, double* res,
               double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 66.0 * G.cells();
  {
    Arg<REAL> ap(p, n, true, true), ab(b, n, true, false);
    MafArgs M(X, Y, Z, sz, *g);
    launch_psor_sweep_maf(R.launch(), G, ap.dev(), ab.dev(), M.axes(), *omg, 0, false);
  }
  *res += R.read_acc(0);
}

// ------------------------------------------------------------------ cz_maf.f90:301-438
void psor2sma_core_maf_(REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, int* ip, int* color, REAL* omg,
                        REAL* b, double* res, REAL* tmp, double* flop) {
  (void)tmp;   // used by the reference only under _SVR
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 66.0 * 0.5 * G.cells();
  {
    Arg<REAL> ap(p, n, true, true), ab(b, n, true, false);
    MafArgs M(X, Y, Z, sz, *g);
    launch_rb_color_maf(R.launch(), G, ap.dev(), ab.dev(), M.axes(), *omg, *ip + *color, 0, false);
  }
  *res += R.read_acc(0);
}

// The five MAF line routines: the work vectors a, c, d, aw, cw, dw and tmp of the reference are scratch (tmp is an
// output nobody reads, cz_Poisson.cpp:565-567); the line solve lives in shared memory here.
// ------------------------------------------------------------------ cz_maf.f90:442-668
void pcr_rb_maf_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, REAL* x, REAL* msk, REAL* rhs, REAL* XX,
                 REAL* YY, REAL* ZZ, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, REAL*,
                 double* flop) {
  (void)ofst;   // ip = ofst + color is computed but never used (cz_maf.f90:507, 526)
  *flop += pcr_maf_flop(idx, *pn, 11.0, true);
  pcr_maf_call(sz, idx, g, pn, *color, x, msk, rhs, XX, YY, ZZ, omg, res);
}

// ------------------------------------------------------------------ cz_maf.f90:1339-1560
void pcr_rb_esa_maf_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, int* s, REAL* x, REAL* msk, REAL* rhs,
                     REAL* XX, REAL* YY, REAL* ZZ, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res,
                     REAL*, double* flop) {
  (void)ofst; (void)s;
  *flop += pcr_maf_flop(idx, *pn, 11.0, true);
  pcr_maf_call(sz, idx, g, pn, *color, x, msk, rhs, XX, YY, ZZ, omg, res);
}

// ------------------------------------------------------------------ cz_maf.f90:672-892
void pcr_maf_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL* XX, REAL* YY, REAL* ZZ, REAL*,
              REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, REAL*, double* flop) {
  *flop += pcr_maf_flop(idx, *pn, 11.0, false);
  pcr_maf_call(sz, idx, g, pn, -1, x, msk, rhs, XX, YY, ZZ, omg, res);
}

// ------------------------------------------------------------------ cz_maf.f90:896-1113
void pcr_eda_maf_(int* sz, int* idx, int* g, int* pn, REAL* x, REAL* msk, REAL* rhs, REAL* XX, REAL* YY, REAL* ZZ,
                  REAL*, REAL*, REAL*, REAL* omg, double* res, REAL*, double* flop) {
  *flop += pcr_maf_flop(idx, *pn, 9.0, false);
  pcr_maf_call(sz, idx, g, pn, -1, x, msk, rhs, XX, YY, ZZ, omg, res);
}

// ------------------------------------------------------------------ cz_maf.f90:1117-1335
void pcr_esa_maf_(int* sz, int* idx, int* g, int* pn, int* s, REAL* x, REAL* msk, REAL* rhs, REAL* XX, REAL* YY,
                  REAL* ZZ, REAL*, REAL*, REAL*, REAL*, REAL*, REAL*, REAL* omg, double* res, REAL*, double* flop) {
  (void)s;
  *flop += pcr_maf_flop(idx, *pn, 9.0, false);
  pcr_maf_call(sz, idx, g, pn, -1, x, msk, rhs, XX, YY, ZZ, omg, res);
}

// ------------------------------------------------------------------ cz_blas.f90:738-829
void calc_rk_maf_(REAL* r, REAL* p, REAL* b, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, REAL* pvt,
                  double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 63.0 * G.cells();
  Arg<REAL> ar(r, n, true, true), ap(p, n, true, false), ab(b, n, true, false), av(pvt, n, true, false);
  MafArgs M(X, Y, Z, sz, *g);
  launch_calc_rk_maf(R.launch(), G, ar.dev(), ap.dev(), ab.dev(), M.axes(), av.dev());
}

// ------------------------------------------------------------------ cz_blas.f90:845-935
void calc_ax_maf_(REAL* ap, REAL* p, int* sz, int* idx, int* g, REAL* X, REAL* Y, REAL* Z, REAL* pvt, double* flop) {
  Runtime& R = rt();
  const Geo G = make_geo(sz, idx, *g);
  const size_t n = (size_t)G.total;
  *flop += 63.0 * G.cells();
  Arg<REAL> aa(ap, n, true, true), a_p(p, n, true, false), av(pvt, n, true, false);
  MafArgs M(X, Y, Z, sz, *g);
  launch_calc_ax_maf(R.launch(), G, aa.dev(), a_p.dev(), M.axes(), av.dev(), nullptr, 0);
}

}  // extern "C"

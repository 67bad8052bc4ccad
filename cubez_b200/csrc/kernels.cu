// cubez_b200/csrc/kernels.cu -- baseline ("v1") sm_100a kernels for the CubeZ hot path:
// point sweeps, BLAS-1 + 7-point operator, boundary/utility and loop-control kernels.
// One thread per (k,i) position marching a chunk of j-planes; k (the contiguous axis) maps to
// threadIdx.x so every warp load/store is a contiguous 128 B run.  Arithmetic mirrors the
// reference statement by statement (same operand order, IEEE division; the library is built
// with -fmad=false so no FMA contraction changes a rounding) => per-cell results are bit
// identical to the reference's un-contracted CPU arithmetic.  Residual / dot sums are REAL
// products accumulated in double through the deterministic grid_reduce of common.cuh.
//
// The bandwidth-optimised fused kernels (TMA-staged, both colours per pass) live in fused.cu.
#include "kernels.cuh"
#include "cells.cuh"
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

namespace czb {

static long long g_launches = 0;
void count_launch(int n) { g_launches += n; }
long long launch_count() { return g_launches; }

namespace {

// ------------------------------------------------------------------ stencil helpers
struct Coef { REAL c1, c2, c3, c4, c5, c6, dd; };
inline Coef make_coef(const REAL* cf) { return Coef{cf[0], cf[1], cf[2], cf[3], cf[4], cf[5], cf[6]}; }

// ss exactly as cz_solver.f90:339-344 / cz_blas.f90:626-631 (left-to-right)
__device__ __forceinline__ REAL stencil_ss(const Geo& G, const Coef& c, const REAL* __restrict__ p, long long off) {
  return c.c1 * p[off + G.nk] + c.c2 * p[off - G.nk] + c.c3 * p[off + G.plane] + c.c4 * p[off - G.plane] +
         c.c5 * p[off + 1] + c.c6 * p[off - 1];
}

struct OpJacobi {
  const REAL* p; REAL* pn; const REAL* b; Coef c; REAL omg;
  __device__ void operator()(const Geo& G, int, int, int, long long off, double* acc) const {
    const REAL pp = p[off];
    const REAL bb = b[off];
    const REAL ss = stencil_ss(G, c, p, off);
    const REAL dp = ((ss - bb) / c.dd - pp) * omg;
    pn[off] = pp + dp;
    acc[0] += (double)(dp * dp);
  }
};

struct OpCopy {
  REAL* dst; const REAL* src;
  __device__ void operator()(const Geo&, int, int, int, long long off, double*) const { dst[off] = src[off]; }
};

struct OpAx {
  REAL* ap; const REAL* p; Coef c; const REAL* w;
  __device__ void operator()(const Geo& G, int, int, int, long long off, double* acc) const {
    const REAL ss = stencil_ss(G, c, p, off);
    const REAL v = ss - c.dd * p[off];
    ap[off] = v;
    if (w) acc[0] += (double)(v * w[off]);
  }
};

struct OpRk {
  REAL* r; const REAL* p; const REAL* b; Coef c;
  __device__ void operator()(const Geo& G, int, int, int, long long off, double*) const {
    const REAL ss = stencil_ss(G, c, p, off);
    r[off] = b[off] - (ss - c.dd * p[off]);
  }
};

struct OpDot1 {
  const REAL* p;
  __device__ void operator()(const Geo&, int, int, int, long long off, double* acc) const {
    const REAL q = p[off]; acc[0] += (double)(q * q);
  }
};
struct OpDot2 {
  const REAL* p; const REAL* q;
  __device__ void operator()(const Geo&, int, int, int, long long off, double* acc) const {
    acc[0] += (double)(p[off] * q[off]);
  }
};
struct OpTriad {   // z = a*x + y   (cz_blas.f90:296)
  REAL* z; const REAL* x; const REAL* y; REAL a; const double* a_dev;
  __device__ void operator()(const Geo&, int, int, int, long long off, double*) const {
    const REAL aa = a_dev ? (REAL)(*a_dev) : a;
    z[off] = aa * x[off] + y[off];
  }
};
struct OpBicg1 {   // p = r + beta*(p - omg*q)   (cz_blas.f90:489)
  REAL* p; const REAL* r; const REAL* q; REAL be, om; const double* be_dev; const double* om_dev;
  __device__ void operator()(const Geo&, int, int, int, long long off, double*) const {
    const REAL b = be_dev ? (REAL)(*be_dev) : be;
    const REAL o = om_dev ? (REAL)(*om_dev) : om;
    p[off] = r[off] + b * (p[off] - o * q[off]);
  }
};
struct OpBicg2 {   // z = a*x + b*y + z   (cz_blas.f90:553)
  REAL* z; const REAL* x; const REAL* y; REAL a, b; const double* a_dev; const double* b_dev;
  __device__ void operator()(const Geo&, int, int, int, long long off, double*) const {
    const REAL aa = a_dev ? (REAL)(*a_dev) : a;
    const REAL bb = b_dev ? (REAL)(*b_dev) : b;
    z[off] = aa * x[off] + bb * y[off] + z[off];
  }
};
// PBiCGstab tail in one pass (cz_Poisson.cpp:470-488): x += alpha*p_ + omega*s_ (blas_bicg_2, cz_blas.f90:553),
// r = s - omega*t (blas_triad :296), acc[0] = (r, r) (blas_dot1) and acc[1] = (r, r0) (the blas_dot2 that opens
// the next iteration, :376) -- 6 reads + 2 writes per cell instead of 10 + 3 in four kernels.
struct OpBicgTail {
  REAL* x; const REAL* p_; const REAL* s_; REAL* r; const REAL* t; const REAL* s; const REAL* r0;
  const double* al_dev; const double* om_dev; const double* nom_dev;
  __device__ void operator()(const Geo&, int, int, int, long long off, double* acc) const {
    const REAL al = (REAL)(*al_dev), om = (REAL)(*om_dev), nom = (REAL)(*nom_dev);
    x[off] = al * p_[off] + om * s_[off] + x[off];
    const REAL rr = nom * t[off] + s[off];
    r[off] = rr;
    acc[0] += (double)(rr * rr);
    acc[1] += (double)(rr * r0[off]);
  }
};
struct FinStore2 {   // acc[slot0] = total0 ; acc[slot1] = total1
  LoopState* st; int slot0, slot1;
  __device__ void operator()(double (&tot)[2]) const {
    if (st) { st->acc[slot0] = tot[0]; st->acc[slot1] = tot[1]; }
  }
};
struct OpOne {
  REAL* x;
  __device__ void operator()(const Geo&, int, int, int, long long off, double*) const { x[off] = (REAL)1.0; }
};

// One colour of red-black SOR, in place.  Thread t of a line handles the cell
// k = kst + mod(i+j+kp,2) + 2t  (cz_solver.f90:466), so all lanes are active.
template <class Fin>
__global__ void __launch_bounds__(256) k_rb_color(Geo G, int jlen, const LoopState* st, REAL* __restrict__ p,
                                                  const REAL* __restrict__ b, Coef c, REAL omg, int kp,
                                                  RedBuf rb, Fin fin) {
  if (st && st->done) return;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = G.ist + blockIdx.y * blockDim.y + threadIdx.y;
  const int j0 = G.jst + blockIdx.z * jlen;
  const int j1 = min(j0 + jlen - 1, G.jed);
  double acc[1] = {0.0};
  if (i <= G.ied) {
    for (int j = j0; j <= j1; ++j) {
      const int k = G.kst + ((i + j + kp) & 1) + 2 * t;
      if (k <= G.ked) {
        const long long off = G.at(k, i, j);
        const REAL pp = p[off];
        const REAL bb = b[off];
        const REAL ss = stencil_ss(G, c, p, off);
        const REAL dp = ((ss - bb) / c.dd - pp) * omg;
        p[off] = pp + dp;
        acc[0] += (double)(dp * dp);
      }
    }
  }
  grid_reduce<1>(acc, rb, fin);
}

// Lexicographic point SOR (psor, cz_solver.f90:207-269) reproduced exactly by hyperplane
// wavefronts: cells with i+j+k == c depend only on plane c-1 (already updated) and c+1 (old),
// so one launch per hyperplane gives the reference's 1-thread result.  (i,j) -> thread, k = c-i-j.
template <class Fin>
__global__ void __launch_bounds__(256) k_psor_plane(Geo G, const LoopState* st, REAL* __restrict__ p,
                                                    const REAL* __restrict__ b, Coef c, REAL omg, int hp,
                                                    RedBuf rb, Fin fin) {
  if (st && st->done) return;
  const int i = G.ist + blockIdx.x * blockDim.x + threadIdx.x;
  const int j = G.jst + blockIdx.y * blockDim.y + threadIdx.y;
  double acc[1] = {0.0};
  const int k = hp - i - j;
  if (i <= G.ied && j <= G.jed && k >= G.kst && k <= G.ked) {
    const long long off = G.at(k, i, j);
    const REAL pp = p[off];
    const REAL bb = b[off];
    const REAL ss = stencil_ss(G, c, p, off);
    const REAL dp = ((ss - bb) / c.dd - pp) * omg;
    p[off] = pp + dp;
    acc[0] += (double)(dp * dp);
  }
  grid_reduce<1>(acc, rb, fin);
}

// ------------------------------------------------------------------ boundary / utility
// bc_k (cz_solver.f90:22-191).  sin(pi*x)*sin(pi*y) in REAL; the float build evaluates sin
// in double and rounds (correctly rounded float sine, which is what glibc's sinf returns in
// all but ~1e-7 of cases) so that boundary values agree with the CPU reference to the bit
// w.h.p.; tests allow 1 ulp.
__device__ __forceinline__ REAL real_sin(REAL a) {
#ifdef _REAL_IS_DOUBLE_
  return sin(a);
#else
  return (float)sin((double)a);
#endif
}
__device__ __forceinline__ REAL real_sinh(REAL a) {
#ifdef _REAL_IS_DOUBLE_
  return sinh(a);
#else
  return (float)sinh((double)a);
#endif
}
__device__ __forceinline__ REAL real_pi() {
#ifdef _REAL_IS_DOUBLE_
  return 2.0 * asin(1.0);
#else
  return 2.0f * (float)asin(1.0);   // == 2.0f*asinf(1.0f) == 0x40490FDB
#endif
}

struct BcArgs { REAL dh, ox, oy; int face[6]; const REAL* tbl; int joff, ioff; };

// bc_k (cz_solver.f90:22-191).  K faces: sin(pi x) sin(pi y); I/J faces: 0.  `tbl` (ix*jx values,
// see bc_table below) holds the K-face Dirichlet data; tbl == null evaluates it on the device.
__global__ void k_bc(Geo G, BcArgs a, const LoopState* st, REAL* __restrict__ p) {
  if (st && st->done) return;
  const int u = blockIdx.x * blockDim.x + threadIdx.x;   // fastest index of the face
  const int v = blockIdx.y * blockDim.y + threadIdx.y;
  const int f = blockIdx.z;                               // face id, cz_fparam.fi order
  if (!a.face[f]) return;
  if (f >= 4) {          // K faces: (i,j) loops 1..ix, 1..jx
    const int i = u + 1, j = v + 1;
    if (i > G.ix || j > G.jx) return;
    // the reference applies the K faces first and the I/J faces (zero) afterwards; all faces
    // run concurrently here, so edge nodes shared with an active I/J face are left to it.
    if ((i == 1 && a.face[0]) || (i == G.ix && a.face[1]) || (j == 1 && a.face[2]) || (j == G.jx && a.face[3])) return;
    REAL val;
    if (a.tbl) {
      val = a.tbl[(size_t)(j - 1) * G.ix + (i - 1)];
    } else {
      const REAL pi = real_pi();
      const REAL x = a.ox + a.dh * (REAL)(i + a.ioff - 1);
      const REAL y = a.oy + a.dh * (REAL)(j + a.joff - 1);
      val = real_sin(pi * x) * real_sin(pi * y);
    }
    p[G.at(f == 4 ? 1 : G.kx, i, j)] = val;
  } else if (f < 2) {    // I faces: (k,j)
    const int k = u + 1, j = v + 1;
    if (k > G.kx || j > G.jx) return;
    p[G.at(k, f == 0 ? 1 : G.ix, j)] = (REAL)0.0;
  } else {               // J faces: (k,i)
    const int k = u + 1, i = v + 1;
    if (k > G.kx || i > G.ix) return;
    p[G.at(k, i, f == 2 ? 1 : G.jx)] = (REAL)0.0;
  }
}

// exact_t (cz_utility.f90:52-82) over 1..sz
__global__ void k_exact(Geo G, REAL dh, REAL ox, REAL oy, REAL oz, REAL* __restrict__ e) {
  const int k = 1 + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = 1 + blockIdx.y * blockDim.y + threadIdx.y;
  const int j = 1 + blockIdx.z;
  if (k > G.kx || i > G.ix) return;
  const REAL r2 = sqrt((REAL)2.0);
  const REAL pi = real_pi();
  const REAL x = ox + dh * (REAL)(i - 1);
  const REAL y = oy + dh * (REAL)(j - 1);
  const REAL z = oz + dh * (REAL)(k - 1);
  e[G.at(k, i, j)] = real_sin(pi * x) * real_sin(pi * y) / real_sinh(r2 * pi) *
                     (real_sinh(r2 * pi * z) - real_sinh(r2 * pi * (z - (REAL)1.0)));
}

// err_t (cz_utility.f90:86-129): e = p - e ; max |e| and the first location in (j,i,k) order.
// Key = (double bits of |r|) compared first, then smaller loop-order index wins.
__global__ void __launch_bounds__(256) k_err(Geo G, int jlen, const REAL* __restrict__ p, REAL* __restrict__ e,
                                             unsigned long long* out2) {
  const int k = G.kst + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = G.ist + blockIdx.y * blockDim.y + threadIdx.y;
  const int j0 = G.jst + blockIdx.z * jlen;
  const int j1 = min(j0 + jlen - 1, G.jed);
  double best = -1.0; unsigned long long bidx = ~0ull;
  if (k <= G.ked && i <= G.ied) {
    const unsigned long long nkk = G.ked - G.kst + 1, nii = G.ied - G.ist + 1;
    for (int j = j0; j <= j1; ++j) {
      const long long off = G.at(k, i, j);
      const REAL r = p[off] - e[off];
      e[off] = r;
      const double q = fabs((double)r);
      const unsigned long long ord = ((unsigned long long)(j - G.jst) * nii + (i - G.ist)) * nkk + (k - G.kst);
      if (q > best || (q == best && ord < bidx)) { best = q; bidx = ord; }
    }
  }
  // block reduce through shared memory, then a global CAS loop on the (value, index) pair
  __shared__ double s_v[256]; __shared__ unsigned long long s_i[256];
  const int tid = threadIdx.x + blockDim.x * threadIdx.y, n = blockDim.x * blockDim.y;
  s_v[tid] = best; s_i[tid] = bidx;
  __syncthreads();
  if (tid == 0) {
    for (int t = 1; t < n; ++t)
      if (s_v[t] > best || (s_v[t] == best && s_i[t] < bidx)) { best = s_v[t]; bidx = s_i[t]; }
    if (best >= 0.0) {
      // spin lock on out2[2]
      while (atomicCAS(&out2[2], 0ull, 1ull) != 0ull) {}
      __threadfence();
      const double cur = __longlong_as_double((long long)out2[0]);
      const unsigned long long ci = out2[1];
      // reference semantic: strict '<' in loop order => location updated only when the max grows;
      // an all-zero field leaves loc = (-1,-1,-1) -> ci stays ~0
      if (best > cur || (best == cur && best > 0.0 && bidx < ci)) {
        out2[0] = (unsigned long long)__double_as_longlong(best);
        out2[1] = bidx;
      }
      __threadfence();
      atomicExch(&out2[2], 0ull);
    }
  }
}

// dst <- src on every cell that is NOT an inner cell (guard cells, boundary nodes): what the second buffer of a
// ping-pong loop must share with the first before the first sweep (the sweeps store every inner cell and carry the
// other cells of the units they store over from their source).  The whole planes below jst / above jed are two
// contiguous blocks (copied by the launcher); this kernel does the planes between: one block per plane, a THREAD per
// row for the few cells left and right of the inner k range (so that a plane costs one memory round trip, not one per
// row), all threads together for the whole rows below ist / above ied, which are contiguous at both ends of a plane.
__global__ void __launch_bounds__(256) k_copy_shell(Geo G, REAL* __restrict__ dst, const REAL* __restrict__ src) {
  const int klo = G.kst + G.g - 1, khi = G.ked + G.g - 1;      // first / last inner C index along k
  const int ilo = G.ist + G.g - 1, ihi = G.ied + G.g - 1;      // first / last inner C row
  const long long nlow = (long long)ilo * G.nk, nhigh = (long long)(G.ni - 1 - ihi) * G.nk;
  for (int jF = G.jst + blockIdx.x; jF <= G.jed; jF += gridDim.x) {
    const long long pl = (long long)(jF + G.g - 1) * G.plane;
    for (int ic = ilo + threadIdx.x; ic <= ihi; ic += blockDim.x) {
      const long long row = pl + (long long)ic * G.nk;
      for (int k = 0; k < klo; ++k) dst[row + k] = src[row + k];
      for (int k = khi + 1; k < G.nk; ++k) dst[row + k] = src[row + k];
    }
    const long long hi0 = pl + (long long)(ihi + 1) * G.nk;
    for (long long e = threadIdx.x; e < nlow; e += blockDim.x) dst[pl + e] = src[pl + e];
    for (long long e = threadIdx.x; e < nhigh; e += blockDim.x) dst[hi0 + e] = src[hi0 + e];
  }
}

__global__ void k_state_reset(LoopState* st) {
  st->done = 0; st->itr = 0; st->status = 0; st->pad = 0; st->res = 0.0;
  for (int a = 0; a < 8; ++a) { st->acc[a] = 0.0; st->scal[a] = 0.0; }
}

__global__ void k_finish_iteration(LoopState* st, int slot, double res_normal, double eps, double* hist,
                                   int hist_len, int slot2) {
  if (st->done) return;
  double res = st->acc[slot];
  if (slot2 >= 0) res += st->acc[slot2];     // edge launch + interior launch of a j-slab
  res *= res_normal;          // cz_Poisson.cpp:69-70
  res = sqrt(res);
  st->res = res;
  const int it = st->itr;
  if (hist && it < hist_len) hist[it] = res;
  st->itr = it + 1;
  if (res < eps) st->done = 1;
}

}  // namespace

// ============================================================================ launchers
void launch_jacobi(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b, const REAL* cf,
                   REAL omg, int slot, bool add) {
  run_cells<1>(L, G, OpJacobi{p, pnew, b, make_coef(cf), omg}, FinStore{L.acc_st, slot, add ? 1 : 0});
}

void launch_rb_color(const Launch& L, const Geo& G, REAL* p, const REAL* b, const REAL* cf, REAL omg, int kp,
                     int slot, bool add) {
  const int kcells = (G.ked - G.kst + 2) / 2;
  Tile t = make_tile(G, kcells, L.rb_blocks);
  k_rb_color<<<t.grid, t.block, 0, L.s>>>(G, t.jlen, L.st, p, b, make_coef(cf), omg, kp, L.rb,
                                           FinStore{L.acc_st, slot, add ? 1 : 0});
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

void launch_copy_inner(const Launch& L, const Geo& G, REAL* dst, const REAL* src) {
  run_cells<0>(L, G, OpCopy{dst, src}, FinNone{});
}

void launch_calc_ax(const Launch& L, const Geo& G, REAL* ap, const REAL* p, const REAL* cf, const REAL* dotwith,
                    int slot) {
  if (dotwith) run_cells<1>(L, G, OpAx{ap, p, make_coef(cf), dotwith}, FinStore{L.acc_st, slot, 0});
  else         run_cells<0>(L, G, OpAx{ap, p, make_coef(cf), nullptr}, FinNone{});
}

void launch_calc_rk(const Launch& L, const Geo& G, REAL* r, const REAL* p, const REAL* b, const REAL* cf) {
  run_cells<0>(L, G, OpRk{r, p, b, make_coef(cf)}, FinNone{});
}

void launch_dot1(const Launch& L, const Geo& G, const REAL* p, int slot) {
  run_cells<1>(L, G, OpDot1{p}, FinStore{L.acc_st, slot, 0});
}
void launch_dot2(const Launch& L, const Geo& G, const REAL* p, const REAL* q, int slot) {
  run_cells<1>(L, G, OpDot2{p, q}, FinStore{L.acc_st, slot, 0});
}
void launch_triad(const Launch& L, const Geo& G, REAL* z, const REAL* x, const REAL* y, REAL a, const double* a_dev) {
  run_cells<0>(L, G, OpTriad{z, x, y, a, a_dev}, FinNone{});
}
void launch_bicg1(const Launch& L, const Geo& G, REAL* p, const REAL* r, const REAL* q, REAL beta, REAL omg,
                  const double* beta_dev, const double* omg_dev) {
  run_cells<0>(L, G, OpBicg1{p, r, q, beta, omg, beta_dev, omg_dev}, FinNone{});
}
void launch_bicg2(const Launch& L, const Geo& G, REAL* z, const REAL* x, const REAL* y, REAL a, REAL b,
                  const double* a_dev, const double* b_dev) {
  run_cells<0>(L, G, OpBicg2{z, x, y, a, b, a_dev, b_dev}, FinNone{});
}

void launch_bicg_tail(const Launch& L, const Geo& G, REAL* x, const REAL* p_, const REAL* s_, REAL* r, const REAL* t,
                      const REAL* s, const REAL* r0, const double* al_dev, const double* om_dev, const double* nom_dev,
                      int slot_rr, int slot_rho) {
  run_cells<2>(L, G, OpBicgTail{x, p_, s_, r, t, s, r0, al_dev, om_dev, nom_dev}, FinStore2{L.acc_st, slot_rr, slot_rho});
}

void launch_imask(const Launch& L, const Geo& G, REAL* x) {   // cz_blas.f90:24-103
  CZB_CHECK(cudaMemsetAsync(x, 0, sizeof(REAL) * (size_t)G.total, L.s));
  Launch L2 = L; L2.st = nullptr;
  run_cells<0>(L2, G, OpOne{x}, FinNone{});
}

void launch_copy_in(const Launch& L, const Geo& G, REAL* y, const REAL* x) {   // cz_blas.f90:205-242
  Geo H = G; H.ist = 1; H.ied = G.ix; H.jst = 1; H.jed = G.jx; H.kst = 1; H.ked = G.kx;
  run_cells<0>(L, H, OpCopy{y, x}, FinNone{});
}

// Dirichlet data of the K faces.  The reference evaluates sin() with the HOST libm
// (cz_solver.f90:52-58 -> sinf/sin), whose float sine is not correctly rounded (glibc 2.39:
// ~1.3 % of arguments differ from the correctly rounded value, and the result depends on the
// CPU's ifunc variant), and a 1-ulp change of a boundary value moves the late residual history
// by ~1e-4 relative.  To make the boundary DATA bit-identical to the reference on the same box,
// the ix*jx table is evaluated once per (size, dh, origin) with the same libm calls and uploaded;
// writing the faces every iteration stays a device kernel.  CZB_BC_SIN=device evaluates
// sin() on the device instead (differs by <= 1 ulp at a few boundary nodes).
struct BcTable { int ix, jx, joff, ioff; REAL dh, ox, oy; REAL* dev; };
static std::vector<BcTable> g_bc_tables;

static const REAL* bc_table(const Geo& G, REAL dh, REAL ox, REAL oy, int joff, int ioff, cudaStream_t s) {
  static int mode = -1;
  if (mode < 0) { const char* e = getenv("CZB_BC_SIN"); mode = (e && !strcmp(e, "device")) ? 1 : 0; }
  if (mode == 1) return nullptr;
  for (const BcTable& t : g_bc_tables)
    if (t.ix == G.ix && t.jx == G.jx && t.joff == joff && t.ioff == ioff && t.dh == dh && t.ox == ox && t.oy == oy) return t.dev;
  std::vector<REAL> h((size_t)G.ix * G.jx);
  const REAL pi = (REAL)2.0 * std::asin((REAL)1.0);
  std::vector<REAL> sx(G.ix);
  for (int i = 1; i <= G.ix; ++i) { const REAL x = ox + dh * (REAL)(i + ioff - 1); sx[i - 1] = std::sin(pi * x); }   // ioff = 0 in the reference; bricks use global i
  for (int j = 1; j <= G.jx; ++j) {
    const REAL y = oy + dh * (REAL)(j + joff - 1);   // joff = 0 in the reference; slabs use global j
    const REAL sy = std::sin(pi * y);
    for (int i = 1; i <= G.ix; ++i) h[(size_t)(j - 1) * G.ix + (i - 1)] = sx[i - 1] * sy;
  }
  BcTable t{G.ix, G.jx, joff, ioff, dh, ox, oy, nullptr};
  CZB_CHECK(cudaMalloc(&t.dev, sizeof(REAL) * h.size()));
  CZB_CHECK(cudaMemcpyAsync(t.dev, h.data(), sizeof(REAL) * h.size(), cudaMemcpyHostToDevice, s));
  CZB_CHECK(cudaStreamSynchronize(s));
  if (g_bc_tables.size() >= 8) { cudaFree(g_bc_tables.front().dev); g_bc_tables.erase(g_bc_tables.begin()); }
  g_bc_tables.push_back(t);
  return t.dev;
}

void launch_bc_k(const Launch& L, const Geo& G, REAL* p, REAL dh, const REAL* org, const int* nID, int joff, int ioff) {
  BcArgs a; a.dh = dh; a.ox = org[0]; a.oy = org[1]; a.tbl = nullptr; a.joff = joff; a.ioff = ioff;
  bool any = false;
  for (int f = 0; f < 6; ++f) { a.face[f] = nID[f] < 0 ? 1 : 0; any |= a.face[f] != 0; }
  if (!any) return;
  if (a.face[4] || a.face[5]) a.tbl = bc_table(G, dh, org[0], org[1], joff, ioff, L.s);
  int m = G.ix; if (G.jx > m) m = G.jx; if (G.kx > m) m = G.kx;
  dim3 block(32, 8, 1), grid((m + 31) / 32, (m + 7) / 8, 6);
  k_bc<<<grid, block, 0, L.s>>>(G, a, L.st, p);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

void launch_exact(const Launch& L, const Geo& G, REAL* e, REAL dh, const REAL* org) {
  dim3 block(64, 4, 1), grid((G.kx + 63) / 64, (G.ix + 3) / 4, G.jx);
  k_exact<<<grid, block, 0, L.s>>>(G, dh, org[0], org[1], org[2], e);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

void launch_err(const Launch& L, const Geo& G, const REAL* p, REAL* e, unsigned long long* out2) {
  const int kcells = G.ked - G.kst + 1;
  Tile t = make_tile(G, kcells, (size_t)1 << 30);
  // out2[0] = bits of 0.0 (dm starts at 0, cz_utility.f90), out2[1] = ~0 (no location), out2[2] = lock
  unsigned long long init[3] = {0ull, ~0ull, 0ull};
  CZB_CHECK(cudaMemcpyAsync(out2, init, sizeof(init), cudaMemcpyHostToDevice, L.s));
  CZB_CHECK(cudaStreamSynchronize(L.s));
  k_err<<<t.grid, t.block, 0, L.s>>>(G, t.jlen, p, e, out2);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

void launch_psor_sweep(const Launch& L, const Geo& G, REAL* p, const REAL* b, const REAL* cf, REAL omg, int slot,
                       bool add) {
  const int ni = G.ied - G.ist + 1, nj = G.jed - G.jst + 1;
  dim3 block(32, 8, 1), grid((ni + 31) / 32, (nj + 7) / 8, 1);
  bool first = true;
  for (int hp = G.ist + G.jst + G.kst; hp <= G.ied + G.jed + G.ked; ++hp) {
    k_psor_plane<<<grid, block, 0, L.s>>>(G, L.st, p, b, make_coef(cf), omg, hp, L.rb,
                                          FinStore{L.acc_st, slot, (add || !first) ? 1 : 0});
    first = false;
    count_launch();
  }
  CZB_CHECK(cudaGetLastError());
}

void launch_copy_shell(const Launch& L, const Geo& G, REAL* dst, const REAL* src) {
  // whole planes: C planes 0 .. jst+g-2 and jed+g .. nj-1
  const long long lo = (long long)(G.jst + G.g - 1) * G.plane, hi0 = (long long)(G.jed + G.g) * G.plane;
  if (lo > 0) CZB_CHECK(cudaMemcpyAsync(dst, src, sizeof(REAL) * lo, cudaMemcpyDeviceToDevice, L.s));
  if (G.total > hi0) CZB_CHECK(cudaMemcpyAsync(dst + hi0, src + hi0, sizeof(REAL) * (G.total - hi0), cudaMemcpyDeviceToDevice, L.s));
  const int npj = G.jed - G.jst + 1;
  if (npj > 0) {
    const int grid = npj < 4096 ? npj : 4096;
    k_copy_shell<<<grid, 256, 0, L.s>>>(G, dst, src);
    CZB_CHECK(cudaGetLastError());
    count_launch();
  }
}

void launch_state_reset(const Launch& L, LoopState* st) {
  k_state_reset<<<1, 1, 0, L.s>>>(st);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

void launch_finish_iteration(const Launch& L, LoopState* st, int slot, double res_normal, double eps, double* hist,
                             int hist_len, int slot2) {
  k_finish_iteration<<<1, 1, 0, L.s>>>(st, slot, res_normal, eps, hist, hist_len, slot2);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

}  // namespace czb

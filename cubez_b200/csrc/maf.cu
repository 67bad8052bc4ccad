// cubez_b200/csrc/maf.cu -- stretched-grid ("MAF") twins of the sweeps: reference
// src/cz_f90/cz_maf.f90 (psor_maf :23, jacobi_maf :131, psor2sma_core_maf :301, pcr_rb_maf :442, pcr_maf :672,
// pcr_eda_maf :896, pcr_esa_maf :1117, pcr_rb_esa_maf :1339) and cz_blas.f90 (calc_rk_maf :738, calc_ax_maf :845,
// search_pivot :947).
//
// Same data movement as the uniform-grid kernels plus three coordinate vectors X(-1:sz(1)+2), Y, Z (Fortran index
// m -> offset m+1; a few KB, served by L1).  The metric coefficients C1..C9 are products of factors from all three
// axes and are therefore recomputed per cell in the reference's statement order (they round per cell; a per-axis
// table would not be bit-exact); the library is built with -fmad=false and IEEE division, so results equal the
// reference's un-contracted CPU arithmetic bit for bit.  The per-thread parts (the i- and k-differences of a thread
// marching over j) are loop invariant.  For the line solvers the PCR coefficients a, c depend on the line, so they
// are carried through the stages in shared memory next to d (no shared table as in pcr.cu).
#include "kernels.cuh"
#include "cells.cuh"

namespace czb {
namespace {

struct MafCell { REAL C1, C2, C3, C7, C8, C9; };

// cz_maf.f90:70-99 (the identical block opens every point routine)
__device__ __forceinline__ MafCell maf_cell(const MafAxes& A, int i, int j, int k) {
  const REAL half = (REAL)0.5, two = (REAL)2.0;
  const REAL xp = __ldg(A.X + i + 2), xc = __ldg(A.X + i + 1), xm = __ldg(A.X + i);
  const REAL yp = __ldg(A.Y + j + 2), yc = __ldg(A.Y + j + 1), ym = __ldg(A.Y + j);
  const REAL zp = __ldg(A.Z + k + 2), zc = __ldg(A.Z + k + 1), zm = __ldg(A.Z + k);
  const REAL XG = half * (xp - xm);
  const REAL YE = half * (yp - ym);
  const REAL ZT = half * (zp - zm);
  const REAL XGG = xp - two * xc + xm;
  const REAL YEE = yp - two * yc + ym;
  const REAL ZTT = zp - two * zc + zm;
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

// the six neighbour products, left to right (cz_maf.f90:103-108)
__device__ __forceinline__ REAL maf_offdiag(const Geo& G, const MafCell& m, const REAL* __restrict__ p, long long off) {
  const REAL half = (REAL)0.5;
  return (m.C1 + half * m.C7) * p[off + G.nk] + (m.C1 - half * m.C7) * p[off - G.nk] +
         (m.C2 + half * m.C8) * p[off + G.plane] + (m.C2 - half * m.C8) * p[off - G.plane] +
         (m.C3 + half * m.C9) * p[off + 1] + (m.C3 - half * m.C9) * p[off - 1];
}

// dp of the point sweeps (cz_maf.f90:102-110); the source term is ADDED ("+ bb")
__device__ __forceinline__ REAL maf_dp(const Geo& G, const MafCell& m, const REAL* __restrict__ p, long long off,
                                       REAL bb, REAL pp, REAL omg) {
  const REAL dd = (REAL)2.0 * (m.C1 + m.C2 + m.C3);
  const REAL rp = maf_offdiag(G, m, p, off) + bb;
  return (rp / dd - pp) * omg;
}

struct OpJacobiMaf {   // cz_maf.f90:131-282 without the copy-back pass
  const REAL* p; REAL* pn; const REAL* b; MafAxes A; REAL omg;
  __device__ void operator()(const Geo& G, int k, int i, int j, long long off, double* acc) const {
    const REAL bb = b[off], pp = p[off];
    const REAL dp = maf_dp(G, maf_cell(A, i, j, k), p, off, bb, pp, omg);
    pn[off] = pp + dp;
    acc[0] += (double)(dp * dp);
  }
};

struct OpAxMaf {   // cz_blas.f90:845-935
  REAL* ap; const REAL* p; const REAL* pvt; MafAxes A; const REAL* w;
  __device__ void operator()(const Geo& G, int k, int i, int j, long long off, double* acc) const {
    const MafCell m = maf_cell(A, i, j, k);
    const REAL v = (maf_offdiag(G, m, p, off) - (REAL)2.0 * (m.C1 + m.C2 + m.C3) * p[off]) * pvt[off];
    ap[off] = v;
    if (w) acc[0] += (double)(v * w[off]);
  }
};

struct OpRkMaf {   // cz_blas.f90:738-829: every neighbour product subtracted in turn
  REAL* r; const REAL* p; const REAL* b; const REAL* pvt; MafAxes A;
  __device__ void operator()(const Geo& G, int k, int i, int j, long long off, double*) const {
    const MafCell m = maf_cell(A, i, j, k);
    const REAL half = (REAL)0.5;
    r[off] = (b[off] + (REAL)2.0 * (m.C1 + m.C2 + m.C3) * p[off] - (m.C1 + half * m.C7) * p[off + G.nk] -
              (m.C1 - half * m.C7) * p[off - G.nk] - (m.C2 + half * m.C8) * p[off + G.plane] -
              (m.C2 - half * m.C8) * p[off - G.plane] - (m.C3 + half * m.C9) * p[off + 1] -
              (m.C3 - half * m.C9) * p[off - 1]) * pvt[off];
  }
};

struct OpPivot {   // cz_blas.f90:947-1039
  REAL* pvt; MafAxes A;
  __device__ void operator()(const Geo&, int k, int i, int j, long long off, double*) const {
    const MafCell m = maf_cell(A, i, j, k);
    const REAL half = (REAL)0.5;
    const REAL s1 = fabs(m.C1 + half * m.C7), s2 = fabs(m.C1 - half * m.C7);
    const REAL s3 = fabs(m.C2 + half * m.C8), s4 = fabs(m.C2 - half * m.C8);
    const REAL s5 = fabs(m.C3 + half * m.C9), s6 = fabs(m.C3 - half * m.C9);
    const REAL s7 = fabs((REAL)2.0 * (m.C1 + m.C2 + m.C3));
    const REAL ss = fmax(fmax(fmax(s1, s2), fmax(s3, s4)), fmax(fmax(s5, s6), s7));
    pvt[off] = (REAL)1.0 / ss;
  }
};

// One colour of red-black SOR in place (cz_maf.f90:301-438): k = kst + mod(i+j+kp,2) + 2t, as k_rb_color.
template <class Fin>
__global__ void __launch_bounds__(256) k_rb_color_maf(Geo G, int jlen, const LoopState* st, REAL* __restrict__ p,
                                                      const REAL* __restrict__ b, MafAxes A, REAL omg, int kp,
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
        const REAL pp = p[off], bb = b[off];
        const REAL dp = maf_dp(G, maf_cell(A, i, j, k), p, off, bb, pp, omg);
        p[off] = pp + dp;
        acc[0] += (double)(dp * dp);
      }
    }
  }
  grid_reduce<1>(acc, rb, fin);
}

// Lexicographic point SOR (cz_maf.f90:23-114) in the reference's 1-thread order: hyperplane i+j+k == hp, as
// k_psor_plane.
template <class Fin>
__global__ void __launch_bounds__(256) k_psor_plane_maf(Geo G, const LoopState* st, REAL* __restrict__ p,
                                                        const REAL* __restrict__ b, MafAxes A, REAL omg, int hp,
                                                        RedBuf rb, Fin fin) {
  if (st && st->done) return;
  const int i = G.ist + blockIdx.x * blockDim.x + threadIdx.x;
  const int j = G.jst + blockIdx.y * blockDim.y + threadIdx.y;
  double acc[1] = {0.0};
  const int k = hp - i - j;
  if (i <= G.ied && j <= G.jed && k >= G.kst && k <= G.ked) {
    const long long off = G.at(k, i, j);
    const REAL pp = p[off], bb = b[off];
    const REAL dp = maf_dp(G, maf_cell(A, i, j, k), p, off, bb, pp, omg);
    p[off] = pp + dp;
    acc[0] += (double)(dp * dp);
  }
  grid_reduce<1>(acc, rb, fin);
}

// ------------------------------------------------------------------------------------------------ line SOR
// One CTA per line at a time (persistent), a / c / d ping-pong in shared memory: 6 rows of n REALs.
// cz_maf.f90:522-647 (pcr_rb_maf); pcr_maf :744-870, pcr_eda_maf :981-1093, pcr_esa_maf :1194-1314 and
// pcr_rb_esa_maf :1419-1539 run the same statements (pn-1 PCR stages + the 2x2 last stage in all five); their
// clamped (kst-1 / ked+1) and zero-padded neighbour indices both read an exact zero coefficient, `zero` here.
// mode: 0/1 = lines with mod(i+j,2) == mode, 3 = lines with i+j == diag (wavefront for the lexicographic ones).
struct MafLineSel { int mode, diag, jlo, nlines, per_j; };

template <class Fin>
__global__ void __launch_bounds__(256) k_pcr_lines_maf(Geo G, int n, int pn, MafLineSel sel, MafAxes A, REAL* x,
                                                       const REAL* __restrict__ msk, const REAL* __restrict__ rhs,
                                                       REAL omg, const LoopState* st, RedBuf rb, Fin fin) {
  if (st && st->done) return;
  extern __shared__ double sm_raw_maf[];
  REAL* a0 = reinterpret_cast<REAL*>(sm_raw_maf);
  REAL* c0 = a0 + n; REAL* d0 = c0 + n; REAL* a1 = d0 + n; REAL* c1 = a1 + n; REAL* d1 = c1 + n;
  const int tid = threadIdx.x, NT = blockDim.x;
  const REAL zero = (REAL)0.0, one = (REAL)1.0, half = (REAL)0.5, two = (REAL)2.0;
  double acc[1] = {0.0};
  for (int line = blockIdx.x; line < sel.nlines; line += gridDim.x) {
    int i, j;
    if (sel.mode <= 1) {          // mod(i+j,2) == color  (cz_maf.f90:526)
      const int jj = line / sel.per_j, m = line - jj * sel.per_j;
      j = G.jst + jj;
      i = G.ist + ((G.ist + j + sel.mode) & 1) + 2 * m;
    } else {                      // wavefront i+j == diag
      j = sel.jlo + line; i = sel.diag - j;
    }
    if (i > G.ied) continue;      // block-uniform
    const long long base = G.at(G.kst, i, j);
    // per-line metrics (cz_maf.f90:528-537)
    const REAL xp = __ldg(A.X + i + 2), xc = __ldg(A.X + i + 1), xm = __ldg(A.X + i);
    const REAL yp = __ldg(A.Y + j + 2), yc = __ldg(A.Y + j + 1), ym = __ldg(A.Y + j);
    const REAL GX = two / (xp - xm);
    const REAL EY = two / (yp - ym);
    const REAL C1 = GX * GX;
    const REAL C2 = EY * EY;
    const REAL C7 = -(xp - two * xc + xm) * C1 * GX;
    const REAL C8 = -(yp - two * yc + ym) * C2 * EY;
    const REAL dd1 = C1 + half * C7, dd2 = C1 - half * C7;
    const REAL cc1 = C2 + half * C8, cc2 = C2 - half * C8;
    // coefficients (:540-557), source term (:560-568) and Dirichlet ends (:572-573)
    for (int k = tid; k < n; k += NT) {
      const int kf = G.kst + k;                       // Fortran k
      const REAL f1 = __ldg(A.Z + kf + 2), f2 = __ldg(A.Z + kf);
      const REAL TZ = two / (f1 - f2);
      const REAL ZTT = f1 - two * __ldg(A.Z + kf + 1) + f2;
      const REAL f3 = TZ * TZ;
      const REAL aw = f3;                             // C3
      const REAL cw = -ZTT * f3 * TZ;                 // C9
      const REAL dw = half / (C1 + C2 + f3);          // 1/R7
      a0[k] = (k == 0) ? zero : -(aw - half * cw) * dw;
      c0[k] = (k == n - 1) ? zero : -(aw + half * cw) * dw;
      const long long off = base + k;
      const REAL mk = msk[off];
      REAL d = (dd1 * x[off + G.nk] + dd2 * x[off - G.nk] + cc1 * x[off + G.plane] + cc2 * x[off - G.plane] - rhs[off]) *
               dw * mk;
      if (k == 0) d = (d + (aw - half * cw) * dw * x[off - 1]) * mk;
      if (k == n - 1) d = (d + (aw + half * cw) * dw * x[off + 1]) * mk;
      d0[k] = d;
    }
    __syncthreads();
    REAL *pa = a0, *pc = c0, *pd = d0, *qa = a1, *qc = c1, *qd = d1;
    for (int p = 0; p < pn - 1; ++p) {                // :580-601
      const int s = 1 << p;
      for (int k = tid; k < n; k += NT) {
        const REAL ap = pa[k], cp = pc[k];
        const bool lo = k >= s, hi = k + s < n;
        const REAL al = lo ? pa[k - s] : zero, cl = lo ? pc[k - s] : zero, dl = lo ? pd[k - s] : zero;
        const REAL ar = hi ? pa[k + s] : zero, cr = hi ? pc[k + s] : zero, dr = hi ? pd[k + s] : zero;
        const REAL e = one / (one - ap * cl - cp * ar);
        qa[k] = -e * ap * al;
        qc[k] = -e * cp * cr;
        qd[k] = e * (pd[k] - ap * dl - cp * dr);
      }
      __syncthreads();
      REAL* t;
      t = pa; pa = qa; qa = t;
      t = pc; pc = qc; qc = t;
      t = pd; pd = qd; qd = t;
    }
    // 2x2 last stage (:605-620) folded into the relaxation (:630-637)
    const int s2 = 1 << (pn - 1);
    for (int k = tid; k < n; k += NT) {
      REAL sol;
      if (k < s2) {
        const bool hi = k + s2 < n;
        const REAL cc = pc[k], aa = hi ? pa[k + s2] : zero;
        const REAL f1 = pd[k], f2 = hi ? pd[k + s2] : zero;
        const REAL jj = one / (one - aa * cc);
        sol = (f1 - cc * f2) * jj;
      } else {
        const int kk = k - s2;
        const REAL cc = pc[kk], aa = pa[k];
        const REAL f1 = pd[kk], f2 = pd[k];
        const REAL jj = one / (one - aa * cc);
        sol = (f2 - aa * f1) * jj;
      }
      const long long off = base + k;
      const REAL pp = x[off];
      const REAL dp = (sol - pp) * omg * msk[off];
      x[off] = pp + dp;
      acc[0] += (double)(dp * dp);
    }
    __syncthreads();
  }
  grid_reduce<1>(acc, rb, fin);
}

// The five MAF line routines add every dp*dp ONCE: `tmp(k) = tmp(k) + dp*dp` (summed into res1 afterwards) under
// -D_SVR, `res1 = res1 + dp*dp` otherwise -- the two statements are the branches of one #ifdef (cz_maf.f90:642-646,
// 659-663).  (Round 1 read both as unconditional and doubled the sum; found by executing the reference's Fortran.)

}  // namespace

void launch_jacobi_maf(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b, const MafAxes& A,
                       REAL omg, int slot, bool add) {
  run_cells<1>(L, G, OpJacobiMaf{p, pnew, b, A, omg}, FinStore{L.acc_st, slot, add ? 1 : 0});
}

void launch_rb_color_maf(const Launch& L, const Geo& G, REAL* p, const REAL* b, const MafAxes& A, REAL omg, int kp,
                         int slot, bool add) {
  const int kcells = (G.ked - G.kst + 2) / 2;
  Tile t = make_tile(G, kcells, L.rb_blocks);
  k_rb_color_maf<<<t.grid, t.block, 0, L.s>>>(G, t.jlen, L.st, p, b, A, omg, kp, L.rb,
                                               FinStore{L.acc_st, slot, add ? 1 : 0});
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

void launch_psor_sweep_maf(const Launch& L, const Geo& G, REAL* p, const REAL* b, const MafAxes& A, REAL omg, int slot,
                           bool add) {
  const int ni = G.ied - G.ist + 1, nj = G.jed - G.jst + 1;
  dim3 block(32, 8, 1), grid((ni + 31) / 32, (nj + 7) / 8, 1);
  bool first = true;
  for (int hp = G.ist + G.jst + G.kst; hp <= G.ied + G.jed + G.ked; ++hp) {
    k_psor_plane_maf<<<grid, block, 0, L.s>>>(G, L.st, p, b, A, omg, hp, L.rb,
                                              FinStore{L.acc_st, slot, (add || !first) ? 1 : 0});
    first = false;
    count_launch();
  }
  CZB_CHECK(cudaGetLastError());
}

void launch_calc_ax_maf(const Launch& L, const Geo& G, REAL* ap, const REAL* p, const MafAxes& A, const REAL* pvt,
                        const REAL* dotwith, int slot) {
  if (dotwith) run_cells<1>(L, G, OpAxMaf{ap, p, pvt, A, dotwith}, FinStore{L.acc_st, slot, 0});
  else         run_cells<0>(L, G, OpAxMaf{ap, p, pvt, A, nullptr}, FinNone{});
}

void launch_calc_rk_maf(const Launch& L, const Geo& G, REAL* r, const REAL* p, const REAL* b, const MafAxes& A,
                        const REAL* pvt) {
  run_cells<0>(L, G, OpRkMaf{r, p, b, pvt, A}, FinNone{});
}

void launch_search_pivot(const Launch& L, const Geo& G, REAL* pvt, const MafAxes& A) {
  run_cells<0>(L, G, OpPivot{pvt, A}, FinNone{});
}

void launch_pcr_lines_maf(const Launch& L, const Geo& G, int pn, int mode, int diag, const MafAxes& A, REAL* x,
                          const REAL* msk, const REAL* rhs, REAL omg, int slot, bool add) {
  const int n = G.ked - G.kst + 1;
  const int ni = G.ied - G.ist + 1, nj = G.jed - G.jst + 1;
  if (n < 2 || pn < 1 || pn > 24 || (1 << pn) <= n) {   // CZ::getNumStage (cz.h:293-300) gives 2^(pn-1) <= n < 2^pn
    fprintf(stderr, "czb200: MAF line SOR needs n >= 2 and n < 2^pn (n=%d, pn=%d)\n", n, pn);
    abort();
  }
  MafLineSel sel; sel.mode = mode; sel.diag = diag; sel.jlo = 0;
  if (mode <= 1) { sel.per_j = (ni + 1) / 2; sel.nlines = sel.per_j * nj; }
  else {
    int jlo = diag - G.ied; if (jlo < G.jst) jlo = G.jst;
    int jhi = diag - G.ist; if (jhi > G.jed) jhi = G.jed;
    sel.jlo = jlo; sel.per_j = 1; sel.nlines = jhi >= jlo ? jhi - jlo + 1 : 0;
  }
  if (sel.nlines <= 0) return;
  int nt = 32; while (nt < n && nt < 256) nt <<= 1;
  int grid = sel.nlines < 148 * 8 ? sel.nlines : 148 * 8;
  if ((size_t)grid > L.rb_blocks) grid = (int)L.rb_blocks;
  const size_t smem = 6 * (size_t)n * sizeof(REAL);
  typedef FinStore F;
  static size_t sm_set = 0;
  if (smem > sm_set) {
    CZB_CHECK(cudaFuncSetAttribute(k_pcr_lines_maf<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    sm_set = smem;
  }
  k_pcr_lines_maf<F><<<grid, nt, smem, L.s>>>(G, n, pn, sel, A, x, msk, rhs, omg, L.st, L.rb, F{L.acc_st, slot, add ? 1 : 0});
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

}  // namespace czb

// cubez_b200/csrc/pcr.cu -- line SOR with parallel-cyclic-reduction (PCR) tridiagonal solves
// along k (reference cz_solver.f90: pcr_rb :497-662, pcr_rb_esa :1261-1469, pcr_j_esa
// :1473-1676, and the lexicographic pcr/pcr_esa/pcr_eda :666-1257 through line wavefronts).
//
// On the uniform grid the coefficient recurrences a(k), c(k), e(k) of every PCR stage are the
// same for every line (cz_solver.f90:556-564 re-initialises them to constants per line), so
// they are evaluated ONCE per (n, pn, last-stage kind) by k_pcr_tables with exactly the
// reference's operation order, and the per-line kernel reduces only the right-hand side d:
//   d1(k) = e * ( d(k) - ap*d(k-s) - cp*d(k+s) )          (cz_solver.f90:602)
// => bitwise the same d as the reference, ~3x less shared-memory traffic, no divisions.
// Out-of-range neighbours are 0 (their coefficient is exactly 0).
#include "kernels.cuh"

namespace czb {
namespace {

struct FinStoreP {
  LoopState* st; int slot; int add;
  __device__ void operator()(double (&tot)[1]) const {
    if (st) st->acc[slot] = (add ? st->acc[slot] : 0.0) + tot[0];
  }
};

// One block.  wa = scratch 4*n REAL (a, c, a1, c1).
__global__ void __launch_bounds__(1024) k_pcr_tables(int n, int pn, int stages, int last4, REAL* __restrict__ tab,
                                                     REAL* __restrict__ fin, REAL* __restrict__ wa) {
  REAL* a = wa; REAL* c = wa + n; REAL* a1 = wa + 2 * (size_t)n; REAL* c1 = wa + 3 * (size_t)n;
  const REAL r = (REAL)1.0 / (REAL)6.0;
  const REAL one = (REAL)1.0, zero = (REAL)0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) {   // cz_solver.f90:556-564
    a[k] = (k == 0) ? zero : -r;
    c[k] = (k == n - 1) ? zero : -r;
  }
  __syncthreads();
  for (int p = 0; p < stages; ++p) {                    // cz_solver.f90:584-608
    const int s = 1 << p;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
      const REAL ap = a[k], cp = c[k];
      const REAL al = (k >= s) ? a[k - s] : zero, cl = (k >= s) ? c[k - s] : zero;
      const REAL ar = (k + s < n) ? a[k + s] : zero, cr = (k + s < n) ? c[k + s] : zero;
      const REAL e = one / (one - ap * cl - cp * ar);
      a1[k] = -e * ap * al;
      c1[k] = -e * cp * cr;
      tab[((size_t)p * 3 + 0) * n + k] = ap;
      tab[((size_t)p * 3 + 1) * n + k] = cp;
      tab[((size_t)p * 3 + 2) * n + k] = e;
    }
    __syncthreads();
    REAL* t;
    t = a; a = a1; a1 = t;
    t = c; c = c1; c1 = t;
  }
  if (!last4) {                                         // cz_solver.f90:612-630
    const int s = 1 << (pn - 1);
    for (int k = threadIdx.x; k < s; k += blockDim.x) {
      const REAL cc1 = (k < n) ? c[k] : zero;
      const REAL aa2 = (k + s < n) ? a[k + s] : zero;
      fin[0 * (size_t)s + k] = cc1;
      fin[1 * (size_t)s + k] = aa2;
      fin[2 * (size_t)s + k] = one / (one - aa2 * cc1);
    }
  } else {                                              // cz_solver.f90:788-844
    const int s = 1 << (pn - 2);
    for (int k = threadIdx.x; k < s; k += blockDim.x) {
      const REAL cc1 = (k < n) ? c[k] : zero;
      const REAL cc2 = (k + s < n) ? c[k + s] : zero;
      const REAL cc3 = (k + 2 * s < n) ? c[k + 2 * s] : zero;
      const REAL aa2 = (k + s < n) ? a[k + s] : zero;
      const REAL aa3 = (k + 2 * s < n) ? a[k + 2 * s] : zero;
      const REAL aa4 = (k + 3 * s < n) ? a[k + 3 * s] : zero;
      fin[0 * (size_t)s + k] = cc1; fin[1 * (size_t)s + k] = cc2; fin[2 * (size_t)s + k] = cc3;
      fin[3 * (size_t)s + k] = aa2; fin[4 * (size_t)s + k] = aa3; fin[5 * (size_t)s + k] = aa4;
      fin[6 * (size_t)s + k] = one / (one - aa4 * cc3 - aa3 * cc2 - aa2 * cc1 * (one - cc3 * aa4));
    }
  }
}

struct LineSel { int mode, ofs, diag, jlo, nlines, per_j; };

// Persistent blocks, one line at a time per block, d in shared memory (ping-pong).
template <int LAST4>
__global__ void __launch_bounds__(256) k_pcr_lines(Geo G, int n, int pn, int stages, const REAL* __restrict__ tab,
                                                   const REAL* __restrict__ fin, LineSel sel,
                                                   const REAL* x, REAL* dst, const REAL* __restrict__ msk,
                                                   const REAL* __restrict__ rhs, REAL omg, const LoopState* st,
                                                   RedBuf rb, FinStoreP finst) {
  if (st && st->done) return;
  extern __shared__ double sm_raw[];
  REAL* dA = reinterpret_cast<REAL*>(sm_raw);
  REAL* dB = dA + n;
  const int tid = threadIdx.x, NT = blockDim.x;
  const REAL r = (REAL)1.0 / (REAL)6.0;
  const REAL zero = (REAL)0.0;
  double acc[1] = {0.0};
  for (int line = blockIdx.x; line < sel.nlines; line += gridDim.x) {
    int i, j;
    if (sel.mode <= 1) {          // mod(i+j+ofs,2) == mode  (cz_solver.f90:549)
      const int jj = line / sel.per_j, m = line - jj * sel.per_j;
      j = G.jst + jj;
      i = G.ist + ((G.ist + j + sel.ofs + sel.mode) & 1) + 2 * m;
    } else if (sel.mode == 2) {   // all lines
      const int jj = line / sel.per_j, m = line - jj * sel.per_j;
      j = G.jst + jj; i = G.ist + m;
    } else {                      // wavefront i+j == diag
      j = sel.jlo + line; i = sel.diag - j;
    }
    if (i > G.ied) continue;      // block-uniform
    const long long base = G.at(G.kst, i, j);
    // source term (cz_solver.f90:569-579)
    for (int k = tid; k < n; k += NT) {
      const long long off = base + k;
      const REAL mk = msk[off];
      REAL d = ((x[off - G.plane] + x[off + G.plane] + x[off - G.nk] + x[off + G.nk] - rhs[off]) * r) * mk;
      if (k == 0) d = (d + x[off - 1] * r) * mk;
      if (k == n - 1) d = (d + x[off + 1] * r) * mk;
      dA[k] = d;
    }
    __syncthreads();
    REAL* da = dA; REAL* db = dB;
    for (int p = 0; p < stages; ++p) {
      const int s = 1 << p;
      const REAL* tp = tab + (size_t)p * 3 * n;
      for (int k = tid; k < n; k += NT) {
        const REAL ap = tp[k], cp = tp[n + k], e = tp[2 * (size_t)n + k];
        const REAL dl = (k >= s) ? da[k - s] : zero;
        const REAL dr = (k + s < n) ? da[k + s] : zero;
        db[k] = e * (da[k] - ap * dl - cp * dr);
      }
      __syncthreads();
      REAL* t = da; da = db; db = t;
    }
    if (LAST4) {
      const int s = 1 << (pn - 2);
      for (int k = tid; k < s; k += NT) {
        const REAL cc1 = fin[k], cc2 = fin[(size_t)s + k], cc3 = fin[2 * (size_t)s + k];
        const REAL aa2 = fin[3 * (size_t)s + k], aa3 = fin[4 * (size_t)s + k], aa4 = fin[5 * (size_t)s + k];
        const REAL inv_detA = fin[6 * (size_t)s + k];
        const int kl = k + s, km = k + 2 * s, kr = k + 3 * s;
        const REAL dd1 = (k < n) ? da[k] : zero;
        const REAL dd2 = (kl < n) ? da[kl] : zero;
        const REAL dd3 = (km < n) ? da[km] : zero;
        const REAL dd4 = (kr < n) ? da[kr] : zero;
        const REAL detA1 = -cc3 * (aa4 * dd1 + cc1 * cc2 * dd4 - aa4 * cc1 * dd2)
                         + dd1 + cc1 * cc2 * dd3 - aa3 * cc2 * dd1 - cc1 * dd2;
        const REAL detA2 = dd2 + cc2 * cc3 * dd4 - aa4 * cc3 * dd2 - cc2 * dd3
                         - aa2 * (dd1 - aa4 * cc3 * dd1);
        const REAL detA3 = dd3 - cc3 * dd4 - aa3 * dd2
                         - aa2 * (cc1 * dd3 - cc1 * cc3 * dd4 - aa3 * dd1);
        const REAL detA4 = dd4 + aa3 * aa4 * dd2 - aa4 * dd3 - aa3 * cc2 * dd4
                         - aa2 * (cc1 * dd4 + aa3 * aa4 * dd1 - aa4 * cc1 * dd3);
        if (k < n) db[k] = detA1 * inv_detA;
        if (kl < n) db[kl] = detA2 * inv_detA;
        if (km < n) db[km] = detA3 * inv_detA;
        if (kr < n) db[kr] = detA4 * inv_detA;
      }
      __syncthreads();
    }
    // last 2x2 stage (cz_solver.f90:612-630) folded into the relaxation (:642-648)
    const int s2 = 1 << (pn - 1);
    for (int k = tid; k < n; k += NT) {
      REAL d1;
      if (LAST4) {
        d1 = db[k];
      } else if (k < s2) {
        const REAL f1 = da[k];
        const REAL f2 = (k + s2 < n) ? da[k + s2] : zero;
        d1 = (f1 - fin[k] * f2) * fin[2 * (size_t)s2 + k];
      } else {
        const int kk = k - s2;
        const REAL f1 = da[kk];
        const REAL f2 = da[k];
        d1 = (f2 - fin[(size_t)s2 + kk] * f1) * fin[2 * (size_t)s2 + kk];
      }
      const long long off = base + k;
      const REAL pp = x[off];
      const REAL dp = (d1 - pp) * omg * msk[off];
      dst[off] = pp + dp;
      acc[0] += (double)(dp * dp);
    }
    __syncthreads();
  }
  grid_reduce<1>(acc, rb, finst);
}

}  // namespace

void pcr_build_tables(cudaStream_t s, PcrTables& T, int n, int pn, bool last4) {
  if (T.tab && T.n == n && T.pn == pn && T.last4 == (last4 ? 1 : 0)) return;
  pcr_free_tables(T);
  if (last4 && pn < 2) { fprintf(stderr, "czb200: pcr 4x4 variant needs pn >= 2 (n=%d)\n", n); abort(); }
  T.n = n; T.pn = pn; T.last4 = last4 ? 1 : 0;
  T.stages = last4 ? pn - 2 : pn - 1;
  const size_t sfin = last4 ? (size_t)7 << (pn - 2) : (size_t)3 << (pn - 1);
  const size_t ntab = (size_t)(T.stages > 0 ? T.stages : 1) * 3 * n;
  CZB_CHECK(cudaMalloc(&T.tab, sizeof(REAL) * ntab));
  CZB_CHECK(cudaMalloc(&T.fin, sizeof(REAL) * sfin));
  REAL* wa = nullptr;
  CZB_CHECK(cudaMalloc(&wa, sizeof(REAL) * 4 * (size_t)n));
  k_pcr_tables<<<1, 1024, 0, s>>>(n, pn, T.stages, T.last4, T.tab, T.fin, wa);
  CZB_CHECK(cudaGetLastError());
  count_launch();
  CZB_CHECK(cudaStreamSynchronize(s));
  CZB_CHECK(cudaFree(wa));
}

void pcr_free_tables(PcrTables& T) {
  if (T.tab) cudaFree(T.tab);
  if (T.fin) cudaFree(T.fin);
  T.tab = nullptr; T.fin = nullptr; T.n = 0;
}

void launch_pcr_lines(const Launch& L, const Geo& G, const PcrTables& T, int mode, int ofs, int diag, const REAL* x,
                      REAL* dst, const REAL* msk, const REAL* rhs, REAL omg, int slot, bool add) {
  const int n = G.ked - G.kst + 1;
  const int ni = G.ied - G.ist + 1, nj = G.jed - G.jst + 1;
  if (n != T.n) { fprintf(stderr, "czb200: pcr tables built for n=%d, called with n=%d\n", T.n, n); abort(); }
  LineSel sel; sel.mode = mode; sel.ofs = ofs; sel.diag = diag; sel.jlo = 0;
  if (mode <= 1)      { sel.per_j = (ni + 1) / 2; sel.nlines = sel.per_j * nj; }
  else if (mode == 2) { sel.per_j = ni;           sel.nlines = ni * nj; }
  else {
    int jlo = diag - G.ied; if (jlo < G.jst) jlo = G.jst;
    int jhi = diag - G.ist; if (jhi > G.jed) jhi = G.jed;
    sel.jlo = jlo; sel.per_j = 1; sel.nlines = jhi >= jlo ? jhi - jlo + 1 : 0;
  }
  if (sel.nlines <= 0) return;
  int nt = 32; while (nt < n && nt < 256) nt <<= 1;
  int grid = sel.nlines < 148 * 16 ? sel.nlines : 148 * 16;
  if ((size_t)grid > L.rb_blocks) grid = (int)L.rb_blocks;
  const size_t smem = 2 * (size_t)n * sizeof(REAL);
  FinStoreP fs{L.acc_st, slot, add ? 1 : 0};
  if (T.last4) {
    CZB_CHECK(cudaFuncSetAttribute(k_pcr_lines<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_pcr_lines<1><<<grid, nt, smem, L.s>>>(G, n, T.pn, T.stages, T.tab, T.fin, sel, x, dst, msk, rhs, omg, L.st, L.rb, fs);
  } else {
    CZB_CHECK(cudaFuncSetAttribute(k_pcr_lines<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_pcr_lines<0><<<grid, nt, smem, L.s>>>(G, n, T.pn, T.stages, T.tab, T.fin, sel, x, dst, msk, rhs, omg, L.st, L.rb, fs);
  }
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

}  // namespace czb

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
#include <type_traits>

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


// ------------------------------------------------------------------------------------------------
// One WARP per line, the right-hand side d in registers.  Lane l owns the EPL consecutive cells
// k = EPL*l .. EPL*l+EPL-1 ("blocked" layout), so a PCR stage with stride s gets its neighbours d(k-s),
// d(k+s) from the lane's own registers (s < EPL, except the s cells next to a lane boundary) or with one
// warp shuffle from lane l -+ s/EPL (s >= EPL): no block-wide barrier, no shared-memory traffic for d.
// The coefficient tables (identical for every line, see above) are staged once per CTA in shared memory
// in a lane-interleaved layout that makes every lane's 128-bit read conflict free.  Loads and stores of
// the fields are done in the "cyclic" layout k = lane + 32*m (coalesced 128 B per instruction); the two
// layouts are exchanged through a small per-warp shared buffer.  Arithmetic and its order are those of
// k_pcr_lines (= cz_solver.f90:569-648), so results are bit-identical.
// four consecutive REALs of a lane: one 128-bit (float) or two 128-bit (double) shared-memory accesses
#ifdef _REAL_IS_DOUBLE_
struct alignas(16) R4 { double x, y, z, w; };
#else
typedef float4 R4;
#endif

template <int EPL>
struct PcrLay {
  static constexpr int NQ = EPL / 4;            // 16-byte chunks per lane
  static constexpr int QSTR = 32 * 4 + 8;       // floats between chunk q and q+1 of a row (padded: conflict-free both ways)
  static constexpr int ROW = NQ * QSTR;         // floats per interleaved row of 32*EPL cells
  __device__ static __forceinline__ int pos(int k) {   // position of cell k inside an interleaved row
    const int L = k / EPL, r = k - L * EPL;
    return (r >> 2) * QSTR + L * 4 + (r & 3);
  }
};

template <int EPL>
__global__ void __launch_bounds__(EPL == 32 ? 512 : 256, (EPL == 32 || sizeof(REAL) == 8) ? 1 : 2) k_pcr_warp(Geo G, int n, int pn, int stages, const REAL* __restrict__ tab,
                                                  const REAL* __restrict__ fin, LineSel sel, const REAL* x, REAL* dst,
                                                  const REAL* __restrict__ msk, const REAL* __restrict__ rhs, REAL omg,
                                                  const LoopState* st, RedBuf rb, FinStoreP finst) {
  if (st && st->done) return;
  typedef PcrLay<EPL> LY;
  constexpr int NQ = LY::NQ, ROW = LY::ROW, NP = 32 * EPL;
  extern __shared__ double sm_raw[];
  REAL* ts = reinterpret_cast<REAL*>(sm_raw);          // [stages][3][ROW] ap, cp, e ; then F1, F2 ; then per-warp buffers
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  REAL* F1 = ts + (size_t)stages * 3 * ROW;
  REAL* F2 = F1 + ROW;
  REAL* wb = F2 + ROW + (size_t)warp * ROW;            // this warp's layout-exchange buffer
  const REAL zero = (REAL)0.0;
  const int s2 = 1 << (pn - 1);
  // ---- stage the tables (zero beyond n: cells k >= n then stay exactly 0 through every stage) ----
  for (int idx = tid; idx < stages * 3 * NP; idx += blockDim.x) {
    const int row = idx / NP, k = idx - row * NP;
    ts[(size_t)row * ROW + LY::pos(k)] = (k < n) ? tab[(size_t)row * n + k] : zero;
  }
  for (int k = tid; k < NP; k += blockDim.x) {          // last 2x2 stage, cz_solver.f90:612-630, per own cell
    REAL f1 = zero, f2 = zero;
    if (k < n) {
      if (k < s2) { f1 = fin[k]; f2 = fin[2 * (size_t)s2 + k]; }
      else        { f1 = fin[(size_t)s2 + (k - s2)]; f2 = fin[2 * (size_t)s2 + (k - s2)]; }
    }
    F1[LY::pos(k)] = f1; F2[LY::pos(k)] = f2;
  }
  __syncthreads();
  const REAL r = (REAL)1.0 / (REAL)6.0;
  const int t2 = s2 / EPL;                              // lane distance of the 2x2 partner (>= 32: none)
  double acc[1] = {0.0};
  for (int line = blockIdx.x * nwarp + warp; line < sel.nlines; line += gridDim.x * nwarp) {
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
    if (i > G.ied) continue;      // warp-uniform
    const long long base = G.at(G.kst, i, j);
    // ---- source term in the cyclic layout (cz_solver.f90:569-579) -> exchange buffer ----
    // loads are issued CH cells at a time before anything is stored, so that they overlap (x and dst may
    // alias, the compiler must keep loads and stores in program order)
    constexpr int CH = EPL < 8 ? EPL : 8;
#pragma unroll
    for (int m0 = 0; m0 < EPL; m0 += CH) {
      REAL vs[CH], vn[CH], vw[CH], ve[CH], vb[CH], vm[CH];
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int k = lane + 32 * (m0 + c);
        const long long off = base + (k < n ? k : 0);
        vs[c] = x[off - G.plane]; vn[c] = x[off + G.plane]; vw[c] = x[off - G.nk]; ve[c] = x[off + G.nk];
        vb[c] = rhs[off]; vm[c] = msk ? msk[off] : (REAL)1.0;
      }
      const REAL xlo = x[base - 1], xhi = x[base + n];
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int k = lane + 32 * (m0 + c);
        REAL d = ((vs[c] + vn[c] + vw[c] + ve[c] - vb[c]) * r) * vm[c];
        if (k == 0) d = (d + xlo * r) * vm[c];
        if (k == n - 1) d = (d + xhi * r) * vm[c];
        wb[LY::pos(k)] = (k < n) ? d : zero;
      }
    }
    __syncwarp();
    REAL d[EPL];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      const R4 v = *reinterpret_cast<const R4*>(wb + q * LY::QSTR + lane * 4);
      d[4 * q] = v.x; d[4 * q + 1] = v.y; d[4 * q + 2] = v.z; d[4 * q + 3] = v.w;
    }
    // ---- PCR stages: d1(k) = e*(d(k) - ap*d(k-s) - cp*d(k+s))  (cz_solver.f90:602) ----
#pragma unroll
    for (int p = 0; p < 5 + (EPL == 4 ? 2 : EPL == 8 ? 3 : EPL == 16 ? 4 : 5); ++p) {
      if (p < stages) {
        const int s = 1 << p;
        const int t = s / EPL;                       // lane distance when s >= EPL
        const REAL* tp = ts + (size_t)p * 3 * ROW + lane * 4;
        REAL dn[EPL];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          const R4 ap4 = *reinterpret_cast<const R4*>(tp + q * LY::QSTR);
          const R4 cp4 = *reinterpret_cast<const R4*>(tp + ROW + q * LY::QSTR);
          const R4 e4 = *reinterpret_cast<const R4*>(tp + 2 * ROW + q * LY::QSTR);
          const REAL apv[4] = {ap4.x, ap4.y, ap4.z, ap4.w}, cpv[4] = {cp4.x, cp4.y, cp4.z, cp4.w};
          const REAL ev[4] = {e4.x, e4.y, e4.z, e4.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int m = 4 * q + e;
            REAL dl, dr;
            if (s < EPL) {
              // left neighbour k-s: own register m-s, or register m-s+EPL of the lane below; right alike
              if (m >= s) dl = d[m - s];
              else { const REAL v = __shfl_up_sync(0xffffffffu, d[(m - s + EPL) & (EPL - 1)], 1); dl = lane >= 1 ? v : zero; }
              if (m + s < EPL) dr = d[m + s];
              else { const REAL v = __shfl_down_sync(0xffffffffu, d[(m + s - EPL) & (EPL - 1)], 1); dr = lane < 31 ? v : zero; }
            } else {
              const REAL vl = __shfl_up_sync(0xffffffffu, d[m], t), vr = __shfl_down_sync(0xffffffffu, d[m], t);
              dl = lane >= t ? vl : zero;
              dr = lane + t < 32 ? vr : zero;
            }
            dn[m] = ev[e] * (d[m] - apv[e] * dl - cpv[e] * dr);
          }
        }
#pragma unroll
        for (int m = 0; m < EPL; ++m) d[m] = dn[m];
      }
    }
    // ---- last 2x2 stage (cz_solver.f90:612-630): d1 = (d_own - F1*d_partner)*F2, partner = k -+ s2 ----
    __syncwarp();
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      const R4 f1 = *reinterpret_cast<const R4*>(F1 + q * LY::QSTR + lane * 4);
      const R4 f2 = *reinterpret_cast<const R4*>(F2 + q * LY::QSTR + lane * 4);
      R4 o;
      REAL pr[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const REAL v = __shfl_xor_sync(0xffffffffu, d[4 * q + e], t2 & 31);
        pr[e] = (t2 < 32) ? v : zero;
      }
      o.x = (d[4 * q] - f1.x * pr[0]) * f2.x;
      o.y = (d[4 * q + 1] - f1.y * pr[1]) * f2.y;
      o.z = (d[4 * q + 2] - f1.z * pr[2]) * f2.z;
      o.w = (d[4 * q + 3] - f1.w * pr[3]) * f2.w;
      *reinterpret_cast<R4*>(wb + q * LY::QSTR + lane * 4) = o;
    }
    __syncwarp();
    // ---- relaxation in the cyclic layout (cz_solver.f90:642-648) ----
#pragma unroll
    for (int m0 = 0; m0 < EPL; m0 += CH) {
      REAL pp[CH], mk[CH];
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int k = lane + 32 * (m0 + c);
        const long long off = base + (k < n ? k : 0);
        pp[c] = x[off]; mk[c] = msk ? msk[off] : (REAL)1.0;
      }
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int k = lane + 32 * (m0 + c);
        if (k < n) {
          const REAL dp = (wb[LY::pos(k)] - pp[c]) * omg * mk[c];
          dst[base + k] = pp[c] + dp;
          acc[0] += (double)(dp * dp);
        }
      }
    }
    __syncwarp();
  }
  grid_reduce<1>(acc, rb, finst);
}


#ifndef _REAL_IS_DOUBLE_
// ------------------------------------------------------------------------------------------------
// Building blocks of the solve-loop kernel k_pcr_warp_tm below (msk == 1 on all line cells, so it is not read):
// packed PCR stage with shared-memory tables, source-term loads of the NEXT line issued during the stages
// of the current one.  Arithmetic identical to k_pcr_warp.
// Packed variant: two neighbouring cells of a lane travel as one 64-bit register pair and are reduced by
// Blackwell's FADD2 / FMUL2 / FFMA2.  ptxas contracts mul.rn.f32x2 + sub.rn.f32x2 into one FFMA2 (one rounding
// instead of two) even with --fmad false, so a product that feeds a subtraction is written fma(a, b, -0) with
// the -0 pair coming from a kernel argument: a*b + (-0) rounds exactly like a*b and cannot be contracted.
typedef unsigned long long pk_t;
__device__ __forceinline__ pk_t pk2(float x, float y) { pk_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y)); return r; }
__device__ __forceinline__ float lo2(pk_t v) { return __uint_as_float((unsigned)v); }
__device__ __forceinline__ float hi2(pk_t v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ pk_t add2(pk_t a, pk_t b) { pk_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ pk_t sub2(pk_t a, pk_t b) { pk_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ pk_t mul2(pk_t a, pk_t b) { pk_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ pk_t fma2(pk_t a, pk_t b, pk_t c) { pk_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

template <int EPL>
__device__ __forceinline__ void pcr_stage_pk(pk_t (&D)[EPL / 2], const int p, const REAL* ts, const int lane, const pk_t nz) {
  typedef PcrLay<EPL> LY;
  constexpr int NQ = LY::NQ, ROW = LY::ROW, NH = EPL / 2;
  const int s = 1 << p;
  const int t = s / EPL;
  const REAL* tp = ts + (size_t)p * 3 * ROW + lane * 4;
  pk_t DN[NH];
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const ulonglong2 ap4 = *reinterpret_cast<const ulonglong2*>(tp + q * LY::QSTR);
    const ulonglong2 cp4 = *reinterpret_cast<const ulonglong2*>(tp + ROW + q * LY::QSTR);
    const ulonglong2 e4 = *reinterpret_cast<const ulonglong2*>(tp + 2 * ROW + q * LY::QSTR);
    const pk_t apv[2] = {ap4.x, ap4.y}, cpv[2] = {cp4.x, cp4.y}, ev[2] = {e4.x, e4.y};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int j = 2 * q + h;                   // pair j = cells 2j, 2j+1 of the lane
      pk_t dl, dr;
      if (s == 1) {                              // neighbours straddle the pairs
        const float l0 = (j >= 1) ? hi2(D[j - 1]) : __shfl_up_sync(0xffffffffu, hi2(D[NH - 1]), 1);
        const float r1 = (j + 1 < NH) ? lo2(D[j + 1]) : __shfl_down_sync(0xffffffffu, lo2(D[0]), 1);
        dl = pk2(l0, lo2(D[j]));
        dr = pk2(hi2(D[j]), r1);
      } else if (s < EPL) {
        const int hs = s / 2;
        if (j >= hs) dl = D[j - hs];
        else dl = __shfl_up_sync(0xffffffffu, D[(j - hs + NH) & (NH - 1)], 1);
        if (j + hs < NH) dr = D[j + hs];
        else dr = __shfl_down_sync(0xffffffffu, D[(j + hs - NH) & (NH - 1)], 1);
      } else {
        dl = __shfl_up_sync(0xffffffffu, D[j], t);
        dr = __shfl_down_sync(0xffffffffu, D[j], t);
      }
      DN[j] = mul2(ev[h], sub2(sub2(D[j], fma2(apv[h], dl, nz)), fma2(cpv[h], dr, nz)));
    }
  }
#pragma unroll
  for (int j = 0; j < NH; ++j) D[j] = DN[j];
}

// source term of cells c0 .. c0+NC-1 (cyclic layout k = lane + 32 c) of a line: pcr_src_load issues the
// loads, pcr_src_put turns them into d (msk == 1) in the exchange buffer.  PcrSrc holds the addresses of
// the lane's cell k = lane in the four neighbour lines and in the source array, so every load is
// pointer + immediate offset; xl / xh point at the two Dirichlet-side neighbours x(kst-1), x(ked+1).
struct PcrSrc { const REAL *ps, *pn, *pw, *pe, *pb, *xl; };
template <int NC>
__device__ __forceinline__ void pcr_src_load(REAL (&v)[6][NC], const PcrSrc& P, const int c0, const int n, const int lane) {
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int cc = c0 + c;
    const bool ok = lane < n - 32 * cc;          // k = lane + 32 cc < n
    v[0][c] = v[1][c] = v[2][c] = v[3][c] = v[4][c] = v[5][c] = (REAL)0.0;
    if (ok) {
      v[0][c] = P.ps[32 * cc]; v[1][c] = P.pn[32 * cc]; v[2][c] = P.pw[32 * cc]; v[3][c] = P.pe[32 * cc];
      v[4][c] = P.pb[32 * cc];
    }
    if (cc == 0 && lane == 0) v[5][c] = P.xl[0];
    if (lane == n - 1 - 32 * cc) v[5][c] = P.xl[n + 1];
  }
}
template <int EPL, int NC>
__device__ __forceinline__ void pcr_src_put(const REAL (&v)[6][NC], REAL* wbl, const int c0, const int n, const int lane, const REAL r) {
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int cc = c0 + c;
    REAL d = (v[0][c] + v[1][c] + v[2][c] + v[3][c] - v[4][c]) * r;
    if ((cc == 0 && lane == 0) || lane == n - 1 - 32 * cc) d = d + v[5][c] * r;
    wbl[(128 / EPL) * cc] = d;                   // wbl = wb + pos(lane); pos(lane + 32 cc) = pos(lane) + (32/EPL)*4*cc
  }
}

// ------------------------------------------------------------------------------------------------
// Coefficient tables in TENSOR MEMORY.  With the tables in shared memory the LSU pipe is the limiter (ncu:
// l1tex 90 % busy, 2/3 of its wavefronts are the table reads: 3 values per cell and stage).  The tables
// are per-lane private data ("lane l needs the coefficients of its own EPL cells"), which is exactly the
// shape of Blackwell's tensor memory: 128 lanes x 512 columns x 32 bit per SM, read with tcgen05.ld
// (32x32b: thread i of a warp reads TMEM lane 32*(warp%4)+i, N consecutive columns) at many times the
// shared-memory bandwidth and off the LSU pipe.  Row R of the table set (3 per stage: ap, cp, e; then
// the two rows of the last 2x2 stage) occupies columns R*EPL .. R*EPL+EPL-1; the four lane quarters
// hold identical copies, written once per CTA with tcgen05.st.  No MMA is involved.  EPL = 32 (lines of
// 513..1024 cells) has 29 rows x 32 columns > 512: its stages >= 5 and the 2x2 rows stay in shared memory.
struct Tm24 { uint32_t a[8], c[8], e[8]; };
__device__ __forceinline__ void tm_ld8(uint32_t (&r)[8], const uint32_t addr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(addr));
}
template <int EPL>
__device__ __forceinline__ void tm_issue(Tm24& t, const uint32_t tl, const int p, const int h) {
  const uint32_t col = tl + (uint32_t)(3 * p * EPL + 8 * h);
  tm_ld8(t.a, col); tm_ld8(t.c, col + EPL); tm_ld8(t.e, col + 2 * EPL);
}
// the registers of t are valid only after this wait; passing them through makes the dependence explicit
__device__ __forceinline__ void tm_wait(Tm24& t) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(t.a[0]), "+r"(t.a[1]), "+r"(t.a[2]), "+r"(t.a[3]), "+r"(t.a[4]), "+r"(t.a[5]), "+r"(t.a[6]), "+r"(t.a[7]),
                 "+r"(t.c[0]), "+r"(t.c[1]), "+r"(t.c[2]), "+r"(t.c[3]), "+r"(t.c[4]), "+r"(t.c[5]), "+r"(t.c[6]), "+r"(t.c[7]),
                 "+r"(t.e[0]), "+r"(t.e[1]), "+r"(t.e[2]), "+r"(t.e[3]), "+r"(t.e[4]), "+r"(t.e[5]), "+r"(t.e[6]), "+r"(t.e[7])
               :: "memory");
}
__device__ __forceinline__ pk_t pku(uint32_t lo, uint32_t hi) { pk_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi)); return r; }

template <int EPL, int WPB, bool BULK, bool L4>
__global__ void __launch_bounds__(WPB * 32, 1) k_pcr_warp_tm(Geo G, int n, int pn, int stages, const REAL* __restrict__ tab,
                                                  const REAL* __restrict__ fin, LineSel sel, const REAL* x, REAL* dst,
                                                  const REAL* __restrict__ rhs, REAL omg, pk_t nz,
                                                  const LoopState* st, RedBuf rb, FinStoreP finst) {
  if (st && st->done) return;
  typedef PcrLay<EPL> LY;
  constexpr int NQ = LY::NQ, ROW = LY::ROW, NP = 32 * EPL, NH = EPL / 2, GH = EPL / 8;
  constexpr int CPS = EPL == 32 ? 4 : 2;
  constexpr int MAXST = 5 + (EPL == 8 ? 3 : EPL == 16 ? 4 : 5);
  constexpr int TP = EPL == 32 ? 5 : MAXST;              // stages whose tables live in tensor memory
  constexpr bool FTM = EPL != 32;                        // 2x2 rows in tensor memory
  static_assert(EPL >= 8 && WPB % 4 == 0, "tensor-memory variant: EPL 8, 16, 32");
  static_assert(!L4 || (BULK && EPL != 32), "4x4 last stage: lines of <= 512 cells with bulk staging");
  extern __shared__ __align__(128) unsigned char sm_tm[];   // TMA bulk destinations need 16-byte alignment
  __shared__ uint32_t tm_base_s;
  __shared__ unsigned long long mbar_s[WPB];
  REAL* sm = reinterpret_cast<REAL*>(sm_tm);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const REAL zero = (REAL)0.0;
  const int s2 = 1 << (pn - 1);
  // L4 (pcr_rb_esa & co, cz_solver.f90:1391-1438): pn-2 stages, then 4x4 systems by Cramer's rule; their 7
  // coefficient rows (cc1 cc2 cc3 aa2 aa3 aa4 1/det) occupy 32 TMEM columns after the stage tables
  const int rows = L4 ? 3 * stages : 3 * stages + 2;
  const int rows_tm = FTM ? rows : (stages < TP ? 3 * stages : 3 * TP);
  const int rows_sm = rows - rows_tm;                    // shared-memory rows: rows_tm .. rows-1
  REAL* ts_s = sm - (size_t)rows_tm * ROW;               // so that row R sits at ts_s + R*ROW
  // per warp: BULK: five raw neighbour / source rows (TMA bulk copies) + one exchange buffer; else two exchange buffers
  const int rawf = BULK ? 5 * G.nk : ROW;
  REAL* wb0 = sm + (size_t)rows_sm * ROW + (size_t)warp * (rawf + ROW) + (BULK ? rawf : 0);
  REAL* raw = wb0 - rawf;
  const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&mbar_s[warp]);
  if (BULK && lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mbar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  auto table = [&](int R, int k) -> REAL {               // entry of row R for cell k (0 beyond the line)
    if (k >= n) return zero;
    if (R < 3 * stages) return tab[(size_t)R * n + k];
    if (R == 3 * stages) return (k < s2) ? fin[k] : fin[(size_t)s2 + (k - s2)];
    return (k < s2) ? fin[2 * (size_t)s2 + k] : fin[2 * (size_t)s2 + (k - s2)];
  };
  // ---- tensor memory: allocate all 512 columns (one CTA per SM), fill the four lane quarters ----
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"((uint32_t)__cvta_generic_to_shared(&tm_base_s)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int idx = tid; idx < rows_sm * NP; idx += WPB * 32) {
    const int row = idx / NP, k = idx - row * NP;
    sm[(size_t)row * ROW + LY::pos(k)] = table(rows_tm + row, k);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // this warp's lane quarter (the shuffle tells the compiler that the address is warp-uniform)
  const uint32_t tl = __shfl_sync(0xffffffffu, tm_base_s + ((uint32_t)((warp & 3) * 32) << 16), 0);
  {   // every warp writes part of the rows of its own lane quarter
#pragma unroll 2
    for (int R = warp >> 2; R < rows_tm; R += WPB / 4) {
      uint32_t v[EPL];
#pragma unroll
      for (int m = 0; m < EPL; ++m) v[m] = __float_as_uint(table(R, EPL * lane + m));
#pragma unroll
      for (int m0 = 0; m0 < EPL; m0 += 4)
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
                     :: "r"(tl + (uint32_t)(R * EPL + m0)), "r"(v[m0]), "r"(v[m0 + 1]), "r"(v[m0 + 2]), "r"(v[m0 + 3]) : "memory");
    }
    if (L4 && warp < 4) {
      // lane (L = lane & 7, g = lane >> 3) solves the systems k = EPL*L + SPL*g + jj, jj < SPL = EPL/4 (k < s = 8*EPL):
      // column SPL*r + jj of the block holds fin[r][k]
      const int s4 = 1 << (pn - 2);
      constexpr int SPL = EPL / 4;
      const int k0 = EPL * (lane & 7) + SPL * (lane >> 3);
      uint32_t v[32];
#pragma unroll
      for (int c = 0; c < 32; ++c) v[c] = (c < 7 * SPL) ? __float_as_uint(fin[(size_t)(c / SPL) * s4 + k0 + (c % SPL)]) : 0u;
#pragma unroll
      for (int c = 0; c < 8 * SPL; c += 4)
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
                     :: "r"(tl + (uint32_t)(3 * stages * EPL + c)), "r"(v[c]), "r"(v[c + 1]), "r"(v[c + 2]), "r"(v[c + 3]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  const REAL r = (REAL)1.0 / (REAL)6.0;
  const int t2 = s2 / EPL;
  double acc[1] = {0.0};
  auto line_base = [&](int line) -> long long {
    if (line >= sel.nlines) return -1;
    int i, j;
    if (sel.mode <= 1) {
      const int jj = line / sel.per_j, m = line - jj * sel.per_j;
      j = G.jst + jj;
      i = G.ist + ((G.ist + j + sel.ofs + sel.mode) & 1) + 2 * m;
    } else if (sel.mode == 2) {
      const int jj = line / sel.per_j, m = line - jj * sel.per_j;
      j = G.jst + jj; i = G.ist + m;
    } else {
      j = sel.jlo + line; i = sel.diag - j;
    }
    return (i > G.ied) ? -1 : G.at(G.kst, i, j);
  };
  const int posl = LY::pos(lane);
  auto src_of = [&](long long b) -> PcrSrc {
    const REAL* px = x + b + lane;
    return PcrSrc{px - G.plane, px + G.plane, px - G.nk, px + G.nk, rhs + b + lane, x + b - 1};
  };
  auto full_load = [&](long long b, REAL* wb) {
    constexpr int CH = 8;
    const PcrSrc P = src_of(b);
#pragma unroll
    for (int m0 = 0; m0 < EPL; m0 += CH) {
      REAL v[6][CH];
      pcr_src_load<CH>(v, P, m0, n, lane);
      pcr_src_put<EPL, CH>(v, wb + posl, m0, n, lane, r);
    }
  };
  const int lstride = gridDim.x * WPB;
  if constexpr (BULK) {
    // ---- rows of the next line arrive by TMA bulk copies while the current line is reduced ----
    const int koff = G.kst + G.g - 1;                     // first line cell inside a row of nk
    const uint32_t rowb = (uint32_t)G.nk * (uint32_t)sizeof(REAL);
    const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(raw);
    auto next_valid = [&](int l, long long& b) -> int {   // first existing line >= l (stride lstride), or nlines
      while (l < sel.nlines) { b = line_base(l); if (b >= 0) return l; l += lstride; }
      b = -1; return sel.nlines;
    };
    REAL elo = zero, ehi = zero;                          // Dirichlet-side neighbours x(kst-1), x(ked+1) of the staged line
    auto issue = [&](long long b) {
      if (lane == 0) {
        const REAL* px = x + b - koff;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mbar), "r"(5u * rowb) : "memory");
        const REAL* src[5] = {px - G.plane, px + G.plane, px - G.nk, px + G.nk, rhs + b - koff};
#pragma unroll
        for (int q = 0; q < 5; ++q)
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       :: "r"(raw_s + (uint32_t)q * rowb), "l"(src[q]), "r"(rowb), "r"(mbar) : "memory");
      }
      if (lane == 0) elo = x[b - 1];
      if (lane == ((n - 1) & 31)) ehi = x[b + n];
    };
    __syncwarp();
    long long base;
    int line = next_valid(blockIdx.x * WPB + warp, base);
    if (line < sel.nlines) issue(base);
    uint32_t ph = 0;
    while (line < sel.nlines) {
      long long nbase;
      const int nline = next_valid(line + lstride, nbase);
      REAL* wc = wb0;
      // ---- source term (cz_solver.f90:569-579) from the staged rows -> exchange buffer ----
      {
        uint32_t ok = 0;
        while (!ok) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                 : "=r"(ok) : "r"(mbar), "r"(ph) : "memory");
        ph ^= 1u;
        const REAL* rk = raw + koff + lane;
#pragma unroll
        for (int c = 0; c < EPL; ++c) {
          const bool in = lane < n - 32 * c;
          REAL d = zero;
          if (in) d = (rk[32 * c] + rk[G.nk + 32 * c] + rk[2 * G.nk + 32 * c] + rk[3 * G.nk + 32 * c] - rk[4 * G.nk + 32 * c]) * r;
          if (c == 0 && lane == 0) d = d + elo * r;
          if (lane == n - 1 - 32 * c) d = d + ehi * r;
          wc[posl + (128 / EPL) * c] = d;
        }
      }
      __syncwarp();
      if (nline < sel.nlines) issue(nbase);
      pk_t D[NH];
      Tm24 T[2];
      tm_issue<EPL>(T[0], tl, 0, 0);
      const REAL* px = x + base + lane;
      REAL* pd = dst + base + lane;
      REAL pp[EPL];
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(wc + q * LY::QSTR + lane * 4);
        D[2 * q] = v.x; D[2 * q + 1] = v.y;
      }
#pragma unroll
      for (int p = 0; p < MAXST; ++p) {
        if (p < stages) {
          if (p == stages - 1) {   // own-line values for the relaxation: in flight during the last stage and the 2x2 stage
#pragma unroll
            for (int c = 0; c < EPL; ++c) pp[c] = (lane < n - 32 * c) ? px[32 * c] : zero;
          }
          if (p < TP) {
            const int s = 1 << p;
            const int t = s / EPL;
            pk_t DN[NH];
#pragma unroll
            for (int h = 0; h < GH; ++h) {
              const int g = p * GH + h;
              Tm24& cur = T[g & 1];
              tm_wait(cur);
              const int p1 = (h + 1 < GH) ? p : p + 1, h1 = (h + 1 < GH) ? h + 1 : 0;
              if (p1 < TP && p1 < stages) tm_issue<EPL>(T[(g + 1) & 1], tl, p1, h1);
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) {
                const int j = 4 * h + jj;
                pk_t dl, dr;
                if (s == 1) {
                  const float l0 = (j >= 1) ? hi2(D[j - 1]) : __shfl_up_sync(0xffffffffu, hi2(D[NH - 1]), 1);
                  const float r1 = (j + 1 < NH) ? lo2(D[j + 1]) : __shfl_down_sync(0xffffffffu, lo2(D[0]), 1);
                  dl = pk2(l0, lo2(D[j]));
                  dr = pk2(hi2(D[j]), r1);
                } else if (s < EPL) {
                  const int hs = s / 2;
                  if (j >= hs) dl = D[j - hs];
                  else dl = __shfl_up_sync(0xffffffffu, D[(j - hs + NH) & (NH - 1)], 1);
                  if (j + hs < NH) dr = D[j + hs];
                  else dr = __shfl_down_sync(0xffffffffu, D[(j + hs - NH) & (NH - 1)], 1);
                } else {
                  dl = __shfl_up_sync(0xffffffffu, D[j], t);
                  dr = __shfl_down_sync(0xffffffffu, D[j], t);
                }
                const pk_t ap = pku(cur.a[2 * jj], cur.a[2 * jj + 1]), cp = pku(cur.c[2 * jj], cur.c[2 * jj + 1]);
                const pk_t ee = pku(cur.e[2 * jj], cur.e[2 * jj + 1]);
                DN[j] = mul2(ee, sub2(sub2(D[j], fma2(ap, dl, nz)), fma2(cp, dr, nz)));
              }
            }
#pragma unroll
            for (int j = 0; j < NH; ++j) D[j] = DN[j];
          } else {
            pcr_stage_pk<EPL>(D, p, ts_s, lane, nz);
          }
        }
      }
      // ---- last 2x2 stage (cz_solver.f90:612-630) ----
      if constexpr (L4) {
        // ---- 4x4 systems {k, k+s, k+2s, k+3s}, s = 8*EPL: their members sit in lanes L, L+8, L+16, L+24 at the
        // same register.  Through the exchange buffer every lane takes SPL = EPL/4 complete systems (the cells
        // SPL*g .. SPL*g+SPL-1 of the lanes L+8j; L = lane & 7, g = lane >> 3), solves them as packed pairs with
        // the reference's expression order and writes the results back in place. ----
        constexpr int SPL = EPL / 4, NPR = SPL / 2;
        const uint32_t fcol = tl + (uint32_t)(3 * stages * EPL);
        tm_ld8(T[0].a, fcol); tm_ld8(T[0].c, fcol + 8);
        if (SPL == 4) { tm_ld8(T[0].e, fcol + 16); tm_ld8(T[1].a, fcol + 24); }
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          ulonglong2 o; o.x = D[2 * q]; o.y = D[2 * q + 1];
          *reinterpret_cast<ulonglong2*>(wc + q * LY::QSTR + lane * 4) = o;
        }
        __syncwarp();
        const int m0 = SPL * (lane >> 3);
        REAL* sys = wc + (m0 >> 2) * LY::QSTR + (lane & 7) * 4 + (m0 & 3);   // member j at + 32*j floats
        pk_t dd[4][NPR];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int h = 0; h < NPR; ++h) dd[j][h] = *reinterpret_cast<const pk_t*>(sys + 32 * j + 2 * h);
        if (SPL != 4) {
#pragma unroll
          for (int e = 0; e < 8; ++e) T[0].e[e] = 0u;
        }
        tm_wait(T[0]);
        if (SPL == 4) {
#pragma unroll
          for (int e = 0; e < 8; ++e) { T[1].c[e] = 0u; T[1].e[e] = 0u; }
          tm_wait(T[1]);
        }
        auto F = [&](int c) -> uint32_t {   // column c of the coefficient block (c is a constant after unrolling)
          return c < 8 ? T[0].a[c & 7] : c < 16 ? T[0].c[c & 7] : c < 24 ? T[0].e[c & 7] : T[1].a[c & 7];
        };
        auto P = [&](pk_t u, pk_t v) -> pk_t { return fma2(u, v, nz); };   // product that may feed an add / sub
        auto neg = [&](pk_t u) -> pk_t { return u ^ 0x8000000080000000ull; };
#pragma unroll
        for (int h = 0; h < NPR; ++h) {
          const pk_t cc1 = pku(F(0 * SPL + 2 * h), F(0 * SPL + 2 * h + 1)), cc2 = pku(F(1 * SPL + 2 * h), F(1 * SPL + 2 * h + 1));
          const pk_t cc3 = pku(F(2 * SPL + 2 * h), F(2 * SPL + 2 * h + 1)), aa2 = pku(F(3 * SPL + 2 * h), F(3 * SPL + 2 * h + 1));
          const pk_t aa3 = pku(F(4 * SPL + 2 * h), F(4 * SPL + 2 * h + 1)), aa4 = pku(F(5 * SPL + 2 * h), F(5 * SPL + 2 * h + 1));
          const pk_t inv = pku(F(6 * SPL + 2 * h), F(6 * SPL + 2 * h + 1));
          const pk_t dd1 = dd[0][h], dd2 = dd[1][h], dd3 = dd[2][h], dd4 = dd[3][h];
          // cz_solver.f90:1391-1438, evaluated left to right as k_pcr_lines<1> does
          pk_t in1 = sub2(add2(P(aa4, dd1), P(P(cc1, cc2), dd4)), P(P(aa4, cc1), dd2));
          pk_t d1 = add2(P(neg(cc3), in1), dd1);
          d1 = add2(d1, P(P(cc1, cc2), dd3)); d1 = sub2(d1, P(P(aa3, cc2), dd1)); d1 = sub2(d1, P(cc1, dd2));
          pk_t d2 = add2(dd2, P(P(cc2, cc3), dd4));
          d2 = sub2(d2, P(P(aa4, cc3), dd2)); d2 = sub2(d2, P(cc2, dd3));
          d2 = sub2(d2, P(aa2, sub2(dd1, P(P(aa4, cc3), dd1))));
          pk_t d3 = sub2(dd3, P(cc3, dd4)); d3 = sub2(d3, P(aa3, dd2));
          d3 = sub2(d3, P(aa2, sub2(sub2(P(cc1, dd3), P(P(cc1, cc3), dd4)), P(aa3, dd1))));
          pk_t d4 = add2(dd4, P(P(aa3, aa4), dd2)); d4 = sub2(d4, P(aa4, dd3)); d4 = sub2(d4, P(P(aa3, cc2), dd4));
          d4 = sub2(d4, P(aa2, sub2(add2(P(cc1, dd4), P(P(aa3, aa4), dd1)), P(P(aa4, cc1), dd3))));
          *reinterpret_cast<pk_t*>(sys + 2 * h) = mul2(d1, inv);
          *reinterpret_cast<pk_t*>(sys + 32 + 2 * h) = mul2(d2, inv);
          *reinterpret_cast<pk_t*>(sys + 64 + 2 * h) = mul2(d3, inv);
          *reinterpret_cast<pk_t*>(sys + 96 + 2 * h) = mul2(d4, inv);
        }
      } else
      if (FTM) {
        const uint32_t fcol = tl + (uint32_t)(3 * stages * EPL);
        tm_ld8(T[0].a, fcol); tm_ld8(T[0].c, fcol + EPL);
#pragma unroll
        for (int h = 0; h < GH; ++h) {
          Tm24& F = T[h & 1];
          tm_wait(F);
          if (h + 1 < GH) { tm_ld8(T[(h + 1) & 1].a, fcol + 8 * (h + 1)); tm_ld8(T[(h + 1) & 1].c, fcol + EPL + 8 * (h + 1)); }
#pragma unroll
          for (int jj = 0; jj < 4; jj += 2) {
            ulonglong2 o;
            const pk_t pr0 = __shfl_xor_sync(0xffffffffu, D[4 * h + jj], 16);       // partner k -+ s2: lane distance s2/EPL = 16
            const pk_t pr1 = __shfl_xor_sync(0xffffffffu, D[4 * h + jj + 1], 16);
            o.x = mul2(sub2(D[4 * h + jj], fma2(pku(F.a[2 * jj], F.a[2 * jj + 1]), pr0, nz)), pku(F.c[2 * jj], F.c[2 * jj + 1]));
            o.y = mul2(sub2(D[4 * h + jj + 1], fma2(pku(F.a[2 * jj + 2], F.a[2 * jj + 3]), pr1, nz)), pku(F.c[2 * jj + 2], F.c[2 * jj + 3]));
            *reinterpret_cast<ulonglong2*>(wc + (2 * h + jj / 2) * LY::QSTR + lane * 4) = o;
          }
        }
      } else {
        const REAL* F1 = ts_s + (size_t)(3 * stages) * ROW;
        const REAL* F2 = F1 + ROW;
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          const ulonglong2 f1 = *reinterpret_cast<const ulonglong2*>(F1 + q * LY::QSTR + lane * 4);
          const ulonglong2 f2 = *reinterpret_cast<const ulonglong2*>(F2 + q * LY::QSTR + lane * 4);
          pk_t prt[2];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const pk_t v = __shfl_xor_sync(0xffffffffu, D[2 * q + h], t2 & 31);
            prt[h] = (t2 < 32) ? v : 0ull;
          }
          ulonglong2 o;
          o.x = mul2(sub2(D[2 * q], fma2(f1.x, prt[0], nz)), f2.x);
          o.y = mul2(sub2(D[2 * q + 1], fma2(f1.y, prt[1], nz)), f2.y);
          *reinterpret_cast<ulonglong2*>(wc + q * LY::QSTR + lane * 4) = o;
        }
      }
      __syncwarp();
      // ---- relaxation in the cyclic layout (cz_solver.f90:642-648), msk == 1; cells k >= n hold 0 - 0 ----
#pragma unroll
      for (int c = 0; c < EPL; ++c) {
        const REAL dp = (wc[posl + (128 / EPL) * c] - pp[c]) * omg;
        if (lane < n - 32 * c) pd[32 * c] = pp[c] + dp;
        acc[0] += (double)(dp * dp);
      }
      __syncwarp();
      line = nline; base = nbase;
    }
  } else {
  int line = blockIdx.x * WPB + warp;
  long long base = line_base(line);
  int buf = 0;
  if (base >= 0) full_load(base, wb0);
  while (line < sel.nlines) {
    const int nline = line + lstride;
    const long long nbase = line_base(nline);
    REAL* wc = wb0 + buf * ROW;
    REAL* wn = wb0 + (buf ^ 1) * ROW;
    __syncwarp();
    if (base >= 0) {
      const PcrSrc P = src_of(nbase >= 0 ? nbase : base);
      pk_t D[NH];
      Tm24 T[2];
      tm_issue<EPL>(T[0], tl, 0, 0);
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(wc + q * LY::QSTR + lane * 4);
        D[2 * q] = v.x; D[2 * q + 1] = v.y;
      }
#pragma unroll
      for (int p = 0; p < MAXST; ++p) {
        if (p < stages) {
          REAL v[6][CPS];
          const bool pf = (nbase >= 0) && (p * CPS < EPL);
          if (pf) pcr_src_load<CPS>(v, P, p * CPS, n, lane);
          if (p < TP) {
            const int s = 1 << p;
            const int t = s / EPL;
            pk_t DN[NH];
#pragma unroll
            for (int h = 0; h < GH; ++h) {
              const int g = p * GH + h;
              Tm24& cur = T[g & 1];
              tm_wait(cur);
              const int p1 = (h + 1 < GH) ? p : p + 1, h1 = (h + 1 < GH) ? h + 1 : 0;
              if (p1 < TP && p1 < stages) tm_issue<EPL>(T[(g + 1) & 1], tl, p1, h1);
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) {
                const int j = 4 * h + jj;
                pk_t dl, dr;
                if (s == 1) {
                  const float l0 = (j >= 1) ? hi2(D[j - 1]) : __shfl_up_sync(0xffffffffu, hi2(D[NH - 1]), 1);
                  const float r1 = (j + 1 < NH) ? lo2(D[j + 1]) : __shfl_down_sync(0xffffffffu, lo2(D[0]), 1);
                  dl = pk2(l0, lo2(D[j]));
                  dr = pk2(hi2(D[j]), r1);
                } else if (s < EPL) {
                  const int hs = s / 2;
                  if (j >= hs) dl = D[j - hs];
                  else dl = __shfl_up_sync(0xffffffffu, D[(j - hs + NH) & (NH - 1)], 1);
                  if (j + hs < NH) dr = D[j + hs];
                  else dr = __shfl_down_sync(0xffffffffu, D[(j + hs - NH) & (NH - 1)], 1);
                } else {
                  dl = __shfl_up_sync(0xffffffffu, D[j], t);
                  dr = __shfl_down_sync(0xffffffffu, D[j], t);
                }
                const pk_t ap = pku(cur.a[2 * jj], cur.a[2 * jj + 1]), cp = pku(cur.c[2 * jj], cur.c[2 * jj + 1]);
                const pk_t ee = pku(cur.e[2 * jj], cur.e[2 * jj + 1]);
                DN[j] = mul2(ee, sub2(sub2(D[j], fma2(ap, dl, nz)), fma2(cp, dr, nz)));
              }
            }
#pragma unroll
            for (int j = 0; j < NH; ++j) D[j] = DN[j];
          } else {
            pcr_stage_pk<EPL>(D, p, ts_s, lane, nz);
          }
          if (pf) pcr_src_put<EPL, CPS>(v, wn + posl, p * CPS, n, lane, r);
        }
      }
      // own-line values for the relaxation, in flight during the 2x2 stage
      const REAL* px = x + base + lane;
      REAL* pd = dst + base + lane;
      REAL pp[EPL];
#pragma unroll
      for (int c = 0; c < EPL; ++c) pp[c] = (lane < n - 32 * c) ? px[32 * c] : zero;
      // ---- last 2x2 stage (cz_solver.f90:612-630) ----
      if (FTM) {
        const uint32_t fcol = tl + (uint32_t)(3 * stages * EPL);
#pragma unroll
        for (int h = 0; h < GH; ++h) {
          Tm24 F;
          tm_ld8(F.a, fcol + 8 * h); tm_ld8(F.c, fcol + EPL + 8 * h);
#pragma unroll
          for (int e = 0; e < 8; ++e) F.e[e] = 0u;
          tm_wait(F);
#pragma unroll
          for (int jj = 0; jj < 4; jj += 2) {
            ulonglong2 o;
            pk_t prt[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const pk_t v = __shfl_xor_sync(0xffffffffu, D[4 * h + jj + u], t2 & 31);
              prt[u] = (t2 < 32) ? v : 0ull;
            }
            o.x = mul2(sub2(D[4 * h + jj], fma2(pku(F.a[2 * jj], F.a[2 * jj + 1]), prt[0], nz)), pku(F.c[2 * jj], F.c[2 * jj + 1]));
            o.y = mul2(sub2(D[4 * h + jj + 1], fma2(pku(F.a[2 * jj + 2], F.a[2 * jj + 3]), prt[1], nz)), pku(F.c[2 * jj + 2], F.c[2 * jj + 3]));
            *reinterpret_cast<ulonglong2*>(wc + (2 * h + jj / 2) * LY::QSTR + lane * 4) = o;
          }
        }
      } else {
        const REAL* F1 = ts_s + (size_t)(3 * stages) * ROW;
        const REAL* F2 = F1 + ROW;
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          const ulonglong2 f1 = *reinterpret_cast<const ulonglong2*>(F1 + q * LY::QSTR + lane * 4);
          const ulonglong2 f2 = *reinterpret_cast<const ulonglong2*>(F2 + q * LY::QSTR + lane * 4);
          pk_t prt[2];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const pk_t v = __shfl_xor_sync(0xffffffffu, D[2 * q + h], t2 & 31);
            prt[h] = (t2 < 32) ? v : 0ull;
          }
          ulonglong2 o;
          o.x = mul2(sub2(D[2 * q], fma2(f1.x, prt[0], nz)), f2.x);
          o.y = mul2(sub2(D[2 * q + 1], fma2(f1.y, prt[1], nz)), f2.y);
          *reinterpret_cast<ulonglong2*>(wc + q * LY::QSTR + lane * 4) = o;
        }
      }
      __syncwarp();
      // ---- relaxation in the cyclic layout (cz_solver.f90:642-648), msk == 1; cells k >= n hold 0 - 0 ----
#pragma unroll
      for (int c = 0; c < EPL; ++c) {
        const REAL dp = (wc[posl + (128 / EPL) * c] - pp[c]) * omg;
        if (lane < n - 32 * c) pd[32 * c] = pp[c] + dp;
        acc[0] += (double)(dp * dp);
      }
    } else if (nbase >= 0) {
      full_load(nbase, wn);
    }
    line = nline; base = nbase; buf ^= 1;
  }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tm_base_s), "r"(512u) : "memory");
  }
  grid_reduce<1>(acc, rb, finst);
}
#endif  // !_REAL_IS_DOUBLE_

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
                      REAL* dst, const REAL* msk, const REAL* rhs, REAL omg, int slot, bool add, bool mask_is_inner) {
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
#ifndef _REAL_IS_DOUBLE_
  static const bool no_warp = getenv("CZB_PCR_BLOCK") != nullptr;
  static const bool no_tm = getenv("CZB_PCR_NOTMEM") != nullptr;
  static const bool no_bulk = getenv("CZB_PCR_NOBULK") != nullptr;
  const int epl = n <= 128 ? 4 : n <= 256 ? 8 : n <= 512 ? 16 : 32;      // cells per lane of the warp kernels
  // tensor-memory kernel: solve loops (mask not read), lines of 129..1024 cells that fill more than half of the
  // 32*epl slots (n = 2^m has one stage more and its 2x2 partner outside the warp: k_pcr_warp)
  const bool tm_ok = mask_is_inner && !no_tm && n > 128 && (1 << T.pn) == 32 * epl;
  // rows staged by TMA bulk copies: row pitch and base addresses 16-byte aligned, 5 rows + exchange buffer per warp
  const bool bulk = !no_bulk && n <= 512 && G.nk % 4 == 0 && ((uintptr_t)x % 16) == 0 && ((uintptr_t)rhs % 16) == 0;
  // 4x4 last stage (pcr_rb_esa & co): only on the tensor-memory kernel with bulk staging, else k_pcr_lines<1>
  const bool l4_ok = T.last4 && tm_ok && bulk;
  if ((!T.last4 || l4_ok) && !no_warp && n >= 17 && n <= 1024) {   // warp-per-line kernels
    auto go_tm = [&](auto epl, auto wpb_, auto bulk_, auto l4_) {   // prefetching variants with the tables in tensor memory
      constexpr int EPL = decltype(epl)::value, WPB = decltype(wpb_)::value;
      constexpr bool BULK = decltype(bulk_)::value, L4 = decltype(l4_)::value;
      typedef PcrLay<EPL> LY;
      const int rows = L4 ? 3 * T.stages : 3 * T.stages + 2;
      const int rows_tm = EPL != 32 ? rows : (T.stages < 5 ? 3 * T.stages : 15);
      const size_t per_warp = BULK ? (size_t)5 * G.nk + LY::ROW : (size_t)2 * LY::ROW;
      const size_t sm = sizeof(REAL) * ((size_t)(rows - rows_tm) * LY::ROW + (size_t)WPB * per_warp);
      int g = (sel.nlines + WPB - 1) / WPB;
      if (g > 148) g = 148;
      if ((size_t)g > L.rb_blocks) g = (int)L.rb_blocks;
      static size_t sm_set = 0;
      if (sm > sm_set) { CZB_CHECK(cudaFuncSetAttribute(k_pcr_warp_tm<EPL, WPB, BULK, L4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); sm_set = sm; }
      k_pcr_warp_tm<EPL, WPB, BULK, L4><<<g, WPB * 32, sm, L.s>>>(G, n, T.pn, T.stages, T.tab, T.fin, sel, x, dst, rhs, omg, 0x8000000080000000ull, L.st, L.rb, fs);
    };
    if (T.last4) {
      using std::integral_constant;
      if (n <= 256) go_tm(integral_constant<int, 8>{}, integral_constant<int, 16>{}, std::true_type{}, std::true_type{});
      else          go_tm(integral_constant<int, 16>{}, integral_constant<int, 16>{}, std::true_type{}, std::true_type{});
      CZB_CHECK(cudaGetLastError());
      count_launch();
      return;
    }
    // without bulk staging the source term is prefetched through registers, (epl == 32 ? 4 : 2) cells per stage
    if (tm_ok && (epl == 32 ? 4 : 2) * T.stages >= epl) {
      using std::integral_constant;
      if (n <= 256)      { if (bulk) go_tm(integral_constant<int, 8>{}, integral_constant<int, 16>{}, std::true_type{}, std::false_type{});
                           else      go_tm(integral_constant<int, 8>{}, integral_constant<int, 16>{}, std::false_type{}, std::false_type{}); }
      else if (n <= 512) { if (bulk) go_tm(integral_constant<int, 16>{}, integral_constant<int, 16>{}, std::true_type{}, std::false_type{});
                           else      go_tm(integral_constant<int, 16>{}, integral_constant<int, 16>{}, std::false_type{}, std::false_type{}); }
      else               go_tm(integral_constant<int, 32>{}, integral_constant<int, 12>{}, std::false_type{}, std::false_type{});
      CZB_CHECK(cudaGetLastError());
      count_launch();
      return;
    }
    auto go = [&](auto epl) {
      constexpr int EPL = decltype(epl)::value;
      typedef PcrLay<EPL> LY;
      const int wpb = (EPL == 32) ? 16 : 8;   // warps (= lines in flight) per CTA; one table copy per CTA
      const size_t sm = sizeof(REAL) * ((size_t)T.stages * 3 * LY::ROW + 2 * LY::ROW + (size_t)wpb * LY::ROW);
      int g = (sel.nlines + wpb - 1) / wpb;
      const int cap = 148 * (int)(sm > 110 * 1024 ? 1 : 2);
      if (g > cap) g = cap;
      if ((size_t)g > L.rb_blocks) g = (int)L.rb_blocks;
      static size_t sm_set = 0;
      if (sm > sm_set) { CZB_CHECK(cudaFuncSetAttribute(k_pcr_warp<EPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); sm_set = sm; }
      // mask_is_inner: msk is imask_k of this geometry, i.e. exactly 1 on every cell of every line -> not read
      k_pcr_warp<EPL><<<g, wpb * 32, sm, L.s>>>(G, n, T.pn, T.stages, T.tab, T.fin, sel, x, dst, mask_is_inner ? nullptr : msk, rhs, omg, L.st, L.rb, fs);
    };
    if (n <= 128)      go(std::integral_constant<int, 4>{});
    else if (n <= 256) go(std::integral_constant<int, 8>{});
    else if (n <= 512) go(std::integral_constant<int, 16>{});
    else               go(std::integral_constant<int, 32>{});
    CZB_CHECK(cudaGetLastError());
    count_launch();
    return;
  }
#else
  // double precision: the warp-per-line kernel with the tables in shared memory (2x2 last stage, 17..512 cells;
  // 1024-cell lines would need 390 KB of tables + exchange buffers)
  static const bool no_warp = getenv("CZB_PCR_BLOCK") != nullptr;
  if (!T.last4 && !no_warp && n >= 17 && n <= 512) {
    auto go = [&](auto epl) {
      constexpr int EPL = decltype(epl)::value;
      typedef PcrLay<EPL> LY;
      const int wpb = 8;
      const size_t sm = sizeof(REAL) * ((size_t)T.stages * 3 * LY::ROW + 2 * LY::ROW + (size_t)wpb * LY::ROW);
      int g = (sel.nlines + wpb - 1) / wpb;
      const int cap = 148 * (int)(sm > 110 * 1024 ? 1 : 2);
      if (g > cap) g = cap;
      if ((size_t)g > L.rb_blocks) g = (int)L.rb_blocks;
      static size_t sm_set = 0;
      if (sm > sm_set) { CZB_CHECK(cudaFuncSetAttribute(k_pcr_warp<EPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); sm_set = sm; }
      k_pcr_warp<EPL><<<g, wpb * 32, sm, L.s>>>(G, n, T.pn, T.stages, T.tab, T.fin, sel, x, dst, mask_is_inner ? nullptr : msk, rhs, omg, L.st, L.rb, fs);
    };
    if (n <= 128)      go(std::integral_constant<int, 4>{});
    else if (n <= 256) go(std::integral_constant<int, 8>{});
    else               go(std::integral_constant<int, 16>{});
    CZB_CHECK(cudaGetLastError());
    count_launch();
    return;
  }
#endif
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

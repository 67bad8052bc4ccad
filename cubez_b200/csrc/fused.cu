// cubez_b200/csrc/fused.cu -- bandwidth-optimised plane-marching kernels (sm_100a).
//
// One CTA owns a (k,i) tile of the cube and marches along j (the slowest axis).  Every j-plane
// of the tile (+2 halo cells per side) is fetched ONCE by a TMA tiled load
// (cp.async.bulk.tensor.3d -> UTMALDG) into a ring of shared-memory slots, several planes
// ahead of the compute; completion is tracked with mbarriers.  Each thread owns CPT
// "pair columns" (two k-adjacent cells, one red one black) whose last four planes live in
// registers, so j-neighbours never touch memory again; in-plane neighbours come from a
// de-interleaved (even-k / odd-k) shared copy of two planes, which makes every neighbour read
// stride-1 across the warp (no bank conflicts).
//
//   MODE_RB2 : BOTH colours of one red-black SOR iteration (2 x psor2sma_core,
//              cz_solver.f90:404-493) in ONE pass over memory: at step s the CTA relaxes the
//              red cells of plane s+1 and then the black cells of plane s (which need the new
//              reds of planes s-1, s, s+1).  Reds of the 1-cell ring around the tile and of the
//              planes next to the j-chunk are recomputed redundantly, so CTAs never exchange
//              data; results go to a second array (ping-pong), which removes the in-place
//              hazard.  HBM traffic: read p, read b, write p = 12 B/cell.iteration (float)
//              against 24 B for two in-place stride-2 passes.
//   MODE_JAC : relaxed Jacobi sweep (cz_solver.f90:284-387) without the copy-back pass.
//   MODE_AX  : 7-point operator ap = ss - dd*p (cz_blas.f90:579-644) fused with (ap, w).
//
// Arithmetic is the reference's, operation for operation (see kernels.cu), so every cell is
// bit-identical to the baseline kernels and to the oracle.
#include "kernels.cuh"
#include <cuda.h>
#include <map>
#include <tuple>
#include <type_traits>

namespace czb {

namespace {

enum { MODE_RB2 = 0, MODE_JAC = 1, MODE_AX = 2 };

constexpr int PR = 64;          // pairs per tile row  -> box width EKB = 128 cells
constexpr int EKB = 2 * PR;
constexpr int NSP = 3;          // raw TMA slots for p  (planes s+2, s+3, s+4)
constexpr int NSB = 3;          // raw TMA slots for b  (planes s+1, s+2, s+3)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a lost TMA must abort the kernel, never hang the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
#pragma unroll 1
  for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (ok) return;
  }
  __trap();
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

struct Coef7 { REAL c1, c2, c3, c4, c5, c6, dd; };

struct MarchArgs {
  Geo G;
  int ei;                 // tile rows (box height); owned rows = ei - 4
  int ntk, nti, njc;      // tiles along k, i and chunks along j
  int ngy;                // chunks launched by this call (grid.y)
  int jl;                 // planes per j-chunk
  int jr0, jr1;           // planes on which reds are relaxed (Fortran j)
  int jb0, jb1;           // owned planes (blacks relaxed, results stored, residual counted)
  int ip;                 // colour offset: cell is red iff (k-kst + i + j + ip) is even
  Coef7 c; REAL omg;
  REAL* out;              // MODE_RB2/JAC: new field ; MODE_AX: ap
  const REAL* w;          // MODE_AX: dot operand (may be null)
  const LoopState* st;
  LoopState* acc_st; int slot; int add;
  RedBuf rb;
  FinishSpec fin;         // convergence test folded into the reduction epilogue (single rank)
  int slot2;              // packed MODE_AX: second accumulator (ap, ap), -1 = none
  int zsrc;               // packed MODE_RB2: the source field is all zero and is not read
  int jc_first, jc_step;  // chunk of a CTA = jc_first + blockIdx.y * jc_step (all chunks: 0, 1; see `part` in march())
};

struct FinStoreF {
  LoopState* st; int slot; int add;
  FinishSpec fin;
  __device__ void operator()(double (&tot)[1]) const {
    if (!st) return;
    const double sum = (add ? st->acc[slot] : 0.0) + tot[0];
    st->acc[slot] = sum;
    if (fin.enable) {             // = k_finish_iteration (kernels.cu), cz_Poisson.cpp:69-77
      double res = sum;
      res *= fin.res_normal;
      res = sqrt(res);
      st->res = res;
      const int it = st->itr;
      if (fin.hist && it < fin.hist_len) fin.hist[it] = res;
      st->itr = it + 1;
      if (res < fin.eps) st->done = 1;
    }
  }
};

// One relaxation, exactly cz_solver.f90:339-347.  UNIT: all six off-diagonal coefficients are
// 1.0 (the CubeZ problem, cz.h:169-172); 1.0*x == x bit for bit, so the products are skipped.
template <bool UNIT>
__device__ __forceinline__ REAL relax(const Coef7& c, REAL omg, REAL pp, REAL bb, REAL ie, REAL iw, REAL jn, REAL js,
                                      REAL kt, REAL kb, REAL& dp) {
  REAL ss;
  if (UNIT) ss = ie + iw + jn + js + kt + kb;
  else ss = c.c1 * ie + c.c2 * iw + c.c3 * jn + c.c4 * js + c.c5 * kt + c.c6 * kb;
  dp = ((ss - bb) / c.dd - pp) * omg;
  return pp + dp;
}
template <bool UNIT>
__device__ __forceinline__ REAL apply_ax(const Coef7& c, REAL pp, REAL ie, REAL iw, REAL jn, REAL js, REAL kt, REAL kb) {
  REAL ss;
  if (UNIT) ss = ie + iw + jn + js + kt + kb;
  else ss = c.c1 * ie + c.c2 * iw + c.c3 * jn + c.c4 * js + c.c5 * kt + c.c6 * kb;
  return ss - c.dd * pp;
}

struct R2 { REAL x, y; };   // a pair column value: x = even-k cell (A), y = odd-k cell (B)

__device__ __forceinline__ R2 lds2(const REAL* p) {
#ifdef _REAL_IS_DOUBLE_
  double2 v = *reinterpret_cast<const double2*>(p); return R2{v.x, v.y};
#else
  float2 v = *reinterpret_cast<const float2*>(p); return R2{v.x, v.y};
#endif
}
__device__ __forceinline__ void stg2(REAL* p, R2 v) {
#ifdef _REAL_IS_DOUBLE_
  *reinterpret_cast<double2*>(p) = make_double2(v.x, v.y);
#else
  *reinterpret_cast<float2*>(p) = make_float2(v.x, v.y);
#endif
}

// ---- shared-memory access with 32-bit addresses and immediate offsets ---------------------
#ifdef _REAL_IS_DOUBLE_
#define CZB_LD "ld.shared.f64"
#define CZB_ST "st.shared.f64"
#define CZB_LD2 "ld.shared.v2.f64"
#define CZB_RC "d"
#else
#define CZB_LD "ld.shared.f32"
#define CZB_ST "st.shared.f32"
#define CZB_LD2 "ld.shared.v2.f32"
#define CZB_RC "f"
#endif
template <int OFF>
__device__ __forceinline__ REAL lds(uint32_t addr) {
  REAL v;
  asm volatile(CZB_LD " %0, [%1+%2];" : "=" CZB_RC(v) : "r"(addr), "n"(OFF) : "memory");
  return v;
}
template <int OFF>
__device__ __forceinline__ void sts(uint32_t addr, REAL v) {
  asm volatile(CZB_ST " [%0+%1], %2;" ::"r"(addr), "n"(OFF), CZB_RC(v) : "memory");
}
// store only where `flag` != 0, as a predicated instruction (a C++ `if` around inline asm becomes a branch)
template <int OFF>
__device__ __forceinline__ void sts_if(uint32_t addr, REAL v, unsigned flag) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p " CZB_ST " [%0+%1], %2;\n\t}" ::"r"(addr), "n"(OFF), CZB_RC(v), "r"(flag) : "memory");
}
__device__ __forceinline__ void stg2_if(REAL* p, R2 v, unsigned flag) {
#ifdef _REAL_IS_DOUBLE_
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p st.global.v2.f64 [%0], {%1, %2};\n\t}" ::"l"(p), "d"(v.x), "d"(v.y), "r"(flag) : "memory");
#else
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p st.global.v2.f32 [%0], {%1, %2};\n\t}" ::"l"(p), "f"(v.x), "f"(v.y), "r"(flag) : "memory");
#endif
}
__device__ __forceinline__ R2 lds_pair(uint32_t addr) {
  R2 v;
  asm volatile(CZB_LD2 " {%0, %1}, [%2];" : "=" CZB_RC(v.x), "=" CZB_RC(v.y) : "r"(addr) : "memory");
  return v;
}

// (ss - bb)/dd.  DIVK: dd == 6 (the CubeZ operator).  a/6 is evaluated as q = a*y, r = fma(-6,q,a),
// q' = fma(r,y,q) with y = RN(1/6): for float this was checked EXHAUSTIVELY against IEEE division
// for every a with biased exponent in [27,247] (0 mismatches of 3.7e9; tests/test_div6.py keeps a
// sampled version); zeros keep their sign; anything else (subnormal / huge / non-finite) takes
// the IEEE division.  The IEEE sequence costs ~45 instructions when an operand is zero or
// subnormal, which is the common case while the solution front is still spreading from the walls.
// Out-of-line IEEE division for the rare operands outside the guarded range; __noinline__ keeps
// the compiler from speculating the ~45-instruction sequence onto the common path.
#ifdef _REAL_IS_DOUBLE_
__device__ __noinline__ REAL div_rare(REAL a, REAL dd) { return a / dd; }
#else
// float: the quotient is formed in double and rounded once more; double rounding is innocuous for
// division when the wide format has >= 2p+2 bits (53 >= 50), also for subnormal results, and the
// double division never meets a subnormal operand (the float IEEE slow path would).
__device__ __noinline__ REAL div_rare(REAL a, REAL dd) { return (float)((double)a / (double)dd); }
#endif

// Exact division by 6 without the IEEE sequence (which costs ~45 instructions as soon as an operand
// is zero or subnormal -- the common case while the solution front is still spreading from the
// walls).  float: as = a*2^64 (exact), q = as*y, r = fma(-6,q,as), q' = fma(r,y,q) [Markstein: the
// correctly rounded quotient], result q'*2^-64 (exact while the quotient is a normal number, and
// verified also for the subnormal operands).  EXHAUSTIVELY checked against a/6 for every float
// with biased exponent 0, 1 and 4..183 (tests/tools/div6_check.c: 3.1e9 values, 0 mismatches).
// Operands outside that set (exponents 2,3: the quotient would be rounded twice; |a| >= 2^57;
// -0; non-finite) are flagged `rare` and patched by the IEEE division behind ONE branch per
// group of quotients (div_fix).  double: same sequence unscaled, guarded to |a| in [2^-900,2^900].
#ifdef _REAL_IS_DOUBLE_
__device__ __forceinline__ bool div_guard_fail(double a) {
  const unsigned e = ((unsigned)__double2hiint(a) >> 20) & 0x7ffu;
  return (e - 123u > 1800u) && (__double_as_longlong(a) != 0ll);   // +0 is exact on the fast path, -0 is not
}
#else
__device__ __forceinline__ bool div_guard_fail(float a) {
  const unsigned bits = __float_as_uint(a);
  const unsigned e = (bits >> 23) & 0xffu;
  return ((e - 4u > 179u) && (e > 1u)) || (bits == 0x80000000u);
}
#endif
template <bool DIVK>
__device__ __forceinline__ REAL div_fast(REAL a, REAL dd, bool& rare) {
  if (!DIVK) return a / dd;
  const REAL y = (REAL)1.0 / (REAL)6.0;
#ifdef _REAL_IS_DOUBLE_
  const REAL q = a * y;
  const REAL r = fma(-6.0, q, a);
  const REAL res = fma(r, y, q);
#else
  const REAL as = a * 0x1p64f;
  const REAL q = as * y;
  const REAL r = __fmaf_rn(-6.0f, q, as);
  const REAL res = __fmaf_rn(r, y, q) * 0x1p-64f;
#endif
  rare = rare || div_guard_fail(a);
  return res;
}
__device__ __forceinline__ REAL div_fix(REAL a, REAL q, REAL dd) {
  return div_guard_fail(a) ? div_rare(a, dd) : q;
}
template <bool DIVK>
__device__ __forceinline__ REAL div_dd(REAL a, REAL dd) {
  bool rare = false;
  REAL q = div_fast<DIVK>(a, dd, rare);
  if (DIVK && rare) q = div_rare(a, dd);
  return q;
}

template <bool UNIT>
__device__ __forceinline__ REAL sum6(const Coef7& c, REAL ie, REAL iw, REAL jn, REAL js, REAL kt, REAL kb) {
  // cz_solver.f90:339-344 order: i+1, i-1, j+1, j-1, k+1, k-1
  if (UNIT) return ie + iw + jn + js + kt + kb;
  return c.c1 * ie + c.c2 * iw + c.c3 * jn + c.c4 * js + c.c5 * kt + c.c6 * kb;
}

// Per-column constant data.
struct Col {
  uint32_t araw;   // byte offset of the pair inside a raw (interleaved) tile slot
  uint32_t ad;     // byte offset of the pair inside the even-k half of a de-interleaved plane
  int gcol;        // element offset of cell A inside a j-plane of the global array
  unsigned fl;     // bit0/1 red-ok A/B, bit2/3 owned A/B, bit4 store pair
};

template <int MODE, int NT, int CPT, bool UNIT, bool DIVK>
__global__ void __launch_bounds__(NT, 1)
k_march(const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_b, const MarchArgs a) {
  if (a.st && a.st->done) return;
  constexpr int RP = NT / PR;              // tile rows covered by one pass of the CTA
  constexpr int EI = RP * CPT;             // tile rows
  constexpr int T = EI * EKB;              // cells per tile plane
  constexpr int SZ = (int)sizeof(REAL);
  constexpr int TB = T * SZ;               // bytes per tile plane
  constexpr int HPB = TB / 2;              // bytes per de-interleaved half plane (even-k, then odd-k)
  constexpr int PRB = PR * SZ;             // bytes per row of a half plane
  constexpr int NB = (MODE == MODE_AX) ? 0 : NSB;
  const Geo& G = a.G;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t sm0 = smem_u32(smem_raw);
  const uint32_t raw_p = sm0;                          // [NSP][TB]
  const uint32_t raw_b = raw_p + NSP * TB;             // [NB][TB]
  const uint32_t dei = raw_b + NB * TB;                // [3][TB]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)(NSP + NB + 3) * TB);   // [NSP + NSB]
  const int tid = threadIdx.x;

  // ---- tile / chunk geometry ------------------------------------------------------------
  const int tk = blockIdx.x % a.ntk, ti = blockIdx.x / a.ntk, jc = a.jc_first + (int)blockIdx.y * a.jc_step;
  const int kst_c = G.kst + G.g - 1, ist_c = G.ist + G.g - 1;     // C offsets of the first inner cell
  const int P0 = (kst_c >> 1) - 1 + tk * (PR - 2);                 // first (halo) pair of the box
  const int R0 = ist_c - 2 + ti * (EI - 4);                        // first (halo) row of the box
  const int ja = a.jb0 + jc * a.jl;                                // owned planes of this chunk
  const int jb = min(ja + a.jl - 1, a.jb1);
  const int pl0 = (MODE == MODE_RB2) ? ja - 2 : ja - 1;            // first p plane fetched
  const int bpl0 = (MODE == MODE_RB2) ? ja - 1 : ja;               // first b plane fetched
  const int nsteps = (MODE == MODE_RB2) ? (jb - ja + 3) : (jb - ja + 1);
  const int npl = nsteps + 2;                                      // p planes fetched
  const int nbpl = nsteps;                                         // b planes fetched

  if (tid == 0) {
    for (int s = 0; s < NSP + NSB; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue_p = [&](int t) {   // plane pl0+t -> slot t % NSP
    uint64_t* bar = &bars[t % NSP];
    mbar_expect_tx(bar, (uint32_t)TB);
    tma_load_3d(smem_raw + (size_t)(t % NSP) * TB, &tm_p, bar, 2 * P0, R0, pl0 + t + G.g - 1);
  };
  auto issue_b = [&](int t) {
    uint64_t* bar = &bars[NSP + t % NSB];
    mbar_expect_tx(bar, (uint32_t)TB);
    tma_load_3d(smem_raw + (size_t)(NSP + t % NSB) * TB, &tm_b, bar, 2 * P0, R0, bpl0 + t + G.g - 1);
  };
  if (tid == 0) {
    for (int t = 0; t < NSP && t < npl; ++t) issue_p(t);
    if (MODE != MODE_AX) for (int t = 0; t < NSB && t < nbpl; ++t) issue_b(t);
  }

  // ---- per-column constants -----------------------------------------------------------------
  // RP is even, so all CPT columns of a thread have the same row parity and therefore the same
  // colour phase: `par` (parity of cell A at plane 0) is a per-thread (in fact per-warp) value.
  Col col[CPT];
  const int q = tid % PR;
  const int kcA = 2 * (P0 + q);
  const int kFA = kcA - G.g + 1;
  const unsigned par = (unsigned)(kFA - G.kst + (R0 + tid / PR - G.g + 1) + a.ip) & 1u;
#pragma unroll
  for (int m = 0; m < CPT; ++m) {
    const int r = tid / PR + RP * m;
    Col c; c.araw = (uint32_t)(r * EKB + 2 * q) * SZ; c.ad = (uint32_t)(r * PR + q) * SZ; c.fl = 0; c.gcol = 0;
    const int ic = R0 + r;
    if (ic < G.ni && kcA + 1 < G.nk) {
      const int iF = ic - G.g + 1, kFB = kFA + 1;
      const bool rowin = (iF >= G.ist && iF <= G.ied);
      const bool inA = rowin && kFA >= G.kst && kFA <= G.ked, inB = rowin && kFB >= G.kst && kFB <= G.ked;
      const int dr = min(r, EI - 1 - r);
      const bool redA = inA && dr >= 1 && q >= 1;                     // needs k-1 inside the box
      const bool redB = inB && dr >= 1 && q <= PR - 2;                // needs k+1 inside the box
      const bool own = (dr >= 2) && (q >= 1) && (q <= PR - 2);
      c.fl = (redA ? 1u : 0u) | (redB ? 2u : 0u) | ((own && inA) ? 4u : 0u) | ((own && inB) ? 8u : 0u) |
             ((own && (inA || inB)) ? 16u : 0u);
      c.gcol = kcA + G.nk * ic;
    }
    col[m] = c;
  }

  R2 cm1[CPT], c0[CPT], c1[CPT], c2[CPT], bprev[CPT], bcur[CPT];
#pragma unroll
  for (int m = 0; m < CPT; ++m) { cm1[m] = c0[m] = c1[m] = c2[m] = bprev[m] = bcur[m] = R2{(REAL)0, (REAL)0}; }
  double acc[1] = {0.0};
  const Coef7 cf = a.c;
  const REAL omg = a.omg;

  // ---- prologue: planes pl0 (-> c0) and pl0+1 (-> c1 + de-interleaved copy in ring slot 1) -----
  mbar_wait(&bars[0], 0);
#pragma unroll
  for (int m = 0; m < CPT; ++m) c0[m] = lds_pair(raw_p + col[m].araw);
  mbar_wait(&bars[1], 0);
#pragma unroll
  for (int m = 0; m < CPT; ++m) {
    c1[m] = lds_pair(raw_p + TB + col[m].araw);
    sts<0>(dei + TB + col[m].ad, c1[m].x); sts<HPB>(dei + TB + col[m].ad, c1[m].y);
  }
  __syncthreads();
  if (tid == 0) { if (NSP < npl) issue_p(NSP); if (NSP + 1 < npl) issue_p(NSP + 1); }

  // ---- main loop ----------------------------------------------------------------------------------
  // One step.  The four plane registers rotate roles every step; instead of moving 32 registers
  // the step body is instantiated four times with the roles permuted (see the 4x unrolled loop).
  int ring = 0;          // u % 3 : ring slot of de-interleaved plane s; raw b slot of this step
  int phase = 0;         // (u / 3) & 1 : mbarrier parity of the b slot; p slot (u+2)%3 has parity pphase
  auto step = [&](int u, R2 (&wm1)[CPT], R2 (&w0)[CPT], R2 (&w1)[CPT], R2 (&w2)[CPT], R2 (&bp)[CPT], R2 (&bc)[CPT]) {
    const int s = pl0 + u;                 // RB2: blacks of plane s, reds of s+1 ; JAC/AX: plane s+1
    const int r0 = ring, r1 = (ring == 2) ? 0 : ring + 1, r2 = (ring == 0) ? 2 : ring - 1;   // u%3, (u+1)%3, (u+2)%3
    const uint32_t e0 = dei + (uint32_t)r0 * TB, e1 = dei + (uint32_t)r1 * TB, e2 = dei + (uint32_t)r2 * TB;
    const int pphase = (ring == 0) ? phase : phase ^ 1;        // ((u+2)/3)&1
    mbar_wait(&bars[r2], (uint32_t)pphase);
    {
      const uint32_t rp = raw_p + (uint32_t)r2 * TB;
#pragma unroll
      for (int m = 0; m < CPT; ++m) w2[m] = lds_pair(rp + col[m].araw);
#pragma unroll
      for (int m = 0; m < CPT; ++m) { sts<0>(e2 + col[m].ad, w2[m].x); sts<HPB>(e2 + col[m].ad, w2[m].y); }
    }
    if (MODE != MODE_AX) {
      mbar_wait(&bars[NSP + r0], (uint32_t)phase);
      const uint32_t rbp = raw_b + (uint32_t)r0 * TB;
#pragma unroll
      for (int m = 0; m < CPT; ++m) bc[m] = lds_pair(rbp + col[m].araw);
    }
    const long long gpl1 = (long long)(s + 1 + G.g - 1) * G.plane;   // global offset of plane s+1
    const long long gpl0 = (long long)(s + G.g - 1) * G.plane;       // global offset of plane s

    if (MODE == MODE_RB2) {
      const bool do_red = (s + 1 >= a.jr0 && s + 1 <= a.jr1);      // CTA-uniform
      const bool red_owned = (s + 1 >= ja && s + 1 <= jb);
      const bool do_black = (s >= ja && s <= jb);
      // A is the red cell of plane s+1 iff (par + s + 1) is even  (warp-uniform)
      const bool xa = ((par + (unsigned)(s + 1)) & 1u) == 0u;
      // One colour phase over the CPT columns of the thread, X = A (xb false) or B (xb true):
      // straight-line code for all columns, a single rare branch for operands of the division
      // outside its guarded range.
      auto both = [&](auto xb) {
        constexpr bool XB = decltype(xb)::value;
        REAL kn1[CPT], ie1[CPT], iw1[CPT], kn0[CPT], ie0[CPT], iw0[CPT];
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
          const uint32_t a1 = e1 + col[m].ad, a0 = e0 + col[m].ad;
          if (!XB) {   // cell A: k-1 = odd half [q-1], k+1 = own B, i+-1 = even half rows r+-1
            kn1[m] = lds<HPB - SZ>(a1); ie1[m] = lds<PRB>(a1); iw1[m] = lds<-PRB>(a1);
            kn0[m] = lds<HPB - SZ>(a0); ie0[m] = lds<PRB>(a0); iw0[m] = lds<-PRB>(a0);
          } else {     // cell B: k-1 = own A, k+1 = even half [q+1], i+-1 = odd half rows r+-1
            kn1[m] = lds<SZ>(a1); ie1[m] = lds<HPB + PRB>(a1); iw1[m] = lds<HPB - PRB>(a1);
            kn0[m] = lds<SZ>(a0); ie0[m] = lds<HPB + PRB>(a0); iw0[m] = lds<HPB - PRB>(a0);
          }
        }
        if (do_red) {
          REAL num[CPT], quo[CPT];
          bool rare = false;
#pragma unroll
          for (int m = 0; m < CPT; ++m) {
            const REAL ss = XB ? sum6<UNIT>(cf, ie1[m], iw1[m], w2[m].y, w0[m].y, kn1[m], w1[m].x)
                               : sum6<UNIT>(cf, ie1[m], iw1[m], w2[m].x, w0[m].x, w1[m].y, kn1[m]);
            num[m] = ss - (XB ? bc[m].y : bc[m].x);
            quo[m] = div_fast<DIVK>(num[m], cf.dd, rare);
          }
          if (DIVK && rare) {
#pragma unroll
            for (int m = 0; m < CPT; ++m) quo[m] = div_fix(num[m], quo[m], cf.dd);
          }
#pragma unroll
          for (int m = 0; m < CPT; ++m) {
            const unsigned fl = col[m].fl;
            const REAL pp = XB ? w1[m].y : w1[m].x;
            const REAL dp = (quo[m] - pp) * omg;
            const bool ok = (fl & (XB ? 2u : 1u)) != 0u;
            const REAL nv = ok ? pp + dp : pp;
            if (XB) w1[m].y = nv; else w1[m].x = nv;
            sts_if<XB ? HPB : 0>(e1 + col[m].ad, nv, fl & (XB ? 2u : 1u));
            const REAL d2 = (ok && red_owned && (fl & (XB ? 8u : 4u))) ? dp * dp : (REAL)0;
            acc[0] += (double)d2;
          }
        }
        if (do_black) {
          REAL num[CPT], quo[CPT];
          bool rare = false;
#pragma unroll
          for (int m = 0; m < CPT; ++m) {
            const REAL ss = XB ? sum6<UNIT>(cf, ie0[m], iw0[m], w1[m].y, wm1[m].y, kn0[m], w0[m].x)
                               : sum6<UNIT>(cf, ie0[m], iw0[m], w1[m].x, wm1[m].x, w0[m].y, kn0[m]);
            num[m] = ss - (XB ? bp[m].y : bp[m].x);
            quo[m] = div_fast<DIVK>(num[m], cf.dd, rare);
          }
          if (DIVK && rare) {
#pragma unroll
            for (int m = 0; m < CPT; ++m) quo[m] = div_fix(num[m], quo[m], cf.dd);
          }
#pragma unroll
          for (int m = 0; m < CPT; ++m) {
            const unsigned fl = col[m].fl;
            const REAL pp = XB ? w0[m].y : w0[m].x;
            const REAL dp = (quo[m] - pp) * omg;
            const bool ok = (fl & (XB ? 8u : 4u)) != 0u;
            const REAL nv = ok ? pp + dp : pp;
            if (XB) w0[m].y = nv; else w0[m].x = nv;
            acc[0] += (double)(ok ? dp * dp : (REAL)0);
            stg2_if(a.out + gpl0 + col[m].gcol, w0[m], fl & 16u);
          }
        }
      };
      if (xa) both(std::false_type{}); else both(std::true_type{});
    } else {
      // Jacobi / operator: plane s+1, both cells of the pair, old values only
      REAL kmA[CPT], ieA[CPT], iwA[CPT], ktB[CPT], ieB[CPT], iwB[CPT];
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        const uint32_t a1 = e1 + col[m].ad;
        kmA[m] = lds<HPB - SZ>(a1); ieA[m] = lds<PRB>(a1); iwA[m] = lds<-PRB>(a1);
        ktB[m] = lds<SZ>(a1); ieB[m] = lds<HPB + PRB>(a1); iwB[m] = lds<HPB - PRB>(a1);
      }
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        const unsigned fl = col[m].fl;
        const REAL ssA = sum6<UNIT>(cf, ieA[m], iwA[m], w2[m].x, w0[m].x, w1[m].y, kmA[m]);
        const REAL ssB = sum6<UNIT>(cf, ieB[m], iwB[m], w2[m].y, w0[m].y, ktB[m], w1[m].x);
        const bool okA = (fl & 4u) != 0u, okB = (fl & 8u) != 0u;
        if (MODE == MODE_JAC) {
          const REAL dpA = (div_dd<DIVK>(ssA - bc[m].x, cf.dd) - w1[m].x) * omg;
          const REAL dpB = (div_dd<DIVK>(ssB - bc[m].y, cf.dd) - w1[m].y) * omg;
          R2 o;
          o.x = okA ? w1[m].x + dpA : w1[m].x;
          o.y = okB ? w1[m].y + dpB : w1[m].y;
          acc[0] += (double)(okA ? dpA * dpA : (REAL)0);
          acc[0] += (double)(okB ? dpB * dpB : (REAL)0);
          stg2_if(a.out + gpl1 + col[m].gcol, o, fl & 16u);
        } else {
          // MODE_AX (solver-internal): non-inner cells of a stored pair are boundary nodes of
          // q / t_, which the reference leaves at their initial 0 (blas_clear / zero-filled
          // allocation) -> store 0 there.
          if (fl & 16u) {
            R2 v;
            v.x = okA ? ssA - cf.dd * w1[m].x : (REAL)0;
            v.y = okB ? ssB - cf.dd * w1[m].y : (REAL)0;
            if (a.w) {
              const R2 wv = lds2(a.w + gpl1 + col[m].gcol);
              acc[0] += (double)(okA ? v.x * wv.x : (REAL)0);
              acc[0] += (double)(okB ? v.y * wv.y : (REAL)0);
            }
            stg2(a.out + gpl1 + col[m].gcol, v);
          }
        }
      }
    }
    __syncthreads();
    if (tid == 0) {   // both slots consumed by every thread: refill them three planes ahead
      if (u + 2 + NSP < npl) issue_p(u + 2 + NSP);
      if (MODE != MODE_AX && u + NSB < nbpl) issue_b(u + NSB);
    }
    if (ring == 2) { ring = 0; phase ^= 1; } else ++ring;
  };
  for (int u = 0; u < nsteps; u += 4) {
    step(u, cm1, c0, c1, c2, bprev, bcur);
    if (u + 1 >= nsteps) break;
    step(u + 1, c0, c1, c2, cm1, bcur, bprev);
    if (u + 2 >= nsteps) break;
    step(u + 2, c1, c2, cm1, c0, bprev, bcur);
    if (u + 3 >= nsteps) break;
    step(u + 3, c2, cm1, c0, c1, bcur, bprev);
  }
  grid_reduce<1>(acc, a.rb, FinStoreF{a.acc_st, a.slot, a.add, a.fin});
}

#ifndef _REAL_IS_DOUBLE_
#include "fused_pk.cuh"
#endif

}  // namespace

// ============================================================================ host side
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CZB_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
    if (!p || qres != cudaDriverEntryPointSuccess) { fprintf(stderr, "czb200: cuTensorMapEncodeTiled unavailable\n"); abort(); }
    fn = (EncodeTiledFn)p;
  }
  return fn;
}

typedef std::tuple<const void*, int, int, int, int> TmKey;
std::map<TmKey, CUtensorMap> g_tmaps;

// Returned by value: a later call may evict the cache, a reference into it would dangle.
CUtensorMap tmap_for(const REAL* base, const Geo& G, int ei) {
  TmKey key(base, G.nk, G.ni, G.nj, ei);
  auto it = g_tmaps.find(key);
  if (it != g_tmaps.end()) return it->second;
  CUtensorMap tm;
  cuuint64_t gdim[3] = {(cuuint64_t)G.nk, (cuuint64_t)G.ni, (cuuint64_t)G.nj};
  cuuint64_t gstr[2] = {(cuuint64_t)G.nk * sizeof(REAL), (cuuint64_t)G.plane * sizeof(REAL)};
  cuuint32_t box[3] = {(cuuint32_t)EKB, (cuuint32_t)ei, 1u};
  cuuint32_t estr[3] = {1u, 1u, 1u};
  CUresult r = encode_fn()(&tm, sizeof(REAL) == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3,
                           (void*)base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { fprintf(stderr, "czb200: cuTensorMapEncodeTiled failed (%d)\n", (int)r); abort(); }
  if (g_tmaps.size() > 256) g_tmaps.clear();
  return g_tmaps.emplace(key, tm).first->second;
}

template <int MODE, int NT, int CPT>
void launch_march_t(const Launch& L, const MarchArgs& a, const CUtensorMap& tp, const CUtensorMap& tb, bool fast, size_t smem) {
  dim3 grid(a.ntk * a.nti, a.ngy, 1);
  if (fast) {   // unit off-diagonals and dd == 6: the CubeZ operator (cz.h:169-172)
    static bool set = false;
    if (!set) { CZB_CHECK(cudaFuncSetAttribute(k_march<MODE, NT, CPT, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); set = true; }
    k_march<MODE, NT, CPT, true, true><<<grid, NT, smem, L.s>>>(tp, tb, a);
  } else {
    static bool set = false;
    if (!set) { CZB_CHECK(cudaFuncSetAttribute(k_march<MODE, NT, CPT, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); set = true; }
    k_march<MODE, NT, CPT, false, false><<<grid, NT, smem, L.s>>>(tp, tb, a);
  }
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

// Tile configuration: NT threads, CPT pair columns per thread -> tile rows EI = NT/64*CPT.
//   f32: cfg 0 = 512 x 4 (128x32 tile, 1 CTA/SM, 16 warps)   cfg 1 = 1024 x 2 (same tile, 32 warps)
//   f64: cfg 0 = 256 x 4 (128x16 tile of doubles, 8 warps)   cfg 1 = 512 x 2 (same tile, 16 warps)
template <int MODE>
void launch_march_cfg(int cfg, const Launch& L, const MarchArgs& a, const CUtensorMap& tp, const CUtensorMap& tb, bool fast, size_t smem) {
#ifdef _REAL_IS_DOUBLE_
  if (cfg == 1) launch_march_t<MODE, 512, 2>(L, a, tp, tb, fast, smem);
  else          launch_march_t<MODE, 256, 4>(L, a, tp, tb, fast, smem);
#else
  if (cfg == 1) launch_march_t<MODE, 1024, 2>(L, a, tp, tb, fast, smem);
  else          launch_march_t<MODE, 512, 4>(L, a, tp, tb, fast, smem);
#endif
}

#ifndef _REAL_IS_DOUBLE_
template <int NT, int CPT, int NSP_, int NSB_>
void launch_rb2_pk_t(const Launch& L, const MarchArgs& a, const REAL* p, const REAL* b) {
  constexpr int EI = NT / 32 * CPT;
  dim3 grid(a.ntk * a.nti, a.ngy, 1);
  const size_t smem = (size_t)(NSP_ + NSB_ + NDE) * EI * EKB * sizeof(float) + 1024;
  static bool set = false;
  if (!set) {
    CZB_CHECK(cudaFuncSetAttribute(k_rb2_pk<NT, CPT, NSP_, NSB_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set = true;
  }
  const CUtensorMap tp = tmap_for(p, a.G, EI);
  const CUtensorMap tb = tmap_for(b, a.G, EI);
  k_rb2_pk<NT, CPT, NSP_, NSB_><<<grid, NT, smem, L.s>>>(tp, tb, a);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}
#endif

#ifndef _REAL_IS_DOUBLE_
template <int MODE, int NT, int CPT, int NSP_, int NSB_>
void launch_lin_pk_t(const Launch& L, const MarchArgs& a, const REAL* p, const REAL* b) {
  constexpr int EI = NT / 32 * CPT;
  dim3 grid(a.ntk * a.nti, a.ngy, 1);
  const size_t smem = (size_t)(NSP_ + (MODE == MODE_JAC ? NSB_ : 0) + 2) * EI * EKB * sizeof(float) + 1024;
  static bool set = false;
  if (!set) { CZB_CHECK(cudaFuncSetAttribute(k_lin_pk<MODE, NT, CPT, NSP_, NSB_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); set = true; }
  const CUtensorMap& tp = tmap_for(p, a.G, EI);
  const CUtensorMap& tb = (MODE == MODE_JAC) ? tmap_for(b, a.G, EI) : tp;
  k_lin_pk<MODE, NT, CPT, NSP_, NSB_><<<grid, NT, smem, L.s>>>(tp, tb, a);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}
#endif

int march_cfg() {
  static int cfg = -1;
#ifdef _REAL_IS_DOUBLE_
  if (cfg < 0) { const char* e = getenv("CZB_MARCH_CFG"); cfg = e ? atoi(e) : 1; }   // 512 x 2: +18 % over 256 x 4
#else
  if (cfg < 0) { const char* e = getenv("CZB_MARCH_CFG"); cfg = e ? atoi(e) : 4; }
#endif
  return cfg;
}
int march_rows(int cfg) {
#ifdef _REAL_IS_DOUBLE_
  (void)cfg; return 16;
#else
  (void)cfg; return 32;
#endif
}

// returns 0 = not eligible, 1 = done, 2 = done including the second accumulator slot2 (packed MODE_AX)
int march(int mode, const Launch& L, const Geo& G, const REAL* p, const REAL* b_or_w, REAL* out, const REAL* cf,
          REAL omg, int ip, int jr0, int jr1, int slot, bool add, int slot2 = -1, bool zsrc = false, int part = 0,
          bool dry = false) {
  if (!fused_eligible(G)) return 0;
  MarchArgs a;
  a.G = G;
  const int cfg = march_cfg();
  a.ei = march_rows(cfg);
  const bool unit = cf[0] == (REAL)1 && cf[1] == (REAL)1 && cf[2] == (REAL)1 && cf[3] == (REAL)1 && cf[4] == (REAL)1 &&
                    cf[5] == (REAL)1 && cf[6] == (REAL)6;
  const int kst_c = G.kst + G.g - 1, ked_c = G.ked + G.g - 1;
#ifndef _REAL_IS_DOUBLE_
  // packed-f32x2 kernel (fused_pk.cuh): the CubeZ operator only; needs 16-byte aligned units
  const bool packed = unit && cfg >= 2 && (G.nk % 4 == 0);
  const bool lin = packed && mode != MODE_RB2;
#else
  const bool packed = false;
  const bool lin = false;
#endif
  if (packed) {
    const int npairs = (ked_c >> 1) - ((kst_c >> 1) & ~1) + 1;     // tiles start at an even pair
    a.ntk = (npairs + PR - 5) / (PR - 4);
  } else {
    const int npairs = (ked_c >> 1) - (kst_c >> 1) + 1;
    a.ntk = (npairs + PR - 3) / (PR - 2);
  }
  // part: 0 = every j-chunk; 1 = the first and the last chunk (the planes a j-slab's neighbours wait for);
  // 2 = the chunks between.  1 and 2 together launch exactly the CTAs of 0 (chunks are independent of each other:
  // each recomputes the reds next to it), so a slab can relax its edges on one stream and its interior on another.
  const int jlo = G.jst, jhi = G.jed;
  const int ni = G.ied - G.ist + 1, nj = jhi - jlo + 1;
  a.nti = lin ? (ni + a.ei - 3) / (a.ei - 2) : (ni + a.ei - 5) / (a.ei - 4);
  // j-chunks.  Every CTA marches one chunk of jl planes and pays a fixed price per chunk: the pipeline fill (3 extra
  // steps for the two-colour sweep, 0 for the one-plane kernels) plus ~6 steps' worth of start-up (barrier set-up, first
  // TMA round trip, reduction epilogue).  With R = tiles * njc / 148 rounds over the SMs the launch takes about
  //     (R <= 1 ? 1 : R + 0.5) * (jl + fill + 6)   steps
  // (one round exactly while every CTA is resident at once; beyond that the hardware scheduler evens the tail out to
  // about half a round).  The chunk count that minimises it is chosen, >= 8 planes per chunk; a j-slab that splits
  // its sweep into an edge and an interior launch (part != 0) needs at least three chunks.  Measured on B200
  // (profiles/r02_chunks.md): 1024^3 -> 4-5 chunks, 512^3 -> 6, 256^3 -> 4, a 128-plane slab of an 8-GPU run -> 3.
  int njc = 1;
  {
    const long long tiles = (long long)a.ntk * a.nti;
    const int fill = (mode == MODE_RB2) ? 3 : 0;
    double best = -1.0;
    const int nmax = nj / 8 > 1 ? (nj / 8 < 256 ? nj / 8 : 256) : 1;
    const int nmin = (part != 0 && nmax >= 3) ? 3 : 1;
    for (int c = nmin; c <= nmax; ++c) {
      const int jl = (nj + c - 1) / c;
      const int cc = (nj + jl - 1) / jl;                       // chunks really needed with this length
      if (cc < nmin) continue;
      const double r = (double)(tiles * cc) / 148.0;
      const double cost = (r <= 1.0 ? 1.0 : r + 0.5) * (double)(jl + fill + 6);
      if (best < 0.0 || cost < best) { best = cost; njc = cc; }
    }
  }
  { static const char* e = getenv("CZB_NJC"); if (e) njc = atoi(e); }   // experiments
  if (njc < 1) njc = 1;
  a.jl = (nj + njc - 1) / njc;
  a.njc = (nj + a.jl - 1) / a.jl;
  a.jc_first = 0; a.jc_step = 1; a.ngy = a.njc;
  if (part == 1 && a.njc >= 2) { a.jc_step = a.njc - 1; a.ngy = 2; }
  if (part == 2) { a.jc_first = 1; a.ngy = a.njc - 2; }
  if ((size_t)a.ntk * a.nti * a.njc > L.rb_blocks) return 0;
  a.jr0 = jr0; a.jr1 = jr1; a.jb0 = jlo; a.jb1 = jhi;
  a.ip = ip;
  a.c = Coef7{cf[0], cf[1], cf[2], cf[3], cf[4], cf[5], cf[6]};
  a.omg = omg;
  a.out = out; a.w = (mode == MODE_AX) ? b_or_w : nullptr;
  a.st = L.st; a.acc_st = L.acc_st; a.slot = slot; a.add = add ? 1 : 0;
  a.rb = L.rb;
  a.fin = L.fin;
  a.slot2 = lin ? slot2 : -1;
  a.zsrc = (zsrc && packed && !lin) ? 1 : 0;
  if (zsrc && !a.zsrc) return 0;   // only the packed two-colour kernel can skip the source
  if (dry) return 1;               // eligibility probe: every early-out above has been passed, nothing is launched
  if (a.ngy <= 0) return 3;        // part 2 of a slab with one or two chunks: nothing between the edge chunks
  const size_t T = (size_t)a.ei * EKB;
  const CUtensorMap& tp = tmap_for(p, G, a.ei);
  const CUtensorMap& tb = (mode == MODE_AX) ? tp : tmap_for(b_or_w, G, a.ei);
  // + 1 KB: halo columns read (and discard) neighbours up to one half-plane row past the last plane
  const size_t smem = ((mode == MODE_AX ? NSP : NSP + NSB) + 3) * T * sizeof(REAL) + 8 * (NSP + NSB) + 1024;
#ifndef _REAL_IS_DOUBLE_
  if (lin) {
    if (mode == MODE_JAC) launch_lin_pk_t<MODE_JAC, 512, 2, 5, 4>(L, a, p, b_or_w);
    else                  launch_lin_pk_t<MODE_AX, 512, 2, 6, 0>(L, a, p, b_or_w);
    return (mode == MODE_AX && slot2 >= 0) ? 2 : 1;
  }
  if (packed) {   // look-ahead (raw TMA slots for p, b): cfg 2 = 3+3, 3 = 4+3, 4 = 5+4, 5 = 5+5
    if (cfg == 3)      launch_rb2_pk_t<512, 2, 4, 3>(L, a, p, b_or_w);
    else if (cfg == 4) launch_rb2_pk_t<512, 2, 5, 4>(L, a, p, b_or_w);
    else if (cfg == 5) launch_rb2_pk_t<512, 2, 5, 5>(L, a, p, b_or_w);
    else               launch_rb2_pk_t<512, 2, 3, 3>(L, a, p, b_or_w);
    return 1;
  }
#endif
  if (mode == MODE_RB2)      launch_march_cfg<MODE_RB2>(cfg, L, a, tp, tb, unit, smem);
  else if (mode == MODE_JAC) launch_march_cfg<MODE_JAC>(cfg, L, a, tp, tb, unit, smem);
  else                       launch_march_cfg<MODE_AX>(cfg, L, a, tp, tb, unit, smem);
  return 1;
}

}  // namespace

bool fused_eligible(const Geo& G) {
  // TMA needs 16-byte row pitch; pair columns need an even row length
  if (((size_t)G.nk * sizeof(REAL)) % 16 != 0 || (G.nk & 1)) return false;
  if (G.g < 2) return false;
  if (G.ked < G.kst || G.ied < G.ist || G.jed < G.jst) return false;
  return true;
}

bool launch_rb2_fused(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b, const REAL* cf, REAL omg,
                      int ip, int jr0, int jr1, int slot, bool add, bool zero_src, int part) {
  return march(MODE_RB2, L, G, p, b, pnew, cf, omg, ip, jr0, jr1, slot, add, -1, zero_src, part) != 0;
}
// Would launch_rb2_fused with these options launch?  (Same early-outs as march(), nothing enqueued.)  The solve
// loops decide ONCE, before choosing the fused / overlapped loop, so an iteration can never become a silent no-op.
bool rb2_fused_launchable(const Launch& L, const Geo& G, const REAL* cf, bool zero_src, int part) {
  return march(MODE_RB2, L, G, nullptr, nullptr, nullptr, cf, (REAL)1, 0, G.jst, G.jed, 0, false, -1, zero_src, part, true) != 0;
}
bool rb2_zero_src_supported(const Geo& G, const REAL* cf) {
#ifdef _REAL_IS_DOUBLE_
  (void)G; (void)cf; return false;
#else
  const bool unit = cf[0] == 1.f && cf[1] == 1.f && cf[2] == 1.f && cf[3] == 1.f && cf[4] == 1.f && cf[5] == 1.f && cf[6] == 6.f;
  return fused_eligible(G) && unit && march_cfg() >= 2 && (G.nk % 4 == 0);
#endif
}
bool jacobi_fused_launchable(const Launch& L, const Geo& G, const REAL* cf, int part) {
  return march(MODE_JAC, L, G, nullptr, nullptr, nullptr, cf, (REAL)1, 0, G.jst, G.jed, 0, false, -1, false, part, true) != 0;
}
bool launch_jacobi_fused(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b, const REAL* cf,
                         REAL omg, int slot, bool add, int part) {
  return march(MODE_JAC, L, G, p, b, pnew, cf, omg, 0, G.jst, G.jed, slot, add, -1, false, part) != 0;
}
int launch_ax_fused(const Launch& L, const Geo& G, REAL* ap, const REAL* p, const REAL* cf, const REAL* dotwith,
                    int slot, int slot2, int part) {
  return march(MODE_AX, L, G, p, dotwith, ap, cf, (REAL)0, 0, G.jst, G.jed, slot, false, slot2, false, part);
}

}  // namespace czb

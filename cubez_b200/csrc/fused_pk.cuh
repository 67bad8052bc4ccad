// cubez_b200/csrc/fused_pk.cuh -- packed-f32x2 plane-marching kernels (float build only).
// Included by fused.cu inside namespace czb::<anonymous>, after its PTX helpers.
//
// Same algorithm, data flow and arithmetic as k_march (fused.cu) -- TMA-staged j-planes, both
// red-black colours per pass, results bit-identical per cell -- but organised so that one
// instruction works on TWO cells:
//   * a thread owns "units" of 4 consecutive k cells  A0 B0 A1 B1  (A = even k, B = odd k); the
//     same-colour cells of a unit travel as one 64-bit register pair  PA = (A0,A1), PB = (B0,B1)
//     and are relaxed by Blackwell's packed FP32 instructions (add/mul/fma.rn.f32x2 -> SASS
//     FADD2/FMUL2/FFMA2: IEEE round-to-nearest per lane, subnormals kept);
//   * the i+-1 neighbours of a pair are ONE 64-bit shared load from the de-interleaved plane
//     copy, the k+-1 neighbours are the thread's own other-colour pair plus one warp shuffle;
//   * raw TMA planes are read with 128-bit loads, new planes stored with 128-bit stores;
//   * the rare-operand test of the exact division by 6 is done once per group of 4 cells on
//     max/min/or of the numerators.
// A warp covers one full tile row (32 units = 128 cells), NT/32 rows per pass, CPT passes.
// Tile halo: 2 pairs (one unit) on each k side, 2 rows on each i side.
//
// Reference arithmetic: psor2sma_core cz_solver.f90:404-493 (ss order :339-344 / :468-473,
// dp :474, update :475, residual :476).

typedef unsigned long long F2;   // two packed floats: low word = first cell, high word = second

__device__ __forceinline__ F2 pk(float x, float y) { return (F2)__float_as_uint(x) | ((F2)__float_as_uint(y) << 32); }
__device__ __forceinline__ float lo(F2 v) { return __uint_as_float((unsigned)v); }
__device__ __forceinline__ float hi(F2 v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ F2 add2(F2 a, F2 b) { F2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ F2 sub2(F2 a, F2 b) { F2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ F2 mul2(F2 a, F2 b) { F2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ F2 fma2(F2 a, F2 b, F2 c) { F2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// bitwise helpers on the two lanes of a pair: (v & m) and (m ? a : b)
__device__ __forceinline__ F2 and2(F2 v, uint2 m) {
  return pk(__uint_as_float(__float_as_uint(lo(v)) & m.x), __uint_as_float(__float_as_uint(hi(v)) & m.y));
}
__device__ __forceinline__ F2 sel2(F2 a, F2 b, uint2 m) {
  const unsigned a0 = __float_as_uint(lo(a)), a1 = __float_as_uint(hi(a)), b0 = __float_as_uint(lo(b)), b1 = __float_as_uint(hi(b));
  return pk(__uint_as_float((a0 & m.x) | (b0 & ~m.x)), __uint_as_float((a1 & m.y) | (b1 & ~m.y)));
}

template <int OFF>
__device__ __forceinline__ F2 lds64(uint32_t addr) {
  F2 v;
  asm volatile("ld.shared.b64 %0, [%1+%2];" : "=l"(v) : "r"(addr), "n"(OFF) : "memory");
  return v;
}
template <int OFF>
__device__ __forceinline__ void sts64(uint32_t addr, F2 v) {
  asm volatile("st.shared.b64 [%0+%1], %2;" ::"r"(addr), "n"(OFF), "l"(v) : "memory");
}
struct U4 { F2 a, b; };   // one unit of one plane: a = (A0,A1) even-k cells, b = (B0,B1) odd-k cells
__device__ __forceinline__ U4 lds_unit(uint32_t addr) {   // 16 B of a raw (interleaved) plane: A0 B0 A1 B1
  float a0, b0, a1, b1;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a0), "=f"(b0), "=f"(a1), "=f"(b1) : "r"(addr) : "memory");
  return U4{pk(a0, a1), pk(b0, b1)};
}
__device__ __forceinline__ void stg_unit_if(float* p, U4 v, unsigned flag) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %5, 0;\n\t@p st.global.v4.f32 [%0], {%1, %2, %3, %4};\n\t}" ::"l"(p),
               "f"(lo(v.a)), "f"(lo(v.b)), "f"(hi(v.a)), "f"(hi(v.b)), "r"(flag) : "memory");
}
// Exact a/6 for two cells (see div_fast in fused.cu: scaled Markstein sequence, exhaustively
// verified for biased exponents 0, 1 and 4..183).
__device__ __forceinline__ F2 div6_2(F2 a) {
  const F2 S = pk(0x1p64f, 0x1p64f), SI = pk(0x1p-64f, 0x1p-64f);
  const F2 Y = pk(1.0f / 6.0f, 1.0f / 6.0f), M6 = pk(-6.0f, -6.0f);
  const F2 as = mul2(a, S);
  const F2 q = mul2(as, Y);
  const F2 r = fma2(M6, q, as);
  return mul2(fma2(r, Y, q), SI);
}
// Patch the (rare) cells whose numerator is outside the verified range with the IEEE division.
// Out of line (one call per group of 4 cells) so that the hot loop stays small.
__device__ __noinline__ float4 div6_fix4(float4 n, float4 q) {
  if (div_guard_fail(n.x)) q.x = div_rare(n.x, 6.0f);
  if (div_guard_fail(n.y)) q.y = div_rare(n.y, 6.0f);
  if (div_guard_fail(n.z)) q.z = div_rare(n.z, 6.0f);
  if (div_guard_fail(n.w)) q.w = div_rare(n.w, 6.0f);
  return q;
}

// ss = p(i+1) + p(i-1) + p(j+1) + p(j-1) + p(k+1) + p(k-1) in exactly this order (cz_solver.f90:339-344,
// :468-473; cz_blas.f90:626-631) for the two X cells (XB false: A0,A1 ; true: B0,B1) of one unit.
// an = shared address of the unit in the de-interleaved copy of the plane, wc = the plane (own
// other-colour pairs are the k neighbours, plus one shuffle), jn / js = the X pairs of planes j+1 / j-1.
template <bool XB, int HPB_, int PRB_>
__device__ __forceinline__ F2 sum6_pk(uint32_t an, const U4& wc, F2 jn, F2 js) {
  constexpr int XO = XB ? HPB_ : 0;
  const F2 ie = lds64<XO + PRB_>(an), iw = lds64<XO - PRB_>(an);
  F2 ss = add2(add2(add2(ie, iw), jn), js);
  if (!XB) {
    // cells (A0,A1): k+1 = own (B0,B1) ; k-1 = (B1 of the unit to the left, own B0)
    const float bl = __shfl_up_sync(0xffffffffu, hi(wc.b), 1);
    ss = add2(ss, wc.b);
    ss = pk(lo(ss) + bl, hi(ss) + lo(wc.b));
  } else {
    // cells (B0,B1): k+1 = (own A1, A0 of the unit to the right) ; k-1 = own (A0,A1)
    const float ar = __shfl_down_sync(0xffffffffu, lo(wc.a), 1);
    ss = pk(lo(ss) + hi(wc.a), hi(ss) + ar);
    ss = add2(ss, wc.a);
  }
  return ss;
}

constexpr int PKH = 2;          // halo pairs on each k side of a tile (one unit)
constexpr int NDE = 4;          // de-interleaved plane copies kept in shared memory (planes s-1 .. s+2)

__device__ __forceinline__ void mbar_wait_a(uint32_t addr, uint32_t parity) {   // bounded wait, 32-bit smem address
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

// Both colours of one red-black SOR iteration, unit coefficients / dd = 6 (the CubeZ operator).
//
// Software pipeline: at step u (s = first plane + u) the CTA relaxes the RED cells of plane s+1 and the
// BLACK cells of plane s-1.  The blacks of plane s-1 need the new reds of planes s-2, s-1, s -- all
// finished in earlier steps -- so the two relaxations of a step are independent instruction
// streams (the blacks trail the reds by two planes instead of one), which doubles the ILP of the
// dependent add/divide chains.  Per step: one TMA plane of p and one of b arrive (NSP_/NSB_ slots
// of look-ahead), one plane is stored, one __syncthreads.
template <int NT, int CPT, int NSP_, int NSB_>
__global__ void __launch_bounds__(NT, 1)
k_rb2_pk(const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_b, const MarchArgs a) {
  constexpr int RPW = NT / 32;             // tile rows covered by one pass of the CTA (one warp per row)
  constexpr int EI = RPW * CPT;            // tile rows
  constexpr int T = EI * EKB;              // cells per tile plane
  constexpr int TB = T * 4;                // bytes per tile plane
  constexpr int HPB = TB / 2;              // bytes per de-interleaved half plane (even-k, then odd-k)
  constexpr int PRB = PR * 4;              // bytes per row of a half plane
  static_assert(RPW % 2 == 0, "all units of a thread must share the row parity");
  static_assert(CPT == 2, "flag layout below assumes two units per thread");
  const Geo& G = a.G;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t sm0 = smem_u32(smem_raw);
  const uint32_t raw_p = sm0;                          // [NSP_][TB]
  const uint32_t raw_b = raw_p + NSP_ * TB;            // [NSB_][TB]
  const uint32_t dei = raw_b + NSB_ * TB;              // [NDE][TB]
  const uint32_t bar0 = dei + NDE * TB + 512;          // [NSP_ + NSB_] mbarriers (after a pad row)
  const int tid = threadIdx.x;
  const int lane = tid & 31, wrow = tid >> 5;

  // ---- tile / chunk geometry ------------------------------------------------------------
  // Chunk of this CTA: jc_first + blockIdx.y * jc_step.  A j-slab of a multi-GPU run launches its first and last
  // chunk (jc_step = njc - 1) on a high-priority stream and the chunks between (jc_first = 1) on the main stream, so
  // that the slab's edge planes can travel to the neighbours while the interior is still being relaxed (solver.cu).
  const int jc = a.jc_first + (int)blockIdx.y * a.jc_step;
  if (a.st && a.st->done) return;
  const int tk = blockIdx.x % a.ntk, ti = blockIdx.x / a.ntk;
  const int kst_c = G.kst + G.g - 1, ist_c = G.ist + G.g - 1;     // C offsets of the first inner cell
  const int P0 = (((kst_c >> 1) & ~1) - PKH) + tk * (PR - 2 * PKH);   // first (halo) pair of the box, even
  const int R0 = ist_c - 2 + ti * (EI - 4);                        // first (halo) row of the box
  // owned planes of this chunk
  const int ja = a.jb0 + jc * a.jl;
  const int jb = min(ja + a.jl - 1, a.jb1);
  const int pl0 = ja - 2;                                          // first p plane fetched
  const int bpl0 = ja - 1;                                         // first b plane fetched
  const int nsteps = jb - ja + 4;                                  // s = ja-2 .. jb+1
  const int npl = nsteps + 1;                                      // p planes ja-2 .. jb+2
  const int nbpl = nsteps - 1;                                     // b planes ja-1 .. jb+1

  if (tid == 0) {
    for (int s = 0; s < NSP_ + NSB_; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u * s) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue_plane = [&](uint32_t dst, uint32_t bar, const CUtensorMap* tm, int plane) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)TB) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(tm), "r"(bar), "r"(2 * P0), "r"(R0), "r"(plane) : "memory");
  };
  auto issue_p = [&](int t) {   // plane pl0+t -> slot t % NSP_
    issue_plane(raw_p + (uint32_t)(t % NSP_) * TB, bar0 + 8u * (uint32_t)(t % NSP_), &tm_p, pl0 + t + G.g - 1);
  };
  auto issue_b = [&](int t) {
    issue_plane(raw_b + (uint32_t)(t % NSB_) * TB, bar0 + 8u * (uint32_t)(NSP_ + t % NSB_), &tm_b, bpl0 + t + G.g - 1);
  };
  // a.zsrc: the source field is known to be all zero (first preconditioner sweep, cz_Poisson.cpp:405-409):
  // its planes are not fetched at all.
  const bool zsrc = a.zsrc != 0;
  if (tid == 0) {
    if (!zsrc) for (int t = 0; t < NSP_ && t < npl; ++t) issue_p(t);
    for (int t = 0; t < NSB_ && t < nbpl; ++t) issue_b(t);
  }

  // ---- per-thread constants ---------------------------------------------------------------
  // Flag word: bit x*4+m*2+e     : the X cell (x = 0 A / 1 B) of pair e of unit m may be relaxed as a RED
  //                                cell here (inner cell, row and k neighbours inside the box);
  //            bit 8+x*4+m*2+e   : that cell is owned by this tile (blacks relaxed, residual counted);
  //            bit 16+m          : unit m is stored.
  // Masks are applied with selects, so whatever the never-referenced guard cells hold (even NaN)
  // cannot leak into a kept value or the residual.
  const float omg = a.omg;
  const F2 omg2 = pk(omg, omg);
  const int kc0 = 2 * (P0 + 2 * lane);            // C index of cell A0 of the thread's units
  const bool unit_in = kc0 >= 0 && kc0 + 3 < G.nk;
  // parity of cell A0 at plane 0 (Fortran indices, cz_solver.f90:466): all units of a thread share it
  const unsigned par = (unsigned)((kc0 - G.g + 1) - G.kst + (R0 + wrow - G.g + 1) + a.ip) & 1u;
  uint32_t araw[CPT], ad[CPT];
  unsigned fl = 0u;
  float* outp[CPT];
#pragma unroll
  for (int m = 0; m < CPT; ++m) {
    const int r = wrow + RPW * m;
    araw[m] = (uint32_t)(r * EKB + 4 * lane) * 4u;
    ad[m] = (uint32_t)(r * PR + 2 * lane) * 4u;
    const int ic = R0 + r, iF = ic - G.g + 1;
    const bool rowin = ic >= 0 && ic < G.ni && iF >= G.ist && iF <= G.ied;
    const int dr = min(r, EI - 1 - r);
    const bool rR = rowin && dr >= 1, rO = rowin && dr >= 2;
    bool kany = false;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int q = 2 * lane + e;
      const bool ownq = (q >= PKH) && (q <= PR - 1 - PKH);
#pragma unroll
      for (int x = 0; x < 2; ++x) {
        const int kc = 2 * (P0 + q) + x, kF = kc - G.g + 1;
        const bool kin = kc >= 0 && kc < G.nk && kF >= G.kst && kF <= G.ked;
        const bool nb = x ? (q <= PR - 2) : (q >= 1);
        if (kin && nb && rR) fl |= 1u << (x * 4 + m * 2 + e);
        if (kin && ownq && rO) fl |= 1u << (8 + x * 4 + m * 2 + e);
        kany = kany || (kin && ownq);
      }
    }
    const bool st = rO && kany && unit_in;
    if (st) fl |= 1u << (16 + m);
    // running store pointer: unit (r, lane) of plane s-1 (advanced by one plane per step)
    outp[m] = a.out + ((long long)(pl0 - 1 + G.g - 1) * G.plane + (long long)G.nk * (st ? ic : 0) + (st ? kc0 : 0));
  }

  U4 P[4][CPT];          // planes s-1, s, s+1, s+2 (roles rotate with the step, see STEP below)
  F2 hb[2][CPT];         // right-hand side of the black cells of planes s-1 and s
#pragma unroll
  for (int m = 0; m < CPT; ++m) {
#pragma unroll
    for (int c = 0; c < 4; ++c) P[c][m] = U4{0ull, 0ull};
    hb[0][m] = hb[1][m] = 0ull;
  }
  double acc[1] = {0.0};

  // ---- prologue: planes pl0 (-> P[0], "s") and pl0+1 (-> P[1], "s+1", + de-interleaved copy) ---------
  if (!zsrc) {
    mbar_wait_a(bar0, 0);
#pragma unroll
    for (int m = 0; m < CPT; ++m) P[0][m] = lds_unit(raw_p + araw[m]);
    mbar_wait_a(bar0 + 8, 0);
#pragma unroll
    for (int m = 0; m < CPT; ++m) P[1][m] = lds_unit(raw_p + TB + araw[m]);
  }
#pragma unroll
  for (int m = 0; m < CPT; ++m) { sts64<0>(dei + TB + ad[m], P[1][m].a); sts64<HPB>(dei + TB + ad[m], P[1][m].b); }
  __syncthreads();
  if (tid == 0 && !zsrc) { if (NSP_ < npl) issue_p(NSP_); if (NSP_ + 1 < npl) issue_p(NSP_ + 1); }

  // ---- main loop ----------------------------------------------------------------------------------
  int sp = 2 % NSP_, pp = (2 / NSP_) & 1;   // raw p slot / mbarrier parity of plane index u+2
  int sb = 0, pb = 0;                       // raw b slot / parity of plane index u

  // Numerators ss - b of the X cells (xb: false = A, true = B) of one plane for all units of the thread.
  // wc = the plane (own other-colour pairs give the k neighbours, plus one shuffle), en = its
  // de-interleaved copy (i neighbours), jn/js = the X pairs of planes j+1 / j-1, bb = right-hand side.
  // Order of the sum: cz_solver.f90:468-473 (i+1, i-1, j+1, j-1, k+1, k-1).
  auto numer = [&](auto xb, const uint32_t en, const U4 (&wc)[CPT], const F2 (&jn)[CPT], const F2 (&js)[CPT],
                   const F2 (&bb)[CPT], F2 (&num)[CPT]) {
    constexpr bool XB = decltype(xb)::value;
#pragma unroll
    for (int m = 0; m < CPT; ++m) num[m] = sub2(sum6_pk<XB, HPB, PRB>(en + ad[m], wc[m], jn[m], js[m]), bb[m]);
  };
  auto bit = [&](unsigned f, int b) -> bool { return ((f >> b) & 1u) != 0u; };
  auto sel = [&](bool c, F2 x, F2 y, bool c2) -> F2 {   // (c ? x.lo : y.lo, c2 ? x.hi : y.hi)
    return pk(c ? lo(x) : lo(y), c2 ? hi(x) : hi(y));
  };

  // One step.  C = u % 4 (which register set holds which plane), XB = the red cells of plane s+1 (and the
  // black cells of plane s-1... of the OTHER position) -- both compile-time, see the unrolled loop below.
  // RO ("reds only"): the first three steps of a chunk have no black plane yet (s-1 < ja); their instantiation
  // leaves the black half out instead of computing it under all-zero flags (2 of the 46 steps of a 43-plane chunk).
  auto step = [&](int u, auto cc, auto xb, auto ro) {
    constexpr int C = decltype(cc)::value;
    constexpr bool XB = decltype(xb)::value;       // position of the reds of plane s+1; blacks of plane s-1 are !XB
    constexpr bool RO = decltype(ro)::value;
    constexpr int XO = XB ? HPB : 0, YO = XB ? 0 : HPB;
    U4 (&wm1)[CPT] = P[(C + 3) & 3];
    U4 (&w0)[CPT] = P[C];
    U4 (&w1)[CPT] = P[(C + 1) & 3];
    U4 (&w2)[CPT] = P[(C + 2) & 3];
    const int s = pl0 + u;
    const uint32_t em1 = dei + (uint32_t)((C + 3) & 3) * TB, e1 = dei + (uint32_t)((C + 1) & 3) * TB,
                   e2 = dei + (uint32_t)((C + 2) & 3) * TB;
    // j-1 neighbours of the blacks of plane s-1: reds of plane s-2, still in the registers that
    // receive plane s+2 below
    F2 jsk[CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) jsk[m] = XB ? w2[m].a : w2[m].b;
    if (u + 2 < npl) {
      if (!zsrc) {
        mbar_wait_a(bar0 + 8u * (uint32_t)sp, (uint32_t)pp);
        const uint32_t rp = raw_p + (uint32_t)sp * TB;
#pragma unroll
        for (int m = 0; m < CPT; ++m) w2[m] = lds_unit(rp + araw[m]);
      } else {
#pragma unroll
        for (int m = 0; m < CPT; ++m) w2[m] = U4{0ull, 0ull};
      }
#pragma unroll
      for (int m = 0; m < CPT; ++m) { sts64<0>(e2 + ad[m], w2[m].a); sts64<HPB>(e2 + ad[m], w2[m].b); }
    }
    U4 bc[CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) bc[m] = U4{0ull, 0ull};
    if (u < nbpl) {
      mbar_wait_a(bar0 + 8u * (uint32_t)(NSP_ + sb), (uint32_t)pb);
      const uint32_t rbp = raw_b + (uint32_t)sb * TB;
#pragma unroll
      for (int m = 0; m < CPT; ++m) bc[m] = lds_unit(rbp + araw[m]);
    }
    // step-uniform activity: reds of plane s+1, blacks of plane s-1
    const bool do_red = (s + 1 >= a.jr0 && s + 1 <= a.jr1);
    const bool red_owned = (s + 1 >= ja && s + 1 <= jb);
    const bool do_black = !RO && (s - 1 >= ja && s - 1 <= jb);
    const unsigned fR = do_red ? (XB ? (fl >> 4) : fl) & 0xfu : 0u;                 // bit m*2+e
    const unsigned fRO = (do_red && red_owned) ? (XB ? (fl >> 12) : (fl >> 8)) & 0xfu : 0u;
    const unsigned fK = do_black ? (XB ? (fl >> 8) : (fl >> 12)) & 0xfu : 0u;      // blacks sit at !XB
    const unsigned fS = do_black ? (fl >> 16) & 0x3u : 0u;

    F2 numR[CPT], numK[CPT], jnR[CPT], jsR[CPT], bbR[CPT], jnK[CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) {
      jnR[m] = XB ? w2[m].b : w2[m].a;  jsR[m] = XB ? w0[m].b : w0[m].a;  bbR[m] = XB ? bc[m].b : bc[m].a;
      jnK[m] = XB ? w0[m].a : w0[m].b;
    }
    numer(xb, e1, w1, jnR, jsR, bbR, numR);
    if constexpr (!RO) numer(std::integral_constant<bool, !XB>{}, em1, wm1, jnK, jsk, hb[C & 1], numK);
    else {
#pragma unroll
      for (int m = 0; m < CPT; ++m) numK[m] = numR[m];     // keeps the range test below one code path
    }
    F2 quoR[CPT], quoK[CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) { quoR[m] = div6_2(numR[m]); quoK[m] = RO ? quoR[m] : div6_2(numK[m]); }
    {  // numerators outside the verified range of the fast division (tiny, huge, -0, non-finite): rare
      float mx = 0.0f, mn = 3.0e38f;
      unsigned orb = 0u;
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        const float n0 = lo(numR[m]), n1 = hi(numR[m]), n2 = lo(numK[m]), n3 = hi(numK[m]);
        mx = fmaxf(fmaxf(mx, fmaxf(fabsf(n0), fabsf(n1))), fmaxf(fabsf(n2), fabsf(n3)));
        mn = fminf(fminf(mn, fminf(fabsf(n0), fabsf(n1))), fminf(fabsf(n2), fabsf(n3)));
        orb |= __float_as_uint(n0) | __float_as_uint(n1) | __float_as_uint(n2) | __float_as_uint(n3);
      }
      if (!((mn >= 0x1p-123f && mx < 0x1p57f) || orb == 0u)) {
        const float4 f = div6_fix4(make_float4(lo(numR[0]), hi(numR[0]), lo(numR[1]), hi(numR[1])),
                                   make_float4(lo(quoR[0]), hi(quoR[0]), lo(quoR[1]), hi(quoR[1])));
        quoR[0] = pk(f.x, f.y); quoR[1] = pk(f.z, f.w);
        if constexpr (!RO) {
          const float4 h = div6_fix4(make_float4(lo(numK[0]), hi(numK[0]), lo(numK[1]), hi(numK[1])),
                                     make_float4(lo(quoK[0]), hi(quoK[0]), lo(quoK[1]), hi(quoK[1])));
          quoK[0] = pk(h.x, h.y); quoK[1] = pk(h.z, h.w);
        }
      }
    }
    // A cell that must not move gets dp = -0.0: p + (-0) == p bit for bit (also for p = +-0) and
    // (-0)^2 adds nothing to the residual, so one select per cell serves both.
    const F2 nz2 = pk(-0.0f, -0.0f);
#pragma unroll
    for (int m = 0; m < CPT; ++m) {
      // reds of plane s+1:  dp = ((ss-b)/dd - p)*omg ; p += dp   (cz_solver.f90:474-476)
      const F2 ppR = XB ? w1[m].b : w1[m].a;
      const F2 dpR = mul2(sub2(quoR[m], ppR), omg2);
      const F2 nvR = add2(ppR, sel(bit(fR, m * 2), dpR, nz2, bit(fR, m * 2 + 1)));
      if (XB) w1[m].b = nvR; else w1[m].a = nvR;
      sts64<XO>(e1 + ad[m], nvR);
      const F2 dkR = sel(bit(fRO, m * 2), dpR, nz2, bit(fRO, m * 2 + 1));
      const F2 d2R = mul2(dkR, dkR);
      acc[0] += (double)lo(d2R);
      acc[0] += (double)hi(d2R);
      if constexpr (!RO) {
        // blacks of plane s-1
        const F2 ppK = XB ? wm1[m].a : wm1[m].b;
        const F2 dkK = sel(bit(fK, m * 2), mul2(sub2(quoK[m], ppK), omg2), nz2, bit(fK, m * 2 + 1));
        const F2 nvK = add2(ppK, dkK);
        if (XB) wm1[m].a = nvK; else wm1[m].b = nvK;
        const F2 d2K = mul2(dkK, dkK);
        acc[0] += (double)lo(d2K);
        acc[0] += (double)hi(d2K);
        stg_unit_if(outp[m], wm1[m], (fS >> m) & 1u);
      }
      outp[m] += G.plane;
      hb[C & 1][m] = XB ? bc[m].a : bc[m].b;     // right-hand side of the blacks of plane s+1 (used at step u+2)
    }
    (void)YO;
    __syncthreads();
    if (tid == 0) {   // the two raw slots read in this step are free: refill them
      if (!zsrc && u + 2 + NSP_ < npl) issue_p(u + 2 + NSP_);
      if (u + NSB_ < nbpl) issue_b(u + NSB_);
    }
    if (++sp == NSP_) { sp = 0; pp ^= 1; }
    if (++sb == NSB_) { sb = 0; pb ^= 1; }
  };
  // A is the red cell of plane s+1 iff (par + s + 1) is even (warp-uniform); it alternates with the step,
  // so the 4x unrolled body exists in two variants (A-B-A-B and B-A-B-A) chosen once.
  using I0 = std::integral_constant<int, 0>; using I1 = std::integral_constant<int, 1>;
  using I2 = std::integral_constant<int, 2>; using I3 = std::integral_constant<int, 3>;
  auto run = [&](auto x0) {                       // X0 = position of the reds in step 0
    constexpr bool X0 = decltype(x0)::value;
    using TX = std::integral_constant<bool, X0>; using TY = std::integral_constant<bool, !X0>;
    // nsteps = jl + 3 >= 4: the first three steps relax reds only
    step(0, I0{}, TX{}, std::true_type{});
    step(1, I1{}, TY{}, std::true_type{});
    step(2, I2{}, TX{}, std::true_type{});
    step(3, I3{}, TY{}, std::false_type{});
    for (int u = 4; u < nsteps; u += 4) {
      step(u, I0{}, TX{}, std::false_type{});
      if (u + 1 >= nsteps) break;
      step(u + 1, I1{}, TY{}, std::false_type{});
      if (u + 2 >= nsteps) break;
      step(u + 2, I2{}, TX{}, std::false_type{});
      if (u + 3 >= nsteps) break;
      step(u + 3, I3{}, TY{}, std::false_type{});
    }
  };
  if (((par + (unsigned)(pl0 + 1)) & 1u) == 0u) run(std::false_type{}); else run(std::true_type{});
  grid_reduce<1>(acc, a.rb, FinStoreF{a.acc_st, a.slot, a.add, a.fin});
}

// ================================================================================================
// One-plane-per-step packed kernels: every cell of plane j from the OLD planes j-1, j, j+1.
//   MODE_JAC : relaxed Jacobi sweep (cz_solver.f90:284-387) without the copy-back pass, residual fused
//   MODE_AX  : ap = ss - dd*p (blas_calc_ax, cz_blas.f90:579-644) fused with the dots (ap, w) and,
//              when a.slot2 >= 0, (ap, ap)  -> PBiCGstab's q = A p^ + (q, r0) and t = A s^ + (t, s) + (t, t)
// Same tile, unit and TMA machinery as k_rb2_pk; tile halo is one row on each i side.
struct FinStore2F {
  LoopState* st; int slot; int slot2;
  __device__ void operator()(double (&tot)[2]) const {
    if (st) { st->acc[slot] = tot[0]; if (slot2 >= 0) st->acc[slot2] = tot[1]; }
  }
};

template <int MODE, int NT, int CPT, int NSP_, int NSB_>
__global__ void __launch_bounds__(NT, 1)
k_lin_pk(const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_b, const MarchArgs a) {
  if (a.st && a.st->done) return;
  constexpr int RPW = NT / 32;
  constexpr int EI = RPW * CPT;
  constexpr int T = EI * EKB;
  constexpr int TB = T * 4;
  constexpr int HPB = TB / 2;
  constexpr int PRB = PR * 4;
  constexpr int NB = (MODE == MODE_JAC) ? NSB_ : 0;
  static_assert(CPT == 2, "flag layout below assumes two units per thread");
  const Geo& G = a.G;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t sm0 = smem_u32(smem_raw);
  const uint32_t raw_p = sm0;                          // [NSP_][TB]
  const uint32_t raw_b = raw_p + NSP_ * TB;            // [NB][TB]
  const uint32_t dei = raw_b + NB * TB;                // [2][TB]
  const uint32_t bar0 = dei + 2 * TB + 512;            // [NSP_ + NB] mbarriers (after a pad row)
  const int tid = threadIdx.x;
  const int lane = tid & 31, wrow = tid >> 5;

  const int tk = blockIdx.x % a.ntk, ti = blockIdx.x / a.ntk, jc = a.jc_first + (int)blockIdx.y * a.jc_step;
  const int kst_c = G.kst + G.g - 1, ist_c = G.ist + G.g - 1;
  const int P0 = (((kst_c >> 1) & ~1) - PKH) + tk * (PR - 2 * PKH);
  const int R0 = ist_c - 1 + ti * (EI - 2);
  const int ja = a.jb0 + jc * a.jl;
  const int jb = min(ja + a.jl - 1, a.jb1);
  const int pl0 = ja - 1;                                          // first p plane fetched
  const int nsteps = jb - ja + 1;                                  // plane j = ja + u
  const int npl = nsteps + 2;

  if (tid == 0) {
    for (int s = 0; s < NSP_ + NB; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u * s) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue_plane = [&](uint32_t dst, uint32_t bar, const CUtensorMap* tm, int plane) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)TB) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(tm), "r"(bar), "r"(2 * P0), "r"(R0), "r"(plane) : "memory");
  };
  auto issue_p = [&](int t) {
    issue_plane(raw_p + (uint32_t)(t % NSP_) * TB, bar0 + 8u * (uint32_t)(t % NSP_), &tm_p, pl0 + t + G.g - 1);
  };
  auto issue_b = [&](int t) {   // right-hand side of plane ja + t
    constexpr int NBD = NSB_ > 0 ? NSB_ : 1;
    issue_plane(raw_b + (uint32_t)(t % NBD) * TB, bar0 + 8u * (uint32_t)(NSP_ + t % NBD), &tm_b, ja + t + G.g - 1);
  };
  if (tid == 0) {
    for (int t = 0; t < NSP_ && t < npl; ++t) issue_p(t);
    if (MODE == MODE_JAC) for (int t = 0; t < NSB_ && t < nsteps; ++t) issue_b(t);
  }

  // flag word: bit x*4+m*2+e = the X cell of pair e of unit m is an inner cell owned by this tile;
  //            bit 8+m = unit m is stored
  const float omg = a.omg;
  const F2 omg2 = pk(omg, omg);
  const int kc0 = 2 * (P0 + 2 * lane);
  const bool unit_in = kc0 >= 0 && kc0 + 3 < G.nk;
  uint32_t araw[CPT], ad[CPT];
  unsigned fl = 0u;
  float* outp[CPT];
  const float* wp[CPT];
#pragma unroll
  for (int m = 0; m < CPT; ++m) {
    const int r = wrow + RPW * m;
    araw[m] = (uint32_t)(r * EKB + 4 * lane) * 4u;
    ad[m] = (uint32_t)(r * PR + 2 * lane) * 4u;
    const int ic = R0 + r, iF = ic - G.g + 1;
    const bool rowin = ic >= 0 && ic < G.ni && iF >= G.ist && iF <= G.ied;
    const bool rO = rowin && r >= 1 && r <= EI - 2;
    bool kany = false;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int q = 2 * lane + e;
      const bool ownq = (q >= PKH) && (q <= PR - 1 - PKH);
#pragma unroll
      for (int x = 0; x < 2; ++x) {
        const int kc = 2 * (P0 + q) + x, kF = kc - G.g + 1;
        const bool kin = kc >= 0 && kc < G.nk && kF >= G.kst && kF <= G.ked;
        if (kin && ownq && rO) fl |= 1u << (x * 4 + m * 2 + e);
        kany = kany || (kin && ownq);
      }
    }
    const bool st = rO && kany && unit_in;
    if (st) fl |= 1u << (8 + m);
    const long long off = (long long)(ja + G.g - 1) * G.plane + (long long)G.nk * (st ? ic : 0) + (st ? kc0 : 0);
    outp[m] = a.out + off;
    wp[m] = (MODE == MODE_AX && a.w) ? a.w + off : nullptr;
  }
  auto bit = [&](int b) -> bool { return ((fl >> b) & 1u) != 0u; };
  auto sel = [&](bool c, F2 x, F2 y, bool c2) -> F2 { return pk(c ? lo(x) : lo(y), c2 ? hi(x) : hi(y)); };

  U4 P[3][CPT];
#pragma unroll
  for (int m = 0; m < CPT; ++m) { P[0][m] = P[1][m] = P[2][m] = U4{0ull, 0ull}; }
  double acc[2] = {0.0, 0.0};
  float4 wv[CPT];      // MODE_AX: dot operand of the plane computed in this step (loaded one step ahead)
#pragma unroll
  for (int m = 0; m < CPT; ++m) wv[m] = make_float4(0.f, 0.f, 0.f, 0.f);

  // prologue: plane ja-1 -> P[0], plane ja -> P[1] + de-interleaved copy (slot 1)
  mbar_wait_a(bar0, 0);
#pragma unroll
  for (int m = 0; m < CPT; ++m) P[0][m] = lds_unit(raw_p + araw[m]);
  mbar_wait_a(bar0 + 8, 0);
#pragma unroll
  for (int m = 0; m < CPT; ++m) {
    P[1][m] = lds_unit(raw_p + TB + araw[m]);
    sts64<0>(dei + TB + ad[m], P[1][m].a); sts64<HPB>(dei + TB + ad[m], P[1][m].b);
    if (MODE == MODE_AX && wp[m] && bit(8 + m)) wv[m] = __ldg(reinterpret_cast<const float4*>(wp[m]));
  }
  __syncthreads();
  if (tid == 0) { if (NSP_ < npl) issue_p(NSP_); if (NSP_ + 1 < npl) issue_p(NSP_ + 1); }

  int sp = 2 % NSP_, pp = (2 / NSP_) & 1;   // raw p slot / parity of plane index u+2
  int sb = 0, pb = 0;                       // raw b slot / parity of plane index u
  auto step = [&](int u, auto cc) {
    constexpr int C = decltype(cc)::value;
    U4 (&w0)[CPT] = P[C];
    U4 (&w1)[CPT] = P[(C + 1) % 3];
    U4 (&w2)[CPT] = P[(C + 2) % 3];
    const uint32_t ecur = dei + (uint32_t)((u + 1) & 1) * TB, enext = dei + (uint32_t)(u & 1) * TB;
    // plane j+1 (always exists: npl = nsteps + 2)
    mbar_wait_a(bar0 + 8u * (uint32_t)sp, (uint32_t)pp);
    {
      const uint32_t rp = raw_p + (uint32_t)sp * TB;
#pragma unroll
      for (int m = 0; m < CPT; ++m) w2[m] = lds_unit(rp + araw[m]);
#pragma unroll
      for (int m = 0; m < CPT; ++m) { sts64<0>(enext + ad[m], w2[m].a); sts64<HPB>(enext + ad[m], w2[m].b); }
    }
    U4 bc[CPT];
    float4 wn[CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) { bc[m] = U4{0ull, 0ull}; wn[m] = make_float4(0.f, 0.f, 0.f, 0.f); }
    if (MODE == MODE_JAC) {
      mbar_wait_a(bar0 + 8u * (uint32_t)(NSP_ + sb), (uint32_t)pb);
      const uint32_t rbp = raw_b + (uint32_t)sb * TB;
#pragma unroll
      for (int m = 0; m < CPT; ++m) bc[m] = lds_unit(rbp + araw[m]);
    } else {
#pragma unroll
      for (int m = 0; m < CPT; ++m)   // dot operand of the NEXT plane
        if (wp[m] && bit(8 + m) && u + 1 < nsteps) wn[m] = __ldg(reinterpret_cast<const float4*>(wp[m] + G.plane));
    }
    F2 ssA[CPT], ssB[CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) {
      ssA[m] = sum6_pk<false, HPB, PRB>(ecur + ad[m], w1[m], w2[m].a, w0[m].a);
      ssB[m] = sum6_pk<true, HPB, PRB>(ecur + ad[m], w1[m], w2[m].b, w0[m].b);
    }
    if (MODE == MODE_JAC) {
      F2 numA[CPT], numB[CPT], quoA[CPT], quoB[CPT];
      float mx = 0.0f, mn = 3.0e38f;
      unsigned orb = 0u;
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        numA[m] = sub2(ssA[m], bc[m].a); numB[m] = sub2(ssB[m], bc[m].b);
        quoA[m] = div6_2(numA[m]); quoB[m] = div6_2(numB[m]);
        const float n0 = lo(numA[m]), n1 = hi(numA[m]), n2 = lo(numB[m]), n3 = hi(numB[m]);
        mx = fmaxf(fmaxf(mx, fmaxf(fabsf(n0), fabsf(n1))), fmaxf(fabsf(n2), fabsf(n3)));
        mn = fminf(fminf(mn, fminf(fabsf(n0), fabsf(n1))), fminf(fabsf(n2), fabsf(n3)));
        orb |= __float_as_uint(n0) | __float_as_uint(n1) | __float_as_uint(n2) | __float_as_uint(n3);
      }
      if (!((mn >= 0x1p-123f && mx < 0x1p57f) || orb == 0u)) {
        const float4 f = div6_fix4(make_float4(lo(numA[0]), hi(numA[0]), lo(numA[1]), hi(numA[1])),
                                   make_float4(lo(quoA[0]), hi(quoA[0]), lo(quoA[1]), hi(quoA[1])));
        quoA[0] = pk(f.x, f.y); quoA[1] = pk(f.z, f.w);
        const float4 h = div6_fix4(make_float4(lo(numB[0]), hi(numB[0]), lo(numB[1]), hi(numB[1])),
                                   make_float4(lo(quoB[0]), hi(quoB[0]), lo(quoB[1]), hi(quoB[1])));
        quoB[0] = pk(h.x, h.y); quoB[1] = pk(h.z, h.w);
      }
      const F2 nz2 = pk(-0.0f, -0.0f);
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        // dp = ((ss-b)/dd - p)*omg ; wk2 = p + dp ; res += dp*dp   (cz_solver.f90:345-348)
        const F2 dA = sel(bit(m * 2), mul2(sub2(quoA[m], w1[m].a), omg2), nz2, bit(m * 2 + 1));
        const F2 dB = sel(bit(4 + m * 2), mul2(sub2(quoB[m], w1[m].b), omg2), nz2, bit(4 + m * 2 + 1));
        const U4 o = U4{add2(w1[m].a, dA), add2(w1[m].b, dB)};
        const F2 qA = mul2(dA, dA), qB = mul2(dB, dB);
        acc[0] += (double)lo(qA); acc[0] += (double)hi(qA);
        acc[0] += (double)lo(qB); acc[0] += (double)hi(qB);
        stg_unit_if(outp[m], o, (fl >> (8 + m)) & 1u);
        outp[m] += G.plane;
      }
    } else {
      const F2 zero2 = 0ull;
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        // ap = ss - dd*p on inner cells (cz_blas.f90:626-632); the other cells of a stored unit are
        // boundary nodes of q / t_, which the reference leaves at their initial 0.
        // dd*p with dd = 6 is formed as (4p) + (2p): doublings are exact, so the single rounding of the
        // sum is RN(6p) -- and there is no multiply that ptxas could contract with the subtraction
        // into an FFMA2 (it does so for mul.rn.f32x2 + sub.rn.f32x2, which changes the rounding).
        const F2 a2 = add2(w1[m].a, w1[m].a), b2 = add2(w1[m].b, w1[m].b);
        const F2 a6 = add2(add2(a2, a2), a2), b6 = add2(add2(b2, b2), b2);
        const F2 vA = sel(bit(m * 2), sub2(ssA[m], a6), zero2, bit(m * 2 + 1));
        const F2 vB = sel(bit(4 + m * 2), sub2(ssB[m], b6), zero2, bit(4 + m * 2 + 1));
        if (wp[m]) {
          const F2 dA = mul2(vA, pk(wv[m].x, wv[m].z)), dB = mul2(vB, pk(wv[m].y, wv[m].w));
          acc[0] += (double)lo(dA); acc[0] += (double)hi(dA);
          acc[0] += (double)lo(dB); acc[0] += (double)hi(dB);
          wp[m] += G.plane;
        }
        if (a.slot2 >= 0) {
          const F2 qA = mul2(vA, vA), qB = mul2(vB, vB);
          acc[1] += (double)lo(qA); acc[1] += (double)hi(qA);
          acc[1] += (double)lo(qB); acc[1] += (double)hi(qB);
        }
        stg_unit_if(outp[m], U4{vA, vB}, (fl >> (8 + m)) & 1u);
        outp[m] += G.plane;
        wv[m] = wn[m];
      }
    }
    __syncthreads();
    if (tid == 0) {
      if (u + 2 + NSP_ < npl) issue_p(u + 2 + NSP_);
      if (MODE == MODE_JAC && u + NSB_ < nsteps) issue_b(u + NSB_);
    }
    if (++sp == NSP_) { sp = 0; pp ^= 1; }
    if (MODE == MODE_JAC) { if (++sb == NSB_) { sb = 0; pb ^= 1; } }
  };
  using I0 = std::integral_constant<int, 0>; using I1 = std::integral_constant<int, 1>;
  using I2 = std::integral_constant<int, 2>;
  for (int u = 0; u < nsteps; u += 3) {
    step(u, I0{});
    if (u + 1 >= nsteps) break;
    step(u + 1, I1{});
    if (u + 2 >= nsteps) break;
    step(u + 2, I2{});
  }
  if (MODE == MODE_JAC) {
    double a1[1] = {acc[0]};
    grid_reduce<1>(a1, a.rb, FinStoreF{a.acc_st, a.slot, a.add, a.fin});
  } else {
    grid_reduce<2>(acc, a.rb, FinStore2F{a.acc_st, a.slot, a.slot2});
  }
}

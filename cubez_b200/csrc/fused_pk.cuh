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

__device__ __forceinline__ F2 pk(float x, float y) { F2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y)); return r; }
__device__ __forceinline__ float lo(F2 v) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v)); return x; }
__device__ __forceinline__ float hi(F2 v) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v)); return y; }
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

constexpr int PKH = 2;          // halo pairs on each k side of a tile (one unit)

// Both colours of one red-black SOR iteration, unit coefficients / dd = 6 (the CubeZ operator).
template <int NT, int CPT>
__global__ void __launch_bounds__(NT, 1)
k_rb2_pk(const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_b, const MarchArgs a) {
  if (a.st && a.st->done) return;
  constexpr int RPW = NT / 32;             // tile rows covered by one pass of the CTA (one warp per row)
  constexpr int EI = RPW * CPT;            // tile rows
  constexpr int T = EI * EKB;              // cells per tile plane
  constexpr int TB = T * 4;                // bytes per tile plane
  constexpr int HPB = TB / 2;              // bytes per de-interleaved half plane (even-k, then odd-k)
  constexpr int PRB = PR * 4;              // bytes per row of a half plane
  static_assert(RPW % 2 == 0, "all units of a thread must share the row parity");
  const Geo& G = a.G;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t sm0 = smem_u32(smem_raw);
  const uint32_t raw_p = sm0;                          // [NSP][TB]
  const uint32_t raw_b = raw_p + NSP * TB;             // [NSB][TB]
  const uint32_t dei = raw_b + NSB * TB;               // [3][TB]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)(NSP + NSB + 3) * TB);   // [NSP + NSB]
  const int tid = threadIdx.x;
  const int lane = tid & 31, wrow = tid >> 5;

  // ---- tile / chunk geometry ------------------------------------------------------------
  const int tk = blockIdx.x % a.ntk, ti = blockIdx.x / a.ntk, jc = blockIdx.y;
  const int kst_c = G.kst + G.g - 1, ist_c = G.ist + G.g - 1;     // C offsets of the first inner cell
  const int P0 = (((kst_c >> 1) & ~1) - PKH) + tk * (PR - 2 * PKH);   // first (halo) pair of the box, even
  const int R0 = ist_c - 2 + ti * (EI - 4);                        // first (halo) row of the box
  const int ja = a.jb0 + jc * a.jl;                                // owned planes of this chunk
  const int jb = min(ja + a.jl - 1, a.jb1);
  const int pl0 = ja - 2;                                          // first p plane fetched
  const int bpl0 = ja - 1;                                         // first b plane fetched
  const int nsteps = jb - ja + 3;
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
    for (int t = 0; t < NSB && t < nbpl; ++t) issue_b(t);
  }

  // ---- per-thread constants ---------------------------------------------------------------
  // Bit masks (all ones / zero) per cell of the thread's units: MR = the red cell may be relaxed here
  // (inner cell, row and k neighbours inside the box), MK = the cell is owned by this tile (blacks are
  // relaxed, residual counted, results stored).  Masks are applied with bitwise selects, so whatever
  // the never-referenced guard cells hold (even NaN) cannot leak into a kept value or the residual.
  const float omg = a.omg;
  const F2 omg2 = pk(omg, omg);
  bool kRb[2][2], kOb[2][2];   // [X = A/B][e]
  bool kany = false;
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int q = 2 * lane + e;
    const int kcA = 2 * (P0 + q);
    const bool ownq = (q >= PKH) && (q <= PR - 1 - PKH);
#pragma unroll
    for (int x = 0; x < 2; ++x) {
      const int kc = kcA + x, kF = kc - G.g + 1;
      const bool kin = kc >= 0 && kc < G.nk && kF >= G.kst && kF <= G.ked;
      const bool nb = x ? (q <= PR - 2) : (q >= 1);
      kRb[x][e] = kin && nb;
      kOb[x][e] = kin && ownq;
      kany = kany || (kin && ownq);
    }
  }
  const int kc0 = 2 * (P0 + 2 * lane);            // C index of cell A0 of the thread's units
  const bool unit_in = kc0 >= 0 && kc0 + 3 < G.nk;
  // parity of cell A0 at plane 0 (Fortran indices, cz_solver.f90:466): all units of a thread share it
  const unsigned par = (unsigned)((kc0 - G.g + 1) - G.kst + (R0 + wrow - G.g + 1) + a.ip) & 1u;
  uint32_t araw[CPT], ad[CPT];
  unsigned stf[CPT];
  uint2 MR[2][CPT], MK[2][CPT];
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
#pragma unroll
    for (int x = 0; x < 2; ++x) {
      MR[x][m] = make_uint2((kRb[x][0] && rR) ? ~0u : 0u, (kRb[x][1] && rR) ? ~0u : 0u);
      MK[x][m] = make_uint2((kOb[x][0] && rO) ? ~0u : 0u, (kOb[x][1] && rO) ? ~0u : 0u);
    }
    stf[m] = (rO && kany && unit_in) ? 1u : 0u;
    // running store pointer: unit (r, lane) of plane pl0 (advanced by one plane per step)
    outp[m] = a.out + ((long long)(pl0 + G.g - 1) * G.plane + (long long)G.nk * (stf[m] ? ic : 0) + (stf[m] ? kc0 : 0));
  }

  U4 cm1[CPT], c0[CPT], c1[CPT], c2[CPT], bprev[CPT], bcur[CPT];
#pragma unroll
  for (int m = 0; m < CPT; ++m) { cm1[m] = c0[m] = c1[m] = c2[m] = bprev[m] = bcur[m] = U4{0ull, 0ull}; }
  double acc[1] = {0.0};

  // ---- prologue: planes pl0 (-> c0) and pl0+1 (-> c1 + de-interleaved copy in ring slot 1) -----
  mbar_wait(&bars[0], 0);
#pragma unroll
  for (int m = 0; m < CPT; ++m) c0[m] = lds_unit(raw_p + araw[m]);
  mbar_wait(&bars[1], 0);
#pragma unroll
  for (int m = 0; m < CPT; ++m) {
    c1[m] = lds_unit(raw_p + TB + araw[m]);
    sts64<0>(dei + TB + ad[m], c1[m].a); sts64<HPB>(dei + TB + ad[m], c1[m].b);
  }
  __syncthreads();
  if (tid == 0) { if (NSP < npl) issue_p(NSP); if (NSP + 1 < npl) issue_p(NSP + 1); }

  // ---- main loop ----------------------------------------------------------------------------------
  // Step u: reds of plane s+1, then blacks of plane s (s = pl0+u).  The four plane registers rotate
  // roles every step; the body is instantiated four times with permuted roles (no register moves).
  int ring = 0;          // u % 3 : ring slot of de-interleaved plane s; raw b slot of this step
  int phase = 0;         // (u / 3) & 1 : mbarrier parity of the b slot; p slot (u+2)%3 has parity pphase
  auto step = [&](int u, U4 (&wm1)[CPT], U4 (&w0)[CPT], U4 (&w1)[CPT], U4 (&w2)[CPT], U4 (&bp)[CPT], U4 (&bc)[CPT]) {
    const int s = pl0 + u;
    const int r0 = ring, r1 = (ring == 2) ? 0 : ring + 1, r2 = (ring == 0) ? 2 : ring - 1;   // u%3, (u+1)%3, (u+2)%3
    const uint32_t e0 = dei + (uint32_t)r0 * TB, e1 = dei + (uint32_t)r1 * TB, e2 = dei + (uint32_t)r2 * TB;
    const int pphase = (ring == 0) ? phase : phase ^ 1;        // ((u+2)/3)&1
    mbar_wait(&bars[r2], (uint32_t)pphase);
    {
      const uint32_t rp = raw_p + (uint32_t)r2 * TB;
#pragma unroll
      for (int m = 0; m < CPT; ++m) w2[m] = lds_unit(rp + araw[m]);
#pragma unroll
      for (int m = 0; m < CPT; ++m) { sts64<0>(e2 + ad[m], w2[m].a); sts64<HPB>(e2 + ad[m], w2[m].b); }
    }
    mbar_wait(&bars[NSP + r0], (uint32_t)phase);
    {
      const uint32_t rbp = raw_b + (uint32_t)r0 * TB;
#pragma unroll
      for (int m = 0; m < CPT; ++m) bc[m] = lds_unit(rbp + araw[m]);
    }
    const bool do_red = (s + 1 >= a.jr0 && s + 1 <= a.jr1);      // CTA-uniform
    const bool red_owned = (s + 1 >= ja && s + 1 <= jb);
    const bool do_black = (s >= ja && s <= jb);
    // A is the red cell of plane s+1 (and the black cell of plane s) iff (par + s + 1) is even  (warp-uniform)
    const bool xa = ((par + (unsigned)(s + 1)) & 1u) == 0u;

    // One relaxation of the X cells (X = A: xb false, X = B: xb true) of plane `wc` for all units of the
    // thread.  en = de-interleaved copy of that plane, wn/ws = planes j+1 / j-1, bb = right-hand side.
    auto relax = [&](auto xb, const uint32_t en, U4 (&wc)[CPT], const U4 (&wn)[CPT], const U4 (&ws)[CPT], const U4 (&bb)[CPT],
                     const uint2 (&msk)[CPT], F2 (&dp)[CPT]) {
      constexpr bool XB = decltype(xb)::value;
      constexpr int XO = XB ? HPB : 0;
      F2 num[CPT], quo[CPT];
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        const uint32_t an = en + ad[m];
        const F2 ie = lds64<XO + PRB>(an), iw = lds64<XO - PRB>(an);
        F2 ss = add2(ie, iw);                                    // i+1, i-1
        if (!XB) {
          // cells (A0,A1): k+1 = own (B0,B1) ; k-1 = (B1 of the unit to the left, own B0)
          const float bl = __shfl_up_sync(0xffffffffu, hi(wc[m].b), 1);
          ss = add2(add2(add2(ss, wn[m].a), ws[m].a), wc[m].b);  // j+1, j-1, k+1
          ss = pk(lo(ss) + bl, hi(ss) + lo(wc[m].b));            // k-1
          num[m] = sub2(ss, bb[m].a);
        } else {
          // cells (B0,B1): k+1 = (own A1, A0 of the unit to the right) ; k-1 = own (A0,A1)
          const float ar = __shfl_down_sync(0xffffffffu, lo(wc[m].a), 1);
          ss = add2(add2(ss, wn[m].b), ws[m].b);                 // j+1, j-1
          ss = pk(lo(ss) + hi(wc[m].a), hi(ss) + ar);            // k+1
          ss = add2(ss, wc[m].a);                                // k-1
          num[m] = sub2(ss, bb[m].b);
        }
        quo[m] = div6_2(num[m]);
      }
      // numerators outside the verified range of the fast division (tiny, huge, -0, non-finite): rare
      float mx = 0.0f, mn = 3.0e38f;
      unsigned orb = 0u;
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        const float n0 = lo(num[m]), n1 = hi(num[m]);
        mx = fmaxf(mx, fmaxf(fabsf(n0), fabsf(n1)));
        mn = fminf(mn, fminf(fabsf(n0), fabsf(n1)));
        orb |= __float_as_uint(n0) | __float_as_uint(n1);
      }
      if (!(orb == 0u || (mn >= 0x1p-123f && mx < 0x1p57f))) {
        static_assert(CPT % 2 == 0, "rare-division fix works on groups of two units");
#pragma unroll
        for (int m = 0; m < CPT; m += 2) {
          const float4 f = div6_fix4(make_float4(lo(num[m]), hi(num[m]), lo(num[m + 1]), hi(num[m + 1])),
                                     make_float4(lo(quo[m]), hi(quo[m]), lo(quo[m + 1]), hi(quo[m + 1])));
          quo[m] = pk(f.x, f.y); quo[m + 1] = pk(f.z, f.w);
        }
      }
#pragma unroll
      for (int m = 0; m < CPT; ++m) {
        const F2 pp = XB ? wc[m].b : wc[m].a;
        dp[m] = mul2(sub2(quo[m], pp), omg2);                    // ((ss-b)/dd - p)*omg
        const F2 nv = sel2(add2(pp, dp[m]), pp, msk[m]);         // p + dp where the cell may move
        if (XB) wc[m].b = nv; else wc[m].a = nv;
      }
    };
    auto both = [&](auto xb) {
      constexpr bool XB = decltype(xb)::value;
      constexpr int XO = XB ? HPB : 0;
      F2 dp[CPT];
      if (do_red) {
        relax(xb, e1, w1, w2, w0, bc, MR[XB], dp);
#pragma unroll
        for (int m = 0; m < CPT; ++m) sts64<XO>(e1 + ad[m], XB ? w1[m].b : w1[m].a);
        if (red_owned) {
#pragma unroll
          for (int m = 0; m < CPT; ++m) {
            const F2 dk = and2(dp[m], MK[XB][m]);
            const F2 d2 = mul2(dk, dk);
            acc[0] += (double)lo(d2);
            acc[0] += (double)hi(d2);
          }
        }
      }
      if (do_black) {
        relax(xb, e0, w0, w1, wm1, bp, MK[XB], dp);
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
          const F2 dk = and2(dp[m], MK[XB][m]);
          const F2 d2 = mul2(dk, dk);
          acc[0] += (double)lo(d2);
          acc[0] += (double)hi(d2);
          stg_unit_if(outp[m], w0[m], stf[m]);
        }
      }
    };
    if (xa) both(std::false_type{}); else both(std::true_type{});
#pragma unroll
    for (int m = 0; m < CPT; ++m) outp[m] += G.plane;
    __syncthreads();
    if (tid == 0) {   // both slots consumed by every thread: refill them three planes ahead
      if (u + 2 + NSP < npl) issue_p(u + 2 + NSP);
      if (u + NSB < nbpl) issue_b(u + NSB);
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
  grid_reduce<1>(acc, a.rb, FinStoreF{a.acc_st, a.slot, a.add});
}

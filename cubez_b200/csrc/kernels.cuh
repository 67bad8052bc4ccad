// cubez_b200/csrc/kernels.cuh -- host-side launchers of the sm_100a kernels (kernels.cu,
// pcr.cu, fused.cu).  All pointers are DEVICE pointers; everything is enqueued on `s`.
// `st` may be null (per-call ABI); when non-null the kernels are no-ops once st->done is set.
#pragma once
#include "common.cuh"

namespace czb {

// Convergence test folded into the sweep (single rank): the CTA that finishes the grid-wide reduction also does
// what k_finish_iteration does -- res = sqrt(sum * res_normal), history, ++itr, done = res < eps
// (cz_Poisson.cpp:69-77) -- which saves one launch per iteration.  Honoured by the fused marching kernels only.
struct FinishSpec { int enable; double res_normal, eps; double* hist; int hist_len; };

struct Launch {           // what every launcher needs
  cudaStream_t s;
  RedBuf rb;              // reduction scratch (sized for rb_blocks blocks)
  size_t rb_blocks;
  const LoopState* st;    // done-flag source (may be null)
  LoopState* acc_st;      // where reduction totals go: acc_st->acc[slot]
  FinishSpec fin{0, 0.0, 0.0, nullptr, 0};
};

void count_launch(int n = 1);

// ---- stationary sweeps (cz_solver.f90) ------------------------------------------------
// jacobi (cz_solver.f90:284-387): pnew = p + omg*((ss-b)/dd - p) on inner cells;
// acc[slot] (+)= sum dp^2.
void launch_jacobi(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b,
                   const REAL* cf, REAL omg, int slot, bool add);
// psor2sma_core (cz_solver.f90:404-493): one colour, in place. kp = ofst + color.
void launch_rb_color(const Launch& L, const Geo& G, REAL* p, const REAL* b, const REAL* cf,
                     REAL omg, int kp, int slot, bool add);
// psor (cz_solver.f90:207-269): lexicographic SOR, exact order through hyperplane wavefronts
void launch_psor_sweep(const Launch& L, const Geo& G, REAL* p, const REAL* b, const REAL* cf, REAL omg,
                       int slot, bool add);
// copy inner cells src -> dst (jacobi copy-back cz_solver.f90:369-375, pcr_j_esa :1659-1665)
void launch_copy_inner(const Launch& L, const Geo& G, REAL* dst, const REAL* src);
// copy every cell that is NOT an inner cell (guards + boundary nodes): the shell a ping-pong partner must share
void launch_copy_shell(const Launch& L, const Geo& G, REAL* dst, const REAL* src);

// ---- BLAS-1 / operator (cz_blas.f90) --------------------------------------------------
void launch_calc_ax(const Launch& L, const Geo& G, REAL* ap, const REAL* p, const REAL* cf,
                    const REAL* dotwith, int slot);   // dotwith!=null: acc[slot] = (ap, dotwith)
void launch_calc_rk(const Launch& L, const Geo& G, REAL* r, const REAL* p, const REAL* b,
                    const REAL* cf);
void launch_dot1(const Launch& L, const Geo& G, const REAL* p, int slot);
void launch_dot2(const Launch& L, const Geo& G, const REAL* p, const REAL* q, int slot);
// z = a*x + y ; a read from device scalar `a_dev` (double holding a REAL value) if non-null
void launch_triad(const Launch& L, const Geo& G, REAL* z, const REAL* x, const REAL* y, REAL a,
                  const double* a_dev);
void launch_bicg1(const Launch& L, const Geo& G, REAL* p, const REAL* r, const REAL* q, REAL beta,
                  REAL omg, const double* beta_dev, const double* omg_dev);
void launch_bicg2(const Launch& L, const Geo& G, REAL* z, const REAL* x, const REAL* y, REAL a,
                  REAL b, const double* a_dev, const double* b_dev);
// x += alpha*p_ + omega*s_ ; r = s - omega*t ; acc[slot_rr] = (r,r) ; acc[slot_rho] = (r,r0).  The scalars are
// read from device memory (LoopState.scal, doubles holding REAL values).
void launch_bicg_tail(const Launch& L, const Geo& G, REAL* x, const REAL* p_, const REAL* s_, REAL* r, const REAL* t,
                      const REAL* s, const REAL* r0, const double* al_dev, const double* om_dev, const double* nom_dev,
                      int slot_rr, int slot_rho);
void launch_imask(const Launch& L, const Geo& G, REAL* x);
void launch_copy_in(const Launch& L, const Geo& G, REAL* y, const REAL* x);

// ---- boundary / utility (cz_solver.f90:22-191, cz_utility.f90) --------------------------
// joff: added to j before evaluating y (0 = reference semantics; j-slabs pass head_j-1 with the GLOBAL
// origin so that the Dirichlet data are bit-identical to the single-domain ones); ioff: the same for i (bricks)
void launch_bc_k(const Launch& L, const Geo& G, REAL* p, REAL dh, const REAL* org, const int* nID, int joff = 0, int ioff = 0);
void launch_exact(const Launch& L, const Geo& G, REAL* e, REAL dh, const REAL* org);
// err_t: e <- p - e on inner cells; out[0] = max |p-e| as double bits, out[1] = order index of
// the first maximum in (j,i,k) loop order.
void launch_err(const Launch& L, const Geo& G, const REAL* p, REAL* e, unsigned long long* out2);

// ---- line SOR with PCR (cz_solver.f90:497-662, 1261-1469, 1473-1676) --------------------
struct PcrTables {          // device tables for one (n, pn, last-stage) combination
  int n, pn, stages, last4;
  REAL* tab;                // [stages][3][n]  ap, cp, e
  REAL* fin;                // last-stage tables, see pcr.cu
};
void pcr_build_tables(cudaStream_t s, PcrTables& T, int n, int pn, bool last4);
void pcr_free_tables(PcrTables& T);
// mode: 0/1 = lines with mod(i+j+ofs,2)==mode ; 2 = all lines (line Jacobi, dst != x) ;
//       3 = lines with i+j == diag (wavefront for the lexicographic variants)
void launch_pcr_lines(const Launch& L, const Geo& G, const PcrTables& T, int mode, int ofs, int diag,
                      const REAL* x, REAL* dst, const REAL* msk, const REAL* rhs, REAL omg, int slot,
                      bool add, bool mask_is_inner = false);
// mask_is_inner: the caller guarantees msk == imask_k of G (1 on the inner range = on every cell of every
// line); multiplying by exactly 1 changes nothing, so the warp-per-line kernel does not read it.

// ---- stretched-grid (MAF) twins (maf.cu; cz_maf.f90, cz_blas.f90:738-1039) ------------------------
// Device pointers to the coordinate vectors X(-1:sz(1)+2), Y(-1:sz(2)+2), Z(-1:sz(3)+2): Fortran index m -> [m+1].
struct MafAxes { const REAL* X; const REAL* Y; const REAL* Z; };
void launch_jacobi_maf(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b, const MafAxes& A,
                       REAL omg, int slot, bool add);
void launch_rb_color_maf(const Launch& L, const Geo& G, REAL* p, const REAL* b, const MafAxes& A, REAL omg, int kp,
                         int slot, bool add);
void launch_psor_sweep_maf(const Launch& L, const Geo& G, REAL* p, const REAL* b, const MafAxes& A, REAL omg, int slot,
                           bool add);
void launch_calc_ax_maf(const Launch& L, const Geo& G, REAL* ap, const REAL* p, const MafAxes& A, const REAL* pvt,
                        const REAL* dotwith, int slot);   // dotwith != null: acc[slot] = (ap, dotwith)
void launch_calc_rk_maf(const Launch& L, const Geo& G, REAL* r, const REAL* p, const REAL* b, const MafAxes& A,
                        const REAL* pvt);
void launch_search_pivot(const Launch& L, const Geo& G, REAL* pvt, const MafAxes& A);
// mode 0/1: lines with mod(i+j,2) == mode; mode 3: lines with i+j == diag.  acc[slot] (+)= sum dp^2
// (cz_maf.f90:642-646: one of the two accumulation statements per build, never both).
void launch_pcr_lines_maf(const Launch& L, const Geo& G, int pn, int mode, int diag, const MafAxes& A, REAL* x,
                          const REAL* msk, const REAL* rhs, REAL omg, int slot, bool add);

// ---- fused TMA-staged marching kernels (fused.cu); return false when the geometry is not
// eligible (row pitch not a multiple of 16 B) and the caller must use the baseline kernels ----
bool fused_eligible(const Geo& G);
// both colours of one red-black SOR iteration, p -> pnew (ping-pong). Reds are relaxed on planes
// jr0..jr1 (includes halo planes owned by a neighbour rank), blacks/stores on jst..jed.
// part: 0 = every j-chunk of the launch, 1 = the first and the last chunk only, 2 = the chunks between (1 + 2 launch
// exactly the CTAs of 0; a j-slab relaxes its edge chunks on a high-priority stream and sends their outermost planes
// to the neighbours while the interior chunks run).  part 2 of a slab with fewer than three chunks launches nothing.
bool launch_rb2_fused(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b, const REAL* cf,
                      REAL omg, int ip, int jr0, int jr1, int slot, bool add, bool zero_src = false, int part = 0);
// zero_src: p is known to be all zero and is not read at all (packed float kernel only; check first)
bool rb2_zero_src_supported(const Geo& G, const REAL* cf);
bool rb2_fused_launchable(const Launch& L, const Geo& G, const REAL* cf, bool zero_src, int part);
bool jacobi_fused_launchable(const Launch& L, const Geo& G, const REAL* cf, int part);
bool launch_jacobi_fused(const Launch& L, const Geo& G, const REAL* p, REAL* pnew, const REAL* b, const REAL* cf,
                         REAL omg, int slot, bool add, int part = 0);
// ap = A p on inner cells, 0 on the other cells of stored pairs; acc[slot] = (ap, dotwith).
// Returns 0 = not eligible (use launch_calc_ax), 1 = done, 2 = done and acc[slot2] = (ap, ap) as well
// (slot2 < 0: not wanted).
int launch_ax_fused(const Launch& L, const Geo& G, REAL* ap, const REAL* p, const REAL* cf, const REAL* dotwith,
                    int slot, int slot2 = -1, int part = 0);

// ---- persistent cooperative sweeps for L2-resident grids (persist.cu) --------------------------------------
// nfields = number of fields the loop touches (Jacobi: X, WRK, B; red-black SOR in place: X, B)
bool persist_eligible(const Geo& G, int nfields);
// Runs iterations it_first..it_last of CZ::JACOBI (mode 0: ping-pong, iteration `it` reads X when it is odd) or of
// CZ::RBSOR (mode 1: both colours in place on X, kp = colour offset) in ONE cooperative launch, convergence test,
// history and LoopState update included.  Returns at once when st->done is already set.
void launch_persist(const Launch& L, const Geo& G, int mode, REAL* X, REAL* W, const REAL* B, const REAL* cf, REAL omg,
                    int kp, int it_first, int it_last, LoopState* st, double res_normal, double eps, double* hist,
                    int hist_len);

// ---- loop control (tiny kernels) --------------------------------------------------------
void launch_state_reset(const Launch& L, LoopState* st);
// res = sqrt(acc[slot]*res_normal); hist[itr] = res; ++itr; done |= res < eps
void launch_finish_iteration(const Launch& L, LoopState* st, int slot, double res_normal, double eps,
                             double* hist, int hist_len, int slot2 = -1);   // slot2 >= 0: acc[slot] += acc[slot2] first

}  // namespace czb

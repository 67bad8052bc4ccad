// cubez_b200/csrc/solver.cu -- device-resident iteration loops reproducing the control flow of
// the reference's CZ::JACOBI / PSOR / RBSOR / PBiCGSTAB / LSOR_PCR_* (src/cz_cpp/cz_Poisson.cpp)
// plus the czb_ctx_* / czb_solve entry points of include/czb200.h.
//
// Difference from the reference that matters for speed, not for results: the convergence test
// `res < eps` is evaluated ON THE DEVICE (k_finish_iteration / k_bicg_scalar set LoopState.done
// and every later kernel of the stream returns immediately), so the host enqueues iterations in
// batches and only polls a pinned copy of the state; iteration counts and residual histories
// are exactly those of a loop that checks after every iteration.
#include "solver.h"
#include "../../include/czb200.h"
#include <cfloat>
#include <cmath>
#include <strings.h>
#include <cuda.h>
#include <nvtx3/nvToolsExt.h>   // header-only; a no-op unless a profiler is attached (reference: PUSH_RANGE / POP_RANGE, cz.h:46-70)

using namespace czb;

namespace {

// BiCGstab scalar recurrences (cz_Poisson.cpp:376-500), all in REAL like the reference's
// REAL_TYPE locals; dot results are rounded to REAL first (Fdot1/Fdot2 return REAL_TYPE).
__global__ void k_bicg_scalar(LoopState* st, int phase, double res_normal, double eps, double* hist, int hist_len) {
  if (st->done) return;
  if (phase == -1) {           // cz_Poisson.cpp:368-371
    st->scal[SC_RHO_OLD] = 1.0; st->scal[SC_ALPHA] = 0.0; st->scal[SC_OMEGA] = 1.0;
    st->scal[SC_NEG_OMEGA] = -1.0; st->scal[SC_BETA] = 0.0; st->scal[SC_RHO] = 0.0;
  } else if (phase == 0) {     // rho, breakdown test, beta  (:376-393)
    const REAL rho = (REAL)st->acc[0];
    st->scal[SC_RHO] = (double)rho;
    if (fabs((double)rho) < (double)FLT_MIN) { st->done = 1; st->status = 1; return; }
    if (st->itr >= 1) {
      const REAL rho_old = (REAL)st->scal[SC_RHO_OLD], alpha = (REAL)st->scal[SC_ALPHA], omega = (REAL)st->scal[SC_OMEGA];
      const REAL beta = rho / rho_old * alpha / omega;
      st->scal[SC_BETA] = (double)beta;
    }
  } else if (phase == 1) {     // alpha = rho / (q, r0)  (:427-431)
    const REAL rho = (REAL)st->scal[SC_RHO];
    const REAL alpha = rho / (REAL)st->acc[1];
    st->scal[SC_ALPHA] = (double)alpha;
    st->scal[SC_NEG_ALPHA] = (double)(-alpha);
  } else if (phase == 2) {     // omega = (t, s) / (t, t)  (:464-465)
    const REAL omega = (REAL)st->acc[2] / (REAL)st->acc[3];
    st->scal[SC_OMEGA] = (double)omega;
    st->scal[SC_NEG_OMEGA] = (double)(-omega);
  } else {                     // res = Fdot1(r); normalise; history; test; rho_old = rho  (:481-500)
    double res = (double)(REAL)st->acc[4];
    res *= res_normal;
    res = sqrt(res);
    st->res = res;
    const int it = st->itr;
    if (hist && it < hist_len) hist[it] = res;
    st->itr = it + 1;
    if (res < eps) { st->done = 1; return; }
    st->scal[SC_RHO_OLD] = st->scal[SC_RHO];
  }
}

// bc_k inside a solve loop = the reference's "BoundaryCondition" section (cz_Poisson.cpp:73-75 and alike)
void phase_bc_k(const Launch& L, const Geo& G, REAL* p, REAL dh, const REAL* org, const int* nID, int joff, int ioff) {
  Phase ph("BoundaryCondition");
  launch_bc_k(L, G, p, dh, org, nID, joff, ioff);
}

int getNumStage(int n) {   // cz.h:293-300
  int b = 1;
  for (int i = 1; i < 20; i++) { b *= 2; if (n < b) return i; }
  return -1;
}

inline double ipow2(int e) { return (double)(1LL << e); }

struct NvtxRange {   // scoped PUSH_RANGE / POP_RANGE
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

}  // namespace

int czb_parse_solver(const char* q) {
  if (!q) return 0;
  struct { const char* n; int v; } tab[] = {
    {"jacobi", LS_JACOBI}, {"psor", LS_PSOR}, {"sor2sma", LS_SOR2SMA}, {"pbicgstab", LS_BICGSTAB},
    {"pcr_rb", LS_PCR_RB}, {"pcr_rb_esa", LS_PCR_RB_ESA}, {"pcr_j_esa", LS_PCR_J_ESA}, {"pcr", LS_PCR},
    {"pcr_eda", LS_PCR_EDA}, {"pcr_esa", LS_PCR_ESA},
    {"jacobi_maf", LS_JACOBI_MAF}, {"psor_maf", LS_PSOR_MAF}, {"sor2sma_maf", LS_SOR2SMA_MAF},
    {"pbicgstab_maf", LS_BICGSTAB_MAF}, {"pcr_rb_maf", LS_PCR_RB_MAF}, {"pcr_rb_esa_maf", LS_PCR_RB_ESA_MAF},
    {"pcr_maf", LS_PCR_MAF}, {"pcr_eda_maf", LS_PCR_EDA_MAF}, {"pcr_esa_maf", LS_PCR_ESA_MAF}};
  for (auto& t : tab) if (!strcasecmp(q, t.n)) return t.v;
  return 0;
}

// ============================================================================ helpers
REAL* czb_ctx::field(REAL*& slot) {
  if (!slot) {
    slot = (REAL*)czb_malloc(sizeof(REAL) * field_elems());
    if (!slot) { fprintf(stderr, "czb200: out of device memory for a work field\n"); abort(); }
  }
  return slot;
}

Launch czb_ctx::launch(bool with_state) {
  Runtime& R = rt();
  Launch L = R.launch(with_state ? d_state : nullptr, d_state);
  return L;
}

void czb_ctx::Comm_S(REAL* f, int width) {
  if (comm && comm->nranks > 1) { Phase ph("Comm_Poisson"); comm_halo(comm, f, G, width, rt().stream); }
}

void czb_ctx::Comm_SUM(int slot, int count) {
  if (comm && comm->nranks > 1) { Phase ph("Comm_Res_Poisson"); comm_allreduce_sum(comm, &d_state->acc[slot], count, rt().stream); }
}

// ---- stretched-grid (MAF) state ---------------------------------------------------------------
// Default mesh = the reference driver's: xc[i] = (REAL)(i-1)*pitch for i = 0..size+2*gc-1 (cz_Evaluate.cpp:342-362).
// On j-slabs the GLOBAL index is used (i-1 + head-1), so that a slab sees the coordinates of the single domain
// (same reasoning as bc_origin).
void czb_ctx::set_mesh(const REAL* x, const REAL* y, const REAL* z) {
  Runtime& R = rt();
  REAL** dst[3] = {&xc, &yc, &zc};
  const REAL* src[3] = {x, y, z};
  for (int a = 0; a < 3; ++a) {
    const size_t n = (size_t)size[a] + 2 * gc;
    std::vector<REAL> h(n);
    for (size_t i = 0; i < n; ++i)
      h[i] = src[a] ? src[a][i] : (REAL)((int)i - 1 + head[a] - 1) * pitch;
    if (!*dst[a]) CZB_CHECK(cudaMalloc(dst[a], sizeof(REAL) * n));
    CZB_CHECK(cudaMemcpyAsync(*dst[a], h.data(), sizeof(REAL) * n, cudaMemcpyHostToDevice, R.stream));
    R.sync();   // h goes out of scope
  }
  mesh_user = (x || y || z);
  if (pvt) launch_search_pivot(launch(false), G, pvt, MafAxes{xc, yc, zc});
}

MafAxes czb_ctx::axes() {
  if (!xc) set_mesh(nullptr, nullptr, nullptr);
  return MafAxes{xc, yc, zc};
}

REAL* czb_ctx::pivot() {
  if (!pvt) {
    const MafAxes A = axes();
    field(pvt);                                        // zero outside the inner cells, as czAllocR_S3D leaves it
    launch_search_pivot(launch(false), G, pvt, A);
  }
  return pvt;
}

void czb_ctx::ensure_comm_stream() {
  if (cs) return;
  int lo = 0, hi = 0;
  CZB_CHECK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CZB_CHECK(cudaStreamCreateWithPriority(&cs, cudaStreamNonBlocking, hi));
  CZB_CHECK(cudaStreamCreateWithPriority(&ov.es, cudaStreamNonBlocking, hi));
  cudaEvent_t* evs[] = {&ev_fin[0], &ev_fin[1], &ev_join, &ov.ev_a[0], &ov.ev_a[1], &ov.ev_b[0], &ov.ev_b[1], &ov.ev_start, &ov.ev_join_e};
  for (cudaEvent_t* e : evs) CZB_CHECK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  Runtime& R = rt();
  CZB_CHECK(cudaMalloc(&ov.rb_edge.partial, sizeof(double) * R.rb_blocks * CZB_MAX_ACC));
  CZB_CHECK(cudaMalloc(&ov.rb_edge.ticket, sizeof(unsigned int)));
  CZB_CHECK(cudaMemset(ov.rb_edge.ticket, 0, sizeof(unsigned int)));
}

// ---- SlabOverlap ---------------------------------------------------------------------------------------------------
bool czb_ctx::slab_overlap_begin(bool local_ok, int w) {
  if (!comm || comm->nranks <= 1) return false;
  if (comm->brick) return false;      // i / k faces need the packed exchange: plain loop (the flag is collective)
  static const bool no_ovl = getenv("CZB_NO_OVERLAP") != nullptr;
  if (no_ovl) return false;
  // local_ok depends on the solver and on the geometry in k and i only (tile counts, row pitch), which every slab of a
  // run shares; a solver that cannot split its sweeps stays on the plain loop without a collective
  if (!local_ok) return false;
  Runtime& R = rt();
  if (!ov.decided) {   // one collective decision per context: a ring that mixes the two loop forms would hang
    ov.p2p = comm_p2p_setup(comm, (size_t)G.plane, R.stream, local_ok);     // true only if every rank is able AND willing
    ov.all_ok = ov.p2p;
    if (!ov.p2p) {     // peer transport unavailable or refused: do all ranks at least want the split launches (NCCL exchange)?
      double* flag = nullptr;
      CZB_CHECK(cudaMalloc((void**)&flag, sizeof(double)));
      const double f = local_ok ? 0.0 : 1.0;
      CZB_CHECK(cudaMemcpyAsync(flag, &f, sizeof(double), cudaMemcpyHostToDevice, R.stream));
      comm_allreduce_sum(comm, flag, 1, R.stream);
      double bad = 1.0;
      CZB_CHECK(cudaMemcpyAsync(&bad, flag, sizeof(double), cudaMemcpyDeviceToHost, R.stream));
      R.sync();
      cudaFree(flag);
      ov.all_ok = (bad == 0.0);
    }
    ov.decided = true;
  }
  if (!ov.all_ok) return false;
  (void)w;
  ensure_comm_stream();
  ov.deferred = false;
  return true;
}

void czb_ctx::slab_exchange(REAL* dst, int w, int it) {
  Phase ph("Comm_Poisson");
  if (ov.p2p) comm_p2p_push(comm, dst + (long long)G.g * G.plane, dst + (long long)(G.jx - w + G.g) * G.plane, w, it & 1, ov.es);
  else        comm_halo(comm, dst, G, w, ov.es);
}

void czb_ctx::slab_overlap_edges_begin(int it, REAL* src, int w, bool converge_check, Launch& LE, Launch& LM) {
  Runtime& R = rt();
  LM = launch();
  LE = launch(); LE.s = ov.es; LE.rb = ov.rb_edge;
  if (it == 1) {   // everything queued on the solve stream so far (state reset, initial exchanges, WRK copy) comes first
    CZB_CHECK(cudaEventRecord(ov.ev_start, R.stream));
    CZB_CHECK(cudaStreamWaitEvent(ov.es, ov.ev_start, 0));
    CZB_CHECK(cudaStreamWaitEvent(cs, ov.ev_start, 0));
  } else {
    // the edge chunks read two planes of the interior launch of iteration it-1; their guard planes come from the
    // neighbours' edge launches of iteration it-1
    CZB_CHECK(cudaStreamWaitEvent(ov.es, ov.ev_b[(it - 1) & 1], 0));
    if (ov.p2p) {
      comm_p2p_wait(comm, ov.es);
      comm_p2p_fetch(comm, src + (long long)(G.g - w) * G.plane, src + (long long)(G.jx + G.g) * G.plane, w, (it - 1) & 1, ov.es);
    }
    CZB_CHECK(cudaStreamWaitEvent(R.stream, ov.ev_a[(it - 1) & 1], 0));   // the interior chunks read the edge chunks' planes
  }
  if (converge_check && it >= 3) {   // the test of iteration it-2 decides whether this one still runs
    CZB_CHECK(cudaStreamWaitEvent(ov.es, ev_fin[(it - 2) & 1], 0));
    CZB_CHECK(cudaStreamWaitEvent(R.stream, ev_fin[(it - 2) & 1], 0));
  }
}

void czb_ctx::slab_overlap_edges_end(int it, REAL* dst, int w, bool defer_exchange) {
  CZB_CHECK(cudaEventRecord(ov.ev_a[it & 1], ov.es));
  ov.deferred = defer_exchange;
  if (!defer_exchange) slab_exchange(dst, w, it);
}

void czb_ctx::slab_overlap_iteration_end(int it, REAL* dst, int w, bool converge_check, int slot_a, int slot_b) {
  Runtime& R = rt();
  CZB_CHECK(cudaEventRecord(ov.ev_b[it & 1], R.stream));
  if (ov.deferred) {
    CZB_CHECK(cudaStreamWaitEvent(ov.es, ov.ev_b[it & 1], 0));
    slab_exchange(dst, w, it);
    ov.deferred = false;
  }
  if (converge_check) {
    CZB_CHECK(cudaStreamWaitEvent(cs, ov.ev_a[it & 1], 0));
    CZB_CHECK(cudaStreamWaitEvent(cs, ov.ev_b[it & 1], 0));
    static const bool exp_noar = getenv("CZB_EXP_NOALLREDUCE") != nullptr;   // timing experiment only (wrong results)
    if (!exp_noar) comm_allreduce_sum(comm, &d_state->acc[slot_a], slot_b == slot_a + 1 ? 2 : 1, cs);
    Launch Lc = launch(); Lc.s = cs;
    launch_finish_iteration(Lc, d_state, slot_a, res_normal, eps, d_hist, hist_cap, slot_b == slot_a + 1 ? slot_b : -1);
    CZB_CHECK(cudaEventRecord(ev_fin[it & 1], cs));
    cs_used = true;
  }
}

void czb_ctx::slab_overlap_finish(REAL* X, int w, int done_iters) {
  Runtime& R = rt();
  CZB_CHECK(cudaEventRecord(ov.ev_join_e, ov.es));
  CZB_CHECK(cudaStreamWaitEvent(R.stream, ov.ev_join_e, 0));
  join_comm_stream();
  // a launch that ran while `done` became visible may have left its reduction ticket behind (see end_loop)
  CZB_CHECK(cudaMemsetAsync(ov.rb_edge.ticket, 0, sizeof(unsigned int), R.stream));
  if (ov.p2p) {
    // Quiesce the ring (every push we are owed has landed, so the neighbours have also finished reading what we sent),
    // then give X the halo of its final state, as a Comm_S(X, w) would have.
    comm_p2p_wait(comm, R.stream);
    if (done_iters >= 1)
      comm_p2p_fetch(comm, X + (long long)(G.g - w) * G.plane, X + (long long)(G.jx + G.g) * G.plane, w, done_iters & 1, R.stream);
  } else if (done_iters & 1) {
    Comm_S(X, w);   // the last exchange filled WRK's guard planes
  }
}

// The persistent cooperative loop (persist.cu) serves single-rank solves of grids whose fields fit L2; the phase
// profile wants one section per kernel call and keeps the per-iteration launches.
bool czb_ctx::persist_ok(int nfields) {
  if (comm && comm->nranks > 1) return false;
  if (rt().prof.on || maf || variant != 0) return false;      // variant 1 = the baseline per-call kernels
  return persist_eligible(G, nfields);
}

void czb_ctx::join_comm_stream() {
  if (!cs || !cs_used) return;
  CZB_CHECK(cudaEventRecord(ev_join, cs));
  CZB_CHECK(cudaStreamWaitEvent(rt().stream, ev_join, 0));
  cs_used = false;
}

void czb_ctx::begin_loop(int itr_max) {
  Runtime& R = rt();
  if (hist_cap < itr_max + 2) {
    if (d_hist) { R.sync(); cudaFree(d_hist); }
    hist_cap = itr_max + 2;
    CZB_CHECK(cudaMalloc(&d_hist, sizeof(double) * hist_cap));
  }
  launch_state_reset(launch(false), d_state);
  CZB_CHECK(cudaEventRecord(ev_t0, R.stream));
}

int czb_ctx::end_loop(int itr_max, double& res, bool bicg) {
  Runtime& R = rt();
  join_comm_stream();
  // an iteration started before `done` became visible may have left the reduction ticket of a partly
  // skipped launch behind: re-arm it
  CZB_CHECK(cudaMemsetAsync(R.rb.ticket, 0, sizeof(unsigned int), R.stream));
  CZB_CHECK(cudaEventRecord(ev_t1, R.stream));
  CZB_CHECK(cudaMemcpyAsync(&h_state[0], d_state, sizeof(LoopState), cudaMemcpyDeviceToHost, R.stream));
  R.sync();
  float ms = 0.f;
  CZB_CHECK(cudaEventElapsedTime(&ms, ev_t0, ev_t1));
  last_ms = ms;
  const LoopState& S = h_state[0];
  res = S.res;
  if (S.status != 0) return 0;                      // breakdown: "Solver error" (cz_Poisson.cpp:379-383)
  if (S.done) return S.itr;
  return bicg ? itr_max : itr_max + 1;              // loop bounds cz_Poisson.cpp:39,373
}

template <class Body>
int czb_ctx::run_checked(int itr_max, double& res, Body body) {
  Runtime& R = rt();
  begin_loop(itr_max);
  const int batch = 16;
  int launched = 0, nb = 0;
  bool stop = false;
  // Single rank: the second batch (iterations 17..32; the first one carries the bc_k launches of iterations 1, 2) is
  // captured into a CUDA graph and REPLAYED for every later full batch.  body(it) depends on `it` only through its
  // parity (ping-pong roles) once it > 2 and every batch starts at an odd iteration, so the captured launches are
  // exactly those a direct enqueue would make; the kernels read their iteration state (done, itr) on the device.
  static const bool no_graph = getenv("CZB_NO_GRAPH") != nullptr;
  bool graphs = !no_graph && !rt().prof.on && !(comm && comm->nranks > 1) && itr_max >= 3 * batch;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t gexec = nullptr;
  long long graph_launches = 0;
  while (launched < itr_max && !stop) {
    const int m = (itr_max - launched < batch) ? itr_max - launched : batch;
    if (graphs && m == batch && nb >= 1 && !gexec &&
        cudaStreamBeginCapture(R.stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
      cudaGetLastError();
      graphs = false;            // stream capture unavailable (e.g. the stream is already being captured by the caller)
    }
    if (graphs && m == batch && nb >= 1) {
      if (!gexec) {
        const long long l0 = launch_count();
        for (int q = 0; q < m; ++q) body(launched + q + 1);
        CZB_CHECK(cudaStreamEndCapture(R.stream, &graph));
        graph_launches = launch_count() - l0;
        CZB_CHECK(cudaGraphInstantiate(&gexec, graph, 0));
      } else {
        count_launch((int)graph_launches);
      }
      CZB_CHECK(cudaGraphLaunch(gexec, R.stream));
    } else {
      for (int q = 0; q < m; ++q) body(launched + q + 1);
    }
    launched += m;
    const int slot = nb & 1;
    // Convergence tests that run on the communication stream must be part of the snapshot: every rank has to see
    // the SAME state after the same batch, or the ranks would enqueue different numbers of iterations (and of
    // halo exchanges / signals) and hang.
    join_comm_stream();
    CZB_CHECK(cudaMemcpyAsync(&h_state[1 + slot], d_state, sizeof(LoopState), cudaMemcpyDeviceToHost, R.stream));
    CZB_CHECK(cudaEventRecord(ev[slot], R.stream));
    if (nb >= 1) {   // look at the batch before this one while this one runs
      const int ps = (nb - 1) & 1;
      CZB_CHECK(cudaEventSynchronize(ev[ps]));
      if (h_state[1 + ps].done) stop = true;
    }
    ++nb;
  }
  const int itr = end_loop(itr_max, res, false);
  if (gexec) cudaGraphExecDestroy(gexec);
  if (graph) cudaGraphDestroy(graph);
  return itr;
}

// ============================================================================ loops
// CZ::JACOBI (cz_Poisson.cpp:30-82).  The reference's copy-back pass (cz_solver.f90:369-375) is
// replaced by ping-pong between X and WRK; both hold the same boundary/guard values.
int czb_ctx::JACOBI(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check) {
  Runtime& R = rt();
  REAL* W = field(WRK);
  if (converge_check) {
    launch_copy_shell(launch(false), G, W, X);   // WRK only needs X's guard cells and boundary nodes: every inner cell is stored by sweep 1
    wrk_guards_zero = false;
  } else if (!wrk_guards_zero) {   // preconditioner call: xx is all-zero outside the inner cells
    CZB_CHECK(cudaMemsetAsync(W, 0, sizeof(REAL) * field_elems(), R.stream));
    wrk_guards_zero = true;
  }
  const int rs = converge_check ? 0 : 7;   // reduction slot (7 = scratch inside the preconditioner)
  // multi-rank: edge chunks + one plane to each neighbour while the interior chunks run (SlabOverlap, see RBSOR)
  const bool ovl = slab_overlap_begin(!maf && variant == 0 && fused_eligible(G) && jacobi_fused_launchable(launch(), G, cf, 0) &&
                                      jacobi_fused_launchable(launch(), G, cf, 1), 1);
  auto body_ovl = [&](int it) {
    REAL* src = (it & 1) ? X : W;
    REAL* dst = (it & 1) ? W : X;
    const int sa = converge_check ? ((it & 1) ? 0 : 5) : rs, sb = converge_check ? sa + 1 : rs;
    Launch LE, LM;
    slab_overlap_edges_begin(it, src, 1, converge_check, LE, LM);
    {
      Phase ph("JACOBI_kernel", 18.0 * G.cells());
      if (!launch_jacobi_fused(LE, G, src, dst, B, cf, ac1, sa, false, 1)) { fprintf(stderr, "czb200: fused Jacobi sweep refused a launch\n"); abort(); }
    }
    slab_overlap_edges_end(it, dst, 1, converge_check && it <= 2);
    launch_jacobi_fused(LM, G, src, dst, B, cf, ac1, sb, false, 2);
    if (converge_check && it <= 2) {
      CZB_CHECK(cudaStreamWaitEvent(R.stream, ov.ev_a[it & 1], 0));
      phase_bc_k(launch(false), G, dst, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
    }
    slab_overlap_iteration_end(it, dst, 1, converge_check, sa, sb);
  };
  auto body = [&](int it) {
    if (ovl) { body_ovl(it); return; }
    REAL* src = (it & 1) ? X : W;
    REAL* dst = (it & 1) ? W : X;
    Launch L = launch();
    // single rank: the convergence test rides in the sweep's reduction epilogue (FinishSpec), no extra launch
    const bool fold = converge_check && !(comm && comm->nranks > 1);
    Launch Lf = L;
    if (fold) Lf.fin = FinishSpec{1, res_normal, eps, d_hist, hist_cap};
    bool folded = false;
    {
      Phase ph(maf ? "JACOBI_MAF_kernel" : "JACOBI_kernel", (maf ? 66.0 : 18.0) * G.cells());   // cz_Poisson.cpp:48-59
      if (maf) launch_jacobi_maf(L, G, src, dst, B, axes(), ac1, rs, false);   // cz_Poisson.cpp:45-52
      else if (variant == 0 && launch_jacobi_fused(Lf, G, src, dst, B, cf, ac1, rs, false)) folded = fold;
      else launch_jacobi(L, G, src, dst, B, cf, ac1, rs, false);
    }
    Comm_S(dst, 1);
    if (converge_check) {
      Comm_SUM(0);
      if (!folded) launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
      // bc_k (cz_Poisson.cpp:73) is idempotent once both ping-pong buffers hold the Dirichlet values (see RBSOR).
      // It is launched WITHOUT the done gate (here and in every loop below): the reference calls it before
      // `if (res < eps) break` (cz_Poisson.cpp:73-77, 495-497), i.e. also on the converging iteration.
      if (it <= 2) phase_bc_k(launch(false), G, dst, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
    }
  };
  int itr, done_iters;
  if (converge_check && persist_ok(3)) {
    // L2-resident grid: iterations 1 and 2 as below (they carry the bc_k calls that give both ping-pong buffers the
    // Dirichlet data), every later one inside ONE cooperative launch (persist.cu)
    begin_loop(itr_max);
    for (int it = 1; it <= 2 && it <= itr_max; ++it) body(it);
    { Phase ph("JACOBI_kernel"); launch_persist(launch(), G, 0, X, W, B, cf, ac1, 0, 3, itr_max, d_state, res_normal, eps, d_hist, hist_cap); }
    itr = end_loop(itr_max, res, false);
    done_iters = h_state[0].itr;
  } else if (converge_check) {
    itr = run_checked(itr_max, res, body);
    done_iters = h_state[0].itr;
  } else {
    for (int it = 1; it <= itr_max; ++it) body(it);
    itr = itr_max + 1; done_iters = itr_max;
  }
  if (!converge_check) join_comm_stream();
  if (done_iters & 1) {
    launch_copy_inner(launch(false), G, X, W);
    if (converge_check) phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);   // X missed the bc_k of an odd last iteration
  }
  if (ovl) slab_overlap_finish(X, 1, done_iters);
  else if (done_iters & 1) Comm_S(X, 1);   // the last exchange filled WRK's guard planes; X leaves with the halo of its final state (cz_Poisson.cpp:63)
  flop += (maf ? 66.0 : 18.0) * G.cells() * done_iters;
  return itr;
}

// CZ::PSOR (cz_Poisson.cpp:95-146): lexicographic SOR in the reference's 1-thread order.
int czb_ctx::PSOR(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check) {
  const int rs = converge_check ? 0 : 7;   // reduction slot (7 = scratch inside the preconditioner)
  auto body = [&](int it) {
    Launch L = launch();
    {
      Phase ph(maf ? "SOR_MAF_kernel" : "SOR_kernel", (maf ? 66.0 : 18.0) * G.cells());   // cz_Poisson.cpp:110-120
      if (maf) launch_psor_sweep_maf(L, G, X, B, axes(), ac1, rs, false);   // cz_Poisson.cpp:108-114
      else     launch_psor_sweep(L, G, X, B, cf, ac1, rs, false);
    }
    Comm_S(X, 1);
    if (converge_check) {
      Comm_SUM(0);
      launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
      if (it <= 1) phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);   // in place: boundary nodes never change afterwards
    }
  };
  int itr, done_iters;
  if (converge_check) { itr = run_checked(itr_max, res, body); done_iters = h_state[0].itr; }
  else { for (int it = 1; it <= itr_max; ++it) body(it); itr = itr_max + 1; done_iters = itr_max; }
  flop += (maf ? 66.0 : 18.0) * G.cells() * done_iters;
  return itr;
}

// CZ::RBSOR (cz_Poisson.cpp:159-235).  Multi-rank: halo exchange after EACH colour so that the
// result equals the single-domain one (the reference's MPI build exchanges once per iteration
// and therefore differs from its own serial result).
int czb_ctx::RBSOR(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check) {
  const int ip = parity_ip();
  const int rs = converge_check ? 0 : 7;   // reduction slot (7 = scratch inside the preconditioner)
  // Eligibility of every launch form the loop below uses is decided ONCE, here (march() refuses e.g. grids with
  // more tiles than reduction slots); a refused launch inside the loop would make the iteration a silent no-op.
  const bool zs_wanted = !converge_check && pre_zero_src;
  // bricks: the two-colour kernel would have to recompute the neighbours' reds in the i and k guard layers as it does
  // along j -> the per-colour kernels with an exchange after each colour
  const bool fused = !maf && (variant == 0) && !(comm && comm->brick) && fused_eligible(G) &&
                     rb2_fused_launchable(launch(), G, cf, false, 0) &&
                     (!zs_wanted || rb2_fused_launchable(launch(), G, cf, true, 0));
  auto must = [](bool ok) {
    if (!ok) { fprintf(stderr, "czb200: the fused red-black sweep refused a launch it had accepted at set-up\n"); abort(); }
  };
  int itr, done_iters;
  if (converge_check && persist_ok(2)) {
    // L2-resident grid: iteration 1 with the per-colour kernels (it carries the bc_k call, cz_Poisson.cpp:226), every
    // later one inside ONE cooperative launch that relaxes both colours in place (persist.cu)
    begin_loop(itr_max);
    {
      Launch L = launch();
      for (int color = 0; color < 2; ++color) {
        Phase ph("SOR2SMA_kernel", 9.0 * G.cells());
        launch_rb_color(L, G, X, B, cf, ac1, ip + color, 0, color == 1);
      }
      launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
      phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
    }
    { Phase ph("SOR2SMA_kernel"); launch_persist(launch(), G, 1, X, nullptr, B, cf, ac1, ip, 2, itr_max, d_state, res_normal, eps, d_hist, hist_cap); }
    itr = end_loop(itr_max, res, false);
    done_iters = h_state[0].itr;
  } else if (fused) {
    // Both colours in one pass (fused.cu MODE_RB2), ping-pong between X and WRK.  Multi-rank: one
    // exchange of TWO planes per iteration; the reds of the neighbour's edge plane are recomputed
    // locally from them, which keeps the result equal to the single-domain one.
    Runtime& R = rt();
    REAL* W = field(WRK);
    const bool mr = comm && comm->nranks > 1;
    if (converge_check) {
      if (mr) { Phase ph("Comm_Poisson"); REAL* both[2] = {X, B}; comm_halo_n(comm, both, 2, G, 2, R.stream); }   // X and B in one NCCL group
      launch_copy_shell(launch(false), G, W, X);   // WRK only needs X's guard cells and boundary nodes: every inner cell is stored by sweep 1
      wrk_guards_zero = false;
    } else {
      if (!wrk_guards_zero) {
        CZB_CHECK(cudaMemsetAsync(W, 0, sizeof(REAL) * field_elems(), R.stream));
        wrk_guards_zero = true;
      }
      Comm_S(B, 2);
    }
    const int jr0 = G.jst - (nID[2] >= 0 ? 1 : 0), jr1 = G.jed + (nID[3] >= 0 ? 1 : 0);
    // Multi-rank overlap (SlabOverlap below): the first and the last j-chunk of the sweep run on a high-priority
    // stream and their outermost planes travel to the neighbours -- copy engines over NVLink peer memory, or NCCL --
    // while the chunks between run on the main stream.  Both launches are the SAME kernel a single GPU runs.  The
    // residual all-reduce and the convergence test of iteration k run on a third stream next to iteration k+1, which
    // only READS the converged field (it writes the other ping-pong buffer); iteration k+2 waits for the test of k,
    // so iteration counts, histories and the returned field are those of a loop that tests every iteration.
    static const bool no_ovl = getenv("CZB_NO_OVERLAP") != nullptr;
    const bool ovl_local = mr && !no_ovl && rb2_fused_launchable(launch(), G, cf, false, 1) &&
                           (!rb2_zero_src_supported(G, cf) || rb2_fused_launchable(launch(), G, cf, true, 1));
    const bool ovl = slab_overlap_begin(ovl_local, 2);          // collective decision, opens the edge stream
    auto body_ovl = [&](int it) {
      REAL* src = (it & 1) ? X : W;
      REAL* dst = (it & 1) ? W : X;
      const int sa = converge_check ? ((it & 1) ? 0 : 5) : rs, sb = converge_check ? sa + 1 : rs;
      const bool zs = !converge_check && it == 1 && pre_zero_src;
      Launch LE, LM;
      slab_overlap_edges_begin(it, src, 2, converge_check, LE, LM);
      {
        Phase ph("SOR2SMA_kernel", 18.0 * G.cells());
        must(launch_rb2_fused(LE, G, src, dst, B, cf, ac1, ip, jr0, jr1, sa, false, zs, 1));
      }
      slab_overlap_edges_end(it, dst, 2, converge_check && it <= 2);
      must(launch_rb2_fused(LM, G, src, dst, B, cf, ac1, ip, jr0, jr1, sb, false, zs, 2));
      if (converge_check && it <= 2) {   // bc_k after both launches, before the edge planes leave (see SlabOverlap)
        CZB_CHECK(cudaStreamWaitEvent(R.stream, ov.ev_a[it & 1], 0));
        phase_bc_k(launch(false), G, dst, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
      }
      slab_overlap_iteration_end(it, dst, 2, converge_check, sa, sb);
    };
    auto body = [&](int it) {
      if (ovl) { body_ovl(it); return; }
      REAL* src = (it & 1) ? X : W;
      REAL* dst = (it & 1) ? W : X;
      Launch L = launch();
      const bool fold = converge_check && !mr;   // convergence test inside the sweep (FinishSpec)
      if (fold) L.fin = FinishSpec{1, res_normal, eps, d_hist, hist_cap};
      {
        Phase ph("SOR2SMA_kernel", 18.0 * G.cells());   // cz_Poisson.cpp:203-210 (both colours)
        // first preconditioner sweep: xx is all zero by contract (cz_Poisson.cpp:405,441) and is not read
        must(launch_rb2_fused(L, G, src, dst, B, cf, ac1, ip, jr0, jr1, rs, false, !converge_check && it == 1 && pre_zero_src));
      }
      Comm_S(dst, 2);
      if (converge_check) {
        Comm_SUM(0);
        if (!fold) launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
        // bc_k (cz_Poisson.cpp:226) is idempotent: the sweep never changes a boundary node and carries
        // the boundary nodes it stores over from src, so once both ping-pong buffers have received the
        // Dirichlet values (iterations 1 and 2) every later call would rewrite identical data.
        if (it <= 2) phase_bc_k(launch(false), G, dst, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
      }
    };
    if (converge_check) { itr = run_checked(itr_max, res, body); done_iters = h_state[0].itr; }
    else { for (int it = 1; it <= itr_max; ++it) body(it); join_comm_stream(); itr = itr_max + 1; done_iters = itr_max; }
    if (done_iters & 1) {
      launch_copy_inner(launch(false), G, X, W);
      if (converge_check) phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
    }
    if (ovl) slab_overlap_finish(X, 2, done_iters);          // X leaves with the halo of its final state
    else if (done_iters & 1) Comm_S(X, 2);                   // the last exchange filled WRK's guard planes
  } else {
    auto body = [&](int it) {
      Launch L = launch();
      for (int color = 0; color < 2; ++color) {
        {
          Phase ph(maf ? "SOR2SMA_MAF_kernel" : "SOR2SMA_kernel", (maf ? 33.0 : 9.0) * G.cells());
          if (maf) launch_rb_color_maf(L, G, X, B, axes(), ac1, ip + color, rs, color == 1);   // cz_Poisson.cpp:190-200
          else     launch_rb_color(L, G, X, B, cf, ac1, ip + color, rs, color == 1);
        }
        Comm_S(X, 1);
      }
      if (converge_check) {
        Comm_SUM(0);
        launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
        if (it <= 1) phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);   // in place: boundary nodes never change afterwards
      }
    };
    if (converge_check) { itr = run_checked(itr_max, res, body); done_iters = h_state[0].itr; }
    else { for (int it = 1; it <= itr_max; ++it) body(it); itr = itr_max + 1; done_iters = itr_max; }
  }
  flop += (maf ? 66.0 : 18.0) * G.cells() * done_iters;
  return itr;
}

static double pcr_iter_flop(const Geo& G, int pn, bool last4, double per4) {
  const double nk = G.ked - G.kst + 1;
  const double lines = (double)(G.jed - G.jst + 1) * (G.ied - G.ist + 1);
  const double f = last4 ? (nk * 6.0 + nk * (pn - 2) * 14.0 + ipow2(pn - 2) * per4 + nk * 6.0 + 6.0)
                         : (nk * 6.0 + nk * (pn - 1) * 14.0 + ipow2(pn - 1) * 9.0 + nk * 6.0 + 6.0);
  return lines * f;
}

// cz_maf.f90:493-505 (per iteration = both colours / all lines); per2 = 11 (pcr_rb_maf, pcr_rb_esa_maf, pcr_maf)
// or 9 (pcr_eda_maf, pcr_esa_maf)
static double pcr_maf_iter_flop(const Geo& G, int pn, double per2) {
  const double nk = G.ked - G.kst + 1;
  const double lines = (double)(G.jed - G.jst + 1) * (G.ied - G.ist + 1);
  return lines * ((24.0 + 3.0 * 2.0 + 12.0) + nk * (11.0 + 10.0) + (nk - 2.0) * 6.0 + nk * (pn - 1) * 16.0 +
                  ipow2(pn - 1) * per2 + nk * 6.0);
}

static REAL* ensure_mask(czb_ctx* c) {
  if (!c->MSK) {
    c->field(c->MSK);
    launch_imask(c->launch(false), c->G, c->MSK);   // cz_Evaluate.cpp:390
  }
  return c->MSK;
}

// CZ::LSOR_PCR_RB (cz_Poisson.cpp:518-608) and LSOR_PCR_RB_ESA (:621-733, last4)
int czb_ctx::LSOR_PCR_RB(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool last4, bool converge_check) {
  const int n = G.ked - G.kst + 1;
  const int pn = getNumStage(n);
  if (pn < 0) { printf("error : number of stage\n"); return 0; }
  PcrTables& T = last4 ? pcr4 : pcr2;
  if (!maf) pcr_build_tables(rt().stream, T, n, pn, last4);
  REAL* M = ensure_mask(this);
  const int ofs = (head[0] + head[1]) & 1;   // global (i+j) parity; 0 on a single domain
  const int rs = converge_check ? 0 : 7;   // reduction slot (7 = scratch inside the preconditioner)
  auto body = [&](int it) {
    Launch L = launch();
    for (int color = 0; color < 2; ++color) {
      // MAF (cz_Poisson.cpp:559-570, 684-695): line-dependent coefficients, 2x2 last stage in both variants
      {
        Phase ph(maf ? "PCR_RB_MAF" : "PCR_RB", 0.5 * (maf ? pcr_maf_iter_flop(G, pn, 11.0) : pcr_iter_flop(G, pn, last4, 78.0)));
        if (maf) launch_pcr_lines_maf(L, G, pn, color ^ ofs, 0, axes(), X, M, B, ac1, rs, color == 1);
        else     launch_pcr_lines(L, G, T, color, ofs, 0, X, X, M, B, ac1, rs, color == 1, true);
      }
      Comm_S(X, 1);
    }
    if (converge_check) {
      Comm_SUM(0);
      launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
      if (it <= 1) phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);   // in place: boundary nodes never change afterwards
    }
  };
  int itr, done_iters;
  if (converge_check) { itr = run_checked(itr_max, res, body); done_iters = h_state[0].itr; }
  else { for (int it = 1; it <= itr_max; ++it) body(it); itr = itr_max + 1; done_iters = itr_max; }
  flop += (maf ? pcr_maf_iter_flop(G, pn, 11.0) : pcr_iter_flop(G, pn, last4, 78.0)) * done_iters;
  return itr;
}

// CZ::LSOR_PCR (:745-816, 4x4), LSOR_PCR_ESA (:910-1005, 4x4), LSOR_PCR_EDA (:828-898, 2x2):
// all lines in (j,i) order in place = Gauss-Seidel between lines; reproduced in the reference's
// 1-thread order by wavefronts over the anti-diagonals i+j = const.
int czb_ctx::LSOR_PCR(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool last4, bool converge_check) {
  const int n = G.ked - G.kst + 1;
  const int pn = getNumStage(n);
  if (pn < 0) { printf("error : number of stage\n"); return 0; }
  PcrTables& T = last4 ? pcr4 : pcr2;
  if (!maf) pcr_build_tables(rt().stream, T, n, pn, last4);
  REAL* M = ensure_mask(this);
  const int rs = converge_check ? 0 : 7;   // reduction slot (7 = scratch inside the preconditioner)
  auto body = [&](int it) {
    Launch L = launch();
    bool first = true;
    {
      Phase ph(maf ? "PCR_MAF" : "PCR", maf ? pcr_maf_iter_flop(G, pn, last4 ? 11.0 : 9.0) : pcr_iter_flop(G, pn, last4, 74.0));
      for (int d = G.ist + G.jst; d <= G.ied + G.jed; ++d) {
        if (maf) launch_pcr_lines_maf(L, G, pn, 3, d, axes(), X, M, B, ac1, rs, !first);   // cz_Poisson.cpp:775-782, 858-865, 963-971
        else     launch_pcr_lines(L, G, T, 3, 0, d, X, X, M, B, ac1, rs, !first, true);
        first = false;
      }
    }
    Comm_S(X, 1);
    if (converge_check) {
      Comm_SUM(0);
      launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
      if (it <= 1) phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);   // in place: boundary nodes never change afterwards
    }
  };
  int itr, done_iters;
  if (converge_check) { itr = run_checked(itr_max, res, body); done_iters = h_state[0].itr; }
  else { for (int it = 1; it <= itr_max; ++it) body(it); itr = itr_max + 1; done_iters = itr_max; }
  flop += (maf ? pcr_maf_iter_flop(G, pn, last4 ? 11.0 : 9.0) : pcr_iter_flop(G, pn, last4, 74.0)) * done_iters;
  return itr;
}

// CZ::LSOR_PCR_J_ESA (cz_Poisson.cpp:1008-1095): line Jacobi
int czb_ctx::LSOR_PCR_J_ESA(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check) {
  const int n = G.ked - G.kst + 1;
  const int pn = getNumStage(n);
  if (pn < 0) { printf("error : number of stage\n"); return 0; }
  pcr_build_tables(rt().stream, pcr2, n, pn, false);
  REAL* M = ensure_mask(this);
  REAL* W = field(WRK);
  wrk_guards_zero = false;
  const int rs = converge_check ? 0 : 7;   // reduction slot (7 = scratch inside the preconditioner)
  int itr, done_iters;
  if (converge_check) {
    // The reference writes the new lines to wrk and copies them back (cz_solver.f90:1632-1665); here X and WRK
    // alternate as source and destination (both hold the same boundary / guard values), as in JACOBI.
    launch_copy_shell(launch(false), G, W, X);   // guard cells + boundary nodes; the lines' cells are all stored by the first pass
    auto body = [&](int it) {
      REAL* src = (it & 1) ? X : W;
      REAL* dst = (it & 1) ? W : X;
      Launch L = launch();
      { Phase ph("PCR_J", pcr_iter_flop(G, pn, false, 0.0)); launch_pcr_lines(L, G, pcr2, 2, 0, 0, src, dst, M, B, ac1, rs, false, true); }
      Comm_S(dst, 1);
      Comm_SUM(0);
      launch_finish_iteration(L, d_state, 0, res_normal, eps, d_hist, hist_cap);
      if (it <= 2) phase_bc_k(launch(false), G, dst, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
    };
    itr = run_checked(itr_max, res, body);
    done_iters = h_state[0].itr;
    if (done_iters & 1) {
      launch_copy_inner(launch(false), G, X, W);
      phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);
      Comm_S(X, 1);   // the last exchange filled WRK's guard planes
    }
  } else {
    for (int it = 1; it <= itr_max; ++it) {
      Launch L = launch();
      launch_pcr_lines(L, G, pcr2, 2, 0, 0, X, W, M, B, ac1, rs, false, true);
      launch_copy_inner(L, G, X, W);
      Comm_S(X, 1);
    }
    itr = itr_max + 1; done_iters = itr_max;
  }
  flop += pcr_iter_flop(G, pn, false, 0.0) * done_iters;
  return itr;
}

// CZ::Preconditioner (cz_Poisson.cpp:273-322): lc_max = 8 sweeps without convergence check,
// starting from the caller-cleared xx; unknown / "none" -> copy.
void czb_ctx::Preconditioner(REAL* xx, REAL* bb, double& flop, int pc_type) {
  NvtxRange range("Preconditioner");
  double res = 0.0;
  const int lc_max = 8;
  const bool outer_maf = maf;
  maf = pc_type >= LS_PSOR_MAF;   // the preconditioner's own kind decides, whatever the outer operator is (:282-317)
  switch (pc_type) {
    case LS_JACOBI:     case LS_JACOBI_MAF:     JACOBI(res, xx, bb, lc_max, flop, false); break;
    case LS_PSOR:       case LS_PSOR_MAF:       PSOR(res, xx, bb, lc_max, flop, false); break;
    case LS_SOR2SMA:    case LS_SOR2SMA_MAF:    RBSOR(res, xx, bb, lc_max, flop, false); break;
    case LS_PCR:        case LS_PCR_MAF:        LSOR_PCR(res, xx, bb, lc_max, flop, true, false); break;
    case LS_PCR_EDA:    case LS_PCR_EDA_MAF:    LSOR_PCR(res, xx, bb, lc_max, flop, false, false); break;
    case LS_PCR_RB:     case LS_PCR_RB_MAF:     LSOR_PCR_RB(res, xx, bb, lc_max, flop, false, false); break;
    case LS_PCR_RB_ESA: case LS_PCR_RB_ESA_MAF: LSOR_PCR_RB(res, xx, bb, lc_max, flop, true, false); break;
    default:
      CZB_CHECK(cudaMemcpyAsync(xx, bb, sizeof(REAL) * field_elems(), cudaMemcpyDeviceToDevice, rt().stream));
  }
  maf = outer_maf;
}

// CZ::PBiCGSTAB (cz_Poisson.cpp:332-504).  Scalars rho/alpha/beta/omega live in LoopState.scal
// on the device (k_bicg_scalar), the vector kernels read them from there.
int czb_ctx::PBiCGSTAB(double& res, REAL* X, REAL* B, int ItrMax, double& flop, int pc_type, bool maf_op) {
  Runtime& R = rt();
  // pbicgstab_maf (:350-358, 415-419, 451-455): the operator is calc_ax_maf / calc_rk_maf scaled by pvt
  const MafAxes MA = maf_op ? axes() : MafAxes{nullptr, nullptr, nullptr};
  const REAL* PV = maf_op ? pivot() : nullptr;
  const size_t bytes = sizeof(REAL) * field_elems();
  REAL *p = field(pcg_p), *p_ = field(pcg_p_), *r = field(pcg_r), *r0 = field(pcg_r0), *q = field(pcg_q),
       *s = field(pcg_s), *s_ = field(pcg_s_), *t_ = field(pcg_t_);
  double* sc = d_state->scal;
  double pflop = 0.0;
  // With the packed two-colour kernel as preconditioner the first of the 8 sweeps treats its source as zero
  // without reading it, so p_ / s_ need no clearing: every cell the sweeps do not store stays at the zero
  // the field was allocated with (czb_malloc), every stored cell is rewritten by sweep 2.
  const bool zs = (pc_type == LS_SOR2SMA) && variant == 0 && !(comm && comm->brick) && rb2_zero_src_supported(G, cf);   // bricks sweep per colour and read p_

  begin_loop(ItrMax);
  {
    Launch L = launch();
    const double nc = G.cells();
    { Phase ph("Blas_Clear"); CZB_CHECK(cudaMemsetAsync(q, 0, bytes, R.stream)); }   // :344
    {
      Phase ph("Blas_Residual", (maf_op ? 63.0 : 14.0) * nc);
      if (maf_op) launch_calc_rk_maf(L, G, r, X, B, MA, PV);                 // :352
      else        launch_calc_rk(L, G, r, X, B, cf);                         // :356
    }
    Comm_S(r, 1);                                                            // :362
    { Phase ph("Blas_Copy"); CZB_CHECK(cudaMemcpyAsync(r0, r, bytes, cudaMemcpyDeviceToDevice, R.stream)); }   // :365
    k_bicg_scalar<<<1, 1, 0, R.stream>>>(d_state, -1, res_normal, eps, d_hist, hist_cap);
    count_launch();
  }
  auto scalar = [&](int phase) {
    k_bicg_scalar<<<1, 1, 0, R.stream>>>(d_state, phase, res_normal, eps, d_hist, hist_cap);
    count_launch();
  };
  auto body = [&](int itr) {
    Launch L = launch();
    const double nc = G.cells(), f_axp = maf_op ? 63.0 : 13.0;
    // The sections carry the reference's labels (cz_Poisson.cpp:376-477).  Where this library fuses several of the
    // reference's passes into one kernel the time is booked under the FIRST of them and the flop of all of them:
    // Blas_AX includes its dot products (Dot2 / Dot1), Blas_BiCG_2 includes the following TRIAD, Dot1 and Dot2.
    if (itr == 1 || variant != 0) { Phase ph("Dot2", 2.0 * nc); launch_dot2(L, G, r, r0, 0); }   // rho :376-383 (later: fused into the tail)
    Comm_SUM(0); scalar(0);
    if (itr == 1) {
      Phase ph("Blas_Copy");
      CZB_CHECK(cudaMemcpyAsync(p, r, bytes, cudaMemcpyDeviceToDevice, R.stream));   // :388
    } else {
      Phase ph("Blas_BiCG_1", 4.0 * nc);
      launch_bicg1(L, G, p, r, q, 0, 0, &sc[SC_BETA], &sc[SC_OMEGA]);        // :397
    }
    Comm_S(p, 1);                                                            // :402
    if (!zs) { Phase ph("Blas_Clear"); CZB_CHECK(cudaMemsetAsync(p_, 0, bytes, R.stream)); }   // :405 (zs: first sweep does not read p_)
    pre_zero_src = zs; Preconditioner(p_, p, pflop, pc_type); pre_zero_src = false;   // :409
    {
      Phase ph("Blas_AX", (f_axp + 2.0) * nc);
      if (maf_op) launch_calc_ax_maf(L, G, q, p_, MA, PV, r0, 1);            // :417
      else if (!(variant == 0 && launch_ax_fused(L, G, q, p_, cf, r0, 1))) launch_calc_ax(L, G, q, p_, cf, r0, 1);
    }
    Comm_SUM(1); scalar(1);          // q = A p_, alpha :421-431
    { Phase ph("Blas_TRIAD", 2.0 * nc); launch_triad(L, G, s, q, r, 0, &sc[SC_NEG_ALPHA]); }   // s = r - alpha q :434
    Comm_S(s, 1);                                                            // :438
    if (!zs) { Phase ph("Blas_Clear"); CZB_CHECK(cudaMemsetAsync(s_, 0, bytes, R.stream)); }   // :441
    pre_zero_src = zs; Preconditioner(s_, s, pflop, pc_type); pre_zero_src = false;   // :445
    {   // t = A s_, (t,s), (t,t) :457-464
      Phase ph("Blas_AX", (f_axp + 4.0) * nc);
      const int f = (variant == 0 && !maf_op) ? launch_ax_fused(L, G, t_, s_, cf, s, 2, 3) : 0;
      if (maf_op) launch_calc_ax_maf(L, G, t_, s_, MA, PV, s, 2);            // :453
      else if (f == 0) launch_calc_ax(L, G, t_, s_, cf, s, 2);
      if (f != 2) launch_dot1(L, G, t_, 3);
    }
    Comm_SUM(2, 2); scalar(2);                                               // omega
    if (variant == 0) {   // x += alpha p_ + omega s_ :470, r = s - omega t :476, (r,r) :481, (r,r0) :376 in one pass
      Phase ph("Blas_BiCG_2", (4.0 + 2.0 + 2.0 + 2.0) * nc);
      launch_bicg_tail(L, G, X, p_, s_, r, t_, s, r0, &sc[SC_ALPHA], &sc[SC_OMEGA], &sc[SC_NEG_OMEGA], 4, 0);
    } else {
      { Phase ph("Blas_BiCG_2", 4.0 * nc); launch_bicg2(L, G, X, p_, s_, 0, 0, &sc[SC_ALPHA], &sc[SC_OMEGA]); }
      { Phase ph("Blas_TRIAD", 2.0 * nc); launch_triad(L, G, r, t_, s, 0, &sc[SC_NEG_OMEGA]); }
      { Phase ph("Dot1", 2.0 * nc); launch_dot1(L, G, r, 4); }
    }
    Comm_S(X, 1); Comm_SUM(4);                                               // :486-488
    scalar(3);                                                               // :490-500
    phase_bc_k(launch(false), G, X, pitch, bc_origin, nID, head[1] - 1, head[0] - 1);                                // :495
  };
  // for (itr=1; itr<ItrMax; itr++)  (cz_Poisson.cpp:373)
  const int n_it = ItrMax - 1;
  const int batch = 4;
  int launched = 0, nb = 0;
  bool stop = false;
  while (launched < n_it && !stop) {
    const int m = (n_it - launched < batch) ? n_it - launched : batch;
    for (int k = 0; k < m; ++k) body(launched + k + 1);
    launched += m;
    const int slot = nb & 1;
    CZB_CHECK(cudaMemcpyAsync(&h_state[1 + slot], d_state, sizeof(LoopState), cudaMemcpyDeviceToHost, R.stream));
    CZB_CHECK(cudaEventRecord(ev[slot], R.stream));
    if (nb >= 1) {
      const int ps = (nb - 1) & 1;
      CZB_CHECK(cudaEventSynchronize(ev[ps]));
      if (h_state[1 + ps].done) stop = true;
    }
    ++nb;
  }
  const int itr = end_loop(ItrMax, res, true);
  const int done_iters = h_state[0].itr;
  const double f_ax = maf_op ? 63.0 : 13.0, f_rk = maf_op ? 63.0 : 14.0;
  flop += (f_rk + (double)done_iters * (2 + 4 + f_ax + 2 + 2 + f_ax + 4 + 4 + 2 + 2)) * G.cells();
  if (launched > 0) flop += pflop * ((double)done_iters / (double)launched);
  return itr;
}

// ============================================================================ C entry points
extern "C" {

int czb_slab_decompose(const int gsz[3], int nranks, int rank, int sz[3], int head[3], int nID[6], int idx[6]) {
  return czb::slab_decompose(gsz, nranks, rank, sz, head, nID, idx);
}

int czb_brick_decompose(const int gsz[3], const int gdv[3], int rank, int sz[3], int head[3], int nID[6], int idx[6]) {
  return czb::brick_decompose(gsz, gdv, rank, sz, head, nID, idx);
}

czb_ctx* czb_ctx_create(const int sz[3], const int idx[6], const int nID[6], CZB_REAL dh, const CZB_REAL org[3]) {
  Runtime& R = rt();
  czb_ctx* c = new czb_ctx();
  memset((void*)c, 0, sizeof(czb_ctx));
  c->gc = 2;
  for (int a = 0; a < 3; ++a) { c->size[a] = sz[a]; c->head[a] = 1; c->origin[a] = org ? org[a] : (REAL)0; c->bc_origin[a] = c->origin[a]; }
  for (int f = 0; f < 6; ++f) c->nID[f] = nID ? nID[f] : -1;
  if (idx) { for (int a = 0; a < 6; ++a) c->innerFidx[a] = idx[a]; }
  else {   // CZ::range_inner_index, cz_miscel.cpp:20-52
    c->innerFidx[0] = 2; c->innerFidx[2] = 2; c->innerFidx[4] = 2;
    c->innerFidx[1] = c->nID[1] < 0 ? sz[0] - 1 : sz[0];
    c->innerFidx[3] = c->nID[3] < 0 ? sz[1] - 1 : sz[1];
    c->innerFidx[5] = c->nID[5] < 0 ? sz[2] - 1 : sz[2];
  }
  c->pitch = dh;
  for (int a = 0; a < 6; ++a) c->cf[a] = (REAL)1.0;   // cz.h:169-172
  c->cf[6] = (REAL)6.0;
  c->eps = 1.0e-5;                                    // cz.h:162
  c->G = make_geo(c->size, c->innerFidx, c->gc);
  c->res_normal = 1.0 / c->G.cells();
  c->wrk_guards_zero = false;
  c->maf = false; c->mesh_user = false; c->xc = c->yc = c->zc = c->pvt = nullptr;
  c->pre_zero_src = false; c->cs = nullptr; c->cs_used = false;
  memset((void*)&c->ov, 0, sizeof(c->ov));
  CZB_CHECK(cudaMalloc(&c->d_state, sizeof(LoopState)));
  CZB_CHECK(cudaMemset(c->d_state, 0, sizeof(LoopState)));
  CZB_CHECK(cudaMallocHost(&c->h_state, sizeof(LoopState) * 4));
  CZB_CHECK(cudaEventCreate(&c->ev[0])); CZB_CHECK(cudaEventCreate(&c->ev[1]));
  CZB_CHECK(cudaEventCreate(&c->ev_t0)); CZB_CHECK(cudaEventCreate(&c->ev_t1));
  (void)R;
  return c;
}

void czb_ctx_destroy(czb_ctx* c) {
  if (!c) return;
  rt().sync();
  REAL* fields[] = {c->WRK, c->MSK, c->pcg_p, c->pcg_p_, c->pcg_r, c->pcg_r0, c->pcg_q, c->pcg_s, c->pcg_s_,
                    c->pcg_t_, c->Xd, c->Bd, c->pvt, c->xc, c->yc, c->zc};
  for (REAL* f : fields) if (f) cudaFree(f);
  pcr_free_tables(c->pcr2); pcr_free_tables(c->pcr4);
  if (c->d_hist) cudaFree(c->d_hist);
  cudaFree(c->d_state); cudaFreeHost(c->h_state);
  cudaEventDestroy(c->ev[0]); cudaEventDestroy(c->ev[1]); cudaEventDestroy(c->ev_t0); cudaEventDestroy(c->ev_t1);
  if (c->cs) {
    cudaStreamSynchronize(c->cs);
    cudaStreamSynchronize(c->ov.es);
    cudaEvent_t evs[] = {c->ev_fin[0], c->ev_fin[1], c->ev_join, c->ov.ev_a[0], c->ov.ev_a[1], c->ov.ev_b[0], c->ov.ev_b[1],
                         c->ov.ev_start, c->ov.ev_join_e};
    for (cudaEvent_t e : evs) cudaEventDestroy(e);
    cudaStreamDestroy(c->cs); cudaStreamDestroy(c->ov.es);
    cudaFree(c->ov.rb_edge.partial); cudaFree(c->ov.rb_edge.ticket);
  }
  comm_destroy(c->comm);
  delete c;
}

void czb_ctx_set_eps(czb_ctx* c, double eps) { c->eps = eps; }
void czb_ctx_set_res_normal(czb_ctx* c, double v) { c->res_normal = v; }
void czb_ctx_set_head(czb_ctx* c, const int head[3]) {
  for (int a = 0; a < 3; ++a) c->head[a] = head[a];
  // Dirichlet data in global coordinates: y = G_origin + dh*(j_global-1) with G_origin = 0 (cz.h),
  // instead of the reference's origin + dh*(j_local-1), whose rounding depends on the decomposition
  c->bc_origin[1] = (head[1] > 1) ? (REAL)0 : c->origin[1];
  c->bc_origin[0] = (head[0] > 1) ? (REAL)0 : c->origin[0];   // the same along i on bricks
}

void czb_ctx_bc_k(czb_ctx* c, CZB_REAL* p) {
  Runtime& R = rt();
  Arg<REAL> ap(p, c->field_elems(), true, true);
  launch_bc_k(R.launch(), c->G, ap.dev(), c->pitch, c->bc_origin, c->nID, c->head[1] - 1, c->head[0] - 1);
}
void czb_ctx_set_variant(czb_ctx* c, int v) { c->variant = v; }
void czb_ctx_set_mesh(czb_ctx* c, const CZB_REAL* x, const CZB_REAL* y, const CZB_REAL* z) { c->set_mesh(x, y, z); }
double czb_last_solve_ms(czb_ctx* c) { return c->last_ms; }

int czb_comm_unique_id(void* id128) { return comm_unique_id(id128); }

int czb_ctx_comm_init(czb_ctx* c, int nranks, int rank, const void* id128) {
  rt();
  if (c->comm) comm_destroy(c->comm);
  c->comm = comm_create(nranks, rank, id128, c->nID);
  comm_warmup(c->comm, rt().stream);
  return 0;
}

static double read_state_acc(czb_ctx* c, int slot) {
  Runtime& R = rt();
  CZB_CHECK(cudaMemcpyAsync(&c->h_state[0], c->d_state, sizeof(LoopState), cudaMemcpyDeviceToHost, R.stream));
  R.sync();
  return c->h_state[0].acc[slot];
}

int czb_sweep_rbsor(czb_ctx* c, const CZB_REAL* x, CZB_REAL* xnew, const CZB_REAL* b, CZB_REAL omg, double* res) {
  Launch L = c->launch(false);
  const int jr0 = c->G.jst - (c->nID[2] >= 0 ? 1 : 0), jr1 = c->G.jed + (c->nID[3] >= 0 ? 1 : 0);
  if (!launch_rb2_fused(L, c->G, x, xnew, b, c->cf, omg, c->parity_ip(), jr0, jr1, 0, false)) return 0;
  if (res) *res = read_state_acc(c, 0);
  return 1;
}

int czb_sweep_jacobi(czb_ctx* c, const CZB_REAL* x, CZB_REAL* xnew, const CZB_REAL* b, CZB_REAL omg, double* res) {
  Launch L = c->launch(false);
  if (!launch_jacobi_fused(L, c->G, x, xnew, b, c->cf, omg, 0, false)) return 0;
  if (res) *res = read_state_acc(c, 0);
  return 1;
}

int czb_apply_ax(czb_ctx* c, CZB_REAL* ap, const CZB_REAL* p, const CZB_REAL* w, double* dot_w, double* dot_self) {
  Launch L = c->launch(false);
  const int f = launch_ax_fused(L, c->G, ap, p, c->cf, w, 0, dot_self ? 1 : -1);
  if (f == 0) return 0;
  if (dot_self && f != 2) launch_dot1(L, c->G, ap, 1);
  if (dot_w && w) *dot_w = read_state_acc(c, 0);
  if (dot_self) *dot_self = read_state_acc(c, 1);
  return 1;
}

int czb_solve(czb_ctx* c, const char* solver, const char* precond, CZB_REAL* X, CZB_REAL* B, int itr_max,
              CZB_REAL omg, double* res_out, double* hist, int hist_len, double* flop_out) {
  Runtime& R = rt();
  const int ls = czb_parse_solver(solver);
  if (ls == 0) { printf("Invalid solver\n"); return 0; }           // CZ::setLS, cz_Evaluate.cpp:799-802
  int pc = 0;
  if (ls == LS_BICGSTAB || ls == LS_BICGSTAB_MAF) {
    pc = czb_parse_solver(precond);
    if (pc == 0 && precond) printf("precon=%s\n", precond);        // CZ::setStrPre falls through to copy
    // names CZ::setStrPre does not know (cz_Evaluate.cpp:571-619) leave pc_type unset -> copy
    if (pc == LS_BICGSTAB || pc == LS_BICGSTAB_MAF || pc == LS_PCR_ESA || pc == LS_PCR_J_ESA || pc == LS_PCR_ESA_MAF) pc = 0;
  }
  {   // line solvers need whole k-lines: a division along z cuts them (the reference has no distributed PCR either)
    auto is_line = [](int v) { return v == LS_PCR || v == LS_PCR_EDA || v == LS_PCR_ESA || v == LS_PCR_RB || v == LS_PCR_RB_ESA ||
                                      v == LS_PCR_J_ESA || v == LS_PCR_MAF || v == LS_PCR_EDA_MAF || v == LS_PCR_ESA_MAF ||
                                      v == LS_PCR_RB_MAF || v == LS_PCR_RB_ESA_MAF; };
    if ((c->nID[4] >= 0 || c->nID[5] >= 0) && (is_line(ls) || is_line(pc))) {
      printf("\tLine solvers (pcr*) need undivided k-lines: use gdv_z = 1.\n");
      return 0;
    }
  }
  c->maf = (ls >= LS_PSOR_MAF && ls != LS_BICGSTAB_MAF);
  NvtxRange range(solver);
  c->ac1 = omg;
  const size_t bytes = sizeof(REAL) * c->field_elems();
  REAL* Xdev = X; REAL* Bdev = B;
  const bool xh = !is_device_ptr(X), bh = !is_device_ptr(B);
  if (xh) { Xdev = c->field(c->Xd); CZB_CHECK(cudaMemcpyAsync(Xdev, X, bytes, cudaMemcpyHostToDevice, R.stream)); R.h2d_bytes += bytes; }
  if (bh) { Bdev = c->field(c->Bd); CZB_CHECK(cudaMemcpyAsync(Bdev, B, bytes, cudaMemcpyHostToDevice, R.stream)); R.h2d_bytes += bytes; }
  double res = 0.0, flop = 0.0;
  int itr = 0;
  switch (ls) {
    case LS_JACOBI:     itr = c->JACOBI(res, Xdev, Bdev, itr_max, flop); break;
    case LS_PSOR:       itr = c->PSOR(res, Xdev, Bdev, itr_max, flop); break;
    case LS_SOR2SMA:    itr = c->RBSOR(res, Xdev, Bdev, itr_max, flop); break;
    case LS_BICGSTAB:   itr = c->PBiCGSTAB(res, Xdev, Bdev, itr_max, flop, pc); break;
    case LS_PCR:        itr = c->LSOR_PCR(res, Xdev, Bdev, itr_max, flop, true); break;
    case LS_PCR_ESA:    itr = c->LSOR_PCR(res, Xdev, Bdev, itr_max, flop, true); break;
    case LS_PCR_EDA:    itr = c->LSOR_PCR(res, Xdev, Bdev, itr_max, flop, false); break;
    case LS_PCR_RB:     itr = c->LSOR_PCR_RB(res, Xdev, Bdev, itr_max, flop, false); break;
    case LS_PCR_RB_ESA: itr = c->LSOR_PCR_RB(res, Xdev, Bdev, itr_max, flop, true); break;
    case LS_PCR_J_ESA:  itr = c->LSOR_PCR_J_ESA(res, Xdev, Bdev, itr_max, flop); break;
    // stretched-grid twins (cz_Evaluate.cpp:418-480): same loops, MAF kernels (c->maf)
    case LS_JACOBI_MAF:     itr = c->JACOBI(res, Xdev, Bdev, itr_max, flop); break;
    case LS_PSOR_MAF:       itr = c->PSOR(res, Xdev, Bdev, itr_max, flop); break;
    case LS_SOR2SMA_MAF:    itr = c->RBSOR(res, Xdev, Bdev, itr_max, flop); break;
    case LS_BICGSTAB_MAF:   itr = c->PBiCGSTAB(res, Xdev, Bdev, itr_max, flop, pc, true); break;
    case LS_PCR_MAF:        itr = c->LSOR_PCR(res, Xdev, Bdev, itr_max, flop, true); break;
    case LS_PCR_ESA_MAF:    itr = c->LSOR_PCR(res, Xdev, Bdev, itr_max, flop, false); break;
    case LS_PCR_EDA_MAF:    itr = c->LSOR_PCR(res, Xdev, Bdev, itr_max, flop, false); break;
    case LS_PCR_RB_MAF:     itr = c->LSOR_PCR_RB(res, Xdev, Bdev, itr_max, flop, false); break;
    case LS_PCR_RB_ESA_MAF: itr = c->LSOR_PCR_RB(res, Xdev, Bdev, itr_max, flop, true); break;
    default: break;
  }
  c->maf = false;
  const int n_hist = c->h_state[0].itr;
  if (hist && hist_len > 0 && n_hist > 0) {
    const int m = n_hist < hist_len ? n_hist : hist_len;
    CZB_CHECK(cudaMemcpyAsync(hist, c->d_hist, sizeof(double) * m, cudaMemcpyDeviceToHost, R.stream));
    R.d2h_bytes += sizeof(double) * m;
  }
  if (xh) { CZB_CHECK(cudaMemcpyAsync(X, Xdev, bytes, cudaMemcpyDeviceToHost, R.stream)); R.d2h_bytes += bytes; }
  R.sync();
  if (res_out) *res_out = res;
  if (flop_out) *flop_out += flop;
  return itr;
}

}  // extern "C"

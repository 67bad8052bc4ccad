// cubez_b200/csrc/solver.h -- czb_ctx: host-side mirror of the reference's `class CZ`
// (src/cz_cpp/cz.h:80-588) restricted to the state the iteration loops of
// src/cz_cpp/cz_Poisson.cpp need.  Method names and argument meaning follow the reference.
#pragma once
#include "runtime.h"
#include "comm.h"
#include <string>
#include <vector>

// reference src/cz_cpp/cz_Define.h:68-89
enum LinearSolver {
  LS_PSOR = 1, LS_SOR2SMA, LS_BICGSTAB, LS_JACOBI, LS_PCR, LS_PCR_EDA, LS_PCR_ESA, LS_PCR_RB, LS_PCR_RB_ESA,
  LS_PCR_J_ESA, LS_PSOR_MAF, LS_SOR2SMA_MAF, LS_BICGSTAB_MAF, LS_JACOBI_MAF, LS_PCR_MAF, LS_PCR_EDA_MAF,
  LS_PCR_ESA_MAF, LS_PCR_RB_MAF, LS_PCR_RB_ESA_MAF
};

struct czb_ctx {
  // ---- CZ state (cz.h) ----
  int size[3];            // local size
  int innerFidx[6];
  int nID[6];
  int head[3];
  int gc;                 // GUIDE = 2 (cz_Define.h:40)
  REAL pitch;             // dh
  REAL origin[3];
  REAL bc_origin[3];      // origin used for the Dirichlet data (global y origin on j-slabs)
  REAL cf[7];             // {1,1,1,1,1,1,6} (cz.h:169-172)
  REAL ac1;               // omega (cz_Evaluate.cpp:99)
  double eps;             // 1e-5 (cz.h:162)
  double res_normal;      // 1 / (global inner cells) (cz_Evaluate.cpp:222-224)
  int variant;            // 0 auto, 1 baseline kernels only

  czb::Geo G;
  czb::Comm* comm;

  // ---- device work arrays (allocated on first use) ----
  REAL* WRK; REAL* MSK;
  REAL *pcg_p, *pcg_p_, *pcg_r, *pcg_r0, *pcg_q, *pcg_s, *pcg_s_, *pcg_t_;
  REAL* Xd; REAL* Bd;     // staging when the caller passes host pointers
  bool wrk_guards_zero;
  bool pre_zero_src;      // the next preconditioner call may skip reading its (all-zero) start vector
  czb::PcrTables pcr2, pcr4;
  // ---- stretched-grid (MAF) twins: coordinate vectors xc, yc, zc (cz_Evaluate.cpp:246-248, 342-362) and the
  // pivot field pvt (:243, 369) on the device; `maf` selects the MAF kernels inside the loops below
  bool maf;
  REAL *xc, *yc, *zc, *pvt;
  bool mesh_user;         // mesh given through czb_ctx_set_mesh (else the reference driver's uniform mesh)

  czb::LoopState* d_state;
  czb::LoopState* h_state;   // pinned, 2 slots
  double* d_hist; int hist_cap;
  cudaEvent_t ev[2], ev_t0, ev_t1;
  double last_ms;
  // multi-rank overlap (RBSOR): halo exchange / residual all-reduce run on a second, high-priority stream
  cudaStream_t cs; cudaEvent_t ev_fin[2], ev_join;
  bool cs_used;
  // multi-rank overlap of ping-pong sweeps (see slab_overlap_* below)
  struct SlabOverlap {
    cudaStream_t es;             // edge stream (high priority): first + last j-chunk of a sweep, then the halo traffic
    czb::RedBuf rb_edge;         // reduction scratch of the edge launch (it runs next to the interior launch)
    cudaEvent_t ev_a[2], ev_b[2], ev_start, ev_join_e;   // edge / interior launch of iteration parity done
    bool decided, all_ok;        // every rank can split its sweeps (decided once, collectively)
    bool p2p;                    // halo planes travel through NVLink peer memory (comm.h), else NCCL send/recv
    bool deferred;               // this iteration's exchange waits for the interior launch (bc_k in between)
  } ov;
  int hist_len_req;

  // ---- helpers ----
  size_t field_elems() const { return (size_t)G.total; }
  REAL* field(REAL*& slot);
  czb::Launch launch(bool with_state = true);
  void Comm_S(REAL* f, int width = 1);            // cz_comm.cpp:23-38
  void Comm_SUM(int slot, int count = 1);         // cz_comm.cpp:74-140 on device scalars
  // cz_Poisson.cpp:181 for kst = 2; the sweeps start a line at kst + mod(i+j+kp, 2) (cz_solver.f90:466), so a brick whose
  // k range starts at kst = 1 (lower k neighbour) shifts by one to keep the GLOBAL colour i+j+k
  int parity_ip() const { return (head[0] + head[1] + head[2] + 1 + (innerFidx[4] - 2)) & 1; }
  czb::MafAxes axes();                            // device mesh, created on first use
  void set_mesh(const REAL* x, const REAL* y, const REAL* z);   // host vectors of size[a] + 2*gc entries
  REAL* pivot();                                  // search_pivot (cz_blas.f90:947-1039), computed on first use

  // ---- iteration loops (cz_Poisson.cpp) ----
  int JACOBI(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check = true);
  int PSOR(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check = true);
  int RBSOR(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check = true);
  int LSOR_PCR_RB(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool last4, bool converge_check = true);
  int LSOR_PCR(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool last4, bool converge_check = true);
  int LSOR_PCR_J_ESA(double& res, REAL* X, REAL* B, int itr_max, double& flop, bool converge_check = true);
  void Preconditioner(REAL* xx, REAL* bb, double& flop, int pc_type);
  int PBiCGSTAB(double& res, REAL* X, REAL* B, int ItrMax, double& flop, int pc_type, bool maf_op = false);

  // generic driver: enqueue `body(it)` for it = 1..itr_max in batches, polling the device
  // state without stalling the queue.  Returns the reference's iteration count.
  template <class Body> int run_checked(int itr_max, double& res, Body body);
  void ensure_comm_stream();
  bool persist_ok(int nfields);                    // single rank, plain kernels, fields fit L2 (persist.cu)
  void join_comm_stream();                        // the solve stream waits for everything queued on cs
  // ---- SlabOverlap: a ping-pong sweep of a j-slab as two launches of the single-GPU kernel -----------------------
  // edge chunks (part 1) on ov.es, then the `w` outermost planes of dst go to the neighbours (peer DMA or NCCL) while
  // the interior chunks (part 2) run on the solve stream; the residual all-reduce + convergence test run on cs.
  bool slab_overlap_begin(bool local_ok, int w);                       // collective; false = use the plain loop
  void slab_overlap_edges_begin(int it, REAL* src, int w, bool converge_check, czb::Launch& LE, czb::Launch& LM);
  void slab_overlap_edges_end(int it, REAL* dst, int w, bool defer_exchange);
  void slab_overlap_iteration_end(int it, REAL* dst, int w, bool converge_check, int slot_a, int slot_b);
  void slab_overlap_finish(REAL* X, int w, int done_iters);
  void slab_exchange(REAL* dst, int w, int it);
  void begin_loop(int itr_max);
  int end_loop(int itr_max, double& res, bool bicg);
};

int czb_parse_solver(const char* name);   // CZ::setLS / setStrPre names -> LinearSolver, 0 unknown

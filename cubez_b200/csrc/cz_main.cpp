// cubez_b200/csrc/cz_main.cpp -- `cz` driver: the reference's command line and console/history
// output (src/main.cpp:15-59, src/cz_cpp/cz_Evaluate.cpp:21-567) over libczb200.  All fields
// live in device memory for the whole run; the solver loop is czb_solve (device-resident).
//
//   ./cz gsz_x gsz_y gsz_z linear_solver IterationMax acc_coef [precond] [gdv_x gdv_y gdv_z] [--gpus N]
//
// Multi-GPU (the reference binary is the multi-rank entry: MPI_Init in main.cpp:31-36, CBrick division in
// cz_Evaluate.cpp:103-159): `--gpus N` cuts the grid into N j-slabs, the reference's own division arguments
// `gdv_x gdv_y gdv_z` into bricks (`1 N 1` = the same j-slabs), one PROCESS per GPU.  The driver is its own launcher: the parent makes the NCCL id, re-executes itself N times
// (CZB_RANK / CZB_NRANKS / CZB_NCCL_ID / CZB_DEVICE in the environment), and prints the global "Error max" from the
// ranks' local maxima; rank 0 writes the console output and the history file exactly as the serial run does.
#include "czb200.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <strings.h>
#include <vector>
#include <signal.h>
#include <unistd.h>
#include <sys/wait.h>

typedef CZB_REAL REAL_TYPE;
#define GUIDE 2   // reference src/cz_cpp/cz_Define.h:40

static const char* solver_names[] = {"jacobi", "psor", "sor2sma", "pbicgstab", "pcr_rb", "pcr_rb_esa", "pcr_j_esa",
                                     "pcr", "pcr_eda", "pcr_esa",
                                     // stretched-grid twins on the driver's uniform mesh (cz_Evaluate.cpp:741-797)
                                     "jacobi_maf", "psor_maf", "sor2sma_maf", "pbicgstab_maf", "pcr_rb_maf",
                                     "pcr_rb_esa_maf", "pcr_maf", "pcr_eda_maf", "pcr_esa_maf", nullptr};

// ---- multi-GPU launcher ---------------------------------------------------------------------------------------------
static std::string hex128(const unsigned char* b) {
  static const char* d = "0123456789abcdef";
  std::string s(256, '0');
  for (int i = 0; i < 128; ++i) { s[2 * i] = d[b[i] >> 4]; s[2 * i + 1] = d[b[i] & 15]; }
  return s;
}
static void unhex128(const char* s, unsigned char* b) {
  auto v = [](char c) { return c <= '9' ? c - '0' : c - 'a' + 10; };
  for (int i = 0; i < 128; ++i) b[i] = (unsigned char)(v(s[2 * i]) * 16 + v(s[2 * i + 1]));
}

// Parent of a `--gpus N` run: spawn one child per GPU, collect "errmax i j k" from each through a pipe.
static int launch_ranks(int nranks, char* argv[]) {
  unsigned char id[128];
  if (czb_comm_unique_id(id) != 0) { printf("\tcannot create the NCCL id\n"); return -1; }
  const std::string idhex = hex128(id);
  std::vector<pid_t> pids(nranks);
  std::vector<int> rfd(nranks);
  for (int r = 0; r < nranks; ++r) {
    int fd[2];
    if (pipe(fd) != 0) { perror("pipe"); return -1; }
    const pid_t pid = fork();
    if (pid < 0) { perror("fork"); return -1; }
    if (pid == 0) {
      close(fd[0]);
      char buf[64];
      snprintf(buf, sizeof(buf), "%d", r); setenv("CZB_RANK", buf, 1); setenv("CZB_DEVICE", buf, 1);
      snprintf(buf, sizeof(buf), "%d", nranks); setenv("CZB_NRANKS", buf, 1);
      snprintf(buf, sizeof(buf), "%d", fd[1]); setenv("CZB_RESULT_FD", buf, 1);
      setenv("CZB_NCCL_ID", idhex.c_str(), 1);
      execv("/proc/self/exe", argv);
      perror("execv"); _exit(127);
    }
    close(fd[1]);
    pids[r] = pid; rfd[r] = fd[0];
  }
  // Reap the ranks as they end.  A rank that fails (no such device, out of memory, a refused configuration) would leave
  // its peers waiting in an NCCL rendezvous or exchange for ever -- mpirun's job under the reference; here the launcher
  // gives them CZB_RANK_GRACE_S seconds (default 10) and then ends them.
  int rc = 0, live = nranks;
  double emax = -1.0; int eloc[3] = {0, 0, 0};
  std::vector<char> done(nranks, 0);
  const char* ge = getenv("CZB_RANK_GRACE_S");
  const double grace = ge ? atof(ge) : 10.0;
  double waited = -1.0;                       // seconds since the first failure, < 0: none yet
  bool terminated = false;
  while (live > 0) {
    bool progress = false;
    for (int r = 0; r < nranks; ++r) {
      if (done[r]) continue;
      int st = 0;
      const pid_t w = waitpid(pids[r], &st, WNOHANG);
      if (w == 0) continue;
      done[r] = 1; --live; progress = true;
      char buf[256]; size_t n = 0; ssize_t k;   // the writer is gone: this reads what it left and then end-of-file
      while (n + 1 < sizeof(buf) && (k = read(rfd[r], buf + n, sizeof(buf) - 1 - n)) > 0) n += (size_t)k;
      buf[n] = 0; close(rfd[r]);
      if (w < 0 || !WIFEXITED(st) || WEXITSTATUS(st) != 0) {
        if (rc == 0 && !terminated) fprintf(stderr, "cz: rank %d ended abnormally (status 0x%x)\n", r, st);
        rc = -1;
        if (waited < 0) waited = 0.0;
      }
      double e; int a, b, c;
      if (sscanf(buf, "%lf %d %d %d", &e, &a, &b, &c) == 4 && e > emax) { emax = e; eloc[0] = a; eloc[1] = b; eloc[2] = c; }
    }
    if (live == 0) break;
    if (!progress) {
      usleep(20000);
      if (waited >= 0) waited += 0.02;
    }
    if (waited >= grace && !terminated) {
      for (int r = 0; r < nranks; ++r) if (!done[r]) kill(pids[r], SIGTERM);
      terminated = true; waited = 0.0;
    } else if (terminated && waited >= 5.0) {
      for (int r = 0; r < nranks; ++r) if (!done[r]) kill(pids[r], SIGKILL);
      waited = 0.0;
    }
  }
  if (rc != 0) { printf("\n\tSolver error.\n\n"); return -1; }
  // global node indices (the reference prints LOCAL indices of the rank that holds the maximum, cz_Evaluate.cpp:556-559)
  printf("\nError max = %e at (%d %d %d)\n\n", emax, eloc[0], eloc[1], eloc[2]);
  return 0;
}

int main(int argc, char* argv[]) {
  // `--gpus N` (anywhere) and the reference's trailing `gdv_x gdv_y gdv_z` (main.cpp:19-29, cz_Evaluate.cpp:72-84)
  int gpus = 0;
  std::vector<char*> av;
  for (int i = 0; i < argc; ++i) {
    if (!strcmp(argv[i], "--gpus") && i + 1 < argc) { gpus = atoi(argv[++i]); continue; }
    av.push_back(argv[i]);
  }
  int ac = (int)av.size();
  int gdv[3] = {1, 0, 1};   // gdv[1] = 0: j-slabs over `gpus` processes
  if (ac == 10 || ac == 11) {
    const int gx = atoi(av[ac - 3]), gy = atoi(av[ac - 2]), gz = atoi(av[ac - 1]);
    if (gx < 1 || gy < 1 || gz < 1) { printf("\tInvalid division.\n"); return 0; }
    if (gpus != 0 && gpus != gx * gy * gz) { printf("\tThe number of proceees does not agree with the division size.\n"); return 0; }   // cz_Evaluate.cpp:79-83
    gpus = gx * gy * gz;
    gdv[0] = gx; gdv[1] = gy; gdv[2] = gz;
    ac -= 3;
  }
  if (gpus < 1) gpus = 1;
  if (gdv[1] == 0) gdv[1] = gpus;
  const char* rk = getenv("CZB_RANK");
  const int nranks = rk ? atoi(getenv("CZB_NRANKS")) : 1;
  const int myRank = rk ? atoi(rk) : 0;
  if (ac == 7 || ac == 8) {
    if (gpus > 1 && !rk) return launch_ranks(gpus, argv);
  }
#define Hostonly_ if (myRank == 0)
  argc = ac;
  argv = av.data();
  if (argc != 7 && argc != 8) {   // main.cpp:19-29
    printf("\tUsage : ./cz gsz_x, gsz_y, gsz_z, linear_solver, IterationMax, acc_coef [precond] [gdv_x, gdv_y, gdv_z] [--gpus N]\n");
    printf("\t\tlinear_solver = {jacobi | psor | sor2sma | pbicgstab | pcr | pcr_eda | pcr_esa | pcr_rb | pcr_rb_esa | pcr_j_esa\n");
    printf("\t\t                 | jacobi_maf | psor_maf | sor2sma_maf | pbicgstab_maf | pcr_maf | pcr_eda_maf | pcr_esa_maf | pcr_rb_maf | pcr_rb_esa_maf}\n");
    printf("\t\tprecond = {none | jacobi | psor | sor2sma | pcr | pcr_eda | pcr_rb | pcr_rb_esa | <the same>_maf}\n\n");
    printf("\t$ ./cz 64 64 64 jacobi 4000 0.8\n");
    printf("\t$ ./cz 64 64 64 pbicgstab 4000 1.1 sor2sma\n");
    return 0;
  }
  int G_size[3] = {atoi(argv[1]), atoi(argv[2]), atoi(argv[3])};
  const char* q = argv[4];
  const char* precon = nullptr;
  const bool bicg = !strcasecmp(q, "pbicgstab") || !strcasecmp(q, "pbicgstab_maf");   // cz_Evaluate.cpp:63
  if (bicg) {
    if (argc != 8) { Hostonly_ printf("command line error : pbicgstab\n"); exit(0); }   // cz_Evaluate.cpp:63-70
    precon = argv[7];
  }
  bool known = false;
  for (int i = 0; solver_names[i]; ++i) known |= !strcasecmp(q, solver_names[i]);
  if (!known) { Hostonly_ printf("Invalid solver\n"); exit(0); }              // cz_Evaluate.cpp:799-802
  const int ItrMax = atoi(argv[5]);
  const REAL_TYPE ac1 = (REAL_TYPE)atof(argv[6]);                             // cz_Evaluate.cpp:99
  REAL_TYPE pitch[3];
  pitch[0] = pitch[1] = pitch[2] = (REAL_TYPE)(1.0 / (REAL_TYPE)(G_size[2] - 1));   // :88
  // sub-domain of this rank (cz_Evaluate.cpp:103-159: CBrick setDivision / findOptimalDivision -> czb_brick_decompose;
  // without division arguments the grid is cut along j only, the decomposition the fused sweeps overlap)
  int size[3] = {G_size[0], G_size[1], G_size[2]};
  int head[3] = {1, 1, 1};
  int nID[6] = {-1, -1, -1, -1, -1, -1};
  int innerFidx[6] = {2, size[0] - 1, 2, size[1] - 1, 2, size[2] - 1};          // cz_miscel.cpp:20-52
  if (nranks > 1 && czb_brick_decompose(G_size, gdv, myRank, size, head, nID, innerFidx) != 0) {
    Hostonly_ printf("\tThe grid has fewer inner nodes along an axis than divisions.\n");
    return -1;
  }
  REAL_TYPE origin[3];                                                         // :136-138, G_origin = 0
  for (int a = 0; a < 3; ++a) origin[a] = (REAL_TYPE)0 + (REAL_TYPE)(head[a] - 1) * pitch[a];
  int gc = GUIDE;

  czb_init(-1);                                  // CZB_DEVICE selects the GPU of this rank
  Hostonly_ printf("Iterative Mehtod = %s\n", q);
  Hostonly_ { if (precon) printf("Preconditioner = %s\n", precon); }
  Hostonly_ { if (nranks > 1) printf("Division = %d x %d x %d (%s, one process per GPU)\n", gdv[0], gdv[1], gdv[2], (gdv[0] == 1 && gdv[2] == 1) ? "j-slabs" : "bricks"); }

  std::string fname = std::string(q) + ".txt";                                 // CZ::setLS
  for (auto& ch : fname) ch = (char)tolower(ch);
  FILE* fph = nullptr;
  Hostonly_ {
    fph = fopen(fname.c_str(), "w");
    if (!fph) { printf("\tSorry, can't open file.\n"); exit(0); }
    fprintf(fph, "Itration      Residual\n");                                  // :217
  }

  const double array_size = (double)(size[0] + 2 * GUIDE) * (size[1] + 2 * GUIDE) * (size[2] + 2 * GUIDE);
  const size_t bytes = (size_t)array_size * sizeof(REAL_TYPE);
  const int nfields = 3 + (bicg ? 9 : 1);
  Hostonly_ printf("\n----------\n\n");
  Hostonly_ printf("\t>> Memory required for Solver (device%s) : %6.2f (MB)\n\n", nranks > 1 ? ", rank 0" : "",
                   array_size * sizeof(REAL_TYPE) * nfields / 1.0e6);

  REAL_TYPE* RHS = (REAL_TYPE*)czb_malloc(bytes);
  REAL_TYPE* P = (REAL_TYPE*)czb_malloc(bytes);
  if (!RHS || !P) { printf("\n\tSolver error.\n\n"); return -1; }

  czb_ctx* ctx = czb_ctx_create(size, innerFidx, nID, pitch[0], origin);
  if (nranks > 1) {
    czb_ctx_set_head(ctx, head);
    czb_ctx_set_res_normal(ctx, 1.0 / ((double)(G_size[0] - 2) * (G_size[1] - 2) * (G_size[2] - 2)));   // :222-224
    unsigned char id[128];
    const char* idhex = getenv("CZB_NCCL_ID");
    if (!idhex || strlen(idhex) != 256) { printf("\tCZB_NCCL_ID missing\n"); return -1; }
    unhex128(idhex, id);
    czb_ctx_comm_init(ctx, nranks, myRank, id);
  }
  czb_ctx_bc_k(ctx, P);                                                        // :375  (global y on slabs)
  czb_ctx_bc_k(ctx, RHS);                                                      // :383

  const char* eps_env = getenv("CZ_EPS");
  if (eps_env) czb_ctx_set_eps(ctx, atof(eps_env));
  std::vector<double> hist((size_t)(ItrMax > 0 ? ItrMax : 1) + 2, 0.0);
  double res = 0.0, flop = 0.0;
  const int itr = czb_solve(ctx, q, precon, P, RHS, ItrMax, ac1, &res, hist.data(), (int)hist.size(), &flop);
  if (itr == 0) { Hostonly_ printf("\n\tSolver error.\n\n"); return -1; }     // main.cpp:45-52

  int nh = itr;
  if (bicg) { if (nh > ItrMax - 1) nh = ItrMax - 1; } else { if (nh > ItrMax) nh = ItrMax; }
  Hostonly_ {
    for (int i = 0; i < nh; ++i) fprintf(fph, "%6d, %13.6e\n", i + 1, hist[i]);   // cz_Poisson.cpp:71
    fclose(fph);
    printf("\n=================================\n");
    printf("Iter = %d  Res = %e\n", itr, res);
    printf("=================================\n");
    const double ms = czb_last_solve_ms(ctx);
    const double cells = (double)(G_size[0] - 2) * (G_size[1] - 2) * (G_size[2] - 2);
    printf("\nSolver loop : %.3f ms on device, %.3f GLUPS (%d iterations)\n", ms, ms > 0 ? cells * nh / ms * 1e-6 : 0.0, nh);
    // the reference reports its analytic flop counters per phase through PMlib (profiling.txt, cz_Evaluate.cpp:492-541)
    printf("Flop count  : %.6e (the reference's analytic formulas%s), %.2f GFLOP/s\n", flop,
           nranks > 1 ? ", rank 0" : "", ms > 0 ? flop / ms * 1e-6 : 0.0);
  }

  // per-phase report, the reference's profiling.txt (cz_Evaluate.cpp:506-544): CZB_PROFILE=1
  { const char* pe = getenv("CZB_PROFILE");
    if (pe && atoi(pe) != 0) { Hostonly_ { char t[64]; snprintf(t, sizeof(t), "cz %s (czb200)", q); czb_profile_report("profiling.txt", t); czb_profile_report(nullptr, t); } } }

  // debug_mode==1 (always on, main.cpp:38-42): error against the analytic solution, :550-563
  REAL_TYPE* ERR = (REAL_TYPE*)czb_malloc(bytes);
  double errmax = 0.0; int loc[3];
  char tmp_fname[30];
  sprintf(tmp_fname, "p_%05d.sph", myRank);
  fileout_t_(size, &gc, P, pitch, origin, tmp_fname);                          // :554-555 (written when CZB_FILEOUT=1)
  exact_t_(size, &gc, ERR, pitch, origin);
  err_t_(size, innerFidx, &gc, &errmax, P, ERR, loc);
  if (nranks == 1) {
    printf("\nError max = %e at (%d %d %d)\n\n", errmax, loc[0], loc[1], loc[2]);
  } else {   // the launcher prints the maximum over ranks
    const char* fd = getenv("CZB_RESULT_FD");
    if (fd) {
      char buf[128];
      const int n = snprintf(buf, sizeof(buf), "%.17g %d %d %d\n", errmax, loc[0] + head[0] - 1, loc[1] + head[1] - 1, loc[2] + head[2] - 1);
      if (write(atoi(fd), buf, (size_t)n) != n) perror("write");
    }
  }
  sprintf(tmp_fname, "e_%05d.sph", myRank);
  fileout_t_(size, &gc, ERR, pitch, origin, tmp_fname);                        // :560-561

  czb_free(ERR); czb_free(P); czb_free(RHS);
  czb_ctx_destroy(ctx);
  return 0;
}

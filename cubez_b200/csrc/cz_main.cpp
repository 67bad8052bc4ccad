// cubez_b200/csrc/cz_main.cpp -- `cz` driver: the reference's command line and console/history
// output (src/main.cpp:15-59, src/cz_cpp/cz_Evaluate.cpp:21-567) over libczb200.  All fields
// live in device memory for the whole run; the solver loop is czb_solve (device-resident).
//
//   ./cz gsz_x gsz_y gsz_z linear_solver IterationMax acc_coef [precond]
#include "czb200.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <strings.h>
#include <vector>

typedef CZB_REAL REAL_TYPE;
#define GUIDE 2   // reference src/cz_cpp/cz_Define.h:40

static const char* solver_names[] = {"jacobi", "psor", "sor2sma", "pbicgstab", "pcr_rb", "pcr_rb_esa", "pcr_j_esa",
                                     "pcr", "pcr_eda", "pcr_esa",
                                     // stretched-grid twins on the driver's uniform mesh (cz_Evaluate.cpp:741-797)
                                     "jacobi_maf", "psor_maf", "sor2sma_maf", "pbicgstab_maf", "pcr_rb_maf",
                                     "pcr_rb_esa_maf", "pcr_maf", "pcr_eda_maf", "pcr_esa_maf", nullptr};

int main(int argc, char* argv[]) {
  if (argc != 7 && argc != 8) {   // main.cpp:19-29 (gdv_x/y/z = CBrick division: not supported here)
    printf("\tUsage : ./cz gsz_x, gsz_y, gsz_z, linear_solver, IterationMax, acc_coef [precond]\n");
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
    if (argc != 8) { printf("command line error : pbicgstab\n"); exit(0); }   // cz_Evaluate.cpp:63-70
    precon = argv[7];
  }
  bool known = false;
  for (int i = 0; solver_names[i]; ++i) known |= !strcasecmp(q, solver_names[i]);
  if (!known) { printf("Invalid solver\n"); exit(0); }                         // cz_Evaluate.cpp:799-802
  const int ItrMax = atoi(argv[5]);
  const REAL_TYPE ac1 = (REAL_TYPE)atof(argv[6]);                             // cz_Evaluate.cpp:99
  REAL_TYPE pitch[3];
  pitch[0] = pitch[1] = pitch[2] = (REAL_TYPE)(1.0 / (REAL_TYPE)(G_size[2] - 1));   // :88
  REAL_TYPE origin[3] = {0, 0, 0};
  int size[3] = {G_size[0], G_size[1], G_size[2]};
  int nID[6] = {-1, -1, -1, -1, -1, -1};
  int gc = GUIDE;

  czb_init(-1);
  printf("Iterative Mehtod = %s\n", q);
  if (precon) printf("Preconditioner = %s\n", precon);

  std::string fname = std::string(q) + ".txt";                                 // CZ::setLS
  for (auto& ch : fname) ch = (char)tolower(ch);
  FILE* fph = fopen(fname.c_str(), "w");
  if (!fph) { printf("\tSorry, can't open file.\n"); exit(0); }
  fprintf(fph, "Itration      Residual\n");                                    // :217

  int innerFidx[6] = {2, size[0] - 1, 2, size[1] - 1, 2, size[2] - 1};          // cz_miscel.cpp:20-52
  const double array_size = (double)(size[0] + 2 * GUIDE) * (size[1] + 2 * GUIDE) * (size[2] + 2 * GUIDE);
  const size_t bytes = (size_t)array_size * sizeof(REAL_TYPE);
  const int nfields = 3 + (bicg ? 9 : 1);
  printf("\n----------\n\n");
  printf("\t>> Memory required for Solver (device) : %6.2f (MB)\n\n", array_size * sizeof(REAL_TYPE) * nfields / 1.0e6);

  REAL_TYPE* RHS = (REAL_TYPE*)czb_malloc(bytes);
  REAL_TYPE* P = (REAL_TYPE*)czb_malloc(bytes);
  if (!RHS || !P) { printf("\n\tSolver error.\n\n"); return -1; }

  bc_k_(size, &gc, P, pitch, origin, nID);                                     // :375
  bc_k_(size, &gc, RHS, pitch, origin, nID);                                   // :383

  czb_ctx* ctx = czb_ctx_create(size, innerFidx, nID, pitch[0], origin);
  const char* eps_env = getenv("CZ_EPS");
  if (eps_env) czb_ctx_set_eps(ctx, atof(eps_env));
  std::vector<double> hist((size_t)(ItrMax > 0 ? ItrMax : 1) + 2, 0.0);
  double res = 0.0, flop = 0.0;
  const int itr = czb_solve(ctx, q, precon, P, RHS, ItrMax, ac1, &res, hist.data(), (int)hist.size(), &flop);
  if (itr == 0) { printf("\n\tSolver error.\n\n"); return -1; }                // main.cpp:45-52

  int nh = itr;
  if (bicg) { if (nh > ItrMax - 1) nh = ItrMax - 1; } else { if (nh > ItrMax) nh = ItrMax; }
  for (int i = 0; i < nh; ++i) fprintf(fph, "%6d, %13.6e\n", i + 1, hist[i]);   // cz_Poisson.cpp:71
  fclose(fph);

  printf("\n=================================\n");
  printf("Iter = %d  Res = %e\n", itr, res);
  printf("=================================\n");
  const double ms = czb_last_solve_ms(ctx);
  const double cells = (double)(size[0] - 2) * (size[1] - 2) * (size[2] - 2);
  printf("\nSolver loop : %.3f ms on device, %.3f GLUPS (%d iterations)\n", ms, ms > 0 ? cells * nh / ms * 1e-6 : 0.0, nh);
  // the reference reports its analytic flop counters per phase through PMlib (profiling.txt, cz_Evaluate.cpp:492-541)
  printf("Flop count  : %.6e (the reference's analytic formulas), %.2f GFLOP/s\n", flop, ms > 0 ? flop / ms * 1e-6 : 0.0);

  // debug_mode==1 (always on, main.cpp:38-42): error against the analytic solution, :550-563
  REAL_TYPE* ERR = (REAL_TYPE*)czb_malloc(bytes);
  double errmax = 0.0; int loc[3];
  char tmp_fname[30];
  sprintf(tmp_fname, "p_%05d.sph", 0);
  fileout_t_(size, &gc, P, pitch, origin, tmp_fname);                          // :554-555 (written when CZB_FILEOUT=1)
  exact_t_(size, &gc, ERR, pitch, origin);
  err_t_(size, innerFidx, &gc, &errmax, P, ERR, loc);
  printf("\nError max = %e at (%d %d %d)\n\n", errmax, loc[0], loc[1], loc[2]);
  sprintf(tmp_fname, "e_%05d.sph", 0);
  fileout_t_(size, &gc, ERR, pitch, origin, tmp_fname);                        // :560-561

  czb_free(ERR); czb_free(P); czb_free(RHS);
  czb_ctx_destroy(ctx);
  return 0;
}

/* include/czb200.h -- C ABI of libczb200_f32.so / libczb200_f64.so
 *
 * B200 (sm_100a) implementation of the iterative-solver hot path of kenoogl/CubeZ behind the
 * reference's own kernel boundary.  One library per precision, exactly like the reference
 * (REAL_TYPE = float unless -D_REAL_IS_DOUBLE_, reference src/cz_cpp/cz_Define.h:28-37).
 *
 * PART 1 are the symbols of the reference's Fortran kernel layer, src/cz_cpp/cz_Ffunc.h:16-578:
 * same names (lower case + trailing underscore = the Fortran mangling the reference links by),
 * same argument order, every argument by pointer, void return.  They are callable from C/C++
 * and from Fortran (interface blocks with bind(C,name='jacobi_') and non-value dummies, see
 * INTEGRATION.md).  Array arguments may be DEVICE (cudaMalloc / czb_malloc / managed) pointers
 * -> zero-copy, or plain HOST pointers (what the unmodified reference driver passes) -> staged
 * through a device mirror for the duration of the call.  Scalar in/out arguments (res, flop,
 * dot results) are host memory, as in the reference.  There is no CPU fallback: without a
 * CUDA device every entry point aborts with a message.
 *
 * PART 2 are extensions the reference ABI lacks and a GPU drop-in needs: device allocation,
 * stream selection, and fused device-resident solve loops that reproduce the control flow of
 * CZ::JACOBI / RBSOR / PBiCGSTAB / LSOR_PCR_* (src/cz_cpp/cz_Poisson.cpp) without a host
 * round-trip per kernel.
 */
#ifndef CZB200_H_
#define CZB200_H_
#include <stddef.h>

#ifdef _REAL_IS_DOUBLE_
typedef double CZB_REAL;
#else
typedef float CZB_REAL;
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* ===================================================================== PART 1
 * sz[3] = {ix,jx,kx} local size without guard cells; idx[6] = {ist,ied,jst,jed,kst,ked}
 * inclusive Fortran indices of the computed cells; g = guard width (2); arrays are
 * (kx+2g, ix+2g, jx+2g) with k contiguous (cz_solver.f90:218); cf[7] = {c1..c6, dd};
 * res and flop are ACCUMULATED into, dot results are overwritten.                         */

/* cz_Ffunc.h:16-21  <- cz_solver.f90:22-191   Dirichlet faces where nID[face] < 0 */
void bc_k_(int* sz, int* g, CZB_REAL* p, CZB_REAL* dh, CZB_REAL* org, int* nID);
/* cz_Ffunc.h:23-32  <- cz_solver.f90:284-387  relaxed Jacobi sweep, wk2 = scratch field */
void jacobi_(CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* cf, CZB_REAL* omg, CZB_REAL* b,
             double* res, CZB_REAL* wk2, double* flop);
/* cz_Ffunc.h:35-43  <- cz_solver.f90:207-269  lexicographic point SOR (hyperplane wavefront) */
void psor_(CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* cf, CZB_REAL* omg, CZB_REAL* b,
           double* res, double* flop);
/* cz_Ffunc.h:48-58  <- cz_solver.f90:404-493  one colour of red-black SOR */
void psor2sma_core_(CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* cf, int* ip, int* color,
                    CZB_REAL* omg, CZB_REAL* b, double* res, double* flop);
/* cz_Ffunc.h:60-77  <- cz_solver.f90:497-662  red-black line SOR, PCR pn-1 stages + 2x2 */
void pcr_rb_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, CZB_REAL* x,
             CZB_REAL* msk, CZB_REAL* rhs, CZB_REAL* WA, CZB_REAL* WC, CZB_REAL* WD,
             CZB_REAL* WAA, CZB_REAL* WCC, CZB_REAL* WDD, CZB_REAL* omg, double* res, double* flop);
/* cz_Ffunc.h:79-97  <- cz_solver.f90:1261-1469  as pcr_rb_, pn-2 stages + 4x4 Cramer */
void pcr_rb_esa_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, int* s, CZB_REAL* x,
                 CZB_REAL* msk, CZB_REAL* rhs, CZB_REAL* WA, CZB_REAL* WC, CZB_REAL* WD,
                 CZB_REAL* WAA, CZB_REAL* WCC, CZB_REAL* WDD, CZB_REAL* omg, double* res,
                 double* flop);
/* cz_Ffunc.h:99-114 <- cz_solver.f90:666-878  lexicographic line SOR (line wavefront), 4x4 */
void pcr_(int* sz, int* idx, int* g, int* pn, CZB_REAL* x, CZB_REAL* msk, CZB_REAL* rhs,
          CZB_REAL* WA, CZB_REAL* WC, CZB_REAL* WD, CZB_REAL* WAA, CZB_REAL* WCC, CZB_REAL* WDD,
          CZB_REAL* omg, double* res, double* flop);
/* cz_Ffunc.h:116-128 <- cz_solver.f90:883-1045  as pcr_, pn-1 stages + 2x2 */
void pcr_eda_(int* sz, int* idx, int* g, int* pn, CZB_REAL* x, CZB_REAL* msk, CZB_REAL* rhs,
              CZB_REAL* WA, CZB_REAL* WC, CZB_REAL* WD, CZB_REAL* omg, double* res, double* flop);
/* cz_Ffunc.h:130-146 <- cz_solver.f90:1050-1257 as pcr_ with zero-padded work arrays */
void pcr_esa_(int* sz, int* idx, int* g, int* pn, int* s, CZB_REAL* x, CZB_REAL* msk,
              CZB_REAL* rhs, CZB_REAL* SA, CZB_REAL* SC, CZB_REAL* SD, CZB_REAL* WA, CZB_REAL* WC,
              CZB_REAL* WD, CZB_REAL* omg, double* res, double* flop);
/* cz_Ffunc.h:148-166 <- cz_solver.f90:1473-1676 line Jacobi, pn-1 stages + 2x2 */
void pcr_j_esa_(int* sz, int* idx, int* g, int* pn, int* s, CZB_REAL* x, CZB_REAL* msk,
                CZB_REAL* rhs, CZB_REAL* SA, CZB_REAL* SC, CZB_REAL* SD, CZB_REAL* WA,
                CZB_REAL* WC, CZB_REAL* WD, CZB_REAL* SRC, CZB_REAL* WRK, CZB_REAL* omg,
                double* res, double* flop);

/* cz_Ffunc.h:442-578 <- cz_blas.f90 */
void imask_k_(CZB_REAL* x, int* sz, int* idx, int* g);                       /* :24-103  */
void blas_clear_(CZB_REAL* x, int* sz, int* g);                              /* :112-149 */
void blas_copy_(CZB_REAL* dst, CZB_REAL* src, int* sz, int* g);              /* :159-195 */
void blas_copy_in_(CZB_REAL* dst, CZB_REAL* src, int* sz, int* g);           /* :205-242 */
void blas_triad_(CZB_REAL* z, CZB_REAL* x, CZB_REAL* y, CZB_REAL* a, int* sz, int* idx, int* g,
                 double* flop);                                              /* :255-308 */
void blas_dot1_(CZB_REAL* r, CZB_REAL* p, int* sz, int* idx, int* g, double* flop);   /* :320-373 */
void blas_dot2_(CZB_REAL* r, CZB_REAL* p, CZB_REAL* q, int* sz, int* idx, int* g,
                double* flop);                                               /* :386-437 */
void blas_bicg_1_(CZB_REAL* p, CZB_REAL* r, CZB_REAL* q, CZB_REAL* beta, CZB_REAL* omg, int* sz,
                  int* idx, int* g, double* flop);                           /* :452-502 */
void blas_bicg_2_(CZB_REAL* z, CZB_REAL* x, CZB_REAL* y, CZB_REAL* a, CZB_REAL* b, int* sz,
                  int* idx, int* g, double* flop);                           /* :517-566 */
void blas_calc_ax_(CZB_REAL* ap, CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* cf,
                   double* flop);                                            /* :579-644 */
void blas_calc_rk_(CZB_REAL* r, CZB_REAL* p, CZB_REAL* b, int* sz, int* idx, int* g,
                   CZB_REAL* cf, double* flop);                              /* :658-723 */
void search_pivot_(CZB_REAL* pvt, int* sz, int* idx, int* g, CZB_REAL* X, CZB_REAL* Y,
                   CZB_REAL* Z);                                             /* :947-1039 */

/* cz_Ffunc.h <- cz_utility.f90:17-129 */
void fileout_t_(int* sz, int* g, CZB_REAL* s, CZB_REAL* dh, CZB_REAL* org, char* fname);
void exact_t_(int* sz, int* g, CZB_REAL* e, CZB_REAL* dh, CZB_REAL* org);
void err_t_(int* sz, int* idx, int* g, double* d, CZB_REAL* p, CZB_REAL* e, int* loc);

/* Stretched-grid ("MAF") twins: cz_Ffunc.h <- cz_maf.f90 and cz_blas.f90:738-935.
 * X, Y, Z are the coordinate vectors X(-1:sz(1)+2), Y(-1:sz(2)+2), Z(-1:sz(3)+2) (host or device); the
 * metric coefficients are evaluated per cell in the reference's statement order.  The 1-D work vectors
 * (a, c, d, aw, cw, dw) and tmp are the reference's scratch space and are not touched.  The five line
 * routines add 2 * sum dp^2 to *res, as the reference does (cz_maf.f90:635-636, 657-660). */
void jacobi_maf_(CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* X, CZB_REAL* Y, CZB_REAL* Z, CZB_REAL* omg,
                 CZB_REAL* b, double* res, CZB_REAL* wk2, CZB_REAL* tmp, double* flop);        /* :131-282 */
void psor_maf_(CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* X, CZB_REAL* Y, CZB_REAL* Z, CZB_REAL* omg,
               CZB_REAL* b, double* res, double* flop);                                        /* :23-114 */
void psor2sma_core_maf_(CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* X, CZB_REAL* Y, CZB_REAL* Z, int* ip,
                        int* color, CZB_REAL* omg, CZB_REAL* b, double* res, CZB_REAL* tmp,
                        double* flop);                                                         /* :301-438 */
void pcr_rb_maf_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, CZB_REAL* x, CZB_REAL* msk,
                 CZB_REAL* rhs, CZB_REAL* XX, CZB_REAL* YY, CZB_REAL* ZZ, CZB_REAL* a, CZB_REAL* c, CZB_REAL* d,
                 CZB_REAL* aw, CZB_REAL* cw, CZB_REAL* dw, CZB_REAL* omg, double* res, CZB_REAL* tmp,
                 double* flop);                                                                /* :442-668 */
void pcr_rb_esa_maf_(int* sz, int* idx, int* g, int* pn, int* ofst, int* color, int* s, CZB_REAL* x,
                     CZB_REAL* msk, CZB_REAL* rhs, CZB_REAL* XX, CZB_REAL* YY, CZB_REAL* ZZ, CZB_REAL* a,
                     CZB_REAL* c, CZB_REAL* d, CZB_REAL* aw, CZB_REAL* cw, CZB_REAL* dw, CZB_REAL* omg,
                     double* res, CZB_REAL* tmp, double* flop);                                /* :1339-1560 */
void pcr_maf_(int* sz, int* idx, int* g, int* pn, CZB_REAL* x, CZB_REAL* msk, CZB_REAL* rhs, CZB_REAL* XX,
              CZB_REAL* YY, CZB_REAL* ZZ, CZB_REAL* a, CZB_REAL* c, CZB_REAL* d, CZB_REAL* aw, CZB_REAL* cw,
              CZB_REAL* dw, CZB_REAL* omg, double* res, CZB_REAL* tmp, double* flop);          /* :672-892 */
void pcr_eda_maf_(int* sz, int* idx, int* g, int* pn, CZB_REAL* x, CZB_REAL* msk, CZB_REAL* rhs, CZB_REAL* XX,
                  CZB_REAL* YY, CZB_REAL* ZZ, CZB_REAL* aw, CZB_REAL* cw, CZB_REAL* dw, CZB_REAL* omg,
                  double* res, CZB_REAL* tmp, double* flop);                                   /* :896-1113 */
void pcr_esa_maf_(int* sz, int* idx, int* g, int* pn, int* s, CZB_REAL* x, CZB_REAL* msk, CZB_REAL* rhs,
                  CZB_REAL* XX, CZB_REAL* YY, CZB_REAL* ZZ, CZB_REAL* a, CZB_REAL* c, CZB_REAL* d, CZB_REAL* aw,
                  CZB_REAL* cw, CZB_REAL* dw, CZB_REAL* omg, double* res, CZB_REAL* tmp,
                  double* flop);                                                               /* :1117-1335 */
void calc_rk_maf_(CZB_REAL* r, CZB_REAL* p, CZB_REAL* b, int* sz, int* idx, int* g, CZB_REAL* X, CZB_REAL* Y,
                  CZB_REAL* Z, CZB_REAL* pvt, double* flop);                     /* cz_blas.f90:738-829 */
void calc_ax_maf_(CZB_REAL* ap, CZB_REAL* p, int* sz, int* idx, int* g, CZB_REAL* X, CZB_REAL* Y, CZB_REAL* Z,
                  CZB_REAL* pvt, double* flop);                                  /* cz_blas.f90:845-935 */

/* ===================================================================== PART 2 */

/* Runtime ------------------------------------------------------------------- */
int czb_real_bytes(void);                 /* 4 or 8 */
const char* czb_version(void);
int czb_init(int device);                 /* select device + create stream; 0 on success.  Aborts
                                             with a message when no CUDA device is present. */
void czb_set_stream(void* cuda_stream);   /* run on a caller-owned cudaStream_t (e.g. torch's) */
void czb_sync(void);
void* czb_malloc(size_t bytes);           /* device memory, zero-filled */
void czb_free(void* p);
void czb_memcpy_h2d(void* dst, const void* src, size_t bytes);   /* synchronous wrt the stream */
void czb_memcpy_d2h(void* dst, const void* src, size_t bytes);
long long czb_launch_count(void);         /* kernels launched by this library so far */
void czb_staged_bytes(double* h2d, double* d2h);   /* bytes staged for host-pointer arguments */
/* Per-phase report, the counterpart of the reference's PMlib profiling.txt (cz_Evaluate.cpp:506-544): device time,
 * calls and the reference's analytic flop counts per section, keyed by the reference's labels (cz_miscel.cpp:177-262:
 * JACOBI_kernel, SOR2SMA_kernel, PCR_RB, Blas_AX, BoundaryCondition, Comm_Poisson ...).  czb_profile(1) resets and
 * enables it (CZB_PROFILE=1 does so at start-up), czb_profile_report writes the table to `path` (NULL = stdout) and
 * returns the number of sections. */
void czb_profile(int on);
int czb_profile_report(const char* path, const char* title);

/* Slab decomposition along j (the slowest axis), SURVEY.md §8(e).  Pure host logic. ------ */
/* For global size gsz[3] and `nranks` slabs: local size, 1-based head index, neighbour table
 * nID[6] (cz_fparam.fi order; -1 = physical boundary) and inner index range of `rank`.
 * Returns 0, or -1 when nranks exceeds the number of inner planes. */
int czb_slab_decompose(const int gsz[3], int nranks, int rank, int sz[3], int head[3],
                       int nID[6], int idx[6]);
/* General 3-D brick decomposition inside the node, the role of CBrick in the reference (cz_Evaluate.cpp:103-159,
 * `gdv_x gdv_y gdv_z` of main.cpp:19-29): gdv[3] = divisions along x (i), y (j), z (k); every axis is cut like the j
 * axis above; rank = ri + gdv[0]*(rj + gdv[1]*rk), so gdv = {1, N, 1} is czb_slab_decompose.  Halo exchange: j faces
 * as contiguous planes, i and k faces packed by a kernel, all in one NCCL group.  The fused two-colour sweep and the
 * overlapped exchange are j-slab features; bricks with i / k neighbours run the per-colour kernels with an exchange
 * after each colour.  Line solvers (pcr*) need gdv[2] = 1.  Returns 0, or -1 when an axis has fewer inner nodes than
 * divisions. */
int czb_brick_decompose(const int gsz[3], const int gdv[3], int rank, int sz[3], int head[3],
                        int nID[6], int idx[6]);

/* Solver context = the state CZ::Evaluate sets up (cz_Evaluate.cpp:56-390) -------------- */
typedef struct czb_ctx czb_ctx;
/* idx may be NULL: computed like CZ::range_inner_index (cz_miscel.cpp:20-52) from nID. */
czb_ctx* czb_ctx_create(const int sz[3], const int idx[6], const int nID[6], CZB_REAL dh,
                        const CZB_REAL org[3]);
void czb_ctx_destroy(czb_ctx* c);
void czb_ctx_set_eps(czb_ctx* c, double eps);                /* default 1e-5, cz.h:162 */
void czb_ctx_set_res_normal(czb_ctx* c, double res_normal);  /* default 1/inner cells */
void czb_ctx_set_head(czb_ctx* c, const int head[3]);        /* global colouring offset, global Dirichlet / mesh index */
void czb_ctx_set_variant(czb_ctx* c, int variant);
/* Coordinate vectors of the stretched-grid (`*_maf`) solvers: host arrays of sz[a]+4 entries (= Fortran
 * X(-1:sz+2)); a NULL vector, or never calling this, gives the reference driver's uniform mesh
 * x[i] = (REAL)(i-1)*dh (cz_Evaluate.cpp:342-362; global index on j-slabs). */
void czb_ctx_set_mesh(czb_ctx* c, const CZB_REAL* x, const CZB_REAL* y, const CZB_REAL* z);
/* bc_k with the context's geometry.  On a j-slab (head[1] > 1) y is evaluated from the GLOBAL plane
 * index, y = dh*(j_global-1), so that the Dirichlet data do not depend on the decomposition. */
void czb_ctx_bc_k(czb_ctx* c, CZB_REAL* p);           /* 0 = auto (fused), 1 = baseline kernels */
/* Multi-rank: NCCL communicator over the j-slab ring.  id = 128-byte ncclUniqueId made by
 * czb_comm_unique_id on rank 0 and distributed by the caller (torch.distributed / MPI). */
int czb_comm_unique_id(void* id128);
int czb_ctx_comm_init(czb_ctx* c, int nranks, int rank, const void* id128);

/* Single fused sweeps on DEVICE fields: the kernels the solve loops below are built from (fused.cu /
 * fused_pk.cuh), exposed so that they can be checked one call at a time against the per-colour reference
 * kernels.  All use the context's geometry and the CubeZ coefficients cf = {1,1,1,1,1,1,6}.  *res / *dot
 * are overwritten.  Return 1 when the fused TMA kernel ran, 0 when the geometry is not eligible (row
 * pitch not a multiple of 16 bytes) and nothing was done.
 *   czb_sweep_rbsor : xnew = both colours of psor2sma_core_ (colour 0 then 1, cz_Poisson.cpp:197-203)
 *                     applied to x; *res = sum dp^2 of both colours.  x itself is not modified.
 *   czb_sweep_jacobi: xnew = jacobi_ applied to x (without the copy-back), *res = sum dp^2.
 *   czb_apply_ax    : ap = blas_calc_ax_(p) on inner cells, 0 elsewhere in the stored range;
 *                     *dot_w = (ap, w) if w != NULL, *dot_self = (ap, ap) if dot_self != NULL. */
int czb_sweep_rbsor(czb_ctx* c, const CZB_REAL* x, CZB_REAL* xnew, const CZB_REAL* b, CZB_REAL omg, double* res);
int czb_sweep_jacobi(czb_ctx* c, const CZB_REAL* x, CZB_REAL* xnew, const CZB_REAL* b, CZB_REAL omg, double* res);
int czb_apply_ax(czb_ctx* c, CZB_REAL* ap, const CZB_REAL* p, const CZB_REAL* w, double* dot_w, double* dot_self);

/* Allocation-free inner product.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same sum as blas_dot2_ above -- REAL products of x and y over the inner cells idx = {ist,ied,jst,jed,kst,ked} of a
 * field with sz[3] nodes and g guard cells, accumulated in double -- but the per-block partial sums and the ticket live
 * in a file-scope static __device__ array (64.25 KB, held by every GPU that loads the library until exit) instead of
 * the runtime's cudaMalloc'ed scratch: the call performs no cudaMalloc / cudaFree / allocating library routine and
 * does not touch the library's runtime object.  x, y (inputs) and out_dev (one double, output) are DEVICE pointers
 * supplied by the caller; everything is enqueued on `stream` (a cudaStream_t, NULL = default stream), nothing is
 * synchronised.  Limit: the inner (k,i) plane may need at most CZB_NOALLOC_MAX_BLOCKS blocks of 256 cells (about 2.1
 * million cells per plane, e.g. 1446 x 1446; the j axis is chunked to fit); larger grids are rejected with a message.
 * All calls share the static array and it keeps the last call's contents (the ticket is cleared on `stream` before
 * each launch): calls that overlap on the GPU -- different streams without an ordering between them -- race.
 * Returns 0, -1 grid beyond the limit, -2 bad argument (null / host pointer), -3 CUDA error or no device. */
#define CZB_NOALLOC_MAX_BLOCKS 8192
int czb_dot2_noalloc(const CZB_REAL* x, const CZB_REAL* y, const int sz[3], const int idx[6], int g,
                     double* out_dev, void* stream);

/* Fused solve loop.  solver/precond names follow CZ::setLS / setStrPre
 * (cz_Evaluate.cpp:571-803): jacobi psor sor2sma pbicgstab pcr pcr_eda pcr_esa pcr_rb
 * pcr_rb_esa pcr_j_esa and the stretched-grid twins jacobi_maf psor_maf sor2sma_maf pbicgstab_maf
 * pcr_maf pcr_eda_maf pcr_esa_maf pcr_rb_maf pcr_rb_esa_maf.  X (in/out) and B are device or host pointers, (sz+2g)^3 REALs.
 * hist (host, may be NULL) receives the normalised residual of each iteration
 * (the reference's history file, cz_Poisson.cpp:71); *res the last residual.
 * Returns the reference's iteration count: the converging iteration, itr_max+1 (stationary)
 * or itr_max (pbicgstab) when not converged, 0 on breakdown / unknown solver. */
int czb_solve(czb_ctx* c, const char* solver, const char* precond, CZB_REAL* X, CZB_REAL* B,
              int itr_max, CZB_REAL omg, double* res, double* hist, int hist_len, double* flop);
/* Wall-clock milliseconds (CUDA events on the solve stream) of the last czb_solve loop. */
double czb_last_solve_ms(czb_ctx* c);

#ifdef __cplusplus
}
#endif
#endif /* CZB200_H_ */

"""ctypes signatures of the reference's Fortran-kernel C ABI (reference
src/cz_cpp/cz_Ffunc.h:16-578) plus the czb_* extensions of include/czb200.h.

The same table binds libczb200_{f32,f64}.so (the product) and -- in tests only -- the CPU
oracle, because both export the reference's symbols.
"""
import ctypes as C

P = C.c_void_p          # every array / scalar argument is passed by pointer
PI = C.POINTER(C.c_int)
PD = C.POINTER(C.c_double)

# name -> number of pointer arguments (all void return), in cz_Ffunc.h order
REFERENCE_SYMBOLS = {
    "bc_k_": 6, "jacobi_": 10, "psor_": 9, "psor2sma_core_": 11, "pcr_rb_": 18, "pcr_rb_esa_": 19,
    "pcr_": 16, "pcr_eda_": 13, "pcr_esa_": 17, "pcr_j_esa_": 19,
    "imask_k_": 4, "blas_clear_": 3, "blas_copy_": 4, "blas_copy_in_": 4, "blas_triad_": 8,
    "blas_dot1_": 6, "blas_dot2_": 7, "blas_bicg_1_": 9, "blas_bicg_2_": 9, "blas_calc_ax_": 7,
    "blas_calc_rk_": 8, "search_pivot_": 7, "fileout_t_": 6, "exact_t_": 5, "err_t_": 7,
    # stretched-grid (MAF) twins, SURVEY.md §8f-2
    "jacobi_maf_": 13, "psor_maf_": 11, "psor2sma_core_maf_": 14, "pcr_rb_maf_": 22,
    "pcr_rb_esa_maf_": 23, "pcr_maf_": 20, "pcr_eda_maf_": 17, "pcr_esa_maf_": 21,
    "calc_rk_maf_": 11, "calc_ax_maf_": 10,
}

CZB_SYMBOLS = [
    "czb_real_bytes", "czb_version", "czb_init", "czb_set_stream", "czb_sync", "czb_malloc", "czb_free",
    "czb_memcpy_h2d", "czb_memcpy_d2h", "czb_launch_count", "czb_staged_bytes", "czb_slab_decompose", "czb_brick_decompose",
    "czb_ctx_create", "czb_ctx_destroy", "czb_ctx_set_eps", "czb_ctx_set_res_normal", "czb_ctx_set_head",
    "czb_ctx_set_variant", "czb_ctx_set_mesh", "czb_ctx_bc_k", "czb_comm_unique_id", "czb_ctx_comm_init", "czb_solve", "czb_last_solve_ms",
    "czb_sweep_rbsor", "czb_sweep_jacobi", "czb_apply_ax", "czb_profile", "czb_profile_report", "czb_dot2_noalloc",
]


def bind_reference_symbols(dll):
    """Attach argtypes/restype for the reference ABI to a loaded library."""
    for name, nargs in REFERENCE_SYMBOLS.items():
        fn = getattr(dll, name)
        fn.argtypes = [P] * nargs
        fn.restype = None
    return dll


def bind_czb_symbols(dll, real_ctype):
    dll.czb_real_bytes.restype = C.c_int
    dll.czb_version.restype = C.c_char_p
    dll.czb_init.argtypes = [C.c_int]; dll.czb_init.restype = C.c_int
    dll.czb_set_stream.argtypes = [P]; dll.czb_set_stream.restype = None
    dll.czb_sync.restype = None
    dll.czb_malloc.argtypes = [C.c_size_t]; dll.czb_malloc.restype = P
    dll.czb_free.argtypes = [P]; dll.czb_free.restype = None
    dll.czb_memcpy_h2d.argtypes = [P, P, C.c_size_t]; dll.czb_memcpy_h2d.restype = None
    dll.czb_memcpy_d2h.argtypes = [P, P, C.c_size_t]; dll.czb_memcpy_d2h.restype = None
    dll.czb_launch_count.restype = C.c_longlong
    dll.czb_staged_bytes.argtypes = [PD, PD]; dll.czb_staged_bytes.restype = None
    dll.czb_slab_decompose.argtypes = [PI, C.c_int, C.c_int, PI, PI, PI, PI]; dll.czb_slab_decompose.restype = C.c_int
    dll.czb_brick_decompose.argtypes = [PI, PI, C.c_int, PI, PI, PI, PI]; dll.czb_brick_decompose.restype = C.c_int
    dll.czb_ctx_create.argtypes = [PI, PI, PI, real_ctype, C.POINTER(real_ctype)]; dll.czb_ctx_create.restype = P
    dll.czb_ctx_destroy.argtypes = [P]; dll.czb_ctx_destroy.restype = None
    dll.czb_ctx_set_eps.argtypes = [P, C.c_double]; dll.czb_ctx_set_eps.restype = None
    dll.czb_ctx_set_res_normal.argtypes = [P, C.c_double]; dll.czb_ctx_set_res_normal.restype = None
    dll.czb_ctx_set_head.argtypes = [P, PI]; dll.czb_ctx_set_head.restype = None
    dll.czb_ctx_set_variant.argtypes = [P, C.c_int]; dll.czb_ctx_set_variant.restype = None
    dll.czb_ctx_set_mesh.argtypes = [P, P, P, P]; dll.czb_ctx_set_mesh.restype = None
    dll.czb_ctx_bc_k.argtypes = [P, P]; dll.czb_ctx_bc_k.restype = None
    dll.czb_comm_unique_id.argtypes = [P]; dll.czb_comm_unique_id.restype = C.c_int
    dll.czb_ctx_comm_init.argtypes = [P, C.c_int, C.c_int, P]; dll.czb_ctx_comm_init.restype = C.c_int
    dll.czb_solve.argtypes = [P, C.c_char_p, C.c_char_p, P, P, C.c_int, real_ctype, PD, PD, C.c_int, PD]
    dll.czb_solve.restype = C.c_int
    dll.czb_last_solve_ms.argtypes = [P]; dll.czb_last_solve_ms.restype = C.c_double
    dll.czb_sweep_rbsor.argtypes = [P, P, P, P, real_ctype, PD]; dll.czb_sweep_rbsor.restype = C.c_int
    dll.czb_sweep_jacobi.argtypes = [P, P, P, P, real_ctype, PD]; dll.czb_sweep_jacobi.restype = C.c_int
    dll.czb_apply_ax.argtypes = [P, P, P, P, PD, PD]; dll.czb_apply_ax.restype = C.c_int
    dll.czb_profile.argtypes = [C.c_int]; dll.czb_profile.restype = None
    dll.czb_dot2_noalloc.argtypes = [P, P, PI, PI, C.c_int, P, P]; dll.czb_dot2_noalloc.restype = C.c_int
    dll.czb_profile_report.argtypes = [C.c_char_p, C.c_char_p]; dll.czb_profile_report.restype = C.c_int
    return dll

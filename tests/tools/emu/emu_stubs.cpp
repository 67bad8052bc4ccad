// tests/tools/emu/emu_stubs.cpp -- TEST INFRASTRUCTURE: host entry points of fused.cu / pcr.cu for the CPU-emulated
// library.  Those files are TMA / tensor-memory / inline-PTX kernels that only exist on the device, so the emulated
// library reports "not eligible" (the solve loops then take the per-colour baseline kernels, as they do on the GPU
// for geometries the fused kernels do not support) and refuses the table-driven line solver.
#include "emu_prelude.h"
#include "kernels.cuh"

namespace czb {
bool fused_eligible(const Geo&) { return false; }
bool launch_rb2_fused(const Launch&, const Geo&, const REAL*, REAL*, const REAL*, const REAL*, REAL, int, int, int, int, bool,
                      bool, int, unsigned int*, int*) { return false; }
bool rb2_zero_src_supported(const Geo&, const REAL*) { return false; }
bool rb2_parts_supported(const Geo&, const REAL*) { return false; }
bool launch_jacobi_fused(const Launch&, const Geo&, const REAL*, REAL*, const REAL*, const REAL*, REAL, int, bool) { return false; }
int launch_ax_fused(const Launch&, const Geo&, REAL*, const REAL*, const REAL*, const REAL*, int, int) { return 0; }
void pcr_build_tables(cudaStream_t, PcrTables&, int, int, bool) {}
void pcr_free_tables(PcrTables&) {}
void launch_pcr_lines(const Launch&, const Geo&, const PcrTables&, int, int, int, const REAL*, REAL*, const REAL*, const REAL*,
                      REAL, int, bool, bool) {
  fprintf(stderr, "emu: the table-driven line-SOR kernels (pcr.cu) are not part of the CPU emulation\n");
  abort();
}
}  // namespace czb

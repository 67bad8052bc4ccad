// tests/tools/emu/emu_stubs.cpp -- TEST INFRASTRUCTURE: host entry points of fused.cu / persist.cu for the CPU-emulated library.
// Its kernels are TMA / inline-PTX code that only exists on the device, so the emulated library reports "not eligible"
// and the solve loops take the per-colour baseline kernels, as they do on the GPU for geometries the fused kernels do
// not support.
#include "emu_prelude.h"
#include "kernels.cuh"

namespace czb {
bool fused_eligible(const Geo&) { return false; }
bool launch_rb2_fused(const Launch&, const Geo&, const REAL*, REAL*, const REAL*, const REAL*, REAL, int, int, int, int, bool,
                      bool, int) { return false; }
bool rb2_zero_src_supported(const Geo&, const REAL*) { return false; }
bool rb2_fused_launchable(const Launch&, const Geo&, const REAL*, bool, int) { return false; }
bool jacobi_fused_launchable(const Launch&, const Geo&, const REAL*, int) { return false; }
bool launch_jacobi_fused(const Launch&, const Geo&, const REAL*, REAL*, const REAL*, const REAL*, REAL, int, bool, int) { return false; }
int launch_ax_fused(const Launch&, const Geo&, REAL*, const REAL*, const REAL*, const REAL*, int, int, int) { return 0; }
// persist.cu: a grid barrier needs co-resident CTAs; the emulation runs blocks one after the other
bool persist_eligible(const Geo&, int) { return false; }
void launch_persist(const Launch&, const Geo&, int, REAL*, REAL*, const REAL*, const REAL*, REAL, int, int, int, LoopState*,
                    double, double, double*, int) { abort(); }
}  // namespace czb

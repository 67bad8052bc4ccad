// tests/tools/emu/simt_line.cpp -- TEST INFRASTRUCTURE: the shared-memory line kernel of maf.cu (k_pcr_lines_maf) run
// with REAL concurrency on the CPU: every CUDA thread of a block is a std::thread, __syncthreads() is a std::barrier,
// __shared__ variables are storage shared by those threads.  Built with -fsanitize=thread, so a missing or misplaced
// barrier (a data race on the a / c / d ping-pong rows) is reported by ThreadSanitizer and fails the test, and the
// fields must still equal the oracle bit for bit.  Blocks run one after the other with a small grid (the kernel's
// line loop is grid-stride).  The grid-wide reduction of common.cuh (warp shuffles; verified on the GPU by the plain
// kernels) is replaced by a mutex-protected sum.
#include <cuda_runtime.h>
#include <algorithm>
#include <barrier>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

static thread_local uint3 threadIdx;
static uint3 blockIdx;
static dim3 blockDim, gridDim;
static std::barrier<>* g_barrier = nullptr;
using std::min;
using std::max;

#define __CUDACC__ 1
#undef __global__
#undef __device__
#undef __host__
#undef __forceinline__
#undef __launch_bounds__
#undef __shared__
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static       // one block at a time: a function-local static IS the block's shared memory
#define __syncthreads() g_barrier->arrive_and_wait()
#define __threadfence() ((void)0)
template <class T> static inline T __ldg(const T* p) { return *p; }
template <class T> static inline T __ldcg(const T* p) { return *p; }
static inline double __shfl_xor_sync(unsigned, double v, int) { return v; }   // only inside the replaced grid_reduce
static inline unsigned int atomicAdd(unsigned int* p, unsigned int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
#define cudaGetLastError() cudaSuccess
#define cudaFuncSetAttribute(...) cudaSuccess
#define cudaGetErrorName(e) "emu"
#define cudaGetErrorString(e) "emu"

#define grid_reduce grid_reduce_of_the_device     // keep common.cuh's version out of the way (never instantiated)
#include "common.cuh"
#undef grid_reduce

namespace czb {
static std::mutex g_sum_mutex;
static double g_sum[CZB_MAX_ACC];
static std::function<void()> g_finish;
template <int NACC, class Fin>
inline void grid_reduce(double (&v)[NACC], RedBuf, Fin fin) {
  std::lock_guard<std::mutex> lock(g_sum_mutex);
  for (int a = 0; a < NACC; ++a) g_sum[a] += v[a];
  g_finish = [fin]() { double t[NACC]; for (int a = 0; a < NACC; ++a) t[a] = g_sum[a]; fin(t); };
}
}  // namespace czb

static int g_grid_cap = 3;      // grid-stride kernels only
template <class Body>
static void emu_launch(dim3 g, dim3 b, Body body) {
  if (g.y != 1 || g.z != 1) { fprintf(stderr, "simt: 1-D grids only\n"); abort(); }
  gridDim = dim3(std::min<unsigned>(g.x, g_grid_cap), 1, 1);
  blockDim = b;
  const unsigned nt = b.x * b.y * b.z;
  for (int a = 0; a < CZB_MAX_ACC; ++a) czb::g_sum[a] = 0.0;
  for (unsigned bx = 0; bx < gridDim.x; ++bx) {
    blockIdx.x = bx; blockIdx.y = blockIdx.z = 0;
    std::barrier<> bar((std::ptrdiff_t)nt);
    g_barrier = &bar;
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; ++t)
      th.emplace_back([&, t] { threadIdx.x = t % b.x; threadIdx.y = (t / b.x) % b.y; threadIdx.z = t / (b.x * b.y); body(); });
    for (auto& x : th) x.join();
  }
  if (czb::g_finish) { czb::g_finish(); czb::g_finish = nullptr; }
}

#include "maf_simt_gen.inc"     // maf.cu + cells.cuh with the launches rewritten (emu_build.py)

namespace czb { void count_launch(int) {} }
using namespace czb;

extern "C" {
int simt_real_bytes() { return (int)sizeof(REAL); }

// color >= 0: the lines of one colour; color < 0: all lines through the i+j wavefronts.  Returns 2 * sum dp^2.
double simt_pcr_maf(REAL* x, REAL* msk, REAL* rhs, int* sz, int* idx, int pn, int color, REAL* X, REAL* Y, REAL* Z, REAL omg) {
  static LoopState st;
  st = LoopState();
  Launch L; L.s = nullptr; L.rb.partial = nullptr; L.rb.ticket = nullptr; L.rb_blocks = 1 << 16; L.st = nullptr; L.acc_st = &st;
  const Geo G = make_geo(sz, idx, 2);
  const MafAxes A{X, Y, Z};
  if (color >= 0) {
    launch_pcr_lines_maf(L, G, pn, color, 0, A, x, msk, rhs, omg, 0, false);
  } else {
    bool first = true;
    for (int d = G.ist + G.jst; d <= G.ied + G.jed; ++d) {
      launch_pcr_lines_maf(L, G, pn, 3, d, A, x, msk, rhs, omg, 0, !first);
      first = false;
    }
  }
  return st.acc[0];
}
}

// Stand-alone driver (a TSan-instrumented library cannot be loaded into an uninstrumented Python):
//   simt_line <in.bin> <out.bin>
// in : int sz[3], idx[6], pn, color ; REAL omg ; REAL x[], msk[], rhs[] (full fields) ; REAL X[], Y[], Z[]
// out: REAL x[] ; double res
int main(int argc, char** argv) {
  if (argc != 3) return 2;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 3;
  int hdr[11];
  REAL omg;
  if (fread(hdr, sizeof(int), 11, f) != 11 || fread(&omg, sizeof(REAL), 1, f) != 1) return 4;
  int* sz = hdr; int* idx = hdr + 3; const int pn = hdr[9], color = hdr[10];
  const size_t total = (size_t)(sz[0] + 4) * (sz[1] + 4) * (sz[2] + 4);
  std::vector<REAL> x(total), msk(total), rhs(total), X(sz[0] + 4), Y(sz[1] + 4), Z(sz[2] + 4);
  bool ok = fread(x.data(), sizeof(REAL), total, f) == total && fread(msk.data(), sizeof(REAL), total, f) == total &&
            fread(rhs.data(), sizeof(REAL), total, f) == total && fread(X.data(), sizeof(REAL), X.size(), f) == X.size() &&
            fread(Y.data(), sizeof(REAL), Y.size(), f) == Y.size() && fread(Z.data(), sizeof(REAL), Z.size(), f) == Z.size();
  fclose(f);
  if (!ok) return 5;
  const double res = simt_pcr_maf(x.data(), msk.data(), rhs.data(), sz, idx, pn, color, X.data(), Y.data(), Z.data(), omg);
  f = fopen(argv[2], "wb");
  fwrite(x.data(), sizeof(REAL), total, f);
  fwrite(&res, sizeof(double), 1, f);
  fclose(f);
  return 0;
}

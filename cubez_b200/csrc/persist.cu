// cubez_b200/csrc/persist.cu -- persistent cooperative sweeps for grids that live in L2.
//
// On a 64^3 .. ~160^3 grid one iteration of CZ::JACOBI / CZ::RBSOR (cz_Poisson.cpp:30-82, 159-235) moves a few MB:
// the plane-marching TMA kernels of fused.cu spend their time filling pipelines and every iteration is a kernel
// launch (19 - 31 us per Jacobi iteration at 64^3 although the data of a sweep would stream from HBM in 0.3 us).  Here ONE
// cooperative launch (one CTA per SM, all co-resident) runs the whole iteration loop: a sweep over the CTA's rows with
// loads that bypass L1 (ld.global.cg: the fields stay in the 126 MB L2, no stale line can be hit), a grid barrier, the
// residual summed by every CTA from the per-CTA partials in the same fixed order -- so all CTAs take the same
// `res < eps` decision without a second barrier -- and the next sweep.  Red-black SOR runs in place with one more
// barrier between the colours.  Arithmetic: the statements of cz_solver.f90:339-348 / 468-476 in the reference's
// operand order with IEEE division (the library is built with -fmad=false), i.e. the same per-cell results as
// k_cells<OpJacobi> / k_rb_color (kernels.cu) and as the oracle; the sums are REAL products accumulated in double in
// an order fixed by the launch shape.
#include "kernels.cuh"
#include <cstdlib>

namespace czb {
namespace {

struct PersistArgs {
  Geo G;
  REAL* X; REAL* W; const REAL* B;
  REAL c1, c2, c3, c4, c5, c6, dd, omg;
  int mode;                 // 0 = Jacobi (ping-pong X / W), 1 = red-black SOR (in place on X)
  int kp;                   // colour offset of the first colour (cz_Poisson.cpp:181)
  int it_first, it_last;    // iterations run by this launch (1-based, inclusive)
  LoopState* st;
  double res_normal, eps;
  double* hist; int hist_len;
  double* partial;          // [2][gridDim.x] per-CTA sums, double buffered over the iterations
  unsigned int* bar;        // grid barrier counter (zeroed before the launch)
};

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// All CTAs of the (cooperative, hence co-resident) grid meet here; `target` counts the arrivals expected so far.
__device__ __forceinline__ void grid_barrier(unsigned int* ctr, unsigned int& target) {
  __syncthreads();
  target += gridDim.x;
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
    while (ld_acquire_u32(ctr) < target) { }
    __threadfence();
  }
  __syncthreads();
}

// One relaxation, cz_solver.f90:339-347 (left to right), neighbours read around L1.
__device__ __forceinline__ void relax_cell(const PersistArgs& a, const REAL* src, REAL* dst, long long off, double& acc) {
  const Geo& G = a.G;
  const REAL pp = __ldcg(src + off);
  const REAL bb = __ldg(a.B + off);
  const REAL ss = a.c1 * __ldcg(src + off + G.nk) + a.c2 * __ldcg(src + off - G.nk) + a.c3 * __ldcg(src + off + G.plane) +
                  a.c4 * __ldcg(src + off - G.plane) + a.c5 * __ldcg(src + off + 1) + a.c6 * __ldcg(src + off - 1);
  const REAL dp = ((ss - bb) / a.dd - pp) * a.omg;
  dst[off] = pp + dp;
  acc += (double)(dp * dp);
}

// color < 0: every cell of the row; else the cells k = kst + mod(i+j+kp+color, 2) + 2t (cz_solver.f90:466)
__device__ __forceinline__ void sweep_rows(const PersistArgs& a, const REAL* src, REAL* dst, int color, double& acc) {
  const Geo& G = a.G;
  const int nki = G.ked - G.kst + 1, nii = G.ied - G.ist + 1, nji = G.jed - G.jst + 1;
  const int nrows = nii * nji;
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int gw = blockIdx.x * wpb + (threadIdx.x >> 5), tw = gridDim.x * wpb;
  for (int r = gw; r < nrows; r += tw) {
    const int i = G.ist + r % nii, j = G.jst + r / nii;
    const long long row = G.at(G.kst, i, j);
    if (color < 0) {
      for (int k = lane; k < nki; k += 32) relax_cell(a, src, dst, row + k, acc);
    } else {
      for (int k = ((i + j + a.kp + color) & 1) + 2 * lane; k < nki; k += 64) relax_cell(a, src, dst, row + k, acc);
    }
  }
}

__global__ void __launch_bounds__(1024, 1) k_persist(const PersistArgs a) {
  if (a.st->done) return;                       // the same value for every CTA: nothing runs concurrently with this launch
  __shared__ double s_w[32];
  __shared__ int s_stop;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const unsigned int nb = gridDim.x;
  unsigned int target = 0u;
  for (int it = a.it_first; it <= a.it_last; ++it) {
    double acc = 0.0;
    if (a.mode == 0) {
      const REAL* src = (it & 1) ? a.X : a.W;   // ping-pong roles of CZ::JACOBI in solver.cu
      REAL* dst = (it & 1) ? a.W : a.X;
      sweep_rows(a, src, dst, -1, acc);
    } else {
      sweep_rows(a, a.X, a.X, 0, acc);
      grid_barrier(a.bar, target);              // the second colour reads the first colour's new values
      sweep_rows(a, a.X, a.X, 1, acc);
    }
    // CTA sum -> partial[it & 1][blockIdx.x]
    acc = warp_sum(acc);
    if (lane == 0) s_w[wid] = acc;
    __syncthreads();
    if (wid == 0) {
      double w = lane < nw ? s_w[lane] : 0.0;
      w = warp_sum(w);
      if (lane == 0) a.partial[(size_t)(it & 1) * nb + blockIdx.x] = w;
    }
    grid_barrier(a.bar, target);
    // every CTA adds the partials in the same order -> the same res, the same decision (cz_Poisson.cpp:69-77)
    if (wid == 0) {
      double w = 0.0;
      for (unsigned int b = lane; b < nb; b += 32) w += __ldcg(&a.partial[(size_t)(it & 1) * nb + b]);
      w = warp_sum(w);
      if (lane == 0) {
        double res = w;
        res *= a.res_normal;
        res = sqrt(res);
        const int stop = res < a.eps ? 1 : 0;
        s_stop = stop;
        if (blockIdx.x == 0) {
          a.st->acc[0] = w;
          a.st->res = res;
          if (a.hist && it - 1 < a.hist_len) a.hist[it - 1] = res;
          a.st->itr = it;
          if (stop) a.st->done = 1;
        }
      }
    }
    __syncthreads();
    if (s_stop) break;
  }
}

struct PersistDev {
  bool probed = false, ok = false;
  int sms = 0;
  unsigned int* bar = nullptr;
};
PersistDev& pdev() { static PersistDev d; return d; }

bool probe() {
  PersistDev& d = pdev();
  if (d.probed) return d.ok;
  d.probed = true;
  int dev = 0, coop = 0, nblk = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nblk, k_persist, 1024, 0) != cudaSuccess) { cudaGetLastError(); return false; }
  if (!coop || nblk < 1 || d.sms < 1) return false;
  CZB_CHECK(cudaMalloc((void**)&d.bar, 256));
  d.ok = true;
  return true;
}

}  // namespace

// Where the persistent loop wins (B200, f32, per-iteration launches of the TMA kernels vs this loop): Jacobi 64^3 30.8 ->
// 5.2 us / iteration, 128^3 25.0 -> 14.5, 160^3 33.9 -> 22.9; red-black SOR 64^3 21.2 -> 7.3, 128^3 23.3 -> 18.6,
// 160^3 29.4 -> 30.3 (break-even), 200^3 39.8 -> 51.8.  Hence: fields (X, WRK, B for Jacobi; X, B for red-black SOR) of
// at most 72 MB resp. 30 MB in total, i.e. ~180^3 resp. ~155^3 cells in float.  CZB_PERSIST_MB overrides the limit
// (0 switches the path off); it is read per solve so that tests can flip it.
bool persist_eligible(const Geo& G, int nfields) {
  const char* e = getenv("CZB_PERSIST_MB");
  const double mb = e ? atof(e) : (nfields >= 3 ? 72.0 : 30.0);
  if (mb <= 0.0) return false;
  if ((double)G.total * sizeof(REAL) * nfields > mb * 1.0e6) return false;
  if (G.ked < G.kst || G.ied < G.ist || G.jed < G.jst) return false;
  return probe();
}

void launch_persist(const Launch& L, const Geo& G, int mode, REAL* X, REAL* W, const REAL* B, const REAL* cf, REAL omg,
                    int kp, int it_first, int it_last, LoopState* st, double res_normal, double eps, double* hist,
                    int hist_len) {
  if (it_first > it_last) return;
  if (!probe()) { fprintf(stderr, "czb200: persistent sweep launched on a device that cannot run it\n"); abort(); }
  PersistDev& d = pdev();
  const int nrows = (G.ied - G.ist + 1) * (G.jed - G.jst + 1);
  int grid = (nrows + 31) / 32;          // 32 warps (= rows in flight) per CTA
  if (grid > d.sms) grid = d.sms;
  if (grid < 1) grid = 1;
  if ((size_t)grid * 2 > L.rb_blocks * CZB_MAX_ACC) { fprintf(stderr, "czb200: reduction scratch too small\n"); abort(); }
  PersistArgs a;
  a.G = G; a.X = X; a.W = W; a.B = B;
  a.c1 = cf[0]; a.c2 = cf[1]; a.c3 = cf[2]; a.c4 = cf[3]; a.c5 = cf[4]; a.c6 = cf[5]; a.dd = cf[6]; a.omg = omg;
  a.mode = mode; a.kp = kp; a.it_first = it_first; a.it_last = it_last;
  a.st = st; a.res_normal = res_normal; a.eps = eps; a.hist = hist; a.hist_len = hist_len;
  a.partial = L.rb.partial; a.bar = d.bar;
  CZB_CHECK(cudaMemsetAsync(d.bar, 0, sizeof(unsigned int), L.s));
  void* args[] = {(void*)&a};
  CZB_CHECK(cudaLaunchCooperativeKernel((const void*)k_persist, dim3(grid), dim3(1024), args, 0, L.s));
  count_launch();
}

}  // namespace czb

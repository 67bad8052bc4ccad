// cubez_b200/csrc/common.cuh -- shared device/host definitions for the sm_100a kernels.
//
// Field layout (identical to the reference so device pointers are a zero-copy drop-in):
// Fortran p(1-g:sz(3)+g, 1-g:sz(1)+g, 1-g:sz(2)+g), index order (k,i,j), k contiguous
// (reference src/cz_f90/cz_solver.f90:218).  All kernels address cells by their Fortran
// indices through Geo::at().
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

#ifdef _REAL_IS_DOUBLE_
typedef double REAL;
#define CZB_REAL_NAME "f64"
#else
typedef float REAL;
#define CZB_REAL_NAME "f32"
#endif

namespace czb {

// Geometry of one (sub)domain array.  Fortran index ranges are inclusive.
struct Geo {
  int ix, jx, kx, g;               // sz(1), sz(2), sz(3), guide
  int nk, ni, nj;                  // padded extents: kx+2g, ix+2g, jx+2g
  long long plane;                 // nk*ni  (distance between j-planes)
  long long total;                 // nk*ni*nj
  int ist, ied, jst, jed, kst, ked;  // inner (computed) range, Fortran indices

  __host__ __device__ inline long long at(int k, int i, int j) const {
    return (long long)(k + g - 1) + (long long)nk * ((long long)(i + g - 1) + (long long)ni * (long long)(j + g - 1));
  }
  __host__ __device__ inline double cells() const {
    return (double)(ied - ist + 1) * (double)(jed - jst + 1) * (double)(ked - kst + 1);
  }
};

inline Geo make_geo(const int* sz, const int* idx, int g) {
  Geo G;
  G.ix = sz[0]; G.jx = sz[1]; G.kx = sz[2]; G.g = g;
  G.nk = sz[2] + 2 * g; G.ni = sz[0] + 2 * g; G.nj = sz[1] + 2 * g;
  G.plane = (long long)G.nk * G.ni;
  G.total = G.plane * G.nj;
  if (idx) { G.ist = idx[0]; G.ied = idx[1]; G.jst = idx[2]; G.jed = idx[3]; G.kst = idx[4]; G.ked = idx[5]; }
  else     { G.ist = 1; G.ied = sz[0]; G.jst = 1; G.jed = sz[1]; G.kst = 1; G.ked = sz[2]; }
  return G;
}

// Device-resident state of an iteration loop (one per solver context).  Kernels of an
// iteration exit immediately once `done` is set, so the host can enqueue iterations in
// batches without reading anything back.
struct LoopState {
  int done;          // 1: converged / broken down -> remaining launches are no-ops
  int itr;           // iterations completed so far
  int status;        // 0 ok, 1 BiCGstab breakdown |rho| < FLT_MIN (cz_Poisson.cpp:379)
  int pad;
  double res;        // last normalised residual sqrt(sum * res_normal)
  double acc[8];     // raw reduction results (sums over inner cells)
  double scal[8];    // BiCGstab scalars, stored as double but rounded to REAL (Fdot1/Fdot2)
};

enum { SC_RHO = 0, SC_RHO_OLD, SC_ALPHA, SC_BETA, SC_OMEGA, SC_NEG_ALPHA, SC_NEG_OMEGA };

// Scratch for the deterministic two-stage reduction: per-block partials + ticket.
struct RedBuf {
  double* partial;     // [max_blocks * CZB_MAX_ACC]
  unsigned int* ticket;
};
#define CZB_MAX_ACC 4

#define CZB_CHECK(call)                                                                      \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      fprintf(stderr, "czb200: CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, \
              __LINE__, cudaGetErrorString(e_));                                             \
      abort();                                                                               \
    }                                                                                        \
  } while (0)

#ifdef __CUDACC__
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-level sum of NACC doubles followed by a grid-level deterministic finish: every block
// stores its partial, the last block to arrive (ticket) sums all partials in block order and
// hands the totals to `fin`.  Order of additions is fixed by (grid, block) shape only, so the
// result is reproducible run to run.  All threads of the block must call this.
template <int NACC, class Fin>
__device__ __forceinline__ void grid_reduce(double (&v)[NACC], RedBuf rb, Fin fin) {
  __shared__ double s_part[NACC][32];
  __shared__ bool s_last;
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  const int nthr = blockDim.x * blockDim.y * blockDim.z;
  const int lane = tid & 31, wid = tid >> 5, nw = (nthr + 31) >> 5;
#pragma unroll
  for (int a = 0; a < NACC; ++a) {
    double w = warp_sum(v[a]);
    if (lane == 0) s_part[a][wid] = w;
  }
  __syncthreads();
  const unsigned int bid = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  const unsigned int nblk = gridDim.x * gridDim.y * gridDim.z;
  if (wid == 0) {
#pragma unroll
    for (int a = 0; a < NACC; ++a) {
      double w = (lane < nw) ? s_part[a][lane] : 0.0;
      w = warp_sum(w);
      if (lane == 0) rb.partial[(size_t)bid * NACC + a] = w;
    }
    if (lane == 0) {
      __threadfence();
      unsigned int t = atomicAdd(rb.ticket, 1u);
      s_last = (t == nblk - 1);
    }
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    double tot[NACC];
#pragma unroll
    for (int a = 0; a < NACC; ++a) {
      double w = 0.0;
      for (unsigned int b = tid; b < nblk; b += nthr) w += __ldcg(&rb.partial[(size_t)b * NACC + a]);
      tot[a] = w;
    }
    __syncthreads();
#pragma unroll
    for (int a = 0; a < NACC; ++a) {
      double w = warp_sum(tot[a]);
      if (lane == 0) s_part[a][wid] = w;
    }
    __syncthreads();
    if (wid == 0) {
#pragma unroll
      for (int a = 0; a < NACC; ++a) {
        double w = (lane < nw) ? s_part[a][lane] : 0.0;
        tot[a] = warp_sum(w);
      }
      if (lane == 0) {
        *rb.ticket = 0u;  // re-arm for the next launch (stream-ordered)
        fin(tot);
      }
    }
  }
}
#endif  // __CUDACC__

}  // namespace czb

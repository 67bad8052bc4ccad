// cubez_b200/csrc/noalloc.cu -- czb_dot2_noalloc: the inner product of blas_dot2_ (cz_blas.f90: blas_dot2) for callers
// that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// NO COUNTERPART IN THE REFERENCE.
//
// blas_dot2_ (abi.cu -> launch_dot2 -> k_cells<OpDot2> + grid_reduce) keeps its per-block partial sums and its ticket
// in scratch memory the runtime cudaMallocs at start-up (runtime.cu).  This variant keeps the same two things in a
// file-scope `static __device__` array of fixed size, so the entry point calls no cudaMalloc / cudaFree / allocating
// library routine at all and does not touch the library's runtime object (whose first use allocates):
//   * scratch = 256-byte header (word 0 = ticket) + CZB_NOALLOC_MAX_BLOCKS doubles = 64.25 KB, held by every GPU that
//     loads the library until the process exits;
//   * in HOST code the array's name is only a stand-in, not a device address: the address is looked up once per device,
//     with that device current, by cudaGetSymbolAddress (checked) and passed to the kernel as an argument;
//   * every call shares the array and it keeps the last call's contents: the ticket must start at zero and is cleared
//     on the caller's stream before each launch (the kernel also re-arms it); the partials are all rewritten before
//     they are read.  Calls that OVERLAP on the GPU (different streams, no ordering between them) race on it;
//   * a (k, i) plane needs one block per 256 inner cells and the j axis is chunked until the grid fits: grids whose
//     inner (k, i) plane alone needs more than CZB_NOALLOC_MAX_BLOCKS blocks are rejected with a message (-1).
// Arithmetic as blas_dot2: REAL products accumulated in double, order fixed by the launch shape.
#include "common.cuh"
#include "../../include/czb200.h"
#include <mutex>

namespace {

constexpr int kHeader = 256;                                   // bytes before the partials (ticket in word 0)
constexpr int kMaxBlocks = CZB_NOALLOC_MAX_BLOCKS;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kHeader + sizeof(double) * kMaxBlocks];

// one thread per (k, i) column of the tile, marching jlen planes; block (bx, by), k on threadIdx.x (coalesced)
__global__ void __launch_bounds__(256) k_dot2_noalloc(czb::Geo G, int jlen, const REAL* __restrict__ x,
                                                      const REAL* __restrict__ y, unsigned char* scratch,
                                                      double* __restrict__ out) {
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  double* partial = reinterpret_cast<double*>(scratch + kHeader);
  __shared__ double s_part[32];
  __shared__ bool s_last;
  const int k = G.kst + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = G.ist + blockIdx.y * blockDim.y + threadIdx.y;
  const int j0 = G.jst + blockIdx.z * jlen;
  const int j1 = min(j0 + jlen - 1, G.jed);
  double acc = 0.0;
  if (k <= G.ked && i <= G.ied) {
    long long off = G.at(k, i, j0);
    for (int j = j0; j <= j1; ++j, off += G.plane) acc += (double)(x[off] * y[off]);   // REAL product, double sum
  }
  const int tid = threadIdx.x + blockDim.x * threadIdx.y, nthr = blockDim.x * blockDim.y;
  const int lane = tid & 31, wid = tid >> 5, nw = (nthr + 31) >> 5;
  acc = czb::warp_sum(acc);
  if (lane == 0) s_part[wid] = acc;
  __syncthreads();
  const unsigned int bid = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  const unsigned int nblk = gridDim.x * gridDim.y * gridDim.z;
  if (wid == 0) {
    double w = (lane < nw) ? s_part[lane] : 0.0;
    w = czb::warp_sum(w);
    if (lane == 0) {
      partial[bid] = w;
      __threadfence();
      s_last = (atomicAdd(ticket, 1u) == nblk - 1);
    }
  }
  __syncthreads();
  if (s_last) {            // the last block to arrive adds the partials in a fixed order
    __threadfence();
    double w = 0.0;
    for (unsigned int b = tid; b < nblk; b += nthr) w += __ldcg(&partial[b]);
    w = czb::warp_sum(w);
    __syncthreads();
    if (lane == 0) s_part[wid] = w;
    __syncthreads();
    if (wid == 0) {
      double t = (lane < nw) ? s_part[lane] : 0.0;
      t = czb::warp_sum(t);
      if (lane == 0) { *out = t; *ticket = 0u; }
    }
  }
}

// device address of the scratch array, looked up once per device with that device current
unsigned char* scratch_address(int dev) {
  static std::mutex mu;
  static unsigned char* addr[kMaxDevices] = {nullptr};
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      fprintf(stderr, "czb200: czb_dot2_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
      cudaGetLastError();
      return nullptr;
    }
    addr[dev] = static_cast<unsigned char*>(p);
  }
  return addr[dev];
}

bool on_device(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

}  // namespace

extern "C" int czb_dot2_noalloc(const CZB_REAL* x, const CZB_REAL* y, const int sz[3], const int idx[6], int g,
                                double* out_dev, void* stream) {
  if (!x || !y || !sz || !out_dev || g < 0) { fprintf(stderr, "czb200: czb_dot2_noalloc: null argument\n"); return -2; }
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    fprintf(stderr, "czb200: czb_dot2_noalloc: no current CUDA device (there is no CPU path)\n");
    cudaGetLastError();
    return -3;
  }
  if (!on_device(x) || !on_device(y) || !on_device(out_dev)) {
    fprintf(stderr, "czb200: czb_dot2_noalloc: x, y and out_dev must be DEVICE pointers (staging host arrays would allocate)\n");
    return -2;
  }
  const czb::Geo G = czb::make_geo(sz, idx, g);
  const int nk = G.ked - G.kst + 1, ni = G.ied - G.ist + 1, nj = G.jed - G.jst + 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (nk <= 0 || ni <= 0 || nj <= 0) {   // empty inner range: the sum is 0
    return cudaMemsetAsync(out_dev, 0, sizeof(double), s) == cudaSuccess ? 0 : -3;
  }
  int bx = 32;
  while (bx < nk && bx < 256) bx += 32;
  int by = 256 / bx; if (by < 1) by = 1;
  if (by > ni) by = ni;
  const long long gx = (nk + bx - 1) / bx, gy = (ni + by - 1) / by;
  if (gx * gy > kMaxBlocks) {
    fprintf(stderr, "czb200: czb_dot2_noalloc: an inner (k,i) plane of %d x %d cells needs %lld blocks, the static scratch holds "
                    "%d (about %d cells per plane); use blas_dot2_ for this grid\n", nk, ni, gx * gy, kMaxBlocks, kMaxBlocks * 256);
    return -1;
  }
  // j chunks: as many as fit the scratch, at most ~64 blocks per SM of a B200 (few, fat blocks keep the last block's sum short)
  long long cap = 148LL * 64;
  if (cap > kMaxBlocks) cap = kMaxBlocks;
  int jlen = 1;
  while (gx * gy * ((nj + jlen - 1) / jlen) > cap) ++jlen;
  const int gz = (nj + jlen - 1) / jlen;
  unsigned char* scratch = scratch_address(dev);
  if (!scratch) return -3;
  // the ticket must start at zero whatever an earlier (possibly aborted) call left behind
  if (cudaMemsetAsync(scratch, 0, sizeof(unsigned int), s) != cudaSuccess) { cudaGetLastError(); return -3; }
  const dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)gz), block(bx, by, 1);
  k_dot2_noalloc<<<grid, block, 0, s>>>(G, jlen, x, y, scratch, out_dev);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { fprintf(stderr, "czb200: czb_dot2_noalloc: launch failed: %s\n", cudaGetErrorString(e)); return -3; }
  return 0;
}

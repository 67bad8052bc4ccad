// tests/tools/emu/emu_prelude.h -- TEST INFRASTRUCTURE: CUDA device built-ins for compiling the library's .cu sources
// with g++ (see emu_build.py).  Every launch is executed as grid*block blocks of ONE thread, sequentially.
#pragma once
#include <cuda_runtime.h>   // host view: dim3, uint3, cudaStream_t, cudaError_t, API prototypes (stubbed in emu_cudart.cpp)
#include <cuda.h>
#if __has_include(<nccl.h>)
#include <nccl.h>           // comm.cpp; pulled in here, before __CUDACC__ is defined, so that its CUDA headers stay in host mode
#endif
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

static uint3 threadIdx, blockIdx;
static dim3 blockDim, gridDim;
using std::min;
using std::max;

#define CZB_EMU_BUILD 1
#define __CUDACC__ 1            // let common.cuh define warp_sum / grid_reduce
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
#define __shared__              // one thread per block: shared variables are plain locals
#define __syncthreads() ((void)0)
#define __threadfence() ((void)0)
#define __threadfence_system() ((void)0)
template <class T> static inline T __ldg(const T* p) { return *p; }
template <class T> static inline T __ldcg(const T* p) { return *p; }
static inline double __shfl_xor_sync(unsigned, double, int) { return 0.0; }   // the other 31 lanes do not exist
static inline unsigned int atomicAdd(unsigned int* p, unsigned int v) { unsigned int o = *p; *p += v; return o; }
static inline unsigned int atomicAdd_system(unsigned int* p, unsigned int v) { return atomicAdd(p, v); }
static inline unsigned long long atomicCAS(unsigned long long* p, unsigned long long c, unsigned long long v) {
  unsigned long long o = *p; if (o == c) *p = v; return o;
}
static inline unsigned long long atomicExch(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; *p = v; return o; }
static inline double __longlong_as_double(long long v) { double d; memcpy(&d, &v, 8); return d; }
static inline long long __double_as_longlong(double d) { long long v; memcpy(&v, &d, 8); return v; }

#define cudaFuncSetAttribute(...) cudaSuccess   // kernels are plain functions here

template <class Body>
static void emu_launch(dim3 g, dim3 b, Body body) {
  const dim3 G(g.x * b.x, g.y * b.y, g.z * b.z);
  if ((long long)G.x * G.y * G.z > (1LL << 20)) { fprintf(stderr, "emu: grid too large for the reduction scratch\n"); abort(); }
  gridDim = G; blockDim = dim3(1, 1, 1);
  threadIdx.x = threadIdx.y = threadIdx.z = 0;
  for (unsigned z = 0; z < G.z; ++z)
    for (unsigned y = 0; y < G.y; ++y)
      for (unsigned x = 0; x < G.x; ++x) { blockIdx.x = x; blockIdx.y = y; blockIdx.z = z; body(); }
}

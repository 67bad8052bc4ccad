// tests/tools/emu/emu_cudart.cpp -- TEST INFRASTRUCTURE: the few CUDA runtime calls the library makes, on host memory
// (device memory = malloc, copies = memcpy, streams / events = tokens; everything is synchronous).
#include <cuda_runtime.h>
#include <cstdlib>
#include <cstring>
#include <map>

namespace {
std::map<const char*, size_t>& allocs() { static std::map<const char*, size_t> m; return m; }
int g_token;
}

extern "C" {
cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return cudaSuccess; }
cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaDeviceSynchronize(void) { return cudaSuccess; }
cudaError_t cudaGetLastError(void) { return cudaSuccess; }
const char* cudaGetErrorName(cudaError_t) { return "emu"; }
const char* cudaGetErrorString(cudaError_t) { return "emu"; }
cudaError_t cudaMalloc(void** p, size_t n) {
  *p = malloc(n ? n : 1);
  if (!*p) return cudaErrorMemoryAllocation;
  allocs()[(const char*)*p] = n ? n : 1;
  return cudaSuccess;
}
cudaError_t cudaFree(void* p) { if (p) { allocs().erase((const char*)p); free(p); } return cudaSuccess; }
cudaError_t cudaMallocHost(void** p, size_t n) { *p = calloc(1, n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
cudaError_t cudaHostAlloc(void** p, size_t n, unsigned int) { return cudaMallocHost(p, n); }   // cuda_runtime.h's C++ cudaMallocHost overload
cudaError_t cudaMemGetInfo(size_t* fr, size_t* tot) { *fr = *tot = (size_t)8 << 30; return cudaSuccess; }
// a `static __device__` array is a plain host array here: its own address is its "device" address
cudaError_t cudaGetSymbolAddress(void** p, const void* sym) { *p = const_cast<void*>(sym); return cudaSuccess; }
cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t*, void*) { return cudaErrorNotSupported; }
cudaError_t cudaIpcOpenMemHandle(void**, cudaIpcMemHandle_t, unsigned int) { return cudaErrorNotSupported; }
cudaError_t cudaIpcCloseMemHandle(void*) { return cudaSuccess; }
// kernels run at "launch" time in the emulation, so there is nothing a graph could replay: capture is refused
cudaError_t cudaStreamBeginCapture(cudaStream_t, cudaStreamCaptureMode) { return cudaErrorNotSupported; }
cudaError_t cudaStreamEndCapture(cudaStream_t, cudaGraph_t*) { return cudaErrorNotSupported; }
cudaError_t cudaGraphInstantiate(cudaGraphExec_t*, cudaGraph_t, unsigned long long) { return cudaErrorNotSupported; }
cudaError_t cudaGraphLaunch(cudaGraphExec_t, cudaStream_t) { return cudaErrorNotSupported; }
cudaError_t cudaGraphExecDestroy(cudaGraphExec_t) { return cudaSuccess; }
cudaError_t cudaGraphDestroy(cudaGraph_t) { return cudaSuccess; }
cudaError_t cudaFreeHost(void* p) { free(p); return cudaSuccess; }
cudaError_t cudaMemset(void* p, int v, size_t n) { memset(p, v, n); return cudaSuccess; }
cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { memset(p, v, n); return cudaSuccess; }
cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { memmove(d, s, n); return cudaSuccess; }
cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned int) { *s = (cudaStream_t)&g_token; return cudaSuccess; }
cudaError_t cudaStreamCreateWithPriority(cudaStream_t* s, unsigned int, int) { *s = (cudaStream_t)&g_token; return cudaSuccess; }
cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned int) { return cudaSuccess; }
cudaError_t cudaDeviceGetStreamPriorityRange(int* lo, int* hi) { *lo = 0; *hi = 0; return cudaSuccess; }
cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = (cudaEvent_t)&g_token; return cudaSuccess; }
cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned int) { *e = (cudaEvent_t)&g_token; return cudaSuccess; }
cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return cudaSuccess; }
cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }
cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
cudaError_t cudaPointerGetAttributes(cudaPointerAttributes* at, const void* p) {
  memset(at, 0, sizeof(*at));
  at->type = cudaMemoryTypeUnregistered;
  auto& m = allocs();
  auto it = m.upper_bound((const char*)p);
  if (it != m.begin()) {
    --it;
    if ((const char*)p < it->first + it->second) at->type = cudaMemoryTypeDevice;
  }
  return cudaSuccess;
}
cudaError_t cudaGetDriverEntryPoint(const char*, void** fn, unsigned long long, cudaDriverEntryPointQueryResult* q) {
  *fn = nullptr; if (q) *q = cudaDriverEntryPointSymbolNotFound; return cudaSuccess;
}
}

// cubez_b200/csrc/runtime.cu -- runtime singleton + the czb_* utility entry points of
// include/czb200.h (device selection, stream, allocation, counters).
#include "runtime.h"
#include "../../include/czb200.h"

namespace czb {

static Runtime g_rt;

void Runtime::init(int dev) {
  if (inited) return;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0) {
    fprintf(stderr,
            "czb200: no CUDA device available (%s). libczb200 is a GPU-only implementation of the CubeZ "
            "kernel layer; there is no CPU fallback.\n",
            e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    abort();
  }
  if (dev < 0) {
    const char* env = getenv("CZB_DEVICE");
    dev = env ? atoi(env) : -1;
    if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) dev = 0; }
  }
  device = dev;
  CZB_CHECK(cudaSetDevice(device));
  CZB_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  own_stream = true;
#ifdef CZB_EMU_BUILD   // tests/tools/emu: one emulated block per CUDA thread, so grids are blockDim times larger
  rb_blocks = (size_t)1 << 20;
#else
  rb_blocks = (size_t)1 << 17;
#endif
  CZB_CHECK(cudaMalloc(&rb.partial, sizeof(double) * rb_blocks * CZB_MAX_ACC));
  CZB_CHECK(cudaMalloc(&rb.ticket, sizeof(unsigned int)));
  CZB_CHECK(cudaMemset(rb.ticket, 0, sizeof(unsigned int)));
  CZB_CHECK(cudaMalloc(&d_scratch, sizeof(LoopState)));
  CZB_CHECK(cudaMemset(d_scratch, 0, sizeof(LoopState)));
  CZB_CHECK(cudaMallocHost(&h_state, sizeof(LoopState) * 4));
  CZB_CHECK(cudaMalloc(&d_err, sizeof(unsigned long long) * 4));
  CZB_CHECK(cudaDeviceSynchronize());
  { const char* e = getenv("CZB_PROFILE"); prof.on = e && atoi(e) != 0; }
  inited = true;
}

Runtime& rt() {
  if (!g_rt.inited) g_rt.init(-1);
  return g_rt;
}

double Runtime::read_acc(int slot) {
  CZB_CHECK(cudaMemcpyAsync(h_state, d_scratch, sizeof(LoopState), cudaMemcpyDeviceToHost, stream));
  CZB_CHECK(cudaStreamSynchronize(stream));
  return h_state->acc[slot];
}

void* Runtime::mirror_acquire(const void* host, size_t bytes) {
  if (mirror_cap == 0) {
    const char* e = getenv("CZB_MIRROR_CAP_MB");
    if (e) mirror_cap = (size_t)atoll(e) << 20;
    else { size_t fr = 0, tot = 0; CZB_CHECK(cudaMemGetInfo(&fr, &tot)); mirror_cap = tot / 4; }
    if (mirror_cap == 0) mirror_cap = 1;
  }
  auto it = mirrors.find(host);
  if (it != mirrors.end() && it->second.bytes < bytes && it->second.pins == 0) {   // grown: replace
    CZB_CHECK(cudaStreamSynchronize(stream));
    cudaFree(it->second.dev); mirror_bytes -= it->second.bytes; mirrors.erase(it); it = mirrors.end();
  }
  if (it != mirrors.end() && it->second.bytes < bytes) {
    fprintf(stderr, "czb200: host array %p is passed twice with different sizes in one call\n", host); abort();
  }
  if (it == mirrors.end()) {
    // evict least-recently-used, unpinned mirrors until the new one fits under the cap
    while (!mirrors.empty() && (mirror_bytes + bytes > mirror_cap || mirrors.size() >= 64)) {
      auto victim = mirrors.end();
      for (auto m = mirrors.begin(); m != mirrors.end(); ++m)
        if (m->second.pins == 0 && (victim == mirrors.end() || m->second.stamp < victim->second.stamp)) victim = m;
      if (victim == mirrors.end()) break;          // everything left is in use by the current call
      CZB_CHECK(cudaStreamSynchronize(stream));
      cudaFree(victim->second.dev); mirror_bytes -= victim->second.bytes; mirrors.erase(victim);
    }
    void* d = nullptr;
    CZB_CHECK(cudaMalloc(&d, bytes));
    it = mirrors.emplace(host, Mirror{d, bytes, 0ull, 0}).first;
    mirror_bytes += bytes;
  }
  it->second.stamp = ++mirror_clock;
  it->second.pins++;
  return it->second.dev;
}

void Runtime::mirror_release(const void* host) {
  auto it = mirrors.find(host);
  if (it != mirrors.end() && it->second.pins > 0) it->second.pins--;
}

// ---- phase profile ---------------------------------------------------------------------------------------------
int PhaseProf::id(const char* label) {
  for (size_t i = 0; i < acc.size(); ++i) if (acc[i].label == label) return (int)i;
  Acc a; a.label = label;
  acc.push_back(a);
  return (int)acc.size() - 1;
}
void PhaseProf::begin(const char* label, double flop, cudaStream_t s) {
  if (cur >= 0) return;                 // sections are exclusive: an inner one is folded into the running one
  if (open.size() >= 8192) drain();
  Rec r; r.label = id(label); r.flop = flop;
  for (cudaEvent_t* e : {&r.e0, &r.e1}) {
    if (!pool.empty()) { *e = pool.back(); pool.pop_back(); }
    else CZB_CHECK(cudaEventCreate(e));
  }
  CZB_CHECK(cudaEventRecord(r.e0, s));
  open.push_back(r);
  cur = (int)open.size() - 1;
}
void PhaseProf::end(cudaStream_t s) {
  if (cur < 0) return;
  CZB_CHECK(cudaEventRecord(open[cur].e1, s));
  cur = -1;
}
void PhaseProf::drain() {
  for (const Rec& r : open) {
    float ms = 0.f;
    CZB_CHECK(cudaEventSynchronize(r.e1));
    CZB_CHECK(cudaEventElapsedTime(&ms, r.e0, r.e1));
    Acc& a = acc[r.label];
    a.ms += ms; a.flop += r.flop; a.calls += 1;
    pool.push_back(r.e0); pool.push_back(r.e1);
  }
  open.clear();
  cur = -1;
}
void PhaseProf::reset() { drain(); acc.clear(); }
void PhaseProf::report(FILE* f, const char* title) {
  drain();
  double tot = 0.0;
  for (const Acc& a : acc) tot += a.ms;
  fprintf(f, "# %s -- device time per section (CUDA events on the solve stream; labels = the reference's PMlib sections,\n", title);
  fprintf(f, "# src/cz_cpp/cz_miscel.cpp:177-262; flop = the reference's analytic counts)\n");
  fprintf(f, "%-22s %10s %12s %8s %14s %10s\n", "Section", "call", "time[ms]", "[%]", "flop", "GFLOP/s");
  for (const Acc& a : acc)
    fprintf(f, "%-22s %10lld %12.3f %8.2f %14.6e %10.2f\n", a.label.c_str(), a.calls, a.ms, tot > 0 ? 100.0 * a.ms / tot : 0.0,
            a.flop, a.ms > 0 ? a.flop / a.ms * 1.0e-6 : 0.0);
  fprintf(f, "%-22s %10s %12.3f %8.2f\n", "Total (exclusive)", "", tot, 100.0);
}

bool is_device_ptr(const void* p) {
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

}  // namespace czb

using namespace czb;

extern "C" {

int czb_real_bytes(void) { return (int)sizeof(REAL); }
const char* czb_version(void) { return "czb200 0.1 (" CZB_REAL_NAME ", sm_100a)"; }

int czb_init(int device) {
  if (g_rt.inited) {
    if (device >= 0 && device != g_rt.device) {
      fprintf(stderr, "czb200: already initialised on device %d (requested %d)\n", g_rt.device, device);
      return -1;
    }
    return 0;
  }
  g_rt.init(device);
  return 0;
}

void czb_set_stream(void* cuda_stream) {
  Runtime& R = rt();
  CZB_CHECK(cudaStreamSynchronize(R.stream));
  if (R.own_stream) { cudaStreamDestroy(R.stream); R.own_stream = false; }
  R.stream = (cudaStream_t)cuda_stream;
}

void czb_sync(void) { rt().sync(); }

void* czb_malloc(size_t bytes) {
  Runtime& R = rt();
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes ? bytes : 1);
  if (e != cudaSuccess) {
    fprintf(stderr, "czb200: cudaMalloc(%zu) failed: %s\n", bytes, cudaGetErrorString(e));
    return nullptr;
  }
  CZB_CHECK(cudaMemsetAsync(p, 0, bytes, R.stream));
  return p;
}

void czb_free(void* p) {
  if (!p) return;
  rt().sync();
  cudaFree(p);
}

void czb_memcpy_h2d(void* dst, const void* src, size_t bytes) {
  Runtime& R = rt();
  CZB_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, R.stream));
  R.sync();
}
void czb_memcpy_d2h(void* dst, const void* src, size_t bytes) {
  Runtime& R = rt();
  CZB_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, R.stream));
  R.sync();
}

long long czb_launch_count(void) { return czb::launch_count(); }

void czb_profile(int on) {
  Runtime& R = rt();
  R.prof.reset();
  R.prof.on = on != 0;
}

int czb_profile_report(const char* path, const char* title) {
  Runtime& R = rt();
  R.sync();
  FILE* f = path ? fopen(path, "w") : stdout;
  if (!f) return -1;
  R.prof.report(f, title ? title : "czb200");
  if (path) fclose(f);
  return (int)R.prof.acc.size();
}

void czb_staged_bytes(double* h2d, double* d2h) {
  Runtime& R = rt();
  if (h2d) *h2d = (double)R.h2d_bytes;
  if (d2h) *d2h = (double)R.d2h_bytes;
}

}  // extern "C"

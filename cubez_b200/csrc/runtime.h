// cubez_b200/csrc/runtime.h -- process-wide runtime of libczb200: device, stream, reduction
// scratch, and the host-pointer staging used by the per-call reference ABI.
#pragma once
#include "kernels.cuh"
#include <unordered_map>
#include <cstring>
#include <string>
#include <vector>

namespace czb {

long long launch_count();

// Device mirror of a caller's host array (per-call reference ABI).  The cache is bounded: least recently used
// mirrors that no call is using are freed once the total exceeds Runtime::mirror_cap bytes or 64 entries.
struct Mirror { void* dev; size_t bytes; unsigned long long stamp; int pins; };

// Per-phase device times keyed by the reference's PMlib labels (src/cz_cpp/cz_miscel.cpp:177-262, the sections its
// TIMING_start/TIMING_stop bracket in cz_Poisson.cpp) -- the data of its profiling.txt (cz_Evaluate.cpp:506-544).
// Off by default (two event records per section); czb_profile(1) or CZB_PROFILE=1 switches it on, which also keeps
// the solve loops from replaying CUDA graphs (events cannot be timed inside a captured batch).
struct PhaseProf {
  bool on = false;
  struct Acc { std::string label; double ms = 0.0, flop = 0.0; long long calls = 0; };
  struct Rec { int label; cudaEvent_t e0, e1; double flop; };
  std::vector<Acc> acc;                 // first-use order
  std::vector<Rec> open;                // sections whose elapsed time has not been read yet
  std::vector<cudaEvent_t> pool;
  int cur = -1;                         // index into `open` of the running section
  int id(const char* label);
  void begin(const char* label, double flop, cudaStream_t s);
  void end(cudaStream_t s);
  void drain();                         // synchronises the events and accumulates
  void reset();
  void report(FILE* f, const char* title);
};

struct Runtime {
  bool inited = false;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  RedBuf rb{nullptr, nullptr};
  size_t rb_blocks = 0;
  LoopState* d_scratch = nullptr;     // device LoopState for per-call ABI reductions
  LoopState* h_state = nullptr;       // pinned host copy
  unsigned long long* d_err = nullptr;
  std::unordered_map<const void*, Mirror> mirrors;
  size_t mirror_bytes = 0, mirror_cap = 0;   // cap: CZB_MIRROR_CAP_MB, default a quarter of the device memory
  unsigned long long mirror_clock = 0;
  void* mirror_acquire(const void* host, size_t bytes);   // pins the mirror until mirror_release
  void mirror_release(const void* host);
  PcrTables pcr2{0, 0, 0, 0, nullptr, nullptr};
  PcrTables pcr4{0, 0, 0, 0, nullptr, nullptr};
  size_t h2d_bytes = 0, d2h_bytes = 0;   // staged traffic (per-call ABI with host pointers)
  PhaseProf prof;

  void init(int dev);
  Launch launch(const LoopState* st = nullptr, LoopState* acc = nullptr) {
    Launch L; L.s = stream; L.rb = rb; L.rb_blocks = rb_blocks; L.st = st; L.acc_st = acc ? acc : d_scratch;
    return L;
  }
  // read acc[slot] of the scratch state back to the host (synchronises the stream)
  double read_acc(int slot);
  void sync() { CZB_CHECK(cudaStreamSynchronize(stream)); }
};

Runtime& rt();   // initialises lazily on the current device (or CZB_DEVICE)

// Scoped section of the phase profile on the runtime's stream; costs one branch when profiling is off.
struct Phase {
  bool live;
  Phase(const char* label, double flop = 0.0) : live(rt().prof.on) { if (live) rt().prof.begin(label, flop, rt().stream); }
  ~Phase() { if (live) rt().prof.end(rt().stream); }
  Phase(const Phase&) = delete;
  Phase& operator=(const Phase&) = delete;
};

// Is `p` usable directly by kernels (device or managed memory)?
bool is_device_ptr(const void* p);

// Argument staging for the reference ABI (cz_Ffunc.h): arrays allocated by the caller with
// new[] are host memory; they are mirrored on the device for the duration of one call
// (H2D before when `in`, D2H after when `out`).  Device / managed pointers pass through.
template <class T>
class Arg {
 public:
  Arg(T* host_or_dev, size_t count, bool in, bool out) : p_(host_or_dev), n_(count), out_(out) {
    if (!p_ || is_device_ptr(p_)) { d_ = p_; staged_ = false; return; }
    Runtime& R = rt();
    const size_t bytes = sizeof(T) * n_;
    d_ = (T*)R.mirror_acquire((const void*)p_, bytes);
    staged_ = true;
    if (in) {
      CZB_CHECK(cudaMemcpyAsync(d_, p_, bytes, cudaMemcpyHostToDevice, R.stream));
      R.h2d_bytes += bytes;
    }
  }
  ~Arg() {
    if (!staged_) return;
    Runtime& R = rt();
    if (out_) {
      CZB_CHECK(cudaMemcpyAsync((void*)p_, d_, sizeof(T) * n_, cudaMemcpyDeviceToHost, R.stream));
      CZB_CHECK(cudaStreamSynchronize(R.stream));
      R.d2h_bytes += sizeof(T) * n_;
    }
    R.mirror_release((const void*)p_);
  }
  Arg(const Arg&) = delete;
  Arg& operator=(const Arg&) = delete;
  T* dev() const { return d_; }
  bool staged() const { return staged_; }
 private:
  T* p_; T* d_; size_t n_; bool out_; bool staged_;
};

}  // namespace czb

// tests/tools/emu/nccl_emu.cpp -- TEST INFRASTRUCTURE: the nine NCCL entry points cubez_b200/csrc/comm.cpp uses, over
// files in a shared directory (env CZB_NCCL_EMU_DIR), for world_size-2 runs of the CPU-emulated library: one OS
// process per rank, exactly as on the GPUs.  Loaded through comm.cpp's own CZB_NCCL_LIB switch.  Everything is
// synchronous; sends never block (a message is a file), receives posted inside a group are completed at ncclGroupEnd.
#include <cuda_runtime.h>
#include <nccl.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <map>
#include <unistd.h>
#include <sys/stat.h>
#include <chrono>

struct ncclComm {
  std::string dir;
  int n, rank;
  std::map<int, long> sent, received;   // per peer sequence numbers
  long ar_seq = 0;
};

namespace {
int g_group = 0;
struct Pending { void* buf; size_t bytes; int peer; ncclComm* c; };
std::vector<Pending> g_pending;

size_t type_bytes(ncclDataType_t t) { return t == ncclFloat64 ? 8 : t == ncclFloat32 ? 4 : 0; }

void write_file(const std::string& path, const void* data, size_t bytes) {
  const std::string tmp = path + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) { fprintf(stderr, "nccl_emu: cannot write %s\n", tmp.c_str()); abort(); }
  fwrite(data, 1, bytes, f);
  fclose(f);
  rename(tmp.c_str(), path.c_str());   // atomic: a reader never sees a partial message
}

void read_file_blocking(const std::string& path, void* data, size_t bytes) {
  const auto t0 = std::chrono::steady_clock::now();
  struct stat st;
  while (stat(path.c_str(), &st) != 0) {
    usleep(100);
    if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(120)) {
      fprintf(stderr, "nccl_emu: timed out waiting for %s\n", path.c_str());
      abort();
    }
  }
  FILE* f = fopen(path.c_str(), "rb");
  if (!f || fread(data, 1, bytes, f) != bytes) { fprintf(stderr, "nccl_emu: short read %s\n", path.c_str()); abort(); }
  fclose(f);
}

void do_recv(const Pending& p) {
  char name[256];
  snprintf(name, sizeof(name), "/p2p_%d_%d_%ld.bin", p.peer, p.c->rank, p.c->received[p.peer]++);
  read_file_blocking(p.c->dir + name, p.buf, p.bytes);
  unlink((p.c->dir + name).c_str());
}
}  // namespace

extern "C" {

ncclResult_t ncclGetUniqueId(ncclUniqueId* id) {
  memset(id, 0, sizeof(*id));
  snprintf(id->internal, sizeof(id->internal), "emu_%d_%ld", (int)getpid(), (long)time(nullptr));
  return ncclSuccess;
}

ncclResult_t ncclCommInitRank(ncclComm_t* comm, int n, ncclUniqueId id, int rank) {
  const char* base = getenv("CZB_NCCL_EMU_DIR");
  if (!base) { fprintf(stderr, "nccl_emu: CZB_NCCL_EMU_DIR not set\n"); return ncclInvalidUsage; }
  ncclComm* c = new ncclComm;
  c->dir = std::string(base) + "/" + id.internal;
  c->n = n; c->rank = rank;
  mkdir(c->dir.c_str(), 0700);
  *comm = c;
  return ncclSuccess;
}

ncclResult_t ncclCommDestroy(ncclComm_t c) { delete c; return ncclSuccess; }

ncclResult_t ncclSend(const void* buf, size_t count, ncclDataType_t t, int peer, ncclComm_t c, cudaStream_t) {
  char name[256];
  snprintf(name, sizeof(name), "/p2p_%d_%d_%ld.bin", c->rank, peer, c->sent[peer]++);
  write_file(c->dir + name, buf, count * type_bytes(t));
  return ncclSuccess;
}

ncclResult_t ncclRecv(void* buf, size_t count, ncclDataType_t t, int peer, ncclComm_t c, cudaStream_t) {
  Pending p{buf, count * type_bytes(t), peer, c};
  if (g_group > 0) g_pending.push_back(p); else do_recv(p);
  return ncclSuccess;
}

ncclResult_t ncclGroupStart() { ++g_group; return ncclSuccess; }
ncclResult_t ncclGroupEnd() {
  if (--g_group == 0) { for (const Pending& p : g_pending) do_recv(p); g_pending.clear(); }
  return ncclSuccess;
}

// sum in rank order on every rank: deterministic and identical everywhere
ncclResult_t ncclAllReduce(const void* send, void* recv, size_t count, ncclDataType_t t, ncclRedOp_t op, ncclComm_t c,
                           cudaStream_t) {
  if (t != ncclFloat64 || op != ncclSum) return ncclInvalidArgument;
  char name[256];
  snprintf(name, sizeof(name), "/ar_%ld_%d.bin", c->ar_seq, c->rank);
  write_file(c->dir + name, send, count * 8);
  std::vector<double> acc(count, 0.0), part(count);
  for (int r = 0; r < c->n; ++r) {
    snprintf(name, sizeof(name), "/ar_%ld_%d.bin", c->ar_seq, r);
    read_file_blocking(c->dir + name, part.data(), count * 8);
    for (size_t i = 0; i < count; ++i) acc[i] += part[i];
  }
  memcpy(recv, acc.data(), count * 8);
  ++c->ar_seq;
  return ncclSuccess;
}

const char* ncclGetErrorString(ncclResult_t) { return "nccl_emu error"; }
}

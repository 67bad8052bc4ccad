// cubez_b200/csrc/comm.cpp -- see comm.h.  Replaces reference src/cz_cpp/cz_comm.cpp:
// Comm_S (CBrick 6-face halo, :23-38) -> ncclSend/ncclRecv of contiguous j-planes over NVLink;
// Comm_SUM_1 (MPI_Allreduce, :74-140) -> ncclAllReduce of device-resident doubles.
#include "comm.h"
#include <nccl.h>
#include <cuda.h>
#include "kernels.cuh"
#include <dlfcn.h>
#include <cstring>

namespace czb {

int slab_decompose(const int gsz[3], int nranks, int rank, int sz[3], int head[3], int nID[6], int idx[6]) {
  const int n = gsz[1] - 2;   // inner planes
  if (nranks < 1 || rank < 0 || rank >= nranks || n < nranks) return -1;
  const int base = n / nranks, rem = n % nranks;
  const int cnt = base + (rank < rem ? 1 : 0);
  const int first = 2 + rank * base + (rank < rem ? rank : rem);   // global j of first owned inner plane
  const bool lo_edge = (rank == 0), hi_edge = (rank == nranks - 1);
  sz[0] = gsz[0]; sz[2] = gsz[2];
  sz[1] = cnt + (lo_edge ? 1 : 0) + (hi_edge ? 1 : 0);
  head[0] = 1; head[2] = 1;
  head[1] = lo_edge ? 1 : first;
  for (int f = 0; f < 6; ++f) nID[f] = -1;
  nID[2] = lo_edge ? -1 : rank - 1;   // J_MINUS  (cz_fparam.fi:10-16)
  nID[3] = hi_edge ? -1 : rank + 1;   // J_PLUS
  idx[0] = 2; idx[1] = sz[0] - 1;
  idx[2] = lo_edge ? 2 : 1;
  idx[3] = hi_edge ? sz[1] - 1 : sz[1];
  idx[4] = 2; idx[5] = sz[2] - 1;
  return 0;
}

// General 3-D brick decomposition (the CBrick role, cz_Evaluate.cpp:103-159): every axis is cut like the j axis of
// slab_decompose; rank = ri + gdv[0] * (rj + gdv[1] * rk), so gdv = {1, N, 1} gives exactly the j-slabs.
int brick_decompose(const int gsz[3], const int gdv[3], int rank, int sz[3], int head[3], int nID[6], int idx[6]) {
  const int nranks = gdv[0] * gdv[1] * gdv[2];
  if (gdv[0] < 1 || gdv[1] < 1 || gdv[2] < 1 || rank < 0 || rank >= nranks) return -1;
  const int r3[3] = {rank % gdv[0], (rank / gdv[0]) % gdv[1], rank / (gdv[0] * gdv[1])};
  const int stride[3] = {1, gdv[0], gdv[0] * gdv[1]};
  for (int a = 0; a < 3; ++a) {
    const int n = gsz[a] - 2, parts = gdv[a], r = r3[a];
    if (n < parts) return -1;
    const int base = n / parts, rem = n % parts;
    const int cnt = base + (r < rem ? 1 : 0);
    const int first = 2 + r * base + (r < rem ? r : rem);   // global index of the first owned inner node
    const bool lo_edge = (r == 0), hi_edge = (r == parts - 1);
    sz[a] = cnt + (lo_edge ? 1 : 0) + (hi_edge ? 1 : 0);
    head[a] = lo_edge ? 1 : first;
    nID[2 * a] = lo_edge ? -1 : rank - stride[a];            // X_MINUS, X_PLUS, Y_MINUS, ... (cz_fparam.fi:10-16)
    nID[2 * a + 1] = hi_edge ? -1 : rank + stride[a];
    idx[2 * a] = lo_edge ? 2 : 1;
    idx[2 * a + 1] = hi_edge ? sz[a] - 1 : sz[a];
  }
  return 0;
}

namespace {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi& api() {
  static NcclApi A;
  if (A.lib) return A;
  const char* names[] = {getenv("CZB_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    if (!nm) continue;
    A.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (A.lib) break;
  }
  if (!A.lib) { fprintf(stderr, "czb200: cannot load NCCL (libnccl.so.2): %s\n", dlerror()); abort(); }
#define CZB_SYM(field, name)                                                     \
  *(void**)(&A.field) = dlsym(A.lib, name);                                      \
  if (!A.field) { fprintf(stderr, "czb200: NCCL symbol %s missing\n", name); abort(); }
  CZB_SYM(GetUniqueId, "ncclGetUniqueId");
  CZB_SYM(CommInitRank, "ncclCommInitRank");
  CZB_SYM(CommDestroy, "ncclCommDestroy");
  CZB_SYM(Send, "ncclSend");
  CZB_SYM(Recv, "ncclRecv");
  CZB_SYM(AllReduce, "ncclAllReduce");
  CZB_SYM(GroupStart, "ncclGroupStart");
  CZB_SYM(GroupEnd, "ncclGroupEnd");
  CZB_SYM(GetErrorString, "ncclGetErrorString");
#undef CZB_SYM
  return A;
}

#define CZB_NCCL(call)                                                                         \
  do {                                                                                         \
    ncclResult_t r_ = (call);                                                                  \
    if (r_ != ncclSuccess) {                                                                   \
      fprintf(stderr, "czb200: NCCL error at %s:%d: %s\n", __FILE__, __LINE__, api().GetErrorString(r_)); \
      abort();                                                                                 \
    }                                                                                          \
  } while (0)

inline ncclDataType_t real_type() { return sizeof(REAL) == 8 ? ncclFloat64 : ncclFloat32; }

}  // namespace

int comm_unique_id(void* id128) {
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  CZB_NCCL(api().GetUniqueId(&id));
  memcpy(id128, &id, 128);
  return 0;
}

Comm* comm_create(int nranks, int rank, const void* id128, const int nID[6]) {
  Comm* c = new Comm;
  c->nranks = nranks; c->rank = rank; c->lo = nID[2]; c->hi = nID[3];
  for (int f = 0; f < 6; ++f) c->nb[f] = nID[f];
  if (nranks > 1) {
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclComm_t comm;
    CZB_NCCL(api().CommInitRank(&comm, nranks, id, rank));
    c->nccl = (void*)comm;
  }
  return c;
}

// First use of a NCCL peer connection / of the all-reduce ring costs hundreds of milliseconds (transport set-up);
// pay it when the communicator is made, not inside the first timed solver loop.  The all-reduce also carries the
// one collective fact every rank must agree on: whether the decomposition has i or k neighbours anywhere (bricks).
void comm_warmup(Comm* c, cudaStream_t s) {
  if (!c || c->nranks <= 1 || !c->nccl) return;
  NcclApi& A = api();
  ncclComm_t comm = (ncclComm_t)c->nccl;
  double* d = nullptr;
  CZB_CHECK(cudaMalloc((void**)&d, 16 * sizeof(double)));
  CZB_CHECK(cudaMemsetAsync(d, 0, 16 * sizeof(double), s));
  const double mine = (c->nb[0] >= 0 || c->nb[1] >= 0 || c->nb[4] >= 0 || c->nb[5] >= 0) ? 1.0 : 0.0;
  CZB_CHECK(cudaMemcpyAsync(d + 12, &mine, sizeof(double), cudaMemcpyHostToDevice, s));
  CZB_NCCL(A.GroupStart());
  for (int f = 0; f < 6; ++f)
    if (c->nb[f] >= 0) {
      CZB_NCCL(A.Send(d + 2 * f, 1, ncclFloat64, c->nb[f], comm, s));
      CZB_NCCL(A.Recv(d + 2 * f + 1, 1, ncclFloat64, c->nb[f], comm, s));
    }
  CZB_NCCL(A.GroupEnd());
  CZB_NCCL(A.AllReduce(d + 12, d + 12, 1, ncclFloat64, ncclSum, comm, s));
  double any = 0.0;
  CZB_CHECK(cudaMemcpyAsync(&any, d + 12, sizeof(double), cudaMemcpyDeviceToHost, s));
  CZB_CHECK(cudaStreamSynchronize(s));
  c->brick = any > 0.0;
  cudaFree(d);
}

void comm_destroy(Comm* c) {
  if (!c) return;
  for (int d = 0; d < 2; ++d) if (c->peer_block[d]) cudaIpcCloseMemHandle(c->peer_block[d]);
  if (c->block) cudaFree(c->block);
  if (c->stage) cudaFree(c->stage);
  if (c->nccl) api().CommDestroy((ncclComm_t)c->nccl);
  delete c;
}

// Header of the exported block; the inbox starts at the next 1 KB boundary (TMA needs 16-byte aligned planes).
static const size_t kP2pHeader = 1024;

bool comm_p2p_setup(Comm* c, size_t plane_elems, cudaStream_t s, bool want) {
  if (!c || c->nranks <= 1 || !c->nccl) return false;
  if (c->p2p_tried) return c->p2p && c->plane_elems == plane_elems;   // decided once per communicator (collectively)
  c->p2p_tried = true;
  static const bool off = getenv("CZB_NO_P2P") != nullptr;
  NcclApi& A = api();
  ncclComm_t comm = (ncclComm_t)c->nccl;
  int ok = (off || !want) ? 0 : 1;
  struct Wire { cudaIpcMemHandle_t h; int ok; int pad[15]; };   // 128 bytes
  static_assert(sizeof(Wire) == 128, "wire record");
  Wire mine; memset(&mine, 0, sizeof(mine));
  const size_t bytes = kP2pHeader + 8 * plane_elems * sizeof(REAL);
  if (ok && cudaMalloc(&c->block, bytes) != cudaSuccess) { cudaGetLastError(); c->block = nullptr; ok = 0; }
  if (ok) {
    CZB_CHECK(cudaMemsetAsync(c->block, 0, bytes, s));
    if (cudaIpcGetMemHandle(&mine.h, c->block) != cudaSuccess) { cudaGetLastError(); ok = 0; }
  }
  mine.ok = ok;
  // exchange the records with both neighbours (device staging: NCCL moves device memory)
  Wire* d = nullptr;
  CZB_CHECK(cudaMalloc((void**)&d, 3 * sizeof(Wire)));
  CZB_CHECK(cudaMemsetAsync(d, 0, 3 * sizeof(Wire), s));
  CZB_CHECK(cudaMemcpyAsync(d, &mine, sizeof(Wire), cudaMemcpyHostToDevice, s));
  CZB_NCCL(A.GroupStart());
  if (c->lo >= 0) { CZB_NCCL(A.Send(d, sizeof(Wire), ncclChar, c->lo, comm, s)); CZB_NCCL(A.Recv(d + 1, sizeof(Wire), ncclChar, c->lo, comm, s)); }
  if (c->hi >= 0) { CZB_NCCL(A.Send(d, sizeof(Wire), ncclChar, c->hi, comm, s)); CZB_NCCL(A.Recv(d + 2, sizeof(Wire), ncclChar, c->hi, comm, s)); }
  CZB_NCCL(A.GroupEnd());
  Wire got[3];
  CZB_CHECK(cudaMemcpyAsync(got, d, 3 * sizeof(Wire), cudaMemcpyDeviceToHost, s));
  CZB_CHECK(cudaStreamSynchronize(s));
  const int nb[2] = {c->lo, c->hi};
  for (int side = 0; side < 2 && ok; ++side) {
    if (nb[side] < 0) continue;
    if (!got[1 + side].ok) { ok = 0; break; }
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, got[1 + side].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
    c->peer_block[side] = p;
  }
  // all ranks or none: a ring with one NCCL link would deadlock the stream waits
  double* flag = nullptr;
  CZB_CHECK(cudaMalloc((void**)&flag, sizeof(double)));
  const double f = ok ? 0.0 : 1.0;
  CZB_CHECK(cudaMemcpyAsync(flag, &f, sizeof(double), cudaMemcpyHostToDevice, s));
  CZB_NCCL(A.AllReduce(flag, flag, 1, ncclFloat64, ncclSum, comm, s));
  double bad = 0.0;
  CZB_CHECK(cudaMemcpyAsync(&bad, flag, sizeof(double), cudaMemcpyDeviceToHost, s));
  CZB_CHECK(cudaStreamSynchronize(s));
  cudaFree(flag); cudaFree(d);
  if (bad != 0.0) {
    for (int side = 0; side < 2; ++side) if (c->peer_block[side]) { cudaIpcCloseMemHandle(c->peer_block[side]); c->peer_block[side] = nullptr; }
    if (c->block) { cudaFree(c->block); c->block = nullptr; }
    if (!off && c->rank == 0) fprintf(stderr, "czb200: peer transport unavailable (IPC mapping failed on some rank); halo exchange stays on NCCL\n");
    return false;
  }
  c->plane_elems = plane_elems;
  c->mbox = (unsigned int*)c->block;
  c->inbox = (REAL*)((char*)c->block + kP2pHeader);
  for (int side = 0; side < 2; ++side) {
    if (!c->peer_block[side]) continue;
    // we are the UPPER neighbour of our lower neighbour: our signals count in its from_hi (word 1), and vice versa
    c->peer_mbox[side] = (unsigned int*)c->peer_block[side] + (side == 0 ? 1 : 0);
    c->peer_inbox[side] = (REAL*)((char*)c->peer_block[side] + kP2pHeader);
  }
  c->tgt[0] = c->tgt[1] = 0u;
  c->p2p = true;
  return true;
}

namespace {
// one thread: make the preceding copies of this stream visible system-wide, then count the push in the neighbours' mailboxes
__global__ void k_post(unsigned int* lo, unsigned int* hi) {
  __threadfence_system();
  if (lo) atomicAdd_system(lo, 1u);
  if (hi) atomicAdd_system(hi, 1u);
}
}  // namespace

void stream_wait_geq(cudaStream_t s, unsigned int* dev, unsigned int value) {
  typedef CUresult (*Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
  static Fn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CZB_CHECK(cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q));
    if (!p || q != cudaDriverEntryPointSuccess) { fprintf(stderr, "czb200: cuStreamWaitValue32 unavailable\n"); abort(); }
    fn = (Fn)p;
  }
  const CUresult r = fn((CUstream)s, (CUdeviceptr)(uintptr_t)dev, value, CU_STREAM_WAIT_VALUE_GEQ);
  if (r != CUDA_SUCCESS) { fprintf(stderr, "czb200: cuStreamWaitValue32 failed (%d)\n", (int)r); abort(); }
}

void comm_p2p_push(Comm* c, const REAL* lo_planes, const REAL* hi_planes, int w, int slot, cudaStream_t s) {
  const size_t P = c->plane_elems, bytes = sizeof(REAL) * P * (size_t)w;
  // inbox plane index = slot*4 + side*2 + p ; we are the UPPER neighbour (side 1) of our lower neighbour and vice versa
  if (c->lo >= 0) { CZB_CHECK(cudaMemcpyAsync(c->peer_inbox[0] + (size_t)(slot * 4 + 2) * P, lo_planes, bytes, cudaMemcpyDeviceToDevice, s)); c->tgt[0] += 1u; }
  if (c->hi >= 0) { CZB_CHECK(cudaMemcpyAsync(c->peer_inbox[1] + (size_t)(slot * 4 + 0) * P, hi_planes, bytes, cudaMemcpyDeviceToDevice, s)); c->tgt[1] += 1u; }
  k_post<<<1, 1, 0, s>>>(c->lo >= 0 ? c->peer_mbox[0] : nullptr, c->hi >= 0 ? c->peer_mbox[1] : nullptr);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

void comm_p2p_wait(Comm* c, cudaStream_t s) {
  if (c->lo >= 0) stream_wait_geq(s, c->mbox + 0, c->tgt[0]);
  if (c->hi >= 0) stream_wait_geq(s, c->mbox + 1, c->tgt[1]);
}

void comm_p2p_fetch(Comm* c, REAL* lo_guard, REAL* hi_guard, int w, int slot, cudaStream_t s) {
  const size_t P = c->plane_elems, bytes = sizeof(REAL) * P * (size_t)w;
  // the neighbour's planes are stored from plane 0 of its region: with w = 1 the plane next to us is plane 0 of
  // side 1 (its first owned plane) but plane 0 of side 0 is its LAST owned plane only if it pushed w = 1 -- the
  // sender always passes the w planes adjacent to the interface in ascending j order
  if (c->lo >= 0) CZB_CHECK(cudaMemcpyAsync(lo_guard, c->inbox + (size_t)(slot * 4 + 0) * P, bytes, cudaMemcpyDeviceToDevice, s));
  if (c->hi >= 0) CZB_CHECK(cudaMemcpyAsync(hi_guard, c->inbox + (size_t)(slot * 4 + 2) * P, bytes, cudaMemcpyDeviceToDevice, s));
}

namespace {
// Box [k0, k0+nkb) x [i0, i0+nib) x all nj planes of f  <->  dense buffer buf[(jc * nib + ir) * nkb + kq]
__global__ void k_box(Geo G, REAL* f, REAL* buf, int k0, int nkb, int i0, int nib, int to_buf) {
  const long long total = (long long)nkb * nib * G.nj;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int kq = (int)(e % nkb);
    const long long r = e / nkb;
    const int ir = (int)(r % nib), jc = (int)(r / nib);
    const long long off = (long long)jc * G.plane + (long long)(i0 + ir) * G.nk + (k0 + kq);
    if (to_buf) buf[e] = f[off]; else f[off] = buf[e];
  }
}

void box_copy(const Geo& G, REAL* f, REAL* buf, int k0, int nkb, int i0, int nib, int to_buf, cudaStream_t s) {
  const long long total = (long long)nkb * nib * G.nj;
  int grid = (int)((total + 255) / 256);
  if (grid > 1184) grid = 1184;
  if (grid < 1) grid = 1;
  k_box<<<grid, 256, 0, s>>>(G, f, buf, k0, nkb, i0, nib, to_buf);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}
}  // namespace

void comm_halo_n(Comm* c, REAL* const* fields, int nf, const Geo& G, int w, cudaStream_t s) {
  if (!c || c->nranks <= 1) return;
  bool any = false;
  for (int f = 0; f < 6; ++f) any = any || c->nb[f] >= 0;
  if (!any) return;
  NcclApi& A = api();
  ncclComm_t comm = (ncclComm_t)c->nccl;
  const size_t cnt = (size_t)w * (size_t)G.plane;
  const long long P = G.plane;
  // i faces: w rows of nk cells in every plane; k faces: w cells of every row -- strided, so they travel through
  // dense staging buffers: [field][face X- X+ Z- Z+][send, recv]
  const size_t ei = (size_t)w * G.nk * G.nj, ek = (size_t)w * G.ni * G.nj;
  const bool ik = c->nb[0] >= 0 || c->nb[1] >= 0 || c->nb[4] >= 0 || c->nb[5] >= 0;
  const size_t per_field = 4 * ei + 4 * ek;
  if (ik && c->stage_elems < per_field * (size_t)nf) {
    if (c->stage) { CZB_CHECK(cudaStreamSynchronize(s)); cudaFree(c->stage); }
    c->stage_elems = per_field * (size_t)nf;
    CZB_CHECK(cudaMalloc((void**)&c->stage, sizeof(REAL) * c->stage_elems));
  }
  auto sbuf = [&](int q, int face) -> REAL* {   // face: 0 X-, 1 X+, 2 Z-, 3 Z+
    return c->stage + (size_t)q * per_field + (face < 2 ? (size_t)face * 2 * ei : 4 * ei + (size_t)(face - 2) * 2 * ek);
  };
  const int g = G.g;
  if (ik)
    for (int q = 0; q < nf; ++q) {   // pack the owned edge layers
      if (c->nb[0] >= 0) box_copy(G, fields[q], sbuf(q, 0), 0, G.nk, g, w, 1, s);
      if (c->nb[1] >= 0) box_copy(G, fields[q], sbuf(q, 1), 0, G.nk, G.ix + g - w, w, 1, s);
      if (c->nb[4] >= 0) box_copy(G, fields[q], sbuf(q, 2), g, w, 0, G.ni, 1, s);
      if (c->nb[5] >= 0) box_copy(G, fields[q], sbuf(q, 3), G.kx + g - w, w, 0, G.ni, 1, s);
    }
  CZB_NCCL(A.GroupStart());          // one group = one NCCL launch for all fields and all neighbours
  for (int q = 0; q < nf; ++q) {
    REAL* f = fields[q];
    if (c->lo >= 0) {   // my planes 1..w -> lower neighbour's guard; its top planes -> my planes 1-w..0
      CZB_NCCL(A.Send(f + (long long)G.g * P, cnt, real_type(), c->lo, comm, s));
      CZB_NCCL(A.Recv(f + (long long)(G.g - w) * P, cnt, real_type(), c->lo, comm, s));
    }
    if (c->hi >= 0) {   // my planes jx-w+1..jx -> upper neighbour's guard; its planes 1..w -> my jx+1..jx+w
      CZB_NCCL(A.Send(f + (long long)(G.jx - w + G.g) * P, cnt, real_type(), c->hi, comm, s));
      CZB_NCCL(A.Recv(f + (long long)(G.jx + G.g) * P, cnt, real_type(), c->hi, comm, s));
    }
    const int fnb[4] = {c->nb[0], c->nb[1], c->nb[4], c->nb[5]};
    for (int face = 0; face < 4; ++face)
      if (fnb[face] >= 0) {
        const size_t e = face < 2 ? ei : ek;
        CZB_NCCL(A.Send(sbuf(q, face), e, real_type(), fnb[face], comm, s));
        CZB_NCCL(A.Recv(sbuf(q, face) + e, e, real_type(), fnb[face], comm, s));
      }
  }
  CZB_NCCL(A.GroupEnd());
  if (ik)
    for (int q = 0; q < nf; ++q) {   // unpack into the guard layers
      if (c->nb[0] >= 0) box_copy(G, fields[q], sbuf(q, 0) + ei, 0, G.nk, g - w, w, 0, s);
      if (c->nb[1] >= 0) box_copy(G, fields[q], sbuf(q, 1) + ei, 0, G.nk, G.ix + g, w, 0, s);
      if (c->nb[4] >= 0) box_copy(G, fields[q], sbuf(q, 2) + ek, g - w, w, 0, G.ni, 0, s);
      if (c->nb[5] >= 0) box_copy(G, fields[q], sbuf(q, 3) + ek, G.kx + g, w, 0, G.ni, 0, s);
    }
}

void comm_halo(Comm* c, REAL* f, const Geo& G, int w, cudaStream_t s) { comm_halo_n(c, &f, 1, G, w, s); }

void comm_allreduce_sum(Comm* c, double* dev, int count, cudaStream_t s) {
  if (!c || c->nranks <= 1) return;
  CZB_NCCL(api().AllReduce(dev, dev, (size_t)count, ncclFloat64, ncclSum, (ncclComm_t)c->nccl, s));
}

}  // namespace czb

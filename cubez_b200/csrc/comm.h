// cubez_b200/csrc/comm.h -- j-slab decomposition and the NCCL transport that replaces the
// reference's CBrick halo exchange + MPI_Allreduce (src/cz_cpp/cz_comm.cpp:23-38, 74-325).
// NCCL is loaded with dlopen only when a multi-rank context is created, so the single-GPU
// library has no NCCL dependency.
#pragma once
#include "common.cuh"

namespace czb {

// Pure host logic (testable without a GPU).  Global nodes 1..gsz[1] along j are cut into
// contiguous slabs; the physical boundary planes j=1 and j=gsz[1] belong to the end ranks,
// the gsz[1]-2 inner planes are spread as evenly as possible (remainder to the first ranks).
// Local index range follows CZ::range_inner_index (cz_miscel.cpp:20-52) except that a rank
// with a lower neighbour starts at jst=1 (its local plane 1 is an owned inner plane; the
// neighbour's plane lives in the guard layer).
int slab_decompose(const int gsz[3], int nranks, int rank, int sz[3], int head[3], int nID[6], int idx[6]);

// 3-D bricks: gdv = divisions along i (x), j (y), k (z); see comm.cpp
int brick_decompose(const int gsz[3], const int gdv[3], int rank, int sz[3], int head[3], int nID[6], int idx[6]);

struct Comm {
  int nranks = 1, rank = 0;
  int lo = -1, hi = -1;     // neighbour ranks along j (-1: none)
  int nb[6] = {-1, -1, -1, -1, -1, -1};   // all six neighbours: X_MINUS X_PLUS Y_MINUS Y_PLUS Z_MINUS Z_PLUS (i, j, k axes)
  bool brick = false;       // some rank of the communicator has an i or k neighbour (decided collectively in comm_warmup)
  REAL* stage = nullptr;    // pack / unpack buffers of the i and k faces (strided in memory; j planes are contiguous)
  size_t stage_elems = 0;
  void* nccl = nullptr;     // ncclComm_t
  // ---- peer transport (NVLink peer memory through the copy engines, no NCCL and no SM in the halo path) ----------
  // Every rank owns one IPC-exported block:  mailbox {from_lo, from_hi} (32-bit counters the NEIGHBOURS increment
  // with a system-scope atomic after each push) followed by the halo inbox  [2 slots][2 sides][2 planes]  of
  // plane_elems REALs each (side 0 = planes owned by the lower neighbour, 1 = by the upper one).  Iteration `it`
  // of a solve loop pushes the slab's edge planes into the neighbours' slot it&1 with cudaMemcpyAsync over the
  // mapped pointers (DMA, overlaps the interior kernel), then posts; the consumer's STREAM waits on its mailbox
  // (cuStreamWaitValue32 -- no kernel spins) and copies the planes from its inbox into its guard planes.
  bool p2p = false;
  bool p2p_tried = false;
  size_t plane_elems = 0;
  void* block = nullptr;                    // local block (cudaMalloc)
  unsigned int* mbox = nullptr;             // local counters {from_lo, from_hi}
  REAL* inbox = nullptr;                    // local inbox, 8 planes
  void* peer_block[2] = {nullptr, nullptr};  // mapped blocks of the lower / upper neighbour
  unsigned int* peer_mbox[2] = {nullptr, nullptr};   // the word of THEIR mailbox that counts OUR pushes
  REAL* peer_inbox[2] = {nullptr, nullptr};
  unsigned int tgt[2] = {0u, 0u};           // value from_lo / from_hi reach once every push expected so far has landed
};

int comm_unique_id(void* id128);
Comm* comm_create(int nranks, int rank, const void* id128, const int nID[6]);
void comm_destroy(Comm* c);
void comm_warmup(Comm* c, cudaStream_t s);   // one tiny exchange with each neighbour + one all-reduce (connection set-up)
// Exchange `width` guard layers with the neighbours (guard cells <- neighbour's edge cells): j-planes directly; on
// bricks also the i and k faces, packed into dense buffers by a kernel, all in one NCCL group.  Faces only -- the
// 7-point stencil reads no edge or corner guard cell.
void comm_halo(Comm* c, REAL* field, const Geo& G, int width, cudaStream_t s);
// the same for several fields in ONE NCCL group (one launch): the exchanges that open a solve loop
void comm_halo_n(Comm* c, REAL* const* fields, int nf, const Geo& G, int width, cudaStream_t s);
// Set up the peer transport for fields with planes of `plane_elems` REALs: allocate and export the local block,
// exchange the IPC handles with the ring neighbours, map theirs.  Collective over the communicator; returns true
// only if EVERY rank succeeded (otherwise all ranks keep the NCCL transport).  CZB_NO_P2P=1 disables it.
bool comm_p2p_setup(Comm* c, size_t plane_elems, cudaStream_t s, bool want = true);
// `want`: this rank can use the transport (kernel eligibility); the result is the AND over all ranks.
// Push `w` (1 or 2) planes starting at lo_planes (-> lower neighbour) and hi_planes (-> upper neighbour) into the
// neighbours' inbox slot, then post to their mailboxes; everything on stream s.  The neighbours run the same program,
// so every push also raises the value comm_p2p_wait will wait for.
void comm_p2p_push(Comm* c, const REAL* lo_planes, const REAL* hi_planes, int w, int slot, cudaStream_t s);
// Stream s waits until every push expected so far has landed in our inbox.
void comm_p2p_wait(Comm* c, cudaStream_t s);
// Copy `w` planes per side from our inbox slot into the guard planes lo_guard / hi_guard (device pointers).
void comm_p2p_fetch(Comm* c, REAL* lo_guard, REAL* hi_guard, int w, int slot, cudaStream_t s);
// cuStreamWaitValue32(GEQ) through the runtime's driver entry point (no link dependency on libcuda)
void stream_wait_geq(cudaStream_t s, unsigned int* dev, unsigned int value);
// In-place sum over ranks of `count` doubles in device memory.
void comm_allreduce_sum(Comm* c, double* dev, int count, cudaStream_t s);

}  // namespace czb

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

struct Comm {
  int nranks = 1, rank = 0;
  int lo = -1, hi = -1;     // neighbour ranks along j (-1: none)
  void* nccl = nullptr;     // ncclComm_t
};

int comm_unique_id(void* id128);
Comm* comm_create(int nranks, int rank, const void* id128, int lo, int hi);
void comm_destroy(Comm* c);
// Exchange `width` j-planes with both neighbours (guard planes <- neighbour's edge planes).
void comm_halo(Comm* c, REAL* field, const Geo& G, int width, cudaStream_t s);
// In-place sum over ranks of `count` doubles in device memory.
void comm_allreduce_sum(Comm* c, double* dev, int count, cudaStream_t s);

}  // namespace czb

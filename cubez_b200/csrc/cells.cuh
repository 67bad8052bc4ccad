// cubez_b200/csrc/cells.cuh -- the generic per-inner-cell kernel shared by kernels.cu and maf.cu: one thread
// per (k,i) position marching a chunk of j-planes, k (the contiguous axis) on threadIdx.x so that every warp
// load/store is a contiguous 128 B run; deterministic grid reduction of up to CZB_MAX_ACC double sums.
#pragma once
#include "kernels.cuh"

namespace czb {
namespace {

struct Tile {
  dim3 grid, block;
  int jlen;
};

// (k,i) tile of ~256 threads, k fastest; j split into chunks so that the grid has at least
// ~8 blocks per SM even for small cubes.
Tile make_tile(const Geo& G, int kcells, size_t max_blocks) {
  Tile t;
  int bx = 32;
  while (bx < kcells && bx < 256) bx += 32;
  int by = 256 / bx; if (by < 1) by = 1;
  const int ni = G.ied - G.ist + 1, nj = G.jed - G.jst + 1;
  if (by > ni) by = ni > 0 ? ni : 1;
  t.block = dim3(bx, by, 1);
  const int gx = (kcells + bx - 1) / bx, gy = (ni + by - 1) / by;
  // longest j-chunks that still give ~64 blocks per SM: few, fat blocks keep the last-block
  // reduction short while leaving enough blocks to balance the tail.
  const long long target = 148LL * 64;
  int jlen = nj > 0 ? nj : 1;
  while (jlen > 4 && (long long)gx * gy * ((nj + jlen - 1) / jlen) < target) jlen = (jlen + 1) / 2;
  int gz = (nj + jlen - 1) / jlen; if (gz < 1) gz = 1;
  while ((size_t)gx * gy * gz > max_blocks && jlen < nj) { jlen <<= 1; gz = (nj + jlen - 1) / jlen; }
  t.jlen = jlen;
  t.grid = dim3(gx > 0 ? gx : 1, gy > 0 ? gy : 1, gz);
  return t;
}

struct FinStore {   // acc[slot] (+)= total
  LoopState* st; int slot; int add;
  __device__ void operator()(double (&tot)[1]) const {
    if (st) st->acc[slot] = (add ? st->acc[slot] : 0.0) + tot[0];
  }
};

// Generic per-inner-cell kernel.  Op::operator()(G,k,i,j,off,acc) is called for every inner
// cell of the thread's (k,i) column in its j-chunk.
template <int NACC, class Op, class Fin>
__global__ void __launch_bounds__(256) k_cells(Geo G, int jlen, const LoopState* st, Op op, RedBuf rb, Fin fin) {
  if (st && st->done) return;
  const int k = G.kst + blockIdx.x * blockDim.x + threadIdx.x;
  const int i = G.ist + blockIdx.y * blockDim.y + threadIdx.y;
  const int j0 = G.jst + blockIdx.z * jlen;
  const int j1 = min(j0 + jlen - 1, G.jed);
  double acc[NACC > 0 ? NACC : 1];
#pragma unroll
  for (int a = 0; a < (NACC > 0 ? NACC : 1); ++a) acc[a] = 0.0;
  if (k <= G.ked && i <= G.ied) {
    long long off = G.at(k, i, j0);
    for (int j = j0; j <= j1; ++j, off += G.plane) op(G, k, i, j, off, acc);
  }
  if constexpr (NACC > 0) grid_reduce<NACC>(acc, rb, fin);
}

template <int NACC, class Op, class Fin>
void run_cells(const Launch& L, const Geo& G, Op op, Fin fin) {
  const int kcells = G.ked - G.kst + 1;
  Tile t = make_tile(G, kcells, L.rb_blocks);
  k_cells<NACC><<<t.grid, t.block, 0, L.s>>>(G, t.jlen, L.st, op, L.rb, fin);
  CZB_CHECK(cudaGetLastError());
  count_launch();
}

struct FinNone { __device__ void operator()(double (&)[1]) const {} };

}  // namespace
}  // namespace czb

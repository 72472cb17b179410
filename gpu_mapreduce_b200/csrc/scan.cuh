// Prefix-sum building blocks used for KV/KMV page packing: every "emit" in the reference is a sequential
// append (KeyValue::add, src/keyvalue.cpp:343-392); on the GPU the append position of record i is the
// exclusive prefix sum of the record sizes before it, so packing = sizes -> scan -> coalesced scatter.
//
// scan_apply(n, value(i), sink(i, exclusive_prefix, value)) runs as three launches: per-tile reduce, a
// single-block scan of the tile totals, and a per-tile scan that hands each element its prefix.  The input
// functor is evaluated twice (once per pass) instead of materialising an n-sized temporary in HBM.
#pragma once

#include "common.cuh"

namespace mrgpu {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_IPT = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_IPT;

__device__ __forceinline__ unsigned long long warp_incl_scan(unsigned long long v, int lane)
{
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// block-wide exclusive scan of one value per thread (blockDim.x <= 1024); returns exclusive prefix, and the
// block total through `total`
__device__ __forceinline__ unsigned long long block_excl_scan(unsigned long long v, unsigned long long &total,
                                                              unsigned long long *sh /*[32]*/)
{
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  unsigned long long incl = warp_incl_scan(v, lane);
  if (lane == 31) sh[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = (lane < nw) ? sh[lane] : 0;
    unsigned long long wi = warp_incl_scan(w, lane);
    sh[lane] = wi - w;
    if (lane == 31) sh[32] = wi;
  }
  __syncthreads();
  unsigned long long r = sh[warp] + incl - v;
  total = sh[32];
  __syncthreads();
  return r;
}

template <typename F>
__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(uint64_t n, F value, unsigned long long *tile_sums)
{
  __shared__ unsigned long long sh[33];
  uint64_t base = (uint64_t) blockIdx.x * SCAN_TILE;
  unsigned long long s = 0;
#pragma unroll
  for (int j = 0; j < SCAN_IPT; j++) {
    uint64_t i = base + (uint64_t) j * SCAN_THREADS + threadIdx.x;
    if (i < n) s += value(i);
  }
  unsigned long long total;
  block_excl_scan(s, total, sh);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// in-place exclusive scan of m values by one block; writes the grand total to *total
__global__ void __launch_bounds__(1024) scan_tiles_kernel(unsigned long long *v, uint64_t m, unsigned long long *total)
{
  __shared__ unsigned long long sh[33];
  unsigned long long carry = 0;
  for (uint64_t base = 0; base < m; base += 1024) {
    uint64_t i = base + threadIdx.x;
    unsigned long long x = (i < m) ? v[i] : 0, t;
    unsigned long long e = block_excl_scan(x, t, sh);
    if (i < m) v[i] = carry + e;
    carry += t;
  }
  if (threadIdx.x == 0) *total = carry;
}

template <typename F, typename G>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_apply_kernel(uint64_t n, F value, G sink, const unsigned long long *__restrict__ tile_excl)
{
  __shared__ unsigned long long sh[33];
  // blocked arrangement: thread t owns SCAN_IPT consecutive elements, so prefixes are exact in index order
  uint64_t base = (uint64_t) blockIdx.x * SCAN_TILE + (uint64_t) threadIdx.x * SCAN_IPT;
  unsigned long long x[SCAN_IPT], s = 0;
#pragma unroll
  for (int j = 0; j < SCAN_IPT; j++) {
    uint64_t i = base + j;
    x[j] = (i < n) ? value(i) : 0;
    s += x[j];
  }
  unsigned long long total;
  unsigned long long e = block_excl_scan(s, total, sh) + tile_excl[blockIdx.x];
#pragma unroll
  for (int j = 0; j < SCAN_IPT; j++) {
    uint64_t i = base + j;
    if (i < n) sink(i, e, x[j]);
    e += x[j];
  }
}

// Two-phase form: scan_tiles() reduces every tile and scans the tile totals (grand total -> *d_total), so
// the host can size an output from the total; scan_sink() then hands each element its exclusive prefix.
// A ScanTiles can feed several scan_sink() passes over the same value functor.
struct ScanTiles {
  Buf<unsigned long long> sums;
  uint64_t ntiles = 0, n = 0;
};

template <typename F>
void scan_tiles(Pool &pool, uint64_t n, F value, ScanTiles &t, unsigned long long *d_total, const char *tag = "scan_reduce", double bytes = 0)
{
  cudaStream_t s = pool.stream;
  t.n = n;
  t.ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  if (n == 0) {
    MR_CUDA(cudaMemsetAsync(d_total, 0, 8, s));
    return;
  }
  t.sums.alloc(pool, t.ntiles);
  ProfScope ps(pool, tag, bytes, 2);
  scan_reduce_kernel<<<(unsigned) t.ntiles, SCAN_THREADS, 0, s>>>(n, value, t.sums.p);
  MR_CUDA(cudaGetLastError());
  scan_tiles_kernel<<<1, 1024, 0, s>>>(t.sums.p, t.ntiles, d_total);
  MR_CUDA(cudaGetLastError());
}

template <typename F, typename G>
void scan_sink(Pool &pool, const ScanTiles &t, F value, G sink, const char *tag = "scan_apply", double bytes = 0)
{
  if (t.n == 0) return;
  ProfScope ps(pool, tag, bytes, 1);
  scan_apply_kernel<<<(unsigned) t.ntiles, SCAN_THREADS, 0, pool.stream>>>(t.n, value, sink, t.sums.p);
  MR_CUDA(cudaGetLastError());
}

// One-shot exclusive scan.  d_total receives the sum (device u64).
template <typename F, typename G>
void scan_apply(Pool &pool, uint64_t n, F value, G sink, unsigned long long *d_total, const char *tag = "scan", double bytes = 0)
{
  ScanTiles t;
  scan_tiles(pool, n, value, t, d_total, tag, 0);
  scan_sink(pool, t, value, sink, tag, bytes);
}

}  // namespace mrgpu

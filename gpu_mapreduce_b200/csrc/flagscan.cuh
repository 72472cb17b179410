// Prefix count of 0/1 flags with coalesced (warp-striped) element access.
//
// Run detection, word-start detection and every "does element i open a new output record" question is a
// prefix sum of flags.  For flags a warp needs no shuffle scan: one ballot per 32 elements and popc give
// every lane its exclusive count, and lane l of round j touches element base + 32 j + l, so the loads and
// the scattered stores issued by the sink are contiguous across the warp (the generic scan_apply hands each
// thread 8 consecutive elements, which turns every 4-byte access into a 32-byte-strided one).
//
// flag_tiles(n, flag) counts per tile and scans the tile totals; flag_sink(n, flag, sink) calls
// sink(i, exclusive_count, flag) for every element.  flag(i) is evaluated once per element per pass.
#pragma once

#include "scan.cuh"

namespace mrgpu {

constexpr int FS_THREADS = 256;
constexpr int FS_IPT = 16;
constexpr int FS_TILE = FS_THREADS * FS_IPT;

template <typename F>
__global__ void __launch_bounds__(FS_THREADS) flag_count_kernel(uint64_t n, F flag, unsigned long long *tile_counts)
{
  __shared__ unsigned wsum[FS_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t wbase = (uint64_t) blockIdx.x * FS_TILE + (uint64_t) warp * (32 * FS_IPT);
  unsigned cnt = 0;
#pragma unroll
  for (int j = 0; j < FS_IPT; j++) {
    uint64_t i = wbase + j * 32 + lane;
    bool f = i < n && flag(i);
    cnt += __popc(__ballot_sync(0xFFFFFFFFu, f));
  }
  if (lane == 0) wsum[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned t = 0;
#pragma unroll
    for (int w = 0; w < FS_THREADS / 32; w++) t += wsum[w];
    tile_counts[blockIdx.x] = t;
  }
}

template <typename F, typename G>
__global__ void __launch_bounds__(FS_THREADS)
flag_sink_kernel(uint64_t n, F flag, G sink, const unsigned long long *__restrict__ tile_excl)
{
  __shared__ unsigned wsum[FS_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t wbase = (uint64_t) blockIdx.x * FS_TILE + (uint64_t) warp * (32 * FS_IPT);
  unsigned bal[FS_IPT], cnt = 0;
#pragma unroll
  for (int j = 0; j < FS_IPT; j++) {
    uint64_t i = wbase + j * 32 + lane;
    bool f = i < n && flag(i);
    bal[j] = __ballot_sync(0xFFFFFFFFu, f);
    cnt += __popc(bal[j]);
  }
  if (lane == 0) wsum[warp] = cnt;
  __syncthreads();
  unsigned long long run = tile_excl[blockIdx.x];
#pragma unroll
  for (int w = 0; w < FS_THREADS / 32; w++)
    if (w < warp) run += wsum[w];
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (int j = 0; j < FS_IPT; j++) {
    uint64_t i = wbase + j * 32 + lane;
    if (i < n) sink(i, run + __popc(bal[j] & lt), (unsigned long long) ((bal[j] >> lane) & 1));
    run += __popc(bal[j]);
  }
}

struct FlagTiles {
  Buf<unsigned long long> sums;
  uint64_t ntiles = 0, n = 0;
};

template <typename F>
void flag_tiles(Pool &pool, uint64_t n, F flag, FlagTiles &t, unsigned long long *d_total, const char *tag = "flag_count", double bytes = 0)
{
  cudaStream_t s = pool.stream;
  t.n = n;
  t.ntiles = (n + FS_TILE - 1) / FS_TILE;
  if (n == 0) {
    MR_CUDA(cudaMemsetAsync(d_total, 0, 8, s));
    return;
  }
  t.sums.alloc(pool, t.ntiles);
  ProfScope ps(pool, tag, bytes, 2);
  flag_count_kernel<<<(unsigned) t.ntiles, FS_THREADS, 0, s>>>(n, flag, t.sums.p);
  MR_CUDA(cudaGetLastError());
  scan_tiles_kernel<<<1, 1024, 0, s>>>(t.sums.p, t.ntiles, d_total);
  MR_CUDA(cudaGetLastError());
}

template <typename F, typename G>
void flag_sink(Pool &pool, const FlagTiles &t, F flag, G sink, const char *tag = "flag_sink", double bytes = 0)
{
  if (t.n == 0) return;
  ProfScope ps(pool, tag, bytes, 1);
  flag_sink_kernel<<<(unsigned) t.ntiles, FS_THREADS, 0, pool.stream>>>(t.n, flag, sink, t.sums.p);
  MR_CUDA(cudaGetLastError());
}

}  // namespace mrgpu

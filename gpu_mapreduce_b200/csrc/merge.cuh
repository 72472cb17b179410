// Merge of two sorted key arrays, and the detection of a long sorted run inside an unsorted one.
//
// An iterative MapReduce (R-MAT: generate, collate, cull, repeat -- reference examples/rmat.cpp:137-153) hands
// convert() a KV whose bulk is the previous round's output, still in key order, with a few percent of new pairs
// before and after it.  Sorting everything again costs one radix pass per varying key byte over the whole array;
// sorting only the new keys and merging them with the run costs one read and one write of it.
//
// merge_sorted(): merge path (the diagonal of the |A| x |B| grid that output position d lies on is searched for
// the point where A[i-1] <= B[d-i] and B[d-i-1] < A[i]); one search per tile in global memory, one per thread in
// shared memory, then MG_IPT sequential steps per thread and a coalesced store through shared memory.
#pragma once

#include "common.cuh"
#include "flagscan.cuh"

namespace mrgpu {

constexpr int MG_THREADS = 256;
constexpr int MG_IPT = 8;
constexpr int MG_TILE = MG_THREADS * MG_IPT;

// number of A elements among the first `diag` outputs; equal keys take A's first
template <typename K, typename I>
__host__ __device__ __forceinline__ I merge_path(const K *A, I na, const K *B, I nb, I diag)
{
  I lo = diag > nb ? diag - nb : 0, hi = diag < na ? diag : na;
  while (lo < hi) {
    const I mid = (lo + hi) >> 1;
    if (A[mid] <= B[diag - 1 - mid]) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// one thread's share of a tile: MG_IPT outputs starting at diagonal d of the tile's two runs (sa, sb)
template <typename K>
__host__ __device__ __forceinline__ void merge_steps(const K *sa, unsigned la, const K *sb, unsigned lb, unsigned d, K (&r)[MG_IPT])
{
  unsigned i = merge_path<K, unsigned>(sa, la, sb, lb, d), j = d - i;
#pragma unroll
  for (int q = 0; q < MG_IPT; q++) {
    const bool ha = i < la, hb = j < lb;
    const K x = ha ? sa[i] : K(0), y = hb ? sb[j] : K(0);
    const bool ta = ha && (!hb || x <= y);
    r[q] = ta ? x : y;
    i += ta;
    j += !ta && hb;
  }
}

template <typename K>
__global__ void __launch_bounds__(256)
merge_split_kernel(const K *__restrict__ A, uint64_t na, const K *__restrict__ B, uint64_t nb, uint64_t ntiles, uint64_t *__restrict__ split)
{
  const uint64_t t = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (t > ntiles) return;
  const uint64_t n = na + nb;
  uint64_t d = t * MG_TILE;
  if (d > n) d = n;
  split[t] = merge_path<K, uint64_t>(A, na, B, nb, d);
}

template <typename K>
__global__ void __launch_bounds__(MG_THREADS)
merge_tiles_kernel(const K *__restrict__ A, uint64_t na, const K *__restrict__ B, uint64_t nb, const uint64_t *__restrict__ split,
                   K *__restrict__ out)
{
  __shared__ K sh[MG_TILE];
  const uint64_t n = na + nb, t = blockIdx.x;
  const uint64_t d0 = t * MG_TILE, d1 = d0 + MG_TILE < n ? d0 + MG_TILE : n;
  const uint64_t a0 = split[t], a1 = split[t + 1], b0 = d0 - a0, b1 = d1 - a1;
  const unsigned la = (unsigned) (a1 - a0), lb = (unsigned) (b1 - b0), m = la + lb;
  for (unsigned i = threadIdx.x; i < m; i += MG_THREADS) sh[i] = i < la ? A[a0 + i] : B[b0 + (i - la)];
  __syncthreads();
  const K *sa = sh, *sb = sh + la;
  unsigned d = threadIdx.x * MG_IPT;
  if (d > m) d = m;
  K r[MG_IPT];
  merge_steps<K>(sa, la, sb, lb, d, r);
  __syncthreads();
#pragma unroll
  for (int q = 0; q < MG_IPT; q++)
    if (threadIdx.x * MG_IPT + q < m) sh[threadIdx.x * MG_IPT + q] = r[q];
  __syncthreads();
  for (unsigned q = threadIdx.x; q < m; q += MG_THREADS) out[d0 + q] = sh[q];
}

template <typename K>
static void merge_sorted(Pool &pool, const K *A, uint64_t na, const K *B, uint64_t nb, K *out)
{
  const uint64_t n = na + nb;
  if (!n) return;
  cudaStream_t s = pool.stream;
  const uint64_t ntiles = (n + MG_TILE - 1) / MG_TILE;
  Buf<uint64_t> split(pool, ntiles + 1);
  ProfScope ps(pool, "merge_sorted", 2.0 * n * sizeof(K));
  merge_split_kernel<K><<<(unsigned) ((ntiles + 1 + 255) / 256), 256, 0, s>>>(A, na, B, nb, ntiles, split.p);
  merge_tiles_kernel<K><<<(unsigned) ntiles, MG_THREADS, 0, s>>>(A, na, B, nb, split.p, out);
  MR_CUDA(cudaGetLastError());
}

// ---- where is the order already there ----------------------------------------------------------------------

// descents inside `nblk` windows of `win` consecutive keys spread evenly over the array: a cheap test that tells
// random input (every second step is a descent) from input that is mostly in order
template <typename K>
__global__ void __launch_bounds__(256)
descent_sample_kernel(const K *__restrict__ k, uint64_t n, uint64_t win, unsigned long long *count)
{
  const uint64_t base = (n - win) / (gridDim.x > 1 ? gridDim.x - 1 : 1) * blockIdx.x;
  unsigned c = 0;
  for (uint64_t i = threadIdx.x + 1; i < win; i += blockDim.x) c += k[base + i] < k[base + i - 1];
  c = __reduce_add_sync(0xFFFFFFFFu, c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, (unsigned long long) c);
}

// One pass over the keys: how many descents there are (k[i] < k[i-1]), and the run that holds the middle element --
// it starts at the last descent at or before n/2 and ends at the first one after it.  A run that covers at least half
// of the keys contains the middle element, so no list of descents is needed to find it.
// out[0] += descents, out[1] = max(start), out[2] = min(end); the caller sets out = {0, 0, n}.
template <typename K>
__global__ void __launch_bounds__(256)
descent_scan_kernel(const K *__restrict__ k, uint64_t n, uint64_t mid, unsigned long long *out)
{
  __shared__ unsigned long long sh[3][8];
  unsigned long long cnt = 0, a = 0, b = n;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += stride)
    if (__ldg(k + i) < __ldg(k + i - 1)) {
      cnt++;
      if (i <= mid) a = i > a ? i : a;
      else b = i < b ? i : b;
    }
  for (int o = 16; o; o >>= 1) {
    cnt += __shfl_down_sync(0xFFFFFFFFu, cnt, o);
    const unsigned long long ta = __shfl_down_sync(0xFFFFFFFFu, a, o), tb = __shfl_down_sync(0xFFFFFFFFu, b, o);
    a = ta > a ? ta : a;
    b = tb < b ? tb : b;
  }
  if ((threadIdx.x & 31) == 0) sh[0][threadIdx.x >> 5] = cnt, sh[1][threadIdx.x >> 5] = a, sh[2][threadIdx.x >> 5] = b;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; w++) {
      cnt += sh[0][w];
      a = sh[1][w] > a ? sh[1][w] : a;
      b = sh[2][w] < b ? sh[2][w] : b;
    }
    if (cnt) {
      atomicAdd(out, cnt);
      atomicMax(out + 1, a);
      atomicMin(out + 2, b);
    }
  }
}

}  // namespace mrgpu

// Stable LSD radix sort of (key, payload) pairs -- the engine under convert() (group-by-key) and
// sort_keys().  It replaces the reference's per-page indirect qsort + 2-way spool merge
// (src/mapreduce.cpp:2462-2633) and the chained-hash passes of convert (src/keymultivalue.cpp:645-789,
// 1139-1209): on a GPU a stable partition by key digits is streaming HBM traffic, a chained table is not.
//
// Structure ("onesweep"): one histogram kernel counts all key bytes in a single read of the keys; then one
// kernel per *varying* byte does rank + decoupled look-back + scatter in a single pass over the data
// (read n, write n).  Bytes whose histogram has one non-empty bin are skipped entirely -- IntCount's 2^24
// distinct keys cost 3 passes, not 4; R-MAT vertices at scale 24 cost 3, not 8.
//
// HBM-bound, no tensor cores.  Tile = 256 threads x IPT keys staged through shared memory so that each
// digit's run leaves the SM as one contiguous, coalesced burst.
#pragma once

#include "common.cuh"

namespace mrgpu {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int SORT_THREADS = 256;  // == RADIX: thread d owns digit d in the block-wide steps
constexpr int SORT_WARPS = SORT_THREADS / 32;

// ---------------------------------------------------------------------------------------------
// histogram of every byte of every key: hist[byte * 256 + digit]

template <typename K>
__device__ __forceinline__ void hist_accumulate(K key, unsigned (&prev)[sizeof(K)], unsigned (&run)[sizeof(K)], unsigned *sh)
{
  // run-length pre-aggregation: a byte that is constant over the data (or over a hot Zipf key) costs one
  // shared atomic per run instead of one per key, so skewed digits do not serialise the shared-memory ALU.
#pragma unroll
  for (int b = 0; b < (int) sizeof(K); b++) {
    unsigned d = (unsigned) (key >> (8 * b)) & 0xFF;
    if (d != prev[b]) {
      if (run[b]) atomicAdd(&sh[b * RADIX + prev[b]], run[b]);
      prev[b] = d;
      run[b] = 0;
    }
    run[b]++;
  }
}

template <typename K>
__device__ __forceinline__ void hist_flush(unsigned (&prev)[sizeof(K)], unsigned (&run)[sizeof(K)], unsigned *sh,
                                           unsigned long long *hist)
{
#pragma unroll
  for (int b = 0; b < (int) sizeof(K); b++)
    if (run[b]) atomicAdd(&sh[b * RADIX + prev[b]], run[b]);
  __syncthreads();
  for (int i = threadIdx.x; i < (int) sizeof(K) * RADIX; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], (unsigned long long) sh[i]);
}

template <typename K>
__global__ void __launch_bounds__(256) radix_hist_kernel(const K *__restrict__ keys, uint64_t n, unsigned long long *hist)
{
  __shared__ unsigned sh[sizeof(K) * RADIX];
  for (int i = threadIdx.x; i < (int) sizeof(K) * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  unsigned prev[sizeof(K)], run[sizeof(K)];
#pragma unroll
  for (int b = 0; b < (int) sizeof(K); b++) prev[b] = 0, run[b] = 0;
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    hist_accumulate<K>(keys[i], prev, run, sh);
  hist_flush<K>(prev, run, sh, hist);
}

// ---------------------------------------------------------------------------------------------
// one onesweep pass over byte `shift/8`

constexpr unsigned long long LB_AGG = 1ull << 62;   // tile-local count published
constexpr unsigned long long LB_INC = 2ull << 62;   // inclusive prefix published
constexpr unsigned long long LB_VAL = (1ull << 62) - 1;

struct NoPayload {};

template <typename K, typename P, int IPT>
struct SortSmem {
  static constexpr int TILE = SORT_THREADS * IPT;
  static constexpr int XBYTES = (sizeof(K) > sizeof(P) ? sizeof(K) : sizeof(P)) * TILE;
};

template <typename K, typename P, bool HASP, int IPT>
__global__ void __launch_bounds__(SORT_THREADS)
onesweep_kernel(const K *__restrict__ kin, K *__restrict__ kout, const P *__restrict__ pin, P *__restrict__ pout, uint64_t n,
                int shift, const unsigned long long *__restrict__ gexcl /*[256] exclusive global digit offsets*/,
                unsigned long long *status /*[ntiles*256]*/, unsigned *tile_counter)
{
  constexpr int TILE = SORT_THREADS * IPT;
  __shared__ unsigned wc[SORT_WARPS][RADIX];
  __shared__ __align__(16) unsigned char xbuf[SortSmem<K, P, IPT>::XBYTES];
  __shared__ unsigned long long gbase[RADIX];
  __shared__ unsigned dstart[RADIX];
  __shared__ unsigned warp_tot[SORT_WARPS];
  __shared__ unsigned s_tile;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // tiles are claimed in launch order so that every tile a look-back waits on is already running
  if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
#pragma unroll
  for (int w = 0; w < SORT_WARPS; w++) wc[w][tid] = 0;
  __syncthreads();
  const uint64_t tile = s_tile;
  const uint64_t tbase = tile * TILE;
  const unsigned nvalid = (unsigned) ((n - tbase) < (uint64_t) TILE ? (n - tbase) : (uint64_t) TILE);

  // ---- load: warp-striped, coalesced; out-of-range slots hold the maximum key
  K key[IPT];
  const uint64_t wbase = tbase + (uint64_t) warp * (32 * IPT);
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    uint64_t i = wbase + j * 32 + lane;
    key[j] = (i < n) ? kin[i] : (K) ~(K) 0;
  }

  // ---- rank inside the warp, digit by digit, in memory order (stable)
  unsigned short rank[IPT];
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    unsigned d = (unsigned) (key[j] >> shift) & 0xFF;
    unsigned peers = __match_any_sync(0xFFFFFFFFu, d);
    unsigned lt = __popc(peers & lanemask_lt());
    unsigned base = 0;
    if (lt == 0) {
      base = wc[warp][d];
      wc[warp][d] = base + __popc(peers);
    }
    base = __shfl_sync(0xFFFFFFFFu, base, __ffs(peers) - 1);
    rank[j] = (unsigned short) (base + lt);
    __syncwarp();
  }
  __syncthreads();

  // ---- thread d: digit d's count in this tile, and each warp's exclusive offset for it
  unsigned cnt = 0;
#pragma unroll
  for (int w = 0; w < SORT_WARPS; w++) {
    unsigned c = wc[w][tid];
    wc[w][tid] = cnt;
    cnt += c;
  }
  // publish for the tiles behind us as early as possible
  unsigned long long *st = status + tile * RADIX + tid;
  if (tile == 0) atomicExch(st, LB_INC | cnt);
  else atomicExch(st, LB_AGG | cnt);

  // exclusive scan of cnt over the 256 digits -> start of each digit's run inside the tile
  unsigned incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  unsigned woff = 0;
#pragma unroll
  for (int w = 0; w < SORT_WARPS; w++)
    if (w < warp) woff += warp_tot[w];
  const unsigned my_start = woff + incl - cnt;
  dstart[tid] = my_start;

  // ---- decoupled look-back for digit d
  unsigned long long excl = 0;
  if (tile > 0) {
    long long t = (long long) tile - 1;
    while (true) {
      unsigned long long v;
      const volatile unsigned long long *pv = status + (uint64_t) t * RADIX + tid;
      do {
        v = *pv;
      } while ((v >> 62) == 0);
      excl += v & LB_VAL;
      if ((v >> 62) == 2) break;
      t--;
    }
    atomicExch(st, LB_INC | (excl + cnt));
  }
  gbase[tid] = gexcl[tid] + excl - my_start;
  __syncthreads();

  // ---- stage keys in tile-sorted order
  K *xk = reinterpret_cast<K *>(xbuf);
  unsigned short pos[IPT];
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    unsigned d = (unsigned) (key[j] >> shift) & 0xFF;
    pos[j] = (unsigned short) (dstart[d] + wc[warp][d] + rank[j]);
    xk[pos[j]] = key[j];
  }
  __syncthreads();

  // ---- scatter: consecutive threads write consecutive slots of a digit's run
  K okey[IPT];
#pragma unroll
  for (int k = 0; k < IPT; k++) {
    unsigned s = k * SORT_THREADS + tid;
    if (s < nvalid) {
      okey[k] = xk[s];
      unsigned d = (unsigned) (okey[k] >> shift) & 0xFF;
      kout[gbase[d] + s] = okey[k];
    }
  }
  if constexpr (HASP) {
    P pay[IPT];
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      uint64_t i = wbase + j * 32 + lane;
      if (i < n) pay[j] = pin[i];
    }
    __syncthreads();
    P *xp = reinterpret_cast<P *>(xbuf);
#pragma unroll
    for (int j = 0; j < IPT; j++) xp[pos[j]] = pay[j];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < IPT; k++) {
      unsigned s = k * SORT_THREADS + tid;
      if (s < nvalid) {
        unsigned d = (unsigned) (okey[k] >> shift) & 0xFF;
        pout[gbase[d] + s] = xp[s];
      }
    }
  }
}

// exclusive scan of each byte's 256-bin histogram, in place; one block per byte
__global__ void radix_scan_hist_kernel(unsigned long long *hist)
{
  __shared__ unsigned long long wt[8];
  unsigned long long *h = hist + blockIdx.x * RADIX;
  int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long v = h[tid], incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) wt[warp] = incl;
  __syncthreads();
  unsigned long long off = 0;
  for (int w = 0; w < warp; w++) off += wt[w];
  h[tid] = off + incl - v;
}

// ---------------------------------------------------------------------------------------------
// host driver

struct SortScratch {
  // all sizes in bytes; see sort_scratch_bytes()
  unsigned long long *hist;     // [8*256] raw counts, then exclusive offsets
  unsigned long long *status;   // [ntiles*256] per pass (re-zeroed between passes)
  unsigned *counter;            // [1]
};

template <typename K>
constexpr int sort_ipt() { return 16; }

inline uint64_t sort_tiles(uint64_t n, int ipt) { return (n + (uint64_t) SORT_THREADS * ipt - 1) / ((uint64_t) SORT_THREADS * ipt); }

// Sorts n pairs.  Data starts in (k0,p0); (k1,p1) is the ping-pong partner.  Returns 0 if the sorted data
// ended in buffer 0, 1 if in buffer 1.  `byte_mask` selects which key bytes may need sorting (others are
// known constant); bytes whose histogram shows a single non-empty bin are skipped as well.
// host_hist (optional, [sizeof(K)*256]) = already computed raw histogram on the device in scratch.hist.
template <typename K, typename P, bool HASP>
int radix_sort_pairs(Pool &pool, K *k0, K *k1, P *p0, P *p1, uint64_t n, unsigned byte_mask, bool hist_ready,
                     unsigned long long *d_hist /*[sizeof(K)*256] device, zeroed or filled*/)
{
  constexpr int IPT = sort_ipt<K>();
  cudaStream_t s = pool.stream;
  if (n == 0) return 0;
  if (!hist_ready) {
    MR_CUDA(cudaMemsetAsync(d_hist, 0, sizeof(K) * RADIX * 8, s));
    ProfScope ps(pool, "radix_hist", (double) n * sizeof(K));
    radix_hist_kernel<K><<<grid_for(n, 256 * 16, 8), 256, 0, s>>>(k0, n, d_hist);
    MR_CUDA(cudaGetLastError());
  }
  unsigned long long h[sizeof(K) * RADIX];
  MR_CUDA(cudaMemcpyAsync(h, d_hist, sizeof(h), cudaMemcpyDeviceToHost, s));
  MR_CUDA(cudaStreamSynchronize(s));
  unsigned need = 0;
  for (int b = 0; b < (int) sizeof(K); b++) {
    if (!((byte_mask >> b) & 1)) continue;
    bool single = false;
    for (int d = 0; d < RADIX; d++)
      if (h[b * RADIX + d] == n) single = true;
    if (!single) need |= 1u << b;
  }
  if (!need) return 0;
  {
    ProfScope ps(pool, "radix_scan_hist");
    radix_scan_hist_kernel<<<(unsigned) sizeof(K), RADIX, 0, s>>>(d_hist);
  }
  MR_CUDA(cudaGetLastError());

  uint64_t ntiles = sort_tiles(n, IPT);
  Buf<unsigned long long> status(pool, ntiles * RADIX);
  Buf<unsigned> counter(pool, 1);
  int cur = 0;
  for (int b = 0; b < (int) sizeof(K); b++) {
    if (!((need >> b) & 1)) continue;
    MR_CUDA(cudaMemsetAsync(status.p, 0, ntiles * RADIX * 8, s));
    MR_CUDA(cudaMemsetAsync(counter.p, 0, 4, s));
    const K *ki = cur ? k1 : k0;
    K *ko = cur ? k0 : k1;
    const P *pi = cur ? p1 : p0;
    P *po = cur ? p0 : p1;
    ProfScope ps(pool, "onesweep_pass", 2.0 * n * (sizeof(K) + (HASP ? sizeof(P) : 0)));
    onesweep_kernel<K, P, HASP, IPT><<<(unsigned) ntiles, SORT_THREADS, 0, s>>>(ki, ko, pi, po, n, 8 * b, d_hist + b * RADIX,
                                                                              status.p, counter.p);
    MR_CUDA(cudaGetLastError());
    cur ^= 1;
  }
  return cur;
}

}  // namespace mrgpu

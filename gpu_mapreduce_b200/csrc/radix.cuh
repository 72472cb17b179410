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

// lanes of the warp holding the same 8-bit digit.  Eight ballots + eight logic ops: measured far cheaper
// on sm_100a than one MATCH.ANY per key (ncu showed ~half of all stall samples waiting on MATCH results
// and a pass bound at ~10k cycles per 4096-key tile by the match unit).
// `sh` = 31 - (bit position of the digit's lowest bit): bit b of the digit is moved to the sign position
// and smeared, so the per-bit work is shift, shift, setp, vote, one 3-input logic op.
__device__ __forceinline__ unsigned match_digit(unsigned key_lo32, int sh)
{
  unsigned m = 0xFFFFFFFFu;
#pragma unroll
  for (int b = 0; b < RADIX_BITS; b++) {
    int sx = ((int) (key_lo32 << (sh - b))) >> 31;  // 0 or -1
    unsigned bal = __ballot_sync(0xFFFFFFFFu, sx < 0);
    m &= ~(bal ^ (unsigned) sx);                      // keep lanes whose bit equals mine
  }
  return m;
}

template <typename K, typename P, int IPT>
struct SortSmem {
  static constexpr int TILE = SORT_THREADS * IPT;
  static constexpr int XBYTES = (sizeof(K) > sizeof(P) ? sizeof(K) : sizeof(P)) * TILE;
};

// The tile body.  FULL = every slot of the tile holds a key (all tiles but the last): no bounds checks.
// Global positions fit 32 bits (callers cap n below 2^32), so scatter indices are u32 in shared memory and
// registers.  Instruction count per key matters here: at 148 SMs x 4 issue slots x 1.965 GHz against
// ~6.5 TB/s a pass that moves 16 B per key has a budget of ~90 thread-instructions per key before it turns
// issue-bound instead of HBM-bound.
template <typename K, typename P, bool HASP, int IPT, bool FULL>
__device__ __forceinline__ void onesweep_tile(const K *__restrict__ kin, K *__restrict__ kout, const P *__restrict__ pin, P *__restrict__ pout,
                                              uint64_t n, int shift, const unsigned long long *__restrict__ gexcl, unsigned long long *status,
                                              uint64_t tile, unsigned (*wc)[RADIX], unsigned char *xbuf, unsigned *gbase, unsigned *warp_tot)
{
  constexpr int TILE = SORT_THREADS * IPT;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t tbase = tile * TILE;
  const unsigned nvalid = FULL ? (unsigned) TILE : (unsigned) (n - tbase);
  const unsigned woff0 = warp * (32 * IPT) + lane;  // this lane's first slot in the tile

  // ---- load: warp-striped, coalesced; out-of-range slots hold the maximum key
  K key[IPT];
  const K *kp = kin + tbase + woff0;
#pragma unroll
  for (int j = 0; j < IPT; j++) key[j] = (FULL || woff0 + j * 32 < nvalid) ? kp[j * 32] : (K) ~(K) 0;

  // ---- rank inside the warp, digit by digit, in memory order (stable).
  // (1) all match masks, independent of each other; (2) each group's lowest lane bumps the warp's digit
  // counter (plain load + store; __syncwarp() orders successive rounds across lanes) and remembers the
  // old value = the group's base; (3) the base is broadcast from that lane.
  unsigned peers[IPT], wbase_r[IPT];
#pragma unroll
  for (int j = 0; j < IPT; j++) peers[j] = match_digit((unsigned) (key[j] >> (shift & 32)), 31 - (shift & 31));
  const unsigned lt_mask = lanemask_lt();
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    wbase_r[j] = 0;
    if ((peers[j] & lt_mask) == 0) {
      unsigned *c = &wc[warp][(unsigned) (key[j] >> shift) & 0xFF];
      unsigned b = *c;
      *c = b + __popc(peers[j]);
      wbase_r[j] = b;
    }
    __syncwarp();
  }
  unsigned rank[IPT];
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    unsigned b = __shfl_sync(0xFFFFFFFFu, wbase_r[j], __ffs(peers[j]) - 1);
    rank[j] = b + __popc(peers[j] & lt_mask);
  }
  __syncthreads();

  // ---- thread d: digit d's count in this tile, and each warp's exclusive offset for it
  unsigned cnt = 0, wex[SORT_WARPS];
#pragma unroll
  for (int w = 0; w < SORT_WARPS; w++) {
    wex[w] = cnt;
    cnt += wc[w][tid];
  }
  // publish for the tiles behind us as early as possible
  unsigned long long *st = status + tile * RADIX + tid;
  atomicExch(st, (tile == 0 ? LB_INC : LB_AGG) | cnt);

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
  // wc[w][d] becomes the tile slot where warp w's keys of digit d start
#pragma unroll
  for (int w = 0; w < SORT_WARPS; w++) wc[w][tid] = my_start + wex[w];

  // ---- decoupled look-back for digit d.  Predecessor states are read four at a time (independent loads in
  // flight) and consumed in order; an entry that is not published yet is polled again.
  unsigned long long excl = 0;
  if (tile > 0) {
    long long t = (long long) tile - 1;
    bool done = false;
    while (!done) {
      constexpr int W = 4;
      unsigned long long v[W];
#pragma unroll
      for (int q = 0; q < W; q++) {
        long long tt = t - q;
        v[q] = tt >= 0 ? *(const volatile unsigned long long *) (status + (uint64_t) tt * RADIX + tid) : LB_INC;
      }
#pragma unroll
      for (int q = 0; q < W; q++) {
        if (done) break;
        unsigned f = (unsigned) (v[q] >> 62);
        if (f == 0) break;  // not there yet: re-read from this entry
        excl += v[q] & LB_VAL;
        t--;
        if (f == 2) done = true;
      }
    }
    atomicExch(st, LB_INC | (excl + cnt));
  }
  gbase[tid] = (unsigned) (gexcl[tid] + excl) - my_start;
  __syncthreads();

  // ---- stage keys in tile-sorted order
  K *xk = reinterpret_cast<K *>(xbuf);
  unsigned (&pos)[IPT] = rank;  // same registers: rank -> slot in the tile
#pragma unroll
  for (int j = 0; j < IPT; j++) {
    pos[j] = wc[warp][(unsigned) (key[j] >> shift) & 0xFF] + rank[j];
    xk[pos[j]] = key[j];
  }
  __syncthreads();

  // ---- scatter: consecutive threads write consecutive slots of a digit's run
  unsigned gidx[IPT];
#pragma unroll
  for (int k = 0; k < IPT; k++) {
    unsigned s = k * SORT_THREADS + tid;
    if (FULL || s < nvalid) {
      K kk = xk[s];
      gidx[k] = gbase[(unsigned) (kk >> shift) & 0xFF] + s;
      kout[gidx[k]] = kk;
    }
  }
  if constexpr (HASP) {
    P pay[IPT];
    const P *pp = pin + tbase + woff0;
#pragma unroll
    for (int j = 0; j < IPT; j++)
      if (FULL || woff0 + j * 32 < nvalid) pay[j] = pp[j * 32];
    __syncthreads();
    P *xp = reinterpret_cast<P *>(xbuf);
#pragma unroll
    for (int j = 0; j < IPT; j++) xp[pos[j]] = pay[j];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < IPT; k++) {
      unsigned s = k * SORT_THREADS + tid;
      if (FULL || s < nvalid) pout[gidx[k]] = xp[s];
    }
  }
}

template <typename K, typename P, bool HASP, int IPT>
__global__ void __launch_bounds__(SORT_THREADS, (sizeof(K) == 4 ? 3 : 2))
onesweep_kernel(const K *__restrict__ kin, K *__restrict__ kout, const P *__restrict__ pin, P *__restrict__ pout, uint64_t n,
                int shift, const unsigned long long *__restrict__ gexcl /*[256] exclusive global digit offsets*/,
                unsigned long long *status /*[ntiles*256]*/, unsigned *tile_counter)
{
  constexpr int TILE = SORT_THREADS * IPT;
  __shared__ unsigned wc[SORT_WARPS][RADIX];
  __shared__ __align__(16) unsigned char xbuf[SortSmem<K, P, IPT>::XBYTES];
  __shared__ unsigned gbase[RADIX];
  __shared__ unsigned warp_tot[SORT_WARPS];
  __shared__ unsigned s_tile;
  const int tid = threadIdx.x;
  // tiles are claimed in launch order so that every tile a look-back waits on is already running
  if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
#pragma unroll
  for (int w = 0; w < SORT_WARPS; w++) wc[w][tid] = 0;
  __syncthreads();
  const uint64_t tile = s_tile;
  if ((tile + 1) * TILE <= n)
    onesweep_tile<K, P, HASP, IPT, true>(kin, kout, pin, pout, n, shift, gexcl, status, tile, wc, xbuf, gbase, warp_tot);
  else
    onesweep_tile<K, P, HASP, IPT, false>(kin, kout, pin, pout, n, shift, gexcl, status, tile, wc, xbuf, gbase, warp_tot);
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

// convert() fast path: dense small-integer keys whose pairs are indistinguishable.
//
// A stable sort groups pairs so that each key's values appear in arrival order.  When every pair carries the
// same value bytes (or none) two pairs with the same key cannot be told apart, so the KMV is fully determined
// by HOW MANY pairs each key has: grouping needs a count per key, not data movement.  If in addition the keys
// are small integers (uniform 4- or 8-byte keys whose value is below 2^26) the counts fit a direct-mapped
// table that stays in the 126 MB L2, and convert() becomes
//     one read of the KV (count)  ->  scan of the table  ->  one write of the KMV (fill)
// instead of extract + one radix pass per key byte + run detection + scatter.  IntCount (count of 2^30 Zipf keys
// over 2^24 values) and OINK's degree (vertex ids, NULL values) are exactly this case; anything else --
// a value that differs, a key outside the table -- raises a flag and the general sort path runs instead.
// Reference semantics are unchanged (KeyMultiValue::convert src/keymultivalue.cpp:486-638): one KMV pair per
// distinct key holding every value; keys come out in ascending order like the sort path.
//
// Skew: Zipf's hottest key is 11.5 % of all pairs, i.e. it is in practically every warp.  Plain global atomics
// would serialise on its address, so every block keeps a shared-memory direct-mapped cache of (key, count)
// that absorbs the frequent keys (a slot is claimed by the first key that hashes to it, which is
// frequency-weighted) and only misses go to the table in L2.  Included by mrgpu.cu.
#pragma once

#include "dense_hash.cuh"  // DENSE_BITS / DENSE_CACHE / DENSE_MULT, dense_slot, dense_key_of (host-testable)
// Per-block cache: DENSE_CACHE 4-byte entries = 224 KB of shared memory, one 1024-thread block per SM.
// An entry is {count : 20 | tag : 12}.  Keys are < 2^26; h = key * odd mod 2^26 is a bijection, the slot is
// floor(h * DENSE_CACHE / 2^26), and the at most 1171 values of h that share a slot differ in their low 11
// bits, so tag = (h & 2047) + 1 identifies the key exactly (0 = empty) and the key can be recovered from
// (slot, tag) when the entry is flushed.  A hit is one shared load + one shared atomic add whose result is not
// needed (the count sits in the high bits: it cannot carry into the tag, and a block flushes its cache to the
// table in L2 every DENSE_ROUND < 2^20 keys, so it cannot wrap either).
// History (ncu + A/B runs on the IntCount workload): the L2 atomics of cache MISSES bound the first versions
// (64-entry cache 125 ms, 512: 50 ms, 8192: 7.7 ms, 16384: 5.4 ms, 32768 entries of tag + 15-bit count: 4.4 ms on
// 16-byte records, 2.95 ms on header-less keys); L2 atomics on one address are slow (forwarding the hottest
// key every 32 hits cost 20 ms).  That version issued at 1.6 IPC with every pipe under 60 %: latency-bound on
// the load -> compare -> atomic -> check chain of one key at a time, hence the batches of four below.
// A round = the keys a block takes between two sweeps of its cache.  MRGPU_DENSE_LAZY (default): a sweep sends to the
// table, and resets, only what it must -- entries it evicts and entries whose 20-bit count could overflow before the
// next sweep (count >= 2^19, rounds of at most 2^19 keys); all others keep counting and reach the table once, at the
// end.  With 0 every sweep flushes every entry (rounds of < 2^20 keys): on the IntCount stream one RED in four was
// such a flush (ncu: 2.46e8 RED sectors for 2^30 pairs; simulated, tools/cache_sim.c: 18.0 % misses + 4.9 % flushes
// against 16.2 % + 4.3 % with the lazy sweep).
#ifndef MRGPU_DENSE_LAZY
#define MRGPU_DENSE_LAZY 1
#endif
constexpr uint32_t DENSE_ROUND_LIMIT = MRGPU_DENSE_LAZY ? (1u << 19) : (1u << 20) - 1;
constexpr uint32_t DENSE_FLUSH_AT = 1u << 19;
#ifndef MRGPU_DENSE_KEEP
#define MRGPU_DENSE_KEEP 2
#endif
constexpr uint32_t DENSE_KEEP = MRGPU_DENSE_KEEP;  // count per round below which a cache entry is evicted at the flush

struct DenseArgs {
  const uint8_t *data;
  uint64_t n;
  uint32_t rs, koff, voff, kb, vb;
};

#ifndef MRGPU_DENSE_EXPERIMENT
#define MRGPU_DENSE_EXPERIMENT 0
#endif
#ifndef MRGPU_DENSE_THREADS
#define MRGPU_DENSE_THREADS 1024
#endif


// A/B variant, off by default (measured: no gain).  ncu shows the L1 data pipe 87 % busy in this kernel, 45 % of its
// wavefronts from shared atomics and two thirds of those counted as bank-conflict replays; with Zipf's hottest key in
// 3.7 of a warp's 32 lanes, a model in which lanes on the same word are served one per pass gives 6.8 passes per
// atomic instruction against 3.7 for distinct addresses alone.  MRGPU_DENSE_DEDUP=1 lets lanes that hold one of two
// keys the warp has lately seen several times per instruction (learned from the stream) elect a leader that adds for
// all of them; =2 fixes the two keys to the stream's hottest (0 and 1).  Fused pass at 2^30 keys: 2.511 ms without,
// 2.452 with the fixed keys, 2.591 with learning -- so lanes on the same word are evidently merged by the hardware
// and the replays are the ordinary bank conflicts of 32 random words (3.7 passes, twice as many wavefronts for an
// atomic as for a load).
#ifndef MRGPU_DENSE_DEDUP
#define MRGPU_DENSE_DEDUP 0
#endif
struct DenseHot {
  uint32_t k0 = 0xFFFFFFFFu, k1 = 0xFFFFFFFFu;  // tracked keys (no valid key is >= 2^26)
  uint32_t s0 = 0, s1 = 0;                      // lanes they were found in since the last review
  uint32_t tick = 0;
};

// four keys at a time: the four tag loads are independent, then the four updates
__device__ __forceinline__ void dense_count4(const uint32_t (&key)[4], unsigned valid_mask, unsigned *__restrict__ cache, unsigned *__restrict__ table,
                                             DenseHot &hot)
{
  uint32_t slot[4], tag[4], e[4];
#pragma unroll
  for (int q = 0; q < 4; q++) dense_slot(key[q], slot[q], tag[q]);
#if MRGPU_DENSE_DEDUP
  uint32_t wgt[4];
  {
    const unsigned act = __activemask(), lt = lanemask_lt();
    // the lanes taking part agree on the two keys by construction, but nothing else in the kernel depends on that:
    // take them from the first lane, so that a group can never mix lanes that compared against different keys
#if MRGPU_DENSE_DEDUP == 2  // timing experiment: the two hottest keys of the IntCount stream, no learning
    hot.k0 = 0, hot.k1 = 1;
#else
    hot.k0 = __shfl_sync(act, hot.k0, __ffs(act) - 1);
    hot.k1 = __shfl_sync(act, hot.k1, __ffs(act) - 1);
#endif
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const bool v = (valid_mask >> q) & 1u;
      const bool is0 = v && key[q] == hot.k0, is1 = v && key[q] == hot.k1;
      const unsigned m0 = __ballot_sync(act, is0), m1 = __ballot_sync(act, is1);
      hot.s0 += __popc(m0);
      hot.s1 += __popc(m1);
      const unsigned mine = is0 ? m0 : (is1 ? m1 : 0u);
      wgt[q] = mine ? ((mine & lt) ? 0u : (uint32_t) __popc(mine)) : 1u;
    }
    // review every 16 calls: the key of the lowest active lane challenges the weaker of the two
    if (MRGPU_DENSE_DEDUP != 2 && (++hot.tick & 15u) == 0) {
      const uint32_t x = __shfl_sync(act, key[0], __ffs(act) - 1);
      const uint32_t c = (uint32_t) __popc(__ballot_sync(act, (valid_mask & 1u) && key[0] == x));
      if (x != hot.k0 && x != hot.k1 && c >= 2) {  // seen c times in one instruction; s counts 64 instructions
        if (hot.s0 <= hot.s1) {
          if (c * 32u > hot.s0) hot.k0 = x;
        } else if (c * 32u > hot.s1) hot.k1 = x;
      }
      hot.s0 = hot.s1 = 0;
    }
  }
#endif
#if MRGPU_DENSE_EXPERIMENT == 4
  // A/B variant (measured: 2.74 ms against 2.61 for the fused pass, 2.262 against 2.261 for the count alone, so not
  // the default): one RETURNING shared atomic per key instead of a load plus an atomic.  The add is speculative: the
  // value it returns carries the slot's tag; a key that finds another key's tag takes its add back and goes to the
  // table, one that finds the slot empty tries to stamp its tag on exactly its own count.
#pragma unroll
  for (int q = 0; q < 4; q++) e[q] = (valid_mask >> q & 1) ? atomicAdd(&cache[slot[q]], 0x1000u) : 0u;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    if (!(valid_mask >> q & 1)) continue;
    const uint32_t t = e[q] & 0xFFFu;
    if (t == tag[q]) continue;  // hit: already counted
    if (t == 0 && atomicCAS(&cache[slot[q]], e[q] + 0x1000u, (e[q] + 0x1000u) | tag[q]) == e[q] + 0x1000u) continue;  // claimed with my 1
    atomicSub(&cache[slot[q]], 0x1000u);
    atomicAdd(&table[key[q]], 1u);
  }
  return;
#endif
#pragma unroll
  for (int q = 0; q < 4; q++) e[q] = cache[slot[q]];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    if (!(valid_mask >> q & 1)) continue;
    uint32_t t = e[q] & 0xFFFu;
    if (t == 0) {  // empty: claim it (first come; the hot keys come first more often than not)
      const uint32_t old = atomicCAS(&cache[slot[q]], 0u, tag[q]);
      t = old ? (old & 0xFFFu) : tag[q];
    }
#if MRGPU_DENSE_EXPERIMENT == 1   // timing experiment: misses are dropped
    if (t == tag[q]) atomicAdd(&cache[slot[q]], 0x1000u);
#elif MRGPU_DENSE_EXPERIMENT == 2  // timing experiment: hits are dropped
    if (t != tag[q]) atomicAdd(&table[key[q]], 1u);
#elif MRGPU_DENSE_EXPERIMENT == 3  // timing experiment: only the tag loads
    if (t == 0xFFFu) atomicAdd(&table[key[q]], 1u);
#elif MRGPU_DENSE_DEDUP
    if (wgt[q]) {  // 0: another lane of the warp adds for this pair
      if (t == tag[q]) atomicAdd(&cache[slot[q]], wgt[q] << 12);
      else atomicAdd(&table[key[q]], wgt[q]);
    }
#else
    if (t == tag[q]) atomicAdd(&cache[slot[q]], 0x1000u);
    else atomicAdd(&table[key[q]], 1u);
#endif
  }
}

// cache -> table; counts back to zero, tags of keys that earned their slot stay
__device__ __forceinline__ void dense_flush(unsigned *__restrict__ cache, unsigned *__restrict__ table, bool final)
{
  for (uint32_t i = threadIdx.x; i < DENSE_CACHE; i += blockDim.x) {
    const uint32_t e = cache[i], c = e >> 12;
    if (c && (!MRGPU_DENSE_LAZY || final || c < DENSE_KEEP || c >= DENSE_FLUSH_AT)) {
      atomicAdd(&table[dense_key_of(i, e & 0xFFFu)], c);
      // a slot whose key came up only once in a whole round is given back: the next key to arrive there is more
      // likely a frequent one (simulated on the IntCount stream: 25.6 % -> 22.7 % of the pairs end as REDs)
      cache[i] = c >= DENSE_KEEP ? (e & 0xFFFu) : 0u;
    }
  }
}

// MODE 0: header-less rows that are just a 4-byte key (a 16-byte load brings four pairs)
// MODE 3: the same with 32-byte loads/stores marked evict-first in L2 (needs 32-byte aligned pointers)
// MODE 1: 16-byte records {kb, vb, w2, w3} (one load per pair)      MODE 2: any other word-aligned layout
template <int MODE>
__global__ void __launch_bounds__(MRGPU_DENSE_THREADS)
dense_count_kernel(DenseArgs a, unsigned *__restrict__ table, unsigned long long *abort_flag, unsigned long long *key_bits, int4 *__restrict__ copy_out)
{
  extern __shared__ unsigned cache[];
  uint32_t kor = 0;  // OR of this thread's keys: bounds the part of the table that has to be scanned afterwards
  for (uint32_t i = threadIdx.x; i < DENSE_CACHE; i += blockDim.x) cache[i] = 0;
  // value of record 0 (compared word-wise; vb is 0, 4 or 8 here)
  uint32_t v0a = 0, v0b = 0;
  if (a.vb >= 4) v0a = *(const uint32_t *) (a.data + a.voff);
  if (a.vb == 8) v0b = *(const uint32_t *) (a.data + a.voff + 4);
  bool bad = false;
  DenseHot hot;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  const uint64_t tid0 = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  // keys a block takes per loop iteration; a round ends before any cache count can reach 2^20
  constexpr uint32_t KPI = (MODE == 0 || MODE == 3) ? 16 : (MODE == 1 ? 8 : 4);
  const uint32_t iters_per_round = DENSE_ROUND_LIMIT / (KPI * blockDim.x);
  const uint64_t units = MODE == 0 ? a.n / 4 : (MODE == 3 ? a.n / 8 : a.n);  // loop units: 16- / 32-byte vectors or pairs
  constexpr int U = MODE == 0 ? 4 : (MODE == 3 ? 2 : (MODE == 1 ? 8 : 4));   // units in flight per thread
  const uint64_t iters = (units + stride * U - 1) / (stride * U);   // same for every thread of the grid
  __syncthreads();
  for (uint64_t it0 = 0; it0 < iters; it0 += iters_per_round) {
    const uint64_t it1 = min(it0 + iters_per_round, iters);
    for (uint64_t it = it0; it < it1 && (!bad || copy_out); it++) {  // a copy goes on after the count is given up
      const uint64_t base = it * stride * U + tid0;
      if (MODE == 3) {
        uint32_t r[U][8];
#pragma unroll
        for (int u = 0; u < U; u++) {
          const uint64_t i = base + (uint64_t) u * stride;
          if (i < units) ld_stream256(a.data + i * 32, r[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
          const uint64_t i = base + (uint64_t) u * stride;
          if (i < units) {
            if (copy_out) st_stream256((uint8_t *) copy_out + i * 32, r[u]);  // emit + count in one pass (map_intcount)
            const uint32_t any = r[u][0] | r[u][1] | r[u][2] | r[u][3] | r[u][4] | r[u][5] | r[u][6] | r[u][7];
            kor |= any;
            if (any >= DENSE_SLOTS) bad = true;
            else if (!bad) {
              const uint32_t k0[4] = {r[u][0], r[u][1], r[u][2], r[u][3]}, k1[4] = {r[u][4], r[u][5], r[u][6], r[u][7]};
              dense_count4(k0, 0xFu, cache, table, hot);
              dense_count4(k1, 0xFu, cache, table, hot);
            }
          }
        }
      } else if (MODE == 0) {
        int4 r[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
          const uint64_t i = base + (uint64_t) u * stride;
          if (i < units) r[u] = ld_stream((const int4 *) a.data + i);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
          const uint64_t i = base + (uint64_t) u * stride;
          if (i < units) {
            if (copy_out) st_stream(copy_out + i, r[u]);  // emit + count in one pass (map_intcount)
            const uint32_t k[4] = {(uint32_t) r[u].x, (uint32_t) r[u].y, (uint32_t) r[u].z, (uint32_t) r[u].w};
            kor |= k[0] | k[1] | k[2] | k[3];
            if ((k[0] | k[1] | k[2] | k[3]) >= DENSE_SLOTS) bad = true;
            else if (!bad) dense_count4(k, 0xFu, cache, table, hot);
          }
        }
      } else if (MODE == 1) {
        int4 r[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
          const uint64_t i = base + (uint64_t) u * stride;
          if (i < units) r[u] = ld_stream((const int4 *) a.data + i);
        }
#pragma unroll
        for (int g = 0; g < U; g += 4) {
          uint32_t k[4];
          unsigned vm = 0;
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const uint64_t i = base + (uint64_t) (g + q) * stride;
            k[q] = 0;
            if (i < units) {
              k[q] = (uint32_t) r[g + q].z;
              const uint32_t w3 = (uint32_t) r[g + q].w;
              // kb 8: w3 is the key's high word (must be 0); kb 4, vb 4: w3 is the value
              if (k[q] >= DENSE_SLOTS || (a.kb == 8 ? w3 != 0 : (a.vb == 4 && w3 != v0a))) bad = true;
              else vm |= 1u << q;
              kor |= k[q];
            }
          }
          if (!bad) dense_count4(k, vm, cache, table, hot);
        }
      } else {
        uint32_t k[4];
        unsigned vm = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const uint64_t i = base + (uint64_t) q * stride;
          k[q] = 0;
          if (i < units) {
            const uint8_t *r = a.data + i * a.rs;
            k[q] = *(const uint32_t *) (r + a.koff);
            uint32_t khi = 0, va = v0a, vb2 = v0b;
            if (a.kb == 8) khi = *(const uint32_t *) (r + a.koff + 4);
            if (a.vb >= 4) va = *(const uint32_t *) (r + a.voff);
            if (a.vb == 8) vb2 = *(const uint32_t *) (r + a.voff + 4);
            if (khi != 0 || k[q] >= DENSE_SLOTS || va != v0a || vb2 != v0b) bad = true;
            else vm |= 1u << q;
            kor |= k[q];
          }
        }
        if (!bad) dense_count4(k, vm, cache, table, hot);
      }
    }
    if (__syncthreads_or(bad)) {
      if (threadIdx.x == 0) *abort_flag = 1;
      if (!copy_out) return;  // the table is abandoned; the caller falls back to the sort path
      bad = true;
      continue;
    }
    dense_flush(cache, table, it1 >= iters);
    __syncthreads();
  }
  if (MODE == 0 || MODE == 3) {  // the few keys after the last full vector
    const uint64_t i = units * (MODE == 0 ? 4 : 8) + tid0;
    if (i < a.n) {
      const uint32_t k = ((const uint32_t *) a.data)[i];
      if (copy_out) ((uint32_t *) copy_out)[i] = k;
      kor |= k;
      if (k >= DENSE_SLOTS) *abort_flag = 1;
      else atomicAdd(&table[k], 1u);
    }
  }
  kor = __reduce_or_sync(0xFFFFFFFFu, kor);
  if ((threadIdx.x & 31) == 0 && kor) atomicOr(key_bits, (unsigned long long) kor);
}

// One thread per 4-byte word of the KMV stream: coalesced stores, no reads of pair data at all.
// A block owns a contiguous chunk of the stream, finds the KMV pairs overlapping it (offs[] is sorted) and
// every thread locates its word inside its pair by binary search in shared memory.
constexpr int FILL_THREADS = 256;
constexpr int FILL_WORDS_PER_THREAD = 16;
constexpr int FILL_CHUNK_WORDS = FILL_THREADS * FILL_WORDS_PER_THREAD;  // 16 KB of output per block
constexpr int FILL_MAX_PAIRS = FILL_CHUNK_WORDS / 4 + 2;               // a KMV pair is at least 16 bytes

__global__ void __launch_bounds__(FILL_THREADS)
dense_fill_kernel(uint32_t *__restrict__ out, uint64_t total_words, const uint64_t *__restrict__ offs, const uint32_t *__restrict__ ukey,
                  uint32_t kw, const uint32_t *__restrict__ ucnt, uint64_t U, Align al, uint32_t kb, uint32_t vb, uint32_t cv0, uint32_t cv1,
                  uint64_t chunks_per_block)
{
  __shared__ uint64_t s_off[FILL_MAX_PAIRS + 1];
  __shared__ uint32_t s_cnt[FILL_MAX_PAIRS];
  __shared__ uint64_t s_u0;
  __shared__ int s_np;
  const uint64_t nchunks = (total_words + FILL_CHUNK_WORDS - 1) / FILL_CHUNK_WORDS;
  const uint64_t c_begin = (uint64_t) blockIdx.x * chunks_per_block;
  const uint64_t c_end = min(c_begin + chunks_per_block, nchunks);
  if (c_begin >= c_end) return;
  // a block walks consecutive chunks, so only its first chunk needs a search in global memory for the pair
  // that contains the chunk start; afterwards the position carries over
  if (threadIdx.x == 0) {
    const uint64_t b0 = c_begin * FILL_CHUNK_WORDS * 4;
    uint64_t lo = 0, hi = U;  // first pair whose end lies beyond b0
    while (lo < hi) {
      uint64_t mid = (lo + hi) >> 1;
      if (offs[mid + 1] > b0) hi = mid;
      else lo = mid + 1;
    }
    s_u0 = lo;
  }
  for (uint64_t c = c_begin; c < c_end; c++) {
    __syncthreads();
    const uint64_t w0 = c * FILL_CHUNK_WORDS;
    const uint64_t b1 = min((w0 + FILL_CHUNK_WORDS) * 4, total_words * 4);
    const uint64_t u0 = s_u0;
    // load the pairs that start before the chunk end, a batch of FILL_THREADS at a time (usually one batch)
    int np_acc = 0;
    for (int base = 0; base <= FILL_MAX_PAIRS; base += FILL_THREADS) {
      const int q = base + threadIdx.x;
      const uint64_t u = u0 + q;
      const uint64_t o = (q <= FILL_MAX_PAIRS && u <= U) ? offs[u] : ~0ull;
      if (q <= FILL_MAX_PAIRS) {
        s_off[q] = o;
        if (q < FILL_MAX_PAIRS && u < U) s_cnt[q] = ucnt[u];
      }
      const int c = __syncthreads_count(o < b1);
      np_acc += c;
      if (c < FILL_THREADS) break;  // block-uniform: the first pair starting at or after b1 is loaded too
    }
    if (threadIdx.x == 0) s_np = np_acc;
    __syncthreads();
    const int np = s_np;
    // each thread produces 16 consecutive words (four 16-byte stores): one binary search in shared memory
    // for its first word, then it walks forward through the pairs
    const uint64_t wfirst = w0 + (uint64_t) threadIdx.x * FILL_WORDS_PER_THREAD;
    if (wfirst < total_words) {
      int lo = 0, hi = np - 1;
      {
        const uint64_t b = wfirst * 4;
        while (lo < hi) {
          int mid = (lo + hi + 1) >> 1;
          if (s_off[mid] <= b) lo = mid;
          else hi = mid - 1;
        }
      }
      uint64_t pbase = s_off[lo], pnext = s_off[lo + 1];
      uint64_t nv = s_cnt[lo], ko = kmv_key_off(al, nv), vo = kmv_val_off(al, nv, kb);
      uint32_t wbuf[FILL_WORDS_PER_THREAD];
      // most of the stream is the inside of a long pair's size array or value array: 64 bytes of one constant
      const uint64_t sb = wfirst * 4, se = sb + 4 * FILL_WORDS_PER_THREAD, r0 = sb - pbase;
      const bool in_sizes = se <= pnext && r0 >= 12 && r0 + 4 * FILL_WORDS_PER_THREAD <= 12 + 4 * nv;
      const bool in_values = se <= pnext && r0 >= vo && r0 + 4 * FILL_WORDS_PER_THREAD <= vo + nv * vb;
      if (in_sizes || in_values) {
        const bool odd = in_values && vb == 8 && ((r0 - vo) & 4);
        const uint32_t e = in_sizes ? vb : (odd ? cv1 : cv0), o = in_sizes ? vb : (vb == 8 ? (odd ? cv0 : cv1) : cv0);
#pragma unroll
        for (int k = 0; k < FILL_WORDS_PER_THREAD; k += 2) {
          wbuf[k] = e;
          wbuf[k + 1] = o;
        }
      } else
#pragma unroll
      for (int k = 0; k < FILL_WORDS_PER_THREAD; k++) {
        const uint64_t b = (wfirst + k) * 4;
        if (b >= pnext && lo + 1 < np) {  // next pair
          lo++;
          pbase = pnext;
          pnext = s_off[lo + 1];
          nv = s_cnt[lo];
          ko = kmv_key_off(al, nv);
          vo = kmv_val_off(al, nv, kb);
        }
        const uint64_t r = b - pbase;
        uint32_t word = 0;
        if (r >= 12 && r < 12 + 4 * nv) word = vb;
        else if (r >= vo && r < vo + nv * vb) word = (vb == 8 && ((r - vo) & 4)) ? cv1 : cv0;
        else if (r == 0) word = (uint32_t) nv;
        else if (r == 4) word = kb;
        else if (r == 8) word = (uint32_t) (nv * vb);
        else if (r >= ko && r < ko + kb) word = __ldg(ukey + (u0 + lo) * kw + (uint32_t) ((r - ko) >> 2));
        wbuf[k] = word;
      }
      if (wfirst + FILL_WORDS_PER_THREAD <= total_words && (((uintptr_t) (out + wfirst)) & 15) == 0) {
#pragma unroll
        for (int k = 0; k < FILL_WORDS_PER_THREAD; k += 4)
          st_stream((int4 *) (out + wfirst + k), make_int4((int) wbuf[k], (int) wbuf[k + 1], (int) wbuf[k + 2], (int) wbuf[k + 3]));
      } else {
#pragma unroll
        for (int k = 0; k < FILL_WORDS_PER_THREAD; k++)
          if (wfirst + k < total_words) out[wfirst + k] = wbuf[k];
      }
    }
    __syncthreads();
    // the pair that contains the next chunk's first byte: the last one seen if it runs past b1, else the next
    if (threadIdx.x == 0) s_u0 = u0 + (uint64_t) np - (s_off[np] > b1 ? 1 : 0);
  }
}

// factored KMV -> the reference's record stream (kmv->data, kmv->offs); no-op for any other KMV
static void kmv_expand_fvar(mrgpu_mr *mr, DevKMV *kmv);  // ops_convert_hash.cuh

static void kmv_expand(mrgpu_mr *mr, DevKMV *kmv)
{
  if (kmv && kmv->fvar) {
    kmv_expand_fvar(mr, kmv);
    return;
  }
  if (!kmv || !kmv->factored) return;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t U = kmv->nkmv;
  const uint32_t kb = (uint32_t) kmv->ukb, vb = (uint32_t) kmv->uvb;
  kmv->offs = (uint64_t *) pool.alloc((U + 1) * 8);
  kmv->offs_cap = U + 1;
  uint64_t *offs = kmv->offs;
  const uint32_t *pc = kmv->fcnt;
  scan_apply(
      pool, U, [=] __device__(uint64_t u) -> unsigned long long {
        uint64_t nv = pc[u];
        return kmv_rec_size(al, nv, kb, nv * vb);
      },
      [=] __device__(uint64_t u, unsigned long long e, unsigned long long) { offs[u] = e; }, mr->d_small, "convert_kmv_sizes", 12.0 * U);
  fetch_small(mr, 1);
  const uint64_t total = mr->h_small[0];
  MR_CUDA(cudaMemcpyAsync(offs + U, &total, 8, cudaMemcpyHostToDevice, s));
  kmv->data = (uint8_t *) pool.alloc(total);
  kmv->cap = total;
  kmv->nbytes = total;
  if (total) {
    ProfScope ps(pool, "convert_dense_fill", (double) total + 16.0 * U);
    const uint64_t words = total / 4;
    const uint64_t nchunks = (words + FILL_CHUNK_WORDS - 1) / FILL_CHUNK_WORDS;
    const uint64_t nblk = std::min<uint64_t>(nchunks, (uint64_t) kSMs * 16);
    const uint64_t cpb = (nchunks + nblk - 1) / nblk;
    dense_fill_kernel<<<(unsigned) ((nchunks + cpb - 1) / cpb), FILL_THREADS, 0, s>>>((uint32_t *) kmv->data, words, offs, kmv->fkey, kmv->fkw, kmv->fcnt, U, al, kb, vb,
                                                                                     kmv->fval[0], kmv->fval[1], cpb);
    MR_CUDA(cudaGetLastError());
  }
  sync(mr);
  pool.free(kmv->fkey, kmv->fcap * 4 * kmv->fkw);
  pool.free(kmv->fcnt, kmv->fcap * 4);
  kmv->fkey = kmv->fcnt = nullptr;
  kmv->fcap = 0;
  kmv->factored = false;
  kmv->h_offs_valid = false;
}

static void dense_prepare(mrgpu_mr *mr)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  // the table lives as long as the handle and is all zero between calls: the compaction in convert_dense puts
  // back a zero wherever it finds a count, so no 256 MB memset per convert
  if (!mr->dense_table) {
    mr->dense_table = (unsigned *) pool.alloc((uint64_t) DENSE_SLOTS * 4);
    MR_CUDA(cudaMemsetAsync(mr->dense_table, 0, (uint64_t) DENSE_SLOTS * 4, s));
  } else if (mr->dense_owner) {  // counts made ahead for a KV that never reached convert()
    MR_CUDA(cudaMemsetAsync(mr->dense_table, 0, (uint64_t) DENSE_SLOTS * 4, s));
  }
  // from here on the caller counts into the table: it is "dirty, owned by nobody" until the caller hands it to a KV
  // (emit_and_count) or the compaction in convert_dense has put the zeros back -- an error in between must not leave
  // stale counts behind a clean flag
  mr->dense_owner = ~0ull;
  static std::atomic<bool> attr_set[64];
  if (mr->device < 64 && !attr_set[mr->device].load()) {
    MR_CUDA(cudaFuncSetAttribute(dense_count_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, DENSE_CACHE * 4));
    MR_CUDA(cudaFuncSetAttribute(dense_count_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, DENSE_CACHE * 4));
    MR_CUDA(cudaFuncSetAttribute(dense_count_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, DENSE_CACHE * 4));
    MR_CUDA(cudaFuncSetAttribute(dense_count_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, DENSE_CACHE * 4));
    attr_set[mr->device].store(true);
  }
}

// Keep the busy part of the count table resident in L2 while a count kernel streams gigabytes past it: an access
// policy window on the launching stream (persisting for the table, streaming for everything else) over an L2
// set-aside of the same size.  Measured on IntCount (fused emit + count, 2^30 keys): no window 3.27 ms, 32 MiB
// 3.01, 40 MiB 2.88, 48 MiB 2.72, 56 MiB 2.56, 60 MiB 2.57, 64 MiB 3.25, 96 MiB 3.59 -- the table's touched part is
// 64 MiB there, but a set-aside beyond half of the 126 MiB L2 starves the streams.
// This touches DEVICE-WIDE state (cudaLimitPersistingL2CacheSize) for the duration of one count launch, so it is a
// setting ("l2pin", MiB, default 56, 0 = leave the device alone; include/mrgpu.h) and it is scoped: the guard's
// destructor removes the stream's window and puts the set-aside back however the scope is left (error included).
// A host application that runs its own persisting lines (a non-zero limit found on entry) keeps them: the limit is
// only ever raised, the persisting lines are not reset, and the limit it had is restored.
struct L2Window {
  mrgpu_mr *mr;
  bool active = false;
  static std::mutex &mu()
  {
    static std::mutex m;
    return m;
  }
  static size_t *saved() { static size_t v[64] = {0}; return v; }
  static int *depth() { static int v[64] = {0}; return v; }
  explicit L2Window(mrgpu_mr *m) : mr(m)
  {
    if (mr->l2pin <= 0 || !mr->dense_table) return;
    const int dev = mr->device < 64 ? mr->device : 0;
    size_t bytes = (size_t) mr->l2pin << 20;
    int maxp = 0, maxw = 0;
    cudaDeviceGetAttribute(&maxp, cudaDevAttrMaxPersistingL2CacheSize, mr->device);
    cudaDeviceGetAttribute(&maxw, cudaDevAttrMaxAccessPolicyWindowSize, mr->device);
    bytes = std::min(bytes, std::min((size_t) maxp, (size_t) maxw));
    if (bytes == 0) {
      cudaGetLastError();
      return;
    }
    {
      // handles on other host threads may be inside their own window on the same device: the set-aside is saved
      // by the first one in and restored by the last one out
      std::lock_guard<std::mutex> lk(mu());
      if (depth()[dev]++ == 0) {
        cudaDeviceGetLimit(&saved()[dev], cudaLimitPersistingL2CacheSize);
        if (saved()[dev] < bytes) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, bytes);
      }
    }
    set_window(bytes);
    active = true;
    cudaGetLastError();  // a device without the feature: run without it
  }
  void set_window(size_t bytes)
  {
    cudaStreamAttrValue av;
    memset(&av, 0, sizeof(av));
    av.accessPolicyWindow.base_ptr = mr->dense_table;
    av.accessPolicyWindow.num_bytes = bytes;
    av.accessPolicyWindow.hitRatio = 1.0f;
    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    cudaStreamSetAttribute(mr->pool.stream, cudaStreamAttributeAccessPolicyWindow, &av);
  }
  L2Window(const L2Window &) = delete;
  L2Window &operator=(const L2Window &) = delete;
  ~L2Window()
  {
    if (!active) return;
    const int dev = mr->device < 64 ? mr->device : 0;
    set_window(0);
    cudaStreamSynchronize(mr->pool.stream);  // the kernel that used the window is done before the set-aside goes
    std::lock_guard<std::mutex> lk(mu());
    if (depth()[dev] > 0 && --depth()[dev] == 0) {
      if (saved()[dev] == 0) cudaCtxResetPersistingL2Cache();  // nobody else's persisting lines can exist
      cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, saved()[dev]);
    }
    cudaGetLastError();
  }
};

// map_intcount on one rank (ops_map.cuh): write the key column AND count it -- convert() then starts at the
// compaction.  If the KV changes before convert() (append, add, ...) the counts are dropped.
static void emit_and_count(mrgpu_mr *mr, DevKV *kv, const uint32_t *keys, uint64_t n)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  dense_prepare(mr);
  unsigned long long *flag = mr->d_small + 8, *key_bits = mr->d_small + 9;
  MR_CUDA(cudaMemsetAsync(flag, 0, 16, s));
  unsigned long long h[2] = {1, 0};
  {
    L2Window win(mr);
    // algorithmic bytes: map (input + KV out, 20 B/pair) + convert's read of the KV (16 B/pair); moved: 4 in, 4 out
    {
    ProfScope ps(pool, "map_intcount_emit_count", (double) n * 36, 1, (double) n * 8);
    DenseArgs a{(const uint8_t *) keys, n, 4, 0, 4, 4, 0};
    if (((((uintptr_t) keys) | ((uintptr_t) kv->data)) & 31) == 0 && !getenv("MRGPU_DENSE_NO256"))
      dense_count_kernel<3><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, (int4 *) kv->data);
    else
      dense_count_kernel<0><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, (int4 *) kv->data);
    MR_CUDA(cudaGetLastError());
    }
    MR_CUDA(cudaMemcpyAsync(h, flag, 16, cudaMemcpyDeviceToHost, s));
    sync(mr);
  }
  kv->counted = true;
  kv->counted_bad = h[0] != 0;
  kv->counted_bits = h[1];
  kv->ahead_n = n;
  mr->dense_owner = kv->uid;
}

// map_intcount with the keys in host memory: chunked DMA into the key column on a second stream; when `count`,
// each chunk is counted as soon as it has landed, so only the last chunk's count is not hidden behind PCIe.
static void emit_from_host(mrgpu_mr *mr, DevKV *kv, const uint32_t *hkeys, uint64_t n, bool count)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  uint32_t *dst = (uint32_t *) kv->data + kv->nkv;
  if (!mr->copy_stream) {
    MR_CUDA(cudaStreamCreateWithFlags(&mr->copy_stream, cudaStreamNonBlocking));
    for (auto &e : mr->copy_ev) MR_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  unsigned long long *flag = mr->d_small + 8, *key_bits = mr->d_small + 9;
  const bool prepared = count;
  if (count) {
    dense_prepare(mr);
    MR_CUDA(cudaMemsetAsync(flag, 0, 16, s));
  }
  // the copy stream starts after everything already queued on the handle's stream (allocation reuse)
  MR_CUDA(cudaEventRecord(mr->copy_ev[0], s));
  MR_CUDA(cudaStreamWaitEvent(mr->copy_stream, mr->copy_ev[0], 0));
  // 64 M keys per chunk: the cache flush at the end of every count launch is 148 x 57 344 REDs, small against
  // the 16 M misses a chunk of this workload has anyway
  const uint64_t chunk = 1ull << 26;
  std::vector<cudaEvent_t> landed;
  bool counted_any = false;
  for (uint64_t o = 0; o < n; o += chunk) {
    const uint64_t m = std::min(chunk, n - o);
    MR_CUDA(cudaMemcpyAsync(dst + o, hkeys + o, m * 4, cudaMemcpyHostToDevice, mr->copy_stream));
    cudaEvent_t e;
    MR_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    landed.push_back(e);
    MR_CUDA(cudaEventRecord(e, mr->copy_stream));
    MR_CUDA(cudaStreamWaitEvent(s, e, 0));
    if (count && (((uintptr_t) (dst + o)) & 15) == 0) {
      ProfScope ps(pool, "map_intcount_count_chunk", (double) m * 16, 1, (double) m * 4);
      DenseArgs a{(const uint8_t *) (dst + o), m, 4, 0, 4, 4, 0};
      dense_count_kernel<0><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
      MR_CUDA(cudaGetLastError());
      counted_any = true;
    } else
      count = false;  // an unaligned destination: convert() will count the whole KV itself
  }
  unsigned long long h[2] = {1, 0};
  if (count) MR_CUDA(cudaMemcpyAsync(h, flag, 16, cudaMemcpyDeviceToHost, s));
  sync(mr);
  for (auto e : landed) cudaEventDestroy(e);
  if (count) {
    kv->counted = true;
    kv->counted_bad = h[0] != 0;
    kv->counted_bits = h[1];
    kv->ahead_n = n;
    mr->dense_owner = kv->uid;
  } else if (counted_any) {
    // counting was abandoned half way: the table holds counts of the first chunks
    mr->dense_owner = ~0ull;  // "dirty, owned by nobody": dense_prepare wipes it before the next use
  } else if (prepared)
    mr->dense_owner = 0;  // prepared, but no chunk was ever counted: the table is still all zero
}

// Weighted rows [key | multiplicity] (store.cuh) into the table: one RED per row.  Rows arrive from every source rank
// with each key at most once per source, so no address is hit more than nranks times: no cache needed.
__global__ void __launch_bounds__(256)
dense_add_rows_kernel(const uint2 *__restrict__ rows, uint64_t nrows, unsigned *__restrict__ table, unsigned long long *abort_flag,
                      unsigned long long *key_bits)
{
  uint32_t kor = 0;
  bool bad = false;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += stride) {
    const uint2 r = __ldg(rows + i);
    kor |= r.x;
    if (r.x >= DENSE_SLOTS) bad = true;
    else atomicAdd(&table[r.x], r.y);
  }
  kor = __reduce_or_sync(0xFFFFFFFFu, kor);
  if ((threadIdx.x & 31) == 0 && kor) atomicOr(key_bits, (unsigned long long) kor);
  if (bad) *abort_flag = 1;
}

__global__ void __launch_bounds__(256) sum_multiplicities_kernel(const uint2 *__restrict__ rows, uint64_t nrows, unsigned long long *out)
{
  unsigned long long s = 0;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += stride) s += __ldg(&rows[i].y);
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

// Count the keys of a header-less KV of 4-byte keys that share one value into the handle's table (the first pass of
// convert's counting path, also what aggregate()'s combiner starts from).  Afterwards the counts belong to that KV
// (kv->counted ...) exactly as if map_intcount had made them while it wrote the keys.
static void dense_count_shared4(mrgpu_mr *mr, DevKV *kv)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const uint64_t n = kv->nkv;
  dense_prepare(mr);
  unsigned long long *flag = mr->d_small + 8, *key_bits = mr->d_small + 9;
  MR_CUDA(cudaMemsetAsync(flag, 0, 16, s));
  unsigned long long h[2] = {1, 0};
  {
    L2Window win(mr);
    {
      ProfScope ps(pool, "convert_dense_count", (double) kv->nbytes, 1, (double) kv->phys_bytes());
      DenseArgs a{kv->data, n, 4, 0, 4, 4, 0};
      if (((uintptr_t) kv->data % 32) == 0 && !getenv("MRGPU_DENSE_NO256"))
        dense_count_kernel<3><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
      else if (((uintptr_t) kv->data % 16) == 0)
        dense_count_kernel<0><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
      else
        dense_count_kernel<2><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
      MR_CUDA(cudaGetLastError());
    }
    MR_CUDA(cudaMemcpyAsync(h, flag, 16, cudaMemcpyDeviceToHost, s));
    sync(mr);
  }
  kv->counted = true;
  kv->counted_bad = h[0] != 0;
  kv->counted_bits = h[1];
  kv->ahead_n = n;
  mr->dense_owner = kv->uid;
}

// The table's non-empty slots as rows [key | count] (ascending keys); every slot read is put back to zero, so the
// table is clean afterwards.  key_bits = OR of all counted keys (bounds the scan).  Returns the number of rows.
static uint64_t dense_table_to_rows(mrgpu_mr *mr, uint64_t key_bits, Buf<uint2> &rows)
{
  Pool &pool = mr->pool;
  uint64_t span = 1;
  while (span <= key_bits) span <<= 1;
  mr->dense_owner = ~0ull;  // dirty until the sink below is enqueued
  unsigned *tb = mr->dense_table;
  auto nonzero = [=] __device__(uint64_t k) -> bool { return tb[k] != 0; };
  FlagTiles bins;
  flag_tiles(pool, span, nonzero, bins, mr->d_small, "aggregate_combine_bins", (double) span * 4);
  fetch_small(mr, 1);
  const uint64_t U = mr->h_small[0];
  rows.alloc(pool, U ? U : 1);
  uint2 *pr = rows.p;
  flag_sink(pool, bins, nonzero, [=] __device__(uint64_t k, unsigned long long e, unsigned long long h) {
    if (h) {
      pr[e] = make_uint2((uint32_t) k, tb[k]);
      tb[k] = 0;
    }
  }, "aggregate_combine_bins", (double) span * 4 + 8.0 * U);
  mr->dense_owner = 0;
  return U;
}

// returns true if the dense path produced mr->kmv; false = not applicable, nothing changed
static bool convert_dense(mrgpu_mr *mr)
{
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t n = kv->nkv;
  const uint32_t kb = kv->kb, vb = kv->vb;
  if (!kv->uniform || (kb != 4 && kb != 8) || (vb != 0 && vb != 4 && vb != 8)) return false;
  const bool weighted = kv->weighted;
  if ((al.k % 4) || (al.v % 4) || (n < (1u << 16) && !weighted)) return false;  // small inputs: the table scan would dominate
  // header-less rows (store.cuh): key at 0, value (if stored per pair) right after it
  const bool shared_val = kv->packed && kv->cval_on;
  const uint32_t rs = kv->packed ? kv->prow() : kv->recsize, koff = kv->packed ? 0 : kv_key_off(al), voff = kv->packed ? kb : kv_val_off(al, kb);
  const uint32_t svb = shared_val ? 0 : vb;  // value bytes the count kernel has to compare per pair

  // counts made by map_intcount while it wrote this very KV?
  const bool ahead = !weighted && kv->counted && kv->ahead_n == n && mr->dense_owner == kv->uid && shared_val && kb == 4;
  unsigned long long *flag = mr->d_small + 8, *key_bits = mr->d_small + 9;
  if (weighted) {  // rows [key | multiplicity] delivered by aggregate()'s combiner: add the multiplicities
    dense_prepare(mr);
    MR_CUDA(cudaMemsetAsync(flag, 0, 16, s));
    if (kv->nrows) {
      ProfScope ps(pool, "convert_dense_add_rows", (double) kv->nbytes, 1, (double) kv->phys_bytes());
      dense_add_rows_kernel<<<grid_for(kv->nrows, 256 * 4, 16), 256, 0, s>>>((const uint2 *) kv->data, kv->nrows, mr->dense_table, flag, key_bits);
      MR_CUDA(cudaGetLastError());
    }
  } else if (!ahead) {
    dense_prepare(mr);
    MR_CUDA(cudaMemsetAsync(flag, 0, 16, s));
    L2Window win(mr);
    ProfScope ps(pool, "convert_dense_count", (double) kv->nbytes, 1, (double) kv->phys_bytes());
    DenseArgs a{kv->data, n, rs, koff, voff, kb, svb};
    const bool a16 = ((uintptr_t) kv->data % 16) == 0;
    if (rs == 4 && koff == 0 && ((uintptr_t) kv->data % 32) == 0 && !getenv("MRGPU_DENSE_NO256"))
      dense_count_kernel<3><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
    else if (rs == 4 && koff == 0 && a16) dense_count_kernel<0><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
    else if (rs == 16 && koff == 8 && a16 && svb != 8) dense_count_kernel<1><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
    else dense_count_kernel<2><<<kSMs, MRGPU_DENSE_THREADS, DENSE_CACHE * 4, s>>>(a, mr->dense_table, flag, key_bits, nullptr);
    MR_CUDA(cudaGetLastError());
  }
  // The table holds counts from here until the compaction below has put the zeros back (or the abort path has wiped
  // it).  Anything that throws in between (pool.alloc out of memory at 2^30 keys, a CUDA error) leaves the handle
  // usable, so the table must read "dirty": the next dense_prepare() then wipes it instead of trusting it.
  mr->dense_owner = ~0ull;
  unsigned *table = mr->dense_table;
  unsigned long long h_flag[2] = {kv->counted_bad ? 1ull : 0ull, kv->counted_bits};
  uint32_t cv[2] = {0, 0};
  if (!ahead || weighted) MR_CUDA(cudaMemcpyAsync(h_flag, flag, 16, cudaMemcpyDeviceToHost, s));
  if (shared_val) cv[0] = kv->cval[0], cv[1] = kv->cval[1];
  else if (vb) MR_CUDA(cudaMemcpyAsync(cv, kv->data + voff, vb, cudaMemcpyDeviceToHost, s));
  sync(mr);
  if (h_flag[0]) {  // a key or value the table cannot take: clean up, the sort path runs
    MR_CUDA(cudaMemsetAsync(table, 0, (uint64_t) DENSE_SLOTS * 4, s));
    mr->dense_owner = 0;
    return false;
  }
  // every key is <= the OR of all keys: only that much of the table can hold counts
  uint64_t span = 1;
  while (span <= h_flag[1]) span <<= 1;

  // non-empty bins -> U, ukey[u], ucnt[u]
  unsigned *tb = table;
  auto nonzero = [=] __device__(uint64_t k) -> bool { return tb[k] != 0; };
  FlagTiles bins;
  flag_tiles(pool, span, nonzero, bins, mr->d_small, "convert_dense_bins", (double) span * 4);
  fetch_small(mr, 1);
  const uint64_t U = mr->h_small[0];
  DevKMV *kmv = new DevKMV(&pool);
  try {
    const uint32_t kw = kb / 4;
    kmv->fkw = kw;
    kmv->fkey = (uint32_t *) pool.alloc((U ? U : 1) * 4 * kw);
    kmv->fcnt = (uint32_t *) pool.alloc((U ? U : 1) * 4);
    kmv->fcap = U ? U : 1;
    uint32_t *pk = kmv->fkey, *pc = kmv->fcnt;
    flag_sink(pool, bins, nonzero, [=] __device__(uint64_t k, unsigned long long e, unsigned long long h) {
      if (h) {
        pk[e * kw] = (uint32_t) k;
        if (kw == 2) pk[e * 2 + 1] = 0;  // 8-byte keys on this path have a zero high word
        pc[e] = tb[k];
        tb[k] = 0;
      }
    }, "convert_dense_bins", (double) span * 4 + 12.0 * U);
    mr->dense_owner = 0;  // the sink that re-zeroes every counted slot is enqueued: clean again
    kmv->factored = true;
    kmv->ascending = true;  // table order = ascending keys
    kmv->fval[0] = cv[0];
    kmv->fval[1] = cv[1];
    kmv->nkmv = U;
    kmv->nvalue = n;
    kmv->ksize = U * (uint64_t) kb;
    kmv->vsize = kv->vsize;
    kmv->ukb = kb;
    kmv->uvb = vb;
    if (!mr->compact) kmv_expand(mr, kmv);
    sync(mr);
  } catch (...) {
    delete kmv;
    throw;
  }
  mr->kmv = kmv;
  return true;
}

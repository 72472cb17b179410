// convert(): KeyValue -> KeyMultiValue (group by key).
//
// Reference: KeyMultiValue::convert src/keymultivalue.cpp:486-638 = chained hash of Unique records
// (kv2unique :645-789), page layout of the KMV pairs (unique2kmv_all :800-966), then a second full pass
// re-hashing every pair to append its value (kv2kmv :1139-1209); overflow goes through Spool partitions.
// Semantics kept: one KMV pair per distinct key; its values in ARRIVAL order (the stable sort below
// preserves it -- reduce(cull) and the postings lists see the same value order as the reference).
// Semantics not kept (none of the reference's callers can observe it after the canonical sort the spec
// allows): KMV pairs come out in sort-key order instead of first-appearance order.
//
// GPU design: grouping = stable LSD radix sort on an exact, compact sort key, then run detection and a
// coalesced scatter into KMV records.  Chained tables and the two re-hash passes are random HBM traffic;
// the sort is streaming.
//   uniform KV, <= 8 varying key bytes: sort key = those bytes (bijective), payload = the value itself
//       (4/8-byte values), nothing (NULL values) or the record index (gathered afterwards);
//   everything else: sort key = 64-bit fingerprint (hashlittle(key,kb,0) : second hash), payload = record
//       index; runs are then VERIFIED key-by-key against their neighbour and the whole grouping is redone
//       with a new seed if two different keys ever share a fingerprint, so the result is exact.
// Included by mrgpu.cu.
#pragma once

// f(i) for i in [0, n)
template <typename F>
__global__ void __launch_bounds__(256) for_each_kernel(uint64_t n, F f)
{
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) f(i);
}

#include "compact_key.cuh"  // struct CompactKey; 16-byte key <-> sort key of its varying bytes (host-testable)

// ---- uniform records: extract sort key (+ payload) and histogram in one read of the KV -----------------

// KEY16: 16-byte keys at a 4-byte aligned offset (R-MAT's EDGE): the key is read as words (one 16-byte load when
// `vec16`: rows and key offset are multiples of 16) and the varying bytes are picked out of two 64-bit halves with
// shifts by amounts that are the same in every thread -- instead of one byte load per varying byte.
template <typename K, int PMODE /*0 none, 1 val32, 2 val64, 3 idx*/, bool IDENT /*key bytes 0..kb-1, kb<=8, word loads*/, bool KEY16 = false>
__global__ void __launch_bounds__(256)
extract_uniform_kernel(const uint8_t *__restrict__ data, uint64_t n, uint32_t rs, uint32_t koff, uint32_t voff, uint32_t kb, CompactKey ck,
                       K *__restrict__ keys, void *__restrict__ pay, unsigned long long *hist, unsigned long long *vdiff, bool vec16 = false,
                       unsigned *wmask = nullptr /* KEY16: OR of (key ^ first key), four words */)
{
  __shared__ unsigned sh[sizeof(K) * RADIX];
  for (int i = threadIdx.x; i < (int) sizeof(K) * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  // value of record 0: if every record carries the same value bytes the sort need not move them at all
  uint64_t v0 = 0;
  bool vd = false;
  if (PMODE == 1) v0 = *(const uint32_t *) (data + voff);
  if (PMODE == 2) v0 = (uint64_t) ((const uint32_t *) (data + voff))[0] | ((uint64_t) ((const uint32_t *) (data + voff))[1] << 32);
  unsigned prev[sizeof(K)], run[sizeof(K)];
#pragma unroll
  for (int b = 0; b < (int) sizeof(K); b++) prev[b] = 0, run[b] = 0;
  unsigned long long f_lo = 0, f_hi = 0, acc_lo = 0, acc_hi = 0;  // KEY16 with wmask: the first key, and what differs from it
  if (KEY16 && wmask) {
    const uint32_t *w = (const uint32_t *) (data + koff);
    f_lo = (unsigned long long) w[0] | ((unsigned long long) w[1] << 32);
    f_hi = (unsigned long long) w[2] | ((unsigned long long) w[3] << 32);
  }
  // A thread's successive records lie 256 apart inside one tile of 4096, not a whole grid apart: in input that is in
  // key order already (the bulk of an iterative job's later rounds) the high digits then repeat from one record of a
  // thread to the next and stay in its run counter; with a grid-wide stride every record ended a run, and the 32
  // lanes of a warp -- neighbouring records -- all hit the same histogram bin (4.2 ms instead of 2.8 ms at 2^28 keys).
  constexpr uint64_t XT = 256 * 16;
  for (uint64_t tile = blockIdx.x; tile * XT < n; tile += gridDim.x)
  for (uint64_t i = tile * XT + threadIdx.x; i < n && i < (tile + 1) * XT; i += 256) {
    const uint8_t *r = data + i * rs;
    K key = 0;
    if (IDENT) {
      const uint32_t *w = (const uint32_t *) (r + koff);
      unsigned long long k64 = 0;
      if (kb > 0) k64 = w[0];
      if (kb > 4) k64 |= (unsigned long long) w[1] << 32;
      if (kb & 3) {  // partial last word: mask bytes beyond the key
        unsigned long long m = kb >= 8 ? ~0ull : ((1ull << (8 * kb)) - 1);
        k64 &= m;
      }
      key = (K) k64;
    } else if (KEY16) {
      unsigned long long lo, hi;
      if (vec16) {
        const uint4 v = *(const uint4 *) (r + koff);
        lo = (unsigned long long) v.x | ((unsigned long long) v.y << 32);
        hi = (unsigned long long) v.z | ((unsigned long long) v.w << 32);
      } else {
        const uint32_t *w = (const uint32_t *) (r + koff);
        lo = (unsigned long long) w[0] | ((unsigned long long) w[1] << 32);
        hi = (unsigned long long) w[2] | ((unsigned long long) w[3] << 32);
      }
      acc_lo |= lo ^ f_lo;
      acc_hi |= hi ^ f_hi;
      key = (K) compact_of_key16(lo, hi, ck);
    } else {
      unsigned long long k64 = 0;
      for (int q = 0; q < ck.nvb; q++) k64 |= (unsigned long long) r[koff + ck.pos[q]] << (8 * q);
      key = (K) k64;
    }
    keys[i] = key;
    if (PMODE == 1) {
      uint32_t v = *(const uint32_t *) (r + voff);
      ((uint32_t *) pay)[i] = v;
      vd |= v != (uint32_t) v0;
    }
    if (PMODE == 2) {
      const uint32_t *v = (const uint32_t *) (r + voff);
      uint64_t vv = (uint64_t) v[0] | ((uint64_t) v[1] << 32);
      ((uint64_t *) pay)[i] = vv;
      vd |= vv != v0;
    }
    if (PMODE == 3) ((uint32_t *) pay)[i] = (uint32_t) i;
    hist_accumulate<K>(key, prev, run, sh);
  }
  hist_flush<K>(prev, run, sh, hist);
  if (__syncthreads_or(vd) && threadIdx.x == 0) *vdiff = 1;
  if (KEY16 && wmask) {
    const uint32_t a[4] = {(uint32_t) acc_lo, (uint32_t) (acc_lo >> 32), (uint32_t) acc_hi, (uint32_t) (acc_hi >> 32)};
#pragma unroll
    for (int w = 0; w < 4; w++) {
      const uint32_t r = __reduce_or_sync(0xFFFFFFFFu, a[w]);
      if ((threadIdx.x & 31) == 0 && r) atomicOr(&wmask[w], r);
    }
  }
}

// IntCount layout {kb=4, vb=4, key, value} in one 16-byte load per record
__global__ void __launch_bounds__(256)
extract16_kernel(const int4 *__restrict__ recs, uint64_t n, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals, unsigned long long *hist,
                 unsigned long long *vdiff)
{
  __shared__ unsigned sh[4 * RADIX];
  for (int i = threadIdx.x; i < 4 * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t v0 = (uint32_t) __ldg((const int *) recs + 3);
  bool vd = false;
  unsigned prev[4] = {0, 0, 0, 0}, run[4] = {0, 0, 0, 0};
  constexpr int U = 4;
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x * U;
  for (uint64_t base = (uint64_t) blockIdx.x * blockDim.x * U + threadIdx.x; base < n; base += stride) {
    int4 r[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      uint64_t i = base + (uint64_t) u * blockDim.x;
      if (i < n) r[u] = ld_stream(recs + i);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      uint64_t i = base + (uint64_t) u * blockDim.x;
      if (i < n) {
        keys[i] = (uint32_t) r[u].z;
        vals[i] = (uint32_t) r[u].w;
        vd |= (uint32_t) r[u].w != v0;
        hist_accumulate<uint32_t>((uint32_t) r[u].z, prev, run, sh);
      }
    }
  }
  hist_flush<uint32_t>(prev, run, sh, hist);
  if (__syncthreads_or(vd) && threadIdx.x == 0) *vdiff = 1;
}

// OR of (key XOR first key) over all records: which key bytes vary at all (uniform keys longer than 8 bytes)
__global__ void key_diff_kernel(const uint8_t *__restrict__ data, uint64_t n, uint32_t rs, uint32_t koff, uint32_t kb, unsigned *mask /*[kb]*/)
{
  extern __shared__ unsigned shm[];
  for (uint32_t q = threadIdx.x; q < kb; q += blockDim.x) shm[q] = 0;
  __syncthreads();
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint8_t *r = data + i * rs + koff;
    for (uint32_t q = 0; q < kb; q++) {
      unsigned d = r[q] ^ data[koff + q];
      if (d && !shm[q]) shm[q] = 1;  // benign race: any writer stores 1
    }
  }
  __syncthreads();
  for (uint32_t q = threadIdx.x; q < kb; q += blockDim.x)
    if (shm[q]) mask[q] = 1;
}

// the same for word-aligned keys of at most 8 words: OR of (word XOR first key's word), reduced per warp
__global__ void __launch_bounds__(256)
key_diff_words_kernel(const uint32_t *__restrict__ data, uint64_t n, uint32_t rw, uint32_t kow, uint32_t kw, unsigned *wmask /*[kw]*/)
{
  uint32_t first[8], acc[8];
#pragma unroll
  for (int w = 0; w < 8; w++) first[w] = w < (int) kw ? data[kow + w] : 0, acc[w] = 0;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t *r = data + i * rw + kow;
#pragma unroll
    for (int w = 0; w < 8; w++)
      if (w < (int) kw) acc[w] |= r[w] ^ first[w];
  }
#pragma unroll
  for (int w = 0; w < 8; w++) {
    const uint32_t a = __reduce_or_sync(0xFFFFFFFFu, acc[w]);
    if ((threadIdx.x & 31) == 0 && a && w < (int) kw) atomicOr(&wmask[w], a);
  }
}

// ---- general records: fingerprint + index ------------------------------------------------------------------

__global__ void __launch_bounds__(256)
extract_fp_kernel(const uint8_t *__restrict__ data, const uint64_t *__restrict__ offs, uint64_t recsize, uint64_t n, Align al, uint32_t seed,
                  uint64_t *__restrict__ keys, uint32_t *__restrict__ idx, unsigned long long *hist)
{
  __shared__ unsigned sh[8 * RADIX];
  for (int i = threadIdx.x; i < 8 * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  unsigned prev[8], run[8];
#pragma unroll
  for (int b = 0; b < 8; b++) prev[b] = 0, run[b] = 0;
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint8_t *r = data + rec_off(offs, recsize, i);
    uint32_t kb = *(const uint32_t *) r;
    const uint8_t *k = r + kv_key_off(al);
    uint64_t fp = key_fingerprint(k, kb, seed);
    keys[i] = fp;
    idx[i] = (uint32_t) i;
    hist_accumulate<uint64_t>(fp, prev, run, sh);
  }
  hist_flush<uint64_t>(prev, run, sh, hist);
}

__device__ __forceinline__ bool keys_equal(const uint8_t *data, const uint64_t *offs, uint64_t recsize, Align al, uint32_t i, uint32_t j)
{
  const uint8_t *a = data + rec_off(offs, recsize, i), *b = data + rec_off(offs, recsize, j);
  uint32_t ka = *(const uint32_t *) a, kbb = *(const uint32_t *) b;
  if (ka != kbb) return false;
  a += kv_key_off(al);
  b += kv_key_off(al);
  uint32_t q = 0;
  if (((((uintptr_t) a) | ((uintptr_t) b)) & 3) == 0)  // word-aligned keys (any alignment setting that is a multiple of 4)
    for (; q + 4 <= ka; q += 4)
      if (*(const uint32_t *) (a + q) != *(const uint32_t *) (b + q)) return false;
  for (; q < ka; q++)
    if (a[q] != b[q]) return false;
  return true;
}

// warp-aggregated add of per-lane contributions to one global counter
__device__ __forceinline__ void warp_add(unsigned long long *ctr, unsigned long long v)
{
  unsigned act = __activemask();
  // sum over active lanes with shuffles restricted to the active set
  for (int o = 16; o; o >>= 1) {
    unsigned long long t = __shfl_down_sync(act, v, o);
    int src = (threadIdx.x & 31) + o;
    if (src < 32 && ((act >> src) & 1)) v += t;
  }
  if ((threadIdx.x & 31) == __ffs(act) - 1) atomicAdd(ctr, v);
}

// ------------------------------------------------------------------------------------------------------------

// Keys that are mostly in order already (merge.cuh): if one ascending run covers at least half of them, sort the
// others and merge.  Returns the index (0/1) of the array that holds the sorted keys, or -1 when the input is not of
// that kind (nothing was changed then).  The setting "presorted" (default 1) switches the test off.
template <typename K>
static int sort_keys_presorted(mrgpu_mr *mr, K *k0, K *k1, uint64_t n, unsigned mask)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  constexpr uint64_t WIN = 4096, NBLK = 64;
  if (!mr->presorted || n < (1u << 18) || n >= (1ull << 32)) return -1;
  MR_CUDA(cudaMemsetAsync(mr->d_small, 0, 16, s));
  {
    ProfScope ps(pool, "presorted_sample", (double) WIN * NBLK * sizeof(K));
    descent_sample_kernel<K><<<NBLK, 256, 0, s>>>(k0, n, WIN, mr->d_small);
    MR_CUDA(cudaGetLastError());
  }
  fetch_small(mr, 1);
  if (mr->h_small[0] > WIN * NBLK / 16) return -1;

  const unsigned long long init[3] = {0ull, 0ull, (unsigned long long) n};
  {
    MR_CUDA(cudaMemcpyAsync(mr->d_small, init, sizeof(init), cudaMemcpyHostToDevice, s));
    ProfScope ps(pool, "presorted_scan", (double) n * sizeof(K));
    descent_scan_kernel<K><<<grid_for(n, 256 * 16, 16), 256, 0, s>>>(k0, n, n / 2, mr->d_small);
    MR_CUDA(cudaGetLastError());
  }
  fetch_small(mr, 3);
  const uint64_t D = mr->h_small[0];
  if (D == 0) return 0;
  if (D > n / 16) return -1;
  // the run that holds the middle element: [a, b)
  const uint64_t a = mr->h_small[1], b = mr->h_small[2], L = b - a, R = n - L;
  if (L < n / 2) return -1;
  Buf<K> r0(pool, R), r1(pool, R);
  if (a) MR_CUDA(cudaMemcpyAsync(r0.p, k0, a * sizeof(K), cudaMemcpyDeviceToDevice, s));
  if (n > b) MR_CUDA(cudaMemcpyAsync(r0.p + a, k0 + b, (n - b) * sizeof(K), cudaMemcpyDeviceToDevice, s));
  Buf<unsigned long long> h2(pool, 8 * RADIX);
  const int c = radix_sort_pairs<K, uint32_t, false>(pool, r0.p, r1.p, (uint32_t *) nullptr, (uint32_t *) nullptr, R, mask, false, h2.p);
  merge_sorted<K>(pool, k0 + a, L, c ? r1.p : r0.p, R, k1);
  return 1;
}

// verify: `ck` is a guess (the bytes that varied last time).  The extract pass then also ORs every key against the
// first one; if a byte outside the guess varies, false is returned and nothing has changed.  key0 is filled in here.
template <typename K, int PMODE>
static bool convert_uniform(mrgpu_mr *mr, const CompactKey &ck, bool ident, std::vector<uint8_t> &key0, bool verify = false)
{
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t n = kv->nkv;
  const RowLayout lay = kv_layout(mr, kv);  // records, or header-less rows [key | value]
  const uint32_t kb = kv->kb, vb = kv->vb, rs = lay.rs, koff = lay.koff, voff = lay.voff;
  using P = typename std::conditional<PMODE == 2, uint64_t, uint32_t>::type;
  constexpr bool HASP = PMODE != 0;

  Buf<K> k0(pool, n), k1(pool, n);
  Buf<P> p0(pool, HASP ? n : 1), p1(pool, HASP ? n : 1);
  Buf<unsigned long long> hist(pool, 8 * RADIX + 8);
  MR_CUDA(cudaMemsetAsync(hist.p, 0, (8 * RADIX + 8) * 8, s));
  unsigned long long *vdiff = hist.p + 8 * RADIX;
  Buf<unsigned> wmask(pool, 4);
  if (verify) MR_CUDA(cudaMemsetAsync(wmask.p, 0, 16, s));

  // 1. one read of the KV: sort keys, payload, byte histograms
  {
  ProfScope ps(pool, "convert_extract", (double) n * (8 + kb + vb + sizeof(K) + (HASP ? sizeof(P) : 0)));
  if (ident && PMODE == 1 && sizeof(K) == 4 && kb == 4 && rs == 16 && koff == 8 && voff == 12 && ((uintptr_t) kv->data % 16) == 0)
    extract16_kernel<<<grid_for(n, 256 * 4, 8), 256, 0, s>>>((const int4 *) kv->data, n, (uint32_t *) k0.p, (uint32_t *) p0.p, hist.p, vdiff);
  else if (ident)
    extract_uniform_kernel<K, PMODE, true><<<grid_for(n, 256 * 8, 8), 256, 0, s>>>(kv->data, n, rs, koff, voff, kb, ck, k0.p, p0.p, hist.p, vdiff);
  else if (kb == 16 && rs % 4 == 0 && koff % 4 == 0 && ((uintptr_t) kv->data % 16) == 0)
    extract_uniform_kernel<K, PMODE, false, true><<<grid_for(n, 256 * 8, 8), 256, 0, s>>>(kv->data, n, rs, koff, voff, kb, ck, k0.p, p0.p, hist.p, vdiff,
                                                                                         rs % 16 == 0 && koff % 16 == 0, verify ? wmask.p : nullptr);
  else if (verify) MR_FAIL(MRGPU_EINVAL, "internal: convert on a guessed byte set needs the 16-byte-key extract");
  else
    extract_uniform_kernel<K, PMODE, false><<<grid_for(n, 256 * 8, 8), 256, 0, s>>>(kv->data, n, rs, koff, voff, kb, ck, k0.p, p0.p, hist.p, vdiff);
  MR_CUDA(cudaGetLastError());
  }

  if (verify) {
    unsigned hm[4];
    key0.resize(kb);
    MR_CUDA(cudaMemcpyAsync(hm, wmask.p, 16, cudaMemcpyDeviceToHost, s));
    MR_CUDA(cudaMemcpyAsync(key0.data(), kv->data + koff, kb, cudaMemcpyDeviceToHost, s));
    sync(mr);
    for (uint32_t q = 0; q < 16; q++)
      if ((hm[q / 4] >> (8 * (q % 4))) & 0xFFu) {
        bool in = false;
        for (int j = 0; j < ck.nvb; j++) in = in || ck.pos[j] == q;
        if (!in) return false;  // a byte the guess does not cover varies: the caller scans the keys and comes back
      }
  }

  // 2. stable LSD radix sort on the varying bytes only
  //    The same economy applies to values: when every pair carries identical value bytes (IntCount's int 1) the
  //    values are not moved through the sort; the KMV writer fills them in.
  unsigned mask = ck.nvb >= 8 ? 0xFF : ((1u << ck.nvb) - 1);
  bool constval = false;
  uint64_t cval = 0;
  if (PMODE == 1 || PMODE == 2) {
    unsigned long long flag = 1;
    MR_CUDA(cudaMemcpyAsync(&flag, vdiff, 8, cudaMemcpyDeviceToHost, s));
    MR_CUDA(cudaMemcpyAsync(&cval, kv->data + voff, vb, cudaMemcpyDeviceToHost, s));
    sync(mr);
    constval = flag == 0;
  }
  int cur;
  if (constval) {
    p0.release();
    p1.release();
    cur = sort_keys_presorted<K>(mr, k0.p, k1.p, n, mask);
    if (cur < 0) cur = radix_sort_pairs<K, P, false>(pool, k0.p, k1.p, (P *) nullptr, (P *) nullptr, n, mask, true, hist.p);
  } else {
    cur = PMODE == 0 ? sort_keys_presorted<K>(mr, k0.p, k1.p, n, mask) : -1;
    if (cur < 0) cur = radix_sort_pairs<K, P, HASP>(pool, k0.p, k1.p, p0.p, p1.p, n, mask, true, hist.p);
  }
  const K *sk = cur ? k1.p : k0.p;
  const P *sp = constval ? (const P *) nullptr : (cur ? p1.p : p0.p);
  (cur ? k0 : k1).release();
  if (!constval) (cur ? p0 : p1).release();

  // 3. run heads -> number of unique keys U and start[u]
  // read-only arrays go through the non-coherent path so that loads of later rounds may be hoisted above the
  // KMV stores of earlier rounds (the compiler must otherwise assume the stores alias them)
  auto head = [=] __device__(uint64_t i) -> bool { return i == 0 || __ldg(sk + i) != __ldg(sk + i - 1); };
  FlagTiles heads;
  flag_tiles(pool, n, head, heads, mr->d_small, "convert_heads_count", (double) n * sizeof(K));
  fetch_small(mr, 1);
  const uint64_t U = mr->h_small[0];
  Buf<uint32_t> start(pool, U + 1);
  {
    uint32_t *st = start.p;
    flag_sink(pool, heads, head, [=] __device__(uint64_t i, unsigned long long e, unsigned long long h) {
      if (h) st[e] = (uint32_t) i;
    }, "convert_heads_compact", (double) n * sizeof(K) + 4.0 * U);
    uint32_t n32 = (uint32_t) n;
    MR_CUDA(cudaMemcpyAsync(start.p + U, &n32, 4, cudaMemcpyHostToDevice, s));
  }

  // 4'. pairs that cannot be told apart (no value, or the same value everywhere): the KMV is (key, count) per
  //     distinct key -- the factored form (store.cuh); its record stream is written only if somebody reads it
  if (mr->compact && (PMODE == 0 || constval) && kb % 4 == 0 && kb > 0 && (vb == 0 || vb == 4 || vb == 8) && al.k % 4 == 0 && al.v % 4 == 0) {
    DevKMV *kmv = new DevKMV(&pool);
    try {
      const uint32_t kw = kb / 4;
      kmv->fkw = kw;
      kmv->fkey = (uint32_t *) pool.alloc((U ? U : 1) * 4 * kw);
      kmv->fcnt = (uint32_t *) pool.alloc((U ? U : 1) * 4);
      kmv->fcap = U ? U : 1;
      Buf<uint32_t> dkey0(pool, kw);
      if (!ident) MR_CUDA(cudaMemcpyAsync(dkey0.p, key0.data(), kb, cudaMemcpyHostToDevice, s));
      const uint8_t *k0b = (const uint8_t *) dkey0.p;
      uint8_t *fk = (uint8_t *) kmv->fkey;
      uint32_t *fc = kmv->fcnt;
      const uint32_t *st = start.p;
      {
        ProfScope ps(pool, "convert_factored", (double) U * (sizeof(K) + 8 + kb));
        // one thread per key: the first key's words with the varying bytes replaced, stored 16 bytes at a time
        const CompactKey ckc = ck;
        const uint32_t *k0w = (const uint32_t *) k0b;
        uint32_t *fkw32 = (uint32_t *) fk;
        auto word = [=] __device__(unsigned long long key, uint32_t w) -> uint32_t {
          if (ident) return (uint32_t) (key >> (32 * (w & 1)));
          uint32_t x = __ldg(k0w + w);
#pragma unroll
          for (int q = 0; q < 8; q++)
            if (q < ckc.nvb && (uint32_t) (ckc.pos[q] >> 2) == w) {
              const uint32_t sh = (ckc.pos[q] & 3) * 8;
              x = (x & ~(0xFFu << sh)) | ((uint32_t) ((key >> (8 * q)) & 0xFF) << sh);
            }
          return x;
        };
        if (kw == 4 && !ident) {
          // 16-byte keys: the first key with its varying bytes cleared, as two 64-bit halves; byte q of the sort key
          // is shifted into its place (the position is the same in every thread: no divergence, no table)
          unsigned long long h0[2] = {0, 0};
          memcpy(h0, key0.data(), 16);
          key16_clear_varying(h0[0], h0[1], ck);
          const unsigned long long lo0 = h0[0], hi0 = h0[1];
          auto mk = [=] __device__(uint64_t u) {
            const uint32_t s0 = __ldg(st + u);
            fc[u] = __ldg(st + u + 1) - s0;
            const unsigned long long key = (unsigned long long) __ldg(sk + s0);
            unsigned long long lo, hi;
            key16_of_compact(key, lo0, hi0, ckc, lo, hi);
            ((ulonglong2 *) fkw32)[u] = make_ulonglong2(lo, hi);
          };
          for_each_kernel<<<grid_for(U, 256, 64), 256, 0, s>>>(U, mk);
        } else if (kw % 4 == 0) {
          auto mk = [=] __device__(uint64_t u) {
            const uint32_t s0 = __ldg(st + u);
            fc[u] = __ldg(st + u + 1) - s0;
            const unsigned long long key = (unsigned long long) __ldg(sk + s0);
            for (uint32_t w = 0; w < kw; w += 4)
              ((uint4 *) fkw32)[(u * kw + w) >> 2] = make_uint4(word(key, w), word(key, w + 1), word(key, w + 2), word(key, w + 3));
          };
          for_each_kernel<<<grid_for(U, 256, 64), 256, 0, s>>>(U, mk);
        } else {
          auto mk = [=] __device__(uint64_t u) {
            const uint32_t s0 = __ldg(st + u);
            fc[u] = __ldg(st + u + 1) - s0;
            const unsigned long long key = (unsigned long long) __ldg(sk + s0);
            for (uint32_t w = 0; w < kw; w++) fkw32[u * kw + w] = word(key, w);
          };
          for_each_kernel<<<grid_for(U, 256, 64), 256, 0, s>>>(U, mk);
        }
        MR_CUDA(cudaGetLastError());
      }
      kmv->factored = true;
      kmv->fval[0] = (uint32_t) cval;
      kmv->fval[1] = (uint32_t) (cval >> 32);
      kmv->nkmv = U;
      kmv->nvalue = n;
      kmv->ksize = U * (uint64_t) kb;
      kmv->vsize = kv->vsize;
      kmv->ukb = kb;
      kmv->uvb = vb;
      sync(mr);
    } catch (...) {
      delete kmv;
      throw;
    }
    mr->kmv = kmv;
    return true;
  }

  // 4. KMV record sizes -> offs[u]
  DevKMV *kmv = new DevKMV(&pool);
  try {
    kmv->offs = (uint64_t *) pool.alloc((U + 1) * 8);
    kmv->offs_cap = U + 1;
    uint64_t *offs = kmv->offs;
    const uint32_t *st = start.p;
    scan_apply(
        pool, U, [=] __device__(uint64_t u) -> unsigned long long {
          uint64_t nv = st[u + 1] - st[u];
          return kmv_rec_size(al, nv, kb, nv * vb);
        },
        [=] __device__(uint64_t u, unsigned long long e, unsigned long long) { offs[u] = e; }, mr->d_small, "convert_kmv_sizes", 16.0 * U);
    fetch_small(mr, 1);
    const uint64_t total = mr->h_small[0];
    MR_CUDA(cudaMemcpyAsync(offs + U, &total, 8, cudaMemcpyHostToDevice, s));
    kmv->data = (uint8_t *) pool.alloc(total);
    kmv->cap = total;
    kmv->nbytes = total;

    // 5. scatter: element i of run u, rank r -> sizes[r], value r; the run head also writes header + key
    uint8_t *out = kmv->data;
    const uint8_t *kvdata = kv->data;
    // first key's bytes (constant bytes of every key) for the compact mode
    Buf<uint8_t> dkey0(pool, kb ? kb : 1);
    if (!ident && kb) MR_CUDA(cudaMemcpyAsync(dkey0.p, key0.data(), kb, cudaMemcpyHostToDevice, s));
    const uint8_t *k0b = dkey0.p;
    flag_sink(pool, heads, head, [=] __device__(uint64_t i, unsigned long long e, unsigned long long h) {
      uint64_t u = e + h - 1;
      uint32_t s0 = __ldg(st + u);
      uint64_t nv = __ldg(st + u + 1) - s0, r = i - s0;
      uint8_t *rec = out + __ldg((const unsigned long long *) offs + u);
      uint64_t ko = kmv_key_off(al, nv), vo = kmv_val_off(al, nv, kb);
      if (h) {
        uint32_t *hd = (uint32_t *) rec;
        hd[0] = (uint32_t) nv;
        hd[1] = kb;
        hd[2] = (uint32_t) (nv * vb);
        for (uint64_t q = 12 + 4 * nv; q < ko; q++) rec[q] = 0;
        K key = __ldg(sk + i);
        if (ident) {
          for (uint32_t q = 0; q < kb; q++) rec[ko + q] = (uint8_t) ((unsigned long long) key >> (8 * q));
        } else {
          for (uint32_t q = 0; q < kb; q++) rec[ko + q] = k0b[q];
          for (int q = 0; q < ck.nvb; q++) rec[ko + ck.pos[q]] = (uint8_t) ((unsigned long long) key >> (8 * q));
        }
        for (uint64_t q = ko + kb; q < vo; q++) rec[q] = 0;
        uint64_t end = kmv_rec_size(al, nv, kb, nv * vb);
        for (uint64_t q = vo + nv * vb; q < end; q++) rec[q] = 0;
      }
      ((uint32_t *) (rec + 12))[r] = vb;
      if (PMODE == 1) *(uint32_t *) (rec + vo + r * 4) = constval ? (uint32_t) cval : (uint32_t) __ldg(sp + i);
      if (PMODE == 2) {
        uint64_t v = constval ? cval : (uint64_t) __ldg(sp + i);
        uint32_t *d = (uint32_t *) (rec + vo + r * 8);
        d[0] = (uint32_t) v;
        d[1] = (uint32_t) (v >> 32);
      }
      if (PMODE == 3) {
        const uint8_t *src = kvdata + (uint64_t) __ldg(sp + i) * rs + voff;
        uint8_t *d = rec + vo + r * vb;
        for (uint32_t q = 0; q < vb; q++) d[q] = src[q];
      }
    }, "convert_kmv_write", (double) n * (sizeof(K) + ((HASP && !constval) ? sizeof(P) : 0)) + (double) total);
    kmv->nkmv = U;
    kmv->nvalue = n;
    kmv->ksize = U * (uint64_t) kb;
    kmv->vsize = kv->vsize;
    kmv->ukb = kb;
    kmv->uvb = vb;
    sync(mr);
  } catch (...) {
    delete kmv;
    throw;
  }
  mr->kmv = kmv;
  return true;
}

static void convert_general(mrgpu_mr *mr)
{
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t n = kv->nkv;
  const uint8_t *data = kv->data;
  const uint64_t *ko = kv->uniform ? nullptr : kv->offs;
  const uint64_t rs = kv->recsize;

  Buf<uint64_t> k0(pool, n), k1(pool, n);
  Buf<uint32_t> p0(pool, n), p1(pool, n);
  Buf<unsigned long long> hist(pool, 8 * RADIX);
  const uint64_t *sk = nullptr;
  const uint32_t *sp = nullptr;
  FlagTiles heads;
  uint64_t U = 0;
  for (uint32_t attempt = 0;; attempt++) {
    if (attempt == 8) MR_FAIL(MRGPU_EINVAL, "convert: fingerprint collisions persisted over 8 seeds");
    MR_CUDA(cudaMemsetAsync(hist.p, 0, 8 * RADIX * 8, s));
    // seed 0 makes the high half exactly the reference's bucket hash hashlittle(key,kb,0) (keymultivalue.cpp:1428)
    {
      ProfScope ps(pool, "convert_fp_extract", (double) kv->nbytes + 20.0 * n);
      extract_fp_kernel<<<grid_for(n, 256 * 4, 8), 256, 0, s>>>(data, ko, rs, n, al, attempt, k0.p, p0.p, hist.p);
      MR_CUDA(cudaGetLastError());
    }
    int cur = radix_sort_pairs<uint64_t, uint32_t, true>(pool, k0.p, k1.p, p0.p, p1.p, n, 0xFF, true, hist.p);
    sk = cur ? k1.p : k0.p;
    sp = cur ? p1.p : p0.p;
    const uint64_t *skc = sk;
    const uint32_t *spc = sp;
    unsigned long long *flag = mr->d_small + 1;
    MR_CUDA(cudaMemsetAsync(flag, 0, 8, s));
    // run heads by fingerprint; inside a run every element must equal its predecessor byte for byte
    auto head = [=] __device__(uint64_t i) -> bool {
      if (i == 0) return true;
      if (__ldg(skc + i) != __ldg(skc + i - 1)) return true;
      if (!keys_equal(data, ko, rs, al, __ldg(spc + i), __ldg(spc + i - 1))) *flag = 1;
      return false;
    };
    FlagTiles ht;
    flag_tiles(pool, n, head, ht, mr->d_small, "convert_fp_heads", 12.0 * n);
    fetch_small(mr, 2);
    U = mr->h_small[0];
    if (mr->h_small[1] == 0) {
      heads = static_cast<FlagTiles &&>(ht);
      break;
    }
  }
  const uint64_t *skc = sk;
  const uint32_t *spc = sp;
  auto head = [=] __device__(uint64_t i) -> bool { return i == 0 || __ldg(skc + i) != __ldg(skc + i - 1); };

  // start[u], and G[i] = value bytes before sorted element i
  Buf<uint32_t> start(pool, U + 1);
  Buf<uint64_t> G(pool, n + 1);
  uint32_t *st = start.p;
  uint64_t *g = G.p;
  flag_sink(pool, heads, head, [=] __device__(uint64_t i, unsigned long long e, unsigned long long h) {
    if (h) st[e] = (uint32_t) i;
  }, "convert_fp_starts", 8.0 * n);
  uint32_t n32 = (uint32_t) n;
  MR_CUDA(cudaMemcpyAsync(st + U, &n32, 4, cudaMemcpyHostToDevice, s));
  // value bytes of every pair as a column in arrival order (one coalesced pass), gathered in sorted order: 4-byte
  // reads from 4n bytes instead of from the records themselves
  Buf<uint32_t> vbs(pool, n);
  {
    uint32_t *pv = vbs.p;
    ProfScope ps(pool, "convert_fp_valuebytes", 12.0 * n);
    for_each_kernel<<<grid_for(n, 256, 64), 256, 0, s>>>(n, [=] __device__(uint64_t i) { pv[i] = ((const uint32_t *) (data + rec_off(ko, rs, i)))[1]; });
    MR_CUDA(cudaGetLastError());
  }
  const uint32_t *cvb = vbs.p;
  scan_apply(
      pool, n, [=] __device__(uint64_t i) -> unsigned long long { return __ldg(cvb + __ldg(spc + i)); },
      [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { g[i] = e; }, mr->d_small + 2, "convert_fp_valuebytes", 16.0 * n);
  MR_CUDA(cudaMemcpyAsync(g + n, mr->d_small + 2, 8, cudaMemcpyDeviceToDevice, s));

  DevKMV *kmv = new DevKMV(&pool);
  try {
    kmv->offs = (uint64_t *) pool.alloc((U + 1) * 8);
    kmv->offs_cap = U + 1;
    uint64_t *offs = kmv->offs;
    unsigned long long *ksum = mr->d_small + 3;
    MR_CUDA(cudaMemsetAsync(ksum, 0, 8, s));
    scan_apply(
        pool, U,
        [=] __device__(uint64_t u) -> unsigned long long {
          uint32_t s0 = st[u], s1 = st[u + 1];
          uint32_t kb = *(const uint32_t *) (data + rec_off(ko, rs, spc[s0]));
          return kmv_rec_size(al, s1 - s0, kb, g[s1] - g[s0]);
        },
        [=] __device__(uint64_t u, unsigned long long e, unsigned long long) {
          offs[u] = e;
          warp_add(ksum, *(const uint32_t *) (data + rec_off(ko, rs, spc[st[u]])));
        },
        mr->d_small, "convert_fp_sizes", 32.0 * U);
    fetch_small(mr, 4);
    const uint64_t total = mr->h_small[0];
    MR_CUDA(cudaMemcpyAsync(offs + U, &total, 8, cudaMemcpyHostToDevice, s));
    kmv->data = (uint8_t *) pool.alloc(total);
    kmv->cap = total;
    kmv->nbytes = total;
    kmv->ksize = mr->h_small[3];
    uint8_t *out = kmv->data;
    flag_sink(pool, heads, head, [=] __device__(uint64_t i, unsigned long long e, unsigned long long h) {
      uint64_t u = e + h - 1;
      uint32_t s0 = st[u], s1 = st[u + 1];
      uint64_t nv = s1 - s0, r = i - s0;
      const uint8_t *src = data + rec_off(ko, rs, spc[i]);
      uint32_t kb = ((const uint32_t *) src)[0], vb = ((const uint32_t *) src)[1];
      uint64_t mvb = g[s1] - g[s0];
      uint8_t *rec = out + offs[u];
      uint64_t koff = kmv_key_off(al, nv), vo = kmv_val_off(al, nv, kb);
      if (h) {
        uint32_t *hd = (uint32_t *) rec;
        hd[0] = (uint32_t) nv;
        hd[1] = kb;
        hd[2] = (uint32_t) mvb;
        for (uint64_t q = 12 + 4 * nv; q < koff; q++) rec[q] = 0;
        const uint8_t *k = src + kv_key_off(al);
        for (uint32_t q = 0; q < kb; q++) rec[koff + q] = k[q];
        for (uint64_t q = koff + kb; q < vo; q++) rec[q] = 0;
        uint64_t end = kmv_rec_size(al, nv, kb, mvb);
        for (uint64_t q = vo + mvb; q < end; q++) rec[q] = 0;
      }
      ((uint32_t *) (rec + 12))[r] = vb;
      const uint8_t *v = src + kv_val_off(al, kb);
      uint8_t *d = rec + vo + (g[i] - g[s0]);
      for (uint32_t q = 0; q < vb; q++) d[q] = v[q];
    }, "convert_fp_write", (double) kv->nbytes + (double) total);
    kmv->nkmv = U;
    kmv->nvalue = n;
    kmv->vsize = kv->vsize;
    kmv->ukb = kv->uniform ? (int64_t) kv->kb : -1;
    kmv->uvb = kv->uniform ? (int64_t) kv->vb : -1;
    sync(mr);
  } catch (...) {
    delete kmv;
    throw;
  }
  mr->kmv = kmv;
}

static bool convert_dense(mrgpu_mr *mr);
static bool convert_hash(mrgpu_mr *mr);

static void convert_resident(mrgpu_mr *mr)
{
  freeze(mr);
  DevKV *kv = mr->kv;
  if (kv->nkv >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "convert: more than 2^32-2 pairs on one GPU is not supported yet");
  delete mr->kmv;
  mr->kmv = nullptr;
  const uint64_t n = kv->nkv;
  if (n == 0) {
    DevKMV *kmv = new DevKMV(&mr->pool);
    kmv->offs = (uint64_t *) mr->pool.alloc(8);
    kmv->offs_cap = 1;
    MR_CUDA(cudaMemsetAsync(kmv->offs, 0, 8, mr->pool.stream));
    mr->kmv = kmv;
  } else {
    bool done = convert_dense(mr);  // small integer keys + indistinguishable pairs: count, do not sort
    // the uniform sort path reads header-less rows [key | value] as they are; rows with a shared value or with
    // multiplicities become records first
    if (!done && kv->packed && (kv->weighted || kv->cval_on)) kv_unpack(mr, kv);
    const Align al = mr->al;
    const RowLayout lay = kv_layout(mr, kv);
    // exact path needs word-aligned records/keys and a value that is empty, 4 or 8 bytes, or gatherable
    if (!done && kv->uniform && (al.k % 4) == 0) {
      CompactKey ck{};
      bool ident = kv->kb <= 8;
      std::vector<uint8_t> key0(kv->kb);
      bool ok = true;
      // 16-byte keys again (an iterative job's next round): extract on the byte set of last time, checked on the way
      if (!ident && kv->kb == 16 && mr->guess_nvb >= 0 && lay.rs % 4 == 0 && lay.koff % 4 == 0 && ((uintptr_t) kv->data % 16) == 0 && n >= (1u << 16)) {
        ck.nvb = mr->guess_nvb;
        memcpy(ck.pos, mr->guess_pos, 8);
        const int pm = kv->vb == 0 ? 0 : ((kv->vb == 4 && lay.voff % 4 == 0) ? 1 : ((kv->vb == 8 && lay.voff % 4 == 0) ? 2 : 3));
#define CONV_GUESS(K, PM) done = convert_uniform<K, PM>(mr, ck, false, key0, true)
        if (ck.nvb <= 4) {
          if (pm == 0) CONV_GUESS(uint32_t, 0);
          else if (pm == 1) CONV_GUESS(uint32_t, 1);
          else if (pm == 2) CONV_GUESS(uint32_t, 2);
          else CONV_GUESS(uint32_t, 3);
        } else {
          if (pm == 0) CONV_GUESS(uint64_t, 0);
          else if (pm == 1) CONV_GUESS(uint64_t, 1);
          else if (pm == 2) CONV_GUESS(uint64_t, 2);
          else CONV_GUESS(uint64_t, 3);
        }
#undef CONV_GUESS
        ck = CompactKey{};
      }
      if (done) {
      } else if (ident) {
        ck.nvb = (int) kv->kb;
        for (int q = 0; q < 8; q++) ck.pos[q] = (unsigned char) q;
      } else if (kv->kb <= 1024) {
        Buf<unsigned> mask(mr->pool, kv->kb);
        MR_CUDA(cudaMemsetAsync(mask.p, 0, kv->kb * 4, mr->pool.stream));
        const bool words = kv->kb % 4 == 0 && kv->kb <= 32 && lay.rs % 4 == 0 && lay.koff % 4 == 0;
        {
          ProfScope ps(mr->pool, "convert_key_diff", (double) n * lay.rs);
          if (words)
            key_diff_words_kernel<<<grid_for(n, 256 * 8, 8), 256, 0, mr->pool.stream>>>((const uint32_t *) kv->data, n, lay.rs / 4, lay.koff / 4,
                                                                                         kv->kb / 4, mask.p);
          else
            key_diff_kernel<<<grid_for(n, 256 * 8, 4), 256, kv->kb * 4, mr->pool.stream>>>(kv->data, n, lay.rs, lay.koff, kv->kb, mask.p);
          MR_CUDA(cudaGetLastError());
        }
        std::vector<unsigned> hm(kv->kb);
        MR_CUDA(cudaMemcpyAsync(hm.data(), mask.p, kv->kb * 4, cudaMemcpyDeviceToHost, mr->pool.stream));
        MR_CUDA(cudaMemcpyAsync(key0.data(), kv->data + lay.koff, kv->kb, cudaMemcpyDeviceToHost, mr->pool.stream));
        sync(mr);
        if (words) {  // word ORs -> one flag per key byte
          std::vector<unsigned> bytes(kv->kb);
          for (uint32_t q = 0; q < kv->kb; q++) bytes[q] = (hm[q / 4] >> (8 * (q % 4))) & 0xFF;
          hm = bytes;
        }
        int nv = 0;
        for (uint32_t q = 0; q < kv->kb; q++)
          if (hm[q]) {
            if (nv < 8) ck.pos[nv] = (unsigned char) q;
            nv++;
          }
        if (nv > 8 || kv->kb > 256) ok = false;
        ck.nvb = nv;
        if (ok && kv->kb == 16) {
          mr->guess_nvb = nv;
          memcpy(mr->guess_pos, ck.pos, 8);
        }
      } else
        ok = false;
      if (ok && !done) {
        bool small = ck.nvb <= 4;
        bool v32 = kv->vb == 4 && (lay.voff % 4) == 0;
        bool v64 = kv->vb == 8 && (lay.voff % 4) == 0;
        int pm = kv->vb == 0 ? 0 : (v32 ? 1 : (v64 ? 2 : 3));
#define CONV_CASE(K, PM) convert_uniform<K, PM>(mr, ck, ident, key0)
        if (small) {
          if (pm == 0) CONV_CASE(uint32_t, 0);
          else if (pm == 1) CONV_CASE(uint32_t, 1);
          else if (pm == 2) CONV_CASE(uint32_t, 2);
          else CONV_CASE(uint32_t, 3);
        } else {
          if (pm == 0) CONV_CASE(uint64_t, 0);
          else if (pm == 1) CONV_CASE(uint64_t, 1);
          else if (pm == 2) CONV_CASE(uint64_t, 2);
          else CONV_CASE(uint64_t, 3);
        }
#undef CONV_CASE
        done = true;
      }
    }
    // variable-length keys without values (words): count by hashing; falls through to the fingerprint sort when the
    // table overflows (too many distinct keys) or two keys share a fingerprint
    if (!done && !kv->uniform) done = convert_hash(mr);
    if (!done) {
      kv_unpack(mr, kv);  // the general path reads records
      convert_general(mr);
    }
  }
  // convert consumes the KV (src/mapreduce.cpp:876-878: "delete kv; kv = NULL")
  delete mr->kv;
  mr->kv = nullptr;
}

// convert(): of the KV in HBM, or -- when part of it lives in host memory (out-of-core tier, ops_spool.cuh) -- set by set
static void convert_impl(mrgpu_mr *mr)
{
  freeze(mr);
  spool_clear_kmv(mr);
  if (!mr->kv_spool.empty()) convert_spooled(mr);
  else convert_resident(mr);
}

extern "C" int mrgpu_convert(mrgpu_mr *mr, uint64_t *nkmv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot convert without KeyValue");
  mr->g_what = G_NONE;
  convert_impl(mr);
  sync(mr);
  if (nkmv) *nkmv = global_count(mr, kmv_total(mr), G_KMV, false);
  API_END(mr)
}

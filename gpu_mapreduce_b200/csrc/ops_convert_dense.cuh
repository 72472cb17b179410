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

constexpr uint32_t DENSE_BITS = 26;                 // table of 2^26 u32 counts = 256 MB (touched part stays in L2)
constexpr uint32_t DENSE_SLOTS = 1u << DENSE_BITS;
constexpr int DENSE_CACHE = 8192;                   // per-block cache entries (64 KB of shared memory)
constexpr uint32_t DENSE_EMPTY = 0xFFFFFFFFu;

struct DenseArgs {
  const uint8_t *data;
  uint64_t n;
  uint32_t rs, koff, voff, kb, vb;
};

__global__ void __launch_bounds__(512)
dense_count_kernel(DenseArgs a, unsigned *__restrict__ table, unsigned long long *abort_flag)
{
  extern __shared__ unsigned dsm[];
  unsigned *ckey = dsm, *ccnt = dsm + DENSE_CACHE;
  for (int i = threadIdx.x; i < DENSE_CACHE; i += blockDim.x) {
    ckey[i] = DENSE_EMPTY;
    ccnt[i] = 0;
  }
  __syncthreads();
  // value of record 0 (compared word-wise; vb is 0, 4 or 8 here)
  uint32_t v0a = 0, v0b = 0;
  if (a.vb >= 4) v0a = *(const uint32_t *) (a.data + a.voff);
  if (a.vb == 8) v0b = *(const uint32_t *) (a.data + a.voff + 4);
  bool bad = false;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  const bool rec16 = a.rs == 16 && ((uintptr_t) a.data % 16) == 0;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    uint32_t key, khi = 0, va = v0a, vb2 = v0b;
    if (rec16) {  // one 16-byte load per record: {kb, vb, w2, w3}
      int4 r = ld_stream((const int4 *) a.data + i);
      key = (uint32_t) r.z;
      if (a.kb == 8) khi = (uint32_t) r.w;
      else if (a.vb == 4) va = (uint32_t) r.w;
    } else {
      const uint8_t *r = a.data + i * a.rs;
      key = *(const uint32_t *) (r + a.koff);
      if (a.kb == 8) khi = *(const uint32_t *) (r + a.koff + 4);
      if (a.vb >= 4) va = *(const uint32_t *) (r + a.voff);
      if (a.vb == 8) vb2 = *(const uint32_t *) (r + a.voff + 4);
    }
    if (khi != 0 || key >= DENSE_SLOTS || va != v0a || vb2 != v0b) {
      bad = true;
      break;
    }
    unsigned slot = (key * 0x9E3779B1u) >> (32 - 13);  // 8192 slots
    unsigned old = ckey[slot];
    if (old == DENSE_EMPTY) old = atomicCAS(&ckey[slot], DENSE_EMPTY, key);
    if (old == DENSE_EMPTY || old == key) atomicAdd(&ccnt[slot], 1u);
    else atomicAdd(&table[key], 1u);
  }
  if (__syncthreads_or(bad)) {
    if (threadIdx.x == 0) *abort_flag = 1;
    return;  // the table is abandoned; the caller falls back to the sort path
  }
  for (int i = threadIdx.x; i < DENSE_CACHE; i += blockDim.x)
    if (ccnt[i]) atomicAdd(&table[ckey[i]], ccnt[i]);
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
                  const uint32_t *__restrict__ ucnt, uint64_t U, Align al, uint32_t kb, uint32_t vb, uint32_t cv0, uint32_t cv1)
{
  __shared__ uint64_t s_off[FILL_MAX_PAIRS + 1];
  __shared__ uint32_t s_key[FILL_MAX_PAIRS], s_cnt[FILL_MAX_PAIRS];
  __shared__ uint64_t s_u0;
  __shared__ int s_np;
  const uint64_t w0 = (uint64_t) blockIdx.x * FILL_CHUNK_WORDS;
  const uint64_t b0 = w0 * 4, b1 = min((w0 + FILL_CHUNK_WORDS) * 4, total_words * 4);
  if (threadIdx.x == 0) {
    // first pair whose end lies beyond the chunk start: upper_bound(offs[1..U], b0)
    uint64_t lo = 0, hi = U;
    while (lo < hi) {
      uint64_t mid = (lo + hi) >> 1;
      if (offs[mid + 1] > b0) hi = mid;
      else lo = mid + 1;
    }
    s_u0 = lo;
  }
  __syncthreads();
  const uint64_t u0 = s_u0;
  // pairs u0 .. while offs[u] < b1
  for (int q = threadIdx.x; q <= FILL_MAX_PAIRS; q += FILL_THREADS) {
    uint64_t u = u0 + q;
    s_off[q] = u <= U ? offs[u] : ~0ull;
    if (q < FILL_MAX_PAIRS && u < U) {
      s_key[q] = ukey[u];
      s_cnt[q] = ucnt[u];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int np = 0;
    while (np < FILL_MAX_PAIRS && s_off[np] < b1) np++;
    s_np = np;  // pairs 0..np-1 start before the chunk end
  }
  __syncthreads();
  const int np = s_np;
#pragma unroll 4
  for (int k = 0; k < FILL_WORDS_PER_THREAD; k++) {
    uint64_t w = w0 + (uint64_t) k * FILL_THREADS + threadIdx.x;
    if (w >= total_words) break;
    uint64_t b = w * 4;
    // last pair with s_off <= b
    int lo = 0, hi = np - 1;
    while (lo < hi) {
      int mid = (lo + hi + 1) >> 1;
      if (s_off[mid] <= b) lo = mid;
      else hi = mid - 1;
    }
    const uint64_t r = b - s_off[lo];
    const uint64_t nv = s_cnt[lo];
    const uint64_t ko = kmv_key_off(al, nv), vo = kmv_val_off(al, nv, kb);
    uint32_t word = 0;
    if (r == 0) word = (uint32_t) nv;
    else if (r == 4) word = kb;
    else if (r == 8) word = (uint32_t) (nv * vb);
    else if (r < 12 + 4 * nv) word = vb;
    else if (r >= ko && r < ko + kb) word = (r - ko) == 0 ? s_key[lo] : 0u;  // 8-byte keys: high word is 0 here
    else if (r >= vo && r < vo + nv * vb) word = (vb == 8 && ((r - vo) & 4)) ? cv1 : cv0;
    out[w] = word;
  }
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
  if ((al.k % 4) || (al.v % 4) || n < (1u << 16)) return false;  // small inputs: the table scan would dominate
  const uint32_t rs = kv->recsize, koff = kv_key_off(al), voff = kv_val_off(al, kb);

  Buf<unsigned> table(pool, DENSE_SLOTS);
  unsigned long long *flag = mr->d_small + 8;
  MR_CUDA(cudaMemsetAsync(table.p, 0, (uint64_t) DENSE_SLOTS * 4, s));
  MR_CUDA(cudaMemsetAsync(flag, 0, 8, s));
  static bool attr_set[64] = {false};
  if (mr->device < 64 && !attr_set[mr->device]) {
    MR_CUDA(cudaFuncSetAttribute(dense_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DENSE_CACHE * 8));
    attr_set[mr->device] = true;
  }
  {
    ProfScope ps(pool, "convert_dense_count", (double) kv->nbytes);
    DenseArgs a{kv->data, n, rs, koff, voff, kb, vb};
    dense_count_kernel<<<kSMs * 3, 512, DENSE_CACHE * 8, s>>>(a, table.p, flag);
    MR_CUDA(cudaGetLastError());
  }
  unsigned long long h_flag = 1;
  uint32_t cv[2] = {0, 0};
  MR_CUDA(cudaMemcpyAsync(&h_flag, flag, 8, cudaMemcpyDeviceToHost, s));
  if (vb) MR_CUDA(cudaMemcpyAsync(cv, kv->data + voff, vb, cudaMemcpyDeviceToHost, s));
  sync(mr);
  if (h_flag) return false;

  // non-empty bins -> U, ukey[u], ucnt[u]
  const unsigned *tb = table.p;
  auto nonzero = [=] __device__(uint64_t k) -> bool { return __ldg(tb + k) != 0; };
  FlagTiles bins;
  flag_tiles(pool, DENSE_SLOTS, nonzero, bins, mr->d_small, "convert_dense_bins", (double) DENSE_SLOTS * 4);
  fetch_small(mr, 1);
  const uint64_t U = mr->h_small[0];
  Buf<uint32_t> ukey(pool, U), ucnt(pool, U);
  {
    uint32_t *pk = ukey.p, *pc = ucnt.p;
    flag_sink(pool, bins, nonzero, [=] __device__(uint64_t k, unsigned long long e, unsigned long long h) {
      if (h) {
        pk[e] = (uint32_t) k;
        pc[e] = __ldg(tb + k);
      }
    }, "convert_dense_bins", (double) DENSE_SLOTS * 4 + 8.0 * U);
  }
  table.release();

  DevKMV *kmv = new DevKMV(&pool);
  try {
    kmv->offs = (uint64_t *) pool.alloc((U + 1) * 8);
    kmv->offs_cap = U + 1;
    uint64_t *offs = kmv->offs;
    const uint32_t *pc = ucnt.p;
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
    {
      ProfScope ps(pool, "convert_dense_fill", (double) total + 16.0 * U);
      const uint64_t words = total / 4;
      dense_fill_kernel<<<(unsigned) ((words + FILL_CHUNK_WORDS - 1) / FILL_CHUNK_WORDS), FILL_THREADS, 0, s>>>(
          (uint32_t *) kmv->data, words, offs, ukey.p, ucnt.p, U, al, kb, vb, cv[0], cv[1]);
      MR_CUDA(cudaGetLastError());
    }
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

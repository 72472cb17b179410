// convert() by hashing: group-by-key for pairs that carry no value bytes, keys of any length.
//
// Reference: KeyMultiValue::convert hashes every key into a chained table of "Unique" records (kv2unique
// src/keymultivalue.cpp:645-789, find :1403, hash :1428) and later re-hashes every pair to append its value
// (kv2kmv :1139-1209).  When no pair has a value the KMV is fully described by (key, number of pairs) per distinct
// key, so grouping is counting and a hash table is the right tool: no pair moves at all.  (Pairs WITH values keep the
// stable sort of ops_convert.cuh: it delivers each key's values in arrival order, which atomics cannot.)
//
// One pass over the records:
//   * fingerprint of the key: the two result words of one lookup3 pass (key_fingerprint, hash.cuh; the high half is the
//     reference's own bucket hash);
//   * block-local open-addressing table in shared memory (linear probing, claimed with a 64-bit CAS): the keys that
//     come up again and again -- a text corpus is Zipf distributed -- are counted there and never reach L2;
//   * what does not fit goes to an open-addressing table in global memory small enough to stay in L2 (CAS to claim,
//     RED to count); each block adds its shared entries at the end;
//   * exactness: an entry remembers one record of its group (`rep`, kept with atomicMin, which hands back the previous
//     one); whoever joins a group compares its key BYTE FOR BYTE with the record it displaced or found there.  By
//     induction every member equals the first, so two different keys can never be merged; if two ever share a
//     fingerprint, or the global table fills up (more distinct keys than it holds), a flag is raised and the caller
//     falls back to the fingerprint sort.
// The non-empty entries are then compacted, ordered by fingerprint (the order the sort path produces) and become
// a factored KMV with variable-length keys (DevKMV::fvar): key bytes once, a count per key.  Included by mrgpu.cu.
#pragma once

constexpr int HG_THREADS = 768;
constexpr uint32_t HG_SLOTS = 2560;             // per block: 2560 x (16 + 8 + 4 + 4 + 4) bytes = 90 KB, two blocks per SM
constexpr uint32_t HG_PROBES = 8;
constexpr uint32_t HG_GLOBAL_LOG2 = 21;         // 2^21 entries x 32 bytes = 64 MB; the part in use stays in the 126 MB L2
constexpr uint32_t HG_GLOBAL_PROBES = 64;
constexpr uint32_t HG_UNSET = 0xFFFFFFFFu;

// One 32-byte sector per entry.  Keys of up to 12 bytes (an 11-letter word and its NUL) are kept in the entry, so
// joining a group in the global table is one 16-byte read of (key words, len) and one RED on the count; without them
// every join walked entry -> representative's offset -> representative's record, three dependent L2 reads.
struct __align__(32) HgEntry {
  unsigned long long fp;  // 0 = empty
  unsigned cnt;
  unsigned rep;           // a record of this group (the smallest index seen so far)
  unsigned kw[3];         // the key, zero padded, when it fits
  unsigned len;           // 0 = not published yet; else keybytes + 1, bit 31 set when the key is not in kw
};

struct HgRecs {
  const uint8_t *data;
  const uint64_t *offs;   // nullptr for uniform records
  uint64_t recsize;
  Align al;
};

__device__ __forceinline__ const uint8_t *hg_key(const HgRecs &r, uint64_t i, uint32_t &kb)
{
  const uint8_t *p = r.data + rec_off(r.offs, r.recsize, i);
  kb = *(const uint32_t *) p;
  return p + kv_key_off(r.al);
}

__device__ __forceinline__ bool hg_same_key(const HgRecs &r, uint32_t i, uint32_t j)
{
  uint32_t ka, kb;
  const uint8_t *a = hg_key(r, i, ka), *b = hg_key(r, j, kb);
  if (ka != kb) return false;
  uint32_t q = 0;
  if (((((uintptr_t) a) | ((uintptr_t) b)) & 3) == 0)
    for (; q + 4 <= ka; q += 4)
      if (*(const uint32_t *) (a + q) != *(const uint32_t *) (b + q)) return false;
  for (; q < ka; q++)
    if (a[q] != b[q]) return false;
  return true;
}

// add `cnt` pairs represented by record `rep` to the global table; flags: 1 = table full, 2 = fingerprint collision.
// kw / len: the key's first words (zero padded) and keybytes + 1; in12 = the key fits the entry's 12 bytes.
// Exactness: whoever claims an entry also enters the representative chain (atomicMin on rep hands back the previous
// representative, against which the newcomer compares its key byte for byte), so the claimer's key equals every key
// that joined through the chain; a key that joins by comparing with the published words equals the claimer's.
__device__ __forceinline__ void hg_global_add(HgEntry *__restrict__ g, uint32_t gmask, unsigned long long fp, unsigned cnt, unsigned rep,
                                              const HgRecs &r, unsigned long long *flags, uint4 kw, uint32_t len, bool in12)
{
  uint32_t slot = (uint32_t) (fp >> 20) & gmask;
  const uint32_t mylen = in12 ? len : (len | 0x80000000u);
  for (uint32_t p = 0; p < HG_GLOBAL_PROBES; p++) {
    unsigned long long cur = __ldcg(&g[slot].fp);
    bool mine = false;
    if (cur == 0) {
      cur = atomicCAS(&g[slot].fp, 0ull, fp);
      mine = cur == 0;
    }
    if (mine || cur == fp) {
      atomicAdd(&g[slot].cnt, cnt);
      if (mine) {  // publish the key: words first, then the length
        if (in12) g[slot].kw[0] = kw.x, g[slot].kw[1] = kw.y, g[slot].kw[2] = kw.z;
        __threadfence();
        atomicExch(&g[slot].len, mylen);
      } else if (in12) {
        const uint4 e = __ldcg((const uint4 *) &g[slot].kw[0]);  // (kw0, kw1, kw2, len) in one read
        if (e.w == mylen) {
          if (e.x != kw.x || e.y != kw.y || e.z != kw.z) atomicOr(flags, 2ull);
          return;
        }
        if (e.w != 0) {  // same fingerprint, a key of another length
          atomicOr(flags, 2ull);
          return;
        }
        // claimed but not published yet: the representative chain below decides
      }
      const unsigned old = atomicMin(&g[slot].rep, rep);
      if (old != HG_UNSET && old != rep && !hg_same_key(r, rep, old)) atomicOr(flags, 2ull);
      return;
    }
    slot = (slot + 1) & gmask;
  }
  atomicOr(flags, 1ull);
}

// Keys of up to 16 bytes -- nearly every word of a text -- are kept INSIDE the block's table entry next to the
// fingerprint, so a hit is decided by comparing four words in shared memory.  (The first version verified every hit
// against the representative record in global memory: ncu counted 12 GB of L1 misses per 31 M words and 15.6 of 23
// stall cycles per issue on the long scoreboard.)  Longer keys still verify against the representative.
// Entry life cycle: fp 0 -> claimed by a 64-bit CAS on fp -> the owner writes key words and rep, fences, publishes
// len (kb + 1; bit 31 set for keys that are not inline).  A reader that finds its fingerprint but no published length
// yet (the owner is a few instructions behind, possibly in the same warp) does not wait: it sends its pair to the
// global table, which is always correct.
__global__ void __launch_bounds__(HG_THREADS)
hash_group_kernel(HgRecs r, uint64_t n, uint32_t seed, HgEntry *__restrict__ g, uint32_t gmask, unsigned long long *flags)
{
  extern __shared__ __align__(16) unsigned char hg_sm[];
  uint4 *s_w = (uint4 *) hg_sm;                                              // inline key words
  unsigned long long *s_fp = (unsigned long long *) (s_w + HG_SLOTS);
  unsigned *s_cnt = (unsigned *) (s_fp + HG_SLOTS), *s_rep = s_cnt + HG_SLOTS;
  volatile unsigned *s_len = s_rep + HG_SLOTS;
  for (uint32_t i = threadIdx.x; i < HG_SLOTS; i += blockDim.x) s_fp[i] = 0, s_cnt[i] = 0, s_rep[i] = HG_UNSET, s_len[i] = 0;
  __syncthreads();
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  // every lane of a warp makes the same number of trips and the warp is brought back together at the top of each:
  // the probe / verify branches below split it, and without the barrier the pieces kept running apart -- each
  // hashing its next word alone (ncu: 1 to 4 active threads per instruction in the hash code)
  for (uint64_t base = (uint64_t) blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < n; base += stride) {
    __syncwarp();
    const uint64_t i = base + (threadIdx.x & 31u);
    if (i < n) {
      uint32_t kb;
      const uint8_t *k = hg_key(r, i, kb);
      unsigned long long fp = key_fingerprint(k, kb, seed);
      if (fp == 0) fp = 1;
      // the key as four words, bytes past its end zeroed (only meaningful when it fits: kb <= 16 and word aligned)
      const bool inl = kb <= 16 && ((((uintptr_t) k) & 3) == 0);
      uint4 kw = make_uint4(0, 0, 0, 0);
      if (inl) {
        const uint32_t *w = (const uint32_t *) k;
        const uint32_t nw = (kb + 3) >> 2, tail = kb & 3u, m = tail ? (0xFFFFFFFFu >> (32 - 8 * tail)) : 0xFFFFFFFFu;
        if (nw > 0) kw.x = w[0] & (nw == 1 ? m : 0xFFFFFFFFu);
        if (nw > 1) kw.y = w[1] & (nw == 2 ? m : 0xFFFFFFFFu);
        if (nw > 2) kw.z = w[2] & (nw == 3 ? m : 0xFFFFFFFFu);
        if (nw > 3) kw.w = w[3] & m;
      }
      const uint32_t mylen = (kb + 1) | (inl ? 0u : 0x80000000u);
      uint32_t slot = __umulhi((uint32_t) (fp ^ (fp >> 29)), HG_SLOTS);
      bool done = false, to_global = false;
#pragma unroll 1
      for (uint32_t p = 0; p < HG_PROBES && !done && !to_global; p++) {
        unsigned long long cur = s_fp[slot];
        if (cur == 0) {
          cur = atomicCAS(&s_fp[slot], 0ull, fp);
          if (cur == 0) {  // mine: fill in, then publish
            s_w[slot] = kw;
            s_rep[slot] = (unsigned) i;
            atomicAdd(&s_cnt[slot], 1u);
            __threadfence_block();
            s_len[slot] = mylen;
            done = true;
            break;
          }
        }
        if (cur == fp) {
          const uint32_t len = s_len[slot];
          __threadfence_block();  // the key words and rep below were written before len was published
          if (len == 0) to_global = true;  // claimed but not yet published
          else if (len != mylen) atomicOr(flags, 2ull), done = true;  // same fingerprint, another key
          else if (inl) {
            const uint4 e = s_w[slot];
            if (e.x == kw.x && e.y == kw.y && e.z == kw.z && e.w == kw.w) atomicAdd(&s_cnt[slot], 1u);
            else atomicOr(flags, 2ull);
            done = true;
          } else {
            atomicAdd(&s_cnt[slot], 1u);
            if (!hg_same_key(r, (uint32_t) i, s_rep[slot])) atomicOr(flags, 2ull);
            done = true;
          }
        } else {
          slot++;
          if (slot == HG_SLOTS) slot = 0;
        }
      }
      if (!done) hg_global_add(g, gmask, fp, 1u, (unsigned) i, r, flags, kw, kb + 1, inl && kb <= 12);  // the block's table is taken by more frequent keys
    }
  }
  __syncthreads();
  for (uint32_t s = threadIdx.x; s < HG_SLOTS; s += blockDim.x)
    if (s_fp[s]) {
      const uint32_t len = s_len[s];
      hg_global_add(g, gmask, s_fp[s], s_cnt[s], s_rep[s], r, flags, s_w[s], len & 0x7FFFFFFFu, !(len & 0x80000000u) && (len & 0x7FFFFFFFu) <= 13);
    }
}

// factored KMV with variable-length keys -> the reference's record stream (every value is empty: the size array of a
// pair is nv zeros, there are no value bytes)
__global__ void __launch_bounds__(256)
fvar_expand_kernel(const uint8_t *__restrict__ vkey, const uint64_t *__restrict__ vkoff, const uint32_t *__restrict__ fcnt, uint64_t U, Align al,
                   uint8_t *__restrict__ out, const uint64_t *__restrict__ offs)
{
  const uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  const uint32_t kb = (uint32_t) (vkoff[u + 1] - vkoff[u]);
  const uint64_t nv = fcnt[u];
  uint8_t *rec = out + offs[u];
  ((uint32_t *) rec)[0] = (uint32_t) nv;
  ((uint32_t *) rec)[1] = kb;
  ((uint32_t *) rec)[2] = 0;
  uint8_t *k = rec + kmv_key_off(al, nv);
  const uint8_t *src = vkey + vkoff[u];
  for (uint32_t q = 0; q < kb; q++) k[q] = src[q];
}

static void kmv_expand_fvar(mrgpu_mr *mr, DevKMV *kmv)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t U = kmv->nkmv;
  kmv->offs = (uint64_t *) pool.alloc((U + 1) * 8);
  kmv->offs_cap = U + 1;
  uint64_t *offs = kmv->offs;
  const uint32_t *pc = kmv->fcnt;
  const uint64_t *ko = kmv->vkoff;
  scan_apply(
      pool, U, [=] __device__(uint64_t u) -> unsigned long long { return kmv_rec_size(al, pc[u], (uint32_t) (ko[u + 1] - ko[u]), 0); },
      [=] __device__(uint64_t u, unsigned long long e, unsigned long long) { offs[u] = e; }, mr->d_small, "convert_kmv_sizes", 12.0 * U);
  fetch_small(mr, 1);
  const uint64_t total = mr->h_small[0];
  MR_CUDA(cudaMemcpyAsync(offs + U, &total, 8, cudaMemcpyHostToDevice, s));
  kmv->data = (uint8_t *) pool.alloc(total ? total : 16);
  kmv->cap = total ? total : 16;
  kmv->nbytes = total;
  if (total) {
    ProfScope ps(pool, "convert_fvar_fill", (double) total);
    MR_CUDA(cudaMemsetAsync(kmv->data, 0, total, s));  // value sizes (all zero) and padding
    fvar_expand_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kmv->vkey, kmv->vkoff, kmv->fcnt, U, al, kmv->data, offs);
    MR_CUDA(cudaGetLastError());
  }
  sync(mr);
  pool.free(kmv->vkey, 0);
  pool.free(kmv->vkoff, 0);
  pool.free(kmv->fcnt, 0);
  kmv->vkey = nullptr;
  kmv->vkoff = nullptr;
  kmv->fcnt = nullptr;
  kmv->fcap = 0;
  kmv->fvar = false;
  kmv->h_offs_valid = false;
}

// returns true if the hash path produced mr->kmv; false = not applicable or abandoned (nothing changed)
static bool convert_hash(mrgpu_mr *mr)
{
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t n = kv->nkv;
  static const bool disabled = getenv("MRGPU_NO_HASH_CONVERT") && atoi(getenv("MRGPU_NO_HASH_CONVERT"));
  if (disabled || !mr->compact || kv->packed || kv->vsize != 0 || n < (1u << 12) || (al.k % 4)) return false;
  const uint32_t gslots = 1u << HG_GLOBAL_LOG2;
  Buf<HgEntry> table(pool, gslots);
  unsigned long long *flags = mr->d_small + 6;
  HgRecs r{kv->data, kv->uniform ? nullptr : kv->offs, kv->recsize, al};
  // empty table: fp 0, count 0, rep UNSET (0xFFFFFFFF): two fills
  static std::atomic<bool> attr[64];
  const size_t smem = (size_t) HG_SLOTS * 36;
  if (mr->device < 64 && !attr[mr->device].load()) {
    MR_CUDA(cudaFuncSetAttribute(hash_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    attr[mr->device].store(true);
  }
  HgEntry *g = table.p;
  {
    ProfScope ps(pool, "convert_hash_clear", (double) gslots * sizeof(HgEntry));
    for_each_kernel<<<grid_for(gslots, 256, 16), 256, 0, s>>>((uint64_t) gslots, [=] __device__(uint64_t i) { g[i] = HgEntry{0ull, 0u, HG_UNSET, {0u, 0u, 0u}, 0u}; });
    MR_CUDA(cudaGetLastError());
  }
  MR_CUDA(cudaMemsetAsync(flags, 0, 8, s));
  {
    ProfScope ps(pool, "convert_hash_group", (double) kv->nbytes + (kv->uniform ? 0.0 : 8.0 * n));
    hash_group_kernel<<<kSMs * 2, HG_THREADS, smem, s>>>(r, n, 0u, g, gslots - 1, flags);
    MR_CUDA(cudaGetLastError());
  }
  // non-empty entries -> (fingerprint, count, rep) arrays
  auto used = [=] __device__(uint64_t i) -> bool { return g[i].fp != 0; };
  FlagTiles ft;
  flag_tiles(pool, gslots, used, ft, mr->d_small, "convert_hash_compact", (double) gslots * sizeof(HgEntry));
  MR_CUDA(cudaMemcpyAsync(mr->h_small + 6, flags, 8, cudaMemcpyDeviceToHost, s));
  fetch_small(mr, 1);
  if (mr->h_small[6]) return false;  // table full or a fingerprint collision: the sort path decides
  const uint64_t U = mr->h_small[0];
  Buf<uint64_t> fp0(pool, U), fp1(pool, U);
  Buf<uint32_t> ix0(pool, U), ix1(pool, U), cnt(pool, U), rep(pool, U);
  {
    uint64_t *pf = fp0.p;
    uint32_t *pi = ix0.p, *pc = cnt.p, *pr = rep.p;
    flag_sink(pool, ft, used, [=] __device__(uint64_t i, unsigned long long e, unsigned long long h) {
      if (h) {
        pf[e] = g[i].fp;
        pi[e] = (uint32_t) e;
        pc[e] = g[i].cnt;
        pr[e] = g[i].rep;
      }
    }, "convert_hash_compact", (double) gslots * sizeof(HgEntry) + 20.0 * U);
  }
  table.release();
  // KMV pairs come out in fingerprint order, like the sort path's
  Buf<unsigned long long> hist(pool, 8 * RADIX);
  const int cur = radix_sort_pairs<uint64_t, uint32_t, true>(pool, fp0.p, fp1.p, ix0.p, ix1.p, U, 0xFF, false, hist.p);
  const uint32_t *perm = cur ? ix1.p : ix0.p;

  DevKMV *kmv = new DevKMV(&pool);
  try {
    kmv->vkoff = (uint64_t *) pool.alloc((U + 1) * 8);
    kmv->fcnt = (uint32_t *) pool.alloc((U ? U : 1) * 4);
    kmv->fcap = U ? U : 1;
    uint64_t *vko = kmv->vkoff;
    uint32_t *fc = kmv->fcnt;
    const uint32_t *pc = cnt.p, *pr = rep.p;
    scan_apply(
        pool, U,
        [=] __device__(uint64_t u) -> unsigned long long {
          uint32_t kb;
          hg_key(r, pr[perm[u]], kb);
          return kb;
        },
        [=] __device__(uint64_t u, unsigned long long e, unsigned long long) {
          vko[u] = e;
          fc[u] = pc[perm[u]];
        },
        mr->d_small, "convert_hash_keys", 24.0 * U);
    fetch_small(mr, 1);
    const uint64_t kbytes = mr->h_small[0];
    MR_CUDA(cudaMemcpyAsync(vko + U, &kbytes, 8, cudaMemcpyHostToDevice, s));
    kmv->vkey = (uint8_t *) pool.alloc(kbytes ? kbytes : 16);
    uint8_t *vk = kmv->vkey;
    for_each_kernel<<<grid_for(U, 256, 64), 256, 0, s>>>(U, [=] __device__(uint64_t u) {
      uint32_t kb;
      const uint8_t *k = hg_key(r, pr[perm[u]], kb);
      uint8_t *d = vk + vko[u];
      for (uint32_t q = 0; q < kb; q++) d[q] = k[q];
    });
    MR_CUDA(cudaGetLastError());
    kmv->fvar = true;
    kmv->nkmv = U;
    kmv->nvalue = n;
    kmv->ksize = kbytes;
    kmv->vsize = 0;
    kmv->ukb = kv->uniform ? (int64_t) kv->kb : -1;
    kmv->uvb = 0;
    sync(mr);
  } catch (...) {
    delete kmv;
    throw;
  }
  mr->kmv = kmv;
  return true;
}

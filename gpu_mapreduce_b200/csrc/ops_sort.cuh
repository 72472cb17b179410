// sort_keys(flag): per-rank sort of a KeyValue by key.
//
// Reference: MapReduce::sort_keys(int) src/mapreduce.cpp:2102-2128 picks a compare function
// (:2692-2802), sort_kv(0) :2359-2453 qsorts an index per page through it and 2-way merges the pages
// through spool files.  glibc's qsort is a merge sort here, i.e. stable.
// Here: an order-preserving transform of the key into an unsigned integer, one stable LSD radix sort of
// (transformed key, record index), then one gather of the records in the new order.
//   +-1 int      : x ^ 0x80000000                       (signed compare)
//   +-2 uint64_t : identity
//   +-3 float    : IEEE total order with -0 folded onto +0 (the reference's < and > call them equal)
//   +-4 double   : same, 64-bit
//   +-5 strcmp   : bytes up to the first NUL, big-endian packed 8 per chunk, chunks sorted last-to-first
//   +-6 strncmp(min len): same as 5 on the key's bytes (identical whenever keys carry their NUL, as every
//                  caller in the reference does; for NUL-less prefix-related keys the reference's order is
//                  qsort-implementation defined)
//   flag < 0     : bitwise complement of the transformed key -> reverse order, ties keep input order like
//                  the reference's stable sort with the reversed comparator.
// Included by mrgpu.cu.
#pragma once

#include "sort_order.cuh"  // f32_order, f64_order, str_chunk: host-callable, checked against the C comparisons on the host

template <typename K>
__global__ void __launch_bounds__(256)
sortkey_extract_kernel(const uint8_t *__restrict__ data, const uint64_t *__restrict__ offs, uint64_t rs, uint64_t n, Align al, int absflag,
                       bool reverse, uint32_t chunk, const uint32_t *__restrict__ perm_in, K *__restrict__ keys, uint32_t *__restrict__ perm,
                       unsigned *bad, bool by_value)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t rec = perm_in ? perm_in[i] : (uint32_t) i;
  const uint8_t *r = data + rec_off(offs, rs, rec);
  // the compared field: the key, or (sort_values) the value -- from here on "k, kb" is that field
  uint32_t kb = *(const uint32_t *) r;
  const uint8_t *k = r + kv_key_off(al);
  if (by_value) {
    k = r + kv_val_off(al, kb);
    kb = ((const uint32_t *) r)[1];
  }
  uint64_t v = 0;
  uint32_t need = (absflag == 1 || absflag == 3) ? 4 : ((absflag == 2 || absflag == 4) ? 8 : 0);
  if (kb < need) {
    *bad = 1;
  } else if (absflag == 1) {
    v = le_bytes(k, 4) ^ 0x80000000u;
  } else if (absflag == 3) {
    v = f32_order(le_bytes(k, 4));
  } else if (absflag == 2 || absflag == 4) {
    uint64_t b = (uint64_t) le_bytes(k, 4) | ((uint64_t) le_bytes(k + 4, 4) << 32);
    v = absflag == 2 ? b : f64_order(b);
  } else {
    v = str_chunk(k, kb, chunk);
  }
  K out = (K) v;
  if (reverse) out = ~out;
  keys[i] = out;
  perm[i] = rec;
}

__global__ void __launch_bounds__(256)
gather_uniform_kernel(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, const uint32_t *__restrict__ perm, uint64_t n, uint32_t words)
{
  // one thread per 4-byte word of the output: consecutive threads write consecutive words
  uint64_t t = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * words) return;
  uint64_t i = t / words;
  uint32_t w = (uint32_t) (t - i * words);
  dst[t] = src[(uint64_t) perm[i] * words + w];
}

__global__ void __launch_bounds__(256)
gather_general_kernel(const uint8_t *__restrict__ src, const uint64_t *__restrict__ soffs, uint8_t *__restrict__ dst,
                      const uint64_t *__restrict__ doffs, const uint32_t *__restrict__ perm, uint64_t n)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t so = soffs[perm[i]], sz = soffs[perm[i] + 1] - so;
  const uint32_t *s = (const uint32_t *) (src + so);
  uint32_t *d = (uint32_t *) (dst + doffs[i]);
  for (uint64_t q = 0; q < sz / 4; q++) d[q] = s[q];
}

__global__ void max_keybytes_kernel(const uint8_t *__restrict__ data, const uint64_t *__restrict__ offs, uint64_t n, unsigned long long *mx, int word)
{
  unsigned long long m = 0;
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    m = max(m, (unsigned long long) ((const uint32_t *) (data + offs[i]))[word]);  // word 0: keybytes, 1: valuebytes
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(mx, m);
}

// number of adjacent pairs out of order (0 = the stable sort would move nothing)
template <typename K>
__global__ void __launch_bounds__(256) sortkey_descents_kernel(const K *__restrict__ k, uint64_t n, unsigned long long *out)
{
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  unsigned c = 0;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n; i += stride) c += k[i] > k[i + 1];
  c = __syncthreads_count(c != 0);
  if (threadIdx.x == 0 && c) atomicAdd(out, (unsigned long long) c);
}

// returns false when the pairs already are in sorted order (perm_out untouched)
template <typename K>
static bool sort_keys_typed(mrgpu_mr *mr, int absflag, bool reverse, uint32_t nchunks, Buf<uint32_t> &perm_out, bool by_value)
{
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const uint64_t n = kv->nkv;
  Buf<K> k0(pool, n), k1(pool, n);
  Buf<uint32_t> p0(pool, n), p1(pool, n);
  Buf<unsigned long long> hist(pool, 8 * RADIX);
  unsigned *bad = (unsigned *) (mr->d_small + 1);
  MR_CUDA(cudaMemsetAsync(bad, 0, 8, s));
  // multi-chunk keys (strings): LSD over chunks, last chunk first; each round re-extracts the chunk of the
  // records in their current order and stable-sorts (chunk, record index).  Round c's extract reads the
  // permutation produced by round c+1 and writes into the other buffer pair.
  K *kb[2] = {k0.p, k1.p};
  uint32_t *pb[2] = {p0.p, p1.p};
  const uint32_t *perm_in = nullptr;
  int w = 0;
  for (int c = (int) nchunks - 1; c >= 0; c--) {
    {
    ProfScope ps(pool, "sort_extract", (double) n * (8 + sizeof(K) + 4));
    sortkey_extract_kernel<K><<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(kv->data, kv->uniform ? nullptr : kv->offs, kv->recsize, n, mr->al,
                                                                         absflag, reverse, (uint32_t) c, perm_in, kb[w], pb[w], bad, by_value);
    MR_CUDA(cudaGetLastError());
    }
    if (nchunks == 1) {
      // adaptive: keys that are already in order (the output of the counting convert + reduce is) need no sort;
      // one read of the extracted keys decides
      unsigned long long *desc = mr->d_small + 2;
      MR_CUDA(cudaMemsetAsync(desc, 0, 8, s));
      {
        ProfScope ps(pool, "sort_check", (double) n * sizeof(K));
        sortkey_descents_kernel<K><<<grid_for(n, 256 * 8, 8), 256, 0, s>>>(kb[w], n, desc);
        MR_CUDA(cudaGetLastError());
      }
      fetch_small(mr, 3);
      if (mr->h_small[1] & 0xFFFFFFFFull) MR_FAIL(MRGPU_EINVAL, by_value ? "sort_values: value shorter than the compared type" : "sort_keys: key shorter than the compared type");
      if (mr->h_small[2] == 0) return false;
    }
    int r = radix_sort_pairs<K, uint32_t, true>(pool, kb[w], kb[w ^ 1], pb[w], pb[w ^ 1], n, 0xFF, false, hist.p);
    int res = w ^ r;
    perm_in = pb[res];
    w = res ^ 1;
  }
  fetch_small(mr, 2);
  if (mr->h_small[1] & 0xFFFFFFFFull) MR_FAIL(MRGPU_EINVAL, by_value ? "sort_values: value shorter than the compared type" : "sort_keys: key shorter than the compared type");
  // hand the permutation buffer to the caller
  if (perm_in == p0.p) perm_out = static_cast<Buf<uint32_t> &&>(p0);
  else perm_out = static_cast<Buf<uint32_t> &&>(p1);
  return true;
}

static void sort_keys_impl(mrgpu_mr *mr, int flag, bool by_value = false)
{
  int absflag = flag < 0 ? -flag : flag;
  if (absflag < 1 || absflag > 6) MR_FAIL(MRGPU_EINVAL, by_value ? "Invalid compare method for sort values" : "Invalid compare method for sort keys");
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const uint64_t n = kv->nkv;
  if (n >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "sort_keys: more than 2^32-2 pairs on one GPU is not supported yet");
  if (n < 2) return;
  if (!by_value && kv->known_sorted(flag)) return;  // reduce() of a counting convert left them in this very order
  uint32_t nchunks = 1;
  if (absflag >= 5) {
    uint64_t maxkb = by_value ? kv->vb : kv->kb;
    if (!kv->uniform) {
      unsigned long long *mx = mr->d_small;
      MR_CUDA(cudaMemsetAsync(mx, 0, 8, s));
      max_keybytes_kernel<<<grid_for(n, 256, 8), 256, 0, s>>>(kv->data, kv->offs, n, mx, by_value ? 1 : 0);
      MR_CUDA(cudaGetLastError());
      fetch_small(mr, 1);
      maxkb = mr->h_small[0];
    }
    nchunks = (uint32_t) std::max<uint64_t>(1, (maxkb + 7) / 8);
  }
  Buf<uint32_t> perm;
  bool moved;
  if (absflag == 1 || absflag == 3) moved = sort_keys_typed<uint32_t>(mr, absflag, flag < 0, 1, perm, by_value);
  else moved = sort_keys_typed<uint64_t>(mr, absflag, flag < 0, nchunks, perm, by_value);
  if (!moved) return;

  // gather the records in sorted order into a new KV
  DevKV *nk = new DevKV(&pool);
  try {
    nk->reserve_bytes(kv->nbytes);
    if (kv->uniform) {
      uint32_t words = kv->recsize / 4;
      uint64_t total = n * words;
      ProfScope ps(pool, "sort_gather", (double) n * (4 + 2.0 * kv->recsize));
      gather_uniform_kernel<<<(unsigned) ((total + 255) / 256), 256, 0, s>>>((const uint32_t *) kv->data, (uint32_t *) nk->data, perm.p, n, words);
      MR_CUDA(cudaGetLastError());
    } else {
      nk->reserve_offs(n + 1);
      const uint64_t *so = kv->offs;
      const uint32_t *pp = perm.p;
      uint64_t *dof = nk->offs;
      scan_apply(
          pool, n, [=] __device__(uint64_t i) -> unsigned long long { return so[pp[i] + 1] - so[pp[i]]; },
          [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { dof[i] = e; }, mr->d_small);
      MR_CUDA(cudaMemcpyAsync(dof + n, mr->d_small, 8, cudaMemcpyDeviceToDevice, s));
      gather_general_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(kv->data, kv->offs, nk->data, nk->offs, perm.p, n);
      MR_CUDA(cudaGetLastError());
    }
    nk->uniform = kv->uniform;
    nk->kb = kv->kb;
    nk->vb = kv->vb;
    nk->recsize = kv->recsize;
    nk->nbytes = kv->nbytes;
    nk->nkv = n;
    nk->ksize = kv->ksize;
    nk->vsize = kv->vsize;
    sync(mr);
  } catch (...) {
    delete nk;
    throw;
  }
  delete mr->kv;
  mr->kv = nk;
}

extern "C" int mrgpu_sort_keys(mrgpu_mr *mr, int flag, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot sort_keys without KeyValue");
  freeze(mr);
  spool_make_resident(mr);
  kv_unpack(mr, mr->kv);
  sort_keys_impl(mr, flag);
  sync(mr);
  if (nkv) *nkv = global_count(mr, mr->kv->nkv, G_KV, true);
  API_END(mr)
}

extern "C" int mrgpu_sort_values(mrgpu_mr *mr, int flag, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot sort_values without KeyValue");
  freeze(mr);
  spool_make_resident(mr);
  kv_unpack(mr, mr->kv);
  sort_keys_impl(mr, flag, true);
  sync(mr);
  if (nkv) *nkv = global_count(mr, mr->kv->nkv, G_KV, true);
  API_END(mr)
}

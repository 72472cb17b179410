// map(): emit into KV pages.  The reference calls a CPU callback per task which calls KeyValue::add per
// pair (src/mapreduce.cpp:1102-1225, src/keyvalue.cpp:343-392).  Here the stock callbacks are device
// functors and "add" is a coalesced store at a position known from the record index (uniform records) or
// from a prefix sum of record sizes (variable records).  Included by mrgpu.cu.
#pragma once

// ---------------------------------------------------------------------------------------------
// KV bookkeeping shared by all map flavours

// map()'s preamble, src/mapreduce.cpp:1114-1131: an existing KMV is dropped; addflag = 0 starts a new KV,
// addflag = 1 re-opens the current one for append (KeyValue::append, src/keyvalue.cpp:185-209).
static void kv_unpack(mrgpu_mr *mr, DevKV *kv);
static void emit_and_count(mrgpu_mr *mr, DevKV *kv, const uint32_t *keys, uint64_t n);  // ops_convert_dense.cuh
static void emit_and_plan(mrgpu_mr *mr, DevKV *kv, const uint32_t *keys, uint64_t n);   // ops_partition.cuh
static void emit_from_host(mrgpu_mr *mr, DevKV *kv, const uint32_t *hkeys, uint64_t n, bool count);  // ops_convert_dense.cuh

static void kv_begin_map(mrgpu_mr *mr, int addflag, bool keep_packed = false)
{
  freeze(mr);
  mr->g_what = G_NONE;
  delete mr->kmv;
  mr->kmv = nullptr;
  spool_clear_kmv(mr);
  if (!addflag) spool_clear_kv(mr);
  if (!addflag || !mr->kv) {
    delete mr->kv;
    mr->kv = new DevKV(&mr->pool);
  } else {
    mr->kv->sorted_flag = 0;  // pairs are about to be appended
    if (!keep_packed) kv_unpack(mr, mr->kv);
  }
}

// ---------------------------------------------------------------------------------------------
// header-less uniform KV (store.cuh) <-> the reference's record stream

// shapes that may be kept header-less: default alignment (record = 8 + kb + vb, no padding), word-sized fields
static bool kv_can_pack(const mrgpu_mr *mr, uint32_t kb, uint32_t vb)
{
  return mr->compact && mr->al.k == 4 && mr->al.v == 4 && kb > 0 && kb % 4 == 0 && vb % 4 == 0;
}

// where key and value sit in what a uniform KV stores per pair: records {kb, vb, key, value}, or header-less rows
struct RowLayout {
  uint32_t rs, koff, voff;
};
static RowLayout kv_layout(const mrgpu_mr *mr, const DevKV *kv)
{
  if (kv->packed) return RowLayout{kv->prow(), 0, kv->kb};
  return RowLayout{kv->recsize, kv_key_off(mr->al), kv_val_off(mr->al, kv->kb)};
}

// a map that emits value-less (kb, 0) pairs may write header-less rows when the KV is empty or already holds such rows
static bool kv_rows_ok(const mrgpu_mr *mr, const DevKV *kv, uint32_t kb)
{
  return kv_can_pack(mr, kb, 0) && (kv->nkv == 0 || (kv->uniform && kv->packed && !kv->weighted && !kv->cval_on && kv->kb == kb && kv->vb == 0));
}
static void kv_rows_begin(DevKV *kv, uint32_t kb, uint32_t recsize)
{
  if (kv->nkv) return;
  kv->uniform = kv->packed = true;
  kv->cval_on = kv->weighted = false;
  kv->cval[0] = kv->cval[1] = 0;
  kv->kb = kb;
  kv->vb = 0;
  kv->recsize = recsize;
}

// 4-byte key + 4-byte value (stored or shared): one 16-byte record per thread
__global__ void __launch_bounds__(256) kv_unpack16_kernel(const uint32_t *__restrict__ rows, uint64_t n, bool cval_on, uint32_t c0, int4 *__restrict__ out)
{
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t k, v;
    if (cval_on) k = __ldg(rows + i), v = c0;
    else {
      uint2 r = __ldg((const uint2 *) rows + i);
      k = r.x, v = r.y;
    }
    st_stream(out + i, make_int4(4, 4, (int) k, (int) v));
  }
}

// any word-sized shape: one thread per output word
__global__ void __launch_bounds__(256)
kv_unpack_kernel(const uint32_t *__restrict__ rows, uint64_t n, uint32_t kw, uint32_t vw, bool cval_on, uint32_t c0, uint32_t c1, uint32_t *__restrict__ out)
{
  const uint32_t rw = 2 + kw + vw, roww = kw + (cval_on ? 0 : vw);
  const uint64_t total = n * rw, stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t w = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; w < total; w += stride) {
    const uint64_t i = w / rw;
    const uint32_t q = (uint32_t) (w - i * rw);
    uint32_t x;
    if (q == 0) x = kw * 4;
    else if (q == 1) x = vw * 4;
    else if (q < 2 + kw || !cval_on) x = rows[i * roww + (q - 2)];
    else x = (q - 2 - kw) ? c1 : c0;
    out[w] = x;
  }
}

// records [first, first + count) of a header-less KV written to out (count * recsize bytes, 16-byte aligned)
static void kv_unpack_rows(mrgpu_mr *mr, const DevKV *kv, uint64_t first, uint64_t count, uint8_t *out)
{
  if (!count) return;
  Pool &pool = mr->pool;
  const uint32_t *rows = (const uint32_t *) (kv->data + first * kv->prow());
  ProfScope ps(pool, "kv_unpack", (double) count * (kv->prow() + kv->recsize));
  if (kv->kb == 4 && kv->vb == 4 && (!kv->cval_on ? ((uintptr_t) rows % 8) == 0 : true))
    kv_unpack16_kernel<<<grid_for(count, 256, 16), 256, 0, pool.stream>>>(rows, count, kv->cval_on, kv->cval[0], (int4 *) out);
  else
    kv_unpack_kernel<<<grid_for(count * (kv->recsize / 4), 256, 16), 256, 0, pool.stream>>>(rows, count, kv->kb / 4, kv->vb / 4, kv->cval_on, kv->cval[0],
                                                                                           kv->cval[1], (uint32_t *) out);
  MR_CUDA(cudaGetLastError());
}

// weighted rows [key | multiplicity] -> records: output record i belongs to the row r with start[r] <= i < start[r+1]
// (start = exclusive prefix sum of the multiplicities); one thread per record, binary search over start[].
// Not on any hot path: a weighted KV is expanded only when somebody asks for its bytes.
__global__ void __launch_bounds__(256)
kv_expand_weighted_kernel(const uint2 *__restrict__ rows, const uint64_t *__restrict__ start, uint64_t nrows, uint64_t total, uint32_t vw, uint32_t c0,
                          uint32_t c1, uint32_t *__restrict__ out)
{
  const uint32_t rw = 3 + vw;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    uint64_t lo = 0, hi = nrows - 1;
    while (lo < hi) {
      const uint64_t mid = (lo + hi + 1) >> 1;
      if (__ldg(start + mid) <= i) lo = mid;
      else hi = mid - 1;
    }
    const uint32_t key = __ldg(&rows[lo].x);
    uint32_t *o = out + i * rw;
    if (vw == 1) *(int4 *) o = make_int4(4, 4, (int) key, (int) c0);
    else {
      o[0] = 4;
      o[1] = vw * 4;
      o[2] = key;
      if (vw == 2) o[3] = c0, o[4] = c1;
    }
  }
}

static void kv_unpack_weighted(mrgpu_mr *mr, DevKV *kv)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const uint64_t R = kv->nrows, total = kv->nkv;
  uint8_t *nd = (uint8_t *) pool.alloc(kv->nbytes ? kv->nbytes : 16);
  if (R && total) {
    Buf<uint64_t> start(pool, R + 1);
    const uint2 *rows = (const uint2 *) kv->data;
    uint64_t *st = start.p;
    scan_apply(
        pool, R, [=] __device__(uint64_t r) -> unsigned long long { return rows[r].y; },
        [=] __device__(uint64_t r, unsigned long long e, unsigned long long) { st[r] = e; }, mr->d_small, "kv_unpack_weighted_scan", 8.0 * R);
    ProfScope ps(pool, "kv_unpack_weighted", (double) R * 8 + (double) kv->nbytes);
    kv_expand_weighted_kernel<<<grid_for(total, 256, 32), 256, 0, s>>>(rows, st, R, total, kv->vb / 4, kv->cval[0], kv->cval[1], (uint32_t *) nd);
    MR_CUDA(cudaGetLastError());
    MR_CUDA(cudaStreamSynchronize(s));  // start[] is released on return; same stream, but keep the expansion simple to reason about
  }
  pool.free(kv->data, kv->cap);
  kv->data = nd;
  kv->cap = kv->nbytes ? kv->nbytes : 16;
  kv->packed = kv->cval_on = kv->weighted = false;
  kv->nrows = 0;
  kv->cval[0] = kv->cval[1] = 0;
}

static void kv_unpack(mrgpu_mr *mr, DevKV *kv)
{
  if (!kv || !kv->packed) return;
  if (kv->weighted) {
    kv_unpack_weighted(mr, kv);
    return;
  }
  if (kv->nkv) {
    Pool &pool = mr->pool;
    uint8_t *nd = (uint8_t *) pool.alloc(kv->nbytes);
    kv_unpack_rows(mr, kv, 0, kv->nkv, nd);
    pool.free(kv->data, kv->cap);
    kv->data = nd;
    kv->cap = kv->nbytes;
  }
  kv->packed = kv->cval_on = false;
  kv->cval[0] = kv->cval[1] = 0;
}

__global__ void iota_offsets_kernel(uint64_t *offs, uint64_t n_plus_1, uint64_t recsize)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_plus_1) offs[i] = i * recsize;
}

__global__ void affine_offsets_kernel(uint64_t *o, uint64_t cnt, uint64_t base, uint64_t rs)
{
  uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (j < cnt) o[j] = base + j * rs;
}

// give a uniform KV an explicit offset index so that records of another shape can be appended
static void kv_make_general(mrgpu_mr *mr, DevKV *kv)
{
  if (!kv->uniform || kv->nkv == 0) {
    if (kv->nkv == 0) {
      kv->uniform = false;
      kv->reserve_offs(1);
      MR_CUDA(cudaMemsetAsync(kv->offs, 0, 8, mr->pool.stream));
    }
    return;
  }
  kv->reserve_offs(kv->nkv + 1);
  iota_offsets_kernel<<<(unsigned) ((kv->nkv + 1 + 255) / 256), 256, 0, mr->pool.stream>>>(kv->offs, kv->nkv + 1, kv->recsize);
  MR_CUDA(cudaGetLastError());
  kv->uniform = false;
  kv->h_offs_valid = false;
}

// Prepare `kv` to receive nnew uniform records (kb, vb); returns the byte offset where they go.
// Keeps the KV uniform when possible, otherwise extends the offset index.
static uint64_t kv_append_uniform(mrgpu_mr *mr, DevKV *kv, uint32_t kb, uint32_t vb, uint64_t nnew)
{
  uint32_t rs = kv_rec_size(mr->al, kb, vb);
  if (rs > mr->pagesize) MR_FAIL(MRGPU_ELIMIT, "Single key/value pair exceeds page size");
  uint64_t base = kv->nbytes;
  bool fits = kv->nkv == 0 ? true : (kv->uniform && kv->kb == kb && kv->vb == vb);
  if (kv->nkv == 0 && !kv->uniform && nnew) kv->uniform = true;  // empty general KV can become uniform again
  if (fits) {
    kv->uniform = true;
    kv->kb = kb;
    kv->vb = vb;
    kv->recsize = rs;
  } else {
    kv_make_general(mr, kv);
    kv->reserve_offs(kv->nkv + nnew + 1);
    uint64_t *o = kv->offs + kv->nkv;
    uint64_t cnt = nnew + 1;
    affine_offsets_kernel<<<(unsigned) ((cnt + 255) / 256), 256, 0, mr->pool.stream>>>(o, cnt, base, rs);
    MR_CUDA(cudaGetLastError());
  }
  kv->reserve_bytes(base + nnew * rs);
  kv->h_offs_valid = false;
  return base;
}

static void kv_commit_uniform(DevKV *kv, uint32_t kb, uint32_t vb, uint32_t rs, uint64_t nnew)
{
  kv->nbytes += nnew * rs;
  kv->nkv += nnew;
  kv->ksize += nnew * kb;
  kv->vsize += nnew * vb;
}

// Host input.  Pinned (page-locked) host memory is read by the kernel directly over PCIe through its device
// alias: the transfer and the emit are one pass and no staging buffer exists.  Pageable memory has no device
// alias and is staged with one cudaMemcpyAsync.  Either way the transfer happens inside the call.
template <typename T>
static const T *stage_input(mrgpu_mr *mr, const T *p, uint64_t n, Buf<T> &tmp, bool allow_zero_copy = true)
{
  if (n == 0) return p;
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) cudaGetLastError();
  if (e == cudaSuccess && (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged)) return p;
  if (e == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer && allow_zero_copy) return (const T *) at.devicePointer;
  tmp.alloc(mr->pool, n);
  MR_CUDA(cudaMemcpyAsync(tmp.p, p, n * sizeof(T), cudaMemcpyHostToDevice, mr->pool.stream));
  return tmp.p;
}

// ---------------------------------------------------------------------------------------------
// IntCount emit: cpu/IntCount.cpp:177-180   kv->add(text+i*4, 4, (char*)&one, 4)

// default alignment: the record is exactly one 16-byte vector {4, 4, key, 1}
__global__ void __launch_bounds__(256) intcount_emit16_kernel(const uint32_t *__restrict__ keys, uint64_t n, int4 *__restrict__ out)
{
  constexpr int U = 4;
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x * U;
  for (uint64_t base = (uint64_t) blockIdx.x * blockDim.x * U + threadIdx.x; base < n; base += stride) {
    uint32_t k[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      uint64_t i = base + (uint64_t) u * blockDim.x;
      k[u] = (i < n) ? __ldg(keys + i) : 0;
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      uint64_t i = base + (uint64_t) u * blockDim.x;
      if (i < n) st_stream(out + i, make_int4(4, 4, (int) k[u], 1));
    }
  }
}

// header-less KV: the emit is a copy of the key column (16 bytes per thread per step, four steps in flight)
__global__ void __launch_bounds__(256) copy_words_kernel(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, uint64_t n)
{
  constexpr int U = 4;
  const bool vec = (((uintptr_t) src | (uintptr_t) dst) & 15) == 0;
  const uint64_t nv = vec ? n / 4 : 0;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t base = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; base < nv; base += stride * U) {
    int4 r[U];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (base + u * stride < nv) r[u] = ld_stream((const int4 *) src + base + u * stride);
#pragma unroll
    for (int u = 0; u < U; u++)
      if (base + u * stride < nv) st_stream((int4 *) dst + base + u * stride, r[u]);
  }
  for (uint64_t i = nv * 4 + (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
}

// any alignment: word-wise stores, padding zeroed
__global__ void intcount_emit_generic_kernel(const uint32_t *__restrict__ keys, uint64_t n, uint8_t *out, Align al, uint32_t rs)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t *r = (uint32_t *) (out + i * rs);
  for (uint32_t w = 0; w < rs / 4; w++) r[w] = 0;
  r[0] = 4;
  r[1] = 4;
  r[kv_key_off(al) / 4] = keys[i];
  r[kv_val_off(al, 4) / 4] = 1;
}

extern "C" int mrgpu_map_intcount(mrgpu_mr *mr, const uint32_t *keys, uint64_t n, int addflag, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (n && !keys) MR_FAIL(MRGPU_EINVAL, "null input");
  kv_begin_map(mr, addflag, true);
  DevKV *kv = mr->kv;
  Buf<uint32_t> tmp;
  // header-less KV from host memory: the key column is filled by DMA (55.6 GB/s measured against 51.7 GB/s for a
  // kernel reading pinned memory through its device alias), in chunks so that counting overlaps the transfer
  if (n && !is_device_ptr(keys) && kv_can_pack(mr, 4, 4) &&
      (kv->nkv == 0 || (kv->packed && !kv->weighted && kv->cval_on && kv->kb == 4 && kv->vb == 4 && kv->cval[0] == 1))) {
    if (kv->nkv == 0) {
      kv->uniform = kv->packed = kv->cval_on = true;
      kv->kb = kv->vb = 4;
      kv->recsize = 16;
      kv->cval[0] = 1;
      kv->cval[1] = 0;
    }
    kv->reserve_bytes((kv->nkv + n) * 4);
    const bool count = kv->nkv == 0 && mr->ahead && (mr->nranks == 1 || mr->combine) && n >= (1u << 16);
    emit_from_host(mr, kv, keys, n, count);
    kv_commit_uniform(kv, 4, 4, 16, n);
    kv->h_offs_valid = false;
    spool_after_map(mr);
    if (nkv) {
      sync(mr);
      *nkv = global_count(mr, kv_total(mr), G_KV, false);
    }
    return MRGPU_OK;
  }
  const uint32_t *dk = stage_input(mr, keys, n, tmp);
  // header-less: every pair is {4, 4, key, 1}, so the KV is the key column plus one shared value
  const bool pack = kv_can_pack(mr, 4, 4) &&
                    (kv->nkv == 0 || (kv->packed && !kv->weighted && kv->cval_on && kv->kb == 4 && kv->vb == 4 && kv->cval[0] == 1));
  if (pack) {
    if (kv->nkv == 0) {
      kv->uniform = kv->packed = kv->cval_on = true;
      kv->kb = kv->vb = 4;
      kv->recsize = 16;
      kv->cval[0] = 1;
      kv->cval[1] = 0;
    }
    kv->reserve_bytes((kv->nkv + n) * 4);
    // The pass that writes the key column can do the first pass of the op that follows in every pipeline: with
    // one rank aggregate() is a no-op and convert() counts keys; with several, aggregate() needs every pair's
    // destination and the per-tile counts (1.89 ms fused against 1.38 + 1.82 apart at two ranks; a first, word-
    // per-lane version of the fused pass had been slower than the two apart: 4.19 against 3.87 ms).
    const bool fresh = kv->nkv == 0 && mr->ahead && n >= (1u << 16) && ((((uintptr_t) dk) | ((uintptr_t) kv->data)) & 15) == 0;
    // (with several ranks and the combiner on, aggregate() starts from the same key counts: ops_aggregate.cuh)
    if (fresh && (mr->nranks == 1 || mr->combine)) {
      emit_and_count(mr, kv, dk, n);
    } else if (fresh && mr->nranks > 1 && mr->nranks <= 256 && !getenv("MRGPU_NO_EMIT_PLAN")) {
      emit_and_plan(mr, kv, dk, n);
    } else if (n) {
      ProfScope ps(mr->pool, "map_intcount_emit", (double) n * (4 + 8 + 4 + 4), 1, (double) n * 8);
      copy_words_kernel<<<grid_for(n, 256 * 16, 16), 256, 0, mr->pool.stream>>>(dk, (uint32_t *) kv->data + kv->nkv, n);
      MR_CUDA(cudaGetLastError());
    }
    kv_commit_uniform(kv, 4, 4, 16, n);
    kv->h_offs_valid = false;
    tmp.release();
    spool_after_map(mr);
    if (nkv) {
      sync(mr);
      *nkv = global_count(mr, kv_total(mr), G_KV, false);
    }
    return MRGPU_OK;
  }
  kv_unpack(mr, kv);
  uint64_t base = kv_append_uniform(mr, kv, 4, 4, n);
  uint32_t rs = kv_rec_size(mr->al, 4, 4);
  if (n) {
    ProfScope ps(mr->pool, "map_intcount_emit", (double) n * (4 + 8 + 4 + 4));
    if (rs == 16 && kv_key_off(mr->al) == 8 && (base & 15) == 0)
      intcount_emit16_kernel<<<grid_for(n, 256 * 4, 16), 256, 0, mr->pool.stream>>>(dk, n, (int4 *) (kv->data + base));
    else
      intcount_emit_generic_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, mr->pool.stream>>>(dk, n, kv->data + base, mr->al, rs);
    MR_CUDA(cudaGetLastError());
  }
  kv_commit_uniform(kv, 4, 4, rs, n);
  tmp.release();
  spool_after_map(mr);
  if (nkv) {
    sync(mr);
    *nkv = global_count(mr, kv_total(mr), G_KV, false);
  }
  API_END(mr)
}

// ---------------------------------------------------------------------------------------------
// map(mr, edge_to_vertices | edge_to_vertex | invert)   src/mapreduce.cpp:1560-1642

__device__ __forceinline__ uint64_t rec_off(const uint64_t *offs, uint64_t recsize, uint64_t i) { return offs ? offs[i] : i * recsize; }

// key = first (and second) u64 of the source key, value = NULL: oink/map_edge_to_vertices.cpp:14-20
__global__ void __launch_bounds__(256)
edge_to_vertices_kernel(const uint8_t *__restrict__ src, const uint64_t *__restrict__ offs, uint64_t recsize, uint32_t koff, uint64_t n,
                        int both, uint8_t *__restrict__ dst, uint32_t out_rs, uint32_t out_koff)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t *r = src + rec_off(offs, recsize, i) + koff;
  uint32_t w[4];
  w[0] = ((const uint32_t *) r)[0];
  w[1] = ((const uint32_t *) r)[1];
  if (both) {
    w[2] = ((const uint32_t *) r)[2];
    w[3] = ((const uint32_t *) r)[3];
  }
  int nout = both ? 2 : 1;
  for (int e = 0; e < nout; e++) {
    uint32_t *o = (uint32_t *) (dst + (i * nout + e) * out_rs);
    if (out_koff == 0) {  // header-less row
      *(uint2 *) o = make_uint2(w[2 * e], w[2 * e + 1]);
    } else if (out_rs == 16 && out_koff == 8) {
      *(int4 *) o = make_int4(8, 0, (int) w[2 * e], (int) w[2 * e + 1]);
    } else {
      for (uint32_t q = 0; q < out_rs / 4; q++) o[q] = 0;
      o[0] = 8;
      o[1] = 0;
      o[out_koff / 4] = w[2 * e];
      o[out_koff / 4 + 1] = w[2 * e + 1];
    }
  }
}

// invert: key <-> value, oink/map_invert.cpp:11-15; byte-wise copy, one thread per record
__global__ void invert_kernel(const uint8_t *__restrict__ src, const uint64_t *__restrict__ soffs, uint64_t srs, uint64_t n,
                              uint8_t *__restrict__ dst, const uint64_t *__restrict__ doffs, uint64_t drs, Align al)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t *r = src + rec_off(soffs, srs, i);
  uint32_t kb = ((const uint32_t *) r)[0], vb = ((const uint32_t *) r)[1];
  const uint8_t *k = r + kv_key_off(al), *v = r + kv_val_off(al, kb);
  uint8_t *o = dst + rec_off(doffs, drs, i);
  uint32_t sz = kv_rec_size(al, vb, kb);
  for (uint32_t q = 0; q < sz / 4; q++) ((uint32_t *) o)[q] = 0;
  ((uint32_t *) o)[0] = vb;
  ((uint32_t *) o)[1] = kb;
  uint8_t *ok = o + kv_key_off(al), *ov = o + kv_val_off(al, vb);
  for (uint32_t q = 0; q < vb; q++) ok[q] = v[q];
  for (uint32_t q = 0; q < kb; q++) ov[q] = k[q];
}

extern "C" int mrgpu_map_mr(mrgpu_mr *mr, mrgpu_mr *srcmr, int functor, int addflag, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!srcmr) MR_FAIL(MRGPU_EINVAL, "null source");
  if (!srcmr->kv) MR_FAIL(MRGPU_ENOKV, "MapReduce passed to map() does not have KeyValue pairs");
  spool_make_resident(srcmr);  // the functors walk the source in HBM
  if (srcmr->device != mr->device) MR_FAIL(MRGPU_EINVAL, "source and destination live on different devices");
  freeze(mr);
  freeze(srcmr);
  // src == dst: the reference detaches the old KV, builds the new one, then deletes the old (:1584-1602, 1634)
  DevKV *src = srcmr->kv;
  // edge_to_vertex(es) read keys only and take header-less rows as they are; the other functors read records
  const bool e2v = functor == MRGPU_MAP_EDGE_TO_VERTICES || functor == MRGPU_MAP_EDGE_TO_VERTEX;
  if (!(e2v && src->packed && !src->weighted)) kv_unpack(srcmr, src);
  order_after(mr, srcmr);  // the kernels below run on mr's stream and read what srcmr's stream wrote
  bool self = (srcmr == mr);
  if (self) mr->kv = nullptr;
  try {
    kv_begin_map(mr, self ? 0 : addflag, e2v);
    DevKV *kv = mr->kv;
    const Align al = mr->al;
    const uint64_t n = src->nkv;
    cudaStream_t s = mr->pool.stream;
    if (e2v) {
      int both = functor == MRGPU_MAP_EDGE_TO_VERTICES;
      uint32_t need = both ? 16 : 8;
      if (n && src->uniform && src->kb < need) MR_FAIL(MRGPU_EINVAL, "edge_to_vertices: source keys are shorter than an EDGE");
      if (n && (srcmr->al.k % 4)) MR_FAIL(MRGPU_EINVAL, "edge_to_vertices needs keyalign >= 4");
      uint64_t nout = n * (both ? 2 : 1);
      const bool rows = kv_rows_ok(mr, kv, 8);  // the vertices as header-less 8-byte rows
      if (!rows) kv_unpack(mr, kv);
      uint32_t rs = kv_rec_size(al, 8, 0);
      uint64_t base;
      if (rows) {
        kv_rows_begin(kv, 8, rs);
        kv->reserve_bytes((kv->nkv + nout) * 8);
        kv->h_offs_valid = false;
        base = kv->nkv * 8;
      } else
        base = kv_append_uniform(mr, kv, 8, 0, nout);
      if (n) {
        const RowLayout sl = src->uniform ? kv_layout(srcmr, src) : RowLayout{0, kv_key_off(srcmr->al), 0};
        ProfScope ps(mr->pool, "map_edge_to_vertices", (double) n * (src->uniform ? sl.rs : 8 + src->kb + src->vb) + (double) nout * (rows ? 8 : rs));
        edge_to_vertices_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(src->data, src->uniform ? nullptr : src->offs, sl.rs, sl.koff, n, both,
                                                                           kv->data + base, rows ? 8 : rs, rows ? 0 : kv_key_off(al));
        MR_CUDA(cudaGetLastError());
      }
      kv_commit_uniform(kv, 8, 0, rs, nout);
    } else if (functor == MRGPU_MAP_INVERT) {
      if (srcmr->al.k != al.k || srcmr->al.v != al.v) MR_FAIL(MRGPU_EINVAL, "invert: alignments differ");
      if (src->uniform || n == 0) {
        uint64_t base = kv_append_uniform(mr, kv, src->vb, src->kb, n);
        uint32_t rs = kv_rec_size(al, src->vb, src->kb);
        if (n) {
          invert_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(src->data, nullptr, src->recsize, n, kv->data + base, nullptr, rs, al);
          MR_CUDA(cudaGetLastError());
        }
        kv_commit_uniform(kv, src->vb, src->kb, rs, n);
      } else {
        // sizes -> scan -> write
        kv_make_general(mr, kv);
        kv->reserve_offs(kv->nkv + n + 1);
        const uint8_t *sd = src->data;
        const uint64_t *so = src->offs;
        uint64_t *doffs = kv->offs + kv->nkv;
        uint64_t base = kv->nbytes;
        scan_apply(
            mr->pool, n,
            [=] __device__(uint64_t i) -> unsigned long long {
              const uint32_t *r = (const uint32_t *) (sd + so[i]);
              return kv_rec_size(al, r[1], r[0]);
            },
            [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { doffs[i] = base + e; }, mr->d_small);
        fetch_small(mr, 1);
        uint64_t total = mr->h_small[0];
        uint64_t end = base + total;
        MR_CUDA(cudaMemcpyAsync(doffs + n, &end, 8, cudaMemcpyHostToDevice, s));
        kv->reserve_bytes(end);
        invert_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(src->data, src->offs, 0, n, kv->data, doffs, 0, al);
        MR_CUDA(cudaGetLastError());
        sync(mr);
        kv->nbytes = end;
        kv->nkv += n;
        kv->ksize += src->vsize;
        kv->vsize += src->ksize;
        kv->h_offs_valid = false;
      }
    } else
      MR_FAIL(MRGPU_EINVAL, "unknown map functor");
  } catch (...) {
    if (self) delete src;
    throw;
  }
  if (self) delete src;
  // another stream read srcmr's store: it must be done before srcmr (whose pool reuses memory in ITS stream's order)
  // is used again
  spool_after_map(mr);
  if (nkv || mr->pool.stream != srcmr->pool.stream) sync(mr);
  if (nkv) *nkv = global_count(mr, kv_total(mr), G_KV, false);
  API_END(mr)
}

// ---------------------------------------------------------------------------------------------
// R-MAT edge generator: oink/map_rmat_generate.cpp:14-67 with glibc drand48 restated
//   X' = (0x5DEECE66D X + 0xB) mod 2^48, value = X' / 2^48.
// The reference draws sequentially from one per-rank stream; edge m consumes draws
// [m*D, (m+1)*D) with D = nlevels (fraction == 0) or 5*nlevels, so thread m jumps the LCG ahead by m*D
// steps in O(log) multiplications and then replays exactly the reference's arithmetic.  Double adds and
// compares only when fraction == 0; for fraction > 0 every multiply/add goes through the round-to-nearest
// intrinsics so that no FMA contraction changes a bit.

#include "rmat_lcg.cuh"  // LCG constants, lcg_affine / lcg_jump, rmat_threshold (host-testable: tests/cpp/rmat_host_test.cu)

__device__ __forceinline__ double lcg_next(unsigned long long &x)
{
  x = (LCG_A * x + LCG_C) & LCG_M;
  return (double) (long long) x * (1.0 / 281474976710656.0);
}

struct RmatParams {
  unsigned long long order;
  int nlevels;
  double a, b, c, d, fraction;
  // fraction == 0: the quadrant tests "drand48() < a", "< a+b", "< a+b+c" on the raw 48-bit state.  drand48() is
  // X / 2^48 exactly (X < 2^48 fits a double's mantissa) and scaling by a power of two is exact, so
  // X / 2^48 < t  <=>  X < t * 2^48  <=>  X < ceil(t * 2^48) for an integer X: the same decisions without the
  // int-to-double conversion and the double compares.  ta/tab/tabc are those ceilings, clamped to [0, 2^48].
  unsigned long long ta, tab, tabc;
  // one grid step ahead in the stream: thread m handles edges m, m + T, m + 2T, ... (T = threads in the grid); its
  // first edge costs a full O(log) jump, every later one this single affine step of T * D draws
  unsigned long long sa, sc;
};

template <bool INT>
__global__ void __launch_bounds__(256)
rmat_generate_kernel(RmatParams p, unsigned long long state0, uint64_t ngenerate, uint8_t *out, uint32_t rs, uint32_t koff, int fast)
{
  const uint64_t T = (uint64_t) gridDim.x * blockDim.x;
  uint64_t m = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= ngenerate) return;
  const int D = p.fraction > 0.0 ? 5 * p.nlevels : p.nlevels;
  unsigned long long xs = state0;
  lcg_jump(m * (unsigned long long) D, xs);
  for (; m < ngenerate; m += T, xs = (p.sa * xs + p.sc) & LCG_M) {
    unsigned long long x = xs, i = 0, j = 0;
    if (INT) {
      // level l decides bit nlevels-1-l of i and j: shift them in, most significant first
      for (int l = 0; l < p.nlevels; l++) {
        x = (LCG_A * x + LCG_C) & LCG_M;
        const bool bi = x >= p.tab, bj = (x >= p.ta && x < p.tab) || x >= p.tabc;
        i = (i << 1) | (unsigned long long) bi;
        j = (j << 1) | (unsigned long long) bj;
      }
    } else {
      unsigned long long delta = p.order >> 1;
      double a1 = p.a, b1 = p.b, c1 = p.c, d1 = p.d;
      for (int l = 0; l < p.nlevels; l++) {
        double rn = lcg_next(x);
        double ab = __dadd_rn(a1, b1);
        if (rn < a1) {
        } else if (rn < ab) {
          j += delta;
        } else if (rn < __dadd_rn(ab, c1)) {
          i += delta;
        } else {
          i += delta;
          j += delta;
        }
        delta >>= 1;
        if (p.fraction > 0.0) {
          // a1 += a1*fraction * (drand48() - 0.5): ((a1*fraction) * (r-0.5)) then add, left to right as in C
          a1 = __dadd_rn(a1, __dmul_rn(__dmul_rn(a1, p.fraction), __dadd_rn(lcg_next(x), -0.5)));
          b1 = __dadd_rn(b1, __dmul_rn(__dmul_rn(b1, p.fraction), __dadd_rn(lcg_next(x), -0.5)));
          c1 = __dadd_rn(c1, __dmul_rn(__dmul_rn(c1, p.fraction), __dadd_rn(lcg_next(x), -0.5)));
          d1 = __dadd_rn(d1, __dmul_rn(__dmul_rn(d1, p.fraction), __dadd_rn(lcg_next(x), -0.5)));
          double total = __dadd_rn(__dadd_rn(__dadd_rn(a1, b1), c1), d1);
          a1 = __ddiv_rn(a1, total);
          b1 = __ddiv_rn(b1, total);
          c1 = __ddiv_rn(c1, total);
          d1 = __ddiv_rn(d1, total);
        }
      }
    }
    uint32_t *r = (uint32_t *) (out + m * rs);
    if (fast == 2) {
      // header-less row = the 16 key bytes
      *(ulonglong2 *) r = make_ulonglong2(i, j);
    } else if (fast) {
      // 24-byte record = three 8-byte stores (records are 8-byte aligned)
      ((uint2 *) r)[0] = make_uint2(16, 0);
      ((unsigned long long *) r)[1] = i;
      ((unsigned long long *) r)[2] = j;
    } else {
      for (uint32_t q = 0; q < rs / 4; q++) r[q] = 0;
      r[0] = 16;
      r[1] = 0;
      r[koff / 4] = (uint32_t) i;
      r[koff / 4 + 1] = (uint32_t) (i >> 32);
      r[koff / 4 + 2] = (uint32_t) j;
      r[koff / 4 + 3] = (uint32_t) (j >> 32);
    }
  }
}

extern "C" int mrgpu_map_rmat(mrgpu_mr *mr, uint64_t ngenerate, int nlevels, double a, double b, double c, double d, double fraction,
                              uint64_t *state48, int addflag, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!state48) MR_FAIL(MRGPU_EINVAL, "null state");
  if (nlevels < 1 || nlevels > 62) MR_FAIL(MRGPU_EINVAL, "Illegal rmat command");
  if (fraction >= 1.0) MR_FAIL(MRGPU_EINVAL, "RMAT fraction must be < 1");
  kv_begin_map(mr, addflag, true);
  DevKV *kv = mr->kv;
  if (mr->al.k % 4) MR_FAIL(MRGPU_EINVAL, "rmat_generate needs keyalign >= 4");
  // header-less rows (the 16 key bytes; store.cuh) when the KV is empty or holds such rows already: what reduce(cull)
  // leaves behind in R-MAT's loop
  const bool rows = kv_rows_ok(mr, kv, 16);
  if (!rows) kv_unpack(mr, kv);
  uint64_t base;
  uint32_t rs = kv_rec_size(mr->al, 16, 0);
  if (rows) {
    kv_rows_begin(kv, 16, rs);
    kv->reserve_bytes((kv->nkv + ngenerate) * 16);
    kv->h_offs_valid = false;
    base = kv->nkv * 16;
  } else
    base = kv_append_uniform(mr, kv, 16, 0, ngenerate);
  RmatParams p{1ull << nlevels, nlevels, a, b, c, d, fraction};
  if (ngenerate) {
    const int fast = rows ? 2 : (rs == 24 && kv_key_off(mr->al) == 8 && (base % 8) == 0);
    const uint32_t out_rs = rows ? 16 : rs;
    const unsigned grid = grid_for(ngenerate, 256, 8);
    const unsigned long long Dd = fraction > 0.0 ? 5ull * nlevels : (unsigned long long) nlevels;
    lcg_affine((unsigned long long) grid * 256 * Dd, p.sa, p.sc);
    // the sums as the reference forms them, in double: a1+b1, then (a1+b1)+c1 (oink/map_rmat_generate.cpp:42-50)
    const double ab = a + b, abc = ab + c;
    p.ta = rmat_threshold(a), p.tab = rmat_threshold(ab), p.tabc = rmat_threshold(abc);
    // the integer form assumes a <= a+b <= a+b+c, as for any sane parameters; anything else takes the double path
    const bool ints = !(fraction > 0.0) && a <= ab && ab <= abc;
    ProfScope ps(mr->pool, "map_rmat_generate", (double) ngenerate * out_rs);
    if (ints)
      rmat_generate_kernel<true><<<grid, 256, 0, mr->pool.stream>>>(p, *state48 & LCG_M, ngenerate, kv->data + base, out_rs, kv_key_off(mr->al), fast);
    else
      rmat_generate_kernel<false><<<grid, 256, 0, mr->pool.stream>>>(p, *state48 & LCG_M, ngenerate, kv->data + base, out_rs, kv_key_off(mr->al), fast);
    MR_CUDA(cudaGetLastError());
  }
  kv_commit_uniform(kv, 16, 0, rs, ngenerate);
  unsigned long long x = *state48 & LCG_M;
  unsigned long long D = fraction > 0.0 ? 5ull * nlevels : (unsigned long long) nlevels;
  lcg_jump(ngenerate * D, x);
  *state48 = x;
  spool_after_map(mr);
  if (nkv) {
    sync(mr);
    *nkv = global_count(mr, kv_total(mr), G_KV, false);
  }
  API_END(mr)
}

// ---------------------------------------------------------------------------------------------
// MapReduce::add(MapReduce*) src/mapreduce.cpp:348-374: append the other object's KV to mine

__global__ void shift_offsets_kernel(uint64_t *dst, const uint64_t *src, uint64_t n_plus_1, uint64_t delta)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_plus_1) dst[i] = src[i] + delta;
}

// append the pairs of record-stream KV `o` to record-stream KV `kv` (both on this device; the copy runs on mr's stream)
static void kv_concat(mrgpu_mr *mr, DevKV *kv, const DevKV *o)
{
  cudaStream_t s = mr->pool.stream;
  if (!o->nkv) return;
  kv->sorted_flag = 0;
  uint64_t base = kv->nbytes;
  bool fits = o->uniform && (kv->nkv == 0 || (kv->uniform && kv->kb == o->kb && kv->vb == o->vb));
  if (fits) {
    kv_append_uniform(mr, kv, o->kb, o->vb, o->nkv);
  } else {
    kv_make_general(mr, kv);
    kv->reserve_offs(kv->nkv + o->nkv + 1);
    if (o->uniform) {
      kv_append_uniform(mr, kv, o->kb, o->vb, o->nkv);  // general branch fills the offsets
    } else {
      shift_offsets_kernel<<<(unsigned) ((o->nkv + 1 + 255) / 256), 256, 0, s>>>(kv->offs + kv->nkv, o->offs, o->nkv + 1, base);
      MR_CUDA(cudaGetLastError());
      kv->reserve_bytes(base + o->nbytes);
    }
  }
  MR_CUDA(cudaMemcpyAsync(kv->data + base, o->data, o->nbytes, cudaMemcpyDeviceToDevice, s));
  kv->nbytes += o->nbytes;
  kv->nkv += o->nkv;
  kv->ksize += o->ksize;
  kv->vsize += o->vsize;
  kv->h_offs_valid = false;
}

extern "C" int mrgpu_add(mrgpu_mr *mr, mrgpu_mr *other, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!other || !other->kv) MR_FAIL(MRGPU_ENOKV, "MapReduce passed to add() does not have KeyValue pairs");
  if (mr == other) MR_FAIL(MRGPU_EINVAL, "Cannot add to self");
  if (mr->kmv) MR_FAIL(MRGPU_EINVAL, "Cannot add to a MapReduce with KeyMultiValue pairs");  // :351-353
  mr->g_what = G_NONE;
  freeze(mr);
  freeze(other);
  if (mr->al.k != other->al.k || mr->al.v != other->al.v) MR_FAIL(MRGPU_EINVAL, "add: alignments differ");
  if (mr->device != other->device) MR_FAIL(MRGPU_EINVAL, "add: the two MapReduce objects live on different devices");
  if (!mr->kv) mr->kv = new DevKV(&mr->pool);
  spool_make_resident(mr);
  spool_make_resident(other);
  DevKV *kv = mr->kv, *o = other->kv;
  kv_unpack(mr, kv);
  kv_unpack(other, o);   // enqueued on other's stream ...
  order_after(mr, other);  // ... and read below on mine
  kv_concat(mr, kv, o);
  sync(mr);
  if (nkv) *nkv = global_count(mr, kv->nkv, G_KV, false);
  API_END(mr)
}

// reduce(): KeyMultiValue -> KeyValue with the stock callbacks as device functors.
//
// Reference: MapReduce::reduce src/mapreduce.cpp:1769-1867 walks the KMV pages, decodes each record and
// calls myreduce(key,keybytes,multivalue,nvalues,valuebytes,kv,ptr); the callback emits with kv->add.
//   count  oink/reduce_count.cpp:13-20   emit (key, int nvalues_total)      [multi-block aware]
//   sum    examples/wordfreq.cpp:137-141 emit (key, int nvalues)
//   cull   oink/reduce_cull.cpp:12-20    emit (key, first value)
// Every one of them emits exactly one pair per KMV pair and needs only the KMV header (+ first value), so
// the kernel is one thread per KMV pair reading the header through offs[u] and writing one KV record; the
// multivalue bytes are never touched (less HBM traffic than the algorithmic KMV size).  Included by mrgpu.cu.
#pragma once

__device__ __forceinline__ void write_kv_record(uint8_t *dst, Align al, uint32_t kb, const uint8_t *key, uint32_t vb, const uint8_t *val)
{
  uint32_t ko = kv_key_off(al), vo = kv_val_off(al, kb), end = kv_rec_size(al, kb, vb);
  ((uint32_t *) dst)[0] = kb;
  ((uint32_t *) dst)[1] = vb;
  for (uint32_t q = 8; q < ko; q++) dst[q] = 0;
  for (uint32_t q = 0; q < kb; q++) dst[ko + q] = key[q];
  for (uint32_t q = ko + kb; q < vo; q++) dst[q] = 0;
  for (uint32_t q = 0; q < vb; q++) dst[vo + q] = val[q];
  for (uint32_t q = vo + vb; q < end; q++) dst[q] = 0;
}

struct KmvHead {
  uint32_t nv, kb, mvb;
  const uint8_t *key, *sizes, *values;
};

__device__ __forceinline__ KmvHead kmv_head(const uint8_t *rec, Align al)
{
  KmvHead h;
  const uint32_t *w = (const uint32_t *) rec;
  h.nv = w[0];
  h.kb = w[1];
  h.mvb = w[2];
  h.sizes = rec + 12;
  h.key = rec + kmv_key_off(al, h.nv);
  h.values = rec + kmv_val_off(al, h.nv, h.kb);
  return h;
}

// output key / value sizes of functor f for a KMV pair:
//   count / sum / histo -> (key, int nvalues);  cull -> (key, first value);  nonzero -> (first half of the key, NULL);
//   degree -> (int nvalues, NULL);  postings -> (key, every value with its terminating NUL turned into ' ', then one NUL)
__device__ __forceinline__ uint32_t reduce_out_vb(int functor, const KmvHead &h)
{
  if (functor == MRGPU_REDUCE_CULL) return h.nv ? *(const uint32_t *) h.sizes : 0;
  if (functor == MRGPU_REDUCE_NONZERO || functor == MRGPU_REDUCE_DEGREE) return 0;
  if (functor == MRGPU_REDUCE_POSTINGS) return h.mvb + 1;
  return 4;
}
__device__ __forceinline__ uint32_t reduce_out_kb(int functor, const KmvHead &h)
{
  if (functor == MRGPU_REDUCE_NONZERO) return h.kb / 2;
  if (functor == MRGPU_REDUCE_DEGREE) return 4;
  return h.kb;
}

// fast path: 4-byte keys, default alignment, count/sum -> one 16-byte store per pair
__global__ void __launch_bounds__(256)
reduce_count16_kernel(const uint8_t *__restrict__ kmv, const uint64_t *__restrict__ offs, uint64_t U, int4 *__restrict__ out)
{
  uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  const uint32_t *w = (const uint32_t *) (kmv + offs[u]);
  uint32_t nv = w[0];
  uint32_t key = w[3 + nv];  // key follows the nv value sizes (kalign 4)
  st_stream(out + u, make_int4(4, 4, (int) key, (int) nv));
}

// factored KMV (store.cuh): pair u is (fkey[u], fcnt[u] values, all equal to the shared value)
__global__ void __launch_bounds__(256)
reduce_factored16_kernel(const uint32_t *__restrict__ fkey, const uint32_t *__restrict__ fcnt, uint64_t U, int4 *__restrict__ out)
{
  uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  st_stream(out + u, make_int4(4, 4, (int) __ldg(fkey + u), (int) __ldg(fcnt + u)));
}

// default alignment (record = 8 + kb + vb, all words): one thread per output word, coalesced both ways
__global__ void __launch_bounds__(256)
reduce_factored_words_kernel(const uint32_t *__restrict__ fkey, const uint32_t *__restrict__ fcnt, uint64_t U, int functor, uint32_t kw, uint32_t vw,
                             uint32_t cv0, uint32_t cv1, uint32_t *__restrict__ out)
{
  // count / sum emit (key, int count); cull emits (key, the shared value)
  const uint32_t ow = functor == MRGPU_REDUCE_CULL ? vw : 1, rw = 2 + kw + ow;
  const uint64_t total = U * rw, stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t t = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const uint64_t u = t / rw;
    const uint32_t w = (uint32_t) (t - u * rw);
    uint32_t x;
    if (w == 0) x = kw * 4;
    else if (w == 1) x = ow * 4;
    else if (w < 2 + kw) x = __ldg(fkey + u * kw + (w - 2));
    else if (functor == MRGPU_REDUCE_CULL) x = (w - 2 - kw) ? cv1 : cv0;
    else x = __ldg(fcnt + u);
    out[t] = x;
  }
}

__global__ void __launch_bounds__(256)
reduce_factored_kernel(const uint32_t *__restrict__ fkey, const uint32_t *__restrict__ fcnt, uint64_t U, int functor, Align al, uint32_t kb,
                       uint32_t vb, uint32_t cv0, uint32_t cv1, uint8_t *out, uint64_t ors)
{
  uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  const uint8_t *key = (const uint8_t *) (fkey + u * (kb / 4));
  const uint32_t cnt = fcnt[u], cv[2] = {cv0, cv1};
  if (functor == MRGPU_REDUCE_CULL) write_kv_record(out + u * ors, al, kb, key, vb, (const uint8_t *) cv);
  else write_kv_record(out + u * ors, al, kb, key, 4, (const uint8_t *) &cnt);
}

__global__ void __launch_bounds__(256)
reduce_generic_kernel(const uint8_t *__restrict__ kmv, const uint64_t *__restrict__ offs, uint64_t U, int functor, Align al, uint8_t *out,
                      const uint64_t *__restrict__ ooffs, uint64_t ors)
{
  uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  KmvHead h = kmv_head(kmv + offs[u], al);
  uint8_t *dst = out + (ooffs ? ooffs[u] : u * ors);
  if (functor == MRGPU_REDUCE_CULL) {
    write_kv_record(dst, al, h.kb, h.key, *(const uint32_t *) h.sizes, h.values);
  } else if (functor == MRGPU_REDUCE_NONZERO) {
    write_kv_record(dst, al, h.kb / 2, h.key, 0, h.key);
  } else if (functor == MRGPU_REDUCE_DEGREE) {
    uint32_t cnt = h.nv;
    write_kv_record(dst, al, 4, (const uint8_t *) &cnt, 0, h.key);
  } else {
    uint32_t cnt = h.nv;
    write_kv_record(dst, al, h.kb, h.key, 4, (const uint8_t *) &cnt);
  }
}

// postings (cpu/InvertedIndex.cpp:335-385 prints, per KMV pair, `key \t v1 v2 ... vn \n`: each value as a C string followed
// by a space).  The values of a KMV pair are contiguous, each a string with its NUL, so the postings text is the
// multivalue with every NUL turned into a space, closed by one NUL.  Two kernels, neither of which cares how the bytes
// are spread over the pairs (Zipf: the hottest URL alone has millions of postings, tens of MB):
//   head: one thread per pair writes lengths, key, the closing NUL and the padding;
//   body: the multivalues of all pairs, end to end, are one byte space; a thread takes 16 consecutive bytes of it,
//         finds their pair with a binary search over the running sum of the multivalue sizes and copies them across.
__global__ void __launch_bounds__(256)
reduce_postings_head_kernel(const uint8_t *__restrict__ kmv, const uint64_t *__restrict__ offs, uint64_t U, Align al, uint8_t *__restrict__ out,
                            const uint64_t *__restrict__ ooffs)
{
  const uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  KmvHead h = kmv_head(kmv + offs[u], al);
  uint8_t *dst = out + ooffs[u];
  const uint32_t vb = h.mvb + 1, ko = kv_key_off(al), vo = kv_val_off(al, h.kb), end = kv_rec_size(al, h.kb, vb);
  ((uint32_t *) dst)[0] = h.kb;
  ((uint32_t *) dst)[1] = vb;
  for (uint32_t q = 8; q < ko; q++) dst[q] = 0;
  for (uint32_t q = 0; q < h.kb; q++) dst[ko + q] = h.key[q];
  for (uint32_t q = ko + h.kb; q < vo; q++) dst[q] = 0;
  for (uint32_t q = vo + h.mvb; q < end; q++) dst[q] = 0;  // the closing NUL and the padding
}

__global__ void __launch_bounds__(256)
reduce_postings_body_kernel(const uint8_t *__restrict__ kmv, const uint64_t *__restrict__ offs, uint64_t U, Align al, uint8_t *__restrict__ out,
                            const uint64_t *__restrict__ ooffs, const uint64_t *__restrict__ cum /*[U+1] multivalue bytes before pair u*/,
                            uint64_t total)
{
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x * 16;
  for (uint64_t x0 = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) * 16; x0 < total; x0 += stride) {
    uint64_t lo = 0, hi = U - 1;  // last pair whose multivalue starts at or before x0
    while (lo < hi) {
      const uint64_t mid = (lo + hi + 1) >> 1;
      if (__ldg(cum + mid) <= x0) lo = mid;
      else hi = mid - 1;
    }
    uint64_t u = lo, cnext = __ldg(cum + u + 1);
    while (cnext <= x0) u++, cnext = __ldg(cum + u + 1);  // pairs without values share a start
    KmvHead h = kmv_head(kmv + offs[u], al);
    const uint8_t *src = h.values - __ldg(cum + u);                 // src[x] = byte x of the space
    uint8_t *dst = out + ooffs[u] + kv_val_off(al, h.kb) - __ldg(cum + u);
    const uint64_t x1 = x0 + 16 < total ? x0 + 16 : total;
    for (uint64_t x = x0; x < x1; x++) {
      while (x >= cnext) {  // the 16 bytes run into the next pair(s)
        u++;
        cnext = __ldg(cum + u + 1);
        h = kmv_head(kmv + offs[u], al);
        src = h.values - __ldg(cum + u);
        dst = out + ooffs[u] + kv_val_off(al, h.kb) - __ldg(cum + u);
      }
      const uint8_t c = src[x];
      dst[x] = c ? c : (uint8_t) ' ';
    }
  }
}

// factored KMV (key words, count, shared value) -> nonzero / degree records, one thread per pair
__global__ void __launch_bounds__(256)
reduce_factored_shape_kernel(const uint32_t *__restrict__ fkey, const uint32_t *__restrict__ fcnt, uint64_t U, int functor, Align al, uint32_t kb,
                             uint8_t *out, uint64_t ors)
{
  uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  const uint8_t *key = (const uint8_t *) (fkey + u * (kb / 4));
  const uint32_t cnt = fcnt[u];
  if (functor == MRGPU_REDUCE_NONZERO) write_kv_record(out + u * ors, al, kb / 2, key, 0, key);
  else write_kv_record(out + u * ors, al, 4, (const uint8_t *) &cnt, 0, key);
}

// factored KMV with variable-length keys (convert's hash path): count / sum -> (key, int count), one thread per pair
__global__ void __launch_bounds__(256)
reduce_fvar_kernel(const uint8_t *__restrict__ vkey, const uint64_t *__restrict__ vkoff, const uint32_t *__restrict__ fcnt, uint64_t U, Align al,
                   uint8_t *__restrict__ out, const uint64_t *__restrict__ ooffs)
{
  const uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  const uint32_t cnt = fcnt[u];
  write_kv_record(out + ooffs[u], al, (uint32_t) (vkoff[u + 1] - vkoff[u]), vkey + vkoff[u], 4, (const uint8_t *) &cnt);
}

static void reduce_impl(mrgpu_mr *mr, int functor)
{
  if (functor < 0 || functor > MRGPU_REDUCE_HISTO) MR_FAIL(MRGPU_EINVAL, "unknown reduce functor");
  if (functor == MRGPU_REDUCE_HISTO) functor = MRGPU_REDUCE_COUNT;  // examples/rmat.cpp:289-295 emits (key, int nvalues): the same pair
  DevKMV *kmv = mr->kmv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t U = kmv->nkmv;
  if (kmv->fvar && !(functor == MRGPU_REDUCE_COUNT || functor == MRGPU_REDUCE_SUM)) kmv_expand(mr, kmv);
  if (functor == MRGPU_REDUCE_POSTINGS) kmv_expand(mr, kmv);  // reads the values
  DevKV *kv = new DevKV(&pool);
  try {
    if (kmv->fvar && U) {
      const uint64_t *vko = kmv->vkoff;
      kv->uniform = false;
      kv->reserve_offs(U + 1);
      uint64_t *oo = kv->offs;
      scan_apply(
          pool, U, [=] __device__(uint64_t u) -> unsigned long long { return kv_rec_size(al, (uint32_t) (vko[u + 1] - vko[u]), 4); },
          [=] __device__(uint64_t u, unsigned long long e, unsigned long long) { oo[u] = e; }, mr->d_small, "reduce_sizes", 16.0 * U);
      fetch_small(mr, 1);
      const uint64_t total = mr->h_small[0];
      MR_CUDA(cudaMemcpyAsync(oo + U, &total, 8, cudaMemcpyHostToDevice, s));
      kv->reserve_bytes(total);
      {
        ProfScope ps(pool, "reduce", (double) kmv->ksize + 12.0 * U + (double) total);
        reduce_fvar_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kmv->vkey, kmv->vkoff, kmv->fcnt, U, al, kv->data, oo);
        MR_CUDA(cudaGetLastError());
      }
      kv->nbytes = total;
      kv->nkv = U;
      kv->ksize = kmv->ksize;
      kv->vsize = 4 * U;
      sync(mr);
      delete mr->kmv;
      mr->kmv = nullptr;
      delete mr->kv;
      mr->kv = kv;
      return;
    }
    const uint8_t *kd = kmv->data;
    const uint64_t *ko = kmv->offs;
    bool cull = functor == MRGPU_REDUCE_CULL;
    const bool shape = functor == MRGPU_REDUCE_NONZERO || functor == MRGPU_REDUCE_DEGREE;
    if (functor == MRGPU_REDUCE_NONZERO && kmv->ukb >= 0 && (kmv->ukb % 2)) MR_FAIL(MRGPU_EINVAL, "reduce(nonzero): the key is not a pair of vertices");
    bool uni = kmv->ukb >= 0 && (!cull || kmv->uvb >= 0) && functor != MRGPU_REDUCE_POSTINGS;
    if (U == 0) {
      // empty KV
    } else if (uni) {
      const uint32_t ikb = (uint32_t) kmv->ukb;  // key bytes of the KMV pairs
      uint32_t kb = functor == MRGPU_REDUCE_NONZERO ? ikb / 2 : (functor == MRGPU_REDUCE_DEGREE ? 4 : ikb);
      uint32_t vb = shape ? 0 : (cull ? (uint32_t) kmv->uvb : 4);
      uint32_t rs = kv_rec_size(al, kb, vb);
      // cull = (key, first value) of every key.  The factored KMV's key array IS that KV in header-less form (rows =
      // keys, the one shared value kept aside: store.cuh), so the array changes owner and nothing is copied.
      const bool handover = cull && kmv->factored && kmv->fkey && kmv->fkw * 4 == kb && kv_can_pack(mr, kb, vb);
      if (handover) {
        kv->data = (uint8_t *) kmv->fkey;
        kv->cap = kmv->fcap * 4 * kmv->fkw;
        kmv->fkey = nullptr;
        kv->packed = true;
        kv->cval_on = vb > 0;
        kv->cval[0] = kmv->fval[0];
        kv->cval[1] = kmv->fval[1];
      } else
        kv->reserve_bytes(U * rs);
      ProfScope ps(pool, "reduce", handover ? 0.0 : (double) U * (12 + 4 + ikb + rs), handover ? 0 : 1);
      if (handover) {
      } else if (shape && kmv->factored)
        reduce_factored_shape_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kmv->fkey, kmv->fcnt, U, functor, al, ikb, kv->data, rs);
      else if (shape)
        reduce_generic_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kd, ko, U, functor, al, kv->data, nullptr, rs);
      else if (kmv->factored) {
        if (!cull && kb == 4 && rs == 16 && kv_key_off(al) == 8 && al.k == 4)
          reduce_factored16_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kmv->fkey, kmv->fcnt, U, (int4 *) kv->data);
        else if (al.k == 4 && al.v == 4 && rs == 8 + kb + vb)
          reduce_factored_words_kernel<<<grid_for(U * (rs / 4), 256, 64), 256, 0, s>>>(kmv->fkey, kmv->fcnt, U, functor, kb / 4, (uint32_t) kmv->uvb / 4,
                                                                                      kmv->fval[0], kmv->fval[1], (uint32_t *) kv->data);
        else
          reduce_factored_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kmv->fkey, kmv->fcnt, U, functor, al, kb, (uint32_t) kmv->uvb,
                                                                              kmv->fval[0], kmv->fval[1], kv->data, rs);
      } else if (!cull && kb == 4 && rs == 16 && kv_key_off(al) == 8 && al.k == 4)
        reduce_count16_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kd, ko, U, (int4 *) kv->data);
      else
        reduce_generic_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kd, ko, U, functor, al, kv->data, nullptr, rs);
      MR_CUDA(cudaGetLastError());
      kv->uniform = true;
      kv->kb = kb;
      kv->vb = vb;
      kv->recsize = rs;
      kv->nbytes = U * rs;
      kv->nkv = U;
      kv->ksize = U * kb;
      kv->vsize = U * vb;
      // one pair per key, in the KMV's order, key unchanged: keys that came out of the counting convert in ascending
      // order are still in it, so sort_keys(1) on int keys / sort_keys(2) on uint64_t keys has nothing to do
      if (kmv->factored && kmv->ascending && !shape && (kb == 4 || kb == 8)) kv->mark_sorted(kb == 4 ? 1 : 2);
    } else {
      kv->uniform = false;
      kv->reserve_offs(U + 1);
      uint64_t *oo = kv->offs;
      unsigned long long *vsum = mr->d_small + 1, *toobig = mr->d_small + 5;
      MR_CUDA(cudaMemsetAsync(vsum, 0, 8, s));
      MR_CUDA(cudaMemsetAsync(toobig, 0, 8, s));
      const uint64_t pagesize = mr->pagesize;
      scan_apply(
          pool, U,
          [=] __device__(uint64_t u) -> unsigned long long {
            KmvHead h = kmv_head(kd + ko[u], al);
            const uint64_t sz = round_up(round_up(round_up(8, al.k) + reduce_out_kb(functor, h), al.v) + (uint64_t) reduce_out_vb(functor, h), al.t);
            if (sz > pagesize || sz > 0x7FFFFFFFull) *toobig = 1;  // KeyValue::add: "Single key/value pair exceeds page size"
            return sz;
          },
          [=] __device__(uint64_t u, unsigned long long e, unsigned long long) {
            oo[u] = e;
            KmvHead h = kmv_head(kd + ko[u], al);
            warp_add(vsum, reduce_out_vb(functor, h));
          },
          mr->d_small);
      fetch_small(mr, 6);
      if (mr->h_small[5]) MR_FAIL(MRGPU_ELIMIT, "Single key/value pair exceeds page size");
      uint64_t total = mr->h_small[0];
      MR_CUDA(cudaMemcpyAsync(oo + U, &total, 8, cudaMemcpyHostToDevice, s));
      kv->reserve_bytes(total);
      if (functor == MRGPU_REDUCE_POSTINGS) {
        Buf<uint64_t> cum(pool, U + 1);
        uint64_t *pcum = cum.p;
        scan_apply(
            pool, U, [=] __device__(uint64_t u) -> unsigned long long { return ((const uint32_t *) (kd + ko[u]))[2]; },
            [=] __device__(uint64_t u, unsigned long long e, unsigned long long) { pcum[u] = e; }, mr->d_small + 2, "reduce_postings_scan", 16.0 * U);
        MR_CUDA(cudaMemcpyAsync(pcum + U, mr->d_small + 2, 8, cudaMemcpyDeviceToDevice, s));
        fetch_small(mr, 3);
        const uint64_t vtotal = mr->h_small[2];
        ProfScope ps(pool, "reduce_postings", (double) kmv->nbytes + (double) total, 2);
        reduce_postings_head_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kd, ko, U, al, kv->data, oo);
        if (vtotal) reduce_postings_body_kernel<<<grid_for((vtotal + 15) / 16, 256, 32), 256, 0, s>>>(kd, ko, U, al, kv->data, oo, pcum, vtotal);
        MR_CUDA(cudaGetLastError());
        sync(mr);  // cum is released on return
      } else
        reduce_generic_kernel<<<(unsigned) ((U + 255) / 256), 256, 0, s>>>(kd, ko, U, functor, al, kv->data, oo, 0);
      MR_CUDA(cudaGetLastError());
      kv->nbytes = total;
      kv->nkv = U;
      kv->ksize = functor == MRGPU_REDUCE_NONZERO ? kmv->ksize / 2 : (functor == MRGPU_REDUCE_DEGREE ? 4 * U : kmv->ksize);
      kv->vsize = mr->h_small[1];
    }
    sync(mr);
  } catch (...) {
    delete kv;
    throw;
  }
  // reduce consumes the KMV (src/mapreduce.cpp:1853-1855)
  delete mr->kmv;
  mr->kmv = nullptr;
  delete mr->kv;
  mr->kv = kv;
}

extern "C" int mrgpu_reduce(mrgpu_mr *mr, int functor, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "Cannot reduce without KeyMultiValue");
  const bool had = mr->g_what == G_KMV;  // every stock functor emits exactly one pair per KMV pair
  mr->g_what = G_NONE;
  if (!mr->kmv_spool.empty()) reduce_spooled(mr, functor);
  else reduce_impl(mr, functor);
  if (had) mr->g_what = G_KV;
  if (nkv) *nkv = global_count(mr, kv_total(mr), G_KV, true);
  API_END(mr)
}

// compress(): convert + reduce on this rank's own pairs, no exchange (src/mapreduce.cpp:749-851) -- the local
// combiner a caller runs before aggregate() to shrink what crosses the fabric
extern "C" int mrgpu_compress(mrgpu_mr *mr, int functor, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot compress without KeyValue");
  if (functor < 0 || functor > MRGPU_REDUCE_CULL) MR_FAIL(MRGPU_EINVAL, "unknown reduce functor");
  mr->g_what = G_NONE;
  convert_impl(mr);
  if (!mr->kmv_spool.empty()) reduce_spooled(mr, functor);
  else reduce_impl(mr, functor);
  if (nkv) *nkv = allreduce_sum(mr, kv_total(mr));
  API_END(mr)
}

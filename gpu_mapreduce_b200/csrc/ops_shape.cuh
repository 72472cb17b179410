// clone() and collapse(): a KeyValue re-shaped into a KeyMultiValue without grouping.
//
//   clone     src/mapreduce.cpp:631-672 -> KeyMultiValue::clone src/keymultivalue.cpp:344-374
//             every KV pair becomes a KMV pair with one value (nvalue 1, the pair's key, its value)
//   collapse  src/mapreduce.cpp:681-702 -> KeyMultiValue::collapse/collapse_one src/keymultivalue.cpp:384-468
//             all KV pairs become ONE KMV pair: key = the caller's key, values = key_1, value_1, key_2, value_2, ...
//             (2n values, sizes alternate keybytes/valuebytes, bytes packed without padding)
// Both are rank-local and keep the pairs in KV order.  Included by mrgpu.cu.
#pragma once

// (keybytes, valuebytes) of KV record i
__device__ __forceinline__ uint2 kv_lens(const uint8_t *__restrict__ data, const uint64_t *__restrict__ offs, uint64_t recsize, uint64_t i)
{
  const uint32_t *r = (const uint32_t *) (data + rec_off(offs, recsize, i));
  return make_uint2(r[0], r[1]);
}

__global__ void __launch_bounds__(256)
clone_write_kernel(const uint8_t *__restrict__ kvd, const uint64_t *__restrict__ kvo, uint64_t recsize, uint64_t n, Align al, uint8_t *__restrict__ out,
                   const uint64_t *__restrict__ offs)
{
  const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t *r = kvd + rec_off(kvo, recsize, i);
  const uint32_t kb = ((const uint32_t *) r)[0], vb = ((const uint32_t *) r)[1];
  const uint8_t *key = r + kv_key_off(al), *val = r + kv_val_off(al, kb);
  uint8_t *d = out + offs[i];
  const uint64_t ko = kmv_key_off(al, 1), vo = kmv_val_off(al, 1, kb), end = kmv_rec_size(al, 1, kb, vb);
  uint32_t *h = (uint32_t *) d;
  h[0] = 1;
  h[1] = kb;
  h[2] = vb;
  h[3] = vb;  // valuesizes[0]
  for (uint64_t q = 16; q < ko; q++) d[q] = 0;
  for (uint32_t q = 0; q < kb; q++) d[ko + q] = key[q];
  for (uint64_t q = ko + kb; q < vo; q++) d[q] = 0;
  for (uint32_t q = 0; q < vb; q++) d[vo + q] = val[q];
  for (uint64_t q = vo + vb; q < end; q++) d[q] = 0;
}

static DevKMV *empty_kmv(mrgpu_mr *mr)
{
  DevKMV *kmv = new DevKMV(&mr->pool);
  kmv->offs = (uint64_t *) mr->pool.alloc(8);
  kmv->offs_cap = 1;
  MR_CUDA(cudaMemsetAsync(kmv->offs, 0, 8, mr->pool.stream));
  return kmv;
}

static void clone_impl(mrgpu_mr *mr)
{
  freeze(mr);
  DevKV *kv = mr->kv;
  kv_unpack(mr, kv);
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t n = kv->nkv;
  delete mr->kmv;
  mr->kmv = nullptr;
  spool_clear_kmv(mr);
  DevKMV *kmv = empty_kmv(mr);
  try {
    if (n) {
      pool.free(kmv->offs, kmv->offs_cap * 8);
      kmv->offs = (uint64_t *) pool.alloc((n + 1) * 8);
      kmv->offs_cap = n + 1;
      uint64_t *offs = kmv->offs;
      const uint8_t *kvd = kv->data;
      const uint64_t *kvo = kv->uniform ? nullptr : kv->offs;
      const uint64_t rs = kv->recsize;
      scan_apply(
          pool, n, [=] __device__(uint64_t i) -> unsigned long long {
            const uint2 l = kv_lens(kvd, kvo, rs, i);
            return kmv_rec_size(al, 1, l.x, l.y);
          },
          [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { offs[i] = e; }, mr->d_small, "clone_sizes", 16.0 * n);
      fetch_small(mr, 1);
      const uint64_t total = mr->h_small[0];
      MR_CUDA(cudaMemcpyAsync(offs + n, &total, 8, cudaMemcpyHostToDevice, s));
      kmv->data = (uint8_t *) pool.alloc(total);
      kmv->cap = kmv->nbytes = total;
      {
        ProfScope ps(pool, "clone_write", (double) kv->nbytes + (double) total);
        clone_write_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(kvd, kvo, rs, n, al, kmv->data, offs);
        MR_CUDA(cudaGetLastError());
      }
      kmv->nkmv = kmv->nvalue = n;
      kmv->ksize = kv->ksize;
      kmv->vsize = kv->vsize;
      kmv->ukb = kv->uniform ? (int64_t) kv->kb : -1;
      kmv->uvb = kv->uniform ? (int64_t) kv->vb : -1;
      sync(mr);
    }
  } catch (...) {
    delete kmv;
    throw;
  }
  mr->kmv = kmv;
  delete mr->kv;  // :665-666
  mr->kv = nullptr;
}

// value bytes of the collapsed pair: record i contributes its key then its value at pos[i]
__global__ void __launch_bounds__(256)
collapse_write_kernel(const uint8_t *__restrict__ kvd, const uint64_t *__restrict__ kvo, uint64_t recsize, uint64_t n, Align al,
                      const uint64_t *__restrict__ pos, uint32_t *__restrict__ sizes, uint8_t *__restrict__ values)
{
  const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t *r = kvd + rec_off(kvo, recsize, i);
  const uint32_t kb = ((const uint32_t *) r)[0], vb = ((const uint32_t *) r)[1];
  const uint8_t *key = r + kv_key_off(al), *val = r + kv_val_off(al, kb);
  sizes[2 * i] = kb;
  sizes[2 * i + 1] = vb;
  uint8_t *d = values + pos[i];
  for (uint32_t q = 0; q < kb; q++) d[q] = key[q];
  for (uint32_t q = 0; q < vb; q++) d[kb + q] = val[q];
}

static void collapse_impl(mrgpu_mr *mr, const void *key, int keybytes)
{
  freeze(mr);
  DevKV *kv = mr->kv;
  kv_unpack(mr, kv);
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  const uint64_t n = kv->nkv, mvb = kv->ksize + kv->vsize;
  const uint64_t ko = kmv_key_off(al, 2 * n), vo = kmv_val_off(al, 2 * n, (uint32_t) keybytes), total = kmv_rec_size(al, 2 * n, (uint32_t) keybytes, mvb);
  // src/keymultivalue.cpp:399-402, 475-478: one page, int-sized counts; the reference has no multi-block collapse either
  if (2 * n > 0x7FFFFFFFull || mvb > 0x7FFFFFFFull || total > mr->pagesize) MR_FAIL(MRGPU_ELIMIT, "Collapse of multi-block KMV not yet supported");
  delete mr->kmv;
  mr->kmv = nullptr;
  spool_clear_kmv(mr);
  DevKMV *kmv = new DevKMV(&pool);
  try {
    kmv->offs = (uint64_t *) pool.alloc(16);
    kmv->offs_cap = 2;
    const uint64_t ho[2] = {0, total};
    MR_CUDA(cudaMemcpyAsync(kmv->offs, ho, 16, cudaMemcpyHostToDevice, s));
    kmv->data = (uint8_t *) pool.alloc(total);
    kmv->cap = kmv->nbytes = total;
    MR_CUDA(cudaMemsetAsync(kmv->data, 0, total, s));  // header gaps and padding
    const uint32_t head[3] = {(uint32_t) (2 * n), (uint32_t) keybytes, (uint32_t) mvb};
    MR_CUDA(cudaMemcpyAsync(kmv->data, head, 12, cudaMemcpyHostToDevice, s));
    if (keybytes) MR_CUDA(cudaMemcpyAsync(kmv->data + ko, key, keybytes, cudaMemcpyHostToDevice, s));
    sync(mr);  // ho/head are stack variables
    if (n) {
      Buf<uint64_t> pos(pool, n);
      uint64_t *pp = pos.p;
      const uint8_t *kvd = kv->data;
      const uint64_t *kvo = kv->uniform ? nullptr : kv->offs;
      const uint64_t rs = kv->recsize;
      scan_apply(
          pool, n, [=] __device__(uint64_t i) -> unsigned long long {
            const uint2 l = kv_lens(kvd, kvo, rs, i);
            return (unsigned long long) l.x + l.y;
          },
          [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { pp[i] = e; }, mr->d_small, "collapse_sizes", 16.0 * n);
      ProfScope ps(pool, "collapse_write", (double) kv->nbytes + (double) total);
      collapse_write_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(kvd, kvo, rs, n, al, pp, (uint32_t *) (kmv->data + 12), kmv->data + vo);
      MR_CUDA(cudaGetLastError());
    }
    kmv->nkmv = 1;
    kmv->nvalue = 2 * n;
    kmv->ksize = (uint64_t) keybytes;
    kmv->vsize = mvb;
    kmv->ukb = keybytes;
    kmv->uvb = -1;
    sync(mr);
  } catch (...) {
    delete kmv;
    throw;
  }
  mr->kmv = kmv;
  delete mr->kv;  // :695-696
  mr->kv = nullptr;
}

extern "C" int mrgpu_clone(mrgpu_mr *mr, uint64_t *nkmv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot clone without KeyValue");
  mr->g_what = G_NONE;
  spool_make_resident(mr);
  clone_impl(mr);
  if (nkmv) *nkmv = allreduce_sum(mr, mr->kmv->nkmv);
  API_END(mr)
}

extern "C" int mrgpu_collapse(mrgpu_mr *mr, const void *key, int keybytes, uint64_t *nkmv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot collapse without KeyValue");
  if (keybytes < 0 || (keybytes && !key)) MR_FAIL(MRGPU_EINVAL, "bad key");
  mr->g_what = G_NONE;
  spool_make_resident(mr);
  collapse_impl(mr, key, keybytes);
  if (nkmv) *nkmv = allreduce_sum(mr, mr->kmv->nkmv);
  API_END(mr)
}

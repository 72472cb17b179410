// Stats and page access: the host-visible side of the HBM stores.
//
// The reference hands callers one `memsize` page at a time (KeyValue::request_info/request_page
// src/keyvalue.cpp:277-308, KeyMultiValue::request_page src/keymultivalue.cpp:236-260) and scan()/print()/
// host reduce callbacks walk records inside it.  The stores here are one contiguous record stream, so a
// "page" is the slice of it that the reference's greedy fill (src/keyvalue.cpp:362-380,
// src/keymultivalue.cpp:853-868) would have put in page i; the table of slices is built on first request
// from the record-offset index.  Included by mrgpu.cu.
#pragma once

struct PageSlice {
  uint64_t first, nkey, begin, end;  // record range and byte range
  int extended;
};

static void fetch_offs(mrgpu_mr *mr, const uint64_t *d_offs, uint64_t count, std::vector<uint64_t> &h)
{
  h.resize(count);
  if (count) MR_CUDA(cudaMemcpyAsync(h.data(), d_offs, count * 8, cudaMemcpyDeviceToHost, mr->pool.stream));
  sync(mr);
}

static constexpr uint64_t kIntMax = 0x7FFFFFFFull;

// page i of a KV
static PageSlice kv_page(mrgpu_mr *mr, uint64_t ipage, uint64_t *npage_out)
{
  DevKV *kv = mr->kv;
  const uint64_t ps = mr->pagesize;
  PageSlice sl{0, 0, 0, 0, 0};
  if (kv->nkv == 0) {  // the reference always has one (empty) page
    if (npage_out) *npage_out = 1;
    return sl;
  }
  if (kv->uniform) {
    uint64_t rpp = std::min<uint64_t>(ps / kv->recsize, kIntMax);
    uint64_t np = (kv->nkv + rpp - 1) / rpp;
    if (npage_out) *npage_out = np;
    if (ipage < np) {
      sl.first = ipage * rpp;
      sl.nkey = std::min(rpp, kv->nkv - sl.first);
      sl.begin = sl.first * kv->recsize;
      sl.end = sl.begin + sl.nkey * kv->recsize;
    }
    return sl;
  }
  if (!kv->h_offs_valid) {
    fetch_offs(mr, kv->offs, kv->nkv + 1, kv->h_offs);
    kv->h_offs_valid = true;
  }
  const std::vector<uint64_t> &o = kv->h_offs;
  uint64_t first = 0, page = 0;
  while (first < kv->nkv) {
    // last record whose end is within first's start + pagesize
    uint64_t limit = o[first] + ps;
    uint64_t hi = std::upper_bound(o.begin() + first + 1, o.begin() + kv->nkv + 1, limit) - o.begin() - 1;
    if (hi - first > kIntMax) hi = first + kIntMax;
    if (page == ipage) {
      sl.first = first;
      sl.nkey = hi - first;
      sl.begin = o[first];
      sl.end = o[hi];
    }
    first = hi;
    page++;
  }
  if (npage_out) *npage_out = page;
  return sl;
}

// page i of a KMV; a pair that alone exceeds the page (or INTMAX values / bytes) is an "extended" pair on a
// page run of its own, as in unique2kmv_all src/keymultivalue.cpp:840-928
static PageSlice kmv_page(mrgpu_mr *mr, uint64_t ipage, uint64_t *npage_out)
{
  DevKMV *kmv = mr->kmv;
  kmv_expand(mr, kmv);  // page boundaries and bytes are those of the record stream
  const uint64_t ps = mr->pagesize;
  PageSlice sl{0, 0, 0, 0, 0};
  if (kmv->nkmv == 0) {
    if (npage_out) *npage_out = 1;
    return sl;
  }
  if (!kmv->h_offs_valid) {
    fetch_offs(mr, kmv->offs, kmv->nkmv + 1, kmv->h_offs);
    kmv->h_offs_valid = true;
  }
  const std::vector<uint64_t> &o = kmv->h_offs;
  uint64_t first = 0, page = 0;
  while (first < kmv->nkmv) {
    uint64_t hi;
    int ext = 0;
    if (o[first + 1] - o[first] > ps) {
      hi = first + 1;
      ext = 1;
    } else {
      uint64_t limit = o[first] + ps;
      hi = std::upper_bound(o.begin() + first + 1, o.begin() + kmv->nkmv + 1, limit) - o.begin() - 1;
      if (hi - first > kIntMax) hi = first + kIntMax;
    }
    if (page == ipage) {
      sl.first = first;
      sl.nkey = hi - first;
      sl.begin = o[first];
      sl.end = o[hi];
      sl.extended = ext;
    }
    first = hi;
    page++;
  }
  if (npage_out) *npage_out = page;
  return sl;
}

extern "C" int mrgpu_kv_stats(mrgpu_mr *mr, mrgpu_stats *out)
{
  API_BEGIN(mr)
  if (!out) MR_FAIL(MRGPU_EINVAL, "null out");
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot print stats without KeyValue");
  freeze(mr);
  DevKV *kv = mr->kv;
  out->npair = kv->nkv;
  out->nvalue = kv->nkv;
  out->keysize = kv->ksize;
  out->valuesize = kv->vsize;
  out->exactsize = 8 * kv->nkv + kv->ksize + kv->vsize;  // src/keyvalue.cpp:671-672
  out->alignsize = kv->nbytes;
  kv_page(mr, ~0ull, &out->npage);
  API_END(mr)
}

extern "C" int mrgpu_kmv_stats(mrgpu_mr *mr, mrgpu_stats *out)
{
  API_BEGIN(mr)
  if (!out) MR_FAIL(MRGPU_EINVAL, "null out");
  if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "Cannot print stats without KeyMultiValue");
  freeze(mr);
  DevKMV *kmv = mr->kmv;
  kmv_expand(mr, kmv);
  out->npair = kmv->nkmv;
  out->nvalue = kmv->nvalue;
  out->keysize = kmv->ksize;
  out->valuesize = kmv->vsize;
  out->exactsize = 12 * kmv->nkmv + 4 * kmv->nvalue + kmv->ksize + kmv->vsize;  // src/keymultivalue.cpp:1460-1461
  out->alignsize = kmv->nbytes;
  kmv_page(mr, ~0ull, &out->npage);
  API_END(mr)
}

// the pair count summed over the ranks, as MapReduce::kv_stats / kmv_stats return it (src/mapreduce.cpp:2945, 2988).
// COLLECTIVE (one all-reduce unless the last operation already produced the number).
extern "C" int mrgpu_global_count(mrgpu_mr *mr, int which, uint64_t *n)
{
  API_BEGIN(mr)
  if (!n) MR_FAIL(MRGPU_EINVAL, "null out");
  if (which == 0) {
    if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot print stats without KeyValue");
    *n = global_count(mr, mr->kv->nkv, G_KV, true);
  } else {
    if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "Cannot print stats without KeyMultiValue");
    *n = global_count(mr, mr->kmv->nkmv, G_KMV, true);
  }
  API_END(mr)
}

extern "C" int mrgpu_kv_page_info(mrgpu_mr *mr, uint64_t ipage, uint64_t *nkey, uint64_t *keysize, uint64_t *valuesize, uint64_t *alignsize)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot request a page without KeyValue");
  freeze(mr);
  uint64_t np;
  PageSlice sl = kv_page(mr, ipage, &np);
  if (ipage >= np) MR_FAIL(MRGPU_EINVAL, "page index out of range");
  DevKV *kv = mr->kv;
  if (nkey) *nkey = sl.nkey;
  if (alignsize) *alignsize = sl.end - sl.begin;
  if (keysize || valuesize) {
    uint64_t ks = 0, vs = 0;
    if (kv->uniform) {
      ks = sl.nkey * kv->kb;
      vs = sl.nkey * kv->vb;
    } else if (sl.nkey) {
      MR_CUDA(cudaMemsetAsync(mr->d_small, 0, 16, mr->pool.stream));
      sum_kv_bytes_kernel<<<grid_for(sl.nkey, 256 * 4, 8), 256, 0, mr->pool.stream>>>(kv->data, kv->offs + sl.first, sl.nkey, mr->d_small,
                                                                                   mr->d_small + 1);
      MR_CUDA(cudaGetLastError());
      fetch_small(mr, 2);
      ks = mr->h_small[0];
      vs = mr->h_small[1];
    }
    if (keysize) *keysize = ks;
    if (valuesize) *valuesize = vs;
  }
  API_END(mr)
}

extern "C" int mrgpu_kv_page_copy(mrgpu_mr *mr, uint64_t ipage, void *dst)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot request a page without KeyValue");
  freeze(mr);
  uint64_t np;
  PageSlice sl = kv_page(mr, ipage, &np);
  if (ipage >= np) MR_FAIL(MRGPU_EINVAL, "page index out of range");
  if (sl.end > sl.begin) {
    if (!dst) MR_FAIL(MRGPU_EINVAL, "null destination");
    if (mr->kv->weighted) kv_unpack(mr, mr->kv);  // rows [key | count] -> records, once; pages are then plain slices
    if (mr->kv->packed) {  // header-less rows: write this page's records out, the KV itself stays as it is
      Buf<uint8_t> tmp(mr->pool, sl.end - sl.begin);
      kv_unpack_rows(mr, mr->kv, sl.first, sl.nkey, tmp.p);
      MR_CUDA(cudaMemcpyAsync(dst, tmp.p, sl.end - sl.begin, cudaMemcpyDeviceToHost, mr->pool.stream));
      sync(mr);
    } else
      MR_CUDA(cudaMemcpyAsync(dst, mr->kv->data + sl.begin, sl.end - sl.begin, cudaMemcpyDeviceToHost, mr->pool.stream));
  }
  sync(mr);
  API_END(mr)
}

extern "C" int mrgpu_kmv_page_info(mrgpu_mr *mr, uint64_t ipage, uint64_t *nkey, uint64_t *alignsize, int *is_extended)
{
  API_BEGIN(mr)
  if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "Cannot request a page without KeyMultiValue");
  freeze(mr);
  uint64_t np;
  PageSlice sl = kmv_page(mr, ipage, &np);
  if (ipage >= np) MR_FAIL(MRGPU_EINVAL, "page index out of range");
  if (nkey) *nkey = sl.nkey;
  if (alignsize) *alignsize = sl.end - sl.begin;
  if (is_extended) *is_extended = sl.extended;
  API_END(mr)
}

extern "C" int mrgpu_kmv_page_copy(mrgpu_mr *mr, uint64_t ipage, void *dst)
{
  API_BEGIN(mr)
  if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "Cannot request a page without KeyMultiValue");
  freeze(mr);
  uint64_t np;
  PageSlice sl = kmv_page(mr, ipage, &np);
  if (ipage >= np) MR_FAIL(MRGPU_EINVAL, "page index out of range");
  if (sl.end > sl.begin) {
    if (!dst) MR_FAIL(MRGPU_EINVAL, "null destination");
    MR_CUDA(cudaMemcpyAsync(dst, mr->kmv->data + sl.begin, sl.end - sl.begin, cudaMemcpyDeviceToHost, mr->pool.stream));
  }
  sync(mr);
  API_END(mr)
}

extern "C" int mrgpu_kv_device_view(mrgpu_mr *mr, const void **data, uint64_t *nbytes, const uint64_t **offsets, uint64_t *record_bytes)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "no KeyValue");
  DevKV *kv = mr->kv;
  kv_unpack(mr, kv);  // the view is of the record stream
  if (data) *data = kv->data;
  if (nbytes) *nbytes = kv->nbytes;
  if (offsets) *offsets = (kv->uniform || kv->nkv == 0) ? nullptr : kv->offs;
  if (record_bytes) *record_bytes = (kv->uniform && kv->nkv) ? kv->recsize : 0;
  API_END(mr)
}

extern "C" int mrgpu_kmv_device_view(mrgpu_mr *mr, const void **data, uint64_t *nbytes, const uint64_t **offsets)
{
  API_BEGIN(mr)
  if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "no KeyMultiValue");
  kmv_expand(mr, mr->kmv);
  if (data) *data = mr->kmv->data;
  if (nbytes) *nbytes = mr->kmv->nbytes;
  if (offsets) *offsets = mr->kmv->offs;
  API_END(mr)
}

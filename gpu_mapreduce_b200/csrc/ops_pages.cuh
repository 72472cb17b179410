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

// page i of a KV given by its record count, shape and (host) offset index: the reference's greedy fill
static PageSlice kv_slice(uint64_t nkv, bool uniform, uint64_t recsize, const uint64_t *o, uint64_t ps, uint64_t ipage, uint64_t *npage_out)
{
  PageSlice sl{0, 0, 0, 0, 0};
  if (nkv == 0) {  // the reference always has one (empty) page
    if (npage_out) *npage_out = 1;
    return sl;
  }
  if (uniform) {
    uint64_t rpp = std::min<uint64_t>(ps / recsize, kIntMax);
    uint64_t np = (nkv + rpp - 1) / rpp;
    if (npage_out) *npage_out = np;
    if (ipage < np) {
      sl.first = ipage * rpp;
      sl.nkey = std::min(rpp, nkv - sl.first);
      sl.begin = sl.first * recsize;
      sl.end = sl.begin + sl.nkey * recsize;
    }
    return sl;
  }
  uint64_t first = 0, page = 0;
  while (first < nkv) {
    // last record whose end is within first's start + pagesize
    uint64_t limit = o[first] + ps;
    uint64_t hi = std::upper_bound(o + first + 1, o + nkv + 1, limit) - o - 1;
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

// page i of the part of the KV that is in HBM
static PageSlice kv_page(mrgpu_mr *mr, uint64_t ipage, uint64_t *npage_out)
{
  DevKV *kv = mr->kv;
  if (kv->nkv && !kv->uniform && !kv->h_offs_valid) {
    fetch_offs(mr, kv->offs, kv->nkv + 1, kv->h_offs);
    kv->h_offs_valid = true;
  }
  return kv_slice(kv->nkv, kv->uniform, kv->recsize, kv->h_offs.data(), mr->pagesize, ipage, npage_out);
}

// A page of the whole KV = host pieces (out-of-core tier, ops_spool.cuh) in order, then the part in HBM.  Every piece
// paginates on its own: a piece boundary ends a page, as a spill does in the reference.  piece = -1: the HBM part.
struct PageLoc {
  int piece;
  PageSlice sl;
};
static PageLoc kv_page_locate(mrgpu_mr *mr, uint64_t ipage, uint64_t *npage_out)
{
  PageLoc loc{-1, PageSlice{0, 0, 0, 0, 0}};
  uint64_t before = 0;
  bool found = false;
  for (size_t i = 0; i < mr->kv_spool.size(); i++) {
    const HostKV &h = mr->kv_spool[i];
    uint64_t np = 0;
    PageSlice sl = kv_slice(h.nkv, h.uniform, h.recsize, h.offs, mr->pagesize, ipage >= before ? ipage - before : ~0ull, &np);
    if (!found && ipage >= before && ipage - before < np) loc = PageLoc{(int) i, sl}, found = true;
    before += np;
  }
  if (mr->kv_spool.empty() || (mr->kv && mr->kv->nkv)) {
    uint64_t np = 0;
    PageSlice sl = kv_page(mr, ipage >= before ? ipage - before : ~0ull, &np);
    if (!found && ipage >= before && ipage - before < np) loc = PageLoc{-1, sl}, found = true;
    before += np;
  }
  if (npage_out) *npage_out = before;
  return loc;
}

static PageSlice kmv_slice(uint64_t nkmv, const uint64_t *o, uint64_t ps, uint64_t ipage, uint64_t *npage_out)
{
  PageSlice sl{0, 0, 0, 0, 0};
  if (nkmv == 0) {
    if (npage_out) *npage_out = 1;
    return sl;
  }
  uint64_t first = 0, page = 0;
  while (first < nkmv) {
    uint64_t hi;
    int ext = 0;
    if (o[first + 1] - o[first] > ps) {
      hi = first + 1;
      ext = 1;
    } else {
      uint64_t limit = o[first] + ps;
      hi = std::upper_bound(o + first + 1, o + nkmv + 1, limit) - o - 1;
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

// page i of the part of the KMV that is in HBM; a pair that alone exceeds the page (or INTMAX values / bytes) is an
// "extended" pair on a page run of its own, as in unique2kmv_all src/keymultivalue.cpp:840-928
static PageSlice kmv_page(mrgpu_mr *mr, uint64_t ipage, uint64_t *npage_out)
{
  DevKMV *kmv = mr->kmv;
  kmv_expand(mr, kmv);  // page boundaries and bytes are those of the record stream
  if (kmv->nkmv && !kmv->h_offs_valid) {
    fetch_offs(mr, kmv->offs, kmv->nkmv + 1, kmv->h_offs);
    kmv->h_offs_valid = true;
  }
  return kmv_slice(kmv->nkmv, kmv->h_offs.data(), mr->pagesize, ipage, npage_out);
}

static PageLoc kmv_page_locate(mrgpu_mr *mr, uint64_t ipage, uint64_t *npage_out)
{
  PageLoc loc{-1, PageSlice{0, 0, 0, 0, 0}};
  uint64_t before = 0;
  bool found = false;
  for (size_t i = 0; i < mr->kmv_spool.size(); i++) {
    const HostKMV &h = mr->kmv_spool[i];
    uint64_t np = 0;
    PageSlice sl = kmv_slice(h.nkmv, h.offs, mr->pagesize, ipage >= before ? ipage - before : ~0ull, &np);
    if (!found && ipage >= before && ipage - before < np) loc = PageLoc{(int) i, sl}, found = true;
    before += np;
  }
  if (mr->kmv_spool.empty() || (mr->kmv && mr->kmv->nkmv)) {
    uint64_t np = 0;
    PageSlice sl = kmv_page(mr, ipage >= before ? ipage - before : ~0ull, &np);
    if (!found && ipage >= before && ipage - before < np) loc = PageLoc{-1, sl}, found = true;
    before += np;
  }
  if (npage_out) *npage_out = before;
  return loc;
}

extern "C" int mrgpu_kv_stats(mrgpu_mr *mr, mrgpu_stats *out)
{
  API_BEGIN(mr)
  if (!out) MR_FAIL(MRGPU_EINVAL, "null out");
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot print stats without KeyValue");
  freeze(mr);
  DevKV *kv = mr->kv;
  out->npair = kv->nkv;
  out->keysize = kv->ksize;
  out->valuesize = kv->vsize;
  out->alignsize = kv->nbytes;
  for (const HostKV &h : mr->kv_spool) {  // pieces in host memory (out-of-core tier)
    out->npair += h.nkv;
    out->keysize += h.ksize;
    out->valuesize += h.vsize;
    out->alignsize += h.nbytes;
  }
  out->nvalue = out->npair;
  out->exactsize = 8 * out->npair + out->keysize + out->valuesize;  // src/keyvalue.cpp:671-672
  kv_page_locate(mr, ~0ull, &out->npage);
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
  out->alignsize = kmv->nbytes;
  for (const HostKMV &h : mr->kmv_spool) {
    out->npair += h.nkmv;
    out->nvalue += h.nvalue;
    out->keysize += h.ksize;
    out->valuesize += h.vsize;
    out->alignsize += h.nbytes;
  }
  out->exactsize = 12 * out->npair + 4 * out->nvalue + out->keysize + out->valuesize;  // src/keymultivalue.cpp:1460-1461
  kmv_page_locate(mr, ~0ull, &out->npage);
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
    *n = global_count(mr, kv_total(mr), G_KV, true);
  } else {
    if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "Cannot print stats without KeyMultiValue");
    *n = global_count(mr, kmv_total(mr), G_KMV, true);
  }
  API_END(mr)
}

extern "C" int mrgpu_kv_page_info(mrgpu_mr *mr, uint64_t ipage, uint64_t *nkey, uint64_t *keysize, uint64_t *valuesize, uint64_t *alignsize)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot request a page without KeyValue");
  freeze(mr);
  uint64_t np;
  const PageLoc loc = kv_page_locate(mr, ipage, &np);
  const PageSlice sl = loc.sl;
  if (ipage >= np) MR_FAIL(MRGPU_EINVAL, "page index out of range");
  DevKV *kv = mr->kv;
  if (nkey) *nkey = sl.nkey;
  if (alignsize) *alignsize = sl.end - sl.begin;
  if (loc.piece >= 0) {  // a piece in host memory: its records are right here
    const HostKV &h = mr->kv_spool[loc.piece];
    uint64_t ks = 0, vs = 0;
    if (h.uniform) ks = sl.nkey * h.kb, vs = sl.nkey * h.vb;
    else
      for (uint64_t i = sl.first; i < sl.first + sl.nkey; i++) {
        const uint32_t *r = (const uint32_t *) (h.data + h.offs[i]);
        ks += r[0], vs += r[1];
      }
    if (keysize) *keysize = ks;
    if (valuesize) *valuesize = vs;
  } else if (keysize || valuesize) {
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
  const PageLoc loc = kv_page_locate(mr, ipage, &np);
  const PageSlice sl = loc.sl;
  if (ipage >= np) MR_FAIL(MRGPU_EINVAL, "page index out of range");
  if (loc.piece >= 0) {
    if (sl.end > sl.begin && !dst) MR_FAIL(MRGPU_EINVAL, "null destination");
    if (sl.end > sl.begin) memcpy(dst, mr->kv_spool[loc.piece].data + sl.begin, sl.end - sl.begin);
  } else if (sl.end > sl.begin) {
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
  const PageSlice sl = kmv_page_locate(mr, ipage, &np).sl;
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
  const PageLoc loc = kmv_page_locate(mr, ipage, &np);
  const PageSlice sl = loc.sl;
  if (ipage >= np) MR_FAIL(MRGPU_EINVAL, "page index out of range");
  if (loc.piece >= 0) {
    if (sl.end > sl.begin && !dst) MR_FAIL(MRGPU_EINVAL, "null destination");
    if (sl.end > sl.begin) memcpy(dst, mr->kmv_spool[loc.piece].data + sl.begin, sl.end - sl.begin);
  } else if (sl.end > sl.begin) {
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
  spool_make_resident(mr);
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
  spool_make_resident(mr);
  kmv_expand(mr, mr->kmv);
  if (data) *data = mr->kmv->data;
  if (nbytes) *nbytes = mr->kmv->nbytes;
  if (offsets) *offsets = mr->kmv->offs;
  API_END(mr)
}

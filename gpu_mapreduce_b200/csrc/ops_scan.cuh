// scan(): read-only walk of the KeyValue / KeyMultiValue with a HOST callback, and the multi-block protocol a
// callback uses for a KMV pair that is bigger than a page.
//
// Reference: MapReduce::scan (KV) src/mapreduce.cpp:1933-1976, (KMV) :1984-2065; multivalue_blocks :1874,
// multivalue_block_select :1887, multivalue_block :1901.  The reference hands the callback pointers into its
// current page; here the page is staged device -> host first (the stores live in HBM), one reference page at a
// time, so the callback sees the same bytes under the same lifetime rule: pointers are valid only during the
// callback.  Included by mrgpu.cu.
#pragma once

struct ScanBlocks {
  bool valid = false;
  int nblock = 0;
  uint64_t nvalue_total = 0, capacity = 0;
  const int *sizes = nullptr;     // value sizes of the whole extended pair (staged)
  const char *values = nullptr;   // its value bytes (staged)
};

static std::map<mrgpu_mr *, ScanBlocks> g_scan_blocks;
static std::mutex g_scan_mu;

static ScanBlocks *scan_blocks(mrgpu_mr *mr, bool create)
{
  std::lock_guard<std::mutex> lk(g_scan_mu);
  auto it = g_scan_blocks.find(mr);
  if (it != g_scan_blocks.end()) return &it->second;
  if (!create) return nullptr;
  return &g_scan_blocks[mr];
}

static void scan_forget(mrgpu_mr *mr)
{
  std::lock_guard<std::mutex> lk(g_scan_mu);
  g_scan_blocks.erase(mr);
}

extern "C" int mrgpu_scan_kv(mrgpu_mr *mr, mrgpu_scan_kv_fn fn, void *ptr, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot scan without KeyValue");
  if (!fn) MR_FAIL(MRGPU_EINVAL, "null callback");
  freeze(mr);
  const Align al = mr->al;
  uint64_t np = 0;
  kv_page_locate(mr, ~0ull, &np);
  std::vector<char> page;
  for (uint64_t ip = 0; ip < np; ip++) {
    uint64_t nkey = 0, asz = 0;
    if (mrgpu_kv_page_info(mr, ip, &nkey, nullptr, nullptr, &asz) != MRGPU_OK) MR_FAIL(MRGPU_EINVAL, mr->err);
    page.resize(asz ? asz : 1);
    if (mrgpu_kv_page_copy(mr, ip, page.data()) != MRGPU_OK) MR_FAIL(MRGPU_ECUDA, mr->err);
    uint64_t p = 0;
    for (uint64_t i = 0; i < nkey; i++) {
      const int kb = *(const int *) &page[p], vb = *(const int *) &page[p + 4];
      const uint64_t k0 = round_up(p + 8, al.k), v0 = round_up(k0 + kb, al.v);
      fn(&page[k0], kb, &page[v0], vb, ptr);
      p = round_up(v0 + vb, al.t);
    }
  }
  if (nkv) *nkv = global_count(mr, kv_total(mr), G_KV, true);
  API_END(mr)
}

extern "C" int mrgpu_scan_kmv(mrgpu_mr *mr, mrgpu_scan_kmv_fn fn, void *ptr, uint64_t *nkmv)
{
  API_BEGIN(mr)
  if (!mr->kmv) MR_FAIL(MRGPU_ENOKMV, "Cannot scan without KeyMultiValue");
  if (!fn) MR_FAIL(MRGPU_EINVAL, "null callback");
  freeze(mr);
  const Align al = mr->al;
  uint64_t np = 0;
  kmv_page_locate(mr, ~0ull, &np);
  std::vector<char> page;
  ScanBlocks *sb = scan_blocks(mr, true);
  for (uint64_t ip = 0; ip < np; ip++) {
    uint64_t nkey = 0, asz = 0;
    int ext = 0;
    if (mrgpu_kmv_page_info(mr, ip, &nkey, &asz, &ext) != MRGPU_OK) MR_FAIL(MRGPU_EINVAL, mr->err);
    page.resize(asz ? asz : 1);
    if (mrgpu_kmv_page_copy(mr, ip, page.data()) != MRGPU_OK) MR_FAIL(MRGPU_ECUDA, mr->err);
    uint64_t p = 0;
    for (uint64_t i = 0; i < nkey; i++) {
      const int nv = *(const int *) &page[p], kb = *(const int *) &page[p + 4], mvb = *(const int *) &page[p + 8];
      int *sizes = (int *) &page[p + 12];
      const uint64_t k0 = round_up(p + 12 + 4 * (uint64_t) (unsigned) nv, al.k), v0 = round_up(k0 + kb, al.v);
      if (!ext) fn(&page[k0], kb, &page[v0], nv, sizes, ptr);
      else {
        // a pair bigger than a page: the callback gets multivalue == NULL, nvalues == 0 and the handle in the
        // valuesizes slot (the reference passes its `this` there, src/mapreduce.cpp:2041-2053) and fetches the
        // values block by block; a block holds pagesize / 2 / sizeof(int) - 1 values (src/keymultivalue.cpp:1241-1242)
        sb->valid = true;
        sb->nvalue_total = (uint64_t) (unsigned) nv;
        sb->capacity = std::max<uint64_t>(1, mr->pagesize / 2 / sizeof(int) - 1);
        sb->nblock = (int) ((sb->nvalue_total + sb->capacity - 1) / sb->capacity);
        sb->sizes = sizes;
        sb->values = &page[v0];
        fn(&page[k0], kb, nullptr, 0, (int *) mr, ptr);
        sb->valid = false;
      }
      p = round_up(v0 + (uint64_t) (unsigned) mvb, al.t);
    }
  }
  if (nkmv) *nkmv = global_count(mr, kmv_total(mr), G_KMV, true);
  API_END(mr)
}

extern "C" int mrgpu_multivalue_blocks(mrgpu_mr *mr, int *nblock, uint64_t *nvalue_total)
{
  API_BEGIN(mr)
  ScanBlocks *sb = scan_blocks(mr, false);
  if (!sb || !sb->valid) MR_FAIL(MRGPU_EINVAL, "Invalid call to multivalue_blocks()");
  if (nblock) *nblock = sb->nblock;
  if (nvalue_total) *nvalue_total = sb->nvalue_total;
  API_END(mr)
}

extern "C" int mrgpu_multivalue_block_select(mrgpu_mr *mr, int which)
{
  API_BEGIN(mr)
  ScanBlocks *sb = scan_blocks(mr, false);
  if (!sb || !sb->valid) MR_FAIL(MRGPU_EINVAL, "Invalid call to multivalue_block_select()");
  if (which != 1 && which != 2) MR_FAIL(MRGPU_EINVAL, "Invalid arg to multivalue_block_select()");
  API_END(mr)
}

extern "C" int mrgpu_multivalue_block(mrgpu_mr *mr, int iblock, char **multivalue, int **valuesizes, int *nvalues)
{
  API_BEGIN(mr)
  ScanBlocks *sb = scan_blocks(mr, false);
  if (!sb || !sb->valid) MR_FAIL(MRGPU_EINVAL, "Invalid call to multivalue_block()");
  if (iblock < 0 || iblock >= sb->nblock) MR_FAIL(MRGPU_EINVAL, "Invalid page request to multivalue_block()");
  const uint64_t first = (uint64_t) iblock * sb->capacity;
  const uint64_t n = std::min<uint64_t>(sb->capacity, sb->nvalue_total - first);
  uint64_t off = 0;
  for (uint64_t i = 0; i < first; i++) off += (uint64_t) (unsigned) sb->sizes[i];
  if (valuesizes) *valuesizes = (int *) (sb->sizes + first);
  if (multivalue) *multivalue = (char *) sb->values + off;
  if (nvalues) *nvalues = (int) n;
  API_END(mr)
}

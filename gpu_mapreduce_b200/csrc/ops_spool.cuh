// Out-of-core tier: KeyValue / KeyMultiValue pieces in pinned HOST memory when a store outgrows its HBM budget.
//
// Reference: MR-MPI keeps one page of a KV/KMV in memory and spills the rest to files (KeyValue::create_page /
// write_page src/keyvalue.cpp:216-255,343-392; Spool src/spool.cpp:76-195); convert() of a KV that does not fit
// splits it by key hash into partitions that do (KeyMultiValue::convert src/keymultivalue.cpp:486-638:
// kv2unique spills, partition2sets :1056, one kv2kmv per set :1139).  Here HBM is the page pool and the next tier is
// pinned host memory (the box has more of it than HBM and the link is 55 GB/s; NVMe would be a third tier):
//   * setting "hbm_budget" (MB, 0 = unlimited; `outofcore` = 1 forces one page): after a map call, a KV whose
//     device part exceeds the budget moves to the host as a piece (`HostKV`); map calls with addflag keep appending
//     to a fresh device part, so a KV = host pieces in order + device tail;
//   * convert() of such a KV: every piece comes back once, is split stably by hashlittle(key) % nsets into
//     per-set slices that go back to the host; then each set -- small enough for HBM by construction -- is
//     converted by the ordinary paths and its KMV moves to the host (`HostKMV`).  Values of a key keep their arrival
//     order (pieces in order, stable split); KMV pairs come out set by set;
//   * reduce() with a stock functor walks the KMV pieces the same way; its output KV spills again if it must;
//   * statistics, pages and scan() read host pieces directly (they are already where the caller wants them);
//   * every other operation needs the whole store in HBM and brings it back first (spool_make_resident), failing
//     with the allocator's out-of-memory error if it cannot.
// Included by mrgpu.cu after ops_aggregate.cuh (it reuses the partition kernels).
#pragma once

static void host_free(void *p)
{
  if (p) cudaFreeHost(p);
}

static void *host_alloc(uint64_t bytes)
{
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 8) != cudaSuccess) {
    cudaGetLastError();
    MR_FAIL(MRGPU_ECUDA, "out of pinned host memory for the out-of-core tier (" + std::to_string(bytes) + " bytes)");
  }
  return p;
}

static void spool_free_kv(HostKV &h)
{
  host_free(h.data);
  host_free(h.offs);
  h.data = nullptr;
  h.offs = nullptr;
}

static void spool_free_kmv(HostKMV &h)
{
  host_free(h.data);
  host_free(h.offs);
  h.data = nullptr;
  h.offs = nullptr;
}

static void spool_clear_kv(mrgpu_mr *mr)
{
  for (auto &h : mr->kv_spool) spool_free_kv(h);
  mr->kv_spool.clear();
  mr->spool_kv_pairs = 0;
}

static void spool_clear_kmv(mrgpu_mr *mr)
{
  for (auto &h : mr->kmv_spool) spool_free_kmv(h);
  mr->kmv_spool.clear();
  mr->spool_kmv_pairs = 0;
}

static uint64_t spool_budget(const mrgpu_mr *mr)
{
  if (mr->hbm_budget > 0) return (uint64_t) mr->hbm_budget << 20;
  if (mr->outofcore == 1) return mr->pagesize;  // the reference's "force data out-of-core": one page in memory
  return 0;
}

// device KV -> host piece (the device store is left empty)
static void kv_to_host(mrgpu_mr *mr, DevKV *kv, HostKV &h)
{
  cudaStream_t s = mr->pool.stream;
  kv_unpack(mr, kv);  // pieces hold the reference's record stream
  h.nkv = kv->nkv, h.nbytes = kv->nbytes, h.ksize = kv->ksize, h.vsize = kv->vsize;
  h.uniform = kv->uniform, h.kb = kv->kb, h.vb = kv->vb, h.recsize = kv->recsize;
  h.data = (uint8_t *) host_alloc(h.nbytes);
  if (h.nbytes) MR_CUDA(cudaMemcpyAsync(h.data, kv->data, h.nbytes, cudaMemcpyDeviceToHost, s));
  if (!h.uniform) {
    h.offs = (uint64_t *) host_alloc((h.nkv + 1) * 8);
    MR_CUDA(cudaMemcpyAsync(h.offs, kv->offs, (h.nkv + 1) * 8, cudaMemcpyDeviceToHost, s));
  }
  sync(mr);
}

static DevKV *kv_from_host(mrgpu_mr *mr, const HostKV &h)
{
  cudaStream_t s = mr->pool.stream;
  DevKV *kv = new DevKV(&mr->pool);
  try {
    kv->reserve_bytes(h.nbytes ? h.nbytes : 512);
    if (h.nbytes) MR_CUDA(cudaMemcpyAsync(kv->data, h.data, h.nbytes, cudaMemcpyHostToDevice, s));
    kv->uniform = h.uniform, kv->kb = h.kb, kv->vb = h.vb, kv->recsize = h.recsize;
    if (!h.uniform) {
      kv->reserve_offs(h.nkv + 1);
      MR_CUDA(cudaMemcpyAsync(kv->offs, h.offs, (h.nkv + 1) * 8, cudaMemcpyHostToDevice, s));
    }
    kv->nkv = h.nkv, kv->nbytes = h.nbytes, kv->ksize = h.ksize, kv->vsize = h.vsize;
    sync(mr);
  } catch (...) {
    delete kv;
    throw;
  }
  return kv;
}

// the device part of the handle's KV becomes the next host piece
static void kv_spill(mrgpu_mr *mr)
{
  DevKV *kv = mr->kv;
  if (!kv || kv->nkv == 0) return;
  HostKV h;
  kv_to_host(mr, kv, h);
  mr->kv_spool.push_back(h);
  mr->spool_kv_pairs += h.nkv;
  delete mr->kv;
  mr->kv = new DevKV(&mr->pool);
}

// called at the end of every map-type operation
static void spool_after_map(mrgpu_mr *mr)
{
  const uint64_t budget = spool_budget(mr);
  if (budget && mr->kv && mr->kv->phys_bytes() > budget) kv_spill(mr);
}

static void kmv_to_host(mrgpu_mr *mr, DevKMV *kmv, HostKMV &h)
{
  cudaStream_t s = mr->pool.stream;
  kmv_expand(mr, kmv);
  h.nkmv = kmv->nkmv, h.nvalue = kmv->nvalue, h.ksize = kmv->ksize, h.vsize = kmv->vsize, h.nbytes = kmv->nbytes;
  h.ukb = kmv->ukb, h.uvb = kmv->uvb;
  h.data = (uint8_t *) host_alloc(h.nbytes);
  h.offs = (uint64_t *) host_alloc((h.nkmv + 1) * 8);
  if (h.nbytes) MR_CUDA(cudaMemcpyAsync(h.data, kmv->data, h.nbytes, cudaMemcpyDeviceToHost, s));
  MR_CUDA(cudaMemcpyAsync(h.offs, kmv->offs, (h.nkmv + 1) * 8, cudaMemcpyDeviceToHost, s));
  sync(mr);
}

static DevKMV *kmv_from_host(mrgpu_mr *mr, const HostKMV &h)
{
  cudaStream_t s = mr->pool.stream;
  DevKMV *kmv = new DevKMV(&mr->pool);
  try {
    kmv->data = (uint8_t *) mr->pool.alloc(h.nbytes ? h.nbytes : 16);
    kmv->cap = h.nbytes ? h.nbytes : 16;
    kmv->offs = (uint64_t *) mr->pool.alloc((h.nkmv + 1) * 8);
    kmv->offs_cap = h.nkmv + 1;
    if (h.nbytes) MR_CUDA(cudaMemcpyAsync(kmv->data, h.data, h.nbytes, cudaMemcpyHostToDevice, s));
    MR_CUDA(cudaMemcpyAsync(kmv->offs, h.offs, (h.nkmv + 1) * 8, cudaMemcpyHostToDevice, s));
    kmv->nkmv = h.nkmv, kmv->nvalue = h.nvalue, kmv->ksize = h.ksize, kmv->vsize = h.vsize, kmv->nbytes = h.nbytes;
    kmv->ukb = h.ukb, kmv->uvb = h.uvb;
    sync(mr);
  } catch (...) {
    delete kmv;
    throw;
  }
  return kmv;
}

// ---------------------------------------------------------------------------------------------------------
// operations that need the whole store in HBM: bring the pieces back (in order) in front of the device part

static void spool_make_resident(mrgpu_mr *mr)
{
  if (!mr->kv_spool.empty()) {
    DevKV *tail = mr->kv;
    DevKV *all = new DevKV(&mr->pool);
    try {
      uint64_t total = tail ? tail->nbytes : 0;
      for (auto &h : mr->kv_spool) total += h.nbytes;
      all->reserve_bytes(total ? total : 512);
      for (auto &h : mr->kv_spool) {
        DevKV *piece = kv_from_host(mr, h);
        kv_concat(mr, all, piece);
        sync(mr);
        delete piece;
      }
      if (tail && tail->nkv) {
        kv_unpack(mr, tail);
        kv_concat(mr, all, tail);
        sync(mr);
      }
    } catch (...) {
      delete all;
      throw;
    }
    delete mr->kv;
    mr->kv = all;
    spool_clear_kv(mr);
  }
  if (!mr->kmv_spool.empty()) MR_FAIL(MRGPU_ELIMIT, "this operation needs the KeyMultiValue in HBM and it is out of core (raise hbm_budget)");
}

// ---------------------------------------------------------------------------------------------------------
// stable split of a device KV (record stream) into nsets groups by hashlittle(key, keybytes, nsets) % nsets:
// sendbuf = the records grouped by set, scount / sbytes per set, ssizes = per-record sizes in sendbuf order when the
// KV is not uniform.  Same kernels as aggregate()'s general local half.
static void spool_split(mrgpu_mr *mr, DevKV *kv, int nsets, Buf<uint8_t> &sendbuf, Buf<uint32_t> &ssizes, std::vector<uint64_t> &scount,
                        std::vector<uint64_t> &sbytes)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const uint64_t n = kv->nkv;
  scount.assign(nsets, 0);
  sbytes.assign(nsets, 0);
  if (!n) return;
  sendbuf.alloc(pool, kv->nbytes);
  Buf<uint32_t> d0(pool, n), d1(pool, n), i0(pool, n), i1(pool, n);
  Buf<unsigned long long> hist(pool, 8 * RADIX);
  MR_CUDA(cudaMemsetAsync(hist.p, 0, 8 * RADIX * 8, s));
  {
    ProfScope ps(pool, "spool_split_dest", (double) kv->nbytes + 8.0 * n);
    partition_dest_kernel<<<grid_for(n, 256 * 4, 8), 256, 0, s>>>(kv->data, kv->uniform ? nullptr : kv->offs, kv->recsize, n, mr->al, (uint32_t) nsets,
                                                                d0.p, i0.p, hist.p);
    MR_CUDA(cudaGetLastError());
  }
  std::vector<unsigned long long> cnt(RADIX);
  MR_CUDA(cudaMemcpyAsync(cnt.data(), hist.p, RADIX * 8, cudaMemcpyDeviceToHost, s));
  sync(mr);
  MR_CUDA(cudaMemsetAsync(hist.p + RADIX, 0, 3 * RADIX * 8, s));
  const int cur = radix_sort_pairs<uint32_t, uint32_t, true>(pool, d0.p, d1.p, i0.p, i1.p, n, 0x1, true, hist.p);
  const uint32_t *perm = cur ? i1.p : i0.p;
  std::vector<uint64_t> srec(nsets + 1, 0);
  for (int p = 0; p < nsets; p++) {
    scount[p] = cnt[p];
    srec[p + 1] = srec[p] + cnt[p];
  }
  ProfScope ps(pool, "spool_split_pack", 2.0 * kv->nbytes + 4.0 * n);
  if (kv->uniform) {
    const uint32_t words = kv->recsize / 4;
    gather_uniform_kernel<<<(unsigned) ((n * words + 255) / 256), 256, 0, s>>>((const uint32_t *) kv->data, (uint32_t *) sendbuf.p, perm, n, words);
    MR_CUDA(cudaGetLastError());
    for (int p = 0; p < nsets; p++) sbytes[p] = scount[p] * kv->recsize;
  } else {
    ssizes.alloc(pool, n);
    Buf<uint64_t> soffs(pool, n + 1);
    record_sizes_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(kv->offs, perm, n, ssizes.p);
    MR_CUDA(cudaGetLastError());
    const uint32_t *sz = ssizes.p;
    uint64_t *so = soffs.p;
    scan_apply(
        pool, n, [=] __device__(uint64_t i) -> unsigned long long { return sz[i]; },
        [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { so[i] = e; }, mr->d_small);
    MR_CUDA(cudaMemcpyAsync(so + n, mr->d_small, 8, cudaMemcpyDeviceToDevice, s));
    gather_general_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(kv->data, kv->offs, sendbuf.p, soffs.p, perm, n);
    MR_CUDA(cudaGetLastError());
    std::vector<uint64_t> edge(nsets + 1);
    for (int p = 0; p <= nsets; p++) MR_CUDA(cudaMemcpyAsync(&edge[p], soffs.p + srec[p], 8, cudaMemcpyDeviceToHost, s));
    sync(mr);
    for (int p = 0; p < nsets; p++) sbytes[p] = edge[p + 1] - edge[p];
  }
  sync(mr);
}

// one slice of one piece that belongs to one set, on the host
struct SetSlice {
  uint8_t *data = nullptr;
  uint32_t *sizes = nullptr;  // per record, when the piece was not uniform
  uint64_t nbytes = 0, nrec = 0;
  bool uniform = true;
  uint32_t kb = 0, vb = 0, recsize = 0;
};

__global__ void __launch_bounds__(256) spool_sum_lengths_kernel(const uint8_t *__restrict__ data, const uint64_t *__restrict__ offs, uint64_t recsize, uint64_t n,
                                                                unsigned long long *ksum, unsigned long long *vsum)
{
  unsigned long long k = 0, v = 0;
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t *r = (const uint32_t *) (data + rec_off(offs, recsize, i));
    k += r[0];
    v += r[1];
  }
  for (int o = 16; o; o >>= 1) {
    k += __shfl_xor_sync(0xFFFFFFFFu, k, o);
    v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (k) atomicAdd(ksum, k);
    if (v) atomicAdd(vsum, v);
  }
}

// the KV of one set, rebuilt in HBM from its slices (in piece order: arrival order is kept)
static DevKV *spool_build_set(mrgpu_mr *mr, std::vector<SetSlice> &slices)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  uint64_t total = 0, nrec = 0;
  bool uniform = true, have = false;
  uint32_t kb = 0, vb = 0, rs = 0;
  for (auto &sl : slices) {
    if (!sl.nrec) continue;
    total += sl.nbytes, nrec += sl.nrec;
    if (!sl.uniform) uniform = false;
    else if (have && (sl.kb != kb || sl.vb != vb)) uniform = false;
    else kb = sl.kb, vb = sl.vb, rs = sl.recsize, have = true;
  }
  DevKV *kv = new DevKV(&pool);
  try {
    kv->reserve_bytes(total ? total : 512);
    uint64_t at = 0;
    for (auto &sl : slices)
      if (sl.nbytes) {
        MR_CUDA(cudaMemcpyAsync(kv->data + at, sl.data, sl.nbytes, cudaMemcpyHostToDevice, s));
        at += sl.nbytes;
      }
    kv->nkv = nrec, kv->nbytes = total;
    if (uniform) {
      kv->uniform = true, kv->kb = kb, kv->vb = vb, kv->recsize = rs;
      kv->ksize = nrec * kb, kv->vsize = nrec * vb;
    } else if (nrec) {
      kv->uniform = false;
      kv->reserve_offs(nrec + 1);
      Buf<uint32_t> sz(pool, nrec);
      uint64_t ar = 0;
      for (auto &sl : slices) {
        if (!sl.nrec) continue;
        if (sl.uniform) {
          fill_u32_kernel<<<(unsigned) ((sl.nrec + 255) / 256), 256, 0, s>>>(sz.p + ar, sl.nrec, sl.recsize);
          MR_CUDA(cudaGetLastError());
        } else
          MR_CUDA(cudaMemcpyAsync(sz.p + ar, sl.sizes, sl.nrec * 4, cudaMemcpyHostToDevice, s));
        ar += sl.nrec;
      }
      const uint32_t *psz = sz.p;
      uint64_t *oo = kv->offs;
      scan_apply(
          pool, nrec, [=] __device__(uint64_t i) -> unsigned long long { return psz[i]; },
          [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { oo[i] = e; }, mr->d_small);
      MR_CUDA(cudaMemcpyAsync(oo + nrec, mr->d_small, 8, cudaMemcpyDeviceToDevice, s));
      MR_CUDA(cudaMemsetAsync(mr->d_small + 1, 0, 16, s));
      spool_sum_lengths_kernel<<<grid_for(nrec, 256 * 4, 8), 256, 0, s>>>(kv->data, kv->offs, 0, nrec, mr->d_small + 1, mr->d_small + 2);
      MR_CUDA(cudaGetLastError());
      fetch_small(mr, 3);
      kv->ksize = mr->h_small[1], kv->vsize = mr->h_small[2];
    }
    sync(mr);
  } catch (...) {
    delete kv;
    throw;
  }
  for (auto &sl : slices) {
    host_free(sl.data);
    host_free(sl.sizes);
  }
  slices.clear();
  return kv;
}

static void convert_resident(mrgpu_mr *mr);  // ops_convert.cuh: convert of the KV that is in HBM

// convert() of a KV with host pieces (KeyMultiValue::convert's partition path, src/keymultivalue.cpp:522-638)
static void convert_spooled(mrgpu_mr *mr)
{
  const uint64_t budget = spool_budget(mr);
  kv_spill(mr);  // the device tail becomes the last piece: everything is on the host, in order
  uint64_t total = 0;
  for (auto &h : mr->kv_spool) total += h.nbytes;
  // a set must fit HBM together with convert's scratch (sort keys, payloads, the KMV): a third of the budget
  uint64_t want = budget ? (3 * total + budget - 1) / budget : 2;
  const int nsets = (int) std::min<uint64_t>(std::max<uint64_t>(want, 2), RADIX);
  std::vector<std::vector<SetSlice>> sets(nsets);
  try {
    for (auto &h : mr->kv_spool) {
      DevKV *piece = kv_from_host(mr, h);
      spool_free_kv(h);
      try {
        Buf<uint8_t> sendbuf;
        Buf<uint32_t> ssizes;
        std::vector<uint64_t> scount, sbytes;
        spool_split(mr, piece, nsets, sendbuf, ssizes, scount, sbytes);
        uint64_t ab = 0, ar = 0;
        for (int p = 0; p < nsets; p++) {
          SetSlice sl;
          sl.nbytes = sbytes[p], sl.nrec = scount[p];
          sl.uniform = piece->uniform, sl.kb = piece->kb, sl.vb = piece->vb, sl.recsize = piece->recsize;
          if (sl.nrec) {
            sl.data = (uint8_t *) host_alloc(sl.nbytes);
            MR_CUDA(cudaMemcpyAsync(sl.data, sendbuf.p + ab, sl.nbytes, cudaMemcpyDeviceToHost, mr->pool.stream));
            if (!sl.uniform) {
              sl.sizes = (uint32_t *) host_alloc(sl.nrec * 4);
              MR_CUDA(cudaMemcpyAsync(sl.sizes, ssizes.p + ar, sl.nrec * 4, cudaMemcpyDeviceToHost, mr->pool.stream));
            }
          }
          sets[p].push_back(sl);
          ab += sl.nbytes, ar += sl.nrec;
        }
        sync(mr);
      } catch (...) {
        delete piece;
        throw;
      }
      delete piece;
    }
    mr->kv_spool.clear();
    mr->spool_kv_pairs = 0;
    for (int p = 0; p < nsets; p++) {
      DevKV *setkv = spool_build_set(mr, sets[p]);
      delete mr->kv;
      mr->kv = setkv;
      if (setkv->nkv == 0) continue;
      convert_resident(mr);  // -> mr->kmv, consumes mr->kv
      HostKMV hk;
      kmv_to_host(mr, mr->kmv, hk);
      mr->kmv_spool.push_back(hk);
      mr->spool_kmv_pairs += hk.nkmv;
      delete mr->kmv;
      mr->kmv = nullptr;
    }
  } catch (...) {
    for (auto &v : sets)
      for (auto &sl : v) {
        host_free(sl.data);
        host_free(sl.sizes);
      }
    spool_clear_kv(mr);   // pieces already consumed are gone: the handle is left without a KeyValue rather than with half of one
    spool_clear_kmv(mr);
    delete mr->kv;
    mr->kv = nullptr;
    throw;
  }
  delete mr->kv;
  mr->kv = nullptr;
  mr->kmv = empty_kmv(mr);  // the KMV is the host pieces; the device part is empty
}

static void reduce_impl(mrgpu_mr *mr, int functor);

// reduce() over KMV pieces: each piece comes back, is reduced by the ordinary kernels, and its output joins the new
// KV (which spills again if it outgrows the budget)
static void reduce_spooled(mrgpu_mr *mr, int functor)
{
  const uint64_t budget = spool_budget(mr);
  DevKV *acc = new DevKV(&mr->pool);
  delete mr->kmv;  // the (empty) device part
  mr->kmv = nullptr;
  delete mr->kv;
  mr->kv = nullptr;
  try {
    for (auto &hk : mr->kmv_spool) {
      mr->kmv = kmv_from_host(mr, hk);
      spool_free_kmv(hk);
      reduce_impl(mr, functor);  // -> mr->kv, consumes mr->kmv
      DevKV *piece = mr->kv;
      mr->kv = nullptr;
      kv_unpack(mr, piece);
      kv_concat(mr, acc, piece);
      sync(mr);
      delete piece;
      if (budget && acc->phys_bytes() > budget) {
        mr->kv = acc;
        kv_spill(mr);  // leaves a fresh empty device KV in mr->kv
        acc = mr->kv;
        mr->kv = nullptr;
      }
    }
  } catch (...) {
    delete acc;
    spool_clear_kmv(mr);
    throw;
  }
  mr->kmv_spool.clear();
  mr->spool_kmv_pairs = 0;
  mr->kv = acc;
}

// ---------------------------------------------------------------------------------------------------------
// MapReduce::copy() src/mapreduce.cpp:290-340: a second object with the same settings and a deep copy of the KV or KMV

extern "C" int mrgpu_copy(mrgpu_mr *mr, mrgpu_mr *src)
{
  API_BEGIN(mr)
  if (!src || src == mr) MR_FAIL(MRGPU_EINVAL, "copy needs another MapReduce");
  if (src->device != mr->device) MR_FAIL(MRGPU_EINVAL, "copy: the two MapReduce objects live on different devices");
  freeze(src);
  if (!mr->frozen) {  // settings travel with the copy (src/mapreduce.cpp:296-312)
    mr->memsize = src->memsize, mr->keyalign = src->keyalign, mr->valuealign = src->valuealign;
    mr->mapstyle = src->mapstyle, mr->all2all = src->all2all, mr->verbosity = src->verbosity, mr->timer = src->timer;
    mr->minpage = src->minpage, mr->maxpage = src->maxpage, mr->outofcore = src->outofcore, mr->hbm_budget = src->hbm_budget;
  }
  freeze(mr);
  if (mr->al.k != src->al.k || mr->al.v != src->al.v) MR_FAIL(MRGPU_EINVAL, "copy: alignments differ");
  mr->g_what = G_NONE;
  delete mr->kv;
  delete mr->kmv;
  mr->kv = nullptr, mr->kmv = nullptr;
  spool_clear_kv(mr);
  spool_clear_kmv(mr);
  spool_make_resident(src);
  cudaStream_t s = mr->pool.stream;
  if (src->kv) {
    kv_unpack(src, src->kv);
    order_after(mr, src);
    mr->kv = new DevKV(&mr->pool);
    kv_concat(mr, mr->kv, src->kv);
  }
  if (src->kmv) {
    DevKMV *k = src->kmv;
    kmv_expand(src, k);
    order_after(mr, src);
    DevKMV *c = new DevKMV(&mr->pool);
    mr->kmv = c;
    c->data = (uint8_t *) mr->pool.alloc(k->nbytes ? k->nbytes : 16);
    c->cap = k->nbytes ? k->nbytes : 16;
    c->offs = (uint64_t *) mr->pool.alloc((k->nkmv + 1) * 8);
    c->offs_cap = k->nkmv + 1;
    if (k->nbytes) MR_CUDA(cudaMemcpyAsync(c->data, k->data, k->nbytes, cudaMemcpyDeviceToDevice, s));
    MR_CUDA(cudaMemcpyAsync(c->offs, k->offs, (k->nkmv + 1) * 8, cudaMemcpyDeviceToDevice, s));
    c->nkmv = k->nkmv, c->nvalue = k->nvalue, c->ksize = k->ksize, c->vsize = k->vsize, c->nbytes = k->nbytes;
    c->ukb = k->ukb, c->uvb = k->uvb;
  }
  sync(mr);
  API_END(mr)
}

// HBM-resident KeyValue / KeyMultiValue stores and the per-handle state.
//
// Layout in HBM.  A KeyValue is ONE contiguous stream of records in the reference's own record layout
// (int keybytes | int valuebytes | pad | key | pad | value | pad, src/keyvalue.cpp:345-390), i.e. the
// concatenation of the reference's pages without their unused tails.  The reference's page boundaries
// (greedy fill to `memsize`, src/keyvalue.cpp:362-380) are metadata computed on demand for the page
// accessors; kernels see one stream.  Two things make the stream random-accessible without parsing:
//   - "uniform" KVs (every record has the same keybytes/valuebytes -- IntCount, R-MAT edges, vertices):
//     record i starts at i * recsize, nothing else is stored;
//   - otherwise a u64 offset index offs[nkv+1] is kept beside the data.
// A KeyMultiValue is likewise one stream of records (int nvalue | int keybytes | int mvaluebytes |
// int valuesizes[nvalue] | pad | key | pad | values | pad, src/keymultivalue.cpp:830-893) plus offs[nkmv+1].
#pragma once

#include <algorithm>
#include <atomic>
#include <vector>

#include "common.cuh"

namespace mrgpu {

// aggregate()'s partition, state between its two passes (ops_partition.cuh)
struct PartPlan {
  int rw = 0, kw = 0;
  uint64_t ntiles = 0;
  Buf<unsigned> toffs;  // [nranks * ntiles] slot of each tile's first pair per destination, destination-major
  Buf<uint8_t> dest;    // destination rank of every pair (nranks <= 256), written by pass 1 for pass 2
};

struct DevKV {
  Pool *pool = nullptr;
  uint8_t *data = nullptr;
  uint64_t cap = 0;      // bytes allocated
  uint64_t nbytes = 0;   // bytes used (sum of the reference's per-page alignsize)
  uint64_t nkv = 0, ksize = 0, vsize = 0;
  bool uniform = true;   // meaningful when nkv > 0
  uint32_t kb = 0, vb = 0, recsize = 0;
  uint64_t *offs = nullptr;  // [nkv+1] when !uniform
  uint64_t offs_cap = 0;     // entries allocated
  std::vector<uint64_t> h_offs;  // host copy of offs (page table / page copies), filled on demand
  bool h_offs_valid = false;
  // Header-less form of a uniform KV (default alignment, key and value multiples of 4 bytes): the two length
  // ints of every record are the same and are not stored; data holds nkv rows of [key | value], and when every
  // pair carries the same value (IntCount's int 1) that value is kept once in cval and rows are just the key.
  // nbytes/recsize keep describing the reference's record stream, which kv_unpack() writes out whenever its
  // bytes are needed (pages, device views, ops without a header-less kernel).
  bool packed = false;
  bool cval_on = false;
  uint32_t cval[2] = {0, 0};
  // Weighted form (what aggregate()'s combiner delivers, ops_aggregate.cuh): pairs that share one value cannot be
  // told apart once their keys are equal, so a run of m equal pairs is kept as ONE row [key | m].  packed and
  // cval_on are set; data holds nrows rows of kb + 4 bytes; nkv = sum of the multiplicities, and nbytes / ksize /
  // vsize / the page table describe the nkv records of the reference's stream, which kv_unpack() writes out when
  // its bytes are needed.  Rows are grouped by source rank; within a source they are in ascending key order.
  bool weighted = false;
  uint64_t nrows = 0;
  // Work done ahead at emit time (map_intcount on one rank, ops_map.cuh): the pass that wrote the key column
  // also counted the keys into the handle's table, which is convert's first pass.  Valid only while nkv == ahead_n.
  uint64_t uid = 0;            // identity of this KV among the handle's stores
  uint64_t ahead_n = 0;
  bool counted = false, counted_bad = false;
  uint64_t counted_bits = 0;
  // ... or, with several ranks, aggregate()'s first pass (destination of every pair, per-tile counts)
  PartPlan *plan = nullptr;
  int plan_nranks = 0;
  std::vector<uint64_t> plan_scount;
  // The pairs are known to be in ascending order under sort_keys(sorted_flag) (1: int keys, 2: uint64_t keys): set by
  // reduce() when its input came from the counting convert, which emits keys in ascending order.  Valid only while the
  // store is the same (nkv and data pointer as recorded): sort_keys(flag) is then the identity and returns at once.
  int sorted_flag = 0;
  uint64_t sorted_n = 0;
  const uint8_t *sorted_ptr = nullptr;
  void mark_sorted(int flag)
  {
    sorted_flag = flag;
    sorted_n = nkv;
    sorted_ptr = data;
  }
  bool known_sorted(int flag) const { return flag > 0 && sorted_flag == flag && sorted_n == nkv && sorted_ptr == data && uniform; }
  uint32_t prow() const { return weighted ? kb + 4 : kb + (cval_on ? 0 : vb); }                  // bytes per stored row
  uint64_t phys_bytes() const { return packed ? (weighted ? nrows : nkv) * prow() : nbytes; }   // bytes of data in use

  explicit DevKV(Pool *p) : pool(p)
  {
    static std::atomic<uint64_t> next_uid{0};
    uid = ++next_uid;
  }
  ~DevKV();
  void reserve_bytes(uint64_t need)
  {
    if (need <= cap) return;
    uint64_t ncap = cap ? cap : 0;
    ncap = need > 2 * ncap ? need : 2 * ncap;
    // large stores are sized exactly when first filled; one that is being appended to (map with addflag, add,
    // gather) grows by half again so that a run of appends does not copy the whole store every time
    // (doubling up to 4 GiB: a file-by-file map copies the store once over in total instead of twice)
    if (need > (1ull << 28)) ncap = (cap && nbytes) ? std::max<uint64_t>(need, need > (4ull << 30) ? cap + cap / 2 : 2 * cap) : need;
    if (data && phys_bytes()) {  // the bytes behind the store are free: grow in place, nothing is copied
      if (pool->try_extend(data, ncap)) {
        cap = ncap;
        return;
      }
      if (ncap > need && pool->try_extend(data, need)) {
        cap = need;
        return;
      }
    }
    uint8_t *nd = (uint8_t *) pool->alloc(ncap);
    if (phys_bytes()) {
      ProfScope ps(*pool, "kv_grow", 2.0 * phys_bytes(), 0);
      MR_CUDA(cudaMemcpyAsync(nd, data, phys_bytes(), cudaMemcpyDeviceToDevice, pool->stream));
    }
    pool->free(data, cap);
    data = nd;
    cap = ncap;
  }
  void reserve_offs(uint64_t entries)
  {
    if (entries <= offs_cap) return;
    uint64_t ncap = entries > 2 * offs_cap ? entries : 2 * offs_cap;
    if (entries > (1ull << 25)) ncap = (offs_cap && nkv) ? std::max<uint64_t>(entries, entries > (1ull << 29) ? offs_cap + offs_cap / 2 : 2 * offs_cap) : entries;
    if (offs && !uniform && nkv) {
      if (pool->try_extend(offs, ncap * 8)) {
        offs_cap = ncap;
        return;
      }
      if (ncap > entries && pool->try_extend(offs, entries * 8)) {
        offs_cap = entries;
        return;
      }
    }
    uint64_t *no = (uint64_t *) pool->alloc(ncap * 8);
    if (offs && !uniform && nkv) {
      ProfScope ps(*pool, "kv_grow", 16.0 * (nkv + 1), 0);
      MR_CUDA(cudaMemcpyAsync(no, offs, (nkv + 1) * 8, cudaMemcpyDeviceToDevice, pool->stream));
    }
    pool->free(offs, offs_cap * 8);
    offs = no;
    offs_cap = ncap;
  }
};

inline DevKV::~DevKV()
{
  delete plan;
  pool->free(data, cap);
  pool->free(offs, offs_cap * 8);
}

struct DevKMV {
  Pool *pool = nullptr;
  uint8_t *data = nullptr;
  uint64_t cap = 0, nbytes = 0;
  uint64_t nkmv = 0, nvalue = 0, ksize = 0, vsize = 0;
  uint64_t *offs = nullptr;  // [nkmv+1]
  uint64_t offs_cap = 0;
  int64_t ukb = -1, uvb = -1;  // keybytes / valuebytes shared by every pair, or -1
  std::vector<uint64_t> h_offs;
  bool h_offs_valid = false;
  // Factored form (convert's dense path): when all pairs of a key are indistinguishable the KMV is fully
  // described by (key, number of values) per distinct key plus the one shared value.  The record stream
  // (data/offs above) is then written only when somebody needs its bytes (kmv_expand: pages, device views, a
  // functor that reads multivalues); reduce's stock functors read this form directly.
  bool factored = false;
  bool ascending = false;  // factored keys are in ascending order as unsigned integers (the counting convert's table order)
  uint32_t *fkey = nullptr, *fcnt = nullptr;  // keys: [nkmv * fkw] words (ukb = 4 * fkw bytes each); counts: [nkmv]
  uint32_t fkw = 1;
  uint64_t fcap = 0;                          // pairs allocated
  uint32_t fval[2] = {0, 0};  // the shared value (uvb bytes)
  // ... and its variable-length-key form (convert's hash path, ops_convert_hash.cuh; every value is empty):
  // key u = vkey[vkoff[u] .. vkoff[u+1]), fcnt[u] pairs
  bool fvar = false;
  uint8_t *vkey = nullptr;
  uint64_t *vkoff = nullptr;

  explicit DevKMV(Pool *p) : pool(p) {}
  ~DevKMV()
  {
    pool->free(data, cap);
    pool->free(offs, offs_cap * 8);
    pool->free(fkey, fcap * 4 * fkw);
    pool->free(fcnt, fcap * 4);
    pool->free(vkey, 0);
    pool->free(vkoff, 0);
  }
};

// Out-of-core tier (ops_spool.cuh): a piece of a KeyValue / KeyMultiValue that lives in pinned host memory, in the
// reference's record-stream layout (what a page copy returns), with the offset index when records differ in size
struct HostKV {
  uint8_t *data = nullptr;
  uint64_t *offs = nullptr;  // [nkv + 1], nullptr when uniform
  uint64_t nbytes = 0, nkv = 0, ksize = 0, vsize = 0;
  bool uniform = true;
  uint32_t kb = 0, vb = 0, recsize = 0;
};
struct HostKMV {
  uint8_t *data = nullptr;
  uint64_t *offs = nullptr;  // [nkmv + 1]
  uint64_t nbytes = 0, nkmv = 0, nvalue = 0, ksize = 0, vsize = 0;
  int64_t ukb = -1, uvb = -1;
};

// host-side staging of pairs added one by one (KeyValue::add from a host callback)
struct Staging {
  bool open = false;
  std::vector<uint8_t> data;
  std::vector<uint64_t> offs;  // record starts, plus end
  uint64_t nkv = 0, ksize = 0, vsize = 0;
  bool uniform = true;
  uint32_t kb = 0, vb = 0;
  void clear()
  {
    data.clear();
    offs.clear();
    nkv = ksize = vsize = 0;
    uniform = true;
    kb = vb = 0;
  }
};

}  // namespace mrgpu

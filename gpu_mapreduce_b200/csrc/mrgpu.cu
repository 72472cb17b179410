// C-ABI entry points (include/mrgpu.h) and the host orchestration of the sm_100a kernels.
// One translation unit; kernels live in the .cuh files next to it.
//
// Reference map (paths relative to the reference tree):
//   map emit / KV packing      src/keyvalue.cpp:343-428      -> ops_map.cuh
//   convert (group by key)     src/keymultivalue.cpp:486-1209 -> ops_convert.cuh (radix.cuh, scan.cuh)
//   reduce                     src/mapreduce.cpp:1769-1867   -> ops_reduce.cuh
//   sort_keys                  src/mapreduce.cpp:2102-2633   -> ops_sort.cuh
//   aggregate + Irregular      src/mapreduce.cpp:385-563, src/irregular.cpp:95-301 -> ops_aggregate.cuh
#include "../../include/mrgpu.h"

#include <dlfcn.h>

#include <algorithm>
#include <exception>
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "hash.cuh"
#include "radix.cuh"
#include "scan.cuh"
#include "flagscan.cuh"
#include "merge.cuh"
#include "store.cuh"

using namespace mrgpu;

static thread_local std::string g_create_error;

struct mrgpu_mr {
  Pool pool;
  int device = 0;
  int64_t memsize = 64, keyalign = 4, valuealign = 4, verbosity = 0, timer = 0, all2all = 1, mapstyle = 0, minpage = 0,
          maxpage = 0, freepage = 1, outofcore = 0, zeropage = 0;  // defaults(): src/mapreduce.cpp:196-262
  // compact = 1 (default): stores may keep a lossless compact form (factored KMV, header-less uniform KV) and
  // write the reference's record stream only when its bytes are asked for; 0: always the record stream
  int64_t compact = 1;
  Align al;
  bool frozen = false;
  uint64_t pagesize = 64ull << 20;
  DevKV *kv = nullptr;
  DevKMV *kmv = nullptr;
  Staging stg;
  int stg_addflag = 0;
  std::string err;
  int rank = 0, nranks = 1;
  void *comm = nullptr;            // ncclComm_t
  unsigned long long *d_small = nullptr;  // [64] device scalars (totals, flags)
  unsigned long long *h_small = nullptr;  // pinned mirror
  unsigned *dense_table = nullptr;  // convert's 2^26-slot count table: allocated at first use, all zero between calls
  cudaStream_t copy_stream = nullptr;  // host input: DMA on this stream overlaps the kernels that consume the chunks
  cudaEvent_t copy_ev[2] = {nullptr, nullptr};
  uint64_t dense_owner = 0;         // uid of the KV whose counts map_intcount already put in the table (0 = none)
  int64_t ahead = 1;                // setting: map_intcount may do the next op's first pass while it writes the keys
  // setting "l2pin": MiB of convert's count table pinned in L2 (persisting access-policy window on the handle's stream
  // + device-wide persisting set-aside) for the duration of one count launch; 0 = never touch that device state
  int64_t l2pin = getenv("MRGPU_DENSE_L2PIN") ? atol(getenv("MRGPU_DENSE_L2PIN")) : 56;
  // setting "combine": aggregate() ships one row [key | count] per distinct key instead of every pair when the pairs
  // cannot be told apart (ops_aggregate.cuh); 0 = always the plain exchange in the reference's delivery order
  int64_t combine = 1;
  // setting "presorted": convert() looks for a long ascending run among the sort keys (the previous round's output of
  // an iterative job) and, if one covers half of them, sorts only the rest and merges (merge.cuh); 0 = always sort
  int64_t presorted = 1;
  // which bytes of a 16-byte uniform key varied in the last convert(): the next one extracts its sort keys on that
  // guess while it checks it, instead of scanning the keys beforehand (ops_convert.cuh)
  int guess_nvb = -1;
  unsigned char guess_pos[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  // setting "hbm_budget" (MB, 0 = unlimited): a KeyValue whose device part outgrows it moves to pinned host memory
  // piece by piece; convert / reduce / stats / pages / scan work on such stores (ops_spool.cuh)
  int64_t hbm_budget = 0;
  std::vector<HostKV> kv_spool;     // host pieces that precede the device part mr->kv, in order
  std::vector<HostKMV> kmv_spool;   // host pieces of the KMV (the device part mr->kmv is then empty)
  uint64_t spool_kv_pairs = 0, spool_kmv_pairs = 0;
  int64_t debug_fail_alloc = -1;    // test hook: the n-th next pool allocation fails (-1 = off); mirrors pool.fail_after
  cudaEvent_t order_ev = nullptr;   // orders this handle's stream behind another handle's (add, map(mr))
  // The global pair count of the current store when the last collective call computed it (1 = of the KV, 2 = of the
  // KMV, 0 = unknown).  Ops that cannot change it (aggregate: pairs only move; sort; reduce with a stock functor: one
  // pair per KMV pair) return it without another all-reduce.  Evolves identically on every rank because every rank
  // makes the same calls in the same order (SURVEY.md 8b "Threading").
  int g_what = 0;
  uint64_t g_count = 0;
  bool p2p_used = false;  // set identically on every rank
  std::set<std::string> ipc_agreed;        // chunks (by IPC handle) that every rank is known to have mapped
  std::map<std::string, void *> ipc_open;  // peers' receive chunks mapped by the peer-store exchange (CUDA IPC handle -> base)
};

// A failed NCCL call leaves the communicator unusable and the peers inside a collective.  The reference ends the
// job there (Error::one -> MPI_Abort, src/error.cpp:47-57); the library's part of that is ncclCommAbort, which
// releases this rank's side so that nothing of it keeps the peers' kernels spinning; the caller (the C++ mirror's
// Error::one, or the launcher when the process exits non-zero) ends the job.
static void comm_abort_on_error(mrgpu_mr *mr);

#define API_BEGIN(mr)                     \
  if (!(mr)) return MRGPU_EINVAL;         \
  try {                                   \
    MR_CUDA(cudaSetDevice((mr)->device));

#define API_END(mr)                       \
  }                                       \
  catch (const Error &e)                  \
  {                                       \
    (mr)->err = e.msg;                    \
    (mr)->g_what = 0;                     \
    if (e.code == MRGPU_ECOMM) comm_abort_on_error(mr); \
    return e.code;                        \
  }                                       \
  catch (const std::exception &e)         \
  {                                       \
    (mr)->err = e.what();                 \
    return MRGPU_EINVAL;                  \
  }                                       \
  return MRGPU_OK;

// freeze alignment + page size at first use, like MapReduce::allocate() src/mapreduce.cpp:3318-3357
static void freeze(mrgpu_mr *mr)
{
  if (mr->frozen) return;
  auto pow2 = [](int64_t a) { return a >= 1 && (a & (a - 1)) == 0; };
  if (!pow2(mr->keyalign) || !pow2(mr->valuealign)) MR_FAIL(MRGPU_EINVAL, "Invalid alignment setting");
  if (mr->memsize == 0) MR_FAIL(MRGPU_EINVAL, "Invalid memsize setting");
  if (mr->minpage < 0) MR_FAIL(MRGPU_EINVAL, "Invalid minpage setting");
  if (mr->maxpage && mr->maxpage < mr->minpage) MR_FAIL(MRGPU_EINVAL, "Invalid maxpage setting");
  mr->al.k = (int) mr->keyalign;
  mr->al.v = (int) mr->valuealign;
  mr->al.t = std::max(std::max(mr->al.k, mr->al.v), 4);
  mr->pagesize = mr->memsize > 0 ? (uint64_t) mr->memsize << 20 : (uint64_t) (-mr->memsize);
  if (mr->pagesize < 512) MR_FAIL(MRGPU_EINVAL, "Page size smaller than ALIGNFILE");
  mr->frozen = true;
}

static void sync(mrgpu_mr *mr) { MR_CUDA(cudaStreamSynchronize(mr->pool.stream)); }

// read k device scalars d_small[0..k) to the host
static void fetch_small(mrgpu_mr *mr, int k)
{
  MR_CUDA(cudaMemcpyAsync(mr->h_small, mr->d_small, 8 * k, cudaMemcpyDeviceToHost, mr->pool.stream));
  sync(mr);
}

static bool is_device_ptr(const void *p)
{
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// global sum over ranks of a per-rank count (the MPI_Allreduce at the end of every reference op)
static uint64_t allreduce_sum(mrgpu_mr *mr, uint64_t v);
enum { G_NONE = 0, G_KV = 1, G_KMV = 2 };
// pairs of the handle's KV / KMV on this rank: host pieces (out-of-core tier) + the part in HBM
static uint64_t kv_total(const mrgpu_mr *mr) { return (mr->kv ? mr->kv->nkv : 0) + mr->spool_kv_pairs; }
static uint64_t kmv_total(const mrgpu_mr *mr) { return (mr->kmv ? mr->kmv->nkmv : 0) + mr->spool_kmv_pairs; }
static void spool_clear_kv(mrgpu_mr *mr);
static void spool_clear_kmv(mrgpu_mr *mr);
static void spool_after_map(mrgpu_mr *mr);
static void spool_make_resident(mrgpu_mr *mr);
static void convert_spooled(mrgpu_mr *mr);
static void reduce_spooled(mrgpu_mr *mr, int functor);
static uint64_t global_count(mrgpu_mr *mr, uint64_t local, int what, bool reuse)
{
  if (reuse && mr->g_what == what) return mr->g_count;
  mr->g_what = G_NONE;
  mr->g_count = allreduce_sum(mr, local);
  mr->g_what = what;
  return mr->g_count;
}

// Two handles may have been created on different streams.  Work that `mr` is about to enqueue and that reads
// memory written by work already enqueued on `other`'s stream is ordered behind it with an event (no host wait).
static void order_after(mrgpu_mr *mr, mrgpu_mr *other)
{
  if (mr == other || mr->pool.stream == other->pool.stream) return;
  if (!mr->order_ev) MR_CUDA(cudaEventCreateWithFlags(&mr->order_ev, cudaEventDisableTiming));
  MR_CUDA(cudaEventRecord(mr->order_ev, other->pool.stream));
  MR_CUDA(cudaStreamWaitEvent(mr->pool.stream, mr->order_ev, 0));
}

#include "ops_map.cuh"
#include "ops_convert.cuh"
#include "ops_convert_dense.cuh"
#include "ops_convert_hash.cuh"
#include "ops_text.cuh"
#include "ops_reduce.cuh"
#include "ops_sort.cuh"
#include "ops_shape.cuh"
#include "ops_partition.cuh"
#include "ops_aggregate.cuh"
#include "ops_spool.cuh"
#include "ops_pages.cuh"
#include "ops_scan.cuh"

// ================================================================================================
// lifetime, settings

extern "C" int mrgpu_create(int device, void *stream, mrgpu_mr **out)
{
  if (!out) return MRGPU_EINVAL;
  *out = nullptr;
  mrgpu_mr *mr = nullptr;
  try {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
      cudaGetLastError();
      MR_FAIL(MRGPU_ECUDA, "no CUDA device: this library has no CPU path");
    }
    if (device < 0 || device >= ndev) MR_FAIL(MRGPU_EINVAL, "invalid device ordinal");
    MR_CUDA(cudaSetDevice(device));
    mr = new mrgpu_mr();
    mr->device = device;
    mr->pool.init(device, (cudaStream_t) stream);
    MR_CUDA(cudaMalloc(&mr->d_small, 64 * 8));
    MR_CUDA(cudaMallocHost(&mr->h_small, 64 * 8));
    *out = mr;
  } catch (const Error &e) {
    g_create_error = e.msg;
    delete mr;
    return e.code;
  }
  return MRGPU_OK;
}

static int destroy_impl(mrgpu_mr *mr, bool collective)
{
  if (!mr) return MRGPU_EINVAL;
  cudaSetDevice(mr->device);
  delete mr->kv;
  delete mr->kmv;
  cudaStreamSynchronize(mr->pool.stream);
  if (mr->comm) {
    // like MapReduce::~MapReduce's MPI_Comm_free (src/mapreduce.cpp:283), collective once ranks share memory:
    // unmap the peers' buffers, then wait until every peer has unmapped mine before the pool frees it.
    // mrgpu_abandon skips the barrier: my mappings are closed, and the chunks peers may still map are never freed.
    if (mr->p2p_used) {
      if (collective) peer_store_teardown(mr);
      else {
        peer_store_unmap(mr);
        mr->pool.abandon = true;
      }
    }
    comm_destroy(mr);
  }
  scan_forget(mr);
  spool_clear_kv(mr);
  spool_clear_kmv(mr);
  if (mr->order_ev) cudaEventDestroy(mr->order_ev);
  if (mr->copy_stream) cudaStreamDestroy(mr->copy_stream);
  for (auto e : mr->copy_ev)
    if (e) cudaEventDestroy(e);
  cudaFree(mr->d_small);
  cudaFreeHost(mr->h_small);
  delete mr;
  return MRGPU_OK;
}

extern "C" int mrgpu_destroy(mrgpu_mr *mr) { return destroy_impl(mr, true); }
extern "C" int mrgpu_abandon(mrgpu_mr *mr) { return destroy_impl(mr, false); }

extern "C" const char *mrgpu_last_error(const mrgpu_mr *mr) { return mr ? mr->err.c_str() : g_create_error.c_str(); }

static int64_t *setting(mrgpu_mr *mr, const char *name)
{
  struct {
    const char *n;
    int64_t *p;
  } tab[] = {{"memsize", &mr->memsize},     {"keyalign", &mr->keyalign},   {"valuealign", &mr->valuealign},
             {"verbosity", &mr->verbosity}, {"timer", &mr->timer},         {"all2all", &mr->all2all},
             {"mapstyle", &mr->mapstyle},   {"minpage", &mr->minpage},     {"maxpage", &mr->maxpage},
             {"freepage", &mr->freepage},   {"outofcore", &mr->outofcore}, {"zeropage", &mr->zeropage},   {"compact", &mr->compact},     {"ahead", &mr->ahead},
             {"l2pin", &mr->l2pin},         {"combine", &mr->combine},     {"hbm_budget", &mr->hbm_budget}, {"presorted", &mr->presorted},
                    {"debug_fail_alloc", &mr->debug_fail_alloc}};
  for (auto &t : tab)
    if (!strcmp(t.n, name)) return t.p;
  return nullptr;
}

extern "C" int mrgpu_set(mrgpu_mr *mr, const char *name, int64_t value)
{
  API_BEGIN(mr)
  int64_t *p = setting(mr, name);
  if (!p) MR_FAIL(MRGPU_EINVAL, std::string("unknown setting ") + name);
  if (mr->frozen && (!strcmp(name, "keyalign") || !strcmp(name, "valuealign") || !strcmp(name, "memsize")) && *p != value)
    MR_FAIL(MRGPU_EINVAL, "memsize/keyalign/valuealign are frozen after first use (src/mapreduce.cpp:3318)");
  if (!strcmp(name, "l2pin") && (value < 0 || value > 1024)) MR_FAIL(MRGPU_EINVAL, "l2pin is a size in MiB between 0 and 1024");
  *p = value;
  if (p == &mr->debug_fail_alloc) mr->pool.fail_after = value;
  API_END(mr)
}

extern "C" int mrgpu_get(const mrgpu_mr *mr, const char *name, int64_t *value)
{
  if (!mr || !value || !name) return MRGPU_EINVAL;
  // read-only: the reference's msize / msizemax (src/mapreduce.h:54), in bytes of HBM held by the handle's page pool
  if (!strcmp(name, "msize")) return *value = (int64_t) mr->pool.live, MRGPU_OK;
  if (!strcmp(name, "msizemax")) return *value = (int64_t) mr->pool.hiwater, MRGPU_OK;
  if (!strcmp(name, "reserved")) return *value = (int64_t) mr->pool.reserved, MRGPU_OK;
  int64_t *p = setting(const_cast<mrgpu_mr *>(mr), name);
  if (!p) return MRGPU_EINVAL;
  *value = *p;
  return MRGPU_OK;
}

extern "C" int mrgpu_comm_rank(const mrgpu_mr *mr, int *rank, int *nranks)
{
  if (!mr) return MRGPU_EINVAL;
  if (rank) *rank = mr->rank;
  if (nranks) *nranks = mr->nranks;
  return MRGPU_OK;
}

// ================================================================================================
// per-kernel timing (the reference's `timer` setting prints per-op seconds, src/mapreduce.cpp:3112-3128)

extern "C" int mrgpu_prof(mrgpu_mr *mr, int enable)
{
  API_BEGIN(mr)
  sync(mr);
  mr->pool.prof.reset();
  mr->pool.prof.on = enable != 0;
  API_END(mr)
}

extern "C" int mrgpu_prof_report(mrgpu_mr *mr, char *buf, uint64_t cap, uint64_t *launches)
{
  API_BEGIN(mr)
  sync(mr);
  if (launches) *launches = mr->pool.prof.launches;
  struct Acc {
    const char *tag;
    double ms, bytes;
    int launches, calls;
    double moved;
  };
  std::vector<Acc> acc;
  for (auto &e : mr->pool.prof.ev) {
    float ms = 0;
    MR_CUDA(cudaEventElapsedTime(&ms, e.a, e.b));
    Acc *a = nullptr;
    for (auto &x : acc)
      if (!strcmp(x.tag, e.tag)) a = &x;
    if (!a) {
      acc.push_back(Acc{e.tag, 0, 0, 0, 0, 0});
      a = &acc.back();
    }
    a->ms += ms;
    a->bytes += e.bytes;
    a->moved += e.moved;
    a->launches += e.launches;
    a->calls++;
  }
  std::string js = "[";
  for (size_t i = 0; i < acc.size(); i++) {
    char line[256];
    snprintf(line, sizeof(line), "%s{\"tag\": \"%s\", \"ms\": %.6f, \"bytes\": %.0f, \"moved\": %.0f, \"launches\": %d, \"calls\": %d}",
             i ? ", " : "", acc[i].tag, acc[i].ms, acc[i].bytes, acc[i].moved, acc[i].launches, acc[i].calls);
    js += line;
  }
  js += "]";
  if (buf && cap) {
    size_t m = js.size() < cap - 1 ? js.size() : cap - 1;
    memcpy(buf, js.data(), m);
    buf[m] = 0;
  }
  API_END(mr)
}

// ================================================================================================
// host staging: KeyValue::add from a host callback

extern "C" int mrgpu_kv_open(mrgpu_mr *mr, int addflag)
{
  API_BEGIN(mr)
  freeze(mr);
  if (mr->stg.open) MR_FAIL(MRGPU_EINVAL, "KeyValue already open");
  mr->stg.clear();
  mr->stg.open = true;
  mr->stg_addflag = addflag;
  API_END(mr)
}

extern "C" int mrgpu_kv_add(mrgpu_mr *mr, uint64_t n, const void *keys, const int32_t *keybytes, int kfix, const void *values,
                            const int32_t *valuebytes, int vfix)
{
  API_BEGIN(mr)
  if (!mr->stg.open) MR_FAIL(MRGPU_EINVAL, "mrgpu_kv_add outside mrgpu_kv_open/close");
  Staging &s = mr->stg;
  const uint8_t *kp = (const uint8_t *) keys, *vp = (const uint8_t *) values;
  const Align al = mr->al;
  for (uint64_t i = 0; i < n; i++) {
    int64_t kb = keybytes ? keybytes[i] : kfix, vb = valuebytes ? valuebytes[i] : vfix;
    if (kb < 0 || vb < 0) MR_FAIL(MRGPU_EINVAL, "negative key/value length");
    uint64_t rs = round_up(round_up(round_up(8, al.k) + kb, al.v) + vb, al.t);
    if (rs > 0x7FFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "Single key/value pair exceeds int size");
    if (rs > mr->pagesize) MR_FAIL(MRGPU_ELIMIT, "Single key/value pair exceeds page size");
    if (s.nkv == 0) {
      s.kb = (uint32_t) kb;
      s.vb = (uint32_t) vb;
    } else if (s.kb != (uint32_t) kb || s.vb != (uint32_t) vb) s.uniform = false;
    uint64_t base = s.data.size();
    s.offs.push_back(base);
    s.data.resize(base + rs, 0);
    uint8_t *r = &s.data[base];
    int32_t k32 = (int32_t) kb, v32 = (int32_t) vb;
    memcpy(r, &k32, 4);
    memcpy(r + 4, &v32, 4);
    if (kb) memcpy(r + kv_key_off(al), kp, kb);
    if (vb) memcpy(r + kv_val_off(al, (uint32_t) kb), vp, vb);
    kp += kb;
    vp += vb;
    s.nkv++;
    s.ksize += kb;
    s.vsize += vb;
  }
  API_END(mr)
}

extern "C" int mrgpu_kv_close(mrgpu_mr *mr, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->stg.open) MR_FAIL(MRGPU_EINVAL, "KeyValue is not open");
  Staging &s = mr->stg;
  s.open = false;
  kv_begin_map(mr, mr->stg_addflag);
  DevKV *kv = mr->kv;
  if (s.nkv) {
    bool fits_uniform = s.uniform && (kv->nkv == 0 || (kv->uniform && kv->kb == s.kb && kv->vb == s.vb));
    if (!fits_uniform) kv_make_general(mr, kv);
    kv->reserve_bytes(kv->nbytes + s.data.size());
    MR_CUDA(cudaMemcpyAsync(kv->data + kv->nbytes, s.data.data(), s.data.size(), cudaMemcpyHostToDevice, mr->pool.stream));
    if (fits_uniform) {
      kv->uniform = true;
      kv->kb = s.kb;
      kv->vb = s.vb;
      kv->recsize = kv_rec_size(mr->al, s.kb, s.vb);
    } else {
      kv->reserve_offs(kv->nkv + s.nkv + 1);
      std::vector<uint64_t> o(s.nkv + 1);
      for (uint64_t i = 0; i < s.nkv; i++) o[i] = kv->nbytes + s.offs[i];
      o[s.nkv] = kv->nbytes + s.data.size();
      MR_CUDA(cudaMemcpyAsync(kv->offs + kv->nkv, o.data(), (s.nkv + 1) * 8, cudaMemcpyHostToDevice, mr->pool.stream));
      sync(mr);  // o goes out of scope
    }
    kv->nbytes += s.data.size();
    kv->nkv += s.nkv;
    kv->ksize += s.ksize;
    kv->vsize += s.vsize;
    kv->h_offs_valid = false;
  }
  sync(mr);
  s.clear();
  spool_after_map(mr);
  if (nkv) *nkv = allreduce_sum(mr, kv_total(mr));
  API_END(mr)
}

// ================================================================================================
// stand-alone kernels

extern "C" int mrgpu_sort_pairs(int device, void *stream, void *keys, void *payload, uint64_t n, int key_bytes, int payload_bytes)
{
  try {
    MR_CUDA(cudaSetDevice(device));
    if ((key_bytes != 4 && key_bytes != 8) || (payload_bytes != 0 && payload_bytes != 4 && payload_bytes != 8)) return MRGPU_EINVAL;
    if (n >= (1ull << 32)) return MRGPU_ELIMIT;
    Pool pool;
    pool.init(device, (cudaStream_t) stream);
    Buf<uint8_t> k1(pool, n * key_bytes), p1(pool, n * (payload_bytes ? payload_bytes : 1));
    Buf<unsigned long long> hist(pool, 8 * RADIX);
    int cur = 0;
#define SORT_CASE(K, P, HASP) \
  cur = radix_sort_pairs<K, P, HASP>(pool, (K *) keys, (K *) k1.p, (P *) payload, (P *) p1.p, n, 0xFF, false, hist.p)
    if (key_bytes == 4 && payload_bytes == 0) SORT_CASE(uint32_t, uint32_t, false);
    else if (key_bytes == 4 && payload_bytes == 4) SORT_CASE(uint32_t, uint32_t, true);
    else if (key_bytes == 4 && payload_bytes == 8) SORT_CASE(uint32_t, uint64_t, true);
    else if (key_bytes == 8 && payload_bytes == 0) SORT_CASE(uint64_t, uint32_t, false);
    else if (key_bytes == 8 && payload_bytes == 4) SORT_CASE(uint64_t, uint32_t, true);
    else SORT_CASE(uint64_t, uint64_t, true);
#undef SORT_CASE
    if (cur) {
      MR_CUDA(cudaMemcpyAsync(keys, k1.p, n * key_bytes, cudaMemcpyDeviceToDevice, pool.stream));
      if (payload_bytes) MR_CUDA(cudaMemcpyAsync(payload, p1.p, n * payload_bytes, cudaMemcpyDeviceToDevice, pool.stream));
    }
    k1.release();
    p1.release();
    hist.release();
    MR_CUDA(cudaStreamSynchronize(pool.stream));
  } catch (const Error &e) {
    g_create_error = e.msg;
    return e.code;
  }
  return MRGPU_OK;
}

__global__ void hashlittle_kernel(const uint8_t *keys, const uint64_t *offs, uint64_t n, uint32_t initval, uint32_t *out)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = hashlittle(keys + offs[i], (uint32_t) (offs[i + 1] - offs[i]), initval);
}

extern "C" int mrgpu_hashlittle(int device, const void *keys, const uint64_t *offsets, uint64_t n, uint32_t initval, uint32_t *out)
{
  try {
    MR_CUDA(cudaSetDevice(device));
    if (n == 0) return MRGPU_OK;
    std::vector<uint64_t> ho(n + 1);
    bool dev_in = is_device_ptr(keys);
    uint8_t *dk = nullptr;
    uint64_t *dofs = nullptr;
    uint32_t *dout = nullptr;
    uint64_t nbytes;
    if (dev_in) MR_CUDA(cudaMemcpy(&nbytes, offsets + n, 8, cudaMemcpyDeviceToHost));
    else nbytes = offsets[n];
    MR_CUDA(cudaMalloc(&dk, nbytes ? nbytes : 1));
    MR_CUDA(cudaMalloc(&dofs, (n + 1) * 8));
    MR_CUDA(cudaMalloc(&dout, n * 4));
    MR_CUDA(cudaMemcpy(dk, keys, nbytes, cudaMemcpyDefault));
    MR_CUDA(cudaMemcpy(dofs, offsets, (n + 1) * 8, cudaMemcpyDefault));
    hashlittle_kernel<<<(unsigned) ((n + 255) / 256), 256>>>(dk, dofs, n, initval, dout);
    MR_CUDA(cudaGetLastError());
    MR_CUDA(cudaMemcpy(out, dout, n * 4, cudaMemcpyDefault));
    cudaFree(dk);
    cudaFree(dofs);
    cudaFree(dout);
  } catch (const Error &e) {
    g_create_error = e.msg;
    return e.code;
  }
  return MRGPU_OK;
}

// ================================================================================================
// Sorting without allocating
//
// mrgpu_sort_pairs() takes its ping-pong buffers, histograms and look-back state from the heap of the GPU
// (cudaMalloc).  Some programs that want this sort are not in a position to do that:
//   - a host application (a database engine, an inference server, a solver) that grabbed practically all of
//     the GPU's memory for its own arena at start-up -- a driver allocation made on its behalf later can fail
//     at a moment nobody can predict, and succeeds only by fragmenting what is left;
//   - code on a latency-critical path: cudaMalloc and cudaFree synchronise the whole device, so an allocation
//     in the middle of a request stalls every other stream of the process.
// For them the scratch lives in the image of this library instead: one file-scope __device__ array, present
// on every GPU that loads libmrgpu.so from the moment the module is loaded until the process exits (about
// 16.5 MiB per GPU), big enough for kNoAllocMaxPairs pairs.  The caller still owns the inputs/outputs.

constexpr uint64_t kNoAllocMaxPairs = 1ull << 20;
constexpr int kNoAllocIpt = 16;
constexpr uint64_t kNaTiles = kNoAllocMaxPairs / ((uint64_t) SORT_THREADS * kNoAllocIpt);
constexpr uint64_t kNaOffKeys = 0;                                              // second key buffer, 8 B per pair
constexpr uint64_t kNaOffPay = kNaOffKeys + 8 * kNoAllocMaxPairs;               // second payload buffer, 8 B per pair
constexpr uint64_t kNaOffStatus = kNaOffPay + 8 * kNoAllocMaxPairs;             // look-back state, one pass at a time
constexpr uint64_t kNaOffHist = kNaOffStatus + kNaTiles * RADIX * 8;            // byte histograms / digit offsets
constexpr uint64_t kNaOffCounter = kNaOffHist + 8 * RADIX * 8;                  // tile ticket
constexpr uint64_t kNaBytes = kNaOffCounter + 256;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address a kernel can use; the
// device address is looked up once per device and kept here.
static unsigned char *g_noalloc_dev[64];
static std::mutex g_noalloc_mu;

static unsigned char *noalloc_scratch(int device)
{
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (device < 0 || device >= 64) MR_FAIL(MRGPU_EINVAL, "invalid device ordinal");
  if (!g_noalloc_dev[device]) {
    unsigned char *p = nullptr;
    MR_CUDA(cudaGetSymbolAddress((void **) &p, g_noalloc_scratch));  // the current device's copy
    g_noalloc_dev[device] = p;
  }
  return g_noalloc_dev[device];
}

template <typename K, typename P, bool HASP>
static void sort_pairs_static(cudaStream_t s, K *k0, P *p0, uint64_t n, unsigned char *scr)
{
  K *k1 = (K *) (scr + kNaOffKeys);
  P *p1 = (P *) (scr + kNaOffPay);
  unsigned long long *status = (unsigned long long *) (scr + kNaOffStatus);
  unsigned long long *hist = (unsigned long long *) (scr + kNaOffHist);
  unsigned *counter = (unsigned *) (scr + kNaOffCounter);
  const uint64_t ntiles = sort_tiles(n, kNoAllocIpt);
  // the array still holds whatever the previous call left: everything that is accumulated into starts at zero
  MR_CUDA(cudaMemsetAsync(hist, 0, sizeof(K) * RADIX * 8, s));
  radix_hist_kernel<K><<<grid_for(n, 256 * 16, 8), 256, 0, s>>>(k0, n, hist);
  MR_CUDA(cudaGetLastError());
  radix_scan_hist_kernel<<<(unsigned) sizeof(K), RADIX, 0, s>>>(hist);
  MR_CUDA(cudaGetLastError());
  // every key byte is sorted (no host round trip to find constant bytes): sizeof(K) passes, an even number,
  // so the sorted pairs end up back in the caller's arrays
  for (int b = 0; b < (int) sizeof(K); b++) {
    MR_CUDA(cudaMemsetAsync(status, 0, ntiles * RADIX * 8, s));
    MR_CUDA(cudaMemsetAsync(counter, 0, 4, s));
    const K *ki = (b & 1) ? k1 : k0;
    K *ko = (b & 1) ? k0 : k1;
    const P *pi = (b & 1) ? p1 : p0;
    P *po = (b & 1) ? p0 : p1;
    onesweep_kernel<K, P, HASP, kNoAllocIpt><<<(unsigned) ntiles, SORT_THREADS, 0, s>>>(ki, ko, pi, po, n, 8 * b, hist + b * RADIX, status, counter);
    MR_CUDA(cudaGetLastError());
  }
}

extern "C" int mrgpu_sort_pairs_noalloc(int device, void *stream, void *keys, void *payload, uint64_t n, int key_bytes, int payload_bytes)
{
  try {
    if ((key_bytes != 4 && key_bytes != 8) || (payload_bytes != 0 && payload_bytes != 4 && payload_bytes != 8))
      MR_FAIL(MRGPU_EINVAL, "mrgpu_sort_pairs_noalloc: key_bytes must be 4 or 8, payload_bytes 0, 4 or 8");
    if (n > kNoAllocMaxPairs)
      MR_FAIL(MRGPU_ELIMIT, "mrgpu_sort_pairs_noalloc: " + std::to_string(n) + " pairs exceed the built-in scratch, which holds at most " +
                                std::to_string(kNoAllocMaxPairs) + " pairs; use mrgpu_sort_pairs (it allocates) for more");
    if (n && (!keys || (payload_bytes && !payload))) MR_FAIL(MRGPU_EINVAL, "mrgpu_sort_pairs_noalloc: null pointer");
    MR_CUDA(cudaSetDevice(device));
    unsigned char *scr = noalloc_scratch(device);
    cudaStream_t s = (cudaStream_t) stream;
    if (n) {
      if (key_bytes == 4 && payload_bytes == 0) sort_pairs_static<uint32_t, uint32_t, false>(s, (uint32_t *) keys, (uint32_t *) nullptr, n, scr);
      else if (key_bytes == 4 && payload_bytes == 4) sort_pairs_static<uint32_t, uint32_t, true>(s, (uint32_t *) keys, (uint32_t *) payload, n, scr);
      else if (key_bytes == 4 && payload_bytes == 8) sort_pairs_static<uint32_t, uint64_t, true>(s, (uint32_t *) keys, (uint64_t *) payload, n, scr);
      else if (key_bytes == 8 && payload_bytes == 0) sort_pairs_static<uint64_t, uint32_t, false>(s, (uint64_t *) keys, (uint32_t *) nullptr, n, scr);
      else if (key_bytes == 8 && payload_bytes == 4) sort_pairs_static<uint64_t, uint32_t, true>(s, (uint64_t *) keys, (uint32_t *) payload, n, scr);
      else sort_pairs_static<uint64_t, uint64_t, true>(s, (uint64_t *) keys, (uint64_t *) payload, n, scr);
    }
    MR_CUDA(cudaStreamSynchronize(s));
  } catch (const Error &e) {
    g_create_error = e.msg;
    return e.code;
  }
  return MRGPU_OK;
}

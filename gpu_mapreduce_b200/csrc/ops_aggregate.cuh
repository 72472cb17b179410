// aggregate(): hash-partition the KeyValue and exchange partitions between ranks (one rank = one GPU).
//
// Reference: MapReduce::aggregate src/mapreduce.cpp:385-563 computes, page by page,
// proclist[i] = hashlittle(key,keybytes,nprocs) % nprocs (:469-472), then Irregular::setup/exchange
// (src/irregular.cpp:95-301) packs pairs by destination and calls MPI_Alltoall (counts) + MPI_Alltoallv
// (payload) under a flow-control loop that keeps every message under INTMAX bytes and every receive
// under two pages.  Received pairs are ordered by source rank, original order within a source.
//
// Here: one kernel hashes every key (bit-exact lookup3), one stable single-digit radix pass groups record
// indices by destination, one gather packs the send buffer, counts travel with an ncclAllGather, payload
// with one ncclGroupStart/ncclSend/ncclRecv/ncclGroupEnd all-to-all over NVLink.  Receive buffers are
// sized from the counts (HBM holds them whole), so the reference's INTMAX / two-page throttling and its
// retry loop disappear; delivery order is the reference's.  Included by mrgpu.cu.
#pragma once

#include <nccl.h>  // types and prototypes only; the library is bound at run time with dlopen

struct NcclApi {
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *);
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*CommAbort)(ncclComm_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  const char *(*GetErrorString)(ncclResult_t);
};

static NcclApi &nccl()
{
  static NcclApi api;
  if (!api.lib) {
    // if the host program already loaded an NCCL (torch does), dlopen returns that copy
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
      api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (!api.lib) MR_FAIL(MRGPU_ECOMM, std::string("cannot load libnccl.so.2: ") + dlerror());
#define NCCL_SYM(field, name)                                      \
  *(void **) (&api.field) = dlsym(api.lib, name);                  \
  if (!api.field) MR_FAIL(MRGPU_ECOMM, "libnccl is missing symbol " name);
    NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    NCCL_SYM(CommInitRank, "ncclCommInitRank")
    NCCL_SYM(CommDestroy, "ncclCommDestroy")
    NCCL_SYM(CommAbort, "ncclCommAbort")
    NCCL_SYM(GroupStart, "ncclGroupStart")
    NCCL_SYM(GroupEnd, "ncclGroupEnd")
    NCCL_SYM(Send, "ncclSend")
    NCCL_SYM(Recv, "ncclRecv")
    NCCL_SYM(AllGather, "ncclAllGather")
    NCCL_SYM(AllReduce, "ncclAllReduce")
    NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef NCCL_SYM
  }
  return api;
}

#define MR_NCCL(expr)                                                                                     \
  do {                                                                                                    \
    ncclResult_t _r = (expr);                                                                             \
    if (_r != ncclSuccess) throw ::mrgpu::Error{MRGPU_ECOMM, std::string(#expr " failed: ") + nccl().GetErrorString(_r)}; \
  } while (0)

// One NCCL communicator per unique id per process, shared by every handle initialised with that id -- the
// reference creates any number of MapReduce objects on one MPI communicator (oink/degree.cpp:45-46 uses two),
// while an NCCL unique id can be turned into a communicator only once.
struct CommRef {
  ncclComm_t comm;
  int refs;
};
static std::map<std::string, CommRef> g_comms;
static std::mutex g_comms_mu;

// A handle only drops its reference; the communicator itself stays cached for the life of the process, because
// the same unique id may be presented again by a later handle and NCCL accepts an id once.
static void comm_destroy(mrgpu_mr *mr)
{
  if (!mr->comm) return;
  std::lock_guard<std::mutex> lk(g_comms_mu);
  for (auto &kvp : g_comms)
    if (kvp.second.comm == (ncclComm_t) mr->comm && kvp.second.refs > 0) kvp.second.refs--;
  mr->comm = nullptr;
}

static void comm_abort_on_error(mrgpu_mr *mr)
{
  if (!mr->comm) return;
  std::lock_guard<std::mutex> lk(g_comms_mu);
  ncclComm_t c = (ncclComm_t) mr->comm;
  for (auto it = g_comms.begin(); it != g_comms.end();)
    if (it->second.comm == c) it = g_comms.erase(it);
    else ++it;
  try {
    nccl().CommAbort(c);
  } catch (...) {
  }
  mr->comm = nullptr;  // every later collective on this handle fails with "communicator not initialised"
  mr->err += " -- communicator aborted; the job cannot continue (the reference calls MPI_Abort here, src/error.cpp:47-57)";
}

extern "C" int mrgpu_comm_unique_id(void *id128)
{
  if (!id128) return MRGPU_EINVAL;
  try {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    MR_NCCL(nccl().GetUniqueId(&id));
    memcpy(id128, &id, 128);
  } catch (const Error &e) {
    g_create_error = e.msg;
    return e.code;
  }
  return MRGPU_OK;
}

extern "C" int mrgpu_comm_init(mrgpu_mr *mr, int rank, int nranks, const void *id128)
{
  API_BEGIN(mr)
  if (nranks < 1 || rank < 0 || rank >= nranks) MR_FAIL(MRGPU_EINVAL, "bad rank / nranks");
  if (nranks > 256) MR_FAIL(MRGPU_ELIMIT, "more than 256 ranks is not supported");
  if (mr->comm) comm_destroy(mr);
  mr->rank = rank;
  mr->nranks = nranks;
  if (nranks > 1) {
    if (!id128) MR_FAIL(MRGPU_EINVAL, "null unique id");
    std::lock_guard<std::mutex> lk(g_comms_mu);
    std::string key((const char *) id128, 128);
    key += std::to_string(mr->device);
    auto it = g_comms.find(key);
    if (it != g_comms.end()) {
      it->second.refs++;
      mr->comm = it->second.comm;
    } else {
      ncclUniqueId id;
      memcpy(&id, id128, 128);
      ncclComm_t c;
      MR_NCCL(nccl().CommInitRank(&c, nranks, id, rank));
      g_comms[key] = CommRef{c, 1};
      mr->comm = c;
    }
  }
  API_END(mr)
}

static uint64_t allreduce_sum(mrgpu_mr *mr, uint64_t v)
{
  if (mr->nranks == 1) return v;
  if (!mr->comm) MR_FAIL(MRGPU_ECOMM, "communicator not initialised");
  cudaStream_t s = mr->pool.stream;
  unsigned long long *d = mr->d_small + 32;
  MR_CUDA(cudaMemcpyAsync(d, &v, 8, cudaMemcpyHostToDevice, s));
  MR_NCCL(nccl().AllReduce(d, d + 1, 1, ncclUint64, ncclSum, (ncclComm_t) mr->comm, s));
  MR_CUDA(cudaMemcpyAsync(mr->h_small + 32, d + 1, 8, cudaMemcpyDeviceToHost, s));
  sync(mr);
  return mr->h_small[32];
}

// dest rank of every record, as a sort key whose only varying byte is the lowest one
__global__ void __launch_bounds__(256)
partition_dest_kernel(const uint8_t *__restrict__ data, const uint64_t *__restrict__ offs, uint64_t rs, uint64_t n, Align al, uint32_t nprocs,
                      uint32_t *__restrict__ dest, uint32_t *__restrict__ idx, unsigned long long *hist)
{
  __shared__ unsigned sh[RADIX];
  sh[threadIdx.x] = 0;
  __syncthreads();
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint8_t *r = data + rec_off(offs, rs, i);
    uint32_t kb = *(const uint32_t *) r;
    uint32_t d = hashlittle(r + kv_key_off(al), kb, nprocs) % nprocs;  // src/mapreduce.cpp:472
    dest[i] = d;
    idx[i] = (uint32_t) i;
    atomicAdd(&sh[d], 1u);
  }
  __syncthreads();
  if (sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], (unsigned long long) sh[threadIdx.x]);
}

__global__ void fill_u32_kernel(uint32_t *p, uint64_t n, uint32_t v)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

__global__ void record_sizes_kernel(const uint64_t *__restrict__ offs, const uint32_t *__restrict__ perm, uint64_t n, uint32_t *__restrict__ sizes)
{
  uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sizes[i] = (uint32_t) (offs[perm[i] + 1] - offs[perm[i]]);
}

__global__ void sum_kv_bytes_kernel(const uint8_t *__restrict__ data, const uint64_t *__restrict__ offs, uint64_t n, unsigned long long *ksum,
                                    unsigned long long *vsum)
{
  unsigned long long k = 0, v = 0;
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t *r = (const uint32_t *) (data + offs[i]);
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

// Host-side exchange plan (the part of Irregular::setup src/irregular.cpp:95-242 that survives): from the
// all-gathered table -- row q = rank q's {records, bytes} for every destination, then its record shape
// {kind 0 empty / 1 variable / 2 uniform, keybytes<<32|valuebytes} -- rank `me` derives what it receives from
// each source, the displacements in source-rank order (= the reference's delivery order) and whether the
// result is still a uniform KV.  Returns 0 = nothing anywhere, 1 = variable records, 2 = uniform (shape set).
static int plan_exchange(int P, int me, const unsigned long long *table, uint64_t *rcount, uint64_t *rbytes, uint64_t *rdisp /*[P+1]*/,
                         uint64_t *rrec /*[P+1]*/, uint64_t *shape)
{
  const int W = 2 * P + 2;
  bool all_uniform = true, have_shape = false;
  unsigned long long sh0 = 0;
  rdisp[0] = rrec[0] = 0;
  for (int q = 0; q < P; q++) {
    rcount[q] = table[(size_t) q * W + 2 * me];
    rbytes[q] = table[(size_t) q * W + 2 * me + 1];
    rdisp[q + 1] = rdisp[q] + rbytes[q];
    rrec[q + 1] = rrec[q] + rcount[q];
    unsigned long long kind = table[(size_t) q * W + 2 * P], sh = table[(size_t) q * W + 2 * P + 1];
    if (kind == 1) all_uniform = false;
    if (kind == 2) {
      if (have_shape && sh != sh0) all_uniform = false;
      sh0 = sh;
      have_shape = true;
    }
  }
  if (shape) *shape = sh0;
  if (!all_uniform) return 1;
  return have_shape ? 2 : 0;
}

// the plan as a C entry point: pure host arithmetic, usable (and tested) without a GPU
extern "C" int mrgpu_plan_exchange(int nranks, int rank, const uint64_t *table, uint64_t *rcount, uint64_t *rbytes, uint64_t *rdisp,
                                   uint64_t *rrec, uint64_t *shape, int *kind)
{
  if (nranks < 1 || rank < 0 || rank >= nranks || !table || !rcount || !rbytes || !rdisp || !rrec) return MRGPU_EINVAL;
  int k = plan_exchange(nranks, rank, (const unsigned long long *) table, rcount, rbytes, rdisp, rrec, shape);
  if (kind) *kind = k;
  return MRGPU_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Peer-store exchange: partition pass 2 writes every pair straight into its destination rank's receive buffer
// over NVLink -- the compute step (hash partition) and the collective (all-to-all) are one kernel; no local
// send buffer, no separate transfer.  Needs: every rank holds uniform records of one shape (so slices are
// record-count arithmetic), and the receive buffers are mapped into every peer with CUDA IPC (one rank = one
// process).  Protocol, all on the handle's stream:
//   1. every rank allocates its receive buffer (size known from the gathered count table);
//   2. all-gather of {ok flag, IPC handle of the cudaMalloc chunk holding it, offset} -- which is also the barrier
//      that guarantees every buffer exists and its previous users have drained;
//   3. each rank maps its peers' chunks (cached per handle), agrees by all-reduce that everyone succeeded;
//   4. scatter kernel: destination d's slice = peer d's buffer + (bytes rank d receives from lower ranks);
//   5. all-reduce = barrier: when it completes every scatter kernel has finished, so the buffer is complete.
// Returns false (nothing written) when any rank cannot take part; the caller then uses ncclSend/ncclRecv.

static bool exchange_peer_store(mrgpu_mr *mr, DevKV *kv, DevKV *nk, PartPlan *pl, uint64_t n, const std::vector<unsigned long long> &table, int W,
                                const std::vector<uint64_t> &scount, const std::vector<uint64_t> &rdisp, uint64_t brecv)
{
  static const bool enabled = !(getenv("MRGPU_P2P") && atoi(getenv("MRGPU_P2P")) == 0);
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const int P = mr->nranks, me = mr->rank;
  ncclComm_t comm = (ncclComm_t) mr->comm;
  // row: [0] ok, [1..8] handle, [9] offset of my receive buffer inside the chunk
  constexpr int RW64 = 10;
  std::vector<unsigned long long> row(RW64, 0), rows((size_t) RW64 * P, 0);
  uint8_t *cbase = nullptr;
  uint64_t csize = 0;
  bool ok = enabled && (n == 0 || pl != nullptr) && pool.chunk_of(nk->data, &cbase, &csize);
  if (ok) {
    cudaIpcMemHandle_t h;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    if (cudaIpcGetMemHandle(&h, cbase) != cudaSuccess) {
      cudaGetLastError();
      ok = false;
    } else {
      memcpy(&row[1], &h, 64);
      row[9] = (unsigned long long) (nk->data - cbase);
    }
  }
  row[0] = ok ? 1 : 0;
  Buf<unsigned long long> drow(pool, RW64), drows(pool, (uint64_t) RW64 * P);
  MR_CUDA(cudaMemcpyAsync(drow.p, row.data(), RW64 * 8, cudaMemcpyHostToDevice, s));
  MR_NCCL(nccl().AllGather(drow.p, drows.p, RW64, ncclUint64, comm, s));
  MR_CUDA(cudaMemcpyAsync(rows.data(), drows.p, (size_t) RW64 * P * 8, cudaMemcpyDeviceToHost, s));
  sync(mr);
  for (int q = 0; q < P; q++)
    if (!rows[(size_t) q * RW64]) return false;  // same answer on every rank
  mr->p2p_used = true;
  // map the peers' chunks
  DestPtrs dp;
  bool opened = true;
  for (int q = 0; q < P && opened; q++) {
    uint8_t *base = nullptr;
    if (q == me) base = cbase;
    else {
      std::string key((const char *) &rows[(size_t) q * RW64 + 1], 64);
      auto it = mr->ipc_open.find(key);
      if (it != mr->ipc_open.end()) base = (uint8_t *) it->second;
      else {
        cudaIpcMemHandle_t h;
        memcpy(&h, &rows[(size_t) q * RW64 + 1], 64);
        void *ptr = nullptr;
        if (cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          cudaGetLastError();
          opened = false;
        } else {
          mr->ipc_open[key] = ptr;
          base = (uint8_t *) ptr;
        }
      }
    }
    if (!opened) break;
    // rank q receives my pairs after those of the ranks below me: sum over sources < me of bytes sent to q
    uint64_t disp = 0;
    for (int src = 0; src < me; src++) disp += table[(size_t) src * W + 2 * q + 1];
    dp.base[q] = base + rows[(size_t) q * RW64 + 9] + disp;
  }
  // Did every rank manage to map its peers?  Every rank sees the same gathered rows and has the same history, so when
  // all of these chunks were already part of a round that ended in agreement, everybody knows the answer is yes
  // (steady state: the pool hands out the same chunks step after step) and the all-reduce is skipped.
  bool known = true;
  for (int q = 0; q < P; q++)
    if (!mr->ipc_agreed.count(std::string((const char *) &rows[(size_t) q * RW64 + 1], 64))) known = false;
  if (!known) {
    if (allreduce_sum(mr, opened ? 0 : 1) != 0) return false;  // somebody could not map a peer: everyone falls back
    for (int q = 0; q < P; q++) mr->ipc_agreed.insert(std::string((const char *) &rows[(size_t) q * RW64 + 1], 64));
  }
  (void) rdisp;
  (void) brecv;
  // bytes this rank puts on / takes off NVLink (its own share stays local): the report carries both
  uint64_t wire = 0, wire_in = 0;
  for (int q = 0; q < P; q++)
    if (q != me) {
      wire += scount[q] * (kv->packed ? kv->prow() : kv->recsize);
      wire_in += table[(size_t) q * W + 2 * me + 1];
    }
  {
    ProfScope ps(pool, "aggregate_exchange_nvlink", (double) wire, 0, (double) wire_in);  // brackets the scatter + the completion barrier
    if (n) partition_scatter(mr, kv, *pl, dp, "aggregate_partition_peer_store", true);
    allreduce_sum(mr, 1);  // barrier: every rank's stores have landed
  }
  return true;
}

// Do all ranks that hold pairs hold them header-less, in the same shape, with the same shared value?
// desc = {has pairs, packed | cval_on << 1 | can_combine << 2 | kb << 8 | vb << 32, cval}; on true, desc is that of the
// agreeing ranks.  *combine (in: this rank could run the combiner; out: every rank that holds pairs can).
static bool agree_packed(mrgpu_mr *mr, DevKV *kv, unsigned long long desc[3], bool *combine)
{
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const int P = mr->nranks;
  unsigned long long mine[3] = {kv->nkv ? 1ull : 0ull, 0, 0};
  if (kv->nkv && kv->packed) {
    mine[1] = 1ull | ((unsigned long long) kv->cval_on << 1) | ((unsigned long long) (*combine ? 1 : 0) << 2) | ((unsigned long long) kv->kb << 8) |
              ((unsigned long long) kv->vb << 32);
    mine[2] = kv->cval_on ? ((unsigned long long) kv->cval[1] << 32) | kv->cval[0] : 0;
  }
  Buf<unsigned long long> dm(pool, 3), da(pool, 3ull * P);
  std::vector<unsigned long long> all(3 * (size_t) P);
  MR_CUDA(cudaMemcpyAsync(dm.p, mine, 24, cudaMemcpyHostToDevice, s));
  MR_NCCL(nccl().AllGather(dm.p, da.p, 3, ncclUint64, (ncclComm_t) mr->comm, s));
  MR_CUDA(cudaMemcpyAsync(all.data(), da.p, 24 * (size_t) P, cudaMemcpyDeviceToHost, s));
  sync(mr);
  bool have = false, ok = true, all_can = true;
  for (int q = 0; q < P && ok; q++) {
    const unsigned long long *r = &all[3 * (size_t) q];
    if (!r[0]) continue;
    if (!(r[1] & 4)) all_can = false;
    if (!(r[1] & 1)) ok = false;
    else if (!have) {
      desc[0] = 1, desc[1] = r[1], desc[2] = r[2];
      have = true;
    } else if (r[1] != desc[1] || r[2] != desc[2])
      ok = false;
  }
  *combine = ok && have && all_can;  // decided from the gathered rows only: the same answer on every rank
  return ok && have;
}

static void peer_store_unmap(mrgpu_mr *mr)
{
  for (auto &kvp : mr->ipc_open) cudaIpcCloseMemHandle(kvp.second);
  mr->ipc_open.clear();
  cudaGetLastError();
}

static void peer_store_teardown(mrgpu_mr *mr)
{
  peer_store_unmap(mr);
  try {
    allreduce_sum(mr, 1);
  } catch (...) {
  }
}

// destinations chosen by the caller (a user hash function evaluated on the host, src/mapreduce.cpp:466-470):
// the same outputs as partition_dest_kernel; a destination outside [0, nprocs) raises the flag
__global__ void __launch_bounds__(256)
user_dest_kernel(const int32_t *__restrict__ udest, uint64_t n, uint32_t nprocs, uint32_t *__restrict__ dest, uint32_t *__restrict__ idx,
                 unsigned long long *hist, unsigned long long *bad)
{
  __shared__ unsigned sh[RADIX];
  sh[threadIdx.x] = 0;
  __syncthreads();
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int32_t d = udest[i];
    const uint32_t du = (d < 0 || (uint32_t) d >= nprocs) ? 0u : (uint32_t) d;
    if (du != (uint32_t) d) *bad = 1;
    dest[i] = du;
    idx[i] = (uint32_t) i;
    atomicAdd(&sh[du], 1u);
  }
  __syncthreads();
  if (sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], (unsigned long long) sh[threadIdx.x]);
}

static void aggregate_impl(mrgpu_mr *mr, const int32_t *user_dest = nullptr)
{
  freeze(mr);
  const int P = mr->nranks;
  if (P == 1) return;  // src/mapreduce.cpp:403-406
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  if (kv->nkv >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "aggregate: more than 2^32-2 pairs on one GPU is not supported yet");
  ncclComm_t comm = (ncclComm_t) mr->comm;

  // header-less rows travel as they are when every rank that holds pairs holds the same header-less shape;
  // otherwise they are written out as records first
  if (kv->weighted) kv_unpack(mr, kv);  // a second aggregate() in a row: start from the records again
  if (kv->packed && (user_dest || !partition_eligible(mr, kv))) kv_unpack(mr, kv);

  // The combiner.  Pairs that share one value (IntCount's int 1, OINK's NULL values) cannot be told apart once their
  // keys are equal, so what the destination needs from this rank is, per distinct key, how many it had.  When every
  // rank that holds pairs holds them as 4-byte integer keys below 2^26 with one shared value, each rank counts its
  // own keys (convert's counting kernel -- or nothing at all when map_intcount already counted them while it wrote
  // them), turns the non-empty table slots into rows [key | count] and THOSE are partitioned and exchanged: Zipf's
  // hottest key crosses NVLink once per rank instead of 10^8 times, so neither the busiest port nor the rank that
  // owns the key sees the skew.  It is the reference's own remedy -- compress() before aggregate()
  // (src/mapreduce.cpp:749-851; cpu/IntCount.cpp:78 has it commented out) -- applied where it cannot change any
  // result.  The receiver holds the rows as a weighted KV (store.cuh): same pair count, statistics and page table
  // as the plain exchange; the pairs of one source come out grouped by key instead of in arrival order.
  // Setting "combine" = 0 keeps the plain exchange (exact delivery order of the reference).
  bool combine = !user_dest && mr->combine && kv->nkv && kv->packed && kv->cval_on && kv->kb == 4 && (kv->vb == 0 || kv->vb == 4 || kv->vb == 8) &&
                 ((uintptr_t) kv->data % 16) == 0;
  if (combine) {
    if (!(kv->counted && kv->ahead_n == kv->nkv && mr->dense_owner == kv->uid)) dense_count_shared4(mr, kv);
    if (kv->counted_bad) combine = false;  // a key the table cannot take
  }
  unsigned long long wire_desc[3];
  const bool packed_wire = agree_packed(mr, kv, wire_desc, &combine);
  if (kv->packed && !packed_wire) kv_unpack(mr, kv);
  DevKV *const orig = kv;
  DevKV *ck = nullptr;  // the rows [key | count] that stand in for this rank's pairs
  struct CkGuard {
    DevKV *&p;
    ~CkGuard() { delete p; }
  } ck_guard{ck};
  if (combine && kv->nkv) {
    Buf<uint2> rows;
    const uint64_t U = dense_table_to_rows(mr, kv->counted_bits, rows);
    ck = new DevKV(&pool);
    ck->data = (uint8_t *) rows.detach();
    ck->cap = (U ? U : 1) * 8;
    ck->uniform = ck->packed = true;
    ck->cval_on = false;
    ck->kb = ck->vb = 4;
    ck->recsize = kv_rec_size(al, 4, 4);
    ck->nkv = U;
    ck->nbytes = U * ck->recsize;
    ck->ksize = ck->vsize = 4 * U;
    kv = ck;
  }
  const uint64_t n = kv->nkv;
  const uint64_t row = kv->packed ? kv->prow() : kv->recsize;  // bytes per pair of a uniform KV as stored and sent

  Buf<uint8_t> sendbuf;
  Buf<uint32_t> ssizes;
  Buf<uint64_t> soffs;
  const bool uni = kv->uniform;
  std::vector<uint64_t> scount(P), sbytes(P), sdisp(P + 1, 0), srec(P + 1, 0);
  // uniform records: pass 1 of the single-pass partition (ops_partition.cuh) gives the per-destination counts now;
  // pass 2 writes the pairs later -- straight into the peers' receive buffers when the peer-store exchange is
  // possible, else into a local send buffer for NCCL
  const bool fast = n && !user_dest && partition_eligible(mr, kv);
  PartPlan pl_local;
  // pass 1 may have been done by map_intcount while it wrote this KV (emit_and_plan)
  const bool planned = fast && kv->plan && kv->ahead_n == n && kv->plan_nranks == P;
  PartPlan &pl = planned ? *kv->plan : pl_local;
  if (fast) {
    if (planned) scount = kv->plan_scount;
    else partition_counts(mr, kv, pl, scount);
    for (int p = 0; p < P; p++) {
      srec[p + 1] = srec[p] + scount[p];
      sbytes[p] = scount[p] * row;
    }
  } else {
    sendbuf.alloc(pool, kv->nbytes);
    // 1. destinations, grouped stably
    Buf<uint32_t> d0(pool, n), d1(pool, n), i0(pool, n), i1(pool, n);
    Buf<unsigned long long> hist(pool, 8 * RADIX);
    MR_CUDA(cudaMemsetAsync(hist.p, 0, 8 * RADIX * 8, s));
    unsigned long long h_bad = 0;
    if (n && user_dest) {
      MR_CUDA(cudaMemsetAsync(mr->d_small + 4, 0, 8, s));
      user_dest_kernel<<<grid_for(n, 256 * 4, 8), 256, 0, s>>>(user_dest, n, (uint32_t) P, d0.p, i0.p, hist.p, mr->d_small + 4);
      MR_CUDA(cudaGetLastError());
      MR_CUDA(cudaMemcpyAsync(&h_bad, mr->d_small + 4, 8, cudaMemcpyDeviceToHost, s));
    } else if (n) {
      ProfScope ps(pool, "aggregate_partition", (double) kv->nbytes + 8.0 * n);
      partition_dest_kernel<<<grid_for(n, 256 * 4, 8), 256, 0, s>>>(kv->data, kv->uniform ? nullptr : kv->offs, kv->recsize, n, al, (uint32_t) P, d0.p,
                                                                  i0.p, hist.p);
      MR_CUDA(cudaGetLastError());
    }
    std::vector<unsigned long long> cnt(RADIX);
    MR_CUDA(cudaMemcpyAsync(cnt.data(), hist.p, RADIX * 8, cudaMemcpyDeviceToHost, s));
    sync(mr);
    // (the reference indexes its send counts with whatever the hash returned; a rank outside the communicator is
    //  a crash there -- here the pair goes to rank 0 and the call reports the misuse after the exchange)
    if (h_bad) mr->err = "aggregate: the user hash returned a processor outside [0, nprocs)";
    // bytes 1..3 of the dest "key" are all zero: mark them as known-constant
    MR_CUDA(cudaMemsetAsync(hist.p + RADIX, 0, 3 * RADIX * 8, s));
    int cur = radix_sort_pairs<uint32_t, uint32_t, true>(pool, d0.p, d1.p, i0.p, i1.p, n, 0x1, true, hist.p);
    const uint32_t *perm = cur ? i1.p : i0.p;

    // 2. pack the send buffer in destination order
    for (int p = 0; p < P; p++) {
      scount[p] = cnt[p];
      srec[p + 1] = srec[p] + cnt[p];
    }
    if (n) {
      if (uni) {
        uint32_t words = kv->recsize / 4;
        ProfScope ps(pool, "aggregate_pack", 2.0 * kv->nbytes + 4.0 * n);
        gather_uniform_kernel<<<(unsigned) ((n * words + 255) / 256), 256, 0, s>>>((const uint32_t *) kv->data, (uint32_t *) sendbuf.p, perm, n, words);
        MR_CUDA(cudaGetLastError());
        for (int p = 0; p < P; p++) sbytes[p] = scount[p] * kv->recsize;
      } else {
        ssizes.alloc(pool, n);
        soffs.alloc(pool, n + 1);
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
        std::vector<uint64_t> edge(P + 1);
        for (int p = 0; p <= P; p++) MR_CUDA(cudaMemcpyAsync(&edge[p], soffs.p + srec[p], 8, cudaMemcpyDeviceToHost, s));
        sync(mr);
        for (int p = 0; p < P; p++) sbytes[p] = edge[p + 1] - edge[p];
      }
    }
  }
  for (int p = 0; p < P; p++) sdisp[p + 1] = sdisp[p] + sbytes[p];

  // 3. counts: every rank learns the whole P x P table (records, bytes) and every rank's record shape
  //    (replaces MPI_Alltoall + MPI_Reduce_scatter of src/irregular.cpp:144,178)
  const int W = 2 * P + 2;
  std::vector<unsigned long long> mine(W);
  for (int p = 0; p < P; p++) {
    mine[2 * p] = scount[p];
    mine[2 * p + 1] = sbytes[p];
  }
  mine[2 * P] = n == 0 ? 0 : (uni ? 2 : 1);
  mine[2 * P + 1] = uni ? (((unsigned long long) kv->kb << 32) | kv->vb) : 0;
  Buf<unsigned long long> dmine(pool, W), dall(pool, (uint64_t) W * P);
  MR_CUDA(cudaMemcpyAsync(dmine.p, mine.data(), W * 8, cudaMemcpyHostToDevice, s));
  MR_NCCL(nccl().AllGather(dmine.p, dall.p, W, ncclUint64, comm, s));
  std::vector<unsigned long long> all((size_t) W * P);
  MR_CUDA(cudaMemcpyAsync(all.data(), dall.p, (size_t) W * P * 8, cudaMemcpyDeviceToHost, s));
  sync(mr);
  std::vector<uint64_t> rcount(P), rbytes(P), rdisp(P + 1, 0), rrec(P + 1, 0);
  uint64_t shape = 0;
  int kind = plan_exchange(P, mr->rank, all.data(), rcount.data(), rbytes.data(), rdisp.data(), rrec.data(), &shape);
  const bool all_uniform = kind != 1, have_shape = kind == 2;
  const uint64_t nrecv = rrec[P], brecv = rdisp[P];
  if (nrecv >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "aggregate: a rank would receive more than 2^32-2 pairs");

  // 4. payload all-to-all over NVLink (replaces MPI_Alltoallv, src/irregular.cpp:292-300)
  DevKV *nk = new DevKV(&pool);
  try {
    nk->reserve_bytes(brecv > 512 ? brecv : 512);
    bool delivered = false;
    if (kind == 2) delivered = exchange_peer_store(mr, kv, nk, fast ? &pl : nullptr, n, all, W, scount, rdisp, brecv);
    if (!delivered && fast) {  // NCCL path needs the pairs packed locally
      sendbuf.alloc(pool, kv->phys_bytes());
      DestPtrs dp;
      for (int p = 0; p < P; p++) dp.base[p] = sendbuf.p + srec[p] * row;
      partition_scatter(mr, kv, pl, dp, "aggregate_partition_scatter");
    }
    Buf<uint32_t> rsizes;
    bool need_sizes = !all_uniform;
    if (need_sizes) {
      rsizes.alloc(pool, nrecv);
      if (uni && n) {  // peers need per-record sizes even if mine are all equal
        ssizes.alloc(pool, n);
        fill_u32_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(ssizes.p, n, kv->recsize);
        MR_CUDA(cudaGetLastError());
      }
    }
    uint64_t wire = 0, wire_in = 0;  // sent / received over NVLink
    for (int p = 0; p < P; p++)
      if (p != mr->rank) wire += sbytes[p], wire_in += rbytes[p];
    if (!delivered) {
    ProfScope psx(pool, "aggregate_exchange_nvlink", (double) wire, 1, (double) wire_in);
    MR_NCCL(nccl().GroupStart());
    for (int p = 0; p < P; p++) {
      if (p == mr->rank) continue;
      if (sbytes[p]) MR_NCCL(nccl().Send(sendbuf.p + sdisp[p], sbytes[p], ncclInt8, p, comm, s));
      if (rbytes[p]) MR_NCCL(nccl().Recv(nk->data + rdisp[p], rbytes[p], ncclInt8, p, comm, s));
      if (need_sizes) {
        if (scount[p]) MR_NCCL(nccl().Send(ssizes.p + srec[p], scount[p], ncclUint32, p, comm, s));
        if (rcount[p]) MR_NCCL(nccl().Recv(rsizes.p + rrec[p], rcount[p], ncclUint32, p, comm, s));
      }
    }
    MR_NCCL(nccl().GroupEnd());
    int me = mr->rank;
    if (sbytes[me])
      MR_CUDA(cudaMemcpyAsync(nk->data + rdisp[me], sendbuf.p + sdisp[me], sbytes[me], cudaMemcpyDeviceToDevice, s));
    if (need_sizes && scount[me])
      MR_CUDA(cudaMemcpyAsync(rsizes.p + rrec[me], ssizes.p + srec[me], scount[me] * 4, cudaMemcpyDeviceToDevice, s));
    }

    nk->nbytes = brecv;
    nk->nkv = nrecv;
    if (have_shape) {
      nk->uniform = true;
      nk->kb = (uint32_t) (shape >> 32);
      nk->vb = (uint32_t) shape;
      nk->recsize = kv_rec_size(al, nk->kb, nk->vb);
      if (packed_wire) {
        nk->packed = true;
        nk->cval_on = (wire_desc[1] >> 1) & 1;
        nk->cval[0] = (uint32_t) wire_desc[2];
        nk->cval[1] = (uint32_t) (wire_desc[2] >> 32);
        nk->nbytes = nrecv * nk->recsize;
      }
      nk->ksize = nrecv * nk->kb;
      nk->vsize = nrecv * nk->vb;
      if (combine) {
        // what arrived are rows [key | count]: the KV they stand for has the senders' shape and shared value
        nk->weighted = true;
        nk->nrows = nrecv;
        nk->kb = (uint32_t) ((wire_desc[1] >> 8) & 0xFFFFFF);
        nk->vb = (uint32_t) (wire_desc[1] >> 32);
        nk->recsize = kv_rec_size(al, nk->kb, nk->vb);
        MR_CUDA(cudaMemsetAsync(mr->d_small, 0, 8, s));
        if (nrecv) {
          sum_multiplicities_kernel<<<grid_for(nrecv, 256 * 8, 8), 256, 0, s>>>((const uint2 *) nk->data, nrecv, mr->d_small);
          MR_CUDA(cudaGetLastError());
        }
        fetch_small(mr, 1);
        nk->nkv = mr->h_small[0];
        if (nk->nkv >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "aggregate: a rank would receive more than 2^32-2 pairs");
        nk->nbytes = nk->nkv * nk->recsize;
        nk->ksize = nk->nkv * nk->kb;
        nk->vsize = nk->nkv * nk->vb;
      }
    } else if (nrecv) {
      nk->uniform = false;
      nk->reserve_offs(nrecv + 1);
      const uint32_t *sz = rsizes.p;
      uint64_t *oo = nk->offs;
      scan_apply(
          pool, nrecv, [=] __device__(uint64_t i) -> unsigned long long { return sz[i]; },
          [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { oo[i] = e; }, mr->d_small);
      MR_CUDA(cudaMemcpyAsync(oo + nrecv, mr->d_small, 8, cudaMemcpyDeviceToDevice, s));
      MR_CUDA(cudaMemsetAsync(mr->d_small + 1, 0, 16, s));
      sum_kv_bytes_kernel<<<grid_for(nrecv, 256 * 4, 8), 256, 0, s>>>(nk->data, nk->offs, nrecv, mr->d_small + 1, mr->d_small + 2);
      MR_CUDA(cudaGetLastError());
      fetch_small(mr, 3);
      nk->ksize = mr->h_small[1];
      nk->vsize = mr->h_small[2];
    }
    sync(mr);
  } catch (...) {
    delete nk;
    throw;
  }
  (void) orig;
  delete mr->kv;
  mr->kv = nk;
}

extern "C" int mrgpu_aggregate(mrgpu_mr *mr, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot aggregate without KeyValue");
  if (mr->nranks > 1 && !mr->comm) MR_FAIL(MRGPU_ECOMM, "communicator not initialised");
  if (mr->nranks > 1) spool_make_resident(mr);  // the exchange works on the store in HBM (one rank: nothing moves)
  aggregate_impl(mr);
  sync(mr);
  if (nkv) *nkv = global_count(mr, kv_total(mr), G_KV, true);  // pairs only moved: the global count is the one before
  API_END(mr)
}

// aggregate(myhash) src/mapreduce.cpp:385-563 with proclist[i] = myhash(key,keybytes) % nprocs (:466-470) evaluated by the
// CALLER: dest[i] = destination rank of pair i (host or device memory, one int per pair of this rank's KV in order).
extern "C" int mrgpu_aggregate_dest(mrgpu_mr *mr, const int32_t *dest, uint64_t ndest, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot aggregate without KeyValue");
  if (mr->nranks > 1 && !mr->comm) MR_FAIL(MRGPU_ECOMM, "communicator not initialised");
  if (ndest != kv_total(mr)) MR_FAIL(MRGPU_EINVAL, "aggregate: one destination per pair is required");
  if (ndest && !dest) MR_FAIL(MRGPU_EINVAL, "null destinations");
  if (mr->nranks > 1) {
    Buf<int32_t> tmp;
    // an empty rank still takes part in the collective calls
    static const int32_t none = 0;
    const int32_t *dd = ndest ? stage_input(mr, dest, ndest, tmp, false) : &none;
    mr->err.clear();
    spool_make_resident(mr);
    aggregate_impl(mr, dd);
    sync(mr);
    if (!mr->err.empty()) MR_FAIL(MRGPU_EINVAL, mr->err);
  }
  if (nkv) *nkv = global_count(mr, kv_total(mr), G_KV, true);
  API_END(mr)
}

extern "C" int mrgpu_collate(mrgpu_mr *mr, uint64_t *nkmv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot collate without KeyValue");
  if (mr->nranks > 1 && !mr->comm) MR_FAIL(MRGPU_ECOMM, "communicator not initialised");
  if (mr->nranks > 1) spool_make_resident(mr);
  aggregate_impl(mr);  // src/mapreduce.cpp:726-727: aggregate(hash); convert();
  mr->g_what = G_NONE;
  convert_impl(mr);
  sync(mr);
  if (nkmv) *nkmv = global_count(mr, kmv_total(mr), G_KMV, false);
  API_END(mr)
}

// ---------------------------------------------------------------------------------------------------------
// gather(numprocs)   src/mapreduce.cpp:893-1036
// Ranks >= numprocs hand their whole KV to rank (me % numprocs) and end up with an empty one; a receiver appends
// what it gets to its own pairs, senders in ascending rank order (me+numprocs, me+2*numprocs, ...), each
// sender's pairs in their order.  The reference moves page by page with a handshake per INTMAX chunk; here
// every rank learns all sizes from one all-gather and the payload moves in one grouped send/recv.
static void gather_impl(mrgpu_mr *mr, int numprocs)
{
  freeze(mr);
  const int P = mr->nranks, me = mr->rank;
  if (P == 1 || numprocs == P) return;  // :905-910
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  ncclComm_t comm = (ncclComm_t) mr->comm;
  DevKV *kv = mr->kv;
  kv_unpack(mr, kv);
  // every rank's {pairs, bytes, kind 0 empty / 1 variable / 2 uniform, keybytes<<32|valuebytes, keysize, valuesize}
  constexpr int W = 6;
  unsigned long long mine[W] = {kv->nkv, kv->nbytes, kv->nkv == 0 ? 0ull : (kv->uniform ? 2ull : 1ull),
                                kv->uniform ? (((unsigned long long) kv->kb << 32) | kv->vb) : 0ull, kv->ksize, kv->vsize};
  std::vector<unsigned long long> all((size_t) W * P);
  {
    Buf<unsigned long long> dm(pool, W), da(pool, (uint64_t) W * P);
    MR_CUDA(cudaMemcpyAsync(dm.p, mine, W * 8, cudaMemcpyHostToDevice, s));
    MR_NCCL(nccl().AllGather(dm.p, da.p, W, ncclUint64, comm, s));
    MR_CUDA(cudaMemcpyAsync(all.data(), da.p, (size_t) W * P * 8, cudaMemcpyDeviceToHost, s));
    sync(mr);
  }
  auto row = [&](int q) { return &all[(size_t) q * W]; };
  if (me >= numprocs) {
    const int dest = me % numprocs;
    if (kv->nkv) {
      MR_NCCL(nccl().GroupStart());
      MR_NCCL(nccl().Send(kv->data, kv->nbytes, ncclInt8, dest, comm, s));
      if (!kv->uniform) MR_NCCL(nccl().Send(kv->offs, kv->nkv + 1, ncclUint64, dest, comm, s));
      MR_NCCL(nccl().GroupEnd());
      sync(mr);
    }
    delete mr->kv;  // :1021-1022: the sender starts over with an empty KV
    mr->kv = new DevKV(&pool);
    return;
  }
  // receiver: what arrives, and whether the result can stay a uniform KV
  uint64_t add_n = 0, add_b = 0, add_k = 0, add_v = 0;
  bool uniform = kv->nkv == 0 || kv->uniform;
  unsigned long long shape = kv->nkv ? mine[3] : 0;
  bool have_shape = kv->nkv != 0;
  for (int q = me + numprocs; q < P; q += numprocs) {
    const unsigned long long *r = row(q);
    if (!r[0]) continue;
    add_n += r[0], add_b += r[1], add_k += r[4], add_v += r[5];
    if (r[2] != 2) uniform = false;
    else if (have_shape && r[3] != shape) uniform = false;
    else shape = r[3], have_shape = true;
  }
  if (!add_n) return;
  if (kv->nkv + add_n >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "gather: a rank would hold more than 2^32-2 pairs");
  if (uniform) {
    kv->uniform = true;
    kv->kb = (uint32_t) (shape >> 32);
    kv->vb = (uint32_t) shape;
    kv->recsize = kv_rec_size(mr->al, kv->kb, kv->vb);
  } else {
    kv_make_general(mr, kv);
    kv->reserve_offs(kv->nkv + add_n + 1);
  }
  kv->reserve_bytes(kv->nbytes + add_b);
  std::vector<Buf<uint64_t>> roffs;  // offsets of the variable-length senders, as they index their own stream
  MR_NCCL(nccl().GroupStart());
  uint64_t at_b = kv->nbytes;
  for (int q = me + numprocs; q < P; q += numprocs) {
    const unsigned long long *r = row(q);
    if (!r[0]) continue;
    MR_NCCL(nccl().Recv(kv->data + at_b, r[1], ncclInt8, q, comm, s));
    if (r[2] == 1) {
      roffs.emplace_back(pool, r[0] + 1);
      MR_NCCL(nccl().Recv(roffs.back().p, r[0] + 1, ncclUint64, q, comm, s));
    }
    at_b += r[1];
  }
  MR_NCCL(nccl().GroupEnd());
  if (!uniform) {
    uint64_t at_n = kv->nkv;
    at_b = kv->nbytes;
    size_t vi = 0;
    for (int q = me + numprocs; q < P; q += numprocs) {
      const unsigned long long *r = row(q);
      if (!r[0]) continue;
      const uint64_t cnt = r[0] + 1;
      if (r[2] == 1) shift_offsets_kernel<<<(unsigned) ((cnt + 255) / 256), 256, 0, s>>>(kv->offs + at_n, roffs[vi++].p, cnt, at_b);
      else affine_offsets_kernel<<<(unsigned) ((cnt + 255) / 256), 256, 0, s>>>(kv->offs + at_n, cnt, at_b, kv_rec_size(mr->al, (uint32_t) (r[3] >> 32), (uint32_t) r[3]));
      MR_CUDA(cudaGetLastError());
      at_n += r[0];
      at_b += r[1];
    }
  }
  kv->sorted_flag = 0;
  kv->nkv += add_n;
  kv->nbytes += add_b;
  kv->ksize += add_k;
  kv->vsize += add_v;
  kv->h_offs_valid = false;
  sync(mr);
}

extern "C" int mrgpu_gather(mrgpu_mr *mr, int numprocs, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot gather without KeyValue");
  if (numprocs < 1 || numprocs > mr->nranks) MR_FAIL(MRGPU_EINVAL, "Invalid processor count for gather");
  if (mr->nranks > 1 && !mr->comm) MR_FAIL(MRGPU_ECOMM, "communicator not initialised");
  mr->g_what = G_NONE;
  spool_make_resident(mr);
  gather_impl(mr, numprocs);
  if (nkv) *nkv = allreduce_sum(mr, mr->kv->nkv);
  API_END(mr)
}

// ---------------------------------------------------------------------------------------------------------
// broadcast(root)   src/mapreduce.cpp:569-625: every rank ends with a copy of root's KeyValue (its own is dropped).
// The reference broadcasts page by page; here root's sizes travel in one all-gather-sized message and the
// payload in one grouped send per peer.
static void broadcast_impl(mrgpu_mr *mr, int root)
{
  freeze(mr);
  const int P = mr->nranks, me = mr->rank;
  if (P == 1) return;  // :582-587
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  ncclComm_t comm = (ncclComm_t) mr->comm;
  if (me == root) kv_unpack(mr, mr->kv);
  DevKV *kv = mr->kv;
  constexpr int W = 6;  // {pairs, bytes, uniform, keybytes<<32|valuebytes, keysize, valuesize} of root
  unsigned long long h[W] = {0, 0, 0, 0, 0, 0};
  if (me == root) {
    h[0] = kv->nkv, h[1] = kv->nbytes, h[2] = (kv->nkv == 0 || kv->uniform) ? 1 : 0;
    h[3] = ((unsigned long long) kv->kb << 32) | kv->vb, h[4] = kv->ksize, h[5] = kv->vsize;
  }
  Buf<unsigned long long> d(pool, W);
  MR_CUDA(cudaMemcpyAsync(d.p, h, W * 8, cudaMemcpyHostToDevice, s));
  MR_NCCL(nccl().GroupStart());
  if (me == root) {
    for (int q = 0; q < P; q++)
      if (q != root) MR_NCCL(nccl().Send(d.p, W, ncclUint64, q, comm, s));
  } else
    MR_NCCL(nccl().Recv(d.p, W, ncclUint64, root, comm, s));
  MR_NCCL(nccl().GroupEnd());
  MR_CUDA(cudaMemcpyAsync(h, d.p, W * 8, cudaMemcpyDeviceToHost, s));
  sync(mr);
  if (me != root) {
    delete mr->kv;  // :590-592
    kv = mr->kv = new DevKV(&pool);
    kv->uniform = h[2] != 0;
    kv->kb = (uint32_t) (h[3] >> 32);
    kv->vb = (uint32_t) h[3];
    kv->recsize = kv->uniform ? kv_rec_size(mr->al, kv->kb, kv->vb) : 0;
    kv->reserve_bytes(h[1] > 512 ? h[1] : 512);
    if (!kv->uniform) kv->reserve_offs(h[0] + 1);
  }
  if (h[0]) {
    MR_NCCL(nccl().GroupStart());
    if (me == root) {
      for (int q = 0; q < P; q++) {
        if (q == root) continue;
        MR_NCCL(nccl().Send(kv->data, h[1], ncclInt8, q, comm, s));
        if (!kv->uniform) MR_NCCL(nccl().Send(kv->offs, h[0] + 1, ncclUint64, q, comm, s));
      }
    } else {
      MR_NCCL(nccl().Recv(kv->data, h[1], ncclInt8, root, comm, s));
      if (!kv->uniform) MR_NCCL(nccl().Recv(kv->offs, h[0] + 1, ncclUint64, root, comm, s));
    }
    MR_NCCL(nccl().GroupEnd());
  } else if (me != root && !kv->uniform)
    MR_CUDA(cudaMemsetAsync(kv->offs, 0, 8, s));
  if (me != root) {
    kv->nkv = h[0];
    kv->nbytes = h[1];
    kv->ksize = h[4];
    kv->vsize = h[5];
    kv->h_offs_valid = false;
  }
  sync(mr);
}

extern "C" int mrgpu_broadcast(mrgpu_mr *mr, int root, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot broadcast without KeyValue");
  if (root < 0 || root >= mr->nranks) MR_FAIL(MRGPU_EINVAL, "Invalid root for broadcast");
  if (mr->nranks > 1 && !mr->comm) MR_FAIL(MRGPU_ECOMM, "communicator not initialised");
  mr->g_what = G_NONE;
  spool_make_resident(mr);
  broadcast_impl(mr, root);
  if (nkv) *nkv = allreduce_sum(mr, mr->kv->nkv);
  API_END(mr)
}

// scrunch(numprocs, key, keybytes) = gather(numprocs) then collapse(key)   src/mapreduce.cpp:2075-2093
extern "C" int mrgpu_scrunch(mrgpu_mr *mr, int numprocs, const void *key, int keybytes, uint64_t *nkmv)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot scrunch without KeyValue");
  if (numprocs < 1 || numprocs > mr->nranks) MR_FAIL(MRGPU_EINVAL, "Invalid processor count for gather");
  if (keybytes < 0 || (keybytes && !key)) MR_FAIL(MRGPU_EINVAL, "bad key");
  if (mr->nranks > 1 && !mr->comm) MR_FAIL(MRGPU_ECOMM, "communicator not initialised");
  mr->g_what = G_NONE;
  spool_make_resident(mr);
  gather_impl(mr, numprocs);
  collapse_impl(mr, key, keybytes);
  if (nkmv) *nkmv = allreduce_sum(mr, mr->kmv->nkmv);
  API_END(mr)
}

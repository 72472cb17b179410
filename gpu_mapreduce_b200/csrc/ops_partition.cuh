// aggregate(), local half for uniform records: hash-partition and pack in ONE pass that moves whole records.
//
// Reference: MapReduce::aggregate computes proclist[i] = hashlittle(key,keybytes,nprocs) % nprocs per pair
// (src/mapreduce.cpp:453-473) and Irregular::exchange copies each pair into its destination's slice of the
// send buffer (src/irregular.cpp:279-288).  The first GPU version did this as hash pass + radix pass on record
// indices + gather (36 ms of a 178 ms step at 8 GPUs); this is the single-pass form: a counting pre-pass gives
// every destination's slice, then each tile of records is ranked by destination inside the warp (ballots on
// the few destination bits), staged through shared memory in destination order and written out as contiguous
// runs; a decoupled look-back carries each destination's running offset from tile to tile.  Stable: a
// destination receives this rank's pairs in their original order, as the reference delivers them.
// Included by mrgpu.cu.
#pragma once

constexpr int PART_THREADS = 256;
constexpr int PART_WARPS = PART_THREADS / 32;
constexpr int PART_IPT = 8;
constexpr int PART_TILE = PART_THREADS * PART_IPT;
constexpr int PART_MAXP = 256;
constexpr int PART_DIRECT_MAXP = 16;  // up to this many ranks records are written without shared-memory staging

// hashlittle of a key held in registers as KW 32-bit words (keys are 4-byte aligned multiples of 4 bytes here)
template <int KW>
__device__ __forceinline__ uint32_t hashlittle_regs(const uint32_t (&w)[KW], uint32_t initval)
{
  uint32_t a, b, c;
  a = b = c = 0xdeadbeefu + (uint32_t) (KW * 4) + initval;
  int i = 0;
#pragma unroll
  for (int rem = KW; rem > 3; rem -= 3) {
    a += w[i];
    b += w[i + 1];
    c += w[i + 2];
    MR_MIX(a, b, c);
    i += 3;
  }
  constexpr int tail = KW == 0 ? 0 : ((KW - 1) % 3) + 1;
  if (tail == 0) return c;
  a += w[i];
  if (tail > 1) b += w[i + 1];
  if (tail > 2) c += w[i + 2];
  MR_FINAL(a, b, c);
  return c;
}

// the key starts at word 2 of the record (after the two length ints; keyalign 4 or 8 with 8-byte header):
// a compile-time index keeps the record in registers (a run-time index sent it to local memory -- ncu showed
// STL.128 spills as the top stall of the first version)
// (KOFF = 2); a header-less row (store.cuh) starts with its key (KOFF = 0)
template <int RW, int KW, int KOFF>
__device__ __forceinline__ uint32_t record_dest(const uint32_t (&rec)[RW], uint32_t nprocs)
{
  uint32_t k[KW];
#pragma unroll
  for (int q = 0; q < KW; q++) k[q] = rec[KOFF + q];
  const uint32_t h = hashlittle_regs<KW>(k, nprocs);
  // rank counts are mostly powers of two: a mask instead of the ~20-instruction division by a run-time value
  return (nprocs & (nprocs - 1)) == 0 ? (h & (nprocs - 1)) : (h % nprocs);
}

template <int RW>
__device__ __forceinline__ void load_record(const uint32_t *__restrict__ p, uint32_t (&rec)[RW])
{
  if (RW % 4 == 0) {
#pragma unroll
    for (int q = 0; q < RW; q += 4) {
      int4 v = ld_stream((const int4 *) (p + q));
      rec[q] = v.x, rec[q + 1] = v.y, rec[q + 2] = v.z, rec[q + 3] = v.w;
    }
  } else if (RW % 2 == 0) {
#pragma unroll
    for (int q = 0; q < RW; q += 2) {
      uint2 v = *(const uint2 *) (p + q);
      rec[q] = v.x, rec[q + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int q = 0; q < RW; q++) rec[q] = p[q];
  }
}

// Pass 1: per-tile, per-destination pair counts, written destination-major: tcount[d * ntiles + tile].
// One exclusive scan over that array (destination-major = output order) then gives every tile the exact
// output slot of its first pair for each destination -- the scatter pass needs no look-back between tiles
// (the first version used a decoupled look-back; ncu showed the tiles waiting on it: 5.3 barrier + 7.5
// long-scoreboard stall cycles per issue, 2.3 TB/s).
template <int RW, int KW, int KOFF>
__global__ void __launch_bounds__(PART_THREADS)
partition_count_kernel(const uint32_t *__restrict__ data, uint64_t n, uint32_t nprocs, uint64_t ntiles, unsigned *__restrict__ tcount,
                       uint8_t *__restrict__ dest_out)
{
  __shared__ unsigned sh[PART_MAXP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    sh[tid] = 0;
    __syncthreads();
    const uint64_t tbase = tile * PART_TILE;
    const unsigned nvalid = (unsigned) ((n - tbase) < (uint64_t) PART_TILE ? (n - tbase) : (uint64_t) PART_TILE);
    const unsigned woff0 = warp * (32 * PART_IPT) + lane;
    uint32_t dest[PART_IPT];
#pragma unroll
    for (int j = 0; j < PART_IPT; j++) {
      const unsigned sidx = woff0 + j * 32;
      dest[j] = 0xFFFFFFFFu;
      if (sidx < nvalid) {
        uint32_t rec[RW];
        load_record<RW>(data + (tbase + sidx) * RW, rec);
        dest[j] = record_dest<RW, KW, KOFF>(rec, nprocs);
        dest_out[tbase + sidx] = (uint8_t) dest[j];  // pass 2 does not hash again
      }
    }
    if (nprocs <= 32) {
      unsigned mine = 0;  // lane d accumulates destination d
      for (uint32_t q = 0; q < nprocs; q++) {
        unsigned c = 0;
#pragma unroll
        for (int j = 0; j < PART_IPT; j++) c += __popc(__ballot_sync(0xFFFFFFFFu, dest[j] == q));
        if (lane == (int) q) mine = c;
      }
      if (lane < (int) nprocs && mine) atomicAdd(&sh[lane], mine);
    } else {
#pragma unroll
      for (int j = 0; j < PART_IPT; j++)
        if (dest[j] != 0xFFFFFFFFu) atomicAdd(&sh[dest[j]], 1u);
    }
    __syncthreads();
    if (tid < (int) nprocs) tcount[(uint64_t) tid * ntiles + tile] = sh[tid];
    __syncthreads();
  }
}

// Pass 1 for header-less 4-byte rows (IntCount's key column): a thread takes four consecutive keys with one 16-byte
// load, writes their four destination bytes with one 4-byte store and -- when map_intcount runs this pass while
// it emits (emit_and_plan) -- copies the keys into the KV with one 16-byte store.  Counting does not care which
// thread holds which pair, so nothing here has to be warp-striped like pass 2.
__global__ void __launch_bounds__(PART_THREADS)
partition_count4_kernel(const uint32_t *__restrict__ keys, uint64_t n, uint32_t nprocs, uint64_t ntiles, unsigned *__restrict__ tcount,
                        uint8_t *__restrict__ dest_out, uint32_t *__restrict__ copy_out)
{
  __shared__ unsigned sh[PART_MAXP];
  const int tid = threadIdx.x, lane = tid & 31;
  constexpr int VPT = PART_IPT / 4;  // 16-byte vectors per thread per tile
  const bool pow2 = (nprocs & (nprocs - 1)) == 0;
  for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    sh[tid] = 0;
    __syncthreads();
    const uint64_t tbase = tile * PART_TILE;
    uint32_t dest[PART_IPT];
    int4 v[VPT];
#pragma unroll
    for (int j = 0; j < VPT; j++) {
      const uint64_t i = tbase + (uint64_t) (j * PART_THREADS + tid) * 4;
      v[j] = make_int4(0, 0, 0, 0);
      if (i + 4 <= n) v[j] = ld_stream((const int4 *) (keys + i));
      else
        for (int q = 0; q < 4; q++)
          if (i + q < n) (&v[j].x)[q] = (int) keys[i + q];
    }
#pragma unroll
    for (int j = 0; j < VPT; j++) {
      const uint64_t i = tbase + (uint64_t) (j * PART_THREADS + tid) * 4;
      uint32_t packed = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const uint32_t k[1] = {(uint32_t) (&v[j].x)[q]};
        const uint32_t h = hashlittle_regs<1>(k, nprocs);
        const uint32_t d = pow2 ? (h & (nprocs - 1)) : (h % nprocs);
        dest[j * 4 + q] = i + q < n ? d : 0xFFFFFFFFu;
        packed |= d << (8 * q);
      }
      if (i + 4 <= n) {
        *(uint32_t *) (dest_out + i) = packed;
        if (copy_out) st_stream((int4 *) (copy_out + i), v[j]);
      } else
        for (int q = 0; q < 4; q++)
          if (i + q < n) {
            dest_out[i + q] = (uint8_t) (packed >> (8 * q));
            if (copy_out) copy_out[i + q] = (uint32_t) (&v[j].x)[q];
          }
    }
    if (nprocs <= 32) {
      unsigned mine = 0;  // lane d accumulates destination d
      for (uint32_t q = 0; q < nprocs; q++) {
        unsigned c = 0;
#pragma unroll
        for (int j = 0; j < PART_IPT; j++) c += __popc(__ballot_sync(0xFFFFFFFFu, dest[j] == q));
        if (lane == (int) q) mine = c;
      }
      if (lane < (int) nprocs && mine) atomicAdd(&sh[lane], mine);
    } else {
#pragma unroll
      for (int j = 0; j < PART_IPT; j++)
        if (dest[j] != 0xFFFFFFFFu) atomicAdd(&sh[dest[j]], 1u);
    }
    __syncthreads();
    if (tid < (int) nprocs) tcount[(uint64_t) tid * ntiles + tile] = sh[tid];
    __syncthreads();
  }
}

// lanes of the warp with the same destination: one ballot per destination bit
__device__ __forceinline__ unsigned match_dest(uint32_t d, int nbits)
{
  unsigned m = 0xFFFFFFFFu;
  for (int b = 0; b < nbits; b++) {
    int sx = ((int) (d << (31 - b))) >> 31;
    unsigned bal = __ballot_sync(0xFFFFFFFFu, sx < 0);
    m &= ~(bal ^ (unsigned) sx);
  }
  return m;
}

// where each destination's pairs go: a slice of the local send buffer, or -- peer-store exchange -- the
// slice of that rank's receive buffer reserved for this source, mapped through CUDA IPC (NVLink stores)
struct DestPtrs {
  uint8_t *base[PART_MAXP];
};

// Pass 2: records -> destination slices.  toffs[d * ntiles + tile] = slot (counted over the whole destination-major
// order) of this tile's first pair for destination d; slots inside a destination's slice are relative to
// toffs[d * ntiles].
template <int RW, int KW, int KOFF>
__global__ void __launch_bounds__(PART_THREADS)
partition_scatter_kernel(const uint32_t *__restrict__ data, DestPtrs dp, uint64_t n, uint32_t nprocs, int nbits, uint64_t ntiles,
                         const unsigned *__restrict__ toffs, bool direct, const uint8_t *__restrict__ dest_in)
{
  extern __shared__ __align__(16) unsigned char psm[];
  uint32_t *xrec = (uint32_t *) psm;                                // [PART_TILE * RW] records in destination order
  unsigned char *xdst = psm + (size_t) PART_TILE * RW * 4;          // [PART_TILE] destination of each staged record
  __shared__ unsigned wc[PART_WARPS][PART_MAXP];
  __shared__ unsigned gbase[PART_MAXP];
  __shared__ unsigned warp_tot[PART_WARPS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int w = 0; w < PART_WARPS; w++) wc[w][tid] = 0;
  __syncthreads();
  const uint64_t tile = blockIdx.x, tbase = tile * PART_TILE;
  const unsigned nvalid = (unsigned) ((n - tbase) < (uint64_t) PART_TILE ? (n - tbase) : (uint64_t) PART_TILE);
  const unsigned woff0 = warp * (32 * PART_IPT) + lane;

  uint32_t rec[PART_IPT][RW];
  uint32_t dest[PART_IPT];
  // all loads first, then all hashes: eight independent 16-byte loads in flight per thread (interleaved, each
  // hash waited for its own load: half of all stall samples in ncu)
#pragma unroll
  for (int j = 0; j < PART_IPT; j++) {
    const unsigned sidx = woff0 + j * 32;
    if (sidx < nvalid) load_record<RW>(data + (tbase + sidx) * RW, rec[j]);
  }
#pragma unroll
  for (int j = 0; j < PART_IPT; j++) {
    const unsigned sidx = woff0 + j * 32;
    dest[j] = nprocs;  // sentinel for empty slots: ranks after every real destination
    if (sidx < nvalid) dest[j] = dest_in[tbase + sidx];
  }
  // rank inside the warp by destination, in memory order
  const unsigned lt_mask = lanemask_lt();
  unsigned rank[PART_IPT];
  if (nprocs <= 32) {
    // few destinations: lane q keeps the warp's running count for destination q in a register; a round is
    // nprocs ballots, the base of a pair's group comes from a shuffle.  No shared-memory round trip per round
    // (the version with a counter in shared memory stalled on that load: 25 % of the samples in ncu).
    unsigned running = 0;
#pragma unroll
    for (int j = 0; j < PART_IPT; j++) {
      unsigned mine = 0, add = 0;
      for (uint32_t q = 0; q < nprocs; q++) {
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, dest[j] == q);
        if (dest[j] == q) mine = bal;
        if (lane == (int) q) add = __popc(bal);
      }
      const unsigned base = __shfl_sync(0xFFFFFFFFu, running, dest[j] & 31);
      rank[j] = base + __popc(mine & lt_mask);
      running += add;
    }
    if (lane < (int) nprocs) wc[warp][lane] = running;
  } else {
#pragma unroll
    for (int j = 0; j < PART_IPT; j++) {
      const bool valid = woff0 + j * 32 < nvalid;
      unsigned peers = match_dest(dest[j], nbits) & __ballot_sync(0xFFFFFFFFu, valid);
      unsigned b = 0;
      if (valid && (peers & lt_mask) == 0) {
        unsigned *c = &wc[warp][dest[j]];
        b = *c;
        *c = b + __popc(peers);
      }
      __syncwarp();
      b = __shfl_sync(0xFFFFFFFFu, b, valid ? __ffs(peers) - 1 : 0);
      rank[j] = b + __popc(peers & lt_mask);
    }
  }
  __syncthreads();

  // thread d < nprocs: this tile's count for destination d and each warp's exclusive offset
  unsigned cnt = 0, wex[PART_WARPS];
#pragma unroll
  for (int w = 0; w < PART_WARPS; w++) {
    wex[w] = cnt;
    cnt += wc[w][tid];
  }
  unsigned incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  unsigned woff = 0;
#pragma unroll
  for (int w = 0; w < PART_WARPS; w++)
    if (w < warp) woff += warp_tot[w];
  const unsigned my_start = woff + incl - cnt;
#pragma unroll
  for (int w = 0; w < PART_WARPS; w++) wc[w][tid] = my_start + wex[w];

  if (tid < (int) nprocs) gbase[tid] = toffs[(uint64_t) tid * ntiles + tile] - toffs[(uint64_t) tid * ntiles] - my_start;
  __syncthreads();

  // stage records in destination order, then write contiguous runs -- in the widest vectors the record size
  // allows (16-byte records: one 128-bit shared store, one 128-bit shared load, one 128-bit global store)
  constexpr int VW = (RW % 4 == 0) ? 4 : ((RW % 2 == 0) ? 2 : 1);
  constexpr int NV = RW / VW;
  using V = typename std::conditional<VW == 4, uint4, typename std::conditional<VW == 2, uint2, uint32_t>::type>::type;
  if (direct) {
    // few destinations: each lane writes its record straight to its slot.  Lanes of a warp round that share a
    // destination hold consecutive slots, so a round produces at most nprocs contiguous pieces (>= 64 bytes
    // each on average for 16-byte records and 8 ranks) and successive rounds extend the same runs; no
    // shared-memory staging, no second barrier.
#pragma unroll
    for (int j = 0; j < PART_IPT; j++) {
      if (woff0 + j * 32 < nvalid) {
        const uint64_t slot = (uint64_t) (gbase[dest[j]] + wc[warp][dest[j]] + rank[j]) * NV;
#pragma unroll
        for (int q = 0; q < NV; q++) {
          V v;
          if constexpr (VW == 4) v = make_uint4(rec[j][4 * q], rec[j][4 * q + 1], rec[j][4 * q + 2], rec[j][4 * q + 3]);
          else if constexpr (VW == 2) v = make_uint2(rec[j][2 * q], rec[j][2 * q + 1]);
          else v = rec[j][q];
          reinterpret_cast<V *>(dp.base[dest[j]])[slot + q] = v;
        }
      }
    }
    return;
  }
  V *xv = reinterpret_cast<V *>(xrec);
#pragma unroll
  for (int j = 0; j < PART_IPT; j++) {
    if (woff0 + j * 32 < nvalid) {
      const unsigned pos = wc[warp][dest[j]] + rank[j];
#pragma unroll
      for (int q = 0; q < NV; q++) {
        V v;
        if constexpr (VW == 4) v = make_uint4(rec[j][4 * q], rec[j][4 * q + 1], rec[j][4 * q + 2], rec[j][4 * q + 3]);
        else if constexpr (VW == 2) v = make_uint2(rec[j][2 * q], rec[j][2 * q + 1]);
        else v = rec[j][q];
        xv[pos * NV + q] = v;
      }
      xdst[pos] = (unsigned char) dest[j];
    }
  }
  __syncthreads();
  if (VW < 4 && nprocs <= PART_DIRECT_MAXP) {
    // narrow rows, few destinations: each destination's run leaves as 16-byte stores (4-byte keys over NVLink:
    // 476 GB/s with one word per lane against 654 GB/s measured for 16-byte records)
    for (uint32_t d = 0; d < nprocs; d++) {
      const unsigned rs = wc[0][d], re = (d + 1 < nprocs) ? wc[0][d + 1] : nvalid;
      const unsigned w0 = rs * RW, len = (re - rs) * RW;  // words in shared memory
      uint32_t *G = (uint32_t *) dp.base[d] + (uint64_t) (gbase[d] + rs) * RW;
      unsigned head = (unsigned) ((16 - ((uintptr_t) G & 15)) & 15) / 4;
      if (head > len) head = len;
      const unsigned nv4 = (len - head) / 4, tail0 = head + nv4 * 4;
      if (tid < (int) head) G[tid] = xrec[w0 + tid];
      for (unsigned v = tid; v < nv4; v += PART_THREADS) {
        const uint32_t *x = xrec + w0 + head + 4 * v;
        *reinterpret_cast<int4 *>(G + head + 4 * v) = make_int4((int) x[0], (int) x[1], (int) x[2], (int) x[3]);
      }
      if (tid < (int) (len - tail0)) G[tail0 + tid] = xrec[w0 + tail0 + tid];
    }
    return;
  }
  const unsigned nvec = nvalid * NV;
  for (unsigned v = tid; v < nvec; v += PART_THREADS) {
    const unsigned s = v / NV, q = v - s * NV;
    reinterpret_cast<V *>(dp.base[xdst[s]])[(uint64_t) (gbase[xdst[s]] + s) * NV + q] = xv[v];
  }
}

// state between the two passes: struct PartPlan, store.cuh

template <int RW, int KW, int KOFF>
static void partition_counts_t(mrgpu_mr *mr, DevKV *kv, PartPlan &pl, std::vector<uint64_t> &scount, const uint32_t *src = nullptr, uint64_t nsrc = 0)
{
  // src != nullptr (header-less 4-byte rows only): the rows come from src and the same pass writes them to kv->data
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const int P = mr->nranks;
  const uint64_t n = src ? nsrc : kv->nkv;
  const uint64_t ntiles = (n + PART_TILE - 1) / PART_TILE;
  pl.rw = RW;
  pl.kw = KW;
  pl.ntiles = ntiles;
  Buf<unsigned> tcount(pool, ntiles * P);
  pl.toffs.alloc(pool, ntiles * P);
  pl.dest.alloc(pool, n);
  {
    // emit + count: map's 20 B/pair plus aggregate's read of the KV (16 B/pair) in the reference layout
    ProfScope ps(pool, src ? "map_intcount_emit_plan" : "aggregate_count", src ? (double) n * 36 : (double) kv->nbytes, 1,
                 src ? (double) n * 9 : (double) kv->phys_bytes() + (double) n);
    const unsigned grid = (unsigned) std::min<uint64_t>(ntiles, (uint64_t) kSMs * 16);
    const uint32_t *in = src ? src : (const uint32_t *) kv->data;
    if (RW == 1 && KW == 1 && KOFF == 0 && ((uintptr_t) in % 16) == 0 && ((uintptr_t) kv->data % 16) == 0)
      partition_count4_kernel<<<grid, PART_THREADS, 0, s>>>(in, n, (uint32_t) P, ntiles, tcount.p, pl.dest.p, src ? (uint32_t *) kv->data : nullptr);
    else if (src) MR_FAIL(MRGPU_EINVAL, "internal: emit_and_plan needs 16-byte aligned header-less rows");
    else partition_count_kernel<RW, KW, KOFF><<<grid, PART_THREADS, 0, s>>>(in, n, (uint32_t) P, ntiles, tcount.p, pl.dest.p);
    MR_CUDA(cudaGetLastError());
  }
  // exclusive scan in destination-major order = slot of every (destination, tile)
  const unsigned *tc = tcount.p;
  unsigned *to = pl.toffs.p;
  scan_apply(
      pool, ntiles * P, [=] __device__(uint64_t i) -> unsigned long long { return tc[i]; },
      [=] __device__(uint64_t i, unsigned long long e, unsigned long long) { to[i] = (unsigned) e; }, mr->d_small, "aggregate_tile_scan",
      8.0 * ntiles * P);
  std::vector<unsigned> first(P);
  for (int p = 0; p < P; p++) MR_CUDA(cudaMemcpyAsync(&first[p], pl.toffs.p + (uint64_t) p * ntiles, 4, cudaMemcpyDeviceToHost, s));
  sync(mr);
  for (int p = 0; p < P; p++) scount[p] = (p + 1 < P ? first[p + 1] : n) - first[p];
}

template <int RW, int KW, int KOFF>
static void partition_scatter_t(mrgpu_mr *mr, DevKV *kv, PartPlan &pl, const DestPtrs &dp, const char *tag, bool remote)
{
  // MRGPU_PART_DIRECT / MRGPU_PEER_DIRECT (0/1) override the choice between per-lane stores and staged runs
  static const char *e_loc = getenv("MRGPU_PART_DIRECT"), *e_rem = getenv("MRGPU_PEER_DIRECT");
  const char *e = remote ? e_rem : e_loc;
  // stores that cross NVLink go out as staged runs (full 128-byte lines): 654 vs 556 GB/s at 2 ranks, 302 vs 228 at 8
  const bool direct = e ? atoi(e) != 0 : (!remote && mr->nranks <= PART_DIRECT_MAXP);
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const int P = mr->nranks;
  int nbits = 0;
  while ((1 << nbits) < P) nbits++;
  const size_t smem = (size_t) PART_TILE * RW * 4 + PART_TILE;
  static std::atomic<bool> attr[64];  // per instantiation and device; the attribute call itself is idempotent
  if (mr->device < 64 && !attr[mr->device].load()) {
    MR_CUDA(cudaFuncSetAttribute(partition_scatter_kernel<RW, KW, KOFF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    attr[mr->device].store(true);
  }
  ProfScope ps(pool, tag, 2.0 * kv->nbytes, 1, 2.0 * kv->phys_bytes());
  partition_scatter_kernel<RW, KW, KOFF><<<(unsigned) pl.ntiles, PART_THREADS, direct ? 0 : smem, s>>>((const uint32_t *) kv->data, dp, kv->nkv, (uint32_t) P, nbits, pl.ntiles,
                                                                                  pl.toffs.p, direct, pl.dest.p);
  MR_CUDA(cudaGetLastError());
}

// rows the kernels are instantiated for: {words per row, key words, key offset}
#define PART_DISPATCH(FN, ...)                                                                  \
  do {                                                                                          \
    const int _kw = (int) (kv->kb / 4), _ko = kv->packed ? 0 : 2;                               \
    const int _rw = (int) ((kv->packed ? kv->prow() : kv->recsize) / 4);                        \
    if (_ko == 2 && _rw == 3 && _kw == 1) FN<3, 1, 2>(__VA_ARGS__);                             \
    else if (_ko == 2 && _rw == 4 && _kw == 1) FN<4, 1, 2>(__VA_ARGS__);                        \
    else if (_ko == 2 && _rw == 4 && _kw == 2) FN<4, 2, 2>(__VA_ARGS__);                        \
    else if (_ko == 2 && _rw == 5 && _kw == 2) FN<5, 2, 2>(__VA_ARGS__);                        \
    else if (_ko == 2 && _rw == 6 && _kw == 2) FN<6, 2, 2>(__VA_ARGS__);                        \
    else if (_ko == 2 && _rw == 6 && _kw == 4) FN<6, 4, 2>(__VA_ARGS__);                        \
    else if (_ko == 0 && _rw == 1 && _kw == 1) FN<1, 1, 0>(__VA_ARGS__);                        \
    else if (_ko == 0 && _rw == 2 && _kw == 1) FN<2, 1, 0>(__VA_ARGS__);                        \
    else if (_ko == 0 && _rw == 2 && _kw == 2) FN<2, 2, 0>(__VA_ARGS__);                        \
    else if (_ko == 0 && _rw == 4 && _kw == 4) FN<4, 4, 0>(__VA_ARGS__);                        \
    else MR_FAIL(MRGPU_EINVAL, "internal: partition kernel not instantiated for this row");    \
  } while (0)

// record shapes the single-pass partition handles: 4-byte-aligned uniform records, key right after the header
// (or header-less rows)
static bool partition_eligible(mrgpu_mr *mr, DevKV *kv)
{
  const Align al = mr->al;
  if (!kv->uniform || kv->nkv == 0 || mr->nranks > PART_MAXP || (al.k % 4) || (kv->kb % 4) || kv->kb == 0 || (kv->recsize % 4)) return false;
  if ((uintptr_t) kv->data % 16) return false;
  const int kw = kv->kb / 4;
  if (kv->packed) {
    const int rw = kv->prow() / 4;
    return (rw == 1 && kw == 1) || (rw == 2 && kw == 1) || (rw == 2 && kw == 2) || (rw == 4 && kw == 4);
  }
  if (kv_key_off(al) != 8) return false;
  const int rw = kv->recsize / 4;
  return (rw == 3 && kw == 1) || (rw == 4 && kw == 1) || (rw == 4 && kw == 2) || (rw == 5 && kw == 2) || (rw == 6 && kw == 2) || (rw == 6 && kw == 4);
}

static void partition_counts(mrgpu_mr *mr, DevKV *kv, PartPlan &pl, std::vector<uint64_t> &scount) { PART_DISPATCH(partition_counts_t, mr, kv, pl, scount); }
// map_intcount on several ranks (ops_map.cuh): the pass that writes the key column also does aggregate()'s
// pass 1 (destination of every pair, per-tile counts).  aggregate() uses the plan if the KV is still the same.
static void emit_and_plan(mrgpu_mr *mr, DevKV *kv, const uint32_t *keys, uint64_t n)
{
  delete kv->plan;
  kv->plan = new PartPlan();
  kv->plan_scount.assign(mr->nranks, 0);
  partition_counts_t<1, 1, 0>(mr, kv, *kv->plan, kv->plan_scount, keys, n);
  kv->plan_nranks = mr->nranks;
  kv->ahead_n = n;
}

static void partition_scatter(mrgpu_mr *mr, DevKV *kv, PartPlan &pl, const DestPtrs &dp, const char *tag, bool remote = false)
{
  PART_DISPATCH(partition_scatter_t, mr, kv, pl, dp, tag, remote);
}

// both passes into one local buffer: destination d's slice starts at record slot sum(scount[0..d))
static bool partition_uniform(mrgpu_mr *mr, DevKV *kv, uint8_t *sendbuf, std::vector<uint64_t> &scount)
{
  if (!partition_eligible(mr, kv) || ((uintptr_t) sendbuf % 16)) return false;
  PartPlan pl;
  partition_counts(mr, kv, pl, scount);
  DestPtrs dp;
  uint64_t off = 0;
  for (int p = 0; p < mr->nranks; p++) {
    dp.base[p] = sendbuf + off;
    off += scount[p] * (kv->packed ? kv->prow() : kv->recsize);
  }
  partition_scatter(mr, kv, pl, dp, "aggregate_partition_scatter");
  sync(mr);
  return true;
}

// Test / profiling hook: the local half of aggregate() for a VIRTUAL rank count on one GPU -- the KV is
// replaced by the same records grouped by destination hashlittle(key,kb,nprocs) % nprocs (stable) and
// counts[d] receives each destination's pair count.  No exchange.  Uniform records only.
extern "C" int mrgpu_debug_partition(mrgpu_mr *mr, int nprocs, uint64_t *counts)
{
  API_BEGIN(mr)
  if (!mr->kv) MR_FAIL(MRGPU_ENOKV, "Cannot aggregate without KeyValue");
  if (nprocs < 1 || nprocs > PART_MAXP) MR_FAIL(MRGPU_EINVAL, "bad nprocs");
  freeze(mr);
  spool_make_resident(mr);
  if (mr->kv->weighted) kv_unpack(mr, mr->kv);
  DevKV *kv = mr->kv;
  DevKV *nk = new DevKV(&mr->pool);
  int saved = mr->nranks;
  try {
    nk->reserve_bytes(kv->phys_bytes());
    std::vector<uint64_t> sc(nprocs, 0);
    mr->nranks = nprocs;
    bool ok = kv->nkv && partition_uniform(mr, kv, nk->data, sc);
    mr->nranks = saved;
    if (!ok) MR_FAIL(MRGPU_EINVAL, "mrgpu_debug_partition: record shape not handled by the single-pass partition");
    for (int p = 0; p < nprocs; p++)
      if (counts) counts[p] = sc[p];
    nk->uniform = true;
    nk->kb = kv->kb;
    nk->vb = kv->vb;
    nk->recsize = kv->recsize;
    nk->packed = kv->packed;
    nk->cval_on = kv->cval_on;
    nk->cval[0] = kv->cval[0];
    nk->cval[1] = kv->cval[1];
    nk->nbytes = kv->nbytes;
    nk->nkv = kv->nkv;
    nk->ksize = kv->ksize;
    nk->vsize = kv->vsize;
  } catch (...) {
    mr->nranks = saved;
    delete nk;
    throw;
  }
  delete mr->kv;
  mr->kv = nk;
  API_END(mr)
}

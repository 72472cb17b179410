// Shared helpers for the sm_100a kernels: error plumbing, the stream-ordered HBM pool, vector
// load/store wrappers and record-layout arithmetic (KV: src/keyvalue.cpp:345-352, KMV:
// src/keymultivalue.cpp:830-838 of the reference).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <map>
#include <mutex>
#include <set>
#include <thread>
#include <string>
#include <vector>

namespace mrgpu {

constexpr int kSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

struct Error {
  int code;
  std::string msg;
};

#define MR_CUDA(expr)                                                                          \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      throw ::mrgpu::Error{-2, std::string(#expr " failed: ") + cudaGetErrorString(_e)};       \
  } while (0)

#define MR_FAIL(code, text) throw ::mrgpu::Error{(code), std::string(text)}

// ---------------------------------------------------------------------------------------------
// Per-kernel timing, the analogue of the reference's `timer` setting (src/mapreduce.cpp:3112-3128 prints
// "<Op> time (secs)").  When enabled every tagged launch is bracketed by CUDA events on the launching
// stream; launches are always counted.
struct ProfEntry {
  const char *tag;
  cudaEvent_t a, b;
  double bytes;  // algorithmic bytes of this launch: the reference's exact sizes, 0 if not stated
  int launches;
  double moved;  // bytes this launch has to move given the layout actually in HBM (= bytes unless a compact store is involved)
};
struct Prof {
  bool on = false;
  std::vector<ProfEntry> ev;
  std::vector<cudaEvent_t> spare;
  uint64_t launches = 0;
  cudaEvent_t get()
  {
    cudaEvent_t e;
    if (!spare.empty()) {
      e = spare.back();
      spare.pop_back();
    } else
      cudaEventCreate(&e);
    return e;
  }
  void reset()
  {
    for (auto &x : ev) {
      spare.push_back(x.a);
      spare.push_back(x.b);
    }
    ev.clear();
    launches = 0;
  }
  ~Prof()
  {
    reset();
    for (auto e : spare) cudaEventDestroy(e);
  }
};

// HBM page pool.  The reference keeps a tagged pool of `memsize` pages and frees unused ones after every
// operation (src/mapreduce.cpp:3318-3517); here HBM is the page pool and KV/KMV/scratch buffers of any size
// are carved from cached chunks, so that after warm-up an operation never calls the driver allocator.
struct Pool {
  cudaStream_t stream = nullptr;
  int device = 0;
  uint64_t live = 0, hiwater = 0;  // the reference's msize / msizemax (src/mapreduce.h:54)
  uint64_t reserved = 0;           // bytes obtained from cudaMalloc
  int64_t fail_after = -1;         // test hook ("debug_fail_alloc" setting): the fail_after-th next alloc() throws
  bool abandon = false;            // mrgpu_abandon(): peers may still map exported chunks, so those are never freed
  Prof prof;

  // Caching sub-allocator over big cudaMalloc'ed chunks.  The driver's own stream-ordered pool
  // (cudaMallocAsync) re-maps physical pages whenever a request does not match a cached block, which
  // measured ~6 ms per GiB on every operation; an operation here allocates tens of GiB of KV/KMV/sort
  // buffers, so blocks are cached and split/coalesced in user space instead.  All work of a handle is issued
  // on one stream, so a block freed after its last kernel was enqueued can be handed out again at once:
  // the next user is ordered behind it on the same stream.
  struct Block {
    uint64_t size;
    int chunk;
  };
  std::map<uint8_t *, Block> freeb;                 // free blocks by address (for coalescing)
  std::multimap<uint64_t, uint8_t *> by_size;       // free blocks by size (best fit)
  std::map<uint8_t *, Block> used;                  // live blocks
  std::vector<std::pair<uint8_t *, uint64_t>> chunks;
  static constexpr uint64_t kGran = 512;

  // Every handle has its own pool, and a pool keeps what it once obtained (that is the point of it).  When one
  // pool's cudaMalloc fails, the memory is often sitting idle in ANOTHER handle's cache (R-MAT's edge handle holds
  // 100 GB of freed sort scratch while the degree handle asks for its first 34 GB): the pools of this host thread
  // on this device are asked to give their wholly-free chunks back before the allocation is declared failed.
  // (Only this thread's: a handle is not thread-safe, and another thread may be inside its own pool.)
  std::thread::id owner;
  static std::mutex &registry_mu()
  {
    static std::mutex m;
    return m;
  }
  static std::set<Pool *> &registry()
  {
    static std::set<Pool *> r;
    return r;
  }
  void init(int dev, cudaStream_t s)
  {
    device = dev;
    stream = s;
    owner = std::this_thread::get_id();
    std::lock_guard<std::mutex> lk(registry_mu());
    registry().insert(this);
  }
  void trim_siblings()
  {
    std::lock_guard<std::mutex> lk(registry_mu());
    for (Pool *p : registry())
      if (p != this && p->device == device && p->owner == owner) p->trim();
  }
  void erase_free(std::map<uint8_t *, Block>::iterator it)
  {
    auto r = by_size.equal_range(it->second.size);
    for (auto j = r.first; j != r.second; ++j)
      if (j->second == it->first) {
        by_size.erase(j);
        break;
      }
    freeb.erase(it);
  }
  void add_free(uint8_t *p, Block b)
  {
    freeb[p] = b;
    by_size.emplace(b.size, p);
  }
  // give wholly-free chunks back to the driver (called when cudaMalloc fails, and at destruction)
  void trim()
  {
    for (size_t c = 0; c < chunks.size(); c++) {
      if (!chunks[c].first || exported.count((int) c)) continue;  // a peer process may hold a mapping of an exported chunk
      auto it = freeb.find(chunks[c].first);
      if (it != freeb.end() && it->second.size == chunks[c].second) {
        erase_free(it);
        cudaStreamSynchronize(stream);
        cudaFree(chunks[c].first);
        reserved -= chunks[c].second;
        chunks[c].first = nullptr;
      }
    }
  }
  void *alloc(uint64_t bytes)
  {
    if (fail_after >= 0 && fail_after-- == 0)
      throw ::mrgpu::Error{-2, "out of GPU memory: allocation failure injected by the debug_fail_alloc setting"};
    if (bytes == 0) bytes = kGran;
    bytes = (bytes + kGran - 1) / kGran * kGran;
    auto it = by_size.lower_bound(bytes);
    uint8_t *p;
    Block b;
    if (it != by_size.end()) {
      p = it->second;
      auto f = freeb.find(p);
      b = f->second;
      erase_free(f);
    } else {
      uint64_t csize = (bytes + (2ull << 20) - 1) / (2ull << 20) * (2ull << 20);
      void *q = nullptr;
      cudaError_t e = cudaMalloc(&q, csize);
      if (e != cudaSuccess) {
        cudaGetLastError();
        trim();
        e = cudaMalloc(&q, csize);
      }
      if (e != cudaSuccess) {
        cudaGetLastError();
        trim_siblings();
        e = cudaMalloc(&q, csize);
      }
      if (e != cudaSuccess) {
        cudaGetLastError();
        throw ::mrgpu::Error{-2, "out of GPU memory: cudaMalloc(" + std::to_string(csize) + " bytes) failed with " +
                                     std::to_string(reserved) + " bytes already reserved by this handle"};
      }
      p = (uint8_t *) q;
      chunks.emplace_back(p, csize);
      reserved += csize;
      b = Block{csize, (int) chunks.size() - 1};
    }
    if (b.size - bytes >= (1ull << 16)) {  // split, keep the tail free
      add_free(p + bytes, Block{b.size - bytes, b.chunk});
      b.size = bytes;
    }
    used[p] = b;
    live += b.size;
    if (live > hiwater) hiwater = live;
    return p;
  }
  void free(void *ptr, uint64_t)
  {
    if (!ptr) return;
    uint8_t *p = (uint8_t *) ptr;
    auto u = used.find(p);
    if (u == used.end()) return;
    Block b = u->second;
    used.erase(u);
    live -= b.size;
    // coalesce with free neighbours of the same chunk
    auto nx = freeb.find(p + b.size);
    if (nx != freeb.end() && nx->second.chunk == b.chunk) {
      b.size += nx->second.size;
      erase_free(nx);
    }
    auto pv = freeb.lower_bound(p);
    if (pv != freeb.begin()) {
      --pv;
      if (pv->first + pv->second.size == p && pv->second.chunk == b.chunk) {
        p = pv->first;
        b.size += pv->second.size;
        erase_free(pv);
      }
    }
    add_free(p, b);
  }
  // grow a live block in place when the bytes behind it are free (same cudaMalloc allocation): a store that is being
  // appended to then keeps its address and nothing is copied
  bool try_extend(void *ptr, uint64_t bytes)
  {
    if (!ptr || fail_after >= 0) return false;
    bytes = (bytes + kGran - 1) / kGran * kGran;
    auto u = used.find((uint8_t *) ptr);
    if (u == used.end()) return false;
    Block &b = u->second;
    if (bytes <= b.size) return true;
    auto nx = freeb.find((uint8_t *) ptr + b.size);
    if (nx == freeb.end() || nx->second.chunk != b.chunk || b.size + nx->second.size < bytes) return false;
    const uint64_t take = bytes - b.size, left = nx->second.size - take;
    uint8_t *np = nx->first;
    const int chunk = nx->second.chunk;
    erase_free(nx);
    uint64_t got = take;
    if (left >= (1ull << 16)) add_free(np + take, Block{left, chunk});
    else got += left;
    b.size += got;
    live += got;
    if (live > hiwater) hiwater = live;
    return true;
  }
  // the cudaMalloc allocation a live block belongs to (for CUDA IPC, which shares whole allocations)
  std::set<int> exported;  // chunks whose CUDA IPC handle went to peer ranks: kept until the pool dies
  bool chunk_of(const void *ptr, uint8_t **base, uint64_t *size)
  {
    auto u = used.find((uint8_t *) ptr);
    if (u == used.end()) return false;
    exported.insert(u->second.chunk);
    *base = chunks[u->second.chunk].first;
    *size = chunks[u->second.chunk].second;
    return true;
  }
  ~Pool()
  {
    {
      std::lock_guard<std::mutex> lk(registry_mu());
      registry().erase(this);
    }
    if (chunks.empty()) return;
    cudaStreamSynchronize(stream);
    for (size_t c = 0; c < chunks.size(); c++)
      if (chunks[c].first && !(abandon && exported.count((int) c))) cudaFree(chunks[c].first);
  }
};

// brackets the kernel launches issued while it is alive
struct ProfScope {
  Pool &p;
  int idx = -1;
  ProfScope(Pool &pool, const char *tag, double bytes = 0, int nlaunch = 1, double moved = -1) : p(pool)
  {
    p.prof.launches += nlaunch;
    if (p.prof.on) {
      ProfEntry e{tag, p.prof.get(), p.prof.get(), bytes, nlaunch, moved < 0 ? bytes : moved};
      cudaEventRecord(e.a, p.stream);
      idx = (int) p.prof.ev.size();
      p.prof.ev.push_back(e);
    }
  }
  ~ProfScope()
  {
    if (idx >= 0) cudaEventRecord(p.prof.ev[idx].b, p.stream);
  }
};

// RAII scratch buffer on the pool
template <typename T>
struct Buf {
  Pool *pool = nullptr;
  T *p = nullptr;
  uint64_t n = 0;
  Buf() {}
  Buf(Pool &pl, uint64_t count) { alloc(pl, count); }
  Buf(const Buf &) = delete;
  Buf &operator=(const Buf &) = delete;
  Buf(Buf &&o) noexcept { *this = static_cast<Buf &&>(o); }
  Buf &operator=(Buf &&o) noexcept
  {
    if (this != &o) {
      release();
      pool = o.pool;
      p = o.p;
      n = o.n;
      o.p = nullptr;
      o.n = 0;
    }
    return *this;
  }
  void alloc(Pool &pl, uint64_t count)
  {
    release();
    pool = &pl;
    n = count;
    p = (T *) pl.alloc(count * sizeof(T));
  }
  void release()
  {
    if (p && pool) pool->free(p, n * sizeof(T));
    p = nullptr;
    n = 0;
  }
  T *detach()
  {
    T *r = p;
    p = nullptr;
    return r;
  }
  ~Buf() { release(); }
};

// ---------------------------------------------------------------------------------------------
// record layout

struct Align {
  int k = 4, v = 4, t = 4;  // kalign, valign, talign = max(k, v, sizeof(int)) src/mapreduce.cpp:3333-3335
};

__host__ __device__ inline uint64_t round_up(uint64_t x, uint32_t a) { return (x + a - 1) & ~(uint64_t) (a - 1); }

// KV: int keybytes | int valuebytes | pad->k | key | pad->v | value | pad->t
__host__ __device__ inline uint32_t kv_key_off(Align al) { return (uint32_t) round_up(8, al.k); }
__host__ __device__ inline uint32_t kv_val_off(Align al, uint32_t kb) { return (uint32_t) round_up(kv_key_off(al) + kb, al.v); }
__host__ __device__ inline uint32_t kv_rec_size(Align al, uint32_t kb, uint32_t vb)
{
  return (uint32_t) round_up(kv_val_off(al, kb) + vb, al.t);
}

// KMV: int nvalue | int keybytes | int mvbytes | int sizes[nvalue] | pad->k | key | pad->v | values | pad->t
__host__ __device__ inline uint64_t kmv_key_off(Align al, uint64_t nv) { return round_up(12 + 4 * nv, al.k); }
__host__ __device__ inline uint64_t kmv_val_off(Align al, uint64_t nv, uint32_t kb) { return round_up(kmv_key_off(al, nv) + kb, al.v); }
__host__ __device__ inline uint64_t kmv_rec_size(Align al, uint64_t nv, uint32_t kb, uint64_t mvb)
{
  return round_up(kmv_val_off(al, nv, kb) + mvb, al.t);
}

// ---------------------------------------------------------------------------------------------
// device helpers

__device__ __forceinline__ unsigned lanemask_lt()
{
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// streaming 128-bit accesses: data that is read or written exactly once should not displace L1 lines
__device__ __forceinline__ int4 ld_stream(const int4 *p)
{
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(int4 *p, int4 v)
{
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// 32-byte streaming accesses (sm_100 has 256-bit global loads/stores) marked evict-first in L2: a stream that is
// read or written once should not push out what the kernel keeps coming back to (convert's count table)
__device__ __forceinline__ void ld_stream256(const void *p, uint32_t (&q)[8])
{
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7])
               : "l"(p));
}
__device__ __forceinline__ void st_stream256(void *p, const uint32_t (&q)[8])
{
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(q[0]), "r"(q[1]), "r"(q[2]),
               "r"(q[3]), "r"(q[4]), "r"(q[5]), "r"(q[6]), "r"(q[7])
               : "memory");
}

inline unsigned grid_for(uint64_t work_items, unsigned per_block, unsigned max_waves = 32)
{
  uint64_t b = (work_items + per_block - 1) / per_block;
  uint64_t cap = (uint64_t) kSMs * max_waves;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned) b;
}

}  // namespace mrgpu

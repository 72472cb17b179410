// Host-only test of the HBM pool and of a store's growth (common.cuh, store.cuh): the CUDA runtime calls they make are
// replaced by host stand-ins, so the allocator's bookkeeping -- best fit, split, coalesce, growth in place, trimming --
// runs on a machine without a GPU.  Built and run by tests/test_host_logic.py (nvcc compiles it: the headers hold device
// code; no kernel is launched).
#include <cuda_runtime.h>

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static int g_mallocs = 0, g_frees = 0, g_copies = 0;
static cudaError_t host_malloc(void **p, size_t n)
{
  *p = aligned_alloc(512, (n + 511) / 512 * 512);
  g_mallocs++;
  return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
static cudaError_t host_free(void *p)
{
  free(p);
  g_frees++;
  return cudaSuccess;
}
static cudaError_t host_copy(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t)
{
  memmove(d, s, n);
  g_copies++;
  return cudaSuccess;
}
#define cudaMalloc host_malloc
#define cudaFree host_free
#define cudaMemcpyAsync host_copy
#define cudaStreamSynchronize(s) cudaSuccess
#define cudaEventRecord(e, s) cudaSuccess

#include "../../gpu_mapreduce_b200/csrc/store.cuh"

using namespace mrgpu;

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

int main()
{
  const uint64_t MB = 1ull << 20;
  {
    Pool pool;
    // first block: a 2 MB chunk is obtained, 1 MB used, 1 MB stays free behind it
    uint8_t *a = (uint8_t *) pool.alloc(MB);
    CHECK(g_mallocs == 1 && pool.reserved == 2 * MB && pool.live == MB);
    // growth in place: takes half of the free tail, keeps the address
    CHECK(pool.try_extend(a, MB + MB / 2));
    CHECK(pool.live == MB + MB / 2 && g_mallocs == 1);
    // the rest of the tail serves the next request; a is now boxed in
    uint8_t *b = (uint8_t *) pool.alloc(MB / 2);
    CHECK(b == a + MB + MB / 2 && g_mallocs == 1);
    CHECK(!pool.try_extend(a, 2 * MB));
    CHECK(pool.try_extend(a, MB));  // shrinking or equal requests are trivially fine
    // b goes back: a can take all of it (a leftover below 64 KB is absorbed rather than split off)
    pool.free(b, 0);
    CHECK(pool.try_extend(a, 2 * MB - 1024));
    CHECK(pool.live == 2 * MB);
    // another chunk; blocks of different chunks never merge
    uint8_t *c = (uint8_t *) pool.alloc(3 * MB);
    CHECK(g_mallocs == 2 && pool.reserved == 6 * MB);
    CHECK(!pool.try_extend(a, 3 * MB));
    pool.free(a, 0);
    pool.free(c, 0);
    CHECK(pool.live == 0);
    // best fit: a 1 MB request takes the 2 MB chunk, not the 4 MB one
    uint8_t *d = (uint8_t *) pool.alloc(MB);
    CHECK(d == a && g_mallocs == 2);
    pool.free(d, 0);
    // wholly-free chunks go back to the driver on trim()
    pool.trim();
    CHECK(g_frees == 2 && pool.reserved == 0);
    // the injected allocation failure (debug_fail_alloc) counts alloc() calls and switches growth in place off
    uint8_t *e = (uint8_t *) pool.alloc(MB);
    pool.fail_after = 1;
    CHECK(!pool.try_extend(e, MB + 512));
    uint8_t *f = (uint8_t *) pool.alloc(512);
    bool threw = false;
    try {
      pool.alloc(512);
    } catch (const Error &) {
      threw = true;
    }
    CHECK(threw && pool.fail_after < 0);
    pool.free(e, 0);
    pool.free(f, 0);
  }
  {
    // a store that is appended to: grows in place while the bytes behind it are free, moves (one copy) when they are not
    Pool pool;
    g_copies = 0;
    DevKV kv(&pool);
    kv.uniform = true;
    kv.kb = 4, kv.vb = 4, kv.recsize = 16;
    kv.reserve_bytes(MB);  // empty store: sized as asked
    CHECK(kv.cap == MB && g_copies == 0);
    memset(kv.data, 0xAB, MB);
    kv.nbytes = MB, kv.nkv = MB / 16;
    const uint8_t *before = kv.data;
    kv.reserve_bytes(MB + 4096);  // the chunk's free tail is right behind it: doubled in place
    CHECK(kv.data == before && kv.cap == 2 * MB && g_copies == 0);
    kv.nbytes = 2 * MB, kv.nkv = 2 * MB / 16;
    memset(kv.data + MB, 0xCD, MB);
    kv.reserve_bytes(2 * MB + 4096);  // nothing behind it any more: a new block, one copy, contents kept
    CHECK(kv.data != before && kv.cap == 4 * MB && g_copies == 1);
    CHECK(kv.data[0] == 0xAB && kv.data[MB - 1] == 0xAB && kv.data[MB] == 0xCD && kv.data[2 * MB - 1] == 0xCD);
    // the order marker is tied to pair count and address
    kv.mark_sorted(1);
    CHECK(kv.known_sorted(1) && !kv.known_sorted(2) && !kv.known_sorted(-1));
    kv.nkv += 1;
    CHECK(!kv.known_sorted(1));
  }
  printf("POOL_HOST_OK\n");
  return 0;
}

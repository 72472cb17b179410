// MapReduce / KeyValue: host-side mirror of the reference's class surface (src/mapreduce.h:22-131,
// src/keyvalue.h:23-59) over the C-ABI CUDA library.  See mapreduce.h for what runs where.
// Written from scratch against the reference's documented behaviour (doc/Interface_c++.txt and the
// file:line sites cited per method); no reference code is reused.
#include "mapreduce.h"

#include <arpa/inet.h>
#include <dirent.h>
#include <netinet/in.h>
#include <stdio.h>
#include <sys/socket.h>
#include <unistd.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/mrgpu.h"
#include "keyvalue.h"
#include "keymultivalue.h"
#include "mrgpu_functors.h"

#ifndef MRMPI_MEMSIZE
#define MRMPI_MEMSIZE 64  // src/mapreduce.cpp:80
#endif

namespace MAPREDUCE_NS {

static KeyMultiValue g_kmv_marker;  // opaque: callers only test MapReduce::kmv for NULL

MapReduce *MapReduce::mrptr = NULL;
int MapReduce::instances_now = 0;
int MapReduce::instances_ever = 0;
int MapReduce::mpi_finalize_flag = 0;
uint64_t MapReduce::msize = 0, MapReduce::msizemax = 0;
uint64_t MapReduce::rsize = 0, MapReduce::wsize = 0;
uint64_t MapReduce::cssize = 0, MapReduce::crsize = 0;
double MapReduce::commtime = 0.0;

// ---------------------------------------------------------------------------------------------- stock functors
// These bodies are never executed: their ADDRESSES select device functors.
namespace gpu {
void count(char *, int, char *, int, int *, KeyValue *, void *) {}
void sum(char *, int, char *, int, int *, KeyValue *, void *) {}
void cull(char *, int, char *, int, int *, KeyValue *, void *) {}
void postings(char *, int, char *, int, int *, KeyValue *, void *) {}
void nonzero(char *, int, char *, int, int *, KeyValue *, void *) {}
void degree(char *, int, char *, int, int *, KeyValue *, void *) {}
void histo(char *, int, char *, int, int *, KeyValue *, void *) {}
void edge_to_vertices(uint64_t, char *, int, char *, int, KeyValue *, void *) {}
void edge_to_vertex(uint64_t, char *, int, char *, int, KeyValue *, void *) {}
void invert(uint64_t, char *, int, char *, int, KeyValue *, void *) {}
void intcount(int, KeyValue *, void *) {}
void rmat_generate(int, KeyValue *, void *) {}
void fileread(int, char *, KeyValue *, void *) {}
void href(int, char *, KeyValue *, void *) {}
}  // namespace gpu

typedef void (*ReduceFn)(char *, int, char *, int, int *, KeyValue *, void *);
typedef void (*MapMrFn)(uint64_t, char *, int, char *, int, KeyValue *, void *);

// callbacks of the application tied to device functors by address (MapReduce::device_reduce / _map_mr / _map_task)
typedef void (*MapTaskFn)(int, KeyValue *, void *);
// (function-local statics: registration happens from static initialisers of other translation units)
static std::vector<std::pair<ReduceFn, int> > &reg_reduce()
{
  static std::vector<std::pair<ReduceFn, int> > v;
  return v;
}
static std::vector<std::pair<MapMrFn, int> > &reg_mapmr()
{
  static std::vector<std::pair<MapMrFn, int> > v;
  return v;
}
static std::vector<std::pair<MapTaskFn, MapReduce::DeviceMapTask> > &reg_maptask()
{
  static std::vector<std::pair<MapTaskFn, MapReduce::DeviceMapTask> > v;
  return v;
}

void MapReduce::device_reduce(ReduceFn host, int functor) { reg_reduce().push_back(std::make_pair(host, functor)); }
void MapReduce::device_map_mr(MapMrFn host, int functor) { reg_mapmr().push_back(std::make_pair(host, functor)); }
void MapReduce::device_map_task(MapTaskFn host, DeviceMapTask dev) { reg_maptask().push_back(std::make_pair(host, dev)); }

static int stock_reduce(ReduceFn f)
{
  for (size_t i = 0; i < reg_reduce().size(); i++)
    if (reg_reduce()[i].first == f) return reg_reduce()[i].second;
  if (f == gpu::count) return MRGPU_REDUCE_COUNT;
  if (f == gpu::sum) return MRGPU_REDUCE_SUM;
  if (f == gpu::cull) return MRGPU_REDUCE_CULL;
  if (f == gpu::postings) return MRGPU_REDUCE_POSTINGS;
  if (f == gpu::nonzero) return MRGPU_REDUCE_NONZERO;
  if (f == gpu::degree) return MRGPU_REDUCE_DEGREE;
  if (f == gpu::histo) return MRGPU_REDUCE_HISTO;
  return -1;
}
static int stock_mapmr(MapMrFn f)
{
  for (size_t i = 0; i < reg_mapmr().size(); i++)
    if (reg_mapmr()[i].first == f) return reg_mapmr()[i].second;
  if (f == gpu::edge_to_vertices) return MRGPU_MAP_EDGE_TO_VERTICES;
  if (f == gpu::edge_to_vertex) return MRGPU_MAP_EDGE_TO_VERTEX;
  if (f == gpu::invert) return MRGPU_MAP_INVERT;
  return -1;
}

// ---------------------------------------------------------------------------------------------- KeyValue

static const size_t kFlushBytes = 32u << 20;

KeyValue::KeyValue(mrgpu_mr *h) : nkv(0), ksize(0), vsize(0), esize(0), fsize(0), msize(0), page(NULL), memtag(0), npage(0), handle(h) {}

void KeyValue::flush()
{
  if (klen.empty()) return;
  int rc = mrgpu_kv_add(handle, klen.size(), kbuf.data(), klen.data(), -1, vbuf.data(), vlen.data(), -1);
  if (rc != MRGPU_OK) {
    printf("ERROR: %s\n", mrgpu_last_error(handle));  // Error::one text from the library, e.g. pair exceeds page
    fflush(stdout);
    exit(1);
  }
  kbuf.clear();
  vbuf.clear();
  klen.clear();
  vlen.clear();
}

static void kv_fatal(mrgpu_mr *h)
{
  printf("ERROR: %s\n", mrgpu_last_error(h));
  fflush(stdout);
  exit(1);
}

void KeyValue::append()
{
  if (mrgpu_kv_open(handle, 1) != MRGPU_OK) kv_fatal(handle);
}

void KeyValue::complete()
{
  flush();
  if (mrgpu_kv_close(handle, NULL) != MRGPU_OK) kv_fatal(handle);
  mrgpu_stats s;
  if (mrgpu_kv_stats(handle, &s) != MRGPU_OK) kv_fatal(handle);
  nkv = s.npair;
  ksize = s.keysize;
  vsize = s.valuesize;
  esize = s.exactsize;
  npage = (int) s.npage;
}

void KeyValue::add(char *key, int keybytes, char *value, int valuebytes)
{
  kbuf.insert(kbuf.end(), key, key + keybytes);
  if (valuebytes) vbuf.insert(vbuf.end(), value, value + valuebytes);
  klen.push_back(keybytes);
  vlen.push_back(valuebytes);
  if (kbuf.size() + vbuf.size() > kFlushBytes) flush();
}

void KeyValue::add(int n, char *key, int keybytes, char *value, int valuebytes)
{
  for (int i = 0; i < n; i++) add(key + (size_t) i * keybytes, keybytes, value + (size_t) i * valuebytes, valuebytes);
}

void KeyValue::add(int n, char *key, int *keybytes, char *value, int *valuebytes)
{
  for (int i = 0; i < n; i++) {
    add(key, keybytes[i], value, valuebytes[i]);
    key += keybytes[i];
    value += valuebytes[i];
  }
}

// ---------------------------------------------------------------------------------------------- life cycle

#ifdef MRGPU_MPI_STUB
// Without MPI (cpp/mpi.h): the one thing the ranks have to share before NCCL can take over is the 128-byte unique
// id.  Rank 0 makes it and serves it on a loopback TCP socket -- port MRGPU_PORT, else MASTER_PORT + 1 (next to the
// launcher's own rendezvous port), else 29517 -- once per process; every MapReduce object of the process reuses it
// (the library keeps one communicator per id).  One node only, which is what one process per GPU of one box needs.
static bool bootstrap_id(int me, int nprocs, char id[128], std::string &why)
{
  static bool have = false;
  static char cached[128];
  if (have) {
    memcpy(id, cached, 128);
    return true;
  }
  int port = 29517;
  if (getenv("MRGPU_PORT")) port = atoi(getenv("MRGPU_PORT"));
  else if (getenv("MASTER_PORT")) port = atoi(getenv("MASTER_PORT")) + 1;
  struct sockaddr_in addr;
  memset(&addr, 0, sizeof(addr));
  addr.sin_family = AF_INET;
  addr.sin_port = htons((uint16_t) port);
  addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
  if (me == 0) {
    if (mrgpu_comm_unique_id(cached) != MRGPU_OK) {
      why = mrgpu_last_error(NULL);
      return false;
    }
    int srv = socket(AF_INET, SOCK_STREAM, 0), yes = 1;
    setsockopt(srv, SOL_SOCKET, SO_REUSEADDR, &yes, sizeof(yes));
    if (srv < 0 || bind(srv, (struct sockaddr *) &addr, sizeof(addr)) != 0 || listen(srv, nprocs) != 0) {
      why = "rank 0 cannot listen on 127.0.0.1:" + std::to_string(port) + " to hand out the NCCL id (set MRGPU_PORT)";
      if (srv >= 0) ::close(srv);
      return false;
    }
    for (int served = 0; served < nprocs - 1; served++) {
      int c = accept(srv, NULL, NULL);
      if (c < 0) {
        why = "accept() failed while handing out the NCCL id";
        ::close(srv);
        return false;
      }
      size_t off = 0;
      while (off < 128) {
        ssize_t w = ::write(c, cached + off, 128 - off);
        if (w <= 0) break;
        off += (size_t) w;
      }
      ::close(c);
    }
    ::close(srv);
  } else {
    bool got = false;
    for (int attempt = 0; attempt < 2400 && !got; attempt++) {  // up to two minutes for rank 0 to come up
      int c = socket(AF_INET, SOCK_STREAM, 0);
      if (c >= 0 && connect(c, (struct sockaddr *) &addr, sizeof(addr)) == 0) {
        size_t off = 0;
        while (off < 128) {
          ssize_t r = ::read(c, cached + off, 128 - off);
          if (r <= 0) break;
          off += (size_t) r;
        }
        got = off == 128;
      }
      if (c >= 0) ::close(c);
      if (!got) usleep(50000);
    }
    if (!got) {
      why = "could not fetch the NCCL id from rank 0 at 127.0.0.1:" + std::to_string(port);
      return false;
    }
  }
  have = true;
  memcpy(id, cached, 128);
  return true;
}
#endif

void MapReduce::init(MPI_Comm caller)
{
  comm = caller;
  MPI_Comm_rank(comm, &me);
  MPI_Comm_size(comm, &nprocs);
  instances_now++;
  instances_ever++;
  // defaults(): src/mapreduce.cpp:196-262
  mapstyle = 0;
  all2all = 1;
  verbosity = 0;
  timer = 0;
  memsize = MRMPI_MEMSIZE;
  minpage = 0;
  maxpage = 0;
  freepage = 1;
  outofcore = 0;
  zeropage = 0;
  keyalign = valuealign = 4;
  fpath = strdup(".");
  mapfilecount = 0;
  kv = NULL;
  kmv = NULL;
  settings_pushed = 0;
  kmv_block_valid = 0;
  blk_buf = NULL;
  time_start = 0;
  int device = 0;
  if (getenv("MRGPU_DEVICE")) device = atoi(getenv("MRGPU_DEVICE"));
  else if (getenv("LOCAL_RANK")) device = atoi(getenv("LOCAL_RANK"));
  gpu = NULL;
  if (mrgpu_create(device, NULL, &gpu) != MRGPU_OK) {
    printf("ERROR: %s\n", mrgpu_last_error(NULL));
    fflush(stdout);
    exit(1);
  }
  lib_blocks = 0;
#ifdef MRGPU_MPI_STUB
  if (nprocs > 1) {
    char id[128];
    std::string why;
    if (!bootstrap_id(me, nprocs, id, why)) all(why.c_str());
    if (mrgpu_comm_init(gpu, me, nprocs, id) != MRGPU_OK) all(mrgpu_last_error(gpu));
  }
#else
  if (nprocs > 1) {  // real MPI: NCCL bootstrap = one broadcast of 128 bytes
    char id[128];
    if (me == 0 && mrgpu_comm_unique_id(id) != MRGPU_OK) all(mrgpu_last_error(NULL));
    MPI_Bcast(id, 128, MPI_BYTE, 0, comm);
    if (mrgpu_comm_init(gpu, me, nprocs, id) != MRGPU_OK) all(mrgpu_last_error(gpu));
  }
#endif
}

MapReduce::MapReduce(MPI_Comm caller) { init(caller); }
MapReduce::MapReduce()
{
  int flag;
  MPI_Initialized(&flag);
  if (!flag) {
    int argc = 0;
    char **argv = NULL;
    MPI_Init(&argc, &argv);
  }
  init(MPI_COMM_WORLD);
}
MapReduce::MapReduce(double)
{
  int flag;
  MPI_Initialized(&flag);
  if (!flag) {
    int argc = 0;
    char **argv = NULL;
    MPI_Init(&argc, &argv);
  }
  init(MPI_COMM_WORLD);
  mpi_finalize_flag = 1;
}

MapReduce::~MapReduce()
{
  delete kv;
  free(fpath);
  free(blk_buf);
  if (gpu) mrgpu_destroy(gpu);
  instances_now--;
  if (mpi_finalize_flag && instances_now == 0) MPI_Finalize();
}

void MapReduce::all(const char *str)
{
  if (me == 0) {
    printf("ERROR: %s\n", str);
    fflush(stdout);
  }
  MPI_Finalize();
  exit(1);
}

void MapReduce::one(const char *str)
{
  char name[MPI_MAX_PROCESSOR_NAME + 1];
  int len;
  MPI_Get_processor_name(name, &len);
  name[len] = '\0';
  printf("ERROR on proc %d (%s): %s\n", me, name, str);
  fflush(stdout);
  MPI_Abort(comm, 1);
}

uint64_t MapReduce::ck(int rc, uint64_t n)
{
  if (rc != MRGPU_OK) all(mrgpu_last_error(gpu));
  return n;
}

// settings are public ints the caller may change until first use (src/mapreduce.cpp:3318-3357)
void MapReduce::push_settings()
{
  if (settings_pushed) return;
  settings_pushed = 1;
  ck(mrgpu_set(gpu, "memsize", memsize), 0);
  ck(mrgpu_set(gpu, "keyalign", keyalign), 0);
  ck(mrgpu_set(gpu, "valuealign", valuealign), 0);
  ck(mrgpu_set(gpu, "outofcore", outofcore), 0);  // 1: stores beyond one page live in host memory (the library's out-of-core tier)
}

void MapReduce::set_fpath(const char *str)
{
  free(fpath);
  fpath = strdup(str);  // kept for compatibility; nothing is spilled to files
}

void MapReduce::start_timer()
{
  if (timer) MPI_Barrier(comm);
  time_start = MPI_Wtime();
}

// src/mapreduce.cpp:3112-3165: "<Op> time (secs) = t" then, with verbosity, the size line of the result
void MapReduce::stats(const char *heading, int which)
{
  if (timer) {
    MPI_Barrier(comm);
    if (me == 0) printf("%s time (secs) = %g\n", heading, MPI_Wtime() - time_start);
  }
  if (verbosity == 0) return;
  if (me == 0) printf("%s: ", heading);
  if (which == 0) kv_stats(1);
  else kmv_stats(1);
}

void MapReduce::sync_kv_fields()
{
  if (!kv) return;
  mrgpu_stats s;
  if (mrgpu_kv_stats(gpu, &s) != MRGPU_OK) return;
  kv->nkv = s.npair;
  kv->ksize = s.keysize;
  kv->vsize = s.valuesize;
  kv->esize = s.exactsize;
  kv->npage = (int) s.npage;
}

uint64_t MapReduce::kv_stats(int level)
{
  if (kv == NULL) all("Cannot print stats without KeyValue");
  mrgpu_stats s;
  ck(mrgpu_kv_stats(gpu, &s), 0);
  uint64_t nall = s.npair;  // the reference returns (and prints) the pair count summed over all ranks (src/mapreduce.cpp:2945-2958)
  ck(mrgpu_global_count(gpu, 0, &nall), 0);
  const double mbyte = 1024.0 * 1024.0;
  if (level >= 1 && me == 0)
    printf("%lu pairs, %.3g Mb keys, %.3g Mb values, %.3g Mb, %d pages\n", (unsigned long) nall, s.keysize / mbyte, s.valuesize / mbyte,
           s.exactsize / mbyte, (int) s.npage);
  return nall;
}

uint64_t MapReduce::kmv_stats(int level)
{
  if (kmv == NULL) all("Cannot print stats without KeyMultiValue");
  mrgpu_stats s;
  ck(mrgpu_kmv_stats(gpu, &s), 0);
  uint64_t nall = s.npair;
  ck(mrgpu_global_count(gpu, 1, &nall), 0);
  const double mbyte = 1024.0 * 1024.0;
  if (level >= 1 && me == 0)
    printf("%lu pairs, %.3g Mb keys, %.3g Mb values, %.3g Mb, %d pages\n", (unsigned long) nall, s.keysize / mbyte, s.valuesize / mbyte,
           s.exactsize / mbyte, (int) s.npage);
  return nall;
}

void MapReduce::cummulative_stats(int, int) {}

// ---------------------------------------------------------------------------------------------- page walks

static inline uint64_t ru(uint64_t x, int a) { return (x + a - 1) / a * a; }

// calls f(key, keybytes, value, valuebytes) for every pair of handle h, page by page (src/mapreduce.cpp:1949-1967)
template <typename F>
void MapReduce::walk_kv(mrgpu_mr *h, F f)
{
  mrgpu_stats s;
  ck(mrgpu_kv_stats(h, &s), 0);
  const int ka = keyalign, va = valuealign, ta = std::max(std::max(ka, va), 4);
  std::vector<char> page;
  for (uint64_t ip = 0; ip < s.npage; ip++) {
    uint64_t nkey, asz;
    ck(mrgpu_kv_page_info(h, ip, &nkey, NULL, NULL, &asz), 0);
    page.resize(asz ? asz : 1);
    ck(mrgpu_kv_page_copy(h, ip, page.data()), 0);
    uint64_t p = 0;
    for (uint64_t i = 0; i < nkey; i++) {
      int kb = *(int *) &page[p], vb = *(int *) &page[p + 4];
      uint64_t k0 = ru(p + 8, ka), v0 = ru(k0 + kb, va);
      f(&page[k0], kb, &page[v0], vb);
      p = ru(v0 + vb, ta);
    }
  }
}

// calls f(key, keybytes, multivalue, nvalues, valuesizes) for every KMV pair; a pair bigger than a page is
// presented through the reference's block protocol (multivalue == NULL, nvalues == 0, valuesizes == (int*)this;
// src/mapreduce.cpp:1828-1848)
template <typename F>
void MapReduce::walk_kmv(mrgpu_mr *h, F f)
{
  mrgpu_stats s;
  ck(mrgpu_kmv_stats(h, &s), 0);
  const int ka = keyalign, va = valuealign, ta = std::max(std::max(ka, va), 4);
  int64_t ms = memsize;
  const uint64_t pagesize = ms > 0 ? (uint64_t) ms << 20 : (uint64_t) (-ms);
  std::vector<char> page;
  for (uint64_t ip = 0; ip < s.npage; ip++) {
    uint64_t nkey, asz;
    int ext;
    ck(mrgpu_kmv_page_info(h, ip, &nkey, &asz, &ext), 0);
    page.resize(asz ? asz : 1);
    ck(mrgpu_kmv_page_copy(h, ip, page.data()), 0);
    uint64_t p = 0;
    for (uint64_t i = 0; i < nkey; i++) {
      int nv = *(int *) &page[p], kb = *(int *) &page[p + 4], mvb = *(int *) &page[p + 8];
      int *sizes = (int *) &page[p + 12];
      uint64_t k0 = ru(p + 12 + 4 * (uint64_t) nv, ka), v0 = ru(k0 + kb, va);
      if (!ext) {
        f(&page[k0], kb, &page[v0], nv, sizes);
      } else {
        kmv_block_valid = 1;
        kmv_nvalue_total = (uint64_t) nv;
        blk_capacity = std::max<uint64_t>(1, pagesize / 2 / sizeof(int) - 1);  // src/keymultivalue.cpp:1241-1242
        kmv_nblock = (int) ((kmv_nvalue_total + blk_capacity - 1) / blk_capacity);
        blk_sizes_base = (const char *) sizes;
        blk_values_base = &page[v0];
        f(&page[k0], kb, (char *) NULL, 0, (int *) this);
        kmv_block_valid = 0;
      }
      p = ru(v0 + (uint64_t) mvb, ta);
    }
  }
}

// While the LIBRARY walks the KMV (scan), the block state lives there (mrgpu_multivalue_*); while this class walks it
// (reduce / compress with a host callback) it lives here.
uint64_t MapReduce::multivalue_blocks(int &nblock)
{
  if (!kmv_block_valid && lib_blocks) {
    uint64_t total = 0;
    if (mrgpu_multivalue_blocks(gpu, &nblock, &total) != MRGPU_OK) one(mrgpu_last_error(gpu));
    return total;
  }
  if (!kmv_block_valid) one("Invalid call to multivalue_blocks()");
  nblock = kmv_nblock;
  return kmv_nvalue_total;
}

void MapReduce::multivalue_block_select(int which)
{
  if (!kmv_block_valid && lib_blocks) {
    if (mrgpu_multivalue_block_select(gpu, which) != MRGPU_OK) one(mrgpu_last_error(gpu));
    return;
  }
  if (!kmv_block_valid) one("Invalid call to multivalue_block_select()");
  if (which != 1 && which != 2) one("Invalid arg to multivalue_block_select()");
}

int MapReduce::multivalue_block(int iblock, char **pmultivalue, int **pvaluesizes)
{
  if (!kmv_block_valid && lib_blocks) {
    int nv = 0;
    if (mrgpu_multivalue_block(gpu, iblock, pmultivalue, pvaluesizes, &nv) != MRGPU_OK) one(mrgpu_last_error(gpu));
    return nv;
  }
  if (!kmv_block_valid) one("Invalid call to multivalue_block()");
  if (iblock < 0 || iblock >= kmv_nblock) one("Invalid page request to multivalue_block()");
  const int *sizes = (const int *) blk_sizes_base;
  uint64_t first = (uint64_t) iblock * blk_capacity;
  uint64_t n = std::min<uint64_t>(blk_capacity, kmv_nvalue_total - first);
  uint64_t off = 0;
  for (uint64_t i = 0; i < first; i++) off += sizes[i];
  *pvaluesizes = (int *) (sizes + first);
  *pmultivalue = (char *) blk_values_base + off;
  return (int) n;
}

// ---------------------------------------------------------------------------------------------- emit helpers

KeyValue *MapReduce::begin_emit(mrgpu_mr *target, int addflag)
{
  if (mrgpu_kv_open(target, addflag) != MRGPU_OK) all(mrgpu_last_error(target));
  return new KeyValue(target);
}

uint64_t MapReduce::end_emit(mrgpu_mr *target, KeyValue *k)
{
  k->flush();
  delete k;
  uint64_t n = 0;
  if (mrgpu_kv_close(target, &n) != MRGPU_OK) all(mrgpu_last_error(target));
  return n;
}

// a second handle on the same device with the same layout settings: host callbacks that consume this
// object's KV/KMV emit into it, then the result is moved over (the reference builds a new KV beside the old
// one in the same way, e.g. src/mapreduce.cpp:1584-1602)
static mrgpu_mr *make_side_handle(MapReduce *mr, int memsize, int keyalign, int valuealign)
{
  mrgpu_mr *t = NULL;
  int device = 0;
  if (getenv("MRGPU_DEVICE")) device = atoi(getenv("MRGPU_DEVICE"));
  else if (getenv("LOCAL_RANK")) device = atoi(getenv("LOCAL_RANK"));
  if (mrgpu_create(device, NULL, &t) != MRGPU_OK) return NULL;
  mrgpu_set(t, "memsize", memsize);
  mrgpu_set(t, "keyalign", keyalign);
  mrgpu_set(t, "valuealign", valuealign);
  (void) mr;
  return t;
}

// ---------------------------------------------------------------------------------------------- map

// task ids of this rank: mapstyle 0 = contiguous chunk, 1 = strided (src/mapreduce.cpp:1136-1170);
// 2 (master/slave) degenerates to chunks -- tasks are independent, only their placement differs
static void my_tasks(int nmap, int me, int nprocs, int mapstyle, std::vector<int> &out)
{
  if (mapstyle == 1) {
    for (int i = me; i < nmap; i += nprocs) out.push_back(i);
  } else {
    uint64_t lo = (uint64_t) me * nmap / nprocs, hi = (uint64_t) (me + 1) * nmap / nprocs;
    for (uint64_t i = lo; i < hi; i++) out.push_back((int) i);
  }
}

// whole file -> pinned-free host buffer (the device functors take host or device memory)
static bool slurp(const char *fname, std::vector<char> &buf)
{
  FILE *fp = fopen(fname, "rb");
  if (!fp) return false;
  fseek(fp, 0, SEEK_END);
  long sz = ftell(fp);
  fseek(fp, 0, SEEK_SET);
  buf.resize(sz > 0 ? (size_t) sz : 0);
  size_t got = sz > 0 ? fread(buf.data(), 1, (size_t) sz, fp) : 0;
  fclose(fp);
  buf.resize(got);
  return true;
}

uint64_t MapReduce::map(int nmap, void (*appmap)(int, KeyValue *, void *), void *ptr, int addflag)
{
  push_settings();
  for (size_t i = 0; i < reg_maptask().size(); i++)
    if (reg_maptask()[i].first == appmap) return reg_maptask()[i].second(this, nmap, ptr, addflag);
  start_timer();
  if (appmap == gpu::intcount || appmap == gpu::rmat_generate) {
    // stock callback = device functor: one library call per task of this rank, no pair touches the host.  The calls
    // with a NULL count do not communicate; the last one (zero extra pairs) returns the global count.
    std::vector<int> tasks;
    my_tasks(nmap, me, nprocs, mapstyle, tasks);
    int flag = addflag;
    for (size_t i = 0; i < tasks.size(); i++) {
      if (appmap == gpu::intcount) {
        const gpu::IntCountInput *in = (const gpu::IntCountInput *) ptr;
        ck(mrgpu_map_intcount(gpu, in->task_keys[tasks[i]], in->task_nkeys[tasks[i]], flag, NULL), 0);
      } else {
        gpu::RmatInput *in = (gpu::RmatInput *) ptr;
        ck(mrgpu_map_rmat(gpu, in->ngenerate, in->nlevels, in->a, in->b, in->c, in->d, in->fraction, &in->state48, flag, NULL), 0);
      }
      flag = 1;
    }
    uint64_t n = 0;  // zero more pairs of the same shape: the call that returns the global count
    if (appmap == gpu::intcount) ck(mrgpu_map_intcount(gpu, NULL, 0, flag, &n), 0);
    else {
      gpu::RmatInput *in = (gpu::RmatInput *) ptr;
      ck(mrgpu_map_rmat(gpu, 0, in->nlevels, in->a, in->b, in->c, in->d, in->fraction, &in->state48, flag, &n), 0);
    }
    delete kv;
    kv = new KeyValue(gpu);
    kmv = NULL;
    sync_kv_fields();
    stats("Map", 0);
    return n;
  }
  KeyValue *out = begin_emit(gpu, addflag);
  std::vector<int> tasks;
  my_tasks(nmap, me, nprocs, mapstyle, tasks);
  for (size_t i = 0; i < tasks.size(); i++) appmap(tasks[i], out, ptr);
  uint64_t n = end_emit(gpu, out);
  delete kv;
  kv = new KeyValue(gpu);
  kmv = NULL;
  sync_kv_fields();
  stats("Map", 0);
  return n;
}

static void list_files(const std::string &path, int recurse, std::vector<std::string> &out)
{
  struct stat st;
  if (stat(path.c_str(), &st) != 0) return;
  if (S_ISDIR(st.st_mode)) {
    DIR *d = opendir(path.c_str());
    if (!d) return;
    std::vector<std::string> names;
    while (struct dirent *e = readdir(d)) {
      if (e->d_name[0] == '.') continue;
      names.push_back(e->d_name);
    }
    closedir(d);
    std::sort(names.begin(), names.end());
    for (size_t i = 0; i < names.size(); i++) {
      std::string child = path + "/" + names[i];
      struct stat cs;
      if (stat(child.c_str(), &cs) != 0) continue;
      if (S_ISDIR(cs.st_mode)) {
        if (recurse) list_files(child, recurse, out);
      } else
        out.push_back(child);
    }
  } else
    out.push_back(path);
}

// map(nstr, strings, self, recurse, readfile, mymap, ptr, addflag): src/mapreduce.cpp:1060-1092.  Each string
// is a file, a directory (expanded, recursively if asked) or, with readfile, a file listing one name per line.
uint64_t MapReduce::map(int nstr, char **strings, int self, int recurse, int readfile, void (*appmap)(int, char *, KeyValue *, void *), void *ptr,
                        int addflag)
{
  push_settings();
  start_timer();
  std::vector<std::string> files;
  for (int i = 0; i < nstr; i++) {
    if (readfile) {
      FILE *fp = fopen(strings[i], "r");
      if (!fp) one("Could not open file of file names");
      char line[4096];
      while (fgets(line, sizeof(line), fp)) {
        char *w = strtok(line, " \t\n\f\r");
        if (w) list_files(w, recurse, files);
      }
      fclose(fp);
    } else
      list_files(strings[i], recurse, files);
  }
  (void) self;
  mapfilecount = (int) files.size();
  if (appmap == gpu::fileread || appmap == gpu::href) {  // stock callback = device functor (tokenizer / href scanner)
    std::vector<int> tasks;
    my_tasks((int) files.size(), me, nprocs, mapstyle, tasks);
    int flag = addflag;
    std::vector<char> buf;
    for (size_t i = 0; i < tasks.size(); i++) {
      const std::string &f = files[tasks[i]];
      if (!slurp(f.c_str(), buf)) one("Could not open file");
      if (appmap == gpu::fileread) ck(mrgpu_map_tokenize(gpu, buf.data(), buf.size(), flag, NULL), 0);
      else {
        size_t slash = f.find_last_of('/');  // cpu/InvertedIndex.cpp:133: the value is the file's base name
        ck(mrgpu_map_href(gpu, buf.data(), buf.size(), f.c_str() + (slash == std::string::npos ? 0 : slash + 1), flag, NULL), 0);
      }
      flag = 1;
    }
    uint64_t n = 0;  // zero more bytes: the call that returns the global count
    if (appmap == gpu::fileread) ck(mrgpu_map_tokenize(gpu, NULL, 0, flag, &n), 0);
    else ck(mrgpu_map_href(gpu, NULL, 0, "", flag, &n), 0);
    delete kv;
    kv = new KeyValue(gpu);
    kmv = NULL;
    sync_kv_fields();
    stats("Map", 0);
    return n;
  }
  KeyValue *out = begin_emit(gpu, addflag);
  std::vector<int> tasks;
  my_tasks((int) files.size(), me, nprocs, mapstyle, tasks);
  for (size_t i = 0; i < tasks.size(); i++) appmap(tasks[i], (char *) files[tasks[i]].c_str(), out, ptr);
  uint64_t n = end_emit(gpu, out);
  delete kv;
  kv = new KeyValue(gpu);
  kmv = NULL;
  sync_kv_fields();
  stats("Map", 0);
  return n;
}

// map(nmap, nstr, strings, recurse, readflag, sepchar | sepstr, delta, mymap, ptr, addflag): the files are cut into
// nmap chunks at separators and mymap(itask, str, size, kv, ptr) sees one chunk (src/mapreduce.cpp:1232-1536).  The
// rules that decide what a task sees are the reference's: tasks per file proportional to file size (at least one,
// adjusted until they sum to nmap, reduced where a chunk would not exceed delta); task j of n of a file reads bytes
// [j*size/n, (j+1)*size/n + delta); a task that is not the file's first starts after (char) / at (string) the first
// separator found within delta; one that is not the last ends after (char) / before (string) the first separator at
// or beyond its nominal end.  The chunk is NUL terminated and size counts the NUL.
uint64_t MapReduce::map_chunks(int nmap, int nstr, char **strings, int recurse, int readflag, int sepwhich, char sepchar, const char *sepstr,
                               int delta, void (*appmap)(int, char *, int, KeyValue *, void *), void *ptr, int addflag)
{
  push_settings();
  start_timer();
  std::vector<std::string> files;
  for (int i = 0; i < nstr; i++) {
    if (readflag) {
      FILE *fp = fopen(strings[i], "r");
      if (!fp) one("Could not open file of file names");
      char line[4096];
      while (fgets(line, sizeof(line), fp)) {
        char *w = strtok(line, " \t\n\f\r");
        if (w) list_files(w, recurse, files);
      }
      fclose(fp);
    } else
      list_files(strings[i], recurse, files);
  }
  const int nfile = (int) files.size();
  mapfilecount = nfile;
  if (nfile > nmap) nmap = nfile;
  std::vector<uint64_t> fsize(nfile);
  uint64_t ntotal = 0;
  for (int i = 0; i < nfile; i++) {
    struct stat st;
    if (stat(files[i].c_str(), &st) < 0) one("Could not query file size");
    fsize[i] = (uint64_t) st.st_size;
    ntotal += fsize[i];
  }
  const uint64_t nideal = std::max<uint64_t>(1, nmap ? ntotal / nmap : 1);
  std::vector<int> per(nfile);
  int ntasks = 0;
  for (int i = 0; i < nfile; i++) {
    per[i] = (int) std::max<uint64_t>(1, fsize[i] / nideal);
    ntasks += per[i];
  }
  while (ntasks < nmap) {
    const int before = ntasks;
    for (int i = 0; i < nfile && ntasks < nmap; i++)
      if (fsize[i] > nideal) per[i]++, ntasks++;
    if (ntasks == before) break;  // no file is big enough to take another task
  }
  while (ntasks > nmap) {
    const int before = ntasks;
    for (int i = 0; i < nfile && ntasks > nmap; i++)
      if (per[i] > 1) per[i]--, ntasks--;
    if (ntasks == before) break;
  }
  nmap = ntasks;
  bool shrunk = false;
  for (int i = 0; i < nfile; i++) {
    if (fsize[i] / per[i] > (uint64_t) delta) continue;
    shrunk = true;
    while (per[i] > 1) {
      per[i]--;
      nmap--;
      if (fsize[i] / per[i] > (uint64_t) delta) break;
    }
  }
  if (shrunk && me == 0) printf("WARNING: File(s) too small for file delta - decreased map tasks to %d\n", nmap);
  std::vector<int> whichfile, whichtask;
  for (int i = 0; i < nfile; i++)
    for (int j = 0; j < per[i]; j++) whichfile.push_back(i), whichtask.push_back(j);

  KeyValue *out = begin_emit(gpu, addflag);
  std::vector<int> tasks;
  my_tasks(nmap, me, nprocs, mapstyle, tasks);
  std::vector<char> str;
  for (size_t t = 0; t < tasks.size(); t++) {
    const int imap = tasks[t], f = whichfile[imap], itask = whichtask[imap], ntask = per[f];
    const uint64_t filesize = fsize[f];
    const uint64_t readstart = (uint64_t) itask * filesize / ntask, readnext = (uint64_t) (itask + 1) * filesize / ntask;
    if (readnext - readstart + delta + 1 > 0x7FFFFFFFull) one("Single file read exceeds int size");
    uint64_t readsize = std::min<uint64_t>(readnext - readstart + delta, filesize - readstart);
    str.assign(readsize + 1, 0);
    FILE *fp = fopen(files[f].c_str(), "rb");
    if (!fp) one("Could not open map file");
    fseek(fp, (long) readstart, SEEK_SET);
    const size_t got = fread(str.data(), 1, readsize, fp);
    fclose(fp);
    str[got] = 0;
    size_t strstart = 0;
    if (itask > 0) {
      char *p = sepwhich ? strchr(str.data(), sepchar) : strstr(str.data(), sepstr);
      if (p == NULL || p - str.data() > delta) one("Could not find file separator within delta");
      strstart = (size_t) (p - str.data()) + (sepwhich ? 1 : 0);
    }
    size_t strstop = readsize;
    if (itask < ntask - 1) {
      char *from = &str[readnext - readstart];
      char *p = sepwhich ? strchr(from, sepchar) : strstr(from, sepstr);
      if (p == NULL) one("Could not find file separator within delta");
      if (sepwhich) p++;
      *p = 0;
      strstop = (size_t) (p - str.data());
    }
    appmap(imap, &str[strstart], (int) (strstop - strstart + 1), out, ptr);
  }
  uint64_t n = end_emit(gpu, out);
  delete kv;
  kv = new KeyValue(gpu);
  kmv = NULL;
  sync_kv_fields();
  stats("Map", 0);
  return n;
}

uint64_t MapReduce::map(int nmap, int nstr, char **strings, int recurse, int readflag, char sepchar, int delta,
                        void (*appmap)(int, char *, int, KeyValue *, void *), void *ptr, int addflag)
{
  return map_chunks(nmap, nstr, strings, recurse, readflag, 1, sepchar, NULL, delta, appmap, ptr, addflag);
}
uint64_t MapReduce::map(int nmap, int nstr, char **strings, int recurse, int readflag, char *sepstr, int delta,
                        void (*appmap)(int, char *, int, KeyValue *, void *), void *ptr, int addflag)
{
  return map_chunks(nmap, nstr, strings, recurse, readflag, 0, 0, sepstr, delta, appmap, ptr, addflag);
}

// map(mr2, mymap, ptr, addflag): src/mapreduce.cpp:1560-1642
uint64_t MapReduce::map(MapReduce *mr, MapMrFn appmap, void *ptr, int addflag)
{
  if (mr->kv == NULL) all("MapReduce passed to map() does not have KeyValue pairs");
  push_settings();
  mr->push_settings();
  start_timer();
  uint64_t n = 0;
  int f = stock_mapmr(appmap);
  if (f >= 0) {
    ck(mrgpu_map_mr(gpu, mr->gpu, f, addflag, &n), 0);
  } else if (mr == this) {
    mrgpu_mr *side = make_side_handle(this, memsize, keyalign, valuealign);
    if (!side) all(mrgpu_last_error(NULL));
    KeyValue *out = begin_emit(side, 0);
    uint64_t itask = 0;
    walk_kv(gpu, [&](char *k, int kb, char *v, int vb) { appmap(itask++, k, kb, v, vb, out, ptr); });
    end_emit(side, out);
    ck(mrgpu_kv_open(gpu, 0), 0);  // drops the old KV
    ck(mrgpu_kv_close(gpu, &n), 0);
    ck(mrgpu_add(gpu, side, &n), 0);
    mrgpu_destroy(side);
  } else {
    KeyValue *out = begin_emit(gpu, addflag);
    uint64_t itask = 0;
    walk_kv(mr->gpu, [&](char *k, int kb, char *v, int vb) { appmap(itask++, k, kb, v, vb, out, ptr); });
    n = end_emit(gpu, out);
  }
  delete kv;
  kv = new KeyValue(gpu);
  kmv = NULL;
  sync_kv_fields();
  stats("Map", 0);
  return n;
}

// open()/close(): src/mapreduce.cpp:1650-1669 -- pairs added to mr->kv between the two calls
void MapReduce::open(int addflag)
{
  push_settings();
  ck(mrgpu_kv_open(gpu, addflag), 0);
  delete kv;
  kv = new KeyValue(gpu);
  kmv = NULL;
}

uint64_t MapReduce::close()
{
  if (!kv) all("Cannot close without KeyValue");
  kv->flush();
  uint64_t n = 0;
  ck(mrgpu_kv_close(gpu, &n), 0);
  sync_kv_fields();
  return n;
}

// ---------------------------------------------------------------------------------------------- the hot path

uint64_t MapReduce::add(MapReduce *mr)
{
  if (mr->kv == NULL) all("MapReduce passed to add() does not have KeyValue pairs");
  if (mr == this) all("Cannot add to self");
  if (kmv) all("Cannot add to a MapReduce with KeyMultiValue pairs");
  push_settings();
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_add(gpu, mr->gpu, &n), 0);
  if (!kv) kv = new KeyValue(gpu);
  sync_kv_fields();
  stats("Add", 0);
  return n;
}

uint64_t MapReduce::aggregate(int (*hash)(char *, int))
{
  if (kv == NULL) all("Cannot aggregate without KeyValue");
  start_timer();
  uint64_t n = 0;
  if (hash && nprocs > 1) {
    // the user's hash is host code: evaluate it for every pair (src/mapreduce.cpp:466-470: hash(key,keybytes) % nprocs),
    // the exchange itself runs on the GPUs with these destinations
    std::vector<int32_t> dest;
    walk_kv(gpu, [&](char *k, int kb, char *, int) { dest.push_back(hash(k, kb) % nprocs); });
    ck(mrgpu_aggregate_dest(gpu, dest.data(), dest.size(), &n), 0);
  } else
    ck(mrgpu_aggregate(gpu, &n), 0);
  sync_kv_fields();
  stats("Aggregate", 0);
  return n;
}

uint64_t MapReduce::convert()
{
  if (kv == NULL) all("Cannot convert without KeyValue");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_convert(gpu, &n), 0);
  delete kv;
  kv = NULL;
  kmv = &g_kmv_marker;
  stats("Convert", 1);
  return n;
}

uint64_t MapReduce::collate(int (*hash)(char *, int))
{
  if (kv == NULL) all("Cannot collate without KeyValue");
  start_timer();
  uint64_t n = 0;
  if (hash && nprocs > 1) {  // src/mapreduce.cpp:726-727: aggregate(hash); convert();
    int vsave = verbosity, tsave = timer;
    verbosity = timer = 0;
    aggregate(hash);
    verbosity = vsave, timer = tsave;
    ck(mrgpu_convert(gpu, &n), 0);
  } else
    ck(mrgpu_collate(gpu, &n), 0);
  delete kv;
  kv = NULL;
  kmv = &g_kmv_marker;
  stats("Collate", 1);
  return n;
}

uint64_t MapReduce::reduce_impl(ReduceFn appreduce, void *ptr, const char *heading)
{
  uint64_t n = 0;
  int f = stock_reduce(appreduce);
  if (f >= 0) {
    ck(mrgpu_reduce(gpu, f, &n), 0);
  } else {
    mrgpu_mr *side = make_side_handle(this, memsize, keyalign, valuealign);
    if (!side) all(mrgpu_last_error(NULL));
    KeyValue *out = begin_emit(side, 0);
    walk_kmv(gpu, [&](char *k, int kb, char *mv, int nv, int *sizes) { appreduce(k, kb, mv, nv, sizes, out, ptr); });
    end_emit(side, out);
    ck(mrgpu_kv_open(gpu, 0), 0);  // drops the KMV, leaves an empty KV
    ck(mrgpu_kv_close(gpu, &n), 0);
    ck(mrgpu_add(gpu, side, &n), 0);
    mrgpu_destroy(side);
  }
  kmv = NULL;
  delete kv;
  kv = new KeyValue(gpu);
  sync_kv_fields();
  stats(heading, 0);
  return n;
}

uint64_t MapReduce::reduce(ReduceFn appreduce, void *ptr)
{
  if (kmv == NULL) all("Cannot reduce without KeyMultiValue");
  start_timer();
  return reduce_impl(appreduce, ptr, "Reduce");
}

// compress(): convert + reduce on this rank's pairs only, no communication (src/mapreduce.cpp:749-851)
uint64_t MapReduce::compress(ReduceFn appcompress, void *ptr)
{
  if (kv == NULL) all("Cannot compress without KeyValue");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_convert(gpu, &n), 0);
  delete kv;
  kv = NULL;
  kmv = &g_kmv_marker;
  return reduce_impl(appcompress, ptr, "Compress");
}

// scan(): the library stages the pages and calls back (mrgpu_scan_kv / mrgpu_scan_kmv); the trampolines only restore
// the reference's convention of passing the MapReduce object in the valuesizes slot of a multi-block pair
struct ScanCtxKV {
  void (*fn)(char *, int, char *, int, void *);
  void *ptr;
};
struct ScanCtxKMV {
  void (*fn)(char *, int, char *, int, int *, void *);
  void *ptr;
  MapReduce *self;
  mrgpu_mr *handle;
};
static void scan_tramp_kv(char *k, int kb, char *v, int vb, void *p)
{
  ScanCtxKV *c = (ScanCtxKV *) p;
  c->fn(k, kb, v, vb, c->ptr);
}
static void scan_tramp_kmv(char *k, int kb, char *mv, int nv, int *sizes, void *p)
{
  ScanCtxKMV *c = (ScanCtxKMV *) p;
  if (mv == NULL && nv == 0 && (void *) sizes == (void *) c->handle) sizes = (int *) c->self;
  c->fn(k, kb, mv, nv, sizes, c->ptr);
}

uint64_t MapReduce::scan(void (*appscan)(char *, int, char *, int, void *), void *ptr)
{
  if (kv == NULL) all("Cannot scan without KeyValue");
  start_timer();
  ScanCtxKV c = {appscan, ptr};
  uint64_t n = 0;
  ck(mrgpu_scan_kv(gpu, scan_tramp_kv, &c, &n), 0);
  stats("Scan", 0);
  return n;
}

uint64_t MapReduce::scan(void (*appscan)(char *, int, char *, int, int *, void *), void *ptr)
{
  if (kmv == NULL) all("Cannot scan without KeyMultiValue");
  start_timer();
  ScanCtxKMV c = {appscan, ptr, this, gpu};
  uint64_t n = 0;
  lib_blocks = 1;
  int rc = mrgpu_scan_kmv(gpu, scan_tramp_kmv, &c, &n);
  lib_blocks = 0;
  ck(rc, 0);
  stats("Scan", 1);
  return n;
}

// ---------------------------------------------------------------------------------------------- sorting

uint64_t MapReduce::sort_keys(int flag)
{
  if (kv == NULL) all("Cannot sort_keys without KeyValue");
  int a = flag < 0 ? -flag : flag;
  if (a < 1 || a > 6) all("Invalid compare method for sort keys");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_sort_keys(gpu, flag, &n), 0);
  sync_kv_fields();
  stats("Sort_keys", 0);
  return n;
}

// A user compare function is host code: the pairs come to the host, are stably sorted (the reference's
// qsort is a merge sort here) and go back.  by_value = 0 sorts by key, 1 by value.
uint64_t MapReduce::host_sort(int (*appcompare)(char *, int, char *, int), int by_value)
{
  struct Rec {
    size_t koff, voff;
    int kb, vb;
  };
  std::vector<char> pool;
  std::vector<Rec> recs;
  walk_kv(gpu, [&](char *k, int kb, char *v, int vb) {
    Rec r;
    r.koff = pool.size();
    pool.insert(pool.end(), k, k + kb);
    pool.push_back(0);  // callers' compare functions may read a terminating NUL that sits in the padding
    r.voff = pool.size();
    pool.insert(pool.end(), v, v + vb);
    pool.push_back(0);
    r.kb = kb;
    r.vb = vb;
    recs.push_back(r);
  });
  std::stable_sort(recs.begin(), recs.end(), [&](const Rec &x, const Rec &y) {
    if (by_value) return appcompare(&pool[x.voff], x.vb, &pool[y.voff], y.vb) < 0;
    return appcompare(&pool[x.koff], x.kb, &pool[y.koff], y.kb) < 0;
  });
  KeyValue *out = begin_emit(gpu, 0);
  for (size_t i = 0; i < recs.size(); i++) out->add(&pool[recs[i].koff], recs[i].kb, &pool[recs[i].voff], recs[i].vb);
  uint64_t n = end_emit(gpu, out);
  sync_kv_fields();
  return n;
}

uint64_t MapReduce::sort_keys(int (*appcompare)(char *, int, char *, int))
{
  if (kv == NULL) all("Cannot sort_keys without KeyValue");
  start_timer();
  uint64_t n = host_sort(appcompare, 0);
  stats("Sort_keys", 0);
  return n;
}

uint64_t MapReduce::sort_values(int (*appcompare)(char *, int, char *, int))
{
  if (kv == NULL) all("Cannot sort_values without KeyValue");
  start_timer();
  uint64_t n = host_sort(appcompare, 1);
  stats("Sort_values", 0);
  return n;
}

uint64_t MapReduce::sort_values(int flag)
{
  if (kv == NULL) all("Cannot sort_values without KeyValue");
  int a = flag < 0 ? -flag : flag;
  if (a < 1 || a > 6) all("Invalid compare method for sort values");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_sort_values(gpu, flag, &n), 0);
  sync_kv_fields();
  stats("Sort_values", 0);
  return n;
}

// sort_multivalues(): the values of every KMV pair sorted among themselves (src/mapreduce.cpp:2189-2350).  The compare
// is host code (a function pointer, or one of the flag functions below), so the pairs come to the host page by
// page; each pair's values are sorted and go back as (key, value) pairs in sorted order, and convert() -- a STABLE
// group-by: values keep their arrival order -- rebuilds the KMV.  Pairs bigger than a page arrive block by block.
static int cmp_flag(int flag, const char *a, int alen, const char *b, int blen)
{
  const int f = flag < 0 ? -flag : flag;
  int r = 0;
  if (f == 1) {
    int x, y;
    memcpy(&x, a, 4), memcpy(&y, b, 4);
    r = x < y ? -1 : (x > y ? 1 : 0);
  } else if (f == 2) {
    uint64_t x, y;
    memcpy(&x, a, 8), memcpy(&y, b, 8);
    r = x < y ? -1 : (x > y ? 1 : 0);
  } else if (f == 3) {
    float x, y;
    memcpy(&x, a, 4), memcpy(&y, b, 4);
    r = x < y ? -1 : (x > y ? 1 : 0);
  } else if (f == 4) {
    double x, y;
    memcpy(&x, a, 8), memcpy(&y, b, 8);
    r = x < y ? -1 : (x > y ? 1 : 0);
  } else if (f == 5) r = strcmp(a, b);
  else r = strncmp(a, b, alen < blen ? alen : blen);
  return flag < 0 ? -r : r;
}

uint64_t MapReduce::sort_multivalues_impl(int flag, int (*appcompare)(char *, int, char *, int))
{
  if (kmv == NULL) all("Cannot sort_multivalues without KeyMultiValue");
  if (!appcompare && (flag == 0 || flag > 6 || flag < -6)) all("Invalid compare method for sort multivalues");
  start_timer();
  mrgpu_mr *side = make_side_handle(this, memsize, keyalign, valuealign);
  if (!side) all(mrgpu_last_error(NULL));
  KeyValue *out = begin_emit(side, 0);
  struct Val {
    const char *p;
    int len;
  };
  std::vector<Val> vals;
  std::vector<char> pool;
  walk_kmv(gpu, [&](char *k, int kb, char *mv, int nv, int *sizes) {
    vals.clear();
    pool.clear();
    if (mv) {
      for (int i = 0; i < nv; i++) {
        vals.push_back(Val{mv, sizes[i]});
        mv += sizes[i];
      }
    } else {  // a pair bigger than a page: gather its blocks (pointers die with each block, so the bytes are copied)
      int nblock;
      multivalue_blocks(nblock);
      std::vector<std::pair<size_t, int> > where;
      for (int b = 0; b < nblock; b++) {
        char *bm;
        int *bs;
        const int n = multivalue_block(b, &bm, &bs);
        for (int i = 0; i < n; i++) {
          where.push_back(std::make_pair(pool.size(), bs[i]));
          pool.insert(pool.end(), bm, bm + bs[i]);
          pool.push_back(0);
          bm += bs[i];
        }
      }
      for (size_t i = 0; i < where.size(); i++) vals.push_back(Val{&pool[where[i].first], where[i].second});
    }
    std::stable_sort(vals.begin(), vals.end(), [&](const Val &x, const Val &y) {
      return appcompare ? appcompare((char *) x.p, x.len, (char *) y.p, y.len) < 0 : cmp_flag(flag, x.p, x.len, y.p, y.len) < 0;
    });
    for (size_t i = 0; i < vals.size(); i++) out->add(k, kb, (char *) vals[i].p, vals[i].len);
  });
  end_emit(side, out);
  uint64_t n = 0;
  ck(mrgpu_kv_open(gpu, 0), 0);  // drops the KMV
  ck(mrgpu_kv_close(gpu, &n), 0);
  ck(mrgpu_add(gpu, side, &n), 0);
  mrgpu_destroy(side);
  ck(mrgpu_convert(gpu, &n), 0);
  delete kv;
  kv = NULL;
  kmv = &g_kmv_marker;
  stats("Sort_multivalues", 1);
  return n;
}

uint64_t MapReduce::sort_multivalues(int flag) { return sort_multivalues_impl(flag, NULL); }
uint64_t MapReduce::sort_multivalues(int (*appcompare)(char *, int, char *, int)) { return sort_multivalues_impl(0, appcompare); }

// ---------------------------------------------------------------------------------------------- the rest

// gather(n): with one rank every pair already lives on rank 0 (src/mapreduce.cpp:893-905 returns early)
uint64_t MapReduce::gather(int numprocs)
{
  if (kv == NULL) all("Cannot gather without KeyValue");
  if (numprocs < 1 || numprocs > nprocs) all("Invalid processor count for gather");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_gather(gpu, numprocs, &n), 0);
  sync_kv_fields();
  stats("Gather", 0);
  return n;
}

MapReduce *MapReduce::copy()
{
  MapReduce *m = new MapReduce(comm);
  m->mapstyle = mapstyle;
  m->all2all = all2all;
  m->verbosity = verbosity;
  m->timer = timer;
  m->memsize = memsize;
  m->minpage = minpage;
  m->maxpage = maxpage;
  m->keyalign = keyalign;
  m->valuealign = valuealign;
  m->outofcore = outofcore;
  push_settings();
  m->push_settings();
  if (kv || kmv) {  // deep copy of whichever store this object holds (src/mapreduce.cpp:314-338)
    ck(mrgpu_copy(m->gpu, gpu), 0);
    if (kv) {
      m->kv = new KeyValue(m->gpu);
      m->sync_kv_fields();
    } else
      m->kmv = &g_kmv_marker;
  }
  return m;
}

uint64_t MapReduce::broadcast(int root)
{
  if (kv == NULL) all("Cannot broadcast without KeyValue");
  if (root < 0 || root >= nprocs) all("Invalid root for broadcast");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_broadcast(gpu, root, &n), 0);
  sync_kv_fields();
  stats("Broadcast", 0);
  return n;
}

// KV -> KMV on the GPU; the host-side objects follow as in convert()
uint64_t MapReduce::clone()
{
  if (kv == NULL) all("Cannot clone without KeyValue");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_clone(gpu, &n), 0);
  delete kv;
  kv = NULL;
  kmv = &g_kmv_marker;
  stats("Clone", 1);
  return n;
}

uint64_t MapReduce::collapse(char *key, int keybytes)
{
  if (kv == NULL) all("Cannot collapse without KeyValue");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_collapse(gpu, key, keybytes, &n), 0);
  delete kv;
  kv = NULL;
  kmv = &g_kmv_marker;
  stats("Collapse", 1);
  return n;
}

uint64_t MapReduce::scrunch(int numprocs, char *key, int keybytes)
{
  if (kv == NULL) all("Cannot scrunch without KeyValue");
  if (numprocs < 1 || numprocs > nprocs) all("Invalid processor count for gather");
  start_timer();
  uint64_t n = 0;
  ck(mrgpu_scrunch(gpu, numprocs, key, keybytes, &n), 0);
  delete kv;
  kv = NULL;
  kmv = &g_kmv_marker;
  stats("Scrunch", 1);
  return n;
}
// ---- print(): the reference's text form of the pairs (KeyValue::print src/keyvalue.cpp:773-830,
// KeyMultiValue::print src/keymultivalue.cpp:1646-1760), produced on the host from the page walks

// one field in the format selected by flag 0..7 (NULL, int, uint64, float, double, string, 2 ints, 2 uint64)
static void print_field(FILE *fp, int flag, const char *p)
{
  int i2[2];
  uint64_t u2[2];
  float f;
  double d;
  switch (flag) {
    case 0: fputs("NULL", fp); break;
    case 1: memcpy(i2, p, 4); fprintf(fp, "%d", i2[0]); break;
    case 2: memcpy(u2, p, 8); fprintf(fp, "%lu", (unsigned long) u2[0]); break;
    case 3: memcpy(&f, p, 4); fprintf(fp, "%g", f); break;
    case 4: memcpy(&d, p, 8); fprintf(fp, "%g", d); break;
    case 5: fprintf(fp, "%s", p); break;
    case 6: memcpy(i2, p, 8); fprintf(fp, "%d %d", i2[0], i2[1]); break;
    default: memcpy(u2, p, 16); fprintf(fp, "%lu %lu", (unsigned long) u2[0], (unsigned long) u2[1]); break;
  }
}

void MapReduce::print_pairs(FILE *fp, int nstride, int kflag, int vflag)
{
  int istride = 0;
  if (kv) {
    walk_kv(gpu, [&](char *k, int kb, char *v, int vb) {
      if (++istride != nstride) return;
      istride = 0;
      fprintf(fp, "KV pair: proc %d, sizes %d %d, key ", me, kb, vb);
      print_field(fp, kflag, k);
      fputs(", value ", fp);
      print_field(fp, vflag, v);
      fputc('\n', fp);
    });
  }
  if (kmv) {
    walk_kmv(gpu, [&](char *k, int kb, char *mv, int nv, int *sizes) {
      if (mv == NULL && nv == 0 && (void *) sizes == (void *) this) {  // a multi-block pair: never strided, never listed
        fprintf(fp, "KMV pair: proc %d, nvalues %lu, size %d, key ", me, (unsigned long) kmv_nvalue_total, kb);
        print_field(fp, kflag, k);
        fputs(", too many values to print\n", fp);
        return;
      }
      if (++istride != nstride) return;
      istride = 0;
      int mvb = 0;
      for (int j = 0; j < nv; j++) mvb += sizes[j];
      fprintf(fp, "KMV pair: proc %d, nvalues %d, sizes %d %d, key ", me, nv, kb, mvb);
      print_field(fp, kflag, k);
      fputs(", values ", fp);
      for (int j = 0; j < nv; j++) {
        print_field(fp, vflag, mv);
        fputc(' ', fp);
        mv += sizes[j];
      }
      fputc('\n', fp);
    });
  }
}

// print to the screen (src/mapreduce.cpp:1671-1711): proc >= 0 only that rank prints, proc < 0 every rank in turn
void MapReduce::print(int proc, int nstride, int kflag, int vflag)
{
  if (kv == NULL && kmv == NULL) all("Cannot print without KeyValue or KeyMultiValue");
  if (kflag < 0 || vflag < 0 || kflag > 7 || vflag > 7) all("Invalid print args");
  if (proc == me) print_pairs(stdout, nstride, kflag, vflag);
  if (proc >= 0) return;
  for (int r = 0; r < nprocs; r++) {
    MPI_Barrier(comm);
    if (r == me) {
      print_pairs(stdout, nstride, kflag, vflag);
      fflush(stdout);
    }
  }
  MPI_Barrier(comm);
}

// print to file(s) (src/mapreduce.cpp:1719-1762): one rank -> `file`; all ranks -> `file` appended in rank order
// (fflag 0) or `file.<rank>` each (fflag 1)
void MapReduce::print(char *file, int fflag, int proc, int nstride, int kflag, int vflag)
{
  if (kv == NULL && kmv == NULL) all("Cannot print without KeyValue or KeyMultiValue");
  if (kflag < 0 || vflag < 0 || kflag > 7 || vflag > 7) all("Invalid print args");
  auto to_file = [&](const std::string &name, const char *mode) {
    FILE *fp = fopen(name.c_str(), mode);
    if (fp == NULL) one("Could not open print file");
    print_pairs(fp, nstride, kflag, vflag);
    fclose(fp);
  };
  if (proc == me) to_file(file, "w");
  if (proc >= 0) return;
  if (fflag == 1) {
    to_file(std::string(file) + "." + std::to_string(me), "w");
    return;
  }
  for (int r = 0; r < nprocs; r++) {
    MPI_Barrier(comm);
    if (r == me) to_file(file, "a");
  }
  MPI_Barrier(comm);
}

}  // namespace MAPREDUCE_NS

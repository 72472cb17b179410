// One caller, two libraries, several ranks.  Written against the MR-MPI class API only.  oracle/Makefile links it
// with the unmodified reference library (serial mpistubs build: 1 rank); __graft_entry__.build() links the same
// source with the GPU library's mirror classes, and tests/test_cpp_dropin.py runs THAT build as 2 processes
// (RANK / WORLD_SIZE / LOCAL_RANK in the environment, one GPU each, no MPI).  Everything printed (by rank 0, after
// gather(1)) must be identical: results do not depend on the number of ranks, and with mapstyle 0 the values of a
// key arrive in task order either way (rank order = task order).
#include "mpi.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "mapreduce.h"
#include "keyvalue.h"
#if __has_include("mrgpu_functors.h")
#include "mrgpu_functors.h"
#define HAVE_DEVICE_FUNCTORS 1
#endif

using namespace MAPREDUCE_NS;

// task-deterministic generator: what a task emits does not depend on which rank runs it
static void gen(int itask, KeyValue *kv, void *ptr)
{
  int n = *(int *) ptr;
  uint64_t st = 0x9E3779B97F4A7C15ULL * (uint64_t) (itask + 1);
  for (int i = 0; i < n; i++) {
    st = st * 6364136223846793005ULL + 1442695040888963407ULL;
    int w = (int) ((st >> 33) % 211);
    char word[32];
    int len = sprintf(word, "k%d_%d", w % 5, w);
    int val = itask * 100000 + i;
    kv->add(word, len + 1, (char *) &val, sizeof(int));
  }
}

// (key) -> (key, {nvalues, sum, first value, last value}): sensitive to the arrival order of the values
static void summarize(char *key, int keybytes, char *multivalue, int nvalues, int *valuebytes, KeyValue *kv, void *)
{
  int64_t out[4] = {nvalues, 0, 0, 0};
  for (int i = 0; i < nvalues; i++) {
    int v;
    memcpy(&v, multivalue, sizeof(int));
    out[1] += v;
    if (i == 0) out[2] = v;
    out[3] = v;
    multivalue += valuebytes[i];
  }
  kv->add(key, keybytes, (char *) out, sizeof(out));
}

static void countvalues(char *key, int keybytes, char *, int nvalues, int *, KeyValue *kv, void *)
{
  kv->add(key, keybytes, (char *) &nvalues, sizeof(int));
}

static int firstchar_hash(char *key, int keybytes) { return keybytes > 1 ? (unsigned char) key[1] * 7 + keybytes : 0; }

static void show4(char *key, int, char *value, int, void *)
{
  int64_t v[4];
  memcpy(v, value, sizeof(v));
  printf("%s %lld %lld %lld %lld\n", key, (long long) v[0], (long long) v[1], (long long) v[2], (long long) v[3]);
}

static void show1(char *key, int, char *value, int, void *)
{
  int v;
  memcpy(&v, value, sizeof(int));
  printf("%s %d\n", key, v);
}

static void show_int(char *key, int, char *value, int, void *)
{
  int k, v;
  memcpy(&k, key, 4);
  memcpy(&v, value, 4);
  printf("%d %d\n", k, v);
}

int main(int narg, char **args)
{
  MPI_Init(&narg, &args);
  int me;
  MPI_Comm_rank(MPI_COMM_WORLD, &me);
  int n = 3000;

  // 1. default partition: map -> collate -> reduce(host callback) -> gather(1) -> sort_keys -> scan
  MapReduce *mr = new MapReduce(MPI_COMM_WORLD);
  uint64_t npairs = mr->map(8, gen, &n);
  uint64_t nuniq = mr->collate(NULL);
  uint64_t nred = mr->reduce(summarize, NULL);
  mr->gather(1);
  mr->sort_keys(5);
  if (me == 0) printf("== default hash: pairs %lu unique %lu reduced %lu\n", (unsigned long) npairs, (unsigned long) nuniq, (unsigned long) nred);
  mr->scan(show4, NULL);
  fflush(stdout);

  // 2. user hash function: aggregate(myhash) -> convert -> reduce -> gather -> sort -> scan
  MapReduce *mu = new MapReduce(MPI_COMM_WORLD);
  mu->map(8, gen, &n);
  uint64_t nagg = mu->aggregate(firstchar_hash);
  uint64_t nconv = mu->convert();
  mu->reduce(summarize, NULL);
  mu->gather(1);
  mu->sort_keys(5);
  if (me == 0) printf("== user hash: pairs %lu unique %lu\n", (unsigned long) nagg, (unsigned long) nconv);
  mu->scan(show4, NULL);
  fflush(stdout);

  // 3. stock reduce callback (a device functor on the GPU library, the same host function on the reference)
  MapReduce *mc = new MapReduce(MPI_COMM_WORLD);
  mc->map(8, gen, &n);
  mc->collate(NULL);
#ifdef HAVE_DEVICE_FUNCTORS
  mc->reduce(gpu::count, NULL);
#else
  mc->reduce(countvalues, NULL);
#endif
  mc->gather(1);
  mc->sort_keys(5);
  if (me == 0) printf("== count\n");
  mc->scan(show1, NULL);
  fflush(stdout);

  // 4. IntCount: (uint32 key, int 1) per key -> collate -> count -> gather -> sort_keys(1)
  std::vector<std::vector<uint32_t> > shards(4);
  std::vector<const uint32_t *> ptrs(4);
  std::vector<uint64_t> lens(4);
  for (int t = 0; t < 4; t++) {
    uint64_t st = 777 + t;
    for (int i = 0; i < 90000; i++) {
      st = st * 6364136223846793005ULL + 1442695040888963407ULL;
      uint32_t r = (uint32_t) (st >> 40);
      shards[t].push_back((r % 1000) * (r % 7 == 0 ? 1 : 0) + (r % 97));   // skewed
    }
    ptrs[t] = shards[t].data();
    lens[t] = shards[t].size();
  }
  MapReduce *mi = new MapReduce(MPI_COMM_WORLD);
#ifdef HAVE_DEVICE_FUNCTORS
  gpu::IntCountInput in = {ptrs.data(), lens.data()};
  uint64_t nic = mi->map(4, gpu::intcount, &in);
#else
  struct Emit {
    static void run(int itask, KeyValue *kv, void *p)
    {
      std::vector<std::vector<uint32_t> > *sh = (std::vector<std::vector<uint32_t> > *) p;
      int one = 1;
      for (size_t i = 0; i < (*sh)[itask].size(); i++) kv->add((char *) &(*sh)[itask][i], 4, (char *) &one, 4);
    }
  };
  uint64_t nic = mi->map(4, Emit::run, &shards);
#endif
  uint64_t nicu = mi->collate(NULL);
#ifdef HAVE_DEVICE_FUNCTORS
  mi->reduce(gpu::count, NULL);
#else
  mi->reduce(countvalues, NULL);
#endif
  mi->gather(1);
  mi->sort_keys(1);
  if (me == 0) printf("== intcount: pairs %lu unique %lu\n", (unsigned long) nic, (unsigned long) nicu);
  mi->scan(show_int, NULL);
  fflush(stdout);

  delete mi;
  delete mc;
  delete mu;
  delete mr;
  MPI_Finalize();
  return 0;
}

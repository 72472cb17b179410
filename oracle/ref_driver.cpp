// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// Driver for the UNMODIFIED reference library (MR-MPI as found under /root/reference/src, built
// serially against /root/reference/mpistubs by oracle/Makefile into oracle/_ref/).  It runs the
// reference's own CPU implementation of map -> aggregate -> convert -> reduce -> sort_keys on files
// in a simple interchange format so that tests/golden/make_golden.py can pin the C restatement in
// oracle/mr_oracle.c and the CUDA path against the real thing, and so that bench.py can time the
// reference CPU path ("cpu_baseline.kind" = "reference", `bench.py --impl reference`).
//
// The callbacks whose bodies define parity (wordfreq fileread/sum, the InvertedIndex href FSM) are
// NOT restated here: oracle/Makefile extracts them from the reference sources at build time into
// oracle/_ref/.build/*.inc and this file #includes them, so the driver executes the reference text
// verbatim.  OINK callbacks (rmat_generate, count, cull, edge_to_vertices, edge_to_vertex) are
// compiled from /root/reference/oink/*.cpp as they lie and linked in.
//
// Interchange formats (little-endian, unpadded):
//   kvfile : u64 n ; n x { i32 keybytes ; i32 valuebytes ; key ; value }
//   kmvfile: u64 n ; n x { i32 keybytes ; u64 nvalue ; key ; nvalue x { i32 valuebytes ; value } }
//
// Sub-commands are listed in usage().

#include "mpi.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <string>
#include <vector>
#include "mapreduce.h"
#include "keyvalue.h"
#include "keymultivalue.h"
#include "hash.h"
#include "typedefs.h"
#include "map_rmat_generate.h"

using namespace MAPREDUCE_NS;

// OINK callbacks compiled from the reference as-is (oink/map_rmat_generate.cpp etc.)
void rmat_generate(int, KeyValue *, void *);
void count(char *, int, char *, int, int *, KeyValue *, void *);
void cull(char *, int, char *, int, int *, KeyValue *, void *);
void edge_to_vertices(uint64_t, char *, int, char *, int, KeyValue *, void *);
void edge_to_vertex(uint64_t, char *, int, char *, int, KeyValue *, void *);

// reference callbacks extracted verbatim at build time
#include "wordfreq_callbacks.inc"   // fileread(), sum()   examples/wordfreq.cpp:104-141

static double now()
{
  struct timeval tv;
  gettimeofday(&tv, NULL);
  return tv.tv_sec + 1e-6 * tv.tv_usec;
}

static void die(const char *msg)
{
  fprintf(stderr, "refdrv: %s\n", msg);
  exit(2);
}

// ---------------------------------------------------------------- kvfile IO

struct Blob {
  std::vector<char> data;
};

static void read_file(const char *path, std::vector<char> &out)
{
  FILE *fp = fopen(path, "rb");
  if (!fp) die("cannot open input file");
  fseek(fp, 0, SEEK_END);
  long n = ftell(fp);
  fseek(fp, 0, SEEK_SET);
  out.resize(n);
  if (n && fread(&out[0], 1, n, fp) != (size_t) n) die("short read");
  fclose(fp);
}

static void load_kvfile(int, KeyValue *kv, void *ptr)
{
  std::vector<char> *buf = (std::vector<char> *) ptr;
  const char *p = &(*buf)[0];
  uint64_t n;
  memcpy(&n, p, 8);
  p += 8;
  for (uint64_t i = 0; i < n; i++) {
    int kb, vb;
    memcpy(&kb, p, 4);
    memcpy(&vb, p + 4, 4);
    p += 8;
    kv->add((char *) p, kb, (char *) (p + kb), vb);
    p += kb + vb;
  }
}

struct DumpKV {
  FILE *fp;
  uint64_t n;
};

static void dump_kv_cb(char *key, int kb, char *val, int vb, void *ptr)
{
  DumpKV *d = (DumpKV *) ptr;
  fwrite(&kb, 4, 1, d->fp);
  fwrite(&vb, 4, 1, d->fp);
  fwrite(key, 1, kb, d->fp);
  fwrite(val, 1, vb, d->fp);
  d->n++;
}

static void dump_kv(MapReduce *mr, const char *path)
{
  DumpKV d;
  d.fp = fopen(path, "wb");
  if (!d.fp) die("cannot open output");
  d.n = 0;
  uint64_t zero = 0;
  fwrite(&zero, 8, 1, d.fp);
  mr->scan(dump_kv_cb, &d);
  fseek(d.fp, 0, SEEK_SET);
  fwrite(&d.n, 8, 1, d.fp);
  fclose(d.fp);
}

// KMV dump is done from a reduce callback so that multi-block ("extended") KMV pairs are walked with
// the reference's own block protocol (src/mapreduce.cpp:1874-1925).
static void dump_kmv_cb(char *key, int kb, char *mv, int nv, int *vbytes, KeyValue *, void *ptr)
{
  DumpKV *d = (DumpKV *) ptr;
  fwrite(&kb, 4, 1, d->fp);
  if (mv) {
    uint64_t n64 = nv;
    fwrite(&n64, 8, 1, d->fp);
    fwrite(key, 1, kb, d->fp);
    for (int i = 0; i < nv; i++) {
      fwrite(&vbytes[i], 4, 1, d->fp);
      fwrite(mv, 1, vbytes[i], d->fp);
      mv += vbytes[i];
    }
  } else {
    MapReduce *mr = (MapReduce *) vbytes;
    int nblocks;
    uint64_t total = mr->multivalue_blocks(nblocks);
    fwrite(&total, 8, 1, d->fp);
    fwrite(key, 1, kb, d->fp);
    for (int ib = 0; ib < nblocks; ib++) {
      int n = mr->multivalue_block(ib, &mv, &vbytes);
      for (int i = 0; i < n; i++) {
        fwrite(&vbytes[i], 4, 1, d->fp);
        fwrite(mv, 1, vbytes[i], d->fp);
        mv += vbytes[i];
      }
    }
  }
  d->n++;
}

static void dump_kmv(MapReduce *mr, const char *path)
{
  DumpKV d;
  d.fp = fopen(path, "wb");
  if (!d.fp) die("cannot open output");
  d.n = 0;
  uint64_t zero = 0;
  fwrite(&zero, 8, 1, d.fp);
  mr->reduce(dump_kmv_cb, &d);
  fseek(d.fp, 0, SEEK_SET);
  fwrite(&d.n, 8, 1, d.fp);
  fclose(d.fp);
}

// raw page images: u64 npage ; npage x { i32 nkey ; u64 alignsize ; bytes }
static void dump_kv_pages(MapReduce *mr, const char *path)
{
  FILE *fp = fopen(path, "wb");
  KeyValue *kv = mr->kv;
  if (!kv) die("pagekv: no KeyValue");
  kv->allocate();   // as MapReduce::scan does before walking pages, src/mapreduce.cpp:1945
  char *page;
  int npage = kv->request_info(&page);
  uint64_t np = npage;
  fwrite(&np, 8, 1, fp);
  for (int i = 0; i < npage; i++) {
    uint64_t ks, vs, as;
    int nkey = kv->request_page(i, ks, vs, as);
    fwrite(&nkey, 4, 1, fp);
    fwrite(&as, 8, 1, fp);
    fwrite(page, 1, as, fp);
  }
  kv->deallocate(0);
  fclose(fp);
}

static void dump_kmv_pages(MapReduce *mr, const char *path)
{
  FILE *fp = fopen(path, "wb");
  KeyMultiValue *kmv = mr->kmv;
  if (!kmv) die("pagekmv: no KeyMultiValue");
  kmv->allocate();
  char *page;
  int npage = kmv->request_info(&page);
  uint64_t np = npage;
  fwrite(&np, 8, 1, fp);
  for (int i = 0; i < npage; i++) {
    uint64_t ks, vs, as;
    int nkey = kmv->request_page(i, 0, ks, vs, as);
    fwrite(&nkey, 4, 1, fp);
    fwrite(&as, 8, 1, fp);
    fwrite(page, 1, as, fp);
  }
  kmv->deallocate(0);
  fclose(fp);
}

// ---------------------------------------------------------------- callbacks that are ours

struct U32File {
  const uint32_t *keys;
  uint64_t n;
};

// IntCount emit: same statement as cpu/IntCount.cpp:177-180 (key = 4 input bytes, value = int 1).
static void intcount_emit(int, KeyValue *kv, void *ptr)
{
  U32File *f = (U32File *) ptr;
  int one = 1;
  for (uint64_t i = 0; i < f->n; i++) kv->add((char *) &f->keys[i], 4, (char *) &one, 4);
}

// InvertedIndex map: file handling is ours (the reference hard-codes /data/bzhaoad/...); the scanner
// between the two markers is cpu/InvertedIndex.cpp:196-270 included verbatim.
#define START 0x00
#define IN_TAG 0x01
#define IN_ATAG 0x02
#define FOUND_HREF 0x03
#define START_LINK 0x04

static void invindex_file(int, char *fname, KeyValue *kv, void *)
{
  struct stat stbuf;
  if (stat(fname, &stbuf) < 0) die("invindex: cannot stat file");
  int filesize = stbuf.st_size;
  FILE *fp = fopen(fname, "r");
  char *text = new char[filesize + 1];
  int nchar = fread(text, 1, filesize, fp);
  text[nchar] = '\0';
  fclose(fp);

  char filename[256];
  const char *slash = strrchr(fname, '/');
  const char *base = slash ? slash + 1 : fname;
  int namelen = strlen(base);
  memcpy(filename, base, namelen + 1);

#include "invindex_fsm.inc"   // cpu/InvertedIndex.cpp:196-270

  delete[] text;
}

// ---------------------------------------------------------------- rmat with virtual ranks

struct VRanks {
  RMAT_struct rmat;
  int nprocs;
  uint64_t nremain;
  unsigned short (*state)[3];
};

// One reference rank = one drand48 stream seeded srand48(seed+me) (oink/rmat.cpp:94) that keeps
// running across iterations.  The serial stub has a single rank, so each virtual rank's 48-bit state
// is swapped in around the unmodified rmat_generate().
static void rmat_vrank(int itask, KeyValue *kv, void *ptr)
{
  VRanks *v = (VRanks *) ptr;
  v->rmat.ngenerate = v->nremain / v->nprocs;                      // oink/rmat.cpp:55
  if ((uint64_t) itask < v->nremain % v->nprocs) v->rmat.ngenerate++;   // oink/rmat.cpp:56
  seed48(v->state[itask]);
  rmat_generate(itask, kv, &v->rmat);
  unsigned short dummy[3] = {0, 0, 0};
  unsigned short *cur = seed48(dummy);
  memcpy(v->state[itask], cur, 6);
}

// ---------------------------------------------------------------- op interpreter

static void timing(const char *op, double t, uint64_t n)
{
  printf("{\"op\": \"%s\", \"secs\": %.6f, \"n\": %llu}\n", op, t, (unsigned long long) n);
  fflush(stdout);
}

static void run_ops(MapReduce *mr, int narg, char **arg, const char *out)
{
  for (int i = 0; i < narg; i++) {
    const char *op = arg[i];
    double t0 = now();
    uint64_t n = 0;
    if (!strcmp(op, "aggregate")) n = mr->aggregate(NULL);
    else if (!strcmp(op, "convert")) n = mr->convert();
    else if (!strcmp(op, "collate")) n = mr->collate(NULL);
    else if (!strcmp(op, "reduce_count")) n = mr->reduce(count, NULL);
    else if (!strcmp(op, "reduce_sum")) n = mr->reduce(sum, NULL);
    else if (!strcmp(op, "reduce_cull")) n = mr->reduce(cull, NULL);
    else if (!strncmp(op, "sort_keys:", 10)) n = mr->sort_keys(atoi(op + 10));
    else if (!strncmp(op, "sort_values:", 12)) n = mr->sort_values(atoi(op + 12));
    else if (!strcmp(op, "clone")) n = mr->clone();
    else if (!strncmp(op, "collapse:", 9)) n = mr->collapse((char *) op + 9, (int) strlen(op + 9) + 1);  // key = text + NUL
    else if (!strcmp(op, "compress_count")) n = mr->compress(count, NULL);
    else if (!strcmp(op, "compress_sum")) n = mr->compress(sum, NULL);
    else if (!strcmp(op, "compress_cull")) n = mr->compress(cull, NULL);
    else if (!strcmp(op, "map_e2v")) n = mr->map(mr, edge_to_vertices, NULL);
    else if (!strcmp(op, "map_e2v1")) n = mr->map(mr, edge_to_vertex, NULL);
    else if (!strcmp(op, "dumpkv")) dump_kv(mr, out);
    else if (!strcmp(op, "dumpkmv")) dump_kmv(mr, out);
    else if (!strcmp(op, "pagekv")) dump_kv_pages(mr, out);
    else if (!strcmp(op, "pagekmv")) dump_kmv_pages(mr, out);
    else die("unknown op");
    timing(op, now() - t0, n);
  }
}

static void usage()
{
  fprintf(stderr,
          "refdrv hash                                  (stdin: '<initval> <hexkey>' per line)\n"
          "refdrv run <in.kv> <out> <memsize> [align] ops...   ops: aggregate convert collate reduce_count\n"
          "        reduce_sum reduce_cull sort_keys:<flag> sort_values:<flag> compress_count compress_sum compress_cull\n"
          "        clone collapse:<key text> map_e2v map_e2v1 dumpkv dumpkmv pagekv pagekmv\n"
          "refdrv intcount <in.u32> <out.kv> <memsize>  map emit, aggregate, convert, reduce count, sort_keys 1\n"
          "refdrv wordfreq <out> <memsize> <mapout|final> files...\n"
          "refdrv invindex <out> <memsize> <mapout|kmv> files...\n"
          "refdrv rmat <N> <ef> <a> <b> <c> <d> <frac> <seed> <P> <memsize> <edges.kv> <degree.kv> <dup>\n"
          "memsize: MB per page if > 0, bytes if < 0 (src/mapreduce.cpp:3351).  align: key/value alignment\n");
  exit(2);
}

static MapReduce *new_mr(int memsize)
{
  MapReduce *mr = new MapReduce(MPI_COMM_WORLD);
  mr->memsize = memsize;
  mr->verbosity = 0;
  mr->timer = 0;
  mr->set_fpath(getenv("REFDRV_FPATH") ? getenv("REFDRV_FPATH") : "/tmp");
  return mr;
}

int main(int narg, char **arg)
{
  MPI_Init(&narg, &arg);
  if (narg < 2) usage();
  std::string cmd = arg[1];

  if (cmd == "hash") {
    char line[8192];
    while (fgets(line, sizeof(line), stdin)) {
      unsigned initval;
      char hex[8192];
      hex[0] = 0;
      int nf = sscanf(line, "%u %s", &initval, hex);
      if (nf < 1) continue;
      int len = strlen(hex) / 2;
      std::vector<char> key(len + 8);
      for (int i = 0; i < len; i++) {
        unsigned b;
        sscanf(hex + 2 * i, "%2x", &b);
        key[i] = (char) b;
      }
      printf("%08x\n", hashlittle(&key[0], len, initval));
    }
  } else if (cmd == "run") {
    if (narg < 6) usage();
    std::vector<char> buf;
    read_file(arg[2], buf);
    MapReduce *mr = new_mr(atoi(arg[4]));
    int first = 5;
    if (arg[5][0] >= '0' && arg[5][0] <= '9') {
      mr->keyalign = mr->valuealign = atoi(arg[5]);
      first = 6;
    }
    double t0 = now();
    uint64_t n = mr->map(1, load_kvfile, &buf);
    timing("load", now() - t0, n);
    run_ops(mr, narg - first, arg + first, arg[3]);
    delete mr;
  } else if (cmd == "intcount") {
    if (narg < 5) usage();
    std::vector<char> buf;
    read_file(arg[2], buf);
    U32File f;
    f.keys = (const uint32_t *) &buf[0];
    f.n = buf.size() / 4;
    MapReduce *mr = new_mr(atoi(arg[4]));
    double t0 = now();
    uint64_t n = mr->map(1, intcount_emit, &f);
    timing("map", now() - t0, n);
    const char *ops[] = {"aggregate", "convert", "reduce_count", "sort_keys:1", "dumpkv"};
    run_ops(mr, 5, (char **) ops, arg[3]);
    delete mr;
  } else if (cmd == "wordfreq") {
    if (narg < 6) usage();
    MapReduce *mr = new_mr(atoi(arg[3]));
    double t0 = now();
    uint64_t n = mr->map(narg - 5, &arg[5], 0, 1, 0, fileread, NULL);   // examples/wordfreq.cpp:64
    timing("map", now() - t0, n);
    if (!strcmp(arg[4], "mapout")) {
      const char *ops[] = {"dumpkv"};
      run_ops(mr, 1, (char **) ops, arg[2]);
    } else {
      const char *ops[] = {"collate", "reduce_sum", "dumpkv"};         // examples/wordfreq.cpp:65-66
      run_ops(mr, 3, (char **) ops, arg[2]);
    }
    delete mr;
  } else if (cmd == "invindex") {
    if (narg < 6) usage();
    MapReduce *mr = new_mr(atoi(arg[3]));
    double t0 = now();
    uint64_t n = mr->map(narg - 5, &arg[5], 0, 0, 0, invindex_file, NULL);
    timing("map", now() - t0, n);
    if (!strcmp(arg[4], "mapout")) {
      const char *ops[] = {"dumpkv"};
      run_ops(mr, 1, (char **) ops, arg[2]);
    } else {
      const char *ops[] = {"aggregate", "convert", "dumpkmv"};         // cpu/InvertedIndex.cpp:80-91
      run_ops(mr, 3, (char **) ops, arg[2]);
    }
    delete mr;
  } else if (cmd == "rmat") {
    if (narg < 15) usage();
    VRanks v;
    v.rmat.nlevels = atoi(arg[2]);
    v.rmat.nnonzero = atoi(arg[3]);
    v.rmat.a = atof(arg[4]);
    v.rmat.b = atof(arg[5]);
    v.rmat.c = atof(arg[6]);
    v.rmat.d = atof(arg[7]);
    v.rmat.fraction = atof(arg[8]);
    int seed = atoi(arg[9]);
    v.nprocs = atoi(arg[10]);
    int memsize = atoi(arg[11]);
    if (v.rmat.a + v.rmat.b + v.rmat.c + v.rmat.d != 1.0) die("RMAT a,b,c,d must sum to 1");
    v.rmat.order = 1 << v.rmat.nlevels;                                // oink/rmat.cpp:95
    v.state = new unsigned short[v.nprocs][3];
    for (int r = 0; r < v.nprocs; r++) {                               // srand48(seed+me) oink/rmat.cpp:94
      srand48(seed + r);
      unsigned short dummy[3] = {0, 0, 0};
      unsigned short *cur = seed48(dummy);
      memcpy(v.state[r], cur, 6);
    }
    MapReduce *mr = new_mr(memsize);
    int niterate = 0;
    uint64_t ntotal = v.rmat.order * v.rmat.nnonzero;
    v.nremain = ntotal;
    while (v.nremain) {                                                // oink/rmat.cpp:53-61
      niterate++;
      double t0 = now();
      uint64_t n = mr->map(v.nprocs, rmat_vrank, &v, 1);
      timing("generate", now() - t0, n);
      t0 = now();
      uint64_t nunique = mr->collate(NULL);
      timing("collate", now() - t0, nunique);
      t0 = now();
      mr->reduce(cull, &v.rmat);
      timing("cull", now() - t0, nunique);
      v.nremain = ntotal - nunique;
    }
    printf("{\"op\": \"rmat_done\", \"iterations\": %d, \"edges\": %llu}\n", niterate,
           (unsigned long long) ntotal);
    dump_kv(mr, arg[12]);
    // Degree::run oink/degree.cpp:43-50
    MapReduce *mrv = new_mr(memsize);
    int dup = atoi(arg[14]);
    double t0 = now();
    uint64_t n;
    if (dup == 1) n = mrv->map(mr, edge_to_vertex, NULL);
    else n = mrv->map(mr, edge_to_vertices, NULL);
    timing("degree_map", now() - t0, n);
    t0 = now();
    n = mrv->collate(NULL);
    timing("degree_collate", now() - t0, n);
    t0 = now();
    n = mrv->reduce(count, NULL);
    timing("degree_reduce", now() - t0, n);
    dump_kv(mrv, arg[13]);
    delete mrv;
    delete mr;
  } else usage();

  MPI_Finalize();
  return 0;
}

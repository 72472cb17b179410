/* MapReduce: source-compatible mirror of the reference's public class (src/mapreduce.h:22-131) over the
 * C-ABI CUDA library (include/mrgpu.h).  Same namespace, public settings, method names, signatures, return
 * values (global pair counts) and error behaviour ("ERROR: msg" + exit(1), src/error.cpp:33-67), so callers
 * written for MR-MPI compile unchanged.  The public declarations are the reference's interface, restated (that is the
 * source-compatibility contract); everything behind them is new.
 *
 * Where the work runs:
 *   aggregate / convert / collate / gather / broadcast / clone / collapse / scrunch / sort_keys(flag) /
 *   sort_values(flag) / add / kv_stats / kmv_stats
 *                                                                                    always on the GPU
 *   map / reduce / compress / scan with one of the STOCK callbacks declared in
 *   mrgpu_functors.h (count, sum, cull, edge_to_vertices, ...)                       on the GPU (device functor)
 *   map / reduce / compress / scan / sort_*(compare fn) with any other function ptr  the callback runs on the
 *     host, one page at a time, exactly as in the reference; pages are staged D2H, emitted pairs H2D
 *
 * sort_multivalues and the two file-chunk map() variants are host code over staged pages / files, like any other
 * host callback. */
#ifndef MAP_REDUCE_H
#define MAP_REDUCE_H

#include "mpi.h"
#include "stdint.h"
#include <stdio.h>

struct mrgpu_mr;

namespace MAPREDUCE_NS {

class KeyValue;
class KeyMultiValue;

class MapReduce {
 public:
  int mapstyle;      // 0 = chunks, 1 = strided, 2 = master/slave
  int all2all;       // 0 = irregular comm, 1 = use MPI_Alltoallv()
  int verbosity;     // 0 = none, 1 = totals, 2 = proc histograms
  int timer;         // 0 = none, 1 = summary, 2 = proc histograms
  int memsize;       // # of Mbytes per page
  int minpage;       // # of pages that will be pre-allocated per proc >= 0
  int maxpage;       // max # of pages that can be allocated per proc, 0 = inf
  int freepage;      // 1 to free unused pages after every operation, 0 if keep
  int outofcore;     // 1 to force data out-of-core, 0 = only if exceeds 1 pg
  int zeropage;      // 1 to init allocated pages to 0, 0 if don't bother
  int keyalign;      // align keys to this byte count
  int valuealign;    // align values to this byte count
  char *fpath;       // prefix path added to intermediate out-of-core files
  int mapfilecount;  // number of files processed by map file variants

  class KeyValue *kv;        // single KV stored by MR
  class KeyMultiValue *kmv;  // single KMV stored by MR (opaque marker: non-NULL while a KMV exists)

  static MapReduce *mrptr;
  static int instances_now;
  static int instances_ever;
  static int mpi_finalize_flag;
  static uint64_t msize, msizemax;
  static uint64_t rsize, wsize;
  static uint64_t cssize, crsize;
  static double commtime;

  MapReduce(MPI_Comm);
  MapReduce();
  MapReduce(double);
  ~MapReduce();

  MapReduce *copy();

  uint64_t add(MapReduce *);
  uint64_t aggregate(int (*)(char *, int));
  uint64_t broadcast(int);
  uint64_t clone();
  uint64_t close();
  uint64_t collapse(char *, int);
  uint64_t collate(int (*)(char *, int));
  uint64_t compress(void (*)(char *, int, char *, int, int *, class KeyValue *, void *), void *);
  uint64_t convert();
  uint64_t gather(int);

  uint64_t map(int, void (*)(int, class KeyValue *, void *), void *, int addflag = 0);
  uint64_t map(int, char **, int, int, int, void (*)(int, char *, class KeyValue *, void *), void *, int addflag = 0);
  uint64_t map(int, int, char **, int, int, char, int, void (*)(int, char *, int, class KeyValue *, void *), void *,
               int addflag = 0);
  uint64_t map(int, int, char **, int, int, char *, int, void (*)(int, char *, int, class KeyValue *, void *), void *,
               int addflag = 0);
  uint64_t map(MapReduce *, void (*)(uint64_t, char *, int, char *, int, class KeyValue *, void *), void *, int addflag = 0);

  void open(int addflag = 0);
  void print(int, int, int, int);
  void print(char *, int, int, int, int, int);
  uint64_t reduce(void (*)(char *, int, char *, int, int *, class KeyValue *, void *), void *);
  uint64_t scan(void (*)(char *, int, char *, int, void *), void *);
  uint64_t scan(void (*)(char *, int, char *, int, int *, void *), void *);
  uint64_t scrunch(int, char *, int);

  uint64_t multivalue_blocks(int &);
  void multivalue_block_select(int);
  int multivalue_block(int, char **, int **);

  uint64_t sort_keys(int);
  uint64_t sort_keys(int (*)(char *, int, char *, int));
  uint64_t sort_values(int);
  uint64_t sort_values(int (*)(char *, int, char *, int));
  uint64_t sort_multivalues(int);
  uint64_t sort_multivalues(int (*)(char *, int, char *, int));

  uint64_t kv_stats(int);
  uint64_t kmv_stats(int);
  void cummulative_stats(int, int);

  void set_fpath(const char *);

  MPI_Comm communicator() { return comm; }
  int num_procs() { return nprocs; }
  int my_proc() { return me; }

  // extension: the C-ABI handle, for callers that mix in mrgpu_* calls (device functors of their own)
  mrgpu_mr *gpu_handle() { return gpu; }

  // extension: an application's OWN callback can be tied to a device functor by address, without touching the call
  // sites: whenever map() / reduce() / compress() is then called with that function pointer, the device functor runs
  // instead of the host function (cpp/oink_gpu_register.cpp ties OINK's rmat_generate, cull, count, edge_to_vertices
  // ... this way, so the unmodified OINK commands run their hot loops on the GPU).
  //   device_reduce / device_map_mr: the MRGPU_REDUCE_* / MRGPU_MAP_* functor that has the same effect
  //   device_map_task: a function that performs the whole map() call (it typically converts `ptr` and calls
  //                    map(nmap, gpu::<stock functor>, ...)) and returns the global pair count
  typedef uint64_t (*DeviceMapTask)(MapReduce *mr, int nmap, void *ptr, int addflag);
  static void device_reduce(void (*host)(char *, int, char *, int, int *, class KeyValue *, void *), int functor);
  static void device_map_mr(void (*host)(uint64_t, char *, int, char *, int, class KeyValue *, void *), int functor);
  static void device_map_task(void (*host)(int, class KeyValue *, void *), DeviceMapTask dev);

 private:
  MPI_Comm comm;
  int me, nprocs;
  mrgpu_mr *gpu;
  double time_start;
  int settings_pushed;

  // multi-block KMV state while a host reduce callback runs (src/mapreduce.h:189-194)
  int lib_blocks;  // the library is walking the KMV (scan): multivalue_*() forward to it
  int kmv_block_valid, kmv_nblock;
  uint64_t kmv_nvalue_total;
  const char *blk_sizes_base, *blk_values_base;  // inside the staged extended pair
  uint64_t blk_capacity;
  char *blk_buf;

  void init(MPI_Comm);
  void push_settings();
  void all(const char *);   // Error::all
  void one(const char *);   // Error::one
  uint64_t ck(int rc, uint64_t n);
  void start_timer();
  void stats(const char *, int which);
  void sync_kv_fields();
  KeyValue *begin_emit(mrgpu_mr *target, int addflag);
  uint64_t end_emit(mrgpu_mr *target, KeyValue *k);
  template <typename F> void walk_kv(mrgpu_mr *h, F f);
  template <typename F> void walk_kmv(mrgpu_mr *h, F f);
  void print_pairs(FILE *fp, int nstride, int kflag, int vflag);
  uint64_t host_sort(int (*)(char *, int, char *, int), int by_value);
  uint64_t sort_multivalues_impl(int flag, int (*)(char *, int, char *, int));
  uint64_t map_chunks(int nmap, int nstr, char **strings, int recurse, int readflag, int sepwhich, char sepchar, const char *sepstr, int delta,
                      void (*)(int, char *, int, class KeyValue *, void *), void *, int addflag);
  uint64_t reduce_impl(void (*)(char *, int, char *, int, int *, class KeyValue *, void *), void *, const char *heading);
};

}  // namespace MAPREDUCE_NS

#endif

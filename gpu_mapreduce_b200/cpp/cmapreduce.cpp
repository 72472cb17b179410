// C interface (cmapreduce.h): every entry point forwards to the mirror class.  The reference does the same with its
// own class (src/cmapreduce.cpp:23-462); callbacks written in C take `void *KVptr` where the C++ signature has
// `KeyValue *`, which is the same pointer, so the function pointers are converted, not wrapped.
#include "cmapreduce.h"

#include "keyvalue.h"
#include "mapreduce.h"

using namespace MAPREDUCE_NS;

typedef void (*MapTaskFn)(int, KeyValue *, void *);
typedef void (*MapFileFn)(int, char *, KeyValue *, void *);
typedef void (*MapStrFn)(int, char *, int, KeyValue *, void *);
typedef void (*MapMrFn)(uint64_t, char *, int, char *, int, KeyValue *, void *);
typedef void (*ReduceFn)(char *, int, char *, int, int *, KeyValue *, void *);

static inline MapReduce *M(void *p) { return (MapReduce *) p; }

void *MR_create(MPI_Comm comm) { return new MapReduce(comm); }
void *MR_create_mpi() { return new MapReduce(); }
void *MR_create_mpi_finalize() { return new MapReduce(0.0); }
void MR_destroy(void *MRptr) { delete M(MRptr); }
void *MR_copy(void *MRptr) { return M(MRptr)->copy(); }

uint64_t MR_add(void *MRptr, void *MRptr2) { return M(MRptr)->add(M(MRptr2)); }
uint64_t MR_aggregate(void *MRptr, int (*myhash)(char *, int)) { return M(MRptr)->aggregate(myhash); }
uint64_t MR_broadcast(void *MRptr, int root) { return M(MRptr)->broadcast(root); }
uint64_t MR_clone(void *MRptr) { return M(MRptr)->clone(); }
uint64_t MR_close(void *MRptr) { return M(MRptr)->close(); }
uint64_t MR_collapse(void *MRptr, char *key, int keybytes) { return M(MRptr)->collapse(key, keybytes); }
uint64_t MR_collate(void *MRptr, int (*myhash)(char *, int)) { return M(MRptr)->collate(myhash); }
uint64_t MR_compress(void *MRptr, void (*mycompress)(char *, int, char *, int, int *, void *, void *), void *APPptr)
{
  return M(MRptr)->compress((ReduceFn) mycompress, APPptr);
}
uint64_t MR_convert(void *MRptr) { return M(MRptr)->convert(); }
uint64_t MR_gather(void *MRptr, int numprocs) { return M(MRptr)->gather(numprocs); }

uint64_t MR_map(void *MRptr, int nmap, void (*mymap)(int, void *, void *), void *APPptr) { return M(MRptr)->map(nmap, (MapTaskFn) mymap, APPptr); }
uint64_t MR_map_add(void *MRptr, int nmap, void (*mymap)(int, void *, void *), void *APPptr, int addflag)
{
  return M(MRptr)->map(nmap, (MapTaskFn) mymap, APPptr, addflag);
}
uint64_t MR_map_file(void *MRptr, int nstr, char **strings, int self, int recurse, int readfile, void (*mymap)(int, char *, void *, void *),
                     void *APPptr)
{
  return M(MRptr)->map(nstr, strings, self, recurse, readfile, (MapFileFn) mymap, APPptr);
}
uint64_t MR_map_file_add(void *MRptr, int nstr, char **strings, int self, int recurse, int readfile,
                         void (*mymap)(int, char *, void *, void *), void *APPptr, int addflag)
{
  return M(MRptr)->map(nstr, strings, self, recurse, readfile, (MapFileFn) mymap, APPptr, addflag);
}
uint64_t MR_map_file_char(void *MRptr, int nmap, int nstr, char **strings, int recurse, int readflag, char sepchar, int delta,
                          void (*mymap)(int, char *, int, void *, void *), void *APPptr)
{
  return M(MRptr)->map(nmap, nstr, strings, recurse, readflag, sepchar, delta, (MapStrFn) mymap, APPptr);
}
uint64_t MR_map_file_char_add(void *MRptr, int nmap, int nstr, char **strings, int recurse, int readflag, char sepchar, int delta,
                              void (*mymap)(int, char *, int, void *, void *), void *APPptr, int addflag)
{
  return M(MRptr)->map(nmap, nstr, strings, recurse, readflag, sepchar, delta, (MapStrFn) mymap, APPptr, addflag);
}
uint64_t MR_map_file_str(void *MRptr, int nmap, int nstr, char **strings, int recurse, int readflag, char *sepstr, int delta,
                         void (*mymap)(int, char *, int, void *, void *), void *APPptr)
{
  return M(MRptr)->map(nmap, nstr, strings, recurse, readflag, sepstr, delta, (MapStrFn) mymap, APPptr);
}
uint64_t MR_map_file_str_add(void *MRptr, int nmap, int nstr, char **strings, int recurse, int readflag, char *sepstr, int delta,
                             void (*mymap)(int, char *, int, void *, void *), void *APPptr, int addflag)
{
  return M(MRptr)->map(nmap, nstr, strings, recurse, readflag, sepstr, delta, (MapStrFn) mymap, APPptr, addflag);
}
uint64_t MR_map_mr(void *MRptr, void *MRptr2, void (*mymap)(uint64_t, char *, int, char *, int, void *, void *), void *APPptr)
{
  return M(MRptr)->map(M(MRptr2), (MapMrFn) mymap, APPptr);
}
uint64_t MR_map_mr_add(void *MRptr, void *MRptr2, void (*mymap)(uint64_t, char *, int, char *, int, void *, void *), void *APPptr,
                       int addflag)
{
  return M(MRptr)->map(M(MRptr2), (MapMrFn) mymap, APPptr, addflag);
}

void MR_open(void *MRptr) { M(MRptr)->open(); }
void MR_open_add(void *MRptr, int addflag) { M(MRptr)->open(addflag); }
void MR_print(void *MRptr, int proc, int nstride, int kflag, int vflag) { M(MRptr)->print(proc, nstride, kflag, vflag); }
void MR_print_file(void *MRptr, char *file, int fflag, int proc, int nstride, int kflag, int vflag)
{
  M(MRptr)->print(file, fflag, proc, nstride, kflag, vflag);
}

uint64_t MR_reduce(void *MRptr, void (*myreduce)(char *, int, char *, int, int *, void *, void *), void *APPptr)
{
  return M(MRptr)->reduce((ReduceFn) myreduce, APPptr);
}
uint64_t MR_multivalue_blocks(void *MRptr, int *pnblock)
{
  int nblock = 0;
  const uint64_t n = M(MRptr)->multivalue_blocks(nblock);
  *pnblock = nblock;
  return n;
}
void MR_multivalue_block_select(void *MRptr, int which) { M(MRptr)->multivalue_block_select(which); }
int MR_multivalue_block(void *MRptr, int iblock, char **ptr_multivalue, int **ptr_valuesizes)
{
  return M(MRptr)->multivalue_block(iblock, ptr_multivalue, ptr_valuesizes);
}
uint64_t MR_scan_kv(void *MRptr, void (*myscan)(char *, int, char *, int, void *), void *APPptr) { return M(MRptr)->scan(myscan, APPptr); }
uint64_t MR_scan_kmv(void *MRptr, void (*myscan)(char *, int, char *, int, int *, void *), void *APPptr) { return M(MRptr)->scan(myscan, APPptr); }

uint64_t MR_scrunch(void *MRptr, int numprocs, char *key, int keybytes) { return M(MRptr)->scrunch(numprocs, key, keybytes); }
uint64_t MR_sort_keys(void *MRptr, int (*mycompare)(char *, int, char *, int)) { return M(MRptr)->sort_keys(mycompare); }
uint64_t MR_sort_keys_flag(void *MRptr, int flag) { return M(MRptr)->sort_keys(flag); }
uint64_t MR_sort_values(void *MRptr, int (*mycompare)(char *, int, char *, int)) { return M(MRptr)->sort_values(mycompare); }
uint64_t MR_sort_values_flag(void *MRptr, int flag) { return M(MRptr)->sort_values(flag); }
uint64_t MR_sort_multivalues(void *MRptr, int (*mycompare)(char *, int, char *, int)) { return M(MRptr)->sort_multivalues(mycompare); }
uint64_t MR_sort_multivalues_flag(void *MRptr, int flag) { return M(MRptr)->sort_multivalues(flag); }

uint64_t MR_kv_stats(void *MRptr, int level) { return M(MRptr)->kv_stats(level); }
uint64_t MR_kmv_stats(void *MRptr, int level) { return M(MRptr)->kmv_stats(level); }
void MR_cummulative_stats(void *MRptr, int level, int reset) { M(MRptr)->cummulative_stats(level, reset); }

void MR_set_mapstyle(void *MRptr, int value) { M(MRptr)->mapstyle = value; }
void MR_set_all2all(void *MRptr, int value) { M(MRptr)->all2all = value; }
void MR_set_verbosity(void *MRptr, int value) { M(MRptr)->verbosity = value; }
void MR_set_timer(void *MRptr, int value) { M(MRptr)->timer = value; }
void MR_set_memsize(void *MRptr, int value) { M(MRptr)->memsize = value; }
void MR_set_minpage(void *MRptr, int value) { M(MRptr)->minpage = value; }
void MR_set_maxpage(void *MRptr, int value) { M(MRptr)->maxpage = value; }
void MR_set_keyalign(void *MRptr, int value) { M(MRptr)->keyalign = value; }
void MR_set_valuealign(void *MRptr, int value) { M(MRptr)->valuealign = value; }
void MR_set_fpath(void *MRptr, char *str) { M(MRptr)->set_fpath(str); }

void MR_kv_add(void *KVptr, char *key, int keybytes, char *value, int valuebytes) { ((KeyValue *) KVptr)->add(key, keybytes, value, valuebytes); }
void MR_kv_add_multi_static(void *KVptr, int n, char *key, int keybytes, char *value, int valuebytes)
{
  ((KeyValue *) KVptr)->add(n, key, keybytes, value, valuebytes);
}
void MR_kv_add_multi_dynamic(void *KVptr, int n, char *key, int *keybytes, char *value, int *valuebytes)
{
  ((KeyValue *) KVptr)->add(n, key, keybytes, value, valuebytes);
}

/* mrgpu.h -- C ABI of the B200 (sm_100a) implementation of the MR-MPI data-parallel hot path:
 *   map emit -> KeyValue pages in HBM -> aggregate (hash partition + all-to-all) -> convert/collate
 *   (group by key -> KeyMultiValue) -> reduce / sort_keys / scan.
 *
 * This is the drop-in boundary: plain pointers, sizes and opaque handles; no
 * C++, torch or CUDA types in any signature (a stream crosses as void*).  The C++ classes
 * MAPREDUCE_NS::MapReduce / KeyValue in gpu_mapreduce_b200/cpp/ are a source-compatible mirror of the
 * reference's src/mapreduce.h:22-131 and src/keyvalue.h:23-59 that forwards here; a host language with
 * an FFI binds these symbols directly (INTEGRATION.md shows the stubs).
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference tree).
 *
 * Conventions
 *  - Every function returns 0 on success, a negative MRGPU_E* code on failure; mrgpu_last_error()
 *    returns the message.  The reference has no error codes: it prints "ERROR: msg" and exits
 *    (src/error.cpp:33-67); the C++ mirror restores exactly that behaviour from these codes and keeps
 *    the reference's message texts ("Cannot convert without KeyValue", ...).
 *  - Like the reference (src/mapreduce.h:43-44) an MR handle owns at most one KeyValue (KV) or one
 *    KeyMultiValue (KMV) at a time; operations consume one and produce the other.
 *  - Counts returned through `uint64_t *n` are the GLOBAL pair counts the reference returns from every
 *    call (MPI_Allreduce SUM, e.g. src/mapreduce.cpp:558), summed over the ranks of the communicator set
 *    with mrgpu_comm_init() (one rank = one process = one GPU).
 *  - Data pointers may be host or device pointers (detected with cudaPointerGetAttributes); host input is
 *    copied to the GPU inside the call.
 *  - All work is issued on the stream given at creation; calls that return counts synchronise it.
 *  - Not thread-safe per handle, same as the reference (SURVEY.md 8b "Threading").
 *  - There is no CPU implementation behind any of these calls: without a CUDA device every call fails
 *    with MRGPU_ECUDA.
 */
#ifndef MRGPU_H
#define MRGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mrgpu_mr mrgpu_mr;

enum {
  MRGPU_OK = 0,
  MRGPU_EINVAL = -1,   /* bad argument / misuse (reference: Error::all) */
  MRGPU_ECUDA = -2,    /* CUDA runtime failure, incl. no device */
  MRGPU_ENOKV = -3,    /* "Cannot <op> without KeyValue" */
  MRGPU_ENOKMV = -4,   /* "Cannot <op> without KeyMultiValue" */
  MRGPU_ELIMIT = -5,   /* exceeds a documented limit (pair > page, > 2^32-1 pairs per sort, ...) */
  MRGPU_ECOMM = -6     /* NCCL failure / communicator misuse */
};

/* device functors standing in for the reference's CPU callbacks (no CPU callback in the hot loop) */
enum {
  MRGPU_REDUCE_COUNT = 0, /* oink/reduce_count.cpp:13-20   value = int number of values            */
  MRGPU_REDUCE_SUM = 1,   /* examples/wordfreq.cpp:137-141 value = int nvalues                     */
  MRGPU_REDUCE_CULL = 2,  /* oink/reduce_cull.cpp:12-20    value = first value                     */
  /* cpu/InvertedIndex.cpp:335-385 (the InvertedIndex reduce writes `key \t v1 v2 ... vn \n` to a file): emits
   * (key, postings) where postings = every value, its terminating NUL turned into a space, in arrival order, closed
   * by one NUL -- the pair's text form `key \t value \n` is the reference's output line.  Values are
   * NUL-terminated strings (what the href scanner emits): every NUL byte of the multivalue becomes the space.
   * A postings string that does not fit a page is MRGPU_ELIMIT. */
  MRGPU_REDUCE_POSTINGS = 3,
  MRGPU_REDUCE_NONZERO = 4, /* examples/rmat.cpp:266-271  (first vertex of the edge key, NULL)             */
  MRGPU_REDUCE_DEGREE = 5,  /* examples/rmat.cpp:279-283  (int nvalues, NULL)                               */
  MRGPU_REDUCE_HISTO = 6    /* examples/rmat.cpp:291-295  (key, int nvalues)                                */
};
enum {
  MRGPU_MAP_EDGE_TO_VERTICES = 0, /* oink/map_edge_to_vertices.cpp:14-20  emit vi and vj           */
  MRGPU_MAP_EDGE_TO_VERTEX = 1,   /* oink/map_edge_to_vertex.cpp          emit vi                  */
  MRGPU_MAP_INVERT = 2            /* oink/map_invert.cpp:11-15            swap key and value       */
};

/* ---- lifetime and settings ------------------------------------------------------------------- */

/* MapReduce::MapReduce(MPI_Comm) src/mapreduce.cpp:93 + defaults() :196-262.  device = CUDA ordinal,
 * stream = cudaStream_t (NULL = the legacy default stream). */
int mrgpu_create(int device, void *stream, mrgpu_mr **out);
/* MapReduce::~MapReduce src/mapreduce.cpp:169.  Collective over the communicator once aggregate() has exchanged
 * pairs by peer stores (the ranks unmap each other's receive buffers, then meet in a barrier before any is freed),
 * like the reference destructor's MPI_Comm_free (src/mapreduce.cpp:283): every rank must call it, in the same
 * order relative to its other collective calls on that communicator. */
int mrgpu_destroy(mrgpu_mr *mr);
/* No counterpart in the reference.  Non-collective teardown for places that cannot promise the above -- a garbage
 * collector's finaliser, an error path on one rank: closes this rank's mappings of its peers' buffers, takes no
 * barrier, and leaves the buffers its peers may still have mapped allocated until the process exits (freeing a
 * buffer another process still maps is undefined in CUDA IPC).  Everything else is released. */
int mrgpu_abandon(mrgpu_mr *mr);
/* message of the last failure on this handle (or of the last failed mrgpu_create when mr == NULL) */
const char *mrgpu_last_error(const mrgpu_mr *mr);

/* Public settings of the reference object, src/mapreduce.h:28-39: "memsize" (MB per page if > 0, bytes if
 * < 0, src/mapreduce.cpp:3351), "keyalign", "valuealign", "verbosity", "timer", "all2all", "mapstyle",
 * "minpage", "maxpage", "freepage", "outofcore", "zeropage".  Alignments freeze at first use like the
 * reference's allocate() (src/mapreduce.cpp:3318-3341).
 * One setting of this library's own: "compact" (default 1).  With 1 a store may be held in HBM in a lossless
 * compact form -- a uniform KeyValue without its per-record length ints (and with a value shared by all
 * pairs kept once), a KeyMultiValue of indistinguishable pairs as (key, count) -- and the reference's record
 * stream is written when its bytes are asked for (page copies, device views, ops that read records).  With 0
 * every store always holds the record stream.  Results and page images are identical either way.
 * "ahead" (default 1): mrgpu_map_intcount's pass over the keys also does the first pass of the op that follows
 * it in every pipeline on one rank (convert's key count); the work
 * is dropped if the KeyValue changes before that op runs.  0 turns this off.
 * "combine" (default 1): aggregate() on several ranks, when every rank's pairs are 4-byte integer keys below 2^26
 * that all carry the same value (what mrgpu_map_intcount emits): such pairs cannot be told apart once their keys are
 * equal, so each rank sends one row (key, how many) per distinct key instead of every pair -- the reference's own
 * remedy for skewed keys, compress() before aggregate() (src/mapreduce.cpp:749-851, cpu/IntCount.cpp:78), applied
 * where it cannot change a result.  Pair counts, statistics, page table and every later result are those of the
 * plain exchange; the one observable difference is the ORDER of the pairs inside the KeyValue after aggregate():
 * still by source rank, but within a source by ascending key instead of arrival order.  0 = always the plain
 * exchange, in the reference's delivery order.  Must be the same on every rank.
 * "presorted" (default 1): convert() of pairs that cannot be told apart looks for one long ascending run among the
 * sort keys -- what an iterative job (generate, collate, cull, repeat) hands it: the previous round's output plus a
 * few new pairs -- and, when a run covers at least half of them, sorts only the others and merges.  The result is the
 * same KeyMultiValue; 0 = always sort everything.
 * "hbm_budget" (MB, default 0 = unlimited) and the reference's "outofcore" (1 = force data out of core; here: a budget
 * of one page): the out-of-core tier.  After a map-type call a KeyValue whose part in HBM exceeds the budget moves to
 * pinned host memory as a piece (the analogue of the reference's page spill files, src/keyvalue.cpp:216-255,
 * src/spool.cpp:76-195); later calls with addflag append to a fresh part in HBM.  convert() of such a KeyValue splits
 * every piece by key hash into sets that fit (KeyMultiValue::convert's partitions, src/keymultivalue.cpp:522-638,
 * 1056), converts set by set and leaves the KeyMultiValue in host pieces; reduce() with a stock functor walks them;
 * statistics, pages and scan read the pieces where they are.  Values of a key keep their arrival order.  Every other
 * operation brings the whole store back into HBM first.  The budget is checked between calls: one call's output is
 * the unit that moves.
 * "l2pin" (MiB, default 56; 0 = off): convert()'s counting path keeps the busy part of its count table resident in
 * L2 while it streams the keys past it.  For the duration of that ONE kernel launch it (a) puts a persisting
 * access-policy window on the handle's stream and (b) raises the DEVICE-WIDE limit cudaLimitPersistingL2CacheSize
 * to this many MiB if it is lower; when the launch has finished the window is removed and the limit restored, also
 * on an error.  If the limit was 0 on entry the persisting lines are reset (cudaCtxResetPersistingL2Cache); if the
 * host application had a limit of its own, its persisting lines are left alone.  A host that manages L2 persistence
 * itself and does not want it touched sets l2pin = 0 (the count then runs about 25 % slower on the IntCount workload).
 * "debug_fail_alloc" (test hook, default -1 = off): the n-th next allocation from the handle's HBM pool fails.
 * A failed NCCL call aborts the handle's communicator (ncclCommAbort) before the call returns MRGPU_ECOMM: the job
 * cannot continue, as after the reference's Error::one -> MPI_Abort (src/error.cpp:47-57). */
/* mrgpu_get also answers three read-only names, in bytes: "msize" / "msizemax" (HBM in use by the handle's page pool now /
 * at its high-water mark: the reference's msize, msizemax of src/mapreduce.h:54) and "reserved" (obtained from cudaMalloc). */
int mrgpu_set(mrgpu_mr *mr, const char *name, int64_t value);
int mrgpu_get(const mrgpu_mr *mr, const char *name, int64_t *value);

/* ---- communicator (replaces the MPI_Comm of src/mapreduce.cpp:93 and src/irregular.cpp) ---------- */

/* ncclGetUniqueId: fills 128 bytes; rank 0 calls it and the host program hands the bytes to all ranks. */
int mrgpu_comm_unique_id(void *id128);
/* ncclCommInitRank.  Without this call the handle is a single-rank communicator (like mpistubs). */
int mrgpu_comm_init(mrgpu_mr *mr, int rank, int nranks, const void *id128);
/* MapReduce::my_proc / num_procs src/mapreduce.h:125-126 */
int mrgpu_comm_rank(const mrgpu_mr *mr, int *rank, int *nranks);

/* ---- map: emit into KV pages ------------------------------------------------------------------- */

/* KeyValue::add(int n, char *key, int *keybytes, char *value, int *valuebytes) src/keyvalue.cpp:419-428
 * as called from a host map()/reduce() callback: appends n pairs given as packed key bytes / value bytes
 * plus per-pair lengths.  keybytes_fixed / valuebytes_fixed >= 0 with NULL arrays is the fixed-length form
 * src/keyvalue.cpp:399-412.  mrgpu_kv_open/close bracket a batch like MapReduce::open/close
 * (src/mapreduce.cpp:1650-1669); addflag as in map(). */
int mrgpu_kv_open(mrgpu_mr *mr, int addflag);
int mrgpu_kv_add(mrgpu_mr *mr, uint64_t n, const void *keys, const int32_t *keybytes, int keybytes_fixed,
                 const void *values, const int32_t *valuebytes, int valuebytes_fixed);
int mrgpu_kv_close(mrgpu_mr *mr, uint64_t *nkv);

/* map(nmap, mymap, ptr, addflag) src/mapreduce.cpp:1044 with the IntCount emitter cpu/IntCount.cpp:177-180
 * as a device functor: for each of the n uint32 in `keys`, key = its 4 bytes, value = int 1. */
int mrgpu_map_intcount(mrgpu_mr *mr, const uint32_t *keys, uint64_t n, int addflag, uint64_t *nkv);

/* map(files..., fileread) src/mapreduce.cpp:1060 with examples/wordfreq.cpp:104-130 as a device functor:
 * one call per file; text is cut at its first NUL, split on " \t\n\f\r"; key = word + NUL, no value. */
int mrgpu_map_tokenize(mrgpu_mr *mr, const void *text, uint64_t nbytes, int addflag, uint64_t *nkv);

/* map with the InvertedIndex scanner cpu/InvertedIndex.cpp:196-270 as a device functor: one call per
 * file; key = URL bytes (no NUL), value = fname including its NUL. */
int mrgpu_map_href(mrgpu_mr *mr, const void *text, uint64_t nbytes, const char *fname, int addflag,
                   uint64_t *nkv);

/* map(nprocs, rmat_generate, &rmat, addflag) oink/rmat.cpp:57 with oink/map_rmat_generate.cpp:14-67 as a
 * device functor.  *state48 is this rank's drand48 state (srand48(seed+me): (seed+me)<<16 | 0x330E,
 * oink/rmat.cpp:94); it is advanced past the draws consumed, as the reference's global stream is. */
int mrgpu_map_rmat(mrgpu_mr *mr, uint64_t ngenerate, int nlevels, double a, double b, double c, double d,
                   double fraction, uint64_t *state48, int addflag, uint64_t *nkv);

/* map(MapReduce *mr2, mymap, ptr, addflag) src/mapreduce.cpp:1560-1642 with a stock OINK callback
 * (MRGPU_MAP_*) as device functor.  src may equal dst (src/mapreduce.cpp:1584-1602). */
int mrgpu_map_mr(mrgpu_mr *dst, mrgpu_mr *src, int functor, int addflag, uint64_t *nkv);

/* ---- the hot path -------------------------------------------------------------------------------- */

/* MapReduce::add(MapReduce *) src/mapreduce.cpp:348-374 */
int mrgpu_add(mrgpu_mr *mr, mrgpu_mr *other, uint64_t *nkv);
/* MapReduce::copy() src/mapreduce.cpp:290-340: `mr` (freshly created) takes src's settings and a deep copy of its
 * KeyValue or KeyMultiValue */
int mrgpu_copy(mrgpu_mr *mr, mrgpu_mr *src);
/* MapReduce::aggregate(NULL) src/mapreduce.cpp:385-563 + Irregular src/irregular.cpp:95-301: pairs move to
 * rank hashlittle(key,keybytes,nprocs) % nprocs; arrival order = (source rank, original order). */
int mrgpu_aggregate(mrgpu_mr *mr, uint64_t *nkv);
/* MapReduce::aggregate(myhash) src/mapreduce.cpp:385-563 with a user hash function: the reference computes
 * proclist[i] = myhash(key,keybytes) % nprocs on the host (:466-470); here the caller does exactly that (the hash is
 * host code) and passes one destination rank per pair of this rank's KeyValue, in KV order, in host or device memory.
 * Delivery order as always: by source rank, original order within a source.  A destination outside [0, nprocs) is
 * reported as MRGPU_EINVAL (the reference would index out of bounds). */
int mrgpu_aggregate_dest(mrgpu_mr *mr, const int32_t *dest, uint64_t ndest, uint64_t *nkv);
/* The host half of Irregular::setup (src/irregular.cpp:95-242): given the all-gathered table of what every
 * rank sends -- row q (2*nranks+2 u64): {records, bytes} per destination, then {kind 0 empty / 1 variable / 2
 * uniform, keybytes<<32|valuebytes} -- compute what `rank` receives from each source (rcount, rbytes),
 * displacements in source-rank order (rdisp, rrec: nranks+1 entries) and the record kind/shape of the result.
 * Pure host arithmetic; mrgpu_aggregate uses it internally. */
int mrgpu_plan_exchange(int nranks, int rank, const uint64_t *table, uint64_t *rcount, uint64_t *rbytes,
                        uint64_t *rdisp, uint64_t *rrec, uint64_t *shape, int *kind);
/* Test / profiling hook for the local half of aggregate() (hash src/mapreduce.cpp:469-472 + pack
 * src/irregular.cpp:279-288) with a virtual rank count on one GPU: the KV is replaced by the same pairs grouped
 * stably by destination hashlittle(key,keybytes,nprocs) % nprocs; counts[d] = pairs for destination d. */
int mrgpu_debug_partition(mrgpu_mr *mr, int nprocs, uint64_t *counts);
/* MapReduce::convert src/mapreduce.cpp:861-886 -> KeyMultiValue::convert src/keymultivalue.cpp:486-638 */
int mrgpu_convert(mrgpu_mr *mr, uint64_t *nkmv);
/* MapReduce::collate(NULL) src/mapreduce.cpp:710-738 */
int mrgpu_collate(mrgpu_mr *mr, uint64_t *nkmv);
/* MapReduce::reduce(myreduce, ptr) src/mapreduce.cpp:1769-1867 with a stock callback (MRGPU_REDUCE_*) */
int mrgpu_reduce(mrgpu_mr *mr, int functor, uint64_t *nkv);
/* MapReduce::sort_keys(int flag) src/mapreduce.cpp:2102-2128; flags +-1..+-6 select the compare functions of
 * src/mapreduce.cpp:2692-2802 */
int mrgpu_sort_keys(mrgpu_mr *mr, int flag, uint64_t *nkv);
/* MapReduce::sort_values(int flag) src/mapreduce.cpp:2156-2182: the same compare functions applied to the values */
int mrgpu_sort_values(mrgpu_mr *mr, int flag, uint64_t *nkv);
/* MapReduce::compress(mycompress, ptr) src/mapreduce.cpp:749-851 with a stock callback (MRGPU_REDUCE_*): convert +
 * reduce on this rank's own pairs, no exchange; the KeyValue is replaced by one pair per distinct local key */
int mrgpu_compress(mrgpu_mr *mr, int functor, uint64_t *nkv);
/* MapReduce::gather(int numprocs) src/mapreduce.cpp:893-1036: ranks >= numprocs send their KeyValue to rank
 * (me % numprocs), which appends the senders' pairs in ascending rank order; senders are left with an empty KeyValue */
int mrgpu_gather(mrgpu_mr *mr, int numprocs, uint64_t *nkv);
/* MapReduce::broadcast(int root) src/mapreduce.cpp:569-625: every rank's KeyValue is replaced by a copy of root's */
int mrgpu_broadcast(mrgpu_mr *mr, int root, uint64_t *nkv);
/* MapReduce::clone() src/mapreduce.cpp:631-672: every KeyValue pair becomes a KeyMultiValue pair with one value */
int mrgpu_clone(mrgpu_mr *mr, uint64_t *nkmv);
/* MapReduce::collapse(key, keybytes) src/mapreduce.cpp:681-702: this rank's pairs become ONE KeyMultiValue pair whose
 * values are key_1, value_1, key_2, value_2, ...; like the reference (src/keymultivalue.cpp:475-478) it refuses a
 * result that does not fit one page */
int mrgpu_collapse(mrgpu_mr *mr, const void *key, int keybytes, uint64_t *nkmv);
/* MapReduce::scrunch(numprocs, key, keybytes) src/mapreduce.cpp:2075-2093: gather(numprocs) then collapse(key) */
int mrgpu_scrunch(mrgpu_mr *mr, int numprocs, const void *key, int keybytes, uint64_t *nkmv);

/* ---- stats and page access (scan, host callbacks, printing) --------------------------------------- */

typedef struct {
  uint64_t npair;      /* nkv / nkmv on this rank (KeyValue::nkv src/keyvalue.h:27)                  */
  uint64_t nvalue;     /* KMV only: total number of values                                           */
  uint64_t keysize;    /* exact key bytes    (ksize)                                                  */
  uint64_t valuesize;  /* exact value bytes  (vsize)                                                  */
  uint64_t exactsize;  /* esize: KV 8n+k+v (src/keyvalue.cpp:671), KMV 12n+4nv+k+v (keymultivalue.cpp:1460) */
  uint64_t alignsize;  /* bytes held in HBM in the reference page layout                              */
  uint64_t npage;      /* pages of `memsize` the reference would split this into                      */
} mrgpu_stats;

/* MapReduce::kv_stats / kmv_stats src/mapreduce.cpp:2937-3010 (this rank's part) */
int mrgpu_kv_stats(mrgpu_mr *mr, mrgpu_stats *out);
int mrgpu_kmv_stats(mrgpu_mr *mr, mrgpu_stats *out);

/* The return value of MapReduce::kv_stats / kmv_stats (src/mapreduce.cpp:2945-2958, 2988-3001): the pair count of the
 * KeyValue (which = 0) or KeyMultiValue (which = 1) summed over all ranks.  COLLECTIVE, unlike the two calls above. */
int mrgpu_global_count(mrgpu_mr *mr, int which, uint64_t *n);

/* KeyValue::request_info / request_page src/keyvalue.cpp:277-308: page ipage of the KV in the reference's
 * byte layout (int keybytes | int valuebytes | pad | key | pad | value | pad).  mrgpu_kv_page_info gives
 * nkey and sizes; mrgpu_kv_page_copy writes alignsize bytes to host memory `dst` (padding zeroed). */
int mrgpu_kv_page_info(mrgpu_mr *mr, uint64_t ipage, uint64_t *nkey, uint64_t *keysize, uint64_t *valuesize,
                       uint64_t *alignsize);
int mrgpu_kv_page_copy(mrgpu_mr *mr, uint64_t ipage, void *dst);
/* KeyMultiValue::request_info / request_page src/keymultivalue.cpp:223-260: layout int nvalue | int keybytes |
 * int mvaluebytes | int valuesizes[nvalue] | pad | key | pad | values | pad.  A pair larger than a page
 * occupies a page run of its own (is_extended = 1, nkey = 1) that the host mirror serves through the
 * reference's multivalue_blocks()/multivalue_block() protocol (src/mapreduce.cpp:1874-1925). */
int mrgpu_kmv_page_info(mrgpu_mr *mr, uint64_t ipage, uint64_t *nkey, uint64_t *alignsize, int *is_extended);
int mrgpu_kmv_page_copy(mrgpu_mr *mr, uint64_t ipage, void *dst);

/* MapReduce::scan(myscan, ptr) src/mapreduce.cpp:1933-1976 (KV form) and :1984-2065 (KMV form): read-only walk that
 * calls a HOST function for every pair.  Pages are staged to the host one at a time; the pointers a callback gets are
 * valid only during the callback (the reference's rule).  Returns the global pair count like the reference.
 * A KMV pair bigger than a page reaches the callback as multivalue == NULL, nvalues == 0 with the handle in the
 * valuesizes slot (the reference passes its MapReduce* there, src/mapreduce.cpp:2041-2053); inside the callback
 * mrgpu_multivalue_blocks / _block_select / _block give the values block by block:
 * MapReduce::multivalue_blocks src/mapreduce.cpp:1874, multivalue_block_select :1887, multivalue_block :1901. */
typedef void (*mrgpu_scan_kv_fn)(char *key, int keybytes, char *value, int valuebytes, void *ptr);
typedef void (*mrgpu_scan_kmv_fn)(char *key, int keybytes, char *multivalue, int nvalues, int *valuesizes, void *ptr);
int mrgpu_scan_kv(mrgpu_mr *mr, mrgpu_scan_kv_fn fn, void *ptr, uint64_t *nkv);
int mrgpu_scan_kmv(mrgpu_mr *mr, mrgpu_scan_kmv_fn fn, void *ptr, uint64_t *nkmv);
int mrgpu_multivalue_blocks(mrgpu_mr *mr, int *nblock, uint64_t *nvalue_total);
int mrgpu_multivalue_block_select(mrgpu_mr *mr, int which);
int mrgpu_multivalue_block(mrgpu_mr *mr, int iblock, char **multivalue, int **valuesizes, int *nvalues);

/* No counterpart in the reference (its pages live in host memory: KeyValue::request_page src/keyvalue.cpp:288).
 * Device views for callers that stay on the GPU (device functors of their own, torch tensors, bench):
 * base pointer of the packed record stream, its byte length, and the u64 record-offset index
 * (NULL when all records have one size: then record i starts at i * record_bytes). */
int mrgpu_kv_device_view(mrgpu_mr *mr, const void **data, uint64_t *nbytes, const uint64_t **offsets,
                         uint64_t *record_bytes);
int mrgpu_kmv_device_view(mrgpu_mr *mr, const void **data, uint64_t *nbytes, const uint64_t **offsets);

/* ---- per-kernel timing ---------------------------------------------------------------------------- */

/* The reference's `timer` setting makes every op print "<Op> time (secs)" (src/mapreduce.cpp:3112-3128,
 * 3230-3234).  mrgpu_prof(mr, 1) resets the counters and brackets every tagged kernel launch with CUDA events
 * on the handle's stream; mrgpu_prof_report() synchronises and writes a JSON array
 * [{"tag", "ms", "bytes" (algorithmic bytes), "launches", "calls"}] into buf and the number of
 * kernel launches since the last mrgpu_prof() call into *launches (counted whether or not timing is on). */
int mrgpu_prof(mrgpu_mr *mr, int enable);
int mrgpu_prof_report(mrgpu_mr *mr, char *buf, uint64_t cap, uint64_t *launches);

/* ---- stand-alone kernels exposed for tests and profiling ---------------------------------------- */

/* hashlittle src/hash.cpp:129 over n keys given as packed bytes + u64 offsets[n+1]; out[i] = hash. */
int mrgpu_hashlittle(int device, const void *keys, const uint64_t *offsets, uint64_t n, uint32_t initval,
                     uint32_t *out);
/* Stable LSD radix sort of n (key, payload) pairs, the engine under convert and sort_keys (replaces the
 * qsort + merge of src/mapreduce.cpp:2462-2633).  key_bytes 4 or 8, payload_bytes 0, 4 or 8; device
 * pointers; sorted data is written back to keys/payload. */
int mrgpu_sort_pairs(int device, void *stream, void *keys, void *payload, uint64_t n, int key_bytes,
                     int payload_bytes);

/* No counterpart in the reference (its sort is a host qsort, nothing there distinguishes callers by whether
 * they may allocate GPU memory).
 *
 * Same result as mrgpu_sort_pairs(), for callers that must not allocate on the GPU: a host application
 * whose own arena already owns nearly all device memory (a later driver allocation may fail at any time), or a
 * latency-critical path that cannot take the device-wide synchronisation of cudaMalloc/cudaFree.  This call
 * makes no cudaMalloc/cudaFree and uses no allocating library routine: the second key/payload buffers, the
 * digit histograms and the look-back state live in one static __device__ array inside libmrgpu.so (about
 * 16.5 MiB, resident on every GPU that loads the library until the process exits).  keys/payload stay yours.
 *  - Limit: n <= 1 048 576 pairs; a larger n fails with MRGPU_ELIMIT and a message that states the limit.
 *  - The first call on a device looks up the array's device address (one cudaGetSymbolAddress, cached
 *    afterwards).  With lazy module loading that first call is also where the driver loads this library's GPU
 *    module and gives the array its 16.5 MiB, so make it during start-up, while memory is still available.
 *  - Every call takes a host mutex and ends in cudaStreamSynchronize: it must not be made while a stream capture
 *    is in progress, the first call or any later one.
 *  - All calls share the one array per GPU and it keeps the previous call's contents; the call clears what it
 *    needs cleared.  Two calls whose GPU work overlaps on the same device (different streams or threads) race
 *    on the scratch and corrupt each other: serialise them.
 *  - Returns after the stream has drained, like mrgpu_sort_pairs() (see the capture note above). */
int mrgpu_sort_pairs_noalloc(int device, void *stream, void *keys, void *payload, uint64_t n, int key_bytes,
                             int payload_bytes);

#ifdef __cplusplus
}
#endif
#endif

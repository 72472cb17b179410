/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference MR-MPI hot path.
 *
 * Nothing in the product (gpu_mapreduce_b200/, include/, the C-ABI library) links, imports or calls
 * this file.  It is used by tests/, by __graft_entry__.smoke() as the checker, and by bench.py's
 * cpu_baseline / --impl reference legs when oracle/_ref (the compiled reference itself) is absent.
 *
 * Parity status: PINNED.  Every function below is checked in tests/test_oracle_*.py against outputs
 * of the unmodified reference library built by oracle/Makefile into oracle/_ref/refdrv (live when the
 * binary is present, and through the fixtures under tests/golden/ generated from it by
 * tests/golden/make_golden.py).  The reference ships no tests or golden vectors of its own
 * (SURVEY.md section 4), so those recorded reference outputs are the ground truth.
 *
 * Data crosses this API in two unpadded little-endian interchange formats (the same ones
 * oracle/ref_driver.cpp reads and writes):
 *   kvfile : u64 n ; n x { i32 keybytes ; i32 valuebytes ; key ; value }
 *   kmvfile: u64 n ; n x { i32 keybytes ; u64 nvalue ; key ; nvalue x { i32 valuebytes ; value } }
 * Output buffers are malloc'ed by the oracle and released with orc_free().
 */
#ifndef MR_ORACLE_H
#define MR_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

void orc_free(void *p);

/* Jenkins lookup3 hashlittle -- src/hash.cpp:129-333 (mix :56-64, final :91-100). */
uint32_t orc_hashlittle(const void *key, size_t length, uint32_t initval);

/* Destination rank of a key in aggregate(): hashlittle(key,keybytes,nprocs) % nprocs --
 * src/mapreduce.cpp:469-472.  out[i] for every record of the kvfile. */
int orc_partition(const uint8_t *kv, int nprocs, int32_t *out);

/* Aligned size of one KV record -- src/keyvalue.cpp:345-352. */
uint64_t orc_kv_record_size(int keybytes, int valuebytes, int kalign, int valign);
/* Aligned size of one KMV record -- src/keymultivalue.cpp:830-838. */
uint64_t orc_kmv_record_size(uint64_t nvalue, int keybytes, uint64_t mvbytes, int kalign, int valign);

/* KV page images in the reference layout, greedy page fill (src/keyvalue.cpp:343-392).  Output format =
 * refdrv "pagekv": u64 npage ; npage x { i32 nkey ; u64 alignsize ; bytes }.  Padding bytes are zero. */
int orc_kv_pages(const uint8_t *kv, uint64_t pagesize, int kalign, int valign, uint8_t **out,
                 uint64_t *outbytes);
/* KMV page images for single-page KMV pairs (src/keymultivalue.cpp:800-893,1139-1209); returns -1 if a pair
 * would need the multi-block form.  Same container format as orc_kv_pages. */
int orc_kmv_pages(const uint8_t *kmv, uint64_t pagesize, int kalign, int valign, uint8_t **out,
                  uint64_t *outbytes);

/* convert(): group by key; unique keys in first-appearance order, values in arrival order --
 * src/keymultivalue.cpp:486-638 (kv2unique :645-789, kv2kmv :1139-1209), in-core case. */
int orc_convert(const uint8_t *kv, uint8_t **kmv_out, uint64_t *outbytes);

/* reduce() with the stock callbacks.  which: 0 = count (oink/reduce_count.cpp:13-20, value = int total
 * number of values), 1 = sum (examples/wordfreq.cpp:137-141, identical for single-page pairs),
 * 2 = cull (oink/reduce_cull.cpp:12-20, value = first value). */
int orc_reduce(const uint8_t *kmv, int which, uint8_t **kv_out, uint64_t *outbytes);

/* sort_keys(flag), flags +-1..+-6 -- src/mapreduce.cpp:2102-2149 with the compare functions at
 * :2692-2802; stable (glibc qsort is a merge sort here). */
int orc_sort_keys(const uint8_t *kv, int flag, uint8_t **kv_out, uint64_t *outbytes);

/* map(mr, edge_to_vertices) / edge_to_vertex -- oink/map_edge_to_vertices.cpp:14-20,
 * oink/map_edge_to_vertex.cpp.  both = 1 emits vi and vj, 0 emits vi only. */
int orc_map_edge_to_vertices(const uint8_t *kv, int both, uint8_t **kv_out, uint64_t *outbytes);

/* wordfreq fileread tokenizer -- examples/wordfreq.cpp:104-130: strtok on " \t\n\f\r", text ends at the
 * first NUL, key = word + NUL, no value. */
int orc_tokenize(const uint8_t *text, uint64_t nchar, uint8_t **kv_out, uint64_t *outbytes);

/* InvertedIndex href scanner -- cpu/InvertedIndex.cpp:196-270 restated as the 8-state DFA of SURVEY.md
 * B.8: key = URL bytes without NUL, value = fname including its NUL. */
int orc_href(const uint8_t *text, uint64_t nchar, const char *fname, uint8_t **kv_out,
             uint64_t *outbytes);

/* glibc drand48: X' = (0x5DEECE66D X + 0xB) mod 2^48, value X'/2^48; srand48(s): X = (s<<16)|0x330E. */
uint64_t orc_srand48(uint32_t seed);
double orc_drand48(uint64_t *state);

/* rmat_generate -- oink/map_rmat_generate.cpp:14-67.  Appends ngenerate edges (vi,vj as u64 pairs) to
 * edges[] and advances *state. */
void orc_rmat_generate(uint64_t *state, uint64_t ngenerate, int nlevels, double a, double b, double c,
                       double d, double fraction, uint64_t *edges);

/* IntCount emit -- cpu/IntCount.cpp:177-180: for each u32 of the input, key = its 4 bytes, value = int 1. */
int orc_intcount_map(const uint32_t *keys, uint64_t n, uint8_t **kv_out, uint64_t *outbytes);

#ifdef __cplusplus
}
#endif
#endif

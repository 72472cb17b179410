/* Stock callbacks with the reference's callback signatures.  Passing one of THESE function pointers to
 * MapReduce::reduce / compress / map(MapReduce*, ...) selects the device functor of the same name -- no CPU
 * callback runs in the hot loop.  (A caller's own function with the same body runs on the host instead.)
 *   count / sum / cull            oink/reduce_count.cpp:13-20, examples/wordfreq.cpp:137-141, oink/reduce_cull.cpp:12-20
 *   edge_to_vertices / _vertex    oink/map_edge_to_vertices.cpp:14-20, oink/map_edge_to_vertex.cpp
 *   invert                        oink/map_invert.cpp:11-15
 *
 * Map callbacks work the same way: MapReduce::map(nmap, gpu::intcount, &input) / map(nstr, strings, ..., gpu::fileread, ...)
 * run the device functor for every task this rank owns (mapstyle 0/1 task split as in the reference):
 *   intcount        cpu/IntCount.cpp:177-180            task i emits (key, int 1) for every uint32 of its key array
 *   fileread        examples/wordfreq.cpp:104-130       task = one file: words split on " \t\n\f\r", key = word + NUL
 *   href            cpu/InvertedIndex.cpp:196-270       task = one file: key = URL of every <a href="...">, value = file name + NUL
 *   rmat_generate   oink/map_rmat_generate.cpp:14-67    task emits ngenerate R-MAT edges from this rank's drand48 stream */
#ifndef MRGPU_FUNCTORS_H
#define MRGPU_FUNCTORS_H

#include <stdint.h>

namespace MAPREDUCE_NS {
class KeyValue;
namespace gpu {
// ptr of map(nmap, gpu::intcount, ptr): per-task key arrays (host or device memory); only the entries of the tasks
// this rank runs are read
struct IntCountInput {
  const uint32_t *const *task_keys;
  const uint64_t *task_nkeys;
};
// ptr of map(nmap, gpu::rmat_generate, ptr): what oink/rmat.cpp:53-57 passes, plus the rank's drand48 state
// ((seed + me) << 16 | 0x330E after srand48(seed + me), oink/rmat.cpp:94), which is advanced past the draws used
struct RmatInput {
  uint64_t ngenerate;
  int nlevels;
  double a, b, c, d, fraction;
  uint64_t state48;
};
void intcount(int, KeyValue *, void *);
void rmat_generate(int, KeyValue *, void *);
void fileread(int, char *, KeyValue *, void *);
void href(int, char *, KeyValue *, void *);
void count(char *, int, char *, int, int *, KeyValue *, void *);
void sum(char *, int, char *, int, int *, KeyValue *, void *);
void cull(char *, int, char *, int, int *, KeyValue *, void *);
// cpu/InvertedIndex.cpp:335-385 as a pair instead of a file line: (key, "v1 v2 ... vn " + NUL); examples/rmat.cpp:266-295
void postings(char *, int, char *, int, int *, KeyValue *, void *);
void nonzero(char *, int, char *, int, int *, KeyValue *, void *);
void degree(char *, int, char *, int, int *, KeyValue *, void *);
void histo(char *, int, char *, int, int *, KeyValue *, void *);
void edge_to_vertices(uint64_t, char *, int, char *, int, KeyValue *, void *);
void edge_to_vertex(uint64_t, char *, int, char *, int, KeyValue *, void *);
void invert(uint64_t, char *, int, char *, int, KeyValue *, void *);
}  // namespace gpu
}  // namespace MAPREDUCE_NS

#endif

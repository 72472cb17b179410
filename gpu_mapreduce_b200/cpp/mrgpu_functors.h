/* Stock callbacks with the reference's callback signatures.  Passing one of THESE function pointers to
 * MapReduce::reduce / compress / map(MapReduce*, ...) selects the device functor of the same name -- no CPU
 * callback runs in the hot loop.  (A caller's own function with the same body runs on the host instead.)
 *   count / sum / cull            oink/reduce_count.cpp:13-20, examples/wordfreq.cpp:137-141, oink/reduce_cull.cpp:12-20
 *   edge_to_vertices / _vertex    oink/map_edge_to_vertices.cpp:14-20, oink/map_edge_to_vertex.cpp
 *   invert                        oink/map_invert.cpp:11-15 */
#ifndef MRGPU_FUNCTORS_H
#define MRGPU_FUNCTORS_H

#include <stdint.h>

namespace MAPREDUCE_NS {
class KeyValue;
namespace gpu {
void count(char *, int, char *, int, int *, KeyValue *, void *);
void sum(char *, int, char *, int, int *, KeyValue *, void *);
void cull(char *, int, char *, int, int *, KeyValue *, void *);
void edge_to_vertices(uint64_t, char *, int, char *, int, KeyValue *, void *);
void edge_to_vertex(uint64_t, char *, int, char *, int, KeyValue *, void *);
void invert(uint64_t, char *, int, char *, int, KeyValue *, void *);
}  // namespace gpu
}  // namespace MAPREDUCE_NS

#endif

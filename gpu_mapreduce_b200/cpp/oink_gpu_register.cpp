// Ties OINK's stock callbacks to device functors, by address.
//
// OINK (the reference's script front-end, oink/*.cpp) is compiled UNMODIFIED against the mirror classes; its commands
// call mr->map(nprocs, rmat_generate, &rmat, 1), mr->reduce(cull, ...), mrv->map(mre, edge_to_vertices, NULL),
// mrv->reduce(count, NULL) ... (oink/rmat.cpp:53-59, oink/degree.cpp:43-50) with OINK's own host functions.  This
// one extra translation unit, linked into the OINK executable, registers those function addresses with
// MapReduce::device_* (cpp/mapreduce.h): from then on the calls above run the device functor of the same meaning and
// no pair goes through a host callback.  Callbacks that are not listed here keep running on the host, as in the
// reference.  Written against OINK's headers (compiled in the scratch copy __graft_entry__.build() makes).
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "mapreduce.h"
#include "keyvalue.h"
#include "mrgpu_functors.h"
#include "mrgpu.h"

#include "map_rmat_generate.h"  // OINK: struct RMAT_struct (oink/map_rmat_generate.h)

using namespace MAPREDUCE_NS;

// OINK's callbacks (global functions; oink/map_*.cpp, oink/reduce_*.cpp)
void rmat_generate(int, KeyValue *, void *);
void cull(char *, int, char *, int, int *, KeyValue *, void *);
void count(char *, int, char *, int, int *, KeyValue *, void *);
void edge_to_vertices(uint64_t, char *, int, char *, int, KeyValue *, void *);
void edge_to_vertex(uint64_t, char *, int, char *, int, KeyValue *, void *);
void invert(uint64_t, char *, int, char *, int, KeyValue *, void *);

// oink/map_rmat_generate.cpp:14-67 draws from the C library's drand48 stream that RMAT::params seeded with
// srand48(seed + me) (oink/rmat.cpp:94).  The device functor takes that stream's 48-bit state, generates the same edges
// from the same draws, and hands the advanced state back, so host code that draws afterwards continues the sequence.
static uint64_t rmat_on_device(MapReduce *mr, int nmap, void *ptr, int addflag)
{
  RMAT_struct *r = (RMAT_struct *) ptr;
  unsigned short zero[3] = {0, 0, 0}, cur[3];
  unsigned short *old = seed48(zero);  // returns the state it replaces: the only way to read it
  memcpy(cur, old, sizeof(cur));
  gpu::RmatInput in;
  in.ngenerate = r->ngenerate;
  in.nlevels = r->nlevels;
  in.a = r->a, in.b = r->b, in.c = r->c, in.d = r->d, in.fraction = r->fraction;
  in.state48 = (uint64_t) cur[0] | ((uint64_t) cur[1] << 16) | ((uint64_t) cur[2] << 32);
  const uint64_t n = mr->map(nmap, gpu::rmat_generate, &in, addflag);
  unsigned short next[3] = {(unsigned short) (in.state48 & 0xFFFF), (unsigned short) ((in.state48 >> 16) & 0xFFFF),
                            (unsigned short) ((in.state48 >> 32) & 0xFFFF)};
  seed48(next);
  return n;
}

namespace {
struct Register {
  Register()
  {
    MapReduce::device_map_task(rmat_generate, rmat_on_device);
    MapReduce::device_reduce(cull, MRGPU_REDUCE_CULL);
    MapReduce::device_reduce(count, MRGPU_REDUCE_COUNT);
    MapReduce::device_map_mr(edge_to_vertices, MRGPU_MAP_EDGE_TO_VERTICES);
    MapReduce::device_map_mr(edge_to_vertex, MRGPU_MAP_EDGE_TO_VERTEX);
    MapReduce::device_map_mr(invert, MRGPU_MAP_INVERT);
  }
} register_oink_callbacks;
}  // namespace

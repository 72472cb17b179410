// Compact sort keys of uniform keys longer than 8 bytes (ops_convert.cuh): only the key bytes that vary at all take
// part in the sort.  For 16-byte keys (R-MAT's EDGE) the key is handled as two 64-bit halves and every shift amount is
// the same in all threads.  In a header of its own so that a host program can check the two directions against each
// other and against the order they must preserve (tests/cpp/compact_host_test.cu).
#pragma once

struct CompactKey {
  int nvb;               // number of varying key bytes (<= 8)
  unsigned char pos[8];  // their positions inside the key, ascending
};

// sort key: byte q = key byte pos[q].  Positions ascend, so comparing sort keys as integers compares the keys as
// little-endian integers over the varying bytes -- and the other bytes are equal in every key.
__host__ __device__ __forceinline__ unsigned long long compact_of_key16(unsigned long long lo, unsigned long long hi, const CompactKey &ck)
{
  unsigned long long k64 = 0;
#pragma unroll
  for (int q = 0; q < 8; q++)
    if (q < ck.nvb) {
      const unsigned p = ck.pos[q];
      k64 |= (((p < 8 ? lo : hi) >> (8 * (p & 7))) & 0xFFull) << (8 * q);
    }
  return k64;
}

// the first key's halves with the varying bytes cleared: what every key has in common
inline void key16_clear_varying(unsigned long long &lo, unsigned long long &hi, const CompactKey &ck)
{
  for (int q = 0; q < ck.nvb; q++) {
    unsigned long long &h = ck.pos[q] < 8 ? lo : hi;
    h &= ~(0xFFull << (8 * (ck.pos[q] & 7)));
  }
}

// the key back from its sort key: byte q of the sort key goes to position pos[q] of (lo0, hi0)
__host__ __device__ __forceinline__ void key16_of_compact(unsigned long long key, unsigned long long lo0, unsigned long long hi0, const CompactKey &ck,
                                                          unsigned long long &lo, unsigned long long &hi)
{
  lo = lo0, hi = hi0;
#pragma unroll
  for (int q = 0; q < 8; q++)
    if (q < ck.nvb) {
      const unsigned p = ck.pos[q];
      const unsigned long long b = ((key >> (8 * q)) & 0xFFull) << (8 * (p & 7));
      if (p < 8) lo |= b;
      else hi |= b;
    }
}

// Order-preserving transforms of sort_keys / sort_values (ops_sort.cuh): a float, a double or 8 bytes of a C string
// become an unsigned integer that compares the way the reference's compare functions do (src/mapreduce.cpp:2692-2802).
// Plain C apart from the qualifiers, in a header of its own so that a host program can check them against the C
// comparisons themselves (tests/cpp/sortorder_host_test.cu).
#pragma once

#include <stdint.h>

__host__ __device__ __forceinline__ uint32_t f32_order(uint32_t b)
{
  if ((b << 1) == 0) b = 0;  // -0 -> +0
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ uint64_t f64_order(uint64_t b)
{
  if ((b << 1) == 0) b = 0;
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}

// chunk c (8 bytes, big-endian) of the NUL-terminated string stored in key[0..kb)
__host__ __device__ __forceinline__ uint64_t str_chunk(const uint8_t *key, uint32_t kb, uint32_t c)
{
  uint64_t v = 0;
  uint32_t lo = c * 8;
  // bytes after the first NUL compare as 0; find whether a NUL occurs before this chunk
  bool ended = false;
  for (uint32_t q = 0; q < lo && q < kb; q++)
    if (key[q] == 0) {
      ended = true;
      break;
    }
  for (uint32_t q = 0; q < 8; q++) {
    uint32_t p = lo + q;
    uint8_t b = 0;
    if (!ended && p < kb) {
      b = key[p];
      if (b == 0) ended = true;
    }
    v = (v << 8) | b;
  }
  return v;
}


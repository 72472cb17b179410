// The count kernel's cache addressing (ops_convert_dense.cuh), in a header of its own so that a host program can
// check it over the whole key range (tests/cpp/dense_host_test.cu): key -> (slot, tag) and back.
// Keys are < 2^26; h = key * odd mod 2^26 is a bijection, the slot is floor(h * DENSE_CACHE / 2^26), and the at most
// 1171 values of h that share a slot differ in their low 11 bits, so tag = (h & 2047) + 1 identifies the key exactly
// (0 = empty) and the key is recovered from (slot, tag) when an entry is flushed.
#pragma once

#include <stdint.h>

constexpr uint32_t DENSE_BITS = 26;                 // table of 2^26 u32 counts = 256 MB (touched part stays in L2)
constexpr uint32_t DENSE_SLOTS = 1u << DENSE_BITS;
constexpr uint32_t DENSE_MASK = DENSE_SLOTS - 1;
constexpr uint32_t DENSE_CACHE = 57344;
constexpr uint32_t DENSE_MULT = 0x278DC01u;         // odd; spreads consecutive integers over the 26-bit range
constexpr uint32_t dense_inv26(uint32_t a)
{
  uint32_t x = a;  // Newton: each step doubles the number of correct low bits
  for (int i = 0; i < 6; i++) x *= 2u - a * x;
  return x & DENSE_MASK;
}
constexpr uint32_t DENSE_MULT_INV = dense_inv26(DENSE_MULT);
static_assert(((DENSE_MULT * DENSE_MULT_INV) & DENSE_MASK) == 1u, "inverse of the hash multiplier");
static_assert((DENSE_SLOTS + DENSE_CACHE - 1) / DENSE_CACHE <= 2048, "an 11-bit tag must separate the keys of a slot");

__host__ __device__ __forceinline__ void dense_slot(uint32_t key, uint32_t &slot, uint32_t &tag)
{
  const uint32_t h = (key * DENSE_MULT) & DENSE_MASK;
#ifdef __CUDA_ARCH__
  slot = __umulhi(h << (32 - DENSE_BITS), DENSE_CACHE);
#else
  slot = (uint32_t) (((uint64_t) (h << (32 - DENSE_BITS)) * DENSE_CACHE) >> 32);
#endif
  tag = (h & 2047u) + 1u;
}

// the key whose entry sits in `slot` with this tag: h lies in [lo, lo + 1171) where lo = ceil(slot * 2^26 / DENSE_CACHE),
// and its low 11 bits are tag - 1
__host__ __device__ __forceinline__ uint32_t dense_key_of(uint32_t slot, uint32_t tag)
{
  const uint32_t lo = (uint32_t) ((((uint64_t) slot << DENSE_BITS) + DENSE_CACHE - 1) / DENSE_CACHE);
  const uint32_t h = lo + ((tag - 1u - lo) & 2047u);
  return (h * DENSE_MULT_INV) & DENSE_MASK;
}

// Bob Jenkins' lookup3 hashlittle (public domain), device version of the reference's src/hash.cpp:129-333.
// The reference picks one of three read strategies by pointer alignment; on little-endian hardware all
// three return the same value, so the device code assembles the 32-bit words from bytes (keys inside KV
// pages are only guaranteed keyalign-aligned).  Bit-exactness matters: aggregate() places a pair on rank
// hashlittle(key,keybytes,nprocs) % nprocs (src/mapreduce.cpp:469-472) and results must land where the
// reference would put them.
#pragma once

#include "common.cuh"

namespace mrgpu {

__host__ __device__ __forceinline__ uint32_t rot32(uint32_t x, int k) { return (x << k) | (x >> (32 - k)); }

#define MR_MIX(a, b, c)  \
  {                      \
    a -= c;              \
    a ^= rot32(c, 4);    \
    c += b;              \
    b -= a;              \
    b ^= rot32(a, 6);    \
    a += c;              \
    c -= b;              \
    c ^= rot32(b, 8);    \
    b += a;              \
    a -= c;              \
    a ^= rot32(c, 16);   \
    c += b;              \
    b -= a;              \
    b ^= rot32(a, 19);   \
    a += c;              \
    c -= b;              \
    c ^= rot32(b, 4);    \
    b += a;              \
  }

#define MR_FINAL(a, b, c) \
  {                       \
    c ^= b;               \
    c -= rot32(b, 14);    \
    a ^= c;               \
    a -= rot32(c, 11);    \
    b ^= a;               \
    b -= rot32(a, 25);    \
    c ^= b;               \
    c -= rot32(b, 16);    \
    a ^= c;               \
    a -= rot32(c, 4);     \
    b ^= a;               \
    b -= rot32(a, 14);    \
    c ^= b;               \
    c -= rot32(b, 24);    \
  }

// little-endian load of up to 4 bytes k[0..m)
__host__ __device__ __forceinline__ uint32_t le_bytes(const uint8_t *k, int m)
{
  uint32_t v = 0;
  for (int i = 0; i < m; i++) v |= (uint32_t) k[i] << (8 * i);
  return v;
}

__host__ __device__ inline uint32_t hashlittle(const uint8_t *k, uint32_t length, uint32_t initval)
{
  uint32_t a, b, c;
  a = b = c = 0xdeadbeefu + length + initval;
  while (length > 12) {
    a += le_bytes(k, 4);
    b += le_bytes(k + 4, 4);
    c += le_bytes(k + 8, 4);
    MR_MIX(a, b, c);
    length -= 12;
    k += 12;
  }
  if (length == 0) return c;  // src/hash.cpp "case 0: return c" -- zero-length tail skips final()
  // tail of 1..12 bytes: missing bytes contribute zero, exactly like the reference's switch fall-through
  int la = length < 4 ? (int) length : 4;
  int lb = length < 4 ? 0 : (length < 8 ? (int) length - 4 : 4);
  int lc = length < 8 ? 0 : (int) length - 8;
  a += le_bytes(k, la);
  if (lb) b += le_bytes(k + 4, lb);
  if (lc) c += le_bytes(k + 8, lc);
  MR_FINAL(a, b, c);
  return c;
}

// fixed-length fast paths used for uniform records (keys are 4-byte aligned inside pages when keyalign>=4)
__host__ __device__ __forceinline__ uint32_t hashlittle_words(const uint32_t *w, int nwords /*1..3*/, uint32_t initval)
{
  uint32_t a, b, c;
  a = b = c = 0xdeadbeefu + (uint32_t) (nwords * 4) + initval;
  a += w[0];
  if (nwords > 1) b += w[1];
  if (nwords > 2) c += w[2];
  MR_FINAL(a, b, c);
  return c;
}

// 64-bit grouping fingerprint of convert(): (c << 32) | b of ONE lookup3 pass (Jenkins' hashlittle2 returns the same two
// words as a 64-bit hash); for seed 0 the high half is exactly hashlittle(key, kb, 0), the reference's bucket hash
// (src/keymultivalue.cpp:1428).  It never decides results: groups formed by fingerprint are verified byte for byte.
// Two bodies with the same value: keys that start on a 4-byte boundary (every key of a page with keyalign >= 4) are
// read as whole words and the bytes past the key are masked off -- no branch depends on the key LENGTH beyond the
// 12-byte rounds, which matters on a GPU: the byte-wise tail takes a different path for every length and a warp of
// words of 11 different lengths ran it 11 times over (ncu: 1-4 active threads per instruction, 4000 thread
// instructions per word).  Whole-word reads stay inside the record: the key field is padded to a multiple of 4.
__host__ __device__ __forceinline__ unsigned long long key_fingerprint(const uint8_t *k, uint32_t length, uint32_t seed)
{
  uint32_t a, b, c;
  a = b = c = 0xdeadbeefu + length + seed;
  if ((((uintptr_t) k) & 3) == 0) {
    const uint32_t *w = (const uint32_t *) k;
    uint32_t rem = length;
    while (rem > 12) {
      a += w[0];
      b += w[1];
      c += w[2];
      MR_MIX(a, b, c);
      w += 3;
      rem -= 12;
    }
    if (rem) {
      uint32_t w0 = w[0], w1 = rem > 4 ? w[1] : 0u, w2 = rem > 8 ? w[2] : 0u;
      const uint32_t tailbytes = rem & 3u;
      const uint32_t m = tailbytes ? (0xFFFFFFFFu >> (32 - 8 * tailbytes)) : 0xFFFFFFFFu;
      w0 &= rem <= 4 ? m : 0xFFFFFFFFu;
      w1 &= (rem > 4 && rem <= 8) ? m : 0xFFFFFFFFu;
      w2 &= rem > 8 ? m : 0xFFFFFFFFu;
      a += w0;
      b += w1;
      c += w2;
      MR_FINAL(a, b, c);
    }
  } else {
    while (length > 12) {
      a += le_bytes(k, 4);
      b += le_bytes(k + 4, 4);
      c += le_bytes(k + 8, 4);
      MR_MIX(a, b, c);
      length -= 12;
      k += 12;
    }
    if (length) {
      int la = length < 4 ? (int) length : 4;
      int lb = length < 4 ? 0 : (length < 8 ? (int) length - 4 : 4);
      int lc = length < 8 ? 0 : (int) length - 8;
      a += le_bytes(k, la);
      if (lb) b += le_bytes(k + 4, lb);
      if (lc) c += le_bytes(k + 8, lc);
      MR_FINAL(a, b, c);
    }
  }
  return ((unsigned long long) c << 32) | b;
}

}  // namespace mrgpu

// The tokenizer's 16-bytes-at-a-time arithmetic (ops_text.cuh): whitespace and NUL masks by SWAR compares, word starts
// and ends as bit operations on the masks.  Plain C apart from the qualifiers, in a header of its own so that a host
// program can compare it with a byte loop (tests/cpp/tok_host_test.cu).
#pragma once

#include <stdint.h>

__host__ __device__ __forceinline__ bool tok_ws(uint8_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\f' || c == '\r'; }

// bit i set iff byte i of the 16 is one of " \t\n\f\r"; bytes at or after `end` count as whitespace
__host__ __device__ __forceinline__ uint32_t tok_ws_mask16(const uint8_t *__restrict__ t, uint64_t p, uint64_t end, uint32_t *nul = nullptr)
{
  if (nul) *nul = 0;
  if (p >= end) return 0xFFFFu;
  uint32_t w[4] = {0, 0, 0, 0};
  if (end - p >= 16) {
    const uint4 v = *reinterpret_cast<const uint4 *>(t + p);
    w[0] = v.x, w[1] = v.y, w[2] = v.z, w[3] = v.w;
  } else  // the last, partial vector: byte loads, never past the text
    for (uint32_t i = 0; i < (uint32_t) (end - p); i++) w[i >> 2] |= (uint32_t) t[p + i] << (8 * (i & 3));
  uint32_t m = 0;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    // zero-byte tests on x ^ c for the five delimiters (exact form: the 7 low bits are tested without borrow,
    // then the high bit)
    uint32_t hit = 0;
    const uint32_t delim[5] = {0x20202020u, 0x09090909u, 0x0A0A0A0Au, 0x0C0C0C0Cu, 0x0D0D0D0Du};
#pragma unroll
    for (int d = 0; d < 5; d++) {
      const uint32_t x = w[q] ^ delim[d];
      hit |= ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;  // 0x80 in every byte of x that is zero
    }
    // gather the four flag bits (bit 7 of each byte) into a nibble
    m |= (((hit >> 7) * 0x00204081u) >> 21 & 0xFu) << (4 * q);
    if (nul) {  // the same test for NUL bytes (strtok's text ends at the first one)
      const uint32_t z = ~(((w[q] & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | w[q]) & 0x80808080u;
      *nul |= (((z >> 7) * 0x00204081u) >> 21 & 0xFu) << (4 * q);
    }
  }
  if (end - p < 16) {
    m |= 0xFFFFu << (uint32_t) (end - p);
    if (nul) *nul &= ~(0xFFFFu << (uint32_t) (end - p));
  }
  return m & 0xFFFFu;
}

__host__ __device__ __forceinline__ void tokv_masks(const uint8_t *__restrict__ t, uint64_t p, uint64_t end, uint32_t &S, uint32_t &E, uint32_t *nul = nullptr)
{
  const uint32_t m = tok_ws_mask16(t, p, end, nul);
  const uint32_t prev = (p == 0 || p > end) ? 1u : (tok_ws(t[p - 1]) ? 1u : 0u);
  const uint32_t next = (p + 16 >= end) ? 1u : (tok_ws(t[p + 16]) ? 1u : 0u);
  const uint32_t nm = ~m & 0xFFFFu;
  S = nm & ((m << 1) | prev);
  E = nm & ((m >> 1) | (next << 15));
}


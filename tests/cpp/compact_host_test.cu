// Host-only test of the compact sort keys of 16-byte keys (gpu_mapreduce_b200/csrc/compact_key.cuh: the functions the
// extract pass and the factored-key writer call): for random sets of up to 8 varying byte positions and random keys
// that differ only there, (1) the sort key holds exactly those bytes, (2) the key comes back from its sort key and
// the common bytes, (3) sort keys compare like the keys compared as 128-bit little-endian integers -- the order the
// merge of an iterative job's rounds relies on.  Built and run by tests/test_host_logic.py; nothing runs on a device.
#include <cuda_runtime.h>

#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../gpu_mapreduce_b200/csrc/compact_key.cuh"

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

static uint64_t rng = 0x853C49E6748FEA9Bull;
static uint64_t xs()
{
  rng ^= rng << 13;
  rng ^= rng >> 7;
  rng ^= rng << 17;
  return rng;
}

static int cmp128(const uint8_t *a, const uint8_t *b)  // as little-endian integers: byte 15 is the most significant
{
  for (int i = 15; i >= 0; i--)
    if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1;
  return 0;
}

int main()
{
  long n = 0;
  for (int trial = 0; trial < 20000; trial++) {
    CompactKey ck;
    ck.nvb = (int) (xs() % 9);
    bool used[16] = {false};
    for (int q = 0; q < ck.nvb;) {
      const int p = (int) (xs() % 16);
      if (!used[p]) used[p] = true, q++;
    }
    int q = 0;
    for (int p = 0; p < 16; p++)
      if (used[p]) ck.pos[q++] = (unsigned char) p;  // ascending, as convert() lists them
    for (; q < 8; q++) ck.pos[q] = 0;
    uint8_t key0[16], a[16], b[16];
    for (int i = 0; i < 16; i++) key0[i] = (uint8_t) xs();
    memcpy(a, key0, 16);
    memcpy(b, key0, 16);
    for (int i = 0; i < ck.nvb; i++) {
      a[ck.pos[i]] = (uint8_t) (xs() % 3 ? xs() : key0[ck.pos[i]]);
      b[ck.pos[i]] = (uint8_t) (xs() % 3 ? xs() : a[ck.pos[i]]);
    }
    unsigned long long alo, ahi, blo, bhi, c0[2];
    memcpy(&alo, a, 8), memcpy(&ahi, a + 8, 8), memcpy(&blo, b, 8), memcpy(&bhi, b + 8, 8), memcpy(c0, key0, 16);
    const unsigned long long ka = compact_of_key16(alo, ahi, ck), kb = compact_of_key16(blo, bhi, ck);
    for (int i = 0; i < 8; i++) CHECK(((ka >> (8 * i)) & 0xFF) == (i < ck.nvb ? a[ck.pos[i]] : 0u));  // (1)
    key16_clear_varying(c0[0], c0[1], ck);
    unsigned long long rlo, rhi;
    key16_of_compact(ka, c0[0], c0[1], ck, rlo, rhi);
    CHECK(rlo == alo && rhi == ahi);  // (2)
    const int want = cmp128(a, b);
    CHECK((ka < kb ? -1 : (ka > kb ? 1 : 0)) == want);  // (3)
    n++;
  }
  printf("COMPACT_HOST_OK %ld\n", n);
  return 0;
}

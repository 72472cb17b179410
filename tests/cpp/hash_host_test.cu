// Host-only test of the library's lookup3 restatement (gpu_mapreduce_b200/csrc/hash.cuh: __host__ __device__ code, the
// very functions the partition and convert kernels call) against known answers recorded from the reference's
// src/hash.cpp (tests/golden, passed in as "initval hexkey hash" lines).  Built and run by tests/test_host_logic.py.
#include <cuda_runtime.h>

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../gpu_mapreduce_b200/csrc/hash.cuh"

using namespace mrgpu;

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

int main(int argc, char **argv)
{
  CHECK(argc == 2);
  FILE *fp = fopen(argv[1], "r");
  CHECK(fp);
  static char hex[1 << 16];
  alignas(16) static uint8_t buf[1 << 15];
  unsigned iv, want;
  int n = 0;
  while (fscanf(fp, "%u %65535s %u", &iv, hex, &want) == 3) {
    const size_t len = strcmp(hex, "-") ? strlen(hex) / 2 : 0;
    for (int off = 0; off < 4; off++) {  // every alignment of the key's first byte
      uint8_t *k = buf + 16 + off;
      for (size_t i = 0; i < len; i++) {
        unsigned b;
        sscanf(hex + 2 * i, "%2x", &b);
        k[i] = (uint8_t) b;
      }
      k[len] = 0xEE, k[len + 1] = 0xEE, k[len + 2] = 0xEE;  // bytes past the key must not matter
      CHECK(hashlittle(k, (uint32_t) len, iv) == want);
      // the grouping fingerprint: its high half is one lookup3 pass' c, i.e. hashlittle itself
      const unsigned long long f = key_fingerprint(k, (uint32_t) len, iv);
      CHECK((uint32_t) (f >> 32) == want);
      // word-wise body (aligned keys) and byte-wise body agree on both halves
      const unsigned long long f0 = key_fingerprint(buf + 16, 0, iv);
      (void) f0;
      if (off) {
        uint8_t *ka = buf + 8192;
        memset(ka, 0, len + 8);  // the word-wise body reads whole words: the key field is zero padded in a page
        memcpy(ka, k, len);
        CHECK(key_fingerprint(ka, (uint32_t) len, iv) == f);
      }
      if (off == 0 && (len == 4 || len == 8 || len == 12)) CHECK(hashlittle_words((const uint32_t *) k, (int) len / 4, iv) == want);
    }
    n++;
  }
  fclose(fp);
  CHECK(n >= 40);
  printf("HASH_HOST_OK %d\n", n);
  return 0;
}

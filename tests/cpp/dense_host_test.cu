// Host-only test of the count kernel's cache addressing (gpu_mapreduce_b200/csrc/dense_hash.cuh) over the WHOLE key
// range the counting convert accepts: every key below 2^26 maps to a slot inside the cache and a tag in 1..2048, and
// (slot, tag) gives the key back -- so an entry that is flushed always adds its count to the right key, and two keys
// can never share (slot, tag).  Built and run by tests/test_host_logic.py; nothing is launched on a device.
#include <cuda_runtime.h>

#include <stdint.h>
#include <stdio.h>

#include <vector>

#include "../../gpu_mapreduce_b200/csrc/dense_hash.cuh"

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

int main()
{
  std::vector<uint32_t> per_slot(DENSE_CACHE, 0);
  uint32_t maxslot = 0;
  for (uint32_t key = 0; key < DENSE_SLOTS; key++) {
    uint32_t slot, tag;
    dense_slot(key, slot, tag);
    CHECK(slot < DENSE_CACHE);
    CHECK(tag >= 1 && tag <= 2048);  // fits the entry's 12 tag bits, 0 stays "empty"
    CHECK(dense_key_of(slot, tag) == key);
    per_slot[slot]++;
    if (slot > maxslot) maxslot = slot;
  }
  CHECK(maxslot == DENSE_CACHE - 1);
  uint32_t lo = ~0u, hi = 0;
  for (uint32_t c : per_slot) {
    if (c < lo) lo = c;
    if (c > hi) hi = c;
  }
  CHECK(lo >= 1170 && hi <= 1171);  // 2^26 / 57344 = 1170.3 keys per slot, evenly
  printf("DENSE_HOST_OK %u..%u keys per slot\n", lo, hi);
  return 0;
}

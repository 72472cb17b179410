// Host-only helper: the library's record-layout arithmetic (common.cuh: kv_rec_size, kmv_rec_size and the offsets inside
// a record) evaluated for cases read from stdin, one result line per case.  tests/test_host_logic.py compares the
// lines with the oracle's restatement of src/keyvalue.cpp:345-352 and src/keymultivalue.cpp:830-838, which is pinned
// against page images recorded from the reference.  Nothing is launched on a device.
#include <cuda_runtime.h>

#include <stdio.h>

#include <algorithm>

#include "../../gpu_mapreduce_b200/csrc/common.cuh"

using namespace mrgpu;

int main()
{
  char what[8];
  int ka, va;
  unsigned long long nv, kb, vb;
  while (scanf("%7s %d %d %llu %llu %llu", what, &ka, &va, &nv, &kb, &vb) == 6) {
    Align al;
    al.k = ka, al.v = va, al.t = std::max(std::max(ka, va), 4);  // src/mapreduce.cpp:3333-3335
    if (what[1] == 'v')  // "kv": nv unused
      printf("%u %u %u\n", kv_key_off(al), kv_val_off(al, (uint32_t) kb), kv_rec_size(al, (uint32_t) kb, (uint32_t) vb));
    else
      printf("%llu %llu %llu\n", (unsigned long long) kmv_key_off(al, nv), (unsigned long long) kmv_val_off(al, nv, (uint32_t) kb),
             (unsigned long long) kmv_rec_size(al, nv, (uint32_t) kb, vb));
  }
  return 0;
}

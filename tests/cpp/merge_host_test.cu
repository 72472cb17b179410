// Host-only test of the merge path arithmetic (gpu_mapreduce_b200/csrc/merge.cuh): the per-tile split search and the
// per-thread steps are __host__ __device__ functions; here they are driven exactly as the two kernels drive them
// (one split per tile of MG_TILE outputs, one diagonal per thread, MG_IPT steps each) and the result is compared with
// std::merge -- duplicates inside and across the runs, empty runs, lengths around tile boundaries.
// Built and run by tests/test_host_logic.py; nothing is launched on a device.
#include <cuda_runtime.h>

#include <stdint.h>
#include <stdio.h>

#include <algorithm>
#include <vector>

#include "../../gpu_mapreduce_b200/csrc/merge.cuh"

using namespace mrgpu;

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

static uint64_t rng = 0x2545F4914F6CDD1Dull;
static uint64_t xs()
{
  rng ^= rng << 13;
  rng ^= rng >> 7;
  rng ^= rng << 17;
  return rng;
}

template <typename K>
static bool one_case(uint64_t na, uint64_t nb, uint64_t range)
{
  std::vector<K> A(na), B(nb);
  for (auto &x : A) x = (K) (xs() % range);
  for (auto &x : B) x = (K) (xs() % range);
  std::sort(A.begin(), A.end());
  std::sort(B.begin(), B.end());
  const uint64_t n = na + nb, ntiles = (n + MG_TILE - 1) / MG_TILE;
  std::vector<uint64_t> split(ntiles + 1);
  for (uint64_t t = 0; t <= ntiles; t++) {  // merge_split_kernel
    uint64_t d = t * MG_TILE;
    if (d > n) d = n;
    split[t] = merge_path<K, uint64_t>(A.data(), na, B.data(), nb, d);
  }
  std::vector<K> out(n), sh(MG_TILE);
  for (uint64_t t = 0; t < ntiles; t++) {  // merge_tiles_kernel, one block
    const uint64_t d0 = t * MG_TILE, d1 = d0 + MG_TILE < n ? d0 + MG_TILE : n;
    const uint64_t a0 = split[t], a1 = split[t + 1], b0 = d0 - a0, b1 = d1 - a1;
    if (a1 < a0 || b1 < b0 || a1 > na || b1 > nb) return false;
    const unsigned la = (unsigned) (a1 - a0), lb = (unsigned) (b1 - b0), m = la + lb;
    if (m != d1 - d0) return false;
    for (unsigned i = 0; i < m; i++) sh[i] = i < la ? A[a0 + i] : B[b0 + (i - la)];
    for (unsigned th = 0; th < MG_THREADS; th++) {
      unsigned d = th * MG_IPT;
      if (d > m) d = m;
      K r[MG_IPT];
      merge_steps<K>(sh.data(), la, sh.data() + la, lb, d, r);
      for (int q = 0; q < MG_IPT; q++)
        if (th * MG_IPT + q < m) out[d0 + th * MG_IPT + q] = r[q];
    }
  }
  std::vector<K> want(n);
  std::merge(A.begin(), A.end(), B.begin(), B.end(), want.begin());
  return out == want;
}

int main()
{
  const uint64_t lens[] = {0, 1, 7, MG_IPT, MG_TILE - 1, MG_TILE, MG_TILE + 1, 3 * MG_TILE + 5, 10000};
  const uint64_t ranges[] = {1, 2, 50, 1000000, ~0ull};
  int n = 0;
  for (uint64_t na : lens)
    for (uint64_t nb : lens)
      for (uint64_t range : ranges) {
        if (na + nb == 0) continue;
        CHECK(one_case<uint64_t>(na, nb, range));
        CHECK(one_case<uint32_t>(na, nb, range > 0xFFFFFFFFull ? 0xFFFFFFFFull : range));
        n += 2;
      }
  // the shape the library meets: one long run and a short sorted rest
  CHECK(one_case<uint64_t>(200000, 3000, 1ull << 40));
  CHECK(one_case<uint64_t>(3000, 200000, 1ull << 40));
  printf("MERGE_HOST_OK %d\n", n + 2);
  return 0;
}

// Host-only test of the R-MAT generator's arithmetic (gpu_mapreduce_b200/csrc/rmat_lcg.cuh) against the C library
// itself: the LCG is glibc's drand48, the k-step jump equals k single steps, and the integer threshold decides
// "drand48() < t" exactly as the double comparison of the reference does (oink/map_rmat_generate.cpp:40-50).
// Built and run by tests/test_host_logic.py; nothing is launched on a device.
#include <cuda_runtime.h>

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../gpu_mapreduce_b200/csrc/rmat_lcg.cuh"

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

static uint64_t rng = 0x9E3779B97F4A7C15ull;
static uint64_t xs()
{
  rng ^= rng << 13;
  rng ^= rng >> 7;
  rng ^= rng << 17;
  return rng;
}

// the C library's 48-bit state, read the only way there is: seed48() returns the state it replaces
static uint64_t libc_state()
{
  unsigned short zero[3] = {0, 0, 0};
  unsigned short *old = seed48(zero);
  unsigned short cur[3] = {old[0], old[1], old[2]};
  seed48(cur);
  return (uint64_t) cur[0] | ((uint64_t) cur[1] << 16) | ((uint64_t) cur[2] << 32);
}

int main()
{
  // 1. one step of the restated generator = one drand48() of the C library, value and state
  srand48(12345);
  uint64_t x = libc_state();
  CHECK(x == (((uint64_t) 12345 << 16) | 0x330E));
  for (int i = 0; i < 100000; i++) {
    const double want = drand48();
    x = (LCG_A * x + LCG_C) & LCG_M;
    const double got = (double) (long long) x * (1.0 / 281474976710656.0);
    CHECK(got == want);
  }
  CHECK(x == libc_state());

  // 2. the jump: k steps at once, for small k against the loop, for any k by composition
  for (int t = 0; t < 2000; t++) {
    const uint64_t x0 = xs() & LCG_M, k = xs() % 5000;
    unsigned long long a = x0, b = x0;
    for (uint64_t i = 0; i < k; i++) a = (LCG_A * a + LCG_C) & LCG_M;
    lcg_jump(k, b);
    CHECK(a == b);
  }
  for (int t = 0; t < 20000; t++) {
    const uint64_t x0 = xs() & LCG_M, k1 = xs() >> (xs() % 40), k2 = xs() >> (xs() % 40);
    unsigned long long a = x0, b = x0;
    lcg_jump(k1, a);
    lcg_jump(k2, a);
    lcg_jump(k1 + k2, b);  // (mod 2^64 in k: the generator's period 2^48 divides it)
    CHECK(a == b);
    unsigned long long ra, rc, c = x0;
    lcg_affine(k1, ra, rc);
    lcg_jump(k1, c);
    CHECK(((ra * x0 + rc) & LCG_M) == c);
  }

  // 3. "drand48() < t"  <=>  "X < rmat_threshold(t)", for thresholds of every kind and states around the boundary
  double ts[64];
  int nt = 0;
  const double fixed[] = {0.57, 0.19, 0.57 + 0.19, (0.57 + 0.19) + 0.19, 0.05, 0.25, 0.5, 0.75, 1.0, 0.0, -0.3, 1.5, 1e-15, 1e-300, 1.0 - 1e-16,
                          3.5527136788005009e-15 /* 2^-48 */, 7.1054273576010019e-15 /* 2^-47 */, 0.45, 0.45 + 0.15, (0.45 + 0.15) + 0.15};
  for (double f : fixed) ts[nt++] = f;
  while (nt < 64) ts[nt++] = (double) (xs() >> 11) * (1.0 / 9007199254740992.0);
  for (int i = 0; i < nt; i++) {
    const double t = ts[i];
    const unsigned long long T = rmat_threshold(t);
    CHECK(T <= (1ull << 48));
    for (int j = 0; j < 4000; j++) {
      uint64_t X;
      if (j < 64) X = (T + (uint64_t) j - 32) & LCG_M;                 // around the threshold
      else if (j < 80) X = j - 64;                                     // smallest states
      else if (j < 96) X = LCG_M - (uint64_t) (j - 80);                // largest states
      else X = xs() & LCG_M;
      const double r = (double) (long long) X * (1.0 / 281474976710656.0);
      CHECK((r < t) == (X < T));
    }
  }
  const double nan = strtod("nan", nullptr);
  CHECK(rmat_threshold(nan) == 0);  // r < NaN is false for every r
  printf("RMAT_HOST_OK\n");
  return 0;
}

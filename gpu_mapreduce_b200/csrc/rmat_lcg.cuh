// glibc's drand48 generator restated (X' = (0x5DEECE66D X + 0xB) mod 2^48, value = X' / 2^48), its k-step jump, and the
// integer form of the test "drand48() < t".  Kept in a header of their own, free of anything else in the library, so
// that a host program can check them against the C library itself (tests/cpp/rmat_host_test.cu).  Used by the R-MAT
// generator in ops_map.cuh.
#pragma once

#include <math.h>

constexpr unsigned long long LCG_A = 0x5DEECE66Dull, LCG_C = 0xBull, LCG_M = (1ull << 48) - 1;

// the k-step jump as an affine map: x_k = ra * x_0 + rc (mod 2^48)
__host__ __device__ inline void lcg_affine(unsigned long long k, unsigned long long &ra, unsigned long long &rc)
{
  unsigned long long ba = LCG_A, bc = LCG_C;
  ra = 1, rc = 0;
  while (k) {
    if (k & 1) {
      ra = (ba * ra) & LCG_M;
      rc = (ba * rc + bc) & LCG_M;
    }
    bc = (ba * bc + bc) & LCG_M;
    ba = (ba * ba) & LCG_M;
    k >>= 1;
  }
}

__host__ __device__ inline void lcg_jump(unsigned long long k, unsigned long long &x)
{
  unsigned long long ra, rc;
  lcg_affine(k, ra, rc);
  x = (ra * x + rc) & LCG_M;
}

// "drand48() < t" decided on the generator's integer state: drand48() is X / 2^48 exactly (X < 2^48 fits a double's
// mantissa) and scaling by a power of two is exact, so X / 2^48 < t  <=>  X < t * 2^48  <=>  X < ceil(t * 2^48) for an
// integer X.  The ceiling, clamped to [0, 2^48]; a NaN compares false with everything, like the double test.
static inline unsigned long long rmat_threshold(double t)
{
  const double s = ceil(t * 281474976710656.0);
  if (!(s > 0.0)) return 0;
  if (s >= 281474976710656.0) return 1ull << 48;
  return (unsigned long long) s;
}


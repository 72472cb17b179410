// Host-only test of sort_keys' order-preserving transforms (gpu_mapreduce_b200/csrc/sort_order.cuh) against the C
// comparisons the reference's compare functions are made of (src/mapreduce.cpp:2692-2802: a < b -> -1, a > b -> 1,
// else 0 for int / uint64_t / float / double; strcmp for strings): the transformed values must compare the same way.
// Built and run by tests/test_host_logic.py; nothing runs on a device.
#include <cuda_runtime.h>

#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../gpu_mapreduce_b200/csrc/sort_order.cuh"

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

static uint64_t rng = 0xC2B2AE3D27D4EB4Full;
static uint64_t xs()
{
  rng ^= rng << 13;
  rng ^= rng >> 7;
  rng ^= rng << 17;
  return rng;
}
template <typename T>
static int sgn3(T a, T b) { return a < b ? -1 : (a > b ? 1 : 0); }

int main()
{
  // floats and doubles: random bit patterns (NaNs left out: the reference's compare calls a NaN equal to everything,
  // which gives qsort no order to preserve) plus the special values
  const float fs[] = {0.0f, -0.0f, 1.0f, -1.0f, INFINITY, -INFINITY, 1e-45f, -1e-45f, 3.4e38f, -3.4e38f, 1.17549435e-38f};
  const double ds[] = {0.0, -0.0, 1.0, -1.0, INFINITY, -INFINITY, 4.9e-324, -4.9e-324, 1.7e308, -1.7e308, 2.2250738585072014e-308};
  for (int t = 0; t < 400000; t++) {
    float a, b;
    uint32_t ua = (uint32_t) xs(), ub = (uint32_t) xs();
    if (t % 7 == 0) a = fs[xs() % 11], memcpy(&ua, &a, 4);
    if (t % 5 == 0) b = fs[xs() % 11], memcpy(&ub, &b, 4);
    if (t % 11 == 0) ub = ua ^ (uint32_t) (1u << (xs() % 32));
    memcpy(&a, &ua, 4), memcpy(&b, &ub, 4);
    if (isnan(a) || isnan(b)) continue;
    CHECK(sgn3(f32_order(ua), f32_order(ub)) == sgn3(a, b));
    double x, y;
    uint64_t ux = xs(), uy = xs();
    if (t % 7 == 0) x = ds[xs() % 11], memcpy(&ux, &x, 8);
    if (t % 5 == 0) y = ds[xs() % 11], memcpy(&uy, &y, 8);
    if (t % 11 == 0) uy = ux ^ (1ull << (xs() % 64));
    memcpy(&x, &ux, 8), memcpy(&y, &uy, 8);
    if (isnan(x) || isnan(y)) continue;
    CHECK(sgn3(f64_order(ux), f64_order(uy)) == sgn3(x, y));
  }
  // strings: keys of up to 40 bytes over a small alphabet incl. bytes >= 0x80, with their NUL somewhere (or nowhere:
  // then strcmp would run past the key, so such keys get one appended for the C side, which is what "the bytes of the
  // key, then nothing" means)
  for (int t = 0; t < 200000; t++) {
    uint8_t a[48], b[48];
    memset(a, 0, sizeof a), memset(b, 0, sizeof b);
    const uint32_t ka = 1 + (uint32_t) (xs() % 40), kb = 1 + (uint32_t) (xs() % 40);
    const uint8_t alpha[] = {'a', 'b', 'c', 0x80, 0xFF, 0x01, 'z'};
    for (uint32_t i = 0; i < ka; i++) a[i] = alpha[xs() % 7];
    for (uint32_t i = 0; i < kb; i++) b[i] = (xs() % 4) ? a[i] ? a[i] : alpha[xs() % 7] : alpha[xs() % 7];
    if (xs() % 2) a[xs() % ka] = 0;
    if (xs() % 2) b[xs() % kb] = 0;
    const uint32_t nch = (40 + 7) / 8;
    int got = 0;
    for (uint32_t c = 0; c < nch && !got; c++) got = sgn3(str_chunk(a, ka, c), str_chunk(b, kb, c));
    uint8_t ca[48], cb[48];
    memcpy(ca, a, 48), memcpy(cb, b, 48);
    ca[ka] = 0, cb[kb] = 0;
    const int w = strcmp((const char *) ca, (const char *) cb);
    CHECK(got == (w < 0 ? -1 : (w > 0 ? 1 : 0)));
  }
  printf("SORTORDER_HOST_OK\n");
  return 0;
}

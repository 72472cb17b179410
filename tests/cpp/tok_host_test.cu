// Host-only test of the tokenizer's SWAR arithmetic (gpu_mapreduce_b200/csrc/tok_swar.cuh: the functions the device
// passes call for every 16 bytes of text) against the definitions spelled out byte by byte: whitespace mask, NUL mask,
// word-start and word-end bits, for random text with every delimiter, NUL and high-bit bytes, at every text length
// around the vector boundaries.  Built and run by tests/test_host_logic.py; nothing is launched on a device.
#include <cuda_runtime.h>

#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../gpu_mapreduce_b200/csrc/tok_swar.cuh"

#define CHECK(c)                                                    \
  do {                                                              \
    if (!(c)) {                                                     \
      printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c);         \
      return 1;                                                     \
    }                                                               \
  } while (0)

static uint64_t rng = 0xD1B54A32D192ED03ull;
static uint64_t xs()
{
  rng ^= rng << 13;
  rng ^= rng >> 7;
  rng ^= rng << 17;
  return rng;
}

static bool ws(uint8_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\f' || c == '\r'; }

int main()
{
  alignas(16) static uint8_t t[256 + 32];
  const uint8_t pool[] = {' ', '\t', '\n', '\f', '\r', 0, 'a', 'b', 'Z', '0', 0x7F, 0x80, 0xA0, 0x8A, 0xFF, 0x1F, 0x21, 0x0B, 0x0E, 0x08};
  long checked = 0;
  for (int trial = 0; trial < 3000; trial++) {
    const int density = 1 + (int) (xs() % 6);
    for (int i = 0; i < 256; i++) t[i] = (xs() % density) ? pool[6 + xs() % (sizeof(pool) - 6)] : pool[xs() % 6];
    const uint64_t ends[] = {0, 1, 15, 16, 17, 31, 32, 33, 100, 255, 256, xs() % 257, xs() % 257};
    for (uint64_t end : ends)
      for (uint64_t p = 0; p < 256; p += 16) {
        uint32_t nul = 0xDEAD, S, E;
        const uint32_t m = tok_ws_mask16(t, p, end, &nul);
        uint32_t wm = 0, wn = 0, ws_ = 0, we = 0;
        for (int b = 0; b < 16; b++) {
          const uint64_t i = p + b;
          const bool in = i < end;
          if (!in || ws(t[i])) wm |= 1u << b;
          if (in && t[i] == 0) wn |= 1u << b;
          if (in && !ws(t[i]) && (i == 0 || ws(t[i - 1]))) ws_ |= 1u << b;
          if (in && !ws(t[i]) && (i + 1 >= end || ws(t[i + 1]))) we |= 1u << b;
        }
        CHECK(m == wm);
        CHECK(nul == wn);
        CHECK(tok_ws_mask16(t, p, end) == wm);  // without the NUL mask
        if (p <= end) {  // the kernels only ask for vectors that start inside the text (or right at its end)
          tokv_masks(t, p, end, S, E);
          CHECK(S == ws_);
          CHECK(E == we);
        }
        checked++;
      }
  }
  printf("TOK_HOST_OK %ld vectors\n", checked);
  return 0;
}

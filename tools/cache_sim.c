// Simulation of the count kernel's per-block cache on one block's share of the IntCount stream (Zipf(1.1) over 2^24
// keys, 2^30 / 148 keys per block, rounds of < 2^20 keys): which fraction of the pairs ends as a RED to the table.
// Policies: direct-mapped (what ops_convert_dense.cuh does) and two-choice variants.  Development tool, not built by
// the package:  gcc -O2 -o tools/_bin/cache_sim tools/cache_sim.c -lm && tools/_bin/cache_sim [1 = scramble rank -> key]
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define NKEYS (1u << 24)
#define CACHE 57344u
static double *cdf;
static uint64_t rng = 88172645463325252ull;
static inline uint64_t xs(void) { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return rng; }
static uint32_t zipf(void)
{
  double u = (xs() >> 11) * (1.0 / 9007199254740992.0) * cdf[NKEYS - 1];
  uint32_t lo = 0, hi = NKEYS - 1;
  while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (cdf[mid] < u) lo = mid + 1; else hi = mid; }
  return lo;
}
static inline uint32_t slot1(uint32_t key) { uint32_t h = (key * 0x278DC01u) & ((1u << 26) - 1); return (uint32_t) (((uint64_t) (h << 6) * CACHE) >> 32); }
static inline uint32_t slot2(uint32_t key) { uint32_t h = (key * 0x3C6EF35u + 0x1234567u) & ((1u << 26) - 1); return (uint32_t) (((uint64_t) (h << 6) * CACHE) >> 32); }

int main(int argc, char **argv)
{
  const uint64_t n = (1ull << 30) / 148, round = 1000000;
  cdf = malloc(sizeof(double) * NKEYS);
  double acc = 0;
  for (uint32_t r = 0; r < NKEYS; r++) { acc += pow(r + 1.0, -1.1); cdf[r] = acc; }
  uint32_t *keys = malloc(n * 4);
  // scramble rank -> key id so that slots are not correlated with rank
  const int scramble = argc > 1 && atoi(argv[1]);
  for (uint64_t i = 0; i < n; i++) { uint32_t r = zipf(); keys[i] = scramble ? (uint32_t) (((uint64_t) r * 0x9E3779B1u) & (NKEYS - 1)) : r; }
  uint32_t *tag = malloc(CACHE * 4), *cnt = malloc(CACHE * 4);
  // direct-mapped, as in the kernel.  mode 0: every round ends with a full flush (each non-empty entry costs one RED),
  // entries that saw fewer than `keep` pairs are evicted.  mode 1: a round only evicts (and flushes) those entries, and
  // flushes the ones whose count could overflow before the next sweep; everything else keeps counting until the end.
  for (int mode = 0; mode < 2; mode++)
    for (int keep = 1; keep <= 4; keep++)
      for (uint64_t rnd = 1u << 18; rnd <= (1u << 20); rnd <<= 1) {
        if (mode == 0 && rnd != (1u << 20) && rnd != (1u << 19)) continue;
        memset(tag, 0xFF, CACHE * 4);
        memset(cnt, 0, CACHE * 4);
        uint64_t miss = 0, flush = 0;
        for (uint64_t i = 0; i < n; i++) {
          const uint32_t k = keys[i], a = slot1(k);
          if (tag[a] == k) cnt[a]++;
          else if (tag[a] == 0xFFFFFFFFu) { tag[a] = k; cnt[a] = 1; }
          else miss++;
          if ((i + 1) % rnd == 0 || i + 1 == n)
            for (uint32_t s = 0; s < CACHE; s++) {
              const int last = i + 1 == n;
              if (!cnt[s]) continue;
              if (mode == 0 || last) { flush++; if (cnt[s] < (uint32_t) keep) tag[s] = 0xFFFFFFFFu; cnt[s] = 0; }
              else if (cnt[s] < (uint32_t) keep) { flush++; tag[s] = 0xFFFFFFFFu; cnt[s] = 0; }
              else if (cnt[s] + rnd >= (1u << 20)) { flush++; cnt[s] = 0; }
              // mode 1 keeps the count: an entry is evicted only if it saw < keep pairs since it was claimed... and a kept
              // entry's count keeps growing, so it is never evicted later (it stays above keep)
            }
        }
        printf("mode %d keep %d round %7llu: misses %.4f + flush REDs %.4f = %.4f of the pairs\n", mode, keep, (unsigned long long) rnd, (double) miss / n,
               (double) flush / n, (double) (miss + flush) / n);
      }
  // lane-group-private caches: the warp's lanes are split into G groups, each with CACHE / G entries in its own banks
  // (fewer bank conflicts per instruction, smaller caches): RED fraction of one group's share of the block's stream
  for (uint32_t G = 1; G <= 32; G <<= 1) {
    const uint32_t C = CACHE / G;
    const uint64_t ng = n / G, rnd = (1u << 19) / G;
    memset(tag, 0xFF, CACHE * 4);
    memset(cnt, 0, CACHE * 4);
    uint64_t miss = 0, flush = 0;
    for (uint64_t i = 0; i < ng; i++) {
      const uint32_t k = keys[i * G];
      const uint32_t h = (k * 0x278DC01u) & ((1u << 26) - 1);
      const uint32_t a = (uint32_t) (((uint64_t) (h << 6) * C) >> 32);
      if (tag[a] == k) cnt[a]++;
      else if (tag[a] == 0xFFFFFFFFu) { tag[a] = k; cnt[a] = 1; }
      else miss++;
      if ((i + 1) % rnd == 0 || i + 1 == ng)
        for (uint32_t s2 = 0; s2 < C; s2++) {
          if (!cnt[s2]) continue;
          if (i + 1 == ng) { flush++; cnt[s2] = 0; }
          else if (cnt[s2] < 2) { flush++; tag[s2] = 0xFFFFFFFFu; cnt[s2] = 0; }
        }
    }
    printf("groups %2u (%5u entries each): misses %.4f + flush REDs %.4f = %.4f of the pairs\n", G, C, (double) miss / ng, (double) flush / ng,
           (double) (miss + flush) / ng);
  }
  return 0;
}

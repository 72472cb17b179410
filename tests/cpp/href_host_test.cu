// Host-only helper: walk a text (file named on the command line) with the library's href automaton
// (gpu_mapreduce_b200/csrc/href_dfa.cuh -- the transition function the device table is built from) and print one
// "start end" line per URL: a URL starts after the quote that takes HREF to LINK and ends at the quote that takes LINK
// back to START, exactly what the device passes extract.  tests/test_host_logic.py compares the spans with the
// reference's own scanner (the oracle's href(), whose body is cut out of cpu/InvertedIndex.cpp at build time).
#include <stdio.h>
#include <stdlib.h>

#include "../../gpu_mapreduce_b200/csrc/href_dfa.cuh"

int main(int argc, char **argv)
{
  if (argc != 2) return 2;
  FILE *fp = fopen(argv[1], "rb");
  if (!fp) return 2;
  fseek(fp, 0, SEEK_END);
  const long n = ftell(fp);
  fseek(fp, 0, SEEK_SET);
  unsigned char *t = (unsigned char *) malloc(n ? n : 1);
  if (fread(t, 1, n, fp) != (size_t) n) return 2;
  fclose(fp);
  int st = HS_START;
  long start = -1;
  for (long i = 0; i < n; i++) {
    const int nx = href_step(st, t[i]);
    if (st == HS_HREF && nx == HS_LINK) start = i + 1;
    if (st == HS_LINK && nx == HS_START) printf("%ld %ld\n", start, i);
    st = nx;
  }
  free(t);
  return 0;
}

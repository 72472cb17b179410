// The InvertedIndex href scanner's automaton (cpu/InvertedIndex.cpp:196-270) as an 8-state DFA without look-ahead:
// the states and the transition function, host code from which ops_text.cuh builds the per-byte table its kernels use.
// In a header of its own so that a host program can walk a text with it and compare the URLs with the reference's own
// scanner (tests/cpp/href_host_test.cu).
#pragma once

enum { HS_START, HS_TAG, HS_ATAG, HS_H1, HS_H2, HS_H3, HS_HREF, HS_LINK };

static inline int href_start_on(int c) { return c == '<' ? HS_TAG : HS_START; }

static inline int href_step(int st, int c)
{
  switch (st) {
    case HS_START: return href_start_on(c);
    case HS_TAG: return c == 'a' ? HS_ATAG : (c == ' ' ? HS_TAG : HS_START);     // cpu/InvertedIndex.cpp:211-221
    case HS_ATAG: return c == 'h' ? HS_H1 : (c == ' ' ? HS_ATAG : HS_START);      // :222-243
    case HS_H1: return c == 'r' ? HS_H2 : href_start_on(c);                       // failed peek: START at the 'h',
    case HS_H2: return c == 'e' ? HS_H3 : href_start_on(c);                       // then these bytes are rescanned
    case HS_H3: return c == 'f' ? HS_HREF : href_start_on(c);                     // from START
    case HS_HREF: return (c == ' ' || c == '=') ? HS_HREF : (c == '"' ? HS_LINK : HS_START);  // :244-253
    default: return c == '"' ? HS_START : HS_LINK;                                // :254-267
  }
}


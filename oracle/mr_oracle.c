/* TEST INFRASTRUCTURE ONLY -- see mr_oracle.h for the contract and the parity status (PINNED against
 * oracle/_ref/refdrv = the unmodified reference library, and tests/golden/ generated from it).
 * Plain C restatement of the reference algorithms; every function cites the reference lines it follows. */
#include "mr_oracle.h"

#include <stdlib.h>
#include <string.h>

void orc_free(void *p) { free(p); }

/* ------------------------------------------------------------------ growable byte buffer */

typedef struct {
  uint8_t *p;
  uint64_t n, cap;
} buf_t;

static void buf_put(buf_t *b, const void *src, uint64_t n)
{
  if (b->n + n > b->cap) {
    uint64_t cap = b->cap ? b->cap : 4096;
    while (cap < b->n + n) cap *= 2;
    b->p = (uint8_t *) realloc(b->p, cap);
    b->cap = cap;
  }
  if (n) memcpy(b->p + b->n, src, n);
  b->n += n;
}

static void buf_zero(buf_t *b, uint64_t n)
{
  static const uint8_t z[64] = {0};
  while (n) {
    uint64_t m = n < 64 ? n : 64;
    buf_put(b, z, m);
    n -= m;
  }
}

static void buf_i32(buf_t *b, int32_t v) { buf_put(b, &v, 4); }
static void buf_u64(buf_t *b, uint64_t v) { buf_put(b, &v, 8); }

static int buf_finish(buf_t *b, uint8_t **out, uint64_t *outbytes)
{
  if (!b->p) b->p = (uint8_t *) malloc(8);
  *out = b->p;
  *outbytes = b->n;
  return 0;
}

/* record view over a kvfile */
typedef struct {
  const uint8_t *key, *val;
  int32_t kb, vb;
} rec_t;

static uint64_t kv_count(const uint8_t *kv)
{
  uint64_t n;
  memcpy(&n, kv, 8);
  return n;
}

static rec_t *kv_index(const uint8_t *kv, uint64_t *n_out)
{
  uint64_t n = kv_count(kv);
  rec_t *r = (rec_t *) malloc((n ? n : 1) * sizeof(rec_t));
  const uint8_t *p = kv + 8;
  for (uint64_t i = 0; i < n; i++) {
    memcpy(&r[i].kb, p, 4);
    memcpy(&r[i].vb, p + 4, 4);
    r[i].key = p + 8;
    r[i].val = r[i].key + r[i].kb;
    p = r[i].val + r[i].vb;
  }
  *n_out = n;
  return r;
}

static void kv_emit(buf_t *b, const void *key, int32_t kb, const void *val, int32_t vb)
{
  buf_i32(b, kb);
  buf_i32(b, vb);
  buf_put(b, key, kb);
  buf_put(b, val, vb);
}

static void kv_begin(buf_t *b)
{
  memset(b, 0, sizeof(*b));
  buf_u64(b, 0);
}

static void kv_end(buf_t *b, uint64_t n) { memcpy(b->p, &n, 8); }

/* ------------------------------------------------------------------ hashlittle
 * src/hash.cpp:129-333.  The reference has three read strategies (aligned 32-bit, aligned 16-bit,
 * bytewise); on a little-endian machine all three produce the same value, so only the
 * bytewise one (src/hash.cpp:283-328) is restated. */

#define ROT(x, k) (((x) << (k)) | ((x) >> (32 - (k))))

#define MIX(a, b, c)     \
  {                      \
    a -= c;              \
    a ^= ROT(c, 4);      \
    c += b;              \
    b -= a;              \
    b ^= ROT(a, 6);      \
    a += c;              \
    c -= b;              \
    c ^= ROT(b, 8);      \
    b += a;              \
    a -= c;              \
    a ^= ROT(c, 16);     \
    c += b;              \
    b -= a;              \
    b ^= ROT(a, 19);     \
    a += c;              \
    c -= b;              \
    c ^= ROT(b, 4);      \
    b += a;              \
  }

#define FINAL(a, b, c)   \
  {                      \
    c ^= b;              \
    c -= ROT(b, 14);     \
    a ^= c;              \
    a -= ROT(c, 11);     \
    b ^= a;              \
    b -= ROT(a, 25);     \
    c ^= b;              \
    c -= ROT(b, 16);     \
    a ^= c;              \
    a -= ROT(c, 4);      \
    b ^= a;              \
    b -= ROT(a, 14);     \
    c ^= b;              \
    c -= ROT(b, 24);     \
  }

uint32_t orc_hashlittle(const void *key, size_t length, uint32_t initval)
{
  const uint8_t *k = (const uint8_t *) key;
  uint32_t a, b, c;
  a = b = c = 0xdeadbeef + ((uint32_t) length) + initval;
  while (length > 12) {
    a += k[0] + ((uint32_t) k[1] << 8) + ((uint32_t) k[2] << 16) + ((uint32_t) k[3] << 24);
    b += k[4] + ((uint32_t) k[5] << 8) + ((uint32_t) k[6] << 16) + ((uint32_t) k[7] << 24);
    c += k[8] + ((uint32_t) k[9] << 8) + ((uint32_t) k[10] << 16) + ((uint32_t) k[11] << 24);
    MIX(a, b, c);
    length -= 12;
    k += 12;
  }
  switch (length) {
    case 12: c += ((uint32_t) k[11]) << 24; /* fall through */
    case 11: c += ((uint32_t) k[10]) << 16; /* fall through */
    case 10: c += ((uint32_t) k[9]) << 8;   /* fall through */
    case 9: c += k[8];                      /* fall through */
    case 8: b += ((uint32_t) k[7]) << 24;   /* fall through */
    case 7: b += ((uint32_t) k[6]) << 16;   /* fall through */
    case 6: b += ((uint32_t) k[5]) << 8;    /* fall through */
    case 5: b += k[4];                      /* fall through */
    case 4: a += ((uint32_t) k[3]) << 24;   /* fall through */
    case 3: a += ((uint32_t) k[2]) << 16;   /* fall through */
    case 2: a += ((uint32_t) k[1]) << 8;    /* fall through */
    case 1: a += k[0]; break;
    case 0: return c;
  }
  FINAL(a, b, c);
  return c;
}

/* src/mapreduce.cpp:469-472 */
int orc_partition(const uint8_t *kv, int nprocs, int32_t *out)
{
  uint64_t n;
  rec_t *r = kv_index(kv, &n);
  for (uint64_t i = 0; i < n; i++)
    out[i] = (int32_t) (orc_hashlittle(r[i].key, r[i].kb, (uint32_t) nprocs) % (uint32_t) nprocs);
  free(r);
  return 0;
}

/* ------------------------------------------------------------------ record layout */

static uint64_t roundup(uint64_t x, int align) { return (x + align - 1) / align * align; }
static int imax(int a, int b) { return a > b ? a : b; }

/* src/keyvalue.cpp:345-352: two ints, pad to kalign, key, pad to valign, value, pad to talign where
 * talign = max(kalign, valign, sizeof(int)) (src/mapreduce.cpp:3333-3335). */
uint64_t orc_kv_record_size(int kb, int vb, int kalign, int valign)
{
  int talign = imax(imax(kalign, valign), 4);
  uint64_t o = roundup(8, kalign) + kb;
  o = roundup(o, valign) + vb;
  return roundup(o, talign);
}

/* src/keymultivalue.cpp:830-838: three ints, nvalue ints, pad, key, pad, values, pad. */
uint64_t orc_kmv_record_size(uint64_t nvalue, int kb, uint64_t mvbytes, int kalign, int valign)
{
  int talign = imax(imax(kalign, valign), 4);
  uint64_t o = roundup(12 + 4 * nvalue, kalign) + kb;
  o = roundup(o, valign) + mvbytes;
  return roundup(o, talign);
}

/* Greedy page fill of KeyValue::add, src/keyvalue.cpp:343-392: a record that does not fit closes the page
 * (:362-380); pages are independent byte arrays whose records start at offset 0. */
int orc_kv_pages(const uint8_t *kv, uint64_t pagesize, int kalign, int valign, uint8_t **out,
                 uint64_t *outbytes)
{
  uint64_t n;
  rec_t *r = kv_index(kv, &n);
  buf_t b;
  memset(&b, 0, sizeof(b));
  buf_u64(&b, 0);
  uint64_t npage = 0, i = 0;
  do {
    uint64_t hdr = b.n;
    buf_i32(&b, 0);
    buf_u64(&b, 0);
    uint64_t start = b.n, nkey = 0;
    while (i < n) {
      uint64_t sz = orc_kv_record_size(r[i].kb, r[i].vb, kalign, valign);
      if ((b.n - start) + sz > pagesize) {
        if (b.n == start) {
          free(r);
          free(b.p);
          return -1; /* "Single key/value pair exceeds page size" src/keyvalue.cpp:369 */
        }
        break;
      }
      uint64_t base = b.n;
      buf_i32(&b, r[i].kb);
      buf_i32(&b, r[i].vb);
      buf_zero(&b, roundup(8, kalign) - 8);
      buf_put(&b, r[i].key, r[i].kb);
      buf_zero(&b, roundup(b.n - base, valign) - (b.n - base));
      buf_put(&b, r[i].val, r[i].vb);
      buf_zero(&b, base + sz - b.n);
      nkey++;
      i++;
    }
    int32_t nk = (int32_t) nkey;
    uint64_t asz = b.n - start;
    memcpy(b.p + hdr, &nk, 4);
    memcpy(b.p + hdr + 4, &asz, 8);
    npage++;
  } while (i < n);
  memcpy(b.p, &npage, 8);
  free(r);
  return buf_finish(&b, out, outbytes);
}

/* ------------------------------------------------------------------ kmvfile walking */

typedef struct {
  const uint8_t *key;
  int32_t kb;
  uint64_t nvalue;
  const uint8_t *values; /* nvalue x { i32 vb ; bytes } */
  const uint8_t *next;
} kmvrec_t;

static kmvrec_t kmv_read(const uint8_t *p)
{
  kmvrec_t k;
  memcpy(&k.kb, p, 4);
  memcpy(&k.nvalue, p + 4, 8);
  k.key = p + 12;
  k.values = k.key + k.kb;
  const uint8_t *q = k.values;
  for (uint64_t i = 0; i < k.nvalue; i++) {
    int32_t vb;
    memcpy(&vb, q, 4);
    q += 4 + vb;
  }
  k.next = q;
  return k;
}

/* Single-page KMV layout: src/keymultivalue.cpp:830-893 (header + key written by unique2kmv_all) and
 * :1139-1209 (kv2kmv appends each value's size to soffset[] and its bytes at voffset in arrival order).
 * Sets (= pages) close when the next pair does not fit (:853-868). */
int orc_kmv_pages(const uint8_t *kmv, uint64_t pagesize, int kalign, int valign, uint8_t **out,
                  uint64_t *outbytes)
{
  uint64_t n = kv_count(kmv);
  const uint8_t *p = kmv + 8;
  buf_t b;
  memset(&b, 0, sizeof(b));
  buf_u64(&b, 0);
  uint64_t npage = 0, i = 0;
  do {
    uint64_t hdr = b.n;
    buf_i32(&b, 0);
    buf_u64(&b, 0);
    uint64_t start = b.n, nkey = 0;
    while (i < n) {
      kmvrec_t k = kmv_read(p);
      uint64_t mvbytes = 0;
      const uint8_t *q = k.values;
      for (uint64_t j = 0; j < k.nvalue; j++) {
        int32_t vb;
        memcpy(&vb, q, 4);
        mvbytes += vb;
        q += 4 + vb;
      }
      uint64_t sz = orc_kmv_record_size(k.nvalue, k.kb, mvbytes, kalign, valign);
      if (sz > pagesize || k.nvalue > 0x7FFFFFFF || mvbytes > 0x7FFFFFFF) {
        free(b.p);
        return -1; /* multi-block pair, src/keymultivalue.cpp:845-847 */
      }
      if ((b.n - start) + sz > pagesize) break;
      uint64_t base = b.n;
      buf_i32(&b, (int32_t) k.nvalue);
      buf_i32(&b, k.kb);
      buf_i32(&b, (int32_t) mvbytes);
      q = k.values;
      for (uint64_t j = 0; j < k.nvalue; j++) {
        buf_put(&b, q, 4);
        int32_t vb;
        memcpy(&vb, q, 4);
        q += 4 + vb;
      }
      buf_zero(&b, roundup(b.n - base, kalign) - (b.n - base));
      buf_put(&b, k.key, k.kb);
      buf_zero(&b, roundup(b.n - base, valign) - (b.n - base));
      q = k.values;
      for (uint64_t j = 0; j < k.nvalue; j++) {
        int32_t vb;
        memcpy(&vb, q, 4);
        buf_put(&b, q + 4, vb);
        q += 4 + vb;
      }
      buf_zero(&b, base + sz - b.n);
      nkey++;
      i++;
      p = k.next;
    }
    int32_t nk = (int32_t) nkey;
    uint64_t asz = b.n - start;
    memcpy(b.p + hdr, &nk, 4);
    memcpy(b.p + hdr + 4, &asz, 8);
    npage++;
  } while (i < n);
  memcpy(b.p, &npage, 8);
  return buf_finish(&b, out, outbytes);
}

/* ------------------------------------------------------------------ convert
 * src/keymultivalue.cpp:486-638.  Pass 1 = kv2unique (:645-789): bucket = hashlittle(key,kb,0) & mask
 * (:1428-1433), chain walk with length + memcmp match (find, :1403-1422); a miss appends a Unique in
 * first-appearance order, a hit bumps nvalue / mvbytes.  Pass 2 = kv2kmv (:1139-1209): every pair is
 * re-hashed, found, and its value appended, so values keep arrival order. */

int orc_convert(const uint8_t *kv, uint8_t **kmv_out, uint64_t *outbytes)
{
  uint64_t n;
  rec_t *r = kv_index(kv, &n);
  uint64_t nb = 1024;
  while (nb < 2 * n) nb <<= 1;
  int64_t *bucket = (int64_t *) malloc(nb * sizeof(int64_t));
  for (uint64_t i = 0; i < nb; i++) bucket[i] = -1;
  uint64_t *uid = (uint64_t *) malloc((n ? n : 1) * sizeof(uint64_t));
  /* per unique: first record, chain link, nvalue, mvbytes */
  uint64_t *ufirst = (uint64_t *) malloc((n ? n : 1) * sizeof(uint64_t));
  int64_t *unext = (int64_t *) malloc((n ? n : 1) * sizeof(int64_t));
  uint64_t *unv = (uint64_t *) malloc((n ? n : 1) * sizeof(uint64_t));
  uint64_t *umv = (uint64_t *) malloc((n ? n : 1) * sizeof(uint64_t));
  uint64_t nu = 0;
  for (uint64_t i = 0; i < n; i++) {
    uint64_t ib = orc_hashlittle(r[i].key, r[i].kb, 0) & (nb - 1);
    int64_t u = bucket[ib], last = -1;
    while (u >= 0) {
      const rec_t *f = &r[ufirst[u]];
      if (f->kb == r[i].kb && memcmp(f->key, r[i].key, r[i].kb) == 0) break;
      last = u;
      u = unext[u];
    }
    if (u < 0) {
      u = (int64_t) nu++;
      ufirst[u] = i;
      unext[u] = -1;
      unv[u] = 0;
      umv[u] = 0;
      if (last < 0) bucket[ib] = u;
      else unext[last] = u;
    }
    unv[u]++;
    umv[u] += r[i].vb;
    uid[i] = (uint64_t) u;
  }
  /* offsets of each unique's region in the kmvfile */
  uint64_t *off = (uint64_t *) malloc((nu + 1) * sizeof(uint64_t));
  uint64_t total = 8;
  for (uint64_t u = 0; u < nu; u++) {
    off[u] = total;
    total += 12 + r[ufirst[u]].kb + 4 * unv[u] + umv[u];
  }
  uint8_t *o = (uint8_t *) malloc(total);
  memcpy(o, &nu, 8);
  uint64_t *cur = (uint64_t *) malloc((nu ? nu : 1) * sizeof(uint64_t));
  for (uint64_t u = 0; u < nu; u++) {
    const rec_t *f = &r[ufirst[u]];
    memcpy(o + off[u], &f->kb, 4);
    memcpy(o + off[u] + 4, &unv[u], 8);
    memcpy(o + off[u] + 12, f->key, f->kb);
    cur[u] = off[u] + 12 + f->kb;
  }
  for (uint64_t i = 0; i < n; i++) {
    uint64_t u = uid[i];
    memcpy(o + cur[u], &r[i].vb, 4);
    memcpy(o + cur[u] + 4, r[i].val, r[i].vb);
    cur[u] += 4 + r[i].vb;
  }
  free(cur);
  free(off);
  free(umv);
  free(unv);
  free(unext);
  free(ufirst);
  free(uid);
  free(bucket);
  free(r);
  *kmv_out = o;
  *outbytes = total;
  return 0;
}

/* ------------------------------------------------------------------ reduce with the stock callbacks
 * Record walk: src/mapreduce.cpp:1799-1850.  count: oink/reduce_count.cpp:13-20 (total over blocks cast to
 * int); sum: examples/wordfreq.cpp:137-141; cull: oink/reduce_cull.cpp:12-20 (first value). */

int orc_reduce(const uint8_t *kmv, int which, uint8_t **kv_out, uint64_t *outbytes)
{
  uint64_t n = kv_count(kmv);
  const uint8_t *p = kmv + 8;
  buf_t b;
  kv_begin(&b);
  for (uint64_t i = 0; i < n; i++) {
    kmvrec_t k = kmv_read(p);
    if (which == 0 || which == 1) {
      int32_t cnt = (int32_t) k.nvalue;
      kv_emit(&b, k.key, k.kb, &cnt, 4);
    } else {
      int32_t vb;
      memcpy(&vb, k.values, 4);
      kv_emit(&b, k.key, k.kb, k.values + 4, vb);
    }
    p = k.next;
  }
  kv_end(&b, n);
  return buf_finish(&b, kv_out, outbytes);
}

/* ------------------------------------------------------------------ sort_keys
 * src/mapreduce.cpp:2102-2149 -> sort_kv(0) :2359-2453; compare functions :2692-2802. */

static int cmp_flag(int flag, const rec_t *x, const rec_t *y)
{
  int a = flag < 0 ? -flag : flag, c = 0;
  switch (a) {
    case 1: {
      int32_t i, j;
      memcpy(&i, x->key, 4);
      memcpy(&j, y->key, 4);
      c = (i < j) ? -1 : (i > j);
      break;
    }
    case 2: {
      uint64_t i, j;
      memcpy(&i, x->key, 8);
      memcpy(&j, y->key, 8);
      c = (i < j) ? -1 : (i > j);
      break;
    }
    case 3: {
      float i, j;
      memcpy(&i, x->key, 4);
      memcpy(&j, y->key, 4);
      c = (i < j) ? -1 : (i > j);
      break;
    }
    case 4: {
      double i, j;
      memcpy(&i, x->key, 8);
      memcpy(&j, y->key, 8);
      c = (i < j) ? -1 : (i > j);
      break;
    }
    case 5: c = strcmp((const char *) x->key, (const char *) y->key); break;
    case 6: c = strncmp((const char *) x->key, (const char *) y->key, x->kb < y->kb ? x->kb : y->kb); break;
  }
  return flag < 0 ? -c : c;
}

static void msort(rec_t *a, rec_t *tmp, uint64_t lo, uint64_t hi, int flag)
{
  if (hi - lo < 2) return;
  uint64_t mid = lo + (hi - lo) / 2;
  msort(a, tmp, lo, mid, flag);
  msort(a, tmp, mid, hi, flag);
  uint64_t i = lo, j = mid, k = lo;
  while (i < mid && j < hi) tmp[k++] = (cmp_flag(flag, &a[j], &a[i]) < 0) ? a[j++] : a[i++];
  while (i < mid) tmp[k++] = a[i++];
  while (j < hi) tmp[k++] = a[j++];
  memcpy(a + lo, tmp + lo, (hi - lo) * sizeof(rec_t));
}

int orc_sort_keys(const uint8_t *kv, int flag, uint8_t **kv_out, uint64_t *outbytes)
{
  int a = flag < 0 ? -flag : flag;
  if (a < 1 || a > 6) return -1; /* "Invalid compare method for sort keys" :2128 */
  uint64_t n;
  rec_t *r = kv_index(kv, &n);
  rec_t *tmp = (rec_t *) malloc((n ? n : 1) * sizeof(rec_t));
  msort(r, tmp, 0, n, flag);
  buf_t b;
  kv_begin(&b);
  for (uint64_t i = 0; i < n; i++) kv_emit(&b, r[i].key, r[i].kb, r[i].val, r[i].vb);
  kv_end(&b, n);
  free(tmp);
  free(r);
  return buf_finish(&b, kv_out, outbytes);
}

/* ------------------------------------------------------------------ map(mr, edge_to_vertices) */

int orc_map_edge_to_vertices(const uint8_t *kv, int both, uint8_t **kv_out, uint64_t *outbytes)
{
  uint64_t n;
  rec_t *r = kv_index(kv, &n);
  buf_t b;
  kv_begin(&b);
  uint64_t m = 0;
  for (uint64_t i = 0; i < n; i++) { /* record walk: src/mapreduce.cpp:1610-1629 */
    kv_emit(&b, r[i].key, 8, NULL, 0);
    m++;
    if (both) {
      kv_emit(&b, r[i].key + 8, 8, NULL, 0);
      m++;
    }
  }
  kv_end(&b, m);
  free(r);
  return buf_finish(&b, kv_out, outbytes);
}

/* ------------------------------------------------------------------ wordfreq tokenizer
 * examples/wordfreq.cpp:104-130: the file is NUL-terminated after nchar bytes and cut by
 * strtok(text, " \t\n\f\r") -- so the text effectively ends at its first NUL byte; each token is emitted
 * with its terminating NUL (strlen+1) and no value. */

static int is_ws(uint8_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\f' || c == '\r'; }

int orc_tokenize(const uint8_t *text, uint64_t nchar, uint8_t **kv_out, uint64_t *outbytes)
{
  buf_t b;
  kv_begin(&b);
  uint64_t end = 0, m = 0;
  while (end < nchar && text[end]) end++;
  uint64_t i = 0;
  while (i < end) {
    while (i < end && is_ws(text[i])) i++;
    if (i >= end) break;
    uint64_t s = i;
    while (i < end && !is_ws(text[i])) i++;
    buf_i32(&b, (int32_t) (i - s + 1));
    buf_i32(&b, 0);
    buf_put(&b, text + s, i - s);
    buf_zero(&b, 1);
    m++;
  }
  kv_end(&b, m);
  return buf_finish(&b, kv_out, outbytes);
}

/* ------------------------------------------------------------------ href scanner
 * cpu/InvertedIndex.cpp:196-270 as the look-ahead-free 8-state DFA of SURVEY.md B.8.  The reference peeks
 * "href" at the 'h' (:224-236) and jumps 3 bytes; a failed peek drops to START at the 'h' and the
 * following bytes are scanned from START, hence startOn() in H1..H3.  The closing-quote search
 * (:254-262) runs to the NUL the reference appends at text[nchar]; a quote found only there (or beyond)
 * emits nothing.  An embedded NUL inside a link does not stop the reference's search, and does not stop
 * this one. */

enum { S_START, S_TAG, S_ATAG, S_H1, S_H2, S_H3, S_HREF, S_LINK };

static int start_on(uint8_t c) { return c == '<' ? S_TAG : S_START; }

int orc_href(const uint8_t *text, uint64_t nchar, const char *fname, uint8_t **kv_out, uint64_t *outbytes)
{
  buf_t b;
  kv_begin(&b);
  int32_t namelen = (int32_t) strlen(fname) + 1;
  int st = S_START;
  uint64_t link_start = 0, m = 0;
  for (uint64_t j = 0; j < nchar; j++) {
    uint8_t c = text[j];
    switch (st) {
      case S_START: st = start_on(c); break;
      case S_TAG: st = (c == 'a') ? S_ATAG : (c == ' ') ? S_TAG : S_START; break;
      case S_ATAG: st = (c == 'h') ? S_H1 : (c == ' ') ? S_ATAG : S_START; break;
      case S_H1: st = (c == 'r') ? S_H2 : start_on(c); break;
      case S_H2: st = (c == 'e') ? S_H3 : start_on(c); break;
      case S_H3: st = (c == 'f') ? S_HREF : start_on(c); break;
      case S_HREF:
        if (c == ' ' || c == '=') st = S_HREF;
        else if (c == '"') {
          st = S_LINK;
          link_start = j + 1;
        } else st = S_START;
        break;
      case S_LINK:
        if (c == '"') {
          kv_emit(&b, text + link_start, (int32_t) (j - link_start), fname, namelen);
          m++;
          st = S_START;
        }
        break;
    }
  }
  kv_end(&b, m);
  return buf_finish(&b, kv_out, outbytes);
}

/* ------------------------------------------------------------------ drand48 + rmat_generate */

uint64_t orc_srand48(uint32_t seed) { return (((uint64_t) seed) << 16) | 0x330E; }

double orc_drand48(uint64_t *state)
{
  *state = (0x5DEECE66DULL * *state + 0xBULL) & 0xFFFFFFFFFFFFULL;
  return (double) *state / 281474976710656.0; /* exact: 48-bit integer scaled by 2^-48 */
}

/* oink/map_rmat_generate.cpp:14-67 */
void orc_rmat_generate(uint64_t *state, uint64_t ngenerate, int nlevels, double a, double b, double c,
                       double d, double fraction, uint64_t *edges)
{
  uint64_t order = 1ULL << nlevels;
  for (uint64_t m = 0; m < ngenerate; m++) {
    uint64_t delta = order >> 1, i = 0, j = 0;
    double a1 = a, b1 = b, c1 = c, d1 = d;
    for (int ilevel = 0; ilevel < nlevels; ilevel++) {
      double rn = orc_drand48(state);
      if (rn < a1) {
      } else if (rn < a1 + b1) {
        j += delta;
      } else if (rn < a1 + b1 + c1) {
        i += delta;
      } else {
        i += delta;
        j += delta;
      }
      delta /= 2;
      if (fraction > 0.0) {
        a1 += a1 * fraction * (orc_drand48(state) - 0.5);
        b1 += b1 * fraction * (orc_drand48(state) - 0.5);
        c1 += c1 * fraction * (orc_drand48(state) - 0.5);
        d1 += d1 * fraction * (orc_drand48(state) - 0.5);
        double total = a1 + b1 + c1 + d1;
        a1 /= total;
        b1 /= total;
        c1 /= total;
        d1 /= total;
      }
    }
    edges[2 * m] = i;
    edges[2 * m + 1] = j;
  }
}

/* ------------------------------------------------------------------ IntCount emit, cpu/IntCount.cpp:177-180 */

int orc_intcount_map(const uint32_t *keys, uint64_t n, uint8_t **kv_out, uint64_t *outbytes)
{
  uint64_t total = 8 + 16 * n;
  uint8_t *o = (uint8_t *) malloc(total);
  memcpy(o, &n, 8);
  int32_t rec[4] = {4, 4, 0, 1};
  for (uint64_t i = 0; i < n; i++) {
    rec[2] = (int32_t) keys[i];
    memcpy(o + 8 + 16 * i, rec, 16);
  }
  *kv_out = o;
  *outbytes = total;
  return 0;
}

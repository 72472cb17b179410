// Text map functors: the wordfreq tokenizer and the InvertedIndex href scanner.
//
// Both reference callbacks are sequential byte loops on the CPU (strtok in examples/wordfreq.cpp:104-130,
// a 5-state switch with look-ahead in cpu/InvertedIndex.cpp:196-270) that call KeyValue::add per hit.  On
// the GPU each becomes: classify bytes in parallel -> prefix sums give every hit its output slot ->
// coalesced pack into KV records.  The fork's own GPU attempt (cuda/InvertedIndex.cu: `mark` +
// thrust::copy_if + compute_url_length + a host kv->add loop) is NOT what is done here: it matches only the
// literal `<a href="`, writes a 4-byte mask per input byte and packs records on the host.
// Included by mrgpu.cu.
#pragma once

// ---------------------------------------------------------------------------------------------
// shared: append variable-length records to a KV

struct VarAppend {
  uint64_t base;     // byte offset of the first new record
  uint64_t *offs;    // device: offs for the new records (kv->offs + old nkv)
};

static VarAppend kv_append_var_begin(mrgpu_mr *mr, DevKV *kv, uint64_t nnew)
{
  kv_make_general(mr, kv);
  kv->reserve_offs(kv->nkv + nnew + 1);
  kv->h_offs_valid = false;
  return VarAppend{kv->nbytes, kv->offs + kv->nkv};
}

// after offs[0..nnew) hold base + exclusive sizes and `total` bytes were computed
static void kv_append_var_reserve(mrgpu_mr *mr, DevKV *kv, const VarAppend &va, uint64_t nnew, uint64_t total)
{
  uint64_t end = va.base + total;
  MR_CUDA(cudaMemcpyAsync(va.offs + nnew, &end, 8, cudaMemcpyHostToDevice, mr->pool.stream));
  kv->reserve_bytes(end);
  sync(mr);  // `end` is a stack variable
}

// ---------------------------------------------------------------------------------------------
// wordfreq tokenizer, examples/wordfreq.cpp:104-130
//   text[nchar] = 0; strtok(text, " \t\n\f\r") ...  kv->add(word, strlen(word)+1, NULL, 0)
// strtok stops at the first NUL, so the effective text is [0, first NUL).

#include "tok_swar.cuh"  // tok_ws, tok_ws_mask16, tokv_masks: host-callable, checked against a byte loop on the host

__global__ void first_nul_kernel(const uint8_t *__restrict__ t, uint64_t first, uint64_t n, unsigned long long *pos)
{
  uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  unsigned long long best = ~0ull;
  for (uint64_t i = first + (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    if (t[i] == 0) {
      best = i;
      break;
    }
  for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
  if ((threadIdx.x & 31) == 0 && best != ~0ull) atomicMin(pos, best);
}

__global__ void __launch_bounds__(256)
tokenize_write_kernel(const uint8_t *__restrict__ t, const uint32_t *__restrict__ starts, const uint32_t *__restrict__ ends, uint64_t nw,
                      uint8_t *__restrict__ out, const uint64_t *__restrict__ offs, Align al)
{
  uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nw) return;
  uint32_t s = starts[j], len = ends[j] - s, kb = len + 1;
  uint8_t *r = out + offs[j];
  uint32_t ko = kv_key_off(al), end = kv_rec_size(al, kb, 0);
  ((uint32_t *) r)[0] = kb;
  ((uint32_t *) r)[1] = 0;
  for (uint32_t q = 8; q < ko; q++) r[q] = 0;
  for (uint32_t q = 0; q < len; q++) r[ko + q] = t[s + q];
  for (uint32_t q = ko + len; q < end; q++) r[q] = 0;  // the NUL and the padding
}

// ---- vectorised form (16 bytes of text per thread) ------------------------------------------------------
// The per-byte scans above read one byte per scan element (25.6 GB/s of text on a B200).  Here a thread takes
// 16 bytes, turns them into a 16-bit whitespace mask with SWAR compares, and word starts / ends are bit
// operations on the masks:  start = ~ws & (ws << 1 | previous byte is ws),  end = ~ws & (ws >> 1 | next byte is ws).
// A 4 KB chunk (one block iteration) reports how many words start in it and where its last start is; a small
// scan turns that into each chunk's first word index and the start position carried in from the left (a word
// may begin chunks earlier); a second pass pairs every END with its start and writes both positions.
constexpr int TOKV_THREADS = 256;
constexpr uint32_t TOKV_CHUNK = TOKV_THREADS * 16;

// per chunk: words starting in it, and (position + 1) of its last start (0 = none)
// NUL: also report the position of the first NUL byte (atomicMin into *nulpos): the caller counts as if there were
// none and counts again up to it in the rare case that there is one -- a pass of its own over the text otherwise
template <bool NUL>
__global__ void __launch_bounds__(TOKV_THREADS)
tokv_count_kernel(const uint8_t *__restrict__ t, uint64_t end, uint64_t nchunks, uint32_t *__restrict__ cnt, uint32_t *__restrict__ last,
                  unsigned long long *nulpos)
{
  __shared__ uint32_t s_cnt[TOKV_THREADS / 32], s_last[TOKV_THREADS / 32];
  for (uint64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const uint64_t p = c * TOKV_CHUNK + (uint64_t) threadIdx.x * 16;
    uint32_t S, E, Z = 0;
    tokv_masks(t, p, end, S, E, NUL ? &Z : nullptr);
    if (NUL && __any_sync(0xFFFFFFFFu, Z != 0)) {
      unsigned long long z = Z ? p + (__ffs(Z) - 1) : ~0ull;
      for (int o = 16; o; o >>= 1) z = min(z, __shfl_xor_sync(0xFFFFFFFFu, z, o));
      if ((threadIdx.x & 31) == 0) atomicMin(nulpos, z);
    }
    uint32_t n = __popc(S), l = S ? (uint32_t) p + (31 - __clz(S)) + 1 : 0;
    n = __reduce_add_sync(0xFFFFFFFFu, n);
    l = __reduce_max_sync(0xFFFFFFFFu, l);
    if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = n, s_last[threadIdx.x >> 5] = l;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t a = 0, b = 0;
      for (int w = 0; w < TOKV_THREADS / 32; w++) a += s_cnt[w], b = max(b, s_last[w]);
      cnt[c] = a;
      last[c] = b;
    }
    __syncthreads();
  }
}

// one block: base[c] = words before chunk c, carry[c] = (position + 1) of the last start before chunk c
__global__ void __launch_bounds__(1024)
tokv_scan_kernel(const uint32_t *__restrict__ cnt, const uint32_t *__restrict__ last, uint64_t nchunks, uint32_t *__restrict__ base,
                 uint32_t *__restrict__ carry, unsigned long long *total)
{
  __shared__ uint32_t ws[32], wm[32];
  __shared__ uint32_t run_sum, run_max;
  if (threadIdx.x == 0) run_sum = 0, run_max = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint64_t c0 = 0; c0 < nchunks; c0 += 1024) {
    const uint64_t c = c0 + threadIdx.x;
    const uint32_t v = c < nchunks ? cnt[c] : 0, l = c < nchunks ? last[c] : 0;
    uint32_t is = v, im = l;  // inclusive scans inside the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t a = __shfl_up_sync(0xFFFFFFFFu, is, o), b = __shfl_up_sync(0xFFFFFFFFu, im, o);
      if (lane >= o) is += a, im = max(im, b);
    }
    if (lane == 31) ws[warp] = is, wm[warp] = im;
    __syncthreads();
    uint32_t ps = run_sum, pm = run_max;
    for (int w = 0; w < warp; w++) ps += ws[w], pm = max(pm, wm[w]);
    uint32_t em = __shfl_up_sync(0xFFFFFFFFu, im, 1);  // exclusive max inside the warp
    if (lane == 0) em = 0;
    if (c < nchunks) {
      base[c] = ps + is - v;
      carry[c] = max(pm, em);  // everything before this chunk
    }
    __syncthreads();
    if (threadIdx.x == 1023) run_sum = ps + is, run_max = max(pm, im);
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = run_sum;
}

// every END finds its word's index and start and writes (start, end) of that word
__global__ void __launch_bounds__(TOKV_THREADS)
tokv_pairs_kernel(const uint8_t *__restrict__ t, uint64_t end, uint64_t nchunks, const uint32_t *__restrict__ base,
                  const uint32_t *__restrict__ carry, uint32_t *__restrict__ starts, uint32_t *__restrict__ ends, unsigned long long *keybytes)
{
  __shared__ uint32_t s_cnt[TOKV_THREADS / 32], s_last[TOKV_THREADS / 32];
  __shared__ unsigned long long s_kb;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long kbytes = 0;  // key bytes (word + NUL) of the words this thread closes: one sum per block at the end
  if (threadIdx.x == 0) s_kb = 0;
  for (uint64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const uint64_t p = c * TOKV_CHUNK + (uint64_t) threadIdx.x * 16;
    uint32_t S, E;
    tokv_masks(t, p, end, S, E);
    const uint32_t n = __popc(S), l = S ? (uint32_t) p + (31 - __clz(S)) + 1 : 0;
    uint32_t is = n, im = l;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t a = __shfl_up_sync(0xFFFFFFFFu, is, o), b = __shfl_up_sync(0xFFFFFFFFu, im, o);
      if (lane >= o) is += a, im = max(im, b);
    }
    if (lane == 31) s_cnt[warp] = is, s_last[warp] = im;
    uint32_t em = __shfl_up_sync(0xFFFFFFFFu, im, 1);
    if (lane == 0) em = 0;
    __syncthreads();
    uint32_t before = base[c], lastb = carry[c];  // starts / last start before this thread's 16 bytes
    for (int w = 0; w < warp; w++) before += s_cnt[w], lastb = max(lastb, s_last[w]);
    before += is - n;
    lastb = max(lastb, em);
    uint32_t e = E;
    while (e) {
      const int bit = __ffs(e) - 1;
      e &= e - 1;
      const uint32_t sle = S & ((2u << bit) - 1);  // starts at or before this end, inside the 16 bytes
      const uint32_t w = before + __popc(sle) - 1;
      const uint32_t ws = sle ? (uint32_t) p + (31 - __clz(sle)) : lastb - 1, we = (uint32_t) p + bit + 1;
      starts[w] = ws;
      ends[w] = we;
      kbytes += we - ws + 1;
    }
    __syncthreads();
  }
  for (int o = 16; o; o >>= 1) kbytes += __shfl_down_sync(0xFFFFFFFFu, kbytes, o);
  if (lane == 0 && kbytes) atomicAdd(&s_kb, kbytes);
  __syncthreads();
  if (threadIdx.x == 0 && s_kb) atomicAdd(keybytes, s_kb);
}

// one thread per word, 4-byte stores; the key bytes come from two aligned 32-bit loads and a funnel shift
__global__ void __launch_bounds__(256)
tokv_write_kernel(const uint8_t *__restrict__ t, uint64_t tbytes, const uint32_t *__restrict__ starts, const uint32_t *__restrict__ ends, uint64_t nw,
                  uint8_t *__restrict__ out, const uint64_t *__restrict__ offs, Align al)
{
  const uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nw) return;
  const uint32_t s = starts[j], len = ends[j] - s, kb = len + 1;
  uint32_t *r = reinterpret_cast<uint32_t *>(out + offs[j]);
  const uint32_t ko = kv_key_off(al) / 4, nwords = kv_rec_size(al, kb, 0) / 4;
  r[0] = kb;
  r[1] = 0;
  for (uint32_t q = 2; q < ko; q++) r[q] = 0;
  const uint32_t *t32 = reinterpret_cast<const uint32_t *>(t);
  const uint32_t sh = (s & 3) * 8;
  const uint64_t safe = tbytes / 4;  // aligned words wholly inside the text buffer
  for (uint32_t q = ko, b = 0; q < nwords; q++, b += 4) {
    uint32_t word = 0;
    if (b < len) {
      const uint64_t a = ((uint64_t) s + b) >> 2;
      if (a + 1 < safe) word = __funnelshift_r(t32[a], t32[a + 1], sh);
      else  // the last bytes of the buffer: byte loads, never past the end
        for (uint32_t i = 0; i < 4 && b + i < len; i++) word |= (uint32_t) t[(uint64_t) s + b + i] << (8 * i);
      if (len - b < 4) word &= (1u << (8 * (len - b))) - 1;  // the NUL and the padding
    }
    r[q] = word;
  }
}

// position of the first NUL, 16 bytes per load (t 16-byte aligned); the tail is left to first_nul_kernel
__global__ void __launch_bounds__(256) first_nul16_kernel(const uint8_t *__restrict__ t, uint64_t nvec, unsigned long long *pos)
{
  const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
  unsigned long long best = ~0ull;
  for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nvec && best == ~0ull; i += stride) {
    const uint4 v = *reinterpret_cast<const uint4 *>(t + i * 16);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int q = 3; q >= 0; q--) {
      const uint32_t z = ~(((w[q] & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | w[q]) & 0x80808080u;  // 0x80 where a byte is zero
      if (z) best = i * 16 + q * 4 + ((__ffs(z) - 1) >> 3);
    }
  }
  for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
  if ((threadIdx.x & 31) == 0 && best != ~0ull) atomicMin(pos, best);
}

extern "C" int mrgpu_map_tokenize(mrgpu_mr *mr, const void *text, uint64_t nbytes, int addflag, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (nbytes && !text) MR_FAIL(MRGPU_EINVAL, "null input");
  if (nbytes >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "map_tokenize: one call handles a file below 4 GiB (the reference reads int filesize)");
  kv_begin_map(mr, addflag);
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  Buf<uint8_t> tmp;
  const uint8_t *t = stage_input(mr, (const uint8_t *) text, nbytes, tmp, false);  // read several times: stage it
  uint64_t nwords = 0;
  if (nbytes) {
    // 16 bytes per thread when the text is 16-byte aligned (MRGPU_TOK_BYTEWISE=1 forces the per-byte form)
    static const bool force_bytewise = getenv("MRGPU_TOK_BYTEWISE") && atoi(getenv("MRGPU_TOK_BYTEWISE"));
    const bool vec = ((uintptr_t) t % 16) == 0 && !force_bytewise;
    uint64_t end = nbytes;
    if (!vec) {
      unsigned long long *d = mr->d_small;
      MR_CUDA(cudaMemsetAsync(d, 0xFF, 8, s));
      {
        ProfScope ps(pool, "tok_first_nul", (double) nbytes, 2);
        const uint64_t nvec = ((uintptr_t) t % 16) == 0 ? nbytes / 16 : 0;
        if (nvec) first_nul16_kernel<<<grid_for(nvec, 256 * 4, 8), 256, 0, s>>>(t, nvec, d);
        if (nbytes > nvec * 16) first_nul_kernel<<<grid_for(nbytes - nvec * 16, 256 * 16, 8), 256, 0, s>>>(t, nvec * 16, nbytes, d);
        MR_CUDA(cudaGetLastError());
      }
      fetch_small(mr, 1);
      end = std::min<uint64_t>(mr->h_small[0], nbytes);
    }
    Buf<uint32_t> starts, ends;
    auto isstart = [=] __device__(uint64_t i) -> unsigned long long { return (!tok_ws(t[i]) && (i == 0 || tok_ws(t[i - 1]))) ? 1ull : 0ull; };
    if (vec) {
      uint64_t nchunks = (end + TOKV_CHUNK - 1) / TOKV_CHUNK;
      Buf<uint32_t> cnt(pool, nchunks), last(pool, nchunks), cbase(pool, nchunks), carry(pool, nchunks);
      // the count pass looks for the first NUL as well and counts as if there were none; a text that does hold one
      // (strtok stops there, examples/wordfreq.cpp:104-130) is counted again up to it
      MR_CUDA(cudaMemsetAsync(mr->d_small, 0, 8, s));
      MR_CUDA(cudaMemsetAsync(mr->d_small + 2, 0xFF, 8, s));
      {
        ProfScope ps(pool, "tok_count", (double) end, 2);
        tokv_count_kernel<true><<<grid_for(nchunks, 1, 16), TOKV_THREADS, 0, s>>>(t, end, nchunks, cnt.p, last.p, mr->d_small + 2);
        MR_CUDA(cudaGetLastError());
        tokv_scan_kernel<<<1, 1024, 0, s>>>(cnt.p, last.p, nchunks, cbase.p, carry.p, mr->d_small);
        MR_CUDA(cudaGetLastError());
      }
      fetch_small(mr, 3);
      if (mr->h_small[2] < nbytes) {
        end = mr->h_small[2];
        nchunks = (end + TOKV_CHUNK - 1) / TOKV_CHUNK;
        MR_CUDA(cudaMemsetAsync(mr->d_small, 0, 8, s));
        if (nchunks) {
          ProfScope ps(pool, "tok_count", (double) end, 2);
          tokv_count_kernel<false><<<grid_for(nchunks, 1, 16), TOKV_THREADS, 0, s>>>(t, end, nchunks, cnt.p, last.p, nullptr);
          MR_CUDA(cudaGetLastError());
          tokv_scan_kernel<<<1, 1024, 0, s>>>(cnt.p, last.p, nchunks, cbase.p, carry.p, mr->d_small);
          MR_CUDA(cudaGetLastError());
        }
        fetch_small(mr, 1);
      }
      if (nchunks) {
        nwords = mr->h_small[0];
        if (nwords) {
          starts.alloc(pool, nwords);
          ends.alloc(pool, nwords);
          ProfScope ps(pool, "tok_pairs", (double) end + 8.0 * nwords);
          MR_CUDA(cudaMemsetAsync(mr->d_small + 1, 0, 8, s));
          tokv_pairs_kernel<<<grid_for(nchunks, 1, 16), TOKV_THREADS, 0, s>>>(t, end, nchunks, cbase.p, carry.p, starts.p, ends.p, mr->d_small + 1);
          MR_CUDA(cudaGetLastError());
        }
      }
    } else {
      ScanTiles st;
      scan_tiles(pool, end, isstart, st, mr->d_small);
      fetch_small(mr, 1);
      nwords = mr->h_small[0];
      if (nwords) {
        starts.alloc(pool, nwords);
        ends.alloc(pool, nwords);
        uint32_t *ps = starts.p, *pe = ends.p;
        scan_sink(pool, st, isstart, [=] __device__(uint64_t i, unsigned long long e, unsigned long long v) {
          if (v) ps[e] = (uint32_t) i;
          if (!tok_ws(t[i]) && (i + 1 == end || tok_ws(t[i + 1]))) pe[e + v - 1] = (uint32_t) (i + 1);
        });
      }
    }
    if (nwords) {
      uint32_t *ps = starts.p, *pe = ends.p;
      VarAppend va = kv_append_var_begin(mr, kv, nwords);
      uint64_t *oo = va.offs;
      uint64_t base = va.base;
      // total key bytes by a reduction of its own (one atomic per warp on a single counter took 2 of this step's 3.2 ms
      // per GiB of text: four million same-address atomics)
      if (!vec) {  // (the 16-bytes-per-thread pairs kernel has summed them already)
        ScanTiles kt;
        scan_tiles(pool, nwords, [=] __device__(uint64_t j) -> unsigned long long { return pe[j] - ps[j] + 1; }, kt, mr->d_small + 1, "tok_sizes", 8.0 * nwords);
      }
      scan_apply(
          pool, nwords, [=] __device__(uint64_t j) -> unsigned long long { return kv_rec_size(al, pe[j] - ps[j] + 1, 0); },
          [=] __device__(uint64_t j, unsigned long long e, unsigned long long) { oo[j] = base + e; }, mr->d_small, "tok_sizes", 16.0 * nwords);
      fetch_small(mr, 2);
      uint64_t total = mr->h_small[0], ksize = mr->h_small[1];
      kv_append_var_reserve(mr, kv, va, nwords, total);
      {
        ProfScope psw(pool, "tok_write", (double) end + 16.0 * nwords + (double) total);
        if (vec) tokv_write_kernel<<<(unsigned) ((nwords + 255) / 256), 256, 0, s>>>(t, nbytes, ps, pe, nwords, kv->data, kv->offs + kv->nkv, al);
        else tokenize_write_kernel<<<(unsigned) ((nwords + 255) / 256), 256, 0, s>>>(t, ps, pe, nwords, kv->data, kv->offs + kv->nkv, al);
        MR_CUDA(cudaGetLastError());
      }
      kv->nbytes = va.base + total;
      kv->nkv += nwords;
      kv->ksize += ksize;
      sync(mr);
    }
  }
  tmp.release();
  sync(mr);
  spool_after_map(mr);
  if (nkv) *nkv = allreduce_sum(mr, kv_total(mr));
  API_END(mr)
}

// ---------------------------------------------------------------------------------------------
// InvertedIndex href scanner, cpu/InvertedIndex.cpp:196-270
//
// The reference's switch has 5 states plus a 4-byte look-ahead ("href" peeked at the 'h', then j += 3) and
// an inner loop that runs to the closing quote.  Made explicit it is an 8-state DFA without look-ahead
// (verified against the verbatim reference code): START, TAG, ATAG, H1, H2, H3, HREF, LINK.
// A DFA over bytes parallelises as a prefix "sum" of per-byte transition functions under composition:
// a function on 8 states packs into 24 bits.  Pass 1 composes each thread's chunk and each block; a tiny
// sequential pass threads the state through the blocks; pass 2 re-walks every chunk from its now-known
// start state and counts LINK entries / exits; a prefix sum places them; pass 3 writes their positions.
// URL k = text[entry_k, exit_k): entries and exits alternate, so the k-th of each pair up.

#include "href_dfa.cuh"  // the eight states and href_step(state, byte): host code, checked against the reference on the host

// The DFA as data.  A transition function (state -> state) of the 8 states is kept in two forms:
//   byte form  HF{lo, hi}: byte st of the 64-bit pair = f(st).  Composition "f then g" is two PRMT instructions: the
//              selector of a byte permute is four nibbles, each naming one of the eight bytes of (g.lo, g.hi) -- exactly
//              g[f[st]] when the nibbles are f's values (hf_sel packs four bytes into four nibbles);
//   nibble form u32: nibble st = f(st), for walking from ONE known state (shift, mask).
// Per byte value the nibble form sits in a 256-entry table that every block copies into shared memory (the first
// version read a __constant__ table with a different index in every lane -- serialised -- and fetched the text one
// byte per load, lanes 32 bytes apart: 38.6 GB/s of text).
struct HF {
  uint32_t lo, hi;
};
__device__ uint32_t g_href_nib[256];
constexpr uint32_t HF_IDENT_LO = 0x03020100u, HF_IDENT_HI = 0x07060504u;
constexpr int HREF_CHUNK = 32;  // bytes per thread: two 16-byte loads, a warp reads 1 KiB contiguous

static void href_upload_table()
{
  static std::atomic<bool> done[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 64 && done[dev].load()) return;
  uint32_t nib[256];
  for (int c = 0; c < 256; c++) {
    uint32_t f = 0;
    for (int st = 0; st < 8; st++) f |= (uint32_t) href_step(st, c) << (4 * st);
    nib[c] = f;
  }
  MR_CUDA(cudaMemcpyToSymbol(g_href_nib, nib, sizeof(nib)));
  if (dev < 64) done[dev].store(true);
}

__device__ __forceinline__ uint32_t hf_sel(uint32_t x)
{
  const uint32_t y = (x | (x >> 4)) & 0x00FF00FFu;
  return (y | (y >> 8)) & 0xFFFFu;
}
// f first, then g
__device__ __forceinline__ HF hf_then(HF f, HF g) { return HF{__byte_perm(g.lo, g.hi, hf_sel(f.lo)), __byte_perm(g.lo, g.hi, hf_sel(f.hi))}; }
__device__ __forceinline__ int hf_at(HF f, int st) { return (int) (((st < 4 ? f.lo : f.hi) >> (8 * (st & 3))) & 7u); }

__device__ __forceinline__ HF hf_shfl_up(HF f, int o) { return HF{__shfl_up_sync(0xFFFFFFFFu, f.lo, o), __shfl_up_sync(0xFFFFFFFFu, f.hi, o)}; }

// Scan of one function per thread under composition (earlier thread first), NW warps per block: inclusive scan inside
// each warp by shuffles, the warp totals scanned by warp 0, two barriers in all.  (The first version was a
// Hillis-Steele scan in shared memory: log2(threads) rounds of two barriers each -- ncu's top stall of the compose
// kernel was the barrier.)  Returns the EXCLUSIVE prefix of the calling thread; *total = the block's composition
// (valid in every thread).  sh must hold NW entries.
template <int NW>
__device__ __forceinline__ HF hf_block_scan(HF f, uint2 *sh, HF *total)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const HF g = hf_shfl_up(f, o);
    if (lane >= o) f = hf_then(g, f);
  }
  if (lane == 31) sh[warp] = make_uint2(f.lo, f.hi);
  __syncthreads();
  if (warp == 0) {
    const uint2 t = lane < NW ? sh[lane] : make_uint2(HF_IDENT_LO, HF_IDENT_HI);
    HF w{t.x, t.y};
#pragma unroll
    for (int o = 1; o < NW; o <<= 1) {
      const HF g = hf_shfl_up(w, o);
      if (lane >= o) w = hf_then(g, w);
    }
    if (lane < NW) sh[lane] = make_uint2(w.lo, w.hi);  // inclusive over warps
  }
  __syncthreads();
  HF ex = hf_shfl_up(f, 1);  // inclusive of the previous lane = exclusive of this one, inside the warp
  if (lane == 0) ex = HF{HF_IDENT_LO, HF_IDENT_HI};
  if (warp) {
    const uint2 p = sh[warp - 1];
    ex = hf_then(HF{p.x, p.y}, ex);
  }
  const uint2 t = sh[NW - 1];
  *total = HF{t.x, t.y};
  return ex;
}

// the 32 bytes of chunk g (zero beyond the text); returns how many of them are text
__device__ __forceinline__ int href_load32(const uint8_t *__restrict__ t, uint64_t n, uint64_t c0, bool aligned, uint32_t (&w)[8])
{
#pragma unroll
  for (int q = 0; q < 8; q++) w[q] = 0;
  if (c0 >= n) return 0;
  if (aligned && c0 + HREF_CHUNK <= n) {
    const uint4 a = *reinterpret_cast<const uint4 *>(t + c0), b = *reinterpret_cast<const uint4 *>(t + c0 + 16);
    w[0] = a.x, w[1] = a.y, w[2] = a.z, w[3] = a.w, w[4] = b.x, w[5] = b.y, w[6] = b.z, w[7] = b.w;
    return HREF_CHUNK;
  }
  const int m = (int) (n - c0 < (uint64_t) HREF_CHUNK ? n - c0 : (uint64_t) HREF_CHUNK);
  // static register indices only (a run-time index would send w[] to local memory)
#pragma unroll
  for (int j = 0; j < 8; j++) {
    uint32_t x = 0;
#pragma unroll
    for (int b = 0; b < 4; b++)
      if (4 * j + b < m) x |= (uint32_t) t[c0 + 4 * j + b] << (8 * b);
    w[j] = x;
  }
  return m;
}

// Function of one 32-byte chunk.  Folding from the RIGHT keeps the accumulated function in byte form and takes the
// table in selector (nibble) form: Q_k = Q_(k+1) after T_k, i.e. byte st of the result = Q_(k+1)[T_k[st]] = two PRMTs of
// the accumulator with T_k's nibbles as selectors -- no conversion between the forms (the left fold needed one per
// byte: 26 instructions per byte, ALU-bound at 0.9 TB/s of text in ncu).
__global__ void __launch_bounds__(256)
href_compose_kernel(const uint8_t *__restrict__ t, uint64_t n, bool aligned, uint2 *__restrict__ tfun /*exclusive prefix per thread-chunk*/,
                    uint2 *__restrict__ bfun /*per block*/)
{
  __shared__ uint32_t s_nib[256];
  __shared__ uint2 sh[8];
  s_nib[threadIdx.x] = g_href_nib[threadIdx.x];
  __syncthreads();
  const uint64_t c0 = ((uint64_t) blockIdx.x * 256 + threadIdx.x) * HREF_CHUNK;
  uint32_t w[8];
  const int m = href_load32(t, n, c0, aligned, w);
  HF f{HF_IDENT_LO, HF_IDENT_HI};
#pragma unroll
  for (int q = HREF_CHUNK - 1; q >= 0; q--)
    if (q < m) {
      const uint32_t sel = s_nib[(w[q >> 2] >> (8 * (q & 3))) & 0xFFu];
      f = HF{__byte_perm(f.lo, f.hi, sel), __byte_perm(f.lo, f.hi, sel >> 16)};
    }
  HF total;
  const HF ex = hf_block_scan<8>(f, sh, &total);
  tfun[(uint64_t) blockIdx.x * 256 + threadIdx.x] = make_uint2(ex.lo, ex.hi);
  if (threadIdx.x == 255) bfun[blockIdx.x] = make_uint2(total.lo, total.hi);
}

// thread the state through the blocks: bstate[b] = state at the first byte of block b.  One block; a thread composes
// 16 consecutive block functions, the 1024 partial results are scanned, every thread then walks its 16 blocks from
// the state it starts in (the first version walked all blocks with one thread: 0.25 ms per 128 MiB of text, more
// than the rest of the scanner together).
constexpr int HREF_BS_PER = 16;
__global__ void __launch_bounds__(1024) href_blockstate_kernel(const uint2 *__restrict__ bfun, uint64_t nblocks, uint8_t *__restrict__ bstate)
{
  __shared__ uint2 sh[32];
  __shared__ int s_state;
  if (threadIdx.x == 0) s_state = HS_START;
  __syncthreads();
  for (uint64_t base = 0; base < nblocks; base += 1024 * HREF_BS_PER) {
    const uint64_t i0 = base + (uint64_t) threadIdx.x * HREF_BS_PER;
    HF f{HF_IDENT_LO, HF_IDENT_HI};
    for (int j = 0; j < HREF_BS_PER; j++)
      if (i0 + j < nblocks) {
        const uint2 g = bfun[i0 + j];
        f = hf_then(f, HF{g.x, g.y});
      }
    HF total;
    const HF ex = hf_block_scan<32>(f, sh, &total);
    const int st0 = s_state;
    int st = hf_at(ex, st0);
    for (int j = 0; j < HREF_BS_PER; j++)
      if (i0 + j < nblocks) {
        bstate[i0 + j] = (uint8_t) st;
        const uint2 g = bfun[i0 + j];
        st = hf_at(HF{g.x, g.y}, st);
      }
    __syncthreads();
    if (threadIdx.x == 0) s_state = hf_at(total, st0);
    __syncthreads();
  }
}

// The walk from the chunk's now-known start state.  Its result is kept as two 32-bit masks per chunk -- bit q of the
// high word: a URL's first byte is q + 1 bytes into the chunk (the opening quote is byte q); bit q of the low word: a
// closing quote at byte q -- so that the pass that writes the positions needs neither the text nor a second walk
// (it was a third read of the text and as expensive as this kernel: nearly every warp holds a chunk with a link).
__global__ void __launch_bounds__(256)
href_count_kernel(const uint8_t *__restrict__ t, uint64_t n, bool aligned, const uint2 *__restrict__ tfun, const uint8_t *__restrict__ bstate,
                  unsigned long long *__restrict__ masks)
{
  __shared__ uint32_t s_nib[256];
  s_nib[threadIdx.x] = g_href_nib[threadIdx.x];
  __syncthreads();
  const uint64_t g = (uint64_t) blockIdx.x * 256 + threadIdx.x;
  const uint2 tf = tfun[g];
  const int st = hf_at(HF{tf.x, tf.y}, bstate[blockIdx.x]);
  uint32_t w[8];
  const uint64_t c0 = g * HREF_CHUNK;
  const int m = href_load32(t, n, c0, aligned, w);
  // L: bit q = the state after byte q is LINK.  LINK is entered only from HREF (by the opening quote) and left only to
  // START (by the closing quote), so entries and exits are the rising and falling edges of L.
  unsigned L = 0;
  int cur = st;
#pragma unroll
  for (int q = 0; q < HREF_CHUNK; q++)
    if (q < m) {
      cur = (int) ((s_nib[(w[q >> 2] >> (8 * (q & 3))) & 0xFFu] >> (4 * cur)) & 7u);
      L |= (cur == HS_LINK ? 1u : 0u) << q;
    }
  const unsigned valid = m >= 32 ? 0xFFFFFFFFu : ((1u << m) - 1u), prev = (L << 1) | (st == HS_LINK ? 1u : 0u);
  const unsigned em = L & ~prev & valid, xm = ~L & prev & valid;
  masks[g] = ((unsigned long long) em << 32) | xm;
}

// every entry / exit position goes to its slot (exclusive counts from the scan of the masks' bit counts)
__global__ void __launch_bounds__(256)
href_sink_kernel(uint64_t nthr, const unsigned long long *__restrict__ masks, const unsigned long long *__restrict__ excl, uint32_t *__restrict__ entry,
                 uint32_t *__restrict__ exitp, unsigned long long *keybytes /* += sum of exits - sum of entries (mod 2^64) */)
{
  __shared__ unsigned long long s_sum;
  if (threadIdx.x == 0) s_sum = 0;
  __syncthreads();
  const uint64_t g = (uint64_t) blockIdx.x * 256 + threadIdx.x;
  const unsigned long long mk = g < nthr ? masks[g] : 0ull;
  unsigned long long sum = 0;
  if (mk) {  // nothing starts or ends in most chunks
    uint32_t em = (uint32_t) (mk >> 32), xm = (uint32_t) mk;
    const unsigned long long e = excl[g];
    uint32_t ie = (uint32_t) (e >> 32), ix = (uint32_t) e;
    const uint32_t c0 = (uint32_t) (g * HREF_CHUNK);
    while (em) {
      const uint32_t p = c0 + (uint32_t) __ffs(em);  // bit q: first URL byte at c0 + q + 1
      entry[ie++] = p;
      sum -= p;
      em &= em - 1;
    }
    while (xm) {
      const uint32_t p = c0 + (uint32_t) __ffs(xm) - 1;
      exitp[ix++] = p;
      sum += p;
      xm &= xm - 1;
    }
  }
  for (int o = 16; o; o >>= 1) sum += __shfl_down_sync(0xFFFFFFFFu, sum, o);
  if ((threadIdx.x & 31) == 0 && sum) atomicAdd(&s_sum, sum);
  __syncthreads();
  if (threadIdx.x == 0 && s_sum) atomicAdd(keybytes, s_sum);
}

// eight lanes per record, one 4-byte store per lane and trip (a warp writes four records' worth of contiguous words);
// one thread per record with byte stores took 0.135 ms per 128 MiB file for 30 MB of records
__global__ void __launch_bounds__(256)
href_write_kernel(const uint8_t *__restrict__ t, const uint32_t *__restrict__ entry, const uint32_t *__restrict__ exitp, uint64_t nurl,
                  const uint8_t *__restrict__ fname, uint32_t fnb, uint8_t *__restrict__ out, const uint64_t *__restrict__ offs, Align al)
{
  const uint64_t j = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const uint32_t l = threadIdx.x & 7u;
  if (j >= nurl) return;
  const uint32_t s = entry[j], kb = exitp[j] - s;
  uint8_t *r = out + offs[j];
  const uint32_t ko = kv_key_off(al), vo = kv_val_off(al, kb), end = kv_rec_size(al, kb, fnb);
  auto byte_at = [&](uint32_t pos) -> uint32_t {
    if (pos < 4) return (kb >> (8 * pos)) & 0xFFu;
    if (pos < 8) return (fnb >> (8 * (pos - 4))) & 0xFFu;
    if (pos >= ko && pos < ko + kb) return t[s + (pos - ko)];
    if (pos >= vo && pos < vo + fnb) return fname[pos - vo];
    return 0u;
  };
  if ((((uintptr_t) r) & 3) == 0) {
    const uint32_t nwords = end >> 2;
    for (uint32_t wq = l; wq < nwords; wq += 8) {
      const uint32_t p = 4 * wq;
      ((uint32_t *) r)[wq] = byte_at(p) | (byte_at(p + 1) << 8) | (byte_at(p + 2) << 16) | (byte_at(p + 3) << 24);
    }
    for (uint32_t p = 4 * nwords + l; p < end; p += 8) r[p] = (uint8_t) byte_at(p);
  } else
    for (uint32_t p = l; p < end; p += 8) r[p] = (uint8_t) byte_at(p);
}

extern "C" int mrgpu_map_href(mrgpu_mr *mr, const void *text, uint64_t nbytes, const char *fname, int addflag, uint64_t *nkv)
{
  API_BEGIN(mr)
  if (nbytes && !text) MR_FAIL(MRGPU_EINVAL, "null input");
  if (!fname) MR_FAIL(MRGPU_EINVAL, "null file name");
  if (nbytes >= 0xFFFFFFFFull) MR_FAIL(MRGPU_ELIMIT, "map_href: one call handles a file below 4 GiB (the reference reads int filesize)");
  kv_begin_map(mr, addflag);
  DevKV *kv = mr->kv;
  Pool &pool = mr->pool;
  cudaStream_t s = pool.stream;
  const Align al = mr->al;
  href_upload_table();
  Buf<uint8_t> tmp;
  const uint8_t *t = stage_input(mr, (const uint8_t *) text, nbytes, tmp, false);  // read several times: stage it
  const uint32_t fnb = (uint32_t) strlen(fname) + 1;  // value = file name with its NUL, cpu/InvertedIndex.cpp:262
  if (nbytes) {
    const uint64_t n = nbytes;
    const uint64_t nblk = (n + 256 * HREF_CHUNK - 1) / (256 * HREF_CHUNK), nthr = nblk * 256;
    Buf<uint2> tfun(pool, nthr), bfun(pool, nblk);
    Buf<uint8_t> bstate(pool, nblk);
    Buf<unsigned long long> cnt(pool, nthr), excl(pool, nthr);
    const bool aligned = ((uintptr_t) t % 16) == 0;
    {
      ProfScope ps(pool, "href_compose", (double) n, 2);
      href_compose_kernel<<<(unsigned) nblk, 256, 0, s>>>(t, n, aligned, tfun.p, bfun.p);
      MR_CUDA(cudaGetLastError());
      href_blockstate_kernel<<<1, 1024, 0, s>>>(bfun.p, nblk, bstate.p);
      MR_CUDA(cudaGetLastError());
    }
    {
      ProfScope ps(pool, "href_count", (double) n);
      href_count_kernel<<<(unsigned) nblk, 256, 0, s>>>(t, n, aligned, tfun.p, bstate.p, cnt.p);
      MR_CUDA(cudaGetLastError());
    }
    const unsigned long long *pc = cnt.p;
    auto val = [=] __device__(uint64_t g) -> unsigned long long {  // (entries, exits) of the chunk, packed like their totals
      const unsigned long long mk = pc[g];
      return ((unsigned long long) __popc((unsigned) (mk >> 32)) << 32) | (unsigned) __popc((unsigned) mk);
    };
    ScanTiles st;
    scan_tiles(pool, nthr, val, st, mr->d_small);
    fetch_small(mr, 1);
    const uint64_t nentry = mr->h_small[0] >> 32, nurl = mr->h_small[0] & 0xFFFFFFFFull;
    if (nurl) {
      Buf<uint32_t> entry(pool, nentry), exitp(pool, nurl);
      uint32_t *pe = entry.p, *px = exitp.p;
      unsigned long long *pex = excl.p;
      scan_sink(pool, st, val, [=] __device__(uint64_t g, unsigned long long e, unsigned long long) { pex[g] = e; });
      {
        ProfScope ps(pool, "href_positions", 16.0 * nthr + 8.0 * nurl);
        MR_CUDA(cudaMemsetAsync(mr->d_small + 1, 0, 8, s));
        href_sink_kernel<<<(unsigned) nblk, 256, 0, s>>>(nthr, cnt.p, excl.p, pe, px, mr->d_small + 1);
        MR_CUDA(cudaGetLastError());
      }
      VarAppend va = kv_append_var_begin(mr, kv, nurl);
      uint64_t *oo = va.offs;
      uint64_t base = va.base;
      if (nentry != nurl) {  // a URL still open at the end of the text: the sums of the positions pass do not pair up
        ScanTiles kt;        // total key bytes by a reduction of its own
        scan_tiles(pool, nurl, [=] __device__(uint64_t j) -> unsigned long long { return px[j] - pe[j]; }, kt, mr->d_small + 1, "href_sizes", 8.0 * nurl);
      }
      scan_apply(
          pool, nurl, [=] __device__(uint64_t j) -> unsigned long long { return kv_rec_size(al, px[j] - pe[j], fnb); },
          [=] __device__(uint64_t j, unsigned long long e, unsigned long long) { oo[j] = base + e; }, mr->d_small, "href_sizes", 16.0 * nurl);
      fetch_small(mr, 2);
      uint64_t total = mr->h_small[0], ksize = mr->h_small[1];
      kv_append_var_reserve(mr, kv, va, nurl, total);
      Buf<uint8_t> dname(pool, fnb);
      MR_CUDA(cudaMemcpyAsync(dname.p, fname, fnb, cudaMemcpyHostToDevice, s));
      {
        ProfScope ps(pool, "href_write", (double) total + 8.0 * nurl);
        href_write_kernel<<<(unsigned) ((nurl * 8 + 255) / 256), 256, 0, s>>>(t, pe, px, nurl, dname.p, fnb, kv->data, kv->offs + kv->nkv, al);
        MR_CUDA(cudaGetLastError());
      }
      kv->nbytes = va.base + total;
      kv->nkv += nurl;
      kv->ksize += ksize;
      kv->vsize += nurl * fnb;
      sync(mr);
    }
  }
  tmp.release();
  sync(mr);
  spool_after_map(mr);
  if (nkv) *nkv = allreduce_sum(mr, kv_total(mr));
  API_END(mr)
}

"""Synthetic inputs for the BASELINE.json configs.

The reference ships no data (its `examples/words/text.txt` is a missing blob and the `cpu/` apps read
hard-coded `/data/bzhaoad/...` files), so every workload here is generated from a fixed seed.  All
generators are counter based: element i depends only on (seed, i), so any shard [lo, hi) can be made
independently on any rank, on the CPU (numpy) or on the GPU (torch), with identical results.
"""
from __future__ import annotations

import numpy as np

_M64 = (1 << 64) - 1


def splitmix64_np(idx: np.ndarray, seed: int) -> np.ndarray:
    """splitmix64 finaliser of (seed + idx * golden) as uint64 (numpy, wraps mod 2^64)."""
    with np.errstate(over="ignore"):
        z = idx.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(seed & _M64)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def _lsr_t(x, s):
    """logical shift right on a torch int64 tensor (torch's >> is arithmetic)."""
    return (x >> s) & ((1 << (64 - s)) - 1)


def _to_i64(c: int) -> int:
    c &= _M64
    return c - (1 << 64) if c >= (1 << 63) else c


def splitmix64_torch(idx, seed: int):
    """Same function as splitmix64_np on a torch int64 tensor (bit pattern identical)."""
    z = idx * _to_i64(0x9E3779B97F4A7C15) + _to_i64(seed)
    z = (z ^ _lsr_t(z, 30)) * _to_i64(0xBF58476D1CE4E5B9)
    z = (z ^ _lsr_t(z, 27)) * _to_i64(0x94D049BB133111EB)
    z = z ^ _lsr_t(z, 31)
    return z


def zipf_cdf(ndistinct: int, exponent: float) -> np.ndarray:
    """float64 CDF of Zipf(exponent) over ranks 1..ndistinct (computed on the CPU, deterministic)."""
    w = np.arange(1, ndistinct + 1, dtype=np.float64) ** (-exponent)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    cdf[-1] = 1.0
    return cdf


INTCOUNT_SEED = 88172645463325252


def intcount_keys_np(lo: int, hi: int, ndistinct: int = 1 << 24, exponent: float = 1.1,
                     seed: int = INTCOUNT_SEED, cdf: np.ndarray | None = None) -> np.ndarray:
    """uint32 keys [lo, hi) of the IntCount workload: key = 0-based Zipf rank of a uniform draw."""
    if cdf is None:
        cdf = zipf_cdf(ndistinct, exponent)
    z = splitmix64_np(np.arange(lo, hi, dtype=np.uint64), seed)
    u = (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    k = np.searchsorted(cdf, u, side="right")
    return np.minimum(k, ndistinct - 1).astype(np.uint32)


def intcount_keys_torch(lo: int, hi: int, device, ndistinct: int = 1 << 24, exponent: float = 1.1,
                        seed: int = INTCOUNT_SEED, cdf_t=None, chunk: int = 1 << 26):
    """Same keys as intcount_keys_np, generated with torch ops on `device` (int32 tensor holding the
    uint32 bit patterns; keys are < 2^31 so the values agree as well)."""
    import torch

    if cdf_t is None:
        cdf_t = torch.from_numpy(zipf_cdf(ndistinct, exponent)).to(device)
    out = torch.empty(hi - lo, dtype=torch.int32, device=device)
    for s in range(lo, hi, chunk):
        e = min(hi, s + chunk)
        idx = torch.arange(s, e, dtype=torch.int64, device=device)
        z = splitmix64_torch(idx, seed)
        u = _lsr_t(z, 11).to(torch.float64) * (1.0 / 9007199254740992.0)
        k = torch.searchsorted(cdf_t, u, right=True)
        out[s - lo:e - lo] = torch.clamp(k, max=ndistinct - 1).to(torch.int32)
    return out


# ---------------------------------------------------------------------------------------- text

WORDS_SEED = 0x9E3779B97F4A7C15


def vocabulary(nvocab: int, seed: int = WORDS_SEED) -> list[bytes]:
    """nvocab distinct lowercase words, lengths 2..12."""
    out, seen, i = [], set(), 0
    while len(out) < nvocab:
        z = int(splitmix64_np(np.array([i], dtype=np.uint64), seed ^ 0x5555)[0])
        i += 1
        ln = 2 + z % 11
        z >>= 4
        w = bytearray()
        for _ in range(ln):
            w.append(97 + z % 26)
            z //= 26
        w = bytes(w)
        if w not in seen:
            seen.add(w)
            out.append(w)
    return out


def words_text(nwords: int, nvocab: int = 1 << 17, exponent: float = 1.0, seed: int = WORDS_SEED,
               start: int = 0) -> bytes:
    """Synthetic wordfreq corpus: Zipf-drawn words, single-space separated, newline every 16 words."""
    vocab = vocabulary(nvocab, seed)
    cdf = zipf_cdf(nvocab, exponent)
    z = splitmix64_np(np.arange(start, start + nwords, dtype=np.uint64), seed)
    u = (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    k = np.minimum(np.searchsorted(cdf, u, side="right"), nvocab - 1)
    parts = []
    for i, w in enumerate(k):
        parts.append(vocab[int(w)])
        parts.append(b"\n" if (i % 16) == 15 else b" ")
    return b"".join(parts)


# ---------------------------------------------------------------------------------------- html

HTML_SEED = 0xD1B54A32D192ED03

_FILLER = [b"<p>lorem ipsum dolor sit amet</p>", b"<div class=\"x\">text</div>", b"plain words here ",
           b"<br/>", b"<span>consectetur</span> ", b"<h1>title</h1>\n", b"<img src=\"i.png\">",
           b"<ul><li>item</li></ul>\n", b"<b>bold</b> and <i>it</i> ", b"&amp; entity; "]

ADVERSARIAL = [b'<a  href = "u2">', b'<ab href="u3">', b'<a hrex="u4"> <a href="u5">', b'<<a href="u6">',
               b'< a href="u7">', b'<a class="c" href="u8">', b"<a href='u9'> <a href=\"u10\">",
               b'<a href="">empty <a href="u11">', b'<a hhref="u12">', b'<a href=x"u13">',
               b'<a\thref="u14">', b'<A HREF="u15">', b'<a href="u16"><a href="u17">',
               b'<a hr<a href="u18">', b'<a hre<a href="u19">', b'<a h<a href="u20">',
               b'<a href<a href="u21">', b'<a href=="u22">', b'<a  h ref="u23">', b'<a a href="u24">']


def html_text(nbytes: int, nurls: int = 1 << 22, seed: int = HTML_SEED, start: int = 0,
              adversarial: bool = False, unterminated_tail: bool = False) -> bytes:
    """Synthetic HTML with an <a href="URL"> about every 256 bytes; URLs Zipf(1.0) over nurls."""
    cdf = zipf_cdf(nurls, 1.0)
    out, size, i = [], 0, start
    while size < nbytes:
        z = int(splitmix64_np(np.array([i], dtype=np.uint64), seed)[0])
        i += 1
        u = (z >> 11) * (1.0 / 9007199254740992.0)
        k = min(int(np.searchsorted(cdf, u, side="right")), nurls - 1)
        frag = [b'<a href="http://h%04d.example/p/%d">link</a>' % (k % 10000, k)]
        z2 = z
        for _ in range(6):
            frag.append(_FILLER[z2 % len(_FILLER)])
            z2 //= len(_FILLER)
        if adversarial:
            frag.append(ADVERSARIAL[z2 % len(ADVERSARIAL)])
        frag = b"".join(frag)
        out.append(frag)
        size += len(frag)
    if unterminated_tail:
        out.append(b'<a href="never closed')
    return b"".join(out)

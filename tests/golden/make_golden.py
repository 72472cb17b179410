#!/usr/bin/env python
"""Generate tests/golden/golden.npz from the UNMODIFIED reference library.

Run in the build container (needs /root/reference, via `make -C oracle ref`):
    python tests/golden/make_golden.py
Every array in golden.npz is either an input we authored (suffix .in, or params in manifest) or bytes
written by oracle/_ref/refdrv, i.e. by the reference's own src/*.cpp + oink callbacks + example
callbacks.  The reference has no golden vectors of its own; these recorded outputs
pin oracle/mr_oracle.c and, through it and directly, the CUDA path.  /root/reference is not needed to
*use* the fixtures.
"""
import json
import os
import struct
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import mrhelpers as H  # noqa: E402
from gpu_mapreduce_b200 import workloads as W  # noqa: E402


def b2a(b: bytes) -> np.ndarray:
    return np.frombuffer(b, dtype=np.uint8).copy()


def rnd(seed):
    return np.random.RandomState(seed)


def main():
    assert H.have_ref(), "build the reference first: make -C oracle ref"
    G, M = {}, {}

    # 1. hashlittle known answers: lengths 0..40, several initvals (src/hash.cpp:129)
    pairs = []
    r = rnd(1)
    for ln in range(0, 41):
        key = bytes(r.randint(0, 256, size=ln).astype(np.uint8))
        for iv in (0, 1, 2, 4, 8, 0xDEADBEEF):
            pairs.append((iv, key))
    pairs += [(0, b"hello world!"), (8, b"hello world!"), (0, struct.pack("<I", 12345)),
              (0, struct.pack("<QQ", 3, 5))]
    hv = H.ref_hash(pairs)
    M["hash"] = [[iv, k.hex(), h] for (iv, k), h in zip(pairs, hv)]

    # 2. page images, SURVEY.md B.2
    recs = [(b"a\0", b""), (b"bb\0", struct.pack("<i", 1)), (b"a\0", b""), (b"ccc\0", struct.pack("<i", 3)),
            (b"bb\0", b""), (b"a\0", struct.pack("<i", 5))]
    kv = H.kv_encode(recs)
    G["pages_b2.in"] = b2a(kv)
    for al in (4, 8):
        G[f"pages_b2.kv.a{al}"] = b2a(H.ref_ops(kv, ["pagekv"], align=al))
        G[f"pages_b2.kmv.a{al}"] = b2a(H.ref_ops(kv, ["convert", "pagekmv"], align=al))
    G["pages_b2.reduce_sum"] = b2a(H.ref_ops(kv, ["convert", "reduce_sum", "pagekv"]))

    # 3. multi-page KV with variable-length records, 32 KiB pages
    r = rnd(3)
    recs = []
    for i in range(3000):
        k = bytes(r.randint(1, 256, size=r.randint(1, 24)).astype(np.uint8))
        v = bytes(r.randint(0, 256, size=r.randint(0, 13)).astype(np.uint8))
        recs.append((k, v))
    kv = H.kv_encode(recs)
    G["pages_multi.in"] = b2a(kv)
    G["pages_multi.kv"] = b2a(H.ref_ops(kv, ["pagekv"], memsize=-32768))

    # 4. convert order + reduce callbacks on variable-length data with repeated keys
    r = rnd(4)
    keys = [bytes(r.randint(1, 256, size=r.randint(1, 20)).astype(np.uint8)) for _ in range(150)]
    recs = []
    for i in range(2500):
        k = keys[int(r.zipf(1.3)) % len(keys)]
        v = bytes(r.randint(0, 256, size=r.randint(0, 9)).astype(np.uint8))
        recs.append((k, v))
    kv = H.kv_encode(recs)
    G["convert.in"] = b2a(kv)
    G["convert.kmv"] = b2a(H.ref_ops(kv, ["convert", "dumpkmv"]))
    G["convert.kmvpages"] = b2a(H.ref_ops(kv, ["convert", "pagekmv"]))
    for cb in ("count", "sum", "cull"):
        G[f"convert.reduce_{cb}"] = b2a(H.ref_ops(kv, ["collate", f"reduce_{cb}", "dumpkv"]))

    # 5. IntCount small: cpu/IntCount.cpp pipeline + reduce count + sort_keys(1)
    keys = W.intcount_keys_np(0, 6000, ndistinct=512)
    M["intcount"] = {"n": 6000, "ndistinct": 512}
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "k.u32"), os.path.join(td, "o.kv")
        keys.tofile(fin)
        H.ref_run(["intcount", fin, fout, 64])
        G["intcount.out"] = b2a(open(fout, "rb").read())
    G["intcount.in"] = keys.view(np.uint8).copy()

    # 6. extended (multi-block) KMV, SURVEY.md B.3: key 7 x 18000 among 20000 pairs, 32 KiB pages
    recs = []
    for i in range(20000):
        k = 7 if (i % 10) != 3 else 100 + (i // 10) % 7
        recs.append((struct.pack("<i", k), struct.pack("<i", 1)))
    recs[0] = (struct.pack("<i", 100), struct.pack("<i", 1))
    kv = H.kv_encode(recs)
    M["extended"] = {"hot": 7}
    G["extended.in"] = b2a(kv)
    G["extended.count"] = b2a(H.ref_ops(kv, ["convert", "reduce_count", "dumpkv"], memsize=-32768))

    # 7. wordfreq: examples/wordfreq.cpp fileread + collate + sum
    t1 = W.words_text(4000, nvocab=300)
    t2 = b"  lead\t\ttabs\fform\rcr\vvt  multi   space\n\n\nend" + W.words_text(500, nvocab=50, start=99) + \
         b"tail-without-newline"
    t3 = b"before nul \x00 after nul is ignored by strtok"
    with tempfile.TemporaryDirectory() as td:
        names = []
        for i, t in enumerate((t1, t2, t3)):
            p = os.path.join(td, f"part-{i:05d}")
            open(p, "wb").write(t)
            names.append(p)
            G[f"wordfreq.in{i}"] = b2a(t)
        fout = os.path.join(td, "o")
        H.ref_run(["wordfreq", fout, 64, "mapout"] + names)
        G["wordfreq.mapout"] = b2a(open(fout, "rb").read())
        H.ref_run(["wordfreq", fout, 64, "final"] + names)
        G["wordfreq.final"] = b2a(open(fout, "rb").read())

    # 8. InvertedIndex: href FSM (cpu/InvertedIndex.cpp:196-270 verbatim) + aggregate + convert
    h1 = W.html_text(12000, nurls=40, adversarial=True)
    h2 = W.html_text(8000, nurls=40, start=1000, adversarial=True, unterminated_tail=True)
    h3 = b"".join(W.ADVERSARIAL) + b'<a href="last"'
    h4 = b'<a href="x\x00y">nul inside link</a><a href="z">'
    with tempfile.TemporaryDirectory() as td:
        names = []
        for i, t in enumerate((h1, h2, h3, h4)):
            p = os.path.join(td, f"part-{i:05d}")
            open(p, "wb").write(t)
            names.append(p)
            G[f"invindex.in{i}"] = b2a(t)
        fout = os.path.join(td, "o")
        H.ref_run(["invindex", fout, 64, "mapout"] + names)
        G["invindex.mapout"] = b2a(open(fout, "rb").read())
        H.ref_run(["invindex", fout, 64, "kmv"] + names)
        G["invindex.kmv"] = b2a(open(fout, "rb").read())

    # 9. R-MAT + degree through the unmodified oink callbacks
    M["rmat"] = []
    cases = [(8, 4, .57, .19, .19, .05, 0.0, 12345, 1, 2), (8, 4, .57, .19, .19, .05, 0.0, 12345, 2, 2),
             (7, 3, .25, .25, .25, .25, 0.0, 7, 3, 1), (6, 4, .45, .15, .15, .25, 0.1, 99, 2, 2),
             # 4 and 8 virtual ranks (SCALE runs: the same ranks on real GPUs, bench_configs.py rmat parity)
             (10, 8, .57, .19, .19, .05, 0.0, 12345, 4, 2), (10, 8, .57, .19, .19, .05, 0.0, 12345, 8, 2)]
    with tempfile.TemporaryDirectory() as td:
        for ci, c in enumerate(cases):
            fe, fd = os.path.join(td, "e"), os.path.join(td, "d")
            N, ef, a, b, cc, d, fr, seed, P, dup = c
            t, _ = H.ref_run(["rmat", N, ef, a, b, cc, d, fr, seed, P, 64, fe, fd, dup])
            it = [x for x in t if x.get("op") == "rmat_done"][0]["iterations"]
            M["rmat"].append({"args": list(c), "iterations": it})
            G[f"rmat{ci}.edges"] = b2a(open(fe, "rb").read())
            G[f"rmat{ci}.degree"] = b2a(open(fd, "rb").read())

    # 10. sort_keys flags +-1..+-6 (compare functions src/mapreduce.cpp:2692-2802)
    r = rnd(10)
    n = 700
    ints = r.randint(-50, 50, size=n).astype(np.int32)
    ints[:5] = [-2 ** 31, 2 ** 31 - 1, 0, -1, 1]
    u64 = r.randint(0, 40, size=n).astype(np.uint64)
    u64[:4] = [2 ** 64 - 1, 2 ** 63, 0, 2 ** 32]
    f32 = (r.randint(-30, 30, size=n) * 0.25).astype(np.float32)
    f32[:4] = [np.float32("inf"), np.float32("-inf"), -0.0, 0.0]
    f64 = (r.randint(-30, 30, size=n) * 0.125).astype(np.float64)
    f64[:4] = [np.inf, -np.inf, -0.0, 0.0]
    words = [bytes(r.randint(97, 101, size=r.randint(0, 6)).astype(np.uint8)) for _ in range(n)]
    words[0] = b"\xff\x80"
    words[1] = b"\x7f"

    def with_vals(keys):
        return [(k, struct.pack("<i", i)) for i, k in enumerate(keys)]

    srt = {1: with_vals([x.tobytes() for x in ints]), 2: with_vals([x.tobytes() for x in u64]),
           3: with_vals([x.tobytes() for x in f32]), 4: with_vals([x.tobytes() for x in f64]),
           5: with_vals([w + b"\0" for w in words]), 6: with_vals([w + b"\0" for w in words])}
    for fl, recs in srt.items():
        kv = H.kv_encode(recs)
        G[f"sort{fl}.in"] = b2a(kv)
        for sgn in (1, -1):
            G[f"sort{fl}.out{'p' if sgn > 0 else 'm'}"] = b2a(H.ref_ops(kv, [f"sort_keys:{sgn * fl}", "dumpkv"]))

    # 11. sort_values flags +-1..+-6 (src/mapreduce.cpp:2156-2182): the records of 10 with key and value swapped
    for fl, recs in srt.items():
        kv = H.kv_encode([(v, k) for k, v in recs])
        G[f"sortv{fl}.in"] = b2a(kv)
        for sgn in (1, -1):
            G[f"sortv{fl}.out{'p' if sgn > 0 else 'm'}"] = b2a(H.ref_ops(kv, [f"sort_values:{sgn * fl}", "dumpkv"]))

    # 12. compress = local convert + reduce in one call (src/mapreduce.cpp:749-851), on the data of 4
    kv4 = G["convert.in"].tobytes()
    for cb in ("count", "sum", "cull"):
        G[f"compress.{cb}"] = b2a(H.ref_ops(kv4, [f"compress_{cb}", "dumpkv"]))

    # 13. clone (src/mapreduce.cpp:631) and collapse (:681) on the data of 4 and of 2; collapse's key is "all\0"
    for name, kvb in (("convert", kv4), ("pages_b2", G["pages_b2.in"].tobytes())):
        G[f"clone.{name}"] = b2a(H.ref_ops(kvb, ["clone", "dumpkmv"]))
        G[f"collapse.{name}"] = b2a(H.ref_ops(kvb, ["collapse:all", "dumpkmv"]))
    for al in (4, 8):
        G[f"collapse.pages_b2.page.a{al}"] = b2a(H.ref_ops(G["pages_b2.in"].tobytes(), ["collapse:all", "pagekmv"], align=al))
        G[f"clone.pages_b2.page.a{al}"] = b2a(H.ref_ops(G["pages_b2.in"].tobytes(), ["clone", "pagekmv"], align=al))

    np.savez_compressed(os.path.join(HERE, "golden.npz"), **G)
    with open(os.path.join(HERE, "manifest.json"), "w") as f:
        json.dump(M, f, indent=0)
    print("wrote", len(G), "arrays,", os.path.getsize(os.path.join(HERE, "golden.npz")), "bytes")


if __name__ == "__main__":
    main()

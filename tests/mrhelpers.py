"""Shared test infrastructure: interchange codecs, the ctypes view of oracle/libmr_oracle.so, the runner
for oracle/_ref/refdrv (the unmodified reference, when it is built) and canonicalisers.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

import ctypes
import hashlib
import json
import os
import struct
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libmr_oracle.so")
REFDRV = os.path.join(ORACLE_DIR, "_ref", "refdrv")
GOLDEN = os.path.join(ROOT, "tests", "golden")


# ------------------------------------------------------------------------------------ codecs

def kv_encode(recs) -> bytes:
    out = [struct.pack("<Q", len(recs))]
    for k, v in recs:
        out.append(struct.pack("<ii", len(k), len(v)))
        out.append(k)
        out.append(v)
    return b"".join(out)


def kv_decode(buf: bytes):
    (n,) = struct.unpack_from("<Q", buf, 0)
    p, out = 8, []
    for _ in range(n):
        kb, vb = struct.unpack_from("<ii", buf, p)
        p += 8
        out.append((bytes(buf[p:p + kb]), bytes(buf[p + kb:p + kb + vb])))
        p += kb + vb
    assert p == len(buf), (p, len(buf))
    return out


def kmv_decode(buf: bytes):
    (n,) = struct.unpack_from("<Q", buf, 0)
    p, out = 8, []
    for _ in range(n):
        kb, nv = struct.unpack_from("<iQ", buf, p)
        p += 12
        key = bytes(buf[p:p + kb])
        p += kb
        vals = []
        for _ in range(nv):
            (vb,) = struct.unpack_from("<i", buf, p)
            p += 4
            vals.append(bytes(buf[p:p + vb]))
            p += vb
        out.append((key, vals))
    assert p == len(buf), (p, len(buf))
    return out


def kmv_encode(recs) -> bytes:
    out = [struct.pack("<Q", len(recs))]
    for k, vals in recs:
        out.append(struct.pack("<iQ", len(k), len(vals)))
        out.append(k)
        for v in vals:
            out.append(struct.pack("<i", len(v)))
            out.append(v)
    return b"".join(out)


def pages_decode(buf: bytes):
    """refdrv pagekv/pagekmv container -> [(nkey, bytes)]"""
    (n,) = struct.unpack_from("<Q", buf, 0)
    p, out = 8, []
    for _ in range(n):
        nkey, asz = struct.unpack_from("<iQ", buf, p)
        p += 12
        out.append((nkey, bytes(buf[p:p + asz])))
        p += asz
    assert p == len(buf)
    return out


def canon_kv(recs):
    """canonical order: by key bytes then value bytes (qsort/hash order is not part of parity)"""
    return sorted(recs)


def canon_kmv(recs, sort_values=False):
    return sorted((k, sorted(v) if sort_values else list(v)) for k, v in recs)


def sha(obj) -> str:
    return hashlib.sha256(repr(obj).encode()).hexdigest()


# ------------------------------------------------------------------------------------ oracle

_lib = None


def oracle_lib():
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "mr_oracle.c")
        if (not os.path.exists(ORACLE_SO)) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"], stdout=subprocess.DEVNULL)
        lib = ctypes.CDLL(ORACLE_SO)
        lib.orc_hashlittle.restype = ctypes.c_uint32
        lib.orc_hashlittle.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_uint32]
        lib.orc_kv_record_size.restype = ctypes.c_uint64
        lib.orc_kmv_record_size.restype = ctypes.c_uint64
        lib.orc_kmv_record_size.argtypes = [ctypes.c_uint64, ctypes.c_int, ctypes.c_uint64, ctypes.c_int,
                                            ctypes.c_int]
        lib.orc_srand48.restype = ctypes.c_uint64
        lib.orc_drand48.restype = ctypes.c_double
        lib.orc_drand48.argtypes = [ctypes.POINTER(ctypes.c_uint64)]
        lib.orc_free.argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def _call_out(fn, *args) -> bytes:
    out = ctypes.c_void_p()
    n = ctypes.c_uint64()
    rc = fn(*args, ctypes.byref(out), ctypes.byref(n))
    if rc != 0:
        raise RuntimeError(f"oracle call failed rc={rc}")
    data = ctypes.string_at(out, n.value)
    oracle_lib().orc_free(out)
    return data


class Oracle:
    """kvfile-bytes in, kvfile/kmvfile-bytes out."""

    def __init__(self):
        self.lib = oracle_lib()

    def hashlittle(self, key: bytes, initval: int) -> int:
        return self.lib.orc_hashlittle(key, len(key), initval)

    def partition(self, kv: bytes, nprocs: int):
        n = struct.unpack_from("<Q", kv, 0)[0]
        out = (ctypes.c_int32 * max(n, 1))()
        self.lib.orc_partition(kv, nprocs, out)
        return list(out[:n])

    def kv_pages(self, kv: bytes, pagesize: int, kalign=4, valign=4) -> bytes:
        return _call_out(self.lib.orc_kv_pages, kv, ctypes.c_uint64(pagesize), kalign, valign)

    def kmv_pages(self, kmv: bytes, pagesize: int, kalign=4, valign=4) -> bytes:
        return _call_out(self.lib.orc_kmv_pages, kmv, ctypes.c_uint64(pagesize), kalign, valign)

    def convert(self, kv: bytes) -> bytes:
        return _call_out(self.lib.orc_convert, kv)

    def reduce(self, kmv: bytes, which: str) -> bytes:
        return _call_out(self.lib.orc_reduce, kmv, {"count": 0, "sum": 1, "cull": 2}[which])

    def sort_keys(self, kv: bytes, flag: int) -> bytes:
        return _call_out(self.lib.orc_sort_keys, kv, flag)

    def sort_values(self, kv: bytes, flag: int) -> bytes:
        """src/mapreduce.cpp:2156-2182: the compare functions of sort_keys applied to the value; the sort is the same
        stable one, so it is sort_keys on the pairs with key and value swapped, swapped back"""
        swapped = self.sort_keys(kv_encode([(v, k) for k, v in kv_decode(kv)]), flag)
        return kv_encode([(k, v) for v, k in kv_decode(swapped)])

    def clone(self, kv: bytes) -> bytes:
        """src/keymultivalue.cpp:344-374: every pair becomes a KMV pair holding its one value, in KV order"""
        return kmv_encode([(k, [v]) for k, v in kv_decode(kv)])

    def collapse(self, kv: bytes, key: bytes) -> bytes:
        """src/keymultivalue.cpp:384-468: ONE KMV pair, values = key_1, value_1, key_2, value_2, ... in KV order"""
        return kmv_encode([(key, [x for k, v in kv_decode(kv) for x in (k, v)])])

    def compress(self, kv: bytes, which: str) -> bytes:
        """src/mapreduce.cpp:749-851: KeyMultiValue::convert of the local pairs, then the callback per distinct key"""
        return self.reduce(self.convert(kv), which)

    def edge_to_vertices(self, kv: bytes, both=True) -> bytes:
        return _call_out(self.lib.orc_map_edge_to_vertices, kv, 1 if both else 0)

    def tokenize(self, text: bytes) -> bytes:
        return _call_out(self.lib.orc_tokenize, text, ctypes.c_uint64(len(text)))

    def href(self, text: bytes, fname: str) -> bytes:
        return _call_out(self.lib.orc_href, text, ctypes.c_uint64(len(text)), fname.encode())

    def intcount_map(self, keys) -> bytes:
        import numpy as np

        keys = np.ascontiguousarray(keys, dtype=np.uint32)
        return _call_out(self.lib.orc_intcount_map, keys.ctypes.data_as(ctypes.c_void_p),
                         ctypes.c_uint64(len(keys)))

    def rmat_generate(self, state: int, ngenerate: int, nlevels: int, a, b, c, d, fraction):
        import numpy as np

        st = ctypes.c_uint64(state)
        edges = np.empty(2 * max(ngenerate, 1), dtype=np.uint64)
        f = self.lib.orc_rmat_generate
        f.restype = None
        f.argtypes = [ctypes.POINTER(ctypes.c_uint64), ctypes.c_uint64, ctypes.c_int] + [ctypes.c_double] * 5 + [
            ctypes.c_void_p]
        f(ctypes.byref(st), ngenerate, nlevels, a, b, c, d, fraction, edges.ctypes.data_as(ctypes.c_void_p))
        return st.value, edges[:2 * ngenerate].reshape(-1, 2)

    def rmat(self, nlevels, nnonzero, a, b, c, d, fraction, seed, nprocs=1):
        """Whole RMAT::run loop (oink/rmat.cpp:37-61) with nprocs rank streams; returns (edge set in
        KV order, iterations)."""
        import numpy as np

        states = [self.lib.orc_srand48(seed + r) for r in range(nprocs)]
        ntotal = (1 << nlevels) * nnonzero
        nremain, kv, niter = ntotal, kv_encode([]), 0
        while nremain:
            niter += 1
            recs = kv_decode(kv)
            for r in range(nprocs):
                ng = nremain // nprocs + (1 if r < nremain % nprocs else 0)
                states[r], e = self.rmat_generate(states[r], ng, nlevels, a, b, c, d, fraction)
                recs += [(e[i].tobytes(), b"") for i in range(len(e))]
            kmv = self.convert(kv_encode(recs))
            kv = self.reduce(kmv, "cull")
            nremain = ntotal - struct.unpack_from("<Q", kv, 0)[0]
        return kv, niter


# ------------------------------------------------------------------------------------ reference driver

def have_ref() -> bool:
    return os.path.exists(REFDRV) and os.access(REFDRV, os.X_OK)


def ref_run(args, stdin: bytes | None = None, timeout=600):
    """Run oracle/_ref/refdrv; returns (list of timing dicts, raw stdout)."""
    env = dict(os.environ)
    with tempfile.TemporaryDirectory() as td:
        env["REFDRV_FPATH"] = td
        p = subprocess.run([REFDRV] + [str(a) for a in args], input=stdin, capture_output=True, timeout=timeout,
                           env=env)
    if p.returncode != 0:
        raise RuntimeError(f"refdrv {args} failed: {p.stdout[-400:]!r} {p.stderr[-400:]!r}")
    timings = []
    for line in p.stdout.decode(errors="replace").splitlines():
        line = line.strip()
        if line.startswith("{"):
            try:
                timings.append(json.loads(line))
            except ValueError:
                pass
    return timings, p.stdout


def ref_ops(kv: bytes, ops, memsize=64, align=None) -> bytes:
    """Load a kvfile into the reference library, apply ops, return the bytes the last dump op wrote."""
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.kv"), os.path.join(td, "out.bin")
        with open(fin, "wb") as f:
            f.write(kv)
        args = ["run", fin, fout, memsize] + ([align] if align else []) + list(ops)
        ref_run(args)
        with open(fout, "rb") as f:
            return f.read()


def ref_hash(pairs):
    """pairs = [(initval, key bytes)] -> list of hashlittle values from src/hash.cpp as compiled."""
    text = "".join(f"{iv} {k.hex()}\n" for iv, k in pairs).encode()
    _, out = ref_run(["hash"], stdin=text)
    return [int(x, 16) for x in out.decode().split()]


# ------------------------------------------------------------------------------------ page parsers

def _ru(x, a):
    return (x + a - 1) // a * a


def kv_page_records(page: bytes, nkey: int, kalign=4, valign=4):
    """Decode a KV page image in the reference layout (src/keyvalue.cpp:345-390).  Padding bytes are
    skipped, not compared: the reference never writes them."""
    talign = max(kalign, valign, 4)
    p, out = 0, []
    for _ in range(nkey):
        kb, vb = struct.unpack_from("<ii", page, p)
        k0 = _ru(p + 8, kalign)
        v0 = _ru(k0 + kb, valign)
        out.append((bytes(page[k0:k0 + kb]), bytes(page[v0:v0 + vb])))
        p = _ru(v0 + vb, talign)
    assert p == len(page), (p, len(page))
    return out


def kmv_page_records(page: bytes, nkey: int, kalign=4, valign=4):
    """Decode a KMV page image (src/keymultivalue.cpp:830-893): [(key, [values])]."""
    talign = max(kalign, valign, 4)
    p, out = 0, []
    for _ in range(nkey):
        nv, kb, mvb = struct.unpack_from("<iii", page, p)
        sizes = struct.unpack_from(f"<{nv}i", page, p + 12)
        k0 = _ru(p + 12 + 4 * nv, kalign)
        v0 = _ru(k0 + kb, valign)
        vals, q = [], v0
        for s in sizes:
            vals.append(bytes(page[q:q + s]))
            q += s
        assert q - v0 == mvb
        out.append((bytes(page[k0:k0 + kb]), vals))
        p = _ru(v0 + mvb, talign)
    assert p == len(page), (p, len(page))
    return out

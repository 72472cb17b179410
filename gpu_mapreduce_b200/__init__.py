"""B200-native MR-MPI hot path: Python host-side mirror over the C-ABI library (include/mrgpu.h).

`MapReduce` mirrors the method names, argument meaning and error behaviour of the reference's
`MAPREDUCE_NS::MapReduce` (src/mapreduce.h:22-131) for the data-parallel path
map -> aggregate -> convert/collate -> reduce -> sort_keys -> scan.  User callbacks of the reference are
replaced by device functors selected by name ("count", "sum", "cull", "edge_to_vertices", ...); pairs
can also be added from the host with `kv_add` (the `KeyValue::add` of a host callback).

There is no CPU implementation here: importing works anywhere, but constructing a `MapReduce` without the
CUDA library and a GPU raises.  All compute is in gpu_mapreduce_b200/libmrgpu.so (built in-tree from
csrc/ by __graft_entry__.build()).
"""
from __future__ import annotations

import ctypes as C
import os
import struct

__all__ = ["MapReduce", "MRError", "lib", "LIB_PATH", "Stats"]

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MRGPU_LIB", os.path.join(_HERE, "libmrgpu.so"))  # MRGPU_LIB: an alternative build, for A/B timing

REDUCE = {"count": 0, "sum": 1, "cull": 2, "postings": 3, "nonzero": 4, "degree": 5, "histo": 6}
MAPMR = {"edge_to_vertices": 0, "edge_to_vertex": 1, "invert": 2}


class MRError(RuntimeError):
    """Raised where the reference would print `ERROR: msg` and exit (src/error.cpp:33-67)."""

    def __init__(self, code, msg):
        super().__init__(f"ERROR: {msg}")
        self.code = code
        self.msg = msg


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("npair", "nvalue", "keysize", "valuesize", "exactsize", "alignsize", "npage")]

    def asdict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


# host callbacks of scan() (src/mapreduce.h:103-105): (key, keybytes, value, valuebytes, ptr) / (key, keybytes, multivalue, nvalues, valuesizes, ptr)
SCAN_KV_FN = C.CFUNCTYPE(None, C.POINTER(C.c_char), C.c_int, C.POINTER(C.c_char), C.c_int, C.c_void_p)
SCAN_KMV_FN = C.CFUNCTYPE(None, C.POINTER(C.c_char), C.c_int, C.POINTER(C.c_char), C.c_int, C.POINTER(C.c_int), C.c_void_p)

_lib = None


def lib():
    """Load libmrgpu.so (fails loudly if it has not been built: there is no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). This package has no CPU or PyTorch fallback.")
    L = C.CDLL(LIB_PATH)
    u64p, vp, i32p = C.POINTER(C.c_uint64), C.c_void_p, C.POINTER(C.c_int32)
    sig = {
        "mrgpu_create": [C.c_int, vp, C.POINTER(vp)],
        "mrgpu_destroy": [vp],
        "mrgpu_abandon": [vp],
        "mrgpu_set": [vp, C.c_char_p, C.c_int64],
        "mrgpu_get": [vp, C.c_char_p, C.POINTER(C.c_int64)],
        "mrgpu_comm_unique_id": [vp],
        "mrgpu_comm_init": [vp, C.c_int, C.c_int, vp],
        "mrgpu_comm_rank": [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)],
        "mrgpu_kv_open": [vp, C.c_int],
        "mrgpu_kv_add": [vp, C.c_uint64, vp, i32p, C.c_int, vp, i32p, C.c_int],
        "mrgpu_kv_close": [vp, u64p],
        "mrgpu_map_intcount": [vp, vp, C.c_uint64, C.c_int, u64p],
        "mrgpu_map_tokenize": [vp, vp, C.c_uint64, C.c_int, u64p],
        "mrgpu_map_href": [vp, vp, C.c_uint64, C.c_char_p, C.c_int, u64p],
        "mrgpu_map_rmat": [vp, C.c_uint64, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, u64p,
                           C.c_int, u64p],
        "mrgpu_map_mr": [vp, vp, C.c_int, C.c_int, u64p],
        "mrgpu_add": [vp, vp, u64p],
        "mrgpu_copy": [vp, vp],
        "mrgpu_aggregate": [vp, u64p],
        "mrgpu_aggregate_dest": [vp, vp, C.c_uint64, u64p],
        "mrgpu_scan_kv": [vp, SCAN_KV_FN, vp, u64p],
        "mrgpu_scan_kmv": [vp, SCAN_KMV_FN, vp, u64p],
        "mrgpu_multivalue_blocks": [vp, C.POINTER(C.c_int), u64p],
        "mrgpu_multivalue_block_select": [vp, C.c_int],
        "mrgpu_multivalue_block": [vp, C.c_int, C.POINTER(C.POINTER(C.c_char)), C.POINTER(C.POINTER(C.c_int)), C.POINTER(C.c_int)],
        "mrgpu_plan_exchange": [C.c_int, C.c_int, u64p, u64p, u64p, u64p, u64p, u64p, C.POINTER(C.c_int)],
        "mrgpu_debug_partition": [vp, C.c_int, u64p],
        "mrgpu_convert": [vp, u64p],
        "mrgpu_collate": [vp, u64p],
        "mrgpu_reduce": [vp, C.c_int, u64p],
        "mrgpu_sort_keys": [vp, C.c_int, u64p],
        "mrgpu_sort_values": [vp, C.c_int, u64p],
        "mrgpu_compress": [vp, C.c_int, u64p],
        "mrgpu_gather": [vp, C.c_int, u64p],
        "mrgpu_broadcast": [vp, C.c_int, u64p],
        "mrgpu_clone": [vp, u64p],
        "mrgpu_collapse": [vp, C.c_char_p, C.c_int, u64p],
        "mrgpu_scrunch": [vp, C.c_int, C.c_char_p, C.c_int, u64p],
        "mrgpu_global_count": [vp, C.c_int, u64p],
        "mrgpu_kv_stats": [vp, C.POINTER(Stats)],
        "mrgpu_kmv_stats": [vp, C.POINTER(Stats)],
        "mrgpu_kv_page_info": [vp, C.c_uint64, u64p, u64p, u64p, u64p],
        "mrgpu_kv_page_copy": [vp, C.c_uint64, vp],
        "mrgpu_kmv_page_info": [vp, C.c_uint64, u64p, u64p, C.POINTER(C.c_int)],
        "mrgpu_kmv_page_copy": [vp, C.c_uint64, vp],
        "mrgpu_kv_device_view": [vp, C.POINTER(vp), u64p, C.POINTER(vp), u64p],
        "mrgpu_kmv_device_view": [vp, C.POINTER(vp), u64p, C.POINTER(vp)],
        "mrgpu_prof": [vp, C.c_int],
        "mrgpu_prof_report": [vp, C.c_char_p, C.c_uint64, u64p],
        "mrgpu_hashlittle": [C.c_int, vp, vp, C.c_uint64, C.c_uint32, vp],
        "mrgpu_sort_pairs": [C.c_int, vp, vp, vp, C.c_uint64, C.c_int, C.c_int],
        "mrgpu_sort_pairs_noalloc": [C.c_int, vp, vp, vp, C.c_uint64, C.c_int, C.c_int],
    }
    for name, args in sig.items():
        f = getattr(L, name)
        f.argtypes = args
        f.restype = C.c_int
    L.mrgpu_last_error.argtypes = [vp]
    L.mrgpu_last_error.restype = C.c_char_p
    _lib = L
    return L


def _ptr(x):
    """data pointer of a torch tensor / numpy array / bytes-like / int"""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if hasattr(x, "ctypes"):
        return x.ctypes.data
    if isinstance(x, (bytes, bytearray, memoryview)):
        import numpy as np

        a = np.frombuffer(x, dtype=np.uint8)  # zero-copy view; valid while the caller holds x
        return a.ctypes.data if a.size else None
    raise TypeError(type(x))


class MapReduce:
    """One MR object = one rank's KeyValue or KeyMultiValue in HBM (src/mapreduce.h:43-44)."""

    _SETTINGS = ("mapstyle", "all2all", "verbosity", "timer", "memsize", "minpage", "maxpage", "freepage",
                 "outofcore", "zeropage", "keyalign", "valuealign",
                 "compact", "ahead", "l2pin", "combine", "presorted", "hbm_budget", "debug_fail_alloc")  # not reference members: 0 = stores always hold the reference's record streams (see mrgpu.h)

    def __init__(self, device: int = 0, stream: int | None = None, rank: int = 0, nranks: int = 1,
                 unique_id: bytes | None = None):
        object.__setattr__(self, "_h", None)
        L = lib()
        h = C.c_void_p()
        rc = L.mrgpu_create(device, stream, C.byref(h))
        if rc != 0:
            raise MRError(rc, L.mrgpu_last_error(None).decode())
        object.__setattr__(self, "_h", h)
        object.__setattr__(self, "device", device)
        if nranks > 1:
            self._ck(L.mrgpu_comm_init(h, rank, nranks, unique_id))

    # -- plumbing
    def _ck(self, rc):
        if rc != 0:
            raise MRError(rc, lib().mrgpu_last_error(self._h).decode())

    def __setattr__(self, name, value):
        if name in self._SETTINGS:  # public int members of the reference object (src/mapreduce.h:28-39)
            self._ck(lib().mrgpu_set(self._h, name.encode(), int(value)))
        else:
            object.__setattr__(self, name, value)

    def __getattr__(self, name):
        if name in MapReduce._SETTINGS:
            v = C.c_int64()
            self._ck(lib().mrgpu_get(self._h, name.encode(), C.byref(v)))
            return v.value
        raise AttributeError(name)

    def destroy(self):
        """~MapReduce (src/mapreduce.cpp:169).  COLLECTIVE once aggregate() has exchanged pairs between ranks, like the
        reference's destructor (MPI_Comm_free): call it explicitly, on every rank, in the same order."""
        if self._h:
            lib().mrgpu_destroy(self._h)
            object.__setattr__(self, "_h", None)

    def __del__(self):
        # The garbage collector runs finalisers in no particular order across ranks, so it must not enter a
        # collective: mrgpu_abandon releases this rank's side without a barrier (see include/mrgpu.h).
        try:
            if self._h:
                lib().mrgpu_abandon(self._h)
                object.__setattr__(self, "_h", None)
        except Exception:
            pass

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        rc = lib().mrgpu_comm_unique_id(buf)
        if rc != 0:
            raise MRError(rc, lib().mrgpu_last_error(None).decode())
        return buf.raw

    def pool_hiwater(self):
        """msizemax (src/mapreduce.h:54): high-water mark, in bytes, of the HBM this handle's page pool had in use"""
        v = C.c_int64()
        self._ck(lib().mrgpu_get(self._h, b"msizemax", C.byref(v)))
        return v.value

    def my_proc(self):
        r, n = C.c_int(), C.c_int()
        lib().mrgpu_comm_rank(self._h, C.byref(r), C.byref(n))
        return r.value

    def num_procs(self):
        r, n = C.c_int(), C.c_int()
        lib().mrgpu_comm_rank(self._h, C.byref(r), C.byref(n))
        return n.value

    def _count(self, fn, *args):
        n = C.c_uint64()
        self._ck(fn(self._h, *args, C.byref(n)))
        return n.value

    # -- map flavours (device functors)
    def map_intcount(self, keys, n=None, addflag=0):
        """map(nmap, mymap, ptr, addflag) with cpu/IntCount.cpp:177-180 as the emitter; `keys` = uint32 array
        (torch CUDA/CPU tensor, numpy array or raw pointer + n)."""
        if n is None:
            n = keys.numel() if hasattr(keys, "numel") else len(keys)
        return self._count(lib().mrgpu_map_intcount, _ptr(keys), n, addflag)

    def map_tokenize(self, text, nbytes=None, addflag=0):
        """map(files, fileread) with examples/wordfreq.cpp:104-130; one call per file."""
        keep = text
        if nbytes is None:
            nbytes = text.numel() if hasattr(text, "numel") else len(text)
        return self._count(lib().mrgpu_map_tokenize, _ptr(keep), nbytes, addflag)

    def map_href(self, text, fname: str, nbytes=None, addflag=0):
        """map with the InvertedIndex scanner cpu/InvertedIndex.cpp:196-270; one call per file."""
        keep = text
        if nbytes is None:
            nbytes = text.numel() if hasattr(text, "numel") else len(text)
        return self._count(lib().mrgpu_map_href, _ptr(keep), nbytes, fname.encode(), addflag)

    def map_rmat(self, ngenerate, nlevels, a, b, c, d, fraction, state48: int, addflag=1):
        """map(nprocs, rmat_generate, &rmat, 1) oink/rmat.cpp:57; returns (global nkv, new drand48 state)."""
        st = C.c_uint64(state48)
        n = C.c_uint64()
        self._ck(lib().mrgpu_map_rmat(self._h, ngenerate, nlevels, a, b, c, d, fraction, C.byref(st), addflag, C.byref(n)))
        return n.value, st.value

    def map_mr(self, src: "MapReduce", functor: str, addflag=0):
        """map(mr2, mymap, ptr, addflag) src/mapreduce.cpp:1560 with a stock OINK callback."""
        return self._count(lib().mrgpu_map_mr, src._h, MAPMR[functor], addflag)

    # -- host staging (KeyValue::add from a host callback)
    def kv_add(self, pairs, addflag=0):
        """open(addflag); kv->add(key,kb,value,vb) for each (key, value) bytes pair; close()."""
        L = lib()
        self._ck(L.mrgpu_kv_open(self._h, addflag))
        if pairs:
            keys = b"".join(k for k, _ in pairs)
            vals = b"".join(v for _, v in pairs)
            kb = (C.c_int32 * len(pairs))(*[len(k) for k, _ in pairs])
            vb = (C.c_int32 * len(pairs))(*[len(v) for _, v in pairs])
            self._ck(L.mrgpu_kv_add(self._h, len(pairs), keys, kb, -1, vals, vb, -1))
        n = C.c_uint64()
        self._ck(L.mrgpu_kv_close(self._h, C.byref(n)))
        return n.value

    # -- the hot path
    def add(self, other: "MapReduce"):
        return self._count(lib().mrgpu_add, other._h)

    def copy(self):
        """copy() src/mapreduce.cpp:290: a new object with the same settings and a deep copy of the KV / KMV"""
        m = MapReduce(device=self.device)
        m._ck(lib().mrgpu_copy(m._h, self._h))
        return m

    def aggregate(self, hash=None):
        """aggregate(myhash) src/mapreduce.cpp:385.  hash = None: the default hashlittle partition, on the GPU.  A callable
        hash(key: bytes, keybytes: int) -> int is host code like the reference's function pointer: it is evaluated here for
        every pair (destination = hash % nprocs, src/mapreduce.cpp:470) and the exchange runs with those destinations."""
        if hash is None:
            return self._count(lib().mrgpu_aggregate)
        import numpy as np

        P = self.num_procs()
        dest = np.array([hash(k, len(k)) % P for k, _ in self.scan_kv()], dtype=np.int32)
        return self._count(lib().mrgpu_aggregate_dest, dest.ctypes.data if dest.size else None, dest.size)

    def debug_partition(self, nprocs):
        """local half of aggregate() for a virtual rank count (test / profiling hook); returns per-destination counts"""
        c = (C.c_uint64 * nprocs)()
        self._ck(lib().mrgpu_debug_partition(self._h, nprocs, c))
        return list(c)

    def convert(self):
        return self._count(lib().mrgpu_convert)

    def collate(self, hash=None):
        """collate(myhash) src/mapreduce.cpp:710-738 = aggregate(myhash) + convert()"""
        if hash is not None:
            self.aggregate(hash)
            return self.convert()
        return self._count(lib().mrgpu_collate)

    def reduce(self, functor: str):
        return self._count(lib().mrgpu_reduce, REDUCE[functor])

    def sort_keys(self, flag: int):
        return self._count(lib().mrgpu_sort_keys, int(flag))

    def sort_values(self, flag: int):
        """sort_values(flag) src/mapreduce.cpp:2156: per-rank stable sort of the pairs by value."""
        return self._count(lib().mrgpu_sort_values, int(flag))

    def compress(self, functor: str):
        """compress(mycompress) src/mapreduce.cpp:749 with a stock callback: local convert + reduce, no exchange."""
        return self._count(lib().mrgpu_compress, REDUCE[functor])

    def broadcast(self, root: int):
        """broadcast(root) src/mapreduce.cpp:569: every rank gets a copy of root's KeyValue."""
        return self._count(lib().mrgpu_broadcast, int(root))

    def clone(self):
        """clone() src/mapreduce.cpp:631: KV -> KMV, one value per pair."""
        return self._count(lib().mrgpu_clone)

    def collapse(self, key: bytes):
        """collapse(key) src/mapreduce.cpp:681: all pairs -> one KMV pair (key_1, value_1, key_2, ...)."""
        return self._count(lib().mrgpu_collapse, key, len(key))

    def scrunch(self, numprocs: int, key: bytes):
        """scrunch(numprocs, key) src/mapreduce.cpp:2075: gather then collapse."""
        return self._count(lib().mrgpu_scrunch, int(numprocs), key, len(key))

    def gather(self, numprocs: int):
        """gather(numprocs) src/mapreduce.cpp:893: pairs of ranks >= numprocs move to rank me % numprocs."""
        return self._count(lib().mrgpu_gather, int(numprocs))

    # -- stats, pages, scan
    def kv_stats(self, level=0):
        s = Stats()
        self._ck(lib().mrgpu_kv_stats(self._h, C.byref(s)))
        return s

    def kmv_stats(self, level=0):
        s = Stats()
        self._ck(lib().mrgpu_kmv_stats(self._h, C.byref(s)))
        return s

    def kv_pages(self):
        """[(nkey, bytes)] in the reference's page layout (KeyValue::request_page)."""
        L = lib()
        out = []
        for ip in range(self.kv_stats().npage):
            nk, ks, vs, asz = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint64()
            self._ck(L.mrgpu_kv_page_info(self._h, ip, C.byref(nk), C.byref(ks), C.byref(vs), C.byref(asz)))
            buf = C.create_string_buffer(max(asz.value, 1))
            self._ck(L.mrgpu_kv_page_copy(self._h, ip, buf))
            out.append((nk.value, buf.raw[:asz.value]))
        return out

    def kmv_pages(self):
        """[(nkey, bytes, is_extended)] in the reference's KMV page layout."""
        L = lib()
        out = []
        for ip in range(self.kmv_stats().npage):
            nk, asz, ext = C.c_uint64(), C.c_uint64(), C.c_int()
            self._ck(L.mrgpu_kmv_page_info(self._h, ip, C.byref(nk), C.byref(asz), C.byref(ext)))
            buf = C.create_string_buffer(max(asz.value, 1))
            self._ck(L.mrgpu_kmv_page_copy(self._h, ip, buf))
            out.append((nk.value, buf.raw[:asz.value], ext.value))
        return out

    def scan_kv(self, fn=None):
        """scan(myscan, ptr) src/mapreduce.cpp:1933-1976 through mrgpu_scan_kv: fn(key, keybytes, value, valuebytes) for every
        pair of this rank; returns the list of (key, value) when fn is None."""
        out = []

        def cb(k, kb, v, vb, _):
            kk, vv = C.string_at(k, kb), C.string_at(v, vb)
            if fn is None:
                out.append((kk, vv))
            else:
                fn(kk, kb, vv, vb)

        self._ck(lib().mrgpu_scan_kv(self._h, SCAN_KV_FN(cb), None, None))
        return out

    def scan_kmv(self, fn=None):
        """scan(myscan, ptr) KMV form src/mapreduce.cpp:1984-2065 through mrgpu_scan_kmv: [(key, [values])].  A pair bigger
        than a page arrives through the multivalue_blocks()/multivalue_block() protocol and is put together here."""
        L = lib()
        out = []

        def split(mv, sizes, nv):
            vals, q = [], 0
            raw = C.string_at(mv, sum(sizes[i] for i in range(nv)))
            for i in range(nv):
                vals.append(raw[q:q + sizes[i]])
                q += sizes[i]
            return vals

        def cb(k, kb, mv, nv, sizes, _):
            key = C.string_at(k, kb)
            if mv:
                vals = split(mv, sizes, nv)
            else:   # multivalue == NULL: CHECK_FOR_BLOCKS / BEGIN_BLOCK_LOOP of oink/blockmacros.h:12-32
                nblock, total = C.c_int(), C.c_uint64()
                self._ck(L.mrgpu_multivalue_blocks(self._h, C.byref(nblock), C.byref(total)))
                vals = []
                for ib in range(nblock.value):
                    pm, ps, n = C.POINTER(C.c_char)(), C.POINTER(C.c_int)(), C.c_int()
                    self._ck(L.mrgpu_multivalue_block(self._h, ib, C.byref(pm), C.byref(ps), C.byref(n)))
                    vals += split(pm, ps, n.value)
                assert len(vals) == total.value
            if fn is None:
                out.append((key, vals))
            else:
                fn(key, kb, b"".join(vals), len(vals), [len(v) for v in vals])

        self._ck(L.mrgpu_scan_kmv(self._h, SCAN_KMV_FN(cb), None, None))
        return out

    def prof(self, enable=True):
        """reset launch counters; with enable, time every tagged kernel (cf. the reference's `timer`)"""
        self._ck(lib().mrgpu_prof(self._h, 1 if enable else 0))

    def prof_report(self):
        """(list of {tag, ms, bytes, launches, calls}, total kernel launches since prof())"""
        import json

        buf = C.create_string_buffer(1 << 16)
        n = C.c_uint64()
        self._ck(lib().mrgpu_prof_report(self._h, buf, len(buf), C.byref(n)))
        return json.loads(buf.value.decode()), n.value

    def kv_device_view(self):
        d, n, o, r = C.c_void_p(), C.c_uint64(), C.c_void_p(), C.c_uint64()
        self._ck(lib().mrgpu_kv_device_view(self._h, C.byref(d), C.byref(n), C.byref(o), C.byref(r)))
        return d.value, n.value, o.value, r.value

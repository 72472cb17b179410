"""GPU parity tests (run on the B200 box with -m gpu): the CUDA path, called through the C-ABI of
gpu_mapreduce_b200/libmrgpu.so, against the oracle and the fixtures recorded from the reference library.
Integer / byte work: everything is compared bit-exact."""
import ctypes
import struct

import numpy as np
import pytest

import mrhelpers as H
from gpu_mapreduce_b200 import workloads as W

pytestmark = pytest.mark.gpu


@pytest.fixture()
def MR():
    import gpu_mapreduce_b200 as g

    made = []

    def make(**settings):
        mr = g.MapReduce(device=0)
        for k, v in settings.items():
            setattr(mr, k, v)
        made.append(mr)
        return mr

    yield make
    for mr in made:
        mr.destroy()


def _torch():
    import torch

    return torch


# ------------------------------------------------------------------------------------------- primitives

@pytest.mark.parametrize("n", [0, 1, 31, 4095, 4096, 4097, 100_003, 1_500_000])
@pytest.mark.parametrize("kbytes,pbytes", [(4, 0), (4, 4), (4, 8), (8, 0), (8, 4), (8, 8)])
def test_sort_pairs_matches_stable_argsort(n, kbytes, pbytes):
    import gpu_mapreduce_b200 as g

    torch = _torch()
    r = np.random.RandomState(n % 1000 + kbytes + pbytes)
    kd = np.uint32 if kbytes == 4 else np.uint64
    # mixture: few distinct values (long runs), constant high bytes, some full-range keys
    keys = r.randint(0, 1 << 20, size=n).astype(kd)
    if n > 10:
        keys[::7] = r.randint(0, 5, size=len(keys[::7]))
        keys[::13] = np.iinfo(kd).max
        keys[::11] = r.randint(0, 1 << 31, size=len(keys[::11])).astype(kd) << (kbytes * 8 - 31)
    pay = None
    if pbytes:
        pd = np.uint32 if pbytes == 4 else np.uint64
        pay = np.arange(n, dtype=pd) * 3 + 1
    tk = torch.from_numpy(keys.view(np.int32 if kbytes == 4 else np.int64).copy()).cuda()
    tp = torch.from_numpy(pay.view(np.int32 if pbytes == 4 else np.int64).copy()).cuda() if pbytes else None
    rc = g.lib().mrgpu_sort_pairs(0, None, tk.data_ptr(), tp.data_ptr() if pbytes else None, n, kbytes, pbytes)
    assert rc == 0
    order = np.argsort(keys, kind="stable")
    got_k = tk.cpu().numpy().view(kd)
    assert np.array_equal(got_k, keys[order])
    if pbytes:
        assert np.array_equal(tp.cpu().numpy().view(pay.dtype), pay[order])


def test_hashlittle_kernel_matches_reference_kat(golden, oracle):
    import gpu_mapreduce_b200 as g

    by_iv = {}
    for iv, hexkey, h in golden.manifest["hash"]:
        by_iv.setdefault(iv, []).append((bytes.fromhex(hexkey), h))
    r = np.random.RandomState(5)
    for iv, items in by_iv.items():
        keys = [k for k, _ in items] + [bytes(r.randint(0, 256, size=r.randint(0, 90)).astype(np.uint8)) for _ in range(200)]
        want = [h for _, h in items] + [oracle.hashlittle(k, iv) for k in keys[len(items):]]
        offs = np.zeros(len(keys) + 1, dtype=np.uint64)
        offs[1:] = np.cumsum([len(k) for k in keys])
        blob = np.frombuffer(b"".join(keys) + b"\0", dtype=np.uint8)
        out = np.zeros(len(keys), dtype=np.uint32)
        rc = g.lib().mrgpu_hashlittle(0, blob.ctypes.data, offs.ctypes.data, len(keys), iv, out.ctypes.data)
        assert rc == 0
        assert out.tolist() == want


# ------------------------------------------------------------------------------------------- IntCount

def _intcount_gpu(mr, keys, device_input):
    torch = _torch()
    if device_input:
        t = torch.from_numpy(keys.view(np.int32).copy()).cuda()
        n = mr.map_intcount(t)
    else:
        n = mr.map_intcount(keys)
    assert n == len(keys)
    return n


@pytest.mark.parametrize("device_input", [True, False])
def test_intcount_golden_pipeline(MR, golden, device_input):
    mr = MR()
    keys = np.frombuffer(golden["intcount.in"], dtype=np.uint32)
    _intcount_gpu(mr, keys, device_input)
    # map output: byte-identical KV page vs the oracle's page image (records are fully specified, no padding)
    pages = mr.kv_pages()
    orc = H.Oracle()
    assert pages == H.pages_decode(orc.kv_pages(orc.intcount_map(keys), 64 << 20))
    assert mr.aggregate() == len(keys)
    nu = mr.convert()
    st = mr.kmv_stats()
    assert st.npair == nu and st.nvalue == len(keys)
    assert mr.reduce("count") == nu
    assert mr.sort_keys(1) == nu
    assert H.kv_encode(mr.scan_kv()) == golden["intcount.out"]


@pytest.mark.parametrize("n,nd", [(1, 1), (2, 1), (5000, 1), (70_000, 3000), (300_000, 1 << 16)])
def test_intcount_vs_oracle_sizes(MR, oracle, n, nd):
    mr = MR()
    keys = W.intcount_keys_np(0, n, ndistinct=nd, seed=n)
    if n > 100:
        keys = keys.copy()
        keys[::97] = 0xFFFFFFF0 + (np.arange(len(keys[::97])) % 7).astype(np.uint32)  # keys with the sign bit set
    _intcount_gpu(mr, keys, True)
    mr.convert()
    # KMV content: canonical key order, values in arrival order (all 1s here, sizes all 4)
    kmv = mr.scan_kmv()
    want = H.canon_kmv(H.kmv_decode(oracle.convert(oracle.intcount_map(keys))))
    assert H.canon_kmv(kmv) == want
    mr.reduce("count")
    mr.sort_keys(1)
    ref = oracle.sort_keys(oracle.reduce(oracle.convert(oracle.intcount_map(keys)), "count"), 1)
    assert H.kv_encode(mr.scan_kv()) == ref


def test_intcount_empty_and_errors(MR):
    import gpu_mapreduce_b200 as g

    mr = MR()
    with pytest.raises(g.MRError, match="Cannot convert without KeyValue"):
        mr.convert()
    with pytest.raises(g.MRError, match="Cannot reduce without KeyMultiValue"):
        mr.reduce("count")
    with pytest.raises(g.MRError, match="Cannot aggregate without KeyValue"):
        mr.aggregate()
    assert mr.map_intcount(np.zeros(0, dtype=np.uint32)) == 0
    assert mr.kv_stats().npage == 1
    assert mr.convert() == 0
    assert mr.reduce("count") == 0
    assert mr.sort_keys(1) == 0
    assert mr.scan_kv() == []
    with pytest.raises(g.MRError, match="Invalid compare method"):
        mr.sort_keys(9)


def test_intcount_alignment8_and_pages(MR, oracle):
    keys = W.intcount_keys_np(0, 5000, ndistinct=100)
    mr = MR(keyalign=8, valuealign=8, memsize=-32768)
    mr.map_intcount(keys)
    got = mr.kv_pages()
    want = H.pages_decode(oracle.kv_pages(oracle.intcount_map(keys), 32768, 8, 8))
    assert len(want) > 1
    assert got == want
    mr.convert()
    gk = []
    for nkey, page, ext in mr.kmv_pages():
        assert not ext
        gk += H.kmv_page_records(page, nkey, 8, 8)
    assert H.canon_kmv(gk) == H.canon_kmv(H.kmv_decode(oracle.convert(oracle.intcount_map(keys))))
    mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(oracle.reduce(oracle.convert(oracle.intcount_map(keys)), "count")))


def test_intcount_addflag_appends(MR, oracle):
    a = W.intcount_keys_np(0, 3000, ndistinct=50, seed=1)
    b = W.intcount_keys_np(0, 2000, ndistinct=50, seed=2)
    mr = MR()
    mr.map_intcount(a)
    assert mr.map_intcount(b, addflag=1) == 5000
    assert mr.scan_kv() == H.kv_decode(oracle.intcount_map(np.concatenate([a, b])))
    assert mr.map_intcount(b, addflag=0) == 2000


# ------------------------------------------------------------------------------------------- convert / reduce

def _load(mr, kvfile):
    recs = H.kv_decode(kvfile)
    assert mr.kv_add(recs) == len(recs)
    return recs


def test_kv_add_pages_match_reference_layout(MR, golden, oracle):
    mr = MR(memsize=-32768)
    recs = _load(mr, golden["pages_multi.in"])
    got = mr.kv_pages()
    ref = H.pages_decode(golden["pages_multi.kv"])
    assert [(n, len(b)) for n, b in got] == [(n, len(b)) for n, b in ref]
    out = []
    for n, pg in got:
        out += H.kv_page_records(pg, n)
    assert out == recs
    st = mr.kv_stats()
    assert st.npair == len(recs) and st.npage == len(ref)
    assert st.exactsize == 8 * len(recs) + sum(len(k) + len(v) for k, v in recs)


def test_convert_general_golden(MR, golden):
    """variable-length keys and values: fingerprint path; key order canonical, value order = arrival"""
    mr = MR()
    _load(mr, golden["convert.in"])
    ref = H.kmv_decode(golden["convert.kmv"])
    assert mr.convert() == len(ref)
    assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(ref)
    st = mr.kmv_stats()
    assert st.keysize == sum(len(k) for k, _ in ref)
    assert st.valuesize == sum(len(v) for _, vs in ref for v in vs)


@pytest.mark.parametrize("cb", ["count", "sum", "cull"])
def test_reduce_callbacks_golden(MR, golden, cb):
    mr = MR()
    _load(mr, golden["convert.in"])
    mr.collate()
    ref = H.kv_decode(golden[f"convert.reduce_{cb}"])
    assert mr.reduce(cb) == len(ref)
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(ref)


def test_b2_page_images(MR, golden):
    mr = MR()
    _load(mr, golden["pages_b2.in"])
    assert mr.kv_pages() == H.pages_decode(golden["pages_b2.kv.a4"])
    mr.convert()
    (nkey, page, ext), = mr.kmv_pages()
    ref_nkey, ref_page = H.pages_decode(golden["pages_b2.kmv.a4"])[0]
    # same records; our key order is canonical, the reference's is first appearance ("a","bb","ccc" in both here)
    assert H.canon_kmv(H.kmv_page_records(page, nkey)) == H.canon_kmv(H.kmv_page_records(ref_page, ref_nkey))
    assert len(page) == len(ref_page)


UNIFORM_CASES = [(4, 0), (4, 4), (4, 8), (8, 0), (8, 4), (3, 5), (16, 0), (16, 4), (12, 12), (1, 0), (24, 8)]


@pytest.mark.parametrize("kb,vb", UNIFORM_CASES)
def test_convert_uniform_shapes(MR, oracle, kb, vb):
    """every uniform (keybytes, valuebytes) route: exact 32/64-bit keys, compacted long keys, gathered values"""
    r = np.random.RandomState(kb * 100 + vb)
    n = 6000
    base = r.randint(0, 256, size=kb).astype(np.uint8)
    recs = []
    for i in range(n):
        k = base.copy()
        # vary at most 3 byte positions so long keys stay compactable; kb=24 case varies 10 -> general path
        nvary = min(kb, 10 if kb == 24 else 3)
        for q in range(nvary):
            k[(q * 5) % kb] = r.randint(0, 6)
        v = struct.pack("<I", i)[:vb] if vb <= 4 else (struct.pack("<I", i) * 3)[:vb]
        recs.append((bytes(k), v))
    mr = MR()
    mr.kv_add(recs)
    want = H.kmv_decode(oracle.convert(H.kv_encode(recs)))
    assert mr.convert() == len(want)
    assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(want)
    mr.reduce("cull" if vb else "count")
    wantr = H.kv_decode(oracle.reduce(oracle.convert(H.kv_encode(recs)), "cull" if vb else "count"))
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(wantr)


@pytest.mark.parametrize("kb,vb", [(4, 4), (8, 8), (4, 8), (16, 4)])
def test_convert_constant_values_are_not_sorted_but_still_written(MR, oracle, kb, vb):
    """every pair carries the same value bytes: the sort skips the payload, the KMV must still hold every value"""
    r = np.random.RandomState(kb + vb)
    val = bytes(r.randint(1, 255, size=vb).astype(np.uint8))
    recs = [(bytes(r.randint(0, 3, size=kb).astype(np.uint8)), val) for _ in range(9000)]
    mr = MR()
    mr.kv_add(recs)
    want = H.kmv_decode(oracle.convert(H.kv_encode(recs)))
    assert mr.convert() == len(want)
    assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(want)
    # one differing value at the very end must switch the payload back on
    recs[-1] = (recs[-1][0], bytes(vb))
    mr2 = MR()
    mr2.kv_add(recs)
    mr2.convert()
    assert H.canon_kmv(mr2.scan_kmv()) == H.canon_kmv(H.kmv_decode(oracle.convert(H.kv_encode(recs))))


@pytest.mark.parametrize("kb,vb", [(4, 4), (4, 0), (8, 0), (8, 8), (4, 8)])
def test_convert_dense_integer_keys(MR, oracle, kb, vb):
    """small integer keys + identical values: grouping by counting (no sort); result must equal the general path's"""
    r = np.random.RandomState(100 + kb + vb)
    n = 150_000
    kd = "<I" if kb == 4 else "<Q"
    keys = (r.zipf(1.2, size=n) % 5000).astype(np.int64)
    keys[:10] = [0, 1, (1 << 26) - 1, 7, 7, 7, 12345, 0, 4999, 2]
    val = bytes(r.randint(1, 255, size=vb).astype(np.uint8))
    recs = [(struct.pack(kd, int(k)), val) for k in keys]
    mr = MR()
    mr.kv_add(recs)
    want = H.kmv_decode(oracle.convert(H.kv_encode(recs)))
    assert mr.convert() == len(want)
    st = mr.kmv_stats()
    assert st.nvalue == n and st.keysize == kb * len(want) and st.valuesize == vb * n
    assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(want)
    mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(oracle.reduce(oracle.convert(H.kv_encode(recs)), "count")))


@pytest.mark.parametrize("spoil", ["bigkey", "value", "highword"])
def test_convert_dense_falls_back_to_the_sort(MR, oracle, spoil):
    """one key outside the table / one differing value anywhere: the counting path must hand over to the sort"""
    r = np.random.RandomState(7)
    n = 120_000
    kb = 8 if spoil == "highword" else 4
    kd = "<I" if kb == 4 else "<Q"
    recs = [(struct.pack(kd, int(k)), struct.pack("<i", 1)) for k in r.randint(0, 3000, size=n)]
    if spoil == "bigkey":
        recs[n // 2] = (struct.pack("<I", 1 << 26), recs[0][1])
    elif spoil == "value":
        recs[n - 1] = (recs[n - 1][0], struct.pack("<i", 2))
    else:
        recs[17] = (struct.pack("<Q", 5 | (1 << 40)), recs[0][1])
    mr = MR()
    mr.kv_add(recs)
    want = H.kmv_decode(oracle.convert(H.kv_encode(recs)))
    assert mr.convert() == len(want)
    assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(want)


def test_convert_hot_key_bigger_than_page(MR, golden, oracle):
    """one key with ~18000 values and 32 KiB pages -> the pair gets its own (extended) page"""
    mr = MR(memsize=-32768)
    _load(mr, golden["extended.in"])
    mr.convert()
    pages = mr.kmv_pages()
    assert any(ext for _, _, ext in pages)
    mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(golden["extended.count"]))


@pytest.mark.parametrize("kb,vb", [(4, 4), (4, 0), (8, 0), (8, 4), (8, 8), (16, 0)])
@pytest.mark.parametrize("P", [2, 3, 8, 33])
def test_single_pass_partition_matches_reference_placement(MR, oracle, kb, vb, P):
    """local half of aggregate(): every pair in the slice of rank hashlittle(key,kb,P) % P, original order kept"""
    r = np.random.RandomState(kb * 7 + vb + P)
    n = 5000 if P == 33 else 21000
    recs = [(bytes(r.randint(0, 256, size=kb).astype(np.uint8)) if i % 3 else struct.pack("<I", 7) * (kb // 4),
             struct.pack("<I", i)[:vb] if vb <= 4 else struct.pack("<II", i, ~i & 0xFFFFFFFF)) for i in range(n)]
    mr = MR()
    mr.kv_add(recs)
    counts = mr.debug_partition(P)
    dest = oracle.partition(H.kv_encode(recs), P)
    want = [rec for p in range(P) for rec, d in zip(recs, dest) if d == p]
    assert counts == [dest.count(p) for p in range(P)]
    assert mr.scan_kv() == want


def test_convert_dense_table_is_clean_between_calls(MR, oracle):
    """the 2^26-slot count table belongs to the handle: a success, an abandoned attempt and another success in a row"""
    mr = MR()
    r = np.random.RandomState(3)
    for rnd, hi in enumerate([5000, 1 << 27, 70000]):
        keys = r.randint(0, 4000 + 1000 * rnd, size=100_000).astype(np.uint32)
        keys[12345] = hi                                   # round 1: a key the table cannot take
        mr.map_intcount(keys)
        want = H.kv_decode(oracle.reduce(oracle.convert(oracle.intcount_map(keys)), "count"))
        assert mr.convert() == len(want)
        mr.reduce("count")
        assert H.canon_kv(mr.scan_kv()) == H.canon_kv(want)


@pytest.mark.parametrize("ahead", [0, 1])
@pytest.mark.parametrize("nth", [0, 1, 2])
def test_convert_dense_survives_an_allocation_failure(MR, oracle, ahead, nth):
    """an out-of-memory between the count and the compaction that re-zeroes the table (injected: the nth next pool
    allocation fails) must not leave stale counts behind: the handle stays usable and the next convert is exact"""
    import gpu_mapreduce_b200 as g

    mr = MR(ahead=ahead)
    r = np.random.RandomState(11 + nth)
    keys = r.randint(0, 9000, size=120_000).astype(np.uint32)
    mr.map_intcount(_torch().from_numpy(keys.view(np.int32)).cuda())
    mr.debug_fail_alloc = nth + (0 if ahead else 0)
    with pytest.raises(g.MRError, match="out of GPU memory"):
        mr.convert()
    mr.debug_fail_alloc = -1
    keys2 = r.randint(3000, 12000, size=150_000).astype(np.uint32)    # overlaps the first key range
    mr.map_intcount(_torch().from_numpy(keys2.view(np.int32)).cuda())
    want = H.kv_decode(oracle.reduce(oracle.convert(oracle.intcount_map(keys2)), "count"))
    assert mr.convert() == len(want)
    mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(want)


def test_sort_keys_leaves_sorted_pairs_alone_and_still_sorts_the_rest(MR, oracle):
    recs = [(struct.pack("<i", k // 3 - 20), struct.pack("<I", k)) for k in range(30_000)]   # ascending, with ties
    mr = MR()
    mr.kv_add(recs)
    assert mr.sort_keys(1) == len(recs)
    assert mr.scan_kv() == recs                                                              # stable: nothing moves
    assert mr.sort_keys(-1) == len(recs)
    assert H.kv_encode(mr.scan_kv()) == oracle.sort_keys(H.kv_encode(recs), -1)
    swapped = recs[:]
    swapped[-1], swapped[-2] = (struct.pack("<i", -99), b"abcd"), swapped[-1]                  # one pair out of place
    mr2 = MR()
    mr2.kv_add(swapped)
    mr2.sort_keys(1)
    assert H.kv_encode(mr2.scan_kv()) == oracle.sort_keys(H.kv_encode(swapped), 1)


def test_sort_keys_after_the_counting_convert_is_free_until_the_kv_changes(MR, oracle):
    """reduce() after the counting convert leaves keys ascending and says so; sort_keys(1) then returns at once -- but
    not after an append, an add, a descending sort, or for another compare flag"""
    keys = W.intcount_keys_np(0, 200_000, ndistinct=5000)

    def tags_of(mr, fn):
        mr.prof(True)
        fn()
        return {t["tag"] for t in mr.prof_report()[0]}

    mr = MR()
    mr.map_intcount(keys)
    mr.collate()
    nu = mr.reduce("count")
    want = oracle.sort_keys(oracle.reduce(oracle.convert(oracle.intcount_map(keys)), "count"), 1)
    assert tags_of(mr, lambda: mr.sort_keys(1)) == set()          # nothing launched
    assert H.kv_encode(mr.scan_kv()) == want
    assert "sort_extract" in tags_of(mr, lambda: mr.sort_keys(3))  # another compare type: the usual path
    mr.sort_keys(1)
    # append one small key: the order is gone and sort_keys must find that out
    mr.kv_add([(struct.pack("<I", 0), struct.pack("<i", 7))], addflag=1)
    assert "sort_extract" in tags_of(mr, lambda: mr.sort_keys(1))
    got = mr.scan_kv()
    assert [k for k, _ in got] == sorted((k for k, _ in got), key=lambda b: struct.unpack("<i", b)[0]) and len(got) == nu + 1
    # add() of another object, and a descending sort in between
    a, b = MR(), MR()
    for m in (a, b):
        m.map_intcount(keys)
        m.collate()
        m.reduce("count")
    a.add(b)
    a.sort_keys(1)
    ks = [struct.unpack("<i", k)[0] for k, _ in a.scan_kv()]
    assert ks == sorted(ks) and len(ks) == 2 * nu
    b.sort_keys(-1)
    b.sort_keys(1)
    assert H.kv_encode(b.scan_kv()) == want


# ------------------------------------------------------------------------------------------- compact stores

@pytest.mark.parametrize("compact", [0, 1])
@pytest.mark.parametrize("cb", ["count", "sum", "cull"])
def test_compact_stores_give_the_same_pairs_and_pages(MR, oracle, compact, cb):
    """header-less KV + factored KMV (compact=1) vs record streams everywhere (compact=0): same pages, same results"""
    n = 200_000
    keys = W.intcount_keys_np(0, n, ndistinct=3000, seed=5)
    kv_ref = oracle.intcount_map(keys)
    mr = MR(compact=compact, memsize=-(1 << 20))
    assert mr.map_intcount(keys[:120_000]) == 120_000
    assert mr.map_intcount(keys[120_000:], addflag=1) == n          # append keeps the form
    st = mr.kv_stats()
    assert (st.npair, st.keysize, st.valuesize, st.alignsize) == (n, 4 * n, 4 * n, 16 * n)
    assert mr.kv_pages() == H.pages_decode(oracle.kv_pages(kv_ref, 1 << 20))   # page copies write the records out
    assert mr.convert() == len(np.unique(keys))                     # counting path, straight from the rows
    kmv_ref = oracle.convert(kv_ref)
    if cb == "count":                                               # reading the KMV writes its record stream out
        st = mr.kmv_stats()
        assert (st.npair, st.nvalue, st.valuesize) == (len(np.unique(keys)), n, 4 * n)
        assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(H.kmv_decode(kmv_ref))
    mr.reduce(cb)                                                   # ... or the functor reads the factored form
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(oracle.reduce(kmv_ref, cb)))


@pytest.mark.parametrize("ahead", [0, 1])
@pytest.mark.parametrize("where", ["device", "pinned", "pageable"])
def test_counts_made_at_emit_time_follow_the_kv(MR, oracle, ahead, where):
    """map_intcount counts while it writes (ahead=1; host input: chunked DMA with the count overlapped); the counts
    must be used only for that very KV, unchanged"""
    torch = _torch()
    a = W.intcount_keys_np(0, 150_000, ndistinct=2000, seed=11)
    b = W.intcount_keys_np(0, 90_000, ndistinct=500, seed=12)
    want = lambda keys: H.canon_kv(H.kv_decode(oracle.reduce(oracle.convert(oracle.intcount_map(keys)), "count")))

    class _Loc:  # hands map_intcount the keys where the test wants them
        def __init__(self, mr):
            self.mr, self.keep = mr, []

        def map_intcount(self, keys, addflag=0):
            if where == "pageable":
                return self.mr.map_intcount(keys, addflag=addflag)
            t = torch.from_numpy(keys.view(np.int32).copy())
            t = t.cuda() if where == "device" else t.pin_memory()
            self.keep.append(t)
            return self.mr.map_intcount(t, addflag=addflag)

        def __getattr__(self, name):
            return getattr(self.mr, name)

    _MR = MR
    MR = lambda **kw: _Loc(_MR(**kw))
    mr = MR(ahead=ahead)
    mr.map_intcount(a)                                   # straight to convert: the counts are the ones made ahead
    mr.convert(), mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == want(a)
    mr.map_intcount(a)
    mr.map_intcount(b)                                   # a's counts are stale now
    mr.convert(), mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == want(b)
    mr.map_intcount(a)
    mr.map_intcount(b, addflag=1)                        # appended: recount
    mr.convert(), mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == want(np.concatenate([a, b]))
    bad = a.copy()
    bad[77_777] = 1 << 30                                # emit-time count gives up, the copy must still be complete
    mr.map_intcount(bad)
    assert mr.scan_kv() == H.kv_decode(oracle.intcount_map(bad))
    mr.convert(), mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == want(bad)
    mr.map_intcount(a)                                   # and the table is clean again afterwards
    mr.convert(), mr.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == want(a)
    mr2 = MR(ahead=ahead)                                # two handles do not share counts
    mr.map_intcount(a)
    mr2.map_intcount(b)
    mr2.convert(), mr.convert()
    mr.reduce("count"), mr2.reduce("count")
    assert H.canon_kv(mr.scan_kv()) == want(a) and H.canon_kv(mr2.scan_kv()) == want(b)


@pytest.mark.parametrize("P", [2, 8, 33])
def test_partition_of_headerless_rows(MR, oracle, P):
    keys = W.intcount_keys_np(0, 70_001, ndistinct=5000, seed=P)
    mr = MR()
    mr.map_intcount(keys)
    counts = mr.debug_partition(P)
    recs = H.kv_decode(oracle.intcount_map(keys))
    dest = oracle.partition(H.kv_encode(recs), P)
    assert counts == [dest.count(p) for p in range(P)]
    assert mr.scan_kv() == [rec for p in range(P) for rec, d in zip(recs, dest) if d == p]


def test_partition_staged_runs_variant():
    """the staged (shared-memory runs, 16-byte stores) form of pass 2 is what crosses NVLink; on one GPU it is
    chosen by MRGPU_PART_DIRECT=0, which is read once per process: rerun the placement tests in a child"""
    import os
    import subprocess
    import sys

    if os.environ.get("MRGPU_PART_DIRECT") == "0":
        pytest.skip("already inside the child run")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_parity.py"), "-q", "-x", "-m", "gpu", "-k",
                        "single_pass_partition or partition_of_headerless_rows"], env=dict(os.environ, MRGPU_PART_DIRECT="0"),
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "passed" in p.stdout and "failed" not in p.stdout


def test_headerless_kv_meets_other_shapes(MR, oracle):
    """ops without a header-less kernel (add, map(mr), sort_keys, appending another shape) see the records"""
    keys = W.intcount_keys_np(0, 3000, ndistinct=40, seed=9)
    recs = H.kv_decode(oracle.intcount_map(keys))
    mr = MR()
    mr.map_intcount(keys)
    extra = [(b"abc", b"z"), (b"", b"")]
    mr.kv_add(extra, addflag=1)
    assert mr.scan_kv() == recs + extra
    a, b = MR(), MR()
    a.map_intcount(keys)
    b.map_intcount(keys[:100])
    assert b.add(a) == 3100
    assert b.scan_kv() == recs[:100] + recs
    c = MR()
    c.map_mr(a, "invert")
    assert c.scan_kv() == [(v, k) for k, v in recs]
    assert a.sort_keys(1) == 3000
    assert H.kv_encode(a.scan_kv()) == oracle.sort_keys(oracle.intcount_map(keys), 1)


def test_headerless_edge_and_vertex_rows_meet_other_ops(MR, oracle):
    """rmat_generate, reduce(cull) and edge_to_vertices keep value-less pairs as header-less rows (store.cuh); every
    other op sees the reference's records: pages, scan, add, append of another shape, invert, sort, clone, copy"""
    st0 = oracle.lib.orc_srand48(7)
    _, edges = oracle.rmat_generate(st0, 3000, 9, .57, .19, .19, .05, 0.0)
    recs = [(e.tobytes(), b"") for e in edges]

    def gen(**settings):
        mr = MR(**settings)
        assert mr.map_rmat(3000, 9, .57, .19, .19, .05, 0.0, st0, addflag=0)[0] == 3000
        return mr

    a = gen()
    assert a.scan_kv() == recs
    (nk, page), = a.kv_pages()
    assert nk == 3000 and page == b"".join(struct.pack("<II", 16, 0) + k for k, _ in recs)
    st = a.kv_stats()
    assert (st.npair, st.keysize, st.valuesize) == (3000, 48000, 0)
    # append pairs of another shape: the rows become records first
    a.kv_add([(b"xy", b"z")], addflag=1)
    assert a.scan_kv() == recs + [(b"xy", b"z")]
    # add() in both directions, copy, invert
    b, c = gen(), MR()
    c.kv_add([(b"k" * 16, b"")])
    assert c.add(b) == 3001 and c.scan_kv() == [(b"k" * 16, b"")] + recs
    assert b.add(c) == 6001 and b.scan_kv() == recs + [(b"k" * 16, b"")] + recs
    d = gen().copy()
    assert d.scan_kv() == recs
    e = MR()
    e.map_mr(d, "invert")
    assert e.scan_kv() == [(v, k) for k, v in recs]
    # cull hands the KMV's key array over as rows; a second round appends rows to it; compact=0 does none of this
    out = {}
    for compact in (1, 0):
        m = gen(compact=compact)
        m.collate()
        nu = m.reduce("cull")
        assert H.canon_kv(m.scan_kv()) == H.canon_kv(sorted(set(recs)))
        m.map_rmat(500, 9, .57, .19, .19, .05, 0.0, st0 + 12345, addflag=1)
        _, more = oracle.rmat_generate(st0 + 12345, 500, 9, .57, .19, .19, .05, 0.0)
        assert sorted(m.scan_kv()) == sorted(list(set(recs)) + [(x.tobytes(), b"") for x in more])
        assert m.kv_stats().npair == nu + 500
        v = MR(compact=compact)
        assert v.map_mr(m, "edge_to_vertices") == 2 * (nu + 500)
        assert v.map_mr(m, "edge_to_vertex", addflag=1) == 3 * (nu + 500)
        out[compact] = (H.kv_encode(v.scan_kv()), [p for _, p in v.kv_pages()])
        v.sort_keys(2)
        ks = [int.from_bytes(k, "little") for k, _ in v.scan_kv()]
        assert ks == sorted(ks) and len(ks) == 3 * (nu + 500)
        m.clone()
        assert len(m.scan_kmv()) == nu + 500
    assert out[1] == out[0]


# ------------------------------------------------------------------------------------------- sort_keys

@pytest.mark.parametrize("flag", [1, 2, 3, 4, 5, 6])
def test_sort_values_flags_golden(MR, golden, flag):
    """sort_values(flag) against pages recorded from the reference (tests/golden section 11)"""
    for sgn, tag in ((1, "p"), (-1, "m")):
        mr = MR()
        _load(mr, golden[f"sortv{flag}.in"])
        n = mr.sort_values(sgn * flag)
        assert n == len(H.kv_decode(golden[f"sortv{flag}.in"]))
        assert H.kv_encode(mr.scan_kv()) == golden[f"sortv{flag}.out{tag}"]


@pytest.mark.parametrize("name", ["convert", "pages_b2"])
def test_clone_and_collapse_golden(MR, golden, oracle, name):
    """clone (src/mapreduce.cpp:631) and collapse (:681) against KMVs dumped by the reference (tests/golden section 13)"""
    kvb = golden["convert.in"] if name == "convert" else golden["pages_b2.in"]
    recs = H.kv_decode(kvb)
    mr = MR()
    _load(mr, kvb)
    assert mr.clone() == len(recs)
    assert mr.scan_kmv() == H.kmv_decode(golden[f"clone.{name}"])            # same pairs, same order
    assert mr.reduce("count") == len(recs)
    assert mr.scan_kv() == [(k, struct.pack("<i", 1)) for k, _ in recs]
    mr = MR()
    _load(mr, kvb)
    assert mr.collapse(b"all\0") == 1
    st = mr.kmv_stats()
    assert (st.npair, st.nvalue, st.keysize, st.valuesize) == (1, 2 * len(recs), 4, sum(len(k) + len(v) for k, v in recs))
    assert mr.scan_kmv() == H.kmv_decode(golden[f"collapse.{name}"])
    assert mr.reduce("count") == 1 and mr.scan_kv() == [(b"all\0", struct.pack("<i", 2 * len(recs)))]


@pytest.mark.parametrize("al", [4, 8])
def test_clone_and_collapse_page_images(MR, golden, al):
    for op in ("clone", "collapse"):
        mr = MR(keyalign=al, valuealign=al)
        _load(mr, golden["pages_b2.in"])
        mr.clone() if op == "clone" else mr.collapse(b"all\0")
        want = H.pages_decode(golden[f"{op}.pages_b2.page.a{al}"])
        got = [(nkey, page) for nkey, page, ext in mr.kmv_pages()]
        assert len(got) == len(want) == 1 and got[0][0] == want[0][0]
        kal = (al, al)
        assert H.kmv_page_records(got[0][1], got[0][0], *kal) == H.kmv_page_records(want[0][1], want[0][0], *kal)


def test_collapse_limits_and_empty(MR):
    import gpu_mapreduce_b200 as g

    mr = MR(memsize=-4096)
    mr.kv_add([(struct.pack("<I", i), b"payload!") for i in range(400)])     # 400 * (8 + 12) bytes of values > one 4 KiB page
    with pytest.raises(g.MRError, match="Collapse of multi-block KMV not yet supported"):
        mr.collapse(b"k")
    with pytest.raises(g.MRError, match="Cannot clone without KeyValue"):
        MR().clone()
    with pytest.raises(g.MRError, match="Cannot collapse without KeyValue"):
        MR().collapse(b"k")
    e = MR()
    e.kv_add([])
    assert e.clone() == 0 and e.scan_kmv() == []
    e2 = MR()
    e2.kv_add([])
    assert e2.collapse(b"kk") == 1 and e2.scan_kmv() == [(b"kk", [])]
    b = MR()
    b.kv_add([(b"a", b"b")])
    assert b.broadcast(0) == 1 and b.scan_kv() == [(b"a", b"b")]           # one rank: nothing moves
    assert b.scrunch(1, b"s\0") == 1 and b.scan_kmv() == [(b"s\0", [b"a", b"b"])]


def test_widened_ops_error_paths(MR):
    import gpu_mapreduce_b200 as g

    mr = MR()
    for call, msg in ((lambda: mr.sort_values(1), "Cannot sort_values without KeyValue"),
                      (lambda: mr.gather(1), "Cannot gather without KeyValue"),
                      (lambda: mr.compress("count"), "Cannot compress without KeyValue")):
        with pytest.raises(g.MRError, match=msg):
            call()
    mr.kv_add([(b"k", b"ab"), (b"j", b"a")])
    with pytest.raises(g.MRError, match="Invalid compare method for sort values"):
        mr.sort_values(7)
    with pytest.raises(g.MRError, match="value shorter than the compared type"):
        mr.sort_values(1)                                  # 2-byte values compared as int
    with pytest.raises(g.MRError, match="Invalid processor count for gather"):
        mr.gather(2)
    assert mr.gather(1) == 2 and mr.scan_kv() == [(b"k", b"ab"), (b"j", b"a")]
    assert mr.sort_values(6) == 2 and mr.scan_kv() == [(b"j", b"a"), (b"k", b"ab")]


@pytest.mark.parametrize("cb", ["count", "sum", "cull"])
def test_compress_golden(MR, golden, cb):
    """compress = local convert + reduce (src/mapreduce.cpp:749-851), against the reference's output"""
    mr = MR()
    _load(mr, golden["convert.in"])
    want = H.kv_decode(golden[f"compress.{cb}"])
    assert mr.compress(cb) == len(want)
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(want)
    with pytest.raises(Exception, match="Cannot compress without KeyValue"):
        mr2 = MR()
        mr2.compress(cb)
    big = MR()                                         # the counting path under compress
    keys = W.intcount_keys_np(0, 100_000, ndistinct=700, seed=3)
    big.map_intcount(keys)
    assert big.compress("count") == len(np.unique(keys))
    u, c = np.unique(keys, return_counts=True)
    assert sorted((struct.unpack("<I", k)[0], struct.unpack("<i", v)[0]) for k, v in big.scan_kv()) == list(zip(u.tolist(), c.tolist()))
    assert big.gather(1) == len(u)                     # one rank: nothing moves (src/mapreduce.cpp:905-910)


@pytest.mark.parametrize("flag", [1, 2, 3, 4, 5, 6])
def test_sort_keys_flags_golden(MR, golden, flag):
    for sgn, name in ((1, "outp"), (-1, "outm")):
        mr = MR()
        _load(mr, golden[f"sort{flag}.in"])
        mr.sort_keys(sgn * flag)
        assert H.kv_encode(mr.scan_kv()) == golden[f"sort{flag}.{name}"]


# ------------------------------------------------------------------------------------------- text functors

def test_wordfreq_golden(MR, golden):
    mr = MR()
    total = 0
    for i in range(3):
        total = mr.map_tokenize(golden[f"wordfreq.in{i}"], addflag=1 if i else 0)
    ref_map = H.kv_decode(golden["wordfreq.mapout"])
    assert total == len(ref_map)
    assert mr.scan_kv() == ref_map          # same pairs in the same order as the reference's fileread
    mr.collate()
    mr.reduce("sum")
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(golden["wordfreq.final"]))


@pytest.mark.parametrize("kalign", [4, 8])
def test_convert_by_hashing_variable_length_keys(MR, oracle, kalign):
    """value-less pairs with variable-length keys (words) are grouped by the hash path (ops_convert_hash.cuh): same KMV --
    pairs, order, statistics, pages -- as the fingerprint sort that compact=0 takes, and as the oracle after canonical sort"""
    r = np.random.RandomState(8)
    vocab = [bytes(r.randint(97, 123, size=r.randint(1, 20)).astype(np.uint8)) + b"\0" for _ in range(5000)]
    vocab += [b"", b"\0", b"ab", b"ab\0", b"ab\0\0"]                       # keys that differ only in length / trailing NULs
    idx = r.zipf(1.2, size=150_000) % len(vocab)
    recs = [(vocab[i], b"") for i in idx]
    want = H.kmv_decode(oracle.convert(H.kv_encode(recs)))
    out = {}
    for compact in (1, 0):
        mr = MR(compact=compact, keyalign=kalign)
        mr.kv_add(recs)
        assert mr.convert() == len(want)
        st = mr.kmv_stats().asdict()
        kmv = mr.scan_kmv()
        assert H.canon_kmv(kmv) == H.canon_kmv(want)
        out[compact] = (st, kmv, mr.kmv_pages())
    assert out[1] == out[0]
    for cb in ("count", "sum", "cull"):
        mr = MR(keyalign=kalign)
        mr.kv_add(recs)
        mr.convert()
        mr.reduce(cb)
        assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(oracle.reduce(oracle.convert(H.kv_encode(recs)), cb)))


def test_convert_by_hashing_gives_up_on_too_many_distinct_keys(MR):
    """more distinct keys than the L2-resident table holds: the hash pass raises its flag and the fingerprint sort runs"""
    n = 2_300_000
    text = b" ".join(b"%d" % i for i in range(n)) + b" 7 7 7"
    mr = MR()
    assert mr.map_tokenize(text) == n + 3
    assert mr.convert() == n
    st = mr.kmv_stats()
    assert st.nvalue == n + 3 and st.keysize == sum(len(b"%d" % i) + 1 for i in range(n))
    assert mr.reduce("count") == n
    mr.sort_keys(5)
    recs = mr.scan_kv()
    assert len(recs) == n and dict(recs)[b"7\0"] == struct.pack("<i", 4) and dict(recs)[b"8\0"] == struct.pack("<i", 1)


def test_tokenize_edge_cases(MR, oracle):
    for text in [b"", b" ", b"a", b" a", b"a ", b"\t\n\f\r", b"one\x00two three", b"\x00abc", b"x" * 5000, b"a b " * 3000 + b"z"]:
        mr = MR()
        n = mr.map_tokenize(text)
        want = H.kv_decode(oracle.tokenize(text))
        assert n == len(want)
        assert mr.scan_kv() == want


def test_tokenize_chunk_boundaries_vs_oracle(MR, oracle):
    """the 16-bytes-per-thread tokenizer: words that straddle 16-byte vectors, warps and 4 KB chunks, a word longer than
    several chunks, every delimiter and \\v (not a delimiter), text cut at a NUL, lengths around the vector size"""
    r = np.random.RandomState(42)
    texts = []
    for ln in (15, 16, 17, 31, 32, 33, 4095, 4096, 4097, 8191, 8193, 70_001):
        a = r.choice(np.frombuffer(b"ab \t\n\f\r\x0bz", dtype=np.uint8), size=ln, p=[.3, .2, .2, .04, .04, .04, .04, .04, .1])
        texts.append(bytes(a))
    texts.append(b"w" * 13_000 + b" " + b"q" * 9_000)                     # words spanning three chunks
    texts.append(b" " * 4096 + b"x" * 4096 + b" " * 4096 + b"y")          # a word that is exactly one chunk
    texts.append(b"ab cd" * 1000 + b"\x00" + b" tail words" * 100)         # strtok stops at the first NUL
    texts.append(bytes(r.randint(1, 256, size=50_000).astype(np.uint8)))   # arbitrary bytes, no NUL
    for al in (4, 8):
        for text in texts:
            mr = MR(keyalign=al, valuealign=al)
            want = H.kv_decode(oracle.tokenize(text))
            assert mr.map_tokenize(text) == len(want)
            assert mr.scan_kv() == want
            assert mr.map_tokenize(text, addflag=1) == 2 * len(want)       # append keeps the earlier records
            assert mr.scan_kv() == want + want


def test_tokenize_bytewise_variant():
    """the per-byte form of the tokenizer (used for text that is not 16-byte aligned) is chosen by
    MRGPU_TOK_BYTEWISE=1, read once per process: rerun the tokenizer tests in a child"""
    import os
    import subprocess
    import sys

    if os.environ.get("MRGPU_TOK_BYTEWISE") == "1":
        pytest.skip("already inside the child run")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_parity.py"), "-q", "-x", "-m", "gpu", "-k",
                        "wordfreq_golden or tokenize_edge or tokenize_chunk"], env=dict(os.environ, MRGPU_TOK_BYTEWISE="1"),
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "passed" in p.stdout and "failed" not in p.stdout


def test_invindex_golden(MR, golden):
    mr = MR()
    for i in range(4):
        mr.map_href(golden[f"invindex.in{i}"], f"part-{i:05d}", addflag=1 if i else 0)
    assert mr.scan_kv() == H.kv_decode(golden["invindex.mapout"])
    mr.aggregate()
    mr.convert()
    assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(H.kmv_decode(golden["invindex.kmv"]))


def _postings_line(values):
    """cpu/InvertedIndex.cpp:352-361: every value printed as a C string followed by a space"""
    return b"".join(v[:v.index(b"\0")] + b" " for v in values) + b"\0"


@pytest.mark.parametrize("compact", [1, 0])
def test_reduce_postings_golden(MR, golden, compact):
    """InvertedIndex end to end: href scanner -> aggregate -> convert -> reduce(postings); the postings string of every URL
    is the reference's output line (values of the reference's own KMV fixture, in ITS arrival order, printed by its rule)"""
    mr = MR(compact=compact)
    for i in range(4):
        mr.map_href(golden[f"invindex.in{i}"], f"part-{i:05d}", addflag=1 if i else 0)
    mr.aggregate()
    nu = mr.convert()
    want = {k: _postings_line(vals) for k, vals in H.kmv_decode(golden["invindex.kmv"])}
    assert mr.reduce("postings") == nu == len(want)
    got = mr.scan_kv()
    assert dict(got) == want and len(got) == len(want)
    st = mr.kv_stats()
    assert st.keysize == sum(len(k) for k in want) and st.valuesize == sum(len(v) for v in want.values())


def test_reduce_postings_mixed_lengths_and_page_limit(MR):
    import gpu_mapreduce_b200 as g

    r = np.random.RandomState(21)
    docs = [b"d%d" % i + b"x" * (i % 9) + b"\0" for i in range(40)]
    recs = [(b"url%03d" % r.randint(0, 60), docs[r.randint(0, 40)]) for _ in range(5000)]
    recs += [(b"hot", docs[i % 40]) for i in range(3000)]           # one key with thousands of postings (warp loop over > 32 values)
    mr = MR()
    mr.kv_add(recs)
    mr.convert()
    want = {}
    for k, v in recs:
        want.setdefault(k, []).append(v)
    mr.reduce("postings")
    assert dict(mr.scan_kv()) == {k: _postings_line(v) for k, v in want.items()}
    small = MR(memsize=-4096)                                        # 4 KiB pages: the hot key's line cannot be a pair
    small.kv_add(recs)
    small.convert()
    with pytest.raises(g.MRError, match="exceeds page size"):
        small.reduce("postings")


@pytest.mark.parametrize("compact", [1, 0])
def test_rmat_example_reduces_nonzero_degree_histo(MR, compact):
    """examples/rmat.cpp:266-295 (32-bit vertices): nonzero -> collate -> degree -> collate -> histo"""
    from collections import Counter

    r = np.random.RandomState(5)
    edges = {(int(a), int(b)) for a, b in r.zipf(1.5, size=(6000, 2)) % 200}
    mr = MR(compact=compact)
    mr.kv_add([(struct.pack("<ii", a, b), b"") for a, b in edges] * 2)       # every edge twice, as before the cull
    assert mr.collate() == len(edges)
    assert mr.reduce("nonzero") == len(edges)
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv([(struct.pack("<i", a), b"") for a, _ in edges])
    rows = Counter(a for a, _ in edges)
    assert mr.collate() == len(rows)
    assert mr.reduce("degree") == len(rows)
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv([(struct.pack("<i", c), b"") for c in rows.values()])
    hist = Counter(rows.values())
    assert mr.collate() == len(hist)
    assert mr.reduce("histo") == len(hist)
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv([(struct.pack("<i", c), struct.pack("<i", n)) for c, n in hist.items()])


def test_href_fsm_random_vs_oracle(MR, oracle):
    r = np.random.RandomState(77)
    toks = [b"<", b"a", b" ", b"h", b"r", b"e", b"f", b"=", b'"', b"x", b"href", b"<a href=\"", b"u", b">", b"\t"]
    for trial in range(30):
        text = b"".join(toks[r.randint(len(toks))] for _ in range(r.randint(1, 4000)))
        mr = MR()
        n = mr.map_href(text, "f.html")
        want = H.kv_decode(oracle.href(text, "f.html"))
        assert n == len(want)
        assert mr.scan_kv() == want


# ------------------------------------------------------------------------------------------- R-MAT + degree

def _gpu_rmat(mr, N, ef, a, b, c, d, fr, seed, P):
    states = [((seed + r) << 16) | 0x330E for r in range(P)]
    ntotal, nremain, niter = (1 << N) * ef, (1 << N) * ef, 0
    while nremain:
        niter += 1
        for r in range(P):
            ng = nremain // P + (1 if r < nremain % P else 0)
            _, states[r] = mr.map_rmat(ng, N, a, b, c, d, fr, states[r], addflag=1)
        nunique = mr.collate()
        mr.reduce("cull")
        nremain = ntotal - nunique
    return niter


def test_rmat_degree_golden(MR, golden):
    for ci, case in enumerate(golden.manifest["rmat"]):
        N, ef, a, b, c, d, fr, seed, P, dup = case["args"]
        mr = MR()
        assert _gpu_rmat(mr, N, ef, a, b, c, d, fr, seed, P) == case["iterations"]
        assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(golden[f"rmat{ci}.edges"]))
        mrv = MR()
        mrv.map_mr(mr, "edge_to_vertex" if dup == 1 else "edge_to_vertices")
        mrv.collate()
        mrv.reduce("count")
        assert H.canon_kv(mrv.scan_kv()) == H.canon_kv(H.kv_decode(golden[f"rmat{ci}.degree"]))


def _edges(mr):
    return np.sort(np.frombuffer(b"".join(k for k, _ in mr.scan_kv()), dtype=np.dtype([("i", "<u8"), ("j", "<u8")])), order=("j", "i"))


def test_rmat_later_rounds_merge_instead_of_sorting(MR):
    """From the second round on, collate's input is the previous round's edges (in key order) plus the new ones: only
    the new ones are sorted and merged in.  Same edges and rounds as with the merge switched off."""
    runs = {}
    for presorted in (1, 0):
        mr = MR()
        mr.presorted = presorted
        mr.prof(True)
        niter = _gpu_rmat(mr, 17, 8, .57, .19, .19, .05, 0.1, 77, 1)
        tags = {t["tag"] for t in mr.prof_report()[0]}
        assert ("merge_sorted" in tags) == bool(presorted)
        runs[presorted] = (niter, _edges(mr))
    assert runs[1][0] == runs[0][0] > 2
    assert len(runs[1][1]) == (1 << 17) * 8 and np.array_equal(runs[1][1], runs[0][1])
    assert len(np.unique(runs[1][1])) == len(runs[1][1])


def test_convert_on_guessed_key_bytes_falls_back_when_more_bytes_vary(MR, oracle):
    """the second convert() of 16-byte keys on one handle extracts its sort keys on the byte set of the first while
    checking it; keys that vary in other bytes as well must take the scan and come out right, and so must keys that
    vary in fewer"""
    mr = MR()
    mr.prof(True)
    for nl, n in ((7, 100_000), (20, 150_000), (20, 120_000), (5, 90_000)):
        st0 = oracle.lib.orc_srand48(nl + n)
        _, edges = oracle.rmat_generate(st0, n, nl, .57, .19, .19, .05, 0.0)
        assert mr.map_rmat(n, nl, .57, .19, .19, .05, 0.0, st0, addflag=0)[0] == n
        want = np.unique(edges, axis=0)
        assert mr.convert() == len(want)
        assert mr.reduce("cull") == len(want)
        got = np.frombuffer(b"".join(k for k, _ in mr.scan_kv()), dtype=np.uint64).reshape(-1, 2)
        assert np.array_equal(np.unique(got, axis=0), want) and len(got) == len(want)
    calls = {t["tag"]: t["calls"] for t in mr.prof_report()[0]}
    assert calls["convert_key_diff"] == 2  # the first call and the one whose keys outgrew the guess


def test_presorted_run_in_the_middle_and_all_sorted(MR, oracle):
    """the long run may sit anywhere (after an exchange it follows what lower ranks sent); an input with no descent at
    all is not sorted again; 4- and 8-byte keys"""
    r = np.random.RandomState(5)
    for dt, hi in ((np.uint32, 1 << 30), (np.uint64, 1 << 40)):
        bulk = np.unique(r.randint(0, hi, size=400_000).astype(dt))
        head, tail = r.randint(0, hi, size=3000).astype(dt), r.randint(0, hi, size=5000).astype(dt)
        for parts in ((head, bulk, tail), (bulk,), (bulk, tail), (head, bulk)):
            keys = np.concatenate(parts)
            mr = MR()
            mr.prof(True)
            import gpu_mapreduce_b200 as g
            L = g.lib()
            mr._ck(L.mrgpu_kv_open(mr._h, 0))
            mr._ck(L.mrgpu_kv_add(mr._h, len(keys), keys.ctypes.data, None, keys.itemsize, None, None, 0))
            cnt = ctypes.c_uint64()
            mr._ck(L.mrgpu_kv_close(mr._h, ctypes.byref(cnt)))
            u, c = np.unique(keys, return_counts=True)
            assert mr.convert() == len(u)
            tags = {t["tag"] for t in mr.prof_report()[0]}
            assert "presorted_scan" in tags and ("merge_sorted" in tags) == (len(parts) > 1)
            assert "onesweep_pass" not in tags or len(parts) > 1
            mr.reduce("count")
            got = dict((int.from_bytes(k, "little"), int.from_bytes(v, "little")) for k, v in mr.scan_kv())
            assert got == dict(zip(u.tolist(), c.tolist()))


def test_rmat_generate_stream_matches_oracle(MR, oracle):
    """one rmat_generate call = the reference's sequential drand48 stream, edge for edge, state included"""
    for fr in (0.0, 0.2):
        st0 = oracle.lib.orc_srand48(4242)
        st1, edges = oracle.rmat_generate(st0, 5000, 14, .57, .19, .19, .05, fr)
        mr = MR()
        n, st = mr.map_rmat(5000, 14, .57, .19, .19, .05, fr, st0, addflag=0)
        assert n == 5000 and st == st1
        got = np.frombuffer(b"".join(k for k, _ in mr.scan_kv()), dtype=np.uint64).reshape(-1, 2)
        assert np.array_equal(got, edges)
    # more edges than one grid of threads: a thread's later edges are reached by the precomputed grid-wide jump;
    # fraction == 0 decides the quadrants on the integer state (thresholds ceil(t * 2^48)), incl. degenerate a, b, c, d
    for n, nl, (a, b, c, d), fr in ((700_001, 20, (.57, .19, .19, .05), 0.0), (650_000, 9, (.25, .25, .25, .25), 0.0),
                                    (610_000, 31, (0.0, 1.0, 0.0, 0.0), 0.0), (620_000, 12, (1e-9, .5, .5 - 1e-9, 0.0), 0.0),
                                    (640_000, 16, (.45, .15, .15, .25), 0.05)):
        st0 = oracle.lib.orc_srand48(99 + nl)
        st1, edges = oracle.rmat_generate(st0, n, nl, a, b, c, d, fr)
        mr = MR()
        cnt, st = mr.map_rmat(n, nl, a, b, c, d, fr, st0, addflag=0)
        assert cnt == n and st == st1
        pages = [np.frombuffer(pg, dtype=np.uint64).reshape(nk, 3)[:, 1:] for nk, pg in mr.kv_pages()]
        assert np.array_equal(np.concatenate(pages), edges)


def test_map_mr_self_and_invert(MR, oracle):
    r = np.random.RandomState(3)
    recs = [(struct.pack("<QQ", r.randint(0, 50), r.randint(0, 50)), b"") for _ in range(3000)]
    mr = MR()
    mr.kv_add(recs)
    assert mr.map_mr(mr, "edge_to_vertices") == 6000      # src == dst, src/mapreduce.cpp:1584-1602
    assert mr.scan_kv() == H.kv_decode(oracle.edge_to_vertices(H.kv_encode(recs), True))
    recs2 = [(bytes(r.randint(0, 256, size=r.randint(1, 9)).astype(np.uint8)), bytes(r.randint(0, 256, size=r.randint(0, 5)).astype(np.uint8)))
             for _ in range(500)]
    m2, m3 = MR(), MR()
    m2.kv_add(recs2)
    m3.map_mr(m2, "invert")
    assert m3.scan_kv() == [(v, k) for k, v in recs2]


def test_copy_global_count_and_abandon(MR, oracle):
    """mrgpu_copy (MapReduce::copy, src/mapreduce.cpp:290): deep copies of a KV (incl. the header-less form) and of a KMV
    (incl. the factored form) that outlive and do not follow the original; mrgpu_global_count; mrgpu_abandon"""
    import ctypes as C

    import gpu_mapreduce_b200 as g

    r = np.random.RandomState(41)
    keys = r.randint(0, 5000, size=100_000).astype(np.uint32)
    a = MR()
    a.map_intcount(keys)                                  # header-less KV
    b = a.copy()
    want_kv = [(struct.pack("<I", int(k)), struct.pack("<i", 1)) for k in keys]
    assert b.scan_kv() == want_kv
    n = C.c_uint64()
    assert g.lib().mrgpu_global_count(b._h, 0, C.byref(n)) == 0 and n.value == len(keys)
    assert g.lib().mrgpu_global_count(b._h, 1, C.byref(n)) == g.lib().mrgpu_global_count(b._h, 1, C.byref(n)) != 0   # no KMV: an error, not a crash
    nu = a.convert()                                      # factored KMV
    c = a.copy()
    a.reduce("count")                                     # the original moves on; the copy still holds the KMV
    assert g.lib().mrgpu_global_count(c._h, 1, C.byref(n)) == 0 and n.value == nu
    want = H.kmv_decode(oracle.convert(H.kv_encode(want_kv)))
    assert H.canon_kmv(c.scan_kmv()) == H.canon_kmv(want)
    assert c.reduce("count") == nu and H.canon_kv(c.scan_kv()) == H.canon_kv(a.scan_kv())
    assert b.scan_kv() == want_kv                         # untouched by everything above
    for h in (b, c):
        assert g.lib().mrgpu_abandon(h._h) == 0           # non-collective teardown (what the Python finaliser uses)
        object.__setattr__(h, "_h", None)


def test_add_appends_other_mr(MR):
    a, b = MR(), MR()
    ra = [(b"k%d" % i, b"v") for i in range(100)]
    rb = [(struct.pack("<I", i), struct.pack("<I", 1)) for i in range(50)]
    a.kv_add(ra)
    b.kv_add(rb)
    assert a.add(b) == 150
    assert a.scan_kv() == ra + rb


def test_add_and_map_mr_across_streams():
    """two handles on distinct non-blocking streams: add() / map(mr) write the other handle's header-less KV out as
    records on ITS stream and read them on mine -- the result must not depend on which stream runs first"""
    import gpu_mapreduce_b200 as g

    torch = _torch()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    n = 1 << 22
    keys = torch.randint(0, 1 << 20, (n,), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    a = g.MapReduce(device=0, stream=s1.cuda_stream)
    b = g.MapReduce(device=0, stream=s2.cuda_stream)
    c = g.MapReduce(device=0, stream=s1.cuda_stream)
    hk = keys.cpu().numpy()

    def records(mr):
        return np.concatenate([np.frombuffer(page, dtype=np.int32) for _, page in mr.kv_pages()]).reshape(-1, 4)

    try:
        for rep in range(3):
            assert b.map_intcount(keys) == n           # header-less (packed) KV on stream 2
            a.kv_add([(struct.pack("<I", 7), struct.pack("<i", 1))])
            assert a.add(b) == n + 1                   # unpack on stream 2, copy on stream 1
            rec = records(a)
            assert rec.shape[0] == n + 1 and rec[0, 2] == 7
            assert np.array_equal(rec[1:, 2], hk) and np.all(rec[:, 3] == 1) and np.all(rec[:, :2] == 4)
            b.map_intcount(keys)
            assert c.map_mr(b, "invert") == n           # map(mr) from a header-less source on the other stream
            rec = records(c)
            assert np.array_equal(rec[:, 3], hk) and np.all(rec[:, 2] == 1)
    finally:
        for h in (a, b, c):
            h.destroy()


# ------------------------------------------------------------------------------------------- out-of-core tier

def test_out_of_core_intcount_pipeline(MR, oracle):
    """hbm_budget = 1 MB: the KV moves to host memory piece by piece while map appends; stats / pages / scan read the
    pieces; convert splits them into sets by key hash; reduce walks the KMV pieces: same results as in core"""
    r = np.random.RandomState(31)
    parts = [(r.zipf(1.3, size=200_000) % 50_000).astype(np.uint32) for _ in range(5)]
    allkeys = np.concatenate(parts)
    want_kv = [(struct.pack("<I", int(k)), struct.pack("<i", 1)) for k in allkeys]
    res = {}
    for budget in (1, 0):
        mr = MR(hbm_budget=budget)
        n = 0
        for i, p in enumerate(parts):
            n = mr.map_intcount(p, addflag=1 if i else 0)
        assert n == len(allkeys)
        st = mr.kv_stats()
        assert st.npair == len(allkeys) and st.keysize == 4 * len(allkeys) and st.alignsize == 16 * len(allkeys)
        assert mr.scan_kv() == want_kv                              # pieces in order, then the part in HBM
        assert mr.aggregate() == len(allkeys)                       # one rank: nothing moves, nothing comes back
        nu = mr.convert()
        ks = mr.kmv_stats()
        assert ks.npair == nu and ks.nvalue == len(allkeys) and ks.exactsize == 12 * nu + 4 * len(allkeys) + 4 * nu + 4 * len(allkeys)
        kmv = mr.scan_kmv()
        assert mr.reduce("count") == nu
        res[budget] = (nu, H.canon_kmv(kmv), H.canon_kv(mr.scan_kv()))
    assert res[1] == res[0]
    want = H.kv_decode(oracle.reduce(oracle.convert(oracle.intcount_map(allkeys)), "count"))
    assert res[1][2] == H.canon_kv(want)


def test_out_of_core_keeps_value_arrival_order(MR, oracle):
    """variable-length keys and values added in batches that spill: every key's values stay in arrival order across
    pieces and sets (the reference's KMV value order), cull takes the first one"""
    r = np.random.RandomState(32)
    vocab = [bytes(r.randint(97, 123, size=r.randint(1, 12)).astype(np.uint8)) for _ in range(700)]
    batches = [[(vocab[r.randint(len(vocab))], b"v%d" % (1000 * b + i)) for i in range(20_000)] for b in range(4)]
    flat = [rec for b in batches for rec in b]
    mr = MR(memsize=-(256 << 10), outofcore=1)                      # the reference's switch: one 256 KiB page stays in memory
    for i, b in enumerate(batches):
        mr.kv_add(b, addflag=1 if i else 0)
    assert mr.kv_stats().npair == len(flat) and mr.kv_stats().npage > 4
    want = H.kmv_decode(oracle.convert(H.kv_encode(flat)))
    assert mr.convert() == len(want)
    assert H.canon_kmv(mr.scan_kmv()) == H.canon_kmv(want)          # value order inside a key is compared as is
    mr.reduce("cull")
    assert H.canon_kv(mr.scan_kv()) == H.canon_kv(H.kv_decode(oracle.reduce(oracle.convert(H.kv_encode(flat)), "cull")))


def test_out_of_core_store_comes_back_for_other_ops(MR, oracle):
    r = np.random.RandomState(33)
    parts = [r.randint(0, 1 << 20, size=150_000).astype(np.uint32) for _ in range(3)]
    mr, other = MR(hbm_budget=1), MR()
    for i, p in enumerate(parts):
        mr.map_intcount(p, addflag=1 if i else 0)
    allkeys = np.concatenate(parts)
    assert mr.sort_keys(1) == len(allkeys)                          # needs the whole KV in HBM: it is brought back
    got = np.array([struct.unpack("<I", k)[0] for k, _ in mr.scan_kv()], dtype=np.uint32)
    assert np.array_equal(got, np.sort(allkeys, kind="stable"))
    other.kv_add([(b"x", b"y")])
    assert other.add(mr) == len(allkeys) + 1


# ------------------------------------------------------------------------------------------- scale properties

def test_intcount_16M_properties(MR):
    """2^24 pairs (beyond what the oracle does in seconds): size-independent properties + numpy bincount"""
    torch = _torch()
    n, nd = 1 << 24, 1 << 20
    keys = W.intcount_keys_torch(0, n, "cuda", ndistinct=nd)
    mr = MR()
    assert mr.map_intcount(keys) == n
    nu = mr.convert()
    st = mr.kmv_stats()
    assert st.nvalue == n and st.exactsize == 12 * nu + 4 * n + 4 * nu + 4 * n
    assert mr.reduce("count") == nu
    assert mr.sort_keys(1) == nu
    (nkey, page), = mr.kv_pages()
    rec = np.frombuffer(page, dtype=np.int32).reshape(-1, 4)
    assert nkey == nu and np.all(rec[:, 0] == 4) and np.all(rec[:, 1] == 4)
    assert np.all(np.diff(rec[:, 2]) > 0)                       # sorted, unique
    assert int(rec[:, 3].astype(np.int64).sum()) == n           # checksum of counts
    bc = np.bincount(keys.cpu().numpy().astype(np.int64), minlength=nd)
    nz = np.nonzero(bc)[0]
    assert np.array_equal(rec[:, 2], nz) and np.array_equal(rec[:, 3], bc[nz])

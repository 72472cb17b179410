"""CPU: randomised cross-check of oracle/mr_oracle.c against the live reference binary oracle/_ref/refdrv.
Skipped where the reference has not been built (it is built in the container by `make -C oracle ref`
and travels to the GPU box as a prebuilt file)."""
import struct

import numpy as np
import pytest

import mrhelpers as H

pytestmark = pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref/refdrv not built")


def _random_kv(seed, n, nkeys, maxk=18, maxv=7):
    r = np.random.RandomState(seed)
    keys = [bytes(r.randint(0, 256, size=r.randint(1, maxk)).astype(np.uint8)) for _ in range(nkeys)]
    return [(keys[r.randint(nkeys)], bytes(r.randint(0, 256, size=r.randint(0, maxv)).astype(np.uint8)))
            for _ in range(n)]


@pytest.mark.parametrize("seed", [11, 12, 13])
def test_hash_random(oracle, seed):
    r = np.random.RandomState(seed)
    pairs = [(int(r.randint(0, 2 ** 31)), bytes(r.randint(0, 256, size=r.randint(0, 70)).astype(np.uint8)))
             for _ in range(300)]
    assert [oracle.hashlittle(k, iv) for iv, k in pairs] == H.ref_hash(pairs)


@pytest.mark.parametrize("seed,n,nkeys", [(21, 1, 1), (22, 4000, 37), (23, 6000, 5000)])
def test_convert_reduce_random(oracle, seed, n, nkeys):
    kv = H.kv_encode(_random_kv(seed, n, nkeys))
    assert oracle.convert(kv) == H.ref_ops(kv, ["convert", "dumpkmv"])
    for cb in ("count", "cull"):
        assert oracle.reduce(oracle.convert(kv), cb) == H.ref_ops(kv, ["collate", f"reduce_{cb}", "dumpkv"])


def test_convert_out_of_core_same_set(oracle):
    # 32 KiB pages force the reference through spool partitions / sets: key order changes, the set does not
    kv = H.kv_encode(_random_kv(31, 9000, 2500))
    ref = H.kmv_decode(H.ref_ops(kv, ["convert", "dumpkmv"], memsize=-32768))
    got = H.kmv_decode(oracle.convert(kv))
    assert H.canon_kmv(got) == H.canon_kmv(ref)


def test_empty_kv(oracle):
    kv = H.kv_encode([])
    assert H.kmv_decode(H.ref_ops(kv, ["convert", "dumpkmv"])) == []
    assert H.kmv_decode(oracle.convert(kv)) == []


@pytest.mark.parametrize("align", [4, 8, 16])
def test_pages_alignment(oracle, align):
    recs = _random_kv(41, 300, 40)
    kv = H.kv_encode(recs)
    ref = H.pages_decode(H.ref_ops(kv, ["pagekv"], align=align))
    got = H.pages_decode(oracle.kv_pages(kv, 64 << 20, align, align))
    assert [(n, len(b)) for n, b in got] == [(n, len(b)) for n, b in ref]
    assert H.kv_page_records(ref[0][1], ref[0][0], align, align) == recs
    refm = H.pages_decode(H.ref_ops(kv, ["convert", "pagekmv"], align=align))
    gotm = H.pages_decode(oracle.kmv_pages(oracle.convert(kv), 64 << 20, align, align))
    assert [(n, len(b)) for n, b in gotm] == [(n, len(b)) for n, b in refm]
    assert H.kmv_page_records(gotm[0][1], gotm[0][0], align, align) == \
        H.kmv_page_records(refm[0][1], refm[0][0], align, align)


def test_rmat_more_params(oracle):
    import os
    import tempfile

    for args in [(9, 2, .57, .19, .19, .05, 0.0, 4242, 4, 2), (5, 8, .3, .3, .2, .2, 0.3, 1, 1, 1)]:
        N, ef, a, b, c, d, fr, seed, P, dup = args
        with tempfile.TemporaryDirectory() as td:
            fe, fd = os.path.join(td, "e"), os.path.join(td, "d")
            t, _ = H.ref_run(["rmat", N, ef, a, b, c, d, fr, seed, P, 64, fe, fd, dup])
            ref_e, ref_d = open(fe, "rb").read(), open(fd, "rb").read()
        kv, niter = oracle.rmat(N, ef, a, b, c, d, fr, seed, P)
        assert niter == [x for x in t if x.get("op") == "rmat_done"][0]["iterations"]
        assert H.canon_kv(H.kv_decode(kv)) == H.canon_kv(H.kv_decode(ref_e))
        verts = oracle.edge_to_vertices(ref_e, both=(dup != 1))
        assert oracle.reduce(oracle.convert(verts), "count") == ref_d


@pytest.mark.parametrize("seed,n,nkeys", [(51, 1, 1), (52, 3000, 40), (53, 2500, 2500)])
def test_clone_collapse_compress_random(oracle, seed, n, nkeys):
    """the ops widened in round 1: the Python restatements against the live reference on random pairs"""
    kv = H.kv_encode(_random_kv(seed, n, nkeys))
    assert oracle.clone(kv) == H.ref_ops(kv, ["clone", "dumpkmv"])
    assert oracle.collapse(kv, b"all\0") == H.ref_ops(kv, ["collapse:all", "dumpkmv"])
    for cb in ("count", "cull"):
        assert oracle.compress(kv, cb) == H.ref_ops(kv, [f"compress_{cb}", "dumpkv"])


@pytest.mark.parametrize("flag", [1, -1, 2, -2, 3, -3, 4, -4, 5, -5, 6, -6])
def test_sort_values_random(oracle, flag):
    r = np.random.RandomState(60 + abs(flag))
    a = abs(flag)
    n = 900
    if a == 1:
        vals = [struct.pack("<i", int(x)) for x in r.randint(-40, 40, size=n)]
    elif a == 2:
        vals = [struct.pack("<Q", int(x)) for x in r.randint(0, 30, size=n)]
    elif a == 3:
        vals = [struct.pack("<f", float(x) * 0.5) for x in r.randint(-20, 20, size=n)]
    elif a == 4:
        vals = [struct.pack("<d", float(x) * 0.25) for x in r.randint(-20, 20, size=n)]
    else:
        vals = [bytes(r.randint(97, 100, size=r.randint(0, 5)).astype(np.uint8)) + b"\0" for _ in range(n)]
    kv = H.kv_encode([(struct.pack("<I", i), v) for i, v in enumerate(vals)])   # keys record the input order: ties must keep it
    assert oracle.sort_values(kv, flag) == H.ref_ops(kv, [f"sort_values:{flag}", "dumpkv"])

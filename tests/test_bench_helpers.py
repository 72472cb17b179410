"""CPU: the pieces of bench.py that the parity block rests on are themselves checked -- its independent lookup3
restatement (used for the placement check `hashlittle(key, 4, P) % P == rank` at every N) against the known answers
recorded from the reference's src/hash.cpp, and the vectorised form against the scalar one."""
import importlib.util
import os
import struct

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_bench_hash_restatement_matches_the_reference_kats(golden):
    b = _bench()
    for iv, hexkey, h in golden.manifest["hash"]:
        assert b.hashlittle_py(bytes.fromhex(hexkey), iv) == h, (iv, hexkey)


def test_bench_vectorised_hash_matches_the_scalar_one():
    b = _bench()
    r = np.random.RandomState(11)
    keys = np.concatenate([r.randint(0, 1 << 32, size=5000, dtype=np.uint64).astype(np.uint32),
                           np.array([0, 1, 2, (1 << 24) - 1, (1 << 26) - 1, 0xFFFFFFFF], dtype=np.uint32)])
    for iv in (0, 1, 2, 4, 8, 33):
        got = b.hashlittle_u32_np(keys, iv)
        want = np.array([b.hashlittle_py(struct.pack("<I", int(k)), iv) for k in keys], dtype=np.uint64)
        assert np.array_equal(np.asarray(got, dtype=np.uint64), want), iv

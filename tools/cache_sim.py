"""CPU simulation (numpy) of the count kernel's per-block cache on the IntCount key stream: which fraction of the
pairs ends as a RED to the table in L2 under a given cache size / hash / eviction rule.  This is where the figures
quoted in DESIGN.md 4 and in ops_convert_dense.cuh come from (first-come 25.6 %, evict-below-2 22.7 %, best key per
slot ~18 %).     python tools/cache_sim.py

One block of the real kernel sees 2^30 / 148 keys of the stream in a strided order; statistically that is a
contiguous sample of the same length, which is what is simulated."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_mapreduce_b200 import workloads as w  # noqa: E402

PER_BLOCK = (1 << 30) // 148
MULT = 0x278DC01  # DENSE_MULT
keys = w.intcount_keys_np(0, PER_BLOCK).astype(np.uint64)
k = keys.astype(np.int64)


def slots(m, mult=MULT):
    h = (keys * mult) & ((1 << 26) - 1)
    return ((h * m) >> 26).astype(np.int64)


def run(m, keep, rnd, mult=MULT):
    """REDs per pair: misses + non-empty entries flushed at the end of every round of `rnd` keys; an entry with fewer
    than `keep` hits in a round is evicted at the flush (keep = 0: first come, never evicted)."""
    slot = slots(m, mult)
    owner = np.full(m, -1, dtype=np.int64)
    reds = 0
    for r0 in range(0, len(k), rnd):
        ks, ss = k[r0:r0 + rnd], slot[r0:r0 + rnd]
        idx = np.nonzero(owner[ss] < 0)[0]
        if len(idx):
            _, fi = np.unique(ss[idx], return_index=True)
            owner[ss[idx[fi]]] = ks[idx[fi]]
        hit = owner[ss] == ks
        cnt = np.bincount(ss[hit], minlength=m)
        reds += int((~hit).sum()) + int((cnt > 0).sum())
        owner[cnt < keep] = -1
    return reds / len(k)


if __name__ == "__main__":
    u, c = np.unique(keys, return_counts=True)
    order = np.argsort(-c)
    print(f"block stream: {len(keys)} keys, {len(u)} distinct")
    for m in (32768, 57344):
        top = np.sort(c)[::-1][:m].sum() / len(keys)
        s2 = slots(m)[np.unique(keys, return_index=True)[1]]          # slot of every distinct key (u order)
        best = np.zeros(m, dtype=np.int64)
        np.maximum.at(best, s2, c)
        print(f"{m} entries: ideal (top-m keys) hits {top:.4f}; best key per slot hits {best.sum() / len(keys):.4f}")
    rnd = 63 * 16384  # keys per block per round in dense_count_kernel<0/3>
    for keep in (0, 2, 3, 4, 8):
        print(f"57344 entries, round {rnd}, evict below {keep}: REDs per pair {run(57344, keep, rnd):.4f}")
    for r in (32 * 16384, 16 * 16384):
        print(f"57344 entries, round {r}, evict below 2: REDs per pair {run(57344, 2, r):.4f}")

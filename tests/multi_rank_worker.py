"""Worker for the multi-rank tests.  Launched by torch.distributed.run with >= 2 ranks.

  backend nccl (default): one GPU per rank, the real aggregate()/collate() through the C-ABI + NCCL.
  MR_WORKER_BACKEND=gloo : CPU only -- exercises the host-side exchange plan (mrgpu_plan_exchange) with the
                           payload moved by gloo, against the reference's delivery rule.
Prints MULTI_OK on rank 0 when every check passed."""
import ctypes as C
import os
import struct
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import mrhelpers as H  # noqa: E402
import gpu_mapreduce_b200 as g  # noqa: E402
from gpu_mapreduce_b200 import workloads as W  # noqa: E402

BACKEND = os.environ.get("MR_WORKER_BACKEND", "nccl")
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", "0"))


def shard_records(case, r):
    """records rank r starts with"""
    rs = np.random.RandomState(1000 * case + r)
    if case == 0:   # uniform 4/4 (IntCount-like), skewed keys
        keys = W.intcount_keys_np(r * 20000, (r + 1) * 20000, ndistinct=3000)
        return [(struct.pack("<I", int(k)), struct.pack("<i", 1)) for k in keys]
    if case == 1:   # variable-length keys and values
        vocab = [bytes(rs.randint(97, 123, size=rs.randint(1, 14)).astype(np.uint8)) + b"\0" for _ in range(400)]
        return [(vocab[rs.randint(len(vocab))], bytes(rs.randint(0, 256, size=rs.randint(0, 6)).astype(np.uint8))) for _ in range(7000)]
    if case == 2:   # mixed: even ranks uniform, odd ranks variable
        if r % 2 == 0:
            return [(struct.pack("<Q", int(x)), b"") for x in rs.randint(0, 500, size=5000)]
        return [(bytes(rs.randint(0, 256, size=rs.randint(1, 12)).astype(np.uint8)), b"v") for _ in range(3000)]
    if case == 3:   # one rank empty
        return [] if r == 0 else [(struct.pack("<QQ", int(a), int(b)), b"") for a, b in rs.randint(0, 40, size=(4000, 2))]
    raise ValueError(case)


def expected_after_aggregate(case, me, orc):
    """reference delivery rule: pairs with hashlittle(key,kb,P) % P == me, by source rank, original order"""
    out = []
    for q in range(world):
        recs = shard_records(case, q)
        dest = orc.partition(H.kv_encode(recs), world)
        out += [rec for rec, d in zip(recs, dest) if d == me]
    return out


def main_nccl():
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    t = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        t.copy_(torch.frombuffer(bytearray(g.MapReduce.unique_id()), dtype=torch.uint8))
    dist.broadcast(t, 0)
    uid = bytes(t.cpu().numpy().tobytes())
    orc = H.Oracle()
    for case in range(4):
        mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
        mine = shard_records(case, rank)
        total = sum(len(shard_records(case, q)) for q in range(world))
        assert mr.kv_add(mine) == total            # every op returns the GLOBAL count
        assert mr.aggregate() == total
        assert mr.scan_kv() == expected_after_aggregate(case, rank, orc), f"case {case}: wrong pairs/order on rank {rank}"
        nuniq = mr.convert()
        mr.reduce("count")
        got = mr.scan_kv()
        gathered = [None] * world
        dist.all_gather_object(gathered, got)
        if rank == 0:
            allrecs = [rec for q in range(world) for rec in shard_records(case, q)]
            want = H.kv_decode(orc.reduce(orc.convert(H.kv_encode(allrecs)), "count"))
            flat = [rec for part in gathered for rec in part]
            assert len(flat) == nuniq == len(want)
            assert H.canon_kv(flat) == H.canon_kv(want), f"case {case}: global counts differ"
            for q, part in enumerate(gathered):   # keys live on the rank the reference would put them on
                for k, _ in part:
                    assert orc.hashlittle(k, world) % world == q
        mr.destroy()

    # aggregate(myhash), src/mapreduce.cpp:466-470: destination = myhash(key, keybytes) % nprocs, evaluated on the host
    def myhash(key, keybytes):
        return (key[0] if keybytes else 0) * 31 + keybytes

    for case in range(4):
        mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
        total = sum(len(shard_records(case, q)) for q in range(world))
        mr.kv_add(shard_records(case, rank))
        assert mr.aggregate(myhash) == total
        want = [rec for q in range(world) for rec in shard_records(case, q) if myhash(rec[0], len(rec[0])) % world == rank]
        assert mr.scan_kv() == want, f"user hash, case {case}: wrong pairs/order on rank {rank}"
        mr.destroy()

    # gather(numprocs), src/mapreduce.cpp:893-1036: rank r < numprocs ends with its own pairs followed by those of
    # ranks r+numprocs, r+2*numprocs, ... ; the others end empty
    for case in range(4):
        for numprocs in sorted({1, max(1, world - 1), world}):
            mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
            total = sum(len(shard_records(case, q)) for q in range(world))
            assert mr.kv_add(shard_records(case, rank)) == total
            assert mr.gather(numprocs) == total
            want = [rec for q in range(rank, world, numprocs) for rec in shard_records(case, q)] if rank < numprocs else []
            assert mr.scan_kv() == want, f"gather({numprocs}) case {case}: wrong pairs on rank {rank}"
            st = mr.kv_stats()
            assert st.npair == len(want) and st.keysize == sum(len(k) for k, _ in want)
            mr.sort_keys(5)                    # the gathered KV is a working KV (collective: every rank calls)
            assert H.canon_kv(mr.scan_kv()) == H.canon_kv(want)
            mr.destroy()

    # broadcast(root), src/mapreduce.cpp:569-625: every rank ends with root's pairs; scrunch = gather + collapse (:2075)
    for case in range(4):
        for root in (0, world - 1):
            mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
            mr.kv_add(shard_records(case, rank))
            want = shard_records(case, root)
            assert mr.broadcast(root) == world * len(want)
            assert mr.scan_kv() == want, f"broadcast({root}) case {case}: wrong pairs on rank {rank}"
            mr.destroy()
        mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
        mr.kv_add(shard_records(case, rank))
        assert mr.scrunch(1, b"all\0") == world            # one KMV pair per rank, only rank 0's has values
        flat = [x for q in range(world) for k, v in shard_records(case, q) for x in (k, v)] if rank == 0 else []
        assert mr.scan_kmv() == [(b"all\0", flat)]
        mr.destroy()

    # IntCount through map_intcount: header-less rows cross the wire when every rank holds them; with one rank
    # set to compact=0 (or one rank empty / holding another shape) the ranks must agree on records instead
    def ic_keys(q, variant):
        return W.intcount_keys_np(q * 80000, (q + 1) * 80000, ndistinct=2000) if not (variant == 2 and q == 0) else np.zeros(0, np.uint32)

    # With the combiner on (default) and every rank holding such rows, what crosses the wire is one row (key, count) per
    # distinct key: same pairs on the same ranks, grouped by key inside each source's segment (include/mrgpu.h "combine").
    for variant, combine in [(v, c) for v in range(4) for c in (1, 0)]:
        mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
        mr.combine = combine
        if variant == 1 and rank == 1:
            mr.compact = 0
        keys = ic_keys(rank, variant)
        shards = [[(struct.pack("<I", int(k)), struct.pack("<i", 1)) for k in ic_keys(q, variant)] for q in range(world)]
        total = sum(len(x) for x in shards)
        assert mr.map_intcount(keys) == total
        if variant == 3:                            # another shape appended on one rank (kv_add is collective: close() sums)
            extra = [(b"wide-key", b"")] if rank == world - 1 else []
            assert mr.kv_add(extra, addflag=1) == total + 1
            shards[world - 1] = shards[world - 1] + [(b"wide-key", b"")]
            total += 1
        assert mr.aggregate() == total
        combined = combine == 1 and variant in (0, 2)
        want = []
        for q in range(world):
            dest = orc.partition(H.kv_encode(shards[q]), world) if shards[q] else []
            seg = [rec for rec, d in zip(shards[q], dest) if d == rank]
            if combined:
                seg.sort(key=lambda kv_: struct.unpack("<I", kv_[0])[0])
            want += seg
        st = mr.kv_stats()                           # statistics describe the full record stream either way
        assert st.npair == len(want) and st.keysize == sum(len(k) for k, _ in want) and st.alignsize == 16 * len(want)
        assert mr.scan_kv() == want, f"intcount variant {variant} combine {combine}: wrong pairs/order on rank {rank}"
        if variant == 0:                             # once from the expanded records (scan above), once from the rows as delivered,
            mr.ahead = 0                             # device input and no counts made at emit time: aggregate() counts itself
            assert mr.map_intcount(torch.from_numpy(keys.view(np.int32)).cuda()) == total
            assert mr.aggregate() == total
        nuniq = mr.convert()
        assert mr.reduce("count") == nuniq
        got = mr.scan_kv()
        bc = {}
        for q in range(world):
            for k, _ in shards[q]:
                bc[k] = bc.get(k, 0) + 1
        mine = {k: c for k, c in bc.items() if orc.hashlittle(k, world) % world == rank}
        assert {k: struct.unpack("<i", v)[0] for k, v in got} == mine, f"intcount variant {variant} combine {combine}: counts differ on rank {rank}"
        counts = [None] * world
        dist.all_gather_object(counts, sum(struct.unpack("<i", v)[0] for _, v in got))
        assert sum(counts) == total
        mr.destroy()

    # reduce(cull) of a factored KMV hands the key array over as header-less rows (the shared value kept once); a second
    # aggregate() sends such rows as they are: 8- and 16-byte keys, with and without a value
    def cull_shard(q, kb, vb):
        rs = np.random.RandomState(77 * q + kb + vb)
        return [(rs.randint(0, 40, size=kb // 4).astype("<u4").tobytes(), b"\x07" * vb) for _ in range(3000)]

    for kb, vb in ((8, 4), (16, 8), (16, 0), (8, 0)):
        mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
        mr.kv_add(cull_shard(rank, kb, vb))
        mr.collate()
        nu = mr.reduce("cull")
        want = sorted({rec for q in range(world) for rec in cull_shard(q, kb, vb) if orc.hashlittle(rec[0], world) % world == rank})
        assert sorted(mr.scan_kv()) == want, f"cull rows kb {kb} vb {vb}: wrong pairs on rank {rank}"
        # every rank appends the same pairs once more; the next exchange carries rows and records mixed per rank
        if rank % 2 == 0:
            mr.kv_add(cull_shard(rank, kb, vb)[:500], addflag=1)
        else:
            mr.kv_add([], addflag=1)
        extra = sum(500 for q in range(world) if q % 2 == 0)
        assert mr.aggregate() == nu + extra
        want2 = sorted(want + [rec for q in range(0, world, 2) for rec in cull_shard(q, kb, vb)[:500] if orc.hashlittle(rec[0], world) % world == rank])
        assert sorted(mr.scan_kv()) == want2, f"cull rows + append kb {kb} vb {vb}: wrong pairs on rank {rank}"
        assert mr.convert() == nu
        mr.reduce("count")
        assert sorted(k for k, _ in mr.scan_kv()) == [k for k, _ in want]
        mr.destroy()

    # R-MAT with one drand48 stream per rank (oink/rmat.cpp:94 srand48(seed+me)) + degree, vs the reference run with P virtual ranks
    import json

    man = json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))
    gold = np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))
    for ci, case in enumerate(man["rmat"]):
        N, ef, a, b, c, d, fr, seed, P, dup = case["args"]
        if P != world:
            continue
        mr = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
        state = ((seed + rank) << 16) | 0x330E
        ntotal = nremain = (1 << N) * ef
        niter = 0
        while nremain:
            niter += 1
            ng = nremain // world + (1 if rank < nremain % world else 0)
            _, state = mr.map_rmat(ng, N, a, b, c, d, fr, state, addflag=1)
            nunique = mr.collate()
            mr.reduce("cull")
            nremain = ntotal - nunique
        assert niter == case["iterations"]
        mrv = g.MapReduce(device=local, rank=rank, nranks=world, unique_id=uid)
        mrv.map_mr(mr, "edge_to_vertex" if dup == 1 else "edge_to_vertices")
        mrv.collate()
        mrv.reduce("count")
        ge, gd = [None] * world, [None] * world
        dist.all_gather_object(ge, mr.scan_kv())
        dist.all_gather_object(gd, mrv.scan_kv())
        if rank == 0:
            assert H.canon_kv([r for p in ge for r in p]) == H.canon_kv(H.kv_decode(gold[f"rmat{ci}.edges"].tobytes()))
            assert H.canon_kv([r for p in gd for r in p]) == H.canon_kv(H.kv_decode(gold[f"rmat{ci}.degree"].tobytes()))
        mr.destroy()
        mrv.destroy()
    dist.barrier()
    if rank == 0:
        print("MULTI_OK", flush=True)
    dist.destroy_process_group()


def main_gloo():
    dist.init_process_group("gloo")
    orc = H.Oracle()
    L = g.lib()
    for case in range(4):
        recs = shard_records(case, rank)
        dest = orc.partition(H.kv_encode(recs), world)
        # what aggregate() does on the device, restated with the oracle: stable grouping by destination
        by_dest = [[rec for rec, d in zip(recs, dest) if d == p] for p in range(world)]
        pages = [b"".join(H.pages_decode(orc.kv_pages(H.kv_encode(x), 1 << 30))[0][1:2]) if x else b"" for x in by_dest]
        shapes = {(len(k), len(v)) for k, v in recs}
        kind = 0 if not recs else (2 if len(shapes) == 1 else 1)
        shape = 0
        if kind == 2:
            kb, vb = next(iter(shapes))
            shape = (kb << 32) | vb
        row = []
        for p in range(world):
            row += [len(by_dest[p]), len(pages[p])]
        row += [kind, shape]
        rows = [None] * world
        dist.all_gather_object(rows, row)
        table = (C.c_uint64 * (len(row) * world))(*[x for r in rows for x in r])
        rc_, rb_, rd_, rr_ = [(C.c_uint64 * (world + 1))() for _ in range(4)]
        sh, kd = C.c_uint64(), C.c_int()
        assert L.mrgpu_plan_exchange(world, rank, table, rc_, rb_, rd_, rr_, C.byref(sh), C.byref(kd)) == 0
        # move the payload with gloo and lay it out by the plan
        sent = [None] * world
        dist.all_gather_object(sent, pages)
        recvbuf = bytearray(rd_[world])
        for q in range(world):
            chunk = sent[q][rank]
            assert len(chunk) == rb_[q]
            recvbuf[rd_[q]:rd_[q] + rb_[q]] = chunk
        want = expected_after_aggregate(case, rank, orc)
        assert rr_[world] == len(want)
        assert H.kv_page_records(bytes(recvbuf), len(want)) == want, f"case {case}: plan delivers wrong order"
        allshapes = {(len(k), len(v)) for q in range(world) for k, v in shard_records(case, q)}
        assert kd.value == (0 if not allshapes else (2 if len(allshapes) == 1 else 1))
        if kd.value == 2:
            kb, vb = next(iter(allshapes))
            assert sh.value == (kb << 32) | vb
    dist.barrier()
    if rank == 0:
        print("MULTI_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main_gloo() if BACKEND == "gloo" else main_nccl()

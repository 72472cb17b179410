#!/usr/bin/env python
"""Headline benchmark: KV pairs/s through map -> aggregate -> convert -> reduce(count) -> sort_keys(1) on the
IntCount workload of BASELINE.json (configs[1]: 2^30 uint32 keys per GPU, Zipf(1.1) over 2^24 distinct).

  python bench.py [--gpus N] [--steps K] [--warmup W]            this repo's CUDA path (one rank per GPU)
  python bench.py --impl reference [...]                          the reference's own CPU path (oracle/_ref)

One step = one full pass of the hot path over one batch of 2^30 synthetic keys per GPU.
  value : inputs already resident in HBM when the timed region starts (CUDA events, max over ranks)
  e2e   : same call sequence through the C-ABI with the keys in pinned HOST memory (H2D inside the timed
          region) and the final (key,count) pages read back to the host (D2H inside the timed region)
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "KV pairs/sec map->aggregate->convert->reduce"
UNIT = "pairs/s"
NDISTINCT = 1 << 24
EXPONENT = 1.1
REF_SAMPLE_LOG2 = 23  # per-process bounded CPU sample of the same key stream (a few seconds of reference CPU work each)


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------ clocks

class ClockSampler:
    """SM clock and throttle reasons sampled while the timed region runs (B200_PROFILING.md recipe).  A step is a
    few milliseconds, so the sampler reads NVML in-process every 2 ms (nvidia-smi -lms cannot start that fast);
    nvidia-smi at 100 ms is the fallback when pynvml is not usable."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    BITS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.path = None
        self.thread = None
        self.samples = []
        self.stop_flag = False

    def _nvml_loop(self, nv, h):
        import time

        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((sm, r))
            except Exception:
                pass
            time.sleep(0.002)

    def start_nvml(self):
        import threading

        import pynvml as nv
        import torch

        nv.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(self.idx).uuid)
            h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = nv.nvmlDeviceGetHandleByIndex(self.idx)
        self.sm_max = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        self.thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
        self.thread.start()

    def stop_nvml(self):
        self.stop_flag = True
        self.thread.join(timeout=2)
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.samples:
            sm = sorted(x for x, _ in self.samples)
            bits = 0
            for _, r in self.samples:
                bits |= int(r)
            out.update(sm_mhz=float(sm[len(sm) // 2]), sm_max_mhz=float(self.sm_max), reasons=[n for b, n in self.BITS if bits & b],
                       samples=len(sm), source="nvml, 2 ms period")
        return out

    def start(self):
        try:
            self.start_nvml()
            return
        except Exception:
            self.thread = None
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.thread:
            return self.stop_nvml()
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx.append(float(f[2]))
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------ reference arm

def run_reference_sample(log2n_each: int, steps: int, warmup: int, nprocs: int | None = None):
    """Times the UNMODIFIED reference library (oracle/_ref/refdrv, built from /root/reference/src by
    oracle/Makefile) on a bounded sample of the same workload: map emit, aggregate, convert, reduce(count),
    sort_keys(1).  The reference is single-threaded per MPI rank and there is no MPI in the image, so "all
    host cores" = C independent single-rank processes (serial mpistubs build) on disjoint shards of the key
    stream, no shuffle between them: an embarrassingly-parallel upper bound for the reference.  Falls back
    to the C restatement (oracle/libmr_oracle.so, kind "port", 1 core) only if the reference binary is absent."""
    from concurrent.futures import ThreadPoolExecutor

    from gpu_mapreduce_b200 import workloads

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import mrhelpers as H

    n = 1 << log2n_each
    kind = "reference" if H.have_ref() else "port"
    C = 1 if kind == "port" else (nprocs or max(1, min(os.cpu_count() or 1, 32)))
    OPS = ("map", "aggregate", "convert", "reduce_count", "sort_keys:1")
    cdf = workloads.zipf_cdf(NDISTINCT, EXPONENT)
    walls, singles = [], []
    with tempfile.TemporaryDirectory() as td:
        shards = []
        for c in range(C):
            keys = workloads.intcount_keys_np(c * n, (c + 1) * n, ndistinct=NDISTINCT, exponent=EXPONENT, cdf=cdf)
            fin = os.path.join(td, f"keys{c}.u32")
            keys.tofile(fin)
            shards.append((fin, os.path.join(td, f"out{c}.kv"), keys if kind == "port" else None))

        def one(sh):
            fin, fout, keys = sh
            if kind == "reference":
                # memsize 1024 MB keeps the shard in core (the reference's out-of-core path is disk I/O)
                t, _ = H.ref_run(["intcount", fin, fout, 1024], timeout=3600)
                return sum(x["secs"] for x in t if x["op"] in OPS)
            orc = H.Oracle()
            t0 = time.perf_counter()
            orc.sort_keys(orc.reduce(orc.convert(orc.intcount_map(keys)), "count"), 1)
            return time.perf_counter() - t0

        for it in range(warmup + steps):
            with ThreadPoolExecutor(max_workers=C) as ex:
                secs = list(ex.map(one, shards))
            if it >= warmup:
                walls.append(max(secs))               # all shards run concurrently: the job ends with the slowest
                singles.append(n / (sum(secs) / len(secs)))
    per_step = sum(walls) / len(walls)
    return {"value": C * n / per_step, "unit": UNIT, "cores": C, "kind": kind,
            "single_core_value": sum(singles) / len(singles),
            "sample": f"{C} concurrent single-rank processes x 2^{log2n_each} keys (disjoint shards of the same Zipf({EXPONENT}) over 2^24 "
                      f"stream), in core (memsize 1024), library op times map+aggregate+convert+reduce(count)+sort_keys(1), job time = slowest "
                      f"shard; no shuffle between shards (no MPI in the image): embarrassingly-parallel upper bound; {os.cpu_count()} host cores present"}, per_step, C * n


def main_reference(args, rank, world):
    if rank != 0:
        return 0
    steps, warmup = args.steps, min(args.warmup, 1)
    cb, per_step, npairs = run_reference_sample(REF_SAMPLE_LOG2, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 keys / byte records", "data": "synthetic",
        "config": {"workload": "IntCount: uint32 keys Zipf(1.1) over 2^24 distinct, count via collate+reduce (BASELINE configs[1])",
                   "pairs_per_step": npairs, "pipeline": "map,aggregate,convert,reduce(count),sort_keys(1)"},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------ parity checks inside the bench
# The driver's test box has one GPU, so the scaling run is the only place where the N > 1 path meets a checker
# the driver can see.  Everything here runs OUTSIDE the timed regions and uses neither oracle/ nor the library's
# own hash: placement is checked against an independent restatement of Jenkins' lookup3 hashlittle()
# (src/hash.cpp:129-333; the partition rule is hashlittle(key,keybytes,nprocs) % nprocs, src/mapreduce.cpp:469-472).

_M32 = 0xFFFFFFFF


def _rot(x, k):
    return ((x << k) | (x >> (32 - k))) & _M32


def hashlittle_py(key: bytes, initval: int) -> int:
    n = len(key)
    a = b = c = (0xDEADBEEF + n + initval) & _M32
    i = 0
    while n > 12:
        a = (a + int.from_bytes(key[i:i + 4], "little")) & _M32
        b = (b + int.from_bytes(key[i + 4:i + 8], "little")) & _M32
        c = (c + int.from_bytes(key[i + 8:i + 12], "little")) & _M32
        a = (a - c) & _M32; a ^= _rot(c, 4); c = (c + b) & _M32
        b = (b - a) & _M32; b ^= _rot(a, 6); a = (a + c) & _M32
        c = (c - b) & _M32; c ^= _rot(b, 8); b = (b + a) & _M32
        a = (a - c) & _M32; a ^= _rot(c, 16); c = (c + b) & _M32
        b = (b - a) & _M32; b ^= _rot(a, 19); a = (a + c) & _M32
        c = (c - b) & _M32; c ^= _rot(b, 4); b = (b + a) & _M32
        i += 12
        n -= 12
    if n == 0:
        return c
    t = key[i:] + b"\0" * (12 - n)
    a = (a + int.from_bytes(t[0:4], "little")) & _M32
    b = (b + int.from_bytes(t[4:8], "little")) & _M32
    c = (c + int.from_bytes(t[8:12], "little")) & _M32
    c ^= b; c = (c - _rot(b, 14)) & _M32
    a ^= c; a = (a - _rot(c, 11)) & _M32
    b ^= a; b = (b - _rot(a, 25)) & _M32
    c ^= b; c = (c - _rot(b, 16)) & _M32
    a ^= c; a = (a - _rot(c, 4)) & _M32
    b ^= a; b = (b - _rot(a, 14)) & _M32
    c ^= b; c = (c - _rot(b, 24)) & _M32
    return c


def hashlittle_u32_np(keys, initval: int):
    """hashlittle of every 4-byte little-endian key of a uint32 array (numpy, wraps mod 2^32)."""
    import numpy as np

    def rot(x, k):
        return (x << np.uint32(k)) | (x >> np.uint32(32 - k))

    with np.errstate(over="ignore"):
        init = np.uint32((0xDEADBEEF + 4 + initval) & _M32)
        a = keys.astype(np.uint32) + init
        b = np.full_like(a, init)
        c = np.full_like(a, init)
        c ^= b; c -= rot(b, 14)
        a ^= c; a -= rot(c, 11)
        b ^= a; b -= rot(a, 25)
        c ^= b; c -= rot(b, 16)
        a ^= c; a -= rot(c, 4)
        b ^= a; b -= rot(a, 14)
        c ^= b; c -= rot(b, 24)
    return c


assert hashlittle_py(b"abcd", 0) == 0xB5F4889C and hashlittle_py(b"hello world!!", 8) == 0x677DEF32 and hashlittle_py(b"", 0) == 0xDEADBEEF


def _case_records(case, r):
    """small deterministic shards for the multi-rank cases run during warm-up (same shapes as tests/multi_rank_worker.py)"""
    import struct

    import numpy as np

    rs = np.random.RandomState(7000 + 100 * case + r)
    if case == 0:    # uniform 4-byte keys / 4-byte values (records, added from the host)
        return [(struct.pack("<I", int(k)), struct.pack("<i", int(v))) for k, v in zip(rs.zipf(1.3, 6000) % 3000, rs.randint(0, 5, 6000))]
    if case == 1:    # variable-length keys and values
        vocab = [bytes(rs.randint(97, 123, size=rs.randint(1, 14)).astype(np.uint8)) + b"\0" for _ in range(300)]
        return [(vocab[rs.randint(len(vocab))], bytes(rs.randint(0, 256, size=rs.randint(0, 6)).astype(np.uint8))) for _ in range(4000)]
    if case == 2:    # mixed: even ranks uniform, odd ranks variable
        if r % 2 == 0:
            return [(struct.pack("<Q", int(x)), b"") for x in rs.randint(0, 500, size=3000)]
        return [(bytes(rs.randint(0, 256, size=rs.randint(1, 12)).astype(np.uint8)), b"v") for _ in range(2000)]
    if case == 3:    # one rank empty, 16-byte keys (R-MAT edges)
        return [] if r == 0 else [(struct.pack("<QQ", int(a), int(b)), b"") for a, b in rs.randint(0, 40, size=(3000, 2))]
    raise ValueError(case)


def multirank_cases(MapReduce, local, rank, world, uid, stream):
    """aggregate()'s placement + delivery order, collate + reduce(count), gather and broadcast on small shards of four
    record shapes, every result compared with a host computation.  Returns the number of cases checked; raises on
    the first difference.  Collective: every rank calls it."""
    ncase = 0
    for case in range(4):
        shards = [_case_records(case, q) for q in range(world)]
        total = sum(len(x) for x in shards)
        dest = [[hashlittle_py(k, world) % world for k, _ in sh] for sh in shards]
        mr = MapReduce(device=local, stream=stream, rank=rank, nranks=world, unique_id=uid)
        assert mr.kv_add(shards[rank]) == total
        assert mr.aggregate() == total
        want = [rec for q in range(world) for rec, d in zip(shards[q], dest[q]) if d == rank]   # by source rank, original order
        assert mr.scan_kv() == want, f"multi-rank case {case}: aggregate() delivered wrong pairs/order on rank {rank}"
        nu = mr.convert()
        assert mr.reduce("count") == nu
        cnt = {}
        for k, _ in want:
            cnt[k] = cnt.get(k, 0) + 1
        got = {k: int.from_bytes(v, "little") for k, v in mr.scan_kv()}
        assert got == cnt, f"multi-rank case {case}: counts differ on rank {rank}"
        mr.destroy()
        ncase += 1
        mr = MapReduce(device=local, stream=stream, rank=rank, nranks=world, unique_id=uid)
        mr.kv_add(shards[rank])
        assert mr.gather(1) == total            # src/mapreduce.cpp:893-1036: everything on rank 0, in rank order
        assert mr.scan_kv() == ([rec for sh in shards for rec in sh] if rank == 0 else [])
        mr.destroy()
        mr = MapReduce(device=local, stream=stream, rank=rank, nranks=world, unique_id=uid)
        mr.kv_add(shards[rank])
        root = world - 1
        assert mr.broadcast(root) == world * len(shards[root])   # src/mapreduce.cpp:569-625
        assert mr.scan_kv() == shards[root]
        mr.destroy()
        ncase += 2
    return ncase


def bind_near_gpu(local: int):
    """Run this rank on the CPU cores next to its GPU (NVML's ideal affinity), so that the pinned host buffers it
    allocates afterwards are on that socket's memory: at 8 ranks the end-to-end path is eight concurrent 4 GiB DMAs,
    and buffers on the far socket cross the inter-socket link.  Host placement only; returns what it did."""
    try:
        import pynvml as nv
        import torch

        nv.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(local).uuid)
            h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = nv.nvmlDeviceGetHandleByIndex(local)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} cores near GPU {local}"
    except Exception as ex:  # placement is an optimisation, never a requirement
        return f"not bound ({type(ex).__name__})"
    return "not bound"


# ------------------------------------------------------------------------------------------ CUDA arm

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2n", type=int, default=30, help="pairs per GPU per step = 2^log2n (BASELINE: 30)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-reference-layout", action="store_true", help="skip the compact=0/ahead=0/combine=0 run of the same pipeline")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return main_reference(args, rank, world)
    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps

    # libraries (NCCL prints its version) may write to stdout: keep fd 1 for the one JSON line only
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import numpy as np
    import torch

    from gpu_mapreduce_b200 import MapReduce, lib, workloads
    import ctypes as C

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    all_cpus = os.sched_getaffinity(0)
    host_binding = bind_near_gpu(local)
    uid = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        t = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(MapReduce.unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().numpy().tobytes())

    stream = torch.cuda.current_stream().cuda_stream
    mr = MapReduce(device=local, stream=stream, rank=rank, nranks=world, unique_id=uid)
    n = 1 << args.log2n
    cdf_t = torch.from_numpy(workloads.zipf_cdf(NDISTINCT, EXPONENT)).to(dev)
    keys = workloads.intcount_keys_torch(rank * n, (rank + 1) * n, dev, ndistinct=NDISTINCT, exponent=EXPONENT, cdf_t=cdf_t)
    del cdf_t
    torch.cuda.empty_cache()
    L = lib()
    peak, peak_src = measured_peak_gbs()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def make_step(h):
        def step(inp):
            h.map_intcount(inp, n)
            h.aggregate()
            h.convert()
            nu = h.reduce("count")
            h.sort_keys(1)
            return nu
        return step

    step = make_step(mr)

    def timed(nsteps, fn):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(nsteps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        barrier()
        return float(ms.item())

    def allmax(x):
        t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def read_back(h, pinned):
        off = 0
        for ip in range(h.kv_stats().npage):
            asz = C.c_uint64()
            h._ck(L.mrgpu_kv_page_info(h._h, ip, None, None, None, C.byref(asz)))
            h._ck(L.mrgpu_kv_page_copy(h._h, ip, pinned.data_ptr() + off))
            off += asz.value
        return off

    # ---- ground truth for the parity check: bincount of every rank's input shard, summed over ranks (outside any timed region)
    truth = torch.zeros(NDISTINCT, dtype=torch.int64, device=dev)
    for lo in range(0, n, 1 << 28):
        truth += torch.bincount(keys[lo:lo + (1 << 28)], minlength=NDISTINCT)
    if world > 1:
        dist.all_reduce(truth)

    def parity(h, what):
        """every (key, count) this rank holds after the pipeline == the global count of that key; every key sits on rank
        hashlittle(key,4,P) % P; the ranks together hold every key that occurs.  Full comparison, no sampling."""
        st = h.kv_stats()
        pinned = torch.empty(max(st.alignsize, 16), dtype=torch.uint8, pin_memory=True)
        nb = read_back(h, pinned)
        rec = pinned[:nb].numpy().view(np.int32).reshape(-1, 4)
        shape_ok = bool((rec[:, 0] == 4).all() and (rec[:, 1] == 4).all())
        sorted_ok = bool((np.diff(rec[:, 2]) > 0).all())
        k = torch.from_numpy(rec[:, 2].astype(np.int64)).to(dev)
        c = torch.from_numpy(rec[:, 3].astype(np.int64)).to(dev)
        in_range = bool(((k >= 0) & (k < NDISTINCT)).all().item()) if k.numel() else True
        mism = int((truth[k] != c).sum().item()) if in_range else int(k.numel())
        place = hashlittle_u32_np(rec[:, 2].astype(np.uint32), world) % np.uint32(world) if world > 1 else np.zeros(len(rec), np.uint32)
        misplaced = int((place != rank).sum())
        tot = torch.tensor([len(rec), int(rec[:, 3].astype(np.int64).sum()), mism, misplaced, 0 if (shape_ok and sorted_ok and in_range) else 1],
                           device=dev, dtype=torch.int64)
        if world > 1:
            dist.all_reduce(tot)
        nkeys, npairs, mism, misplaced, malformed = (int(x) for x in tot.tolist())
        expect_keys = int((truth > 0).sum().item())
        ok = mism == 0 and misplaced == 0 and malformed == 0 and nkeys == expect_keys and npairs == world * n
        return {"what": what, "n_keys": nkeys, "n_keys_expected": expect_keys, "pairs": npairs, "mismatches": mism + abs(nkeys - expect_keys),
                "placement_ok": misplaced == 0, "sorted_unique_per_rank": malformed == 0, "ok": ok}

    # ---- device-resident input
    nunique = 0
    cases = 0
    for w in range(args.warmup):
        nunique = step(keys)
        if w == 0 and world > 1:   # the other record shapes / exchange paths / gather / broadcast, checked on small shards
            cases = multirank_cases(MapReduce, local, rank, world, uid, stream)
    mr.prof(False)  # resets the launch counter
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed(args.steps, lambda: step(keys))
    clocks = sampler.stop()
    _, launches = mr.prof_report()
    value = world * n * args.steps / (ms / 1e3)
    step_ms = ms / args.steps
    res_bytes = mr.kv_stats().alignsize
    checks = [parity(mr, "headline pipeline, last timed step")]

    # ---- per-kernel timing (one extra step, events around every tagged launch)
    def kernel_table(h, fn):
        h.prof(True)
        fn()
        kern, _ = h.prof_report()
        h.prof(False)
        for k in kern:
            k["ms_per_launch"] = k["ms"] / max(k["calls"], 1)
            k["gbs"] = (k["bytes"] / 1e9) / (k["ms"] / 1e3) if k["ms"] > 0 and k["bytes"] > 0 else None
            k["moved_gbs"] = (k.get("moved", 0) / 1e9) / (k["ms"] / 1e3) if k["ms"] > 0 and k.get("moved", 0) > 0 else None
        return kern

    def shuffle_block(kern):
        """max over GPUs of max(bytes sent, bytes received over NVLink) / exchange time"""
        x = [k for k in kern if k["tag"] == "aggregate_exchange_nvlink" and k["ms"] > 0]
        sent = sum(k["bytes"] for k in x)
        recv = sum(k.get("moved", 0) for k in x)
        ms_x = sum(k["ms"] for k in x)
        bmax, tmax = allmax(max(sent, recv)), allmax(ms_x)
        if tmax <= 0:
            return None
        return {"bus_gbs": bmax / 1e9 / (tmax / 1e3), "bytes_max_over_ranks": bmax, "ms_max_over_ranks": tmax, "nominal_gbs_per_direction": 900.0,
                "frac_of_nominal": bmax / 1e9 / (tmax / 1e3) / 900.0, "rank0": {"sent": sent, "received": recv, "ms": ms_x},
                "definition": "max over ranks of max(sent, received) NVLink bytes / max over ranks of the exchange time (scatter kernel + completion barrier)"}

    kern = kernel_table(mr, lambda: step(keys))
    shuffle = None
    if world > 1:
        shuffle = {"headline": shuffle_block(kern)}
        if shuffle["headline"]:
            shuffle["headline"]["note"] = ("combiner on: each rank ships one (key,count) row per distinct key, so the exchange is latency- not "
                                           "bandwidth-bound; the plain exchange of every pair is measured below")
        # the plain exchange (every pair crosses, reference delivery order) on the same input, one untimed + one profiled step
        mr.combine = 0
        step(keys)
        kern_plain = kernel_table(mr, lambda: step(keys))
        checks.append(parity(mr, "combine=0 (plain exchange of every pair)"))
        shuffle["plain"] = shuffle_block(kern_plain)
        if shuffle["plain"]:
            shuffle["plain"]["step_ms_rank0"] = sum(k["ms"] for k in kern_plain if k["tag"] != "aggregate_exchange_nvlink")
            shuffle["plain"]["note"] = "combine=0: 4 bytes per pair cross NVLink; Zipf's hottest keys make one rank's ingress the bound"
        mr.combine = 1
    dom = max([k for k in kern if k["tag"] != "aggregate_exchange_nvlink"], key=lambda k: k["ms"]) if kern else None
    roofline = None
    traffic, traffic_src = None, None
    try:  # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (not re-measured in this run)
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if dom and dom["tag"] in tj and tj[dom["tag"]]["pairs"] == n:
            traffic = tj[dom["tag"]]["dram_bytes"]
            traffic_src = "not re-measured in this run: " + tj[dom["tag"]].get("capture", "committed ncu capture")
    except Exception:
        traffic = None

    def roofline_of(dom, kern, step_ms, use_moved):
        # achieved = bytes per launch / launch time.  use_moved: the bytes the kernel has to move with the layout that is
        # in HBM (header-less KV); else the reference record layout's exact sizes
        g = (dom["moved_gbs"] or dom["gbs"]) if use_moved else dom["gbs"]
        return {"bound": "hbm", "kernel": dom["tag"], "achieved": g, "peak": peak, "unit": "GB/s", "frac": (g / peak) if g else None,
                "peak_source": peak_src, "launches_per_step": dom["calls"], "ms_per_launch": dom["ms_per_launch"],
                "share_of_step": dom["ms"] / step_ms,
                "algorithmic_bytes_per_launch": ((dom.get("moved", 0) or dom["bytes"]) if use_moved else dom["bytes"]) / max(dom["calls"], 1),
                "kernels": [{"tag": k["tag"], "ms": round(k["ms"], 4), "calls": k["calls"],
                             "gbs": None if k["gbs"] is None else round(k["gbs"], 1),
                             "moved_gbs": None if k["moved_gbs"] is None else round(k["moved_gbs"], 1)}
                            for k in sorted(kern, key=lambda k: -k["ms"])]}

    if dom:
        roofline = roofline_of(dom, kern, step_ms, True)
        roofline["traffic"] = traffic
        roofline["traffic_source"] = traffic_src
        roofline["reference_layout"] = {"algorithmic_bytes_per_launch": dom["bytes"] / max(dom["calls"], 1), "achieved": dom["gbs"],
                                        "frac": (dom["gbs"] / peak) if dom["gbs"] else None}

    # ---- end to end: pinned host input, H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        hkeys = torch.empty(n, dtype=torch.int32, pin_memory=True)
        hkeys.copy_(keys)
        torch.cuda.synchronize()
        pinned_out = torch.empty(max(res_bytes, 16), dtype=torch.uint8, pin_memory=True)

        def e2e_step():
            step(hkeys)
            read_back(mr, pinned_out)

        e2e_step()
        ems = timed(args.steps, e2e_step)
        e2e = {"value": world * n * args.steps / (ems / 1e3), "unit": UNIT, "h2d_bytes_per_step": 4 * n, "d2h_bytes_per_step": int(res_bytes),
               "ms_per_step": ems / args.steps, "host_placement": host_binding}
        checks.append(parity(mr, "end-to-end (host keys in, result pages out), last timed step"))
        del hkeys, pinned_out
    mr.destroy()
    del mr

    # ---- the general path on the reference's record layout: compact=0 (every store holds the reference's record stream),
    # ahead=0 (no pass fused across ops), combine=0 (every pair crosses the wire): what SURVEY.md 8d's byte counts describe
    ref_layout = None
    if not args.no_reference_layout:
        torch.cuda.empty_cache()
        mr2 = MapReduce(device=local, stream=stream, rank=rank, nranks=world, unique_id=uid)
        mr2.compact, mr2.ahead, mr2.combine = 0, 0, 0
        step2 = make_step(mr2)
        rsteps = max(1, min(args.steps, 3))
        step2(keys)
        step2(keys)
        rms = timed(rsteps, lambda: step2(keys)) / rsteps
        checks.append(parity(mr2, "reference record layout run (compact=0, ahead=0, combine=0)"))
        kern2 = kernel_table(mr2, lambda: step2(keys))
        dom2 = max([k for k in kern2 if k["tag"] != "aggregate_exchange_nvlink"], key=lambda k: k["ms"]) if kern2 else None
        U = nunique / world
        kmvb = 16 * U + 8 * n
        alg = n * 20 + (32 * n if world > 1 else 0) + (16 * n + kmvb) + (kmvb + 16 * U) + 2 * 16 * U
        ref_layout = {"settings": {"compact": 0, "ahead": 0, "combine": 0}, "steps": rsteps, "ms_per_step": rms,
                      "value": world * n / (rms / 1e3), "unit": UNIT,
                      "algorithmic_bytes_per_step_per_gpu": alg, "pipeline_gbs_per_gpu": alg / 1e9 / (rms / 1e3),
                      "pipeline_frac_of_peak": alg / 1e9 / (rms / 1e3) / peak,
                      "roofline": roofline_of(dom2, kern2, rms, False) if dom2 else None,
                      "shuffle": shuffle_block(kern2) if world > 1 else None,
                      "note": "bytes are SURVEY.md 8d's exact sizes: map 20 B/pair, aggregate 32 (P>1), convert 16 + KMVb, reduce KMVb + 16U, sort 2*16U"}
        mr2.destroy()
        del mr2

    os.sched_setaffinity(0, all_cpus)  # the CPU baseline below uses every host core
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu_baseline, _, _ = run_reference_sample(REF_SAMPLE_LOG2, 1, 0)
        except Exception as ex:  # the baseline is reported, never required for the GPU number
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 1, "kind": "reference", "sample": f"failed: {ex}"}

    all_ok = all(c["ok"] for c in checks)
    if rank == 0:
        # algorithmic bytes of the whole step: map 20 B/pair, aggregate 32 (P>1), convert 16 + KMV, reduce KMV + KV
        U = nunique / world
        kmvb = 16 * U + 8 * n
        alg = n * 20 + (32 * n if world > 1 else 0) + (16 * n + kmvb) + (kmvb + 16 * U) + 2 * 16 * U
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 keys / byte records", "data": "synthetic",
            "config": {"workload": "IntCount: 2^%d uint32 keys per GPU, Zipf(1.1) over 2^24 distinct, count via collate+reduce "
                                   "(BASELINE configs[1])" % args.log2n,
                       "pairs_per_gpu_per_step": n, "pipeline": "map,aggregate,convert,reduce(count),sort_keys(1)",
                       "unique_keys": int(nunique), "l2": "inputs (4 GiB keys, 16 GiB KV) exceed the 126 MB L2; no flush needed",
                       "parallelism": f"hash-partitioned ranks x{world}" if world > 1 else "1 rank"},
            "gpu_launches": int(launches), "clocks": clocks, "e2e": e2e, "roofline": roofline, "cpu_baseline": cpu_baseline,
            "shuffle": shuffle,
            "parity_check": {"ok": all_ok, "n_keys": checks[0]["n_keys"], "mismatches": sum(c["mismatches"] for c in checks),
                             "placement_ok": all(c["placement_ok"] for c in checks), "cases": cases, "runs": checks,
                             "how": "every (key,count) of every rank against torch.bincount of all input shards (all-reduced); key placement "
                                    "against an independent lookup3 restatement; at N>1 also 4 record shapes x (aggregate order, collate+count, "
                                    "gather, broadcast) on small shards"},
            "reference_layout_run": ref_layout,
            # what the whole step would have to move in the reference's record layout and the rate that
            # corresponds to; with the compact stores the step moves a fraction of it, so this is not a utilisation
            "pipeline_reference_layout_bytes_per_step_per_gpu": alg,
            "pipeline_reference_layout_gbs_per_gpu": alg / 1e9 / (step_ms / 1e3),
        }
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.destroy_process_group()
    if not all_ok:
        sys.stderr.write("PARITY CHECK FAILED: " + json.dumps(checks) + "\n")
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())

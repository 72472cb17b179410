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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(inp):
        mr.map_intcount(inp, n)
        mr.aggregate()
        mr.convert()
        nu = mr.reduce("count")
        mr.sort_keys(1)
        return nu

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

    # ---- device-resident input
    nunique = 0
    for _ in range(args.warmup):
        nunique = step(keys)
    mr.prof(False)  # resets the launch counter
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed(args.steps, lambda: step(keys))
    clocks = sampler.stop()
    _, launches = mr.prof_report()
    value = world * n * args.steps / (ms / 1e3)

    # ---- checks on the last result (size-independent properties; full parity lives in tests/)
    st = mr.kv_stats()
    local_sum = 0
    L = lib()
    res_bytes = st.alignsize
    pinned_out = torch.empty(max(res_bytes, 16), dtype=torch.uint8, pin_memory=True)

    def read_back():
        off = 0
        for ip in range(mr.kv_stats().npage):
            asz = C.c_uint64()
            mr._ck(L.mrgpu_kv_page_info(mr._h, ip, None, None, None, C.byref(asz)))
            mr._ck(L.mrgpu_kv_page_copy(mr._h, ip, pinned_out.data_ptr() + off))
            off += asz.value
        return off

    nb = read_back()
    rec = pinned_out[:nb].numpy().view(np.int32).reshape(-1, 4)
    assert (rec[:, 0] == 4).all() and (rec[:, 1] == 4).all()
    assert (np.diff(rec[:, 2]) > 0).all(), "sort_keys(1) output is not strictly increasing"
    local_sum = int(rec[:, 3].astype(np.int64).sum())
    tot = torch.tensor([local_sum], device=dev, dtype=torch.int64)
    if world > 1:
        dist.all_reduce(tot)
    assert int(tot.item()) == world * n, f"checksum of counts {int(tot.item())} != pairs {world * n}"

    # ---- per-kernel timing (one extra step, events around every tagged launch)
    mr.prof(True)
    step(keys)
    kern, _ = mr.prof_report()
    mr.prof(False)
    peak, peak_src = measured_peak_gbs()
    for k in kern:
        k["ms_per_launch"] = k["ms"] / max(k["calls"], 1)
        k["gbs"] = (k["bytes"] / 1e9) / (k["ms"] / 1e3) if k["ms"] > 0 and k["bytes"] > 0 else None
        k["moved_gbs"] = (k.get("moved", 0) / 1e9) / (k["ms"] / 1e3) if k["ms"] > 0 and k.get("moved", 0) > 0 else None
    shuffle = None
    for k in kern:
        if k["tag"] == "aggregate_exchange_nvlink" and k["ms"] > 0:
            # bytes = max(sent, received) over NVLink on this rank; reported for rank 0
            shuffle = {"bus_gbs": k["gbs"], "ms": k["ms"], "bytes_over_nvlink": k["bytes"], "nominal_gbs_per_direction": 900.0,
                       "measured_peer_copy_gbs": 770.0, "frac_of_nominal": (k["gbs"] or 0) / 900.0, "note": "rank 0; Zipf skew makes ranks unequal"}
    dom = max([k for k in kern if k["tag"] != "aggregate_exchange_nvlink"], key=lambda k: k["ms"]) if kern else None
    step_ms = ms / args.steps
    roofline = None
    traffic = None
    try:  # measured DRAM bytes per launch of the dominant kernel, from the committed ncu --set full capture
        tj = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_ncu_traffic.json")))
        if dom and dom["tag"] in tj and tj[dom["tag"]]["pairs"] == n:
            traffic = tj[dom["tag"]]["dram_bytes"]
    except Exception:
        traffic = None
    if dom:
        # achieved = the bytes the kernel has to move with the layout that is in HBM (header-less KV,
        # 8 B/pair for the dominant kernel) / its time.  The reference record layout's figure for the same work
        # (SURVEY.md 8d exact sizes, 36 B/pair) is kept beside it: against that one the rate exceeds the HBM peak,
        # which is the point of the layout, not a measurement of the memory system.
        mgbs = dom["moved_gbs"] or dom["gbs"]
        roofline = {"bound": "hbm", "kernel": dom["tag"], "achieved": mgbs, "peak": peak, "unit": "GB/s",
                    "frac": (mgbs / peak) if mgbs else None, "traffic": traffic,
                    "peak_source": peak_src, "launches_per_step": dom["calls"], "ms_per_launch": dom["ms_per_launch"],
                    "share_of_step": dom["ms"] / step_ms,
                    "algorithmic_bytes_per_launch": (dom.get("moved", 0) or dom["bytes"]) / max(dom["calls"], 1),
                    "reference_layout": {"algorithmic_bytes_per_launch": dom["bytes"] / max(dom["calls"], 1), "achieved": dom["gbs"],
                                         "frac": (dom["gbs"] / peak) if dom["gbs"] else None},
                    "kernels": [{"tag": k["tag"], "ms": round(k["ms"], 4), "calls": k["calls"],
                                 "gbs": None if k["gbs"] is None else round(k["gbs"], 1),
                                 "moved_gbs": None if k["moved_gbs"] is None else round(k["moved_gbs"], 1)}
                                for k in sorted(kern, key=lambda k: -k["ms"])]}

    # ---- end to end: pinned host input, H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        hkeys = torch.empty(n, dtype=torch.int32, pin_memory=True)
        hkeys.copy_(keys)
        torch.cuda.synchronize()

        def e2e_step():
            step(hkeys)
            read_back()

        e2e_step()
        ems = timed(args.steps, e2e_step)
        e2e = {"value": world * n * args.steps / (ems / 1e3), "unit": UNIT, "h2d_bytes_per_step": 4 * n, "d2h_bytes_per_step": int(res_bytes),
               "ms_per_step": ems / args.steps}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu_baseline, _, _ = run_reference_sample(REF_SAMPLE_LOG2, 1, 0)
        except Exception as ex:  # the baseline is reported, never required for the GPU number
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 1, "kind": "reference", "sample": f"failed: {ex}"}

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
            # what the whole step would have to move in the reference's record layout and the rate that
            # corresponds to; with the compact stores the step moves a fraction of it, so this is not a utilisation
            "pipeline_reference_layout_bytes_per_step_per_gpu": alg,
            "pipeline_reference_layout_gbs_per_gpu": alg / 1e9 / (step_ms / 1e3),
        }
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    mr.destroy()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

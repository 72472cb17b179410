"""Microbenchmark (not a test): the fused emit+count pass and the count-only pass of convert() on the IntCount key
stream, with the result checked against torch.bincount.   python tools/countbench.py <log2n> [exponent]
MRGPU_LIB=<variant .so> selects an A/B build (e.g. -DMRGPU_DENSE_HOT=0|4|8)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_mapreduce_b200 import MapReduce, workloads  # noqa: E402

log2n = int(sys.argv[1])
expo = float(sys.argv[2]) if len(sys.argv) > 2 else 1.1
n = 1 << log2n
dev = torch.device("cuda", 0)
keys = workloads.intcount_keys_torch(0, n, dev, exponent=expo)
truth = torch.zeros(1 << 24, dtype=torch.int64, device=dev)
for lo in range(0, n, 1 << 28):
    truth += torch.bincount(keys[lo:lo + (1 << 28)], minlength=1 << 24)
out = {}
for ahead in (1, 0):
    mr = MapReduce(device=0, stream=torch.cuda.current_stream().cuda_stream)
    mr.ahead = ahead
    best = {}
    for it in range(6):
        mr.prof(it >= 2)
        mr.map_intcount(keys, n)
        nu = mr.convert()
        if it >= 2:
            for k in mr.prof_report()[0]:
                best[k["tag"]] = min(best.get(k["tag"], 1e9), k["ms"])
    mr.reduce("count")
    import numpy as np
    rec = np.concatenate([np.frombuffer(p, dtype=np.int32) for _, p in mr.kv_pages()]).reshape(-1, 4)
    k = torch.from_numpy(rec[:, 2].astype(np.int64)).to(dev)
    c = torch.from_numpy(rec[:, 3].astype(np.int64)).to(dev)
    ok = bool((truth[k] == c).all().item()) and int(c.sum().item()) == n and len(rec) == int((truth > 0).sum().item())
    out[f"ahead={ahead}"] = {t: round(v, 4) for t, v in best.items() if v > 0.02}
    out[f"ahead={ahead}"]["exact"] = ok
    mr.destroy()
print(os.environ.get("MRGPU_LIB", "default"), f"2^{log2n} zipf {expo}", out, flush=True)

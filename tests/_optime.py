import sys, os, time
sys.path.insert(0, "/root/repo")
import torch
from gpu_mapreduce_b200 import MapReduce, workloads
log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 28
n = 1 << log2n
dev = torch.device("cuda", 0)
mr = MapReduce(device=0, stream=torch.cuda.current_stream().cuda_stream)
keys = workloads.intcount_keys_torch(0, n, dev)
def ev(): 
    e = torch.cuda.Event(enable_timing=True); e.record(); return e
for it in range(5):
    t0 = time.perf_counter()
    e = [ev()]
    mr.map_intcount(keys, n); e.append(ev())
    mr.aggregate(); e.append(ev())
    mr.convert(); e.append(ev())
    mr.reduce("count"); e.append(ev())
    mr.sort_keys(1); e.append(ev())
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    names = ["map", "aggregate", "convert", "reduce", "sort_keys"]
    print(it, "wall %.2f ms |" % ((t1 - t0) * 1e3), " ".join("%s %.2f" % (nm, e[i].elapsed_time(e[i + 1])) for i, nm in enumerate(names)), flush=True)

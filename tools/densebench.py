"""Microbenchmark (not a test): time convert()'s counting kernel on the IntCount key stream.
   python tools/densebench.py <log2n>   [MRGPU_LIB=<variant .so> for A/B runs]"""
import sys, torch
sys.path.insert(0, "/root/repo")
from gpu_mapreduce_b200 import MapReduce, workloads
log2n = int(sys.argv[1])
n = 1 << log2n
mr = MapReduce(device=0, stream=torch.cuda.current_stream().cuda_stream)
keys = workloads.intcount_keys_torch(0, n, torch.device("cuda", 0))
for it in range(4):
    mr.map_intcount(keys, n)
    mr.prof(it == 3)
    nu = mr.convert()
    if it == 3:
        print("unique", nu)
        for k in mr.prof_report()[0]:
            print(k["tag"], round(k["ms"], 4))

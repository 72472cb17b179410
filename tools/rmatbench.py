"""Microbenchmark (not a test): per-kernel times of one R-MAT generate -> collate -> cull iteration.
   python tools/rmatbench.py <nlevels>"""
import sys, torch
sys.path.insert(0, "/root/repo")
from gpu_mapreduce_b200 import MapReduce
N = int(sys.argv[1])
mr = MapReduce(device=0, stream=torch.cuda.current_stream().cuda_stream)
state = (12345 << 16) | 0x330E
n = (1 << N) * 16
for it in range(2):
    mr.prof(it == 1)
    _, st2 = mr.map_rmat(n, N, .57, .19, .19, .05, 0.0, state, addflag=0)
    nu = mr.collate()
    mr.reduce("cull")
tot = 0
for k in sorted(mr.prof_report()[0], key=lambda k: -k["ms"]):
    print(k["tag"], round(k["ms"], 3), k["calls"]); tot += k["ms"]
print("edges", n, "unique", nu, "kernel ms", round(tot, 2))

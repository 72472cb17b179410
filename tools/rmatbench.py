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

# degree: map(mr, edge_to_vertices) -> collate -> reduce(count), run twice (cold, warm) with wall times
import time
for it in range(2):
    mrv = MapReduce(device=0, stream=torch.cuda.current_stream().cuda_stream)
    mrv.prof(True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    mrv.map_mr(mr, "edge_to_vertices")
    torch.cuda.synchronize(); t1 = time.perf_counter()
    nv = mrv.collate()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    mrv.reduce("count")
    torch.cuda.synchronize(); t3 = time.perf_counter()
    print("degree run", it, "wall ms map/collate/reduce", round((t1 - t0) * 1e3, 2), round((t2 - t1) * 1e3, 2), round((t3 - t2) * 1e3, 2), "vertices", nv)
    for k in sorted(mrv.prof_report()[0], key=lambda k: -k["ms"]):
        print("   ", k["tag"], round(k["ms"], 3), k["calls"])
    mrv.destroy()

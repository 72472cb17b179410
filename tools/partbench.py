import sys, torch
sys.path.insert(0, "/root/repo")
from gpu_mapreduce_b200 import MapReduce, workloads
log2n, P = int(sys.argv[1]), int(sys.argv[2])
n = 1 << log2n
mr = MapReduce(device=0, stream=torch.cuda.current_stream().cuda_stream)
keys = workloads.intcount_keys_torch(0, n, torch.device("cuda", 0))
for it in range(4):
    mr.map_intcount(keys, n)
    mr.prof(it == 3)
    c = mr.debug_partition(P)
    if it == 3:
        print(c[:8], sum(c) == n)
        for k in mr.prof_report()[0]: print(k)

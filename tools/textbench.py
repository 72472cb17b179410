"""Microbenchmark (not a test): per-kernel times of map_tokenize on synthetic words.  python tools/textbench.py <log2 bytes>"""
import sys, torch
sys.path.insert(0, "/root/repo")
from gpu_mapreduce_b200 import MapReduce
n = 1 << int(sys.argv[1])
g = torch.Generator(device="cuda"); g.manual_seed(1)
# words of 2..12 letters separated by one space: draw letters, then put a space where a counter hits
t = torch.randint(97, 123, (n,), dtype=torch.uint8, device="cuda", generator=g)
gap = torch.randint(3, 14, (n // 8,), device="cuda", generator=g).cumsum(0)
t[gap[gap < n]] = 32
mr = MapReduce(device=0, stream=torch.cuda.current_stream().cuda_stream)
for it in range(3):
    mr.prof(it == 2)
    nw = mr.map_tokenize(t)
tot = 0
for k in mr.prof_report()[0]:
    print(k["tag"], round(k["ms"], 4)); tot += k["ms"]
print("words", nw, "kernel ms", round(tot, 3), "text GB/s", round(n / tot / 1e6, 1))

# collate + reduce(sum) on the words just emitted (general path: variable-length keys)
for it in range(2):
    mr.map_tokenize(t)
    mr.prof(it == 1)
    nu = mr.collate()
    mr.reduce("sum")
tot = 0
for k in sorted(mr.prof_report()[0], key=lambda k: -k["ms"]):
    print("   ", k["tag"], round(k["ms"], 3), k["calls"]); tot += k["ms"]
print("unique", nu, "collate+reduce kernel ms", round(tot, 2))

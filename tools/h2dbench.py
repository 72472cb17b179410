import torch, time
n = 1 << 30
h = torch.empty(n, dtype=torch.int32, pin_memory=True); h.fill_(7)
d = torch.empty(n, dtype=torch.int32, device="cuda")
for chunk in (n, n // 4, n // 16, n // 64):
    for rep in range(2):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for o in range(0, n, chunk):
            d[o:o + chunk].copy_(h[o:o + chunk], non_blocking=True)
        e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print("H2D chunk", chunk * 4 >> 20, "MiB:", round(ms, 2), "ms", round(n * 4 / ms / 1e6, 1), "GB/s")
hh = torch.empty(1 << 26, dtype=torch.int32, pin_memory=True)
torch.cuda.synchronize(); e0.record(); hh.copy_(d[:1 << 26], non_blocking=True); e1.record(); torch.cuda.synchronize()
print("D2H 256 MiB", round((1 << 28) / e0.elapsed_time(e1) / 1e6, 1), "GB/s")

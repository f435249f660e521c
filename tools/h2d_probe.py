import torch, time
n = 5_502_926_848
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for chunk in (n, n // 4, n // 16, n // 64, n // 342):
    torch.cuda.synchronize()
    best = 1e9
    for r in range(3):
        t0 = time.perf_counter()
        for o in range(0, n - chunk + 1, chunk):
            d[o:o + chunk].copy_(h[o:o + chunk], non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    print(f"chunk {chunk/1e6:9.1f} MB: {n/best/1e9:6.2f} GB/s  ({best*1e3:.1f} ms)")

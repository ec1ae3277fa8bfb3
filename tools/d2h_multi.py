"""Development aid: aggregate device->host bandwidth of several GPUs copying at once (does the host side scale with the
number of PCIe links?).  python tools/d2h_multi.py [MB per GPU]"""
import sys, time, torch
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 128
n = torch.cuda.device_count()
src = [torch.empty(mb << 20, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
dst = [torch.empty(mb << 20, dtype=torch.uint8).pin_memory() for i in range(n)]
streams = [torch.cuda.Stream(device=i) for i in range(n)]
def run(devs, reps=10):
    for _ in range(2):
        for i in devs:
            with torch.cuda.stream(streams[i]): dst[i].copy_(src[i], non_blocking=True)
    for i in devs: streams[i].synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in devs:
            with torch.cuda.stream(streams[i]): dst[i].copy_(src[i], non_blocking=True)
    for i in devs: streams[i].synchronize()
    dt = time.perf_counter() - t0
    return len(devs) * mb * reps / 1024 / dt
sets = [[0]] + [[0, k] for k in range(1, n)] + ([list(range(4))] if n >= 4 else []) + ([[0, 2, 4, 6]] if n >= 8 else []) + ([list(range(n))] if n > 2 else [])
for s in sets:
    print(f"GPUs {s}: {run(s):.1f} GiB/s aggregate", flush=True)

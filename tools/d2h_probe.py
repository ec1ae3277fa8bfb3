"""Development aid: device->pinned-host copy rate of a 132.7 MB frame, alone and while the trace kernels run."""
import importlib, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes; H = rtb.host
n = 3840 * 2160 * 16
src = torch.empty(n, dtype=torch.uint8, device="cuda"); dst = torch.empty(n, dtype=torch.uint8).pin_memory()
cs = torch.cuda.Stream()
def copy_ms(chunks):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(cs):
        e0.record()
        step = n // chunks
        for k in range(chunks): dst[k * step:(k + 1) * step].copy_(src[k * step:(k + 1) * step], non_blocking=True)
        e1.record()
    return e0, e1
for chunks in (1, 7, 16):
    for _ in range(2): e0, e1 = copy_ms(chunks); torch.cuda.synchronize()
    print(f"alone, {chunks} chunks: {e0.elapsed_time(e1):.3f} ms = {n / e0.elapsed_time(e1) / 1e6:.1f} GB/s")
s, p = sc.cfg3_scene(), sc.cfg3_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_scene(s)
for _ in range(3): rt.render_device(p, H.RT_OUT_RGBA32F, sync_stats=True)
for rep in range(3):
    torch.cuda.synchronize()
    for _ in range(3): rt.render_device(p, H.RT_OUT_RGBA32F, sync_stats=False)     # ~7 ms of kernels queued
    e0, e1 = copy_ms(7)
    torch.cuda.synchronize()
    print(f"under load, 7 chunks: {e0.elapsed_time(e1):.3f} ms = {n / e0.elapsed_time(e1) / 1e6:.1f} GB/s")

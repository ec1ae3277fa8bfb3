"""Development aid: where the time of one rt_render_device call on a 1/8 partition goes — CUDA events around the call on the caller's
stream (what bench.py times), the library's own ev0..ev1 (rt_stats.ms_render), host time of the call."""
import importlib, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
S = rtb.scenes
s, p = S.cfg3_scene(), S.cfg3_params()
part = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_partition(3 % part, part, 32)
stream = torch.cuda.Stream(); rt.set_stream(stream.cuda_stream)
rt.set_scene(s)
want = rtb.host.RT_OUT_RGBA32F
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(4): rt.render_device(p, want)
torch.cuda.synchronize()
res = []
for i in range(12):
    with torch.cuda.stream(stream): flush.zero_()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record(stream); rt.render_device(p, want, sync_stats=False); b.record(stream)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    st = rt.stats()
    res.append((a.elapsed_time(b), st["ms_render"], 1e3 * (t1 - t0)))
r = np.array(res[2:])
print(f"part 1/{part}: events around the call {r[:,0].mean():.3f} ms, library ev0..ev1 {r[:,1].mean():.3f} ms, host time of the call {r[:,2].mean():.3f} ms")

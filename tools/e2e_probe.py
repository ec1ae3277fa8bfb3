"""Development aid: wall time of rt_render (host Color[] out, pinned) vs copy_waves, next to the device-only time."""
import importlib, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes; H = rtb.host
s, p = sc.cfg3_scene(), sc.cfg3_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_scene(s)
out = rt.alloc_host_outputs(p, H.RT_OUT_RGBA32F, pinned=True)
for cw in (1, 2, 4, 8, 16, 32):
    rt.set_option("copy_waves", cw)
    ts = []
    for i in range(8):
        t0 = time.perf_counter(); r = rt.render(p, H.RT_OUT_RGBA32F, out=out); ts.append(time.perf_counter() - t0)
    st = r["stats"]
    print(f"copy_waves={cw:2d} wall_ms={1e3*np.median(ts[2:]):.3f} device_ms={st['ms_render']:.3f} d2h_tail_ms={st['ms_d2h']:.3f} launches={st['kernel_launches']}", flush=True)

rt.set_option("copy_waves", 8)
for lanes in (1, 2):
    for waves in (1, 2, 3, 4, 6, 8):
        rt.set_option("lanes", lanes); rt.set_option("waves", waves)
        ms = []
        for i in range(6):
            st = rt.render_device(p, H.RT_OUT_RGBA32F, sync_stats=True); ms.append(st["ms_render"])
        print(f"render_device lanes={lanes} waves={waves} device_ms={np.median(ms[2:]):.3f} launches={st['kernel_launches']}", flush=True)

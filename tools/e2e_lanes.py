"""Development aid: wall time of rt_render (host Color[] out, pinned) vs lanes x copy_waves."""
import importlib, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes; H = rtb.host
s, p = sc.cfg3_scene(), sc.cfg3_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_scene(s)
out = rt.alloc_host_outputs(p, H.RT_OUT_RGBA32F, pinned=True)
for lanes in (2, 3, 4):
    for cw in (6, 8, 12, 16):
        rt.set_option("copy_waves", cw); rt.set_option("lanes", lanes)
        ts = []
        for i in range(10):
            t0 = time.perf_counter(); r = rt.render(p, H.RT_OUT_RGBA32F, out=out); ts.append(time.perf_counter() - t0)
        st = r["stats"]
        print(f"lanes={lanes} copy_waves={cw:2d} wall_ms={1e3*np.median(ts[3:]):.3f} device_ms={st['ms_render']:.3f} d2h_tail_ms={st['ms_d2h']:.3f}", flush=True)

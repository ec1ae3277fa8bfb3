"""Development aid: wall time of rt_render (host Color[] out, pinned) for different wave plans of the pipelined copy-out."""
import importlib, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes; H = rtb.host
cfg = sys.argv.pop(1)
s, p = (sc.cfg3_scene(), sc.cfg3_params()) if cfg == "cfg3" else (sc.rt_unity(True), sc.rt_unity_params(3840, 2160, 4))
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_scene(s)
out = rt.alloc_host_outputs(p, H.RT_OUT_RGBA32F, pinned=True)
for lanes in (2,):
    rt.set_option("lanes", lanes)
    for plan in [int(a) for a in sys.argv[1:]]:
        rt.set_option("copy_plan", plan)
        ts = []
        for i in range(12):
            t0 = time.perf_counter(); r = rt.render(p, H.RT_OUT_RGBA32F, out=out); ts.append(time.perf_counter() - t0)
        st = r["stats"]
        print(f"lanes={lanes} plan={plan} wall_ms={1e3*np.median(ts[4:]):.3f} device_ms={st['ms_render']:.3f}", flush=True)

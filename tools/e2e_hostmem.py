"""Development aid: rt_render wall time into a pageable, a registered (rt_host_register) and a pinned host Color[]."""
import importlib, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes; H = rtb.host
s, p = sc.cfg3_scene(), sc.cfg3_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_scene(s)
want = H.RT_OUT_RGBA32F
def run(out, label):
    ts = []
    for i in range(10):
        t0 = time.perf_counter(); r = rt.render(p, want, out=out); ts.append(time.perf_counter() - t0)
    print(f"{label}: wall_ms={1e3*np.median(ts[3:]):.3f} min={1e3*min(ts[3:]):.3f} device_ms={r['stats']['ms_render']:.3f}", flush=True)
pinned = rt.alloc_host_outputs(p, want, pinned=True); run(pinned, "pinned (cudaHostAlloc)")
page = rt.alloc_host_outputs(p, want); page["rgba32f"][:] = 1
run(page, "pageable")
t0 = time.perf_counter(); rt.host_register(page["rgba32f"]); print(f"rt_host_register: {1e3*(time.perf_counter()-t0):.2f} ms")
run(page, "registered")
print("same:", np.array_equal(page["rgba32f"], pinned["rgba32f"]))
rt.host_unregister(page["rgba32f"])
run(page, "pageable again")

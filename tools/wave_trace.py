"""Development aid: per-wave render/copy completion times of rt_render's pipelined copy-out (needs a -DRT_WAVE_TRACE build:
   RTB200_EXTRA_NVCC_FLAGS=-DRT_WAVE_TRACE RTB200_OUT=/root/repo/gpurun_in/librtb200_wt.so python unity-raytracer_b200/build.py --force)
usage: RTB200_WAVE_TRACE=1 python tools/wave_trace.py <lib.so> [copy_waves]"""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes; H = rtb.host
H.LIB_PATH = sys.argv[1]
s, p = sc.cfg3_scene(), sc.cfg3_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_scene(s)
if len(sys.argv) > 2: rt.set_option("copy_waves", int(sys.argv[2]))
out = rt.alloc_host_outputs(p, H.RT_OUT_RGBA32F, pinned=True)
os.environ.pop("RTB200_WAVE_TRACE", None)
for i in range(5): rt.render(p, H.RT_OUT_RGBA32F, out=out)
os.environ["RTB200_WAVE_TRACE"] = "1"
t0 = time.perf_counter(); r = rt.render(p, H.RT_OUT_RGBA32F, out=out); print(f"wall {1e3 * (time.perf_counter() - t0):.3f} ms, device {r['stats']['ms_render']:.3f} ms")

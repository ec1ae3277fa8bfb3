"""Development aid: per-tile start/duration of k_trace_primary (needs a -DRT_TILE_TIMING build of the library:
   RTB200_EXTRA_NVCC_FLAGS=-DRT_TILE_TIMING RTB200_OUT=/tmp/librtb200_dbg.so python unity-raytracer_b200/build.py --force)"""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
H = rtb.host
H.LIB_PATH = sys.argv[1]
part = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rank = int(sys.argv[3]) if len(sys.argv) > 3 else 0
sc = rtb.scenes
s, p = sc.cfg3_scene(), sc.cfg3_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_option("waves", 1)
if part > 1: rt.set_partition(rank, part, 32)
rt.set_scene(s)
L = H.load_library()
L.rt_debug_tile_times.argtypes = [C.c_int, C.c_void_p]
nt = 480 * 540
for i in range(3):
    rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)
L.rt_debug_tile_times(nt, None)
st = rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)
buf = np.zeros((nt, 2), np.uint64)
L.rt_debug_tile_times(nt, buf.ctypes.data)
v = buf[buf[:, 1] > 0]
start = (v[:, 0] - v[:, 0].min()).astype(np.float64) / 1e3; dur = v[:, 1].astype(np.float64) / 1e3
end = start + dur
print(f"part={part} rank={rank} tiles={len(v)} ms_render={st['ms_render']:.3f} kernel span={end.max():.1f} us")
print("duration us: mean %.1f  p50 %.1f  p90 %.1f  p99 %.1f  p99.9 %.1f  max %.1f" % (dur.mean(), *np.percentile(dur, [50, 90, 99, 99.9]), dur.max()))
print("sum of durations / (4736 warp slots) = %.1f us" % (dur.sum() / 4736))
order = np.argsort(end)[-10:]
for i in order: print(f"  late tile: start {start[i]:.1f} dur {dur[i]:.1f} end {end[i]:.1f}")
h, e = np.histogram(start, bins=10); print("tile starts per 10% of span:", h.tolist())

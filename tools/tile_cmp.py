"""Development aid: per-tile duration of k_trace_primary with packet vs per-lane traversal (needs a -DRT_TILE_TIMING build, see tile_times.py)."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
H = rtb.host
H.LIB_PATH = sys.argv[1]
sc = rtb.scenes
s, p = sc.cfg3_scene(), sc.cfg3_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_option("waves", 1)
rt.set_scene(s)
L = H.load_library()
L.rt_debug_tile_times.argtypes = [C.c_int, C.c_void_p]
nt = 480 * 540
res = {}
for packet in (1, 0):
    rt.set_option("packet", packet)
    for i in range(3):
        rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)
    L.rt_debug_tile_times(nt, None)
    st = rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)
    buf = np.zeros((nt, 2), np.uint64)
    L.rt_debug_tile_times(nt, buf.ctypes.data)
    res[packet] = buf[:, 1].astype(np.float64) / 1e3
    print(f"packet={packet} ms_render={st['ms_render']:.3f} sum={res[packet].sum()/1e3:.1f} ms max={res[packet].max():.1f} us")
a, b = res[1], res[0]
order = np.argsort(-a)
print("heaviest packet tiles: tile, packet us, lane us")
for i in order[:30]: print(f"  {i:7d} (x{i % 480:3d} y{i // 480:3d}) {a[i]:8.1f} {b[i]:8.1f}")
for lo, hi in [(0, 5), (5, 10), (10, 20), (20, 40), (40, 80), (80, 1e9)]:
    m = (a >= lo) & (a < hi)
    if m.any(): print(f"packet us in [{lo},{hi}): n={m.sum():7d} packet sum {a[m].sum()/1e3:8.2f} ms lane sum {b[m].sum()/1e3:8.2f} ms  lane max {b[m].max():.1f}")
np.savez_compressed("gpurun_out/tile_cmp.npz", packet=a.astype(np.float32), lane=b.astype(np.float32))

"""Development aid: where the frame time goes, as a coarse map of per-tile costs of both trace kernels (needs a -DRT_TILE_TIMING build:
   RTB200_EXTRA_NVCC_FLAGS=-DRT_TILE_TIMING RTB200_OUT=/root/repo/gpurun_in/librtb200_dbg.so python unity-raytracer_b200/build.py --force)
usage: python tools/tile_map.py <lib.so> [cfg]"""
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
tx, ty = 480, 540
nt = tx * ty
for i in range(3): rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)
L.rt_debug_tile_times(nt, None)
st = rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)
buf = np.zeros(3 * nt, np.uint64)
L.rt_debug_tile_times(nt, buf.ctypes.data)
prim = buf[:2 * nt].reshape(nt, 2)[:, 1].astype(np.float64).reshape(ty, tx) / 1e3      # us
shad = buf[2 * nt:].astype(np.float64).reshape(ty, tx) / 1965.0                        # cycles -> us at 1965 MHz
print(f"ms_render {st['ms_render']:.3f}; primary tile-us total {prim.sum():.0f}, shadow tile-us total {shad.sum():.0f}")
def coarse(a, by, bx): return a.reshape(ty // by, by, tx // bx, bx).sum(axis=(1, 3))
for name, a in (("primary", prim), ("shadow", shad)):
    c = coarse(a, 30, 30)
    print(name, "share of the kernel's tile time per 30x30-tile block (rows top->bottom), percent:")
    for r in c: print(" ".join(f"{100 * v / a.sum():4.1f}" for v in r))
    d = np.sort(a.ravel())[::-1]
    print(name, "heaviest 1%% of tiles hold %.1f%% of the time, heaviest 10%%: %.1f%%; max %.1f us, mean %.2f us" % (100 * d[:nt // 100].sum() / d.sum(), 100 * d[:nt // 10].sum() / d.sum(), d[0], d.mean()))
np.save("gpurun_out/tile_prim.npy", prim.astype(np.float32)); np.save("gpurun_out/tile_shad.npy", shad.astype(np.float32))

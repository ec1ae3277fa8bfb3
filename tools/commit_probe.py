"""Development aid: where the wall time of a scene (re-)commit goes, call by call (host.py's set_scene unrolled)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
H = rtb.host; sc = rtb.scenes
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
opts = dict(a.split("=") for a in sys.argv[2:])
s = {"cfg3": sc.cfg3_scene, "cfg4": sc.cfg4_scene, "cfg5": sc.cfg5_scene}[cfg]()
rt = rtb.RayTracer((0,))
for k, v in opts.items(): rt.set_option(k, float(v))
L, h = rt.L, rt.h
f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
for rep in range(3):
    T = {}
    def tick(name, t0): T[name] = T.get(name, 0.0) + (time.perf_counter() - t0)
    t00 = time.perf_counter()
    t0 = time.perf_counter(); rt._chk(L.rt_scene_begin(h)); tick("begin", t0)
    if s.spheres.shape[0]:
        t0 = time.perf_counter(); a, m = f32(s.spheres), np.ascontiguousarray(s.sphere_mats); tick("py", t0)
        t0 = time.perf_counter(); rt._chk(L.rt_scene_spheres(h, H._fptr(a), m.ctypes.data, a.shape[0])); tick("spheres", t0)
    if s.tris.shape[0]:
        t0 = time.perf_counter(); v, n, m = f32(s.tris), f32(s.tri_normals), np.ascontiguousarray(s.tri_mats); tick("py", t0)
        t0 = time.perf_counter(); rt._chk(L.rt_scene_triangles(h, H._fptr(v), H._fptr(n), m.ctypes.data, v.shape[0])); tick("triangles", t0)
    for (v, n, mat, aabb) in s.meshes:
        t0 = time.perf_counter(); v, n, aabb, mat = f32(v), f32(n), f32(aabb), np.ascontiguousarray(mat); tick("py", t0)
        t0 = time.perf_counter(); rt._chk(L.rt_scene_mesh(h, H._fptr(v), H._fptr(n), v.shape[0], mat.ctypes.data, H._fptr(aabb))); tick("mesh", t0)
    li, amb = f32(s.lights), f32(s.ambient)
    t0 = time.perf_counter(); rt._chk(L.rt_scene_lights(h, H._fptr(li), li.shape[0], H._fptr(amb))); tick("lights", t0)
    t0 = time.perf_counter(); rt._chk(L.rt_scene_commit(h, None)); tick("commit", t0)
    tot = time.perf_counter() - t00
    st = rt.stats()
    print(f"{cfg} rep {rep}: total {1e3*tot:.2f} ms ({s.n_prims/tot/1e6:.1f} Mprims/s) | " + ", ".join(f"{k} {1e3*v:.2f}" for k, v in T.items()) +
          f" | device: upload {st['ms_commit_upload']:.2f} build {st['ms_commit_build']:.2f}", flush=True)
# dirty-tracking path: only the spheres moved (cfg5) / only the lights moved (others), then rt_scene_recommit
for rep in range(3):
    t0 = time.perf_counter()
    if s.spheres.shape[0]:
        rt.update_spheres(0, s.spheres)
        what = f"update {s.spheres.shape[0]} spheres"
    else:
        rt.update_lights(s.lights, s.ambient)
        what = "update lights"
    t1 = time.perf_counter()
    rt.recommit()
    t2 = time.perf_counter()
    st = rt.stats()
    print(f"{cfg} {what} {1e3*(t1-t0):.2f} ms + recommit {1e3*(t2-t1):.2f} ms ({s.n_prims/(t2-t0)/1e6:.1f} Mprims/s rebuilt) | device: upload {st['ms_commit_upload']:.2f} build {st['ms_commit_build']:.2f}", flush=True)
# refit path: the whole mesh moved a little (all vertices re-sent), tree kept, boxes recomputed; then the frame time with the refitted tree
if s.meshes:
    p = {"cfg3": sc.cfg3_params, "cfg4": sc.cfg4_params, "cfg5": sc.cfg5_params}[cfg]()
    base = rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)["ms_render"]
    base = rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)["ms_render"]
    v0, n0, mat, aabb = s.meshes[0]
    for rep, amp in enumerate((0.0, 0.01, 0.01, 0.05)):
        rng = np.random.default_rng(rep)
        v = (v0 + amp * rng.standard_normal(v0.shape).astype(np.float32)).astype(np.float32) if amp else v0
        lo, hi = v.reshape(-1, 3).min(axis=0), v.reshape(-1, 3).max(axis=0)
        t0 = time.perf_counter(); rt.update_mesh(0, v, n0, np.concatenate([lo, hi])); t1 = time.perf_counter()
        rt.refit(); t2 = time.perf_counter()
        st = rt.stats()
        ms = [rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)["ms_render"] for _ in range(3)][-1]
        print(f"{cfg} jitter {amp}: update mesh 0 ({v.shape[0]} tris) {1e3*(t1-t0):.2f} ms + refit {1e3*(t2-t1):.2f} ms | device: build {st['ms_commit_build']:.3f} | frame {ms:.3f} ms (before: {base:.3f})", flush=True)

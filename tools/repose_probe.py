"""Development aid: wall time of re-posing every instance of cfg4-as-instances (rt_scene_update_instance x 125 + rt_scene_recommit)."""
import importlib, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
S = rtb.scenes
s, p = S.cfg4_instanced(), S.cfg4_params(1920, 1080)
r = rtb.RayTracer((0,)); r.set_option("builder", 1)
t0 = time.perf_counter(); r.set_scene(s); print(f"commit {1e3*(time.perf_counter()-t0):.2f} ms")
r.render_device(p, 2)
boxes = [S.bake_tris(s.base_meshes[m.base], m.m3x4)[2] for m in s.meshes]
for rep in range(6):
    torch.cuda.synchronize()                              # (the frame rendered at the end of the previous round is asynchronous)
    t0 = time.perf_counter()
    for i, m in enumerate(s.meshes):
        r.update_instance(i, m.m3x4, boxes[i])
    t1 = time.perf_counter()
    r.recommit()
    t2 = time.perf_counter()
    st = r.stats()
    print(f"rep {rep}: 125 updates {1e3*(t1-t0):.3f} ms, recommit {1e3*(t2-t1):.3f} ms (device: upload {st['ms_commit_upload']:.3f} + build {st['ms_commit_build']:.3f})", flush=True)
    r.render_device(p, 2)
idx, mm, bb = np.arange(len(s.meshes), dtype=np.int32), np.stack([m.m3x4 for m in s.meshes]), np.stack(boxes)
for rep in range(6):                                     # the same through ONE rt_scene_update_instances call
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r.update_instances(idx, mm, bb)
    t1 = time.perf_counter()
    r.recommit()
    t2 = time.perf_counter()
    print(f"batched rep {rep}: 1 call {1e3*(t1-t0):.3f} ms, recommit {1e3*(t2-t1):.3f} ms", flush=True)
    r.render_device(p, 2)
for rep in range(4):                                     # ... and with Mesh.AABB left to the library (device reduction over the resident base mesh)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r.update_instances(idx, mm, None)
    t1 = time.perf_counter()
    r.recommit()
    t2 = time.perf_counter()
    print(f"batched, AABBs computed, rep {rep}: 1 call {1e3*(t1-t0):.3f} ms, recommit {1e3*(t2-t1):.3f} ms", flush=True)
    r.render_device(p, 2)
t0 = time.perf_counter(); r.set_scene(s); print(f"full re-declaration (BLAS cached) {1e3*(time.perf_counter()-t0):.2f} ms")
# phases of the re-declaration (what host.set_scene does, call by call)
from ctypes import byref
L, h = r.L, r.h
f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
fp = rtb.host._fptr
for rep in range(3):
    torch.cuda.synchronize()
    t = [time.perf_counter()]
    L.rt_scene_begin(h); t.append(time.perf_counter())
    for bv in s.base_meshes:
        bv = f32(bv); L.rt_scene_base_mesh(h, fp(bv), bv.shape[0])
    t.append(time.perf_counter())
    for m in s.meshes:
        mm, mat = f32(m.m3x4), np.ascontiguousarray(m.mat)
        L.rt_scene_mesh_instance(h, int(m.base), fp(mm), mat.ctypes.data, None)
    t.append(time.perf_counter())
    li, amb = f32(s.lights), f32(s.ambient)
    L.rt_scene_lights(h, fp(li), li.shape[0], fp(amb)); t.append(time.perf_counter())
    L.rt_scene_commit(h, None); t.append(time.perf_counter())
    d = [1e3 * (b - a) for a, b in zip(t, t[1:])]
    print(f"re-declaration phases: begin {d[0]:.3f}, base mesh {d[1]:.3f}, 125 instances {d[2]:.3f}, lights {d[3]:.3f}, commit {d[4]:.3f} ms", flush=True)
    r.render_device(p, 2)

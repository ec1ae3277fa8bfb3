"""Development aid: the reference's frame loop is UpdateScene() then CastPixelRays() (RayTracingSetup.cs:171-199).  With one
context the two run back to back; with two contexts on one GPU a caller that accepts one frame of latency can commit scene
k+1 (one host thread) while frame k renders (another).  Prints ms per frame for both, and checks the frames are the same."""
import importlib, os, sys, threading, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
H = rtb.host; sc = rtb.scenes
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 12
s = {"cfg3": sc.cfg3_scene, "cfg4": sc.cfg4_scene, "cfg5": sc.cfg5_scene}[cfg]()
p = {"cfg3": sc.cfg3_params, "cfg4": sc.cfg4_params, "cfg5": sc.cfg5_params}[cfg]()
want = H.RT_OUT_RGBA32F
opts = dict(a.split("=") for a in sys.argv[3:])
pinned = int(opts.pop("pinned", 0))
refit = int(opts.pop("refit", 0))                             # 1: per-frame update = rt_scene_update_mesh of every mesh + rt_scene_refit (tree kept)
if pinned:                                                    # caller arrays page-locked (a shim that keeps NativeArrays pinned)
    import torch
    def pin(a):
        tt = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        keep.append(tt)
        return tt.numpy()
    keep = []
    s.meshes = [(pin(v), pin(n), m, ab) for (v, n, m, ab) in s.meshes]
ctx = [rtb.RayTracer((0,)), rtb.RayTracer((0,))]
for c in ctx:
    for k, v in opts.items(): c.set_option(k, float(v))
outs = [rtb.RayTracer.alloc_host_outputs(p, want, pinned=True) for _ in range(2)]
def commit(c):
    if not refit:
        c.set_scene(s)
        return
    for i, (v, n, m, ab) in enumerate(s.meshes):
        c.update_mesh(i, v, n, ab)
    c.refit()
for c, o in zip(ctx, outs):
    c.set_scene(s); c.render(p, want, out=o); c.render(p, want, out=o)
ref = outs[0]["rgba32f"].copy()
# (a) one context, back to back
t0 = time.perf_counter()
for k in range(frames):
    commit(ctx[0]); ctx[0].render(p, want, out=outs[0])
seq = (time.perf_counter() - t0) / frames
assert np.array_equal(ref, outs[0]["rgba32f"], equal_nan=True)
t0 = time.perf_counter()
for k in range(frames): commit(ctx[0])
only_commit = (time.perf_counter() - t0) / frames
t0 = time.perf_counter()
for k in range(frames): ctx[0].render(p, want, out=outs[0])
only_render = (time.perf_counter() - t0) / frames
print(f"{cfg}: commit alone {1e3*only_commit:.2f} ms, render alone {1e3*only_render:.2f} ms", flush=True)
# (b) two contexts: commit of scene k+1 on a second thread while frame k renders
ready = [threading.Semaphore(0), threading.Semaphore(0)]      # scene committed into ctx[i]
free = [threading.Semaphore(1), threading.Semaphore(1)]       # ctx[i] rendered, may take the next scene
def committer():
    for k in range(frames):
        i = k & 1
        free[i].acquire(); commit(ctx[i]); ready[i].release()
th = threading.Thread(target=committer)
t0 = time.perf_counter(); th.start()
ok = True
for k in range(frames):
    i = k & 1
    ready[i].acquire(); ctx[i].render(p, want, out=outs[i]); free[i].release()
th.join()
pipe = (time.perf_counter() - t0) / frames
ok = np.array_equal(ref, outs[0]["rgba32f"], equal_nan=True) and np.array_equal(ref, outs[1]["rgba32f"], equal_nan=True)
print(f"{cfg}: one context commit+render {1e3*seq:.2f} ms/frame | two contexts pipelined {1e3*pipe:.2f} ms/frame | frames identical: {ok}", flush=True)
for c in ctx: c.close()

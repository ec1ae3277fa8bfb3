"""Quick on-GPU probe: commit + render timings for a config (not the bench contract; a development aid)."""
import importlib, sys, time, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
opts = dict(a.split("=") for a in sys.argv[3:])
timing = int(opts.pop("timing", 1))
flush = int(opts.pop("flush", 0))
part = int(opts.pop("part", 1))
host_time = int(opts.pop("host", 0))
recommit = int(opts.pop("recommit", 0))
if "lib" in opts: rtb.host.LIB_PATH = opts.pop("lib")       # alternative build of the library (development)
t0 = time.time()
if cfg == "cfg2": s, p = sc.rt_unity(True), sc.rt_unity_params(3840, 2160, 4)
elif cfg == "cfg3": s, p = sc.cfg3_scene(), sc.cfg3_params()
elif cfg == "cfg4": s, p = sc.cfg4_scene(), sc.cfg4_params()
elif cfg == "cfg4i": s, p = sc.cfg4_instanced(), sc.cfg4_params()            # one base mesh + 125 instances
elif cfg == "cfg4ib": s, p = sc.cfg4_instanced().baked(), sc.cfg4_params()    # its baked twin (bit-identical frame)
elif cfg == "cfg5": s, p = sc.cfg5_scene(), sc.cfg5_params()
else: s, p = sc.cfg3_scene(n=int(cfg)), sc.cfg3_params()
print(f"scene {cfg}: {s.n_prims} prims, generated in {time.time()-t0:.1f}s", flush=True)
rt = rtb.RayTracer((0,))
if part > 1: rt.set_partition(int(opts.pop('rank', 0)), part, 32)
for k, v in opts.items(): rt.set_option(k, float(v))
rt.set_option("kernel_timing", timing)
t0 = time.time(); rt.set_scene(s); print(f"commit wall {time.time()-t0:.3f}s", flush=True)
for i in range(recommit):                                  # steady-state re-commit (buffers already allocated), as the reference's per-frame UpdateScene
    t0 = time.perf_counter(); rt.set_scene(s); w = time.perf_counter() - t0
    st = rt.stats()
    print(f"recommit {i}: wall {1e3*w:.2f} ms, upload {st['ms_commit_upload']:.2f} ms, build {st['ms_commit_build']:.2f} ms, {s.n_prims/w/1e6:.1f} Mprims/s", flush=True)
fl = None
if flush:
    import torch
    fl = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for i in range(reps):
    if fl is not None:
        fl.zero_(); torch.cuda.synchronize()
    if host_time:                                            # host enqueue time of one frame (no sync inside the call)
        import torch
        torch.cuda.synchronize(); h0 = time.perf_counter()
        rt.render_device(p, rtb.host.RT_OUT_RGBA8, sync_stats=False)
        h1 = time.perf_counter(); torch.cuda.synchronize(); h2 = time.perf_counter()
        print(f"host enqueue {1e3*(h1-h0):.3f} ms, enqueue+drain {1e3*(h2-h0):.3f} ms", flush=True)
    st = rt.render_device(p, rtb.host.RT_OUT_RGBA8, sync_stats=True)
    rays = st["rays_primary"] + st["rays_shadow"] + st["rays_mirror"]
    print(json.dumps(dict(ms=round(st["ms_render"], 3), mrays=round(rays / st["ms_render"] / 1e3, 1), rays=rays, launches=st["kernel_launches"],
                          nodes=st["n_nodes"], commit_up=round(st["ms_commit_upload"], 2), commit_build=round(st["ms_commit_build"], 2), k={k: round(v, 3) for k, v in st["ms_kernel"].items() if v > 0})), flush=True)

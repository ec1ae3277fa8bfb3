"""Work donation in the primary kernel — an experiment that is NOT in the tree: apply profiles/r2_donate_experiment.patch first (it adds the
option `donate`; write-up in profiles/r2_experiments.md).  Frames with and without it must be bit-identical; prints the kernel times of
both on the full frame and on band partitions.  Development aid.
    python tools/donate_probe.py [cfg3|cfg2|cfg5] [lib=path] [worlds=1,8] [reps=4]"""
import importlib, sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes
cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
opts = dict(a.split("=") for a in sys.argv[2:])
if "lib" in opts: rtb.host.LIB_PATH = opts.pop("lib")
worlds = [int(x) for x in opts.pop("worlds", "1,8").split(",")]
reps = int(opts.pop("reps", 4))
ranks = [int(x) for x in opts.pop("ranks", "").split(",") if x]
if cfg == "cfg2": s, p = sc.rt_unity(True), sc.rt_unity_params(3840, 2160, 4)
elif cfg == "cfg5": s, p = sc.cfg5_scene(), sc.cfg5_params()
elif cfg == "cfg1": s, p = sc.rt_unity(False), sc.rt_unity_params(1280, 720, 0)
else: s, p = sc.cfg3_scene(), sc.cfg3_params()
WANT = rtb.host.RT_OUT_ALL
ok = True
for world in worlds:
    for rank in ((ranks or range(world)) if world > 1 else [0]):
        frames = {}
        for don in (0, 1):
            rt = rtb.RayTracer((0,))
            if world > 1: rt.set_partition(rank, world, 32)
            rt.set_option("donate", don)
            rt.set_option("kernel_timing", int(os.environ.get("DON_TIMING", "1")))
            for k, v in opts.items(): rt.set_option(k, float(v))
            rt.set_scene(s)
            best = None
            for i in range(reps):
                st = rt.render_device(p, WANT, sync_stats=True)
                k = st["ms_kernel"]
                if int(os.environ.get("DON_VERBOSE", "0")): print(f"      frame {i}: {st['ms_render']:.3f} ms {json.dumps({a: round(b, 3) for a, b in k.items() if b > 0})}", flush=True)
                if best is None or st["ms_render"] < best[0]: best = (st["ms_render"], {a: round(b, 3) for a, b in k.items() if b > 0})
            frames[don] = rt.read_outputs(p, WANT)
            print(f"world {world} rank {rank} donate {don}: {best[0]:.3f} ms  {json.dumps(best[1])}", flush=True)
            rt.close()
        same = all(np.array_equal(frames[0][k].view(np.uint8), frames[1][k].view(np.uint8)) for k in ("rgba32f", "rgba8", "hit_id", "hit_t"))
        if not same:
            ok = False
            for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
                a, b = frames[0][k].view(np.uint8).ravel(), frames[1][k].view(np.uint8).ravel()
                print(f"   DIFFERENT {k}: {int((a != b).sum())} bytes", flush=True)
        else: print("   identical", flush=True)
print("ALL IDENTICAL" if ok else "MISMATCH")

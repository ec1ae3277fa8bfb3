"""Development aid: per-ray node visits / primitive tests of the traversal kernels (n_node, n_prim for the L2 and
issue rooflines).  Needs the instrumented build of the library:
   RTB200_EXTRA_NVCC_FLAGS=-DRT_COUNT_VISITS RTB200_OUT=/root/repo/gpurun_in_cnt.so python unity-raytracer_b200/build.py --force
usage: python tools/visit_counts.py <lib.so> <workload> [out.json]     (run on the GPU box; the counts are deterministic per BVH)"""
import ctypes as C, importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtb = importlib.import_module("unity-raytracer_b200")
H = rtb.host
H.LIB_PATH = sys.argv[1]
wl = sys.argv[2] if len(sys.argv) > 2 else "cfg3"
sc = rtb.scenes
if wl == "cfg2": s, p = sc.rt_unity(True), sc.rt_unity_params(3840, 2160, 4)
elif wl == "cfg3": s, p = sc.cfg3_scene(), sc.cfg3_params()
elif wl == "cfg4": s, p = sc.cfg4_scene(), sc.cfg4_params()
elif wl == "cfg4i": s, p = sc.cfg4_instanced(), sc.cfg4_params()
elif wl == "cfg4ib": s, p = sc.cfg4_instanced().baked(), sc.cfg4_params()
else: s, p = sc.cfg5_scene(), sc.cfg5_params()
L = H.load_library()
L.rt_debug_visit_counts.argtypes = [C.c_void_p, C.c_int]
names = ["lane_rays", "lane_node_visits", "lane_prim_tests", "packet_warps", "packet_rays", "packet_node_fetches", "packet_lane_node_visits",
         "packet_prim_fetches", "packet_lane_prim_tests", "packet_pops_culled"]
out = {"workload": wl}
builders = [int(b) for b in os.environ.get("VC_BUILDERS", "1,0").split(",")]
for builder in builders:
    for packet in (1, 0):
        rt = rtb.RayTracer((0,)); rt.set_option("builder", builder); rt.set_option("packet", packet); rt.set_option("hints", 0)
        rt.set_scene(s)
        L.rt_debug_visit_counts(None, 1)
        st = rt.render_device(p, H.RT_OUT_RGBA8, sync_stats=True)
        c = np.zeros(20, np.uint64)
        L.rt_debug_visit_counts(c.ctypes.data, 1)
        rec = {"n_nodes": st["n_nodes"], "n_prims": st["n_prims"], "rays_primary": st["rays_primary"], "rays_shadow": st["rays_shadow"], "rays_mirror": st["rays_mirror"]}
        for cls, base in (("closest", 0), ("any", 10)):
            d = {n: int(c[base + i]) for i, n in enumerate(names) if n != "_"}
            rays = d["lane_rays"] + d["packet_rays"]
            if rays:
                # per ray: what one ray's own traversal does (per-lane runs) / what its lane executes inside a packet
                d["node_visits_per_ray"] = (d["lane_node_visits"] + d["packet_lane_node_visits"]) / rays
                d["prim_tests_per_ray"] = (d["lane_prim_tests"] + d["packet_lane_prim_tests"]) / rays
                # fetches: a per-lane visit is its own 64-B node / 48-B primitive request, a packet visit is ONE request per warp
                d["node_fetches_per_ray"] = (d["lane_node_visits"] + d["packet_node_fetches"]) / rays
                d["prim_fetches_per_ray"] = (d["lane_prim_tests"] + d["packet_prim_fetches"]) / rays
            rec[cls] = d
        out[f"builder={'sah' if builder else 'lbvh'},packet={packet}"] = rec
        print(f"builder={builder} packet={packet}:", {k: {m: round(v, 2) for m, v in rec[k].items() if m.endswith('per_ray')} for k in ("closest", "any")}, flush=True)
        rt.close()
dst = sys.argv[3] if len(sys.argv) > 3 else "gpurun_out/visit_counts_%s.json" % wl
with open(dst, "w") as f:
    json.dump(out, f, indent=1)

import importlib, sys, os, json
sys.path.insert(0, "/root/repo")
rtb = importlib.import_module("unity-raytracer_b200")
sc = rtb.scenes
s, p = sc.cfg5_scene(), sc.cfg5_params()
rt = rtb.RayTracer((0,)); rt.set_option("builder", 1); rt.set_option("kernel_timing", 1)
rt.set_scene(s)
for i in range(2):
    st = rt.render_device(p, rtb.host.RT_OUT_RGBA8, sync_stats=True)
print({k: st[k] for k in ("rays_primary", "rays_shadow", "rays_mirror", "ms_render", "kernel_launches")}, {k: round(v, 2) for k, v in st["ms_kernel"].items() if v > 0})

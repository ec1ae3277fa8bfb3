"""-m gpu: the CUDA path, called through the C-ABI (ctypes -> librtb200.so), against the CPU oracle on the same inputs."""
import numpy as np
import pytest

from conftest import compare_renders

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rt(rtb):
    r = rtb.RayTracer((0,))
    yield r
    r.close()


def _check(cmp, exact_float=True):
    assert cmp["id"] == 0 and cmp["t"] == 0, cmp
    assert cmp["u8_gt1"] == 0, cmp
    assert cmp["u8_gt0"] <= max(1, cmp["n"] // 1000), cmp          # >= 99.9 % of pixels identical after quantisation
    if exact_float:
        # float colour may differ only through pow() (CUDA's double pow vs glibc's): such a pixel is off by last bits of its sum,
        # never more (f32_far, conftest.float_far), and only a specular pixel can be one
        assert cmp["f32_far"] == 0, cmp
        assert cmp["f32"] <= max(2, cmp["n"] // 200), cmp


def test_kat_sphere_reference_test(rt, scenes):
    """Tests/RayTracerTests.cs:8-27: ray (0,0,0)->(1,0,0) vs sphere C=(300,0,0) r2=1 hits at t = 299."""
    s = scenes.SceneArrays()
    s.add_sphere((300, 0, 0), 1.0, scenes.material())
    rt.set_scene(s)
    ids, t = rt.intersect([(0, 0, 0)], [(1, 0, 0)])
    assert ids[0].tolist() == [1, 0, -1] and t[0] == np.float32(299.0)


def test_cfg1_rt_unity(rt, scenes, oracle):
    s, p = scenes.rt_unity(), scenes.rt_unity_params()
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    rt.set_scene(s)
    got = rt.render(p)
    _check(compare_renders(ref, got))
    st = got["stats"]
    assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == ref["stats"]
    # KAT-2: centre ray hits the sphere at 19.6 and shades to (244.60649, 15, 15)
    ids, t = rt.intersect([(0, 0, 0)], [(0, 0, 1)])
    assert ids[0].tolist() == [1, 0, -1] and t[0] == np.float32(19.6)


def test_cfg2_mirror_depth4(rt, scenes, oracle):
    s, p = scenes.rt_unity(mirror_sphere=True), scenes.rt_unity_params(480, 270, 4)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    rt.set_scene(s)
    got = rt.render(p)
    _check(compare_renders(ref, got))
    assert got["stats"]["rays_mirror"] == ref["stats"][2] > 0
    assert got["stats"]["rays_shadow"] == ref["stats"][1]


@pytest.mark.parametrize("seed,spp", [(7, 1), (3, 2)])
def test_random_soup_vs_faithful_scan(rt, scenes, oracle, seed, spp):
    s, p = scenes.random_soup(seed=seed), scenes.soup_params(96, 64, 3, spp)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    rt.set_scene(s)
    got = rt.render(p)
    _check(compare_renders(ref, got))
    st = got["stats"]
    assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == ref["stats"]


def test_many_lights_and_unfused_shading(rtb, scenes, oracle):
    """The light loop (RayTracingSetup.cs:333-355) beyond the fused kernel's 32-light limit: 40 lights go through the separate
    occlusion-mask + shading kernels, checked against the faithful scan; with 2 lights the unfused and the fused kernels,
    and backcull off, must give the fused frame bit for bit."""
    from conftest import many_lights_scene
    s, p = many_lights_scene(scenes), scenes.soup_params(80, 48, 2, 1)
    two = scenes.random_soup(seed=13, n_tris=600, n_spheres=60)
    assert s.lights.shape[0] == 40
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    r = rtb.RayTracer((0,)); r.set_scene(s)
    got = r.render(p)
    _check(compare_renders(ref, got))
    st = got["stats"]
    assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == ref["stats"]
    r.set_scene(two)
    fused = r.render(p)
    for key in ("fuse_shade", "backcull"):
        q = rtb.RayTracer((0,)); q.set_option(key, 0); q.set_scene(two)
        other = q.render(p); q.close()
        for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
            assert np.array_equal(fused[k], other[k]), (key, k)
    r.close()


def test_reference_driver_interface_and_device_resident_api(rtb, scenes, oracle):
    """The reference's own call order (RayTracingSetup.Update, RayTracingSetup.cs:171-199): set ImagePlane / BackgroundColor /
    MaxReflectionBounces, UpdateScene(), CastPixelRays(CameraData) -> PixelColors (Color[] index x + y*X).  Then the
    device-resident calls under it: rt_render_device + rt_read_outputs, the device pointers of rt_device_outputs (read back
    through torch, no library copy), and a caller-owned stream (rt_set_stream) give the same frame."""
    import torch
    s, p = scenes.rt_unity(mirror_sphere=True), scenes.rt_unity_params(200, 112, 3)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    setup = rtb.RayTracingSetup((0,))
    setup.ImagePlane = rtb.ImagePlane(rtb.Resolution(200, 112), p["dist"], p["half_h"], p["half_v"])
    setup.BackgroundColor = tuple(p["background_rgb"])
    setup.MaxReflectionBounces = 3
    setup.UpdateScene(s)
    cam = rtb.CameraData(p["pos"], p["fwd"], p["right"], p["up"])
    colors = setup.CastPixelRays(cam)
    assert colors is setup.PixelColors and colors.shape == (200 * 112, 4)
    assert np.array_equal(colors, ref["rgba32f"], equal_nan=True)
    r = setup._rt
    st = r.render_device(p, rtb.host.RT_OUT_ALL, sync_stats=True)
    assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == ref["stats"]
    back = r.read_outputs(p)
    _check(compare_renders(ref, back))
    ptrs = r.device_outputs()
    assert all(ptrs)
    n = 200 * 112

    class _Dev:                                             # __cuda_array_interface__ view of a library-owned device buffer
        def __init__(self, ptr, shape, typestr):
            self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}
    f32 = torch.as_tensor(_Dev(ptrs[0], (n, 4), "<f4"), device="cuda:0").cpu().numpy()
    ids = torch.as_tensor(_Dev(ptrs[2], (n, 3), "<i4"), device="cuda:0").cpu().numpy()
    assert np.array_equal(f32, ref["rgba32f"], equal_nan=True) and np.array_equal(ids, back["hit_id"])
    stream = torch.cuda.Stream(device=0)
    r.set_stream(stream.cuda_stream)
    r.render_device(scenes.rt_unity_params(200, 112, 0), rtb.host.RT_OUT_ALL)      # a different frame in between
    r.render_device(p, rtb.host.RT_OUT_ALL)
    stream.synchronize()
    again = r.read_outputs(p)
    for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
        assert np.array_equal(back[k], again[k], equal_nan=True), k
    r.set_stream(None)
    r.close()


def test_error_codes_leave_the_context_usable(rtb, scenes):
    """Every misuse returns its documented code (include/rtb200.h RT_E_*) with a message in rt_last_error, launches nothing,
    and the context renders the next valid frame unchanged."""
    E = rtb.host.RtError
    INVALID, STATE = -1, -4                                            # include/rtb200.h:31-35
    r = rtb.RayTracer((0,))
    s, p = scenes.rt_unity(), scenes.rt_unity_params(64, 40, 2)

    def fails(code, text, fn, *a):
        with pytest.raises(E) as ei:
            fn(*a)
        assert ei.value.code == code and text in str(ei.value), (ei.value.code, str(ei.value))

    fails(STATE, "before rt_scene_commit", r.render, p)
    fails(STATE, "before rt_scene_commit", r.intersect, [(0, 0, 0)], [(0, 0, 1)])
    fails(STATE, "no committed scene", r.recommit)
    fails(INVALID, "unknown option", r.set_option, "no_such_option", 1)
    fails(INVALID, "lanes must be 1..8", r.set_option, "lanes", 9)
    fails(INVALID, "leaf_max must be 1..8", r.set_option, "leaf_max", 0)
    fails(INVALID, "bad arguments", r.set_partition, 2, 2)
    r.set_scene(s)
    want = r.render(p)
    fails(INVALID, "negative resolution", r.render_device, dict(p, res_x=-1))
    fails(INVALID, "spp_grid above 64", r.render, dict(p, spp_grid=65))
    fails(INVALID, "bad range", r.update_spheres, s.spheres.shape[0], s.spheres[:1])
    fails(INVALID, "bad arguments", r.update_mesh, 7, s.meshes[0][0], s.meshes[0][1], s.meshes[0][3])
    r.render_device(p, rtb.host.RT_OUT_RGBA8)
    fails(STATE, "was not rendered", r.read_outputs, p, rtb.host.RT_OUT_HIT_T)
    got = r.render(p)                                                  # failed updates left the scene clean and unchanged
    for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
        assert np.array_equal(want[k], got[k]), k
    r.close()


def test_degenerate_geometry_matches_the_scan(rtb, scenes, oracle):
    """conftest.degenerate_scene (zero-area / repeated / huge / NaN / infinite primitives, camera inside a sphere): device LBVH
    and host SAH builder, packet and per-lane traversal, against the faithful scan."""
    from conftest import degenerate_scene
    for nan in (False, True):
        s, p = degenerate_scene(scenes, nan), scenes.soup_params(96, 64, 2, 1)
        ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
        for builder, packet in ((0, 1), (0, 0), (1, 1)):
            r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_option("packet", packet); r.set_scene(s)
            got = r.render(p); r.close()
            _check(compare_renders(ref, got))
            assert (got["stats"]["rays_primary"], got["stats"]["rays_shadow"], got["stats"]["rays_mirror"]) == ref["stats"], (nan, builder, packet)


def test_cameras_off_axis(rt, scenes, oracle):
    """conftest.camera_set (every view axis, an oblique rotation, a sheared non-unit basis): a soup with mirrors against the
    faithful scan, and the reduced config-3 mesh (packet traversal in every octant) against the accelerated oracle."""
    from conftest import camera_set
    s = scenes.random_soup(seed=4, n_tris=500, n_spheres=60)
    ora = oracle.OracleScene(s)
    rt.set_scene(s)
    for i, cam in enumerate(camera_set()):
        p = scenes.make_params(112, 80, 10.0, 8.0, 6.0, (0.1, 0.2, 0.3), 2, 1, cam=cam)
        ref = ora.render(p, oracle.FAITHFUL)
        got = rt.render(p)
        _check(compare_renders(ref, got))
        assert (got["stats"]["rays_primary"], got["stats"]["rays_shadow"], got["stats"]["rays_mirror"]) == ref["stats"], i
    s = scenes.cfg3_scene(n=48)
    ora = oracle.OracleScene(s)
    rt.set_scene(s)
    hits = 0
    for i, cam in enumerate(camera_set()):
        p = scenes.make_params(320, 180, 10.0, 8.0, 4.5, (0.1, 0.1, 0.2), 0, 1, cam=cam)
        ref = ora.render(p, oracle.ACCELERATED)
        got = rt.render(p)
        _check(compare_renders(ref, got))
        hits += int((got["hit_id"][:, 0] != 0).sum())
    assert hits > 7 * 320 * 180 // 20                       # the views do look at the mesh


def test_deep_mirror_recursion(rtb, scenes, oracle):
    """Two parallel mirror walls around the camera: the reflection chain (RayTracingSetup.cs:357-372) runs to the
    bounce limit on most pixels.  Depths 30 and 64 against the faithful scan; 65 is refused (a clamp would change the
    picture) unless the scene has no mirror material, where the depth cannot matter."""
    from conftest import mirror_walls_scene
    s = mirror_walls_scene(scenes)
    ora = oracle.OracleScene(s)
    r = rtb.RayTracer((0,)); r.set_scene(s)
    for depth in (30, 64):
        p = scenes.make_params(96, 64, 10.0, 8.0, 5.0, (0.05, 0.05, 0.1), depth, 1)
        ref = ora.render(p, oracle.FAITHFUL)
        got = r.render(p)
        _check(compare_renders(ref, got))
        assert (got["stats"]["rays_primary"], got["stats"]["rays_shadow"], got["stats"]["rays_mirror"]) == ref["stats"]
        assert got["stats"]["rays_mirror"] > depth * 96 * 64 // 2            # the chains really are that long
    with pytest.raises(rtb.host.RtError) as ei:
        r.render(dict(p, max_reflection_bounces=65))
    assert ei.value.code == -5 and "above 64" in str(ei.value)
    plain = scenes.rt_unity()                                                # no mirror material: any depth is the same frame
    r.set_scene(plain)
    a = r.render(scenes.rt_unity_params(64, 40, 1000)); b = r.render(scenes.rt_unity_params(64, 40, 0))
    assert np.array_equal(a["rgba32f"], b["rgba32f"])
    r.close()


def test_two_contexts_commit_and_render_concurrently(rtb, scenes, oracle):
    """Contexts share no state: one host thread re-commits scenes into context B (alternating two different scenes) while
    another renders context A and vice versa — the pipelined frame loop of bench.py's `frame_loop`.  Every frame must equal
    the frame of its scene rendered alone."""
    import threading
    sa, sb = scenes.cfg3_scene(n=40), scenes.random_soup(seed=17, n_tris=3000, n_spheres=300)
    pa, pb = scenes.cfg3_params(480, 270), scenes.soup_params(320, 200, 3, 1)
    solo = rtb.RayTracer((0,))
    want = {}
    for key, (s, p) in (("a", (sa, pa)), ("b", (sb, pb))):
        solo.set_scene(s); want[key] = solo.render(p)
    solo.close()
    _check(compare_renders(oracle.OracleScene(sb).render(pb, oracle.FAITHFUL), want["b"]))
    ctx = [rtb.RayTracer((0,)), rtb.RayTracer((0,))]
    frames = 12
    ready = [threading.Semaphore(0), threading.Semaphore(0)]
    free = [threading.Semaphore(1), threading.Semaphore(1)]
    plan = [("a", sa, pa) if (k // 2) % 2 == 0 else ("b", sb, pb) for k in range(frames)]
    errors = []

    def committer():
        try:
            for k in range(frames):
                free[k & 1].acquire(); ctx[k & 1].set_scene(plan[k][1]); ready[k & 1].release()
        except Exception as e:                                      # noqa: BLE001
            errors.append(e)
            for sem in ready:
                sem.release()
    th = threading.Thread(target=committer); th.start()
    for k in range(frames):
        ready[k & 1].acquire()
        if errors:
            break
        got = ctx[k & 1].render(plan[k][2])
        free[k & 1].release()
        for key in ("hit_id", "hit_t", "rgba8", "rgba32f"):
            assert np.array_equal(want[plan[k][0]][key], got[key], equal_nan=True), (k, key)
    th.join()
    assert not errors, errors
    for c in ctx:
        c.close()


def test_pipelined_driver_returns_the_previous_scene(rtb, scenes):
    """PipelinedRayTracingSetup.Update(scene k, cam): frame of scene k-1 (scene 0 on the first call) while scene k is being
    committed on the worker thread; Flush() shows the last one.  Each frame equals that scene rendered alone."""
    seq = [scenes.rt_unity(mirror_sphere=True), scenes.random_soup(seed=17, n_tris=3000, n_spheres=300), scenes.cfg3_scene(n=32),
           scenes.rt_unity(), scenes.random_soup(seed=18, n_tris=500, n_spheres=50)]
    p = scenes.make_params(240, 136, 10.0, 16.0, 9.0, (0.1, 0.1, 0.2), 3, 1)
    cam = rtb.CameraData(p["pos"], p["fwd"], p["right"], p["up"])

    def configure(d):
        d.ImagePlane = rtb.ImagePlane(rtb.Resolution(240, 136), p["dist"], p["half_h"], p["half_v"])
        d.BackgroundColor, d.MaxReflectionBounces = tuple(p["background_rgb"]), 3
    solo = rtb.RayTracingSetup((0,)); configure(solo)
    want = []
    for s in seq:
        solo.UpdateScene(s); want.append(solo.CastPixelRays(cam).copy())
    solo.close()
    assert not np.array_equal(want[0], want[1])
    pipe = rtb.PipelinedRayTracingSetup((0,)); configure(pipe)
    for rep in range(2):                                     # second round: both contexts re-committed with other scenes
        for k, s in enumerate(seq):
            got = pipe.Update(s, cam)
            shown = 0 if (rep == 0 and k == 0) else (k - 1) % len(seq)
            assert np.array_equal(got, want[shown], equal_nan=True), (rep, k)
    assert np.array_equal(pipe.Flush(cam), want[-1], equal_nan=True)
    pipe.close()


def test_intersect_random_rays(rt, scenes, oracle):
    s = scenes.random_soup(seed=11)
    rng = np.random.default_rng(5)
    o = rng.uniform(-40, 40, (20000, 3)).astype(np.float32)
    d = rng.normal(size=(20000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    rid, rt_ = oracle.OracleScene(s).intersect(o, d, oracle.FAITHFUL)
    rt.set_scene(s)
    gid, gt = rt.intersect(o, d)
    assert (rid[:, :2] == gid[:, :2]).all() and (rt_ == gt).all()
    mesh = rid[:, 0] == 3
    assert (rid[mesh, 2] == gid[mesh, 2]).all()


def test_cfg3_reduced(rt, scenes, oracle):
    """~130k-triangle displaced sphere, 640x360: full frame vs the accelerated oracle, pixel subset vs the faithful scan."""
    s, p = scenes.cfg3_scene(n=80), scenes.cfg3_params(640, 360)
    ora = oracle.OracleScene(s)
    ref = ora.render(p, oracle.ACCELERATED)
    rt.set_scene(s)
    got = rt.render(p)
    _check(compare_renders(ref, got))
    px = np.random.default_rng(1).choice(640 * 360, 1024, replace=False)
    sub = ora.render(p, oracle.FAITHFUL, pixels=px)
    gsub = {k: got[k][px] for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}
    _check(compare_renders(sub, gsub))


@pytest.mark.parametrize("phong", [20.0, 1.0, 64.0, 3.0, 0.0, 20.5, 100.0])
def test_phong_exponent_paths(rt, scenes, oracle, phong):
    """Blinn-Phong pow (RayTracingSetup.cs:399): whole exponents <= 64 take the device's exponentiation-by-squaring path
    (rt_hd.h um_pow), everything else the library pow; both must give the oracle's float colours."""
    s, p = scenes.cfg3_scene(n=40, phong=phong), scenes.cfg3_params(480, 270)
    ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED)
    rt.set_scene(s)
    cmp = compare_renders(ref, rt.render(p))
    _check(cmp)
    assert cmp["f32"] <= 2, cmp


def test_edge_cases(rt, scenes):
    # empty scene: everything misses, background colour, FLT_MAX distance (Scene.cs:45-61)
    s = scenes.SceneArrays()
    rt.set_scene(s)
    p = scenes.make_params(33, 17, 10.0, 8.0, 5.0, (0.25, 0.5, 1.0))
    got = rt.render(p)
    assert (got["hit_id"] == np.array([0, -1, -1])).all()
    assert (got["hit_t"] == np.finfo(np.float32).max).all()
    assert (got["rgba8"] == np.array([64, 128, 255, 255], np.uint8)).all()
    # 0x0 image plane (Assets/Scenes/SampleScene.unity ships one): a no-op
    got = rt.render(scenes.make_params(0, 0, 10.0, 8.0, 5.0))
    assert got["rgba32f"].shape == (0, 4) and got["stats"]["rays_primary"] == 0
    # no lights: ambient only
    s = scenes.SceneArrays(); s.add_sphere((0, 0, 20), 25.0, scenes.material(ambient=(1, 0.5, 0.25))); s.ambient[:] = (100, 100, 100)
    rt.set_scene(s)
    got = rt.render(scenes.make_params(9, 9, 10.0, 1.0, 1.0))
    assert (got["rgba8"][40] == np.array([100, 50, 25, 255], np.uint8)).all()


def test_partition_union_equals_full_frame(rtb, scenes):
    """rank r of w renders only its row bands; the union over ranks is bit-identical to the 1-rank frame."""
    s, p = scenes.random_soup(seed=5), scenes.soup_params(100, 77, 2, 1)
    full = rtb.RayTracer((0,)); full.set_scene(s)
    ref = full.render(p)
    for world in (2, 3):
        acc = {k: np.zeros_like(ref[k]) for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}
        for r in range(world):
            part = rtb.RayTracer((0,)); part.set_partition(r, world, 8); part.set_scene(s)
            g = part.render(p)
            rows = np.arange(77); mine = ((rows // 8) % world) == r
            mask = np.repeat(mine, 100)
            for k in acc:
                acc[k][mask] = g[k][mask]
            part.close()
        for k in acc:
            assert (acc[k] == ref[k]).all() or k == "rgba32f" and np.array_equal(acc[k], ref[k], equal_nan=True), (world, k)
    full.close()


def test_small_waves_equal_one_wave(rt, scenes):
    s, p = scenes.random_soup(seed=9), scenes.soup_params(96, 64, 3, 1)
    rt.set_scene(s)
    a = rt.render(p)
    rt.set_option("chunk_pixels", 512)
    b = rt.render(p)
    rt.set_option("chunk_pixels", 1 << 23)
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k


@pytest.mark.parametrize("lanes,deep", [(1, 1), (2, 2), (3, 8), (8, 8)])
def test_lane_count_does_not_change_the_frame(rtb, scenes, lanes, deep):
    """Waves rotate over 1..8 lanes (streams + buffer sets; 8 for frames with mirror depths and spp passes): the per-pixel
    accumulation order of the spp passes and every result stay the same, through rt_render_device and through rt_render."""
    s, p = scenes.random_soup(seed=9), scenes.soup_params(160, 96, 4, 2)          # depth 4, 2x2 spp: 4 passes
    r = rtb.RayTracer((0,)); r.set_scene(s)
    r.set_option("lanes", 1); r.set_option("deep_lanes", 1); r.set_option("chunk_pixels", 160 * 96)
    a = r.render(p)
    r.set_option("lanes", lanes); r.set_option("deep_lanes", deep)
    for chunk in (160 * 96, 2048):                                                 # 1 wave per pass / 8 waves per pass
        r.set_option("chunk_pixels", chunk)
        for _ in range(2):                                                        # second frame: hints active
            b = r.render(p)
            for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
                assert np.array_equal(a[k], b[k], equal_nan=True), (chunk, k)
        assert a["stats"]["rays_mirror"] == b["stats"]["rays_mirror"] and a["stats"]["rays_shadow"] == b["stats"]["rays_shadow"]
    r.close()


def test_host_sah_builder_parity(rtb, scenes, oracle):
    """option builder=1 (host binned-SAH BVH, csrc/rt_sah.h): same results as the reference's scan."""
    r = rtb.RayTracer((0,)); r.set_option("builder", 1)
    try:
        s, p = scenes.random_soup(seed=7), scenes.soup_params(96, 64, 3, 1)
        ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
        r.set_scene(s)
        _check(compare_renders(ref, r.render(p)))
        s, p = scenes.cfg3_scene(n=80), scenes.cfg3_params(640, 360)
        ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED)
        r.set_scene(s)
        got = r.render(p)
        _check(compare_renders(ref, got))
        assert (got["stats"]["rays_primary"], got["stats"]["rays_shadow"], got["stats"]["rays_mirror"]) == ref["stats"]
        # builder options that change the tree, never the result: SAH leaf termination, reference pre-splitting (a triangle
        # then sits in several leaves; more slots than primitives), and a refit of such a tree
        n_nodes = r.stats()["n_nodes"]
        r.set_option("sah_term", 1.0); r.set_option("presplit", 2.5); r.set_option("presplit_depth", 3)
        r.set_scene(s)
        assert r.stats()["n_nodes"] > n_nodes
        g2 = r.render(p)
        for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
            assert np.array_equal(got[k], g2[k], equal_nan=True), k
        r.update_lights(s.lights, s.ambient); r.refit()
        g3 = r.render(p)
        for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
            assert np.array_equal(got[k], g3[k], equal_nan=True), k
    finally:
        r.close()


def test_golden_fixtures_on_gpu(rt, scenes):
    """committed golden vectors (tests/golden/*.npz) vs the CUDA path: ids/t bit-exact, u8 within 1 LSB, ray counts equal."""
    import os
    from golden import make_golden as G
    for name, mk in G.CASES.items():
        s, p = mk(scenes)
        g = np.load(os.path.join(os.path.dirname(os.path.abspath(G.__file__)), name + ".npz"))
        rt.set_scene(s)
        got = rt.render(p)
        assert np.array_equal(got["hit_id"][:, :2], g["hit_id"][:, :2]), name
        assert np.array_equal(got["hit_t"], g["hit_t"]), name
        assert np.abs(got["rgba8"].astype(int) - g["rgba8"].astype(int)).max() <= 1, name
        st = got["stats"]
        assert [st["rays_primary"], st["rays_shadow"], st["rays_mirror"]] == [int(v) for v in g["stats"]], name


def test_packet_and_per_lane_traversal_agree(rtb, scenes):
    """warp-coherent packet traversal (rt_packet.cuh) vs the per-lane while-while loop: bit-identical frames."""
    s, p = scenes.cfg3_scene(n=48), scenes.cfg3_params(480, 270)
    r = rtb.RayTracer((0,)); r.set_scene(s)
    try:
        a = r.render(p)
        r.set_option("packet", 0)
        b = r.render(p)
        for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
            assert np.array_equal(a[k], b[k], equal_nan=True), k
    finally:
        r.close()


def test_in_process_multi_device_equals_single(rtb, scenes):
    """rt_create with 2 devices: interleaved bands + peer-store gather into device 0 == the 1-device frame, bit for bit."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s, p = scenes.random_soup(seed=5), scenes.soup_params(200, 150, 2, 1)
    one = rtb.RayTracer((0,)); one.set_scene(s); a = one.render(p); one.close()
    two = rtb.RayTracer((0, 1)); two.set_partition(0, 1, 8); two.set_scene(s); b = two.render(p)
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k
    assert a["stats"]["rays_shadow"] == b["stats"]["rays_shadow"] and a["stats"]["rays_mirror"] == b["stats"]["rays_mirror"]
    # dirty-tracking updates reach every device of the context
    s2 = _moved_copy(s, (0.5, 0.25, -0.5))
    one = rtb.RayTracer((0,)); one.set_scene(s2); a2 = one.render(p); one.close()
    two.update_spheres(0, s2.spheres); two.update_triangles(0, s2.tris, s2.tri_normals)
    two.update_mesh(0, s2.meshes[0][0], s2.meshes[0][1], s2.meshes[0][3]); two.update_lights(s2.lights, s2.ambient)
    two.recommit()
    b2 = two.render(p)
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a2[k], b2[k], equal_nan=True), k
    # ... and so does a refit (tree kept on every device): back to the first layout
    two.update_spheres(0, s.spheres); two.update_triangles(0, s.tris, s.tri_normals)
    two.update_mesh(0, s.meshes[0][0], s.meshes[0][1], s.meshes[0][3]); two.update_lights(s.lights, s.ambient)
    two.refit()
    b3 = two.render(p); two.close()
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a[k], b3[k], equal_nan=True), k
    # host SAH builder on two devices: mesh updates go to the devices only, the host rebuild fetches the vertices back
    two = rtb.RayTracer((0, 1)); two.set_option("builder", 1); two.set_partition(0, 1, 8); two.set_scene(s)
    two.update_spheres(0, s2.spheres); two.update_triangles(0, s2.tris, s2.tri_normals)
    two.update_mesh(0, s2.meshes[0][0], s2.meshes[0][1], s2.meshes[0][3]); two.update_lights(s2.lights, s2.ambient)
    two.refit(); b4 = two.render(p)
    two.recommit(); b5 = two.render(p); two.close()
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a2[k], b4[k], equal_nan=True) and np.array_equal(a2[k], b5[k], equal_nan=True), k


def test_render_full_size_properties(rt, scenes, oracle):
    """BASELINE config 2 at its full size (3840x2160, depth 4): size-independent checks — the 8x down-sampled lattice of
    the 4K frame equals the same pixels rendered as a 480x270 frame only where rays coincide, so instead check
    (a) determinism (two renders bit-identical), (b) ray-count identities, (c) every primary miss is background."""
    s, p = scenes.rt_unity(mirror_sphere=True), scenes.rt_unity_params(3840, 2160, 4)
    rt.set_scene(s)
    a = rt.render(p, 2 | 4 | 8)
    b = rt.render(p, 2 | 4 | 8)
    for k in ("rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a[k], b[k]), k
    st = a["stats"]
    hits = int((a["hit_id"][:, 0] != 0).sum())
    assert st["rays_primary"] == 3840 * 2160
    assert hits <= st["rays_shadow"] <= hits + st["rays_mirror"]        # one light: one shadow query per hit event
    miss = a["hit_id"][:, 0] == 0
    assert (a["rgba8"][miss] == np.array([0, 0, 0, 255], np.uint8)).all()
    assert (a["hit_t"][miss] == np.finfo(np.float32).max).all()
    # (d) 4 096 random pixels of the 4K frame against the faithful linear scan (mirror depth 4, Phong 20), float colour included
    px = np.random.default_rng(2).choice(3840 * 2160, 4096, replace=False)
    c = rt.render(p, 15)
    ref = oracle.OracleScene(s, build_bvh=False).render(p, oracle.FAITHFUL, pixels=px)
    _check(compare_renders(ref, {k: c[k][px] for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}))


def test_pipelined_copy_out_equals_plain(rt, scenes):
    """rt_render overlaps the copy-out of finished waves with the next waves' kernels (copy_waves): same bytes out."""
    s, p = scenes.random_soup(seed=9), scenes.soup_params(333, 201, 2, 1)
    rt.set_scene(s)
    rt.set_option("copy_waves", 1)
    a = rt.render(p)
    for w in (2, 8, 64):
        rt.set_option("copy_waves", w)
        b = rt.render(p)
        for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
            assert np.array_equal(a[k], b[k], equal_nan=True), (w, k)
    rt.set_option("copy_waves", 8)


def test_cfg3_full_size(rtb, scenes, oracle):
    """BASELINE config 3 at full size (1 003 522 primitives, 3840x2160): (a) frames from the device LBVH and from the host SAH
    BVH are bit-identical (the result may not depend on the tree), (b) 4 096 random pixels equal the oracle (CPU BVH mode),
    (c) ray counts: one shadow query per (hit, light)."""
    s, p = scenes.cfg3_scene(), scenes.cfg3_params()
    want = 1 | 2 | 4 | 8
    frames = []
    for builder in (0, 1):
        r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_scene(s)
        frames.append(r.render(p, want))
        frames.append(r.render(p, want))                 # second frame runs in heavy-first tile order: same bytes
        r.close()
    for f in frames[1:]:
        for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
            assert np.array_equal(frames[0][k], f[k]), k
    st = frames[0]["stats"]
    hits = int((frames[0]["hit_id"][:, 0] != 0).sum())
    assert st["rays_primary"] == 3840 * 2160 and st["rays_shadow"] == 2 * hits and st["rays_mirror"] == 0
    px = np.random.default_rng(3).choice(3840 * 2160, 4096, replace=False)
    ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED, pixels=px)
    got = {k: frames[0][k][px] for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}
    _check(compare_renders(ref, got))                    # ids, t, u8 and float colour (Phong 20 on every triangle)


def test_cfg4_reduced_many_meshes(rt, scenes, oracle):
    """BASELINE config 4 scaled down (2x2x2 baked instances of a 5 120-triangle sphere): per-mesh AABB gates + per-mesh materials."""
    s, p = scenes.cfg4_scene(n=16, grid=2), scenes.cfg4_params(480, 270)
    ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED)
    rt.set_scene(s)
    got = rt.render(p)
    _check(compare_renders(ref, got))
    assert (got["hit_id"][:, 0] == 3).any() and len(np.unique(got["hit_id"][got["hit_id"][:, 0] == 3, 2])) > 3


def test_cfg5_reduced_spheres_mirrors_spp(rt, scenes, oracle):
    """BASELINE config 5 scaled down (2 000 spheres with 10 % mirrors + 8 000-triangle sphere, 2x2 spp, depth 4)."""
    s, p = scenes.cfg5_scene(n=20, n_spheres=2000), scenes.cfg5_params(240, 135, 2, 4)
    ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED)
    rt.set_scene(s)
    got = rt.render(p)
    _check(compare_renders(ref, got))
    st = got["stats"]
    assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == ref["stats"] and st["rays_mirror"] > 0


def test_streaming_commit_equals_staged_commit_and_recommits(rtb, scenes, oracle):
    """Commit path: rt_scene_mesh uploads straight from the caller's arrays (stream_upload, the default
    with the device builder) — same frame as the host-staged commit, also when scenes of different sizes follow each other on
    one context, as the reference's per-frame UpdateScene does (RayTracingSetup.cs:120-128,184)."""
    big, pb = scenes.cfg4_scene(n=16, grid=2), scenes.cfg4_params(240, 135)
    small, ps = scenes.rt_unity(), scenes.rt_unity_params()
    want = {}
    for name, (s, p) in {"big": (big, pb), "small": (small, ps)}.items():
        r = rtb.RayTracer((0,)); r.set_option("stream_upload", 0); r.set_scene(s)
        want[name] = r.render(p); r.close()
    _check(compare_renders(oracle.OracleScene(big).render(pb, oracle.ACCELERATED), want["big"]))
    r = rtb.RayTracer((0,))
    for name in ("small", "big", "small", "big", "big"):           # grow, shrink, grow again, identical re-commit
        s, p = (big, pb) if name == "big" else (small, ps)
        r.set_scene(s)
        got = r.render(p)
        for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
            assert np.array_equal(want[name][k], got[k]), (name, k)
        assert r.stats()["h2d_bytes"] >= 48 * s.n_prims - 32 * s.spheres.shape[0]
    r.close()


def _moved_copy(s, d):
    """the same scene after every sphere, loose triangle, mesh 0 and the lights moved by d (a translation keeps the normals)"""
    import copy
    d = np.asarray(d, np.float32)
    m = copy.deepcopy(s)
    m.spheres[:, :3] += d
    m.tris += np.tile(d, 3)
    v, n, mat, aabb = m.meshes[0]
    m.meshes[0] = (v + np.tile(d, 3), n, mat, aabb + np.tile(d, 2))
    m.lights[:, :3] += d
    return m


@pytest.mark.parametrize("builder", [0, 1])
def test_dirty_updates_plus_recommit_equal_full_commit(rtb, scenes, oracle, builder):
    """rt_scene_update_* + rt_scene_recommit = committing the moved scene from scratch = the oracle."""
    s, p = scenes.cfg5_scene(n=12, n_spheres=500), scenes.cfg5_params(160, 90, 1, 2)
    s2 = _moved_copy(s, (0.75, -0.5, 1.25))
    ref = oracle.OracleScene(s2).render(p, oracle.ACCELERATED)
    fresh = rtb.RayTracer((0,)); fresh.set_option("builder", builder); fresh.set_scene(s2)
    want = fresh.render(p); fresh.close()
    _check(compare_renders(ref, want))
    r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_scene(s)
    r.render(p)
    half = s.spheres.shape[0] // 2
    r.update_spheres(0, s2.spheres[:half]); r.update_spheres(half, s2.spheres[half:])
    r.update_triangles(0, s2.tris, s2.tri_normals)
    r.update_mesh(0, s2.meshes[0][0], s2.meshes[0][1], s2.meshes[0][3])
    r.update_lights(s2.lights, s2.ambient)
    with pytest.raises(RuntimeError):
        r.render(p)                                             # dirty scene: RT_E_STATE until rt_scene_recommit
    r.recommit()
    got = r.render(p)
    for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
        assert np.array_equal(want[k], got[k]), k
    st = r.stats()
    dirty = 48 * (s.tris.shape[0] + s.meshes[0][0].shape[0]) + 16 * s.spheres.shape[0]
    assert dirty <= st["h2d_bytes"] < dirty + 4096                  # the dirty arrays + the small tables; materials stay on the device
    with pytest.raises(RuntimeError):
        r.update_spheres(s.spheres.shape[0] - 1, s2.spheres[:2])    # beyond the committed count
    r.recommit()
    got = r.render(p)
    assert np.array_equal(want["hit_id"], got["hit_id"])
    r.close()


@pytest.mark.parametrize("builder", [0, 1])
def test_refit_equals_full_commit(rtb, scenes, oracle, builder):
    """rt_scene_update_* + rt_scene_refit (tree kept, boxes recomputed bottom-up) = committing the moved scene from scratch = the
    oracle, for both builders' trees, repeatedly, and for a deformation the old tree was not built for."""
    s, p = scenes.cfg5_scene(n=12, n_spheres=500), scenes.cfg5_params(160, 90, 1, 2)
    r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_scene(s)
    r.render(p)
    n_nodes = r.stats()["n_nodes"]
    rng = np.random.default_rng(5)
    for step, shift in enumerate([(0.75, -0.5, 1.25), (-2.0, 1.0, 0.5)]):
        s2 = _moved_copy(s, shift)
        if step == 1:                                            # scramble: every sphere jumps somewhere else, the mesh is squashed
            s2.spheres[:, :3] = rng.uniform(-30, 30, s2.spheres[:, :3].shape).astype(np.float32) + np.float32([0, 0, 60])
            v = s2.meshes[0][0].copy(); v.reshape(-1, 3, 3)[:, :, 1] *= np.float32(0.5)
            lo, hi = v.reshape(-1, 3).min(axis=0), v.reshape(-1, 3).max(axis=0)
            s2.meshes[0] = (v, s2.meshes[0][1], s2.meshes[0][2], np.concatenate([lo, hi]).astype(np.float32))
        ref = oracle.OracleScene(s2).render(p, oracle.ACCELERATED)
        fresh = rtb.RayTracer((0,)); fresh.set_option("builder", builder); fresh.set_scene(s2)
        want = fresh.render(p); fresh.close()
        _check(compare_renders(ref, want))
        r.update_spheres(0, s2.spheres)
        r.update_triangles(0, s2.tris, s2.tri_normals)
        r.update_mesh(0, s2.meshes[0][0], s2.meshes[0][1], s2.meshes[0][3])
        r.update_lights(s2.lights, s2.ambient)
        with pytest.raises(RuntimeError):
            r.render(p)                                         # dirty until refit / recommit
        r.refit()
        got = r.render(p)
        for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
            assert np.array_equal(want[k], got[k]), (step, k)
        assert r.stats()["n_nodes"] == n_nodes                  # same tree
    r.recommit()                                                # and a rebuild afterwards still works
    got = r.render(p)
    assert np.array_equal(want["hit_id"], got["hit_id"]) and np.array_equal(want["rgba32f"], got["rgba32f"])
    r.close()
    e = rtb.RayTracer((0,))
    with pytest.raises(RuntimeError):
        e.refit()                                               # nothing committed
    e.close()


def test_light_split_and_hints_do_not_change_the_frame(rtb, scenes):
    """Scheduling only: heavy-first tile order (hints, active from the second frame on) and per-light work items for heavy
    tiles (light_split, combined by the last item through atomics) must reproduce the first frame bit for bit."""
    s, p = scenes.cfg3_scene(n=48), scenes.cfg3_params(960, 540)
    base = rtb.RayTracer((0,)); base.set_option("hints", 0); base.set_scene(s)
    want = base.render(p); base.close()
    for ls in (1, 3, 6):
        r = rtb.RayTracer((0,)); r.set_option("light_split", ls); r.set_scene(s)
        for frame in range(3):
            got = r.render(p)
            for k in ("hit_id", "hit_t", "rgba8", "rgba32f"):
                assert np.array_equal(want[k], got[k]), (ls, frame, k)
            assert got["stats"]["rays_shadow"] == want["stats"]["rays_shadow"]
        r.close()


def test_cfg4_full_size(rtb, scenes, oracle):
    """BASELINE config 4 at full size (125 meshes, 10 240 000 triangles, 7680x4320): (a) the second frame (heavy-first tile
    order) equals the first, (b) the union of a 4-rank band partition (quarter-frame partitions: one wave, per-light items)
    equals the 1-rank frame, (c) 2 048 random pixels equal the oracle (CPU BVH mode), (d) one shadow query per (hit, light)."""
    s, p = scenes.cfg4_scene(), scenes.cfg4_params()
    want = 1 | 2 | 4 | 8
    px = np.random.default_rng(4).choice(7680 * 4320, 2048, replace=False)
    r = rtb.RayTracer((0,)); r.set_scene(s)
    f0 = r.render(p, want)
    ids, t, rgba = f0["hit_id"].copy(), f0["hit_t"].copy(), f0["rgba8"].copy()
    rgbaf = f0["rgba32f"][px].copy()
    f1 = r.render(p, want, out=f0)
    assert np.array_equal(ids, f1["hit_id"]) and np.array_equal(t, f1["hit_t"]) and np.array_equal(rgba, f1["rgba8"])
    assert np.array_equal(rgbaf, f1["rgba32f"][px])
    st = f1["stats"]
    hits = int((ids[:, 0] != 0).sum())
    assert st["rays_primary"] == 7680 * 4320 and st["rays_shadow"] == 2 * hits and st["rays_mirror"] == 0
    assert len(np.unique(ids[ids[:, 0] == 3, 2])) >= 50                      # most of the 125 meshes are visible
    r.close()
    rows = np.arange(4320)
    for rank in range(4):
        part = rtb.RayTracer((0,)); part.set_partition(rank, 4, 32); part.set_scene(s)
        g = None
        for frame in range(2):                                               # second frame: hints + per-light work items
            g = part.render(p, 2 | 8, out=g)
        mine = np.repeat(((rows // 32) % 4) == rank, 7680)
        assert np.array_equal(g["rgba8"][mine], rgba[mine]) and np.array_equal(g["hit_t"][mine], t[mine]), rank
        part.close()
    ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED, pixels=px)
    _check(compare_renders(ref, dict(hit_id=ids[px], hit_t=t[px], rgba8=rgba[px], rgba32f=rgbaf)))


def test_cfg5_full_size(rtb, scenes, oracle):
    """BASELINE config 5 at full size (100 000 spheres + 5 000 000 triangles, 3840x2160, 4x4 spp, depth 4): (a) device LBVH and
    host SAH BVH give the same frame, (b) 512 random pixels (16 samples each, mirrors included) equal the oracle,
    (c) ray counts: primary = 16 per pixel, mirror rays present."""
    s, p = scenes.cfg5_scene(), scenes.cfg5_params()
    want = 1 | 2 | 4 | 8
    frames = []
    for builder in (0, 1):
        r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_scene(s)
        f = r.render(p, want)
        frames.append({k: f[k].copy() if k != "stats" else f[k] for k in f})
        r.close()
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(frames[0][k], frames[1][k]), k
    st = frames[0]["stats"]
    assert st["rays_primary"] == 16 * 3840 * 2160 and st["rays_mirror"] > 0 and st["rays_shadow"] > st["rays_primary"]
    assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == tuple(frames[1]["stats"][k] for k in ("rays_primary", "rays_shadow", "rays_mirror"))
    px = np.random.default_rng(5).choice(3840 * 2160, 512, replace=False)
    ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED, pixels=px)
    _check(compare_renders(ref, {k: frames[0][k][px] for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}))
    # the sphere silhouettes of this scene are where the reference's binary32 discriminant decides (rt_sphere_box_radius): the
    # accelerated oracle itself against the faithful scan on some of those pixels
    sph = px[ref["hit_id"][:, 0] == 1][:24]
    if len(sph):
        fa = oracle.OracleScene(s, build_bvh=False).render(p, oracle.FAITHFUL, pixels=sph)
        _check(compare_renders(fa, {k: frames[0][k][sph] for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}))


def test_registered_host_buffer_gives_the_same_frame(rtb, scenes):
    """rt_host_register page-locks a caller array that is reused across frames (the PixelColors array): same bytes as the
    pageable and the pipelined paths, and unregistering leaves the array usable."""
    s, p = scenes.cfg3_scene(n=32), scenes.cfg3_params(960, 544)
    r = rtb.RayTracer((0,)); r.set_scene(s)
    want = r.render(p)
    out = r.alloc_host_outputs(p, 1 | 2)
    r.host_register(out["rgba32f"]); r.host_register(out["rgba8"])
    for frame in range(2):
        out["rgba32f"][:] = -1; out["rgba8"][:] = 7
        r.render(p, 1 | 2, out=out)
        assert np.array_equal(want["rgba32f"], out["rgba32f"]) and np.array_equal(want["rgba8"], out["rgba8"])
    r.host_unregister(out["rgba32f"]); r.host_unregister(out["rgba8"])
    r.render(p, 1 | 2, out=out)
    assert np.array_equal(want["rgba32f"], out["rgba32f"])
    with pytest.raises(RuntimeError):
        r.host_unregister(out["rgba8"])                     # not registered any more: CUDA error surfaces as a status code
    r.close()


# ---- round 2 -------------------------------------------------------------------------------------------------------------
def test_c_smoke_test_of_the_boundary(rtb, tmp_path):
    """tests/abi_smoke.c: the header compiled as strict C99 and the whole call sequence of a C caller (create -> scene ->
    commit -> rt_render / rt_intersect -> KAT-1..4 -> destroy), SURVEY §8(b)."""
    import subprocess
    from test_abi import build_abi_smoke
    exe = build_abi_smoke(str(tmp_path))
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and "abi_smoke ok" in r.stdout, r.stdout + r.stderr


def test_lights_appearing_after_a_frame_without_lights(rtb, scenes, oracle):
    """a frame with 0 point lights launches no shadow kernel, so it records no tile costs for it; the
    next frame (lights switched on through rt_scene_update_lights, same layout otherwise) must not take those empty class
    lists for hints — the fused kernel would hand out no tile at all and leave the frame unwritten."""
    import copy
    s = scenes.random_soup(seed=5, n_tris=900, n_spheres=80)
    p = scenes.soup_params(192, 128, 0, 1)
    lit = copy.deepcopy(s)
    dark = copy.deepcopy(s); dark.lights = dark.lights[:0]
    for builder in (0, 1):
        r = rtb.RayTracer((0,)); r.set_option("builder", builder)
        r.set_scene(dark)
        ref_dark = oracle.OracleScene(dark).render(p, oracle.FAITHFUL)
        for _ in range(2):
            _check(compare_renders(ref_dark, r.render(p)))
        r.update_lights(lit.lights, lit.ambient); r.recommit()
        ref_lit = oracle.OracleScene(lit).render(p, oracle.FAITHFUL)
        for _ in range(3):                                  # first frame without hints, then with the hints it recorded
            got = r.render(p)
            _check(compare_renders(ref_lit, got))
            assert got["stats"]["rays_shadow"] == ref_lit["stats"][1] > 0
        r.set_option("hints", 0)
        _check(compare_renders(ref_lit, r.render(p)))
        r.update_lights(dark.lights, dark.ambient); r.refit()               # and back, through the refit path
        r.set_option("hints", 1)
        for _ in range(2):
            _check(compare_renders(ref_dark, r.render(p)))
        r.close()


@pytest.mark.parametrize("distance", [2e3, 2e5, 3e6])
def test_far_camera_keeps_the_scan_result(rtb, scenes, oracle, distance):
    """a camera 1e2 .. 1e5 scene extents away (k_trace_primary<FAR>, sphere boxes inflated for the camera's
    distance through the automatic refit): same fixture as tests/test_emu.py, here on the GPU, both builders and both loops."""
    from test_emu import far_camera_case
    s, p = far_camera_case(scenes, distance)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    assert (ref["hit_id"][:, 0] != 0).sum() > 100
    for builder, packet in ((0, 1), (1, 1), (1, 0)):
        r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_option("packet", packet)
        r.set_scene(s)
        for _ in range(2):
            _check(compare_renders(ref, r.render(p)))
        r.close()


def test_sphere_silhouettes_match_the_scan(rtb, scenes, oracle):
    """Small spheres far from the ray origin: the reference's binary32 discriminant hits them out to ~1.2 radii
    (rt_build_steps.h rt_sphere_box_radius).  Explicit rays through rt_intersect, from inside and far outside the scene,
    and a rendered frame whose camera sits outside the scene (automatic refit of the sphere boxes), vs the faithful scan."""
    from test_emu import grazing_sphere_rays
    rng = np.random.default_rng(11)
    s = scenes.SceneArrays()
    m = scenes.material(diffuse=(0.7, 0.7, 0.7), ambient=(1, 1, 1), mirror=(0.5, 0.5, 0.5), is_mirror=1)
    n_s = 1500
    centers = rng.uniform(-40, 40, (n_s, 3)) + np.array((0, 0, 100.0))
    radii = rng.uniform(0.1, 0.4, n_s)
    for c, r in zip(centers, radii):
        s.add_sphere(tuple(np.float32(c)), float(np.float32(r * r)), m)
    t = np.array((-60, -45, 160, 60, -45, 160, 0, 60, 160), np.float32)
    s.add_triangle(t, scenes.triangle_normal(t)[0], m)
    s.add_light((0, 80, 0), 1e5); s.add_light((-70, 10, 60), 1e5); s.ambient[:] = (20, 20, 20)
    centers = s.spheres[:, :3].astype(np.float64); radii = np.sqrt(s.spheres[:, 3].astype(np.float64))
    ref_scene = oracle.OracleScene(s)
    p = scenes.make_params(256, 192, 10.0, 4.5, 3.4, background=(0.1, 0.1, 0.1), bounces=2)     # camera at the origin, outside the sphere cloud
    ref = ref_scene.render(p, oracle.FAITHFUL)
    for builder in (0, 1):
        r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_scene(s)
        for box in (((-1, -1, -1), (1, 1, 1)), ((-40, -40, 60), (40, 40, 140)), ((-900, -900, -2000), (900, 900, -1000))):
            O, D = grazing_sphere_rays(rng, centers, radii, np.asarray(box, np.float64), 16)
            rid, rt_ = ref_scene.intersect(O, D, oracle.FAITHFUL)
            gid, gt = r.intersect(O, D)
            assert np.array_equal(rid[:, :2], gid[:, :2]) and np.array_equal(rt_, gt), (builder, box, int((rid[:, :2] != gid[:, :2]).any(axis=1).sum()))
        for _ in range(2):
            _check(compare_renders(ref, r.render(p)))
        # a direction that is not normalised: the reference's sphere test then describes another shape; rt_intersect scans
        O = np.zeros((64, 3), np.float32); D = (centers[:64] * 1.7).astype(np.float32)
        rid, rt_ = ref_scene.intersect(O, D, oracle.FAITHFUL)
        gid, gt = r.intersect(O, D)
        assert np.array_equal(rid[:, :2], gid[:, :2]) and np.array_equal(rt_, gt, equal_nan=True)
        r.close()


@pytest.mark.parametrize("world,band_rows,res", [(2, 32, (640, 360)), (4, 32, (512, 777)), (8, 16, (300, 1100)), (3, 8, (129, 431))])
def test_ranks_copy_their_bands_into_one_host_frame(rtb, scenes, world, band_rows, res):
    """rt_render on a partition: the rank renders into its own framebuffer and copies ONLY its bands, wave by wave, straight
    into the caller's full-frame arrays (one shared host frame for all ranks — bench.py's N-GPU e2e path, where it is a POSIX
    shared-memory segment).  After every rank has rendered into the same arrays they hold the 1-rank frame, bit for bit, for
    every output, whatever the wave count; rows of other ranks are never touched."""
    s = scenes.random_soup(seed=5)
    p = scenes.soup_params(res[0], res[1], 2, 1)
    full = rtb.RayTracer((0,)); full.set_scene(s); ref = full.render(p); full.close()
    for waves in (0, 1, 3):
        frame = rtb.RayTracer.alloc_host_outputs(p)
        frame["rgba32f"][:] = -7.0; frame["hit_t"][:] = -7.0; frame["hit_id"][:] = -7; frame["rgba8"][:] = 7
        for r in range(world):
            part = rtb.RayTracer((0,)); part.set_partition(r, world, band_rows); part.set_option("copy_waves_part", waves)
            part.set_scene(s)
            before = {k: frame[k].copy() for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}
            part.render(p, out=frame)
            rows = np.arange(res[1]); mine = np.repeat(((rows // band_rows) % world) == r, res[0])
            for k in before:
                assert np.array_equal(frame[k][~mine], before[k][~mine], equal_nan=True), (world, r, k, "foreign rows written")
            with pytest.raises(rtb.host.RtError):            # nothing was gathered: rt_read_outputs must say so
                part.read_outputs(p)
            part.close()
        for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
            assert np.array_equal(frame[k], ref[k], equal_nan=True), (world, waves, k)


def test_hand_derived_kats_on_the_gpu(rtb, scenes):
    """KAT-5 (Blinn-Phong), KAT-6 (depth-1 mirror) — tests/kats.py, exact rational evaluation of the reference's formulas —
    through rt_render; KAT-7 (slab test, 0 * inf) through the scene gate of rt_intersect; both builders."""
    import exact32 as X
    import kats
    for builder in (0, 1):
        for kat in (kats.kat5_blinn_phong, kats.kat6_mirror_depth1):
            k = kat()
            r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_scene(k["scene"])
            got = r.render(k["params"])
            assert tuple(got["hit_id"][0][:2]) == k["hit"] and got["hit_t"][0] == k["t"]
            expect = np.array([X.div(c, np.float32(255)) for c in k["value"]], np.float32)
            # (float)pow: CUDA's double pow (or the exact squaring path) against the hand value's glibc pow — at most last bits
            assert np.allclose(got["rgba32f"][0][:3], expect, rtol=3e-7, atol=0), (got["rgba32f"][0], expect)
            assert np.array_equal(got["rgba32f"][0][1:3], expect[1:3]) or kat is kats.kat6_mirror_depth1
            st = got["stats"]
            assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == (1, k["shadow_rays"], k["mirror_rays"])
            r.close()
        # KAT-7: Scene.IntersectRay gates on RayAABBIntersection(ray, Scene.AABB) (Scene.cs:54): a sphere inside the box, the box
        # given as Scene.AABB; the ray can only report the sphere when the slab test lets it through
        s = scenes.SceneArrays()
        s.add_sphere((1.2, 0.0, 5.5), 0.3 * 0.3, scenes.material(diffuse=(1, 1, 1)))      # on the path of the rays at x = 1
        s.add_sphere((1.9, 0.0, 5.5), 0.7 * 0.7, scenes.material(diffuse=(1, 1, 1)))      # on the path of x = 2 AND of x = 2.5 (outside the box)
        s.scene_aabb = np.array((1, -1, 5, 2, 1, 6), np.float32)
        r = rtb.RayTracer((0,)); r.set_option("builder", builder); r.set_scene(s)
        for o, d, _box, want in kats.kat7_slab_zero_direction():
            ids, t = r.intersect([o], [d])
            assert (ids[0][0] == 1) == (want and d[2] > 0), (o, d, ids[0])               # x = 2.5 would hit the second sphere but for the gate
        r.close()


def test_sorted_mirror_queues_change_nothing(rtb, scenes, oracle):
    """Option sort_queue (rt_render.cu k_qs_*): mirror-ray queues counting-sorted by (direction bin, origin cell) before each
    depth is traced.  A queue entry carries its parent index, so the order may change the time only: frames with the sort forced
    on, forced off and on auto are identical (all four outputs) and equal the oracle — mirrors between mirrors (depth 64 chains),
    a soup with 30 % mirror materials at 2x2 spp over several waves and lanes, cfg2 at 480x270."""
    from conftest import mirror_walls_scene
    cases = [(mirror_walls_scene(scenes), scenes.make_params(96, 64, 10.0, 8.0, 5.0, (0.1, 0.2, 0.3), 64, 1), {}),
             (scenes.random_soup(seed=3), scenes.soup_params(160, 96, 3, 2), {"chunk_pixels": 2048, "deep_lanes": 4}),
             (scenes.rt_unity(mirror_sphere=True), scenes.rt_unity_params(480, 270, 4), {})]
    for s, p, opts in cases:
        ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
        frames = []
        for mode in (1, 0, -1):
            r = rtb.RayTracer((0,))
            r.set_option("sort_queue", mode)
            for k, v in opts.items():
                r.set_option(k, v)
            r.set_scene(s)
            for _ in range(2):
                got = r.render(p)
            frames.append(got)
            r.close()
        _check(compare_renders(ref, frames[0]))
        st = frames[0]["stats"]
        assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == ref["stats"] and st["rays_mirror"] > 0
        assert st["kernel_launches"] > frames[1]["stats"]["kernel_launches"]          # the sort kernels did run
        for f in frames[1:]:
            for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
                assert np.array_equal(frames[0][k], f[k], equal_nan=True), k

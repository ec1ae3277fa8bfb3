"""Instanced meshes (rt_scene_base_mesh / rt_scene_mesh_instance / rt_scene_update_instance).

The reference bakes localToWorldMatrix into world-space vertices for every mesh on every frame (SceneMesh.cs:21-31) and
its scan sees baked triangles only.  The claim tested here: an instanced mesh renders BIT-IDENTICALLY to the same mesh baked
with the reference's arithmetic (scenes.bake_tris: Matrix4x4.MultiplyPoint3x4, -Triangle.Normal) — hit ids, distances and float
colours, no edge exemption — because only the culling happens in object space.  So the oracle consumes `scene.baked()`.

not gpu: the product's device headers on the CPU (tests/emu) against the faithful oracle.
gpu    : through the C-ABI; instanced frame == baked frame on the GPU (all four outputs) and == oracle."""
import numpy as np
import pytest

from conftest import compare_renders


def _exact(cmp):
    assert cmp["id"] == 0 and cmp["t"] == 0 and cmp["u8_gt0"] == 0 and cmp["f32"] == 0, cmp


def _rot(axis, deg):
    a = np.asarray(axis, np.float64); a /= np.linalg.norm(a)
    c, s = np.cos(np.radians(deg)), np.sin(np.radians(deg))
    K = np.array(((0, -a[2], a[1]), (a[2], 0, -a[0]), (-a[1], a[0], 0)))
    return np.eye(3) * c + s * K + (1 - c) * np.outer(a, a)


def _fixtures(scenes):
    return [("cfg1 cube as an instance", scenes.rt_unity(instanced=True), scenes.rt_unity_params()),
            ("cfg2 cube as an instance", scenes.rt_unity(True, True), scenes.rt_unity_params(96, 54, 4)),
            ("instances + baked meshes + spheres + loose triangles", scenes.instanced_soup(), scenes.soup_params(96, 64, 3)),
            ("instances only, 2x2 spp", _instances_only(scenes), scenes.soup_params(80, 48, 2, 2))]


def _instances_only(scenes):
    s = scenes.instanced_soup(seed=5, n_bases=2, n_instances=14, n_baked=0)
    s.spheres, s.sphere_mats = s.spheres[:0], s.sphere_mats[:0]
    s.tris, s.tri_normals, s.tri_mats = s.tris[:0], s.tri_normals[:0], s.tri_mats[:0]
    return s


# ---------------------------------------------------------------------------------------------------------------
# CPU: the device headers (rt_core.h two-level traversal, rt_instances.h BLAS / TLAS packing) against the oracle
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("builder", [0, 1])
def test_emulated_instances_equal_the_baked_scene(emu, scenes, oracle, builder):
    for name, s, p in _fixtures(scenes):
        ref = oracle.OracleScene(s.baked()).render(p, oracle.FAITHFUL)
        e = emu.EmuTracer(builder=builder)
        e.set_scene(s)
        got = e.render(p)
        _exact(compare_renders(ref, got))
        assert (got["stats"]["rays_primary"], got["stats"]["rays_shadow"], got["stats"]["rays_mirror"]) == ref["stats"], name
        assert (got["hit_id"][:, 0] == 3).sum() > 50, name                   # instanced triangles are actually being hit
        # the baked twin through the same code is the same frame too
        b = emu.EmuTracer(builder=builder)
        b.set_scene(s.baked())
        gb = b.render(p)
        for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
            assert np.array_equal(got[k], gb[k], equal_nan=True), (name, k)


def test_emulated_golden_cfg1_with_the_cube_instanced(emu, scenes):
    """the committed golden fixture of cfg1 (the oracle's frame of the BAKED RT.unity scene) is reproduced with the cube as an instance"""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cfg1_rt_unity_50x50.npz"))
    e = emu.EmuTracer()
    e.set_scene(scenes.rt_unity(instanced=True))
    got = e.render(scenes.rt_unity_params())
    assert np.array_equal(got["hit_id"][:, :2], g["hit_id"][:, :2]) and np.array_equal(got["hit_t"], g["hit_t"])
    assert np.array_equal(got["rgba8"], g["rgba8"])


def test_emulated_repose_and_explicit_rays(emu, scenes, oracle):
    """rt_scene_update_instance + recommit == a fresh commit of the re-posed scene == oracle; rt_intersect through instances,
    normalised (two-level walk) and not (linear scan over every instanced triangle)."""
    s, p = scenes.instanced_soup(seed=9, n_instances=7), scenes.soup_params(72, 48, 2)
    e = emu.EmuTracer()
    e.set_scene(s)
    rng = np.random.default_rng(3)
    inst = [i for i, m in enumerate(s.meshes) if isinstance(m, scenes.MeshInstance)]
    for i in inst[::2]:                                                     # move every other instance, scale and all
        m = s.meshes[i]
        M = m.m3x4.copy()
        M[:, :3] = (_rot(rng.normal(size=3), rng.uniform(0, 360)) @ M[:, :3].astype(np.float64) * rng.uniform(0.7, 1.4)).astype(np.float32)
        M[:, 3] += rng.uniform(-6, 6, 3).astype(np.float32)
        m.m3x4 = M
        e.update_instance(i, M, scenes.bake_tris(s.base_meshes[m.base], M)[2])
    e.recommit()
    got = e.render(p)
    ora = oracle.OracleScene(s.baked())
    _exact(compare_renders(ora.render(p, oracle.FAITHFUL), got))
    o = rng.uniform(-40, 40, (3000, 3)).astype(np.float32); o[:, 2] += 75
    d = rng.normal(size=(3000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    d[::5] *= np.float32(1.7)                                               # not normalised: rt_scan_all
    rid, rt_ = ora.intersect(o, d, oracle.FAITHFUL)
    gid, gt = e.intersect(o, d)
    assert np.array_equal(rid[:, :2], gid[:, :2]) and np.array_equal(rt_, gt)
    mesh = rid[:, 0] == 3
    assert mesh.sum() > 30 and np.array_equal(rid[mesh, 2], gid[mesh, 2])


def _extreme_scene(scenes):
    """instances scaled so that object coordinates are ~1e3 / ~1e-3, sheared, mirrored (negative determinant), and an empty base mesh"""
    s = scenes.SceneArrays()
    rng = np.random.default_rng(17)
    k = 120
    base = (rng.uniform(-1, 1, (k, 1, 3)) + rng.normal(scale=0.3, size=(k, 3, 3))).reshape(k, 9).astype(np.float32)
    b = s.add_base_mesh(base)
    empty = s.add_base_mesh(np.zeros((0, 9), np.float32))
    mats = [scenes.material((0.9, 0.3, 0.2), (1, 1, 1), (0, 0, 0), (1, 1, 1), 12.0, 0), scenes.material((0.2, 0.6, 0.9), (1, 1, 1), (0.8, 0.8, 0.8), (1, 1, 1), 30.0, 1)]

    def M(lin, t):
        out = np.zeros((3, 4), np.float32); out[:, :3] = np.asarray(lin, np.float32); out[:, 3] = t
        return out
    s.add_instance(b, M(np.eye(3) * 6.0, (0, 0, 40)), mats[0])
    big, small = s.add_base_mesh(base * np.float32(1000.0)), s.add_base_mesh(base * np.float32(0.001))
    s.add_instance(big, M(_rot((1, 1, 0), 40) * 4e-3, (9, 2, 46)), mats[1])                        # object coordinates ~1e3, scale 4e-3
    s.add_instance(empty, M(np.eye(3), (0, 0, 30)), mats[0])
    s.add_instance(b, M(np.diag((-3.0, 3.0, 3.0)), (-10, -3, 50)), mats[0])                       # mirrored
    s.add_instance(b, M(np.array(((4, 1.5, 0), (0, 4, 0.7), (0.3, 0, 4))), (4, 8, 55)), mats[1])  # sheared
    s.add_instance(small, M(_rot((0, 1, 1), 70) * 4e3, (2, -8, 44)), mats[0])                       # object coordinates ~1e-3, scale 4e3
    s.add_light((30, 40, 0), 2e5); s.add_light((-30, 10, 20), 1e5)
    s.ambient = np.array((15, 15, 15), np.float32)
    return s


def _cameras_near_and_far(scenes):
    """the scene seen from inside, from 30 000 and from 60 000 units away (hundreds of scene extents: the FAR form of the walk)"""
    out = []
    for pos in ((0.0, 0.0, 0.0), (0.0, 0.0, -30000.0), (52000.0, 31000.0, 44.0)):
        fwd = np.array((0, 0, 46.0)) - np.array(pos); dist = np.linalg.norm(fwd); fwd /= dist
        right = np.cross((0, 1, 0), fwd); right /= np.linalg.norm(right)
        up = np.cross(fwd, right)
        cam = dict(pos=tuple(np.float32(x) for x in pos), fwd=tuple(fwd.astype(np.float32)), right=tuple(right.astype(np.float32)), up=tuple(up.astype(np.float32)))
        out.append((pos, scenes.make_params(64, 48, float(max(10.0, dist - 30.0)), 16.0, 12.0, (0.1, 0.1, 0.2), 2, 1, cam=cam)))
    return out


def test_emulated_far_camera_and_extreme_instances(emu, scenes, oracle):
    """(a) a camera hundreds of scene extents away (the FAR form pads the object-space slabs per ray), (b) instances whose object
    coordinates are ~1e3 and ~1e-3, sheared, mirrored, and an empty base mesh: culling stays conservative, results exact."""
    s = _extreme_scene(scenes)
    ora = oracle.OracleScene(s.baked())
    for pos, p in _cameras_near_and_far(scenes):
        for builder in (0, 1):
            e = emu.EmuTracer(builder=builder)
            e.set_scene(s)
            got = e.render(p)
            _exact(compare_renders(ora.render(p, oracle.FAITHFUL), got))
            assert (got["hit_id"][:, 0] == 3).sum() > 100, pos


def test_bake_arithmetic_is_the_references():
    """scenes.bake_tris (the fixture-side bake, SceneMesh.cs:21-52) against a scalar restatement: MultiplyPoint3x4 left to right."""
    import importlib
    scenes = importlib.import_module("unity-raytracer_b200").scenes
    rng = np.random.default_rng(1)
    F = np.float32
    M = rng.normal(size=(3, 4)).astype(F)
    v = rng.normal(size=(5, 9)).astype(F)
    tris, normals, aabb = scenes.bake_tris(v, M)
    for t in range(5):
        for k in range(3):
            x, y, z = v[t, 3 * k: 3 * k + 3]
            for r in range(3):
                want = F(F(F(M[r, 0] * x) + F(M[r, 1] * y)) + F(M[r, 2] * z)) + M[r, 3]
                assert tris[t, 3 * k + r] == F(want)
    assert np.array_equal(aabb[:3], tris.reshape(-1, 3).min(0)) and np.array_equal(aabb[3:], tris.reshape(-1, 3).max(0))


# ---------------------------------------------------------------------------------------------------------------
# GPU, through the C-ABI
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("builder", [0, 1])
def test_gpu_instances_equal_baked_and_oracle(rtb, scenes, oracle, builder):
    for name, s, p in _fixtures(scenes):
        ref = oracle.OracleScene(s.baked()).render(p, oracle.FAITHFUL)
        frames = []
        for scene in (s, s.baked()):
            for packet in (1, 0):
                r = rtb.RayTracer((0,))
                r.set_option("builder", builder); r.set_option("packet", packet)
                r.set_scene(scene)
                for _ in range(2):                                          # second frame: heavy-first tile order
                    got = r.render(p)
                frames.append(got)
                r.close()
        c = compare_renders(ref, frames[0])
        assert c["id"] == 0 and c["t"] == 0 and c["u8_gt1"] == 0 and c["f32_far"] == 0 and c["f32"] <= 2, (name, c)   # pow() only
        st = frames[0]["stats"]
        assert (st["rays_primary"], st["rays_shadow"], st["rays_mirror"]) == ref["stats"], name
        for f in frames[1:]:                                                # instanced / baked, packet / per-lane: one frame
            for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
                assert np.array_equal(frames[0][k], f[k], equal_nan=True), (name, k)


@pytest.mark.gpu
def test_gpu_instance_updates_errors_and_explicit_rays(rtb, scenes, oracle):
    s, p = scenes.instanced_soup(seed=9, n_instances=7), scenes.soup_params(72, 48, 2)
    r = rtb.RayTracer((0,))
    r.set_scene(s)
    r.render(p)
    rng = np.random.default_rng(3)
    inst = [i for i, m in enumerate(s.meshes) if isinstance(m, scenes.MeshInstance)]
    baked_mesh = [i for i, m in enumerate(s.meshes) if not isinstance(m, scenes.MeshInstance)]
    for rep in range(2):
        batch = []
        for i in inst[rep::2]:
            m = s.meshes[i]
            M = m.m3x4.copy()
            M[:, :3] = (_rot(rng.normal(size=3), rng.uniform(0, 360)) @ M[:, :3].astype(np.float64) * rng.uniform(0.7, 1.4)).astype(np.float32)
            M[:, 3] += rng.uniform(-6, 6, 3).astype(np.float32)
            m.m3x4 = M
            if rep == 0: r.update_instance(i, M, scenes.bake_tris(s.base_meshes[m.base], M)[2])          # one by one, Mesh.AABB given
            else: batch.append((i, M))
        if batch:                                                           # rt_scene_update_instances: one call, AABBs computed by the library
            r.update_instances([i for i, _ in batch], np.stack([M for _, M in batch]), None)
        with pytest.raises(rtb.host.RtError) as ei:                         # dirty scene: rendering needs the recommit
            r.render(p)
        assert ei.value.code == -4
        (r.recommit if rep == 0 else r.refit)()
        got = r.render(p)
        ora = oracle.OracleScene(s.baked())
        ref = ora.render(p, oracle.FAITHFUL)
        c = compare_renders(ref, got)
        assert c["id"] == 0 and c["t"] == 0 and c["u8_gt1"] == 0 and c["f32_far"] == 0 and c["f32"] <= 2, c
        fresh = rtb.RayTracer((0,)); fresh.set_scene(s)
        gf = fresh.render(p); fresh.close()
        for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
            assert np.array_equal(got[k], gf[k], equal_nan=True), k
    o = rng.uniform(-40, 40, (20000, 3)).astype(np.float32); o[:, 2] += 75
    d = rng.normal(size=(20000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    d[::5] *= np.float32(1.7)
    rid, rt_ = ora.intersect(o, d, oracle.FAITHFUL)
    for fn in (r.intersect, r.intersect_static):
        gid, gt = fn(o, d)
        assert np.array_equal(rid[:, :2], gid[:, :2]) and np.array_equal(rt_, gt)
        mesh = rid[:, 0] == 3
        assert mesh.sum() > 100 and np.array_equal(rid[mesh, 2], gid[mesh, 2])
    # errors: every one leaves the scene as it was
    ident = np.eye(3, 4, dtype=np.float32)
    for bad, code in (((inst[0], np.zeros((3, 4), np.float32), None), -5), ((inst[0], ident * np.float32(np.nan), None), -5),
                      ((baked_mesh[0], ident, None), -1), ((len(s.meshes), ident, None), -1)):
        with pytest.raises(rtb.host.RtError) as ei:
            r.update_instance(*bad)
        assert ei.value.code == code, bad
    # the batched call is all or nothing: a bad LAST entry leaves the poses of the good ones before it untouched (scene not even dirty)
    moved = ident.copy(); moved[:, 3] = 50.0
    for bad_idx, bad_m, code in ((baked_mesh[0], ident, -1), (inst[1], np.zeros((3, 4), np.float32), -5), (-1, ident, -1)):
        with pytest.raises(rtb.host.RtError) as ei:
            r.update_instances([inst[0], bad_idx], np.stack([moved, bad_m]), None)
        assert ei.value.code == code and "entry 1" in str(ei.value), (bad_idx, str(ei.value))
    r.update_instances([], np.zeros((0, 12), np.float32), None)             # n = 0: a no-op, not an error, nothing dirty
    m0 = s.meshes[baked_mesh[0]]
    with pytest.raises(rtb.host.RtError):
        r.update_mesh(inst[0], m0[0], m0[1], m0[3])                         # an instance is not re-sent as triangles
    again = r.render(p)
    assert np.array_equal(again["rgba32f"], got["rgba32f"], equal_nan=True)
    # the whole scene declared again (what a shim that re-sends everything per frame does): the unchanged base meshes get their BLAS
    # back from the previous scene, a changed one is rebuilt; frames as from a fresh context
    r.set_scene(s)
    assert np.array_equal(r.render(p)["rgba32f"], got["rgba32f"], equal_nan=True)
    s.base_meshes[0] = (s.base_meshes[0] * np.float32(1.25)).astype(np.float32)
    r.set_scene(s)
    fresh = rtb.RayTracer((0,)); fresh.set_scene(s)
    assert np.array_equal(r.render(p)["rgba32f"], fresh.render(p)["rgba32f"], equal_nan=True)
    fresh.close()
    # the packed BLAS sections are kept across a re-declaration only while they were emitted from exactly these base meshes: other base
    # meshes declared in a scene WITHOUT instances (nothing is packed for it) and then again in a scene with instances must be emitted
    sa, sa_plain = scenes.instanced_soup(seed=11, n_instances=7), scenes.instanced_soup(seed=11, n_instances=7)
    sa_plain.meshes = [m for m in sa_plain.meshes if not isinstance(m, scenes.MeshInstance)]
    r.set_scene(sa_plain); r.render(p)
    r.set_scene(sa)
    fresh = rtb.RayTracer((0,)); fresh.set_scene(sa)
    a, f = r.render(p), fresh.render(p)
    for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
        assert np.array_equal(a[k], f[k], equal_nan=True), k
    r.set_scene(sa)                                                         # and unchanged once more: nothing of the base meshes is re-sent
    assert np.array_equal(r.render(p)["rgba32f"], f["rgba32f"], equal_nan=True)
    fresh.close()
    r.close()


@pytest.mark.gpu
def test_gpu_mesh_aabb_of_reposed_instances_computed_on_the_device(rtb, scenes):
    """rt_scene_update_instance(s) with aabb = NULL on a committed scene: Mesh.AABB (SceneMesh.cs:22-31) comes from a reduction over the
    base mesh resident on the device.  A scene of ONE instance has Scene.AABB == that Mesh.AABB (Scene.cs:17-41), which must equal the
    min / max of the baked vertices bit for bit — for rotated, scaled and sheared poses, and for poses that put an extremum at exactly
    0 (the sign of a zero depends on the vertex order: the library re-derives those on the host)."""
    base, _ = scenes.displaced_icosphere(40, 3.0, (0.0, 0.0, 0.0), 77)              # 32 000 triangles, 96 000 vertices ...
    base = np.concatenate([base, base[:2000] * np.float32(1.01)]).astype(np.float32)   # ... + 6 000: above the device path's threshold
    assert 3 * len(base) >= 100000
    s = scenes.SceneArrays()
    b = s.add_base_mesh(base)
    M0 = np.eye(3, 4, dtype=np.float32); M0[:, 3] = (0, 0, 30)
    s.add_instance(b, M0, scenes.material((0.5, 0.5, 0.5), (1, 1, 1)))
    s.add_light((10, 20, 0), 1e4)
    p = scenes.make_params(64, 48, 10.0, 4.0, 3.0)
    r = rtb.RayTracer((0,))
    r.set_scene(s)
    rng = np.random.default_rng(5)
    vmin = base.reshape(-1, 3).min(axis=0)
    poses = []
    for k in range(6):
        M = np.zeros((3, 4), np.float32)
        M[:, :3] = (_rot(rng.normal(size=3), rng.uniform(0, 360)) * rng.uniform(0.5, 2.0)).astype(np.float32)
        if k == 4: M[0, 1] += np.float32(0.3)                                           # shear
        M[:, 3] = rng.uniform(-5, 5, 3).astype(np.float32) + np.float32((0, 0, 30))
        poses.append(M)
    Z = np.eye(3, 4, dtype=np.float32); Z[:, 3] = -vmin                                 # x + (-xmin) == 0 exactly: zero extrema on all three axes
    poses.append(Z)
    for k, M in enumerate(poses):
        if k % 2: r.update_instance(0, M, None)
        else: r.update_instances([0], M[None], None)
        r.recommit()
        want = scenes.bake_tris(base, M)[2]
        got = r.scene_aabb()
        assert np.array_equal(got.view(np.uint32), np.asarray(want, np.float32).view(np.uint32)), (k, got, want)
    # and the frame is the one of a context that was GIVEN the box
    s.meshes[0].m3x4 = poses[2]; s.meshes[0].aabb = scenes.bake_tris(base, poses[2])[2]
    r.update_instance(0, poses[2], None); r.recommit()
    fresh = rtb.RayTracer((0,)); fresh.set_scene(s)
    a, f = r.render(p), fresh.render(p)
    for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
        assert np.array_equal(a[k], f[k], equal_nan=True), k
    assert (a["hit_id"][:, 0] == 3).sum() > 100
    fresh.close(); r.close()


@pytest.mark.gpu
def test_gpu_far_camera_and_extreme_instances(rtb, scenes, oracle):
    """the FAR instantiation of the instanced primary kernel and the padding of strongly scaled instances, on the GPU"""
    s = _extreme_scene(scenes)
    ora = oracle.OracleScene(s.baked())
    for pos, p in _cameras_near_and_far(scenes):
        ref = ora.render(p, oracle.FAITHFUL)
        for packet in (1, 0):
            r = rtb.RayTracer((0,)); r.set_option("packet", packet); r.set_scene(s)
            got = r.render(p)
            c = compare_renders(ref, got)
            assert c["id"] == 0 and c["t"] == 0 and c["u8_gt1"] == 0 and c["f32_far"] == 0 and c["f32"] <= 2, (pos, packet, c)
            r.close()


@pytest.mark.gpu
def test_gpu_instanced_band_partition_and_host_copy_out(rtb, scenes):
    """the union of a 3-rank band partition of an instanced scene == the 1-rank frame; rt_render (pipelined copy-out) == rt_render_device"""
    s, p = scenes.instanced_soup(seed=2, n_instances=12), scenes.soup_params(320, 256, 2)
    one = rtb.RayTracer((0,)); one.set_scene(s)
    full = one.render(p)
    rows = np.arange(256)
    for rank in range(3):
        part = rtb.RayTracer((0,)); part.set_partition(rank, 3, 32); part.set_scene(s)
        g = part.render(p)
        mine = np.repeat(((rows // 32) % 3) == rank, 320)
        for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
            assert np.array_equal(g[k][mine], full[k][mine], equal_nan=True), (rank, k)
        part.close()
    one.close()


@pytest.mark.gpu
def test_gpu_cfg4_instanced_full_size(rtb, scenes, oracle):
    """BASELINE config 4 as an instanced-mesh scene at full size: ONE 81 920-triangle base mesh, 125 instances, 7680x4320.
    (a) the instanced frame == the frame of the 10 240 000 baked triangles, all four outputs, every pixel; (b) 2 048 random
    pixels == the oracle on the baked scene; (c) the device holds megabytes, not the 0.9 GB of the baked scene;
    (d) re-posing all 125 instances + recommit: sub-millisecond, and the new frame again equals the baked one."""
    s, p = scenes.cfg4_instanced(), scenes.cfg4_params()
    want = 1 | 2 | 4 | 8
    import torch
    free0 = torch.cuda.mem_get_info()[0]
    r = rtb.RayTracer((0,)); r.set_option("builder", 1); r.set_scene(s)
    scene_bytes = free0 - torch.cuda.mem_get_info()[0]
    fi = r.render(p, want)
    fi = {k: (v.copy() if k != "stats" else v) for k, v in fi.items()}
    sb = s.baked()
    rb = rtb.RayTracer((0,)); rb.set_option("builder", 1); rb.set_scene(sb)
    fb = rb.render(p, want)
    for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
        assert np.array_equal(fi[k], fb[k], equal_nan=True), k
    assert fi["stats"]["rays_shadow"] == fb["stats"]["rays_shadow"] and fi["stats"]["rays_primary"] == 7680 * 4320
    assert len(np.unique(fi["hit_id"][fi["hit_id"][:, 0] == 3, 2])) >= 50
    px = np.random.default_rng(4).choice(7680 * 4320, 2048, replace=False)
    ref = oracle.OracleScene(sb).render(p, oracle.ACCELERATED, pixels=px)
    c = compare_renders(ref, {k: fi[k][px] for k in ("hit_id", "hit_t", "rgba8", "rgba32f")})
    assert c["id"] == 0 and c["t"] == 0 and c["u8_gt1"] == 0 and c["f32_far"] == 0 and c["f32"] <= max(2, c["n"] // 200), c
    assert scene_bytes < 64 << 20, scene_bytes                               # (allocator granularity included; the baked scene: ~0.9 GB)
    # re-pose everything
    import time
    rng = np.random.default_rng(8)
    for m in s.meshes:
        M = m.m3x4.copy()
        M[:, :3] = (_rot(rng.normal(size=3), rng.uniform(0, 360)) @ M[:, :3].astype(np.float64)).astype(np.float32)
        M[:, 3] += rng.uniform(-1.5, 1.5, 3).astype(np.float32)
        m.m3x4 = M
    boxes = [scenes.bake_tris(s.base_meshes[m.base], m.m3x4)[2] for m in s.meshes]
    t0 = time.perf_counter()
    for i, m in enumerate(s.meshes):
        r.update_instance(i, m.m3x4, boxes[i])
    r.recommit()
    ms = (time.perf_counter() - t0) * 1e3
    f2 = r.render(p, want)
    sb2 = s.baked()
    for i, m in enumerate(sb2.meshes):
        rb.update_mesh(i, m[0], m[1], m[3])
    rb.recommit()
    fb2 = rb.render(p, want)
    for k in ("hit_id", "hit_t", "rgba32f", "rgba8"):
        assert np.array_equal(f2[k], fb2[k], equal_nan=True), k
    assert not np.array_equal(f2["hit_t"], fi["hit_t"])
    assert ms < 5.0, ms                                                     # 125 ctypes calls + TLAS rebuild + 16 KB upload (the baked twin: ~1 s)
    r.close(); rb.close()


@pytest.mark.gpu
def test_gpu_instances_on_an_in_process_two_device_context(rtb, scenes):
    """rt_create with 2 devices and an instanced scene: both output modes (rt_render: per-device copy-out; rt_render_device: peer-store
    gather) equal the 1-device frame; re-posing reaches every device."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s, p = scenes.instanced_soup(seed=2, n_instances=12), scenes.soup_params(320, 256, 2)
    one = rtb.RayTracer((0,)); one.set_scene(s); a = one.render(p)
    two = rtb.RayTracer((0, 1)); two.set_partition(0, 1, 8); two.set_scene(s)
    b = two.render(p)
    two.render_device(p, 15); c = two.read_outputs(p, 15)
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a[k], b[k], equal_nan=True) and np.array_equal(a[k], c[k], equal_nan=True), k
    inst = [i for i, m in enumerate(s.meshes) if isinstance(m, scenes.MeshInstance)]
    for i in inst[::3]:
        M = s.meshes[i].m3x4.copy(); M[:, 3] += np.float32(2.5); s.meshes[i].m3x4 = M
        one.update_instance(i, M); two.update_instance(i, M)
    one.recommit(); two.recommit()
    a2, b2 = one.render(p), two.render(p)
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(a2[k], b2[k], equal_nan=True), k
    assert not np.array_equal(a2["hit_t"], a["hit_t"])
    one.close(); two.close()

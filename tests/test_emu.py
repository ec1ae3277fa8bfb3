"""-m "not gpu": the product's host/device headers (rt_core.h traversal + shading, rt_build_steps.h LBVH build,
rt_host_scene.h flattening) instantiated serially on the CPU by tests/emu and compared with the oracle.

This checks the *logic* the CUDA kernels run — LBVH build, BVH-vs-linear-scan equivalence incl. tie-break and AABB gates,
wave/depth bookkeeping, pixel<->tile mapping, band partition — where no GPU exists.  It is not a CPU path of the product:
tests/emu is test infrastructure and librtb200.so never links it (tests/test_abi.py checks the product fails without CUDA).
"""
import numpy as np
import pytest

from conftest import compare_renders


def _strict(cmp):
    # same libm on both sides here, so even the float colour must be bit-identical
    assert cmp["id"] == 0 and cmp["t"] == 0 and cmp["u8_gt0"] == 0 and cmp["f32"] == 0, cmp


def test_cfg1_rt_unity(emu, oracle, scenes):
    s, p = scenes.rt_unity(), scenes.rt_unity_params()
    e = emu.EmuTracer(); e.set_scene(s)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    got = e.render(p)
    _strict(compare_renders(ref, got))
    assert (got["stats"]["rays_primary"], got["stats"]["rays_shadow"], got["stats"]["rays_mirror"]) == ref["stats"]
    assert np.array_equal(e.scene_aabb(), oracle.OracleScene(s).aabb())


def test_cfg2_mirror_recursion_order(emu, oracle, scenes):
    """Iterative wave + back-to-front fold == the reference's recursion (RayTracingSetup.cs:358-363), bit for bit."""
    s, p = scenes.rt_unity(mirror_sphere=True), scenes.rt_unity_params(160, 90, 4)
    e = emu.EmuTracer(); e.set_scene(s)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    got = e.render(p)
    _strict(compare_renders(ref, got))
    assert got["stats"]["rays_mirror"] == ref["stats"][2] > 0


@pytest.mark.parametrize("leaf_max,builder", [(1, 0), (4, 0), (8, 0), (1, 1), (4, 1), (8, 1)])
def test_bvh_equals_linear_scan(emu, oracle, scenes, leaf_max, builder):
    """device-LBVH logic (builder 0) and the host binned-SAH builder (builder 1, rt_sah.h) vs the reference's scan"""
    s, p = scenes.random_soup(seed=7), scenes.soup_params(64, 48, 3, 1)
    e = emu.EmuTracer(leaf_max=leaf_max, builder=builder); e.set_scene(s)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    _strict(compare_renders(ref, e.render(p)))
    assert e.max_stack() < 64


@pytest.mark.parametrize("builder", [0, 1])
def test_refit_keeps_the_tree_and_the_result(emu, oracle, scenes, builder):
    """rt_scene_refit's logic (rt_node_parents + bottom-up rt_ref_box over the emitted node array, rt_build_steps.h): after
    the primitives moved — here: scrambled, the worst case for the old tree — the refitted tree still answers every query like
    the reference's scan of the moved scene."""
    import copy
    s, p = scenes.random_soup(seed=7), scenes.soup_params(64, 48, 3, 1)
    e = emu.EmuTracer(leaf_max=4, builder=builder); e.set_scene(s)
    _strict(compare_renders(oracle.OracleScene(s).render(p, oracle.FAITHFUL), e.render(p)))
    n_nodes = e.n_nodes()
    rng = np.random.default_rng(2)
    m = copy.deepcopy(s)
    m.spheres[:, :3] += rng.uniform(-6, 6, m.spheres[:, :3].shape).astype(np.float32)
    m.tris += np.tile(rng.uniform(-6, 6, (m.tris.shape[0], 3)).astype(np.float32), 3)
    for k, (v, n, mat, aabb) in enumerate(m.meshes):
        v = (v + np.tile(rng.uniform(-3, 3, (v.shape[0], 3)).astype(np.float32), 3)).astype(np.float32)
        lo, hi = v.reshape(-1, 3).min(axis=0), v.reshape(-1, 3).max(axis=0)
        m.meshes[k] = (v, n, mat, np.concatenate([lo, hi]).astype(np.float32))
    m.scene_aabb = None
    e.set_scene(m, refit=True)
    assert e.n_nodes() == n_nodes
    _strict(compare_renders(oracle.OracleScene(m).render(p, oracle.FAITHFUL), e.render(p)))


@pytest.mark.parametrize("ratio,depth", [(2.5, 3), (1.0, 5)])
def test_presplit_references_keep_the_result(emu, oracle, scenes, ratio, depth):
    """Reference pre-splitting of the host SAH builder (rt_sah.h RtPresplit): thin diagonal triangles enter the tree as several
    clipped pieces, so a triangle sits in several leaves — the query result (argmin over (t, scan key), any-hit) must not
    change, nor must a refit of the split tree."""
    import copy
    for s, p in ((scenes.random_soup(seed=7), scenes.soup_params(64, 48, 3, 1)), (scenes.cfg3_scene(n=8), scenes.cfg3_params(96, 54))):
        plain = emu.EmuTracer(leaf_max=4, builder=1); plain.set_scene(s)
        e = emu.EmuTracer(leaf_max=4, builder=1); e.set_presplit(ratio, depth); e.set_scene(s)
        assert e.n_slots() > plain.n_slots() == s.n_prims                  # something was split
        ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
        _strict(compare_renders(ref, e.render(p)))
        m = copy.deepcopy(s)
        m.tris += np.float32(0.25)
        m.meshes = [(v + np.float32(0.25), n, mat, aabb + np.float32(0.25)) for (v, n, mat, aabb) in m.meshes]
        m.scene_aabb = None
        e.set_scene(m, refit=True)
        _strict(compare_renders(oracle.OracleScene(m).render(p, oracle.FAITHFUL), e.render(p)))


def test_spp_grid_extension(emu, oracle, scenes):
    s, p = scenes.random_soup(seed=3), scenes.soup_params(48, 32, 2, 2)
    e = emu.EmuTracer(); e.set_scene(s)
    _strict(compare_renders(oracle.OracleScene(s).render(p, oracle.FAITHFUL), e.render(p)))


def test_intersect_random_rays(emu, oracle, scenes):
    s = scenes.random_soup(seed=11)
    rng = np.random.default_rng(5)
    o = rng.uniform(-40, 40, (4000, 3)).astype(np.float32)
    d = rng.normal(size=(4000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    rid, rt_ = oracle.OracleScene(s).intersect(o, d, oracle.FAITHFUL)
    e = emu.EmuTracer(); e.set_scene(s)
    gid, gt = e.intersect(o, d)
    assert (rid[:, :2] == gid[:, :2]).all() and (rt_ == gt).all()


def test_duplicate_primitives_tie_break(emu, scenes):
    """Identical primitives (equal Morton codes, equal t): the lowest scan index must win (Scene.cs strict '>')."""
    s = scenes.SceneArrays()
    tri = np.array((-1, -1, 5, 1, -1, 5, 0, 1, 5), np.float32)
    n = scenes.triangle_normal(tri)[0]
    for _ in range(9):
        s.add_triangle(tri, n, scenes.material())
    s.add_sphere((0, 0, 6), 1.0, scenes.material())          # front pole also at t = 5: sphere (class 1) beats loose tris (class 2)
    for leaf_max, builder in ((1, 0), (4, 0), (1, 1), (4, 1)):
        e = emu.EmuTracer(leaf_max=leaf_max, builder=builder); e.set_scene(s)
        ids, t = e.intersect([(0, 0, 0)], [(0, 0, 1)])
        assert t[0] == np.float32(5.0) and ids[0, :2].tolist() == [1, 0]
    s.spheres = s.spheres[:0]; s.sphere_mats = s.sphere_mats[:0]
    e = emu.EmuTracer(leaf_max=2); e.set_scene(s)
    assert e.intersect([(0, 0, 0)], [(0, 0, 1)])[0][0, :2].tolist() == [2, 0]


def test_mesh_aabb_gate_is_semantic(emu, oracle, scenes):
    """Scene.cs:67: a mesh whose (caller-supplied) AABB the ray misses is skipped even if a triangle would be hit."""
    s = scenes.SceneArrays()
    tri = np.array((-1, -1, 5, 1, -1, 5, 0, 1, 5), np.float32).reshape(1, 9)
    nm, _ = scenes.mesh_from_world_tris(tri)
    s.add_mesh(tri, nm, scenes.material(), np.array((10, 10, 10, 11, 11, 11), np.float32))     # deliberately wrong box
    s.scene_aabb = np.array((-20, -20, -20, 20, 20, 20), np.float32)
    e = emu.EmuTracer(); e.set_scene(s)
    ids, t = e.intersect([(0, 0, 0)], [(0, 0, 1)])
    rid, rt_ = oracle.OracleScene(s).intersect([(0, 0, 0)], [(0, 0, 1)])
    assert ids[0].tolist() == rid[0].tolist() == [0, -1, -1] and t[0] == rt_[0]


def test_scene_aabb_gate_is_semantic(emu, oracle, scenes):
    """Scene.cs:54: a ray that misses the (caller-supplied) scene AABB misses everything."""
    s = scenes.rt_unity()
    s.scene_aabb = np.array((100, 100, 100, 101, 101, 101), np.float32)
    e = emu.EmuTracer(); e.set_scene(s)
    p = scenes.rt_unity_params(20, 20)
    got, ref = e.render(p), oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    assert (got["hit_id"][:, 0] == 0).all()
    _strict(compare_renders(ref, got))


def test_partition_union_and_waves(emu, scenes):
    s, p = scenes.random_soup(seed=5), scenes.soup_params(100, 77, 2, 1)
    e = emu.EmuTracer(); e.set_scene(s)
    ref = e.render(p)
    e.set_option("chunk_pixels", 256)
    small = e.render(p)
    for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
        assert np.array_equal(ref[k], small[k], equal_nan=True), k
    e.set_option("chunk_pixels", 1 << 23)
    for world in (2, 3, 8):
        acc = {k: np.zeros_like(ref[k]) for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}
        rays = 0
        for r in range(world):
            e.set_partition(r, world, 8)
            g = e.render(p)
            rays += g["stats"]["rays_primary"]
            mine = np.repeat(((np.arange(77) // 8) % world) == r, 100)
            for k in acc:
                acc[k][mine] = g[k][mine]
        e.set_partition(0, 1, 32)
        assert rays == 100 * 77
        for k in acc:
            assert np.array_equal(acc[k], ref[k], equal_nan=True), (world, k)


def test_empty_and_degenerate(emu, scenes):
    e = emu.EmuTracer(); e.set_scene(scenes.SceneArrays())
    got = e.render(scenes.make_params(5, 3, 10.0, 8.0, 5.0, (0.25, 0.5, 1.0)))
    assert (got["hit_id"] == np.array([0, -1, -1])).all() and (got["rgba8"] == np.array([64, 128, 255, 255], np.uint8)).all()
    got = e.render(scenes.make_params(0, 0, 10.0, 8.0, 5.0))
    assert got["rgba32f"].shape == (0, 4)
    # one primitive: root is a leaf
    s = scenes.SceneArrays(); s.add_sphere((0, 0, 20), 25.0, scenes.material(ambient=(1, 0.5, 0.25))); s.ambient[:] = (100, 100, 100)
    e = emu.EmuTracer(); e.set_scene(s)
    got = e.render(scenes.make_params(9, 9, 10.0, 1.0, 1.0))
    assert (got["rgba8"][40] == np.array([100, 50, 25, 255], np.uint8)).all()


@pytest.mark.parametrize("builder", [0, 1])
def test_cfg3_reduced(emu, oracle, scenes, builder):
    s, p = scenes.cfg3_scene(n=24), scenes.cfg3_params(160, 90)
    e = emu.EmuTracer(builder=builder); e.set_scene(s)
    ora = oracle.OracleScene(s)
    _strict(compare_renders(ora.render(p, oracle.ACCELERATED), e.render(p)))


def test_ray_tri_division_free_rejections_match_reference(emu, oracle):
    """rt_ray_tri decides most rejections on undivided numerators; every outcome must equal RMath.cs:29-73 (the oracle),
    including rays aimed exactly at edges/vertices (u, v, u+v on their thresholds), tiny and huge triangles, and NaN/inf."""
    rng = np.random.default_rng(42)
    n_checked = n_hit = 0
    for it in range(6000):
        scale = 10.0 ** rng.uniform(-2.5, 2.5)
        v = (rng.normal(size=(3, 3)) * scale + rng.normal(size=3) * 20).astype(np.float32)
        o = (rng.normal(size=3) * 30).astype(np.float32)
        mode = it % 4
        if mode == 0:                                      # random interior/exterior point
            w = rng.dirichlet((1, 1, 1)) if rng.uniform() < 0.7 else rng.normal(size=3)
        elif mode == 1:                                    # on an edge
            w = np.array([rng.uniform(), 0.0, 0.0]); w[1] = 1 - w[0]; w = np.roll(w, it % 3)
        elif mode == 2:                                    # on a vertex
            w = np.roll(np.array([1.0, 0.0, 0.0]), it % 3)
        else:                                              # just outside / inside an edge
            w = np.array([rng.uniform(), 0.0, 0.0]); w[1] = 1 - w[0] + rng.normal() * 1e-6; w = np.roll(w, it % 3)
        target = (w[:, None] * v.astype(np.float64)).sum(0) / max(w.sum(), 1e-9) if abs(w.sum()) > 1e-6 else v[0]
        d = target - o
        d = (d / np.linalg.norm(d)).astype(np.float32)
        if it % 7 == 0:
            d = -d
        rh, rt_ = oracle.ray_triangle(o, d, v.reshape(9))
        gh, gt = emu.ray_triangle(o, d, v.reshape(9))
        assert rh == gh and (not rh or rt_ == gt), (it, rh, gh, rt_, gt)
        n_checked += 1; n_hit += rh
        if rh:                                             # tmax rule: a cull distance beyond the hit never rejects it ...
            gh2, gt2 = emu.ray_triangle(o, d, v.reshape(9), float(rt_) * 1.000125)
            assert gh2 and gt2 == rt_
    assert n_hit > 1000
    for bad in (np.nan, np.inf):
        v = np.array((0, 0, 5, 1, 0, 5, 0, 1, 5), np.float32); v[4] = bad
        assert emu.ray_triangle((0.2, 0.2, 0), (0, 0, 1), v)[0] == oracle.ray_triangle((0.2, 0.2, 0), (0, 0, 1), v)[0]


def test_tile_mapping_division_by_multiplication(emu, scenes):
    """rt_item_to_pixel divides the tile index by tiles_x with a multiply + fix-up: odd widths/heights, all partitions."""
    s = scenes.SceneArrays(); s.add_sphere((0, 0, 20), 25.0, scenes.material(ambient=(1, 1, 1))); s.ambient[:] = (50, 50, 50)
    e = emu.EmuTracer(); e.set_scene(s)
    for (rx, ry) in ((1, 1), (7, 5), (9, 33), (1023, 3), (130, 131)):
        p = scenes.make_params(rx, ry, 10.0, 8.0, 5.0)
        full = e.render(p)
        assert full["stats"]["rays_primary"] == rx * ry
        for world, band in ((3, 4), (2, 12), (5, 32)):            # power-of-two and non-power-of-two band heights
            acc = np.zeros_like(full["hit_t"]); cnt = 0
            for r in range(world):
                e.set_partition(r, world, band)
                g = e.render(p)
                mine = np.repeat(((np.arange(ry) // band) % world) == r, rx)
                acc[mine] = g["hit_t"][mine]; cnt += g["stats"]["rays_primary"]
            e.set_partition(0, 1, 32)
            assert cnt == rx * ry and np.array_equal(acc, full["hit_t"]), (rx, ry, world, band)


@pytest.mark.parametrize("case", ["plain", "kd_inf", "neg_zero_ambient", "huge_light", "light_on_surface"])
def test_backfacing_light_shadow_skip_is_exact(emu, oracle, scenes, case):
    """rt_shadow_irrelevant: shadow queries of lights behind the surface are answered without traversal.  The reference
    (RayTracingSetup.cs:333-355,384-391) adds kd*0*E + 0 for such a light when it is visible and nothing when it is occluded:
    identical unless kd*0*E is not a zero or the colour is -0.0 — those scenes must fall back to the full query."""
    s = scenes.SceneArrays()
    kd = (np.inf, 0.5, 0.5) if case == "kd_inf" else (0.9, 0.5, 0.5)
    ka = (0.0, 1.0, 1.0)
    big = scenes.material(kd, ka, (0, 0, 0), (1, 1, 1), 20.0, 0)
    # a wall facing the camera (normal -z): one light in front of it, one BEHIND it (back-facing, visible: nothing in between),
    # one behind it and occluded by a second triangle
    s.add_triangle((-30, -30, 40, 0, 30, 40, 30, -30, 40), (0, 0, -1), big)
    s.add_triangle((-30, -30, 60, 0, 30, 60, 30, -30, 60), (0, 0, -1), scenes.material((0.2, 0.9, 0.2), (1, 1, 1)))
    s.add_sphere((12, 0, 30), 16.0, scenes.material((0.3, 0.3, 0.9), (1, 1, 1), (0, 0, 0), (1, 1, 1), 8.0, 0))
    s.add_light((0, 5, 0), 1e38 if case == "huge_light" else 2e5)
    s.add_light((5, 0, 50), 1e5)                       # between the walls: behind the first, in front of the second
    s.add_light((0, 0, 80), 1e5)                       # behind both
    if case == "light_on_surface":
        s.add_light((0, 0, 40), 1e5)                   # d2 -> 0 for the centre pixel: E = inf
    s.ambient = np.array((-1.0, 15, 15) if case == "neg_zero_ambient" else (15, 15, 15), np.float32)   # -1 * 0 = -0.0
    p = scenes.make_params(96, 64, 10.0, 8.0, 5.0, (0.1, 0.1, 0.1), 0, 1)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    e = emu.EmuTracer(); e.set_scene(s)
    got = e.render(p)
    assert np.array_equal(ref["hit_id"], got["hit_id"])
    assert np.array_equal(ref["rgba32f"].view(np.uint32), got["rgba32f"].view(np.uint32))     # bit pattern: NaN and -0.0 included
    assert got["stats"]["rays_shadow"] == ref["stats"][1]                                      # skipped queries are still counted


@pytest.mark.parametrize("nan", [False, True])
def test_degenerate_geometry_matches_the_scan(emu, oracle, scenes, nan):
    """conftest.degenerate_scene: both builders, leaf sizes 1 and 4, against the faithful scan."""
    from conftest import degenerate_scene
    s, p = degenerate_scene(scenes, nan), scenes.soup_params(64, 40, 2, 1)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    for builder in (0, 1):
        for leaf_max in (1, 4):
            e = emu.EmuTracer(leaf_max=leaf_max, builder=builder); e.set_scene(s)
            cmp = compare_renders(ref, e.render(p))
            assert cmp["id"] == 0 and cmp["t"] == 0 and cmp["u8_gt0"] == 0 and cmp["f32"] == 0, (builder, leaf_max, cmp)


def test_cameras_off_axis(emu, oracle, scenes):
    """conftest.camera_set: every view direction / octant and a sheared basis, soup with mirrors, against the faithful scan."""
    from conftest import camera_set
    s = scenes.random_soup(seed=4, n_tris=500, n_spheres=60)
    e = emu.EmuTracer(); e.set_scene(s)
    ora = oracle.OracleScene(s)
    for i, cam in enumerate(camera_set()):
        p = scenes.make_params(56, 40, 10.0, 8.0, 6.0, (0.1, 0.2, 0.3), 2, 1, cam=cam)
        cmp = compare_renders(ora.render(p, oracle.FAITHFUL), e.render(p))
        assert cmp["id"] == 0 and cmp["t"] == 0 and cmp["u8_gt0"] == 0 and cmp["f32"] == 0, (i, cmp)


def test_golden_fixtures_in_the_emulator(emu, scenes):
    """every committed golden vector (tests/golden/*.npz, edge cases included) through the device headers compiled for the CPU:
    ids and t bit-exact, u8 within 1 LSB — the check the GPU suite runs (test_golden_fixtures_on_gpu), without a GPU."""
    import os
    from golden import make_golden as G
    for name, mk in G.CASES.items():
        s, p = mk(scenes)
        g = np.load(os.path.join(os.path.dirname(os.path.abspath(G.__file__)), name + ".npz"))
        e = emu.EmuTracer(); e.set_scene(s)
        got = e.render(p)
        assert np.array_equal(got["hit_id"][:, :2], g["hit_id"][:, :2]), name
        assert np.array_equal(got["hit_t"], g["hit_t"]), name
        assert np.abs(got["rgba8"].astype(int) - g["rgba8"].astype(int)).max() <= 1, name


def far_camera_case(scenes, distance):
    """an axis-aligned cube (thin node boxes: a face seen edge-on is a zero-thickness slab) plus a few spheres, seen
    from `distance` away through a very narrow image plane, so that the slab test's rounding (|O/d| * 2^-24) exceeds the
    node padding (max|coordinate| * 2^-16)."""
    s = scenes.SceneArrays()
    h = 5.0
    corners = np.array([(x, y, z) for x in (-h, h) for y in (-h, h) for z in (-h, h)], np.float32) + np.array((1.0, -2.0, 3.0), np.float32)
    quads = ((0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3))
    tris = []
    for a, b, c, d in quads:
        tris.append(np.concatenate([corners[a], corners[b], corners[c]])); tris.append(np.concatenate([corners[a], corners[c], corners[d]]))
    tris = np.asarray(tris, np.float32)
    nrm, box = scenes.mesh_from_world_tris(tris)
    s.add_mesh(tris, nrm, scenes.material(diffuse=(0.2, 0.8, 0.7), ambient=(1, 1, 1), specular=(0.4, 0.4, 0.4), phong=12.0), box)
    s.add_sphere((9.0, 4.0, 1.0), 4.0, scenes.material(diffuse=(0.9, 0.2, 0.2), ambient=(1, 1, 1)))
    s.add_sphere((-8.0, -6.0, 6.0), 2.25, scenes.material(diffuse=(0.2, 0.2, 0.9), ambient=(1, 1, 1), mirror=(0.8, 0.8, 0.8), is_mirror=1))
    s.add_light((30.0, 40.0, -20.0), 4e4); s.ambient[:] = (12, 12, 12)
    # looking along an oblique direction so that no ray direction component is zero, from far away
    fwd = np.array((0.36, -0.48, 0.8), np.float64); fwd /= np.linalg.norm(fwd)
    right = np.cross((0, 1, 0), fwd); right /= np.linalg.norm(right)
    up = np.cross(fwd, right)
    pos = np.array((1.0, -2.0, 3.0)) - fwd * distance
    cam = dict(pos=tuple(np.float32(pos)), fwd=tuple(np.float32(fwd)), right=tuple(np.float32(right)), up=tuple(np.float32(up)))
    # the plane at distance `distance` spans ~ +-16 world units: the scene fills the frame
    p = scenes.make_params(64, 48, float(distance), 16.0, 12.0, background=(0.1, 0.1, 0.2), bounces=2, cam=cam)
    return s, p


@pytest.mark.parametrize("builder", [0, 1])
@pytest.mark.parametrize("distance", [2e3, 2e4, 2e5, 3e6])
def test_far_camera_keeps_the_scan_result(emu, oracle, scenes, builder, distance):
    """A camera 1e2 .. 1e5 scene extents away: the BVH walk must still return what the reference's scan returns
    (rt_origin_slack, rt_core.h); primary rays of such a view take the per-lane loop on the GPU (RtView::far_origin)."""
    s, p = far_camera_case(scenes, distance)
    e = emu.EmuTracer(leaf_max=1, builder=builder); e.set_scene(s)
    ref = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    got = e.render(p)
    _strict(compare_renders(ref, got))
    assert (ref["hit_id"][:, 0] != 0).sum() > 100          # the fixture does see geometry


def grazing_sphere_rays(rng, centers, radii, origin_box, n_per=64):
    """normalised binary32 rays from random origins that pass each sphere centre at 0.98 .. 1.3 radii — where the reference's
    cancelling discriminant (RMath.cs:88-93) decides differently from the exact geometry"""
    O, D = [], []
    for c, r in zip(centers, radii):
        o = rng.uniform(origin_box[0], origin_box[1], (n_per, 3))
        v = c - o
        vh = v / np.linalg.norm(v, axis=1, keepdims=True)
        a = rng.normal(size=(n_per, 3)); a -= (a * vh).sum(1, keepdims=True) * vh; a /= np.linalg.norm(a, axis=1, keepdims=True)
        tgt = c + a * (r * rng.uniform(0.98, 1.3, (n_per, 1)))
        d = (tgt - o).astype(np.float32)
        O.append(o.astype(np.float32)); D.append(d)
    O, D = np.concatenate(O), np.concatenate(D)
    dd = (D[:, 0] * D[:, 0] + D[:, 1] * D[:, 1]) + D[:, 2] * D[:, 2]                         # math.normalize = rsqrt(dot) * x
    D = (np.float32(1) / np.sqrt(dd))[:, None] * D
    return O, D.astype(np.float32)


@pytest.mark.parametrize("builder", [0, 1])
def test_sphere_silhouettes_match_the_scan(emu, oracle, scenes, builder):
    """Small spheres seen from far away: rays that geometrically MISS a sphere by up to ~20 % of its radius are hits for the
    reference's binary32 discriminant.  The sphere boxes are inflated for exactly that (rt_sphere_box_radius), otherwise the
    BVH would cull what the scan hits.  Explicit rays (rt_intersect's path) from inside and far outside the scene, and the
    accelerated oracle, against the faithful scan."""
    rng = np.random.default_rng(11)
    s = scenes.SceneArrays()
    m = scenes.material(diffuse=(0.7, 0.7, 0.7), ambient=(1, 1, 1))
    n_s = 400
    centers = rng.uniform(-40, 40, (n_s, 3)) + np.array((0, 0, 100.0))
    radii = rng.uniform(0.1, 0.4, n_s)
    for c, r in zip(centers, radii):
        s.add_sphere(tuple(np.float32(c)), float(np.float32(r * r)), m)
    t = np.array((-60, -45, 160, 60, -45, 160, 0, 60, 160), np.float32)
    s.add_triangle(t, scenes.triangle_normal(t)[0], m)
    s.add_light((0, 80, 0), 1e5)
    centers = s.spheres[:, :3].astype(np.float64); radii = np.sqrt(s.spheres[:, 3].astype(np.float64))
    ref_scene = oracle.OracleScene(s)
    e = emu.EmuTracer(leaf_max=2, builder=builder); e.set_scene(s)
    n_diff_geom = 0
    for box in (((-1, -1, -1), (1, 1, 1)), ((-40, -40, 60), (40, 40, 140)), ((-900, -900, -2000), (900, 900, -1000))):
        O, D = grazing_sphere_rays(rng, centers, radii, np.asarray(box, np.float64), 48)
        rid, rt_ = ref_scene.intersect(O, D, oracle.FAITHFUL)
        gid, gt = e.intersect(O, D)
        assert np.array_equal(rid[:, :2], gid[:, :2]) and np.array_equal(rt_, gt), (box, int((rid[:, :2] != gid[:, :2]).any(axis=1).sum()))
        aid, at = ref_scene.intersect(O, D, oracle.ACCELERATED)
        assert np.array_equal(rid[:, :2], aid[:, :2]) and np.array_equal(rt_, at), box
        # how many of the reference's hits are geometric misses by more than the node padding?
        k = np.repeat(np.arange(n_s), 48)
        oc = O.astype(np.float64) - centers[k]
        u = (D.astype(np.float64) * oc).sum(1)
        perp = np.sqrt(np.maximum((oc ** 2).sum(1) - u * u, 0))
        n_diff_geom += int(((rid[:, 0] == 1) & (rid[:, 1] == k) & (perp > radii[k] + 0.01)).sum())
    assert n_diff_geom > 50, n_diff_geom          # the fixture does exercise the inflated region

"""-m "not gpu": pins the CPU oracle (oracle/rt_oracle.cpp) to every known answer available for this path.

The reference's own tests pin ONE value (Tests/RayTracerTests.cs:8-27, KAT-1).  KAT-2..4 are hand-derived from the cited
formulas; tests/golden/*.npz are outputs of the oracle itself frozen by tests/golden/make_golden.py so
that later edits of the oracle cannot silently drift.  The reference is C#/Unity and cannot execute here: parity for
everything but KAT-1 is "unpinned by the reference" (DESIGN.md §4).
"""
import os

import numpy as np
import pytest

F = np.float32
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_kat1_reference_unit_test(oracle):
    """Tests/RayTracerTests.cs:11-26: O=(0,0,0) D=(1,0,0), sphere C=(300,0,0) r2=1 -> hit, t = 299 exactly."""
    hit, t = oracle.ray_sphere((0, 0, 0), (1, 0, 0), (300, 0, 0, 1))
    assert hit and t == F(299.0)


def test_kat2_centre_ray_of_rt_unity(oracle, scenes):
    s = scenes.rt_unity()
    ora = oracle.OracleScene(s)
    ids, t = ora.intersect([(0, 0, 0)], [(0, 0, 1)])
    assert ids[0, 0] == 1 and ids[0, 1] == 0 and t[0] == F(19.6)
    rgb = ora.shade_ray(scenes.rt_unity_params(), (0, 0, 0), (0, 0, 1))
    # hand-derived: E = 1e5/417.6841 = 239.41537, cos = 0.95902985 -> 15 + 229.606 = 244.60649
    assert abs(float(rgb[0]) - 244.60649) < 2e-4 and rgb[1] == F(15) and rgb[2] == F(15)


def test_kat3_pixel_25_25(oracle, scenes):
    o, d, pp = oracle.primary_ray(scenes.rt_unity_params(), 25, 25)
    assert np.allclose(pp, (0.39999962, -0.19999981, 10.0), atol=1e-6)
    assert np.allclose(d, (0.03996002, -0.01998001, 0.99900156), atol=1e-7)
    assert (o == 0).all()


def test_kat4_triangle(oracle, scenes):
    s = scenes.rt_unity()
    hit, t = oracle.ray_triangle((17.1, 0, 0), (0, 0, 1), s.tris[0])
    assert hit and t == F(15.0)
    assert np.array_equal(oracle.triangle_normal(s.tris[0]), np.array((0, 0, -1), F))


def test_slab_test_semantics(oracle):
    """RMath.cs:12-26: branch-free slab test, NaN from 0*inf neutralised by math.min/max's first-arg rule."""
    box = (-1, -1, -1, 1, 1, 1)
    assert oracle.ray_aabb((0, 0, -5), (0, 0, 1), box)
    assert not oracle.ray_aabb((0, 0, -5), (0, 0, -1), box)          # box behind the ray
    assert oracle.ray_aabb((0, 0, 0), (1, 0, 0), box)                # origin inside
    assert not oracle.ray_aabb((2, 0, -5), (0, 0, 1), box)           # parallel to x slabs, outside them
    assert oracle.ray_aabb((1, 0, -5), (0, 0, 1), box)               # origin on a slab plane: 0*inf = NaN neutralised


def test_triangle_epsilon_rules(oracle):
    """RMath.cs:39 |a| < 1e-5 -> miss (absolute); RMath.cs:66 t must exceed 1e-5; two-sided."""
    tri = (0, 0, 5, 1, 0, 5, 0, 1, 5)
    assert oracle.ray_triangle((0.2, 0.2, 0), (0, 0, 1), tri)[0]
    assert oracle.ray_triangle((0.2, 0.2, 10), (0, 0, -1), tri)[0]                    # back face hits too
    assert not oracle.ray_triangle((0.2, 0.2, 5), (0, 0, 1), tri)[0]                  # t = 0 <= eps
    tiny = (0, 0, 5, 1e-3, 0, 5, 0, 1e-3, 5)                                          # 2*area = 1e-6 < eps -> "parallel"
    assert not oracle.ray_triangle((2e-4, 2e-4, 0), (0, 0, 1), tiny)[0]
    assert not oracle.ray_triangle((0.8, 0.8, 0), (0, 0, 1), tri)[0]                  # u + v > 1


def test_sphere_inside_and_behind(oracle):
    """RMath.cs:81-108: from inside returns the far root; no epsilon; sphere behind the origin misses."""
    hit, t = oracle.ray_sphere((0, 0, 0), (0, 0, 1), (0, 0, 0, 4))
    assert hit and t == F(2.0)
    assert not oracle.ray_sphere((0, 0, 0), (0, 0, 1), (0, 0, -10, 4))[0]


def test_miss_result(oracle, scenes):
    """Scene.cs:45-61: miss = {None(0), -1, -1, float.MaxValue}."""
    ora = oracle.OracleScene(scenes.rt_unity())
    ids, t = ora.intersect([(0, 0, 0)], [(0, 0, -1)])
    assert ids[0].tolist() == [0, -1, -1] and t[0] == np.finfo(F).max


def test_scan_order_tie_break(oracle, scenes):
    """Scene.cs:64-115: on exactly equal t the earlier-tested primitive wins: mesh tris < spheres < loose tris."""
    s = scenes.SceneArrays()
    tri = np.array((-1, -1, 5, 1, -1, 5, 0, 1, 5), F)
    n = scenes.triangle_normal(tri)[0]
    s.add_triangle(tri, n, scenes.material())
    s.add_triangle(tri, n, scenes.material())
    ora = oracle.OracleScene(s)
    assert ora.intersect([(0, 0, 0)], [(0, 0, 1)])[0][0, :2].tolist() == [2, 0]
    for mode in (oracle.FAITHFUL, oracle.ACCELERATED):
        assert ora.intersect([(0, 0, 0)], [(0, 0, 1)], mode)[0][0, :2].tolist() == [2, 0]
    nm, aabb = scenes.mesh_from_world_tris(tri.reshape(1, 9))
    s.add_mesh(tri.reshape(1, 9), nm, scenes.material(), aabb)
    ora = oracle.OracleScene(s)
    for mode in (oracle.FAITHFUL, oracle.ACCELERATED):
        assert ora.intersect([(0, 0, 0)], [(0, 0, 1)], mode)[0][0].tolist() == [3, 0, 0]


def test_scene_aabb(oracle, scenes):
    """Scene.CalculateAABB, Scene.cs:17-41 == min/max over mesh AABBs, loose vertices, sphere boxes."""
    s = scenes.rt_unity()
    bb = oracle.OracleScene(s).aabb()
    lo = np.minimum.reduce([s.meshes[0][3][:3], s.tris.reshape(-1, 3).min(0), s.spheres[0, :3] - F(10)])
    hi = np.maximum.reduce([s.meshes[0][3][3:], s.tris.reshape(-1, 3).max(0), s.spheres[0, :3] + F(10)])
    assert np.array_equal(bb, np.concatenate([lo, hi]).astype(F))


def test_specular_cutoff_constant(oracle):
    """rt_hd.h RT_SPEC_CUTOFF: largest float x with degrees(acos(x)) > 90 (RayTracingSetup.cs:384-391)."""
    assert oracle.specular_cutoff() == F(float.fromhex("-0x1.bbbd3p-24"))


@pytest.mark.parametrize("seed", [7, 3])
def test_accelerated_equals_faithful(oracle, scenes, seed):
    s, p = scenes.random_soup(seed=seed), scenes.soup_params(64, 48, 3, 1)
    ora = oracle.OracleScene(s)
    a, b = ora.render(p, oracle.FAITHFUL), ora.render(p, oracle.ACCELERATED)
    for k in ("rgba8", "hit_t"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["hit_id"][:, :2], b["hit_id"][:, :2])
    assert np.array_equal(a["rgba32f"], b["rgba32f"], equal_nan=True)
    assert a["stats"] == b["stats"]


def test_ray_budget_cfg1(oracle, scenes):
    """2 500 primary rays; one shadow ray per (hit, light); no mirror rays in the shipped scene (D7)."""
    r = oracle.OracleScene(scenes.rt_unity()).render(scenes.rt_unity_params(), oracle.FAITHFUL)
    pr, sh, mi = r["stats"]
    assert pr == 2500 and mi == 0 and sh == int((r["hit_id"][:, 0] != 0).sum())


def test_inline_invariants(oracle, scenes):
    """The reference's Debug.Asserts as properties: primary directions are unit length (RTS:306, RMath.cs:115-123)."""
    p = scenes.rt_unity_params()
    for x, y in ((0, 0), (49, 49), (25, 0), (3, 44)):
        _, d, _ = oracle.primary_ray(p, x, y)
        assert abs(float(d[0]) ** 2 + float(d[1]) ** 2 + float(d[2]) ** 2 - 1.0) < 1e-5


@pytest.mark.parametrize("name", ["cfg1_rt_unity_50x50", "cfg2_mirror_96x54_d4", "soup7_64x48_d3", "soup3_48x32_d2_spp2", "degenerate_nan_64x40_d2",
                                  "soup4_oblique_camera_56x40_d2", "soup4_sheared_camera_56x40_d2", "soup13_40_lights_48x32_d2", "mirror_walls_48x32_d64"])
def test_golden_fixtures(oracle, scenes, name):
    """Frozen oracle outputs (tests/golden/make_golden.py).  u8/ids/t bit-exact; float colour bit-exact too (same libm)."""
    from golden import make_golden as G
    s, p = G.CASES[name](scenes)
    g = np.load(os.path.join(GOLD, name + ".npz"))
    r = oracle.OracleScene(s).render(p, oracle.FAITHFUL)
    assert np.array_equal(r["hit_id"][:, :2], g["hit_id"][:, :2])
    assert np.array_equal(r["hit_t"], g["hit_t"])
    assert np.abs(r["rgba8"].astype(int) - g["rgba8"].astype(int)).max() <= 1
    assert tuple(int(v) for v in g["stats"]) == r["stats"]

"""-m "not gpu": a second opinion on the oracle ("next" item 2).

Both sides of every GPU parity test descend from oracle/rt_oracle.cpp, and the reference itself holds exactly one number
(Tests/RayTracerTests.cs:8-27, t = 299).  Here the oracle meets two things that were NOT derived from it:
  * tests/csharp_restatement.py — the hot path restated a second time, straight from the C#, scalar numpy float32, no BVH;
  * tests/kats.py — hand-derived known answers evaluated in exact rational arithmetic rounded per operation.
Which function is pinned by which vector: DESIGN.md §4.
"""
import numpy as np
import pytest

import csharp_restatement as R
import exact32 as X
import kats
from conftest import degenerate_scene


def _same(ref, got, what):
    """ids (MeshIndex only for mesh triangles), t and float colour, bit for bit (NaN == NaN)"""
    rid, gid = ref["hit_id"], got["hit_id"]
    bad = (rid[:, :2] != gid[:, :2]).any(axis=1) | ((rid[:, 0] == 3) & (rid[:, 2] != gid[:, 2]))
    assert not bad.any(), (what, "ids", int(bad.sum()))
    assert np.array_equal(ref["hit_t"], got["hit_t"], equal_nan=True), (what, "t", int((ref["hit_t"] != got["hit_t"]).sum()))
    a, b = ref["rgba32f"], got["rgba32f"]
    assert np.array_equal(a.view(np.uint32) | (np.isnan(a) * 0xffffffff).astype(np.uint32), b.view(np.uint32) | (np.isnan(b) * 0xffffffff).astype(np.uint32)), \
        (what, "rgba32f", int(((a != b) & ~(np.isnan(a) & np.isnan(b))).any(axis=1).sum()))
    assert tuple(ref["stats"]) == tuple(got["stats"]), (what, ref["stats"], got["stats"])


def test_restatement_vs_oracle_cfg1_every_pixel(oracle, scenes):
    """RayTracing.unity as shipped, all 2 500 pixels: ray generation, slab gates, triangle + sphere tests, scan order, shading"""
    s, p = scenes.rt_unity(), scenes.rt_unity_params()
    _same(R.render(s, p), oracle.OracleScene(s).render(p, oracle.FAITHFUL), "cfg1")


def test_restatement_vs_oracle_cfg2_mirror_depth4(oracle, scenes):
    """the mirror sphere restored (cfg2), 96x54, depth 4: Blinn-Phong, Reflect, recursion and its accumulation order.
    numpy's float64 power and math.acos are glibc's pow/acos, the very functions the oracle calls, so even the float colour
    must agree bit for bit."""
    s, p = scenes.rt_unity(mirror_sphere=True), scenes.rt_unity_params(96, 54, 4)
    a = R.render(s, p)
    assert a["stats"][2] > 0
    _same(a, oracle.OracleScene(s).render(p, oracle.FAITHFUL), "cfg2")


def test_restatement_vs_oracle_degenerate_and_nan_geometry(oracle, scenes):
    """zero-area / repeated / NaN / infinite primitives, r2 = 0, camera inside a sphere: NaN handling of math.min/max, the
    strict '>' tie-break, `disc < 0` with NaN"""
    s = degenerate_scene(scenes, nan=True)
    p = scenes.soup_params(40, 26, 2, 1)
    rng = np.random.default_rng(3)
    px = np.sort(rng.choice(40 * 26, 260, replace=False))
    _same(R.render(s, p, px), oracle.OracleScene(s, build_bvh=False).render(p, oracle.FAITHFUL, pixels=px), "degenerate")


def test_restatement_vs_oracle_soup_with_meshes_and_mirrors(oracle, scenes):
    """meshes (per-mesh AABB gate, one material per mesh, negated normals), loose triangles, spheres, mirrors, 2 lights"""
    s = scenes.random_soup(seed=11, n_tris=60, n_spheres=12, n_meshes=2)
    p = scenes.soup_params(36, 24, 3, 1)
    _same(R.render(s, p), oracle.OracleScene(s, build_bvh=False).render(p, oracle.FAITHFUL), "soup")
    cams = __import__("conftest").camera_set()
    q = dict(p); q.update(cams[-1])                              # the sheared, non-unit camera basis
    px = np.arange(0, 36 * 24, 7)
    _same(R.render(s, q, px), oracle.OracleScene(s, build_bvh=False).render(q, oracle.FAITHFUL, pixels=px), "soup, sheared camera")


@pytest.mark.parametrize("kat", [kats.kat5_blinn_phong, kats.kat6_mirror_depth1])
def test_hand_derived_shading_kats(oracle, scenes, kat):
    """KAT-5 (Blinn-Phong specular, RTS:382-399) and KAT-6 (depth-1 mirror term and its accumulation, RTS:358-362): exact
    rational evaluation of the cited formulas vs the oracle AND the numpy restatement, bit for bit."""
    k = kat()
    want = np.array([float(c) / 255.0 for c in k["value"]], np.float64)
    for name, got in (("oracle", oracle.OracleScene(k["scene"]).render(k["params"], oracle.FAITHFUL)), ("restatement", R.render(k["scene"], k["params"]))):
        assert tuple(got["hit_id"][0][:2]) == k["hit"], name
        assert got["hit_t"][0] == k["t"], (name, got["hit_t"][0], k["t"])
        expect = np.array([X.div(c, np.float32(255)) for c in k["value"]], np.float32)       # Rgb.Color = Value / 255f, Rgb.cs:13
        assert np.array_equal(got["rgba32f"][0][:3], expect), (name, got["rgba32f"][0], expect, want)
        assert got["rgba32f"][0][3] == 1.0
        assert tuple(got["stats"]) == (1, k["shadow_rays"], k["mirror_rays"]), name


def test_kat7_slab_test_with_zero_direction_on_a_slab_plane(oracle):
    for o, d, box, want in kats.kat7_slab_zero_direction():
        assert oracle.ray_aabb(o, d, box) is want, (o, d)
        assert R.ray_aabb(R.f3(o), R.f3(d), R.f3(box[:3]), R.f3(box[3:])) is want, (o, d)


def test_kat8_normalize_is_rsqrt_times_x(oracle):
    import ctypes as C
    for v, rsqrt_form, div_form in kats.kat8_normalize_is_rsqrt_times_x():
        out = np.zeros(3, np.float32)
        vin = np.array(v, np.float32)
        oracle.lib().ora_normalize(vin.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_float)))
        assert tuple(out) == tuple(rsqrt_form) and tuple(out) != tuple(div_form), (v, out, rsqrt_form, div_form)
        assert tuple(R.normalize(R.f3(v))) == tuple(rsqrt_form)


def test_exact32_rounding_agrees_with_the_fpu():
    """the hand-derivation helper itself: exact-rational round-to-nearest-even == numpy float32 on random operands"""
    rng = np.random.default_rng(5)
    a = (rng.standard_normal(400) * 10.0 ** rng.integers(-20, 20, 400)).astype(np.float32)
    b = (rng.standard_normal(400) * 10.0 ** rng.integers(-20, 20, 400)).astype(np.float32)
    with np.errstate(all="ignore"):
        for x, y in zip(a, b):
            assert X.add(x, y) == x + y and X.mul(x, y) == x * y and X.div(x, y) == x / y and X.sub(x, y) == x - y, (x, y)
            assert X.sqrt(abs(x)) == np.float32(np.sqrt(np.float64(abs(x))))

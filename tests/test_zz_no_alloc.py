"""rt_intersect_static — the allocation-free form of rt_intersect (include/rtb200.h; no counterpart in the reference, added
for callers that cannot call cudaMalloc).  Collected last on purpose.  Through the C-ABI:
its output must equal rt_intersect's bit for bit (ids and distances are integer / exact-arithmetic results: tolerance 0, as
in test_gpu_parity.py), on two calls in a row — the second with other rays and another count, since the static scratch
array keeps the first call's contents."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rays(n, seed, spread):
    rng = np.random.default_rng(seed)
    o = rng.uniform(-spread, spread, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d.astype(np.float32)


def _same(a, b):
    (ia, ta), (ib, tb) = a, b
    return np.array_equal(ia, ib) and np.array_equal(ta.view(np.uint32), tb.view(np.uint32))


def test_intersect_static_equals_intersect_twice_in_a_row(rtb, scenes, oracle):
    r = rtb.RayTracer((0,))
    try:
        r.set_scene(scenes.random_soup(seed=11))           # spheres + loose triangles + meshes
        # call 1: a full-size batch, origins inside and well outside the scene (outside the committed origin box the
        # static form takes the linear scan instead of re-fitting the sphere boxes, which would allocate)
        o1, d1 = _rays(20000, 5, 40.0)
        o1[::7] *= 6.0
        got1 = r.intersect_static(o1, d1)
        # call 2, straight after: fewer rays, other values — stale scratch contents beyond n must not leak into the result
        o2, d2 = _rays(777, 6, 30.0)
        got2 = r.intersect_static(o2, d2)
        want1, want2 = r.intersect(o1, d1), r.intersect(o2, d2)
        assert _same(got1, want1), "rt_intersect_static differs from rt_intersect (first call)"
        assert _same(got2, want2), "rt_intersect_static differs from rt_intersect (second call)"
        assert _same(r.intersect_static(o1, d1), want1)    # and again after rt_intersect grew the origin box
        assert (want1[0][:, 0] != 0).sum() > 300 and (want1[0][:, 0] == 0).sum() > 300    # hits and misses both exercised
        # and against the reference's algorithm itself (the oracle's faithful scan) on a sample
        rid, rt_ = oracle.OracleScene(scenes.random_soup(seed=11)).intersect(o2, d2, oracle.FAITHFUL)
        assert (rid[:, :2] == got2[0][:, :2]).all() and (rt_ == got2[1]).all()
        # the limit is a clear error, not a truncation; zero rays are a no-op
        big_o, big_d = _rays(r.INTERSECT_STATIC_MAX + 1, 7, 10.0)
        with pytest.raises(rtb.host.RtError) as ei:
            r.intersect_static(big_o, big_d)
        assert ei.value.code == -1 and "RT_INTERSECT_STATIC_MAX" in str(ei.value)
        full_o, full_d = big_o[:-1], big_d[:-1]
        assert _same(r.intersect_static(full_o, full_d), r.intersect(full_o, full_d))        # exactly the maximum
        ids0, t0 = r.intersect_static(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
        assert ids0.shape == (0, 3) and t0.shape == (0,)
    finally:
        r.close()

import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def rtb():
    return importlib.import_module("unity-raytracer_b200")


CHECK_KINDS = ["stack overflow", "node index", "leaf slot range", "TLAS leaf range", "tile index", "queue capacity", "virtual slot decode",
               "pixel index", "hit slot", "mesh index", "light index", "bake leaf size"]
_check_totals = {"tests": 0, "violations": 0}


@pytest.fixture(autouse=True)
def _self_checking_build(request, rtb):
    """With RTB200_LIB pointing at a -DRT_CHECKED build (tools/checked_run.sh; compute-sanitizer is closed on this GPU pool) every
    GPU test also asserts that none of the library's index checks fired while it ran."""
    yield
    if not os.environ.get("RTB200_LIB") or "gpu" not in request.keywords:
        return
    import ctypes as C
    L = rtb.load_library()
    if not hasattr(L, "rt_debug_checks"):
        return
    out = (C.c_uint * 16)()
    assert L.rt_debug_checks(out, 1) == 0
    bad = {CHECK_KINDS[i] if i < len(CHECK_KINDS) else str(i): int(v) for i, v in enumerate(out) if v}
    _check_totals["tests"] += 1
    _check_totals["violations"] += sum(bad.values())
    assert not bad, f"RT_CHECKED build: index checks fired: {bad}"


def pytest_terminal_summary(terminalreporter):
    if os.environ.get("RTB200_LIB") and _check_totals["tests"]:
        terminalreporter.write_line(f"RT_CHECKED build: {_check_totals['tests']} GPU tests ran with {len(CHECK_KINDS)} kinds of index checks compiled in; "
                                    f"violations counted: {_check_totals['violations']}")


@pytest.fixture(scope="session")
def scenes(rtb):
    return rtb.scenes


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def emu():
    from emu import emu as E
    E.lib()
    return E


def compare_renders(ref, got, check_float=True):
    """Parity rule of BASELINE.json north_star: ids/t bit-exact, u8 within +-1 LSB on >= 99.9 % of pixels.
    Returns a dict of mismatch counts; MeshIndex is compared only for MeshTriangle hits (Scene.cs:96-97 keeps a stale
    value for the other types)."""
    rid, gid = ref["hit_id"], got["hit_id"]
    idm = (rid[:, :2] != gid[:, :2]).any(axis=1)
    idm |= (rid[:, 0] == 3) & (rid[:, 2] != gid[:, 2])
    tm = ref["hit_t"] != got["hit_t"]
    d8 = np.abs(ref["rgba8"].astype(np.int32) - got["rgba8"].astype(np.int32)).max(axis=1)
    out = dict(n=len(idm), id=int(idm.sum()), t=int(tm.sum()), u8_gt0=int((d8 > 0).sum()), u8_gt1=int((d8 > 1).sum()))
    if check_float:
        a, b = ref["rgba32f"], got["rgba32f"]
        out["f32"] = int(((a != b) & ~(np.isnan(a) & np.isnan(b))).any(axis=1).sum())
        out["f32_far"] = float_far(a, b)
    return out


def float_far(a, b, ulps=8):
    """pixels whose float colour differs by MORE than a last-bits difference of one term (|a - b| > ulps * 2^-24 * max|.|):
    the only legitimate source of a float mismatch is the Blinn-Phong pow() (CUDA's double pow vs glibc's, <= 1 ulp of a
    double apart, which can flip the rounding of (float)pow and then moves the pixel by an ulp or two of the sum)"""
    a, b = a.astype(np.float64), b.astype(np.float64)
    with np.errstate(invalid="ignore"):
        tol = ulps * 2.0 ** -24 * np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-30)
        bad = (np.abs(a - b) > tol) & ~(np.isnan(a) & np.isnan(b)) & ~((a == b))
        bad |= np.isnan(a) != np.isnan(b)
    return int(bad.any(axis=1).sum())


def degenerate_scene(scenes, nan=False):
    """A soup plus the inputs the reference's scan takes without complaint: zero-area triangles, repeated primitives, a
    sphere with r2 = 0, the camera inside a huge sphere, a triangle edge-on to the centre ray, coordinates near 1e19 and
    (nan=True) NaN / infinite vertices and centres — whatever Scene.cs:43-122 answers for them is the required answer."""
    s = scenes.random_soup(seed=21, n_tris=300, n_spheres=40)
    m = scenes.material()
    flat = np.array((0, 0, -1), np.float32)
    for tri in ((1, 1, 12, 1, 1, 12, 1, 1, 12), (-5, 0, 15, 0, 0, 15, 5, 0, 15)):          # a point, a collinear triangle
        s.add_triangle(np.array(tri, np.float32), flat, m)
    t = np.array((-3, -3, 18, 3, -3, 18, 0, 3, 18), np.float32)
    for _ in range(5):
        s.add_triangle(t, scenes.triangle_normal(t)[0], m)
    for _ in range(3):
        s.add_sphere((4, 4, 22), 9.0, m)
    s.add_sphere((0, 0, 10), 0.0, m)
    s.add_sphere((0, 0, 0), 1e6, m)
    t = np.array((0, -2, 5, 0, 2, 5, 0, 0, 30), np.float32)
    s.add_triangle(t, scenes.triangle_normal(t)[0], m)
    s.add_triangle(np.array((-1e18, -1e18, 1e19, 1e18, -1e18, 1e19, 0, 1e18, 1e19), np.float32), flat, m)
    if nan:
        s.add_triangle(np.array((np.nan, 0, 9, 1, 0, 9, 0, 1, 9), np.float32), flat, m)
        s.add_triangle(np.array((np.inf, 0, 9, 1, 0, 9, 0, 1, 9), np.float32), flat, m)
        s.add_sphere((np.nan, 0, 9), 1.0, m)
    return s


def camera_set():
    """CameraData as Unity hands it over (transform.position / forward / right / up, CameraData.cs:5-10): looking down every
    axis, an oblique rotation off the origin, and a basis that is neither unit length nor orthogonal (the reference uses the
    vectors as given, RayTracingSetup.cs:282-300)."""
    def rot(axis, deg):
        a = np.asarray(axis, np.float64); a /= np.linalg.norm(a)
        c, s = np.cos(np.radians(deg)), np.sin(np.radians(deg))
        K = np.array(((0, -a[2], a[1]), (a[2], 0, -a[0]), (-a[1], a[0], 0)))
        return np.eye(3) * c + s * K + (1 - c) * np.outer(a, a)
    cams = []
    for R, pos in ((rot((0, 1, 0), 180), (0, 0, 60)), (rot((0, 1, 0), 90), (-60, 0, 20)), (rot((0, 1, 0), -90), (60, 0, 20)),
                   (rot((1, 0, 0), 90), (0, 60, 20)), (rot((1, 0, 0), -90), (0, -60, 20)), (rot((1, 2, 3), 37), (7, -11, -25))):
        R = R.astype(np.float32)
        cams.append(dict(pos=tuple(np.float32(x) for x in pos), right=tuple(R[:, 0]), up=tuple(R[:, 1]), fwd=tuple(R[:, 2])))
    cams.append(dict(pos=(3.0, -2.0, -30.0), fwd=(0.1, 0.05, 1.7), right=(0.8, 0.1, 0.0), up=(0.2, 1.3, -0.1)))
    return cams


def many_lights_scene(scenes, n_lights=40):
    """a soup lit by n_lights point lights (RayTracingSetup.cs:333-355 loops over all of them)"""
    s = scenes.random_soup(seed=13, n_tris=600, n_spheres=60)
    rng = np.random.default_rng(2)
    for _ in range(n_lights - s.lights.shape[0]):
        s.add_light(tuple(rng.uniform(-45, 45, 3)), float(rng.uniform(2e3, 2e4)))
    return s


def mirror_walls_scene(scenes):
    """two parallel 8000 x 8000 mirror walls around the camera: reflection chains (RayTracingSetup.cs:357-372) run to the limit"""
    s = scenes.SceneArrays()
    mir = scenes.material(diffuse=(0.2, 0.1, 0.1), ambient=(0.1, 0.1, 0.1), mirror=(0.97, 0.95, 0.9), specular=(0.3, 0.3, 0.3), phong=8.0, is_mirror=1)
    for z, side in ((20.0, 1.0), (-20.0, -1.0)):
        for tri in ((-4e3, -4e3, z, 4e3, -4e3, z, 4e3, 4e3, z), (-4e3, -4e3, z, 4e3, 4e3, z, -4e3, 4e3, z)):
            s.add_triangle(np.array(tri, np.float32), np.array((0, 0, -side), np.float32), mir)
    s.add_sphere((6, 2, 4), 2.0, scenes.material(diffuse=(0.8, 0.3, 0.2), ambient=(0.2, 0.2, 0.2), specular=(0.5, 0.5, 0.5), phong=20.0))
    t = np.array((-9, -6, 2, -3, -6, 5, -6, 4, 3), np.float32)
    s.add_triangle(t, scenes.triangle_normal(t)[0], scenes.material(diffuse=(0.2, 0.7, 0.3), ambient=(0.1, 0.1, 0.1)))
    s.add_light((0, 8, 0), 3e3); s.add_light((5, -7, 1), 2e3); s.ambient[:] = (30, 30, 30)
    return s

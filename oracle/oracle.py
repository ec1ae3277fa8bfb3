"""ctypes loader for the CPU oracle (oracle/rt_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.

The argument layouts are the same flat arrays the product C-ABI (include/rtb200.h) takes, so one
`SceneArrays` description (unity-raytracer_b200/scenes.py) feeds both sides identically.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.environ.get("RT_ORACLE_LIB") or os.path.join(_HERE, "librt_oracle.so")     # override: sanitizer builds (tools/sanitize_host.sh)

FAITHFUL, ACCELERATED = 0, 1


class Params(C.Structure):
    """Same layout as rt_render_params (include/rtb200.h)."""
    _fields_ = [("pos", C.c_float * 3), ("fwd", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("res_x", C.c_int), ("res_y", C.c_int), ("dist", C.c_float), ("half_h", C.c_float), ("half_v", C.c_float),
                ("background_rgb", C.c_float * 3), ("max_reflection_bounces", C.c_int), ("spp_grid", C.c_int)]


def build(force=False):
    src = os.path.join(_HERE, "rt_oracle.cpp")
    if os.environ.get("RT_ORACLE_LIB"):
        return _LIB
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
        L.ora_scene_new.restype = C.c_void_p
        L.ora_scene_free.argtypes = [C.c_void_p]
        L.ora_scene_spheres.argtypes = [C.c_void_p, fp, fp, C.c_int]
        L.ora_scene_triangles.argtypes = [C.c_void_p, fp, fp, fp, C.c_int]
        L.ora_scene_mesh.argtypes = [C.c_void_p, fp, fp, C.c_int, fp, fp]
        L.ora_scene_lights.argtypes = [C.c_void_p, fp, C.c_int, fp]
        L.ora_scene_commit.argtypes = [C.c_void_p, fp, C.c_int]
        L.ora_scene_get_aabb.argtypes = [C.c_void_p, fp]
        L.ora_ray_aabb.argtypes = [fp, fp, fp]
        L.ora_ray_triangle.argtypes = [fp, fp, fp, fp]
        L.ora_ray_sphere.argtypes = [fp, fp, fp, fp]
        L.ora_triangle_normal.argtypes = [fp, fp]
        L.ora_normalize.argtypes = [fp, fp]
        L.ora_intersect.argtypes = [C.c_void_p, fp, fp, C.c_int, C.c_int, ip, fp]
        L.ora_shade_ray.argtypes = [C.c_void_p, C.POINTER(Params), fp, fp, C.c_int, fp]
        L.ora_primary_ray.argtypes = [C.POINTER(Params), C.c_int, C.c_int, C.c_int, C.c_int, fp, fp, fp]
        L.ora_render.argtypes = [C.c_void_p, C.POINTER(Params), C.c_int, C.c_int, C.POINTER(C.c_int64), C.c_int64,
                                 fp, C.POINTER(C.c_uint8), ip, fp, C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        L.ora_specular_cutoff.restype = C.c_float
        _lib = L
    return _lib


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


def make_params(p):
    """p: dict with the rt_render_params fields."""
    q = Params()
    for k in ("pos", "fwd", "right", "up", "background_rgb"):
        setattr(q, k, (C.c_float * 3)(*[float(v) for v in p[k]]))
    q.res_x, q.res_y = int(p["res_x"]), int(p["res_y"])
    q.dist, q.half_h, q.half_v = float(p["dist"]), float(p["half_h"]), float(p["half_v"])
    q.max_reflection_bounces = int(p.get("max_reflection_bounces", 0))
    q.spp_grid = int(p.get("spp_grid", 1))
    return q


def ray_aabb(o, d, box6):
    return bool(lib().ora_ray_aabb(_f(o)[1], _f(d)[1], _f(box6)[1]))


def ray_triangle(o, d, v9):
    t = C.c_float(0)
    hit = lib().ora_ray_triangle(_f(o)[1], _f(d)[1], _f(v9)[1], C.byref(t))
    return bool(hit), np.float32(t.value)


def ray_sphere(o, d, c_r2):
    t = C.c_float(0)
    hit = lib().ora_ray_sphere(_f(o)[1], _f(d)[1], _f(c_r2)[1], C.byref(t))
    return bool(hit), np.float32(t.value)


def triangle_normal(v9):
    out = np.zeros(3, np.float32)
    lib().ora_triangle_normal(_f(v9)[1], out.ctypes.data_as(C.POINTER(C.c_float)))
    return out


def specular_cutoff():
    return np.float32(lib().ora_specular_cutoff())


class OracleScene:
    """Oracle-side scene built from a SceneArrays-like object (fields: spheres, sphere_mats, tris, tri_normals,
    tri_mats, meshes=[(verts, normals, mat, aabb)], lights, ambient, scene_aabb or None)."""

    def __init__(self, sa, build_bvh=True):
        L = lib()
        self._h = C.c_void_p(L.ora_scene_new())
        self._keep = []
        if sa.spheres.shape[0]:
            L.ora_scene_spheres(self._h, _f(sa.spheres)[1], _f(sa.sphere_mats.view(np.float32))[1], sa.spheres.shape[0])
        if sa.tris.shape[0]:
            L.ora_scene_triangles(self._h, _f(sa.tris)[1], _f(sa.tri_normals)[1], _f(sa.tri_mats.view(np.float32))[1],
                                  sa.tris.shape[0])
        for (v, n, mat, aabb) in sa.meshes:
            L.ora_scene_mesh(self._h, _f(v)[1], _f(n)[1], v.shape[0], _f(mat.view(np.float32))[1], _f(aabb)[1])
        amb = _f(sa.ambient)
        L.ora_scene_lights(self._h, _f(sa.lights)[1], sa.lights.shape[0], amb[1])
        if sa.scene_aabb is not None:
            L.ora_scene_commit(self._h, _f(sa.scene_aabb)[1], int(build_bvh))
        else:
            L.ora_scene_commit(self._h, None, int(build_bvh))

    def __del__(self):
        try:
            lib().ora_scene_free(self._h)
        except Exception:
            pass

    def aabb(self):
        out = np.zeros(6, np.float32)
        lib().ora_scene_get_aabb(self._h, out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def intersect(self, origins, dirs, mode=FAITHFUL):
        o, po = _f(np.atleast_2d(origins))
        d, pd = _f(np.atleast_2d(dirs))
        n = o.shape[0]
        ids = np.zeros((n, 3), np.int32)
        t = np.zeros(n, np.float32)
        rc = lib().ora_intersect(self._h, po, pd, n, mode, ids.ctypes.data_as(C.POINTER(C.c_int32)),
                                 t.ctypes.data_as(C.POINTER(C.c_float)))
        assert rc == 0
        return ids, t

    def shade_ray(self, params, o, d, mode=FAITHFUL):
        q = make_params(params)
        out = np.zeros(3, np.float32)
        rc = lib().ora_shade_ray(self._h, C.byref(q), _f(o)[1], _f(d)[1], mode, out.ctypes.data_as(C.POINTER(C.c_float)))
        assert rc == 0
        return out

    def render(self, params, mode=FAITHFUL, threads=0, pixels=None):
        """Returns dict(rgba32f, rgba8, hit_id, hit_t, stats=(primary,shadow,mirror), seconds)."""
        q = make_params(params)
        if pixels is None:
            n = q.res_x * q.res_y
            pp, np_ = None, 0
        else:
            pixels = np.ascontiguousarray(pixels, dtype=np.int64)
            n = pixels.shape[0]
            pp, np_ = pixels.ctypes.data_as(C.POINTER(C.c_int64)), n
        rgba = np.zeros((n, 4), np.float32)
        u8 = np.zeros((n, 4), np.uint8)
        ids = np.zeros((n, 3), np.int32)
        t = np.zeros(n, np.float32)
        st = (C.c_uint64 * 3)()
        sec = C.c_double(0)
        rc = lib().ora_render(self._h, C.byref(q), mode, threads, pp, np_,
                              rgba.ctypes.data_as(C.POINTER(C.c_float)), u8.ctypes.data_as(C.POINTER(C.c_uint8)),
                              ids.ctypes.data_as(C.POINTER(C.c_int32)), t.ctypes.data_as(C.POINTER(C.c_float)),
                              st, C.byref(sec))
        assert rc == 0, rc
        return dict(rgba32f=rgba, rgba8=u8, hit_id=ids, hit_t=t, stats=tuple(int(x) for x in st), seconds=sec.value)


def primary_ray(params, x, y, i=0, j=0):
    q = make_params(params)
    o = np.zeros(3, np.float32); d = np.zeros(3, np.float32); pp = np.zeros(3, np.float32)
    fp = C.POINTER(C.c_float)
    lib().ora_primary_ray(C.byref(q), x, y, i, j, o.ctypes.data_as(fp), d.ctypes.data_as(fp), pp.ctypes.data_as(fp))
    return o, d, pp

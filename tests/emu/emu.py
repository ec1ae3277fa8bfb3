"""ctypes loader for tests/emu/librt_emu.so — TEST HARNESS ONLY (see rt_emu.cpp): the product's device headers
instantiated serially on the CPU, with the same method names as host.RayTracer so tests can run one body on both."""
import ctypes as C
import importlib
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.environ.get("RT_EMU_LIB") or os.path.join(_HERE, "librt_emu.so")          # override: sanitizer builds (tools/sanitize_host.sh)
host = importlib.import_module("unity-raytracer_b200.host")

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.environ.get("RT_EMU_LIB"):
            subprocess.check_call(["make", "-s", "-C", _HERE])
        L = C.CDLL(_LIB)
        fp, ip, vp = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_void_p
        L.emu_new.restype = vp
        L.emu_free.argtypes = [vp]
        L.emu_scene_spheres.argtypes = [vp, fp, vp, C.c_int]
        L.emu_scene_triangles.argtypes = [vp, fp, fp, vp, C.c_int]
        L.emu_scene_mesh.argtypes = [vp, fp, fp, C.c_int, vp, fp]
        L.emu_scene_lights.argtypes = [vp, fp, C.c_int, fp]
        L.emu_scene_base_mesh.argtypes = [vp, fp, C.c_int]
        L.emu_scene_mesh_instance.argtypes = [vp, C.c_int, fp, vp, fp]
        L.emu_scene_update_instance.argtypes = [vp, C.c_int, fp, fp]
        L.emu_scene_recommit.argtypes = [vp, fp]
        L.emu_scene_commit.argtypes = [vp, fp, C.c_int, C.c_double, C.c_double, C.c_int]
        L.emu_scene_clear.argtypes = [vp]
        L.emu_scene_refit.argtypes = [vp, fp, C.c_int, C.c_double, C.c_double]
        L.emu_set_presplit.argtypes = [vp, C.c_double, C.c_int]
        L.emu_n_slots.argtypes = [vp]
        L.emu_get_aabb.argtypes = [vp, fp]
        L.emu_n_nodes.argtypes = [vp]
        L.emu_max_stack.argtypes = [vp]
        L.emu_ray_triangle.argtypes = [fp, fp, fp, C.c_float, fp]
        L.emu_intersect.argtypes = [vp, fp, fp, C.c_int, ip, fp]
        L.emu_render.argtypes = [vp, C.POINTER(host.RenderParams), C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp,
                                 C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class EmuTracer:
    def __init__(self, leaf_max=4, pad_scale=1.0 / 65536.0, t_slack=1.000125, builder=0):
        self.L = lib()
        self.h = C.c_void_p(self.L.emu_new())
        self.opts = (leaf_max, pad_scale, t_slack, builder)
        self.rank, self.world, self.band_rows, self.chunk = 0, 1, 32, 1 << 23

    def __del__(self):
        try:
            self.L.emu_free(self.h)
        except Exception:
            pass

    def set_scene(self, sa, refit=False):
        L, h = self.L, self.h
        if refit:
            L.emu_scene_clear(h)
        f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        if sa.spheres.shape[0]:
            s, m = f32(sa.spheres), np.ascontiguousarray(sa.sphere_mats)
            L.emu_scene_spheres(h, _fp(s), m.ctypes.data, s.shape[0])
        if sa.tris.shape[0]:
            v, n, m = f32(sa.tris), f32(sa.tri_normals), np.ascontiguousarray(sa.tri_mats)
            L.emu_scene_triangles(h, _fp(v), _fp(n), m.ctypes.data, v.shape[0])
        scenes = host.scenes
        for bv in getattr(sa, "base_meshes", []):
            bv = f32(bv)
            assert L.emu_scene_base_mesh(h, _fp(bv), bv.shape[0]) >= 0
        for m in sa.meshes:
            if isinstance(m, scenes.MeshInstance):                  # (the emulation wants the AABB: compute it as the bake does)
                mm, mat = f32(m.m3x4), np.ascontiguousarray(m.mat)
                bb = f32(m.aabb if m.aabb is not None else scenes.bake_tris(sa.base_meshes[m.base], m.m3x4)[2])
                assert L.emu_scene_mesh_instance(h, int(m.base), _fp(mm), mat.ctypes.data, _fp(bb)) == 0
                continue
            v, n, mat, aabb = m
            v, n, aabb, mat = f32(v), f32(n), f32(aabb), np.ascontiguousarray(mat)
            L.emu_scene_mesh(h, _fp(v), _fp(n), v.shape[0], mat.ctypes.data, _fp(aabb))
        li, amb = f32(sa.lights), f32(sa.ambient)
        L.emu_scene_lights(h, _fp(li), li.shape[0], _fp(amb))
        bb = _fp(f32(sa.scene_aabb)) if sa.scene_aabb is not None else None
        if refit:
            assert L.emu_scene_refit(h, bb, self.opts[0], self.opts[1], self.opts[2]) == 0
            return
        L.emu_scene_commit(h, bb, self.opts[0], self.opts[1], self.opts[2], self.opts[3])

    def update_instance(self, mesh_index, m3x4, aabb):
        mm, bb = np.ascontiguousarray(m3x4, dtype=np.float32), np.ascontiguousarray(aabb, dtype=np.float32)
        assert self.L.emu_scene_update_instance(self.h, int(mesh_index), _fp(mm), _fp(bb)) == 0

    def recommit(self, scene_aabb=None):
        bb = None if scene_aabb is None else _fp(np.ascontiguousarray(scene_aabb, dtype=np.float32))
        assert self.L.emu_scene_recommit(self.h, bb) == 0

    def scene_aabb(self):
        out = np.zeros(6, np.float32)
        self.L.emu_get_aabb(self.h, _fp(out))
        return out

    def set_presplit(self, ratio, depth=3):
        self.L.emu_set_presplit(self.h, float(ratio), int(depth))

    def n_slots(self):
        return self.L.emu_n_slots(self.h)

    def n_nodes(self):
        return self.L.emu_n_nodes(self.h)

    def max_stack(self):
        return self.L.emu_max_stack(self.h)

    def set_partition(self, rank, world, band_rows=32):
        self.rank, self.world, self.band_rows = rank, world, band_rows

    def set_option(self, key, value):
        if key == "chunk_pixels":
            self.chunk = int(value)

    def intersect(self, origins, dirs):
        o = np.ascontiguousarray(np.atleast_2d(origins), dtype=np.float32)
        d = np.ascontiguousarray(np.atleast_2d(dirs), dtype=np.float32)
        n = o.shape[0]
        ids = np.zeros((n, 3), np.int32); t = np.zeros(n, np.float32)
        self.L.emu_intersect(self.h, _fp(o), _fp(d), n, ids.ctypes.data_as(C.POINTER(C.c_int32)), _fp(t))
        return ids, t

    def render(self, params, want=host.RT_OUT_ALL, out=None):
        q = host.make_params(params)
        out = out if out is not None else host.RayTracer.alloc_host_outputs(params, want)
        ptr = lambda k: out[k].ctypes.data if k in out else None
        st = (C.c_uint64 * 3)()
        self.L.emu_render(self.h, C.byref(q), self.rank, self.world, self.band_rows, self.chunk, ptr("rgba32f"), ptr("rgba8"),
                          ptr("hit_id"), ptr("hit_t"), st)
        out["stats"] = dict(rays_primary=int(st[0]), rays_shadow=int(st[1]), rays_mirror=int(st[2]))
        return out


def ray_triangle(o, d, v9, tmax=float("inf")):
    """rt_core.h rt_ray_tri (division-free rejections + exact accept path) for one ray/triangle pair."""
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    o, d, v9 = f(o), f(d), f(v9)
    t = C.c_float(0)
    hit = lib().emu_ray_triangle(_fp(o), _fp(d), _fp(v9), C.c_float(tmax), C.byref(t))
    return bool(hit), np.float32(t.value)

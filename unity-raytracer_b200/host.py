"""Host side above the C-ABI (include/rtb200.h): a ctypes binding of librtb200.so plus a thin mirror of the reference's
own driver interface (`RayTracingSetup`: ImagePlane / BackgroundColor / MaxReflectionBounces / Scene, UpdateScene(),
CastPixelRays(CameraData) -> PixelColors; RayTracingSetup.cs:21-40,120-128,275-302), so tests read like reference code.

The reference's host language is C# (Unity); no .NET/Mono toolchain exists in this image, so the managed shim is shipped
as source only (unity/RayTracerNative.cs) and this module is the executable host side.

There is NO CPU path here: if librtb200.so is missing or no CUDA device is present, construction raises.
"""
import ctypes as C
import os
import numpy as np

from . import scenes

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB200_LIB") or os.path.join(_HERE, "librtb200.so")      # RTB200_LIB: an instrumented development build (tools/checked_run.sh)

RT_OUT_RGBA32F, RT_OUT_RGBA8, RT_OUT_HIT_ID, RT_OUT_HIT_T = 1, 2, 4, 8
RT_OUT_ALL = 15

ABI_SYMBOLS = ["rt_create", "rt_destroy", "rt_last_error", "rt_abi_version", "rt_scene_begin", "rt_scene_spheres",
               "rt_scene_triangles", "rt_scene_mesh", "rt_scene_base_mesh", "rt_scene_mesh_instance", "rt_scene_update_instance", "rt_scene_update_instances", "rt_scene_lights", "rt_scene_commit", "rt_scene_get_aabb", "rt_scene_update_mesh",
               "rt_scene_update_triangles", "rt_scene_update_spheres", "rt_scene_update_lights", "rt_scene_recommit", "rt_scene_refit", "rt_render",
               "rt_render_device", "rt_read_outputs", "rt_device_outputs", "rt_intersect", "rt_intersect_static", "rt_get_stats", "rt_set_stream",
               "rt_set_partition", "rt_host_register", "rt_host_unregister", "rt_alloc_outputs", "rt_export_outputs", "rt_import_outputs", "rt_set_option"]


class RenderParams(C.Structure):
    """rt_render_params"""
    _fields_ = [("pos", C.c_float * 3), ("fwd", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("res_x", C.c_int32), ("res_y", C.c_int32), ("dist", C.c_float), ("half_h", C.c_float), ("half_v", C.c_float),
                ("background_rgb", C.c_float * 3), ("max_reflection_bounces", C.c_int32), ("spp_grid", C.c_int32)]


class Stats(C.Structure):
    """rt_stats"""
    _fields_ = [("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_mirror", C.c_uint64),
                ("ms_render", C.c_float), ("ms_d2h", C.c_float), ("ms_commit_upload", C.c_float), ("ms_commit_build", C.c_float),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("n_prims", C.c_uint32), ("n_nodes", C.c_uint32),
                ("kernel_launches", C.c_uint32), ("reserved", C.c_uint32),
                ("ms_kernel", C.c_float * 8), ("launches_kernel", C.c_uint32 * 8), ("rays_shadow_skipped", C.c_uint64)]

    KERNELS = ("trace_primary", "shade_primary", "trace_queue", "shade_queue", "fold", "finalize", "shadow_primary", "shadow_queue")

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k not in ("reserved", "ms_kernel", "launches_kernel")}
        d["ms_kernel"] = {n: float(self.ms_kernel[i]) for i, n in enumerate(self.KERNELS)}
        d["launches_kernel"] = {n: int(self.launches_kernel[i]) for i, n in enumerate(self.KERNELS)}
        return d


def make_params(p):
    q = RenderParams()
    for k in ("pos", "fwd", "right", "up", "background_rgb"):
        setattr(q, k, (C.c_float * 3)(*[float(v) for v in p[k]]))
    q.res_x, q.res_y = int(p["res_x"]), int(p["res_y"])
    q.dist, q.half_h, q.half_v = float(p["dist"]), float(p["half_h"]), float(p["half_v"])
    q.max_reflection_bounces = int(p.get("max_reflection_bounces", 0))
    q.spp_grid = int(p.get("spp_grid", 1))
    return q


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"librtb200 error {code}: {msg}")
        self.code = code


_lib = None


def load_library():
    """dlopen librtb200.so and declare the ABI. Raises if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python unity-raytracer_b200/build.py` "
                           "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    fp, ip, vp = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_void_p
    L.rt_create.argtypes = [C.POINTER(vp), ip, C.c_int]
    L.rt_destroy.argtypes = [vp]; L.rt_destroy.restype = None
    L.rt_last_error.argtypes = [vp]; L.rt_last_error.restype = C.c_char_p
    L.rt_abi_version.restype = C.c_int
    L.rt_scene_begin.argtypes = [vp]
    L.rt_scene_spheres.argtypes = [vp, fp, vp, C.c_int]
    L.rt_scene_triangles.argtypes = [vp, fp, fp, vp, C.c_int]
    L.rt_scene_mesh.argtypes = [vp, fp, fp, C.c_int, vp, fp]
    L.rt_scene_lights.argtypes = [vp, fp, C.c_int, fp]
    L.rt_scene_base_mesh.argtypes = [vp, fp, C.c_int]
    L.rt_scene_mesh_instance.argtypes = [vp, C.c_int, fp, vp, fp]
    L.rt_scene_update_instance.argtypes = [vp, C.c_int, fp, fp]
    L.rt_scene_update_instances.argtypes = [vp, C.POINTER(C.c_int32), fp, fp, C.c_int]
    L.rt_scene_commit.argtypes = [vp, fp]
    L.rt_scene_get_aabb.argtypes = [vp, fp]
    L.rt_scene_update_mesh.argtypes = [vp, C.c_int, fp, fp, fp]
    L.rt_scene_update_triangles.argtypes = [vp, C.c_int, C.c_int, fp, fp]
    L.rt_scene_update_spheres.argtypes = [vp, C.c_int, C.c_int, fp]
    L.rt_scene_update_lights.argtypes = [vp, fp, C.c_int, fp]
    L.rt_scene_recommit.argtypes = [vp, fp]
    L.rt_scene_refit.argtypes = [vp, fp]
    L.rt_render.argtypes = [vp, C.POINTER(RenderParams), vp, vp, vp, vp, C.POINTER(Stats)]
    L.rt_render_device.argtypes = [vp, C.POINTER(RenderParams), C.c_int, C.POINTER(Stats)]
    L.rt_read_outputs.argtypes = [vp, vp, vp, vp, vp]
    L.rt_device_outputs.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.rt_intersect.argtypes = [vp, fp, fp, C.c_int, ip, fp]
    L.rt_intersect_static.argtypes = [vp, fp, fp, C.c_int, ip, fp]
    L.rt_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.rt_set_stream.argtypes = [vp, vp]
    L.rt_set_partition.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    L.rt_alloc_outputs.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    L.rt_export_outputs.argtypes = [vp, vp]
    L.rt_import_outputs.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int]
    L.rt_set_option.argtypes = [vp, C.c_char_p, C.c_double]
    L.rt_host_register.argtypes = [vp, vp, C.c_size_t]
    L.rt_host_unregister.argtypes = [vp, vp]
    _lib = L
    return L


def _fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class RayTracer:
    """One rt_ctx. Methods map 1:1 onto the C-ABI."""

    def __init__(self, devices=(0,)):
        self.L = load_library()
        h = C.c_void_p()
        ids = (C.c_int32 * len(devices))(*devices)
        rc = self.L.rt_create(C.byref(h), ids, len(devices))
        if rc != 0:
            raise RtError(rc, self.L.rt_last_error(None).decode())
        self.h = h
        self.n_devices = len(devices)

    def close(self):
        if getattr(self, "h", None):
            self.L.rt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise RtError(rc, self.L.rt_last_error(self.h).decode())

    # ---- scene (UpdateScene's result, RayTracingSetup.cs:120-128) ----
    def set_scene(self, sa):
        L, h = self.L, self.h
        self._chk(L.rt_scene_begin(h))
        f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        if sa.spheres.shape[0]:
            s, m = f32(sa.spheres), np.ascontiguousarray(sa.sphere_mats)
            self._chk(L.rt_scene_spheres(h, _fptr(s), m.ctypes.data, s.shape[0]))
        if sa.tris.shape[0]:
            v, n, m = f32(sa.tris), f32(sa.tri_normals), np.ascontiguousarray(sa.tri_mats)
            self._chk(L.rt_scene_triangles(h, _fptr(v), _fptr(n), m.ctypes.data, v.shape[0]))
        for bv in getattr(sa, "base_meshes", []):                  # object-space meshes the instances refer to
            bv = f32(bv)
            rc = L.rt_scene_base_mesh(h, _fptr(bv), bv.shape[0])
            if rc < 0:
                self._chk(rc)
        for m in sa.meshes:                                         # call order == MeshIndex, baked meshes and instances alike
            if isinstance(m, scenes.MeshInstance):
                mm, mat = f32(m.m3x4), np.ascontiguousarray(m.mat)
                bb = None if m.aabb is None else f32(m.aabb)
                self._chk(L.rt_scene_mesh_instance(h, int(m.base), _fptr(mm), mat.ctypes.data, _fptr(bb) if bb is not None else None))
                continue
            v, n, mat, aabb = m
            v, n, aabb, mat = f32(v), f32(n), f32(aabb), np.ascontiguousarray(mat)
            self._chk(L.rt_scene_mesh(h, _fptr(v), _fptr(n), v.shape[0], mat.ctypes.data, _fptr(aabb)))
        li, amb = f32(sa.lights), f32(sa.ambient)
        self._chk(L.rt_scene_lights(h, _fptr(li), li.shape[0], _fptr(amb)))
        if sa.scene_aabb is not None:
            bb = f32(sa.scene_aabb)
            self._chk(L.rt_scene_commit(h, _fptr(bb)))
        else:
            self._chk(L.rt_scene_commit(h, None))

    # ---- dirty-tracking updates (rt_scene_update_* + rt_scene_recommit): only what moved is re-sent ----
    def update_mesh(self, mesh_index, v, n, aabb):
        v, n, aabb = (np.ascontiguousarray(a, dtype=np.float32) for a in (v, n, aabb))
        self._chk(self.L.rt_scene_update_mesh(self.h, int(mesh_index), _fptr(v), _fptr(n), _fptr(aabb)))

    def update_instance(self, mesh_index, m3x4, aabb=None):
        """rt_scene_update_instance: new localToWorldMatrix (and Mesh.AABB, None = computed) of an instanced mesh."""
        mm = np.ascontiguousarray(m3x4, dtype=np.float32)
        bb = None if aabb is None else np.ascontiguousarray(aabb, dtype=np.float32)
        self._chk(self.L.rt_scene_update_instance(self.h, int(mesh_index), _fptr(mm), _fptr(bb) if bb is not None else None))

    def update_instances(self, mesh_indices, m3x4s, aabbs=None):
        """rt_scene_update_instances: many instances re-posed in one call (all or nothing); m3x4s [n, 12], aabbs [n, 6] or None."""
        idx = np.ascontiguousarray(mesh_indices, dtype=np.int32).reshape(-1)
        mm = np.ascontiguousarray(m3x4s, dtype=np.float32).reshape(-1)
        bb = None if aabbs is None else np.ascontiguousarray(aabbs, dtype=np.float32).reshape(-1)
        if mm.size != 12 * idx.size or (bb is not None and bb.size != 6 * idx.size): raise ValueError("update_instances: need 12 (and 6) floats per index")
        self._chk(self.L.rt_scene_update_instances(self.h, idx.ctypes.data_as(C.POINTER(C.c_int32)), _fptr(mm), _fptr(bb) if bb is not None else None, int(idx.size)))

    def update_triangles(self, first, v, n):
        v, n = np.ascontiguousarray(v, dtype=np.float32), np.ascontiguousarray(n, dtype=np.float32)
        self._chk(self.L.rt_scene_update_triangles(self.h, int(first), int(v.shape[0]), _fptr(v), _fptr(n)))

    def update_spheres(self, first, center_r2):
        s = np.ascontiguousarray(center_r2, dtype=np.float32)
        self._chk(self.L.rt_scene_update_spheres(self.h, int(first), int(s.shape[0]), _fptr(s)))

    def update_lights(self, lights, ambient=None):
        li = np.ascontiguousarray(lights, dtype=np.float32)
        amb = None if ambient is None else np.ascontiguousarray(ambient, dtype=np.float32)
        self._chk(self.L.rt_scene_update_lights(self.h, _fptr(li), int(li.shape[0]), _fptr(amb) if amb is not None else None))

    def recommit(self, scene_aabb=None):
        bb = None if scene_aabb is None else np.ascontiguousarray(scene_aabb, dtype=np.float32)
        self._chk(self.L.rt_scene_recommit(self.h, _fptr(bb) if bb is not None else None))

    def refit(self, scene_aabb=None):
        """rt_scene_refit: keep the committed tree, recompute its boxes from the updated primitives"""
        bb = None if scene_aabb is None else np.ascontiguousarray(scene_aabb, dtype=np.float32)
        self._chk(self.L.rt_scene_refit(self.h, _fptr(bb) if bb is not None else None))

    def host_register(self, arr):
        """page-lock a numpy array that is reused across frames (rt_host_register)"""
        self._chk(self.L.rt_host_register(self.h, arr.ctypes.data, arr.nbytes))

    def host_unregister(self, arr):
        self._chk(self.L.rt_host_unregister(self.h, arr.ctypes.data))

    def scene_aabb(self):
        out = np.zeros(6, np.float32)
        self._chk(self.L.rt_scene_get_aabb(self.h, _fptr(out)))
        return out

    # ---- render (CastPixelRays, RayTracingSetup.cs:275-302) ----
    @staticmethod
    def alloc_host_outputs(params, want=RT_OUT_ALL, pinned=False):
        n = int(params["res_x"]) * int(params["res_y"])
        spec = {"rgba32f": (RT_OUT_RGBA32F, (n, 4), np.float32), "rgba8": (RT_OUT_RGBA8, (n, 4), np.uint8),
                "hit_id": (RT_OUT_HIT_ID, (n, 3), np.int32), "hit_t": (RT_OUT_HIT_T, (n,), np.float32)}
        out = {}
        for k, (bit, shape, dt) in spec.items():
            if not (want & bit):
                continue
            if pinned:
                import torch
                tdt = {np.float32: torch.float32, np.uint8: torch.uint8, np.int32: torch.int32}[dt]
                out[k] = torch.empty(shape, dtype=tdt).pin_memory().numpy()
            else:
                out[k] = np.zeros(shape, dt)
        return out

    def render(self, params, want=RT_OUT_ALL, out=None):
        """rt_render: host buffers in, host buffers out (D2H inside the call)."""
        q = make_params(params)
        out = out if out is not None else self.alloc_host_outputs(params, want)
        ptr = lambda k: out[k].ctypes.data if k in out else None
        st = Stats()
        self._chk(self.L.rt_render(self.h, C.byref(q), ptr("rgba32f"), ptr("rgba8"), ptr("hit_id"), ptr("hit_t"), C.byref(st)))
        out["stats"] = st.as_dict()
        return out

    def render_device(self, params, want=RT_OUT_RGBA8, sync_stats=False):
        q = make_params(params)
        if sync_stats:
            st = Stats()
            self._chk(self.L.rt_render_device(self.h, C.byref(q), want, C.byref(st)))
            return st.as_dict()
        self._chk(self.L.rt_render_device(self.h, C.byref(q), want, None))
        return None

    def read_outputs(self, params, want=RT_OUT_ALL, out=None):
        out = out if out is not None else self.alloc_host_outputs(params, want)
        ptr = lambda k: out[k].ctypes.data if k in out else None
        self._chk(self.L.rt_read_outputs(self.h, ptr("rgba32f"), ptr("rgba8"), ptr("hit_id"), ptr("hit_t")))
        return out

    def device_outputs(self):
        p = [C.c_void_p() for _ in range(4)]
        self._chk(self.L.rt_device_outputs(self.h, *[C.byref(x) for x in p]))
        return [x.value for x in p]

    def stats(self):
        st = Stats()
        self._chk(self.L.rt_get_stats(self.h, C.byref(st)))
        return st.as_dict()

    def intersect(self, origins, dirs):
        """Scene.IntersectRay for n rays (Scene.cs:43-122) -> (ids (n,3) int32, t (n,) float32)."""
        o = np.ascontiguousarray(np.atleast_2d(origins), dtype=np.float32)
        d = np.ascontiguousarray(np.atleast_2d(dirs), dtype=np.float32)
        n = o.shape[0]
        ids = np.zeros((n, 3), np.int32); t = np.zeros(n, np.float32)
        self._chk(self.L.rt_intersect(self.h, _fptr(o), _fptr(d), n, ids.ctypes.data_as(C.POINTER(C.c_int32)), _fptr(t)))
        return ids, t

    INTERSECT_STATIC_MAX = 65536          # RT_INTERSECT_STATIC_MAX

    def intersect_static(self, origins, dirs):
        """rt_intersect_static: the same query without any device allocation (staging in a static device array of the
        library, at most INTERSECT_STATIC_MAX rays per call; no counterpart in the reference)."""
        o = np.ascontiguousarray(np.atleast_2d(origins), dtype=np.float32)
        d = np.ascontiguousarray(np.atleast_2d(dirs), dtype=np.float32)
        n = o.shape[0]
        ids = np.zeros((n, 3), np.int32); t = np.zeros(n, np.float32)
        self._chk(self.L.rt_intersect_static(self.h, _fptr(o), _fptr(d), n, ids.ctypes.data_as(C.POINTER(C.c_int32)), _fptr(t)))
        return ids, t

    # ---- plumbing ----
    def set_stream(self, cuda_stream_ptr):
        self._chk(self.L.rt_set_stream(self.h, cuda_stream_ptr))

    def set_partition(self, rank, world, band_rows=32):
        self._chk(self.L.rt_set_partition(self.h, rank, world, band_rows))

    def set_option(self, key, value):
        self._chk(self.L.rt_set_option(self.h, key.encode(), float(value)))

    def alloc_outputs(self, res_x, res_y, want):
        self._chk(self.L.rt_alloc_outputs(self.h, res_x, res_y, want))

    def export_outputs(self):
        buf = C.create_string_buffer(4 * 64)
        self._chk(self.L.rt_export_outputs(self.h, buf))
        return buf.raw

    def import_outputs(self, handles, res_x, res_y, want):
        buf = C.create_string_buffer(handles, 4 * 64)
        self._chk(self.L.rt_import_outputs(self.h, buf, res_x, res_y, want))


# --------------------------------------------------------------------------------------------------------------------
# Mirror of the reference's driver object (RayTracingSetup.cs), minus everything Unity-editor (gizmos, Debug.Draw*).
# --------------------------------------------------------------------------------------------------------------------
class Resolution:                      # Data/Camera/Resolution.cs:6-9
    def __init__(self, X=0, Y=0):
        self.X, self.Y = X, Y


class ImagePlane:                      # Data/Camera/ImagePlane.cs:11-24
    def __init__(self, Resolution=None, DistanceToCamera=0.0, HalfHorizontalLength=0.0, HalfVerticalLength=0.0):
        self.Resolution = Resolution or globals()["Resolution"]()
        self.DistanceToCamera = DistanceToCamera
        self.HalfHorizontalLength = HalfHorizontalLength
        self.HalfVerticalLength = HalfVerticalLength


class CameraData:                      # Data/Camera/CameraData.cs:5-10
    def __init__(self, Position=(0, 0, 0), Forward=(0, 0, 1), Right=(1, 0, 0), Up=(0, 1, 0)):
        self.Position, self.Forward, self.Right, self.Up = Position, Forward, Right, Up


class RayTracingSetup:
    """Fields and call order of RayTracingSetup.Update() (RayTracingSetup.cs:171-199): UpdateScene() then
    CastPixelRays(CameraData), which fills PixelColors (Color[] as an (X*Y,4) float32 array, index x + y*X)."""

    def __init__(self, devices=(0,)):
        self.ImagePlane = ImagePlane()
        self.BackgroundColor = (0.0, 0.0, 0.0)      # Color.black, RTS:22
        self.MaxReflectionBounces = 0               # RTS:23
        self.PixelColors = np.zeros((0, 4), np.float32)
        self.Scene = None
        self._rt = RayTracer(devices)

    def UpdateScene(self, scene_arrays):            # RTS:120-128 (the Fetch* calls are the caller's: Unity components)
        self.Scene = scene_arrays
        self._rt.set_scene(scene_arrays)

    def params(self, cam):
        ip = self.ImagePlane
        return dict(pos=cam.Position, fwd=cam.Forward, right=cam.Right, up=cam.Up, res_x=ip.Resolution.X, res_y=ip.Resolution.Y,
                    dist=ip.DistanceToCamera, half_h=ip.HalfHorizontalLength, half_v=ip.HalfVerticalLength,
                    background_rgb=self.BackgroundColor, max_reflection_bounces=self.MaxReflectionBounces, spp_grid=1)

    def CastPixelRays(self, cameraData):            # RTS:275-302
        out = self._rt.render(self.params(cameraData), RT_OUT_RGBA32F)
        self.PixelColors = out["rgba32f"]
        return self.PixelColors

    def close(self):
        self._rt.close()


class PipelinedRayTracingSetup(RayTracingSetup):
    """RayTracingSetup.Update() (RayTracingSetup.cs:171-199) for callers that accept ONE frame of latency: two contexts on the
    same devices; Update(scene, cam) hands `scene` to a worker thread that commits it into the idle context while this thread
    renders the scene handed over by the previous Update with `cam`.  The first Update has nothing older to show, so it
    commits and renders `scene` itself.  Contexts share no state (include/rtb200.h, conventions), so the two calls overlap:
    the host-bound upload runs under the device-bound frame (bench.py `frame_loop`)."""

    def __init__(self, devices=(0,)):
        import threading
        super().__init__(devices)
        self._ctx = [self._rt, RayTracer(devices)]
        self._front = None                      # index of the context holding the newest committed scene
        self._pending = None                    # (thread, index, error list) of the commit in flight
        self._threading = threading

    def set_option(self, key, value):
        for c in self._ctx:
            c.set_option(key, value)

    def _join(self):
        if self._pending is not None:
            th, idx, err = self._pending
            th.join()
            self._pending = None
            if err:
                raise err[0]
            self._front = idx

    def Update(self, scene_arrays, cameraData):
        self._join()
        if self._front is None:                 # first frame: nothing to overlap with
            self._ctx[0].set_scene(scene_arrays)
            self._front = 0
            self.Scene = scene_arrays
        else:
            idx, err = 1 - self._front, []

            def work(c=self._ctx[idx], s=scene_arrays, e=err):
                try:
                    c.set_scene(s)
                except Exception as ex:         # noqa: BLE001 - re-raised on the caller's thread by _join
                    e.append(ex)
            th = self._threading.Thread(target=work)
            th.start()
            self._pending = (th, idx, err)
        out = self._ctx[self._front].render(self.params(cameraData), RT_OUT_RGBA32F)
        self.PixelColors = out["rgba32f"]
        if self._pending is not None:
            self.Scene = scene_arrays
        return self.PixelColors

    def Flush(self, cameraData):
        """renders the scene handed over by the last Update (ends the one-frame latency, e.g. before a screenshot)"""
        self._join()
        out = self._ctx[self._front].render(self.params(cameraData), RT_OUT_RGBA32F)
        self.PixelColors = out["rgba32f"]
        return self.PixelColors

    def close(self):
        try:
            self._join()
        finally:
            for c in self._ctx:
                c.close()

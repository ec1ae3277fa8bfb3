"""A SECOND, independent restatement of the reference's hot path — test infrastructure only.

Written straight from the C# under /root/reference/Assets/RayTracer (RMath.cs, Scene.cs, AABB.cs, Sphere.cs, ImagePlane.cs,
Rgb.cs, Demo-RayTracing/RayTracingSetup.cs:275-455) without looking at oracle/rt_oracle.cpp or csrc/rt_core.h: scalar numpy
float32, no BVH, recursion as the C# has it.  Its job is to disagree with the C++ oracle wherever the two readings of the C#
differ (tests/test_independent.py).  Every float32 operation is one numpy scalar operation, so nothing is contracted or widened.
Unity.Mathematics 1.2.6 semantics used: dot = x*x' + y*y' + z*z' left to right; cross(a, b) = (a*b.yzx - a.yzx*b).yzx;
normalize(v) = rsqrt(dot(v, v)) * v with rsqrt(x) = 1 / sqrt(x); min(x, y) = (isnan(y) || x < y) ? x : y (max alike);
sqrt / acos / pow go through System.Math in double and are rounded back to float.
"""
import math
import numpy as np

F = np.float32
EPS = F(0.00001)          # RMath.Epsilon
SHADOW_EPS = F(0.0001)    # RayTracingSetup.ShadowRayEpsilon
FLT_MAX = np.finfo(np.float32).max
NONE, SPHERE, TRIANGLE, MESH_TRIANGLE = 0, 1, 2, 3     # ObjectType.cs


def f3(v):
    return (F(v[0]), F(v[1]), F(v[2]))


def add(a, b): return (a[0] + b[0], a[1] + b[1], a[2] + b[2])
def sub(a, b): return (a[0] - b[0], a[1] - b[1], a[2] - b[2])
def mul(a, b): return (a[0] * b[0], a[1] * b[1], a[2] * b[2])
def scale(a, s): return (a[0] * s, a[1] * s, a[2] * s)          # float3 * float
def lscale(s, a): return (s * a[0], s * a[1], s * a[2])         # float * float3
def div(a, s): return (a[0] / s, a[1] / s, a[2] / s)
def dot(a, b): return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]
def um_min(x, y): return x if (np.isnan(y) or x < y) else y
def um_max(x, y): return x if (np.isnan(y) or x > y) else y
def um_sqrt(x): return F(np.sqrt(np.float64(x)))


def cross(a, b):
    t = (a[0] * b[1] - a[1] * b[0], a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2])     # (a*b.yzx - a.yzx*b)
    return (t[1], t[2], t[0])                                                                # .yzx


def normalize(v):
    return lscale(F(1.0) / um_sqrt(dot(v, v)), v)


def ray_aabb(o, d, mn, mx):                                      # RMath.cs:12-26
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        inv = (F(1.0) / d[0], F(1.0) / d[1], F(1.0) / d[2])
        tmin, tmax = F(0.0), F(np.inf)
        for i in range(3):
            t1 = (mn[i] - o[i]) * inv[i]
            t2 = (mx[i] - o[i]) * inv[i]
            tmin = um_min(um_max(t1, tmin), um_max(t2, tmin))
            tmax = um_max(um_min(t1, tmax), um_min(t2, tmax))
        return bool(tmin <= tmax)


def ray_triangle(o, d, v0, v1, v2):                              # RMath.cs:29-73 -> (hit, t)
    with np.errstate(all="ignore"):
        e1, e2 = sub(v1, v0), sub(v2, v0)
        h = cross(d, e2)
        a = dot(e1, h)
        if a > -EPS and a < EPS:
            return False, F(0)
        f = F(1.0) / a
        s = sub(o, v0)
        u = f * dot(s, h)
        if u < F(0) or u > F(1):
            return False, F(0)
        q = cross(s, e1)
        v = f * dot(d, q)
        if v < F(0) or u + v > F(1):
            return False, F(0)
        t = f * dot(e2, q)
        return (True, t) if t > EPS else (False, F(0))


def ray_sphere(o, d, c, r2):                                     # RMath.cs:81-108 -> (hit, t)
    with np.errstate(all="ignore"):
        oc = sub(o, c)
        uoc = dot(d, oc)
        disc = uoc * uoc - (dot(oc, oc) - r2)
        if disc < 0:
            return False, F(0)
        sq = um_sqrt(disc)
        big = -uoc + sq
        if big < 0:
            return False, F(0)
        small = -uoc - sq
        return True, (big if small < 0 else small)


class Scene:
    """Scene.cs + the private half of RayTracingSetup.cs, fed from a SceneArrays-like object (flat arrays of the C-ABI)."""

    def __init__(self, sa):
        self.spheres = [(f3(s[:3]), F(s[3])) for s in sa.spheres]
        self.sphere_mats = list(sa.sphere_mats)
        self.tris = [(f3(t[0:3]), f3(t[3:6]), f3(t[6:9])) for t in sa.tris]
        self.tri_normals = [f3(n) for n in sa.tri_normals]
        self.tri_mats = list(sa.tri_mats)
        self.meshes = []
        for v, n, mat, box in sa.meshes:
            self.meshes.append(([(f3(t[0:3]), f3(t[3:6]), f3(t[6:9])) for t in v], [f3(x) for x in n], mat[0], (f3(box[:3]), f3(box[3:]))))
        self.lights = [(f3(l[:3]), f3(l[3:])) for l in sa.lights]
        self.ambient = f3(sa.ambient)
        if sa.scene_aabb is not None:
            self.aabb = (f3(sa.scene_aabb[:3]), f3(sa.scene_aabb[3:]))
        else:
            self.aabb = self.calculate_aabb()
        self.counts = [0, 0, 0]                                  # primary, shadow, mirror IntersectRay calls

    def calculate_aabb(self):                                    # Scene.cs:17-41, AABB.cs:10-20, Sphere.cs:17-22
        mn, mx = [F(FLT_MAX)] * 3, [F(-FLT_MAX)] * 3             # float.MinValue is -float.MaxValue
        for _t, _n, _m, (bmn, bmx) in self.meshes:               # Encapsulate(AABB): min(Min, other.Min)
            for k in range(3):
                mn[k] = um_min(mn[k], bmn[k]); mx[k] = um_max(mx[k], bmx[k])
        for tri in self.tris:                                    # Encapsulate(point): min(point, Min)
            for p in tri:
                for k in range(3):
                    mn[k] = um_min(p[k], mn[k]); mx[k] = um_max(p[k], mx[k])
        for c, r2 in self.spheres:
            r = um_sqrt(r2)
            for k in range(3):
                mn[k] = um_min(mn[k], c[k] - r); mx[k] = um_max(mx[k], c[k] + r)
        return (tuple(mn), tuple(mx))

    def intersect_ray(self, o, d):                               # Scene.cs:43-122 -> (type, index, mesh_index, distance)
        best, typ, index, mesh_index = F(FLT_MAX), NONE, -1, -1
        if not ray_aabb(o, d, *self.aabb):
            return typ, index, mesh_index, best
        for mi, (tris, _n, _m, box) in enumerate(self.meshes):
            if ray_aabb(o, d, *box):
                for ti, tri in enumerate(tris):
                    hit, t = ray_triangle(o, d, *tri)
                    if hit and best > t:
                        best, typ, index, mesh_index = t, MESH_TRIANGLE, ti, mi
        for si, (c, r2) in enumerate(self.spheres):
            hit, t = ray_sphere(o, d, c, r2)
            if hit and best > t:
                best, typ, index = t, SPHERE, si                 # MeshIndex keeps its value
        for ti, tri in enumerate(self.tris):
            hit, t = ray_triangle(o, d, *tri)
            if hit and best > t:
                best, typ, index = t, TRIANGLE, ti
        return typ, index, mesh_index, best

    def normal_and_material(self, p, typ, index, mesh_index):    # RayTracingSetup.cs:402-436
        if typ == SPHERE:
            return normalize(sub(p, self.spheres[index][0])), self.sphere_mats[index]
        if typ == TRIANGLE:
            return self.tri_normals[index], self.tri_mats[index]
        return self.meshes[mesh_index][1][index], self.meshes[mesh_index][2]

    def shade(self, o, d, bounce, max_bounces, background255, primary_out=None):     # RayTracingSetup.cs:304-366
        typ, index, mesh_index, dist = self.intersect_ray(o, d)
        if primary_out is not None:
            primary_out.extend((typ, index, mesh_index, dist))
        if typ == NONE:
            return background255
        with np.errstate(all="ignore"):
            p = add(o, scale(d, dist))                           # Ray.GetPoint
            n, mat = self.normal_and_material(p, typ, index, mesh_index)
            color = mul(self.ambient, f3(mat["ambient"]))        # CalculateAmbient(ambientReflectance, ambientRadiance): radiance * reflectance
            view = normalize(sub(o, p))
            for lpos, lint in self.lights:
                ldir = normalize(sub(lpos, p))
                so = add(p, scale(n, SHADOW_EPS))
                self.counts[1] += 1
                styp, _i, _m, sdist = self.intersect_ray(so, ldir)
                dl = sub(lpos, p)                                # distancesq(x, y) = lengthsq(y - x)
                d2 = dot(dl, dl)
                if styp != NONE and sdist * sdist < d2:
                    continue
                e = div(lint, d2)
                cosnl = um_max(F(0), dot(ldir, n))               # CalculateDiffuse
                diffuse = mul(scale(f3(mat["diffuse"]), cosnl), e)
                ldn = dot(ldir, n)                               # CalculateSpecular
                angle = F(math.acos(np.float64(ldn))) * F(57.295779513) if abs(ldn) <= 1 else F(np.nan)     # degrees(x) = x * 57.29578f
                if angle > F(90):
                    specular = (F(0), F(0), F(0))
                else:
                    v = add(ldir, view)
                    h = div(v, um_sqrt(dot(v, v)))
                    cosnh = um_max(F(0), dot(n, h))
                    pw = F(np.float64(cosnh) ** np.float64(mat["phong"])) if not (cosnh == 0 and mat["phong"] == 0) else F(1)
                    specular = mul(scale(f3(mat["specular"]), pw), e)
                color = add(color, add(diffuse, specular))
            if mat["is_mirror"] != 0 and bounce < max_bounces:
                ro = add(p, scale(n, SHADOW_EPS))                # Reflect
                rd = sub(mul(lscale(F(2), n), (dot(view, n),) * 3), view)
                self.counts[2] += 1
                child = self.shade(ro, rd, bounce + 1, max_bounces, background255)
                color = add(color, mul(f3(mat["mirror"]), child))
            return color


def render(sa, params, pixels=None):
    """CastPixelRays (RayTracingSetup.cs:275-302) for the pixel indices `pixels` (default: all): dict of float colour
    (Color = Value / 255, alpha 1), primary hit ids and distances, ray counts."""
    sc = Scene(sa)
    pos, fwd, right, up = (f3(params[k]) for k in ("pos", "fwd", "right", "up"))
    rx, ry = int(params["res_x"]), int(params["res_y"])
    dist, hh, hv = F(params["dist"]), F(params["half_h"]), F(params["half_v"])
    center = add(pos, scale(fwd, dist))                          # ImagePlane.Center
    top_left = add(sub(center, scale(right, hh)), scale(up, hv)) # center - halfRight + halfUp
    hlen, vlen = hh * F(2), hv * F(2)
    bg = scale(f3(params["background_rgb"]), F(255))             # new Rgb(Color) = rgb * 255
    px = range(rx * ry) if pixels is None else [int(i) for i in pixels]
    out = dict(rgba32f=np.zeros((len(px), 4), F), hit_id=np.zeros((len(px), 3), np.int32), hit_t=np.zeros(len(px), F))
    for k, i in enumerate(px):
        x, y = i % rx, i // rx
        right_move = (F(x) + F(0.5)) * hlen / F(rx)
        down_move = (F(y) + F(0.5)) * vlen / F(ry)
        pp = sub(add(top_left, lscale(right_move, right)), scale(up, down_move))
        d = normalize(sub(pp, pos))
        sc.counts[0] += 1
        prim = []
        c = sc.shade(pos, d, 0, int(params.get("max_reflection_bounces", 0)), bg, prim)
        out["rgba32f"][k] = (c[0] / F(255), c[1] / F(255), c[2] / F(255), F(1))
        out["hit_id"][k] = prim[:3]; out["hit_t"][k] = prim[3]
    out["stats"] = tuple(sc.counts)
    return out

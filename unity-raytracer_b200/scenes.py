"""Scene descriptions in the flat layouts the C-ABI takes (include/rtb200.h), the transcription of the reference's
only ray-traced scene, and the deterministic synthetic scenes BASELINE.json's configs name.

Host-side only (numpy).  Everything here is the *producer* side of the boundary — the part that in the reference is
played by SceneComponents/*.cs (SceneSphere.cs:9-22, SceneTriangle.cs:16-22, SceneMesh.cs:11-53,
ScenePointLight.cs:9-13) and by the Unity scene file Demo-RayTracing/RayTracing.unity.  The arrays built here are
handed unchanged to both the CUDA library and the CPU oracle.
"""
import numpy as np

# rt_material: MaterialData.cs:7-14 with bool IsMirror widened to int32 (not blittable otherwise) = 56 bytes
MAT_DTYPE = np.dtype([("diffuse", "<f4", 3), ("ambient", "<f4", 3), ("mirror", "<f4", 3), ("specular", "<f4", 3),
                      ("phong", "<f4"), ("is_mirror", "<i4")])
assert MAT_DTYPE.itemsize == 56

F = np.float32


def material(diffuse=(0, 0, 0), ambient=(0, 0, 0), mirror=(0, 0, 0), specular=(0, 0, 0), phong=0.0, is_mirror=0):
    m = np.zeros(1, MAT_DTYPE)
    m["diffuse"], m["ambient"], m["mirror"], m["specular"] = diffuse, ambient, mirror, specular
    m["phong"], m["is_mirror"] = phong, is_mirror
    return m


class MeshInstance:
    """An instanced mesh of a SceneArrays: base-mesh index, 3x4 localToWorldMatrix, material, Mesh.AABB or None."""

    def __init__(self, base, m3x4, mat, aabb):
        self.base, self.m3x4, self.mat, self.aabb = base, m3x4, mat, aabb


class SceneArrays:
    """The reference's `Scene` struct (Data/Objects/Scene.cs:8-15) as flat arrays.

    spheres      (n,4)  center.xyz, RadiusSquared                     Sphere.cs:9-10
    sphere_mats  (n,)   MAT_DTYPE                                     SphereData.cs
    tris         (n,9)  Vertex0, Vertex1, Vertex2 (loose triangles)   Triangle.cs:9-11
    tri_normals  (n,3)  TriangleData.Normals                          TriangleData.cs:11
    tri_mats     (n,)   MAT_DTYPE
    meshes       list of (verts (n,9), normals (n,3), mat (1,) MAT_DTYPE, aabb (6,))   Mesh.cs:9-12; order == MeshIndex
    lights       (n,6)  Position, Intensity                           PointLightData.cs:9-10
    ambient      (3,)   AmbientLightData.Radiance
    scene_aabb   (6,) or None -> recomputed per Scene.cs:17-41
    """

    def __init__(self):
        self.spheres = np.zeros((0, 4), F)
        self.sphere_mats = np.zeros(0, MAT_DTYPE)
        self.tris = np.zeros((0, 9), F)
        self.tri_normals = np.zeros((0, 3), F)
        self.tri_mats = np.zeros(0, MAT_DTYPE)
        self.meshes = []
        self.lights = np.zeros((0, 6), F)
        self.ambient = np.zeros(3, F)
        self.scene_aabb = None

    def add_sphere(self, center, r2, mat):
        self.spheres = np.vstack([self.spheres, np.array([[*center, r2]], F)])
        self.sphere_mats = np.concatenate([self.sphere_mats, mat])

    def add_triangle(self, v9, normal, mat):
        self.tris = np.vstack([self.tris, np.asarray(v9, F).reshape(1, 9)])
        self.tri_normals = np.vstack([self.tri_normals, np.asarray(normal, F).reshape(1, 3)])
        self.tri_mats = np.concatenate([self.tri_mats, mat])

    def add_mesh(self, verts9, normals, mat, aabb):
        self.meshes.append((np.ascontiguousarray(verts9, F), np.ascontiguousarray(normals, F), mat,
                            np.ascontiguousarray(aabb, F)))

    # ---- instanced meshes (rt_scene_base_mesh / rt_scene_mesh_instance): SceneMesh.cs:11-53 without the per-frame bake ----
    def add_base_mesh(self, verts9_object):
        """Object-space triangles (n,9) of a mesh that instances refer to; returns its index."""
        if not hasattr(self, "base_meshes"):
            self.base_meshes = []
        self.base_meshes.append(np.ascontiguousarray(verts9_object, F).reshape(-1, 9))
        return len(self.base_meshes) - 1

    def add_instance(self, base, m3x4, mat, aabb=None):
        """One occurrence of base mesh `base` under the 3x4 localToWorldMatrix m3x4; takes one MeshIndex like add_mesh.
        aabb None = let the library compute Mesh.AABB as SceneMesh.cs:22-31 does."""
        self.meshes.append(MeshInstance(base, np.ascontiguousarray(m3x4, F).reshape(3, 4), mat,
                                        None if aabb is None else np.ascontiguousarray(aabb, F)))

    def baked(self):
        """The same scene with every instance baked the way the reference does it (bake_tris): what the oracle consumes."""
        out = SceneArrays()
        out.spheres, out.sphere_mats, out.tris, out.tri_normals, out.tri_mats = self.spheres, self.sphere_mats, self.tris, self.tri_normals, self.tri_mats
        out.lights, out.ambient, out.scene_aabb = self.lights, self.ambient, self.scene_aabb
        for m in self.meshes:
            if isinstance(m, MeshInstance):
                tris, normals, aabb = bake_tris(self.base_meshes[m.base], m.m3x4)
                out.add_mesh(tris, normals, m.mat, aabb if m.aabb is None else m.aabb)
            else:
                out.meshes.append(m)
        return out

    def add_light(self, pos, intensity):
        inten = np.broadcast_to(np.asarray(intensity, F), (3,))        # ScenePointLight.cs:12 scalar -> float3
        self.lights = np.vstack([self.lights, np.array([[*pos, *inten]], F)])

    @property
    def n_mesh_tris(self):
        return sum((self.base_meshes[m.base].shape[0] if isinstance(m, MeshInstance) else m[0].shape[0]) for m in self.meshes)

    @property
    def n_prims(self):
        return self.n_mesh_tris + self.spheres.shape[0] + self.tris.shape[0]


# ----------------------------------------------------------------------------------------------------------------
# fp32 restatements of the producer-side math (strict binary32: numpy float32 ufuncs do not contract)
# ----------------------------------------------------------------------------------------------------------------
def um_cross(a, b):
    """Unity.Mathematics cross(x,y) = (x*y.yzx - x.yzx*y).yzx, float32, rows."""
    a = np.asarray(a, F); b = np.asarray(b, F)
    yzx = [1, 2, 0]
    r = a * b[..., yzx] - a[..., yzx] * b
    return r[..., yzx]


def um_dot(a, b):
    a = np.asarray(a, F); b = np.asarray(b, F)
    return (a[..., 0] * b[..., 0] + a[..., 1] * b[..., 1]) + a[..., 2] * b[..., 2]


def triangle_normal(v9):
    """Triangle.Normal, Triangle.cs:13-21: cross(V2-V0, V1-V0) / length(...)."""
    v = np.asarray(v9, F).reshape(-1, 3, 3)
    c = um_cross(v[:, 2] - v[:, 0], v[:, 1] - v[:, 0])
    ln = np.sqrt(um_dot(c, c)).astype(F)
    return (c / ln[:, None]).astype(F)


def trs_matrix(position, quat_xyzw, scale):
    """Transform.localToWorldMatrix of a root object = Matrix4x4.TRS (Unity native; restated in fp32)."""
    qx, qy, qz, qw = [F(v) for v in quat_xyzw]
    x, y, z = qx * F(2), qy * F(2), qz * F(2)
    xx, yy, zz = qx * x, qy * y, qz * z
    xy, xz, yz = qx * y, qx * z, qy * z
    wx, wy, wz = qw * x, qw * y, qw * z
    one = F(1)
    R = np.array([[one - (yy + zz), xy - wz, xz + wy],
                  [xy + wz, one - (xx + zz), yz - wx],
                  [xz - wy, yz + wx, one - (xx + yy)]], F)
    M = np.zeros((3, 4), F)
    M[:, :3] = R * np.asarray(scale, F)[None, :]
    M[:, 3] = np.asarray(position, F)
    return M


def multiply_point_3x4(M, p):
    """Matrix4x4.MultiplyPoint3x4: m00*x + m01*y + m02*z + m03, left to right, fp32."""
    p = np.asarray(p, F)
    return (((M[None, :, 0] * p[:, 0:1]) + (M[None, :, 1] * p[:, 1:2])) + (M[None, :, 2] * p[:, 2:3])) + M[None, :, 3]


def bake_mesh(local_verts, indices, M):
    """SceneMesh.Mesh getter, SceneMesh.cs:11-53: world-bake, AABB, triangles, NEGATED normals."""
    w = multiply_point_3x4(M, local_verts).astype(F)
    aabb = np.concatenate([w.min(axis=0), w.max(axis=0)]).astype(F)
    idx = np.asarray(indices, np.int64).reshape(-1, 3)
    tris = w[idx].reshape(-1, 9)
    normals = (-triangle_normal(tris)).astype(F)
    return tris, normals, aabb


def bake_tris(tris9_object, M):
    """SceneMesh.cs:21-52 for an un-indexed triangle list: MultiplyPoint3x4 of every vertex, AABB of the baked vertices,
    negated Triangle.Normal.  This is the arithmetic rt_scene_mesh_instance reproduces on the device."""
    w = multiply_point_3x4(np.asarray(M, F), np.asarray(tris9_object, F).reshape(-1, 3)).astype(F)
    if w.shape[0] == 0:                                             # aabb.Min = float.MaxValue, aabb.Max = float.MinValue (SceneMesh.cs:23-24)
        return np.zeros((0, 9), F), np.zeros((0, 3), F), np.array([np.finfo(F).max] * 3 + [-np.finfo(F).max] * 3, F)
    aabb = np.concatenate([w.min(axis=0), w.max(axis=0)]).astype(F)
    tris = w.reshape(-1, 9)
    return tris, (-triangle_normal(tris)).astype(F), aabb


# Unity built-in Cube (default-resources fileID 10202): 24 verts, 12 tris. Not in the repo; restated.
_CUBE_V = np.array([(.5, -.5, .5), (-.5, -.5, .5), (.5, .5, .5), (-.5, .5, .5), (.5, .5, -.5), (-.5, .5, -.5),
                    (.5, -.5, -.5), (-.5, -.5, -.5), (.5, .5, .5), (-.5, .5, .5), (.5, .5, -.5), (-.5, .5, -.5),
                    (.5, -.5, -.5), (.5, -.5, .5), (-.5, -.5, .5), (-.5, -.5, -.5), (-.5, -.5, .5), (-.5, .5, .5),
                    (-.5, .5, -.5), (-.5, -.5, -.5), (.5, -.5, -.5), (.5, .5, -.5), (.5, .5, .5), (.5, -.5, .5)], F)
_CUBE_I = [0, 2, 3, 0, 3, 1, 8, 4, 5, 8, 5, 9, 10, 6, 7, 10, 7, 11, 12, 13, 14, 12, 14, 15, 16, 17, 18, 16, 18, 19,
           20, 21, 22, 20, 22, 23]


def camera_identity():
    return dict(pos=(0, 0, 0), fwd=(0, 0, 1), right=(1, 0, 0), up=(0, 1, 0))


def make_params(res_x, res_y, dist, half_h, half_v, background=(0, 0, 0), bounces=0, spp_grid=1, cam=None):
    p = dict(cam or camera_identity())
    p.update(res_x=res_x, res_y=res_y, dist=dist, half_h=half_h, half_v=half_v, background_rgb=background,
             max_reflection_bounces=bounces, spp_grid=spp_grid)
    return p


def rt_unity(mirror_sphere=False, instanced=False):
    """Demo-RayTracing/RayTracing.unity transcribed (RT.unity:244-684 + Prefabs/*.prefab).

    mirror_sphere=False: as shipped (the Sphere instance overrides IsMirror=0 and ks=0, RT.unity:445-458).
    mirror_sphere=True : Sphere prefab defaults restored (Prefabs/Sphere.prefab:99-117) = BASELINE config 2.
    instanced=True     : the Cube goes in as an instance (object-space cube + localToWorldMatrix) instead of baked triangles.
    """
    s = SceneArrays()
    # SceneSphere.cs:13-21: radius = localScale.x * 0.5, stores r^2. Scale 20 -> r 10 -> r2 100.
    radius = F(20) * F(0.5)
    if mirror_sphere:
        smat = material((1, 0, 0), (1, 1, 1), (1, 1, 1), (1, 1, 1), 20.0, 1)
    else:
        smat = material((1, 0, 0), (1, 1, 1), (1, 1, 1), (0, 0, 0), 20.0, 0)
    s.add_sphere((0, 0, 29.6), radius * radius, smat)
    # SceneTriangle.cs:16-22: position + Offset{0,1,2}; normal = Triangle.Normal (NOT negated, RayTracingSetup.cs:166)
    offs = np.array([(0, 10, 0), (-10, -10, 0), (10, -10, 0)], F)
    for pos, kd in (((17.1, 0, 15), (0, 1, 0)), ((14.16, 0, 21.45), (1, 0, 1))):
        v = (np.asarray(pos, F)[None, :] + offs).reshape(9)
        s.add_triangle(v, triangle_normal(v)[0], material(kd, (1, 1, 1)))
    # Cube mesh: RT.unity:380-436, Prefabs/Cube.prefab (scale 28.664,10,10)
    M = trs_matrix((-24.7, 1.5497656e-6, 27.6), (-0.37513673, 0.13105033, 0.3026398, 0.8663183), (28.664, 10, 10))
    tris, normals, aabb = bake_mesh(_CUBE_V, _CUBE_I, M)
    if instanced:
        s.add_instance(s.add_base_mesh(_CUBE_V[np.asarray(_CUBE_I, np.int64).reshape(-1, 3)].reshape(-1, 9)), M, material((0, 1, 1)), aabb)
    else:
        s.add_mesh(tris, normals, material((0, 1, 1)), aabb)
    s.add_light((5.79, 0, 0), 100000.0)
    s.ambient = np.array((15, 15, 15), F)
    return s


def rt_unity_params(res_x=50, res_y=50, bounces=5, spp_grid=1):
    """RT.unity:346-354: 50x50, Dist 10, HalfH 20, HalfV 10, background black, MaxReflectionBounces 5."""
    return make_params(res_x, res_y, 10.0, 20.0, 10.0, (0, 0, 0), bounces, spp_grid)


# ----------------------------------------------------------------------------------------------------------------
# synthetic scenes (BASELINE.json configs 3-5)
# ----------------------------------------------------------------------------------------------------------------
def splitmix64(seed, idx):
    """splitmix64 keyed on (seed, index) -> uint64; vectorised."""
    with np.errstate(over="ignore"):
        z = (np.uint64(seed) + (np.asarray(idx, np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform01(seed, idx):
    return (splitmix64(seed, idx) >> np.uint64(40)).astype(np.float64) * (1.0 / (1 << 24))


_ICO_CACHE = {}


def icosphere(n):
    """Class-I geodesic sphere of frequency n: 10 n^2 + 2 unit vertices (float64), 20 n^2 CCW-outward faces."""
    if n in _ICO_CACHE:
        return _ICO_CACHE[n]
    t = (1.0 + 5 ** 0.5) / 2.0
    V = np.array([(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
                  (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)], np.float64)
    V /= np.linalg.norm(V, axis=1, keepdims=True)
    Fc = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
          (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
          (8, 6, 7), (9, 8, 1)]
    ii, jj = np.meshgrid(np.arange(n + 1), np.arange(n + 1), indexing="ij")
    keep = (ii + jj) <= n
    gi, gj = ii[keep], jj[keep]                                   # (n+1)(n+2)/2 grid points
    lid = -np.ones((n + 1, n + 1), np.int64)
    lid[gi, gj] = np.arange(gi.size)
    up = (ii + jj) <= n - 1
    ui, uj = ii[up], jj[up]
    dn = (ii + jj) <= n - 2
    di, dj = ii[dn], jj[dn]
    tri_local = np.concatenate([np.stack([lid[ui, uj], lid[ui + 1, uj], lid[ui, uj + 1]], 1),
                                np.stack([lid[di + 1, dj], lid[di + 1, dj + 1], lid[di, dj + 1]], 1)])
    pts, faces = [], []
    for f, (a, b, c) in enumerate(Fc):
        P = (V[a][None] * (n - gi - gj)[:, None] + V[b][None] * gi[:, None] + V[c][None] * gj[:, None]) / n
        P /= np.linalg.norm(P, axis=1, keepdims=True)
        pts.append(P)
        faces.append(tri_local + f * gi.size)
    P = np.concatenate(pts); Fi = np.concatenate(faces)
    q = np.floor((P + 1.000000123) * (1 << 20) + 0.37).astype(np.int64)      # weld shared edge/corner vertices
    key = (q[:, 0] << 42) | (q[:, 1] << 21) | q[:, 2]
    _, first, inv = np.unique(key, return_index=True, return_inverse=True)
    verts = P[first]
    Fi = inv[Fi]
    assert verts.shape[0] == 10 * n * n + 2, (verts.shape, n)
    assert Fi.shape[0] == 20 * n * n
    _ICO_CACHE[n] = (verts, Fi)
    return verts, Fi


def displaced_icosphere(n, radius, center, seed, amp=0.02, rot=None):
    """World-space triangles (n_tris,9) float32 of a radially displaced geodesic sphere; vertex k is displaced by
    radius*amp*U(-1,1) with U from splitmix64 keyed on (seed, k), so shared vertices stay welded."""
    verts, faces = icosphere(n)
    u = uniform01(seed, np.arange(verts.shape[0])) * 2.0 - 1.0
    p = verts * (radius * (1.0 + amp * u))[:, None]
    if rot is not None:
        p = p @ np.asarray(rot, np.float64).T
    w = (p + np.asarray(center, np.float64)[None]).astype(F)
    return w[faces].reshape(-1, 9), w


def mesh_from_world_tris(tris9, world_verts=None):
    """Producer-side per SceneMesh.cs:22-52 for already-baked vertices: negated normals + vertex AABB."""
    normals = (-triangle_normal(tris9)).astype(F)
    v = world_verts if world_verts is not None else tris9.reshape(-1, 3)
    aabb = np.concatenate([v.min(axis=0), v.max(axis=0)]).astype(F)
    return normals, aabb


def add_ground(s, y, half, center_xz=(0.0, 0.0), kd=(0.6, 0.6, 0.6)):
    """Two loose triangles whose Triangle.Normal (cross(V2-V0, V1-V0)) points +y."""
    cx, cz = center_xz
    a, b, c, d = (cx - half, y, cz - half), (cx + half, y, cz - half), (cx + half, y, cz + half), (cx - half, y, cz + half)
    for v in ((a, b, d), (c, d, b)):
        v9 = np.asarray(v, F).reshape(9)
        s.add_triangle(v9, triangle_normal(v9)[0], material(kd, (1, 1, 1), (0, 0, 0), (0, 0, 0), 0.0, 0))


def random_rotation(seed, k):
    """Deterministic rotation matrix from 3 uniforms (Shoemake)."""
    u1, u2, u3 = [float(uniform01(seed, 3 * k + i)) for i in range(3)]
    q = np.array([np.sqrt(1 - u1) * np.sin(2 * np.pi * u2), np.sqrt(1 - u1) * np.cos(2 * np.pi * u2),
                  np.sqrt(u1) * np.sin(2 * np.pi * u3), np.sqrt(u1) * np.cos(2 * np.pi * u3)])
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


SEED = 12345
ASPECT_HALF_H = 8.8889   # 16:9 with Dist 10, HalfV 5


def cfg3_scene(n=224, radius=20.0, phong=20.0):
    """~1M-triangle displaced geodesic sphere (20*224^2 = 1 003 520) + ground quad, 2 lights."""
    s = SceneArrays()
    tris, w = displaced_icosphere(n, radius, (0, 0, 60), SEED)
    normals, aabb = mesh_from_world_tris(tris, w)
    s.add_mesh(tris, normals, material((0.8, 0.8, 0.8), (1, 1, 1), (0, 0, 0), (1, 1, 1), phong, 0), aabb)
    add_ground(s, -radius - 2.0, 80.0, (0.0, 60.0))
    s.add_light((40, 40, 0), 2e5)
    s.add_light((-40, 40, 0), 2e5)
    s.ambient = np.array((15, 15, 15), F)
    return s


def cfg3_params(res_x=3840, res_y=2160):
    return make_params(res_x, res_y, 10.0, ASPECT_HALF_H, 5.0, (0.1, 0.1, 0.2), 0, 1)


def cfg4_scene(n=64, grid=5, radius=4.0):
    """grid^3 baked 'instances' of a displaced geodesic sphere (n=64: 81 920 tris each; 125 -> 10 240 000)."""
    s = SceneArrays()
    spacing = 3.0 * radius
    k = 0
    half = (grid - 1) / 2.0
    for ix in range(grid):
        for iy in range(grid):
            for iz in range(grid):
                c = ((ix - half) * spacing, (iy - half) * spacing, 70.0 + (iz - half) * spacing)
                tris, w = displaced_icosphere(n, radius, c, SEED, rot=random_rotation(SEED + 1, k))
                normals, aabb = mesh_from_world_tris(tris, w)
                kd = tuple(0.2 + 0.8 * float(uniform01(SEED + 2, 3 * k + j)) for j in range(3))
                s.add_mesh(tris, normals, material(kd, (1, 1, 1), (0, 0, 0), (1, 1, 1), 20.0, 0), aabb)
                k += 1
    s.add_light((40, 40, 0), 2e5)
    s.add_light((-40, 40, 0), 2e5)
    s.ambient = np.array((15, 15, 15), F)
    return s


def cfg4_instanced(n=64, grid=5, radius=4.0):
    """BASELINE config 4 as what its name says, an instanced-mesh scene: ONE displaced geodesic sphere in object space
    (n=64: 81 920 triangles) and grid^3 instances of it, each with its own rotation, position and diffuse colour.  `.baked()`
    gives the equivalent scene of grid^3 baked meshes (the reference's SceneMesh.cs:21-31 arithmetic), bit-comparable with it."""
    s = SceneArrays()
    base_tris, _ = displaced_icosphere(n, radius, (0.0, 0.0, 0.0), SEED)
    b = s.add_base_mesh(base_tris)
    spacing = 3.0 * radius
    half = (grid - 1) / 2.0
    k = 0
    for ix in range(grid):
        for iy in range(grid):
            for iz in range(grid):
                M = np.zeros((3, 4), F)
                M[:, :3] = random_rotation(SEED + 1, k).astype(F)
                M[:, 3] = ((ix - half) * spacing, (iy - half) * spacing, 70.0 + (iz - half) * spacing)
                kd = tuple(0.2 + 0.8 * float(uniform01(SEED + 2, 3 * k + j)) for j in range(3))
                s.add_instance(b, M, material(kd, (1, 1, 1), (0, 0, 0), (1, 1, 1), 20.0, 0))
                k += 1
    s.add_light((40, 40, 0), 2e5)
    s.add_light((-40, 40, 0), 2e5)
    s.ambient = np.array((15, 15, 15), F)
    return s


def instanced_soup(seed=21, n_bases=3, n_instances=9, n_baked=2, extent=30.0, mirror_frac=0.3):
    """Parity fixture for instancing: a few small object-space meshes, instances of them under random rotations, non-uniform
    scales and translations (some of them mirrors), interleaved in MeshIndex order with baked meshes, plus the spheres and
    loose triangles of random_soup."""
    rng = np.random.default_rng(seed)
    s = random_soup(n_tris=300, n_spheres=60, n_meshes=0, seed=seed + 1, extent=extent, mirror_frac=mirror_frac)
    bases = []
    for b in range(n_bases):
        k = int(rng.integers(20, 200))
        c = rng.uniform(-3, 3, (k, 1, 3))
        tr = (c + rng.normal(scale=1.2, size=(k, 3, 3))).reshape(k, 9).astype(F)
        bases.append(s.add_base_mesh(tr))

    def rand_matrix():
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        x, y, z, w = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        M = np.zeros((3, 4), F)
        M[:, :3] = (R * rng.uniform(0.4, 2.5, 3)[None, :]).astype(F)
        M[:, 3] = rng.uniform(-extent * 0.7, extent * 0.7, 3) + np.array([0, 0, 2.5 * extent])
        return M

    order = ["i"] * n_instances + ["b"] * n_baked
    rng.shuffle(order)
    for kind in order:
        mir = rng.uniform() < mirror_frac
        mat = material(tuple(rng.uniform(0.1, 1.0, 3)), (1, 1, 1), (0.7, 0.7, 0.7) if mir else (0, 0, 0), (1, 1, 1), 20.0, 1 if mir else 0)
        if kind == "i":
            s.add_instance(bases[int(rng.integers(0, n_bases))], rand_matrix(), mat)
        else:
            k = int(rng.integers(10, 80))
            c = rng.uniform(-extent * 0.7, extent * 0.7, (1, 1, 3)) + np.array([0, 0, 2.5 * extent])
            tr = (c + rng.normal(scale=2.0, size=(k, 3, 3))).reshape(k, 9).astype(F)
            normals, aabb = mesh_from_world_tris(tr)
            s.add_mesh(tr, normals, mat, aabb)
    return s


def cfg4_params(res_x=7680, res_y=4320):
    return make_params(res_x, res_y, 10.0, ASPECT_HALF_H, 5.0, (0.1, 0.1, 0.2), 0, 1)


def cfg5_scene(n=500, n_spheres=100000, radius=30.0):
    """n_spheres small spheres (10% mirrors) in an 80^3 box + 20*n^2-triangle displaced sphere + ground quad."""
    s = SceneArrays()
    i = np.arange(n_spheres)
    c = np.stack([uniform01(SEED + 10, 3 * i + a) for a in range(3)], 1) * 80.0 - 40.0
    c[:, 2] += 70.0
    r = 0.1 + 0.3 * uniform01(SEED + 11, i)
    s.spheres = np.concatenate([c, (r * r)[:, None]], 1).astype(F)
    m = np.zeros(n_spheres, MAT_DTYPE)
    m["diffuse"] = np.stack([uniform01(SEED + 12, 3 * i + a) for a in range(3)], 1).astype(F)
    m["ambient"] = 1.0
    m["specular"] = 1.0
    m["phong"] = 20.0
    mir = uniform01(SEED + 13, i) < 0.1
    m["is_mirror"] = mir.astype(np.int32)
    m["mirror"] = np.where(mir[:, None], 0.8, 0.0).astype(F)
    s.sphere_mats = m
    tris, w = displaced_icosphere(n, radius, (0, 0, 70), SEED)
    normals, aabb = mesh_from_world_tris(tris, w)
    s.add_mesh(tris, normals, material((0.8, 0.8, 0.8), (1, 1, 1), (0, 0, 0), (1, 1, 1), 20.0, 0), aabb)
    add_ground(s, -radius - 12.0, 90.0, (0.0, 70.0))
    s.add_light((40, 40, 0), 2e5)
    s.add_light((-40, 40, 0), 2e5)
    s.ambient = np.array((15, 15, 15), F)
    return s


def cfg5_params(res_x=3840, res_y=2160, spp_grid=4, bounces=4):
    return make_params(res_x, res_y, 10.0, ASPECT_HALF_H, 5.0, (0.1, 0.1, 0.2), bounces, spp_grid)


def random_soup(n_tris=2000, n_spheres=200, n_meshes=3, seed=7, extent=30.0, mirror_frac=0.3):
    """Small random mixed scene for parity tests: loose triangles, spheres, a few meshes of random triangles."""
    rng = np.random.default_rng(seed)
    s = SceneArrays()

    def rand_tris(k, size=2.0):
        c = rng.uniform(-extent, extent, (k, 1, 3)) + np.array([0, 0, 2.5 * extent])
        return (c + rng.uniform(-size, size, (k, 3, 3))).astype(F).reshape(k, 9)

    def rand_mats(k):
        m = np.zeros(k, MAT_DTYPE)
        m["diffuse"] = rng.uniform(0, 1, (k, 3)); m["ambient"] = rng.uniform(0, 1, (k, 3))
        m["specular"] = rng.uniform(0, 1, (k, 3)); m["phong"] = rng.uniform(1, 40, k)
        mir = rng.uniform(0, 1, k) < mirror_frac
        m["is_mirror"] = mir; m["mirror"] = rng.uniform(0, 1, (k, 3)) * mir[:, None]
        return m

    per = max(1, n_tris // (2 * max(1, n_meshes)))
    for _ in range(n_meshes):
        t = rand_tris(per)
        normals, aabb = mesh_from_world_tris(t)
        s.add_mesh(t, normals, rand_mats(1), aabb)
    cs = rng.uniform(-extent, extent, (n_spheres, 3)) + np.array([0, 0, 2.5 * extent])
    rs = rng.uniform(0.3, 2.5, n_spheres)
    s.spheres = np.concatenate([cs, (rs * rs)[:, None]], 1).astype(F)
    s.sphere_mats = rand_mats(n_spheres)
    nl = n_tris - per * n_meshes
    s.tris = rand_tris(nl)
    s.tri_normals = triangle_normal(s.tris)
    s.tri_mats = rand_mats(nl)
    s.add_light((40, 40, 0), 2e5)
    s.add_light((-30, 10, 20), 1e5)
    s.ambient = np.array((15, 15, 15), F)
    return s


def soup_params(res_x=96, res_y=64, bounces=3, spp_grid=1):
    return make_params(res_x, res_y, 10.0, 8.0, 5.0, (0.2, 0.3, 0.4), bounces, spp_grid)

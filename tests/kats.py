"""Hand-derived known answers (KAT-5 .. KAT-8) for what the reference's own test does not pin.

Each function follows ONE straight-line path through the reference's formulas — cited line by line — in exact rational
arithmetic rounded per operation (tests/exact32.py).  Nothing here calls the oracle, the numpy restatement or the product.
The expected values are compared with all three (tests/test_independent.py on the CPU, tests/test_gpu_parity.py on the GPU).
"""
import importlib

import numpy as np

import exact32 as X

F = np.float32
_sc = importlib.import_module("unity-raytracer_b200.scenes")


def _one_pixel_params(bounces):
    """RayTracing.unity's image plane (RT.unity:346-352) at resolution 1x1: the only pixel centre lies on the view axis, so
    the primary ray is normalize((0,0,10)) = rsqrt(100) * (0,0,10) = (0, 0, fl(0.1f * 10)) = (0,0,1) (RTS:291-297)."""
    return _sc.make_params(1, 1, 10.0, 20.0, 10.0, background=(0, 0, 0), bounces=bounces)


def _light_terms(P, N, V, lpos, lint, kd, ks, phong):
    """RayTracingSetup.cs:330,334,350-355 + CalculateDiffuse :443-455 + CalculateSpecular :375-400 for an unshadowed light
    that lies in front of the surface; returns diffuse + specular (per channel)."""
    ldir = X.normalize(X.vsub(lpos, P))                            # :330
    dl = X.vsub(lpos, P)
    d2 = X.dot(dl, dl)                                             # distancesq(surfacePoint, lightPosition), :334
    E = X.vdiv(lint, d2)                                           # :350
    ldn = X.dot(ldir, N)
    assert ldn > 0 and X.acos_deg(ldn) <= F(90)                    # max(0, .) inactive (:449), angle <= 90 (:388)
    diffuse = X.vmul(X.vscale(kd, ldn), E)                         # diffuseReflectance * cos * receivedIrradiance, :453
    vv = X.vadd(ldir, V)                                           # :393
    H = X.vdiv(vv, X.sqrt(X.dot(vv, vv)))                          # v / length(v), :394
    cnh = X.dot(N, H)
    assert cnh > 0                                                 # max(0, .) inactive, :398
    specular = X.vmul(X.vscale(ks, X.powf(cnh, phong)), E)         # :399
    return X.vadd(diffuse, specular)                               # diffuseRgb + specularRgb, :355


def kat5_blinn_phong():
    """KAT-5: the cfg2 sphere (Prefabs/Sphere.prefab defaults: ks = (1,1,1), Phong 20) under the centre ray, no bounce.
    Hit: oc = (0,0,-29.6), uoc = -29.6, disc = fl(uoc^2) - (fl(oc.oc) - 100) = 100, t = 29.6 - 10 (RMath.cs:85-106)."""
    s = _sc.SceneArrays()
    kd, ka, ks = X.v(1, 0, 0), X.v(1, 1, 1), X.v(1, 1, 1)
    s.add_sphere((0, 0, 29.6), 100.0, _sc.material(diffuse=kd, ambient=ka, specular=ks, phong=20.0))
    lpos, lint, amb = X.v(5.79, 0, 0), X.v(1e5, 1e5, 1e5), X.v(15, 15, 15)
    s.add_light(lpos, 1e5); s.ambient[:] = amb
    O, D, C = X.v(0, 0, 0), X.v(0, 0, 1), X.v(0, 0, 29.6)
    oc = X.vsub(O, C)
    uoc = X.dot(D, oc)
    disc = X.sub(X.mul(uoc, uoc), X.sub(X.dot(oc, oc), F(100)))
    t = X.sub(-uoc, X.sqrt(disc))                                  # smallRoot = -uoc - sqrt(disc) >= 0
    P = X.vadd(O, X.vscale(D, t))                                  # Ray.GetPoint, Ray.cs:18-21
    N = X.normalize(X.vsub(P, C))                                  # GetSphereNormal, RTS:402-407
    V = X.normalize(X.vsub(O, P))                                  # RTS:325
    color = X.vmul(amb, ka)                                        # CalculateAmbient, RTS:438-441
    color = X.vadd(color, _light_terms(P, N, V, lpos, lint, kd, ks, F(20)))
    return dict(scene=s, params=_one_pixel_params(0), hit=(1, 0), t=t, value=color, shadow_rays=1, mirror_rays=0)


def kat6_mirror_depth1():
    """KAT-6: a mirror sphere in front of the camera and a green sphere behind it; MaxReflectionBounces 1.  The centre ray hits
    the mirror at P0 = (0,0,t0); Reflect (RTS:368-373) gives origin P0 + N*1e-4 and direction 2*N*dot(V,N) - V = (0,0,-1), which
    hits the second sphere; colour = local(P0) + km * Shade(reflectRay, 1).Value (RTS:358-362)."""
    s = _sc.SceneArrays()
    kd0, ka0, ks0, km0 = X.v(1, 0, 0), X.v(1, 1, 1), X.v(1, 1, 1), X.v(0.5, 0.25, 0.75)
    kd1, ka1, ks1 = X.v(0, 1, 0), X.v(0.5, 0.5, 0.5), X.v(0, 0, 0)
    C0, C1 = X.v(0, 0, 29.6), X.v(0, 0, -30)
    s.add_sphere(C0, 100.0, _sc.material(diffuse=kd0, ambient=ka0, specular=ks0, mirror=km0, phong=20.0, is_mirror=1))
    s.add_sphere(C1, 100.0, _sc.material(diffuse=kd1, ambient=ka1, specular=ks1, phong=3.0))
    lpos, lint, amb = X.v(5.79, 0, 0), X.v(1e5, 1e5, 1e5), X.v(15, 15, 15)
    s.add_light(lpos, 1e5); s.ambient[:] = amb

    def sphere_t(O, D, C):                                         # RMath.cs:85-106, small root
        oc = X.vsub(O, C)
        uoc = X.dot(D, oc)
        disc = X.sub(X.mul(uoc, uoc), X.sub(X.dot(oc, oc), F(100)))
        assert disc >= 0
        small = X.sub(-uoc, X.sqrt(disc))
        assert small >= 0
        return small
    O, D = X.v(0, 0, 0), X.v(0, 0, 1)
    t0 = sphere_t(O, D, C0)
    P0 = X.vadd(O, X.vscale(D, t0))
    N0 = X.normalize(X.vsub(P0, C0))
    V0 = X.normalize(X.vsub(O, P0))
    c0 = X.vadd(X.vmul(amb, ka0), _light_terms(P0, N0, V0, lpos, lint, kd0, ks0, F(20)))
    RO = X.vadd(P0, X.vscale(N0, F(0.0001)))                       # RTS:370
    vn = X.dot(V0, N0)
    RD = X.vsub(X.vmul(tuple(X.mul(F(2), n) for n in N0), (vn, vn, vn)), V0)     # 2 * surfaceNormal * dot(rayDirection, surfaceNormal) - rayDirection, RTS:371
    t1 = sphere_t(RO, RD, C1)
    P1 = X.vadd(RO, X.vscale(RD, t1))
    N1 = X.normalize(X.vsub(P1, C1))
    V1 = X.normalize(X.vsub(RO, P1))
    ldir1 = X.normalize(X.vsub(lpos, P1))
    ldn1 = X.dot(ldir1, N1)
    assert ldn1 > 0
    dl = X.vsub(lpos, P1)
    d21 = X.dot(dl, dl)
    E1 = X.vdiv(lint, d21)
    diffuse1 = X.vmul(X.vscale(kd1, ldn1), E1)
    vv = X.vadd(ldir1, V1)
    H1 = X.vdiv(vv, X.sqrt(X.dot(vv, vv)))
    spec1 = X.vmul(X.vscale(ks1, X.powf(X.dot(N1, H1), F(3))), E1)  # ks = 0: the term is computed and adds +0 (RTS:399)
    c1 = X.vadd(X.vmul(amb, ka1), X.vadd(diffuse1, spec1))
    color = X.vadd(c0, X.vmul(km0, c1))                            # color += new Rgb(mirrorReflectance * Shade(...).Value), RTS:362
    return dict(scene=s, params=_one_pixel_params(1), hit=(1, 0), t=t0, value=color, shadow_rays=2, mirror_rays=1)


def kat7_slab_zero_direction():
    """KAT-7: RMath.RayAABBIntersection (RMath.cs:12-26) with a zero direction component while the origin lies ON a slab plane:
    t1 = (Min.x - O.x) * rcp(0) = 0 * inf = NaN.  math.max(NaN, tmin) returns tmin (NaN is the first argument, the comparison is
    false), math.min(NaN, tmax) returns tmax: the NaN is neutral, the ray counts as inside that slab.
    Returns (origin, direction, box6, expected) rows, worked out by hand in the comments."""
    box = (1, -1, 5, 2, 1, 6)
    return [
        # x: t1 = 0*inf = NaN, t2 = 1*inf = inf -> tmin = min(max(NaN,0)=0, max(inf,0)=inf) = 0, tmax = max(min(NaN,inf)=inf, inf) = inf
        # y: t1 = -inf, t2 = inf -> tmin = min(0, inf) = 0, tmax = max(-inf, inf) = inf;  z: t1 = 5, t2 = 6 -> tmin 5, tmax 6: hit
        ((1, 0, 0), (0, 0, 1), box, True),
        # origin on the Max.x plane: t1 = -inf, t2 = NaN -> tmin = min(0, max(NaN,0)=0) = 0, tmax = max(-inf, min(NaN,inf)=inf) = inf: hit
        ((2, 0, 0), (0, 0, 1), box, True),
        # origin outside the x slab: t1 = t2 = -inf -> tmax = -inf < tmin = 0: miss
        ((2.5, 0, 0), (0, 0, 1), box, False),
        # negative zero: rcp(-0) = -inf; t1 = 0*-inf = NaN, t2 = -inf -> tmin = min(0, max(-inf,0)=0) = 0, tmax = max(min(NaN,inf)=inf, min(-inf,inf)=-inf) = inf: hit
        ((1, 0, 0), (-0.0, 0, 1), box, True),
        # box behind the ray: z: t1 = -5, t2 = -6... Min.z 5, Max.z 6, D.z = -1: t1 = -5, t2 = -6 -> tmin = min(max(-5,0), max(-6,0)) = 0, tmax = max(-5, -6) = -5: miss
        ((1.5, 0, 0), (0, 0, -1), box, False),
    ]


def kat8_normalize_is_rsqrt_times_x():
    """KAT-8: math.normalize(v) = rsqrt(dot(v, v)) * v with rsqrt(x) = 1 / sqrt(x) (Unity.Mathematics 1.2.6): two roundings
    more than v / length(v), and a different last bit for many vectors.  Returns vectors where the two forms differ, with the
    rsqrt form (what RTS:297,325,330,405 compute) as the expected result."""
    out = []
    for vec in ((1, 2, 3), (3, 1, 7), (5, -2, 11), (0.3, 0.7, -0.2), (13, 17, 19), (1, 1, 1), (2, 3, 6), (7, -9, 4)):
        v = X.v(*vec)
        a = X.normalize(v)
        ln = X.sqrt(X.dot(v, v))
        b = tuple(X.div(x, ln) for x in v)
        if any(x != y for x, y in zip(a, b)):
            out.append((v, a, b))
    assert len(out) >= 3
    return out

// rt_oracle.cpp — CPU restatement of the unity-raytracer per-pixel hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and there
// only as the checker / the timed CPU baseline — never as the thing shipped.
//
// What it restates (all paths under /root/reference/Assets/RayTracer/):
//   Math/RMath.cs:9-108                      slab test, Moller-Trumbore, analytic sphere
//   Data/Objects/Scene.cs:17-122             CalculateAABB, IntersectRay (linear scan, strict '>' tie-break)
//   Data/Camera/ImagePlane.cs:26-44          Center, GetRect (TopLeft only is consumed)
//   Demo-RayTracing/RayTracingSetup.cs:275-455  CastPixelRays, Shade, Reflect, CalculateSpecular/Diffuse/Ambient
//   Data/Shading/Rgb.cs:8-28                 0..255 float colour, .Color = Value/255
//   Data/Objects/{Triangle,Sphere}.cs        Normal, Sphere.AABB
//
// The reference is C# on UnityEngine + com.unity.mathematics 1.2.6 (Packages/manifest.json:8); neither a
// .NET/Mono runtime nor the package source is present, so the reference itself cannot be executed here.
// Arithmetic semantics of Unity.Mathematics are restated from its published math.cs (um_* below).
// PARITY PINNING: the reference's own tests pin exactly one value on this path (Tests/RayTracerTests.cs:8-27,
// ray-sphere t = 299).  Everything else (triangle test, slab test, scan order, shading, ray-gen) is
// "parity unpinned" by the reference: the known answers in tests/golden are hand-derived from the cited
// formulas.  Floating-point mode is DEFINED as strict IEEE binary32 (no FMA contraction, no widening):
// build with -O2 -ffp-contract=off and never -ffast-math.
//
// Two query modes share the same primitive tests and the same (t, scan-order) tie-break:
//   faithful    : the reference's linear scan, Scene.cs:43-122 line by line   (parity authority)
//   accelerated : CPU BVH2 over all primitives, conservative padded boxes, argmin over (t, scan key),
//                 exact scene-/mesh-AABB gates.  Validated against `faithful` in tests.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cfloat>
#include <vector>
#include <thread>
#include <atomic>
#include <algorithm>
#include <limits>
#include <chrono>

namespace {

struct f3 { float x, y, z; };
static inline f3 mk(float x, float y, float z) { return f3{x, y, z}; }
static inline f3 bc(float s) { return f3{s, s, s}; }
static inline f3 operator+(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline f3 operator-(f3 a, f3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline f3 operator*(f3 a, f3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline f3 operator*(f3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
static inline f3 operator*(float s, f3 a) { return mk(s * a.x, s * a.y, s * a.z); }
static inline f3 operator/(f3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
static inline float idx(const f3& a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// ---- Unity.Mathematics 1.2.6 semantics (math.cs) -------------------------
static inline float um_min(float x, float y) { return (std::isnan(y) || x < y) ? x : y; }
static inline float um_max(float x, float y) { return (std::isnan(y) || x > y) ? x : y; }
static inline f3 um_min3(f3 x, f3 y) { return mk(um_min(x.x, y.x), um_min(x.y, y.y), um_min(x.z, y.z)); }
static inline f3 um_max3(f3 x, f3 y) { return mk(um_max(x.x, y.x), um_max(x.y, y.y), um_max(x.z, y.z)); }
static inline float um_dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// cross(x,y) = (x * y.yzx - x.yzx * y).yzx
static inline f3 um_cross(f3 a, f3 b) {
    f3 byzx = mk(b.y, b.z, b.x), ayzx = mk(a.y, a.z, a.x);
    f3 r = a * byzx - ayzx * b;
    return mk(r.y, r.z, r.x);
}
static inline float um_sqrt(float x) { return (float)std::sqrt((double)x); }   // (float)System.Math.Sqrt
static inline float um_rsqrt(float x) { return 1.0f / um_sqrt(x); }
static inline float um_lengthsq(f3 a) { return um_dot(a, a); }
static inline float um_length(f3 a) { return um_sqrt(um_dot(a, a)); }
static inline f3 um_normalize(f3 a) { return um_rsqrt(um_dot(a, a)) * a; }
static inline float um_distancesq(f3 a, f3 b) { return um_lengthsq(b - a); }
static inline f3 um_rcp(f3 a) { return mk(1.0f / a.x, 1.0f / a.y, 1.0f / a.z); }
static inline float um_acos(float x) { return (float)std::acos((double)x); }
static inline float um_pow(float x, float y) { return (float)std::pow((double)x, (double)y); }
static inline float um_degrees(float x) { return x * 57.295779513f; }

struct Ray { f3 o, d; };
struct AABB { f3 mn, mx; };
struct Tri { f3 v0, v1, v2; };
struct Sph { f3 c; float r2; };
struct Mat { f3 kd, ka, km, ks; float phong; int mirror; };   // MaterialData.cs:7-14 (bool -> int)
struct Light { f3 pos, intensity; };                           // PointLightData.cs:7-10

constexpr float kEps = 0.00001f;          // RMath.cs:9
constexpr float kShadowEps = 0.0001f;     // RayTracingSetup.cs:42
enum { T_NONE = 0, T_SPHERE = 1, T_TRI = 2, T_MESHTRI = 3 };   // ObjectType.cs:3-9

// RMath.cs:12-26
static inline bool ray_aabb(const Ray& r, const AABB& b) {
    f3 inv = um_rcp(r.d);
    float tmin = 0.0f, tmax = INFINITY;
    for (int i = 0; i < 3; ++i) {
        float t1 = (idx(b.mn, i) - idx(r.o, i)) * idx(inv, i);
        float t2 = (idx(b.mx, i) - idx(r.o, i)) * idx(inv, i);
        tmin = um_min(um_max(t1, tmin), um_max(t2, tmin));
        tmax = um_max(um_min(t1, tmax), um_min(t2, tmax));
    }
    return tmin <= tmax;
}

// RMath.cs:29-73
static inline bool ray_tri(const Ray& r, const Tri& tr, float& dist) {
    f3 e1 = tr.v1 - tr.v0, e2 = tr.v2 - tr.v0;
    f3 h = um_cross(r.d, e2);
    float a = um_dot(e1, h);
    if (a > -kEps && a < kEps) { dist = 0.f; return false; }
    float f = 1.0f / a;
    f3 s = r.o - tr.v0;
    float u = f * um_dot(s, h);
    if (u < 0.0f || u > 1.0f) { dist = 0.f; return false; }
    f3 q = um_cross(s, e1);
    float v = f * um_dot(r.d, q);
    if (v < 0.0f || u + v > 1.0f) { dist = 0.f; return false; }
    float t = f * um_dot(e2, q);
    if (t > kEps) { dist = t; return true; }
    dist = 0.f;
    return false;
}

// RMath.cs:81-108
static inline bool ray_sphere(const Ray& r, const Sph& s, float& dist) {
    f3 oc = r.o - s.c;
    float uoc = um_dot(r.d, oc);
    float disc = uoc * uoc - (um_lengthsq(oc) - s.r2);
    if (disc < 0) { dist = 0.f; return false; }
    float sq = um_sqrt(disc);
    float big = -uoc + sq;
    if (big < 0) { dist = 0.f; return false; }
    float small = -uoc - sq;
    dist = small < 0 ? big : small;
    return true;
}

struct Mesh {
    std::vector<Tri> tris;
    std::vector<f3> normals;
    Mat mat;
    AABB aabb;
};

struct Hit { int type, index, mesh; float dist; };

struct BNode { AABB box; int left, right, first, count; };   // count>0 -> leaf over prim_order[first..]

struct Scene {
    std::vector<Sph> spheres;  std::vector<Mat> sphere_mats;
    std::vector<Tri> tris;     std::vector<f3> tri_normals;  std::vector<Mat> tri_mats;
    std::vector<Mesh> meshes;
    std::vector<Light> lights;
    f3 ambient{0, 0, 0};
    AABB aabb{};
    bool committed = false;
    // accelerated mode
    std::vector<BNode> nodes;
    std::vector<uint32_t> order;          // scan keys in BVH leaf order
    std::vector<uint32_t> mesh_base;      // scan key of mesh m's triangle 0; mesh_base[n_meshes] = first sphere key
    uint32_t n_meshtris = 0;
    bool bvh_built = false;

    // Scene.cs:17-41
    void calc_aabb() {
        AABB a; a.mn = bc(FLT_MAX); a.mx = bc(-FLT_MAX);      // float.MinValue == -FLT_MAX
        for (auto& m : meshes) { a.mn = um_min3(a.mn, m.aabb.mn); a.mx = um_max3(a.mx, m.aabb.mx); }   // AABB.cs:16-20
        for (auto& t : tris) {
            for (const f3* p : {&t.v0, &t.v1, &t.v2}) { a.mn = um_min3(*p, a.mn); a.mx = um_max3(*p, a.mx); }  // AABB.cs:10-14
        }
        for (auto& s : spheres) {                                // Sphere.cs:17-22
            float r = um_sqrt(s.r2);
            AABB sb{s.c - bc(r), s.c + bc(r)};
            a.mn = um_min3(a.mn, sb.mn); a.mx = um_max3(a.mx, sb.mx);
        }
        aabb = a;
    }

    // Scene.cs:43-122, line by line (including the stale MeshIndex on sphere / loose-triangle hits).
    Hit intersect_faithful(const Ray& ray) const {
        float best = FLT_MAX;
        Hit h{T_NONE, -1, -1, best};
        if (!ray_aabb(ray, aabb)) return h;
        for (int mi = 0; mi < (int)meshes.size(); ++mi) {
            const Mesh& m = meshes[mi];
            if (ray_aabb(ray, m.aabb)) {
                for (int ti = 0; ti < (int)m.tris.size(); ++ti) {
                    float d;
                    if (ray_tri(ray, m.tris[ti], d)) {
                        if (best > d) { best = d; h.type = T_MESHTRI; h.index = ti; h.mesh = mi; }
                    }
                }
            }
        }
        for (int si = 0; si < (int)spheres.size(); ++si) {
            float d;
            if (ray_sphere(ray, spheres[si], d)) {
                if (best > d) { best = d; h.type = T_SPHERE; h.index = si; }
            }
        }
        for (int ti = 0; ti < (int)tris.size(); ++ti) {
            float d;
            if (ray_tri(ray, tris[ti], d)) {
                if (best > d) { best = d; h.type = T_TRI; h.index = ti; }
            }
        }
        h.dist = best;
        return h;
    }

    // ---- accelerated mode ---------------------------------------------------------------------------
    // scan key: position of the primitive in the reference's test order (Scene.cs:64-115):
    //   mesh triangles by (meshIndex, triIndex) < spheres by index < loose triangles by index.
    uint32_t n_prims() const { return n_meshtris + (uint32_t)spheres.size() + (uint32_t)tris.size(); }

    void key_decode(uint32_t key, int& type, int& index, int& mesh) const {
        if (key < n_meshtris) {
            int m = int(std::upper_bound(mesh_base.begin(), mesh_base.end(), key) - mesh_base.begin()) - 1;
            type = T_MESHTRI; mesh = m; index = int(key - mesh_base[m]);
        } else if (key < n_meshtris + spheres.size()) {
            type = T_SPHERE; index = int(key - n_meshtris); mesh = -1;
        } else {
            type = T_TRI; index = int(key - n_meshtris - spheres.size()); mesh = -1;
        }
    }

    AABB prim_box(uint32_t key) const {
        int type, index, mesh; key_decode(key, type, index, mesh);
        AABB b;
        if (type == T_SPHERE) {
            // RaySphereIntersection (RMath.cs:81-108) computes disc = uoc*uoc - (lengthsq(oc) - r2) in binary32: two numbers of
            // size |oc|^2 cancel down to ~r2, so `disc < 0 -> miss` accepts rays that pass the centre at up to
            // sqrt(r2 + ~23 * 2^-24 * |oc|^2) (measured: 9.5 * 2^-24; a sphere of radius 0.1 seen from 100 away is "hit" out to
            // 1.21 r).  That is the reference's answer, so the box must hold those rays: r_box^2 = r2 + 2^-19 * Rmax^2, Rmax =
            // distance from the centre to the farthest corner of the box all ray origins lie in (obox).
            const Sph& s = spheres[index];
            double rr = (double)s.r2;
            if (obox_valid) {
                double m2 = 0.0;
                for (int k = 0; k < 3; ++k) {
                    double d = std::fmax(std::fabs((double)idx(s.c, k) - idx(obox.mn, k)), std::fabs((double)idx(s.c, k) - idx(obox.mx, k)));
                    m2 += d * d;
                }
                rr += std::ldexp(m2, -19);
            }
            float r = (rr >= 0.0 && rr < 1e37) ? (float)std::sqrt(rr) * 1.0001f + 1e-6f : std::sqrt(s.r2) * 1.0001f + 1e-6f;
            b.mn = s.c - bc(r); b.mx = s.c + bc(r);
        } else {
            const Tri& t = (type == T_MESHTRI) ? meshes[mesh].tris[index] : tris[index];
            b.mn = mk(std::fmin(t.v0.x, std::fmin(t.v1.x, t.v2.x)), std::fmin(t.v0.y, std::fmin(t.v1.y, t.v2.y)),
                      std::fmin(t.v0.z, std::fmin(t.v1.z, t.v2.z)));
            b.mx = mk(std::fmax(t.v0.x, std::fmax(t.v1.x, t.v2.x)), std::fmax(t.v0.y, std::fmax(t.v1.y, t.v2.y)),
                      std::fmax(t.v0.z, std::fmax(t.v1.z, t.v2.z)));
        }
        return b;
    }

    float pad = 0.f;
    // box of every ray origin the BVH must serve: the primitive bounds (shadow / mirror rays start on surfaces) united with
    // the explicit origins (camera, ora_intersect rays) seen so far; see prim_box
    AABB obox{bc(0.f), bc(0.f)}; bool obox_valid = false;
    AABB extra{bc(0.f), bc(0.f)}; bool extra_valid = false;

    bool origin_covered(const f3& o) const {
        if (spheres.empty() || !bvh_built) return true;
        for (int k = 0; k < 3; ++k) {
            float v = idx(o, k);
            if (!std::isfinite(v)) return true;                   // nothing to cover: such a ray hits nothing
            if (v < idx(obox.mn, k) || v > idx(obox.mx, k)) return false;
        }
        return true;
    }
    // makes the accelerated mode valid for rays starting at o (rebuilds the BVH when the sphere boxes were too small for it)
    void cover_origin(const f3& o) {
        if (origin_covered(o)) return;
        if (!extra_valid) { extra.mn = o; extra.mx = o; extra_valid = true; }
        extra.mn = mk(std::fmin(extra.mn.x, o.x), std::fmin(extra.mn.y, o.y), std::fmin(extra.mn.z, o.z));
        extra.mx = mk(std::fmax(extra.mx.x, o.x), std::fmax(extra.mx.y, o.y), std::fmax(extra.mx.z, o.z));
        build_bvh();
    }

    void build_bvh() {
        mesh_base.clear(); n_meshtris = 0;
        for (auto& m : meshes) { mesh_base.push_back(n_meshtris); n_meshtris += (uint32_t)m.tris.size(); }
        mesh_base.push_back(n_meshtris);
        uint32_t n = n_prims();
        {   // origin box = un-inflated bounds of everything (+ margin) united with the explicit origins
            obox_valid = false;
            AABB b{bc(FLT_MAX), bc(-FLT_MAX)};
            for (uint32_t i = 0; i < n; ++i) {
                AABB pb = prim_box(i);
                b.mn = mk(std::fmin(b.mn.x, pb.mn.x), std::fmin(b.mn.y, pb.mn.y), std::fmin(b.mn.z, pb.mn.z));
                b.mx = mk(std::fmax(b.mx.x, pb.mx.x), std::fmax(b.mx.y, pb.mx.y), std::fmax(b.mx.z, pb.mx.z));
            }
            if (extra_valid) {
                b.mn = mk(std::fmin(b.mn.x, extra.mn.x), std::fmin(b.mn.y, extra.mn.y), std::fmin(b.mn.z, extra.mn.z));
                b.mx = mk(std::fmax(b.mx.x, extra.mx.x), std::fmax(b.mx.y, extra.mx.y), std::fmax(b.mx.z, extra.mx.z));
            }
            float am = 0.f;
            for (int k = 0; k < 3; ++k) { if (std::isfinite(idx(b.mn, k))) am = std::fmax(am, std::fabs(idx(b.mn, k))); if (std::isfinite(idx(b.mx, k))) am = std::fmax(am, std::fabs(idx(b.mx, k))); }
            const float m = 1e-2f * (1.f + am), L = 1e18f;
            auto cl = [&](float v) { return std::fmin(std::fmax(v, -L), L); };
            obox.mn = mk(cl(b.mn.x - m), cl(b.mn.y - m), cl(b.mn.z - m)); obox.mx = mk(cl(b.mx.x + m), cl(b.mx.y + m), cl(b.mx.z + m));
            obox_valid = n > 0 && obox.mn.x <= obox.mx.x && obox.mn.y <= obox.mx.y && obox.mn.z <= obox.mx.z;
        }
        order.resize(n);
        std::vector<AABB> boxes(n);
        std::vector<f3> cent(n);
        float amax = 0.f;
        for (uint32_t i = 0; i < n; ++i) {
            order[i] = i; boxes[i] = prim_box(i);
            cent[i] = (boxes[i].mn + boxes[i].mx) * 0.5f;
            for (int k = 0; k < 3; ++k) {
                amax = std::fmax(amax, std::fabs(idx(boxes[i].mn, k)));
                amax = std::fmax(amax, std::fabs(idx(boxes[i].mx, k)));
            }
        }
        // generous conservative padding (4x what the device BVH uses, amax/65536): the accelerated oracle must never
        // cull a primitive whose computed t could win.  Cost is irrelevant here.
        pad = amax * (1.0f / 16384.0f) + 1e-5f;
        nodes.clear(); nodes.reserve(n ? 2 * n / 3 + 16 : 1);
        if (n == 0) { bvh_built = true; return; }
        struct Job { int node; uint32_t lo, hi; };
        nodes.push_back(BNode{});
        std::vector<Job> stack{{0, 0, n}};
        while (!stack.empty()) {
            Job j = stack.back(); stack.pop_back();
            AABB b{bc(FLT_MAX), bc(-FLT_MAX)}, cb{bc(FLT_MAX), bc(-FLT_MAX)};
            for (uint32_t i = j.lo; i < j.hi; ++i) {
                const AABB& pb = boxes[order[i]];
                b.mn = mk(std::fmin(b.mn.x, pb.mn.x), std::fmin(b.mn.y, pb.mn.y), std::fmin(b.mn.z, pb.mn.z));
                b.mx = mk(std::fmax(b.mx.x, pb.mx.x), std::fmax(b.mx.y, pb.mx.y), std::fmax(b.mx.z, pb.mx.z));
                const f3& c = cent[order[i]];
                cb.mn = mk(std::fmin(cb.mn.x, c.x), std::fmin(cb.mn.y, c.y), std::fmin(cb.mn.z, c.z));
                cb.mx = mk(std::fmax(cb.mx.x, c.x), std::fmax(cb.mx.y, c.y), std::fmax(cb.mx.z, c.z));
            }
            BNode nd{};
            nd.box.mn = b.mn - bc(pad); nd.box.mx = b.mx + bc(pad);
            uint32_t cnt = j.hi - j.lo;
            if (cnt <= 4) { nd.first = (int)j.lo; nd.count = (int)cnt; nodes[j.node] = nd; continue; }
            f3 ext = cb.mx - cb.mn;
            int ax = 0; if (ext.y > ext.x) ax = 1; if (ext.z > idx(ext, ax)) ax = 2;
            float split = idx(cb.mn, ax) + idx(ext, ax) * 0.5f;
            {   // binned surface-area heuristic over the three axes (16 bins); the midpoint of the longest axis is the fallback.
                // Only the tree changes, never a result (same primitive tests, same (t, key) argmin): a fairer CPU baseline than
                // the midpoint tree of round 1 ("reference-arm bias").
                const int NB = 16;
                float best_cost = FLT_MAX;
                for (int a = 0; a < 3; ++a) {
                    const float e = idx(ext, a);
                    if (!(e > 0.f) || !std::isfinite(e)) continue;
                    int bc_[NB] = {0}; AABB bb[NB];
                    for (int k = 0; k < NB; ++k) bb[k] = AABB{bc(FLT_MAX), bc(-FLT_MAX)};
                    const float sc = (float)NB / e, c0 = idx(cb.mn, a);
                    for (uint32_t i = j.lo; i < j.hi; ++i) {
                        int k = (int)((idx(cent[order[i]], a) - c0) * sc); k = k < 0 ? 0 : (k >= NB ? NB - 1 : k);
                        const AABB& pb = boxes[order[i]];
                        ++bc_[k];
                        bb[k].mn = mk(std::fmin(bb[k].mn.x, pb.mn.x), std::fmin(bb[k].mn.y, pb.mn.y), std::fmin(bb[k].mn.z, pb.mn.z));
                        bb[k].mx = mk(std::fmax(bb[k].mx.x, pb.mx.x), std::fmax(bb[k].mx.y, pb.mx.y), std::fmax(bb[k].mx.z, pb.mx.z));
                    }
                    auto area = [](const AABB& q) { f3 d = q.mx - q.mn; return (d.x < 0 || d.y < 0 || d.z < 0) ? 0.f : d.x * d.y + d.y * d.z + d.z * d.x; };
                    float ra[NB]; int rc[NB]; AABB acc{bc(FLT_MAX), bc(-FLT_MAX)}; int c = 0;
                    for (int k = NB - 1; k >= 1; --k) {
                        if (bc_[k]) { acc.mn = mk(std::fmin(acc.mn.x, bb[k].mn.x), std::fmin(acc.mn.y, bb[k].mn.y), std::fmin(acc.mn.z, bb[k].mn.z));
                                      acc.mx = mk(std::fmax(acc.mx.x, bb[k].mx.x), std::fmax(acc.mx.y, bb[k].mx.y), std::fmax(acc.mx.z, bb[k].mx.z)); }
                        c += bc_[k]; ra[k] = area(acc); rc[k] = c;
                    }
                    acc = AABB{bc(FLT_MAX), bc(-FLT_MAX)}; c = 0;
                    for (int k = 0; k < NB - 1; ++k) {
                        if (bc_[k]) { acc.mn = mk(std::fmin(acc.mn.x, bb[k].mn.x), std::fmin(acc.mn.y, bb[k].mn.y), std::fmin(acc.mn.z, bb[k].mn.z));
                                      acc.mx = mk(std::fmax(acc.mx.x, bb[k].mx.x), std::fmax(acc.mx.y, bb[k].mx.y), std::fmax(acc.mx.z, bb[k].mx.z)); }
                        c += bc_[k];
                        if (c == 0 || rc[k + 1] == 0) continue;
                        const float cost = area(acc) * (float)c + ra[k + 1] * (float)rc[k + 1];
                        if (cost < best_cost) { best_cost = cost; ax = a; split = c0 + (float)(k + 1) / sc; }
                    }
                }
            }
            uint32_t mid = uint32_t(std::partition(order.begin() + j.lo, order.begin() + j.hi,
                               [&](uint32_t k) { return idx(cent[k], ax) < split; }) - order.begin());
            if (mid == j.lo || mid == j.hi) {
                mid = j.lo + cnt / 2;
                std::nth_element(order.begin() + j.lo, order.begin() + mid, order.begin() + j.hi,
                                 [&](uint32_t a, uint32_t c) { return idx(cent[a], ax) < idx(cent[c], ax); });
            }
            nd.left = (int)nodes.size(); nd.right = nd.left + 1; nd.count = 0;
            nodes[j.node] = nd;
            nodes.push_back(BNode{}); nodes.push_back(BNode{});
            stack.push_back({nd.left, j.lo, mid});
            stack.push_back({nd.right, mid, j.hi});
        }
        pref.resize(n);                                          // leaf order -> {type, index, mesh}: no per-test key decode
        for (uint32_t i = 0; i < n; ++i) key_decode(order[i], pref[i].type, pref[i].index, pref[i].mesh);
        bvh_built = true;
    }
    struct PRef { int type, index, mesh; };
    std::vector<PRef> pref;

    // conservative slab test for BVH nodes: returns entry distance, or +inf when missed.
    // dl: every plane moved outward by this much for this ray — the rounding of (plane - o) and of the reference's own
    // fl(O - v0) grows with |o|, a fixed padding does not (a camera 1e4 scene extents away; rt_core.h rt_far_pad)
    static inline float node_entry(const f3& o, const f3& inv, const AABB& b, float tbest, float dl) {
        float t0 = 0.f, t1 = tbest;
        for (int i = 0; i < 3; ++i) {
            float a = ((idx(b.mn, i) - dl) - idx(o, i)) * idx(inv, i);
            float c = ((idx(b.mx, i) + dl) - idx(o, i)) * idx(inv, i);
            float lo = std::fmin(a, c), hi = std::fmax(a, c);     // fmin/fmax ignore a NaN operand
            t0 = std::fmax(t0, lo); t1 = std::fmin(t1, hi);
        }
        return (t0 <= t1) ? t0 : INFINITY;
    }

    // closest hit with the reference tie-break: argmin over (t, scan key); the same exact gates.
    // any_tmax2 >= 0 switches to the shadow query: returns as soon as some hit has t*t < any_tmax2
    // (result-equivalent to closest-hit-then-compare, RayTracingSetup.cs:333-345, because every reported t >= 0).
    Hit intersect_accel(const Ray& ray, std::vector<int8_t>& gate_cache, float any_tmax2 = -1.f) const {
        Hit h{T_NONE, -1, -1, FLT_MAX};
        if (!ray_aabb(ray, aabb)) return h;                      // Scene.cs:54
        if (nodes.empty()) return h;
        std::fill(gate_cache.begin(), gate_cache.end(), int8_t(-1));
        f3 inv = um_rcp(ray.d);
        float best = FLT_MAX; uint32_t best_key = 0xffffffffu;
        float cull = FLT_MAX;
        if (any_tmax2 >= 0.f) cull = std::sqrt(any_tmax2) * 1.001f + pad;
        int stack[256]; int sp = 0;
        float dl = std::fmax(std::fmax(std::fabs(ray.o.x), std::fabs(ray.o.y)), std::fabs(ray.o.z)) * 0x1p-17f;
        if (!(dl == dl)) dl = INFINITY;
        if (node_entry(ray.o, inv, nodes[0].box, cull, dl) == INFINITY) return h;
        stack[sp++] = 0;
        while (sp) {
            const BNode& nd = nodes[stack[--sp]];
            float lim = std::fmin(cull, best * 1.001f + pad);
            if (nd.count == 0) {                       // children tested before the push, nearer child popped first
                float el = node_entry(ray.o, inv, nodes[nd.left].box, lim, dl), er = node_entry(ray.o, inv, nodes[nd.right].box, lim, dl);
                if (el != INFINITY && er != INFINITY) {
                    if (el <= er) { stack[sp++] = nd.right; stack[sp++] = nd.left; } else { stack[sp++] = nd.left; stack[sp++] = nd.right; }
                } else if (el != INFINITY) stack[sp++] = nd.left;
                else if (er != INFINITY) stack[sp++] = nd.right;
                continue;
            }
            if (node_entry(ray.o, inv, nd.box, lim, dl) == INFINITY) continue;     // best may have shrunk since the push
            for (int i = 0; i < nd.count; ++i) {
                const uint32_t key = order[nd.first + i];
                const PRef& pr = pref[nd.first + i];
                const int type = pr.type, index = pr.index, mesh = pr.mesh;
                float d; bool hit;
                if (type == T_SPHERE) hit = ray_sphere(ray, spheres[index], d);
                else if (type == T_TRI) hit = ray_tri(ray, tris[index], d);
                else {
                    hit = ray_tri(ray, meshes[mesh].tris[index], d);
                    if (hit) {                                   // Scene.cs:67 mesh gate, exact formula
                        if (gate_cache[mesh] < 0) gate_cache[mesh] = ray_aabb(ray, meshes[mesh].aabb) ? 1 : 0;
                        hit = gate_cache[mesh] == 1;
                    }
                }
                if (!hit) continue;
                if (any_tmax2 >= 0.f) {
                    if (d * d < any_tmax2) { h.type = type; h.index = index; h.mesh = mesh; h.dist = d; return h; }
                    continue;
                }
                if (d < best || (d == best && key < best_key)) { best = d; best_key = key; }
            }
        }
        if (any_tmax2 >= 0.f) return h;          // no occluder in front of the light
        if (best_key != 0xffffffffu) { key_decode(best_key, h.type, h.index, h.mesh); h.dist = best; }
        return h;
    }
};

struct Params {               // mirrors rt_render_params (include/rtb200.h)
    float pos[3], fwd[3], right[3], up[3];
    int res_x, res_y; float dist, half_h, half_v;
    float background_rgb[3];
    int max_reflection_bounces;
    int spp_grid;
};

struct Stats { uint64_t primary, shadow, mirror; };

struct Tracer {
    const Scene& sc; const Params& p; int mode;           // 0 faithful, 1 accelerated
    std::vector<int8_t> gate_cache;
    Stats st{0, 0, 0};
    f3 bg;
    Tracer(const Scene& s, const Params& pp, int m) : sc(s), p(pp), mode(m), gate_cache(s.meshes.size()) {
        bg = mk(p.background_rgb[0], p.background_rgb[1], p.background_rgb[2]) * 255.f;   // Rgb.cs:15-18
    }
    Hit closest(const Ray& r) { return mode == 0 ? sc.intersect_faithful(r) : sc.intersect_accel(r, gate_cache); }

    // RayTracingSetup.cs:304-366
    f3 shade(const Ray& ray, int bounce, Hit* primary_out) {
        Hit hit = closest(ray);
        if (primary_out) *primary_out = hit;
        if (hit.type == T_NONE) return bg;
        f3 P = ray.o + ray.d * hit.dist;                         // Ray.cs:18-21
        f3 N; const Mat* m;
        switch (hit.type) {                                      // RayTracingSetup.cs:409-436
            case T_SPHERE: N = um_normalize(P - sc.spheres[hit.index].c); m = &sc.sphere_mats[hit.index]; break;
            case T_TRI:    N = sc.tri_normals[hit.index]; m = &sc.tri_mats[hit.index]; break;
            default:       N = sc.meshes[hit.mesh].normals[hit.index]; m = &sc.meshes[hit.mesh].mat; break;
        }
        f3 color = sc.ambient * m->ka;                           // :438-441
        f3 V = um_normalize(ray.o - P);                          // :325
        for (const Light& L : sc.lights) {
            f3 Ldir = um_normalize(L.pos - P);
            Ray sray{P + N * kShadowEps, Ldir};
            float d2 = um_distancesq(P, L.pos);
            ++st.shadow;
            if (mode == 0) {
                Hit sh = sc.intersect_faithful(sray);
                if (sh.type != T_NONE) { float hd2 = sh.dist * sh.dist; if (hd2 < d2) continue; }
            } else {
                Hit sh = sc.intersect_accel(sray, gate_cache, d2);
                if (sh.type != T_NONE) continue;
            }
            f3 E = L.intensity / d2;                             // :350
            // CalculateDiffuse :443-455
            float cosNL = um_max(0.f, um_dot(Ldir, N));
            f3 diff = m->kd * cosNL * E;
            // CalculateSpecular :375-400
            f3 spec;
            float ldn = um_dot(Ldir, N);
            float angle = um_degrees(um_acos(ldn));
            if (angle > 90.f) spec = bc(0.f);
            else {
                f3 v = Ldir + V;
                f3 H = v / um_length(v);
                float cosNH = um_max(0.f, um_dot(N, H));
                spec = m->ks * um_pow(cosNH, m->phong) * E;
            }
            color = color + (diff + spec);                       // :355
        }
        if (m->mirror && bounce < p.max_reflection_bounces) {    // :358
            Ray rr{P + N * kShadowEps, 2.f * N * um_dot(V, N) - V};   // :368-373, direction NOT renormalised
            ++st.mirror;
            color = color + m->km * shade(rr, bounce + 1, nullptr);
        }
        return color;
    }
};

static inline uint8_t quant(float v) {           // our definition (reference has none, Rgb.cs:12-13): clamp(rint(v),0,255)
    if (std::isnan(v)) return 0;
    float r = std::nearbyintf(v);                // round-half-even under the default rounding mode
    if (r < 0.f) r = 0.f; if (r > 255.f) r = 255.f;
    return (uint8_t)r;
}

struct Outputs { float* rgba32f; uint8_t* rgba8; int32_t* hit_id; float* hit_t; };

// One pixel: ImagePlane.cs:26-44 + RayTracingSetup.cs:275-302 (+ the spp_grid extension).
static void render_pixel(Tracer& tr, int x, int y, size_t out_slot, const Outputs& o) {
    const Params& p = tr.p;
    f3 pos = mk(p.pos[0], p.pos[1], p.pos[2]), fwd = mk(p.fwd[0], p.fwd[1], p.fwd[2]);
    f3 right = mk(p.right[0], p.right[1], p.right[2]), up = mk(p.up[0], p.up[1], p.up[2]);
    f3 center = pos + fwd * p.dist;
    f3 halfUp = up * p.half_v, halfRight = right * p.half_h;
    f3 topLeft = center - halfRight + halfUp;
    float H = p.half_h * 2.f, Vv = p.half_v * 2.f;
    int g = p.spp_grid < 1 ? 1 : p.spp_grid;
    f3 acc = bc(0.f); bool first = true; Hit ph{T_NONE, -1, -1, FLT_MAX};
    for (int j = 0; j < g; ++j) for (int i = 0; i < g; ++i) {
        float ox = ((float)i + 0.5f) / (float)g, oy = ((float)j + 0.5f) / (float)g;
        float rightMove = ((float)x + ox) * H / (float)p.res_x;
        float downMove = ((float)y + oy) * Vv / (float)p.res_y;
        f3 pp = topLeft + rightMove * right - up * downMove;
        Ray ray{pos, um_normalize(pp - pos)};
        ++tr.st.primary;
        Hit h;
        f3 c = tr.shade(ray, 0, &h);
        if (first) { acc = c; ph = h; first = false; } else acc = acc + c;
    }
    if (g > 1) acc = acc / (float)(g * g);
    if (o.rgba32f) { float* q = o.rgba32f + 4 * out_slot; q[0] = acc.x / 255.f; q[1] = acc.y / 255.f; q[2] = acc.z / 255.f; q[3] = 1.f; }
    if (o.rgba8) { uint8_t* q = o.rgba8 + 4 * out_slot; q[0] = quant(acc.x); q[1] = quant(acc.y); q[2] = quant(acc.z); q[3] = 255; }
    if (o.hit_id) { int32_t* q = o.hit_id + 3 * out_slot; q[0] = ph.type; q[1] = ph.index; q[2] = ph.mesh; }
    if (o.hit_t) o.hit_t[out_slot] = ph.dist;
}

}  // namespace

extern "C" {

void* ora_scene_new() { return new Scene(); }
void ora_scene_free(void* s) { delete (Scene*)s; }

static Mat mat_from(const float* m) {     // 13 floats + int (as 14 x 4 bytes), same as rt_material
    Mat r; r.kd = mk(m[0], m[1], m[2]); r.ka = mk(m[3], m[4], m[5]); r.km = mk(m[6], m[7], m[8]); r.ks = mk(m[9], m[10], m[11]);
    r.phong = m[12]; int32_t im; std::memcpy(&im, m + 13, 4); r.mirror = im; return r;
}

int ora_scene_spheres(void* s_, const float* c_r2, const float* mats, int n) {
    Scene* s = (Scene*)s_;
    for (int i = 0; i < n; ++i) {
        s->spheres.push_back(Sph{mk(c_r2[4 * i], c_r2[4 * i + 1], c_r2[4 * i + 2]), c_r2[4 * i + 3]});
        s->sphere_mats.push_back(mat_from(mats + 14 * i));
    }
    return 0;
}
int ora_scene_triangles(void* s_, const float* v, const float* nrm, const float* mats, int n) {
    Scene* s = (Scene*)s_;
    for (int i = 0; i < n; ++i) {
        const float* q = v + 9 * i;
        s->tris.push_back(Tri{mk(q[0], q[1], q[2]), mk(q[3], q[4], q[5]), mk(q[6], q[7], q[8])});
        s->tri_normals.push_back(mk(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]));
        s->tri_mats.push_back(mat_from(mats + 14 * i));
    }
    return 0;
}
int ora_scene_mesh(void* s_, const float* v, const float* nrm, int n, const float* mat, const float* aabb) {
    Scene* s = (Scene*)s_;
    Mesh m; m.tris.resize(n); m.normals.resize(n);
    for (int i = 0; i < n; ++i) {
        const float* q = v + 9 * (size_t)i;
        m.tris[i] = Tri{mk(q[0], q[1], q[2]), mk(q[3], q[4], q[5]), mk(q[6], q[7], q[8])};
        m.normals[i] = mk(nrm[3 * (size_t)i], nrm[3 * (size_t)i + 1], nrm[3 * (size_t)i + 2]);
    }
    m.mat = mat_from(mat);
    m.aabb = AABB{mk(aabb[0], aabb[1], aabb[2]), mk(aabb[3], aabb[4], aabb[5])};
    s->meshes.push_back(std::move(m));
    return 0;
}
int ora_scene_lights(void* s_, const float* pos_int, int n, const float* ambient) {
    Scene* s = (Scene*)s_;
    s->lights.clear();
    for (int i = 0; i < n; ++i)
        s->lights.push_back(Light{mk(pos_int[6 * i], pos_int[6 * i + 1], pos_int[6 * i + 2]),
                                  mk(pos_int[6 * i + 3], pos_int[6 * i + 4], pos_int[6 * i + 5])});
    s->ambient = ambient ? mk(ambient[0], ambient[1], ambient[2]) : bc(0.f);
    return 0;
}
// scene_aabb == NULL -> Scene.CalculateAABB (Scene.cs:17-41)
int ora_scene_commit(void* s_, const float* scene_aabb, int build_bvh) {
    Scene* s = (Scene*)s_;
    if (scene_aabb) s->aabb = AABB{mk(scene_aabb[0], scene_aabb[1], scene_aabb[2]), mk(scene_aabb[3], scene_aabb[4], scene_aabb[5])};
    else s->calc_aabb();
    if (build_bvh) s->build_bvh();
    s->committed = true;
    return 0;
}
void ora_scene_get_aabb(void* s_, float* out6) {
    Scene* s = (Scene*)s_;
    out6[0] = s->aabb.mn.x; out6[1] = s->aabb.mn.y; out6[2] = s->aabb.mn.z;
    out6[3] = s->aabb.mx.x; out6[4] = s->aabb.mx.y; out6[5] = s->aabb.mx.z;
}

// primitive KAT entry points
int ora_ray_aabb(const float* o, const float* d, const float* box6) {
    return ray_aabb(Ray{mk(o[0], o[1], o[2]), mk(d[0], d[1], d[2])},
                    AABB{mk(box6[0], box6[1], box6[2]), mk(box6[3], box6[4], box6[5])}) ? 1 : 0;
}
int ora_ray_triangle(const float* o, const float* d, const float* v9, float* t) {
    return ray_tri(Ray{mk(o[0], o[1], o[2]), mk(d[0], d[1], d[2])},
                   Tri{mk(v9[0], v9[1], v9[2]), mk(v9[3], v9[4], v9[5]), mk(v9[6], v9[7], v9[8])}, *t) ? 1 : 0;
}
int ora_ray_sphere(const float* o, const float* d, const float* c_r2, float* t) {
    return ray_sphere(Ray{mk(o[0], o[1], o[2]), mk(d[0], d[1], d[2])}, Sph{mk(c_r2[0], c_r2[1], c_r2[2]), c_r2[3]}, *t) ? 1 : 0;
}
// Triangle.Normal (Triangle.cs:13-21): cross(V2-V0, V1-V0) / length
void ora_triangle_normal(const float* v9, float* n3) {
    f3 v0 = mk(v9[0], v9[1], v9[2]), v1 = mk(v9[3], v9[4], v9[5]), v2 = mk(v9[6], v9[7], v9[8]);
    f3 c = um_cross(v2 - v0, v1 - v0);
    f3 n = c / um_length(c);
    n3[0] = n.x; n3[1] = n.y; n3[2] = n.z;
}
void ora_normalize(const float* v, float* out) { f3 n = um_normalize(mk(v[0], v[1], v[2])); out[0] = n.x; out[1] = n.y; out[2] = n.z; }

// Scene.IntersectRay for n rays. mode 0 faithful / 1 accelerated. out_id: 3 ints per ray, out_t: 1 float.
int ora_intersect(void* s_, const float* o, const float* d, int n, int mode, int32_t* out_id, float* out_t) {
    Scene* s = (Scene*)s_;
    if (!s->committed || (mode == 1 && !s->bvh_built)) return -1;
    std::vector<int8_t> gc(s->meshes.size());
    if (mode == 1) for (int i = 0; i < n; ++i) s->cover_origin(mk(o[3 * i], o[3 * i + 1], o[3 * i + 2]));
    for (int i = 0; i < n; ++i) {
        Ray r{mk(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk(d[3 * i], d[3 * i + 1], d[3 * i + 2])};
        Hit h = mode == 0 ? s->intersect_faithful(r) : s->intersect_accel(r, gc);
        out_id[3 * i] = h.type; out_id[3 * i + 1] = h.index; out_id[3 * i + 2] = h.mesh; out_t[i] = h.dist;
    }
    return 0;
}

// Shade one explicit ray (RayTracingSetup.cs:304-366) -> Rgb.Value (0..255 scale, unclamped).
int ora_shade_ray(void* s_, const Params* p, const float* o, const float* d, int mode, float* rgb_value) {
    Scene* s = (Scene*)s_;
    if (!s->committed || (mode == 1 && !s->bvh_built)) return -1;
    if (mode == 1) s->cover_origin(mk(o[0], o[1], o[2]));
    Tracer tr(*s, *p, mode);
    f3 c = tr.shade(Ray{mk(o[0], o[1], o[2]), mk(d[0], d[1], d[2])}, 0, nullptr);
    rgb_value[0] = c.x; rgb_value[1] = c.y; rgb_value[2] = c.z;
    return 0;
}

// Primary ray of pixel (x,y), sample (i,j) of a g x g grid: RayTracingSetup.cs:288-298.
void ora_primary_ray(const Params* p, int x, int y, int i, int j, float* o3, float* d3, float* pixel_pos3) {
    f3 pos = mk(p->pos[0], p->pos[1], p->pos[2]), fwd = mk(p->fwd[0], p->fwd[1], p->fwd[2]);
    f3 right = mk(p->right[0], p->right[1], p->right[2]), up = mk(p->up[0], p->up[1], p->up[2]);
    f3 center = pos + fwd * p->dist;
    f3 topLeft = center - right * p->half_h + up * p->half_v;
    int g = p->spp_grid < 1 ? 1 : p->spp_grid;
    float ox = ((float)i + 0.5f) / (float)g, oy = ((float)j + 0.5f) / (float)g;
    float rightMove = ((float)x + ox) * (p->half_h * 2.f) / (float)p->res_x;
    float downMove = ((float)y + oy) * (p->half_v * 2.f) / (float)p->res_y;
    f3 pp = topLeft + rightMove * right - up * downMove;
    f3 dir = um_normalize(pp - pos);
    o3[0] = pos.x; o3[1] = pos.y; o3[2] = pos.z; d3[0] = dir.x; d3[1] = dir.y; d3[2] = dir.z;
    if (pixel_pos3) { pixel_pos3[0] = pp.x; pixel_pos3[1] = pp.y; pixel_pos3[2] = pp.z; }
}

// Render. pixels == NULL: the full frame, outputs indexed x + y*resX (RayTracingSetup.cs:243-246).
// pixels != NULL: only those n_pixels pixel indices, outputs indexed by position in `pixels`.
// stats3: {primary, shadow, mirror} IntersectRay-equivalent queries issued. Returns wall seconds in *seconds.
int ora_render(void* s_, const Params* p, int mode, int n_threads, const int64_t* pixels, int64_t n_pixels,
               float* rgba32f, uint8_t* rgba8, int32_t* hit_id, float* hit_t, uint64_t* stats3, double* seconds) {
    Scene* s = (Scene*)s_;
    if (!s->committed || (mode == 1 && !s->bvh_built)) return -1;
    if (p->res_x < 0 || p->res_y < 0) return -2;
    if (mode == 1) s->cover_origin(mk(p->pos[0], p->pos[1], p->pos[2]));
    int64_t total = pixels ? n_pixels : (int64_t)p->res_x * p->res_y;
    Outputs o{rgba32f, rgba8, hit_id, hit_t};
    if (n_threads < 1) n_threads = (int)std::thread::hardware_concurrency();
    if (n_threads < 1) n_threads = 1;
    std::atomic<int64_t> next{0};
    std::vector<Stats> sts(n_threads, Stats{0, 0, 0});
    const int64_t grain = pixels ? 16 : std::max<int64_t>(64, std::min<int64_t>(4096, total / (n_threads * 64) + 1));
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](int tid) {
        Tracer tr(*s, *p, mode);
        for (;;) {
            int64_t b = next.fetch_add(grain);
            if (b >= total) break;
            int64_t e = std::min(total, b + grain);
            for (int64_t k = b; k < e; ++k) {
                int64_t pix = pixels ? pixels[k] : k;
                render_pixel(tr, int(pix % p->res_x), int(pix / p->res_x), (size_t)k, o);
            }
        }
        sts[tid] = tr.st;
    };
    if (n_threads == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t) th.emplace_back(work, t);
        for (auto& t : th) t.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (stats3) {
        stats3[0] = stats3[1] = stats3[2] = 0;
        for (auto& st : sts) { stats3[0] += st.primary; stats3[1] += st.shadow; stats3[2] += st.mirror; }
    }
    return 0;
}

// Largest float x for which degrees((float)acos((double)x)) > 90f (RayTracingSetup.cs:384-391) holds, found by
// bisection over the float ordering; the predicate is monotone in x. Used to pin the device-side constant.
float ora_specular_cutoff() {
    auto pred = [](float x) { return um_degrees(um_acos(x)) > 90.f; };
    float lo = -1.f, hi = 1.f;            // pred(lo) true, pred(hi) false
    while (std::nextafterf(lo, hi) != hi) {
        float mid = lo + (hi - lo) * 0.5f;
        if (mid == lo || mid == hi) mid = std::nextafterf(lo, hi);
        if (pred(mid)) lo = mid; else hi = mid;
    }
    return lo;
}

}  // extern "C"

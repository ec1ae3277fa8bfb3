// rt_core.h — the per-ray hot path as host/device inline code: exact primitive tests, BVH traversal, the
// shading/light loop, ray generation and the pixel<->work-item mapping.
//
// The CUDA kernels (rt_render.cu) instantiate these with a shared-memory traversal stack; tests/emu instantiates the
// very same code on the host (g++ -ffp-contract=off) so the logic can be checked against the oracle without a GPU.
// This file is NOT a CPU fallback: nothing in the shipped library calls it on the host.
//
// Reference citations are into /root/reference/Assets/RayTracer/.
#pragma once
#include "rt_hd.h"

#if defined(__CUDA_ARCH__)
#define RT_LDG4(p) __ldg(p)
// wave buffers and framebuffers are written once and read once: streaming (evict-first) accesses keep them from pushing
// the BVH nodes and primitives out of L2
#define RT_ST_STREAM(p, v) __stcs((p), (v))
#define RT_LD_STREAM(p) __ldcs(p)
#else
#define RT_LDG4(p) (*(p))
#define RT_ST_STREAM(p, v) (*(p) = (v))
#define RT_LD_STREAM(p) (*(p))
#endif

// ---------------------------------------------------------------------------------------------------------------
// Device scene (all pointers are device memory; layouts in DESIGN.md §3)
// ---------------------------------------------------------------------------------------------------------------
struct RtScene {
    const float4* nodes;      // BVH2 nodes, 4 x float4 each:
                              //   n0 = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)   n1 = same for child 1
                              //   n2 = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)   n3 = (ref0, ref1, -, -) int bits
                              //   ref >= 0: node index;  ref < 0: leaf, ~ref = first_slot*8 + (count-1)
    const float4* prims;      // 3 x float4 per primitive, in BVH leaf ("slot") order:
                              //   triangle: (v0.xyz, key) (e1.xyz, 0) (e2.xyz, meshIndex|-1)    e1=v1-v0, e2=v2-v0
                              //   sphere  : (c.xyz,  key) (r2,0,0, 1) (0,0,0, -1)
                              //   key = position in the reference's scan order (Scene.cs:64-115), int bits
    const float4* pinfo;      // per slot: (normal.xyz, material id bits)
    const float4* mats;       // 4 x float4 per material: (kd, phong) (ka, is_mirror bits) (km, 0) (ks, 0)
    const float4* mesh_aabb;  // 2 x float4 per mesh: (Min,0) (Max,0) exactly as the caller passed them (Mesh.cs:12)
    const int32_t* mesh_base; // scan key of triangle 0 of mesh m; [n_meshes] = n_meshtris
    const float4* lights;     // 2 x float4 per light: (Position,0) (Intensity,0)
    int32_t root;             // 0 (internal root) or a leaf ref when n_prims <= leaf_max
    int32_t n_prims, n_meshes, n_meshtris, n_spheres, n_lights;
    int32_t n_slots;          // leaf slots = n_prims, or more when the host builder pre-split references (a triangle then sits in several slots)
    f3 aabb_min, aabb_max;    // Scene.AABB (Scene.cs:15) as given / computed per Scene.cs:17-41
    f3 ambient;               // AmbientLight.Radiance
    float t_slack;            // multiplicative slack on the traversal cull distance (conservative)
    int32_t packet;           // 1: coherent warps traverse as one packet (rt_packet.cuh); 0: always per-lane
    int32_t backcull;         // 1: shadow queries of lights behind the surface may be skipped (rt_shadow_irrelevant; set at commit)
    float coherent_spread2;   // packet mode needs the warp's ray origins within this squared distance of each other
    float origin_reach;       // ray origins within this distance of the world origin are covered by the node padding alone
                              // (= RT_FAR_ORIGIN_FACTOR * pad * 2^16, see rt_far_pad); set at commit
    // ---- instanced meshes (rt_scene_mesh_instance; "Instancing" below).  n_inst = 0 / null pointers without them ----
    const float4* inodes;     // TLAS nodes, then the BLAS nodes of every base mesh: same 64-byte node format; child refs are indices
                              // into this array, BLAS leaf refs are slots of iprims
    const float4* iprims;     // BLAS leaf slots, 3 x float4 per OBJECT-space triangle: (v0, triangle index in its mesh) (v1, 0) (v2, 0)
    const float4* inst;       // RT_INST_F4 x float4 per instance, in TLAS leaf order:
                              //   [0..2] rows of the object->world matrix M (m00 m01 m02 m03)      -> the exact world vertices
                              //   [3..5] rows of the world->object matrix (box culling only)
                              //   [6]    int bits: (BLAS root ref, first BLAS slot, MeshIndex, scan key of triangle 0 = mesh_base[MeshIndex])
                              //   [7]    (virtual-slot base as int bits, material id as int bits, far-pad factor |Minv|, 0)
    const int32_t* inst_vbase;// [n_inst + 1] prefix sums of the instances' triangle counts (TLAS order): a hit on slot s of instance j
                              // is reported as virtual slot n_slots + inst_vbase[j] + (s - first BLAS slot of j)
    int32_t tlas_root, n_inst;
    int32_t n_nodes, n_inodes, n_islots;   // array sizes (nodes, inodes, BLAS slots in iprims): read only by the RT_CHECKED build
};
enum { RT_INST_F4 = 8 };

// Self-checking development build (-DRT_CHECKED; tools/checked_run.sh).  compute-sanitizer is closed on this GPU pool, so every
// index the render kernels form is tested against its bound right where it is used; violations are counted per kind in
// g_rt_check[] (rt_debug_checks reads them) — the product build compiles all of it out.
enum { RT_CHK_STACK = 0, RT_CHK_NODE = 1, RT_CHK_LEAF = 2, RT_CHK_TLAS_LEAF = 3, RT_CHK_TILE = 4, RT_CHK_QUEUE = 5, RT_CHK_VSLOT = 6,
       RT_CHK_PIXEL = 7, RT_CHK_HIT_SLOT = 8, RT_CHK_MESH = 9, RT_CHK_LIGHT = 10, RT_CHK_BAKE = 11, RT_CHK_KINDS = 16 };
#if defined(RT_CHECKED) && defined(__CUDACC__)
static __device__ unsigned g_rt_check[RT_CHK_KINDS];
#endif
#if defined(RT_CHECKED) && defined(__CUDA_ARCH__)
#define RT_CHECK(kind, cond) do { if (!(cond)) atomicAdd(&g_rt_check[kind], 1u); } while (0)
#else
#define RT_CHECK(kind, cond) ((void)0)
#endif

struct RtView {
    f3 pos, fwd, right, up;
    int32_t res_x, res_y;
    float dist, half_h, half_v;
    f3 bg255;                 // new Rgb(BackgroundColor) = rgb * 255 (Rgb.cs:15-18)
    int32_t max_bounces;
    int32_t spp_grid;
    // partition of the frame over ranks/devices: this view owns the row bands b with b % world == rank
    int32_t rank, world, band_rows;
    int32_t tiles_x;          // ceil(res_x / 8)
    uint32_t tiles_x_magic;   // floor(2^32 / tiles_x) (0xffffffff for tiles_x == 1): division by multiplication, rt_view_finish()
    int32_t band_shift;       // log2(band_rows) when it is a power of two, else -1
    int32_t n_local_rows;     // rows owned by this rank
    int32_t n_tiles;          // tiles_x * ceil(n_local_rows / 4); one tile = 8x4 pixels = one warp
    int32_t far_origin;       // 1: the camera lies so far outside the geometry that the node padding no longer covers the rounding
                              // of the slab test (rt_far_pad): primary rays take the FAR form of the per-lane loop
};

// ---------------------------------------------------------------------------------------------------------------
// Exact primitive tests (bit-for-bit the reference's arithmetic)
// ---------------------------------------------------------------------------------------------------------------
// RMath.RayAABBIntersection, Math/RMath.cs:12-26.  inv = rcp(ray.Direction) (RMath.cs:14) is computed once per ray
// (rt_rcp3) and shared by the scene gate, the mesh gates and the traversal.
RT_HD f3 rt_rcp3(f3 D) { return mk3(rt_div(1.0f, D.x), rt_div(1.0f, D.y), rt_div(1.0f, D.z)); }
RT_HD bool rt_ray_aabb_inv(f3 O, f3 inv, f3 mn, f3 mx) {
    float tmin = 0.0f, tmax = INFINITY;
    float t1 = (mn.x - O.x) * inv.x, t2 = (mx.x - O.x) * inv.x;
    tmin = um_min(um_max(t1, tmin), um_max(t2, tmin));
    tmax = um_max(um_min(t1, tmax), um_min(t2, tmax));
    t1 = (mn.y - O.y) * inv.y; t2 = (mx.y - O.y) * inv.y;
    tmin = um_min(um_max(t1, tmin), um_max(t2, tmin));
    tmax = um_max(um_min(t1, tmax), um_min(t2, tmax));
    t1 = (mn.z - O.z) * inv.z; t2 = (mx.z - O.z) * inv.z;
    tmin = um_min(um_max(t1, tmin), um_max(t2, tmin));
    tmax = um_max(um_min(t1, tmax), um_min(t2, tmax));
    return tmin <= tmax;
}
RT_HD bool rt_ray_aabb_exact(f3 O, f3 D, f3 mn, f3 mx) { return rt_ray_aabb_inv(O, rt_rcp3(D), mn, mx); }

// RMath.RayTriangleIntersection, Math/RMath.cs:29-73. e1/e2 are the single subtractions v1-v0, v2-v0 (RMath.cs:34-35)
// precomputed at upload, which is bit-identical to computing them per call.
//
// The reference divides first (f = 1/a) and then rejects on u, v, t.  An IEEE division costs ~12 instructions, and almost
// every tested triangle is rejected, so the rejections are decided on the *undivided* numerators wherever the outcome of
// the reference's own comparison is certain; only candidates that survive are divided and re-tested with the reference's
// exact expressions.  With f = fl(1/a) and x = fl(f*n) the relative error of x against n/a is < 2^-22, hence:
//   sign(n) != sign(a), n/a not tiny     =>  x < 0 for sure           (u < 0, v < 0;  for t: x <= 0, never > eps)
//   |n| > |a| * (1 + 2^-19)              =>  x > 1 for sure           (u > 1, u + v > 1)
//   |n_t| > |a| * tmax                   =>  t > tmax / (1 + 2^-22): the caller's tmax carries a 1.25e-4 slack, so such a
//                                            hit can neither beat the current best nor satisfy the any-hit predicate
// tmax = INFINITY disables the last rule.  Returns exactly what the reference returns for every hit with t <= ~tmax.
RT_HD bool rt_ray_tri(f3 O, f3 D, f3 v0, f3 e1, f3 e2, float tmax, float& t_out) {
    f3 h = um_cross(D, e2);
    float a = um_dot(e1, h);
    if (a > -RT_EPS && a < RT_EPS) return false;
    const float aa = fabsf(a);
    const bool aneg = a < 0.0f;
    const bool sane = aa < 1e10f;                           // n/a cannot underflow to -0 when |n| > 1e-20
    f3 s = O - v0;
    float nu = um_dot(s, h);
    const float a1 = aa * 1.000002f;
    if (fabsf(nu) > a1) return false;                       // u > 1 or u < -1
    if ((nu < 0.0f) != aneg && fabsf(nu) > 1e-20f && sane) return false;          // u < 0
    f3 q = um_cross(s, e1);
    float nv = um_dot(D, q);
    if ((nv < 0.0f) != aneg && fabsf(nv) > 1e-20f && sane) return false;          // v < 0
    if (fabsf(nu) + fabsf(nv) > a1 && (nu < 0.0f) == aneg && (nv < 0.0f) == aneg) return false;   // u + v > 1
    float nt = um_dot(e2, q);
    if ((nt < 0.0f) != aneg || nt == 0.0f) return false;    // t <= 0 (or -0): fails t > eps
    if (fabsf(nt) > aa * tmax) return false;                // t beyond the caller's (slackened) cull distance
    float f = rt_div(1.0f, a);
    float u = f * nu;
    if (u < 0.0f || u > 1.0f) return false;
    float v = f * nv;
    if (v < 0.0f || u + v > 1.0f) return false;
    float t = f * nt;
    if (t > RT_EPS) { t_out = t; return true; }
    return false;
}

// RMath.RaySphereIntersection, Math/RMath.cs:81-108 (assumes |D| = 1; no epsilon: t may be 0)
RT_HD bool rt_ray_sphere(f3 O, f3 D, f3 c, float r2, float& t_out) {
    f3 oc = O - c;
    float uoc = um_dot(D, oc);
    float disc = uoc * uoc - (um_dot(oc, oc) - r2);
    if (disc < 0) return false;
    float sq = rt_sqrt(disc);
    float big = -uoc + sq;
    if (big < 0) return false;
    float small = -uoc - sq;
    t_out = small < 0 ? big : small;
    return true;
}

// ---------------------------------------------------------------------------------------------------------------
// Instancing (rt_scene_mesh_instance).  The reference bakes transform.localToWorldMatrix into world-space vertices for every
// mesh on every frame (SceneComponents/SceneMesh.cs:21-31) and Scene.IntersectRay sees only the baked triangles
// (Data/Objects/Mesh.cs:7-12).  An instance keeps the base mesh in OBJECT space plus the 3x4 matrix and produces the same world
// triangle on the fly, with the arithmetic of the bake —
//     Matrix4x4.MultiplyPoint3x4:  w.x = ((m00*x + m01*y) + m02*z) + m03      (binary32, left to right; SceneMesh.cs:28)
//     edge1 = w1 - w0, edge2 = w2 - w0                                         (RMath.cs:34-35)
//     normal = -(cross(w2 - w0, w1 - w0) / length(...))                        (SceneMesh.cs:45, Triangle.cs:13-21)
// — so the exact triangle test sees bit-identical operands: an instanced mesh gives the same hits, distances and colours as
// the same mesh baked by the caller, bit for bit, at 48 bytes per re-posed mesh instead of 48 bytes per triangle.  Object space
// is used for CULLING only: the ray is carried into it approximately and tested against BLAS boxes padded for that
// approximation (rt_instances.h).  t is preserved by an affine map (the direction is not renormalised), so cull distances
// carry over unchanged.
// ---------------------------------------------------------------------------------------------------------------
RT_HD f3 rt_xform_point(float4 r0, float4 r1, float4 r2, f3 p) {             // MultiplyPoint3x4, exact operation order
    return mk3(((r0.x * p.x + r0.y * p.y) + r0.z * p.z) + r0.w,
               ((r1.x * p.x + r1.y * p.y) + r1.z * p.z) + r1.w,
               ((r2.x * p.x + r2.y * p.y) + r2.z * p.z) + r2.w);
}
// culling-side transforms (any rounding will do: the BLAS padding covers it)
RT_HD f3 rt_xform_point_cull(float4 r0, float4 r1, float4 r2, f3 p) {
    return mk3(rt_fma(r0.x, p.x, rt_fma(r0.y, p.y, rt_fma(r0.z, p.z, r0.w))), rt_fma(r1.x, p.x, rt_fma(r1.y, p.y, rt_fma(r1.z, p.z, r1.w))),
               rt_fma(r2.x, p.x, rt_fma(r2.y, p.y, rt_fma(r2.z, p.z, r2.w))));
}
RT_HD f3 rt_xform_dir_cull(float4 r0, float4 r1, float4 r2, f3 d) {
    return mk3(rt_fma(r0.x, d.x, rt_fma(r0.y, d.y, r0.z * d.z)), rt_fma(r1.x, d.x, rt_fma(r1.y, d.y, r1.z * d.z)),
               rt_fma(r2.x, d.x, rt_fma(r2.y, d.y, r2.z * d.z)));
}
// 1 / x for the object-space slab constants: any value within a few ulps will do (one MUFU.RCP instead of an IEEE division)
RT_HD float rt_rcp_cull(float x) {
#if defined(__CUDA_ARCH__)
    float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
#else
    return 1.0f / x;
#endif
}
// -Triangle.Normal of a world triangle (Triangle.cs:13-21: v = cross(V2 - V0, V1 - V0); v / length(v); SceneMesh.cs:45 negates)
RT_HD f3 rt_tri_normal_neg(f3 w0, f3 w1, f3 w2) {
    f3 c = um_cross(w2 - w0, w1 - w0);
    f3 n = c / um_length(c);
    return mk3(-n.x, -n.y, -n.z);
}
// Stack sentinels of the two-level walk (real leaf refs lie above them: at most 2^28 - 8 slots; RT_PACKET_DONE = 0x80000000)
#define RT_SENT_TLAS ((int32_t)0x80000001)    // the world BVH is done: continue at the TLAS root
#define RT_SENT_EXIT ((int32_t)0x80000002)    // the BLAS of the current instance is done: back to the TLAS, world-space slabs again

// ---------------------------------------------------------------------------------------------------------------
// BVH traversal.  Result-equivalent to the reference's linear scan (Scene.cs:43-122):
//   closest: argmin over all primitives that pass the exact test of (t, scan key)  == strict-'>' update in scan order
//   any    : true iff some primitive passes with t*t < d2  == closest-hit-then-compare (RayTracingSetup.cs:333-345),
//            because every reported t is >= 0 so the closest hit satisfies the predicate whenever any hit does.
// Node boxes are padded (rt_build) and the cull distance carries a slack, so a primitive whose *computed* t could win is
// never culled; the per-mesh AABB gate (Scene.cs:67) is applied with the exact formula to any candidate that would be
// accepted.  The scene AABB gate (Scene.cs:54) is applied by the callers via rt_scene_gate().
// STK: push(int), pop() -> int, empty().
// ---------------------------------------------------------------------------------------------------------------
RT_HD bool rt_scene_gate(const RtScene& S, f3 O, f3 inv) { return rt_ray_aabb_inv(O, inv, S.aabb_min, S.aabb_max); }
// traversal's own (conservative) inverse direction: the exact one, with zero components replaced by a huge finite value
// Visit counters of the instrumented development build (-DRT_COUNT_VISITS, tools/visit_counts.py): per query class
// (closest-hit = 0, any-hit = 10): +0 rays, +1 node visits, +2 primitive tests of the per-lane traversal; +3 warps, +4 rays,
// +5 warp-level node fetches, +6 lane-level node visits, +7 warp-level primitive fetches, +8 lane-level primitive tests of
// the packet traversal, +9 stack entries of the packet traversal dropped on pop (rt_packet_pop).  Compiled out of the product build.
#if defined(RT_COUNT_VISITS) && defined(__CUDACC__)
static __device__ unsigned long long g_rt_visits[20];
#endif
#if defined(RT_COUNT_VISITS) && defined(__CUDA_ARCH__)
#define RT_COUNT(i, n) atomicAdd(&g_rt_visits[i], (unsigned long long)(n))
#else
#define RT_COUNT(i, n) ((void)0)
#endif
#ifdef RT_COUNT_VISITS
#define RT_CNT_PARAM , int cbase
#define RT_CNT_ARG(ANY) , ((ANY) ? 10 : 0)
#define RT_CNT_FWD , cbase
#else
#define RT_CNT_FWD
#define RT_CNT_PARAM
#define RT_CNT_ARG(ANY)
#endif

RT_HD float rt_safe_inv(float d, float inv) { return fabsf(d) > 1e-30f ? inv : copysignf(1e30f, d); }
// Ray origins far outside the geometry.  The node test computes fma(plane, 1/d, -fl(O/d)); its rounding, and the rounding
// of fl(O - v0) through which the reference's own triangle test sees the geometry (equivalent to displacing a triangle by
// ~10 u |O - v0| across the ray, u = 2^-24), are covered by the node padding (max|coordinate| * 2^-16, rt_build) only while
// the origin lies within RT_FAR_ORIGIN_FACTOR times max|coordinate| of the world origin — always true for shadow and mirror
// rays, which start on surfaces.  A camera farther out (RtView::far_origin) sends its primary rays through the FAR form of
// rt_traverse, which moves every slab plane outward by rt_far_pad(O) = 32 u * max|O| for that ray: thin boxes (a cube face
// seen edge-on) are then never culled although the reference's exact test hits the primitive inside.  (Spheres need more than
// that, at any distance: rt_sphere_box_radius, rt_build_steps.h.)
#define RT_FAR_ORIGIN_FACTOR 16.0f
RT_HD float rt_far_pad(f3 O) {
    float d = fmaxf(fmaxf(fabsf(O.x), fabsf(O.y)), fabsf(O.z)) * 0x1p-19f;
    return d == d ? d : INFINITY;
}

// INST: the scene has instanced meshes (S.n_inst > 0 possible): after the world BVH the walk continues through the TLAS
// (S.inodes / S.tlas_root) and, per instance, through the base mesh's BLAS with the ray carried into object space for the box
// tests; every triangle test runs in world space on the on-the-fly baked vertices ("Instancing" above).
template <bool ANY, bool FAR = false, bool INST = false, class STK>
RT_HD bool rt_traverse(const RtScene& S, f3 O, f3 D, f3 inv, float any_d2, STK& stk, float& best_t, int32_t& best_slot) {
    best_t = FLT_MAX; best_slot = -1;
    stk.sp = 0;                       // an any-hit query returns early and leaves entries behind
    if (S.n_prims <= 0 && !(INST && S.n_inst > 0)) return false;
    float idx_ = rt_safe_inv(D.x, inv.x), idy_ = rt_safe_inv(D.y, inv.y), idz_ = rt_safe_inv(D.z, inv.z);
    float oodx = O.x * idx_, oody = O.y * idy_, oodz = O.z * idz_;
    // FAR: every slab plane moved outward by dl in space = by dl * |1/d| along the ray (inf * 0 cannot occur: |1/d| > 0)
    const float dl = FAR ? rt_far_pad(O) : 0.0f;
    float ex = dl * fabsf(idx_), ey = dl * fabsf(idy_), ez = dl * fabsf(idz_);
    uint32_t best_key = 0xffffffffu;
    float tcull = ANY ? rt_sqrt(any_d2) * S.t_slack : INFINITY;
    int32_t gate_mesh = -1; bool gate_ok = true;
    int32_t node = S.root;
    const float4* nodes = S.nodes;
    int32_t level = 0, cur = -1;      // INST: 0 = world BVH, 1 = TLAS, 2 = BLAS of instance `cur`
    if (INST && S.n_inst > 0) { if (S.n_prims > 0) stk.push(RT_SENT_TLAS); else node = RT_SENT_TLAS; }
    RT_COUNT((ANY ? 10 : 0) + 0, 1);
    for (;;) {
        while (node >= 0) {
            RT_COUNT((ANY ? 10 : 0) + 1, 1);
            RT_CHECK(RT_CHK_NODE, node < ((INST && level > 0) ? S.n_inodes : S.n_nodes));
            const float4* np = (INST ? nodes : S.nodes) + 4 * (size_t)node;
            float4 n0 = RT_LDG4(np), n1 = RT_LDG4(np + 1), n2 = RT_LDG4(np + 2), n3 = RT_LDG4(np + 3);
            float ax0 = rt_fma(n0.x, idx_, -oodx), ax1 = rt_fma(n0.y, idx_, -oodx);
            float ay0 = rt_fma(n0.z, idy_, -oody), ay1 = rt_fma(n0.w, idy_, -oody);
            float az0 = rt_fma(n2.x, idz_, -oodz), az1 = rt_fma(n2.y, idz_, -oodz);
            float bx0 = rt_fma(n1.x, idx_, -oodx), bx1 = rt_fma(n1.y, idx_, -oodx);
            float by0 = rt_fma(n1.z, idy_, -oody), by1 = rt_fma(n1.w, idy_, -oody);
            float bz0 = rt_fma(n2.z, idz_, -oodz), bz1 = rt_fma(n2.w, idz_, -oodz);
            float amin, amax, bmin, bmax;
            if (FAR) {
                amin = fmaxf(fmaxf(fminf(ax0, ax1) - ex, fminf(ay0, ay1) - ey), fmaxf(fminf(az0, az1) - ez, 0.0f));
                amax = fminf(fminf(fmaxf(ax0, ax1) + ex, fmaxf(ay0, ay1) + ey), fminf(fmaxf(az0, az1) + ez, tcull));
                bmin = fmaxf(fmaxf(fminf(bx0, bx1) - ex, fminf(by0, by1) - ey), fmaxf(fminf(bz0, bz1) - ez, 0.0f));
                bmax = fminf(fminf(fmaxf(bx0, bx1) + ex, fmaxf(by0, by1) + ey), fminf(fmaxf(bz0, bz1) + ez, tcull));
            } else {
                amin = fmaxf(fmaxf(fminf(ax0, ax1), fminf(ay0, ay1)), fmaxf(fminf(az0, az1), 0.0f));
                amax = fminf(fminf(fmaxf(ax0, ax1), fmaxf(ay0, ay1)), fminf(fmaxf(az0, az1), tcull));
                bmin = fmaxf(fmaxf(fminf(bx0, bx1), fminf(by0, by1)), fmaxf(fminf(bz0, bz1), 0.0f));
                bmax = fminf(fminf(fmaxf(bx0, bx1), fmaxf(by0, by1)), fminf(fmaxf(bz0, bz1), tcull));
            }
            bool ha = amin <= amax, hb = bmin <= bmax;
            int32_t ra = f2i(n3.x), rb = f2i(n3.y);
            if (ha && hb) {
                // closest hit: nearer child first; any hit: the child the ray stays inside longer (order cannot change the answer)
                if (ANY ? (bmax - bmin) > (amax - amin) : bmin < amin) { int32_t t = ra; ra = rb; rb = t; }
                stk.push(rb);
                node = ra;
            } else if (ha) node = ra;
            else if (hb) node = rb;
            else {
                if (stk.empty()) return best_slot >= 0;
                node = stk.pop();
            }
        }
        if (INST && node <= RT_SENT_EXIT) {                         // a sentinel, not a leaf
            if (node == RT_SENT_TLAS) { nodes = S.inodes; level = 1; node = S.tlas_root; continue; }
            idx_ = rt_safe_inv(D.x, inv.x); idy_ = rt_safe_inv(D.y, inv.y); idz_ = rt_safe_inv(D.z, inv.z);     // RT_SENT_EXIT
            oodx = O.x * idx_; oody = O.y * idy_; oodz = O.z * idz_;
            if (FAR) { ex = dl * fabsf(idx_); ey = dl * fabsf(idy_); ez = dl * fabsf(idz_); }
            level = 1;
            if (stk.empty()) return best_slot >= 0;
            node = stk.pop();
            continue;
        }
        if (INST && level == 1) {                                   // TLAS leaf: instances [first, first + cnt)
            const int32_t ref = ~node, first = ref >> 3, cnt = (ref & 7) + 1;
            RT_CHECK(RT_CHK_TLAS_LEAF, first >= 0 && first + cnt <= S.n_inst);
            if (cnt > 1) stk.push(~((first + 1) * 8 + (cnt - 2)));  // the rest of the leaf, after this instance
            stk.push(RT_SENT_EXIT);
            cur = first;
            const float4* ip = S.inst + RT_INST_F4 * (size_t)cur;
            const float4 a0 = RT_LDG4(ip + 3), a1 = RT_LDG4(ip + 4), a2 = RT_LDG4(ip + 5);
            const f3 Oo = rt_xform_point_cull(a0, a1, a2, O), Do = rt_xform_dir_cull(a0, a1, a2, D);
            idx_ = rt_safe_inv(Do.x, rt_rcp_cull(Do.x)); idy_ = rt_safe_inv(Do.y, rt_rcp_cull(Do.y)); idz_ = rt_safe_inv(Do.z, rt_rcp_cull(Do.z));
            oodx = Oo.x * idx_; oody = Oo.y * idy_; oodz = Oo.z * idz_;
            if (FAR) { const float dlo = dl * RT_LDG4(ip + 7).z; ex = dlo * fabsf(idx_); ey = dlo * fabsf(idy_); ez = dlo * fabsf(idz_); }
            level = 2;
            node = f2i(RT_LDG4(ip + 6).x);
            continue;
        }
        // leaf
        {
            int32_t ref = ~node;
            int32_t first = ref >> 3, cnt = (ref & 7) + 1;
            RT_CHECK(RT_CHK_LEAF, first >= 0 && first + cnt <= ((INST && level == 2) ? S.n_islots : S.n_slots));
            for (int32_t k = 0; k < cnt; ++k) {
                int32_t slot = first + k;
                const float4* pp = ((INST && level == 2) ? S.iprims : S.prims) + 3 * (size_t)slot;
                float4 p0 = RT_LDG4(pp), p1 = RT_LDG4(pp + 1), p2 = RT_LDG4(pp + 2);
                RT_COUNT((ANY ? 10 : 0) + 2, 1);
                float t; bool hit;
                uint32_t key; int32_t mesh;
                if (INST && level == 2) {                           // object-space triangle of instance `cur`: bake it, test it in world space
                    const float4* ip = S.inst + RT_INST_F4 * (size_t)cur;
                    const float4 m0 = RT_LDG4(ip), m1 = RT_LDG4(ip + 1), m2 = RT_LDG4(ip + 2), ii = RT_LDG4(ip + 6);
                    const f3 w0 = rt_xform_point(m0, m1, m2, mk3(p0.x, p0.y, p0.z)), w1 = rt_xform_point(m0, m1, m2, mk3(p1.x, p1.y, p1.z)),
                             w2 = rt_xform_point(m0, m1, m2, mk3(p2.x, p2.y, p2.z));
                    hit = rt_ray_tri(O, D, w0, w1 - w0, w2 - w0, tcull, t);
                    key = (uint32_t)(f2i(ii.w) + f2i(p0.w)); mesh = f2i(ii.z);
                    slot = S.n_slots + f2i(RT_LDG4(ip + 7).x) + (slot - f2i(ii.y));      // virtual slot (rt_inst_decode)
                } else {
                    if (f2i(p1.w) != 0) hit = rt_ray_sphere(O, D, mk3(p0.x, p0.y, p0.z), p1.x, t);
                    else hit = rt_ray_tri(O, D, mk3(p0.x, p0.y, p0.z), mk3(p1.x, p1.y, p1.z), mk3(p2.x, p2.y, p2.z), tcull, t);
                    key = (uint32_t)f2i(p0.w); mesh = f2i(p2.w);
                }
                if (!hit) continue;
                bool better = ANY ? (t * t < any_d2) : (t < best_t || (t == best_t && key < best_key));
                if (!better) continue;
                if (mesh >= 0) {                                    // Scene.cs:67
                    RT_CHECK(RT_CHK_MESH, mesh < S.n_meshes);
                    if (mesh != gate_mesh) {
                        gate_mesh = mesh;
                        float4 lo = RT_LDG4(S.mesh_aabb + 2 * mesh), hi = RT_LDG4(S.mesh_aabb + 2 * mesh + 1);
                        gate_ok = rt_ray_aabb_inv(O, inv, mk3(lo.x, lo.y, lo.z), mk3(hi.x, hi.y, hi.z));
                    }
                    if (!gate_ok) continue;
                }
                if (ANY) { best_t = t; best_slot = slot; return true; }
                best_t = t; best_key = key; best_slot = slot;
                tcull = t * S.t_slack;
            }
        }
        if (stk.empty()) return best_slot >= 0;
        node = stk.pop();
    }
}

// The same query without the tree: every slot tested, argmin over (t, scan key), per-mesh gate on acceptance.  For explicit
// rays whose direction is not normalised (rt_intersect): the reference's sphere test then no longer describes the sphere the
// boxes were built for, so nothing may be culled.
template <bool INST = false>
RT_HD bool rt_scan_all(const RtScene& S, f3 O, f3 D, f3 inv, float& best_t, int32_t& best_slot) {
    best_t = FLT_MAX; best_slot = -1;
    uint32_t best_key = 0xffffffffu;
    if (INST) for (int32_t j = 0; j < S.n_inst; ++j) {              // instanced meshes: every triangle, baked on the fly
        const float4* ip = S.inst + RT_INST_F4 * (size_t)j;
        const float4 m0 = RT_LDG4(ip), m1 = RT_LDG4(ip + 1), m2 = RT_LDG4(ip + 2), ii = RT_LDG4(ip + 6);
        const int32_t mesh = f2i(ii.z), cnt = S.inst_vbase[j + 1] - S.inst_vbase[j];
        float4 lo = RT_LDG4(S.mesh_aabb + 2 * mesh), hi = RT_LDG4(S.mesh_aabb + 2 * mesh + 1);
        if (!rt_ray_aabb_inv(O, inv, mk3(lo.x, lo.y, lo.z), mk3(hi.x, hi.y, hi.z))) continue;      // Scene.cs:67
        for (int32_t k = 0; k < cnt; ++k) {
            const float4* pp = S.iprims + 3 * (size_t)(f2i(ii.y) + k);
            float4 p0 = RT_LDG4(pp), p1 = RT_LDG4(pp + 1), p2 = RT_LDG4(pp + 2);
            const f3 w0 = rt_xform_point(m0, m1, m2, mk3(p0.x, p0.y, p0.z)), w1 = rt_xform_point(m0, m1, m2, mk3(p1.x, p1.y, p1.z)),
                     w2 = rt_xform_point(m0, m1, m2, mk3(p2.x, p2.y, p2.z));
            float t;
            if (!rt_ray_tri(O, D, w0, w1 - w0, w2 - w0, INFINITY, t)) continue;
            const uint32_t key = (uint32_t)(f2i(ii.w) + f2i(p0.w));
            if (!(t < best_t || (t == best_t && key < best_key))) continue;
            best_t = t; best_key = key; best_slot = S.n_slots + S.inst_vbase[j] + k;
        }
    }
    for (int32_t slot = 0; slot < S.n_slots; ++slot) {
        const float4* pp = S.prims + 3 * (size_t)slot;
        float4 p0 = RT_LDG4(pp), p1 = RT_LDG4(pp + 1), p2 = RT_LDG4(pp + 2);
        float t; bool hit;
        if (f2i(p1.w) != 0) hit = rt_ray_sphere(O, D, mk3(p0.x, p0.y, p0.z), p1.x, t);
        else hit = rt_ray_tri(O, D, mk3(p0.x, p0.y, p0.z), mk3(p1.x, p1.y, p1.z), mk3(p2.x, p2.y, p2.z), INFINITY, t);
        if (!hit) continue;
        const uint32_t key = (uint32_t)f2i(p0.w);
        if (!(t < best_t || (t == best_t && key < best_key))) continue;
        const int32_t mesh = f2i(p2.w);
        if (mesh >= 0) {                                            // Scene.cs:67
            float4 lo = RT_LDG4(S.mesh_aabb + 2 * mesh), hi = RT_LDG4(S.mesh_aabb + 2 * mesh + 1);
            if (!rt_ray_aabb_inv(O, inv, mk3(lo.x, lo.y, lo.z), mk3(hi.x, hi.y, hi.z))) continue;
        }
        best_t = t; best_key = key; best_slot = slot;
    }
    return best_slot >= 0;
}

// slot -> ObjectId {Type, Index, MeshIndex} (ObjectId.cs:7-9)
// virtual slot of an instanced hit -> (instance in TLAS order, BLAS slot)
RT_HD void rt_inst_decode(const RtScene& S, int32_t slot, int32_t& j, int32_t& bslot) {
    const int32_t v = slot - S.n_slots;
    int32_t lo = 0, hi = S.n_inst;                                  // largest j with inst_vbase[j] <= v (skips empty instances)
    while (hi - lo > 1) { const int32_t mid = (lo + hi) >> 1; if (S.inst_vbase[mid] <= v) lo = mid; else hi = mid; }
    j = lo;
    bslot = f2i(RT_LDG4(S.inst + RT_INST_F4 * (size_t)j + 6).y) + (v - S.inst_vbase[j]);
    RT_CHECK(RT_CHK_VSLOT, v >= 0 && j >= 0 && j < S.n_inst && v < S.inst_vbase[S.n_inst] && v < S.inst_vbase[j + 1] && bslot >= 0 && bslot < S.n_islots);
}
// world vertices of an instanced hit, as the reference's bake computes them (SceneMesh.cs:28)
RT_HD void rt_inst_world_tri(const RtScene& S, int32_t j, int32_t bslot, f3& w0, f3& w1, f3& w2) {
    const float4* ip = S.inst + RT_INST_F4 * (size_t)j;
    const float4 m0 = RT_LDG4(ip), m1 = RT_LDG4(ip + 1), m2 = RT_LDG4(ip + 2);
    const float4* pp = S.iprims + 3 * (size_t)bslot;
    const float4 p0 = RT_LDG4(pp), p1 = RT_LDG4(pp + 1), p2 = RT_LDG4(pp + 2);
    w0 = rt_xform_point(m0, m1, m2, mk3(p0.x, p0.y, p0.z)); w1 = rt_xform_point(m0, m1, m2, mk3(p1.x, p1.y, p1.z));
    w2 = rt_xform_point(m0, m1, m2, mk3(p2.x, p2.y, p2.z));
}

template <bool INST = false>
RT_HD void rt_slot_to_id(const RtScene& S, int32_t slot, int32_t& type, int32_t& index, int32_t& mesh) {
    if (slot < 0) { type = 0; index = -1; mesh = -1; return; }
    if (INST && slot >= S.n_slots) {                                // a triangle of an instanced mesh: MeshTriangle, Index within its mesh
        int32_t j, bslot; rt_inst_decode(S, slot, j, bslot);
        type = 3; index = f2i(RT_LDG4(S.iprims + 3 * (size_t)bslot).w); mesh = f2i(RT_LDG4(S.inst + RT_INST_F4 * (size_t)j + 6).z);
        return;
    }
    const float4* pp = S.prims + 3 * (size_t)slot;
    int32_t key = f2i(RT_LDG4(pp).w), kind = f2i(RT_LDG4(pp + 1).w), m = f2i(RT_LDG4(pp + 2).w);
    if (kind != 0) { type = 1; index = key - S.n_meshtris; mesh = -1; }
    else if (m >= 0) { type = 3; index = key - S.mesh_base[m]; mesh = m; }
    else { type = 2; index = key - S.n_meshtris - S.n_spheres; mesh = -1; }
}

// ---------------------------------------------------------------------------------------------------------------
// Shading: RayTracingSetup.Shade minus the recursion (RayTracingSetup.cs:304-366).  Returns the colour of this bounce
// (ambient + light loop, Rgb.Value scale 0..255) and, when the surface is a mirror and bounce < max, the reflection ray
// and MirrorReflectance; the caller adds  km * Shade(reflect)  afterwards, innermost first, like the recursion does.
// ---------------------------------------------------------------------------------------------------------------
struct RtMirror { bool emit; f3 O, D, km; };

// Surface state of one hit between the pieces below (the kernels drive the light loop warp-uniformly so that the shadow
// query of all 32 lanes can run as one packet; rt_shade_hit strings the same pieces together for a single lane).
struct RtSurf {
    f3 P, N, V, sO;           // hit point, normal as stored / sphere normal, view vector, P + N*ShadowRayEpsilon
    f3 kd, ks, km, color;
    float phong; bool is_mirror;
};

// RTS:321-323: surface point and normal (stored normal for triangles, RTS:420-431; GetSphereNormal, RTS:402-407)
template <bool INST = false>
RT_HD void rt_surf_geom(const RtScene& S, f3 O, f3 D, float t, int32_t slot, f3& P, f3& N, int32_t& mat_id) {
    if (INST && slot >= S.n_slots) {                                // instanced triangle: the normal the bake would have stored
        int32_t j, bslot; rt_inst_decode(S, slot, j, bslot);
        f3 w0, w1, w2; rt_inst_world_tri(S, j, bslot, w0, w1, w2);
        P = O + D * t;
        N = rt_tri_normal_neg(w0, w1, w2);
        mat_id = f2i(RT_LDG4(S.inst + RT_INST_F4 * (size_t)j + 7).y);
        return;
    }
    const float4* pp = S.prims + 3 * (size_t)slot;
    float4 p0 = RT_LDG4(pp), p1 = RT_LDG4(pp + 1);
    float4 pi = RT_LDG4(S.pinfo + slot);
    P = O + D * t;                                                           // Ray.GetPoint, Ray.cs:18-21
    if (f2i(p1.w) != 0) N = um_normalize(P - mk3(p0.x, p0.y, p0.z));
    else N = mk3(pi.x, pi.y, pi.z);
    mat_id = f2i(pi.w);
}

// RTS:321-325: + material lookup (RTS:409-436), ambient term (RTS:438-441), view vector
template <bool INST = false>
RT_HD void rt_surf_begin(const RtScene& S, f3 O, f3 D, float t, int32_t slot, RtSurf& sf) {
    int32_t mat_id;
    rt_surf_geom<INST>(S, O, D, t, slot, sf.P, sf.N, mat_id);
    const float4* mp = S.mats + 4 * (size_t)mat_id;
    float4 m0 = RT_LDG4(mp), m1 = RT_LDG4(mp + 1), m2 = RT_LDG4(mp + 2), m3 = RT_LDG4(mp + 3);
    sf.kd = mk3(m0.x, m0.y, m0.z); sf.km = mk3(m2.x, m2.y, m2.z); sf.ks = mk3(m3.x, m3.y, m3.z);
    sf.phong = m0.w; sf.is_mirror = f2i(m1.w) != 0;
    sf.color = S.ambient * mk3(m1.x, m1.y, m1.z);                            // CalculateAmbient, RTS:438-441
    sf.V = um_normalize(O - sf.P);                                           // RTS:325
    sf.sO = sf.P + sf.N * RT_SHADOW_EPS;                                     // RTS:331 (also Reflect's origin, RTS:370)
}

// RTS:329-334: shadow ray direction and lightDistanceSq for light li, from surface point P
RT_HD void rt_light_ray(const RtScene& S, f3 P, int32_t li, f3& Ldir, float& d2) {
    RT_CHECK(RT_CHK_LIGHT, li >= 0 && li < S.n_lights);
    float4 lp = RT_LDG4(S.lights + 2 * li);
    f3 Lpos = mk3(lp.x, lp.y, lp.z);
    Ldir = um_normalize(Lpos - P);                                           // RTS:330
    f3 dl = Lpos - P;
    d2 = um_dot(dl, dl);                                                     // distancesq(P, Lpos), RTS:334
}

// RTS:350-355 for an unshadowed light: color += diffuse + specular
RT_HD void rt_light_add(const RtScene& S, RtSurf& sf, int32_t li, f3 Ldir, float d2) {
    float4 lint = RT_LDG4(S.lights + 2 * li + 1);
    f3 E = mk3(lint.x, lint.y, lint.z) / d2;                                 // RTS:350
    float ldn = um_dot(Ldir, sf.N);
    f3 diff = sf.kd * um_max(0.0f, ldn) * E;                                 // CalculateDiffuse, RTS:443-455
    f3 spec;
    if (ldn <= RT_SPEC_CUTOFF) spec = bc3(0.0f);                             // angle > 90, RTS:384-391
    else {
        f3 v = Ldir + sf.V;
        f3 H = v / um_length(v);
        float cnh = um_max(0.0f, um_dot(sf.N, H));
        spec = sf.ks * um_pow(cnh, sf.phong) * E;                            // RTS:398-399
    }
    sf.color = sf.color + (diff + spec);                                     // RTS:355
}

// RTS:358-362 + Reflect (RTS:368-373, direction not renormalised)
RT_HD void rt_surf_mirror(const RtSurf& sf, int32_t bounce, int32_t max_bounces, RtMirror& mir) {
    mir.emit = sf.is_mirror && bounce < max_bounces;
    if (mir.emit) {
        mir.O = sf.sO;
        mir.D = 2.0f * sf.N * um_dot(sf.V, sf.N) - sf.V;
        mir.km = sf.km;
    }
}

// A shadow query whose answer cannot change the pixel.  For a light behind the surface (dot(L, N) <= RT_SPEC_CUTOFF < 0) the
// reference adds diffuse = kd * max(0, ldn) * E = kd * 0 * E and specular = 0 when the light is visible (RTS:350-355,384-391)
// and nothing when it is occluded: the same colour, bit for bit, provided kd * 0 * E is a zero and the colour it is added to
// is not -0.0.  rt_device_commit sets S.backcull only when that holds for every material and light of the scene (finite kd,
// finite intensities below 1e18, no -0.0 channel in ambient * ka); d2 > 1e-20 then keeps E = I / d2 finite.  The query is
// still counted (the reference issues it) and reported as occluded, so the shading loop skips the light.
RT_HD bool rt_shadow_irrelevant(const RtScene& S, f3 Ldir, f3 N, float d2) {
    return S.backcull != 0 && um_dot(Ldir, N) <= RT_SPEC_CUTOFF && d2 > 1e-20f;
}

// Shadow queries of one hit for lights [light_begin, light_begin + n) (n <= 32), one lane: bit i set = light
// light_begin+i is occluded (RTS:331-345).  The kernels run the same loop warp-uniformly with packet traversal.
template <bool INST = false, class STK>
RT_HD uint32_t rt_shadow_mask(const RtScene& S, f3 O, f3 D, float t, int32_t slot, int32_t light_begin, int32_t n, STK& stk,
                              uint32_t& n_shadow) {
    f3 P, N; int32_t mat_id;
    rt_surf_geom<INST>(S, O, D, t, slot, P, N, mat_id);
    f3 sO = P + N * RT_SHADOW_EPS;                                           // RTS:331
    uint32_t mask = 0;
    for (int32_t i = 0; i < n; ++i) {
        f3 Ldir; float d2;
        rt_light_ray(S, P, light_begin + i, Ldir, d2);
        ++n_shadow;
        if (rt_shadow_irrelevant(S, Ldir, N, d2)) { mask |= 1u << i; continue; }
        f3 inv = rt_rcp3(Ldir);
        if (rt_scene_gate(S, sO, inv)) {                                     // Scene.cs:54 inside the shadow query
            float st; int32_t ss;
            if (rt_traverse<true, false, INST>(S, sO, Ldir, inv, d2, stk, st, ss)) mask |= 1u << i;
        }
    }
    return mask;
}

// Shade (RTS:304-366 minus the recursion) with the shadow results already known: occl[(li >> 5) * stride + q] bit (li & 31).
template <bool INST = false>
RT_HD f3 rt_shade_masked(const RtScene& S, f3 O, f3 D, float t, int32_t slot, int32_t bounce, int32_t max_bounces,
                         const uint32_t* occl, size_t stride, size_t q, RtMirror& mir) {
    RtSurf sf;
    rt_surf_begin<INST>(S, O, D, t, slot, sf);
    uint32_t mask = 0;
    for (int32_t li = 0; li < S.n_lights; ++li) {
        if ((li & 31) == 0) mask = occl[(size_t)(li >> 5) * stride + q];
        if (mask & (1u << (li & 31))) continue;                              // occluder before the light, RTS:337-345
        f3 Ldir; float d2;
        rt_light_ray(S, sf.P, li, Ldir, d2);
        rt_light_add(S, sf, li, Ldir, d2);
    }
    rt_surf_mirror(sf, bounce, max_bounces, mir);
    return sf.color;
}

// ---------------------------------------------------------------------------------------------------------------
// Work-item mapping and ray generation
// ---------------------------------------------------------------------------------------------------------------
// rows owned by `rank`: bands b = rank, rank+world, ...; band b covers rows [b*band_rows, min(res_y,(b+1)*band_rows))
RT_HD int32_t rt_local_rows(int32_t res_y, int32_t rank, int32_t world, int32_t band_rows) {
    int32_t n_bands = (res_y + band_rows - 1) / band_rows, rows = 0;
    for (int32_t b = rank; b < n_bands; b += world) {
        int32_t r = res_y - b * band_rows;
        rows += r < band_rows ? r : band_rows;
    }
    return rows;
}
// fills the derived fields of a view (tiles, local rows, division constants) from res/rank/world/band_rows
RT_HD void rt_view_finish(RtView& V) {
    V.tiles_x = (V.res_x + 7) / 8;
    V.n_local_rows = rt_local_rows(V.res_y, V.rank, V.world, V.band_rows);
    V.n_tiles = V.tiles_x * ((V.n_local_rows + 3) / 4);
    V.tiles_x_magic = V.tiles_x <= 1 ? 0xffffffffu : (uint32_t)(0x100000000ull / (uint64_t)V.tiles_x);
    V.band_shift = -1;
    for (int32_t k = 0; k < 31; ++k) if (V.band_rows == (1 << k)) V.band_shift = k;
}
// sets V.far_origin for a committed scene (see rt_far_pad)
RT_HD void rt_view_far_origin(RtView& V, float origin_reach) {
    const float o = fmaxf(fmaxf(fabsf(V.pos.x), fabsf(V.pos.y)), fabsf(V.pos.z));
    V.far_origin = (o <= origin_reach) ? 0 : 1;                              // NaN positions count as far
}
RT_HD float rt_origin_reach(float pad) { return RT_FAR_ORIGIN_FACTOR * 65536.0f * pad; }
// local work item (tile, lane) -> pixel. One tile is an 8x4 pixel block (one warp) of the rank's local image.
// tile / tiles_x by multiplication: q' = mulhi(tile, floor(2^32/d)) is q or q-1 for every 32-bit tile; one fix-up step.
RT_HD bool rt_item_to_pixel(const RtView& V, int32_t tile, int32_t lane, int32_t& x, int32_t& y) {
    uint32_t ty = (uint32_t)(((uint64_t)(uint32_t)tile * (uint64_t)V.tiles_x_magic) >> 32);
    uint32_t tx = (uint32_t)tile - ty * (uint32_t)V.tiles_x;
    if (tx >= (uint32_t)V.tiles_x) { tx -= (uint32_t)V.tiles_x; ++ty; }
    x = (int32_t)tx * 8 + (lane & 7);
    int32_t r = (int32_t)ty * 4 + (lane >> 3);              // local row
    if (V.world == 1) y = r;
    else if (V.band_shift >= 0) y = ((((r >> V.band_shift) * V.world + V.rank)) << V.band_shift) + (r & (V.band_rows - 1));
    else y = ((r / V.band_rows) * V.world + V.rank) * V.band_rows + (r % V.band_rows);
    return x < V.res_x && r < V.n_local_rows && y < V.res_y;
}

// CastPixelRays body, RayTracingSetup.cs:275-298 with ImagePlane.GetRect (ImagePlane.cs:26-44); sample (si,sj) of the
// spp_grid x spp_grid extension replaces the +0.5f pixel-centre offset (spp_grid 1 -> exactly +0.5f).
RT_HD void rt_primary_ray(const RtView& V, int32_t x, int32_t y, int32_t si, int32_t sj, f3& O, f3& D) {
    f3 center = V.pos + V.fwd * V.dist;
    f3 halfUp = V.up * V.half_v, halfRight = V.right * V.half_h;
    f3 topLeft = center - halfRight + halfUp;
    float H = V.half_h * 2.0f, Vl = V.half_v * 2.0f;
    float g = (float)V.spp_grid;
    float ox = rt_div((float)si + 0.5f, g), oy = rt_div((float)sj + 0.5f, g);
    float rightMove = rt_div(((float)x + ox) * H, (float)V.res_x);
    float downMove = rt_div(((float)y + oy) * Vl, (float)V.res_y);
    f3 pp = topLeft + rightMove * V.right - V.up * downMove;
    O = V.pos;
    D = um_normalize(pp - V.pos);
}

// our u8 quantisation (the reference has none, Rgb.cs:12-13): clamp(round_half_even(Value), 0, 255), NaN -> 0
RT_HD uint8_t rt_quant(float v) {
    if (v != v) return 0;
#if defined(__CUDA_ARCH__)
    float r = rintf(v);
#else
    float r = nearbyintf(v);
#endif
    r = r < 0.0f ? 0.0f : (r > 255.0f ? 255.0f : r);
    return (uint8_t)r;
}

// ---------------------------------------------------------------------------------------------------------------
// Per-item bodies of the non-traversal kernels (shared with tests/emu)
// ---------------------------------------------------------------------------------------------------------------
// color += new Rgb(mirrorReflectance * Shade(reflectRay).Value), RayTracingSetup.cs:362. link = (km.xyz, parent index)
RT_HD void rt_fold_item(uint32_t q, const float4* link, const float4* local_cur, float4* local_prev) {
    float4 l = link[q], c = local_cur[q];
    int32_t parent = f2i(l.w);
    float4 p = local_prev[parent];
    p.x = p.x + l.x * c.x; p.y = p.y + l.y * c.y; p.z = p.z + l.z * c.z;
    local_prev[parent] = p;
}

struct RtOut { float4* rgba32f; uint32_t* rgba8; int32_t* hit_id; float* hit_t; };

// One work item of a finished wave -> framebuffer(s).  pass/n_pass: sample index of the spp_grid^2 extension; colours are
// summed in pass order and divided once, before Rgb.Color (Rgb.cs:13) and the u8 quantisation.
template <bool INST = false>
RT_HD void rt_finalize_value(const RtScene& S, const RtView& V, int32_t tile_begin, uint32_t q, int32_t pass, int32_t n_pass,
                             float4 c, float ht, int32_t hslot, float4* accum, const RtOut& out) {
    int32_t x, y;
    if (!rt_item_to_pixel(V, tile_begin + (int32_t)(q >> 5), (int32_t)(q & 31), x, y)) return;
    size_t pix = (size_t)x + (size_t)y * (size_t)V.res_x;                  // GetPixelIndex, RayTracingSetup.cs:243-246
    RT_CHECK(RT_CHK_PIXEL, x >= 0 && y >= 0 && x < V.res_x && y < V.res_y);
    if (pass == 0) {
        if (out.hit_id) {
            int32_t ty, ix, me; rt_slot_to_id<INST>(S, hslot, ty, ix, me);
            out.hit_id[3 * pix] = ty; out.hit_id[3 * pix + 1] = ix; out.hit_id[3 * pix + 2] = me;
        }
        if (out.hit_t) RT_ST_STREAM(out.hit_t + pix, ht);
    }
    if (n_pass > 1) {
        size_t a = ((size_t)tile_begin << 5) + q;
        if (pass > 0) { float4 p = accum[a]; c.x = p.x + c.x; c.y = p.y + c.y; c.z = p.z + c.z; }
        if (pass < n_pass - 1) { accum[a] = c; return; }
        float np = (float)n_pass;
        c.x = rt_div(c.x, np); c.y = rt_div(c.y, np); c.z = rt_div(c.z, np);
    }
    if (out.rgba32f) RT_ST_STREAM(out.rgba32f + pix, make_float4(rt_div(c.x, 255.f), rt_div(c.y, 255.f), rt_div(c.z, 255.f), 1.f));
    if (out.rgba8) RT_ST_STREAM(out.rgba8 + pix, (uint32_t)rt_quant(c.x) | ((uint32_t)rt_quant(c.y) << 8) | ((uint32_t)rt_quant(c.z) << 16) | 0xff000000u);
}

template <bool INST = false>
RT_HD void rt_finalize_item(const RtScene& S, const RtView& V, int32_t tile_begin, uint32_t q, int32_t pass, int32_t n_pass,
                            const float4* local0, const float* hit_t, const int32_t* hit_slot, float4* accum, const RtOut& out) {
    rt_finalize_value<INST>(S, V, tile_begin, q, pass, n_pass, local0[q], hit_t[q], hit_slot[q], accum, out);
}

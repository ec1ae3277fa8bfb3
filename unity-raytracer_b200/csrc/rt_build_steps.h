// rt_build_steps.h — per-element bodies of the device LBVH build (Morton codes -> radix sort -> Karras 2012 hierarchy ->
// bottom-up refit -> leaf collapse -> compact node emission), written host/device so tests/emu can run the identical
// logic serially on the CPU.  rt_build.cu wraps each body in a grid-stride kernel.
//
// The reference's own BVH (Data/Collision/BVH.cs) is unfinished (Subdivide throws, BVH.cs:77-79) and unreferenced, so
// there is no reference build to match; what must match is the *result of the closest-hit query* (Scene.cs:43-122).
// That is guaranteed by traversal (rt_core.h) as long as node boxes are conservative, which the padding here ensures.
#pragma once
#include "rt_hd.h"

struct RtRaw {                    // raw uploaded arrays, in the reference's scan order
    const float* tri_v;           // 9 floats per triangle: all mesh triangles (mesh order), then the loose triangles
    const float* tri_n;           // 3 floats per triangle, same order (Mesh.TriangleNormals / TriangleData.Normals)
    const float4* spheres;        // (Center, RadiusSquared)
    const int32_t* mesh_base;     // [n_meshes+1]
    int32_t n_meshes, n_meshtris, n_spheres, n_loose;
    // Sphere boxes (rt_sphere_box_radius): every ray origin the committed tree must serve lies in [obox_lo, obox_hi]
    // (the primitive bounds, where shadow and mirror rays start, united with the camera positions seen so far);
    // sphere_k = 0 switches the inflation off (the pass that computes the primitive bounds themselves).
    f3 obox_lo, obox_hi;
    float sphere_k;
    // Scenes with instanced meshes (rt_instances.h): the instanced triangles have scan keys (they are meshes of the reference's
    // scan, Scene.cs:64-93) but no entry in the raw arrays, so for everything that IS in the raw arrays the enumeration index
    // used by the builders ("build index": mesh triangles in tri_v, spheres, loose triangles) and the scan key differ:
    //   mesh_base / n_meshtris above index the RAW arrays; mesh_key[m] = scan key of triangle 0 of mesh m; key_tris = scan keys
    //   taken by all mesh triangles, instanced ones included.  Without instances mesh_key == mesh_base and key_tris == n_meshtris.
    const int32_t* mesh_key;
    int32_t key_tris;
};

// Box half-width of a sphere.  RMath.RaySphereIntersection (RMath.cs:81-108) evaluates
//     disc = uoc*uoc - (lengthsq(oc) - r2),  oc = O - C,  uoc = dot(D, oc)
// in binary32, where uoc^2 and lengthsq(oc) are both ~|oc|^2 and cancel down to ~r2: the computed disc differs from the exact
// r2 - dist(ray, C)^2 by up to  (12 u from the products/sums + 2 u from fl(O - C) + 9 u from |D|^2 = 1 +- 9 u of a normalised
// direction) * |oc|^2,  u = 2^-24  (measured on random grazing rays: up to 9.5 u).  `disc < 0 -> miss` therefore ACCEPTS rays
// that pass the centre at up to sqrt(r2 + 23 u |oc|^2): seen from 100 units away a sphere of radius 0.1 is hit out to 1.21 r.
// That is the reference's answer, so the BVH may not cull those rays: the box uses  r_box = sqrt(r2 + RT_SPHERE_K * Rmax^2),
// Rmax = distance from C to the farthest corner of the origin box, RT_SPHERE_K = 32 u = 2^-19 (mirror directions
// 2N(V.N) - V are not renormalised, RayTracingSetup.cs:371, and carry a few u more).  Directions are assumed normalised in
// binary32; rt_intersect handles others by a scan.
#define RT_SPHERE_K 0x1p-19f
RT_HD float rt_sphere_box_radius(const RtRaw& R, float4 s) {
    float r = rt_sqrt(s.w);
    if (R.sphere_k > 0.0f) {
        const double dx = fmax(fabs((double)s.x - R.obox_lo.x), fabs((double)s.x - R.obox_hi.x));
        const double dy = fmax(fabs((double)s.y - R.obox_lo.y), fabs((double)s.y - R.obox_hi.y));
        const double dz = fmax(fabs((double)s.z - R.obox_lo.z), fabs((double)s.z - R.obox_hi.z));
        const double rr = (double)s.w + (double)R.sphere_k * (dx * dx + dy * dy + dz * dz);
        if (rr >= 0.0 && rr < 1e37) r = (float)(sqrt(rr) * 1.000001);       // else: NaN / negative / huge input, box as before
    }
    return r;
}

RT_HD int32_t rt_raw_nprims(const RtRaw& R) { return R.n_meshtris + R.n_spheres + R.n_loose; }

// build index -> (kind, raw index, mesh index, material id, scan key)
RT_HD void rt_key_decode(const RtRaw& R, int32_t key, int32_t& kind, int32_t& raw, int32_t& mesh, int32_t& mat, int32_t& scan_key) {
    if (key < R.n_meshtris) {
        int32_t lo = 0, hi = R.n_meshes;                // largest m with mesh_base[m] <= key (instanced meshes own no raw range)
        while (hi - lo > 1) { int32_t mid = (lo + hi) >> 1; if (R.mesh_base[mid] <= key) lo = mid; else hi = mid; }
        kind = 0; raw = key; mesh = lo; mat = lo;
        scan_key = R.mesh_key[lo] + (key - R.mesh_base[lo]);
    } else if (key < R.n_meshtris + R.n_spheres) {
        kind = 1; raw = key - R.n_meshtris; mesh = -1; mat = R.n_meshes + raw;
        scan_key = R.key_tris + raw;
    } else {
        kind = 0; raw = key - R.n_spheres; mesh = -1; mat = R.n_meshes + R.n_spheres + (key - R.n_meshtris - R.n_spheres);
        scan_key = R.key_tris + (key - R.n_meshtris);
    }
}
RT_HD void rt_key_decode(const RtRaw& R, int32_t key, int32_t& kind, int32_t& raw, int32_t& mesh, int32_t& mat) {
    int32_t sk; rt_key_decode(R, key, kind, raw, mesh, mat, sk);
}

RT_HD void rt_prim_box(const RtRaw& R, int32_t key, f3& lo, f3& hi) {
    int32_t kind, raw, mesh, mat; rt_key_decode(R, key, kind, raw, mesh, mat);
    if (kind == 1) {
        float4 s = R.spheres[raw];
        float r = rt_sphere_box_radius(R, s);
        lo = mk3(s.x - r, s.y - r, s.z - r); hi = mk3(s.x + r, s.y + r, s.z + r);
    } else {
        const float* v = R.tri_v + 9 * (size_t)raw;
        lo = mk3(fminf(v[0], fminf(v[3], v[6])), fminf(v[1], fminf(v[4], v[7])), fminf(v[2], fminf(v[5], v[8])));
        hi = mk3(fmaxf(v[0], fmaxf(v[3], v[6])), fmaxf(v[1], fmaxf(v[4], v[7])), fmaxf(v[2], fmaxf(v[5], v[8])));
    }
}

RT_HD uint64_t rt_expand21(uint64_t v) {
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}
// 63-bit Morton code of the box centre, normalised to the union of all primitive boxes [bmin, bmax]
RT_HD uint64_t rt_morton(f3 lo, f3 hi, f3 bmin, f3 inv_ext) {
    float cx = ((lo.x + hi.x) * 0.5f - bmin.x) * inv_ext.x;
    float cy = ((lo.y + hi.y) * 0.5f - bmin.y) * inv_ext.y;
    float cz = ((lo.z + hi.z) * 0.5f - bmin.z) * inv_ext.z;
    const float S = 2097152.0f;
    float qx = fminf(fmaxf(cx * S, 0.0f), S - 1.0f), qy = fminf(fmaxf(cy * S, 0.0f), S - 1.0f), qz = fminf(fmaxf(cz * S, 0.0f), S - 1.0f);
    if (!(qx == qx)) qx = 0.0f; if (!(qy == qy)) qy = 0.0f; if (!(qz == qz)) qz = 0.0f;
    return (rt_expand21((uint64_t)qx) << 2) | (rt_expand21((uint64_t)qy) << 1) | rt_expand21((uint64_t)qz);
}

RT_HD int32_t rt_clz64(uint64_t v) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)v);
#else
    return v ? __builtin_clzll(v) : 64;
#endif
}
RT_HD int32_t rt_clz32(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __clz((int)v);
#else
    return v ? __builtin_clz(v) : 32;
#endif
}
// Karras' delta: common-prefix length of sorted keys i and j, ties broken by index; -1 outside [0,n)
RT_HD int32_t rt_delta(const uint64_t* keys, int32_t n, int32_t i, int32_t j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + rt_clz32((uint32_t)i ^ (uint32_t)j);
    return rt_clz64(a ^ b);
}

struct RtTopo {                // Karras hierarchy over n sorted leaves: n-1 internal nodes, node 0 is the root
    int32_t* left;             // child codes: >= 0 internal node index, < 0 leaf ~slot
    int32_t* right;
    int32_t* first;            // covered slot range [first, last]
    int32_t* last;
    int32_t* parent_int;       // parent of internal node i (-1 for the root)
    int32_t* parent_leaf;      // parent of leaf slot s
};

RT_HD void rt_karras_node(const uint64_t* keys, int32_t n, int32_t i, const RtTopo& T) {
    int32_t d = (rt_delta(keys, n, i, i + 1) - rt_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int32_t dmin = rt_delta(keys, n, i, i - d);
    int32_t lmax = 2;
    while (rt_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int32_t l = 0;
    for (int32_t t = lmax >> 1; t >= 1; t >>= 1)
        if (rt_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int32_t j = i + l * d;
    int32_t dnode = rt_delta(keys, n, i, j);
    int32_t s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (rt_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int32_t gamma = i + s * d + (d < 0 ? d : 0);
    int32_t lo = i < j ? i : j, hi = i < j ? j : i;
    T.first[i] = lo; T.last[i] = hi;
    if (lo == gamma) { T.left[i] = ~gamma; T.parent_leaf[gamma] = i; }
    else { T.left[i] = gamma; T.parent_int[gamma] = i; }
    if (hi == gamma + 1) { T.right[i] = ~(gamma + 1); T.parent_leaf[gamma + 1] = i; }
    else { T.right[i] = gamma + 1; T.parent_int[gamma + 1] = i; }
    if (i == 0) T.parent_int[0] = -1;
}

// per-slot primitive records (layout: RtScene in rt_core.h)
RT_HD void rt_emit_slot(const RtRaw& R, int32_t slot, int32_t bkey, float4* prims, float4* pinfo, float4* leaf_lo, float4* leaf_hi) {
    int32_t kind, raw, mesh, mat, key; rt_key_decode(R, bkey, kind, raw, mesh, mat, key);
    f3 lo, hi; rt_prim_box(R, bkey, lo, hi);
    leaf_lo[slot] = make_float4(lo.x, lo.y, lo.z, 0.f);
    leaf_hi[slot] = make_float4(hi.x, hi.y, hi.z, 0.f);
    float4* pp = prims + 3 * (size_t)slot;
    if (kind == 1) {
        float4 s = R.spheres[raw];
        pp[0] = make_float4(s.x, s.y, s.z, i2f(key));
        pp[1] = make_float4(s.w, 0.f, 0.f, i2f(1));
        pp[2] = make_float4(0.f, 0.f, 0.f, i2f(-1));
        pinfo[slot] = make_float4(0.f, 0.f, 0.f, i2f(mat));
    } else {
        const float* v = R.tri_v + 9 * (size_t)raw;
        const float* nn = R.tri_n + 3 * (size_t)raw;
        pp[0] = make_float4(v[0], v[1], v[2], i2f(key));
        pp[1] = make_float4(v[3] - v[0], v[4] - v[1], v[5] - v[2], i2f(0));       // edge1 = v1 - v0, RMath.cs:34
        pp[2] = make_float4(v[6] - v[0], v[7] - v[1], v[8] - v[2], i2f(mesh));    // edge2 = v2 - v0, RMath.cs:35
        pinfo[slot] = make_float4(nn[0], nn[1], nn[2], i2f(mat));
    }
}

// child code -> (traversal ref, box). new_index maps live internal nodes to their compacted position.
RT_HD void rt_child_ref(int32_t code, const RtTopo& T, const int32_t* new_index, int32_t leaf_max,
                        const float4* leaf_lo, const float4* leaf_hi, const float4* int_lo, const float4* int_hi,
                        int32_t& ref, float4& lo, float4& hi) {
    if (code < 0) {
        int32_t s = ~code;
        ref = ~(s * 8); lo = leaf_lo[s]; hi = leaf_hi[s];
    } else {
        int32_t cnt = T.last[code] - T.first[code] + 1;
        lo = int_lo[code]; hi = int_hi[code];
        ref = (cnt <= leaf_max) ? ~(T.first[code] * 8 + (cnt - 1)) : new_index[code];
    }
}

RT_HD void rt_emit_node(int32_t i, const RtTopo& T, const int32_t* new_index, int32_t leaf_max, float pad,
                        const float4* leaf_lo, const float4* leaf_hi, const float4* int_lo, const float4* int_hi, float4* nodes) {
    if (T.last[i] - T.first[i] + 1 <= leaf_max) return;          // collapsed into a leaf of its parent (or of the root)
    int32_t ra, rb; float4 alo, ahi, blo, bhi;
    rt_child_ref(T.left[i], T, new_index, leaf_max, leaf_lo, leaf_hi, int_lo, int_hi, ra, alo, ahi);
    rt_child_ref(T.right[i], T, new_index, leaf_max, leaf_lo, leaf_hi, int_lo, int_hi, rb, blo, bhi);
    float4* np = nodes + 4 * (size_t)new_index[i];
    np[0] = make_float4(alo.x - pad, ahi.x + pad, alo.y - pad, ahi.y + pad);
    np[1] = make_float4(blo.x - pad, bhi.x + pad, blo.y - pad, bhi.y + pad);
    np[2] = make_float4(alo.z - pad, ahi.z + pad, blo.z - pad, bhi.z + pad);
    np[3] = make_float4(i2f(ra), i2f(rb), 0.f, 0.f);
}

// ---- refit of an emitted node array (rt_scene_refit): same tree, new boxes ----------------------------------------------
// parent[c] = node that references inner node c (the root keeps -1)
RT_HD void rt_node_parents(const float4* nodes, int32_t i, int32_t* parent) {
    float4 n3 = nodes[4 * (size_t)i + 3];
    int32_t ra = f2i(n3.x), rb = f2i(n3.y);
    if (ra >= 0) parent[ra] = i;
    if (rb >= 0) parent[rb] = i;
}
// box of a child reference: union of the slot boxes of a leaf, or the (already refitted) box of an inner node
RT_HD void rt_ref_box(int32_t ref, const float4* leaf_lo, const float4* leaf_hi, const float4* int_lo, const float4* int_hi, f3& lo, f3& hi) {
    if (ref >= 0) { float4 a = int_lo[ref], b = int_hi[ref]; lo = mk3(a.x, a.y, a.z); hi = mk3(b.x, b.y, b.z); return; }
    int32_t r = ~ref, first = r >> 3, cnt = (r & 7) + 1;
    lo = mk3(FLT_MAX, FLT_MAX, FLT_MAX); hi = mk3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    for (int32_t k = 0; k < cnt; ++k) {
        float4 a = leaf_lo[first + k], b = leaf_hi[first + k];
        lo = mk3(fminf(lo.x, a.x), fminf(lo.y, a.y), fminf(lo.z, a.z)); hi = mk3(fmaxf(hi.x, b.x), fmaxf(hi.y, b.y), fmaxf(hi.z, b.z));
    }
}

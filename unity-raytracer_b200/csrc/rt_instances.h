// rt_instances.h — host side of instanced meshes (rt_scene_base_mesh / rt_scene_mesh_instance, include/rtb200.h): the
// object-space base meshes, the instances (3x4 matrix + MeshIndex), one BLAS per base mesh and the TLAS over the instances,
// packed as the arrays RtScene::inodes / iprims / inst / inst_vbase describe (rt_core.h "Instancing").
//
// What it replaces: SceneMesh.Mesh (SceneComponents/SceneMesh.cs:11-53) bakes transform.localToWorldMatrix into a fresh world-
// space copy of every mesh on every frame, and Scene.IntersectRay (Scene.cs:64-93) scans those copies.  Here a mesh that occurs
// k times is stored once; re-posing an instance costs its 48-byte matrix and a TLAS rebuild over the instance boxes (host, a
// few microseconds per hundred instances) — not a re-upload and re-build of its triangles.  Results are bit-identical to the
// baked scene: only the CULLING happens in object space (rt_core.h), so the tree layout and the padding below can never change
// a hit, a distance or a colour, only the time it takes.
//
// Host only; shared by rt_build.cu and the CPU emulation of the device headers (tests/emu).
#pragma once
#include <atomic>
#include <vector>
#include <cmath>
#include <cstring>
#include <cstdint>
#include "rt_sah.h"

struct RtBaseMesh {
    std::vector<float> v;                 // 9 floats per triangle, OBJECT space (what MeshFilter.sharedMesh holds, SceneMesh.cs:15-17)
    bool built = false;                   // BLAS below is current
    std::vector<float> nodes;             // 16 floats per node, UNPADDED boxes, refs relative to this BLAS (node 0 = root)
    std::vector<int32_t> order;           // BLAS slot -> triangle index
    int32_t root = -1;
    int depth = 0;
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0}, amax = 0.f;     // object-space bounds, max |coordinate|
    // Object-space boxes that together cover the mesh: the BLAS cut off after ~RT_HULL_BOXES nodes.  An instance's world bounds
    // are the union of the TRANSFORMED boxes: for a rotated instance the transformed corners of the one object box reach up to
    // sqrt(3) times too far per axis (cfg4: boxes of neighbouring instances overlapped and every ray entered 3 - 5 instances it
    // could not hit); the pieces of a cut are small, so their rotated boxes hug the surface.
    std::vector<float> hull;              // 6 floats per box (lo, hi)
    uint64_t hash = 0;                    // of the vertex bytes: a base mesh re-declared unchanged after rt_scene_begin keeps its BLAS
    uint64_t uid = 0;                     // identity of the built BLAS (set by build_blas, travels with the cache): what the packed sections were emitted from
    int32_t n_tris() const { return (int32_t)(v.size() / 9); }
    static uint64_t hash_bytes(const void* p, size_t n) {             // four independent lanes: the multiplies of one lane are a dependent chain
        const unsigned char* b = (const unsigned char*)p;
        uint64_t h[4] = {0x9E3779B97F4A7C15ull ^ (uint64_t)n, 0xC2B2AE3D27D4EB4Full, 0x165667B19E3779F9ull, 0x27D4EB2F165667C5ull};
        size_t i = 0;
        for (; i + 32 <= n; i += 32) {
            uint64_t w[4]; std::memcpy(w, b + i, 32);
            for (int k = 0; k < 4; ++k) { h[k] = (h[k] ^ w[k]) * 0xff51afd7ed558ccdull; h[k] ^= h[k] >> 32; }
        }
        uint64_t r = h[0];
        for (int k = 1; k < 4; ++k) { r = (r ^ h[k]) * 0xc4ceb9fe1a85ec53ull; r ^= r >> 29; }
        for (; i < n; ++i) { r = (r ^ b[i]) * 0x100000001b3ull; }
        return r;
    }
};

struct RtInstance {
    int32_t base = -1, mesh = -1;         // base mesh, MeshIndex (position among all rt_scene_mesh / rt_scene_mesh_instance calls)
    float m[12] = {0};                    // rows of the object->world matrix: m00 m01 m02 m03 | m10 .. | m20 ..   (Matrix4x4, 3x4 part)
    float inv[12] = {0};                  // rows of the world->object map: O' = A O + a
    float inv_norm = 0.f;                 // |A|_inf (max absolute row sum): bounds |A x|_inf / |x|_inf
    float wlo[3] = {0, 0, 0}, whi[3] = {0, 0, 0};               // conservative world bounds (transformed corners of the object bounds)
};

struct RtInstanceSet {
    std::vector<RtBaseMesh> bases;
    std::vector<RtBaseMesh> prev_bases;   // the base meshes of the scene before the last clear(): a shim that re-sends its whole scene every
                                          // frame (RayTracingSetup.cs:120-128) re-declares the same meshes and gets their BLAS back (take_cached)
    std::vector<RtInstance> insts;        // call order
    // ---- packed for the device ----
    std::vector<float> inodes;            // [BLAS nodes of base 0 .. B-1][TLAS nodes]
    std::vector<float> iprims;            // 12 floats per BLAS slot
    std::vector<float> inst;              // 32 floats per packed instance (TLAS leaf order)
    std::vector<int32_t> vbase;           // [n_packed + 1]
    std::vector<int32_t> node_base, slot_base;                  // per base mesh
    std::vector<float> blas_pad;          // padding baked into each BLAS section of inodes
    std::vector<uint64_t> packed_uid;     // RtBaseMesh::uid of the BLAS each section was emitted from
    int32_t tlas_base = 0, tlas_root = -1, n_packed = 0;
    int depth = 0;                        // stack entries the two-level walk may need below the world BVH
    bool blas_changed = true;             // iprims / the BLAS part of inodes must be (re)uploaded
    uint64_t generation = 0;              // bumped by every pack()
    uint64_t blas_generation = 0;         // bumped whenever iprims / the BLAS sections were re-emitted: devices compare it with what they hold

    // the built BLAS of an identical base mesh declared in the previous scene, if any (moved out of the cache)
    // (compares the CALLER's array with the cached copies: a mesh re-declared unchanged is neither copied nor scanned again)
    bool take_cached(const float* v, size_t n_floats, uint64_t hash, RtBaseMesh& B) {
        for (size_t i = 0; i < prev_bases.size(); ++i) {
            RtBaseMesh& P = prev_bases[i];
            if (P.built && P.hash == hash && P.v.size() == n_floats && (n_floats == 0 || std::memcmp(P.v.data(), v, n_floats * 4) == 0)) {
                B = std::move(P); P.built = false; P.v.clear();
                return true;
            }
        }
        return false;
    }
    // The packed BLAS sections (iprims, the head of inodes, node_base / slot_base / blas_pad) survive clear(): when the next scene
    // declares the SAME base meshes in the same order (keep_blas_if_same) they are neither emitted nor uploaded again.
    void clear() { prev_bases.clear(); prev_bases.swap(bases); insts.clear(); inst.clear(); vbase.clear();
                   tlas_base = 0; tlas_root = -1; n_packed = 0; depth = 0; blas_changed = true; }
    // at rt_scene_commit, once every base mesh is declared: the packed BLAS sections are still those of exactly this list of BLASes
    void keep_blas_if_same() {
        if (bases.empty() || packed_uid.size() != bases.size() || slot_base.size() != bases.size() + 1 || blas_pad.size() != bases.size()) return;
        for (size_t i = 0; i < bases.size(); ++i) if (!bases[i].built || bases[i].uid == 0 || bases[i].uid != packed_uid[i]) return;
        blas_changed = false;
    }
    bool any() const { return !insts.empty(); }
    int64_t n_tris() const { int64_t n = 0; for (auto& i : insts) n += bases[(size_t)i.base].n_tris(); return n; }

    // object->world matrix of an instance and everything derived from it; false: not invertible / not finite / too ill-conditioned
    // for object-space culling (such a mesh must be baked by the caller and sent through rt_scene_mesh)
    static bool set_matrix(RtInstance& I, const float m[12], const RtBaseMesh& B) {
        for (int k = 0; k < 12; ++k) { if (!std::isfinite(m[k])) return false; I.m[k] = m[k]; }
        const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
        const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
        const double n1 = std::fmax(std::fabs(a) + std::fabs(b) + std::fabs(c), std::fmax(std::fabs(d) + std::fabs(e) + std::fabs(f), std::fabs(g) + std::fabs(h) + std::fabs(i)));
        if (!(std::fabs(det) > 1e-30) || !(n1 > 0)) return false;
        const double A[9] = {(e * i - f * h) / det, (c * h - b * i) / det, (b * f - c * e) / det,
                             (f * g - d * i) / det, (a * i - c * g) / det, (c * d - a * f) / det,
                             (d * h - e * g) / det, (b * g - a * h) / det, (a * e - b * d) / det};
        double ninv = 0;
        for (int r = 0; r < 3; ++r) ninv = std::fmax(ninv, std::fabs(A[3 * r]) + std::fabs(A[3 * r + 1]) + std::fabs(A[3 * r + 2]));
        if (!std::isfinite(ninv) || n1 * ninv > 1e5) return false;                     // condition number bound
        for (int r = 0; r < 3; ++r) {
            for (int k = 0; k < 3; ++k) I.inv[4 * r + k] = (float)A[3 * r + k];
            I.inv[4 * r + 3] = (float)-(A[3 * r] * m[3] + A[3 * r + 1] * m[7] + A[3 * r + 2] * m[11]);
        }
        I.inv_norm = (float)(ninv * 1.0001);
        // world bounds: the 8 corners of every box of the mesh's cover (RtBaseMesh::hull; the one object box when there is none),
        // in double, rounded outward (padding is added by the callers)
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        if (B.n_tris() == 0) { for (int k = 0; k < 3; ++k) { lo[k] = hi[k] = m[4 * k + 3]; } }
        else {
            const size_t nb = B.hull.empty() ? 1 : B.hull.size() / 6;
            for (size_t bx = 0; bx < nb; ++bx) {
                const float* blo = B.hull.empty() ? B.lo : B.hull.data() + 6 * bx;
                const float* bhi = B.hull.empty() ? B.hi : B.hull.data() + 6 * bx + 3;
                // box of a transformed box: centre through the map, half extents through |M| (the same as the min / max over its 8 corners)
                const double cx = 0.5 * ((double)blo[0] + bhi[0]), cy = 0.5 * ((double)blo[1] + bhi[1]), cz = 0.5 * ((double)blo[2] + bhi[2]);
                const double hx = 0.5 * ((double)bhi[0] - blo[0]), hy = 0.5 * ((double)bhi[1] - blo[1]), hz = 0.5 * ((double)bhi[2] - blo[2]);
                for (int r = 0; r < 3; ++r) {
                    const double wc = (double)m[4 * r] * cx + (double)m[4 * r + 1] * cy + (double)m[4 * r + 2] * cz + (double)m[4 * r + 3];
                    const double wh = (std::fabs((double)m[4 * r]) * hx + std::fabs((double)m[4 * r + 1]) * hy + std::fabs((double)m[4 * r + 2]) * hz) * (1.0 + 1e-12);
                    lo[r] = std::fmin(lo[r], wc - wh); hi[r] = std::fmax(hi[r], wc + wh);
                }
            }
        }
        for (int k = 0; k < 3; ++k) {
            if (!std::isfinite(lo[k]) || !std::isfinite(hi[k]) || std::fabs(lo[k]) > 1e30 || std::fabs(hi[k]) > 1e30) return false;
            I.wlo[k] = std::nextafterf((float)lo[k], -FLT_MAX); I.whi[k] = std::nextafterf((float)hi[k], FLT_MAX);
        }
        return true;
    }

    // object-space bounds of a base mesh; false when a vertex is not finite (object-space culling needs finite boxes)
    static bool set_bounds(RtBaseMesh& B) {
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}, am = 0.f;
        const size_t nv = B.v.size() / 3;
        for (size_t i = 0; i < nv; ++i)
            for (int k = 0; k < 3; ++k) {
                const float x = B.v[3 * i + k], ax = std::fabs(x);
                if (!(ax <= 1e30f)) return false;                     // NaN, infinite or huge
                lo[k] = x < lo[k] ? x : lo[k]; hi[k] = x > hi[k] ? x : hi[k]; am = ax > am ? ax : am;
            }
        if (B.v.empty()) for (int k = 0; k < 3; ++k) lo[k] = hi[k] = 0.f;
        for (int k = 0; k < 3; ++k) { B.lo[k] = lo[k]; B.hi[k] = hi[k]; }
        B.amax = am;
        return true;
    }

    // world bounds of all instances united into b6 / amax (the callers' primitive bounds)
    void grow_bounds(float b6[6], float& amax) const {
        for (auto& I : insts) {
            if (bases[(size_t)I.base].n_tris() == 0) continue;
            for (int k = 0; k < 3; ++k) {
                b6[k] = std::fmin(b6[k], I.wlo[k]); b6[3 + k] = std::fmax(b6[3 + k], I.whi[k]);
                amax = std::fmax(amax, std::fmax(std::fabs(I.wlo[k]), std::fabs(I.whi[k])));
            }
        }
    }

    void build_blas(RtBaseMesh& B, int leaf_max, int threads) {
        const int32_t n = B.n_tris();
        std::vector<RtSahBox> pb((size_t)n);
        for (int32_t t = 0; t < n; ++t) {
            const float* v = B.v.data() + 9 * (size_t)t;
            for (int k = 0; k < 3; ++k) { pb[(size_t)t].lo[k] = std::fmin(v[k], std::fmin(v[3 + k], v[6 + k])); pb[(size_t)t].hi[k] = std::fmax(v[k], std::fmax(v[3 + k], v[6 + k])); }
        }
        RtSahBuilder S;
        B.root = S.run(pb.data(), n, leaf_max, 0.0f, threads);
        B.nodes = S.nodes; B.order = S.order; B.depth = S.max_depth.load();
        B.built = true;
        static std::atomic<uint64_t> next_uid{1};
        B.uid = next_uid.fetch_add(1);
        // cover of the mesh: cut the tree breadth-first after RT_HULL_BOXES pieces (child boxes are stored in the parent node)
        B.hull.clear();
        if (B.root >= 0) {
            struct Piece { int32_t ref; float b[6]; };
            std::vector<Piece> cur, next;
            Piece r0; r0.ref = B.root; for (int k = 0; k < 3; ++k) { r0.b[k] = B.lo[k]; r0.b[3 + k] = B.hi[k]; }
            cur.push_back(r0);
            for (int round = 0; round < 16; ++round) {
                size_t inner = 0; for (auto& p : cur) inner += p.ref >= 0 ? 1 : 0;
                if (inner == 0 || cur.size() + inner > RT_HULL_BOXES) break;
                next.clear();
                for (auto& p : cur) {
                    if (p.ref < 0) { next.push_back(p); continue; }
                    const float* np = B.nodes.data() + 16 * (size_t)p.ref;
                    Piece a, c;
                    std::memcpy(&a.ref, np + 12, 4); std::memcpy(&c.ref, np + 13, 4);
                    a.b[0] = np[0]; a.b[3] = np[1]; a.b[1] = np[2]; a.b[4] = np[3]; a.b[2] = np[8]; a.b[5] = np[9];
                    c.b[0] = np[4]; c.b[3] = np[5]; c.b[1] = np[6]; c.b[4] = np[7]; c.b[2] = np[10]; c.b[5] = np[11];
                    next.push_back(a); next.push_back(c);
                }
                cur.swap(next);
            }
            if (cur.size() > 1) for (auto& p : cur) B.hull.insert(B.hull.end(), p.b, p.b + 6);
        }
    }
    static constexpr size_t RT_HULL_BOXES = 128;

    // Padding of a BLAS, in object units.  A box test inside an instance sees (1) the ray carried into object space in binary32
    // (origin ~4u (|A||O| + |a|), direction ~4u relative, i.e. ~4u |A| x path length), (2) the slab arithmetic itself (~4u (|O'| +
    // |plane|)), and has to cover (3) what the exact world-space test tolerates around a triangle (the reference's own rounding,
    // ~10u |O - v0|, which the world BVH covers with amax * pad_scale) and (4) the rounding of the baked vertices (~3u |w|), both
    // mapped into object space by |A|.  With ray origins within 2 amax_w of the world origin (farther cameras take the FAR form,
    // whose padding grows with the ray: RtScene::origin_reach) all of it is below ~90 u |A| amax_w; the padding is
    // pad_scale (= 256 u) x (|A|_inf amax_w + amax_obj), the largest over the instances of the base mesh.
    static float need_pad(const RtInstance& I, const RtBaseMesh& B, float amax_w, double pad_scale) {
        return (float)(pad_scale * ((double)I.inv_norm * amax_w + B.amax)) + 1e-30f;
    }

    // Builds what is missing and packs everything.  amax_w: max |coordinate| over the whole scene (instances included);
    // pad_w: the world BVH's node padding (TLAS boxes get twice that on top of the rounded-out corner bounds).
    // Returns false when the two-level walk could need more stack than the kernels have (err says so).
    bool pack(const std::vector<int32_t>& mesh_base, float amax_w, float pad_w, double pad_scale, int leaf_max, int threads, int stack_entries, const char** err) {
        ++generation;
        const size_t nb = bases.size();
        for (auto& B : bases) if (!B.built) { build_blas(B, leaf_max, threads); blas_changed = true; }
        // padding each BLAS needs now; re-emit the BLAS sections only when one of them grew beyond what is baked in
        std::vector<float> need(nb, 0.f);
        for (auto& I : insts) need[(size_t)I.base] = std::fmax(need[(size_t)I.base], need_pad(I, bases[(size_t)I.base], amax_w, pad_scale));
        if (blas_pad.size() != nb) blas_changed = true;
        if (!blas_changed) for (size_t b = 0; b < nb; ++b) if (need[b] > blas_pad[b] || need[b] < 0.25f * blas_pad[b]) blas_changed = true;
        if (blas_changed) {
            node_base.assign(nb + 1, 0); slot_base.assign(nb + 1, 0); blas_pad.assign(nb, 0.f);
            packed_uid.assign(nb, 0); for (size_t b = 0; b < nb; ++b) packed_uid[b] = bases[b].uid;
            for (size_t b = 0; b < nb; ++b) { node_base[b + 1] = node_base[b] + (int32_t)(bases[b].nodes.size() / 16); slot_base[b + 1] = slot_base[b] + bases[b].n_tris(); }
            iprims.assign((size_t)slot_base[nb] * 12, 0.f);
            inodes.assign((size_t)node_base[nb] * 16, 0.f);
            for (size_t b = 0; b < nb; ++b) {
                const RtBaseMesh& B = bases[b];
                const float pad = need[b] * 1.5f;                         // head room: small pose changes do not re-emit
                blas_pad[b] = pad;
                for (int32_t s = 0; s < B.n_tris(); ++s) {
                    const int32_t t = B.order[(size_t)s];
                    const float* v = B.v.data() + 9 * (size_t)t;
                    float* pp = iprims.data() + 12 * (size_t)(slot_base[b] + s);
                    pp[0] = v[0]; pp[1] = v[1]; pp[2] = v[2]; std::memcpy(pp + 3, &t, 4);
                    pp[4] = v[3]; pp[5] = v[4]; pp[6] = v[5]; pp[7] = 0.f;
                    pp[8] = v[6]; pp[9] = v[7]; pp[10] = v[8]; pp[11] = 0.f;
                }
                for (size_t i = 0; i < B.nodes.size() / 16; ++i) {
                    const float* src = B.nodes.data() + 16 * i;
                    float* dst = inodes.data() + 16 * ((size_t)node_base[b] + i);
                    for (int k = 0; k < 12; ++k) dst[k] = (k & 1) ? src[k] + pad : src[k] - pad;      // (lo, hi) pairs
                    for (int k = 0; k < 2; ++k) {
                        int32_t r; std::memcpy(&r, src + 12 + k, 4);
                        r = fix_ref(r, node_base[b], slot_base[b]);
                        std::memcpy(dst + 12 + k, &r, 4);
                    }
                }
            }
            ++blas_generation;
            blas_changed = false;
        }
        tlas_base = node_base.empty() ? 0 : node_base[nb];
        // ---- TLAS over the instances that have triangles ----
        std::vector<int32_t> live;
        for (size_t i = 0; i < insts.size(); ++i) if (bases[(size_t)insts[i].base].n_tris() > 0) live.push_back((int32_t)i);
        n_packed = (int32_t)live.size();
        std::vector<RtSahBox> ib((size_t)n_packed);
        const float tpad = 2.0f * pad_w;
        for (int32_t k = 0; k < n_packed; ++k) {
            const RtInstance& I = insts[(size_t)live[(size_t)k]];
            for (int a = 0; a < 3; ++a) { ib[(size_t)k].lo[a] = I.wlo[a] - tpad; ib[(size_t)k].hi[a] = I.whi[a] + tpad; }
        }
        RtSahBuilder T;
        int32_t troot = T.run(ib.data(), n_packed, 1, pad_w, 1);
        inodes.resize((size_t)tlas_base * 16 + T.nodes.size());
        for (size_t i = 0; i < T.nodes.size() / 16; ++i) {
            const float* src = T.nodes.data() + 16 * i;
            float* dst = inodes.data() + 16 * ((size_t)tlas_base + i);
            std::memcpy(dst, src, 64);
            for (int k = 0; k < 2; ++k) { int32_t r; std::memcpy(&r, src + 12 + k, 4); r = fix_ref(r, tlas_base, 0); std::memcpy(dst + 12 + k, &r, 4); }
        }
        tlas_root = n_packed > 0 ? fix_ref(troot, tlas_base, 0) : -1;
        inst.assign((size_t)n_packed * 32, 0.f);
        vbase.assign((size_t)n_packed + 1, 0);
        int max_blas_depth = 0;
        for (int32_t k = 0; k < n_packed; ++k) {
            const RtInstance& I = insts[(size_t)live[(size_t)T.order[(size_t)k]]];
            const RtBaseMesh& B = bases[(size_t)I.base];
            float* r = inst.data() + 32 * (size_t)k;
            std::memcpy(r, I.m, 48); std::memcpy(r + 12, I.inv, 48);
            const int32_t ints[4] = {fix_ref(B.root, node_base[(size_t)I.base], slot_base[(size_t)I.base]), slot_base[(size_t)I.base], I.mesh, mesh_base[(size_t)I.mesh]};
            std::memcpy(r + 24, ints, 16);
            const int32_t vb = vbase[(size_t)k], mat = I.mesh;            // material id == MeshIndex (RtHostScene::pack_materials)
            std::memcpy(r + 28, &vb, 4); std::memcpy(r + 29, &mat, 4);
            r[30] = 4.0f * I.inv_norm; r[31] = 0.f;
            vbase[(size_t)k + 1] = vb + B.n_tris();
            max_blas_depth = B.depth > max_blas_depth ? B.depth : max_blas_depth;
        }
        depth = T.max_depth.load() + max_blas_depth + 6;                 // + sentinels and the split leaf entries
        if (depth > stack_entries) { if (err) *err = "TLAS + BLAS deeper than the traversal stack"; return false; }
        return true;
    }

    // BLAS/TLAS-relative traversal ref -> ref into the packed arrays
    static int32_t fix_ref(int32_t r, int32_t node_off, int32_t slot_off) {
        if (r >= 0) return r + node_off;
        const int32_t x = ~r;
        return ~(((x >> 3) + slot_off) * 8 + (x & 7));
    }
};

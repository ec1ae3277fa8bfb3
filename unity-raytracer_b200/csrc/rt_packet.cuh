// rt_packet.cuh — warp-coherent ("packet") BVH traversal, device only.
//
// Same query as rt_traverse (rt_core.h) and the same exact primitive tests, but control flow is warp-uniform: all 32
// lanes walk ONE path, the union of what any lane needs.  Each lane still tests every visited node's child boxes and
// every visited leaf's primitives with its OWN ray and its OWN cull distance, so each lane tests a superset of the
// primitives its private traversal would test.  A superset cannot change argmin over (t, scan key) nor the any-hit
// predicate, hence results are bit-identical to the per-lane traversal (and to the reference's scan).
//
// Why: ncu on the per-lane while-while loop showed 13 of 32 lanes active on average (profiles/r1_ncu_summary.md):
// lanes that reach a leaf early idle while the others descend.  Primary rays of an 8x4 pixel tile, and shadow rays from
// that tile to one point light, are almost parallel, so the union path is barely longer than a single lane's path while
// no lane ever idles; node and primitive loads become warp-uniform (one L1 wavefront instead of up to 32).
// Every lane keeps a private copy of the (identical) stack in its shared-memory column, so a warp can drop to the
// per-lane traversal at any point without conversion.
//
// All 32 lanes of the warp must call; lanes without a ray pass active = false.
#pragma once
#include "rt_core.h"

// Cheap coherence test: packet traversal pays off when the rays are nearly parallel and start close together.
__device__ __forceinline__ bool rt_warp_coherent(f3 O, f3 D, bool active, float max_origin_spread2) {
    const unsigned full = 0xffffffffu;
    unsigned m = __ballot_sync(full, active);
    if (m == 0) return true;
    int src = __ffs(m) - 1;
    float ox = __shfl_sync(full, O.x, src), oy = __shfl_sync(full, O.y, src), oz = __shfl_sync(full, O.z, src);
    float dx = __shfl_sync(full, D.x, src), dy = __shfl_sync(full, D.y, src), dz = __shfl_sync(full, D.z, src);
    float c = D.x * dx + D.y * dy + D.z * dz;
    float l2 = (D.x * D.x + D.y * D.y + D.z * D.z) * (dx * dx + dy * dy + dz * dz);
    float ex = O.x - ox, ey = O.y - oy, ez = O.z - oz;
    bool ok = !active || (c > 0.f && c * c > 0.9f * l2 && (ex * ex + ey * ey + ez * ez) <= max_origin_spread2);   // cos > ~0.95
    return __all_sync(full, ok);
}

// Inner-node loop of the packet traversal, specialised on the direction octant the whole warp shares (OCT bit 0/1/2 = x/y/z
// direction negative; OCT = 8: lanes disagree, generic form).  With the sign known, which slab plane is the near one is
// known too, so the per-axis min/max pairs of the generic slab test disappear: 20 instead of 36 FMNMX/FFMA per node
// (the slab planes are monotone under the FMA, so near <= far always holds and the result is the same interval).
// Returns the next leaf reference (< 0), or RT_PACKET_DONE when the stack ran empty.
#define RT_PACKET_DONE ((int32_t)0x80000000)
#ifdef RT_CHECKED
#define RT_CHK_PARAM , int32_t n_bound
#define RT_CHK_ARG , ((INST && level > 0) ? S.n_inodes : S.n_nodes)
#else
#define RT_CHK_PARAM
#define RT_CHK_ARG
#endif
// Distance-culled pops (RT_POP_CULL, default on).  Every pushed reference carries, per lane, the lane's entry distance into
// that box (+inf when the lane missed it).  Cull distances only shrink (closest hit found; any-hit lanes that have their
// occluder drop to -1), so  entry > cull  now implies the lane's own box test would fail now: an entry no lane can reach is
// dropped on pop without fetching the node or testing the leaf's primitives.  Conservative in the same sense as the box
// test itself, hence the result is unchanged.  STK: push_packet(ref, dist), pop_packet(dist&), empty().
#ifndef RT_POP_CULL
#define RT_POP_CULL 1
#endif
template <class STK>
__device__ __forceinline__ int32_t rt_packet_pop(STK& stk, float tcull RT_CNT_PARAM) {
#if RT_POP_CULL
    for (;;) {
        if (stk.empty()) return RT_PACKET_DONE;
        float d;
        int32_t node = stk.pop_packet(d);
        if (__any_sync(0xffffffffu, d <= tcull)) return node;
#ifdef RT_COUNT_VISITS
        if ((threadIdx.x & 31) == 0) RT_COUNT(cbase + 9, 1);
#endif
    }
#else
    if (stk.empty()) return RT_PACKET_DONE;
    return stk.pop();
#endif
}
// Child order when both are hit (ORD), by majority of the lanes that hit both: 0 = nearer entry first (closest-hit queries:
// the cull distance shrinks fastest); 2 = any-hit queries: the child in which the ray stays LONGER first.  The result of an
// any-hit query does not depend on the order, only how soon the lanes find their occluder: cfg3 shadow kernel -2.8 %, cfg5
// frame -6.5 % against nearest-first.  Rejected: farthest-first (+11 %); "child more lanes still need first" (same gain on
// a full frame but the heaviest tiles get longer: +10 % on the shadow kernel of a 1/8 partition).
#ifndef RT_ANY_ORDER
#define RT_ANY_ORDER 2
#endif
template <int OCT, int ORD, class STK>
__device__ __forceinline__ int32_t rt_packet_descend(const float4* __restrict__ nodes, float idx_, float idy_, float idz_,
                                                     float oodx, float oody, float oodz, float tcull, STK& stk, int32_t node RT_CNT_PARAM RT_CHK_PARAM) {
    const unsigned full = 0xffffffffu;
    while (node >= 0) {
        RT_CHECK(RT_CHK_NODE, node < n_bound);
        const float4* np = nodes + 4 * (size_t)node;
        float4 n0 = RT_LDG4(np), n1 = RT_LDG4(np + 1), n2 = RT_LDG4(np + 2), n3 = RT_LDG4(np + 3);
#ifdef RT_COUNT_VISITS
        { unsigned live = __ballot_sync(full, tcull >= 0.0f);
          if ((threadIdx.x & 31) == 0) { RT_COUNT(cbase + 5, 1); RT_COUNT(cbase + 6, __popc(live)); } }
#endif
        float amin, amax, bmin, bmax;
        if (OCT == 8) {
            float ax0 = rt_fma(n0.x, idx_, -oodx), ax1 = rt_fma(n0.y, idx_, -oodx);
            float ay0 = rt_fma(n0.z, idy_, -oody), ay1 = rt_fma(n0.w, idy_, -oody);
            float az0 = rt_fma(n2.x, idz_, -oodz), az1 = rt_fma(n2.y, idz_, -oodz);
            float bx0 = rt_fma(n1.x, idx_, -oodx), bx1 = rt_fma(n1.y, idx_, -oodx);
            float by0 = rt_fma(n1.z, idy_, -oody), by1 = rt_fma(n1.w, idy_, -oody);
            float bz0 = rt_fma(n2.z, idz_, -oodz), bz1 = rt_fma(n2.w, idz_, -oodz);
            amin = fmaxf(fmaxf(fminf(ax0, ax1), fminf(ay0, ay1)), fmaxf(fminf(az0, az1), 0.0f));
            amax = fminf(fminf(fmaxf(ax0, ax1), fmaxf(ay0, ay1)), fminf(fmaxf(az0, az1), tcull));
            bmin = fmaxf(fmaxf(fminf(bx0, bx1), fminf(by0, by1)), fmaxf(fminf(bz0, bz1), 0.0f));
            bmax = fminf(fminf(fmaxf(bx0, bx1), fmaxf(by0, by1)), fminf(fmaxf(bz0, bz1), tcull));
        } else {
            constexpr bool sx = (OCT & 1) != 0, sy = (OCT & 2) != 0, sz = (OCT & 4) != 0;
            float anx = rt_fma(sx ? n0.y : n0.x, idx_, -oodx), afx = rt_fma(sx ? n0.x : n0.y, idx_, -oodx);
            float any_ = rt_fma(sy ? n0.w : n0.z, idy_, -oody), afy = rt_fma(sy ? n0.z : n0.w, idy_, -oody);
            float anz = rt_fma(sz ? n2.y : n2.x, idz_, -oodz), afz = rt_fma(sz ? n2.x : n2.y, idz_, -oodz);
            float bnx = rt_fma(sx ? n1.y : n1.x, idx_, -oodx), bfx = rt_fma(sx ? n1.x : n1.y, idx_, -oodx);
            float bny = rt_fma(sy ? n1.w : n1.z, idy_, -oody), bfy = rt_fma(sy ? n1.z : n1.w, idy_, -oody);
            float bnz = rt_fma(sz ? n2.w : n2.z, idz_, -oodz), bfz = rt_fma(sz ? n2.z : n2.w, idz_, -oodz);
            amin = fmaxf(fmaxf(anx, any_), fmaxf(anz, 0.0f));
            amax = fminf(fminf(afx, afy), fminf(afz, tcull));
            bmin = fmaxf(fmaxf(bnx, bny), fmaxf(bnz, 0.0f));
            bmax = fminf(fminf(bfx, bfy), fminf(bfz, tcull));
        }
        bool ha = amin <= amax, hb = bmin <= bmax;
        unsigned ma = __ballot_sync(full, ha), mb = __ballot_sync(full, hb);
        int32_t ra = f2i(n3.x), rb = f2i(n3.y);
        if (ma && mb) {
            unsigned both = ma & mb;
            unsigned bnear = ORD == 0 ? __ballot_sync(full, ha && hb && bmin < amin)
                                      : __ballot_sync(full, ha && hb && (bmax - bmin) > (amax - amin));
            const bool swap = 2 * __popc(bnear) > __popc(both);                              // majority vote on the order
#if RT_POP_CULL
            const float da = ha ? amin : INFINITY, db = hb ? bmin : INFINITY;
            stk.push_packet(swap ? ra : rb, swap ? da : db);
#else
            stk.push(swap ? ra : rb);
#endif
            node = swap ? rb : ra;
        } else if (ma) node = ra;
        else if (mb) node = rb;
        else {
            node = rt_packet_pop(stk, tcull RT_CNT_FWD);
            if (node == RT_PACKET_DONE) return node;
        }
    }
    return node;
}

// direction octant shared by every lane that has a ray (8 = mixed); lanes without a ray fail every box test anyway
__device__ __forceinline__ int rt_packet_octant(float idx_, float idy_, float idz_, bool active) {
    const unsigned full = 0xffffffffu;
    int oct = (idx_ < 0.0f ? 1 : 0) | (idy_ < 0.0f ? 2 : 0) | (idz_ < 0.0f ? 4 : 0);
    unsigned am = __ballot_sync(full, active);
    int lead = __shfl_sync(full, oct, am ? __ffs(am) - 1 : 0);
    bool sane = idx_ == idx_ && idy_ == idy_ && idz_ == idz_;
    return __all_sync(full, !active || (oct == lead && sane)) ? lead : 8;
}

// INST: two-level walk for scenes with instanced meshes, as in rt_traverse<.., INST> (rt_core.h "Instancing"): the packet walks
// the world BVH, then the TLAS, and inside an instance the base mesh's BLAS with every lane's ray carried into object space for
// the box tests (an affine map keeps the rays of a packet coherent and their t unchanged); triangle tests stay in world space
// on the vertices baked on the fly.  The level and the current instance are warp-uniform.
// bake: INST only — 9 float4 of shared memory owned by this warp (RT_BAKE_F4): the world-space vertices of the current BLAS leaf.
// A triangle of an instance is the same for all 32 lanes, so its bake (27 multiplications + 27 additions, SceneMesh.cs:28) is done
// ONCE per leaf by the warp — lane i computes coordinate i % 3 of vertex i / 3 with the reference's operation order — instead of
// by every lane for itself: the instanced triangle test then costs what the baked one costs (ncu, cfg4 as instances: the
// per-lane bake was 20 % of the fused shadow kernel's instructions, its triangle tests twice the baked scene's).
#define RT_BAKE_F4 9
#ifndef RT_OCT_GENERIC
#define RT_OCT_GENERIC 0            /* development: 1 = generic node loop in every instantiation */
#endif
#ifndef RT_INST_OCT
#define RT_INST_OCT 0
#endif
template <bool ANY, bool INST = false, class STK>
__device__ __forceinline__ bool rt_traverse_packet(const RtScene& S, f3 O, f3 D, f3 inv, bool active, float any_d2, STK& stk,
                                                   float& best_t, int32_t& best_slot, float4* bake = nullptr) {
    const unsigned full = 0xffffffffu;
    best_t = FLT_MAX; best_slot = -1;
    stk.sp = 0;
    if (S.n_prims <= 0 && !(INST && S.n_inst > 0)) return false;
    float idx_ = rt_safe_inv(D.x, inv.x), idy_ = rt_safe_inv(D.y, inv.y), idz_ = rt_safe_inv(D.z, inv.z);
    float oodx = O.x * idx_, oody = O.y * idy_, oodz = O.z * idz_;
    uint32_t best_key = 0xffffffffu;
    float tcull = ANY ? rt_sqrt(any_d2) * S.t_slack : INFINITY;
    if (!active) tcull = -1.0f;                       // every box test of this lane fails (tmin >= 0 > tcull)
    int32_t gate_mesh = -1; bool gate_ok = true;
    bool found = false;                               // ANY: this lane already has its occluder
    int oct = rt_packet_octant(idx_, idy_, idz_, active);
    const float4* __restrict__ nodes = S.nodes;
    int32_t node = S.root;                            // warp-uniform from here on
    int32_t level = 0, cur = -1;                      // INST: 0 = world BVH, 1 = TLAS, 2 = BLAS of instance `cur` (warp-uniform)
    int32_t cur_mesh = -1, cur_key0 = 0, cur_vslot0 = 0;   // of instance `cur`: MeshIndex, scan key of its triangle 0, virtual slot of BLAS slot 0
    if (INST && S.n_inst > 0) { if (S.n_prims > 0) stk.push_packet(RT_SENT_TLAS, -INFINITY); else node = RT_SENT_TLAS; }
#ifdef RT_COUNT_VISITS
    { unsigned am = __ballot_sync(full, active);
      if ((threadIdx.x & 31) == 0) { RT_COUNT((ANY ? 10 : 0) + 3, 1); RT_COUNT((ANY ? 10 : 0) + 4, __popc(am)); } }
#endif
    for (;;) {
        if (INST && node < 0 && node != RT_PACKET_DONE) {
            if (node <= RT_SENT_EXIT) {                                 // a sentinel, not a leaf
                if (node == RT_SENT_TLAS) { nodes = S.inodes; level = 1; node = S.tlas_root; continue; }
                idx_ = rt_safe_inv(D.x, inv.x); idy_ = rt_safe_inv(D.y, inv.y); idz_ = rt_safe_inv(D.z, inv.z);     // RT_SENT_EXIT
                oodx = O.x * idx_; oody = O.y * idy_; oodz = O.z * idz_;
                oct = rt_packet_octant(idx_, idy_, idz_, active);
                level = 1;
                node = rt_packet_pop(stk, tcull RT_CNT_ARG(ANY));
                if (node == RT_PACKET_DONE) return ANY ? found : best_slot >= 0;
                continue;
            }
            if (level == 1) {                                           // TLAS leaf: instances [first, first + cnt)
                const int32_t ref = ~node, first = ref >> 3, cnt = (ref & 7) + 1;
                RT_CHECK(RT_CHK_TLAS_LEAF, first >= 0 && first + cnt <= S.n_inst);
                if (cnt > 1) stk.push_packet(~((first + 1) * 8 + (cnt - 2)), -INFINITY);     // the rest of the leaf (its box was passed already)
                stk.push_packet(RT_SENT_EXIT, -INFINITY);
                cur = first;
                const float4* ip = S.inst + RT_INST_F4 * (size_t)cur;
                const float4 a0 = RT_LDG4(ip + 3), a1 = RT_LDG4(ip + 4), a2 = RT_LDG4(ip + 5);
                const f3 Oo = rt_xform_point_cull(a0, a1, a2, O), Do = rt_xform_dir_cull(a0, a1, a2, D);
                idx_ = rt_safe_inv(Do.x, rt_rcp_cull(Do.x)); idy_ = rt_safe_inv(Do.y, rt_rcp_cull(Do.y)); idz_ = rt_safe_inv(Do.z, rt_rcp_cull(Do.z));
                oodx = Oo.x * idx_; oody = Oo.y * idy_; oodz = Oo.z * idz_;
                oct = rt_packet_octant(idx_, idy_, idz_, active);
                level = 2;
                const float4 ii = RT_LDG4(ip + 6);
                cur_mesh = f2i(ii.z); cur_key0 = f2i(ii.w); cur_vslot0 = S.n_slots + f2i(RT_LDG4(ip + 7).x) - f2i(ii.y);
                node = f2i(ii.x);
                continue;
            }
        }
        // RT_INST_OCT (INST instantiations): 2 = octant-specialised node loops everywhere, 1 = in the world BVH / TLAS only (inside an
        // instance the octant is the OBJECT-space one, which differs from instance to instance: all eight copies of the loop are then
        // hot at once and compete for the 32 KB instruction cache), 0 = generic loop only
        if ((INST && RT_INST_OCT == 0) || RT_OCT_GENERIC) { if (node >= 0) node = rt_packet_descend<8, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); }
        else if (!INST || node >= 0)
        switch ((INST && RT_INST_OCT == 1 && level == 2) ? 8 : oct) {
            case 0: node = rt_packet_descend<0, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            case 1: node = rt_packet_descend<1, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            case 2: node = rt_packet_descend<2, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            case 3: node = rt_packet_descend<3, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            case 4: node = rt_packet_descend<4, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            case 5: node = rt_packet_descend<5, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            case 6: node = rt_packet_descend<6, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            case 7: node = rt_packet_descend<7, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
            default: node = rt_packet_descend<8, (ANY ? RT_ANY_ORDER : 0)>(nodes, idx_, idy_, idz_, oodx, oody, oodz, tcull, stk, node RT_CNT_ARG(ANY) RT_CHK_ARG); break;
        }
        if (node == RT_PACKET_DONE) return ANY ? found : best_slot >= 0;
        if (INST && (node <= RT_SENT_EXIT || level == 1)) continue;     // sentinel or TLAS leaf: handled at the top of the loop
        {
            int32_t ref = ~node;
            int32_t first = ref >> 3, cnt = (ref & 7) + 1;
            RT_CHECK(RT_CHK_LEAF, first >= 0 && first + cnt <= ((INST && level == 2) ? S.n_islots : S.n_slots));
            RT_CHECK(RT_CHK_BAKE, !(INST && level == 2) || cnt <= 3);
            if (INST && level == 2) {                                   // bake the leaf's (<= 3) triangles once for the whole warp
                const int lane = threadIdx.x & 31;
                __syncwarp(full);                                       // every lane is done with the previous leaf's vertices
                if (lane < 9 * cnt) {
                    const int vtx = lane / 3, comp = lane - 3 * vtx;
                    const float4 pv = RT_LDG4(S.iprims + 3 * (size_t)first + vtx);
                    const float4 mr = RT_LDG4(S.inst + RT_INST_F4 * (size_t)cur + comp);
                    float* wb = (float*)bake;
                    wb[4 * vtx + comp] = ((mr.x * pv.x + mr.y * pv.y) + mr.z * pv.z) + mr.w;      // MultiplyPoint3x4, exact operation order
                    if (comp == 0) wb[4 * vtx + 3] = pv.w;              // (vertex 0 of a triangle carries its index in the mesh)
                }
                __syncwarp(full);
            }
            for (int32_t k = 0; k < cnt; ++k) {
                int32_t slot = first + k;
                float4 p0, p1, p2;
                if (INST && level == 2) { p0 = bake[3 * k]; p1 = bake[3 * k + 1]; p2 = bake[3 * k + 2]; }
                else { const float4* pp = S.prims + 3 * (size_t)slot; p0 = RT_LDG4(pp); p1 = RT_LDG4(pp + 1); p2 = RT_LDG4(pp + 2); }
#ifdef RT_COUNT_VISITS
                { unsigned live = __ballot_sync(full, tcull >= 0.0f);
                  if ((threadIdx.x & 31) == 0) { RT_COUNT((ANY ? 10 : 0) + 7, 1); RT_COUNT((ANY ? 10 : 0) + 8, __popc(live)); } }
#endif
                float t = 0.0f; bool hit = false;
                uint32_t key; int32_t mesh;
                if (INST && level == 2) {                               // triangle of instance `cur`, baked above: tested in world space
                    if (tcull >= 0.0f) {
                        const f3 w0 = mk3(p0.x, p0.y, p0.z);
                        hit = rt_ray_tri(O, D, w0, mk3(p1.x, p1.y, p1.z) - w0, mk3(p2.x, p2.y, p2.z) - w0, tcull, t);      // edge1, edge2: RMath.cs:34-35
                    }
                    key = (uint32_t)(cur_key0 + f2i(p0.w)); mesh = cur_mesh;
                    slot = cur_vslot0 + slot;                           // virtual slot (rt_inst_decode)
                } else {
                if (tcull >= 0.0f) {                                    // else: lane has no ray / is already occluded
                    if (f2i(p1.w) != 0) hit = rt_ray_sphere(O, D, mk3(p0.x, p0.y, p0.z), p1.x, t);
                    else hit = rt_ray_tri(O, D, mk3(p0.x, p0.y, p0.z), mk3(p1.x, p1.y, p1.z), mk3(p2.x, p2.y, p2.z), tcull, t);
                }
                key = (uint32_t)f2i(p0.w); mesh = f2i(p2.w);
                }
                bool cand = hit && (ANY ? (t * t < any_d2) : (t < best_t || (t == best_t && key < best_key)));
                if (!__any_sync(full, cand)) continue;                  // warp-uniform: nobody would accept this primitive
                // Per-mesh AABB gate (Scene.cs:67) of a candidate.  The primitive, hence its mesh, is the same for all lanes, so
                // the gate is evaluated by the WHOLE warp at once (each lane for its own ray, exact formula) and kept until the
                // walk reaches a primitive of another mesh: once per query on a one-mesh scene.  Evaluating it lane by lane as
                // each lane met its first candidate cost up to 32 divergent evaluations per query (ncu: ~5 % of the fused shadow
                // kernel's instructions).  rcp(ray.Direction) is recomputed here — same bits as the caller's `inv` — which frees
                // three registers across the loops.
                if (mesh >= 0) {
                    RT_CHECK(RT_CHK_MESH, mesh < S.n_meshes);
                    if (mesh != gate_mesh) {
                        gate_mesh = mesh;
                        float4 lo = RT_LDG4(S.mesh_aabb + 2 * mesh), hi = RT_LDG4(S.mesh_aabb + 2 * mesh + 1);
                        gate_ok = rt_ray_aabb_inv(O, rt_rcp3(D), mk3(lo.x, lo.y, lo.z), mk3(hi.x, hi.y, hi.z));
                    }
                    cand = cand && gate_ok;
                }
                if (cand) {
                    best_t = t; best_key = key; best_slot = slot;
                    if (ANY) { found = true; tcull = -1.0f; }
                    else tcull = t * S.t_slack;
                }
            }
        }
        if (ANY && __all_sync(full, tcull < 0.0f)) return found;        // every lane is occluded or has no ray
        node = rt_packet_pop(stk, tcull RT_CNT_ARG(ANY));
        if (node == RT_PACKET_DONE) return ANY ? found : best_slot >= 0;
    }
}

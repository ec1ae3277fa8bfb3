// rt_emu.cpp — TEST HARNESS ONLY.  Runs the product's host/device headers (rt_core.h, rt_build_steps.h: the very code the
// CUDA kernels instantiate) serially on the CPU so the BVH build, traversal, shading, wave/depth bookkeeping and pixel
// mapping can be checked against the oracle in the `-m "not gpu"` suite, where no GPU exists.
// It exports only emu_* symbols, is built into tests/emu/ and is never loaded by the product package: librtb200.so has
// no CPU path and fails without a CUDA device.
#include <vector>
#include <algorithm>
#include <numeric>
#include <cstdio>
#include "../../unity-raytracer_b200/csrc/rt_core.h"
#include "../../unity-raytracer_b200/csrc/rt_build_steps.h"
#include "../../unity-raytracer_b200/csrc/rt_host_scene.h"
#include "../../unity-raytracer_b200/csrc/rt_sah.h"
#include "../../unity-raytracer_b200/csrc/rt_instances.h"

namespace {

struct VecStack {
    int buf[256]; int sp = 0; int max_sp = 0;
    void push(int v) { buf[sp++] = v; if (sp > max_sp) max_sp = sp; }
    int pop() { return buf[--sp]; }
    bool empty() const { return sp == 0; }
};

struct Emu {
    RtHostScene host;
    // "device" memory
    std::vector<float> raw_v, raw_n; std::vector<float4> raw_s;
    std::vector<int32_t> mesh_base, mesh_raw;
    std::vector<float4> nodes, prims, pinfo, mats, mesh_aabb, lights;
    std::vector<float4> inodes, iprims, inst; std::vector<int32_t> inst_vbase;      // RtDevice::inodes / iprims / inst / inst_vbase
    std::vector<int32_t> order;               // slot -> scan key of the committed tree (the device keeps it in vals_alt)
    int32_t root = 0;
    double presplit = 0.0; int presplit_depth = 3;   // RtOptions::presplit / presplit_depth
    RtScene S{};
    int n_nodes = 0, max_stack = 0;
    bool has_mirror = false;
    float obox[6] = {0, 0, 0, 0, 0, 0};        // RtDevice::obox
    int opt_leaf_max = 4, opt_builder = 0; double opt_pad_scale = 1.0 / 65536.0, opt_t_slack = 1.000125;   // of the last commit
};

void commit(Emu& E, int leaf_max, double pad_scale, double t_slack, int builder, bool refit = false) {
    RtHostScene& H = E.host;
    const int n_mt = H.n_meshtris(), n_sp = H.n_spheres(), n_lo = H.n_loose(), n_me = H.n_meshes(), n_li = H.n_lights();
    const int n = (int)H.n_prims();
    E.raw_v = H.mesh_v; E.raw_v.insert(E.raw_v.end(), H.loose_v.begin(), H.loose_v.end());
    E.raw_n = H.mesh_n; E.raw_n.insert(E.raw_n.end(), H.loose_n.begin(), H.loose_n.end());
    E.raw_s.resize(n_sp);
    for (int i = 0; i < n_sp; ++i) E.raw_s[i] = make_float4(H.spheres[4 * i], H.spheres[4 * i + 1], H.spheres[4 * i + 2], H.spheres[4 * i + 3]);
    const int n_kt = H.n_keytris();
    E.mesh_base = H.mesh_base; E.mesh_base.push_back(n_kt);
    E.mesh_raw = H.mesh_raw; E.mesh_raw.push_back(n_mt);
    std::vector<float> m = H.pack_materials();
    E.has_mirror = H.has_mirror;
    E.mats.resize(m.size() / 4);
    if (!m.empty()) memcpy(E.mats.data(), m.data(), m.size() * 4);            // (memcpy with a null source is undefined even for 0 bytes: UBSan)
    E.mesh_aabb.resize(2 * n_me);
    for (int k = 0; k < n_me; ++k) {
        E.mesh_aabb[2 * k] = make_float4(H.mesh_aabb[6 * k], H.mesh_aabb[6 * k + 1], H.mesh_aabb[6 * k + 2], 0.f);
        E.mesh_aabb[2 * k + 1] = make_float4(H.mesh_aabb[6 * k + 3], H.mesh_aabb[6 * k + 4], H.mesh_aabb[6 * k + 5], 0.f);
    }
    E.lights.resize(2 * n_li);
    for (int k = 0; k < n_li; ++k) {
        E.lights[2 * k] = make_float4(H.lights[6 * k], H.lights[6 * k + 1], H.lights[6 * k + 2], 0.f);
        E.lights[2 * k + 1] = make_float4(H.lights[6 * k + 3], H.lights[6 * k + 4], H.lights[6 * k + 5], 0.f);
    }
    RtScene& S = E.S; S = RtScene{};
    S.n_prims = n; S.n_meshes = n_me; S.n_meshtris = n_kt; S.n_spheres = n_sp; S.n_lights = n_li;
    S.aabb_min = mk3(H.aabb[0], H.aabb[1], H.aabb[2]); S.aabb_max = mk3(H.aabb[3], H.aabb[4], H.aabb[5]);
    S.ambient = mk3(H.ambient[0], H.ambient[1], H.ambient[2]);
    S.t_slack = (float)t_slack;
    S.backcull = H.backcull_ok() ? 1 : 0;
    S.mats = E.mats.data(); S.mesh_aabb = E.mesh_aabb.data(); S.mesh_base = E.mesh_base.data(); S.lights = E.lights.data();
    if (!refit) E.n_nodes = 0;
    if (n == 0 && !H.inst.any()) return;
    RtRaw R{E.raw_v.data(), E.raw_n.data(), E.raw_s.data(), E.mesh_raw.data(), n_me, n_mt, n_sp, n_lo, mk3(0.f, 0.f, 0.f), mk3(0.f, 0.f, 0.f), 0.0f,
            E.mesh_base.data(), n_kt};
    leaf_max = leaf_max < 1 ? 1 : (leaf_max > 8 ? 8 : leaf_max);
    float b[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = 0; i < n; ++i) {                                  // k_bounds: un-inflated sphere boxes (R.sphere_k = 0)
        f3 lo, hi; rt_prim_box(R, i, lo, hi);
        b[0] = fminf(b[0], lo.x); b[1] = fminf(b[1], lo.y); b[2] = fminf(b[2], lo.z);
        b[3] = fmaxf(b[3], hi.x); b[4] = fmaxf(b[4], hi.y); b[5] = fmaxf(b[5], hi.z);
    }
    float amax = 0.f; for (int k = 0; k < 6; ++k) if (std::isfinite(b[k])) amax = fmaxf(amax, fabsf(b[k]));
    float bo[6]; for (int k = 0; k < 6; ++k) bo[k] = b[k];
    H.inst.grow_bounds(bo, amax);                                  // origin box and padding: instances included (rt_device_commit)
    const float pad = (float)(amax * pad_scale) + 1e-30f;
    S.origin_reach = rt_origin_reach(pad);
    H.origin_box(bo, amax, E.obox);                                // rt_device_commit's set_obox
    if (H.inst.any()) {                                            // rt_device_commit's upload_instances
        const char* why = nullptr;
        if (!H.inst.pack(E.mesh_base, amax, pad, pad_scale, leaf_max > 3 ? 3 : leaf_max, 2, 116, &why)) { fprintf(stderr, "emu: %s\n", why ? why : "pack failed"); abort(); }
        auto as4 = [](const std::vector<float>& v, std::vector<float4>& o) { o.assign(v.size() / 4 + 1, float4{}); if (!v.empty()) memcpy(o.data(), v.data(), v.size() * 4); };
        as4(H.inst.inodes, E.inodes); as4(H.inst.iprims, E.iprims); as4(H.inst.inst, E.inst);
        E.inst_vbase = H.inst.vbase;
        S.inodes = E.inodes.data(); S.iprims = E.iprims.data(); S.inst = E.inst.data(); S.inst_vbase = E.inst_vbase.data();
        S.tlas_root = H.inst.tlas_root; S.n_inst = H.inst.n_packed;
        S.origin_reach = fminf(S.origin_reach, 2.0f * amax);
    }
    if (n == 0) { S.n_slots = 0; return; }
    R.obox_lo = mk3(E.obox[0], E.obox[1], E.obox[2]); R.obox_hi = mk3(E.obox[3], E.obox[4], E.obox[5]); R.sphere_k = RT_SPHERE_K;
    S.n_slots = refit ? (int)E.order.size() : n;
    f3 bmin = mk3(b[0], b[1], b[2]);
    f3 inv_ext = mk3(b[3] > b[0] ? 1.0f / (b[3] - b[0]) : 0.f, b[4] > b[1] ? 1.0f / (b[4] - b[1]) : 0.f, b[5] > b[2] ? 1.0f / (b[5] - b[2]) : 0.f);
    if (refit) {                                                   // rt_scene_refit: k_emit_slots + k_node_parents + k_refit_nodes, serialised
        const int ns = (int)E.order.size();                         // slots (>= n for a pre-split tree)
        E.prims.assign((size_t)3 * ns, float4{}); E.pinfo.assign(ns, float4{});
        std::vector<float4> leaf_lo(ns), leaf_hi(ns);
        for (int s = 0; s < ns; ++s) rt_emit_slot(R, s, E.order[s], E.prims.data(), E.pinfo.data(), leaf_lo.data(), leaf_hi.data());
        const int nn = E.n_nodes;
        std::vector<int32_t> parent(nn, -1), arrived(nn, 0);
        for (int i = 0; i < nn; ++i) rt_node_parents(E.nodes.data(), i, parent.data());
        std::vector<float4> int_lo(nn), int_hi(nn);
        for (int s0 = 0; s0 < nn; ++s0) {
            float4 n3 = E.nodes[4 * (size_t)s0 + 3];
            if (f2i(n3.x) >= 0 || f2i(n3.y) >= 0) continue;
            int i = s0;
            for (;;) {
                float4* np = E.nodes.data() + 4 * (size_t)i;
                f3 alo, ahi, blo, bhi;
                rt_ref_box(f2i(np[3].x), leaf_lo.data(), leaf_hi.data(), int_lo.data(), int_hi.data(), alo, ahi);
                rt_ref_box(f2i(np[3].y), leaf_lo.data(), leaf_hi.data(), int_lo.data(), int_hi.data(), blo, bhi);
                np[0] = make_float4(alo.x - pad, ahi.x + pad, alo.y - pad, ahi.y + pad);
                np[1] = make_float4(blo.x - pad, bhi.x + pad, blo.y - pad, bhi.y + pad);
                np[2] = make_float4(alo.z - pad, ahi.z + pad, blo.z - pad, bhi.z + pad);
                int_lo[i] = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
                int_hi[i] = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
                const int p = parent[i];
                if (p < 0) break;
                float4 p3 = E.nodes[4 * (size_t)p + 3];
                const int need = (f2i(p3.x) >= 0 ? 1 : 0) + (f2i(p3.y) >= 0 ? 1 : 0);
                if (++arrived[p] < need) break;
                i = p;
            }
        }
        S.prims = E.prims.data(); S.pinfo = E.pinfo.data(); S.nodes = E.nodes.data(); S.root = E.root;
        return;
    }
    if (builder == 1) {                                            // host binned-SAH path of rt_device_commit
        std::vector<RtSahBox> pb(n);
        for (int i = 0; i < n; ++i) { f3 lo, hi; rt_prim_box(R, i, lo, hi); pb[i] = RtSahBox{{lo.x, lo.y, lo.z}, {hi.x, hi.y, hi.z}}; }
        RtSahBuilder B;
        std::vector<RtSahBox> rb; std::vector<int32_t> rk;
        int ns = n;
        if (E.presplit > 0.0) {                                     // reference pre-splitting, as rt_device_commit does it
            auto tri = [&](int32_t key) -> const float* {
                if (key < n_mt) return H.mesh_v.data() + 9 * (size_t)key;
                if (key >= n_mt + n_sp) return H.loose_v.data() + 9 * (size_t)(key - n_mt - n_sp);
                return nullptr;
            };
            RtPresplit::run(pb.data(), n, tri, E.presplit, E.presplit_depth, 3, rb, rk);
            ns = (int)rb.size();
        }
        int root = B.run(rk.empty() ? pb.data() : rb.data(), ns, leaf_max, pad, 4);
        if (!rk.empty()) for (int s = 0; s < ns; ++s) B.order[s] = rk[B.order[s]];
        E.prims.assign((size_t)3 * ns, float4{}); E.pinfo.assign(ns, float4{});
        std::vector<float4> leaf_lo(ns), leaf_hi(ns);
        for (int s = 0; s < ns; ++s) rt_emit_slot(R, s, B.order[s], E.prims.data(), E.pinfo.data(), leaf_lo.data(), leaf_hi.data());
        E.n_nodes = (int)(B.nodes.size() / 16);
        E.nodes.assign(std::max<size_t>(4, B.nodes.size() / 4), float4{});
        if (!B.nodes.empty()) memcpy(E.nodes.data(), B.nodes.data(), B.nodes.size() * 4);
        S.prims = E.prims.data(); S.pinfo = E.pinfo.data(); S.nodes = E.nodes.data(); S.root = root;
        S.n_slots = ns;
        E.order = B.order; E.root = root;
        return;
    }
    std::vector<uint64_t> keys(n); std::vector<int32_t> vals(n);
    for (int i = 0; i < n; ++i) { f3 lo, hi; rt_prim_box(R, i, lo, hi); keys[i] = rt_morton(lo, hi, bmin, inv_ext); vals[i] = i; }
    std::stable_sort(vals.begin(), vals.end(), [&](int a, int c) { return keys[a] < keys[c]; });   // radix sort is stable
    std::vector<uint64_t> skeys(n);
    for (int i = 0; i < n; ++i) skeys[i] = keys[vals[i]];
    E.prims.assign((size_t)3 * n, float4{}); E.pinfo.assign(n, float4{});
    std::vector<float4> leaf_lo(n), leaf_hi(n);
    for (int s = 0; s < n; ++s) rt_emit_slot(R, s, vals[s], E.prims.data(), E.pinfo.data(), leaf_lo.data(), leaf_hi.data());
    S.prims = E.prims.data(); S.pinfo = E.pinfo.data();
    E.order = vals;
    if (n <= leaf_max) { S.root = ~(n - 1); E.root = S.root; E.nodes.assign(4, float4{}); S.nodes = E.nodes.data(); return; }
    const int ni = n - 1;
    std::vector<int32_t> tp((size_t)5 * ni + n, 0);
    RtTopo T{tp.data(), tp.data() + ni, tp.data() + 2 * (size_t)ni, tp.data() + 3 * (size_t)ni, tp.data() + 4 * (size_t)ni, tp.data() + 5 * (size_t)ni};
    for (int i = 0; i < ni; ++i) rt_karras_node(skeys.data(), n, i, T);
    std::vector<float4> int_lo(ni), int_hi(ni); std::vector<int32_t> flags(ni, 0);
    for (int s = 0; s < n; ++s) {                                   // k_refit, serialised
        int p = T.parent_leaf[s];
        while (p >= 0) {
            if (flags[p]++ == 0) break;
            int l = T.left[p], r = T.right[p];
            float4 alo = l < 0 ? leaf_lo[~l] : int_lo[l], ahi = l < 0 ? leaf_hi[~l] : int_hi[l];
            float4 blo = r < 0 ? leaf_lo[~r] : int_lo[r], bhi = r < 0 ? leaf_hi[~r] : int_hi[r];
            int_lo[p] = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
            int_hi[p] = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
            p = T.parent_int[p];
        }
    }
    std::vector<int32_t> new_index(ni);
    int run = 0;
    for (int i = 0; i < ni; ++i) { new_index[i] = run; run += (T.last[i] - T.first[i] + 1 > leaf_max) ? 1 : 0; }
    E.n_nodes = run;
    E.nodes.assign((size_t)4 * run, float4{});
    for (int i = 0; i < ni; ++i) rt_emit_node(i, T, new_index.data(), leaf_max, pad, leaf_lo.data(), leaf_hi.data(), int_lo.data(), int_hi.data(), E.nodes.data());
    S.root = 0; S.nodes = E.nodes.data(); E.root = 0;
}

// the product picks the INST instantiation of every kernel when the committed scene has instanced meshes (rt_render.cu)
template <bool ANY, bool FAR>
bool trav(const RtScene& S, f3 O, f3 D, f3 inv, float d2, VecStack& stk, float& bt, int& bs) {
    return S.n_inst > 0 ? rt_traverse<ANY, FAR, true>(S, O, D, inv, d2, stk, bt, bs) : rt_traverse<ANY, FAR, false>(S, O, D, inv, d2, stk, bt, bs);
}

RtView make_view(const rt_render_params* p, int rank, int world, int band_rows) {
    RtView V{};
    V.pos = mk3(p->pos[0], p->pos[1], p->pos[2]); V.fwd = mk3(p->fwd[0], p->fwd[1], p->fwd[2]);
    V.right = mk3(p->right[0], p->right[1], p->right[2]); V.up = mk3(p->up[0], p->up[1], p->up[2]);
    V.res_x = p->res_x; V.res_y = p->res_y; V.dist = p->dist; V.half_h = p->half_h; V.half_v = p->half_v;
    V.bg255 = mk3(p->background_rgb[0] * 255.f, p->background_rgb[1] * 255.f, p->background_rgb[2] * 255.f);
    V.max_bounces = p->max_reflection_bounces < 0 ? 0 : p->max_reflection_bounces;
    V.spp_grid = p->spp_grid < 1 ? 1 : p->spp_grid;
    V.rank = rank; V.world = world; V.band_rows = band_rows;
    rt_view_finish(V);
    return V;
}

// rt_abi.cu cover_camera: a camera outside the committed origin box -> remember it and recompute the sphere boxes (refit)
void cover_origin(Emu& E, const float* pos) {
    if (E.S.n_spheres <= 0 || E.S.n_prims <= 0) return;
    if (!E.host.cover_camera(pos, E.obox)) return;
    commit(E, E.opt_leaf_max, E.opt_pad_scale, E.opt_t_slack, 0, true);
}

}  // namespace

extern "C" {

void* emu_new() { return new Emu(); }
void emu_free(void* e) { delete (Emu*)e; }
int emu_scene_spheres(void* e, const float* c, const rt_material* m, int n) {
    Emu* E = (Emu*)e; E->host.spheres.insert(E->host.spheres.end(), c, c + 4 * (size_t)n); E->host.sphere_mats.insert(E->host.sphere_mats.end(), m, m + n); return 0;
}
int emu_scene_triangles(void* e, const float* v, const float* nn, const rt_material* m, int n) {
    Emu* E = (Emu*)e;
    E->host.loose_v.insert(E->host.loose_v.end(), v, v + 9 * (size_t)n); E->host.loose_n.insert(E->host.loose_n.end(), nn, nn + 3 * (size_t)n);
    E->host.loose_mats.insert(E->host.loose_mats.end(), m, m + n); return 0;
}
int emu_scene_mesh(void* e, const float* v, const float* nn, int n, const rt_material* m, const float* aabb) {
    Emu* E = (Emu*)e;
    E->host.push_mesh(-1, *m, aabb);
    E->host.mesh_v.insert(E->host.mesh_v.end(), v, v + 9 * (size_t)n); E->host.mesh_n.insert(E->host.mesh_n.end(), nn, nn + 3 * (size_t)n);
    E->host.mesh_tris += n;
    return 0;
}
// rt_scene_base_mesh / rt_scene_mesh_instance / rt_scene_update_instance as rt_abi.cu implements them (aabb must be given here)
int emu_scene_base_mesh(void* e, const float* v, int n) {
    Emu* E = (Emu*)e;
    RtBaseMesh B; B.v.assign(v, v + 9 * (size_t)n);
    if (!RtInstanceSet::set_bounds(B)) return -5;
    E->host.inst.build_blas(B, E->opt_leaf_max > 3 ? 3 : (E->opt_leaf_max < 1 ? 1 : E->opt_leaf_max), 2);      // as rt_scene_base_mesh does
    E->host.inst.bases.push_back(std::move(B)); E->host.inst.blas_changed = true;
    return (int)E->host.inst.bases.size() - 1;
}
int emu_scene_mesh_instance(void* e, int base, const float* m12, const rt_material* m, const float* aabb) {
    Emu* E = (Emu*)e; RtHostScene& H = E->host;
    if (base < 0 || base >= (int)H.inst.bases.size()) return -1;
    RtInstance I; I.base = base; I.mesh = H.n_meshes();
    if (!RtInstanceSet::set_matrix(I, m12, H.inst.bases[(size_t)base])) return -5;
    H.push_mesh((int32_t)H.inst.insts.size(), *m, aabb);
    H.inst.insts.push_back(I);
    H.inst_tris += H.inst.bases[(size_t)base].n_tris();
    return 0;
}
int emu_scene_update_instance(void* e, int mesh_index, const float* m12, const float* aabb) {
    Emu* E = (Emu*)e; RtHostScene& H = E->host;
    if (mesh_index < 0 || mesh_index >= H.n_meshes() || H.mesh_inst[(size_t)mesh_index] < 0) return -1;
    RtInstance I = H.inst.insts[(size_t)H.mesh_inst[(size_t)mesh_index]];
    if (!RtInstanceSet::set_matrix(I, m12, H.inst.bases[(size_t)I.base])) return -5;
    H.inst.insts[(size_t)H.mesh_inst[(size_t)mesh_index]] = I;
    for (int k = 0; k < 6; ++k) H.mesh_aabb[6 * (size_t)mesh_index + k] = aabb[k];
    return 0;
}
// rt_scene_recommit after rt_scene_update_instance: same arrays, instance tree rebuilt
int emu_scene_recommit(void* e, const float* aabb) {
    Emu* E = (Emu*)e;
    if (aabb) for (int k = 0; k < 6; ++k) E->host.aabb[k] = aabb[k]; else E->host.calc_aabb();
    commit(*E, E->opt_leaf_max, E->opt_pad_scale, E->opt_t_slack, E->opt_builder);
    return 0;
}
int emu_scene_lights(void* e, const float* pi, int n, const float* amb) {
    Emu* E = (Emu*)e; E->host.lights.assign(pi, pi + 6 * (size_t)n);
    for (int k = 0; k < 3; ++k) E->host.ambient[k] = amb ? amb[k] : 0.f; return 0;
}
int emu_scene_commit(void* e, const float* aabb, int leaf_max, double pad_scale, double t_slack, int builder) {
    Emu* E = (Emu*)e;
    if (aabb) for (int k = 0; k < 6; ++k) E->host.aabb[k] = aabb[k]; else E->host.calc_aabb();
    E->opt_leaf_max = leaf_max; E->opt_pad_scale = pad_scale; E->opt_t_slack = t_slack; E->opt_builder = builder;
    commit(*E, leaf_max, pad_scale, t_slack, builder);
    return 0;
}
// rt_scene_update_* + rt_scene_refit: the scene is fed again (same counts) after emu_scene_clear; the tree and slot order are kept
void emu_scene_clear(void* e) { Emu* E = (Emu*)e; E->host.clear(); E->host.mesh_tris = 0; }
int emu_scene_refit(void* e, const float* aabb, int leaf_max, double pad_scale, double t_slack) {
    Emu* E = (Emu*)e;
    if ((int64_t)E->order.size() < E->host.n_prims() || E->host.n_prims() <= 0) return -1;
    if (aabb) for (int k = 0; k < 6; ++k) E->host.aabb[k] = aabb[k]; else E->host.calc_aabb();
    commit(*E, leaf_max, pad_scale, t_slack, 0, true);
    return 0;
}
void emu_set_presplit(void* e, double ratio, int depth) { ((Emu*)e)->presplit = ratio; ((Emu*)e)->presplit_depth = depth; }
int emu_n_slots(void* e) { return (int)((Emu*)e)->order.size(); }
void emu_get_aabb(void* e, float* out) { for (int k = 0; k < 6; ++k) out[k] = ((Emu*)e)->host.aabb[k]; }
int emu_n_nodes(void* e) { return ((Emu*)e)->n_nodes; }
int emu_max_stack(void* e) { return ((Emu*)e)->max_stack; }

// rt_ray_tri exactly as the kernels call it (e1/e2 precomputed like rt_emit_slot does); tmax = the caller's cull distance
int emu_ray_triangle(const float* o, const float* d, const float* v, float tmax, float* t) {
    f3 v0 = mk3(v[0], v[1], v[2]);
    f3 e1 = mk3(v[3] - v[0], v[4] - v[1], v[5] - v[2]), e2 = mk3(v[6] - v[0], v[7] - v[1], v[8] - v[2]);
    float tt = 0.f;
    bool hit = rt_ray_tri(mk3(o[0], o[1], o[2]), mk3(d[0], d[1], d[2]), v0, e1, e2, tmax, tt);
    *t = tt;
    return hit ? 1 : 0;
}

int emu_intersect(void* e, const float* o, const float* d, int n, int32_t* out_id, float* out_t) {
    Emu* E = (Emu*)e; const RtScene& S = E->S;
    VecStack stk;
    if (S.n_spheres > 0 && n > 0) {                                 // rt_intersect: the explicit origins join the origin box
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) { const float v = o[3 * i + k]; if (std::isfinite(v)) { lo[k] = fminf(lo[k], v); hi[k] = fmaxf(hi[k], v); } }
        for (int k = 0; k < 3; ++k) if (!(lo[k] <= hi[k])) { lo[k] = 0.f; hi[k] = 0.f; }
        cover_origin(*E, lo); cover_origin(*E, hi);
    }
    for (int i = 0; i < n; ++i) {
        f3 O = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), D = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        float bt = FLT_MAX; int bs = -1;
        { f3 inv = rt_rcp3(D);
          if (rt_scene_gate(S, O, inv)) {                              // k_intersect
              const float dd = um_dot(D, D);
              if (S.n_spheres > 0 && !(fabsf(dd - 1.0f) <= 0x1p-20f)) { if (S.n_inst > 0) rt_scan_all<true>(S, O, D, inv, bt, bs); else rt_scan_all<false>(S, O, D, inv, bt, bs); }
              else { stk.sp = 0; trav<false, true>(S, O, D, inv, 0.f, stk, bt, bs); }
          } }
        int32_t ty, ix, me; if (S.n_inst > 0) rt_slot_to_id<true>(S, bs, ty, ix, me); else rt_slot_to_id<false>(S, bs, ty, ix, me);
        out_id[3 * i] = ty; out_id[3 * i + 1] = ix; out_id[3 * i + 2] = me; out_t[i] = bt;
    }
    if (stk.max_sp > E->max_stack) E->max_stack = stk.max_sp;
    return 0;
}

// Mirrors rt_device_render (rt_render.cu): waves of tiles, depth loop with queues, back-to-front fold, finalize.
int emu_render(void* e, const rt_render_params* p, int rank, int world, int band_rows, int chunk_pixels,
               float* rgba32f, uint8_t* rgba8, int32_t* hit_id, float* hit_t, uint64_t* stats3) {
    Emu* E = (Emu*)e; const RtScene& S = E->S;
    cover_origin(*E, p->pos);
    RtView V = make_view(p, rank, world, band_rows);
    rt_view_far_origin(V, S.origin_reach);
    int maxd = (E->has_mirror && V.max_bounces > 0) ? (V.max_bounces > 64 ? 64 : V.max_bounces) : 0;
    V.max_bounces = maxd;
    uint64_t n_shadow = 0, n_mirror = 0;
    if (stats3) stats3[0] = stats3[1] = stats3[2] = 0;
    if (V.n_tiles <= 0 || V.res_x <= 0) return 0;
    const int g = V.spp_grid, n_pass = g * g;
    int wave_tiles = std::max(1, chunk_pixels / 32); if (wave_tiles > V.n_tiles) wave_tiles = V.n_tiles;
    const size_t W = (size_t)wave_tiles * 32;
    std::vector<float> hit0_t(W), hitq_t(W); std::vector<int32_t> hit0_slot(W), hitq_slot(W);
    std::vector<std::vector<float4>> local(maxd + 1, std::vector<float4>(W)), link(maxd + 2, std::vector<float4>(W));
    std::vector<float4> rays[2] = {std::vector<float4>(2 * W), std::vector<float4>(2 * W)};
    std::vector<float4> accum((size_t)V.n_tiles * 32);
    std::vector<uint32_t> occl(W * (size_t)std::max(1, (S.n_lights + 31) / 32));
    RtOut out{(float4*)rgba32f, (uint32_t*)rgba8, hit_id, hit_t};
    VecStack stk;
    for (int sj = 0, pass = 0; sj < g; ++sj) for (int si = 0; si < g; ++si, ++pass)
        for (int tb = 0; tb < V.n_tiles; tb += wave_tiles) {
            const int nt = std::min(wave_tiles, V.n_tiles - tb);
            std::vector<unsigned> count(maxd + 2, 0);
            for (int t = 0; t < nt; ++t) for (int lane = 0; lane < 32; ++lane) {           // k_trace_primary
                int x, y; bool valid = rt_item_to_pixel(V, tb + t, lane, x, y);
                float bt = FLT_MAX; int bs = -1;
                if (valid) { f3 O, D; rt_primary_ray(V, x, y, si, sj, O, D); { f3 inv = rt_rcp3(D); if (rt_scene_gate(S, O, inv)) { stk.sp = 0;
                    if (V.far_origin) trav<false, true>(S, O, D, inv, 0.f, stk, bt, bs); else trav<false, false>(S, O, D, inv, 0.f, stk, bt, bs); } } }
                size_t q = (size_t)t * 32 + lane; hit0_t[q] = bt; hit0_slot[q] = valid ? bs : -2;
            }
            for (int d = 0; d <= maxd; ++d) {
                const unsigned n = d == 0 ? (unsigned)nt * 32u : count[d];
                const std::vector<float4>& rin = rays[(d + 1) & 1]; std::vector<float4>& rout = rays[d & 1];
                std::vector<float>& ht = d == 0 ? hit0_t : hitq_t; std::vector<int32_t>& hs = d == 0 ? hit0_slot : hitq_slot;
                if (d > 0) for (unsigned q = 0; q < n; ++q) {                                  // k_trace_queue
                    f3 O = mk3(rin[2 * q].x, rin[2 * q].y, rin[2 * q].z), D = mk3(rin[2 * q + 1].x, rin[2 * q + 1].y, rin[2 * q + 1].z);
                    float bt = FLT_MAX; int bs = -1;
                    { f3 inv = rt_rcp3(D); if (rt_scene_gate(S, O, inv)) { stk.sp = 0; trav<false, false>(S, O, D, inv, 0.f, stk, bt, bs); } }
                    ht[q] = bt; hs[q] = bs;
                }
                const int n_chunks = (S.n_lights + 31) / 32;
                auto ray_of = [&](unsigned q, f3& O, f3& D) {
                    if (d == 0) { int x, y; rt_item_to_pixel(V, tb + (int)(q >> 5), (int)(q & 31), x, y); rt_primary_ray(V, x, y, si, sj, O, D); }
                    else { O = mk3(rin[2 * q].x, rin[2 * q].y, rin[2 * q].z); D = mk3(rin[2 * q + 1].x, rin[2 * q + 1].y, rin[2 * q + 1].z); }
                };
                for (int c = 0; c < n_chunks; ++c)                                              // k_trace_shadow (per-lane form)
                    for (unsigned q = 0; q < n; ++q) {
                        uint32_t mask = 0;
                        if (hs[q] >= 0) {
                            f3 O, D; ray_of(q, O, D);
                            uint32_t ns = 0;
                            mask = S.n_inst > 0 ? rt_shadow_mask<true>(S, O, D, ht[q], hs[q], c * 32, std::min(32, S.n_lights - c * 32), stk, ns)
                                                : rt_shadow_mask<false>(S, O, D, ht[q], hs[q], c * 32, std::min(32, S.n_lights - c * 32), stk, ns);
                            n_shadow += ns;
                        }
                        occl[(size_t)c * W + q] = mask;
                    }
                for (unsigned q = 0; q < n; ++q) {                                              // k_shade
                    RtMirror mir; mir.emit = false;
                    int slot = hs[q];
                    if (slot == -1) local[d][q] = make_float4(V.bg255.x, V.bg255.y, V.bg255.z, 0.f);
                    else if (slot >= 0) {
                        f3 O, D; ray_of(q, O, D);
                        f3 c = S.n_inst > 0 ? rt_shade_masked<true>(S, O, D, ht[q], slot, d, V.max_bounces, occl.data(), W, q, mir)
                                            : rt_shade_masked<false>(S, O, D, ht[q], slot, d, V.max_bounces, occl.data(), W, q, mir);
                        local[d][q] = make_float4(c.x, c.y, c.z, 0.f);
                    }
                    if (mir.emit) {
                        size_t o = count[d + 1]++; ++n_mirror;
                        rout[2 * o] = make_float4(mir.O.x, mir.O.y, mir.O.z, 0.f); rout[2 * o + 1] = make_float4(mir.D.x, mir.D.y, mir.D.z, 0.f);
                        link[d + 1][o] = make_float4(mir.km.x, mir.km.y, mir.km.z, i2f((int)q));
                    }
                }
            }
            for (int d = maxd; d >= 1; --d) for (unsigned q = 0; q < count[d]; ++q) rt_fold_item(q, link[d].data(), local[d].data(), local[d - 1].data());
            for (unsigned q = 0; q < (unsigned)nt * 32u; ++q) {
                if (S.n_inst > 0) rt_finalize_item<true>(S, V, tb, q, pass, n_pass, local[0].data(), hit0_t.data(), hit0_slot.data(), accum.data(), out);
                else rt_finalize_item<false>(S, V, tb, q, pass, n_pass, local[0].data(), hit0_t.data(), hit0_slot.data(), accum.data(), out);
            }
        }
    if (stk.max_sp > E->max_stack) E->max_stack = stk.max_sp;
    if (stats3) { stats3[0] = (uint64_t)V.res_x * V.n_local_rows * n_pass; stats3[1] = n_shadow; stats3[2] = n_mirror; }
    return 0;
}

}  // extern "C"

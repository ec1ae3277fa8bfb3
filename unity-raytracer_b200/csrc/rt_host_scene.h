// rt_host_scene.h — host-side accumulation of the caller's scene arrays (rt_scene_* calls) into the flat, scan-ordered
// staging arrays that are uploaded verbatim.  Mirrors what UpdateScene() leaves in `Scene`
// (RayTracingSetup.cs:120-128; Scene.cs:8-15) and Scene.CalculateAABB (Scene.cs:17-41).
#pragma once
#include <vector>
#include <cstring>
#include <cstdint>
#include <cmath>
#include <cfloat>
#include "rt_hd.h"
#include "rt_instances.h"
#include "../../include/rtb200.h"

struct RtHostScene {
    // scan order: mesh triangles (mesh order, triangle order), spheres, loose triangles  (Scene.cs:64-115)
    std::vector<float> mesh_v, mesh_n;          // 9 / 3 floats per mesh triangle
    std::vector<int32_t> mesh_base;             // first scan key of each mesh (+ total at the end after finish())
    // Instanced meshes (rt_scene_mesh_instance) are meshes of the scan like any other — MeshIndex, scan keys, material, AABB —
    // but own no range of mesh_v / mesh_n: mesh_raw[m] = first triangle of mesh m in the RAW arrays (== mesh_base[m] as long as
    // no instanced mesh precedes it), mesh_inst[m] = its index in inst.insts or -1, inst_tris = scan keys taken by instances.
    std::vector<int32_t> mesh_raw, mesh_inst;
    int64_t inst_tris = 0;
    RtInstanceSet inst;
    std::vector<float> mesh_aabb;               // 6 floats per mesh, verbatim (Mesh.cs:12)
    std::vector<rt_material> mesh_mats;
    std::vector<float> spheres;                 // 4 floats each
    std::vector<rt_material> sphere_mats;
    std::vector<float> loose_v, loose_n;        // 9 / 3 floats per loose triangle
    std::vector<rt_material> loose_mats;
    std::vector<float> lights;                  // 6 floats each
    float ambient[3] = {0, 0, 0};
    float aabb[6] = {0, 0, 0, 0, 0, 0};
    bool has_mirror = false;
    // Streaming commit (single device, device-built BVH): rt_scene_mesh copies the caller's arrays straight into the device's
    // raw buffers and mesh_v / mesh_n stay empty — nothing on the host needs the mesh vertices (the mesh AABB is the caller's).
    bool mesh_on_device = false;
    // rt_scene_update_mesh sends new vertices to the devices only; mesh_v / mesh_n (read by the host SAH builder alone) are
    // then out of date and are fetched back from device 0 if a host rebuild ever needs them (rt_abi.cu commit_impl)
    bool mesh_stale = false;
    int64_t mesh_tris = 0;                      // mesh triangles so far (== mesh_v.size() / 9 when staged on the host)
    int builder = 0;                            // RtOptions::builder latched at rt_scene_begin
    // camera positions the context has rendered from (viewer state, survives clear()): part of the origin box the sphere
    // boxes are inflated for (rt_build_steps.h rt_sphere_box_radius)
    bool cam_valid = false;
    float cam_box[6] = {0, 0, 0, 0, 0, 0};

    // Origin box of a commit (RtRaw::obox_*): the un-inflated primitive bounds b[6] — shadow and mirror rays start on surfaces,
    // RayTracingSetup.cs:331,370 — with a small margin, united with the camera positions the context has rendered from.
    void origin_box(const float b[6], float amax, float out[6]) const {
        const float m = 1e-3f * (1.0f + amax), L = 1e18f;
        for (int k = 0; k < 3; ++k) {
            float lo = b[k] - m, hi = b[3 + k] + m;
            if (!(lo <= hi)) { lo = 0.f; hi = 0.f; }             // empty / NaN bounds
            if (cam_valid) { lo = fminf(lo, cam_box[k]); hi = fmaxf(hi, cam_box[3 + k]); }
            out[k] = fminf(fmaxf(lo, -L), L); out[3 + k] = fminf(fmaxf(hi, -L), L);
        }
    }
    // A camera at `pos` outside the committed origin box `obox`: remember it, with room to move (an orbiting camera settles
    // after a few frames).  Returns true when the sphere boxes have to be recomputed (refit) for the grown box.
    bool cover_camera(const float pos[3], const float obox[6]) {
        bool inside = true, finite = true;
        for (int k = 0; k < 3; ++k) { finite = finite && std::isfinite(pos[k]); inside = inside && pos[k] >= obox[k] && pos[k] <= obox[3 + k]; }
        if (inside || !finite) return false;
        for (int k = 0; k < 3; ++k) {
            float lo = cam_valid ? fminf(cam_box[k], pos[k]) : pos[k], hi = cam_valid ? fmaxf(cam_box[3 + k], pos[k]) : pos[k];
            const float room = 0.25f * fmaxf(fmaxf(hi, obox[3 + k]) - fminf(lo, obox[k]), 0.f);
            if (pos[k] < obox[k]) lo = pos[k] - room;
            if (pos[k] > obox[3 + k]) hi = pos[k] + room;
            cam_box[k] = lo; cam_box[3 + k] = hi;
        }
        cam_valid = true;
        return true;
    }

    // keeps the capacity of the staging vectors: a scene that is re-committed every frame (RayTracingSetup.cs:120-128) must
    // not pay for fresh pages each time (48 MB of first-touch page faults cost more than the copy itself)
    void clear() {
        mesh_raw.clear(); mesh_inst.clear(); inst_tris = 0; inst.clear();
        mesh_v.clear(); mesh_n.clear(); mesh_base.clear(); mesh_aabb.clear(); mesh_mats.clear(); spheres.clear(); sphere_mats.clear();
        loose_v.clear(); loose_n.clear(); loose_mats.clear(); lights.clear();
        for (int k = 0; k < 3; ++k) ambient[k] = 0.f;
        for (int k = 0; k < 6; ++k) aabb[k] = 0.f;
        has_mirror = false; mesh_on_device = false; mesh_stale = false; mesh_tris = 0; builder = 0;
    }
    int32_t n_meshes() const { return (int32_t)mesh_mats.size(); }
    int32_t n_meshtris() const { return (int32_t)mesh_tris; }                  // triangles in the raw arrays (baked meshes)
    int32_t n_keytris() const { return (int32_t)(mesh_tris + inst_tris); }      // scan keys taken by mesh triangles, instanced ones included
    // one more mesh of the scan (rt_scene_mesh: inst_index -1; rt_scene_mesh_instance: its index in inst.insts)
    void push_mesh(int32_t inst_index, const rt_material& mat, const float aabb6[6]) {
        mesh_base.push_back(n_keytris()); mesh_raw.push_back(n_meshtris()); mesh_inst.push_back(inst_index);
        mesh_mats.push_back(mat); mesh_aabb.insert(mesh_aabb.end(), aabb6, aabb6 + 6);
    }
    int32_t n_spheres() const { return (int32_t)(spheres.size() / 4); }
    int32_t n_loose() const { return (int32_t)(loose_v.size() / 9); }
    int64_t n_prims() const { return (int64_t)n_meshtris() + n_spheres() + n_loose(); }
    int32_t n_lights() const { return (int32_t)(lights.size() / 6); }

    // Scene.CalculateAABB, Scene.cs:17-41 (AABB.Encapsulate, AABB.cs:10-20; Sphere.AABB, Sphere.cs:17-22)
    void calc_aabb() {
        float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};   // float.MinValue == -FLT_MAX
        for (int32_t m = 0; m < n_meshes(); ++m)
            for (int k = 0; k < 3; ++k) { mn[k] = um_min(mn[k], mesh_aabb[6 * m + k]); mx[k] = um_max(mx[k], mesh_aabb[6 * m + 3 + k]); }
        for (size_t t = 0; t < loose_v.size() / 9; ++t)
            for (int v = 0; v < 3; ++v)
                for (int k = 0; k < 3; ++k) { float p = loose_v[9 * t + 3 * v + k]; mn[k] = um_min(p, mn[k]); mx[k] = um_max(p, mx[k]); }
        for (size_t s = 0; s < spheres.size() / 4; ++s) {
            float r = sqrtf(spheres[4 * s + 3]);
            for (int k = 0; k < 3; ++k) { mn[k] = um_min(mn[k], spheres[4 * s + k] - r); mx[k] = um_max(mx[k], spheres[4 * s + k] + r); }
        }
        for (int k = 0; k < 3; ++k) { aabb[k] = mn[k]; aabb[3 + k] = mx[k]; }
    }

    // preconditions of rt_shadow_irrelevant (rt_core.h), over every material and light of the scene: kd finite (kd * 0 is a
    // zero), intensities finite and below 1e18 (E = I / d2 finite for d2 > 1e-20), no -0.0 channel in ambient * ka (adding the
    // +0.0 contribution of an unoccluded back-facing light would flip its sign)
    bool backcull_ok() const {
        auto neg0 = [](float x) { return x == 0.0f && std::signbit(x); };
        auto mat_ok = [&](const rt_material& m) {
            for (int k = 0; k < 3; ++k) {
                if (!std::isfinite(m.diffuse[k])) return false;
                if (neg0(ambient[k] * m.ambient[k])) return false;
            }
            return true;
        };
        for (auto& m : mesh_mats) if (!mat_ok(m)) return false;
        for (auto& m : sphere_mats) if (!mat_ok(m)) return false;
        for (auto& m : loose_mats) if (!mat_ok(m)) return false;
        for (size_t l = 0; l < lights.size() / 6; ++l)
            for (int k = 3; k < 6; ++k) if (!std::isfinite(lights[6 * l + k]) || !(std::fabs(lights[6 * l + k]) < 1e18f)) return false;
        return true;
    }

    // all materials in material-id order: meshes, spheres, loose triangles; 4 x float4 each (layout: RtScene::mats)
    std::vector<float> pack_materials() {
        std::vector<float> out;
        has_mirror = false;
        auto put = [&](const rt_material& m) {
            float rec[16] = {m.diffuse[0], m.diffuse[1], m.diffuse[2], m.phong,
                             m.ambient[0], m.ambient[1], m.ambient[2], 0.f,
                             m.mirror[0], m.mirror[1], m.mirror[2], 0.f,
                             m.specular[0], m.specular[1], m.specular[2], 0.f};
            int32_t im = m.is_mirror ? 1 : 0;
            std::memcpy(&rec[7], &im, 4);
            if (im) has_mirror = true;
            out.insert(out.end(), rec, rec + 16);
        };
        for (auto& m : mesh_mats) put(m);
        for (auto& m : sphere_mats) put(m);
        for (auto& m : loose_mats) put(m);
        return out;
    }
};

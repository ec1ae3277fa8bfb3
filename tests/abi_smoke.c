/* abi_smoke.c — the boundary seen from plain C (C99): what a [DllImport] caller binds is this C view of include/rtb200.h,
 * not the C++ one (call site it replaces: Assets/RayTracer/Demo-RayTracing/RayTracingSetup.cs:198).
 *
 *   gcc -std=c99 -Wall -Wextra -Werror -I include tests/abi_smoke.c -o abi_smoke -L unity-raytracer_b200 -lrtb200 -lm
 *   ./abi_smoke            with a GPU   : create -> scene -> commit -> rt_render / rt_intersect -> known answers -> destroy
 *   ./abi_smoke --no-gpu   without one  : struct layouts + rt_create must FAIL with RT_E_CUDA (there is no CPU fallback)
 *
 * Known answers (derived from the reference's formulas, not from the library):
 *   KAT-1  Tests/RayTracerTests.cs:8-27        ray (0,0,0)->(1,0,0) vs sphere C=(300,0,0) r2=1        t = 299
 *   KAT-2  RayTracing.unity sphere + light     ray (0,0,0)->(0,0,1): Sphere 0, t = 19.6, Value = (244.60649, 15, 15)
 *   KAT-3  RayTracingSetup.cs:288-298          pixel (25,25) of the 50x50 plane: Direction = (0.03996002, -0.01998001, 0.99900156)
 *   KAT-4  RMath.cs:29-73                      ray (17.1,0,0)->(0,0,1) vs the "Triangle" instance             t = 15
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "rtb200.h"

/* C99 has no static_assert: a negative array size stops the compiler */
typedef char assert_params_88[(sizeof(rt_render_params) == 88) ? 1 : -1];
typedef char assert_material_56[(sizeof(rt_material) == 56) ? 1 : -1];
typedef char assert_stats_144[(sizeof(rt_stats) == 144) ? 1 : -1];

static int failures = 0;
#define CHECK(cond, ...) do { if (!(cond)) { ++failures; fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); } } while (0)
#define CALL(ctx, expr) do { int rc_ = (expr); if (rc_ != RT_OK) { fprintf(stderr, "FAIL %s:%d: %s -> %d (%s)\n", __FILE__, __LINE__, #expr, rc_, rt_last_error(ctx)); return 1; } } while (0)

static rt_material material(float dr, float dg, float db, float amb) {
    rt_material m;
    memset(&m, 0, sizeof m);
    m.diffuse[0] = dr; m.diffuse[1] = dg; m.diffuse[2] = db;
    m.ambient[0] = m.ambient[1] = m.ambient[2] = amb;
    return m;
}

static rt_render_params plane(int rx, int ry) {          /* RayTracing.unity:346-354, camera at the origin, identity rotation */
    rt_render_params p;
    memset(&p, 0, sizeof p);
    p.fwd[2] = 1.f; p.right[0] = 1.f; p.up[1] = 1.f;
    p.res_x = rx; p.res_y = ry; p.dist = 10.f; p.half_h = 20.f; p.half_v = 10.f;
    p.max_reflection_bounces = 5; p.spp_grid = 1;
    return p;
}

static int with_gpu(void) {
    rt_ctx* ctx = NULL;
    int dev = 0;
    CALL(NULL, rt_create(&ctx, &dev, 1));
    CHECK(rt_abi_version() == RTB200_ABI_VERSION, "abi version %d", rt_abi_version());

    /* ---- KAT-1 ---- */
    {
        const float sph[4] = {300.f, 0.f, 0.f, 1.f};
        rt_material m = material(1.f, 1.f, 1.f, 1.f);
        const float o[3] = {0.f, 0.f, 0.f}, d[3] = {1.f, 0.f, 0.f};
        int32_t id[3]; float t = 0.f;
        CALL(ctx, rt_scene_begin(ctx));
        CALL(ctx, rt_scene_spheres(ctx, sph, &m, 1));
        CALL(ctx, rt_scene_lights(ctx, NULL, 0, NULL));
        CALL(ctx, rt_scene_commit(ctx, NULL));
        CALL(ctx, rt_intersect(ctx, o, d, 1, id, &t));
        CHECK(id[0] == RT_TYPE_SPHERE && id[1] == 0 && id[2] == -1 && t == 299.0f, "KAT-1: id {%d,%d,%d} t %.9g", id[0], id[1], id[2], (double)t);
    }
    /* ---- the sphere, the two loose triangles and the light of RayTracing.unity (SURVEY.md §8(c) fixture table) ---- */
    {
        const float sph[4] = {0.f, 0.f, 29.6f, 100.f};
        rt_material sm = material(1.f, 0.f, 0.f, 1.f);
        const float tri[18] = {17.1f, 10.f, 15.f, 7.1f, -10.f, 15.f, 27.1f, -10.f, 15.f,
                               14.16f, 10.f, 21.45f, 4.16f, -10.f, 21.45f, 24.16f, -10.f, 21.45f};
        const float nrm[6] = {0.f, 0.f, -1.f, 0.f, 0.f, -1.f};
        rt_material tm[2];
        const float light[6] = {5.79f, 0.f, 0.f, 1e5f, 1e5f, 1e5f}, ambient[3] = {15.f, 15.f, 15.f};
        float aabb[6];
        tm[0] = material(0.f, 1.f, 0.f, 1.f); tm[1] = material(1.f, 0.f, 1.f, 1.f);
        CALL(ctx, rt_scene_begin(ctx));
        CALL(ctx, rt_scene_spheres(ctx, sph, &sm, 1));
        CALL(ctx, rt_scene_triangles(ctx, tri, nrm, tm, 2));
        CALL(ctx, rt_scene_lights(ctx, light, 1, ambient));
        CALL(ctx, rt_scene_commit(ctx, NULL));
        CALL(ctx, rt_scene_get_aabb(ctx, aabb));
        CHECK(aabb[0] == -10.f && fabsf(aabb[5] - 39.6f) < 1e-5f && aabb[3] == 27.1f, "Scene.CalculateAABB: x min %.9g z max %.9g", (double)aabb[0], (double)aabb[5]);
        /* KAT-4 */
        {
            const float o[3] = {17.1f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f};
            int32_t id[3]; float t = 0.f;
            CALL(ctx, rt_intersect(ctx, o, d, 1, id, &t));
            CHECK(id[0] == RT_TYPE_TRIANGLE && id[1] == 0 && t == 15.0f, "KAT-4: id {%d,%d,%d} t %.9g", id[0], id[1], id[2], (double)t);
        }
        /* KAT-2 through rt_render: a 1x1 plane puts its only pixel centre on the view axis */
        {
            rt_render_params p = plane(1, 1);
            float rgba[4], t = 0.f; uint8_t q[4]; int32_t id[3]; rt_stats st;
            CALL(ctx, rt_render(ctx, &p, rgba, q, id, &t, &st));
            CHECK(id[0] == RT_TYPE_SPHERE && id[1] == 0 && id[2] == -1, "KAT-2: id {%d,%d,%d}", id[0], id[1], id[2]);
            CHECK(t == 19.6f, "KAT-2: t %.9g", (double)t);
            CHECK(fabsf(rgba[0] * 255.f - 244.60649f) < 2e-4f && rgba[1] * 255.f == 15.f && rgba[2] * 255.f == 15.f && rgba[3] == 1.f,
                  "KAT-2: Value (%.9g, %.9g, %.9g)", (double)(rgba[0] * 255.f), (double)(rgba[1] * 255.f), (double)(rgba[2] * 255.f));
            CHECK(q[0] == 245 && q[1] == 15 && q[2] == 15 && q[3] == 255, "KAT-2: u8 (%d,%d,%d,%d)", q[0], q[1], q[2], q[3]);
            CHECK(st.rays_primary == 1 && st.rays_shadow == 1 && st.rays_mirror == 0, "KAT-2: ray counts");
        }
        /* KAT-3 + frame indexing: pixel (25,25) of the 50x50 frame equals Scene.IntersectRay of the hand-derived direction */
        {
            rt_render_params p = plane(50, 50);
            const size_t n = 2500;
            float* rgba = (float*)malloc(n * 4 * sizeof(float));
            float* ht = (float*)malloc(n * sizeof(float));
            int32_t* hid = (int32_t*)malloc(n * 3 * sizeof(int32_t));
            const float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.03996002f, -0.01998001f, 0.99900156f};
            int32_t id[3]; float t = 0.f;
            const size_t pix = 25 + 25 * 50;                                  /* GetPixelIndex, RayTracingSetup.cs:243-246 */
            if (!rgba || !ht || !hid) return 1;
            CALL(ctx, rt_render(ctx, &p, rgba, NULL, hid, ht, NULL));
            CALL(ctx, rt_intersect(ctx, o, d, 1, id, &t));
            CHECK(hid[3 * pix] == RT_TYPE_SPHERE && hid[3 * pix] == id[0] && hid[3 * pix + 1] == id[1], "KAT-3: pixel id {%d,%d} ray id {%d,%d}", hid[3 * pix], hid[3 * pix + 1], id[0], id[1]);
            CHECK(fabsf(ht[pix] - t) <= 4e-6f * t, "KAT-3: pixel t %.9g, ray t %.9g", (double)ht[pix], (double)t);
            CHECK(hid[0] == RT_TYPE_NONE && hid[1] == -1 && ht[0] == 3.402823466e+38f && rgba[3] == 1.f, "top-left pixel must miss");
            free(rgba); free(ht); free(hid);
        }
    }
    /* ---- an instanced mesh (SceneMesh.cs:11-53 without the per-frame bake): KAT-4's triangle stored in OBJECT space around the
     * origin and placed by its localToWorldMatrix; re-posed through rt_scene_update_instances; the batch is all or nothing ---- */
    {
        const float obj[9] = {0.f, 10.f, 0.f, -10.f, -10.f, 0.f, 10.f, -10.f, 0.f};        /* KAT-4's triangle minus (17.1, 0, 15) */
        const float place[12] = {1.f, 0.f, 0.f, 17.f, 0.f, 1.f, 0.f, 0.f, 0.f, 0.f, 1.f, 15.f};
        float moved[24];
        const int32_t which[2] = {0, 0}, bad[2] = {0, 7};
        const float o[3] = {17.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f};
        rt_material m = material(0.f, 1.f, 0.f, 1.f);
        int32_t id[3]; float t = 0.f, aabb[6];
        int base;
        CALL(ctx, rt_scene_begin(ctx));
        base = rt_scene_base_mesh(ctx, obj, 1);
        CHECK(base == 0, "rt_scene_base_mesh: %d (%s)", base, rt_last_error(ctx));
        CALL(ctx, rt_scene_mesh_instance(ctx, base, place, &m, NULL));                         /* Mesh.AABB computed as SceneMesh.cs:22-31 */
        CALL(ctx, rt_scene_lights(ctx, NULL, 0, NULL));
        CALL(ctx, rt_scene_commit(ctx, NULL));
        CALL(ctx, rt_scene_get_aabb(ctx, aabb));
        CHECK(aabb[0] == 7.f && aabb[3] == 27.f && aabb[2] == 15.f && aabb[5] == 15.f, "instance AABB: x [%.9g, %.9g] z [%.9g, %.9g]", (double)aabb[0], (double)aabb[3], (double)aabb[2], (double)aabb[5]);
        CALL(ctx, rt_intersect(ctx, o, d, 1, id, &t));
        CHECK(id[0] == RT_TYPE_MESH_TRIANGLE && id[1] == 0 && id[2] == 0 && t == 15.0f, "instanced triangle: id {%d,%d,%d} t %.9g", id[0], id[1], id[2], (double)t);
        memcpy(moved, place, sizeof(place)); memcpy(moved + 12, place, sizeof(place));
        moved[11] = 99.f; moved[12 + 11] = 20.f;                                                /* the same mesh twice: its later pose stands */
        CHECK(rt_scene_update_instances(ctx, bad, moved, NULL, 2) == RT_E_INVALID && strstr(rt_last_error(ctx), "entry 1") != NULL, "bad batch: %s", rt_last_error(ctx));
        CALL(ctx, rt_intersect(ctx, o, d, 1, id, &t));                                           /* nothing moved, the scene is not even dirty */
        CHECK(t == 15.0f, "after the rejected batch: t %.9g", (double)t);
        CALL(ctx, rt_scene_update_instances(ctx, which, moved, NULL, 2));
        CHECK(rt_intersect(ctx, o, d, 1, id, &t) == RT_E_STATE, "a dirty scene must be recommitted first");
        CALL(ctx, rt_scene_recommit(ctx, NULL));
        CALL(ctx, rt_intersect(ctx, o, d, 1, id, &t));
        CHECK(id[0] == RT_TYPE_MESH_TRIANGLE && t == 20.0f, "re-posed instance: id {%d,%d,%d} t %.9g", id[0], id[1], id[2], (double)t);
    }
    /* ---- error behaviour: NULL arguments and call order never crash, 0x0 is a no-op (SampleScene.unity) ---- */
    {
        rt_render_params p = plane(0, 0);
        CHECK(rt_render(ctx, NULL, NULL, NULL, NULL, NULL, NULL) == RT_E_INVALID, "NULL params");
        CHECK(rt_render(ctx, &p, NULL, NULL, NULL, NULL, NULL) == RT_OK, "0x0 plane");
        CHECK(rt_scene_commit(ctx, NULL) == RT_E_STATE, "commit without begin");
        CHECK(rt_set_option(ctx, "no_such_option", 1.0) == RT_E_INVALID, "unknown option");
        CHECK(strlen(rt_last_error(ctx)) > 0, "rt_last_error empty");
    }
    rt_destroy(ctx);
    return 0;
}

int main(int argc, char** argv) {
    if (argc > 1 && strcmp(argv[1], "--no-gpu") == 0) {
        rt_ctx* ctx = (rt_ctx*)0x1;
        int rc = rt_create(&ctx, NULL, 1);
        CHECK(rc == RT_E_CUDA && ctx == NULL, "rt_create without a device: rc %d", rc);
        CHECK(strstr(rt_last_error(NULL), "no CPU fallback") != NULL, "message: %s", rt_last_error(NULL));
        CHECK(rt_create(NULL, NULL, 1) == RT_E_INVALID, "NULL out");
        CHECK(rt_create(&ctx, NULL, 0) == RT_E_INVALID && rt_create(&ctx, NULL, 17) == RT_E_INVALID, "device count range");
        rt_destroy(NULL);
    } else if (with_gpu() != 0) ++failures;
    if (failures) { fprintf(stderr, "abi_smoke: %d failure(s)\n", failures); return 1; }
    printf("abi_smoke ok\n");
    return 0;
}

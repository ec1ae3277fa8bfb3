/* rtb200.h — C-ABI of librtb200.so, the B200-native replacement for the reference's per-pixel hot path.
 *
 * The reference (vectorized-runner/unity-raytracer) has NO native/FFI interface of its own: the path is the private
 * managed call  CastPixelRays(CameraData)  (Assets/RayTracer/Demo-RayTracing/RayTracingSetup.cs:198, body :275-302)
 * which reads this.Scene / this.ImagePlane / BackgroundColor / MaxReflectionBounces and writes Color[] PixelColors.
 * This header therefore DEFINES the boundary; every entry point cites the reference member it replaces or mirrors.
 * All paths below are relative to Assets/RayTracer/ of the reference.
 *
 * Conventions
 *   - plain C, no exceptions, no callbacks; every function returns RT_OK (0) or a negative rt_status.
 *   - all array arguments are caller-owned HOST memory, read/written only during the call (the library copies);
 *     rt_ctx is library-owned. Counts of 0 and 0x0 resolutions are valid no-ops (Assets/Scenes/SampleScene.unity
 *     ships a 0x0 ImagePlane).
 *   - one caller thread per rt_ctx (Unity main thread); calls are blocking unless stated; contexts are independent.
 *   - there is no CPU fallback: with no CUDA device rt_create fails with RT_E_CUDA.
 */
#ifndef RTB200_H
#define RTB200_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RTB200_ABI_VERSION 2

typedef struct rt_ctx rt_ctx;

typedef enum {
    RT_OK = 0,
    RT_E_INVALID = -1,      /* bad argument */
    RT_E_CUDA = -2,         /* CUDA runtime error / no device */
    RT_E_OOM = -3,          /* device or host allocation failed */
    RT_E_STATE = -4,        /* call order violated (e.g. render before commit) */
    RT_E_UNSUPPORTED = -5
} rt_status;

/* ObjectType, Data/Objects/ObjectType.cs:3-9 */
enum { RT_TYPE_NONE = 0, RT_TYPE_SPHERE = 1, RT_TYPE_TRIANGLE = 2, RT_TYPE_MESH_TRIANGLE = 3 };

/* MaterialData, Data/Shading/MaterialData.cs:7-14. `bool IsMirror` is widened to int32 so the struct is blittable. */
typedef struct {
    float diffuse[3];       /* DiffuseReflectance  */
    float ambient[3];       /* AmbientReflectance  */
    float mirror[3];        /* MirrorReflectance   */
    float specular[3];      /* SpecularReflectance */
    float phong;            /* PhongExponent       */
    int32_t is_mirror;      /* IsMirror            */
} rt_material;              /* 56 bytes */

/* CameraData (Data/Camera/CameraData.cs:7-10) + ImagePlane (Data/Camera/ImagePlane.cs:13-21) +
 * RayTracingSetup.BackgroundColor / MaxReflectionBounces (RayTracingSetup.cs:22-23). */
typedef struct {
    float pos[3], fwd[3], right[3], up[3];   /* Position; Forward/Right/Up already normalised by the caller (RTS:173-181) */
    int32_t res_x, res_y;                    /* Resolution.X, .Y */
    float dist, half_h, half_v;              /* DistanceToCamera, HalfHorizontalLength, HalfVerticalLength */
    float background_rgb[3];                 /* 0..1; multiplied by 255 inside (Rgb.cs:15-18) */
    int32_t max_reflection_bounces;          /* RTS:23; above 64 in a scene that has mirror materials: RT_E_UNSUPPORTED */
    int32_t spp_grid;                        /* 1 = reference behaviour (one ray at the pixel centre, RTS:291-292);
                                                g>1 = g*g stratified sample centres, averaged in fixed (j,i) order */
} rt_render_params;                          /* 88 bytes */

typedef struct {
    uint64_t rays_primary, rays_shadow, rays_mirror;   /* Scene.IntersectRay-equivalent queries issued by the last render */
    float ms_render;            /* device time of the last render, ray-gen .. final pixel write (CUDA events) */
    float ms_d2h;               /* host copy-out time of the last rt_render (0 for rt_render_device) */
    float ms_commit_upload;     /* last rt_scene_commit / rt_scene_recommit: host->device (incl. the uploads inside rt_scene_mesh / rt_scene_update_*) */
    float ms_commit_build;      /* last rt_scene_commit: device BVH build */
    uint64_t h2d_bytes;         /* bytes uploaded by the last commit */
    uint64_t d2h_bytes;         /* bytes copied out by the last rt_render */
    uint32_t n_prims, n_nodes;  /* device BVH size */
    uint32_t kernel_launches;   /* kernels of this library launched by the last render */
    uint32_t reserved;
    /* per-kernel device time of the last render, summed over launches; filled only when option "kernel_timing" is 1
     * (CUDA events around every launch on the launching stream). Order: RT_K_* below. */
    float ms_kernel[8];
    uint32_t launches_kernel[8];
    uint64_t rays_shadow_skipped;   /* of rays_shadow: queries for lights behind the surface answered without traversal (option "backcull") */
} rt_stats;
enum { RT_K_TRACE_PRIMARY = 0, RT_K_SHADE_PRIMARY = 1, RT_K_TRACE_QUEUE = 2, RT_K_SHADE_QUEUE = 3, RT_K_FOLD = 4, RT_K_FINALIZE = 5,
       RT_K_SHADOW_PRIMARY = 6, RT_K_SHADOW_QUEUE = 7, RT_K_COUNT = 8 };

/* ---- lifetime ---------------------------------------------------------------------------------------------- */
/* n_devices in 1..16 (one box: typically 1, 2, 4 or 8); device_ids NULL = devices 0..n_devices-1.  n_devices > 1 is in-process
 * multi-GPU: screen-space row bands interleaved over the devices (any count works, the band owner is band % n_devices), scene
 * replicated.  rt_render_device gathers the frame on device_ids[0] (every device stores its bands straight into that device's
 * buffers over NVLink peer mappings); rt_render copies every device's bands to the caller's host arrays over that device's own
 * PCIe link.  Devices that cannot reach device_ids[0] as a peer: RT_E_UNSUPPORTED. */
int rt_create(rt_ctx** out, const int* device_ids, int n_devices);
void rt_destroy(rt_ctx* ctx);
const char* rt_last_error(const rt_ctx* ctx);        /* valid until the next call on the same ctx; ctx may be NULL */
int rt_abi_version(void);

/* ---- scene: replaces UpdateScene()'s result, RayTracingSetup.cs:120-128. index in array == ObjectId.Index ---- */
int rt_scene_begin(rt_ctx* ctx);                      /* Scene.*Data.Clear(), RTS:161,250,260 */
/* SphereData (Data/Objects/SphereData.cs:7-16): center_r2 = 4 floats each = Sphere{Center,RadiusSquared} (Sphere.cs:9-10) */
int rt_scene_spheres(rt_ctx* ctx, const float* center_r2, const rt_material* materials, int n);
/* TriangleData (Data/Objects/TriangleData.cs:8-19): v = 9 floats each (Triangle.cs:9-11), normals = 3 each */
int rt_scene_triangles(rt_ctx* ctx, const float* v, const float* normals, const rt_material* materials, int n);
/* one Mesh (Data/Objects/Mesh.cs:7-12); call order == ObjectId.MeshIndex; aabb = {Min.xyz, Max.xyz} as the caller
 * computed it (SceneMesh.cs:22-31) — it is a *semantic* cull in Scene.IntersectRay (Scene.cs:67) and is used verbatim. */
int rt_scene_mesh(rt_ctx* ctx, const float* v, const float* normals, int n_tris, const rt_material* material,
                  const float aabb[6]);
/* Instanced meshes: SceneMesh.Mesh (SceneComponents/SceneMesh.cs:11-53) WITHOUT the per-frame world bake.  The reference
 * multiplies every vertex of every mesh by transform.localToWorldMatrix on every frame (SceneMesh.cs:21-31) and hands the scan
 * baked triangles (Data/Objects/Mesh.cs:7-12).  Here the mesh stays in object space, is stored once however often it occurs, and
 * each occurrence is a 3x4 matrix: the library bakes triangles on the fly with the bake's own arithmetic —
 *   world = Matrix4x4.MultiplyPoint3x4(v): ((m00*x + m01*y) + m02*z) + m03 per row, binary32, left to right (SceneMesh.cs:28),
 *   TriangleNormals[i] = -Triangle.Normal of the world triangle (SceneMesh.cs:45, Data/Objects/Triangle.cs:13-21)
 * — so an instanced mesh renders bit-identically to the same mesh baked by the caller with that formula and sent through
 * rt_scene_mesh (ids, distances, colours; no edge exemption), while re-posing it costs 48 bytes instead of 48 bytes per triangle.
 * rt_scene_base_mesh: v = 9 floats per triangle in OBJECT space (MeshFilter.sharedMesh vertices gathered by its index buffer,
 *   SceneMesh.cs:15-17,36-41; finite values only).  Returns the base-mesh index (>= 0, call order) or a negative rt_status.  A base
 *   mesh is not part of the scan by itself.
 * rt_scene_mesh_instance: one occurrence; takes the place of ONE rt_scene_mesh call — instances and baked meshes share the call-order
 *   numbering that is ObjectId.MeshIndex, in any mix.  m3x4 = rows of localToWorldMatrix (m00 m01 m02 m03, m10 .., m20 ..); must be
 *   finite and invertible (condition number below 1e5), else RT_E_UNSUPPORTED: bake such a mesh.  aabb = Mesh.AABB as the caller
 *   computed it (the semantic cull of Scene.cs:67, used verbatim) or NULL -> computed as SceneMesh.cs:22-31 does (min / max of the
 *   baked vertices; costs a pass over the base mesh on the host).
 * rt_scene_update_instance: re-pose an instance of a committed scene (dirty-tracking update: follow with rt_scene_recommit or
 *   rt_scene_refit; both rebuild the instance tree, which takes microseconds, and leave the base meshes alone).
 * rt_scene_update_instances: the same for n instances in one call — what a frame of the reference does, which re-reads every
 *   SceneMesh's localToWorldMatrix (RTS:120-128, SceneMesh.cs:21): mesh_indices[n], m3x4 = 12 floats per entry, aabbs = 6 floats per
 *   entry or NULL (all computed).  All or nothing: when an entry is rejected (RT_E_INVALID: not an instanced mesh; RT_E_UNSUPPORTED:
 *   its matrix) no pose has changed; rt_last_error names the entry.  A mesh index that appears twice ends with its later pose. */
int rt_scene_base_mesh(rt_ctx* ctx, const float* v, int n_tris);
int rt_scene_mesh_instance(rt_ctx* ctx, int base_mesh, const float m3x4[12], const rt_material* material, const float aabb[6]);
int rt_scene_update_instance(rt_ctx* ctx, int mesh_index, const float m3x4[12], const float aabb[6]);
int rt_scene_update_instances(rt_ctx* ctx, const int32_t* mesh_indices, const float* m3x4, const float* aabbs, int n);
/* Scene.PointLights (6 floats each = PointLightData{Position,Intensity}, PointLightData.cs:9-10) + AmbientLight.Radiance;
 * ambient_radiance NULL = default(AmbientLightData) = 0 (RTS:132) */
int rt_scene_lights(rt_ctx* ctx, const float* pos_intensity, int n, const float ambient_radiance[3]);
/* scene_aabb = Scene.AABB as the caller computed it, or NULL -> recomputed per Scene.CalculateAABB (Scene.cs:17-41).
 * Uploads, replicates to every device, builds the device BVH. */
int rt_scene_commit(rt_ctx* ctx, const float scene_aabb[6]);
int rt_scene_get_aabb(rt_ctx* ctx, float out_aabb[6]);

/* Dirty-tracking updates of a committed scene.  The reference rebuilds its whole Scene every frame (UpdateScene(),
 * RayTracingSetup.cs:120-128, called from Update() at :184) although usually only a few transforms moved; these entry points let
 * the shim re-send only what changed.  Counts and materials stay as committed; the arrays replace the object's data in place
 * (same layouts as rt_scene_mesh / rt_scene_triangles / rt_scene_spheres / rt_scene_lights; mesh_index and first are the
 * reference's MeshIndex / Index).  After one or more updates rt_scene_recommit rebuilds the BVH from the device-resident
 * arrays (scene_aabb as for rt_scene_commit); rendering between an update and the recommit returns RT_E_STATE.  An update
 * rejected for its arguments (RT_E_INVALID) changes nothing and leaves a clean scene clean. */
int rt_scene_update_mesh(rt_ctx* ctx, int mesh_index, const float* vertices, const float* normals, const float aabb[6]);
int rt_scene_update_triangles(rt_ctx* ctx, int first, int n, const float* vertices, const float* normals);
int rt_scene_update_spheres(rt_ctx* ctx, int first, int n, const float* center_r2);
int rt_scene_update_lights(rt_ctx* ctx, const float* pos_intensity, int n, const float ambient_radiance[3]);
int rt_scene_recommit(rt_ctx* ctx, const float scene_aabb[6]);
/* Same contract as rt_scene_recommit, but the committed tree (either builder's) and its leaf order are KEPT and only the boxes
 * are recomputed bottom-up from the moved primitives: the cheap path for deforming / slowly moving geometry, and the way to keep
 * a host-built SAH tree on a dynamic scene (1M triangles: ~0.2 ms instead of a 0.3 s host build).  Results are identical to a
 * fresh commit by construction (the query result does not depend on the tree, only on the boxes being conservative); frame
 * time degrades as the geometry moves away from the layout the tree was built for, until the next commit / recommit. */
int rt_scene_refit(rt_ctx* ctx, const float scene_aabb[6]);

/* ---- render: replaces CastPixelRays(CameraData), RayTracingSetup.cs:198,275-302 ---------------------------- */
/* out_rgba32f : X*Y*4 floats == UnityEngine.Color[] PixelColors (RTS:40,300; Rgb.Color = Value/255, alpha 1, Rgb.cs:13)
 * out_rgba8   : X*Y*4 bytes, clamp(rint(Rgb.Value),0,255), alpha 255 (our definition: the reference never quantises)
 * out_hit_id  : X*Y*3 int32 {Type, Index, MeshIndex} of the PRIMARY hit (IntersectionResult.ObjectId, ObjectId.cs:7-9);
 *               MeshIndex is -1 unless Type == RT_TYPE_MESH_TRIANGLE; miss = {0,-1,-1}
 * out_hit_t   : X*Y floats, IntersectionResult.Distance of the primary hit, FLT_MAX on a miss (Scene.cs:45)
 * pixel index = x + y*res_x, y = 0 is the top row (RTS:243-246). Any output may be NULL.
 * Several devices (rt_create with n_devices > 1): every device copies the bands it rendered to the caller's arrays over its
 * own PCIe link, wave by wave while it still renders — a 4K Color[] frame is 2.3 ms of ONE link.  A ctx that holds a
 * partition of the frame (rt_set_partition, one process per GPU) writes ONLY its own bands: all ranks pass the same
 * full-frame arrays (shared memory) and the frame is complete when every rank has returned. */
int rt_render(rt_ctx* ctx, const rt_render_params* params, float* out_rgba32f, uint8_t* out_rgba8,
              int32_t* out_hit_id, float* out_hit_t, rt_stats* stats);

/* Same render, but the result stays in device memory (the framebuffers owned by the ctx on device_ids[0]);
 * `want` is a bit mask of RT_OUT_*. Asynchronous with respect to the host only if rt_set_stream gave a stream. */
enum { RT_OUT_RGBA32F = 1, RT_OUT_RGBA8 = 2, RT_OUT_HIT_ID = 4, RT_OUT_HIT_T = 8 };
int rt_render_device(rt_ctx* ctx, const rt_render_params* params, int want, rt_stats* stats);
/* copy the device framebuffers of the last rt_render_device out to host memory (any pointer may be NULL); RT_E_STATE after
 * an rt_render of a multi-device / partitioned ctx, which gathers nothing on the device */
int rt_read_outputs(rt_ctx* ctx, float* out_rgba32f, uint8_t* out_rgba8, int32_t* out_hit_id, float* out_hit_t);
/* raw device pointers of those framebuffers (NULL when not allocated): for zero-copy consumers / NCCL gathers */
int rt_device_outputs(rt_ctx* ctx, void** rgba32f, void** rgba8, void** hit_id, void** hit_t);

/* Scene.IntersectRay(Ray) for n explicit rays (Scene.cs:43-122): out_id 3 int32 each, out_t 1 float each. */
int rt_intersect(rt_ctx* ctx, const float* origins, const float* directions, int n, int32_t* out_id, float* out_t);

/* rt_intersect without any device allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, same results as rt_intersect, but the rays and results are staged in a file-scope device array of the library
 * (RT_INTERSECT_STATIC_MAX rays x 40 bytes = 2.5 MiB, held by every GPU that loads librtb200.so until the process exits) instead
 * of per-call cudaMalloc'ed buffers; the call makes no cudaMalloc / cudaFree / allocating library call.  n above
 * RT_INTERSECT_STATIC_MAX: RT_E_INVALID (split the batch).  Every call on a device shares that one array, which keeps the last
 * call's contents (nothing in it needs clearing: each call overwrites what it reads back): calls that overlap on the same GPU —
 * two contexts, or two host threads — race and must be serialised by the caller.  Single device (device_ids[0]).  Because it may not
 * re-fit the sphere boxes (that allocates), a ray whose origin lies outside the origin box of the last commit (scene bounds united
 * with the camera positions rendered from so far) in a scene that has spheres takes the reference's linear scan over all primitives:
 * identical result, O(primitives) time. */
#define RT_INTERSECT_STATIC_MAX 65536
int rt_intersect_static(rt_ctx* ctx, const float* origins, const float* directions, int n, int32_t* out_id, float* out_t);

int rt_get_stats(rt_ctx* ctx, rt_stats* stats);

/* ---- one-process-per-GPU plumbing (torchrun / NCCL launches) ------------------------------------------------- */
/* Use the caller's CUDA stream (cudaStream_t) for all work on device_ids[0]; NULL restores the library's own stream. */
int rt_set_stream(rt_ctx* ctx, void* cuda_stream);
/* This ctx renders only the row bands b with b % world == rank (band = band_rows consecutive rows, band_rows % 4 == 0).
 * Default rank 0 / world 1 / band_rows 32. Outputs keep full-frame indexing. */
int rt_set_partition(rt_ctx* ctx, int rank, int world, int band_rows);
/* Peer gather without a collective: rank 0 exports its framebuffers (one 64-byte cudaIpcMemHandle_t each, order
 * rgba32f, rgba8, hit_id, hit_t; zeroed when not allocated) after rt_alloc_outputs; other ranks import them and
 * from then on write their bands straight into rank 0's memory over NVLink. */
int rt_alloc_outputs(rt_ctx* ctx, int res_x, int res_y, int want);
int rt_export_outputs(rt_ctx* ctx, void* handles_4x64);
int rt_import_outputs(rt_ctx* ctx, const void* handles_4x64, int res_x, int res_y, int want);

/* Page-locks caller memory that outlives the call (cudaHostRegister) — typically the PixelColors array (allocated once per
 * resolution, RayTracingSetup.cs:40,206-210; the managed side pins it with GCHandle) or mesh arrays that are re-sent every
 * frame.  Copies from / to registered memory run at PCIe speed (a 4K Color[] frame: 2.3 ms instead of ~13 ms through the
 * driver's pageable staging) and overlap the kernels.  Purely an optimisation: unregistered memory works everywhere. */
int rt_host_register(rt_ctx* ctx, void* ptr, size_t bytes);
int rt_host_unregister(rt_ctx* ctx, void* ptr);

/* Tuning knobs (all optional; results never depend on them).  Unknown keys return RT_E_INVALID.
 *   builder        0 = device LBVH (default, 0.65 ms per million primitives once the buffers exist), 1 = host binned-SAH BVH2 (slower commit, fewer node
 *                  visits per ray); takes effect at the next rt_scene_begin.  leaf_max (1..8, default 4), pad_scale, t_slack likewise.
 *   packet         1 = warp-coherent packet traversal for coherent warps (default), 0 = always the per-lane loop
 *   fuse_shade     1 = shading (and the pixel write at depth 0) inside the shadow-trace kernel when the scene has <= 32 lights (default)
 *   backcull       1 = shadow queries of lights behind the surface are answered without traversal when that is exact for every
 *                  material and light of the scene (default; decided at commit; the queries are still counted in rays_shadow)
 *   light_split    fused shadow kernel: tiles of cost class >= this are handed out as one work item per light (0 = off, -1 = auto,
 *                  the default: class 6 on partitions of a quarter frame or less); needs hints
 *   hints          1 = heavy-first tile order from the previous frame's measured tile costs (default)
 *   sort_queue     mirror-ray queues are re-ordered by (direction bin, origin cell) before they are traced, so that the lanes of a warp
 *                  walk similar paths: 1 = always, 0 = never, -1 = auto (default: scenes of >= 65 536 primitives)
 *   waves, lanes   minimum number of waves rt_render_device cuts a pass into (default 0 = auto: 2, or 1 on a quarter-frame partition or less) and the 1..8 streams they rotate over
 *                  (default 2: the drain of one wave overlaps the start of the next)
 *   deep_lanes     lanes of a frame with mirror depths and >= 4 waves (spp passes), whose many small queue kernels then run side by side (1..8, default 8)
 *   presplit, presplit_depth   host SAH builder: > 0 = triangles whose box half-area exceeds this multiple of their area enter the tree as up to
 *                  2^depth clipped pieces (reference pre-splitting; measured slower for packet traversal, default 0 = off)
 *   sah_term       host SAH builder: > 0 = SAH-driven leaf termination (node step cost / primitive test cost) for ranges of <= 8 primitives (default 0 = split down to leaf_max)
 *   copy_waves     number of waves rt_render splits a frame into so that the copy-out of one wave overlaps the kernels of the
 *                  next (default 8, 1 = off); copy_waves_part: the same for a device that renders a band partition (0 = auto:
 *                  4 on a half frame, 2 below)
 *   chunk_pixels   upper bound of a wave in pixels (bounds the wave buffers)
 *   ctas_per_sm    cap on the resident CTAs per SM of the persistent kernels (0 = occupancy limit)
 *   prefetch, prefetch_max_mb   bulk L2 prefetch of nodes + primitives at the start of a frame when they fit (default off)
 *   stream_upload  1 = rt_scene_mesh uploads straight from the caller's arrays, no host staging (default; single device, builder 0)
 *   upload_threads host threads that fill the pinned bounce buffers of an upload from pageable memory (default 4; 1 = cudaMemcpyAsync)
 *   kernel_timing  1 = CUDA events around every launch, reported in rt_stats.ms_kernel (serialises the lanes) */
int rt_set_option(rt_ctx* ctx, const char* key, double value);

#ifdef __cplusplus
}
#endif
#endif /* RTB200_H */

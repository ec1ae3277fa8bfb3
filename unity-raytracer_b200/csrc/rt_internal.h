// rt_internal.h — host-side state of librtb200 (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include <cstdio>
#include "rt_core.h"
#include "rt_build_steps.h"
#include "rt_host_scene.h"

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t ensure(size_t n) {
        if (n <= bytes && p) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; bytes = 0; }
        if (n == 0) n = 16;
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    // like ensure, but the first `keep` bytes survive a reallocation (grown by half to amortise repeated appends)
    cudaError_t ensure_keep(size_t n, size_t keep, cudaStream_t st) {
        if (n <= bytes && p) return cudaSuccess;
        if (!p || keep == 0) return ensure(n);
        void* q = nullptr;
        const size_t cap = n + n / 2;
        cudaError_t e = cudaMalloc(&q, cap);
        if (e != cudaSuccess) return e;
        e = cudaMemcpyAsync(q, p, keep < bytes ? keep : bytes, cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { cudaFree(q); return e; }
        cudaFree(p); p = q; bytes = cap;
        return cudaSuccess;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
    template <class T> T* as() const { return (T*)p; }
};

struct RtOptions {
    int chunk_pixels = 1 << 23;        // work items (pixel samples) per wave; multiple of 32
    int leaf_max = 4;                  // LBVH subtrees with <= leaf_max primitives become one leaf (<= 8)
    double presplit = 0.0;             // host SAH builder: > 0 = triangles whose box half-area exceeds this multiple of their own area enter the
                                       // builder as clipped pieces (rt_sah.h RtPresplit); 0 = off
    int presplit_depth = 3;            // at most 2^depth pieces per triangle
    double sah_term = 0.0;             // host SAH builder: > 0 = ranges of <= 8 primitives are split by an exact sweep and only while the
                                       // SAH says so (node step cost / primitive test cost = this ratio); 0 = split down to leaf_max
    double pad_scale = 1.0 / 65536.0;  // node-box padding = max |coordinate| * pad_scale
    double t_slack = 1.000125;         // multiplicative slack on the traversal cull distance
    int block = 128;
    int ctas_per_sm = 0;               // 0 = occupancy query
    int packet = 1;                    // warp-coherent packet traversal for coherent warps
    int builder = 0;                   // 0 = device LBVH (fast commit), 1 = host binned-SAH (rt_sah.h; slower commit, fewer node visits)
    int waves = 0;                     // rt_render_device: minimum number of waves per pass; 0 = auto: 2 (two lanes: tails overlap),
                                       // 1 on partitions of a quarter frame or less (smaller kernels only lengthen the critical path)
    int lanes = 2;                     // 1 = all waves on one stream; 2.. = consecutive waves rotate over this many streams (tail overlap)
    int deep_lanes = 8;                // lanes of a frame with mirror depths AND >= 4 waves (spp passes): its queue kernels trace few,
                                       // incoherent rays each and last as long as their longest ray, so many of them run side by side
    long long copy_plan = 0;           // development: wave plan of the pipelined copy-out as decimal digits (sixteenths of the rows per wave, sum 16); 0 = 1244221
    int copy_waves = 8;                // rt_render (host outputs): a device that renders the whole frame splits it into this many waves
                                       // so that the device->host copy of wave k overlaps the kernels of wave k+1 (1 = no overlap)
    int copy_waves_part = 0;           // the same for a device that renders a band partition (several devices / ranks, each copying its
                                       // own bands out over its own PCIe link); 0 = auto: 4 waves on a half frame, 2 below
    int fuse_shade = 1;                // shading (+ pixel write at depth 0) inside the shadow-trace kernel when <= 32 lights
    int prefetch = 0;                  // bulk L2 prefetch of nodes + primitives at the start of every frame ...
    int prefetch_max_mb = 96;          // ... when they total at most this many MiB (L2 is 126 MB)
    int upload_threads = 4;            // host threads that fill the pinned bounce buffers of a pageable upload (1 = plain cudaMemcpyAsync)
    int stream_upload = 1;             // single device + builder 0: rt_scene_mesh uploads straight from the caller's arrays (no host staging)
    int backcull = 1;                  // skip shadow traversals whose answer cannot change the pixel (rt_shadow_irrelevant), when exact
    int light_split = -1;              // fused depth-0 shadow kernel: tiles of cost class >= this become one work item per light on
                                       // separate warps (needs hints).  0 = off, -1 = auto: class 6 (>= 128K cycles) on partitions of
                                       // 1/4 frame or less, off otherwise — it shortens the heaviest tiles (1/8 partition: shadow
                                       // kernel -5..11 %) but costs 4 % on a full frame (duplicated set-up, lost cache sharing)
    int hints = 1;                     // heavy-first tile order from the previous frame's measured tile costs
    int sort_queue = 0;                // mirror-ray queues are re-ordered by (direction bin, origin cell) before they are traced (rt_render.cu
                                       // k_qs_*): 1 = always, 0 = never, -1 = auto (scenes of >= 65 536 primitives, where incoherent lanes cost)
    int kernel_timing = 0;             // 1 = CUDA events around every kernel launch (rt_stats.ms_kernel)
};

enum { RT_MAX_DEPTH = 64 };           // max_reflection_bounces is clamped to this
enum { RT_STACK_ENTRIES = 116 };      // traversal stack entries per thread (rt_render.cu: 12 in shared memory + 104 in local memory);
                                      // bounds the depth of any tree the builders may emit (LBVH: 96 key bits; SAH: rt_sah.h MAX_SAH_DEPTH)

struct RtDevice {
    int device = -1;
    int sm_count = 0;
    cudaStream_t stream = nullptr;      // stream in use
    cudaStream_t own_stream = nullptr;  // created by the library
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // ---- scene ----
    DevBuf raw_v, raw_n, raw_s, nodes, prims, pinfo, mats, mesh_aabb, mesh_base, lights;
    // instanced meshes (rt_instances.h): first raw triangle per mesh, BLAS + TLAS nodes, object-space triangles, instance records
    DevBuf mesh_raw, inodes, iprims, inst, inst_vbase;
    uint64_t inst_blas_gen = 0;         // RtInstanceSet::blas_generation this device holds in iprims / the BLAS part of inodes
    RtScene S{};
    uint32_t n_nodes = 0;
    bool has_mirror = false;
    float obox[6] = {0, 0, 0, 0, 0, 0};  // origin box the committed sphere boxes were inflated for (rt_sphere_box_radius)
    // refit (rt_scene_refit): the committed tree is kept and only its boxes are recomputed
    int32_t bvh_prims = -1, bvh_slots = 0, bvh_root = 0;   // primitive count / root ref of the tree in `nodes` (+ slot order in vals_alt); -1: none
    bool bvh_parents = false;           // node_parent holds the parents of this tree
    DevBuf node_parent;
    // ---- build scratch ----
    DevBuf keys, keys_alt, vals, vals_alt, cub_tmp, bounds, topo_i32, leaf_lo, leaf_hi, int_lo, int_hi, flags, new_index;
    // ---- frame work buffers (sized to the wave capacity W), one set per lane ----
    // A frame is cut into waves; consecutive waves alternate between two lanes (stream + buffer set) so that the drain
    // of one wave's persistent kernels is filled by the next wave's kernels instead of idling the SMs.
    static constexpr int RT_MAX_LANES = 8;
    struct Lane {
        cudaStream_t own_stream = nullptr;   // lane 0 runs on D.stream, the others on their own stream
        cudaEvent_t done = nullptr;
        DevBuf hit0_t, hit0_slot, hitq_t, hitq_slot, rays[2], counters, occl, split_occl, split_acc;
        DevBuf link_tmp, qs_hist;           // option sort_queue: emission target of the links, bin counters of the queue sort
        std::vector<DevBuf> local;          // [depth] float4 per entry
        std::vector<DevBuf> link;           // [depth>=1] float4 (km.xyz, parent index bits)
    };
    Lane lanes[RT_MAX_LANES];
    DevBuf accum, stats;
    // heavy-first tile order (rt_render.cu "cost hints"): per tile-indexed trace kernel, double-buffered class lists
    struct Hints { DevBuf lists[2], counts[2]; };
    Hints hint_primary, hint_shadow;
    int hint_cur = 0; bool hint_ready = false;
    bool hint_shadow_ready = false;          // the frame that recorded the hints launched its depth-0 shadow kernel(s) (it had lights)
    static constexpr int HINT_SIG = 12;
    long long hint_sig[HINT_SIG] = {0};
    bool split_dirty = false;                // a frame was abandoned half-way: the per-light combine buffers may be non-zero
    std::vector<cudaEvent_t> fin_ev;        // spp passes: finalize of (pass p, wave i) follows finalize of (pass p-1, wave i)
    // ---- outputs: full-frame buffers. Owned on the gather device; other devices/ranks hold peer pointers ----
    DevBuf own_rgba32f, own_rgba8, own_hit_id, own_hit_t;
    void* o_rgba32f = nullptr; void* o_rgba8 = nullptr; void* o_hit_id = nullptr; void* o_hit_t = nullptr;   // targets of the NEXT render
    bool imported[4] = {false, false, false, false};
    void* imp[4] = {nullptr, nullptr, nullptr, nullptr};   // gather targets imported from another process (rt_import_outputs)
    int out_res_x = 0, out_res_y = 0;
    // partition
    int rank = 0, world = 1;
    // per-launch timing events (option kernel_timing): (kernel class, begin, end)
    std::vector<cudaEvent_t> kev_pool;
    std::vector<int> kev_class;        // class of launch i; its events are kev_pool[2i], kev_pool[2i+1]
    // pipelined copy-out (rt_render with host buffers): one event per finished wave + the pixel rows it completed
    cudaStream_t copy_stream = nullptr;
    // host->device staging ring of rt_upload (rt_build.cu): pinned bounce buffers filled by several host threads
    void* up_ring = nullptr; size_t up_slot_bytes = 0; int up_slots = 0; std::vector<cudaEvent_t> up_ev;
    std::vector<cudaEvent_t> wave_ev;
    struct WaveMark { int row0, row1; };
    std::vector<WaveMark> wave_marks;
};

#define RT_CUDA(call)                                                                                     \
    do {                                                                                                  \
        cudaError_t _e = (call);                                                                          \
        if (_e != cudaSuccess) {                                                                          \
            char _b[512];                                                                                 \
            snprintf(_b, sizeof(_b), "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
            err = _b;                                                                                     \
            return (_e == cudaErrorMemoryAllocation) ? RT_E_OOM : RT_E_CUDA;                              \
        }                                                                                                 \
    } while (0)

// rt_build.cu
// host -> device copy of caller memory that may be pageable: pinned sources go straight to the copy engine, pageable ones
// through a ring of pinned bounce buffers filled by `threads` host threads (the driver's own pageable path is one thread)
int rt_upload(RtDevice& D, void* dst, const void* src, size_t bytes, int threads, std::string& err);
void rt_upload_release(RtDevice& D);
// streaming commit: append n_tris mesh triangles (caller's arrays, read during the call only) at triangle offset `first`
int rt_device_stage_mesh(RtDevice& D, const float* v, const float* normals, int64_t first, int64_t n_tris, int threads, std::string& err);
int rt_device_update_raw(RtDevice& D, int what, int64_t first, int64_t n, const float* a, const float* b, int threads, std::string& err);
int rt_device_commit(RtDevice& D, const RtHostScene& H, const RtOptions& opt, std::string& err, float* ms_upload,
                     float* ms_build, uint64_t* h2d_bytes, bool raw_resident = false, bool refit = false);
// Mesh.AABB (SceneMesh.cs:22-31) of n instances (indices into IS.insts) from the object-space vertices resident on D: out6 = 6 floats
// each.  RT_E_STATE: D does not hold the current base meshes (the caller computes on the host).
int rt_device_instance_aabbs(RtDevice& D, const RtInstanceSet& IS, const int32_t* inst_idx, int n, float* out6, std::string& err);
// rt_render.cu
struct RtFrameStats { uint64_t primary, shadow, mirror, shadow_skipped; uint32_t launches; float ms_kernel[RT_K_COUNT]; uint32_t launches_kernel[RT_K_COUNT]; };
int rt_device_render(RtDevice& D, const RtView& V, int want, const RtOptions& opt, std::string& err, RtFrameStats* fs, int split = 0);
int rt_device_intersect(RtDevice& D, const float* d_origins, const float* d_dirs, int n, int32_t* d_ids, float* d_t,
                        const RtOptions& opt, std::string& err);
int rt_device_read_stats(RtDevice& D, RtFrameStats* fs, std::string& err);
// allocation-free rt_intersect: host arrays in and out, staging in a file-scope device array of rt_render.cu
int rt_device_intersect_static(RtDevice& D, const float* h_origins, const float* h_dirs, int n, int32_t* h_ids, float* h_t, std::string& err);

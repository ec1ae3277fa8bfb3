// rt_build.cu — scene upload + device LBVH build for sm_100a.
//   upload (scan-ordered staging arrays) -> primitive boxes + bounds -> 63-bit Morton codes -> radix sort (CUB, the
//   toolkit's library sort; commit is not the timed hot path) -> Karras hierarchy -> atomic bottom-up refit ->
//   leaf collapse + live-node compaction -> 64-byte BVH2 nodes with padded child boxes.
// Per-element logic lives in rt_build_steps.h (host/device) so tests can run it on the CPU.
#include "rt_internal.h"
#include "rt_sah.h"
#include <thread>
#include <cstring>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace {

constexpr int BT = 256;
inline int grid_for(int64_t n) { int64_t g = (n + BT - 1) / BT; return (int)(g < 1 ? 1 : (g > 148 * 32 ? 148 * 32 : g)); }

__device__ __forceinline__ uint32_t f_ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
inline float f_unord(uint32_t u) { u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u; float f; memcpy(&f, &u, 4); return f; }

// bounds[0..2] = min (ordered uint), bounds[3..5] = max
__global__ void k_bounds(RtRaw R, int32_t n, uint32_t* bounds) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        f3 a, b; rt_prim_box(R, i, a, b);
        lo[0] = fminf(lo[0], a.x); lo[1] = fminf(lo[1], a.y); lo[2] = fminf(lo[2], a.z);
        hi[0] = fmaxf(hi[0], b.x); hi[1] = fmaxf(hi[1], b.y); hi[2] = fmaxf(hi[2], b.z);
    }
    for (int k = 0; k < 3; ++k) {
        for (int o = 16; o; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        for (int k = 0; k < 3; ++k) { atomicMin(bounds + k, f_ord(lo[k])); atomicMax(bounds + 3 + k, f_ord(hi[k])); }
    }
}

__global__ void k_morton(RtRaw R, int32_t n, f3 bmin, f3 inv_ext, uint64_t* keys, int32_t* vals) {
    for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        f3 a, b; rt_prim_box(R, i, a, b);
        keys[i] = rt_morton(a, b, bmin, inv_ext);
        vals[i] = i;
    }
}

__global__ void k_emit_slots(RtRaw R, int32_t n, const int32_t* sorted_keys, float4* prims, float4* pinfo, float4* leaf_lo, float4* leaf_hi) {
    for (int32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x)
        rt_emit_slot(R, s, sorted_keys[s], prims, pinfo, leaf_lo, leaf_hi);
}

__global__ void k_karras(const uint64_t* keys, int32_t n, RtTopo T) {
    for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x) rt_karras_node(keys, n, i, T);
}

__device__ __forceinline__ float4 ld_cg4(const float4* p) { return __ldcg(p); }

// bottom-up refit: the second thread to reach a node computes its box (Karras 2012 §4)
__global__ void k_refit(int32_t n, RtTopo T, const float4* leaf_lo, const float4* leaf_hi, float4* int_lo, float4* int_hi, int32_t* flags) {
    for (int32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x) {
        int32_t p = T.parent_leaf[s];
        while (p >= 0) {
            __threadfence();
            if (atomicAdd(flags + p, 1) == 0) break;            // first arrival: the sibling subtree is not done yet
            int32_t l = T.left[p], r = T.right[p];
            float4 alo = l < 0 ? leaf_lo[~l] : ld_cg4(int_lo + l), ahi = l < 0 ? leaf_hi[~l] : ld_cg4(int_hi + l);
            float4 blo = r < 0 ? leaf_lo[~r] : ld_cg4(int_lo + r), bhi = r < 0 ? leaf_hi[~r] : ld_cg4(int_hi + r);
            float4 lo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
            float4 hi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
            __stcg(int_lo + p, lo); __stcg(int_hi + p, hi);
            p = T.parent_int[p];
        }
    }
}

__global__ void k_live(int32_t n_int, RtTopo T, int32_t leaf_max, int32_t* live) {
    for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_int; i += gridDim.x * blockDim.x)
        live[i] = (T.last[i] - T.first[i] + 1 > leaf_max) ? 1 : 0;
}

__global__ void k_emit_nodes(int32_t n_int, RtTopo T, const int32_t* new_index, int32_t leaf_max, float pad, const float4* leaf_lo,
                             const float4* leaf_hi, const float4* int_lo, const float4* int_hi, float4* nodes) {
    for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_int; i += gridDim.x * blockDim.x)
        rt_emit_node(i, T, new_index, leaf_max, pad, leaf_lo, leaf_hi, int_lo, int_hi, nodes);
}

__global__ void k_node_parents(const float4* nodes, int32_t n_nodes, int32_t* parent) {
    for (int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += gridDim.x * blockDim.x) rt_node_parents(nodes, i, parent);
}

// Refit of the emitted node array, bottom-up: a thread starts at every node whose children are both leaves and climbs; a
// node with inner children is processed by the LAST of them to arrive (arrival counter), after a fence, so it reads finished
// boxes (same scheme as k_refit on the Karras tree).  int_lo/int_hi hold the unpadded box per node; they are written with
// st.cg and read with ld.cg by another thread.
__global__ void k_refit_nodes(int32_t n_nodes, float pad, const int32_t* parent, const float4* leaf_lo, const float4* leaf_hi,
                              float4* int_lo, float4* int_hi, int32_t* arrived, float4* nodes) {
    for (int32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n_nodes; s += gridDim.x * blockDim.x) {
        float4 n3 = nodes[4 * (size_t)s + 3];
        if (f2i(n3.x) >= 0 || f2i(n3.y) >= 0) continue;            // has an inner child: reached from below
        int32_t i = s;
        for (;;) {
            float4* np = nodes + 4 * (size_t)i;
            float4 r3 = np[3];
            f3 alo, ahi, blo, bhi;
            const int32_t ra = f2i(r3.x), rb = f2i(r3.y);
            if (ra >= 0) { float4 a = ld_cg4(int_lo + ra), b = ld_cg4(int_hi + ra); alo = mk3(a.x, a.y, a.z); ahi = mk3(b.x, b.y, b.z); }
            else rt_ref_box(ra, leaf_lo, leaf_hi, int_lo, int_hi, alo, ahi);
            if (rb >= 0) { float4 a = ld_cg4(int_lo + rb), b = ld_cg4(int_hi + rb); blo = mk3(a.x, a.y, a.z); bhi = mk3(b.x, b.y, b.z); }
            else rt_ref_box(rb, leaf_lo, leaf_hi, int_lo, int_hi, blo, bhi);
            np[0] = make_float4(alo.x - pad, ahi.x + pad, alo.y - pad, ahi.y + pad);
            np[1] = make_float4(blo.x - pad, bhi.x + pad, blo.y - pad, bhi.y + pad);
            np[2] = make_float4(alo.z - pad, ahi.z + pad, blo.z - pad, bhi.z + pad);
            __stcg(int_lo + i, make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f));
            __stcg(int_hi + i, make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f));
            const int32_t p = parent[i];
            if (p < 0) break;
            float4 p3 = nodes[4 * (size_t)p + 3];
            const int32_t need = (f2i(p3.x) >= 0 ? 1 : 0) + (f2i(p3.y) >= 0 ? 1 : 0);
            __threadfence();
            if (atomicAdd(arrived + p, 1) + 1 < need) break;        // the other inner child is not done yet
            i = p;
        }
    }
}

}  // namespace

// ---- pageable host -> device upload --------------------------------------------------------------------------------------
// cudaMemcpyAsync from pageable memory is staged by the driver on ONE thread (measured 11.8 GB/s on this box: 4.1 ms for
// cfg3's 48 MB, 86 % of a steady-state re-commit).  Here `threads` host threads copy 2 MiB chunks into their own two pinned
// slots and queue the DMA of each chunk behind it, so the memcpys run in parallel with each other and with PCIe:
// 4 threads 21 GB/s (2.2 ms); 8 threads no better — the box's host memcpy bandwidth (14 GB/s per thread, ~30 GB/s total) binds.
void rt_upload_release(RtDevice& D) {
    for (auto e : D.up_ev) cudaEventDestroy(e);
    D.up_ev.clear();
    if (D.up_ring) cudaFreeHost(D.up_ring);
    D.up_ring = nullptr; D.up_slots = 0; D.up_slot_bytes = 0;
}

int rt_upload(RtDevice& D, void* dst, const void* src, size_t bytes, int threads, std::string& err) {
    if (bytes == 0) return RT_OK;
    cudaStream_t st = D.stream;
    cudaPointerAttributes at{};
    bool pinned = cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();                                        // unregistered host memory reports an error on old runtimes
    constexpr size_t SLOT = 2u << 20;                          // 0.5 / 1 / 4 / 8 MiB measured slower (2.6 / 2.8 / 3.2 / 3.3 ms vs 2.2 ms for 48 MB)
    if (threads > 16) threads = 16;
    if (pinned || threads < 2 || bytes < 4 * SLOT) {
        RT_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
        return RT_OK;
    }
    const int slots = 2 * threads;
    if (D.up_slots < slots || D.up_slot_bytes != SLOT) {
        RT_CUDA(cudaStreamSynchronize(st));
        rt_upload_release(D);
        RT_CUDA(cudaHostAlloc(&D.up_ring, (size_t)slots * SLOT, cudaHostAllocDefault));
        D.up_slots = slots; D.up_slot_bytes = SLOT;
        for (int i = 0; i < slots; ++i) { cudaEvent_t e; RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); D.up_ev.push_back(e); }
    }
    const size_t n_chunks = (bytes + SLOT - 1) / SLOT;
    std::vector<cudaError_t> rc((size_t)threads, cudaSuccess);
    auto work = [&](int t) {
        cudaError_t e = cudaSetDevice(D.device);
        int turn = 0;
        for (size_t c = (size_t)t; c < n_chunks && e == cudaSuccess; c += (size_t)threads, ++turn) {
            const int slot = 2 * t + (turn & 1);
            char* bounce = (char*)D.up_ring + (size_t)slot * SLOT;
            const size_t off = c * SLOT, len = bytes - off < SLOT ? bytes - off : SLOT;
            e = cudaEventSynchronize(D.up_ev[slot]);       // the DMA that last read this slot (this or an earlier upload) is done
            if (e != cudaSuccess) break;
            std::memcpy(bounce, (const char*)src + off, len);
            e = cudaMemcpyAsync((char*)dst + off, bounce, len, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaEventRecord(D.up_ev[slot], st);
        }
        rc[(size_t)t] = e;
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    for (cudaError_t e : rc) RT_CUDA(e);
    return RT_OK;                                              // all DMAs are queued on D.stream; the bounce slots stay valid
}

int rt_device_stage_mesh(RtDevice& D, const float* v, const float* normals, int64_t first, int64_t n_tris, int threads, std::string& err) {
    RT_CUDA(cudaSetDevice(D.device));
    cudaStream_t st = D.stream;
    RT_CUDA(D.raw_v.ensure_keep((size_t)(first + n_tris) * 36, (size_t)first * 36, st));
    RT_CUDA(D.raw_n.ensure_keep((size_t)(first + n_tris) * 12, (size_t)first * 12, st));
    if (n_tris > 0) {
        int rc = rt_upload(D, (char*)D.raw_v.p + (size_t)first * 36, v, (size_t)n_tris * 36, threads, err);
        if (rc == RT_OK) rc = rt_upload(D, (char*)D.raw_n.p + (size_t)first * 12, normals, (size_t)n_tris * 12, threads, err);
        if (rc) return rc;
        RT_CUDA(cudaStreamSynchronize(st));          // the caller's arrays are only valid during the call (pinned callers too)
    }
    return RT_OK;
}

// in-place replacement of a range of the device-resident raw arrays of a committed scene (rt_scene_update_*)
int rt_device_update_raw(RtDevice& D, int what, int64_t first, int64_t n, const float* a, const float* b, int threads, std::string& err) {
    RT_CUDA(cudaSetDevice(D.device));
    if (n <= 0) return RT_OK;
    int rc = RT_OK;
    if (what == 0) {                                            // triangles: a = 9 floats each, b = 3 floats each
        if ((size_t)(first + n) * 36 > D.raw_v.bytes || (size_t)(first + n) * 12 > D.raw_n.bytes) { err = "update beyond the committed arrays"; return RT_E_INVALID; }
        rc = rt_upload(D, (char*)D.raw_v.p + (size_t)first * 36, a, (size_t)n * 36, threads, err);
        if (rc == RT_OK) rc = rt_upload(D, (char*)D.raw_n.p + (size_t)first * 12, b, (size_t)n * 12, threads, err);
    } else {                                                    // spheres: a = 4 floats each
        if ((size_t)(first + n) * 16 > D.raw_s.bytes) { err = "update beyond the committed arrays"; return RT_E_INVALID; }
        rc = rt_upload(D, (char*)D.raw_s.p + (size_t)first * 16, a, (size_t)n * 16, threads, err);
    }
    if (rc) return rc;
    RT_CUDA(cudaStreamSynchronize(D.stream));
    return RT_OK;
}

// raw_resident: the triangle / sphere arrays of this scene are already on the device (rt_scene_recommit after
// rt_scene_update_*): only the small tables are uploaded before the BVH is rebuilt.
// ---- Mesh.AABB of instances on the device (rt_scene_mesh_instance / rt_scene_update_instance(s) with aabb = NULL) -------------
// SceneMesh.cs:22-31: min / max over the baked vertices.  The object-space vertices are resident (iprims, the BLAS leaf slots),
// so a re-posed instance costs one pass over them at L2 bandwidth instead of one host thread's pass (cfg4: 125 instances of
// 81 920 triangles, 8 ms on 16 host threads).  Min / max of finite values do not depend on the order, so the parallel reduction
// gives the bits of the reference's sequential loop — except the sign of a zero extremum (math.min/max keep the LATER of +0 / -0),
// which the caller re-derives on the host (rt_abi.cu resolve_pending_aabbs).
struct RtAabbReq { int32_t first_slot, n_tris; float m[12]; };
__global__ void k_instance_aabbs(const float4* __restrict__ iprims, const RtAabbReq* __restrict__ req, uint32_t* __restrict__ out) {
    const RtAabbReq R = req[blockIdx.x];
    const float4 r0 = make_float4(R.m[0], R.m[1], R.m[2], R.m[3]), r1 = make_float4(R.m[4], R.m[5], R.m[6], R.m[7]), r2 = make_float4(R.m[8], R.m[9], R.m[10], R.m[11]);
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};     // float.MaxValue / float.MinValue
    const int64_t n_vert = 3 * (int64_t)R.n_tris;
    for (int64_t i = (int64_t)blockIdx.y * blockDim.x + threadIdx.x; i < n_vert; i += (int64_t)gridDim.y * blockDim.x) {
        const float4 p = iprims[3 * (size_t)R.first_slot + (size_t)i];
        const f3 w = rt_xform_point(r0, r1, r2, mk3(p.x, p.y, p.z));                         // MultiplyPoint3x4, exact operation order
        lo[0] = fminf(lo[0], w.x); lo[1] = fminf(lo[1], w.y); lo[2] = fminf(lo[2], w.z);     // (a NaN is skipped, as um_min / um_max skip it)
        hi[0] = fmaxf(hi[0], w.x); hi[1] = fmaxf(hi[1], w.y); hi[2] = fmaxf(hi[2], w.z);
    }
    for (int k = 0; k < 3; ++k)
        for (int o = 16; o > 0; o >>= 1) { lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o)); hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o)); }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 3; ++k) { atomicMin(out + 6 * (size_t)blockIdx.x + k, f_ord(lo[k])); atomicMax(out + 6 * (size_t)blockIdx.x + 3 + k, f_ord(hi[k])); }
}

int rt_device_instance_aabbs(RtDevice& D, const RtInstanceSet& IS, const int32_t* inst_idx, int n, float* out6, std::string& err) {
    if (n <= 0) return RT_OK;
    if (D.inst_blas_gen != IS.blas_generation || IS.blas_changed || !D.iprims.p || IS.slot_base.size() != IS.bases.size() + 1) return RT_E_STATE;   // the device does not hold these base meshes
    RT_CUDA(cudaSetDevice(D.device));
    cudaStream_t st = D.stream;
    std::vector<RtAabbReq> req((size_t)n);
    std::vector<uint32_t> box((size_t)n * 6);
    int32_t max_tris = 1;
    for (int i = 0; i < n; ++i) {
        const RtInstance& I = IS.insts[(size_t)inst_idx[i]];
        req[(size_t)i].first_slot = IS.slot_base[(size_t)I.base]; req[(size_t)i].n_tris = IS.bases[(size_t)I.base].n_tris();
        memcpy(req[(size_t)i].m, I.m, sizeof(I.m));
        if (req[(size_t)i].n_tris > max_tris) max_tris = req[(size_t)i].n_tris;
        for (int k = 0; k < 3; ++k) { box[6 * (size_t)i + k] = 0xffffffffu; box[6 * (size_t)i + 3 + k] = 0u; }
    }
    const size_t req_bytes = (size_t)n * sizeof(RtAabbReq), box_off = (req_bytes + 255) & ~(size_t)255, box_bytes = (size_t)n * 24;
    RT_CUDA(D.cub_tmp.ensure(box_off + box_bytes));                     // (build scratch: free between commits)
    RT_CUDA(cudaMemcpyAsync(D.cub_tmp.p, req.data(), req_bytes, cudaMemcpyHostToDevice, st));
    RT_CUDA(cudaMemcpyAsync((char*)D.cub_tmp.p + box_off, box.data(), box_bytes, cudaMemcpyHostToDevice, st));
    int parts = (D.sm_count * 8 + n - 1) / n;
    const int max_parts = (3 * max_tris + 255) / 256;
    parts = parts < 1 ? 1 : (parts > max_parts ? max_parts : parts);
    if (parts > 65535) parts = 65535;
    k_instance_aabbs<<<dim3((unsigned)n, (unsigned)parts), 256, 0, st>>>(D.iprims.as<float4>(), (const RtAabbReq*)D.cub_tmp.p, (uint32_t*)((char*)D.cub_tmp.p + box_off));
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(box.data(), (char*)D.cub_tmp.p + box_off, box_bytes, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    for (size_t k = 0; k < box.size(); ++k) out6[k] = f_unord(box[k]);
    return RT_OK;
}

int rt_device_commit(RtDevice& D, const RtHostScene& H, const RtOptions& opt, std::string& err, float* ms_upload,
                     float* ms_build, uint64_t* h2d_bytes, bool raw_resident, bool refit) {
    RT_CUDA(cudaSetDevice(D.device));
    cudaStream_t st = D.stream;
    const int32_t n_mt = H.n_meshtris(), n_sp = H.n_spheres(), n_lo = H.n_loose(), n_me = H.n_meshes(), n_li = H.n_lights();
    const int64_t n64 = H.n_prims();
    if (n64 > (int64_t)(1 << 28) - 8) { err = "too many primitives (limit 2^28)"; return RT_E_UNSUPPORTED; }
    const int32_t n = (int32_t)n64;
    uint64_t up = 0;
    RT_CUDA(cudaEventRecord(D.ev0, st));
    // ---- upload ----
    const size_t ntri = (size_t)n_mt + n_lo;
    const size_t keep_tris = raw_resident ? ntri : (H.mesh_on_device ? (size_t)n_mt : 0);   // streamed by rt_scene_mesh already
    if (raw_resident && (D.raw_v.bytes < ntri * 36 || D.raw_n.bytes < ntri * 12 || (n_sp && D.raw_s.bytes < (size_t)n_sp * 16))) {
        err = "rt_scene_recommit: no committed scene of this size on the device"; return RT_E_STATE;
    }
    RT_CUDA(D.raw_v.ensure_keep(ntri * 9 * 4, keep_tris * 36, st)); RT_CUDA(D.raw_n.ensure_keep(ntri * 3 * 4, keep_tris * 12, st));
    if (!raw_resident) RT_CUDA(D.raw_s.ensure((size_t)n_sp * 16));
    if (n_mt && !H.mesh_on_device && !raw_resident) {
        int rc = rt_upload(D, D.raw_v.p, H.mesh_v.data(), (size_t)n_mt * 36, opt.upload_threads, err);
        if (rc == RT_OK) rc = rt_upload(D, D.raw_n.p, H.mesh_n.data(), (size_t)n_mt * 12, opt.upload_threads, err);
        if (rc) return rc;
    }
    if (n_lo && !raw_resident) {
        int rc = rt_upload(D, (char*)D.raw_v.p + (size_t)n_mt * 36, H.loose_v.data(), (size_t)n_lo * 36, opt.upload_threads, err);
        if (rc == RT_OK) rc = rt_upload(D, (char*)D.raw_n.p + (size_t)n_mt * 12, H.loose_n.data(), (size_t)n_lo * 12, opt.upload_threads, err);
        if (rc) return rc;
    }
    if (n_sp && !raw_resident) RT_CUDA(cudaMemcpyAsync(D.raw_s.p, H.spheres.data(), (size_t)n_sp * 16, cudaMemcpyHostToDevice, st));
    up += (ntri - keep_tris) * 48 + (raw_resident ? 0 : (uint64_t)n_sp * 16);
    std::vector<float> mats;
    if (!raw_resident) {                                       // materials cannot change through rt_scene_update_*
        RtHostScene& Hm = const_cast<RtHostScene&>(H);
        mats = Hm.pack_materials();
        D.has_mirror = Hm.has_mirror;
        RT_CUDA(D.mats.ensure(mats.size() * 4));
        if (!mats.empty()) RT_CUDA(cudaMemcpyAsync(D.mats.p, mats.data(), mats.size() * 4, cudaMemcpyHostToDevice, st));
    }
    std::vector<float> maabb((size_t)n_me * 8, 0.f);
    for (int m = 0; m < n_me; ++m) { for (int k = 0; k < 3; ++k) { maabb[8 * m + k] = H.mesh_aabb[6 * m + k]; maabb[8 * m + 4 + k] = H.mesh_aabb[6 * m + 3 + k]; } }
    RT_CUDA(D.mesh_aabb.ensure(maabb.size() * 4));
    if (n_me) RT_CUDA(cudaMemcpyAsync(D.mesh_aabb.p, maabb.data(), maabb.size() * 4, cudaMemcpyHostToDevice, st));
    const int32_t n_kt = H.n_keytris();                        // scan keys of all mesh triangles, instanced ones included
    if ((int64_t)n64 + H.inst_tris > (int64_t)(1 << 28) - 8) { err = "too many primitives (limit 2^28)"; return RT_E_UNSUPPORTED; }
    std::vector<int32_t> mbase(H.mesh_base); mbase.push_back(n_kt);
    RT_CUDA(D.mesh_base.ensure(mbase.size() * 4));
    RT_CUDA(cudaMemcpyAsync(D.mesh_base.p, mbase.data(), mbase.size() * 4, cudaMemcpyHostToDevice, st));
    std::vector<int32_t> mraw(H.mesh_raw); mraw.push_back(n_mt);
    RT_CUDA(D.mesh_raw.ensure(mraw.size() * 4));
    RT_CUDA(cudaMemcpyAsync(D.mesh_raw.p, mraw.data(), mraw.size() * 4, cudaMemcpyHostToDevice, st));
    std::vector<float> lights((size_t)n_li * 8, 0.f);
    for (int l = 0; l < n_li; ++l) { for (int k = 0; k < 3; ++k) { lights[8 * l + k] = H.lights[6 * l + k]; lights[8 * l + 4 + k] = H.lights[6 * l + 3 + k]; } }
    RT_CUDA(D.lights.ensure(lights.size() * 4));
    if (n_li) RT_CUDA(cudaMemcpyAsync(D.lights.p, lights.data(), lights.size() * 4, cudaMemcpyHostToDevice, st));
    up += mats.size() * 4 + maabb.size() * 4 + mbase.size() * 8 + lights.size() * 4;
    RT_CUDA(cudaStreamSynchronize(st));        // staging vectors above go out of scope; also separates upload/build timing
    RT_CUDA(cudaEventRecord(D.ev1, st));
    RT_CUDA(cudaEventSynchronize(D.ev1));
    if (ms_upload) RT_CUDA(cudaEventElapsedTime(ms_upload, D.ev0, D.ev1));
    if (h2d_bytes) *h2d_bytes = up;

    // ---- scene view (pointers filled below) ----
    RtScene& S = D.S;
    D.hint_ready = false;                      // tile-cost hints describe the previous scene's frames
    const int32_t old_prims = D.bvh_prims, old_root = D.bvh_root, old_slots = D.bvh_slots;
    const uint32_t old_nodes = D.n_nodes;
    if (refit && n <= 0 && H.inst.any()) refit = false;       // nothing but instances: a refit is the TLAS rebuild every commit does
    if (refit && (old_prims != n || n <= 0 || old_slots < n || D.vals_alt.bytes < (size_t)old_slots * 4)) { err = "rt_scene_refit: no committed tree over this many primitives"; return RT_E_STATE; }
    D.bvh_prims = -1;
    S = RtScene{};
    S.n_prims = n; S.n_meshes = n_me; S.n_meshtris = n_kt; S.n_spheres = n_sp; S.n_lights = n_li;
    S.aabb_min = mk3(H.aabb[0], H.aabb[1], H.aabb[2]); S.aabb_max = mk3(H.aabb[3], H.aabb[4], H.aabb[5]);
    S.ambient = mk3(H.ambient[0], H.ambient[1], H.ambient[2]);
    S.t_slack = (float)opt.t_slack;
    S.packet = opt.packet;
    {   // packet traversal: origins within 5 % of the scene diagonal
        float dx = H.aabb[3] - H.aabb[0], dy = H.aabb[4] - H.aabb[1], dz = H.aabb[5] - H.aabb[2];
        float d2 = dx * dx + dy * dy + dz * dz;
        S.coherent_spread2 = std::isfinite(d2) ? 0.0025f * d2 : 0.f;
    }
    S.backcull = (opt.backcull != 0 && H.backcull_ok()) ? 1 : 0;
    S.mats = D.mats.as<float4>(); S.mesh_aabb = D.mesh_aabb.as<float4>(); S.mesh_base = D.mesh_base.as<int32_t>();
    S.lights = D.lights.as<float4>();
    D.n_nodes = 0;
    RT_CUDA(cudaEventRecord(D.ev0, st));
    // Instanced meshes: BLAS / TLAS / instance records (rt_instances.h), packed once per commit (device 0 of the ctx; bounds and
    // builders are deterministic, so the other devices upload the same arrays) once the scene's extent — which sizes the
    // padding — is known.  The BLAS part travels only when it changed; a re-pose sends the TLAS and 128 bytes per instance.
    RtInstanceSet& IS = const_cast<RtHostScene&>(H).inst;
    auto upload_instances = [&](float amax_w, float pad_w) -> int {
        if (!IS.any()) return RT_OK;
        if (D.rank == 0) {
            const char* why = nullptr;
            int threads = (int)std::thread::hardware_concurrency();
            const int leaf_max_i = opt.leaf_max < 1 ? 1 : (opt.leaf_max > 3 ? 3 : opt.leaf_max);     // <= 3 triangles per BLAS leaf: one cooperative bake (rt_packet.cuh)
            if (!IS.pack(mbase, amax_w, pad_w, opt.pad_scale, leaf_max_i, threads < 1 ? 1 : threads, RT_STACK_ENTRIES, &why)) { err = why ? why : "instance packing failed"; return RT_E_UNSUPPORTED; }
        }
        if ((int64_t)S.n_slots + IS.n_tris() > (int64_t)0x7fffff00) { err = "too many instanced triangles (virtual slots exceed 2^31)"; return RT_E_UNSUPPORTED; }
        const size_t blas_floats = (size_t)IS.tlas_base * 16;
        const bool grow = D.inodes.bytes < IS.inodes.size() * 4 + 64;
        RT_CUDA(D.inodes.ensure(IS.inodes.size() * 4 + 64)); RT_CUDA(D.iprims.ensure(IS.iprims.size() * 4 + 48));
        RT_CUDA(D.inst.ensure(IS.inst.size() * 4 + 16)); RT_CUDA(D.inst_vbase.ensure(IS.vbase.size() * 4));
        if (grow || D.inst_blas_gen != IS.blas_generation) {
            if (blas_floats) RT_CUDA(cudaMemcpyAsync(D.inodes.p, IS.inodes.data(), blas_floats * 4, cudaMemcpyHostToDevice, st));
            if (!IS.iprims.empty()) RT_CUDA(cudaMemcpyAsync(D.iprims.p, IS.iprims.data(), IS.iprims.size() * 4, cudaMemcpyHostToDevice, st));
            D.inst_blas_gen = IS.blas_generation;
        }
        if (IS.inodes.size() > blas_floats)
            RT_CUDA(cudaMemcpyAsync((char*)D.inodes.p + blas_floats * 4, IS.inodes.data() + blas_floats, (IS.inodes.size() - blas_floats) * 4, cudaMemcpyHostToDevice, st));
        if (!IS.inst.empty()) RT_CUDA(cudaMemcpyAsync(D.inst.p, IS.inst.data(), IS.inst.size() * 4, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaMemcpyAsync(D.inst_vbase.p, IS.vbase.data(), IS.vbase.size() * 4, cudaMemcpyHostToDevice, st));
        RT_CUDA(cudaStreamSynchronize(st));                      // the packed arrays may be re-packed by the next commit
        S.inodes = D.inodes.as<float4>(); S.iprims = D.iprims.as<float4>(); S.inst = D.inst.as<float4>(); S.inst_vbase = D.inst_vbase.as<int32_t>();
        S.tlas_root = IS.tlas_root; S.n_inst = IS.n_packed;
        S.n_inodes = (int32_t)(IS.inodes.size() / 16); S.n_islots = (int32_t)(IS.iprims.size() / 12);
        // cameras beyond 2 x the scene's extent take the FAR form (ray-dependent padding): the BLAS padding assumes it (need_pad)
        S.origin_reach = fminf(S.origin_reach, 2.0f * amax_w);
        return RT_OK;
    };
    if (n <= 0 && IS.any()) {                                   // nothing but instances: no world BVH
        float b6[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}, amax = 0.f;
        IS.grow_bounds(b6, amax);
        const float pad = (float)(amax * opt.pad_scale) + 1e-30f;
        S.origin_reach = rt_origin_reach(pad);
        H.origin_box(b6, amax, D.obox);
        S.n_slots = 0;
        int rc = upload_instances(amax, pad);
        if (rc) return rc;
        D.bvh_prims = 0; D.bvh_slots = 0; D.bvh_root = 0; D.bvh_parents = false;
    }
    if (n > 0) {
        RtRaw R{D.raw_v.as<float>(), D.raw_n.as<float>(), D.raw_s.as<float4>(), D.mesh_raw.as<int32_t>(), n_me, n_mt, n_sp, n_lo,
                mk3(0.f, 0.f, 0.f), mk3(0.f, 0.f, 0.f), 0.0f, D.mesh_base.as<int32_t>(), n_kt};   // sphere boxes un-inflated until the origin box is known (set_obox)
        auto set_obox = [&](const float b6[6], float amax_) {
            H.origin_box(b6, amax_, D.obox);
            R.obox_lo = mk3(D.obox[0], D.obox[1], D.obox[2]); R.obox_hi = mk3(D.obox[3], D.obox[4], D.obox[5]);
            R.sphere_k = RT_SPHERE_K;
        };
        S.n_slots = n;
        const int leaf_max = opt.leaf_max < 1 ? 1 : (opt.leaf_max > 8 ? 8 : opt.leaf_max);
        if (refit) {
            // ---- refit: keep the tree (either builder's) and the slot order, recompute every box from the moved primitives ----
            RT_CUDA(D.bounds.ensure(6 * 4));
            uint32_t binit[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
            RT_CUDA(cudaMemcpyAsync(D.bounds.p, binit, sizeof(binit), cudaMemcpyHostToDevice, st));
            k_bounds<<<grid_for(n) < D.sm_count * 4 ? grid_for(n) : D.sm_count * 4, BT, 0, st>>>(R, n, D.bounds.as<uint32_t>());
            uint32_t bord[6];
            RT_CUDA(cudaMemcpyAsync(bord, D.bounds.p, sizeof(bord), cudaMemcpyDeviceToHost, st));
            // (slots of a pre-split tree get the whole triangle's box again: conservative, the clipping is lost until the next commit)
            RT_CUDA(D.leaf_lo.ensure((size_t)old_slots * 16)); RT_CUDA(D.leaf_hi.ensure((size_t)old_slots * 16));
            RT_CUDA(cudaStreamSynchronize(st));
            float amax = 0.f, b6[6];
            for (int k = 0; k < 6; ++k) { b6[k] = f_unord(bord[k]); if (std::isfinite(b6[k])) amax = fmaxf(amax, fabsf(b6[k])); }
            IS.grow_bounds(b6, amax);
            const float pad = (float)(amax * opt.pad_scale) + 1e-30f;
            S.origin_reach = rt_origin_reach(pad);
            set_obox(b6, amax);
            S.n_slots = old_slots;
            { int rci = upload_instances(amax, pad); if (rci) return rci; }
            k_emit_slots<<<grid_for(old_slots), BT, 0, st>>>(R, old_slots, D.vals_alt.as<int32_t>(), D.prims.as<float4>(), D.pinfo.as<float4>(), D.leaf_lo.as<float4>(), D.leaf_hi.as<float4>());
            D.n_nodes = old_nodes;
            if (old_nodes > 0) {
                const int32_t nn = (int32_t)old_nodes;
                RT_CUDA(D.int_lo.ensure((size_t)nn * 16)); RT_CUDA(D.int_hi.ensure((size_t)nn * 16));
                RT_CUDA(D.flags.ensure((size_t)nn * 4)); RT_CUDA(D.node_parent.ensure((size_t)nn * 4));
                if (!D.bvh_parents) {
                    RT_CUDA(cudaMemsetAsync(D.node_parent.p, 0xff, (size_t)nn * 4, st));
                    k_node_parents<<<grid_for(nn), BT, 0, st>>>(D.nodes.as<float4>(), nn, D.node_parent.as<int32_t>());
                    D.bvh_parents = true;
                }
                RT_CUDA(cudaMemsetAsync(D.flags.p, 0, (size_t)nn * 4, st));
                k_refit_nodes<<<grid_for(nn), BT, 0, st>>>(nn, pad, D.node_parent.as<int32_t>(), D.leaf_lo.as<float4>(), D.leaf_hi.as<float4>(),
                                                          D.int_lo.as<float4>(), D.int_hi.as<float4>(), D.flags.as<int32_t>(), D.nodes.as<float4>());
            }
            RT_CUDA(cudaGetLastError());
            S.prims = D.prims.as<float4>(); S.pinfo = D.pinfo.as<float4>(); S.nodes = D.nodes.as<float4>();
            S.root = old_root;
            D.bvh_prims = n; D.bvh_slots = old_slots; D.bvh_root = old_root;
            S.n_nodes = (int32_t)D.n_nodes;
            RT_CUDA(cudaEventRecord(D.ev1, st));
            RT_CUDA(cudaEventSynchronize(D.ev1));
            if (ms_build) RT_CUDA(cudaEventElapsedTime(ms_build, D.ev0, D.ev1));
            return RT_OK;
        }
        if (H.builder == 1) {
            // ---- host binned-SAH build (rt_sah.h); the device only re-lays the primitives out in leaf order ----
            std::vector<RtSahBox> pb((size_t)n);
            float amax = 0.f;
            if (n_sp > 0) {                                        // origin box first: the un-inflated bounds of everything
                float b6[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}, am = 0.f;
                auto grow = [&](float v, int k) { b6[k] = fminf(b6[k], v); b6[3 + k] = fmaxf(b6[3 + k], v); if (std::isfinite(v)) am = fmaxf(am, fabsf(v)); };
                for (size_t i = 0; i < (size_t)n_mt * 9; ++i) grow(H.mesh_v[i], (int)(i % 3));
                for (size_t i = 0; i < (size_t)n_lo * 9; ++i) grow(H.loose_v[i], (int)(i % 3));
                for (int32_t i = 0; i < n_sp; ++i) { const float* sp = H.spheres.data() + 4 * (size_t)i; const float r = sqrtf(sp[3]); for (int k = 0; k < 3; ++k) { grow(sp[k] - r, k); grow(sp[k] + r, k); } }
                IS.grow_bounds(b6, am);
                set_obox(b6, am);
            }
            for (int32_t key = 0; key < n; ++key) {
                RtSahBox& x = pb[key];
                if (key >= n_mt && key < n_mt + n_sp) {
                    const float* sp = H.spheres.data() + 4 * (size_t)(key - n_mt);
                    const float r = rt_sphere_box_radius(R, make_float4(sp[0], sp[1], sp[2], sp[3]));
                    for (int k = 0; k < 3; ++k) { x.lo[k] = sp[k] - r; x.hi[k] = sp[k] + r; }
                } else {
                    const float* v = key < n_mt ? H.mesh_v.data() + 9 * (size_t)key : H.loose_v.data() + 9 * (size_t)(key - n_mt - n_sp);
                    for (int k = 0; k < 3; ++k) { x.lo[k] = fminf(v[k], fminf(v[3 + k], v[6 + k])); x.hi[k] = fmaxf(v[k], fmaxf(v[3 + k], v[6 + k])); }
                }
                for (int k = 0; k < 3; ++k) {
                    if (std::isfinite(x.lo[k])) amax = fmaxf(amax, fabsf(x.lo[k]));
                    if (std::isfinite(x.hi[k])) amax = fmaxf(amax, fabsf(x.hi[k]));
                }
            }
            { float bi[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}; IS.grow_bounds(bi, amax); }
            const float pad = (float)(amax * opt.pad_scale) + 1e-30f;
            S.origin_reach = rt_origin_reach(pad);
            RtSahBuilder B;
            B.term = (float)opt.sah_term;
            int threads = (int)std::thread::hardware_concurrency();
            // reference pre-splitting (rt_sah.h RtPresplit): thin diagonal triangles enter the builder as several clipped pieces
            std::vector<RtSahBox> rb; std::vector<int32_t> rk;
            const RtSahBox* items = pb.data();
            int32_t ns = n;                                           // leaf slots = BVH references (>= n with pre-splitting)
            if (opt.presplit > 0.0) {
                auto tri = [&](int32_t key) -> const float* {
                    if (key < n_mt) return H.mesh_v.data() + 9 * (size_t)key;
                    if (key >= n_mt + n_sp) return H.loose_v.data() + 9 * (size_t)(key - n_mt - n_sp);
                    return nullptr;
                };
                RtPresplit::run(pb.data(), n, tri, opt.presplit, opt.presplit_depth < 0 ? 0 : (opt.presplit_depth > 6 ? 6 : opt.presplit_depth),
                                threads < 1 ? 1 : threads, rb, rk);
                if (rb.size() > (size_t)(1 << 28) - 8) { err = "too many BVH references after pre-splitting"; return RT_E_UNSUPPORTED; }
                items = rb.data(); ns = (int32_t)rb.size();
            }
            const int32_t root = B.run(items, ns, leaf_max, pad, threads < 1 ? 1 : threads);
            if (B.max_depth.load() + 2 > RT_STACK_ENTRIES) { err = "host SAH tree deeper than the traversal stack"; return RT_E_UNSUPPORTED; }
            if (!rk.empty()) for (int32_t s2 = 0; s2 < ns; ++s2) B.order[(size_t)s2] = rk[(size_t)B.order[(size_t)s2]];      // slot -> scan key
            RT_CUDA(D.vals_alt.ensure((size_t)ns * 4));
            RT_CUDA(cudaMemcpyAsync(D.vals_alt.p, B.order.data(), (size_t)ns * 4, cudaMemcpyHostToDevice, st));
            D.n_nodes = (uint32_t)(B.nodes.size() / 16);
            RT_CUDA(D.nodes.ensure(B.nodes.size() * 4 + 64));
            if (!B.nodes.empty()) RT_CUDA(cudaMemcpyAsync(D.nodes.p, B.nodes.data(), B.nodes.size() * 4, cudaMemcpyHostToDevice, st));
            RT_CUDA(D.prims.ensure((size_t)ns * 48)); RT_CUDA(D.pinfo.ensure((size_t)ns * 16));
            RT_CUDA(D.leaf_lo.ensure((size_t)ns * 16)); RT_CUDA(D.leaf_hi.ensure((size_t)ns * 16));
            k_emit_slots<<<grid_for(ns), BT, 0, st>>>(R, ns, D.vals_alt.as<int32_t>(), D.prims.as<float4>(), D.pinfo.as<float4>(), D.leaf_lo.as<float4>(), D.leaf_hi.as<float4>());
            RT_CUDA(cudaGetLastError());
            RT_CUDA(cudaStreamSynchronize(st));                   // B.order / B.nodes go out of scope
            S.prims = D.prims.as<float4>(); S.pinfo = D.pinfo.as<float4>(); S.nodes = D.nodes.as<float4>();
            S.root = root;
            S.n_slots = ns;
            { int rci = upload_instances(amax, pad); if (rci) return rci; }
            D.bvh_prims = n; D.bvh_slots = ns; D.bvh_root = root; D.bvh_parents = false;
            S.n_nodes = (int32_t)D.n_nodes;
            RT_CUDA(cudaEventRecord(D.ev1, st));
            RT_CUDA(cudaEventSynchronize(D.ev1));
            if (ms_build) RT_CUDA(cudaEventElapsedTime(ms_build, D.ev0, D.ev1));
            return RT_OK;
        }
        // bounds
        RT_CUDA(D.bounds.ensure(6 * 4));
        uint32_t binit[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
        RT_CUDA(cudaMemcpyAsync(D.bounds.p, binit, sizeof(binit), cudaMemcpyHostToDevice, st));
        // one CTA per SM-quarter: the grid-stride loop reduces in registers, so only ~5k warps touch the 6 atomics (was 31k: 0.13 ms)
        k_bounds<<<grid_for(n) < D.sm_count * 4 ? grid_for(n) : D.sm_count * 4, BT, 0, st>>>(R, n, D.bounds.as<uint32_t>());
        uint32_t bord[6];
        RT_CUDA(cudaMemcpyAsync(bord, D.bounds.p, sizeof(bord), cudaMemcpyDeviceToHost, st));
        RT_CUDA(cudaStreamSynchronize(st));
        float b[6]; for (int k = 0; k < 6; ++k) b[k] = f_unord(bord[k]);
        float amax = 0.f; for (int k = 0; k < 6; ++k) if (std::isfinite(b[k])) amax = fmaxf(amax, fabsf(b[k]));
        { float bo[6]; for (int k = 0; k < 6; ++k) bo[k] = b[k];          // origin box and padding: instances included; Morton grid: world primitives only
          IS.grow_bounds(bo, amax);
          const float pad0 = (float)(amax * opt.pad_scale) + 1e-30f;
          S.origin_reach = rt_origin_reach(pad0);
          set_obox(bo, amax); }
        const float pad = (float)(amax * opt.pad_scale) + 1e-30f;
        f3 bmin = mk3(b[0], b[1], b[2]);
        f3 inv_ext = mk3(b[3] > b[0] ? 1.0f / (b[3] - b[0]) : 0.f, b[4] > b[1] ? 1.0f / (b[4] - b[1]) : 0.f, b[5] > b[2] ? 1.0f / (b[5] - b[2]) : 0.f);
        // morton + sort
        RT_CUDA(D.keys.ensure((size_t)n * 8)); RT_CUDA(D.keys_alt.ensure((size_t)n * 8));
        RT_CUDA(D.vals.ensure((size_t)n * 4)); RT_CUDA(D.vals_alt.ensure((size_t)n * 4));
        k_morton<<<grid_for(n), BT, 0, st>>>(R, n, bmin, inv_ext, D.keys.as<uint64_t>(), D.vals.as<int32_t>());
        size_t tmp_bytes = 0;
        RT_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, D.keys.as<uint64_t>(), D.keys_alt.as<uint64_t>(), D.vals.as<int32_t>(),
                                                D.vals_alt.as<int32_t>(), n, 0, 63, st));
        size_t scan_bytes = 0;
        RT_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int32_t*)nullptr, (int32_t*)nullptr, n, st));
        RT_CUDA(D.cub_tmp.ensure(tmp_bytes > scan_bytes ? tmp_bytes : scan_bytes));
        RT_CUDA(cub::DeviceRadixSort::SortPairs(D.cub_tmp.p, tmp_bytes, D.keys.as<uint64_t>(), D.keys_alt.as<uint64_t>(), D.vals.as<int32_t>(),
                                                D.vals_alt.as<int32_t>(), n, 0, 63, st));
        const uint64_t* skeys = D.keys_alt.as<uint64_t>();
        const int32_t* svals = D.vals_alt.as<int32_t>();
        // slots
        RT_CUDA(D.prims.ensure((size_t)n * 48)); RT_CUDA(D.pinfo.ensure((size_t)n * 16));
        RT_CUDA(D.leaf_lo.ensure((size_t)n * 16)); RT_CUDA(D.leaf_hi.ensure((size_t)n * 16));
        k_emit_slots<<<grid_for(n), BT, 0, st>>>(R, n, svals, D.prims.as<float4>(), D.pinfo.as<float4>(), D.leaf_lo.as<float4>(), D.leaf_hi.as<float4>());
        S.prims = D.prims.as<float4>(); S.pinfo = D.pinfo.as<float4>();
        if (n <= leaf_max) {
            S.root = ~(0 * 8 + (n - 1));
            RT_CUDA(D.nodes.ensure(64));
            S.nodes = D.nodes.as<float4>();
        } else {
            const int32_t ni = n - 1;
            RT_CUDA(D.topo_i32.ensure(((size_t)ni * 5 + n) * 4));
            int32_t* tp = D.topo_i32.as<int32_t>();
            RtTopo T{tp, tp + ni, tp + 2 * (size_t)ni, tp + 3 * (size_t)ni, tp + 4 * (size_t)ni, tp + 5 * (size_t)ni};
            k_karras<<<grid_for(ni), BT, 0, st>>>(skeys, n, T);
            RT_CUDA(D.int_lo.ensure((size_t)ni * 16)); RT_CUDA(D.int_hi.ensure((size_t)ni * 16));
            RT_CUDA(D.flags.ensure((size_t)ni * 4)); RT_CUDA(D.new_index.ensure((size_t)ni * 4));
            RT_CUDA(cudaMemsetAsync(D.flags.p, 0, (size_t)ni * 4, st));
            k_refit<<<grid_for(n), BT, 0, st>>>(n, T, D.leaf_lo.as<float4>(), D.leaf_hi.as<float4>(), D.int_lo.as<float4>(), D.int_hi.as<float4>(),
                                                 D.flags.as<int32_t>());
            k_live<<<grid_for(ni), BT, 0, st>>>(ni, T, leaf_max, D.flags.as<int32_t>());
            RT_CUDA(cub::DeviceScan::ExclusiveSum(D.cub_tmp.p, scan_bytes, D.flags.as<int32_t>(), D.new_index.as<int32_t>(), ni, st));
            int32_t last_idx = 0, last_flag = 0;
            RT_CUDA(cudaMemcpyAsync(&last_idx, D.new_index.as<int32_t>() + (ni - 1), 4, cudaMemcpyDeviceToHost, st));
            RT_CUDA(cudaMemcpyAsync(&last_flag, D.flags.as<int32_t>() + (ni - 1), 4, cudaMemcpyDeviceToHost, st));
            RT_CUDA(cudaStreamSynchronize(st));
            D.n_nodes = (uint32_t)(last_idx + last_flag);
            RT_CUDA(D.nodes.ensure((size_t)D.n_nodes * 64));
            k_emit_nodes<<<grid_for(ni), BT, 0, st>>>(ni, T, D.new_index.as<int32_t>(), leaf_max, pad, D.leaf_lo.as<float4>(), D.leaf_hi.as<float4>(),
                                                      D.int_lo.as<float4>(), D.int_hi.as<float4>(), D.nodes.as<float4>());
            S.root = 0;
            S.nodes = D.nodes.as<float4>();
        }
        RT_CUDA(cudaGetLastError());
        { int rci = upload_instances(amax, pad); if (rci) return rci; }
        D.bvh_prims = n; D.bvh_slots = n; D.bvh_root = S.root; D.bvh_parents = false;
    }
    S.n_nodes = (int32_t)D.n_nodes;
            RT_CUDA(cudaEventRecord(D.ev1, st));
    RT_CUDA(cudaEventSynchronize(D.ev1));
    if (ms_build) RT_CUDA(cudaEventElapsedTime(ms_build, D.ev0, D.ev1));
    return RT_OK;
}

// rt_render.cu — the per-pixel hot path on sm_100a: persistent-thread wavefront kernels.
//
//   k_trace_primary : ray generation (CastPixelRays, RTS:275-298) fused with closest-hit BVH traversal
//   k_shade         : shading + light loop (Shade, RTS:304-366) with the shadow query traced inline as an any-hit
//                     traversal; appends mirror rays to the next depth's queue (warp-aggregated ballot/popc append)
//   k_trace_queue   : closest-hit traversal of a queued (mirror) ray wave
//   k_fold          : colour_{d-1} += km * colour_d, deepest first — the recursion's evaluation order (RTS:358-363)
//   k_finalize      : Rgb.Color (Rgb.cs:13) / u8 quantisation / primary hit id + distance, written straight into the
//                     gather device's full-frame buffers (local or NVLink peer memory)
// All traversal kernels are persistent: 148 x k resident CTAs fetch 32-item batches (one 8x4 pixel tile = one warp) from
// a global counter; the traversal stack lives in shared memory, strided so a warp's accesses are conflict-free.
// Compiled with -fmad=false: see rt_hd.h.
#include "rt_internal.h"

namespace {

constexpr int STK_CAP = 24;      // shared-memory stack entries per thread
constexpr int STK_OVF = 104;     // local-memory overflow entries (LBVH depth bound: 64 key bits + 32 index bits)

struct SmemStack {
    int* base; int stride; int sp;
    int ovf[STK_OVF];
    __device__ __forceinline__ SmemStack(int* b, int s) : base(b), stride(s), sp(0) {}
    __device__ __forceinline__ void push(int v) {
        if (sp < STK_CAP) base[sp * stride] = v;
        else if (sp - STK_CAP < STK_OVF) ovf[sp - STK_CAP] = v;
        ++sp;
    }
    __device__ __forceinline__ int pop() { --sp; return sp < STK_CAP ? base[sp * stride] : ovf[sp - STK_CAP < STK_OVF ? sp - STK_CAP : STK_OVF - 1]; }
    __device__ __forceinline__ bool empty() const { return sp == 0; }
};

extern __shared__ int g_smem_stack[];

__device__ __forceinline__ int warp_fetch(unsigned* counter, int lane) {
    int t = 0;
    if (lane == 0) t = (int)atomicAdd(counter, 1u);
    return __shfl_sync(0xffffffffu, t, 0);
}

// counters layout (unsigned): [0 .. D]: queue counts per depth; [C_WORK + 2*d]: trace work, [C_WORK + 2*d + 1]: shade work
constexpr int C_WORK = RT_MAX_DEPTH + 2;
constexpr int C_TOTAL = C_WORK + 2 * (RT_MAX_DEPTH + 2);

__global__ void __launch_bounds__(128) k_trace_primary(RtScene S, RtView V, int tile_begin, int n_tiles, int si, int sj,
                                                        float* __restrict__ hit_t, int* __restrict__ hit_slot, unsigned* work) {
    SmemStack stk(g_smem_stack + threadIdx.x, blockDim.x);
    const int lane = threadIdx.x & 31;
    for (;;) {
        int t = warp_fetch(work, lane);
        if (t >= n_tiles) break;
        int x, y;
        bool valid = rt_item_to_pixel(V, tile_begin + t, lane, x, y);
        float bt = FLT_MAX; int bs = -1;
        if (valid) {
            f3 O, D; rt_primary_ray(V, x, y, si, sj, O, D);
            if (rt_scene_gate(S, O, D)) { stk.sp = 0; rt_traverse<false>(S, O, D, 0.f, stk, bt, bs); }
        }
        size_t q = (size_t)t * 32 + lane;
        hit_t[q] = bt; hit_slot[q] = valid ? bs : -2;
    }
}

__global__ void __launch_bounds__(128) k_trace_queue(RtScene S, const unsigned* __restrict__ n_ptr, const float4* __restrict__ rays,
                                                      float* __restrict__ hit_t, int* __restrict__ hit_slot, unsigned* work) {
    SmemStack stk(g_smem_stack + threadIdx.x, blockDim.x);
    const int lane = threadIdx.x & 31;
    const unsigned n = *n_ptr;
    for (;;) {
        unsigned q = (unsigned)warp_fetch(work, lane) * 32u + lane;
        if (q - lane >= n) break;
        if (q < n) {
            float4 ro = rays[2 * (size_t)q], rd = rays[2 * (size_t)q + 1];
            f3 O = mk3(ro.x, ro.y, ro.z), D = mk3(rd.x, rd.y, rd.z);
            float bt = FLT_MAX; int bs = -1;
            if (rt_scene_gate(S, O, D)) { stk.sp = 0; rt_traverse<false>(S, O, D, 0.f, stk, bt, bs); }
            hit_t[q] = bt; hit_slot[q] = bs;
        }
    }
}

// DEPTH0: items are pixel samples of tiles [tile_begin, tile_begin+n_tiles); rays are regenerated, not stored.
template <bool DEPTH0>
__global__ void __launch_bounds__(128) k_shade(RtScene S, RtView V, int depth, int tile_begin, int n_tiles, int si, int sj,
                                                const unsigned* __restrict__ n_ptr, const float4* __restrict__ rays_in,
                                                const float* __restrict__ hit_t, const int* __restrict__ hit_slot, float4* __restrict__ local_out,
                                                float4* __restrict__ rays_out, float4* __restrict__ link_out, unsigned* next_count, unsigned* work,
                                                unsigned long long* stats) {
    SmemStack stk(g_smem_stack + threadIdx.x, blockDim.x);
    const int lane = threadIdx.x & 31;
    const unsigned n = DEPTH0 ? (unsigned)n_tiles * 32u : *n_ptr;
    unsigned my_shadow = 0;
    for (;;) {
        unsigned q = (unsigned)warp_fetch(work, lane) * 32u + lane;
        if (q - lane >= n) break;
        RtMirror mir; mir.emit = false;
        if (q < n) {
            int slot = hit_slot[q];
            if (slot == -1) local_out[q] = make_float4(V.bg255.x, V.bg255.y, V.bg255.z, 0.f);   // RTS:310-311
            else if (slot >= 0) {
                f3 O, D;
                if (DEPTH0) { int x, y; rt_item_to_pixel(V, tile_begin + (int)(q >> 5), lane, x, y); rt_primary_ray(V, x, y, si, sj, O, D); }
                else { float4 ro = rays_in[2 * (size_t)q], rd = rays_in[2 * (size_t)q + 1]; O = mk3(ro.x, ro.y, ro.z); D = mk3(rd.x, rd.y, rd.z); }
                f3 c = rt_shade_hit(S, O, D, hit_t[q], slot, depth, V.max_bounces, stk, mir, my_shadow);
                local_out[q] = make_float4(c.x, c.y, c.z, 0.f);
            }
        }
        // warp-aggregated append of the reflection rays to the next depth's queue
        unsigned m = __ballot_sync(0xffffffffu, mir.emit);
        if (m) {
            unsigned base = 0;
            if (lane == 0) { base = atomicAdd(next_count, (unsigned)__popc(m)); atomicAdd(stats + 2, (unsigned long long)__popc(m)); }
            base = __shfl_sync(0xffffffffu, base, 0);
            if (mir.emit) {
                size_t o = base + __popc(m & ((1u << lane) - 1u));
                rays_out[2 * o] = make_float4(mir.O.x, mir.O.y, mir.O.z, 0.f);
                rays_out[2 * o + 1] = make_float4(mir.D.x, mir.D.y, mir.D.z, 0.f);
                link_out[o] = make_float4(mir.km.x, mir.km.y, mir.km.z, __int_as_float((int)q));
            }
        }
    }
    my_shadow = __reduce_add_sync(0xffffffffu, my_shadow);
    if (lane == 0 && my_shadow) atomicAdd(stats + 1, (unsigned long long)my_shadow);
}

__global__ void k_fold(const unsigned* __restrict__ n_ptr, const float4* __restrict__ link, const float4* __restrict__ local_cur, float4* local_prev) {
    const unsigned n = *n_ptr;
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) rt_fold_item(q, link, local_cur, local_prev);
}

__global__ void k_finalize(RtScene S, RtView V, int tile_begin, int n_tiles, int pass, int n_pass, const float4* __restrict__ local0,
                           const float* __restrict__ hit_t, const int* __restrict__ hit_slot, float4* accum, RtOut out) {
    const unsigned n = (unsigned)n_tiles * 32u;
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x)
        rt_finalize_item(S, V, tile_begin, q, pass, n_pass, local0, hit_t, hit_slot, accum, out);
}

__global__ void __launch_bounds__(128) k_intersect(RtScene S, int n, const float* __restrict__ o, const float* __restrict__ d,
                                                    int32_t* __restrict__ out_id, float* __restrict__ out_t) {
    SmemStack stk(g_smem_stack + threadIdx.x, blockDim.x);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        f3 O = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), D = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        float bt = FLT_MAX; int bs = -1;
        if (rt_scene_gate(S, O, D)) { stk.sp = 0; rt_traverse<false>(S, O, D, 0.f, stk, bt, bs); }
        int32_t ty, ix, me; rt_slot_to_id(S, bs, ty, ix, me);
        out_id[3 * i] = ty; out_id[3 * i + 1] = ix; out_id[3 * i + 2] = me; out_t[i] = bt;
    }
}

template <class K>
int persistent_grid(K kernel, int block, size_t smem, int sm_count, int ctas_per_sm) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    if (ctas_per_sm > 0 && ctas_per_sm < per_sm) per_sm = ctas_per_sm;
    return sm_count * per_sm;
}

}  // namespace

int rt_device_render(RtDevice& D, const RtView& V_in, int want, const RtOptions& opt, std::string& err, RtFrameStats* fs) {
    RT_CUDA(cudaSetDevice(D.device));
    RtView V = V_in;
    {   // reflection depth actually needed: none without a mirror material (RTS:358); clamp to the buffer limit
        int md = (D.has_mirror && V.max_bounces > 0) ? (V.max_bounces > RT_MAX_DEPTH ? RT_MAX_DEPTH : V.max_bounces) : 0;
        V.max_bounces = md;
    }
    cudaStream_t st = D.stream;
    uint32_t launches = 0;
    RT_CUDA(D.stats.ensure(4 * 8));
    RT_CUDA(cudaMemsetAsync(D.stats.p, 0, 4 * 8, st));
    RT_CUDA(cudaEventRecord(D.ev0, st));
    const int g = V.spp_grid, n_pass = g * g;
    uint64_t primary = 0;
    if (V.n_tiles > 0 && V.res_x > 0) {
        const int block = 128;
        const size_t smem = (size_t)STK_CAP * block * sizeof(int);
        int wave_tiles = (opt.chunk_pixels / 32) < 1 ? 1 : (opt.chunk_pixels / 32);
        if (wave_tiles > V.n_tiles) wave_tiles = V.n_tiles;
        const size_t W = (size_t)wave_tiles * 32;
        const int maxd = V.max_bounces;
        // work buffers
        RT_CUDA(D.hit0_t.ensure(W * 4)); RT_CUDA(D.hit0_slot.ensure(W * 4));
        RT_CUDA(D.counters.ensure(C_TOTAL * 4));
        if ((int)D.local.size() < maxd + 1) D.local.resize(maxd + 1);
        if ((int)D.link.size() < maxd + 1) D.link.resize(maxd + 1);
        for (int d = 0; d <= maxd; ++d) RT_CUDA(D.local[d].ensure(W * 16));
        if (maxd > 0) {
            RT_CUDA(D.hitq_t.ensure(W * 4)); RT_CUDA(D.hitq_slot.ensure(W * 4));
            RT_CUDA(D.rays[0].ensure(W * 32)); RT_CUDA(D.rays[1].ensure(W * 32));
            for (int d = 1; d <= maxd; ++d) RT_CUDA(D.link[d].ensure(W * 16));
        }
        if (n_pass > 1) RT_CUDA(D.accum.ensure((size_t)V.n_tiles * 32 * 16));
        const int g_trace = persistent_grid(k_trace_primary, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_traceq = persistent_grid(k_trace_queue, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_shade0 = persistent_grid(k_shade<true>, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_shaden = persistent_grid(k_shade<false>, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_flat = D.sm_count * 8;
        unsigned* C = D.counters.as<unsigned>();
        unsigned long long* ST = D.stats.as<unsigned long long>();
        RtOut out{(want & RT_OUT_RGBA32F) ? (float4*)D.o_rgba32f : nullptr, (want & RT_OUT_RGBA8) ? (uint32_t*)D.o_rgba8 : nullptr,
                  (want & RT_OUT_HIT_ID) ? (int32_t*)D.o_hit_id : nullptr, (want & RT_OUT_HIT_T) ? (float*)D.o_hit_t : nullptr};
        for (int sj = 0, pass = 0; sj < g; ++sj)
            for (int si = 0; si < g; ++si, ++pass)
                for (int tb = 0; tb < V.n_tiles; tb += wave_tiles) {
                    const int nt = (V.n_tiles - tb) < wave_tiles ? (V.n_tiles - tb) : wave_tiles;
                    RT_CUDA(cudaMemsetAsync(C, 0, C_TOTAL * 4, st));
                    k_trace_primary<<<g_trace, block, smem, st>>>(D.S, V, tb, nt, si, sj, D.hit0_t.as<float>(), D.hit0_slot.as<int>(), C + C_WORK);
                    k_shade<true><<<g_shade0, block, smem, st>>>(D.S, V, 0, tb, nt, si, sj, nullptr, nullptr, D.hit0_t.as<float>(), D.hit0_slot.as<int>(),
                                                                  D.local[0].as<float4>(), maxd > 0 ? D.rays[0].as<float4>() : nullptr,
                                                                  maxd > 0 ? D.link[1].as<float4>() : nullptr, C + 1, C + C_WORK + 1, ST);
                    launches += 2;
                    for (int d = 1; d <= maxd; ++d) {
                        float4* rin = D.rays[(d - 1) & 1].as<float4>();
                        float4* rout = D.rays[d & 1].as<float4>();
                        k_trace_queue<<<g_traceq, block, smem, st>>>(D.S, C + d, rin, D.hitq_t.as<float>(), D.hitq_slot.as<int>(), C + C_WORK + 2 * d);
                        k_shade<false><<<g_shaden, block, smem, st>>>(D.S, V, d, 0, 0, 0, 0, C + d, rin, D.hitq_t.as<float>(), D.hitq_slot.as<int>(),
                                                                       D.local[d].as<float4>(), d < maxd ? rout : nullptr,
                                                                       d < maxd ? D.link[d + 1].as<float4>() : nullptr, C + d + 1, C + C_WORK + 2 * d + 1, ST);
                        launches += 2;
                    }
                    for (int d = maxd; d >= 1; --d) {
                        k_fold<<<g_flat, 256, 0, st>>>(C + d, D.link[d].as<float4>(), D.local[d].as<float4>(), D.local[d - 1].as<float4>());
                        ++launches;
                    }
                    k_finalize<<<g_flat, 256, 0, st>>>(D.S, V, tb, nt, pass, n_pass, D.local[0].as<float4>(), D.hit0_t.as<float>(), D.hit0_slot.as<int>(),
                                                        n_pass > 1 ? D.accum.as<float4>() : nullptr, out);
                    ++launches;
                }
        RT_CUDA(cudaGetLastError());
        primary = (uint64_t)V.res_x * (uint64_t)V.n_local_rows * (uint64_t)n_pass;
    }
    RT_CUDA(cudaEventRecord(D.ev1, st));
    if (fs) { fs->primary = primary; fs->shadow = 0; fs->mirror = 0; fs->launches = launches; }
    return RT_OK;
}

int rt_device_read_stats(RtDevice& D, RtFrameStats* fs, std::string& err) {
    RT_CUDA(cudaSetDevice(D.device));
    unsigned long long h[4] = {0, 0, 0, 0};
    if (D.stats.p) {
        RT_CUDA(cudaMemcpyAsync(h, D.stats.p, sizeof(h), cudaMemcpyDeviceToHost, D.stream));
        RT_CUDA(cudaStreamSynchronize(D.stream));
    }
    fs->shadow = h[1]; fs->mirror = h[2];
    return RT_OK;
}

int rt_device_intersect(RtDevice& D, const float* d_origins, const float* d_dirs, int n, int32_t* d_ids, float* d_t,
                        const RtOptions& opt, std::string& err) {
    RT_CUDA(cudaSetDevice(D.device));
    if (n <= 0) return RT_OK;
    const int block = 128;
    const size_t smem = (size_t)STK_CAP * block * sizeof(int);
    int grid = (n + block - 1) / block;
    if (grid > D.sm_count * 16) grid = D.sm_count * 16;
    k_intersect<<<grid, block, smem, D.stream>>>(D.S, n, d_origins, d_dirs, d_ids, d_t);
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

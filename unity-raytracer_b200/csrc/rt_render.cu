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
#include <mutex>
#include <utility>
#include <vector>
#include "rt_internal.h"
#include "rt_packet.cuh"

namespace {

// Resident 128-thread CTAs per SM the trace kernels are compiled for.  The traversal is a dependent pointer chase, so
// more resident warps hide more load latency: 8 -> 10 CTAs (64 -> 48 registers, a few spills outside the node loop)
// measured 2.66 -> 2.50 ms on cfg3; 12 (40 registers) helps only the lean primary kernel.
#ifndef RT_MIN_CTAS
#define RT_MIN_CTAS 10
#endif
#ifndef RT_MIN_CTAS_PRIMARY
#define RT_MIN_CTAS_PRIMARY 12
#endif
// the instantiations for scenes with instanced meshes (INST) carry the two-level state and bake triangles on the fly
#ifndef RT_MIN_CTAS_INST
#define RT_MIN_CTAS_INST 10
#endif
#ifndef RT_STK_CAP
#define RT_STK_CAP 12
#endif
constexpr int STK_CAP = RT_STK_CAP;   // shared-memory stack entries per thread
constexpr int STK_OVF = 104;     // local-memory overflow entries (LBVH depth bound: 64 key bits + 32 index bits)
static_assert(RT_STK_CAP != 12 || STK_CAP + STK_OVF == RT_STACK_ENTRIES, "rt_internal.h RT_STACK_ENTRIES is what the builders are checked against");

// sp lives in a register; the (rarely used) overflow array is a separate local object so that indexing it dynamically
// does not drag the whole stack object into local memory (ncu on the first version: one STL per push/pop for sp).
// The shared-memory column is addressed through a precomputed 32-bit shared-window address with explicit st/ld.shared:
// the generic-pointer form made the compiler rebuild the window base (S2R SR_CgaCtaId, LEA, 2 IMAD) on every push and pop.
// An entry is TWO words per thread: the reference and, for the packet traversal (rt_packet.cuh), the lane's own entry
// distance into the pushed box (+inf: this lane missed it) — one st/ld.shared.v2 per push/pop off a single address register
// (a separate warp-shared reference column needed a second address, which the compiler rematerialised on every node step at
// the 40/48-register occupancy cap: +8 instructions per step).  A popped entry that no lane can still reach (entry distance >
// the lane's current cull distance; lanes that found their occluder have cull distance -1) is dropped without fetching it:
// rt_packet_pop.  The per-lane traversal uses the reference word only.
constexpr int RT_BLOCK = 128;         // threads per CTA of every kernel that owns a traversal stack
struct SmemStack {
    unsigned base; int sp; int* ovf; float* ovfd;
    static constexpr unsigned stride = (unsigned)RT_BLOCK * 8u;
    __device__ __forceinline__ SmemStack(int* b, int* o, float* od) : base((unsigned)__cvta_generic_to_shared(b)), sp(0), ovf(o), ovfd(od) {
        // opaque to the compiler: at the register cap it otherwise REBUILDS the address (S2UR CgaCtaId, ULEA, S2R tid, LEA) on
        // every node step instead of keeping one register live
        asm volatile("" : "+r"(base));
    }
    __device__ __forceinline__ void push(int v) {
        RT_CHECK(RT_CHK_STACK, sp >= 0 && sp < STK_CAP + STK_OVF);
        if (sp < STK_CAP) asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + (unsigned)sp * stride), "r"(v) : "memory");
        else if (sp - STK_CAP < STK_OVF) ovf[sp - STK_CAP] = v;
        ++sp;
    }
    __device__ __forceinline__ int pop() {
        --sp;
        int v;
        if (sp < STK_CAP) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(base + (unsigned)sp * stride) : "memory");
        else v = ovf[sp - STK_CAP < STK_OVF ? sp - STK_CAP : STK_OVF - 1];
        return v;
    }
    __device__ __forceinline__ bool empty() const { return sp == 0; }
    __device__ __forceinline__ void push_packet(int v, float dist) {
        RT_CHECK(RT_CHK_STACK, sp >= 0 && sp < STK_CAP + STK_OVF);
        if (sp < STK_CAP) asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(base + (unsigned)sp * stride), "r"(v), "r"(__float_as_int(dist)) : "memory");
        else if (sp - STK_CAP < STK_OVF) { ovf[sp - STK_CAP] = v; ovfd[sp - STK_CAP] = dist; }
        ++sp;
    }
    __device__ __forceinline__ int pop_packet(float& dist) {
        --sp;
        int v, d;
        if (sp < STK_CAP) asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v), "=r"(d) : "r"(base + (unsigned)sp * stride) : "memory");
        else { int i = sp - STK_CAP < STK_OVF ? sp - STK_CAP : STK_OVF - 1; v = ovf[i]; d = __float_as_int(ovfd[i]); }
        dist = __int_as_float(d);
        return v;
    }
};
constexpr size_t RT_STACK_SMEM = (size_t)STK_CAP * RT_BLOCK * 8;
// INST instantiations: behind the stacks, RT_BAKE_F4 float4 per warp for the vertices of the BLAS leaf the warp bakes (rt_packet.cuh)
constexpr size_t RT_BAKE_SMEM = (size_t)(RT_BLOCK / 32) * RT_BAKE_F4 * 16;
#define RT_WARP_BAKE() (reinterpret_cast<float4*>(g_smem_stack + RT_STACK_SMEM / 4) + (threadIdx.x >> 5) * RT_BAKE_F4)
#define RT_DECL_STACK(name) int name##_ovf[STK_OVF]; float name##_ovfd[STK_OVF]; \
    SmemStack name(g_smem_stack + 2 * threadIdx.x, name##_ovf, name##_ovfd)

extern __shared__ int g_smem_stack[];

__device__ __forceinline__ int warp_fetch(unsigned* counter, int lane) {
    int t = 0;
    if (lane == 0) t = (int)atomicAdd(counter, 1u);
    return __shfl_sync(0xffffffffu, t, 0);
}

// ---- heavy-first tile order ("cost hints") ----------------------------------------------------------------------------
// Tile costs are extremely skewed (cfg3, ncu/clock64: median 2.5 us, p99 70 us, max 190 us — silhouette tiles whose rays
// graze the displaced shell), so a persistent kernel that hands tiles out in image order ends with a long tail: the last
// heavy tiles start late and the other warps idle (17 % of k_trace_primary on the full 4K frame, > 50 % on a 1/8 band
// partition).  Each tile-indexed trace kernel therefore measures every tile (clock64) and files it under one of 8 cost
// classes (log2 of its cycles); the NEXT frame hands the tiles of the same wave out class by class, most expensive first
// (longest-processing-time-first list scheduling), falling back to image order when no hints exist (first frame, changed
// resolution / partition / wave layout).  Hints only reorder work: results are unaffected, stale hints are harmless.
constexpr int RT_NCLS = 8;
// lists[c * stride + i], counts[c]; null = image order.  lsplit: cost classes >= lsplit are handed out as one item per light
// (fused depth-0 shadow kernel only; RT_NCLS = off)
struct RtHintIn { const int* lists; const unsigned* counts; size_t stride; int lsplit; };
struct RtHintOut { int* lists; unsigned* counts; size_t stride; };                // null = do not record

// (tile times are taken with the 32-bit cycle counter: a tile lasts < 1 ms, differences are wrap-safe, and the start value
// occupies one register across the traversal instead of two)
__device__ __forceinline__ int rt_cost_class(unsigned cycles) {
    int c = 31 - __clz((int)((cycles >> 11) | 1u));             // < 2K cycles: 0, 2-4K: 1, ... >= 256K: 7
    return c > RT_NCLS - 1 ? RT_NCLS - 1 : c;
}
// k-th work item of the wave -> tile index within the wave, heaviest class first
__device__ __forceinline__ int rt_hint_tile(const RtHintIn& h, int k, int lane) {
    if (!h.lists) return k;
    unsigned cnt = lane < RT_NCLS ? h.counts[RT_NCLS - 1 - lane] : 0u;          // lane 0 holds the heaviest class
    unsigned incl = cnt;
    for (int o = 1; o < RT_NCLS; o <<= 1) { unsigned v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
    unsigned m = __ballot_sync(0xffffffffu, lane < RT_NCLS && (unsigned)k < incl);
    int src = __ffs(m) - 1;                                // first class whose inclusive prefix exceeds k
    if (src < 0) return k;                                 // inconsistent hints: cannot happen when the host validated them
    unsigned excl = __shfl_sync(0xffffffffu, incl - cnt, src);
    return h.lists[(size_t)(RT_NCLS - 1 - src) * h.stride + ((unsigned)k - excl)];
}
// Per-light items for heavy tiles.  On a small partition a trace kernel lasts as long as its heaviest tile (a dependent chain
// that no amount of idle warps shortens), and the fused shadow kernel walks a tile's lights one after the other: cfg3's
// heaviest tiles take 2 x 150 us.  The shadow queries of different lights are independent, so tiles of the classes >=
// h.lsplit become n_l work items (tile, light) that run on n_l warps; each ORs its occlusion bit into a per-item word and the
// LAST item of a tile to finish (per-tile arrival counter) shades and writes the pixel with the combined mask — exactly the
// mask the single-warp loop would have built.  Each extra item regenerates the ray and the surface point and the lights of
// a tile no longer share their cached nodes: 4 % on a full frame, so the host enables the split only on partitions of a
// quarter frame or less (RtOptions::light_split), where the critical path is what counts (shadow kernel -5..11 %).
// Returns false when k is beyond the last item; light = -1: all lights in this item.
__device__ __forceinline__ bool rt_hint_item_lights(const RtHintIn& h, int k, int lane, int n_tiles, int n_l, int& tile, int& light) {
    light = -1;
    if (!h.lists) { tile = k; return k < n_tiles; }
    const int cls = RT_NCLS - 1 - lane;                                          // lane 0 holds the heaviest class
    const unsigned mult = (lane < RT_NCLS && cls >= h.lsplit && n_l > 1) ? (unsigned)n_l : 1u;
    const unsigned cnt = lane < RT_NCLS ? h.counts[cls] * mult : 0u;
    unsigned incl = cnt;
    for (int o = 1; o < RT_NCLS; o <<= 1) { unsigned v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
    const unsigned m = __ballot_sync(0xffffffffu, lane < RT_NCLS && (unsigned)k < incl);
    const int src = __ffs(m) - 1;                                                // first class whose inclusive prefix exceeds k
    if (src < 0) return false;                                                   // k >= number of items
    unsigned j = (unsigned)k - __shfl_sync(0xffffffffu, incl - cnt, src);
    if (__shfl_sync(0xffffffffu, mult, src) > 1u) { light = (int)(j % (unsigned)n_l); j /= (unsigned)n_l; }
    tile = h.lists[(size_t)(RT_NCLS - 1 - src) * h.stride + j];
    return true;
}
__device__ __forceinline__ void rt_hint_record(const RtHintOut& h, int tile, unsigned cycles, int lane) {
    if (!h.lists || lane != 0) return;
    int c = rt_cost_class(cycles);
    unsigned pos = atomicAdd(h.counts + c, 1u);
    h.lists[(size_t)c * h.stride + pos] = tile;
}

// counters layout (unsigned): [0 .. D]: queue counts per depth; work counters of depth d at C_WORK + 8*d:
//   +0 closest-hit trace, +1..+7 shadow-trace launches (one per chunk of 32 lights)
constexpr int C_WORK = RT_MAX_DEPTH + 2;
constexpr int C_TOTAL = C_WORK + 8 * (RT_MAX_DEPTH + 2);

#ifdef RT_TILE_TIMING
__device__ unsigned long long* g_rt_tile_dbg = nullptr;      // per tile: (start ns, duration ns); development aid only
#endif
// FAR (RtView::far_origin, a camera many scene extents away): per-lane loop with ray-dependent node padding (rt_far_pad) —
// a kernel of its own so that its registers do not count against the packet loop's
template <bool FAR, bool INST>
__global__ void __launch_bounds__(128, INST ? RT_MIN_CTAS_INST : RT_MIN_CTAS_PRIMARY) k_trace_primary(RtScene S, RtView V, int tile_begin, int n_tiles, int si, int sj,
                                                        float* __restrict__ hit_t, int* __restrict__ hit_slot, unsigned* work,
                                                        RtHintIn hin, RtHintOut hout) {
    RT_DECL_STACK(stk);
    const int lane = threadIdx.x & 31;
    for (;;) {
        int t = warp_fetch(work, lane);
        if (t >= n_tiles) break;
        t = rt_hint_tile(hin, t, lane);
        RT_CHECK(RT_CHK_TILE, t >= 0 && t < n_tiles);
        const unsigned c0 = (unsigned)clock();
#ifdef RT_TILE_TIMING
        unsigned long long dbg_t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
#endif
        int x, y;
        bool valid = rt_item_to_pixel(V, tile_begin + t, lane, x, y);
        float bt = FLT_MAX; int bs = -1;
        f3 O = V.pos, D = V.fwd;
        bool active = false;
        if (valid) rt_primary_ray(V, x, y, si, sj, O, D);
        const f3 inv = rt_rcp3(D);
        if (valid) active = rt_scene_gate(S, O, inv);                                               // Scene.cs:54
        if (FAR) { if (active) rt_traverse<false, true, INST>(S, O, D, inv, 0.f, stk, bt, bs); }
        else if (S.packet) {              // rays of one 8x4 pixel tile share the origin and are almost parallel
            if (__any_sync(0xffffffffu, active)) rt_traverse_packet<false, INST>(S, O, D, inv, active, 0.f, stk, bt, bs, INST ? RT_WARP_BAKE() : nullptr);
        } else if (active) rt_traverse<false, false, INST>(S, O, D, inv, 0.f, stk, bt, bs);
        size_t q = (size_t)t * 32 + lane;
        RT_ST_STREAM(hit_t + q, bt); RT_ST_STREAM(hit_slot + q, valid ? bs : -2);
        rt_hint_record(hout, t, (unsigned)clock() - c0, lane);
#ifdef RT_TILE_TIMING
        unsigned long long dbg_t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t1));
        if (lane == 0 && g_rt_tile_dbg) { g_rt_tile_dbg[2 * (size_t)(tile_begin + t)] = dbg_t0; g_rt_tile_dbg[2 * (size_t)(tile_begin + t) + 1] = dbg_t1 - dbg_t0; }
#endif
    }
}

template <bool INST>
__global__ void __launch_bounds__(128, INST ? RT_MIN_CTAS_INST : RT_MIN_CTAS) k_trace_queue(RtScene S, const unsigned* __restrict__ n_ptr, const float4* __restrict__ rays,
                                                      float* __restrict__ hit_t, int* __restrict__ hit_slot, unsigned* work) {
    RT_DECL_STACK(stk);
    const int lane = threadIdx.x & 31;
    const unsigned n = *n_ptr;
    for (;;) {
        unsigned q = (unsigned)warp_fetch(work, lane) * 32u + lane;
        if (q - lane >= n) break;
        f3 O = mk3(0.f, 0.f, 0.f), D = mk3(0.f, 0.f, 1.f);
        bool active = false;
        if (q < n) {
            float4 ro = rays[2 * (size_t)q], rd = rays[2 * (size_t)q + 1];
            O = mk3(ro.x, ro.y, ro.z); D = mk3(rd.x, rd.y, rd.z);
        }
        const f3 inv = rt_rcp3(D);
        if (q < n) active = rt_scene_gate(S, O, inv);
        float bt = FLT_MAX; int bs = -1;
        if (S.packet && rt_warp_coherent(O, D, active, S.coherent_spread2)) {
            if (__any_sync(0xffffffffu, active)) rt_traverse_packet<false, INST>(S, O, D, inv, active, 0.f, stk, bt, bs, INST ? RT_WARP_BAKE() : nullptr);
        } else if (active) rt_traverse<false, false, INST>(S, O, D, inv, 0.f, stk, bt, bs);
        if (q < n) { hit_t[q] = bt; hit_slot[q] = bs; }
    }
}

// item -> the ray that produced its hit. DEPTH0: regenerated from the pixel (never stored); else read from the queue.
template <bool DEPTH0>
__device__ __forceinline__ void item_ray(const RtView& V, int tile_begin, unsigned q, int lane, int si, int sj,
                                         const float4* __restrict__ rays_in, f3& O, f3& D) {
    if (DEPTH0) { int x, y; rt_item_to_pixel(V, tile_begin + (int)(q >> 5), lane, x, y); rt_primary_ray(V, x, y, si, sj, O, D); }
    else { float4 ro = rays_in[2 * (size_t)q], rd = rays_in[2 * (size_t)q + 1]; O = mk3(ro.x, ro.y, ro.z); D = mk3(rd.x, rd.y, rd.z); }
}

// Shadow queries (RTS:329-345) of every hit of a wave for lights [light_begin, light_begin+n_l), n_l <= 32, as any-hit
// traversals; bit i of occl[q] = light light_begin+i is occluded.  The light loop is warp-uniform so the 32 queries towards
// one light go as one packet when the warp's hits lie close together.
// FUSE (scenes with <= 32 lights): the rest of Shade (RTS:321-362: ambient, diffuse, Blinn-Phong, mirror-ray emission) runs
// right here once the occlusion mask is known, and with `fin` set (no deeper wave pending) so does the pixel write — the hit
// record, surface point and mask never travel through memory and two launches per depth disappear.
struct RtFuse {
    int depth;                         // bounce index of this wave
    float4* local_out;                 // colour of this bounce per item
    float4* rays_out; float4* link_out; unsigned* next_count;      // mirror-ray queue of the next depth (null at max depth)
    int fin;                           // 1: also finalize (depth 0 wave, no mirror depth above it)
    int pass, n_pass; float4* accum; RtOut out;
    uint32_t* split_occl;              // per item: occlusion bits of a tile split by light (zero between uses)
    unsigned long long* split_acc;     // per tile: (cycles << 6 | arrivals) of its per-light items (zero between uses)
    unsigned cap;                      // capacity of the queue buffers in entries (read only by the RT_CHECKED build)
};

// (A "mega" variant that also ran k_trace_primary's body here, camera ray to final colour in registers, was measured
// SLOWER on cfg3 — 2.49 -> 3.06 ms, instruction-cache pressure of the two traversal instantiations in one loop — and removed.)
template <bool DEPTH0, bool FUSE, bool INST>
__global__ void __launch_bounds__(128, INST ? RT_MIN_CTAS_INST : RT_MIN_CTAS) k_trace_shadow(RtScene S, RtView V, int tile_begin, int n_tiles, int si, int sj,
                                                          const unsigned* __restrict__ n_ptr, const float4* __restrict__ rays_in,
                                                          const float* __restrict__ hit_t, const int* __restrict__ hit_slot,
                                                          int light_begin, int n_l, uint32_t* __restrict__ occl, unsigned* work,
                                                          unsigned long long* stats, RtFuse F, RtHintIn hin, RtHintOut hout) {
    RT_DECL_STACK(stk);
    const int lane = threadIdx.x & 31;
    const unsigned full = 0xffffffffu;
    const unsigned n = DEPTH0 ? (unsigned)n_tiles * 32u : *n_ptr;
    unsigned my_shadow = 0, my_mirror = 0, my_skipped = 0;
    for (;;) {
        int wt = warp_fetch(work, lane), one_light = -1;
        if (DEPTH0 && FUSE) { if (!rt_hint_item_lights(hin, wt, lane, n_tiles, n_l, wt, one_light)) break; }
        else {
            if ((unsigned)wt * 32u >= n) break;
            if (DEPTH0) wt = rt_hint_tile(hin, wt, lane);
        }
        RT_CHECK(RT_CHK_TILE, wt >= 0 && (unsigned)wt * 32u < n + 32u);
        unsigned c0 = (unsigned)clock();
        const unsigned q = (unsigned)wt * 32u + lane;
        f3 O = mk3(0.f, 0.f, 0.f), D = mk3(0.f, 0.f, 1.f);
        float ht = 0.f;
        int slot;
        slot = q < n ? RT_LD_STREAM(hit_slot + q) : -2;
        RT_CHECK(RT_CHK_HIT_SLOT, slot >= -2 && slot < S.n_slots + (INST && S.n_inst > 0 ? S.inst_vbase[S.n_inst] : 0));
        const bool has = slot >= 0;
        uint32_t mask = 0;
        if (__any_sync(full, has)) {
            f3 P = mk3(0.f, 0.f, 0.f), sO = P, N = mk3(0.f, 0.f, 1.f);
            if (has) {
                int mat_id;
                item_ray<DEPTH0>(V, tile_begin, q, lane, si, sj, rays_in, O, D);
                ht = RT_LD_STREAM(hit_t + q);
                rt_surf_geom<INST>(S, O, D, ht, slot, P, N, mat_id);
                sO = P + N * RT_SHADOW_EPS;                                                        // RTS:331
            }
            const int i0 = one_light >= 0 ? one_light : 0, i1 = one_light >= 0 ? one_light + 1 : n_l;
            for (int i = i0; i < i1; ++i) {
                f3 Ldir = mk3(0.f, 0.f, 1.f); float d2 = 0.f;
                bool act = false;
                bool occluded = false;
                if (has) {
                    rt_light_ray(S, P, light_begin + i, Ldir, d2); ++my_shadow;
                    occluded = rt_shadow_irrelevant(S, Ldir, N, d2);                                   // light behind the surface
                    my_skipped += occluded ? 1u : 0u;
                }
                const f3 inv = rt_rcp3(Ldir);
                if (has && !occluded) act = rt_scene_gate(S, sO, inv);                                 // Scene.cs:54
                float st; int ss;
                if (S.packet && rt_warp_coherent(sO, Ldir, act, S.coherent_spread2)) {
                    if (__any_sync(full, act) && rt_traverse_packet<true, INST>(S, sO, Ldir, inv, act, d2, stk, st, ss, INST ? RT_WARP_BAKE() : nullptr)) occluded = true;
                } else if (act) occluded = rt_traverse<true, false, INST>(S, sO, Ldir, inv, d2, stk, st, ss);
                if (occluded) mask |= 1u << i;
            }
        }
        if (!FUSE) { if (q < n) occl[q] = mask; if (DEPTH0) rt_hint_record(hout, wt, (unsigned)clock() - c0, lane); continue; }
        if (DEPTH0 && one_light >= 0) {                        // one of the n_l items of a tile split by light
            if (mask) atomicOr(F.split_occl + q, mask);
            __threadfence();
            __syncwarp(full);
            unsigned long long old = 0ull;
            if (lane == 0) old = atomicAdd(F.split_acc + wt, ((unsigned long long)((unsigned)clock() - c0) << 6) | 1ull);
            old = __shfl_sync(full, old, 0);
            if ((int)(old & 63ull) != n_l - 1) continue;        // another item finishes the tile
            __threadfence();
            mask = atomicExch(F.split_occl + q, 0u);            // everybody's bits; leaves the word zero for the next use
            if (lane == 0) F.split_acc[wt] = 0ull;
            c0 -= (unsigned)(old >> 6);                         // the tile's cost = the sum over its items
        }
        // ---- fused shading (k_shade's body) ----
        RtMirror mir; mir.emit = false;
        float4 col = make_float4(V.bg255.x, V.bg255.y, V.bg255.z, 0.f);                            // RTS:310-311
        if (has) {
            f3 c = rt_shade_masked<INST>(S, O, D, ht, slot, F.depth, V.max_bounces, &mask, 0, 0, mir);
            col = make_float4(c.x, c.y, c.z, 0.f);
        }
        if (!(DEPTH0 && F.fin) && slot >= -1) RT_ST_STREAM(F.local_out + q, col);
        unsigned m = __ballot_sync(full, mir.emit);
        if (m) {
            unsigned base = 0;
            if (lane == 0) { base = atomicAdd(F.next_count, (unsigned)__popc(m)); my_mirror += (unsigned)__popc(m); }
            base = __shfl_sync(full, base, 0);
            if (mir.emit) {
                size_t o = base + __popc(m & ((1u << lane) - 1u));
                RT_CHECK(RT_CHK_QUEUE, o < (size_t)F.cap);
                F.rays_out[2 * o] = make_float4(mir.O.x, mir.O.y, mir.O.z, 0.f);
                F.rays_out[2 * o + 1] = make_float4(mir.D.x, mir.D.y, mir.D.z, 0.f);
                F.link_out[o] = make_float4(mir.km.x, mir.km.y, mir.km.z, __int_as_float((int)q));
            }
        }
        if (DEPTH0 && F.fin && slot >= -1) rt_finalize_value<INST>(S, V, tile_begin, q, F.pass, F.n_pass, col, has ? ht : FLT_MAX, slot, F.accum, F.out);
        if (DEPTH0) rt_hint_record(hout, wt, (unsigned)clock() - c0, lane);
#ifdef RT_TILE_TIMING
        // shadow-kernel tile costs (cycles; per-light items of a split tile add up) behind the primary kernel's records
        if (DEPTH0 && lane == 0 && g_rt_tile_dbg) atomicAdd(g_rt_tile_dbg + 2 * (size_t)V.n_tiles + (size_t)(tile_begin + wt), (unsigned long long)((unsigned)clock() - c0));
#endif
    }
    my_shadow = __reduce_add_sync(full, my_shadow);
    if (lane == 0 && my_shadow) atomicAdd(stats + 1, (unsigned long long)my_shadow);
    my_skipped = __reduce_add_sync(full, my_skipped);
    if (lane == 0 && my_skipped) atomicAdd(stats + 3, (unsigned long long)my_skipped);
    if (FUSE && lane == 0 && my_mirror) atomicAdd(stats + 2, (unsigned long long)my_mirror);
}

// Shading + light loop with the shadow results known (rt_shade_masked); appends the reflection rays of mirror hits to
// the next depth's queue with a warp-aggregated (ballot/popc) reservation.  Pure streaming kernel: no traversal.
template <bool DEPTH0, bool INST>
__global__ void __launch_bounds__(256) k_shade(RtScene S, RtView V, int depth, int tile_begin, int n_tiles, int si, int sj,
                                                const unsigned* __restrict__ n_ptr, const float4* __restrict__ rays_in,
                                                const float* __restrict__ hit_t, const int* __restrict__ hit_slot,
                                                const uint32_t* __restrict__ occl, size_t occl_stride, float4* __restrict__ local_out,
                                                float4* __restrict__ rays_out, float4* __restrict__ link_out, unsigned* next_count,
                                                unsigned long long* stats) {
    const int lane = threadIdx.x & 31;
    const unsigned full = 0xffffffffu;
    const unsigned n = DEPTH0 ? (unsigned)n_tiles * 32u : *n_ptr;
    const unsigned n32 = (n + 31u) & ~31u;                       // whole warps, so the ballots below are convergent
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n32; q += gridDim.x * blockDim.x) {
        RtMirror mir; mir.emit = false;
        const int slot = q < n ? hit_slot[q] : -2;
        if (slot == -1) local_out[q] = make_float4(V.bg255.x, V.bg255.y, V.bg255.z, 0.f);       // RTS:310-311
        else if (slot >= 0) {
            f3 O, D;
            item_ray<DEPTH0>(V, tile_begin, q, lane, si, sj, rays_in, O, D);
            f3 c = rt_shade_masked<INST>(S, O, D, hit_t[q], slot, depth, V.max_bounces, occl, occl_stride, q, mir);
            local_out[q] = make_float4(c.x, c.y, c.z, 0.f);
        }
        unsigned m = __ballot_sync(full, mir.emit);
        if (m) {
            unsigned base = 0;
            if (lane == 0) { base = atomicAdd(next_count, (unsigned)__popc(m)); atomicAdd(stats + 2, (unsigned long long)__popc(m)); }
            base = __shfl_sync(full, base, 0);
            if (mir.emit) {
                size_t o = base + __popc(m & ((1u << lane) - 1u));
                rays_out[2 * o] = make_float4(mir.O.x, mir.O.y, mir.O.z, 0.f);
                rays_out[2 * o + 1] = make_float4(mir.D.x, mir.D.y, mir.D.z, 0.f);
                link_out[o] = make_float4(mir.km.x, mir.km.y, mir.km.z, __int_as_float((int)q));
            }
        }
    }
}

__global__ void k_fold(const unsigned* __restrict__ n_ptr, const float4* __restrict__ link, const float4* __restrict__ local_cur, float4* local_prev) {
    const unsigned n = *n_ptr;
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) rt_fold_item(q, link, local_cur, local_prev);
}

template <bool INST>
__global__ void k_finalize(RtScene S, RtView V, int tile_begin, int n_tiles, int pass, int n_pass, const float4* __restrict__ local0,
                           const float* __restrict__ hit_t, const int* __restrict__ hit_slot, float4* accum, RtOut out) {
    const unsigned n = (unsigned)n_tiles * 32u;
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x)
        rt_finalize_item<INST>(S, V, tile_begin, q, pass, n_pass, local0, hit_t, hit_slot, accum, out);
}

template <bool INST>
__global__ void __launch_bounds__(128) k_intersect(RtScene S, int n, const float* __restrict__ o, const float* __restrict__ d,
                                                    int32_t* __restrict__ out_id, float* __restrict__ out_t) {
    RT_DECL_STACK(stk);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        f3 O = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), D = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        float bt = FLT_MAX; int bs = -1;
        const f3 inv = rt_rcp3(D);
        if (rt_scene_gate(S, O, inv)) {
            // explicit rays come from anywhere: always the ray-padded walk.  A direction that is not normalised in binary32
            // breaks what the sphere boxes assume (rt_sphere_box_radius), so such a ray scans the primitives like the reference.
            const float dd = um_dot(D, D);
            if (S.n_spheres > 0 && !(fabsf(dd - 1.0f) <= 0x1p-20f)) rt_scan_all<INST>(S, O, D, inv, bt, bs);
            else { stk.sp = 0; rt_traverse<false, true, INST>(S, O, D, inv, 0.f, stk, bt, bs); }
        }
        int32_t ty, ix, me; rt_slot_to_id<INST>(S, bs, ty, ix, me);
        out_id[3 * i] = ty; out_id[3 * i + 1] = ix; out_id[3 * i + 2] = me; out_t[i] = bt;
    }
}

// ---- re-ordering of a mirror-ray queue (option "sort_queue") -------------------------------------------------------------------
// Mirror rays leave the shading kernel in the order their warps happened to reserve queue space: neighbours in the queue start
// anywhere in the scene and point anywhere (cfg5: reflections off 10 000 small mirror spheres), so the 32 lanes of a queue warp
// walk 32 unrelated paths — the per-lane loop at a third of its lanes, every node fetch its own L1 wavefront (17 M mirror rays
// took 80 ms, 0.2 Grays/s, against 7 Grays/s for primary rays).  Before a depth is traced its queue is therefore counting-sorted
// by (direction bin, origin cell): 6 cube-map faces x 4 x 4 direction bins (major key: the lanes of a warp then share their
// octant and mostly their path) and an 8^3 grid over Scene.AABB in Morton order (minor key).  A queue entry is (ray, link) and
// the link carries the parent index (rt_fold_item), each parent has at most one reflection ray, and entries are independent:
// their order can change nothing but the time.  Three small kernels per depth and wave; the sorted copy lives in the queue
// buffer the previous depth no longer needs.
constexpr int RT_QS_CELLS = 8, RT_QS_DIRS = 96;
constexpr int RT_QS_BINS = RT_QS_DIRS * RT_QS_CELLS * RT_QS_CELLS * RT_QS_CELLS;       // 49 152
struct RtQsGrid { f3 lo, scale; };
__device__ __forceinline__ unsigned rt_qs_bin(const RtQsGrid& G, float4 ro, float4 rd) {
    const float ax = fabsf(rd.x), ay = fabsf(rd.y), az = fabsf(rd.z);
    int face; float m, u, v;
    if (ax >= ay && ax >= az) { face = rd.x < 0.f ? 1 : 0; m = ax; u = rd.y; v = rd.z; }
    else if (ay >= az) { face = rd.y < 0.f ? 3 : 2; m = ay; u = rd.x; v = rd.z; }
    else { face = rd.z < 0.f ? 5 : 4; m = az; u = rd.x; v = rd.y; }
    const float im = m > 0.f ? __frcp_rn(m) : 0.f;
    int iu = (int)fminf(fmaxf((u * im + 1.0f) * 2.0f, 0.0f), 3.0f), iv = (int)fminf(fmaxf((v * im + 1.0f) * 2.0f, 0.0f), 3.0f);   // NaN -> 0
    const unsigned cx = (unsigned)fminf(fmaxf((ro.x - G.lo.x) * G.scale.x, 0.0f), (float)(RT_QS_CELLS - 1));
    const unsigned cy = (unsigned)fminf(fmaxf((ro.y - G.lo.y) * G.scale.y, 0.0f), (float)(RT_QS_CELLS - 1));
    const unsigned cz = (unsigned)fminf(fmaxf((ro.z - G.lo.z) * G.scale.z, 0.0f), (float)(RT_QS_CELLS - 1));
    unsigned mo = 0;
#pragma unroll
    for (int b = 0; b < 3; ++b) mo |= (((cx >> b) & 1u) << (3 * b)) | (((cy >> b) & 1u) << (3 * b + 1)) | (((cz >> b) & 1u) << (3 * b + 2));
    return (unsigned)(face * 16 + iu * 4 + iv) * (unsigned)(RT_QS_CELLS * RT_QS_CELLS * RT_QS_CELLS) + mo;
}
__global__ void k_qs_hist(const unsigned* __restrict__ n_ptr, const float4* __restrict__ rays, RtQsGrid G, unsigned* hist) {
    const unsigned n = *n_ptr;
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x)
        atomicAdd(hist + rt_qs_bin(G, rays[2 * (size_t)q], rays[2 * (size_t)q + 1]), 1u);
}
// exclusive scan of the RT_QS_BINS counters in place: one CTA of 1024 threads walks the array in coalesced tiles of 1024
// (warp-shuffle scan inside a warp, 32 warp totals through shared memory, running carry between tiles)
__global__ void __launch_bounds__(1024) k_qs_scan(unsigned* hist) {
    __shared__ unsigned wsum[32];
    __shared__ unsigned carry;
    static_assert(RT_QS_BINS % 1024 == 0, "whole tiles");
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0u;
    __syncthreads();
    for (int tile = 0; tile < RT_QS_BINS / 1024; ++tile) {
        const unsigned v = hist[tile * 1024 + tid];
        unsigned incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned n = __shfl_up_sync(full, incl, o); if (lane >= o) incl += n; }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const unsigned w = wsum[lane];
            unsigned wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned n = __shfl_up_sync(full, wi, o); if (lane >= o) wi += n; }
            wsum[lane] = wi - w;                                // exclusive prefix of the warp totals
        }
        __syncthreads();
        const unsigned base = carry + wsum[warp];
        hist[tile * 1024 + tid] = base + incl - v;
        __syncthreads();
        if (tid == 1023) carry = base + incl;
        __syncthreads();
    }
}
__global__ void k_qs_scatter(const unsigned* __restrict__ n_ptr, const float4* __restrict__ rays_in, const float4* __restrict__ link_in, RtQsGrid G,
                             unsigned* offsets, float4* __restrict__ rays_out, float4* __restrict__ link_out) {
    const unsigned n = *n_ptr;
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        const float4 ro = rays_in[2 * (size_t)q], rd = rays_in[2 * (size_t)q + 1];
        const unsigned pos = atomicAdd(offsets + rt_qs_bin(G, ro, rd), 1u);
        rays_out[2 * (size_t)pos] = ro; rays_out[2 * (size_t)pos + 1] = rd;
        link_out[pos] = link_in[q];
    }
}

// Frame prologue (option "prefetch"): pull the BVH (nodes + primitive records) into L2 at HBM streaming speed.  Traversal
// is a dependent pointer chase, so on a cold L2 every first touch of a node costs a DRAM round trip on the critical path
// of some warp; one bulk pass over the ~70 MB of a 1M-triangle scene takes ~15 us.
__global__ void k_prefetch_l2(const char* __restrict__ a, size_t na, const char* __restrict__ b, size_t nb) {
    const size_t stride = (size_t)gridDim.x * blockDim.x * 128u;
    for (size_t o = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 128u; o < na; o += stride) asm volatile("prefetch.global.L2 [%0];" ::"l"(a + o));
    for (size_t o = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 128u; o < nb; o += stride) asm volatile("prefetch.global.L2 [%0];" ::"l"(b + o));
}

// resident CTAs per SM of a persistent kernel (occupancy query, cached per kernel FUNCTION — instantiations of one template
// share their pointer type, so the key is the pointer, not K; the query costs ~10 us of host time and a frame on a 1/8
// partition lasts only 0.5 ms).  Contexts on different host threads may get here at once.
template <class K>
int persistent_grid(K kernel, int block, size_t smem, int sm_count, int ctas_per_sm) {
    static std::mutex mu;
    static std::vector<std::pair<const void*, int>> seen;
    int cached = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        for (auto& e : seen) if (e.first == (const void*)kernel) cached = e.second;
        if (cached < 1) {
            int per_sm = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
            cached = per_sm;
            seen.emplace_back((const void*)kernel, cached);
        }
    }
    int per_sm = cached;
    if (ctas_per_sm > 0 && ctas_per_sm < per_sm) per_sm = ctas_per_sm;
    return sm_count * per_sm;
}

}  // namespace

// split < 0: cut the frame into |split| waves of whole pixel rows AND record one event per finished wave (rt_render's
// pipelined copy-out); split > 1: just cut into that many waves; else one wave per chunk_pixels.
int rt_device_render(RtDevice& D, const RtView& V_in, int want, const RtOptions& opt, std::string& err, RtFrameStats* fs, int split) {
    RT_CUDA(cudaSetDevice(D.device));
    RtView V = V_in;
    {   // reflection depth actually needed: none without a mirror material (RTS:358); clamp to the buffer limit
        int md = (D.has_mirror && V.max_bounces > 0) ? (V.max_bounces > RT_MAX_DEPTH ? RT_MAX_DEPTH : V.max_bounces) : 0;
        V.max_bounces = md;
    }
    cudaStream_t st = D.stream;                 // stream of the wave being enqueued (one of the lanes')
    uint32_t launches = 0;
    D.kev_class.clear();
    cudaError_t kev_err = cudaSuccess;          // first failure of the timing events (checked once after the launches)
    auto kbegin = [&](int cls) {
        if (!opt.kernel_timing || kev_err != cudaSuccess) return;
        size_t i = D.kev_class.size();
        while (D.kev_pool.size() < 2 * (i + 1)) { cudaEvent_t e; if ((kev_err = cudaEventCreate(&e)) != cudaSuccess) return; D.kev_pool.push_back(e); }
        D.kev_class.push_back(cls);
        kev_err = cudaEventRecord(D.kev_pool[2 * i], st);
    };
    auto kend = [&]() {
        if (!opt.kernel_timing || kev_err != cudaSuccess || D.kev_class.empty()) return;
        kev_err = cudaEventRecord(D.kev_pool[2 * (D.kev_class.size() - 1) + 1], st);
    };
    RT_CUDA(D.stats.ensure(4 * 8));
    RT_CUDA(cudaMemsetAsync(D.stats.p, 0, 4 * 8, D.stream));
    RT_CUDA(cudaEventRecord(D.ev0, D.stream));
    const int g = V.spp_grid, n_pass = g * g;
    uint64_t primary = 0;
    D.wave_marks.clear();
    if (V.n_tiles > 0 && V.res_x > 0) {
        const int block = RT_BLOCK;
        int wave_tiles = (opt.chunk_pixels / 32) < 1 ? 1 : (opt.chunk_pixels / 32);
        if (wave_tiles > V.n_tiles) wave_tiles = V.n_tiles;
        const int chunk_tiles = wave_tiles;                                 // capacity limit set by chunk_pixels
        bool mark = split < 0;
        const int n_split = split < 0 ? -split : split;
        if (n_split > 1) {
            // waves of whole tile rows (4 pixel rows each): a finished wave is a contiguous block of the (1-rank) frame
            const int tile_rows = V.n_tiles / V.tiles_x;
            int rows_per_wave = (tile_rows + n_split - 1) / n_split;
            const int cap_rows = wave_tiles / V.tiles_x;
            if (cap_rows >= 1) {
                if (rows_per_wave > cap_rows) rows_per_wave = cap_rows;
                if (rows_per_wave < 1) rows_per_wave = 1;
                wave_tiles = rows_per_wave * V.tiles_x;
            } else mark = false;
        } else mark = false;
        // wave plan: first tile of every wave (+ end).  Uniform, except for the pipelined copy-out with the default
        // copy_waves = 8, where small waves sit at both ends: 1,2,4,4,2,2,1 sixteenths of the rows.  The copy engine needs
        // as long as the SMs (2.3 ms of PCIe for a 4K Color[] frame), so the call ends at
        // max(first wave + copy of everything, render of everything + copy of the last wave): both ends want to be small,
        // while every extra wave costs a drain, so the middle ones stay big.
        std::vector<int> wave_tb;
        {
            const int tile_rows = V.n_tiles / V.tiles_x;
            const int cap_rows = chunk_tiles / V.tiles_x;
            int plan16[16] = {1, 2, 4, 4, 2, 2, 1}, n_plan = 7, plan_max = 4;
            if (opt.copy_plan > 0) {                                 // decimal digits = sixteenths of the rows per wave (must sum to 16)
                int digits[16], nd = 0, sum = 0, mx = 0;
                for (long long v = opt.copy_plan; v > 0 && nd < 16; v /= 10) { digits[nd++] = (int)(v % 10); }
                for (int k = 0; k < nd; ++k) { sum += digits[k]; mx = digits[k] > mx ? digits[k] : mx; }
                if (sum == 16 && nd >= 2) { n_plan = nd; plan_max = mx; for (int k = 0; k < nd; ++k) plan16[k] = digits[nd - 1 - k]; }
            }
            if (mark && n_split == 8 && tile_rows >= 64 && cap_rows >= tile_rows * plan_max / 16 + 2) {
                int acc = 0;
                for (int k = 0; k < n_plan; ++k) { if (plan16[k] > 0) wave_tb.push_back((int)((long long)tile_rows * acc / 16) * V.tiles_x); acc += plan16[k]; }
                wave_tiles = ((tile_rows * plan_max + 15) / 16 + 1) * V.tiles_x;     // capacity: the largest wave
                if (wave_tiles > V.n_tiles) wave_tiles = V.n_tiles;
            } else {
                for (int tb = 0; tb < V.n_tiles; tb += wave_tiles) wave_tb.push_back(tb);
            }
            wave_tb.push_back(V.n_tiles);
        }
        const int waves_per_pass = (int)wave_tb.size() - 1;
        int n_lanes = 1;
        if (!opt.kernel_timing && waves_per_pass * n_pass > 1) {
            n_lanes = opt.lanes;
            // (not for the pipelined copy-out of a single pass: there the waves must FINISH one after the other, early ones early,
            // so that their copies can start — with all of them side by side cfg2's rt_render went from 2.9 to 3.2 ms)
            if (V.max_bounces > 0 && waves_per_pass * n_pass >= 4 && opt.deep_lanes > n_lanes && (!mark || n_pass > 1)) n_lanes = opt.deep_lanes;
            if (n_lanes > waves_per_pass * n_pass) n_lanes = waves_per_pass * n_pass;
            if (n_lanes > RtDevice::RT_MAX_LANES) n_lanes = RtDevice::RT_MAX_LANES;
            if (n_lanes < 1) n_lanes = 1;
        }
        const size_t W = (size_t)wave_tiles * 32;
        const int maxd = V.max_bounces;
        // mirror queues are re-ordered before they are traced (k_qs_*) when the scene is big enough for ray incoherence to cost something
        const bool sort_queue = maxd > 0 && (opt.sort_queue > 0 || (opt.sort_queue < 0 && (int64_t)D.S.n_prims + (D.S.n_inst > 0 ? 65536 : 0) >= 65536));
        // Sorted neighbours are similar, not parallel: they pass the packet's coherence test (rt_warp_coherent: within ~18 degrees) and
        // would then walk the UNION of 32 paths that really differ — measured 3x slower than unsorted.  Sorted queues are therefore
        // traced by the per-lane loop, which profits from the order through convergent lanes and shared cache lines (sort_queue = 2
        // keeps the packet test: development).
        RtScene Sq = D.S;
        if (sort_queue && opt.sort_queue != 2) Sq.packet = 0;
        RtQsGrid qsg;
        {
            const f3 lo = D.S.aabb_min, hi = D.S.aabb_max;
            auto sc = [](float a, float b) { const float e = b - a; return (e > 0.f && std::isfinite(e)) ? (float)RT_QS_CELLS / e : 0.f; };
            qsg.lo = mk3(std::isfinite(lo.x) ? lo.x : 0.f, std::isfinite(lo.y) ? lo.y : 0.f, std::isfinite(lo.z) ? lo.z : 0.f);
            qsg.scale = mk3(sc(lo.x, hi.x), sc(lo.y, hi.y), sc(lo.z, hi.z));
        }
        const int n_chunks = (D.S.n_lights + 31) / 32;                      // shadow launches per depth
        for (int li = 0; li < n_lanes; ++li) {                              // work buffers of each lane
            RtDevice::Lane& L = D.lanes[li];
            RT_CUDA(L.hit0_t.ensure(W * 4)); RT_CUDA(L.hit0_slot.ensure(W * 4));
            RT_CUDA(L.counters.ensure(C_TOTAL * 4));
            if ((int)L.local.size() < maxd + 1) L.local.resize(maxd + 1);
            if ((int)L.link.size() < maxd + 1) L.link.resize(maxd + 1);
            for (int d = 0; d <= maxd; ++d) RT_CUDA(L.local[d].ensure(W * 16));
            if (maxd > 0) {
                RT_CUDA(L.hitq_t.ensure(W * 4)); RT_CUDA(L.hitq_slot.ensure(W * 4));
                RT_CUDA(L.rays[0].ensure(W * 32)); RT_CUDA(L.rays[1].ensure(W * 32));
                for (int d = 1; d <= maxd; ++d) RT_CUDA(L.link[d].ensure(W * 16));
                if (sort_queue) { RT_CUDA(L.link_tmp.ensure(W * 16)); RT_CUDA(L.qs_hist.ensure((size_t)RT_QS_BINS * 4)); }
            }
            RT_CUDA(L.occl.ensure(W * 4 * (size_t)(n_chunks < 1 ? 1 : n_chunks)));
            if (li > 0) {
                if (!L.own_stream) RT_CUDA(cudaStreamCreateWithFlags(&L.own_stream, cudaStreamNonBlocking));
                if (!L.done) RT_CUDA(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
                RT_CUDA(cudaStreamWaitEvent(L.own_stream, D.ev0, 0));        // after everything queued before this frame
            }
            if (opt.light_split != 0 && opt.fuse_shade && n_chunks == 1) {
                // Combine buffers of the per-light items.  Invariant: all zero between frames (the last item of a tile leaves
                // its words zero).  They are zeroed when (re)allocated and after a frame that did not run to its end, on the
                // stream of the lane that uses them, so the order against that lane's kernels never depends on another lane.
                const void* p0 = L.split_occl.p; const void* p1 = L.split_acc.p;
                RT_CUDA(L.split_occl.ensure(W * 4)); RT_CUDA(L.split_acc.ensure((size_t)wave_tiles * 8));
                cudaStream_t ls = li ? L.own_stream : D.stream;
                if (L.split_occl.p != p0 || D.split_dirty) RT_CUDA(cudaMemsetAsync(L.split_occl.p, 0, L.split_occl.bytes, ls));
                if (L.split_acc.p != p1 || D.split_dirty) RT_CUDA(cudaMemsetAsync(L.split_acc.p, 0, L.split_acc.bytes, ls));
            }
        }
        D.split_dirty = true;                                               // until this frame's launches are all queued
        if (n_pass > 1) {
            RT_CUDA(D.accum.ensure((size_t)V.n_tiles * 32 * 16));
            while (n_lanes > 1 && (int)D.fin_ev.size() < waves_per_pass) { cudaEvent_t e; RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); D.fin_ev.push_back(e); }
        }
        // ---- cost hints of the previous frame (same layout only) / recording for the next ----
        // Hints are used only when the previous frame RECORDED them under the same layout: tiles, wave plan, resolution,
        // partition, scene size, and what the shadow pass looks like (a frame without lights launches no shadow kernel and
        // leaves that kernel's class counts at zero — with them the fused kernel would hand out no tile at all).
        const bool fuse = opt.fuse_shade && n_chunks == 1;                  // shade (+ finalize) inside the shadow kernel
        long long plan_hash = 1469598103934665603ll;
        for (int tbk : wave_tb) plan_hash = (plan_hash ^ tbk) * 1099511628211ll;
        const long long sig[RtDevice::HINT_SIG] = {V.n_tiles, (long long)wave_tiles * 64 + waves_per_pass, V.res_x, V.res_y, V.rank, V.world, V.band_rows,
                                                   (long long)D.S.n_prims + ((long long)D.S.n_inst << 32), D.S.n_lights, fuse ? 1 : 0, plan_hash, n_pass};
        bool hint_in = opt.hints && D.hint_ready;
        for (int k = 0; k < RtDevice::HINT_SIG; ++k) hint_in = hint_in && D.hint_sig[k] == sig[k];
        const bool hint_in_shadow = hint_in && D.hint_shadow_ready;
        const bool hint_rec = opt.hints != 0;
        const int hb_in = D.hint_cur, hb_out = 1 - D.hint_cur;
        const size_t hstride = (size_t)V.n_tiles;
        if (hint_rec) {
            for (auto* hh : {&D.hint_primary, &D.hint_shadow}) {
                for (int k = 0; k < 2; ++k) { RT_CUDA(hh->lists[k].ensure(hstride * RT_NCLS * 4)); RT_CUDA(hh->counts[k].ensure((size_t)waves_per_pass * RT_NCLS * 4)); }
                RT_CUDA(cudaMemsetAsync(hh->counts[hb_out].p, 0, (size_t)waves_per_pass * RT_NCLS * 4, D.stream));
            }
            for (int li = 1; li < n_lanes; ++li) {                           // the lanes must see the cleared counters
                RT_CUDA(cudaEventRecord(D.lanes[li].done, D.stream));
                RT_CUDA(cudaStreamWaitEvent(D.lanes[li].own_stream, D.lanes[li].done, 0));
            }
        }
        const int lsplit = opt.light_split >= 0 ? opt.light_split : (V.world >= 4 ? 6 : 0);
        auto hin_of = [&](RtDevice::Hints& hh, int tb, int wi) {
            RtHintIn h{nullptr, nullptr, 0, RT_NCLS};
            const bool shadow = &hh == &D.hint_shadow;
            if (shadow ? hint_in_shadow : hint_in) h = RtHintIn{hh.lists[hb_in].as<int>() + tb, hh.counts[hb_in].as<unsigned>() + (size_t)wi * RT_NCLS, hstride,
                                                                (shadow && lsplit > 0) ? lsplit : RT_NCLS};
            return h;
        };
        auto hout_of = [&](RtDevice::Hints& hh, int tb, int wi, bool on) {
            RtHintOut h{nullptr, nullptr, 0};
            if (hint_rec && on) h = RtHintOut{hh.lists[hb_out].as<int>() + tb, hh.counts[hb_out].as<unsigned>() + (size_t)wi * RT_NCLS, hstride};
            return h;
        };
        // the instantiations for this frame: INST when the scene has instanced meshes, FAR for a camera beyond the node padding's reach
        const bool inst = D.S.n_inst > 0;
        const auto kp_primary = inst ? (V.far_origin ? k_trace_primary<true, true> : k_trace_primary<false, true>)
                                     : (V.far_origin ? k_trace_primary<true, false> : k_trace_primary<false, false>);
        const auto kp_queue = inst ? k_trace_queue<true> : k_trace_queue<false>;
        const auto kp_shadow0 = inst ? (fuse ? k_trace_shadow<true, true, true> : k_trace_shadow<true, false, true>)
                                     : (fuse ? k_trace_shadow<true, true, false> : k_trace_shadow<true, false, false>);
        const auto kp_shadown = inst ? (fuse ? k_trace_shadow<false, true, true> : k_trace_shadow<false, false, true>)
                                     : (fuse ? k_trace_shadow<false, true, false> : k_trace_shadow<false, false, false>);
        const auto kp_shade0 = inst ? k_shade<true, true> : k_shade<true, false>;
        const auto kp_shaden = inst ? k_shade<false, true> : k_shade<false, false>;
        const auto kp_finalize = inst ? k_finalize<true> : k_finalize<false>;
        const size_t smem = RT_STACK_SMEM + (inst ? RT_BAKE_SMEM : 0);
        const int g_trace = persistent_grid(kp_primary, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_traceq = persistent_grid(kp_queue, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_shadow0 = persistent_grid(kp_shadow0, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_shadown = persistent_grid(kp_shadown, block, smem, D.sm_count, opt.ctas_per_sm);
        const int g_flat = D.sm_count * 8;
        unsigned long long* ST = D.stats.as<unsigned long long>();
        RtOut out{(want & RT_OUT_RGBA32F) ? (float4*)D.o_rgba32f : nullptr, (want & RT_OUT_RGBA8) ? (uint32_t*)D.o_rgba8 : nullptr,
                  (want & RT_OUT_HIT_ID) ? (int32_t*)D.o_hit_id : nullptr, (want & RT_OUT_HIT_T) ? (float*)D.o_hit_t : nullptr};
        if (opt.prefetch) {                                                 // L2 prologue (k_prefetch_l2), before the lanes fork
            const size_t nb_nodes = (size_t)D.n_nodes * 64, nb_prims = (size_t)D.S.n_prims * 48;
            if (D.S.nodes && D.S.prims && nb_nodes + nb_prims > 0 && nb_nodes + nb_prims <= ((size_t)opt.prefetch_max_mb << 20)) {
                k_prefetch_l2<<<D.sm_count * 4, 256, 0, D.stream>>>((const char*)D.S.nodes, nb_nodes, (const char*)D.S.prims, nb_prims);
                ++launches;
            }
        }
        int wave_seq = 0;
        for (int sj = 0, pass = 0; sj < g; ++sj)
            for (int si = 0; si < g; ++si, ++pass)
                for (int wi = 0; wi < waves_per_pass; ++wi, ++wave_seq) {
                    const int tb = wave_tb[wi], nt = wave_tb[wi + 1] - tb;
                    RtDevice::Lane& L = D.lanes[wave_seq % n_lanes];
                    st = (wave_seq % n_lanes) ? L.own_stream : D.stream;
                    unsigned* C = L.counters.as<unsigned>();
                    uint32_t* OC = L.occl.as<uint32_t>();
                    RT_CUDA(cudaMemsetAsync(C, 0, C_TOTAL * 4, st));
                    {
                    kbegin(RT_K_TRACE_PRIMARY);
                    kp_primary<<<g_trace, block, smem, st>>>(D.S, V, tb, nt, si, sj, L.hit0_t.as<float>(), L.hit0_slot.as<int>(), C + C_WORK,
                                                              hin_of(D.hint_primary, tb, wi), hout_of(D.hint_primary, tb, wi, pass == 0));
                    kend(); ++launches;
                    }
                    for (int d = 0; d <= maxd; ++d) {
                        // queue buffers.  Unsorted: depth d reads rays[(d-1)&1] and emits into rays[d&1] / link[d+1].  Sorted: every depth
                        // emits into rays[0] / link_tmp, the next depth sorts that into rays[1] / link[d] and reads it from there.
                        const float4* rin = d ? L.rays[sort_queue ? 1 : ((d - 1) & 1)].as<float4>() : nullptr;
                        float4* rout = d < maxd ? L.rays[sort_queue ? 0 : (d & 1)].as<float4>() : nullptr;
                        float4* lout = d < maxd ? (sort_queue ? L.link_tmp.as<float4>() : L.link[d + 1].as<float4>()) : nullptr;
                        const float* ht = d ? L.hitq_t.as<float>() : L.hit0_t.as<float>();
                        const int* hs = d ? L.hitq_slot.as<int>() : L.hit0_slot.as<int>();
                        unsigned* wk = C + C_WORK + 8 * d;
                        if (d > 0) {
                            kbegin(RT_K_TRACE_QUEUE);
                            if (sort_queue) {
                                unsigned* hist = L.qs_hist.as<unsigned>();
                                RT_CUDA(cudaMemsetAsync(hist, 0, (size_t)RT_QS_BINS * 4, st));
                                k_qs_hist<<<g_flat, 256, 0, st>>>(C + d, L.rays[0].as<float4>(), qsg, hist);
                                k_qs_scan<<<1, 1024, 0, st>>>(hist);
                                k_qs_scatter<<<g_flat, 256, 0, st>>>(C + d, L.rays[0].as<float4>(), L.link_tmp.as<float4>(), qsg, hist, L.rays[1].as<float4>(), L.link[d].as<float4>());
                                launches += 3;
                            }
                            kp_queue<<<g_traceq, block, smem, st>>>(Sq, C + d, rin, L.hitq_t.as<float>(), L.hitq_slot.as<int>(), wk);
                            kend(); ++launches;
                        }
                        if (fuse) {
                            RtFuse F{d, L.local[d].as<float4>(), rout, lout, C + d + 1, (d == 0 && maxd == 0) ? 1 : 0, pass, n_pass,
                                     n_pass > 1 ? D.accum.as<float4>() : nullptr, out, L.split_occl.as<uint32_t>(), L.split_acc.as<unsigned long long>(), (unsigned)W};
                            if (F.fin && n_pass > 1 && n_lanes > 1 && pass > 0) RT_CUDA(cudaStreamWaitEvent(st, D.fin_ev[wi], 0));
                            kbegin(d ? RT_K_SHADOW_QUEUE : RT_K_SHADOW_PRIMARY);
                            if (d == 0) kp_shadow0<<<g_shadow0, block, smem, st>>>(D.S, V, tb, nt, si, sj, nullptr, nullptr, ht, hs, 0, D.S.n_lights, OC, wk + 1, ST, F,
                                                                                  hin_of(D.hint_shadow, tb, wi), hout_of(D.hint_shadow, tb, wi, pass == 0));
                            else kp_shadown<<<g_shadown, block, smem, st>>>((opt.sort_queue == 3 ? Sq : D.S), V, 0, 0, 0, 0, C + d, rin, ht, hs, 0, D.S.n_lights, OC, wk + 1, ST, F, RtHintIn{nullptr, nullptr, 0, RT_NCLS}, RtHintOut{nullptr, nullptr, 0});
                            kend(); ++launches;
                            continue;
                        }
                        for (int c = 0; c < n_chunks; ++c) {
                            unsigned* wc = wk + 1 + (c % 7);
                            if (c >= 7) RT_CUDA(cudaMemsetAsync(wc, 0, 4, st));
                            const int lb = c * 32, nl = (D.S.n_lights - lb) < 32 ? (D.S.n_lights - lb) : 32;
                            kbegin(d ? RT_K_SHADOW_QUEUE : RT_K_SHADOW_PRIMARY);
                            if (d == 0) kp_shadow0<<<g_shadow0, block, smem, st>>>(D.S, V, tb, nt, si, sj, nullptr, nullptr, ht, hs, lb, nl, OC + (size_t)c * W, wc, ST, RtFuse{},
                                                                                  hin_of(D.hint_shadow, tb, wi), hout_of(D.hint_shadow, tb, wi, pass == 0 && c == 0));
                            else kp_shadown<<<g_shadown, block, smem, st>>>(D.S, V, 0, 0, 0, 0, C + d, rin, ht, hs, lb, nl, OC + (size_t)c * W, wc, ST, RtFuse{}, RtHintIn{nullptr, nullptr, 0, RT_NCLS}, RtHintOut{nullptr, nullptr, 0});
                            kend(); ++launches;
                        }
                        kbegin(d ? RT_K_SHADE_QUEUE : RT_K_SHADE_PRIMARY);
                        if (d == 0) kp_shade0<<<g_flat, 256, 0, st>>>(D.S, V, 0, tb, nt, si, sj, nullptr, nullptr, ht, hs, OC, W, L.local[0].as<float4>(), rout, lout, C + 1, ST);
                        else kp_shaden<<<g_flat, 256, 0, st>>>(D.S, V, d, 0, 0, 0, 0, C + d, rin, ht, hs, OC, W, L.local[d].as<float4>(), rout, lout, C + d + 1, ST);
                        kend(); ++launches;
                    }
                    for (int d = maxd; d >= 1; --d) {
                        kbegin(RT_K_FOLD);
                        k_fold<<<g_flat, 256, 0, st>>>(C + d, L.link[d].as<float4>(), L.local[d].as<float4>(), L.local[d - 1].as<float4>());
                        kend();
                        ++launches;
                    }
                    if (!(fuse && maxd == 0)) {
                        // spp passes accumulate per pixel: pass p of these tiles must follow pass p-1 of the same tiles
                        if (n_pass > 1 && n_lanes > 1 && pass > 0) RT_CUDA(cudaStreamWaitEvent(st, D.fin_ev[wi], 0));
                        kbegin(RT_K_FINALIZE);
                        kp_finalize<<<g_flat, 256, 0, st>>>(D.S, V, tb, nt, pass, n_pass, L.local[0].as<float4>(), L.hit0_t.as<float>(), L.hit0_slot.as<int>(),
                                                            n_pass > 1 ? D.accum.as<float4>() : nullptr, out);
                        kend();
                        ++launches;
                    }
                    if (n_pass > 1 && n_lanes > 1 && pass < n_pass - 1) RT_CUDA(cudaEventRecord(D.fin_ev[wi], st));
                    if (mark && pass == n_pass - 1) {
                        const size_t i = D.wave_marks.size();
                        #ifdef RT_WAVE_TRACE
                        while (D.wave_ev.size() <= i) { cudaEvent_t e; RT_CUDA(cudaEventCreate(&e)); D.wave_ev.push_back(e); }
#else
                        while (D.wave_ev.size() <= i) { cudaEvent_t e; RT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); D.wave_ev.push_back(e); }
#endif
                        RT_CUDA(cudaEventRecord(D.wave_ev[i], st));
                        const int r0 = (tb / V.tiles_x) * 4, r1 = ((tb + nt) / V.tiles_x) * 4;
                        D.wave_marks.push_back(RtDevice::WaveMark{r0, r1 < V.n_local_rows ? r1 : V.n_local_rows});   // LOCAL rows of this partition
                    }
                }
        for (int li = 1; li < n_lanes; ++li) {                               // join the lanes back into the caller's stream
            RT_CUDA(cudaEventRecord(D.lanes[li].done, D.lanes[li].own_stream));
            RT_CUDA(cudaStreamWaitEvent(D.stream, D.lanes[li].done, 0));
        }
        RT_CUDA(cudaGetLastError());
        RT_CUDA(kev_err);
        if (hint_rec) { D.hint_cur = hb_out; D.hint_ready = true; D.hint_shadow_ready = n_chunks > 0; for (int k = 0; k < RtDevice::HINT_SIG; ++k) D.hint_sig[k] = sig[k]; }
        else D.hint_ready = false;
        D.split_dirty = false;
        primary = (uint64_t)V.res_x * (uint64_t)V.n_local_rows * (uint64_t)n_pass;
    }
    RT_CUDA(cudaEventRecord(D.ev1, D.stream));
    if (fs) { fs->primary = primary; fs->shadow = 0; fs->mirror = 0; fs->shadow_skipped = 0; fs->launches = launches; }
    return RT_OK;
}

#ifdef RT_TILE_TIMING
extern "C" int rt_debug_tile_times(int n_tiles, unsigned long long* host_out) {       // (start, duration) per tile of the last frame
    static unsigned long long* buf = nullptr; static int cap = 0;
    if (host_out && buf) return (int)cudaMemcpy(host_out, buf, (size_t)(n_tiles < cap ? n_tiles : cap) * 24, cudaMemcpyDeviceToHost);   // + shadow-kernel cycles per tile
    if (n_tiles > cap) { if (buf) cudaFree(buf); cudaMalloc(&buf, (size_t)n_tiles * 24); cap = n_tiles; }
    cudaMemset(buf, 0, (size_t)cap * 24);
    return (int)cudaMemcpyToSymbol(g_rt_tile_dbg, &buf, sizeof(buf));
}
#endif

#ifdef RT_CHECKED
extern "C" int rt_debug_checks(unsigned* host_out16, int reset) {                       // violation counters of the RT_CHECKED build (rt_core.h)
    int rc = 0;
    if (host_out16) rc = (int)cudaMemcpyFromSymbol(host_out16, g_rt_check, RT_CHK_KINDS * sizeof(unsigned));
    if (reset) { unsigned z[RT_CHK_KINDS] = {0}; rc |= (int)cudaMemcpyToSymbol(g_rt_check, z, sizeof(z)); }
    return rc;
}
#endif

#ifdef RT_COUNT_VISITS
extern "C" int rt_debug_visit_counts(unsigned long long* host_out20, int reset) {       // counters of rt_core.h's RT_COUNT
    int rc = 0;
    if (host_out20) rc = (int)cudaMemcpyFromSymbol(host_out20, g_rt_visits, 20 * sizeof(unsigned long long));
    if (reset) { unsigned long long z[20] = {0}; rc |= (int)cudaMemcpyToSymbol(g_rt_visits, z, sizeof(z)); }
    return rc;
}
#endif

int rt_device_read_stats(RtDevice& D, RtFrameStats* fs, std::string& err) {
    RT_CUDA(cudaSetDevice(D.device));
    unsigned long long h[4] = {0, 0, 0, 0};
    if (D.stats.p) {
        RT_CUDA(cudaMemcpyAsync(h, D.stats.p, sizeof(h), cudaMemcpyDeviceToHost, D.stream));
        RT_CUDA(cudaStreamSynchronize(D.stream));
    }
    fs->shadow = h[1]; fs->mirror = h[2]; fs->shadow_skipped = h[3];
    for (int k = 0; k < RT_K_COUNT; ++k) { fs->ms_kernel[k] = 0.f; fs->launches_kernel[k] = 0; }
    for (size_t i = 0; i < D.kev_class.size(); ++i) {
        float m = 0.f;
        RT_CUDA(cudaEventSynchronize(D.kev_pool[2 * i + 1]));
        RT_CUDA(cudaEventElapsedTime(&m, D.kev_pool[2 * i], D.kev_pool[2 * i + 1]));
        fs->ms_kernel[D.kev_class[i]] += m; fs->launches_kernel[D.kev_class[i]] += 1;
    }
    return RT_OK;
}

int rt_device_intersect(RtDevice& D, const float* d_origins, const float* d_dirs, int n, int32_t* d_ids, float* d_t,
                        const RtOptions& opt, std::string& err) {
    RT_CUDA(cudaSetDevice(D.device));
    if (n <= 0) return RT_OK;
    const int block = RT_BLOCK;
    const size_t smem = RT_STACK_SMEM;
    int grid = (n + block - 1) / block;
    if (grid > D.sm_count * 16) grid = D.sm_count * 16;
    if (D.S.n_inst > 0) k_intersect<true><<<grid, block, smem, D.stream>>>(D.S, n, d_origins, d_dirs, d_ids, d_t);
    else k_intersect<false><<<grid, block, smem, D.stream>>>(D.S, n, d_origins, d_dirs, d_ids, d_t);
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

// ---- rt_intersect_static: the allocation-free form of rt_intersect (include/rtb200.h) ---------------------------------------
// rt_intersect stages its rays and results in four cudaMalloc'ed buffers per call.  Callers that must not allocate (code under
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the same query with that staging in a
// file-scope device array: RT_INTERSECT_STATIC_MAX rays x 40 B = 2.5 MiB, held by every GPU that loads the library until
// the process exits.  Layout (offsets in bytes, all multiples of 256): origins [0, 12 MAX), directions [12 MAX, 24 MAX),
// ids [24 MAX, 36 MAX), distances [36 MAX, 40 MAX).  Nothing in it has to start at zero: the kernel overwrites ids[i] and t[i]
// for every i < n before they are copied out, and reads only the o[i], d[i] the same call uploaded; what a previous call left
// beyond n is never read.  Every call of every context on a device shares the ONE array: calls that overlap on the GPU
// (two contexts, or one context on two host threads) race.
// In host code the array's NAME is a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress,
// asked once per device with that device current.
__device__ __align__(256) static unsigned char g_intersect_static[(size_t)RT_INTERSECT_STATIC_MAX * 40];

namespace {

// k_intersect with one difference: this entry point may not refit the sphere boxes for origins outside the origin box of the
// last commit (rt_abi.cu cover_origin allocates), so such a ray — like one whose direction is not normalised — takes the
// reference's own linear scan (Scene.cs:43-122), which needs no boxes at all.  Same primitive tests, same (t, key) argmin.
template <bool INST>
__global__ void __launch_bounds__(128) k_intersect_static(RtScene S, int n, f3 obox_lo, f3 obox_hi, const float* __restrict__ o,
                                                           const float* __restrict__ d, int32_t* __restrict__ out_id, float* __restrict__ out_t) {
    RT_DECL_STACK(stk);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        f3 O = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), D = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        float bt = FLT_MAX; int bs = -1;
        const f3 inv = rt_rcp3(D);
        if (rt_scene_gate(S, O, inv)) {
            const float dd = um_dot(D, D);
            const bool covered = O.x >= obox_lo.x && O.x <= obox_hi.x && O.y >= obox_lo.y && O.y <= obox_hi.y && O.z >= obox_lo.z && O.z <= obox_hi.z;
            if (S.n_spheres > 0 && (!covered || !(fabsf(dd - 1.0f) <= 0x1p-20f))) rt_scan_all<INST>(S, O, D, inv, bt, bs);
            else { stk.sp = 0; rt_traverse<false, true, INST>(S, O, D, inv, 0.f, stk, bt, bs); }
        }
        int32_t ty, ix, me; rt_slot_to_id<INST>(S, bs, ty, ix, me);
        out_id[3 * i] = ty; out_id[3 * i + 1] = ix; out_id[3 * i + 2] = me; out_t[i] = bt;
    }
}

// device address of g_intersect_static on the CURRENT device, looked up once per device
int intersect_static_base(int device, unsigned char** out, std::string& err) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static unsigned char* base[MAX_DEV] = {};
    if (device < 0 || device >= MAX_DEV) { err = "rt_intersect_static: device ordinal out of range"; return RT_E_UNSUPPORTED; }
    std::lock_guard<std::mutex> lock(mu);
    if (!base[device]) {
        void* p = nullptr;
        RT_CUDA(cudaGetSymbolAddress(&p, g_intersect_static));
        if (!p) { err = "cudaGetSymbolAddress returned a null address"; return RT_E_CUDA; }
        base[device] = (unsigned char*)p;
    }
    *out = base[device];
    return RT_OK;
}

}  // namespace

// host arrays in, host arrays out; no cudaMalloc / cudaFree / allocating library routine on this path
int rt_device_intersect_static(RtDevice& D, const float* h_origins, const float* h_dirs, int n, int32_t* h_ids, float* h_t, std::string& err) {
    RT_CUDA(cudaSetDevice(D.device));
    if (n <= 0) return RT_OK;
    if (n > RT_INTERSECT_STATIC_MAX) { err = "rt_intersect_static: n above RT_INTERSECT_STATIC_MAX"; return RT_E_INVALID; }
    unsigned char* base = nullptr;
    int rc = intersect_static_base(D.device, &base, err);
    if (rc) return rc;
    constexpr size_t M = (size_t)RT_INTERSECT_STATIC_MAX;
    float* d_o = (float*)base; float* d_d = (float*)(base + 12 * M);
    int32_t* d_ids = (int32_t*)(base + 24 * M); float* d_t = (float*)(base + 36 * M);
    RT_CUDA(cudaMemcpyAsync(d_o, h_origins, (size_t)n * 12, cudaMemcpyHostToDevice, D.stream));
    RT_CUDA(cudaMemcpyAsync(d_d, h_dirs, (size_t)n * 12, cudaMemcpyHostToDevice, D.stream));
    int grid = (n + RT_BLOCK - 1) / RT_BLOCK;
    if (grid > D.sm_count * 16) grid = D.sm_count * 16;
    const f3 olo = mk3(D.obox[0], D.obox[1], D.obox[2]), ohi = mk3(D.obox[3], D.obox[4], D.obox[5]);
    if (D.S.n_inst > 0) k_intersect_static<true><<<grid, RT_BLOCK, RT_STACK_SMEM, D.stream>>>(D.S, n, olo, ohi, d_o, d_d, d_ids, d_t);
    else k_intersect_static<false><<<grid, RT_BLOCK, RT_STACK_SMEM, D.stream>>>(D.S, n, olo, ohi, d_o, d_d, d_ids, d_t);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpyAsync(h_ids, d_ids, (size_t)n * 12, cudaMemcpyDeviceToHost, D.stream));
    RT_CUDA(cudaMemcpyAsync(h_t, d_t, (size_t)n * 4, cudaMemcpyDeviceToHost, D.stream));
    RT_CUDA(cudaStreamSynchronize(D.stream));
    return RT_OK;
}

// rt_sah.h — host-side binned-SAH BVH2 builder (option "builder" = 1), the north star's alternative to the device
// LBVH: "the host-built BVH is re-laid out as a compact node array".  The reference's own builder
// (Data/Collision/BVH.cs:13-79) is a longest-axis midpoint split that stops at `throw new NotImplementedException()`
// (BVH.cs:77-79) and is never called; this finishes the idea with a surface-area heuristic and emits the device node
// format directly (rt_core.h RtScene::nodes: 64-byte nodes holding both child boxes, padded like the LBVH's).
//
// Host only, C++17, std::thread task parallelism over subtrees.  Output:
//   order[s]  = scan key of the primitive in leaf slot s (uploaded; k_emit_slots builds the slot records from it)
//   nodes     = 16 floats per node; node 0 is the root (unless the whole scene is one leaf)
// Result equivalence with the reference's scan does not depend on the tree at all (rt_core.h), only on the boxes being
// conservative, which the padding guarantees exactly as for the LBVH.
#pragma once
#include <vector>
#include <thread>
#include <future>
#include <atomic>
#include <array>
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <cfloat>
#include <cmath>

struct RtSahBox { float lo[3], hi[3]; };

struct RtSahBuilder {
    const RtSahBox* boxes = nullptr;          // per scan key
    int32_t n = 0, leaf_max = 4;
    float pad = 0.f;
    float term = 0.f;                         // > 0: SAH termination for small ranges (RtOptions::sah_term)
    std::vector<int32_t> order;               // leaf-slot order of scan keys
    std::vector<float> nodes;                 // 16 floats per node
    std::atomic<int32_t> n_nodes{0};
    int spawn_depth = 0, threads_ = 1;
    static constexpr int NB = 16;             // bins per axis
    static constexpr int MAX_SAH_DEPTH = 64;  // deeper ranges are halved by count (see build)
    std::atomic<int> max_depth{0};            // deepest leaf of the last run (checked against the stack bound at commit)

    static inline float area(const float* lo, const float* hi) {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return (dx < 0.f || dy < 0.f || dz < 0.f) ? 0.f : dx * dy + dy * dz + dz * dx;
    }
    static inline void grow(float* lo, float* hi, const float* blo, const float* bhi) {
        for (int k = 0; k < 3; ++k) { lo[k] = std::fmin(lo[k], blo[k]); hi[k] = std::fmax(hi[k], bhi[k]); }
    }
    static inline void reset(float* lo, float* hi) { for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; } }

    // Builds the subtree over order[b, e); returns its traversal ref (>= 0 node index, < 0 leaf) and its (unpadded) box.
    int32_t build(int32_t b, int32_t e, int depth, float* out_lo, float* out_hi) {
        const int32_t cnt = e - b;
        float lo[3], hi[3], clo[3], chi[3];
        reset(lo, hi); reset(clo, chi);
        // Ranges near the root are swept by several threads (the subtree tasks have not fanned out yet: without this the top
        // three levels cost as much as all the others together).  min/max and counts are order-independent, so the tree is
        // the same as the serial one.
        const int par = (depth < spawn_depth && cnt >= (1 << 16)) ? (threads_ >> depth > 1 ? threads_ >> depth : 1) : 1;
        auto chunk = [&](int c, int32_t& cb, int32_t& ce) { cb = b + (int32_t)((int64_t)cnt * c / par); ce = b + (int32_t)((int64_t)cnt * (c + 1) / par); };
        auto bounds_of = [&](int32_t cb, int32_t ce, float* l, float* h, float* cl, float* ch) {
            reset(l, h); reset(cl, ch);
            for (int32_t i = cb; i < ce; ++i) {
                const RtSahBox& x = boxes[order[i]];
                grow(l, h, x.lo, x.hi);
                for (int k = 0; k < 3; ++k) { float c = 0.5f * (x.lo[k] + x.hi[k]); cl[k] = std::fmin(cl[k], c); ch[k] = std::fmax(ch[k], c); }
            }
        };
        if (par == 1) bounds_of(b, e, lo, hi, clo, chi);
        else {
            std::vector<std::array<float, 12>> part((size_t)par);
            std::vector<std::future<void>> fs;
            for (int c = 0; c < par; ++c) {
                int32_t cb, ce; chunk(c, cb, ce);
                float* q = part[(size_t)c].data();
                fs.push_back(std::async(std::launch::async, [=, &bounds_of]() { bounds_of(cb, ce, q, q + 3, q + 6, q + 9); }));
            }
            for (auto& f : fs) f.get();
            for (int c = 0; c < par; ++c) { const float* q = part[(size_t)c].data(); grow(lo, hi, q, q + 3); grow(clo, chi, q + 6, q + 9); }
        }
        for (int k = 0; k < 3; ++k) { out_lo[k] = lo[k]; out_hi[k] = hi[k]; }
        { int seen = max_depth.load(std::memory_order_relaxed); while (depth > seen && !max_depth.compare_exchange_weak(seen, depth, std::memory_order_relaxed)) {} }
        if (term <= 0.f ? cnt <= leaf_max : cnt == 1) return ~(b * 8 + (cnt - 1));
        if (term > 0.f && cnt <= 8) {
            // ---- small range: exact sweep over the sorted centroids of each axis; leaf unless a split pays ----
            int32_t best_ax = -1, best_k = 0; float best = FLT_MAX;
            int32_t tmp[8], keep[3][8];
            for (int ax = 0; ax < 3; ++ax) {
                for (int32_t i = 0; i < cnt; ++i) tmp[i] = order[b + i];
                std::sort(tmp, tmp + cnt, [&](int32_t x, int32_t y) {
                    const float cx = boxes[x].lo[ax] + boxes[x].hi[ax], cy = boxes[y].lo[ax] + boxes[y].hi[ax];
                    return cx < cy || (cx == cy && x < y);
                });
                for (int32_t i = 0; i < cnt; ++i) keep[ax][i] = tmp[i];
                float ra[8]; float tlo[3], thi[3]; reset(tlo, thi);
                for (int32_t i = cnt - 1; i >= 1; --i) { grow(tlo, thi, boxes[tmp[i]].lo, boxes[tmp[i]].hi); ra[i] = area(tlo, thi); }
                reset(tlo, thi);
                for (int32_t k = 1; k < cnt; ++k) {
                    grow(tlo, thi, boxes[tmp[k - 1]].lo, boxes[tmp[k - 1]].hi);
                    const float cost = area(tlo, thi) * (float)k + ra[k] * (float)(cnt - k);
                    if (cost < best) { best = cost; best_ax = ax; best_k = k; }
                }
            }
            const float pa = area(lo, hi);
            if (cnt <= leaf_max && (best_ax < 0 || !((float)cnt * pa > term * pa + best))) return ~(b * 8 + (cnt - 1));
            if (best_ax < 0) { best_ax = 0; best_k = cnt / 2; }
            for (int32_t i = 0; i < cnt; ++i) order[b + i] = keep[best_ax][i];
            const int32_t mid2 = b + best_k;
            const int32_t me2 = n_nodes.fetch_add(1);
            float llo[3], lhi[3], rlo[3], rhi[3];
            const int32_t rl = build(b, mid2, depth + 1, llo, lhi), rr = build(mid2, e, depth + 1, rlo, rhi);
            float* np = nodes.data() + 16 * (size_t)me2;
            np[0] = llo[0] - pad; np[1] = lhi[0] + pad; np[2] = llo[1] - pad; np[3] = lhi[1] + pad;
            np[4] = rlo[0] - pad; np[5] = rhi[0] + pad; np[6] = rlo[1] - pad; np[7] = rhi[1] + pad;
            np[8] = llo[2] - pad; np[9] = lhi[2] + pad; np[10] = rlo[2] - pad; np[11] = rhi[2] + pad;
            std::memcpy(np + 12, &rl, 4); std::memcpy(np + 13, &rr, 4); np[14] = 0.f; np[15] = 0.f;
            return me2;
        }
        // ---- binned SAH over the three axes ----
        struct Bins { int32_t bc[3][NB]; float blo[3][NB][3], bhi[3][NB][3]; };
        float scale3[3]; bool axis_ok[3];
        for (int ax = 0; ax < 3; ++ax) {
            const float ext = chi[ax] - clo[ax];
            axis_ok[ax] = (ext > 0.f) && std::isfinite(ext);
            scale3[ax] = axis_ok[ax] ? (float)NB / ext : 0.f;
        }
        auto bin_range = [&](int32_t cb, int32_t ce, Bins& B) {
            for (int ax = 0; ax < 3; ++ax) for (int j = 0; j < NB; ++j) { B.bc[ax][j] = 0; reset(B.blo[ax][j], B.bhi[ax][j]); }
            for (int32_t i = cb; i < ce; ++i) {
                const RtSahBox& x = boxes[order[i]];
                for (int ax = 0; ax < 3; ++ax) {
                    if (!axis_ok[ax]) continue;
                    int j = (int)((0.5f * (x.lo[ax] + x.hi[ax]) - clo[ax]) * scale3[ax]);
                    j = j < 0 ? 0 : (j >= NB ? NB - 1 : j);
                    ++B.bc[ax][j]; grow(B.blo[ax][j], B.bhi[ax][j], x.lo, x.hi);
                }
            }
        };
        Bins first;                                              // stack: one per node, no allocation on the serial path
        std::vector<Bins> bins;
        if (par == 1) bin_range(b, e, first);
        else {
            bins.resize((size_t)par);
            std::vector<std::future<void>> fs;
            for (int c = 0; c < par; ++c) {
                int32_t cb, ce; chunk(c, cb, ce);
                Bins* B = &bins[(size_t)c];
                fs.push_back(std::async(std::launch::async, [=, &bin_range]() { bin_range(cb, ce, *B); }));
            }
            for (auto& f : fs) f.get();
            first = bins[0];
            for (int c = 1; c < par; ++c)
                for (int ax = 0; ax < 3; ++ax) for (int j = 0; j < NB; ++j) {
                    first.bc[ax][j] += bins[(size_t)c].bc[ax][j];
                    grow(first.blo[ax][j], first.bhi[ax][j], bins[(size_t)c].blo[ax][j], bins[(size_t)c].bhi[ax][j]);
                }
        }
        int best_axis = -1, best_bin = -1; float best_cost = FLT_MAX;
        for (int ax = 0; ax < 3; ++ax) {
            if (!axis_ok[ax]) continue;
            const int32_t* bc = first.bc[ax];
            float (*blo)[3] = first.blo[ax]; float (*bhi)[3] = first.bhi[ax];
            float ra[NB]; int32_t rc[NB];
            float tlo[3], thi[3]; reset(tlo, thi); int32_t c = 0;
            for (int j = NB - 1; j >= 1; --j) { if (bc[j]) grow(tlo, thi, blo[j], bhi[j]); c += bc[j]; ra[j] = area(tlo, thi); rc[j] = c; }
            reset(tlo, thi); c = 0;
            for (int j = 0; j < NB - 1; ++j) {
                if (bc[j]) grow(tlo, thi, blo[j], bhi[j]);
                c += bc[j];
                if (c == 0 || rc[j + 1] == 0) continue;
                float cost = area(tlo, thi) * (float)c + ra[j + 1] * (float)rc[j + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = j; }
            }
        }
        int32_t mid;
        // Depth bound.  The traversal stack (rt_render.cu SmemStack) holds 12 + 104 entries and an entry is pushed at most once
        // per level, so a tree deeper than that would silently drop subtrees.  Binned splits on geometrically distributed
        // primitives can peel one bin per level without end; from MAX_SAH_DEPTH on the range is halved by count instead, which
        // ends after log2(2^28) = 28 more levels (+ 8 for the exact sweep of small ranges): 64 + 28 + 8 = 100 < 116.
        if (depth >= MAX_SAH_DEPTH) best_axis = -1;
        if (best_axis < 0) {
            mid = b + cnt / 2;                                   // all centroids coincide (or depth bound): split by count
        } else {
            const int ax = best_axis;
            const float scale = (float)NB / (chi[ax] - clo[ax]);
            const float c0 = clo[ax];
            const int bb = best_bin;
            auto it = std::partition(order.begin() + b, order.begin() + e, [&](int32_t k) {
                const RtSahBox& x = boxes[k];
                int j = (int)((0.5f * (x.lo[ax] + x.hi[ax]) - c0) * scale);
                j = j < 0 ? 0 : (j >= NB ? NB - 1 : j);
                return j <= bb;
            });
            mid = (int32_t)(it - order.begin());
            if (mid == b || mid == e) mid = b + cnt / 2;
        }
        const int32_t me = n_nodes.fetch_add(1);
        float llo[3], lhi[3], rlo[3], rhi[3];
        int32_t rl, rr;
        if (depth < spawn_depth && cnt > 8192) {
            auto fut = std::async(std::launch::async, [&]() { return build(b, mid, depth + 1, llo, lhi); });
            rr = build(mid, e, depth + 1, rlo, rhi);
            rl = fut.get();
        } else {
            rl = build(b, mid, depth + 1, llo, lhi);
            rr = build(mid, e, depth + 1, rlo, rhi);
        }
        float* np = nodes.data() + 16 * (size_t)me;              // layout: RtScene::nodes (rt_core.h)
        np[0] = llo[0] - pad; np[1] = lhi[0] + pad; np[2] = llo[1] - pad; np[3] = lhi[1] + pad;
        np[4] = rlo[0] - pad; np[5] = rhi[0] + pad; np[6] = rlo[1] - pad; np[7] = rhi[1] + pad;
        np[8] = llo[2] - pad; np[9] = lhi[2] + pad; np[10] = rlo[2] - pad; np[11] = rhi[2] + pad;
        std::memcpy(np + 12, &rl, 4); std::memcpy(np + 13, &rr, 4); np[14] = 0.f; np[15] = 0.f;
        return me;
    }

    // returns the root ref; the root node (if any) is moved to index 0 because traversal starts at S.root
    int32_t run(const RtSahBox* prim_boxes, int32_t n_prims, int32_t leaf_max_, float pad_, int threads) {
        boxes = prim_boxes; n = n_prims; leaf_max = leaf_max_ < 1 ? 1 : (leaf_max_ > 8 ? 8 : leaf_max_); pad = pad_;
        order.resize(n);
        for (int32_t i = 0; i < n; ++i) order[i] = i;
        nodes.assign((size_t)16 * (n > 1 ? n - 1 : 1), 0.f);
        n_nodes = 0; max_depth = 0;
        spawn_depth = 0; threads_ = threads < 1 ? 1 : threads;
        for (int t = threads_; t > 1; t >>= 1) ++spawn_depth;
        spawn_depth += 1;                                        // a few more tasks than threads, for balance
        if (n == 0) return -1;
        float lo[3], hi[3];
        int32_t root = build(0, n, 0, lo, hi);
        nodes.resize((size_t)16 * n_nodes.load());
        return root;                                             // the first fetch_add happened at the root: root == 0
    }
};

// ---- reference pre-splitting ("early split clipping", Ernst & Greiner 2007) for the host SAH builder --------------------
// A thin triangle that lies diagonally in space has an axis-aligned box far larger than itself (cfg3's displaced shell:
// 0.1-wide facets that run up to 0.8 along the radius): the leaf boxes of neighbours overlap massively and every ray near
// the surface visits many leaves for nothing.  Such a triangle is cut by axis-aligned planes (midpoint of the longest box
// axis) into up to 2^max_depth convex pieces; each piece becomes its own BVH *reference* with the box of the clipped
// polygon, all references carrying the same triangle.  The query result cannot change: a leaf stores the whole triangle
// (exact test as ever), duplicates of a triangle in several leaves give the same (t, scan key), and the piece boxes cover
// the triangle.  Boxes are computed in double, rounded outward and still padded by the builder.
// split while  box half-area / polygon area > ratio  (1 = an axis-aligned rectangle, 2 = an axis-aligned right triangle).
struct RtPresplit {
    struct P3 { double x[3]; };
    static int clip(const P3* in, int n, int ax, double c, bool keep_low, P3* out) {        // Sutherland-Hodgman against x[ax] = c
        int m = 0;
        for (int i = 0; i < n; ++i) {
            const P3& a = in[i]; const P3& b = in[(i + 1) % n];
            const double da = keep_low ? c - a.x[ax] : a.x[ax] - c, db = keep_low ? c - b.x[ax] : b.x[ax] - c;   // >= 0: inside
            if (da >= 0) out[m++] = a;
            if ((da > 0 && db < 0) || (da < 0 && db > 0)) {
                const double t = da / (da - db);
                P3 p; for (int k = 0; k < 3; ++k) p.x[k] = a.x[k] + t * (b.x[k] - a.x[k]);
                p.x[ax] = c;
                out[m++] = p;
            }
        }
        return m;
    }
    static double area(const P3* p, int n) {                                               // polygon area (fan)
        double sx = 0, sy = 0, sz = 0;
        for (int i = 1; i + 1 < n; ++i) {
            const double ux = p[i].x[0] - p[0].x[0], uy = p[i].x[1] - p[0].x[1], uz = p[i].x[2] - p[0].x[2];
            const double vx = p[i + 1].x[0] - p[0].x[0], vy = p[i + 1].x[1] - p[0].x[1], vz = p[i + 1].x[2] - p[0].x[2];
            sx += uy * vz - uz * vy; sy += uz * vx - ux * vz; sz += ux * vy - uy * vx;
        }
        return 0.5 * std::sqrt(sx * sx + sy * sy + sz * sz);
    }
    static void emit(const P3* p, int n, const RtSahBox& clipto, int32_t key, std::vector<RtSahBox>& rb, std::vector<int32_t>& rk) {
        RtSahBox b;
        for (int k = 0; k < 3; ++k) {
            double lo = p[0].x[k], hi = p[0].x[k];
            for (int i = 1; i < n; ++i) { lo = std::fmin(lo, p[i].x[k]); hi = std::fmax(hi, p[i].x[k]); }
            float flo = (float)lo, fhi = (float)hi;
            if ((double)flo > lo) flo = std::nextafterf(flo, -FLT_MAX);
            if ((double)fhi < hi) fhi = std::nextafterf(fhi, FLT_MAX);
            b.lo[k] = std::fmax(flo, clipto.lo[k]); b.hi[k] = std::fmin(fhi, clipto.hi[k]);      // never beyond the triangle's own box
        }
        rb.push_back(b); rk.push_back(key);
    }
    static void split(const P3* p, int n, const RtSahBox& full, int32_t key, int depth, int max_depth, double ratio,
                      std::vector<RtSahBox>& rb, std::vector<int32_t>& rk) {
        double lo[3], hi[3];
        for (int k = 0; k < 3; ++k) { lo[k] = hi[k] = p[0].x[k]; for (int i = 1; i < n; ++i) { lo[k] = std::fmin(lo[k], p[i].x[k]); hi[k] = std::fmax(hi[k], p[i].x[k]); } }
        const double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        const double ha = dx * dy + dy * dz + dz * dx, pa = area(p, n);
        int ax = dx >= dy ? (dx >= dz ? 0 : 2) : (dy >= dz ? 1 : 2);
        const double ext = hi[ax] - lo[ax];
        if (depth >= max_depth || !(ha > ratio * pa) || !(ext > 0) || !std::isfinite(ha)) { emit(p, n, full, key, rb, rk); return; }
        const double c = 0.5 * (lo[ax] + hi[ax]);
        P3 a[12], b[12];
        const int na = clip(p, n, ax, c, true, a), nb = clip(p, n, ax, c, false, b);
        if (na < 3 || nb < 3 || na > 10 || nb > 10) { emit(p, n, full, key, rb, rk); return; }    // degenerate cut: keep the piece whole
        split(a, na, full, key, depth + 1, max_depth, ratio, rb, rk);
        split(b, nb, full, key, depth + 1, max_depth, ratio, rb, rk);
    }
    // pb: box per scan key; tri(key): the 9 vertex floats of a triangle key, nullptr for a sphere (never split).
    // References come out in key order (pieces of one triangle adjacent), independent of the thread count.
    template <class TRI>
    static void run(const RtSahBox* pb, int32_t n, TRI tri, double ratio, int max_depth, int threads,
                    std::vector<RtSahBox>& rb, std::vector<int32_t>& rk) {
        threads = threads < 1 ? 1 : (threads > 64 ? 64 : threads);
        if (n < (1 << 14)) threads = 1;
        std::vector<std::vector<RtSahBox>> tb((size_t)threads); std::vector<std::vector<int32_t>> tk((size_t)threads);
        auto work = [&](int t) {
            const int32_t k0 = (int32_t)((int64_t)n * t / threads), k1 = (int32_t)((int64_t)n * (t + 1) / threads);
            tb[(size_t)t].reserve((size_t)(k1 - k0) * 2); tk[(size_t)t].reserve((size_t)(k1 - k0) * 2);
            for (int32_t key = k0; key < k1; ++key) {
                const float* v = tri(key);
                bool ok = v != nullptr;
                for (int k = 0; ok && k < 9; ++k) ok = std::isfinite(v[k]);
                if (!ok) { tb[(size_t)t].push_back(pb[key]); tk[(size_t)t].push_back(key); continue; }
                P3 p[3];
                for (int i = 0; i < 3; ++i) for (int k = 0; k < 3; ++k) p[i].x[k] = (double)v[3 * i + k];
                split(p, 3, pb[key], key, 0, max_depth, ratio, tb[(size_t)t], tk[(size_t)t]);
            }
        };
        std::vector<std::thread> pool;
        for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
        work(0);
        for (auto& th : pool) th.join();
        rb.clear(); rk.clear();
        for (int t = 0; t < threads; ++t) { rb.insert(rb.end(), tb[(size_t)t].begin(), tb[(size_t)t].end()); rk.insert(rk.end(), tk[(size_t)t].begin(), tk[(size_t)t].end()); }
    }
};

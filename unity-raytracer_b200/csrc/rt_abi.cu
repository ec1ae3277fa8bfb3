// rt_abi.cu — the extern "C" boundary of librtb200.so (include/rtb200.h): context, scene staging, commit, render,
// multi-device band partition and the peer-store framebuffer gather.  No exceptions cross this boundary.
#include "rt_internal.h"
#include <new>
#include <chrono>
#include <thread>
#include <atomic>

struct rt_ctx {
    std::vector<RtDevice> devs;
    RtHostScene host;
    RtOptions opt;
    std::string err;
    rt_stats stats{};
    bool in_scene = false, committed = false;
    bool dirty = false;                                     // rt_scene_update_* since the last (re)commit: rendering needs rt_scene_recommit
    int ext_rank = 0, ext_world = 1, band_rows = 32;
    int last_res_x = 0, last_res_y = 0, last_want = 0;
    bool stats_pending = false;
    bool last_local = false;                                // the last render wrote per-device framebuffers (rt_render), not the gather target
    float stream_ms = 0.f; uint64_t stream_bytes = 0;      // streaming commit: time / bytes spent inside rt_scene_mesh
    std::vector<int32_t> aabb_pending;                      // instanced meshes whose Mesh.AABB the library still has to compute (resolve_pending_aabbs)
    std::vector<RtFrameStats> fstats;
};

static_assert(sizeof(rt_render_params) == 88 && sizeof(rt_material) == 56 && sizeof(rt_stats) == 144, "C-ABI struct layout (include/rtb200.h)");

static thread_local std::string g_err;   // rt_last_error(NULL): errors of rt_create

extern "C" { static int commit_impl(rt_ctx* c, const float scene_aabb[6], bool raw_resident, bool refit); }

namespace {

int fail(rt_ctx* c, int code, const std::string& msg) { if (c) c->err = msg; else g_err = msg; return code; }

void device_release(RtDevice& D) {
    if (D.device < 0) return;
    cudaSetDevice(D.device);
    for (DevBuf* b : {&D.raw_v, &D.raw_n, &D.raw_s, &D.nodes, &D.prims, &D.pinfo, &D.mats, &D.mesh_aabb, &D.mesh_base, &D.lights, &D.keys,
                      &D.mesh_raw, &D.inodes, &D.iprims, &D.inst, &D.inst_vbase,
                      &D.keys_alt, &D.vals, &D.vals_alt, &D.cub_tmp, &D.bounds, &D.topo_i32, &D.leaf_lo, &D.leaf_hi, &D.int_lo, &D.int_hi,
                      &D.flags, &D.new_index, &D.accum, &D.stats, &D.own_rgba32f, &D.own_rgba8, &D.own_hit_id, &D.own_hit_t})
        b->release();
    rt_upload_release(D);
    for (auto& L : D.lanes) {
        for (DevBuf* b : {&L.hit0_t, &L.hit0_slot, &L.hitq_t, &L.hitq_slot, &L.rays[0], &L.rays[1], &L.counters, &L.occl, &L.split_occl, &L.split_acc, &L.link_tmp, &L.qs_hist}) b->release();
        for (auto& b : L.local) b.release();
        for (auto& b : L.link) b.release();
        if (L.done) cudaEventDestroy(L.done);
        if (L.own_stream) cudaStreamDestroy(L.own_stream);
        L.done = nullptr; L.own_stream = nullptr;
    }
    for (auto* hh : {&D.hint_primary, &D.hint_shadow}) for (int k = 0; k < 2; ++k) { hh->lists[k].release(); hh->counts[k].release(); }
    for (cudaEvent_t e : D.fin_ev) cudaEventDestroy(e);
    D.fin_ev.clear();
    for (int k = 0; k < 4; ++k) if (D.imported[k] && D.imp[k]) { cudaIpcCloseMemHandle(D.imp[k]); D.imp[k] = nullptr; D.imported[k] = false; }
    for (cudaEvent_t e : D.kev_pool) cudaEventDestroy(e);
    D.kev_pool.clear();
    if (D.ev0) cudaEventDestroy(D.ev0);
    if (D.ev1) cudaEventDestroy(D.ev1);
    for (cudaEvent_t e : D.wave_ev) cudaEventDestroy(e);
    D.wave_ev.clear();
    if (D.copy_stream) cudaStreamDestroy(D.copy_stream);
    if (D.own_stream) cudaStreamDestroy(D.own_stream);
    D.device = -1;
}

size_t out_bytes(int k, int rx, int ry) {
    size_t px = (size_t)rx * (size_t)ry;
    return k == 0 ? px * 16 : (k == 1 ? px * 4 : (k == 2 ? px * 12 : px * 4));
}

// Where the next render writes.  gather (rt_render_device): the full-frame buffers of device 0 (or the ones imported from
// rank 0's process), every device pointing at them — peer stores over NVLink.  local (rt_render, host outputs): every device
// its OWN full-frame-indexed buffers, of which it fills only its bands and copies them out over its own PCIe link.
int ensure_outputs(rt_ctx* c, int rx, int ry, int want, bool local) {
    std::string& err = c->err;
    for (size_t i = 0; i < c->devs.size(); ++i) {
        RtDevice& D = c->devs[i];
        RtDevice& D0 = c->devs[0];
        DevBuf* own[4] = {&D.own_rgba32f, &D.own_rgba8, &D.own_hit_id, &D.own_hit_t};
        void** ptr[4] = {&D.o_rgba32f, &D.o_rgba8, &D.o_hit_id, &D.o_hit_t};
        void* first[4] = {D0.o_rgba32f, D0.o_rgba8, D0.o_hit_id, D0.o_hit_t};
        for (int k = 0; k < 4; ++k) {
            if (!(want & (1 << k))) continue;
            if (!local && i > 0) { *ptr[k] = first[k]; continue; }
            if (!local && D.imported[k]) { *ptr[k] = D.imp[k]; continue; }
            RT_CUDA(cudaSetDevice(D.device));
            RT_CUDA(own[k]->ensure(out_bytes(k, rx, ry)));
            *ptr[k] = own[k]->p;
        }
        D.out_res_x = rx; D.out_res_y = ry;
    }
    return RT_OK;
}

RtView make_view(const rt_ctx* c, const rt_render_params* p, int dev_index) {
    RtView V{};
    V.pos = mk3(p->pos[0], p->pos[1], p->pos[2]); V.fwd = mk3(p->fwd[0], p->fwd[1], p->fwd[2]);
    V.right = mk3(p->right[0], p->right[1], p->right[2]); V.up = mk3(p->up[0], p->up[1], p->up[2]);
    V.res_x = p->res_x; V.res_y = p->res_y; V.dist = p->dist; V.half_h = p->half_h; V.half_v = p->half_v;
    V.bg255 = mk3(p->background_rgb[0] * 255.f, p->background_rgb[1] * 255.f, p->background_rgb[2] * 255.f);   // Rgb.cs:15-18
    V.max_bounces = p->max_reflection_bounces < 0 ? 0 : p->max_reflection_bounces;
    V.spp_grid = p->spp_grid < 1 ? 1 : p->spp_grid;
    const int nd = (int)c->devs.size();
    V.rank = c->ext_rank * nd + dev_index; V.world = c->ext_world * nd; V.band_rows = c->band_rows;
    rt_view_finish(V);
    rt_view_far_origin(V, c->devs[(size_t)dev_index].S.origin_reach);
    return V;
}

int validate_params(rt_ctx* c, const rt_render_params* p) {
    if (!p) return fail(c, RT_E_INVALID, "params is NULL");
    if (p->res_x < 0 || p->res_y < 0) return fail(c, RT_E_INVALID, "negative resolution");
    if ((int64_t)p->res_x * p->res_y > (int64_t)1 << 30) return fail(c, RT_E_UNSUPPORTED, "resolution above 2^30 pixels");
    if (p->spp_grid > 64) return fail(c, RT_E_INVALID, "spp_grid above 64");
    // the reference recurses without a bound (RTS:357-372); a silent clamp would change the picture, so say so instead
    if (p->max_reflection_bounces > RT_MAX_DEPTH && !c->devs.empty() && c->devs[0].has_mirror)
        return fail(c, RT_E_UNSUPPORTED, "max_reflection_bounces above 64 in a scene with mirrors");
    return RT_OK;
}

int finish_stats(rt_ctx* c) {
    if (!c->stats_pending) return RT_OK;
    std::string& err = c->err;
    uint64_t pr = 0, sh = 0, mi = 0, sk = 0; uint32_t launches = 0; float ms = 0.f;
    for (size_t i = 0; i < c->devs.size(); ++i) {
        RtDevice& D = c->devs[i];
        RtFrameStats fs = c->fstats[i];
        int rc = rt_device_read_stats(D, &fs, err);
        if (rc) return rc;
        RT_CUDA(cudaEventSynchronize(D.ev1));
        float m = 0.f; RT_CUDA(cudaEventElapsedTime(&m, D.ev0, D.ev1));
        ms = m > ms ? m : ms;
        pr += fs.primary; sh += fs.shadow; mi += fs.mirror; sk += fs.shadow_skipped; launches += fs.launches;
        for (int k = 0; k < RT_K_COUNT; ++k) {       // max over devices: they run concurrently
            if (i == 0 || fs.ms_kernel[k] > c->stats.ms_kernel[k]) c->stats.ms_kernel[k] = fs.ms_kernel[k];
            c->stats.launches_kernel[k] = (i == 0 ? 0 : c->stats.launches_kernel[k]) + fs.launches_kernel[k];
        }
    }
    c->stats.rays_primary = pr; c->stats.rays_shadow = sh; c->stats.rays_mirror = mi; c->stats.rays_shadow_skipped = sk;
    c->stats.ms_render = ms; c->stats.kernel_launches = launches;
    c->stats_pending = false;
    return RT_OK;
}

int sync_all(rt_ctx* c);

// Sphere boxes are inflated for ray origins inside the origin box of the last commit (rt_build_steps.h rt_sphere_box_radius).
// A camera outside it makes the context remember the position (with some room to move: an orbiting camera settles after a
// few frames) and recompute the boxes through the refit path — same tree, same result, new sphere boxes.
int cover_origin(rt_ctx* c, const float pos[3]) {
    const RtDevice& D0 = c->devs[0];
    if (D0.S.n_spheres <= 0 || D0.S.n_prims <= 0) return RT_OK;
    RtHostScene& H = c->host;
    if (!H.cover_camera(pos, D0.obox)) return RT_OK;
    int rc = sync_all(c);
    if (rc) return rc;
    float keep[6]; for (int k = 0; k < 6; ++k) keep[k] = H.aabb[k];
    const float up = c->stats.ms_commit_upload, bd = c->stats.ms_commit_build; const uint64_t hb = c->stats.h2d_bytes;
    rc = commit_impl(c, keep, true, true);
    c->stats.ms_commit_upload = up; c->stats.ms_commit_build = bd; c->stats.h2d_bytes = hb;     // keep the caller's commit figures
    return rc;
}

// host destinations of rt_render (NULL members: not wanted); null pointer = rt_render_device
struct HostDst { char* p[4]; };

// rt_render: host buffers out.  Every device renders its bands into its own framebuffer in waves of whole pixel rows and
// copies wave k out on its copy stream — over its OWN PCIe link, straight into the caller's arrays — while wave k+1
// renders.  The 4K Color[] frame is 132.7 MB, i.e. 2.3 ms of one PCIe Gen5 link, as long as the render itself: one device
// hides the render behind the copy; N devices divide the copy by N (a frame gathered on one GPU first would leave through
// that GPU's single link and take the same 2.3 ms however many GPUs rendered it).
// A wave is a range of LOCAL rows of the device's band partition; a band is contiguous in the frame.
int copy_rows_out(RtDevice& D, const RtView& V, int lr0, int lr1, char* const dst[4], std::string& err, uint64_t& bytes) {
    char* src[4] = {(char*)D.o_rgba32f, (char*)D.o_rgba8, (char*)D.o_hit_id, (char*)D.o_hit_t};
    const size_t bpp[4] = {16, 4, 12, 4};
    const size_t rx = (size_t)V.res_x;
    for (int r = lr0; r < lr1;) {
        int y0, n;
        if (V.world == 1) { y0 = r; n = lr1 - r; }
        else {
            const int lb = r / V.band_rows, in_band = r % V.band_rows;
            y0 = (lb * V.world + V.rank) * V.band_rows + in_band;
            n = V.band_rows - in_band;
            if (n > lr1 - r) n = lr1 - r;
        }
        if (y0 + n > V.res_y) n = V.res_y - y0;
        if (n <= 0) break;
        for (int k = 0; k < 4; ++k) {
            if (!dst[k]) continue;
            const size_t off = (size_t)y0 * rx * bpp[k], len = (size_t)n * rx * bpp[k];
            RT_CUDA(cudaMemcpyAsync(dst[k] + off, src[k] + off, len, cudaMemcpyDeviceToHost, D.copy_stream));
            bytes += len;
        }
        r += n;
    }
    return RT_OK;
}

// queued right behind the device's launches (same host thread): waits for each wave's event on the device's copy stream
int enqueue_copy_out(rt_ctx* c, size_t i, const RtView& V, const HostDst& dst, std::string& err, uint64_t& bytes) {
    RtDevice& D = c->devs[i];
    RT_CUDA(cudaSetDevice(D.device));
    if (!D.copy_stream) RT_CUDA(cudaStreamCreateWithFlags(&D.copy_stream, cudaStreamNonBlocking));
    if (D.wave_marks.empty()) {                                  // not split: everything after the frame
        RT_CUDA(cudaStreamWaitEvent(D.copy_stream, D.ev1, 0));
        return copy_rows_out(D, V, 0, V.n_local_rows, dst.p, err, bytes);
    }
    for (size_t w = 0; w < D.wave_marks.size(); ++w) {
        RT_CUDA(cudaStreamWaitEvent(D.copy_stream, D.wave_ev[w], 0));
        int rc = copy_rows_out(D, V, D.wave_marks[w].row0, D.wave_marks[w].row1, dst.p, err, bytes);
        if (rc) return rc;
#ifdef RT_WAVE_TRACE
        {   // development aid (-DRT_WAVE_TRACE): when each wave finished rendering / copying, relative to the frame's first event
            static std::vector<cudaEvent_t> cev;
            while (cev.size() <= w) { cudaEvent_t e; cudaEventCreate(&e); cev.push_back(e); }
            cudaEventRecord(cev[w], D.copy_stream);
            if (w + 1 == D.wave_marks.size()) {
                cudaStreamSynchronize(D.copy_stream); cudaStreamSynchronize(D.stream);
                if (getenv("RTB200_WAVE_TRACE")) for (size_t k = 0; k <= w; ++k) {
                    float tr = 0.f, tc = 0.f; cudaEventElapsedTime(&tr, D.ev0, D.wave_ev[k]); cudaEventElapsedTime(&tc, D.ev0, cev[k]);
                    fprintf(stderr, "dev %zu wave %zu rows %d..%d rendered at %.3f ms, copied at %.3f ms\n", i, k, D.wave_marks[k].row0, D.wave_marks[k].row1, tr, tc);
                }
            }
        }
#endif
    }
    return RT_OK;
}

int render_device_impl(rt_ctx* c, const rt_render_params* p, int want, int split = 0, const HostDst* host = nullptr) {
    int rc = validate_params(c, p);
    if (rc) return rc;
    if (!c->committed) return fail(c, RT_E_STATE, "rt_render before rt_scene_commit");
    if ((rc = cover_origin(c, p->pos))) return rc;
    const bool local = host != nullptr;
    c->last_res_x = p->res_x; c->last_res_y = p->res_y; c->last_want = want;
    c->stats.rays_primary = c->stats.rays_shadow = c->stats.rays_mirror = 0;
    c->stats.ms_render = 0.f; c->stats.kernel_launches = 0; c->stats.ms_d2h = 0.f; c->stats.d2h_bytes = 0;
    if ((int64_t)p->res_x * p->res_y == 0) { c->stats_pending = false; return RT_OK; }     // 0x0 ImagePlane: a no-op
    rc = ensure_outputs(c, p->res_x, p->res_y, want, local);
    if (rc) return rc;
    c->last_local = local;
    const size_t nd = c->devs.size();
    c->fstats.assign(nd, RtFrameStats{});
    std::vector<uint64_t> bytes(nd, 0);
    auto one = [&](size_t i, std::string& err) -> int {
        RtView V = make_view(c, p, (int)i);
        int r = rt_device_render(c->devs[i], V, want, c->opt, err, &c->fstats[i], split ? split : (c->opt.waves > 0 ? c->opt.waves : (V.world >= 4 ? 1 : 2)));
        if (r == RT_OK && host) r = enqueue_copy_out(c, i, V, *host, err, bytes[i]);
        return r;
    };
    if (nd == 1) rc = one(0, c->err);
    else {
        // one host thread per device: queueing a frame's launches takes ~0.1 ms of host time, which run one device after the
        // other would start the last of 8 devices most of a frame late
        std::vector<std::string> errs(nd);
        std::vector<int> rcs(nd, RT_OK);
        std::vector<std::thread> pool;
        for (size_t i = 1; i < nd; ++i) pool.emplace_back([&, i]() { rcs[i] = one(i, errs[i]); });
        rcs[0] = one(0, errs[0]);
        for (auto& t : pool) t.join();
        for (size_t i = 0; i < nd && rc == RT_OK; ++i) if (rcs[i]) { rc = rcs[i]; c->err = errs[i]; }
    }
    if (rc) return rc;
    for (uint64_t b : bytes) c->stats.d2h_bytes += b;
    c->stats_pending = true;
    return RT_OK;
}

int sync_all(rt_ctx* c) {
    std::string& err = c->err;
    for (auto& D : c->devs) { RT_CUDA(cudaSetDevice(D.device)); RT_CUDA(cudaStreamSynchronize(D.stream)); }
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_abi_version(void) { return RTB200_ABI_VERSION; }

const char* rt_last_error(const rt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_err.c_str(); }

int rt_create(rt_ctx** out, const int* device_ids, int n_devices) {
    if (!out) return fail(nullptr, RT_E_INVALID, "out is NULL");
    *out = nullptr;
    if (n_devices < 1 || n_devices > 16) return fail(nullptr, RT_E_INVALID, "n_devices must be in 1..16");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count < 1)
        return fail(nullptr, RT_E_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) + " (librtb200 has no CPU fallback)");
    rt_ctx* c = new (std::nothrow) rt_ctx();
    if (!c) return fail(nullptr, RT_E_OOM, "host allocation failed");
    std::string& err = g_err;
    c->devs.resize(n_devices);
    auto bail = [&](int code) { for (auto& D : c->devs) device_release(D); delete c; return code; };
    for (int i = 0; i < n_devices; ++i) {
        RtDevice& D = c->devs[i];
        int dev = device_ids ? device_ids[i] : i;
        if (dev < 0 || dev >= count) { g_err = "device id out of range"; return bail(RT_E_INVALID); }
        D.device = dev;
        cudaDeviceProp prop;
        if ((e = cudaSetDevice(dev)) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, dev)) != cudaSuccess ||
            (e = cudaStreamCreateWithFlags(&D.own_stream, cudaStreamNonBlocking)) != cudaSuccess ||
            (e = cudaEventCreate(&D.ev0)) != cudaSuccess || (e = cudaEventCreate(&D.ev1)) != cudaSuccess) {
            g_err = std::string("device init failed: ") + cudaGetErrorString(e);
            return bail(RT_E_CUDA);
        }
        D.stream = D.own_stream;
        D.sm_count = prop.multiProcessorCount;
        D.rank = i; D.world = n_devices;
    }
    // peer mappings towards the gather device (device_ids[0]) for the fused framebuffer gather
    for (int i = 1; i < n_devices; ++i) {
        int can = 0;
        cudaDeviceCanAccessPeer(&can, c->devs[i].device, c->devs[0].device);
        if (!can) { g_err = "devices cannot access device_ids[0] as a peer"; return bail(RT_E_UNSUPPORTED); }
        cudaSetDevice(c->devs[i].device);
        e = cudaDeviceEnablePeerAccess(c->devs[0].device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { g_err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); return bail(RT_E_CUDA); }
        cudaGetLastError();
    }
    (void)err;
    *out = c;
    return RT_OK;
}

void rt_destroy(rt_ctx* c) {
    if (!c) return;
    for (auto& D : c->devs) { if (D.device >= 0) { cudaSetDevice(D.device); cudaStreamSynchronize(D.stream); } }
    for (auto& D : c->devs) device_release(D);
    delete c;
}

int rt_scene_begin(rt_ctx* c) {
    if (!c) return RT_E_INVALID;
    int rc = sync_all(c);                       // frames still in flight read the buffers the new scene overwrites
    if (rc) return rc;
    c->host.clear(); c->aabb_pending.clear(); c->in_scene = true; c->committed = false; c->dirty = false;
    c->host.builder = c->opt.builder;           // latched: the builder decides where the mesh arrays are staged
    c->host.mesh_on_device = c->opt.stream_upload && c->opt.builder == 0 && c->devs.size() == 1;
    c->stream_ms = 0.f; c->stream_bytes = 0;
    return RT_OK;
}

int rt_scene_spheres(rt_ctx* c, const float* center_r2, const rt_material* mats, int n) {
    if (!c) return RT_E_INVALID;
    if (!c->in_scene) return fail(c, RT_E_STATE, "rt_scene_spheres outside rt_scene_begin/commit");
    if (n < 0 || (n > 0 && (!center_r2 || !mats))) return fail(c, RT_E_INVALID, "rt_scene_spheres: bad arguments");
    c->host.spheres.insert(c->host.spheres.end(), center_r2, center_r2 + 4 * (size_t)n);
    c->host.sphere_mats.insert(c->host.sphere_mats.end(), mats, mats + n);
    return RT_OK;
}

int rt_scene_triangles(rt_ctx* c, const float* v, const float* normals, const rt_material* mats, int n) {
    if (!c) return RT_E_INVALID;
    if (!c->in_scene) return fail(c, RT_E_STATE, "rt_scene_triangles outside rt_scene_begin/commit");
    if (n < 0 || (n > 0 && (!v || !normals || !mats))) return fail(c, RT_E_INVALID, "rt_scene_triangles: bad arguments");
    c->host.loose_v.insert(c->host.loose_v.end(), v, v + 9 * (size_t)n);
    c->host.loose_n.insert(c->host.loose_n.end(), normals, normals + 3 * (size_t)n);
    c->host.loose_mats.insert(c->host.loose_mats.end(), mats, mats + n);
    return RT_OK;
}

int rt_scene_mesh(rt_ctx* c, const float* v, const float* normals, int n_tris, const rt_material* mat, const float aabb[6]) {
    if (!c) return RT_E_INVALID;
    if (!c->in_scene) return fail(c, RT_E_STATE, "rt_scene_mesh outside rt_scene_begin/commit");
    if (n_tris < 0 || !mat || !aabb || (n_tris > 0 && (!v || !normals))) return fail(c, RT_E_INVALID, "rt_scene_mesh: bad arguments");
    if ((int64_t)c->host.mesh_tris + c->host.inst_tris + n_tris > (int64_t)(1 << 28)) return fail(c, RT_E_UNSUPPORTED, "too many primitives (limit 2^28)");
    c->host.push_mesh(-1, *mat, aabb);
    if (c->host.mesh_on_device) {
        auto t0 = std::chrono::steady_clock::now();
        int rc = rt_device_stage_mesh(c->devs[0], v, normals, c->host.mesh_tris, n_tris, c->opt.upload_threads, c->err);
        if (rc) return rc;
        c->stream_ms += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        c->stream_bytes += (uint64_t)n_tris * 48;
    } else {
        c->host.mesh_v.insert(c->host.mesh_v.end(), v, v + 9 * (size_t)n_tris);
        c->host.mesh_n.insert(c->host.mesh_n.end(), normals, normals + 3 * (size_t)n_tris);
    }
    c->host.mesh_tris += n_tris;
    return RT_OK;
}

// ---- instanced meshes: SceneMesh.cs:11-53 without the per-frame world bake (rt_instances.h) ----
int rt_scene_base_mesh(rt_ctx* c, const float* v, int n_tris) {
    if (!c) return RT_E_INVALID;
    if (!c->in_scene) return fail(c, RT_E_STATE, "rt_scene_base_mesh outside rt_scene_begin/commit");
    if (n_tris < 0 || (n_tris > 0 && !v)) return fail(c, RT_E_INVALID, "rt_scene_base_mesh: bad arguments");
    if (n_tris > (1 << 26)) return fail(c, RT_E_UNSUPPORTED, "rt_scene_base_mesh: more than 2^26 triangles");
    RtBaseMesh B;
    const uint64_t hash = RtBaseMesh::hash_bytes(v, 36 * (size_t)n_tris);
    if (!c->host.inst.take_cached(v, 9 * (size_t)n_tris, hash, B)) {
        B.v.assign(v, v + 9 * (size_t)n_tris);
        if (!RtInstanceSet::set_bounds(B)) return fail(c, RT_E_UNSUPPORTED, "rt_scene_base_mesh: vertices must be finite (bake such a mesh and use rt_scene_mesh)");
        B.hash = hash;
        // the BLAS is built here rather than at commit: the instances declared next take their world bounds from its top levels
        int threads = (int)std::thread::hardware_concurrency();
        c->host.inst.build_blas(B, c->opt.leaf_max < 1 ? 1 : (c->opt.leaf_max > 3 ? 3 : c->opt.leaf_max), threads < 1 ? 1 : threads);
    }
    c->host.inst.bases.push_back(std::move(B));
    c->host.inst.blas_changed = true;
    return (int)c->host.inst.bases.size() - 1;
}

// world AABB of an instance exactly as SceneMesh.cs:22-31 computes it: min / max over the baked vertices (Encapsulate, AABB.cs:10-20)
static void instance_aabb(const RtBaseMesh& B, const float m[12], float out[6]) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};       // float.MaxValue / float.MinValue
    const float4 r0 = make_float4(m[0], m[1], m[2], m[3]), r1 = make_float4(m[4], m[5], m[6], m[7]), r2 = make_float4(m[8], m[9], m[10], m[11]);
    for (size_t i = 0; i < B.v.size() / 3; ++i) {
        const f3 w = rt_xform_point(r0, r1, r2, mk3(B.v[3 * i], B.v[3 * i + 1], B.v[3 * i + 2]));
        mn[0] = um_min(mn[0], w.x); mn[1] = um_min(mn[1], w.y); mn[2] = um_min(mn[2], w.z);
        mx[0] = um_max(mx[0], w.x); mx[1] = um_max(mx[1], w.y); mx[2] = um_max(mx[2], w.z);
    }
    for (int k = 0; k < 3; ++k) { out[k] = mn[k]; out[3 + k] = mx[k]; }
}

int rt_scene_mesh_instance(rt_ctx* c, int base_mesh, const float m3x4[12], const rt_material* mat, const float aabb[6]) {
    if (!c) return RT_E_INVALID;
    if (!c->in_scene) return fail(c, RT_E_STATE, "rt_scene_mesh_instance outside rt_scene_begin/commit");
    RtHostScene& H = c->host;
    if (base_mesh < 0 || base_mesh >= (int)H.inst.bases.size() || !m3x4 || !mat) return fail(c, RT_E_INVALID, "rt_scene_mesh_instance: bad arguments");
    const RtBaseMesh& B = H.inst.bases[(size_t)base_mesh];
    if ((int64_t)H.mesh_tris + H.inst_tris + B.n_tris() > (int64_t)(1 << 28)) return fail(c, RT_E_UNSUPPORTED, "too many primitives (limit 2^28)");
    RtInstance I;
    I.base = base_mesh; I.mesh = H.n_meshes();
    if (!RtInstanceSet::set_matrix(I, m3x4, B))
        return fail(c, RT_E_UNSUPPORTED, "rt_scene_mesh_instance: matrix not finite, singular or too ill-conditioned for an instance (bake the mesh and use rt_scene_mesh)");
    float box[6] = {0, 0, 0, 0, 0, 0};
    if (aabb) for (int k = 0; k < 6; ++k) box[k] = aabb[k];
    else c->aabb_pending.push_back(I.mesh);                      // computed at commit, all pending ones in parallel (a pass over the base mesh each)
    H.push_mesh((int32_t)H.inst.insts.size(), *mat, box);
    H.inst.insts.push_back(I);
    H.inst_tris += B.n_tris();
    return RT_OK;
}

// Mesh.AABB of the instances declared / re-posed with aabb == NULL: SceneMesh.cs:22-31 over the baked vertices, one host thread per
// instance up to the core count (cfg4: 125 instances x 245 760 vertices = 95 ms on one thread)
static void resolve_pending_aabbs(rt_ctx* c) {
    RtHostScene& H = c->host;
    std::vector<int32_t> todo;
    todo.swap(c->aabb_pending);
    if (todo.empty()) return;
    // re-posed instances of a committed scene: the base meshes are resident on the device — one pass at L2 bandwidth instead of a host
    // thread's pass per instance (worth the round trip from about 10^5 vertices on)
    int64_t verts = 0;
    for (int32_t mesh : todo) verts += 3 * (int64_t)H.inst.bases[(size_t)H.inst.insts[(size_t)H.mesh_inst[(size_t)mesh]].base].n_tris();
    if (verts >= 100000 && !c->devs.empty()) {          // (rt_device_instance_aabbs refuses when the device does not hold these base meshes)
        std::vector<int32_t> idx(todo.size());
        for (size_t k = 0; k < todo.size(); ++k) idx[k] = H.mesh_inst[(size_t)todo[k]];
        std::vector<float> box(6 * todo.size());
        std::string err;
        if (rt_device_instance_aabbs(c->devs[0], H.inst, idx.data(), (int)idx.size(), box.data(), err) == RT_OK) {
            std::vector<int32_t> redo;
            for (size_t k = 0; k < todo.size(); ++k) {
                bool plain = true;                                   // finite and no zero extremum (whose sign depends on the vertex order)
                for (int j = 0; j < 6; ++j) plain = plain && std::isfinite(box[6 * k + j]) && box[6 * k + j] != 0.0f;
                if (plain) std::memcpy(H.mesh_aabb.data() + 6 * (size_t)todo[k], box.data() + 6 * k, 24);
                else redo.push_back(todo[k]);
            }
            todo.swap(redo);
            if (todo.empty()) return;
        }
    }
    std::atomic<size_t> next{0};
    auto work = [&]() {
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= todo.size()) return;
            const int32_t mesh = todo[k];
            const RtInstance& I = H.inst.insts[(size_t)H.mesh_inst[(size_t)mesh]];
            instance_aabb(H.inst.bases[(size_t)I.base], I.m, H.mesh_aabb.data() + 6 * (size_t)mesh);
        }
    };
    size_t nt = std::thread::hardware_concurrency();
    nt = nt < 1 ? 1 : (nt > todo.size() ? todo.size() : nt);
    std::vector<std::thread> pool;
    for (size_t t = 1; t < nt; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

int rt_scene_lights(rt_ctx* c, const float* pos_intensity, int n, const float ambient[3]) {
    if (!c) return RT_E_INVALID;
    if (!c->in_scene) return fail(c, RT_E_STATE, "rt_scene_lights outside rt_scene_begin/commit");
    if (n < 0 || (n > 0 && !pos_intensity)) return fail(c, RT_E_INVALID, "rt_scene_lights: bad arguments");
    c->host.lights.assign(pos_intensity, pos_intensity + 6 * (size_t)n);
    for (int k = 0; k < 3; ++k) c->host.ambient[k] = ambient ? ambient[k] : 0.f;
    return RT_OK;
}

static int commit_impl(rt_ctx* c, const float scene_aabb[6], bool raw_resident, bool refit) {
    resolve_pending_aabbs(c);
    if (c->host.mesh_stale && !refit && c->host.builder == 1 && !c->host.mesh_on_device) {   // (the builder latched at rt_scene_begin, as rt_device_commit uses)   // host rebuild after device-only mesh updates
        RtDevice& D = c->devs[0];
        std::string& err = c->err;
        const size_t n_mt = (size_t)c->host.n_meshtris();
        if (c->host.mesh_v.size() < 9 * n_mt || c->host.mesh_n.size() < 3 * n_mt || D.raw_v.bytes < n_mt * 36 || D.raw_n.bytes < n_mt * 12)
            return fail(c, RT_E_STATE, "mesh arrays of the committed scene are gone");
        RT_CUDA(cudaSetDevice(D.device));
        RT_CUDA(cudaMemcpyAsync(c->host.mesh_v.data(), D.raw_v.p, n_mt * 36, cudaMemcpyDeviceToHost, D.stream));
        RT_CUDA(cudaMemcpyAsync(c->host.mesh_n.data(), D.raw_n.p, n_mt * 12, cudaMemcpyDeviceToHost, D.stream));
        RT_CUDA(cudaStreamSynchronize(D.stream));
        c->host.mesh_stale = false;
    }
    if (scene_aabb) for (int k = 0; k < 6; ++k) c->host.aabb[k] = scene_aabb[k];
    else c->host.calc_aabb();
    float up = 0.f, bd = 0.f; uint64_t bytes = 0;
    for (auto& D : c->devs) {
        float u = 0.f, b = 0.f; uint64_t by = 0;
        int rc = rt_device_commit(D, c->host, c->opt, c->err, &u, &b, &by, raw_resident, refit);
        if (rc) return rc;
        up += u; bd += b; bytes += by;
    }
    c->stats.ms_commit_upload = up + c->stream_ms; c->stats.ms_commit_build = bd; c->stats.h2d_bytes = bytes + c->stream_bytes;
    c->stats.n_prims = (uint32_t)(c->host.n_prims() + c->host.inst_tris); c->stats.n_nodes = c->devs[0].n_nodes;
    c->stream_ms = 0.f; c->stream_bytes = 0;
    c->in_scene = false; c->committed = true;
    return RT_OK;
}

int rt_scene_commit(rt_ctx* c, const float scene_aabb[6]) {
    if (!c) return RT_E_INVALID;
    if (!c->in_scene) return fail(c, RT_E_STATE, "rt_scene_commit without rt_scene_begin");
    c->host.inst.keep_blas_if_same();        // the same base meshes as the previous scene, in the same order: nothing of them is re-sent
    return commit_impl(c, scene_aabb, false, false);
}

// ---- dirty-tracking updates of a committed scene ----
static int update_prolog(rt_ctx* c, const char* who) {
    if (!c) return RT_E_INVALID;
    if (c->in_scene) return fail(c, RT_E_STATE, std::string(who) + " inside rt_scene_begin/commit");
    if (!c->committed && !c->dirty) return fail(c, RT_E_STATE, std::string(who) + ": no committed scene");
    return RT_OK;
}
// called once the arguments are known to be good: a rejected update leaves the scene clean
static int update_mark(rt_ctx* c) {
    if (!c->dirty) { int rc = sync_all(c); if (rc) return rc; }          // frames in flight still read the old data
    c->dirty = true; c->committed = false;
    return RT_OK;
}
static int update_raw(rt_ctx* c, int what, int64_t first, int64_t n, const float* a, const float* b) {
    auto t0 = std::chrono::steady_clock::now();
    for (auto& D : c->devs) { int rc = rt_device_update_raw(D, what, first, n, a, b, c->opt.upload_threads, c->err); if (rc) return rc; }
    c->stream_ms += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->stream_bytes += (uint64_t)n * (what == 0 ? 48 : 16) * c->devs.size();
    return RT_OK;
}

int rt_scene_update_mesh(rt_ctx* c, int mesh_index, const float* v, const float* normals, const float aabb[6]) {
    int rc = update_prolog(c, "rt_scene_update_mesh");
    if (rc) return rc;
    RtHostScene& H = c->host;
    if (mesh_index < 0 || mesh_index >= H.n_meshes() || !aabb) return fail(c, RT_E_INVALID, "rt_scene_update_mesh: bad arguments");
    if (H.mesh_inst[(size_t)mesh_index] >= 0) return fail(c, RT_E_INVALID, "rt_scene_update_mesh: mesh is an instance (use rt_scene_update_instance)");
    const int64_t first = H.mesh_raw[mesh_index];
    const int64_t n = (mesh_index + 1 < H.n_meshes() ? (int64_t)H.mesh_raw[mesh_index + 1] : H.mesh_tris) - first;
    if (n > 0 && (!v || !normals)) return fail(c, RT_E_INVALID, "rt_scene_update_mesh: bad arguments");
    if ((rc = update_mark(c))) return rc;
    for (int k = 0; k < 6; ++k) H.mesh_aabb[6 * (size_t)mesh_index + k] = aabb[k];
    // the host copies feed the host SAH builder only; keeping them current would cost a second pass over the caller's arrays
    // on every update (host memory bandwidth is what binds an update), so they are refreshed from the device when needed
    if (!H.mesh_on_device && n > 0) H.mesh_stale = true;
    return update_raw(c, 0, first, n, v, normals);
}

int rt_scene_update_instance(rt_ctx* c, int mesh_index, const float m3x4[12], const float aabb[6]) {
    int rc = update_prolog(c, "rt_scene_update_instance");
    if (rc) return rc;
    RtHostScene& H = c->host;
    if (mesh_index < 0 || mesh_index >= H.n_meshes() || !m3x4 || H.mesh_inst[(size_t)mesh_index] < 0)
        return fail(c, RT_E_INVALID, "rt_scene_update_instance: bad arguments (not an instanced mesh?)");
    RtInstance I = H.inst.insts[(size_t)H.mesh_inst[(size_t)mesh_index]];
    const RtBaseMesh& B = H.inst.bases[(size_t)I.base];
    if (!RtInstanceSet::set_matrix(I, m3x4, B))
        return fail(c, RT_E_UNSUPPORTED, "rt_scene_update_instance: matrix not finite, singular or too ill-conditioned for an instance");
    if ((rc = update_mark(c))) return rc;
    H.inst.insts[(size_t)H.mesh_inst[(size_t)mesh_index]] = I;
    if (aabb) for (int k = 0; k < 6; ++k) H.mesh_aabb[6 * (size_t)mesh_index + k] = aabb[k];
    else c->aabb_pending.push_back(mesh_index);
    (void)B;
    return RT_OK;
}

int rt_scene_update_instances(rt_ctx* c, const int32_t* mesh_indices, const float* m3x4, const float* aabbs, int n) {
    int rc = update_prolog(c, "rt_scene_update_instances");
    if (rc) return rc;
    RtHostScene& H = c->host;
    if (n < 0 || (n > 0 && (!mesh_indices || !m3x4))) return fail(c, RT_E_INVALID, "rt_scene_update_instances: bad arguments");
    if (n == 0) return RT_OK;
    // all or nothing: every pose is validated (and its inverse / bounds computed) before the first one is stored
    std::vector<RtInstance> posed((size_t)n);
    for (int i = 0; i < n; ++i) {
        const int m = mesh_indices[i];
        if (m < 0 || m >= H.n_meshes() || H.mesh_inst[(size_t)m] < 0) {
            char b[128]; snprintf(b, sizeof(b), "rt_scene_update_instances: entry %d: mesh %d is not an instanced mesh", i, m);
            return fail(c, RT_E_INVALID, b);
        }
        posed[(size_t)i] = H.inst.insts[(size_t)H.mesh_inst[(size_t)m]];
        if (!RtInstanceSet::set_matrix(posed[(size_t)i], m3x4 + 12 * (size_t)i, H.inst.bases[(size_t)posed[(size_t)i].base])) {
            char b[160]; snprintf(b, sizeof(b), "rt_scene_update_instances: entry %d (mesh %d): matrix not finite, singular or too ill-conditioned for an instance", i, m);
            return fail(c, RT_E_UNSUPPORTED, b);
        }
    }
    if ((rc = update_mark(c))) return rc;
    for (int i = 0; i < n; ++i) {
        const int m = mesh_indices[i];
        H.inst.insts[(size_t)H.mesh_inst[(size_t)m]] = posed[(size_t)i];
        if (aabbs) for (int k = 0; k < 6; ++k) H.mesh_aabb[6 * (size_t)m + k] = aabbs[6 * (size_t)i + k];
        else c->aabb_pending.push_back(m);
    }
    return RT_OK;
}

int rt_scene_update_triangles(rt_ctx* c, int first, int n, const float* v, const float* normals) {
    int rc = update_prolog(c, "rt_scene_update_triangles");
    if (rc) return rc;
    RtHostScene& H = c->host;
    if (first < 0 || n < 0 || (int64_t)first + n > H.n_loose() || (n > 0 && (!v || !normals))) return fail(c, RT_E_INVALID, "rt_scene_update_triangles: bad range");
    if (n == 0) return RT_OK;
    if ((rc = update_mark(c))) return rc;
    std::memcpy(H.loose_v.data() + 9 * (size_t)first, v, (size_t)n * 36);
    std::memcpy(H.loose_n.data() + 3 * (size_t)first, normals, (size_t)n * 12);
    return update_raw(c, 0, (int64_t)H.n_meshtris() + first, n, v, normals);
}

int rt_scene_update_spheres(rt_ctx* c, int first, int n, const float* center_r2) {
    int rc = update_prolog(c, "rt_scene_update_spheres");
    if (rc) return rc;
    RtHostScene& H = c->host;
    if (first < 0 || n < 0 || (int64_t)first + n > H.n_spheres() || (n > 0 && !center_r2)) return fail(c, RT_E_INVALID, "rt_scene_update_spheres: bad range");
    if (n == 0) return RT_OK;
    if ((rc = update_mark(c))) return rc;
    std::memcpy(H.spheres.data() + 4 * (size_t)first, center_r2, (size_t)n * 16);
    return update_raw(c, 1, first, n, center_r2, nullptr);
}

int rt_scene_update_lights(rt_ctx* c, const float* pos_intensity, int n, const float ambient[3]) {
    int rc = update_prolog(c, "rt_scene_update_lights");
    if (rc) return rc;
    if (n < 0 || (n > 0 && !pos_intensity)) return fail(c, RT_E_INVALID, "rt_scene_update_lights: bad arguments");
    if ((rc = update_mark(c))) return rc;
    c->host.lights.assign(pos_intensity, pos_intensity + 6 * (size_t)n);
    if (ambient) for (int k = 0; k < 3; ++k) c->host.ambient[k] = ambient[k];
    return RT_OK;
}

int rt_scene_recommit(rt_ctx* c, const float scene_aabb[6]) {
    if (!c) return RT_E_INVALID;
    if (c->in_scene) return fail(c, RT_E_STATE, "rt_scene_recommit inside rt_scene_begin/commit");
    if (!c->dirty && !c->committed) return fail(c, RT_E_STATE, "rt_scene_recommit: no committed scene");
    if (!c->dirty) { int rc = sync_all(c); if (rc) return rc; }
    int rc = commit_impl(c, scene_aabb, true, false);
    if (rc == RT_OK) c->dirty = false;
    return rc;
}

int rt_scene_refit(rt_ctx* c, const float scene_aabb[6]) {
    if (!c) return RT_E_INVALID;
    if (c->in_scene) return fail(c, RT_E_STATE, "rt_scene_refit inside rt_scene_begin/commit");
    if (!c->dirty && !c->committed) return fail(c, RT_E_STATE, "rt_scene_refit: no committed scene");
    if (!c->dirty) { int rc = sync_all(c); if (rc) return rc; }
    int rc = commit_impl(c, scene_aabb, true, true);
    if (rc == RT_OK) c->dirty = false;
    return rc;
}

int rt_scene_get_aabb(rt_ctx* c, float out[6]) {
    if (!c || !out) return RT_E_INVALID;
    if (!c->committed) return fail(c, RT_E_STATE, "scene not committed");
    for (int k = 0; k < 6; ++k) out[k] = c->host.aabb[k];
    return RT_OK;
}

int rt_render_device(rt_ctx* c, const rt_render_params* p, int want, rt_stats* stats) {
    if (!c) return RT_E_INVALID;
    int rc = render_device_impl(c, p, want);
    if (rc) return rc;
    if (stats) { rc = rt_get_stats(c, stats); if (rc) return rc; }
    return RT_OK;
}

int rt_read_outputs(rt_ctx* c, float* rgba32f, uint8_t* rgba8, int32_t* hit_id, float* hit_t) {
    if (!c) return RT_E_INVALID;
    std::string& err = c->err;
    const int rx = c->last_res_x, ry = c->last_res_y;
    if ((int64_t)rx * ry == 0) return RT_OK;
    int rc = sync_all(c);
    if (rc) return rc;
    RtDevice& D0 = c->devs[0];
    RT_CUDA(cudaSetDevice(D0.device));
    if (c->last_local && (c->devs.size() > 1 || c->ext_world > 1))
        return fail(c, RT_E_STATE, "rt_read_outputs after rt_render of a partition: the frame was not gathered (use rt_render_device)");
    void* dst[4] = {rgba32f, rgba8, hit_id, hit_t};
    void* src[4] = {D0.o_rgba32f, D0.o_rgba8, D0.o_hit_id, D0.o_hit_t};
    auto t0 = std::chrono::steady_clock::now();
    uint64_t bytes = 0;
    for (int k = 0; k < 4; ++k) {
        if (!dst[k]) continue;
        if (!(c->last_want & (1 << k)) || !src[k]) return fail(c, RT_E_STATE, "requested output was not rendered");
        RT_CUDA(cudaMemcpyAsync(dst[k], src[k], out_bytes(k, rx, ry), cudaMemcpyDeviceToHost, D0.stream));
        bytes += out_bytes(k, rx, ry);
    }
    RT_CUDA(cudaStreamSynchronize(D0.stream));
    c->stats.ms_d2h = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->stats.d2h_bytes = bytes;
    return RT_OK;
}

int rt_render(rt_ctx* c, const rt_render_params* p, float* rgba32f, uint8_t* rgba8, int32_t* hit_id, float* hit_t, rt_stats* stats) {
    if (!c) return RT_E_INVALID;
    int want = (rgba32f ? RT_OUT_RGBA32F : 0) | (rgba8 ? RT_OUT_RGBA8 : 0) | (hit_id ? RT_OUT_HIT_ID : 0) | (hit_t ? RT_OUT_HIT_T : 0);
    const int world = c->ext_world * (int)c->devs.size();
    int cw = world == 1 ? c->opt.copy_waves : (c->opt.copy_waves_part > 0 ? c->opt.copy_waves_part : (world == 2 ? 4 : 2));
    const bool pipelined = want && cw > 1 && p && p->res_y / world >= 64;
    const HostDst dst{{(char*)rgba32f, (char*)rgba8, (char*)hit_id, (char*)hit_t}};
    auto t0 = std::chrono::steady_clock::now();
    int rc = render_device_impl(c, p, want, pipelined ? -cw : 0, &dst);
    if (rc) return rc;
    if (p && (int64_t)p->res_x * p->res_y > 0) {
        std::string& err = c->err;
        for (auto& D : c->devs) {
            RT_CUDA(cudaSetDevice(D.device));
            if (D.copy_stream) RT_CUDA(cudaStreamSynchronize(D.copy_stream));
            RT_CUDA(cudaStreamSynchronize(D.stream));
        }
        c->stats.ms_d2h = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();   // launches + copies, wall
    }
    if (stats) { rc = rt_get_stats(c, stats); if (rc) return rc; }
    return RT_OK;
}

int rt_host_register(rt_ctx* c, void* ptr, size_t bytes) {
    if (!c) return RT_E_INVALID;
    if (!ptr || bytes == 0) return fail(c, RT_E_INVALID, "rt_host_register: bad arguments");
    std::string& err = c->err;
    RT_CUDA(cudaSetDevice(c->devs[0].device));
    RT_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return RT_OK;
}

int rt_host_unregister(rt_ctx* c, void* ptr) {
    if (!c) return RT_E_INVALID;
    if (!ptr) return fail(c, RT_E_INVALID, "rt_host_unregister: bad arguments");
    std::string& err = c->err;
    int rc = sync_all(c);                                   // copies still in flight use the registration
    if (rc) return rc;
    RT_CUDA(cudaHostUnregister(ptr));
    return RT_OK;
}

int rt_device_outputs(rt_ctx* c, void** rgba32f, void** rgba8, void** hit_id, void** hit_t) {
    if (!c) return RT_E_INVALID;
    RtDevice& D0 = c->devs[0];
    if (rgba32f) *rgba32f = D0.o_rgba32f;
    if (rgba8) *rgba8 = D0.o_rgba8;
    if (hit_id) *hit_id = D0.o_hit_id;
    if (hit_t) *hit_t = D0.o_hit_t;
    return RT_OK;
}

int rt_get_stats(rt_ctx* c, rt_stats* stats) {
    if (!c || !stats) return RT_E_INVALID;
    int rc = finish_stats(c);
    if (rc) return rc;
    *stats = c->stats;
    return RT_OK;
}

int rt_intersect(rt_ctx* c, const float* origins, const float* dirs, int n, int32_t* out_id, float* out_t) {
    if (!c) return RT_E_INVALID;
    if (!c->committed) return fail(c, RT_E_STATE, "rt_intersect before rt_scene_commit");
    if (n < 0 || (n > 0 && (!origins || !dirs || !out_id || !out_t))) return fail(c, RT_E_INVALID, "rt_intersect: bad arguments");
    if (n == 0) return RT_OK;
    if (c->devs[0].S.n_spheres > 0) {                          // explicit origins belong to the origin box too (both corners: it is convex)
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int i = 0; i < n; ++i)
            for (int k = 0; k < 3; ++k) { const float v = origins[3 * (size_t)i + k]; if (std::isfinite(v)) { lo[k] = fminf(lo[k], v); hi[k] = fmaxf(hi[k], v); } }
        for (int k = 0; k < 3; ++k) if (!(lo[k] <= hi[k])) { lo[k] = 0.f; hi[k] = 0.f; }
        int rc0 = cover_origin(c, lo);
        if (rc0 == RT_OK) rc0 = cover_origin(c, hi);
        if (rc0) return rc0;
    }
    std::string& err = c->err;
    RtDevice& D = c->devs[0];
    RT_CUDA(cudaSetDevice(D.device));
    DevBuf o, d, ids, t;
    auto cleanup = [&]() { o.release(); d.release(); ids.release(); t.release(); };
    cudaError_t e;
    if ((e = o.ensure((size_t)n * 12)) || (e = d.ensure((size_t)n * 12)) || (e = ids.ensure((size_t)n * 12)) || (e = t.ensure((size_t)n * 4))) {
        cleanup(); return fail(c, RT_E_OOM, cudaGetErrorString(e));
    }
    int rc = RT_OK;
    if ((e = cudaMemcpyAsync(o.p, origins, (size_t)n * 12, cudaMemcpyHostToDevice, D.stream)) ||
        (e = cudaMemcpyAsync(d.p, dirs, (size_t)n * 12, cudaMemcpyHostToDevice, D.stream))) rc = fail(c, RT_E_CUDA, cudaGetErrorString(e));
    if (!rc) rc = rt_device_intersect(D, o.as<float>(), d.as<float>(), n, ids.as<int32_t>(), t.as<float>(), c->opt, c->err);
    if (!rc && ((e = cudaMemcpyAsync(out_id, ids.p, (size_t)n * 12, cudaMemcpyDeviceToHost, D.stream)) ||
                (e = cudaMemcpyAsync(out_t, t.p, (size_t)n * 4, cudaMemcpyDeviceToHost, D.stream)) ||
                (e = cudaStreamSynchronize(D.stream)))) rc = fail(c, RT_E_CUDA, cudaGetErrorString(e));
    cleanup();
    return rc;
}

// no counterpart in the reference (include/rtb200.h): rt_intersect with its staging in a static device array, no allocation
int rt_intersect_static(rt_ctx* c, const float* origins, const float* dirs, int n, int32_t* out_id, float* out_t) {
    if (!c) return RT_E_INVALID;
    if (!c->committed) return fail(c, RT_E_STATE, "rt_intersect_static before rt_scene_commit");
    if (n < 0 || (n > 0 && (!origins || !dirs || !out_id || !out_t))) return fail(c, RT_E_INVALID, "rt_intersect_static: bad arguments");
    if (n > RT_INTERSECT_STATIC_MAX)
        return fail(c, RT_E_INVALID, "rt_intersect_static: n = " + std::to_string(n) + " exceeds RT_INTERSECT_STATIC_MAX = " + std::to_string(RT_INTERSECT_STATIC_MAX) +
                                         " rays (the static scratch array); split the batch or use rt_intersect");
    if (n == 0) return RT_OK;
    return rt_device_intersect_static(c->devs[0], origins, dirs, n, out_id, out_t, c->err);
}

int rt_set_stream(rt_ctx* c, void* cuda_stream) {
    if (!c) return RT_E_INVALID;
    RtDevice& D = c->devs[0];
    std::string& err = c->err;
    RT_CUDA(cudaSetDevice(D.device));
    RT_CUDA(cudaStreamSynchronize(D.stream));                 // work queued on the old stream reads buffers the next call may touch
    D.stream = cuda_stream ? (cudaStream_t)cuda_stream : D.own_stream;
    return RT_OK;
}

int rt_set_partition(rt_ctx* c, int rank, int world, int band_rows) {
    if (!c) return RT_E_INVALID;
    if (world < 1 || rank < 0 || rank >= world || band_rows < 4 || (band_rows % 4) != 0) return fail(c, RT_E_INVALID, "rt_set_partition: bad arguments");
    c->ext_rank = rank; c->ext_world = world; c->band_rows = band_rows;
    return RT_OK;
}

int rt_alloc_outputs(rt_ctx* c, int res_x, int res_y, int want) {
    if (!c) return RT_E_INVALID;
    if (res_x < 0 || res_y < 0) return fail(c, RT_E_INVALID, "negative resolution");
    return ensure_outputs(c, res_x, res_y, want, false);
}

int rt_export_outputs(rt_ctx* c, void* handles) {
    if (!c || !handles) return RT_E_INVALID;
    std::string& err = c->err;
    RtDevice& D0 = c->devs[0];
    RT_CUDA(cudaSetDevice(D0.device));
    memset(handles, 0, 4 * sizeof(cudaIpcMemHandle_t));
    void* ptr[4] = {D0.own_rgba32f.p, D0.own_rgba8.p, D0.own_hit_id.p, D0.own_hit_t.p};
    for (int k = 0; k < 4; ++k) {
        if (!ptr[k] || D0.imported[k]) continue;
        cudaIpcMemHandle_t h;
        RT_CUDA(cudaIpcGetMemHandle(&h, ptr[k]));
        memcpy((char*)handles + k * sizeof(h), &h, sizeof(h));
    }
    return RT_OK;
}

int rt_import_outputs(rt_ctx* c, const void* handles, int res_x, int res_y, int want) {
    if (!c || !handles) return RT_E_INVALID;
    std::string& err = c->err;
    RtDevice& D0 = c->devs[0];
    RT_CUDA(cudaSetDevice(D0.device));
    void** ptr[4] = {&D0.o_rgba32f, &D0.o_rgba8, &D0.o_hit_id, &D0.o_hit_t};
    for (int k = 0; k < 4; ++k) {
        if (!(want & (1 << k))) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)handles + k * sizeof(h), sizeof(h));
        if (D0.imported[k] && D0.imp[k]) { cudaIpcCloseMemHandle(D0.imp[k]); D0.imp[k] = nullptr; *ptr[k] = nullptr; D0.imported[k] = false; }
        void* p = nullptr;
        RT_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        *ptr[k] = p; D0.imp[k] = p; D0.imported[k] = true;
    }
    D0.out_res_x = res_x; D0.out_res_y = res_y;
    for (size_t i = 1; i < c->devs.size(); ++i) {
        RtDevice& D = c->devs[i];
        D.o_rgba32f = D0.o_rgba32f; D.o_rgba8 = D0.o_rgba8; D.o_hit_id = D0.o_hit_id; D.o_hit_t = D0.o_hit_t;
    }
    return RT_OK;
}

int rt_set_option(rt_ctx* c, const char* key, double value) {
    if (!c || !key) return RT_E_INVALID;
    std::string k(key);
    if (k == "chunk_pixels") { if (value < 32) return fail(c, RT_E_INVALID, "chunk_pixels < 32"); c->opt.chunk_pixels = ((int)value / 32) * 32; }
    else if (k == "leaf_max") { if (value < 1 || value > 8) return fail(c, RT_E_INVALID, "leaf_max must be 1..8"); c->opt.leaf_max = (int)value; }
    else if (k == "presplit") { if (value < 0) return fail(c, RT_E_INVALID, "presplit < 0"); c->opt.presplit = value; }
    else if (k == "presplit_depth") { if (value < 0 || value > 6) return fail(c, RT_E_INVALID, "presplit_depth must be 0..6"); c->opt.presplit_depth = (int)value; }
    else if (k == "sah_term") { if (value < 0) return fail(c, RT_E_INVALID, "sah_term < 0"); c->opt.sah_term = value; }
    else if (k == "pad_scale") { if (value < 0) return fail(c, RT_E_INVALID, "pad_scale < 0"); c->opt.pad_scale = value; }
    else if (k == "t_slack") { if (value < 1) return fail(c, RT_E_INVALID, "t_slack < 1"); c->opt.t_slack = value; }
    else if (k == "ctas_per_sm") { c->opt.ctas_per_sm = (int)value; }
    else if (k == "block") { c->opt.block = (int)value; }
    else if (k == "waves") { if (value < 0 || value > 64) return fail(c, RT_E_INVALID, "waves must be 0 (auto) or 1..64"); c->opt.waves = (int)value; }
    else if (k == "lanes") { if (value < 1 || value > RtDevice::RT_MAX_LANES) return fail(c, RT_E_INVALID, "lanes must be 1..8"); c->opt.lanes = (int)value; }
    else if (k == "deep_lanes") { if (value < 1 || value > RtDevice::RT_MAX_LANES) return fail(c, RT_E_INVALID, "deep_lanes must be 1..8"); c->opt.deep_lanes = (int)value; }
    else if (k == "fuse_shade") { c->opt.fuse_shade = value != 0; }
    else if (k == "prefetch") { c->opt.prefetch = value != 0; }
    else if (k == "prefetch_max_mb") { c->opt.prefetch_max_mb = (int)value; }
    else if (k == "upload_threads") { c->opt.upload_threads = (int)value; }
    else if (k == "stream_upload") { c->opt.stream_upload = value != 0; }
    else if (k == "backcull") { c->opt.backcull = value != 0; }
    else if (k == "light_split") { c->opt.light_split = (int)value; }
    else if (k == "hints") { c->opt.hints = value != 0; }
    else if (k == "sort_queue") { c->opt.sort_queue = value < 0 ? -1 : (int)value; }
    else if (k == "copy_plan") { if (value < 0) return fail(c, RT_E_INVALID, "copy_plan < 0"); c->opt.copy_plan = (long long)value; }
    else if (k == "copy_waves_part") { if (value < 0 || value > 64) return fail(c, RT_E_INVALID, "copy_waves_part must be 0 (auto) or 1..64"); c->opt.copy_waves_part = (int)value; }
    else if (k == "copy_waves") { if (value < 1 || value > 64) return fail(c, RT_E_INVALID, "copy_waves must be 1..64"); c->opt.copy_waves = (int)value; }
    else if (k == "builder") { if (value != 0 && value != 1) return fail(c, RT_E_INVALID, "builder must be 0 (device LBVH) or 1 (host SAH)"); c->opt.builder = (int)value; }
    else if (k == "kernel_timing") { c->opt.kernel_timing = value != 0; }
    else if (k == "packet") { c->opt.packet = value != 0; for (auto& D : c->devs) D.S.packet = c->opt.packet; }
    else return fail(c, RT_E_INVALID, "unknown option: " + k);
    return RT_OK;
}

}  // extern "C"

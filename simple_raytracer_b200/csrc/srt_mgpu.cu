// srt_mgpu.cu — several GPUs driven from one host thread through the C ABI (include/srt.h, srt_mgpu_*).
// Host-side plumbing over srt_context / srt_render: no kernels of its own.  Pixels are independent
// (simple_raytracer.cpp:511-523) and frames are independent (:534), so both shardings are exact.
//
// Single frame: the frame lives on the first GPU; GPU r renders the interleaved row blocks r, r+R, ... with the frame's
// planes as its output pointers — the render kernels' stores go through peer-mapped memory (NVLink / NVSwitch) straight
// into the first GPU's buffer.  There is no staging copy and no collective: the only synchronisation is one event per GPU
// that the first GPU's stream waits for.  Without peer access a GPU renders into a local frame and its blocks are moved
// with cudaMemcpyPeerAsync.
#include <cuda_runtime.h>

#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "srt.h"
#include "srt_internal.hpp"

using srti::DevBuf;
using srti::fail;

namespace {
constexpr int kLanesPerDevice = 2;      // image series: frames in flight per GPU
}

struct srt_mgpu {
    struct Gpu {
        int device = 0;
        srt_context* ctx = nullptr;
        srt_scene* scene = nullptr;          // the replicated scene (single-frame mode)
        bool peer = false;                   // its kernels may write into the first GPU's memory
        cudaEvent_t done = nullptr;
        DevBuf rgb, obj, tri, t;             // !peer: local full-size planes
        srt_context* lane_ctx[kLanesPerDevice] = {};      // image series
        srt_scene* lane_scene[kLanesPerDevice] = {};
    };
    std::vector<Gpu> gpus;
    DevBuf rgb, obj, tri, t;                 // the frame on the first GPU (for host output planes)
    cudaEvent_t fork = nullptr, t0 = nullptr, t1 = nullptr;
    bool have_frame = false;
};

extern "C" {

void srt_mgpu_destroy(srt_mgpu* m) {
    if (!m) return;
    for (auto& g : m->gpus) {
        if (!g.ctx) continue;
        cudaSetDevice(g.device);
        cudaStreamSynchronize(g.ctx->stream);
        for (int l = 0; l < kLanesPerDevice; ++l) {
            if (g.lane_scene[l]) srt_scene_destroy(g.lane_scene[l]);
            if (g.lane_ctx[l]) srt_context_destroy(g.lane_ctx[l]);
        }
        if (g.scene) srt_scene_destroy(g.scene);
        g.rgb.release(); g.obj.release(); g.tri.release(); g.t.release();
        if (g.done) cudaEventDestroy(g.done);
    }
    if (!m->gpus.empty() && m->gpus[0].ctx) {
        cudaSetDevice(m->gpus[0].device);
        m->rgb.release(); m->obj.release(); m->tri.release(); m->t.release();
        if (m->fork) cudaEventDestroy(m->fork);
        if (m->t0) cudaEventDestroy(m->t0);
        if (m->t1) cudaEventDestroy(m->t1);
    }
    for (auto& g : m->gpus)
        if (g.ctx) srt_context_destroy(g.ctx);
    delete m;
}

int srt_mgpu_create(const int* devices, int n_devices, srt_mgpu** out) {
    if (!out) return fail(SRT_ERR_INVALID_ARG, "out is NULL");
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64) return fail(SRT_ERR_INVALID_ARG, "n_devices %d outside [1,64]", n_devices);
    srt_mgpu* m = new srt_mgpu();
    m->gpus.resize((size_t)n_devices);
    int rc = SRT_OK;
    for (int r = 0; r < n_devices && rc == SRT_OK; ++r) {
        srt_mgpu::Gpu& g = m->gpus[(size_t)r];
        g.device = devices ? devices[r] : r;
        rc = srt_context_create(g.device, &g.ctx);
        if (rc) break;
        if (cudaEventCreateWithFlags(&g.done, cudaEventDisableTiming) != cudaSuccess) { rc = fail(SRT_ERR_CUDA, "cudaEventCreate failed"); break; }
        const int first = m->gpus[0].device;
        if (g.device == first) {
            g.peer = true;
        } else {
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, g.device, first) != cudaSuccess) { cudaGetLastError(); can = 0; }
            if (can) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(first, 0);      // current device = g.device (srt_context_create)
                if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) g.peer = true;
                cudaGetLastError();
            }
        }
    }
    if (rc == SRT_OK) {
        cudaSetDevice(m->gpus[0].device);
        if (cudaEventCreateWithFlags(&m->fork, cudaEventDisableTiming) != cudaSuccess || cudaEventCreate(&m->t0) != cudaSuccess ||
            cudaEventCreate(&m->t1) != cudaSuccess)
            rc = fail(SRT_ERR_CUDA, "cudaEventCreate failed");
    }
    if (rc != SRT_OK) { const std::string keep = srti::g_err; srt_mgpu_destroy(m); srti::g_err = keep; return rc; }
    *out = m;
    return SRT_OK;
}

int srt_mgpu_info(srt_mgpu* m, int* n_devices, int* peer_writes) {
    if (!m) return fail(SRT_ERR_INVALID_ARG, "mgpu is NULL");
    if (n_devices) *n_devices = (int)m->gpus.size();
    if (peer_writes) {
        int n = 0;
        for (auto& g : m->gpus) n += g.peer ? 1 : 0;
        *peer_writes = n;
    }
    return SRT_OK;
}

int srt_mgpu_scene_upload(srt_mgpu* m, const srt_scene_desc* desc) {
    if (!m || !desc) return fail(SRT_ERR_INVALID_ARG, "NULL argument");
    for (auto& g : m->gpus) {
        const int rc = g.scene ? srt_scene_update(g.ctx, g.scene, desc) : srt_scene_upload(g.ctx, desc, &g.scene);
        if (rc) return rc;
    }
    return SRT_OK;
}

int srt_mgpu_sync(srt_mgpu* m) {
    if (!m) return fail(SRT_ERR_INVALID_ARG, "mgpu is NULL");
    for (auto& g : m->gpus) {
        int rc = srt_sync(g.ctx);
        for (int l = 0; l < kLanesPerDevice && !rc; ++l)
            if (g.lane_ctx[l]) rc = srt_sync(g.lane_ctx[l]);
        if (rc) return rc;
    }
    return SRT_OK;
}

int srt_mgpu_render_frame(srt_mgpu* m, const srt_render_params* params, const srt_frame* out, int row_block) {
    if (!m || !params || !out) return fail(SRT_ERR_INVALID_ARG, "NULL argument");
    if (row_block == 0) row_block = 16;
    if (row_block < 0 || row_block % srtk::kTileH) return fail(SRT_ERR_INVALID_ARG, "row_block %d must be a positive multiple of %d", row_block, srtk::kTileH);
    if (params->width <= 0 || params->height <= 0) return fail(SRT_ERR_INVALID_ARG, "bad image size %dx%d", params->width, params->height);
    const int R = (int)m->gpus.size();
    srt_mgpu::Gpu& g0 = m->gpus[0];
    if (!g0.scene) return fail(SRT_ERR_INVALID_ARG, "srt_mgpu_render_frame: no scene (srt_mgpu_scene_upload)");
    const size_t npix = (size_t)params->width * params->height;
    int rb = params->row_begin, re = params->row_end;
    if (rb == 0 && re == 0) re = params->height;
    if (rb < 0 || re > params->height || rb > re) return fail(SRT_ERR_INVALID_ARG, "bad row range [%d,%d)", rb, re);

    // where the frame lives on the first GPU: the caller's device planes, or internal ones for host planes
    struct Plane { void* user; DevBuf* own; size_t elem; void* dev; bool host; };
    Plane planes[4] = {{out->rgb, &m->rgb, 3, nullptr, false}, {out->obj_id, &m->obj, 4, nullptr, false},
                       {out->tri_id, &m->tri, 4, nullptr, false}, {out->t, &m->t, 4, nullptr, false}};
    CU(cudaSetDevice(g0.device));
    for (auto& pl : planes) {
        if (!pl.user) continue;
        if (srti::is_device_pointer(pl.user)) { pl.dev = pl.user; continue; }
        pl.host = true;
        const int rc = pl.own->reserve(npix * pl.elem);
        if (rc) return rc;
        pl.dev = pl.own->p;
    }
    CU(cudaEventRecord(m->t0, g0.ctx->stream));
    CU(cudaEventRecord(m->fork, g0.ctx->stream));           // every GPU starts after what is queued on the first one's stream
    // One GPU's share: wait for the fork, render its blocks into the frame (or locally + peer copies), record `done`.
    auto issue = [&](int r) -> int {
        srt_mgpu::Gpu& g = m->gpus[(size_t)r];
        CU(cudaSetDevice(g.device));
        if (r > 0) CU(cudaStreamWaitEvent(g.ctx->stream, m->fork, 0));
        srt_render_params p = *params;
        p.row_begin = rb; p.row_end = re;
        p.row_block = row_block; p.shard_index = r; p.shard_count = R;
        srt_frame fr = {(uint8_t*)planes[0].dev, (int32_t*)planes[1].dev, (int32_t*)planes[2].dev, (float*)planes[3].dev};
        DevBuf* local[4] = {&g.rgb, &g.obj, &g.tri, &g.t};
        if (!g.peer) {                                      // no peer mapping: render here, move the blocks afterwards
            void** slots[4] = {(void**)&fr.rgb, (void**)&fr.obj_id, (void**)&fr.tri_id, (void**)&fr.t};
            for (int k = 0; k < 4; ++k) {
                if (!planes[k].dev) continue;
                const int rc = local[k]->reserve(npix * planes[k].elem);
                if (rc) return rc;
                *slots[k] = local[k]->p;
            }
        }
        const int rc = srt_render(g.ctx, g.scene, &p, &fr);
        if (rc) return rc;
        if (!g.peer) {
            for (int r0 = rb + r * row_block; r0 < re; r0 += row_block * R) {
                const int r1 = r0 + row_block < re ? r0 + row_block : re;
                const size_t at = (size_t)r0 * params->width, n = (size_t)(r1 - r0) * params->width;
                for (int k = 0; k < 4; ++k)
                    if (planes[k].dev)
                        CU(cudaMemcpyPeerAsync((char*)planes[k].dev + at * planes[k].elem, g0.device, (char*)local[k]->p + at * planes[k].elem,
                                               g.device, n * planes[k].elem, g.ctx->stream));
            }
        }
        if (r > 0) CU(cudaEventRecord(g.done, g.ctx->stream));
        return SRT_OK;
    };
    // A frame is a dozen launches per GPU: issued from one thread, eight GPUs start one after the other and the last one
    // starts ~0.4 ms late.  Every GPU but the first gets its own short-lived issuing thread (distinct contexts, distinct
    // devices: no shared state but the read-only frame description).
    std::vector<int> rcs((size_t)R, SRT_OK);
    std::vector<std::string> errs((size_t)R);
    std::vector<std::thread> workers;
    for (int r = 1; r < R; ++r)
        workers.emplace_back([&, r] { rcs[(size_t)r] = issue(r); if (rcs[(size_t)r]) errs[(size_t)r] = srti::g_err; });
    rcs[0] = issue(0);
    if (rcs[0]) errs[0] = srti::g_err;
    for (auto& w : workers) w.join();
    for (int r = 0; r < R; ++r)
        if (rcs[(size_t)r]) { srti::g_err = errs[(size_t)r]; return rcs[(size_t)r]; }
    CU(cudaSetDevice(g0.device));
    for (int r = 1; r < R; ++r) CU(cudaStreamWaitEvent(g0.ctx->stream, m->gpus[(size_t)r].done, 0));   // the join
    CU(cudaEventRecord(m->t1, g0.ctx->stream));
    for (auto& pl : planes)
        if (pl.host) {
            const size_t at = (size_t)rb * params->width, n = (size_t)(re - rb) * params->width;
            if (n) CU(cudaMemcpyAsync((char*)pl.user + at * pl.elem, (char*)pl.dev + at * pl.elem, n * pl.elem, cudaMemcpyDeviceToHost, g0.ctx->stream));
        }
    m->have_frame = true;
    return SRT_OK;
}

int srt_mgpu_last_stats(srt_mgpu* m, srt_stats* out) {
    if (!m || !out) return fail(SRT_ERR_INVALID_ARG, "NULL argument");
    if (!m->have_frame) return fail(SRT_ERR_INVALID_ARG, "no srt_mgpu_render_frame has been issued");
    std::memset(out, 0, sizeof *out);
    for (auto& g : m->gpus) {
        srt_stats s;
        const int rc = srt_last_stats(g.ctx, &s);
        if (rc) return rc;
        out->primary_rays += s.primary_rays; out->hit_pixels += s.hit_pixels; out->shadow_rays += s.shadow_rays;
        out->node_visits += s.node_visits; out->tri_tests += s.tri_tests; out->exact_slab_tests += s.exact_slab_tests;
        out->kernel_launches += s.kernel_launches;
    }
    CU(cudaSetDevice(m->gpus[0].device));
    CU(cudaEventSynchronize(m->t1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, m->t0, m->t1));
    out->kernel_ms = ms;
    return SRT_OK;
}

int srt_mgpu_render_series(srt_mgpu* m, const srt_scene_desc* const* scenes, const srt_render_params* params, uint32_t n_frames,
                           uint8_t* const* rgb_out) {
    if (!m || (n_frames && (!scenes || !params || !rgb_out))) return fail(SRT_ERR_INVALID_ARG, "NULL argument");
    const uint32_t R = (uint32_t)m->gpus.size();
    for (uint32_t f = 0; f < n_frames; ++f) {
        if (!scenes[f] || !rgb_out[f]) return fail(SRT_ERR_INVALID_ARG, "frame %u: NULL scene or output", f);
        srt_mgpu::Gpu& g = m->gpus[f % R];                  // frame f -> GPU f mod R (:534: frames are independent)
        const int l = (int)((f / R) % kLanesPerDevice);
        if (!g.lane_ctx[l]) {
            const int rc = srt_context_create(g.device, &g.lane_ctx[l]);
            if (rc) return rc;
            g.lane_ctx[l]->bands = 1;                       // several frames are in flight already
        }
        int rc = g.lane_scene[l] ? srt_scene_update(g.lane_ctx[l], g.lane_scene[l], scenes[f])
                                 : srt_scene_upload(g.lane_ctx[l], scenes[f], &g.lane_scene[l]);
        if (rc) return rc;
        srt_frame fr = {rgb_out[f], nullptr, nullptr, nullptr};
        if ((rc = srt_render(g.lane_ctx[l], g.lane_scene[l], &params[f], &fr))) return rc;
    }
    return SRT_OK;
}

}  // extern "C"

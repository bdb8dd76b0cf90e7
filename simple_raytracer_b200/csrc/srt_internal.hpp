// srt_internal.hpp — definitions shared by the translation units that implement the C ABI (srt_capi.cu: contexts,
// scene upload, render; srt_prep_capi.cu: device scene prep).  Not part of the public interface.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "srt.h"
#include "srt_kernels.cuh"

namespace srti {

extern thread_local std::string g_err;
int fail(int code, const char* fmt, ...);          // records the message for srt_last_error(), returns `code`
bool is_device_pointer(const void* p);

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return srti::fail(e__ == cudaErrorMemoryAllocation ? SRT_ERR_OOM : SRT_ERR_CUDA, "%s: %s",   \
                        #call, cudaGetErrorString(e__));                                           \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes) {
        if (bytes <= cap) return SRT_OK;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) return srti::fail(SRT_ERR_OOM, "cudaMalloc(%zu): %s", want, cudaGetErrorString(e));
        cap = want;
        return SRT_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

}  // namespace srti

struct srt_prep_state;                             // srt_prep_capi.cu
void srti_prep_state_destroy(srt_prep_state* p);

struct srt_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    int sm_count = 0, cc_major = 0, cc_minor = 0;
    size_t l2_bytes = 0;
    unsigned long long* counters = nullptr;      // device, srtk::CNT_WORDS
    unsigned long long* counters_host = nullptr; // pinned
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_copy = nullptr;
    srti::DevBuf out_rgb, out_obj, out_tri, out_t;     // staging for HOST output pointers
    srti::DevBuf hit_t, hit_tri, tile_interest, tile_list, tile_order, tile_done, shadow_mask;
    uint32_t frame_seq = 0;   // scratch of the phased kernel
    srti::DevBuf sparse_rgb, sparse_seg, sparse_points, sparse_colors, sparse_total;   // srt_render_sparse
    unsigned long long* sparse_total_host = nullptr;   // pinned
    uint64_t sparse_count = 0;
    bool have_sparse = false, sparse_pending = false;
    int last_launches = 0;
    // last render
    bool have_last = false;
    // Two-band rendering (srt_render): the lower half of the rows is rendered by a child context on its own stream, so
    // that the stages of the two halves overlap (one half's shadow pass runs in the other half's primary-pass tail).
    static constexpr int kMaxBands = 4;
    srt_context* band[kMaxBands - 1] = {};   // children (owned) for bands 1..; their `parent` points back
    srt_context* parent = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join[kMaxBands - 1] = {};
    int last_bands = 1;               // bands of the last render (1 = one piece)
    bool last_banded = false;
    int bands = 2;                    // srt_context_set_bands (SRT_DEFAULT_BANDS overrides the default)
    cudaEvent_t stage_ev[5] = {};     // srt_context_set_stage_timing: around classify / primary / shadow / shade
    bool stage_timing = false, have_stage = false;
    uint64_t last_primary = 0;
    int last_lights = 0;
    uint32_t last_objects = 0;
};

struct srt_scene {
    srt_context* ctx = nullptr;
    srti::DevBuf nodes, geom, attr, flat_geom, flat_attr, objects, texdesc;
    srti::DevBuf vnormal, flat_vnormal;           // optional vertex normals (smooth_normals)
    bool has_vnormal = false, has_flat_vnormal = false;
    srti::DevBuf wide, root_ref;                        // device-internal traversal layout (srt_phased.cu)
    // `wide` with plane pairs sorted near-first for one direction octant (the shadow pass's fast path).  Derived data: built on
    // the owning context's stream by srt_render when the scene's contents or the light's octant changed (-1: stale).
    srti::DevBuf pwide;                                 // four quadrant-sorted copies of `wide` for the primary pass (built with it)
    mutable srti::DevBuf swide;
    mutable int swide_octant = -1;
    bool phased_ok = false;                       // leaf sizes / index ranges fit the packed traversal refs
    float coord_bound = 0.f;                      // max |coordinate| over the root boxes
    std::vector<srti::DevBuf> tex;
    std::vector<srtk::TexDev> tex_host;
    struct TexKey { const void* host; int32_t width, height; uint64_t fingerprint; };
    std::vector<TexKey> tex_key;                  // what each resident slot was uploaded from (srt_scene_update compares)
    uint64_t n_nodes = 0, n_tris = 0, n_flat = 0;
    uint32_t n_objects = 0, n_textures = 0;
    uint32_t max_depth = 0;
    srt_prep_state* prep = nullptr;              // device scene prep (srt_scene_prep): scratch, plan, status
    const float* coord_bound_dev = nullptr;       // set by srt_scene_prep: the bound lives on the device
};

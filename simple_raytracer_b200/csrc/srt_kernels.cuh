// srt_kernels.cuh — launch-side view of the render kernels (shared by srt_kernels.cu / srt_capi.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "srt.h"

namespace srtk {

struct TexDev {
    const uint8_t* rgb;
    int32_t width, height;
};

// Slots of the per-context device counter block.
enum { CNT_TILE = 0, CNT_HIT = 1, CNT_NODE = 2, CNT_TRI = 3, CNT_LIST = 4, CNT_SHADOW = 5, CNT_SHADE = 6,
       CNT_EXACT = 7, CNT_HEAVY = 8, CNT_LIGHT = 9, CNT_PDONE = 10, CNT_TIMEOUT = 11, CNT_WORDS = 16 };

// Everything one frame needs, passed by value as a __grid_constant__ kernel parameter.
struct KArgs {
    const float4* nodes;        // srt_node as 2 x float4
    const float4* geom;         // srt_tri_geom as 3 x float4 (leaf order)
    const float4* attr;         // srt_tri_attr as 3 x float4
    const float4* flat_geom;    // load order (legacy one-box path) or null
    const float4* flat_attr;
    const float4* vnormal;      // srt_tri_vnormal as 3 x float4 (leaf order) or null
    const float4* flat_vnormal;
    int32_t smooth_normals;     // 1: interpolateNormal :132-140 instead of the flat normal
    int32_t normalise_lights;   // 1: light sum / light_count before tone mapping
    int32_t texture_mode;       // SRT_TEX_*
    float shadow_t_lo, shadow_t_hi;   // an occluder counts for t in [lo, hi]: (-inf, +inf) as shipped, (eps, 1) with shadow_segment
    const srt_object* objects;
    const TexDev* textures;
    uint32_t n_objects;
    int32_t width, height, row_begin, row_end;
    int32_t i0, j0;             // first values of the loops :511 / :513 (float -> int truncation of -W/2, -H/2)
    float focal;
    int32_t light_count;
    float light_color[3];
    float shadow_div, tone_k, gamma;
    uint8_t sky[3];
    uint8_t no_sky_fill;
    int32_t use_bvh;
    int32_t tiles_x, tiles_y;   // 8x4-pixel warp tiles this launch renders (its shard of [row_begin,row_end))
    int32_t block_tile_rows, shard_index, shard_count;   // row-block sharding: tile rows per block; 0 = off
    uint8_t* rgb;
    int32_t* obj_id;
    int32_t* tri_id;
    float* t;
    unsigned long long* counters;   // CNT_* slots, zeroed before launch
    float lights[3 * SRT_MAX_LIGHTS];   // staircase positions, softShadow :363-383
    float light_lo[3], light_hi[3];     // bounding box of the light positions (bundle test of the primary pass)
    // ---- phased kernel only (device-internal data built at upload, scratch owned by the context) ----
    const float4* wide;         // per interior node i: the two child boxes + child refs, 4 x float4 at wide[4*i]
    const uint32_t* root_ref;   // per object: traversal ref of its root
    const float4* pwide;        // four copies of `wide`, planes sorted near / far for the primary rays' quadrants 0..3 (or null)
    size_t pwide_stride;        // float4s per copy
    const float4* swide;        // `wide` with every box's planes sorted near / far for rays of octant `shadow_octant` (or null)
    uint32_t shadow_octant;     // bit k: the shadow rays' direction component k is negative (all lights lie far outside the scene)
    float* hit_t;               // [height][width] scratch: closest-hit parameter
    int32_t* hit_tri;           // [height][width] scratch: closest-hit triangle (leaf order) or -1
    uint32_t* tile_interest;    // [tiles] lanes whose shadow rays can reach another object's root box
    unsigned long long* tile_list;   // published tiles with tile_interest != 0: (frame_seq << 32) | tile
    uint32_t* tile_done;        // [tiles] shadow items completed
    uint32_t frame_seq;         // tag of this launch's tile_list entries (no clearing between frames)
    uint32_t* tile_order;       // [tiles] phase-P queue order: expected-heavy tiles first
    uint32_t* shadow_mask;      // [tiles][light_count] ballot of occluded lanes
    float coord_bound;          // max |coordinate| over all root boxes (margin of the filtered slab test)
    const float* coord_bound_dev;   // non-null: read the bound from device memory instead (device-prepared scenes)
    int32_t stack_cap;          // per-thread traversal stack entries (shared memory)
    int32_t debug_timeline;     // 1: phase P overwrites the t / obj_id / tri_id planes with per-tile cycles / SM / start ns
};

constexpr int kTileW = 8, kTileH = 4;          // one warp = 8x4 pixels

// First pixel row of local tile row `ty` (interleaved row-block sharding, see srt_render_params).
__host__ __device__ inline int tile_row_to_py(int row_begin, int block_tile_rows, int shard_index, int shard_count, int ty) {
    if (block_tile_rows > 0) {
        const int blk = ty / block_tile_rows, off = ty - blk * block_tile_rows;
        ty = (blk * shard_count + shard_index) * block_tile_rows + off;
    }
    return row_begin + ty * kTileH;
}
constexpr int kMaxStack = 64;                  // >= deepest BVH the upload accepts

// variant: SRT_VARIANT_*; returns cudaError_t from the launch.
cudaError_t launch_render(const KArgs& a, int variant, int sm_count, cudaStream_t stream);

// ---- phased kernel (srt_phased.cu) ----
constexpr uint32_t kRefNone = 0x7FFFFFFFu;      // traversal refs: < kRefNone interior node index; bit 31 = leaf
constexpr uint32_t kLeafMaxCount = 15, kLeafFirstBits = 27;
// Builds wide[] / root_ref[] from nodes[] / objects[] on the device (after every scene upload).
cudaError_t launch_build_wide(const float4* nodes, const srt_object* objects, uint32_t n_nodes, uint32_t n_objects,
                              float4* wide, uint32_t* root_ref, cudaStream_t stream);
// swide[] from wide[] for one ray-direction octant (the shadow pass's fast path; rebuilt when the scene or the octant changes).
cudaError_t launch_build_shadow_wide(const float4* wide, uint32_t n_nodes, uint32_t octant, float4* swide, cudaStream_t stream);
// Shared memory the phased kernel needs per CTA for `stack_cap`, and whether it fits.
int phased_stack_cap(int max_depth, int n_objects);   // stack of interior nodes + room for the leaf list
size_t phased_smem_bytes(int stack_cap);
// (launch_render_phased: the staged pipeline, its overlap variant, or the fused dataflow kernel.)
// Sparse (ImageData-style) output of a rendered RGB8 frame: srt_render_sparse.  seg_counts: sparse_segments() words.
size_t sparse_segments(int width, int rows);
cudaError_t launch_sparse_compact(const uint8_t* rgb, int width, int row_begin, int row_end, uint32_t* seg_counts, float* points,
                                  float* colors, unsigned long long* total, cudaStream_t stream);

// srt_sparse_from_frame_static: the same compaction with its scratch in a file-scope __device__ array (srt_kernels.cu);
// `device` must be current.  Allocates nothing.
cudaError_t launch_sparse_compact_static(int device, const uint8_t* rgb, int width, int row_begin, int row_end, float* points,
                                         float* colors, unsigned long long* total, cudaStream_t stream);

// Leaf evaluation hook (srt_eval_leaf): `in` / `out` are device arrays.
int leaf_in_width(int op);
int leaf_out_width(int op);
cudaError_t launch_eval_leaf(int op, const float* in, float* out, size_t n, cudaStream_t stream);

cudaError_t launch_render_phased(const KArgs& a, int variant, int sm_count, cudaStream_t stream, int* launches, cudaEvent_t* stage_ev = nullptr);

}  // namespace srtk

// srt_prep.cuh — launch-side view of the device scene prep (srt_prep.cu): the per-frame work the reference does
// on the host before it renders — transformTriangles (Object.cpp:183-190), createBoundingHierarchy
// (Object.cpp:205-284) and the flattening of srt_flatten.hpp — as sm_100a kernels that produce the SAME arrays.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "srt.h"

namespace srtp {

// The tree topology depends on the triangle count alone (halves n/2 | n - n/2, recursion iff > 8,
// Object.cpp:254-267; the root is always split, :282): one table per mesh, depth-first pre-order.
struct TopoNode {
    uint32_t start, count;     // range of the node in the object's (sorted) triangle order
    uint32_t right;            // index of the right child relative to the object's root; 0 = leaf (left child = self + 1)
    uint32_t depth;
};

struct PrepObject {            // one object of the frame (device copy of srt_prep_object + placement in the scene arrays)
    const float4*   pts;       // mesh: [n_tris][3] points as loaded
    const float2*   uv;        // mesh: [n_tris][3] texel coordinates or null
    const int32_t*  tex;       // mesh: [n_tris] texture index or null
    const float*    vn;        // mesh: [n_tris][9] vertex normals as loaded, or null (zeros, like a mesh without vn lines)
    const TopoNode* topo;      // mesh: [n_nodes]
    uint32_t n_tris, n_nodes;
    uint32_t tri_base, node_base, flat_first;
    uint32_t n_matrices, mat_first;
    float color[3], ambient, specular, shininess;
};

struct WorkItem { uint32_t obj, node, start, count; };   // one node to split (node / start relative to its object)

struct PrepArgs {
    const PrepObject* objects;
    uint32_t n_objects;
    const float* matrices;     // 16 floats each, column-major
    // scene arrays (outputs)
    float4* nodes; float4* geom; float4* attr; float4* flat_geom; float4* flat_attr;
    float4* vnormal; float4* flat_vnormal;     // vertex normals in leaf / load order (copied, never transformed)
    srt_object* scene_objects;
    float* coord_bound;        // max |coordinate| over the root boxes
    uint32_t* status;          // bit 0: a transformed coordinate is not finite
    // scratch, indexed by scene triangle position (tri_base + i)
    float4* xpts;              // [N][3] transformed points
    float* lite;               // [N][6] per-triangle box: lo xyz, hi xyz
    uint32_t* order;           // [N] object-relative triangle ids in the current BVH order
    float* key; uint32_t* idx; uint32_t* posl; uint32_t* posr;   // [N] each: sort workspace of nodes too large for shared memory
    uint32_t* segq;            // [2][N/8 + n][3] segment queues of those nodes
    uint32_t segq_stride;      // entries per queue
    unsigned long long* rootacc;   // [n_objects][6] packed (value code, position) minima of the object boxes
    unsigned long long* dbg;       // optional timing slots of the finishing CTAs (SRT_PREP_DEBUG), else null
};

// ---- large nodes: the introsort loop of a node spread over many CTAs (one launch per step of a round) ----
struct WideSeg {               // a segment being partitioned: [lo, hi) are scene positions, key[lo] holds the pivot
    uint32_t lo, hi, d;
    float pivot;
    uint32_t tile_first, ntiles;
    uint32_t K;                // pairs swapped (written by the CTA that holds the last swapping pair)
    uint32_t done;             // tiles that finished the swap step
};
struct WideRound { uint32_t nsegs, ntiles; };
struct WideSmall { uint32_t lo, hi, d; };    // a segment short enough for one CTA's shared memory (or out of depth / rounds)

struct WideArgs {
    WideSeg* segs;   uint32_t max_segs;      // [2][max_segs], by round parity
    uint2* tile_seg; uint32_t max_tiles;     // [2][max_tiles]: (segment, tile index within the segment)
    uint2* tile_cnt;                         // [max_tiles]: elements not less / not greater than the pivot
    WideRound* rounds;                       // [kWideMaxRounds + 1]
    WideSmall* fin;  uint32_t max_fin;       // segments left to one CTA each
    uint32_t* fin_count;
    unsigned long long* childacc;            // [items][2][6] packed minima of the child boxes
};
constexpr int kWideMaxRounds = 40;
constexpr uint32_t kWideTile = 1024;         // elements per tile CTA
constexpr uint32_t kWideSmall = 12288;       // a segment at most this long is finished by one CTA in shared memory

cudaError_t launch_wide_level(const PrepArgs& a, const WideArgs& w, const WorkItem* items, uint32_t n_items, uint32_t max_count,
                              uint32_t rounds, uint32_t tile_grid, int sm_count, cudaStream_t stream, int* launches);

constexpr int kSubtreeMax = 4096;     // a node with at most this many triangles is built, with its whole subtree, by one CTA in shared memory
constexpr int kBuildThreads = 1024;

size_t build_smem_bytes(bool subtree);
cudaError_t prepare_kernels();        // one-time function attributes (shared-memory opt-in)

cudaError_t launch_transform(const PrepArgs& a, uint32_t max_tris, cudaStream_t stream);
cudaError_t launch_roots(const PrepArgs& a, cudaStream_t stream);
// items: device array.  subtree = false: every CTA splits ONE node (sort + child boxes); true: the node and its whole subtree.
cudaError_t launch_build(const PrepArgs& a, const WorkItem* items, uint32_t n_items, bool subtree, cudaStream_t stream);
cudaError_t launch_flatten(const PrepArgs& a, uint32_t max_tris, bool flat_lists, cudaStream_t stream);

}  // namespace srtp

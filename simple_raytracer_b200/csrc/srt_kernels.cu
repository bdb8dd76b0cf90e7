// srt_kernels.cu — the render kernels (sm_100a).  See DESIGN.md §Kernels.
//
// K1 "simple": persistent grid, one warp = one 8x4 pixel tile pulled from an atomic tile counter,
// one thread = one pixel: primary ray -> stack traversal of every object's flattened BVH with the
// reference's slab test -> Moeller-Trumbore on every leaf the line pierces -> closest hit with
// the reference's first-visited-wins rule -> N staircase lights, each with an any-hit shadow ray
// through the OTHER objects -> Phong -> tone map / gamma -> 8-bit quantise -> sky fill.
// Replaces simple_raytracer.cpp:505-525 (pixel loop), :405-457, :296-317, :321-401, and the
// 8-bit / sky rules of drawImage :461-498.
#include <mutex>
#include "srt_kernels.cuh"
#include "srt_device.cuh"
#include "srt_shade.cuh"

namespace srtk {
using namespace srtd;

namespace {


// boundingBoxIntersection :296-317 merged with the candidate loop of rayIntersection :424-431:
// depth-first, left child first, every leaf whose box passes the slab test; strict '<' keeps the
// first-visited triangle on equal t (S5, S6).
template <bool COUNT>
__device__ __forceinline__ void closest_in_tree(const KArgs& a, uint32_t root, int32_t obj, V3 o, V3 d, Hit& h,
                                                unsigned& nodes_seen, unsigned& tris_seen) {
    uint32_t stack[kMaxStack];
    int sp = 0;
    uint32_t ni = root;
    for (;;) {
        const float4 n0 = ldg4(a.nodes + 2 * (size_t)ni);
        const float4 n1 = ldg4(a.nodes + 2 * (size_t)ni + 1);
        if (COUNT) ++nodes_seen;
        bool descend = false;
        if (slab_exact(o, d, n0, n1)) {
            const uint32_t na = __float_as_uint(n0.w), nb = __float_as_uint(n1.w);
            if (nb & SRT_LEAF_BIT) {
                const uint32_t cnt = nb & ~SRT_LEAF_BIT;
                for (uint32_t k = 0; k < cnt; ++k) {
                    const float4* g = a.geom + 3 * (size_t)(na + k);
                    const float t = moller_trumbore(o, d, ldg4(g), ldg4(g + 1), ldg4(g + 2));
                    if (COUNT) ++tris_seen;
                    if (t != -INFINITY && t < h.t) { h.t = t; h.tri = (int32_t)(na + k); h.obj = obj; }
                }
            } else {
                stack[sp++] = na;          // right child later
                ni = ni + 1;               // left child first
                descend = true;
            }
        }
        if (!descend) {
            if (sp == 0) break;
            ni = stack[--sp];
        }
    }
}

// Legacy one-box path :415-420.
template <bool COUNT>
__device__ __forceinline__ void closest_in_flat(const KArgs& a, const srt_object& ob, int32_t obj, V3 d, Hit& h,
                                                unsigned& nodes_seen, unsigned& tris_seen) {
    const V3 o = v3(0.f, 0.f, 0.f);
    if (COUNT) ++nodes_seen;
    const float4 lo = make_float4(ob.box_min[0], ob.box_min[1], ob.box_min[2], 0.f);
    const float4 hi = make_float4(ob.box_max[0], ob.box_max[1], ob.box_max[2], 0.f);
    if (!slab_exact(o, d, lo, hi)) return;                                   // intersectRayAabb :417
    for (uint32_t k = 0; k < ob.tri_count; ++k) {
        const float4* g = a.flat_geom + 3 * (size_t)(ob.flat_first + k);
        const float t = moller_trumbore(o, d, ldg4(g), ldg4(g + 1), ldg4(g + 2));
        if (COUNT) ++tris_seen;
        if (t != -INFINITY && t < h.t) { h.t = t; h.tri = (int32_t)(ob.flat_first + k); h.obj = obj; }
    }
}

// shadowIntersection :321-342 restricted to one object's tree: true as soon as ANY candidate
// triangle returns a hit (any t >= 0, no upper bound at the light, no epsilon — S11).
template <bool COUNT>
__device__ __forceinline__ bool any_hit_in_tree(const KArgs& a, uint32_t root, V3 o, V3 d, unsigned& nodes_seen,
                                                unsigned& tris_seen) {
    uint32_t stack[kMaxStack];
    int sp = 0;
    uint32_t ni = root;
    for (;;) {
        const float4 n0 = ldg4(a.nodes + 2 * (size_t)ni);
        const float4 n1 = ldg4(a.nodes + 2 * (size_t)ni + 1);
        if (COUNT) ++nodes_seen;
        bool descend = false;
        if (slab_exact(o, d, n0, n1)) {
            const uint32_t na = __float_as_uint(n0.w), nb = __float_as_uint(n1.w);
            if (nb & SRT_LEAF_BIT) {
                const uint32_t cnt = nb & ~SRT_LEAF_BIT;
                for (uint32_t k = 0; k < cnt; ++k) {
                    const float4* g = a.geom + 3 * (size_t)(na + k);
                    if (COUNT) ++tris_seen;
                    const float ts = moller_trumbore(o, d, ldg4(g), ldg4(g + 1), ldg4(g + 2));
                    if (ts != -INFINITY && !(ts < a.shadow_t_lo) && !(ts > a.shadow_t_hi)) return true;
                }
            } else {
                stack[sp++] = na;
                ni = ni + 1;
                descend = true;
            }
        }
        if (!descend) {
            if (sp == 0) break;
            ni = stack[--sp];
        }
    }
    return false;
}

template <bool COUNT, bool TIMELINE = false>
__global__ void __launch_bounds__(128) render_simple(const __grid_constant__ KArgs a) {
    const unsigned lane = threadIdx.x & 31;
    const int n_tiles = a.tiles_x * a.tiles_y;
    unsigned hits = 0, nodes_seen = 0, tris_seen = 0;
    const bool use_flat = !a.use_bvh;
    for (;;) {
        unsigned tile = 0;
        if (lane == 0) tile = (unsigned)atomicAdd(&a.counters[CNT_TILE], 1ull);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if ((int)tile >= n_tiles) break;
        const int tx = (int)(tile % (unsigned)a.tiles_x), ty = (int)(tile / (unsigned)a.tiles_x);
        const int px = tx * kTileW + (int)(lane & 7);
        const int py = tile_row_to_py(a.row_begin, a.block_tile_rows, a.shard_index, a.shard_count, ty) + (int)(lane >> 3);
        long long tl_c0 = 0; unsigned long long tl_ns = 0;
        if (TIMELINE) { tl_c0 = clock64(); asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tl_ns)); }
        if (px >= a.width || py >= a.row_end) continue;
        // ray.direction = (i, j, focal), i = px + i0, j = py + j0            :511-515 (S1)
        const V3 o = v3(0.f, 0.f, 0.f);
        const V3 d = v3((float)(px + a.i0), (float)(py + a.j0), a.focal);
        Hit h; h.t = INFINITY; h.tri = -1; h.obj = -1;
        for (uint32_t ob = 0; ob < a.n_objects; ++ob) {                       // :409 (S2 order)
            if (use_flat) closest_in_flat<COUNT>(a, a.objects[ob], (int32_t)ob, d, h, nodes_seen, tris_seen);
            else closest_in_tree<COUNT>(a, a.objects[ob].root, (int32_t)ob, o, d, h, nodes_seen, tris_seen);
        }
        V3 color = v3(0.f, 0.f, 0.f);
        const bool hit = h.tri >= 0;
        if (hit) {
            ++hits;
            const PhongPixel pp = prepare_pixel(a, h, d, use_flat);
            const V3 so = v3(mul(d.x, h.t), mul(d.y, h.t), mul(d.z, h.t));    // shadowRay.origin :326
            const V3 lc = v3(a.light_color[0], a.light_color[1], a.light_color[2]);
            for (int i = 0; i < a.light_count; ++i) {                         // :366
                const V3 L = v3(a.lights[3 * i], a.lights[3 * i + 1], a.lights[3 * i + 2]);
                bool shadowed = false;
                const V3 sd = v3(sub(L.x, so.x), sub(L.y, so.y), sub(L.z, so.z));   // :325
                for (uint32_t ob = 0; ob < a.n_objects && !shadowed; ++ob) {  // :322
                    if ((int32_t)ob == h.obj) continue;                       // :331
                    shadowed = any_hit_in_tree<COUNT>(a, a.objects[ob].root, so, sd, nodes_seen, tris_seen);
                }
                V3 c = phong(pp, L, lc);                                      // :368
                if (shadowed) { c.x = dvd(c.x, a.shadow_div); c.y = dvd(c.y, a.shadow_div); c.z = dvd(c.z, a.shadow_div); }  // :369
                color.x = add(color.x, c.x); color.y = add(color.y, c.y); color.z = add(color.z, c.z);   // :370
            }
            if (a.normalise_lights) color = average_lights(color, a.light_count);
        }
        store_pixel(a, px, py, hit, color, h);
        if (TIMELINE) {   // debugging aid: per-tile cost map instead of the id / t planes
            __syncwarp();
            const long long dc = clock64() - tl_c0;
            unsigned smid; asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
            const size_t at = (size_t)py * a.width + px;
            if (a.t) a.t[at] = (float)dc;
            if (a.obj_id) a.obj_id[at] = (int32_t)smid;
            if (a.tri_id) a.tri_id[at] = (int32_t)(tl_ns & 0x7fffffffull);
        }
    }
    // one atomic per warp for the statistics
    for (int s = 16; s > 0; s >>= 1) {
        hits += __shfl_xor_sync(0xffffffffu, hits, s);
        if (COUNT) { nodes_seen += __shfl_xor_sync(0xffffffffu, nodes_seen, s); tris_seen += __shfl_xor_sync(0xffffffffu, tris_seen, s); }
    }
    if (lane == 0) {
        atomicAdd(&a.counters[CNT_HIT], (unsigned long long)hits);
        if (COUNT) { atomicAdd(&a.counters[CNT_NODE], (unsigned long long)nodes_seen); atomicAdd(&a.counters[CNT_TRI], (unsigned long long)tris_seen); }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Sparse output: the reference's ImageData (simple_raytracer.cpp:28, :505-525) — only the pixels whose quantised
// colour is not (0,0,0) (:518), in its column-major scan order (x outer :511, y inner :513), positions and colours
// as floats.  Three small kernels over the rendered RGB8 frame: count per (column, 32-row segment), exclusive scan
// in (column-major) segment order, scatter.
namespace {
constexpr int kSegRows = 32;

__device__ __forceinline__ bool drawn(const uint8_t* rgb, size_t at) { return (rgb[3 * at] | rgb[3 * at + 1] | rgb[3 * at + 2]) != 0; }

__global__ void sparse_count_kernel(const uint8_t* __restrict__ rgb, int width, int row_begin, int row_end, int n_seg,
                                    uint32_t* __restrict__ counts) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, seg = blockIdx.y;
    if (x >= width) return;
    const int y0 = row_begin + seg * kSegRows, y1 = min(y0 + kSegRows, row_end);
    uint32_t n = 0;
    for (int y = y0; y < y1; ++y) n += drawn(rgb, (size_t)y * width + x) ? 1u : 0u;
    counts[(size_t)x * n_seg + seg] = n;
}

// One block: exclusive scan of counts[0..n) in place, total to *total.
__global__ void sparse_scan_kernel(uint32_t* __restrict__ counts, uint32_t n, unsigned long long* __restrict__ total) {
    __shared__ uint32_t part[1024];
    const uint32_t per = (n + blockDim.x - 1) / blockDim.x;
    const uint32_t b = min(threadIdx.x * per, n), e = min(b + per, n);
    uint32_t sum = 0;
    for (uint32_t i = b; i < e; ++i) sum += counts[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (uint32_t off = 1; off < blockDim.x; off <<= 1) {               // Hillis-Steele over the per-thread sums
        const uint32_t v = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - sum;                             // exclusive prefix of this thread's chunk
    for (uint32_t i = b; i < e; ++i) { const uint32_t c = counts[i]; counts[i] = run; run += c; }
    if (threadIdx.x == blockDim.x - 1) *total = part[threadIdx.x];
}

__global__ void sparse_write_kernel(const uint8_t* __restrict__ rgb, int width, int row_begin, int row_end, int n_seg,
                                    const uint32_t* __restrict__ offsets, float* __restrict__ points, float* __restrict__ colors) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, seg = blockIdx.y;
    if (x >= width) return;
    const int y0 = row_begin + seg * kSegRows, y1 = min(y0 + kSegRows, row_end);
    size_t o = offsets[(size_t)x * n_seg + seg];
    for (int y = y0; y < y1; ++y) {
        const size_t at = (size_t)y * width + x;
        if (!drawn(rgb, at)) continue;
        points[2 * o] = (float)x; points[2 * o + 1] = (float)y;                       // glm::vec2(x, y)          :516
        colors[3 * o] = (float)rgb[3 * at]; colors[3 * o + 1] = (float)rgb[3 * at + 1]; colors[3 * o + 2] = (float)rgb[3 * at + 2];
        ++o;
    }
}
}  // namespace

size_t sparse_segments(int width, int rows) { return (size_t)width * (size_t)((rows + kSegRows - 1) / kSegRows); }

cudaError_t launch_sparse_compact(const uint8_t* rgb, int width, int row_begin, int row_end, uint32_t* seg_counts, float* points,
                                  float* colors, unsigned long long* total, cudaStream_t stream) {
    const int rows = row_end - row_begin;
    if (width <= 0 || rows <= 0) return cudaMemsetAsync(total, 0, sizeof(unsigned long long), stream);
    const int n_seg = (rows + kSegRows - 1) / kSegRows;
    const dim3 grid((unsigned)((width + 127) / 128), (unsigned)n_seg);
    sparse_count_kernel<<<grid, 128, 0, stream>>>(rgb, width, row_begin, row_end, n_seg, seg_counts);
    sparse_scan_kernel<<<1, 1024, 0, stream>>>(seg_counts, (uint32_t)((size_t)width * n_seg), total);
    sparse_write_kernel<<<grid, 128, 0, stream>>>(rgb, width, row_begin, row_end, n_seg, seg_counts, points, colors);
    return cudaGetLastError();
}

// ---- srt_sparse_from_frame_static: the ImageData compaction with its scratch in a file-scope device array --------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops).  The three compaction launches above, with the per-(column,
// 32-row segment) count / offset words in g_sparse_static_scratch instead of a context buffer.  Every GPU that loads
// the library holds the array (1 MiB) until exit; every call on a device shares it, so calls that overlap on the GPU race.
// Nothing has to start at zero: the count pass writes every word [0, width * segments) the scan and write passes read,
// and the scan pass writes the total — whatever an earlier call left behind is never read.
constexpr size_t kSparseStaticWords = SRT_SPARSE_STATIC_MAX_SEGMENTS;
static __device__ __align__(256) uint32_t g_sparse_static_scratch[kSparseStaticWords];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_sparse_static_mutex;
uint32_t* g_sparse_static_ptr[kMaxDevices] = {};      // per device: the array's DEVICE address (the symbol itself is a host stand-in)

__global__ void store_u64_kernel(unsigned long long* p, unsigned long long v) { *p = v; }
}  // namespace

// Device address of the scratch array on `device`, which must be current.  Looked up once per device.
cudaError_t sparse_static_scratch(int device, uint32_t** out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_sparse_static_mutex);
    if (!g_sparse_static_ptr[device]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_sparse_static_scratch);
        if (e != cudaSuccess) return e;
        if (!ptr) return cudaErrorInvalidSymbol;
        g_sparse_static_ptr[device] = (uint32_t*)ptr;
    }
    *out = g_sparse_static_ptr[device];
    return cudaSuccess;
}

cudaError_t launch_sparse_compact_static(int device, const uint8_t* rgb, int width, int row_begin, int row_end, float* points,
                                         float* colors, unsigned long long* total, cudaStream_t stream) {
    const int rows = row_end - row_begin;
    if (width <= 0 || rows <= 0) {                    // (no cudaMemset: `total` may be pinned host memory)
        store_u64_kernel<<<1, 1, 0, stream>>>(total, 0ull);
        return cudaGetLastError();
    }
    if (sparse_segments(width, rows) > kSparseStaticWords) return cudaErrorInvalidValue;   // the entry point reports this first
    uint32_t* scratch = nullptr;
    const cudaError_t e = sparse_static_scratch(device, &scratch);
    if (e != cudaSuccess) return e;
    const int n_seg = (rows + kSegRows - 1) / kSegRows;
    const dim3 grid((unsigned)((width + 127) / 128), (unsigned)n_seg);
    sparse_count_kernel<<<grid, 128, 0, stream>>>(rgb, width, row_begin, row_end, n_seg, scratch);
    sparse_scan_kernel<<<1, 1024, 0, stream>>>(scratch, (uint32_t)((size_t)width * n_seg), total);
    sparse_write_kernel<<<grid, 128, 0, stream>>>(rgb, width, row_begin, row_end, n_seg, scratch, points, colors);
    return cudaGetLastError();
}

cudaError_t launch_render(const KArgs& a, int variant, int sm_count, cudaStream_t stream) {
    const int n_tiles = a.tiles_x * a.tiles_y;
    if (n_tiles <= 0) return cudaSuccess;
    const int warps_per_block = 4;
    int blocks = sm_count * 8;                       // persistent: 8 resident CTAs of 4 warps per SM
    const int need = (n_tiles + warps_per_block - 1) / warps_per_block;
    if (blocks > need) blocks = need;
    if (variant == 100) render_simple<true><<<blocks, 128, 0, stream>>>(a);      // instrumented (node / triangle counts)
    else if (variant == 101) render_simple<false, true><<<blocks, 128, 0, stream>>>(a);   // per-tile timeline in the id / t planes
    else render_simple<false><<<blocks, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace srtk

// srt_prep.cu — device scene prep: transformTriangles, createBoundingHierarchy and the flattener as sm_100a kernels
// that reproduce the reference's host results bit for bit.
//
// What has to match (Object.cpp):
//   :183-190  every point goes through one mat4 * vec4 per transformTriangles call, each rounded to fp32
//             (GLM: (m0*x + m1*y) + (m2*z + m3*w)), never a pre-multiplied matrix;
//   :205-221  boxes are sequential folds of glm::min / glm::max over (triangle, point) in order: the VALUE is
//             order-independent, the sign of a zero result is that of the first zero met — carried here as
//             (value, position) minima;
//   :225-272  a node's triangles are sorted with std::sort on pointOne.<axis> ('<'), the axis being the longest
//             box side (strict, else z); halves n/2 | rest; recursion iff a half has more than 8 triangles;
//             children sort what the parent's sort left behind.
// std::sort is not stable, and triangles that share their first vertex have equal keys, so the PERMUTATION
// libstdc++'s introsort produces decides which triangles land in which leaf.  It is reproduced exactly:
//   * __introsort_loop is a sequence of independent Hoare partitions (median of first+1 / mid / last-1 moved to
//     first, __unguarded_partition).  One partition pairs the k-th element from the left that is not less than
//     the pivot with the k-th element from the right that is not greater, swaps the pairs while they have not
//     crossed, and cuts where they cross.  Ranking both kinds of elements (ballots + prefix counts) gives the
//     same swaps without the sequential scan; the two sides of the cut are new segments, to depth 2*lg(n);
//   * segments of at most 16 elements are left to __final_insertion_sort, which is a stable sort in which no
//     element leaves its segment: every element counts the larger keys among the 15 positions before it and the
//     smaller keys among the 15 after it;
//   * a segment that exhausts the depth limit is heap-sorted by one thread with libstdc++'s exact
//     __make_heap / __sort_heap sequence (rare: adversarial key patterns only).
// The comparison outcomes are the host's (same fp32 keys), hence the same permutation, the same tree.
#include <cfloat>
#include <type_traits>

#include "srt_prep.cuh"

namespace srtp {
namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr uint32_t kNoPos = 0xFFFFFFFFu;
constexpr int kNLMax = 512;           // nodes of one level inside a subtree CTA (kSubtreeMax / 9 < 512)
constexpr int kSmemQueue = 256;       // segments longer than 16 inside kSubtreeMax elements: < 4096 / 17
static_assert(2 * kNLMax <= kBuildThreads, "one thread per child in the level compaction");

// ---- monotone code of a float for (value, position) minima; +0 and -0 share one code (they compare equal) ----
__device__ __forceinline__ uint32_t fcode(float v) {
    uint32_t b = __float_as_uint(v);
    if ((b << 1) == 0) b = 0;
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ unsigned long long pack_min(float v, uint32_t pos) { return ((unsigned long long)fcode(v) << 32) | pos; }
__device__ __forceinline__ unsigned long long pack_max(float v, uint32_t pos) { return ((unsigned long long)(~fcode(v)) << 32) | pos; }
__device__ __forceinline__ unsigned long long seed_of(int q) { return q < 3 ? pack_min(FLT_MAX, kNoPos) : pack_max(-FLT_MAX, kNoPos); }
__device__ __forceinline__ unsigned long long min64(unsigned long long a, unsigned long long b) { return a < b ? a : b; }

__device__ __forceinline__ unsigned long long warp_min64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min64(v, __shfl_xor_sync(kFull, v, o));
    return v;
}

// glm::min(x, y) = (y < x) ? y : x ; glm::max(x, y) = (x < y) ? y : x
__device__ __forceinline__ float glm_min(float x, float y) { return (y < x) ? y : x; }
__device__ __forceinline__ float glm_max(float x, float y) { return (x < y) ? y : x; }

// mat4 * vec4, glm/detail/type_mat4x4.inl: (m0*v.x + m1*v.y) + (m2*v.z + m3*v.w), every operation rounded.
__device__ __forceinline__ float4 mat_vec(const float* __restrict__ m, float4 v) {
    float4 r;
    r.x = __fadd_rn(__fadd_rn(__fmul_rn(m[0], v.x), __fmul_rn(m[4], v.y)), __fadd_rn(__fmul_rn(m[8], v.z), __fmul_rn(m[12], v.w)));
    r.y = __fadd_rn(__fadd_rn(__fmul_rn(m[1], v.x), __fmul_rn(m[5], v.y)), __fadd_rn(__fmul_rn(m[9], v.z), __fmul_rn(m[13], v.w)));
    r.z = __fadd_rn(__fadd_rn(__fmul_rn(m[2], v.x), __fmul_rn(m[6], v.y)), __fadd_rn(__fmul_rn(m[10], v.z), __fmul_rn(m[14], v.w)));
    r.w = __fadd_rn(__fadd_rn(__fmul_rn(m[3], v.x), __fmul_rn(m[7], v.y)), __fadd_rn(__fmul_rn(m[11], v.z), __fmul_rn(m[15], v.w)));
    return r;
}

__device__ __forceinline__ bool finite4(float4 v) { return isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w); }

// ------------------------------------------------------------------------------------------------------------
// transformTriangles for every pending matrix of every object + per-triangle boxes + the object boxes
// (createBoundingHierarchy :279 / calculateBoundingBoxes :205-221).
__global__ void __launch_bounds__(256) transform_kernel(const PrepArgs a) {
    const PrepObject o = a.objects[blockIdx.y];
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    const bool valid = i < o.n_tris;
    if (blockIdx.x * 256u >= o.n_tris) return;
    unsigned long long acc[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) acc[q] = seed_of(q);
    bool bad = false;
    if (valid) {
        float4 p[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) p[k] = o.pts[3 * (size_t)i + k];
        for (uint32_t m = 0; m < o.n_matrices; ++m) {
            const float* M = a.matrices + 16 * (size_t)(o.mat_first + m);
#pragma unroll
            for (int k = 0; k < 3; ++k) p[k] = mat_vec(M, p[k]);
        }
        const size_t g = (size_t)o.tri_base + i;
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            a.xpts[3 * g + k] = p[k];
            bad = bad || !finite4(p[k]);
            lo[0] = glm_min(lo[0], p[k].x); lo[1] = glm_min(lo[1], p[k].y); lo[2] = glm_min(lo[2], p[k].z);
            hi[0] = glm_max(hi[0], p[k].x); hi[1] = glm_max(hi[1], p[k].y); hi[2] = glm_max(hi[2], p[k].z);
        }
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            a.lite[6 * g + q] = lo[q];
            a.lite[6 * g + 3 + q] = hi[q];
            if (lo[q] == lo[q] && lo[q] < FLT_MAX) acc[q] = pack_min(lo[q], i);
            if (hi[q] == hi[q] && hi[q] > -FLT_MAX) acc[3 + q] = pack_max(hi[q], i);
        }
        a.order[g] = i;
    }
    if (__any_sync(kFull, bad) && (threadIdx.x & 31) == 0) atomicOr(a.status, 1u);
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        const unsigned long long v = warp_min64(acc[q]);
        if ((threadIdx.x & 31) == 0 && v != seed_of(q)) atomicMin(a.rootacc + 6 * (size_t)blockIdx.y + q, v);
    }
}

// Value behind a packed minimum: the winning triangle's own box component (keeps the sign of a zero).
__device__ __forceinline__ float decode_load_order(const PrepArgs& a, const PrepObject& o, unsigned long long v, int q) {
    const uint32_t pos = (uint32_t)v;
    if (pos == kNoPos) return q < 3 ? FLT_MAX : -FLT_MAX;
    return a.lite[6 * ((size_t)o.tri_base + pos) + q];
}
__device__ __forceinline__ float decode_sorted(const PrepArgs& a, const PrepObject& o, unsigned long long v, int q) {
    const uint32_t pos = (uint32_t)v;
    if (pos == kNoPos) return q < 3 ? FLT_MAX : -FLT_MAX;
    const uint32_t tri = a.order[(size_t)o.tri_base + pos];
    return a.lite[6 * ((size_t)o.tri_base + tri) + q];
}

__device__ __forceinline__ void write_node(const PrepArgs& a, const PrepObject& o, uint32_t node_rel, const float* box) {
    const TopoNode t = o.topo[node_rel];
    uint32_t na, nb;
    if (t.right == 0) { na = o.tri_base + t.start; nb = SRT_LEAF_BIT | t.count; }
    else { na = o.node_base + t.right; nb = 0; }
    float4* n = a.nodes + 2 * ((size_t)o.node_base + node_rel);
    n[0] = make_float4(box[0], box[1], box[2], __uint_as_float(na));
    n[1] = make_float4(box[3], box[4], box[5], __uint_as_float(nb));
}

// One thread per object: the object record, the root node (always an interior node, :282) and the coordinate bound.
__global__ void roots_kernel(const PrepArgs a) {
    const uint32_t obj = blockIdx.x * blockDim.x + threadIdx.x;
    if (obj >= a.n_objects) return;
    const PrepObject o = a.objects[obj];
    float box[6];
    for (int q = 0; q < 6; ++q) box[q] = decode_load_order(a, o, a.rootacc[6 * (size_t)obj + q], q);
    srt_object so;
    so.color[0] = o.color[0]; so.color[1] = o.color[1]; so.color[2] = o.color[2];
    so.ambient = o.ambient; so.specular = o.specular; so.shininess = o.shininess;
    so.root = o.node_base; so.tri_first = o.tri_base; so.tri_count = o.n_tris; so.flat_first = o.flat_first;
    for (int k = 0; k < 3; ++k) { so.box_min[k] = box[k]; so.box_max[k] = box[3 + k]; }
    a.scene_objects[obj] = so;
    write_node(a, o, 0, box);
    float bound = 0.f;
    for (int q = 0; q < 6; ++q) bound = fmaxf(bound, fabsf(box[q]));
    atomicMax((unsigned int*)a.coord_bound, __float_as_uint(bound));     // non-negative floats order like their bits
}

// ------------------------------------------------------------------------------------------------------------
// std::sort emulation
struct Seg { uint32_t lo, hi, d; };

// Where the sort workspace lives.  Shared memory is addressed through the shared window (ld.shared / st.shared):
// the sort is a chain of dependent small accesses, and a generic load of shared memory costs several times an LDS.
__device__ __forceinline__ float lds_f(uint32_t a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_f(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void sts_u(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

struct SharedMem {                 // byte addresses in the shared window
    static constexpr uint32_t kLong = 2048;        // CTA-wide partitions run one after the other: measured slower at 768
    uint32_t key_, idx_, posl_, posr_, q_[2];
    __device__ __forceinline__ float key(uint32_t i) const { return lds_f(key_ + 4u * i); }
    __device__ __forceinline__ uint32_t idx(uint32_t i) const { return lds_u(idx_ + 4u * i); }
    __device__ __forceinline__ uint32_t posl(uint32_t i) const { return lds_u(posl_ + 4u * i); }
    __device__ __forceinline__ uint32_t posr(uint32_t i) const { return lds_u(posr_ + 4u * i); }
    __device__ __forceinline__ void set_key(uint32_t i, float v) const { sts_f(key_ + 4u * i, v); }
    __device__ __forceinline__ void set_idx(uint32_t i, uint32_t v) const { sts_u(idx_ + 4u * i, v); }
    __device__ __forceinline__ void set_posl(uint32_t i, uint32_t v) const { sts_u(posl_ + 4u * i, v); }
    __device__ __forceinline__ void set_posr(uint32_t i, uint32_t v) const { sts_u(posr_ + 4u * i, v); }
    __device__ __forceinline__ Seg seg(int nq, uint32_t e) const {
        const uint32_t a = q_[nq] + 12u * e;
        return Seg{lds_u(a), lds_u(a + 4), lds_u(a + 8)};
    }
    __device__ __forceinline__ void set_seg(int nq, uint32_t e, Seg v) const {
        const uint32_t a = q_[nq] + 12u * e;
        sts_u(a, v.lo); sts_u(a + 4, v.hi); sts_u(a + 8, v.d);
    }
};
struct GlobalMem {                 // global scratch (nodes too large for shared memory)
    static constexpr uint32_t kLong = 2048;
    float* key_; uint32_t* idx_; uint32_t* posl_; uint32_t* posr_; Seg* q_[2];
    __device__ __forceinline__ float key(uint32_t i) const { return key_[i]; }
    __device__ __forceinline__ uint32_t idx(uint32_t i) const { return idx_[i]; }
    __device__ __forceinline__ uint32_t posl(uint32_t i) const { return posl_[i]; }
    __device__ __forceinline__ uint32_t posr(uint32_t i) const { return posr_[i]; }
    __device__ __forceinline__ void set_key(uint32_t i, float v) const { key_[i] = v; }
    __device__ __forceinline__ void set_idx(uint32_t i, uint32_t v) const { idx_[i] = v; }
    __device__ __forceinline__ void set_posl(uint32_t i, uint32_t v) const { posl_[i] = v; }
    __device__ __forceinline__ void set_posr(uint32_t i, uint32_t v) const { posr_[i] = v; }
    __device__ __forceinline__ Seg seg(int nq, uint32_t e) const { return q_[nq][e]; }
    __device__ __forceinline__ void set_seg(int nq, uint32_t e, Seg v) const { q_[nq][e] = v; }
};

template <class M>
struct SortSpace {
    M m;
    uint32_t* qcount;              // [4] in shared memory: entries of q[0], q[1]; of those, segments of at least M::kLong
    uint32_t* warp_cnt;            // [64] in shared memory (CTA-wide partition)
    uint32_t* sh_k;                // in shared memory
    unsigned long long* dbg;       // optional: per round (segments << 40 | cycles), 64 slots (SRT_PREP_DEBUG)
};

template <class M>
__device__ __forceinline__ void swap_pair(const SortSpace<M>& s, uint32_t x, uint32_t y) {
    const float kx = s.m.key(x), ky = s.m.key(y);
    const uint32_t ix = s.m.idx(x), iy = s.m.idx(y);
    s.m.set_key(x, ky); s.m.set_key(y, kx);
    s.m.set_idx(x, iy); s.m.set_idx(y, ix);
}

// std::__move_median_to_first(first, first + 1, mid, last - 1) — bits/stl_algo.h
template <class M>
__device__ __forceinline__ void median_to_first(const SortSpace<M>& s, uint32_t first, uint32_t last) {
    const uint32_t A = first + 1, B = first + (last - first) / 2, C = last - 1;
    const float ka = s.m.key(A), kb = s.m.key(B), kc = s.m.key(C);
    uint32_t m;
    if (ka < kb) { if (kb < kc) m = B; else if (ka < kc) m = C; else m = A; }
    else if (ka < kc) m = A;
    else if (kb < kc) m = C;
    else m = B;
    swap_pair(s, first, m);
}

template <class M>
__device__ __forceinline__ void push_segment(const SortSpace<M>& s, int nq, uint32_t lo, uint32_t hi, uint32_t d) {
    if (hi - lo <= 16) return;
    const uint32_t e = atomicAdd(s.qcount + nq, 1u);
    s.m.set_seg(nq, e, Seg{lo, hi, d});
    if (hi - lo >= M::kLong && d > 0) atomicAdd(s.qcount + 2 + nq, 1u);
}
template <class M>
__device__ __forceinline__ void push_children(const SortSpace<M>& s, int nq, uint32_t first, uint32_t cut, uint32_t last, uint32_t d) {
    push_segment(s, nq, cut, last, d);
    push_segment(s, nq, first, cut, d);
}

// One Hoare partition (std::__unguarded_partition_pivot) by one warp.
template <class M>
__device__ void warp_partition(const SortSpace<M>& s, Seg sg, int nq, unsigned lane) {
    const uint32_t first = sg.lo, last = sg.hi;
    if (lane == 0) median_to_first(s, first, last);
    __syncwarp();
    const float p = s.m.key(first);
    const unsigned lt = (1u << lane) - 1u;
    uint32_t TL = 0, TR = 0;
    uint32_t base = first + 1;
    for (; base + 128 <= last; base += 128) {                            // four chunks per trip: the loads go out together
        float k[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) k[u] = s.m.key(base + 32u * u + lane);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t x = base + 32u * u + lane;
            const bool fl = !(k[u] < p), fr = !(p < k[u]);
            const unsigned bl = __ballot_sync(kFull, fl), br = __ballot_sync(kFull, fr);
            if (fl) s.m.set_posl(first + TL + __popc(bl & lt), x);
            if (fr) s.m.set_posr(first + TR + __popc(br & lt), x);
            TL += __popc(bl); TR += __popc(br);
        }
    }
    for (; base < last; base += 32) {
        const uint32_t x = base + lane;
        const bool valid = x < last;
        const float k = valid ? s.m.key(x) : 0.f;
        const bool fl = valid && !(k < p), fr = valid && !(p < k);
        const unsigned bl = __ballot_sync(kFull, fl), br = __ballot_sync(kFull, fr);
        if (fl) s.m.set_posl(first + TL + __popc(bl & lt), x);
        if (fr) s.m.set_posr(first + TR + __popc(br & lt), x);
        TL += __popc(bl); TR += __popc(br);
    }
    __syncwarp();
    const uint32_t mn = TL < TR ? TL : TR;
    uint32_t K = 0;
    for (uint32_t kb = 0; kb < mn; kb += 32) {
        const uint32_t k = kb + lane;
        bool sw = false;
        uint32_t l = 0, r = 0;
        if (k < mn) { l = s.m.posl(first + k); r = s.m.posr(first + TR - 1 - k); sw = l < r; }
        const unsigned b = __ballot_sync(kFull, sw);
        if (sw) swap_pair(s, l, r);
        K += __popc(b);
        if (b != kFull) break;
    }
    __syncwarp();
    if (lane == 0) {
        uint32_t cut = kNoPos;
        if (K < TL) cut = s.m.posl(first + K);
        if (K >= 1) { const uint32_t r = s.m.posr(first + TR - K); cut = r < cut ? r : cut; }
        push_children(s, nq, first, cut, last, sg.d - 1);
    }
    __syncwarp();
}

// position of the n-th (n >= 1) set bit of mask counted from bit 0 / from bit 31; mask has at least n bits set
__device__ __forceinline__ uint32_t nth_set_lo(uint32_t mask, uint32_t n) {
    uint32_t pos = 0;
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1)
        if ((uint32_t)__popc(mask & ((1u << (pos + step)) - 1u)) < n) pos += step;
    return pos;
}
__device__ __forceinline__ uint32_t nth_set_hi(uint32_t mask, uint32_t n) { return 31u - nth_set_lo(__brev(mask), n); }

template <class M>
__device__ void serial_heapsort(const SortSpace<M>& s, uint32_t f, long len);

// The rest of the introsort loop of a segment of at most 32 elements, one element per lane, in registers: the same
// median / pivot / pairing rules as warp_partition with ballots instead of position lists.
template <class M>
__device__ void warp_small_sort(const SortSpace<M>& s, Seg sg, unsigned lane) {
    const uint32_t first = sg.lo, n = sg.hi - sg.lo;
    float k = lane < n ? s.m.key(first + lane) : 0.f;
    uint32_t id = lane < n ? s.m.idx(first + lane) : 0u;
    const unsigned lt = (1u << lane) - 1u, gt = ~((2u << lane) - 1u);
    uint32_t lo = 0, hi = n, d = sg.d;                     // at most one side of a cut is longer than 16: no stack needed
    while (hi - lo > 16) {
        if (d == 0) {                                       // depth limit: libstdc++'s heap sort, by one lane, in memory
            if (lane >= lo && lane < hi) { s.m.set_key(first + lane, k); s.m.set_idx(first + lane, id); }
            __syncwarp();
            if (lane == 0) serial_heapsort(s, first + lo, (long)(hi - lo));
            __syncwarp();
            if (lane >= lo && lane < hi) { k = s.m.key(first + lane); id = s.m.idx(first + lane); }
            break;
        }
        const uint32_t A = lo + 1, B = lo + (hi - lo) / 2, C = hi - 1;
        const float ka = __shfl_sync(kFull, k, A), kb = __shfl_sync(kFull, k, B), kc = __shfl_sync(kFull, k, C);
        uint32_t m;
        if (ka < kb) { if (kb < kc) m = B; else if (ka < kc) m = C; else m = A; }
        else if (ka < kc) m = A;
        else if (kb < kc) m = C;
        else m = B;
        {
            const float k_lo = __shfl_sync(kFull, k, lo), k_m = __shfl_sync(kFull, k, m);
            const uint32_t i_lo = __shfl_sync(kFull, id, lo), i_m = __shfl_sync(kFull, id, m);
            if (lane == lo) { k = k_m; id = i_m; } else if (lane == m) { k = k_lo; id = i_lo; }
        }
        const float p = __shfl_sync(kFull, k, lo);
        const bool in = lane > lo && lane < hi;
        const bool fl = in && !(k < p), fr = in && !(p < k);
        const unsigned bl = __ballot_sync(kFull, fl), br = __ballot_sync(kFull, fr);
        const uint32_t TL = __popc(bl), TR = __popc(br);
        uint32_t partner = lane;
        bool swl = false;
        if (fl) {                                           // k-th from the left not less than the pivot <-> k-th from the right not greater
            const uint32_t rk = __popc(bl & lt) + 1;
            if (rk <= TR) { const uint32_t r = nth_set_hi(br, rk); if (lane < r) { partner = r; swl = true; } }
        }
        if (fr && !swl) {
            const uint32_t rk = __popc(br & gt) + 1;
            if (rk <= TL) { const uint32_t l = nth_set_lo(bl, rk); if (l < lane) partner = l; }
        }
        const uint32_t K = __popc(__ballot_sync(kFull, swl));
        k = __shfl_sync(kFull, k, partner); id = __shfl_sync(kFull, id, partner);
        uint32_t cut = kNoPos;
        if (K < TL) cut = nth_set_lo(bl, K + 1);
        if (K >= 1) { const uint32_t r = nth_set_hi(br, K); cut = r < cut ? r : cut; }
        --d;
        if (hi - cut > 16) lo = cut; else hi = cut;          // continue with the side that is still longer than 16 (if any)
    }
    if (lane < n) { s.m.set_key(first + lane, k); s.m.set_idx(first + lane, id); }
    __syncwarp();
}

// The same partition by the whole CTA (kBuildThreads threads; every thread calls it with the same segment).
template <class M>
__device__ void cta_partition(const SortSpace<M>& s, Seg sg, int nq) {
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t first = sg.lo, last = sg.hi;
    if (tid == 0) { median_to_first(s, first, last); *s.sh_k = 0; }
    __syncthreads();
    const float p = s.m.key(first);
    const uint32_t lo = first + 1, len = last - lo;
    const uint32_t per = ((len + 31u) / 32u + 31u) & ~31u;          // slice per warp, a multiple of 32
    const uint32_t s0 = lo + warp * per < last ? lo + warp * per : last;
    const uint32_t s1 = s0 + per < last ? s0 + per : last;
    uint32_t tl = 0, tr = 0;
    for (uint32_t x = s0 + lane; x < s1; x += 32) {
        const float k = s.m.key(x);
        tl += !(k < p); tr += !(p < k);
    }
    tl = __reduce_add_sync(kFull, tl); tr = __reduce_add_sync(kFull, tr);
    if (lane == 0) { s.warp_cnt[warp] = tl; s.warp_cnt[32 + warp] = tr; }
    __syncthreads();
    const uint32_t all_l = s.warp_cnt[lane], all_r = s.warp_cnt[32 + lane];
    const uint32_t TL = __reduce_add_sync(kFull, all_l), TR = __reduce_add_sync(kFull, all_r);
    uint32_t bl0 = __reduce_add_sync(kFull, lane < warp ? all_l : 0u), br0 = __reduce_add_sync(kFull, lane < warp ? all_r : 0u);
    const unsigned lt = (1u << lane) - 1u;
    for (uint32_t base = s0; base < s1; base += 128) {
        float k[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { const uint32_t x = base + 32u * u + lane; k[u] = x < s1 ? s.m.key(x) : 0.f; }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t x = base + 32u * u + lane;
            const bool valid = x < s1;
            const bool fl = valid && !(k[u] < p), fr = valid && !(p < k[u]);
            const unsigned bl = __ballot_sync(kFull, fl), br = __ballot_sync(kFull, fr);
            if (fl) s.m.set_posl(first + bl0 + __popc(bl & lt), x);
            if (fr) s.m.set_posr(first + br0 + __popc(br & lt), x);
            bl0 += __popc(bl); br0 += __popc(br);
        }
    }
    __syncthreads();
    const uint32_t mn = TL < TR ? TL : TR;
    for (uint32_t k = tid; k < mn; k += kBuildThreads) {
        const uint32_t l = s.m.posl(first + k), r = s.m.posr(first + TR - 1 - k);
        if (l < r) {
            swap_pair(s, l, r);
            const bool last_pair = (k + 1 == mn) || !(s.m.posl(first + k + 1) < s.m.posr(first + TR - 2 - k));
            if (last_pair) *s.sh_k = k + 1;
        }
    }
    __syncthreads();
    if (tid == 0) {
        const uint32_t K = *s.sh_k;
        uint32_t cut = kNoPos;
        if (K < TL) cut = s.m.posl(first + K);
        if (K >= 1) { const uint32_t r = s.m.posr(first + TR - K); cut = r < cut ? r : cut; }
        push_children(s, nq, first, cut, last, sg.d - 1);
    }
    __syncthreads();
}

// std::__adjust_heap + std::__push_heap (bits/stl_heap.h) on key/idx pairs starting at `f`.
template <class M>
__device__ void heap_adjust(const SortSpace<M>& s, uint32_t f, long hole, long len, float vk, uint32_t vi) {
    const long top = hole;
    long sc = hole;
    while (sc < (len - 1) / 2) {
        sc = 2 * (sc + 1);
        if (s.m.key(f + sc) < s.m.key(f + sc - 1)) sc--;
        s.m.set_key(f + hole, s.m.key(f + sc)); s.m.set_idx(f + hole, s.m.idx(f + sc));
        hole = sc;
    }
    if ((len & 1) == 0 && sc == (len - 2) / 2) {
        sc = 2 * (sc + 1);
        s.m.set_key(f + hole, s.m.key(f + sc - 1)); s.m.set_idx(f + hole, s.m.idx(f + sc - 1));
        hole = sc - 1;
    }
    long parent = (hole - 1) / 2;
    while (hole > top && s.m.key(f + parent) < vk) {
        s.m.set_key(f + hole, s.m.key(f + parent)); s.m.set_idx(f + hole, s.m.idx(f + parent));
        hole = parent;
        parent = (hole - 1) / 2;
    }
    s.m.set_key(f + hole, vk); s.m.set_idx(f + hole, vi);
}

// std::__partial_sort(first, last, last) = __make_heap + __sort_heap: the depth-limit fallback of __introsort_loop.
template <class M>
__device__ void serial_heapsort(const SortSpace<M>& s, uint32_t f, long len) {
    if (len >= 2) {
        long parent = (len - 2) / 2;
        while (true) {
            heap_adjust(s, f, parent, len, s.m.key(f + parent), s.m.idx(f + parent));
            if (parent == 0) break;
            parent--;
        }
    }
    long last = len;
    while (last > 1) {
        --last;
        const float vk = s.m.key(f + last); const uint32_t vi = s.m.idx(f + last);
        s.m.set_key(f + last, s.m.key(f)); s.m.set_idx(f + last, s.m.idx(f));
        heap_adjust(s, f, 0, last, vk, vi);
    }
}

// Runs the introsort loop of every segment queued in q[0] (all threads of the CTA).
template <class M>
__device__ void sort_rounds(const SortSpace<M>& s) {
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    int cur = 0, round = 0;
    long long t_prev = s.dbg ? clock64() : 0;
    while (true) {
        __syncthreads();
        const uint32_t n = s.qcount[cur];
        if (s.dbg && tid == 0 && round < 64) {
            const long long t = clock64();
            s.dbg[round] = ((unsigned long long)n << 40) | (unsigned long long)(t - t_prev);
            t_prev = t;
            ++round;
        }
        if (n == 0) break;
        const uint32_t n_long = s.qcount[2 + cur];
        if (tid == 0) { s.qcount[cur ^ 1] = 0; s.qcount[2 + (cur ^ 1)] = 0; }
        __syncthreads();
        for (uint32_t e = 0; n_long != 0 && e < n; ++e) {        // the long segments, one after the other, CTA-wide
            const Seg sg = s.m.seg(cur, e);
            if (sg.hi - sg.lo >= M::kLong && sg.d > 0) cta_partition(s, sg, cur ^ 1);
        }
        for (uint32_t e = warp; e < n; e += kBuildThreads / 32) {     // the others, one warp each
            const Seg sg = s.m.seg(cur, e);
            if (sg.d == 0) { if (lane == 0) serial_heapsort(s, sg.lo, (long)(sg.hi - sg.lo)); }
            else if (sg.hi - sg.lo <= 32) warp_small_sort(s, sg, lane);
            else if (sg.hi - sg.lo < M::kLong) warp_partition(s, sg, cur ^ 1, lane);
        }
        cur ^= 1;
    }
}

__device__ __forceinline__ uint32_t lg2_floor(uint32_t n) { return 31u - __clz(n); }

// ------------------------------------------------------------------------------------------------------------
// Large nodes: the same introsort loop, one node's segment spread over many CTAs.  A round is three launches over
// the tiles (1024 elements) of all segments longer than kWideSmall: count the elements not less / not greater than
// the pivot per tile; rank them (a tile adds the counts of the tiles before it) and record their positions; swap the
// pairs and, in the CTA that finishes a segment last, cut it and set up its two sides for the next round.  Shorter
// segments, and whatever is left after the planned number of rounds, are finished by one CTA each.
struct SortShared {
    uint32_t qcount[4];
    uint32_t warp_cnt[64];
    uint32_t sh_k;
};

// median of (first + 1, mid, last - 1) to first, reading around L1 (other CTAs of this launch wrote these keys);
// returns the pivot.
__device__ float median_to_first_cg(float* key, uint32_t* idx, uint32_t first, uint32_t last) {
    const uint32_t A = first + 1, B = first + (last - first) / 2, C = last - 1;
    const float ka = __ldcg(key + A), kb = __ldcg(key + B), kc = __ldcg(key + C);
    uint32_t m; float km;
    if (ka < kb) { if (kb < kc) { m = B; km = kb; } else if (ka < kc) { m = C; km = kc; } else { m = A; km = ka; } }
    else if (ka < kc) { m = A; km = ka; }
    else if (kb < kc) { m = C; km = kc; }
    else { m = B; km = kb; }
    const float kf = __ldcg(key + first);
    const uint32_t jf = __ldcg(idx + first), jm = __ldcg(idx + m);
    key[first] = km; key[m] = kf;
    idx[first] = jm; idx[m] = jf;
    return km;
}

// One thread: queue [lo, hi) for round `round`, or hand it to the finishing CTAs.
__device__ void wide_emit(const PrepArgs& a, const WideArgs& w, uint32_t lo, uint32_t hi, uint32_t d, uint32_t round, bool no_more_rounds) {
    const uint32_t len = hi - lo;
    if (len <= 16) return;
    if (len <= kWideSmall || d == 0 || no_more_rounds) {
        const uint32_t e = atomicAdd(w.fin_count, 1u);
        if (e < w.max_fin) w.fin[e] = WideSmall{lo, hi, d}; else atomicOr(a.status, 2u);
        return;
    }
    const uint32_t nt = (len - 1 + kWideTile - 1) / kWideTile;
    const uint32_t t0 = atomicAdd(&w.rounds[round].ntiles, nt);
    const uint32_t sid = atomicAdd(&w.rounds[round].nsegs, 1u);
    if (sid >= w.max_segs || t0 + nt > w.max_tiles) { atomicOr(a.status, 2u); return; }
    const float pivot = median_to_first_cg(a.key, a.idx, lo, hi);
    w.segs[(size_t)(round & 1u) * w.max_segs + sid] = WideSeg{lo, hi, d, pivot, t0, nt, 0u, 0u};
    uint2* ts = w.tile_seg + (size_t)(round & 1u) * w.max_tiles + t0;
    for (uint32_t i = 0; i < nt; ++i) ts[i] = make_uint2(sid, i);
}

// keys of the level's nodes: pointOne.<axis> of every triangle in its current order
__global__ void __launch_bounds__(256) wide_begin_kernel(const PrepArgs a, const WorkItem* __restrict__ items) {
    const WorkItem it = items[blockIdx.y];
    if (blockIdx.x * kWideTile >= it.count) return;
    const PrepObject o = a.objects[it.obj];
    const float4* n = a.nodes + 2 * ((size_t)o.node_base + it.node);
    const float4 b0 = n[0], b1 = n[1];
    const float sx = fabsf(__fsub_rn(b1.x, b0.x)), sy = fabsf(__fsub_rn(b1.y, b0.y)), sz = fabsf(__fsub_rn(b1.z, b0.z));
    const uint32_t ax = (sx > sy && sx > sz) ? 0u : ((sy > sx && sy > sz) ? 1u : 2u);
    const size_t gbase = (size_t)o.tri_base + it.start;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const uint32_t li = blockIdx.x * kWideTile + u * 256u + threadIdx.x;
        if (li < it.count) {
            const uint32_t tri = a.order[gbase + li];
            const float4 p1 = a.xpts[3 * ((size_t)o.tri_base + tri)];
            a.idx[gbase + li] = tri;
            a.key[gbase + li] = ax == 0 ? p1.x : (ax == 1 ? p1.y : p1.z);
        }
    }
}

__global__ void wide_init_kernel(const PrepArgs a, const WideArgs w, const WorkItem* __restrict__ items, const uint32_t n_items, const uint32_t rounds) {
    for (uint32_t r = threadIdx.x; r <= (uint32_t)kWideMaxRounds; r += blockDim.x) w.rounds[r] = WideRound{0u, 0u};
    if (threadIdx.x == 0) *w.fin_count = 0;
    for (uint32_t i = threadIdx.x; i < n_items * 12u; i += blockDim.x) w.childacc[i] = seed_of((int)(i % 6u));
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n_items; i += blockDim.x) {
        const WorkItem it = items[i];
        const uint32_t lo = a.objects[it.obj].tri_base + it.start;
        if (it.count > 16) wide_emit(a, w, lo, lo + it.count, 2u * lg2_floor(it.count), 0, rounds == 0);
    }
}

__global__ void __launch_bounds__(256) wide_count_kernel(const PrepArgs a, const WideArgs w, const uint32_t round) {
    const uint32_t t = blockIdx.x;
    if (t >= w.rounds[round].ntiles) return;
    const uint2 ts = w.tile_seg[(size_t)(round & 1u) * w.max_tiles + t];
    const WideSeg S = w.segs[(size_t)(round & 1u) * w.max_segs + ts.x];
    const uint32_t x0 = S.lo + 1 + ts.y * kWideTile, x1 = x0 + kWideTile < S.hi ? x0 + kWideTile : S.hi;
    const float p = S.pivot;
    uint32_t cl = 0, cr = 0;
    for (uint32_t x = x0 + threadIdx.x; x < x1; x += 256) { const float k = a.key[x]; cl += !(k < p); cr += !(p < k); }
    cl = __reduce_add_sync(kFull, cl); cr = __reduce_add_sync(kFull, cr);
    __shared__ uint32_t sl[8], sr[8];
    if ((threadIdx.x & 31u) == 0) { sl[threadIdx.x >> 5] = cl; sr[threadIdx.x >> 5] = cr; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tl = 0, tr = 0;
        for (int i = 0; i < 8; ++i) { tl += sl[i]; tr += sr[i]; }
        w.tile_cnt[t] = make_uint2(tl, tr);
    }
}

__global__ void __launch_bounds__(256) wide_scatter_kernel(const PrepArgs a, const WideArgs w, const uint32_t round) {
    const uint32_t t = blockIdx.x;
    if (t >= w.rounds[round].ntiles) return;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint2 ts = w.tile_seg[(size_t)(round & 1u) * w.max_tiles + t];
    const WideSeg S = w.segs[(size_t)(round & 1u) * w.max_segs + ts.x];
    const uint32_t x0 = S.lo + 1 + ts.y * kWideTile + warp * 128u, xe = S.lo + 1 + (ts.y + 1) * kWideTile;
    const uint32_t x1 = xe < S.hi ? xe : S.hi;
    const float p = S.pivot;
    __shared__ uint32_t sl[8], sr[8], base[2];
    unsigned bl[4], br[4];
    uint32_t wl = 0, wr = 0;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const uint32_t x = x0 + 32u * u + lane;
        const bool valid = x < x1;
        const float k = valid ? a.key[x] : 0.f;
        bl[u] = __ballot_sync(kFull, valid && !(k < p));
        br[u] = __ballot_sync(kFull, valid && !(p < k));
        wl += __popc(bl[u]); wr += __popc(br[u]);
    }
    if (lane == 0) { sl[warp] = wl; sr[warp] = wr; }
    if (warp == 0) {                                             // ranks before this tile: the counts of the segment's earlier tiles
        uint32_t pl = 0, pr = 0;
        for (uint32_t i = lane; i < ts.y; i += 32) { const uint2 c = w.tile_cnt[S.tile_first + i]; pl += c.x; pr += c.y; }
        pl = __reduce_add_sync(kFull, pl); pr = __reduce_add_sync(kFull, pr);
        if (lane == 0) { base[0] = pl; base[1] = pr; }
    }
    __syncthreads();
    uint32_t offl = base[0], offr = base[1];
    for (unsigned i = 0; i < warp; ++i) { offl += sl[i]; offr += sr[i]; }
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const uint32_t x = x0 + 32u * u + lane;
        if ((bl[u] >> lane) & 1u) a.posl[S.lo + offl + __popc(bl[u] & lt)] = x;
        if ((br[u] >> lane) & 1u) a.posr[S.lo + offr + __popc(br[u] & lt)] = x;
        offl += __popc(bl[u]); offr += __popc(br[u]);
    }
}

__global__ void __launch_bounds__(256) wide_swap_kernel(const PrepArgs a, const WideArgs w, const uint32_t round, const uint32_t last_round) {
    const uint32_t t = blockIdx.x;
    if (t >= w.rounds[round].ntiles) return;
    const unsigned lane = threadIdx.x & 31u;
    const uint2 ts = w.tile_seg[(size_t)(round & 1u) * w.max_tiles + t];
    WideSeg* sp = w.segs + (size_t)(round & 1u) * w.max_segs + ts.x;
    const uint32_t lo = sp->lo, hi = sp->hi, d = sp->d, tile_first = sp->tile_first, ntiles = sp->ntiles;
    __shared__ uint32_t tot[2];
    __shared__ int is_last;
    if (threadIdx.x < 32) {
        uint32_t pl = 0, pr = 0;
        for (uint32_t i = lane; i < ntiles; i += 32) { const uint2 c = w.tile_cnt[tile_first + i]; pl += c.x; pr += c.y; }
        pl = __reduce_add_sync(kFull, pl); pr = __reduce_add_sync(kFull, pr);
        if (lane == 0) { tot[0] = pl; tot[1] = pr; }
    }
    __syncthreads();
    const uint32_t TL = tot[0], TR = tot[1], mn = TL < TR ? TL : TR;
    const uint32_t k1 = (ts.y + 1) * kWideTile < mn ? (ts.y + 1) * kWideTile : mn;
    for (uint32_t k = ts.y * kWideTile + threadIdx.x; k < k1; k += 256) {
        const uint32_t l = a.posl[lo + k], r = a.posr[lo + TR - 1 - k];
        if (l < r) {
            const float kl = a.key[l], kr = a.key[r];
            const uint32_t il = a.idx[l], ir = a.idx[r];
            a.key[l] = kr; a.key[r] = kl; a.idx[l] = ir; a.idx[r] = il;
            const bool last_pair = (k + 1 == mn) || !(a.posl[lo + k + 1] < a.posr[lo + TR - 2 - k]);
            if (last_pair) sp->K = k + 1;
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(&sp->done, 1u) == ntiles - 1;
    __syncthreads();
    if (!is_last || threadIdx.x != 0) return;
    __threadfence();
    const uint32_t K = __ldcg(&sp->K);
    uint32_t cut = kNoPos;
    if (K < TL) cut = a.posl[lo + K];
    if (K >= 1) { const uint32_t r = a.posr[lo + TR - K]; cut = r < cut ? r : cut; }
    wide_emit(a, w, cut, hi, d - 1, round + 1, last_round != 0);
    wide_emit(a, w, lo, cut, d - 1, round + 1, last_round != 0);
}

// The segments left to one CTA each: the rest of their introsort loop in shared memory (global memory if longer).
__global__ void __launch_bounds__(kBuildThreads, 1) wide_finish_kernel(const PrepArgs a, const WideArgs w) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortShared& sh = *reinterpret_cast<SortShared*>(smem_raw);
    Seg* q0 = reinterpret_cast<Seg*>(smem_raw + 512);
    constexpr uint32_t kQ = kWideSmall / 16;
    float* wsp = reinterpret_cast<float*>(smem_raw + 512 + 2 * kQ * sizeof(Seg));
    const uint32_t n = *w.fin_count < w.max_fin ? *w.fin_count : w.max_fin;
    const uint32_t sh_base = (uint32_t)__cvta_generic_to_shared(wsp);
    for (uint32_t e = blockIdx.x; e < n; e += gridDim.x) {
        const WideSmall f = w.fin[e];
        const uint32_t len = f.hi - f.lo;
        unsigned long long* dbg = (a.dbg && e < 8) ? a.dbg + 64 * e : nullptr;
        if (dbg && threadIdx.x == 0) for (int i = 0; i < 64; ++i) dbg[i] = 0;
        if (threadIdx.x == 0) sh.qcount[0] = sh.qcount[1] = sh.qcount[2] = sh.qcount[3] = 0;
        if (len <= kWideSmall) {
            SortSpace<SharedMem> s;
            s.qcount = sh.qcount; s.warp_cnt = sh.warp_cnt; s.sh_k = &sh.sh_k; s.dbg = dbg;
            s.m.key_ = sh_base; s.m.idx_ = sh_base + 4u * kWideSmall; s.m.posl_ = sh_base + 8u * kWideSmall; s.m.posr_ = sh_base + 12u * kWideSmall;
            s.m.q_[0] = (uint32_t)__cvta_generic_to_shared(q0); s.m.q_[1] = s.m.q_[0] + 12u * kQ;
            for (uint32_t i = threadIdx.x; i < len; i += kBuildThreads) { s.m.set_key(i, a.key[f.lo + i]); s.m.set_idx(i, a.idx[f.lo + i]); }
            if (threadIdx.x == 0) push_segment(s, 0, 0u, len, f.d);
            sort_rounds(s);
            for (uint32_t i = threadIdx.x; i < len; i += kBuildThreads) { a.key[f.lo + i] = s.m.key(i); a.idx[f.lo + i] = s.m.idx(i); }
        } else {
            SortSpace<GlobalMem> s;
            s.qcount = sh.qcount; s.warp_cnt = sh.warp_cnt; s.sh_k = &sh.sh_k; s.dbg = dbg;
            s.m.key_ = a.key; s.m.idx_ = a.idx; s.m.posl_ = a.posl; s.m.posr_ = a.posr;
            s.m.q_[0] = reinterpret_cast<Seg*>(a.segq) + f.lo / 8;
            s.m.q_[1] = reinterpret_cast<Seg*>(a.segq) + a.segq_stride + f.lo / 8;
            if (threadIdx.x == 0) push_segment(s, 0, f.lo, f.hi, f.d);
            sort_rounds(s);
        }
        __syncthreads();
    }
}

// __final_insertion_sort of the level's nodes (stable, window of 15) -> order[], and the child boxes as packed minima.
__global__ void __launch_bounds__(256) wide_final_kernel(const PrepArgs a, const WideArgs w, const WorkItem* __restrict__ items) {
    const WorkItem it = items[blockIdx.y];
    if (blockIdx.x * kWideTile >= it.count) return;
    const PrepObject o = a.objects[it.obj];
    const uint32_t ns = o.tri_base + it.start, ne = ns + it.count, half = it.count / 2;
    const unsigned lane = threadIdx.x & 31u;
#pragma unroll 1
    for (int u = 0; u < 4; ++u) {
        const uint32_t x = ns + blockIdx.x * kWideTile + u * 256u + threadIdx.x;
        const bool active = x < ne;
        unsigned long long acc[6];
#pragma unroll
        for (int q = 0; q < 6; ++q) acc[q] = seed_of(q);
        uint32_t child = 0;
        if (active) {
            const float k = a.key[x];
            uint32_t dst = x;
            const uint32_t y0 = x >= ns + 15 ? x - 15 : ns, y1 = x + 16 < ne ? x + 16 : ne;
            for (uint32_t y = y0; y < x; ++y) dst -= (a.key[y] > k);
            for (uint32_t y = x + 1; y < y1; ++y) dst += (a.key[y] < k);
            const uint32_t tri = a.idx[x];
            a.order[dst] = tri;
            child = (dst - ns) < half ? 0u : 1u;
            const uint32_t pos = dst - o.tri_base;                      // position in the object's order
            const float2* lp = reinterpret_cast<const float2*>(a.lite + 6 * ((size_t)o.tri_base + tri));
            const float2 v0 = lp[0], v1 = lp[1], v2 = lp[2];
            const float v[6] = {v0.x, v0.y, v1.x, v1.y, v2.x, v2.y};
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                if (v[q] == v[q] && v[q] < FLT_MAX) acc[q] = pack_min(v[q], pos);
                if (v[3 + q] == v[3 + q] && v[3 + q] > -FLT_MAX) acc[3 + q] = pack_max(v[3 + q], pos);
            }
        }
        const unsigned m_act = __ballot_sync(kFull, active), m_one = __ballot_sync(kFull, active && child == 1u);
        if (m_act == 0) continue;
        unsigned long long* dstacc = w.childacc + (size_t)blockIdx.y * 12u;
        if (m_one == 0 || m_one == m_act) {
#pragma unroll
            for (int q = 0; q < 6; ++q) {
                const unsigned long long v = warp_min64(acc[q]);
                if (lane == 0 && v != seed_of(q)) atomicMin(dstacc + (m_one ? 6 : 0) + q, v);
            }
        } else if (active) {
#pragma unroll
            for (int q = 0; q < 6; ++q) if (acc[q] != seed_of(q)) atomicMin(dstacc + child * 6u + q, acc[q]);
        }
    }
}

__global__ void wide_children_kernel(const PrepArgs a, const WideArgs w, const WorkItem* __restrict__ items, const uint32_t n_items) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * n_items) return;
    const WorkItem it = items[i >> 1];
    const PrepObject o = a.objects[it.obj];
    float box[6];
    for (int q = 0; q < 6; ++q) box[q] = decode_sorted(a, o, w.childacc[(size_t)i * 6u + q], q);
    const uint32_t child = (i & 1u) ? o.topo[it.node].right : it.node + 1;
    write_node(a, o, child, box);
}

// ------------------------------------------------------------------------------------------------------------
struct Level {                     // nodes of the current level inside the CTA's range (shared memory), ascending by start
    uint32_t node[kNLMax], start[kNLMax], count[kNLMax];     // start: position relative to the CTA's range
    uint32_t axis[kNLMax];
};

__device__ __forceinline__ int find_node(const Level& L, uint32_t m, uint32_t li) {
    uint32_t lo = 0, hi = m;                                  // largest j with start[j] <= li
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (L.start[mid] <= li) lo = mid; else hi = mid; }
    return (li >= L.start[lo] && li - L.start[lo] < L.count[lo]) ? (int)lo : -1;
}

struct BuildShared {
    Level lev[2];
    uint32_t qcount[4];
    uint32_t warp_cnt[64];
    uint32_t sh_k;
    uint32_t next_m;
    unsigned long long partial[kBuildThreads / 32][6];
    Seg smem_q[2][kSmemQueue];
};

// Box of child `c` (0 = left, 1 = right of level node c >> 1) from warp `g` of `G` cooperating warps.
__device__ __forceinline__ void child_partial(const PrepArgs& a, const PrepObject& o, const Level& L, uint32_t c, uint32_t g, uint32_t G,
                                              uint32_t range_start, unsigned lane, unsigned long long* out6) {
    const uint32_t j = c >> 1, s0 = L.start[j], cnt = L.count[j], half = cnt / 2;
    const uint32_t cs = (c & 1u) ? s0 + half : s0, cc = (c & 1u) ? cnt - half : half;
    unsigned long long acc[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) acc[q] = seed_of(q);
    for (uint32_t i = g * 32u + lane; i < cc; i += G * 32u) {
        const uint32_t pos = range_start + cs + i;                       // position in the object's order
        const uint32_t tri = a.order[(size_t)o.tri_base + pos];
        const float2* lp = reinterpret_cast<const float2*>(a.lite + 6 * ((size_t)o.tri_base + tri));
        const float2 v0 = lp[0], v1 = lp[1], v2 = lp[2];
        const float v[6] = {v0.x, v0.y, v1.x, v1.y, v2.x, v2.y};
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            if (v[q] == v[q] && v[q] < FLT_MAX) acc[q] = min64(acc[q], pack_min(v[q], pos));
            if (v[3 + q] == v[3 + q] && v[3 + q] > -FLT_MAX) acc[3 + q] = min64(acc[3 + q], pack_max(v[3 + q], pos));
        }
    }
#pragma unroll
    for (int q = 0; q < 6; ++q) acc[q] = warp_min64(acc[q]);
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < 6; ++q) out6[q] = acc[q];
    }
}

// Lanes 0..5 hold the six packed minima of child c: decode and write the child's node.
__device__ __forceinline__ void child_finish(const PrepArgs& a, const PrepObject& o, const Level& L, uint32_t c, unsigned long long v,
                                             unsigned lane) {
    const float val = lane < 6 ? decode_sorted(a, o, v, (int)lane) : 0.f;
    float box[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) box[q] = __shfl_sync(kFull, val, q);
    if (lane == 0) {
        const uint32_t parent = L.node[c >> 1];
        const uint32_t child = (c & 1u) ? o.topo[parent].right : parent + 1;
        write_node(a, o, child, box);
    }
}

// One CTA per work item.  subtree = 0: split ONE node (workspace in global scratch); 1: the node and everything below it
// (workspace in shared memory).
template <bool SUBTREE>
__global__ void __launch_bounds__(kBuildThreads, 1) build_kernel(const PrepArgs a, const WorkItem* __restrict__ items) {
    constexpr int subtree = SUBTREE ? 1 : 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BuildShared& sh = *reinterpret_cast<BuildShared*>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const WorkItem it = items[blockIdx.x];
    const PrepObject o = a.objects[it.obj];
    const size_t gbase = (size_t)o.tri_base + it.start;              // scene position of the CTA's range

    SortSpace<typename std::conditional<SUBTREE, SharedMem, GlobalMem>::type> s;
    if constexpr (SUBTREE) {
        const uint32_t w = (uint32_t)__cvta_generic_to_shared(smem_raw + ((sizeof(BuildShared) + 15) & ~size_t(15)));
        s.m.key_ = w; s.m.idx_ = w + 4u * kSubtreeMax; s.m.posl_ = w + 8u * kSubtreeMax; s.m.posr_ = w + 12u * kSubtreeMax;
        s.m.q_[0] = (uint32_t)__cvta_generic_to_shared(sh.smem_q[0]); s.m.q_[1] = (uint32_t)__cvta_generic_to_shared(sh.smem_q[1]);
    } else {
        s.m.key_ = a.key + gbase; s.m.idx_ = a.idx + gbase; s.m.posl_ = a.posl + gbase; s.m.posr_ = a.posr + gbase;
        s.m.q_[0] = reinterpret_cast<Seg*>(a.segq) + gbase / 8;
        s.m.q_[1] = reinterpret_cast<Seg*>(a.segq) + a.segq_stride + gbase / 8;
    }
    s.qcount = sh.qcount; s.warp_cnt = sh.warp_cnt; s.sh_k = &sh.sh_k; s.dbg = nullptr;

    int cl = 0;
    uint32_t m = 1;
    if (tid == 0) { sh.lev[0].node[0] = it.node; sh.lev[0].start[0] = 0; sh.lev[0].count[0] = it.count; }
    __syncthreads();

    while (m > 0) {
        Level& L = sh.lev[cl];
        // axis of every node: the longest side of its box, strictly, else z (:240-248)
        for (uint32_t j = tid; j < m; j += kBuildThreads) {
            const float4* n = a.nodes + 2 * ((size_t)o.node_base + L.node[j]);
            const float4 b0 = n[0], b1 = n[1];
            const float sx = fabsf(__fsub_rn(b1.x, b0.x)), sy = fabsf(__fsub_rn(b1.y, b0.y)), sz = fabsf(__fsub_rn(b1.z, b0.z));
            L.axis[j] = (sx > sy && sx > sz) ? 0u : ((sy > sx && sy > sz) ? 1u : 2u);
        }
        if (tid == 0) { sh.qcount[0] = 0; sh.qcount[1] = 0; sh.qcount[2] = 0; sh.qcount[3] = 0; }
        __syncthreads();
        // keys: pointOne.<axis> of the node's triangles in their current order (:241-247)
        for (uint32_t li = tid; li < it.count; li += kBuildThreads) {
            const int j = find_node(L, m, li);
            if (j < 0) continue;
            const uint32_t tri = a.order[gbase + li];
            const float4 p1 = a.xpts[3 * ((size_t)o.tri_base + tri)];
            const uint32_t ax = L.axis[j];
            s.m.set_idx(li, tri);
            s.m.set_key(li, ax == 0 ? p1.x : (ax == 1 ? p1.y : p1.z));
        }
        for (uint32_t j = tid; j < m; j += kBuildThreads) {
            const uint32_t cnt = L.count[j];
            if (cnt > 16) push_segment(s, 0, L.start[j], L.start[j] + cnt, 2u * lg2_floor(cnt));
        }
        sort_rounds(s);            // starts and ends with a CTA barrier
        // __final_insertion_sort: stable, nothing leaves its (at most 16 long) segment
        for (uint32_t li = tid; li < it.count; li += kBuildThreads) {
            const int j = find_node(L, m, li);
            if (j < 0) continue;
            const uint32_t ns = L.start[j], ne = ns + L.count[j];
            const float k = s.m.key(li);
            uint32_t dst = li;
            const uint32_t y0 = li >= ns + 15 ? li - 15 : ns, y1 = li + 16 < ne ? li + 16 : ne;
            for (uint32_t y = y0; y < li; ++y) dst -= (s.m.key(y) > k);
            for (uint32_t y = li + 1; y < y1; ++y) dst += (s.m.key(y) < k);
            a.order[gbase + dst] = s.m.idx(li);
        }
        __syncthreads();
        // child boxes (:254-259) and the children's nodes
        const uint32_t nchild = 2 * m;
        if (nchild <= kBuildThreads / 32) {
            const uint32_t G = (kBuildThreads / 32) / nchild, c = warp / G, g = warp % G;
            if (c < nchild) child_partial(a, o, L, c, g, G, it.start, lane, sh.partial[warp]);
            __syncthreads();
            if (c < nchild && g == 0) {
                unsigned long long v = 0;
                if (lane < 6) { v = sh.partial[warp][lane]; for (uint32_t gg = 1; gg < G; ++gg) v = min64(v, sh.partial[warp + gg][lane]); }
                child_finish(a, o, L, c, v, lane);
            }
        } else {
            for (uint32_t c = warp; c < nchild; c += kBuildThreads / 32) {
                child_partial(a, o, L, c, 0, 1, it.start, lane, sh.partial[warp]);
                __syncwarp();
                const unsigned long long v = lane < 6 ? sh.partial[warp][lane] : 0ull;
                __syncwarp();
                child_finish(a, o, L, c, v, lane);
            }
        }
        if (!subtree) break;
        // next level: the children with more than 8 triangles (:261-267), in order
        Level& N = sh.lev[cl ^ 1];
        __syncthreads();
        {
            const uint32_t c = tid;
            bool keep = false;
            uint32_t cs = 0, cc = 0, child = 0;
            if (c < nchild) {
                const uint32_t j = c >> 1, cnt = L.count[j], half = cnt / 2;
                cs = (c & 1u) ? L.start[j] + half : L.start[j];
                cc = (c & 1u) ? cnt - half : half;
                child = (c & 1u) ? o.topo[L.node[j]].right : L.node[j] + 1;
                keep = cc > 8;
            }
            const unsigned b = __ballot_sync(kFull, keep);
            if (lane == 0) sh.warp_cnt[warp] = __popc(b);
            __syncthreads();
            const uint32_t mine = sh.warp_cnt[lane];
            const uint32_t before = __reduce_add_sync(kFull, lane < warp ? mine : 0u);
            const uint32_t total = __reduce_add_sync(kFull, mine);
            if (keep) {
                const uint32_t at = before + __popc(b & ((1u << lane) - 1u));
                N.node[at] = child; N.start[at] = cs; N.count[at] = cc;
            }
            if (tid == 0) sh.next_m = total;
            __syncthreads();
            m = sh.next_m;
        }
        cl ^= 1;
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------------------
// srt_flatten.hpp fill_triangle in leaf order (order[] after the build) or load order (identity).
__device__ __forceinline__ void fill_triangle(const PrepArgs& a, const PrepObject& o, uint32_t obj, uint32_t tri, float4* geom, float4* attr) {
    const float4* xp = a.xpts + 3 * ((size_t)o.tri_base + tri);
    const float4 P1 = xp[0], P2 = xp[1], P3 = xp[2];
    // vec3(P) / P.w and the two edges, rayTriangleIntersection :45-51
    const float p1x = __fdiv_rn(P1.x, P1.w), p1y = __fdiv_rn(P1.y, P1.w), p1z = __fdiv_rn(P1.z, P1.w);
    const float p2x = __fdiv_rn(P2.x, P2.w), p2y = __fdiv_rn(P2.y, P2.w), p2z = __fdiv_rn(P2.z, P2.w);
    const float p3x = __fdiv_rn(P3.x, P3.w), p3y = __fdiv_rn(P3.y, P3.w), p3z = __fdiv_rn(P3.z, P3.w);
    const int32_t tex = o.tex ? o.tex[tri] : -1;
    geom[0] = make_float4(p1x, p1y, p1z, __int_as_float((int32_t)obj));
    geom[1] = make_float4(__fsub_rn(p2x, p1x), __fsub_rn(p2y, p1y), __fsub_rn(p2z, p1z), __int_as_float(tex));
    geom[2] = make_float4(__fsub_rn(p3x, p1x), __fsub_rn(p3y, p1y), __fsub_rn(p3z, p1z), __uint_as_float(0u));
    // calculateTriangleNormal :32-37 on the raw xyz; glm::cross, glm::normalize = v * (1 / sqrt(dot))
    const float v1x = __fsub_rn(P2.x, P1.x), v1y = __fsub_rn(P2.y, P1.y), v1z = __fsub_rn(P2.z, P1.z);
    const float v2x = __fsub_rn(P3.x, P1.x), v2y = __fsub_rn(P3.y, P1.y), v2z = __fsub_rn(P3.z, P1.z);
    const float cx = __fsub_rn(__fmul_rn(v1y, v2z), __fmul_rn(v2y, v1z));
    const float cy = __fsub_rn(__fmul_rn(v1z, v2x), __fmul_rn(v2z, v1x));
    const float cz = __fsub_rn(__fmul_rn(v1x, v2y), __fmul_rn(v2x, v1y));
    const float dd = __fadd_rn(__fadd_rn(__fmul_rn(cx, cx), __fmul_rn(cy, cy)), __fmul_rn(cz, cz));
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(dd));
    float2 t0 = make_float2(0.f, 0.f), t1 = t0, t2 = t0;
    if (o.uv) { t0 = o.uv[3 * (size_t)tri]; t1 = o.uv[3 * (size_t)tri + 1]; t2 = o.uv[3 * (size_t)tri + 2]; }
    attr[0] = make_float4(__fmul_rn(cx, inv), __fmul_rn(cy, inv), __fmul_rn(cz, inv), 0.f);
    attr[1] = make_float4(t0.x, t0.y, t1.x, t1.y);
    attr[2] = make_float4(t2.x, t2.y, 0.f, 0.f);
}

__device__ __forceinline__ void fill_vnormal(const PrepObject& o, uint32_t tri, float4* out) {
    float n[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (o.vn) for (int k = 0; k < 9; ++k) n[k] = o.vn[9 * (size_t)tri + k];
    out[0] = make_float4(n[0], n[1], n[2], 0.f);
    out[1] = make_float4(n[3], n[4], n[5], 0.f);
    out[2] = make_float4(n[6], n[7], n[8], 0.f);
}

__global__ void __launch_bounds__(256) flatten_kernel(const PrepArgs a, const int flat_lists) {
    const PrepObject o = a.objects[blockIdx.y];
    const uint32_t j = blockIdx.x * 256u + threadIdx.x;
    if (j >= o.n_tris) return;
    const size_t g = (size_t)o.tri_base + j;
    const uint32_t tri = a.order[g];
    fill_triangle(a, o, blockIdx.y, tri, a.geom + 3 * g, a.attr + 3 * g);
    fill_vnormal(o, tri, a.vnormal + 3 * g);
    if (flat_lists) {
        const size_t f = (size_t)o.flat_first + j;
        fill_triangle(a, o, blockIdx.y, j, a.flat_geom + 3 * f, a.flat_attr + 3 * f);
        fill_vnormal(o, j, a.flat_vnormal + 3 * f);
    }
}

}  // namespace

size_t build_smem_bytes(bool subtree) {
    const size_t base = (sizeof(BuildShared) + 15) & ~size_t(15);
    return subtree ? base + 4 * sizeof(float) * (size_t)kSubtreeMax : base;
}

size_t wide_finish_smem_bytes() { return 512 + 2 * (size_t)(kWideSmall / 16) * sizeof(Seg) + 4 * sizeof(float) * (size_t)kWideSmall; }

cudaError_t launch_wide_level(const PrepArgs& a, const WideArgs& w, const WorkItem* items, uint32_t n_items, uint32_t max_count,
                              uint32_t rounds, uint32_t tile_grid, int sm_count, cudaStream_t stream, int* launches) {
    if (n_items == 0) return cudaSuccess;
    const dim3 tiles((max_count + kWideTile - 1) / kWideTile, n_items);
    wide_begin_kernel<<<tiles, 256, 0, stream>>>(a, items);
    wide_init_kernel<<<1, 256, 0, stream>>>(a, w, items, n_items, rounds);
    for (uint32_t r = 0; r < rounds; ++r) {
        wide_count_kernel<<<tile_grid, 256, 0, stream>>>(a, w, r);
        wide_scatter_kernel<<<tile_grid, 256, 0, stream>>>(a, w, r);
        wide_swap_kernel<<<tile_grid, 256, 0, stream>>>(a, w, r, r + 1 == rounds ? 1u : 0u);
    }
    wide_finish_kernel<<<sm_count, kBuildThreads, wide_finish_smem_bytes(), stream>>>(a, w);
    wide_final_kernel<<<tiles, 256, 0, stream>>>(a, w, items);
    wide_children_kernel<<<(2 * n_items + 63) / 64, 64, 0, stream>>>(a, w, items, n_items);
    if (launches) *launches += 5 + 3 * (int)rounds;
    return cudaGetLastError();
}

cudaError_t prepare_kernels() {
    cudaError_t e = cudaFuncSetAttribute(build_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)build_smem_bytes(true));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(wide_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wide_finish_smem_bytes());
}

cudaError_t launch_transform(const PrepArgs& a, uint32_t max_tris, cudaStream_t stream) {
    if (a.n_objects == 0 || max_tris == 0) return cudaSuccess;
    transform_kernel<<<dim3((max_tris + 255) / 256, a.n_objects), 256, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_roots(const PrepArgs& a, cudaStream_t stream) {
    if (a.n_objects == 0) return cudaSuccess;
    roots_kernel<<<(a.n_objects + 63) / 64, 64, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_build(const PrepArgs& a, const WorkItem* items, uint32_t n_items, bool subtree, cudaStream_t stream) {
    if (n_items == 0) return cudaSuccess;
    if (subtree) build_kernel<true><<<n_items, kBuildThreads, build_smem_bytes(true), stream>>>(a, items);
    else build_kernel<false><<<n_items, kBuildThreads, build_smem_bytes(false), stream>>>(a, items);
    return cudaGetLastError();
}

cudaError_t launch_flatten(const PrepArgs& a, uint32_t max_tris, bool flat_lists, cudaStream_t stream) {
    if (a.n_objects == 0 || max_tris == 0) return cudaSuccess;
    flatten_kernel<<<dim3((max_tris + 255) / 256, a.n_objects), 256, 0, stream>>>(a, flat_lists ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace srtp

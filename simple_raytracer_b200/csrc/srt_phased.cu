// srt_phased.cu — the production render kernels (sm_100a): three stages, each a persistent grid of warps that
// pull work items from atomic queues.  See DESIGN.md §4.
//
//   P  primary   work item = 8x4 pixel tile (one warp, lane = pixel): primary ray (:505-517), closest hit
//                over every object's BVH (:405-431, :296-317, :42-75) -> hit record (t, triangle) in
//                scratch; then, with the hit still in registers, the root-box test of each of the N
//                shadow rays against every OTHER object (:321-328): lanes none of whose shadow rays
//                can enter another object's hierarchy are finished (never occluded); tiles that have
//                interested lanes are appended to a compact list.
//   S  shadow    work item = (listed tile, light), light-major: any-hit traversal (:321-342) of the
//                other objects; result = one ballot word per (tile, light).  Items are 1/N of a
//                tile's shadow work, so the heaviest tile no longer dictates the kernel's tail
//                (profiles/r01_simple_ncu_summary.txt: 40 % of SM cycles idle with whole-pixel items).
//   C  shade     work item = tile: base colour / texture (:350-361, :437-443), N x Phong (:144-200)
//                with the /5 of occluded lights (:369), tone map, gamma, quantise, sky (:391-398,
//                :447-449, :476-487).
//
// The default path launches the stages as separate, stream-ordered kernels (render_phase<PHASE, OVERLAP>; OVERLAP
// lets the shadow pass start in the primary pass's tail through programmatic dependent launch); render_dataflow
// is the same three stage bodies fused into one kernel (SRT_VARIANT_FUSED).
//
// Arithmetic: everything that reaches an output (Moeller-Trumbore, hit point, Phong, tone map) is the
// exact operation sequence of srt_device.cuh.  The slab test is evaluated with a FILTER: an
// approximate test (one FFMA per plane with a per-ray reciprocal) whose rounding error is bounded;
// when the approximate near/far values are further apart than the bound, the exact test
// (true divisions, reference compare order, :252-293) provably gives the same answer; otherwise —
// and for every ray with a zero / denormal / non-finite direction component — the exact test is run.
// The set of leaves visited is therefore exactly the reference's (S3/S5), only cheaper.
// Closest hit uses the order-independent rule min over (t, visit index) (S6), so traversal order is free.
#include <mutex>

#include "srt_kernels.cuh"
#include "srt_device.cuh"
#include "srt_shade.cuh"

namespace srtk {
using namespace srtd;

namespace {

constexpr int kThreads = 128;       // 4 warps per CTA (64 / 32 measured: +-1 %)
constexpr int kFusedMinBlocks = 8;  // the fused kernel: 64 registers
// queue entries a warp takes per atomic: consecutive tiles / consecutive lights of one tile (locality in the SM's L1)
constexpr unsigned kFull = 0xffffffffu;

// ---------------------------------------------------------------------------------------------
// Filtered slab test.
//
// Reference value per plane: t* = RN(RN(b - o) / d).  Filter value: t~ = RN(b*i + n) with
// i = rcp.approx(d) (|rel err| <= 2^-22) and n = RN(-o*i).  With q = (b-o)/d:
//   |t* - q| <= 2*2^-24 |q|,   |t~ - q| <= 1.5*2^-22 |q| + 2^-24 |n|   (plus 2^-149 absolute)
// so |t~ - t*| <= 2^-24 (8.1 |q| + |n|).  Every |q| of a ray is <= A = max_axis(B*|i| + |n|) (1 + 2^-20),
// B = largest |coordinate| of any box.  The reference's decision is  max(near*) <= min(far*)
// (its swap / merge / compare sequence reduces to that when no NaN is involved; NaN needs 0/0 or
// inf-inf, which only rays with a zero or non-finite component produce — those are `degenerate`).
// Hence with  m = 2^-24 (20 A + 3 max|n|) + 1e-37 :  far~ - near~ >= m  => reference says HIT,
// far~ - near~ < -m => reference says MISS; in between the exact test decides.
struct FRay {
    V3 o, d;          // the exact ray (reference values)
    float ix, iy, iz; // approximate reciprocals
    float nx, ny, nz; // -o * i
    float m;          // decision margin; NaN = exact test always (both comparisons below are then false)
};

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// NaN-propagating min / max (FMNMX.NAN): a NaN plane value (0 * inf) must reach the decision, where
// it makes both comparisons false and so selects the exact test.
__device__ __forceinline__ float minn(float a, float b) { float r; asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float maxn(float a, float b) { float r; asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }

// A zero direction component with a zero origin component (the centre row / column of the primary
// rays, S3) stays on the filter: i = rcp(+-0) = +-inf and n = 0 give b*i = +-inf, which IS the
// reference's b/0 (same sign rules), and 0*inf = NaN where the reference has 0/0.  The margin only
// has to cover the axes with finite values.
// The bound is a launch constant for uploaded scenes; a scene prepared on the device (srt_scene_prep) keeps it in
// device memory, because no host code ever sees the boxes.
__device__ __forceinline__ float coord_bound_of(const KArgs& a) {
    return a.coord_bound_dev ? __ldg(a.coord_bound_dev) : a.coord_bound;
}

__device__ __forceinline__ FRay make_fray(V3 o, V3 d, float coord_bound) {
    FRay r;
    r.o = o; r.d = d;
    r.ix = rcp_approx(d.x); r.iy = rcp_approx(d.y); r.iz = rcp_approx(d.z);
    r.nx = o.x == 0.f ? 0.f : -(o.x * r.ix);
    r.ny = o.y == 0.f ? 0.f : -(o.y * r.iy);
    r.nz = o.z == 0.f ? 0.f : -(o.z * r.iz);
    const float inf = __int_as_float(0x7f800000);
    const bool zx = d.x == 0.f && o.x == 0.f, zy = d.y == 0.f && o.y == 0.f, zz = d.z == 0.f && o.z == 0.f;
    const float ax = zx ? 0.f : fabsf(r.ix), ay = zy ? 0.f : fabsf(r.iy), az = zz ? 0.f : fabsf(r.iz);
    const float bx = fabsf(r.nx), by = fabsf(r.ny), bz = fabsf(r.nz);
    const float A = fmaxf(fmaxf(fmaf(coord_bound, ax, bx), fmaf(coord_bound, ay, by)), fmaf(coord_bound, az, bz));
    const float nmax = fmaxf(fmaxf(bx, by), bz);
    r.m = fmaf(A, 20.0f * 5.9604645e-8f, fmaf(nmax, 3.0f * 5.9604645e-8f, 1e-37f));
    // denormal / huge direction components, a zero component with a non-zero origin, or anything
    // that made the margin non-finite: always the exact test
    const float lo = 1e-30f, hi = 1e30f;
    const bool okx = zx ? (fabsf(r.ix) == inf) : (ax > lo && ax < hi);
    const bool oky = zy ? (fabsf(r.iy) == inf) : (ay > lo && ay < hi);
    const bool okz = zz ? (fabsf(r.iz) == inf) : (az > lo && az < hi);
    if (!(okx && oky && okz && r.m < 1e30f) || (zx && zy && zz)) r.m = __int_as_float(0x7fc00000);
    return r;
}

__device__ __forceinline__ bool slab_exact_call(V3 o, V3 d, float4 lo, float4 hi) { return slab_exact(o, d, lo, hi); }

// Returns the reference's decision for this box.  `tn_out` = approximate entry parameter (ordering only).
__device__ __forceinline__ bool slab_filtered(const FRay& r, float lox, float loy, float loz, float hix, float hiy,
                                              float hiz, float& tn_out, float& tf_out, unsigned& exact_count) {
    const float x0 = fmaf(lox, r.ix, r.nx), x1 = fmaf(hix, r.ix, r.nx);
    const float y0 = fmaf(loy, r.iy, r.ny), y1 = fmaf(hiy, r.iy, r.ny);
    const float z0 = fmaf(loz, r.iz, r.nz), z1 = fmaf(hiz, r.iz, r.nz);
    const float tn = maxn(maxn(minn(x0, x1), minn(y0, y1)), minn(z0, z1));
    const float tf = minn(minn(maxn(x0, x1), maxn(y0, y1)), maxn(z0, z1));
    tn_out = tn; tf_out = tf;
    const float diff = tf - tn;
    if (diff >= r.m) return true;
    if (diff < -r.m) return false;
    ++exact_count;
    return slab_exact_call(r.o, r.d, make_float4(lox, loy, loz, 0.f), make_float4(hix, hiy, hiz, 0.f));
}

// Both child boxes of a wide node (w0..w2 = lo/hi of the left box, lo/hi of the right box): the twelve
// plane values are independent, and ONE rarely taken branch covers the uncertain cases of both boxes.
__device__ __forceinline__ void slab_filtered_pair(const FRay& r, const float4& w0, const float4& w1, const float4& w2,
                                                   bool& hl, bool& hr, float& tnl, float& tfl, float& tnr, float& tfr,
                                                   unsigned& exact_count) {
    {
        const float x0 = fmaf(w0.x, r.ix, r.nx), x1 = fmaf(w0.w, r.ix, r.nx);
        const float y0 = fmaf(w0.y, r.iy, r.ny), y1 = fmaf(w1.x, r.iy, r.ny);
        const float z0 = fmaf(w0.z, r.iz, r.nz), z1 = fmaf(w1.y, r.iz, r.nz);
        tnl = maxn(maxn(minn(x0, x1), minn(y0, y1)), minn(z0, z1));
        tfl = minn(minn(maxn(x0, x1), maxn(y0, y1)), maxn(z0, z1));
    }
    {
        const float x0 = fmaf(w1.z, r.ix, r.nx), x1 = fmaf(w2.y, r.ix, r.nx);
        const float y0 = fmaf(w1.w, r.iy, r.ny), y1 = fmaf(w2.z, r.iy, r.ny);
        const float z0 = fmaf(w2.x, r.iz, r.nz), z1 = fmaf(w2.w, r.iz, r.nz);
        tnr = maxn(maxn(minn(x0, x1), minn(y0, y1)), minn(z0, z1));
        tfr = minn(minn(maxn(x0, x1), maxn(y0, y1)), maxn(z0, z1));
    }
    const float dl = tfl - tnl, dr = tfr - tnr;
    hl = dl >= r.m; hr = dr >= r.m;
    const bool ul = !(hl || dl < -r.m), ur = !(hr || dr < -r.m);
    if (ul || ur) {
        if (ul) { ++exact_count; hl = slab_exact_call(r.o, r.d, make_float4(w0.x, w0.y, w0.z, 0.f), make_float4(w0.w, w1.x, w1.y, 0.f)); }
        if (ur) { ++exact_count; hr = slab_exact_call(r.o, r.d, make_float4(w1.z, w1.w, w2.x, 0.f), make_float4(w2.y, w2.z, w2.w, 0.f)); }
    }
}

// The same pair test on a node of `swide`: the planes of every box are stored near-first for the octant of the ray's
// direction (slot "lo" = the plane the ray meets first), so min(x0, x1) IS x0 and max(x0, x1) IS x1 — bit for bit the
// values slab_filtered_pair computes (lo <= hi and the sign of i make lo*i+n and hi*i+n ordered, fma is monotone) — and
// the six min / max per box that sort the plane values disappear: 16 instead of 28 arithmetic instructions per node.
// Only for rays whose three direction components are non-zero with the signs of `oct`.  The exact test needs the real
// box: the planes are swapped back (rare path).
__device__ __forceinline__ void slab_sorted_pair(const FRay& r, uint32_t oct, const float4& w0, const float4& w1, const float4& w2,
                                                 bool& hl, bool& hr, float& tnl, float& tfl, float& tnr, float& tfr,
                                                 unsigned& exact_count) {
    tnl = maxn(maxn(fmaf(w0.x, r.ix, r.nx), fmaf(w0.y, r.iy, r.ny)), fmaf(w0.z, r.iz, r.nz));
    tfl = minn(minn(fmaf(w0.w, r.ix, r.nx), fmaf(w1.x, r.iy, r.ny)), fmaf(w1.y, r.iz, r.nz));
    tnr = maxn(maxn(fmaf(w1.z, r.ix, r.nx), fmaf(w1.w, r.iy, r.ny)), fmaf(w2.x, r.iz, r.nz));
    tfr = minn(minn(fmaf(w2.y, r.ix, r.nx), fmaf(w2.z, r.iy, r.ny)), fmaf(w2.w, r.iz, r.nz));
    const float dl = tfl - tnl, dr = tfr - tnr;
    hl = dl >= r.m; hr = dr >= r.m;
    const bool ul = !(hl || dl < -r.m), ur = !(hr || dr < -r.m);
    if (ul || ur) {
        const bool sx = oct & 1u, sy = oct & 2u, sz = oct & 4u;
        if (ul) {
            ++exact_count;
            hl = slab_exact_call(r.o, r.d, make_float4(sx ? w0.w : w0.x, sy ? w1.x : w0.y, sz ? w1.y : w0.z, 0.f),
                                 make_float4(sx ? w0.x : w0.w, sy ? w0.y : w1.x, sz ? w0.z : w1.y, 0.f));
        }
        if (ur) {
            ++exact_count;
            hr = slab_exact_call(r.o, r.d, make_float4(sx ? w2.y : w1.z, sy ? w2.z : w1.w, sz ? w2.w : w2.x, 0.f),
                                 make_float4(sx ? w1.z : w2.y, sy ? w1.w : w2.z, sz ? w2.x : w2.w, 0.f));
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Traversal stack + leaf list: per thread, in shared memory, one column per thread (bank = tid % 32).
// The stack (interior nodes whose own box passed) grows from the bottom of the column, the list of leaves
// waiting for their triangle tests from the top; the column has kLeafRoom entries more than the deepest stack.
constexpr int kLeafRoom = 8;   // 0 / 2 / 4 / 8 / 16 measured: 8 (the list also uses whatever the stack leaves free)
struct Stack {
    uint32_t* base;   // &smem[threadIdx.x]
    int sp;
    __device__ __forceinline__ void push(uint32_t v) { base[sp * kThreads] = v; ++sp; }
    __device__ __forceinline__ uint32_t pop() { if (sp == 0) return kRefNone; --sp; return base[sp * kThreads]; }
    __device__ __forceinline__ void put_leaf(int cap, int& nl, uint32_t ref) { base[(cap - 1 - nl) * kThreads] = ref; ++nl; }
    __device__ __forceinline__ uint32_t get_leaf(int cap, int i) const { return base[(cap - 1 - i) * kThreads]; }
};

__device__ __forceinline__ bool is_leaf(uint32_t ref) { return (ref >> 31) != 0u; }
__device__ __forceinline__ uint32_t leaf_first(uint32_t ref) { return ref & ((1u << kLeafFirstBits) - 1u); }
__device__ __forceinline__ uint32_t leaf_count(uint32_t ref) { return (ref >> kLeafFirstBits) & 15u; }

// The closest-hit traversal separates box tests from triangle tests.  In a plain while-while loop a lane that
// reaches a leaf idles until every other lane of the warp has reached one too, and lanes without a leaf idle
// through the triangle tests (ncu: 22/32 active threads).  Here a leaf child is put on the lane's leaf list the
// moment its box passes and the lane keeps descending; the triangle tests then run as ONE flat loop over
// (leaf, triangle) per lane: 26.5/32 active threads, -20 % time for the primary pass.  (The same scheme as
// speculative traversal for the any-hit rays — keep descending with up to 2-4 postponed leaves — raised their
// thread efficiency from 15.7 to 19.5 of 32 but executed 15 % more instructions and needed either the larger
// shared-memory column, halving the L1 hit rate, or registers; measured slower in both forms and not kept.)

// Closest hit over everything on the stack (refs of nodes whose own box already passed).  The reference visits
// EVERY reachable leaf (no pruning by distance, S5) and the closest hit is the order-independent min over
// (t, visit index) (S6), so collecting leaves first changes nothing but the schedule.
// SORTED: every ray of the warp points into the quadrant (sign of d.x, sign of d.y, d.z > 0) whose pre-sorted copy of the nodes
// `nodes` is (a.pwide): the pair test needs no min / max to sort the plane values (slab_sorted_pair).
template <bool SORTED>
__device__ __forceinline__ void trace_closest(const KArgs& a, const FRay& r, Stack& st, Hit& h, unsigned& exact_count,
                                              const float4* nodes, uint32_t oct) {
    const int cap = a.stack_cap;
    int nl = 0;
    uint32_t ref = st.pop();
    for (;;) {
        while (ref != kRefNone) {
            if (is_leaf(ref)) {                                   // a root that is itself a leaf
                st.put_leaf(cap, nl, ref);
                ref = st.pop();
            } else {
                const float4* w = nodes + 4 * (size_t)ref;
                const float4 w0 = ldg4(w), w1 = ldg4(w + 1), w2 = ldg4(w + 2), w3 = ldg4(w + 3);
                float tnl, tfl, tnr, tfr;
                bool hl, hr;
                if (SORTED) slab_sorted_pair(r, oct, w0, w1, w2, hl, hr, tnl, tfl, tnr, tfr, exact_count);
                else slab_filtered_pair(r, w0, w1, w2, hl, hr, tnl, tfl, tnr, tfr, exact_count);
                const uint32_t rl = __float_as_uint(w3.x), rr = __float_as_uint(w3.y);
                if (hl && is_leaf(rl)) { st.put_leaf(cap, nl, rl); hl = false; }
                if (hr && is_leaf(rr)) { st.put_leaf(cap, nl, rr); hr = false; }
                if (hl) { if (hr) st.push(rr); ref = rl; }
                else if (hr) ref = rr;
                else ref = st.pop();
            }
            if (st.sp + nl + 3 > cap) break;                     // room for one push and two leaves, or test the leaves first
        }
        uint32_t cur = 0, end = 0;
        int li = 0;
        for (;;) {
            while (cur == end && li < nl) { const uint32_t l = st.get_leaf(cap, li++); cur = leaf_first(l); end = cur + leaf_count(l); }
            if (cur == end) break;
            const float4* g = a.geom + 3 * (size_t)cur;
            const float t = moller_trumbore(r.o, r.d, ldg4(g), ldg4(g + 1), ldg4(g + 2));
            // S6: first visited wins ties == smallest visit index wins ties
            if (t != -INFINITY && (t < h.t || (t == h.t && (int32_t)cur < h.tri))) { h.t = t; h.tri = (int32_t)cur; }
            ++cur;
        }
        nl = 0;
        if (ref == kRefNone) break;
    }
}

// Any hit (shadowIntersection :321-342): true as soon as one candidate triangle reports t >= 0.
// Order is free for a boolean; children are visited nearest-entry-first, boxes wholly behind the
// origin last, so occluded rays usually stop after a few leaves.
// RANGED (srt_render_params::shadow_segment): only hits with shadow_t_lo <= t <= shadow_t_hi count; the box tests stay
// the unbounded ones (a superset).  Compiled separately so that the shipped rule pays nothing for the switch.
// SORTED: the warp's rays all point into octant a.shadow_octant — nodes come from a.swide and the pair test skips the sort.
template <bool COUNT = false, bool RANGED = false, bool SORTED = false>
__device__ __forceinline__ bool trace_any(const KArgs& a, const FRay& r, Stack& st, unsigned& exact_count, unsigned* boxes = nullptr) {
    uint32_t ref = st.pop();
    const float4* const nodes = SORTED ? a.swide : a.wide;
    while (ref != kRefNone) {
        while (ref < kRefNone) {
            const float4* w = nodes + 4 * (size_t)ref;
            const float4 w0 = ldg4(w), w1 = ldg4(w + 1), w2 = ldg4(w + 2), w3 = ldg4(w + 3);
            float tnl, tfl, tnr, tfr;
            bool hl, hr;
            if (SORTED) slab_sorted_pair(r, a.shadow_octant, w0, w1, w2, hl, hr, tnl, tfl, tnr, tfr, exact_count);
            else slab_filtered_pair(r, w0, w1, w2, hl, hr, tnl, tfl, tnr, tfr, exact_count);
            if (COUNT) *boxes += 2;
            const uint32_t rl = __float_as_uint(w3.x), rr = __float_as_uint(w3.y);
            if (hl && hr) {
                // the child the ray enters first (raw entry parameter; measured against "clamped to the origin, boxes behind it last",
                // nearer exit first and always-left: -1.7 % / +1.5 % / +9.5 % shadow time at 180 degrees)
                if (tnr < tnl) { st.push(rl); ref = rr; } else { st.push(rr); ref = rl; }
            } else if (hl) ref = rl;
            else if (hr) ref = rr;
            else ref = st.pop();
        }
        if (is_leaf(ref)) {
            const uint32_t first = leaf_first(ref), cnt = leaf_count(ref);
            bool found = false;
            for (uint32_t k = 0; k < cnt; ++k) {
                const float4* g = a.geom + 3 * (size_t)(first + k);
                const float ts = moller_trumbore_filtered(r.o, r.d, ldg4(g), ldg4(g + 1), ldg4(g + 2));   // misses cost no division
                if (ts != -INFINITY && (!RANGED || (!(ts < a.shadow_t_lo) && !(ts > a.shadow_t_hi)))) found = true;
            }
            if (found) return true;
            ref = st.pop();
        }
    }
    return false;
}

// ---------------------------------------------------------------------------------------------
// Work distribution: one warp = one item, pulled from a global counter.
__device__ __forceinline__ unsigned next_item(unsigned long long* counter, unsigned lane, unsigned chunk = 1) {
    unsigned v = 0;
    if (lane == 0) v = (unsigned)atomicAdd(counter, (unsigned long long)chunk);
    return __shfl_sync(kFull, v, 0);
}

struct Pixel { int px, py; bool valid; };

__device__ __forceinline__ Pixel pixel_of(const KArgs& a, unsigned tile, unsigned lane) {
    const int tx = (int)(tile % (unsigned)a.tiles_x), ty = (int)(tile / (unsigned)a.tiles_x);
    Pixel p;
    p.px = tx * kTileW + (int)(lane & 7);
    p.py = tile_row_to_py(a.row_begin, a.block_tile_rows, a.shard_index, a.shard_count, ty) + (int)(lane >> 3);
    p.valid = p.px < a.width && p.py < a.row_end;
    return p;
}

// ray.direction = (i, j, focal), i = px + i0, j = py + j0            :511-515 (S1)
__device__ __forceinline__ V3 primary_dir(const KArgs& a, const Pixel& p) {
    return v3((float)(p.px + a.i0), (float)(p.py + a.j0), a.focal);
}

// Root-box tests of one ray against the objects (all, or all but `skip`); passing roots are pushed
// in reverse so object 0 is traversed first.  Returns true if anything was pushed.
__device__ __forceinline__ bool push_roots(const KArgs& a, const FRay& r, Stack& st, int32_t skip, unsigned& exact_count) {
    bool any = false;
    for (int ob = (int)a.n_objects - 1; ob >= 0; --ob) {
        if (ob == skip) continue;                                             // :331
        const float4* n = a.nodes + 2 * (size_t)a.objects[ob].root;
        const float4 n0 = ldg4(n), n1 = ldg4(n + 1);
        float tn, tf;
        if (slab_filtered(r, n0.x, n0.y, n0.z, n1.x, n1.y, n1.z, tn, tf, exact_count)) {
            st.push(__ldg(a.root_ref + ob));
            any = true;
        }
    }
    return any;
}

// shadowRay of light i for the hit point of `d`, `t`                  :325-326
__device__ __forceinline__ void shadow_ray(const KArgs& a, V3 d, float t, int i, V3& so, V3& sd) {
    so = v3(mul(d.x, t), mul(d.y, t), mul(d.z, t));
    sd = v3(sub(a.lights[3 * i], so.x), sub(a.lights[3 * i + 1], so.y), sub(a.lights[3 * i + 2], so.z));
}

// Can ANY of the shadow rays of one hit point — origin `so`, directions RN(L - so) for the light positions L inside the box
// [llo, lhi] — pass the reference's slab test (:252-293) for this box?  Returns true only when NONE can, decided once for the
// whole bundle: per axis the rays' plane parameters RN(RN(b - so) / d) share their numerators (one origin) and are monotone
// in d (rounding is monotone; d = RN(L - so) is monotone in L and must not reach zero), so they lie between the values at
// the two ends of the direction interval — computed here with the very same operations.  If the hull intervals of two axes
// are disjoint, every ray has maxA < minB for that pair, which the reference's compare sequence turns into a miss whichever
// pair it is.  Anything else (a direction interval that touches zero, a NaN) answers "don't know".
__device__ __forceinline__ bool bundle_misses_box(V3 so, const float* llo, const float* lhi, float4 bmin, float4 bmax) {
    const float o[3] = {so.x, so.y, so.z}, lo[3] = {bmin.x, bmin.y, bmin.z}, hi[3] = {bmax.x, bmax.y, bmax.z};
    float tmin[3], tmax[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float d0 = sub(llo[k], o[k]), d1 = sub(lhi[k], o[k]);
        if (!(d0 > 0.f || d1 < 0.f)) return false;                          // the interval touches zero (or NaN): don't know
        const float n0 = sub(lo[k], o[k]), n1 = sub(hi[k], o[k]);
        const float q00 = dvd(n0, d0), q01 = dvd(n0, d1), q10 = dvd(n1, d0), q11 = dvd(n1, d1);
        tmin[k] = fminf(fminf(q00, q01), fminf(q10, q11));
        tmax[k] = fmaxf(fmaxf(q00, q01), fmaxf(q10, q11));
        if (!(tmin[k] <= tmax[k])) return false;                           // NaN somewhere: don't know
    }
    return tmax[0] < tmin[1] || tmax[1] < tmin[0] || tmax[0] < tmin[2] || tmax[2] < tmin[0] || tmax[1] < tmin[2] || tmax[2] < tmin[1];
}

// ---- phase 0: queue order -------------------------------------------------------------------------
// Tiles whose rays enter the root box of a large object cost 10-50x more than ground / sky tiles
// (profiles/r01_*: 100-400 k cycles against 6-20 k).  To keep them out of the tail they are queued
// first.  Heuristic only (lane = tile, its four corner rays, approximate test): a wrong class moves
// a tile in the queue, never changes a result.
constexpr uint32_t kBigObjectTris = 64;

__device__ __forceinline__ void phase_classify(const KArgs& a) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned n_tiles = (unsigned)(a.tiles_x * a.tiles_y);
    const unsigned warps = gridDim.x * (kThreads / 32), gw = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    for (unsigned base = gw * 32; base < n_tiles; base += warps * 32) {
        const unsigned tile = base + lane;
        bool heavy = false;
        if (tile < n_tiles) {
            const int tx = (int)(tile % (unsigned)a.tiles_x), ty = (int)(tile / (unsigned)a.tiles_x);
            for (int c = 0; c < 4 && !heavy; ++c) {
                const int px = tx * kTileW + ((c & 1) ? kTileW - 1 : 0);
                const int py = tile_row_to_py(a.row_begin, a.block_tile_rows, a.shard_index, a.shard_count, ty) + ((c & 2) ? kTileH - 1 : 0);
                const V3 d = v3((float)(px + a.i0), (float)(py + a.j0), a.focal);
                const float ix = rcp_approx(d.x), iy = rcp_approx(d.y), iz = rcp_approx(d.z);
                for (uint32_t ob = 0; ob < a.n_objects && !heavy; ++ob) {
                    if (a.objects[ob].tri_count < kBigObjectTris) continue;
                    const float4* n = a.nodes + 2 * (size_t)a.objects[ob].root;
                    const float4 n0 = ldg4(n), n1 = ldg4(n + 1);
                    const float x0 = n0.x * ix, x1 = n1.x * ix, y0 = n0.y * iy, y1 = n1.y * iy, z0 = n0.z * iz, z1 = n1.z * iz;
                    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fminf(z0, z1));
                    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
                    heavy = tn <= tf;
                }
            }
        }
        const unsigned hb = __ballot_sync(kFull, heavy);
        const unsigned lb = __ballot_sync(kFull, tile < n_tiles && !heavy);
        unsigned hbase = 0, lbase = 0;
        if (lane == 0) {
            if (hb) hbase = (unsigned)atomicAdd(&a.counters[CNT_HEAVY], (unsigned long long)__popc(hb));
            if (lb) lbase = (unsigned)atomicAdd(&a.counters[CNT_LIGHT], (unsigned long long)__popc(lb));
        }
        hbase = __shfl_sync(kFull, hbase, 0);
        lbase = __shfl_sync(kFull, lbase, 0);
        const unsigned below = (1u << lane) - 1u;
        if (heavy) a.tile_order[hbase + __popc(hb & below)] = tile;                       // front of the queue
        else if (tile < n_tiles) a.tile_order[n_tiles - 1 - (lbase + __popc(lb & below))] = tile;   // back of the queue
    }
}

// ---- per-item bodies (shared by the dataflow kernel and the per-phase profiling kernels) ------------
// Reads of data another warp wrote during this launch (hit records, interest words, shadow masks)
// go through L2 (__ldcg): L1 is not coherent between SMs.

// Closest hit of one tile (lane = pixel) + the hit record + "can any shadow ray of this pixel reach
// another object's root box" (the root test of :328).  Returns the ballot of interested lanes.
template <bool TL = false>
__device__ __forceinline__ unsigned primary_tile(const KArgs& a, Stack& st, unsigned tile, unsigned lane, Pixel& p, V3& d,
                                                 Hit& h, unsigned& hit_ballot, unsigned& exact_count) {
    long long tl_c0 = 0; unsigned long long tl_ns = 0;
    if (TL && a.debug_timeline == 1) { tl_c0 = clock64(); asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tl_ns)); }
    p = pixel_of(a, tile, lane);
    d = primary_dir(a, p);
    const FRay r = make_fray(v3(0.f, 0.f, 0.f), d, coord_bound_of(a));
    h.t = INFINITY; h.tri = -1; h.obj = -1;
    st.sp = 0;
    if (p.valid) push_roots(a, r, st, -1, exact_count);                       // :409 (S2 order via the S6 rule)
    // primary rays of a tile share the signs of d.x and d.y unless the tile straddles the centre column / row (d.z = focal)
    const uint32_t quad = (d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u);
    const uint32_t quad0 = __shfl_sync(kFull, quad, 0);
    const bool fits = !p.valid || (quad == quad0 && d.x != 0.f && d.y != 0.f);
    if (a.pwide != nullptr && a.focal > 0.f && __all_sync(kFull, fits) && __shfl_sync(kFull, p.valid ? 1 : 0, 0))
        trace_closest<true>(a, r, st, h, exact_count, a.pwide + (size_t)quad0 * a.pwide_stride, quad0);
    else
        trace_closest<false>(a, r, st, h, exact_count, a.wide, 0u);
    bool interested = false;
    if (p.valid) {
        const size_t at = (size_t)p.py * a.width + p.px;
        if (h.tri >= 0) h.obj = __float_as_int(ldg4(a.geom + 3 * (size_t)h.tri).w);
        a.hit_t[at] = h.t;
        a.hit_tri[at] = h.tri;
        if (a.obj_id) a.obj_id[at] = h.obj;
        if (a.tri_id) a.tri_id[at] = h.tri;
        if (a.t && a.debug_timeline != 2) a.t[at] = h.t;
        if (h.tri >= 0 && a.n_objects >= 2) {
            // all light samples at once first: most pixels of most views cannot reach another object's box with ANY of their
            // shadow rays, and saying so per light costs a ray set-up and a root test each (16 x ~90 instructions per pixel)
            bool maybe = a.light_count <= 2;
            if (!maybe) {
                const V3 so = v3(mul(d.x, h.t), mul(d.y, h.t), mul(d.z, h.t));                 // shadow_ray's origin (:326)
                for (int ob = 0; ob < (int)a.n_objects && !maybe; ++ob) {
                    if (ob == h.obj) continue;
                    const float4* n = a.nodes + 2 * (size_t)a.objects[ob].root;
                    maybe = !bundle_misses_box(so, a.light_lo, a.light_hi, ldg4(n), ldg4(n + 1));
                }
            }
            for (int i = 0; i < a.light_count && maybe && !interested; ++i) {
                V3 so, sd;
                shadow_ray(a, d, h.t, i, so, sd);
                const FRay s = make_fray(so, sd, coord_bound_of(a));
                st.sp = 0;
                interested = push_roots(a, s, st, h.obj, exact_count);
            }
            st.sp = 0;
        }
    }
    hit_ballot = __ballot_sync(kFull, p.valid && h.tri >= 0);
    const unsigned int_ballot = __ballot_sync(kFull, interested);
    if (TL && a.debug_timeline == 1 && p.valid) {
        __syncwarp();
        const size_t at = (size_t)p.py * a.width + p.px;
        unsigned smid; asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
        if (a.t) a.t[at] = (float)(clock64() - tl_c0);
        if (a.obj_id) a.obj_id[at] = (int32_t)smid;
        if (a.tri_id) a.tri_id[at] = (int32_t)(tl_ns & 0x7fffffffull);
    }
    return int_ballot;
}

// Occlusion of one light for the interested lanes of one tile -> one ballot word (:321-342).
template <bool COUNT = false>
__device__ __forceinline__ void shadow_item(const KArgs& a, Stack& st, unsigned tile, int light, unsigned lane,
                                            unsigned& exact_count, unsigned* box_occluded = nullptr, unsigned* box_free = nullptr) {
    const unsigned interest = __ldcg(a.tile_interest + tile);
    const Pixel p = pixel_of(a, tile, lane);
    st.sp = 0;
    V3 so = v3(0.f, 0.f, 0.f), sd = v3(1.f, 1.f, 1.f);
    int32_t obj = -1;
    const bool mine = (interest >> lane) & 1u;
    if (mine) {
        const size_t at = (size_t)p.py * a.width + p.px;
        const float t = __ldcg(a.hit_t + at);
        const int32_t tri = __ldcg(a.hit_tri + at);
        obj = __float_as_int(ldg4(a.geom + 3 * (size_t)tri).w);
        shadow_ray(a, primary_dir(a, p), t, light, so, sd);
    }
    const FRay s = make_fray(so, sd, coord_bound_of(a));
    if (mine) push_roots(a, s, st, obj, exact_count);
    unsigned boxes = 0;
    const bool ranged = a.shadow_t_hi != INFINITY;     // uniform over the launch
    // fast path: every traversing lane's ray has three non-zero direction components with the launch's octant signs
    const uint32_t my_oct = (sd.x < 0.f ? 1u : 0u) | (sd.y < 0.f ? 2u : 0u) | (sd.z < 0.f ? 4u : 0u);
    const bool fits = !mine || (my_oct == a.shadow_octant && sd.x != 0.f && sd.y != 0.f && sd.z != 0.f);
    const bool sorted = !COUNT && !ranged && a.swide != nullptr && __all_sync(kFull, fits);
    const bool occluded = sorted ? trace_any<false, false, true>(a, s, st, exact_count, &boxes)
                                 : (ranged ? trace_any<COUNT, true>(a, s, st, exact_count, &boxes) : trace_any<COUNT, false>(a, s, st, exact_count, &boxes));
    if (COUNT) { if (occluded) *box_occluded += boxes; else *box_free += boxes; }
    const unsigned ballot = __ballot_sync(kFull, mine && occluded);
    if (lane == 0) a.shadow_mask[(size_t)tile * a.light_count + light] = ballot;
}

// The 8-bit pixel of a shaded colour: tone map, gamma, truncating quantisation (:391-398, :447-449, :470-472) and the sky
// rule (:518, :476-487), packed as r | g << 8 | b << 16.
__device__ __forceinline__ uint32_t quantise_pixel(const KArgs& a, bool hit, V3 color) {
    int r = 0, g = 0, b = 0;
    if (hit) {
        r = tone_quantise(color.x, a.tone_k, a.gamma) & 0xff;
        g = tone_quantise(color.y, a.tone_k, a.gamma) & 0xff;
        b = tone_quantise(color.z, a.tone_k, a.gamma) & 0xff;
    }
    if ((r | g | b) == 0 && !a.no_sky_fill) { r = a.sky[0]; g = a.sky[1]; b = a.sky[2]; }
    return (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16);
}

// RGB8 store of a whole 8x4 tile (all 32 lanes call it; lane = pixel, c = its packed colour).  A tile row is 24
// contiguous bytes = six 32-bit words; word w of a row takes its bytes from pixels (w/3)*4 + w%3 and the next one, so two
// shuffles and a funnel shift build it and lanes 0..5 of every row store one word each: 24 STG.32 per tile instead of 96
// STG.U8 — a quarter of the store instructions, and whole 32-byte sectors per row for frames that live in a peer GPU's
// memory (srt_mgpu_render_frame writes them over NVLink).  Needs 4-byte aligned rows (width % 4 == 0, aligned base) and a
// tile that lies inside the frame; anything else takes the byte path.
__device__ __forceinline__ void store_tile_rgb(const KArgs& a, const Pixel& p, uint32_t c, unsigned lane) {
    if (!a.rgb) return;
    const bool aligned = ((a.width & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.rgb) & 3u) == 0);
    const bool full = __all_sync(kFull, p.valid);
    if (aligned && full) {
        const unsigned col = lane & 7u, wl = col % 3u, base = (col / 3u) * 4u + (lane & ~7u);
        const uint32_t ca = __shfl_sync(kFull, c, (int)(base + wl));
        const uint32_t cb = __shfl_sync(kFull, c, (int)((base + wl + 1u) & 31u));
        if (col < 6u) {
            const uint32_t word = __funnelshift_r(ca | (cb << 24), cb >> 8, 8u * wl);   // bytes wl..2 of ca, then the low bytes of cb
            uint8_t* row = a.rgb + ((size_t)p.py * a.width + (size_t)(p.px - (int)col)) * 3;
            reinterpret_cast<uint32_t*>(row)[col] = word;
        }
        return;
    }
    if (p.valid) {
        uint8_t* q = a.rgb + ((size_t)p.py * a.width + p.px) * 3;
        q[0] = (uint8_t)c; q[1] = (uint8_t)(c >> 8); q[2] = (uint8_t)(c >> 16);
    }
}

// Shading of one pixel (:350-401, :437-449, :476-487).  `mine`: this lane has shadow masks to read.  Called by all 32
// lanes of the tile's warp (the store is a warp-wide operation).
__device__ __forceinline__ void shade_pixel(const KArgs& a, const Pixel& p, const Hit& h, V3 d, bool mine, unsigned tile,
                                            unsigned lane) {
    V3 color = v3(0.f, 0.f, 0.f);
    const bool hit = p.valid && h.tri >= 0;
    if (hit) {
        const V3 lc = v3(a.light_color[0], a.light_color[1], a.light_color[2]);
        const PhongPixel pp = prepare_pixel(a, h, d, false);
        const uint32_t* masks = a.shadow_mask + (size_t)tile * a.light_count;
        for (int i = 0; i < a.light_count; ++i) {                             // :366
            const V3 L = v3(a.lights[3 * i], a.lights[3 * i + 1], a.lights[3 * i + 2]);
            const bool shadowed = mine && ((__ldcg(masks + i) >> lane) & 1u); // :367
            V3 c = phong(pp, L, lc);                                          // :368
            if (shadowed) { c.x = dvd(c.x, a.shadow_div); c.y = dvd(c.y, a.shadow_div); c.z = dvd(c.z, a.shadow_div); }  // :369
            color.x = add(color.x, c.x); color.y = add(color.y, c.y); color.z = add(color.z, c.z);   // :370
        }
        if (a.normalise_lights) color = average_lights(color, a.light_count);
    }
    store_tile_rgb(a, p, quantise_pixel(a, hit, color), lane);   // ids / t planes were written with the hit record
}

// Shading of a tile from its stored hit records (after its shadow items).
__device__ __forceinline__ void shade_tile_from_memory(const KArgs& a, unsigned tile, unsigned lane) {
    const Pixel p = pixel_of(a, tile, lane);
    Hit h; h.t = INFINITY; h.tri = -1; h.obj = -1;
    if (p.valid) {
        const size_t at = (size_t)p.py * a.width + p.px;
        h.t = __ldcg(a.hit_t + at);
        h.tri = __ldcg(a.hit_tri + at);
        if (h.tri >= 0) h.obj = __float_as_int(ldg4(a.geom + 3 * (size_t)h.tri).w);
    }
    const bool mine = (__ldcg(a.tile_interest + tile) >> lane) & 1u;
    shade_pixel(a, p, h, primary_dir(a, p), mine, tile, lane);
}

// ---- the three stages as separate passes (the production pipeline) -------------------------------------
// OVERLAP (SRT_VARIANT_OVERLAP): the shadow pass is launched with programmatic stream serialization,
// so its CTAs become resident as soon as CTAs of the primary pass retire and fill the SMs that pass's long
// tail tiles (bunny silhouette: 100-400 k cycles each) would leave idle.  The passes are then ordered by
// data, not by the launch boundary: a primary tile is PUBLISHED (hit records + interest word, fence,
// tagged list entry, fence, "tiles decided" count) and a shadow item waits for its list entry.  The
// primary CTAs are all resident before the first shadow CTA exists (the grid is the resident capacity and
// every CTA triggers the dependent launch first thing), and they never wait, so the waits cannot deadlock.
// Measured: +2 % on the orbit average, -5 % on C3 itself (short shadow pass, the publication protocol for nothing),
// so the default keeps the launches strictly ordered.
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}

template <bool OVERLAP>
__device__ __forceinline__ void phase_primary(const KArgs& a, uint32_t* smem, unsigned& exact_count) {
    if (OVERLAP) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const unsigned lane = threadIdx.x & 31;
    const unsigned n_tiles = (unsigned)(a.tiles_x * a.tiles_y);
    Stack st; st.base = smem + threadIdx.x; st.sp = 0;
    unsigned hits = 0, decided = 0;
    for (;;) {
        const unsigned q = next_item(&a.counters[CNT_TILE], lane);
        if (q >= n_tiles) break;
        const unsigned tile = a.tile_order[q];
        Pixel p; V3 d; Hit h; unsigned hit_ballot;
        const unsigned int_ballot = primary_tile(a, st, tile, lane, p, d, h, hit_ballot, exact_count);
        ++decided;
        if (OVERLAP && int_ballot) __syncwarp();                 // the lanes' hit records happen-before lane 0's fence
        if (lane == 0) {
            hits += __popc(hit_ballot);
            a.tile_interest[tile] = int_ballot;
            if (int_ballot) {
                const unsigned slot = (unsigned)atomicAdd(&a.counters[CNT_LIST], 1ull);
                const unsigned long long entry = ((unsigned long long)a.frame_seq << 32) | tile;
                if (OVERLAP) {
                    __threadfence();                             // (cumulative) hit records + interest word before the entry
                    *reinterpret_cast<volatile unsigned long long*>(a.tile_list + slot) = entry;
                } else {
                    a.tile_list[slot] = entry;
                }
            }
        }
    }
    if (OVERLAP) {                                               // this warp's tiles are decided: entries before the count
        __syncwarp();
        if (lane == 0) { __threadfence(); atomicAdd(&a.counters[CNT_PDONE], (unsigned long long)decided); }
    }
    if (lane == 0 && hits) atomicAdd(&a.counters[CNT_HIT], (unsigned long long)hits);
}

template <bool OVERLAP>
__device__ __forceinline__ void phase_shadow(const KArgs& a, uint32_t* smem, unsigned& exact_count) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned N = (unsigned)a.light_count;
    Stack st; st.base = smem + threadIdx.x; st.sp = 0;
    if (!OVERLAP) {
        const unsigned n_items = (unsigned)a.counters[CNT_LIST] * N;
        for (;;) {
            const unsigned item = next_item(&a.counters[CNT_SHADOW], lane);
            if (item >= n_items) break;
            const unsigned slot = item / N;
            shadow_item(a, st, (unsigned)(a.tile_list[slot] & 0xffffffffull), (int)(item - slot * N), lane, exact_count);
        }
        return;
    }
    const unsigned n_tiles = (unsigned)(a.tiles_x * a.tiles_y);
    const unsigned long long seq = a.frame_seq;
    bool all_published = false;                                    // warp-uniform: the primary pass has finished
    unsigned n_items = 0;
    for (;;) {
        const unsigned item = next_item(&a.counters[CNT_SHADOW], lane);
        const unsigned slot = item / N;
        if (!all_published) {
            unsigned long long done = 0;
            if (lane == 0) done = ld_volatile_u64(a.counters + CNT_PDONE);
            if (__shfl_sync(kFull, done, 0) >= n_tiles) {          // from here on: plain loads, no waiting, no fences
                __threadfence();
                all_published = true;
                n_items = (unsigned)ld_volatile_u64(a.counters + CNT_LIST) * N;
            }
        }
        if (all_published) {
            if (item >= n_items) break;
            shadow_item(a, st, (unsigned)(__ldcg(a.tile_list + slot) & 0xffffffffull), (int)(item - slot * N), lane, exact_count);
            continue;
        }
        // the primary pass is still running: wait for this item's list entry
        unsigned long long e = 0;
        if (lane == 0 && slot < n_tiles) {
            unsigned ns = 250;
            for (unsigned spins = 0;; ++spins) {
                e = ld_volatile_u64(a.tile_list + slot);
                if ((e >> 32) == seq) break;
                if (spins > (1u << 20)) { atomicAdd(&a.counters[CNT_TIMEOUT], 1ull); break; }   // > 1 s: give up loudly
                if (ld_volatile_u64(a.counters + CNT_PDONE) >= n_tiles) {          // every tile decided: final answer
                    __threadfence();
                    e = ld_volatile_u64(a.tile_list + slot);
                    break;
                }
                __nanosleep(ns);                                   // back off: thousands of waiting warps poll two lines
                if (ns < 2000) ns *= 2;
            }
        }
        e = __shfl_sync(kFull, e, 0);
        if ((e >> 32) != seq) break;                               // beyond the last published tile: no more work
        __threadfence();
        shadow_item(a, st, (unsigned)(e & 0xffffffffull), (int)(item - slot * N), lane, exact_count);
    }
}

__device__ __forceinline__ void phase_shade(const KArgs& a) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned n_tiles = (unsigned)(a.tiles_x * a.tiles_y);
    for (;;) {
        const unsigned tile = next_item(&a.counters[CNT_SHADE], lane);
        if (tile >= n_tiles) break;
        shade_tile_from_memory(a, tile, lane);
    }
}

// ---- the production kernel: one persistent grid, dataflow between the stages ------------------------
// Every warp first pulls primary tiles (expected-heavy first).  A tile none of whose lanes can be
// shadowed is shaded on the spot, from registers.  Otherwise the tile is PUBLISHED: its hit records
// and interest word are made visible (fence), then its id tagged with this frame's sequence number
// goes into the next slot of tile_list.  When the tile queue is exhausted a warp turns to shadow
// items (slot, light); an item whose slot is not published yet is waited for — the publishers are
// running warps that never wait themselves, so this cannot deadlock — and the warp that completes a
// tile's last shadow item shades that tile.  No grid-wide barrier: the long primary tiles at the end
// of the queue (bunny silhouette: 400 us against a 300 us phase) overlap with shadow and shading
// work instead of idling the machine (profiles/r01_phased_fused: 35 % of warp samples sat in barriers).

// Out-of-line shading for the fused kernel: the shading code needs 72 registers, the traversal loops 56; inlined
// into one kernel under the 64-register cap the traversal loops would carry the spills.
__device__ __noinline__ void shade_pixel_call(const KArgs& a, Pixel p, Hit h, V3 d, unsigned tile, unsigned lane) {
    shade_pixel(a, p, h, d, false, tile, lane);
}
__device__ __noinline__ void shade_tile_call(const KArgs& a, unsigned tile, unsigned lane) { shade_tile_from_memory(a, tile, lane); }

// COUNT / TL: the instrumented builds (variants 104 / 102, 103); the product kernel carries none of their state.
template <bool COUNT, bool TL>
__global__ void __launch_bounds__(kThreads, kFusedMinBlocks) render_dataflow(const __grid_constant__ KArgs a) {
    extern __shared__ uint32_t smem[];
    const unsigned lane = threadIdx.x & 31;
    const unsigned n_tiles = (unsigned)(a.tiles_x * a.tiles_y);
    const unsigned N = (unsigned)a.light_count;
    const unsigned long long seq = a.frame_seq;
    Stack st; st.base = smem + threadIdx.x; st.sp = 0;
    unsigned hits = 0, exact_count = 0;
    // debug_timeline == 2: per-warp timestamps (start, primary queue empty, end) + item counts in the t plane
    unsigned long long ts0 = 0, ts1 = 0, ts2 = 0; unsigned n_p = 0, n_s = 0, n_c = 0;
    unsigned box_occluded = 0, box_free = 0;   // COUNT: shadow-stage box tests by rays that ended occluded / unoccluded
    if (TL && a.debug_timeline == 2) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ts0));

    for (;;) {
      const unsigned q = next_item(&a.counters[CNT_TILE], lane);
      if (q >= n_tiles) break;
      {
        ++n_p;
        const unsigned tile = a.tile_order[q];
        Pixel p; V3 d; Hit h; unsigned hit_ballot;
        const unsigned int_ballot = primary_tile<TL>(a, st, tile, lane, p, d, h, hit_ballot, exact_count);
        if (int_ballot == 0) {
            shade_pixel_call(a, p, h, d, tile, lane);
        } else {
            __threadfence();                                   // this lane's hit record before the publication
            __syncwarp();
        }
        if (lane == 0) {
            hits += __popc(hit_ballot);
            if (int_ballot) {
                a.tile_interest[tile] = int_ballot;
                a.tile_done[tile] = 0u;
                const unsigned slot = (unsigned)atomicAdd(&a.counters[CNT_LIST], 1ull);
                __threadfence();
                *reinterpret_cast<volatile unsigned long long*>(a.tile_list + slot) = (seq << 32) | tile;
                __threadfence();
            }
            atomicAdd(&a.counters[CNT_PDONE], 1ull);
        }
      }
    }

    if (TL && a.debug_timeline == 2) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ts1));
    if (a.n_objects >= 2 && N > 0) {
        bool finished = false;
        while (!finished) {
          const unsigned s = next_item(&a.counters[CNT_SHADOW], lane);
          {
            const unsigned slot = s / N;
            const int light = (int)(s - slot * N);
            unsigned long long e = 0;
            if (lane == 0 && slot < n_tiles) {
                for (unsigned spins = 0;; ++spins) {
                    e = ld_volatile_u64(a.tile_list + slot);
                    if ((e >> 32) == seq) break;
                    if (spins > (1u << 23)) { atomicAdd(&a.counters[CNT_TIMEOUT], 1ull); break; }   // > 1 s: give up loudly
                    if (ld_volatile_u64(a.counters + CNT_PDONE) >= n_tiles) {      // every tile decided: final answer
                        __threadfence();
                        e = ld_volatile_u64(a.tile_list + slot);
                        break;
                    }
                    __nanosleep(200);
                }
            }
            e = __shfl_sync(kFull, e, 0);
            if ((e >> 32) != seq) { finished = true; break; }  // beyond the last published tile: no more work
            __threadfence();
            const unsigned tile = (unsigned)(e & 0xffffffffull);
            shadow_item<COUNT>(a, st, tile, light, lane, exact_count, &box_occluded, &box_free);
            ++n_s;
            unsigned last = 0;
            if (lane == 0) {
                __threadfence();                               // the mask before the completion count
                last = atomicAdd(a.tile_done + tile, 1u) == N - 1 ? 1u : 0u;
            }
            last = __shfl_sync(kFull, last, 0);
            if (last) {
                __threadfence();
                shade_tile_call(a, tile, lane);
                ++n_c;
            }
          }
        }
    }

    if (TL && a.debug_timeline == 2 && a.t) {
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ts2));
        __syncwarp();
        if (lane == 0) {
            unsigned long long* out = reinterpret_cast<unsigned long long*>(a.t) + 8 * (size_t)(blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5));
            out[0] = ts0; out[1] = ts1; out[2] = ts2; out[3] = n_p; out[4] = n_s; out[5] = n_c;
        }
    }
    if (COUNT) {
        for (int s = 16; s > 0; s >>= 1) { box_occluded += __shfl_xor_sync(kFull, box_occluded, s); box_free += __shfl_xor_sync(kFull, box_free, s); }
        if (lane == 0) { atomicAdd(&a.counters[CNT_NODE], (unsigned long long)box_occluded); atomicAdd(&a.counters[CNT_TRI], (unsigned long long)box_free); }
    }
    for (int s = 16; s > 0; s >>= 1) exact_count += __shfl_xor_sync(kFull, exact_count, s);
    if (lane == 0) {
        if (hits) atomicAdd(&a.counters[CNT_HIT], (unsigned long long)hits);
        if (exact_count) atomicAdd(&a.counters[CNT_EXACT], (unsigned long long)exact_count);
    }
}

__global__ void __launch_bounds__(kThreads) classify_kernel(const __grid_constant__ KArgs a) { phase_classify(a); }

// The stages as separate launches.
template <int PHASE, bool OVERLAP>
__global__ void __launch_bounds__(kThreads) render_phase(const __grid_constant__ KArgs a) {
    extern __shared__ uint32_t smem[];
    unsigned exact_count = 0;
    if (PHASE == 0) phase_primary<OVERLAP>(a, smem, exact_count);
    if (PHASE == 1) phase_shadow<OVERLAP>(a, smem, exact_count);
    if (PHASE == 2) phase_shade(a);
    for (int s = 16; s > 0; s >>= 1) exact_count += __shfl_xor_sync(kFull, exact_count, s);
    if ((threadIdx.x & 31) == 0 && exact_count) atomicAdd(&a.counters[CNT_EXACT], (unsigned long long)exact_count);
}

// wide[4*i .. 4*i+3] for interior node i: {Lmin.xyz, Lmax.x} {Lmax.yz, Rmin.xy} {Rmin.z, Rmax.xyz} {Lref, Rref, 0, 0}
__device__ __forceinline__ uint32_t ref_of(const float4* nodes, uint32_t i) {
    const uint32_t na = __float_as_uint(nodes[2 * (size_t)i].w), nb = __float_as_uint(nodes[2 * (size_t)i + 1].w);
    if (nb & SRT_LEAF_BIT) return SRT_LEAF_BIT | ((nb & 15u) << kLeafFirstBits) | (na & ((1u << kLeafFirstBits) - 1u));
    return i;
}

__global__ void build_wide_kernel(const float4* nodes, const srt_object* objects, uint32_t n_nodes, uint32_t n_objects,
                                  float4* wide, uint32_t* root_ref) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_objects) root_ref[i] = ref_of(nodes, objects[i].root);
    if (i >= n_nodes) return;
    const float4 n0 = nodes[2 * (size_t)i], n1 = nodes[2 * (size_t)i + 1];
    if (__float_as_uint(n1.w) & SRT_LEAF_BIT) return;
    const uint32_t l = i + 1, r = __float_as_uint(n0.w);
    const float4 l0 = nodes[2 * (size_t)l], l1 = nodes[2 * (size_t)l + 1];
    const float4 r0 = nodes[2 * (size_t)r], r1 = nodes[2 * (size_t)r + 1];
    float4* w = wide + 4 * (size_t)i;
    w[0] = make_float4(l0.x, l0.y, l0.z, l1.x);
    w[1] = make_float4(l1.y, l1.z, r0.x, r0.y);
    w[2] = make_float4(r0.z, r1.x, r1.y, r1.z);
    w[3] = make_float4(__uint_as_float(ref_of(nodes, l)), __uint_as_float(ref_of(nodes, r)), 0.f, 0.f);
}

// swide[4*i..]: node i of `wide` with the planes of both child boxes stored near-first for direction octant `oct`.
__global__ void build_shadow_wide_kernel(const float4* __restrict__ wide, uint32_t n_nodes, uint32_t oct, float4* __restrict__ swide) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const float4 w0 = wide[4 * (size_t)i], w1 = wide[4 * (size_t)i + 1], w2 = wide[4 * (size_t)i + 2], w3 = wide[4 * (size_t)i + 3];
    const bool sx = oct & 1u, sy = oct & 2u, sz = oct & 4u;
    // left box: lo = (w0.x, w0.y, w0.z), hi = (w0.w, w1.x, w1.y); right box: lo = (w1.z, w1.w, w2.x), hi = (w2.y, w2.z, w2.w)
    float4 o0, o1, o2;
    o0.x = sx ? w0.w : w0.x; o0.y = sy ? w1.x : w0.y; o0.z = sz ? w1.y : w0.z;
    o0.w = sx ? w0.x : w0.w; o1.x = sy ? w0.y : w1.x; o1.y = sz ? w0.z : w1.y;
    o1.z = sx ? w2.y : w1.z; o1.w = sy ? w2.z : w1.w; o2.x = sz ? w2.w : w2.x;
    o2.y = sx ? w1.z : w2.y; o2.z = sy ? w1.w : w2.z; o2.w = sz ? w2.x : w2.w;
    float4* o = swide + 4 * (size_t)i;
    o[0] = o0; o[1] = o1; o[2] = o2; o[3] = w3;
}

}  // namespace

cudaError_t launch_build_shadow_wide(const float4* wide, uint32_t n_nodes, uint32_t octant, float4* swide, cudaStream_t stream) {
    if (n_nodes == 0) return cudaSuccess;
    build_shadow_wide_kernel<<<(n_nodes + 255) / 256, 256, 0, stream>>>(wide, n_nodes, octant, swide);
    return cudaGetLastError();
}

cudaError_t launch_build_wide(const float4* nodes, const srt_object* objects, uint32_t n_nodes, uint32_t n_objects,
                              float4* wide, uint32_t* root_ref, cudaStream_t stream) {
    const uint32_t n = n_nodes > n_objects ? n_nodes : n_objects;
    if (n == 0) return cudaSuccess;
    build_wide_kernel<<<(n + 255) / 256, 256, 0, stream>>>(nodes, objects, n_nodes, n_objects, wide, root_ref);
    return cudaGetLastError();
}

int phased_stack_cap(int max_depth, int n_objects) { return max_depth + n_objects + 2 + kLeafRoom; }
size_t phased_smem_bytes(int stack_cap) { return (size_t)stack_cap * kThreads * sizeof(uint32_t); }

// ---------------------------------------------------------------------------------------------
// srt_eval_leaf: one thread per case, the SAME device functions the render kernels call.
namespace {
__global__ void eval_leaf_kernel(int op, const float* __restrict__ in, float* __restrict__ out, size_t n, int wi, int wo) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* x = in + i * wi;
    float* y = out + i * wo;
    switch (op) {
    case SRT_LEAF_POWF: y[0] = srt_powf(x[0], x[1]); break;
    case SRT_LEAF_SLAB: {
        const V3 o = v3(x[0], x[1], x[2]), d = v3(x[3], x[4], x[5]);
        float bound = 0.f;
        for (int k = 6; k < 12; ++k) bound = fmaxf(bound, fabsf(x[k]));
        const FRay r = make_fray(o, d, bound);
        float tn, tf; unsigned exact_count = 0;
        const bool f = slab_filtered(r, x[6], x[7], x[8], x[9], x[10], x[11], tn, tf, exact_count);
        y[0] = slab_exact(o, d, make_float4(x[6], x[7], x[8], 0.f), make_float4(x[9], x[10], x[11], 0.f)) ? 1.f : 0.f;
        y[1] = f ? 1.f : 0.f;
        y[2] = (float)exact_count;
        break;
    }
    case SRT_LEAF_TRIANGLE:
        y[0] = moller_trumbore(v3(x[0], x[1], x[2]), v3(x[3], x[4], x[5]), make_float4(x[6], x[7], x[8], 0.f),
                               make_float4(x[9], x[10], x[11], 0.f), make_float4(x[12], x[13], x[14], 0.f));
        break;
    case SRT_LEAF_TRIANGLE_FILTERED: {
        const V3 o = v3(x[0], x[1], x[2]), d = v3(x[3], x[4], x[5]);
        const float4 g0 = make_float4(x[6], x[7], x[8], 0.f), g1 = make_float4(x[9], x[10], x[11], 0.f), g2 = make_float4(x[12], x[13], x[14], 0.f);
        y[0] = moller_trumbore_filtered(o, d, g0, g1, g2);
        y[1] = moller_trumbore(o, d, g0, g1, g2);
        y[2] = moller_trumbore_took_tail(o, d, g0, g1, g2) ? 1.f : 0.f;
        break;
    }
    case SRT_LEAF_PHONG: {
        PhongPixel px;
        const V3 d = v3(x[3], x[4], x[5]);
        const float t = x[18];
        px.n = v3(x[0], x[1], x[2]);
        px.P = v3(add(0.f, mul(t, d.x)), add(0.f, mul(t, d.y)), add(0.f, mul(t, d.z)));   // :156 with origin (0,0,0)
        px.v = normalize3(v3(-d.x, -d.y, -d.z));                                           // :190
        px.oc = v3(x[12], x[13], x[14]);
        px.ambientStrength = x[15]; px.specularStrength = x[16]; px.shininess = x[17];
        const V3 c = phong(px, v3(x[6], x[7], x[8]), v3(x[9], x[10], x[11]));
        y[0] = c.x; y[1] = c.y; y[2] = c.z;
        break;
    }
    case SRT_LEAF_TONE: y[0] = (float)tone_quantise(x[0], x[1], x[2]); break;
    case SRT_LEAF_SMOOTH_NORMAL: {
        const V3 n = smooth_normal(v3(x[0], x[1], x[2]), v3(x[3], x[4], x[5]), v3(x[6], x[7], x[8]), make_float4(x[9], x[10], x[11], 0.f),
                                   make_float4(x[12], x[13], x[14], 0.f), make_float4(x[15], x[16], x[17], 0.f), v3(x[18], x[19], x[20]));
        y[0] = n.x; y[1] = n.y; y[2] = n.z;
        break;
    }
    default: break;
    }
}
}  // namespace

int leaf_in_width(int op) {
    switch (op) {
    case SRT_LEAF_POWF: return 2;
    case SRT_LEAF_SLAB: return 12;
    case SRT_LEAF_TRIANGLE: case SRT_LEAF_TRIANGLE_FILTERED: return 15;
    case SRT_LEAF_PHONG: return 19;
    case SRT_LEAF_TONE: return 3;
    case SRT_LEAF_SMOOTH_NORMAL: return 21;
    default: return 0;
    }
}
int leaf_out_width(int op) {
    switch (op) {
    case SRT_LEAF_POWF: case SRT_LEAF_TRIANGLE: case SRT_LEAF_TONE: return 1;
    case SRT_LEAF_SLAB: case SRT_LEAF_PHONG: case SRT_LEAF_SMOOTH_NORMAL: case SRT_LEAF_TRIANGLE_FILTERED: return 3;
    default: return 0;
    }
}
cudaError_t launch_eval_leaf(int op, const float* in, float* out, size_t n, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    eval_leaf_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(op, in, out, n, leaf_in_width(op), leaf_out_width(op));
    return cudaGetLastError();
}

namespace {
template <class K>
cudaError_t persistent_blocks(K kernel, size_t smem, int sm_count, int n_tiles, int* blocks) {
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    *blocks = sm_count * per_sm;                                   // persistent grid: exactly the resident capacity
    const int need = (n_tiles + kThreads / 32 - 1) / (kThreads / 32);
    if (*blocks > need) *blocks = need;
    return cudaSuccess;
}
}  // namespace

cudaError_t launch_render_phased(const KArgs& a, int variant, int sm_count, cudaStream_t stream, int* launches, cudaEvent_t* stage_ev) {
    const int n_tiles = a.tiles_x * a.tiles_y;
    if (launches) *launches = 0;
    if (n_tiles <= 0) return cudaSuccess;
    const size_t smem = phased_smem_bytes(a.stack_cap);
    const size_t smem_s = phased_smem_bytes(a.stack_cap - kLeafRoom);   // the shadow pass keeps its leaves in registers
    // The opt-in to more than 48 KB of dynamic shared memory is a per-DEVICE attribute of each kernel: once per device.
    {
        static std::mutex attr_mutex;
        static bool attr_set[64] = {};
        int dev = 0;
        cudaError_t ea = cudaGetDevice(&dev);
        if (ea != cudaSuccess) return ea;
        std::lock_guard<std::mutex> lock(attr_mutex);
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            const void* kernels[] = {(const void*)render_dataflow<false, false>, (const void*)render_dataflow<true, false>,
                                     (const void*)render_dataflow<false, true>, (const void*)render_phase<0, false>,
                                     (const void*)render_phase<1, false>, (const void*)render_phase<0, true>, (const void*)render_phase<1, true>};
            for (const void* k : kernels)
                if ((ea = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024)) != cudaSuccess) return ea;
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
    }
    int cblocks = (n_tiles + kThreads - 1) / kThreads;
    if (cblocks > sm_count * 8) cblocks = sm_count * 8;
    // stage_ev (srt_context_set_stage_timing): five events around the four passes of the default pipeline
    const bool timed = stage_ev && (variant == SRT_VARIANT_DEFAULT);
    if (timed) cudaEventRecord(stage_ev[0], stream);
    classify_kernel<<<cblocks, kThreads, 0, stream>>>(a);
    if (timed) cudaEventRecord(stage_ev[1], stream);
    cudaError_t e;
    const bool shadows = a.n_objects >= 2 && a.light_count > 0;
    if (variant == SRT_VARIANT_DEFAULT || variant == SRT_VARIANT_OVERLAP) {
        const bool overlap = variant == SRT_VARIANT_OVERLAP;
        int bp = 0, bs = 0, bc = 0;
        if ((e = persistent_blocks(render_phase<0, false>, smem, sm_count, n_tiles, &bp)) != cudaSuccess) return e;
        if ((e = persistent_blocks(render_phase<1, false>, smem_s, sm_count, n_tiles, &bs)) != cudaSuccess) return e;
        if ((e = persistent_blocks(render_phase<2, false>, 0, sm_count, n_tiles, &bc)) != cudaSuccess) return e;
        int n = 1;
        if (overlap && shadows) {
            render_phase<0, true><<<bp, kThreads, smem, stream>>>(a);
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)bs); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem_s; cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            if ((e = cudaLaunchKernelEx(&cfg, render_phase<1, true>, a)) != cudaSuccess) return e;
            n += 2;
        } else {
            render_phase<0, false><<<bp, kThreads, smem, stream>>>(a);
            ++n;
            if (timed) cudaEventRecord(stage_ev[2], stream);
            if (shadows) { render_phase<1, false><<<bs, kThreads, smem_s, stream>>>(a); ++n; }
            if (timed) cudaEventRecord(stage_ev[3], stream);
        }
        render_phase<2, false><<<bc, kThreads, 0, stream>>>(a);
        if (timed) cudaEventRecord(stage_ev[4], stream);
        ++n;
        if (launches) *launches = n;
        return cudaGetLastError();
    }
    int blocks = 0;
    if ((e = persistent_blocks(render_dataflow<false, false>, smem, sm_count, n_tiles, &blocks)) != cudaSuccess) return e;
    if (a.debug_timeline == 3) render_dataflow<true, false><<<blocks, kThreads, smem, stream>>>(a);   // node_visits / tri_tests = shadow box tests of occluded / free rays
    else if (a.debug_timeline != 0) render_dataflow<false, true><<<blocks, kThreads, smem, stream>>>(a);
    else render_dataflow<false, false><<<blocks, kThreads, smem, stream>>>(a);
    if (launches) *launches = 2;
    return cudaGetLastError();
}

}  // namespace srtk

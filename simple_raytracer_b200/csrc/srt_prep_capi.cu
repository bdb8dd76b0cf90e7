// srt_prep_capi.cu — C ABI of the device scene prep (include/srt.h: srt_mesh_*, srt_scene_prep, …).
// Host-side plumbing only: mesh upload, the tree topology table (a function of the triangle count alone), the
// launch plan, scratch memory.  All geometry arithmetic runs in srt_prep.cu on the device; there is no CPU path.
#include <cuda_runtime.h>

#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "srt.h"
#include "srt_internal.hpp"
#include "srt_prep.cuh"

using srti::DevBuf;
using srti::fail;

struct srt_mesh {
    srt_context* ctx = nullptr;
    DevBuf pts, uv, tex, topo, vn;
    bool has_uv = false, has_tex = false, has_vn = false;
    uint64_t n_tris = 0;
    std::vector<srtp::TopoNode> topo_host;
    uint32_t max_depth = 0;
    int32_t max_tex = -1;
};

struct srt_prep_state {
    DevBuf xpts, lite, order, key, idx, posl, posr, segq, blob, misc, wide;
    int last_launches = 0;
    cudaGraphExec_t graph = nullptr;           // the frame's launch sequence, captured once per scene composition
    std::vector<uint64_t> graph_sig;
    bool graph_failed = false;                 // capture was refused once (e.g. a stream that cannot be captured)
    std::vector<unsigned char> blob_host;
    // the recipe goes through pinned staging (a pageable cudaMemcpyAsync would first wait for the stream): two
    // buffers in turn, each guarded by the event of its last copy
    unsigned char* pin[2] = {nullptr, nullptr};
    size_t pin_cap[2] = {0, 0};
    cudaEvent_t pin_ev[2] = {nullptr, nullptr};
    int turn = 0;
    uint32_t* status_host = nullptr;       // pinned
    bool kernels_ready = false;
};

namespace {

// Object.cpp:254-267 / :282 — halves n/2 | n - n/2, a half recurses iff it has more than 8 triangles, the root always splits.
uint32_t topo_rec(std::vector<srtp::TopoNode>& out, uint32_t start, uint32_t count, uint32_t depth, bool force, uint32_t& max_depth) {
    const uint32_t me = (uint32_t)out.size();
    out.push_back(srtp::TopoNode{start, count, 0u, depth});
    if (depth > max_depth) max_depth = depth;
    if (force || count > 8) {
        const uint32_t nl = count / 2;
        topo_rec(out, start, nl, depth + 1, false, max_depth);
        out[me].right = (uint32_t)out.size();
        topo_rec(out, start + nl, count - nl, depth + 1, false, max_depth);
    }
    return me;
}

int upload(srt_context* c, DevBuf& dst, const void* src, size_t bytes) {
    if (bytes == 0) return SRT_OK;
    int rc = dst.reserve(bytes);
    if (rc) return rc;
    CU(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
    return SRT_OK;
}

template <class T>
size_t blob_put(std::vector<unsigned char>& b, const T* src, size_t n) {
    const size_t at = (b.size() + 15) & ~size_t(15);
    b.resize(at + n * sizeof(T));
    if (n) std::memcpy(b.data() + at, src, n * sizeof(T));
    return at;
}

}  // namespace

void srti_prep_state_destroy(srt_prep_state* p) {
    if (!p) return;
    p->xpts.release(); p->lite.release(); p->order.release(); p->key.release(); p->idx.release();
    p->posl.release(); p->posr.release(); p->segq.release(); p->blob.release(); p->misc.release(); p->wide.release();
    if (p->graph) cudaGraphExecDestroy(p->graph);
    if (p->status_host) cudaFreeHost(p->status_host);
    for (int k = 0; k < 2; ++k) {
        if (p->pin[k]) cudaFreeHost(p->pin[k]);
        if (p->pin_ev[k]) cudaEventDestroy(p->pin_ev[k]);
    }
    delete p;
}

extern "C" {

int srt_mesh_upload(srt_context* c, const srt_mesh_desc* d, srt_mesh** out) {
    if (!c || !d || !out) return fail(SRT_ERR_INVALID_ARG, "NULL argument");
    *out = nullptr;
    if (d->n_tris && !d->points) return fail(SRT_ERR_INVALID_ARG, "srt_mesh_upload: points is NULL");
    if (d->n_tris >= (1ull << 27)) return fail(SRT_ERR_UNSUPPORTED, "srt_mesh_upload: mesh too large (%llu triangles)", (unsigned long long)d->n_tris);
    CU(cudaSetDevice(c->device));
    srt_mesh* m = new srt_mesh();
    m->ctx = c;
    m->n_tris = d->n_tris;
    topo_rec(m->topo_host, 0, (uint32_t)d->n_tris, 0, true, m->max_depth);
    int rc = upload(c, m->pts, d->points, d->n_tris * 12 * sizeof(float));
    if (!rc && d->uv) { rc = upload(c, m->uv, d->uv, d->n_tris * 6 * sizeof(float)); m->has_uv = d->n_tris > 0; }
    if (!rc && d->tex) {
        rc = upload(c, m->tex, d->tex, d->n_tris * sizeof(int32_t));
        m->has_tex = d->n_tris > 0;
        for (uint64_t i = 0; i < d->n_tris; ++i) {
            if (d->tex[i] < -1) { rc = fail(SRT_ERR_SCENE, "srt_mesh_upload: triangle %llu has texture index %d", (unsigned long long)i, d->tex[i]); break; }
            if (d->tex[i] > m->max_tex) m->max_tex = d->tex[i];
        }
    }
    if (!rc && d->normals) { rc = upload(c, m->vn, d->normals, d->n_tris * 9 * sizeof(float)); m->has_vn = d->n_tris > 0; }
    if (!rc) rc = upload(c, m->topo, m->topo_host.data(), m->topo_host.size() * sizeof(srtp::TopoNode));
    if (!rc) {
        cudaError_t e = cudaStreamSynchronize(c->stream);           // the host arrays may be freed on return
        if (e != cudaSuccess) rc = fail(SRT_ERR_CUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(e));
    }
    if (rc) { srt_mesh_destroy(m); return rc; }
    *out = m;
    return SRT_OK;
}

void srt_mesh_destroy(srt_mesh* m) {
    if (!m) return;
    if (m->ctx) { cudaSetDevice(m->ctx->device); cudaStreamSynchronize(m->ctx->stream); }
    m->pts.release(); m->uv.release(); m->tex.release(); m->topo.release(); m->vn.release();
    delete m;
}

int srt_mesh_info(const srt_mesh* m, uint64_t* n_tris, uint64_t* n_nodes, uint32_t* max_depth) {
    if (!m) return fail(SRT_ERR_INVALID_ARG, "mesh is NULL");
    if (n_tris) *n_tris = m->n_tris;
    if (n_nodes) *n_nodes = m->topo_host.size();
    if (max_depth) *max_depth = m->max_depth;
    return SRT_OK;
}

int srt_scene_prep(srt_context* c, const srt_prep_object* objs, uint32_t n_objects, const srt_texture_desc* textures,
                   uint32_t n_textures, int with_flat_lists, srt_scene** scene) {
    if (!c || !scene || (n_objects && !objs)) return fail(SRT_ERR_INVALID_ARG, "NULL argument");
    if (n_textures && !textures && !*scene) return fail(SRT_ERR_INVALID_ARG, "textures is NULL");
    CU(cudaSetDevice(c->device));

    // ---- placement of the objects in the scene arrays, validation ----
    uint64_t n_tris = 0, n_nodes = 0, n_mats = 0;
    uint32_t max_depth = 0, max_tris = 0;
    for (uint32_t i = 0; i < n_objects; ++i) {
        const srt_prep_object& po = objs[i];
        // a mesh is immutable after its (synchronous) upload: any context of the same device may read it
        if (!po.mesh || !po.mesh->ctx || po.mesh->ctx->device != c->device)
            return fail(SRT_ERR_INVALID_ARG, "object %u: mesh is NULL or lives on another device", i);
        if (po.n_matrices && !po.matrices) return fail(SRT_ERR_INVALID_ARG, "object %u: matrices is NULL", i);
        n_tris += po.mesh->n_tris; n_nodes += po.mesh->topo_host.size(); n_mats += po.n_matrices;
        if (po.mesh->max_depth > max_depth) max_depth = po.mesh->max_depth;
        if (po.mesh->n_tris > max_tris) max_tris = (uint32_t)po.mesh->n_tris;
    }
    if (n_tris >= (1ull << srtk::kLeafFirstBits) || n_nodes >= 0x40000000ull)
        return fail(SRT_ERR_UNSUPPORTED, "scene too large for the packed traversal references");
    if (max_depth + 2 > (uint32_t)srtk::kMaxStack)
        return fail(SRT_ERR_UNSUPPORTED, "BVH depth %u exceeds the kernel's traversal stack (%d)", max_depth, srtk::kMaxStack);

    srt_scene* s = *scene;
    const bool fresh = s == nullptr;
    if (fresh) { s = new srt_scene(); s->ctx = c; }
    else if (s->ctx != c) return fail(SRT_ERR_INVALID_ARG, "scene does not belong to this context");
    auto bail = [&](int rc) { if (fresh) srt_scene_destroy(s); return rc; };
    if (!s->prep) s->prep = new srt_prep_state();
    srt_prep_state& ps = *s->prep;
    if (!ps.kernels_ready) {
        cudaError_t e = srtp::prepare_kernels();
        if (e != cudaSuccess) return bail(fail(SRT_ERR_CUDA, "scene prep kernels: %s", cudaGetErrorString(e)));
        ps.kernels_ready = true;
    }
    if (!ps.status_host) {
        cudaError_t e = cudaMallocHost(&ps.status_host, sizeof(uint32_t));
        if (e != cudaSuccess) return bail(fail(SRT_ERR_OOM, "cudaMallocHost: %s", cudaGetErrorString(e)));
    }

    // ---- textures: uploaded with a new scene, or when the caller passes a new set ----
    int rc;
    if (fresh || (textures && n_textures)) {
        for (auto& t : s->tex) t.release();
        s->tex.assign(n_textures, DevBuf());
        s->tex_host.assign(n_textures, srtk::TexDev());
        for (uint32_t t = 0; t < n_textures; ++t) {
            if (!textures[t].rgb || textures[t].width <= 0 || textures[t].height <= 0)
                return bail(fail(SRT_ERR_SCENE, "texture %u has no data (the reference dereferences textureData[...] at simple_raytracer.cpp:354-358)", t));
            const size_t bytes = (size_t)textures[t].width * textures[t].height * 3;
            if ((rc = upload(c, s->tex[t], textures[t].rgb, bytes))) return bail(rc);
            s->tex_host[t].rgb = (const uint8_t*)s->tex[t].p;
            s->tex_host[t].width = textures[t].width; s->tex_host[t].height = textures[t].height;
        }
        s->n_textures = n_textures;
        if ((rc = upload(c, s->texdesc, s->tex_host.data(), s->tex_host.size() * sizeof(srtk::TexDev)))) return bail(rc);
        cudaError_t e = cudaStreamSynchronize(c->stream);           // s->tex_host is reused; texture arrays may be freed on return
        if (e != cudaSuccess) return bail(fail(SRT_ERR_CUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(e)));
    }
    for (uint32_t i = 0; i < n_objects; ++i)
        if (objs[i].mesh->max_tex >= (int32_t)s->n_textures)
            return bail(fail(SRT_ERR_SCENE, "object %u: textured triangle references texture %d, the scene has %u (textured triangle without texture, simple_raytracer.cpp:354)",
                             i, objs[i].mesh->max_tex, s->n_textures));

    // ---- device memory ----
    const size_t N = (size_t)n_tris;
    const size_t n_flat = with_flat_lists ? N : 0;
    if ((rc = s->nodes.reserve(n_nodes * sizeof(srt_node)))) return bail(rc);
    if ((rc = s->geom.reserve(N * sizeof(srt_tri_geom)))) return bail(rc);
    if ((rc = s->attr.reserve(N * sizeof(srt_tri_attr)))) return bail(rc);
    if ((rc = s->flat_geom.reserve(n_flat * sizeof(srt_tri_geom)))) return bail(rc);
    if ((rc = s->flat_attr.reserve(n_flat * sizeof(srt_tri_attr)))) return bail(rc);
    if ((rc = s->vnormal.reserve(N * sizeof(srt_tri_vnormal)))) return bail(rc);
    if ((rc = s->flat_vnormal.reserve(n_flat * sizeof(srt_tri_vnormal)))) return bail(rc);
    if ((rc = s->objects.reserve((size_t)n_objects * sizeof(srt_object)))) return bail(rc);
    if ((rc = s->wide.reserve(n_nodes * 64))) return bail(rc);
    if ((rc = s->pwide.reserve(n_nodes * 64 * 4))) return bail(rc);            // quadrant-sorted copies for the primary pass
    if ((rc = s->root_ref.reserve((size_t)n_objects * sizeof(uint32_t)))) return bail(rc);
    const size_t segq_stride = N / 8 + 8;
    if ((rc = ps.xpts.reserve(N * 48))) return bail(rc);
    if ((rc = ps.lite.reserve(N * 24))) return bail(rc);
    if ((rc = ps.order.reserve(N * 4))) return bail(rc);
    if ((rc = ps.key.reserve(N * 4))) return bail(rc);
    if ((rc = ps.idx.reserve(N * 4))) return bail(rc);
    if ((rc = ps.posl.reserve(N * 4))) return bail(rc);
    if ((rc = ps.posr.reserve(N * 4))) return bail(rc);
    if ((rc = ps.segq.reserve(2 * segq_stride * 12))) return bail(rc);

    // ---- the frame's recipe: objects, matrices, work lists, accumulator seeds — one upload ----
    std::vector<srtp::PrepObject> pobj(n_objects);
    std::vector<float> mats(16 * (size_t)n_mats);
    std::vector<std::vector<srtp::WorkItem>> big;                  // per depth: nodes too large for one CTA's shared memory
    std::vector<srtp::WorkItem> sub;                               // subtree roots
    uint32_t tri_base = 0, node_base = 0, mat_first = 0;
    for (uint32_t i = 0; i < n_objects; ++i) {
        const srt_prep_object& po = objs[i];
        const srt_mesh& m = *po.mesh;
        srtp::PrepObject& q = pobj[i];
        q.pts = (const float4*)m.pts.p;
        q.uv = m.has_uv ? (const float2*)m.uv.p : nullptr;
        q.tex = m.has_tex ? (const int32_t*)m.tex.p : nullptr;
        q.vn = m.has_vn ? (const float*)m.vn.p : nullptr;
        q.topo = (const srtp::TopoNode*)m.topo.p;
        q.n_tris = (uint32_t)m.n_tris; q.n_nodes = (uint32_t)m.topo_host.size();
        q.tri_base = tri_base; q.node_base = node_base;
        q.flat_first = with_flat_lists ? tri_base : SRT_NO_INDEX;
        q.n_matrices = po.n_matrices; q.mat_first = mat_first;
        for (int k = 0; k < 3; ++k) q.color[k] = po.color[k];
        q.ambient = po.ambient; q.specular = po.specular; q.shininess = po.shininess;
        if (po.n_matrices) std::memcpy(mats.data() + 16 * (size_t)mat_first, po.matrices, 16 * sizeof(float) * po.n_matrices);
        // launch plan: walk down from the root while the nodes are too large for one CTA
        std::vector<uint32_t> frontier{0u};
        for (uint32_t depth = 0; !frontier.empty(); ++depth) {
            std::vector<uint32_t> next;
            for (uint32_t n : frontier) {
                const srtp::TopoNode& t = m.topo_host[n];
                if (t.right == 0) continue;                        // leaf: its box comes from its parent's split
                const srtp::WorkItem w{i, n, t.start, t.count};
                if (t.count > (uint32_t)srtp::kSubtreeMax) {
                    if (big.size() <= depth) big.resize(depth + 1);
                    big[depth].push_back(w);
                    next.push_back(n + 1); next.push_back(t.right);
                } else {
                    sub.push_back(w);
                }
            }
            frontier.swap(next);
        }
        tri_base += q.n_tris; node_base += q.n_nodes; mat_first += po.n_matrices;
    }
    std::vector<unsigned long long> seeds(6 * (size_t)n_objects + 2);
    {
        // packed (value code, position) minima, see srt_prep.cu: code(FLT_MAX) for the lower, ~code(-FLT_MAX) for the upper corners
        const float fmx = FLT_MAX;
        uint32_t bits; std::memcpy(&bits, &fmx, 4);
        const unsigned long long lo_seed = ((unsigned long long)(bits | 0x80000000u) << 32) | 0xFFFFFFFFull;
        const uint32_t neg_code = ~(bits | 0x80000000u);           // code(-FLT_MAX) = ~bits(-FLT_MAX)
        const unsigned long long hi_seed = ((unsigned long long)(~neg_code) << 32) | 0xFFFFFFFFull;
        for (uint32_t i = 0; i < n_objects; ++i)
            for (int k = 0; k < 6; ++k) seeds[6 * (size_t)i + k] = k < 3 ? lo_seed : hi_seed;
        seeds[6 * (size_t)n_objects] = 0;                          // coord_bound (float 0) + status word
        seeds[6 * (size_t)n_objects + 1] = 0;
    }
    std::vector<unsigned char>& blob = ps.blob_host;
    blob.clear();
    const size_t at_seeds = blob_put(blob, seeds.data(), seeds.size());
    const size_t at_obj = blob_put(blob, pobj.data(), pobj.size());
    const size_t at_mat = blob_put(blob, mats.data(), mats.size());
    std::vector<size_t> at_big(big.size());
    for (size_t d = 0; d < big.size(); ++d) at_big[d] = blob_put(blob, big[d].data(), big[d].size());
    const size_t at_sub = blob_put(blob, sub.data(), sub.size());
    if ((rc = ps.blob.reserve(blob.size() + 16))) return bail(rc);
    unsigned char* dblob = (unsigned char*)ps.blob.p;

    srtp::PrepArgs a;
    std::memset(&a, 0, sizeof a);
    a.objects = (const srtp::PrepObject*)(dblob + at_obj);
    a.n_objects = n_objects;
    a.matrices = (const float*)(dblob + at_mat);
    a.nodes = (float4*)s->nodes.p; a.geom = (float4*)s->geom.p; a.attr = (float4*)s->attr.p;
    a.flat_geom = (float4*)s->flat_geom.p; a.flat_attr = (float4*)s->flat_attr.p;
    a.vnormal = (float4*)s->vnormal.p; a.flat_vnormal = (float4*)s->flat_vnormal.p;
    a.scene_objects = (srt_object*)s->objects.p;
    a.rootacc = (unsigned long long*)(dblob + at_seeds);
    a.coord_bound = (float*)(dblob + at_seeds + 6 * (size_t)n_objects * 8);
    a.status = (uint32_t*)(dblob + at_seeds + 6 * (size_t)n_objects * 8 + 8);
    a.xpts = (float4*)ps.xpts.p; a.lite = (float*)ps.lite.p; a.order = (uint32_t*)ps.order.p;
    a.key = (float*)ps.key.p; a.idx = (uint32_t*)ps.idx.p; a.posl = (uint32_t*)ps.posl.p; a.posr = (uint32_t*)ps.posr.p;
    a.segq = (uint32_t*)ps.segq.p; a.segq_stride = (uint32_t)segq_stride;
    static const bool debug_timing = std::getenv("SRT_PREP_DEBUG") != nullptr;
    if (debug_timing) { if ((rc = ps.misc.reserve(8 * 64 * 8))) return bail(rc); a.dbg = (unsigned long long*)ps.misc.p; }

    // workspace of the many-CTA path for the large nodes
    size_t max_items = 1, n_big_total = 0;
    for (const auto& lv : big) { if (lv.size() > max_items) max_items = lv.size(); n_big_total += lv.size(); }
    srtp::WideArgs w;
    std::memset(&w, 0, sizeof w);
    w.max_segs = (uint32_t)(N / 4096 + 2 * max_items + 64);
    w.max_tiles = (uint32_t)(N / srtp::kWideTile + w.max_segs + 64);
    w.max_fin = (uint32_t)(N / 16 + max_items + 64);
    {
        size_t off = 0;
        auto carve = [&](size_t bytes) { const size_t at = off; off = (off + bytes + 255) & ~size_t(255); return at; };
        const size_t o_segs = carve(2 * (size_t)w.max_segs * sizeof(srtp::WideSeg));
        const size_t o_tseg = carve(2 * (size_t)w.max_tiles * sizeof(uint2));
        const size_t o_tcnt = carve((size_t)w.max_tiles * sizeof(uint2));
        const size_t o_rnd = carve((srtp::kWideMaxRounds + 1) * sizeof(srtp::WideRound));
        const size_t o_fin = carve((size_t)w.max_fin * sizeof(srtp::WideSmall));
        const size_t o_cnt = carve(sizeof(uint32_t));
        const size_t o_acc = carve(max_items * 12 * sizeof(unsigned long long));
        if (n_big_total && (rc = ps.wide.reserve(off))) return bail(rc);
        unsigned char* wb = (unsigned char*)ps.wide.p;
        w.segs = (srtp::WideSeg*)(wb + o_segs); w.tile_seg = (uint2*)(wb + o_tseg); w.tile_cnt = (uint2*)(wb + o_tcnt);
        w.rounds = (srtp::WideRound*)(wb + o_rnd); w.fin = (srtp::WideSmall*)(wb + o_fin); w.fin_count = (uint32_t*)(wb + o_cnt);
        w.childacc = (unsigned long long*)(wb + o_acc);
    }
    static const bool single_cta_big = std::getenv("SRT_PREP_SINGLE_CTA") != nullptr;     // A/B: the one-CTA-per-node path
    static const bool use_graph = std::getenv("SRT_PREP_NO_GRAPH") == nullptr;
    int launches = 0;
    // Everything the frame enqueues: the recipe upload, the kernels, the status read-back.
    auto enqueue = [&](const unsigned char* pinned) -> cudaError_t {
        cudaError_t e = cudaMemcpyAsync(ps.blob.p, pinned, blob.size(), cudaMemcpyHostToDevice, c->stream);
        if (e == cudaSuccess) e = srtp::launch_transform(a, max_tris, c->stream);
        if (e == cudaSuccess) e = srtp::launch_roots(a, c->stream);
        launches = 2;
        for (size_t d = 0; d < big.size() && e == cudaSuccess; ++d) {
            if (big[d].empty()) continue;
            if (single_cta_big) {
                e = srtp::launch_build(a, (const srtp::WorkItem*)(dblob + at_big[d]), (uint32_t)big[d].size(), false, c->stream);
                ++launches;
                continue;
            }
            uint32_t maxc = 0;
            size_t tiles = 8, total = 0;
            for (const srtp::WorkItem& it : big[d]) { if (it.count > maxc) maxc = it.count; total += it.count; tiles += it.count / srtp::kWideTile + 1; }
            tiles += 2 * (total / srtp::kWideSmall + big[d].size());
            // planned many-CTA rounds: until the segments fit one CTA if every pivot split evenly, plus three (a median of
            // three rarely does; a round with nothing left to do costs three empty launches, a segment left too long costs
            // a one-CTA sort in global memory)
            uint32_t rounds = 0;
            static const int slack = std::getenv("SRT_PREP_SLACK") ? std::atoi(std::getenv("SRT_PREP_SLACK")) : 3;
            if (maxc > srtp::kWideSmall) { rounds = (uint32_t)slack; for (uint64_t v = srtp::kWideSmall; v < maxc; v *= 2) ++rounds; }
            if (rounds > (uint32_t)srtp::kWideMaxRounds - 1) rounds = srtp::kWideMaxRounds - 1;
            if (tiles > w.max_tiles) tiles = w.max_tiles;
            e = srtp::launch_wide_level(a, w, (const srtp::WorkItem*)(dblob + at_big[d]), (uint32_t)big[d].size(), maxc, rounds,
                                        (uint32_t)tiles, c->sm_count, c->stream, &launches);
        }
        if (e == cudaSuccess) e = srtp::launch_build(a, (const srtp::WorkItem*)(dblob + at_sub), (uint32_t)sub.size(), true, c->stream);
        if (e == cudaSuccess) e = srtp::launch_flatten(a, max_tris, with_flat_lists != 0, c->stream);
        if (e == cudaSuccess && n_objects > 0)
            e = srtk::launch_build_wide((const float4*)s->nodes.p, (const srt_object*)s->objects.p, (uint32_t)n_nodes, n_objects,
                                        (float4*)s->wide.p, (uint32_t*)s->root_ref.p, c->stream);
        for (uint32_t q = 0; q < 4 && e == cudaSuccess && n_objects > 0; ++q)
            e = srtk::launch_build_shadow_wide((const float4*)s->wide.p, (uint32_t)n_nodes, q, (float4*)s->pwide.p + (size_t)q * n_nodes * 4, c->stream);
        launches += 7;
        if (e == cudaSuccess) e = cudaMemcpyAsync(ps.status_host, a.status, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream);
        return e;
    };
    auto stage = [&](int k) -> int {                               // recipe -> pinned buffer k (its previous copy must be over)
        if (!ps.pin_ev[k]) CU(cudaEventCreateWithFlags(&ps.pin_ev[k], cudaEventDisableTiming));
        else CU(cudaEventSynchronize(ps.pin_ev[k]));
        if (ps.pin_cap[k] < blob.size()) {
            if (ps.pin[k]) cudaFreeHost(ps.pin[k]);
            ps.pin[k] = nullptr; ps.pin_cap[k] = 0;
            const size_t want = blob.size() + blob.size() / 2 + 4096;
            cudaError_t pe = cudaMallocHost(&ps.pin[k], want);
            if (pe != cudaSuccess) return fail(SRT_ERR_OOM, "cudaMallocHost(%zu): %s", want, cudaGetErrorString(pe));
            ps.pin_cap[k] = want;
        }
        std::memcpy(ps.pin[k], blob.data(), blob.size());
        return SRT_OK;
    };
    cudaError_t e = cudaSuccess;
    // (the legacy default stream cannot be captured: a context bound to it launches kernel by kernel)
    const bool graph_ok = use_graph && !ps.graph_failed && c->stream != nullptr && c->stream != cudaStreamLegacy &&
                          c->stream != cudaStreamPerThread;
    bool launched = false;
    if (graph_ok) {
        // The launch sequence of a frame depends on the composition of the scene only (which meshes, how many
        // matrices, where the buffers live), not on its contents: it is captured once into a CUDA graph and replayed
        // — one launch instead of about eighty.  The recipe always travels through pinned buffer 0.
        std::vector<uint64_t> sig;
        auto put = [&](const void* ptr) { sig.push_back((uint64_t)(uintptr_t)ptr); };
        put(s->nodes.p); put(s->geom.p); put(s->attr.p); put(s->flat_geom.p); put(s->flat_attr.p); put(s->objects.p);
        put(s->wide.p); put(s->pwide.p); put(s->root_ref.p); put(s->vnormal.p); put(s->flat_vnormal.p); put(ps.xpts.p); put(ps.lite.p); put(ps.order.p); put(ps.key.p); put(ps.idx.p);
        put(ps.posl.p); put(ps.posr.p); put(ps.segq.p); put(ps.blob.p); put(ps.wide.p); put(ps.misc.p); put(c->stream);
        sig.push_back(blob.size()); sig.push_back(n_objects); sig.push_back((uint64_t)(with_flat_lists != 0));
        sig.push_back((uint64_t)single_cta_big);
        for (uint32_t i = 0; i < n_objects; ++i) { put(objs[i].mesh); sig.push_back(objs[i].n_matrices); }
        if ((rc = stage(0))) return bail(rc);
        put(ps.pin[0]);
        if (!ps.graph || sig != ps.graph_sig) {
            if (ps.graph) { cudaGraphExecDestroy(ps.graph); ps.graph = nullptr; }
            cudaGraph_t g = nullptr;
            e = cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal);
            if (e == cudaSuccess) {
                const cudaError_t e1 = enqueue(ps.pin[0]);
                const cudaError_t e2 = cudaStreamEndCapture(c->stream, &g);
                e = e1 != cudaSuccess ? e1 : e2;
            }
            if (e == cudaSuccess) e = cudaGraphInstantiate(&ps.graph, g, 0);
            if (g) cudaGraphDestroy(g);
            if (e != cudaSuccess) {                                // not capturable here: plain launches from now on
                cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
                if (cudaStreamIsCapturing(c->stream, &st) == cudaSuccess && st != cudaStreamCaptureStatusNone) {
                    cudaGraph_t dead = nullptr;
                    cudaStreamEndCapture(c->stream, &dead);
                    if (dead) cudaGraphDestroy(dead);
                }
                cudaGetLastError();
                ps.graph = nullptr; ps.graph_sig.clear(); ps.graph_failed = true;
                e = cudaSuccess;
            } else {
                ps.graph_sig = sig;
            }
        } else {
            launches = ps.last_launches;
        }
        if (ps.graph) {
            e = cudaGraphLaunch(ps.graph, c->stream);
            if (e == cudaSuccess) e = cudaEventRecord(ps.pin_ev[0], c->stream);
            launched = true;
        }
    }
    if (!launched) {
        const int k = ps.turn;
        ps.turn ^= 1;
        if ((rc = stage(k))) return bail(rc);
        e = enqueue(ps.pin[k]);
        if (e == cudaSuccess) e = cudaEventRecord(ps.pin_ev[k], c->stream);
    }
    if (e != cudaSuccess) return bail(fail(SRT_ERR_CUDA, "scene prep launch: %s", cudaGetErrorString(e)));
    ps.last_launches = launches;

    s->n_nodes = n_nodes; s->n_tris = n_tris; s->n_flat = n_flat; s->n_objects = n_objects;
    s->max_depth = max_depth;
    s->phased_ok = n_objects > 0;                                  // leaves hold at most 8 triangles by construction
    s->swide_octant = -1;                                          // derived from `wide`: the next render rebuilds it
    s->has_vnormal = N > 0; s->has_flat_vnormal = n_flat > 0;      // (zeros for a mesh uploaded without normals)
    s->coord_bound = 0.f;
    s->coord_bound_dev = a.coord_bound;
    *scene = s;
    return SRT_OK;
}

int srt_scene_prep_status(srt_context* c, srt_scene* s) {
    if (!c || !s || s->ctx != c) return fail(SRT_ERR_INVALID_ARG, "scene does not belong to this context");
    if (!s->prep || !s->prep->status_host) return fail(SRT_ERR_INVALID_ARG, "scene was not prepared with srt_scene_prep");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    if (std::getenv("SRT_PREP_DEBUG") && s->prep->misc.p) {       // last level's finishing CTAs: per round, segments and cycles
        unsigned long long h[8 * 64];
        cudaMemcpy(h, s->prep->misc.p, sizeof h, cudaMemcpyDeviceToHost);
        for (int e = 0; e < 8; ++e) {
            std::fprintf(stderr, "finish cta %d:", e);
            for (int r = 0; r < 64 && h[64 * e + r]; ++r) std::fprintf(stderr, " %llu/%llu", h[64 * e + r] >> 40, h[64 * e + r] & ((1ull << 40) - 1));
            std::fprintf(stderr, "\n");
        }
    }
    if (*s->prep->status_host & 2u)
        return fail(SRT_ERR_CUDA, "srt_scene_prep: internal work-list overflow (please report)");
    if (*s->prep->status_host & 1u)
        return fail(SRT_ERR_UNSUPPORTED, "srt_scene_prep: a transformed coordinate is not finite (std::sort's comparison is not a strict weak order on NaN keys); prepare this frame on the host");
    return SRT_OK;
}

int srt_scene_counts(const srt_scene* s, uint64_t* n_nodes, uint64_t* n_tris, uint64_t* n_flat, uint32_t* n_objects, uint32_t* max_depth) {
    if (!s) return fail(SRT_ERR_INVALID_ARG, "scene is NULL");
    if (n_nodes) *n_nodes = s->n_nodes;
    if (n_tris) *n_tris = s->n_tris;
    if (n_flat) *n_flat = s->n_flat;
    if (n_objects) *n_objects = s->n_objects;
    if (max_depth) *max_depth = s->max_depth;
    return SRT_OK;
}

int srt_scene_download_vnormals(srt_context* c, const srt_scene* s, srt_tri_vnormal* vnormal, srt_tri_vnormal* flat_vnormal) {
    if (!c || !s || s->ctx != c) return fail(SRT_ERR_INVALID_ARG, "scene does not belong to this context");
    CU(cudaSetDevice(c->device));
    if (vnormal && s->n_tris) {
        if (!s->has_vnormal) return fail(SRT_ERR_INVALID_ARG, "scene has no vertex normals");
        CU(cudaMemcpyAsync(vnormal, s->vnormal.p, s->n_tris * sizeof(srt_tri_vnormal), cudaMemcpyDeviceToHost, c->stream));
    }
    if (flat_vnormal && s->n_flat) {
        if (!s->has_flat_vnormal) return fail(SRT_ERR_INVALID_ARG, "scene has no load-order vertex normals");
        CU(cudaMemcpyAsync(flat_vnormal, s->flat_vnormal.p, s->n_flat * sizeof(srt_tri_vnormal), cudaMemcpyDeviceToHost, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    return SRT_OK;
}

int srt_scene_download(srt_context* c, const srt_scene* s, srt_node* nodes, srt_tri_geom* geom, srt_tri_attr* attr,
                       srt_tri_geom* flat_geom, srt_tri_attr* flat_attr, srt_object* objects) {
    if (!c || !s || s->ctx != c) return fail(SRT_ERR_INVALID_ARG, "scene does not belong to this context");
    CU(cudaSetDevice(c->device));
    if (nodes && s->n_nodes) CU(cudaMemcpyAsync(nodes, s->nodes.p, s->n_nodes * sizeof(srt_node), cudaMemcpyDeviceToHost, c->stream));
    if (geom && s->n_tris) CU(cudaMemcpyAsync(geom, s->geom.p, s->n_tris * sizeof(srt_tri_geom), cudaMemcpyDeviceToHost, c->stream));
    if (attr && s->n_tris) CU(cudaMemcpyAsync(attr, s->attr.p, s->n_tris * sizeof(srt_tri_attr), cudaMemcpyDeviceToHost, c->stream));
    if (flat_geom && s->n_flat) CU(cudaMemcpyAsync(flat_geom, s->flat_geom.p, s->n_flat * sizeof(srt_tri_geom), cudaMemcpyDeviceToHost, c->stream));
    if (flat_attr && s->n_flat) CU(cudaMemcpyAsync(flat_attr, s->flat_attr.p, s->n_flat * sizeof(srt_tri_attr), cudaMemcpyDeviceToHost, c->stream));
    if (objects && s->n_objects) CU(cudaMemcpyAsync(objects, s->objects.p, (size_t)s->n_objects * sizeof(srt_object), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return SRT_OK;
}

}  // extern "C"

// srt_flatten.hpp — host-side flattener: reference scene objects -> the plain arrays of srt.h.
//
// Header-only and duck-typed: it works on ANY type that has the member names of the reference's
// ObjectManager / Node / Triangle (Object.h:15-89) — the reference's own classes when a
// maintainer drops this core into the reference, or this repo's scene mirror
// (srt_scene.hpp).  It reads the scene, never changes it.
//
// What it must reproduce exactly:
//   S2  object order  = iteration order of the objTriangles unordered_map (simple_raytracer.cpp:409)
//   S5  triangle order = depth-first, left child first, concatenation of leaf->triangles (:296-317)
//   S7  p0/e1/e2      = vec3(P)/P.w and the two edge subtractions of rayTriangleIntersection :45-51
//   S12 flat normal   = calculateTriangleNormal :32-37 from the UN-divided xyz
// All arithmetic below is single fp32 operations in the reference's order.  Compile the
// including translation unit WITHOUT floating-point contraction (-ffp-contract=off; the x86-64
// baseline ISA has no FMA, so a default g++ build of the reference is fine).
#ifndef SRT_FLATTEN_HPP_
#define SRT_FLATTEN_HPP_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "srt.h"

#if defined(__FMA__) && defined(__FP_FAST_FMAF) && !defined(SRT_FLATTEN_ALLOW_FMA_ISA)
#error "srt_flatten.hpp: build with -ffp-contract=off (and define SRT_FLATTEN_ALLOW_FMA_ISA) when the target ISA has FMA"
#endif

namespace srt {

struct FlatScene {
    std::vector<srt_node>     nodes;
    std::vector<srt_tri_geom> geom;       // leaf (candidate) order
    std::vector<srt_tri_attr> attr;
    std::vector<srt_tri_geom> flat_geom;  // objTriangles load order (legacy one-box path), optional
    std::vector<srt_tri_attr> flat_attr;
    std::vector<srt_tri_vnormal> vnormal, flat_vnormal;   // vertex normals (smooth_normals), same orders
    std::vector<srt_object>   objects;
    std::vector<srt_texture_desc> textures;
    std::vector<std::string>  object_names;
    std::vector<std::string>  texture_names;
    uint32_t max_depth = 0;

    srt_scene_desc desc() const {
        srt_scene_desc d;
        std::memset(&d, 0, sizeof d);
        d.nodes = nodes.data();          d.n_nodes = nodes.size();
        d.geom = geom.data();            d.attr = attr.data();   d.n_tris = geom.size();
        d.flat_geom = flat_geom.empty() ? nullptr : flat_geom.data();
        d.flat_attr = flat_attr.empty() ? nullptr : flat_attr.data();
        d.n_flat = flat_geom.size();
        d.objects = objects.data();      d.n_objects = (uint32_t)objects.size();
        d.textures = textures.empty() ? nullptr : textures.data();
        d.n_textures = (uint32_t)textures.size();
        d.vnormal = vnormal.size() == geom.size() && !vnormal.empty() ? vnormal.data() : nullptr;
        d.flat_vnormal = flat_vnormal.size() == flat_geom.size() && !flat_vnormal.empty() ? flat_vnormal.data() : nullptr;
        return d;
    }
};

namespace detail {

// glm::vec3(v4) / v4.w  — component-wise true division (glm type_vec3.inl operator/)
template <class V4>
inline void cart(const V4& p, float out[3]) {
    out[0] = p.x / p.w; out[1] = p.y / p.w; out[2] = p.z / p.w;
}

template <class Tri>
inline void fill_triangle(const Tri& t, int32_t obj, int32_t tex, srt_tri_geom& g, srt_tri_attr& a) {
    float p1[3], p2[3], p3[3];
    cart(t.pointOne, p1); cart(t.pointTwo, p2); cart(t.pointThree, p3);
    for (int k = 0; k < 3; ++k) {
        g.p0[k] = p1[k];
        g.e1[k] = p2[k] - p1[k];                       // p1p2  :50
        g.e2[k] = p3[k] - p1[k];                       // p1p3  :51
    }
    g.obj = obj; g.tex = tex; g.reserved = 0;

    // calculateTriangleNormal :32-37 — vec3(vec4 - vec4): xyz of the raw points, no /w.
    const float v1x = t.pointTwo.x - t.pointOne.x, v1y = t.pointTwo.y - t.pointOne.y, v1z = t.pointTwo.z - t.pointOne.z;
    const float v2x = t.pointThree.x - t.pointOne.x, v2y = t.pointThree.y - t.pointOne.y, v2z = t.pointThree.z - t.pointOne.z;
    // glm::cross (func_geometric.inl:73-83)
    const float cx = v1y * v2z - v2y * v1z;
    const float cy = v1z * v2x - v2z * v1x;
    const float cz = v1x * v2y - v2x * v1y;
    // glm::normalize = v * inversesqrt(dot(v,v)); inversesqrt(x) = 1/sqrt(x); dot = (x+y)+z
    const float dd = cx * cx + cy * cy + cz * cz;
    const float inv = 1.0f / std::sqrt(dd);
    a.n[0] = cx * inv; a.n[1] = cy * inv; a.n[2] = cz * inv;
    a.pad0 = 0.0f;
    a.uv[0][0] = t.colorOneCoordinate.x;   a.uv[0][1] = t.colorOneCoordinate.y;
    a.uv[1][0] = t.colorTwoCoordinate.x;   a.uv[1][1] = t.colorTwoCoordinate.y;
    a.uv[2][0] = t.colorThreeCoordinate.x; a.uv[2][1] = t.colorThreeCoordinate.y;
    a.pad1[0] = a.pad1[1] = 0.0f;
}

// Triangle::normalOne/Two/Three as stored (never transformed by the reference).
template <class Tri>
inline void fill_vnormal(const Tri& t, srt_tri_vnormal& v) {
    v.n[0][0] = t.normalOne.x;   v.n[0][1] = t.normalOne.y;   v.n[0][2] = t.normalOne.z;   v.n[0][3] = 0.f;
    v.n[1][0] = t.normalTwo.x;   v.n[1][1] = t.normalTwo.y;   v.n[1][2] = t.normalTwo.z;   v.n[1][3] = 0.f;
    v.n[2][0] = t.normalThree.x; v.n[2][1] = t.normalThree.y; v.n[2][2] = t.normalThree.z; v.n[2][3] = 0.f;
}

template <class OM>
struct Flattener {
    OM& om;
    FlatScene& fs;
    std::unordered_map<std::string, int32_t> tex_index;
    std::string err;

    Flattener(OM& o, FlatScene& f) : om(o), fs(f) {}

    // Texture slot for a triangle; -1 = untextured; -2 = error (reference would dereference a
    // null/absent textureData entry at simple_raytracer.cpp:354-358).
    int32_t texture_of(const std::string& name) {
        if (name.empty()) return -1;
        auto it = tex_index.find(name);
        if (it != tex_index.end()) return it->second;
        auto td = om.textureData.find(name);
        auto dim = om.textureDimensions.find(name);
        if (td == om.textureData.end() || !td->second || dim == om.textureDimensions.end() ||
            dim->second.x <= 0 || dim->second.y <= 0) {
            err = "textured triangle references texture '" + name + "' that is not loaded";
            return -2;
        }
        srt_texture_desc d;
        d.rgb = td->second; d.width = dim->second.x; d.height = dim->second.y;
        int32_t idx = (int32_t)fs.textures.size();
        fs.textures.push_back(d);
        fs.texture_names.push_back(name);
        tex_index.emplace(name, idx);
        return idx;
    }

    template <class NodeT>
    bool walk(const NodeT* n, int32_t obj, uint32_t depth) {
        if (depth > fs.max_depth) fs.max_depth = depth;
        const uint32_t me = (uint32_t)fs.nodes.size();
        fs.nodes.emplace_back();
        {
            srt_node& nd = fs.nodes[me];
            nd.bmin[0] = n->minBox.x; nd.bmin[1] = n->minBox.y; nd.bmin[2] = n->minBox.z;
            nd.bmax[0] = n->maxBox.x; nd.bmax[1] = n->maxBox.y; nd.bmax[2] = n->maxBox.z;
        }
        if (!n->left && !n->right) {                        // leaf test of :303
            const uint32_t first = (uint32_t)fs.geom.size();
            for (const auto& t : n->triangles) {
                int32_t tex = texture_of(t.textureName);
                if (tex == -2) return false;
                fs.geom.emplace_back(); fs.attr.emplace_back(); fs.vnormal.emplace_back();
                fill_triangle(t, obj, tex, fs.geom.back(), fs.attr.back());
                fill_vnormal(t, fs.vnormal.back());
            }
            fs.nodes[me].a = first;
            fs.nodes[me].b = SRT_LEAF_BIT | (uint32_t)n->triangles.size();
            return true;
        }
        if (!n->left || !n->right) {                        // :308-309 would dereference null
            err = "BVH node with exactly one child";
            return false;
        }
        if (!walk(n->left, obj, depth + 1)) return false;   // left child == me + 1
        fs.nodes[me].a = (uint32_t)fs.nodes.size();         // right child
        fs.nodes[me].b = 0;
        return walk(n->right, obj, depth + 1);
    }
};

// Texture slots were handed out in first-encounter order of the leaf walk, and that order moves with the camera (the
// hierarchy is rebuilt from view-space vertices every frame).  Renumber them by texture NAME so that the same set of
// textures always gets the same slots: a scene updated in place (srt_scene_update) then keeps a valid texture table.
inline void canonical_texture_slots(FlatScene& fs) {
    const size_t n = fs.textures.size();
    if (n < 2) return;
    std::vector<int32_t> order(n);
    for (size_t i = 0; i < n; ++i) order[i] = (int32_t)i;
    std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return fs.texture_names[(size_t)a] < fs.texture_names[(size_t)b]; });
    std::vector<int32_t> slot_of(n);
    std::vector<srt_texture_desc> tex(n);
    std::vector<std::string> names(n);
    for (size_t k = 0; k < n; ++k) {
        slot_of[(size_t)order[k]] = (int32_t)k;
        tex[k] = fs.textures[(size_t)order[k]];
        names[k] = fs.texture_names[(size_t)order[k]];
    }
    fs.textures.swap(tex);
    fs.texture_names.swap(names);
    for (auto& g : fs.geom) if (g.tex >= 0) g.tex = slot_of[(size_t)g.tex];
    for (auto& g : fs.flat_geom) if (g.tex >= 0) g.tex = slot_of[(size_t)g.tex];
}

template <class M, class K>
inline bool lookup3(M& m, const K& key, float out[3]) {
    auto it = m.find(key);
    if (it == m.end()) return false;
    out[0] = it->second.x; out[1] = it->second.y; out[2] = it->second.z;
    return true;
}

}  // namespace detail

// Flatten `om`.  Returns SRT_OK or SRT_ERR_SCENE (message in *err) for scenes on which the
// reference itself would crash — nothing is "repaired" silently, with one
// exception that mirrors the reference's own behaviour: an object that has no objColors entry
// gets (0,0,0), which is what objColors[name] default-inserts at simple_raytracer.cpp:439.
template <class OM>
int flatten(OM& om, FlatScene& fs, bool with_flat_lists = false, std::string* err = nullptr) {
    fs = FlatScene();
    detail::Flattener<OM> fl(om, fs);
    int32_t obj = 0;
    for (const auto& pair : om.objTriangles) {              // S2: the map's own iteration order
        const std::string& name = pair.first;
        auto bvh = om.boundingVolumeHierarchy.find(name);
        if (bvh == om.boundingVolumeHierarchy.end() || !bvh->second) {
            if (err) *err = "object '" + name + "' has no bounding hierarchy (null Node* at simple_raytracer.cpp:298)";
            return SRT_ERR_SCENE;
        }
        srt_object o;
        std::memset(&o, 0, sizeof o);
        float c[3] = {0.f, 0.f, 0.f};
        detail::lookup3(om.objColors, name, c);
        o.color[0] = c[0]; o.color[1] = c[1]; o.color[2] = c[2];
        float pr[3];
        if (!detail::lookup3(om.objProperties, name, pr)) {  // :368 would default-insert (0,0,0)
            pr[0] = pr[1] = pr[2] = 0.f;
        }
        o.ambient = pr[0]; o.specular = pr[1]; o.shininess = pr[2];
        float bmin[3] = {0, 0, 0}, bmax[3] = {0, 0, 0};
        detail::lookup3(om.minBox, name, bmin);
        detail::lookup3(om.maxBox, name, bmax);
        for (int k = 0; k < 3; ++k) { o.box_min[k] = bmin[k]; o.box_max[k] = bmax[k]; }
        o.root = (uint32_t)fs.nodes.size();
        o.tri_first = (uint32_t)fs.geom.size();
        if (!fl.walk(bvh->second, obj, 0)) {
            if (err) *err = "object '" + name + "': " + fl.err;
            return SRT_ERR_SCENE;
        }
        o.tri_count = (uint32_t)fs.geom.size() - o.tri_first;
        o.flat_first = SRT_NO_INDEX;
        if (with_flat_lists) {
            o.flat_first = (uint32_t)fs.flat_geom.size();
            for (const auto& t : pair.second) {
                int32_t tex = fl.texture_of(t.textureName);
                if (tex == -2) {
                    if (err) *err = "object '" + name + "': " + fl.err;
                    return SRT_ERR_SCENE;
                }
                fs.flat_geom.emplace_back(); fs.flat_attr.emplace_back(); fs.flat_vnormal.emplace_back();
                detail::fill_triangle(t, obj, tex, fs.flat_geom.back(), fs.flat_attr.back());
                detail::fill_vnormal(t, fs.flat_vnormal.back());
            }
            if (pair.second.size() != o.tri_count) {
                // createBoundingHierarchy copies objTriangles[name] (Object.cpp:277-281); a later
                // transformTriangles without rebuilding would desynchronise the two.
                if (err) *err = "object '" + name + "': BVH holds a different number of triangles than objTriangles";
                return SRT_ERR_SCENE;
            }
        }
        fs.objects.push_back(o);
        fs.object_names.push_back(name);
        ++obj;
    }
    detail::canonical_texture_slots(fs);
    return SRT_OK;
}

}  // namespace srt
#endif  // SRT_FLATTEN_HPP_

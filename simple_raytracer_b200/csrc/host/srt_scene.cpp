// srt_scene.cpp — ObjectManager / Transformation of this repository (see srt_scene.hpp).
#include <memory>
#include <mutex>
#include <cstring>
#include <cstdlib>

#include "srt_scene.hpp"

#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <cmath>
#include <cstdlib>
#include <stdexcept>

namespace srt {

// ------------------------------------------------------------------------------------------------
// Transformation — same matrices as Transformation.cpp:6-90 (hand-filled, column-major m[col][row]).
namespace {
struct Entry { int col, row; float v; };
template <size_t N>
mat4 filled(float diagonal, const Entry (&e)[N]) {
    mat4 m(diagonal);
    for (const Entry& x : e) m[x.col][x.row] = x.v;
    return m;
}
}  // namespace

mat4 Transformation::scaleObj(float sx, float sy, float sz) {
    const Entry e[] = {{0, 0, sx}, {1, 1, sy}, {2, 2, sz}, {3, 3, 1.0f}};
    return filled(0.0f, e);
}
mat4 Transformation::rotateObjX(float a) {          // Transformation.cpp:15-25
    const float c = std::cos(a), s = std::sin(a);
    const Entry e[] = {{0, 0, 1.0f}, {1, 1, c}, {1, 2, -s}, {2, 1, s}, {2, 2, c}, {3, 3, 1.0f}};
    return filled(0.0f, e);
}
mat4 Transformation::rotateObjY(float a) {          // Transformation.cpp:27-36
    const float c = std::cos(a), s = std::sin(a);
    const Entry e[] = {{0, 0, c}, {0, 2, s}, {1, 1, 1.0f}, {2, 0, -s}, {2, 2, c}, {3, 3, 1.0f}};
    return filled(0.0f, e);
}
mat4 Transformation::rotateObjZ(float a) {          // Transformation.cpp:38-47
    const float c = std::cos(a), s = std::sin(a);
    const Entry e[] = {{0, 0, c}, {0, 1, -s}, {1, 0, s}, {1, 1, c}, {2, 2, 1.0f}, {3, 3, 1.0f}};
    return filled(0.0f, e);
}
mat4 Transformation::mirrorObj(bool mx, bool my, bool mz) {
    mat4 m(1.0f);
    if (mx) m[0][0] = -1.0f;
    if (my) m[1][1] = -1.0f;
    if (mz) m[2][2] = -1.0f;
    return m;
}
mat4 Transformation::shearObj(float xy, float xz, float yx, float yz, float zx, float zy) {   // Transformation.cpp:65-76
    const Entry e[] = {{1, 0, xy}, {2, 0, xz}, {0, 1, yx}, {2, 1, yz}, {0, 2, zx}, {1, 2, zy}};
    return filled(1.0f, e);
}
mat4 Transformation::changeObjPosition(vec3 p) {
    mat4 m(1.0f);
    m[3] = vec4(p, 1.0f);
    return m;
}
mat4 Transformation::createViewMatrix(vec3 position, vec3 rotation) {     // T * Rz * Ry * Rx, Transformation.cpp:84-90
    mat4 m = changeObjPosition(position);
    m = m * rotateObjZ(rotation.z);
    m = m * rotateObjY(rotation.y);
    m = m * rotateObjX(rotation.x);
    return m;
}

// ------------------------------------------------------------------------------------------------
namespace {

void free_tree(Node* n) {
    while (n) {                       // iterative on the right spine, recursive on the left
        free_tree(n->left);
        Node* r = n->right;
        delete n;
        n = r;
    }
}

struct Box { vec3 lo, hi; };

inline void grow(Box& b, const vec4& p) {          // glm::min / glm::max per component, Object.cpp:211-218
    b.lo.x = glm_min(b.lo.x, p.x); b.lo.y = glm_min(b.lo.y, p.y); b.lo.z = glm_min(b.lo.z, p.z);
}
inline void grow_hi(Box& b, const vec4& p) {
    b.hi.x = glm_max(b.hi.x, p.x); b.hi.y = glm_max(b.hi.y, p.y); b.hi.z = glm_max(b.hi.z, p.z);
}

// calculateBoundingBoxes, Object.cpp:205-221: raw xyz of the vec4 points (no /w), FLT_MAX seeds.
Box bounds(const std::vector<Triangle>& tris, const uint32_t* order, size_t n) {
    Box b;
    b.lo = vec3(FLT_MAX); b.hi = vec3(-FLT_MAX);
    for (size_t i = 0; i < n; ++i) {
        const Triangle& t = tris[order[i]];
        grow(b, t.pointOne); grow(b, t.pointTwo); grow(b, t.pointThree);
        grow_hi(b, t.pointOne); grow_hi(b, t.pointTwo); grow_hi(b, t.pointThree);
    }
    return b;
}

struct SortKey { float key; uint32_t index; };

// Per-triangle data the build touches at every level, packed (the Triangle objects are 180 bytes apart):
// pointOne.xyz (the sort keys, Object.cpp:193-201) and the triangle's own box.
struct TriLite { float key[3]; float lo[3], hi[3]; };

struct BuildInput {
    const std::vector<Triangle>* tris;
    std::vector<TriLite> lite;
    bool finite = true;        // no NaN / inf coordinate: min / max over per-triangle boxes is order-independent
};

// calculateBoundingBoxes over an index range via the per-triangle boxes.  min / max of finite floats do not depend
// on the evaluation order except for the sign of a zero result (glm::min keeps the first of +0 / -0), so a zero
// component — and any input with non-finite coordinates — takes the reference's own loop.
Box bounds_fast(const BuildInput& in, const uint32_t* order, size_t n) {
    if (!in.finite) return bounds(*in.tris, order, n);
    float lo0 = FLT_MAX, lo1 = FLT_MAX, lo2 = FLT_MAX, hi0 = -FLT_MAX, hi1 = -FLT_MAX, hi2 = -FLT_MAX;
    for (size_t i = 0; i < n; ++i) {
        const TriLite& t = in.lite[order[i]];
        lo0 = t.lo[0] < lo0 ? t.lo[0] : lo0; lo1 = t.lo[1] < lo1 ? t.lo[1] : lo1; lo2 = t.lo[2] < lo2 ? t.lo[2] : lo2;
        hi0 = hi0 < t.hi[0] ? t.hi[0] : hi0; hi1 = hi1 < t.hi[1] ? t.hi[1] : hi1; hi2 = hi2 < t.hi[2] ? t.hi[2] : hi2;
    }
    if (lo0 == 0.0f || lo1 == 0.0f || lo2 == 0.0f || hi0 == 0.0f || hi1 == 0.0f || hi2 == 0.0f) return bounds(*in.tris, order, n);
    Box b;
    b.lo = vec3(lo0, lo1, lo2); b.hi = vec3(hi0, hi1, hi2);
    return b;
}

Node* make_leaf(const std::vector<Triangle>& tris, const uint32_t* order, size_t n, const Box& b) {
    Node* leaf = new Node();
    leaf->minBox = b.lo; leaf->maxBox = b.hi;
    leaf->triangles.reserve(n);
    for (size_t i = 0; i < n; ++i) leaf->triangles.push_back(tris[order[i]]);
    return leaf;
}

// splitTrianglesForBox, Object.cpp:225-272, on an index range instead of Triangle vectors.
// Same tree: the axis rule (:240-248), std::sort on pointOne.axis with '<' (same introsort, same
// comparison outcomes => same permutation), halves n/2 | rest (:254-255), recursion iff size > 8.
Node* split(const BuildInput& in, uint32_t* order, size_t n, const Box& box, std::vector<SortKey>& scratch) {
    const std::vector<Triangle>& tris = *in.tris;
    Node* node = new Node();
    node->minBox = box.lo; node->maxBox = box.hi;
    const float sx = std::fabs(box.hi.x - box.lo.x);
    const float sy = std::fabs(box.hi.y - box.lo.y);
    const float sz = std::fabs(box.hi.z - box.lo.z);
    const int axis = (sx > sy && sx > sz) ? 0 : ((sy > sx && sy > sz) ? 1 : 2);
    scratch.resize(n);
    for (size_t i = 0; i < n; ++i) {
        scratch[i].key = in.lite[order[i]].key[axis];
        scratch[i].index = order[i];
    }
    std::sort(scratch.begin(), scratch.end(), [](const SortKey& a, const SortKey& b) { return a.key < b.key; });
    for (size_t i = 0; i < n; ++i) order[i] = scratch[i].index;
    const size_t nl = n / 2, nr = n - nl;
    const Box bl = bounds_fast(in, order, nl);
    const Box br = bounds_fast(in, order + nl, nr);
    const size_t stop = 8;                                    // triangleSizeStop, Object.cpp:261
    node->left = nl > stop ? split(in, order, nl, bl, scratch) : make_leaf(tris, order, nl, bl);
    node->right = nr > stop ? split(in, order + nl, nr, br, scratch) : make_leaf(tris, order + nl, nr, br);
    return node;
}

}  // namespace

ObjectManager::~ObjectManager() {
    for (auto& kv : boundingVolumeHierarchy) free_tree(kv.second);
    for (auto& kv : textureData) std::free(kv.second);
}

const std::vector<Triangle>& ObjectManager::getTriangles(const std::string& objFilename) const {
    return objTriangles.at(objFilename);
}
void ObjectManager::setTriangles(const std::string& objFilename, const std::vector<Triangle>& triangles) {
    objTriangles[objFilename] = triangles;
}
void ObjectManager::setColor(const std::string& objFilename, const vec3& color) { objColors[objFilename] = color; }
vec3 ObjectManager::getColor(const std::string& objFilename) const { return objColors.at(objFilename); }

void ObjectManager::transformTriangles(const std::string& objFilename, const mat4& matrix) {
    std::vector<Triangle>& triangles = objTriangles[objFilename];   // operator[]: creates the object like the reference
    for (Triangle& t : triangles) {
        t.pointOne = matrix * t.pointOne;
        t.pointTwo = matrix * t.pointTwo;
        t.pointThree = matrix * t.pointThree;
    }
}

void ObjectManager::createBoundingHierarchy(const std::string& objFilename) {
    const std::vector<Triangle>& triangles = objTriangles[objFilename];
    std::vector<uint32_t> order(triangles.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = (uint32_t)i;
    const Box all = bounds(triangles, order.data(), order.size());
    minBox[objFilename] = all.lo;
    maxBox[objFilename] = all.hi;
    std::vector<SortKey> scratch;
    BuildInput in;
    in.tris = &triangles;
    in.lite.resize(triangles.size());
    for (size_t i = 0; i < triangles.size(); ++i) {
        const Triangle& t = triangles[i];
        TriLite& l = in.lite[i];
        const float px[3] = {t.pointOne.x, t.pointTwo.x, t.pointThree.x};
        const float py[3] = {t.pointOne.y, t.pointTwo.y, t.pointThree.y};
        const float pz[3] = {t.pointOne.z, t.pointTwo.z, t.pointThree.z};
        l.key[0] = px[0]; l.key[1] = py[0]; l.key[2] = pz[0];
        l.lo[0] = std::min(px[0], std::min(px[1], px[2])); l.hi[0] = std::max(px[0], std::max(px[1], px[2]));
        l.lo[1] = std::min(py[0], std::min(py[1], py[2])); l.hi[1] = std::max(py[0], std::max(py[1], py[2]));
        l.lo[2] = std::min(pz[0], std::min(pz[1], pz[2])); l.hi[2] = std::max(pz[0], std::max(pz[1], pz[2]));
        for (int k = 0; k < 3; ++k)
            if (!(std::isfinite(px[k]) && std::isfinite(py[k]) && std::isfinite(pz[k]))) in.finite = false;
    }
    Node* root = split(in, order.data(), order.size(), all, scratch);   // the root is always split once (:282)
    Node*& slot = boundingVolumeHierarchy[objFilename];
    free_tree(slot);
    slot = root;
}

// ---- asset cache (not in the reference, which re-parses every OBJ and re-decodes every texture on every
// frame of its loop, simple_raytracer.cpp:537, :554).  Parsing and decoding are deterministic functions of the
// files, so serving the parsed mesh / decoded pixels from memory gives bit-identical scenes.  Off by default.
namespace {
std::mutex g_cache_mutex;
bool g_cache_enabled = false;
struct CachedMesh { ObjMesh mesh; bool ok; };
std::unordered_map<std::string, std::shared_ptr<const CachedMesh>> g_mesh_cache;
struct CachedImage { std::vector<unsigned char> rgb; int w, h; };
std::unordered_map<std::string, std::shared_ptr<const CachedImage>> g_image_cache;

std::shared_ptr<const CachedMesh> cached_read_obj(const std::string& path) {
    {
        std::lock_guard<std::mutex> lock(g_cache_mutex);
        auto it = g_mesh_cache.find(path);
        if (g_cache_enabled && it != g_mesh_cache.end()) return it->second;
    }
    auto m = std::make_shared<CachedMesh>();
    m->ok = read_obj(path, m->mesh);
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    if (g_cache_enabled) g_mesh_cache[path] = m;
    return m;
}

unsigned char* cached_decode(TextureDecoder decode, const std::string& path, int* w, int* h) {
    {
        std::lock_guard<std::mutex> lock(g_cache_mutex);
        auto it = g_image_cache.find(path);
        if (g_cache_enabled && it != g_image_cache.end()) {
            const CachedImage& im = *it->second;
            unsigned char* copy = (unsigned char*)std::malloc(im.rgb.size() ? im.rgb.size() : 1);
            if (!copy) return nullptr;
            std::memcpy(copy, im.rgb.data(), im.rgb.size());
            *w = im.w; *h = im.h;
            return copy;
        }
    }
    unsigned char* data = decode(path.c_str(), w, h);
    if (data && g_cache_enabled) {
        auto im = std::make_shared<CachedImage>();
        im->rgb.assign(data, data + (size_t)(*w) * (size_t)(*h) * 3);
        im->w = *w; im->h = *h;
        std::lock_guard<std::mutex> lock(g_cache_mutex);
        g_image_cache[path] = im;
    }
    return data;
}
}  // namespace

void set_asset_cache(bool enabled) {
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    g_cache_enabled = enabled;
    if (!enabled) { g_mesh_cache.clear(); g_image_cache.clear(); }
}

// loadObjFile, Object.cpp:25-170.
void ObjectManager::loadObjFile(const std::string& objFilename) {
    objColors[objFilename] = vec3(1.f, 0.f, 0.f);                 // :29
    objProperties[objFilename] = vec3(0.2f, 0.5f, 15.0f);         // :31-34
    lastWarning.clear();
    const std::shared_ptr<const CachedMesh> parsed = cached_read_obj(resolvePath(objFilename));
    const ObjMesh& mesh = parsed->mesh;
    if (!parsed->ok) lastWarning = "ObjReader: " + mesh.error;    // the reference prints and continues (:35-39)
    TextureDecoder decode = textureDecoder ? textureDecoder : &decode_image_rgb8;
    for (const ObjMaterial& material : mesh.materials) {          // :52-68
        if (material.diffuse_texname.empty()) continue;
        const std::string& texturePath = material.diffuse_texname;   // as written in the MTL, relative to the CWD
        if (textureData.find(texturePath) != textureData.end()) continue;
        int width = 0, height = 0;
        unsigned char* data = cached_decode(decode, resolvePath(texturePath), &width, &height);
        if (data) {
            textureData[texturePath] = data;
            textureDimensions[texturePath] = ivec2(width, height);
        } else {
            lastWarning += "Failed to load texture: " + texturePath + "\n";
        }
    }
    std::vector<Triangle> triangles;
    const size_t nfaces = mesh.indices.size() / 3;
    triangles.reserve(nfaces);
    for (size_t f = 0; f < nfaces; ++f) {
        Triangle tria;
        for (int v = 0; v < 3; ++v) {
            const ObjIndex idx = mesh.indices[3 * f + v];
            std::string texturePath;
            vec2 texCoordinate(0.f, 0.f);
            vec4 vertex(0.f, 0.f, 0.f, 1.f);
            vec3 vnormal(0.f, 0.f, 0.f);
            vec3 vcolor(1.f, 1.f, 1.f);
            vertex.x = mesh.vertices[3 * (size_t)idx.v + 0];
            vertex.y = mesh.vertices[3 * (size_t)idx.v + 1];
            vertex.z = mesh.vertices[3 * (size_t)idx.v + 2];
            if (idx.vn >= 0) {
                vnormal = vec3(mesh.normals[3 * (size_t)idx.vn], mesh.normals[3 * (size_t)idx.vn + 1],
                               mesh.normals[3 * (size_t)idx.vn + 2]);
            }
            if (idx.vt >= 0) {
                const float tx = mesh.texcoords[2 * (size_t)idx.vt], ty = mesh.texcoords[2 * (size_t)idx.vt + 1];
                const int materialID = mesh.material_ids[f];
                if (materialID >= 0 && materialID < (int)mesh.materials.size()) {
                    texturePath = mesh.materials[(size_t)materialID].diffuse_texname;
                    auto td = textureData.find(texturePath);
                    if (!texturePath.empty() && td != textureData.end()) {
                        const unsigned char* texData = td->second;
                        const ivec2 texDim = textureDimensions[texturePath];
                        if (texDim.x > 0 && texDim.y > 0 && texData) {
                            // texel coordinate with wrap, Object.cpp:113-119
                            int u = static_cast<int>(std::floor(tx * texDim.x)) % texDim.x;
                            int w = static_cast<int>(std::floor((1.0f - ty) * texDim.y)) % texDim.y;
                            u = (u + texDim.x) % texDim.x;
                            w = (w + texDim.y) % texDim.y;
                            texCoordinate = vec2((float)u, (float)w);
                            const size_t texIndex = ((size_t)w * (size_t)texDim.x + (size_t)u) * 3;
                            vcolor = vec3(texData[texIndex] / 255.0f, texData[texIndex + 1] / 255.0f,
                                          texData[texIndex + 2] / 255.0f);
                        }
                    }
                }
            }
            if (v == 0) {
                tria.pointOne = vertex; tria.normalOne = vnormal; tria.colorOneCoordinate = texCoordinate;
                tria.color = vcolor;
                tria.textureName = texturePath;       // set even when the texture failed to load (:149)
            } else if (v == 1) {
                tria.pointTwo = vertex; tria.normalTwo = vnormal; tria.colorTwoCoordinate = texCoordinate;
            } else {
                tria.pointThree = vertex; tria.normalThree = vnormal; tria.colorThreeCoordinate = texCoordinate;
            }
        }
        triangles.push_back(std::move(tria));
    }
    objTriangles[objFilename] = std::move(triangles);             // :169
}

}  // namespace srt

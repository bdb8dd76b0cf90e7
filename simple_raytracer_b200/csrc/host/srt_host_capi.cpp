// srt_host_capi.cpp — C entry points of include/srt_host.h over srt::ObjectManager / Transformation.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "srt_flatten.hpp"
#include "srt_host.h"
#include "srt_scene.hpp"

struct srt_host_scene {
    srt::ObjectManager om;
    srt::FlatScene flat;
    srt_scene_desc desc;
    bool flat_valid = false;
    std::string error;
};

namespace {
srt::mat4 load16(const float* m16) {
    srt::mat4 m;
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) m[c][r] = m16[c * 4 + r];
    return m;
}
void store16(const srt::mat4& m, float* out16) {
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) out16[c * 4 + r] = m[c][r];
}
}  // namespace

extern "C" {

srt_host_scene* srt_host_scene_create(void) { return new srt_host_scene(); }
void srt_host_scene_destroy(srt_host_scene* s) { delete s; }
const char* srt_host_last_error(srt_host_scene* s) { return s ? s->error.c_str() : "null scene"; }
const char* srt_host_last_warning(srt_host_scene* s) { return s ? s->om.lastWarning.c_str() : ""; }

void srt_host_scene_set_base(srt_host_scene* s, const char* dir) {
    if (!s) return;
    s->om.baseDir = dir ? dir : "";
    while (s->om.baseDir.size() > 1 && s->om.baseDir.back() == '/') s->om.baseDir.pop_back();
}

int srt_host_load_obj(srt_host_scene* s, const char* path) {
    s->flat_valid = false;
    s->om.loadObjFile(path);
    return (int)s->om.objTriangles[path].size();
}
void srt_host_set_color(srt_host_scene* s, const char* name, float r, float g, float b) {
    s->om.setColor(name, srt::vec3(r, g, b));
}
void srt_host_set_properties(srt_host_scene* s, const char* name, float a, float sp, float sh) {
    s->om.objProperties[name] = srt::vec3(a, sp, sh);
}
int srt_host_get_properties(srt_host_scene* s, const char* name, float* out3) {
    auto it = s->om.objProperties.find(name);
    if (it == s->om.objProperties.end()) return -1;
    out3[0] = it->second.x; out3[1] = it->second.y; out3[2] = it->second.z;
    return 0;
}
int srt_host_clone(srt_host_scene* s, const char* src, const char* dst, int copy_properties) {
    s->flat_valid = false;
    try {
        s->om.objTriangles[dst] = s->om.getTriangles(src);          // main :565
        if (copy_properties) s->om.objProperties[dst] = s->om.objProperties[src];   // main :566
    } catch (const std::exception& e) { s->error = e.what(); return -1; }
    return 0;
}
void srt_host_transform(srt_host_scene* s, const char* name, const float* m16) {
    s->flat_valid = false;
    s->om.transformTriangles(name, load16(m16));
}
void srt_host_create_bounding_hierarchy(srt_host_scene* s, const char* name) {
    s->flat_valid = false;
    s->om.createBoundingHierarchy(name);
}
int srt_host_object_count(srt_host_scene* s) { return (int)s->om.objTriangles.size(); }
int srt_host_object_name(srt_host_scene* s, int i, char* buf, int buflen) {
    int k = 0;
    for (const auto& pair : s->om.objTriangles) {
        if (k == i) { std::snprintf(buf, (size_t)buflen, "%s", pair.first.c_str()); return (int)pair.first.size(); }
        ++k;
    }
    return -1;
}
int srt_host_triangle_count(srt_host_scene* s, const char* name) {
    auto it = s->om.objTriangles.find(name);
    return it == s->om.objTriangles.end() ? -1 : (int)it->second.size();
}
int srt_host_get_triangles(srt_host_scene* s, const char* name, float* out30, uint8_t* tex_flags) {
    auto it = s->om.objTriangles.find(name);
    if (it == s->om.objTriangles.end()) return -1;
    size_t i = 0;
    for (const srt::Triangle& t : it->second) {
        float* o = out30 + 30 * i;
        const srt::vec4* P[3] = {&t.pointOne, &t.pointTwo, &t.pointThree};
        const srt::vec3* N[3] = {&t.normalOne, &t.normalTwo, &t.normalThree};
        const srt::vec2* C[3] = {&t.colorOneCoordinate, &t.colorTwoCoordinate, &t.colorThreeCoordinate};
        for (int k = 0; k < 3; ++k) { o[4*k] = P[k]->x; o[4*k+1] = P[k]->y; o[4*k+2] = P[k]->z; o[4*k+3] = P[k]->w; }
        for (int k = 0; k < 3; ++k) { o[12+3*k] = N[k]->x; o[12+3*k+1] = N[k]->y; o[12+3*k+2] = N[k]->z; }
        for (int k = 0; k < 3; ++k) { o[21+2*k] = C[k]->x; o[21+2*k+1] = C[k]->y; }
        o[27] = t.color.x; o[28] = t.color.y; o[29] = t.color.z;
        if (tex_flags) tex_flags[i] = t.textureName.empty() ? 0 : 1;
        ++i;
    }
    return (int)i;
}
int srt_host_decode_image(const char* path, int use_sidecar, uint8_t* out, size_t cap, int* width, int* height) {
    int w = 0, h = 0;
    unsigned char* px = srt::decode_image_file(path, &w, &h, use_sidecar != 0);
    if (!px) return -1;
    if (width) *width = w;
    if (height) *height = h;
    int rc = 0;
    if (out) {
        const size_t bytes = (size_t)w * (size_t)h * 3;
        if (cap < bytes) rc = -2; else std::memcpy(out, px, bytes);
    }
    std::free(px);
    return rc;
}

int srt_host_set_texture(srt_host_scene* s, const char* name, const uint8_t* rgb, int width, int height) {
    if (!rgb || width <= 0 || height <= 0) { s->error = "bad texture"; return -1; }
    const size_t bytes = (size_t)width * (size_t)height * 3;
    unsigned char* copy = (unsigned char*)std::malloc(bytes);
    if (!copy) { s->error = "out of memory"; return -1; }
    std::memcpy(copy, rgb, bytes);
    auto it = s->om.textureData.find(name);
    if (it != s->om.textureData.end()) std::free(it->second);
    s->om.textureData[name] = copy;
    s->om.textureDimensions[name] = srt::ivec2(width, height);
    s->flat_valid = false;
    return 0;
}

// Texture slot of every triangle of `name` (slots numbered by first appearance in load order, -1 = untextured);
// the slot names go to names_buf, one per line.  Returns the number of slots, -1 for an unknown object, -2 if
// names_buf is too small.
int srt_host_triangle_textures(srt_host_scene* s, const char* name, int32_t* slot_out, char* names_buf, int buflen) {
    auto it = s->om.objTriangles.find(name);
    if (it == s->om.objTriangles.end()) return -1;
    std::unordered_map<std::string, int32_t> slot;
    std::string names;
    size_t i = 0;
    for (const srt::Triangle& t : it->second) {
        int32_t k = -1;
        if (!t.textureName.empty()) {
            auto f = slot.find(t.textureName);
            if (f == slot.end()) { k = (int32_t)slot.size(); slot.emplace(t.textureName, k); names += t.textureName; names += '\n'; }
            else k = f->second;
        }
        if (slot_out) slot_out[i] = k;
        ++i;
    }
    if (names_buf) {
        if ((int)names.size() + 1 > buflen) return -2;
        std::memcpy(names_buf, names.c_str(), names.size() + 1);
    }
    return (int)slot.size();
}
int srt_host_texture_data(srt_host_scene* s, const char* texture_name, const uint8_t** rgb, int* width, int* height) {
    auto td = s->om.textureData.find(texture_name);
    auto dim = s->om.textureDimensions.find(texture_name);
    if (td == s->om.textureData.end() || !td->second || dim == s->om.textureDimensions.end()) return -1;
    *rgb = td->second; *width = dim->second.x; *height = dim->second.y;
    return 0;
}
// Iteration order of an unordered_map<std::string, ...> whose keys were inserted in the given sequence (S2: the
// object order of simple_raytracer.cpp:409 / :322 is the iteration order of objTriangles).  order_out[k] = index
// into names of the k-th key visited.
int srt_host_object_order(const char* const* names, int n, int* order_out) {
    std::unordered_map<std::string, std::vector<srt::Triangle>> m;     // same key / mapped types as ObjectManager::objTriangles
    std::unordered_map<std::string, int> index;
    for (int i = 0; i < n; ++i) { m[names[i]]; index.emplace(names[i], i); }
    int k = 0;
    for (const auto& pair : m) order_out[k++] = index[pair.first];
    return k;
}

void srt_host_xf_scale(float sx, float sy, float sz, float* o) { store16(srt::Transformation::scaleObj(sx, sy, sz), o); }
void srt_host_xf_rotate_x(float a, float* o) { store16(srt::Transformation::rotateObjX(a), o); }
void srt_host_xf_rotate_y(float a, float* o) { store16(srt::Transformation::rotateObjY(a), o); }
void srt_host_xf_rotate_z(float a, float* o) { store16(srt::Transformation::rotateObjZ(a), o); }
void srt_host_xf_mirror(int mx, int my, int mz, float* o) { store16(srt::Transformation::mirrorObj(mx, my, mz), o); }
void srt_host_xf_shear(const float* s6, float* o) {
    store16(srt::Transformation::shearObj(s6[0], s6[1], s6[2], s6[3], s6[4], s6[5]), o);
}
void srt_host_xf_translate(float x, float y, float z, float* o) {
    store16(srt::Transformation::changeObjPosition(srt::vec3(x, y, z)), o);
}
void srt_host_xf_view(const float* p, const float* r, float* o) {
    store16(srt::Transformation::createViewMatrix(srt::vec3(p[0], p[1], p[2]), srt::vec3(r[0], r[1], r[2])), o);
}
void srt_host_mat_inverse(const float* m16, float* o) { store16(srt::inverse(load16(m16)), o); }
void srt_host_mat_mul(const float* a, const float* b, float* o) { store16(load16(a) * load16(b), o); }
void srt_host_mat_vec(const float* m16, const float* v4, float* out4) {
    const srt::vec4 r = load16(m16) * srt::vec4(v4[0], v4[1], v4[2], v4[3]);
    out4[0] = r.x; out4[1] = r.y; out4[2] = r.z; out4[3] = r.w;
}
float srt_host_radians(float degrees) { return srt::radians(degrees); }

int srt_host_flatten(srt_host_scene* s, int with_flat_lists) {
    std::string err;
    const int rc = srt::flatten(s->om, s->flat, with_flat_lists != 0, &err);
    s->flat_valid = (rc == 0);
    if (rc != 0) s->error = err;
    else s->desc = s->flat.desc();
    return rc;
}
const srt_scene_desc* srt_host_flat_desc(srt_host_scene* s) { return s->flat_valid ? &s->desc : nullptr; }
int srt_host_flat_max_depth(srt_host_scene* s) { return (int)s->flat.max_depth; }
int srt_host_flat_texture_name(srt_host_scene* s, int i, char* buf, int buflen) {
    if (i < 0 || i >= (int)s->flat.texture_names.size()) return -1;
    std::snprintf(buf, (size_t)buflen, "%s", s->flat.texture_names[(size_t)i].c_str());
    return (int)s->flat.texture_names[(size_t)i].size();
}

int srt_host_write_png(const char* path, const uint8_t* rgb, int width, int height) {
    return srt::write_png(path, rgb, width, height) ? 0 : -1;
}

int srt_host_write_bmp(const char* path, const uint8_t* rgb, int width, int height) {
    return srt::write_bmp(path, rgb, width, height) ? 0 : -1;
}

void srt_host_set_asset_cache(int enabled) { srt::set_asset_cache(enabled != 0); }

int srt_host_parse_double(const char* text, double* out) {
    const char* end = text + std::strcspn(text, " \t\r");
    return srt::parse_obj_double(text, end, out) ? 1 : 0;
}

}  // extern "C"

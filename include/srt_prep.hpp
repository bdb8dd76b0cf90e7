// srt_prep.hpp — the per-frame scene prep of the reference's frame loop, on the GPU.
//
// The reference's main() re-loads, re-transforms and re-builds every object for every frame (simple_raytracer.cpp
// :534-618): loadObjFile, a few transformTriangles calls (the camera is baked into the vertices, :558) and
// createBoundingHierarchy per object, all on one host thread.  With this header the meshes are loaded ONCE; per
// frame the maintainer records the same transformTriangles calls and commits them:
//
//     ObjectManager objManager;                       // once, before the frame loop: load, set colours / properties,
//     objManager.loadObjFile("./obj/cube.obj"); ...   //   clone (:565) — everything EXCEPT the per-frame transforms
//     srt::Renderer gpu(0);
//     srt::DevicePrep prep(gpu.context());
//     prep.loadMeshes(objManager);
//     for (int angleDegree = 0; angleDegree < 360; angleDegree += 10) {
//         ...                                         // viewMatrix as before (:546-551)
//         prep.transformTriangles("./obj/cube.obj", Transformation::scaleObj(35, 35, 35));   // was objManager->...
//         prep.transformTriangles("./obj/cube.obj", Transformation::changeObjPosition(glm::vec3(0, 10, 0)));
//         prep.transformTriangles("./obj/cube.obj", glm::inverse(viewMatrix));
//         ...                                         // (no createBoundingHierarchy calls: commit builds every hierarchy)
//         srt_scene* scene = prep.commit(objManager);
//         ImageData points = gpu.renderPrepared<ImageData>(imageSize, lightPos, scene);
//         drawImage(...);
//     }
//
// srt_scene_prep (srt.h) transforms the vertices (one fp32 mat4 * vec4 per recorded call, like Object.cpp:183-190),
// builds the hierarchy of Object.cpp:205-284 (same std::sort permutation, same boxes) and flattens, on the device,
// bit-identical to what srt_flatten.hpp produces from the reference's own objects after the same calls.
// Header-only, duck-typed on the reference's member names (Object.h:15-89) like srt_flatten.hpp.
#ifndef SRT_PREP_HPP_
#define SRT_PREP_HPP_

#include <cstdint>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "srt.h"

namespace srt {

class DevicePrep {
public:
    explicit DevicePrep(srt_context* ctx) : ctx_(ctx) {}
    ~DevicePrep() {
        if (scene_) srt_scene_destroy(scene_);
        for (auto& kv : meshes_) srt_mesh_destroy(kv.second);
    }
    DevicePrep(const DevicePrep&) = delete;
    DevicePrep& operator=(const DevicePrep&) = delete;

    // Uploads objTriangles[name] of every object as it is NOW (object space: call before any per-frame transform)
    // and every texture the triangles name.  Call again after loading or cloning further objects.
    template <class OM>
    void loadMeshes(OM& om) {
        for (const auto& pair : om.objTriangles) {
            const std::string& name = pair.first;
            const auto& tris = pair.second;
            std::vector<float> pts(tris.size() * 12), uv(tris.size() * 6), vn(tris.size() * 9);
            std::vector<int32_t> tex(tris.size(), -1);
            bool textured = false;
            size_t i = 0;
            for (const auto& t : tris) {
                float* p = &pts[12 * i];
                p[0] = t.pointOne.x;   p[1] = t.pointOne.y;   p[2] = t.pointOne.z;    p[3] = t.pointOne.w;
                p[4] = t.pointTwo.x;   p[5] = t.pointTwo.y;   p[6] = t.pointTwo.z;    p[7] = t.pointTwo.w;
                p[8] = t.pointThree.x; p[9] = t.pointThree.y; p[10] = t.pointThree.z; p[11] = t.pointThree.w;
                float* c = &uv[6 * i];
                c[0] = t.colorOneCoordinate.x;   c[1] = t.colorOneCoordinate.y;
                c[2] = t.colorTwoCoordinate.x;   c[3] = t.colorTwoCoordinate.y;
                c[4] = t.colorThreeCoordinate.x; c[5] = t.colorThreeCoordinate.y;
                float* n = &vn[9 * i];
                n[0] = t.normalOne.x;   n[1] = t.normalOne.y;   n[2] = t.normalOne.z;
                n[3] = t.normalTwo.x;   n[4] = t.normalTwo.y;   n[5] = t.normalTwo.z;
                n[6] = t.normalThree.x; n[7] = t.normalThree.y; n[8] = t.normalThree.z;
                if (!t.textureName.empty()) { tex[i] = textureSlot(om, t.textureName); textured = true; }
                ++i;
            }
            srt_mesh_desc d;
            d.points = pts.data(); d.n_tris = tris.size();
            d.uv = textured ? uv.data() : nullptr;
            d.tex = textured ? tex.data() : nullptr;
            d.normals = vn.data();
            srt_mesh* m = nullptr;
            check(srt_mesh_upload(ctx_, &d, &m));
            auto old = meshes_.find(name);
            if (old != meshes_.end()) { srt_mesh_destroy(old->second); old->second = m; }
            else meshes_.emplace(name, m);
        }
    }

    // Records ObjectManager::transformTriangles(name, m) (Object.cpp:183) for the frame being built.
    template <class Mat4>
    void transformTriangles(const std::string& name, const Mat4& m) {
        std::vector<float>& v = matrices_[name];
        for (int c = 0; c < 4; ++c)
            for (int r = 0; r < 4; ++r) v.push_back(m[c][r]);      // glm::mat4 is column-major: m[col][row]
    }

    // Builds the frame's scene on the device from the recorded transforms, in the iteration order of
    // om.objTriangles (S2), with om's colours and properties; clears the recorded transforms.  The returned scene
    // belongs to this object and is overwritten by the next commit.  Asynchronous; status() reports the outcome.
    template <class OM>
    srt_scene* commit(OM& om, bool with_load_order_lists = false) {
        std::vector<srt_prep_object> objs;
        for (const auto& pair : om.objTriangles) {
            const std::string& name = pair.first;
            auto m = meshes_.find(name);
            if (m == meshes_.end()) throw std::runtime_error("srt::DevicePrep: object '" + name + "' was not uploaded (loadMeshes)");
            srt_prep_object o;
            o.mesh = m->second;
            const std::vector<float>& mats = matrices_[name];
            o.matrices = mats.empty() ? nullptr : mats.data();
            o.n_matrices = (uint32_t)(mats.size() / 16);
            o.color[0] = o.color[1] = o.color[2] = 0.f;            // objColors[name] default-inserts (0,0,0), :439
            auto c = om.objColors.find(name);
            if (c != om.objColors.end()) { o.color[0] = c->second.x; o.color[1] = c->second.y; o.color[2] = c->second.z; }
            o.ambient = o.specular = o.shininess = 0.f;            // objProperties[name], :368
            auto p = om.objProperties.find(name);
            if (p != om.objProperties.end()) { o.ambient = p->second.x; o.specular = p->second.y; o.shininess = p->second.z; }
            objs.push_back(o);
        }
        const bool send_textures = scene_ == nullptr || textures_sent_ != textures_.size();
        check(srt_scene_prep(ctx_, objs.data(), (uint32_t)objs.size(), send_textures ? textures_.data() : nullptr,
                             send_textures ? (uint32_t)textures_.size() : 0u, with_load_order_lists ? 1 : 0, &scene_));
        textures_sent_ = textures_.size();
        for (auto& kv : matrices_) kv.second.clear();
        return scene_;
    }

    // Synchronises; throws if the last commit could not be done on the device (non-finite coordinates).
    void status() { check(srt_scene_prep_status(ctx_, scene_)); }
    srt_scene* scene() { return scene_; }
    const std::vector<std::string>& textureNames() const { return texture_names_; }

private:
    template <class OM>
    int32_t textureSlot(OM& om, const std::string& name) {
        auto it = slot_.find(name);
        if (it != slot_.end()) return it->second;
        auto td = om.textureData.find(name);
        auto dim = om.textureDimensions.find(name);
        if (td == om.textureData.end() || !td->second || dim == om.textureDimensions.end() || dim->second.x <= 0 || dim->second.y <= 0)
            throw std::runtime_error("srt::DevicePrep: textured triangle references texture '" + name +
                                     "' that is not loaded (simple_raytracer.cpp:354-358 would dereference it)");
        srt_texture_desc d;
        d.rgb = td->second; d.width = dim->second.x; d.height = dim->second.y;
        const int32_t s = (int32_t)textures_.size();
        textures_.push_back(d);
        texture_names_.push_back(name);
        slot_.emplace(name, s);
        return s;
    }
    static void check(int rc) {
        if (rc != SRT_OK) throw std::runtime_error(std::string("srt: ") + srt_last_error());
    }

    srt_context* ctx_;
    srt_scene* scene_ = nullptr;
    std::unordered_map<std::string, srt_mesh*> meshes_;
    std::unordered_map<std::string, std::vector<float>> matrices_;
    std::unordered_map<std::string, int32_t> slot_;
    std::vector<srt_texture_desc> textures_;
    std::vector<std::string> texture_names_;
    size_t textures_sent_ = 0;
};

}  // namespace srt
#endif  // SRT_PREP_HPP_

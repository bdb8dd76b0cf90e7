// srt_dropin.hpp — the literal drop-in for the reference's one hot-path entry point
//
//     ImageData sendRaysAndIntersectPointsColors(const glm::vec2& imageSize, const glm::vec4& lightPos,
//                                                ObjectManager* objManager);      simple_raytracer.cpp:505
//
// Header-only C++ over the C ABI (srt.h) and the flattener (srt_flatten.hpp).  It is duck-typed on the
// reference's type and member names (ObjectManager / Node / Triangle, Object.h:15-89; ImageData :28;
// glm::vec2 / vec3 / vec4), so it compiles against the reference's own headers when a maintainer adds
// it to the reference, and against this repository's scene mirror (srt_scene.hpp).
//
//     srt::Renderer gpu(0);                                            // once: one B200
//     ImageData points = gpu.sendRaysAndIntersectPointsColors<ImageData>(imageSize, lightPos, &objManager);
//
// Same result contract as the reference: `imagePoints` / `imageColors` hold ONLY the pixels whose
// quantised colour is not (0,0,0) (:518), in the reference's column-major scan order (:511-513), colours
// as floats 0..255 (:447-449).  The literals of the reference (focal 400, one light, +3 staircase, /5,
// 0.5, 1.1) are the defaults of `params` and can be changed before the call.
// Errors: throws std::runtime_error with srt_last_error() — the reference's hot path has no error path;
// the situations reported here are the ones in which the reference would crash or in
// which no CUDA device exists (there is no CPU fallback).
#ifndef SRT_DROPIN_HPP_
#define SRT_DROPIN_HPP_

#include <cstdint>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "srt.h"
#include "srt_flatten.hpp"

namespace srt {

class Renderer {
public:
    srt_render_params params;                 // the reference's literals; edit between calls if wanted

    explicit Renderer(int device = 0) {
        srt_default_params(&params);
        check(srt_context_create(device, &ctx_));
    }
    ~Renderer() {
        if (scene_) srt_scene_destroy(scene_);
        if (ctx_) srt_context_destroy(ctx_);
    }
    Renderer(const Renderer&) = delete;
    Renderer& operator=(const Renderer&) = delete;

    // Flatten + upload `om` (call again after the scene changed; sendRays... does it on every call,
    // like the reference re-reads its maps on every call).
    template <class OM>
    void uploadScene(OM& om, bool with_load_order_lists = false) {
        std::string err;
        FlatScene fs;
        if (flatten(om, fs, with_load_order_lists, &err) != SRT_OK) throw std::runtime_error("srt::flatten: " + err);
        const srt_scene_desc d = fs.desc();
        if (scene_) check(srt_scene_update(ctx_, scene_, &d));
        else check(srt_scene_upload(ctx_, &d, &scene_));
        check(srt_sync(ctx_));                 // `fs` dies at the end of this function
    }

    // Dense render of the uploaded scene: rgb[h][w][3] (sky filled unless params.no_sky_fill).
    void render(std::vector<uint8_t>& rgb, std::vector<int32_t>* obj_id = nullptr, std::vector<int32_t>* tri_id = nullptr,
                std::vector<float>* t = nullptr) {
        if (!scene_) throw std::runtime_error("srt::Renderer: no scene uploaded");
        const size_t n = (size_t)params.width * (size_t)params.height;
        rgb.resize(n * 3);
        srt_frame fr = {rgb.data(), nullptr, nullptr, nullptr};
        if (obj_id) { obj_id->resize(n); fr.obj_id = obj_id->data(); }
        if (tri_id) { tri_id->resize(n); fr.tri_id = tri_id->data(); }
        if (t) { t->resize(n); fr.t = t->data(); }
        check(srt_render(ctx_, scene_, &params, &fr));
        check(srt_sync(ctx_));
    }

    // The reference's signature.  ImageDataT needs `imagePoints` (vector of 2-float vectors) and
    // `imageColors` (vector of 3-float vectors), like the struct at simple_raytracer.cpp:28.
    template <class ImageDataT, class Vec2, class Vec4, class OM>
    ImageDataT sendRaysAndIntersectPointsColors(const Vec2& imageSize, const Vec4& lightPos, OM* objManager) {
        uploadScene(*objManager, params.use_bvh == 0);
        return renderPrepared<ImageDataT>(imageSize, lightPos, scene_);
    }

    // The same call on a scene that is already on the device — uploaded by uploadScene, or built there by
    // srt::DevicePrep::commit (srt_prep.hpp), which replaces the per-frame transformTriangles /
    // createBoundingHierarchy calls of the reference's frame loop.
    template <class ImageDataT, class Vec2, class Vec4>
    ImageDataT renderPrepared(const Vec2& imageSize, const Vec4& lightPos, srt_scene* scene) {
        if (!scene) throw std::runtime_error("srt::Renderer: no scene");
        params.width = (int32_t)imageSize.x;
        params.height = (int32_t)imageSize.y;
        params.light_pos[0] = lightPos.x; params.light_pos[1] = lightPos.y; params.light_pos[2] = lightPos.z;
        ImageDataT out;
        using P = typename decltype(out.imagePoints)::value_type;
        using Cl = typename decltype(out.imageColors)::value_type;
        if (sizeof(P) == 2 * sizeof(float) && sizeof(Cl) == 3 * sizeof(float) && std::is_trivially_copyable<P>::value &&
            std::is_trivially_copyable<Cl>::value) {
            // glm::vec2 / glm::vec3 are packed floats: the device builds both lists (non-black filter :518,
            // column-major order :511-513) and they are copied straight into the vectors' storage.
            uint64_t n = 0;
            check(srt_render_sparse(ctx_, scene, &params, &n));
            out.imagePoints.resize((size_t)n);
            out.imageColors.resize((size_t)n);
            check(srt_fetch_sparse(ctx_, reinterpret_cast<float*>(out.imagePoints.data()),
                                   reinterpret_cast<float*>(out.imageColors.data()), n));
            return out;
        }
        const int32_t keep = params.no_sky_fill;
        params.no_sky_fill = 1;                // (0,0,0) marks "not drawn" exactly like :518
        const size_t npx = (size_t)params.width * (size_t)params.height;
        std::vector<uint8_t> rgb(npx * 3);
        srt_frame fr = {rgb.data(), nullptr, nullptr, nullptr};
        check(srt_render(ctx_, scene, &params, &fr));
        check(srt_sync(ctx_));
        params.no_sky_fill = keep;
        const int W = params.width, H = params.height;
        for (int x = 0; x < W; ++x)            // :511 outer loop over columns
            for (int y = 0; y < H; ++y) {      // :513
                const uint8_t* c = &rgb[((size_t)y * W + x) * 3];
                if (c[0] == 0 && c[1] == 0 && c[2] == 0) continue;          // :518
                out.imagePoints.push_back(P((float)x, (float)y));
                out.imageColors.push_back(Cl((float)c[0], (float)c[1], (float)c[2]));
            }
        return out;
    }

    srt_stats lastStats() {
        srt_stats s;
        check(srt_last_stats(ctx_, &s));
        return s;
    }
    srt_context* context() { return ctx_; }

private:
    srt_context* ctx_ = nullptr;
    srt_scene* scene_ = nullptr;
    static void check(int rc) {
        if (rc != SRT_OK) throw std::runtime_error(std::string("srt: ") + srt_last_error());
    }
};

}  // namespace srt
#endif  // SRT_DROPIN_HPP_

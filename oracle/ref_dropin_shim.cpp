// oracle/ref_dropin_shim.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// include/srt_dropin.hpp and include/srt_prep.hpp compiled against the REFERENCE's own types (ObjectManager, ImageData,
// glm) -> oracle/_ref/libsrt_ref_dropin.so, the only test library that links the product (libsrt_b200.so).  It is loaded
// by the two drop-in tests alone; the oracle / CPU-baseline library libsrt_ref.so carries no product code.
// Scene handles come from libsrt_ref.so (ref_scene_create): the same RefScene type, compiled from the same header.
#include "ref_scene.hpp"

namespace {
glm::mat4 load(const float* m16) {
    glm::mat4 m;
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) m[c][r] = m16[c * 4 + r];
    return m;
}
}  // namespace

// ---- the drop-in adapter (include/srt_dropin.hpp) compiled against the REFERENCE's own types -------
// Test only: proves that srt::Renderer::sendRaysAndIntersectPointsColors is a drop-in for :505 — same
// ImageData type, same sparse column-major contents — by running it on the reference's ObjectManager.
#include "srt_dropin.hpp"
extern "C" long ref_dropin_render(void* h, int W, int Hh, const float* light4, float* points2, float* colors3, long cap,
                                  char* err, int errlen) {
    RefScene* s = (RefScene*)h;
    try {
        static srt::Renderer* gpu = nullptr;       // one context for the test process
        if (!gpu) gpu = new srt::Renderer(0);
        ImageData d = gpu->sendRaysAndIntersectPointsColors<ImageData>(glm::vec2((float)W, (float)Hh),
                                                                      glm::vec4(light4[0], light4[1], light4[2], light4[3]), &s->om);
        const long n = (long)d.imagePoints.size();
        for (long i = 0; i < n && i < cap; ++i) {
            points2[2 * i] = d.imagePoints[i].x; points2[2 * i + 1] = d.imagePoints[i].y;
            colors3[3 * i] = d.imageColors[i].x; colors3[3 * i + 1] = d.imageColors[i].y; colors3[3 * i + 2] = d.imageColors[i].z;
        }
        return n;
    } catch (const std::exception& e) {
        if (err && errlen > 0) std::snprintf(err, (size_t)errlen, "%s", e.what());
        return -1;
    }
}
// ---- the device scene prep (include/srt_prep.hpp) compiled against the REFERENCE's own types ----------------
// Test only.  `h` holds the meshes as LOADED (no transforms).  The recipe is applied twice: to a copy of the
// reference's ObjectManager by the reference's own transformTriangles / createBoundingHierarchy (-> flatten: the
// expected arrays, in `h`), and through srt::DevicePrep on the GPU (-> the arrays written to the out pointers).
#include "srt_prep.hpp"
extern "C" long ref_device_prep(void* h, int n_calls, const char* const* names, const float* m16s, int W, int Hh,
                                const float* light4, float* points2, float* colors3, long cap, srt_node* nodes,
                                srt_tri_geom* geom, srt_tri_attr* attr, srt_object* objects, char* err, int errlen) {
    RefScene* s = (RefScene*)h;
    try {
        static srt::Renderer* gpu = nullptr;
        if (!gpu) gpu = new srt::Renderer(0);
        srt::DevicePrep prep(gpu->context());
        prep.loadMeshes(s->om);
        for (int i = 0; i < n_calls; ++i) prep.transformTriangles(names[i], load(m16s + 16 * i));
        srt_scene* sc = prep.commit(s->om);
        prep.status();
        if (srt_scene_download(gpu->context(), sc, nodes, geom, attr, nullptr, nullptr, objects) != SRT_OK)
            throw std::runtime_error(srt_last_error());
        ImageData d = gpu->renderPrepared<ImageData>(glm::vec2((float)W, (float)Hh), glm::vec4(light4[0], light4[1], light4[2], light4[3]), sc);
        const long n = (long)d.imagePoints.size();
        for (long i = 0; i < n && i < cap; ++i) {
            points2[2 * i] = d.imagePoints[i].x; points2[2 * i + 1] = d.imagePoints[i].y;
            colors3[3 * i] = d.imageColors[i].x; colors3[3 * i + 1] = d.imageColors[i].y; colors3[3 * i + 2] = d.imageColors[i].z;
        }
        return n;
    } catch (const std::exception& e) {
        if (err && errlen > 0) std::snprintf(err, (size_t)errlen, "%s", e.what());
        return -1;
    }
}

// srt_image.cpp — texture pixels for ObjectManager::loadObjFile (the reference calls
// stbi_load(path, &w, &h, &ch, 3) at Object.cpp:57).
//
// This repository carries no JPEG / PNG decoder of its own yet.  Texture pixels come from a
// pre-decoded sidecar file `<texture path>.rgb8` that oracle/stage_assets.py writes next to every
// texture when it stages the reference's mesh / texture data (it decodes with the reference's own
// stb_image through oracle/_ref, so the pixels are bit-identical to what the reference renders
// with — JPEG decoders differ in their IDCT / upsampling, so "some other decoder" would not do).
// Sidecar layout: 4 bytes "SRT8", int32 width, int32 height (little endian), then
// height*width*3 bytes RGB, rows top-down — exactly the buffer stbi_load(..., 3) returns.
// Without a sidecar the function returns nullptr, which loadObjFile reports the way the reference
// reports a failed stbi_load (Object.cpp:63-65: message, continue); a scene that then references
// the texture is rejected by the flattener (SRT_ERR_SCENE) instead of crashing.
// Callers that already hold decoded RGB8 pixels can install ObjectManager::textureDecoder.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "srt_scene.hpp"

namespace srt {

unsigned char* decode_image_rgb8(const char* path, int* width, int* height) {
    if (width) *width = 0;
    if (height) *height = 0;
    if (!path) return nullptr;
    const std::string side = std::string(path) + ".rgb8";
    FILE* f = std::fopen(side.c_str(), "rb");
    if (!f) return nullptr;
    unsigned char hdr[12];
    int32_t w = 0, h = 0;
    unsigned char* data = nullptr;
    if (std::fread(hdr, 1, 12, f) == 12 && std::memcmp(hdr, "SRT8", 4) == 0) {
        std::memcpy(&w, hdr + 4, 4);
        std::memcpy(&h, hdr + 8, 4);
        if (w > 0 && h > 0 && (int64_t)w * h <= (int64_t)1 << 28) {
            const size_t n = (size_t)w * h * 3;
            data = (unsigned char*)std::malloc(n);
            if (data && std::fread(data, 1, n, f) != n) { std::free(data); data = nullptr; }
        }
    }
    std::fclose(f);
    if (!data) return nullptr;
    if (width) *width = w;
    if (height) *height = h;
    return data;
}

}  // namespace srt

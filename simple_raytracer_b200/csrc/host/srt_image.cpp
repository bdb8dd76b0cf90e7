// srt_image.cpp — texture pixels for ObjectManager::loadObjFile (the reference calls
// stbi_load(path, &w, &h, &ch, 3) at Object.cpp:57).
//
// PNG and JPEG files — the two formats the reference's assets use — are decoded by this repository's own decoders
// (srt_png.cpp, srt_jpeg.cpp), which restate the format rules and stb_image's conversions so that the pixels are
// the ones the reference renders with (JPEG decoders differ in their IDCT / upsampling / colour conversion, so "some
// other decoder" would not do).  Any other format, or a file variant those decoders do not cover, is not decoded: the
// function returns nullptr, which loadObjFile reports the way the reference
// reports a failed stbi_load (Object.cpp:63-65: message, continue); a scene that then references
// the texture is rejected by the flattener (SRT_ERR_SCENE) instead of crashing.
// Callers that already hold decoded RGB8 pixels can install ObjectManager::textureDecoder.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "srt_scene.hpp"

namespace srt {

namespace {
// The file through our own decoders; nullptr when it is neither PNG nor JPEG or cannot be decoded.
unsigned char* decode_known_format(const char* path, int* width, int* height) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return nullptr;
    std::vector<uint8_t> file;
    unsigned char chunk[1 << 16];
    for (size_t got; (got = std::fread(chunk, 1, sizeof chunk, f)) > 0;) file.insert(file.end(), chunk, chunk + got);
    std::fclose(f);
    std::vector<uint8_t> rgb;
    int w = 0, h = 0;
    const bool ok = (file.size() > 8 && file[0] == 0x89) ? decode_png(file.data(), file.size(), &w, &h, rgb)
                  : (file.size() > 4 && file[0] == 0xFF && file[1] == 0xD8) ? decode_jpeg(file.data(), file.size(), &w, &h, rgb)
                  : false;
    if (!ok || w <= 0 || h <= 0 || rgb.size() != (size_t)w * h * 3) return nullptr;
    unsigned char* out = (unsigned char*)std::malloc(rgb.size());
    if (!out) return nullptr;
    std::memcpy(out, rgb.data(), rgb.size());
    if (width) *width = w;
    if (height) *height = h;
    return out;
}
}  // namespace

unsigned char* decode_image_rgb8(const char* path, int* width, int* height) { return decode_image_file(path, width, height, true); }

unsigned char* decode_image_file(const char* path, int* width, int* height, bool /*unused: no sidecar fallback any more*/) {
    if (width) *width = 0;
    if (height) *height = 0;
    if (!path) return nullptr;
    return decode_known_format(path, width, height);
}

// drawImage(..., saveImage = true) (simple_raytracer.cpp:461-498) writes the frame with CImg<unsigned char>::save_bmp.
// The same file from a dense, sky-filled RGB8 frame (what srt_render returns): 54-byte header with the values
// CImg writes (file size, data offset 0x36, info-header size 0x28, width, height, 1 plane, 24 bits, image size, and
// its two 0x0100 resolution fields at 0x26 / 0x2A), then the rows bottom-up as B, G, R, each padded to 4 bytes.
bool write_bmp(const char* path, const unsigned char* rgb, int width, int height) {
    if (!path || !rgb || width <= 0 || height <= 0) return false;
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    const uint32_t w = (uint32_t)width, h = (uint32_t)height;
    const uint32_t align = (4u - (3u * w) % 4u) % 4u, row = 3u * w + align, buf_size = row * h, file_size = 54u + buf_size;
    unsigned char header[54] = {0};
    auto put32 = [&](int at, uint32_t v) { for (int k = 0; k < 4; ++k) header[at + k] = (unsigned char)((v >> (8 * k)) & 0xFF); };
    header[0] = 'B'; header[1] = 'M';
    put32(0x02, file_size);
    header[0x0A] = 0x36;
    header[0x0E] = 0x28;
    put32(0x12, w);
    put32(0x16, h);
    header[0x1A] = 1;
    header[0x1C] = 24;
    put32(0x22, buf_size);
    header[0x27] = 0x1;
    header[0x2B] = 0x1;
    bool ok = std::fwrite(header, 1, 54, f) == 54;
    std::string line(row, '\0');
    for (int y = height - 1; y >= 0 && ok; --y) {
        const unsigned char* src = rgb + (size_t)y * w * 3;
        for (uint32_t x = 0; x < w; ++x) { line[3 * x] = (char)src[3 * x + 2]; line[3 * x + 1] = (char)src[3 * x + 1]; line[3 * x + 2] = (char)src[3 * x]; }
        ok = std::fwrite(line.data(), 1, row, f) == row;
    }
    ok = (std::fclose(f) == 0) && ok;
    return ok;
}

bool write_png(const char* path, const unsigned char* rgb, int width, int height) {
    std::vector<uint8_t> file;
    if (!path || !encode_png(rgb, width, height, file)) return false;
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    bool ok = std::fwrite(file.data(), 1, file.size(), f) == file.size();
    ok = (std::fclose(f) == 0) && ok;
    return ok;
}

}  // namespace srt

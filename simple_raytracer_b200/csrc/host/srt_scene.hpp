// srt_scene.hpp — the host-side scene surface of this repository.
//
// It mirrors the reference's scene interface — class, member and method NAMES, argument meaning
// and error behaviour of Object.h:15-89 and Transformation.h:10-20 — so that a scene script
// written against the reference (its main(), simple_raytracer.cpp:542-778) runs unchanged against
// srt::ObjectManager / srt::Transformation, and so that the duck-typed flattener
// (srt_flatten.hpp) accepts either.  The implementation is this repository's own:
//   * OBJ/MTL reading: srt_obj_reader.cpp (restates the parsing rules of tinyobjloader 2.0 that
//     affect the triangles; no third-party code),
//   * BVH build: index-range median split that yields the SAME tree and leaf order as
//     Object.cpp:225-284 without copying Triangle vectors at every level (only leaves own triangles),
//   * matrices: srt_math.hpp (GLM operation order).
// What is deliberately identical to the reference is listed where it matters for parity
// (unordered_map<std::string,...> containers, same insertion sequence => same
// iteration order with the same libstdc++).
#ifndef SRT_SCENE_HPP_
#define SRT_SCENE_HPP_

#include <cstddef>
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

#include "srt_math.hpp"

namespace srt {

class Triangle {                                    // Object.h:15-36
public:
    vec4 pointOne, pointTwo, pointThree;
    vec3 normalOne, normalTwo, normalThree;
    vec2 colorOneCoordinate, colorTwoCoordinate, colorThreeCoordinate;   // texel coordinates, Object.cpp:113-119
    vec3 color;
    std::string textureName;
};

struct Node {                                       // Object.h:46-57
    vec3 minBox, maxBox;
    std::vector<Triangle> triangles;                // filled for LEAVES only (the reference fills every level)
    Node* left = nullptr;
    Node* right = nullptr;
};

// Decoder hook: returns malloc'ed RGB8 (3 channels forced, like stbi_load(..., 3)) or nullptr.
using TextureDecoder = unsigned char* (*)(const char* path, int* width, int* height);

class ObjectManager {                               // Object.h:59-89
public:
    std::unordered_map<std::string, vec3> minBox;
    std::unordered_map<std::string, vec3> maxBox;
    std::unordered_map<std::string, vec3> objProperties;          // ambient, specular, shininess
    std::unordered_map<std::string, std::vector<Triangle>> objTriangles;
    std::unordered_map<std::string, Node*> boundingVolumeHierarchy;
    std::unordered_map<std::string, vec3> objColors;
    std::unordered_map<std::string, unsigned char*> textureData;
    std::unordered_map<std::string, ivec2> textureDimensions;

    ObjectManager() = default;
    ~ObjectManager();                               // frees Node trees and texture buffers (the reference leaks both)
    ObjectManager(const ObjectManager&) = delete;
    ObjectManager& operator=(const ObjectManager&) = delete;

    void loadObjFile(const std::string& objFilename);                               // Object.cpp:25
    const std::vector<Triangle>& getTriangles(const std::string& objFilename) const;   // throws std::out_of_range, Object.cpp:174
    void setTriangles(const std::string& objFilename, const std::vector<Triangle>& triangles);
    void transformTriangles(const std::string& objFilename, const mat4& matrix);    // Object.cpp:183
    void createBoundingHierarchy(const std::string& objFilename);                   // Object.cpp:275
    void setColor(const std::string& objFilename, const vec3& color);
    vec3 getColor(const std::string& objFilename) const;                            // throws std::out_of_range, Object.cpp:292

    // Not in the reference: where texture files are decoded.  Default: srt::decode_image_rgb8.
    TextureDecoder textureDecoder = nullptr;
    // Not in the reference (which resolves every OBJ / MTL / texture path against the process CWD, Object.cpp:35, :57):
    // a directory that relative paths are resolved against instead.  Map keys stay the names as written.  Lets a
    // multi-threaded host (the series driver's prep pool) load assets without touching the process-wide CWD.
    std::string baseDir;
    std::string resolvePath(const std::string& path) const {
        if (baseDir.empty() || path.empty() || path[0] == '/') return path;
        return baseDir + "/" + path;
    }
    // Not in the reference: last loader diagnostics (the reference prints them to std::cerr and continues).
    std::string lastWarning;
};

class Transformation {                              // Transformation.h:10-20
public:
    static mat4 scaleObj(float sx, float sy, float sz);
    static mat4 rotateObjX(float degree);
    static mat4 rotateObjY(float degree);
    static mat4 rotateObjZ(float degree);
    static mat4 mirrorObj(bool mirrorX = false, bool mirrorY = false, bool mirrorZ = false);
    static mat4 shearObj(float shearXY = 0.0f, float shearXZ = 0.0f, float shearYX = 0.0f, float shearYZ = 0.0f,
                         float shearZX = 0.0f, float shearZY = 0.0f);
    static mat4 changeObjPosition(vec3 position);
    static mat4 createViewMatrix(vec3 position, vec3 rotation);
};

// --- OBJ / MTL reading (srt_obj_reader.cpp) ---------------------------------------------------
struct ObjIndex { int v = -1, vn = -1, vt = -1; };
struct ObjMaterial { std::string name, diffuse_texname; };
struct ObjMesh {
    std::vector<float> vertices, normals, texcoords;       // xyz / xyz / uv, as parsed
    std::vector<ObjIndex> indices;                          // 3 per output triangle (after triangulation)
    std::vector<int> material_ids;                          // per output triangle
    std::vector<ObjMaterial> materials;
    std::string warning, error;
};
// Reads `path` (materials are searched next to it).  Returns false on a parse error (the mesh
// holds whatever was read up to then, like tinyobj::ObjReader).
bool read_obj(const std::string& path, ObjMesh& out);
// The float parser used for every number in OBJ/MTL files (restates tinyobjloader's tryParseDouble).
bool parse_obj_double(const char* s, const char* s_end, double* result);

// Process-wide cache of parsed OBJ/MTL files and decoded textures (thread-safe; off by default; disabling clears it).
void set_asset_cache(bool enabled);

// --- image decoding (srt_image.cpp) -------------------------------------------------------------
unsigned char* decode_image_rgb8(const char* path, int* width, int* height);
unsigned char* decode_image_file(const char* path, int* width, int* height, bool use_sidecar);
// Decoders of our own for the two formats the reference's assets use, to the pixels stbi_load(..., 3) returns
// (srt_png.cpp, srt_jpeg.cpp).  false: not that format, damaged, or a variant the decoder does not cover.
bool decode_png(const uint8_t* data, size_t n, int* width, int* height, std::vector<uint8_t>& rgb);
bool decode_jpeg(const uint8_t* data, size_t n, int* width, int* height, std::vector<uint8_t>& rgb);
// drawImage's saved file (:461-498, CImg save_bmp) from a dense sky-filled RGB8 frame; false on I/O failure.
bool write_bmp(const char* path, const unsigned char* rgb, int width, int height);
// Not in the reference (it saves BMP only): the same frame as an 8-bit RGB PNG (srt_png.cpp).
bool encode_png(const uint8_t* rgb, int width, int height, std::vector<uint8_t>& file);
bool write_png(const char* path, const unsigned char* rgb, int width, int height);

}  // namespace srt
#endif

// srt_image.cpp — texture file decoding for ObjectManager::loadObjFile.
//
// STATUS: not implemented yet — returns nullptr, which loadObjFile reports exactly like the
// reference reports a failed stbi_load (Object.cpp:63-65: message, continue); a scene that then
// references the texture is rejected by the flattener (SRT_ERR_SCENE) instead of crashing.
// Callers that already hold decoded RGB8 pixels can install ObjectManager::textureDecoder.
#include "srt_scene.hpp"

namespace srt {

unsigned char* decode_image_rgb8(const char* path, int* width, int* height) {
    (void)path;
    if (width) *width = 0;
    if (height) *height = 0;
    return nullptr;
}

}  // namespace srt

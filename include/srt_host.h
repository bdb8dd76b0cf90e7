/*
 * srt_host.h — C entry points of the host-side scene surface (srt::ObjectManager /
 * srt::Transformation, simple_raytracer_b200/csrc/host/srt_scene.hpp) for callers that are not
 * C++ (this repository's Python plumbing, tests, bench).  A C++ host uses the classes directly.
 *
 * Every function mirrors one call of the reference's scene surface; the cited lines are the
 * reference's (Object.cpp / Transformation.cpp / simple_raytracer.cpp main()).  Matrices are 16
 * floats, column-major like glm::mat4.  No rendering happens here — see srt.h.
 */
#ifndef SRT_HOST_H_
#define SRT_HOST_H_

#include <stdint.h>
#include "srt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct srt_host_scene srt_host_scene;        /* owns one ObjectManager (main :537) */

srt_host_scene* srt_host_scene_create(void);
void srt_host_scene_destroy(srt_host_scene* s);
const char* srt_host_last_error(srt_host_scene* s);
const char* srt_host_last_warning(srt_host_scene* s);   /* what the reference prints to std::cerr */

/* Directory that relative OBJ / MTL / texture paths are resolved against (NULL or "": the process CWD, like the
 * reference, Object.cpp:35, :57).  Per scene, so host threads can load assets concurrently without chdir. */
void srt_host_scene_set_base(srt_host_scene* s, const char* dir);

/* ObjectManager (Object.h:59-89) */
int  srt_host_load_obj(srt_host_scene* s, const char* path);                          /* loadObjFile; returns #triangles */
void srt_host_set_color(srt_host_scene* s, const char* name, float r, float g, float b);   /* setColor */
void srt_host_set_properties(srt_host_scene* s, const char* name, float ambient, float specular, float shininess);
int  srt_host_get_properties(srt_host_scene* s, const char* name, float* out3);
int  srt_host_clone(srt_host_scene* s, const char* src, const char* dst, int copy_properties);   /* main :565-566 */
void srt_host_transform(srt_host_scene* s, const char* name, const float* m16);       /* transformTriangles */
void srt_host_create_bounding_hierarchy(srt_host_scene* s, const char* name);         /* createBoundingHierarchy */
int  srt_host_object_count(srt_host_scene* s);
int  srt_host_object_name(srt_host_scene* s, int i, char* buf, int buflen);           /* objTriangles iteration order */
int  srt_host_triangle_count(srt_host_scene* s, const char* name);
/* 30 floats per triangle: 3 x vec4 points, 3 x vec3 normals, 3 x vec2 texel coords, vec3 colour */
int  srt_host_get_triangles(srt_host_scene* s, const char* name, float* out30, uint8_t* tex_flags);
/* Install already-decoded RGB8 pixels for a texture name (copied). */
int  srt_host_set_texture(srt_host_scene* s, const char* name, const uint8_t* rgb, int width, int height);

/* For the device scene prep (srt_scene_prep): texture slot of every triangle of `name` (slots numbered by first
 * appearance in load order, -1 = untextured), slot names one per line in names_buf; returns the slot count, -1 for
 * an unknown object, -2 when names_buf is too small. */
int  srt_host_triangle_textures(srt_host_scene* s, const char* name, int32_t* slot_out, char* names_buf, int buflen);
int  srt_host_texture_data(srt_host_scene* s, const char* texture_name, const uint8_t** rgb, int* width, int* height);
/* A texture file to the RGB8 pixels stbi_load(path, &w, &h, &ch, 3) (Object.cpp:57) returns: PNG / JPEG through this
 * library's own decoders; other formats are not decoded (use_sidecar is ignored — the pre-decoded sidecar fallback of
 * round 1 is gone, the product opens nothing the reference's code produced).  out == NULL: sizes only.  0, -1 (cannot decode) or -2 (cap too small). */
int  srt_host_decode_image(const char* path, int use_sidecar, uint8_t* out, size_t cap, int* width, int* height);
/* Iteration order of ObjectManager::objTriangles for keys inserted in the given sequence (the object order of
 * simple_raytracer.cpp:409 / :322): order_out[k] = index into names of the k-th key visited; duplicates count once. */
int  srt_host_object_order(const char* const* names, int n, int* order_out);

/* Transformation (Transformation.h:10-20) and the glm calls of main() */
void srt_host_xf_scale(float sx, float sy, float sz, float* out16);
void srt_host_xf_rotate_x(float radians, float* out16);
void srt_host_xf_rotate_y(float radians, float* out16);
void srt_host_xf_rotate_z(float radians, float* out16);
void srt_host_xf_mirror(int mx, int my, int mz, float* out16);
void srt_host_xf_shear(const float* s6, float* out16);
void srt_host_xf_translate(float x, float y, float z, float* out16);
void srt_host_xf_view(const float* pos3, const float* rot3, float* out16);            /* createViewMatrix */
void srt_host_mat_inverse(const float* m16, float* out16);                            /* glm::inverse */
void srt_host_mat_mul(const float* a16, const float* b16, float* out16);
void srt_host_mat_vec(const float* m16, const float* v4, float* out4);
float srt_host_radians(float degrees);                                                /* glm::radians */

/* Flatten (srt_flatten.hpp) into arrays owned by the scene handle; valid until the next
 * flatten / mutation / destroy.  The returned descriptor can go straight into srt_scene_upload. */
int  srt_host_flatten(srt_host_scene* s, int with_flat_lists);
const srt_scene_desc* srt_host_flat_desc(srt_host_scene* s);
int  srt_host_flat_max_depth(srt_host_scene* s);
int  srt_host_flat_texture_name(srt_host_scene* s, int i, char* buf, int buflen);

/* Process-wide cache of parsed OBJ/MTL files and decoded textures: the reference re-reads them on every frame
 * of its loop (simple_raytracer.cpp:537, :554); loading is deterministic, so cached loads are bit-identical.
 * Thread-safe; off by default; disabling clears the cache. */
/* drawImage(..., saveImage = true) :461-498: the BMP CImg writes, from a dense sky-filled RGB8 frame (srt_render's
 * output); byte-identical to the reference's file.  0 on success, -1 on I/O failure. */
int  srt_host_write_bmp(const char* path, const uint8_t* rgb, int width, int height);
/* Not a reference format: the same dense RGB8 frame as a PNG file (8-bit RGB, lossless). */
int  srt_host_write_png(const char* path, const uint8_t* rgb, int width, int height);
void srt_host_set_asset_cache(int enabled);

/* OBJ number parser (for tests): returns 1 and the value, or 0. */
int  srt_host_parse_double(const char* text, double* out);

#ifdef __cplusplus
}
#endif
#endif

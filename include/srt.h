/*
 * srt.h — C ABI of the B200-native render core for leonlang/simple_raytracer scenes.
 *
 * This is the drop-in boundary for ONE path of the reference: the per-pixel render loop
 *
 *     ImageData sendRaysAndIntersectPointsColors(const glm::vec2& imageSize,
 *                                                const glm::vec4& lightPos,
 *                                                ObjectManager* objManager);
 *                                      (reference simple_raytracer.cpp:505-525, caller :784)
 *
 * and everything it calls (rayIntersection :405-457, boundingBoxIntersection :296-317,
 * intersectRayAabbNoOrigin :252-293, intersectRayAabb :204-248, rayTriangleIntersection :42-75,
 * softShadow :348-401, shadowIntersection :321-342, phongIllumination :144-200,
 * calculateBarycentricCoords :79-117, getTextureCoordinate :121-128) plus the image contract of
 * drawImage :461-498 (8-bit conversion and sky fill).
 *
 * The host keeps the reference's scene objects (ObjectManager / Triangle / Node, Object.h:15-89).
 * `srt_flatten.hpp` turns them into the plain arrays below; the functions here upload those
 * arrays to one B200 and run the persistent sm_100a render kernel.  Plain C, POD only, no C++
 * or torch types cross this boundary.  Every function returns 0 on success or a negative
 * srt_status; srt_last_error() gives the message (thread-local).
 *
 * There is NO CPU fallback: without a CUDA device every compute entry point fails with
 * SRT_ERR_NO_DEVICE.
 */
#ifndef SRT_H_
#define SRT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRT_ABI_VERSION 1
#define SRT_MAX_LIGHTS 64          /* softShadow(lightAmount, …) :348; the reference ships 1, used 8/16/32 */
#define SRT_LEAF_BIT 0x80000000u
#define SRT_NO_INDEX 0xFFFFFFFFu

typedef enum srt_status {
    SRT_OK = 0,
    SRT_ERR_INVALID_ARG = -1,
    SRT_ERR_NO_DEVICE = -2,
    SRT_ERR_CUDA = -3,
    SRT_ERR_OOM = -4,
    SRT_ERR_SCENE = -5,            /* scene violates a precondition the reference would crash on */
    SRT_ERR_UNSUPPORTED = -6
} srt_status;

/* ---------------------------------------------------------------------------------------------
 * Flattened scene (what replaces the pointer-linked Node tree, Object.h:46-57).
 *
 * Nodes of one object are stored in depth-first pre-order: the LEFT child of interior node i is
 * node i+1, `a` holds the index of the RIGHT child.  A leaf has SRT_LEAF_BIT set in `b`; then
 * `a` = first triangle (index into geom/attr), `b & ~SRT_LEAF_BIT` = triangle count (may be 0).
 * Boxes are Node::minBox / Node::maxBox verbatim (computed from the un-divided xyz of the vec4
 * points, Object.cpp:205-221).
 * ------------------------------------------------------------------------------------------- */
typedef struct srt_node {
    float    bmin[3];
    uint32_t a;
    float    bmax[3];
    uint32_t b;
} srt_node;                                   /* 32 B: two 16-byte loads */

/* Triangle geometry in candidate (leaf DFS) order — the order boundingBoxIntersection :296-317
 * concatenates leaves in, so `index in this array` == the reference's visit index within the
 * object.  p0 = vec3(pointOne)/pointOne.w, e1 = P2/w2 - P1/w1, e2 = P3/w3 - P1/w1: the first six
 * operations of rayTriangleIntersection :45-51 and calculateBarycentricCoords :81-90, done once
 * on the host with the same fp32 operations. */
typedef struct srt_tri_geom {
    float    p0[3];
    int32_t  obj;                             /* owning object (index into objects[]) */
    float    e1[3];
    int32_t  tex;                             /* texture index or -1 (Triangle::textureName empty) */
    float    e2[3];
    uint32_t reserved;
} srt_tri_geom;                               /* 48 B: three 16-byte loads */

/* Shading attributes, same order as geom.  n = calculateTriangleNormal :32-37 (flat normal from
 * the UN-divided xyz), uv = colorOne/Two/ThreeCoordinate (pixel-space, Object.cpp:113-119). */
typedef struct srt_tri_attr {
    float n[3];
    float pad0;
    float uv[3][2];
    float pad1[2];
} srt_tri_attr;                               /* 48 B */

/* Optional, same order as geom: the per-vertex normals Triangle::normalOne/Two/Three exactly as loadObjFile stored
 * them (Object.cpp:98-104) — the reference never transforms them (transformTriangles touches the points only).
 * Needed only for srt_render_params::smooth_normals (interpolateNormal :132-140). */
typedef struct srt_tri_vnormal {
    float n[3][4];                            /* xyz of each vertex normal, w unused (0) */
} srt_tri_vnormal;                            /* 48 B: three 16-byte loads */

/* One entry per key of ObjectManager::objTriangles, in that unordered_map's ITERATION order
 * (simple_raytracer.cpp:409, :322) — the order decides exact-tie resolution. */
typedef struct srt_object {
    float    color[3];                        /* objColors[name]                    Object.h:68 */
    float    ambient, specular, shininess;    /* objProperties[name]                Object.h:64 */
    uint32_t root;                            /* root node index                                 */
    uint32_t tri_first, tri_count;            /* range in geom/attr (leaf order)                 */
    uint32_t flat_first;                      /* range start in flat_geom (load order) or SRT_NO_INDEX */
    float    box_min[3];                      /* minBox[name] / maxBox[name]        Object.h:61-62 */
    float    box_max[3];
} srt_object;                                 /* 64 B */

/* RGB8 texture exactly as stbi_load(..., 3) returned it (Object.cpp:57-61). */
typedef struct srt_texture_desc {
    const uint8_t* rgb;
    int32_t width, height;
} srt_texture_desc;

typedef struct srt_scene_desc {
    const srt_node*     nodes;     uint64_t n_nodes;
    const srt_tri_geom* geom;      /* leaf order */
    const srt_tri_attr* attr;      uint64_t n_tris;
    /* Optional (may be NULL / 0): triangles in objTriangles[name] LOAD order, for the legacy
     * one-box-per-object path (simple_raytracer.cpp:415-420). */
    const srt_tri_geom* flat_geom;
    const srt_tri_attr* flat_attr; uint64_t n_flat;
    const srt_object*   objects;   uint32_t n_objects;
    const srt_texture_desc* textures; uint32_t n_textures;
    /* Optional (NULL: the scene cannot be rendered with smooth_normals): vertex normals in leaf order / load order. */
    const srt_tri_vnormal* vnormal;
    const srt_tri_vnormal* flat_vnormal;
} srt_scene_desc;

/* Every literal of the reference's hot path, as a parameter (defaults = the literals). */
typedef struct srt_render_params {
    int32_t width, height;        /* imageSize (600,400)                         :773 */
    float   focal;                /* ray.direction.z = 400.0f                    :506 */
    float   light_pos[3];         /* view-space light, lightPos.xyz              :778 */
    int32_t light_count;          /* softShadow(1, …)                            :445 */
    float   light_step;           /* += 3.0f staircase                           :372-382 */
    float   light_color[3];       /* (1,1,1)                                     :433 */
    float   shadow_div;           /* colorPhong /= 5                             :369 */
    float   tone_k;               /* color / (color + 0.5f)                      :391 */
    float   gamma;                /* pow(color, 1.1f)                            :396 */
    uint8_t sky[3];               /* (173,216,230)                               :476 */
    uint8_t pad0;
    int32_t use_bvh;              /* 1: BVH path :422; 0: legacy one-box path :417 (needs flat_geom) */
    int32_t row_begin, row_end;   /* render rows [row_begin,row_end); 0,0 = all rows (tile-row shard) */
    int32_t variant;              /* SRT_VARIANT_* kernel selection, 0 = default */
    int32_t no_sky_fill;          /* 1: leave miss / black-hit pixels (0,0,0) (rebuilding ImageData :518) */
    /* Interleaved row-block sharding of ONE frame over several GPUs: of the blocks of `row_block` rows
     * (a multiple of 4; 0 = no sharding) that make up [row_begin,row_end), render blocks shard_index,
     * shard_index + shard_count, ...  Pixels are independent (:511-523), so the union of the shards is
     * byte-identical to one full render. */
    int32_t row_block, shard_index, shard_count;
    /* Where the light_count light samples sit (NOT a parity mode unless 0):
     * SRT_LIGHTS_STAIRCASE  the reference's deterministic sequence (:363-383), the default;
     * SRT_LIGHTS_JITTERED   sample i = light_pos + extent * (u_x, u_y, u_z), extent = light_step * ceil(light_count / 3)
     *                       (the box the staircase spans), u in [0,1) from a counter-based generator keyed by
     *                       (light_seed, 3 i + axis): reproducible, independent of evaluation order. */
    int32_t  light_mode;
    uint32_t light_seed;
    /* 1: phongIllumination with the normal the reference ships commented out (:162): n = interpolateNormal(triangle,
     * barycentricCoords) :132-140 — normalize(b.x * normalOne + b.y * normalTwo + b.z * normalThree) — instead of the
     * flat calculateTriangleNormal (:164).  Needs a scene with vertex normals.  0 (default): the shipped behaviour. */
    int32_t  smooth_normals;
    /* Quality switches the reference lacks.  All 0 = the shipped behaviour; none is a parity mode
     * (the oracle restates them, tests compare against that):
     * normalise_lights  1: the sum over the light samples (:370) is divided by light_count before tone mapping.
     * shadow_segment    1: an occluder counts only where it lies between the point and the light sample,
     *                   shadow_eps <= t <= 1 along shadowRay = light - point (:325-327); 0: any t >= 0 (an object
     *                   BEHIND the light darkens the point, as shipped).
     * texture_mode      how the interpolated texel coordinate (x, y) (:123-125; the scene carries the reference's
     *                   per-vertex TEXEL coordinates, Object.cpp:113-119) addresses the image:
     *                   SRT_TEX_TRUNCATE  texel ((int)x, (int)y) through the linear index of :357, index clamped to the image
     *                                     where the reference would read out of bounds;
     *                   SRT_TEX_NEAREST   texel (floor x mod width, floor y mod height): each axis wraps on its own;
     *                   SRT_TEX_BILINEAR  blend of the four texels around (x, y) by its fractional parts, same wrapping. */
    int32_t  normalise_lights;
    int32_t  shadow_segment;
    float    shadow_eps;
    int32_t  texture_mode;
} srt_render_params;
enum { SRT_LIGHTS_STAIRCASE = 0, SRT_LIGHTS_JITTERED = 1 };
enum { SRT_TEX_TRUNCATE = 0, SRT_TEX_NEAREST = 1, SRT_TEX_BILINEAR = 2 };

enum { SRT_VARIANT_DEFAULT = 0,   /* staged pipeline: classify, primary, shadow, shade as stream-ordered launches; falls back to
                                     SIMPLE for the legacy one-box path */
       SRT_VARIANT_SIMPLE = 1,    /* one thread = one pixel, whole pixel per work item */
       SRT_VARIANT_FUSED = 2,     /* the stages fused into one persistent dataflow kernel */
       SRT_VARIANT_OVERLAP = 4,   /* DEFAULT with the shadow pass overlapping the primary pass's tail (programmatic
                                     dependent launch, passes ordered by published tile entries) */
       SRT_VARIANT_COUNT = 100,   /* SIMPLE + node / triangle counters */
       SRT_VARIANT_TIMELINE = 101 /* SIMPLE + per-tile cost map in the id / t planes (debug) */ };

/* Output planes; each pointer may be NULL (plane not wanted), a DEVICE pointer (written in place
 * on the context stream, no copy) or a HOST pointer (rendered to an internal device buffer and
 * copied back on the context stream).  Full-frame addressing: row r starts at r*width.
 * rgb is interleaved RGB8, row-major, top row first, sky already filled (drawImage :476-487). */
typedef struct srt_frame {
    uint8_t* rgb;                 /* [height][width][3] */
    int32_t* obj_id;              /* [height][width], -1 = miss                   */
    int32_t* tri_id;              /* [height][width], global leaf-order (or flat) index, -1 = miss */
    float*   t;                   /* [height][width], hit parameter, +inf = miss  */
} srt_frame;

typedef struct srt_stats {
    uint64_t primary_rays;        /* rows*width */
    uint64_t hit_pixels;
    uint64_t shadow_rays;         /* light_count*hit_pixels if the scene has >= 2 objects else 0 */
    uint64_t node_visits;         /* only counted by the instrumented variant, else 0 */
    uint64_t tri_tests;
    uint64_t exact_slab_tests;    /* phased kernel: slab tests the filter could not decide (ran the exact divisions) */
    float    kernel_ms;           /* CUDA-event time of the render kernel(s) */
    int32_t  kernel_launches;     /* kernels launched by the last srt_render */
} srt_stats;

typedef struct srt_context srt_context;
typedef struct srt_scene   srt_scene;

/* ---- lifetime ---- */
int  srt_abi_version(void);
int  srt_device_count(int* count);
int  srt_context_create(int device, srt_context** out);
void srt_context_destroy(srt_context* ctx);
/* Use an existing cudaStream_t (e.g. torch's current stream) instead of the context's own. */
int  srt_context_set_stream(srt_context* ctx, void* cuda_stream);
/* Row bands per frame of the default pipeline: 2 (default) renders the upper and the lower half of the rows on two
 * streams so that their stages overlap (-11 % time per frame; the same bytes, pixels are independent :511-523) — what a
 * caller that renders ONE frame at a time wants; 1 renders a frame in one piece, for callers that already keep several
 * frames in flight on several contexts (an image series gains nothing from a second band; the series driver sets 1 on
 * its lanes).  3 and 4 are accepted (bands of at least 64 rows) and measured no better than 2 (profiles/).  A context's
 * share of fewer than 128 rows and the other kernel variants are always one piece; a sharded frame (row_block /
 * shard_count) is cut at a multiple of row_block * shard_count, so both bands keep the frame's block ownership.
 * Returns the previous value, or a negative srt_status. */
int  srt_context_set_bands(srt_context* ctx, int bands);
int  srt_context_get_info(srt_context* ctx, int* sm_count, int* cc_major, int* cc_minor,
                          size_t* l2_bytes);

/* ---- scene ---- */
/* Validates (returns SRT_ERR_SCENE where the reference would crash: textured triangle without
 * texture, node/triangle index out of range, object without BVH …), allocates device memory and
 * copies.  Synchronous with respect to the host arrays (they may be freed on return). */
int  srt_scene_upload(srt_context* ctx, const srt_scene_desc* desc, srt_scene** out);
/* Re-upload new contents into an existing scene object (per-frame scenes of an image series);
 * device buffers are reused when large enough.  The copy is asynchronous on the context stream
 * when the host arrays are pinned (srt_host_alloc).
 * Textures: desc->textures == NULL keeps the resident table (n_textures must equal the resident count).  Otherwise the
 * table of `desc` becomes the scene's table: every slot is compared with the resident one (host pointer, size, a sampled
 * fingerprint of the texels) and only slots that differ are copied again — the triangles' `tex` indices of `desc` always
 * refer to the table passed with them.  srt_flatten.hpp numbers textures by name, so the slots of a series do not move. */
int  srt_scene_update(srt_context* ctx, srt_scene* scene, const srt_scene_desc* desc);
void srt_scene_destroy(srt_scene* scene);

/* ---- device scene prep: the per-frame host work of the reference's frame loop, on the device ----
 * The reference re-transforms every object and rebuilds its bounding hierarchy on the host for every frame
 * (main :553-618 -> ObjectManager::transformTriangles Object.cpp:183-190, createBoundingHierarchy Object.cpp:275-284
 * -> splitTrianglesForBox :225-272, calculateBoundingBoxes :205-221).  These entry points keep the object-space
 * mesh resident on the GPU and produce, from a list of matrices per object, exactly the arrays srt_flatten.hpp
 * would produce from the reference's ObjectManager after the same calls: the same vertices (one fp32 mat4 * vec4 per
 * transformTriangles call), the same tree (libstdc++'s std::sort permutation is reproduced, ties included), the same
 * boxes, the same leaf order — bit for bit.  Only the texture slot numbers are the caller's (srt_mesh_desc::tex).
 * Coordinates must be finite (srt_scene_prep_status reports SRT_ERR_UNSUPPORTED otherwise). */
typedef struct srt_mesh srt_mesh;
typedef struct srt_mesh_desc {
    const float*   points;        /* [n_tris][3][4]  pointOne / pointTwo / pointThree (vec4) in objTriangles load order */
    const float*   uv;            /* [n_tris][3][2]  colorOne/Two/ThreeCoordinate (texel space), or NULL = zeros      */
    const int32_t* tex;           /* [n_tris]        texture slot of Triangle::textureName or -1; NULL = all -1       */
    uint64_t       n_tris;
    const float*   normals;       /* [n_tris][3][3]  normalOne / normalTwo / normalThree, or NULL (no smooth_normals)   */
} srt_mesh_desc;
int  srt_mesh_upload(srt_context* ctx, const srt_mesh_desc* desc, srt_mesh** out);
void srt_mesh_destroy(srt_mesh* mesh);
int  srt_mesh_info(const srt_mesh* mesh, uint64_t* n_tris, uint64_t* n_nodes, uint32_t* max_depth);

/* One entry per key of objTriangles, in that map's iteration order (like srt_object). */
typedef struct srt_prep_object {
    const srt_mesh* mesh;         /* the triangles loadObjFile / setTriangles put under this name                    */
    const float*    matrices;     /* n_matrices x 16 floats, column-major (glm::mat4): the transformTriangles calls  */
    uint32_t        n_matrices;   /*   made on this object since it was loaded, in call order                        */
    float           color[3];     /* objColors[name]                                                                */
    float           ambient, specular, shininess;   /* objProperties[name]                                          */
} srt_prep_object;

/* Transforms, builds and flattens on the context stream (asynchronous) into *scene — created when *scene is NULL,
 * otherwise reused (its buffers grow as needed).  textures: uploaded when the scene is created or when non-NULL;
 * pass NULL / 0 afterwards to keep them.  with_flat_lists: also fill the load-order lists of the legacy one-box
 * path (use_bvh = 0). */
int  srt_scene_prep(srt_context* ctx, const srt_prep_object* objects, uint32_t n_objects,
                    const srt_texture_desc* textures, uint32_t n_textures, int with_flat_lists, srt_scene** scene);
/* Synchronises the context; SRT_OK, or SRT_ERR_UNSUPPORTED when the last srt_scene_prep met a non-finite coordinate. */
int  srt_scene_prep_status(srt_context* ctx, srt_scene* scene);
/* Sizes and contents of a device scene (tests, debugging); any output pointer may be NULL. */
int  srt_scene_counts(const srt_scene* scene, uint64_t* n_nodes, uint64_t* n_tris, uint64_t* n_flat, uint32_t* n_objects,
                      uint32_t* max_depth);
int  srt_scene_download(srt_context* ctx, const srt_scene* scene, srt_node* nodes, srt_tri_geom* geom, srt_tri_attr* attr,
                        srt_tri_geom* flat_geom, srt_tri_attr* flat_attr, srt_object* objects);
int  srt_scene_download_vnormals(srt_context* ctx, const srt_scene* scene, srt_tri_vnormal* vnormal, srt_tri_vnormal* flat_vnormal);

/* ---- render: the replacement for sendRaysAndIntersectPointsColors :505 ---- */
/* Asynchronous on the context stream; call srt_sync before reading host outputs. */
int  srt_render(srt_context* ctx, const srt_scene* scene, const srt_render_params* params,
                const srt_frame* out);
int  srt_sync(srt_context* ctx);
/* Statistics of the most recent srt_render on this context (synchronises the context). */
int  srt_last_stats(srt_context* ctx, srt_stats* out);

/* Per-pass device times of the default pipeline (measurement aid).  With stage timing on, srt_render records events
 * around its four launches; srt_last_stage_ms returns {classify, primary, shadow, shade} in ms for the last render, which
 * must have been a one-piece frame (srt_context_set_bands(ctx, 1): with two bands the passes of the halves interleave on
 * two streams and a per-pass time has no meaning).  Synchronises. */
int  srt_context_set_stage_timing(srt_context* ctx, int enabled);
int  srt_last_stage_ms(srt_context* ctx, float* ms4);

/* ---- render with the reference's own result format (struct ImageData :28, filled by :505-525) ----
 * Renders like srt_render, then compacts on the device: only the pixels whose quantised colour is not (0,0,0)
 * (:518), in the reference's column-major scan order (x outer :511, y inner :513); positions (x, y) and colours
 * (r, g, b in 0..255) as floats, i.e. the memory layout of std::vector<glm::vec2> / std::vector<glm::vec3>.
 * srt_render_sparse synchronises and returns the number of entries; srt_fetch_sparse copies up to `capacity`
 * entries to host or device arrays (points_xy[2*i], colors_rgb[3*i]; either may be NULL). */
int  srt_render_sparse(srt_context* ctx, const srt_scene* scene, const srt_render_params* params, uint64_t* count);
/* The same in two steps, for callers that keep several frames in flight (two contexts alternating: one renders while the
 * other's lists travel to the host): _begin enqueues render + compaction + the count's download and returns at once;
 * _count waits for that count alone.  srt_render_sparse = _begin + _count.  srt_fetch_sparse (which waits for a
 * pending count itself) then copies the lists. */
int  srt_render_sparse_begin(srt_context* ctx, const srt_scene* scene, const srt_render_params* params);
int  srt_render_sparse_count(srt_context* ctx, uint64_t* count);
/* srt_fetch_sparse without the wait: the copies are enqueued on the context stream (pinned destinations for real
 * asynchrony) and are complete after srt_sync, or once a LATER srt_render_sparse_count on this context has returned. */
int  srt_fetch_sparse_async(srt_context* ctx, float* points_xy, float* colors_rgb, uint64_t capacity);
int  srt_fetch_sparse(srt_context* ctx, float* points_xy, float* colors_rgb, uint64_t capacity);

/* ---- the same compaction for callers that must not allocate ----
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Builds the ImageData lists of
 * srt_render_sparse from a dense frame the caller already holds: `rgb` = [height][width][3] as srt_render wrote it with
 * no_sky_fill = 1 ((0,0,0) = not drawn, :518); rows [row_begin, row_end) (0, 0 = all) are scanned; `points_xy`
 * (2 floats per entry), `colors_rgb` (3 floats per entry) need room for width * rows entries; `*count` receives the number
 * of entries.  All four pointers must be addresses the DEVICE can reach — device memory, or pinned host memory from
 * srt_host_alloc — because this call makes no copies and allocates nothing: no cudaMalloc / cudaFree, no staging buffer.
 * Its scratch (one word per column and 32-row segment) is a static device array inside the library, 1 MiB per GPU held
 * until exit, which bounds the frame: width * ceil(rows / 32) <= SRT_SPARSE_STATIC_MAX_SEGMENTS (a 3840x2160 frame uses
 * 261,120), else SRT_ERR_UNSUPPORTED.  All calls on one GPU share that array: calls that overlap on the GPU (two contexts
 * / streams of one device) race — serialise them.  Asynchronous on the context stream; srt_sync before reading. */
#define SRT_SPARSE_STATIC_MAX_SEGMENTS 262144
int  srt_sparse_from_frame_static(srt_context* ctx, const uint8_t* rgb, int32_t width, int32_t height, int32_t row_begin,
                                  int32_t row_end, float* points_xy, float* colors_rgb, uint64_t* count);

/* ---- several GPUs from one host thread (SURVEY.md §8b export list, §8e) ----
 * The reference's pixel loop is a pure function of the pixel (:511-523) and its frames are independent (:534), so both
 * shardings are exact: the output of R GPUs is byte-identical to one GPU's.
 *   one frame     interleaved blocks of `row_block` rows, block b -> GPU b mod R (srt_render_params.row_block /
 *                 shard_index / shard_count).  The frame lives on the first GPU; every other GPU's kernels write their
 *                 row blocks STRAIGHT into it through peer-mapped memory (NVLink / NVSwitch stores, no staging copy, no
 *                 collective call); where peer access is unavailable the blocks are rendered locally and moved with
 *                 cudaMemcpyPeerAsync.  One join (events) on the first GPU's stream, then — for host planes — one D2H copy.
 *   image series  frame f -> GPU f mod R, two frames in flight per GPU (upload / render / download overlap), each frame
 *                 downloaded by the GPU that rendered it.
 * devices: CUDA ordinals (NULL: 0 .. n_devices-1); an ordinal may repeat (several shards on one GPU: how the single-GPU
 * tests exercise this path).  All calls are asynchronous; srt_mgpu_sync waits for every GPU.  One host thread at a time. */
typedef struct srt_mgpu srt_mgpu;
int  srt_mgpu_create(const int* devices, int n_devices, srt_mgpu** out);
void srt_mgpu_destroy(srt_mgpu* m);
/* peer_writes: number of GPUs whose kernels write into the first GPU's frame directly (the first GPU included). */
int  srt_mgpu_info(srt_mgpu* m, int* n_devices, int* peer_writes);
/* Replicates the scene on every GPU (first call uploads, later calls update in place like srt_scene_update). */
int  srt_mgpu_scene_upload(srt_mgpu* m, const srt_scene_desc* desc);
/* One frame of the replicated scene over all GPUs.  `out` planes: host pointers, or device pointers on the first GPU.
 * row_block: rows per block, a multiple of 4 (0: 16).  params->row_block / shard_* are ignored (set per GPU here). */
int  srt_mgpu_render_frame(srt_mgpu* m, const srt_render_params* params, const srt_frame* out, int row_block);
/* n_frames frames: scenes[f] (host arrays, pinned for asynchronous uploads) rendered with params[f] into rgb_out[f]
 * (host, width*height*3 bytes each; pinned for asynchronous downloads). */
int  srt_mgpu_render_series(srt_mgpu* m, const srt_scene_desc* const* scenes, const srt_render_params* params,
                            uint32_t n_frames, uint8_t* const* rgb_out);
int  srt_mgpu_sync(srt_mgpu* m);
/* Sums of the per-GPU statistics of the last srt_mgpu_render_frame; kernel_ms = the frame's span on the first GPU's
 * stream, from the fork to the join of all GPUs (before the download).  Synchronises. */
int  srt_mgpu_last_stats(srt_mgpu* m, srt_stats* out);

/* ---- one process per GPU: a frame shared between the processes of one box ----
 * srt_ipc_export writes the 64-byte CUDA IPC handle of a buffer from srt_device_alloc (rank 0's frame); srt_ipc_open maps
 * it in another process for the kernels of that process's context (peer access to the owner's GPU is opened on the way) —
 * the returned address is a valid output plane for srt_render, whose stores then go straight into rank 0's memory over
 * NVLink / NVSwitch (simple_raytracer_b200/multigpu.py: interleaved row blocks per rank, no gather, one small all-reduce
 * as the join).  srt_ipc_close unmaps (synchronises the context first).  The owner must keep the buffer alive. */
int  srt_ipc_export(srt_context* ctx, void* device_ptr, void* handle64);
int  srt_ipc_open(srt_context* ctx, const void* handle64, void** device_ptr);
int  srt_ipc_close(srt_context* ctx, void* device_ptr);

/* ---- helpers ---- */
void srt_default_params(srt_render_params* p);             /* fills the reference's literals */
/* Light sample positions, out[3*i..] for i < count: the staircase of softShadow :363-383 (sequential fp32 adds) or
 * the jittered samples of light_mode. */
int  srt_light_positions(const srt_render_params* p, float* out_xyz);
void* srt_host_alloc(size_t bytes);                        /* pinned host memory */
void  srt_host_free(void* p);
void* srt_device_alloc(srt_context* ctx, size_t bytes);
void  srt_device_free(srt_context* ctx, void* p);
const char* srt_last_error(void);
/* Measurement helper (roofline denominator, SURVEY.md §8d: "L2, to be pinned by a measured L2-read microbenchmark on the
 * box"): every SM streams a `bytes`-sized device buffer `repeats` times with 16-byte loads; returns the read rate of the
 * passes after the first in GB/s (CUDA events on the context's stream).  bytes <= L2 size measures L2 -> SM bandwidth,
 * bytes >> L2 size HBM.  Not on the render path. */
int  srt_measure_read_bandwidth(srt_context* ctx, size_t bytes, int repeats, double* gb_per_s);

/* ---- leaf evaluation (test hook; not on the render path) ----
 * Runs ONE device leaf function of the render kernels on n independent cases so that tests can hold it against
 * the reference's own functions (tests/golden/leaf_kat.npz).  `in` / `out` are HOST arrays, row-major
 * [n][in_width] / [n][out_width]:
 *   SRT_LEAF_POWF      in  x, y                                                            out powf(x, y)          (:196, :398)
 *   SRT_LEAF_SLAB      in  o[3], d[3], bmin[3], bmax[3]                                    out exact, filtered, used_exact
 *                      (intersectRayAabbNoOrigin :252-293; `filtered` is the decision of the kernels' filtered
 *                      slab test, which must equal `exact`; used_exact = 1 when the filter deferred to the exact test)
 *   SRT_LEAF_TRIANGLE  in  o[3], d[3], p0[3], e1[3], e2[3]                                 out t or -inf           (:42-75)
 *   SRT_LEAF_TRIANGLE_FILTERED  same input                                                  out filtered, exact, took_tail
 *                      (the shadow pass's moller_trumbore_filtered, which decides clear misses without the division,
 *                      and the plain function: the two must be the same bits; took_tail = 1 when the reference's own
 *                      operation sequence was run)
 *   SRT_LEAF_PHONG     in  n[3], d[3], light[3], light_color[3], obj_color[3],
 *                          ambient, specular, shininess, t                                 out rgb[3]              (:144-200)
 *   SRT_LEAF_TONE      in  c, tone_k, gamma                                                out q (0..255)          (:391-398, :447-449)
 *   SRT_LEAF_SMOOTH_NORMAL in p0[3], e1[3], e2[3], n1[3], n2[3], n3[3], P[3]                       out n[3]
 *                      (interpolateNormal :132-140 at the barycentric coordinates :79-117 of P)
 */
enum { SRT_LEAF_POWF = 0, SRT_LEAF_SLAB = 1, SRT_LEAF_TRIANGLE = 2, SRT_LEAF_PHONG = 3, SRT_LEAF_TONE = 4,
       SRT_LEAF_SMOOTH_NORMAL = 5, SRT_LEAF_TRIANGLE_FILTERED = 6 };
int  srt_eval_leaf(srt_context* ctx, int op, const float* in, float* out, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* SRT_H_ */

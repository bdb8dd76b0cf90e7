/* oracle/srt_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Plain-C CPU restatement of the reference's per-pixel render loop on the flattened scene of
 * include/srt.h.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / --impl
 * reference legs may load it.  See srt_oracle.c for the reference citations. */
#ifndef SRT_ORACLE_H_
#define SRT_ORACLE_H_
#include <stdint.h>
#include "srt.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Work counters under the REFERENCE's traversal rules (S3/S5/S11), excluding the traversal of
 * the hit object's own BVH that shadowIntersection :328 performs and discards, and counting the
 * final closest hit's shading only. */
typedef struct srt_oracle_counters {
    uint64_t primary_rays;
    uint64_t hit_pixels;
    uint64_t shadow_rays;           /* light_count * hit_pixels when the scene has >= 2 objects */
    uint64_t node_visits_primary;   /* slab tests */
    uint64_t tri_tests_primary;     /* Moeller-Trumbore calls */
    uint64_t node_visits_shadow;
    uint64_t tri_tests_shadow;
    uint64_t lights_evaluated;      /* phongIllumination calls */
    uint64_t texel_fetches;
    uint64_t occluded;              /* shadow rays that found an occluder */
} srt_oracle_counters;

/* Renders columns col_begin, col_begin+col_step, ... (col_step <= 0 -> 1) of rows
 * [row_begin,row_end) with `threads` host threads (0 = all).  Output planes as in srt_frame
 * (host pointers, any may be NULL); untouched outside the sample.  Returns pixels traced or <0. */
long srt_oracle_render(const srt_scene_desc* scene, const srt_render_params* params, const srt_frame* out,
                       int threads, int col_begin, int col_step, srt_oracle_counters* counters,
                       double* seconds);

/* Leaf functions (known-answer tests against the reference's own functions). */
float srt_oracle_moller_trumbore(const float o[3], const float d[3], const float p0[3], const float e1[3],
                                 const float e2[3]);
int   srt_oracle_slab(const float o[3], const float d[3], const float bmin[3], const float bmax[3]);
void  srt_oracle_barycentric(const float p0[3], const float e1[3], const float e2[3], const float point[3],
                             float out[3]);
void  srt_oracle_phong(const float n[3], const float d[3], const float light[3], const float light_color[3],
                       const float obj_color[3], float ambient, float specular, float shininess, float t,
                       float out[3]);
void  srt_oracle_texture_sample(const uint8_t* rgb, int32_t width, int32_t height, float x, float y, int32_t mode, float out[3]);
void  srt_oracle_tonemap_quantise(const float color[3], float tone_k, float gamma, float out[3], int32_t q[3]);
void  srt_oracle_light_positions(const float light_pos[3], int count, float step, float* out_xyz);
void  srt_oracle_smooth_normal(const float p0[3], const float e1[3], const float e2[3], const float n9[9], const float point[3], float out[3]);
void  srt_oracle_light_positions_mode(const srt_render_params* params, float* out_xyz);

#ifdef __cplusplus
}
#endif
#endif

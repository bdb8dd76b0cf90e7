/* oracle/srt_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the reference's per-pixel render loop (leonlang/simple_raytracer,
 * simple_raytracer.cpp:32-457, 505-525) on the flattened scene of include/srt.h.  It exists to
 * check the CUDA kernel and to serve as the timed CPU baseline where oracle/_ref is absent.
 * Parity status: PINNED — tests/test_oracle_golden.py compares every function here with
 * the reference's own compiled functions (oracle/_ref/libsrt_ref.so) bit for bit, and
 * tests/golden/ holds reference-generated vectors for boxes without /root/reference.
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (no -ffast-math, no -march=native): every line below
 * is a sequence of single IEEE fp32 operations in the reference's order.
 * glm formulas restated: dot = (x+y)+z, cross, normalize = v*(1/sqrt(dot)), reflect = I-N*dot(N,I)*2
 * (library/glm-master/glm/detail/func_geometric.inl:48-55, 73-83, 98-106, 120-126).
 */
#define _GNU_SOURCE
#include "srt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#define ORACLE_STACK 128

static inline float dot3(const float a[3], const float b[3]) {
    return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2];
}
static inline void cross3(const float x[3], const float y[3], float o[3]) {
    o[0] = x[1] * y[2] - y[1] * x[2];
    o[1] = x[2] * y[0] - y[2] * x[0];
    o[2] = x[0] * y[1] - y[0] * x[1];
}
/* glm::max(x, y) = (x < y) ? y : x  (func_common.inl) — NaN in x stays NaN, unlike fmaxf. */
static inline float glm_max(float x, float y) { return (x < y) ? y : x; }
static inline void normalize3(const float v[3], float o[3]) {
    const float inv = 1.0f / sqrtf(dot3(v, v));
    o[0] = v[0] * inv; o[1] = v[1] * inv; o[2] = v[2] * inv;
}

/* rayTriangleIntersection, simple_raytracer.cpp:42-75 (p0,e1,e2 = lines :45-51 done by the flattener). */
float srt_oracle_moller_trumbore(const float o[3], const float d[3], const float p0[3], const float e1[3],
                                 const float e2[3]) {
    float pvec[3], tvec[3], qvec[3];
    cross3(d, e2, pvec);                                   /* :54 */
    const float det = dot3(e1, pvec);                      /* :55 */
    if (fabsf(det) < 1e-12f) return -INFINITY;             /* :57 */
    const float invDet = 1.0f / det;                       /* :60 */
    tvec[0] = o[0] - p0[0]; tvec[1] = o[1] - p0[1]; tvec[2] = o[2] - p0[2];   /* :62 */
    const float u = dot3(tvec, pvec) * invDet;             /* :64 */
    if (u < 0.0f || u > 1.0f) return -INFINITY;            /* :65 */
    cross3(tvec, e1, qvec);                                /* :66 */
    const float v = dot3(d, qvec) * invDet;                /* :67 */
    if (v < 0.0f || u + v > 1.0f) return -INFINITY;        /* :68 */
    const float t = dot3(e2, qvec) * invDet;               /* :70 */
    if (t < 0.0f) return -INFINITY;                        /* :73 */
    return t;
}

/* intersectRayAabbNoOrigin, :252-293.  intersectRayAabb :204-248 is the same function with
 * origin (0,0,0): (b - 0.0f) / d == b / d bit for bit, so one routine serves both. */
int srt_oracle_slab(const float o[3], const float d[3], const float bmin[3], const float bmax[3]) {
    float minXT = (bmin[0] - o[0]) / d[0];
    float maxXT = (bmax[0] - o[0]) / d[0];
    if (minXT > maxXT) { float s = minXT; minXT = maxXT; maxXT = s; }
    float minYT = (bmin[1] - o[1]) / d[1];
    float maxYT = (bmax[1] - o[1]) / d[1];
    if (minYT > maxYT) { float s = minYT; minYT = maxYT; maxYT = s; }
    if (maxXT < minYT || maxYT < minXT) return 0;          /* :269 */
    if (minYT > minXT) minXT = minYT;                      /* :274 */
    if (maxYT < maxXT) maxXT = maxYT;                      /* :277 */
    float minZT = (bmin[2] - o[2]) / d[2];
    float maxZT = (bmax[2] - o[2]) / d[2];
    if (minZT > maxZT) { float s = minZT; minZT = maxZT; maxZT = s; }
    if ((minXT > maxZT) || (minZT > maxXT)) return 0;      /* :288 */
    return 1;
}

/* calculateBarycentricCoords, :79-117 (v0 = e1, v1 = e2). */
void srt_oracle_barycentric(const float p0[3], const float e1[3], const float e2[3], const float point[3],
                            float out[3]) {
    float v2[3] = {point[0] - p0[0], point[1] - p0[1], point[2] - p0[2]};
    const float d00 = dot3(e1, e1), d01 = dot3(e1, e2), d11 = dot3(e2, e2);
    const float d20 = dot3(v2, e1), d21 = dot3(v2, e2);
    const float denom = d00 * d11 - d01 * d01;
    const float v = (d11 * d20 - d01 * d21) / denom;
    const float w = (d00 * d21 - d01 * d20) / denom;
    const float u = 1.0f - v - w;
    out[0] = u; out[1] = v; out[2] = w;
}

/* interpolateNormal :132-140 at the barycentric coordinates of `point` (:159): the normal phongIllumination uses when
 * its commented-out line :162 replaces :164.  n9 = normalOne, normalTwo, normalThree as loaded. */
void srt_oracle_smooth_normal(const float p0[3], const float e1[3], const float e2[3], const float n9[9],
                              const float point[3], float out[3]) {
    float bc[3], s[3];
    srt_oracle_barycentric(p0, e1, e2, point, bc);
    for (int c = 0; c < 3; ++c) s[c] = (bc[0] * n9[c] + bc[1] * n9[3 + c]) + bc[2] * n9[6 + c];
    normalize3(s, out);
}

/* phongIllumination, :144-200 with n = calculateTriangleNormal :32-37 precomputed. */
void srt_oracle_phong(const float n[3], const float d[3], const float light[3], const float lc[3],
                      const float oc[3], float ambientStrength, float specularStrength, float shininess,
                      float t, float out[3]) {
    const float rView = 1.0f / 3.14159265358979323846264338327950288f;     /* :153 */
    float P[3] = {0.0f + t * d[0], 0.0f + t * d[1], 0.0f + t * d[2]};      /* :156, origin (0,0,0) */
    float lp[3] = {light[0] - P[0], light[1] - P[1], light[2] - P[2]};
    float l[3];
    normalize3(lp, l);                                                     /* :166 */
    float dp = dot3(n, l);                                                 /* :174 */
    if (dp < 0.0f) dp = -dp;
    const float dpm = glm_max(dp, 0.0f);
    float nd[3] = {-d[0], -d[1], -d[2]}, v[3];
    normalize3(nd, v);                                                     /* :190 */
    float I[3] = {-l[0], -l[1], -l[2]};
    const float k = dot3(n, I);                                            /* reflect :191 */
    float r[3] = {I[0] - n[0] * k * 2.0f, I[1] - n[1] * k * 2.0f, I[2] - n[2] * k * 2.0f};
    const float pw = powf(glm_max(dot3(r, v), 0.0f), shininess);             /* :196 */
    const float amb = rView * ambientStrength;                             /* :184 (1/pi)*ambientStrength */
    for (int c = 0; c < 3; ++c) {
        const float diffuse = rView * oc[c] * lc[c] * dpm;                 /* :178 */
        const float ambient = amb * oc[c] * lc[c];                         /* :184 */
        const float specular = lc[c] * specularStrength * dpm * pw;        /* :196 */
        out[c] = diffuse + specular + ambient;                             /* :199 */
    }
}

/* softShadow :391-398 tone map + gamma, then rayIntersection :447-449 quantisation. */
void srt_oracle_tonemap_quantise(const float color[3], float tone_k, float gamma, float out[3], int32_t q[3]) {
    for (int c = 0; c < 3; ++c) {
        float x = color[c] / (color[c] + tone_k);
        x = powf(x, gamma);
        out[c] = x;
        q[c] = (int32_t)(x * 255);
    }
}

/* Light staircase of softShadow :363-383. */
void srt_oracle_light_positions(const float light_pos[3], int count, float step, float* out) {
    float L[3] = {light_pos[0], light_pos[1], light_pos[2]};
    for (int i = 0; i < count; ++i) {
        out[3 * i] = L[0]; out[3 * i + 1] = L[1]; out[3 * i + 2] = L[2];
        L[i % 3] += step;
    }
}

/* srt_render_params::light_mode (srt.h): the staircase, or the jittered samples (not a reference mode: restated
 * from include/srt.h's definition so that the CUDA path can be checked in that mode too). */
void srt_oracle_light_positions_mode(const srt_render_params* p, float* out) {
    if (p->light_mode == SRT_LIGHTS_JITTERED) {
        const volatile float extent = p->light_step * (float)((p->light_count + 2) / 3);
        for (int i = 0; i < p->light_count; ++i)
            for (int a = 0; a < 3; ++a) {
                uint32_t x = (uint32_t)(3 * i + a) * 0x9E3779B9u + p->light_seed;
                x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
                const volatile float u = (float)(x >> 8) * (1.0f / 16777216.0f);
                const volatile float off = extent * u;
                out[3 * i + a] = p->light_pos[a] + off;
            }
        return;
    }
    srt_oracle_light_positions(p->light_pos, p->light_count, p->light_step, out);
}

/* ------------------------------------------------------------------------------------------- */
typedef struct {
    const srt_scene_desc* sc;
    const srt_render_params* pr;
    float lights[3 * SRT_MAX_LIGHTS];
    int i0, j0;                      /* first values of the loops :511 / :513 */
} oracle_job;

typedef struct { float best; int32_t tri; int32_t obj; } oracle_hit;

/* boundingBoxIntersection :296-317 + the candidate loop of rayIntersection :424-431, merged:
 * depth-first, left child first, every leaf whose box passes the slab test, strict '<'. */
static void closest_in_object(const oracle_job* J, uint32_t obj, const float o[3], const float d[3],
                              oracle_hit* h, srt_oracle_counters* C) {
    const srt_node* nodes = J->sc->nodes;
    const srt_tri_geom* geom = J->sc->geom;
    uint32_t stack[ORACLE_STACK];
    int sp = 0;
    stack[sp++] = J->sc->objects[obj].root;
    while (sp) {
        uint32_t ni = stack[--sp];
        for (;;) {
            const srt_node* n = &nodes[ni];
            C->node_visits_primary++;
            if (!srt_oracle_slab(o, d, n->bmin, n->bmax)) break;
            if (n->b & SRT_LEAF_BIT) {
                const uint32_t cnt = n->b & ~SRT_LEAF_BIT;
                for (uint32_t k = 0; k < cnt; ++k) {
                    const srt_tri_geom* g = &geom[n->a + k];
                    C->tri_tests_primary++;
                    const float t = srt_oracle_moller_trumbore(o, d, g->p0, g->e1, g->e2);
                    if (t != -INFINITY && t < h->best) { h->best = t; h->tri = (int32_t)(n->a + k); h->obj = (int32_t)obj; }
                }
                break;
            }
            stack[sp++] = n->a;      /* right child later */
            ni = ni + 1;             /* left child first  */
        }
    }
}

/* Legacy one-box path :415-420: intersectRayAabb on minBox/maxBox[name], then every triangle of
 * objTriangles[name] in load order. */
static void closest_in_object_flat(const oracle_job* J, uint32_t obj, const float o[3], const float d[3],
                                   oracle_hit* h, srt_oracle_counters* C) {
    const srt_object* ob = &J->sc->objects[obj];
    const float zero[3] = {0.f, 0.f, 0.f};
    C->node_visits_primary++;
    if (!srt_oracle_slab(zero, d, ob->box_min, ob->box_max)) return;
    for (uint32_t k = 0; k < ob->tri_count; ++k) {
        const srt_tri_geom* g = &J->sc->flat_geom[ob->flat_first + k];
        C->tri_tests_primary++;
        const float t = srt_oracle_moller_trumbore(o, d, g->p0, g->e1, g->e2);
        if (t != -INFINITY && t < h->best) { h->best = t; h->tri = (int32_t)(ob->flat_first + k); h->obj = (int32_t)obj; }
    }
}

/* shadowIntersection :321-342: any Moeller-Trumbore hit (any t >= 0) on a candidate triangle of
 * any OTHER object; the own object's traversal (:328, discarded at :331) is skipped.
 * srt_render_params::shadow_segment (not a reference mode): only hits with shadow_eps <= t <= 1 count. */
static int shadowed(const oracle_job* J, int32_t own, const float so[3], const float sd[3], srt_oracle_counters* C) {
    const int segment = J->pr->shadow_segment != 0;
    const float t_lo = J->pr->shadow_eps, t_hi = 1.0f;
    const srt_node* nodes = J->sc->nodes;
    const srt_tri_geom* geom = J->sc->geom;
    for (uint32_t obj = 0; obj < J->sc->n_objects; ++obj) {
        if ((int32_t)obj == own) continue;
        uint32_t stack[ORACLE_STACK];
        int sp = 0;
        stack[sp++] = J->sc->objects[obj].root;
        while (sp) {
            uint32_t ni = stack[--sp];
            for (;;) {
                const srt_node* n = &nodes[ni];
                C->node_visits_shadow++;
                if (!srt_oracle_slab(so, sd, n->bmin, n->bmax)) break;
                if (n->b & SRT_LEAF_BIT) {
                    const uint32_t cnt = n->b & ~SRT_LEAF_BIT;
                    for (uint32_t k = 0; k < cnt; ++k) {
                        const srt_tri_geom* g = &geom[n->a + k];
                        C->tri_tests_shadow++;
                        const float ts = srt_oracle_moller_trumbore(so, sd, g->p0, g->e1, g->e2);
                        if (ts != -INFINITY && (!segment || (!(ts < t_lo) && !(ts > t_hi)))) return 1;
                    }
                    break;
                }
                stack[sp++] = n->a;
                ni = ni + 1;
            }
        }
    }
    return 0;
}

/* srt_render_params::texture_mode NEAREST / BILINEAR (not reference modes; srt.h): (x, y) is the interpolated texel
 * coordinate; every axis wraps on its own (floor, then mod).  Bilinear: the texel values / 255 are blended in x, then in
 * y, with the fractional parts of (x, y): (c00 * (1 - fx) + c10 * fx) * (1 - fy) + (c01 * (1 - fx) + c11 * fx) * fy. */
static int wrap_index(float f, int n) {
    if (!(f > -1.0e9f && f < 1.0e9f)) return 0;          /* NaN / huge: texel 0 */
    int i = (int)f % n;
    if (i < 0) i += n;
    return i;
}
void srt_oracle_texture_sample(const uint8_t* rgb, int32_t width, int32_t height, float x, float y, int32_t mode, float out[3]) {
    const float x0f = floorf(x), y0f = floorf(y);
    const int x0 = wrap_index(x0f, width), y0 = wrap_index(y0f, height);
    if (mode == SRT_TEX_NEAREST) {
        const uint8_t* p = rgb + ((long)y0 * width + x0) * 3;
        out[0] = p[0] / 255.0f; out[1] = p[1] / 255.0f; out[2] = p[2] / 255.0f;
        return;
    }
    const int x1 = x0 + 1 == width ? 0 : x0 + 1, y1 = y0 + 1 == height ? 0 : y0 + 1;
    float fx = x - x0f, fy = y - y0f;
    if (!(fx >= 0.0f && fx <= 1.0f)) fx = 0.0f;
    if (!(fy >= 0.0f && fy <= 1.0f)) fy = 0.0f;
    const float gx = 1.0f - fx, gy = 1.0f - fy;
    const uint8_t* p00 = rgb + ((long)y0 * width + x0) * 3; const uint8_t* p10 = rgb + ((long)y0 * width + x1) * 3;
    const uint8_t* p01 = rgb + ((long)y1 * width + x0) * 3; const uint8_t* p11 = rgb + ((long)y1 * width + x1) * 3;
    for (int c = 0; c < 3; ++c) {
        const float top = (p00[c] / 255.0f) * gx + (p10[c] / 255.0f) * fx;
        const float bot = (p01[c] / 255.0f) * gx + (p11[c] / 255.0f) * fx;
        out[c] = top * gy + bot * fy;
    }
}

/* softShadow :348-401 for the winning hit, then quantisation :447-449. */
static void shade(const oracle_job* J, const oracle_hit* h, const float d[3], int use_flat, float colorPoint[3],
                  srt_oracle_counters* C) {
    const srt_scene_desc* sc = J->sc;
    const srt_render_params* pr = J->pr;
    const srt_tri_geom* g = use_flat ? &sc->flat_geom[h->tri] : &sc->geom[h->tri];
    const srt_tri_attr* a = use_flat ? &sc->flat_attr[h->tri] : &sc->attr[h->tri];
    const srt_object* ob = &sc->objects[h->obj];
    const float t = h->best;
    float base[3] = {ob->color[0], ob->color[1], ob->color[2]};             /* :437-443 */
    if (g->tex >= 0) {                                                       /* :350-361 */
        const srt_texture_desc* tx = &sc->textures[g->tex];
        float P[3] = {0.0f + t * d[0], 0.0f + t * d[1], 0.0f + t * d[2]};
        float bc[3];
        srt_oracle_barycentric(g->p0, g->e1, g->e2, P, bc);
        const float ux = bc[0] * a->uv[0][0] + bc[1] * a->uv[1][0] + bc[2] * a->uv[2][0];   /* :123-125 */
        const float uy = bc[0] * a->uv[0][1] + bc[1] * a->uv[1][1] + bc[2] * a->uv[2][1];
        if (pr->texture_mode == SRT_TEX_TRUNCATE) {
            long idx = ((long)((int)uy * tx->width + (int)ux)) * 3;          /* :357 */
            const long last = ((long)tx->width * tx->height - 1) * 3;
            if (idx < 0) idx = 0;             /* the reference reads out of bounds here (UB); clamp */
            if (idx > last) idx = last;
            base[0] = tx->rgb[idx + 0] / 255.0f; base[1] = tx->rgb[idx + 1] / 255.0f; base[2] = tx->rgb[idx + 2] / 255.0f;
        } else {
            srt_oracle_texture_sample(tx->rgb, tx->width, tx->height, ux, uy, pr->texture_mode, base);
        }
        C->texel_fetches++;
    }
    float nrm[3] = {a->n[0], a->n[1], a->n[2]};                              /* :164 */
    if (pr->smooth_normals) {                                                /* :162 */
        const srt_tri_vnormal* vn = use_flat ? &sc->flat_vnormal[h->tri] : &sc->vnormal[h->tri];
        const float n9[9] = {vn->n[0][0], vn->n[0][1], vn->n[0][2], vn->n[1][0], vn->n[1][1], vn->n[1][2],
                             vn->n[2][0], vn->n[2][1], vn->n[2][2]};
        const float P[3] = {0.0f + t * d[0], 0.0f + t * d[1], 0.0f + t * d[2]};
        srt_oracle_smooth_normal(g->p0, g->e1, g->e2, n9, P, nrm);
    }
    float color[3] = {0.f, 0.f, 0.f};
    const float so[3] = {d[0] * t, d[1] * t, d[2] * t};                      /* :326 */
    for (int i = 0; i < pr->light_count; ++i) {                              /* :366 */
        const float* L = &J->lights[3 * i];
        int sh = 0;
        if (sc->n_objects > 1) {
            const float sd[3] = {L[0] - so[0], L[1] - so[1], L[2] - so[2]};  /* :325 */
            sh = shadowed(J, h->obj, so, sd, C);
            C->shadow_rays++;
            if (sh) C->occluded++;
        }
        float c[3];
        srt_oracle_phong(nrm, d, L, pr->light_color, base, ob->ambient, ob->specular, ob->shininess, t, c);
        C->lights_evaluated++;
        if (sh) { c[0] /= pr->shadow_div; c[1] /= pr->shadow_div; c[2] /= pr->shadow_div; }   /* :369 */
        color[0] += c[0]; color[1] += c[1]; color[2] += c[2];                /* :370 */
    }
    if (pr->normalise_lights && pr->light_count > 0) {                       /* not a reference mode */
        const float n = (float)pr->light_count;
        color[0] /= n; color[1] /= n; color[2] /= n;
    }
    float tm[3]; int32_t q[3];
    srt_oracle_tonemap_quantise(color, pr->tone_k, pr->gamma, tm, q);
    colorPoint[0] = (float)q[0]; colorPoint[1] = (float)q[1]; colorPoint[2] = (float)q[2];
}

typedef struct {
    oracle_job* J;
    const srt_frame* out;
    int col_begin, col_step, row_begin, row_end;
    volatile long* next_col;
    srt_oracle_counters C;
    long traced;
} oracle_worker;

static void* worker_main(void* arg) {
    oracle_worker* w = (oracle_worker*)arg;
    const oracle_job* J = w->J;
    const srt_render_params* pr = J->pr;
    const int W = pr->width;
    const int use_flat = !pr->use_bvh;
    const float o[3] = {0.f, 0.f, 0.f};                                       /* Ray::origin, Object.cpp:17-18 */
    for (;;) {
        const long c = __sync_fetch_and_add(w->next_col, 1);
        const long px = (long)w->col_begin + c * w->col_step;
        if (px >= W) break;
        for (int py = w->row_begin; py < w->row_end; ++py) {
            const float d[3] = {(float)(J->i0 + (int)px), (float)(J->j0 + py), pr->focal};   /* :514-515 */
            oracle_hit h = {INFINITY, -1, -1};
            for (uint32_t obj = 0; obj < J->sc->n_objects; ++obj) {              /* :409 */
                if (use_flat) closest_in_object_flat(J, obj, o, d, &h, &w->C);
                else closest_in_object(J, obj, o, d, &h, &w->C);
            }
            float cp[3] = {0.f, 0.f, 0.f};
            if (h.tri >= 0) { shade(J, &h, d, use_flat, cp, &w->C); w->C.hit_pixels++; }
            w->C.primary_rays++;
            w->traced++;
            const size_t at = (size_t)py * W + (size_t)px;
            if (w->out->rgb) {
                /* :518 drop all-zero; drawImage :469-487 float -> unsigned char, zero -> sky */
                uint8_t b0 = 0, b1 = 0, b2 = 0;
                if (!(cp[0] == 0.f && cp[1] == 0.f && cp[2] == 0.f)) {
                    b0 = (uint8_t)(int32_t)cp[0]; b1 = (uint8_t)(int32_t)cp[1]; b2 = (uint8_t)(int32_t)cp[2];
                }
                if (b0 == 0 && b1 == 0 && b2 == 0 && !pr->no_sky_fill) { b0 = pr->sky[0]; b1 = pr->sky[1]; b2 = pr->sky[2]; }
                uint8_t* p = w->out->rgb + at * 3;
                p[0] = b0; p[1] = b1; p[2] = b2;
            }
            if (w->out->obj_id) w->out->obj_id[at] = h.obj;
            if (w->out->tri_id) w->out->tri_id[at] = h.tri;
            if (w->out->t) w->out->t[at] = h.best;
        }
    }
    return NULL;
}

long srt_oracle_render(const srt_scene_desc* scene, const srt_render_params* params, const srt_frame* out,
                       int threads, int col_begin, int col_step, srt_oracle_counters* counters, double* seconds) {
    if (!scene || !params || !out) return -1;
    if (params->light_count < 0 || params->light_count > SRT_MAX_LIGHTS) return -1;
    if (!params->use_bvh && (!scene->flat_geom || !scene->flat_attr)) return -1;
    if (params->smooth_normals && scene->n_objects > 0 && !(params->use_bvh ? scene->vnormal : scene->flat_vnormal)) return -1;
    oracle_job J;
    J.sc = scene; J.pr = params;
    srt_oracle_light_positions_mode(params, J.lights);
    /* loop bounds of :511/:513: int i = -imageSize.x / 2 (float -> int truncation) */
    J.i0 = (int)(-(float)params->width / 2);
    J.j0 = (int)(-(float)params->height / 2);
    int nt = threads > 0 ? threads : (int)sysconf(_SC_NPROCESSORS_ONLN);
    if (nt < 1) nt = 1;
    if (nt > 1024) nt = 1024;
    int rb = params->row_begin, re = params->row_end;
    if (rb == 0 && re == 0) re = params->height;
    volatile long next = 0;
    oracle_worker* ws = (oracle_worker*)calloc((size_t)nt, sizeof *ws);
    pthread_t* th = (pthread_t*)calloc((size_t)nt, sizeof *th);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int k = 0; k < nt; ++k) {
        ws[k].J = &J; ws[k].out = out; ws[k].col_begin = col_begin > 0 ? col_begin : 0;
        ws[k].col_step = col_step > 0 ? col_step : 1; ws[k].row_begin = rb; ws[k].row_end = re; ws[k].next_col = &next;
        if (k > 0) pthread_create(&th[k], NULL, worker_main, &ws[k]);
    }
    worker_main(&ws[0]);
    for (int k = 1; k < nt; ++k) pthread_join(th[k], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (seconds) *seconds = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    long traced = 0;
    srt_oracle_counters tot;
    memset(&tot, 0, sizeof tot);
    for (int k = 0; k < nt; ++k) {
        traced += ws[k].traced;
        const uint64_t* s = (const uint64_t*)&ws[k].C;
        uint64_t* dst = (uint64_t*)&tot;
        for (size_t q = 0; q < sizeof tot / sizeof(uint64_t); ++q) dst[q] += s[q];
    }
    if (counters) *counters = tot;
    free(ws); free(th);
    return traced;
}

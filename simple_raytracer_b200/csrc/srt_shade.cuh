// srt_shade.cuh — per-pixel shading shared by the render kernels: base colour / texture lookup,
// the light-independent part of Phong, and the final 8-bit store with the sky rule.
#pragma once
#include "srt_kernels.cuh"
#include "srt_device.cuh"

namespace srtk {
using namespace srtd;

struct Hit { float t; int32_t tri; int32_t obj; };

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// interpolateNormal :132-140 at calculateBarycentricCoords(triangle, P) :159: (b.x * n1 + b.y * n2) + b.z * n3, normalised.
__device__ __forceinline__ V3 smooth_normal(V3 p0, V3 e1, V3 e2, float4 n1, float4 n2, float4 n3, V3 P) {
    const V3 bc = barycentric(p0, e1, e2, P);
    const V3 s = v3(add(add(mul(bc.x, n1.x), mul(bc.y, n2.x)), mul(bc.z, n3.x)),
                    add(add(mul(bc.x, n1.y), mul(bc.y, n2.y)), mul(bc.z, n3.y)),
                    add(add(mul(bc.x, n1.z), mul(bc.y, n2.z)), mul(bc.z, n3.z)));
    return normalize3(s);
}

// srt_render_params::normalise_lights: the light sum divided by the number of samples.
__device__ __forceinline__ V3 average_lights(V3 c, int n) {
    const float f = (float)n;
    return n > 0 ? v3(dvd(c.x, f), dvd(c.y, f), dvd(c.z, f)) : c;
}

// srt_render_params::texture_mode NEAREST / BILINEAR (srt.h): wrapped addressing of the texel coordinate (x, y).
__device__ __forceinline__ int wrap_index(float f, int n) {
    if (!(f > -1.0e9f && f < 1.0e9f)) return 0;
    int i = (int)f % n;
    return i < 0 ? i + n : i;
}
__device__ __forceinline__ V3 texel(const TexDev& tx, int x, int y) {
    const uint8_t* p = tx.rgb + ((long long)y * tx.width + x) * 3;
    return v3(dvd((float)p[0], 255.0f), dvd((float)p[1], 255.0f), dvd((float)p[2], 255.0f));
}
__device__ __forceinline__ float blend2(float c0, float w0, float c1, float w1) { return add(mul(c0, w0), mul(c1, w1)); }
__device__ __forceinline__ V3 texture_sample(const TexDev& tx, float x, float y, int mode) {
    const float x0f = floorf(x), y0f = floorf(y);
    const int x0 = wrap_index(x0f, tx.width), y0 = wrap_index(y0f, tx.height);
    if (mode == SRT_TEX_NEAREST) return texel(tx, x0, y0);
    const int x1 = x0 + 1 == tx.width ? 0 : x0 + 1, y1 = y0 + 1 == tx.height ? 0 : y0 + 1;
    float fx = sub(x, x0f), fy = sub(y, y0f);
    if (!(fx >= 0.0f && fx <= 1.0f)) fx = 0.0f;
    if (!(fy >= 0.0f && fy <= 1.0f)) fy = 0.0f;
    const float gx = sub(1.0f, fx), gy = sub(1.0f, fy);
    const V3 c00 = texel(tx, x0, y0), c10 = texel(tx, x1, y0), c01 = texel(tx, x0, y1), c11 = texel(tx, x1, y1);
    return v3(blend2(blend2(c00.x, gx, c10.x, fx), gy, blend2(c01.x, gx, c11.x, fx), fy),
              blend2(blend2(c00.y, gx, c10.y, fx), gy, blend2(c01.y, gx, c11.y, fx), fy),
              blend2(blend2(c00.z, gx, c10.z, fx), gy, blend2(c01.z, gx, c11.z, fx), fy));
}

// Base colour (:437-443, :350-361) and the light-independent part of phongIllumination.
__device__ __forceinline__ PhongPixel prepare_pixel(const KArgs& a, const Hit& h, V3 d, bool use_flat) {
    const float4* gp = (use_flat ? a.flat_geom : a.geom) + 3 * (size_t)h.tri;
    const float4* ap = (use_flat ? a.flat_attr : a.attr) + 3 * (size_t)h.tri;
    const float4 g0 = ldg4(gp), g1 = ldg4(gp + 1), g2 = ldg4(gp + 2);
    const float4 a0 = ldg4(ap);
    const srt_object& ob = a.objects[h.obj];
    PhongPixel px;
    px.n = v3(a0.x, a0.y, a0.z);                                          // calculateTriangleNormal :164
    // ray->origin + distance * ray->direction, origin = (0,0,0)          :156 / :351
    px.P = v3(add(0.0f, mul(h.t, d.x)), add(0.0f, mul(h.t, d.y)), add(0.0f, mul(h.t, d.z)));
    px.v = normalize3(v3(-d.x, -d.y, -d.z));                              // :190
    px.oc = v3(ob.color[0], ob.color[1], ob.color[2]);                    // :439
    px.ambientStrength = ob.ambient; px.specularStrength = ob.specular; px.shininess = ob.shininess;
    if (a.smooth_normals) {                                               // the commented-out line :162
        const float4* np = (use_flat ? a.flat_vnormal : a.vnormal) + 3 * (size_t)h.tri;
        px.n = smooth_normal(v3(g0.x, g0.y, g0.z), v3(g1.x, g1.y, g1.z), v3(g2.x, g2.y, g2.z), ldg4(np), ldg4(np + 1),
                             ldg4(np + 2), px.P);
    }
    const int32_t tex = __float_as_int(g1.w);
    if (tex >= 0) {                                                       // :350
        const float4 a1 = ldg4(ap + 1), a2 = ldg4(ap + 2);
        // srt_tri_attr as three float4: n.xyz pad | uv0.x uv0.y uv1.x uv1.y | uv2.x uv2.y pad pad
        const V3 bc = barycentric(v3(g0.x, g0.y, g0.z), v3(g1.x, g1.y, g1.z), v3(g2.x, g2.y, g2.z), px.P);
        // getTextureCoordinate :123-125: (b.x*A + b.y*B) + b.z*C per component
        const float ux = add(add(mul(bc.x, a1.x), mul(bc.y, a1.z)), mul(bc.z, a2.x));
        const float uy = add(add(mul(bc.x, a1.y), mul(bc.y, a1.w)), mul(bc.z, a2.y));
        const TexDev tx = a.textures[tex];
        if (a.texture_mode == SRT_TEX_TRUNCATE) {
            long long idx = ((long long)(__float2int_rz(uy) * tx.width + __float2int_rz(ux))) * 3;   // :357
            const long long last = ((long long)tx.width * tx.height - 1) * 3;
            idx = idx < 0 ? 0 : (idx > last ? last : idx);   // the reference reads out of bounds here (UB): clamp
            px.oc = v3(dvd((float)tx.rgb[idx + 0], 255.0f), dvd((float)tx.rgb[idx + 1], 255.0f),
                       dvd((float)tx.rgb[idx + 2], 255.0f));              // :358-360
        } else {
            px.oc = texture_sample(tx, ux, uy, a.texture_mode);
        }
    }
    return px;
}

__device__ __forceinline__ void store_pixel(const KArgs& a, int px, int py, bool hit, V3 color, const Hit& h,
                                            bool planes = true) {
    const size_t at = (size_t)py * a.width + px;
    if (a.rgb) {
        int r = 0, g = 0, b = 0;
        if (hit) {
            r = tone_quantise(color.x, a.tone_k, a.gamma) & 0xff;         // :391-398, :447-449, :470-472
            g = tone_quantise(color.y, a.tone_k, a.gamma) & 0xff;
            b = tone_quantise(color.z, a.tone_k, a.gamma) & 0xff;
        }
        if ((r | g | b) == 0 && !a.no_sky_fill) { r = a.sky[0]; g = a.sky[1]; b = a.sky[2]; }   // :518, :476-487
        uint8_t* p = a.rgb + at * 3;
        p[0] = (uint8_t)r; p[1] = (uint8_t)g; p[2] = (uint8_t)b;
    }
    if (!planes) return;
    if (a.obj_id) a.obj_id[at] = h.obj;
    if (a.tri_id) a.tri_id[at] = h.tri;
    if (a.t) a.t[at] = h.t;
}


}  // namespace srtk

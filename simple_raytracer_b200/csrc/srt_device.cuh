// srt_device.cuh — device-side arithmetic of the render core (sm_100a).
//
// Every function states the reference lines it reproduces (leonlang/simple_raytracer,
// simple_raytracer.cpp unless noted) and evaluates the SAME sequence of IEEE-754 binary32
// operations, round-to-nearest, with no FMA contraction: all products/sums go through
// __fmul_rn/__fadd_rn/__fsub_rn (which nvcc never fuses, whatever -fmad says), divisions through
// __fdiv_rn and square roots through __fsqrt_rn.  The only operations that are
// not bit-defined by IEEE are the two powf calls (:196, :398); see srt_powf below.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "srt_powf_tables.cuh"

namespace srtd {

#define SRT_DEV __device__ __forceinline__

SRT_DEV float mul(float a, float b) { return __fmul_rn(a, b); }
SRT_DEV float add(float a, float b) { return __fadd_rn(a, b); }
SRT_DEV float sub(float a, float b) { return __fsub_rn(a, b); }
SRT_DEV float dvd(float a, float b) { return __fdiv_rn(a, b); }

struct V3 { float x, y, z; };

SRT_DEV V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
// glm::dot(vec3) = (x*x' + y*y') + z*z'           glm/detail/func_geometric.inl:48-55
SRT_DEV float dot3(V3 a, V3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
// glm::cross                                       glm/detail/func_geometric.inl:73-83
SRT_DEV V3 cross3(V3 x, V3 y) {
    return v3(sub(mul(x.y, y.z), mul(y.y, x.z)),
              sub(mul(x.z, y.x), mul(y.z, x.x)),
              sub(mul(x.x, y.y), mul(y.x, x.y)));
}
// glm::normalize = v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x)
//                                                  func_geometric.inl:98-106, func_exponential.inl:134-139
SRT_DEV V3 normalize3(V3 v) {
    const float inv = dvd(1.0f, __fsqrt_rn(dot3(v, v)));
    return v3(mul(v.x, inv), mul(v.y, inv), mul(v.z, inv));
}
// glm::max(x, y) = (x < y) ? y : x — keeps a NaN in x, unlike fmaxf   func_common.inl:30-35
SRT_DEV float glm_max(float x, float y) { return (x < y) ? y : x; }

// std::pow(float,float) (glibc powf: evaluated in double, result within 0.52 ulp before the final
// rounding — about 1e-3 of its results are not the correctly rounded value).  libdevice powf is only
// good to a few ulp, which would move ~1e-4 of all 8-bit channel values by one LSB.  Here x^y =
// exp2(y * log2 x) is evaluated in double to a relative error below 2^-44 and rounded once to float
// (tables and error budget: tools/gen_powf_tables.py): the correctly rounded powf except when the
// exact value lies within ~1e-13 relative of a float rounding boundary.  Everything outside the
// plain case (x not positive-normal, y not finite, results near the double range) goes to pow() in
// double, which has the same property and all the special cases of C99 Annex F.
SRT_DEV float srt_powf(float x, float y) {
    namespace T = srtk::powf_tab;
    const uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 1.0f;                                   // pow(1, y) = 1 for every y
    if (ix == 0u && y > 0.0f) return 0.0f;                                // pow(+0, y > 0) = +0 (the clamped specular term)
    if (ix - 0x00800000u < 0x7f000000u && y >= 1.0f && y <= 32.0f && y == truncf(y)) {
        // A small integer exponent — phongIllumination's shininess (15 by default, :196): x^n by binary exponentiation in
        // double, at most nine multiplications, relative error below 2^-49 before the single rounding to float (tighter than
        // the exp2 / log2 path below, at a quarter of its instructions).  Over- and underflow of the double intermediates
        // happen only where the float result is inf or 0 anyway.
        double b = (double)x, acc = 1.0;
        int n = (int)y;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            if (n & 1) acc *= b;
            b *= b;
            n >>= 1;
        }
        return (float)acc;
    }
    if (ix - 0x00800000u < 0x7f000000u && fabsf(y) < 3.0e38f) {           // x positive, normal, finite; y finite
        const uint32_t tmp = ix - T::kOff;
        const double2 c = __ldg(&T::kLog[(tmp >> (23 - T::kLogBits)) & ((1u << T::kLogBits) - 1u)]);
        const double z = (double)__uint_as_float(ix - (tmp & 0xff800000u));
        const double k = (double)((int32_t)tmp >> 23);
        const double r = fma(z, c.x, -1.0);
        double p = fma(T::kL7, r, T::kL6);
        p = fma(p, r, T::kL5); p = fma(p, r, T::kL4); p = fma(p, r, T::kL3); p = fma(p, r, T::kL2); p = fma(p, r, T::kL1);
        const double t = (double)y * fma(p, r, c.y + k);                  // y * log2 x
        if (fabs(t) < 1000.0) {
            const double shift = 6755399441055744.0;                      // 1.5 * 2^52: round to nearest integer
            double kd = fma(t, (double)(1 << T::kExpBits), shift);
            const int32_t ki = __double2loint(kd);
            kd -= shift;
            const double q = fma(kd, -1.0 / (double)(1 << T::kExpBits), t);   // |q| <= 1/64
            double e = fma(T::kE5, q, T::kE4);
            e = fma(e, q, T::kE3); e = fma(e, q, T::kE2); e = fma(e, q, T::kE1); e = fma(e, q, 1.0);
            const double s = __ldg(&T::kExp[ki & ((1 << T::kExpBits) - 1)]);
            const int hi = __double2hiint(s) + ((ki >> T::kExpBits) << 20);   // times 2^(ki div 32)
            return (float)(__hiloint2double(hi, __double2loint(s)) * e);
        }
    }
    return (float)pow((double)x, (double)y);
}

// ---------------------------------------------------------------------------------------------
// intersectRayAabbNoOrigin :252-293 (and intersectRayAabb :204-248, which is the same function
// with origin (0,0,0): (b - 0.0f)/d == b/d bit for bit).  True IEEE division, the reference's
// compare order, NaN (0/0) comparisons false exactly as written.  No t >= 0 clipping (S3).
SRT_DEV bool slab_exact(V3 o, V3 d, float4 lo, float4 hi) {
    float minXT = dvd(sub(lo.x, o.x), d.x);
    float maxXT = dvd(sub(hi.x, o.x), d.x);
    if (minXT > maxXT) { const float s = minXT; minXT = maxXT; maxXT = s; }      // :258
    float minYT = dvd(sub(lo.y, o.y), d.y);
    float maxYT = dvd(sub(hi.y, o.y), d.y);
    if (minYT > maxYT) { const float s = minYT; minYT = maxYT; maxYT = s; }      // :264
    if (maxXT < minYT || maxYT < minXT) return false;                            // :269
    if (minYT > minXT) minXT = minYT;                                            // :274
    if (maxYT < maxXT) maxXT = maxYT;                                            // :277
    float minZT = dvd(sub(lo.z, o.z), d.z);
    float maxZT = dvd(sub(hi.z, o.z), d.z);
    if (minZT > maxZT) { const float s = minZT; minZT = maxZT; maxZT = s; }      // :284
    if ((minXT > maxZT) || (minZT > maxXT)) return false;                        // :288
    return true;
}

// rayTriangleIntersection :42-75.  g0 = (p0.xyz, obj), g1 = (e1.xyz, tex), g2 = (e2.xyz, -)
// with p0/e1/e2 = lines :45-51 precomputed by the flattener.  Returns -INF for a miss.
SRT_DEV float moller_trumbore(V3 o, V3 d, float4 g0, float4 g1, float4 g2) {
    const V3 p0 = v3(g0.x, g0.y, g0.z), e1 = v3(g1.x, g1.y, g1.z), e2 = v3(g2.x, g2.y, g2.z);
    const V3 pvec = cross3(d, e2);                                    // :54
    const float det = dot3(e1, pvec);                                 // :55
    if (fabsf(det) < 1e-12f) return -INFINITY;                        // :57
    const float invDet = dvd(1.0f, det);                              // :60
    const V3 tvec = v3(sub(o.x, p0.x), sub(o.y, p0.y), sub(o.z, p0.z));   // :62
    const float u = mul(dot3(tvec, pvec), invDet);                    // :64
    if (u < 0.0f || u > 1.0f) return -INFINITY;                       // :65
    const V3 qvec = cross3(tvec, e1);                                 // :66
    const float v = mul(dot3(d, qvec), invDet);                       // :67
    if (v < 0.0f || add(u, v) > 1.0f) return -INFINITY;               // :68
    const float t = mul(dot3(e2, qvec), invDet);                      // :70
    if (t < 0.0f) return -INFINITY;                                   // :73
    return t;
}

// The same function with the three early returns decided WITHOUT the division wherever that is certain: the value
// returned is moller_trumbore's, bit for bit, but a triangle the ray clearly misses — nearly every candidate — costs no
// 1/det and no products with it.  det, A = dot(tvec, pvec), B = dot(d, qvec), C = dot(e2, qvec) are the reference's own
// values.  With s = sign(det), a = s*A, D = |det| the reference computes u = RN(A * RN(1/det)), within (1 +- 2^-23) of a/D:
//   u < 0   is certain when a < 0 and |a| > D * 2^-80   (the product cannot round to -0);
//   u > 1   is certain when a > RN(D * (1 + 2^-20))     (u >= (a/D)(1 - 2^-23) > 1);
//   v < 0   likewise with b = s*B;
//   RN(u + v) > 1 is certain when a, b >= 0 and RN(a + b) > RN(D * (1 + 2^-19));
//   t < 0   is certain when c = s*C < 0 and |c| > D * 2^-80.
// A comparison with NaN / inf is false, i.e. "not certain", and everything not certain takes the reference's own
// operations (the tail below).  Rays through an edge within ~1e-6 of the triangle's width and real hits reach the tail.
SRT_DEV float moller_trumbore_filtered(V3 o, V3 d, float4 g0, float4 g1, float4 g2, bool* took_tail = nullptr) {
    if (took_tail) *took_tail = false;
    const V3 p0 = v3(g0.x, g0.y, g0.z), e1 = v3(g1.x, g1.y, g1.z), e2 = v3(g2.x, g2.y, g2.z);
    const V3 pvec = cross3(d, e2);                                    // :54
    const float det = dot3(e1, pvec);                                 // :55
    const float D = fabsf(det);
    if (D < 1e-12f) return -INFINITY;                                 // :57
    const uint32_t sgn = __float_as_uint(det) & 0x80000000u;
    const float tiny = D * 8.2718061e-25f;                            // D * 2^-80 (normal: D >= 1e-12)
    const V3 tvec = v3(sub(o.x, p0.x), sub(o.y, p0.y), sub(o.z, p0.z));   // :62
    const float A = dot3(tvec, pvec);
    const float a = __uint_as_float(__float_as_uint(A) ^ sgn);
    if ((a < 0.0f && -a > tiny) || a > D * 1.00000095367431640625f) return -INFINITY;      // :65, certain
    const V3 qvec = cross3(tvec, e1);                                 // :66
    const float B = dot3(d, qvec);
    const float b = __uint_as_float(__float_as_uint(B) ^ sgn);
    if ((b < 0.0f && -b > tiny) || (a >= 0.0f && b >= 0.0f && a + b > D * 1.0000019073486328125f)) return -INFINITY;   // :68, certain
    const float C = dot3(e2, qvec);
    const float c = __uint_as_float(__float_as_uint(C) ^ sgn);
    if (c < 0.0f && -c > tiny) return -INFINITY;                      // :73, certain
    // not certain: the reference's own sequence
    if (took_tail) *took_tail = true;
    const float invDet = dvd(1.0f, det);                              // :60
    const float u = mul(A, invDet);                                   // :64
    if (u < 0.0f || u > 1.0f) return -INFINITY;                       // :65
    const float v = mul(B, invDet);                                   // :67
    if (v < 0.0f || add(u, v) > 1.0f) return -INFINITY;               // :68
    const float t = mul(C, invDet);                                   // :70
    if (t < 0.0f) return -INFINITY;                                   // :73
    return t;
}

SRT_DEV bool moller_trumbore_took_tail(V3 o, V3 d, float4 g0, float4 g1, float4 g2) {      // test hook (srt_eval_leaf)
    bool tail;
    moller_trumbore_filtered(o, d, g0, g1, g2, &tail);
    return tail;
}

// calculateBarycentricCoords :79-117 (v0 = e1, v1 = e2); returns (u, v, w).
SRT_DEV V3 barycentric(V3 p0, V3 e1, V3 e2, V3 P) {
    const V3 v2 = v3(sub(P.x, p0.x), sub(P.y, p0.y), sub(P.z, p0.z));
    const float d00 = dot3(e1, e1), d01 = dot3(e1, e2), d11 = dot3(e2, e2);
    const float d20 = dot3(v2, e1), d21 = dot3(v2, e2);
    const float denom = sub(mul(d00, d11), mul(d01, d01));            // :105
    const float v = dvd(sub(mul(d11, d20), mul(d01, d21)), denom);    // :109
    const float w = dvd(sub(mul(d00, d21), mul(d01, d20)), denom);    // :110
    const float u = sub(sub(1.0f, v), w);                             // :112
    return v3(u, v, w);
}

// Per-pixel constants of phongIllumination that do not depend on the light.
struct PhongPixel {
    V3 n;        // calculateTriangleNormal :32-37 (precomputed by the flattener)
    V3 P;        // ray->origin + distance * ray->direction        :156
    V3 v;        // normalize(-ray->direction)                     :190
    V3 oc;       // objectColor (object colour or texel)
    float ambientStrength, specularStrength, shininess;
};

// phongIllumination :144-200 for one light.
SRT_DEV V3 phong(const PhongPixel& px, V3 L, V3 lc) {
    const float rView = dvd(1.0f, 3.14159265358979323846264338327950288f);   // 1.0f / glm::pi<float>()   :153 (folded)
    const V3 l = normalize3(v3(sub(L.x, px.P.x), sub(L.y, px.P.y), sub(L.z, px.P.z)));   // :166
    float dp = dot3(px.n, l);                                         // :174
    if (dp < 0.0f) dp = -dp;                                          // :175-177
    const float dpm = glm_max(dp, 0.0f);
    const V3 I = v3(-l.x, -l.y, -l.z);
    const float k = dot3(px.n, I);                                    // reflect(-l, n) :191
    const V3 r = v3(sub(I.x, mul(mul(px.n.x, k), 2.0f)), sub(I.y, mul(mul(px.n.y, k), 2.0f)),
                    sub(I.z, mul(mul(px.n.z, k), 2.0f)));
    const float pw = srt_powf(glm_max(dot3(r, px.v), 0.0f), px.shininess);   // :196
    const float amb = mul(rView, px.ambientStrength);                 // (1/pi) * ambientStrength  :184
    V3 out;
    {
        const float diffuse = mul(mul(mul(rView, px.oc.x), lc.x), dpm);          // :178
        const float ambient = mul(mul(amb, px.oc.x), lc.x);                      // :184
        const float specular = mul(mul(mul(lc.x, px.specularStrength), dpm), pw);   // :196
        out.x = add(add(diffuse, specular), ambient);                            // :199
    }
    {
        const float diffuse = mul(mul(mul(rView, px.oc.y), lc.y), dpm);
        const float ambient = mul(mul(amb, px.oc.y), lc.y);
        const float specular = mul(mul(mul(lc.y, px.specularStrength), dpm), pw);
        out.y = add(add(diffuse, specular), ambient);
    }
    {
        const float diffuse = mul(mul(mul(rView, px.oc.z), lc.z), dpm);
        const float ambient = mul(mul(amb, px.oc.z), lc.z);
        const float specular = mul(mul(mul(lc.z, px.specularStrength), dpm), pw);
        out.z = add(add(diffuse, specular), ambient);
    }
    return out;
}

// softShadow :391-398 (tone map, gamma) + rayIntersection :447-449 (truncating quantisation)
// + drawImage :470-472 (float -> unsigned char).
SRT_DEV int tone_quantise(float c, float tone_k, float gamma) {
    float x = dvd(c, add(c, tone_k));                                 // :391
    x = srt_powf(x, gamma);                                           // :398
    // int(color * 255) on x86 is cvttss2si: NaN / out of range -> INT_MIN, whose low byte is 0.
    const float s = mul(x, 255.0f);
    int q = (s != s || s >= 2147483648.0f || s < -2147483648.0f) ? 0 : __float2int_rz(s);
    return q;
}

}  // namespace srtd

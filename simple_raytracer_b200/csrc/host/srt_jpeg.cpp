// srt_jpeg.cpp — JPEG files to the RGB8 buffer stbi_load(path, &w, &h, &ch, 3) returns for them (Object.cpp:57; the
// reference decodes textures with its vendored stb_image 2.30, library/stb-master/stb_image.h).
//
// JPEG is lossy and its standard leaves the arithmetic of the inverse DCT, of chroma upsampling and of the colour
// conversion to the decoder, so two correct decoders differ by an LSB here and there.  "The pixels the reference renders
// with" therefore means stb_image's arithmetic, restated here on top of our own parser / entropy decoder:
//   * baseline and progressive Huffman streams (SOF0 / SOF1 / SOF2), 8-bit, 1 or 3 components, restart intervals;
//     coefficients are multiplied by the quantiser and truncated to 16 bits (stb_image.h:2226-2261, :3068-3073);
//   * inverse DCT: the 13-bit fixed-point "islow" butterfly with stb_image's constants and rounding — column pass
//     >> 10 after + 512 with the all-AC-zero shortcut dc * 4, row pass >> 17 after + 65536 + (128 << 17), clamp (:2452-2523);
//   * upsampling: 2x1 / 1x2 / 2x2 by the (3 near + far) triangle filters with their edge rules, anything else by
//     replication (:3465-3527, :3646-3655); the vertical phase walks like :3921-3936;
//   * YCbCr -> RGB in 20-bit fixed point with the masked green term (:3660-3683); components named 'R','G','B', or an
//     Adobe segment with transform 0 and no JFIF segment, are copied (:3880); grey is replicated.
// Arithmetic coding, 12-bit samples and four-component (CMYK / YCCK) files are refused (the caller then falls back).
// Tested against the reference's decoder on its texture files, bit for bit (tests/test_host_scene.py).
#include <cstdint>
#include <cstring>
#include <vector>

#include "srt_scene.hpp"

namespace srt {
namespace {

const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                             35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// Entropy-coded segment reader: bytes with 0xFF00 unstuffed; at a marker it stops advancing and supplies zero bits.
struct Bits {
    const uint8_t* p = nullptr; size_t n = 0, at = 0;
    uint32_t hold = 0; int have = 0;
    int marker = 0;                       // the marker that ended the segment (0: none seen yet)
    void start(size_t pos) { at = pos; hold = 0; have = 0; marker = 0; }
    void fill() {
        while (have <= 24) {
            uint32_t b = 0;
            if (!marker && at < n) {
                b = p[at];
                if (b == 0xFF) {
                    size_t q = at + 1;
                    while (q < n && p[q] == 0xFF) ++q;                 // fill bytes
                    if (q < n && p[q] == 0) { at = q + 1; }             // stuffed zero: a data byte 0xFF
                    else { marker = q < n ? p[q] : 0xD9; at = q < n ? q + 1 : n; b = 0; }
                } else {
                    ++at;
                }
            }
            hold |= b << (24 - have); have += 8;
        }
    }
    int bit() { if (have < 1) fill(); const int v = (int)(hold >> 31); hold <<= 1; --have; return v; }
    int bits(int k) {
        if (k == 0) return 0;
        if (have < k) fill();
        const int v = (int)(hold >> (32 - k)); hold <<= k; have -= k;
        return v;
    }
};

struct HuffTable {
    int maxcode[18]; int valptr[17]; int mincode[17]; uint8_t values[256];
    HuffTable() { for (int l = 0; l < 18; ++l) maxcode[l] = -1; std::memset(valptr, 0, sizeof valptr); std::memset(mincode, 0, sizeof mincode); }   // undefined table: every code is invalid
    bool build(const int counts[16], const uint8_t* vals, int nvals) {
        int code = 0, k = 0;
        for (int l = 1; l <= 16; ++l) {
            valptr[l] = k; mincode[l] = code;
            code += counts[l - 1]; k += counts[l - 1];
            if (counts[l - 1] && code - 1 >= (1 << l)) return false;
            maxcode[l] = counts[l - 1] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        if (k != nvals || k > 256) return false;
        std::memcpy(values, vals, nvals);
        return true;
    }
    int decode(Bits& br) const {
        int code = 0;
        for (int l = 1; l <= 16; ++l) {
            code = (code << 1) | br.bit();
            if (maxcode[l] >= 0 && code <= maxcode[l] && code >= mincode[l]) return values[valptr[l] + code - mincode[l]];
        }
        return -1;
    }
};

// The value of `s` additional bits (ITU T.81 F.2.2.1): negative when the first bit is 0.
inline int extend(Bits& br, int s) {
    if (s == 0) return 0;
    const int v = br.bits(s);
    return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v;
}

inline bool fits_short(int64_t v) { return v >= -32768 && v <= 32767; }

struct Component {
    int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0;
    int x = 0, y = 0, w2 = 0, h2 = 0;     // size in samples; padded size (whole MCUs)
    int dc_pred = 0;
    std::vector<uint8_t> data;            // [h2][w2] samples after the inverse DCT
    std::vector<int16_t> coeff;           // progressive: [h2/8][w2/8][64] coefficients in natural order
};

struct Decoder {
    Bits br;
    int width = 0, height = 0, ncomp = 0;
    bool progressive = false, jfif = false;
    int adobe_transform = -1, rgb_ids = 0;
    int restart_interval = 0;
    int h_max = 1, v_max = 1, mcu_x = 0, mcu_y = 0;
    uint16_t dequant[4][64];
    HuffTable dc[4], ac[4];
    Component comp[4];
    int scan_n = 0, order[4] = {0, 0, 0, 0};
    int spec_start = 0, spec_end = 63, succ_high = 0, succ_low = 0;
    int eob_run = 0;

    bool block_baseline(Component& c, int16_t* d);
    bool block_dc_progressive(Component& c, int16_t* d);
    bool block_ac_progressive(Component& c, int16_t* d);
    bool scan();
};

// One block of a sequential scan: coefficient * quantiser, truncated to 16 bits, in natural order.
bool Decoder::block_baseline(Component& c, int16_t* d) {
    const uint16_t* q = dequant[c.tq];
    std::memset(d, 0, 64 * sizeof(int16_t));
    const int t = dc[c.hd].decode(br);
    if (t < 0 || t > 15) return false;
    c.dc_pred += extend(br, t);
    if (!fits_short((int64_t)c.dc_pred * q[0])) return false;          // stb_image refuses such a stream too (:2224)
    d[0] = (int16_t)(c.dc_pred * q[0]);
    for (int k = 1; k < 64;) {
        const int rs = ac[c.ha].decode(br);
        if (rs < 0) return false;
        const int s = rs & 15, r = rs >> 4;
        if (s == 0) { if (rs != 0xF0) break; k += 16; continue; }
        k += r;
        if (k > 63) return false;
        const int zz = kZigzag[k++];
        d[zz] = (int16_t)(extend(br, s) * q[zz]);
    }
    return true;
}

bool Decoder::block_dc_progressive(Component& c, int16_t* d) {
    if (spec_end != 0) return false;
    if (succ_high == 0) {
        std::memset(d, 0, 64 * sizeof(int16_t));
        const int t = dc[c.hd].decode(br);
        if (t < 0 || t > 15) return false;
        c.dc_pred += extend(br, t);
        if (!fits_short((int64_t)c.dc_pred * (1 << succ_low))) return false;
        d[0] = (int16_t)(c.dc_pred * (1 << succ_low));
    } else if (br.bit()) {
        d[0] = (int16_t)(d[0] + (int16_t)(1 << succ_low));
    }
    return true;
}

bool Decoder::block_ac_progressive(Component& c, int16_t* d) {
    if (spec_start == 0) return false;
    if (succ_high == 0) {                                   // first pass over this band
        if (eob_run) { --eob_run; return true; }
        int k = spec_start;
        do {
            const int rs = ac[c.ha].decode(br);
            if (rs < 0) return false;
            const int s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (r < 15) { eob_run = (1 << r); if (r) eob_run += br.bits(r); --eob_run; break; }
                k += 16;
            } else {
                k += r;
                if (k > 63) return false;
                d[kZigzag[k++]] = (int16_t)(extend(br, s) * (1 << succ_low));
            }
        } while (k <= spec_end);
        return true;
    }
    const int16_t bit = (int16_t)(1 << succ_low);           // refinement pass
    auto refine = [&](int16_t* p) {
        if (br.bit() && (*p & bit) == 0) *p = (int16_t)(*p > 0 ? *p + bit : *p - bit);
    };
    if (eob_run) {
        --eob_run;
        for (int k = spec_start; k <= spec_end; ++k) { int16_t* p = &d[kZigzag[k]]; if (*p != 0) refine(p); }
        return true;
    }
    int k = spec_start;
    do {
        const int rs = ac[c.ha].decode(br);
        if (rs < 0) return false;
        int s = rs & 15, r = rs >> 4;
        if (s == 0) {
            if (r < 15) { eob_run = (1 << r) - 1; if (r) eob_run += br.bits(r); r = 64; }     // run to the end of the band
        } else {
            if (s != 1) return false;
            s = br.bit() ? bit : -bit;
        }
        while (k <= spec_end) {                              // skip r zero coefficients, refining the non-zero ones passed
            int16_t* p = &d[kZigzag[k++]];
            if (*p != 0) refine(p);
            else { if (r == 0) { *p = (int16_t)s; break; } --r; }
        }
    } while (k <= spec_end);
    return true;
}

// ---- inverse DCT (stb_image.h:2452-2523) ----------------------------------------------------------
inline uint8_t clamp255(int64_t x) { return (uint8_t)(x < 0 ? 0 : x > 255 ? 255 : x); }
constexpr int fix(double x) { return (int)(x * 4096 + 0.5); }

// 64-bit intermediates: the same values as stb_image's 32-bit ones on every valid stream, and defined behaviour (no
// signed overflow) on damaged ones.
typedef int64_t wide;
struct Butterfly { wide x0, x1, x2, x3, t0, t1, t2, t3; };
inline Butterfly butterfly(wide s0, wide s1, wide s2, wide s3, wide s4, wide s5, wide s6, wide s7) {
    Butterfly b;
    wide p1 = (s2 + s6) * fix(0.5411961f);
    wide t2 = p1 + s6 * fix(-1.847759065f);
    wide t3 = p1 + s2 * fix(0.765366865f);
    wide t0 = (s0 + s4) * 4096, t1 = (s0 - s4) * 4096;
    b.x0 = t0 + t3; b.x3 = t0 - t3; b.x1 = t1 + t2; b.x2 = t1 - t2;
    t0 = s7; t1 = s5; t2 = s3; t3 = s1;
    wide p3 = t0 + t2, p4 = t1 + t3;
    p1 = t0 + t3;
    wide p2 = t1 + t2;
    const wide p5 = (p3 + p4) * fix(1.175875602f);
    t0 = t0 * fix(0.298631336f); t1 = t1 * fix(2.053119869f); t2 = t2 * fix(3.072711026f); t3 = t3 * fix(1.501321110f);
    p1 = p5 + p1 * fix(-0.899976223f); p2 = p5 + p2 * fix(-2.562915447f);
    p3 = p3 * fix(-1.961570560f); p4 = p4 * fix(-0.390180644f);
    b.t3 = t3 + p1 + p4; b.t2 = t2 + p2 + p3; b.t1 = t1 + p2 + p4; b.t0 = t0 + p1 + p3;
    return b;
}

void idct_block(uint8_t* out, int stride, const int16_t* d) {
    wide val[64];
    for (int i = 0; i < 8; ++i) {
        const int16_t* c = d + i; wide* v = val + i;
        if (c[8] == 0 && c[16] == 0 && c[24] == 0 && c[32] == 0 && c[40] == 0 && c[48] == 0 && c[56] == 0) {
            const int dcterm = c[0] * 4;
            for (int r = 0; r < 8; ++r) v[8 * r] = dcterm;
        } else {
            Butterfly b = butterfly(c[0], c[8], c[16], c[24], c[32], c[40], c[48], c[56]);
            b.x0 += 512; b.x1 += 512; b.x2 += 512; b.x3 += 512;
            v[0] = (b.x0 + b.t3) >> 10; v[56] = (b.x0 - b.t3) >> 10;
            v[8] = (b.x1 + b.t2) >> 10; v[48] = (b.x1 - b.t2) >> 10;
            v[16] = (b.x2 + b.t1) >> 10; v[40] = (b.x2 - b.t1) >> 10;
            v[24] = (b.x3 + b.t0) >> 10; v[32] = (b.x3 - b.t0) >> 10;
        }
    }
    for (int i = 0; i < 8; ++i) {
        const wide* v = val + 8 * i; uint8_t* o = out + (size_t)stride * i;
        Butterfly b = butterfly(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
        const wide bias = 65536 + (128 << 17);
        b.x0 += bias; b.x1 += bias; b.x2 += bias; b.x3 += bias;
        o[0] = clamp255((b.x0 + b.t3) >> 17); o[7] = clamp255((b.x0 - b.t3) >> 17);
        o[1] = clamp255((b.x1 + b.t2) >> 17); o[6] = clamp255((b.x1 - b.t2) >> 17);
        o[2] = clamp255((b.x2 + b.t1) >> 17); o[5] = clamp255((b.x2 - b.t1) >> 17);
        o[3] = clamp255((b.x3 + b.t0) >> 17); o[4] = clamp255((b.x3 - b.t0) >> 17);
    }
}

// ---- one scan (stb_image.h:2949-3059) -------------------------------------------------------------
// Returns false on a damaged stream.  A restart interval that is not followed by an RSTn marker ends the scan.
bool Decoder::scan() {
    int todo = restart_interval ? restart_interval : 0x7fffffff;
    eob_run = 0;
    for (int k = 0; k < 4; ++k) comp[k].dc_pred = 0;
    int16_t block[64];
    auto unit_done = [&]() -> bool {                         // true: the scan goes on
        if (--todo > 0) return true;
        if (br.have < 24) br.fill();
        if (br.marker < 0xD0 || br.marker > 0xD7) return false;
        br.start(br.at);
        for (int k = 0; k < 4; ++k) comp[k].dc_pred = 0;
        eob_run = 0;
        todo = restart_interval ? restart_interval : 0x7fffffff;
        return true;
    };
    auto one_block = [&](Component& c, int bx, int by) -> bool {
        if (!progressive) {
            if (!block_baseline(c, block)) return false;
            idct_block(c.data.data() + (size_t)c.w2 * by * 8 + bx * 8, c.w2, block);
            return true;
        }
        int16_t* d = c.coeff.data() + 64 * ((size_t)bx + (size_t)by * (c.w2 / 8));
        return spec_start == 0 ? block_dc_progressive(c, d) : block_ac_progressive(c, d);
    };
    if (scan_n == 1) {                                       // non-interleaved: the component's own block grid
        Component& c = comp[order[0]];
        const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
        for (int j = 0; j < h; ++j)
            for (int i = 0; i < w; ++i) {
                if (!one_block(c, i, j)) return false;
                if (!unit_done()) return true;
            }
        return true;
    }
    if (progressive && spec_start != 0) return false;        // AC scans are never interleaved
    for (int j = 0; j < mcu_y; ++j)
        for (int i = 0; i < mcu_x; ++i) {
            for (int k = 0; k < scan_n; ++k) {
                Component& c = comp[order[k]];
                for (int y = 0; y < c.v; ++y)
                    for (int x = 0; x < c.h; ++x)
                        if (!one_block(c, i * c.h + x, j * c.v + y)) return false;
            }
            if (!unit_done()) return true;
        }
    return true;
}

// ---- upsampling (stb_image.h:3465-3527, :3646-3655) -------------------------------------------------
const uint8_t* upsample_row(uint8_t* out, const uint8_t* near, const uint8_t* far, int w, int hs, int vs) {
    if (hs == 1 && vs == 1) return near;
    if (hs == 1 && vs == 2) { for (int i = 0; i < w; ++i) out[i] = (uint8_t)((3 * near[i] + far[i] + 2) >> 2); return out; }
    if (hs == 2 && vs == 1) {
        if (w == 1) { out[0] = out[1] = near[0]; return out; }
        out[0] = near[0];
        out[1] = (uint8_t)((near[0] * 3 + near[1] + 2) >> 2);
        int i;
        for (i = 1; i < w - 1; ++i) {
            const int n = 3 * near[i] + 2;
            out[i * 2 + 0] = (uint8_t)((n + near[i - 1]) >> 2);
            out[i * 2 + 1] = (uint8_t)((n + near[i + 1]) >> 2);
        }
        out[i * 2 + 0] = (uint8_t)((near[w - 2] * 3 + near[w - 1] + 2) >> 2);
        out[i * 2 + 1] = near[w - 1];
        return out;
    }
    if (hs == 2 && vs == 2) {
        if (w == 1) { out[0] = out[1] = (uint8_t)((3 * near[0] + far[0] + 2) >> 2); return out; }
        int t1 = 3 * near[0] + far[0];
        out[0] = (uint8_t)((t1 + 2) >> 2);
        for (int i = 1; i < w; ++i) {
            const int t0 = t1;
            t1 = 3 * near[i] + far[i];
            out[i * 2 - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
            out[i * 2] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
        }
        out[w * 2 - 1] = (uint8_t)((t1 + 2) >> 2);
        return out;
    }
    for (int i = 0; i < w; ++i) for (int j = 0; j < hs; ++j) out[i * hs + j] = near[i];
    return out;
}

// ---- colour (stb_image.h:3660-3683) ---------------------------------------------------------------
constexpr int fix20(float x) { return ((int)(x * 4096.0f + 0.5f)) << 8; }
inline void ycc_to_rgb(uint8_t* out, int y, int cb, int cr) {
    const int y_fixed = (y << 20) + (1 << 19);
    cr -= 128; cb -= 128;
    int r = y_fixed + cr * fix20(1.40200f);
    int g = y_fixed + (cr * -fix20(0.71414f)) + (int)((unsigned)(cb * -fix20(0.34414f)) & 0xffff0000u);
    int b = y_fixed + cb * fix20(1.77200f);
    r >>= 20; g >>= 20; b >>= 20;
    out[0] = clamp255(r); out[1] = clamp255(g); out[2] = clamp255(b);
}

inline int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

}  // namespace

bool decode_jpeg(const uint8_t* data, size_t n, int* width, int* height, std::vector<uint8_t>& rgb) {
    if (n < 4 || data[0] != 0xFF || data[1] != 0xD8) return false;
    Decoder z;
    z.br.p = data; z.br.n = n;
    std::memset(z.dequant, 0, sizeof z.dequant);
    bool have_frame = false;
    size_t at = 2;
    int pending = 0;                                         // a marker already consumed by the entropy reader
    for (;;) {
        int m = pending;
        pending = 0;
        if (!m) {                                            // next marker: 0xFF (fill bytes) then a non-zero code
            while (at < n && data[at] != 0xFF) ++at;
            while (at < n && data[at] == 0xFF) ++at;
            if (at >= n) return false;
            m = data[at++];
            if (m == 0) continue;
        }
        if (m == 0xD9) break;                                // EOI
        if (m >= 0xD0 && m <= 0xD7) continue;                // a stray RSTn between segments
        if (at + 2 > n) return false;
        const int len = be16(data + at);
        if (len < 2 || at + len > n) return false;
        const uint8_t* seg = data + at + 2;
        const int body = len - 2;
        if (m == 0xC0 || m == 0xC1 || m == 0xC2) {           // frame header
            if (have_frame || body < 6) return false;
            z.progressive = m == 0xC2;
            z.height = be16(seg + 1); z.width = be16(seg + 3); z.ncomp = seg[5];
            if (seg[0] != 8 || z.width == 0 || z.height == 0 || (z.ncomp != 1 && z.ncomp != 3)) return false;
            if ((int64_t)z.width * z.height > ((int64_t)1 << 28)) return false;
            if (body != 6 + 3 * z.ncomp) return false;
            for (int i = 0; i < z.ncomp; ++i) {
                Component& c = z.comp[i];
                c.id = seg[6 + 3 * i]; c.h = seg[7 + 3 * i] >> 4; c.v = seg[7 + 3 * i] & 15; c.tq = seg[8 + 3 * i];
                if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) return false;
                if (z.ncomp == 3 && c.id == "RGB"[i]) ++z.rgb_ids;
                if (c.h > z.h_max) z.h_max = c.h;
                if (c.v > z.v_max) z.v_max = c.v;
            }
            for (int i = 0; i < z.ncomp; ++i) if (z.h_max % z.comp[i].h || z.v_max % z.comp[i].v) return false;
            z.mcu_x = (z.width + z.h_max * 8 - 1) / (z.h_max * 8);
            z.mcu_y = (z.height + z.v_max * 8 - 1) / (z.v_max * 8);
            for (int i = 0; i < z.ncomp; ++i) {
                Component& c = z.comp[i];
                c.x = (z.width * c.h + z.h_max - 1) / z.h_max;
                c.y = (z.height * c.v + z.v_max - 1) / z.v_max;
                c.w2 = z.mcu_x * c.h * 8; c.h2 = z.mcu_y * c.v * 8;
                c.data.assign((size_t)c.w2 * c.h2, 0);
                if (z.progressive) c.coeff.assign((size_t)c.w2 * c.h2, 0);
            }
            have_frame = true;
        } else if (m == 0xC4) {                              // Huffman tables
            int left = body; const uint8_t* p = seg;
            while (left > 0) {
                if (left < 17) return false;
                const int tc = p[0] >> 4, th = p[0] & 15;
                int counts[16], total = 0;
                for (int i = 0; i < 16; ++i) { counts[i] = p[1 + i]; total += counts[i]; }
                if (tc > 1 || th > 3 || total > 256 || left < 17 + total) return false;
                if (!(tc ? z.ac[th] : z.dc[th]).build(counts, p + 17, total)) return false;
                p += 17 + total; left -= 17 + total;
            }
        } else if (m == 0xDB) {                              // quantisers, stored in natural order
            int left = body; const uint8_t* p = seg;
            while (left > 0) {
                const int wide = p[0] >> 4, t = p[0] & 15, need = 1 + (wide ? 128 : 64);
                if (wide > 1 || t > 3 || left < need) return false;
                for (int i = 0; i < 64; ++i) z.dequant[t][kZigzag[i]] = (uint16_t)(wide ? be16(p + 1 + 2 * i) : p[1 + i]);
                p += need; left -= need;
            }
        } else if (m == 0xDD) {
            if (body != 2) return false;
            z.restart_interval = be16(seg);
        } else if (m == 0xE0 && body >= 5) {
            if (!std::memcmp(seg, "JFIF\0", 5)) z.jfif = true;
        } else if (m == 0xEE && body >= 12) {
            if (!std::memcmp(seg, "Adobe\0", 6)) z.adobe_transform = seg[11];
        } else if (m == 0xDA) {                              // scan header + entropy-coded data
            if (!have_frame || body < 1) return false;
            z.scan_n = seg[0];
            if (z.scan_n < 1 || z.scan_n > z.ncomp || body != 4 + 2 * z.scan_n) return false;
            for (int i = 0; i < z.scan_n; ++i) {
                int which = 0;
                while (which < z.ncomp && z.comp[which].id != seg[1 + 2 * i]) ++which;
                if (which == z.ncomp) return false;
                z.comp[which].hd = seg[2 + 2 * i] >> 4; z.comp[which].ha = seg[2 + 2 * i] & 15;
                if (z.comp[which].hd > 3 || z.comp[which].ha > 3) return false;
                z.order[i] = which;
            }
            const uint8_t* tail = seg + 1 + 2 * z.scan_n;
            z.spec_start = tail[0]; z.spec_end = tail[1]; z.succ_high = tail[2] >> 4; z.succ_low = tail[2] & 15;
            if (z.progressive) {
                if (z.spec_start > 63 || z.spec_end > 63 || z.spec_start > z.spec_end || z.succ_high > 13 || z.succ_low > 13) return false;
            } else {
                if (z.spec_start != 0 || z.succ_high != 0 || z.succ_low != 0) return false;
                z.spec_end = 63;
            }
            z.br.start(at + len);
            if (!z.scan()) return false;
            at = z.br.at;
            pending = z.br.marker;                           // 0: the reader stopped inside data; look for the next marker
            continue;
        } else if ((m >= 0xE0 && m <= 0xEF) || m == 0xFE || m == 0xDC) {
            // other application segments, comments, DNL: skipped
        } else {
            return false;                                    // arithmetic coding, lossless, hierarchical ...: not covered
        }
        at += len;
    }
    if (!have_frame) return false;
    if (z.progressive)                                       // stb_image.h:3080-3096: dequantise and transform the visible blocks
        for (int k = 0; k < z.ncomp; ++k) {
            Component& c = z.comp[k];
            const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
            for (int j = 0; j < h; ++j)
                for (int i = 0; i < w; ++i) {
                    int16_t* d = c.coeff.data() + 64 * ((size_t)i + (size_t)j * (c.w2 / 8));
                    for (int e = 0; e < 64; ++e) d[e] = (int16_t)(d[e] * z.dequant[c.tq][e]);
                    idct_block(c.data.data() + (size_t)c.w2 * j * 8 + i * 8, c.w2, d);
                }
        }
    // rows: every component upsampled to full width, then converted (stb_image.h:3889-3990)
    const bool is_rgb = z.ncomp == 3 && (z.rgb_ids == 3 || (z.adobe_transform == 0 && !z.jfif));
    struct Walk { int hs, vs, ystep, w_lores, ypos; const uint8_t* line0; const uint8_t* line1; std::vector<uint8_t> buf; } walk[3];
    for (int k = 0; k < z.ncomp; ++k) {
        Walk& r = walk[k];
        r.hs = z.h_max / z.comp[k].h; r.vs = z.v_max / z.comp[k].v;
        r.ystep = r.vs >> 1; r.w_lores = (z.width + r.hs - 1) / r.hs; r.ypos = 0;
        r.line0 = r.line1 = z.comp[k].data.data();
        r.buf.assign((size_t)z.width + 8, 0);
    }
    rgb.resize((size_t)z.width * z.height * 3);
    for (int j = 0; j < z.height; ++j) {
        const uint8_t* row[3] = {nullptr, nullptr, nullptr};
        for (int k = 0; k < z.ncomp; ++k) {
            Walk& r = walk[k];
            const bool bottom = r.ystep >= (r.vs >> 1);
            row[k] = upsample_row(r.buf.data(), bottom ? r.line1 : r.line0, bottom ? r.line0 : r.line1, r.w_lores, r.hs, r.vs);
            if (++r.ystep >= r.vs) {
                r.ystep = 0;
                r.line0 = r.line1;
                if (++r.ypos < z.comp[k].y) r.line1 += z.comp[k].w2;
            }
        }
        uint8_t* out = rgb.data() + (size_t)j * z.width * 3;
        for (int i = 0; i < z.width; ++i, out += 3) {
            if (z.ncomp == 1) out[0] = out[1] = out[2] = row[0][i];
            else if (is_rgb) { out[0] = row[0][i]; out[1] = row[1][i]; out[2] = row[2][i]; }
            else ycc_to_rgb(out, row[0][i], row[1][i], row[2][i]);
        }
    }
    if (width) *width = z.width;
    if (height) *height = z.height;
    return true;
}

}  // namespace srt

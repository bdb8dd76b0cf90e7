// srt_png.cpp — PNG files to the RGB8 buffer stbi_load(path, &w, &h, &ch, 3) returns for them (Object.cpp:57; the
// reference decodes textures with its vendored stb_image 2.30).
//
// PNG is lossless, so "the same pixels" only needs the format rules plus stb_image's conversions to three channels:
//   * colour types 0 / 2 / 3 / 4 / 6, bit depths 1-16, Adam7 interlacing, filters 0-4 (PNG specification);
//   * 16-bit samples keep their high byte (stbi__convert_16_to_8), grey samples of depth 1 / 2 / 4 are scaled by
//     0xff / 0x55 / 0x11 (stbi__depth_scale_table), palette entries are looked up, alpha (channel or tRNS) is dropped,
//     grey is replicated into R, G, B (stbi__convert_format to req_comp = 3);
//   * zlib stream = concatenated IDAT chunks, inflate with stored / fixed / dynamic blocks (RFC 1950 / 1951).
// CRCs and the Adler-32 are not checked (stb_image does not check them either).  Own implementation; tested against
// the reference's decoder on its texture files and on generated images of every colour type (tests/test_host_scene.py).
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "srt_scene.hpp"

namespace srt {
namespace {

// ---- inflate (RFC 1951) -------------------------------------------------------------------------
struct BitReader {
    const uint8_t* p; size_t n, at = 0;
    uint32_t hold = 0; int have = 0;
    bool overrun = false;
    uint32_t bits(int k) {
        while (have < k) {
            uint32_t b = 0;
            if (at < n) b = p[at++]; else overrun = true;
            hold |= b << have; have += 8;
        }
        const uint32_t v = k ? (hold & ((1u << k) - 1u)) : 0u;
        hold >>= k; have -= k;
        return v;
    }
    void align_byte() { hold = 0; have = 0; }
};

struct Huffman {                 // canonical code: symbols ordered by (length, value)
    uint16_t count[16]; uint16_t symbol[288];
    bool build(const uint8_t* len, int n) {
        std::memset(count, 0, sizeof count);
        for (int i = 0; i < n; ++i) ++count[len[i]];
        count[0] = 0;
        int left = 1;
        for (int l = 1; l < 16; ++l) { left = (left << 1) - count[l]; if (left < 0) return false; }
        uint16_t offs[16]; offs[1] = 0;
        for (int l = 1; l < 15; ++l) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
        for (int i = 0; i < n; ++i) if (len[i]) symbol[offs[len[i]]++] = (uint16_t)i;
        return true;
    }
    int decode(BitReader& br) const {
        int code = 0, first = 0, index = 0;
        for (int l = 1; l < 16; ++l) {
            code |= (int)br.bits(1);
            const int c = count[l];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
};

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

bool inflate_block(BitReader& br, const Huffman& lit, const Huffman& dist, std::vector<uint8_t>& out, size_t limit) {
    for (;;) {
        const int s = lit.decode(br);
        if (s < 0 || br.overrun || out.size() > limit) return false;
        if (s < 256) { out.push_back((uint8_t)s); continue; }
        if (s == 256) return true;
        if (s > 285) return false;
        const int len = kLenBase[s - 257] + (int)br.bits(kLenExtra[s - 257]);
        const int d = dist.decode(br);
        if (d < 0 || d > 29) return false;
        const size_t back = kDistBase[d] + br.bits(kDistExtra[d]);
        if (back > out.size()) return false;
        const size_t from = out.size() - back;
        for (int k = 0; k < len; ++k) out.push_back(out[from + k]);
    }
}

// `limit`: the most bytes the caller can use; a stream that produces more is refused (no decompression bombs).
bool zlib_inflate(const uint8_t* data, size_t n, std::vector<uint8_t>& out, size_t limit) {
    if (n < 2 || (data[0] & 15) != 8 || ((data[0] << 8) | data[1]) % 31 != 0 || (data[1] & 32)) return false;
    BitReader br{data + 2, n - 2};
    Huffman lit, dist;
    for (;;) {
        const uint32_t last = br.bits(1), type = br.bits(2);
        if (type == 0) {
            br.align_byte();
            if (br.at + 4 > br.n) return false;
            const uint32_t len = br.p[br.at] | (br.p[br.at + 1] << 8), nlen = br.p[br.at + 2] | (br.p[br.at + 3] << 8);
            br.at += 4;
            if ((len ^ 0xffffu) != nlen || br.at + len > br.n || out.size() + len > limit) return false;
            out.insert(out.end(), br.p + br.at, br.p + br.at + len);
            br.at += len;
        } else if (type == 1) {
            uint8_t l[288];
            for (int i = 0; i < 288; ++i) l[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
            uint8_t d5[30]; std::memset(d5, 5, sizeof d5);
            lit.build(l, 288); dist.build(d5, 30);
            if (!inflate_block(br, lit, dist, out, limit)) return false;
        } else if (type == 2) {
            const int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
            static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            uint8_t cl[19] = {0};
            for (int i = 0; i < ncode; ++i) cl[order[i]] = (uint8_t)br.bits(3);
            Huffman code;
            if (nlen > 286 || ndist > 30 || !code.build(cl, 19)) return false;
            uint8_t lens[320];
            int i = 0;
            while (i < nlen + ndist) {
                const int s = code.decode(br);
                if (s < 0 || br.overrun) return false;
                if (s < 16) { lens[i++] = (uint8_t)s; continue; }
                int rep; uint8_t v = 0;
                if (s == 16) { if (i == 0) return false; v = lens[i - 1]; rep = 3 + (int)br.bits(2); }
                else if (s == 17) rep = 3 + (int)br.bits(3);
                else rep = 11 + (int)br.bits(7);
                if (i + rep > nlen + ndist) return false;
                while (rep--) lens[i++] = v;
            }
            if (!lit.build(lens, nlen) || !dist.build(lens + nlen, ndist)) return false;
            if (!inflate_block(br, lit, dist, out, limit)) return false;
        } else {
            return false;
        }
        if (br.overrun) return false;
        if (last) return true;
    }
}

// ---- PNG ----------------------------------------------------------------------------------------
inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// One (sub)image of `w` x `h` pixels from the filtered scanlines at `src`: undo the filters in place, then write one
// byte per sample (high byte of 16-bit samples, low-depth samples unpacked, grey scaled) to dst[(y*w + x)*channels + c].
bool unfilter_image(uint8_t* src, size_t avail, uint32_t w, uint32_t h, int channels, int depth, bool scale_grey,
                    std::vector<uint8_t>& dst) {
    const size_t row_bytes = ((size_t)w * channels * depth + 7) / 8;
    const int bpp = channels * depth >= 8 ? channels * depth / 8 : 1;
    if (avail < (row_bytes + 1) * h) return false;
    dst.assign((size_t)w * h * channels, 0);
    const uint8_t* prev = nullptr;
    for (uint32_t y = 0; y < h; ++y) {
        uint8_t* row = src + (row_bytes + 1) * y;
        const int filter = row[0];
        uint8_t* cur = row + 1;
        if (filter > 4) return false;
        for (size_t i = 0; i < row_bytes; ++i) {
            const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = prev ? prev[i] : 0, c = (prev && i >= (size_t)bpp) ? prev[i - bpp] : 0;
            int v = cur[i];
            switch (filter) {
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: break;
            }
            cur[i] = (uint8_t)v;
        }
        prev = cur;
        uint8_t* out = dst.data() + (size_t)y * w * channels;
        const size_t samples = (size_t)w * channels;
        if (depth == 8) std::memcpy(out, cur, samples);
        else if (depth == 16) { for (size_t i = 0; i < samples; ++i) out[i] = cur[2 * i]; }
        else {
            static const uint8_t scale_of[5] = {0, 0xff, 0x55, 0, 0x11};
            const int scale = scale_grey ? scale_of[depth] : 1, mask = (1 << depth) - 1;
            for (size_t i = 0; i < samples; ++i) {
                const size_t bit = i * depth;
                out[i] = (uint8_t)(((cur[bit >> 3] >> (8 - depth - (bit & 7))) & mask) * scale);
            }
        }
    }
    return true;
}

}  // namespace

bool decode_png(const uint8_t* data, size_t n, int* width, int* height, std::vector<uint8_t>& rgb) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (n < 8 || std::memcmp(data, sig, 8) != 0) return false;
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat, palette;
    bool seen_end = false;
    for (size_t at = 8; at + 8 <= n && !seen_end;) {
        const uint32_t len = be32(data + at);
        const uint8_t* type = data + at + 4;
        const uint8_t* body = data + at + 8;
        if (len > n - at - 8) return false;
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len != 13 || ctype >= 0) return false;
            w = be32(body); h = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
            if (w == 0 || h == 0 || w > (1u << 24) || h > (1u << 24) || (uint64_t)w * h > (1ull << 28)) return false;
            if (body[10] != 0 || body[11] != 0 || interlace > 1) return false;
            if (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16) return false;
            if (ctype != 0 && ctype != 2 && ctype != 3 && ctype != 4 && ctype != 6) return false;
            if (ctype == 3 && depth == 16) return false;
            if ((ctype == 2 || ctype == 4 || ctype == 6) && depth < 8) return false;
        } else if (ctype < 0) {
            return false;                                   // IHDR comes first
        } else if (!std::memcmp(type, "PLTE", 4)) {
            if (len > 256 * 3 || len % 3) return false;
            palette.assign(body, body + len);
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            seen_end = true;
        }
        at += 12 + (size_t)len;
    }
    if (ctype < 0 || idat.empty() || (ctype == 3 && palette.empty())) return false;
    std::vector<uint8_t> raw;
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;
    const size_t most = (((size_t)w * channels * depth + 7) / 8 + 1) * h + (interlace ? 7 * (size_t)h + 64 : 0);   // Adam7: a filter byte per sub-row
    if (!zlib_inflate(idat.data(), idat.size(), raw, most + 258)) return false;
    std::vector<uint8_t> samples((size_t)w * h * channels), part;
    if (!interlace) {
        if (!unfilter_image(raw.data(), raw.size(), w, h, channels, depth, ctype == 0, samples)) return false;
    } else {                                                // Adam7: seven sub-images, each filtered on its own
        static const int x0[7] = {0, 4, 0, 2, 0, 1, 0}, y0[7] = {0, 0, 4, 0, 2, 0, 1}, dx[7] = {8, 8, 4, 4, 2, 2, 1}, dy[7] = {8, 8, 8, 4, 4, 2, 2};
        size_t used = 0;
        for (int p = 0; p < 7; ++p) {
            if ((int64_t)w <= x0[p] || (int64_t)h <= y0[p]) continue;
            const uint32_t pw = (w - x0[p] + dx[p] - 1) / dx[p], ph = (h - y0[p] + dy[p] - 1) / dy[p];
            if (pw == 0 || ph == 0) continue;
            if (!unfilter_image(raw.data() + used, raw.size() - used, pw, ph, channels, depth, ctype == 0, part)) return false;
            used += (((size_t)pw * channels * depth + 7) / 8 + 1) * ph;
            for (uint32_t y = 0; y < ph; ++y)
                for (uint32_t x = 0; x < pw; ++x)
                    std::memcpy(&samples[(((size_t)y * dy[p] + y0[p]) * w + (size_t)x * dx[p] + x0[p]) * channels],
                                &part[((size_t)y * pw + x) * channels], channels);
        }
    }
    rgb.resize((size_t)w * h * 3);
    const size_t pixels = (size_t)w * h;
    for (size_t i = 0; i < pixels; ++i) {
        const uint8_t* s = &samples[i * channels];
        uint8_t* d = &rgb[3 * i];
        if (ctype == 3) {
            const size_t e = (size_t)s[0] * 3;
            if (e + 2 < palette.size()) { d[0] = palette[e]; d[1] = palette[e + 1]; d[2] = palette[e + 2]; }
            else { d[0] = d[1] = d[2] = 0; }
        } else if (channels >= 3) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; }
        else { d[0] = d[1] = d[2] = s[0]; }
    }
    if (width) *width = (int)w;
    if (height) *height = (int)h;
    return true;
}

// ---- PNG output ("optional PNG" beside the reference's BMP) ---------------------------------
// 8-bit RGB, non-interlaced; every row takes the filter (0-4) with the smallest sum of absolute residuals; one zlib
// stream of fixed-Huffman blocks with a greedy hash-chain LZ77 (window 32 KiB).  Lossless, so any decoder returns the
// frame bytes; not a reference format (the reference saves BMP, drawImage :488-494).
namespace {

struct BitWriter {
    std::vector<uint8_t>& out; uint32_t hold = 0; int have = 0;
    explicit BitWriter(std::vector<uint8_t>& o) : out(o) {}
    void put(uint32_t v, int n) { hold |= v << have; have += n; while (have >= 8) { out.push_back((uint8_t)hold); hold >>= 8; have -= 8; } }
    void put_code(uint32_t code, int n) { uint32_t r = 0; for (int i = 0; i < n; ++i) r |= ((code >> i) & 1u) << (n - 1 - i); put(r, n); }   // Huffman codes go MSB first
    void flush() { if (have) { out.push_back((uint8_t)hold); hold = 0; have = 0; } }
};

void put_fixed_literal(BitWriter& bw, int s) {               // RFC 1951 3.2.6
    if (s < 144) bw.put_code(0x30 + s, 8);
    else if (s < 256) bw.put_code(0x190 + (s - 144), 9);
    else if (s < 280) bw.put_code(s - 256, 7);
    else bw.put_code(0xC0 + (s - 280), 8);
}

void deflate_fixed(const std::vector<uint8_t>& in, std::vector<uint8_t>& out) {
    BitWriter bw(out);
    bw.put(1, 1); bw.put(1, 2);                              // one final block, fixed codes
    const size_t n = in.size();
    const int kHashBits = 15, kChain = 24;
    std::vector<int32_t> head((size_t)1 << kHashBits, -1), prev(n, -1);
    auto hash = [&](size_t i) { return ((uint32_t)(in[i] | (in[i + 1] << 8) | (in[i + 2] << 16)) * 2654435761u) >> (32 - kHashBits); };
    size_t i = 0;
    while (i < n) {
        int best_len = 0; size_t best_dist = 0;
        if (i + 3 <= n) {
            const uint32_t h = hash(i);
            int32_t c = head[h];
            for (int steps = 0; c >= 0 && steps < kChain && i - (size_t)c <= 32768; c = prev[c], ++steps) {
                int l = 0;
                const int lim = (int)std::min<size_t>(258, n - i);
                while (l < lim && in[c + l] == in[i + l]) ++l;
                if (l > best_len) { best_len = l; best_dist = i - (size_t)c; if (l == lim) break; }
            }
        }
        if (best_len >= 3) {
            int ls = 28; while (kLenBase[ls] > best_len) --ls;
            put_fixed_literal(bw, 257 + ls);
            bw.put((uint32_t)(best_len - kLenBase[ls]), kLenExtra[ls]);
            int ds = 29; while (kDistBase[ds] > best_dist) --ds;
            bw.put_code((uint32_t)ds, 5);
            bw.put((uint32_t)(best_dist - kDistBase[ds]), kDistExtra[ds]);
        } else {
            put_fixed_literal(bw, in[i]);
            best_len = 1;
        }
        for (int k = 0; k < best_len; ++k, ++i)
            if (i + 3 <= n) { const uint32_t h = hash(i); prev[i] = head[h]; head[h] = (int32_t)i; }
    }
    put_fixed_literal(bw, 256);
    bw.flush();
}

uint32_t crc32_of(const uint8_t* p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256]; static bool ready = false;
    if (!ready) { for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; } ready = true; }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 255] ^ (crc >> 8);
    return ~crc;
}

void put_chunk(std::vector<uint8_t>& file, const char* tag, const std::vector<uint8_t>& body) {
    auto be = [&](uint32_t v) { for (int k = 3; k >= 0; --k) file.push_back((uint8_t)(v >> (8 * k))); };
    be((uint32_t)body.size());
    const size_t from = file.size();
    file.insert(file.end(), tag, tag + 4);
    file.insert(file.end(), body.begin(), body.end());
    be(crc32_of(file.data() + from, file.size() - from));
}

}  // namespace

bool encode_png(const uint8_t* rgb, int width, int height, std::vector<uint8_t>& file) {
    if (!rgb || width <= 0 || height <= 0 || (int64_t)width * height > ((int64_t)1 << 28)) return false;
    const size_t row = (size_t)width * 3;
    std::vector<uint8_t> raw; raw.reserve((row + 1) * height);
    std::vector<uint8_t> cand(row), best(row);
    for (int y = 0; y < height; ++y) {
        const uint8_t* cur = rgb + row * y; const uint8_t* up = y ? cur - row : nullptr;
        long best_cost = -1; int best_f = 0;
        for (int f = 0; f < 5; ++f) {
            long cost = 0;
            for (size_t i = 0; i < row; ++i) {
                const int a = i >= 3 ? cur[i - 3] : 0, b = up ? up[i] : 0, c = (up && i >= 3) ? up[i - 3] : 0;
                const int pred = f == 0 ? 0 : f == 1 ? a : f == 2 ? b : f == 3 ? (a + b) >> 1 : paeth(a, b, c);
                cand[i] = (uint8_t)(cur[i] - pred);
                cost += cand[i] < 128 ? cand[i] : 256 - cand[i];
            }
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_f = f; best.swap(cand); }
        }
        raw.push_back((uint8_t)best_f);
        raw.insert(raw.end(), best.begin(), best.end());
    }
    uint32_t s1 = 1, s2 = 0;                                 // Adler-32 of the uncompressed stream
    for (size_t i = 0; i < raw.size();) { const size_t end = std::min(raw.size(), i + 5552); for (; i < end; ++i) { s1 += raw[i]; s2 += s1; } s1 %= 65521; s2 %= 65521; }
    std::vector<uint8_t> z = {0x78, 0x01};
    deflate_fixed(raw, z);
    for (int k = 3; k >= 0; --k) z.push_back((uint8_t)(((s2 << 16) | s1) >> (8 * k)));
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    file.assign(sig, sig + 8);
    std::vector<uint8_t> hdr(13, 0);
    for (int k = 0; k < 4; ++k) { hdr[k] = (uint8_t)((uint32_t)width >> (8 * (3 - k))); hdr[4 + k] = (uint8_t)((uint32_t)height >> (8 * (3 - k))); }
    hdr[8] = 8; hdr[9] = 2;
    put_chunk(file, "IHDR", hdr);
    put_chunk(file, "IDAT", z);
    put_chunk(file, "IEND", {});
    return true;
}

}  // namespace srt

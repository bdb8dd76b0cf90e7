// srt_obj_reader.cpp — Wavefront OBJ / MTL reader of this repository.
//
// Own implementation.  It restates, without sharing code, the parsing RULES of tinyobjloader 2.0
// (the reference's loader, library/tinyobjloader/tiny_obj_loader.h in the reference tree) that
// decide the resulting triangle list, because parity needs the same floats and the same triangles
// in the same order:
//   * numbers: tinyobj's own decimal parser (tryParseDouble, tiny_obj_loader.h:897-1028), which is
//     NOT correctly rounded — integer part by Horner, fraction digit * 10^-k from a short table then
//     pow(10,-k), exponent via ldexp(m * pow(5,e), e) — then a cast to float,
//   * indices: 1-based, negative = relative to the count so far, 0 invalid for `v` (:819-848, :1188-1240),
//   * faces: < 3 vertices dropped; triangles kept; quads split along the SHORTER diagonal
//     (squared length, '<' picks 0-2, else 1-3, :1562-1606); polygons with more vertices are
//     triangulated by srt::triangulate_polygon (srt_polygon.cpp),
//   * the material of a face is the `usemtl` in force when its face group is flushed (:2850-2875),
//   * `mtllib` files are looked up next to the OBJ file (:3453-3475); `map_Kd` takes the rest of
//     the line after its options as the texture name (:1334-1346).
// Face order in the output is file order (groups/objects only split shapes, whose concatenation
// the reference walks in order, Object.cpp:70-166).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <set>
#include <sstream>

#include "srt_scene.hpp"

namespace srt {

bool triangulate_polygon(const std::vector<float>& v, const std::vector<ObjIndex>& face,
                         std::vector<ObjIndex>& out_triangles);   // srt_polygon.cpp

namespace {

inline bool is_digit(char c) { return c >= '0' && c <= '9'; }
inline bool is_space(char c) { return c == ' ' || c == '\t'; }
inline bool is_eol(char c) { return c == '\r' || c == '\n' || c == '\0'; }

}  // namespace

bool parse_obj_double(const char* s, const char* s_end, double* result) {
    if (s >= s_end) return false;
    double mantissa = 0.0;
    int exponent = 0;
    bool negative = false, exp_negative = false, leading_dot = false;
    const char* p = s;
    if (*p == '+' || *p == '-') {
        negative = (*p == '-');
        ++p;
        if (p != s_end && *p == '.') leading_dot = true;
    } else if (*p == '.') {
        leading_dot = true;
    } else if (!is_digit(*p)) {
        return false;
    }
    int read = 0;
    if (!leading_dot) {
        while (p != s_end && is_digit(*p)) { mantissa *= 10; mantissa += (int)(*p - '0'); ++p; ++read; }
        if (read == 0) return false;
    }
    bool more = (p != s_end);
    if (more && *p == '.') {
        static const double lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
        ++p;
        read = 1;
        while (p != s_end && is_digit(*p)) {
            mantissa += (int)(*p - '0') * (read < 8 ? lut[read] : std::pow(10.0, -read));
            ++read; ++p;
        }
        more = (p != s_end);
    } else if (more && !(*p == 'e' || *p == 'E')) {
        more = false;
    }
    if (more && (*p == 'e' || *p == 'E')) {
        ++p;
        if (p != s_end && (*p == '+' || *p == '-')) { exp_negative = (*p == '-'); ++p; }
        else if (p == s_end || !is_digit(*p)) return false;
        read = 0;
        while (p != s_end && is_digit(*p)) {
            if (exponent > 2147483647 / 10) return false;
            exponent = exponent * 10 + (int)(*p - '0');
            ++p; ++read;
        }
        if (exp_negative) exponent = -exponent;
        if (read == 0) return false;
    }
    *result = (negative ? -1 : 1) * (exponent ? std::ldexp(mantissa * std::pow(5.0, exponent), exponent) : mantissa);
    return true;
}

namespace {

// One whitespace-delimited number; missing / malformed -> `fallback` (tinyobj parseReal, :1030-1038).
float next_real(const char*& t, double fallback = 0.0) {
    t += std::strspn(t, " \t");
    const char* end = t + std::strcspn(t, " \t\r");
    double val = fallback;
    parse_obj_double(t, end, &val);
    t = end;
    return static_cast<float>(val);
}

std::string next_word(const char*& t) {
    t += std::strspn(t, " \t");
    const size_t n = std::strcspn(t, " \t\r");
    std::string s(t, t + n);
    t += n;
    return s;
}

// 1-based -> 0-based; negative = relative to `n`; returns false for an unusable index.
bool fix_index(int idx, int n, int* out, bool allow_zero) {
    if (idx > 0) { *out = idx - 1; return true; }
    if (idx == 0) { *out = -1; return allow_zero; }
    *out = n + idx;
    return *out >= 0;
}

// i, i/j, i//k, i/j/k
bool parse_triple(const char*& t, int nv, int nvn, int nvt, ObjIndex* out) {
    ObjIndex vi;
    if (!fix_index(std::atoi(t), nv, &vi.v, false)) return false;
    t += std::strcspn(t, "/ \t\r");
    if (*t != '/') { *out = vi; return true; }
    ++t;
    if (*t == '/') {
        ++t;
        if (!fix_index(std::atoi(t), nvn, &vi.vn, true)) return false;
        t += std::strcspn(t, "/ \t\r");
        *out = vi;
        return true;
    }
    if (!fix_index(std::atoi(t), nvt, &vi.vt, true)) return false;
    t += std::strcspn(t, "/ \t\r");
    if (*t != '/') { *out = vi; return true; }
    ++t;
    if (!fix_index(std::atoi(t), nvn, &vi.vn, true)) return false;
    t += std::strcspn(t, "/ \t\r");
    *out = vi;
    return true;
}

// Reads a line ending in \n, \r\n or \r.
bool get_line(std::istream& in, std::string& line) {
    line.clear();
    if (in.peek() == std::char_traits<char>::eof()) return false;
    std::streambuf* sb = in.rdbuf();
    for (;;) {
        const int c = sb->sbumpc();
        if (c == '\n') return true;
        if (c == '\r') { if (sb->sgetc() == '\n') sb->sbumpc(); return true; }
        if (c == std::char_traits<char>::eof()) { in.setstate(std::ios::eofbit); return true; }
        line.push_back((char)c);
    }
}

void read_mtl(std::istream& in, std::vector<ObjMaterial>& materials, std::map<std::string, int>& by_name) {
    ObjMaterial cur;
    std::string line;
    auto flush = [&]() {
        by_name.insert(std::make_pair(cur.name, (int)materials.size()));   // insert: the first of a name wins
        materials.push_back(cur);
    };
    while (get_line(in, line)) {
        while (!line.empty() && (line.back() == '\n' || line.back() == '\r')) line.pop_back();
        const char* t = line.c_str();
        t += std::strspn(t, " \t");
        if (*t == '\0' || *t == '#') continue;
        if (std::strncmp(t, "newmtl", 6) == 0 && is_space(t[6])) {
            if (!cur.name.empty()) flush();
            cur = ObjMaterial();
            t += 7;
            cur.name = next_word(t);
            continue;
        }
        if (std::strncmp(t, "map_Kd", 6) == 0 && is_space(t[6])) {
            t += 7;
            // options (-o u v w, -s u v w, -bm x, ...) then the file name = the REST of the line
            std::string name;
            while (!is_eol(*t)) {
                t += std::strspn(t, " \t");
                if (*t == '-') {
                    const std::string opt = next_word(t);
                    int nargs = 1;
                    if (opt == "-o" || opt == "-s" || opt == "-t") nargs = 3;
                    else if (opt == "-mm") nargs = 2;
                    for (int k = 0; k < nargs; ++k) {
                        const char* save = t;
                        t += std::strspn(t, " \t");
                        if (opt != "-o" && opt != "-s" && opt != "-t" && opt != "-mm") { next_word(t); break; }
                        double dummy;
                        const char* end = t + std::strcspn(t, " \t\r");
                        if (!parse_obj_double(t, end, &dummy)) { t = save; break; }   // optional v / w components
                        t = end;
                    }
                } else {
                    name = std::string(t);
                    t += name.size();
                }
            }
            if (!name.empty()) cur.diffuse_texname = name;
            continue;
        }
    }
    flush();   // tinyobj pushes the last material unconditionally
}

struct PendingFace { std::vector<ObjIndex> idx; };

}  // namespace

bool read_obj(const std::string& path, ObjMesh& out) {
    out = ObjMesh();
    std::ifstream in(path.c_str());
    if (!in) { out.error = "Cannot open file [" + path + "]\n"; return false; }
    std::string base;
    {
        const size_t pos = path.find_last_of("/\\");
        if (pos != std::string::npos) base = path.substr(0, pos) + "/";
    }
    std::map<std::string, int> material_by_name;
    std::set<std::string> loaded_mtl;
    std::vector<PendingFace> group;
    int material = -1;

    // exportGroupsToShape: triangulate and emit the pending face group with the material in force.
    auto flush = [&]() {
        std::vector<ObjIndex> tris;
        for (const PendingFace& f : group) {
            const size_t n = f.idx.size();
            if (n < 3) { out.warning += "Degenerated face found\n."; continue; }
            if (n == 3) {
                for (int k = 0; k < 3; ++k) out.indices.push_back(f.idx[k]);
                out.material_ids.push_back(material);
            } else if (n == 4) {
                bool ok = true;
                for (int k = 0; k < 4; ++k) ok = ok && (3 * (size_t)f.idx[k].v + 2 < out.vertices.size());
                if (!ok) { out.warning += "Face with invalid vertex index found.\n"; continue; }
                const float* a = &out.vertices[3 * (size_t)f.idx[0].v];
                const float* b = &out.vertices[3 * (size_t)f.idx[1].v];
                const float* c = &out.vertices[3 * (size_t)f.idx[2].v];
                const float* d = &out.vertices[3 * (size_t)f.idx[3].v];
                const float e02x = c[0] - a[0], e02y = c[1] - a[1], e02z = c[2] - a[2];
                const float e13x = d[0] - b[0], e13y = d[1] - b[1], e13z = d[2] - b[2];
                const float sqr02 = e02x * e02x + e02y * e02y + e02z * e02z;
                const float sqr13 = e13x * e13x + e13y * e13y + e13z * e13z;
                static const int split02[6] = {0, 1, 2, 0, 2, 3}, split13[6] = {0, 1, 3, 1, 2, 3};
                const int* order = (sqr02 < sqr13) ? split02 : split13;
                for (int k = 0; k < 6; ++k) out.indices.push_back(f.idx[order[k]]);
                out.material_ids.push_back(material);
                out.material_ids.push_back(material);
            } else {
                tris.clear();
                if (!triangulate_polygon(out.vertices, f.idx, tris)) {
                    out.warning += "polygon could not be triangulated\n";
                    continue;
                }
                for (const ObjIndex& i : tris) out.indices.push_back(i);
                for (size_t k = 0; k < tris.size() / 3; ++k) out.material_ids.push_back(material);
            }
        }
        group.clear();
    };

    std::string line;
    size_t line_no = 0;
    while (get_line(in, line)) {
        ++line_no;
        if (line.empty()) continue;
        const char* t = line.c_str();
        t += std::strspn(t, " \t");
        if (*t == '\0' || *t == '#') continue;
        if (t[0] == 'v' && is_space(t[1])) {
            t += 2;
            const float x = next_real(t), y = next_real(t), z = next_real(t);
            out.vertices.push_back(x); out.vertices.push_back(y); out.vertices.push_back(z);
            continue;
        }
        if (t[0] == 'v' && t[1] == 'n' && is_space(t[2])) {
            t += 3;
            const float x = next_real(t), y = next_real(t), z = next_real(t);
            out.normals.push_back(x); out.normals.push_back(y); out.normals.push_back(z);
            continue;
        }
        if (t[0] == 'v' && t[1] == 't' && is_space(t[2])) {
            t += 3;
            const float u = next_real(t), v = next_real(t);
            out.texcoords.push_back(u); out.texcoords.push_back(v);
            continue;
        }
        if (t[0] == 'f' && is_space(t[1])) {
            t += 2;
            t += std::strspn(t, " \t");
            PendingFace face;
            while (!is_eol(*t)) {
                ObjIndex vi;
                if (!parse_triple(t, (int)(out.vertices.size() / 3), (int)(out.normals.size() / 3),
                                  (int)(out.texcoords.size() / 2), &vi)) {
                    std::ostringstream ss;
                    ss << "Failed to parse `f' line (e.g. a zero value for vertex index or invalid relative vertex index). Line "
                       << line_no << ").\n";
                    out.error += ss.str();
                    return false;
                }
                face.idx.push_back(vi);
                t += std::strspn(t, " \t\r");
            }
            group.push_back(std::move(face));
            continue;
        }
        if (std::strncmp(t, "usemtl", 6) == 0) {
            t += 6;
            const std::string name = next_word(t);
            int id = -1;
            auto it = material_by_name.find(name);
            if (it != material_by_name.end()) id = it->second;
            else out.warning += "material [ '" + name + "' ] not found in .mtl\n";
            if (id != material) { flush(); material = id; }
            continue;
        }
        if (std::strncmp(t, "mtllib", 6) == 0 && is_space(t[6])) {
            t += 7;
            // file names separated by ' ' (backslash escapes a space)
            std::vector<std::string> names;
            std::string cur;
            bool esc = false;
            for (const char* p = t; *p; ++p) {
                if (esc) { cur.push_back(*p); esc = false; }
                else if (*p == '\\') esc = true;
                else if (*p == ' ') { if (!cur.empty()) names.push_back(cur); cur.clear(); }
                else cur.push_back(*p);
            }
            if (!cur.empty()) names.push_back(cur);
            bool found = false;
            for (const std::string& n : names) {
                if (loaded_mtl.count(n)) { found = true; continue; }
                std::ifstream mtl((base + n).c_str());
                if (mtl) {
                    read_mtl(mtl, out.materials, material_by_name);
                    loaded_mtl.insert(n);
                    found = true;
                    break;
                }
                out.warning += "Material file [ " + n + " ] not found in a path : " + base + "\n";
            }
            if (!found) out.warning += "Failed to load material file(s). Use default material.\n";
            continue;
        }
        if ((t[0] == 'g' || t[0] == 'o') && is_space(t[1])) { flush(); continue; }
        // everything else (s, l, p, t, vw, ...) does not change the triangle list
    }
    flush();
    // The reference indexes attrib.vertices / normals / texcoords unchecked (Object.cpp:87-102);
    // an out-of-range index is a crash there and an error here.
    const int nv = (int)(out.vertices.size() / 3), nn = (int)(out.normals.size() / 3), nt = (int)(out.texcoords.size() / 2);
    for (const ObjIndex& i : out.indices) {
        if (i.v < 0 || i.v >= nv || i.vn >= nn || i.vt >= nt) {
            out.error += "face references a vertex / normal / texcoord that does not exist\n";
            out.indices.clear(); out.material_ids.clear();
            return false;
        }
    }
    return true;
}

}  // namespace srt

// srt_polygon.cpp — triangulation of OBJ faces with more than four vertices, producing the triangles (set AND order)
// the reference's loader produces.
//
// The reference parses OBJ files with tinyobjloader built with TINYOBJLOADER_USE_MAPBOX_EARCUT
// (simple_raytracer.cpp:15-16).  For such a face that loader (tiny_obj_loader.h:1618-1724)
//   1. takes the face normal by Newell's method in float, negated and normalised,
//   2. builds a frame (u, v, w = n) from it and projects every vertex to (dot(p,u), dot(p,v)) in float,
//   3. hands the one ring to mapbox::earcut (library/tinyobjloader/mapbox/earcut.hpp, vendored dependency) and emits
//      the returned index triples in order.
// This file restates steps 1-3.  The ear clipper below is our own implementation of the published earcut algorithm
// for ONE ring without holes (OBJ faces have none): ring → clockwise list, ear slicing with the "skip the next
// vertex" rule, then the three fallbacks in the library's order (drop collinear / duplicate points, cure local
// self-intersections, split along a valid diagonal).  Predicates are evaluated in double on the float projections,
// with the library's expressions, so every comparison lands the same way.  Rings of 80 or more vertices make the
// library index its nodes along a z-order curve to find candidate points faster; that changes which points are
// LOOKED AT, not the predicate, and the plain scan used here gives the same triangles (tests/test_host_scene.py checks
// faces of 5-120 vertices — convex, star-shaped, combs, repeated / collinear points, non-planar, self-crossing —
// against the reference's loader, bit for bit; obj/house/house.obj is the one asset of the reference with such faces).
#include <cmath>
#include <cstdint>

#include "srt_scene.hpp"

namespace srt {
namespace {

struct P3 { float x, y, z; };
inline P3 cross3(const P3& a, const P3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float dot3(const P3& a, const P3& b) { return (a.x * b.x + a.y * b.y + a.z * b.z); }
inline float length3(const P3& e) { return std::sqrt(e.x * e.x + e.y * e.y + e.z * e.z); }

// Steps 1-2 (tiny_obj_loader.h:1624-1700).  false: the normal has zero length and the loader drops the face.
bool project_face(const std::vector<float>& v, const std::vector<ObjIndex>& face, std::vector<double>& xs, std::vector<double>& ys) {
    const size_t n = face.size();
    P3 nrm = {0.f, 0.f, 0.f};
    for (size_t k = 0; k < n; ++k) {
        const float* p = &v[3 * (size_t)face[k].v];
        const float* q = &v[3 * (size_t)face[(k + 1) % n].v];
        const P3 a = {p[0] - q[0], p[1] - q[1], p[2] - q[2]};
        const P3 b = {p[0] + q[0], p[1] + q[1], p[2] + q[2]};
        nrm.x += (a.y * b.z);
        nrm.y += (a.z * b.x);
        nrm.z += (a.x * b.y);
    }
    const float len = length3(nrm);
    if (len <= 0) return false;
    const float inv = -1.0f / len;
    const P3 w = {nrm.x * inv, nrm.y * inv, nrm.z * inv};
    const P3 helper = (std::fabs(w.x) > 0.9999999f) ? P3{0.f, 1.f, 0.f} : P3{1.f, 0.f, 0.f};
    const P3 c = cross3(w, helper);
    const float inv_c = 1.0f / length3(c);
    const P3 axis_v = {c.x * inv_c, c.y * inv_c, c.z * inv_c};
    const P3 axis_u = cross3(w, axis_v);
    xs.resize(n); ys.resize(n);
    for (size_t k = 0; k < n; ++k) {
        const float* p = &v[3 * (size_t)face[k].v];
        const P3 pt = {p[0], p[1], p[2]};
        xs[k] = (double)dot3(pt, axis_u);
        ys[k] = (double)dot3(pt, axis_v);
    }
    return true;
}

// Step 3.  Nodes live in parallel arrays and are named by their slot; a node taken out of the ring keeps its own
// links (the library relies on that when it steps on from a removed node, earcut.hpp:808-815).
class EarClipper {
public:
    std::vector<uint32_t> out;     // positions within the face, three per triangle

    void run(const std::vector<double>& xs, const std::vector<double>& ys) {
        const size_t n = xs.size();
        const size_t cap = n * 3 / 2 + 8;
        x_.reserve(cap); y_.reserve(cap); id_.reserve(cap); prev_.reserve(cap); next_.reserve(cap);
        // winding of the ring as given (earcut.hpp:191-207): the list is linked clockwise
        double sum = 0;
        for (size_t i = 0, j = n ? n - 1 : 0; i < n; j = i++) sum += (xs[j] - xs[i]) * (ys[i] + ys[j]);
        int last = -1;
        if (sum > 0) { for (size_t i = 0; i < n; ++i) last = append(i, xs[i], ys[i], last); }
        else         { for (size_t i = n; i-- > 0;)   last = append(i, xs[i], ys[i], last); }
        if (last >= 0 && same_point(last, next_[last])) { unlink(last); last = next_[last]; }
        if (last < 0 || prev_[last] == next_[last]) return;
        slice(last, 0);
    }

private:
    std::vector<double> x_, y_;
    std::vector<uint32_t> id_;
    std::vector<int> prev_, next_;

    int make(uint32_t id, double x, double y) {
        x_.push_back(x); y_.push_back(y); id_.push_back(id); prev_.push_back(-1); next_.push_back(-1);
        return (int)x_.size() - 1;
    }
    int append(size_t id, double x, double y, int last) {
        const int p = make((uint32_t)id, x, y);
        if (last < 0) { prev_[p] = p; next_[p] = p; }
        else { next_[p] = next_[last]; prev_[p] = last; prev_[next_[last]] = p; next_[last] = p; }
        return p;
    }
    void unlink(int p) { prev_[next_[p]] = prev_[p]; next_[prev_[p]] = next_[p]; }

    // signed area of (p, q, r), earcut.hpp:676-679
    double turn(int p, int q, int r) const { return (y_[q] - y_[p]) * (x_[r] - x_[q]) - (x_[q] - x_[p]) * (y_[r] - y_[q]); }
    bool same_point(int a, int b) const { return x_[a] == x_[b] && y_[a] == y_[b]; }
    static int sgn(double v) { return (0.0 < v) - (v < 0.0); }
    bool inside_triangle(int a, int b, int c, int p) const {     // earcut.hpp:660-664
        const double ax = x_[a], ay = y_[a], bx = x_[b], by = y_[b], cx = x_[c], cy = y_[c], px = x_[p], py = y_[p];
        return (cx - px) * (ay - py) - (ax - px) * (cy - py) >= 0 &&
               (ax - px) * (by - py) - (bx - px) * (ay - py) >= 0 &&
               (bx - px) * (cy - py) - (cx - px) * (by - py) >= 0;
    }

    void emit(int a, int b, int c) { out.push_back(id_[a]); out.push_back(id_[b]); out.push_back(id_[c]); }

    // earcut.hpp:309-327: convex corner with no reflex vertex of the rest of the ring inside it
    bool is_ear(int ear) const {
        const int a = prev_[ear], c = next_[ear];
        if (turn(a, ear, c) >= 0) return false;
        for (int p = next_[c]; p != a; p = next_[p])
            if (inside_triangle(a, ear, c, p) && turn(prev_[p], p, next_[p]) >= 0) return false;
        return true;
    }

    // earcut.hpp:230-253: take out duplicate and collinear points, walking back one step after each removal
    int drop_flat_points(int start, int end = -1) {
        if (end < 0) end = start;
        int p = start;
        bool again;
        do {
            again = false;
            if (same_point(p, next_[p]) || turn(prev_[p], p, next_[p]) == 0) {
                unlink(p);
                p = end = prev_[p];
                if (p == next_[p]) break;
                again = true;
            } else {
                p = next_[p];
            }
        } while (again || p != end);
        return end;
    }

    void slice(int ear, int pass);
    int cure_local_crossings(int start);
    void split_and_slice(int start);
    bool segments_cross(int p1, int q1, int p2, int q2) const;
    bool between(int p, int q, int r) const;
    bool crosses_ring(int a, int b) const;
    bool locally_inside(int a, int b) const;
    bool middle_inside(int a, int b) const;
    bool usable_diagonal(int a, int b) const;
    int split_ring(int a, int b);
};

// The slicing loop (earcut.hpp:256-306).  After a cut the walk resumes two vertices on, which avoids slivers; a full
// lap without an ear moves to the next fallback: pass 0 → drop flat points, pass 1 → also cure local crossings,
// pass 2 → split the ring along a diagonal and start over on both halves.
void EarClipper::slice(int ear, int pass) {
    if (ear < 0) return;
    int stop = ear;
    while (prev_[ear] != next_[ear]) {
        const int before = prev_[ear], after = next_[ear];
        if (is_ear(ear)) {
            emit(before, ear, after);
            unlink(ear);
            ear = next_[after];
            stop = next_[after];
            continue;
        }
        ear = after;
        if (ear == stop) {
            if (pass == 0) slice(drop_flat_points(ear), 1);
            else if (pass == 1) slice(cure_local_crossings(drop_flat_points(ear)), 2);
            else if (pass == 2) split_and_slice(ear);
            break;
        }
    }
}

// earcut.hpp:370-394: where edge (a, p) crosses edge (p.next, b), cut the corner a-p-b off and drop p and p.next
int EarClipper::cure_local_crossings(int start) {
    int p = start;
    do {
        const int a = prev_[p], b = next_[next_[p]];
        if (!same_point(a, b) && segments_cross(a, p, next_[p], b) && locally_inside(a, b) && locally_inside(b, a)) {
            emit(a, p, b);
            unlink(p);
            unlink(next_[p]);
            p = start = b;
        }
        p = next_[p];
    } while (p != start);
    return drop_flat_points(p);
}

// earcut.hpp:397-422: first valid diagonal in ring order splits the ring; both parts are sliced from pass 0
void EarClipper::split_and_slice(int start) {
    int a = start;
    do {
        for (int b = next_[next_[a]]; b != prev_[a]; b = next_[b]) {
            if (id_[a] != id_[b] && usable_diagonal(a, b)) {
                int c = split_ring(a, b);
                a = drop_flat_points(a, next_[a]);
                c = drop_flat_points(c, next_[c]);
                slice(a, 0);
                slice(c, 0);
                return;
            }
        }
        a = next_[a];
    } while (a != start);
}

bool EarClipper::between(int p, int q, int r) const {     // q within the box of segment pr (collinear case), :709-715
    return x_[q] <= std::fmax(x_[p], x_[r]) && x_[q] >= std::fmin(x_[p], x_[r]) &&
           y_[q] <= std::fmax(y_[p], y_[r]) && y_[q] >= std::fmin(y_[p], y_[r]);
}

bool EarClipper::segments_cross(int p1, int q1, int p2, int q2) const {     // earcut.hpp:688-705
    const int o1 = sgn(turn(p1, q1, p2)), o2 = sgn(turn(p1, q1, q2));
    const int o3 = sgn(turn(p2, q2, p1)), o4 = sgn(turn(p2, q2, q1));
    if (o1 != o2 && o3 != o4) return true;
    if (o1 == 0 && between(p1, p2, q1)) return true;
    if (o2 == 0 && between(p1, q2, q1)) return true;
    if (o3 == 0 && between(p2, p1, q2)) return true;
    if (o4 == 0 && between(p2, q1, q2)) return true;
    return false;
}

bool EarClipper::crosses_ring(int a, int b) const {     // earcut.hpp:723-733
    int p = a;
    do {
        const int q = next_[p];
        if (id_[p] != id_[a] && id_[q] != id_[a] && id_[p] != id_[b] && id_[q] != id_[b] && segments_cross(p, q, a, b)) return true;
        p = q;
    } while (p != a);
    return false;
}

bool EarClipper::locally_inside(int a, int b) const {     // earcut.hpp:736-741
    return turn(prev_[a], a, next_[a]) < 0
        ? turn(a, b, next_[a]) >= 0 && turn(a, prev_[a], b) >= 0
        : turn(a, b, prev_[a]) < 0 || turn(a, next_[a], b) < 0;
}

bool EarClipper::middle_inside(int a, int b) const {     // earcut.hpp:744-759: crossing-number test of the midpoint
    bool inside = false;
    const double px = (x_[a] + x_[b]) / 2, py = (y_[a] + y_[b]) / 2;
    int p = a;
    do {
        const int q = next_[p];
        if (((y_[p] > py) != (y_[q] > py)) && y_[q] != y_[p] &&
            (px < (x_[q] - x_[p]) * (py - y_[p]) / (y_[q] - y_[p]) + x_[p]))
            inside = !inside;
        p = q;
    } while (p != a);
    return inside;
}

bool EarClipper::usable_diagonal(int a, int b) const {     // earcut.hpp:667-673
    return id_[next_[a]] != id_[b] && id_[prev_[a]] != id_[b] && !crosses_ring(a, b) &&
           ((locally_inside(a, b) && locally_inside(b, a) && middle_inside(a, b) &&
             (turn(prev_[a], a, prev_[b]) != 0.0 || turn(a, prev_[b], b) != 0.0)) ||
            (same_point(a, b) && turn(prev_[a], a, next_[a]) > 0 && turn(prev_[b], b, next_[b]) > 0));
}

// earcut.hpp:765-785: ring … a b … keeps a and b; copies a', b' close the other ring … b' a' …
int EarClipper::split_ring(int a, int b) {
    const int a2 = make(id_[a], x_[a], y_[a]);
    const int b2 = make(id_[b], x_[b], y_[b]);
    const int an = next_[a], bp = prev_[b];
    next_[a] = b;   prev_[b] = a;
    next_[a2] = an; prev_[an] = a2;
    next_[b2] = a2; prev_[a2] = b2;
    next_[bp] = b2; prev_[b2] = bp;
    return b2;
}

}  // namespace

bool triangulate_polygon(const std::vector<float>& v, const std::vector<ObjIndex>& face, std::vector<ObjIndex>& out) {
    if (face.size() < 3) return false;
    for (const ObjIndex& i : face)
        if (i.v < 0 || 3 * (size_t)i.v + 2 >= v.size()) return false;     // the loader asserts here
    std::vector<double> xs, ys;
    if (!project_face(v, face, xs, ys)) return true;     // zero-length normal: the face is dropped without a warning (:1658)
    EarClipper clip;
    clip.run(xs, ys);
    for (uint32_t k : clip.out) out.push_back(face[k]);
    return true;
}

}  // namespace srt

// srt_polygon.cpp — triangulation of OBJ faces with more than four vertices.
//
// STATUS: fan triangulation from the first vertex (exact for convex planar polygons).  The
// reference's loader uses mapbox earcut for these faces (tiny_obj_loader.h:1618 with
// TINYOBJLOADER_USE_MAPBOX_EARCUT, simple_raytracer.cpp:16), whose triangle set / order can differ
// for non-convex faces.  Of the reference's assets only obj/house/house.obj has such faces
// (7 of 2,389).  DESIGN.md lists this as an open host-side item; the render core is unaffected
// (it consumes whatever triangles the host produced).
#include "srt_scene.hpp"

namespace srt {

bool triangulate_polygon(const std::vector<float>& v, const std::vector<ObjIndex>& face,
                         std::vector<ObjIndex>& out) {
    (void)v;
    if (face.size() < 3) return false;
    for (size_t k = 1; k + 1 < face.size(); ++k) {
        out.push_back(face[0]);
        out.push_back(face[k]);
        out.push_back(face[k + 1]);
    }
    return true;
}

}  // namespace srt

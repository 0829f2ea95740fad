/*
 * ref_shapes_glue.cpp — TEST INFRASTRUCTURE.  Wraps the REFERENCE'S OWN C++ shape
 * classes (src/shapes/sphere.hpp:41-60, plane.hpp:44-57, wall.hpp:46-68,
 * triangle.hpp:100-130), compiled from /root/reference in place, behind one C
 * function so the oracle's primitive tests can be cross-checked against them.
 * Built by oracle/build_ref.py into oracle/_ref/libref_shapes.so.
 */
#include <memory>
#include <iostream>
#include <glm/glm.hpp>
#include "shapes/sphere.hpp"
#include "shapes/wall.hpp"
#include "shapes/triangle.hpp"

// type: 0 sphere (a=center, s0=radius) | 1 plane (a=normal, b=point)
//       2 wall (a=start, s0=width, s1=height, b=normal) | 3 triangle (a,b,c; invert flag in s0)
// Returns the reference's IntersectType (0 NONE, 1 INNER, 2 OUTER); hit point for non-NONE.
extern "C" int refshape_intersect(int type, const float* a, const float* b, const float* c, float s0, float s1,
                                  const float* start, const float* dir, float* hit, float* normal_d) {
    glm::vec3 A(a[0], a[1], a[2]), B(b[0], b[1], b[2]), Cc(c[0], c[1], c[2]);
    Ray ray(glm::vec3(start[0], start[1], start[2]), glm::vec3(dir[0], dir[1], dir[2]));
    Intersection r;
    if (type == 0) { Sphere s(A, s0); r = s.get_intersection(ray); }
    else if (type == 1) { Plane p(A, B); r = p.get_intersection(ray);
        if (normal_d) { normal_d[0] = p.m_normal.x; normal_d[1] = p.m_normal.y; normal_d[2] = p.m_normal.z; normal_d[3] = p.d; } }
    else if (type == 2) { Wall w(A, s0, s1, B); r = w.get_intersection(ray);
        if (normal_d) { normal_d[0] = w.m_normal.x; normal_d[1] = w.m_normal.y; normal_d[2] = w.m_normal.z; normal_d[3] = w.d; } }
    else { Triangle t(A, B, Cc); if (s0 != 0.f) t.invert_normal(); r = t.get_intersection(ray);
        if (normal_d) { normal_d[0] = t.m_normal.x; normal_d[1] = t.m_normal.y; normal_d[2] = t.m_normal.z; normal_d[3] = t.d; } }
    hit[0] = r.hit_point.x; hit[1] = r.hit_point.y; hit[2] = r.hit_point.z;
    return (int)r.intersect_type;
}

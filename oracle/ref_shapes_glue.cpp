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

/*
 * The reference's own BoundingBox (src/BoundingBox.hpp:34-95), Wall::end (shapes/wall.hpp:16-31) and Triangle::center
 * (shapes/triangle.hpp:27-29) — the arithmetic buildBVH / split (src/main.cpp:1111-1193) is made of — for pinning the host
 * restatement and the device builder K8 (rt_build.cuh) against the reference's code itself.  mesh.hpp drags in OpenGL;
 * BoundingBox only needs Mesh::mesh2triangles to exist, so the header's include guard is taken and a stand-in declared.
 */
#define MESH_H
class Mesh { public: std::vector<Triangle> mesh2triangles() { return {}; } };
#include "BoundingBox.hpp"

// Shapes i = 0..n-1, types 0 sphere / 2 wall / 3 triangle (1 = infinite plane: a = normal, b = point), parameters as for
// refshape_intersect (12 floats per shape: a, b, c, s0, s1, unused).  Grows ONE reference BoundingBox by all of them in
// order through growToInclude(std::unique_ptr<Shape>&), the overload buildBVH uses.  Per shape also returns what
// serializeShape would put into the flat structure (normal, d: 4 floats) and split's centre (3 floats; planes: untouched).
extern "C" void refbbox_grow(int n, const int* type, const float* par, float box[6], float* normal_d, float* center) {
    BoundingBox bb;
    for (int i = 0; i < n; ++i) {
        const float* p = par + 12 * (size_t)i;
        glm::vec3 A(p[0], p[1], p[2]), B(p[3], p[4], p[5]), Cc(p[6], p[7], p[8]);
        std::unique_ptr<Shape> s;
        if (type[i] == 0) s = std::make_unique<Sphere>(A, p[9]);
        else if (type[i] == 1) s = std::make_unique<Plane>(A, B);
        else if (type[i] == 2) s = std::make_unique<Wall>(A, p[9], p[10], B);
        else s = std::make_unique<Triangle>(A, B, Cc);
        bb.growToInclude(s);
        float* nd = normal_d + 4 * (size_t)i; float* c = center + 3 * (size_t)i;
        if (auto sp = dynamic_cast<Sphere*>(s.get())) { c[0] = sp->m_center.x; c[1] = sp->m_center.y; c[2] = sp->m_center.z; }
        else if (auto w = dynamic_cast<Wall*>(s.get())) {
            const glm::vec3 m = (w->start + w->end()) * 0.5f;                      // main.cpp:1131-1134
            c[0] = m.x; c[1] = m.y; c[2] = m.z;
            nd[0] = w->m_normal.x; nd[1] = w->m_normal.y; nd[2] = w->m_normal.z; nd[3] = w->d;
        } else if (auto t = dynamic_cast<Triangle*>(s.get())) {
            const glm::vec3 m = t->center();
            c[0] = m.x; c[1] = m.y; c[2] = m.z;
            nd[0] = t->m_normal.x; nd[1] = t->m_normal.y; nd[2] = t->m_normal.z; nd[3] = t->d;
        } else if (auto pl = dynamic_cast<Plane*>(s.get())) {
            nd[0] = pl->m_normal.x; nd[1] = pl->m_normal.y; nd[2] = pl->m_normal.z; nd[3] = pl->d;
        }
    }
    box[0] = bb.Min.x; box[1] = bb.Min.y; box[2] = bb.Min.z; box[3] = bb.Max.x; box[4] = bb.Max.y; box[5] = bb.Max.z;
}

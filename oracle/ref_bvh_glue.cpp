/*
 * ref_bvh_glue.cpp — TEST INFRASTRUCTURE.  The REFERENCE'S OWN scene-side code of the path, compiled in place:
 *   class Node                       src/main.cpp:74-90
 *   struct Wheel / wheels[4]         src/main.cpp:103-108
 *   serializeBVH                     src/main.cpp:955-979
 *   updateScene (GL calls stubbed)   src/main.cpp:981-992
 *   serializeShape                   src/main.cpp:994-1066
 *   updateBVH                        src/main.cpp:1068-1077
 *   bounceSphere                     src/main.cpp:1079-1082
 *   updateWheelAnimations            src/main.cpp:1084-1109
 *   split / buildBVH                 src/main.cpp:1111-1193
 * oracle/build_ref.py cuts these definitions out of main.cpp MECHANICALLY (by signature + brace matching, no edits)
 * into oracle/_ref/ref_main_*.inc; this file supplies what main.cpp's other 1000 lines (window, GUI, GL) would have
 * declared around them, and a C interface.  Built into oracle/_ref/libref_bvh.so.  Pins the repo's host builder
 * (host/rt_host.cpp), the device builder K8 (csrc/rt_build.cuh), the host animation and K6 / K7 (csrc/rt_anim.cuh)
 * to upstream bytes: tests/test_ref_builder.py, tests/test_gpu_build.py, tests/test_gpu_parity.py.
 */
#include <cmath>
#include <cstring>
#include <iostream>
#include <memory>
#include <vector>
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>
#include "shapes/shape.hpp"
#include "shapes/sphere.hpp"
#include "shapes/plane.hpp"
#include "shapes/wall.hpp"
#include "shapes/triangle.hpp"
#include "flatStructures.hpp"
// mesh.hpp drags in OpenGL; BoundingBox only needs Mesh::mesh2triangles to exist (as in ref_shapes_glue.cpp)
#define MESH_H
class Mesh { public: std::vector<Triangle> mesh2triangles() { return {}; } };
#include "BoundingBox.hpp"

// what main.cpp declares around the extracted definitions (src/main.cpp:60-131), GL names as inert stand-ins
typedef unsigned GLuint;
#define GL_SHADER_STORAGE_BUFFER 0
#define glBindBuffer(...) ((void)0)
#define glBufferSubData(...) ((void)0)
FlatShape serializeShape(const std::unique_ptr<Shape>& shape);
#include "_ref/ref_main_node.inc"
void updateBVH();
void split(std::unique_ptr<Node>& parentNode, int depth = 15);
int buildBVH(int maxDepth = 15);
struct Scene {
    std::vector<std::unique_ptr<Shape>> shapes;
    std::vector<std::unique_ptr<Node>> bvhNodes;
} scene;
#include "_ref/ref_main_wheel.inc"
std::vector<int> animatedIndices;
float deltaTime = 0.0f;
#include "_ref/ref_main_funcs.inc"

// flatNodes / bvhIndices are the globals flatStructures.hpp:107-108 itself defines
namespace { FlatScene flatScene; }

static_assert(sizeof(FlatShape) == 192 && sizeof(FlatNode) == 48, "std430 sizes");

extern "C" {

// Scene from flat shapes: objects through the reference constructors, then every member set to the flat value (the
// constructors re-normalise normals; the builder must see exactly the bytes the other builders see).
int refbvh_load(const FlatShape* shapes, int n, const unsigned char* animated) {
    scene.shapes.clear(); scene.bvhNodes.clear(); animatedIndices.clear();
    for (auto& w : wheels) { w.shapeIndices.clear(); w.rotationAngle = 0; }
    for (int i = 0; i < n; ++i) {
        const FlatShape& f = shapes[i];
        std::unique_ptr<Shape> s;
        if (f.type == 0) { s = std::make_unique<Sphere>(f.sphereCenter, f.sphereRadius); }
        else if (f.type == 1) { auto p = std::make_unique<Plane>(f.planeNormal, glm::vec3(0)); p->m_normal = f.planeNormal; p->d = f.planeD; s = std::move(p); }
        else if (f.type == 2) { auto w = std::make_unique<Wall>(f.wallStart, f.wallWidth, f.wallHeight, f.planeNormal); w->m_normal = f.planeNormal; w->d = f.planeD; s = std::move(w); }
        else if (f.type == 3) { auto t = std::make_unique<Triangle>(f.triP1, f.triP2, f.triP3); t->m_normal = f.planeNormal; t->d = f.planeD; s = std::move(t); }
        else return -1;
        s->material.color = f.material.color; s->material.fresnelStrength = f.material.fresnelStrength;
        s->material.ambientStrength = f.material.ambientStrength; s->material.diffuseStrength = f.material.diffuseStrength;
        s->material.specularStrength = f.material.specularStrength; s->material.shininess = f.material.shininess;
        if (animated && animated[i]) { s->animated = true; animatedIndices.push_back(i); }
        scene.shapes.push_back(std::move(s));
    }
    flatScene.shapes.assign(shapes, shapes + n);
    return 0;
}

// buildBVH + serializeBVH (main.cpp:1175-1193, 955-979)
void refbvh_build(int maxDepth, int* nNodes, int* nIndices) {
    buildBVH(maxDepth);
    serializeBVH(flatNodes, bvhIndices);
    *nNodes = (int)flatNodes.size(); *nIndices = (int)bvhIndices.size();
}
void refbvh_get(FlatNode* nodes, int* indices) {
    if (nodes && !flatNodes.empty()) std::memcpy(nodes, flatNodes.data(), flatNodes.size() * sizeof(FlatNode));
    if (indices && !bvhIndices.empty()) std::memcpy(indices, bvhIndices.data(), bvhIndices.size() * sizeof(int));
}
void refbvh_get_shapes(FlatShape* out) {
    if (!flatScene.shapes.empty()) std::memcpy(out, flatScene.shapes.data(), flatScene.shapes.size() * sizeof(FlatShape));
}

// wheel k of the car scene as generateScene2 registers it (main.cpp:766-781): axis (0,0,1)
int refbvh_wheel(int k, const int* idx, int n, const float center[3]) {
    if (k < 0 || k >= 4) return -1;
    wheels[k].shapeIndices.assign(idx, idx + n);
    wheels[k].rotationAxis = glm::vec3(0, 0, 1);
    wheels[k].center = glm::vec3(center[0], center[1], center[2]);
    return 0;
}
// the frame step of the render loop (main.cpp:328-344 with the per-scene animation calls in front of it)
int refbvh_bounce(int shape, float elapsed, float amplitude, float frequency) {
    auto* s = dynamic_cast<Sphere*>(scene.shapes[(size_t)shape].get());
    if (!s) return -1;
    bounceSphere(s, elapsed, amplitude, frequency);
    return 0;
}
void refbvh_wheels_step(float elapsed, float dt) { deltaTime = dt; updateWheelAnimations(elapsed); }
void refbvh_frame(int rebuild, int maxDepth, int* nNodes, int* nIndices) {
    updateScene(flatScene, 0);
    if (rebuild) buildBVH(maxDepth); else updateBVH();
    serializeBVH(flatNodes, bvhIndices);
    *nNodes = (int)flatNodes.size(); *nIndices = (int)bvhIndices.size();
}

}  // extern "C"

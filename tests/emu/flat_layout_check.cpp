// flat_layout_check.cpp — TEST INFRASTRUCTURE.  The reference's own src/flatStructures.hpp next to include/rt_flat.h in one
// translation unit: every size and every field offset of the wire format must agree (the C ABI consumes the reference's
// std430 bytes unchanged).  Compiled only where /root/reference is present (tests/test_abi.py).
#include <cstddef>
#include "flatStructures.hpp"            // the reference's (include path: <reference>/src, <reference>/glm)
#include "rt_flat.h"

#define SAME(refT, refF, ourT, ourF) static_assert(offsetof(refT, refF) == offsetof(ourT, ourF), #refT "::" #refF)
static_assert(sizeof(FlatLight) == sizeof(rt_flat_light), "FlatLight");
static_assert(sizeof(FlatCamera) == sizeof(rt_flat_camera), "FlatCamera");
static_assert(sizeof(FlatMaterial) == sizeof(rt_flat_material), "FlatMaterial");
static_assert(sizeof(FlatShape) == sizeof(rt_flat_shape) && alignof(FlatShape) == 16, "FlatShape");
static_assert(sizeof(FlatNode) == sizeof(rt_flat_node), "FlatNode");
SAME(FlatLight, position, rt_flat_light, position); SAME(FlatLight, color, rt_flat_light, color);
SAME(FlatCamera, Position, rt_flat_camera, Position); SAME(FlatCamera, aspectRatio, rt_flat_camera, aspectRatio);
SAME(FlatCamera, Front, rt_flat_camera, Front); SAME(FlatCamera, Up, rt_flat_camera, Up);
SAME(FlatCamera, Right, rt_flat_camera, Right); SAME(FlatCamera, fov, rt_flat_camera, fov);
SAME(FlatMaterial, color, rt_flat_material, color); SAME(FlatMaterial, fresnelStrength, rt_flat_material, fresnelStrength);
SAME(FlatMaterial, ambientStrength, rt_flat_material, ambientStrength); SAME(FlatMaterial, diffuseStrength, rt_flat_material, diffuseStrength);
SAME(FlatMaterial, specularStrength, rt_flat_material, specularStrength); SAME(FlatMaterial, shininess, rt_flat_material, shininess);
SAME(FlatShape, type, rt_flat_shape, type); SAME(FlatShape, material, rt_flat_shape, material);
SAME(FlatShape, sphereCenter, rt_flat_shape, sphereCenter); SAME(FlatShape, sphereRadius, rt_flat_shape, sphereRadius);
SAME(FlatShape, planeNormal, rt_flat_shape, planeNormal); SAME(FlatShape, planeD, rt_flat_shape, planeD);
SAME(FlatShape, wallStart, rt_flat_shape, wallStart); SAME(FlatShape, wallWidth, rt_flat_shape, wallWidth);
SAME(FlatShape, wallHeight, rt_flat_shape, wallHeight);
SAME(FlatShape, triP1, rt_flat_shape, triP1); SAME(FlatShape, triP2, rt_flat_shape, triP2); SAME(FlatShape, triP3, rt_flat_shape, triP3);
SAME(FlatNode, boundsMin, rt_flat_node, boundsMin); SAME(FlatNode, boundsMax, rt_flat_node, boundsMax);
SAME(FlatNode, leftChild, rt_flat_node, leftChild); SAME(FlatNode, rightChild, rt_flat_node, rightChild);
SAME(FlatNode, startShapeIdx, rt_flat_node, startShapeIdx); SAME(FlatNode, numShapes, rt_flat_node, numShapes);
int main() { return 0; }

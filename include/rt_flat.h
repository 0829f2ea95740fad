/*
 * rt_flat.h — byte-exact mirrors of the reference's std430 "flat" structures.
 *
 * These are the wire format the reference uploads to its five SSBOs
 * (reference: src/flatStructures.hpp:7-106, shader-side twins in
 * src/shaders/gpu_shader.comp:4-86, binding contract README.md:77-98).
 * The C ABI in rt_api.h consumes them unchanged.  Plain C, no glm: a
 * reference-side caller passes `&flatScene.shapes[0]`, `flatNodes.data()`,
 * `bvhIndices.data()` straight through.
 *
 * Sizes/offsets were verified against the reference header compiled with
 * g++ 13 and are locked below with static asserts.
 */
#ifndef RT_FLAT_H
#define RT_FLAT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* src/flatStructures.hpp:72-78 (32 B) */
typedef struct rt_flat_light {
    float position[3];
    float padding1;
    float color[3];
    float padding2;
} rt_flat_light;

/* src/flatStructures.hpp:55-70 (80 B on the host; the shader reads bytes 0-67) */
typedef struct rt_flat_camera {
    float Position[3];
    float aspectRatio;
    float Front[3];
    float padding2;
    float Up[3];
    float padding3;
    float Right[3];
    float padding4;
    float fov;
    float padding5[3];
} rt_flat_camera;

/* src/flatStructures.hpp:7-20 (32 B, 16-aligned) */
typedef struct rt_flat_material {
    float   color[3];
    float   fresnelStrength;
    float   ambientStrength;
    float   diffuseStrength;
    float   specularStrength;
    int32_t shininess;
} rt_flat_material;

/* Shape type tags, gpu_shader.comp:40 / README.md:62 */
enum { RT_SHAPE_SPHERE = 0, RT_SHAPE_PLANE = 1, RT_SHAPE_WALL = 2, RT_SHAPE_TRIANGLE = 3 };

/* src/flatStructures.hpp:22-53 (192 B, 16-aligned) */
typedef struct rt_flat_shape {
    int32_t          type;            /*   0 */
    int32_t          _pad0[3];        /*   4 */
    float            padding[3];      /*  16 */
    float            _pad1;           /*  28 */
    rt_flat_material material;        /*  32 */
    float            sphereCenter[3]; /*  64 */
    float            sphereRadius;    /*  76 */
    float            planeNormal[3];  /*  80 */
    float            planeD;          /*  92 */
    float            wallStart[3];    /*  96 */
    float            wallWidth;       /* 108 */
    float            wallHeight;      /* 112 */
    float            _pad2[3];        /* 116 */
    float            padding1[3];     /* 128 */
    float            _pad3;           /* 140 */
    float            triP1[3];        /* 144 */
    float            padding2;        /* 156 */
    float            triP2[3];        /* 160 */
    float            padding3;        /* 172 */
    float            triP3[3];        /* 176 */
    float            padding4;        /* 188 */
} rt_flat_shape;

/* src/flatStructures.hpp:94-106 (48 B, 16-aligned).  Leaf <=> leftChild == -1
 * (gpu_shader.comp:399); numShapes/startShapeIdx are only meaningful on leaves
 * (src/main.cpp:967-976). */
typedef struct rt_flat_node {
    float   boundsMin[3];
    float   padding1;
    float   boundsMax[3];
    float   padding2;
    int32_t leftChild;
    int32_t rightChild;
    int32_t startShapeIdx;
    int32_t numShapes;
} rt_flat_node;

#ifdef __cplusplus
}
#define RT_SA(c) static_assert(c, #c)
#else
#define RT_SA(c) _Static_assert(c, #c)
#endif

RT_SA(sizeof(rt_flat_light) == 32);
RT_SA(offsetof(rt_flat_light, color) == 16);
RT_SA(sizeof(rt_flat_camera) == 80);
RT_SA(offsetof(rt_flat_camera, aspectRatio) == 12);
RT_SA(offsetof(rt_flat_camera, Front) == 16);
RT_SA(offsetof(rt_flat_camera, Up) == 32);
RT_SA(offsetof(rt_flat_camera, Right) == 48);
RT_SA(offsetof(rt_flat_camera, fov) == 64);
RT_SA(sizeof(rt_flat_material) == 32);
RT_SA(offsetof(rt_flat_material, fresnelStrength) == 12);
RT_SA(offsetof(rt_flat_material, shininess) == 28);
RT_SA(sizeof(rt_flat_shape) == 192);
RT_SA(offsetof(rt_flat_shape, material) == 32);
RT_SA(offsetof(rt_flat_shape, sphereCenter) == 64);
RT_SA(offsetof(rt_flat_shape, sphereRadius) == 76);
RT_SA(offsetof(rt_flat_shape, planeNormal) == 80);
RT_SA(offsetof(rt_flat_shape, planeD) == 92);
RT_SA(offsetof(rt_flat_shape, wallStart) == 96);
RT_SA(offsetof(rt_flat_shape, wallWidth) == 108);
RT_SA(offsetof(rt_flat_shape, wallHeight) == 112);
RT_SA(offsetof(rt_flat_shape, triP1) == 144);
RT_SA(offsetof(rt_flat_shape, triP2) == 160);
RT_SA(offsetof(rt_flat_shape, triP3) == 176);
RT_SA(sizeof(rt_flat_node) == 48);
RT_SA(offsetof(rt_flat_node, boundsMax) == 16);
RT_SA(offsetof(rt_flat_node, leftChild) == 32);
RT_SA(offsetof(rt_flat_node, startShapeIdx) == 40);
RT_SA(offsetof(rt_flat_node, numShapes) == 44);
#undef RT_SA

#endif /* RT_FLAT_H */

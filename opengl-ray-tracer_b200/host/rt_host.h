/*
 * rt_host.h — headless stand-in for the reference's CPU side (layers L3/L4 of
 * SURVEY.md §1): deterministic scene generators, the reference's midpoint-split
 * BVH builder / grow-only refit, flattening to the std430 wire format and the
 * scene animations.  It exists to FEED the hot path (librt_b200.so) and the
 * oracle identical bytes; it is not part of the hot path.
 *
 * Restates (never copies): generateScene1/2/3 src/main.cpp:583-804,1196-1229,
 * buildBVH/split/updateBVH :1068-1077,1111-1193, serializeBVH/serializeShape
 * :955-1066, bounceSphere/updateWheelAnimations :1079-1109, BoundingBox.hpp,
 * Camera::LookAt camera.hpp:139-163, shape constructors shapes/*.hpp.
 */
#ifndef RT_HOST_H
#define RT_HOST_H

#include "../../include/rt_api.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rh_scene rh_scene;

/* name: "monkey" (1240 shapes, buildBVH(15)), "car" (4022 tris + 100 spheres,
 * buildBVH(25)), "soup" (n_override triangles, default 1 000 000, + n/100
 * spheres, buildBVH(40)), "triangle" (reference scene 3 + a BVH), "toy"
 * (a handful of shapes of every type incl. the NaN-basis floor wall).
 * The model files are absent upstream (.gitignore:78); meshes are synthetic
 * stand-ins with identical primitive counts.
 * bvh_depth <= 0 selects the scene's default. */
rh_scene* rh_scene_create(const char* name, int width, int height, int n_override, int bvh_depth);

/* Wrap caller-provided flat shapes (any mix of types) and build the reference
 * BVH over them.  Infinite planes (type 1) get no bounds, as upstream. */
rh_scene* rh_scene_from_shapes(const rt_flat_shape* shapes, int nShapes,
                               const rt_flat_light* light, const rt_flat_camera* camera, int bvh_depth);
void rh_scene_destroy(rh_scene* s);

/* A Wavefront OBJ as flat triangles (type 3) with `material_from`'s material (NULL: material.hpp:23's defaults), every vertex
 * offset by origin: the reference's model path (model.hpp:49-168 through assimp, mesh.hpp:163-189, main.cpp:645-675, 729-786)
 * restated without assimp — one mesh per object / group / material run with faces, polygons as fans, one vertex per face
 * corner.  apply_rule = 0 gives what main.cpp puts into its scenes (Triangle(a, b, c): the natural plane of the corners),
 * 1 keeps mesh2triangles' normal rule (mesh.hpp:178-184).  Writes at most cap shapes to out and the first triangle of
 * every mesh (+ the total as the last entry) to mesh_first; returns the number of triangles in the file (call again with a
 * larger buffer if it exceeds cap), -2 if the file cannot be opened, -3 on a malformed `v` / `f` record.  Feed the result,
 * together with spheres and walls, to rh_scene_from_shapes. */
int rh_obj_triangles(const char* path, const float origin[3], const rt_flat_shape* material_from, int apply_rule,
                     rt_flat_shape* out, int cap, int32_t* mesh_first, int mesh_cap, int* n_meshes);

const rt_flat_light*  rh_light(const rh_scene* s);
const rt_flat_camera* rh_camera(const rh_scene* s);
const rt_flat_shape*  rh_shapes(const rh_scene* s, int* n);
const rt_flat_node*   rh_nodes(const rh_scene* s, int* n);
const int32_t*        rh_indices(const rh_scene* s, int* n);

/* BVH statistics: out[0]=nodes out[1]=leaves out[2]=max leaf size out[3]=depth */
void rh_bvh_stats(const rh_scene* s, int out[4]);

/* Advance the scene's animation to absolute time t (seconds) with frame delta
 * dt, re-serialise the animated shapes and update the BVH:
 * mode 0 = grow-only refit (reference behaviour, main.cpp:336-345),
 * mode 1 = full rebuild (buildBVH again; node count may change).
 * Returns the number of animated shapes; their indices via rh_animated. */
int rh_scene_animate(rh_scene* s, float t, float dt, int mode);
const int32_t* rh_animated(const rh_scene* s, int* n);

/* The scene's animation list, for callers that animate on the device (rt_anim_bounce / rt_anim_rotate_z):
 * entry k: kind 0 = bounce (shapes[0] = the sphere; params = origin y, amplitude, frequency),
 *          kind 1 = rotation about z (shapes[0..count) = triangles; params = centre x, y, z, speed).
 * rh_anim_count = number of entries; rh_anim_get returns the shape list. */
int rh_anim_count(const rh_scene* s);
const int32_t* rh_anim_get(const rh_scene* s, int k, int* kind, int* count, float params[4]);

/* Camera helper (camera.hpp:139-163): place + LookAt, fov in degrees. */
void rh_set_camera(rh_scene* s, const float pos[3], const float target[3], float fov, float aspect);

/* The interactive camera of the reference's render loop (camera.hpp:30-165; processInput main.cpp:509-533, mouse_callback
 * :536-551): what the host does to the camera between two frames before serializeCamera (:806-816) and the camera
 * sub-upload (:328-335) -> rt_update_camera.  direction: 0 FORWARD, 1 BACKWARD, 2 LEFT, 3 RIGHT, 4 UP, 5 DOWN
 * (camera.hpp:11-18).  Bit-identical to the reference class compiled in place (tests/test_camera.py). */
typedef struct rh_cam rh_cam;
rh_cam* rh_cam_create(const float pos[3], float aspect);          /* Camera(), then Position / aspectRatio */
void rh_cam_destroy(rh_cam* c);
void rh_cam_look_at(rh_cam* c, const float target[3]);            /* Camera::LookAt            :139-150 */
void rh_cam_keyboard(rh_cam* c, int direction, float deltaTime);  /* Camera::ProcessKeyboard   :75-90   */
void rh_cam_mouse(rh_cam* c, float xoffset, float yoffset, int constrainPitch);   /* ProcessMouseMovement :93-112 */
void rh_cam_scroll(rh_cam* c, float yoffset);                     /* Camera::ProcessMouseScroll :115-122 */
void rh_cam_shift(rh_cam* c, int pressed);                        /* MovementSpeed, main.cpp:528 */
void rh_cam_state(const rh_cam* c, float out[4]);                 /* Yaw, Pitch, Zoom, MovementSpeed */
void rh_cam_serialize(const rh_cam* c, rt_flat_camera* out);      /* serializeCamera main.cpp:806-816 */

#ifdef __cplusplus
}
#endif
#endif

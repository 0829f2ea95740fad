/*
 * ref_camera_glue.cpp — TEST INFRASTRUCTURE.  The REFERENCE'S OWN Camera class (src/camera.hpp:30-165: the
 * constructor, LookAt, ProcessKeyboard, ProcessMouseMovement, ProcessMouseScroll, updateCameraVectors), compiled
 * from /root/reference in place, with serializeCamera's field copies (src/main.cpp:806-816) into the reference's
 * own FlatCamera (src/flatStructures.hpp) behind a C interface — the per-frame camera step of the render loop
 * (processInput main.cpp:509-533, mouse_callback :536-551, the camera sub-upload :328-335) that feeds
 * rt_update_camera.  Pins the headless host's rh_cam_* (opengl-ray-tracer_b200/host/rt_host.cpp).
 * Also the reference's Light (light.hpp) + serializeLight (main.cpp:818-823): the light SSBO's 32 bytes.
 * Built by oracle/build_ref.py into oracle/_ref/libref_camera.so; oracle/stubs/glad/glad.h supplies GLboolean.
 */
#include <cstring>
#include "camera.hpp"
#include "light.hpp"
#include "flatStructures.hpp"

extern "C" {

void* refcam_create(const float pos[3], float aspect) {
    Camera* c = new Camera();                       // as generateScene1/2/3 do: Camera(), then Position / aspectRatio, then LookAt
    c->Position = glm::vec3(pos[0], pos[1], pos[2]);
    c->aspectRatio = aspect;
    return c;
}
void refcam_destroy(void* h) { delete static_cast<Camera*>(h); }
void refcam_look_at(void* h, const float t[3]) { static_cast<Camera*>(h)->LookAt(glm::vec3(t[0], t[1], t[2])); }
void refcam_keyboard(void* h, int direction, float dt) { static_cast<Camera*>(h)->ProcessKeyboard((Camera_Movement)direction, dt); }
void refcam_mouse(void* h, float dx, float dy, int constrain) { static_cast<Camera*>(h)->ProcessMouseMovement(dx, dy, (GLboolean)constrain); }
void refcam_scroll(void* h, float dy) { static_cast<Camera*>(h)->ProcessMouseScroll(dy); }
void refcam_shift(void* h, int pressed) { static_cast<Camera*>(h)->MovementSpeed = pressed ? SPEED * SPEEDAMPLIFIER : SPEED; }   // main.cpp:528
void refcam_state(void* h, float out[4]) {
    const Camera* c = static_cast<Camera*>(h);
    out[0] = c->Yaw; out[1] = c->Pitch; out[2] = c->Zoom; out[3] = c->MovementSpeed;
}
/* serializeCamera (main.cpp:806-816): the 80 bytes glBufferSubData sends (padding zeroed here; upstream leaves it indeterminate) */
void refcam_serialize(void* h, void* out80) {
    const Camera& cam = *static_cast<Camera*>(h);
    FlatCamera f;
    std::memset(&f, 0, sizeof f);
    f.Position = cam.Position;
    f.aspectRatio = cam.aspectRatio;
    f.Front = cam.Front;
    f.Up = cam.Up;
    f.Right = cam.Right;
    f.fov = cam.fov;
    static_assert(sizeof(FlatCamera) == 80, "FlatCamera is 80 bytes (SURVEY 8b)");
    std::memcpy(out80, &f, sizeof f);
}

/* Light (light.hpp:6-35) + serializeLight (main.cpp:818-823): the 32 bytes of the light SSBO / rt_update_light */
void reflight_serialize(const float pos[3], const float base[3], float intensity, void* out32) {
    Light light(glm::vec3(pos[0], pos[1], pos[2]), glm::vec3(base[0], base[1], base[2]), intensity);
    FlatLight f;
    std::memset(&f, 0, sizeof f);
    f.position = light.position;
    f.color = light.color;
    static_assert(sizeof(FlatLight) == 32, "FlatLight is 32 bytes (SURVEY 8b)");
    std::memcpy(out32, &f, sizeof f);
}

}  // extern "C"

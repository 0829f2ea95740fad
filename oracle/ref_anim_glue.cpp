/*
 * ref_anim_glue.cpp — TEST INFRASTRUCTURE.  The expression of the reference's updateWheelAnimations (src/main.cpp:1091-1105)
 * evaluated with the reference's vendored GLM, for pinning opengl-ray-tracer_b200/host/rt_anim_math.h.  cos / sin are passed
 * in so that both sides use the same libm values (glm::rotate calls cos / sin itself: it is fed an angle whose cos / sin the
 * caller recomputes identically — std::cos / std::sin of the same float).
 * Built by oracle/build_ref.py into oracle/_ref/libref_anim.so.
 */
#include <cmath>
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>

extern "C" void refanim_wheel(const float* center, float angle, const float* axis, int n, const float* pts, float* out, float* matrix) {
    glm::vec3 c(center[0], center[1], center[2]);
    glm::mat4 translateToOrigin = glm::translate(glm::mat4(1.0f), -c);
    glm::mat4 translateBack = glm::translate(glm::mat4(1.0f), c);
    glm::mat4 rotationMatrix = glm::rotate(glm::mat4(1.0f), angle, glm::vec3(axis[0], axis[1], axis[2]));
    glm::mat4 transform = translateBack * rotationMatrix * translateToOrigin;
    for (int i = 0; i < n; ++i) {
        glm::vec3 p = glm::vec3(transform * glm::vec4(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], 1.0f));
        out[3 * i] = p.x; out[3 * i + 1] = p.y; out[3 * i + 2] = p.z;
    }
    for (int col = 0; col < 4; ++col) for (int r = 0; r < 4; ++r) matrix[4 * col + r] = transform[col][r];
}

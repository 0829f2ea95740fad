/*
 * rt_anim_math.h — the arithmetic of the reference's wheel animation (updateWheelAnimations, src/main.cpp:1084-1109):
 *     transform = translate(I, center) * rotate(I, angle, axis) * translate(I, -center);   p' = vec3(transform * vec4(p, 1))
 * restated operation by operation in GLM 0.9.8.5's own order (gtc/matrix_transform.inl:10-50, detail/type_mat4x4.inl
 * :526-537 and :595-613), so that it can be evaluated without GLM and — the 4x4 matrix once per wheel on the host, the
 * matrix-vector product per vertex in a kernel — on the device, bit for bit.
 *
 * USED BY the headless host (rt_host.cpp: rotate_wheels) and by rt_animate (rt_api.cu: the 12 matrix entries per wheel that K6,
 * rt_anim.cuh, applies per vertex).  tests/test_host.py pins this header to GLM itself; tests/test_ref_builder.py pins the
 * animated scene to the reference's own updateWheelAnimations compiled in place (oracle/_ref/libref_bvh.so).
 * Column-major float[16] as in GLM: m[4 * col + row].  Pure C++ (host and device).
 */
#ifndef RT_ANIM_MATH_H
#define RT_ANIM_MATH_H

#include <math.h>

#if defined(__CUDACC__)
#define RTA_HD __host__ __device__ inline
#else
#define RTA_HD inline
#endif

RTA_HD void rta_identity(float m[16]) { for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? 1.0f : 0.0f; }

/* glm::translate(m, v): Result = m; Result[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3] */
RTA_HD void rta_translate(const float m[16], const float v[3], float out[16]) {
    for (int i = 0; i < 16; ++i) out[i] = m[i];
    for (int r = 0; r < 4; ++r) out[12 + r] = ((m[r] * v[0] + m[4 + r] * v[1]) + m[8 + r] * v[2]) + m[12 + r];
}

/* glm::rotate(m, angle, v) given c = cos(angle), s = sin(angle) (the caller's libm) */
RTA_HD void rta_rotate(const float m[16], float c, float s, const float v[3], float out[16]) {
    const float inv = 1.0f / sqrtf((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);        /* normalize = v * inversesqrt(dot(v, v)) */
    const float ax[3] = {v[0] * inv, v[1] * inv, v[2] * inv};
    const float omc = 1.0f - c;
    const float t[3] = {omc * ax[0], omc * ax[1], omc * ax[2]};
    float R[3][3];
    R[0][0] = c + t[0] * ax[0];          R[0][1] = t[0] * ax[1] + s * ax[2];  R[0][2] = t[0] * ax[2] - s * ax[1];
    R[1][0] = t[1] * ax[0] - s * ax[2];  R[1][1] = c + t[1] * ax[1];          R[1][2] = t[1] * ax[2] + s * ax[0];
    R[2][0] = t[2] * ax[0] + s * ax[1];  R[2][1] = t[2] * ax[1] - s * ax[0];  R[2][2] = c + t[2] * ax[2];
    for (int col = 0; col < 3; ++col)
        for (int r = 0; r < 4; ++r) out[4 * col + r] = (m[r] * R[col][0] + m[4 + r] * R[col][1]) + m[8 + r] * R[col][2];
    for (int r = 0; r < 4; ++r) out[12 + r] = m[12 + r];
}

/* glm mat4 * mat4: Result[col] = ((A[0] * B[col][0] + A[1] * B[col][1]) + A[2] * B[col][2]) + A[3] * B[col][3] */
RTA_HD void rta_mul(const float a[16], const float b[16], float out[16]) {
    for (int col = 0; col < 4; ++col)
        for (int r = 0; r < 4; ++r)
            out[4 * col + r] = ((a[r] * b[4 * col] + a[4 + r] * b[4 * col + 1]) + a[8 + r] * b[4 * col + 2]) + a[12 + r] * b[4 * col + 3];
}

/* vec3(m * vec4(p, 1)): (m[0] * x + m[1] * y) + (m[2] * z + m[3] * 1) */
RTA_HD void rta_apply(const float m[16], const float p[3], float out[3]) {
    for (int r = 0; r < 3; ++r) out[r] = (m[r] * p[0] + m[4 + r] * p[1]) + (m[8 + r] * p[2] + m[12 + r] * 1.0f);
}

/* the wheel transform of one frame: translate(I, center) * rotate(I, angle, axis) * translate(I, -center) */
RTA_HD void rta_wheel_transform(const float center[3], float c, float s, const float axis[3], float out[16]) {
    float I[16], t0[16], t1[16], rot[16], tmp[16];
    const float neg[3] = {-center[0], -center[1], -center[2]};
    rta_identity(I);
    rta_translate(I, neg, t0);
    rta_translate(I, center, t1);
    rta_rotate(I, c, s, axis, rot);
    rta_mul(t1, rot, tmp);
    rta_mul(tmp, t0, out);
}

#endif /* RT_ANIM_MATH_H */

// anim_math_emu.cpp — TEST INFRASTRUCTURE: rt_anim_math.h behind a C function (tests/test_host.py compares it with GLM).
#include <cmath>
#include "../../opengl-ray-tracer_b200/host/rt_anim_math.h"
extern "C" void rta_wheel(const float* center, float angle, const float* axis, int n, const float* pts, float* out, float* matrix) {
    rta_wheel_transform(center, std::cos(angle), std::sin(angle), axis, matrix);
    for (int i = 0; i < n; ++i) rta_apply(matrix, pts + 3 * i, out + 3 * i);
}

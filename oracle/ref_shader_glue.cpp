/*
 * ref_shader_glue.cpp — TEST INFRASTRUCTURE.  Compiles the REFERENCE SHADER
 * FILE ITSELF (src/shaders/gpu_shader.comp) as C++ against the reference's
 * vendored GLM, so the hand restatement in oracle_rt.cpp can be pinned against
 * "the reference run here" (there is no graphics API on these machines).
 *
 * oracle/build_ref.py produces oracle/_ref/gpu_shader_body.inc from the file
 * where it lies under /root/reference (a mechanical text transform: f-suffix on
 * float literals, `out float` -> `float&`, swizzle store -> vec4 ctor, the
 * layout(...) block -> REF_BUFFERS) and compiles this glue around it into
 * oracle/_ref/libref_shader.so.  Nothing of the reference is copied into the
 * repository: _ref/ is git-ignored.
 *
 * Known, documented deviations of this build from GLSL-spec semantics (they
 * come from GLM 0.9.8.5, glm/glm/detail/func_common.inl): mix(x,y,a) is
 * x + a*(y-x); min/max return the second operand for NaN in the first.
 * A second build with -DREF_GLSL_BUILTINS (libref_shader_glsl.so) defines those
 * built-ins as the GLSL specification words them; tests/test_oracle_golden.py
 * compares the restatement with it on NaN-laden inputs.
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include <glm/glm.hpp>

#include "../include/rt_api.h"

namespace refshader {
using namespace glm;
typedef unsigned int uint;

inline float pow(float a, int b) { return std::pow(a, (float)b); }
inline float pow(float a, float b) { return std::pow(a, b); }

#ifdef REF_GLSL_BUILTINS
// Second build (libref_shader_glsl.so): the same shader text with the built-ins the shader calls defined as the GLSL 4.60
// specification words them (8.3 Common Functions) instead of GLM's forms: min(x,y) = y if y < x, otherwise x; max(x,y) = y
// if x < y, otherwise x; clamp = min(max(x, lo), hi); mix = x*(1-a) + y*a.  The two builds agree wherever no NaN or signed
// zero reaches these functions; with a NaN operand GLM's `x < y ? x : y` / `x > y ? x : y` keep the OTHER operand than the
// specification's selects.  Declared here they hide GLM's templates for unqualified calls from the shader body.
inline float min(float x, float y) { return y < x ? y : x; }
inline float max(float x, float y) { return x < y ? y : x; }
inline vec3 min(vec3 a, vec3 b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(vec3 a, vec3 b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float clamp(float x, float lo, float hi) { return min(max(x, lo), hi); }
inline vec3 mix(vec3 x, vec3 y, float a) { return x * (1.0f - a) + y * a; }
#endif

template <class T>
struct SSBO {
    const T* p = nullptr;
    int n = 0;
    int length() const { return n; }
    const T& operator[](int i) const { return p[i]; }
};

struct Image { float* rgba = nullptr; int W = 0, H = 0; };
inline void imageStore(Image& img, ivec2 c, vec4 v) {
    float* o = img.rgba + 4 * ((size_t)c.y * (size_t)img.W + (size_t)c.x);
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
}

static thread_local ivec2 gl_GlobalInvocationID_xy;
static Image imgOutput;

#define uniform static
#define main shader_main
#define REF_BUFFERS                         \
    static Light light;                     \
    static Camera camera;                   \
    static SSBO<Shape> shapes;              \
    static SSBO<Node> bvhNodes;             \
    static SSBO<int> bvhIndices;

#include "_ref/gpu_shader_body.inc"

#undef main
#undef uniform

static_assert(sizeof(Light) == 32, "Light");
static_assert(sizeof(Camera) == 80, "Camera");
static_assert(sizeof(Material) == 32, "Material");
static_assert(sizeof(Node) == 48, "Node");

static std::vector<Shape> g_shapes;
}  // namespace refshader

extern "C" int ref_render(const rt_flat_light* light, const rt_flat_camera* camera,
                          const rt_flat_shape* shapes, int nShapes,
                          const rt_flat_node* nodes, int nNodes,
                          const int32_t* indices, int nIndices,
                          const rt_params* params, int W, int H,
                          int row_step, int row_phase, float* rgba, int threads) {
    using namespace refshader;
    if (!light || !camera || !params || !rgba || nNodes <= 0) return -1;   // root = length()-1 must exist
    std::memcpy(&refshader::light, light, 32);
    std::memcpy(&refshader::camera, camera, 80);
    g_shapes.assign((size_t)nShapes, Shape());
    for (int i = 0; i < nShapes; ++i) {                   // unpack std430 FlatShape (192 B) -> C++ Shape
        const rt_flat_shape& f = shapes[i];
        Shape& s = g_shapes[(size_t)i];
        s.type = f.type;
        std::memcpy(&s.material, &f.material, 32);
        s.sphereCenter = vec3(f.sphereCenter[0], f.sphereCenter[1], f.sphereCenter[2]);
        s.sphereRadius = f.sphereRadius;
        s.planeNormal = vec3(f.planeNormal[0], f.planeNormal[1], f.planeNormal[2]);
        s.planeD = f.planeD;
        s.wallStart = vec3(f.wallStart[0], f.wallStart[1], f.wallStart[2]);
        s.wallWidth = f.wallWidth;
        s.wallHeight = f.wallHeight;
        s.triP1 = vec3(f.triP1[0], f.triP1[1], f.triP1[2]);
        s.triP2 = vec3(f.triP2[0], f.triP2[1], f.triP2[2]);
        s.triP3 = vec3(f.triP3[0], f.triP3[1], f.triP3[2]);
    }
    refshader::shapes.p = g_shapes.data();  refshader::shapes.n = nShapes;
    refshader::bvhNodes.p = reinterpret_cast<const Node*>(nodes);  refshader::bvhNodes.n = nNodes;
    refshader::bvhIndices.p = indices;  refshader::bvhIndices.n = nIndices;
    screenRes = vec2(params->screenRes[0], params->screenRes[1]);
    maxBounces = params->maxBounces;
    useBVH = params->useBVH != 0;
    useFresnel = params->useFresnel != 0;
    useMollerTrumbore = params->useMollerTrumbore != 0;
    imgOutput.rgba = rgba; imgOutput.W = W; imgOutput.H = H;

    std::vector<int> rows;
    for (int y = 0; y < H; ++y) if (y % row_step == row_phase) rows.push_back(y);
    if (threads < 1) threads = 1;
    auto work = [&](int tid) {
        for (size_t r = (size_t)tid; r < rows.size(); r += (size_t)threads)
            for (int x = 0; x < W; ++x) {
                gl_GlobalInvocationID_xy = ivec2(x, rows[r]);
                shader_main();
            }
    };
    if (threads == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& t : pool) t.join();
    }
    return 0;
}

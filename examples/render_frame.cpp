// render_frame.cpp — the reference's GPU frame (src/main.cpp:328-361) driven through the C ABI from
// C++, headless: build a scene with the host library (the stand-in for generateScene*/buildBVH/
// serializeScene), upload it, render, optionally animate a few frames the way the reference's
// frame loop does, and write the rgba32f image as a PFM and the displayed frame (rt_render_rgba8: K1 + the display quad's
// arithmetic, K5) as <out>.ppm.
//
//   g++ -std=c++17 -O2 -Iinclude examples/render_frame.cpp -o render_frame \
//       opengl-ray-tracer_b200/csrc/librt_b200.so opengl-ray-tracer_b200/host/librt_host.so
//   ./render_frame car 800 600 out.pfm [frames]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../include/rt_api.h"
#include "../opengl-ray-tracer_b200/host/rt_host.h"

static int fail(rt_ctx* c, const char* what, int rc) {
    std::fprintf(stderr, "%s failed (%d): %s\n", what, rc, c ? rt_last_error(c) : "");
    return 1;
}

int main(int argc, char** argv) {
    const char* name = argc > 1 ? argv[1] : "car";
    const int W = argc > 2 ? std::atoi(argv[2]) : 800, H = argc > 3 ? std::atoi(argv[3]) : 600;
    const char* out = argc > 4 ? argv[4] : "frame.pfm";
    const int frames = argc > 5 ? std::atoi(argv[5]) : 1;

    rh_scene* sc = rh_scene_create(name, W, H, 0, 0);                        // generateSceneN + buildBVH + serializeScene
    if (!sc) { std::fprintf(stderr, "unknown scene %s\n", name); return 1; }
    int nShapes = 0, nNodes = 0, nIdx = 0, nAnim = 0;
    const rt_flat_shape* shapes = rh_shapes(sc, &nShapes);
    const rt_flat_node* nodes = rh_nodes(sc, &nNodes);
    const int32_t* idx = rh_indices(sc, &nIdx);

    rt_ctx* rt = nullptr;                                                    // replaces ComputeShader + texture + SSBOs
    int rc = rt_create(&rt, W, H, 0, RT_FLAG_DEFAULT);
    if (rc != RT_OK) return fail(nullptr, "rt_create (no CUDA device? there is no CPU fallback)", rc);
    rc = rt_upload_scene(rt, rh_light(sc), rh_camera(sc), shapes, nShapes, nodes, nNodes, idx, nIdx);
    if (rc != RT_OK) return fail(rt, "rt_upload_scene", rc);

    rt_params p = {{(float)W, (float)H}, 3, 1, 0, 0};                        // main.cpp:118,135-137 defaults
    std::vector<float> frame((size_t)W * H * 4);
    for (int f = 0; f < frames; ++f) {
        if (f > 0) {                                                         // main.cpp:336-345, 438-457
            rh_scene_animate(sc, f / 60.0f, 1.0f / 60.0f, 0);
            const int32_t* anim = rh_animated(sc, &nAnim);
            shapes = rh_shapes(sc, &nShapes); nodes = rh_nodes(sc, &nNodes);
            for (int i = 0; i < nAnim; ++i)
                if ((rc = rt_update_shapes(rt, anim[i], 1, shapes + anim[i])) != RT_OK) return fail(rt, "rt_update_shapes", rc);
            if ((rc = rt_update_nodes(rt, nodes, nNodes)) != RT_OK) return fail(rt, "rt_update_nodes", rc);
        }
        if ((rc = rt_render(rt, &p, frame.data(), nullptr)) != RT_OK) return fail(rt, "rt_render", rc);
    }
    float ms[4];
    rt_last_timings(rt, ms);
    std::printf("%s %dx%d: %d frame(s), last render %.3f ms, d2h %.3f ms, %lld kernel launches, %s\n", name, W, H, frames, ms[2], ms[3],
                (long long)rt_launch_count(rt), rt_build_info());

    if (FILE* fp = std::fopen(out, "wb")) {                                   // PFM is bottom-up; row 0 of the image is the +Up edge
        std::fprintf(fp, "PF\n%d %d\n-1.0\n", W, H);
        for (int y = H - 1; y >= 0; --y)
            for (int x = 0; x < W; ++x) std::fwrite(&frame[((size_t)y * W + x) * 4], sizeof(float), 3, fp);
        std::fclose(fp);
    }
    // the frame as the reference shows it (renderQuad main.cpp:474-501 + shader.frag:8-12): <out>.ppm, row 0 = the +Up edge
    std::vector<uint8_t> shown((size_t)W * H * 4);
    if ((rc = rt_render_rgba8(rt, &p, shown.data(), nullptr)) != RT_OK) return fail(rt, "rt_render_rgba8", rc);
    if (FILE* fp = std::fopen((std::string(out) + ".ppm").c_str(), "wb")) {
        std::fprintf(fp, "P6\n%d %d\n255\n", W, H);
        for (size_t i = 0; i < (size_t)W * H; ++i) std::fwrite(&shown[4 * i], 1, 3, fp);
        std::fclose(fp);
    }
    rt_destroy(rt);
    rh_scene_destroy(sc);
    return 0;
}

// sanitize_walk_main.cpp — TEST INFRASTRUCTURE: the scalar emulation of the render kernel's BVH query (walk_emu.cpp: the layout
// of rt_layout.h walked with the primitive / box tests of rt_device.cuh, the kernel's own source) as a plain program for
// AddressSanitizer / UndefinedBehaviorSanitizer: every slot, node and link the walk touches is checked.  Rays: a camera-like
// fan towards the scene plus rays from inside it, closest-hit and any-hit, both triangle modes, lists A and B.
// Built and run by tests/test_sanitizers.py.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "walk_emu.cpp"
#include "../../opengl-ray-tracer_b200/host/rt_host.h"
int main(int argc, char** argv) {
    const char* name = argc > 1 ? argv[1] : "car"; const int n = argc > 2 ? atoi(argv[2]) : 0;
    rh_scene* sc = rh_scene_create(name, 800, 600, n, 0);
    if (!sc) return 2;
    int ns, nn, ni; const rt_flat_shape* sh = rh_shapes(sc, &ns); const rt_flat_node* nd = rh_nodes(sc, &nn); const int32_t* ix = rh_indices(sc, &ni);
    unsigned long long total = 0;
    for (int flags : {0, 1, 3}) {                       // reference walk / sub-trees / sub-trees + list B
        void* e = emu_create(sh, ns, nd, nn, ix, ni, flags, 200.0);
        if (!e) return 3;
        std::vector<float> o, d, bound;
        unsigned s = 12345u;
        auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (float)(s >> 8) / 16777216.0f; };
        for (int i = 0; i < 4000; ++i) {
            const bool inside = i % 4 == 3;
            const float ox = inside ? (rnd() - 0.5f) * 20.f : (rnd() - 0.5f) * 4.f, oy = inside ? -rnd() * 10.f : -10.f + (rnd() - 0.5f) * 4.f,
                        oz = inside ? (rnd() - 0.5f) * 10.f : 40.f;
            float dx = (rnd() - 0.5f) * (inside ? 2.f : 0.9f), dy = (rnd() - 0.5f) * (inside ? 2.f : 0.6f) + (inside ? 0.f : 0.25f), dz = inside ? rnd() - 0.5f : -1.f;
            const float inv = 1.0f / sqrtf(dx * dx + dy * dy + dz * dz);
            o.insert(o.end(), {ox, oy, oz}); d.insert(d.end(), {dx * inv, dy * inv, dz * inv}); bound.push_back(5.f + 60.f * rnd());
        }
        const int m = (int)bound.size();
        std::vector<int> shape((size_t)m); std::vector<float> dist((size_t)m), pt(3 * (size_t)m);
        uint64_t cnt[2];
        for (int mt = 0; mt < 2; ++mt)
            for (int any = 0; any < 2; ++any) {
                emu_query(e, m, o.data(), d.data(), mt, any, bound.data(), shape.data(), dist.data(), pt.data(), cnt);
                int hits = 0; for (int v : shape) hits += v >= 0;
                total += cnt[0] + cnt[1] + (unsigned long long)hits;
                printf("%s flags %d mt %d any %d: %d rays, %d hits, %llu node visits, %llu primitive tests\n", name, flags, mt, any, m, hits,
                       (unsigned long long)cnt[0], (unsigned long long)cnt[1]);
            }
        emu_destroy(e);
    }
    rh_scene_destroy(sc);
    return total > 0 ? 0 : 4;
}

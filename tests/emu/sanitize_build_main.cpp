// sanitize_build_main.cpp — TEST INFRASTRUCTURE: the emulated K8 / K9 bodies (build_emu.cpp) as a plain program, so that they can
// run under AddressSanitizer / UndefinedBehaviorSanitizer (compute-sanitizer is closed on the GPU pool): the index arithmetic of
// the kernels is exactly this source.  Built and run by tests/test_sanitizers.py.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "build_emu.cpp"
#include "../../opengl-ray-tracer_b200/host/rt_host.h"
int main(int argc, char** argv) {
    const char* name = argv[1]; int n = atoi(argv[2]); int depth = atoi(argv[3]);
    rh_scene* sc = rh_scene_create(name, 800, 600, n, depth);
    int ns, nn, ni; const rt_flat_shape* sh = rh_shapes(sc, &ns); const rt_flat_node* nd = rh_nodes(sc, &nn); const int32_t* ix = rh_indices(sc, &ni);
    std::vector<rt_flat_node> out((size_t)2 * ns); std::vector<int32_t> idx((size_t)ns);
    std::vector<int> src(2 * (size_t)ns), lc(4 * (size_t)ns), ps((size_t)ns), pr((size_t)ns);
    for (unsigned order : {0u, 3u}) {
        int m = emu_build_bvh_list(sh, ns, depth, out.data(), idx.data(), order, src.data(), lc.data(), ps.data(), pr.data());
        bool same = m == nn && memcmp(out.data(), nd, (size_t)nn * sizeof(rt_flat_node)) == 0 && memcmp(idx.data(), ix, (size_t)ni * 4) == 0;
        printf("%s order %u: nodes %d (host %d) identical %d\n", name, order, m, nn, (int)same);
    }
    rh_scene_destroy(sc);
}

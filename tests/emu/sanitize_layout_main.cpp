// sanitize_layout_main.cpp — TEST INFRASTRUCTURE: the host layout builder (rt_layout.h: list A, fat-leaf sub-trees, list B with
// the multi-threaded SAH builder, refit graph) as a plain program for ThreadSanitizer / AddressSanitizer runs.
// Prints the FNV-1a hash of every layout array (the value tests/test_layout_emu.py pins).  Built and run by tests/test_sanitizers.py.
#include <cstdio>
#include <cstdlib>
#include "../../opengl-ray-tracer_b200/csrc/rt_layout.h"
#include "../../opengl-ray-tracer_b200/host/rt_host.h"
int main(int argc, char** argv) {
    const char* name = argc > 1 ? argv[1] : "car"; const int n = argc > 2 ? atoi(argv[2]) : 0;
    const int depth = argc > 3 ? atoi(argv[3]) : 0;            // 0: the scene's own BVH depth
    rh_scene* sc = rh_scene_create(name, 800, 600, n, depth);
    if (!sc) return 2;
    int ns, nn, ni; const rt_flat_shape* sh = rh_shapes(sc, &ns); const rt_flat_node* nd = rh_nodes(sc, &nn); const int32_t* ix = rh_indices(sc, &ni);
    rtl::Layout L;
    if (!rtl::build_layout(nd, nn, ix, (size_t)ni, sh, ns, true, 200.0, L)) return 3;
    rtl::build_top(L, nd, nn);
    rtl::build_refit_graph(L);
    unsigned long long h = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t bytes) { const unsigned char* c = (const unsigned char*)p; for (size_t i = 0; i < bytes; ++i) { h ^= c[i]; h *= 1099511628211ull; } };
    mix(L.src.data(), L.src.size() * 4); mix(L.link.data(), L.link.size() * 4); mix(L.cw.data(), L.cw.size() * 4);
    mix(L.pshape.data(), L.pshape.size() * 4); mix(L.prank.data(), L.prank.size() * 4);
    mix(L.subBox.data(), L.subBox.size() * sizeof(rtl::Box));
    mix(L.srcB.data(), L.srcB.size() * 4); mix(L.linkB.data(), L.linkB.size() * 4); mix(L.cwB.data(), L.cwB.size() * 4);
    mix(L.rfParent.data(), L.rfParent.size() * 4); mix(L.rfChild.data(), L.rfChild.size() * 4);
    mix(L.rfLeaf.data(), L.rfLeaf.size() * 4); mix(L.rfStart.data(), L.rfStart.size() * 4);
    printf("layout hash %016llx\n", h);
    rh_scene_destroy(sc);
    return 0;
}

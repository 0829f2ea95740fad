// build_emu.cpp — TEST INFRASTRUCTURE.  Runs the per-thread bodies of the device BVH builder
// (opengl-ray-tracer_b200/csrc/rt_build.cuh, namespace rt::bld — the very source nvcc compiles into K8)
// on the host, one "thread" at a time and one "kernel" after the other in the order rt_api.cu launches
// them, so that the CPU suite can compare the emitted FlatNode[] / bvhIndices[] with the host builder's
// byte for byte.  `threadOrder` != 0 walks the threads of every kernel in a scrambled order: the result
// must not depend on it (atomics, allocation order of the build ids).  Never loaded by the product.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <random>
#include <vector>

#include "../../opengl-ray-tracer_b200/csrc/rt_build.cuh"
#include "../../opengl-ray-tracer_b200/csrc/rt_build_list.cuh"
#include "../../opengl-ray-tracer_b200/csrc/rt_layout.h"

using namespace rt;

// listSrc .. listPrank (optional, 2n / 2n*2 / n / n ints): K9's list A from the same tables (rt_build_list.cuh)
extern "C" int emu_build_bvh_list(const rt_flat_shape* shapes, int n, int maxDepth, rt_flat_node* outNodes, int32_t* outIdx,
                                  unsigned threadOrder, int* listSrc, int* listLinkCw, int* listPshape, int* listPrank) {
    if (n <= 0) return -1;
    const size_t maxNodes = 2 * (size_t)n;
    std::vector<int> start(maxNodes), count(maxNodes), left(maxNodes), parent(maxNodes), depth(maxNodes), desc(maxNodes), idx(maxNodes), first(maxNodes);
    std::vector<bld::u64> key(6 * maxNodes);
        std::vector<unsigned> arrived(maxNodes, 0u);
    bld::Nodes N{start.data(), count.data(), left.data(), parent.data(), depth.data(), key.data(), desc.data(), arrived.data(), idx.data(), first.data()};
    bld::Ctl ctl;
    std::vector<int> ord[2] = {std::vector<int>((size_t)n), std::vector<int>((size_t)n)};
    std::vector<int> pnode[2] = {std::vector<int>((size_t)n), std::vector<int>((size_t)n)};
    std::vector<int> flag((size_t)n), scanL((size_t)n + 1);
    std::vector<float4> cen((size_t)n), rmn((size_t)n), rmx((size_t)n);
    bld::Records R{cen.data(), rmn.data(), rmx.data()};

    std::mt19937 rng(threadOrder);
    auto threads = [&](int m) {
        std::vector<int> t((size_t)m);
        std::iota(t.begin(), t.end(), 0);
        if (threadOrder) std::shuffle(t.begin(), t.end(), rng);
        return t;
    };

    const int nIds = 2 * n;
    for (int id : threads(nIds)) bld::setup_id(id, N, &ctl, n, maxDepth);
    for (int p : threads(n)) {
        bld::u64 k[6];
        bld::init_position(p, shapes, R, ord[0].data(), pnode[0].data(), k);
        bld::fold_keys(N.key, k);
    }
    for (int level = 0; level < maxDepth; ++level) {
        if (!ctl.active[level]) continue;
        const int cur = level & 1, nxt = cur ^ 1;
        for (int p : threads(n)) flag[(size_t)p] = bld::classify(p, R, ord[cur].data(), pnode[cur].data(), N);
        int run = 0;
        for (int p = 0; p < n; ++p) { scanL[(size_t)p] = run; run += flag[(size_t)p]; }
        scanL[(size_t)n] = run;
        for (int p : threads(n)) {
            int newPos, newNode; bool moved;
            bld::split_position(p, flag[(size_t)p], level, pnode[cur].data(), scanL.data(), N, &ctl, newPos, newNode, moved);
            const int shape = ord[cur][(size_t)p];
            ord[nxt][(size_t)newPos] = shape;
            pnode[nxt][(size_t)newPos] = newNode;
            if (moved) {
                bld::u64 k[6];
                bld::record_keys(rmn[(size_t)shape], rmx[(size_t)shape], (unsigned)newPos, k);
                bld::fold_keys(N.key + 6 * (size_t)newNode, k);
            }
        }
        ctl.parity = nxt;
    }
    for (int id : threads(nIds)) bld::count_descendants(id, N);
    for (int id : threads(nIds)) bld::number_node(id, N, &ctl);
    for (int id : threads(nIds)) bld::emit_node(id, N, ord[ctl.parity].data(), outNodes, outIdx);
    const int nNodes = ctl.nNodes;
    if (listSrc)
        for (int id : threads(nIds))
            bld::thread_node(id, N, ord[ctl.parity].data(), shapes, listSrc, reinterpret_cast<int2*>(listLinkCw), listPshape, listPrank);
    return nNodes;
}

extern "C" int emu_build_bvh(const rt_flat_shape* shapes, int n, int maxDepth, rt_flat_node* outNodes, int32_t* outIdx,
                             unsigned threadOrder) {
    return emu_build_bvh_list(shapes, n, maxDepth, outNodes, outIdx, threadOrder, nullptr, nullptr, nullptr, nullptr);
}

// The host layout builder's list A without sub-trees (the plain reference walk), for comparison with K9.
extern "C" int emu_host_list(const rt_flat_shape* shapes, int n, const rt_flat_node* nodes, int nNodes, const int32_t* idx, int nIdx,
                             int* src, int* linkCw, int* pshape, int* prank) {
    rtl::Layout L;
    if (!rtl::build_layout(nodes, nNodes, idx, (size_t)nIdx, shapes, n, false, 1.0, L)) return -1;
    for (int k = 0; k < L.nodes(); ++k) { src[k] = L.src[(size_t)k]; linkCw[2 * k] = L.link[(size_t)k]; linkCw[2 * k + 1] = L.cw[(size_t)k]; }
    for (int s = 0; s < L.prims(); ++s) { pshape[s] = L.pshape[(size_t)s]; prank[s] = L.prank[(size_t)s]; }
    return L.nodes();
}

// split's centre of every shape as K8 computes it (rt_build.cuh: shape_record), for pinning against the reference's classes
extern "C" void emu_shape_centres(const rt_flat_shape* shapes, int n, float* cen) {
    for (int i = 0; i < n; ++i) {
        float4 c, mn, mx;
        bld::shape_record(shapes + i, c, mn, mx);
        cen[3 * i] = c.x; cen[3 * i + 1] = c.y; cen[3 * i + 2] = c.z;
    }
}

// walk_emu.cpp — TEST INFRASTRUCTURE.  A scalar host emulation of the render kernel's BVH query
// (opengl-ray-tracer_b200/csrc/rt_kernels.cuh: walk) over the layout built by rt_layout.h, using
// the very same primitive/box test source (rt_device.cuh compiled by g++ -ffp-contract=off).
// It lets the CPU suite check, ray by ray against the oracle, that the conservative sub-trees and
// the rank tie-break change no decision.  Never loaded by the product.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../opengl-ray-tracer_b200/csrc/rt_device.cuh"
#include "../../opengl-ray-tracer_b200/csrc/rt_layout.h"

using namespace rt;

namespace {
struct Emu {
    rtl::Layout L;
    std::vector<Prim> prims;
    std::vector<float4> nodes;     // 2 per node (list A)
    std::vector<float4> nodesB;    // list B (own tree over the reference leaves), empty if unused
    const rt_flat_node* raw = nullptr;
    bool boxesFinite = true;
    int nNodes = 0;
    float omax = 0.f;
};

template <bool MT>
void query(const Emu& E, const float* o, const float* d, int anyMode, float bound, int* outShape, float* outDist,
           float* outP, uint64_t* cnt) {
    Ray r = make_ray(mk(o), mk(d), E.boxesFinite, E.omax);
    float best = anyMode ? bound : 1e20f;
    int slot = -1; V3 bp = mk(0.f, 0.f, 0.f);
    // well-behaved rays walk list B when it exists (the kernel decides per warp; per ray here)
    const bool useB = r.sub && !E.nodesB.empty();
    const std::vector<float4>& N = useB ? E.nodesB : E.nodes;
    const std::vector<int>& CW = useB ? E.L.cwB : E.L.cw;
    const std::vector<int>& LNK = useB ? E.L.linkB : E.L.link;
    const int n = (int)CW.size();
    int k = 0;
    while (k < n) {
        const float4 a = N[2 * (size_t)k], b = N[2 * (size_t)k + 1];
        const int cw = CW[(size_t)k], lnk = LNK[(size_t)k];
        cnt[0]++;
        const bool ours = cw == RT_INNER_SUB || (cw >= 0 && (cw & RT_LEAF_SUB_BIT));
        const bool hit = ours ? (!r.sub || slab_fast(r, a, b)) : (r.fast ? slab_fast(r, a, b) : slab(r, a, b));
        if (cw < 0) { k = hit ? k + 1 : lnk; continue; }
        k = k + 1;
        if (!hit) continue;
        const int c = cw & RT_LEAF_COUNT_MASK;
        cnt[1] += (uint64_t)c;
        bool stop = false;
        for (int s = lnk; s < lnk + c && !stop; ++s) {
            V3 hp;
            if (!test_prim<MT>(E.prims[(size_t)s], r, hp)) continue;
            const float dist = distance3(r.o, hp);
            bool take = dist < best;
            if (!anyMode && dist == best && slot >= 0) take = E.L.prank[(size_t)s] < E.L.prank[(size_t)slot];
            if (take) { best = dist; slot = s; bp = hp; if (anyMode) stop = true; }
        }
        if (stop) break;
    }
    *outShape = slot >= 0 ? E.L.pshape[(size_t)slot] : -1;
    *outDist = best;
    if (outP) { outP[0] = bp.x; outP[1] = bp.y; outP[2] = bp.z; }
}
}  // namespace

static void fill_nodes(Emu* E, const rt_flat_node* nodes) {
    auto fill = [&](const std::vector<int>& src, std::vector<float4>& out) {
        out.resize(2 * src.size());
        for (size_t k = 0; k < src.size(); ++k) {
            const int s = src[k];
            if (s >= 0) {
                out[2 * k] = make_float4(nodes[s].boundsMin[0], nodes[s].boundsMin[1], nodes[s].boundsMin[2], 0.f);
                out[2 * k + 1] = make_float4(nodes[s].boundsMax[0], nodes[s].boundsMax[1], nodes[s].boundsMax[2], 0.f);
            } else {
                const rtl::Box& b = E->L.subBox[(size_t)(-1 - s)];
                out[2 * k] = make_float4(b.mn[0], b.mn[1], b.mn[2], 0.f);
                out[2 * k + 1] = make_float4(b.mx[0], b.mx[1], b.mx[2], 0.f);
            }
        }
    };
    fill(E->L.src, E->nodes);
    if (E->L.useB) fill(E->L.srcB, E->nodesB); else E->nodesB.clear();
}

extern "C" {
// rtl::scan_nodes against the three separate passes it replaced (rt_api.cu: same_topology, the safe-coordinate loop, rtl::ref_tree_nested)
int emu_scan_nodes(const rt_flat_node* in, const rt_flat_node* old, int n, int* out) {
    const rtl::NodeScan r = rtl::scan_nodes(in, old, n, 1e37f);
    out[0] = r.sameTopo; out[1] = r.safe; out[2] = r.nested;
    bool same = old != nullptr;
    for (int i = 0; same && i < n; ++i)
        if (in[i].leftChild != old[i].leftChild || in[i].rightChild != old[i].rightChild ||
            (in[i].leftChild == -1 && (in[i].startShapeIdx != old[i].startShapeIdx || in[i].numShapes != old[i].numShapes))) same = false;
    bool safe = true;
    for (int i = 0; i < n && safe; ++i)
        for (int k = 0; k < 3; ++k)
            if (!(std::fabs(in[i].boundsMin[k]) <= 1e37f) || !(std::fabs(in[i].boundsMax[k]) <= 1e37f)) safe = false;
    out[3] = same; out[4] = safe; out[5] = rtl::ref_tree_nested(in, n);
    return 1;
}

void* emu_create(const rt_flat_shape* shapes, int nShapes, const rt_flat_node* nodes, int nNodes,
                 const int32_t* indices, int nIndices, int subtrees, double omax) {
    const bool topTree = (subtrees & 2) != 0;
    subtrees &= 1;
    Emu* E = new Emu();
    for (int i = 0; i < nNodes && E->boxesFinite; ++i)
        for (int k = 0; k < 3; ++k)
            if (!(fabsf(nodes[i].boundsMin[k]) <= RT_SAFE_COORD) || !(fabsf(nodes[i].boundsMax[k]) <= RT_SAFE_COORD)) E->boxesFinite = false;
    E->omax = (float)omax;
    E->nNodes = nNodes;
    if (!rtl::build_layout(nodes, nNodes, indices, (size_t)nIndices, shapes, nShapes, subtrees && E->boxesFinite, omax, E->L)) { delete E; return nullptr; }
    E->prims.resize((size_t)E->L.prims());
    for (int s = 0; s < E->L.prims(); ++s) E->prims[(size_t)s] = prim_from_shape(shapes + E->L.pshape[(size_t)s]);
    if (subtrees && E->boxesFinite && topTree) rtl::build_top(E->L, nodes, nNodes);
    fill_nodes(E, nodes);
    return E;
}
void emu_destroy(void* e) { delete static_cast<Emu*>(e); }

// FNV-1a over every array of the layout (lists A and B, slots, boxes, refit graph): a regression handle on the
// host-side builder — the render kernel's behaviour on the GPU was validated against exactly these layouts.
unsigned long long emu_layout_hash(void* e) {
    Emu* E = static_cast<Emu*>(e);
    rtl::Layout& L = E->L;
    rtl::build_refit_graph(L);
    unsigned long long h = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t bytes) { const unsigned char* c = (const unsigned char*)p; for (size_t i = 0; i < bytes; ++i) { h ^= c[i]; h *= 1099511628211ull; } };
    mix(L.src.data(), L.src.size() * 4); mix(L.link.data(), L.link.size() * 4); mix(L.cw.data(), L.cw.size() * 4);
    mix(L.pshape.data(), L.pshape.size() * 4); mix(L.prank.data(), L.prank.size() * 4);
    mix(L.subBox.data(), L.subBox.size() * sizeof(rtl::Box));
    mix(L.srcB.data(), L.srcB.size() * 4); mix(L.linkB.data(), L.linkB.size() * 4); mix(L.cwB.data(), L.cwB.size() * 4);
    mix(L.rfParent.data(), L.rfParent.size() * 4); mix(L.rfChild.data(), L.rfChild.size() * 4);
    mix(L.rfLeaf.data(), L.rfLeaf.size() * 4); mix(L.rfStart.data(), L.rfStart.size() * 4);
    return h;
}

// info: [0] emitted nodes [1] sub nodes [2] prims [3] fat leaves [4] max leaf size [5] nodes of list B (0 = unused)
void emu_info(void* e, int out[6]) {
    out[5] = static_cast<Emu*>(e)->L.useB ? (int)static_cast<Emu*>(e)->L.srcB.size() : 0;
    const Emu* E = static_cast<Emu*>(e);
    out[0] = E->L.nodes(); out[1] = E->L.nSubA; out[2] = E->L.prims(); out[3] = E->L.nRefLeavesFat; out[4] = 0;
    for (int k = 0; k < E->L.nodes(); ++k) if (E->L.cw[(size_t)k] >= 0 && (E->L.cw[(size_t)k] & RT_LEAF_COUNT_MASK) > out[4]) out[4] = E->L.cw[(size_t)k] & RT_LEAF_COUNT_MASK;
}

// The boxes of the fat-leaf sub-tree nodes (conservative bounds, rtl::prim_bound): 6 floats each; returns their number.
int emu_subboxes(void* e, float* out, int cap) {
    const Emu* E = static_cast<Emu*>(e);
    const int n = E->L.nSubA;
    for (int j = 0; j < n && j < cap; ++j) {
        const rtl::Box& b = E->L.subBox[(size_t)j];
        for (int a = 0; a < 3; ++a) { out[6 * j + a] = b.mn[a]; out[6 * j + 3 + a] = b.mx[a]; }
    }
    return n;
}

// Refit after the caller changed shapes in place (same pointer semantics as rt_update_shapes): returns 1 ok, 0 = rebuild needed.
int emu_refit(void* e, const rt_flat_shape* shapes, const rt_flat_node* nodes) {
    Emu* E = static_cast<Emu*>(e);
    if (!rtl::refit_subboxes(E->L, shapes, (double)E->omax)) return 0;
    for (int s = 0; s < E->L.prims(); ++s) E->prims[(size_t)s] = prim_from_shape(shapes + E->L.pshape[(size_t)s]);
    if (E->L.useB) { if (rtl::ref_tree_nested(nodes, E->nNodes)) rtl::refit_top(E->L, nodes); else E->L.useB = false; }
    fill_nodes(E, nodes);
    return 1;
}

// The refit graph (rtl::build_refit_graph, executed by the device's refit_kernel) replayed sequentially with
// the kernel's arrival rule: returns the number of subBox entries whose box differs from the host refit
// (rtl::refit_subboxes / refit_top), or -1 if some entry never got computed.
int emu_refit_graph_check(void* e, const rt_flat_shape* shapes, const rt_flat_node* nodes) {
    Emu* E = static_cast<Emu*>(e);
    rtl::Layout& L = E->L;
    rtl::build_refit_graph(L);
    const int nE = (int)L.subBox.size();
    std::vector<rtl::Box> box((size_t)nE);
    std::vector<int> arrived((size_t)nE, 0), done((size_t)nE, 0);
    auto childBox = [&](int ref) { return ref >= 0 ? box[(size_t)ref] : rtl::raw_box(nodes[-1 - ref]); };
    for (int st : L.rfStart) {
        int j = st;
        if (j >= 0) {
            rtl::Box b; b.reset();
            for (int q = 0; q < L.rfLeaf[2 * (size_t)j + 1]; ++q) {
                rtl::Box pb;
                if (!rtl::prim_bound(shapes[L.pshape[(size_t)(L.rfLeaf[2 * (size_t)j] + q)]], L.omax, pb)) pb.everything();
                b.grow(pb);
            }
            box[(size_t)j] = b; done[(size_t)j] = 1;
            j = L.rfParent[(size_t)j];
        } else j = -1 - j;
        while (j >= 0) {
            if (arrived[(size_t)j]++ == 0) break;
            rtl::Box b; b.reset();
            b.grow(childBox(L.rfChild[2 * (size_t)j])); b.grow(childBox(L.rfChild[2 * (size_t)j + 1]));
            box[(size_t)j] = b; done[(size_t)j] = 1;
            j = L.rfParent[(size_t)j];
        }
    }
    int bad = 0;
    for (int j = 0; j < nE; ++j) {
        if (!done[(size_t)j]) return -1;
        if (std::memcmp(&box[(size_t)j], &L.subBox[(size_t)j], sizeof(rtl::Box)) != 0) bad++;
    }
    return bad;
}

// n rays; anyMode: bounds[i] is the any-hit limit.  counters: [0] nodes visited [1] primitive tests.
void emu_query(void* e, int n, const float* origins, const float* dirs, int mt, int anyMode, const float* bounds,
               int* outShape, float* outDist, float* outPoint, uint64_t counters[2]) {
    const Emu* E = static_cast<Emu*>(e);
    counters[0] = counters[1] = 0;
    for (int i = 0; i < n; ++i) {
        float* P = outPoint ? outPoint + 3 * (size_t)i : nullptr;
        if (mt) query<true>(*E, origins + 3 * (size_t)i, dirs + 3 * (size_t)i, anyMode, anyMode ? bounds[i] : 0.f, outShape + i, outDist + i, P, counters);
        else    query<false>(*E, origins + 3 * (size_t)i, dirs + 3 * (size_t)i, anyMode, anyMode ? bounds[i] : 0.f, outShape + i, outDist + i, P, counters);
    }
}
}

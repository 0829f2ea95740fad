// rt_api.cu — librt_b200.so: the C ABI of include/rt_api.h on top of the
// kernels in rt_kernels.cuh.  Context = what the reference keeps in GL objects
// (program, 5 SSBOs, output texture; src/main.cpp:182-275).  No CPU fallback.
#include "rt_kernels.cuh"
#include "rt_layout.h"
#include "rt_refit.cuh"
#include "rt_anim.cuh"
#include "rt_build.cuh"
#include "rt_build_list.cuh"
#include "../host/rt_anim_math.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <algorithm>
#include <new>
#include <string>
#include <utility>
#include <vector>

using namespace rt;

#ifndef RT_TOP_TREE_RETRY_EVERY
#define RT_TOP_TREE_RETRY_EVERY 64           // a rejected list B is tried again after this many topology changes
#endif

namespace {

template <class T>
struct DevBuf {                       // capacity-based device array
    T* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 4 + 16;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
template <class T>
struct PinBuf {                       // capacity-based pinned host array (keeps contents on growth)
    T* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t n, size_t keep) {
        if (n <= cap) return cudaSuccess;
        size_t want = n + n / 4 + 16;
        T* q = nullptr;
        cudaError_t e = cudaMallocHost(&q, want * sizeof(T));
        if (e != cudaSuccess) return e;
        if (p && keep) std::memcpy(q, p, keep * sizeof(T));
        if (p) cudaFreeHost(p);
        p = q; cap = want;
        return cudaSuccess;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

}  // namespace

#ifndef RT_MAX_BANDS
#define RT_MAX_BANDS 8
#endif
#ifndef RT_MAX_BANDS_RGBA8
#define RT_MAX_BANDS_RGBA8 3
#endif
#ifndef RT_BAND_ROWS
#define RT_BAND_ROWS 16         // at least this many tile rows per band
#endif
#ifndef RT_DEVICE_LAYOUT_MAX_LEAF
#define RT_DEVICE_LAYOUT_MAX_LEAF 16
#endif

struct rt_ctx {
    int device = 0, W = 0, H = 0;
    unsigned flags = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t ownStream = nullptr;
    cudaEvent_t ev[8] = {};          // h2d0,h2d1, repack0,repack1, render0,render1, d2h0,d2h1
    cudaEvent_t copyDone = nullptr;  // last H2D out of the pinned shadows has finished
    cudaStream_t stream2 = nullptr, copyStream = nullptr;   // banded rt_render: second render stream, D2H stream
    cudaEvent_t evBand[RT_MAX_BANDS + 3] = {};
    bool copyPending = false;
    bool timed[4] = {false, false, false, false};
    std::string err;
    int64_t launches = 0;
    uint64_t lastTally[RT_NTALLY] = {};
    int smCount = 148;
    int fewPerSm = RT_FEW_BLOCKS_PER_SM;   // resident render blocks per SM for launches with few items per warp (0: as many as fit)

    // raw scene, pinned shadow + device copy
    PinBuf<rt_flat_shape> hShapes; DevBuf<rt_flat_shape> dShapes; int nShapes = 0;
    PinBuf<rt_flat_node>  hNodes;  DevBuf<rt_flat_node>  dNodes;  int nNodes = 0;
    std::vector<int32_t> indices;                                 // host only (feeds the visit-order map)
    rt_flat_light light{}; rt_flat_camera cam{};
    bool haveScene = false, haveCam = false, haveLight = false;

    // visit-order maps (host-built, pinned) and the repacked device scene
    PinBuf<int>  hOrder;   DevBuf<int>  dOrder;
    PinBuf<int2> hLinkCnt; DevBuf<int2> dLinkCnt;
    PinBuf<int>  hPshape;  DevBuf<int>  dPshape;
    PinBuf<int>  hPrank;   DevBuf<int>  dPrank;
    PinBuf<int>  hOrderB;   DevBuf<int>  dOrderB;      // list B (own tree over the reference leaves)
    PinBuf<int2> hLinkCntB; DevBuf<int2> dLinkCntB;
    DevBuf<float4> dNodesB;
    int nDfsB = 0;
    PinBuf<float4> hSubBox; DevBuf<float4> dSubBox;   // 2 float4 per sub-tree node (host-computed conservative boxes)
    rtl::Layout layout;               // host copy of the threaded layout
    // refit graph of the boxes that are ours (rt_refit.cuh), uploaded with the layout
    PinBuf<int> hRf; DevBuf<int> dRf;           // [child 2nE | leaf 2nE | parent nE | start nStart]
    DevBuf<unsigned> dRfArrived;                // nE counters + 1 "degraded" flag
    int nRfStart = 0;
    // Flag of the device refit (a primitive could not be bounded any more), in MAPPED pinned memory: K4 stores into it directly.
    // (It used to be a 4-byte D2H copy in the render stream: that copy queued on the D2H engine behind the previous frame's
    // 7.7 MB copy on the copy stream and with it the whole next frame did: the pipelined monkey loop ran at 0.36 instead of 0.27 ms.)
    int* hRfDegraded = nullptr; int* dRfDegraded = nullptr;
    // device-side animation (rt_anim.cuh)
    struct AnimDef { int kind; float originY, amp, freq, cx, cy, cz, speed; };
    std::vector<AnimDef> animDefs; std::vector<AnimItem> animItems;
    PinBuf<AnimEntry> hAnimEntries; DevBuf<AnimEntry> dAnimEntries; DevBuf<AnimItem> dAnimItems;
    DevBuf<unsigned char> dAnimFlag; DevBuf<int> dIndices, dNodeAux; DevBuf<unsigned> dGrowArrived;   // dNodeAux: [parent nNodes | leaves nLeaves]
    int nLeavesRaw = 0;
    DevBuf<int> dGrowSched; DevBuf<int4> dGrowEnt; int growLevels = 0;   // the animated leaves and their ancestors, level by level (anim_grow_path_kernel); 0: grow_kernel
    // device-side BVH build (rt_build.cuh, K8): scratch tables and the emitted FlatNode[] / bvhIndices[]
    DevBuf<int> bInt; DevBuf<bld::u64> bKey; DevBuf<unsigned char> bFlag; DevBuf<bld::Ctl> bCtl;
    DevBuf<rt_flat_node> bNodes; DevBuf<int32_t> bIdx; DevBuf<float4> bRec;
    PinBuf<rt_flat_node> bHostNodes; PinBuf<int32_t> bHostIdx; int bNodeCount = 0, bIdxCount = 0;   // pinned: the tree comes back at PCIe speed
    bool nodesNested = false;         // the nodes of the last set_nodes: parents around their children (rtl::scan_nodes)
    bool h2dStarted = false;          // rt_upload_scene has recorded ev[0] in front of its early shape copy
    bool shapesOnDevice = false;      // the pending host range [shapeLo, shapeHi) has already been copied (rt_build_bvh)
    float lastBuildMs = 0.f;
    cudaEvent_t evBuild[2] = {nullptr, nullptr};
    bool animDirty = true;            // registrations or topology changed: item list / flags / tree maps must go up again
    bool animItemsDirty = true;       // registrations changed: the item list alone (rt_animate_shapes needs no tree maps)
    bool devShapesDirty = false, devNodesDirty = false;   // the device copies were changed by K6 / K7: repack without an upload
    bool hostStale = false;           // the host shadows no longer describe the device scene (shapes and / or node boxes)
    bool hostShapesStale = false, hostNodesStale = false;
    // K9 laid the tree out on the device: the host's node shadow still holds an OLDER tree (boxes AND topology) and the
    // host's index list is stale too; both are fetched on demand (ensure_host_tree)
    bool hostTreeStale = false;
    bool layoutOnDevice = false;
    float rootBox[6] = {0, 0, 0, 0, 0, 0};   // the root's box of the current tree (omax)
    PinBuf<bld::Ctl> hCtl;
    DevBuf<int> kOrder, kPshape, kPrank; DevBuf<int2> kLinkCnt;   // K9 writes here; adopted by a swap when the list can be used
    cudaEvent_t rfDone = nullptr; bool rfPending = false;
    int nSub = 0;
    bool subDirty = false;            // sub-tree boxes must be recomputed (shapes moved / omax grew)
    float omax = 0.f;
    DevBuf<float4> dNodesT, q0, q1, q2, q3;
    DevBuf<int> ptype;
    int nDfs = 0, nPrims = 0;

    // pending work, flushed at render time so per-shape updates coalesce (main.cpp:981-992)
    int shapeLo = 0, shapeHi = 0;     // dirty raw-shape range [lo,hi)
    // the same updates as disjoint intervals: while the shadow is stale (rt_animate moved shapes on the device) only what
    // the host actually rewrote may go up — the gap between two updates holds pre-animation values
    std::vector<std::pair<int, int>> shapeRanges;
    bool nodesDirty = false, topoDirty = false;
    bool boxesFinite = false;

    // output
    DevBuf<float4> image; DevBuf<int> ids;
    // rt_render_async pipelines frames: two device images alternate, so that the D2H of frame i (copy stream)
    // overlaps the render of frame i+1; evFrame[k] = the copy out of image k has finished
    DevBuf<float4> image2; DevBuf<int> ids2;
    DevBuf<uchar4> image8[2];         // rt_render_rgba8: the display quad's RGBA8 frame (K5 behind K1), one per pipelined image
    cudaEvent_t evFrame[2] = {nullptr, nullptr};
    bool framePending[2] = {false, false};
    int imgPar = 0;                   // image the next pipelined frame renders into
    // The D2H copy of a pipelined single-launch frame is DEFERRED until the next frame's render kernel is about to be launched
    // (or until somebody waits for the frame): while a 7.7 MB DMA to the host is in flight every dependent kernel launch of the
    // stream takes ~8 us longer on this box (measured with six empty kernels per frame: 0.192 -> 0.239 ms), and an animated frame
    // begins with five to seven small launches (K6, K7, K4, K2 ...): the copy now runs beside the long render kernel only.
    struct DeferredCopy { bool valid = false; float* hostRgba = nullptr; const float* devRgba = nullptr; int32_t* hostIds = nullptr;
                          const int32_t* devIds = nullptr; uint8_t* host8 = nullptr; const uchar4* dev8 = nullptr; size_t pixels = 0; int par = 0; } deferred;
    cudaEvent_t evDefer = nullptr;
    float4* lastImage = nullptr;
    DevBuf<float4> frame;             // full W*H frame other ranks write into (rt_frame_export)
    float* frameTarget = nullptr;     // where this rank's tiles go row-major (local frame, peer mapping, or caller's)
    void* ipcOpened = nullptr;        // mapping obtained from cudaIpcOpenMemHandle
    std::vector<void*> hostFrames;    // host buffers pinned by rt_host_frame_register (undone by rt_destroy if the caller did not)
    // signalled fused gather (rt_api.h): double-buffered frame + completion words on rank 0, one release word per other rank
    bool syncOn = false;
    unsigned frameSeq = 0;            // frames rendered in signalled mode (frame f -> buffer f & 1)
    DevBuf<unsigned> releaseLocal, ctaDone;
    unsigned* peerRelease[64] = {};   // rank 0: the other ranks' release words (same-process pointers or IPC mappings)
    void* peerIpc[64] = {};
    DevBuf<unsigned*> dPeerRelease;
    int (*waitValue32)(cudaStream_t, unsigned long long, unsigned, unsigned) = nullptr;   // cuStreamWaitValue32 (driver entry point)
    DevBuf<unsigned long long> tally;
    DevBuf<unsigned int> queue;       // work-stealing counter of K1
    DevBuf<unsigned> itemCost, heavyList, lptCtl;   // heavy-first scheduling from the previous frame's item costs
    DevBuf<unsigned char> heavyFlag;
    int lptItems = -1;                // item count the recorded costs belong to (-1: none yet)
    unsigned lptFrame = 0;            // frames rendered with these costs (parity = slot of the control block)
    int gridBlocks[6] = {0, 0, 0, 0, 0, 0}; // persistent grid per kernel variant
    bool noCluster = false;           // a cluster launch of the builder failed on this device once
    bool noCooperative = false;       // ... a cooperative launch
    int bldGridBlocks = 0;            // blocks of bld_grid_kernel that are resident at once
    int topSkipsLeft = 0, topRejectedNodes = 0;   // list B was not worth using for a tree of that many nodes: not tried again for a while
    int shardRank = 0, shardWorld = 1;
};

namespace {

int fail(rt_ctx* c, int code, const char* what, cudaError_t e = cudaSuccess) {
    if (c) {
        c->err = what;
        if (e != cudaSuccess) { c->err += ": "; c->err += cudaGetErrorString(e); }
    }
    return code;
}
#define CU(c, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail((c), RT_ERR_CUDA, #call, e_); } while (0)

inline int tiles_x(const rt_ctx* c) { return (c->W + RT_TILE_W - 1) / RT_TILE_W; }
inline int tiles_y(const rt_ctx* c) { return (c->H + RT_TILE_H - 1) / RT_TILE_H; }
inline int tiles_mine(const rt_ctx* c) {
    int n = tiles_x(c) * tiles_y(c);
    return n > c->shardRank ? (n - c->shardRank + c->shardWorld - 1) / c->shardWorld : 0;
}
// Shards are padded to rank 0's tile count so that every rank gathers the same size.
inline int64_t out_pixels(const rt_ctx* c) {
    if (c->shardWorld == 1) return (int64_t)c->W * c->H;
    const int nT = tiles_x(c) * tiles_y(c);
    return (int64_t)((nT + c->shardWorld - 1) / c->shardWorld) * RT_TILE_W * RT_TILE_H;
}
// The pinned shadows are the source of async copies: wait for the last one before
// the host overwrites them (lets the next frame's update overlap the previous render).
inline void wait_copies(rt_ctx* c) {
    if (c->copyPending) { cudaEventSynchronize(c->copyDone); c->copyPending = false; }
}
// Issue the deferred D2H copy of the last pipelined frame (rt_ctx::deferred).  The copy stream already waits for that frame's
// render kernel; `afterStream` makes it wait for what the context stream holds so far as well (the small launches that open
// the next frame), the callers that are about to wait for the frame pass false.
int flush_deferred(rt_ctx* c, bool afterStream) {
    if (!c->deferred.valid) return RT_OK;
    rt_ctx::DeferredCopy& d = c->deferred;
    d.valid = false;
    if (afterStream) { cudaEventRecord(c->evDefer, c->stream); cudaStreamWaitEvent(c->copyStream, c->evDefer, 0); }
    cudaEventRecord(c->ev[6], c->copyStream);
    if (d.hostRgba) CU(c, cudaMemcpyAsync(d.hostRgba, d.devRgba, d.pixels * sizeof(float4), cudaMemcpyDeviceToHost, c->copyStream));
    if (d.hostIds) CU(c, cudaMemcpyAsync(d.hostIds, d.devIds, d.pixels * sizeof(int), cudaMemcpyDeviceToHost, c->copyStream));
    if (d.host8) CU(c, cudaMemcpyAsync(d.host8, d.dev8, d.pixels * sizeof(uchar4), cudaMemcpyDeviceToHost, c->copyStream));
    cudaEventRecord(c->ev[7], c->copyStream); c->timed[3] = true;
    cudaEventRecord(c->evFrame[d.par], c->copyStream);
    return RT_OK;
}

// Build the threaded layout (rt_layout.h): reference nodes in the reference's visit order (root =
// last node, right child first, gpu_shader.comp:386, :424-425) with fat leaves expanded into
// conservative sub-trees.  Validates what the shader would trust.
double pick_omax(const rt_ctx* c) {
    double m = 1.0;
    if (c->nNodes > 0) {
        for (int k = 0; k < 6; ++k) {
            const double a = std::fabs((double)c->rootBox[k]);
            if (std::isfinite(a) && a > m) m = a;
        }
    }
    if (c->haveCam) for (int k = 0; k < 3; ++k) { const double a = std::fabs((double)c->cam.Position[k]); if (std::isfinite(a) && a > m) m = a; }
    if (c->haveLight) for (int k = 0; k < 3; ++k) { const double a = std::fabs((double)c->light.position[k]); if (std::isfinite(a) && a > m) m = a; }
    return 4.0 * m;
}

// After a device-side layout (K9) the host never saw the tree: bring FlatNode[] / bvhIndices[] into the host shadows for
// the callers that need them there (a host layout build, the animation tables, rt_get_bvh, topology comparisons).
int ensure_host_tree(rt_ctx* c) {
    if (!c->hostTreeStale) return RT_OK;
    CU(c, cudaStreamSynchronize(c->stream));
    wait_copies(c);
    CU(c, c->hNodes.reserve((size_t)c->nNodes, 0));
    if (c->nNodes > 0) CU(c, cudaMemcpy(c->hNodes.p, c->dNodes.p, (size_t)c->nNodes * sizeof(rt_flat_node), cudaMemcpyDeviceToHost));
    c->indices.resize((size_t)c->bIdxCount);
    if (c->bIdxCount > 0) CU(c, cudaMemcpy(c->indices.data(), c->bIdx.p, (size_t)c->bIdxCount * sizeof(int32_t), cudaMemcpyDeviceToHost));
    c->hostTreeStale = false; c->hostNodesStale = false;
    c->hostStale = c->hostShapesStale;
    return RT_OK;
}

void merge_ranges(std::vector<std::pair<int, int>>& v) {
    if (v.size() < 2) return;
    std::sort(v.begin(), v.end());
    size_t w = 0;
    for (size_t i = 1; i < v.size(); ++i) {
        if (v[i].first <= v[w].second) { if (v[i].second > v[w].second) v[w].second = v[i].second; }
        else v[++w] = v[i];
    }
    v.resize(w + 1);
}

void stage_subboxes(rt_ctx* c) {
    const rtl::Layout& L = c->layout;
    for (size_t j = 0; j < L.subBox.size(); ++j) {
        const rtl::Box& b = L.subBox[j];
        c->hSubBox.p[2 * j]     = make_float4(b.mn[0], b.mn[1], b.mn[2], 0.f);
        c->hSubBox.p[2 * j + 1] = make_float4(b.mx[0], b.mx[1], b.mx[2], 0.f);
    }
}

int build_visit_order(rt_ctx* c) {
    rtl::Layout& L = c->layout;
    const bool subtrees = c->boxesFinite && !(c->flags & RT_FLAG_REFERENCE_WALK);
    c->omax = (float)pick_omax(c);
    if (!rtl::build_layout(c->hNodes.p, c->nNodes, c->indices.data(), c->indices.size(), c->hShapes.p, c->nShapes,
                           subtrees, (double)c->omax, L))
        return fail(c, RT_ERR_SCENE, L.error ? L.error : "malformed BVH");
    // List B's tree is tried again only now and then for a context whose last tree of about this size was not worth using
    // (monkey: 0.5 ms of host time per topology change for a list that the SAH estimate rejects every time).  Building it on a
    // second thread beside build_layout was measured too (std::async; the tree depends on the reference leaves only): no gain on
    // the bench box (car cold upload 2.06 -> 2.22 ms), not kept.
    bool wantTop = subtrees && !(c->flags & RT_FLAG_NO_TOP_TREE);
    if (wantTop && c->topSkipsLeft > 0 && std::abs(c->nNodes - c->topRejectedNodes) * 8 <= c->topRejectedNodes) { --c->topSkipsLeft; wantTop = false; }
    if (wantTop) {
        rtl::build_top(L, c->hNodes.p, c->nNodes);
        if (L.useB) c->topSkipsLeft = 0;
        else { c->topRejectedNodes = c->nNodes; c->topSkipsLeft = RT_TOP_TREE_RETRY_EVERY - 1; }
    }
    const int m = L.nodes(), prims = L.prims();
    c->nSub = (int)L.subBox.size();
    c->nDfsB = L.useB ? (int)L.srcB.size() : 0;
    CU(c, c->hOrderB.reserve((size_t)c->nDfsB, 0));
    CU(c, c->hLinkCntB.reserve((size_t)c->nDfsB, 0));
    for (int k = 0; k < c->nDfsB; ++k) { c->hOrderB.p[k] = L.srcB[(size_t)k]; c->hLinkCntB.p[k] = make_int2(L.linkB[(size_t)k], L.cwB[(size_t)k]); }
    CU(c, c->hOrder.reserve((size_t)m, 0));
    CU(c, c->hLinkCnt.reserve((size_t)m, 0));
    CU(c, c->hPshape.reserve((size_t)prims, 0));
    CU(c, c->hPrank.reserve((size_t)prims, 0));
    CU(c, c->hSubBox.reserve((size_t)c->nSub * 2, 0));
    for (int k = 0; k < m; ++k) { c->hOrder.p[k] = L.src[(size_t)k]; c->hLinkCnt.p[k] = make_int2(L.link[(size_t)k], L.cw[(size_t)k]); }
    if (prims > 0) {
        std::memcpy(c->hPshape.p, L.pshape.data(), (size_t)prims * sizeof(int));
        std::memcpy(c->hPrank.p, L.prank.data(), (size_t)prims * sizeof(int));
    }
    stage_subboxes(c);
    rtl::build_refit_graph(L);
    c->nRfStart = (int)L.rfStart.size();
    CU(c, c->hRf.reserve((size_t)c->nSub * 5 + (size_t)c->nRfStart, 0));
    if (c->nSub > 0) {
        std::memcpy(c->hRf.p, L.rfChild.data(), (size_t)c->nSub * 2 * sizeof(int));
        std::memcpy(c->hRf.p + 2 * (size_t)c->nSub, L.rfLeaf.data(), (size_t)c->nSub * 2 * sizeof(int));
        std::memcpy(c->hRf.p + 4 * (size_t)c->nSub, L.rfParent.data(), (size_t)c->nSub * sizeof(int));
        if (c->nRfStart > 0) std::memcpy(c->hRf.p + 5 * (size_t)c->nSub, L.rfStart.data(), (size_t)c->nRfStart * sizeof(int));
    }
    c->nDfs = m; c->nPrims = prims;
    c->subDirty = false;
    return RT_OK;
}

// Bring the device scene up to date: coalesced H2D of whatever changed, then K2.
int flush_scene(rt_ctx* c) {
    // a refit on the device of an earlier frame met a primitive it could not bound any more: rebuild the layout
    if (c->rfPending && cudaEventQuery(c->rfDone) == cudaSuccess) {
        c->rfPending = false;
        if (*c->hRfDegraded) { *c->hRfDegraded = 0; c->topoDirty = true; }
    }
    // the camera may have left the region the sub-tree margins were computed for
    if (c->nSub > 0 && !c->topoDirty && c->haveCam) {
        double m = 0;
        for (int k = 0; k < 3; ++k) m = std::fmax(m, std::fabs((double)c->cam.Position[k]));
        if (!(m <= 0.5 * (double)c->omax)) { c->omax = (float)pick_omax(c); c->subDirty = true; }
    }
    const bool anyCopy = c->shapeHi > c->shapeLo || c->nodesDirty || c->topoDirty || c->subDirty || c->devShapesDirty || c->devNodesDirty;
    if (!anyCopy) return RT_OK;
    if (c->topoDirty && c->hostTreeStale) { int rc = ensure_host_tree(c); if (rc != RT_OK) return rc; }
    if (c->topoDirty && c->hostStale) {
        // the layout is rebuilt from the host shadows: bring them up to date with what the device animated
        CU(c, cudaStreamSynchronize(c->stream));
        wait_copies(c);
        if (c->hostShapesStale && c->nShapes > 0) {        // everything but the intervals the host has just rewritten
            merge_ranges(c->shapeRanges);
            int at = 0;
            for (size_t i = 0; i <= c->shapeRanges.size(); ++i) {
                const int lo = i < c->shapeRanges.size() ? c->shapeRanges[i].first : c->nShapes;
                if (lo > at) CU(c, cudaMemcpy(c->hShapes.p + at, c->dShapes.p + at, (size_t)(lo - at) * sizeof(rt_flat_shape), cudaMemcpyDeviceToHost));
                if (i < c->shapeRanges.size()) at = c->shapeRanges[i].second;
            }
            c->shapeRanges.clear();                        // the shadow is whole again: one coalesced copy below is right
        }
        if (c->hostNodesStale && !c->nodesDirty && c->nNodes > 0)
            CU(c, cudaMemcpy(c->hNodes.p, c->dNodes.p, (size_t)c->nNodes * sizeof(rt_flat_node), cudaMemcpyDeviceToHost));
        c->hostStale = c->hostShapesStale = c->hostNodesStale = false;
    }
    const bool shapesChanged = c->shapeHi > c->shapeLo;
    if ((shapesChanged || c->devShapesDirty) && c->nSub > 0) c->subDirty = true;
    if (!c->h2dStarted) cudaEventRecord(c->ev[0], c->stream);
    c->h2dStarted = false;
    if (shapesChanged && !c->shapesOnDevice) {
        CU(c, c->dShapes.reserve((size_t)c->nShapes));
        if (c->hostShapesStale && !c->shapeRanges.empty()) {
            // the shadow between the updated intervals is older than the device copy (K6 animated it): one copy per interval
            merge_ranges(c->shapeRanges);
            for (const auto& r : c->shapeRanges)
                CU(c, cudaMemcpyAsync(c->dShapes.p + r.first, c->hShapes.p + r.first,
                                      (size_t)(r.second - r.first) * sizeof(rt_flat_shape), cudaMemcpyHostToDevice, c->stream));
        } else {
            CU(c, cudaMemcpyAsync(c->dShapes.p + c->shapeLo, c->hShapes.p + c->shapeLo,
                                  (size_t)(c->shapeHi - c->shapeLo) * sizeof(rt_flat_shape), cudaMemcpyHostToDevice, c->stream));
        }
    }
    if (c->nodesDirty && c->nNodes > 0) {
        CU(c, c->dNodes.reserve((size_t)c->nNodes));
        CU(c, cudaMemcpyAsync(c->dNodes.p, c->hNodes.p, (size_t)c->nNodes * sizeof(rt_flat_node), cudaMemcpyHostToDevice, c->stream));
    }
    bool subUpload = false, devRefit = false;
    if (c->nodesDirty && !c->topoDirty && c->nDfsB > 0) {
        // reference boxes changed under an unchanged topology (grow-only refit): B's union boxes follow;
        // if the reference tree stopped being nested, B is no longer equivalent and is switched off
        if (c->nodesNested) devRefit = true;     // (set_nodes' scan of exactly these nodes)
        else { c->nDfsB = 0; c->layout.useB = false; }
    }
    if (c->devNodesDirty && !c->topoDirty && c->nDfsB > 0) devRefit = true;   // K7 keeps parents around their children
    if (c->subDirty && !c->topoDirty) {
        // shapes moved under a sub-tree (or omax grew): its boxes are refitted on the device (K4)
        devRefit = true;
        c->subDirty = false;
    }
    if (c->topoDirty) {
        wait_copies(c);
        c->layoutOnDevice = false;
        int rc = build_visit_order(c);
        if (rc != RT_OK) return rc;
        CU(c, c->dOrder.reserve((size_t)c->nDfs)); CU(c, c->dLinkCnt.reserve((size_t)c->nDfs));
        CU(c, c->dPshape.reserve((size_t)c->nPrims)); CU(c, c->dPrank.reserve((size_t)c->nPrims));
        CU(c, c->dNodesT.reserve((size_t)c->nDfs * 2));
        // +4 slots of padding: the batched triangle loop loads q0[s..s+3] before masking by the leaf end
        CU(c, c->q0.reserve((size_t)c->nPrims + 4)); CU(c, c->q1.reserve((size_t)c->nPrims + 4));
        CU(c, c->q2.reserve((size_t)c->nPrims + 4)); CU(c, c->q3.reserve((size_t)c->nPrims + 4));
        CU(c, c->ptype.reserve((size_t)c->nPrims));
        if (c->nDfs > 0) {
            CU(c, cudaMemcpyAsync(c->dOrder.p, c->hOrder.p, (size_t)c->nDfs * sizeof(int), cudaMemcpyHostToDevice, c->stream));
            CU(c, cudaMemcpyAsync(c->dLinkCnt.p, c->hLinkCnt.p, (size_t)c->nDfs * sizeof(int2), cudaMemcpyHostToDevice, c->stream));
        }
        if (c->nDfsB > 0) {
            CU(c, c->dOrderB.reserve((size_t)c->nDfsB)); CU(c, c->dLinkCntB.reserve((size_t)c->nDfsB)); CU(c, c->dNodesB.reserve((size_t)c->nDfsB * 2));
            CU(c, cudaMemcpyAsync(c->dOrderB.p, c->hOrderB.p, (size_t)c->nDfsB * sizeof(int), cudaMemcpyHostToDevice, c->stream));
            CU(c, cudaMemcpyAsync(c->dLinkCntB.p, c->hLinkCntB.p, (size_t)c->nDfsB * sizeof(int2), cudaMemcpyHostToDevice, c->stream));
        }
        if (c->nPrims > 0) {
            CU(c, cudaMemcpyAsync(c->dPshape.p, c->hPshape.p, (size_t)c->nPrims * sizeof(int), cudaMemcpyHostToDevice, c->stream));
            CU(c, cudaMemcpyAsync(c->dPrank.p, c->hPrank.p, (size_t)c->nPrims * sizeof(int), cudaMemcpyHostToDevice, c->stream));
        }
        if (c->nSub > 0) {
            CU(c, c->dRf.reserve((size_t)c->nSub * 5 + (size_t)c->nRfStart));
            CU(c, c->dRfArrived.reserve((size_t)c->nSub + 1));
            CU(c, cudaMemsetAsync(c->dRfArrived.p, 0, ((size_t)c->nSub + 1) * sizeof(unsigned), c->stream));   // once per topology: K4 leaves them zero
            CU(c, cudaMemcpyAsync(c->dRf.p, c->hRf.p, ((size_t)c->nSub * 5 + (size_t)c->nRfStart) * sizeof(int), cudaMemcpyHostToDevice, c->stream));
        }
        subUpload = true; devRefit = false;       // the layout build computed every box on the host
        c->animDirty = true;                      // tree maps of the device animation follow the new topology
    }
    if (subUpload && c->nSub > 0) {
        CU(c, c->dSubBox.reserve((size_t)c->nSub * 2));
        CU(c, cudaMemcpyAsync(c->dSubBox.p, c->hSubBox.p, (size_t)c->nSub * 2 * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    }
    cudaEventRecord(c->ev[1], c->stream); c->timed[0] = true;
    cudaEventRecord(c->copyDone, c->stream); c->copyPending = true;
    cudaEventRecord(c->ev[2], c->stream);
    if (devRefit && c->nSub > 0 && c->nRfStart > 0) {
        const int nE = c->nSub;
        refit_kernel<<<(c->nRfStart + 127) / 128, 128, 0, c->stream>>>(c->dShapes.p, c->dNodes.p, c->dPshape.p,
            c->dRf.p + 5 * (size_t)nE, c->nRfStart, c->dRf.p + 4 * (size_t)nE, reinterpret_cast<const int2*>(c->dRf.p),
            reinterpret_cast<const int2*>(c->dRf.p + 2 * (size_t)nE), c->dRfArrived.p, c->dSubBox.p, (double)c->omax,
            c->dRfDegraded);
        c->launches++;
        if (!c->rfPending) { cudaEventRecord(c->rfDone, c->stream); c->rfPending = true; }
        subUpload = true;                          // the node lists take the new boxes below
    }
    {   // K2: both node lists and the primitive slots in one launch
        const bool doNodes = (c->nodesDirty || c->topoDirty || subUpload || c->devNodesDirty) && c->nDfs > 0;
        const bool doPrims = (shapesChanged || c->topoDirty || c->devShapesDirty) && c->nPrims > 0;
        RepackAll R;
        R.raw = c->dNodes.p; R.subBox = c->dSubBox.p;
        R.orderA = c->dOrder.p; R.linkA = c->dLinkCnt.p; R.outA = c->dNodesT.p; R.nA = doNodes ? c->nDfs : 0;
        R.orderB = c->dOrderB.p; R.linkB = c->dLinkCntB.p; R.outB = c->dNodesB.p; R.nB = doNodes ? c->nDfsB : 0;
        R.shapes = c->dShapes.p; R.pshape = c->dPshape.p; R.q0 = c->q0.p; R.q1 = c->q1.p; R.q2 = c->q2.p; R.q3 = c->q3.p; R.ptype = c->ptype.p;
        R.nPrims = doPrims ? c->nPrims : 0;
        R.blocksA = (R.nA + 255) / 256; R.blocksB = (R.nB + 255) / 256;
        const long long blocks = (long long)R.blocksA + R.blocksB + (R.nPrims + 255) / 256;
        if (blocks > 0) {
            repack_all_kernel<<<(unsigned)blocks, 256, 0, c->stream>>>(R);
            c->launches++;
        }
    }
    cudaEventRecord(c->ev[3], c->stream); c->timed[1] = true;
    CU(c, cudaGetLastError());
    c->shapeLo = c->shapeHi = 0; c->nodesDirty = c->topoDirty = false;
    c->shapeRanges.clear();
    c->devShapesDirty = c->devNodesDirty = false;
    c->shapesOnDevice = false;
    return RT_OK;
}


// adopt != NULL: `nodes` is adopt->p, a pinned buffer of this context (the device builder's staging): it BECOMES the
// shadow by a pointer swap instead of being copied into it.
int set_nodes(rt_ctx* c, const void* nodes, int nNodes, PinBuf<rt_flat_node>* adopt = nullptr) {
    const rt_flat_node* in = static_cast<const rt_flat_node*>(nodes);
    // (after a device-side layout the shadow holds an older tree: nothing to compare with)
    const bool comparable = c->haveScene && !c->hostTreeStale && nNodes == c->nNodes && !c->topoDirty;
    c->hostTreeStale = false; c->layoutOnDevice = false;
    wait_copies(c);
    // one read-only pass: the topology compare with the shadow, |bound| <= 1e37 everywhere (the NaN-free slab fast path) and
    // parents around their children (list B stays valid under a refit; flush_scene); then the copy into the shadow
    rtl::NodeScan sc;
    sc = rtl::scan_nodes(in, comparable ? c->hNodes.p : nullptr, nNodes, RT_SAFE_COORD);
    if (adopt) {
        std::swap(c->hNodes, *adopt);
    } else {
        CU(c, c->hNodes.reserve((size_t)nNodes, 0));
        if (nNodes > 0) std::memcpy(c->hNodes.p, in, (size_t)nNodes * sizeof(rt_flat_node));
    }
    const bool sameTopo = comparable && sc.sameTopo;
    c->nNodes = nNodes;
    if (nNodes > 0) for (int k = 0; k < 3; ++k) { c->rootBox[k] = in[nNodes - 1].boundsMin[k]; c->rootBox[3 + k] = in[nNodes - 1].boundsMax[k]; }
    c->boxesFinite = sc.safe;
    c->nodesNested = sc.nested;
    c->nodesDirty = true;
    c->hostNodesStale = false;
    if (!sameTopo) c->topoDirty = true;
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_create(rt_ctx** out, int width, int height, int device, unsigned flags) {
    if (!out || width <= 0 || height <= 0 || device < 0) return RT_ERR_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device >= count) return RT_ERR_CUDA;   // no CPU fallback
    rt_ctx* c = new (std::nothrow) rt_ctx();
    if (!c) return RT_ERR_NOMEM;
    c->device = device; c->W = width; c->H = height; c->flags = flags;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return RT_ERR_CUDA; }
    for (auto& e : c->ev) if (cudaEventCreate(&e) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    if (cudaEventCreateWithFlags(&c->copyDone, cudaEventDisableTiming) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    if (cudaEventCreateWithFlags(&c->rfDone, cudaEventDisableTiming) != cudaSuccess ||
        cudaHostAlloc(&c->hRfDegraded, sizeof(int), cudaHostAllocMapped) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    *c->hRfDegraded = 0;
    if (cudaHostGetDevicePointer(&c->dRfDegraded, c->hRfDegraded, 0) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    cudaDeviceGetAttribute(&c->smCount, cudaDevAttrMultiProcessorCount, device);
    if (const char* e = std::getenv("RT_B200_FEW_BLOCKS_PER_SM")) c->fewPerSm = std::atoi(e);     // kernel-variant experiments (tools/kbench.py)
    c->ownStream = c->stream;
    if (cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    for (auto& e : c->evBand) if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    for (auto& e : c->evFrame) if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    if (cudaEventCreateWithFlags(&c->evDefer, cudaEventDisableTiming) != cudaSuccess) { rt_destroy(c); return RT_ERR_CUDA; }
    *out = c;
    return RT_OK;
}

void rt_destroy(rt_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    c->hShapes.release(); c->dShapes.release(); c->hNodes.release(); c->dNodes.release();
    c->hOrder.release(); c->dOrder.release(); c->hLinkCnt.release(); c->dLinkCnt.release();
    c->hPshape.release(); c->dPshape.release(); c->hPrank.release(); c->dPrank.release();
    c->hSubBox.release(); c->dSubBox.release(); c->dNodesT.release();
    c->hOrderB.release(); c->dOrderB.release(); c->hLinkCntB.release(); c->dLinkCntB.release(); c->dNodesB.release();
    c->q0.release(); c->q1.release(); c->q2.release(); c->q3.release(); c->ptype.release();
    for (void* h : c->hostFrames) cudaHostUnregister(h);     // a registered host frame the caller never unregistered
    c->hostFrames.clear();
    if (c->ipcOpened) cudaIpcCloseMemHandle(c->ipcOpened);
    for (auto& p : c->peerIpc) if (p) cudaIpcCloseMemHandle(p);
    c->releaseLocal.release(); c->ctaDone.release(); c->dPeerRelease.release();
    c->frame.release();
    c->itemCost.release(); c->heavyList.release(); c->lptCtl.release(); c->heavyFlag.release();
    c->image.release(); c->ids.release(); c->tally.release(); c->queue.release();
    for (auto& e : c->ev) if (e) cudaEventDestroy(e);
    if (c->copyDone) cudaEventDestroy(c->copyDone);
    if (c->rfDone) cudaEventDestroy(c->rfDone);
    if (c->hRfDegraded) cudaFreeHost(c->hRfDegraded);
    c->hRf.release(); c->dRf.release(); c->dRfArrived.release(); c->hCtl.release();
    c->kOrder.release(); c->kLinkCnt.release(); c->kPshape.release(); c->kPrank.release();
    c->hAnimEntries.release(); c->dAnimEntries.release(); c->dAnimItems.release(); c->dAnimFlag.release();
    c->dIndices.release(); c->dNodeAux.release(); c->dGrowArrived.release(); c->dGrowSched.release(); c->dGrowEnt.release();
    for (auto& e : c->evBuild) if (e) cudaEventDestroy(e);
    c->bInt.release(); c->bKey.release(); c->bFlag.release(); c->bCtl.release(); c->bNodes.release(); c->bIdx.release(); c->bRec.release(); c->bHostNodes.release(); c->bHostIdx.release();
    for (auto& e : c->evBand) if (e) cudaEventDestroy(e);
    for (auto& e : c->evFrame) if (e) cudaEventDestroy(e);
    if (c->evDefer) cudaEventDestroy(c->evDefer);
    c->image2.release(); c->ids2.release(); c->image8[0].release(); c->image8[1].release();
    if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); }
    if (c->copyStream) { cudaStreamSynchronize(c->copyStream); cudaStreamDestroy(c->copyStream); }
    if (!c->ownStream) c->ownStream = c->stream;      // rt_create failed half-way
    if (c->ownStream) cudaStreamDestroy(c->ownStream);
    delete c;
}

const char* rt_last_error(const rt_ctx* c) { return c ? c->err.c_str() : "null context"; }

int rt_upload_scene(rt_ctx* c, const void* light32, const void* camera80, const void* shapes, int nShapes,
                    const void* nodes, int nNodes, const int32_t* indices, int nIndices) {
    if (!c) return RT_ERR_ARG;
    if ((shapes && nShapes < 0) || (nodes && nNodes < 0) || (indices && nIndices < 0)) return fail(c, RT_ERR_ARG, "negative count");
    CU(c, cudaSetDevice(c->device));
    if (light32)  { std::memcpy(&c->light, light32, sizeof c->light); c->haveLight = true; }
    if (camera80) { std::memcpy(&c->cam, camera80, sizeof c->cam); c->haveCam = true; }
    // A re-upload of an unchanged topology (same shape types, same index list, same node graph) keeps
    // the host-built layout: only the bytes go up again and the sub-tree boxes are refitted.
    if (shapes) {
        wait_copies(c);
        bool sameTypes = c->haveScene && nShapes == c->nShapes && !c->topoDirty;
        const rt_flat_shape* in = static_cast<const rt_flat_shape*>(shapes);
        if (!nodes && !indices && c->haveScene && nShapes < c->nShapes && c->nNodes > 0) {
            // a shorter shape buffer on its own: the tree the context holds would index past its end.  It is dropped
            // (useBVH frames hit nothing) until nodes / indices are supplied or rt_build_bvh builds a new one.
            c->nNodes = 0; c->indices.clear(); c->topoDirty = true; c->nodesDirty = true; c->hostNodesStale = false;
            c->hostTreeStale = false; c->layoutOnDevice = false; c->bIdxCount = 0;
        }
        if (sameTypes) {                   // (without an early exit: the usual answer is "the same")
            unsigned diff = 0u;
            for (int i = 0; i < nShapes; ++i) diff |= (unsigned)(in[i].type != c->hShapes.p[i].type);
            sameTypes = diff == 0u;
        }
        CU(c, c->hShapes.reserve((size_t)nShapes, 0));
        if (nShapes > 0) std::memcpy(c->hShapes.p, shapes, (size_t)nShapes * sizeof(rt_flat_shape));
        c->nShapes = nShapes; c->shapeLo = 0; c->shapeHi = nShapes;
        c->hostShapesStale = false; c->shapesOnDevice = false; c->shapeRanges.clear();
        if (!sameTypes) c->topoDirty = true;
        if (nShapes > 0) {
            // the H2D copy of the shapes starts NOW and runs while the host stages and checks nodes and indices (flush_scene does
            // not repeat it; its copyDone event, recorded behind this copy in stream order, guards the pinned source)
            CU(c, c->dShapes.reserve((size_t)nShapes));
            cudaEventRecord(c->ev[0], c->stream); c->h2dStarted = true;
            CU(c, cudaMemcpyAsync(c->dShapes.p, c->hShapes.p, (size_t)nShapes * sizeof(rt_flat_shape), cudaMemcpyHostToDevice, c->stream));
            c->shapesOnDevice = true;
        }
    }
    if ((indices || nodes) && c->hostTreeStale) { int rc = ensure_host_tree(c); if (rc != RT_OK) return rc; }
    if (indices) {
        const bool same = c->haveScene && (size_t)nIndices == c->indices.size() &&
                          (nIndices == 0 || std::memcmp(indices, c->indices.data(), (size_t)nIndices * sizeof(int32_t)) == 0);
        if (!same) { c->indices.assign(indices, indices + nIndices); c->topoDirty = true; }
    }
    if (nodes) { int rc = set_nodes(c, nodes, nNodes); if (rc != RT_OK) return rc; }
    if (shapes || nodes || indices) c->haveScene = true;
    return flush_scene(c);
}

int rt_update_camera(rt_ctx* c, const void* camera80) {
    if (!c || !camera80) return RT_ERR_ARG;
    std::memcpy(&c->cam, camera80, sizeof c->cam); c->haveCam = true;
    return RT_OK;
}
int rt_update_light(rt_ctx* c, const void* light32) {
    if (!c || !light32) return RT_ERR_ARG;
    std::memcpy(&c->light, light32, sizeof c->light); c->haveLight = true;
    return RT_OK;
}
int rt_update_shapes(rt_ctx* c, int first, int count, const void* shapes) {
    if (!c || !shapes) return RT_ERR_ARG;
    if (first < 0 || count < 0 || first + count > c->nShapes) return fail(c, RT_ERR_ARG, "shape range outside the uploaded buffer");
    if (count == 0) return RT_OK;
    wait_copies(c);
    {   // a type change invalidates the leaf kind tags: rebuild the visit-order maps
        const rt_flat_shape* in = static_cast<const rt_flat_shape*>(shapes);
        for (int i = 0; i < count; ++i) if (in[i].type != c->hShapes.p[first + i].type) { c->topoDirty = true; break; }
    }
    std::memcpy(c->hShapes.p + first, shapes, (size_t)count * sizeof(rt_flat_shape));
    c->shapesOnDevice = false;
    if (c->shapeHi <= c->shapeLo) { c->shapeLo = first; c->shapeHi = first + count; }
    else { if (first < c->shapeLo) c->shapeLo = first; if (first + count > c->shapeHi) c->shapeHi = first + count; }
    if (c->hostShapesStale) {
        if (c->shapeRanges.size() >= 256) merge_ranges(c->shapeRanges);
        c->shapeRanges.emplace_back(first, first + count);
    }
    return RT_OK;
}
int rt_update_nodes(rt_ctx* c, const void* nodes, int nNodes) {
    if (!c || !nodes || nNodes < 0) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    if (c->hostTreeStale) { int rc = ensure_host_tree(c); if (rc != RT_OK) return rc; }
    return set_nodes(c, nodes, nNodes);
}

int rt_set_shard(rt_ctx* c, int rank, int world) {
    if (!c || world < 1 || rank < 0 || rank >= world) return RT_ERR_ARG;
    { const int rc = flush_deferred(c, false); if (rc != RT_OK) return rc; }   // its source image may be reallocated for the new shard size
    c->shardRank = rank; c->shardWorld = world; c->lptItems = -1;
    return RT_OK;
}
int64_t rt_shard_pixels(const rt_ctx* c) { return c ? out_pixels(c) : 0; }

static int render_impl(rt_ctx* c, const rt_params* p, float* dev_rgba, int32_t* dev_ids, bool count,
                       float* host_rgba, int32_t* host_ids, bool pipeline, uint8_t* host_rgba8 = nullptr);

int rt_render_device(rt_ctx* c, const rt_params* p, float* dev_rgba, int32_t* dev_ids) {
    return render_impl(c, p, dev_rgba, dev_ids, false, nullptr, nullptr, false);
}

int rt_render_stats(rt_ctx* c, const rt_params* p, uint64_t out[8], int32_t* host_cost) {
    if (!c || !p || !out) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, c->tally.reserve(RT_NTALLY));
    CU(c, cudaMemsetAsync(c->tally.p, 0, RT_NTALLY * sizeof(unsigned long long), c->stream));
    int32_t* dcost = nullptr;
    if (host_cost) { CU(c, c->ids.reserve((size_t)out_pixels(c))); dcost = c->ids.p; }
    int rc = render_impl(c, p, nullptr, dcost, true, nullptr, nullptr, false);
    if (rc != RT_OK) return rc;
    if (host_cost) CU(c, cudaMemcpyAsync(host_cost, dcost, (size_t)out_pixels(c) * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaMemcpyAsync(c->lastTally, c->tally.p, RT_NTALLY * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 8; ++i) out[i] = c->lastTally[i];
    return RT_OK;
}

int rt_item_stats(rt_ctx* c, uint64_t out[3]) {
    if (!c || !out) return RT_ERR_ARG;
    out[0] = c->lastTally[16]; out[1] = c->lastTally[17]; out[2] = c->lastTally[18];
    return RT_OK;
}

int rt_walk_stats(rt_ctx* c, uint64_t out[8]) {
    if (!c || !out) return RT_ERR_ARG;
    for (int i = 0; i < 8; ++i) out[i] = c->lastTally[8 + i];

    return RT_OK;
}

// host_rgba / host_ids != NULL: BANDED mode — the frame is rendered as up to RT_MAX_BANDS launches over
// consecutive tile rows, alternating between two streams (the next band's CTAs fill the SMs as the
// previous band's retire), and each finished band is copied to the host on a third stream while
// the following bands render, so the D2H of the 16 B/pixel frame overlaps the kernel.
static int render_impl(rt_ctx* c, const rt_params* p, float* dev_rgba, int32_t* dev_ids, bool count,
                       float* host_rgba, int32_t* host_ids, bool pipeline, uint8_t* host_rgba8) {
    if (!c || !p) return RT_ERR_ARG;
    if (!c->haveScene || !c->haveCam || !c->haveLight) return fail(c, RT_ERR_STATE, "render before scene/camera/light upload");
    if (dev_ids && !count && !(c->flags & RT_FLAG_HIT_IDS)) return fail(c, RT_ERR_ARG, "hit ids need RT_FLAG_HIT_IDS");
    CU(c, cudaSetDevice(c->device));
    int rc = flush_scene(c);
    if (rc != RT_OK) return rc;
    const int64_t px = out_pixels(c);
    // fused gather: no explicit output buffer, a frame target set, several ranks
    const bool useFrame = !dev_rgba && !host_rgba && !host_rgba8 && !count && c->frameTarget;
    const bool wantIds = (c->flags & RT_FLAG_HIT_IDS) != 0;
    const bool ownImage = !dev_rgba && !useFrame, ownIds = wantIds && !dev_ids;
    pipeline = pipeline && ownImage && (host_rgba || host_ids || host_rgba8) && !count;
    const int par = pipeline ? c->imgPar : 0;
    if (pipeline) c->imgPar ^= 1;
    if (ownImage) {
        DevBuf<float4>& img = par ? c->image2 : c->image;
        CU(c, img.reserve((size_t)px)); dev_rgba = reinterpret_cast<float*>(img.p);
        c->lastImage = img.p;
    }
    if (ownIds) { DevBuf<int>& idb = par ? c->ids2 : c->ids; CU(c, idb.reserve((size_t)px)); dev_ids = idb.p; }
    uchar4* dev8 = nullptr;
    if (host_rgba8) {
        if (!ownImage || count) return fail(c, RT_ERR_ARG, "an RGBA8 frame is rendered into the context's own image");
        CU(c, c->image8[par].reserve((size_t)px)); dev8 = c->image8[par].p;
    }
    // a frame that is not pipelined, or one that renders into the image of the deferred copy, sends that copy off first
    if (c->deferred.valid && (!pipeline || c->deferred.par == par)) { rc = flush_deferred(c, false); if (rc != RT_OK) return rc; }
    if ((ownImage || ownIds) && c->framePending[par]) {
        // an earlier pipelined frame's copy may still be reading this image: the kernels below wait for it
        cudaStreamWaitEvent(c->stream, c->evFrame[par], 0);
        if (!pipeline) c->framePending[par] = false;
    }

    DeviceScene S;
    S.nodes = c->dNodesT.p; S.nodesB = c->nDfsB > 0 ? c->dNodesB.p : nullptr; S.nNodesB = c->nDfsB; S.q0 = c->q0.p; S.q1 = c->q1.p; S.q2 = c->q2.p; S.q3 = c->q3.p;
    S.ptype = c->ptype.p; S.pshape = c->dPshape.p; S.prank = c->dPrank.p; S.shapes = c->dShapes.p;
    S.nNodes = c->nDfs; S.nPrims = c->nPrims; S.nShapes = c->nShapes; S.boxesFinite = c->boxesFinite ? 1 : 0;
    S.omax = c->omax;

    FrameParams F;
    std::memcpy(F.camP, c->cam.Position, 12); std::memcpy(F.camF, c->cam.Front, 12);
    std::memcpy(F.camU, c->cam.Up, 12); std::memcpy(F.camR, c->cam.Right, 12);
    // getRay's per-frame constants (gpu_shader.comp:156-157), evaluated once on the host
    F.planeH = 2.0f * std::tan((c->cam.fov / 2.0f) * 0.017453292519943295f);
    F.planeW = F.planeH * c->cam.aspectRatio;
    std::memcpy(F.lightPos, c->light.position, 12); std::memcpy(F.lightCol, c->light.color, 12);
    F.resW = p->screenRes[0]; F.resH = p->screenRes[1];
    F.maxBounces = p->maxBounces; F.useBVH = p->useBVH != 0; F.useFresnel = p->useFresnel != 0; F.useMT = p->useMollerTrumbore != 0;
    F.W = c->W; F.H = c->H; F.tilesX = tiles_x(c); F.tilesY = tiles_y(c);
    F.shardRank = c->shardRank; F.shardWorld = c->shardWorld; F.nTilesMine = tiles_mine(c);
    F.writeIds = (count && dev_ids) ? 2 : wantIds ? 1 : 0;
    F.frameTarget = 0; F.doneWord = nullptr; F.ctaDone = nullptr; F.doneSeq = 0;
    if (useFrame) { F.frameTarget = 1; dev_rgba = c->frameTarget; }
    if (useFrame && c->syncOn) {
        const size_t frameFloats = 4 * (size_t)c->W * (size_t)c->H;
        const unsigned seq = ++c->frameSeq;
        dev_rgba = c->frameTarget + (seq & 1u) * frameFloats;
        if (c->shardRank != 0) {
            // buffer seq & 1 held frame seq - 2: wait (in the stream, no kernel spins) until rank 0 has released it
            if (seq > 2 && c->waitValue32(c->stream, (unsigned long long)(uintptr_t)c->releaseLocal.p, seq - 2, 0u /* GEQ */) != 0)
                return fail(c, RT_ERR_CUDA, "cuStreamWaitValue32 (release word)");
            F.doneWord = reinterpret_cast<unsigned*>(c->frameTarget + 2 * frameFloats) + (seq & 1u);   // this buffer's completion counter
            F.ctaDone = c->ctaDone.p; F.doneSeq = seq;
        }
    }

    const int tilesMine = tiles_mine(c);
    const bool toHost = host_rgba || host_ids || host_rgba8;
    // bands: whole tile rows (world == 1: a band is a contiguous block of image rows)
    const int rowTiles = c->shardWorld == 1 ? F.tilesX : 1;
    int nBands = 1;
    // a frame of up to 2 M pixels is ONE launch (heavy-first order and split blocks need the whole frame, and a small frame's
    // copy is short: car 800x600 end to end 0.93 -> 0.63 ms); pipelined, its copy overlaps the NEXT frame's render
#ifndef RT_BAND_MIN_PIXELS
#define RT_BAND_MIN_PIXELS ((int64_t)1 << 21)
#endif
    if (toHost && !count && px > RT_BAND_MIN_PIXELS) {
        const int rows = (tilesMine + rowTiles - 1) / rowTiles;
        nBands = rows / RT_BAND_ROWS;                    // at least RT_BAND_ROWS tile rows per band
        if (nBands > RT_MAX_BANDS) nBands = RT_MAX_BANDS;
        if (nBands < 1) nBands = 1;
        // an RGBA8 frame's copy is a quarter as long: fewer, larger bands (4K car, upload + frame: 2.10 / 2.04 / 1.96 / 1.98 / 2.06 ms with 1 / 2 / 3 / 4 / 8 bands)
        if (host_rgba8 && nBands > RT_MAX_BANDS_RGBA8) nBands = RT_MAX_BANDS_RGBA8;
    }
    CU(c, c->queue.reserve(RT_MAX_BANDS));
    // the walked list is far larger than L1: the instantiation with more resident warps (rt_kernels.cuh: RT_MIN_BLOCKS_BIG)
    const bool big = !count && F.useBVH && (c->nDfsB > 0 ? c->nDfsB : c->nDfs) >= RT_BIG_SCENE_NODES;
    const int variant = big ? 4 + (F.useMT ? 1 : 0) : (F.useMT ? 1 : 0) + (count ? 2 : 0);
    if (tilesMine > 0 && c->gridBlocks[variant] == 0) {       // persistent grid: fill every SM once
        int perSM = 0;
        const void* fn = variant == 0 ? (const void*)render_kernel<false, false, RT_MIN_BLOCKS> : variant == 1 ? (const void*)render_kernel<true, false, RT_MIN_BLOCKS>
                       : variant == 2 ? (const void*)render_kernel<false, true, RT_MIN_BLOCKS> : variant == 3 ? (const void*)render_kernel<true, true, RT_MIN_BLOCKS>
                       : variant == 4 ? (const void*)render_kernel<false, false, RT_MIN_BLOCKS_BIG> : (const void*)render_kernel<true, false, RT_MIN_BLOCKS_BIG>;
        CU(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, fn, RT_BLOCK_THREADS, 0));
        c->gridBlocks[variant] = c->smCount * (perSM > 0 ? perSM : 1);
    }
    // Few items per warp (an eighth of a frame, 800x600): the launch is bound by its slowest block's chain of dependent walks,
    // and an iteration of that chain waits for the issue slots of the other warps of its scheduler — fewer resident blocks per
    // SM shorten the chain at the price of throughput the launch does not need (RT_FEW_BLOCKS_PER_SM, 0 = off).
    int gridCap = c->gridBlocks[variant];
    if (c->fewPerSm > 0 && nBands == 1 && !count && !big && tilesMine * RT_WARPS_PER_TILE < 8 * gridCap * (RT_BLOCK_THREADS / 32)) {
        const int g = c->smCount * c->fewPerSm;
        if (g < gridCap) gridCap = g;
    }
    // heavy-first order (single-launch renders): the list is made from the previous frame's item costs
    F.itemCost = nullptr; F.heavyList = nullptr; F.heavyFlag = nullptr; F.lptCtl = nullptr; F.lptPar = 0;
    unsigned int* queue0 = c->queue.p;                         // claim counter of the plain queue (band 0)
    const bool heavyFirst = nBands == 1 && !count && tilesMine > 0 && tilesMine * RT_WARPS_PER_TILE < (1 << 24) && !(c->flags & RT_FLAG_NO_HEAVY_FIRST);
    cudaEventRecord(c->ev[4], c->stream);
    if (heavyFirst) {
        const int items = tilesMine * RT_WARPS_PER_TILE;
        CU(c, c->itemCost.reserve((size_t)items)); CU(c, c->heavyList.reserve((size_t)items * (1u << RT_SPLIT_MAXLG)));
        CU(c, c->heavyFlag.reserve((size_t)items)); CU(c, c->lptCtl.reserve(8));
        if (c->lptItems != items) {                            // new frame geometry: no costs yet, every slot of the control block zero
            CU(c, cudaMemsetAsync(c->lptCtl.p, 0, 8 * sizeof(unsigned), c->stream));
            c->lptItems = items; c->lptFrame = 0;
        } else {
            const int gridWarps = gridCap * (RT_BLOCK_THREADS / 32);
            lpt_prepare_kernel<<<(items + 255) / 256, 256, 0, c->stream>>>(c->itemCost.p, items, c->lptCtl.p, c->heavyList.p, c->heavyFlag.p,
                gridWarps, items < 5 * gridWarps ? RT_HEAVY_FACTOR_FEW : RT_HEAVY_FACTOR, (int)(c->lptFrame & 1u));
            c->launches += 1;
        }
        F.itemCost = c->itemCost.p; F.heavyList = c->heavyList.p; F.heavyFlag = c->heavyFlag.p; F.lptCtl = c->lptCtl.p;
        F.lptPar = (int)(c->lptFrame & 1u);
        queue0 = c->lptCtl.p + 3;
        c->lptFrame++;
    } else {
        CU(c, cudaMemsetAsync(c->queue.p, 0, RT_MAX_BANDS * sizeof(unsigned int), c->stream));
    }
    // the previous pipelined frame's copy starts here: after the small launches of this frame, beside its render kernel
    if (c->deferred.valid) { rc = flush_deferred(c, true); if (rc != RT_OK) return rc; }
    cudaEvent_t evFork = c->evBand[RT_MAX_BANDS], evJoin = c->evBand[RT_MAX_BANDS + 1], evCopied = c->evBand[RT_MAX_BANDS + 2];
    if (nBands > 1 || toHost) {
        cudaEventRecord(evFork, c->stream);
        if (nBands > 1) cudaStreamWaitEvent(c->stream2, evFork, 0);
    }
    const int rowsTotal = (tilesMine + rowTiles - 1) / rowTiles;
    bool firstCopy = true;
    for (int band = 0; band < nBands && tilesMine > 0; ++band) {
        const int r0 = (int)((long long)rowsTotal * band / nBands), r1 = (int)((long long)rowsTotal * (band + 1) / nBands);
        const int tb = r0 * rowTiles, te = r1 * rowTiles < tilesMine ? r1 * rowTiles : tilesMine;
        if (te <= tb) continue;
        F.tileBegin = tb; F.nTilesMine = te - tb;
        cudaStream_t rs = (band & 1) ? c->stream2 : c->stream;
        const int items = F.nTilesMine * RT_WARPS_PER_TILE;
        const int needed = (items + RT_BLOCK_THREADS / 32 - 1) / (RT_BLOCK_THREADS / 32);
        int grid = needed < gridCap ? needed : gridCap;
        const dim3 blk(RT_BLOCK_THREADS);
        float4* o4 = reinterpret_cast<float4*>(dev_rgba);
        unsigned long long* tl = c->tally.p;
        unsigned int* q = band == 0 ? queue0 : c->queue.p + band;
        if (count) {
            if (F.useMT) render_kernel<true, true, RT_MIN_BLOCKS><<<grid, blk, 0, rs>>>(S, F, o4, dev_ids, tl, q);
            else         render_kernel<false, true, RT_MIN_BLOCKS><<<grid, blk, 0, rs>>>(S, F, o4, dev_ids, tl, q);
        } else {
            if (big) {
                if (F.useMT) render_kernel<true, false, RT_MIN_BLOCKS_BIG><<<grid, blk, 0, rs>>>(S, F, o4, dev_ids, nullptr, q);
                else         render_kernel<false, false, RT_MIN_BLOCKS_BIG><<<grid, blk, 0, rs>>>(S, F, o4, dev_ids, nullptr, q);
            } else {
                if (F.useMT) render_kernel<true, false, RT_MIN_BLOCKS><<<grid, blk, 0, rs>>>(S, F, o4, dev_ids, nullptr, q);
                else         render_kernel<false, false, RT_MIN_BLOCKS><<<grid, blk, 0, rs>>>(S, F, o4, dev_ids, nullptr, q);
            }
        }
        c->launches++;
        if (toHost) {
            // pixels of this band in the output buffer: whole image rows (world == 1) or whole compact tiles
            size_t p0, p1;
            if (c->shardWorld == 1) {
                const size_t y0 = (size_t)r0 * RT_TILE_H, y1 = (size_t)r1 * RT_TILE_H < (size_t)c->H ? (size_t)r1 * RT_TILE_H : (size_t)c->H;
                p0 = y0 * (size_t)c->W; p1 = y1 * (size_t)c->W;
            } else {
                p0 = (size_t)tb * RT_TILE_W * RT_TILE_H; p1 = (size_t)te * RT_TILE_W * RT_TILE_H;
            }
            if (dev8 && p1 > p0) {               // K5 behind K1 on the band's stream: the display quad's arithmetic on these pixels
                const long long n8 = (long long)(p1 - p0);
                present_kernel<<<(unsigned)((n8 + 256 * RT_PRESENT_UNROLL - 1) / (256 * RT_PRESENT_UNROLL)), 256, 0, rs>>>(
                    reinterpret_cast<const float4*>(dev_rgba) + p0, dev8 + p0, n8);
                c->launches++;
            }
            cudaEventRecord(c->evBand[band], rs);
            cudaStreamWaitEvent(c->copyStream, c->evBand[band], 0);
            if (pipeline && nBands == 1 && !(c->flags & RT_FLAG_NO_DEFERRED_COPY)) {
                rt_ctx::DeferredCopy& d = c->deferred;
                d.valid = true; d.hostRgba = host_rgba ? host_rgba + 4 * p0 : nullptr; d.devRgba = dev_rgba + 4 * p0;
                d.hostIds = host_ids ? host_ids + p0 : nullptr; d.devIds = host_ids ? dev_ids + p0 : nullptr; d.pixels = p1 - p0; d.par = par;
                d.host8 = host_rgba8 ? host_rgba8 + 4 * p0 : nullptr; d.dev8 = host_rgba8 ? dev8 + p0 : nullptr;
                c->framePending[par] = true;             // evFrame[par] is recorded when the copy is issued (flush_deferred)
                continue;
            }
            if (firstCopy) { cudaEventRecord(c->ev[6], c->copyStream); firstCopy = false; }
            if (host_rgba) CU(c, cudaMemcpyAsync(host_rgba + 4 * p0, dev_rgba + 4 * p0, (p1 - p0) * sizeof(float4), cudaMemcpyDeviceToHost, c->copyStream));
            if (host_ids) CU(c, cudaMemcpyAsync(host_ids + p0, dev_ids + p0, (p1 - p0) * sizeof(int), cudaMemcpyDeviceToHost, c->copyStream));
            if (host_rgba8) CU(c, cudaMemcpyAsync(host_rgba8 + 4 * p0, dev8 + p0, (p1 - p0) * sizeof(uchar4), cudaMemcpyDeviceToHost, c->copyStream));
        }
    }
    if (nBands > 1) { cudaEventRecord(evJoin, c->stream2); cudaStreamWaitEvent(c->stream, evJoin, 0); }
    cudaEventRecord(c->ev[5], c->stream); c->timed[2] = true;
    if (toHost && !firstCopy) {
        cudaEventRecord(c->ev[7], c->copyStream); c->timed[3] = true;
        if (pipeline) {
            cudaEventRecord(c->evFrame[par], c->copyStream); c->framePending[par] = true;
        } else {
            cudaEventRecord(evCopied, c->copyStream);
            cudaStreamWaitEvent(c->stream, evCopied, 0);  // everything is ordered on the context stream again
        }
    }
    CU(c, cudaGetLastError());
    return RT_OK;
}

int rt_render(rt_ctx* c, const rt_params* p, float* host_rgba, int32_t* host_hit_ids) {
    if (!c || !p) return RT_ERR_ARG;
    if (host_hit_ids && !(c->flags & RT_FLAG_HIT_IDS)) return fail(c, RT_ERR_ARG, "hit ids need RT_FLAG_HIT_IDS");
    int rc = render_impl(c, p, nullptr, nullptr, false, host_rgba, host_hit_ids, false);
    if (rc != RT_OK) return rc;
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_render_async(rt_ctx* c, const rt_params* p, float* host_rgba, int32_t* host_hit_ids) {
    if (!c || !p) return RT_ERR_ARG;
    if (host_hit_ids && !(c->flags & RT_FLAG_HIT_IDS)) return fail(c, RT_ERR_ARG, "hit ids need RT_FLAG_HIT_IDS");
    return render_impl(c, p, nullptr, nullptr, false, host_rgba, host_hit_ids, true);
}

int rt_render_rgba8(rt_ctx* c, const rt_params* p, uint8_t* host_rgba8, int32_t* host_hit_ids) {
    if (!c || !p || !host_rgba8) return RT_ERR_ARG;
    if (host_hit_ids && !(c->flags & RT_FLAG_HIT_IDS)) return fail(c, RT_ERR_ARG, "hit ids need RT_FLAG_HIT_IDS");
    int rc = render_impl(c, p, nullptr, nullptr, false, nullptr, host_hit_ids, false, host_rgba8);
    if (rc != RT_OK) return rc;
    CU(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_render_rgba8_async(rt_ctx* c, const rt_params* p, uint8_t* host_rgba8, int32_t* host_hit_ids) {
    if (!c || !p || !host_rgba8) return RT_ERR_ARG;
    if (host_hit_ids && !(c->flags & RT_FLAG_HIT_IDS)) return fail(c, RT_ERR_ARG, "hit ids need RT_FLAG_HIT_IDS");
    return render_impl(c, p, nullptr, nullptr, false, nullptr, host_hit_ids, true, host_rgba8);
}

int rt_wait_frames(rt_ctx* c, int max_in_flight) {
    if (!c || max_in_flight < 0) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    // at most two frames are ever in flight (two device images); imgPar is the image of the OLDER one
    const int older = c->imgPar, newer = c->imgPar ^ 1;
    if (c->deferred.valid && (max_in_flight == 0 || c->deferred.par == older)) { const int rc = flush_deferred(c, false); if (rc != RT_OK) return rc; }
    if (max_in_flight <= 1 && c->framePending[older]) { CU(c, cudaEventSynchronize(c->evFrame[older])); c->framePending[older] = false; }
    if (max_in_flight == 0 && c->framePending[newer]) { CU(c, cudaEventSynchronize(c->evFrame[newer])); c->framePending[newer] = false; }
    return RT_OK;
}

int rt_set_frame_target(rt_ctx* c, float* dev_frame) {
    if (!c) return RT_ERR_ARG;
    c->frameTarget = dev_frame;
    c->syncOn = false;                 // a new target: signalled mode is switched on again by rt_sync_enable
    return RT_OK;
}

int rt_frame_export(rt_ctx* c, unsigned char handle[64], float** dev_frame) {
    if (!c || !handle) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    // two frames (signalled mode alternates them) + 64 completion words
    const size_t want = 2 * (size_t)c->W * (size_t)c->H + 16;
    if (c->frame.cap < want) { c->frame.release(); CU(c, cudaMalloc(&c->frame.p, want * sizeof(float4))); c->frame.cap = want; }
    CU(c, cudaMemset(c->frame.p + 2 * (size_t)c->W * (size_t)c->H, 0, 16 * sizeof(float4)));
    cudaIpcMemHandle_t h;
    CU(c, cudaIpcGetMemHandle(&h, c->frame.p));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    std::memcpy(handle, &h, 64);
    c->frameTarget = reinterpret_cast<float*>(c->frame.p);
    if (dev_frame) *dev_frame = c->frameTarget;
    return RT_OK;
}

int rt_frame_open(rt_ctx* c, const unsigned char handle[64]) {
    if (!c || !handle) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    if (c->ipcOpened) { cudaIpcCloseMemHandle(c->ipcOpened); c->ipcOpened = nullptr; c->frameTarget = nullptr; }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    void* p = nullptr;
    CU(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->ipcOpened = p;
    c->frameTarget = static_cast<float*>(p);
    return RT_OK;
}

int rt_frame_acquire(rt_ctx* c, float** dev_frame) {
    if (!c) return RT_ERR_ARG;
    if (!c->syncOn || c->shardRank != 0 || !c->frame.p) return fail(c, RT_ERR_STATE, "rt_frame_acquire: rank 0 in signalled mode only");
    CU(c, cudaSetDevice(c->device));
    // One counter per buffer; every other rank's last thread block adds 1 per frame.  Frame f is the ceil(f / 2)-th frame of
    // its buffer: complete at ceil(f / 2) * (world - 1) (compared modulo 2^32).  Per buffer, because a rank may already have
    // finished frame f + 1 (the other buffer) when another has not finished f; nobody starts f + 2 before f was released.
    unsigned* counter = reinterpret_cast<unsigned*>(c->frame.p + 2 * (size_t)c->W * (size_t)c->H) + (c->frameSeq & 1u);
    const unsigned expect = ((c->frameSeq + 1u) / 2u) * (unsigned)(c->shardWorld - 1);
    if (c->shardWorld > 1 && c->waitValue32(c->stream, (unsigned long long)(uintptr_t)counter, expect, 0u /* GEQ */) != 0)
        return fail(c, RT_ERR_CUDA, "cuStreamWaitValue32 (completion counter)");
    if (dev_frame) *dev_frame = reinterpret_cast<float*>(c->frame.p) + (c->frameSeq & 1u) * 4 * (size_t)c->W * (size_t)c->H;
    return RT_OK;
}

int rt_frame_release(rt_ctx* c) {
    if (!c) return RT_ERR_ARG;
    if (!c->syncOn || c->shardRank != 0) return fail(c, RT_ERR_STATE, "rt_frame_release: rank 0 in signalled mode only");
    CU(c, cudaSetDevice(c->device));
    if (c->shardWorld > 1) {
        release_kernel<<<1, 64, 0, c->stream>>>(c->dPeerRelease.p, c->shardWorld, c->frameSeq);
        c->launches++;
        CU(c, cudaGetLastError());
    }
    return RT_OK;
}

int rt_frame_download(rt_ctx* c, float* host_rgba) {
    if (!c || !host_rgba) return RT_ERR_ARG;
    if (!c->frame.p) return fail(c, RT_ERR_STATE, "no exported frame on this context");
    CU(c, cudaSetDevice(c->device));
    const float* src = reinterpret_cast<const float*>(c->frame.p);
    if (c->syncOn) { float* f = nullptr; int rc = rt_frame_acquire(c, &f); if (rc != RT_OK) return rc; src = f; }
    CU(c, cudaMemcpyAsync(host_rgba, src, (size_t)c->W * (size_t)c->H * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    if (c->syncOn) return rt_frame_release(c);
    return RT_OK;
}

int rt_sync_local(rt_ctx* c, unsigned** dev_release_word) {
    if (!c) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, c->releaseLocal.reserve(1));
    CU(c, cudaMemset(c->releaseLocal.p, 0, sizeof(unsigned)));
    if (dev_release_word) *dev_release_word = c->releaseLocal.p;
    return RT_OK;
}

int rt_sync_export(rt_ctx* c, unsigned char handle[64]) {
    if (!c || !handle) return RT_ERR_ARG;
    int rc = rt_sync_local(c, nullptr);
    if (rc != RT_OK) return rc;
    cudaIpcMemHandle_t h;
    CU(c, cudaIpcGetMemHandle(&h, c->releaseLocal.p));
    std::memcpy(handle, &h, 64);
    return RT_OK;
}

int rt_sync_peer(rt_ctx* c, int rank, unsigned* dev_word, const unsigned char handle[64]) {
    if (!c || rank < 1 || rank >= 64 || (!dev_word && !handle)) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    if (c->peerIpc[rank]) { cudaIpcCloseMemHandle(c->peerIpc[rank]); c->peerIpc[rank] = nullptr; }
    if (!dev_word) {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, 64);
        void* p = nullptr;
        CU(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        c->peerIpc[rank] = p; dev_word = static_cast<unsigned*>(p);
    }
    c->peerRelease[rank] = dev_word;
    return RT_OK;
}

int rt_sync_enable(rt_ctx* c, int world) {
    if (!c || world < 0 || world > 64) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    if (world == 0) { c->syncOn = false; return RT_OK; }
    if (world != c->shardWorld) return fail(c, RT_ERR_ARG, "rt_sync_enable: world differs from rt_set_shard's");
    if (!c->frameTarget) return fail(c, RT_ERR_STATE, "rt_sync_enable needs the exported frame (rt_frame_export / rt_frame_open / rt_set_frame_target)");
    if (!c->waitValue32) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn || q != cudaDriverEntryPointSuccess)
            return fail(c, RT_ERR_CUDA, "stream memory operations (cuStreamWaitValue32) are not available");
        c->waitValue32 = reinterpret_cast<int (*)(cudaStream_t, unsigned long long, unsigned, unsigned)>(fn);
    }
    CU(c, cudaStreamSynchronize(c->stream));
    if (c->shardRank == 0) {
        if (!c->frame.p) return fail(c, RT_ERR_STATE, "rt_sync_enable on rank 0 needs rt_frame_export");
        for (int r = 1; r < world; ++r) if (!c->peerRelease[r]) return fail(c, RT_ERR_STATE, "rt_sync_enable: a rank's release word is missing (rt_sync_peer)");
        CU(c, c->dPeerRelease.reserve(64));
        CU(c, cudaMemcpy(c->dPeerRelease.p, c->peerRelease, 64 * sizeof(unsigned*), cudaMemcpyHostToDevice));
        CU(c, cudaMemset(c->frame.p + 2 * (size_t)c->W * (size_t)c->H, 0, 16 * sizeof(float4)));
    } else {
        if (!c->releaseLocal.p) return fail(c, RT_ERR_STATE, "rt_sync_enable: rt_sync_local / rt_sync_export first");
        CU(c, cudaMemset(c->releaseLocal.p, 0, sizeof(unsigned)));
        CU(c, c->ctaDone.reserve(1));
        CU(c, cudaMemset(c->ctaDone.p, 0, sizeof(unsigned)));
    }
    c->frameSeq = 0; c->syncOn = true;
    return RT_OK;
}

int rt_host_frame_register(rt_ctx* c, void* host_frame, size_t bytes, float** dev_alias) {
    if (!c || !host_frame || !dev_alias) return RT_ERR_ARG;
    if (bytes < (size_t)c->W * (size_t)c->H * sizeof(float4)) return fail(c, RT_ERR_ARG, "host frame smaller than W*H float4");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaHostRegister(host_frame, bytes, cudaHostRegisterMapped | cudaHostRegisterPortable));
    void* d = nullptr;
    cudaError_t e = cudaHostGetDevicePointer(&d, host_frame, 0);
    if (e != cudaSuccess) { cudaHostUnregister(host_frame); CU(c, e); }
    *dev_alias = static_cast<float*>(d);
    c->hostFrames.push_back(host_frame);
    return RT_OK;
}

int rt_host_frame_unregister(rt_ctx* c, void* host_frame) {
    if (!c || !host_frame) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    CU(c, cudaHostUnregister(host_frame));
    c->hostFrames.erase(std::remove(c->hostFrames.begin(), c->hostFrames.end(), host_frame), c->hostFrames.end());
    return RT_OK;
}

int rt_anim_clear(rt_ctx* c) {
    if (!c) return RT_ERR_ARG;
    c->animDefs.clear(); c->animItems.clear(); c->animDirty = true; c->animItemsDirty = true;
    return RT_OK;
}

int rt_anim_bounce(rt_ctx* c, int shape, float origin_y, float amplitude, float frequency) {
    if (!c) return RT_ERR_ARG;
    if (shape < 0 || shape >= c->nShapes) return fail(c, RT_ERR_ARG, "animated shape outside the uploaded buffer");
    if (c->hShapes.p[shape].type != RT_SHAPE_SPHERE) return fail(c, RT_ERR_ARG, "rt_anim_bounce wants a sphere");
    c->animItems.push_back(AnimItem{shape, (int)c->animDefs.size()});
    c->animDefs.push_back(rt_ctx::AnimDef{RT_ANIM_BOUNCE, origin_y, amplitude, frequency, 0.f, 0.f, 0.f, 0.f});
    c->animDirty = true; c->animItemsDirty = true;
    return RT_OK;
}

int rt_anim_rotate_z(rt_ctx* c, const int32_t* shapes, int count, const float center[3], float speed) {
    if (!c || !shapes || !center || count < 0) return RT_ERR_ARG;
    for (int i = 0; i < count; ++i) {
        if (shapes[i] < 0 || shapes[i] >= c->nShapes) return fail(c, RT_ERR_ARG, "animated shape outside the uploaded buffer");
        if (c->hShapes.p[shapes[i]].type != RT_SHAPE_TRIANGLE) return fail(c, RT_ERR_ARG, "rt_anim_rotate_z wants triangles");
    }
    for (int i = 0; i < count; ++i) c->animItems.push_back(AnimItem{shapes[i], (int)c->animDefs.size()});
    c->animDefs.push_back(rt_ctx::AnimDef{RT_ANIM_ROTATE_Z, 0.f, 0.f, 0.f, center[0], center[1], center[2], speed});
    c->animDirty = true; c->animItemsDirty = true;
    return RT_OK;
}

namespace {
// Tables of the device animation: items, per-shape flags, bvhIndices and the raw tree's parent / leaf lists.
int upload_anim_tables(rt_ctx* c) {
    if (c->hostTreeStale) { int rc = ensure_host_tree(c); if (rc != RT_OK) return rc; }
    const int nN = c->nNodes, nS = c->nShapes;
    std::vector<int> aux((size_t)nN, -1);            // parent per raw node, then the leaf list
    std::vector<int> leaves;
    for (int i = 0; i < nN; ++i) {
        const rt_flat_node& x = c->hNodes.p[i];
        if (x.leftChild == -1) {
            if (x.numShapes < 0 || x.startShapeIdx < 0 || (size_t)x.startShapeIdx + (size_t)x.numShapes > c->indices.size())
                return fail(c, RT_ERR_SCENE, "leaf index range outside bvhIndices");
            leaves.push_back(i);
            continue;
        }
        const int ch[2] = {x.leftChild, x.rightChild};
        for (int k : ch) {
            if (k < 0 || k >= nN || k == i || aux[(size_t)k] != -1) return fail(c, RT_ERR_SCENE, "device animation needs a proper tree (one parent per node)");
            aux[(size_t)k] = i;
        }
    }
    if (nN > 0 && aux[(size_t)nN - 1] != -1) return fail(c, RT_ERR_SCENE, "the root (last node) has a parent");
    c->nLeavesRaw = (int)leaves.size();
    aux.insert(aux.end(), leaves.begin(), leaves.end());
    std::vector<unsigned char> flag((size_t)nS, 0);
    for (const AnimItem& it : c->animItems) if (it.shape < nS) flag[(size_t)it.shape] = 1;
    CU(c, c->dNodeAux.reserve(aux.size())); CU(c, c->dGrowArrived.reserve((size_t)nN));
    if (nN > 0) CU(c, cudaMemsetAsync(c->dGrowArrived.p, 0, (size_t)nN * sizeof(unsigned), c->stream));   // once per table upload: K7 leaves them zero
    CU(c, c->dAnimFlag.reserve((size_t)nS)); CU(c, c->dIndices.reserve(c->indices.size()));
    CU(c, c->dAnimItems.reserve(c->animItems.size())); CU(c, c->dAnimEntries.reserve(c->animDefs.size()));
    CU(c, c->hAnimEntries.reserve(c->animDefs.size(), 0));
    // pageable sources: these copies return after the data has been staged
    if (!aux.empty()) CU(c, cudaMemcpyAsync(c->dNodeAux.p, aux.data(), aux.size() * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    if (nS > 0) CU(c, cudaMemcpyAsync(c->dAnimFlag.p, flag.data(), (size_t)nS, cudaMemcpyHostToDevice, c->stream));
    if (!c->indices.empty()) CU(c, cudaMemcpyAsync(c->dIndices.p, c->indices.data(), c->indices.size() * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    if (!c->animItems.empty()) CU(c, cudaMemcpyAsync(c->dAnimItems.p, c->animItems.data(), c->animItems.size() * sizeof(AnimItem), cudaMemcpyHostToDevice, c->stream));
    // the nodes that can change: the leaves that hold an animated shape and their ancestors, by height above such a leaf
    std::vector<int> height((size_t)nN, -1);
    int nListed = 0, maxH = -1;
    for (int lf : leaves) {
        const rt_flat_node& x = c->hNodes.p[lf];
        bool any = false;
        for (int q = 0; q < x.numShapes && !any; ++q) {
            const int sIdx = c->indices[(size_t)x.startShapeIdx + (size_t)q];
            any = sIdx >= 0 && sIdx < nS && flag[(size_t)sIdx];
        }
        if (!any) continue;
        int h = 0;
        for (int j = lf; j >= 0 && height[(size_t)j] < h; j = aux[(size_t)j], ++h) {   // aux[0 .. nN): parents
            if (height[(size_t)j] < 0) ++nListed;
            height[(size_t)j] = h;
            if (h > maxH) maxH = h;
            if (h > nN) return fail(c, RT_ERR_SCENE, "device animation needs a proper tree (a cycle of parents)");
        }
    }
    c->growLevels = 0;
    if (nListed > 0 && nListed <= RT_GROW_PATH_MAX && maxH + 1 <= RT_LEVELS_MAX) {
        std::vector<int> sched((size_t)maxH + 2, 0);
        for (int j = 0; j < nN; ++j) if (height[(size_t)j] >= 0) sched[(size_t)height[(size_t)j] + 1]++;
        for (int h = 0; h <= maxH; ++h) sched[(size_t)h + 1] += sched[(size_t)h];
        std::vector<int4> ent((size_t)nListed);
        std::vector<int> fill(sched.begin(), sched.end() - 1);
        for (int j = 0; j < nN; ++j) if (height[(size_t)j] >= 0) {
            const rt_flat_node& x = c->hNodes.p[j];
            ent[(size_t)fill[(size_t)height[(size_t)j]]++] = make_int4(j, x.leftChild, x.rightChild, 0);
        }
        CU(c, c->dGrowSched.reserve(sched.size())); CU(c, c->dGrowEnt.reserve(ent.size()));
        CU(c, cudaMemcpyAsync(c->dGrowSched.p, sched.data(), sched.size() * sizeof(int), cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(c->dGrowEnt.p, ent.data(), ent.size() * sizeof(int4), cudaMemcpyHostToDevice, c->stream));
        c->growLevels = maxH + 1;
    }
    CU(c, cudaStreamSynchronize(c->stream));
    c->animDirty = false; c->animItemsDirty = false;
    return RT_OK;
}
// The item list alone (K6 without K7): what rt_animate_shapes needs.
int upload_anim_items(rt_ctx* c) {
    CU(c, c->dAnimItems.reserve(c->animItems.size())); CU(c, c->dAnimEntries.reserve(c->animDefs.size()));
    CU(c, c->hAnimEntries.reserve(c->animDefs.size(), 0));
    if (!c->animItems.empty()) CU(c, cudaMemcpyAsync(c->dAnimItems.p, c->animItems.data(), c->animItems.size() * sizeof(AnimItem), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->animItemsDirty = false;
    return RT_OK;
}
}  // namespace

static int animate_impl(rt_ctx* c, float t, float dt, bool grow);
int rt_animate(rt_ctx* c, float t, float dt) { return animate_impl(c, t, dt, true); }
int rt_animate_shapes(rt_ctx* c, float t, float dt) { return animate_impl(c, t, dt, false); }

static int animate_impl(rt_ctx* c, float t, float dt, bool grow) {
    if (!c) return RT_ERR_ARG;
    if (!c->haveScene) return fail(c, RT_ERR_STATE, "rt_animate before a scene upload");
    CU(c, cudaSetDevice(c->device));
    int rc = flush_scene(c);                       // pending host updates first: K6 works on the device copies
    if (rc != RT_OK) return rc;
    if (c->animDefs.empty() || c->animItems.empty()) return RT_OK;
    for (const AnimItem& it : c->animItems) if (it.shape >= c->nShapes) return fail(c, RT_ERR_STATE, "animation registered for a shape that is gone");
    if (grow) { if (c->animDirty) { rc = upload_anim_tables(c); if (rc != RT_OK) return rc; } }
    else if (c->animItemsDirty) { rc = upload_anim_items(c); if (rc != RT_OK) return rc; }
    const bool packed = c->animDefs.size() <= RT_ANIM_PACK;      // the frame's entries as a kernel parameter: no copy at all
    AnimPack pack;
    if (!packed) wait_copies(c);                   // the pinned entry table is the source of an async copy
    for (size_t e = 0; e < c->animDefs.size(); ++e) {
        const rt_ctx::AnimDef& d = c->animDefs[e];
        AnimEntry& o = packed ? pack.e[e] : c->hAnimEntries.p[e];
        o.kind = d.kind;
        for (float& x : o.m) x = 0.f;
        if (d.kind == RT_ANIM_BOUNCE) { o.m[0] = d.originY + d.amp * std::sin(d.freq * t); }              // main.cpp:1079-1082
        else {                                                                                             // main.cpp:1084-1109
            const float ang = d.speed * dt, center[3] = {d.cx, d.cy, d.cz}, axis[3] = {0.f, 0.f, 1.f};
            float m[16];
            rta_wheel_transform(center, std::cos(ang), std::sin(ang), axis, m);
            for (int col = 0; col < 4; ++col) for (int r = 0; r < 3; ++r) o.m[3 * col + r] = m[4 * col + r];
        }
    }
    const int nI = (int)c->animItems.size();
    const bool pathGrow = grow && c->nNodes > 0 && c->growLevels > 0 && !(c->flags & RT_FLAG_NO_ONE_BLOCK_UPDATE);
    if (pathGrow) {                                // K6 + K7 in one block: the animated shapes, then the listed nodes level by level
        if (packed) for (size_t e = c->animDefs.size(); e < RT_ANIM_PACK; ++e) { pack.e[e].kind = 0; for (float& x : pack.e[e].m) x = 0.f; }
        else {
            for (auto& pe : pack.e) { pe.kind = 0; for (float& x : pe.m) x = 0.f; }
            CU(c, cudaMemcpyAsync(c->dAnimEntries.p, c->hAnimEntries.p, c->animDefs.size() * sizeof(AnimEntry), cudaMemcpyHostToDevice, c->stream));
            cudaEventRecord(c->copyDone, c->stream); c->copyPending = true;
        }
        anim_grow_path_kernel<<<1, RT_GROW_PATH_THREADS, 0, c->stream>>>(c->dShapes.p, c->dAnimItems.p, nI, packed ? nullptr : c->dAnimEntries.p, pack,
            c->dNodes.p, c->dIndices.p, c->dAnimFlag.p, c->dGrowSched.p, c->dGrowEnt.p, c->growLevels);
        c->launches++;
        c->devNodesDirty = true;
    } else if (packed) {
        for (size_t e = c->animDefs.size(); e < RT_ANIM_PACK; ++e) { pack.e[e].kind = 0; for (float& x : pack.e[e].m) x = 0.f; }
        anim_kernel_packed<<<(nI + 127) / 128, 128, 0, c->stream>>>(c->dShapes.p, c->dAnimItems.p, nI, pack);
    } else {
        CU(c, cudaMemcpyAsync(c->dAnimEntries.p, c->hAnimEntries.p, c->animDefs.size() * sizeof(AnimEntry), cudaMemcpyHostToDevice, c->stream));
        cudaEventRecord(c->copyDone, c->stream); c->copyPending = true;
        anim_kernel<<<(nI + 127) / 128, 128, 0, c->stream>>>(c->dShapes.p, c->dAnimItems.p, nI, c->dAnimEntries.p);
    }
    c->launches++;
    if (grow && !pathGrow && c->nNodes > 0 && c->nLeavesRaw > 0) {
        grow_kernel<<<(c->nLeavesRaw + 127) / 128, 128, 0, c->stream>>>(c->dNodes.p, c->dShapes.p, c->dIndices.p, c->dAnimFlag.p,
            c->dNodeAux.p + c->nNodes, c->nLeavesRaw, c->dNodeAux.p, c->dGrowArrived.p);
        c->launches++;
        c->devNodesDirty = true;
    }
    CU(c, cudaGetLastError());
    c->devShapesDirty = true; c->hostStale = c->hostShapesStale = true;
    if (c->devNodesDirty) c->hostNodesStale = true;
    return RT_OK;
}

int rt_download_scene(rt_ctx* c, void* shapes, void* nodes) {
    if (!c) return RT_ERR_ARG;
    if (!c->haveScene) return fail(c, RT_ERR_STATE, "no scene");
    CU(c, cudaSetDevice(c->device));
    int rc = flush_scene(c);
    if (rc != RT_OK) return rc;
    CU(c, cudaStreamSynchronize(c->stream));
    if (shapes && c->nShapes > 0) CU(c, cudaMemcpy(shapes, c->dShapes.p, (size_t)c->nShapes * sizeof(rt_flat_shape), cudaMemcpyDeviceToHost));
    if (nodes && c->nNodes > 0) CU(c, cudaMemcpy(nodes, c->dNodes.p, (size_t)c->nNodes * sizeof(rt_flat_node), cudaMemcpyDeviceToHost));
    return RT_OK;
}

// K8 (rt_build.cuh): buildBVH + serializeBVH of the reference on the device copy of the shapes.
int rt_build_bvh(rt_ctx* c, int maxDepth, unsigned flags, int* nNodes_out, int* nIndices_out) {
    if (!c) return RT_ERR_ARG;
    if (maxDepth < 0 || maxDepth > 64) return fail(c, RT_ERR_ARG, "rt_build_bvh: maxDepth must be 0 ... 64 (the shader's traversal stack holds 64 entries)");
    if (!c->haveScene || c->hShapes.p == nullptr) return fail(c, RT_ERR_STATE, "rt_build_bvh before the shapes were uploaded");
    if ((size_t)c->nShapes >= (1u << 29)) return fail(c, RT_ERR_ARG, "rt_build_bvh: more than 2^29 shapes");
    CU(c, cudaSetDevice(c->device));
    const int n = c->nShapes;
    if (n == 0) {                                   // buildBVH over nothing: one empty leaf (main.cpp:1175-1193)
        CU(c, c->bHostNodes.reserve(1, 0));
        c->bNodeCount = 1; c->bIdxCount = 0;
        rt_flat_node& r = c->bHostNodes.p[0];
        std::memset(&r, 0, sizeof r);
        for (int k = 0; k < 3; ++k) { r.boundsMin[k] = INFINITY; r.boundsMax[k] = -INFINITY; }
        r.leftChild = r.rightChild = -1;
    } else {
        // the device copy of the shapes must be current: push the pending host range now (flush_scene will not repeat it)
        if (c->shapeHi > c->shapeLo && !c->shapesOnDevice) {
            CU(c, c->dShapes.reserve((size_t)n));
            if (c->hostShapesStale && !c->shapeRanges.empty()) {      // see flush_scene: only what the host rewrote
                merge_ranges(c->shapeRanges);
                for (const auto& r : c->shapeRanges)
                    CU(c, cudaMemcpyAsync(c->dShapes.p + r.first, c->hShapes.p + r.first,
                                          (size_t)(r.second - r.first) * sizeof(rt_flat_shape), cudaMemcpyHostToDevice, c->stream));
            } else {
                CU(c, cudaMemcpyAsync(c->dShapes.p + c->shapeLo, c->hShapes.p + c->shapeLo,
                                      (size_t)(c->shapeHi - c->shapeLo) * sizeof(rt_flat_shape), cudaMemcpyHostToDevice, c->stream));
            }
            cudaEventRecord(c->copyDone, c->stream); c->copyPending = true;
            c->shapesOnDevice = true;
        }
        const size_t maxNodes = 2 * (size_t)n;     // build ids 0 .. 2n - 2
        const int nBlocks = (n + RT_BLD_THREADS - 1) / RT_BLD_THREADS;
        const int nIds = 2 * n, idBlocks = (nIds + RT_BLD_THREADS - 1) / RT_BLD_THREADS;
        // int tables: 8 per id + arrived, 2 orders + 2 node maps + scan (n + 1), block sums
        const size_t nInt = 9 * maxNodes + 4 * (size_t)n + ((size_t)n + 1) + (size_t)nBlocks + 1;
        CU(c, c->bInt.reserve(nInt)); CU(c, c->bKey.reserve(6 * maxNodes));
        CU(c, c->bFlag.reserve((size_t)n)); CU(c, c->bCtl.reserve(1));
        CU(c, c->bNodes.reserve(maxNodes)); CU(c, c->bIdx.reserve((size_t)n)); CU(c, c->bRec.reserve(3 * (size_t)n));
        BuildBufs B;
        int* q = c->bInt.p;
        B.N.start = q; q += maxNodes; B.N.count = q; q += maxNodes; B.N.left = q; q += maxNodes; B.N.parent = q; q += maxNodes;
        B.N.depth = q; q += maxNodes; B.N.desc = q; q += maxNodes; B.N.idx = q; q += maxNodes; B.N.first = q; q += maxNodes;
        B.N.arrived = reinterpret_cast<unsigned*>(q); q += maxNodes;
        B.ord[0] = q; q += n; B.ord[1] = q; q += n; B.pnode[0] = q; q += n; B.pnode[1] = q; q += n;
        B.scanL = q; q += (size_t)n + 1; B.bsum = q;
        B.N.key = c->bKey.p; B.flag = c->bFlag.p; B.ctl = c->bCtl.p;
        B.R.cen = c->bRec.p; B.R.mn = c->bRec.p + n; B.R.mx = c->bRec.p + 2 * (size_t)n;
        const int agg = (flags & RT_BUILD_NO_WARP_AGGREGATION) ? 0 : 1;
        cudaStream_t st = c->stream;
        for (auto& e : c->evBuild) if (!e) CU(c, cudaEventCreate(&e));
        cudaEventRecord(c->evBuild[0], st);
        bool built = false;
        if (n <= RT_BLD_SMALL_MAX && !(flags & RT_BUILD_NO_SINGLE_BLOCK)) {
            // launch-bound sizes: one block runs the whole build and stops at the first level with nothing pending
            bld_small_kernel<<<1, RT_BLD_SMALL_THREADS, 0, st>>>(B, c->dShapes.p, n, maxDepth, c->bNodes.p, c->bIdx.p);
            c->launches += 1;
            built = true;
        } else if (n <= RT_BLD_CLUSTER_MAX && !(flags & RT_BUILD_NO_CLUSTER) && !c->noCluster) {
            // launch-bound still: ONE CLUSTER of 2 ... 16 blocks runs the whole build, cluster barriers in place of kernel boundaries
            // (16 blocks: a non-portable cluster size, available on B200; a thread then owns up to two positions)
            int blocks = 1;
            while (blocks * RT_BLD_SMALL_THREADS < n && blocks < RT_BLD_CLUSTER_MAX_BLOCKS) blocks *= 2;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)blocks); cfg.blockDim = dim3(RT_BLD_SMALL_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = st;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = (unsigned)blocks; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            const rt_flat_shape* shp = c->dShapes.p;
            cudaError_t e = cudaSuccess;
            if (blocks > 8) e = cudaFuncSetAttribute(bld_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            if (e == cudaSuccess) e = cudaLaunchKernelEx(&cfg, bld_cluster_kernel, B, shp, n, maxDepth, c->bNodes.p, c->bIdx.p);
            if (e == cudaSuccess) { c->launches += 1; built = true; }
            else { cudaGetLastError(); c->noCluster = true; }           // this device cannot co-schedule such a cluster: per-phase launches from now on
        }
        if (!built && !(flags & RT_BUILD_NO_COOPERATIVE) && !c->noCooperative && n > RT_BLD_SMALL_MAX && n <= RT_BLD_GRID_MAX) {
            // up to ~120 k shapes the launches still weigh: ONE cooperative launch, a block per SM, grid barriers in place of the kernel boundaries
            if (c->bldGridBlocks == 0) {
                int perSM = 0, coop = 0;
                cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device);
                if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, bld_grid_kernel, RT_BLD_SMALL_THREADS, 0) == cudaSuccess && perSM > 0)
                    c->bldGridBlocks = c->smCount * perSM;
                else { cudaGetLastError(); c->noCooperative = true; }
            }
            if (!c->noCooperative) {
                int blocks = (n + RT_BLD_SMALL_THREADS - 1) / RT_BLD_SMALL_THREADS;
                if (blocks > c->bldGridBlocks) blocks = c->bldGridBlocks;
                const rt_flat_shape* shp = c->dShapes.p;
                rt_flat_node* outN = c->bNodes.p; int32_t* outI = c->bIdx.p;
                int nn = n, md = maxDepth;
                void* args[] = {&B, &shp, &nn, &md, &outN, &outI};
                const cudaError_t e = cudaLaunchCooperativeKernel((const void*)bld_grid_kernel, dim3((unsigned)blocks), dim3(RT_BLD_SMALL_THREADS), args, 0, st);
                if (e == cudaSuccess) { c->launches += 1; built = true; }
                else { cudaGetLastError(); c->noCooperative = true; }
            }
        }
        if (!built) {
            bld_setup_kernel<<<idBlocks, RT_BLD_THREADS, 0, st>>>(B, n, maxDepth);
            bld_init_kernel<<<nBlocks, RT_BLD_THREADS, 0, st>>>(B, c->dShapes.p, n, agg);
            c->launches += 2;
            // Every shape alone is reached after about log2(n) levels of a balanced tree; from a few levels past that
            // on, every 8th level first asks the device whether anything is still pending (a level with nothing
            // pending costs four empty launches, the question one small round trip).
            int firstAsk = 4;
            while ((1 << (firstAsk - 4)) < n && firstAsk < 36) ++firstAsk;
            for (int level = 0; level < maxDepth; ++level) {
                if (level >= firstAsk && (level - firstAsk) % 8 == 0) {
                    int active = 1;
                    CU(c, cudaMemcpyAsync(&active, &c->bCtl.p->active[level], sizeof(int), cudaMemcpyDeviceToHost, st));
                    CU(c, cudaStreamSynchronize(st));
                    if (!active) break;
                }
                bld_classify_kernel<<<nBlocks, RT_BLD_THREADS, 0, st>>>(B, n, level);
                bld_scan_blocks_kernel<<<1, 1024, 0, st>>>(B, n, nBlocks, level);
                bld_scan_apply_kernel<<<nBlocks, RT_BLD_THREADS, 0, st>>>(B, n, level);
                bld_split_kernel<<<nBlocks, RT_BLD_THREADS, 0, st>>>(B, n, level, agg);
                c->launches += 4;
            }
            bld_desc_kernel<<<idBlocks, RT_BLD_THREADS, 0, st>>>(B, nIds);
            bld_number_kernel<<<idBlocks, RT_BLD_THREADS, 0, st>>>(B, nIds);
            bld_emit_kernel<<<idBlocks, RT_BLD_THREADS, 0, st>>>(B, nIds, c->bNodes.p, c->bIdx.p);
            c->launches += 3;
        }
        cudaEventRecord(c->evBuild[1], st);
        CU(c, cudaGetLastError());
        if (flags & RT_BUILD_DEVICE_LAYOUT) {
            // K9: list A, the primitive slots and their ranks straight from the builder's tables, where they are
            CU(c, c->kOrder.reserve(maxNodes)); CU(c, c->kLinkCnt.reserve(maxNodes));
            CU(c, c->kPshape.reserve((size_t)n)); CU(c, c->kPrank.reserve((size_t)n));
            CU(c, c->hCtl.reserve(1, 0));
            bld_list_kernel<<<idBlocks, RT_BLD_THREADS, 0, st>>>(B, nIds, c->dShapes.p, c->bNodes.p, c->kOrder.p, c->kLinkCnt.p, c->kPshape.p, c->kPrank.p);
            c->launches += 1;
            cudaEventRecord(c->evBuild[1], st);
            CU(c, cudaMemcpyAsync(c->hCtl.p, c->bCtl.p, sizeof(bld::Ctl), cudaMemcpyDeviceToHost, st));
            CU(c, cudaStreamSynchronize(st));
            const bld::Ctl& ctl = *c->hCtl.p;
            if (ctl.nNodes < 1 || (size_t)ctl.nNodes >= maxNodes) return fail(c, RT_ERR_CUDA, "rt_build_bvh: node count out of range");
            cudaEventElapsedTime(&c->lastBuildMs, c->evBuild[0], c->evBuild[1]);
            const bool plain = (c->flags & RT_FLAG_REFERENCE_WALK) != 0 || ctl.notFinite != 0;   // the host builds no sub-trees then either
            // a leaf of up to RT_DEVICE_LAYOUT_MAX_LEAF shapes is walked as it is (the host layout would expand leaves of more
            // than RT_FAT_LEAF = 8 shapes into a sub-tree; testing up to 16 shapes in a row costs less than a trip to the host)
            if (plain || ctl.maxLeaf <= RT_DEVICE_LAYOUT_MAX_LEAF) {
                const int nN = ctl.nNodes;
                wait_copies(c);
                std::swap(c->dOrder, c->kOrder); std::swap(c->dLinkCnt, c->kLinkCnt);
                std::swap(c->dPshape, c->kPshape); std::swap(c->dPrank, c->kPrank);
                CU(c, c->dNodes.reserve((size_t)nN));
                CU(c, cudaMemcpyAsync(c->dNodes.p, c->bNodes.p, (size_t)nN * sizeof(rt_flat_node), cudaMemcpyDeviceToDevice, st));
                CU(c, c->dNodesT.reserve((size_t)nN * 2));
                CU(c, c->q0.reserve((size_t)n + 4)); CU(c, c->q1.reserve((size_t)n + 4));
                CU(c, c->q2.reserve((size_t)n + 4)); CU(c, c->q3.reserve((size_t)n + 4));
                CU(c, c->ptype.reserve((size_t)n));
                c->nNodes = nN; c->nDfs = nN; c->nPrims = n; c->nSub = 0; c->nDfsB = 0; c->nRfStart = 0;
                c->bNodeCount = nN; c->bIdxCount = n;
                c->layout = rtl::Layout();
                c->boxesFinite = ctl.notFinite == 0;
                for (int k = 0; k < 6; ++k) c->rootBox[k] = ctl.rootBox[k];
                c->omax = (float)pick_omax(c);
                c->hostTreeStale = true; c->hostNodesStale = true; c->hostStale = true; c->layoutOnDevice = true;
                c->topoDirty = false; c->nodesDirty = false; c->subDirty = false;
                c->devNodesDirty = true; c->devShapesDirty = true;      // flush_scene: repack the lists and the slots, nothing to upload
                c->animDirty = true;
                if (nNodes_out) *nNodes_out = nN;
                if (nIndices_out) *nIndices_out = n;
                return RT_OK;
            }
            // a fat leaf: its sub-tree is built by the host layout builder — continue as without the flag
        }
        int nNodes = 0;                                // the numbering pass knows how many nodes there are
        CU(c, cudaMemcpyAsync(&nNodes, &c->bCtl.p->nNodes, sizeof(int), cudaMemcpyDeviceToHost, st));
        CU(c, cudaStreamSynchronize(st));
        if (nNodes < 1 || (size_t)nNodes >= maxNodes) return fail(c, RT_ERR_CUDA, "rt_build_bvh: node count out of range");
        CU(c, c->bHostNodes.reserve((size_t)nNodes, 0)); CU(c, c->bHostIdx.reserve((size_t)n, 0));
        c->bNodeCount = nNodes; c->bIdxCount = n;
        CU(c, cudaMemcpyAsync(c->bHostNodes.p, c->bNodes.p, (size_t)nNodes * sizeof(rt_flat_node), cudaMemcpyDeviceToHost, st));
        CU(c, cudaMemcpyAsync(c->bHostIdx.p, c->bIdx.p, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CU(c, cudaStreamSynchronize(st));
        cudaEventElapsedTime(&c->lastBuildMs, c->evBuild[0], c->evBuild[1]);
    }
    // hand the tree to the context exactly as an upload of these nodes and indices would
    if (c->hostTreeStale) { c->indices.clear(); }     // the host list belongs to an older tree
    const bool sameIdx = (size_t)c->bIdxCount == c->indices.size() &&
                         (c->bIdxCount == 0 || std::memcmp(c->bHostIdx.p, c->indices.data(), (size_t)c->bIdxCount * sizeof(int32_t)) == 0);
    if (!sameIdx) { c->indices.assign(c->bHostIdx.p, c->bHostIdx.p + c->bIdxCount); c->topoDirty = true; }
    int rc = set_nodes(c, c->bHostNodes.p, c->bNodeCount, &c->bHostNodes);
    if (rc != RT_OK) return rc;
    if (nNodes_out) *nNodes_out = c->bNodeCount;
    if (nIndices_out) *nIndices_out = c->bIdxCount;
    return RT_OK;
}

int rt_get_bvh(rt_ctx* c, void* nodes, int nodeCap, int32_t* indices, int indexCap) {
    if (!c) return RT_ERR_ARG;
    if (!c->haveScene) return fail(c, RT_ERR_STATE, "no scene");
    if (c->hostTreeStale) { CU(c, cudaSetDevice(c->device)); int rc = flush_scene(c); if (rc != RT_OK) return rc; rc = ensure_host_tree(c); if (rc != RT_OK) return rc; }
    if ((nodes && nodeCap < c->nNodes) || (indices && (size_t)indexCap < c->indices.size())) return fail(c, RT_ERR_ARG, "rt_get_bvh: buffer too small");
    if (c->hostNodesStale) {                        // K7 grew the boxes on the device
        CU(c, cudaSetDevice(c->device));
        int rc = flush_scene(c);
        if (rc != RT_OK) return rc;
        CU(c, cudaStreamSynchronize(c->stream));
        if (nodes && c->nNodes > 0) CU(c, cudaMemcpy(nodes, c->dNodes.p, (size_t)c->nNodes * sizeof(rt_flat_node), cudaMemcpyDeviceToHost));
    } else if (nodes && c->nNodes > 0) {
        wait_copies(c);
        std::memcpy(nodes, c->hNodes.p, (size_t)c->nNodes * sizeof(rt_flat_node));
    }
    if (indices && !c->indices.empty()) std::memcpy(indices, c->indices.data(), c->indices.size() * sizeof(int32_t));
    return RT_OK;
}
float rt_last_build_ms(const rt_ctx* c) { return c ? c->lastBuildMs : 0.f; }
int rt_layout_info(rt_ctx* c, int out[4]) {
    if (!c || !out) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    int rc = flush_scene(c);
    if (rc != RT_OK) return rc;
    out[0] = c->nDfs; out[1] = c->nDfsB; out[2] = c->nSub; out[3] = c->layoutOnDevice ? 1 : 0;
    return RT_OK;
}

int rt_untile(rt_ctx* c, const float* dev_shards, int world, float* dev_rgba_out) {
    if (!c || !dev_shards || !dev_rgba_out || world < 1) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    const int nT = tiles_x(c) * tiles_y(c);
    const long long shardPixels = (long long)((nT + world - 1) / world) * RT_TILE_W * RT_TILE_H;   // rank 0's (largest) shard
    const int blocks = (nT + RT_UNTILE_UNROLL - 1) / RT_UNTILE_UNROLL;
    untile_kernel<<<(unsigned)blocks, RT_TILE_W * RT_TILE_H, 0, c->stream>>>(reinterpret_cast<const float4*>(dev_shards),
        reinterpret_cast<float4*>(dev_rgba_out), c->W, c->H, tiles_x(c), nT, world, shardPixels);
    c->launches++;
    CU(c, cudaGetLastError());
    return RT_OK;
}

int rt_present(rt_ctx* c, const float* dev_rgba, uint8_t* dev_rgba8_out, int64_t pixels) {
    if (!c || !dev_rgba || !dev_rgba8_out || pixels < 0) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    if (pixels > 0) {
        present_kernel<<<(unsigned)((pixels + 256 * RT_PRESENT_UNROLL - 1) / (256 * RT_PRESENT_UNROLL)), 256, 0, c->stream>>>(reinterpret_cast<const float4*>(dev_rgba),
            reinterpret_cast<uchar4*>(dev_rgba8_out), (long long)pixels);
        c->launches++;
    }
    CU(c, cudaGetLastError());
    return RT_OK;
}

void* rt_stream(rt_ctx* c) { return c ? (void*)c->stream : nullptr; }
int rt_set_stream(rt_ctx* c, void* s) {
    if (!c) return RT_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    { const int rc = flush_deferred(c, false); if (rc != RT_OK) return rc; }
    CU(c, cudaStreamSynchronize(c->stream));
    c->copyPending = false;
    c->stream = s ? static_cast<cudaStream_t>(s) : c->ownStream;
    return RT_OK;
}
int rt_synchronize(rt_ctx* c) {
    if (!c) return RT_ERR_ARG;
    { const int rc = flush_deferred(c, false); if (rc != RT_OK) return rc; }
    CU(c, cudaStreamSynchronize(c->stream));
    if (c->framePending[0] || c->framePending[1]) {       // pipelined frames finish on the copy stream
        CU(c, cudaStreamSynchronize(c->copyStream));
        c->framePending[0] = c->framePending[1] = false;
    }
    return RT_OK;
}
float* rt_device_image(rt_ctx* c) { return c ? reinterpret_cast<float*>(c->lastImage ? c->lastImage : c->image.p) : nullptr; }

int rt_last_timings(rt_ctx* c, float out_ms[4]) {
    if (!c || !out_ms) return RT_ERR_ARG;
    { const int rc = flush_deferred(c, false); if (rc != RT_OK) return rc; }
    CU(c, cudaStreamSynchronize(c->stream));
    CU(c, cudaStreamSynchronize(c->copyStream));
    for (int i = 0; i < 4; ++i) {
        out_ms[i] = -1.0f;
        if (c->timed[i]) cudaEventElapsedTime(&out_ms[i], c->ev[2 * i], c->ev[2 * i + 1]);
    }
    return RT_OK;
}
int64_t rt_launch_count(const rt_ctx* c) { return c ? c->launches : 0; }
#ifdef RT_FAST_BUILD
const char* rt_build_info(void) { return "librt_b200_fast sm_100a fmad=ON (fast build: ids / colours may differ from the parity build) prec-div prec-sqrt tile=32x8 persistent-warps"; }
#else
const char* rt_build_info(void) { return "librt_b200 sm_100a fmad=off prec-div prec-sqrt tile=32x8 persistent-warps"; }
#endif


// ---- rt_build_bvh_static: K8's single-block build without any allocation -----------------------------------
// No counterpart in the reference; added for callers that may not call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  rt_build_bvh keeps
// its tables in DevBufs that grow on demand; this entry point keeps them in ONE file-scope __device__ array of a
// fixed size instead, so that it calls no cudaMalloc / cudaFree / allocating library routine at all.  Every GPU
// that loads the library holds the array (RT_STATIC_BUILD_SCRATCH_BYTES, about 0.5 MB) until the process exits.
//
// In host code `g_staticBuildScratch` is only the host-side stand-in of the symbol, NOT a device address: the
// address comes from cudaGetSymbolAddress, asked once per device with that device current (static_scratch_ptr),
// and that pointer is what the kernel gets.  All calls share the array and it keeps the last call's contents:
// the kernel's own set-up phase re-initialises every table entry it reads before writing (bld::setup_id: the box
// keys, the arrival counters, the child links and the control words; bld::init_position: order, node map and
// records), so nothing has to be cleared by the host — but two calls that overlap on the GPU (different streams,
// or two host threads) race on it.
}  // extern "C"

namespace {

constexpr size_t static_align(size_t v) { return (v + 255) & ~(size_t)255; }
constexpr size_t kSbN       = RT_STATIC_BUILD_MAX_SHAPES;
constexpr size_t kSbIds     = 2 * kSbN;                                           // build ids 0 .. 2n - 2
constexpr size_t kSbKeyOff  = 0;                                                  // u64  [6 * ids]
constexpr size_t kSbRecOff  = kSbKeyOff + static_align(6 * kSbIds * sizeof(bld::u64));   // float4 [3 * n]
constexpr size_t kSbIntOff  = kSbRecOff + static_align(3 * kSbN * sizeof(float4));       // int [9 * ids + 4 * n + n + 1 + 1]
constexpr size_t kSbIntCnt  = 9 * kSbIds + 4 * kSbN + kSbN + 1 + 1;
constexpr size_t kSbCtlOff  = kSbIntOff + static_align(kSbIntCnt * sizeof(int));         // bld::Ctl
constexpr size_t kSbFlagOff = kSbCtlOff + static_align(sizeof(bld::Ctl));                // unsigned char [n]
constexpr size_t kSbBytes   = kSbFlagOff + static_align(kSbN);
static_assert(kSbBytes == RT_STATIC_BUILD_SCRATCH_BYTES, "include/rt_api.h states the size of the static build scratch");
static_assert(RT_STATIC_BUILD_MAX_SHAPES <= RT_BLD_SMALL_MAX, "the single-block build covers RT_BLD_SMALL_MAX shapes");

static __device__ __align__(256) unsigned char g_staticBuildScratch[kSbBytes];

// the tables of one build inside the scratch array (fixed offsets: every table is sized for the maximum)
__device__ __forceinline__ BuildBufs static_build_bufs(unsigned char* s) {
    BuildBufs B;
    B.N.key = reinterpret_cast<bld::u64*>(s + kSbKeyOff);
    float4* rec = reinterpret_cast<float4*>(s + kSbRecOff);
    B.R.cen = rec; B.R.mn = rec + kSbN; B.R.mx = rec + 2 * kSbN;
    int* q = reinterpret_cast<int*>(s + kSbIntOff);
    B.N.start = q; q += kSbIds; B.N.count = q; q += kSbIds; B.N.left = q; q += kSbIds; B.N.parent = q; q += kSbIds;
    B.N.depth = q; q += kSbIds; B.N.desc = q; q += kSbIds; B.N.idx = q; q += kSbIds; B.N.first = q; q += kSbIds;
    B.N.arrived = reinterpret_cast<unsigned*>(q); q += kSbIds;
    B.ord[0] = q; q += kSbN; B.ord[1] = q; q += kSbN; B.pnode[0] = q; q += kSbN; B.pnode[1] = q; q += kSbN;
    B.scanL = q; q += kSbN + 1; B.bsum = q;
    B.ctl = reinterpret_cast<bld::Ctl*>(s + kSbCtlOff);
    B.flag = s + kSbFlagOff;
    return B;
}

// The variant of bld_small_kernel (rt_build.cuh): the same per-thread bodies (bld::*), phases and barriers, hence the
// same bytes; the tables are carved out of `scratch` here on the device, an empty shape list gives buildBVH's one empty
// leaf (main.cpp:1175-1193) and the counts land in the caller's device memory (no host round trip).
__global__ void __launch_bounds__(RT_BLD_SMALL_THREADS)
bld_small_static_kernel(unsigned char* __restrict__ scratch, const rt_flat_shape* __restrict__ shapes, int n, int maxDepth,
                        rt_flat_node* __restrict__ out, int32_t* __restrict__ indices, int32_t* __restrict__ counts) {
    __shared__ int warpSum[RT_BLD_SMALL_THREADS / 32];
    const int t = threadIdx.x, T = blockDim.x, lane = t & 31, w = t >> 5;
    if (n == 0) {
        if (t == 0) {
            rt_flat_node r;
            memset(&r, 0, sizeof r);
            for (int k = 0; k < 3; ++k) { r.boundsMin[k] = INFINITY; r.boundsMax[k] = -INFINITY; }
            r.leftChild = r.rightChild = -1;
            out[0] = r;
            counts[0] = 1; counts[1] = 0;
        }
        return;
    }
    const BuildBufs B = static_build_bufs(scratch);
    const int nIds = 2 * n;
    for (int id = t; id < nIds; id += T) bld::setup_id(id, B.N, B.ctl, n, maxDepth);      // clears what must start clean
    __syncthreads();
    for (int base = 0; base < n; base += T) {
        const int p = base + t;
        bld::u64 k[6];
        if (p < n) bld::init_position(p, shapes, B.R, B.ord[0], B.pnode[0], k);
        bld_fold_warp(B.N.key, p < n ? 0 : -1, k);
    }
    __syncthreads();
    const int chunk = (n + T - 1) / T, lo = min(n, t * chunk), hi = min(n, lo + chunk);
    int level = 0;
    for (; level < maxDepth; ++level) {
        if (!B.ctl->active[level]) break;
        const int cur = level & 1, nxt = cur ^ 1;
        int mine = 0;
        for (int p = lo; p < hi; ++p) { const int f = bld::classify(p, B.R, B.ord[cur], B.pnode[cur], B.N); B.flag[p] = (unsigned char)f; mine += f; }
        int incl = mine;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += v; }
        if (lane == 31) warpSum[w] = incl;
        __syncthreads();
        if (w == 0) {
            int v = warpSum[lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, off); if (lane >= off) v += u; }
            warpSum[lane] = v;
        }
        __syncthreads();
        int run = (w > 0 ? warpSum[w - 1] : 0) + incl - mine;
        for (int p = lo; p < hi; ++p) { B.scanL[p] = run; run += (int)B.flag[p]; }
        if (t == T - 1) B.scanL[n] = warpSum[31];
        __syncthreads();
        for (int base = 0; base < n; base += T) {
            const int p = base + t;
            bld::u64 k[6];
            int node = -1;
            if (p < n) {
                int newPos, newNode; bool moved;
                bld::split_position(p, (int)B.flag[p], level, B.pnode[cur], B.scanL, B.N, B.ctl, newPos, newNode, moved);
                const int shape = B.ord[cur][p];
                B.ord[nxt][newPos] = shape;
                B.pnode[nxt][newPos] = newNode;
                if (moved) { bld::record_keys(B.R.mn[shape], B.R.mx[shape], (unsigned)newPos, k); node = newNode; }
            }
            bld_fold_warp(B.N.key, node, k);
        }
        if (t == 0) B.ctl->parity = nxt;
        __syncthreads();
    }
    const int* ord = B.ord[B.ctl->parity];
    for (int lev = level; lev >= 0; --lev) {                            // descendant counts level by level, as in bld_small_kernel
        for (int id = t; id < nIds; id += T)
            if (B.N.left[id] != RT_BLD_ABSENT && maxDepth - B.N.depth[id] == lev) bld::desc_from_children(id, B.N);
        __syncthreads();
    }
    for (int id = t; id < nIds; id += T) bld::number_node(id, B.N, B.ctl);
    __syncthreads();
    for (int id = t; id < nIds; id += T) bld::emit_node(id, B.N, ord, out, indices);
    if (t == 0) { counts[0] = B.ctl->nNodes; counts[1] = n; }
}

// The device address of g_staticBuildScratch on the CURRENT device, resolved once per device.
constexpr int kSbMaxDevices = 64;
unsigned char* g_staticBuildPtr[kSbMaxDevices];     // zero-initialised; written once per device (same value if two threads race)

int static_fail(char* err, size_t cap, int code, const char* fmt, ...) {
    if (err && cap > 0) {
        va_list ap; va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

int static_scratch_ptr(unsigned char** out, char* err, size_t cap) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return static_fail(err, cap, RT_ERR_CUDA, "rt_build_bvh_static: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kSbMaxDevices) return static_fail(err, cap, RT_ERR_ARG, "rt_build_bvh_static: device %d out of range", dev);
    unsigned char* p = __atomic_load_n(&g_staticBuildPtr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_staticBuildScratch);
        if (e != cudaSuccess || !p)
            return static_fail(err, cap, RT_ERR_CUDA, "rt_build_bvh_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
        __atomic_store_n(&g_staticBuildPtr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_build_bvh_static(const void* dev_shapes, int nShapes, int maxDepth,
                        void* dev_nodes_out, int nodeCapacity, int32_t* dev_indices_out, int indexCapacity,
                        int32_t* dev_counts_out, void* cuda_stream, char* err, size_t errCapacity) {
    if (err && errCapacity > 0) err[0] = 0;
    if (nShapes < 0) return static_fail(err, errCapacity, RT_ERR_ARG, "rt_build_bvh_static: nShapes = %d", nShapes);
    if (nShapes > RT_STATIC_BUILD_MAX_SHAPES)
        return static_fail(err, errCapacity, RT_ERR_ARG,
                           "rt_build_bvh_static: %d shapes, the static scratch holds the tables of at most %d (RT_STATIC_BUILD_MAX_SHAPES); "
                           "use rt_build_bvh for larger scenes", nShapes, (int)RT_STATIC_BUILD_MAX_SHAPES);
    if (maxDepth < 0 || maxDepth > 64)
        return static_fail(err, errCapacity, RT_ERR_ARG, "rt_build_bvh_static: maxDepth = %d, the shader's traversal stack holds 64 entries", maxDepth);
    unsigned char* scratch = nullptr;
    const int rc = static_scratch_ptr(&scratch, err, errCapacity);
    if (rc != RT_OK) return rc;
    if (nShapes == 0 && !dev_shapes && !dev_nodes_out && !dev_indices_out && !dev_counts_out) return RT_OK;   // set-up only
    const int needNodes = nShapes > 0 ? 2 * nShapes - 1 : 1;
    if (!dev_nodes_out || !dev_counts_out || (nShapes > 0 && (!dev_shapes || !dev_indices_out)))
        return static_fail(err, errCapacity, RT_ERR_ARG, "rt_build_bvh_static: NULL input or output pointer");
    if (nodeCapacity < needNodes || indexCapacity < nShapes)
        return static_fail(err, errCapacity, RT_ERR_ARG,
                           "rt_build_bvh_static: %d shapes need room for %d nodes and %d indices (the tree's size is not known before the build), got %d and %d",
                           nShapes, needNodes, nShapes, nodeCapacity, indexCapacity);
    bld_small_static_kernel<<<1, RT_BLD_SMALL_THREADS, 0, static_cast<cudaStream_t>(cuda_stream)>>>(
        scratch, static_cast<const rt_flat_shape*>(dev_shapes), nShapes, maxDepth,
        static_cast<rt_flat_node*>(dev_nodes_out), dev_indices_out, dev_counts_out);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return static_fail(err, errCapacity, RT_ERR_CUDA, "rt_build_bvh_static: launch: %s", cudaGetErrorString(e));
    return RT_OK;
}

}  // extern "C"

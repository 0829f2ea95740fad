// rt_build.cuh — K8: the reference's BVH builder on the device ("rebuild").
//
// Upstream builds its tree on the host (buildBVH / split, src/main.cpp:1111-1193): a top-down midpoint
// split of the longest box axis, stable in the order of the shape list, stopped by a depth budget or as
// soon as one side of a split is empty; serializeBVH (:955-979) then numbers the nodes in the order the
// recursion pushed them (both sub-trees first, then the left child, then the right child, the root last)
// and concatenates the leaves' shape lists into bvhIndices.  A DIFFERENT tree would change the visit
// order of the shader's traversal and with it which of two equally distant shapes wins, so this builder
// does not invent a new tree: it emits the reference's FlatNode[] and bvhIndices[] BYTE FOR BYTE, level
// by level instead of recursively:
//
//   per level   classify  every shape of a pending node: centre[axis] < splitPos (the host's fp32 operations)
//               scan      exclusive prefix sum of the "left" flags over the whole position array
//               split     every position works out from the scan whether its node splits (both sides non-empty),
//                         where the stable partition puts it and which child (ids follow from the boundary
//                         position, no counter) owns it; the node's first position records the decision; every
//                         shape folds its points into its child's box with 64-bit atomicMin / atomicMax on
//                         (ordered value, sequence) keys: min / max are exact and the key's low word keeps "first
//                         in list order wins among equal values", i.e. the sign of a zero the host's
//                         `x < mn ? x : mn` would keep.  The keys ARE the boxes: readers decode them.
//   then        desc      descendants per node, leaves upwards (second arrival at a parent adds its children)
//               number    the reference's node index and startShapeIdx of every node from the sums along
//                         its path to the root
//               emit      FlatNode[] and bvhIndices[]
//
// No host round trip per level: every launch covers all n positions and a level at which nothing is pending
// exits at once (past log2(n) + 4 levels the host asks every 8th level whether anything is still pending).
// Scenes of up to RT_BLD_SMALL_MAX shapes run the whole build in one block (bld_small_kernel).  Everything that decides a byte of the output lives in the RT_DEV functions
// below; tests/emu/build_emu.cpp compiles exactly those with g++ and tests/test_build_emu.py compares
// their output with the host builder's on the CPU.  The __global__ wrappers add only indexing, the block
// scans and the warp- / block-aggregated form of the box atomics.
#pragma once
#include "rt_device.cuh"
#ifdef __CUDACC__
#include <cooperative_groups.h>
#endif

namespace rt {
namespace bld {

typedef unsigned long long u64;

#define RT_BLD_PENDING (-2)
#define RT_BLD_LEAF (-1)
#define RT_BLD_ABSENT (-3)

#ifdef __CUDACC__
RT_DEV void amin64(u64* p, u64 v) { atomicMin(p, v); }
RT_DEV void amax64(u64* p, u64 v) { atomicMax(p, v); }
RT_DEV unsigned aaddu(unsigned* p, unsigned v) { return atomicAdd(p, v); }
RT_DEV void fence() { __threadfence(); }
RT_DEV int  ld_int(const int* p) { return __ldcg(p); }
RT_DEV unsigned f2u(float f) { return __float_as_uint(f); }
RT_DEV float u2f(unsigned u) { return __uint_as_float(u); }
#else
// host emulation (tests only): one thread at a time
inline void amin64(u64* p, u64 v) { if (v < *p) *p = v; }
inline void amax64(u64* p, u64 v) { if (v > *p) *p = v; }
inline unsigned aaddu(unsigned* p, unsigned v) { unsigned o = *p; *p += v; return o; }
inline void fence() {}
inline int  ld_int(const int* p) { return *p; }
inline unsigned f2u(float f) { unsigned u; memcpy(&u, &f, 4); return u; }
inline float u2f(unsigned u) { float f; memcpy(&f, &u, 4); return f; }
#endif

// Build-time node table, indexed by build id (see setup_id): root = 0, the children of one node are (id, id + 1).
struct Nodes {
    int* start;        // first position of the node's segment in the position array
    int* count;        // shapes in the segment (= numShapes of the flat node, inner nodes too)
    int* left;         // RT_BLD_ABSENT, RT_BLD_PENDING, RT_BLD_LEAF or the build id of the left child
    int* parent;       // build id, -1 at the root
    int* depth;        // depth budget left (split's `depth` argument)
    u64* key;          // 6 per node: min x y z, max x y z as (ordered value << 32 | sequence)
    int* desc;         // number of descendants
    unsigned* arrived; // counters of the upward pass
    int* idx;          // index in the reference's node array
    int* first;        // startShapeIdx
};

// Control words on the device.
struct Ctl {
    int nNodes;        // number of nodes of the finished tree (written by the numbering pass)
    int parity;        // which of the two position arrays holds the final order
    int active[66];    // active[L]: some node is pending at level L
    // written by K9 (rt_build_list.cuh) for the host's decision whether the device-built list can be used as it is
    int maxLeaf;       // shapes in the largest leaf
    int notFinite;     // some node bound is NaN / inf / beyond RT_SAFE_COORD
    float rootBox[6];  // the root's box (min xyz, max xyz)
};

// ---- ordered keys -------------------------------------------------------------------------------------
RT_DEV unsigned ord32(float v) {              // monotone map of the non-NaN floats onto unsigned, -0 and +0 together
    unsigned u = f2u(v);
    if ((u << 1) == 0u) u = 0u;
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
RT_DEV float unord32(unsigned hi, unsigned neg) {
    unsigned u = (hi & 0x80000000u) ? (hi ^ 0x80000000u) : ~hi;
    if (u == 0u && neg) u = 0x80000000u;
    return u2f(u);
}
RT_DEV u64 key_min(float v, unsigned seq) { return ((u64)ord32(v) << 32) | (u64)((seq << 1) | (f2u(v) >> 31)); }
RT_DEV u64 key_max(float v, unsigned seq) { return ((u64)ord32(v) << 32) | (u64)(((0x7fffffffu - seq) << 1) | (f2u(v) >> 31)); }
#define RT_BLD_KEY_MIN_ID 0xff800000ffffffffull   /* +inf, last sequence */
#define RT_BLD_KEY_MAX_ID 0x007fffff00000000ull   /* -inf */
RT_DEV float key_value(u64 k) { return unord32((unsigned)(k >> 32), (unsigned)k & 1u); }

// ---- the shape side of BoundingBox::growToInclude and of split's centre --------------------------------
// The points a shape grows a box by, in the order the host grows them (BoundingBox.hpp:60-95): spheres
// centre + r then centre - r, walls start then the opposite corner (wall.hpp:16-31), triangles their
// vertices unless an x is not finite; infinite planes (and unknown types) none.
RT_DEV V3 wall_end(const rt_flat_shape* __restrict__ s) {
    const V3 n = mk(s->planeNormal), st = mk(s->wallStart);
    V3 t1;
    if (fabsf(n.x) > fabsf(n.y)) t1 = normalize3(mk(-n.z, 0.f, n.x));
    else t1 = normalize3(mk(0.f, -n.z, n.y));
    const V3 t2 = normalize3(cross3(n, t1));
    return (st + (s->wallWidth * t1)) + (s->wallHeight * t2);
}
RT_DEV int shape_points(const rt_flat_shape* __restrict__ s, V3 pt[3]) {
    const int type = s->type;
    if (type == RT_SHAPE_SPHERE) {
        const float r = s->sphereRadius;
        pt[0] = mk(s->sphereCenter[0] + r, s->sphereCenter[1] + r, s->sphereCenter[2] + r);
        pt[1] = mk(s->sphereCenter[0] - r, s->sphereCenter[1] - r, s->sphereCenter[2] - r);
        return 2;
    }
    if (type == RT_SHAPE_WALL) { pt[0] = mk(s->wallStart); pt[1] = wall_end(s); return 2; }
    if (type == RT_SHAPE_TRIANGLE) {
        if (!(isfinite(s->triP1[0]) && isfinite(s->triP2[0]) && isfinite(s->triP3[0]))) return 0;
        pt[0] = mk(s->triP1); pt[1] = mk(s->triP2); pt[2] = mk(s->triP3);
        return 3;
    }
    return 0;
}
RT_DEV float comp(V3 v, int a) { return a == 0 ? v.x : a == 1 ? v.y : v.z; }

// Per-shape record, computed once per build (the level loop never reads the 192-byte shapes again):
//   cen = split's centre of the shape (main.cpp:1128-1142)
//   mn / mx = per axis the value the shape contributes to a box minimum / maximum — the first of its points, in the
//   host's growing order, among those that compare equal — and in .w which point that was (2 bits per axis) and
//   whether the axis has a value at all (bit 6 + axis): NaN coordinates never change a bound (`NaN < mn` is false).
struct Records { float4* cen; float4* mn; float4* mx; };

RT_DEV void shape_record(const rt_flat_shape* __restrict__ s, float4& cen, float4& mn, float4& mx) {
    V3 pt[3];
    const int np = shape_points(s, pt);
    float lo[3] = {0.f, 0.f, 0.f}, hi[3] = {0.f, 0.f, 0.f};
    unsigned metaLo = 0u, metaHi = 0u;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        bool have = false;
        for (int j = 0; j < np; ++j) {
            const float v = comp(pt[j], a);
            if (v != v) continue;
            if (!have) { lo[a] = v; hi[a] = v; metaLo |= ((unsigned)j << (2 * a)) | (1u << (6 + a)); metaHi |= ((unsigned)j << (2 * a)) | (1u << (6 + a)); have = true; continue; }
            if (v < lo[a]) { lo[a] = v; metaLo = (metaLo & ~(3u << (2 * a))) | ((unsigned)j << (2 * a)); }
            if (hi[a] < v) { hi[a] = v; metaHi = (metaHi & ~(3u << (2 * a))) | ((unsigned)j << (2 * a)); }
        }
    }
    V3 c = mk(0.f, 0.f, 0.f);                            // upstream leaves a plane's centre uninitialised; 0 as in rt_host.cpp
    const int type = s->type;
    if (type == RT_SHAPE_SPHERE) c = mk(s->sphereCenter);
    else if (type == RT_SHAPE_WALL) c = (mk(s->wallStart) + wall_end(s)) * 0.5f;
    else if (type == RT_SHAPE_TRIANGLE) c = ((mk(s->triP1) + mk(s->triP2)) + mk(s->triP3)) / 3.0f;
    cen = make_float4(c.x, c.y, c.z, 0.f);
    mn = make_float4(lo[0], lo[1], lo[2], u2f(metaLo));
    mx = make_float4(hi[0], hi[1], hi[2], u2f(metaHi));
}

// The six keys of a shape standing at position `pos` (sequence = pos * 4 + point number); identity where it has none.
RT_DEV void record_keys(float4 mn, float4 mx, unsigned pos, u64 k[6]) {
    const unsigned mLo = f2u(mn.w), mHi = f2u(mx.w);
    const float lo[3] = {mn.x, mn.y, mn.z}, hi[3] = {mx.x, mx.y, mx.z};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        k[a]     = (mLo >> (6 + a)) & 1u ? key_min(lo[a], pos * 4u + ((mLo >> (2 * a)) & 3u)) : RT_BLD_KEY_MIN_ID;
        k[3 + a] = (mHi >> (6 + a)) & 1u ? key_max(hi[a], pos * 4u + ((mHi >> (2 * a)) & 3u)) : RT_BLD_KEY_MAX_ID;
    }
}

// split's side test (main.cpp:1119-1145): does a shape with this centre go to the left child of a node with this box?
RT_DEV bool goes_left(float4 cen, const float* box) {
    const float sx = box[3] - box[0], sy = box[4] - box[1], sz = box[5] - box[2];
    const int axis = sx > (sy < sz ? sz : sy) ? 0 : sy > sz ? 1 : 2;
    const float pos = (box[axis] + box[3 + axis]) * 0.5f;
    const float c = axis == 0 ? cen.x : axis == 1 ? cen.y : cen.z;
    return c < pos;
}

// ---- per-thread bodies of the level loop -------------------------------------------------------------------
// Build ids.  The root is 0.  A node whose segment [start, start + count) splits after nL shapes gets the children
// (2q - 1, 2q) with q = start + nL, the first position of the right child: a boundary between two positions is created
// at most once in a build, so the ids are unique without a counter, lie in [1, 2n - 2], siblings are neighbours, and
// every position of the node can compute its child's id on its own.  Ids nobody created stay RT_BLD_ABSENT.
RT_DEV void setup_id(int id, const Nodes& N, Ctl* ctl, int n, int maxDepth) {
    for (int a = 0; a < 3; ++a) { N.key[6 * (size_t)id + a] = RT_BLD_KEY_MIN_ID; N.key[6 * (size_t)id + 3 + a] = RT_BLD_KEY_MAX_ID; }
    N.arrived[id] = 0u;
    if (id != 0) { N.left[id] = RT_BLD_ABSENT; return; }
    N.start[0] = 0; N.count[0] = n; N.parent[0] = -1; N.depth[0] = maxDepth;
    N.left[0] = maxDepth > 0 ? RT_BLD_PENDING : RT_BLD_LEAF;
    ctl->nNodes = 1; ctl->parity = 0; ctl->maxLeaf = 0; ctl->notFinite = 0;
    for (int l = 0; l < 66; ++l) ctl->active[l] = 0;
    ctl->active[0] = maxDepth > 0 ? 1 : 0;
}

// position p of level 0: the identity order, everything in the root; the shape's record; its keys for the root box
RT_DEV void init_position(int p, const rt_flat_shape* __restrict__ shapes, const Records& R, int* ord,
                          int* pnode, u64 k[6]) {
    ord[p] = p; pnode[p] = 0;
    float4 cen, mn, mx;
    shape_record(shapes + p, cen, mn, mx);
    R.cen[p] = cen; R.mn[p] = mn; R.mx[p] = mx;
    record_keys(mn, mx, (unsigned)p, k);
}

RT_DEV void node_box(const Nodes& N, int node, float box[6]) {
#pragma unroll
    for (int a = 0; a < 6; ++a) box[a] = key_value(N.key[6 * (size_t)node + a]);
}

RT_DEV int classify(int p, const Records& R, const int* ord, const int* pnode, const Nodes& N) {
    const int node = pnode[p];
    if (N.left[node] != RT_BLD_PENDING) return 0;
    float box[6];
    node_box(N, node, box);
    return goes_left(R.cen[ord[p]], box) ? 1 : 0;
}

// scanL: exclusive prefix sums of the flags, n + 1 entries.  Position p of a pending node: does the node split (both
// sides non-empty, :1156-1161), where does p go, which node owns it then.  The node's first position also records
// the decision: a leaf, or the two children (whose keys are still the identity from setup_id; the shapes fold
// their bounds in during this very sweep).
RT_DEV void split_position(int p, int flag, int level, const int* pnode, const int* scanL, const Nodes& N, Ctl* ctl,
                           int& newPos, int& newNode, bool& moved) {
    const int node = pnode[p];
    newPos = p; newNode = node; moved = false;
    if (N.left[node] == RT_BLD_LEAF) return;             // decided at an earlier level (or by its first position just now)
    const int st = N.start[node], cnt = N.count[node];
    const int nL = scanL[st + cnt] - scanL[st];
    const bool first = p == st;
    if (nL == 0 || nL == cnt) { if (first) N.left[node] = RT_BLD_LEAF; return; }
    const int q = st + nL, l = 2 * q - 1;
    const int rankL = scanL[p] - scanL[st];
    if (flag) { newPos = st + rankL; newNode = l; }
    else { newPos = q + ((p - st) - rankL); newNode = l + 1; }
    moved = true;
    if (!first) return;
    const int d = N.depth[node] - 1;
    for (int s = 0; s < 2; ++s) {
        const int id = l + s;
        N.start[id] = s == 0 ? st : q;
        N.count[id] = s == 0 ? nL : cnt - nL;
        N.parent[id] = node;
        N.depth[id] = d;
        N.left[id] = d > 0 ? RT_BLD_PENDING : RT_BLD_LEAF;                 // split(child, 0) makes a leaf (:1113-1118)
    }
    N.left[node] = l;
    if (d > 0) ctl->active[level + 1] = 1;
}

RT_DEV void fold_keys(u64* nodeKeys, const u64 k[6]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (k[a] != RT_BLD_KEY_MIN_ID) amin64(nodeKeys + a, k[a]);
        if (k[3 + a] != RT_BLD_KEY_MAX_ID) amax64(nodeKeys + 3 + a, k[3 + a]);
    }
}

// ---- numbering (over all 2n ids; absent ones do nothing) -------------------------------------------------------
RT_DEV void count_descendants(int id, const Nodes& N) {
    if (N.left[id] != RT_BLD_LEAF) return;        // leaves start
    N.desc[id] = 0;
    for (int j = N.parent[id]; j >= 0; j = N.parent[j]) {
        fence();
        if (aaddu(N.arrived + j, 1u) == 0u) return;
        fence();
        const int l = N.left[j];
        N.desc[j] = ld_int(N.desc + l) + ld_int(N.desc + l + 1) + 2;
    }
}

// The same numbers level by level (one block, a barrier between the levels, deepest level first): a node's children are done.
RT_DEV void desc_from_children(int id, const Nodes& N) {
    const int l = N.left[id];
    N.desc[id] = l >= 0 ? N.desc[l] + N.desc[l + 1] + 2 : 0;
}

RT_DEV int leaf_shapes_below(const Nodes& N, int id) { return N.left[id] >= 0 ? N.count[id] : 0; }   // indices emitted by id's descendants

// The reference pushes, for a node P whose descendants start at node index b and whose indices start at e:
// [descendants of L] [descendants of R] L R.  Hence for a child: index and first bvhIndices entry follow from
// b(P), e(P), and b / e of a node follow from its ancestors: a right child starts after its left sibling's block.
RT_DEV void number_node(int id, const Nodes& N, Ctl* ctl) {
    if (N.left[id] == RT_BLD_ABSENT) return;
    if (id == 0) { N.idx[0] = N.desc[0]; N.first[0] = leaf_shapes_below(N, 0); ctl->nNodes = N.desc[0] + 1; return; }
    const int par = N.parent[id];
    int b = 0, e = 0;
    for (int a = par; a != 0; a = N.parent[a]) {
        const int pa = N.parent[a], sib = N.left[pa];
        if (a != sib) { b += N.desc[sib]; e += leaf_shapes_below(N, sib); }
    }
    const int l = N.left[par], r = l + 1;
    const int idxL = b + N.desc[l] + N.desc[r];
    const int firstL = e + leaf_shapes_below(N, l) + leaf_shapes_below(N, r);
    if (id == l) { N.idx[id] = idxL; N.first[id] = firstL; }
    else { N.idx[id] = idxL + 1; N.first[id] = firstL + (N.left[l] >= 0 ? 0 : N.count[l]); }
}

RT_DEV void emit_node(int id, const Nodes& N, const int* ord, rt_flat_node* out, int32_t* indices) {
    if (N.left[id] == RT_BLD_ABSENT) return;
    rt_flat_node f;
    float b[6];
    node_box(N, id, b);
    f.boundsMin[0] = b[0]; f.boundsMin[1] = b[1]; f.boundsMin[2] = b[2]; f.padding1 = 0.f;
    f.boundsMax[0] = b[3]; f.boundsMax[1] = b[4]; f.boundsMax[2] = b[5]; f.padding2 = 0.f;
    const int l = N.left[id];
    f.leftChild = l >= 0 ? N.idx[l] : -1;
    f.rightChild = l >= 0 ? N.idx[l + 1] : -1;
    f.startShapeIdx = N.first[id];
    f.numShapes = N.count[id];
    out[N.idx[id]] = f;
    if (l < 0) {
        const int st = N.start[id], cnt = N.count[id], o = N.first[id];
        for (int k = 0; k < cnt; ++k) indices[o + k] = ord[st + k];
    }
}

}  // namespace bld

#ifdef __CUDACC__
// ---- kernels ------------------------------------------------------------------------------------------------
#define RT_BLD_THREADS 256

struct BuildBufs {
    bld::Nodes N;
    bld::Records R;
    bld::Ctl* ctl;
    int* ord[2];
    int* pnode[2];
    unsigned char* flag;
    int* scanL;          // n + 1
    int* bsum;           // one per block of RT_BLD_THREADS positions (+1)
};

__global__ void bld_setup_kernel(BuildBufs B, int n, int maxDepth) {      // one thread per build id, 2n of them
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id < 2 * n) bld::setup_id(id, B.N, B.ctl, n, maxDepth);
}

// Warp-wide minimum (isMin) or maximum of 64-bit keys, two 32-bit REDUX steps: the high words, then the low words
// of the lanes that hold the winning high word.
__device__ __forceinline__ bld::u64 bld_warp_best(bld::u64 v, bool isMin) {
    const unsigned full = 0xffffffffu;
    const unsigned hi = (unsigned)(v >> 32), lo = (unsigned)v;
    const unsigned bestHi = isMin ? __reduce_min_sync(full, hi) : __reduce_max_sync(full, hi);
    const unsigned cand = hi == bestHi ? lo : (isMin ? 0xffffffffu : 0u);
    const unsigned bestLo = isMin ? __reduce_min_sync(full, cand) : __reduce_max_sync(full, cand);
    return ((bld::u64)bestHi << 32) | bestLo;
}

// The same for one warp (all 32 lanes call it): when its shapes all go to one pair of siblings the warp reduces each key
// first (REDUX) and issues one atomic per key and side; a mixed warp issues one atomic per shape and key.
__device__ __forceinline__ void bld_fold_warp(bld::u64* keyBase, int node, const bld::u64 k[6]) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int pair = node < 0 ? -1 : (node + 1) >> 1;
    const unsigned have = __ballot_sync(full, node >= 0);
    if (have == 0u) return;
    const int wpair = __shfl_sync(full, pair, __ffs(have) - 1);
    const bool warpUniform = __all_sync(full, node < 0 || pair == wpair);
    if (!warpUniform) { if (node >= 0) bld::fold_keys(keyBase + 6 * (size_t)node, k); return; }
    const int wleft = wpair == 0 ? 0 : 2 * wpair - 1;
#pragma unroll
    for (int side = 0; side < 2; ++side) {
        const bool mine = node == wleft + side;
        if (__ballot_sync(full, mine) == 0u) continue;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            const bool isMin = a < 3;
            const bld::u64 id = isMin ? RT_BLD_KEY_MIN_ID : RT_BLD_KEY_MAX_ID;
            const bld::u64 best = bld_warp_best(mine ? k[a] : id, isMin);
            if (lane == 0 && best != id) {
                if (isMin) bld::amin64(keyBase + 6 * (size_t)(wleft + side) + a, best);
                else bld::amax64(keyBase + 6 * (size_t)(wleft + side) + a, best);
            }
        }
    }
}

// Fold the keys k of this thread's shape into node `node` (node < 0: nothing to fold).  Called by every thread of
// the block.  While the segments are long, all the shapes of a warp — of the whole block — go to the same pair of
// siblings: then the warp reduces each key first (REDUX), the block combines its warps in shared memory, and one
// global atomic per key and side leaves the block (warp, when only the warp is uniform).  Mixed warps fall back to
// one atomic per shape and key.  agg == 0 forces that fallback everywhere (diagnostic).
__device__ __forceinline__ void bld_fold_block(bld::u64* keyBase, int node, const bld::u64 k[6], bool agg) {
    __shared__ bld::u64 sk[12];
    __shared__ int sPair;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int pair = node < 0 ? -1 : (node + 1) >> 1;                 // siblings (l, l + 1), l odd, share (l + 1) / 2; the root is pair 0
    if (!agg) { if (node >= 0) bld::fold_keys(keyBase + 6 * (size_t)node, k); return; }
    if (threadIdx.x == 0) sPair = -1;
    if (threadIdx.x < 12) sk[threadIdx.x] = threadIdx.x % 6 < 3 ? RT_BLD_KEY_MIN_ID : RT_BLD_KEY_MAX_ID;
    __syncthreads();
    if (node >= 0) sPair = pair;                                      // any one of them
    __syncthreads();
    const int pair0 = sPair;
    if (pair0 < 0) return;                                            // nothing to fold in this block (uniform)
    const bool blockUniform = __syncthreads_and(node < 0 || pair == pair0) != 0;
    const int left0 = pair0 == 0 ? 0 : 2 * pair0 - 1;                  // the pair's first node (the root stands alone)
    if (blockUniform) {
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const bool mine = node == left0 + side;
            if (__ballot_sync(full, mine) == 0u) continue;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const bool isMin = a < 3;
                const bld::u64 id = isMin ? RT_BLD_KEY_MIN_ID : RT_BLD_KEY_MAX_ID;
                const bld::u64 best = bld_warp_best(mine ? k[a] : id, isMin);
                if (lane == 0 && best != id) { if (isMin) atomicMin(&sk[side * 6 + a], best); else atomicMax(&sk[side * 6 + a], best); }
            }
        }
        __syncthreads();
        if (threadIdx.x < 12) {
            const int side = threadIdx.x / 6, a = threadIdx.x % 6;
            const bld::u64 v = sk[threadIdx.x];
            if (a < 3) { if (v != RT_BLD_KEY_MIN_ID) bld::amin64(keyBase + 6 * (size_t)(left0 + side) + a, v); }
            else       { if (v != RT_BLD_KEY_MAX_ID) bld::amax64(keyBase + 6 * (size_t)(left0 + side) + a, v); }
        }
        return;
    }
    bld_fold_warp(keyBase, node, k);
}

__global__ void bld_init_kernel(BuildBufs B, const rt_flat_shape* __restrict__ shapes, int n, int agg) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    bld::u64 k[6];
    int node = -1;
    if (p < n) { bld::init_position(p, shapes, B.R, B.ord[0], B.pnode[0], k); node = 0; }
    bld_fold_block(B.N.key, node, k, agg != 0);
}

// classify + block totals of the flags
__global__ void bld_classify_kernel(BuildBufs B, int n, int level) {
    if (!B.ctl->active[level]) return;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int cur = level & 1;
    int f = 0;
    if (p < n) { f = bld::classify(p, B.R, B.ord[cur], B.pnode[cur], B.N); B.flag[p] = (unsigned char)f; }
    const int total = __syncthreads_count(f);
    if (threadIdx.x == 0) B.bsum[blockIdx.x] = total;
}

// one block: exclusive scan of the block totals in place, grand total into scanL[n]
__global__ void bld_scan_blocks_kernel(BuildBufs B, int n, int nBlocks, int level) {
    if (!B.ctl->active[level]) return;
    __shared__ int sh[1024];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nBlocks; base += 1024) {
        const int i = base + (int)threadIdx.x;
        const int v = i < nBlocks ? B.bsum[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {                     // inclusive Hillis-Steele
            const int add = (int)threadIdx.x >= off ? sh[threadIdx.x - off] : 0;
            __syncthreads();
            sh[threadIdx.x] += add;
            __syncthreads();
        }
        const int incl = sh[threadIdx.x];
        const int c = carry;
        if (i < nBlocks) B.bsum[i] = c + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) B.scanL[n] = carry;
}

// scanL[p] = block offset + exclusive prefix of the flags inside the block
__global__ void bld_scan_apply_kernel(BuildBufs B, int n, int level) {
    if (!B.ctl->active[level]) return;
    __shared__ int warpTot[RT_BLD_THREADS / 32];
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int f = p < n ? (int)B.flag[p] : 0;
    const unsigned bal = __ballot_sync(0xffffffffu, f);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) warpTot[w] = __popc(bal);
    __syncthreads();
    int before = 0;
    for (int i = 0; i < w; ++i) before += warpTot[i];
    if (p < n) B.scanL[p] = B.bsum[blockIdx.x] + before + __popc(bal & ((1u << lane) - 1u));
}

// stable partition of every splitting node's segment + the node's decision (first position) + the children's boxes
__global__ void bld_split_kernel(BuildBufs B, int n, int level, int agg) {
    if (!B.ctl->active[level]) return;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int cur = level & 1, nxt = cur ^ 1;
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
    bld_fold_block(B.N.key, node, k, agg != 0);
    if (p == 0) B.ctl->parity = nxt;
}

__global__ void bld_desc_kernel(BuildBufs B, int nIds) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id < nIds) bld::count_descendants(id, B.N);
}
__global__ void bld_number_kernel(BuildBufs B, int nIds) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id < nIds) bld::number_node(id, B.N, B.ctl);
}
__global__ void bld_emit_kernel(BuildBufs B, int nIds, rt_flat_node* __restrict__ out, int32_t* __restrict__ indices) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id < nIds) bld::emit_node(id, B.N, B.ord[B.ctl->parity], out, indices);
}
// ---- small scenes: the whole build in ONE block -----------------------------------------------------------
// Up to RT_BLD_SMALL_MAX shapes (measured: 1240 shapes 0.28 vs 0.49 ms, 4122 shapes 0.80 vs 0.73 ms) the level loop is launch-bound (6 launches of a few microseconds per level), so one
// block of 1024 threads runs every phase itself, separated by __syncthreads(), and stops at the first level with
// nothing pending; the same per-thread bodies, the same tables in global memory (which this block alone touches).
#define RT_BLD_SMALL_MAX 2048
#define RT_BLD_SMALL_THREADS 1024

__global__ void __launch_bounds__(RT_BLD_SMALL_THREADS)
bld_small_kernel(BuildBufs B, const rt_flat_shape* __restrict__ shapes, int n, int maxDepth,
                 rt_flat_node* out, int32_t* indices) {
    __shared__ int warpSum[RT_BLD_SMALL_THREADS / 32];
    const int t = threadIdx.x, T = blockDim.x, lane = t & 31, w = t >> 5;
    const int nIds = 2 * n;
    for (int id = t; id < nIds; id += T) bld::setup_id(id, B.N, B.ctl, n, maxDepth);
    __syncthreads();
    for (int base = 0; base < n; base += T) {                          // whole warps stay together for the REDUX
        const int p = base + t;
        bld::u64 k[6];
        if (p < n) bld::init_position(p, shapes, B.R, B.ord[0], B.pnode[0], k);
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            const bool isMin = a < 3;
            const bld::u64 id = isMin ? RT_BLD_KEY_MIN_ID : RT_BLD_KEY_MAX_ID;
            const bld::u64 best = bld_warp_best(p < n ? k[a] : id, isMin);
            if (lane == 0 && best != id) { if (isMin) bld::amin64(B.N.key + a, best); else bld::amax64(B.N.key + a, best); }
        }
    }
    __syncthreads();
    const int chunk = (n + T - 1) / T, lo = min(n, t * chunk), hi = min(n, lo + chunk);
    int level = 0;
    for (; level < maxDepth; ++level) {
        if (!B.ctl->active[level]) break;                              // written before the last barrier: uniform
        const int cur = level & 1, nxt = cur ^ 1;
        // classify + exclusive scan of the flags: each thread owns the contiguous positions [lo, hi)
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
            warpSum[lane] = v;                                         // inclusive over the warps
        }
        __syncthreads();
        int run = (w > 0 ? warpSum[w - 1] : 0) + incl - mine;
        for (int p = lo; p < hi; ++p) { B.scanL[p] = run; run += (int)B.flag[p]; }
        if (t == T - 1) B.scanL[n] = warpSum[31];
        __syncthreads();
        for (int base = 0; base < n; base += T) {                      // whole warps stay together: the fold is warp-aggregated
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
            bld_fold_warp(B.N.key, node, k);        // (the per-shape atomics of round 1 were 7 k same-address atomics per level at the top)
        }
        if (t == 0) B.ctl->parity = nxt;
        __syncthreads();
    }
    const int* ord = B.ord[B.ctl->parity];
    // Descendant counts level by level, deepest first (the nodes of level `level`, made by the last sweep, are leaves): the
    // fenced second-arrival walk of the multi-block path (count_descendants) took 103 k of this kernel's 410 k cycles on the
    // monkey (clock64 per phase), this takes 22 k.  A node's level is maxDepth minus the depth budget it was made with.
    for (int lev = level; lev >= 0; --lev) {
        for (int id = t; id < nIds; id += T)
            if (B.N.left[id] != RT_BLD_ABSENT && maxDepth - B.N.depth[id] == lev) bld::desc_from_children(id, B.N);
        __syncthreads();
    }
    for (int id = t; id < nIds; id += T) bld::number_node(id, B.N, B.ctl);
    __syncthreads();
    for (int id = t; id < nIds; id += T) bld::emit_node(id, B.N, ord, out, indices);
}

// ---- mid-size scenes: the whole build in ONE THREAD-BLOCK CLUSTER ----------------------------------------------
// Between the single block (<= RT_BLD_SMALL_MAX shapes) and the one-launch-per-phase path lies the range where the latter is
// launch-bound (car-4122: ~105 launches of 4 - 5 us = 0.52 ms for 1 MB of work).  A cluster of up to 16 blocks of 1024 threads
// (sm_90+: the blocks of a cluster are co-scheduled and share a hardware barrier) runs the same phases with cluster.sync() —
// MEMBAR.GPU + the cluster barrier + an L1 invalidate, so the tables in global memory written by one block are what the others
// read after it — in place of the kernel boundaries: one or two positions per thread, the same per-thread bodies (bld::*), the block
// scans joined through one word per block, the descendant counts level by level as in bld_small_kernel.
// Measured on a B200 (tools/ncu_target_build.py, profiles/r2/cluster_build.txt): car-4122 0.536 -> 0.245 ms, 3 k / 6 k / 12 k / 24 k
// shapes 0.37 / 0.55 / 0.56 / 0.60 -> 0.23 / 0.24 / 0.28 / 0.44 ms; 39 k: 0.60 against 0.49, from there on the per-phase launches win
// (a thread would own several positions per sweep); up to 2048 shapes the single block wins (monkey 0.156 against 0.17 - 0.19).
#ifndef RT_BLD_CLUSTER_MAX_BLOCKS
#define RT_BLD_CLUSTER_MAX_BLOCKS 16
#endif
#ifndef RT_BLD_CLUSTER_MAX
#define RT_BLD_CLUSTER_MAX 28000
#endif

__global__ void __launch_bounds__(RT_BLD_SMALL_THREADS)
bld_cluster_kernel(BuildBufs B, const rt_flat_shape* __restrict__ shapes, int n, int maxDepth,
                   rt_flat_node* out, int32_t* indices) {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    __shared__ int warpSum[RT_BLD_SMALL_THREADS / 32];
    const int t = threadIdx.x, T = blockDim.x, lane = t & 31, w = t >> 5;
    const int rank = (int)cl.block_rank(), C = (int)cl.num_blocks();
    const int G = C * T, g = rank * T + t;                             // the cluster is the whole grid
    const int nIds = 2 * n;
    for (int id = g; id < nIds; id += G) bld::setup_id(id, B.N, B.ctl, n, maxDepth);
    cl.sync();
    for (int base = 0; base < n; base += G) {                          // whole warps stay together for the REDUX
        const int p = base + g;
        bld::u64 k[6];
        int node = -1;
        if (p < n) { bld::init_position(p, shapes, B.R, B.ord[0], B.pnode[0], k); node = 0; }
        bld_fold_warp(B.N.key, node, k);
    }
    cl.sync();
    const int chunk = (n + G - 1) / G, lo = min(n, g * chunk), hi = min(n, lo + chunk);   // positions [lo, hi) of this thread: usually one
    int level = 0;
    for (; level < maxDepth; ++level) {
        if (!B.ctl->active[level]) break;                              // written before the last cluster barrier: the same for every thread
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
            warpSum[lane] = v;                                         // inclusive over the warps of this block
            if (lane == 31) B.bsum[rank] = v;                          // this block's total, for the blocks after it
        }
        cl.sync();
        int run = (w > 0 ? warpSum[w - 1] : 0) + incl - mine;
        for (int r = 0; r < rank; ++r) run += B.bsum[r];
        for (int p = lo; p < hi; ++p) { B.scanL[p] = run; run += (int)B.flag[p]; }
        if (g == G - 1) B.scanL[n] = run;                              // the last thread's running sum is the grand total
        cl.sync();
        for (int base = 0; base < n; base += G) {
            const int p = base + g;
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
        if (g == 0) B.ctl->parity = nxt;
        cl.sync();
    }
    const int* ord = B.ord[B.ctl->parity];
    for (int lev = level; lev >= 0; --lev) {                            // descendant counts, deepest level first
        for (int id = g; id < nIds; id += G)
            if (B.N.left[id] != RT_BLD_ABSENT && maxDepth - B.N.depth[id] == lev) bld::desc_from_children(id, B.N);
        cl.sync();
    }
    for (int id = g; id < nIds; id += G) bld::number_node(id, B.N, B.ctl);
    cl.sync();
    for (int id = g; id < nIds; id += G) bld::emit_node(id, B.N, ord, out, indices);
}

// ---- large scenes: the whole build in ONE COOPERATIVE GRID ------------------------------------------------------
// Beyond the cluster's reach the per-phase path pays ~100 launches too (soup-1M: 26 levels x 4 launches of 4 - 5 us = 0.5 of
// its 2.5 ms; 100 k shapes: 0.7 ms, mostly launches).  One block of 1024 threads per SM, launched cooperatively, runs the same
// phases with grid.sync() (block barrier + MEMBAR.GPU + an atomic barrier word + L1 invalidate) where the kernel boundaries
// were.  Every block owns a CONTIGUOUS range of positions for the classify / scan phase and sweeps it in tiles of 1024
// (coalesced, a block scan per tile with a running carry); the block totals are joined by one warp per block; the move / fold
// sweep is strided over the whole grid as in the other paths.  Same per-thread bodies (bld::*), same tree.
// Measured (profiles/r2/cluster_build.txt): 60 k shapes 0.65 -> 0.53 ms, 100 k 0.73 -> 0.59 ms; at 300 k the per-phase launches are
// ahead again (0.93 against 1.12 ms, soup-1M 2.5 against 3.7: 1024-thread blocks, one per SM, hide less latency than the
// per-phase kernels' 256-thread blocks at full occupancy), so the cooperative launch is taken up to RT_BLD_GRID_MAX shapes.
#ifndef RT_BLD_GRID_MAX
#define RT_BLD_GRID_MAX 120000
#endif
__global__ void __launch_bounds__(RT_BLD_SMALL_THREADS)
bld_grid_kernel(BuildBufs B, const rt_flat_shape* __restrict__ shapes, int n, int maxDepth,
                rt_flat_node* out, int32_t* indices) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    __shared__ int warpSum[RT_BLD_SMALL_THREADS / 32];
    __shared__ int blockBase;
    const int t = threadIdx.x, T = blockDim.x, lane = t & 31, w = t >> 5;
    const int rank = (int)blockIdx.x, C = (int)gridDim.x;
    const int G = C * T, g = rank * T + t;
    const int nIds = 2 * n;
    for (int id = g; id < nIds; id += G) bld::setup_id(id, B.N, B.ctl, n, maxDepth);
    grid.sync();
    for (int base = 0; base < n; base += G) {                          // whole warps stay together for the REDUX
        const int p = base + g;
        bld::u64 k[6];
        int node = -1;
        if (p < n) { bld::init_position(p, shapes, B.R, B.ord[0], B.pnode[0], k); node = 0; }
        bld_fold_warp(B.N.key, node, k);
    }
    grid.sync();
    const int per = (((n + C - 1) / C) + T - 1) / T * T;               // positions per block, a whole number of tiles
    const int r0 = min(n, rank * per), r1 = min(n, r0 + per);
    int level = 0;
    for (; level < maxDepth; ++level) {
        if (!B.ctl->active[level]) break;                              // written before the last grid barrier: the same for every thread
        const int cur = level & 1, nxt = cur ^ 1;
        int carry = 0;                                                 // flags set in this block's range before the current tile
        for (int tile = r0; tile < r1; tile += T) {
            const int p = tile + t;
            int f = 0;
            if (p < r1) { f = bld::classify(p, B.R, B.ord[cur], B.pnode[cur], B.N); B.flag[p] = (unsigned char)f; }
            int incl = f;
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
            if (p < r1) B.scanL[p] = carry + (w > 0 ? warpSum[w - 1] : 0) + incl - f;   // exclusive within the block's range
            carry += warpSum[31];
            __syncthreads();                                           // warpSum is reused by the next tile
        }
        if (t == 0) B.bsum[rank] = carry;
        grid.sync();
        if (w == 0) {                                                  // flags set in the blocks before this one
            int s = 0;
            for (int i = lane; i < rank; i += 32) s += B.bsum[i];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
            if (lane == 0) blockBase = s;
        }
        __syncthreads();
        const int base = blockBase;
        for (int p = r0 + t; p < r1; p += T) B.scanL[p] += base;
        if (rank == C - 1 && t == 0) B.scanL[n] = base + carry;        // the grand total
        grid.sync();
        for (int b0 = 0; b0 < n; b0 += G) {
            const int p = b0 + g;
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
        if (g == 0) B.ctl->parity = nxt;
        grid.sync();
    }
    const int* ord = B.ord[B.ctl->parity];
    for (int lev = level; lev >= 0; --lev) {                            // descendant counts, deepest level first
        for (int id = g; id < nIds; id += G)
            if (B.N.left[id] != RT_BLD_ABSENT && maxDepth - B.N.depth[id] == lev) bld::desc_from_children(id, B.N);
        grid.sync();
    }
    for (int id = g; id < nIds; id += G) bld::number_node(id, B.N, B.ctl);
    grid.sync();
    for (int id = g; id < nIds; id += G) bld::emit_node(id, B.N, ord, out, indices);
}
#endif  // __CUDACC__


}  // namespace rt

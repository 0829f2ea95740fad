// rt_anim.cuh — K6 / K7: the reference's per-frame animation step on the device.
//
// Upstream, an animated frame is: move the animated shapes on the host (bounceSphere main.cpp:1079-1082,
// updateWheelAnimations :1084-1109), re-serialize each one and push it with its own 192-byte
// glBufferSubData (:981-992), grow the boxes of every BVH node that holds an animated shape
// (updateBVH :1068-1077, grow-only) and re-upload all nodes (:342-344).  Here the shapes and nodes stay
// on the device:
//   K6 anim_kernel   one thread per animated shape; the per-entry scalars (the bounced y; the 12 entries of the
//                    frame's GLM wheel transform, from cos/sin of the host's libm) are computed on the host
//                    and passed in, the per-vertex mat4 * vec4 runs here in GLM's operation order — the
//                    result is bit-identical to the reference's own updateWheelAnimations compiled in place
//                    (oracle/_ref/libref_bvh.so; tests/test_gpu_parity.py, tests/test_ref_builder.py).
//                    Wheels keep their stale planeNormal / planeD, as upstream does (:1101-1106).
//   K7 grow_kernel   grow-only refit of the raw reference nodes: one thread per leaf grows its box by the
//                    leaf's animated shapes (BoundingBox::growToInclude, BoundingBox.hpp:60-95) and walks
//                    towards the root; the second thread to arrive at an inner node joins its children
//                    into it.  For a tree whose parents contain their children (the reference builder's)
//                    this equals updateBVH, which grows every node by the animated shapes below it.
// K4 (our own boxes) and K2 (repack) follow as after an upload.
#pragma once
#include "rt_kernels.cuh"

namespace rt {

enum { RT_ANIM_BOUNCE = 0, RT_ANIM_ROTATE_Z = 1 };
// bounce: m[0] = new centre y; rotate: m[0..11] = rows 0..2 of the frame's GLM wheel transform, column-major m[3 * col + row]
// (translate(center) * rotate(angle, axis) * translate(-center), computed on the host by host/rt_anim_math.h)
struct AnimEntry { int kind; float m[12]; };
struct AnimItem  { int shape; int entry; };

// Up to RT_ANIM_PACK entries travel as a kernel PARAMETER: a frame of the animated loop then issues no copy-engine
// operation at all (a small H2D copy or memset in the render stream queues behind the previous frame's 7.7 MB D2H copy
// on the copy engines of this box, and the whole next frame with it; see DESIGN: the pipelined monkey loop).
#define RT_ANIM_PACK 8
struct AnimPack { AnimEntry e[RT_ANIM_PACK]; };

__device__ __forceinline__ void anim_one(rt_flat_shape* shapes, const AnimItem it, const AnimEntry* __restrict__ entries) {
    const AnimEntry e = entries[it.entry];
    rt_flat_shape* s = shapes + it.shape;
    if (e.kind == RT_ANIM_BOUNCE) {
        s->sphereCenter[1] = e.m[0];
    } else {
        // vec3(transform * vec4(p, 1)) in GLM's operation order (type_mat4x4.inl:595-613; rta_apply in host/rt_anim_math.h):
        // (m[0] * x + m[1] * y) + (m[2] * z + m[3] * 1); compiled with -fmad=false, so bit-identical to the host's GLM path
        float* P[3] = {s->triP1, s->triP2, s->triP3};
#pragma unroll
        for (int v = 0; v < 3; ++v) {
            const float x = P[v][0], y = P[v][1], z = P[v][2];
#pragma unroll
            for (int r = 0; r < 3; ++r) P[v][r] = (e.m[r] * x + e.m[3 + r] * y) + (e.m[6 + r] * z + e.m[9 + r] * 1.0f);
        }
    }
}
__device__ __forceinline__ void anim_body(rt_flat_shape* __restrict__ shapes, const AnimItem* __restrict__ items, int nItems,
                                          const AnimEntry* __restrict__ entries) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nItems) anim_one(shapes, items[i], entries);
}
__global__ void anim_kernel(rt_flat_shape* __restrict__ shapes, const AnimItem* __restrict__ items, int nItems,
                            const AnimEntry* __restrict__ entries) {
    anim_body(shapes, items, nItems, entries);
}
__global__ void anim_kernel_packed(rt_flat_shape* __restrict__ shapes, const AnimItem* __restrict__ items, int nItems,
                                   const __grid_constant__ AnimPack pack) {
    anim_body(shapes, items, nItems, pack.e);
}

struct GBox { float mn[3], mx[3]; };
__device__ __forceinline__ void g_grow(GBox& b, float x, float y, float z) {       // glm::min / glm::max select order
    b.mn[0] = x < b.mn[0] ? x : b.mn[0]; b.mn[1] = y < b.mn[1] ? y : b.mn[1]; b.mn[2] = z < b.mn[2] ? z : b.mn[2];
    b.mx[0] = b.mx[0] < x ? x : b.mx[0]; b.mx[1] = b.mx[1] < y ? y : b.mx[1]; b.mx[2] = b.mx[2] < z ? z : b.mx[2];
}
// BoundingBox::growToInclude(Shape*) (BoundingBox.hpp:60-95): spheres by centre +- radius, walls by start and
// the opposite corner (wall.hpp:16-31), triangles by their vertices; infinite planes are ignored.
__device__ __forceinline__ void g_grow_shape(GBox& b, const rt_flat_shape* s) {
    const int type = s->type;
    if (type == RT_SHAPE_SPHERE) {
        const float r = s->sphereRadius;
        g_grow(b, s->sphereCenter[0] + r, s->sphereCenter[1] + r, s->sphereCenter[2] + r);
        g_grow(b, s->sphereCenter[0] - r, s->sphereCenter[1] - r, s->sphereCenter[2] - r);
    } else if (type == RT_SHAPE_WALL) {
        const V3 n = mk(s->planeNormal), st = mk(s->wallStart);
        V3 t1;
        if (fabsf(n.x) > fabsf(n.y)) t1 = normalize3(mk(-n.z, 0.f, n.x));
        else t1 = normalize3(mk(0.f, -n.z, n.y));
        const V3 t2 = normalize3(cross3(n, t1));
        const V3 e = (st + (s->wallWidth * t1)) + (s->wallHeight * t2);
        g_grow(b, st.x, st.y, st.z); g_grow(b, e.x, e.y, e.z);
    } else if (type == RT_SHAPE_TRIANGLE) {
        if (isfinite(s->triP1[0]) && isfinite(s->triP2[0]) && isfinite(s->triP3[0])) {
            g_grow(b, s->triP1[0], s->triP1[1], s->triP1[2]);
            g_grow(b, s->triP2[0], s->triP2[1], s->triP2[2]);
            g_grow(b, s->triP3[0], s->triP3[1], s->triP3[2]);
        }
    }
}
__device__ __forceinline__ GBox g_load(const rt_flat_node* n) {
    const float4 a = __ldcg(reinterpret_cast<const float4*>(n)), c = __ldcg(reinterpret_cast<const float4*>(n) + 1);
    GBox b; b.mn[0] = a.x; b.mn[1] = a.y; b.mn[2] = a.z; b.mx[0] = c.x; b.mx[1] = c.y; b.mx[2] = c.z;
    return b;
}
__device__ __forceinline__ void g_store(rt_flat_node* n, const GBox& b) {
    float* p = reinterpret_cast<float*>(n);
    __stcg(p + 0, b.mn[0]); __stcg(p + 1, b.mn[1]); __stcg(p + 2, b.mn[2]);
    __stcg(p + 4, b.mx[0]); __stcg(p + 5, b.mx[1]); __stcg(p + 6, b.mx[2]);
}

// leaves[i]: raw index of leaf i; parent[n]: raw index of n's parent (-1 at the root); arrived: counters that are zero at
// launch and zero again at the end (the second arrival resets its counter: no memset per frame)
__global__ void grow_kernel(rt_flat_node* __restrict__ nodes, const rt_flat_shape* __restrict__ shapes,
                            const int* __restrict__ indices, const unsigned char* __restrict__ animated,
                            const int* __restrict__ leaves, int nLeaves, const int* __restrict__ parent,
                            unsigned* __restrict__ arrived) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nLeaves) return;
    int n = leaves[i];
    {
        GBox b = g_load(nodes + n);
        const int s0 = nodes[n].startShapeIdx, cnt = nodes[n].numShapes;
        bool any = false;
        for (int q = 0; q < cnt; ++q) {
            const int s = indices[s0 + q];
            if (animated[s]) { g_grow_shape(b, shapes + s); any = true; }
        }
        if (any) g_store(nodes + n, b);
    }
    for (int j = parent[n]; j >= 0; j = parent[j]) {
        __threadfence();
        if (atomicAdd(arrived + j, 1u) == 0u) return;
        arrived[j] = 0u;                           // both children are here: nobody touches this counter again in this launch
        __threadfence();
        GBox b = g_load(nodes + j);
        const GBox l = g_load(nodes + nodes[j].leftChild), r = g_load(nodes + nodes[j].rightChild);
#pragma unroll
        for (int a = 0; a < 3; ++a) {            // union (an empty child, min > max, changes nothing)
            b.mn[a] = l.mn[a] < b.mn[a] ? l.mn[a] : b.mn[a]; b.mn[a] = r.mn[a] < b.mn[a] ? r.mn[a] : b.mn[a];
            b.mx[a] = b.mx[a] < l.mx[a] ? l.mx[a] : b.mx[a]; b.mx[a] = b.mx[a] < r.mx[a] ? r.mx[a] : b.mx[a];
        }
        g_store(nodes + j, b);
    }
}

// K6 + K7 in ONE block for the usual case of a few animated shapes (round 2).  Only the nodes that hold an animated shape can
// change (updateBVH grows a node by ITS animated shapes, main.cpp:1068-1077), i.e. the animated leaves and their ancestors;
// the host lists them level by level at registration time (level 0: the animated leaves; level h: the inner nodes whose
// longest way down to an animated leaf has h edges; upload_anim_tables) and this block walks the levels between barriers:
// no atomics, no fences, a chain of `levels` dependent box loads.  grow_kernel (every leaf climbs, the second arrival at a
// node joins its children) took 32 us on the monkey's 1861 nodes for three bouncing spheres; this takes ~5.
// sched[0 .. nLevels]: entry offsets of the levels; ent[e] = {node, left child, right child, -}.
#define RT_GROW_PATH_MAX 16384        // more listed nodes than this: grow_kernel
#define RT_GROW_PATH_THREADS 256
#ifndef RT_LEVELS_MAX
#define RT_LEVELS_MAX 1023          // more levels than this: the general kernels
#endif
__global__ void __launch_bounds__(RT_GROW_PATH_THREADS)
anim_grow_path_kernel(rt_flat_shape* shapes, const AnimItem* __restrict__ items, int nItems, const AnimEntry* __restrict__ entries,
                      const __grid_constant__ AnimPack pack, rt_flat_node* nodes, const int* __restrict__ indices,
                      const unsigned char* __restrict__ animated, const int* __restrict__ sched, const int4* __restrict__ ent, int nLevels) {
    __shared__ int sOff[RT_LEVELS_MAX + 1];
    for (int i = threadIdx.x; i <= nLevels; i += blockDim.x) sOff[i] = sched[i];
    const AnimEntry* en = entries ? entries : pack.e;
    for (int i = threadIdx.x; i < nItems; i += blockDim.x) anim_one(shapes, items[i], en);
    __syncthreads();                                 // the block's own writes to the shapes are visible to it from here on
    // one dependent trip to L2 per level: the entry of the NEXT level is fetched before the barrier that ends this one
    int4 cur = make_int4(0, 0, 0, 0);
    if (nLevels > 0 && sOff[0] + (int)threadIdx.x < sOff[1]) cur = ent[sOff[0] + threadIdx.x];
    for (int lev = 0; lev < nLevels; ++lev) {
        const int e0 = sOff[lev], e1 = sOff[lev + 1];
        int4 nxt = make_int4(0, 0, 0, 0);
        if (lev + 1 < nLevels && e1 + (int)threadIdx.x < sOff[lev + 2]) nxt = ent[e1 + threadIdx.x];
        for (int e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
            const int4 x = e == e0 + (int)threadIdx.x ? cur : ent[e];
            GBox b = g_load(nodes + x.x);
            if (lev == 0) {
                const int s0 = nodes[x.x].startShapeIdx, cnt = nodes[x.x].numShapes;
                for (int q = 0; q < cnt; ++q) {
                    const int s = indices[s0 + q];
                    if (animated[s]) g_grow_shape(b, shapes + s);
                }
            } else {
                const GBox l = g_load(nodes + x.y), r = g_load(nodes + x.z);
#pragma unroll
                for (int a = 0; a < 3; ++a) {        // as grow_kernel
                    b.mn[a] = l.mn[a] < b.mn[a] ? l.mn[a] : b.mn[a]; b.mn[a] = r.mn[a] < b.mn[a] ? r.mn[a] : b.mn[a];
                    b.mx[a] = b.mx[a] < l.mx[a] ? l.mx[a] : b.mx[a]; b.mx[a] = b.mx[a] < r.mx[a] ? r.mx[a] : b.mx[a];
                }
            }
            g_store(nodes + x.x, b);
        }
        __syncthreads();
        cur = nxt;
    }
}

}  // namespace rt

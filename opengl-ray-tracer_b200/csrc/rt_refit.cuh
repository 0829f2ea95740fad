// rt_refit.cuh — K4: device-side refit of the conservative boxes.
//
// When shapes move under an unchanged topology (the reference's animation path: per-shape
// glBufferSubData + updateBVH + node re-upload, src/main.cpp:340-344, :981-992) the boxes that are
// OURS have to follow: the fat-leaf sub-tree boxes (unions of rtl::prim_bound, rt_layout.h) and the
// boxes of list B's top nodes (exact unions of the reference's leaf boxes).  The host used to do
// this (rtl::refit_subboxes / refit_top) before every upload; here one launch does it on the
// freshly uploaded device copies: one thread per sub-tree leaf (primitive bounds in double, the same
// rtl::prim_bound source compiled for the device) and per raw child of a top node, each then walks
// towards the root — the second thread to arrive at a node (atomic counter) computes its union.
// A primitive that stopped being boundable gets an all-covering box (the walk then tests it
// always: still exact, just slower) and raises `degraded` so that the host rebuilds the layout at
// its next upload without ever synchronising on it.
#pragma once
#include "rt_kernels.cuh"
#include "rt_layout.h"

namespace rt {

__device__ __forceinline__ rtl::Box rf_box(const float4* __restrict__ subBox, const rt_flat_node* __restrict__ raw, int ref) {
    float4 mn, mx;
    if (ref >= 0) { mn = __ldcg(subBox + 2 * (size_t)ref); mx = __ldcg(subBox + 2 * (size_t)ref + 1); }   // written in this launch: read at L2
    else { const float4* s = reinterpret_cast<const float4*>(raw + (-1 - ref)); mn = __ldg(s); mx = __ldg(s + 1); }
    rtl::Box b;
    b.mn[0] = mn.x; b.mn[1] = mn.y; b.mn[2] = mn.z; b.mx[0] = mx.x; b.mx[1] = mx.y; b.mx[2] = mx.z;
    return b;
}

__device__ __forceinline__ void rf_store(float4* __restrict__ subBox, int j, const rtl::Box& b) {
    __stcg(subBox + 2 * (size_t)j,     make_float4(b.mn[0], b.mn[1], b.mn[2], 0.f));
    __stcg(subBox + 2 * (size_t)j + 1, make_float4(b.mx[0], b.mx[1], b.mx[2], 0.f));
}

// start[i] >= 0: sub-tree leaf entry; < 0: a raw reference node arriving at top entry -1-start[i]
__global__ void refit_kernel(const rt_flat_shape* __restrict__ shapes, const rt_flat_node* __restrict__ raw,
                             const int* __restrict__ pshape, const int* __restrict__ start, int nStart,
                             const int* __restrict__ parent, const int2* __restrict__ child, const int2* __restrict__ leaf,
                             unsigned* __restrict__ arrived, float4* __restrict__ subBox, double omax, int* __restrict__ degraded) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nStart) return;
    int j = start[i];
    if (j >= 0) {
        const int2 lf = leaf[j];
        rtl::Box b; b.reset();
        for (int q = 0; q < lf.y; ++q) {
            rtl::Box pb;
            if (!rtl::prim_bound(shapes[pshape[lf.x + q]], omax, pb)) { pb.everything(); *degraded = 1; }
            b.grow(pb);
        }
        rf_store(subBox, j, b);
        j = parent[j];
    } else {
        j = -1 - j;
    }
    while (j >= 0) {
        __threadfence();
        if (atomicAdd(arrived + j, 1u) == 0u) return;      // the sibling is not there yet: it will do the union
        arrived[j] = 0u;                                   // zero again for the next launch (no memset per frame)
        __threadfence();
        const int2 ch = child[j];
        rtl::Box b; b.reset();
        b.grow(rf_box(subBox, raw, ch.x));
        b.grow(rf_box(subBox, raw, ch.y));
        rf_store(subBox, j, b);
        j = parent[j];
    }
}

}  // namespace rt

// rt_device.cuh — device-side data layout, exact-arithmetic helpers, primitive
// tests and BVH traversal for the sm_100a ray-tracing kernel.
//
// Written from the behavioural spec in SURVEY.md §8(a) (what gpu_shader.comp
// computes), not translated from the GLSL: the scene is repacked into float4
// SoA in the reference's *visit order*, the BVH becomes a stackless threaded
// list, per-primitive ray-independent terms are precomputed once at upload, and
// shadow rays are any-hit.  Every hit/miss DECISION is still made with exactly
// the IEEE fp32 operations, in exactly the order, of the shader (compiled with
// -fmad=false), so closest-hit shape ids are bit-identical to the oracle.
#pragma once
#include <stdint.h>
#ifdef __CUDACC__
#include <cuda_runtime.h>
#define RT_DEV __device__ __forceinline__
#else
// Host compilation (g++ -ffp-contract=off) is used ONLY by tests/emu: a scalar emulation of the
// walk that checks the layout builder and the conservative sub-tree bounds without a GPU.  The
// product library is always built by nvcc and has no host execution path.
#include <math.h>
#include <vector_types.h>
#include <vector_functions.h>
#define RT_DEV inline
#endif

#include "../../include/rt_api.h"

namespace rt {

// ---------------------------------------------------------------------------
// Device scene (built by the repack kernels, K2)
// ---------------------------------------------------------------------------
// Threaded BVH: nodes in the reference's DFS order (root, right subtree, left
// subtree — gpu_shader.comp:424-425 pushes left then right and pops LIFO).
//   n[2k]   = (boundsMin.xyz, bits(link))   inner: index of the node that follows
//                                           this subtree;  leaf: first prim slot
//   n[2k+1] = (boundsMax.xyz, bits(count))  inner: -1;     leaf: numShapes >= 0
// A box hit on an inner node continues at k+1; a miss jumps to link.
// Leaves' primitives are laid out contiguously in the same visit order, so
// "first improvement wins on ties" (strict <, :411) is preserved by construction.
//
// Primitive slot s (SoA, 16-byte loads):
//   type 0 sphere   q0=(center, r*r)
//   type 1 plane    q0=(n, D)
//   type 2 wall     q0=(n, D) q1=(start, width) q2=(u, height) q3=(v, 0)
//   type 3 triangle q0=(n, D) q1=(P1, d00)      q2=(e1, d01)   q3=(e2, d11)
struct DeviceScene {
    const float4* __restrict__ nodes;   // 2 per node: list A (reference tree order, fat leaves expanded)
    const float4* __restrict__ nodesB;  // list B: own SAH tree over the reference leaves (rt_layout.h build_top), or NULL
    int nNodesB;
    const float4* __restrict__ q0;
    const float4* __restrict__ q1;
    const float4* __restrict__ q2;
    const float4* __restrict__ q3;
    const int*    __restrict__ ptype;   // per slot
    const int*    __restrict__ pshape;  // per slot: original shape index
    const int*    __restrict__ prank;   // per slot: position in the reference's primitive visit sequence
    const rt_flat_shape* __restrict__ shapes;   // raw std430 shapes (materials, brute-force path)
    int nNodes;
    int nPrims;
    int nShapes;
    int boxesFinite;      // every node bound is finite: the NaN-free slab fast path is legal
    float omax;           // sub-tree boxes are conservative for ray origins with |o|_inf <= omax
};

// Leaf "kind" tag kept in bits 28..30 of a leaf's count word (host-computed at upload):
// all primitives of the leaf have one type, so the per-primitive type load is skipped.
// Bit 30 marks a leaf of one of OUR sub-trees (rt_layout.h): its box is a conservative bound, not a
// reference box.  Inner nodes: count word -1 = reference node, -2 = sub-tree node.
#define RT_LEAF_COUNT_MASK 0x0fffffff
#define RT_LEAF_KIND_SHIFT 28
#define RT_LEAF_SUB_BIT (1 << 30)
#define RT_INNER_REF (-1)
#define RT_INNER_SUB (-2)
enum { RT_LEAF_MIXED = 0, RT_LEAF_TRIANGLES = 1, RT_LEAF_SPHERES = 2 };

struct FrameParams {
    float camP[3], camF[3], camU[3], camR[3];
    float planeH, planeW;      // imagePlaneHeight / Width (tan hoisted to the host)
    float lightPos[3], lightCol[3];
    float resW, resH;          // uniform screenRes
    int   maxBounces, useBVH, useFresnel, useMT;
    int   W, H;                // image size in pixels
    int   tilesX, tilesY;      // tile grid
    int   shardRank, shardWorld;
    int   nTilesMine;          // tiles of this rank handled by this launch ...
    int   tileBegin;           // ... starting at this index of the rank's tile list (banded launches)
    int   writeIds;
    // Heavy-first scheduling (LPT with last frame's measured item costs, see lpt_*_kernel): NULL = off
    unsigned* itemCost;        // [items] SM cycles >> 4 of every warp item, written by this launch
    const unsigned* heavyList; // items whose last cost exceeded RT_HEAVY_FACTOR x mean, claimed first
    const unsigned char* heavyFlag;
    // control block of the heavy-first scheduler, 8 words: [0], [1] number of heavy items of even / odd frames, [2] claim counter
    // of the heavy list, [3] claim counter of the plain queue, [4..5], [6..7] 64-bit sums of this launch's item costs (even / odd)
    unsigned* lptCtl;
    int   lptPar;              // parity of this frame: which count / sum slot it uses
    int   frameTarget;         // world > 1: colours go row-major into a full W*H frame (rank 0's, possibly peer memory)
    // signalled fused gather: the last thread block of the launch adds 1 to *doneWord (the completion counter in rank 0's memory)
    unsigned* doneWord;        // NULL = off
    unsigned* ctaDone;         // local counter of finished thread blocks (the last block resets it)
    unsigned  doneSeq;
};

// ---------------------------------------------------------------------------
// Exact helpers (GLSL-spec definitions, listed in DESIGN.md)
// ---------------------------------------------------------------------------
struct V3 { float x, y, z; };
RT_DEV V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
RT_DEV V3 mk(const float* p) { return mk(p[0], p[1], p[2]); }
RT_DEV V3 xyz(float4 q) { return mk(q.x, q.y, q.z); }
RT_DEV V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_DEV V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_DEV V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
RT_DEV V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
RT_DEV V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
RT_DEV V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
RT_DEV V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
RT_DEV float gmin(float x, float y) { return y < x ? y : x; }
RT_DEV float gmax(float x, float y) { return x < y ? y : x; }
RT_DEV float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RT_DEV float length3(V3 v) { return sqrtf(dot3(v, v)); }
RT_DEV float distance3(V3 a, V3 b) { return length3(a - b); }
RT_DEV V3 normalize3(V3 v) { float inv = 1.0f / sqrtf(dot3(v, v)); return v * inv; }
RT_DEV V3 cross3(V3 a, V3 b) {
    return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
RT_DEV V3 reflect3(V3 I, V3 N) { float d = dot3(N, I); return I - N * d * 2.0f; }
RT_DEV float mix1(float x, float y, float a) { return x * (1.0f - a) + y * a; }

struct Ray {
    V3 o, d;
    V3 inv;        // 1/d, hoisted out of the per-node slab test (gpu_shader.comp:365)
    float aa;      // dot(d,d), hoisted out of the sphere test (:250)
    bool fast;     // origin and 1/d all finite: (bound - o) * inv can never be NaN for finite bounds
    bool sub;      // fast and |o|_inf <= omax: the sub-tree boxes may cull for this ray
};
RT_DEV bool bounded3(V3 v, float lim) {      // false for NaN / inf
    return fabsf(v.x) <= lim && fabsf(v.y) <= lim && fabsf(v.z) <= lim;
}
#define RT_SAFE_COORD 1e37f   // |bound|,|origin| <= 1e37: (bound - origin) cannot overflow to inf
RT_DEV Ray make_ray(V3 o, V3 d, bool boxesFinite, float omax) {
    Ray r; r.o = o; r.d = d;
    r.inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    r.aa = dot3(d, d);
    r.fast = boxesFinite && bounded3(o, RT_SAFE_COORD) && bounded3(r.inv, 3.402823466e38f);
    r.sub = r.fast && bounded3(o, omax);
    return r;
}

// One primitive's precomputed record.
struct Prim { float4 q0, q1, q2, q3; int type; };

// Ray-independent terms of a raw std430 shape, with the shader's own operation
// order (gpu_shader.comp:218-229 edges/dots, :305-307 wall basis, :252 r*r).
RT_DEV Prim prim_from_shape(const rt_flat_shape* __restrict__ s) {
    Prim p;
    p.type = s->type;
    p.q0 = p.q1 = p.q2 = p.q3 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.type == RT_SHAPE_SPHERE) {
        float r = s->sphereRadius;
        p.q0 = make_float4(s->sphereCenter[0], s->sphereCenter[1], s->sphereCenter[2], r * r);
    } else if (p.type == RT_SHAPE_PLANE || p.type == RT_SHAPE_WALL || p.type == RT_SHAPE_TRIANGLE) {
        V3 n = mk(s->planeNormal);
        p.q0 = make_float4(n.x, n.y, n.z, s->planeD);
        if (p.type == RT_SHAPE_WALL) {
            V3 u = normalize3(cross3(n, mk(0.f, 1.f, 0.f)));
            if (length3(u) < 1e-5f) u = normalize3(cross3(n, mk(1.f, 0.f, 0.f)));   // false for NaN: kept
            V3 v = normalize3(cross3(n, u));
            p.q1 = make_float4(s->wallStart[0], s->wallStart[1], s->wallStart[2], s->wallWidth);
            p.q2 = make_float4(u.x, u.y, u.z, s->wallHeight);
            p.q3 = make_float4(v.x, v.y, v.z, 0.f);
        } else if (p.type == RT_SHAPE_TRIANGLE) {
            V3 p1 = mk(s->triP1);
            V3 e1 = mk(s->triP2) - p1;
            V3 e2 = mk(s->triP3) - p1;
            p.q1 = make_float4(p1.x, p1.y, p1.z, dot3(e1, e1));
            p.q2 = make_float4(e1.x, e1.y, e1.z, dot3(e1, e2));
            p.q3 = make_float4(e2.x, e2.y, e2.z, dot3(e2, e2));
        }
    }
    return p;
}

// ---------------------------------------------------------------------------
// Primitive tests.  Return true for an INNER hit and set hp (OUTER == no hit for
// every caller in the shader).  a5-a9 of SURVEY.md §8(a).
// ---------------------------------------------------------------------------
RT_DEV bool plane_stage(float4 q0, const Ray& r, V3& hp) {
    V3 n = xyz(q0);
    float np = dot3(n, r.d);
    // INNER needs np > 0 and t > 0 (np == 0 -> NONE, np < 0 -> OUTER/NONE, NaN -> NONE).
    if (!(np > 0.0f)) return false;
    // With np > 0 the quotient t = num/np can only be > 0 if num > 0: the IEEE division is
    // skipped for every primitive behind the ray without changing any decision.
    float num = -(q0.w + dot3(n, r.o));
    if (!(num > 0.0f)) return false;
    float t = num / np;
    if (!(t > 0.0f)) return false;         // underflow to 0
    hp = r.o + t * r.d;
    return true;
}

RT_DEV bool hit_sphere(float4 q0, const Ray& r, V3& hp) {
    V3 oc = r.o - xyz(q0);
    float bb = 2.0f * dot3(r.d, oc);
    float cc = dot3(oc, oc) - q0.w;
    float D = bb * bb - 4.0f * r.aa * cc;
    if (!(D > 0.0f)) return false;
    float sD = sqrtf(D);
    float t1 = (-bb - sD) / (2.0f * r.aa);
    if (!(t1 > 0.0f)) return false;        // t2 > 0 would be OUTER
    hp = r.o + t1 * r.d;
    return true;
}

RT_DEV bool wall_stage2(float4 q1, float4 q2, float4 q3, V3 hp) {
    V3 local = hp - xyz(q1);
    float uProj = dot3(local, xyz(q2));
    float vProj = dot3(local, xyz(q3));
    return !(uProj < 0.0f || uProj > q1.w || vProj < 0.0f || vProj > q2.w);
}

RT_DEV bool bary_stage2(float4 q1, float4 q2, float4 q3, V3 hp) {
    V3 toPoint = hp - xyz(q1);
    V3 e1 = xyz(q2), e2 = xyz(q3);
    float d00 = q1.w, d01 = q2.w, d11 = q3.w;
    float d20 = dot3(toPoint, e1);
    float d21 = dot3(toPoint, e2);
    float denom = d00 * d11 - d01 * d01;
    float X = d11 * d20 - d01 * d21;
    float Y = d00 * d21 - d01 * d20;
    // Sufficient, exact rejects before the two IEEE divisions: for 0 < denom < 1e30 and
    // X < -1e-12 the quotient X/denom is a non-zero negative float (|X/denom| > 1e-42), so
    // v < 0 holds in the shader as well; same for w.  Everything else takes the full path.
    if (denom > 0.0f && denom < 1e30f && (X < -1e-12f || Y < -1e-12f)) return false;
    float v = X / denom;
    float w = Y / denom;
    float u = 1.0f - v - w;
    return !(u < 0.0f || v < 0.0f || w < 0.0f);
}

RT_DEV bool hit_tri_mt(float4 q1, float4 q2, float4 q3, const Ray& r, V3& hp) {
    V3 e1 = xyz(q2), e2 = xyz(q3);
    V3 h = cross3(r.d, e2);
    float a = dot3(e1, h);
    if (fabsf(a) < 1e-5f) return false;
    float f = 1.0f / a;
    V3 s = r.o - xyz(q1);
    float u = f * dot3(s, h);
    if (u < 0.0f || u > 1.0f) return false;
    V3 q = cross3(s, e1);
    float v = f * dot3(r.d, q);
    if (v < 0.0f || u + v > 1.0f) return false;
    float t = f * dot3(e2, q);
    if (!(t > 0.0f)) return false;
    hp = r.o + t * r.d;
    return true;
}

// ---------------------------------------------------------------------------
// Box tests (a4) and the per-type dispatch
// ---------------------------------------------------------------------------
// Slab test with the shader's select semantics for min/max (a4, :364-377).
RT_DEV bool slab(const Ray& r, float4 a, float4 b) {
    float t0x = (a.x - r.o.x) * r.inv.x, t1x = (b.x - r.o.x) * r.inv.x;
    float t0y = (a.y - r.o.y) * r.inv.y, t1y = (b.y - r.o.y) * r.inv.y;
    float t0z = (a.z - r.o.z) * r.inv.z, t1z = (b.z - r.o.z) * r.inv.z;
    float tMin = gmax(gmax(gmin(t0x, t1x), gmin(t0y, t1y)), gmin(t0z, t1z));
    float tMax = gmin(gmin(gmax(t0x, t1x), gmax(t0y, t1y)), gmax(t0z, t1z));
    return tMax >= tMin && tMax > 0.0f;
}
// Same arithmetic with single-instruction FMNMX min/max.  Legal only when no operand can be
// NaN (Ray::fast): then fminf/fmaxf and the GLSL selects return the same value up to the sign
// of zero, which the two comparisons below cannot see.
RT_DEV bool slab_fast(const Ray& r, float4 a, float4 b) {
    float t0x = (a.x - r.o.x) * r.inv.x, t1x = (b.x - r.o.x) * r.inv.x;
    float t0y = (a.y - r.o.y) * r.inv.y, t1y = (b.y - r.o.y) * r.inv.y;
    float t0z = (a.z - r.o.z) * r.inv.z, t1z = (b.z - r.o.z) * r.inv.z;
    float tMin = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    float tMax = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    return tMax >= tMin && tMax > 0.0f;
}

template <bool MT>
RT_DEV bool test_prim(const Prim& p, const Ray& r, V3& hp) {
    if (p.type == RT_SHAPE_TRIANGLE) {
        if (MT) return hit_tri_mt(p.q1, p.q2, p.q3, r, hp);
        return plane_stage(p.q0, r, hp) && bary_stage2(p.q1, p.q2, p.q3, hp);
    }
    if (p.type == RT_SHAPE_SPHERE) return hit_sphere(p.q0, r, hp);
    if (p.type == RT_SHAPE_WALL) return plane_stage(p.q0, r, hp) && wall_stage2(p.q1, p.q2, p.q3, hp);
    if (p.type == RT_SHAPE_PLANE) return plane_stage(p.q0, r, hp);
    return false;
}

}  // namespace rt

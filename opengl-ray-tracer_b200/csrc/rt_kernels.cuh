// rt_kernels.cuh — K1 (render), K2 (repack), K3 (untile), K4 (present).
#pragma once
#include "rt_device.cuh"

#ifndef RT_BLOCK_THREADS
#define RT_BLOCK_THREADS 128
#endif
#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 8          // resident CTAs per SM the register allocation aims for
#endif
// A scene whose walked list does not stay in L1 (the soup: 63 MB of nodes) waits for L2 / HBM, not for issue slots: there
// 10 resident CTAs of 48 registers beat 8 of 64 in spite of their spills (soup-1M at 4K 27.5 -> 26.1 ms, 300 k shapes 10.3 ->
// 9.9, 60 k 2.72 -> 2.67; 11 / 12 CTAs and 9: slower; the car, issue-bound, loses 5 %: profiles/r2/occupancy_sweep.txt).
// The second instantiation is chosen at launch for lists of RT_BIG_SCENE_NODES nodes and more.
#ifndef RT_MIN_BLOCKS_BIG
#define RT_MIN_BLOCKS_BIG 10
#endif
#ifndef RT_FEW_BLOCKS_PER_SM
#define RT_FEW_BLOCKS_PER_SM 6      // resident blocks per SM for launches with fewer than 8 items per warp (rt_api.cu: gridCap); 0 = as many as fit.
                                    // Such a launch is bound by its slowest block's chain of dependent walks, and an iteration of that chain also waits
                                    // for the issue slots of the other warps of its scheduler: 6 blocks instead of 8 give the eighth of the car's 4K frame
                                    // 0.265 -> 0.260 ms, 800x600 0.272 -> 0.259 ms, monkey flat; 7: 0.255 / 0.268; 5 and 4 lose throughput
                                    // (profiles/r2/few_blocks_sweep.txt; frames bit-identical)
#endif
#ifndef RT_BIG_SCENE_NODES
#define RT_BIG_SCENE_NODES 100000
#endif

namespace rt {

// ---------------------------------------------------------------------------
// Traversal
// ---------------------------------------------------------------------------
struct Hit { float dist; int slot; V3 p; };

// Optional work counters (stats launches only; the timed kernel is compiled without them).
// [0] rays (BVH/scan queries) [1] primary [2] shadow [3] reflection [4] nodes visited
// [5] primitive tests [6] closest-hit queries that hit [7] shadow queries that were occluded
// [8..11] primary segments, [12..15] later segments: packet-mode warp iterations, lane-mode warp
// iterations, lanes served in packet mode, lanes served in lane mode (walk efficiency diagnostics)
#define RT_NTALLY 20       // [16] max cycles of one warp item, [17] sum of item cycles, [18] items, [19] kernel cycles (max over warps)
struct Tally { unsigned long long v[RT_NTALLY]; };
#define RT_TALLY(i, n) do { if (COUNT) tl.v[i] += (n); } while (0)

// One BVH query of one lane.  Closest-hit (a3): best starts at 1e20 and every candidate with
// dist < best replaces it (strict '<', first in visit order wins ties, :410-411).  Shadow
// (a12): the shader runs a full closest-hit query and then tests hitDist < lightDist; "some
// INNER hit has dist < 1e20 and dist < lightDist" is the same boolean (the closest hit has
// the minimum dist), so best starts at min(1e20, lightDist) and the first candidate below it
// ends the query.
struct Query {
    Ray r;
    float best;      // current bound: candidates need dist < best
    int slot;        // slot of the accepted candidate, -1 = none
    V3 p;            // its hit point
#if RT_TIMELINE
    unsigned iters;  // diagnostic build: walk iterations of this warp so far
#endif
};
struct Cand { float dist; int slot; V3 p; };

// Candidate acceptance.  The reference keeps the FIRST candidate in its visit order among equal
// distances (strict '<', :410-411).  Slots of our sub-trees are not in that order, so equal
// distances are resolved by the slot's rank in the reference sequence (S.prank) instead.
__device__ __forceinline__ void offer(const DeviceScene& S, Query& q, int s, V3 hp, bool anyMode) {
    const float dist = distance3(q.r.o, hp);
    bool take = dist < q.best;
    if (!anyMode && dist == q.best && q.slot >= 0) take = __ldg(S.prank + s) < __ldg(S.prank + q.slot);
    if (take) { q.best = dist; q.slot = s; q.p = hp; }
}

// Triangle leaf, Barycentric mode.  The plane stage of four consecutive slots is evaluated as
// four independent dependency chains (the loads are issued together; the slot arrays are
// padded so reading past `end` is safe); the survivors then go through the division and the
// barycentric stage in slot order (unrolled: the rolled body with its index selects was 8 % slower
// and spilled more).  Returns true when an any-hit query is satisfied.
__device__ __forceinline__ bool leaf_triangles_bary(const DeviceScene& S, Query& q, int s0, int end, bool anyMode) {
    const Ray& r = q.r;
    for (int s = s0; s < end; s += 4) {
        float np[4], num[4];
        unsigned pm = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 qq = __ldg(S.q0 + s + i);
            const V3 n = xyz(qq);
            np[i] = dot3(n, r.d);
            num[i] = -(qq.w + dot3(n, r.o));
            if ((np[i] > 0.0f) && (num[i] > 0.0f) && (s + i < end)) pm |= 1u << i;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (pm & (1u << i)) {
                const float t = num[i] / np[i];
                if (t > 0.0f) {
                    const V3 hp = r.o + t * r.d;
                    if (bary_stage2(__ldg(S.q1 + s + i), __ldg(S.q2 + s + i), __ldg(S.q3 + s + i), hp)) {
                        const int before = q.slot;
                        offer(S, q, s + i, hp, anyMode);
                        if (anyMode && q.slot != before) return true;
                    }
                }
            }
        }
    }
    return false;
}

// Everything that is not a pure Barycentric-triangle leaf (spheres, walls, planes, mixed leaves,
// Moller-Trumbore mode) goes through one out-of-line routine so that the hot loop stays small.
// By-value in, by-value out: nothing of the caller's state is forced into local memory.
template <bool MT>
__device__ __noinline__ Cand leaf_generic(const float4* __restrict__ q0, const float4* __restrict__ q1,
                                          const float4* __restrict__ q2, const float4* __restrict__ q3,
                                          const int* __restrict__ ptype, const int* __restrict__ prank, Ray r, int s0, int end,
                                          float best, int bestSlot, int anyMode) {
    Cand c; c.dist = best; c.slot = bestSlot; c.p = mk(0.f, 0.f, 0.f);
    for (int s = s0; s < end; ++s) {
        const int type = __ldg(ptype + s);
        V3 hp;
        bool hit = false;
        if (type == RT_SHAPE_TRIANGLE) {
            if (MT) hit = hit_tri_mt(__ldg(q1 + s), __ldg(q2 + s), __ldg(q3 + s), r, hp);
            else hit = plane_stage(__ldg(q0 + s), r, hp) && bary_stage2(__ldg(q1 + s), __ldg(q2 + s), __ldg(q3 + s), hp);
        } else if (type == RT_SHAPE_SPHERE) {
            hit = hit_sphere(__ldg(q0 + s), r, hp);
        } else if (type == RT_SHAPE_WALL) {
            hit = plane_stage(__ldg(q0 + s), r, hp) && wall_stage2(__ldg(q1 + s), __ldg(q2 + s), __ldg(q3 + s), hp);
        } else if (type == RT_SHAPE_PLANE) {
            hit = plane_stage(__ldg(q0 + s), r, hp);
        }
        if (hit) {
            const float dist = distance3(r.o, hp);
            bool take = dist < c.dist;
            if (!anyMode && dist == c.dist && c.slot >= 0) take = __ldg(prank + s) < __ldg(prank + c.slot);
            if (take) { c.dist = dist; c.slot = s; c.p = hp; if (anyMode) break; }
        }
    }
    return c;
}

// Warp-synchronous ("packet") walk of the threaded BVH.  Every lane visits exactly the
// reference's node set for ITS ray in the reference's order (gpu_shader.comp:390-427),
// but the warp advances together: each iteration handles the smallest pending node index
// over the lanes (one REDUX), lanes whose next node is further along wait.  Because the
// list is in visit order every lane's index only grows, so the warp walks the UNION of
// the lanes' node sets once, all node and primitive loads are warp-uniform (one L1
// wavefront), and a fat leaf is looped once for all lanes that hit its box instead of once
// per lane at unrelated iterations.  Must be called by all 32 lanes; `active` masks lanes.
#define RT_FULL 0xffffffffu
#define RT_DONE 0x7fffffff
#ifndef RT_LOW_RUNS
#define RT_LOW_RUNS 2
#endif
#ifndef RT_LANE_SHIFT
#define RT_LANE_SHIFT 2     // switch to lane mode when an iteration serves fewer than 1/2^shift of the query's lanes
#endif
#ifndef RT_SAMPLE
#define RT_SAMPLE 4         // power of two
#endif

template <bool MT, bool COUNT>
__device__ __forceinline__ void walk(const DeviceScene& S, Query& q, bool active, bool anyMode, Tally& tl, int tbase) {
    const Ray& r = q.r;
    int k = active ? 0 : RT_DONE;
    // Rays that cannot use the NaN-free fast slab test (non-finite origin or 1/d) or for which the
    // sub-tree margins do not hold (|o| > omax) are rare: one warp-uniform flag keeps their handling
    // out of the hot loop (r.sub implies r.fast).
    const bool anyOdd = __any_sync(RT_FULL, active && !r.sub);
    // Two modes share this loop and the per-lane cursor k.  PACKET: the warp handles the smallest
    // pending node (uniform loads, the union of the lanes' node sets is walked once) — best for
    // coherent rays.  LANE: every lane handles its own node (divergent loads, no union) — best for
    // incoherent rays, where the union approaches the whole tree.  The walk starts in packet mode;
    // every RT_SAMPLE-th iteration it checks how many lanes the iteration serves and switches for
    // the rest of the query after RT_LOW_RUNS consecutive samples below a quarter of the query's lanes.
    int pm = -1;                          // -1: packet mode, 0: lane mode (a mask: see kcmp below)
    int low = 0, it = 0;
#ifndef RT_LANE_MIN
#define RT_LANE_MIN 3
#endif
    // ... but at least RT_LANE_MIN lanes: a SPARSE query (a few surviving bounce lanes) whose lanes take turns at the warp's
    // smallest node serves one lane per iteration in packet mode and would never fall below a quarter of itself (round 2: the
    // soup's later segments ran 34 M packet iterations at 1.28 lanes each; soup 4K 29.3 -> 27.5 ms, monkey 800x600 0.185 -> 0.176)
    int quarter = (__popc(__ballot_sync(RT_FULL, active)) + (1 << RT_LANE_SHIFT) - 1) >> RT_LANE_SHIFT;
    if (quarter < RT_LANE_MIN) quarter = RT_LANE_MIN;
    // List B (own tree over the reference leaves) is valid for rays with finite origin and 1/d inside
    // omax; a warp with any other ray walks list A, the reference's own tree order.
    const bool useB = S.nNodesB > 0 && !anyOdd;
    int n = useB ? S.nNodesB : S.nNodes;
    const char* __restrict__ nbase = reinterpret_cast<const char*>(useB ? S.nodesB : S.nodes);
    // keep the selected list (base and length) in registers: without these barriers ptxas re-selects list A / B from the
    // constant bank in every iteration and leaves the mode-sampling instructions predicated in the loop (~8 of ~58
    // instructions per iteration; car 4K 1.295 -> 1.214 ms, profiles/r2/design_choices.json)
    asm volatile("" : "+l"(nbase));
    asm volatile("" : "+r"(n));
    for (;;) {
        const int kmin = __reduce_min_sync(RT_FULL, k);
        if (kmin >= n) break;
        const int kcmp = (kmin & pm) | (k & ~pm);       // packet: the warp's node; lane: the lane's own node
        const bool mine = (k == kcmp) && (k < n);
        const int km = mine ? k : kmin;                 // packet mode: kmin for every lane
        const bool packet = pm != 0;
        if (((++it) & (RT_SAMPLE - 1)) == 0 && packet) {
            low = __popc(__ballot_sync(RT_FULL, mine)) < quarter ? low + 1 : 0;
            if (low >= RT_LOW_RUNS) pm = 0;
        }
        const float4* np = reinterpret_cast<const float4*>(nbase + (unsigned)km * 32u);
        const float4 a = __ldg(np);
        const float4 b = __ldg(np + 1);
        const int cw = __float_as_int(b.w);
        const int lnk = __float_as_int(a.w);
        RT_TALLY(4, mine ? 1 : 0);
        if (COUNT) { if ((threadIdx.x & 31) == 0) tl.v[tbase + (packet ? 0 : 1)] += 1; tl.v[tbase + (packet ? 2 : 3)] += mine ? 1 : 0; }
        // The box test is evaluated by ALL lanes without a branch (idle lanes cost no issue slots;
        // a divergent branch around it would).  Reference nodes: the shader's exact slab test; nodes
        // of our own sub-trees (cw == -2 or leaf bit 30): conservative boxes, skipped only for
        // rays the margins were computed for.
        bool h = slab_fast(r, a, b);
        if (anyOdd) {
            const bool ours = cw == RT_INNER_SUB || (cw >= 0 && (cw & RT_LEAF_SUB_BIT));
            if (ours) h = h || !r.sub;
            else if (!r.fast) h = slab(r, a, b);
        }
        const bool inner = cw < 0;
        const int next = (inner && !h) ? lnk : km + 1;  // inner: hit -> next in list, miss -> skip the subtree
        if (mine) k = next;
        if (!inner && mine && h) {
            const int cnt = cw & RT_LEAF_COUNT_MASK, kind = (cw >> RT_LEAF_KIND_SHIFT) & 3;
            const int end = lnk + cnt;
            RT_TALLY(5, cnt);
            bool stop = false;
            if (!MT && kind == RT_LEAF_TRIANGLES) {
                stop = leaf_triangles_bary(S, q, lnk, end, anyMode);
            } else {
                const Cand c = leaf_generic<MT>(S.q0, S.q1, S.q2, S.q3, S.ptype, S.prank, r, lnk, end, q.best, q.slot, anyMode ? 1 : 0);
                if (c.slot != q.slot) { q.best = c.dist; q.slot = c.slot; q.p = c.p; stop = anyMode; }
            }
            if (stop) k = RT_DONE;
        }
    }
#if RT_TIMELINE
    q.iters += (unsigned)it;
#endif
}
// Brute-force twins (a14, useBVH == false): linear scan over the raw shapes.
template <bool MT>
__device__ __forceinline__ void scan_closest(const DeviceScene& S, const Ray& r, Hit& h) {
    h.dist = 1e20f; h.slot = -1; h.p = mk(0.f, 0.f, 0.f);
    for (int i = 0; i < S.nShapes; ++i) {
        Prim p = prim_from_shape(S.shapes + i);
        V3 hp;
        if (test_prim<MT>(p, r, hp)) {
            float dist = distance3(r.o, hp);
            if (dist < h.dist) { h.dist = dist; h.slot = i; h.p = hp; }
        }
    }
}
template <bool MT>
__device__ __forceinline__ bool scan_any(const DeviceScene& S, const Ray& r, float lightDist) {
    for (int i = 0; i < S.nShapes; ++i) {
        Prim p = prim_from_shape(S.shapes + i);
        V3 hp;
        if (test_prim<MT>(p, r, hp)) {
            float dist = distance3(r.o, hp);
            if (dist < lightDist) return true;       // :572-578 (no 1e20 gate in this twin)
        }
    }
    return false;
}

// ---------------------------------------------------------------------------
// Shading (a1, a2, a10-a13)
// ---------------------------------------------------------------------------
struct Material { V3 color; float fresnel, ambient, diffuse, specular; int shininess; };

__device__ __forceinline__ Material load_material(const rt_flat_shape* __restrict__ s) {
    const float4* m = reinterpret_cast<const float4*>(&s->material);
    float4 m0 = __ldg(m), m1 = __ldg(m + 1);
    Material r;
    r.color = mk(m0.x, m0.y, m0.z); r.fresnel = m0.w;
    r.ambient = m1.x; r.diffuse = m1.y; r.specular = m1.z; r.shininess = __float_as_int(m1.w);
    return r;
}

// getNormalFromShape (a10, :64-71)
__device__ __forceinline__ V3 shape_normal(const rt_flat_shape* __restrict__ sh, V3 hp) {
    if (__ldg(&sh->type) == RT_SHAPE_SPHERE)
        return normalize3(hp - mk(__ldg(&sh->sphereCenter[0]), __ldg(&sh->sphereCenter[1]), __ldg(&sh->sphereCenter[2])));
    return mk(__ldg(&sh->planeNormal[0]), __ldg(&sh->planeNormal[1]), __ldg(&sh->planeNormal[2]));
}

__device__ __forceinline__ V3 phong(V3 point, V3 normal, V3 viewDir, V3 lpos, V3 lcol, const Material& m) {
    float distToLight = distance3(lpos, point);
    V3 lightColor = lcol / distToLight;
    V3 ambient = m.ambient * lightColor;
    V3 lightDir = normalize3(lpos - point);
    float diff = gmax(dot3(normal, lightDir), 0.0f);
    V3 diffuse = (m.diffuse * diff) * lightColor;
    V3 specular = mk(0.f, 0.f, 0.f);
    if (diff > 0.0f) {
        V3 reflectDir = reflect3(-lightDir, normal);
        float spec = powf(gmax(dot3(viewDir, reflectDir), 0.0f), (float)m.shininess);
        specular = (m.specular * spec) * lightColor;
    }
    return ((ambient + diffuse) + specular) * m.color;
}

// One pixel per lane, executed by all 32 lanes of the warp in lock step (the packet walk
// needs the whole warp); `alive` masks the lanes that still have work.  The bounce loop
// (gpu_shader.comp:450-517) is run as a two-phase machine — even steps are the closest-hit
// query of the current ray, odd steps the shadow query and the shading that depends on it —
// so that the walk is instantiated ONCE and both kinds of query share its code.
template <bool MT, bool COUNT>
__device__ __forceinline__ float4 shade_pixel(const DeviceScene& S, const FrameParams& P, int x, int y, bool inside,
                                              int& idOut, Tally& tl, bool paired) {
    const bool bf = S.boxesFinite != 0;
    // Live state across the walks is kept small (occupancy): the view direction, the hit point, the hit
    // shape's index, and the two colour accumulators.  Normal and material are re-read after the shadow walk.
    V3 vd, hp = mk(0.f, 0.f, 0.f);
    V3 acc = mk(0.f, 0.f, 0.f), att = mk(1.f, 1.f, 1.f);
    int shape = -1;
    idOut = -1;
    bool alive = inside;
    Query q;
#if RT_TIMELINE
    q.iters = 0;
#endif
    {
        const float ndcX = 2.0f * (float)x / P.resW - 1.0f;
        const float ndcY = 1.0f - 2.0f * (float)y / P.resH;
        const V3 camP = mk(P.camP);
        const V3 ipp = ((camP + mk(P.camF)) + (ndcX * P.planeW / 2.0f) * mk(P.camR)) + (ndcY * P.planeH / 2.0f) * mk(P.camU);
        q.r = make_ray(camP, normalize3(ipp - camP), bf, S.omax);
        vd = q.r.d;
    }

    // PAIRED mode (a row group of a split heavy block: at most 16 pixel lanes, lane ^ 16 is idle): the shadow
    // query of bounce b and the closest-hit query of bounce b+1 do not depend on each other — the reflected ray
    // follows from the hit and the incoming direction alone (:495-497) — so the idle partner lane walks the
    // reflection while the pixel's own lane walks the shadow ray: the pixel's chain of dependent walks shrinks
    // from 2 * maxBounces to maxBounces + 1.  Same queries, same arithmetic, same pixel.
    paired = paired && P.useBVH && !COUNT && P.maxBounces > 0;   // maxBounces == 0: the bounce loop never runs (:450), no step at all
    const int nSteps = paired ? P.maxBounces + 1 : 2 * P.maxBounces;
    for (int step = 0; step < nSteps; ++step) {
        if (!__any_sync(RT_FULL, alive)) break;
        const bool shadowStep = paired ? step > 0 : (step & 1) != 0;
        float lightDist = 0.f;
        bool helper = false;                               // this lane walks its partner's reflected ray in this step
        if (!shadowStep) {
            q.best = 1e20f;                                // q.r is the current ray
        } else {
            const V3 lpos = mk(P.lightPos);
            const V3 nrm = shape >= 0 ? shape_normal(S.shapes + shape, hp) : mk(0.f, 0.f, 1.f);
            lightDist = distance3(lpos, hp);
            q.r = make_ray(hp + nrm * (P.useBVH ? 1e-3f : 1e-5f), normalize3(lpos - hp), bf, S.omax);
            // dist < 1e20 && dist < lightDist  ==  dist < min; a NaN lightDist can never be beaten
            q.best = lightDist < 1e20f ? lightDist : (lightDist >= 1e20f ? 1e20f : -1.0f);
            if (paired) {
                bool cont = false;
                V3 ro = mk(0.f, 0.f, 0.f), rd = ro;
                if (alive && step < P.maxBounces) {         // bounce step-1 continues iff specular > 0 (:495) and a bounce is left
                    cont = __ldg(&S.shapes[shape].material.specularStrength) > 0.0f;
                    ro = hp + nrm * 1e-3f; rd = reflect3(vd, nrm);
                }
                const bool pc = __shfl_xor_sync(RT_FULL, cont ? 1 : 0, 16) != 0;
                const V3 po = mk(__shfl_xor_sync(RT_FULL, ro.x, 16), __shfl_xor_sync(RT_FULL, ro.y, 16), __shfl_xor_sync(RT_FULL, ro.z, 16));
                const V3 pd = mk(__shfl_xor_sync(RT_FULL, rd.x, 16), __shfl_xor_sync(RT_FULL, rd.y, 16), __shfl_xor_sync(RT_FULL, rd.z, 16));
                if (!alive && pc) { helper = true; q.r = make_ray(po, pd, bf, S.omax); q.best = 1e20f; }
            }
        }
        q.slot = -1; q.p = mk(0.f, 0.f, 0.f);
        int nextSlot = -1; V3 nextP = mk(0.f, 0.f, 0.f);   // paired: the partner's closest hit of the reflected ray
        if (P.useBVH) {
            walk<MT, COUNT>(S, q, alive || helper, shadowStep && !helper, tl, step < 2 ? 8 : 12);
            if (paired && shadowStep) {
                nextSlot = __shfl_xor_sync(RT_FULL, q.slot, 16);
                nextP = mk(__shfl_xor_sync(RT_FULL, q.p.x, 16), __shfl_xor_sync(RT_FULL, q.p.y, 16), __shfl_xor_sync(RT_FULL, q.p.z, 16));
            }
        } else if (alive) {
            if (!shadowStep) { Hit h; scan_closest<MT>(S, q.r, h); q.slot = h.slot; q.p = h.p; RT_TALLY(5, S.nShapes); }
            else if (scan_any<MT>(S, q.r, lightDist)) q.slot = 0;
        }
        if (!alive) continue;
        RT_TALLY(0, 1);
        if (!shadowStep) {                                 // closest-hit result (:452-464)
            shape = q.slot < 0 ? -1 : P.useBVH ? __ldg(S.pshape + q.slot) : q.slot;
            RT_TALLY(step == 0 ? 1 : 3, 1); RT_TALLY(6, shape >= 0 ? 1 : 0);
            if (step == 0) idOut = shape;
            if (shape < 0) {
                const float a = (float)y / P.resH;
                const V3 bg = mk(mix1(0.05f, 0.5f, a), mix1(0.07f, 0.7f, a), mix1(0.1f, 1.0f, a));
                acc = acc + att * bg; alive = false; continue;
            }
            hp = q.p;
        } else {                                           // shadow result, shading, next ray (:474-516)
            const bool inShadow = q.slot >= 0;
            RT_TALLY(2, 1); RT_TALLY(7, inShadow ? 1 : 0);
            const rt_flat_shape* sh = S.shapes + shape;
            const Material m = load_material(sh);
            const V3 nrm = shape_normal(sh, hp);
            V3 pc = phong(hp, nrm, vd, mk(P.lightPos), mk(P.lightCol), m);
            if (inShadow) pc = pc * 0.3f;
            acc = acc + att * pc;
            if (!(m.specular > 0.0f)) { alive = false; continue; }
            const V3 rd = reflect3(vd, nrm);
            q.r = make_ray(hp + nrm * 1e-3f, rd, bf, S.omax);
            vd = rd;
            if (P.useFresnel) {
                float fr = powf(1.0f - gmax(dot3(-rd, nrm), 0.0f), 5.0f);
                fr = gmin(gmax(fr, 0.0f), 0.8f);
                const float rw = m.fresnel * fr, mw = 1.0f - rw;
                att = att * mk(mix1(m.color.x, 1.0f, rw), mix1(m.color.y, 1.0f, rw), mix1(m.color.z, 1.0f, rw));
                acc = acc + (mw * m.color) * pc;
            } else {
                att = att * m.specular;
            }
            if (paired) {                                  // the reflected ray's closest hit is already here (:452-464)
                if (step >= P.maxBounces) { alive = false; continue; }   // bounces exhausted: no background (:450)
                shape = nextSlot < 0 ? -1 : __ldg(S.pshape + nextSlot);
                if (shape < 0) {
                    const float a = (float)y / P.resH;
                    const V3 bg = mk(mix1(0.05f, 0.5f, a), mix1(0.07f, 0.7f, a), mix1(0.1f, 1.0f, a));
                    acc = acc + att * bg; alive = false; continue;
                }
                hp = nextP;
            }
        }
    }
#if RT_TIMELINE
    return make_float4(acc.x, acc.y, acc.z, __uint_as_float(q.iters));
#else
    return make_float4(acc.x, acc.y, acc.z, 1.0f);
#endif
}

// ---------------------------------------------------------------------------
// K1: render.  Persistent warps with work stealing: the grid is sized to fill the
// SMs once and every WARP pulls the next 8x4-pixel block from an atomic queue
// (lane 0 claims, __shfl broadcasts), so a heavy block delays one warp, not a CTA.
// Work item w: shard tile t = w / 8 of this rank (tile index rank + t*world), warp
// block b = w % 8 inside the 32x8 tile (4 across, 2 down).
// world == 1 writes row-major W*H; world > 1 writes the compact tile-major shard.
// ---------------------------------------------------------------------------
#define RT_WARPS_PER_TILE ((RT_TILE_W / 8) * (RT_TILE_H / 4))
#ifndef RT_HEAVY_SLOTS
#define RT_HEAVY_SLOTS 24      // warp slots (of 32 resident) that claim heavy blocks first
#endif

template <bool MT, bool COUNT, int MINB>
__global__ void __launch_bounds__(RT_BLOCK_THREADS, MINB)
render_kernel(DeviceScene S, FrameParams P, float4* __restrict__ out, int* __restrict__ ids,
              unsigned long long* __restrict__ tally, unsigned int* __restrict__ queue) {
    Tally tl;
    if (COUNT) { for (int i = 0; i < RT_NTALLY; ++i) tl.v[i] = 0; }
    const int lane = threadIdx.x & 31;
    const int nItems = P.nTilesMine * RT_WARPS_PER_TILE;
    // Heavy-first order: the critical path of a frame is its slowest 8x4 block (an incoherent bounce
    // packet walks hundreds of nodes in sequence); blocks that were slow in the previous frame are
    // claimed first so that their chains overlap the bulk of the work instead of forming the tail.
    // The SM's warp schedulers favour the low warp slots (measured: slots 0-7 complete twice as many
    // blocks as slots 28-31, tools/timeline.py), and a slow block on a starved slot runs three times
    // longer — so only the favoured half of the slots claims from the heavy list while it lasts.
    const int nHeavy = P.lptCtl ? (int)__ldg(P.lptCtl + P.lptPar) : 0;
    unsigned long long costSum = 0;              // this warp's share of the frame's cost sum (one atomic per warp at the end)
    unsigned warpSlot;
    asm("mov.u32 %0, %%warpid;" : "=r"(warpSlot));
    const bool favoured = warpSlot < RT_HEAVY_SLOTS;
    bool heavyLeft = nHeavy > 0, lightLeft = true;
    while (heavyLeft || lightLeft) {             // persistent warp: until both lists are empty (this form of the loop
                                                 // header matters: with for (;;) ptxas spills 154 instead of 54 bytes)
        int w = -1;
        int parts = 1, part = 0;                 // a heavy block may come as 2 or 4 row groups (see lpt_mark_kernel)
        if (heavyLeft && (favoured || !lightLeft)) {
            int hq = 0;
            if (lane == 0) hq = (int)atomicAdd(P.lptCtl + 2, 1u);
            hq = __shfl_sync(0xffffffffu, hq, 0);
            if (hq < nHeavy) {
                const unsigned e = __ldg(P.heavyList + hq);
                w = (int)(e & 0x00ffffffu); part = (int)((e >> 24) & 31u); parts = 1 << (e >> 29);
            } else heavyLeft = false;
        }
        if (w < 0 && lightLeft) {
            int lq = 0;
            if (lane == 0) lq = (int)atomicAdd(queue, 1u);
            lq = __shfl_sync(0xffffffffu, lq, 0);
            if (lq >= nItems) lightLeft = false;
            else if (!(nHeavy > 0 && __ldg(P.heavyFlag + lq))) w = lq;
        }
        if (w < 0) { if (!heavyLeft && !lightLeft) break; continue; }
        const long long itemClk0 = P.itemCost ? clock64() : 0;
#if RT_TIMELINE
        const long long tlClk0 = clock64();
#endif
        const int t = P.tileBegin + w / RT_WARPS_PER_TILE, b = w % RT_WARPS_PER_TILE;
        const int tile = P.shardRank + t * P.shardWorld;
        const int lx = ((b & 3) << 3) + (lane & 7);
        const int ly = ((b >> 2) << 2) + (lane >> 3);
        // tile index -> place in the tile grid.  One GPU: row-major.  Sharded: every tile row is rotated by its row number, so
        // that a rank's tiles (k = rank mod world) form DIAGONALS instead of the vertical stripes they would form whenever the
        // row length is a multiple of world (4K: 120 tiles per row) — the content of a frame varies across columns
        const int trow = tile / P.tilesX, tc0 = tile - trow * P.tilesX;
        const int tcol = P.shardWorld == 1 ? tc0 : (tc0 + trow) % P.tilesX;
        const int x = tcol * RT_TILE_W + lx;
        const int y = trow * RT_TILE_H + ly;
        const bool partOK = ((lane * parts) >> 5) == part;             // the lanes (rows / row pieces of the 8x4 block) this warp renders
        const bool inside = x < P.W && y < P.H && partOK;
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        int id = -1;
        unsigned long long cost0 = 0;
        long long clk0 = 0;
        unsigned long long it0[4] = {0, 0, 0, 0};
        if (COUNT) { cost0 = tl.v[4] + tl.v[5]; clk0 = clock64(); it0[0] = tl.v[8]; it0[1] = tl.v[9]; it0[2] = tl.v[12]; it0[3] = tl.v[13]; }
        c = shade_pixel<MT, COUNT>(S, P, x, y, inside, id, tl, parts >= 2);
        if (COUNT && P.writeIds == 2) id = (int)(tl.v[4] + tl.v[5] - cost0);   // diagnostic: per-pixel work AOV
        if (COUNT) {
            const unsigned long long dt = (unsigned long long)(clock64() - clk0);
            if (lane == 0) { atomicMax(tally + 16, dt); tl.v[17] += dt; tl.v[18] += 1; }
            if (P.writeIds == 2) {                          // walk iterations of this item (lane 0 counts them): pixels 4..7
                const unsigned d0 = __shfl_sync(0xffffffffu, (unsigned)(tl.v[8] - it0[0]), 0), d1 = __shfl_sync(0xffffffffu, (unsigned)(tl.v[9] - it0[1]), 0);
                const unsigned d2 = __shfl_sync(0xffffffffu, (unsigned)(tl.v[12] - it0[2]), 0), d3 = __shfl_sync(0xffffffffu, (unsigned)(tl.v[13] - it0[3]), 0);
                if (lane >= 4 && lane < 8) id = (int)(lane == 4 ? d0 : lane == 5 ? d1 : lane == 6 ? d2 : d3);
            }
            if (P.writeIds == 2 && lane < 4) {              // timeline AOV in the first four pixels of the block
                unsigned long long gt; unsigned smid, wid;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
                asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
                asm volatile("mov.u32 %0, %%warpid;" : "=r"(wid));
                id = lane == 0 ? (int)(gt & 0x7fffffff) : lane == 1 ? (int)(dt >> 2) : lane == 2 ? (int)smid : (int)wid;
            }
        }
#if RT_TIMELINE      // diagnostic build only (tools/timeline2.py): the ids plane carries when and where every part of an item ran
        if (!COUNT && P.writeIds == 1) {
            unsigned long long gt; unsigned smid;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            const int rel = lane - (part << 5) / parts;
            const unsigned long long dtl = (unsigned long long)(clock64() - tlClk0);
            if (rel == 0) id = (int)(gt & 0x7fffffffull);
            else if (rel == 1) id = (int)(dtl >> 2);
            else if (rel == 2) id = (int)(smid | (warpSlot << 8) | ((unsigned)parts << 16) | ((unsigned)part << 24));
            else if (rel == 3) id = (int)__float_as_uint(c.w);
        }
        c.w = 1.0f;
#endif
        __syncwarp();
        if (P.shardWorld == 1) {
            if (inside) {
                const size_t o = (size_t)y * (size_t)P.W + (size_t)x;
                out[o] = c;
                if (P.writeIds) ids[o] = id;
            }
        } else {
            const size_t o = (size_t)t * (RT_TILE_W * RT_TILE_H) + (size_t)(ly * RT_TILE_W + lx);
            // fused gather: the pixel goes straight to its row-major place in the full frame, which
            // may live on another GPU (peer mapping over NVLink); 128-byte row segments per warp
            if (P.frameTarget) { if (inside) out[(size_t)y * (size_t)P.W + (size_t)x] = c; }
            else if (partOK) out[o] = c;
            if (P.writeIds && partOK) ids[o] = id;
        }
        if (P.itemCost && lane == 0) {           // parts add up (lpt_mark_kernel zeroed the entry of a split block)
            const unsigned dtc = (unsigned)((unsigned long long)(clock64() - itemClk0) >> 4);
            if (parts == 1) P.itemCost[w] = dtc; else atomicAdd(P.itemCost + w, dtc);
            costSum += dtc;
        }
    }
    if (P.lptCtl && lane == 0 && costSum) atomicAdd(reinterpret_cast<unsigned long long*>(P.lptCtl + 4 + 2 * P.lptPar), costSum);
    if (COUNT) {
        for (int i = 0; i < RT_NTALLY; ++i) {
            unsigned long long v = tl.v[i];
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && v && i != 16 && i != 19) atomicAdd(tally + i, v);
        }
    }    if (P.doneWord) {
        // compute + signal: every thread makes its pixel stores (peer memory over NVLink) visible system-wide, the last
        // thread block of the launch then adds 1 to the completion counter on rank 0 (a peer atomic over NVLink)
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(P.ctaDone, 1u) == gridDim.x - 1) {
                *P.ctaDone = 0u;                            // ready for the next launch (stream order): no memset per frame
                __threadfence_system();
                atomicAdd_system(P.doneWord, 1u);           // ONE completion counter on rank 0: every other rank adds 1 per frame
            }
        }
    }
}

// ---------------------------------------------------------------------------
// Heavy-first list for this launch from the previous launch's item costs: ONE small launch per frame (round 1 used two
// memsets and two kernels: ~10 us of a 0.3 ms frame).  The render kernel itself sums its item costs (one atomic per warp)
// into the slot of its frame parity; this kernel reads the previous frame's sum, marks and lists the heavy items into this
// parity's count, and resets what the coming render launch and the next prepare launch need: the two claim counters, this
// parity's sum, the other parity's count.
// ---------------------------------------------------------------------------
// A block is HEAVY when its cost exceeds heavyFactor x the mean.  Measured on the car frame at 4K over 1 / 2 / 4 / 8 ranks
// (profiles/r2/heavy_first_sweep.txt): 4 x is as good as 3 x on a whole frame and 6 - 8 % faster on a half or a quarter
// of it; with fewer than five blocks per warp (an eighth of the frame, 800x600) 3 x is better.
#ifndef RT_HEAVY_FACTOR
#define RT_HEAVY_FACTOR 4
#endif
#ifndef RT_HEAVY_FACTOR_FEW
#define RT_HEAVY_FACTOR_FEW 3
#endif
// A block whose cost is a large share of what ONE warp would ideally do in the whole frame (sum / warps of the
// grid) bounds the frame however early it starts — the case of small frames and of many GPUs sharing a frame.
// Such a block is listed as 2^k lane groups (8x2, 8x1, 4x1 ... pixels each, other lanes idle) that different
// warps pick up: the dependent walk of each group is shorter because its lane-mode iterations serve fewer lanes.
#ifndef RT_SPLIT_MAXLG
#define RT_SPLIT_MAXLG 3       // at most 2^this parts
#endif
#ifndef RT_SPLIT_K
#define RT_SPLIT_K 2           // parts ~ K * cost / share, rounded down to a power of two (two parts from cost > share on)
#endif
__global__ void lpt_prepare_kernel(unsigned* __restrict__ cost, int n, unsigned* __restrict__ ctl,
                                   unsigned* __restrict__ heavyList, unsigned char* __restrict__ heavyFlag, int gridWarps, int heavyFactor, int par) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { ctl[2] = 0u; ctl[3] = 0u; ctl[4 + 2 * par] = 0u; ctl[5 + 2 * par] = 0u; ctl[par ^ 1] = 0u; }
    if (i >= n) return;
    const unsigned long long sum = *reinterpret_cast<const unsigned long long*>(ctl + 4 + 2 * (par ^ 1));
    const unsigned long long ci = cost[i];
    const bool heavy = ci * (unsigned long long)n > (unsigned long long)heavyFactor * sum;
    heavyFlag[i] = heavy ? 1 : 0;
    if (!heavy) return;
    const unsigned long long want = sum > 0 ? ci * (unsigned long long)gridWarps * RT_SPLIT_K / sum : 0;   // ~ K * cost / share
    unsigned lg = 0;
    while (lg < RT_SPLIT_MAXLG && (2ull << lg) <= want) ++lg;
    const unsigned parts = 1u << lg;
    const unsigned at = atomicAdd(ctl + par, parts);
    for (unsigned p = 0; p < parts; ++p) heavyList[at + p] = (unsigned)i | (p << 24) | (lg << 29);
    if (parts > 1) cost[i] = 0;                  // the parts add their costs up
}

// ---------------------------------------------------------------------------
// K2: repack.  nodes: gather the raw 48-byte nodes into the threaded layout;
// prims: gather the raw 192-byte shapes into float4 SoA in visit order.
// ---------------------------------------------------------------------------

// K2 as ONE launch (round 2): list A, list B and the primitive slots are independent of each other; the block index selects
// the job (blocks [0, bA): list A; [bA, bA + bB): list B; the rest: primitive slots).  A small scene's frame update paid three
// launch gaps for 13 blocks of work.
struct RepackAll {
    const rt_flat_node* raw; const float4* subBox;
    const int* orderA; const int2* linkA; float4* outA; int nA;
    const int* orderB; const int2* linkB; float4* outB; int nB;
    const rt_flat_shape* shapes; const int* pshape; float4 *q0, *q1, *q2, *q3; int* ptype; int nPrims;
    int blocksA, blocksB;
};
__global__ void __launch_bounds__(256) repack_all_kernel(const __grid_constant__ RepackAll P) {
    int blk = blockIdx.x;
    if (blk < P.blocksA + P.blocksB) {
        const bool inB = blk >= P.blocksA;
        if (inB) blk -= P.blocksA;
        const int k = blk * 256 + threadIdx.x, n = inB ? P.nB : P.nA;
        if (k >= n) return;
        const int o = (inB ? P.orderB : P.orderA)[k];
        const float4* src = o >= 0 ? reinterpret_cast<const float4*>(P.raw + o) : P.subBox + 2 * (size_t)(-1 - o);
        const float4 mn = __ldg(src), mx = __ldg(src + 1);
        const int2 lc = (inB ? P.linkB : P.linkA)[k];
        float4* out = inB ? P.outB : P.outA;
        out[2 * k]     = make_float4(mn.x, mn.y, mn.z, __int_as_float(lc.x));
        out[2 * k + 1] = make_float4(mx.x, mx.y, mx.z, __int_as_float(lc.y));
    } else {
        const int s = (blk - P.blocksA - P.blocksB) * 256 + threadIdx.x;
        if (s >= P.nPrims) return;
        const Prim p = prim_from_shape(P.shapes + P.pshape[s]);
        P.q0[s] = p.q0; P.q1[s] = p.q1; P.q2[s] = p.q2; P.q3[s] = p.q3; P.ptype[s] = p.type;
    }
}

// ---------------------------------------------------------------------------
// K3: untile.  shards = `world` compact buffers of shardPixels float4 each, rank
// after rank (the NCCL gather's landing layout) -> row-major W*H.
// ---------------------------------------------------------------------------
// signalled fused gather: rank 0 tells the other ranks that it has finished reading frame `seq`
__global__ void release_kernel(unsigned* const* __restrict__ words, int n, unsigned seq) {
    const int i = threadIdx.x;
    if (i < n && words[i]) { __threadfence_system(); *reinterpret_cast<volatile unsigned*>(words[i]) = seq; }
}

// One block of 256 threads per 32x8 tile of the OUTPUT, four tiles per block and pass: the tile arithmetic (divisions by the
// runtime tile-row length) is done once per tile on the uniform datapath instead of once per pixel, a tile is one contiguous
// 4 KB read and eight 512-byte row segments of writes, and the four independent loads per thread keep enough bytes in
// flight for HBM (one pixel per thread with per-pixel divisions: 72 % of the measured copy peak at 4K, tools/k35_probe.py).
#define RT_UNTILE_UNROLL 4
__global__ void __launch_bounds__(RT_TILE_W * RT_TILE_H)
untile_kernel(const float4* __restrict__ shards, float4* __restrict__ out, int W, int H,
              int tilesX, int nTiles, int world, long long shardPixels) {
    const int lx = threadIdx.x % RT_TILE_W, ly = threadIdx.x / RT_TILE_W;
    for (int base = blockIdx.x * RT_UNTILE_UNROLL; base < nTiles; base += gridDim.x * RT_UNTILE_UNROLL) {
        float4 v[RT_UNTILE_UNROLL];
        long long dst[RT_UNTILE_UNROLL];
#pragma unroll
        for (int u = 0; u < RT_UNTILE_UNROLL; ++u) {
            const int to = base + u;                       // tile of the output, row-major over the tile grid
            dst[u] = -1;
            if (to >= nTiles) continue;
            const int trow = to / tilesX, tcol = to - trow * tilesX;
            const int tc0 = world == 1 ? tcol : (tcol + tilesX - trow % tilesX) % tilesX;     // inverse of the render kernel's row rotation
            const int tile = trow * tilesX + tc0;
            const int rank = tile % world, t = tile / world;
            const int x = tcol * RT_TILE_W + lx, y = trow * RT_TILE_H + ly;
            if (x < W && y < H) {
                dst[u] = (long long)y * W + x;
                v[u] = __ldg(shards + (long long)rank * shardPixels + (long long)t * (RT_TILE_W * RT_TILE_H) + threadIdx.x);
            }
        }
#pragma unroll
        for (int u = 0; u < RT_UNTILE_UNROLL; ++u) if (dst[u] >= 0) out[dst[u]] = v[u];
    }
}

// ---------------------------------------------------------------------------
// K4: present.  shader.frag:8-12 samples the image 1:1 and writes
// vec4(color.rgb, 1); the default framebuffer clamps to [0,1] and stores RGBA8.
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned quant8(float v) {
    v = gmin(gmax(v, 0.0f), 1.0f);
    if (!(v == v)) v = 0.0f;
    return (unsigned)(int)(v * 255.0f + 0.5f);
}
// Four pixels per thread, a whole pass of the grid apart (every load and store instruction of a warp is contiguous): the
// four independent 16-byte loads keep enough bytes in flight (one pixel per thread: 62 % of the copy peak at 4K).
#define RT_PRESENT_UNROLL 4
__global__ void present_kernel(const float4* __restrict__ in, uchar4* __restrict__ out, long long n) {
    const long long T = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += T * RT_PRESENT_UNROLL) {
        float4 c[RT_PRESENT_UNROLL];
#pragma unroll
        for (int u = 0; u < RT_PRESENT_UNROLL; ++u) if (i + u * T < n) c[u] = __ldg(in + i + u * T);
#pragma unroll
        for (int u = 0; u < RT_PRESENT_UNROLL; ++u)
            if (i + u * T < n) out[i + u * T] = make_uchar4((unsigned char)quant8(c[u].x), (unsigned char)quant8(c[u].y), (unsigned char)quant8(c[u].z), 255);
    }
}

}  // namespace rt

// rt_layout.h — host-side builder of the device scene layout (C++17; no CUDA needed: the test
// emulator compiles it with g++.  Box / prim_bound are also compiled for the device, where the
// refit kernel (rt_kernels.cuh: refit_kernel) recomputes the conservative boxes after shapes moved).
//
// Input: the reference's raw FlatNode[] / bvhIndices[] / FlatShape[] (src/flatStructures.hpp).
// Output: the threaded node list the render kernel walks:
//   * reference nodes in the reference's visit order (root = last raw node, right child first,
//     gpu_shader.comp:386, 424-425), each with a skip link;
//   * FAT leaves (more than RT_FAT_LEAF shapes — the reference builder stops splitting as soon as
//     one side of a midpoint split is empty, src/main.cpp:1156-1161, and leaves reach hundreds of
//     shapes) become inner nodes over a private binned-SAH sub-tree whose boxes are CONSERVATIVE
//     bounds of each primitive's acceptance region (see prim_bound): a ray that misses such a box
//     cannot be accepted by any primitive below it, so skipping the subtree changes no decision.
//     Primitives whose acceptance region cannot be bounded (infinite planes, NaN wall bases,
//     degenerate / sliver triangles, non-finite data) stay in an "always tested" leaf.
//   * every primitive slot carries its RANK = position in the reference's own primitive visit
//     sequence; the kernel resolves equal-distance candidates by smaller rank, which is exactly
//     "first candidate in visit order wins" (strict '<', gpu_shader.comp:410-411) whatever the
//     order the sub-tree presents them in.
#pragma once
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/rt_flat.h"

#if defined(__CUDACC__)
#define RTL_HD __host__ __device__
#else
#define RTL_HD
#endif

namespace rtl {

// math that must compile for host and device alike
RTL_HD inline double dmin(double a, double b) { return a < b ? a : b; }
RTL_HD inline double dmax(double a, double b) { return a > b ? a : b; }
RTL_HD inline double dabs(double a) { return a < 0 ? -a : a; }
RTL_HD inline float  fmin_(float a, float b) { return b < a ? b : a; }      // std::min semantics (keeps a on NaN b)
RTL_HD inline float  fmax_(float a, float b) { return a < b ? b : a; }
RTL_HD inline bool   fin(double x) { return dabs(x) <= 1.7976931348623157e308; }   // false for NaN and +-inf
RTL_HD inline float  f_up(float x) {        // nextafter(x, +inf) for finite x
    if (!(x < INFINITY)) return x;             // +inf, NaN
    if (x == 0.f) return 1.401298464e-45f;
    union { float f; uint32_t u; } v; v.f = x;
    v.u = x > 0.f ? v.u + 1u : v.u - 1u;
    return v.f;
}
RTL_HD inline float  f_down(float x) { return -f_up(-x); }

#ifndef RT_FAT_LEAF
#define RT_FAT_LEAF      8      // reference leaves with more shapes than this get a sub-tree
#endif
#ifndef RT_SUB_LEAF
#define RT_SUB_LEAF      4      // sub-tree leaves hold at most this many primitives
#endif
#define RTL_COUNT_MASK   0x0fffffff
#define RTL_KIND_SHIFT   28
#define RTL_SUB_BIT      (1 << 30)   // leaf count word: box is one of ours (conservative test allowed)
#define RTL_INNER_REF    (-1)        // inner count word: reference node (exact reference slab test)
#define RTL_INNER_SUB    (-2)        // inner count word: sub-tree node
enum { KIND_MIXED = 0, KIND_TRIANGLES = 1, KIND_SPHERES = 2 };

struct Box {
    float mn[3], mx[3];
    RTL_HD void reset() { for (int a = 0; a < 3; ++a) { mn[a] = INFINITY; mx[a] = -INFINITY; } }
    RTL_HD void grow(const Box& b) { for (int a = 0; a < 3; ++a) { mn[a] = fmin_(mn[a], b.mn[a]); mx[a] = fmax_(mx[a], b.mx[a]); } }
    RTL_HD void grow(const double p[3]) { for (int a = 0; a < 3; ++a) { mn[a] = fmin_(mn[a], (float)p[a]); mx[a] = fmax_(mx[a], (float)p[a]); } }
    RTL_HD void everything() { for (int a = 0; a < 3; ++a) { mn[a] = -INFINITY; mx[a] = INFINITY; } }   // a box every finite ray passes
    double area() const {
        double d[3] = {(double)mx[0] - mn[0], (double)mx[1] - mn[1], (double)mx[2] - mn[2]};
        return d[0] < 0 ? 0.0 : 2.0 * (d[0] * d[1] + d[1] * d[2] + d[2] * d[0]);
    }
};

// ---------------------------------------------------------------------------
// Conservative bound of a primitive's ACCEPTANCE REGION.
//
// Needed property: if get_intersection (gpu_shader.comp:242-328, either triangle mode) returns
// INNER for a ray whose origin satisfies |o|_inf <= omax, then the ray's line passes through the
// returned box (with a positive exit parameter).  fp32 rounding makes the accepted set slightly
// larger than the geometric primitive, by an amount that scales with the magnitude of the ray
// origin (e.g. the sphere discriminant bb^2 - 4 aa cc carries an absolute error ~ eps |o-c|^2),
// so the margins below are expressed with S = omax + |primitive coordinates| and E = 2^-23, with
// generous safety factors.  Rays with a larger origin ignore the sub-tree boxes in the kernel.
// Returns false = unbounded (primitive goes to the always-tested leaf).
// ---------------------------------------------------------------------------
RTL_HD inline bool finite3(const float* p) { return fin(p[0]) && fin(p[1]) && fin(p[2]); }
RTL_HD inline double amax3(const float* p) { return dmax(dabs((double)p[0]), dmax(dabs((double)p[1]), dabs((double)p[2]))); }

RTL_HD inline void inflate(Box& b, double m) {
    for (int a = 0; a < 3; ++a) {
        // outward rounding of the double arithmetic to float (callers guarantee finite values below 1e30)
        b.mn[a] = f_down((float)((double)b.mn[a] - m));
        b.mx[a] = f_up((float)((double)b.mx[a] + m));
    }
}

RTL_HD inline bool prim_bound(const rt_flat_shape& s, double omax, Box& out) {
    const double E = 1.1920929e-7;
    out.reset();
    if (s.type == RT_SHAPE_SPHERE) {
        if (!finite3(s.sphereCenter) || !fin(s.sphereRadius)) return false;
        const double S = omax + amax3(s.sphereCenter) + dabs((double)s.sphereRadius);
        const double r = dabs((double)s.sphereRadius);
        // D > 0 can be decided wrongly while p^2 < r^2 + c E |o-c|^2 (p = distance line-centre)
        const double reach = ::sqrt(r * r + 64.0 * E * S * S) + 64.0 * E * S;
        if (!(reach < 1e30)) return false;
        double c[3] = {s.sphereCenter[0], s.sphereCenter[1], s.sphereCenter[2]};
        out.grow(c);
        inflate(out, reach);
        return true;
    }
    if (s.type == RT_SHAPE_WALL) {
        const float* n = s.planeNormal;
        if (!finite3(n) || !finite3(s.wallStart) || !fin(s.planeD) || !fin(s.wallWidth) || !fin(s.wallHeight)) return false;
        const double nn = (double)n[0] * n[0] + (double)n[1] * n[1] + (double)n[2] * n[2];
        if (!(nn > 1e-12) || !(nn < 1e12)) return false;
        // the shader's tangent basis (gpu_shader.comp:305-307), here in double; a degenerate first
        // cross product gives NaN in fp32 (the wall then accepts its whole plane)
        double u[3] = {-(double)n[2], 0.0, (double)n[0]};                 // n x (0,1,0)
        double ul = ::sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        if (!(ul > 1e-3 * ::sqrt(nn))) return false;                    // near the NaN / fallback switch: do not bound
        for (double& x : u) x /= ul;
        double v[3] = {n[1] * u[2] - u[1] * n[2], n[2] * u[0] - u[2] * n[0], n[0] * u[1] - u[0] * n[1]};
        double vl = ::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        if (!(vl > 0)) return false;
        for (double& x : v) x /= vl;
        const double w = s.wallWidth, h = s.wallHeight;
        if (w < 0 || h < 0) { /* empty acceptance set; still bound it by the corner */ }
        const double S = omax + amax3(s.wallStart) + dabs(w) + dabs(h);
        for (int i = 0; i < 4; ++i) {
            double a = (i & 1) ? dmax(w, 0.0) : 0.0, b = (i & 2) ? dmax(h, 0.0) : 0.0, c[3];
            for (int k = 0; k < 3; ++k) c[k] = s.wallStart[k] + a * u[k] + b * v[k];
            // hit points lie on the stored plane (n, D): project the corner onto it along n
            const double off = ((double)n[0] * c[0] + (double)n[1] * c[1] + (double)n[2] * c[2] + (double)s.planeD) / nn;
            for (int k = 0; k < 3; ++k) c[k] -= off * n[k];
            if (!fin(c[0]) || !fin(c[1]) || !fin(c[2])) return false;
            out.grow(c);
        }
        const double m = 128.0 * E * (S + amax3(out.mn) + amax3(out.mx));
        if (!(m < 1e30)) return false;
        inflate(out, m);
        return true;
    }
    if (s.type == RT_SHAPE_TRIANGLE) {
        const float* n = s.planeNormal;
        if (!finite3(n) || !finite3(s.triP1) || !finite3(s.triP2) || !finite3(s.triP3) || !fin(s.planeD)) return false;
        double P[3][3], e1[3], e2[3];
        for (int k = 0; k < 3; ++k) { P[0][k] = s.triP1[k]; P[1][k] = s.triP2[k]; P[2][k] = s.triP3[k]; e1[k] = P[1][k] - P[0][k]; e2[k] = P[2][k] - P[0][k]; }
        const double d00 = e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2], d11 = e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2];
        const double d01 = e1[0] * e2[0] + e1[1] * e2[1] + e1[2] * e2[2];
        const double denom = d00 * d11 - d01 * d01;
        if (!(d00 > 0) || !(d11 > 0)) return false;
        const double sin2 = denom / (d00 * d11);
        if (!(sin2 > 1e-4)) return false;                                  // sliver / degenerate: hits everywhere on its plane
        const double l1 = ::sqrt(d00), l2 = ::sqrt(d11);
        if (dmax(l1, l2) > 1e4 * dmin(l1, l2)) return false;
        double nt[3] = {e1[1] * e2[2] - e2[1] * e1[2], e1[2] * e2[0] - e2[2] * e1[0], e1[0] * e2[1] - e2[0] * e1[1]};
        const double ntl = ::sqrt(nt[0] * nt[0] + nt[1] * nt[1] + nt[2] * nt[2]);
        const double nn = (double)n[0] * n[0] + (double)n[1] * n[1] + (double)n[2] * n[2];
        if (!(nn > 1e-12) || !(nn < 1e12)) return false;
        const double cosv = ((double)n[0] * nt[0] + (double)n[1] * nt[1] + (double)n[2] * nt[2]) / (ntl * ::sqrt(nn));
        // Barycentric mode accepts points of the STORED plane (n, D) whose projection along the
        // triangle normal falls into the triangle; animated triangles carry a stale plane
        // (src/main.cpp:1101-1106), so the region is the triangle projected onto (n, D) along nt.
        if (!(dabs(cosv) > 0.05)) return false;
        const double S = omax + amax3(s.triP1) + amax3(s.triP2) + amax3(s.triP3);
        // barycentric fuzz (relative to the triangle, independent of the ray origin: the origin-dependent
        // error of the hit point itself is the absolute margin m below): accepted while u,v,w >= -delta  ==  triangle scaled by (1+3 delta) about its centroid
        const double delta = 64.0 * E * (1.0 + (l1 + l2) / dmin(l1, l2)) / sin2;
        if (!(delta < 0.5)) return false;
        double cen[3];
        for (int k = 0; k < 3; ++k) cen[k] = (P[0][k] + P[1][k] + P[2][k]) / 3.0;
        const double nd = (double)n[0] * nt[0] + (double)n[1] * nt[1] + (double)n[2] * nt[2];
        for (int i = 0; i < 3; ++i) {
            double q[3];
            for (int k = 0; k < 3; ++k) q[k] = cen[k] + (P[i][k] - cen[k]) * (1.0 + 3.0 * delta);
            out.grow(q);                                                   // Moller-Trumbore mode: the triangle itself
            const double off = ((double)n[0] * q[0] + (double)n[1] * q[1] + (double)n[2] * q[2] + (double)s.planeD) / nd;
            for (int k = 0; k < 3; ++k) q[k] -= off * nt[k];
            if (!fin(q[0]) || !fin(q[1]) || !fin(q[2])) return false;
            out.grow(q);                                                   // Barycentric mode: its image on the stored plane
        }
        const double m = 128.0 * E * (S + amax3(out.mn) + amax3(out.mx)) / ::sqrt(sin2) / dabs(cosv);
        if (!(m < 1e30)) return false;
        inflate(out, m);
        return true;
    }
    return false;   // infinite planes (type 1) and unknown types
}

// ---------------------------------------------------------------------------
// Sub-tree over the bounded primitives of one fat leaf: binned SAH, leaves <= RT_SUB_LEAF.
// ---------------------------------------------------------------------------
struct SubNode { Box box; int left = -1, right = -1; int first = 0, count = 0; };   // leaf: left == -1, prims [first, first+count) of `items`

struct SubBuilder {
    const std::vector<Box>& pb;          // bound per item
    std::vector<int>& items;             // permuted in place
    std::vector<SubNode> nodes;
    int maxLeaf = RT_SUB_LEAF;
    SubBuilder(const std::vector<Box>& b, std::vector<int>& it) : pb(b), items(it) {}

    // The builder works on records (item, bound) that travel with the permutation, so that every sweep over a node's
    // items is a sequential read; the centroid of a bound is recomputed where it is needed (same expression, same value).
    struct Rec { int item; Box b; };
    static constexpr int NB = 16;
    static int bin_of(double c, double lo, double ext) {
        int k = (int)((c - lo) / ext * NB);
        return k < 0 ? 0 : k >= NB ? NB - 1 : k;
    }
    static double centroid(const Box& b, int a) { return 0.5 * ((double)b.mn[a] + b.mx[a]); }

    // One node over recs[first, first + count): bounds, binned-SAH split, stable partition.  Returns the split
    // position, or -1 for a leaf.  `out` receives the node's fields.
    int split_node(Rec* recs, Rec* scratch, int first, int count, SubNode& out) const {
        Box bb; bb.reset();
        Box cb; cb.reset();
        for (int i = first; i < first + count; ++i) {
            const Box& b = recs[i].b;
            bb.grow(b);
            const double c[3] = {centroid(b, 0), centroid(b, 1), centroid(b, 2)};
            cb.grow(c);
        }
        out.box = bb; out.first = first; out.count = count; out.left = out.right = -1;
        if (count <= maxLeaf) return -1;
        // binned SAH over the centroid bounds: one sweep over the items fills the bins of all three axes.  Only the
        // bins that received something are reset, accumulated and evaluated: an empty bin changes neither side of a
        // split, and of the candidates between two occupied bins the first one (the lower bin itself) is the one a
        // scan over all NB - 1 candidates would keep — most nodes of a deep tree hold a handful of items.
        Box bins[3][NB]; int cnt[3][NB];
        double lo[3], ext[3]; unsigned occ[3] = {0u, 0u, 0u};
        for (int a = 0; a < 3; ++a) { lo[a] = cb.mn[a]; ext[a] = (double)cb.mx[a] - lo[a]; }
        for (int i = first; i < first + count; ++i) {
            const Box& b = recs[i].b;
            for (int a = 0; a < 3; ++a) {
                if (!(ext[a] > 0)) continue;
                const int k = bin_of(centroid(b, a), lo[a], ext[a]);
                if (!((occ[a] >> k) & 1u)) { bins[a][k].reset(); cnt[a][k] = 0; occ[a] |= 1u << k; }
                bins[a][k].grow(b); cnt[a][k]++;
            }
        }
        int bestAxis = -1, bestSplit = -1; double bestCost = INFINITY;
        for (int a = 0; a < 3; ++a) {
            if (!(ext[a] > 0)) continue;
            int ks[NB], m = 0;
            for (int k = 0; k < NB; ++k) if ((occ[a] >> k) & 1u) ks[m++] = k;
            double rightArea[NB]; int rightCnt[NB];                  // of the occupied bins i .. m-1
            Box acc; acc.reset(); int n = 0;
            for (int i = m - 1; i > 0; --i) { acc.grow(bins[a][ks[i]]); n += cnt[a][ks[i]]; rightArea[i] = acc.area(); rightCnt[i] = n; }
            acc.reset(); n = 0;
            for (int i = 0; i + 1 < m; ++i) {
                acc.grow(bins[a][ks[i]]); n += cnt[a][ks[i]];
                const double cost = acc.area() * n + rightArea[i + 1] * rightCnt[i + 1];
                if (cost < bestCost) { bestCost = cost; bestAxis = a; bestSplit = ks[i]; }
            }
        }
        int mid;
        if (bestAxis < 0) {
            mid = first + count / 2;                                        // identical centroids: split by count
        } else {
            // stable partition: bins <= bestSplit first, both sides in their old order
            const int a = bestAxis;
            int w = first, r = first;
            for (int i = first; i < first + count; ++i) {
                if (bin_of(centroid(recs[i].b, a), lo[a], ext[a]) <= bestSplit) recs[w++] = recs[i];
                else scratch[r++] = recs[i];
            }
            mid = w;
            std::copy(scratch + first, scratch + r, recs + w);
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        return mid;
    }

    // Subtree over recs[first, first + count) into `out` (ids local to `out`), sequentially.
    int build_rec(Rec* recs, Rec* scratch, int first, int count, std::vector<SubNode>& out) const {
        const int id = (int)out.size();
        out.emplace_back();
        SubNode nd;
        const int mid = split_node(recs, scratch, first, count, nd);
        out[(size_t)id] = nd;
        if (mid < 0) return id;
        const int l = build_rec(recs, scratch, first, mid - first, out);
        const int r = build_rec(recs, scratch, mid, first + count - mid, out);
        out[(size_t)id].left = l; out[(size_t)id].right = r;
        return id;
    }

    // The same subtree with its two halves built side by side while `forks` levels of forking are left: the right half goes
    // to a new thread and into a node list of its own (disjoint record ranges, nothing shared but the read-only bounds),
    // which is appended when both are done.  The sweep over a node's items stays sequential, so the critical path is the
    // root's sweep plus half of it, a quarter ... (two sweeps of the whole range) instead of one sweep per level until the
    // ranges were small enough to hand out (round 2's first form: five levels for a million leaves).
    static constexpr int PAR_MIN = 8192;             // halves smaller than this are not worth a thread
    int build_par(Rec* recs, Rec* scratch, int first, int count, std::vector<SubNode>& out, int forks) const {
        if (forks <= 0 || count < 2 * PAR_MIN) return build_rec(recs, scratch, first, count, out);
        const int id = (int)out.size();
        out.emplace_back();
        SubNode nd;
        const int mid = split_node(recs, scratch, first, count, nd);
        out[(size_t)id] = nd;
        if (mid < 0) return id;
        const int nl = mid - first, nr = first + count - mid;
        int l, r;
        std::vector<SubNode> other;
        std::thread th;
        bool forked = false;
        if (nl >= PAR_MIN && nr >= PAR_MIN) {
            try { th = std::thread([&, this] { build_par(recs, scratch, mid, nr, other, forks - 1); }); forked = true; }
            catch (...) {}                               // no more threads to be had: this one does both halves
        }
        l = build_par(recs, scratch, first, nl, out, forked ? forks - 1 : forks);
        if (forked) {
            th.join();
            const int off = (int)out.size();             // local id j of `other`  ->  off + j; its root is its element 0
            out.reserve(out.size() + other.size());
            for (SubNode x : other) {
                if (x.left >= 0) { x.left += off; x.right += off; }
                out.push_back(x);
            }
            r = off;
        } else {
            r = build_par(recs, scratch, mid, nr, out, forks);
        }
        out[(size_t)id].left = l; out[(size_t)id].right = r;
        return id;
    }

    // maxLeaf == 1: every leaf holds exactly one item and every inner node has two non-empty children, so the subtree over
    // `count` items has exactly 2 * count - 1 nodes: ids are handed out in advance (the node itself, then the left subtree's
    // 2 * nl - 1 ids, then the right subtree's) and the halves write into the one pre-sized list without ever moving a node.
    void build_fixed(Rec* recs, Rec* scratch, int first, int count, SubNode* out, int id, int forks) const {
        SubNode nd;
        const int mid = split_node(recs, scratch, first, count, nd);
        if (mid < 0) { out[id] = nd; return; }
        const int nl = mid - first, nr = first + count - mid;
        nd.left = id + 1; nd.right = id + 2 * nl;
        out[id] = nd;
        std::thread th;
        bool forked = false;
        if (forks > 0 && nl >= PAR_MIN && nr >= PAR_MIN) {
            try { th = std::thread([=] { build_fixed(recs, scratch, mid, nr, out, nd.right, forks - 1); }); forked = true; }
            catch (...) {}
        }
        build_fixed(recs, scratch, first, nl, out, nd.left, forked ? forks - 1 : forks);
        if (forked) th.join();
        else build_fixed(recs, scratch, mid, nr, out, nd.right, forks);
    }

    int build(int first, int count);
};

// Large ranges (the tree over a million reference leaves) are split among the host's cores (build_par).  Node ids differ
// from a sequential build, the TREE (boxes, children, item order) is the same, and only that reaches the layout.
inline int SubBuilder::build(int first, int count) {
    if (count <= 0) return -1;
    std::vector<Rec> recs((size_t)count), scratch((size_t)count);
    for (int i = 0; i < count; ++i) { recs[(size_t)i].item = items[(size_t)(first + i)]; recs[(size_t)i].b = pb[(size_t)recs[(size_t)i].item]; }
    const int base = (int)nodes.size();
    int forks = 0;
    if (count >= (1 << 16)) {
        unsigned hw = std::thread::hardware_concurrency();
        const unsigned nThreads = hw == 0 ? 1 : hw > 16 ? 16 : hw;
        while ((1u << forks) < nThreads) ++forks;
        if (nThreads > 1) forks += 2;                    // four times as many halves as cores: the halves are not equally deep
    }
    int root;
    if (forks > 0 && maxLeaf == 1) {
        nodes.resize((size_t)base + 2 * (size_t)count - 1);
        build_fixed(recs.data(), scratch.data(), 0, count, nodes.data(), base, forks);
        root = base;
    } else if (base == 0) root = build_par(recs.data(), scratch.data(), 0, count, nodes, forks);
    else {                                               // ids local to a list of their own, then appended (as a forked half is)
        std::vector<SubNode> mine;
        build_par(recs.data(), scratch.data(), 0, count, mine, forks);
        root = base;
        for (SubNode x : mine) { if (x.left >= 0) { x.left += base; x.right += base; } nodes.push_back(x); }
    }
    // node ranges are relative to `first`; the permutation goes back into items
    if (first != 0) for (size_t j = (size_t)base; j < nodes.size(); ++j) nodes[j].first += first;
    for (int i = 0; i < count; ++i) items[(size_t)(first + i)] = recs[(size_t)i].item;
    return root;
}

// ---------------------------------------------------------------------------
// The layout.
// ---------------------------------------------------------------------------
struct Layout {
    // per emitted node k
    std::vector<int> src;        // raw node index (reference node) or -1-j for sub node j (box = subBox[j])
    std::vector<int> link;       // inner: first node after the subtree; leaf: first primitive slot
    std::vector<int> cw;         // count word (see RTL_* above)
    // per primitive slot
    std::vector<int> pshape;     // original shape index
    std::vector<int> prank;      // position in the reference's primitive visit sequence
    // sub nodes
    std::vector<Box> subBox;     // box of sub node j
    std::vector<int> subPos;     // emitted position of sub node j
    double omax = 0;             // the margins hold for ray origins with |o|_inf <= omax
    int nRefLeavesFat = 0;
    // reference leaves ("entries") in visit order: position of their node (leaf, or fat leaf turned inner) in the list
    std::vector<int> entryPos, entryRaw;
    // Layout B (see build_top): our own SAH tree over the reference LEAVES, used for well-behaved rays
    std::vector<int> srcB, linkB, cwB;
    std::vector<int> topPos;     // position in B of top node t; its box is subBox[nSubA + t]
    int nSubA = 0;               // sub boxes [0, nSubA) belong to the fat-leaf sub-trees, the rest to B's top nodes
    double costA = 0, costB = 0; // SAH estimates (expected box tests per random ray) of the two node lists
    bool useB = false;
    // refit graph (build_refit_graph): what the device needs to recompute every box of subBox bottom-up
    std::vector<int> rfParent;   // per subBox entry: the entry whose box is the union containing it, -1 = none
    std::vector<int> rfChild;    // 2 per entry: >= 0 = entry, < 0 = raw reference node -1-x; unused for sub-tree leaves
    std::vector<int> rfLeaf;     // 2 per entry: first primitive slot, count (count 0 = inner node)
    std::vector<int> rfStart;    // work items: >= 0 = sub-tree leaf entry; < 0 = a raw child arriving at entry -1-x
    const char* error = nullptr;

    int nodes() const { return (int)src.size(); }
    int prims() const { return (int)pshape.size(); }
};

inline int leaf_kind(const rt_flat_shape* shapes, const int* slots, int n) {
    if (n == 0) return KIND_MIXED;
    int nTri = 0, nSph = 0;
    for (int i = 0; i < n; ++i) { const int t = shapes[slots[i]].type; nTri += t == RT_SHAPE_TRIANGLE; nSph += t == RT_SHAPE_SPHERE; }
    return nTri == n ? KIND_TRIANGLES : nSph == n ? KIND_SPHERES : KIND_MIXED;
}

// Builds the layout.  subtrees = false reproduces the plain reference walk (every leaf kept).
// Returns false and sets L.error on a malformed BVH (what the shader would blindly trust).
inline bool build_layout(const rt_flat_node* nd, int nNodes, const int32_t* indices, size_t nIndices,
                         const rt_flat_shape* shapes, int nShapes, bool subtrees, double omax, Layout& L) {
    L = Layout();
    L.omax = omax;
    if (nNodes <= 0) return true;
    struct Item { int node; int depth; int sub; int parent; };       // node >= 0: raw node; node == -1: entry `sub` of the pending sub list
    // pending sub-tree emissions are described by (builder index, sub node id)
    struct Pending { int kind; int a, b, c; };            // kind 0: always-leaf (slot range a..a+b), kind 1: sub node (tree a, node b)
    std::vector<Pending> pend;
    std::vector<SubBuilder*> trees;
    std::vector<std::vector<int>*> treeItems;
    std::vector<std::vector<Box>*> treeBoxes;
    std::vector<int> treeSlot0;                           // first slot of the tree's bounded primitives
    std::vector<Item> st; st.push_back({nNodes - 1, 1, -1, -1});
    std::vector<int> parent;                              // emitted position of each emitted node's parent
    L.src.reserve((size_t)nNodes); L.link.reserve((size_t)nNodes); L.cw.reserve((size_t)nNodes); parent.reserve((size_t)nNodes);
    L.pshape.reserve(nIndices); L.prank.reserve(nIndices);
    L.entryPos.reserve((size_t)nNodes / 2 + 1); L.entryRaw.reserve((size_t)nNodes / 2 + 1);
    const size_t limit = (size_t)nNodes * 4 + 64;
    size_t refEmitted = 0;
    int rank = 0;
    bool ok = true;
    auto fail = [&](const char* e) { L.error = e; ok = false; };
    while (!st.empty() && ok) {
        const Item it = st.back(); st.pop_back();
        const int k = L.nodes();
        parent.push_back(it.parent);
        if (it.node >= 0) {
            if (it.node >= nNodes) { fail("BVH child index out of range"); break; }
            if (++refEmitted > limit) { fail("BVH is cyclic or shares subtrees excessively"); break; }
            const rt_flat_node& x = nd[it.node];
            if (x.leftChild != -1) {
                if (it.depth + 1 > 64) { fail("BVH deeper than the reference's 64-entry stack"); break; }
                if (x.leftChild < 0 || x.rightChild < 0) { fail("BVH child index out of range"); break; }
                L.src.push_back(it.node); L.link.push_back(0); L.cw.push_back(RTL_INNER_REF);
                st.push_back({x.leftChild, it.depth + 1, -1, k});
                st.push_back({x.rightChild, it.depth + 1, -1, k});
                continue;
            }
            if (x.numShapes < 0 || x.startShapeIdx < 0 || (size_t)x.startShapeIdx + (size_t)x.numShapes > nIndices) { fail("leaf index range outside bvhIndices"); break; }
            if (x.numShapes > RTL_COUNT_MASK) { fail("leaf with more than 2^28 shapes"); break; }
            if ((size_t)L.prims() + (size_t)x.numShapes > 0x7fffffffu) { fail("too many leaf primitives"); break; }
            const int n = x.numShapes, slot0 = L.prims();
            for (int i = 0; i < n; ++i) {
                const int s = indices[(size_t)x.startShapeIdx + (size_t)i];
                if (s < 0 || s >= nShapes) { fail("bvhIndices entry outside the shape buffer"); break; }
                L.pshape.push_back(s); L.prank.push_back(rank++);
            }
            if (!ok) break;
            if (!subtrees || n <= RT_FAT_LEAF) {
                L.entryPos.push_back(k); L.entryRaw.push_back(it.node);
                L.src.push_back(it.node); L.link.push_back(slot0);
                L.cw.push_back(n | (leaf_kind(shapes, L.pshape.data() + slot0, n) << RTL_KIND_SHIFT));
                continue;
            }
            // fat leaf: partition its slots into [unbounded | bounded], build the sub-tree over the bounded ones
            std::vector<Box>* boxes = new std::vector<Box>((size_t)n);
            std::vector<int>* items = new std::vector<int>();
            std::vector<int> unb;
            for (int i = 0; i < n; ++i) {
                if (prim_bound(shapes[L.pshape[(size_t)(slot0 + i)]], omax, (*boxes)[(size_t)i])) items->push_back(i);
                else unb.push_back(i);
            }
            SubBuilder* sb = new SubBuilder(*boxes, *items);
            if (!items->empty()) sb->build(0, (int)items->size());
            // rewrite the slot range in the new order
            std::vector<int> ps((size_t)n), pr((size_t)n);
            int w = 0;
            for (int i : unb) { ps[(size_t)w] = L.pshape[(size_t)(slot0 + i)]; pr[(size_t)w] = L.prank[(size_t)(slot0 + i)]; ++w; }
            for (int i : *items) { ps[(size_t)w] = L.pshape[(size_t)(slot0 + i)]; pr[(size_t)w] = L.prank[(size_t)(slot0 + i)]; ++w; }
            std::copy(ps.begin(), ps.end(), L.pshape.begin() + slot0);
            std::copy(pr.begin(), pr.end(), L.prank.begin() + slot0);
            // boxes re-indexed by position inside the bounded range
            std::vector<Box> nb(items->size());
            for (size_t j = 0; j < items->size(); ++j) nb[j] = (*boxes)[(size_t)(*items)[j]];
            boxes->swap(nb);
            const int t = (int)trees.size();
            trees.push_back(sb); treeItems.push_back(items); treeBoxes.push_back(boxes); treeSlot0.push_back(slot0 + (int)unb.size());
            L.nRefLeavesFat++;
            L.entryPos.push_back(k); L.entryRaw.push_back(it.node);
            // the reference leaf becomes an inner node with the reference box (exact reference test)
            L.src.push_back(it.node); L.link.push_back(0); L.cw.push_back(RTL_INNER_REF);
            // children (popped in reverse push order): always-leaf first, then the sub-tree root
            if (!items->empty()) { pend.push_back({1, t, 0, 0}); st.push_back({-1, it.depth + 1, (int)pend.size() - 1, k}); }
            if (!unb.empty() || items->empty()) { pend.push_back({0, slot0, (int)unb.size(), it.node}); st.push_back({-1, it.depth + 1, (int)pend.size() - 1, k}); }
            continue;
        }
        const Pending p = pend[(size_t)it.sub];
        if (p.kind == 0) {      // always-tested leaf: carries the reference leaf's own box and test
            L.src.push_back(p.c); L.link.push_back(p.a);
            L.cw.push_back(p.b | (leaf_kind(shapes, L.pshape.data() + p.a, p.b) << RTL_KIND_SHIFT));
            continue;
        }
        const SubNode& sn = trees[(size_t)p.a]->nodes[(size_t)p.b];
        const int j = (int)L.subBox.size();
        L.subBox.push_back(sn.box); L.subPos.push_back(k);
        L.src.push_back(-1 - j);
        if (sn.left < 0) {
            const int s0 = treeSlot0[(size_t)p.a] + sn.first;
            L.link.push_back(s0);
            L.cw.push_back(sn.count | (leaf_kind(shapes, L.pshape.data() + s0, sn.count) << RTL_KIND_SHIFT) | RTL_SUB_BIT);
        } else {
            L.link.push_back(0); L.cw.push_back(RTL_INNER_SUB);
            pend.push_back({1, p.a, sn.right, 0}); st.push_back({-1, it.depth + 1, (int)pend.size() - 1, k});
            pend.push_back({1, p.a, sn.left, 0});  st.push_back({-1, it.depth + 1, (int)pend.size() - 1, k});
        }
    }
    for (auto* p : trees) delete p;
    for (auto* p : treeItems) delete p;
    for (auto* p : treeBoxes) delete p;
    if (!ok) return false;
    // skip links of inner nodes: pre-order => link[k] = k + size of the subtree rooted at k
    const int m = L.nodes();
    std::vector<int> size((size_t)m, 1);
    for (int k = m - 1; k >= 1; --k) size[(size_t)parent[(size_t)k]] += size[(size_t)k];
    for (int k = 0; k < m; ++k) if (L.cw[(size_t)k] < 0) L.link[(size_t)k] = k + size[(size_t)k];
    L.nSubA = (int)L.subBox.size();
    return true;
}


// Recompute the sub-tree boxes after shapes moved (rt_update_shapes) or omax changed; topology and
// slot order stay.  Returns false when a primitive below a sub-tree is no longer boundable (it has
// to move to the always-tested leaf): the caller then rebuilds the layout.
inline bool refit_subboxes(Layout& L, const rt_flat_shape* shapes, double omax) {
    L.omax = omax;
    for (int j = L.nSubA - 1; j >= 0; --j) {
        const int k = L.subPos[(size_t)j];
        Box b; b.reset();
        if (L.cw[(size_t)k] >= 0) {
            const int n = L.cw[(size_t)k] & RTL_COUNT_MASK, s0 = L.link[(size_t)k];
            for (int i = 0; i < n; ++i) {
                Box pb;
                if (!prim_bound(shapes[L.pshape[(size_t)(s0 + i)]], omax, pb)) return false;
                b.grow(pb);
            }
        } else {
            const int c1 = k + 1;
            const int c2 = L.cw[(size_t)c1] < 0 ? L.link[(size_t)c1] : c1 + 1;
            b.grow(L.subBox[(size_t)(-1 - L.src[(size_t)c1])]);
            b.grow(L.subBox[(size_t)(-1 - L.src[(size_t)c2])]);
        }
        L.subBox[(size_t)j] = b;
    }
    return true;
}


// ---------------------------------------------------------------------------
// Layout B: our own tree over the reference LEAVES.
//
// When every reference node's box contains its children's boxes coordinate-wise (true for the
// reference builder, whose boxes are unions; checked here), the fp32 slab test is monotone for a
// ray with finite origin and finite 1/d: a ray that passes a leaf's box test passes the test of
// every box containing it.  The reference then tests a leaf's primitives iff the ray passes the
// LEAF's own box test, and any tree of exact union boxes over the leaves selects the same leaves.
// B is a binned-SAH tree over the leaf boxes (the reference's midpoint-split tree is shallow on
// quality: the 1M-triangle soup costs ~290 box tests per ray through it), entries copied from A
// (fat-leaf sub-trees included), candidates ordered by rank as before.  Rays with a non-finite
// origin or 1/d (NaN semantics are not monotone) and rays beyond omax walk list A.
// ---------------------------------------------------------------------------
inline Box raw_box(const rt_flat_node& x) {
    Box b;
    for (int a = 0; a < 3; ++a) { b.mn[a] = x.boundsMin[a]; b.mx[a] = x.boundsMax[a]; }
    return b;
}

inline bool ref_tree_nested(const rt_flat_node* nd, int nNodes) {
    for (int i = 0; i < nNodes; ++i) {
        const rt_flat_node& x = nd[i];
        for (int a = 0; a < 3; ++a) if (!std::isfinite(x.boundsMin[a]) || !std::isfinite(x.boundsMax[a])) return false;
        if (x.leftChild == -1) continue;
        const int ch[2] = {x.leftChild, x.rightChild};
        for (int c : ch) {
            if (c < 0 || c >= nNodes) return false;
            for (int a = 0; a < 3; ++a) {
                // an EMPTY child box (min > max: a leaf without shapes keeps the +inf/-inf initial box) never passes its own test
                if (!(nd[c].boundsMin[a] >= x.boundsMin[a]) || !(nd[c].boundsMax[a] <= x.boundsMax[a])) return false;
            }
        }
    }
    return true;
}

// What a node upload has to know about the new nodes, in ONE pass without early exits (the three separate passes — the
// topology compare, the |bound| <= 1e37 check and ref_tree_nested — were 105 of the 176 us that rt_upload_scene spent on
// the host for the car's 3743 nodes; this pass: ~10 us):
//   sameTopo  child links of every node, and the index range of every leaf, equal those of `old` (old == NULL: false)
//   safe      every bound is a number with |v| <= safeCoord (the NaN-free slab test is legal)
//   nested    ref_tree_nested(in, n)
// Read-only on purpose: with the copy into the shadow in the same loop the loads of in[i + 1] run into the stores to
// dst[i] (same page offsets, 4K aliasing) and the pass is slower than the three it replaces; memcpy afterwards is not.
struct NodeScan { bool sameTopo, safe, nested; };
inline NodeScan scan_nodes(const rt_flat_node* in, const rt_flat_node* old, int n, float safeCoord) {
    unsigned topoBad = old ? 0u : 1u, notNested = 0u;
    uint32_t maxAbs = 0u, safeBits;
    std::memcpy(&safeBits, &safeCoord, 4);
    for (int i = 0; i < n; ++i) {
        const rt_flat_node& x = in[i];
        if (old) {
            const rt_flat_node& o = old[i];
            const unsigned leaf = x.leftChild == -1;
            topoBad |= (unsigned)(x.leftChild != o.leftChild) | (unsigned)(x.rightChild != o.rightChild) |
                       (leaf & ((unsigned)(x.startShapeIdx != o.startShapeIdx) | (unsigned)(x.numShapes != o.numShapes)));
        }
        // |v| as an integer: the order of the bit patterns is the order of the magnitudes, inf and every NaN above all numbers
        for (int a = 0; a < 3; ++a) {
            uint32_t lo, hi;
            std::memcpy(&lo, &x.boundsMin[a], 4); std::memcpy(&hi, &x.boundsMax[a], 4);
            lo &= 0x7fffffffu; hi &= 0x7fffffffu;
            maxAbs = lo > maxAbs ? lo : maxAbs; maxAbs = hi > maxAbs ? hi : maxAbs;
        }
        if (x.leftChild != -1) {
            const int ch[2] = {x.leftChild, x.rightChild};
            for (int c : ch) {
                const bool in_range = c >= 0 && c < n;
                notNested |= (unsigned)!in_range;
                const rt_flat_node& y = in[in_range ? c : i];
                for (int a = 0; a < 3; ++a)
                    notNested |= (unsigned)!(y.boundsMin[a] >= x.boundsMin[a]) | (unsigned)!(y.boundsMax[a] <= x.boundsMax[a]);
            }
        }
    }
    NodeScan r;
    r.sameTopo = topoBad == 0u; r.safe = maxAbs <= safeBits; r.nested = notNested == 0u && maxAbs < 0x7f800000u;
    return r;
}

// Recompute the boxes of B's top nodes from the current raw node boxes (bottom-up) and the SAH estimates.
inline void refit_top(Layout& L, const rt_flat_node* nd) {
    const int mB = (int)L.srcB.size();
    auto boxOf = [&](int k) -> Box { const int s = L.srcB[(size_t)k]; return s >= 0 ? raw_box(nd[s]) : L.subBox[(size_t)(-1 - s)]; };
    for (int t = (int)L.topPos.size() - 1; t >= 0; --t) {
        const int k = L.topPos[(size_t)t];
        const int c1 = k + 1, c2 = L.cwB[(size_t)c1] < 0 ? L.linkB[(size_t)c1] : c1 + 1;
        Box b; b.reset();
        b.grow(boxOf(c1)); b.grow(boxOf(c2));
        L.subBox[(size_t)(L.nSubA + t)] = b;
    }
    // SAH estimate: sum of node areas relative to the root (every node whose parent is hit costs one test)
    auto cost = [&](const std::vector<int>& src, int n) {
        if (n == 0) return 0.0;
        auto bx = [&](int k) { const int s = src[(size_t)k]; return s >= 0 ? raw_box(nd[s]) : L.subBox[(size_t)(-1 - s)]; };
        const double root = bx(0).area();
        double c = 0;
        for (int k = 0; k < n; ++k) c += bx(k).area();
        return root > 0 ? c / root : 0.0;
    };
    L.costA = cost(L.src, L.nodes());
    L.costB = cost(L.srcB, mB);
}

inline void build_top(Layout& L, const rt_flat_node* nd, int nNodes) {
    L.srcB.clear(); L.linkB.clear(); L.cwB.clear(); L.topPos.clear();
    L.subBox.resize((size_t)L.nSubA);
    L.useB = false;
    const int nE = (int)L.entryPos.size();
    if (nE < 4 || !ref_tree_nested(nd, nNodes)) return;
    std::vector<Box> eb((size_t)nE);
    std::vector<int> items((size_t)nE);
    for (int e = 0; e < nE; ++e) { eb[(size_t)e] = raw_box(nd[L.entryRaw[(size_t)e]]); items[(size_t)e] = e; }
    // empty leaves (min > max) can never be hit: leave them out of B
    items.erase(std::remove_if(items.begin(), items.end(), [&](int e) {
        const Box& b = eb[(size_t)e];
        return b.mn[0] > b.mx[0] || b.mn[1] > b.mx[1] || b.mn[2] > b.mx[2];
    }), items.end());
    if (items.size() < 4) return;
    SubBuilder sb(eb, items);
    sb.maxLeaf = 1;
    sb.build(0, (int)items.size());
    // emit B in pre-order with an explicit stack; a top inner node's link is patched when its subtree is complete
    struct Item { int node; int patch; };      // node >= 0: visit; node < 0: patch link of B position `patch`
    {   // B holds A's entries (every node of A that is not a reference inner node, at most all of A) plus one top node per inner node of the tree
        const size_t cap = L.src.size() + sb.nodes.size() / 2 + 1;
        L.srcB.reserve(cap); L.linkB.reserve(cap); L.cwB.reserve(cap);
        L.topPos.reserve(sb.nodes.size() / 2 + 1); L.subBox.reserve(L.subBox.size() + sb.nodes.size() / 2 + 1);
    }
    std::vector<Item> st; st.push_back({0, -1});
    while (!st.empty()) {
        const Item it = st.back(); st.pop_back();
        if (it.node < 0) { L.linkB[(size_t)it.patch] = (int)L.srcB.size(); continue; }
        const SubNode& sn = sb.nodes[(size_t)it.node];
        const int k = (int)L.srcB.size();
        if (sn.left < 0) {                       // one reference leaf: copy its nodes from A
            const int e = items[(size_t)sn.first];
            const int p0 = L.entryPos[(size_t)e];
            const int p1 = L.cw[(size_t)p0] < 0 ? L.link[(size_t)p0] : p0 + 1;
            const int delta = k - p0;
            for (int q = p0; q < p1; ++q) {
                L.srcB.push_back(L.src[(size_t)q]); L.cwB.push_back(L.cw[(size_t)q]);
                L.linkB.push_back(L.cw[(size_t)q] < 0 ? L.link[(size_t)q] + delta : L.link[(size_t)q]);
            }
            continue;
        }
        const int t = (int)L.topPos.size();
        L.topPos.push_back(k);
        L.subBox.emplace_back();
        L.srcB.push_back(-1 - (L.nSubA + t)); L.cwB.push_back(RTL_INNER_REF); L.linkB.push_back(0);
        st.push_back({-1, k});
        st.push_back({sn.right, -1});
        st.push_back({sn.left, -1});
    }
    refit_top(L, nd);
    L.useB = L.costB < 0.9 * L.costA;
    if (!L.useB) { L.srcB.clear(); L.linkB.clear(); L.cwB.clear(); L.topPos.clear(); L.subBox.resize((size_t)L.nSubA); }
}

// The dependency graph of the boxes in subBox, for the device refit (rt_kernels.cuh: refit_kernel):
// a sub-tree leaf's box is the union of its primitives' bounds, every other entry the union of its
// two children (entries, or raw reference nodes under B's top nodes).  Call after build_layout / build_top.
inline void build_refit_graph(Layout& L) {
    const int nE = (int)L.subBox.size();
    L.rfParent.assign((size_t)nE, -1); L.rfChild.assign((size_t)nE * 2, 0); L.rfLeaf.assign((size_t)nE * 2, 0); L.rfStart.clear();
    for (int j = 0; j < L.nSubA; ++j) {
        const int k = L.subPos[(size_t)j];
        if (L.cw[(size_t)k] >= 0) {
            L.rfLeaf[2 * (size_t)j] = L.link[(size_t)k]; L.rfLeaf[2 * (size_t)j + 1] = L.cw[(size_t)k] & RTL_COUNT_MASK;
            L.rfStart.push_back(j);
        } else {
            const int c1 = k + 1, c2 = L.cw[(size_t)c1] < 0 ? L.link[(size_t)c1] : c1 + 1;
            const int e1 = -1 - L.src[(size_t)c1], e2 = -1 - L.src[(size_t)c2];
            L.rfChild[2 * (size_t)j] = e1; L.rfChild[2 * (size_t)j + 1] = e2;
            L.rfParent[(size_t)e1] = j; L.rfParent[(size_t)e2] = j;
        }
    }
    for (int t = 0; t < (int)L.topPos.size(); ++t) {
        const int j = L.nSubA + t, k = L.topPos[(size_t)t];
        const int c1 = k + 1, c2 = L.cwB[(size_t)c1] < 0 ? L.linkB[(size_t)c1] : c1 + 1;
        const int cs[2] = {c1, c2};
        for (int i = 0; i < 2; ++i) {
            const int s = L.srcB[(size_t)cs[i]];
            if (s >= 0) { L.rfChild[2 * (size_t)j + i] = -1 - s; L.rfStart.push_back(-1 - j); }
            else        { L.rfChild[2 * (size_t)j + i] = -1 - s; L.rfParent[(size_t)(-1 - s)] = j; }
        }
    }
}

}  // namespace rtl

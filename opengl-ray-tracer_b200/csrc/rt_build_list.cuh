// rt_build_list.cuh — K9: list A of the device scene layout straight from K8's node table, for reference trees whose
// leaves are all small (no fat-leaf sub-trees), or for the plain reference walk (RT_FLAG_REFERENCE_WALK) of any tree.
// Launched by rt_build_bvh(..., RT_BUILD_DEVICE_LAYOUT): the tree K8 has just built is laid out where it is, no host
// layout call and no copy of the tree to the host (a tree with a fat leaf falls back to the host layout builder).
//
// rt_layout.h builds the threaded node list on the host: the reference nodes in the reference's visit order (root
// first, right child before left, gpu_shader.comp:386, 424-425), a skip link per inner node, and the leaves' shapes in
// visit order as primitive slots with their rank.  On a large scene that is the dominant cost of a topology change
// (soup-1M 0.23 s for list A against 2.5 ms for K8).  For a tree without fat leaves the list is a pure
// renumbering that every node can do on its own from K8's tables, exactly like number_node does for the reference's
// push order:
//     position(root) = 0;  position(right child) = position(parent) + 1;
//     position(left child) = position(parent) + 1 + size(right sibling's subtree)
//     first slot(right child) = first slot(parent);  first slot(left child) = first slot(parent) + shapes under the right sibling
//     link(inner) = position + size of its own subtree;  link(leaf) = first slot;  count word as rtl::build_layout writes it
// The per-thread body below is compiled by tests/emu/build_emu.cpp and compared word for word with
// rtl::build_layout(subtrees = false) in tests/test_build_emu.py.
#pragma once
#include "rt_build.cuh"

namespace rt {
namespace bld {

// out_src[pos] = reference node index, out_linkcw[pos] = (link, count word), pshape / prank per slot
RT_DEV void thread_node(int id, const Nodes& N, const int* ord, const rt_flat_shape* __restrict__ shapes,
                        int* out_src, int2* out_linkcw, int* pshape, int* prank) {
    if (N.left[id] == RT_BLD_ABSENT) return;
    int pos = 0, base = 0;
    for (int a = id; a != 0; a = N.parent[a]) {
        const int pa = N.parent[a], l = N.left[pa];
        pos += 1;
        if (a == l) { pos += N.desc[l + 1] + 1; base += N.count[l + 1]; }      // a left child comes after its right sibling's subtree
    }
    out_src[pos] = N.idx[id];
    if (N.left[id] >= 0) { out_linkcw[pos] = make_int2(pos + N.desc[id] + 1, RT_INNER_REF); return; }
    const int st = N.start[id], cnt = N.count[id];
    int nTri = 0, nSph = 0;
    for (int i = 0; i < cnt; ++i) {
        const int s = ord[st + i];
        pshape[base + i] = s; prank[base + i] = base + i;
        const int t = shapes[s].type;
        nTri += t == RT_SHAPE_TRIANGLE; nSph += t == RT_SHAPE_SPHERE;
    }
    const int kind = cnt == 0 ? RT_LEAF_MIXED : nTri == cnt ? RT_LEAF_TRIANGLES : nSph == cnt ? RT_LEAF_SPHERES : RT_LEAF_MIXED;
    out_linkcw[pos] = make_int2(base, cnt | (kind << RT_LEAF_KIND_SHIFT));
}

}  // namespace bld

#ifdef __CUDACC__
// One thread per build id.  Besides the list it gathers what the host needs to decide whether the list can be used as
// it is: the largest leaf, whether every bound is finite (the NaN-free slab test), the root's box (omax).
__global__ void bld_list_kernel(BuildBufs B, int nIds, const rt_flat_shape* __restrict__ shapes, const rt_flat_node* __restrict__ flat,
                                int* __restrict__ outSrc, int2* __restrict__ outLinkCw, int* __restrict__ pshape, int* __restrict__ prank) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= nIds || B.N.left[id] == RT_BLD_ABSENT) return;
    bld::thread_node(id, B.N, B.ord[B.ctl->parity], shapes, outSrc, outLinkCw, pshape, prank);
    if (B.N.left[id] < 0) atomicMax(&B.ctl->maxLeaf, B.N.count[id]);
    const rt_flat_node& f = flat[B.N.idx[id]];
    bool fin = true;
    for (int k = 0; k < 3; ++k) fin = fin && fabsf(f.boundsMin[k]) <= RT_SAFE_COORD && fabsf(f.boundsMax[k]) <= RT_SAFE_COORD;
    if (!fin) atomicOr(&B.ctl->notFinite, 1);
    if (id == 0) for (int k = 0; k < 3; ++k) { B.ctl->rootBox[k] = f.boundsMin[k]; B.ctl->rootBox[3 + k] = f.boundsMax[k]; }
}
#endif

}  // namespace rt

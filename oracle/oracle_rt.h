/*
 * oracle_rt.h — C interface of the CPU oracle (TEST INFRASTRUCTURE ONLY).
 *
 * The oracle is a scalar fp32 restatement of the reference's compute shader
 * (src/shaders/gpu_shader.comp).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it; the product
 * library librt_b200.so never links or calls it.
 */
#ifndef ORACLE_RT_H
#define ORACLE_RT_H

#include "../include/rt_api.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Deterministic per-frame work counters: the reference
 * traversal is unordered and unculled, so these are a pure function of the
 * scene bytes and the params. */
typedef struct oracle_counters {
    uint64_t rays;          /* intersectScene2-equivalent queries (primary+shadow+reflection) */
    uint64_t primary_rays;
    uint64_t shadow_rays;
    uint64_t reflect_rays;
    uint64_t node_visits;   /* nodes popped, incl. those failing the slab test               */
    uint64_t index_reads;   /* bvhIndices reads                                              */
    uint64_t tests[4];      /* get_intersection calls by shape type                          */
    uint64_t tri_stage2;    /* Barycentric triangles that passed the plane stage             */
    uint64_t ray_hits;      /* queries that returned INNER                                   */
    uint64_t alg_bytes;     /* 48*node_visits + 4*index_reads + sum b(type) + 48*ray_hits    */
} oracle_counters;

/* Render rows y with y % row_step == row_phase (row_step=1,row_phase=0: all
 * rows; other rows of the outputs are left untouched) using `threads` worker
 * threads.  rgba: W*H*4 floats; ids: W*H int32 primary-hit shape index (-1 =
 * miss) or NULL; dists: W*H floats primary closestDist (1e20 on miss) or NULL;
 * counters may be NULL.  Returns 0, or -1 on bad arguments. */
int oracle_render(const rt_flat_light* light, const rt_flat_camera* camera,
                  const rt_flat_shape* shapes, int nShapes,
                  const rt_flat_node* nodes, int nNodes,
                  const int32_t* indices, int nIndices,
                  const rt_params* params, int W, int H,
                  int row_step, int row_phase,
                  float* rgba, int32_t* ids, float* dists,
                  oracle_counters* counters, int threads);

/* Single primitive test, for KATs: returns intersect_type (0 NONE, 1 INNER,
 * 2 OUTER) and, for INNER, the hit point. */
int oracle_intersect_shape(const rt_flat_shape* shape, const float start[3], const float dir[3],
                           int useMollerTrumbore, float hit_point[3]);

/* Single BVH query (intersectScene2): returns shape index or -1. */
int oracle_intersect_scene(const rt_flat_shape* shapes, int nShapes,
                           const rt_flat_node* nodes, int nNodes,
                           const int32_t* indices, int nIndices,
                           const float start[3], const float dir[3], int useMollerTrumbore,
                           float hit_point[3], float hit_normal[3], float* dist);

/* The same query for n rays (starts/dirs: n*3 floats).  out_dist is closestDist (1e20 on a miss). */
void oracle_intersect_scene_batch(const rt_flat_shape* shapes, int nShapes, const rt_flat_node* nodes, int nNodes,
                                  const int32_t* indices, int nIndices, int n, const float* starts, const float* dirs,
                                  int useMollerTrumbore, int32_t* out_shape, float* out_dist, float* out_point);

/* Tie pixels of the sampled rows: primary rays for which a candidate of a DIFFERENT shape lies within 1 ulp of the
 * winning closestDist (gpu_shader.comp:410-411 resolves those by visit order).  ties: W*H bytes (0/1) or NULL.
 * Returns the number of tie pixels, or -1 on bad arguments. */
int64_t oracle_primary_ties(const rt_flat_camera* camera, const rt_flat_shape* shapes, int nShapes,
                            const rt_flat_node* nodes, int nNodes, const int32_t* indices, int nIndices,
                            const rt_params* params, int W, int H, int row_step, int row_phase,
                            uint8_t* ties, int threads);

/* K4 restatement: clamp [0,1], alpha=1, quantise to RGBA8 (round to nearest). */
void oracle_present(const float* rgba, uint8_t* rgba8, int64_t pixels);

#ifdef __cplusplus
}
#endif
#endif

/*
 * rt_api.h — C ABI of librt_b200.so, the B200-native replacement for the
 * reference's compute-shader dispatch.
 *
 * What it replaces in the reference (Junacik99/opengl-ray-tracer):
 *   - class ComputeShader (src/computeShader.hpp:13-114): load/compile/link of
 *     gpu_shader.comp, use(), setBool/setInt/setFloat/setVec2 uniform setters;
 *   - the SSBO/texture plumbing in src/main.cpp: output image creation
 *     (:182-191), the five glGenBuffers/glBufferData/glBindBufferBase blocks
 *     (:238-275, bindings 1..5 = Light, Camera, Shapes, Nodes, Indices),
 *     the per-frame glBufferSubData updates (:328-348, :981-992) and the
 *     dispatch glDispatchCompute(W,H,1) + glMemoryBarrier + uniforms (:352-361);
 *   - gpu_shader.comp:433-624 itself (the per-pixel program).
 *
 * Conventions: every entry point returns RT_OK (0) or a negative rt_status and
 * never prints or aborts (the reference only prints, computeShader.hpp:64-79).
 * The caller keeps ownership of every host pointer; data is copied into pinned
 * staging before the call returns.  A context is single-threaded, like a GL
 * context.  All device work runs on one internal stream.  There is NO CPU
 * fallback: without a CUDA device rt_create fails with RT_ERR_CUDA.
 *
 * Plain C: pointers and sizes only, no torch / glm / STL types.
 */
#ifndef RT_API_H
#define RT_API_H

#include "rt_flat.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_ctx rt_ctx;

typedef enum rt_status {
    RT_OK            =  0,
    RT_ERR_ARG       = -1,  /* NULL / out-of-range argument                          */
    RT_ERR_CUDA      = -2,  /* a CUDA runtime call failed (see rt_last_error)         */
    RT_ERR_SCENE     = -3,  /* malformed BVH: bad child/shape index, cycle, depth>64  */
    RT_ERR_STATE     = -4,  /* call order: render before upload, etc.                 */
    RT_ERR_NOMEM     = -5
} rt_status;

/* The five uniforms of gpu_shader.comp:126-130 (set at src/main.cpp:357-361).
 * Unlike the reference, which sets them after the dispatch (Appendix B of
 * SURVEY.md), they are passed with the render call. */
typedef struct rt_params {
    float screenRes[2];      /* uniform vec2 screenRes  (W, H)                 */
    int   maxBounces;        /* uniform int  maxBounces (GUI range 1..10)      */
    int   useBVH;            /* uniform bool useBVH                            */
    int   useFresnel;        /* uniform bool useFresnel                        */
    int   useMollerTrumbore; /* uniform bool useMollerTrumbore                 */
} rt_params;

/* rt_create flags */
enum {
    RT_FLAG_DEFAULT   = 0u,
    /* Walk exactly the reference's node and primitive sequence.  By default the
     * reference's fat leaves (its builder stops splitting when one side of a
     * midpoint split is empty, src/main.cpp:1156-1161) are expanded at upload
     * into private sub-trees whose boxes are conservative bounds of each
     * primitive's fp32 acceptance region, so no hit/miss decision changes. */
    RT_FLAG_REFERENCE_WALK = 1u << 0,
    /* Keep the reference's tree above its leaves.  By default, when the reference
     * boxes are nested (children inside parents), a binned-SAH tree of exact union
     * boxes over the reference LEAVES replaces it for rays with finite origin and
     * 1/d whenever its SAH estimate is >10% cheaper: the fp32 slab test is monotone
     * for such rays, so the same leaves pass. */
    RT_FLAG_NO_TOP_TREE = 1u << 2,
    /* Claim the 8x4-pixel blocks in plain order.  By default rt_render_device records every block's
     * cost and the next frame claims the blocks that were slow first (frame-to-frame coherence), so
     * the long dependent walks of incoherent bounce packets overlap the bulk instead of forming the tail. */
    RT_FLAG_NO_HEAVY_FIRST = 1u << 3,
    /* Diagnostic: start the host copy of a pipelined frame (rt_render_async) as soon as its render kernel has finished.  By
     * default the copy of a single-launch frame starts when the NEXT frame's render kernel is launched, or when somebody waits
     * for the frame, whichever is first: the small launches that open an animated frame run slower under a device-to-host DMA. */
    RT_FLAG_NO_DEFERRED_COPY = 1u << 4,
    /* Diagnostic: rt_animate always runs K6 and the general K7 (every leaf climbs the tree) as two launches.  By default, when
     * the animated leaves and their ancestors are at most 16384 nodes, ONE block moves the shapes and then grows exactly those
     * nodes level by level (the only ones updateBVH can change, main.cpp:1068-1077). */
    RT_FLAG_NO_ONE_BLOCK_UPDATE = 1u << 5,
    /* Record per-pixel primary-hit shape ids (an AOV the reference lacks;
     * needed by the parity criterion).  Off = the id buffer is not written. */
    RT_FLAG_HIT_IDS   = 1u << 1
};

/* Tile geometry used when a frame is sharded over several GPUs: tile k belongs to rank k % world.  Tile k sits in tile
 * row k / tilesX; its column is k % tilesX for one GPU and (k % tilesX + row) % tilesX when sharded (every row rotated by its
 * number, so that a rank's tiles form diagonals rather than vertical stripes).  A rank's compact shard holds its tiles in
 * increasing k, 32x8 pixels each, row-major inside the tile; rt_untile restores the frame. */
#define RT_TILE_W 32
#define RT_TILE_H 8

/* ---- lifetime ------------------------------------------------------------
 * Replaces ComputeShader::ComputeShader + the output-texture creation
 * (computeShader.hpp:29-83, main.cpp:182-191).  device = CUDA ordinal. */
int  rt_create(rt_ctx** out, int width, int height, int device, unsigned flags);
/* Drains the context's streams, frees everything it owns, unpins host frames still registered (rt_host_frame_register)
 * and closes IPC mappings.  The handle is invalid afterwards (NULL is accepted; a second call on the same handle is not). */
void rt_destroy(rt_ctx* ctx);
const char* rt_last_error(const rt_ctx* ctx);

/* ---- scene upload ----------------------------------------------------------
 * Replaces the five glBufferData blocks (main.cpp:238-275).  Any pointer may
 * be NULL = keep what is resident.  shapes/nodes/indices are the reference's
 * FlatShape[]/FlatNode[]/int[] bytes; the root is nodes[nNodes-1]
 * (gpu_shader.comp:386).  The buffers are staged through pinned memory with
 * cudaMemcpyAsync and repacked on the device into float4 SoA.
 * A SHORTER shape buffer passed on its own (nodes and indices NULL) drops the
 * tree the context holds — its indices would point past the end; useBVH frames
 * then hit nothing until nodes / indices are supplied or rt_build_bvh is called. */
int rt_upload_scene(rt_ctx* ctx,
                    const void* light32, const void* camera80,
                    const void* shapes, int nShapes,
                    const void* nodes, int nNodes,
                    const int32_t* indices, int nIndices);

/* Per-frame partial updates (main.cpp:328-334 camera/light, :981-992 one
 * 192-byte sub-update per animated shape, :343-344 all nodes). */
int rt_update_camera(rt_ctx* ctx, const void* camera80);
int rt_update_light(rt_ctx* ctx, const void* light32);
int rt_update_shapes(rt_ctx* ctx, int first, int count, const void* shapes);
int rt_update_nodes(rt_ctx* ctx, const void* nodes, int nNodes);

/* ---- multi-GPU sharding ----------------------------------------------------
 * This context renders only the RT_TILE_W x RT_TILE_H tiles k with
 * k % world == rank, into a compact tile-major buffer (tile j of this rank at
 * offset j*RT_TILE_W*RT_TILE_H pixels).  rank=0, world=1 restores full frames.
 * rt_shard_pixels returns the number of pixels (incl. padding of edge tiles)
 * in this rank's compact buffer. */
int     rt_set_shard(rt_ctx* ctx, int rank, int world);
int64_t rt_shard_pixels(const rt_ctx* ctx);

/* ---- dispatch --------------------------------------------------------------
 * Replaces use() + glDispatchCompute(W,H,1) + glMemoryBarrier + the uniform
 * setters (main.cpp:352-361).  Renders into the context's device image and,
 * if host_rgba != NULL, copies the row-major W*H float4 image
 * (imgOutput, rgba32f) to the host; host_hit_ids (W*H int32, -1 = miss) needs
 * RT_FLAG_HIT_IDS.  Synchronous.  With a shard set (world>1) the host copies
 * are the compact tile-major shard instead. */
int rt_render(rt_ctx* ctx, const rt_params* params, float* host_rgba, int32_t* host_hit_ids);

/* rt_render without the final wait: the frame lands in host_rgba (pinned memory for a truly
 * asynchronous copy) once rt_synchronize(ctx) returns.  Lets the host prepare the next frame's
 * scene update (main.cpp:438-457 animation step) while this one renders and copies.
 * Frames are pipelined over two device images: the copy of frame i overlaps the render of frame i+1 (for a frame of one
 * launch it is issued together with that render kernel, or by the first rt_wait_frames / rt_synchronize / rt_render* call
 * that needs it, so a frame nobody follows up on or waits for stays on the device).
 * rt_wait_frames(ctx, n) blocks until at most n of the issued frames are still in flight (n = 1: the
 * older one is complete and its host buffer may be read / reused; n = 0: all of them). */
int rt_render_async(rt_ctx* ctx, const rt_params* params, float* host_rgba, int32_t* host_hit_ids);

/* The frame as the reference SHOWS it: the compute dispatch followed by the display quad (renderQuad main.cpp:474-501,
 * shader.vert:7-11, shader.frag:8-12: the rgba32f image sampled 1:1, alpha forced to 1, and the default framebuffer's
 * conversion: clamp to [0,1], quantise to 8 bits) — K1 with K5 (rt_present's kernel) behind it band by band, then the copy of
 * W*H*4 BYTES instead of W*H*16: a 3840x2160 frame leaves over PCIe in 0.6 ms instead of 2.4.  host_rgba8: W*H RGBA8 pixels,
 * row-major (the compact tile-major shard with a shard set).  The rgba32f image stays on the device (rt_device_image).
 * _async: as rt_render_async (two frames in flight, rt_wait_frames). */
int rt_render_rgba8(rt_ctx* ctx, const rt_params* params, uint8_t* host_rgba8, int32_t* host_hit_ids);
int rt_render_rgba8_async(rt_ctx* ctx, const rt_params* params, uint8_t* host_rgba8, int32_t* host_hit_ids);
int rt_wait_frames(rt_ctx* ctx, int max_in_flight);

/* Same, but writes into caller-provided DEVICE buffers (e.g. torch tensors)
 * and returns without synchronising: work is ordered on rt_stream(ctx). */
int rt_render_device(rt_ctx* ctx, const rt_params* params, float* dev_rgba, int32_t* dev_hit_ids);

/* Untimed diagnostic render with work counters (the timed kernel carries none):
 * out[0] rays = BVH queries (primary+shadow+reflection), [1] primary, [2] shadow,
 * [3] reflection, [4] BVH nodes visited, [5] primitive tests, [6] closest-hit
 * queries that hit, [7] occluded shadow queries.  host_cost_or_null (one int32 per
 * output pixel) receives nodes visited + primitive tests of that pixel.  Synchronous. */
int rt_render_stats(rt_ctx* ctx, const rt_params* params, uint64_t out[8], int32_t* host_cost_or_null);

/* Walk-efficiency diagnostics of the last rt_render_stats call: out[0..3] for primary segments,
 * out[4..7] for later (bounce) segments: warp iterations in packet mode, warp iterations in lane
 * mode, lanes served in packet mode, lanes served in lane mode. */
int rt_walk_stats(rt_ctx* ctx, uint64_t out[8]);
/* Same call: out[0] = SM cycles of the slowest 8x4-pixel warp item, out[1] = sum over items, out[2] = items. */
int rt_item_stats(rt_ctx* ctx, uint64_t out[3]);

/* ---- fused gather (multi-GPU) -----------------------------------------------
 * Instead of gathering compact shards and untiling them, every rank's render kernel can store
 * its pixels straight into their row-major place of ONE full W*H frame that lives on rank 0,
 * through a peer mapping over NVLink (the exchange step happens inside the render kernel, tile
 * by tile).  Rank 0 calls rt_frame_export (allocates the frame, returns a 64-byte CUDA IPC
 * handle to ship to the other processes and the frame's device pointer); the others call
 * rt_frame_open.  A single-process driver can hand any device pointer to rt_set_frame_target
 * (NULL switches the mode off).  While a frame target is set,
 * rt_render_device(ctx, params, NULL, NULL) writes there; the frame is complete once every rank
 * has synchronised (a barrier is the only collective left).  RULE for this plain mode: a rank may render the NEXT
 * frame into the same target only after the consumer has finished with the current one — a second barrier after the
 * consumer's read, before any rank renders again (bench.py: "release barrier").  The signalled mode below needs neither. */
int rt_set_frame_target(rt_ctx* ctx, float* dev_frame);
int rt_frame_export(rt_ctx* ctx, unsigned char handle[64], float** dev_frame);
int rt_frame_open(rt_ctx* ctx, const unsigned char handle[64]);
/* Rank 0: asynchronous copy of the exported frame to the host (ordered on rt_stream).  In signalled mode:
 * rt_frame_acquire + the copy + rt_frame_release. */
int rt_frame_download(rt_ctx* ctx, float* host_rgba);

/* ---- signalled fused gather: compute + signal, no collective ------------------------------------------------
 * The exported frame is DOUBLE-BUFFERED (frame f goes to buffer f & 1) and carries one completion counter per buffer; every
 * other rank owns one release word.  In signalled mode
 *   - a rank's render kernel stores its tiles into rank 0's buffer over NVLink and its LAST thread block, after a
 *     system-scope fence, adds 1 to that buffer's completion counter on rank 0 (a peer atomic);
 *   - rank 0's rt_frame_acquire makes its stream wait (ONE stream memory operation on a local word, no kernel spins) until
 *     the counter has reached (number of frames of this buffer) * (world - 1), and returns the buffer: work enqueued afterwards
 *     sees the whole frame;
 *   - rank 0's rt_frame_release stores the frame number into every other rank's release word; a rank's render of frame
 *     f + 2 (same buffer) waits in ITS stream for the release of frame f.  Ranks therefore run at most one frame ahead of
 *     the consumer and never overwrite a frame that is being read: no barrier, no NCCL call per frame.
 * Set-up (once): rank 0 rt_frame_export, others rt_frame_open (or rt_set_frame_target in one process); every rank r > 0
 * rt_sync_local (+ rt_sync_export for another process) and rank 0 rt_sync_peer(r, ...) for each of them; then every rank
 * rt_sync_enable(ctx, world).  Every rank must render the same sequence of frames. */
int rt_sync_local(rt_ctx* ctx, unsigned** dev_release_word);
int rt_sync_export(rt_ctx* ctx, unsigned char handle[64]);
int rt_sync_peer(rt_ctx* ctx, int rank, unsigned* dev_release_word_or_null, const unsigned char handle_or_null[64]);
int rt_sync_enable(rt_ctx* ctx, int world_or_0);
int rt_frame_acquire(rt_ctx* ctx, float** dev_frame);
int rt_frame_release(rt_ctx* ctx);
/* Host frame: pin and map a caller-owned host buffer of W*H float4 (for several ranks: one POSIX
 * shared-memory segment that every process maps) and return its device alias.  Passed to
 * rt_set_frame_target, the render kernel stores its pixels straight into host memory over this
 * GPU's own PCIe link while it renders: no device frame, no gather and no separate D2H copy, and
 * with N ranks the N links carry the frame in parallel.  The pixels are visible to the host after
 * rt_synchronize (all ranks).  rt_host_frame_unregister undoes the pinning. */
int rt_host_frame_register(rt_ctx* ctx, void* host_frame, size_t bytes, float** dev_alias);
int rt_host_frame_unregister(rt_ctx* ctx, void* host_frame);

/* ---- device-side animation (flag-gated addition, the host path above stays the default) ----
 * The reference animates on the host and re-uploads (bounceSphere src/main.cpp:1079-1082,
 * updateWheelAnimations :1084-1109, updateScene :981-992, updateBVH :1068-1077 + node re-upload :342-344).
 * With the animations registered here, rt_animate does the same frame step on the device copies: K6 moves
 * the registered shapes (the per-entry sin / cos come from the host's libm, so the shapes are bit-identical
 * to the host path), K7 grows the reference node boxes (updateBVH semantics), then the library's own boxes
 * are refitted and the scene repacked as after an upload.  Nothing crosses PCIe but a few scalars.
 *   rt_anim_bounce   sphere `shape`: centre.y = origin_y + amplitude * sin(frequency * t)
 *   rt_anim_rotate_z triangles `shapes[0..count)`: vertices rotate about the z axis through `center` by
 *                    speed * dt per frame; planeNormal / planeD stay as they are (stale, as upstream)
 * The host copies kept by the context go stale; rt_download_scene fetches the current device shapes and
 * nodes (parity checks, or to continue on the host). */
int rt_anim_clear(rt_ctx* ctx);
int rt_anim_bounce(rt_ctx* ctx, int shape, float origin_y, float amplitude, float frequency);
int rt_anim_rotate_z(rt_ctx* ctx, const int32_t* shapes, int count, const float center[3], float speed);
int rt_animate(rt_ctx* ctx, float t, float dt);
/* K6 only: moves the registered shapes and leaves the node boxes alone — for a caller that rebuilds the tree afterwards
 * (rt_build_bvh) instead of growing it: per-frame rebuild loops need neither K7 nor its tables of the tree. */
int rt_animate_shapes(rt_ctx* ctx, float t, float dt);
int rt_download_scene(rt_ctx* ctx, void* shapes_or_NULL, void* nodes_or_NULL);

/* ---- device-side BVH build (K8; opt-in, the host-built tree stays the default) ----
 * Replaces buildBVH / split (src/main.cpp:1111-1193) + serializeBVH (:955-979) + the node / index upload
 * (:258-275, :342-344) for callers that only keep the shapes current: the reference's midpoint-split tree
 * over the shapes the context holds (the device copy, i.e. including what rt_animate moved) is built on the
 * GPU level by level and emitted in the reference's own numbering — FlatNode[] and bvhIndices[] are BYTE
 * IDENTICAL to the host builder's, so every traversal decision and tie-break stays what it was.  The tree
 * becomes the context's BVH as if it had been passed to rt_upload_scene (the threaded layout follows at the
 * next render; an unchanged topology keeps it and only refits).  maxDepth = buildBVH's argument (15 / 25 in
 * the reference scenes), 0 ... 64 (the shader's stack; anything else is RT_ERR_ARG).  By size the build runs in one thread block (<= 2048 shapes), in one
 * thread-block cluster of up to 16 blocks (<= 28000), in one cooperative launch of a block per SM (<= 120000) or as one launch per
 * phase and level; the tree is the same on every path.
 * Shapes of unknown type and infinite planes get no
 * bounds, as upstream.  rt_get_bvh copies the context's current nodes / indices out (either may be NULL);
 * rt_last_build_ms = device time of the last build. */
enum { RT_BUILD_DEFAULT = 0u,
       RT_BUILD_NO_WARP_AGGREGATION = 1u << 0,  /* one atomic per shape and bound instead of per warp / block (diagnostic) */
       RT_BUILD_NO_SINGLE_BLOCK = 1u << 1,      /* scenes of <= 2048 shapes do not take the single-block path (diagnostic) */
       RT_BUILD_NO_COOPERATIVE = 1u << 4,       /* scenes of <= 120000 shapes do not take the single cooperative launch (diagnostic) */
       RT_BUILD_NO_CLUSTER = 1u << 3,           /* scenes of <= 28000 shapes do not take the single-cluster path (diagnostic) */
       /* K9: lay the built tree out ON THE DEVICE (threaded list A, primitive slots, ranks straight from the builder's
        * tables): no host layout build and no copy of the tree to the host (rt_get_bvh fetches it on demand).  Applies
        * when the tree has no leaf of more than 16 shapes (such leaves are walked as they are; the host layout would expand
        * leaves of more than 8 shapes into sub-trees, which pays off for the hundreds of shapes the reference builder can
        * leave in a leaf) or the context walks the reference's own sequence (RT_FLAG_REFERENCE_WALK); otherwise the call
        * behaves as without the flag.  The own tree over the reference leaves (list B) is not built for a
        * device-laid-out tree. */
       RT_BUILD_DEVICE_LAYOUT = 1u << 2 };
int rt_build_bvh(rt_ctx* ctx, int maxDepth, unsigned build_flags, int* nNodes_out, int* nIndices_out);
int rt_get_bvh(rt_ctx* ctx, void* nodes_or_NULL, int nodeCapacity, int32_t* indices_or_NULL, int indexCapacity);
/* Layout of the scene the next render walks: out[0] nodes of list A, out[1] nodes of list B (0 = unused), out[2] nodes of
 * fat-leaf sub-trees, out[3] = 1 if the layout was produced on the device (K9), 0 if by the host layout builder. */
int rt_layout_info(rt_ctx* ctx, int out[4]);
float rt_last_build_ms(const rt_ctx* ctx);

/* ---- K8 without any allocation ----
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The single-block form of the K8 build
 * (the same tree: FlatNode[] / bvhIndices[] byte-identical to rt_build_bvh's and to buildBVH + serializeBVH,
 * src/main.cpp:955-979, :1111-1193) with its working tables in a file-scope __device__ array of the library instead of
 * buffers of a context: the call needs no rt_ctx and calls no cudaMalloc / cudaFree / allocating routine.  It only
 * launches one kernel on `cuda_stream` (NULL = the default stream) of the CURRENT device and returns without
 * synchronising; the first call on a device also resolves the array's device address (cudaGetSymbolAddress) — make
 * that call outside a stream capture; with nShapes = 0 and every pointer NULL the call does only this set-up.
 *   dev_shapes        nShapes FlatShape records (192 B each) in device memory
 *   dev_nodes_out     room for nodeCapacity >= max(1, 2 * nShapes - 1) FlatNode records (48 B each), device memory
 *   dev_indices_out   room for indexCapacity >= nShapes int32, device memory
 *   dev_counts_out    2 x int32 in device memory: [0] nodes written, [1] indices written
 *   err_or_NULL       receives the message of a failed call (errCapacity bytes, always terminated)
 * Limit: nShapes <= RT_STATIC_BUILD_MAX_SHAPES; more is rejected with RT_ERR_ARG and a message naming the limit (use
 * rt_build_bvh).  Every GPU that loads the library holds RT_STATIC_BUILD_SCRATCH_BYTES of device memory for the array
 * until the process exits.  All calls share the array: it keeps the last call's contents (the kernel re-initialises
 * what it reads), and two calls that overlap on the GPU — different streams, or two host threads — race on it:
 * order them on one stream or with events. */
#define RT_STATIC_BUILD_MAX_SHAPES 2048
#define RT_STATIC_BUILD_SCRATCH_BYTES 486144
int rt_build_bvh_static(const void* dev_shapes, int nShapes, int maxDepth,
                        void* dev_nodes_out, int nodeCapacity, int32_t* dev_indices_out, int indexCapacity,
                        int32_t* dev_counts_out, void* cuda_stream, char* err_or_NULL, size_t errCapacity);

/* K3: scatter `world` compact tile-major shards, laid out rank after rank in
 * dev_shards (each rt_shard_pixels long, as gathered by NCCL), into one
 * row-major W*H image. */
int rt_untile(rt_ctx* ctx, const float* dev_shards, int world, float* dev_rgba_out);

/* K4: the display quad's arithmetic (shader.frag:8-12 + default-framebuffer
 * conversion): alpha forced to 1, clamp to [0,1], quantise to RGBA8. */
int rt_present(rt_ctx* ctx, const float* dev_rgba, uint8_t* dev_rgba8_out, int64_t pixels);

/* ---- plumbing --------------------------------------------------------------*/
void* rt_stream(rt_ctx* ctx);                 /* cudaStream_t of the context     */
/* Run all further work of this context on a caller-owned cudaStream_t (e.g. the
 * stream a framework times with its own events); NULL restores the internal one.
 * The previous stream is drained first (pinned-shadow copies in flight complete with it).  A context is meant to be
 * driven by ONE host thread at a time: the calls are not internally synchronised. */
int   rt_set_stream(rt_ctx* ctx, void* cuda_stream);
int   rt_synchronize(rt_ctx* ctx);
float* rt_device_image(rt_ctx* ctx);          /* internal float4 image           */
/* CUDA-event timings of the last calls, milliseconds (-1 = not measured yet):
 * [0] H2D staging copy  [1] repack kernel(s)  [2] render kernel  [3] D2H */
int   rt_last_timings(rt_ctx* ctx, float out_ms[4]);
/* Number of kernels launched by this context since creation. */
int64_t rt_launch_count(const rt_ctx* ctx);
/* Build description: "sm_100a fmad=off ..." */
const char* rt_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* RT_API_H */

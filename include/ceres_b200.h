/* ceres_b200.h -- C ABI of the B200-native ceres-raytracer render loop.
 *
 * This is the drop-in boundary for ONE call of the reference:
 *
 *     std::pair<int,int> render(const Camera<Scalar>&, const Vector3<Scalar>& sun_position, const Bvh<Scalar>&,
 *                               const Triangle<Scalar>* triangles, std::array<Vector3<Scalar>,3>* tri_norms,
 *                               Scalar* pixels, size_t width, size_t height);      (include/render.hpp:86-89)
 *
 * as called from the frame loop of include/anim.cpp:105 and from include/static.cpp:130.  The reference has no
 * FFI layer; the functions below are what a binding of that call would bind (see INTEGRATION.md for the C++
 * shim with the identical template signature, include/ceres/render.hpp, and for a ctypes stub).
 *
 * Plain pointers and sizes only.  All input arrays use the reference's own POD layouts (Scalar = float):
 *   nodes      bvh::Bvh<float>::Node, 32 B  {float bounds[6] (minx,maxx,miny,maxy,minz,maxz);
 *                                            uint32 primitive_count; uint32 first_child_or_primitive}
 *                                                                                   (lib/bvh/bvh.hpp:25-30)
 *   prim_idx   size_t[tri_count]            Bvh::primitive_indices                  (lib/bvh/bvh.hpp:94)
 *   tris       bvh::Triangle<float>, 48 B   {p0, e1, e2, n}                          (lib/bvh/triangle.hpp:27)
 *   tri_norms  std::array<Vector3<float>,3>, 36 B                                    (lib/obj_norms.hpp:113-115)
 *
 * There is NO CPU fallback: every entry point that computes fails with CERES_ERR_NO_DEVICE / CERES_ERR_CUDA when
 * no sm_100 device is usable.  Scalar = double (anim -d, include/anim.cpp:154-155): see ceres_render_f64 below.
 * A ceres_scene is not thread-safe; use one per host thread / per GPU.
 */
#ifndef CERES_B200_H
#define CERES_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CERES_B200_ABI_VERSION 2

enum {
    CERES_OK = 0,
    CERES_ERR_INVALID = 1,     /* bad argument / malformed BVH */
    CERES_ERR_CUDA = 2,        /* a CUDA runtime call failed; see ceres_last_error() */
    CERES_ERR_NO_DEVICE = 3,   /* no usable GPU (the library never computes on the CPU) */
    CERES_ERR_UNSUPPORTED = 4
};

/* flags for the render calls */
#define CERES_NO_SHADOWS   1u  /* skip the sun shadow ray (render.hpp:135-137) and shade every primary hit
                                  ("primary rays only", BASELINE.json configs 1-2; not a mode of the reference) */
#define CERES_FRAME_RETAINED 4u /* host frames (ceres_render / ceres_render_shard), OPT-IN contract: "the pixels / ids buffers I pass
                                   are the ones I passed to the previous call with this flag on this scene, and nobody has written
                                   into them since".  The library then knows that everything outside the rectangles it wrote
                                   last time is still zero / CERES_MISS and only zeroes what the previous frame's rectangles
                                   leave uncovered by the new ones, instead of zero-filling ~70 % of a 4K frame from the host on
                                   every call (render.hpp:107,117 overwrite every pixel; the result is the same frame).  The
                                   first call with the flag, a different buffer / size / shard, or a call without the flag in
                                   between re-establishes the whole frame.  Ignored by the paths that do not use it. */
#define CERES_PAYLOAD_NO_FLIP 8u /* ceres_render_ppm8 / ppm16 (+ _shard): output row = frame row, as anim.cpp:117-119 copies pixels[j] into its
                                   image rows in order; default: rows flipped, row 0 = frame row height-1, as static.cpp:135-147 writes them */
#define CERES_COMPACT_ROWS 2u  /* sharded render: write the shard's rows packed, not at their frame position */

#define CERES_MISS 0xFFFFFFFFu /* value of ids[] where the primary ray hit nothing */

typedef struct ceres_scene ceres_scene;

/* Camera<float>, include/render.hpp:16-22 (fov in degrees) */
typedef struct ceres_camera {
    float eye[3];
    float dir[3];
    float up[3];
    float fov;
} ceres_camera;

/* Row-block sharding of one frame across GPUs: rows are grouped in blocks of `block_rows`,
 * block b belongs to shard b % count.  {0,0,1} or NULL = the whole frame. */
typedef struct ceres_shard {
    uint32_t block_rows;
    uint32_t index;
    uint32_t count;
} ceres_shard;

/* Kernel tuning knobs (0 = library default).  Exposed so bench.py/tests can sweep them; results never depend
 * on them. */
typedef struct ceres_tuning {
    uint32_t variant;            /* 0 = the frame kernel (the product).  1-6 = the wave pipelines of round 1, present only in
                                    the experiments build (libceres_b200_exp.so, -DCERES_EXPERIMENTS); see DESIGN.md */
    uint32_t primary_block;      /* experiments: threads per CTA, primary wave */
    uint32_t shadow_block;       /* experiments: threads per CTA, shadow wave */
    uint32_t ctas_per_sm;        /* persistent CTAs per SM */
    uint32_t shadow_order;       /* frame kernel, any-hit shadow rays: 0 default, 1 nearer child first (the reference's
                                    order), 2 left child first (no ordering work; the boolean cannot depend on it) */
    uint32_t reserved[3];        /* [0] frame kernel: shadow-subtree sharing between the lanes of a warp, 0 default / 1 off /
                                    2 on (experiments, variant != 0: lane-refill threshold); [1] row bands per frame;
                                    [2] frame kernel: sharing period | min idle lanes << 8 (experiments: chunk) */
} ceres_tuning;

/* ---- scene ------------------------------------------------------------------------------------------------- */

/* Flattens the reference BVH + triangles + normals and uploads them ONCE to `device` (replaces the per-call
 * pointer arguments bvh/triangles/tri_norms of render.hpp:86-89; the reference also builds them once, outside
 * the frame loop, anim.cpp:46-65).  tri_norms may be NULL (zeros; shading then yields NaN exactly as the
 * reference would with zero normals). */
int ceres_scene_create(const void* nodes, size_t node_count, const size_t* prim_idx, const void* tris,
                       const void* tri_norms, size_t tri_count, int device, ceres_scene** out);
void ceres_scene_destroy(ceres_scene* scene);
int ceres_scene_set_tuning(ceres_scene* scene, const ceres_tuning* tuning);
/* bytes of device memory held by the flattened scene */
size_t ceres_scene_device_bytes(const ceres_scene* scene);
/* entries of traversal stack the scene's tree can need (its depth; <= 64, single_ray_traverser.hpp:14) */
uint32_t ceres_scene_stack_depth(const ceres_scene* scene);
/* nodes of the scene's tree in the reference's numbering (what ceres_scene_download expects) */
size_t ceres_scene_node_count(const ceres_scene* scene);
/* bytes the last host-frame render (ceres_render / _shard / one frame of _sequence) moved over PCIe: the screen
 * rectangle of the scene's rows + the counters; everything else of the frame is zero-filled by host threads */
uint64_t ceres_scene_last_d2h_bytes(const ceres_scene* scene);
/* conservative screen rectangle of the scene in the last frame prepared on this scene: columns [rect[0], rect[1]) x LOCAL
 * rows [rect[2], rect[3]) of the shard; only its pixels are traced (everything outside cannot hit anything) */
int ceres_scene_last_rectangle(const ceres_scene* scene, uint32_t rect[4]);

/* ---- scene preparation on the device (the step before the path) ------------------------------
 * For animated objects: move the triangles that already live on the GPU and refit the BVH there, instead of a host
 * rebuild + re-upload.  Synchronous; must not overlap frames of this scene that are still in flight.
 *   ceres_scene_rotate     rotate_triangles<Axis> (render.hpp:24-44): axis 0/1/2, degrees; cos/sin on the host
 *   ceres_scene_transform  transform_triangles (transform.hpp:106-129): a = Transform::a row-major 3x3, v = Transform::v
 *   ceres_scene_refit      HierarchyRefitter::refit (lib/bvh/hierarchy_refitter.hpp:17-45), leaf box = union of its
 *                          triangles' bounding_box() (triangle.hpp:39-44); also refreshes the culling rectangle's box
 *   ceres_scene_download   the scene as the reference would hold it now: nodes_out = node_count x 32 B reference
 *                          nodes (same numbering as at creation), tris_out = tri_count x 48 B in primitive order
 *                          (either may be NULL).
 * tri_norms are NOT transformed -- neither does the reference (render.hpp:24-44 touches only `triangles`). */
int ceres_scene_rotate(ceres_scene* scene, int axis, float degrees);
int ceres_scene_transform(ceres_scene* scene, const float a[9], const float v[3]);
int ceres_scene_refit(ceres_scene* scene);
int ceres_scene_download(ceres_scene* scene, void* nodes_out, size_t node_count, void* tris_out, size_t tri_count);

/* ---- BVH construction on the GPU ---------------------------------------------------------------
 * Morton codes of the triangle centres, a radix sort, then (default) a level-synchronous top-down binned SAH build over that
 * order (what the reference's BinnedSahBuilder does on the host, static.cpp:106-107) with SAH leaf collapsing -- or
 * locally-ordered clustering, or Karras' radix tree alone (ceres_build_options; the GPU-friendly builders the reference sketches
 * in lib/bvh/locally_ordered_clustering_builder.hpp:186-234 and linear_bvh_builder.hpp:118-185) -- in the REFERENCE's data
 * structure: nodes_out = up to (2 * tri_count - 1) x 32-byte bvh::Bvh<float>::Node with node 0 the root and siblings adjacent
 * (bvh.hpp:25-30), inner nodes in pre-order, prim_idx_out = tri_count x size_t Bvh::primitive_indices in depth-first leaf order,
 * up to max_leaf triangles per leaf.  tris: HOST tri_count x 48-byte
 * Triangle<float>.  Replaces the host build of anim.cpp:46-65 / static.cpp:98-107 where build time matters (10 M
 * triangles); the tree differs from SweepSah's, so exact-t ties may pick another triangle than the SweepSah tree would --
 * frames are bit-exact against the reference's traverser walking THIS tree.  build_ms (nullable): device time of the
 * build kernels without the transfers. */
int ceres_build_lbvh(const void* tris, size_t tri_count, int device, void* nodes_out, size_t* prim_idx_out, double* build_ms);
/* The same build, reporting how many nodes the tree has: with leaf collapsing (below) fewer than 2 tri_count - 1; nodes_out must
 * still hold 2 tri_count - 1 nodes (the tail is zero).  ceres_build_lbvh = this call without the count. */
int ceres_build_bvh(const void* tris, size_t tri_count, int device, void* nodes_out, size_t* node_count_out, size_t* prim_idx_out,
                    double* build_ms);
/* What the build calls make over the Morton-sorted triangles (process-wide; 0 = default for every argument):
 *   algorithm     0 (default) = 2 = top-down binned SAH, one level per pass: every open node bins its members' centroids (16 bins per
 *                 axis over its centroid bounds; 4 for nodes of at most 16 triangles), splits where area x count is smallest over
 *                 the three axes, members never move in memory (binned_sah_builder.hpp on the host; renders 1-5 % FASTER than the
 *                 reference's SweepSah tree, builds 10.5 M triangles in 35 ms);
 *                 3 = parallel locally-ordered clustering (the idea of lib/bvh/locally_ordered_clustering_builder.hpp:186-234:
 *                 every round each cluster merges with the neighbour within `radius` places whose union has the smallest area, if
 *                 the choice is mutual; 3 x faster to build, 6-9 % slower to render than SweepSah's tree);
 *                 1 = the radix tree of the Morton codes alone (linear_bvh_builder.hpp:118-185; fastest to build, ~19 % slower to
 *                 render)
 *   radius        clustering: neighbours searched on either side (default 8)
 *   max_leaf      SAH leaf collapsing: a subtree of at most max_leaf triangles becomes one leaf where area x count x
 *                 triangle_cost <= area + cost(children) (the rule by which SweepSah stops splitting, top_down_builder.hpp:41);
 *                 1 = off: one triangle per leaf, 2 tri_count - 1 nodes (default 8)
 *   triangle_cost cost of one triangle test in node steps (default 1.5: a Moeller-Trumbore test is ~90 instructions of the frame
 *                 kernel, a node-pair step ~60) */
int ceres_build_options(int algorithm, int radius, int max_leaf, float triangle_cost);

/* The same build, kept on the device: a scene whose tree never visits the host (the host-side flatten of
 * ceres_scene_create costs more than the build for 10 M triangles).  tris / norms as for ceres_scene_create
 * (norms nullable); the resulting scene is identical to ceres_scene_create(ceres_build_lbvh(tris)) for rendering,
 * rotate / transform / refit and download (reference node numbering: children of radix-tree node i at 1 + 2 i, 2 + 2 i). */
int ceres_scene_create_lbvh(const void* tris, const void* norms, size_t tri_count, int device, ceres_scene** out, double* build_ms);

/* ---- Scalar = double (`anim -d`, include/anim.cpp:154-155) -------------------------------------
 * render<double>() (render.hpp:86-156) over the double-precision scene the reference holds: nodes = node_count x
 * 64-byte bvh::Bvh<double>::Node {double bounds[6]; uint64 primitive_count; uint64 first_child_or_primitive}
 * (bvh.hpp:20-28), tris = tri_count x 96-byte Triangle<double> (triangle.hpp:27), norms = tri_count x 72 bytes
 * (obj_norms.hpp:113-115; nullable), prim_idx as for ceres_scene_create.  pixels: HOST 3 * width * height doubles,
 * layout 3 * (width * j + i) (render.hpp:107); ids (nullable): HOST width * height primitive indices, CERES_MISS where
 * nothing was hit; counters = the pair of render.hpp:155; flags: CERES_NO_SHADOWS.  kernel_ms (nullable): device time
 * of the frame's kernels.  ceres_render_f64_shard: the float path's row-block sharding (ceres_shard; one scene per GPU,
 * `pixels` / `ids` = the FULL-SIZE frame, of which the call writes only the shard's rows, zeros included; counters = the
 * shard's: rays = own rows * width + own primary hits) -- the shards of a frame add up to ceres_render_f64's frame and
 * counters exactly (pixels are independent, render.hpp:104). */
typedef struct ceres_scene_f64 ceres_scene_f64;
typedef struct ceres_camera_f64 {   /* Camera<double>, include/render.hpp:16-22 (fov in degrees) */
    double eye[3];
    double dir[3];
    double up[3];
    double fov;
} ceres_camera_f64;
int ceres_scene_f64_create(const void* nodes, size_t node_count, const size_t* prim_idx, const void* tris, const void* norms,
                           size_t tri_count, int device, ceres_scene_f64** out);
void ceres_scene_f64_destroy(ceres_scene_f64* scene);
int ceres_render_f64(ceres_scene_f64* scene, const ceres_camera_f64* cam, const double sun[3], size_t width, size_t height,
                     unsigned flags, double* pixels, uint32_t* ids, uint64_t counters[2], double* kernel_ms);
int ceres_render_f64_shard(ceres_scene_f64* scene, const ceres_camera_f64* cam, const double sun[3], size_t width, size_t height,
                           unsigned flags, const ceres_shard* shard, double* pixels, uint32_t* ids, uint64_t counters[2],
                           double* kernel_ms);

/* ---- render: host buffers (the drop-in for render.hpp:86-156) ------------------------------------------------ */

/* pixels: caller-owned HOST buffer of 3*width*height floats, layout pixels[3*(width*j+i)+c], row j=0 is v=-1,
 * every pixel written (render.hpp:107,117,144-149).  ids (nullable): HOST uint32[width*height], primitive index of
 * the primary hit (hit->primitive_index, render.hpp:120) or CERES_MISS.  counters (nullable): [0] = rays
 * (= width*height + primary hits), [1] = hits (= primary hits + occluded shadow rays), the pair render()
 * returns (render.hpp:155).  Blocks until the frame is in `pixels`. */
int ceres_render(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3],
                 size_t width, size_t height, unsigned flags, float* pixels, uint32_t* ids, uint64_t counters[2]);

/* The same frame as the payload of the reference's P6 file (static.cpp:135-147): quantised on the device
 * (uint8(max(min(p*255,255),0)), rows flipped: row 0 = frame row height-1) and only the 8-bit screen rectangle of the
 * scene copied -- a quarter of the float path's PCIe bytes.  rgb8: HOST 3*width*height bytes. */
int ceres_render_ppm8(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3],
                      size_t width, size_t height, unsigned flags, uint8_t* rgb8, uint64_t counters[2]);

/* The 16-bit counterpart for `anim` frames: channel = uint16(max(min(p*65535,65535),0)) -- the x 65535 quantum scaling
 * of anim.cpp:117-119 -- big-endian as a P6 file with maxval 65535 stores it, rows flipped like ceres_render_ppm8.
 * rgb16: HOST 6*width*height bytes; half of the float path's PCIe and host-memory bytes. */
int ceres_render_ppm16(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3],
                       size_t width, size_t height, unsigned flags, uint8_t* rgb16, uint64_t counters[2]);

/* One shard of either payload: only the shard's rows of the FULL-SIZE payload are written (output row height-1-j for frame
 * row j), so the ranks of a box fill one shared payload, each over its own PCIe link, with no gather (counters: the shard's).
 * This is the path on which more GPUs shorten a frame end to end: the 8-bit 4K payload is 24.9 MB of host memory instead of
 * the float frame's 99.5 MB, which no number of GPUs can write faster than the host's DRAM takes it. */
int ceres_render_ppm8_shard(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3], size_t width, size_t height,
                            unsigned flags, const ceres_shard* shard, uint8_t* rgb8, uint64_t counters[2]);
int ceres_render_ppm16_shard(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3], size_t width, size_t height,
                             unsigned flags, const ceres_shard* shard, uint8_t* rgb16, uint64_t counters[2]);

/* Same, one shard of the frame.  Without CERES_COMPACT_ROWS only the shard's rows of pixels/ids are written. */
int ceres_render_shard(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3],
                       size_t width, size_t height, unsigned flags, const ceres_shard* shard,
                       float* pixels, uint32_t* ids, uint64_t counters[2]);

/* One frame on n GPUs from one process: scenes[k] (created on device k from the same arrays) renders the row blocks
 * b = k (mod n) of CERES_TILE_BLOCK_ROWS rows and DMA-writes them in place into the host frame; counters are summed.
 * No collective: every GPU drains its rows over its own PCIe link. */
#define CERES_TILE_BLOCK_ROWS 16u
int ceres_render_multi(ceres_scene* const* scenes, int n, const ceres_camera* camera, const float sun_position[3],
                       size_t width, size_t height, unsigned flags, float* pixels, uint32_t* ids, uint64_t counters[2]);

/* N frames with one upload of the camera path: cams = n_frames x 13 floats (eye[3],dir[3],up[3],fov,sun[3]),
 * the per-frame values of anim.cpp:87-89; pixels_all = n_frames frames back to back; counters_all (nullable)
 * 2 per frame.  Each frame takes the host-frame path of ceres_render (only the scene's screen rectangle crosses PCIe); two
 * frames are in flight at a time, the odd ones rendered from a second thread of the library over the same resident geometry.
 * While the call runs, ceres_render*, ceres_scene_rotate / _transform / _refit on the same scene from another thread return
 * CERES_ERR_INVALID instead of racing with it. */
int ceres_render_sequence(ceres_scene* scene, size_t n_frames, const float* cams, size_t width, size_t height,
                          unsigned flags, float* pixels_all, uint64_t* counters_all);
/* The same sequence as quantised P6 payloads (bytes_per_channel 1: ceres_render_ppm8, 2: ceres_render_ppm16 per frame; flags may
 * carry CERES_PAYLOAD_NO_FLIP, the row order of anim.cpp:117-119): payload_all = n_frames x 3*width*height*bytes_per_channel. */
int ceres_render_sequence_payload(ceres_scene* scene, size_t n_frames, const float* cams, size_t width, size_t height,
                                  unsigned flags, int bytes_per_channel, uint8_t* payload_all, uint64_t* counters_all);

/* ---- render: device buffers (asynchronous; used by bench.py, multi-GPU gather and the host calls above) ------- */

/* Enqueues the frame on `stream` (a cudaStream_t; NULL = the scene's own stream, cudaStreamLegacy (0x1) = the
 * legacy default stream) and returns.  d_pixels / d_ids /
 * d_counters are DEVICE pointers on the scene's device (d_ids, d_counters nullable; d_counters = 2 x uint64).
 * The frames of one scene share its record queue and counters: a frame enqueued on another stream than the scene's previous
 * frame is made to wait for that frame's end (an event recorded after every frame), so frames of one scene never overlap --
 * use one scene per stream (ceres_render_sequence does, over the same resident geometry) to overlap frames. */
int ceres_render_device(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3],
                        size_t width, size_t height, unsigned flags, const ceres_shard* shard, void* stream,
                        float* d_pixels, uint32_t* d_ids, uint64_t* d_counters);

/* ceres_render_device for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same frame, same kernel, same outputs as ceres_render_device(shard = NULL); the difference is where the frame kernel's
 * scratch lives (record queue, its ready[] credits, the counter slot): in a file-scope `static __device__` array of
 * CERES_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit, instead of the scene's grow-on-demand
 * cudaMalloc'ed buffers.  The call performs no cudaMalloc / cudaFree / allocating library routine -- only
 * cudaGetSymbolAddress on the first call per device, then async memset + 2 kernel launches on `stream`: after one warm-up
 * call it can be captured into a CUDA graph and replayed (tests/test_zz_no_alloc.py does).
 * Limits: width * height <= CERES_STATIC_MAX_PIXELS (larger frames are refused with CERES_ERR_INVALID and a message that
 * names the limit); whole frames only (no shard); flags: CERES_NO_SHADOWS only; statistics / timeline modes and tuning
 * variants are refused (CERES_ERR_UNSUPPORTED).  d_ids / d_counters nullable as above.
 * ALL calls on a device share the one array, whatever the scene or stream: the call re-initialises what must start at zero,
 * but two calls that overlap on the GPU (different streams, or different scenes' default streams) RACE -- keep them in
 * one stream or order them with events.  It does not touch the scene's own queue, so it does not wait for, or hold up,
 * the scene's ceres_render_device frames. */
#define CERES_STATIC_MAX_PIXELS (1u << 20)          /* e.g. 1024 x 1024, 1280 x 720 */
#define CERES_STATIC_SCRATCH_BYTES 17367552u        /* 16 MiB of records + 576 KiB of credits + 256 B of counters */
int ceres_render_device_static(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3],
                               size_t width, size_t height, unsigned flags, void* stream,
                               float* d_pixels, uint32_t* d_ids, uint64_t* d_counters);

/* ---- render: one frame on N GPUs, fused with its "gather" over NVLink peer memory ------------
 * One process per GPU.  Rank 0 owns the frame: a device allocation holding the float RGB frame (+ optional ids) and
 * a small control block; the other ranks map it (CUDA IPC) and their kernels store the pixels of their row blocks
 * STRAIGHT INTO rank 0's memory while they shade -- there is no separate gather pass and no collective call.
 * Per frame ("epoch" e, counted by every rank in lockstep) every rank -- rank 0 included -- launches ONE kernel: its frame
 * kernel writes EVERY pixel of its row blocks that lies inside the scene's conservative screen rectangle (zeros for misses and
 * shadowed hits, colours for lit hits) over NVLink into rank 0's memory, fences at system scope and publishes
 * done_epoch[r] = e with its (rays, hits).  The frame is library-owned, so outside the rectangle it already holds zeros:
 * a rank clears only what the PREVIOUS epoch's rectangle left outside the current one (a moving camera; nothing for a
 * repeated view) -- there is no whole-frame clear and no rank waits for another rank's clear.  Rank 0's kernel ends by waiting
 * until done_epoch[r] >= e for every rank and sums the counters into d_counters.  The only cross-rank gate at the start: a
 * rank's first store of epoch e waits until rank 0's kernels of epoch e have begun, so a rank that runs ahead cannot write
 * into a frame rank 0's process is still reading (stream order on rank 0).  When several ranks share ONE GPU (tests), enqueue
 * rank 0 first.  Waits are bounded (about 4 s): on expiry the control block's error word is set and ceres_peer_frame_error
 * reports it; the GPU is never left spinning.  All ranks must call ceres_render_peer the same number of times.
 * Whoever writes into the frame through ceres_peer_frame_pixels/_ids must call ceres_peer_frame_invalidate on every rank
 * before the next ceres_render_peer (the next frame then clears all of the rank's rows once). */
typedef struct ceres_peer_frame ceres_peer_frame;
#define CERES_PEER_HANDLE_BYTES 64
#define CERES_PEER_MAX_RANKS 16
/* rank 0: allocate the frame on the scene's device; handle_out (CERES_PEER_HANDLE_BYTES) is what the other ranks
 * pass to ceres_peer_frame_open (ship it with any host-side channel, e.g. a torch.distributed broadcast). */
int ceres_peer_frame_create(ceres_scene* scene, size_t width, size_t height, int want_ids, int world,
                            ceres_peer_frame** out, unsigned char* handle_out);
/* rank r > 0 (another process, any GPU of the box with peer access to rank 0's): map rank 0's frame. */
int ceres_peer_frame_open(ceres_scene* scene, const unsigned char* handle, size_t width, size_t height, int want_ids,
                          int world, int rank, ceres_peer_frame** out);
/* rank r > 0 living in the SAME process as the owner (tests; a multi-GPU single-process host): shares the mapping. */
int ceres_peer_frame_alias(ceres_peer_frame* owner, ceres_scene* scene, int rank, ceres_peer_frame** out);
void ceres_peer_frame_destroy(ceres_peer_frame* frame);
/* device pointers of the assembled frame (valid in the calling process; meaningful to read on rank 0) */
float* ceres_peer_frame_pixels(ceres_peer_frame* frame);
uint32_t* ceres_peer_frame_ids(ceres_peer_frame* frame);
/* Enqueue this rank's part of the next frame on `stream`.  d_counters (rank 0 only, nullable): DEVICE 2 x uint64,
 * receives the frame totals once every rank has delivered. */
int ceres_render_peer(ceres_scene* scene, const ceres_camera* camera, const float sun_position[3], unsigned flags,
                      void* stream, ceres_peer_frame* frame, uint64_t* d_counters);
/* after the stream has been synchronised: 0 = no wait of any rank expired so far */
int ceres_peer_frame_error(ceres_peer_frame* frame);
/* the frame's memory was written by somebody else: this rank's next frame re-establishes zeros outside its rectangle */
int ceres_peer_frame_invalidate(ceres_peer_frame* frame);

/* Per-frame work statistics of the LAST ceres_render_device call on this scene, read back after the stream has
 * been synchronised: stats[0..7] = {S_miss,T_miss,S_hit,T_hit,S_shadow,T_shadow,rays,hits} where S = node-pair
 * steps and T = triangle tests actually executed by the GPU (shadow rays exit at the first hit, so S_shadow and
 * T_shadow are smaller than the reference's closest-hit figures).  Only filled when the scene was put in
 * statistics mode with ceres_scene_set_stats(scene, 1) (slower kernels). */
int ceres_scene_set_stats(ceres_scene* scene, int enable);
int ceres_scene_get_stats(ceres_scene* scene, uint64_t stats[8]);
/* Statistics mode also fills a histogram of node-pair steps per ray, accumulated over the frames rendered since the
 * last call (the call resets it): hist[b] = primary rays with b = min(steps / 4, 63), hist[64 + b] = shadow rays. */
int ceres_scene_get_step_histogram(ceres_scene* scene, uint64_t hist[128]);
/* Device-side errors of the frames enqueued so far on this scene (synchronises the device): *bits = 0 when none.
 * Bit 0: a bounded wait on the record queue expired.  Checked build only (-DCERES_CHECKED, ceres_build_is_checked()):
 * 0x0100 node-pair index, 0x0200 triangle slot, 0x0400 stack overflow, 0x0800 record queue, 0x1000 ready[], 0x2000 pixel
 * store, 0x4000 stack underflow -- an index was outside its bound (it was replaced by 0 and the run continued);
 * 0x10000 / 0x20000: the per-frame queue state was not re-armed between frames.  Reading clears the word. */
int ceres_scene_frame_error(ceres_scene* scene, uint64_t* bits);
int ceres_build_is_checked(void);

/* Per-warp timeline of the frame kernel (development aid for the per-rank timelines of DESIGN.md).  While enabled, every
 * warp w of the LAST frame kernel launched on this scene leaves CERES_TIMELINE_WORDS words at out[CERES_TIMELINE_WORDS w]:
 * {t_begin, t_primary_end, t_end (globaltimer ns), primary tiles, shadow chunks, longest tile [ns], its first work item,
 * claim time of the last tile, longest chunk [ns], its index, total wait for queue chunks [ns], time the peer gate was passed}; *warps = number of
 * entries written (<= cap_warps; warps that never ran are all-zero).  Costs a few uniform branches per tile / chunk. */
#define CERES_TIMELINE_WORDS 12
int ceres_scene_set_timeline(ceres_scene* scene, int enable);
int ceres_scene_get_timeline(ceres_scene* scene, uint64_t* out, size_t cap_warps, size_t* warps);

/* Per-kernel timing.  While enabled, every ceres_render_device call records CUDA events on its launch stream
 * before/after each of its kernels.  ceres_scene_get_kernel_ms waits for the recorded frames and returns the SUMS
 * over the frames recorded since the last ceres_scene_set_profiling call: ms[0] clear_outside_kernel, ms[1] frame_kernel,
 * ms[2] 0 (experiments build, variants 1-6: primary wave incl. clear, shadow wave, shade pass), ms[3] first-to-last
 * event; *frames (nullable) = number of frames summed. */
int ceres_scene_set_profiling(ceres_scene* scene, int enable);
int ceres_scene_get_kernel_ms(ceres_scene* scene, double ms[4], uint64_t* frames);

/* ---- known-answer entry points (tests) ----------------------------------------------------------------------- */

/* Closest-hit traversal of arbitrary rays through the scene (single_ray_traverser.hpp:142-154 with
 * ClosestPrimitiveIntersector).  rays: HOST n x 8 floats (origin[3], direction[3], tmin, tmax).  out_id: HOST
 * uint32[n]; out_tuv: HOST n x 3 floats (t,u,v; zeros on miss). */
int ceres_trace_rays(ceres_scene* scene, const float* rays, size_t n, uint32_t* out_id, float* out_tuv);

/* Frame setup of render.hpp:91-97 on the host: out = dir[3], image_u[3], image_v[3]. */
int ceres_frame_setup(const ceres_camera* camera, size_t width, size_t height, float out[9]);

/* ---- framebuffer post-processing on the device (the step after the path, static.cpp:135-147) ------------------ */

/* float RGB -> 8-bit, vertically flipped (row height-1 first), uint8(max(min(p*255,255),0)): the payload of the
 * reference's P6 PPM.  d_rgb8: DEVICE 3*width*height bytes. */
int ceres_quantize_device(ceres_scene* scene, const float* d_pixels, size_t width, size_t height, void* stream,
                          uint8_t* d_rgb8);

/* ---- pinned host memory helpers ------------------------------------------------------------------------------ */
/* A framebuffer that is page-locked (allocated here, registered here, or by any cudaHostAlloc) is filled by DMA
 * directly; any other host pointer is accepted and goes through the driver's pageable-copy path (slower).  The
 * library never registers caller memory behind the caller's back. */
void* ceres_host_alloc(size_t bytes);
void ceres_host_free(void* p);
int ceres_host_register(void* p, size_t bytes);   /* cudaHostRegister; caller unregisters before freeing */
int ceres_host_unregister(void* p);
int ceres_host_is_pinned(const void* p);

/* Read bandwidth of this GPU for a working set of the given size, measured with the library's own 128-bit __ldg
 * read kernel (best of 5): a few MB per SM-resident set -> L1, 16-64 MB -> L2, >= 1 GB -> HBM.  bench.py reports
 * the cache-resident scene's traffic against the L2 figure next to the HBM roofline. */
int ceres_probe_read_bandwidth(int device, size_t working_set_bytes, int repeats, double* gb_per_s);

/* Streaming-store bandwidth (GB/s, best of `repeats`) of the library's host-fill pool into `bytes` of page-locked memory.
 * Every byte of a host frame has to be written into host DRAM by somebody -- DMA for the scene's rectangle, this pool
 * for the zeros around it -- so host_frame_bytes / this figure is the floor of a host-frame call (bench.py e2e.floor_ms). */
int ceres_probe_host_fill_bandwidth(size_t bytes, int repeats, double* gb_per_s);

/* ---- misc --------------------------------------------------------------------------------------------------- */
int ceres_abi_version(void);
int ceres_device_count(void);
/* launches of this library's kernels since load (all scenes); bench.py reports the delta as gpu_launches */
uint64_t ceres_kernel_launch_count(void);
/* thread-local description of the last failure */
const char* ceres_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* CERES_B200_H */

/* minipath_b200 — C ABI of the B200-native renderer core for minipath's per-pixel trace path.
 *
 * The reference (bluecube/minipath, pure Rust) has no FFI of its own; the boundary this library sits
 * behind is the crate's public Rust API (`render`, `RenderSettings`, `RenderProgress`, `Camera`,
 * `Scene<TriangleBvh>`; reference src/lib.rs:8-10).  Every entry point below names the reference
 * item (file:line under the reference tree) it replaces.  INTEGRATION.md shows the Rust `-sys` binding.
 *
 * Conventions: plain C types only; every function returns 0 on success and a negative MP_E* code on
 * failure (message via mp_last_error(), thread-local); nothing throws across the ABI; one MpCtx per GPU,
 * calls on one context are serialised by the caller (mirrors one Worker per thread,
 * reference src/renderer/worker.rs:15-20).  There is NO CPU fallback: without a CUDA device every
 * device entry point fails with MP_ENODEVICE.
 */
#ifndef MINIPATH_B200_H
#define MINIPATH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MP_OK 0
#define MP_EINVAL (-1)
#define MP_ECUDA (-2)
#define MP_ENODEVICE (-3)
#define MP_EIO (-4)
#define MP_ESTATE (-5)
#define MP_ENOMEM (-6)

/* ---- node links: reference src/scene/triangle_bvh/mod.rs:57-114 (CompressedNodeLink) ---- */
#define MP_LINK_COUNT_BITS 3u
#define MP_LINK_COUNT_MASK 7u
#define MP_LINK_NULL 0xFFFFFFF8u        /* (u32::MAX >> 3) << 3, mod.rs:69 */
#define MP_LINK_MAX_INDEX 0x1FFFFFFEu   /* (u32::MAX >> 3) - 1, mod.rs:71 */
#define MP_LINK_MAX_COUNT 7u            /* mod.rs:73 */
#define MP_NO_HIT 0xFFFFFFFFu           /* TriangleIdx::default() == usize::MAX, mod.rs:143-147 */
#define MP_INNER_NODE_CHILDREN 8        /* mod.rs:14 */
#define MP_LEAF_PACKET_SIZE 8           /* mod.rs:15 */
#define MP_LEAF_MAX_TRIANGLES 56        /* mod.rs:16-17 */

/* Encode/decode helpers; return MP_EINVAL where the reference panics (mod.rs:78-87,189-236). */
int mp_link_new_leaf(uint32_t packet_index, uint32_t packet_count, uint32_t* out_link);
int mp_link_new_inner(uint32_t node_index, uint32_t* out_link);
/* kind: 0 = null, 1 = inner, 2 = leaf */
int mp_link_decode(uint32_t link, int* out_kind, uint32_t* out_index, uint32_t* out_count);

/* ---- flat BVH: device image of reference TriangleBvh (triangle_bvh/mod.rs:20-53) ---- */

/* Enclosing box of a node as carried on the reference's traversal stack (WorldBoxSized,
 * ray_bvh_intersection.rs:21-23; geometry/aabb.rs:286-303): min and size = max - min. */
typedef struct MpBox {
    float min[3];
    float _pad0;
    float size[3];
    float _pad1;
} MpBox; /* 32 B */

/* InnerNode (mod.rs:33-38): child_bounds = RelativeBox8 {min:{x,y,z}, max:{x,y,z}} of u16x8
 * (compressed_geometry.rs:59-63,117), then 8 links.  128 B = 8 x 16-byte loads. */
typedef struct MpInnerNode {
    uint16_t bounds[6][8]; /* [0..2] = min.x,min.y,min.z ; [3..5] = max.x,max.y,max.z ; [..][lane] */
    uint32_t links[8];
} MpInnerNode;

/* RelativeTriangle8 (compressed_geometry.rs:147): 3 vertices x {x,y,z} x u16x8. 144 B = 9 x 16 B. */
typedef struct MpTrianglePacket {
    uint16_t v[3][3][8]; /* [vertex][axis][lane] */
} MpTrianglePacket;

/* TriangleShadingData (mod.rs:40-45) narrowed to u32: bit 31 of flags_material = flat_shading,
 * low 31 bits = material.  Padding triangles hold {0,0,0,0} like the reference's Default. */
typedef struct MpTriShading {
    uint32_t v[3];
    uint32_t flags_material;
} MpTriShading;
#define MP_TRI_FLAT 0x80000000u

/* VertexShadingData (mod.rs:49-53). */
typedef struct MpVertexShading {
    float normal[3];
    float _pad0;
    float tex[3];
    float _pad1;
} MpVertexShading;

/* Material table entry (extension).  The reference carries `material: usize` through TriangleShadingData and
 * HitRecord (mod.rs:40-45; geometry/mod.rs:70-80) but always writes 0 (building.rs:201) and never reads it; here the id is
 * the low 31 bits of MpTriShading.flags_material and indexes this table.  albedo = Kd of the .mtl entry; MP_MAT_CHECKER
 * halves it on the odd cells of a grid of `checker_scale` cells per unit of the interpolated texture coordinate
 * (triangle.rs:229-244) — a procedural stand-in for map_Kd, whose image files the reference's data/ does not ship. */
typedef struct MpMaterial {
    float albedo[3];
    uint32_t flags; /* MP_MAT_* */
    float checker_scale;
    float _pad[3];
} MpMaterial; /* 32 B */
#define MP_MAT_CHECKER 1u
#define MP_TRI_MATERIAL_MASK 0x7fffffffu
#define MP_DEFAULT_ALBEDO 0.7f        /* the grey of scenes without a table, and of .mtl entries without Kd */
#define MP_DEFAULT_CHECKER_SCALE 4.0f /* cells per texture-coordinate unit given to .mtl entries that have a map_Kd */

typedef struct MpFlatBvh {
    float bbox_min[3];
    float bbox_max[3];
    uint32_t root;        /* CompressedNodeLink */
    uint32_t n_inner;
    uint32_t n_packets;
    uint32_t n_vertices;
    uint32_t n_triangles; /* real (unpadded) triangles */
    uint32_t n_materials; /* 0 = no table: the extension integrator uses its constant grey albedo for every triangle */
    const MpInnerNode* inner;          /* [n_inner] */
    const MpBox* inner_box;            /* [n_inner]   enclosing box each inner node is decoded against */
    const MpTrianglePacket* packets;   /* [n_packets] */
    const MpBox* packet_box;           /* [n_packets] enclosing box of the leaf the packet belongs to */
    const MpTriShading* tri_shading;   /* [n_packets*8] indexed by packet*8+lane */
    const MpVertexShading* vertices;   /* [n_vertices] */
    const MpMaterial* materials;       /* [n_materials] or NULL */
} MpFlatBvh;

/* Host-side scene loading — replaces TriangleBvh::with_obj / build (building.rs:28-120).
 * quad_mode: 0 = skip non-triangles like the reference (building.rs:43-46);
 *            1 = fan-triangulate polygons (documented deviation, needed for data/cube.obj). */
int mp_bvh_build_obj(const char* path, int quad_mode, MpFlatBvh** out);
/* Build from an indexed mesh in memory (positions xyz, optional normals xyz (may be NULL → flat),
 * triangles as 3 vertex indices).  Same builder as above. */
int mp_bvh_build_mesh(const float* positions, const float* normals, uint32_t n_vertices,
                      const uint32_t* triangles, uint64_t n_triangles, MpFlatBvh** out);
/* builder: 0 = reference split (building.rs:238-345, O(G^3) agglomerative merge);
 *          1 = median-split builder emitting the same node format (for large soups; not reference-pinned) */
int mp_bvh_build_mesh_ex(const float* positions, const float* normals, uint32_t n_vertices,
                         const uint32_t* triangles, uint64_t n_triangles, int builder, MpFlatBvh** out);
/* Same with per-vertex texture coordinates (uv pairs, may be NULL → 0), a material id per triangle (may be NULL → 0) and the
 * material table (n_materials = 0 → none; every id must be < max(n_materials, 1)).  `mp_bvh_build_obj` fills the same fields
 * from `mtllib` / `usemtl` / `Kd`: entry 0 is the default material (faces before any `usemtl`), then the .mtl entries in file
 * order; an OBJ without `mtllib` gets no table. */
int mp_bvh_build_mesh_mat(const float* positions, const float* normals, const float* texcoords, uint32_t n_vertices,
                          const uint32_t* triangles, const uint32_t* tri_material, uint64_t n_triangles,
                          const MpMaterial* materials, uint32_t n_materials, int builder, MpFlatBvh** out);
/* The same arguments, built ON THE GPU `device` (csrc/build_device.cu): Morton sort of the centroids, up to 8 children
 * per node by radix splits, exact boxes bottom-up, then the reference's top-down round-out quantisation against the decompressed
 * parent box (building.rs:134-155) and round-to-nearest leaf vertices (:183-195).  The node FORMAT and its quantisation rules are the
 * reference's; the split is not (the reference's agglomerative SAH merge is sequential), so the tree differs from mp_bvh_build_*'s.
 * out_device_ms (may be NULL): device time of the build, inputs resident, read-back excluded. */
int mp_bvh_build_device(int device, const float* positions, const float* normals, const float* texcoords, uint32_t n_vertices,
                        const uint32_t* triangles, const uint32_t* tri_material, uint64_t n_triangles,
                        const MpMaterial* materials, uint32_t n_materials, MpFlatBvh** out, float* out_device_ms);
int mp_bvh_free(MpFlatBvh* bvh);
/* Structural statistics (node / packet counts, depth) — printing.rs:11-70 analogue, numbers only. */
typedef struct MpBvhStats {
    uint32_t n_inner, n_leaves, n_packets, n_triangles;
    uint32_t max_depth, max_stack;
    float mean_leaf_fill, mean_inner_fill;
} MpBvhStats;
int mp_bvh_stats(const MpFlatBvh* bvh, MpBvhStats* out);

/* ---- camera: reference src/camera.rs ---- */
typedef struct MpCamera { /* camera.rs:9-23 */
    float center[3];
    float forward[3];
    float up[3];
    float right[3];
    float focus_distance;
    float sensor_size;    /* metres */
    int32_t sensor_is_width; /* SensorSize::Width / Height */
    float focal_length;
    float f_number;
} MpCamera;
typedef struct MpCameraSampler { /* camera.rs:26-39 */
    float center[3];
    float up[3];
    float right[3];
    float film_origin_offset[3];
    float pixel_scale;
    float lens_radius;
    float lens_weight;
} MpCameraSampler;
int mp_camera_default(MpCamera* out);                                              /* camera.rs:42-52 */
int mp_camera_look_at(MpCamera* cam, const float center[3], const float look_at[3],
                      const float up[3]);                                           /* camera.rs:93-101 */
int mp_camera_look_direction(MpCamera* cam, const float center[3], const float forward[3],
                             const float up[3]);                                    /* camera.rs:104-116 */
int mp_camera_build_sampler(const MpCamera* cam, uint32_t res_x, uint32_t res_y,
                            MpCameraSampler* out);                                  /* camera.rs:123-146 */

/* ---- screen tiling: reference src/screen_block.rs ---- */
typedef struct MpTile { uint32_t min_x, min_y, max_x, max_y; } MpTile; /* ScreenBlock = AABB<Point2<u32>> */
/* tile_ordering (screen_block.rs:46-81): tiles sorted by distance to centre + Exp noise.
 * seed == 0 → noise drawn from the OS like the reference; otherwise deterministic.
 * Writes up to cap tiles, returns the tile count in *out_n. */
int mp_tile_ordering(uint32_t min_x, uint32_t min_y, uint32_t max_x, uint32_t max_y, uint32_t tile_size,
                     uint64_t seed, MpTile* out, size_t cap, size_t* out_n);

/* ---- device context ---- */
typedef struct MpCtx MpCtx;

typedef struct MpRay { /* geometry/mod.rs:34-42 */
    float origin[3];
    float direction[3];
    float inv_direction[3];
} MpRay; /* 36 B */
typedef struct MpHit { /* LeafHitRecord (ray_bvh_intersection.rs:165-171) without the normal */
    uint32_t tri;  /* packet*8+lane, MP_NO_HIT on miss */
    float t, u, v;
} MpHit; /* 16 B */

typedef struct MpHitRecord { /* HitRecord (geometry/mod.rs:70-80) */
    float t;
    float point[3];
    float normal[3];
    uint32_t material;
    float texture_coords[3];
    int32_t hit; /* 0 = miss (Option::None), every other field zero */
} MpHitRecord; /* 48 B */

#define MP_MODE_MEGAKERNEL 0u /* one thread per pixel runs the whole sample loop */
#define MP_MODE_WAVEFRONT 1u  /* ray-gen → trace → shade → compact → … → film accumulate */
/* mp_render only: the library picks the pipeline from the scene and the settings — the megakernel for the reference's own
 * workload (max_bounces = 0) and for scenes that are a single leaf, else the wavefront pipeline with MP_TRACE_HYBRID_POOL_S0
 * (MP_TRACE_POOL when the BVH is deeper than the thread-per-ray kernels' 128-entry stack).  Same film whatever is picked. */
#define MP_MODE_AUTO 2u
#define MP_ORDER_REFERENCE 0u /* exact LIFO / lane order of ray_bvh_intersection.rs:39-64,161 (bit-exact gate) */
#define MP_ORDER_NEAR_FIRST 1u /* children pushed far→near; not bit-pinned, mismatches are reported */

typedef struct MpRenderParams {
    uint32_t sample_count;   /* RenderSettings::sample_count for this call (renderer/mod.rs:9) */
    uint32_t sample_base;    /* index of the first sample (spp split across GPUs / progressive passes) */
    uint32_t max_bounces;    /* 0 = the reference shader (worker.rs:57-64); >0 = extension integrator */
    uint32_t mode;           /* MP_MODE_* */
    uint32_t order;          /* MP_ORDER_* */
    uint32_t tile_size;      /* RenderSettings::tile_size: upper bound of every tile's width and height */
    uint64_t seed;
} MpRenderParams;

typedef struct MpCounters {
    uint64_t rays;          /* ray segments traced (camera + bounce + shadow) */
    uint64_t camera_rays;
    uint64_t shadow_rays;
    uint64_t inner_visits;  /* inner nodes decoded  (128 B each) */
    uint64_t packet_visits; /* triangle packets tested (144 B each) */
    uint64_t leaf_visits;
    uint64_t launches;      /* kernels launched by this library */
} MpCounters;

const char* mp_last_error(void);
int mp_device_count(int* out);
int mp_ctx_create(int device, MpCtx** out);
int mp_ctx_destroy(MpCtx* ctx);
/* Process-wide pool of idle contexts (streams, film, path-state and staging buffers stay allocated between renders).
 * mp_render() takes its per-device worker contexts from here. */
int mp_ctx_pool_acquire(int device, MpCtx** out);
int mp_ctx_pool_release(MpCtx* ctx);
int mp_ctx_pool_clear(void);
/* Run on a caller-owned CUDA stream (cudaStream_t as void*); NULL → the context's own stream. */
int mp_ctx_set_stream(MpCtx* ctx, void* cuda_stream);
int mp_ctx_sync(MpCtx* ctx);
/* The stream the context launches on (cudaStream_t as void*): what a caller hands to its own collectives or copies. */
int mp_ctx_stream(MpCtx* ctx, void** out_cuda_stream);
/* Which traversal kernel serves mp_trace_rays* and the wavefront pipeline (same visiting order, same bits):
 * THREAD = one thread per ray, SoA node rows, local-memory stack (trace.cuh; default);
 * OCTET  = eight lanes per ray, child-major 16-byte records, shared-memory stack, self-refilling octets (octet.cuh). */
#define MP_TRACE_THREAD 0u
#define MP_TRACE_OCTET 1u
/* PERSIST = one thread per ray, persistent lanes that refill themselves, warp-voted inner / packet phases (persist.cuh) */
#define MP_TRACE_PERSIST 2u
/* HYBRID = THREAD for camera rays (coherent), PERSIST for bounce and shadow rays (incoherent) */
#define MP_TRACE_HYBRID 3u
/* POOL = each warp owns a pool of 96 rays (state in shared memory, stacks in a global scratch) and runs 32 work items of ONE
 * kind (inner node / triangle packet) in lockstep per round (pool.cuh); HYBRID_POOL = THREAD for camera rays, POOL after */
#define MP_TRACE_POOL 4u
#define MP_TRACE_HYBRID_POOL 5u
/* HYBRID_POOL_S0 = HYBRID_POOL whose camera-ray kernel also traces the depth-0 sun rays in place (they leave neighbouring
 * pixels' hit points in ONE direction, so they are as coherent as the camera rays): no queue round trip and no pool launch
 * for the largest shadow batch of a pass.  Same film, same counters. */
#define MP_TRACE_HYBRID_POOL_S0 6u
int mp_ctx_set_trace_variant(MpCtx* ctx, uint32_t variant);
/* Node layout the THREAD / PERSIST / POOL kernels and the megakernel read (same results, bit for bit):
 * QUANTISED = the reference records (InnerNode 128 B, RelativeTriangle8 144 B; u16 → f32 on every visit,
 *             compressed_geometry.rs:95-110,160-162), each interleaved with its enclosing box into one 160 B / 192 B record
 *             that a lane gathers with 256-bit loads — fewest bytes per visit, best for incoherent rays (L1-bound gathers);
 * DECODED   = the dequantisation hoisted to upload time: f32 child boxes (256 B per node) and f32 v0 / e1 / e2 per triangle
 *             (288 B per packet), written once by a device kernel with the same fma — ~2x fewer instructions per visit, best
 *             for coherent rays (issue-bound);
 * AUTO      = DECODED for the thread-per-ray kernels on camera rays and the megakernel, QUANTISED for PERSIST / POOL unless the
 *             decoded BVH is at most 1 MB (default). */
#define MP_LAYOUT_QUANTISED 0u
#define MP_LAYOUT_DECODED 1u
#define MP_LAYOUT_AUTO 2u
int mp_ctx_set_node_layout(MpCtx* ctx, uint32_t layout);
/* Count traversal work (adds a few atomics per warp); off by default. */
int mp_ctx_enable_counters(MpCtx* ctx, int on);
int mp_counters(MpCtx* ctx, MpCounters* out);
int mp_counters_reset(MpCtx* ctx);
/* Device time (ms, CUDA events on the launching stream) of the last render/trace call's kernels. */
int mp_last_device_ms(MpCtx* ctx, float* out_ms);

/* Roofline denominators not in MEASURED_PEAKS.json, measured on this device (no reference counterpart):
 * FP32 FMA throughput (TFLOP/s, 2 flop per FMA) and L2 read bandwidth (GB/s) over a buffer of `bytes`. */
int mp_probe_fp32_tflops(MpCtx* ctx, float* out_tflops);
int mp_probe_l2_gbs(MpCtx* ctx, size_t bytes, float* out_gbs);
/* One shape of the L2 probe (1..4 blocks of 512 threads per SM, 16- or 32-byte loads): the caller sweeps shapes and buffer sizes
 * and takes the maximum as the L2 roof. */
int mp_probe_l2_gbs_shape(MpCtx* ctx, size_t bytes, uint32_t blocks_per_sm, uint32_t vec_bytes, float* out_gbs);
/* L1 gather roof of an incoherent traversal: every lane fetches its own 5- or 6-sector record (160 / 192 bytes) with 256-bit
 * loads; result in 10^9 32-byte sectors per second over the whole GPU. */
int mp_probe_l1_gather(MpCtx* ctx, uint32_t n_records, uint32_t sectors_per_record, float* out_gsectors_per_s);

/* Scene<TriangleBvh> → device (borrowed host arrays are copied; caller keeps ownership). */
int mp_scene_upload(MpCtx* ctx, const MpFlatBvh* bvh);
/* CameraSampler for the render (camera.build_sampler(resolution), machinery.rs:65). */
int mp_camera_set(MpCtx* ctx, const MpCameraSampler* sampler);

/* Film = f32x4 accumulator per pixel (the `pixel_sum` of worker.rs:40-43 kept on the device). */
int mp_film_create(MpCtx* ctx, uint32_t width, uint32_t height);
/* Use caller-owned device memory (width*height*4 floats) as the film, e.g. a tensor that a collective reduces. */
int mp_film_bind_external(MpCtx* ctx, void* device_ptr, uint32_t width, uint32_t height);
int mp_film_clear(MpCtx* ctx);
int mp_film_device_ptr(MpCtx* ctx, void** out);

/* Worker::render_tile for n tiles at once (worker.rs:32-49): accumulates sample_count samples per pixel
 * into the film, sequentially per pixel in sample order.  Asynchronous on the context's stream. */
int mp_render_tiles(MpCtx* ctx, const MpTile* tiles, uint32_t n_tiles, const MpRenderParams* params);
/* The render's whole tile_ordering kept on the device (RenderState::tile_ordering, machinery.rs:36-41); batches are then
 * addressed by index range, like the workers' `next_tile_index` (machinery.rs:205-209).  Asynchronous. */
int mp_tiles_upload(MpCtx* ctx, const MpTile* tiles, uint32_t n_tiles, uint32_t tile_size);
int mp_render_tile_range(MpCtx* ctx, uint32_t first, uint32_t n_tiles, const MpRenderParams* params);
/* Tile-packed RGBA8 read-back of a range, double-buffered (slot 0/1): _begin queues resolve + copy into pinned host
 * memory, _end waits for that slot and returns the buffer [n_tiles][tile_size][tile_size] (valid until the slot is reused).
 * Lets the next batch render while the previous one is copied into the image (machinery.rs:78-89). */
int mp_film_resolve_tiles_begin(MpCtx* ctx, uint32_t first, uint32_t n_tiles, uint32_t tile_size, uint32_t spp_total, int slot);
int mp_film_resolve_tiles_end(MpCtx* ctx, int slot, const uint8_t** host_tiles);
/* pixel_sum * (1/spp) then color_to_image (worker.rs:44,69-76) for the whole film → host RGBA8. Synchronises. */
int mp_film_resolve_rgba8(MpCtx* ctx, uint32_t spp_total, uint8_t* host_out);
/* Same, into caller-owned device memory (width*height RGBA8 words); asynchronous on the context's stream. */
int mp_film_resolve_device(MpCtx* ctx, uint32_t spp_total, void* dev_out);
/* Same for n tiles only, into a tile-packed host buffer [n_tiles][tile_size][tile_size] RGBA8 — the worker's
 * `buffer` of machinery.rs:66-67 for a whole batch; only the first (w x h) of each tile slot is written. Synchronises. */
int mp_film_resolve_tiles_rgba8(MpCtx* ctx, const MpTile* tiles, uint32_t n_tiles, uint32_t tile_size, uint32_t spp_total,
                                uint8_t* host_out);
int mp_film_read_f32(MpCtx* ctx, float* host_out);

/* Object::intersect on caller rays (scene/mod.rs:7-10; ray_bvh_intersection.rs:26-64) — the bit-exactness gate. */
int mp_trace_rays(MpCtx* ctx, const MpRay* host_rays, uint64_t n, uint32_t order, MpHit* host_out);
/* Device-resident variant (pointers from mp_dev_alloc or any CUDA allocation). Asynchronous. */
int mp_trace_rays_device(MpCtx* ctx, const MpRay* dev_rays, uint64_t n, uint32_t order, MpHit* dev_out);
/* Object::intersect including the hit finalisation (ray_bvh_intersection.rs:66-95; triangle.rs:229-244): point, shading
 * normal, material id and interpolated texture coordinates per ray — the full HitRecord, bit for bit. */
int mp_intersect_rays(MpCtx* ctx, const MpRay* host_rays, uint64_t n, uint32_t order, MpHitRecord* host_out);
/* mp_trace_rays_device without any allocation inside the call.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and bits as mp_trace_rays_device with MP_TRACE_POOL (dev_rays / dev_out are the caller's device buffers;
 * asynchronous on the context's stream), but the pool kernel's scratch (per-ray traversal stacks, claim counter) lives in ONE
 * static __device__ array of MP_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; the persistent grid is
 * sized to what fits.  Limits, rejected with MP_EINVAL and a message: n <= MP_NOALLOC_MAX_RAYS, and a BVH whose traversal stack
 * needs at most MP_NOALLOC_MAX_STACK entries per ray (MpBvhStats.max_stack; the stand-in needs 108).  The scene must have been
 * uploaded before (mp_scene_upload allocates; this call never does: no cudaMalloc / cudaFree / allocating library routine).
 * All calls on one device share the array: the call clears what must start at zero itself, and calls that OVERLAP on the GPU
 * (two streams, two contexts of one device) race on it — serialise them. */
#define MP_NOALLOC_SCRATCH_BYTES 67108864u /* 64 MiB */
#define MP_NOALLOC_MAX_STACK 128u
#define MP_NOALLOC_MAX_RAYS 0x40000000u
int mp_trace_rays_device_noalloc(MpCtx* ctx, const MpRay* dev_rays, uint64_t n, uint32_t order, MpHit* dev_out);
/* CameraSampler::sample_ray (camera.rs:176-191) for every (pixel, sample) of the tiles, in
 * tile → row-major pixel → sample order; writes n_pixels*sample_count rays to the host. */
int mp_generate_rays(MpCtx* ctx, const MpTile* tiles, uint32_t n_tiles, const MpRenderParams* params,
                     MpRay* host_out, uint64_t cap, uint64_t* out_n);

/* Raw device memory helpers for harnesses that keep inputs resident in HBM. */
int mp_dev_alloc(MpCtx* ctx, size_t bytes, void** out);
int mp_dev_free(MpCtx* ctx, void* ptr);
int mp_dev_upload(MpCtx* ctx, void* dev_dst, const void* host_src, size_t bytes);
int mp_dev_download(MpCtx* ctx, void* host_dst, const void* dev_src, size_t bytes);
int mp_host_alloc_pinned(size_t bytes, void** out);
int mp_host_free_pinned(void* ptr);

/* ---- render(): reference src/renderer/machinery.rs:20-123 ---- */
typedef struct MpRenderSettings { /* renderer/mod.rs:7-13 + extension fields with reference-preserving defaults */
    uint32_t tile_size;
    uint32_t sample_count;
    uint32_t resolution_x, resolution_y;
    uint32_t max_bounces; /* 0 → reference behaviour */
    uint32_t mode;        /* MP_MODE_*, incl. MP_MODE_AUTO */
    uint32_t order;       /* MP_ORDER_* */
    uint32_t tiles_per_batch; /* 0 → default */
    uint32_t trace_kernel;    /* 0 → chosen by the library; otherwise MP_TRACE_* + 1 */
    uint32_t _pad;
    uint64_t seed;        /* 0 → OS-random like SmallRng::from_os_rng (worker.rs:25) */
} MpRenderSettings;
typedef struct MpProgressSnapshot { size_t finished, total; } MpProgressSnapshot; /* machinery.rs:180-183 */
typedef void (*MpStartedTileFn)(void* user, MpTile tile);
typedef void (*MpFinishedTileFn)(void* user, MpTile tile, MpProgressSnapshot progress);
typedef struct MpRenderProgress MpRenderProgress;

/* Starts rendering on background worker threads (one per device in `devices`) and returns immediately.
 * Callbacks fire from worker threads: started before a tile renders, finished after its pixels are in the image. */
int mp_render(const MpFlatBvh* scene, const MpCamera* camera, const MpRenderSettings* settings,
              const int* devices, int n_devices,
              MpStartedTileFn started, MpFinishedTileFn finished, void* user, MpRenderProgress** out);
int mp_progress_snapshot(MpRenderProgress* p, MpProgressSnapshot* out); /* machinery.rs:133-142 */
int mp_progress_is_finished(MpRenderProgress* p, int* out);             /* :144-146 */
int mp_progress_elapsed_ms(MpRenderProgress* p, double* out);           /* :150-157 */
int mp_progress_abort(MpRenderProgress* p);                             /* :161-165 */
int mp_progress_wait(MpRenderProgress* p);                              /* :169-173 */
/* Copies the RGBA8 image (w*h*4) under the image lock (:175-177). */
int mp_progress_image(MpRenderProgress* p, uint8_t* host_out, size_t cap);
/* The image itself (pinned host memory, w*h*4 bytes, valid until mp_progress_free) — what `image()` hands out behind its
 * Mutex (:175-177); tiles appear in it while the render runs, it is complete once mp_progress_wait has returned. */
int mp_progress_image_ptr(MpRenderProgress* p, const uint8_t** out, size_t* out_bytes);
int mp_progress_free(MpRenderProgress* p);

/* Image output (the reference never saves its image, src/cli.rs:44-46): 8-bit RGBA PNG with stored deflate
 * blocks, or binary PPM without alpha.  Headless — bytes to a file. */
int mp_image_write_png(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height);
int mp_image_write_ppm(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height);

/* Cross-process dynamic tile queue (the AtomicUsize of machinery.rs:43,205-209 shared by the ranks of one box). */
typedef struct MpTileQueue MpTileQueue;
int mp_tileq_open(const char* name, int create, MpTileQueue** out);
int mp_tileq_reset(MpTileQueue* q);
int mp_tileq_next(MpTileQueue* q, uint64_t batch, uint64_t* out_first); /* fetch_add(batch) */
int mp_tileq_peek(MpTileQueue* q, uint64_t* out_next);                   /* load() */
/* Barrier over the `world` processes that share the segment (used to rewind the counter between renders); timeout_ms = 0 waits forever. */
int mp_tileq_barrier(MpTileQueue* q, uint32_t world, uint32_t timeout_ms);
int mp_tileq_close(MpTileQueue* q, int unlink);

/* ---- multi-GPU groups: render()'s scheduler over the GPUs of one box (machinery.rs:51-116,205-209) ----
 * Workers = GPUs.  MP_SHARD_TILES: the devices pull batches of consecutive tiles of the tile list from ONE shared counter
 * (a device's first claim is 3/4 of an equal share in one piece, then remaining / (2 * devices) tiles at a time, at least min_batch — 0 = as many tiles
 * as carry 2^20 paths —: coarse while work is plentiful, fine at the tail) — total work
 * fixed, "strong" scaling (BASELINE configs[2]).  MP_SHARD_SPP: every device renders all tiles with its own contiguous slice of
 * the sample indices (BASELINE configs[3]).  Either way each device accumulates into its own zero-initialised full-resolution
 * film and ONE ncclReduce(sum, f32 x 4 per pixel) over NVLink combines them on rank 0, which resolves.  No traffic between GPUs
 * during traversal.  NCCL is dlopen()ed when the first group with more than one device is made. */
typedef struct MpGroup MpGroup;
#define MP_SHARD_TILES 0u
#define MP_SHARD_SPP 1u
#define MP_NCCL_ID_BYTES 128
/* One process, n devices (NULL / 0: all): a host thread per device during mp_group_render, ncclCommInitAll. */
int mp_group_create(const int* devices, int n_devices, MpGroup** out);
/* One process per GPU (e.g. torchrun): rank 0 makes the NCCL id and the caller ships its 128 bytes to the other ranks; the tile
 * counter lives in POSIX shared memory under queue_name (created by rank 0, which must therefore come first: create it before
 * the id is shipped). */
int mp_group_unique_id(uint8_t out[MP_NCCL_ID_BYTES]);
int mp_group_create_rank(int device, int rank, int world, const uint8_t id[MP_NCCL_ID_BYTES], const char* queue_name, MpGroup** out);
int mp_group_destroy(MpGroup* g);
int mp_group_size(MpGroup* g, int* out_local, int* out_world);
int mp_group_ctx(MpGroup* g, int local_index, MpCtx** out);  /* borrowed: counters, variants, streams of one member */
int mp_group_scene_upload(MpGroup* g, const MpFlatBvh* bvh);   /* replicated on every device */
int mp_group_camera_set(MpGroup* g, const MpCameraSampler* sampler);
int mp_group_set_trace_variant(MpGroup* g, uint32_t variant);
int mp_group_film_create(MpGroup* g, uint32_t width, uint32_t height); /* zeroed films; same size again = clear */
/* Launches this process's share of the frame and returns when its claim loops have ended (the last batches may still be running
 * on the devices; everything later is stream-ordered behind them). */
int mp_group_render(MpGroup* g, const MpTile* tiles, uint32_t n_tiles, const MpRenderParams* params, uint32_t shard, uint32_t min_batch);
int mp_group_reduce_film(MpGroup* g);                                  /* ncclReduce(sum) to rank 0, in place, asynchronous */
int mp_group_resolve_rgba8(MpGroup* g, uint32_t spp_total, uint8_t* host_out); /* rank 0: resolve + D2H; other ranks: drain */
int mp_group_read_film(MpGroup* g, int local_index, float* host_out);
int mp_group_sync(MpGroup* g);
/* Device time (CUDA events) from the start of the last mp_group_render to the end of its last batch — or of the reduce, if one
 * followed — maximum over this process's devices. */
int mp_group_last_ms(MpGroup* g, float* out_ms);
int mp_group_member_stats(MpGroup* g, int local_index, uint64_t* out_batches, uint64_t* out_tiles, float* out_ms);

#ifdef __cplusplus
}
#endif
#endif /* MINIPATH_B200_H */

/* oracle/ — CPU restatement of minipath's per-pixel trace path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may load this
 * library; nothing under minipath_b200/ links, imports or calls it.
 *
 * The reference is pure Rust and cannot be compiled in this image (no cargo/rustc), so this is a C++ restatement,
 * function by function, of the reference sources cited at each declaration (paths relative to the reference
 * tree).  Pinning status:
 *   - PINNED by the reference's own known-answer tests (ported in tests/test_oracle_kat.py): AABB slab test
 *     (src/geometry/aabb.rs:374-471), unit-interval quantisation (compressed_geometry.rs:190-200), node-link
 *     encoding (triangle_bvh/mod.rs:189-236), bit_iter (util/mod.rs:40-56), camera orientation
 *     (camera.rs:200-226), fma_dot/fma_cross vs unfused (util/simba.rs:127-176).
 *   - PARITY UNPINNED (no reference test, golden vector or runnable reference exists): Triangle8::intersect,
 *     TriangleBvh::intersect, hit finalisation, render_tile / color_to_image, the builder.  There the oracle IS the
 *     restatement.  What narrows the gap (tests/test_independent.py): the traversal is checked against a brute-force closest
 *     hit over all packets and against plain numpy Moller-Trumbore on the SOURCE mesh; the builder (oracle_build.inc) and the
 *     product's builder are two separate restatements that must agree byte for byte and satisfy the format's invariants.  The reference RNG is OS-seeded (worker.rs:25), so no seed-exact image exists even
 *     reference-vs-reference; oracle and CUDA share a counter-based RNG (below) so that they can be compared
 *     bit for bit.
 *   - max_bounces > 0 is an extension with no reference behaviour at all.
 *
 * Arithmetic contract: f32, round-to-nearest-even, compiled with -ffp-contract=off; fused multiply-adds
 * appear only as explicit fmaf at the sites where the reference calls mul_add / mul_sub.
 */
#ifndef MINIPATH_ORACLE_H
#define MINIPATH_ORACLE_H

#include "../include/minipath_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- small pieces, exposed for the known-answer tests ---- */
/* UnitInterval8::compress_internal / decompress, one lane (compressed_geometry.rs:25-51). rounding: 0 round, 1 floor, 2 ceil */
uint16_t orc_compress_unit(float v, int rounding);
float orc_decompress_unit_rel(uint16_t q);
float orc_decompress_coord(uint16_t q, float size, float min);
/* Ray::new (geometry/mod.rs:44-54) */
void orc_ray_new(const float origin[3], const float direction[3], MpRay* out);
/* AABB<WorldPoint8>::intersect (aabb.rs:257-283) for 8 boxes given SoA: bmin[3][8], bmax[3][8]. */
void orc_aabb8_intersect(const float bmin[3][8], const float bmax[3][8], const MpRay* ray, float max_t, float t1[8],
                         float t2[8]);
/* Triangle<WorldPoint8>::intersect (triangle.rs:187-216): v[vertex][axis][lane]; returns the lane mask. */
uint32_t orc_triangle8_intersect(const float v[3][3][8], const MpRay* ray, float t[8], float u[8], float w[8]);
/* fma_dot / fma_cross (util/simba.rs:57-67), one lane */
float orc_fma_dot(const float a[3], const float b[3]);
void orc_fma_cross(const float a[3], const float b[3], float out[3]);
/* bit_iter (util/mod.rs:6-31): writes set-bit positions low→high, returns their number */
int orc_bit_iter(uint64_t bits, int out[64]);
/* CompressedNodeLink (triangle_bvh/mod.rs:66-114); return 0 ok, -1 where the reference panics */
int orc_link_new_leaf(uint32_t index, uint32_t count, uint32_t* out);
int orc_link_new_inner(uint32_t index, uint32_t* out);
int orc_link_decode(uint32_t link, int* kind, uint32_t* index, uint32_t* count);
/* color_to_image (worker.rs:69-76) for one channel */
uint8_t orc_color_channel_to_u8(float c);

/* ---- counter-based RNG shared with the CUDA path (the reference's SmallRng is OS-seeded) ---- */
uint32_t orc_pcg_hash(uint32_t v);
void orc_rng_init(uint32_t pixel_index, uint32_t sample_index, uint64_t seed, uint32_t rng[2]);
uint32_t orc_rng_next(uint32_t rng[2]);

/* ---- the path ---- */
/* 0 = scalar lane loops, 1 = AVX2+FMA intrinsics (bit-identical; the CPU baseline uses 1) */
void orc_set_simd(int on);
int orc_get_simd(void);

/* TriangleBvh::intersect without finalisation (ray_bvh_intersection.rs:26-64,104-162).  The stack carries the
 * child boxes exactly as the reference does — the bvh's box tables are NOT read here. */
void orc_trace_rays(const MpFlatBvh* bvh, const MpRay* rays, uint64_t n, uint32_t order, MpHit* out, MpCounters* counters,
                    int threads);

typedef struct OrcHitRecord { /* geometry/mod.rs:71-80 */
    float t;
    float point[3];
    float normal[3];
    uint32_t material;
    float texture_coords[3];
    int hit;
} OrcHitRecord;
/* Object::intersect including finalisation (ray_bvh_intersection.rs:66-95) */
void orc_intersect(const MpFlatBvh* bvh, const MpRay* ray, uint32_t order, OrcHitRecord* out);
void orc_intersect_rays(const MpFlatBvh* bvh, const MpRay* rays, uint64_t n, uint32_t order, OrcHitRecord* out, int threads);

/* CameraSampler::sample_ray (camera.rs:176-191) with the shared RNG */
void orc_sample_ray(const MpCameraSampler* cam, uint32_t px, uint32_t py, uint32_t* rng_state, MpRay* out);
/* rays for all (pixel, sample) of the tiles in tile → row-major pixel → sample order */
uint64_t orc_generate_rays(const MpCameraSampler* cam, uint32_t width, const MpTile* tiles, uint32_t n_tiles,
                           const MpRenderParams* params, MpRay* out, uint64_t cap);

/* render(): tile queue on `threads` worker threads (machinery.rs:51-116), Worker::render_tile (worker.rs:32-49)
 * accumulating into film (w*h*4 floats, NOT cleared).  Returns wall seconds. */
double orc_render_tiles(const MpFlatBvh* bvh, const MpCameraSampler* cam, uint32_t width, uint32_t height,
                        const MpTile* tiles, uint32_t n_tiles, const MpRenderParams* params, float* film,
                        MpCounters* counters, int threads);
/* pixel_sum*(1/spp) + color_to_image for the whole film (worker.rs:44,69-76) */
void orc_resolve_rgba8(const float* film, uint32_t width, uint32_t height, uint32_t spp_total, uint8_t* out);

/* ---- scene sources and checks that do not share code with the product (oracle_build.inc) ---- */
/* TriangleBvh::with_obj / build (building.rs:28-207,238-450) restated on its own; quad_mode 1 fan-triangulates polygons.
 * Returns 0, or -1 with orc_build_error() where the reference would panic.  The arrays carry the box tables too, so the same
 * MpFlatBvh can be uploaded to the GPU: the device then traverses a tree the product did not build. */
int orc_bvh_build_obj(const char* path, int quad_mode, MpFlatBvh** out);
/* with texture coordinates (uv pairs or NULL), a material id per triangle (or NULL) and the material table */
int orc_bvh_build_mesh_mat(const float* positions, const float* normals, const float* texcoords, uint32_t n_vertices, const uint32_t* triangles,
                           const uint32_t* tri_material, uint64_t n_triangles, const MpMaterial* materials, uint32_t n_materials, MpFlatBvh** out);
int orc_bvh_build_mesh(const float* positions, const float* normals, uint32_t n_vertices, const uint32_t* triangles,
                       uint64_t n_triangles, MpFlatBvh** out);
void orc_bvh_free(MpFlatBvh* bvh);
const char* orc_build_error(void);
/* Closest hit over ALL packets with Triangle8::intersect, no inner nodes, no traversal (lowest triangle index among equal t). */
void orc_brute_force(const MpFlatBvh* bvh, const MpRay* rays, uint64_t n, MpHit* out, int threads);
/* Camera::look_at + build_sampler (camera.rs:93-116,123-146) and a row-major tile list (screen_block.rs:46-60,146-160): what the
 * CPU arms of bench.py need to run without mapping the product library. */
void orc_camera_sampler(const float center[3], const float target[3], const float up_hint[3], float sensor_size, int sensor_is_width,
                        float focal_length, float f_number, float focus_distance, uint32_t res_x, uint32_t res_y, MpCameraSampler* out);
uint32_t orc_tiles(uint32_t width, uint32_t height, uint32_t tile_size, MpTile* out, uint32_t cap);

#ifdef __cplusplus
}
#endif
#endif

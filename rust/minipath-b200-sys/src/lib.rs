//! Raw bindings to `include/minipath_b200.h` — one declaration per entry point, `#[repr(C)]` structs field for field.
//! Every function returns 0 (`MP_OK`) or a negative `MP_E*` code; the message is `mp_last_error()` (thread-local).
//! Nothing here was compiled in the repository's image (no Rust toolchain there); `tests/test_rust_sys.py` checks this
//! file against the header: same function set, same argument counts, same struct sizes.
#![allow(non_camel_case_types, non_snake_case, clippy::missing_safety_doc)]

use std::os::raw::{c_char, c_int, c_void};

pub const MP_OK: c_int = 0;
pub const MP_EINVAL: c_int = -1;
pub const MP_ECUDA: c_int = -2;
pub const MP_ENODEVICE: c_int = -3;
pub const MP_EIO: c_int = -4;
pub const MP_ESTATE: c_int = -5;
pub const MP_ENOMEM: c_int = -6;

// CompressedNodeLink (src/scene/triangle_bvh/mod.rs:57-114)
pub const MP_LINK_COUNT_BITS: u32 = 3;
pub const MP_LINK_COUNT_MASK: u32 = 7;
pub const MP_LINK_NULL: u32 = 0xFFFF_FFF8;
pub const MP_LINK_MAX_INDEX: u32 = 0x1FFF_FFFE;
pub const MP_LINK_MAX_COUNT: u32 = 7;
pub const MP_NO_HIT: u32 = 0xFFFF_FFFF;
pub const MP_INNER_NODE_CHILDREN: usize = 8;
pub const MP_LEAF_PACKET_SIZE: usize = 8;
pub const MP_LEAF_MAX_TRIANGLES: usize = 56;
pub const MP_TRI_FLAT: u32 = 0x8000_0000;

pub const MP_MODE_MEGAKERNEL: u32 = 0;
pub const MP_MODE_WAVEFRONT: u32 = 1;
pub const MP_MODE_AUTO: u32 = 2;
pub const MP_ORDER_REFERENCE: u32 = 0;
pub const MP_ORDER_NEAR_FIRST: u32 = 1;
pub const MP_TRACE_THREAD: u32 = 0;
pub const MP_TRACE_OCTET: u32 = 1;
pub const MP_TRACE_PERSIST: u32 = 2;
pub const MP_TRACE_HYBRID: u32 = 3;
pub const MP_TRACE_POOL: u32 = 4;
pub const MP_TRACE_HYBRID_POOL: u32 = 5;
pub const MP_TRACE_HYBRID_POOL_S0: u32 = 6;
pub const MP_LAYOUT_QUANTISED: u32 = 0;
pub const MP_LAYOUT_DECODED: u32 = 1;
pub const MP_LAYOUT_AUTO: u32 = 2;
pub const MP_SHARD_TILES: u32 = 0;
pub const MP_SHARD_SPP: u32 = 1;
pub const MP_NCCL_ID_BYTES: usize = 128;

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpBox {
    pub min: [f32; 3],
    pub _pad0: f32,
    pub size: [f32; 3],
    pub _pad1: f32,
}

/// `InnerNode` (mod.rs:33-38): RelativeBox8 {min.x,min.y,min.z,max.x,max.y,max.z} of u16x8, then 8 links.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpInnerNode {
    pub bounds: [[u16; 8]; 6],
    pub links: [u32; 8],
}

/// `RelativeTriangle8` (compressed_geometry.rs:147): [vertex][axis][lane].
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpTrianglePacket {
    pub v: [[[u16; 8]; 3]; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpTriShading {
    pub v: [u32; 3],
    pub flags_material: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpVertexShading {
    pub normal: [f32; 3],
    pub _pad0: f32,
    pub tex: [f32; 3],
    pub _pad1: f32,
}

/// Material table entry (extension): albedo = Kd, MP_MAT_CHECKER halves it on the odd cells of the texture-coordinate grid.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpMaterial {
    pub albedo: [f32; 3],
    pub flags: u32,
    pub checker_scale: f32,
    pub _pad: [f32; 3],
}
pub const MP_MAT_CHECKER: u32 = 1;
pub const MP_NOALLOC_SCRATCH_BYTES: u32 = 67108864;
pub const MP_NOALLOC_MAX_STACK: u32 = 128;
pub const MP_NOALLOC_MAX_RAYS: u64 = 0x40000000;
pub const MP_TRI_MATERIAL_MASK: u32 = 0x7fff_ffff;
pub const MP_DEFAULT_ALBEDO: f32 = 0.7;
pub const MP_DEFAULT_CHECKER_SCALE: f32 = 4.0;

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct MpFlatBvh {
    pub bbox_min: [f32; 3],
    pub bbox_max: [f32; 3],
    pub root: u32,
    pub n_inner: u32,
    pub n_packets: u32,
    pub n_vertices: u32,
    pub n_triangles: u32,
    pub n_materials: u32,
    pub inner: *const MpInnerNode,
    pub inner_box: *const MpBox,
    pub packets: *const MpTrianglePacket,
    pub packet_box: *const MpBox,
    pub tri_shading: *const MpTriShading,
    pub vertices: *const MpVertexShading,
    pub materials: *const MpMaterial,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpBvhStats {
    pub n_inner: u32,
    pub n_leaves: u32,
    pub n_packets: u32,
    pub n_triangles: u32,
    pub max_depth: u32,
    pub max_stack: u32,
    pub mean_leaf_fill: f32,
    pub mean_inner_fill: f32,
}

/// `Camera` (src/camera.rs:9-23).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpCamera {
    pub center: [f32; 3],
    pub forward: [f32; 3],
    pub up: [f32; 3],
    pub right: [f32; 3],
    pub focus_distance: f32,
    pub sensor_size: f32,
    pub sensor_is_width: i32,
    pub focal_length: f32,
    pub f_number: f32,
}

/// `CameraSampler` (src/camera.rs:26-39).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpCameraSampler {
    pub center: [f32; 3],
    pub up: [f32; 3],
    pub right: [f32; 3],
    pub film_origin_offset: [f32; 3],
    pub pixel_scale: f32,
    pub lens_radius: f32,
    pub lens_weight: f32,
}

/// `ScreenBlock` = `AABB<Point2<u32>>` (src/screen_block.rs).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default, PartialEq, Eq)]
pub struct MpTile {
    pub min_x: u32,
    pub min_y: u32,
    pub max_x: u32,
    pub max_y: u32,
}

/// `Ray` (src/geometry/mod.rs:34-42).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpRay {
    pub origin: [f32; 3],
    pub direction: [f32; 3],
    pub inv_direction: [f32; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpHit {
    pub tri: u32,
    pub t: f32,
    pub u: f32,
    pub v: f32,
}

/// HitRecord (geometry/mod.rs:70-80); hit == 0 is Option::None.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpHitRecord {
    pub t: f32,
    pub point: [f32; 3],
    pub normal: [f32; 3],
    pub material: u32,
    pub texture_coords: [f32; 3],
    pub hit: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpRenderParams {
    pub sample_count: u32,
    pub sample_base: u32,
    pub max_bounces: u32,
    pub mode: u32,
    pub order: u32,
    pub tile_size: u32,
    pub seed: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpCounters {
    pub rays: u64,
    pub camera_rays: u64,
    pub shadow_rays: u64,
    pub inner_visits: u64,
    pub packet_visits: u64,
    pub leaf_visits: u64,
    pub launches: u64,
}

/// `RenderSettings` (src/renderer/mod.rs:7-13) plus the extension fields, whose zero defaults keep reference behaviour.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpRenderSettings {
    pub tile_size: u32,
    pub sample_count: u32,
    pub resolution_x: u32,
    pub resolution_y: u32,
    pub max_bounces: u32,
    pub mode: u32,
    pub order: u32,
    pub tiles_per_batch: u32,
    pub trace_kernel: u32,
    pub _pad: u32,
    pub seed: u64,
}

/// `RenderProgressSnapshot` (src/renderer/machinery.rs:180-183).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct MpProgressSnapshot {
    pub finished: usize,
    pub total: usize,
}

#[repr(C)]
pub struct MpCtx {
    _opaque: [u8; 0],
}
#[repr(C)]
pub struct MpRenderProgress {
    _opaque: [u8; 0],
}
#[repr(C)]
pub struct MpTileQueue {
    _opaque: [u8; 0],
}
#[repr(C)]
pub struct MpGroup {
    _opaque: [u8; 0],
}

pub type MpStartedTileFn = Option<unsafe extern "C" fn(user: *mut c_void, tile: MpTile)>;
pub type MpFinishedTileFn = Option<unsafe extern "C" fn(user: *mut c_void, tile: MpTile, progress: MpProgressSnapshot)>;

extern "C" {
    pub fn mp_last_error() -> *const c_char;

    // node links (triangle_bvh/mod.rs:78-109)
    pub fn mp_link_new_leaf(packet_index: u32, packet_count: u32, out_link: *mut u32) -> c_int;
    pub fn mp_link_new_inner(node_index: u32, out_link: *mut u32) -> c_int;
    pub fn mp_link_decode(link: u32, out_kind: *mut c_int, out_index: *mut u32, out_count: *mut u32) -> c_int;

    // host-side scene loading (building.rs:28-120) for callers that do not bring a TriangleBvh of their own
    pub fn mp_bvh_build_obj(path: *const c_char, quad_mode: c_int, out: *mut *mut MpFlatBvh) -> c_int;
    pub fn mp_bvh_build_mesh(positions: *const f32, normals: *const f32, n_vertices: u32, triangles: *const u32, n_triangles: u64, out: *mut *mut MpFlatBvh) -> c_int;
    pub fn mp_bvh_build_mesh_ex(positions: *const f32, normals: *const f32, n_vertices: u32, triangles: *const u32, n_triangles: u64, builder: c_int, out: *mut *mut MpFlatBvh) -> c_int;
    pub fn mp_bvh_build_mesh_mat(positions: *const f32, normals: *const f32, texcoords: *const f32, n_vertices: u32, triangles: *const u32, tri_material: *const u32, n_triangles: u64, materials: *const MpMaterial, n_materials: u32, builder: c_int, out: *mut *mut MpFlatBvh) -> c_int;
    pub fn mp_bvh_build_device(device: c_int, positions: *const f32, normals: *const f32, texcoords: *const f32, n_vertices: u32, triangles: *const u32, tri_material: *const u32, n_triangles: u64, materials: *const MpMaterial, n_materials: u32, out: *mut *mut MpFlatBvh, out_device_ms: *mut f32) -> c_int;
    pub fn mp_bvh_free(bvh: *mut MpFlatBvh) -> c_int;
    pub fn mp_bvh_stats(bvh: *const MpFlatBvh, out: *mut MpBvhStats) -> c_int;

    // camera (camera.rs)
    pub fn mp_camera_default(out: *mut MpCamera) -> c_int;
    pub fn mp_camera_look_at(cam: *mut MpCamera, center: *const f32, look_at: *const f32, up: *const f32) -> c_int;
    pub fn mp_camera_look_direction(cam: *mut MpCamera, center: *const f32, forward: *const f32, up: *const f32) -> c_int;
    pub fn mp_camera_build_sampler(cam: *const MpCamera, res_x: u32, res_y: u32, out: *mut MpCameraSampler) -> c_int;

    // tiling (screen_block.rs:46-81)
    pub fn mp_tile_ordering(min_x: u32, min_y: u32, max_x: u32, max_y: u32, tile_size: u32, seed: u64, out: *mut MpTile, cap: usize, out_n: *mut usize) -> c_int;

    // device contexts
    pub fn mp_device_count(out: *mut c_int) -> c_int;
    pub fn mp_ctx_create(device: c_int, out: *mut *mut MpCtx) -> c_int;
    pub fn mp_ctx_destroy(ctx: *mut MpCtx) -> c_int;
    pub fn mp_ctx_pool_acquire(device: c_int, out: *mut *mut MpCtx) -> c_int;
    pub fn mp_ctx_pool_release(ctx: *mut MpCtx) -> c_int;
    pub fn mp_ctx_pool_clear() -> c_int;
    pub fn mp_ctx_set_stream(ctx: *mut MpCtx, cuda_stream: *mut c_void) -> c_int;
    pub fn mp_ctx_sync(ctx: *mut MpCtx) -> c_int;
    pub fn mp_ctx_stream(ctx: *mut MpCtx, out_cuda_stream: *mut *mut c_void) -> c_int;
    pub fn mp_ctx_set_trace_variant(ctx: *mut MpCtx, variant: u32) -> c_int;
    pub fn mp_ctx_set_node_layout(ctx: *mut MpCtx, layout: u32) -> c_int;
    pub fn mp_ctx_enable_counters(ctx: *mut MpCtx, on: c_int) -> c_int;
    pub fn mp_counters(ctx: *mut MpCtx, out: *mut MpCounters) -> c_int;
    pub fn mp_counters_reset(ctx: *mut MpCtx) -> c_int;
    pub fn mp_last_device_ms(ctx: *mut MpCtx, out_ms: *mut f32) -> c_int;
    pub fn mp_probe_fp32_tflops(ctx: *mut MpCtx, out_tflops: *mut f32) -> c_int;
    pub fn mp_probe_l2_gbs(ctx: *mut MpCtx, bytes: usize, out_gbs: *mut f32) -> c_int;
    pub fn mp_probe_l2_gbs_shape(ctx: *mut MpCtx, bytes: usize, blocks_per_sm: u32, vec_bytes: u32, out_gbs: *mut f32) -> c_int;
    pub fn mp_probe_l1_gather(ctx: *mut MpCtx, n_records: u32, sectors_per_record: u32, out_gsectors_per_s: *mut f32) -> c_int;

    // scene, camera, film
    pub fn mp_scene_upload(ctx: *mut MpCtx, bvh: *const MpFlatBvh) -> c_int;
    pub fn mp_camera_set(ctx: *mut MpCtx, sampler: *const MpCameraSampler) -> c_int;
    pub fn mp_film_create(ctx: *mut MpCtx, width: u32, height: u32) -> c_int;
    pub fn mp_film_bind_external(ctx: *mut MpCtx, device_ptr: *mut c_void, width: u32, height: u32) -> c_int;
    pub fn mp_film_clear(ctx: *mut MpCtx) -> c_int;
    pub fn mp_film_device_ptr(ctx: *mut MpCtx, out: *mut *mut c_void) -> c_int;

    // Worker::render_tile for batches of tiles (worker.rs:32-49)
    pub fn mp_render_tiles(ctx: *mut MpCtx, tiles: *const MpTile, n_tiles: u32, params: *const MpRenderParams) -> c_int;
    pub fn mp_tiles_upload(ctx: *mut MpCtx, tiles: *const MpTile, n_tiles: u32, tile_size: u32) -> c_int;
    pub fn mp_render_tile_range(ctx: *mut MpCtx, first: u32, n_tiles: u32, params: *const MpRenderParams) -> c_int;
    pub fn mp_film_resolve_tiles_begin(ctx: *mut MpCtx, first: u32, n_tiles: u32, tile_size: u32, spp_total: u32, slot: c_int) -> c_int;
    pub fn mp_film_resolve_tiles_end(ctx: *mut MpCtx, slot: c_int, host_tiles: *mut *const u8) -> c_int;
    pub fn mp_film_resolve_rgba8(ctx: *mut MpCtx, spp_total: u32, host_out: *mut u8) -> c_int;
    pub fn mp_film_resolve_device(ctx: *mut MpCtx, spp_total: u32, dev_out: *mut c_void) -> c_int;
    pub fn mp_film_resolve_tiles_rgba8(ctx: *mut MpCtx, tiles: *const MpTile, n_tiles: u32, tile_size: u32, spp_total: u32, host_out: *mut u8) -> c_int;
    pub fn mp_film_read_f32(ctx: *mut MpCtx, host_out: *mut f32) -> c_int;

    // Object::intersect on caller rays (scene/mod.rs:7-10) and CameraSampler::sample_ray dumps (camera.rs:176-191)
    pub fn mp_trace_rays(ctx: *mut MpCtx, host_rays: *const MpRay, n: u64, order: u32, host_out: *mut MpHit) -> c_int;
    pub fn mp_trace_rays_device(ctx: *mut MpCtx, dev_rays: *const MpRay, n: u64, order: u32, dev_out: *mut MpHit) -> c_int;
    pub fn mp_trace_rays_device_noalloc(ctx: *mut MpCtx, dev_rays: *const MpRay, n: u64, order: u32, dev_out: *mut MpHit) -> c_int;
    pub fn mp_intersect_rays(ctx: *mut MpCtx, host_rays: *const MpRay, n: u64, order: u32, host_out: *mut MpHitRecord) -> c_int;
    pub fn mp_generate_rays(ctx: *mut MpCtx, tiles: *const MpTile, n_tiles: u32, params: *const MpRenderParams, host_out: *mut MpRay, cap: u64, out_n: *mut u64) -> c_int;

    // raw memory helpers
    pub fn mp_dev_alloc(ctx: *mut MpCtx, bytes: usize, out: *mut *mut c_void) -> c_int;
    pub fn mp_dev_free(ctx: *mut MpCtx, ptr: *mut c_void) -> c_int;
    pub fn mp_dev_upload(ctx: *mut MpCtx, dev_dst: *mut c_void, host_src: *const c_void, bytes: usize) -> c_int;
    pub fn mp_dev_download(ctx: *mut MpCtx, host_dst: *mut c_void, dev_src: *const c_void, bytes: usize) -> c_int;
    pub fn mp_host_alloc_pinned(bytes: usize, out: *mut *mut c_void) -> c_int;
    pub fn mp_host_free_pinned(ptr: *mut c_void) -> c_int;

    // render() and RenderProgress (machinery.rs:20-178)
    pub fn mp_render(scene: *const MpFlatBvh, camera: *const MpCamera, settings: *const MpRenderSettings, devices: *const c_int, n_devices: c_int, started: MpStartedTileFn, finished: MpFinishedTileFn, user: *mut c_void, out: *mut *mut MpRenderProgress) -> c_int;
    pub fn mp_progress_snapshot(p: *mut MpRenderProgress, out: *mut MpProgressSnapshot) -> c_int;
    pub fn mp_progress_is_finished(p: *mut MpRenderProgress, out: *mut c_int) -> c_int;
    pub fn mp_progress_elapsed_ms(p: *mut MpRenderProgress, out: *mut f64) -> c_int;
    pub fn mp_progress_abort(p: *mut MpRenderProgress) -> c_int;
    pub fn mp_progress_wait(p: *mut MpRenderProgress) -> c_int;
    pub fn mp_progress_image(p: *mut MpRenderProgress, host_out: *mut u8, cap: usize) -> c_int;
    pub fn mp_progress_image_ptr(p: *mut MpRenderProgress, out: *mut *const u8, out_bytes: *mut usize) -> c_int;
    pub fn mp_progress_free(p: *mut MpRenderProgress) -> c_int;

    // image output
    pub fn mp_image_write_png(path: *const c_char, rgba: *const u8, width: u32, height: u32) -> c_int;
    pub fn mp_image_write_ppm(path: *const c_char, rgba: *const u8, width: u32, height: u32) -> c_int;

    // cross-process tile queue (the AtomicUsize of machinery.rs:43,205-209)
    pub fn mp_tileq_open(name: *const c_char, create: c_int, out: *mut *mut MpTileQueue) -> c_int;
    pub fn mp_tileq_reset(q: *mut MpTileQueue) -> c_int;
    pub fn mp_tileq_next(q: *mut MpTileQueue, batch: u64, out_first: *mut u64) -> c_int;
    pub fn mp_tileq_peek(q: *mut MpTileQueue, out_next: *mut u64) -> c_int;
    pub fn mp_tileq_barrier(q: *mut MpTileQueue, world: u32, timeout_ms: u32) -> c_int;
    pub fn mp_tileq_close(q: *mut MpTileQueue, unlink: c_int) -> c_int;

    // multi-GPU groups (machinery.rs:51-116,205-209 over the GPUs of one box; NCCL film reduce)
    pub fn mp_group_create(devices: *const c_int, n_devices: c_int, out: *mut *mut MpGroup) -> c_int;
    pub fn mp_group_unique_id(out: *mut u8) -> c_int;
    pub fn mp_group_create_rank(device: c_int, rank: c_int, world: c_int, id: *const u8, queue_name: *const c_char, out: *mut *mut MpGroup) -> c_int;
    pub fn mp_group_destroy(g: *mut MpGroup) -> c_int;
    pub fn mp_group_size(g: *mut MpGroup, out_local: *mut c_int, out_world: *mut c_int) -> c_int;
    pub fn mp_group_ctx(g: *mut MpGroup, local_index: c_int, out: *mut *mut MpCtx) -> c_int;
    pub fn mp_group_scene_upload(g: *mut MpGroup, bvh: *const MpFlatBvh) -> c_int;
    pub fn mp_group_camera_set(g: *mut MpGroup, sampler: *const MpCameraSampler) -> c_int;
    pub fn mp_group_set_trace_variant(g: *mut MpGroup, variant: u32) -> c_int;
    pub fn mp_group_film_create(g: *mut MpGroup, width: u32, height: u32) -> c_int;
    pub fn mp_group_render(g: *mut MpGroup, tiles: *const MpTile, n_tiles: u32, params: *const MpRenderParams, shard: u32, min_batch: u32) -> c_int;
    pub fn mp_group_reduce_film(g: *mut MpGroup) -> c_int;
    pub fn mp_group_resolve_rgba8(g: *mut MpGroup, spp_total: u32, host_out: *mut u8) -> c_int;
    pub fn mp_group_read_film(g: *mut MpGroup, local_index: c_int, host_out: *mut f32) -> c_int;
    pub fn mp_group_sync(g: *mut MpGroup) -> c_int;
    pub fn mp_group_last_ms(g: *mut MpGroup, out_ms: *mut f32) -> c_int;
    pub fn mp_group_member_stats(g: *mut MpGroup, local_index: c_int, out_batches: *mut u64, out_tiles: *mut u64, out_ms: *mut f32) -> c_int;
}

/// The library's thread-local message for the last failed call on this thread.
pub fn last_error() -> String {
    unsafe {
        let p = mp_last_error();
        if p.is_null() {
            String::new()
        } else {
            std::ffi::CStr::from_ptr(p).to_string_lossy().into_owned()
        }
    }
}

/// `rc == 0` or an error carrying the library's message — what the safe wrapper turns into `anyhow::Error`.
pub fn check(rc: c_int) -> Result<(), (c_int, String)> {
    if rc == MP_OK {
        Ok(())
    } else {
        Err((rc, last_error()))
    }
}

#[cfg(test)]
mod tests {
    use super::*;
    use std::mem::size_of;

    #[test]
    fn layouts_match_the_header() {
        assert_eq!(size_of::<MpBox>(), 32);
        assert_eq!(size_of::<MpInnerNode>(), 128);
        assert_eq!(size_of::<MpTrianglePacket>(), 144);
        assert_eq!(size_of::<MpTriShading>(), 16);
        assert_eq!(size_of::<MpVertexShading>(), 32);
        assert_eq!(size_of::<MpRay>(), 36);
        assert_eq!(size_of::<MpHit>(), 16);
        assert_eq!(size_of::<MpTile>(), 16);
        assert_eq!(size_of::<MpRenderParams>(), 32);
        assert_eq!(size_of::<MpRenderSettings>(), 48);
        assert_eq!(size_of::<MpCounters>(), 56);
    }

    // the reference's own link known-answers (triangle_bvh/mod.rs:189-236), through the ABI
    #[test]
    fn link_encoding() {
        let mut link = 0u32;
        assert_eq!(unsafe { mp_link_new_leaf(5, 3, &mut link) }, MP_OK);
        assert_eq!(link, (5 << 3) | 3);
        assert_eq!(unsafe { mp_link_new_leaf(536_870_911, 1, &mut link) }, MP_EINVAL);
        assert_eq!(unsafe { mp_link_new_inner(536_870_912, &mut link) }, MP_EINVAL);
        let (mut kind, mut index, mut count) = (0, 0u32, 0u32);
        assert_eq!(unsafe { mp_link_decode(MP_LINK_NULL, &mut kind, &mut index, &mut count) }, MP_OK);
        assert_eq!(kind, 0);
    }
}

// Device context and the C ABI around the kernels (include/minipath_b200.h).
// nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -prec-div=true -prec-sqrt=true -lineinfo
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "host_common.h"
#include "kernels.cuh"
#include "wavefront.cuh"
#include "octet.cuh"
#include "persist.cuh"
#include "pool.cuh"
#include "decode.cuh"
#include "probes.cuh"

using namespace mp;

#define CK(call)                                                                                                  \
    do {                                                                                                          \
        cudaError_t _e = (call);                                                                                  \
        if (_e != cudaSuccess) return fail(MP_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

struct MpCtx {
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    bool count_visits = false;
    uint64_t launches = 0;

    // scene
    void* d_scene_blob = nullptr;
    size_t scene_blob_cap = 0;
    mpd::Scene scene{};
    mpd::SceneCM scene_cm{};
    uint32_t trace_variant = MP_TRACE_THREAD;
    uint32_t layout = MP_LAYOUT_AUTO;     // requested node layout
    bool decoded_ok = false;              // the decoded arrays exist and passed the min <= max check
    bool decoded_small = false;           // decoded BVH <= 1 MB: it stays L1/L2-hot and many lanes share a node, so even the pool kernels prefer it
    uint32_t* d_flags = nullptr;
    int sm_count = 148;
    int octet_bps = 0;
    int persist_bps = 0;
    int pool_bps[2] = {0, 0};  // resident pool blocks per SM, [QUANTISED, DECODED] instantiation
    uint2* pool_scratch = nullptr;
    size_t pool_scratch_cap = 0;
    uint32_t* d_claim = nullptr;
    size_t octet_bps_smem = 0;
    bool has_scene = false;
    uint32_t max_stack = 0;
    // camera
    mpd::Camera cam{};
    bool has_cam = false;
    // film
    float4* film = nullptr;
    bool film_owned = false;
    uint32_t width = 0, height = 0;
    uint32_t* d_rgba8 = nullptr;
    uint32_t* d_tile_out = nullptr;
    size_t tile_out_cap = 0;
    struct TileSlot {  // double-buffered tile read-back of mp_film_resolve_tiles_begin / _end
        uint32_t* d = nullptr;
        uint8_t* h = nullptr;  // pinned
        size_t cap = 0;
        cudaEvent_t done = nullptr;
    } slots[2];
    size_t tiles_resident = 0;  // length of the list stored by mp_tiles_upload (0: d_tiles holds a staged batch)
    // counters
    mpd::DevCounters* d_counters = nullptr;
    // tile staging
    MpTile* d_tiles = nullptr;
    size_t tiles_cap = 0;
    size_t tiles_staged = 0;
    MpTile* h_tiles_pinned = nullptr;
    // wavefront state
    mpd::WavefrontBuffers wf{};
};

static int use_device(MpCtx* ctx) {
    if (!ctx) return fail(MP_EINVAL, "null context");
    CK(cudaSetDevice(ctx->device));
    return MP_OK;
}

static uint32_t seed_hash(uint64_t seed) {
    auto h = [](uint32_t v) {
        uint32_t s = v * 747796405u + 2891336453u;
        uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
        return (w >> 22) ^ w;
    };
    return h((uint32_t)seed ^ h((uint32_t)(seed >> 32)));
}

static int stage_tiles(MpCtx* ctx, const MpTile* tiles, uint32_t n) {
    if (n > ctx->tiles_cap) {
        if (ctx->d_tiles) CK(cudaFree(ctx->d_tiles));
        if (ctx->h_tiles_pinned) CK(cudaFreeHost(ctx->h_tiles_pinned));
        ctx->d_tiles = nullptr, ctx->h_tiles_pinned = nullptr;
        size_t cap = std::max<size_t>(n, 1024);
        CK(cudaMalloc(&ctx->d_tiles, cap * sizeof(MpTile)));
        CK(cudaMallocHost(&ctx->h_tiles_pinned, cap * sizeof(MpTile)));
        ctx->tiles_cap = cap;
        ctx->tiles_staged = 0;
        ctx->tiles_resident = 0;
    } else {
        // same tile list as the previous call (a frame re-rendered, the next sample pass): already on the device
        if (n == ctx->tiles_staged && memcmp(ctx->h_tiles_pinned, tiles, n * sizeof(MpTile)) == 0) return MP_OK;
        // the pinned staging buffer may still be in flight from the previous call on this stream
        CK(cudaStreamSynchronize(ctx->stream));
    }
    ctx->tiles_staged = n;
    ctx->tiles_resident = 0;
    memcpy(ctx->h_tiles_pinned, tiles, n * sizeof(MpTile));
    CK(cudaMemcpyAsync(ctx->d_tiles, ctx->h_tiles_pinned, n * sizeof(MpTile), cudaMemcpyHostToDevice, ctx->stream));
    return MP_OK;
}

static mpd::TileMap make_tile_map(MpCtx* ctx, uint32_t n_tiles, uint32_t tile_size, size_t first = 0) {
    mpd::TileMap tm;
    tm.tiles = ctx->d_tiles + first;
    tm.n_tiles = n_tiles;
    tm.mt_x = (tile_size + 7) / 8;
    tm.threads_per_tile = tm.mt_x * ((tile_size + 3) / 4) * 32;
    return tm;
}

static int check_tiles(MpCtx* ctx, const MpTile* tiles, uint32_t n, uint32_t tile_size, bool need_film) {
    if (!tiles && n) return fail(MP_EINVAL, "null tiles");
    if (tile_size == 0) return fail(MP_EINVAL, "tile_size must be non-zero");
    for (uint32_t i = 0; i < n; ++i) {
        const MpTile& t = tiles[i];
        if (t.max_x < t.min_x || t.max_y < t.min_y || t.max_x - t.min_x > tile_size || t.max_y - t.min_y > tile_size)
            return fail(MP_EINVAL, "tile %u larger than tile_size %u", i, tile_size);
        if (need_film && (t.max_x > ctx->width || t.max_y > ctx->height)) return fail(MP_EINVAL, "tile %u outside the %ux%u film", i, ctx->width, ctx->height);
    }
    return MP_OK;
}

extern "C" {

int mp_device_count(int* out) {
    if (!out) return fail(MP_EINVAL, "null out");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *out = 0;
        return fail(MP_ENODEVICE, "no CUDA device: %s", cudaGetErrorString(e));
    }
    *out = n;
    return MP_OK;
}

int mp_ctx_create(int device, MpCtx** out) {
    if (!out) return fail(MP_EINVAL, "null out");
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return fail(MP_ENODEVICE, "no CUDA device available; this library has no CPU fallback");
    if (device < 0 || device >= n) return fail(MP_EINVAL, "device %d out of range (have %d)", device, n);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(MP_ENODEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    MpCtx* ctx = new MpCtx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    ctx->stream = ctx->own_stream;
    CK(cudaEventCreate(&ctx->ev0));
    CK(cudaEventCreate(&ctx->ev1));
    CK(cudaMalloc(&ctx->d_counters, sizeof(mpd::DevCounters)));
    CK(cudaMemset(ctx->d_counters, 0, sizeof(mpd::DevCounters)));
    *out = ctx;
    return MP_OK;
}

int mp_ctx_destroy(MpCtx* ctx) {
    if (!ctx) return MP_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    mpd::wavefront_free(ctx->wf);
    if (ctx->d_scene_blob) cudaFree(ctx->d_scene_blob);
    if (ctx->film && ctx->film_owned) cudaFree(ctx->film);
    if (ctx->d_rgba8) cudaFree(ctx->d_rgba8);
    if (ctx->d_tile_out) cudaFree(ctx->d_tile_out);
    if (ctx->pool_scratch) cudaFree(ctx->pool_scratch);
    if (ctx->d_claim) cudaFree(ctx->d_claim);
    if (ctx->d_flags) cudaFree(ctx->d_flags);
    for (auto& sl : ctx->slots) {
        if (sl.d) cudaFree(sl.d);
        if (sl.h) cudaFreeHost(sl.h);
        if (sl.done) cudaEventDestroy(sl.done);
    }
    if (ctx->d_counters) cudaFree(ctx->d_counters);
    if (ctx->d_tiles) cudaFree(ctx->d_tiles);
    if (ctx->h_tiles_pinned) cudaFreeHost(ctx->h_tiles_pinned);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaStreamDestroy(ctx->own_stream);
    delete ctx;
    return MP_OK;
}

// Process-wide pool of idle contexts, one list per device: mp_render() takes its workers' contexts from here so that
// repeated renders reuse streams, films, path-state buffers and pinned staging instead of re-allocating them
// (the reference keeps no such state, but its per-render cost is a thread spawn, not a 200 MB device allocation).
static std::mutex g_pool_mutex;
static std::vector<MpCtx*> g_pool;

int mp_ctx_pool_acquire(int device, MpCtx** out) {
    if (!out) return fail(MP_EINVAL, "null out");
    {
        std::lock_guard<std::mutex> g(g_pool_mutex);
        for (size_t i = 0; i < g_pool.size(); ++i)
            if (g_pool[i]->device == device) {
                *out = g_pool[i];
                g_pool.erase(g_pool.begin() + i);
                return MP_OK;
            }
    }
    return mp_ctx_create(device, out);
}

int mp_ctx_pool_release(MpCtx* ctx) {
    if (!ctx) return MP_OK;
    if (int rc = use_device(ctx)) return rc;
    cudaStreamSynchronize(ctx->stream);
    ctx->stream = ctx->own_stream;
    ctx->count_visits = false;
    std::lock_guard<std::mutex> g(g_pool_mutex);
    g_pool.push_back(ctx);
    return MP_OK;
}

int mp_ctx_pool_clear(void) {
    std::vector<MpCtx*> all;
    {
        std::lock_guard<std::mutex> g(g_pool_mutex);
        all.swap(g_pool);
    }
    for (MpCtx* c : all) mp_ctx_destroy(c);
    return MP_OK;
}

int mp_ctx_set_stream(MpCtx* ctx, void* cuda_stream) {
    if (int rc = use_device(ctx)) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return MP_OK;
}

int mp_ctx_set_trace_variant(MpCtx* ctx, uint32_t variant) {
    if (!ctx) return fail(MP_EINVAL, "null context");
    if (variant > MP_TRACE_HYBRID_POOL_S0) return fail(MP_EINVAL, "bad trace variant %u", variant);
    ctx->trace_variant = variant;
    return MP_OK;
}

int mp_ctx_set_node_layout(MpCtx* ctx, uint32_t layout) {
    if (!ctx) return fail(MP_EINVAL, "null context");
    if (layout > MP_LAYOUT_AUTO) return fail(MP_EINVAL, "bad node layout %u", layout);
    ctx->layout = layout;
    return MP_OK;
}

int mp_ctx_sync(MpCtx* ctx) {
    if (int rc = use_device(ctx)) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return MP_OK;
}

int mp_ctx_stream(MpCtx* ctx, void** out) {
    if (!ctx || !out) return fail(MP_EINVAL, "null argument");
    *out = (void*)ctx->stream;
    return MP_OK;
}

int mp_ctx_enable_counters(MpCtx* ctx, int on) {
    if (!ctx) return fail(MP_EINVAL, "null context");
    ctx->count_visits = on != 0;
    return MP_OK;
}

int mp_counters(MpCtx* ctx, MpCounters* out) {
    if (int rc = use_device(ctx)) return rc;
    if (!out) return fail(MP_EINVAL, "null out");
    mpd::DevCounters h;
    CK(cudaMemcpyAsync(&h, ctx->d_counters, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    out->rays = h.rays, out->camera_rays = h.camera_rays, out->shadow_rays = h.shadow_rays;
    out->inner_visits = h.inner_visits, out->packet_visits = h.packet_visits, out->leaf_visits = h.leaf_visits;
    out->launches = ctx->launches;
    return MP_OK;
}

int mp_counters_reset(MpCtx* ctx) {
    if (int rc = use_device(ctx)) return rc;
    CK(cudaMemsetAsync(ctx->d_counters, 0, sizeof(mpd::DevCounters), ctx->stream));
    ctx->launches = 0;
    return MP_OK;
}

int mp_last_device_ms(MpCtx* ctx, float* out_ms) {
    if (int rc = use_device(ctx)) return rc;
    if (!out_ms) return fail(MP_EINVAL, "null out");
    if (!ctx->timed) return fail(MP_ESTATE, "no render or trace call has been issued yet");
    CK(cudaEventSynchronize(ctx->ev1));
    CK(cudaEventElapsedTime(out_ms, ctx->ev0, ctx->ev1));
    return MP_OK;
}

int mp_scene_upload(MpCtx* ctx, const MpFlatBvh* b) {
    if (int rc = use_device(ctx)) return rc;
    if (!b || !b->packets || !b->packet_box || !b->tri_shading || !b->vertices || (b->n_inner && (!b->inner || !b->inner_box)))
        return fail(MP_EINVAL, "incomplete MpFlatBvh");
    if (b->n_packets == 0) return fail(MP_EINVAL, "scene has no triangle packets");
    if (b->n_materials && !b->materials) return fail(MP_EINVAL, "n_materials without a table");
    {  // every real triangle's material id must index the table (the kernels do not clamp)
        const uint32_t limit = std::max(b->n_materials, 1u);
        for (size_t i = 0, n = (size_t)b->n_packets * 8; i < n; ++i)
            if ((b->tri_shading[i].flags_material & MP_TRI_MATERIAL_MASK) >= limit)
                return fail(MP_EINVAL, "triangle %zu has material %u, the table has %u entries", i, b->tri_shading[i].flags_material & MP_TRI_MATERIAL_MASK, b->n_materials);
    }
    MpBvhStats st;
    if (int rc = mp_bvh_stats(b, &st)) return rc;
    if (st.max_stack > MPD_OCTET_STACK_MAX) return fail(MP_EINVAL, "BVH depth %u needs a %u-entry stack (limit %d)", st.max_depth, st.max_stack, MPD_OCTET_STACK_MAX);
    CK(cudaStreamSynchronize(ctx->stream));
    // one allocation, every array 256-byte aligned
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t s_inner = al((size_t)b->n_inner * sizeof(MpInnerNode)), s_ibox = al((size_t)b->n_inner * sizeof(MpBox));
    const size_t s_pack = al((size_t)b->n_packets * sizeof(MpTrianglePacket)), s_pbox = al((size_t)b->n_packets * sizeof(MpBox));
    const size_t s_sh = al((size_t)b->n_packets * 8 * sizeof(MpTriShading)), s_v = al((size_t)b->n_vertices * sizeof(MpVertexShading));
    const size_t s_inner_w = al((size_t)b->n_inner * MPD_NODE_W * sizeof(float4)), s_pack_w = al((size_t)b->n_packets * MPD_PACK_W * sizeof(float4));
    const size_t s_inner_r = al((size_t)b->n_inner * 160), s_pack_r = al((size_t)b->n_packets * 192);
    const size_t s_mat = al((size_t)b->n_materials * sizeof(MpMaterial));
    const size_t total = 2 * s_inner + s_ibox + 2 * s_pack + s_pbox + s_sh + s_v + s_inner_w + s_pack_w + s_inner_r + s_pack_r + s_mat + 256;
    if (total > ctx->scene_blob_cap) {  // pooled contexts keep the allocation between renders
        if (ctx->d_scene_blob) CK(cudaFree(ctx->d_scene_blob));
        ctx->d_scene_blob = nullptr, ctx->scene_blob_cap = 0, ctx->has_scene = false;
        CK(cudaMalloc(&ctx->d_scene_blob, total));
        ctx->scene_blob_cap = total;
    }
    char* p = (char*)ctx->d_scene_blob;
    auto up = [&](const void* src, size_t bytes, size_t slot) -> const void* {
        const void* dst = p;
        if (bytes) cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, ctx->stream);
        p += slot;
        return dst;
    };
    ctx->scene.inner = (const uint4*)up(b->inner, (size_t)b->n_inner * sizeof(MpInnerNode), s_inner);
    ctx->scene.inner_box = (const float4*)up(b->inner_box, (size_t)b->n_inner * sizeof(MpBox), s_ibox);
    ctx->scene.packets = (const uint4*)up(b->packets, (size_t)b->n_packets * sizeof(MpTrianglePacket), s_pack);
    ctx->scene.packet_box = (const float4*)up(b->packet_box, (size_t)b->n_packets * sizeof(MpBox), s_pbox);
    ctx->scene.tri_shading = (const uint4*)up(b->tri_shading, (size_t)b->n_packets * 8 * sizeof(MpTriShading), s_sh);
    ctx->scene.vertices = (const float4*)up(b->vertices, (size_t)b->n_vertices * sizeof(MpVertexShading), s_v);
    ctx->scene.materials = b->n_materials ? (const float4*)up(b->materials, (size_t)b->n_materials * sizeof(MpMaterial), s_mat) : nullptr;
    ctx->scene.n_materials = b->n_materials;
    // child-major copies for the octet kernels (octet.cuh), written on the device
    uint4* inner_cm = (uint4*)p;
    p += s_inner;
    uint4* packets_cm = (uint4*)p;
    p += s_pack;
    ctx->scene_cm.inner_cm = inner_cm, ctx->scene_cm.packets_cm = packets_cm;
    if (b->n_inner) {
        mpd::build_inner_cm_kernel<<<(unsigned)(((size_t)b->n_inner * 8 + 255) / 256), 256, 0, ctx->stream>>>(ctx->scene, b->n_inner, inner_cm);
        ctx->launches++;
    }
    mpd::build_packets_cm_kernel<<<(unsigned)(((size_t)b->n_packets * 9 + 255) / 256), 256, 0, ctx->stream>>>(ctx->scene, b->n_packets, packets_cm);
    ctx->launches++;
    // decoded layout (trace.cuh): written on the device from the arrays just uploaded
    float4* inner_w = (float4*)p;
    p += s_inner_w;
    float4* packets_w = (float4*)p;
    p += s_pack_w;
    ctx->scene.inner_w = inner_w, ctx->scene.packets_w = packets_w;
    if (!ctx->d_flags) CK(cudaMalloc(&ctx->d_flags, sizeof(uint32_t)));
    CK(cudaMemsetAsync(ctx->d_flags, 0, sizeof(uint32_t), ctx->stream));
    if (b->n_inner) {
        mpd::decode_inner_kernel<<<(unsigned)(((size_t)b->n_inner * 8 + 255) / 256), 256, 0, ctx->stream>>>(ctx->scene, b->n_inner, inner_w, ctx->d_flags);
        ctx->launches++;
    }
    mpd::decode_packets_kernel<<<(unsigned)(((size_t)b->n_packets * 8 + 255) / 256), 256, 0, ctx->stream>>>(ctx->scene, b->n_packets, packets_w);
    ctx->launches++;
    // quantised records for 256-bit gathers
    uint4* inner_rec = (uint4*)p;
    p += s_inner_r;
    uint4* packets_rec = (uint4*)p;
    p += s_pack_r;
    ctx->scene.inner_rec = inner_rec, ctx->scene.packets_rec = packets_rec;
    if (b->n_inner) {
        mpd::build_inner_records_kernel<<<(unsigned)(((size_t)b->n_inner * 10 + 255) / 256), 256, 0, ctx->stream>>>(ctx->scene, b->n_inner, inner_rec);
        ctx->launches++;
    }
    mpd::build_packet_records_kernel<<<(unsigned)(((size_t)b->n_packets * 12 + 255) / 256), 256, 0, ctx->stream>>>(ctx->scene, b->n_packets, packets_rec);
    ctx->launches++;
    uint32_t h_flags = 0;
    CK(cudaMemcpyAsync(&h_flags, ctx->d_flags, sizeof h_flags, cudaMemcpyDeviceToHost, ctx->stream));
    ctx->scene.root = b->root;
    {  // extension integrator epsilon: 1e-4 * |bbox size|, same unfused arithmetic as the oracle
        V3 s = v3(b->bbox_max[0], b->bbox_max[1], b->bbox_max[2]) - v3(b->bbox_min[0], b->bbox_min[1], b->bbox_min[2]);
        ctx->scene.eps = 1e-4f * std::sqrt(dot(s, s));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaGetLastError());
    ctx->decoded_ok = h_flags == 0;  // a child box with min > max: stay on the quantised records
    ctx->scene.minmax_ok = h_flags == 0 ? 1u : 0u;
    ctx->decoded_small = s_inner_w + s_pack_w <= (1u << 20);
    ctx->max_stack = st.max_stack;
    ctx->has_scene = true;
    return MP_OK;
}

int mp_camera_set(MpCtx* ctx, const MpCameraSampler* s) {
    if (!ctx || !s) return fail(MP_EINVAL, "null argument");
    ctx->cam.center = mpd::F3{s->center[0], s->center[1], s->center[2]};
    ctx->cam.up = mpd::F3{s->up[0], s->up[1], s->up[2]};
    ctx->cam.right = mpd::F3{s->right[0], s->right[1], s->right[2]};
    ctx->cam.film_origin_offset = mpd::F3{s->film_origin_offset[0], s->film_origin_offset[1], s->film_origin_offset[2]};
    ctx->cam.pixel_scale = s->pixel_scale;
    ctx->cam.lens_radius = s->lens_radius;
    ctx->cam.lens_weight = s->lens_weight;
    ctx->has_cam = true;
    return MP_OK;
}

static int film_release(MpCtx* ctx) {
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->film && ctx->film_owned) CK(cudaFree(ctx->film));
    if (ctx->d_rgba8) CK(cudaFree(ctx->d_rgba8));
    ctx->film = nullptr, ctx->d_rgba8 = nullptr;
    return MP_OK;
}

int mp_film_create(MpCtx* ctx, uint32_t width, uint32_t height) {
    if (int rc = use_device(ctx)) return rc;
    if (width == 0 || height == 0 || (uint64_t)width * height > 0x7fffffffull) return fail(MP_EINVAL, "bad film size %ux%u", width, height);
    if (ctx->film && ctx->film_owned && ctx->width == width && ctx->height == height) return mp_film_clear(ctx);  // pooled context: reuse
    if (int rc = film_release(ctx)) return rc;
    const size_t n = (size_t)width * height;
    CK(cudaMalloc(&ctx->film, n * sizeof(float4)));
    CK(cudaMalloc(&ctx->d_rgba8, n * sizeof(uint32_t)));
    CK(cudaMemsetAsync(ctx->film, 0, n * sizeof(float4), ctx->stream));
    ctx->film_owned = true;
    ctx->width = width, ctx->height = height;
    return MP_OK;
}

int mp_film_bind_external(MpCtx* ctx, void* device_ptr, uint32_t width, uint32_t height) {
    if (int rc = use_device(ctx)) return rc;
    if (!device_ptr || ((uintptr_t)device_ptr & 15)) return fail(MP_EINVAL, "film pointer must be a 16-byte aligned device pointer");
    if (width == 0 || height == 0 || (uint64_t)width * height > 0x7fffffffull) return fail(MP_EINVAL, "bad film size %ux%u", width, height);
    if (int rc = film_release(ctx)) return rc;
    ctx->film = (float4*)device_ptr;
    ctx->film_owned = false;
    CK(cudaMalloc(&ctx->d_rgba8, (size_t)width * height * sizeof(uint32_t)));
    ctx->width = width, ctx->height = height;
    return MP_OK;
}

int mp_film_clear(MpCtx* ctx) {
    if (int rc = use_device(ctx)) return rc;
    if (!ctx->film) return fail(MP_ESTATE, "no film");
    CK(cudaMemsetAsync(ctx->film, 0, (size_t)ctx->width * ctx->height * sizeof(float4), ctx->stream));
    return MP_OK;
}

int mp_film_device_ptr(MpCtx* ctx, void** out) {
    if (!ctx || !out) return fail(MP_EINVAL, "null argument");
    if (!ctx->film) return fail(MP_ESTATE, "no film");
    *out = ctx->film;
    return MP_OK;
}

}  // extern "C"

#include "context_render.inc"
#include "context_probes.inc"

// ---------------------------------------------------------------------------------------------------------------------------
// mp_trace_rays_device_noalloc — Object::intersect on device-resident rays WITHOUT any allocation in the call.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The pool traversal (pool.cuh) needs scratch: the per-ray stacks of every resident pool ([entry][slot] columns of 8-byte
// entries) and the launch's claim counter.  mp_trace_rays_device takes both from cudaMalloc'ed buffers sized for the device and
// the scene; this variant takes them from ONE file-scope array of MP_NOALLOC_SCRATCH_BYTES that every GPU loading the library
// holds until exit, and sizes its persistent grid to what fits (a block needs POOL_WARPS * POOL_W * stack_cap * 8 bytes).
// In host code `g_noalloc_scratch` is only a stand-in for the symbol: its device address is fetched with cudaGetSymbolAddress,
// once per device with that device current, and THAT pointer is what the kernel gets.
// Every call shares the array: the claim counter is cleared on the stream before each launch (stack entries are always written
// before they are read, so they need no clearing), and two calls that overlap on the GPU (different streams or contexts of one
// device) race on it — callers serialise them.
static __device__ __align__(256) unsigned char g_noalloc_scratch[MP_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mutex;
unsigned char* g_noalloc_ptr[kNoallocMaxDevices] = {};  // device address of g_noalloc_scratch per device, nullptr = not asked yet

// Device address of the static scratch on `device` (which must be current).
int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= kNoallocMaxDevices) return fail(MP_EINVAL, "device %d outside the static scratch table", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(MP_ECUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", device, cudaGetErrorString(e));
        g_noalloc_ptr[device] = (unsigned char*)p;
    }
    *out = g_noalloc_ptr[device];
    return MP_OK;
}
}  // namespace

extern "C" int mp_trace_rays_device_noalloc(MpCtx* ctx, const MpRay* dev_rays, uint64_t n, uint32_t order, MpHit* dev_out) {
    if (int rc = use_device(ctx)) return rc;
    if (!ctx->has_scene) return fail(MP_ESTATE, "no scene uploaded");
    if (order > MP_ORDER_NEAR_FIRST) return fail(MP_EINVAL, "bad order");
    if (n == 0) return MP_OK;
    if (!dev_rays || !dev_out) return fail(MP_EINVAL, "null buffer");
    if (n > MP_NOALLOC_MAX_RAYS) return fail(MP_EINVAL, "%llu rays in one call; the no-allocation entry point takes at most %llu", (unsigned long long)n, (unsigned long long)MP_NOALLOC_MAX_RAYS);
    const uint32_t cap = std::max<uint32_t>(8, (ctx->max_stack + 3u) & ~3u);
    if (cap > MP_NOALLOC_MAX_STACK)
        return fail(MP_EINVAL, "this BVH needs a %u-entry traversal stack per ray; the static scratch serves at most %u (use mp_trace_rays_device)", cap, (unsigned)MP_NOALLOC_MAX_STACK);
    unsigned char* base = nullptr;
    if (int rc = noalloc_scratch(ctx->device, &base)) return rc;
    uint32_t* claim = reinterpret_cast<uint32_t*>(base);            // first 256 bytes: the claim counter
    uint2* stacks = reinterpret_cast<uint2*>(base + 256);
    const size_t per_block = (size_t)POOL_WARPS * POOL_W * cap * sizeof(uint2);
    const unsigned long long fit = (MP_NOALLOC_SCRATCH_BYTES - 256) / per_block;  // >= 1 by the two limits above
    const bool dec = effective_layout(ctx, false) == MP_LAYOUT_DECODED;
    const unsigned long long resident = (unsigned long long)ctx->sm_count * (dec ? POOL_MIN_BLOCKS_DECODED : POOL_MIN_BLOCKS);
    const unsigned long long need = (n + POOL_WARPS * 32 - 1) / (POOL_WARPS * 32);
    const unsigned blocks = (unsigned)std::max<unsigned long long>(1, std::min(std::min(need, resident), fit));
    const size_t smem = (size_t)POOL_WARPS * POOL_SMEM_PER_WARP;
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    CK(cudaMemsetAsync(claim, 0, 256, ctx->stream));
    const float* r = (const float*)dev_rays;
    uint4* h = (uint4*)dev_out;
    MP_DISPATCH(ctx, false, order, mpd::trace_rays_pool_kernel<L, O, C><<<blocks, 128, smem, ctx->stream>>>(ctx->scene, r, (uint32_t)n, h, ctx->d_counters, claim, stacks, cap));
    ctx->launches++;
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->timed = true;
    CK(cudaGetLastError());
    return MP_OK;
}

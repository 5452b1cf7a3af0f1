// render(): the reference's tile scheduler (src/renderer/machinery.rs:20-210) with GPUs as the workers.
// One host thread per device pulls batches of tiles from one shared atomic counter (next_tile_index,
// machinery.rs:43,205-209), renders them with mp_render_tiles, resolves the batch to RGBA8 and copies it into the
// shared image under its mutex (machinery.rs:78-89), firing the reference's two callbacks from the worker thread.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <random>
#include <thread>
#include <vector>

#include "host_common.h"

using namespace mp;

// Pinned image buffers are expensive to create (cudaMallocHost of a 1080p image ~ 1 ms) and renders repeat: keep a few.
static std::mutex g_img_mutex;
static std::vector<std::pair<uint8_t*, size_t>> g_img_pool;
static uint8_t* image_acquire(size_t bytes) {
    uint8_t* ptr = nullptr;
    {
        std::lock_guard<std::mutex> g(g_img_mutex);
        for (size_t i = 0; i < g_img_pool.size(); ++i)
            if (g_img_pool[i].second == bytes) {
                ptr = g_img_pool[i].first;
                g_img_pool.erase(g_img_pool.begin() + i);
                break;
            }
    }
    if (!ptr && mp_host_alloc_pinned(bytes, (void**)&ptr) != MP_OK) return nullptr;
    memset(ptr, 0, bytes);  // RgbaImage::new (machinery.rs:34)
    return ptr;
}
static void image_release(uint8_t* ptr, size_t bytes) {
    if (!ptr) return;
    {
        std::lock_guard<std::mutex> g(g_img_mutex);
        if (g_img_pool.size() < 4) {
            g_img_pool.emplace_back(ptr, bytes);
            return;
        }
    }
    mp_host_free_pinned(ptr);
}

struct MpRenderProgress {
    ~MpRenderProgress() { image_release(image, image_bytes); }
    // RenderState (machinery.rs:191-203)
    MpRenderSettings settings{};
    MpCameraSampler sampler{};
    const MpFlatBvh* scene = nullptr;
    std::mutex image_mutex;
    uint8_t* image = nullptr;  // RGBA8, w*h*4, pinned (taken from / returned to the process-wide pool below)
    size_t image_bytes = 0;
    std::vector<MpTile> tile_ordering;
    std::atomic<size_t> next_tile_index{0};
    std::chrono::steady_clock::time_point start_time;
    std::mutex end_mutex;
    size_t workers_done = 0;
    bool has_end = false;
    std::chrono::steady_clock::duration end_elapsed{};
    // RenderProgress (machinery.rs:125-129)
    size_t worker_count = 0;
    size_t batch = 0;      // fixed tiles per batch (settings.tiles_per_batch), 0 = sized by the library per claim
    size_t min_batch = 4;  // library-sized batches: never below this many tiles
    std::atomic<size_t> tiles_done{0};  // tiles whose pixels are in `image`
    std::mutex join_mutex;              // wait() / free() may be called from several threads (RenderProgress is Sync)
    uint32_t trace_variant = MP_TRACE_THREAD;
    std::vector<std::thread> threads;
    std::atomic<size_t> threads_finished{0};
    MpStartedTileFn started = nullptr;
    MpFinishedTileFn finished = nullptr;
    void* user = nullptr;
    std::mutex error_mutex;
    std::string error;
    int error_code = MP_OK;
};

static void worker_fail(MpRenderProgress* p, int code) {
    std::lock_guard<std::mutex> g(p->error_mutex);
    if (p->error_code == MP_OK) {
        p->error_code = code;
        p->error = mp_last_error();
    }
    p->next_tile_index.store(p->tile_ordering.size(), std::memory_order_release);  // stop the other workers too
}

// One batch in flight per read-back slot: rendered and resolved on the device, waiting to be copied into the image.
struct Pending {
    size_t first = 0, n = 0;
    bool live = false;
};

// image.lock().copy_from(buffer.view(0,0,w,h), tile.min) (machinery.rs:78-89) + finished callbacks (:91-99)
static int finish_batch(MpRenderProgress* p, MpCtx* ctx, Pending& b, int slot) {
    if (!b.live) return MP_OK;
    b.live = false;
    const MpRenderSettings& s = p->settings;
    const uint8_t* buffer = nullptr;
    if (int rc = mp_film_resolve_tiles_end(ctx, slot, &buffer)) return rc;
    const MpTile* tiles = &p->tile_ordering[b.first];
    const size_t total = p->tile_ordering.size();
    {
        std::lock_guard<std::mutex> g(p->image_mutex);
        for (size_t i = 0; i < b.n; ++i) {
            const MpTile& t = tiles[i];
            const size_t w = t.max_x - t.min_x;
            for (uint32_t y = t.min_y; y < t.max_y; ++y)
                memcpy(&p->image[((size_t)y * s.resolution_x + t.min_x) * 4], &buffer[((i * s.tile_size + (y - t.min_y)) * s.tile_size) * 4], w * 4);
        }
    }
    // the reference estimates `finished` as next_tile_index - worker_count (machinery.rs:96); batches make that estimate too coarse,
    // so the tiles that have reached the image are counted exactly
    const size_t done = p->tiles_done.fetch_add(b.n, std::memory_order_acq_rel) + b.n;
    if (p->finished) {
        MpProgressSnapshot snap{std::min(done, total), total};
        for (size_t i = 0; i < b.n; ++i) p->finished(p->user, tiles[i], snap);
    }
    return MP_OK;
}

static void worker_main(MpRenderProgress* p, int device) {
    const MpRenderSettings& s = p->settings;
    const size_t total = p->tile_ordering.size();
    MpCtx* ctx = nullptr;
    const bool timing = getenv("MINIPATH_B200_TIMING") != nullptr;  // developer aid: where a render's host time goes
    auto t_mark = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[mp_render dev %d] %-14s %.3f ms\n", device, what, std::chrono::duration<double, std::milli>(now - t_mark).count());
        t_mark = now;
    };
    lap("thread start");
    int rc = mp_ctx_pool_acquire(device, &ctx);
    if (!rc) rc = mp_ctx_set_trace_variant(ctx, p->trace_variant);
    lap("ctx acquire");
    if (!rc) rc = mp_scene_upload(ctx, p->scene);
    lap("scene upload");
    if (!rc) rc = mp_camera_set(ctx, &p->sampler);  // camera.build_sampler(settings.resolution) (machinery.rs:65)
    if (!rc) rc = mp_film_create(ctx, s.resolution_x, s.resolution_y);
    if (!rc) rc = mp_tiles_upload(ctx, p->tile_ordering.data(), (uint32_t)total, s.tile_size);
    lap("film+tiles");
    MpRenderParams params{};
    params.sample_count = s.sample_count;
    params.sample_base = 0;
    params.max_bounces = s.max_bounces;
    params.mode = s.mode;
    params.order = s.order;
    params.tile_size = s.tile_size;
    params.seed = s.seed;
    if (rc) worker_fail(p, rc);

    Pending pending[2];
    int slot = 0;
    while (!rc) {
        // get_next_tile (machinery.rs:205-209), a batch at a time.  Library-sized batches: half of an equal share of what is left
        // (guided self-scheduling: coarse while work is plentiful so that every wavefront pass stays wide, fine at the tail so that
        // the devices finish together), never below min_batch, which for one device is "enough tiles for one full-width pass".
        size_t want = p->batch;
        if (want == 0) {
            const size_t cur = p->next_tile_index.load(std::memory_order_acquire);
            if (cur >= total) break;
            want = p->worker_count == 1 ? p->min_batch : std::max(p->min_batch, std::min<size_t>(64, (total - cur) / (2 * p->worker_count)));
        }
        const size_t first = p->next_tile_index.fetch_add(want, std::memory_order_acq_rel);
        if (first >= total) break;
        const size_t n = std::min(want, total - first);
        if (p->started)
            for (size_t i = 0; i < n; ++i) p->started(p->user, p->tile_ordering[first + i]);
        if (first == 0 && n == total) {
            // the whole frame in one batch (default for one device): resolve straight into the pinned image, no tile staging
            rc = mp_render_tile_range(ctx, 0, (uint32_t)n, &params);
            lap("launch");
            if (!rc) rc = mp_ctx_sync(ctx);  // wait for the render OUTSIDE the image lock: image() stays responsive meanwhile
            if (!rc) {
                std::lock_guard<std::mutex> g(p->image_mutex);
                rc = mp_film_resolve_rgba8(ctx, s.sample_count, p->image);
            }
            lap("render+resolve");
            p->tiles_done.fetch_add(n, std::memory_order_acq_rel);
            if (!rc && p->finished) {
                const MpProgressSnapshot snap{total, total};
                for (size_t i = 0; i < n; ++i) p->finished(p->user, p->tile_ordering[i], snap);
            }
            continue;
        }
        rc = finish_batch(p, ctx, pending[slot], slot);  // this slot's previous batch (two batches back) leaves first
        if (!rc) rc = mp_render_tile_range(ctx, (uint32_t)first, (uint32_t)n, &params);  // worker.render_tile (machinery.rs:77)
        if (!rc) rc = mp_film_resolve_tiles_begin(ctx, (uint32_t)first, (uint32_t)n, s.tile_size, s.sample_count, slot);
        if (rc) break;
        pending[slot] = Pending{first, n, true};
        slot ^= 1;
        rc = finish_batch(p, ctx, pending[slot], slot);  // while this batch renders, the previous one goes into the image
    }
    for (int k = 0; k < 2 && !rc; ++k) rc = finish_batch(p, ctx, pending[k ^ slot], k ^ slot);
    if (rc) worker_fail(p, rc);
    if (ctx) {
        if (rc) mp_ctx_destroy(ctx); else mp_ctx_pool_release(ctx);
    }
    lap("release");
    {
        auto elapsed = std::chrono::steady_clock::now() - p->start_time;  // machinery.rs:107-113
        std::lock_guard<std::mutex> g(p->end_mutex);
        p->workers_done += 1;
        if (p->workers_done == p->worker_count) {
            p->has_end = true;
            p->end_elapsed = elapsed;
        }
    }
    p->threads_finished.fetch_add(1, std::memory_order_release);
}

extern "C" {

int mp_render(const MpFlatBvh* scene, const MpCamera* camera, const MpRenderSettings* settings, const int* devices, int n_devices,
              MpStartedTileFn started, MpFinishedTileFn finished, void* user, MpRenderProgress** out) {
    if (!scene || !camera || !settings || !out) return fail(MP_EINVAL, "null argument");
    if (settings->tile_size == 0 || settings->sample_count == 0) return fail(MP_EINVAL, "tile_size and sample_count must be non-zero (NonZeroU32)");
    if (settings->resolution_x == 0 || settings->resolution_y == 0) return fail(MP_EINVAL, "empty resolution");
    if (settings->trace_kernel > MP_TRACE_HYBRID_POOL_S0 + 1) return fail(MP_EINVAL, "bad trace_kernel");
    if (settings->mode > MP_MODE_AUTO || settings->order > MP_ORDER_NEAR_FIRST) return fail(MP_EINVAL, "bad mode/order");
    int have = 0;
    if (int rc = mp_device_count(&have)) return rc;
    if (have == 0) return fail(MP_ENODEVICE, "no CUDA device available; this library has no CPU fallback");
    std::vector<int> devs;
    if (devices && n_devices > 0)
        devs.assign(devices, devices + n_devices);
    else
        for (int i = 0; i < have; ++i) devs.push_back(i);  // one worker per device, like one per core (machinery.rs:31-32)
    for (int d : devs)
        if (d < 0 || d >= have) return fail(MP_EINVAL, "device %d out of range (have %d)", d, have);

    MpRenderProgress* p = new MpRenderProgress();
    p->settings = *settings;
    if (p->settings.seed == 0) {  // SmallRng::from_os_rng (worker.rs:25)
        std::random_device rd;
        p->settings.seed = ((uint64_t)rd() << 32) | rd();
        if (p->settings.seed == 0) p->settings.seed = 1;
    }
    p->scene = scene;
    if (int rc = mp_camera_build_sampler(camera, settings->resolution_x, settings->resolution_y, &p->sampler)) {
        delete p;
        return rc;
    }
    p->image_bytes = (size_t)settings->resolution_x * settings->resolution_y * 4;
    p->image = image_acquire(p->image_bytes);
    if (!p->image) {
        delete p;
        return fail(MP_ENOMEM, "pinned image allocation failed");
    }
    size_t n_tiles = 0;
    mp_tile_ordering(0, 0, settings->resolution_x, settings->resolution_y, settings->tile_size, 0, nullptr, 0, &n_tiles);
    p->tile_ordering.resize(n_tiles);
    if (int rc = mp_tile_ordering(0, 0, settings->resolution_x, settings->resolution_y, settings->tile_size, 0, p->tile_ordering.data(),
                                  n_tiles, &n_tiles)) {
        delete p;
        return rc;
    }
    {  // traversal kernel: thread-per-ray unless the BVH needs a deeper stack than its local array holds
        MpBvhStats st;
        const bool have = mp_bvh_stats(scene, &st) == MP_OK;
        const bool deep = have && st.max_stack > 128;
        // deep trees: the pool kernel keeps its stacks in global scratch (any depth) and is the fast one; OCTET (3-4x the instructions) only serves
        // callers that ask for it.  The megakernel cannot run them (128-entry local stack), so a forced MP_MODE_MEGAKERNEL fails in launch_render.
        if (deep) p->trace_variant = MP_TRACE_POOL;
        if (p->settings.mode == MP_MODE_AUTO) {
            // measured: primary rays only → megakernel (2.6 vs 3.0 ms on the reference's bench workload); a scene that is one
            // leaf → megakernel (cube 13.1 vs 17.2 ms); otherwise the wavefront pipeline with pooled incoherent rays (teapot 32.0 vs 47.1 ms,
            // stand-in 1.98 vs 4.74 s per 256 spp)
            if (!deep && (p->settings.max_bounces == 0 || (have && st.n_inner == 0))) {
                p->settings.mode = MP_MODE_MEGAKERNEL;
            } else {
                p->settings.mode = MP_MODE_WAVEFRONT;
                p->trace_variant = deep ? MP_TRACE_POOL : MP_TRACE_HYBRID_POOL_S0;
            }
        }
        if (settings->trace_kernel) p->trace_variant = settings->trace_kernel - 1;
    }
    p->worker_count = devs.size();
    // Batches (tiles claimed per fetch_add).  tiles_per_batch set: exactly that.  Otherwise the library sizes them: a wavefront pass
    // carries about 32 M paths whatever the batch (context_render.inc: WF_TARGET_PATHS), so a batch of 32 M / (tile pixels * spp)
    // tiles is one full-width pass — batches of at least that size cost nothing and keep abort(), progress() and the tile
    // callbacks at pass granularity (a 4K 1024-spp frame is hundreds of batches, not one).  The megakernel (one launch per batch)
    // gets the same bound: the reference bench's 2048x1536x10 spp frame stays a single launch.
    {
        const size_t paths_per_tile = (size_t)settings->tile_size * settings->tile_size * settings->sample_count;
        const size_t pass_tiles = ((size_t)1 << 25) / std::max<size_t>(1, paths_per_tile);
        // several devices: the tail batches shrink to about a million paths each, so that the devices finish together
        p->min_batch = devs.size() == 1 ? std::max<size_t>(4, pass_tiles) : std::max<size_t>(1, (((size_t)1 << 20) + paths_per_tile - 1) / std::max<size_t>(1, paths_per_tile));
        p->batch = settings->tiles_per_batch;
    }
    p->started = started, p->finished = finished, p->user = user;
    p->start_time = std::chrono::steady_clock::now();
    try {
        for (int d : devs) p->threads.emplace_back(worker_main, p, d);
    } catch (const std::exception& e) {  // anyhow::Result from thread spawn (machinery.rs:116)
        p->next_tile_index.store(n_tiles);
        for (auto& t : p->threads) t.join();
        delete p;
        return fail(MP_ENOMEM, "thread spawn failed: %s", e.what());
    }
    *out = p;
    return MP_OK;
}

int mp_progress_snapshot(MpRenderProgress* p, MpProgressSnapshot* out) {  // machinery.rs:133-142
    if (!p || !out) return fail(MP_EINVAL, "null argument");
    const size_t total = p->tile_ordering.size();
    out->finished = std::min(p->tiles_done.load(std::memory_order_acquire), total);
    out->total = total;
    return MP_OK;
}
int mp_progress_is_finished(MpRenderProgress* p, int* out) {  // machinery.rs:144-146
    if (!p || !out) return fail(MP_EINVAL, "null argument");
    *out = p->threads_finished.load(std::memory_order_acquire) == p->worker_count;
    return MP_OK;
}
int mp_progress_elapsed_ms(MpRenderProgress* p, double* out) {  // machinery.rs:150-157
    if (!p || !out) return fail(MP_EINVAL, "null argument");
    std::lock_guard<std::mutex> g(p->end_mutex);
    auto d = p->has_end ? p->end_elapsed : (std::chrono::steady_clock::now() - p->start_time);
    *out = std::chrono::duration<double, std::milli>(d).count();
    return MP_OK;
}
int mp_progress_abort(MpRenderProgress* p) {  // machinery.rs:161-165
    if (!p) return fail(MP_EINVAL, "null argument");
    p->next_tile_index.store(p->tile_ordering.size(), std::memory_order_release);
    return MP_OK;
}
int mp_progress_wait(MpRenderProgress* p) {  // machinery.rs:169-173
    if (!p) return fail(MP_EINVAL, "null argument");
    {
        std::lock_guard<std::mutex> j(p->join_mutex);
        for (auto& t : p->threads)
            if (t.joinable()) t.join();
        p->threads.clear();
    }
    std::lock_guard<std::mutex> g(p->error_mutex);
    if (p->error_code != MP_OK) return fail(p->error_code, "render worker failed: %s", p->error.c_str());
    return MP_OK;
}
int mp_progress_image(MpRenderProgress* p, uint8_t* host_out, size_t cap) {  // machinery.rs:175-177
    if (!p || !host_out) return fail(MP_EINVAL, "null argument");
    std::lock_guard<std::mutex> g(p->image_mutex);
    if (cap < p->image_bytes) return fail(MP_EINVAL, "image buffer too small: %zu < %zu", cap, p->image_bytes);
    memcpy(host_out, p->image, p->image_bytes);
    return MP_OK;
}
int mp_progress_image_ptr(MpRenderProgress* p, const uint8_t** out, size_t* out_bytes) {  // the &RgbaImage behind the lock, no copy
    if (!p || !out) return fail(MP_EINVAL, "null argument");
    *out = p->image;
    if (out_bytes) *out_bytes = p->image_bytes;
    return MP_OK;
}
int mp_progress_free(MpRenderProgress* p) {
    if (!p) return MP_OK;
    mp_progress_abort(p);
    {
        std::lock_guard<std::mutex> j(p->join_mutex);
        for (auto& t : p->threads)
            if (t.joinable()) t.join();
        p->threads.clear();
    }
    delete p;
    return MP_OK;
}

}  // extern "C"

// Host-side camera (reference src/camera.rs), screen tiling (src/screen_block.rs) and the cross-process
// tile counter (the AtomicUsize of src/renderer/machinery.rs:43,205-209).  Compile with -ffp-contract=off.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

#include "host_common.h"

using namespace mp;

static inline V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }
static inline void st3(float* p, V3 v) { p[0] = v.x, p[1] = v.y, p[2] = v.z; }

extern "C" {

// Default camera: 35 mm (24 mm high sensor), 50 mm f/9, looks along -Z, focus at infinity (camera.rs:41-52).
int mp_camera_default(MpCamera* out) {
    if (!out) return fail(MP_EINVAL, "null camera");
    memset(out, 0, sizeof *out);
    st3(out->center, v3(0, 0, 0));
    st3(out->forward, v3(0, 0, -1));
    st3(out->up, v3(0, 1, 0));
    st3(out->right, v3(1, 0, 0));
    out->focus_distance = INFINITY;
    out->sensor_size = 24e-3f;
    out->sensor_is_width = 0;
    out->focal_length = 50e-3f;
    out->f_number = 9.0f;
    return MP_OK;
}

// nalgebra Isometry3::look_at_rh(eye, target, up).inverse(): rotation columns are
// x = normalize(up × z), y = normalize(z × x), z = normalize(eye − target); the frame is built directly from
// that matrix (the reference round-trips it through a unit quaternion, which can differ in the last ulp).
static void frame_from_direction(MpCamera* cam, V3 center, V3 dir, V3 up) {
    V3 z = normalize(-dir);
    V3 x = normalize(cross(up, z));
    V3 y = normalize(cross(z, x));
    st3(cam->center, center);
    st3(cam->forward, -z);
    st3(cam->up, y);
    st3(cam->right, x);
}

int mp_camera_look_at(MpCamera* cam, const float center[3], const float look_at[3], const float up[3]) {
    if (!cam || !center || !look_at || !up) return fail(MP_EINVAL, "null argument");
    V3 c = ld3(center), t = ld3(look_at);
    frame_from_direction(cam, c, t - c, ld3(up));
    cam->focus_distance = norm(t - c);  // camera.rs:98
    return MP_OK;
}

int mp_camera_look_direction(MpCamera* cam, const float center[3], const float forward[3], const float up[3]) {
    if (!cam || !center || !forward || !up) return fail(MP_EINVAL, "null argument");
    frame_from_direction(cam, ld3(center), ld3(forward), ld3(up));  // focus distance unchanged (camera.rs:112-115)
    return MP_OK;
}

int mp_camera_build_sampler(const MpCamera* cam, uint32_t res_x, uint32_t res_y, MpCameraSampler* out) {  // camera.rs:123-146
    if (!cam || !out) return fail(MP_EINVAL, "null argument");
    if (res_x == 0 || res_y == 0) return fail(MP_EINVAL, "empty resolution");
    float rx = (float)res_x, ry = (float)res_y;
    float pixel_scale = cam->sensor_is_width ? cam->sensor_size / rx : cam->sensor_size / ry;
    float uvx = ((rx - 1.0f) * pixel_scale) / 2.0f;
    float uvy = ((ry - 1.0f) * pixel_scale) / 2.0f;
    V3 forward = ld3(cam->forward), up = ld3(cam->up), right = ld3(cam->right);
    V3 off = ((-forward) * cam->focal_length + right * uvx) - up * uvy;
    st3(out->center, ld3(cam->center));
    st3(out->up, up);
    st3(out->right, right);
    st3(out->film_origin_offset, off);
    out->pixel_scale = pixel_scale;
    out->lens_radius = cam->focal_length / (2.0f * cam->f_number);
    out->lens_weight = cam->focal_length / cam->focus_distance;
    return MP_OK;
}

// divide_range (screen_block.rs:144-160)
static void divide_range(uint32_t start, uint32_t end, uint32_t tile_size, std::vector<std::pair<uint32_t, uint32_t>>& out) {
    uint32_t total = end - start;
    uint32_t full = total / tile_size;
    uint32_t n = full + (full * tile_size != total ? 1 : 0);
    for (uint32_t i = 0; i < n; ++i) {
        uint32_t s = start + i * tile_size;
        out.push_back({s, std::min(end, s + tile_size)});
    }
}

int mp_tile_ordering(uint32_t min_x, uint32_t min_y, uint32_t max_x, uint32_t max_y, uint32_t tile_size, uint64_t seed,
                     MpTile* out, size_t cap, size_t* out_n) {  // screen_block.rs:46-81
    if (!out_n) return fail(MP_EINVAL, "null out_n");
    if (tile_size == 0) return fail(MP_EINVAL, "tile_size must be non-zero (NonZeroU32)");
    *out_n = 0;
    if (!(min_x < max_x && min_y < max_y)) return MP_OK;  // is_empty (screen_block.rs:10-12)
    std::vector<std::pair<uint32_t, uint32_t>> xs, ys;
    divide_range(min_x, max_x, tile_size, xs);
    divide_range(min_y, max_y, tile_size, ys);
    size_t n = xs.size() * ys.size();
    *out_n = n;
    if (!out || cap < n) return out ? fail(MP_EINVAL, "tile buffer too small: %zu < %zu", cap, n) : MP_OK;

    float cx = (float)((min_x + max_x) / 2u), cy = (float)((min_y + max_y) / 2u);  // AABB::center on u32 (aabb.rs:114-120)
    float randomness_scale = std::sqrt(cx * cx + cy * cy) * 0.1f;
    std::mt19937_64 rng(seed ? seed : ((uint64_t)std::random_device{}() << 32 | std::random_device{}()));
    std::uniform_real_distribution<float> uni(0.0f, 1.0f);
    struct Keyed {
        float key;
        MpTile tile;
    };
    std::vector<Keyed> tiles;
    tiles.reserve(n);
    for (auto& y : ys)
        for (auto& x : xs) {
            MpTile t{x.first, y.first, x.second, y.second};
            float tx = cx - (float)((t.min_x + t.max_x) / 2u), ty = cy - (float)((t.min_y + t.max_y) / 2u);
            float noise = randomness_scale > 0.0f ? -std::log(1.0f - uni(rng)) * randomness_scale : 0.0f;  // Exp(1/scale)
            tiles.push_back(Keyed{std::sqrt(tx * tx + ty * ty) + noise, t});
        }
    std::stable_sort(tiles.begin(), tiles.end(), [](const Keyed& a, const Keyed& b) { return a.key < b.key; });
    for (size_t i = 0; i < n; ++i) out[i] = tiles[i].tile;
    return MP_OK;
}

// ---------------------------------------------------------------------------------------------------
// Cross-process tile counter in POSIX shared memory: every rank of one box does fetch_add(batch) on it,
// exactly like workers do on RenderState::next_tile_index (machinery.rs:205-209).
// ---------------------------------------------------------------------------------------------------
struct MpTileQueue {
    std::atomic<uint64_t>* counter;     // byte 0 of the 64-byte segment: next_tile_index
    std::atomic<uint32_t>* arrived;     // byte 8: ranks waiting at the barrier
    std::atomic<uint32_t>* generation;  // byte 12: barrier generation (sense reversal)
    char name[200];
};

int mp_tileq_open(const char* name, int create, MpTileQueue** out) {
    if (!name || !out) return fail(MP_EINVAL, "null argument");
    static_assert(sizeof(std::atomic<uint64_t>) == 8, "lock-free 64-bit atomic expected");
    int fd = shm_open(name, create ? (O_CREAT | O_RDWR) : O_RDWR, 0600);
    if (fd < 0) return fail(MP_EIO, "shm_open(%s) failed: %s", name, strerror(errno));
    if (create && ftruncate(fd, 64) != 0) {
        close(fd);
        return fail(MP_EIO, "ftruncate(%s) failed: %s", name, strerror(errno));
    }
    void* p = mmap(nullptr, 64, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return fail(MP_EIO, "mmap(%s) failed: %s", name, strerror(errno));
    MpTileQueue* q = new MpTileQueue();
    q->counter = reinterpret_cast<std::atomic<uint64_t>*>(p);
    q->arrived = reinterpret_cast<std::atomic<uint32_t>*>((char*)p + 8);
    q->generation = reinterpret_cast<std::atomic<uint32_t>*>((char*)p + 12);
    snprintf(q->name, sizeof q->name, "%s", name);
    if (create) {
        q->counter->store(0, std::memory_order_release);
        q->arrived->store(0, std::memory_order_release);
        q->generation->store(0, std::memory_order_release);
    }
    *out = q;
    return MP_OK;
}
int mp_tileq_reset(MpTileQueue* q) {
    if (!q) return fail(MP_EINVAL, "null queue");
    q->counter->store(0, std::memory_order_release);
    return MP_OK;
}
int mp_tileq_next(MpTileQueue* q, uint64_t batch, uint64_t* out_first) {
    if (!q || !out_first) return fail(MP_EINVAL, "null argument");
    *out_first = q->counter->fetch_add(batch, std::memory_order_acq_rel);
    return MP_OK;
}
int mp_tileq_peek(MpTileQueue* q, uint64_t* out_next) {
    if (!q || !out_next) return fail(MP_EINVAL, "null argument");
    *out_next = q->counter->load(std::memory_order_acquire);
    return MP_OK;
}
// Sense-reversing barrier over the ranks that share the segment: the last arrival bumps the generation, the others spin on it
// (yielding; a render step lasts milliseconds to seconds, a barrier microseconds).  timeout_ms bounds the wait for a rank that died.
int mp_tileq_barrier(MpTileQueue* q, uint32_t world, uint32_t timeout_ms) {
    if (!q || world == 0) return fail(MP_EINVAL, "bad argument");
    if (world == 1) return MP_OK;
    const uint32_t gen = q->generation->load(std::memory_order_acquire);
    if (q->arrived->fetch_add(1, std::memory_order_acq_rel) + 1 == world) {
        q->arrived->store(0, std::memory_order_release);
        q->generation->store(gen + 1, std::memory_order_release);
        return MP_OK;
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (uint32_t spins = 0; q->generation->load(std::memory_order_acquire) == gen; ++spins) {
        if ((spins & 63u) == 63u) {
            std::this_thread::yield();
            if (timeout_ms && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(timeout_ms))
                return fail(MP_ESTATE, "tile-queue barrier timed out after %u ms (%s): a rank is missing", timeout_ms, q->name);
        }
    }
    return MP_OK;
}
int mp_tileq_close(MpTileQueue* q, int unlink_it) {
    if (!q) return MP_OK;
    munmap((void*)q->counter, 64);
    if (unlink_it) shm_unlink(q->name);
    delete q;
    return MP_OK;
}

}  // extern "C"

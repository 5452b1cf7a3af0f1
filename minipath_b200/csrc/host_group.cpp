// mp_group_*: the reference's tile scheduler (src/renderer/machinery.rs:51-116,205-209) over the GPUs of one box, with the
// films combined by ONE NCCL reduce over NVLink — the multi-GPU half of the C ABI.
//
// Two shapes, same entry points:
//   * single process, n devices (mp_group_create): one host thread per device for the duration of a render call, one
//     std::atomic tile counter, ncclCommInitAll;
//   * one process per GPU (mp_group_create_rank, e.g. under torchrun): the counter lives in POSIX shared memory (MpTileQueue),
//     the communicator comes from ncclCommInitRank with an id made by rank 0 (mp_group_unique_id) and shipped by the caller.
// Workers pull BATCHES of consecutive tiles of the render's tile_ordering: batch = remaining / (2 * workers), clamped to
// [min_batch, 64] — large while there is plenty of work (every wavefront pass stays wide), fine at the tail (guided
// self-scheduling).  A worker claims its next batch only when its device has at most one batch queued behind the running one,
// otherwise a fast host thread would drain the counter while its GPU is still busy and nothing would balance.
// No collective runs during traversal: every device owns a zero-initialised full-resolution f32x4 film and writes only its
// tiles (or its samples), so one ncclReduce(sum) to rank 0 is the gather.
//
// NCCL is loaded with dlopen at group creation: the single-GPU library keeps no link-time dependency on it.
#include <cuda_runtime_api.h>
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "host_common.h"

using namespace mp;

namespace {

// ---- the part of nccl.h this file needs (NCCL 2.x ABI; /usr/include/nccl.h:37-38,146-181,260,285,351,493-500) ----
typedef void* NcclComm;
typedef struct { char internal[128]; } NcclUniqueId;
static_assert(sizeof(NcclUniqueId) == MP_NCCL_ID_BYTES, "ncclUniqueId is 128 bytes");
enum { kNcclSuccess = 0, kNcclFloat32 = 7, kNcclSum = 0 };
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*CommInitAll)(NcclComm*, int, const int*) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};
std::mutex g_nccl_mutex;
NcclApi g_nccl;

int load_nccl() {
    std::lock_guard<std::mutex> g(g_nccl_mutex);
    if (g_nccl.handle) return MP_OK;
    // an already loaded libnccl.so.2 (e.g. the one a host framework brought) is reused: same SONAME
    const char* names[] = {getenv("MINIPATH_B200_NCCL"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names)
        if (n && *n && (h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) return fail(MP_ENODEVICE, "NCCL not found (dlopen libnccl.so.2: %s); multi-GPU groups need it", dlerror());
    NcclApi a;
    a.handle = h;
#define MP_NCCL_SYM(field, sym)                                         \
    *(void**)(&a.field) = dlsym(h, sym);                                \
    if (!a.field) return fail(MP_ENODEVICE, "libnccl lacks %s", sym);
    MP_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    MP_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    MP_NCCL_SYM(CommInitAll, "ncclCommInitAll")
    MP_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    MP_NCCL_SYM(Reduce, "ncclReduce")
    MP_NCCL_SYM(GroupStart, "ncclGroupStart")
    MP_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    MP_NCCL_SYM(GetErrorString, "ncclGetErrorString")
    MP_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef MP_NCCL_SYM
    g_nccl = a;
    return MP_OK;
}

#define NCCL_CK(call)                                                                                        \
    do {                                                                                                     \
        int _r = (call);                                                                                     \
        if (_r != kNcclSuccess) return fail(MP_ECUDA, "%s failed: %s", #call, g_nccl.GetErrorString(_r));   \
    } while (0)
#define CU_CK(call)                                                                                          \
    do {                                                                                                     \
        cudaError_t _e = (call);                                                                             \
        if (_e != cudaSuccess) return fail(MP_ECUDA, "%s failed: %s", #call, cudaGetErrorString(_e));       \
    } while (0)

struct Member {  // one local device
    int device = 0;
    MpCtx* ctx = nullptr;
    NcclComm comm = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;  // device time of the last render (+ reduce)
    cudaEvent_t done[2] = {nullptr, nullptr};          // completion of the two most recent batches (back-pressure)
    uint64_t batches = 0, tiles = 0;                   // of the last render
};

}  // namespace

struct MpGroup {
    std::vector<Member> members;
    int rank_base = 0;  // global rank of members[0]
    int world = 1;
    MpTileQueue* queue = nullptr;  // multi-process groups
    std::atomic<uint64_t> next{0};  // single-process groups
    uint32_t width = 0, height = 0;
    bool timed = false;
};

namespace {

// Guided self-scheduling after a worker's first claim: half of an equal share of what is left, at least min_batch, at most 64 tiles.
uint64_t guided_batch(uint64_t remaining, uint32_t workers, uint32_t min_batch) {
    static const uint64_t div_pct = [] {  // share of an equal split a claim takes, as its inverse in percent (200 = half); experiment knob
        const char* e = getenv("MINIPATH_B200_GUIDED_DIV_PCT");
        const long v = e ? atol(e) : 200;
        return (uint64_t)(v >= 100 && v <= 1000 ? v : 200);
    }();
    const uint64_t b = remaining * 100ull / (div_pct * std::max(1u, workers));
    return std::max<uint64_t>(std::max(1u, min_batch), std::min<uint64_t>(64, b));
}

// A worker's FIRST claim of a frame: a fixed share of an equal split in one piece.  Small batches run narrower wavefront passes (measured on
// 8 GPUs, profiles/r2q_timeline_*: 3.47 ms per tile in 64-tile batches, 3.7 ms in the 8..31-tile batches of the guided sequence), while the tail
// only has to absorb the +-5 % by which equal tile counts differ in cost — so most of a worker's tiles come in one wide batch and the
// guided sequence shares out the rest.
uint64_t first_share(uint64_t total, uint32_t workers) {
    static const uint64_t pct = [] {
        const char* e = getenv("MINIPATH_B200_FIRST_SHARE_PCT");
        const long v = e ? atol(e) : 75;
        return (uint64_t)(v >= 0 && v <= 100 ? v : 75);
    }();
    return total * pct / (100ull * std::max(1u, workers));
}

int claim(MpGroup* g, uint64_t total, uint32_t min_batch, bool first_claim, uint64_t* first, uint64_t* n) {
    uint64_t cur = 0;
    if (g->queue) {
        if (int rc = mp_tileq_peek(g->queue, &cur)) return rc;
    } else {
        cur = g->next.load(std::memory_order_acquire);
    }
    if (cur >= total) {
        *first = total, *n = 0;
        return MP_OK;
    }
    uint64_t want = guided_batch(total - cur, (uint32_t)g->world, min_batch);
    if (first_claim) want = std::max(want, std::min(first_share(total, (uint32_t)g->world), total - cur));
    uint64_t f = 0;
    if (g->queue) {
        if (int rc = mp_tileq_next(g->queue, want, &f)) return rc;
    } else {
        f = g->next.fetch_add(want, std::memory_order_acq_rel);  // get_next_tile (machinery.rs:205-209), a batch at a time
    }
    *first = std::min(f, total);
    *n = f >= total ? 0 : std::min(want, total - f);
    return MP_OK;
}

// One device's share of a render call; runs on its own host thread when the group has several local devices.
int member_render(MpGroup* g, Member& m, const MpTile* tiles, uint32_t n_tiles, const MpRenderParams* p, uint32_t shard, uint32_t min_batch, int global_rank) {
    CU_CK(cudaSetDevice(m.device));
    void* stream = nullptr;
    if (int rc = mp_ctx_stream(m.ctx, &stream)) return rc;
    if (int rc = mp_tiles_upload(m.ctx, tiles, n_tiles, p->tile_size)) return rc;
    CU_CK(cudaEventRecord(m.ev_begin, (cudaStream_t)stream));
    m.batches = m.tiles = 0;
    if (shard == MP_SHARD_SPP) {
        // contiguous slice of the sample indices, remainder to the low ranks: distinct RNG counters per device
        const uint32_t q = p->sample_count / (uint32_t)g->world, r = p->sample_count % (uint32_t)g->world;
        const uint32_t count = q + ((uint32_t)global_rank < r ? 1u : 0u);
        const uint32_t base = (uint32_t)global_rank * q + std::min<uint32_t>((uint32_t)global_rank, r);
        if (count) {
            MpRenderParams mine = *p;
            mine.sample_base = p->sample_base + base, mine.sample_count = count;
            if (int rc = mp_render_tile_range(m.ctx, 0, n_tiles, &mine)) return rc;
            m.batches = 1, m.tiles = n_tiles;
        }
    } else {
        for (int k = 0;; k ^= 1) {
            if (m.batches >= 2) CU_CK(cudaEventSynchronize(m.done[k]));  // the batch before the previous one has left the device
            uint64_t first = 0, n = 0;
            if (int rc = claim(g, n_tiles, min_batch, m.batches == 0, &first, &n)) return rc;
            if (n == 0) break;
            if (int rc = mp_render_tile_range(m.ctx, (uint32_t)first, (uint32_t)n, p)) return rc;  // worker.render_tile (machinery.rs:77)
            CU_CK(cudaEventRecord(m.done[k], (cudaStream_t)stream));
            m.batches++, m.tiles += n;
        }
    }
    CU_CK(cudaEventRecord(m.ev_end, (cudaStream_t)stream));
    return MP_OK;
}

int make_member(int device, Member* out) {
    Member m;
    m.device = device;
    if (int rc = mp_ctx_create(device, &m.ctx)) return rc;
    CU_CK(cudaSetDevice(device));
    CU_CK(cudaEventCreate(&m.ev_begin));
    CU_CK(cudaEventCreate(&m.ev_end));
    for (auto& e : m.done) CU_CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    *out = m;
    return MP_OK;
}

int root_member(MpGroup* g, Member** out) {
    *out = g->rank_base == 0 ? &g->members[0] : nullptr;
    return MP_OK;
}

}  // namespace

extern "C" {

int mp_group_unique_id(uint8_t out[MP_NCCL_ID_BYTES]) {
    if (!out) return fail(MP_EINVAL, "null out");
    if (int rc = load_nccl()) return rc;
    NcclUniqueId id;
    NCCL_CK(g_nccl.GetUniqueId(&id));
    memcpy(out, id.internal, MP_NCCL_ID_BYTES);
    return MP_OK;
}

int mp_group_create(const int* devices, int n_devices, MpGroup** out) {
    if (!out) return fail(MP_EINVAL, "null out");
    int have = 0;
    if (int rc = mp_device_count(&have)) return rc;
    if (have == 0) return fail(MP_ENODEVICE, "no CUDA device available; this library has no CPU fallback");
    std::vector<int> devs;
    if (devices && n_devices > 0)
        devs.assign(devices, devices + n_devices);
    else
        for (int i = 0; i < have; ++i) devs.push_back(i);
    for (size_t i = 0; i < devs.size(); ++i) {
        if (devs[i] < 0 || devs[i] >= have) return fail(MP_EINVAL, "device %d out of range (have %d)", devs[i], have);
        for (size_t j = 0; j < i; ++j)
            if (devs[j] == devs[i]) return fail(MP_EINVAL, "device %d listed twice", devs[i]);
    }
    MpGroup* g = new MpGroup();
    g->world = (int)devs.size();
    for (int d : devs) {
        Member m;
        if (int rc = make_member(d, &m)) {
            mp_group_destroy(g);
            return rc;
        }
        g->members.push_back(m);
    }
    if (devs.size() > 1) {  // a one-device group needs no communicator
        int rc = load_nccl();
        if (!rc) {
            std::vector<NcclComm> comms(devs.size(), nullptr);
            const int r = g_nccl.CommInitAll(comms.data(), (int)devs.size(), devs.data());
            if (r != kNcclSuccess) rc = fail(MP_ECUDA, "ncclCommInitAll failed: %s", g_nccl.GetErrorString(r));
            for (size_t i = 0; i < devs.size() && !rc; ++i) g->members[i].comm = comms[i];
        }
        if (rc) {
            mp_group_destroy(g);
            return rc;
        }
    }
    *out = g;
    return MP_OK;
}

int mp_group_create_rank(int device, int rank, int world, const uint8_t id[MP_NCCL_ID_BYTES], const char* queue_name, MpGroup** out) {
    if (!out || world < 1 || rank < 0 || rank >= world) return fail(MP_EINVAL, "bad rank %d of %d", rank, world);
    if (world > 1 && (!id || !queue_name)) return fail(MP_EINVAL, "a multi-process group needs the NCCL id of rank 0 and a tile-queue name");
    MpGroup* g = new MpGroup();
    g->world = world, g->rank_base = rank;
    Member m;
    int rc = make_member(device, &m);
    if (!rc) g->members.push_back(m);
    if (!rc && world > 1) rc = load_nccl();
    // Rank 0 creates the shared tile counter BEFORE it enters ncclCommInitRank and the other ranks attach AFTER they have left it:
    // the communicator cannot come up before rank 0 has joined, so the segment exists by the time anyone opens it.
    if (!rc && world > 1 && rank == 0) rc = mp_tileq_open(queue_name, 1, &g->queue);
    if (!rc && world > 1) {
        NcclUniqueId uid;
        memcpy(uid.internal, id, MP_NCCL_ID_BYTES);
        cudaSetDevice(device);
        const int r = g_nccl.CommInitRank(&g->members[0].comm, world, uid, rank);
        if (r != kNcclSuccess) rc = fail(MP_ECUDA, "ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
    }
    if (!rc && world > 1 && rank != 0) rc = mp_tileq_open(queue_name, 0, &g->queue);
    if (rc) {
        mp_group_destroy(g);
        return rc;
    }
    *out = g;
    return MP_OK;
}

int mp_group_destroy(MpGroup* g) {
    if (!g) return MP_OK;
    for (Member& m : g->members) {
        cudaSetDevice(m.device);
        if (m.ctx) mp_ctx_sync(m.ctx);
        if (m.comm && g_nccl.CommDestroy) g_nccl.CommDestroy(m.comm);
        if (m.ev_begin) cudaEventDestroy(m.ev_begin);
        if (m.ev_end) cudaEventDestroy(m.ev_end);
        for (auto& e : m.done)
            if (e) cudaEventDestroy(e);
        if (m.ctx) mp_ctx_destroy(m.ctx);
    }
    if (g->queue) mp_tileq_close(g->queue, g->rank_base == 0);
    delete g;
    return MP_OK;
}

int mp_group_size(MpGroup* g, int* out_local, int* out_world) {
    if (!g) return fail(MP_EINVAL, "null group");
    if (out_local) *out_local = (int)g->members.size();
    if (out_world) *out_world = g->world;
    return MP_OK;
}

int mp_group_ctx(MpGroup* g, int local_index, MpCtx** out) {
    if (!g || !out || local_index < 0 || local_index >= (int)g->members.size()) return fail(MP_EINVAL, "bad group member %d", local_index);
    *out = g->members[local_index].ctx;
    return MP_OK;
}

int mp_group_scene_upload(MpGroup* g, const MpFlatBvh* bvh) {  // the scene is replicated on every device
    if (!g) return fail(MP_EINVAL, "null group");
    for (Member& m : g->members)
        if (int rc = mp_scene_upload(m.ctx, bvh)) return rc;
    return MP_OK;
}

int mp_group_camera_set(MpGroup* g, const MpCameraSampler* s) {
    if (!g) return fail(MP_EINVAL, "null group");
    for (Member& m : g->members)
        if (int rc = mp_camera_set(m.ctx, s)) return rc;
    return MP_OK;
}

int mp_group_set_trace_variant(MpGroup* g, uint32_t variant) {
    if (!g) return fail(MP_EINVAL, "null group");
    for (Member& m : g->members)
        if (int rc = mp_ctx_set_trace_variant(m.ctx, variant)) return rc;
    return MP_OK;
}

int mp_group_film_create(MpGroup* g, uint32_t width, uint32_t height) {
    if (!g) return fail(MP_EINVAL, "null group");
    for (Member& m : g->members)
        if (int rc = mp_film_create(m.ctx, width, height)) return rc;  // zero-initialised; an existing film of this size is cleared
    g->width = width, g->height = height;
    return MP_OK;
}

int mp_group_render(MpGroup* g, const MpTile* tiles, uint32_t n_tiles, const MpRenderParams* params, uint32_t shard, uint32_t min_batch) {
    if (!g || !params) return fail(MP_EINVAL, "null argument");
    if (shard > MP_SHARD_SPP) return fail(MP_EINVAL, "bad sharding scheme %u", shard);
    if (!g->width) return fail(MP_ESTATE, "no film (mp_group_film_create)");
    if (n_tiles == 0) return fail(MP_EINVAL, "empty tile list");
    if (min_batch == 0) {  // auto: the smallest batch still carries about a million paths (a 64x64 tile at 256 spp is one)
        const uint64_t per_tile = (uint64_t)std::max(1u, params->tile_size) * std::max(1u, params->tile_size) * std::max(1u, params->sample_count);
        min_batch = (uint32_t)std::max<uint64_t>(1, (((uint64_t)1 << 20) + per_tile - 1) / per_tile);
    }
    if (shard == MP_SHARD_TILES) {
        if (g->queue) {  // every rank has left the previous render's claim loop before rank 0 rewinds the counter
            if (int rc = mp_tileq_barrier(g->queue, (uint32_t)g->world, 120000)) return rc;
            if (g->rank_base == 0)
                if (int rc = mp_tileq_reset(g->queue)) return rc;
            if (int rc = mp_tileq_barrier(g->queue, (uint32_t)g->world, 120000)) return rc;
        } else {
            g->next.store(0, std::memory_order_release);
        }
    }
    std::vector<int> rcs(g->members.size(), MP_OK);
    std::vector<std::string> errs(g->members.size());
    auto run = [&](size_t i) {
        rcs[i] = member_render(g, g->members[i], tiles, n_tiles, params, shard, min_batch, g->rank_base + (int)i);
        if (rcs[i]) errs[i] = mp_last_error();  // thread-local message → carried back to the caller's thread
    };
    if (g->members.size() == 1) {
        run(0);
    } else {
        std::vector<std::thread> th;
        for (size_t i = 1; i < g->members.size(); ++i) th.emplace_back(run, i);
        run(0);
        for (auto& t : th) t.join();
    }
    for (size_t i = 0; i < rcs.size(); ++i)
        if (rcs[i]) return fail(rcs[i], "device %d: %s", g->members[i].device, errs[i].c_str());
    g->timed = true;
    return MP_OK;
}

int mp_group_reduce_film(MpGroup* g) {
    if (!g) return fail(MP_EINVAL, "null group");
    if (!g->width) return fail(MP_ESTATE, "no film");
    if (g->world == 1) return MP_OK;
    const size_t count = (size_t)g->width * g->height * 4;
    NCCL_CK(g_nccl.GroupStart());
    for (Member& m : g->members) {
        void *film = nullptr, *stream = nullptr;
        if (int rc = mp_film_device_ptr(m.ctx, &film)) return rc;
        if (int rc = mp_ctx_stream(m.ctx, &stream)) return rc;
        CU_CK(cudaSetDevice(m.device));
        // in place on the root; the other ranks' films are left as they were
        NCCL_CK(g_nccl.Reduce(film, film, count, kNcclFloat32, kNcclSum, 0, m.comm, (cudaStream_t)stream));
    }
    NCCL_CK(g_nccl.GroupEnd());
    for (Member& m : g->members) {
        void* stream = nullptr;
        if (int rc = mp_ctx_stream(m.ctx, &stream)) return rc;
        CU_CK(cudaSetDevice(m.device));
        CU_CK(cudaEventRecord(m.ev_end, (cudaStream_t)stream));
    }
    return MP_OK;
}

int mp_group_sync(MpGroup* g) {
    if (!g) return fail(MP_EINVAL, "null group");
    for (Member& m : g->members)
        if (int rc = mp_ctx_sync(m.ctx)) return rc;
    return MP_OK;
}

int mp_group_resolve_rgba8(MpGroup* g, uint32_t spp_total, uint8_t* host_out) {
    if (!g) return fail(MP_EINVAL, "null group");
    Member* root = nullptr;
    root_member(g, &root);
    for (Member& m : g->members)
        if (&m != root)
            if (int rc = mp_ctx_sync(m.ctx)) return rc;
    if (!root) return MP_OK;  // this process does not hold rank 0: its part of the frame is done once its stream has drained
    return mp_film_resolve_rgba8(root->ctx, spp_total, host_out);
}

int mp_group_read_film(MpGroup* g, int local_index, float* host_out) {
    if (!g || local_index < 0 || local_index >= (int)g->members.size()) return fail(MP_EINVAL, "bad group member %d", local_index);
    return mp_film_read_f32(g->members[local_index].ctx, host_out);
}

int mp_group_last_ms(MpGroup* g, float* out_ms) {
    if (!g || !out_ms) return fail(MP_EINVAL, "null argument");
    if (!g->timed) return fail(MP_ESTATE, "no group render has been issued yet");
    float worst = 0.0f;
    for (Member& m : g->members) {
        CU_CK(cudaSetDevice(m.device));
        CU_CK(cudaEventSynchronize(m.ev_end));
        float ms = 0.0f;
        CU_CK(cudaEventElapsedTime(&ms, m.ev_begin, m.ev_end));
        worst = std::max(worst, ms);
    }
    *out_ms = worst;
    return MP_OK;
}

int mp_group_member_stats(MpGroup* g, int local_index, uint64_t* out_batches, uint64_t* out_tiles, float* out_ms) {
    if (!g || local_index < 0 || local_index >= (int)g->members.size()) return fail(MP_EINVAL, "bad group member %d", local_index);
    Member& m = g->members[local_index];
    if (out_batches) *out_batches = m.batches;
    if (out_tiles) *out_tiles = m.tiles;
    if (out_ms) {
        if (!g->timed) return fail(MP_ESTATE, "no group render has been issued yet");
        CU_CK(cudaSetDevice(m.device));
        CU_CK(cudaEventSynchronize(m.ev_end));
        CU_CK(cudaEventElapsedTime(out_ms, m.ev_begin, m.ev_end));
    }
    return MP_OK;
}

}  // extern "C"

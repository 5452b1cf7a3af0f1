// Wavefront pipeline: ray-gen → [trace → shade (+ light sampling → shadow trace) → compact]* → film accumulate.
// The reference integrates one sample at a time inside Worker::render_tile (src/renderer/worker.rs:32-65); here one
// pass carries one sample of every pixel of the tile batch through the stages as separate kernels, with path state
// in HBM (SoA, 16-byte records) and queues of path slots compacted between bounces by warp ballot + block prefix sum.
// Per-path arithmetic and its order are exactly those of the megakernel (kernels.cuh), so both produce the same film.
#pragma once
#include "kernels.cuh"

namespace mpd {

#define WF_MAX_DEPTH 64
#define WF_STAGE_THREADS 512  // block size of the ray-gen and shade stages (one queue atomic per block)

struct WavefrontBuffers {
    size_t cap = 0;               // path slots
    float4* ray_o = nullptr;      // origin.xyz, throughput T
    float4* ray_d = nullptr;      // unit direction.xyz, radiance L
    uint4* hit = nullptr;         // tri, t, u, v
    uint4* meta = nullptr;        // pixel_index, rng state, depth | alpha << 31, rng increment
    uint32_t* queue[2] = {nullptr, nullptr};
    float4* sh_o = nullptr;       // shadow ray origin.xyz, contribution
    uint32_t* sh_path = nullptr;
    float4* rgb = nullptr;        // scenes with a material table only: T.g, T.b, L.g, L.b (ray_o.w / ray_d.w carry the red channel)
    float2* sh_c = nullptr;       //   "   : green and blue contribution of the shadow ray (sh_o.w carries red)
    uint32_t* counts = nullptr;   // 4 rows of WF_MAX_DEPTH+2: queue sizes per depth, shadow queue sizes, pool claim cursors (trace), (shadow)
};

inline void wavefront_free(WavefrontBuffers& w) {
    cudaFree(w.ray_o), cudaFree(w.ray_d), cudaFree(w.hit), cudaFree(w.meta);
    cudaFree(w.queue[0]), cudaFree(w.queue[1]), cudaFree(w.sh_o), cudaFree(w.sh_path), cudaFree(w.counts);
    cudaFree(w.rgb), cudaFree(w.sh_c);
    w = WavefrontBuffers{};
}

inline cudaError_t wavefront_reserve(WavefrontBuffers& w, size_t cap, bool rgb) {
    if (cap <= w.cap && (!rgb || w.rgb)) return cudaSuccess;
    cap = cap > w.cap ? cap : w.cap;
    wavefront_free(w);
    cudaError_t e;
#define WF_ALLOC(ptr, bytes) \
    if ((e = cudaMalloc(&(ptr), (bytes))) != cudaSuccess) return e;
    WF_ALLOC(w.ray_o, cap * sizeof(float4))
    WF_ALLOC(w.ray_d, cap * sizeof(float4))
    WF_ALLOC(w.hit, cap * sizeof(uint4))
    WF_ALLOC(w.meta, cap * sizeof(uint4))
    WF_ALLOC(w.queue[0], cap * sizeof(uint32_t))
    WF_ALLOC(w.queue[1], cap * sizeof(uint32_t))
    WF_ALLOC(w.sh_o, cap * sizeof(float4))
    WF_ALLOC(w.sh_path, cap * sizeof(uint32_t))
    WF_ALLOC(w.counts, 4 * (WF_MAX_DEPTH + 2) * sizeof(uint32_t))
    if (rgb) {
        WF_ALLOC(w.rgb, cap * sizeof(float4))
        WF_ALLOC(w.sh_c, cap * sizeof(float2))
    }
#undef WF_ALLOC
    w.cap = cap;
    return cudaSuccess;
}

// Stream compaction: every thread of the block calls this; threads with pred get consecutive slots of the queue.
// Warp ballot gives the intra-warp rank, a prefix sum over the warp totals in shared memory gives the intra-block
// offset, and one atomicAdd per block reserves the range.
__device__ __forceinline__ uint32_t block_compact_slot(bool pred, uint32_t* counter) {
    // any block size up to 1024 (a multiple of 32): the 512-thread stages issue a quarter of the single-address atomics of 128-thread blocks
    __shared__ uint32_t warp_base[32];
    __shared__ uint32_t block_base;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const uint32_t ballot = __ballot_sync(0xffffffffu, pred);
    if (lane == 0) warp_base[warp] = __popc(ballot);
    __syncthreads();
    if (warp == 0) {
        const uint32_t c = lane < n_warps ? warp_base[lane] : 0u;
        uint32_t incl = c;  // inclusive scan of the warp totals
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= (uint32_t)d) incl += v;
        }
        if (lane < n_warps) warp_base[lane] = incl - c;
        if (lane == 31) block_base = incl ? atomicAdd(counter, incl) : 0u;
    }
    __syncthreads();
    const uint32_t slot = block_base + warp_base[warp] + __popc(ballot & ((1u << lane) - 1u));
    __syncthreads();  // shared scratch is reused by the next call
    return slot;
}

// Barrier-free variant: one atomicAdd per warp that has anything to append.  Used where the threads of a block finish at very
// different times (a traversal precedes the compaction), so that early warps retire instead of waiting at a block barrier.
__device__ __forceinline__ uint32_t warp_compact_slot(bool pred, uint32_t* counter) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t ballot = __ballot_sync(0xffffffffu, pred);
    uint32_t base = 0u;
    if (ballot != 0u && lane == (uint32_t)(__ffs(ballot) - 1)) base = atomicAdd(counter, (uint32_t)__popc(ballot));
    base = __shfl_sync(0xffffffffu, base, ballot ? __ffs(ballot) - 1 : 0);
    return base + __popc(ballot & ((1u << lane) - 1u));
}

#ifndef WF_PRIMARY_WARP_COMPACT
#define WF_PRIMARY_WARP_COMPACT 1
#endif
// 80 registers / 24 warps per SM and no block barrier after the traversal: 34.1 -> 33.4 ms per default step (profiles/r1f_variants_primary_ab*.json)
#ifndef WF_PRIMARY_MIN_BLOCKS
#define WF_PRIMARY_MIN_BLOCKS 6
#endif
#ifndef WF_PRIMARY_S0_BLOCK
#define WF_PRIMARY_S0_BLOCK 0  // 1: SHADOW0 compacts the block's sun rays into its first lanes (two block barriers) instead of tracing them in the owning lanes
#endif

// Stage 1: CameraSampler::sample_ray (camera.rs:176-191) for `n_samples` consecutive samples of every pixel of the tile batch.
// Path slot = s * batch_threads + (thread index inside the batch): neighbouring slots are neighbouring pixels of one sample.
__global__ void __launch_bounds__(WF_STAGE_THREADS) wf_raygen_kernel(Camera cam, TileMap tm, uint32_t width, uint32_t sample_first, uint32_t n_samples,
                                                        unsigned long long batch_threads, uint32_t seed_hash, float4* __restrict__ ray_o,
                                                        float4* __restrict__ ray_d, uint4* __restrict__ meta, float4* __restrict__ rgb,
                                                        uint32_t* __restrict__ queue, uint32_t* count, DevCounters* dc) {
    // grid = (blocks of the batch, samples of the pass): blockIdx.y is the sample, no division to find it
    const uint32_t s = blockIdx.y, g = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long gid = (unsigned long long)s * batch_threads + g;
    uint32_t x = 0, y = 0, ti = 0;
    const bool active = s < n_samples && g < batch_threads && map_pixel32(tm, g, x, y, ti);
    ThreadCounts tc{};
    if (active) {
        const uint32_t pixel_index = y * width + x;
        Rng rng = rng_init(pixel_index, sample_first + s, seed_hash);
        const Ray r = sample_ray(cam, x, y, rng);
        ray_o[gid] = make_float4(r.o.x, r.o.y, r.o.z, 1.0f);
        ray_d[gid] = make_float4(r.d.x, r.d.y, r.d.z, 0.0f);
        meta[gid] = make_uint4(pixel_index, rng.state, 0u, rng.inc);
        if (rgb) rgb[gid] = make_float4(1.0f, 1.0f, 0.0f, 0.0f);
        tc.camera_rays = 1;
    }
    const uint32_t slot = block_compact_slot(active, count);
    if (active) queue[slot] = (uint32_t)gid;
    flush_counts(dc, tc, false);
}

__device__ __forceinline__ Ray load_ray(const float4 o, const float4 d) {
    Ray r;
    r.o = f3(o.x, o.y, o.z);
    r.d = f3(d.x, d.y, d.z);
    // same expression as Ray::new on the same unit direction (geometry/mod.rs:47)
    r.inv.x = r.d.x == 0.0f ? MPD_INF : __fdiv_rn(1.0f, r.d.x);
    r.inv.y = r.d.y == 0.0f ? MPD_INF : __fdiv_rn(1.0f, r.d.y);
    r.inv.z = r.d.z == 0.0f ? MPD_INF : __fdiv_rn(1.0f, r.d.z);
    return r;
}

// Stage 2: closest-hit traversal of the queued paths.
template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) wf_trace_kernel(Scene sc, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count,
                                                       const float4* __restrict__ ray_o, const float4* __restrict__ ray_d,
                                                       uint4* __restrict__ hit, DevCounters* dc) {
    const uint32_t n = *count;
    if (blockIdx.x * blockDim.x >= n) return;  // whole block idle
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    ThreadCounts tc{};
    if (i < n) {
        const uint32_t path = queue[i];
        const Ray r = load_ray(ray_o[path], ray_d[path]);
        tc.rays = 1;
        const Hit h = trace<LAYOUT, ORDER, COUNT>(sc, r, false, tc.vis);
        hit[path] = make_uint4(h.tri, __float_as_uint(h.t), __float_as_uint(h.u), __float_as_uint(h.v));
    }
    flush_counts(dc, tc, COUNT);
}

// An unoccluded light-sampling ray adds its contribution to the path's radiance (every shadow kernel ends here).
__device__ __forceinline__ void add_sun(float4* __restrict__ ray_d, float4* __restrict__ rgb, const float2* __restrict__ sh_c, uint32_t path, uint32_t i,
                                        float red) {
    ray_d[path].w = __fadd_rn(ray_d[path].w, red);
    if (rgb) {
        const float2 c = sh_c[i];
        float2* gb = reinterpret_cast<float2*>(rgb + path) + 1;  // L.g, L.b
        const float2 l = *gb;
        *gb = make_float2(__fadd_rn(l.x, c.x), __fadd_rn(l.y, c.y));
    }
}

// Stage 3: hit finalisation + shading + light/BSDF sampling; survivors are compacted into the next queue and
// shadow rays into the shadow queue.
template <bool RGB>
__global__ void __launch_bounds__(WF_STAGE_THREADS) wf_shade_kernel(Scene sc, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count,
                                                       uint32_t* __restrict__ next_queue, uint32_t* next_count,
                                                       float4* __restrict__ sh_o, uint32_t* __restrict__ sh_path, uint32_t* sh_count,
                                                       float4* __restrict__ ray_o, float4* __restrict__ ray_d, const uint4* __restrict__ hit,
                                                       uint4* __restrict__ meta, float4* __restrict__ rgb, float2* __restrict__ sh_c, uint32_t max_bounces) {
    const uint32_t n = *count;
    if (blockIdx.x * blockDim.x >= n) return;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool survive = false, shadow = false;
    uint32_t path = 0;
    float4 so = make_float4(0.f, 0.f, 0.f, 0.f);
    float2 sc2 = make_float2(0.f, 0.f);
    if (i < n) {
        path = queue[i];
        const float4 o4 = ray_o[path];
        float4 d4 = ray_d[path];
        const uint4 h4 = hit[path];
        uint4 m = meta[path];
        const Hit h{__uint_as_float(h4.y), __uint_as_float(h4.z), __uint_as_float(h4.w), h4.x};
        const F3 ro = f3(o4.x, o4.y, o4.z), rd = f3(d4.x, d4.y, d4.z);
        float T = o4.w, L = d4.w;  // the red channel, or all three of a scene without materials
        float4 e = make_float4(0.f, 0.f, 0.f, 0.f);  // T.g, T.b, L.g, L.b
        if (RGB) e = rgb[path];
        const uint32_t depth = m.z & 0x7fffffffu;
        if (max_bounces == 0) {  // the reference shader (worker.rs:57-64)
            if (h.tri != MP_NO_HIT) {
                L = fabsf(dotn(rd, shading_normal(sc, h)));
                m.z |= 0x80000000u;
            }
            ray_d[path].w = L;
            if (RGB) rgb[path] = make_float4(e.x, e.y, L, L);
            meta[path].z = m.z;
        } else if (h.tri == MP_NO_HIT) {
            ray_d[path].w = __fadd_rn(L, __fmul_rn(T, EXT_ENV));
            if (RGB) rgb[path] = make_float4(e.x, e.y, __fadd_rn(e.z, __fmul_rn(e.x, EXT_ENV)), __fadd_rn(e.w, __fmul_rn(e.y, EXT_ENV)));
        } else {
            if (depth == 0) m.z |= 0x80000000u;
            if (depth != max_bounces) {
                F3 nrm = shading_normal(sc, h);
                if (dotn(rd, nrm) > 0.0f) nrm = neg(nrm);
                const F3 origin = (ro + rd * h.t) + nrm * sc.eps;
                if (RGB) {
                    const F3 albedo = surface_albedo(sc, h);
                    T = __fmul_rn(T, albedo.x), e.x = __fmul_rn(e.x, albedo.y), e.y = __fmul_rn(e.y, albedo.z);
                } else {
                    T = __fmul_rn(T, EXT_ALBEDO);
                }
                const float c = dotn(nrm, f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
                if (c > 0.0f) {
                    shadow = true;
                    so = make_float4(origin.x, origin.y, origin.z, __fmul_rn(__fmul_rn(T, EXT_SUN_E), c));
                    if (RGB) sc2 = make_float2(__fmul_rn(__fmul_rn(e.x, EXT_SUN_E), c), __fmul_rn(__fmul_rn(e.y, EXT_SUN_E), c));
                }
                survive = true;
                Rng rng{m.y, m.w};
                if (depth >= EXT_RR_START) {
                    const float q = u01(rng);
                    const float ps = RGB ? fmaxf(fmaxf(T, e.x), e.y) : T;
                    if (q >= ps) {
                        survive = false;
                    } else {
                        T = __fdiv_rn(T, ps);
                        if (RGB) e.x = __fdiv_rn(e.x, ps), e.y = __fdiv_rn(e.y, ps);
                    }
                }
                if (survive) {
                    const Ray nr = ray_new(origin, bounce_direction(nrm, rng));
                    m.y = rng.state;
                    ray_o[path] = make_float4(nr.o.x, nr.o.y, nr.o.z, T);
                    ray_d[path] = make_float4(nr.d.x, nr.d.y, nr.d.z, L);
                    if (RGB) rgb[path] = e;
                    m.z = (m.z & 0x80000000u) | (depth + 1u);
                }
            }
            meta[path] = m;
        }
    }
    const uint32_t sslot = block_compact_slot(shadow, sh_count);
    if (shadow) {
        sh_o[sslot] = so;
        sh_path[sslot] = path;
        if (RGB) sh_c[sslot] = sc2;
    }
    const uint32_t slot = block_compact_slot(survive, next_count);
    if (survive) next_queue[slot] = path;
}

// Stage 3b: any-hit traversal of the light-sampling rays; unoccluded ones add their contribution to the path.
template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) wf_shadow_kernel(Scene sc, const float4* __restrict__ sh_o, const uint32_t* __restrict__ sh_path,
                                                        const uint32_t* __restrict__ sh_count, float4* __restrict__ ray_d, float4* __restrict__ rgb,
                                                        const float2* __restrict__ sh_c, DevCounters* dc) {
    const uint32_t n = *sh_count;
    if (blockIdx.x * blockDim.x >= n) return;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    ThreadCounts tc{};
    if (i < n) {
        const float4 o = sh_o[i];
        const Ray r = ray_new(f3(o.x, o.y, o.z), f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
        tc.rays = 1, tc.shadow_rays = 1;
        const Hit h = trace<LAYOUT, ORDER, COUNT>(sc, r, true, tc.vis);
        if (h.tri == MP_NO_HIT) add_sun(ray_d, rgb, sh_c, sh_path[i], i, o.w);
    }
    flush_counts(dc, tc, COUNT);
}

// Stages 1-3 of depth 0 in ONE kernel: camera ray → closest hit (thread per ray: camera rays are coherent) → shade, with the path
// in registers.  As separate stages a camera path crosses HBM four times (ray written, read by the trace, hit written, everything
// read again by the shade) although most camera rays of a sparse scene end right there (three quarters of the teapot's miss); fused,
// a path writes its direction + radiance and its meta word once, and only the survivors write an origin and queue entries.
// The memory left behind is exactly what wf_raygen + wf_trace + wf_shade at depth 0 leave (minus the hit record, which only the
// shade stage of the same depth reads), so the later stages are unchanged.
// SHADOW0: the depth-0 sun ray (wf_shadow_kernel of depth 0) is traced here too instead of being queued: the shadow rays of one warp
// start on neighbouring pixels' hit points and share ONE direction, so the thread-per-ray traversal stays coherent; the contribution
// is added to L before anything else is (the order wf_shadow_kernel of depth 0 adds it in), so the film does not change.
// RGB: the scene has a material table (wf_shade_kernel<true>); the green and blue channels go to `rgb` / `sh_c`.
template <int LAYOUT, int ORDER, bool COUNT, bool SHADOW0, bool RGB>
__global__ void __launch_bounds__(128, WF_PRIMARY_MIN_BLOCKS) wf_primary_kernel(Scene sc, Camera cam, TileMap tm, uint32_t width, uint32_t sample_first, uint32_t n_samples,
                                                         unsigned long long batch_threads, uint32_t seed_hash, uint32_t max_bounces,
                                                         float4* __restrict__ ray_o, float4* __restrict__ ray_d, uint4* __restrict__ meta,
                                                         uint32_t* __restrict__ next_queue, uint32_t* next_count, float4* __restrict__ sh_o,
                                                         uint32_t* __restrict__ sh_path, uint32_t* sh_count, float4* __restrict__ rgb,
                                                         float2* __restrict__ sh_c, DevCounters* dc) {
    const uint32_t s = blockIdx.y, g = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long gid = (unsigned long long)s * batch_threads + g;
    uint32_t x = 0, y = 0, ti = 0;
    const bool active = s < n_samples && g < batch_threads && map_pixel32(tm, g, x, y, ti);
    ThreadCounts tc{};
    bool survive = false, shadow = false;
    float4 so = make_float4(0.f, 0.f, 0.f, 0.f);
    float2 sc2 = make_float2(0.f, 0.f);
    uint4 m = make_uint4(0u, 0u, 0u, 0u);
    float4 out_d = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 e = make_float4(1.0f, 1.0f, 0.0f, 0.0f);  // T.g, T.b, L.g, L.b
    if (active) {
        const uint32_t pixel_index = y * width + x;
        Rng rng = rng_init(pixel_index, sample_first + s, seed_hash);
        const Ray r = sample_ray(cam, x, y, rng);
        tc.camera_rays = 1, tc.rays = 1;
        const Hit h = trace<LAYOUT, ORDER, COUNT>(sc, r, false, tc.vis);
        // ---- wf_shade_kernel at depth 0 (T = 1, L = 0) ----
        float T = 1.0f, L = 0.0f;
        m = make_uint4(pixel_index, rng.state, 0u, rng.inc);
        out_d = make_float4(r.d.x, r.d.y, r.d.z, 0.0f);
        if (max_bounces == 0) {  // the reference shader (worker.rs:57-64)
            if (h.tri != MP_NO_HIT) {
                L = fabsf(dotn(r.d, shading_normal(sc, h)));
                m.z |= 0x80000000u;
            }
            out_d.w = L;
            if (RGB) e.z = L, e.w = L;
        } else if (h.tri == MP_NO_HIT) {
            out_d.w = __fadd_rn(L, __fmul_rn(T, EXT_ENV));
            if (RGB) e.z = __fadd_rn(e.z, __fmul_rn(e.x, EXT_ENV)), e.w = __fadd_rn(e.w, __fmul_rn(e.y, EXT_ENV));
        } else {
            m.z |= 0x80000000u;  // depth 0: alpha = primary hit
            if (0u != max_bounces) {
                F3 nrm = shading_normal(sc, h);
                if (dotn(r.d, nrm) > 0.0f) nrm = neg(nrm);
                const F3 origin = (r.o + r.d * h.t) + nrm * sc.eps;
                if (RGB) {
                    const F3 albedo = surface_albedo(sc, h);
                    T = __fmul_rn(T, albedo.x), e.x = __fmul_rn(e.x, albedo.y), e.y = __fmul_rn(e.y, albedo.z);
                } else {
                    T = __fmul_rn(T, EXT_ALBEDO);
                }
                const float c = dotn(nrm, f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
                if (c > 0.0f) {
                    shadow = true;
                    so = make_float4(origin.x, origin.y, origin.z, __fmul_rn(__fmul_rn(T, EXT_SUN_E), c));
                    if (RGB) sc2 = make_float2(__fmul_rn(__fmul_rn(e.x, EXT_SUN_E), c), __fmul_rn(__fmul_rn(e.y, EXT_SUN_E), c));
                }
                survive = true;  // Russian roulette starts at depth EXT_RR_START >= 1
                static_assert(EXT_RR_START >= 1u, "depth 0 never plays Russian roulette");
                {
                    const Ray nr = ray_new(origin, bounce_direction(nrm, rng));
                    m.y = rng.state;
                    ray_o[gid] = make_float4(nr.o.x, nr.o.y, nr.o.z, T);
                    out_d = make_float4(nr.d.x, nr.d.y, nr.d.z, L);
                    m.z = (m.z & 0x80000000u) | 1u;
                }
            }
        }
#if !WF_PRIMARY_S0_BLOCK
        if (SHADOW0 && shadow) {  // ---- wf_shadow_kernel at depth 0 ----
            const Ray sr = ray_new(f3(so.x, so.y, so.z), f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
            tc.rays += 1, tc.shadow_rays += 1;
            const Hit sh = trace<LAYOUT, ORDER, COUNT>(sc, sr, true, tc.vis);
            if (sh.tri == MP_NO_HIT) {
                out_d.w = __fadd_rn(out_d.w, so.w);
                if (RGB) e.z = __fadd_rn(e.z, sc2.x), e.w = __fadd_rn(e.w, sc2.y);
            }
            shadow = false;
        }
#endif
    }
#if WF_PRIMARY_S0_BLOCK
    if (SHADOW0) {  // the block's sun rays compacted into its first lanes: full warps trace them instead of the sparse lanes that own them
        __shared__ float4 s_so[128];
        __shared__ unsigned char s_lit[128];
        __shared__ uint32_t s_n;
        if (threadIdx.x == 0) s_n = 0u;
        __syncthreads();
        const uint32_t mine = warp_compact_slot(shadow, &s_n);
        if (shadow) s_so[mine] = so;
        __syncthreads();
        if (threadIdx.x < s_n) {
            const float4 o = s_so[threadIdx.x];
            const Ray sr = ray_new(f3(o.x, o.y, o.z), f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
            tc.rays += 1, tc.shadow_rays += 1;
            const Hit sh = trace<LAYOUT, ORDER, COUNT>(sc, sr, true, tc.vis);
            s_lit[threadIdx.x] = sh.tri == MP_NO_HIT ? 1 : 0;
        }
        __syncthreads();
        if (shadow && s_lit[mine]) {
            out_d.w = __fadd_rn(out_d.w, so.w);
            if (RGB) e.z = __fadd_rn(e.z, sc2.x), e.w = __fadd_rn(e.w, sc2.y);
        }
        shadow = false;
    }
#endif
    if (active) {
        ray_d[gid] = out_d;
        meta[gid] = m;
        if (RGB) rgb[gid] = e;
    }
#if WF_PRIMARY_WARP_COMPACT
    const uint32_t sslot = SHADOW0 ? 0u : warp_compact_slot(shadow, sh_count);
    const uint32_t slot = warp_compact_slot(survive, next_count);
#else
    const uint32_t sslot = SHADOW0 ? 0u : block_compact_slot(shadow, sh_count);
    const uint32_t slot = block_compact_slot(survive, next_count);
#endif
    if (!SHADOW0 && shadow) {
        sh_o[sslot] = so;
        sh_path[sslot] = (uint32_t)gid;
        if (RGB) sh_c[sslot] = sc2;
    }
    if (survive) next_queue[slot] = (uint32_t)gid;
    flush_counts(dc, tc, COUNT);
}

// Stage 4: film accumulation (worker.rs:42): film[pixel] += (L, L, L, alpha), the pass's samples in sample order.
__global__ void __launch_bounds__(128) wf_accumulate_kernel(TileMap tm, float4* __restrict__ film, const float4* __restrict__ ray_d,
                                                            const uint4* __restrict__ meta, const float4* __restrict__ rgb, uint32_t n_samples,
                                                            unsigned long long batch_threads) {
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x, y, ti;
    if (gid >= batch_threads || !map_pixel32(tm, gid, x, y, ti)) return;
    const uint32_t pixel = meta[gid].x;
    float4 f = film[pixel];
    for (uint32_t s = 0; s < n_samples; ++s) {
        const unsigned long long slot = s * batch_threads + gid;
        const float L = ray_d[slot].w;
        float Lg = L, Lb = L;
        if (rgb) {
            const float4 e = rgb[slot];
            Lg = e.z, Lb = e.w;
        }
        const float alpha = (meta[slot].z & 0x80000000u) ? 1.0f : 0.0f;
        f.x = __fadd_rn(f.x, L), f.y = __fadd_rn(f.y, Lg), f.z = __fadd_rn(f.z, Lb), f.w = __fadd_rn(f.w, alpha);
    }
    film[pixel] = f;
}

}  // namespace mpd

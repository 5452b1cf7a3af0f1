// Kernels of the trace path: primary-ray dump, caller-ray trace (gate), the per-pixel megakernel and film resolve.
// The wavefront pipeline lives in wavefront.cuh.
#pragma once
#include "trace.cuh"

namespace mpd {

// Extension integrator constants (max_bounces > 0; the reference defines no behaviour there — SURVEY A.7)
#define EXT_ALBEDO 0.7f
#define EXT_ENV 1.0f
#define EXT_SUN_E 1.0f
#define EXT_SUN_X 0.25916052f
#define EXT_SUN_Y 0.86386842f
#define EXT_SUN_Z 0.43193421f
#define EXT_RR_START 3u

// Diffuse reflectance at a hit of the extension integrator (oracle: surface_albedo).  No material table: the constant grey.
// Otherwise the entry of the triangle's material id; MP_MAT_CHECKER halves it on the odd cells of a grid over the texture
// coordinate interpolated like the reference's HitRecord::texture_coords (ray_bvh_intersection.rs:69-94; triangle.rs:229-244).
__device__ __forceinline__ F3 interpolate_tex(const Scene& sc, const uint4 sh, const Hit& h) {
    const float w = __fsub_rn(__fsub_rn(1.0f, h.u), h.v);  // triangle.rs:234
    const float4 t0 = __ldg(sc.vertices + 2 * (size_t)sh.x + 1), t1 = __ldg(sc.vertices + 2 * (size_t)sh.y + 1),
                 t2 = __ldg(sc.vertices + 2 * (size_t)sh.z + 1);
    return (f3(t0.x, t0.y, t0.z) * w + f3(t1.x, t1.y, t1.z) * h.u) + f3(t2.x, t2.y, t2.z) * h.v;
}
__device__ __forceinline__ F3 surface_albedo(const Scene& sc, const Hit& h) {
    if (sc.n_materials == 0u) return f3(EXT_ALBEDO, EXT_ALBEDO, EXT_ALBEDO);
    const uint4 sh = __ldg(sc.tri_shading + h.tri);
    const uint32_t id = sh.w & MP_TRI_MATERIAL_MASK;
    const float4 m0 = __ldg(sc.materials + 2 * (size_t)id), m1 = __ldg(sc.materials + 2 * (size_t)id + 1);
    F3 a = f3(m0.x, m0.y, m0.z);
    if (__float_as_uint(m0.w) & MP_MAT_CHECKER) {
        const F3 tex = interpolate_tex(sc, sh, h);
        const float cell = __fadd_rn(floorf(__fmul_rn(tex.x, m1.x)), floorf(__fmul_rn(tex.y, m1.x)));
        const float half = __fmul_rn(cell, 0.5f);
        if (floorf(half) != half) a = a * 0.5f;  // odd cell
    }
    return a;
}
__device__ __forceinline__ float max3(const F3& v) { return fmaxf(fmaxf(v.x, v.y), v.z); }

struct DevCounters {  // mirrors MpCounters (launches is host-side)
    unsigned long long rays, camera_rays, shadow_rays, inner_visits, packet_visits, leaf_visits;
};

struct ThreadCounts {
    uint32_t rays, camera_rays, shadow_rays;
    Visits vis;
};

__device__ __forceinline__ void flush_counts(DevCounters* dc, const ThreadCounts& c, bool count_visits) {
    // every thread of the warp reaches this point (inactive threads carry zeros)
    uint32_t rays = __reduce_add_sync(0xffffffffu, c.rays);
    uint32_t cam = __reduce_add_sync(0xffffffffu, c.camera_rays);
    uint32_t sh = __reduce_add_sync(0xffffffffu, c.shadow_rays);
    uint32_t vi = 0, vp = 0, vl = 0;
    if (count_visits) {
        vi = __reduce_add_sync(0xffffffffu, c.vis.inner);
        vp = __reduce_add_sync(0xffffffffu, c.vis.packets);
        vl = __reduce_add_sync(0xffffffffu, c.vis.leaves);
    }
    if ((threadIdx.x & 31) == 0) {
        if (rays) atomicAdd(&dc->rays, (unsigned long long)rays);
        if (cam) atomicAdd(&dc->camera_rays, (unsigned long long)cam);
        if (sh) atomicAdd(&dc->shadow_rays, (unsigned long long)sh);
        if (count_visits) {
            atomicAdd(&dc->inner_visits, (unsigned long long)vi);
            atomicAdd(&dc->packet_visits, (unsigned long long)vp);
            atomicAdd(&dc->leaf_visits, (unsigned long long)vl);
        }
    }
}

struct TileMap {  // how threads cover tiles: 8x4-pixel micro-tiles, one per warp
    const MpTile* tiles;
    uint32_t n_tiles;
    uint32_t mt_x;              // micro-tiles per tile row = ceil(tile_size / 8)
    uint32_t threads_per_tile;  // mt_x * ceil(tile_size / 4) * 32
};

__device__ __forceinline__ bool map_pixel(const TileMap& tm, unsigned long long gid, uint32_t& x, uint32_t& y, uint32_t& tile_index) {
    const uint32_t ti = (uint32_t)(gid / tm.threads_per_tile);
    if (ti >= tm.n_tiles) return false;
    const uint32_t rem = (uint32_t)(gid % tm.threads_per_tile);
    const uint32_t mt = rem >> 5, l = rem & 31u;
    const MpTile t = tm.tiles[ti];
    x = t.min_x + (mt % tm.mt_x) * 8u + (l & 7u);
    y = t.min_y + (mt / tm.mt_x) * 4u + (l >> 3);
    tile_index = ti;
    return x < t.max_x && y < t.max_y;
}

// Same mapping for launches whose thread index fits 32 bits (every wavefront batch): no 64-bit division.
__device__ __forceinline__ bool map_pixel32(const TileMap& tm, uint32_t gid, uint32_t& x, uint32_t& y, uint32_t& tile_index) {
    const uint32_t ti = gid / tm.threads_per_tile;
    if (ti >= tm.n_tiles) return false;
    const uint32_t rem = gid - ti * tm.threads_per_tile;
    const uint32_t mt = rem >> 5, l = rem & 31u;
    const MpTile t = tm.tiles[ti];
    const uint32_t my = mt / tm.mt_x;
    x = t.min_x + (mt - my * tm.mt_x) * 8u + (l & 7u);
    y = t.min_y + my * 4u + (l >> 3);
    tile_index = ti;
    return x < t.max_x && y < t.max_y;
}

// Cosine-weighted direction around n: Malley's method on rand_distr's unit disc, frame after Duff et al. 2017 (oracle: render_sample).
__device__ __forceinline__ F3 bounce_direction(const F3& n, Rng& rng) {
    float x, y;
    unit_disc(rng, x, y);
    const float z = __fsqrt_rn(fmaxf(0.0f, __fsub_rn(1.0f, __fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)))));
    const float sign = copysignf(1.0f, n.z);
    const float a = __fdiv_rn(-1.0f, __fadd_rn(sign, n.z));
    const float b = __fmul_rn(__fmul_rn(n.x, n.y), a);
    const F3 t1 = f3(__fadd_rn(1.0f, __fmul_rn(__fmul_rn(sign, __fmul_rn(n.x, n.x)), a)), __fmul_rn(sign, b), __fmul_rn(-sign, n.x));
    const F3 t2 = f3(b, __fadd_rn(sign, __fmul_rn(__fmul_rn(n.y, n.y), a)), -n.y);
    return (t1 * x + t2 * y) + n * z;
}

// The bounce loop of render_sample with a material table (oracle/oracle_render.inc: render_sample).
template <int LAYOUT, int ORDER, bool COUNT>
__device__ __noinline__ float4 render_sample_rgb(const Scene& sc, Ray ray, Rng rng, uint32_t max_bounces, ThreadCounts& tc) {
    F3 T = f3(1.0f, 1.0f, 1.0f), L = f3(0.0f, 0.0f, 0.0f);
    float alpha = 0.0f;
    for (uint32_t depth = 0;; ++depth) {
        tc.rays++;
        const Hit h = trace<LAYOUT, ORDER, COUNT>(sc, ray, false, tc.vis);
        if (h.tri == MP_NO_HIT) {
            L = L + T * EXT_ENV;
            break;
        }
        if (depth == 0) alpha = 1.0f;
        if (depth == max_bounces) break;
        F3 n = shading_normal(sc, h);
        if (dotn(ray.d, n) > 0.0f) n = neg(n);
        const F3 origin = (ray.o + ray.d * h.t) + n * sc.eps;
        const F3 albedo = surface_albedo(sc, h);
        T = f3(__fmul_rn(T.x, albedo.x), __fmul_rn(T.y, albedo.y), __fmul_rn(T.z, albedo.z));
        const F3 sun = f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z);
        const float c = dotn(n, sun);
        if (c > 0.0f) {
            const Ray sray = ray_new(origin, sun);
            tc.rays++, tc.shadow_rays++;
            const Hit sh = trace<LAYOUT, ORDER, COUNT>(sc, sray, true, tc.vis);
            if (sh.tri == MP_NO_HIT) L = L + (T * EXT_SUN_E) * c;
        }
        if (depth >= EXT_RR_START) {
            const float q = u01(rng);
            const float ps = max3(T);
            if (q >= ps) break;
            T = f3(__fdiv_rn(T.x, ps), __fdiv_rn(T.y, ps), __fdiv_rn(T.z, ps));
        }
        ray = ray_new(origin, bounce_direction(n, rng));
    }
    return make_float4(L.x, L.y, L.z, alpha);
}

// Worker::render_sample (worker.rs:51-65) for max_bounces == 0; extension integrator otherwise.
// RGB: the scene has a material table — throughput and radiance per channel; without one the three channels are equal and one
// scalar carries them (the arithmetic per channel is the same, so a grey table gives the film of no table).
template <int LAYOUT, int ORDER, bool COUNT, bool RGB>
__device__ __forceinline__ float4 render_sample(const Scene& sc, const Camera& cam, uint32_t px, uint32_t py, uint32_t pixel_index,
                                                uint32_t sample, uint32_t seed_hash, uint32_t max_bounces, ThreadCounts& tc) {
    Rng rng = rng_init(pixel_index, sample, seed_hash);
    Ray ray = sample_ray(cam, px, py, rng);
    tc.camera_rays++;
    if (max_bounces == 0) {
        tc.rays++;
        const Hit h = trace<LAYOUT, ORDER, COUNT>(sc, ray, false, tc.vis);
        if (h.tri == MP_NO_HIT) return make_float4(0.f, 0.f, 0.f, 0.f);
        const F3 n = shading_normal(sc, h);
        const float d = fabsf(dotn(ray.d, n));
        return make_float4(d, d, d, 1.0f);
    }
    if (RGB) return render_sample_rgb<LAYOUT, ORDER, COUNT>(sc, ray, rng, max_bounces, tc);
    float T = 1.0f, L = 0.0f, alpha = 0.0f;
    for (uint32_t depth = 0;; ++depth) {
        tc.rays++;
        const Hit h = trace<LAYOUT, ORDER, COUNT>(sc, ray, false, tc.vis);
        if (h.tri == MP_NO_HIT) {
            L = __fadd_rn(L, __fmul_rn(T, EXT_ENV));
            break;
        }
        if (depth == 0) alpha = 1.0f;
        if (depth == max_bounces) break;
        F3 n = shading_normal(sc, h);
        if (dotn(ray.d, n) > 0.0f) n = neg(n);
        const F3 origin = (ray.o + ray.d * h.t) + n * sc.eps;
        T = __fmul_rn(T, EXT_ALBEDO);
        const F3 sun = f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z);
        const float c = dotn(n, sun);
        if (c > 0.0f) {
            const Ray sray = ray_new(origin, sun);
            tc.rays++, tc.shadow_rays++;
            const Hit sh = trace<LAYOUT, ORDER, COUNT>(sc, sray, true, tc.vis);
            if (sh.tri == MP_NO_HIT) L = __fadd_rn(L, __fmul_rn(__fmul_rn(T, EXT_SUN_E), c));
        }
        if (depth >= EXT_RR_START) {
            const float q = u01(rng);
            if (q >= T) break;
            T = __fdiv_rn(T, T);
        }
        float x, y;
        unit_disc(rng, x, y);
        const float z = __fsqrt_rn(fmaxf(0.0f, __fsub_rn(1.0f, __fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)))));
        const float sign = copysignf(1.0f, n.z);
        const float a = __fdiv_rn(-1.0f, __fadd_rn(sign, n.z));
        const float b = __fmul_rn(__fmul_rn(n.x, n.y), a);
        const F3 t1 = f3(__fadd_rn(1.0f, __fmul_rn(__fmul_rn(sign, __fmul_rn(n.x, n.x)), a)), __fmul_rn(sign, b), __fmul_rn(-sign, n.x));
        const F3 t2 = f3(b, __fadd_rn(sign, __fmul_rn(__fmul_rn(n.y, n.y), a)), -n.y);
        const F3 dir = (t1 * x + t2 * y) + n * z;
        ray = ray_new(origin, dir);
    }
    return make_float4(L, L, L, alpha);
}

// Worker::render_tile for many tiles (worker.rs:32-49): one thread per pixel, samples summed sequentially.
template <int LAYOUT, int ORDER, bool COUNT, bool RGB>
// 6 blocks (24 warps) per SM = 80 registers with ~300 B of spills: +8 % over the unconstrained 128 registers / 16 warps on teapot and cube
// (profiles/r1f_variants_primary_ab2.json; 5, 7 and 8 blocks measured slower) - the kernel waits on fixed-latency dependencies, not on memory.
#ifndef MEGA_MIN_BLOCKS
#define MEGA_MIN_BLOCKS 6
#endif
__global__ void __launch_bounds__(128, MEGA_MIN_BLOCKS) render_mega_kernel(Scene sc, Camera cam, TileMap tm, float4* __restrict__ film, uint32_t width,
                                                          uint32_t sample_base, uint32_t sample_count, uint32_t max_bounces,
                                                          uint32_t seed_hash, DevCounters* dc) {
    const unsigned long long gid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x = 0, y = 0, ti = 0;
    const bool active = map_pixel(tm, gid, x, y, ti);
    ThreadCounts tc{};
    if (active) {
        const uint32_t pixel_index = y * width + x;
        float4 sum = film[pixel_index];
        for (uint32_t s = 0; s < sample_count; ++s) {
            const float4 c = render_sample<LAYOUT, ORDER, COUNT, RGB>(sc, cam, x, y, pixel_index, sample_base + s, seed_hash, max_bounces, tc);
            sum.x = __fadd_rn(sum.x, c.x), sum.y = __fadd_rn(sum.y, c.y), sum.z = __fadd_rn(sum.z, c.z), sum.w = __fadd_rn(sum.w, c.w);
        }
        film[pixel_index] = sum;
    }
    flush_counts(dc, tc, COUNT);
}

// Object::intersect on caller-provided rays (the bit-exactness gate).  Rays are 36-byte AoS records.
template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) trace_rays_kernel(Scene sc, const float* __restrict__ rays, unsigned long long n,
                                                         uint4* __restrict__ hits, DevCounters* dc) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    ThreadCounts tc{};
    if (i < n) {
        const float* p = rays + 9 * i;
        Ray r;
        r.o = f3(__ldg(p + 0), __ldg(p + 1), __ldg(p + 2));
        r.d = f3(__ldg(p + 3), __ldg(p + 4), __ldg(p + 5));
        r.inv = f3(__ldg(p + 6), __ldg(p + 7), __ldg(p + 8));
        tc.rays = 1;
        const Hit h = trace<LAYOUT, ORDER, COUNT>(sc, r, false, tc.vis);
        hits[i] = make_uint4(h.tri, __float_as_uint(h.t), __float_as_uint(h.u), __float_as_uint(h.v));
    }
    flush_counts(dc, tc, COUNT);
}

// Hit finalisation (ray_bvh_intersection.rs:66-95): LeafHitRecord → HitRecord {t, point, normal, material, texture_coords}.
// 12 words per record in the order of MpHitRecord.
__global__ void __launch_bounds__(128) finalise_hits_kernel(Scene sc, const float* __restrict__ rays, const uint4* __restrict__ hits, unsigned long long n,
                                                            uint32_t* __restrict__ out) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 h4 = hits[i];
    uint32_t* rec = out + 12 * i;
    if (h4.x == MP_NO_HIT) {
#pragma unroll
        for (int k = 0; k < 12; ++k) rec[k] = 0u;
        return;
    }
    const Hit h{__uint_as_float(h4.y), __uint_as_float(h4.z), __uint_as_float(h4.w), h4.x};
    const float* p = rays + 9 * i;
    const F3 o = f3(__ldg(p + 0), __ldg(p + 1), __ldg(p + 2)), d = f3(__ldg(p + 3), __ldg(p + 4), __ldg(p + 5));
    const uint4 sh = __ldg(sc.tri_shading + h.tri);
    const F3 point = o + d * h.t;  // Ray::point_at
    const F3 nrm = shading_normal(sc, h);
    const F3 tex = interpolate_tex(sc, sh, h);
    rec[0] = __float_as_uint(h.t);
    rec[1] = __float_as_uint(point.x), rec[2] = __float_as_uint(point.y), rec[3] = __float_as_uint(point.z);
    rec[4] = __float_as_uint(nrm.x), rec[5] = __float_as_uint(nrm.y), rec[6] = __float_as_uint(nrm.z);
    rec[7] = sh.w & MP_TRI_MATERIAL_MASK;
    rec[8] = __float_as_uint(tex.x), rec[9] = __float_as_uint(tex.y), rec[10] = __float_as_uint(tex.z);
    rec[11] = 1u;
}

// CameraSampler::sample_ray for every (pixel, sample) of the tiles, written in tile → row-major pixel → sample order.
__global__ void __launch_bounds__(128) generate_rays_kernel(Camera cam, TileMap tm, const unsigned long long* __restrict__ tile_offsets,
                                                            uint32_t width, uint32_t sample_base, uint32_t sample_count, uint32_t seed_hash,
                                                            float* __restrict__ out) {
    const unsigned long long gid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x, y, ti;
    if (!map_pixel(tm, gid, x, y, ti)) return;
    const MpTile t = tm.tiles[ti];
    const unsigned long long local = (unsigned long long)(y - t.min_y) * (t.max_x - t.min_x) + (x - t.min_x);
    for (uint32_t s = 0; s < sample_count; ++s) {
        Rng rng = rng_init(y * width + x, sample_base + s, seed_hash);
        const Ray r = sample_ray(cam, x, y, rng);
        float* p = out + 9 * ((tile_offsets[ti] + local) * sample_count + s);
        p[0] = r.o.x, p[1] = r.o.y, p[2] = r.o.z, p[3] = r.d.x, p[4] = r.d.y, p[5] = r.d.z, p[6] = r.inv.x, p[7] = r.inv.y, p[8] = r.inv.z;
    }
}

// pixel_sum * (1/spp) and color_to_image (worker.rs:44,69-76) for the whole film.
__global__ void __launch_bounds__(256) resolve_kernel(const float4* __restrict__ film, uint32_t n_pixels, float inv_spp,
                                                      uint32_t* __restrict__ rgba8) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    const float4 c = film[i];
    rgba8[i] = channel_to_u8(__fmul_rn(c.x, inv_spp)) | channel_to_u8(__fmul_rn(c.y, inv_spp)) << 8 |
               channel_to_u8(__fmul_rn(c.z, inv_spp)) << 16 | channel_to_u8(__fmul_rn(c.w, inv_spp)) << 24;
}

// Same for the pixels of a tile batch, into a tile-packed buffer [tile][tile_size][tile_size] (the worker's tile
// `buffer`, machinery.rs:66-67).
__global__ void __launch_bounds__(128) resolve_tiles_kernel(TileMap tm, const float4* __restrict__ film, uint32_t width, uint32_t tile_size,
                                                            float inv_spp, uint32_t* __restrict__ out) {
    const unsigned long long gid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x, y, ti;
    if (!map_pixel(tm, gid, x, y, ti)) return;
    const MpTile t = tm.tiles[ti];
    const float4 c = film[y * width + x];
    out[((size_t)ti * tile_size + (y - t.min_y)) * tile_size + (x - t.min_x)] =
        channel_to_u8(__fmul_rn(c.x, inv_spp)) | channel_to_u8(__fmul_rn(c.y, inv_spp)) << 8 |
        channel_to_u8(__fmul_rn(c.z, inv_spp)) << 16 | channel_to_u8(__fmul_rn(c.w, inv_spp)) << 24;
}

}  // namespace mpd

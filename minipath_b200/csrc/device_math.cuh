// Device-side arithmetic of the trace path, written so that every f32 result is bit-identical to the reference's
// AVX2+FMA arithmetic: compile with -fmad=false (no contraction), IEEE div/sqrt, no fast-math; fused operations
// appear only as explicit __fmaf_rn at the sites where the reference calls mul_add / mul_sub
// (reference src/util/simba.rs:57-67, src/scene/triangle_bvh/compressed_geometry.rs:103-108).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/minipath_b200.h"

namespace mpd {

#define MPD_INF __int_as_float(0x7f800000)
#define MPD_NINF __int_as_float(0xff800000)
#define MPD_FLT_MAX 3.402823466e+38f

struct F3 {
    float x, y, z;
};
__device__ __forceinline__ F3 f3(float x, float y, float z) { return F3{x, y, z}; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return F3{__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)}; }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return F3{__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)}; }
__device__ __forceinline__ F3 operator*(F3 a, float s) { return F3{__fmul_rn(a.x, s), __fmul_rn(a.y, s), __fmul_rn(a.z, s)}; }
__device__ __forceinline__ F3 operator/(F3 a, float s) { return F3{__fdiv_rn(a.x, s), __fdiv_rn(a.y, s), __fdiv_rn(a.z, s)}; }
__device__ __forceinline__ F3 neg(F3 a) { return F3{-a.x, -a.y, -a.z}; }
// nalgebra 3-vector dot and cross: unfused, (x*x' + y*y') + z*z'
__device__ __forceinline__ float dotn(F3 a, F3 b) { return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z)); }
__device__ __forceinline__ F3 crossn(F3 a, F3 b) {
    return F3{__fsub_rn(__fmul_rn(a.y, b.z), __fmul_rn(a.z, b.y)), __fsub_rn(__fmul_rn(a.z, b.x), __fmul_rn(a.x, b.z)),
              __fsub_rn(__fmul_rn(a.x, b.y), __fmul_rn(a.y, b.x))};
}
// fma_dot / fma_cross (util/simba.rs:57-67)
__device__ __forceinline__ float fma_dot(F3 a, F3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, __fmul_rn(a.x, b.x))); }
__device__ __forceinline__ F3 fma_cross(F3 a, F3 b) {
    return F3{__fmaf_rn(a.y, b.z, -__fmul_rn(a.z, b.y)), __fmaf_rn(a.z, b.x, -__fmul_rn(a.x, b.z)),
              __fmaf_rn(a.x, b.y, -__fmul_rn(a.y, b.x))};
}
__device__ __forceinline__ F3 normalize(F3 v) { return v / __fsqrt_rn(dotn(v, v)); }  // Unit::new_normalize

// vminps / vmaxps(self, rhs): the second operand on NaN or equality (wide fast_min / fast_max, util/simba.rs:69-75)
__device__ __forceinline__ float minps(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float maxps(float a, float b) { return a > b ? a : b; }

// UnitInterval8::decompress for lane `lane` of a u16x8 held in a uint4, then the FMA against the enclosing box
// (compressed_geometry.rs:48-51,95-110).  float(q)*K is evaluated as fma(m, K, -2^23*K) with m = 2^23 + q built by
// one byte-permute: m - 2^23 = q exactly and 2^23*K is exact, so the single rounding of the fma equals the single
// rounding of float(q)*K — bit-identical, without an int→float conversion.
#define MPD_K_INV_U16 (1.0f / 65535.0f)
__device__ __forceinline__ float unit_rel(uint32_t word, int odd) {
    float m = __uint_as_float(__byte_perm(word, 0x4B000000u, odd ? 0x7432 : 0x7410));
    return __fmaf_rn(m, MPD_K_INV_U16, -8388608.0f * MPD_K_INV_U16);
}
__device__ __forceinline__ uint32_t q_word(const uint4& q, int lane) {
    int w = lane >> 1;
    return w == 0 ? q.x : (w == 1 ? q.y : (w == 2 ? q.z : q.w));
}
__device__ __forceinline__ float dequant(const uint4& q, int lane, float size, float min) {
    return __fmaf_rn(size, unit_rel(q_word(q, lane), lane & 1), min);
}

// Ray::new (geometry/mod.rs:44-54)
struct Ray {
    F3 o, d, inv;
};
__device__ __forceinline__ Ray ray_new(F3 origin, F3 direction) {
    Ray r;
    r.o = origin;
    r.d = normalize(direction);
    r.inv.x = r.d.x == 0.0f ? MPD_INF : __fdiv_rn(1.0f, r.d.x);
    r.inv.y = r.d.y == 0.0f ? MPD_INF : __fdiv_rn(1.0f, r.d.y);
    r.inv.z = r.d.z == 0.0f ? MPD_INF : __fdiv_rn(1.0f, r.d.z);
    return r;
}

// Counter-based RNG shared with the CPU oracle (the reference's SmallRng is OS-seeded, worker.rs:25, so only the
// distributions of camera.rs:178-184 are reproduced): PCG-RXS-M-XS 32 with 64 bits per path — the 32-bit LCG state AND its own
// odd increment (the PCG "stream"), two independent hashes of (pixel, sample, seed).  One shared increment would make all paths
// windows of ONE 2^32-long sequence: a 4K x 1024 spp frame has 8.5e9 paths, more than that sequence has starting points.
__device__ __forceinline__ uint32_t pcg_hash(uint32_t v) {
    uint32_t s = v * 747796405u + 2891336453u;
    uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return (w >> 22) ^ w;
}
struct Rng {
    uint32_t state, inc;
};
__device__ __forceinline__ Rng rng_init(uint32_t pixel_index, uint32_t sample_index, uint32_t seed_hash) {
    Rng r;
    r.state = pcg_hash(pixel_index ^ pcg_hash(sample_index ^ seed_hash));
    r.inc = pcg_hash(sample_index ^ pcg_hash(pixel_index ^ ~seed_hash)) | 1u;
    return r;
}
__device__ __forceinline__ uint32_t rng_next(Rng& r) {
    uint32_t s = r.state * 747796405u + r.inc;
    r.state = s;
    uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return (w >> 22) ^ w;
}
__device__ __forceinline__ float u01(Rng& s) { return __fmul_rn((float)(rng_next(s) >> 8), 1.0f / 16777216.0f); }
// rand_distr::UnitDisc: rejection sampling
__device__ __forceinline__ void unit_disc(Rng& s, float& x, float& y) {
    for (;;) {
        float a = __fsub_rn(__fmul_rn(2.0f, u01(s)), 1.0f);
        float b = __fsub_rn(__fmul_rn(2.0f, u01(s)), 1.0f);
        if (__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)) <= 1.0f) {
            x = a, y = b;
            return;
        }
    }
}

struct Camera {  // CameraSampler (camera.rs:26-39)
    F3 center, up, right, film_origin_offset;
    float pixel_scale, lens_radius, lens_weight;
};

// CameraSampler::sample_ray (camera.rs:176-191)
__device__ __forceinline__ Ray sample_ray(const Camera& c, uint32_t px, uint32_t py, Rng& rng) {
    float film_u = __fadd_rn((float)px, __fsub_rn(u01(rng), 0.5f));
    float film_v = __fadd_rn((float)py, __fsub_rn(u01(rng), 0.5f));
    F3 fp = (c.film_origin_offset + c.up * __fmul_rn(film_v, c.pixel_scale)) - c.right * __fmul_rn(film_u, c.pixel_scale);
    float l0, l1;
    unit_disc(rng, l0, l1);
    F3 lens = c.right * __fmul_rn(c.lens_radius, l0) + c.up * __fmul_rn(c.lens_radius, l1);
    F3 dir = lens * c.lens_weight - fp;
    return ray_new(c.center + lens, dir);
}

// color_to_image (worker.rs:69-76): (c*255).round().clamp(0,255) as u8, round = half away from zero
__device__ __forceinline__ uint32_t channel_to_u8(float c) {
    float r = roundf(__fmul_rn(c, 255.0f));
    r = r < 0.0f ? 0.0f : (r > 255.0f ? 255.0f : r);
    return (r != r) ? 0u : (uint32_t)r;
}

}  // namespace mpd

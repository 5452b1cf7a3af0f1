// CPU restatement of minipath's trace path — TEST INFRASTRUCTURE (see oracle.h for scope and pinning status).
// Build: g++ -O2 -mavx2 -mfma -ffp-contract=off -shared -fPIC (top-level Makefile, target oracle/liboracle.so; __graft_entry__.build()).
#include "oracle.h"

#include <immintrin.h>
#include <pthread.h>

#include <atomic>
#include <chrono>
#include <cmath>
#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

namespace {

const float K_INV_U16 = 1.0f / 65535.0f;  // compressed_geometry.rs:49 (folded in f32)

// wide::f32x8::fast_min / fast_max with AVX = vminps / vmaxps(self, rhs): second operand on NaN or equality
inline float minps(float a, float b) { return a < b ? a : b; }
inline float maxps(float a, float b) { return a > b ? a : b; }

struct V3 {
    float x, y, z;
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3 neg(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 ld(const float* p) { return {p[0], p[1], p[2]}; }
inline void st(float* p, V3 v) { p[0] = v.x, p[1] = v.y, p[2] = v.z; }
// nalgebra 3-vector dot / cross: unfused
inline float dotn(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 crossn(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// util/simba.rs:57-67
inline float fma_dot(V3 a, V3 b) { return std::fmaf(a.z, b.z, std::fmaf(a.y, b.y, a.x * b.x)); }
inline V3 fma_cross(V3 a, V3 b) {
    return {std::fmaf(a.y, b.z, -(a.z * b.y)), std::fmaf(a.z, b.x, -(a.x * b.z)), std::fmaf(a.x, b.y, -(a.y * b.x))};
}
inline V3 normalize(V3 v) { return v / std::sqrt(dotn(v, v)); }  // Unit::new_normalize

inline float dequant(uint16_t q, float size, float min) {  // compressed_geometry.rs:48-51,95-110
    float rel = (float)(int32_t)q * K_INV_U16;
    return std::fmaf(size, rel, min);
}

std::atomic<int> g_simd{1};

}  // namespace

extern "C" {

uint16_t orc_compress_unit(float v, int rounding) {  // compressed_geometry.rs:25-46
    float s = v * 65535.0f;
    s = rounding == 0 ? std::nearbyintf(s) : (rounding == 1 ? std::floor(s) : std::ceil(s));
    s = minps(s, 65535.0f);
    s = maxps(s, 0.0f);
    return (uint16_t)(int32_t)s;
}
float orc_decompress_unit_rel(uint16_t q) { return (float)(int32_t)q * K_INV_U16; }
float orc_decompress_coord(uint16_t q, float size, float min) { return dequant(q, size, min); }

void orc_ray_new(const float origin[3], const float direction[3], MpRay* out) {  // geometry/mod.rs:44-54
    V3 d = normalize(ld(direction));
    st(out->origin, ld(origin));
    st(out->direction, d);
    out->inv_direction[0] = d.x == 0.0f ? INFINITY : 1.0f / d.x;
    out->inv_direction[1] = d.y == 0.0f ? INFINITY : 1.0f / d.y;
    out->inv_direction[2] = d.z == 0.0f ? INFINITY : 1.0f / d.z;
}

void orc_aabb8_intersect(const float bmin[3][8], const float bmax[3][8], const MpRay* ray, float max_t, float t1[8],
                         float t2[8]) {  // aabb.rs:257-283
    for (int i = 0; i < 8; ++i) {
        float cmin[3], cmax[3];
        for (int a = 0; a < 3; ++a) {
            float lo = (bmin[a][i] - ray->origin[a]) * ray->inv_direction[a];
            if (std::isnan(lo)) lo = -INFINITY;
            float hi = (bmax[a][i] - ray->origin[a]) * ray->inv_direction[a];
            if (std::isnan(hi)) hi = INFINITY;
            cmin[a] = minps(lo, hi);
            cmax[a] = maxps(lo, hi);
        }
        t1[i] = maxps(maxps(cmin[0], 0.0f), maxps(cmin[1], cmin[2]));
        t2[i] = minps(minps(cmax[0], max_t), minps(cmax[1], cmax[2]));
    }
}

uint32_t orc_triangle8_intersect(const float v[3][3][8], const MpRay* ray, float t[8], float u[8], float w[8]) {  // triangle.rs:187-216
    uint32_t mask = 0;
    V3 o = ld(ray->origin), d = ld(ray->direction);
    for (int i = 0; i < 8; ++i) {
        V3 v0{v[0][0][i], v[0][1][i], v[0][2][i]}, v1{v[1][0][i], v[1][1][i], v[1][2][i]}, v2{v[2][0][i], v[2][1][i], v[2][2][i]};
        V3 e1 = v1 - v0, e2 = v2 - v0;
        V3 p = fma_cross(d, e2);
        float det = fma_dot(e1, p);
        float inv = 1.0f / det;
        V3 s = o - v0;
        u[i] = inv * fma_dot(s, p);
        V3 q = fma_cross(s, e1);
        w[i] = inv * fma_dot(d, q);
        t[i] = inv * fma_dot(e2, q);
        if (u[i] >= 0.0f && w[i] >= 0.0f && (u[i] + w[i]) <= 1.0f) mask |= 1u << i;
    }
    return mask;
}

float orc_fma_dot(const float a[3], const float b[3]) { return fma_dot(ld(a), ld(b)); }
void orc_fma_cross(const float a[3], const float b[3], float out[3]) { st(out, fma_cross(ld(a), ld(b))); }

int orc_bit_iter(uint64_t bits, int out[64]) {  // util/mod.rs:17-26
    int n = 0;
    while (bits) {
        out[n++] = __builtin_ctzll(bits);
        bits &= bits - 1;
    }
    return n;
}

int orc_link_new_leaf(uint32_t index, uint32_t count, uint32_t* out) {  // mod.rs:78-82
    if (count < 1 || count > 7) return -1;
    if (index > MP_LINK_MAX_INDEX) return -1;  // index_vec MAX_INDEX check (mod.rs:137-141)
    *out = index << 3 | count;
    return 0;
}
int orc_link_new_inner(uint32_t index, uint32_t* out) {  // mod.rs:85-87
    if (index > MP_LINK_MAX_INDEX) return -1;
    *out = index << 3;
    return 0;
}
int orc_link_decode(uint32_t link, int* kind, uint32_t* index, uint32_t* count) {  // mod.rs:89-109
    if (link == MP_LINK_NULL) {
        *kind = 0, *index = 0, *count = 0;
        return 0;
    }
    *count = link & 7u;
    *index = link >> 3;
    *kind = *count == 0 ? 1 : 2;
    return 0;
}

uint8_t orc_color_channel_to_u8(float c) {  // worker.rs:71: (c*255).round().clamp(0,255) as u8 ; round = half away from zero
    float r = std::round(c * 255.0f);
    r = r < 0.0f ? 0.0f : (r > 255.0f ? 255.0f : r);  // NaN stays NaN through clamp, `as u8` saturating → 0
    return std::isnan(r) ? 0 : (uint8_t)r;
}

// ---- shared counter-based RNG (PCG-RXS-M-XS 32) ----
uint32_t orc_pcg_hash(uint32_t v) {
    uint32_t s = v * 747796405u + 2891336453u;
    uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return (w >> 22) ^ w;
}
// 64 bits per path: rng[0] = the LCG state, rng[1] = its odd increment (the PCG stream) — two hashes of (pixel, sample, seed)
void orc_rng_init(uint32_t pixel_index, uint32_t sample_index, uint64_t seed, uint32_t rng[2]) {
    uint32_t h = orc_pcg_hash((uint32_t)(seed >> 32));
    h = orc_pcg_hash((uint32_t)seed ^ h);  // the 32-bit seed hash both sides start from
    rng[0] = orc_pcg_hash(pixel_index ^ orc_pcg_hash(sample_index ^ h));
    rng[1] = orc_pcg_hash(sample_index ^ orc_pcg_hash(pixel_index ^ ~h)) | 1u;
}
uint32_t orc_rng_next(uint32_t rng[2]) {
    uint32_t s = rng[0] * 747796405u + rng[1];
    rng[0] = s;
    uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return (w >> 22) ^ w;
}

void orc_set_simd(int on) { g_simd.store(on ? 1 : 0); }
int orc_get_simd(void) { return g_simd.load(); }

}  // extern "C"

#include "oracle_trace.inc"
#include "oracle_render.inc"
#include "oracle_build.inc"

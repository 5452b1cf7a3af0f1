// Shared host-side helpers of libminipath_b200 (error reporting, small vector math).
#pragma once
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/minipath_b200.h"

namespace mp {

// Thread-local error message returned by mp_last_error().
std::string& last_error_ref();
int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));

struct V3 {
    float x, y, z;
};
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
static inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
static inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
// nalgebra dot of 3-vectors: (x*x' + y*y') + z*z', unfused
static inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
static inline float norm(V3 a) { return std::sqrt(dot(a, a)); }
static inline V3 normalize(V3 a) { return a / norm(a); }
static inline V3 vmin(V3 a, V3 b) { return V3{std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z)}; }
static inline V3 vmax(V3 a, V3 b) { return V3{std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z)}; }

struct Box3 {
    V3 min{INFINITY, INFINITY, INFINITY};
    V3 max{-INFINITY, -INFINITY, -INFINITY};
    void extend(V3 p) {
        min = vmin(min, p);
        max = vmax(max, p);
    }
    V3 size() const { return max - min; }
};

}  // namespace mp

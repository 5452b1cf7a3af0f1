// Roofline denominators that MEASURED_PEAKS.json does not carry: FP32 FMA issue rate and L2 read bandwidth,
// measured on the device the context lives on ("L2 BW and FP32 peak must be measured by microbenchmarks").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mpd {

// 8 independent FMA chains per thread: 2 flop per FFMA, ITERS * 8 FFMAs per thread.
__global__ void __launch_bounds__(256) probe_fp32_kernel(float* __restrict__ out, float a, float b, int iters) {
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
    for (int i = 0; i < iters; ++i) {
        x0 = __fmaf_rn(x0, a, b), x1 = __fmaf_rn(x1, a, b), x2 = __fmaf_rn(x2, a, b), x3 = __fmaf_rn(x3, a, b);
        x4 = __fmaf_rn(x4, a, b), x5 = __fmaf_rn(x5, a, b), x6 = __fmaf_rn(x6, a, b), x7 = __fmaf_rn(x7, a, b);
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 12345.678f) out[0] = s;  // keeps the chains alive
}

// Every block streams the whole buffer (resident in L2 after the warm-up launch) with 128-bit loads that bypass L1.
__global__ void __launch_bounds__(512) probe_l2_kernel(const uint4* __restrict__ buf, size_t n_vec, int passes, uint32_t* __restrict__ out) {
    uint32_t acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int p = 0; p < passes; ++p)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
            const uint4 v = __ldcg(buf + i);
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x9e3779b9u) out[0] = acc;
}

// The same stream with 256-bit loads (ld.global.cg.v8: L2 only), for the sweep that looks for the highest L2 read rate.
__global__ void __launch_bounds__(512) probe_l2_v8_kernel(const uint4* __restrict__ buf, size_t n_vec2, int passes, uint32_t* __restrict__ out) {
    uint32_t acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int p = 0; p < passes; ++p)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec2; i += stride) {
            uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
            asm volatile("ld.global.cg.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7)
                         : "l"(buf + 2 * i));
            acc ^= r0 ^ r1 ^ r2 ^ r3 ^ r4 ^ r5 ^ r6 ^ r7;
        }
    if (acc == 0x9e3779b9u) out[0] = acc;
}

// L1 gather roof of an incoherent traversal: every lane fetches ITS OWN record (SECTORS x 32 bytes, 32-byte aligned, picked by a
// per-lane LCG whose next state depends on the loaded data, like a traversal step) with 256-bit read-only loads.  The L1 pipe
// serves such a warp-wide gather one 32-byte sector per lane and cycle whatever the records hold, so this rate — sectors per
// second over the whole GPU — bounds any thread-per-node traversal (profiles/r1b_ubench_l1_patterns.txt, scripts/ubench/l1_patterns.cu).
template <int SECTORS>
__global__ void __launch_bounds__(128) probe_l1_gather_kernel(const uint4* __restrict__ recs, uint32_t n_recs, int iters, uint32_t* __restrict__ out) {
    uint32_t seed = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        seed = seed * 747796405u + 2891336453u;
        const uint32_t w = ((seed >> ((seed >> 28) + 4u)) ^ seed) * 277803737u;
        const uint32_t idx = ((w >> 22) ^ w) % n_recs;
        const uint4* n = recs + (size_t)idx * (2 * SECTORS);
#pragma unroll
        for (int k = 0; k < SECTORS; ++k) {
            uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7)
                         : "l"(n + 2 * k));
            acc ^= r0 ^ r3 ^ r5 ^ r7;
        }
        seed ^= acc & 1u;
    }
    if (acc == 0x12345678u) out[0] = acc;
}

}  // namespace mpd

// Scene decode: writes the MP_LAYOUT_DECODED arrays (trace.cuh) from the reference-layout arrays, once per mp_scene_upload.
// The arithmetic is RelativeBox8::decompress / RelativeTriangle8::decompress (reference
// src/scene/triangle_bvh/compressed_geometry.rs:95-110,133-135,160-162) through the very device function the per-visit
// path uses (`dequant`, device_math.cuh), so hoisting it cannot change a bit.
#pragma once
#include "trace.cuh"

namespace mpd {

// One thread per (node, child).  `flags` bit 0 is raised when a real child has min > max on some axis (impossible for the
// round-out quantisation of the builder, compressed_geometry.rs:122-131, but MpFlatBvh is caller data): the per-ray
// near/far row selection of inner_decoded would then be wrong and the context stays on the quantised layout.
__global__ void __launch_bounds__(256) decode_inner_kernel(Scene sc, uint32_t n_inner, float4* __restrict__ out, uint32_t* flags) {
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i = gid >> 3, c = gid & 7u;
    if (i >= n_inner) return;
    const float4 b0 = sc.inner_box[2 * (size_t)i], b1 = sc.inner_box[2 * (size_t)i + 1];
    const uint4* n = sc.inner + 8 * (size_t)i;
    const uint32_t link = reinterpret_cast<const uint32_t*>(n + 6)[c];
    float v[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const int a = k % 3;
        v[k] = dequant(n[k], (int)c, a == 0 ? b1.x : (a == 1 ? b1.y : b1.z), a == 0 ? b0.x : (a == 1 ? b0.y : b0.z));
    }
    if (link == MP_LINK_NULL) {
        v[0] = v[1] = v[2] = MPD_INF;
        v[3] = v[4] = v[5] = MPD_NINF;
    } else if (!(v[0] <= v[3] && v[1] <= v[4] && v[2] <= v[5])) {
        atomicOr(flags, 1u);
    }
    float* o = reinterpret_cast<float*>(out + MPD_NODE_W * (size_t)i);
#pragma unroll
    for (int k = 0; k < 6; ++k) o[k * 8 + c] = v[k];
    reinterpret_cast<uint32_t*>(o)[48 + c] = link;
    o[56 + c] = 0.0f;
}

// One thread per (packet, lane): v0 and the two edges (triangle.rs:193-194, unfused subtractions).
__global__ void __launch_bounds__(256) decode_packets_kernel(Scene sc, uint32_t n_packets, float4* __restrict__ out) {
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t p = gid >> 3, c = gid & 7u;
    if (p >= n_packets) return;
    const float4 b0 = sc.packet_box[2 * (size_t)p], b1 = sc.packet_box[2 * (size_t)p + 1];
    const uint4* q = sc.packets + 9 * (size_t)p;
    float v[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const int a = k % 3;
        v[k] = dequant(q[k], (int)c, a == 0 ? b1.x : (a == 1 ? b1.y : b1.z), a == 0 ? b0.x : (a == 1 ? b0.y : b0.z));
    }
    const F3 v0 = f3(v[0], v[1], v[2]), e1 = f3(v[3], v[4], v[5]) - v0, e2 = f3(v[6], v[7], v[8]) - v0;
    float* o = reinterpret_cast<float*>(out + MPD_PACK_W * (size_t)p);
    o[0 * 8 + c] = v0.x, o[1 * 8 + c] = v0.y, o[2 * 8 + c] = v0.z;
    o[3 * 8 + c] = e1.x, o[4 * 8 + c] = e1.y, o[5 * 8 + c] = e1.z;
    o[6 * 8 + c] = e2.x, o[7 * 8 + c] = e2.y, o[8 * 8 + c] = e2.z;
}

// Quantised records (trace.cuh: inner_rec / packets_rec): the reference rows, the links and the enclosing box interleaved into one
// 32-byte-aligned record per node / packet so that a lane's gather is 5 (6) 256-bit loads.  One thread per 16-byte word.
__global__ void __launch_bounds__(256) build_inner_records_kernel(Scene sc, uint32_t n_inner, uint4* __restrict__ out) {
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = gid / 10;
    const uint32_t w = (uint32_t)(gid % 10);
    if (i >= n_inner) return;
    out[gid] = w < 8 ? sc.inner[8 * i + w] : reinterpret_cast<const uint4*>(sc.inner_box)[2 * i + (w - 8)];
}
__global__ void __launch_bounds__(256) build_packet_records_kernel(Scene sc, uint32_t n_packets, uint4* __restrict__ out) {
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t p = gid / 12;
    const uint32_t w = (uint32_t)(gid % 12);
    if (p >= n_packets) return;
    out[gid] = w < 9 ? sc.packets[9 * p + w] : (w < 11 ? reinterpret_cast<const uint4*>(sc.packet_box)[2 * p + (w - 9)] : make_uint4(0u, 0u, 0u, 0u));
}

// Child-major copies for the octet kernels (octet.cuh): one 16-byte record per child / per triangle, + the row of eight v2.z.
__global__ void __launch_bounds__(256) build_inner_cm_kernel(Scene sc, uint32_t n_inner, uint4* __restrict__ out) {
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = gid >> 3;
    const uint32_t k = (uint32_t)(gid & 7u);
    if (i >= n_inner) return;
    const uint16_t* b = reinterpret_cast<const uint16_t*>(sc.inner + 8 * i);  // bounds[6][8]
    const uint32_t link = reinterpret_cast<const uint32_t*>(sc.inner + 8 * i + 6)[k];
    out[gid] = make_uint4(b[0 * 8 + k] | (uint32_t)b[1 * 8 + k] << 16, b[2 * 8 + k] | (uint32_t)b[3 * 8 + k] << 16,
                          b[4 * 8 + k] | (uint32_t)b[5 * 8 + k] << 16, link);
}
__global__ void __launch_bounds__(256) build_packets_cm_kernel(Scene sc, uint32_t n_packets, uint4* __restrict__ out) {
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t p = gid / 9;
    const uint32_t k = (uint32_t)(gid % 9);
    if (p >= n_packets) return;
    const uint16_t* v = reinterpret_cast<const uint16_t*>(sc.packets + 9 * p);  // v[vertex*3 + axis][lane]
    if (k < 8)
        out[gid] = make_uint4(v[0 * 8 + k] | (uint32_t)v[1 * 8 + k] << 16, v[2 * 8 + k] | (uint32_t)v[3 * 8 + k] << 16,
                              v[4 * 8 + k] | (uint32_t)v[5 * 8 + k] << 16, v[6 * 8 + k] | (uint32_t)v[7 * 8 + k] << 16);
    else
        out[gid] = sc.packets[9 * p + 8];  // the eight v2.z as they are
}

}  // namespace mpd

// Micro-benchmark: cost of fetching a BVH node per traversal step for different access shapes (divergent node indices).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o l1_patterns l1_patterns.cu ; run on the GPU box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 747796405u + 2891336453u; uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u; return (w >> 22) ^ w; }

__device__ __forceinline__ void ld256(const void* p, uint32_t (&r)[8]) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]) : "l"(p));
}

// PATTERN: 1 thread/node 10xLDG128 (128B node + 32B box) ; 2 thread/node 5xLDG256 (160B record) ; 3 thread/node 14xLDG128 (256B decoded)
//          4 thread/node 7xLDG256 ; 5 octet/node 2xLDG128 per lane (decoded child-major, 2 lines) ; 6 octet/node 1xLDG256 per lane
//          7 octet/node 1xLDG128 per lane (quantised child-major 128B) + 2xLDG128 box broadcast
template <int PATTERN>
__global__ void __launch_bounds__(128) bench(const uint4* __restrict__ nodes, const uint4* __restrict__ boxes, uint32_t n_nodes, int iters, uint32_t* out) {
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t seed = (PATTERN >= 5 ? (gid >> 3) : gid) * 2654435761u + 12345u;
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        const uint32_t idx = lcg(seed) % n_nodes;
        if (PATTERN == 1) {
            const uint4* n = nodes + (size_t)idx * 8;
#pragma unroll
            for (int k = 0; k < 8; ++k) { uint4 v = __ldg(n + k); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
            uint4 b0 = __ldg(boxes + 2 * (size_t)idx), b1 = __ldg(boxes + 2 * (size_t)idx + 1);
            acc ^= b0.x ^ b1.y;
        } else if (PATTERN == 2) {
            const uint4* n = nodes + (size_t)idx * 10;  // 160 B records, 32-byte aligned
#pragma unroll
            for (int k = 0; k < 5; ++k) { uint32_t r[8]; ld256(n + 2 * k, r); acc ^= r[0] ^ r[3] ^ r[5] ^ r[7]; }
        } else if (PATTERN == 3) {
            const uint4* n = nodes + (size_t)idx * 16;
#pragma unroll
            for (int k = 0; k < 14; ++k) { uint4 v = __ldg(n + k); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
        } else if (PATTERN == 4) {
            const uint4* n = nodes + (size_t)idx * 16;
#pragma unroll
            for (int k = 0; k < 7; ++k) { uint32_t r[8]; ld256(n + 2 * k, r); acc ^= r[0] ^ r[3] ^ r[5] ^ r[7]; }
        } else if (PATTERN == 5) {
            const uint4* n = nodes + (size_t)idx * 16;
            uint4 a = __ldg(n + (lane & 7u)), b = __ldg(n + 8 + (lane & 7u));
            acc ^= a.x ^ a.w ^ b.y ^ b.z;
        } else if (PATTERN == 6) {
            const uint4* n = nodes + (size_t)idx * 16;
            uint32_t r[8]; ld256(n + 2 * (lane & 7u), r); acc ^= r[0] ^ r[3] ^ r[5] ^ r[7];
        } else if (PATTERN == 7) {
            const uint4* n = nodes + (size_t)idx * 8;
            uint4 a = __ldg(n + (lane & 7u));
            uint4 b0 = __ldg(boxes + 2 * (size_t)idx), b1 = __ldg(boxes + 2 * (size_t)idx + 1);
            acc ^= a.x ^ a.w ^ b0.x ^ b1.y;
        }
        seed ^= acc & 1u;  // make the next index depend on the loaded data (like a traversal)
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int P>
static void run(const char* name, const uint4* nodes, const uint4* boxes, uint32_t n_nodes, uint32_t* out, int visits_per_warp_iter) {
    int bps = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, bench<P>, 128, 0);
    for (int occ : {2, 4, 8, 16}) {
        int blocks_per_sm = occ < bps ? occ : bps;
        const int iters = 2000;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        bench<P><<<148 * blocks_per_sm, 128>>>(nodes, boxes, n_nodes, 100, out);
        cudaEventRecord(e0);
        bench<P><<<148 * blocks_per_sm, 128>>>(nodes, boxes, n_nodes, iters, out);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double visits = (double)148 * blocks_per_sm * 4 * visits_per_warp_iter * iters;
        printf("%-44s blocks/SM %2d  %.3f ms  %.2f Gvisits/s  %.2f clk/visit/SM\n", name, blocks_per_sm, ms, visits / ms / 1e6, ms * 1e-3 * 1.965e9 * 148 / visits);
    }
}

int main() {
    const uint32_t n_nodes = 40000;  // ~ the stand-in's inner+leaf records; 10 MB at 256 B
    uint4 *nodes, *boxes; uint32_t* out;
    cudaMalloc(&nodes, (size_t)n_nodes * 256); cudaMalloc(&boxes, (size_t)n_nodes * 32); cudaMalloc(&out, 4);
    cudaMemset(nodes, 1, (size_t)n_nodes * 256); cudaMemset(boxes, 2, (size_t)n_nodes * 32);
    run<1>("thread/node 10xLDG.128 (quantised + box)", nodes, boxes, n_nodes, out, 32);
    run<2>("thread/node 5xLDG.256 (160 B record)", nodes, boxes, n_nodes, out, 32);
    run<3>("thread/node 14xLDG.128 (decoded 256 B)", nodes, boxes, n_nodes, out, 32);
    run<4>("thread/node 7xLDG.256 (decoded 256 B)", nodes, boxes, n_nodes, out, 32);
    run<5>("octet/node 2xLDG.128 per lane (decoded cm)", nodes, boxes, n_nodes, out, 4);
    run<6>("octet/node 1xLDG.256 per lane (decoded cm)", nodes, boxes, n_nodes, out, 4);
    run<7>("octet/node LDG.128 + 2 box (quantised cm)", nodes, boxes, n_nodes, out, 4);
    return 0;
}

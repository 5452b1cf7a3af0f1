// Device BVH8 builder — SURVEY §8 f1: TriangleBvh::build (reference src/scene/triangle_bvh/building.rs:83-207) on the GPU, emitting the
// reference node format (MpInnerNode / MpTrianglePacket, include/minipath_b200.h).
//
// What is kept from the reference, bit for bit in its arithmetic:
//   * an inner node has 2..8 children, a leaf 1..7 packets of 8 triangles (mod.rs:14-17), padding lanes all-zero (building.rs:157);
//   * child boxes are quantised ROUND-OUT (floor the min, ceil the max) against the node's enclosing box and every child is then built
//     against the DECOMPRESSED box (building.rs:146-155; compressed_geometry.rs:122-131) — hence a strict top-down pass;
//   * leaf vertices are quantised round-to-nearest-even against the leaf's decompressed box (building.rs:183-195);
//   * parents precede their children in `inner` (building.rs:130-132).
// What is different: the SPLIT.  The reference runs a binned agglomerative SAH merge per node (building.rs:238-345), a sequential
// O(bins^3) loop; here the triangles are sorted once along a 63-bit Morton curve of their centroids and every node cuts its range of
// that order into up to 8 children: three rounds of radix splits (the highest differing Morton bit, Karras 2012) while a node holds
// more than 8 leaves' worth of triangles, equal runs of whole packets when it can be covered by at most 8 leaves (a leaf holds up to
// DEVICE_LEAF_DEFAULT triangles, not the format's 56: a visited leaf tests all its packets).  The tree is therefore NOT the
// reference's tree — its quality is measured against it (visits per ray, tests/test_gpu_builder.py, profiles/) — but any traversal
// of it follows the same rules and gives the same bits on CPU and GPU.
//
// Pipeline (all on the device, breadth-first, one small read-back of two counters per level):
//   prepare  per-triangle box + centroid, scene and centroid bounds           (1 launch)
//   sort     Morton keys, cub::DeviceRadixSort (library call: not part of the trace hot path)
//   split    per level: child ranges → exclusive scans (cub::DeviceScan) → child ids, leaf packet ranges, links
//   boxes    exact leaf boxes, then exact node boxes bottom-up per level
//   quantise per level top-down: round-out child boxes, decompressed boxes handed to the children
//   leaves   packets, per-triangle shading records
// The result is copied into a host MpFlatBvh (the interchange format every other entry point takes).
#include <cub/cub.cuh>

#include <vector>

#include "host_common.h"

namespace mp {
MpFlatBvh* adopt_flat_bvh(const float bbox_min[3], const float bbox_max[3], uint32_t root, uint32_t n_triangles, std::vector<MpInnerNode>& inner,
                          std::vector<MpBox>& inner_box, std::vector<MpTrianglePacket>& packets, std::vector<MpBox>& packet_box,
                          std::vector<MpTriShading>& tri_shading, std::vector<MpVertexShading>& vertices, std::vector<MpMaterial>& materials);
}

namespace mpb {

#define CKB(x)                                                                                         \
    do {                                                                                               \
        cudaError_t e_ = (x);                                                                          \
        if (e_ != cudaSuccess) return mp::fail(MP_ECUDA, "%s: %s", #x, cudaGetErrorString(e_));        \
    } while (0)

struct Range {
    uint32_t begin, end;
};
struct NodeTopo {        // one inner node while the hierarchy is being laid out
    Range child[8];      // ranges of the Morton order; begin == end: no child
    uint32_t link[8];    // CompressedNodeLink of every child once its index is known
    uint32_t n_inner;    // children that are inner nodes
    uint32_t n_packets;  // packets of the children that are leaves
};
struct Leaf {
    uint32_t begin, end, packet_base;
};
struct BoxF {
    float lo[3], hi[3];
};

__device__ __forceinline__ uint32_t float_flip(float f) {  // order-preserving map float -> uint
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float float_unflip(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

// bounds[0..2] scene min, [3..5] scene max, [6..8] centroid min, [9..11] centroid max — as flipped uints for atomicMin / atomicMax
__global__ void __launch_bounds__(256) prepare_kernel(const float* __restrict__ pos, const uint32_t* __restrict__ tri, uint32_t n_tri,
                                                      BoxF* __restrict__ tri_box, float* __restrict__ centroid, uint32_t* bounds) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}, c[3] = {0.f, 0.f, 0.f};
    const bool act = i < n_tri;
    if (act) {
        for (int k = 0; k < 3; ++k) {
            const float* p = pos + 3 * (size_t)tri[3 * (size_t)i + k];
            for (int a = 0; a < 3; ++a) {
                lo[a] = fminf(lo[a], p[a]), hi[a] = fmaxf(hi[a], p[a]);
                c[a] = __fadd_rn(c[a], p[a]);  // triangle.rs:115-120: ((0 + p0) + p1 + p2) / 3
            }
        }
        BoxF b;
        for (int a = 0; a < 3; ++a) b.lo[a] = lo[a], b.hi[a] = hi[a], c[a] = __fdiv_rn(c[a], 3.0f), centroid[3 * (size_t)i + a] = c[a];
        tri_box[i] = b;
    }
    // warp reduction, one atomic per warp and value
    for (int a = 0; a < 3; ++a) {
        uint32_t vlo = act ? float_flip(lo[a]) : 0xffffffffu, vhi = act ? float_flip(hi[a]) : 0u;
        uint32_t clo = act ? float_flip(c[a]) : 0xffffffffu, chi = act ? float_flip(c[a]) : 0u;
        vlo = __reduce_min_sync(0xffffffffu, vlo), vhi = __reduce_max_sync(0xffffffffu, vhi);
        clo = __reduce_min_sync(0xffffffffu, clo), chi = __reduce_max_sync(0xffffffffu, chi);
        if ((threadIdx.x & 31) == 0) {
            atomicMin(bounds + a, vlo), atomicMax(bounds + 3 + a, vhi);
            atomicMin(bounds + 6 + a, clo), atomicMax(bounds + 9 + a, chi);
        }
    }
}

__device__ __forceinline__ unsigned long long spread21(uint32_t v) {  // 21 bits -> every third bit of 63
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(256) morton_kernel(const float* __restrict__ centroid, uint32_t n_tri, const uint32_t* __restrict__ bounds,
                                                     unsigned long long* __restrict__ keys, uint32_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tri) return;
    unsigned long long code = 0;
    for (int a = 0; a < 3; ++a) {
        const float lo = float_unflip(bounds[6 + a]), hi = float_unflip(bounds[9 + a]);
        const float ext = hi - lo;
        float t = ext > 0.0f ? (centroid[3 * (size_t)i + a] - lo) / ext : 0.0f;
        t = fminf(fmaxf(t, 0.0f), 1.0f);
        const uint32_t q = min(2097151u, (uint32_t)(t * 2097152.0f));
        code |= spread21(q) << (2 - a);
    }
    keys[i] = code;
    ids[i] = i;
}

// First index of the upper half of [b, e) when the range is cut at the highest Morton bit in which its ends differ; the middle
// (in whole packets) when all its codes are equal.
__device__ __forceinline__ uint32_t radix_split(const unsigned long long* __restrict__ keys, uint32_t b, uint32_t e) {
    const unsigned long long first = keys[b], last = keys[e - 1];
    if (first == last) {
        uint32_t mid = b + ((((e - b) / 2) + 7u) & ~7u);
        if (mid >= e) mid = b + (e - b) / 2;
        return mid;
    }
    const int prefix = __clzll((long long)(first ^ last));
    uint32_t lo = b, hi = e - 1;  // keys[lo] shares more than `prefix` bits with first, keys[hi] does not
    while (hi - lo > 1) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (__clzll((long long)(first ^ keys[mid])) > prefix)
            lo = mid;
        else
            hi = mid;
    }
    return hi;
}

// Triangles a leaf may hold.  The format allows 56 (7 packets, mod.rs:16-17), but a visited leaf tests ALL its packets: the reference's
// cost model (0.75 per packet against 1 per inner node, building.rs:358-383) ends up near 2 packets per leaf on real scenes; one packet per leaf is
// the measured optimum here (profiles/r2g_device_builder_leaf_size.txt: fewer sectors per ray than the reference split).  MINIPATH_B200_DEVICE_LEAF overrides it (8..56, a multiple of 8).
#define DEVICE_LEAF_DEFAULT 8u

// Children of every node of one level.  node_range[first + k] is the Morton range of the level's k-th node.
__global__ void __launch_bounds__(128) split_kernel(const unsigned long long* __restrict__ keys, const Range* __restrict__ node_range, uint32_t first,
                                                    uint32_t count, uint32_t LEAF_MAX, NodeTopo* __restrict__ topo, uint32_t* __restrict__ n_inner, uint32_t* __restrict__ n_packets) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    const Range r = node_range[first + k];
    const uint32_t n = r.end - r.begin;  // > LEAF_MAX
    Range c[8];
    int m = 0;
    if (n <= 8u * LEAF_MAX) {
        // at most 8 leaves cover the node: equal runs of whole packets (full packets, leaves of one size)
        const uint32_t leaves = (n + LEAF_MAX - 1) / LEAF_MAX;                 // 2..8
        const uint32_t per = ((n + leaves - 1) / leaves + 7u) & ~7u;           // <= LEAF_MAX
        for (uint32_t b = r.begin; b < r.end; b += per) c[m++] = Range{b, min(b + per, r.end)};
    } else {
        c[m++] = r;
        for (int round = 0; round < 3; ++round) {
            Range nxt[8];
            int mm = 0;
            for (int j = 0; j < m; ++j) {
                if (c[j].end - c[j].begin <= LEAF_MAX) {
                    nxt[mm++] = c[j];
                    continue;
                }
                const uint32_t s = radix_split(keys, c[j].begin, c[j].end);
                nxt[mm++] = Range{c[j].begin, s};
                nxt[mm++] = Range{s, c[j].end};
            }
            m = mm;
            for (int j = 0; j < m; ++j) c[j] = nxt[j];
        }
    }
    NodeTopo t;
    uint32_t ni = 0, np = 0;
    for (int j = 0; j < 8; ++j) {
        t.child[j] = j < m ? c[j] : Range{0u, 0u};
        t.link[j] = MP_LINK_NULL;
        if (j < m) {
            const uint32_t cn = c[j].end - c[j].begin;
            if (cn > LEAF_MAX)
                ++ni;
            else
                np += (cn + 7u) / 8u;
        }
    }
    t.n_inner = ni, t.n_packets = np;
    topo[first + k] = t;
    n_inner[k] = ni, n_packets[k] = np;
}

// Child indices from the scanned counts: inner children become the nodes next_first + inner_off[k] .., leaves take packets from
// packet_first + packet_off[k] on and a slot of `leaves` from leaf_first + leaf_off.. (leaf_off is scanned separately).
__global__ void __launch_bounds__(128) assign_kernel(NodeTopo* __restrict__ topo, Range* __restrict__ node_range, uint32_t first, uint32_t count, uint32_t LEAF_MAX,
                                                     const uint32_t* __restrict__ inner_off, const uint32_t* __restrict__ packet_off, uint32_t next_first,
                                                     uint32_t packet_first, Leaf* __restrict__ leaves, uint32_t* __restrict__ leaf_count,
                                                     uint32_t* __restrict__ parent_slot) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    NodeTopo& t = topo[first + k];
    uint32_t ni = next_first + inner_off[k], pk = packet_first + packet_off[k];
    for (int j = 0; j < 8; ++j) {
        const uint32_t cn = t.child[j].end - t.child[j].begin;
        if (cn == 0) continue;
        if (cn > LEAF_MAX) {
            node_range[ni] = t.child[j];
            parent_slot[ni] = (first + k) << 3 | (uint32_t)j;
            t.link[j] = ni << MP_LINK_COUNT_BITS;
            ++ni;
        } else {
            const uint32_t packets = (cn + 7u) / 8u;
            const uint32_t l = atomicAdd(leaf_count, 1u);  // the order of `leaves` is free: nothing indexes it but the link below
            leaves[l] = Leaf{t.child[j].begin, t.child[j].end, pk};
            t.link[j] = pk << MP_LINK_COUNT_BITS | packets;
            pk += packets;
        }
    }
}

// exact box of the triangles of a Morton range
__device__ __forceinline__ BoxF range_box(const BoxF* __restrict__ tri_box, const uint32_t* __restrict__ ids, uint32_t b, uint32_t e) {
    BoxF r{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
    for (uint32_t i = b; i < e; ++i) {
        const BoxF t = tri_box[ids[i]];
        for (int a = 0; a < 3; ++a) r.lo[a] = fminf(r.lo[a], t.lo[a]), r.hi[a] = fmaxf(r.hi[a], t.hi[a]);
    }
    return r;
}

// Exact boxes of the children of one level's nodes (leaf children from their triangles, inner children from the level below, which
// is already done) and of the nodes themselves.
__global__ void __launch_bounds__(128) exact_kernel(const NodeTopo* __restrict__ topo, uint32_t first, uint32_t count, const BoxF* __restrict__ tri_box,
                                                    const uint32_t* __restrict__ ids, BoxF* __restrict__ node_exact, BoxF* __restrict__ child_exact) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (node, child slot)
    const uint32_t k = g >> 3, j = g & 7u;
    BoxF b{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
    if (k < count) {
        const NodeTopo& t = topo[first + k];
        const uint32_t link = t.link[j];
        if (link != MP_LINK_NULL) {
            if ((link & MP_LINK_COUNT_MASK) == 0u)
                b = node_exact[link >> MP_LINK_COUNT_BITS];
            else
                b = range_box(tri_box, ids, t.child[j].begin, t.child[j].end);
            child_exact[(size_t)(first + k) * 8 + j] = b;
        }
    }
    // union over the 8 slots of the node (8 consecutive lanes)
    for (int a = 0; a < 3; ++a)
        for (int d = 1; d < 8; d <<= 1) {
            b.lo[a] = fminf(b.lo[a], __shfl_xor_sync(0xffffffffu, b.lo[a], d));
            b.hi[a] = fmaxf(b.hi[a], __shfl_xor_sync(0xffffffffu, b.hi[a], d));
        }
    if (k < count && j == 0) node_exact[first + k] = b;
}

// UnitInterval8::compress_internal for one lane (compressed_geometry.rs:25-46) and its inverse (:48-51,:103-108)
__device__ __forceinline__ uint32_t compress_unit(float v, int rounding) {  // 0 nearest-even, 1 floor, 2 ceil
    float s = __fmul_rn(v, 65535.0f);
    s = rounding == 0 ? rintf(s) : (rounding == 1 ? floorf(s) : ceilf(s));
    s = (s < 65535.0f) ? s : 65535.0f;  // minps: NaN -> 65535
    s = (s > 0.0f) ? s : 0.0f;
    return (uint32_t)(int32_t)s;
}
__device__ __forceinline__ float decompress_unit(uint32_t q, float size, float mn) {
    return __fmaf_rn(size, __fmul_rn((float)(int32_t)q, 1.0f / 65535.0f), mn);
}

// Round-out quantisation of one level (building.rs:134-155): node k's enclosing box is known (the root's is the scene box, every
// other one was written by its parent); its children's decompressed boxes become THEIR enclosing boxes.
__global__ void __launch_bounds__(128) quantise_kernel(const NodeTopo* __restrict__ topo, uint32_t first, uint32_t count, const BoxF* __restrict__ child_exact,
                                                       BoxF* __restrict__ node_enclosing, BoxF* __restrict__ packet_enclosing, MpInnerNode* __restrict__ inner,
                                                       MpBox* __restrict__ inner_box) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t k = g >> 3, j = g & 7u;
    if (k >= count) return;
    const uint32_t node = first + k;
    const BoxF enc = node_enclosing[node];
    float size[3];
    for (int a = 0; a < 3; ++a) size[a] = __fsub_rn(enc.hi[a], enc.lo[a]);
    const uint32_t link = topo[node].link[j];
    MpInnerNode& out = inner[node];
    out.links[j] = link;
    if (link == MP_LINK_NULL) {
        for (int a = 0; a < 6; ++a) out.bounds[a][j] = 0;
    } else {
        const BoxF ex = child_exact[(size_t)node * 8 + j];
        BoxF dec;
        for (int a = 0; a < 3; ++a) {
            const uint32_t qlo = compress_unit(__fdiv_rn(__fsub_rn(ex.lo[a], enc.lo[a]), size[a]), 1);
            const uint32_t qhi = compress_unit(__fdiv_rn(__fsub_rn(ex.hi[a], enc.lo[a]), size[a]), 2);
            out.bounds[a][j] = (uint16_t)qlo, out.bounds[3 + a][j] = (uint16_t)qhi;
            dec.lo[a] = decompress_unit(qlo, size[a], enc.lo[a]), dec.hi[a] = decompress_unit(qhi, size[a], enc.lo[a]);
        }
        const uint32_t index = link >> MP_LINK_COUNT_BITS, packets = link & MP_LINK_COUNT_MASK;
        if (packets == 0u)
            node_enclosing[index] = dec;
        else
            for (uint32_t p = 0; p < packets; ++p) packet_enclosing[index + p] = dec;
    }
    if (j == 0) {  // From<&AABB> for AABBSized (aabb.rs:292-303)
        MpBox b{};
        for (int a = 0; a < 3; ++a) b.min[a] = enc.lo[a], b.size[a] = size[a];
        inner_box[node] = b;
    }
}

// Leaves (building.rs:170-207): one thread per triangle lane of every packet.
__global__ void __launch_bounds__(256) leaf_kernel(const Leaf* __restrict__ leaves, uint32_t n_leaves, const uint32_t* __restrict__ lane_leaf,
                                                   uint32_t n_lanes, const uint32_t* __restrict__ ids, const float* __restrict__ pos,
                                                   const uint32_t* __restrict__ tri, const uint32_t* __restrict__ tri_material,
                                                   const unsigned char* __restrict__ zero_normal, const BoxF* __restrict__ packet_enclosing,
                                                   MpTrianglePacket* __restrict__ packets, MpBox* __restrict__ packet_box, MpTriShading* __restrict__ shading) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;  // packet * 8 + lane
    if (g >= n_lanes) return;
    const uint32_t packet = g >> 3, lane = g & 7u;
    const Leaf lf = leaves[lane_leaf[packet]];
    const uint32_t i = lf.begin + (packet - lf.packet_base) * 8u + lane;  // position in the Morton order
    const BoxF enc = packet_enclosing[packet];
    float size[3];
    for (int a = 0; a < 3; ++a) size[a] = __fsub_rn(enc.hi[a], enc.lo[a]);
    MpTriShading sh{{0u, 0u, 0u}, 0u};  // TriangleShadingData::default() on padding lanes
    uint16_t q[3][3] = {};              // padding lanes all-zero (building.rs:157)
    if (i < lf.end) {
        const uint32_t t = ids[i];
        bool flat = false;
        for (int k = 0; k < 3; ++k) {
            const uint32_t v = tri[3 * (size_t)t + k];
            for (int a = 0; a < 3; ++a) q[k][a] = (uint16_t)compress_unit(__fdiv_rn(__fsub_rn(pos[3 * (size_t)v + a], enc.lo[a]), size[a]), 0);
            flat |= zero_normal[v] != 0;  // building.rs:200
            sh.v[k] = v;
        }
        sh.flags_material = (flat ? MP_TRI_FLAT : 0u) | (tri_material ? tri_material[t] : 0u);
    }
    for (int k = 0; k < 3; ++k)
        for (int a = 0; a < 3; ++a) packets[packet].v[k][a][lane] = q[k][a];
    shading[g] = sh;
    if (lane == 0) {
        MpBox b{};
        for (int a = 0; a < 3; ++a) b.min[a] = enc.lo[a], b.size[a] = size[a];
        packet_box[packet] = b;
    }
}

__global__ void __launch_bounds__(256) lane_leaf_kernel(const Leaf* __restrict__ leaves, uint32_t n_leaves, uint32_t* __restrict__ lane_leaf) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n_leaves) return;
    const Leaf lf = leaves[l];
    const uint32_t packets = (lf.end - lf.begin + 7u) / 8u;
    for (uint32_t p = 0; p < packets; ++p) lane_leaf[lf.packet_base + p] = l;
}

struct DevBuf {  // frees what it holds when the build returns, on every path
    std::vector<void*> ptrs;
    ~DevBuf() {
        for (void* p : ptrs) cudaFree(p);
    }
    template <class T>
    cudaError_t alloc(T** out, size_t n) {
        cudaError_t e = cudaMalloc((void**)out, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(*out);
        return e;
    }
};

}  // namespace mpb

extern "C" int mp_bvh_build_device(int device, const float* positions, const float* normals, const float* texcoords, uint32_t n_vertices,
                                   const uint32_t* triangles, const uint32_t* tri_material, uint64_t n_triangles, const MpMaterial* materials,
                                   uint32_t n_materials, MpFlatBvh** out, float* out_device_ms) {
    using namespace mpb;
    if (!positions || !triangles || !out || (n_materials && !materials)) return mp::fail(MP_EINVAL, "null argument");
    if (n_triangles == 0) return mp::fail(MP_EINVAL, "scene has no triangles (reference asserts in build_leaf, building.rs:178)");
    if (n_triangles > 0x7fffffffull / 8) return mp::fail(MP_EINVAL, "too many triangles for the 29-bit link index range (mod.rs:71)");
    const uint32_t nt = (uint32_t)n_triangles;
    uint32_t LEAF_MAX = DEVICE_LEAF_DEFAULT;
    if (const char* e = getenv("MINIPATH_B200_DEVICE_LEAF")) {
        const long v = atol(e);
        if (v < 8 || v > MP_LEAF_MAX_TRIANGLES || v % 8) return mp::fail(MP_EINVAL, "MINIPATH_B200_DEVICE_LEAF must be a multiple of 8 in 8..56");
        LEAF_MAX = (uint32_t)v;
    }
    for (uint64_t i = 0; i < 3 * n_triangles; ++i)
        if (triangles[i] >= n_vertices) return mp::fail(MP_EINVAL, "triangle %llu references vertex %u >= %u", (unsigned long long)(i / 3), triangles[i], n_vertices);
    if (tri_material)
        for (uint64_t i = 0; i < n_triangles; ++i)
            if (tri_material[i] >= std::max(n_materials, 1u)) return mp::fail(MP_EINVAL, "triangle material %u outside the table of %u", tri_material[i], n_materials);
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) return mp::fail(MP_ENODEVICE, "no CUDA device %d", device);
    int prev_device = 0;
    CKB(cudaGetDevice(&prev_device));
    struct DeviceGuard {  // the caller's current device is restored on every path
        int d;
        ~DeviceGuard() { cudaSetDevice(d); }
    } device_guard{prev_device};
    CKB(cudaSetDevice(device));

    // host side of the vertex records (building.rs:56-66): normals normalised, a zero normal marks its triangles flat
    std::vector<MpVertexShading> vertices(n_vertices);
    std::vector<unsigned char> zero_normal(n_vertices, 1);
    for (uint32_t i = 0; i < n_vertices; ++i) {
        MpVertexShading v{};
        if (normals) {
            mp::V3 n = mp::v3(normals[3 * (size_t)i], normals[3 * (size_t)i + 1], normals[3 * (size_t)i + 2]);
            if (mp::dot(n, n) != 0.0f) {
                n = mp::normalize(n);
                zero_normal[i] = 0;
            }
            v.normal[0] = n.x, v.normal[1] = n.y, v.normal[2] = n.z;
        }
        if (texcoords) v.tex[0] = texcoords[2 * (size_t)i], v.tex[1] = texcoords[2 * (size_t)i + 1];
        vertices[i] = v;
    }

    cudaStream_t st;
    CKB(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    struct StreamGuard {
        cudaStream_t s;
        ~StreamGuard() { cudaStreamDestroy(s); }
    } sg{st};
    cudaEvent_t ev0, ev1;
    CKB(cudaEventCreate(&ev0));
    CKB(cudaEventCreate(&ev1));
    struct EventGuard {
        cudaEvent_t a, b;
        ~EventGuard() { cudaEventDestroy(a), cudaEventDestroy(b); }
    } eg{ev0, ev1};

    DevBuf dev;
    float *d_pos, *d_centroid;
    uint32_t *d_tri, *d_mat = nullptr, *d_bounds, *d_ids, *d_ids_sorted, *d_cnt_inner, *d_cnt_packets, *d_off_inner, *d_off_packets, *d_leaf_count, *d_parent_slot;
    unsigned char* d_zero;
    unsigned long long *d_keys, *d_keys_sorted;
    BoxF *d_tri_box, *d_node_exact, *d_child_exact, *d_node_enclosing;
    Range* d_node_range;
    NodeTopo* d_topo;
    Leaf* d_leaves;
    const size_t node_cap = nt;  // every inner node has >= 2 children and every leaf >= 1 triangle
    CKB(dev.alloc(&d_pos, 3 * (size_t)n_vertices));
    CKB(dev.alloc(&d_tri, 3 * (size_t)nt));
    CKB(dev.alloc(&d_zero, n_vertices));
    if (tri_material) CKB(dev.alloc(&d_mat, nt));
    CKB(dev.alloc(&d_centroid, 3 * (size_t)nt));
    CKB(dev.alloc(&d_tri_box, nt));
    CKB(dev.alloc(&d_bounds, 12));
    CKB(dev.alloc(&d_keys, nt));
    CKB(dev.alloc(&d_keys_sorted, nt));
    CKB(dev.alloc(&d_ids, nt));
    CKB(dev.alloc(&d_ids_sorted, nt));
    CKB(dev.alloc(&d_node_range, node_cap));
    CKB(dev.alloc(&d_parent_slot, node_cap));
    CKB(dev.alloc(&d_topo, node_cap));
    CKB(dev.alloc(&d_leaves, nt));
    CKB(dev.alloc(&d_leaf_count, 1));
    CKB(dev.alloc(&d_cnt_inner, node_cap));
    CKB(dev.alloc(&d_cnt_packets, node_cap));
    CKB(dev.alloc(&d_off_inner, node_cap + 1));
    CKB(dev.alloc(&d_off_packets, node_cap + 1));
    CKB(cudaMemcpyAsync(d_pos, positions, 3 * (size_t)n_vertices * sizeof(float), cudaMemcpyHostToDevice, st));
    CKB(cudaMemcpyAsync(d_tri, triangles, 3 * (size_t)nt * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    CKB(cudaMemcpyAsync(d_zero, zero_normal.data(), n_vertices, cudaMemcpyHostToDevice, st));
    if (tri_material) CKB(cudaMemcpyAsync(d_mat, tri_material, (size_t)nt * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    const uint32_t init_bounds[12] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    CKB(cudaMemcpyAsync(d_bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice, st));
    CKB(cudaMemsetAsync(d_leaf_count, 0, sizeof(uint32_t), st));
    CKB(cudaEventRecord(ev0, st));

    prepare_kernel<<<(nt + 255) / 256, 256, 0, st>>>(d_pos, d_tri, nt, d_tri_box, d_centroid, d_bounds);
    morton_kernel<<<(nt + 255) / 256, 256, 0, st>>>(d_centroid, nt, d_bounds, d_keys, d_ids);
    {
        size_t tmp_bytes = 0;
        CKB(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_keys, d_keys_sorted, d_ids, d_ids_sorted, (int)nt, 0, 63, st));
        void* tmp = nullptr;
        CKB(dev.alloc((unsigned char**)&tmp, tmp_bytes));
        CKB(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, d_keys, d_keys_sorted, d_ids, d_ids_sorted, (int)nt, 0, 63, st));
    }
    size_t scan_bytes = 0;
    CKB(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_cnt_inner, d_off_inner, (int)node_cap + 1, st));
    void* scan_tmp = nullptr;
    CKB(dev.alloc((unsigned char**)&scan_tmp, scan_bytes));

    uint32_t h_bounds[12];
    CKB(cudaMemcpyAsync(h_bounds, d_bounds, sizeof h_bounds, cudaMemcpyDeviceToHost, st));

    // ---- hierarchy, breadth-first ----
    std::vector<std::pair<uint32_t, uint32_t>> levels;  // (first node, count)
    uint32_t n_inner = 0, n_packets = 0, root = 0;
    if (nt <= LEAF_MAX) {  // the whole scene is one leaf (building.rs:109-113)
        const Leaf lf{0u, nt, 0u};
        CKB(cudaMemcpyAsync(d_leaves, &lf, sizeof lf, cudaMemcpyHostToDevice, st));
        const uint32_t one = 1;
        CKB(cudaMemcpyAsync(d_leaf_count, &one, sizeof one, cudaMemcpyHostToDevice, st));
        n_packets = (nt + 7) / 8;
        root = 0u << MP_LINK_COUNT_BITS | n_packets;
    } else {
        const Range all{0u, nt};
        CKB(cudaMemcpyAsync(d_node_range, &all, sizeof all, cudaMemcpyHostToDevice, st));
        uint32_t first = 0, count = 1;
        n_inner = 1;
        root = 0u;  // inner node 0
        while (count > 0) {
            levels.emplace_back(first, count);
            const unsigned blocks = (count + 127) / 128;
            split_kernel<<<blocks, 128, 0, st>>>(d_keys_sorted, d_node_range, first, count, LEAF_MAX, d_topo, d_cnt_inner, d_cnt_packets);
            // one extra element so that the last entry of the exclusive scan is the level's total
            CKB(cudaMemsetAsync(d_cnt_inner + count, 0, sizeof(uint32_t), st));
            CKB(cudaMemsetAsync(d_cnt_packets + count, 0, sizeof(uint32_t), st));
            CKB(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, d_cnt_inner, d_off_inner, (int)count + 1, st));
            CKB(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, d_cnt_packets, d_off_packets, (int)count + 1, st));
            uint32_t totals[2];
            CKB(cudaMemcpyAsync(&totals[0], d_off_inner + count, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CKB(cudaMemcpyAsync(&totals[1], d_off_packets + count, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CKB(cudaStreamSynchronize(st));
            if ((size_t)n_inner + totals[0] > node_cap) return mp::fail(MP_EINVAL, "device builder: node capacity exceeded");
            assign_kernel<<<blocks, 128, 0, st>>>(d_topo, d_node_range, first, count, LEAF_MAX, d_off_inner, d_off_packets, n_inner, n_packets, d_leaves, d_leaf_count,
                                                  d_parent_slot);
            first = n_inner, count = totals[0];
            n_inner += totals[0], n_packets += totals[1];
            if (levels.size() > 64) return mp::fail(MP_EINVAL, "device builder: hierarchy deeper than 64 levels");
        }
    }
    if (n_inner > MP_LINK_MAX_INDEX || n_packets > MP_LINK_MAX_INDEX) return mp::fail(MP_EINVAL, "scene exceeds the 29-bit link index range (mod.rs:71)");
    uint32_t n_leaves = 0;
    CKB(cudaMemcpyAsync(&n_leaves, d_leaf_count, sizeof n_leaves, cudaMemcpyDeviceToHost, st));
    CKB(cudaStreamSynchronize(st));

    // ---- boxes and records ----
    MpInnerNode* d_inner;
    MpBox *d_inner_box, *d_packet_box;
    MpTrianglePacket* d_packets;
    MpTriShading* d_shading;
    BoxF* d_packet_enclosing;
    uint32_t* d_lane_leaf;
    CKB(dev.alloc(&d_node_exact, n_inner));
    CKB(dev.alloc(&d_child_exact, (size_t)n_inner * 8));
    CKB(dev.alloc(&d_node_enclosing, n_inner));
    CKB(dev.alloc(&d_inner, n_inner));
    CKB(dev.alloc(&d_inner_box, n_inner));
    CKB(dev.alloc(&d_packets, n_packets));
    CKB(dev.alloc(&d_packet_box, n_packets));
    CKB(dev.alloc(&d_shading, (size_t)n_packets * 8));
    CKB(dev.alloc(&d_packet_enclosing, n_packets));
    CKB(dev.alloc(&d_lane_leaf, n_packets));
    BoxF scene;
    for (int a = 0; a < 3; ++a) {
        const uint32_t lo = h_bounds[a], hi = h_bounds[3 + a];
        scene.lo[a] = (lo & 0x80000000u) ? [&] { uint32_t u = lo & 0x7fffffffu; float f; memcpy(&f, &u, 4); return f; }() : [&] { uint32_t u = ~lo; float f; memcpy(&f, &u, 4); return f; }();
        scene.hi[a] = (hi & 0x80000000u) ? [&] { uint32_t u = hi & 0x7fffffffu; float f; memcpy(&f, &u, 4); return f; }() : [&] { uint32_t u = ~hi; float f; memcpy(&f, &u, 4); return f; }();
    }
    for (size_t l = levels.size(); l-- > 0;) {  // exact boxes, deepest level first
        const uint32_t first = levels[l].first, count = levels[l].second;
        exact_kernel<<<(count * 8 + 127) / 128, 128, 0, st>>>(d_topo, first, count, d_tri_box, d_ids_sorted, d_node_exact, d_child_exact);
    }
    if (n_inner) {
        CKB(cudaMemcpyAsync(d_node_enclosing, &scene, sizeof scene, cudaMemcpyHostToDevice, st));  // the root is built against the exact scene box (building.rs:84-85)
        for (size_t l = 0; l < levels.size(); ++l) {
            const uint32_t first = levels[l].first, count = levels[l].second;
            quantise_kernel<<<(count * 8 + 127) / 128, 128, 0, st>>>(d_topo, first, count, d_child_exact, d_node_enclosing, d_packet_enclosing, d_inner, d_inner_box);
        }
    } else {
        std::vector<BoxF> enc(n_packets, scene);
        CKB(cudaMemcpyAsync(d_packet_enclosing, enc.data(), enc.size() * sizeof(BoxF), cudaMemcpyHostToDevice, st));
        CKB(cudaStreamSynchronize(st));
    }
    lane_leaf_kernel<<<(n_leaves + 255) / 256, 256, 0, st>>>(d_leaves, n_leaves, d_lane_leaf);
    leaf_kernel<<<(n_packets * 8 + 255) / 256, 256, 0, st>>>(d_leaves, n_leaves, d_lane_leaf, n_packets * 8, d_ids_sorted, d_pos, d_tri, d_mat, d_zero, d_packet_enclosing,
                                                              d_packets, d_packet_box, d_shading);
    CKB(cudaEventRecord(ev1, st));

    std::vector<MpInnerNode> inner(n_inner);
    std::vector<MpBox> inner_box(n_inner), packet_box(n_packets);
    std::vector<MpTrianglePacket> packets(n_packets);
    std::vector<MpTriShading> shading((size_t)n_packets * 8);
    if (n_inner) {
        CKB(cudaMemcpyAsync(inner.data(), d_inner, (size_t)n_inner * sizeof(MpInnerNode), cudaMemcpyDeviceToHost, st));
        CKB(cudaMemcpyAsync(inner_box.data(), d_inner_box, (size_t)n_inner * sizeof(MpBox), cudaMemcpyDeviceToHost, st));
    }
    CKB(cudaMemcpyAsync(packets.data(), d_packets, (size_t)n_packets * sizeof(MpTrianglePacket), cudaMemcpyDeviceToHost, st));
    CKB(cudaMemcpyAsync(packet_box.data(), d_packet_box, (size_t)n_packets * sizeof(MpBox), cudaMemcpyDeviceToHost, st));
    CKB(cudaMemcpyAsync(shading.data(), d_shading, shading.size() * sizeof(MpTriShading), cudaMemcpyDeviceToHost, st));
    CKB(cudaStreamSynchronize(st));
    CKB(cudaGetLastError());
    if (out_device_ms) CKB(cudaEventElapsedTime(out_device_ms, ev0, ev1));

    std::vector<MpMaterial> mats(materials, materials + (materials ? n_materials : 0));
    *out = mp::adopt_flat_bvh(scene.lo, scene.hi, root, nt, inner, inner_box, packets, packet_box, shading, vertices, mats);
    return MP_OK;
}

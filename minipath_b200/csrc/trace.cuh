// BVH8 traversal + ray/triangle intersection on the device, one thread per ray, in the reference's exact
// visiting order (reference src/scene/triangle_bvh/ray_bvh_intersection.rs:26-64,104-162;
// src/geometry/aabb.rs:257-283; src/geometry/triangle.rs:187-216).
//
// Layout read here (all 16-byte aligned, fetched with 128-bit loads through the read-only path):
//   inner[i]      8 x uint4 : 6 x u16x8 child bounds (min.x,min.y,min.z,max.x,max.y,max.z) + 8 links
//   inner_box[i]  2 x float4: enclosing box (min, size) node i is dequantised against
//   packets[p]    9 x uint4 : v0.x,v0.y,v0.z,v1.x,...,v2.z as u16x8
//   packet_box[p] 2 x float4: enclosing box of the leaf that owns packet p
// The reference carries each child's box on its stack (32-byte entries); the box of a node only depends on its
// parent's box and quantised bounds, so it is precomputed on the host with the same fma and the device stack
// entry shrinks to (link, t1) = 8 bytes.  Bit-neutral.
#pragma once
#include "device_math.cuh"

namespace mpd {

#define MPD_STACK_MAX 128        // thread-per-ray variant: local-memory stack
#define MPD_OCTET_STACK_MAX 384  // octet variant: shared-memory stack, 16 octets x cap x 8 B per block

struct Scene {
    const uint4* __restrict__ inner;
    const float4* __restrict__ inner_box;
    const uint4* __restrict__ packets;
    const float4* __restrict__ packet_box;
    const uint4* __restrict__ tri_shading;
    const float4* __restrict__ vertices;
    const uint4* __restrict__ inner_rec;   // quantised node records: 6 bound rows, 2 link rows, box (min, size) = 160 B = 5 x LDG.256
    const uint4* __restrict__ packets_rec; // quantised packet records: 9 rows, leaf box, pad = 192 B = 6 x LDG.256
    const float4* __restrict__ inner_w;    // decoded nodes (MP_LAYOUT_DECODED), MPD_NODE_W float4 each
    const float4* __restrict__ packets_w;  // decoded packets, MPD_PACK_W float4 each
    uint32_t root;
    float eps;  // extension integrator: origin offset, 1e-4 * scene diagonal
    const float4* __restrict__ materials;  // 2 x float4 per MpMaterial, or null (no table: constant grey albedo)
    uint32_t n_materials;
    uint32_t minmax_ok;  // every real child box has min <= max after dequantisation (checked by decode_inner_kernel at upload)
};

struct Hit {  // LeafHitRecord (ray_bvh_intersection.rs:165-171); the geometric normal is recomputed at finalise
    float t, u, v;
    uint32_t tri;
};

struct Visits {
    uint32_t inner, packets, leaves;
};

// One child lane of InnerNode::intersect: dequantise, slab test clamped to [0, max_t].
// NANPATH is taken only when some inv_direction component is +inf (a zero direction component): only then can
// (b - o) * inv be 0 * inf = NaN, which aabb.rs:262-267 replaces by -inf / +inf.  Without NaNs, vminps/vmaxps equal
// fminf/fmaxf up to the sign of a zero, which no comparison below can observe.
template <bool NANPATH>
__device__ __forceinline__ bool child_slab(const uint4 (&q)[6], int lane, const F3& bmin, const F3& bsize, const Ray& r,
                                           float max_t, float& t1_out) {
    float lox = __fmul_rn(__fsub_rn(dequant(q[0], lane, bsize.x, bmin.x), r.o.x), r.inv.x);
    float loy = __fmul_rn(__fsub_rn(dequant(q[1], lane, bsize.y, bmin.y), r.o.y), r.inv.y);
    float loz = __fmul_rn(__fsub_rn(dequant(q[2], lane, bsize.z, bmin.z), r.o.z), r.inv.z);
    float hix = __fmul_rn(__fsub_rn(dequant(q[3], lane, bsize.x, bmin.x), r.o.x), r.inv.x);
    float hiy = __fmul_rn(__fsub_rn(dequant(q[4], lane, bsize.y, bmin.y), r.o.y), r.inv.y);
    float hiz = __fmul_rn(__fsub_rn(dequant(q[5], lane, bsize.z, bmin.z), r.o.z), r.inv.z);
    float t1, t2;
    if (NANPATH) {
        if (lox != lox) lox = MPD_NINF;
        if (loy != loy) loy = MPD_NINF;
        if (loz != loz) loz = MPD_NINF;
        if (hix != hix) hix = MPD_INF;
        if (hiy != hiy) hiy = MPD_INF;
        if (hiz != hiz) hiz = MPD_INF;
        float cminx = minps(lox, hix), cminy = minps(loy, hiy), cminz = minps(loz, hiz);
        float cmaxx = maxps(lox, hix), cmaxy = maxps(loy, hiy), cmaxz = maxps(loz, hiz);
        t1 = maxps(maxps(cminx, 0.0f), maxps(cminy, cminz));
        t2 = minps(minps(cmaxx, max_t), minps(cmaxy, cmaxz));
    } else {
        float cminx = fminf(lox, hix), cminy = fminf(loy, hiy), cminz = fminf(loz, hiz);
        float cmaxx = fmaxf(lox, hix), cmaxy = fmaxf(loy, hiy), cmaxz = fmaxf(loz, hiz);
        t1 = fmaxf(fmaxf(cminx, 0.0f), fmaxf(cminy, cminz));
        t2 = fminf(fminf(cmaxx, max_t), fminf(cmaxy, cmaxz));
    }
    t1_out = t1;
    return t1 <= t2;
}

// One lane of Triangle<WorldPoint8>::intersect + the t-range mask of intersect_triangles
// (triangle.rs:187-216; ray_bvh_intersection.rs:125).
__device__ __forceinline__ bool tri_lane(const uint4 (&q)[9], int lane, const F3& bmin, const F3& bsize, const Ray& r,
                                         float max_t, float& t_out, float& u_out, float& v_out) {
    F3 v0 = f3(dequant(q[0], lane, bsize.x, bmin.x), dequant(q[1], lane, bsize.y, bmin.y), dequant(q[2], lane, bsize.z, bmin.z));
    F3 v1 = f3(dequant(q[3], lane, bsize.x, bmin.x), dequant(q[4], lane, bsize.y, bmin.y), dequant(q[5], lane, bsize.z, bmin.z));
    F3 v2 = f3(dequant(q[6], lane, bsize.x, bmin.x), dequant(q[7], lane, bsize.y, bmin.y), dequant(q[8], lane, bsize.z, bmin.z));
    F3 e1 = v1 - v0, e2 = v2 - v0;
    F3 p = fma_cross(r.d, e2);
    float det = fma_dot(e1, p);
    float inv = __frcp_rn(det);  // = 1.0f / det correctly rounded (triangle.rs:203); may be infinite, no epsilon test in the reference
    F3 s = r.o - v0;
    float u = __fmul_rn(inv, fma_dot(s, p));
    F3 qv = fma_cross(s, e1);
    float v = __fmul_rn(inv, fma_dot(r.d, qv));
    float t = __fmul_rn(inv, fma_dot(e2, qv));
    t_out = t, u_out = u, v_out = v;
    return u >= 0.0f && v >= 0.0f && __fadd_rn(u, v) <= 1.0f && t >= 0.0f && t <= max_t;
}

// ---- MP_LAYOUT_DECODED: the same nodes with the u16 -> f32 dequantisation hoisted out of the traversal ----
// The traversal kernels are bound by instruction issue, not by bytes (profiles/r1_wf_trace_thread.txt: 70 % issue slots busy,
// L2 at 4 %), and 18 of the ~46 instructions a child costs are its six dequantisations (PRMT + 2 FFMA each).  Every child
// coordinate only depends on the node's enclosing box and its quantised bounds, so `decode_scene_kernel` evaluates
// fma(size, float(q) * K, min) ONCE per scene with the same device function and stores the f32 result; the traversal then reads
//   inner_w[i]   16 x float4: min.x[8] min.y[8] min.z[8] max.x[8] max.y[8] max.z[8] links[8] pad[8]          (256 B)
//   packets_w[p] 18 x float4: v0.x[8] v0.y[8] v0.z[8] e1.x[8] e1.y[8] e1.z[8] e2.x[8] e2.y[8] e2.z[8]        (288 B)
// with e1 = v1 - v0, e2 = v2 - v0 (the unfused subtractions of triangle.rs:193-194).  Same operations on the same operands
// → the same bits as decoding per visit.  NULL children are stored as min = +inf, max = -inf so that they fail the slab test.
#define MPD_NODE_W 16
#define MPD_PACK_W 18

// 256-bit read-only load (LDG.E.256, sm_100+): a divergent gather costs one L1 wavefront per lane and INSTRUCTION whatever its
// width (profiles/r1b_ubench_l1_patterns.txt: 10 x LDG.128 = 9.3 clk per node and SM, 5 x LDG.256 = 4.9), so the quantised
// records are read 32 bytes at a time.
// MPD_LD_NODE / MPD_LD_PACKET pick the L1 eviction policy of the record loads (0 default, 1 evict_last, 2 no_allocate,
// 3 evict_first): an A/B switch for scripts/build_variant.sh, measured in profiles/r2c_variants_pool_levers.txt.
#ifndef MPD_LD_NODE
#define MPD_LD_NODE 0
#endif
#ifndef MPD_LD_PACKET
#define MPD_LD_PACKET 2  // packets are streamed once per leaf visit: keeping them out of L1 leaves it to the node records (+2.0 % on the stand-in, teapot unchanged)
#endif
template <int POL>
__device__ __forceinline__ void ldg256p(const uint4* p, uint4& a, uint4& b) {
#define MPD_LDG256(Q)                                                                                   \
    asm("ld.global.nc" Q ".v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"                                     \
        : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)        \
        : "l"(p))
    if (POL == 1) MPD_LDG256(".L1::evict_last");
    else if (POL == 2) MPD_LDG256(".L1::no_allocate");
    else if (POL == 3) MPD_LDG256(".L1::evict_first");
    else MPD_LDG256("");
#undef MPD_LDG256
}
__device__ __forceinline__ void ldg256(const uint4* p, uint4& a, uint4& b) { ldg256p<0>(p, a, b); }
__device__ __forceinline__ void ldg256n(const uint4* p, uint4& a, uint4& b) { ldg256p<MPD_LD_NODE>(p, a, b); }
__device__ __forceinline__ void ldg256k(const uint4* p, uint4& a, uint4& b) { ldg256p<MPD_LD_PACKET>(p, a, b); }
__device__ __forceinline__ float4 as_f4(const uint4& v) { return make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w)); }

// A ray whose inverse direction is finite and non-NaN crosses the planes of a child box in a known order: (b - o) * inv is
// monotonic in b, so min(lo, hi) is the plane on the side the ray comes from.  Decoded nodes keep min <= max per axis
// (checked by decode_scene_kernel), hence the near / far rows can be picked per RAY instead of per child with 6 min/max.
struct RaySel {
    uint32_t nx, ny, nz, fx, fy, fz;  // float4 row offsets inside a decoded node
};
__device__ __forceinline__ RaySel ray_sel(const Ray& r) {
    RaySel s;
    const bool sx = r.inv.x < 0.0f, sy = r.inv.y < 0.0f, sz = r.inv.z < 0.0f;
    s.nx = sx ? 6u : 0u, s.fx = sx ? 0u : 6u;
    s.ny = sy ? 8u : 2u, s.fy = sy ? 2u : 8u;
    s.nz = sz ? 10u : 4u, s.fz = sz ? 4u : 10u;
    return s;
}
// Rays that need the reference's NaN handling (aabb.rs:262-267) or are not finite at all take the general path.
// So do all rays of a scene whose child boxes are not all min <= max (never the case for a builder-made BVH): the fast path picks
// the near / far planes per ray, which presumes ordered boxes.
__device__ __forceinline__ bool ray_needs_nanpath(const Scene& sc, const Ray& r) {
    return !sc.minmax_ok | !(fabsf(r.inv.x) < MPD_INF) | !(fabsf(r.inv.y) < MPD_INF) | !(fabsf(r.inv.z) < MPD_INF) | !(fabsf(r.o.x) < MPD_INF) |
           !(fabsf(r.o.y) < MPD_INF) | !(fabsf(r.o.z) < MPD_INF);
}

__device__ __forceinline__ float comp(const float4& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w)); }
__device__ __forceinline__ uint32_t comp(const uint4& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w)); }

// Stack views: thread-per-ray kernels index a private array, the pool kernel interleaves the stacks of a warp's rays.
struct LocalStack {
    uint2* p;
    __device__ __forceinline__ uint2& at(int i) const { return p[i]; }
};

// Where the hit children of a node go.  StackSink pushes every one (lanes ascending, so lane 7 pops first, :161; NEAR_FIRST keeps
// the node's entries sorted far → near).  The pool kernel has its own sink (pool.cuh: PoolDeferSink) that keeps the LAST hit child
// in registers instead of pushing it: that entry would be popped next and always survives `t1 > best.t` (:40) because
// t1 <= t2 <= max_t = best.t, so handing it over directly saves a stack store and a load per inner node, same visiting order.
template <int ORDER, class S>
struct StackSink {
    S stk;
    int base, sp;
    __device__ __forceinline__ void hit(uint32_t link, float t1) {
        if (ORDER == MP_ORDER_NEAR_FIRST) {
            int pos = sp;  // ties: lower lane nearer the top
            while (pos > base && __uint_as_float(stk.at(pos - 1).y) <= t1) {
                stk.at(pos) = stk.at(pos - 1);
                --pos;
            }
            stk.at(pos) = make_uint2(link, __float_as_uint(t1));
            ++sp;
        } else {
            stk.at(sp++) = make_uint2(link, __float_as_uint(t1));
        }
    }
};
// InnerNode::intersect on a quantised node record (ray_bvh_intersection.rs:149-162).
//
// MPD_INNER_SPLIT (default): the slab tests of the eight children are ONE straight-line block that ends in a hit mask, and the
// hit children are handed to the sink afterwards, lanes ascending.  The first version tested and pushed child by child: a NULL
// test, a branch around the test and a branch around the push per child — the compiler then re-materialised the dequantisation
// constants behind every branch (6 of the 58 instructions a child cost, profiles/r2_sass_pool_inner_before_after.txt) and could not
// overlap the dependent chains of neighbouring children.  Same operations on the same operands, so the same bits.
// MPD_QSEL: on the NaN-free path the near / far plane of every axis is known per RAY (sign of inv; needs min <= max per child,
// Scene::minmax_ok), so the near and far rows are picked once per node with 24 selects instead of ordering lo / hi per child
// with 48 min / max — what the decoded layout does with row offsets (RaySel).
#ifndef MPD_INNER_SPLIT
#define MPD_INNER_SPLIT 1
#endif
#ifndef MPD_QSEL
#define MPD_QSEL 1
#endif
__device__ __forceinline__ uint4 sel4(bool c, const uint4& a, const uint4& b) { return make_uint4(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z, c ? a.w : b.w); }

// Slab tests of the eight children of a quantised node record: returns the hit mask (bit = child lane), fills links and t1.
template <bool NANPATH>
__device__ __forceinline__ uint32_t inner_quantised_hits(const Scene& sc, const Ray& r, uint32_t index, float max_t, uint32_t (&links)[8], float (&t1s)[8]) {
    const uint4* n = sc.inner_rec + 10 * (size_t)index;
    uint4 q[6], la, lb, b0, b1;
    ldg256n(n + 0, q[0], q[1]);
    ldg256n(n + 2, q[2], q[3]);
    ldg256n(n + 4, q[4], q[5]);
    ldg256n(n + 6, la, lb);
    ldg256n(n + 8, b0, b1);
    const F3 bmin = f3(__uint_as_float(b0.x), __uint_as_float(b0.y), __uint_as_float(b0.z));
    const F3 bsize = f3(__uint_as_float(b1.x), __uint_as_float(b1.y), __uint_as_float(b1.z));
    links[0] = la.x, links[1] = la.y, links[2] = la.z, links[3] = la.w, links[4] = lb.x, links[5] = lb.y, links[6] = lb.z, links[7] = lb.w;
    uint32_t mask = 0u;
    if (!NANPATH && MPD_QSEL) {
        const bool sx = r.inv.x < 0.0f, sy = r.inv.y < 0.0f, sz = r.inv.z < 0.0f;
        const uint4 nx = sel4(sx, q[3], q[0]), fx = sel4(sx, q[0], q[3]);
        const uint4 ny = sel4(sy, q[4], q[1]), fy = sel4(sy, q[1], q[4]);
        const uint4 nz = sel4(sz, q[5], q[2]), fz = sel4(sz, q[2], q[5]);
#pragma unroll
        for (int lane = 0; lane < 8; ++lane) {
            const float lox = __fmul_rn(__fsub_rn(dequant(nx, lane, bsize.x, bmin.x), r.o.x), r.inv.x);
            const float loy = __fmul_rn(__fsub_rn(dequant(ny, lane, bsize.y, bmin.y), r.o.y), r.inv.y);
            const float loz = __fmul_rn(__fsub_rn(dequant(nz, lane, bsize.z, bmin.z), r.o.z), r.inv.z);
            const float hix = __fmul_rn(__fsub_rn(dequant(fx, lane, bsize.x, bmin.x), r.o.x), r.inv.x);
            const float hiy = __fmul_rn(__fsub_rn(dequant(fy, lane, bsize.y, bmin.y), r.o.y), r.inv.y);
            const float hiz = __fmul_rn(__fsub_rn(dequant(fz, lane, bsize.z, bmin.z), r.o.z), r.inv.z);
            const float t1 = fmaxf(fmaxf(lox, loy), fmaxf(loz, 0.0f));  // FMNMX3 + FMNMX
            const float t2 = fminf(fminf(hix, hiy), fminf(hiz, max_t));
            t1s[lane] = t1;
            mask |= (t1 <= t2 && links[lane] != MP_LINK_NULL) ? (1u << lane) : 0u;
        }
    } else {
#pragma unroll
        for (int lane = 0; lane < 8; ++lane) {
            float t1;
            const bool hit = child_slab<NANPATH>(q, lane, bmin, bsize, r, max_t, t1);
            t1s[lane] = t1;
            mask |= (hit && links[lane] != MP_LINK_NULL) ? (1u << lane) : 0u;  // NULL: all-zero bounds, discarded when popped (:49); skipping is neutral
        }
    }
    return mask;
}

template <bool NANPATH, class Sink>
__device__ __forceinline__ void inner_quantised(const Scene& sc, const Ray& r, uint32_t index, float max_t, Sink& sink) {
#if MPD_INNER_SPLIT
    uint32_t links[8];
    float t1s[8];
    const uint32_t mask = inner_quantised_hits<NANPATH>(sc, r, index, max_t, links, t1s);
#pragma unroll
    for (int lane = 0; lane < 8; ++lane)
        if (mask & (1u << lane)) sink.hit(links[lane], t1s[lane]);
#else
    const uint4* n = sc.inner_rec + 10 * (size_t)index;
    uint4 q[6], la, lb, b0, b1;
    ldg256n(n + 0, q[0], q[1]);
    ldg256n(n + 2, q[2], q[3]);
    ldg256n(n + 4, q[4], q[5]);
    ldg256n(n + 6, la, lb);
    ldg256n(n + 8, b0, b1);
    const F3 bmin = f3(__uint_as_float(b0.x), __uint_as_float(b0.y), __uint_as_float(b0.z));
    const F3 bsize = f3(__uint_as_float(b1.x), __uint_as_float(b1.y), __uint_as_float(b1.z));
    const uint32_t links[8] = {la.x, la.y, la.z, la.w, lb.x, lb.y, lb.z, lb.w};
#pragma unroll
    for (int lane = 0; lane < 8; ++lane) {
        if (links[lane] == MP_LINK_NULL) continue;  // all-zero bounds, discarded when popped (:49); skipping is neutral
        float t1;
        if (child_slab<NANPATH>(q, lane, bmin, bsize, r, max_t, t1)) sink.hit(links[lane], t1);
    }
#endif
}

// One quantised packet record: RelativeTriangle8::decompress + Triangle<WorldPoint8>::intersect (ray_bvh_intersection.rs:104-141).
__device__ __forceinline__ void packet_quantised(const Scene& sc, const Ray& r, uint32_t pidx, float max_t, Hit& local) {
    const uint4* pk = sc.packets_rec + 12 * (size_t)pidx;
    uint4 q[9], b0, b1, pad;
    ldg256k(pk + 0, q[0], q[1]);
    ldg256k(pk + 2, q[2], q[3]);
    ldg256k(pk + 4, q[4], q[5]);
    ldg256k(pk + 6, q[6], q[7]);
    ldg256k(pk + 8, q[8], b0);
    ldg256k(pk + 10, b1, pad);
    const F3 bmin = f3(__uint_as_float(b0.x), __uint_as_float(b0.y), __uint_as_float(b0.z));
    const F3 bsize = f3(__uint_as_float(b1.x), __uint_as_float(b1.y), __uint_as_float(b1.z));
#pragma unroll
    for (int lane = 0; lane < 8; ++lane) {
        float t, u, v;
        if (tri_lane(q, lane, bmin, bsize, r, max_t, t, u, v) && t < local.t) {
            local.t = t, local.u = u, local.v = v;
            local.tri = pidx * 8u + (uint32_t)lane;
        }
    }
}

// InnerNode::intersect on a decoded node (ray_bvh_intersection.rs:149-162).
template <bool NANPATH, class Sink>
__device__ __forceinline__ void inner_decoded(const Scene& sc, const Ray& r, const RaySel& sel, uint32_t index, float max_t, Sink& sink) {
    const float4* n = sc.inner_w + MPD_NODE_W * (size_t)index;
#pragma unroll
    for (int h = 0; h < 2; ++h) {  // children 4h .. 4h+3
        const uint4 lk = __ldg(reinterpret_cast<const uint4*>(n + 12 + h));
        if (NANPATH) {
            const float4 ax = __ldg(n + 0 + h), ay = __ldg(n + 2 + h), az = __ldg(n + 4 + h);
            const float4 bx = __ldg(n + 6 + h), by = __ldg(n + 8 + h), bz = __ldg(n + 10 + h);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (comp(lk, k) == MP_LINK_NULL) continue;
                float lox = __fmul_rn(__fsub_rn(comp(ax, k), r.o.x), r.inv.x), loy = __fmul_rn(__fsub_rn(comp(ay, k), r.o.y), r.inv.y);
                float loz = __fmul_rn(__fsub_rn(comp(az, k), r.o.z), r.inv.z), hix = __fmul_rn(__fsub_rn(comp(bx, k), r.o.x), r.inv.x);
                float hiy = __fmul_rn(__fsub_rn(comp(by, k), r.o.y), r.inv.y), hiz = __fmul_rn(__fsub_rn(comp(bz, k), r.o.z), r.inv.z);
                if (lox != lox) lox = MPD_NINF;
                if (loy != loy) loy = MPD_NINF;
                if (loz != loz) loz = MPD_NINF;
                if (hix != hix) hix = MPD_INF;
                if (hiy != hiy) hiy = MPD_INF;
                if (hiz != hiz) hiz = MPD_INF;
                const float cminx = minps(lox, hix), cminy = minps(loy, hiy), cminz = minps(loz, hiz);
                const float cmaxx = maxps(lox, hix), cmaxy = maxps(loy, hiy), cmaxz = maxps(loz, hiz);
                const float t1 = maxps(maxps(cminx, 0.0f), maxps(cminy, cminz));
                const float t2 = minps(minps(cmaxx, max_t), minps(cmaxy, cmaxz));
                if (t1 <= t2) sink.hit(comp(lk, k), t1);
            }
        } else {
            const float4 ax = __ldg(n + sel.nx + h), ay = __ldg(n + sel.ny + h), az = __ldg(n + sel.nz + h);
            const float4 bx = __ldg(n + sel.fx + h), by = __ldg(n + sel.fy + h), bz = __ldg(n + sel.fz + h);
#if MPD_INNER_SPLIT
            float t1s[4];
            uint32_t mask = 0u;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float lox = __fmul_rn(__fsub_rn(comp(ax, k), r.o.x), r.inv.x), loy = __fmul_rn(__fsub_rn(comp(ay, k), r.o.y), r.inv.y);
                const float loz = __fmul_rn(__fsub_rn(comp(az, k), r.o.z), r.inv.z), hix = __fmul_rn(__fsub_rn(comp(bx, k), r.o.x), r.inv.x);
                const float hiy = __fmul_rn(__fsub_rn(comp(by, k), r.o.y), r.inv.y), hiz = __fmul_rn(__fsub_rn(comp(bz, k), r.o.z), r.inv.z);
                const float t1 = fmaxf(fmaxf(lox, loy), fmaxf(loz, 0.0f));    // FMNMX3 + FMNMX
                const float t2 = fminf(fminf(hix, hiy), fminf(hiz, max_t));
                t1s[k] = t1;
                mask |= t1 <= t2 ? (1u << k) : 0u;  // NULL children: t1 = +inf, t2 = -inf
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (mask & (1u << k)) sink.hit(comp(lk, k), t1s[k]);
#else
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float lox = __fmul_rn(__fsub_rn(comp(ax, k), r.o.x), r.inv.x), loy = __fmul_rn(__fsub_rn(comp(ay, k), r.o.y), r.inv.y);
                const float loz = __fmul_rn(__fsub_rn(comp(az, k), r.o.z), r.inv.z), hix = __fmul_rn(__fsub_rn(comp(bx, k), r.o.x), r.inv.x);
                const float hiy = __fmul_rn(__fsub_rn(comp(by, k), r.o.y), r.inv.y), hiz = __fmul_rn(__fsub_rn(comp(bz, k), r.o.z), r.inv.z);
                const float t1 = fmaxf(fmaxf(lox, loy), fmaxf(loz, 0.0f));    // FMNMX3 + FMNMX
                const float t2 = fminf(fminf(hix, hiy), fminf(hiz, max_t));
                if (t1 <= t2) sink.hit(comp(lk, k), t1);  // NULL children: t1 = +inf, t2 = -inf
            }
#endif
        }
    }
}

// Decoded node read with 256-bit loads (NaN-free rays only): a row pair (children 0-3, 4-7) of one plane is 32 contiguous bytes,
// so the six near / far rows and the links are 7 x LDG.256 instead of 14 x LDG.128 — for incoherent gathers the request count is
// what the L1 pipe charges (profiles/r1b_ubench_l1_patterns.txt).  Used by the pool kernels' MIXED layout: decoded inner nodes
// (a few thousand nodes, cache-resident) with quantised triangle packets.
__device__ __forceinline__ uint32_t inner_decoded256_hits(const Scene& sc, const Ray& r, const RaySel& sel, uint32_t index, float max_t, uint32_t (&links)[8], float (&t1s)[8]) {
    const uint4* n = reinterpret_cast<const uint4*>(sc.inner_w + MPD_NODE_W * (size_t)index);
    uint4 ax[2], ay[2], az[2], bx[2], by[2], bz[2], lk[2];
    ldg256n(n + sel.nx, ax[0], ax[1]);
    ldg256n(n + sel.ny, ay[0], ay[1]);
    ldg256n(n + sel.nz, az[0], az[1]);
    ldg256n(n + sel.fx, bx[0], bx[1]);
    ldg256n(n + sel.fy, by[0], by[1]);
    ldg256n(n + sel.fz, bz[0], bz[1]);
    ldg256n(n + 12, lk[0], lk[1]);
    uint32_t mask = 0u;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int h = c >> 2, k = c & 3;
        const float lox = __fmul_rn(__fsub_rn(__uint_as_float(comp(ax[h], k)), r.o.x), r.inv.x), loy = __fmul_rn(__fsub_rn(__uint_as_float(comp(ay[h], k)), r.o.y), r.inv.y);
        const float loz = __fmul_rn(__fsub_rn(__uint_as_float(comp(az[h], k)), r.o.z), r.inv.z), hix = __fmul_rn(__fsub_rn(__uint_as_float(comp(bx[h], k)), r.o.x), r.inv.x);
        const float hiy = __fmul_rn(__fsub_rn(__uint_as_float(comp(by[h], k)), r.o.y), r.inv.y), hiz = __fmul_rn(__fsub_rn(__uint_as_float(comp(bz[h], k)), r.o.z), r.inv.z);
        const float t1 = fmaxf(fmaxf(lox, loy), fmaxf(loz, 0.0f));
        const float t2 = fminf(fminf(hix, hiy), fminf(hiz, max_t));
        t1s[c] = t1;
        links[c] = comp(lk[h], k);
        mask |= t1 <= t2 ? (1u << c) : 0u;  // NULL children: t1 = +inf, t2 = -inf
    }
    return mask;
}
template <class Sink>
__device__ __forceinline__ void inner_decoded256(const Scene& sc, const Ray& r, const RaySel& sel, uint32_t index, float max_t, Sink& sink) {
    uint32_t links[8];
    float t1s[8];
    const uint32_t mask = inner_decoded256_hits(sc, r, sel, index, max_t, links, t1s);
#pragma unroll
    for (int c = 0; c < 8; ++c)
        if (mask & (1u << c)) sink.hit(links[c], t1s[c]);
}

// One decoded packet: Triangle<WorldPoint8>::intersect + the t-range mask (triangle.rs:187-216; ray_bvh_intersection.rs:125-134).
__device__ __forceinline__ void packet_decoded(const Scene& sc, const Ray& r, uint32_t pidx, float max_t, Hit& local) {
    const float4* pk = sc.packets_w + MPD_PACK_W * (size_t)pidx;
#pragma unroll
    for (int h = 0; h < 2; ++h) {  // triangles 4h .. 4h+3
        float4 c[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) c[k] = __ldg(pk + 2 * k + h);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const F3 v0 = f3(comp(c[0], k), comp(c[1], k), comp(c[2], k));
            const F3 e1 = f3(comp(c[3], k), comp(c[4], k), comp(c[5], k));
            const F3 e2 = f3(comp(c[6], k), comp(c[7], k), comp(c[8], k));
            const F3 p = fma_cross(r.d, e2);
            const float det = fma_dot(e1, p);
            const float inv = __frcp_rn(det);  // correctly rounded 1/det = the reference's division; may be infinite
            const F3 s = r.o - v0;
            const float u = __fmul_rn(inv, fma_dot(s, p));
            const F3 qv = fma_cross(s, e1);
            const float v = __fmul_rn(inv, fma_dot(r.d, qv));
            const float t = __fmul_rn(inv, fma_dot(e2, qv));
            if (u >= 0.0f && v >= 0.0f && __fadd_rn(u, v) <= 1.0f && t >= 0.0f && t <= max_t && t < local.t) {
                local.t = t, local.u = u, local.v = v;
                local.tri = pidx * 8u + (uint32_t)(4 * h + k);
            }
        }
    }
}

// One inner node / one packet for one ray in either layout.
template <int LAYOUT, bool NANPATH, class Sink>
__device__ __forceinline__ void visit_inner(const Scene& sc, const Ray& r, const RaySel& sel, uint32_t index, float max_t, Sink& sink) {
    if (LAYOUT == MP_LAYOUT_DECODED)
        inner_decoded<NANPATH>(sc, r, sel, index, max_t, sink);
    else
        inner_quantised<NANPATH>(sc, r, index, max_t, sink);
}
template <int LAYOUT>
__device__ __forceinline__ void visit_packet(const Scene& sc, const Ray& r, uint32_t pidx, float max_t, Hit& local) {
    if (LAYOUT == MP_LAYOUT_DECODED)
        packet_decoded(sc, r, pidx, max_t, local);
    else
        packet_quantised(sc, r, pidx, max_t, local);
}

template <int LAYOUT, int ORDER, bool COUNT, bool NANPATH>
__device__ __forceinline__ Hit trace_impl(const Scene& sc, const Ray& r, bool any_hit, Visits& vis) {
    uint2 stk[MPD_STACK_MAX];  // (link, t1 bits)
    stk[0] = make_uint2(sc.root, __float_as_uint(MPD_NINF));
    StackSink<ORDER, LocalStack> sink{LocalStack{stk}, 0, 1};
    Hit best{MPD_FLT_MAX, 0.0f, 0.0f, MP_NO_HIT};
    RaySel sel{};
    if (LAYOUT == MP_LAYOUT_DECODED && !NANPATH) sel = ray_sel(r);

    while (sink.sp > 0) {
        const uint2 e = stk[--sink.sp];
        if (__uint_as_float(e.y) > best.t) continue;  // ray_bvh_intersection.rs:40
        const uint32_t count = e.x & MP_LINK_COUNT_MASK, index = e.x >> MP_LINK_COUNT_BITS;
        if (count == 0) {
            // ---- inner node (ray_bvh_intersection.rs:149-162) ----
            if (COUNT) vis.inner++;
            sink.base = sink.sp;
            visit_inner<LAYOUT, NANPATH>(sc, r, sel, index, best.t, sink);
        } else {
            // ---- leaf: ≤7 packets of 8 triangles (ray_bvh_intersection.rs:104-141) ----
            if (COUNT) vis.leaves++, vis.packets += count;
            const float max_t = best.t;  // not updated inside the leaf
            Hit local{MPD_INF, 0.0f, 0.0f, MP_NO_HIT};
            for (uint32_t j = 0; j < count; ++j) {
                visit_packet<LAYOUT>(sc, r, index + j, max_t, local);
                if (any_hit && local.t < max_t) {  // shadow rays (extension): occluded, the remaining packets cannot change that
                    if (COUNT) vis.packets -= count - 1u - j;
                    break;
                }
            }
            if (local.t < best.t) {  // :59
                best = local;
                if (any_hit) break;
            }
        }
    }
    return best;
}

template <int LAYOUT, int ORDER, bool COUNT>
__device__ __forceinline__ Hit trace(const Scene& sc, const Ray& r, bool any_hit, Visits& vis) {
    // |inv| is infinite for a zero (or denormal) direction component
    if (ray_needs_nanpath(sc, r)) return trace_impl<LAYOUT, ORDER, COUNT, true>(sc, r, any_hit, vis);
    return trace_impl<LAYOUT, ORDER, COUNT, false>(sc, r, any_hit, vis);
}

// Hit finalisation (ray_bvh_intersection.rs:66-95): shading normal, normalised.
__device__ __forceinline__ F3 shading_normal(const Scene& sc, const Hit& h) {
    const uint4 sh = __ldg(sc.tri_shading + h.tri);
    F3 n;
    if (sh.w & MP_TRI_FLAT) {
        // geometric normal = e1 × e2 of the dequantised triangle, unfused (triangle.rs:141-144)
        const uint32_t p = h.tri >> 3;
        const int lane = (int)(h.tri & 7u);
        const float4 b0 = __ldg(sc.packet_box + 2 * (size_t)p), b1 = __ldg(sc.packet_box + 2 * (size_t)p + 1);
        const uint16_t* q = reinterpret_cast<const uint16_t*>(sc.packets + 9 * (size_t)p);
        float c[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const int a = k % 3;
            const float size = a == 0 ? b1.x : (a == 1 ? b1.y : b1.z), mn = a == 0 ? b0.x : (a == 1 ? b0.y : b0.z);
            const float rel = __fmul_rn((float)(int)q[k * 8 + lane], MPD_K_INV_U16);
            c[k] = __fmaf_rn(size, rel, mn);
        }
        F3 v0 = f3(c[0], c[1], c[2]), v1 = f3(c[3], c[4], c[5]), v2 = f3(c[6], c[7], c[8]);
        n = crossn(v1 - v0, v2 - v0);
    } else {
        const float w = __fsub_rn(__fsub_rn(1.0f, h.u), h.v);  // triangle.rs:234
        const float4 n0 = __ldg(sc.vertices + 2 * (size_t)sh.x), n1 = __ldg(sc.vertices + 2 * (size_t)sh.y),
                     n2 = __ldg(sc.vertices + 2 * (size_t)sh.z);
        n = (f3(n0.x, n0.y, n0.z) * w + f3(n1.x, n1.y, n1.z) * h.u) + f3(n2.x, n2.y, n2.z) * h.v;
    }
    return normalize(n);
}

}  // namespace mpd

// BVH8 traversal with EIGHT LANES PER RAY ("octet"): lane k of an octet owns child k of an inner node and triangle k of
// a packet, so the reference's 8-wide f32x8 arithmetic (src/scene/triangle_bvh/ray_bvh_intersection.rs:104-162,
// src/geometry/aabb.rs:257-283, src/geometry/triangle.rs:187-216) maps lane for lane onto 8 CUDA threads and a warp
// carries 4 rays.  Same visiting order and same per-lane arithmetic as trace.cuh (thread per ray), hence the same bits;
// what changes is the execution shape: divergence is per octet instead of per thread, each octet refills itself with the
// next ray as soon as its stack runs empty, and the per-ray stack lives in shared memory (8-byte entries).
//
// Device layout read here ("child-major", built from MpFlatBvh at upload; one 128-bit load per lane):
//   inner_cm[i*8+k]   uint4 { min.x|min.y<<16, min.z|max.x<<16, max.y|max.z<<16, link }   child k of node i
//   packets_cm[p*9+k] uint4 { v0.x|v0.y<<16, v0.z|v1.x<<16, v1.y|v1.z<<16, v2.x|v2.y<<16 } triangle k of packet p
//   packets_cm[p*9+8] 8 x u16 v2.z of the 8 triangles
//   inner_box / packet_box as in trace.cuh.
#pragma once
#include "wavefront.cuh"

namespace mpd {

struct SceneCM {
    const uint4* __restrict__ inner_cm;
    const uint4* __restrict__ packets_cm;
};

__device__ __forceinline__ float unit_rel_lo(uint32_t w) {
    return __fmaf_rn(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7410)), MPD_K_INV_U16, -8388608.0f * MPD_K_INV_U16);
}
__device__ __forceinline__ float unit_rel_hi(uint32_t w) {
    return __fmaf_rn(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7432)), MPD_K_INV_U16, -8388608.0f * MPD_K_INV_U16);
}

// One lane's share of InnerNode::intersect (same expressions as child_slab in trace.cuh).
template <bool NANPATH>
__device__ __forceinline__ bool octet_child_slab(const uint4 c, const F3& bmin, const F3& bsize, const Ray& r, float max_t, float& t1_out) {
    float lox = __fmul_rn(__fsub_rn(__fmaf_rn(bsize.x, unit_rel_lo(c.x), bmin.x), r.o.x), r.inv.x);
    float loy = __fmul_rn(__fsub_rn(__fmaf_rn(bsize.y, unit_rel_hi(c.x), bmin.y), r.o.y), r.inv.y);
    float loz = __fmul_rn(__fsub_rn(__fmaf_rn(bsize.z, unit_rel_lo(c.y), bmin.z), r.o.z), r.inv.z);
    float hix = __fmul_rn(__fsub_rn(__fmaf_rn(bsize.x, unit_rel_hi(c.y), bmin.x), r.o.x), r.inv.x);
    float hiy = __fmul_rn(__fsub_rn(__fmaf_rn(bsize.y, unit_rel_lo(c.z), bmin.y), r.o.y), r.inv.y);
    float hiz = __fmul_rn(__fsub_rn(__fmaf_rn(bsize.z, unit_rel_hi(c.z), bmin.z), r.o.z), r.inv.z);
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

// One lane's triangle of a packet (same expressions as tri_lane in trace.cuh).
__device__ __forceinline__ bool octet_tri(const uint4 a, uint32_t v2z_word, bool odd, const F3& bmin, const F3& bsize, const Ray& r,
                                          float max_t, float& t_out, float& u_out, float& v_out) {
    const F3 v0 = f3(__fmaf_rn(bsize.x, unit_rel_lo(a.x), bmin.x), __fmaf_rn(bsize.y, unit_rel_hi(a.x), bmin.y), __fmaf_rn(bsize.z, unit_rel_lo(a.y), bmin.z));
    const F3 v1 = f3(__fmaf_rn(bsize.x, unit_rel_hi(a.y), bmin.x), __fmaf_rn(bsize.y, unit_rel_lo(a.z), bmin.y), __fmaf_rn(bsize.z, unit_rel_hi(a.z), bmin.z));
    const float z2 = odd ? unit_rel_hi(v2z_word) : unit_rel_lo(v2z_word);
    const F3 v2 = f3(__fmaf_rn(bsize.x, unit_rel_lo(a.w), bmin.x), __fmaf_rn(bsize.y, unit_rel_hi(a.w), bmin.y), __fmaf_rn(bsize.z, z2, bmin.z));
    const F3 e1 = v1 - v0, e2 = v2 - v0;
    const F3 p = fma_cross(r.d, e2);
    const float det = fma_dot(e1, p);
    const float inv = __fdiv_rn(1.0f, det);
    const F3 s = r.o - v0;
    const float u = __fmul_rn(inv, fma_dot(s, p));
    const F3 qv = fma_cross(s, e1);
    const float v = __fmul_rn(inv, fma_dot(r.d, qv));
    const float t = __fmul_rn(inv, fma_dot(e2, qv));
    t_out = t, u_out = u, v_out = v;
    return u >= 0.0f && v >= 0.0f && __fadd_rn(u, v) <= 1.0f && t >= 0.0f && t <= max_t;
}

struct OctetLane {
    uint32_t sub;    // lane within the octet: child / triangle index
    uint32_t shift;  // first warp lane of the octet (0, 8, 16, 24)
    uint32_t mask;   // warp lane mask of the octet
};
__device__ __forceinline__ OctetLane octet_lane() {
    const uint32_t l = threadIdx.x & 31u;
    return OctetLane{l & 7u, l & 24u, 0xFFu << (l & 24u)};
}

// Warp-synchronous driver.  The four octets of a warp run in lockstep through PHASES so that every ballot / shuffle is
// a full-warp collective in converged code (a first version used octet-scoped masks: the compiler serialised each of
// them per octet — VOTE executed with 8 of 32 threads — see profiles/r1_wf_trace_octet_v1_partialmask.txt):
//   refill  an octet whose ray is finished emits the hit and loads its next ray (static stride over the queue)
//   pop     every octet without pending work pops its stack until an entry survives `t1 > best.t` (:40)
//   inner   octets holding an inner node test its 8 children, one per lane, and push the hits (:149-162)
//   leaf    octets holding a leaf test ONE packet, one triangle per lane (:104-141); a leaf of c packets takes c phases
// Per iteration the phase with more waiting octets runs; the others keep their pending entry.  Per ray the visiting
// order, the `best.t` snapshots and the per-lane arithmetic are exactly those of trace.cuh.
//
// fetch(i, r) loads ray i into r.o, r.d (and r.inv unless INV_FROM_D) — identical values in all 8 lanes;
// emit(i, best) is called by all 8 lanes when ray i is finished (lane sub == 0 stores).
template <int ORDER, bool COUNT, bool INV_FROM_D, class Fetch, class Emit>
__device__ __forceinline__ void octet_run(const Scene& sc, const SceneCM& cm, uint32_t n, uint32_t first, uint32_t stride, bool any_hit,
                                          uint2* __restrict__ stk, Fetch fetch, Emit emit, ThreadCounts& tc) {
    constexpr uint32_t FULL = 0xffffffffu;
    const OctetLane ol = octet_lane();
    uint32_t cur = 0, next = first;
    bool have = false, nanpath = false;
    int sp = 0;
    Ray r{};
    Hit best{MPD_FLT_MAX, 0.0f, 0.0f, MP_NO_HIT};
    bool pend_inner = false;  // an inner node popped and waiting for the inner phase
    uint32_t p_index = 0;
    uint32_t lf_index = 0, lf_rem = 0;  // leaf in progress: next packet, packets left
    float lf_max_t = 0.0f;
    Hit lf_local{MPD_INF, 0.0f, 0.0f, MP_NO_HIT};

    for (;;) {
        // ---- refill ----
        bool refilled = false;
        if (have && !pend_inner && lf_rem == 0 && sp == 0) {
            emit(cur, best);
            have = false;
        }
        if (!have && next < n) {
            cur = next, next += stride;
            fetch(cur, r);
            best = Hit{MPD_FLT_MAX, 0.0f, 0.0f, MP_NO_HIT};
            if (ol.sub == 0) stk[0] = make_uint2(sc.root, __float_as_uint(MPD_NINF));
            sp = 1;
            have = refilled = true;
            tc.rays++;
        }
        if (!__any_sync(FULL, have)) break;
        if (__any_sync(FULL, refilled)) {
            if (INV_FROM_D) {  // Ray::new's inv_direction (geometry/mod.rs:47), one division per lane instead of three
                const uint32_t a = ol.sub % 3u;
                const float dc = a == 0 ? r.d.x : (a == 1 ? r.d.y : r.d.z);
                const float ic = dc == 0.0f ? MPD_INF : __fdiv_rn(1.0f, dc);
                const float ix = __shfl_sync(FULL, ic, ol.shift + 0), iy = __shfl_sync(FULL, ic, ol.shift + 1), iz = __shfl_sync(FULL, ic, ol.shift + 2);
                if (refilled) r.inv = f3(ix, iy, iz);
            }
            if (refilled) nanpath = (fabsf(r.inv.x) == MPD_INF) | (fabsf(r.inv.y) == MPD_INF) | (fabsf(r.inv.z) == MPD_INF);
        }
        __syncwarp();  // stack entries written in the previous iteration (and the root just pushed) are visible

        // ---- pop ----
        bool need = have && !pend_inner && lf_rem == 0;
        while (__any_sync(FULL, need)) {
            if (need) {
                if (sp == 0) {
                    need = false;  // ray finished: emitted at the top of the next iteration
                } else {
                    const uint2 e = stk[--sp];
                    if (!(__uint_as_float(e.y) > best.t)) {  // ray_bvh_intersection.rs:40
                        need = false;
                        const uint32_t count = e.x & MP_LINK_COUNT_MASK, index = e.x >> MP_LINK_COUNT_BITS;
                        if (count == 0) {
                            pend_inner = true, p_index = index;
                        } else {
                            lf_index = index, lf_rem = count, lf_max_t = best.t;  // max_t is not updated inside the leaf
                            lf_local = Hit{MPD_INF, 0.0f, 0.0f, MP_NO_HIT};
                            if (COUNT) tc.vis.leaves++, tc.vis.packets += count;
                        }
                    }
                }
            }
        }

        // ---- choose the phase ----
        const uint32_t n_in = __popc(__ballot_sync(FULL, pend_inner)), n_lf = __popc(__ballot_sync(FULL, lf_rem != 0));
        if (n_in != 0 && n_in >= n_lf) {
            // ---- inner phase: lane k tests child k (ray_bvh_intersection.rs:149-162) ----
            bool hit = false;
            float t1 = 0.0f;
            uint32_t link = MP_LINK_NULL;
            if (pend_inner) {
                if (COUNT) tc.vis.inner++;
                const float4 b0 = __ldg(sc.inner_box + 2 * (size_t)p_index), b1 = __ldg(sc.inner_box + 2 * (size_t)p_index + 1);
                const uint4 c = __ldg(cm.inner_cm + 8 * (size_t)p_index + ol.sub);
                const F3 bmin = f3(b0.x, b0.y, b0.z), bsize = f3(b1.x, b1.y, b1.z);
                hit = nanpath ? octet_child_slab<true>(c, bmin, bsize, r, best.t, t1) : octet_child_slab<false>(c, bmin, bsize, r, best.t, t1);
                link = c.w;
                hit = hit && link != MP_LINK_NULL;  // NULL children are discarded when popped (:49); skipping them is neutral
            }
            const uint32_t hits = (__ballot_sync(FULL, hit) >> ol.shift) & 0xFFu;
            uint32_t pos;
            if (ORDER == MP_ORDER_NEAR_FIRST) {
                // far → near from the base; ties: lower lane nearer the top (same final order as the insertion in trace.cuh)
                pos = 0;
#pragma unroll
                for (int k = 1; k < 8; ++k) {
                    const uint32_t j = (ol.sub + k) & 7u;
                    const float tj = __shfl_sync(FULL, t1, ol.shift + j);
                    if (((hits >> j) & 1u) && (tj > t1 || (tj == t1 && j > ol.sub))) ++pos;
                }
            } else {
                pos = __popc(hits & ((1u << ol.sub) - 1u));  // lane order: lane 7 pops first (:161)
            }
            if (hit) stk[sp + pos] = make_uint2(link, __float_as_uint(t1));
            if (pend_inner) sp += __popc(hits);
            pend_inner = false;
        } else if (n_lf != 0) {
            // ---- leaf phase: lane k tests triangle k of the leaf's next packet (ray_bvh_intersection.rs:104-141) ----
            bool valid = false;
            float t = 0.0f, u = 0.0f, v = 0.0f;
            if (lf_rem != 0) {
                const float4 b0 = __ldg(sc.packet_box + 2 * (size_t)lf_index), b1 = __ldg(sc.packet_box + 2 * (size_t)lf_index + 1);
                const uint4* pk = cm.packets_cm + 9 * (size_t)lf_index;
                const uint4 a = __ldg(pk + ol.sub);
                const uint32_t zw = __ldg(reinterpret_cast<const uint32_t*>(pk + 8) + (ol.sub >> 1));
                valid = octet_tri(a, zw, ol.sub & 1u, f3(b0.x, b0.y, b0.z), f3(b1.x, b1.y, b1.z), r, lf_max_t, t, u, v);
            }
            // first minimum over the 8 lanes in lane order (:127-135): min of t, then the lowest lane that attains it
            const float key = valid ? t : MPD_INF;
            float m = fminf(key, __shfl_xor_sync(FULL, key, 1));
            m = fminf(m, __shfl_xor_sync(FULL, m, 2));
            m = fminf(m, __shfl_xor_sync(FULL, m, 4));
            const uint32_t eq = (__ballot_sync(FULL, valid && key == m) >> ol.shift) & 0xFFu;
            const uint32_t win = eq ? (uint32_t)__ffs(eq) - 1u : 0u;
            const float tw = __shfl_sync(FULL, t, ol.shift + win);
            const float uw = __shfl_sync(FULL, u, ol.shift + win);
            const float vw = __shfl_sync(FULL, v, ol.shift + win);
            if (lf_rem != 0) {
                if (eq && tw < lf_local.t) lf_local = Hit{tw, uw, vw, lf_index * 8u + win};
                ++lf_index;
                if (any_hit && lf_rem > 1 && lf_local.t < lf_max_t) {  // occluded: skip the leaf's remaining packets
                    if (COUNT) tc.vis.packets -= lf_rem - 1u;
                    lf_rem = 1;
                }
                if (--lf_rem == 0 && lf_local.t < best.t) {  // :59
                    best = lf_local;
                    if (any_hit) sp = 0;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// kernels: 128 threads = 16 octets per block; dynamic shared memory = 16 stacks of stack_cap 8-byte entries
// ---------------------------------------------------------------------------------------------------------------
#define OCTETS_PER_BLOCK 16

__device__ __forceinline__ void octet_finish_counts(DevCounters* dc, ThreadCounts& tc, bool count_visits) {
    if ((threadIdx.x & 7u) != 0) tc = ThreadCounts{};  // the 8 lanes of an octet counted the same ray
    flush_counts(dc, tc, count_visits);
}

// Object::intersect on caller-provided rays (the bit-exactness gate), octet variant.
template <int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) trace_rays_octet_kernel(Scene sc, SceneCM cm, const float* __restrict__ rays, uint32_t n,
                                                               uint4* __restrict__ hits, DevCounters* dc, uint32_t stack_cap) {
    extern __shared__ uint2 octet_stacks[];
    uint2* stk = octet_stacks + (threadIdx.x >> 3) * stack_cap;
    ThreadCounts tc{};
    octet_run<ORDER, COUNT, false>(
        sc, cm, n, blockIdx.x * OCTETS_PER_BLOCK + (threadIdx.x >> 3), gridDim.x * OCTETS_PER_BLOCK, false, stk,
        [&](uint32_t i, Ray& r) {
            const float* p = rays + 9 * (size_t)i;
            r.o = f3(__ldg(p + 0), __ldg(p + 1), __ldg(p + 2));
            r.d = f3(__ldg(p + 3), __ldg(p + 4), __ldg(p + 5));
            r.inv = f3(__ldg(p + 6), __ldg(p + 7), __ldg(p + 8));
        },
        [&](uint32_t i, const Hit& h) {
            if ((threadIdx.x & 7u) == 0) hits[i] = make_uint4(h.tri, __float_as_uint(h.t), __float_as_uint(h.u), __float_as_uint(h.v));
        },
        tc);
    octet_finish_counts(dc, tc, COUNT);
}

// Wavefront stage 2 (closest hit), octet variant of wf_trace_kernel.
template <int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) wf_trace_octet_kernel(Scene sc, SceneCM cm, const uint32_t* __restrict__ queue,
                                                             const uint32_t* __restrict__ count, const float4* __restrict__ ray_o,
                                                             const float4* __restrict__ ray_d, uint4* __restrict__ hit, DevCounters* dc,
                                                             uint32_t stack_cap) {
    extern __shared__ uint2 octet_stacks[];
    uint2* stk = octet_stacks + (threadIdx.x >> 3) * stack_cap;
    const OctetLane ol = octet_lane();
    ThreadCounts tc{};
    uint32_t path = 0;
    octet_run<ORDER, COUNT, true>(
        sc, cm, *count, blockIdx.x * OCTETS_PER_BLOCK + (threadIdx.x >> 3), gridDim.x * OCTETS_PER_BLOCK, false, stk,
        [&](uint32_t i, Ray& r) {
            path = queue[i];
            const float4 o = ray_o[path], d = ray_d[path];
            r.o = f3(o.x, o.y, o.z), r.d = f3(d.x, d.y, d.z);
        },
        [&](uint32_t, const Hit& h) {
            if (ol.sub == 0) hit[path] = make_uint4(h.tri, __float_as_uint(h.t), __float_as_uint(h.u), __float_as_uint(h.v));
        },
        tc);
    octet_finish_counts(dc, tc, COUNT);
}

// Wavefront stage 3b (any hit towards the sun), octet variant of wf_shadow_kernel.
template <int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) wf_shadow_octet_kernel(Scene sc, SceneCM cm, const float4* __restrict__ sh_o,
                                                              const uint32_t* __restrict__ sh_path, const uint32_t* __restrict__ sh_count,
                                                              float4* __restrict__ ray_d, float4* __restrict__ rgb, const float2* __restrict__ sh_c,
                                                              DevCounters* dc, uint32_t stack_cap) {
    extern __shared__ uint2 octet_stacks[];
    uint2* stk = octet_stacks + (threadIdx.x >> 3) * stack_cap;
    const OctetLane ol = octet_lane();
    ThreadCounts tc{};
    float contribution = 0.0f;
    const Ray sun = ray_new(f3(0.f, 0.f, 0.f), f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
    octet_run<ORDER, COUNT, false>(
        sc, cm, *sh_count, blockIdx.x * OCTETS_PER_BLOCK + (threadIdx.x >> 3), gridDim.x * OCTETS_PER_BLOCK, true, stk,
        [&](uint32_t i, Ray& r) {
            const float4 o = sh_o[i];
            contribution = o.w;
            r.o = f3(o.x, o.y, o.z), r.d = sun.d, r.inv = sun.inv;  // = ray_new(origin, sun): the direction is the same for all
            tc.shadow_rays++;
        },
        [&](uint32_t i, const Hit& h) {
            if (ol.sub == 0 && h.tri == MP_NO_HIT) add_sun(ray_d, rgb, sh_c, sh_path[i], i, contribution);
        },
        tc);
    octet_finish_counts(dc, tc, COUNT);
}

}  // namespace mpd

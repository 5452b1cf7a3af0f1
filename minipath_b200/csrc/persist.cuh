// Thread-per-ray BVH8 traversal, PERSISTENT and PHASED — the third traversal kernel variant (MP_TRACE_PERSIST).
// Same per-ray visiting order and arithmetic as trace.cuh (reference src/scene/triangle_bvh/ray_bvh_intersection.rs:26-64,
// 104-162), hence the same bits; what changes is how a warp's 32 rays share the instruction stream:
//   * every lane refills itself with the next ray of the queue as soon as its ray is finished (no lane idles until the
//     warp's longest ray ends — the incoherent bounces ran at 7-11 active threads of 32, profiles/r1_wf_trace_thread.txt);
//   * work is cut into uniform items — one inner node (8 child slab tests) or ONE triangle packet (8 triangle tests) —
//     and per iteration the warp votes and runs the item kind that more lanes are waiting for; the minority keeps its
//     popped entry for the next round.  A leaf of c packets takes c packet items, with the leaf-local best hit and the
//     max_t snapshot (:106-107) carried in registers between them.
#pragma once
#include "wavefront.cuh"

namespace mpd {

// fetch(i, r) loads ray i completely (o, d, inv); emit(i, best) stores its result.
template <int LAYOUT, int ORDER, bool COUNT, class Fetch, class Emit>
__device__ __forceinline__ void persist_run(const Scene& sc, uint32_t n, uint32_t first, uint32_t stride, bool any_hit, Fetch fetch, Emit emit,
                                            ThreadCounts& tc) {
    constexpr uint32_t FULL = 0xffffffffu;
    uint2 stk[MPD_STACK_MAX];
    int sp = 0;
    uint32_t cur = 0, next = first;
    bool have = false, nanpath = false;
    Ray r{};
    RaySel sel{};
    Hit best{MPD_FLT_MAX, 0.0f, 0.0f, MP_NO_HIT};
    uint32_t kind = 0;  // 0 nothing popped, 1 inner node pending, 2 leaf in progress
    uint32_t p_index = 0, lf_rem = 0;
    float lf_max_t = 0.0f;
    Hit lf_local{MPD_INF, 0.0f, 0.0f, MP_NO_HIT};

    for (;;) {
        // ---- refill + pop: until this lane holds an item or has run out of rays ----
        while (kind == 0) {
            if (sp == 0) {
                if (have) {
                    emit(cur, best);
                    have = false;
                }
                if (next >= n) break;
                cur = next, next += stride;
                fetch(cur, r);
                nanpath = ray_needs_nanpath(sc, r);
                if (LAYOUT == MP_LAYOUT_DECODED) sel = ray_sel(r);
                best = Hit{MPD_FLT_MAX, 0.0f, 0.0f, MP_NO_HIT};
                stk[0] = make_uint2(sc.root, __float_as_uint(MPD_NINF));
                sp = 1;
                have = true;
                tc.rays++;
            }
            const uint2 e = stk[--sp];
            if (__uint_as_float(e.y) > best.t) continue;  // ray_bvh_intersection.rs:40
            const uint32_t count = e.x & MP_LINK_COUNT_MASK;
            p_index = e.x >> MP_LINK_COUNT_BITS;
            if (count == 0) {
                kind = 1;
            } else {
                kind = 2, lf_rem = count, lf_max_t = best.t;  // max_t is not updated inside the leaf
                lf_local = Hit{MPD_INF, 0.0f, 0.0f, MP_NO_HIT};
                if (COUNT) tc.vis.leaves++, tc.vis.packets += count;
            }
        }
        const uint32_t m_in = __ballot_sync(FULL, kind == 1), m_lf = __ballot_sync(FULL, kind == 2);
        if ((m_in | m_lf) == 0) break;  // every lane of the warp is out of rays
        if (__popc(m_in) >= __popc(m_lf)) {
            if (kind == 1) {
                if (COUNT) tc.vis.inner++;
                StackSink<ORDER, LocalStack> sink{LocalStack{stk}, sp, sp};
                if (nanpath)
                    visit_inner<LAYOUT, true>(sc, r, sel, p_index, best.t, sink);
                else
                    visit_inner<LAYOUT, false>(sc, r, sel, p_index, best.t, sink);
                sp = sink.sp;
                kind = 0;
            }
        } else if (kind == 2) {
            // ---- one packet of the leaf (ray_bvh_intersection.rs:104-141) ----
            visit_packet<LAYOUT>(sc, r, p_index, lf_max_t, lf_local);
            ++p_index;
            if (any_hit && lf_rem > 1 && lf_local.t < lf_max_t) {  // occluded: skip the leaf's remaining packets
                if (COUNT) tc.vis.packets -= lf_rem - 1u;
                lf_rem = 1;
            }
            if (--lf_rem == 0) {
                kind = 0;
                if (lf_local.t < best.t) {  // :59
                    best = lf_local;
                    if (any_hit) sp = 0;
                }
            }
        }
    }
}

// ---- kernels: persistent grid (resident blocks x SMs), lane-strided over the queue ----

template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) trace_rays_persist_kernel(Scene sc, const float* __restrict__ rays, uint32_t n, uint4* __restrict__ hits,
                                                                 DevCounters* dc) {
    ThreadCounts tc{};
    persist_run<LAYOUT, ORDER, COUNT>(
        sc, n, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x, false,
        [&](uint32_t i, Ray& r) {
            const float* p = rays + 9 * (size_t)i;
            r.o = f3(__ldg(p + 0), __ldg(p + 1), __ldg(p + 2));
            r.d = f3(__ldg(p + 3), __ldg(p + 4), __ldg(p + 5));
            r.inv = f3(__ldg(p + 6), __ldg(p + 7), __ldg(p + 8));
        },
        [&](uint32_t i, const Hit& h) { hits[i] = make_uint4(h.tri, __float_as_uint(h.t), __float_as_uint(h.u), __float_as_uint(h.v)); }, tc);
    flush_counts(dc, tc, COUNT);
}

template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) wf_trace_persist_kernel(Scene sc, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count,
                                                               const float4* __restrict__ ray_o, const float4* __restrict__ ray_d,
                                                               uint4* __restrict__ hit, DevCounters* dc) {
    ThreadCounts tc{};
    uint32_t path = 0;
    persist_run<LAYOUT, ORDER, COUNT>(
        sc, *count, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x, false,
        [&](uint32_t i, Ray& r) {
            path = queue[i];
            r = load_ray(ray_o[path], ray_d[path]);
        },
        [&](uint32_t, const Hit& h) { hit[path] = make_uint4(h.tri, __float_as_uint(h.t), __float_as_uint(h.u), __float_as_uint(h.v)); }, tc);
    flush_counts(dc, tc, COUNT);
}

template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128) wf_shadow_persist_kernel(Scene sc, const float4* __restrict__ sh_o, const uint32_t* __restrict__ sh_path,
                                                                const uint32_t* __restrict__ sh_count, float4* __restrict__ ray_d, float4* __restrict__ rgb,
                                                                const float2* __restrict__ sh_c, DevCounters* dc) {
    ThreadCounts tc{};
    float contribution = 0.0f;
    const Ray sun = ray_new(f3(0.f, 0.f, 0.f), f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
    persist_run<LAYOUT, ORDER, COUNT>(
        sc, *sh_count, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x, true,
        [&](uint32_t i, Ray& r) {
            const float4 o = sh_o[i];
            contribution = o.w;
            r.o = f3(o.x, o.y, o.z), r.d = sun.d, r.inv = sun.inv;  // = ray_new(origin, sun)
            tc.shadow_rays++;
        },
        [&](uint32_t i, const Hit& h) {
            if (h.tri == MP_NO_HIT) add_sun(ray_d, rgb, sh_c, sh_path[i], i, contribution);
        },
        tc);
    flush_counts(dc, tc, COUNT);
}

}  // namespace mpd

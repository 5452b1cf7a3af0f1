// BVH8 traversal with a per-warp RAY POOL — the fifth traversal variant (MP_TRACE_POOL).
//
// The thread-per-ray kernels lose most of the machine to SIMT divergence on incoherent rays: at any moment some lanes want
// an inner node (8 slab tests), some a triangle packet (8 Möller–Trumbore tests), some are finished (profiles/r1_wf_trace_thread.txt:
// 7-11 of 32 threads active on bounce rays).  Here a warp does not own 32 rays but a POOL of 96, whose state lives in shared
// memory (ray, best hit, leaf-in-progress; SoA, 4-byte columns) with the stacks in a global scratch area.  Every ray has exactly
// one pending work item — an inner node or ONE packet of a leaf — and the warp keeps three small ring queues of slot ids
// (inner items, packet items, free slots).  Each round the warp takes 32 items OF ONE KIND and runs them in lockstep, so all 32
// lanes execute the same code on real work; then every lane pops its ray's next surviving stack entry (`t1 > best.t`,
// ray_bvh_intersection.rs:40) and files the slot under the right queue, or emits the hit and frees the slot.  With 64..96 rays
// alive, one of the two queues always holds at least 32 items; free slots are refilled 32 at a time with consecutive rays of the
// input queue (claimed with one atomicAdd), which also keeps the ray fetch coalesced.
//
// Per ray, the visiting order, the `best.t` snapshots (max_t at node pop, leaf-local best, :106-107,:59) and the arithmetic are
// exactly those of trace.cuh, hence the same bits.
#pragma once
#include "persist.cuh"

namespace mpd {

#ifndef POOL_W
#define POOL_W 80       // ray slots per warp (round 2: 96 -> 80 together with 7 blocks per SM, profiles/r2o_variants_pool_occupancy.txt)
#endif
#define POOL_RING 128   // ring capacity (power of two >= POOL_W)
#ifndef POOL_VEC_RAY
#define POOL_VEC_RAY 0  // 1: the read-only part of a slot (origin, direction, inv.xy) lives in two float4 groups read with LDS.128 — a 128-bit warp access
                        // is served per quarter-warp, so the bank conflicts of arbitrary slot sets arise among 8 lanes instead of 32, and 8 scalar loads become 2
#endif
#if POOL_VEC_RAY
#define POOL_VECS 2     // float4 groups per slot
#define POOL_COLS 13    // 4-byte state columns per slot
#else
#define POOL_VECS 0
#define POOL_COLS 20    // 4-byte state columns per slot
#endif
#define POOL_STATE_BYTES ((POOL_VECS * 16 + POOL_COLS * 4) * POOL_W)  // per warp; 80 bytes per slot
#define POOL_WARPS 4    // warps per block
#ifndef POOL_MIN_BLOCKS
#define POOL_MIN_BLOCKS 7  // resident blocks per SM the register allocation aims at.  The kernel waits on L2 round trips (long_scoreboard is the largest stall,
                           // 5 warps per scheduler issue 65 % of the cycles): since the inner-node code lost its branches, 72 registers cost 36 bytes of spills and
                           // 28 warps per SM beat 20 (round 1 measured 6 / 8 blocks slower: 80 / 64 registers spilled much more then).  7 blocks x 80-ray pools x 80 B
                           // of state fit the 228 KB shared-memory configuration.
#endif
#ifndef POOL_MIN_BLOCKS_DECODED
#define POOL_MIN_BLOCKS_DECODED 4  // the DECODED instantiations serve BVHs of at most 1 MB (AUTO layout), which live in L1: there 4 blocks (102 registers)
                                   // beat 5 (96): default workload 32.0 -> 30.0 ms, while the stand-in (QUANTISED records) loses 10 % with 4
#endif
#define POOL_BLOCKS_FOR(LAYOUT) ((LAYOUT) == MP_LAYOUT_DECODED ? POOL_MIN_BLOCKS_DECODED : POOL_MIN_BLOCKS)
#ifndef POOL_MIXED
#define POOL_MIXED 0    // 1: the QUANTISED instantiations read DECODED inner nodes (7 x LDG.256) and keep the quantised packets
#endif
#ifndef POOL_PUSH_RANK
#define POOL_PUSH_RANK 1 // hit children go to the stack as predicated stores at sp + rank, the top one stays in registers (no per-child branches)
#endif
#ifndef POOL_STACK_CG
#define POOL_STACK_CG 0 // 1: the global stack columns bypass L1 (ld.global.cg / st.global.cg) so that they do not evict node records from it
#endif
#ifndef POOL_POP2
#define POOL_POP2 1     // pop loads two stack entries per iteration (both in flight before either is tested); 0: one
#endif
#ifndef POOL_SS
#define POOL_SS 0       // entries of every ray's stack kept in shared memory (power of two, 0 = none); the rest lives in the global scratch.
                        // Measured on the stand-in (8 spp, ms per frame): 0 -> 79.4, 2 -> 85.8, 4 -> 86.7: the ring bookkeeping in the unrolled
                        // child loop and the lost occupancy cost more than the L1 requests it saves, so it is off.
#endif
#define POOL_SMEM_PER_WARP (POOL_STATE_BYTES + 3 * POOL_RING + POOL_SS * POOL_W * 8)

struct PoolWarp {  // shared-memory view of one warp's pool
    float4* vec;         // [POOL_VECS][POOL_W]  (POOL_VEC_RAY)
    float* col;          // [POOL_COLS][POOL_W]
    unsigned char* ring; // [3][POOL_RING]: 0 inner items, 1 packet items, 2 free slots
    __device__ __forceinline__ float& f(int c, uint32_t s) const { return col[c * POOL_W + s]; }
    __device__ __forceinline__ uint32_t& u(int c, uint32_t s) const { return reinterpret_cast<uint32_t*>(col)[c * POOL_W + s]; }
};
// columns
enum {
#if !POOL_VEC_RAY
       PC_OX, PC_OY, PC_OZ, PC_DX, PC_DY, PC_DZ, PC_IX, PC_IY,
#endif
       PC_IZ, PC_BT, PC_BU, PC_BV, PC_BTRI, PC_SP, PC_RID, PC_ITEM,  // PC_ITEM: the pending item as a CompressedNodeLink — index << 3 | packets left (0 = inner node)
       PC_LT, PC_LU, PC_LV, PC_LTRI };

// Per-ray traversal stack.  Profiling the first pool kernel showed three quarters of its global L1 requests were stack pushes and
// pops (profiles/r1c_wf_trace_pool_records.txt): every lane touches its own line.  With SHORT the newest POOL_SS entries of a
// ray's stack live in shared memory as a ring ([entry % POOL_SS][slot], 8-byte words, neighbouring slots in neighbouring banks)
// and only entries pushed deeper than that spill to the global scratch ([entry][slot]); `lo` is the first entry held in the ring.
// NEAR_FIRST needs random access for its insertion sort and keeps the whole stack in the scratch (SHORT = false).
template <bool SHORT>
struct PStack {
    static constexpr bool kShort = SHORT;
    uint2* ring;   // shared memory, this ray's column
    uint2* spill;  // global scratch, this ray's column
    int sp, lo;
    __device__ __forceinline__ uint2& at(int i) const { return spill[(size_t)i * POOL_W]; }  // !SHORT only (StackSink's lvalue use)
    __device__ __forceinline__ uint2 ld(int i) const { return POOL_STACK_CG ? __ldcg(spill + (size_t)i * POOL_W) : spill[(size_t)i * POOL_W]; }
    __device__ __forceinline__ void put(int i, uint2 e) const {
        if (POOL_STACK_CG) __stcg(spill + (size_t)i * POOL_W, e);
        else spill[(size_t)i * POOL_W] = e;
    }
    __device__ __forceinline__ void push(uint2 e) {
        if (SHORT) {
            if (sp - lo == POOL_SS) {  // ring full: the oldest entry goes to the scratch
                put(lo, ring[(lo & (POOL_SS - 1)) * POOL_W]);
                ++lo;
            }
            ring[(sp & (POOL_SS - 1)) * POOL_W] = e;
            ++sp;
        } else {
            put(sp, e);
            ++sp;
        }
    }
    __device__ __forceinline__ uint2 pop() {  // sp > 0
        --sp;
        if (SHORT) {
            if (sp >= lo) return ring[(sp & (POOL_SS - 1)) * POOL_W];
            lo = sp;  // ring empty: take the entry back from the scratch
        }
        return ld(sp);
    }
    __device__ __forceinline__ void clear() { sp = 0, lo = 0; }
    __device__ __forceinline__ uint32_t pack() const { return SHORT ? ((uint32_t)sp | (uint32_t)lo << 16) : (uint32_t)sp; }
    __device__ __forceinline__ void unpack(uint32_t v) {
        sp = SHORT ? (int)(v & 0xffffu) : (int)v;
        lo = SHORT ? (int)(v >> 16) : 0;
    }
};

// DeferSink (trace.cuh) on a PStack: every hit child but the last is pushed, the last is handed to pool_advance.
template <class S>
struct PoolDeferSink {
    S& stk;
    uint2 top;
    bool has;
    __device__ __forceinline__ void hit(uint32_t link, float t1) {
        if (has) stk.push(top);
        top = make_uint2(link, __float_as_uint(t1));
        has = true;
    }
};

struct PoolQueues {  // warp-uniform ring cursors
    uint32_t head[3], tail[3];
    __device__ __forceinline__ uint32_t size(int q) const { return tail[q] - head[q]; }
};

// All 32 lanes call this; lanes with pred append their slot to ring q.
__device__ __forceinline__ void pool_push(const PoolWarp& pw, PoolQueues& qs, int q, bool pred, uint32_t slot) {
    const uint32_t m = __ballot_sync(0xffffffffu, pred);
    if (pred) pw.ring[q * POOL_RING + ((qs.tail[q] + __popc(m & ((1u << (threadIdx.x & 31u)) - 1u))) & (POOL_RING - 1))] = (unsigned char)slot;
    qs.tail[q] += __popc(m);
}

// Pop the ray's next surviving entry and file the slot; `active` lanes own `slot` with stack `stk` and current best_t.  `has_top`:
// the entry the inner phase kept in registers instead of pushing it (PoolDeferSink), or the root of a fresh ray — it is the top of
// the stack and needs no `t1 > best.t` test (t1 <= t2 <= max_t = best.t; the root carries -inf).
template <bool COUNT, class S, class Emit>
__device__ __forceinline__ uint32_t pool_next(const PoolWarp& pw, uint32_t slot, S& stk, float best_t, Emit emit, ThreadCounts& tc, bool has_top, uint2 top) {
    uint32_t kind = 0;  // 0 finished, 1 inner, 2 leaf
    while (has_top || stk.sp > 0) {
        uint2 e;
        if (has_top) {
            e = top, has_top = false;
        } else if (S::kShort) {
            e = stk.pop();
            if (__uint_as_float(e.y) > best_t) continue;  // ray_bvh_intersection.rs:40
        } else if (!POOL_POP2) {
            e = stk.pop();
            if (__uint_as_float(e.y) > best_t) continue;  // ray_bvh_intersection.rs:40
        } else {
            // Two entries per iteration, both loads in flight before either is tested: a run of culled entries is otherwise a
            // chain of dependent L2 round trips (13 % of the kernel's stall samples sat on this test, profiles/r1d_wf_pool_32Mpaths.txt).
            const int i = stk.sp - 1;
            const uint2 a = stk.ld(i);
            const uint2 b = stk.ld(i > 0 ? i - 1 : 0);
            stk.sp = i;
            if (!(__uint_as_float(a.y) > best_t)) {  // ray_bvh_intersection.rs:40
                e = a;
            } else if (i > 0 && !(__uint_as_float(b.y) > best_t)) {
                e = b, stk.sp = i - 1;
            } else {
                stk.sp = i > 0 ? i - 1 : 0;
                continue;
            }
        }
        const uint32_t count = e.x & MP_LINK_COUNT_MASK;
        pw.u(PC_ITEM, slot) = e.x;
        if (count == 0) {
            kind = 1;
        } else {
            kind = 2;
            pw.f(PC_LT, slot) = MPD_INF;  // max_t of the leaf is best.t, which does not change until the leaf is finished
            pw.u(PC_LTRI, slot) = MP_NO_HIT;
            if (COUNT) tc.vis.leaves++, tc.vis.packets += count;
        }
        break;
    }
    pw.u(PC_SP, slot) = stk.pack();
    if (kind == 0)
        emit(pw.u(PC_RID, slot), Hit{pw.f(PC_BT, slot), pw.f(PC_BU, slot), pw.f(PC_BV, slot), pw.u(PC_BTRI, slot)});
    return kind;
}

template <bool COUNT, class S, class Emit>
__device__ __forceinline__ void pool_advance(const PoolWarp& pw, PoolQueues& qs, bool active, uint32_t slot, S& stk, float best_t, Emit emit,
                                             ThreadCounts& tc, bool has_top = false, uint2 top = make_uint2(0u, 0u)) {
    uint32_t kind = 0;  // 0 finished, 1 inner, 2 leaf
    if (active) kind = pool_next<COUNT>(pw, slot, stk, best_t, emit, tc, has_top, top);
    pool_push(pw, qs, 0, active && kind == 1, slot);
    pool_push(pw, qs, 1, active && kind == 2, slot);
    pool_push(pw, qs, 2, active && kind == 0, slot);
}

// `fixed`: every ray of the launch has this direction (the sun rays of wf_shadow_pool_kernel) — direction and inverse direction then stay in
// registers and only the origin is kept per slot: 3 instead of 9 state loads per round.
__device__ __forceinline__ Ray pool_load_ray(const PoolWarp& pw, uint32_t s, const Ray* fixed) {
    Ray r;
#if !POOL_VEC_RAY
    if (fixed) {
        r.o = f3(pw.f(PC_OX, s), pw.f(PC_OY, s), pw.f(PC_OZ, s));
        r.d = fixed->d, r.inv = fixed->inv;
        return r;
    }
#endif
#if POOL_VEC_RAY
    const float4 a = pw.vec[s], b = pw.vec[POOL_W + s];
    r.o = f3(a.x, a.y, a.z);
    r.d = f3(a.w, b.x, b.y);
    r.inv = f3(b.z, b.w, pw.f(PC_IZ, s));
#else
    r.o = f3(pw.f(PC_OX, s), pw.f(PC_OY, s), pw.f(PC_OZ, s));
    r.d = f3(pw.f(PC_DX, s), pw.f(PC_DY, s), pw.f(PC_DZ, s));
    r.inv = f3(pw.f(PC_IX, s), pw.f(PC_IY, s), pw.f(PC_IZ, s));
#endif
    return r;
}
__device__ __forceinline__ void pool_store_ray(const PoolWarp& pw, uint32_t s, const Ray& r, const Ray* fixed) {
#if !POOL_VEC_RAY
    if (fixed) {
        pw.f(PC_OX, s) = r.o.x, pw.f(PC_OY, s) = r.o.y, pw.f(PC_OZ, s) = r.o.z;
        return;
    }
#endif
#if POOL_VEC_RAY
    pw.vec[s] = make_float4(r.o.x, r.o.y, r.o.z, r.d.x);
    pw.vec[POOL_W + s] = make_float4(r.d.y, r.d.z, r.inv.x, r.inv.y);
    pw.f(PC_IZ, s) = r.inv.z;
#else
    pw.f(PC_OX, s) = r.o.x, pw.f(PC_OY, s) = r.o.y, pw.f(PC_OZ, s) = r.o.z;
    pw.f(PC_DX, s) = r.d.x, pw.f(PC_DY, s) = r.d.y, pw.f(PC_DZ, s) = r.d.z;
    pw.f(PC_IX, s) = r.inv.x, pw.f(PC_IY, s) = r.inv.y, pw.f(PC_IZ, s) = r.inv.z;
#endif
}

// fetch(i, r) loads ray i completely (o, d, inv); emit(i, hit) stores its result.  `claim` is a zeroed device counter shared by
// the launch: warps claim 32 consecutive rays at a time.  `scratch` holds gridDim.x * POOL_WARPS * POOL_W stacks of stack_cap entries.
template <int LAYOUT, int ORDER, bool COUNT, class Fetch, class Emit>
__device__ __forceinline__ void pool_run(const Scene& sc, uint32_t n, uint32_t* claim, bool any_hit, uint2* __restrict__ scratch, uint32_t stack_cap,
                                         unsigned char* smem, Fetch fetch, Emit emit, ThreadCounts& tc, const Ray* fixed = nullptr) {
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    PoolWarp pw;
    pw.vec = reinterpret_cast<float4*>(smem + (size_t)warp * POOL_SMEM_PER_WARP);
    pw.col = reinterpret_cast<float*>(smem + (size_t)warp * POOL_SMEM_PER_WARP + POOL_VECS * 16 * POOL_W);
    pw.ring = smem + (size_t)warp * POOL_SMEM_PER_WARP + POOL_STATE_BYTES;
    uint2* const wstacks = scratch + ((size_t)blockIdx.x * POOL_WARPS + warp) * POOL_W * stack_cap;
    uint2* const wring = reinterpret_cast<uint2*>(smem + (size_t)warp * POOL_SMEM_PER_WARP + POOL_STATE_BYTES + 3 * POOL_RING);
    constexpr bool SHORT = ORDER == MP_ORDER_REFERENCE && POOL_SS > 0;
    using Stack = PStack<SHORT>;
    PoolQueues qs{};
    for (uint32_t s = lane; s < POOL_W; s += 32) pw.ring[2 * POOL_RING + s] = (unsigned char)s;  // every slot starts free
    qs.tail[2] = POOL_W;
    bool more = true;  // the input queue may still hold rays
    __syncwarp();

    for (;;) {
        const uint32_t n_in = qs.size(0), n_lf = qs.size(1), n_free = qs.size(2);
        // ---- refill: 32 free slots (or the pool is nearly empty) and rays left ----
        if (more && (n_free >= 32 || (n_in < 32 && n_lf < 32 && n_free > 0))) {
            const uint32_t want = n_free < 32 ? n_free : 32;
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(claim, want);
            base = __shfl_sync(FULL, base, 0);
            const uint32_t got = base >= n ? 0 : (n - base < want ? n - base : want);
            if (got < want) more = false;
            const bool act = lane < got;
            uint32_t slot = 0;
            if (act) {
                slot = pw.ring[2 * POOL_RING + ((qs.head[2] + lane) & (POOL_RING - 1))];
                Ray r;
                fetch(base + lane, r);
                pool_store_ray(pw, slot, r, fixed);
                pw.f(PC_BT, slot) = MPD_FLT_MAX, pw.f(PC_BU, slot) = 0.0f, pw.f(PC_BV, slot) = 0.0f, pw.u(PC_BTRI, slot) = MP_NO_HIT;
                pw.u(PC_RID, slot) = base + lane;
                tc.rays++;
            }
            qs.head[2] += got;
            Stack stk{wring + slot, wstacks + slot, 0, 0};
            pool_advance<COUNT>(pw, qs, act, slot, stk, MPD_FLT_MAX, emit, tc, true, make_uint2(sc.root, __float_as_uint(MPD_NINF)));
            __syncwarp();
            continue;
        }
        if (n_in == 0 && n_lf == 0) break;  // nothing pending and nothing left to fetch

        if (n_in >= n_lf) {
            // ---- 32 inner nodes (ray_bvh_intersection.rs:149-162) ----
            const uint32_t take = n_in < 32 ? n_in : 32;
            const bool act = lane < take;
            uint32_t slot = 0;
            float bt = 0.0f;
            Stack stk{wring, wstacks, 0, 0};
            bool has_top = false;
            uint2 top = make_uint2(0u, 0u);
            if (act) {
                slot = pw.ring[0 * POOL_RING + ((qs.head[0] + lane) & (POOL_RING - 1))];
                stk.ring = wring + slot, stk.spill = wstacks + slot;
                stk.unpack(pw.u(PC_SP, slot));
                const Ray r = pool_load_ray(pw, slot, fixed);
                bt = pw.f(PC_BT, slot);
                const uint32_t node = pw.u(PC_ITEM, slot) >> MP_LINK_COUNT_BITS;
                if (COUNT) tc.vis.inner++;
                const bool nanpath = ray_needs_nanpath(sc, r);
                RaySel sel{};
                constexpr bool MIXED = POOL_MIXED && LAYOUT == MP_LAYOUT_QUANTISED;
                if ((MIXED || LAYOUT == MP_LAYOUT_DECODED) && !nanpath) sel = ray_sel(r);
                if (ORDER == MP_ORDER_REFERENCE && POOL_PUSH_RANK && !SHORT && (MIXED || LAYOUT == MP_LAYOUT_QUANTISED)) {
                    // Every hit child but the highest lane goes to the stack at sp + (its rank among the hits), lanes ascending — the
                    // order StackSink / PoolDeferSink produce — and the highest lane, which would be popped next and always survives the
                    // `t1 > best.t` test (t1 <= t2 <= max_t = best.t), is handed to pool_advance in registers.
                    uint32_t links[8];
                    float t1s[8];
                    uint32_t mask;
                    if (nanpath)
                        mask = inner_quantised_hits<true>(sc, r, node, bt, links, t1s);
                    else if (MIXED)
                        mask = inner_decoded256_hits(sc, r, sel, node, bt, links, t1s);
                    else
                        mask = inner_quantised_hits<false>(sc, r, node, bt, links, t1s);
                    has_top = mask != 0u;
#pragma unroll
                    for (int lane = 0; lane < 8; ++lane) {
                        const bool h = (mask >> lane) & 1u;
                        const bool push = lane < 7 && h && (mask >> (lane + 1)) != 0u;  // a higher lane hit too: this one is not the top
                        if (push) {
                            stk.put(stk.sp, make_uint2(links[lane], __float_as_uint(t1s[lane])));
                            ++stk.sp;
                        }
                        if (h) top = make_uint2(links[lane], __float_as_uint(t1s[lane]));
                    }
                } else if (ORDER == MP_ORDER_REFERENCE) {
                    PoolDeferSink<Stack> sink{stk, make_uint2(0u, 0u), false};
                    if (nanpath)
                        visit_inner<LAYOUT, true>(sc, r, sel, node, bt, sink);
                    else if (MIXED)
                        inner_decoded256(sc, r, sel, node, bt, sink);
                    else
                        visit_inner<LAYOUT, false>(sc, r, sel, node, bt, sink);
                    has_top = sink.has, top = sink.top;
                } else {
                    StackSink<ORDER, Stack> sink{stk, stk.sp, stk.sp};
                    if (nanpath)
                        visit_inner<LAYOUT, true>(sc, r, sel, node, bt, sink);
                    else
                        visit_inner<LAYOUT, false>(sc, r, sel, node, bt, sink);
                    stk.sp = sink.sp;
                }
            }
            qs.head[0] += take;
            pool_advance<COUNT>(pw, qs, act, slot, stk, bt, emit, tc, has_top, top);
        } else {
            // ---- 32 triangle packets, one per ray (ray_bvh_intersection.rs:104-141) ----
            const uint32_t take = n_lf < 32 ? n_lf : 32;
            const bool act = lane < take;
            uint32_t slot = 0;
            float bt = 0.0f;
            Stack stk{wring, wstacks, 0, 0};
            bool done = false;  // the leaf's last packet has been tested
            if (act) {
                slot = pw.ring[1 * POOL_RING + ((qs.head[1] + lane) & (POOL_RING - 1))];
                const Ray r = pool_load_ray(pw, slot, fixed);
                const uint32_t item = pw.u(PC_ITEM, slot);
                uint32_t pidx = item >> MP_LINK_COUNT_BITS, rem = item & MP_LINK_COUNT_MASK;
                bt = pw.f(PC_BT, slot);  // = the leaf's max_t: best.t is not updated inside the leaf (:106-107)
                Hit local{pw.f(PC_LT, slot), pw.f(PC_LU, slot), pw.f(PC_LV, slot), pw.u(PC_LTRI, slot)};
                visit_packet<LAYOUT>(sc, r, pidx, bt, local);
                ++pidx, --rem;
                if (any_hit && rem != 0 && local.t < bt) {  // occluded: skip the leaf's remaining packets
                    if (COUNT) tc.vis.packets -= rem;
                    rem = 0;
                }
                pw.u(PC_ITEM, slot) = pidx << MP_LINK_COUNT_BITS | rem;
                if (rem == 0) {
                    done = true;
                    stk.ring = wring + slot, stk.spill = wstacks + slot;
                    stk.unpack(pw.u(PC_SP, slot));
                    if (local.t < bt) {  // :59
                        bt = local.t;
                        pw.f(PC_BT, slot) = local.t, pw.f(PC_BU, slot) = local.u, pw.f(PC_BV, slot) = local.v, pw.u(PC_BTRI, slot) = local.tri;
                        if (any_hit) stk.clear();
                    }
                } else {
                    pw.f(PC_LT, slot) = local.t, pw.f(PC_LU, slot) = local.u, pw.f(PC_LV, slot) = local.v, pw.u(PC_LTRI, slot) = local.tri;
                }
            }
            qs.head[1] += take;
            pool_push(pw, qs, 1, act && !done, slot);  // more packets of the same leaf: back into the packet queue
            pool_advance<COUNT>(pw, qs, act && done, slot, stk, bt, emit, tc);
        }
        __syncwarp();
    }
}

#ifndef POOL_LANE_OWNED
#define POOL_LANE_OWNED 0
#endif
#define POOL_LANE_OWNED_CHECK POOL_LANE_OWNED
#ifndef POOL_OWNED_REFILL
#define POOL_OWNED_REFILL 32u  // free slots in the warp that trigger a refill (one atomicAdd on the launch's claim counter)
#endif
// LANE-OWNED POOL (POOL_LANE_OWNED = 1).  In pool_run a round takes ANY 32 of the warp's 96 slots, so the lanes' state accesses
// (col[c * 96 + slot], bank = slot mod 32) collide: the ncu capture of the stand-in's closest-hit launch counts one bank-conflict replay per
// shared access (profiles/r1g_pool_atrium_kernel.txt: 305 M loads + 285 M stores, 590 M conflicts), on the same L1 data pipe the node
// gathers use.  Here lane L owns the slots L, L + 32, L + 64: every state access of a round hits bank L — conflict-free — and the three ring
// queues with their ballots disappear; a round's kind is voted (lanes that hold an inner item vs lanes that hold a packet item) and every lane
// runs one of ITS rays that has an item of that kind.  Per ray nothing changes (same visiting order, snapshots and arithmetic).
template <int LAYOUT, int ORDER, bool COUNT, class Fetch, class Emit>
__device__ __forceinline__ void pool_run_owned(const Scene& sc, uint32_t n, uint32_t* claim, bool any_hit, uint2* __restrict__ scratch, uint32_t stack_cap,
                                               unsigned char* smem, Fetch fetch, Emit emit, ThreadCounts& tc, const Ray* fixed = nullptr) {
    constexpr uint32_t FULL = 0xffffffffu;
    constexpr int K = POOL_W / 32;  // slots per lane
    static_assert(!POOL_LANE_OWNED_CHECK || (POOL_W % 32 == 0 && K <= 8), "lane-owned pools hold a whole number of slots per lane");
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    PoolWarp pw;
    pw.vec = reinterpret_cast<float4*>(smem + (size_t)warp * POOL_SMEM_PER_WARP);
    pw.col = reinterpret_cast<float*>(smem + (size_t)warp * POOL_SMEM_PER_WARP + POOL_VECS * 16 * POOL_W);
    pw.ring = nullptr;
    uint2* const wstacks = scratch + ((size_t)blockIdx.x * POOL_WARPS + warp) * POOL_W * stack_cap;
    using Stack = PStack<false>;
    uint32_t kinds = 0;  // 2 bits per owned slot: 0 free, 1 inner item pending, 2 packet item pending
    bool more = true;    // the input queue may still hold rays

    for (;;) {
        uint32_t n_free_me = 0;
        bool has_in = false, has_pk = false;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const uint32_t kd = (kinds >> (2 * k)) & 3u;
            n_free_me += kd == 0u, has_in |= kd == 1u, has_pk |= kd == 2u;
        }
        const uint32_t n_in = __popc(__ballot_sync(FULL, has_in)), n_pk = __popc(__ballot_sync(FULL, has_pk));
        // ---- refill: at least 32 free slots in the warp, or the next round would run thin ----
        if (more) {
            uint32_t incl = n_free_me;  // inclusive scan of the lanes' free-slot counts
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t v = __shfl_up_sync(FULL, incl, d);
                if (lane >= (uint32_t)d) incl += v;
            }
            const uint32_t total = __shfl_sync(FULL, incl, 31);
            if (total >= POOL_OWNED_REFILL || (total > 0u && n_in < 16u && n_pk < 16u)) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(claim, total);
                base = __shfl_sync(FULL, base, 0);
                if (base + total >= n) more = false;  // this claim reaches the end of the queue
                uint32_t idx = base + (incl - n_free_me);
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    if (((kinds >> (2 * k)) & 3u) != 0u || idx >= n) continue;
                    const uint32_t slot = (uint32_t)k * 32u + lane;
                    Ray r;
                    fetch(idx, r);
                    pool_store_ray(pw, slot, r, fixed);
                    pw.f(PC_BT, slot) = MPD_FLT_MAX, pw.f(PC_BU, slot) = 0.0f, pw.f(PC_BV, slot) = 0.0f, pw.u(PC_BTRI, slot) = MP_NO_HIT;
                    pw.u(PC_RID, slot) = idx;
                    tc.rays++;
                    ++idx;
                    Stack stk{nullptr, wstacks + slot, 0, 0};
                    const uint32_t kd = pool_next<COUNT>(pw, slot, stk, MPD_FLT_MAX, emit, tc, true, make_uint2(sc.root, __float_as_uint(MPD_NINF)));
                    kinds |= kd << (2 * k);
                }
                continue;
            }
        }
        if (n_in == 0u && n_pk == 0u) break;  // nothing pending; `more` is false here (96 free slots would have been refilled)

        if (n_in >= n_pk) {
            // ---- one inner node per lane that holds one (ray_bvh_intersection.rs:149-162) ----
            if (has_in) {
                int k = 0;
#pragma unroll
                for (int j = K - 1; j >= 0; --j)
                    if (((kinds >> (2 * j)) & 3u) == 1u) k = j;
                const uint32_t slot = (uint32_t)k * 32u + lane;
                Stack stk{nullptr, wstacks + slot, 0, 0};
                stk.unpack(pw.u(PC_SP, slot));
                const Ray r = pool_load_ray(pw, slot, fixed);
                const float bt = pw.f(PC_BT, slot);
                const uint32_t node = pw.u(PC_ITEM, slot) >> MP_LINK_COUNT_BITS;
                if (COUNT) tc.vis.inner++;
                const bool nanpath = ray_needs_nanpath(sc, r);
                RaySel sel{};
                if (LAYOUT == MP_LAYOUT_DECODED && !nanpath) sel = ray_sel(r);
                bool has_top = false;
                uint2 top = make_uint2(0u, 0u);
                if (ORDER == MP_ORDER_REFERENCE) {
                    PoolDeferSink<Stack> sink{stk, make_uint2(0u, 0u), false};
                    if (nanpath)
                        visit_inner<LAYOUT, true>(sc, r, sel, node, bt, sink);
                    else
                        visit_inner<LAYOUT, false>(sc, r, sel, node, bt, sink);
                    has_top = sink.has, top = sink.top;
                } else {
                    StackSink<ORDER, Stack> sink{stk, stk.sp, stk.sp};
                    if (nanpath)
                        visit_inner<LAYOUT, true>(sc, r, sel, node, bt, sink);
                    else
                        visit_inner<LAYOUT, false>(sc, r, sel, node, bt, sink);
                    stk.sp = sink.sp;
                }
                const uint32_t kd = pool_next<COUNT>(pw, slot, stk, bt, emit, tc, has_top, top);
                kinds = (kinds & ~(3u << (2 * k))) | (kd << (2 * k));
            }
        } else if (has_pk) {
            // ---- one triangle packet per lane that holds one (ray_bvh_intersection.rs:104-141) ----
            int k = 0;
#pragma unroll
            for (int j = K - 1; j >= 0; --j)
                if (((kinds >> (2 * j)) & 3u) == 2u) k = j;
            const uint32_t slot = (uint32_t)k * 32u + lane;
            const Ray r = pool_load_ray(pw, slot, fixed);
            const uint32_t item = pw.u(PC_ITEM, slot);
            uint32_t pidx = item >> MP_LINK_COUNT_BITS, rem = item & MP_LINK_COUNT_MASK;
            float bt = pw.f(PC_BT, slot);  // = the leaf's max_t: best.t is not updated inside the leaf (:106-107)
            Hit local{pw.f(PC_LT, slot), pw.f(PC_LU, slot), pw.f(PC_LV, slot), pw.u(PC_LTRI, slot)};
            visit_packet<LAYOUT>(sc, r, pidx, bt, local);
            ++pidx, --rem;
            if (any_hit && rem != 0 && local.t < bt) {  // occluded: skip the leaf's remaining packets
                if (COUNT) tc.vis.packets -= rem;
                rem = 0;
            }
            pw.u(PC_ITEM, slot) = pidx << MP_LINK_COUNT_BITS | rem;
            if (rem == 0) {
                Stack stk{nullptr, wstacks + slot, 0, 0};
                stk.unpack(pw.u(PC_SP, slot));
                if (local.t < bt) {  // :59
                    bt = local.t;
                    pw.f(PC_BT, slot) = local.t, pw.f(PC_BU, slot) = local.u, pw.f(PC_BV, slot) = local.v, pw.u(PC_BTRI, slot) = local.tri;
                    if (any_hit) stk.clear();
                }
                const uint32_t kd = pool_next<COUNT>(pw, slot, stk, bt, emit, tc, false, make_uint2(0u, 0u));
                kinds = (kinds & ~(3u << (2 * k))) | (kd << (2 * k));
            } else {
                pw.f(PC_LT, slot) = local.t, pw.f(PC_LU, slot) = local.u, pw.f(PC_LV, slot) = local.v, pw.u(PC_LTRI, slot) = local.tri;
            }
        }
    }
}

#ifndef POOL_LANE_OWNED
#define POOL_LANE_OWNED 0
#endif
#if POOL_LANE_OWNED
#define POOL_RUN pool_run_owned
#else
#define POOL_RUN pool_run
#endif

// ---- kernels: 128 threads = 4 pools per block, persistent grid ----

template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128, POOL_BLOCKS_FOR(LAYOUT)) trace_rays_pool_kernel(Scene sc, const float* __restrict__ rays, uint32_t n, uint4* __restrict__ hits,
                                                              DevCounters* dc, uint32_t* claim, uint2* scratch, uint32_t stack_cap) {
    extern __shared__ __align__(16) unsigned char pool_smem[];
    ThreadCounts tc{};
    POOL_RUN<LAYOUT, ORDER, COUNT>(
        sc, n, claim, false, scratch, stack_cap, pool_smem,
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
__global__ void __launch_bounds__(128, POOL_BLOCKS_FOR(LAYOUT)) wf_trace_pool_kernel(Scene sc, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count,
                                                            const float4* __restrict__ ray_o, const float4* __restrict__ ray_d,
                                                            uint4* __restrict__ hit, DevCounters* dc, uint32_t* claim, uint2* scratch,
                                                            uint32_t stack_cap) {
    extern __shared__ __align__(16) unsigned char pool_smem[];
    ThreadCounts tc{};
    POOL_RUN<LAYOUT, ORDER, COUNT>(
        sc, *count, claim, false, scratch, stack_cap, pool_smem,
        [&](uint32_t i, Ray& r) {
            const uint32_t path = queue[i];
            r = load_ray(ray_o[path], ray_d[path]);
        },
        [&](uint32_t i, const Hit& h) { hit[queue[i]] = make_uint4(h.tri, __float_as_uint(h.t), __float_as_uint(h.u), __float_as_uint(h.v)); }, tc);
    flush_counts(dc, tc, COUNT);
}

template <int LAYOUT, int ORDER, bool COUNT>
__global__ void __launch_bounds__(128, POOL_BLOCKS_FOR(LAYOUT)) wf_shadow_pool_kernel(Scene sc, const float4* __restrict__ sh_o, const uint32_t* __restrict__ sh_path,
                                                             const uint32_t* __restrict__ sh_count, float4* __restrict__ ray_d, float4* __restrict__ rgb,
                                                             const float2* __restrict__ sh_c, DevCounters* dc, uint32_t* claim, uint2* scratch,
                                                             uint32_t stack_cap) {
    extern __shared__ __align__(16) unsigned char pool_smem[];
    ThreadCounts tc{};
    const Ray sun = ray_new(f3(0.f, 0.f, 0.f), f3(EXT_SUN_X, EXT_SUN_Y, EXT_SUN_Z));
    POOL_RUN<LAYOUT, ORDER, COUNT>(
        sc, *sh_count, claim, true, scratch, stack_cap, pool_smem,
        [&](uint32_t i, Ray& r) {
            const float4 o = sh_o[i];
            r.o = f3(o.x, o.y, o.z), r.d = sun.d, r.inv = sun.inv;  // = ray_new(origin, sun)
            tc.shadow_rays++;
        },
        [&](uint32_t i, const Hit& h) {
            if (h.tri == MP_NO_HIT) add_sun(ray_d, rgb, sh_c, sh_path[i], i, sh_o[i].w);
        },
        tc, &sun);
    flush_counts(dc, tc, COUNT);
}

}  // namespace mpd

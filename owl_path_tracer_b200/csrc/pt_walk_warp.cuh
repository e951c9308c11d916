// pt_walk_warp.cuh — warp-cooperative closest-hit walk (device only).
//
// Same result as walk_closest (pt_bvh8.h): the closest Moller-Trumbore hit with ties to the
// smaller triangle id — a result that does not depend on visiting order, so this walker is
// free to reorder work for SIMD efficiency:
//   * persistent warps fetch rays from the queue with one atomic per refill; a lane whose
//     ray is finished is refilled as soon as kRefill lanes are idle (the first ncu capture of
//     the one-thread-per-ray kernel showed 4.7 of 32 lanes active per issued instruction:
//     traversal lengths differ by 10x inside a warp once camera and bounce rays are mixed);
//   * every iteration the warp votes for ONE phase, node step or triangle test, and only
//     the lanes with that kind of work run it; leaf triangles that lose the vote are pushed
//     on the lane's stack as a triangle group and come back later;
//   * the 8 child boxes of a node are tested without branches (empty slots have meta = 0
//     and contribute no bits).
#pragma once
#include "pt_bvh8.h"

namespace pt {

// Tunables (compile-time; tools/build_variants.sh builds A/B libraries from them).
#ifndef PT_REFILL
#define PT_REFILL 6          // refill when at least this many lanes are idle
#endif
#ifndef PT_TRI_VOTE
#define PT_TRI_VOTE 3        // triangle phase when PT_TRI_VOTE * (lanes with triangles) >= working lanes
#endif
#ifndef PT_TRI_LOOP
#define PT_TRI_LOOP 1        // stay in the triangle phase while the vote still holds (skips the loop head per triangle)
#endif
#ifndef PT_SUSPEND_LANES
#define PT_SUSPEND_LANES 3   // SUSPEND walks: once the queue is drained, a warp left with <= this many rays parks them
#endif
#ifndef PT_SUSPEND_MIN
#define PT_SUSPEND_MIN 100000 // ... only in launches of at least this many rays (smaller ones run every ray to its end)
#endif
#ifndef PT_SUSPEND_STEPS
#define PT_SUSPEND_STEPS 0   // ... and (if > 0) any warp parks whatever it still holds this many loop steps after the queue drained
#endif
#ifndef PT_TIMELINE
#define PT_TIMELINE 0        // development: per-warp timestamps of one chosen k_extend launch (tools/timeline.py)
#endif
#if PT_TIMELINE
__device__ unsigned long long* g_timeline = nullptr;  // 4 words per warp: start, queue drained, exit (globaltimer ns), rays << 32 | loop steps
__device__ int g_timeline_on = 0;
__device__ __forceinline__ unsigned long long timeline_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#endif
#ifndef PT_WALK_BLOCK
#define PT_WALK_BLOCK 128       // threads per block of the persistent walker kernels
#endif
#ifndef PT_SMEM_STACK
#define PT_SMEM_STACK 0         // traversal-stack entries per lane kept in shared memory (the rest in local memory)
#endif
constexpr int kRefill = PT_REFILL;
constexpr int kWalkStack = 2 * kStackSize;  // per tree level at most one node group + one postponed triangle group
constexpr int kSmemStack = PT_SMEM_STACK;

// The entries of a lane's traversal stack BELOW its top (the top lives in registers).  Entry e < kSmemStack sits in
// shared memory at [e][thread] (consecutive threads -> consecutive 8-byte words: conflict-free), deeper entries in the
// lane's local-memory array.  With kSmemStack == 0 this is the plain local array.
struct WalkStack {
    uint2* local;
#if PT_SMEM_STACK > 0
    uint2* shared;   // &s_stack[threadIdx.x]
#endif
    __device__ __forceinline__ uint2 ld(int e) const {
#if PT_SMEM_STACK > 0
        if (e < kSmemStack) return shared[e * PT_WALK_BLOCK];
        return local[e - kSmemStack];
#else
        return local[e];
#endif
    }
    __device__ __forceinline__ void st(int e, uint2 v) const {
#if PT_SMEM_STACK > 0
        if (e < kSmemStack) shared[e * PT_WALK_BLOCK] = v;
        else local[e - kSmemStack] = v;
#else
        local[e] = v;
#endif
    }
};

__device__ __forceinline__ u4 ldg4(const u4* p) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    return u4{v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ f4 ldg4f(const f4* p) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    return f4{v.x, v.y, v.z, v.w};
}

#ifndef PT_TRI_POOL
#define PT_TRI_POOL 1   // 1: the triangle phase pools the pending blocks of all lanes of the warp (see warp_walk);
                        // 0: every lane tests its own blocks, one per step (the walker up to profiles/r02o)
#endif

// The packed two-triangle test of warp_walk's triangle phase for an arbitrary ray: the better accepted hit of the
// block by (t, triangle id), id == kNoHit when neither triangle is hit.  Same operations, same bits.
struct BlockHit {
    float t, u, v;
    uint32_t id;
};
__device__ __forceinline__ BlockHit test_block(const TriPair* __restrict__ tp, const f3 o, const f3 d, const float tmin,
                                               const float tmax, uint32_t& n_tested) {
    const f4 w0 = ldg4f(&tp->w[0]), w1 = ldg4f(&tp->w[1]), w2 = ldg4f(&tp->w[2]), w3 = ldg4f(&tp->w[3]), w4 = ldg4f(&tp->w[4]);
    const float2 v0x = make_float2(w0.x, w0.y), v0y = make_float2(w0.z, w0.w), v0z = make_float2(w1.x, w1.y);
    const float2 e1x = make_float2(w1.z, w1.w), e1y = make_float2(w2.x, w2.y), e1z = make_float2(w2.z, w2.w);
    const float2 e2x = make_float2(w3.x, w3.y), e2y = make_float2(w3.z, w3.w), e2z = make_float2(w4.x, w4.y);
    const float2 dx = make_float2(d.x, d.x), dy = make_float2(d.y, d.y), dz = make_float2(d.z, d.z);
    const float2 ndx = make_float2(-d.x, -d.x), ndy = make_float2(-d.y, -d.y), ndz = make_float2(-d.z, -d.z);
    const float2 px = __ffma2_rn(dy, e2z, __fmul2_rn(ndz, e2y));
    const float2 py = __ffma2_rn(dz, e2x, __fmul2_rn(ndx, e2z));
    const float2 pz = __ffma2_rn(dx, e2y, __fmul2_rn(ndy, e2x));
    const float2 det = __ffma2_rn(e1z, pz, __ffma2_rn(e1y, py, __fmul2_rn(e1x, px)));
    const float2 inv = make_float2(1.0f / det.x, 1.0f / det.y);
    const float2 sx = __fadd2_rn(make_float2(o.x, o.x), make_float2(-v0x.x, -v0x.y));
    const float2 sy = __fadd2_rn(make_float2(o.y, o.y), make_float2(-v0y.x, -v0y.y));
    const float2 sz = __fadd2_rn(make_float2(o.z, o.z), make_float2(-v0z.x, -v0z.y));
    const float2 uu = __fmul2_rn(__ffma2_rn(sz, pz, __ffma2_rn(sy, py, __fmul2_rn(sx, px))), inv);
    const float2 qx = __ffma2_rn(sy, e1z, __fmul2_rn(make_float2(-sz.x, -sz.y), e1y));
    const float2 qy = __ffma2_rn(sz, e1x, __fmul2_rn(make_float2(-sx.x, -sx.y), e1z));
    const float2 qz = __ffma2_rn(sx, e1y, __fmul2_rn(make_float2(-sy.x, -sy.y), e1x));
    const float2 vv = __fmul2_rn(__ffma2_rn(dz, qz, __ffma2_rn(dy, qy, __fmul2_rn(dx, qx))), inv);
    const float2 tt = __fmul2_rn(__ffma2_rn(e2z, qz, __ffma2_rn(e2y, qy, __fmul2_rn(e2x, qx))), inv);
    const uint32_t idA = f2u(w4.z), idB = f2u(w4.w);
    n_tested = 1u + (idB != kNoHit ? 1u : 0u);
    const bool okA = (det.x != 0.0f) && (uu.x >= 0.0f && uu.x <= 1.0f) && (vv.x >= 0.0f && uu.x + vv.x <= 1.0f) &&
                     (tt.x > tmin && tt.x < tmax);
    const bool okB = (det.y != 0.0f) && (uu.y >= 0.0f && uu.y <= 1.0f) && (vv.y >= 0.0f && uu.y + vv.y <= 1.0f) &&
                     (tt.y > tmin && tt.y < tmax);
    BlockHit h{0.0f, 0.0f, 0.0f, kNoHit};
    if (okA) h = BlockHit{tt.x, uu.x, vv.x, idA};
    if (okB && (!okA || tt.y < h.t || (tt.y == h.t && idB < h.id))) h = BlockHit{tt.y, uu.y, vv.y, idB};
    return h;
}

// magic: 0x3F800000 handed in as a RUNTIME value (SceneView::plane_magic, a kernel parameter), so that
//        it sits in a register and the byte-permute selector of plane_q is the immediate.
// ANY:  shadow rays — the first accepted triangle ends the ray (any hit in (tmin, tmax)); the hit record then
//       only says whether something was hit.
// SUSPEND: bounded tails.  A launch ends with every warp waiting for the longest of its ~32 resident rays while
//       its other lanes idle (traversal lengths are heavy-tailed: the last ray of a warp typically takes 5-10x
//       the mean).  With SUSPEND, a warp that is out of new rays and down to PT_SUSPEND_LANES rays parks them:
//       traversal state (closest hit so far, pending groups, stack) goes to a per-lane save area, the hit
//       record says kPending, and the ray continues in the NEXT launch of this walker, refilled into a full
//       warp like any other ray (k_shade passes a pending lane through untouched).  The result of a walk
//       does not depend on where it was interrupted.
// Src:  bool load(uint32_t i, f3& o, f3& d, float& tmin, float& tmax)   (i < n); true = entry i continues a parked walk
//       void restore(i, hit, ng_base, ng_mask, tg_base, tg_mask, sp, top_b, top_m, stack)        (SUSPEND only)
// Sink: void store(uint32_t i, const Hit& h)
//       void suspend(i, hit, ng_base, ng_mask, tg_base, tg_mask, sp, top_b, top_m, stack)        (SUSPEND only)
template <bool COUNT, bool ANY, bool SUSPEND, class Src, class Sink>
__device__ __forceinline__ void warp_walk(const Node8* __restrict__ nodes, const TriPair* __restrict__ tris, uint32_t n,
                                          uint32_t* cursor, Src src, Sink sink, WalkCounters& wc, const uint32_t magic) {
    const unsigned FULL = 0xffffffffu;
    const int lid = threadIdx.x & 31;

    bool has_ray = false, exhausted = false;
    // 0: only an empty warp leaves
    const int park = (SUSPEND && n >= (uint32_t)PT_SUSPEND_MIN) ? PT_SUSPEND_LANES : 0;
    int slice = PT_SUSPEND_STEPS;  // time slice left once the queue is drained (warp-uniform)
    uint32_t ray_index = 0;
    f3 o{0, 0, 0}, d{0, 0, 1};
    float tmin = 0, tmax = 0;
    float idx = 0, idy = 0, idz = 0;
    uint32_t octinv4 = 0;
    Hit hit{0, 0, 0, kNoHit};
    uint32_t ng_base = 0, ng_mask = 0, tg_base = 0, tg_mask = 0;
    // Traversal stack of (base, mask) pairs.  The TOP entry lives in registers (top_b, top_m) and only
    // the entries below it in local memory: a pop hands out the registers at once and starts the
    // load of the next entry, which is not needed before the following pop or push — a whole node or
    // triangle step later (the pop used to wait for its own local load: 10 % of all stall samples).
    uint2 stack_local[kWalkStack - kSmemStack];
#if PT_SMEM_STACK > 0
    __shared__ uint2 s_stack[PT_SMEM_STACK * PT_WALK_BLOCK];
    const WalkStack stack{stack_local, s_stack + threadIdx.x};
#else
    const WalkStack stack{stack_local};
#endif
    uint32_t top_b = 0, top_m = 0;
    int sp = 0;
#define PT_PUSH(B, M)                                         \
    do {                                                      \
        if (sp > 0) stack.st(sp - 1, make_uint2(top_b, top_m)); \
        top_b = (B);                                          \
        top_m = (M);                                          \
        ++sp;                                                 \
    } while (0)

#if PT_TIMELINE
    const bool tl_on = SUSPEND && g_timeline_on != 0;
    unsigned long long tl_start = 0, tl_drained = 0;
    uint32_t tl_rays = 0, tl_steps = 0, tl_node_lanes = 0, tl_tri_lanes = 0, tl_node_steps = 0;
    if (tl_on) tl_start = timeline_now();
#endif
    for (;;) {
#if PT_TIMELINE
        ++tl_steps;
#endif
        // ---- every lane with a ray must hold some work: pop the stack, or retire the ray
        if (has_ray && ng_mask <= 0x00FFFFFFu && tg_mask == 0) {
            if (sp == 0) {
                if (hit.tri == kNoHit) hit.t = tmax;
                sink.store(ray_index, hit);
                has_ray = false;
            } else {
                if (top_m > 0x00FFFFFFu) {
                    ng_base = top_b;
                    ng_mask = top_m;
                } else {
                    tg_base = top_b;
                    tg_mask = top_m;
                }
                --sp;
                if (sp > 0) {
                    const uint2 nxt = stack.ld(sp - 1);
                    top_b = nxt.x;
                    top_m = nxt.y;
                }
            }
        }
        unsigned act = __ballot_sync(FULL, has_ray);
        // ---- refill idle lanes (also the ones that retired just above)
        if (exhausted) {
            // out of new rays: run what is left; a SUSPEND walk parks the last few rays of the warp instead
            if constexpr (SUSPEND) {
                bool leave = __popc(act) <= park;
                if (PT_SUSPEND_STEPS > 0 && park > 0 && --slice < 0) leave = true;
                if (leave && __all_sync(FULL, !has_ray || sp <= kSuspStack + 1)) {
                    if (has_ray) sink.suspend(ray_index, hit, ng_base, ng_mask, tg_base, tg_mask, sp, top_b, top_m, stack);
                    break;
                }
            }
        } else if (__popc(~act) >= kRefill) {
            const unsigned idle = ~act;
            const int cnt = __popc(idle);
            uint32_t base = 0;
            if (lid == 0) base = atomicAdd(cursor, (uint32_t)cnt);
            base = __shfl_sync(FULL, base, 0);
            if (!has_ray) {
                const uint32_t i = base + (uint32_t)__popc(idle & ((1u << lid) - 1u));
                if (i < n) {
                    const bool resumed = src.load(i, o, d, tmin, tmax);
                    ray_index = i;
                    has_ray = true;
                    idx = safe_rcp_dir(d.x);
                    idy = safe_rcp_dir(d.y);
                    idz = safe_rcp_dir(d.z);
                    octinv4 = ((d.x < 0.0f ? 0u : 4u) | (d.y < 0.0f ? 0u : 2u) | (d.z < 0.0f ? 0u : 1u)) * 0x01010101u;
                    hit = Hit{tmax, 0.0f, 0.0f, kNoHit};
                    ng_base = 0;
                    ng_mask = 0x80000000u;
                    tg_mask = 0;
                    sp = 0;
                    if constexpr (SUSPEND) {
                        if (__builtin_expect(resumed, 0)) src.restore(i, hit, ng_base, ng_mask, tg_base, tg_mask, sp, top_b, top_m, stack);
                    }
                }
            }
            if (base + (uint32_t)cnt >= n) exhausted = true;
#if PT_TIMELINE
            if (tl_on) {
                tl_rays += (uint32_t)__popc(__ballot_sync(FULL, has_ray) & idle);
                if (exhausted) tl_drained = timeline_now();
            }
#endif
            act = __ballot_sync(FULL, has_ray);
        }
        if (act == 0) {
            if (exhausted) break;
            continue;
        }
        // ---- vote: triangle phase when at least a third of the working lanes hold triangles.  (A lane
        // with a ray always holds node work or triangles here, so "nobody has node work" already implies
        // that every working lane votes for triangles.)
        const bool want_tri = has_ray && tg_mask != 0;
        const bool can_node = has_ray && ng_mask > 0x00FFFFFFu;
        const int n_tri = __popc(__ballot_sync(FULL, want_tri));
        const bool tri_phase = PT_TRI_VOTE * n_tri >= __popc(act);

#if PT_TIMELINE
        if (tri_phase) tl_tri_lanes += (uint32_t)n_tri;
        else {
            tl_node_lanes += (uint32_t)__popc(__ballot_sync(FULL, can_node));
            ++tl_node_steps;
        }
#endif
        if (tri_phase) {
#if PT_TRI_POOL
            // Pooled triangle phase: the pending blocks of ALL lanes are dealt out over the 32 lanes, whoever owns
            // them (a lane tests block k of the warp with its owner's ray, fetched by shuffles), so a phase that
            // found 12 lanes with two or three blocks each runs as one full step instead of three thinning ones.
            // The closest hit of a ray is the minimum of (t, triangle id) over its blocks whatever the order: owners
            // publish their current (t, id) as one 64-bit key in shared memory, testers atomicMin theirs in, the
            // tester that ends up holding an owner's key hands over (u, v).  Voting by the number of pending blocks
            // (16 / 24 / 32) or leaving a thin last round for the next phase were both slower than this lane vote
            // followed by all rounds (profiles/r02s_*).
            {
                __shared__ unsigned long long s_key[PT_WALK_BLOCK];
                __shared__ float s_u[PT_WALK_BLOCK], s_v[PT_WALK_BLOCK];
                const int wbase = (int)(threadIdx.x & ~31u);
                const uint32_t my_mask = has_ray ? tg_mask : 0u;
                const int c = __popc(my_mask);
                int incl = c;
#pragma unroll
                for (int dl = 1; dl < 32; dl <<= 1) {
                    const int up = __shfl_up_sync(FULL, incl, dl);
                    if (lid >= dl) incl += up;
                }
                const int excl = incl - c;
                const int total = __shfl_sync(FULL, incl, 31);
                const unsigned long long my_key = ((unsigned long long)__float_as_uint(hit.t) << 32) | hit.tri;
                s_key[wbase + lid] = my_key;
                __syncwarp();
                for (int k0 = 0; k0 < total; k0 += 32) {
                    const int k = k0 + lid;
                    int own = 0;  // first lane whose inclusive count exceeds k
#pragma unroll
                    for (int step = 16; step >= 1; step >>= 1) {
                        const int probe = __shfl_sync(FULL, incl, own + step - 1);
                        if (probe <= k) own += step;
                    }
                    const int o_excl = __shfl_sync(FULL, excl, own);
                    const uint32_t o_mask = __shfl_sync(FULL, my_mask, own);
                    const uint32_t o_base = __shfl_sync(FULL, tg_base, own);
                    const f3 bo{__shfl_sync(FULL, o.x, own), __shfl_sync(FULL, o.y, own), __shfl_sync(FULL, o.z, own)};
                    const f3 bd{__shfl_sync(FULL, d.x, own), __shfl_sync(FULL, d.y, own), __shfl_sync(FULL, d.z, own)};
                    const float btmin = __shfl_sync(FULL, tmin, own), btmax = __shfl_sync(FULL, tmax, own);
                    unsigned long long key = ~0ull;
                    float bu = 0.0f, bv = 0.0f;
                    if (k < total) {
                        // the owner's (k - o_excl + 1)-th pending block from the bottom: any one-to-one assignment of
                        // blocks to testers will do (a loop clearing top bits was 3.5 % of the kernel's instructions)
                        const int bit = (int)__fns(o_mask, 0u, k - o_excl + 1);
                        uint32_t n_tested;
                        const BlockHit h = test_block(tris + (o_base + (uint32_t)bit), bo, bd, btmin, btmax, n_tested);
                        if (COUNT) wc.tris += n_tested;
                        if (h.id != kNoHit) {
                            key = ((unsigned long long)__float_as_uint(h.t) << 32) | h.id;
                            bu = h.u;
                            bv = h.v;
                            atomicMin(&s_key[wbase + own], key);
                        }
                    }
                    __syncwarp();
                    if (key != ~0ull && s_key[wbase + own] == key) {
                        s_u[wbase + own] = bu;
                        s_v[wbase + own] = bv;
                    }
                    __syncwarp();
                }
                if (c > 0) {
                    const unsigned long long best = s_key[wbase + lid];
                    if (best != my_key) {
                        hit.t = __uint_as_float((uint32_t)(best >> 32));
                        hit.tri = (uint32_t)best;
                        hit.u = s_u[wbase + lid];
                        hit.v = s_v[wbase + lid];
                    }
                    tg_mask = 0;
                    if (ANY && hit.tri != kNoHit) {  // occluded: drop the rest of the traversal
                        ng_mask = 0;
                        sp = 0;
                    }
                }
                __syncwarp();
            }
#else
#if PT_TRI_LOOP
            const int n_act = __popc(act);
            bool again;
            do {
#endif
            if (has_ray && tg_mask != 0) {
                const int bit = 31 - __clz(tg_mask);
                tg_mask &= ~(1u << bit);
                const TriPair* tp = tris + (tg_base + (uint32_t)bit);
                const f4 w0 = ldg4f(&tp->w[0]), w1 = ldg4f(&tp->w[1]), w2 = ldg4f(&tp->w[2]), w3 = ldg4f(&tp->w[3]),
                         w4 = ldg4f(&tp->w[4]);
                // Both triangles of the block at once: every float2 holds (triangle A, triangle B) and every
                // operation is the packed form of the scalar one in ray_triangle (pt_bvh8.h) — fma.rn.f32x2 /
                // mul.rn.f32x2 / add.rn.f32x2 round each half like the scalar instruction, so the bits are
                // those of two scalar tests (conditions evaluated together, no early outs).
                const float2 v0x = make_float2(w0.x, w0.y), v0y = make_float2(w0.z, w0.w), v0z = make_float2(w1.x, w1.y);
                const float2 e1x = make_float2(w1.z, w1.w), e1y = make_float2(w2.x, w2.y), e1z = make_float2(w2.z, w2.w);
                const float2 e2x = make_float2(w3.x, w3.y), e2y = make_float2(w3.z, w3.w), e2z = make_float2(w4.x, w4.y);
                const float2 dx = make_float2(d.x, d.x), dy = make_float2(d.y, d.y), dz = make_float2(d.z, d.z);
                const float2 ndx = make_float2(-d.x, -d.x), ndy = make_float2(-d.y, -d.y), ndz = make_float2(-d.z, -d.z);
                // p = d x e2        (a x b).x = fma(a.y, b.z, -(a.z b.y)),  -(a.z b.y) = (-a.z) b.y exactly
                const float2 px = __ffma2_rn(dy, e2z, __fmul2_rn(ndz, e2y));
                const float2 py = __ffma2_rn(dz, e2x, __fmul2_rn(ndx, e2z));
                const float2 pz = __ffma2_rn(dx, e2y, __fmul2_rn(ndy, e2x));
                const float2 det = __ffma2_rn(e1z, pz, __ffma2_rn(e1y, py, __fmul2_rn(e1x, px)));
                const float2 inv = make_float2(1.0f / det.x, 1.0f / det.y);
                // s = o - v0
                const float2 sx = __fadd2_rn(make_float2(o.x, o.x), make_float2(-v0x.x, -v0x.y));
                const float2 sy = __fadd2_rn(make_float2(o.y, o.y), make_float2(-v0y.x, -v0y.y));
                const float2 sz = __fadd2_rn(make_float2(o.z, o.z), make_float2(-v0z.x, -v0z.y));
                const float2 uu = __fmul2_rn(__ffma2_rn(sz, pz, __ffma2_rn(sy, py, __fmul2_rn(sx, px))), inv);
                // q = s x e1
                const float2 qx = __ffma2_rn(sy, e1z, __fmul2_rn(make_float2(-sz.x, -sz.y), e1y));
                const float2 qy = __ffma2_rn(sz, e1x, __fmul2_rn(make_float2(-sx.x, -sx.y), e1z));
                const float2 qz = __ffma2_rn(sx, e1y, __fmul2_rn(make_float2(-sy.x, -sy.y), e1x));
                const float2 vv = __fmul2_rn(__ffma2_rn(dz, qz, __ffma2_rn(dy, qy, __fmul2_rn(dx, qx))), inv);
                const float2 tt = __fmul2_rn(__ffma2_rn(e2z, qz, __ffma2_rn(e2y, qy, __fmul2_rn(e2x, qx))), inv);
                const uint32_t idA = f2u(w4.z), idB = f2u(w4.w);
                if (COUNT) wc.tris += 1u + (idB != kNoHit ? 1u : 0u);
                const bool okA = (det.x != 0.0f) && (uu.x >= 0.0f && uu.x <= 1.0f) && (vv.x >= 0.0f && uu.x + vv.x <= 1.0f) &&
                                 (tt.x > tmin && tt.x < tmax);
                const bool okB = (det.y != 0.0f) && (uu.y >= 0.0f && uu.y <= 1.0f) && (vv.y >= 0.0f && uu.y + vv.y <= 1.0f) &&
                                 (tt.y > tmin && tt.y < tmax);
                if (okA && (tt.x < hit.t || (tt.x == hit.t && idA < hit.tri))) {
                    hit.t = tt.x;
                    hit.u = uu.x;
                    hit.v = vv.x;
                    hit.tri = idA;
                }
                if (okB && (tt.y < hit.t || (tt.y == hit.t && idB < hit.tri))) {
                    hit.t = tt.y;
                    hit.u = uu.y;
                    hit.v = vv.y;
                    hit.tri = idB;
                }
                if (ANY && hit.tri != kNoHit) {  // occluded: drop the rest of the traversal
                    ng_mask = 0;
                    tg_mask = 0;
                    sp = 0;
                }
            }
#if PT_TRI_LOOP
            // same vote as above, without the trip through the loop head (no lane gains work in here)
            const int n_more = __popc(__ballot_sync(FULL, has_ray && tg_mask != 0));
            again = PT_TRI_VOTE * n_more >= n_act;
#if PT_TIMELINE
            if (again) {
                tl_tri_lanes += (uint32_t)n_more;
                ++tl_steps;
            }
#endif
            } while (again);
#endif
#endif  // PT_TRI_POOL
        } else if (can_node) {
            if (tg_mask != 0) {  // lost the vote: keep the leaf triangles for later
                PT_PUSH(tg_base, tg_mask);
                tg_mask = 0;
            }
            const int bit = 31 - __clz(ng_mask);
            const uint32_t imask = ng_mask & 0xFFu;
            ng_mask &= ~(1u << bit);
            if (ng_mask > 0x00FFFFFFu) PT_PUSH(ng_base, ng_mask);
            const uint32_t slot = (uint32_t)(bit - 24) ^ (octinv4 & 0xFFu);
            const uint32_t rel = (uint32_t)__popc(imask & ~(0xFFFFFFFFu << slot));
            const Node8* np = nodes + (ng_base + rel);
            const u4 w0 = ldg4(&np->w[0]), w1 = ldg4(&np->w[1]), w2 = ldg4(&np->w[2]), w3 = ldg4(&np->w[3]),
                     w4 = ldg4(&np->w[4]);
            if (COUNT) wc.nodes++;
            const uint32_t e = w0.w;
            float aix, aiy, aiz, aox, aoy, aoz;
            slab_axis(e & 0xFFu, u2f(w0.x), o.x, idx, aix, aox);
            slab_axis((e >> 8) & 0xFFu, u2f(w0.y), o.y, idy, aiy, aoy);
            slab_axis((e >> 16) & 0xFFu, u2f(w0.z), o.z, idz, aiz, aoz);
            const bool nx = d.x < 0.0f, ny = d.y < 0.0f, nz = d.z < 0.0f;
            const float2 aix2 = make_float2(aix, aix), aiy2 = make_float2(aiy, aiy), aiz2 = make_float2(aiz, aiz);
            const float2 aox2 = make_float2(aox, aox), aoy2 = make_float2(aoy, aoy), aoz2 = make_float2(aoz, aoz);
            uint32_t hitmask = 0;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const uint32_t meta4 = half ? w1.w : w1.z;
                const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xFFu;
                const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
                const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
                const uint32_t qlox = half ? w2.y : w2.x, qloy = half ? w2.w : w2.z, qloz = half ? w3.y : w3.x;
                const uint32_t qhix = half ? w3.w : w3.z, qhiy = half ? w4.y : w4.x, qhiz = half ? w4.w : w4.z;
                const uint32_t nearx = nx ? qhix : qlox, farx = nx ? qlox : qhix;
                const uint32_t neary = ny ? qhiy : qloy, fary = ny ? qloy : qhiy;
                const uint32_t nearz = nz ? qhiz : qloz, farz = nz ? qloz : qhiz;
#pragma unroll
                for (int j = 0; j < 4; j += 2) {
                    // two children per instruction: fma.rn.f32x2 (FFMA2, new on sm_100) does both
                    // slab products of an axis with one issue slot; same IEEE result per element
                    const float2 tnx = __ffma2_rn(make_float2(plane_q(nearx, j, magic), plane_q(nearx, j + 1, magic)), aix2, aox2);
                    const float2 tny = __ffma2_rn(make_float2(plane_q(neary, j, magic), plane_q(neary, j + 1, magic)), aiy2, aoy2);
                    const float2 tnz = __ffma2_rn(make_float2(plane_q(nearz, j, magic), plane_q(nearz, j + 1, magic)), aiz2, aoz2);
                    const float2 tfx = __ffma2_rn(make_float2(plane_q(farx, j, magic), plane_q(farx, j + 1, magic)), aix2, aox2);
                    const float2 tfy = __ffma2_rn(make_float2(plane_q(fary, j, magic), plane_q(fary, j + 1, magic)), aiy2, aoy2);
                    const float2 tfz = __ffma2_rn(make_float2(plane_q(farz, j, magic), plane_q(farz, j + 1, magic)), aiz2, aoz2);
                    const float tn0 = fmaxf(fmaxf(tnx.x, tny.x), fmaxf(tnz.x, tmin));
                    const float tf0 = fminf(fminf(tfx.x, tfy.x), fminf(tfz.x, hit.t));
                    const float tn1 = fmaxf(fmaxf(tnx.y, tny.y), fmaxf(tnz.y, tmin));
                    const float tf1 = fminf(fminf(tfx.y, tfy.y), fminf(tfz.y, hit.t));
                    const uint32_t cb0 = (child_bits4 >> (8 * j)) & 0xFFu, bi0 = (bit_index4 >> (8 * j)) & 0xFFu;
                    const uint32_t cb1 = (child_bits4 >> (8 * j + 8)) & 0xFFu, bi1 = (bit_index4 >> (8 * j + 8)) & 0xFFu;
                    hitmask |= (tn0 <= tf0) ? (cb0 << bi0) : 0u;
                    hitmask |= (tn1 <= tf1) ? (cb1 << bi1) : 0u;
                }
            }
            ng_base = w1.x;
            ng_mask = (hitmask & 0xFF000000u) | (e >> 24);
            tg_base = w1.y;
            tg_mask = hitmask & 0x00FFFFFFu;
        }
    }
#undef PT_PUSH
#if PT_TIMELINE
    if (tl_on && lid == 0) {
        const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
        g_timeline[4 * w + 0] = tl_start;
        g_timeline[4 * w + 1] = tl_drained;
        g_timeline[4 * w + 2] = timeline_now();
        g_timeline[4 * w + 3] = ((unsigned long long)tl_rays << 32) | tl_steps;
        g_timeline[4 * 65536 + 2 * w + 0] = ((unsigned long long)tl_node_lanes << 32) | tl_tri_lanes;
        g_timeline[4 * 65536 + 2 * w + 1] = tl_node_steps;
    }
#endif
}

}  // namespace pt

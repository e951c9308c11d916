// pt_walk_batch.cuh — warp walker with DEFERRED, BATCHED triangle tests (build option PT_BATCH_TRI, device only).
//
// Same contract and same result as warp_walk (pt_walk_warp.cuh): the closest Moller-Trumbore hit, ties to the smaller
// triangle id — which does not depend on the order in which boxes and triangles are visited.  What changes is who tests
// the triangles.  In the voted walker a triangle step serves the ~11 lanes of 32 that hold a leaf, and triangle steps
// are 46 % of all steps (tools/timeline.py).  Here a node step does not test its leaf blocks: the lane appends
// (owner lane, block) entries to a per-warp list in shared memory and goes on walking.  Once 32 entries wait, ALL 32
// lanes test one entry each — for whichever ray owns it, ray data and current best hit live in shared memory — and
// fold the results into the owners' hits with a 64-bit atomicMin on (t bits, triangle id): exactly the order
// "smaller t, then smaller id" of the scalar rule.  The price is that an owner's hit.t is a few node steps stale.
#pragma once
#include "pt_walk_warp.cuh"

namespace pt {

#ifndef PT_WALK_BLOCK
#define PT_WALK_BLOCK 128
#endif
#ifndef PT_BATCH_APPEND
#define PT_BATCH_APPEND 3    // blocks a lane may append per loop step (the rest of a big leaf group waits a step)
#endif
#ifndef PT_BATCH_FLUSH
#define PT_BATCH_FLUSH 32    // test triangles once this many entries wait
#endif
#ifndef PT_BATCH_MIN_NODE
#define PT_BATCH_MIN_NODE 1  // ... or as soon as fewer lanes than this can do a node step
#endif
constexpr int kBatchList = 32 + 32 * PT_BATCH_APPEND;

template <bool COUNT, bool ANY, bool SUSPEND, class Src, class Sink>
__device__ __forceinline__ void warp_walk_batch(const Node8* __restrict__ nodes, const TriPair* __restrict__ tris, uint32_t n,
                                                uint32_t* cursor, Src src, Sink sink, WalkCounters& wc, const uint32_t magic,
                                                const bool allow_park = true) {
    constexpr int kWarps = PT_WALK_BLOCK / 32;
    __shared__ uint32_t s_list[kWarps][kBatchList];        // (owner lane << 27) | triangle-pair block
    __shared__ unsigned long long s_best[kWarps][32];      // per lane: (t bits << 32) | triangle id of its best hit
    __shared__ float2 s_uv[kWarps][32];                    // barycentrics of that hit (written by the winning tester)
    __shared__ float4 s_ray[kWarps][32][2];                // origin, tmin | direction, tmax
    __shared__ uint32_t s_pend[kWarps][32];                // entries of this lane still waiting in the list
    const unsigned FULL = 0xffffffffu;
    const int lid = threadIdx.x & 31;
    const int w = (threadIdx.x >> 5) % kWarps;

    bool has_ray = false, exhausted = false;
    const int park = (SUSPEND && allow_park && n >= (uint32_t)PT_SUSPEND_MIN) ? PT_SUSPEND_LANES : 0;
    uint32_t ray_index = 0;
    f3 o{0, 0, 0}, d{0, 0, 1};
    float tmin = 0, tmax = 0;
    float idx = 0, idy = 0, idz = 0;
    uint32_t octinv4 = 0;
    Hit hit{0, 0, 0, kNoHit};
    unsigned long long mykey = ~0ull;
    uint32_t pend = 0;
    int n_list = 0;  // warp-uniform
    uint32_t ng_base = 0, ng_mask = 0, tg_base = 0, tg_mask = 0;
    uint2 stack[kWalkStack];
    uint32_t top_b = 0, top_m = 0;
    int sp = 0;
#define PT_PUSH(B, M)                                        \
    do {                                                     \
        if (sp > 0) stack[sp - 1] = make_uint2(top_b, top_m); \
        top_b = (B);                                         \
        top_m = (M);                                         \
        ++sp;                                                \
    } while (0)

    for (;;) {
        // ---- a lane without node work and without blocks to append pops its stack; with an empty stack it retires
        // once none of its entries waits in the list any more
        if (has_ray && ng_mask <= 0x00FFFFFFu && tg_mask == 0) {
            if (sp == 0) {
                if (pend == 0) {
                    if (hit.tri == kNoHit) hit.t = tmax;
                    sink.store(ray_index, hit);
                    has_ray = false;
                }
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
                    const uint2 nxt = stack[sp - 1];
                    top_b = nxt.x;
                    top_m = nxt.y;
                }
            }
        }
        unsigned act = __ballot_sync(FULL, has_ray);
        bool flush = false;
        if (exhausted) {
            if constexpr (SUSPEND) {
                if (__popc(act) <= park) {
                    if (n_list > 0) {
                        flush = true;   // a parked lane must not leave entries behind
                    } else if (__all_sync(FULL, !has_ray || sp <= kSuspStack + 1)) {
                        if (has_ray) sink.suspend(ray_index, hit, ng_base, ng_mask, tg_base, tg_mask, sp, top_b, top_m, stack);
                        break;
                    }
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
                    mykey = ((unsigned long long)f2u(hit.t) << 32) | hit.tri;
                    pend = 0;
                    s_best[w][lid] = mykey;
                    s_pend[w][lid] = 0u;
                    s_ray[w][lid][0] = make_float4(o.x, o.y, o.z, tmin);
                    s_ray[w][lid][1] = make_float4(d.x, d.y, d.z, tmax);
                }
            }
            if (base + (uint32_t)cnt >= n) exhausted = true;
            act = __ballot_sync(FULL, has_ray);
        }
        if (act == 0) {
            if (exhausted) break;
            continue;
        }
        // ---- append: every lane hands up to PT_BATCH_APPEND of its leaf blocks to the list
        {
            // (nothing is appended while a full batch waits: the list holds at most 31 + 32 * PT_BATCH_APPEND entries)
            const int cnt = (n_list < 32 && has_ray && tg_mask != 0) ? min(__popc(tg_mask), PT_BATCH_APPEND) : 0;
            const unsigned b0 = __ballot_sync(FULL, (cnt & 1) != 0), b1 = __ballot_sync(FULL, (cnt & 2) != 0);
            if (b0 | b1) {
                const unsigned lt = (1u << lid) - 1u;
                int pos = n_list + __popc(b0 & lt) + 2 * __popc(b1 & lt);
                for (int k = 0; k < cnt; ++k) {
                    const int bit = 31 - __clz(tg_mask);
                    tg_mask &= ~(1u << bit);
                    s_list[w][pos++] = ((uint32_t)lid << 27) | (tg_base + (uint32_t)bit);
                }
                if (cnt) {
                    pend += (uint32_t)cnt;
                    s_pend[w][lid] = pend;
                }
                n_list += __popc(b0) + 2 * __popc(b1);
            }
        }
        const bool can_node = has_ray && tg_mask == 0 && ng_mask > 0x00FFFFFFu;
        const unsigned any_node = __ballot_sync(FULL, can_node);
        if (n_list >= PT_BATCH_FLUSH || (n_list > 0 && (flush || __popc(any_node) < PT_BATCH_MIN_NODE))) {
            // ---- batched triangle step: the last min(32, n_list) entries, one per lane
            __syncwarp();
            const int n_take = min(32, n_list), first = n_list - n_take;
            unsigned long long key = ~0ull;
            float2 uv = make_float2(0.0f, 0.0f);
            int owner = 0;
            if (lid < n_take) {
                const uint32_t e = s_list[w][first + lid];
                owner = (int)(e >> 27);
                const TriPair* tp = tris + (e & 0x07FFFFFFu);
                const f4 w0 = ldg4f(&tp->w[0]), w1 = ldg4f(&tp->w[1]), w2 = ldg4f(&tp->w[2]), w3 = ldg4f(&tp->w[3]),
                         w4 = ldg4f(&tp->w[4]);
                const float4 ra = s_ray[w][owner][0], rb = s_ray[w][owner][1];
                // both triangles of the block with the packed operations of warp_walk (same bits as two scalar tests)
                const float2 v0x = make_float2(w0.x, w0.y), v0y = make_float2(w0.z, w0.w), v0z = make_float2(w1.x, w1.y);
                const float2 e1x = make_float2(w1.z, w1.w), e1y = make_float2(w2.x, w2.y), e1z = make_float2(w2.z, w2.w);
                const float2 e2x = make_float2(w3.x, w3.y), e2y = make_float2(w3.z, w3.w), e2z = make_float2(w4.x, w4.y);
                const float2 dx = make_float2(rb.x, rb.x), dy = make_float2(rb.y, rb.y), dz = make_float2(rb.z, rb.z);
                const float2 ndx = make_float2(-rb.x, -rb.x), ndy = make_float2(-rb.y, -rb.y), ndz = make_float2(-rb.z, -rb.z);
                const float2 px = __ffma2_rn(dy, e2z, __fmul2_rn(ndz, e2y));
                const float2 py = __ffma2_rn(dz, e2x, __fmul2_rn(ndx, e2z));
                const float2 pz = __ffma2_rn(dx, e2y, __fmul2_rn(ndy, e2x));
                const float2 det = __ffma2_rn(e1z, pz, __ffma2_rn(e1y, py, __fmul2_rn(e1x, px)));
                const float2 inv = make_float2(1.0f / det.x, 1.0f / det.y);
                const float2 sx = __fadd2_rn(make_float2(ra.x, ra.x), make_float2(-v0x.x, -v0x.y));
                const float2 sy = __fadd2_rn(make_float2(ra.y, ra.y), make_float2(-v0y.x, -v0y.y));
                const float2 sz = __fadd2_rn(make_float2(ra.z, ra.z), make_float2(-v0z.x, -v0z.y));
                const float2 uu = __fmul2_rn(__ffma2_rn(sz, pz, __ffma2_rn(sy, py, __fmul2_rn(sx, px))), inv);
                const float2 qx = __ffma2_rn(sy, e1z, __fmul2_rn(make_float2(-sz.x, -sz.y), e1y));
                const float2 qy = __ffma2_rn(sz, e1x, __fmul2_rn(make_float2(-sx.x, -sx.y), e1z));
                const float2 qz = __ffma2_rn(sx, e1y, __fmul2_rn(make_float2(-sy.x, -sy.y), e1x));
                const float2 vv = __fmul2_rn(__ffma2_rn(dz, qz, __ffma2_rn(dy, qy, __fmul2_rn(dx, qx))), inv);
                const float2 tt = __fmul2_rn(__ffma2_rn(e2z, qz, __ffma2_rn(e2y, qy, __fmul2_rn(e2x, qx))), inv);
                const uint32_t idA = f2u(w4.z), idB = f2u(w4.w);
                if (COUNT) wc.tris += 1u + (idB != kNoHit ? 1u : 0u);
                const float rtmin = ra.w, rtmax = rb.w;
                const bool okA = (det.x != 0.0f) && (uu.x >= 0.0f && uu.x <= 1.0f) && (vv.x >= 0.0f && uu.x + vv.x <= 1.0f) &&
                                 (tt.x > rtmin && tt.x < rtmax);
                const bool okB = (det.y != 0.0f) && (uu.y >= 0.0f && uu.y <= 1.0f) && (vv.y >= 0.0f && uu.y + vv.y <= 1.0f) &&
                                 (tt.y > rtmin && tt.y < rtmax);
                // accepted t is positive, so its bit pattern orders like the value: min over (t bits, id) is the scalar rule
                const unsigned long long kA = okA ? (((unsigned long long)f2u(tt.x) << 32) | idA) : ~0ull;
                const unsigned long long kB = okB ? (((unsigned long long)f2u(tt.y) << 32) | idB) : ~0ull;
                key = kA;
                uv = make_float2(uu.x, vv.x);
                if (kB < kA) {
                    key = kB;
                    uv = make_float2(uu.y, vv.y);
                }
                if (key != ~0ull) atomicMin(&s_best[w][owner], key);
                atomicSub(&s_pend[w][owner], 1u);
            }
            __syncwarp();
            if (key != ~0ull && s_best[w][owner] == key) s_uv[w][owner] = uv;
            __syncwarp();
            n_list = first;
            if (has_ray) {
                const unsigned long long k = s_best[w][lid];
                if (k != mykey) {
                    mykey = k;
                    hit.t = u2f((uint32_t)(k >> 32));
                    hit.tri = (uint32_t)k;
                    const float2 b = s_uv[w][lid];
                    hit.u = b.x;
                    hit.v = b.y;
                }
                pend = s_pend[w][lid];
                if (ANY && hit.tri != kNoHit) {  // occluded: drop the rest of the traversal (entries still listed are tested in vain)
                    ng_mask = 0;
                    tg_mask = 0;
                    sp = 0;
                }
            }
        } else if (can_node) {
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
}

}  // namespace pt

// pt_bvh8.h — compressed 8-wide BVH: node layout, ray/triangle test, closest-hit walk.
//
// This replaces what the reference gets from the OptiX driver (optixAccelBuild +
// optixTrace on RT cores; call sites device.cu:133, application.cpp:131-140).  B200
// has no RT cores, so traversal is ordinary SM code and the tree is laid out for
// bytes: 80-byte nodes with 8-bit quantised child boxes (the layout of Ylitie,
// Karras & Laine, "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide
// BVHs", HPG 2017) and 48-byte triangles stored in node order.
//
// Result definition (the parity contract): the closest hit is the triangle with the
// smallest Moller-Trumbore t in (tmin, tmax); equal t is broken towards the smaller
// global triangle id.  That is independent of the tree and of the visiting order, so
// the scalar oracle (own binary BVH or brute force) must return the same triangle
// and the same (t,u,v) bits.  Child boxes are padded by the builder so that the
// rounding of the slab test can never cull a triangle the test would accept.
#pragma once
#include "pt_defs.h"

namespace pt {

struct alignas(16) u4 {
    uint32_t x, y, z, w;
};
struct alignas(16) f4 {
    float x, y, z, w;
};

PT_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } c;
    c.f = f;
    return c.u;
#endif
}
PT_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c;
    c.u = u;
    return c.f;
#endif
}
PT_HD int popc32(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __popc(v);
#else
    return __builtin_popcount(v);
#endif
}
PT_HD int msb32(uint32_t v) {  // index of highest set bit, v != 0
#if defined(__CUDA_ARCH__)
    return 31 - __clz(v);
#else
    return 31 - __builtin_clz(v);
#endif
}

// 80-byte node as five 16-byte words:
//  w0: px, py, pz, {ex, ey, ez, imask}
//  w1: child_base, tri_base, meta[0..3], meta[4..7]
//  w2: qlo_x[0..7], qlo_y[0..7]
//  w3: qlo_z[0..7], qhi_x[0..7]
//  w4: qhi_y[0..7], qhi_z[0..7]
// meta[i]: 0 = empty; internal child: (1<<5) | (24+i); leaf: (unary block count << 5) |
// offset of its first triangle-pair block from tri_base (1..3 blocks = 1..6 triangles -> 001, 011, 111).
struct Node8 {
    u4 w[5];
};
static_assert(sizeof(Node8) == 80, "Node8 is 80 bytes");

// 80-byte block of TWO triangles A, B (v0 + edges, prepared once so the walk does no index chasing), stored
// component-interleaved: every 16-byte word holds (A, B) pairs that the warp walker feeds to the packed
// fp32x2 instructions of sm_100 (FMUL2 / FFMA2 / FADD2) — one block test costs about what one triangle cost:
//  w0: v0.x_A v0.x_B v0.y_A v0.y_B     w1: v0.z_A v0.z_B e1.x_A e1.x_B     w2: e1.y_A e1.y_B e1.z_A e1.z_B
//  w3: e2.x_A e2.x_B e2.y_A e2.y_B     w4: e2.z_A e2.z_B id_A   id_B       (global triangle ids as bits)
// A leaf of n triangles owns ceil(n/2) consecutive blocks; the unused half of its last block has id = kNoHit
// and all-zero geometry (det = 0: never accepted).
struct TriPair {
    f4 w[5];
};
static_assert(sizeof(TriPair) == 80, "TriPair is 80 bytes");

struct Hit {
    float t, u, v;
    uint32_t tri;  // global triangle id, 0xFFFFFFFF = miss
};
constexpr uint32_t kNoHit = 0xFFFFFFFFu;
constexpr uint32_t kPending = 0xFFFFFFFEu;  // hit.tri of a parked walk (pt_walk_warp.cuh SUSPEND)
constexpr int kSuspStack = 16;               // a walk is parked only while at most this many entries sit below its stack top
constexpr int kStackSize = 32;  // node groups; the builder refuses deeper trees

// Moller-Trumbore with edges e1 = v1 - v0, e2 = v2 - v0.  Expression order is part of
// the parity contract (oracle/oracle.cpp ray_triangle mirrors it):
//   p = d x e2, det = e1 . p, inv = 1/det, s = o - v0, u = (s . p) inv,
//   q = s x e1, v = (d . q) inv, t = (e2 . q) inv;  accept 0<=u<=1, v>=0, u+v<=1,
//   tmin < t < tmax.  Barycentrics follow OptiX: P = (1-u-v) v0 + u v1 + v v2.
// Products are accumulated with EXPLICIT fused multiply-adds (fmaf: one IEEE rounding, the same on
// nvcc and g++ whatever the contraction flags): 30 floating-point instructions per test instead of
// 48 with separate multiplies and adds, and smaller rounding error in det / u / v / t.
//   a . b = fma(a.z, b.z, fma(a.y, b.y, a.x b.x))        (a x b).x = fma(a.y, b.z, -(a.z b.y)), ...
PT_HD float dot_fma(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
PT_HD f3 cross_fma(f3 a, f3 b) {
    return f3{fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x))};
}
PT_HD bool ray_triangle(f3 o, f3 d, f3 v0, f3 e1, f3 e2, float tmin, float tmax, float& t, float& u,
                        float& v) {
    f3 p = cross_fma(d, e2);
    float det = dot_fma(e1, p);
    if (!(det != 0.0f)) return false;
    float inv = 1.0f / det;
    f3 s = o - v0;
    float uu = dot_fma(s, p) * inv;
    if (!(uu >= 0.0f && uu <= 1.0f)) return false;
    f3 q = cross_fma(s, e1);
    float vv = dot_fma(d, q) * inv;
    if (!(vv >= 0.0f && uu + vv <= 1.0f)) return false;
    float tt = dot_fma(e2, q) * inv;
    if (!(tt > tmin && tt < tmax)) return false;
    t = tt;
    u = uu;
    v = vv;
    return true;
}

PT_HD float safe_rcp_dir(float d) {
    float a = fabsf(d) < 1e-30f ? (d < 0.0f ? -1e-30f : 1e-30f) : d;
    return 1.0f / a;
}

// Quantised plane byte j of a packed word as the float 1 + q * 2^-15, built by ONE byte permute
// (0x3F80'qq'00) instead of an int->float conversion: on sm_100 the conversion I2F.U8 runs on the
// quarter-rate XU pipe and 48 of them per node made that pipe the busiest one of the walker
// (profiles/r01_ncu_summary.md).  The slab test then evaluates
//     t = fma(1 + q 2^-15, 2^15 a, b - 2^15 a)  =  q a + b  (+ rounding of the constant term,
// at most 2^-9 of a quantisation step, covered by the builder's 1/64-step outward slack).
// `magic` must hold 0x3F800000; device callers fetch it through plane_magic() so that it lives in
// a register and the permute selector can be the instruction's immediate (with both as
// constants ptxas re-materialises the selector in front of every PRMT).
PT_HD uint32_t plane_magic() {
#if defined(__CUDA_ARCH__)
    uint32_t m;
    asm("mov.b32 %0, 0x3F800000;" : "=r"(m));
    return m;
#else
    return 0x3F800000u;
#endif
}
PT_HD float plane_q(uint32_t word, int j, uint32_t magic) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(word, magic, 0x7604u | ((uint32_t)j << 4)));
#else
    return u2f(magic | (((word >> (8 * j)) & 0xFFu) << 8));
#endif
}
// per axis: a2 = 2^15 * scale / d,  b2 = (p - o) / d - a2
PT_HD void slab_axis(uint32_t exp_byte, float p, float o, float inv_d, float& a2, float& b2) {
    a2 = u2f((exp_byte + 15u) << 23) * inv_d;
    b2 = (p - o) * inv_d - a2;
}

struct WalkCounters {
    uint32_t nodes, tris;
};

// Closest-hit walk.  COUNT=true also counts node fetches and triangle tests.
template <bool COUNT>
PT_HD void walk_closest(const Node8* __restrict__ nodes, const TriPair* __restrict__ tris, f3 o, f3 d,
                        float tmin, float tmax, Hit& hit, WalkCounters* wc) {
    hit.t = tmax;
    hit.u = 0.0f;
    hit.v = 0.0f;
    hit.tri = kNoHit;

    const float idx = safe_rcp_dir(d.x), idy = safe_rcp_dir(d.y), idz = safe_rcp_dir(d.z);
    const uint32_t octinv = (d.x < 0.0f ? 0u : 4u) | (d.y < 0.0f ? 0u : 2u) | (d.z < 0.0f ? 0u : 1u);

    const uint32_t magic = plane_magic();
    uint32_t stack_base[kStackSize];
    uint32_t stack_mask[kStackSize];
    int sp = 0;

    uint32_t ng_base = 0, ng_mask = 0x80000000u;  // node group: root
    for (;;) {
        uint32_t tg_base = 0, tg_mask = 0;
        if (ng_mask > 0x00FFFFFFu) {
            // pop the nearest (octant order) child of the current node group
            const int bit = msb32(ng_mask);
            const uint32_t imask = ng_mask & 0xFFu;
            ng_mask &= ~(1u << bit);
            if (ng_mask > 0x00FFFFFFu) {
                stack_base[sp] = ng_base;
                stack_mask[sp] = ng_mask;
                ++sp;
            }
            const uint32_t slot = (uint32_t)(bit - 24) ^ octinv;
            const uint32_t rel = (uint32_t)popc32(imask & ~(0xFFFFFFFFu << slot));
            const Node8* np = nodes + (ng_base + rel);
            const u4 w0 = np->w[0], w1 = np->w[1], w2 = np->w[2], w3 = np->w[3], w4 = np->w[4];
            if (COUNT) wc->nodes++;

            const uint32_t e = w0.w;
            float aix, aiy, aiz, aox, aoy, aoz;
            slab_axis(e & 0xFFu, u2f(w0.x), o.x, idx, aix, aox);
            slab_axis((e >> 8) & 0xFFu, u2f(w0.y), o.y, idy, aiy, aoy);
            slab_axis((e >> 16) & 0xFFu, u2f(w0.z), o.z, idz, aiz, aoz);
            uint32_t hitmask = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int j = i & 3, sh = j * 8;
                const uint32_t meta = ((i < 4 ? w1.z : w1.w) >> sh) & 0xFFu;
                if (meta == 0) continue;
                const float lox = plane_q(i < 4 ? w2.x : w2.y, j, magic), loy = plane_q(i < 4 ? w2.z : w2.w, j, magic);
                const float loz = plane_q(i < 4 ? w3.x : w3.y, j, magic), hix = plane_q(i < 4 ? w3.z : w3.w, j, magic);
                const float hiy = plane_q(i < 4 ? w4.x : w4.y, j, magic), hiz = plane_q(i < 4 ? w4.z : w4.w, j, magic);
                const float nx = d.x < 0.0f ? hix : lox, fx = d.x < 0.0f ? lox : hix;
                const float ny = d.y < 0.0f ? hiy : loy, fy = d.y < 0.0f ? loy : hiy;
                const float nz = d.z < 0.0f ? hiz : loz, fz = d.z < 0.0f ? loz : hiz;
                const float tn = fmaxf(fmaxf(fmaf(nx, aix, aox), fmaf(ny, aiy, aoy)),
                                       fmaxf(fmaf(nz, aiz, aoz), tmin));
                const float tf = fminf(fminf(fmaf(fx, aix, aox), fmaf(fy, aiy, aoy)),
                                       fminf(fmaf(fz, aiz, aoz), hit.t));
                if (tn <= tf) {
                    const bool inner = (meta & 0x18u) == 0x18u;
                    const uint32_t bidx = inner ? ((meta & 0x1Fu) ^ octinv) : (meta & 0x1Fu);
                    hitmask |= (meta >> 5) << bidx;
                }
            }
            ng_base = w1.x;
            ng_mask = (hitmask & 0xFF000000u) | (e >> 24);
            tg_base = w1.y;
            tg_mask = hitmask & 0x00FFFFFFu;
        } else {
            tg_base = ng_base;
            tg_mask = ng_mask;
            ng_base = 0;
            ng_mask = 0;
        }

        while (tg_mask != 0) {
            const int bit = msb32(tg_mask);
            tg_mask &= ~(1u << bit);
            const TriPair* tp = tris + (tg_base + (uint32_t)bit);
            const f4 w0 = tp->w[0], w1 = tp->w[1], w2 = tp->w[2], w3 = tp->w[3], w4 = tp->w[4];
            // the two triangles of the block one after the other (the warp walker does both at once with
            // packed instructions: same IEEE operations per triangle, same bits)
            for (int half = 0; half < 2; ++half) {
                const uint32_t id = f2u(half ? w4.w : w4.z);
                if (id == kNoHit) continue;
                if (COUNT) wc->tris++;
                const f3 v0 = half ? f3{w0.y, w0.w, w1.y} : f3{w0.x, w0.z, w1.x};
                const f3 e1 = half ? f3{w1.w, w2.y, w2.w} : f3{w1.z, w2.x, w2.z};
                const f3 e2 = half ? f3{w3.y, w3.w, w4.y} : f3{w3.x, w3.z, w4.x};
                float t, u, v;
                if (ray_triangle(o, d, v0, e1, e2, tmin, tmax, t, u, v)) {
                    if (t < hit.t || (t == hit.t && id < hit.tri)) {
                        hit.t = t;
                        hit.u = u;
                        hit.v = v;
                        hit.tri = id;
                    }
                }
            }
        }

        if (ng_mask <= 0x00FFFFFFu) {
            if (sp == 0) break;
            --sp;
            ng_base = stack_base[sp];
            ng_mask = stack_mask[sp];
        }
    }
    if (hit.tri == kNoHit) hit.t = tmax;
}

}  // namespace pt

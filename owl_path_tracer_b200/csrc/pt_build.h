// pt_build.h — per-element steps of the on-device BVH builder.
//
// Replaces optixAccelBuild as driven by the reference through OWL
// (application.cpp:131-140, extern/owl/owl/TrianglesGeomGroup.cpp:121-330).  Pipeline:
//   1. triangle boxes + scene box          (tri_prepare, one thread per triangle)
//   2. 63-bit Morton key of the centroid   (same pass)
//   3. radix sort of (key, triangle)       (cub::DeviceRadixSort in bvh_build.cu)
//   4. binary tree by parallel locally-ordered clustering over the Morton order
//      (ploc_nearest / ploc_flags / scan / ploc_emit, rounds until one cluster is left);
//      every new node gets its box, triangle count and collapse plan when it is created
//  4b. reinsertion passes over that tree (ri_find / ri_bid / ri_settle / ri_hold / ri_apply, then
//      refit_from_leaf for boxes, counts and plans): nodes move to where they cost the least area
//   5. cost-optimal collapse plan per binary node (plan_node, dynamic programme over
//      "subtree in at most i slots", Ylitie et al. 2017)
//   6. top-down collapse to 8-wide nodes following the plan, subtrees of <= 3
//      triangles may become leaves; octant-ordered slot assignment; 8-bit quantisation;
//      triangles re-emitted in node order   (collapse_node, level by level)
//   (radix_tree_node / refit_from_leaf build the plain LBVH tree instead; kept for the
//    builder comparison in tools/bvh_quality.py)
// Each step is a PT_HD function of the element index so that tests/emu can run the
// very same code sequentially on the host.
#pragma once
#include <cstring>

#include "pt_bvh8.h"

namespace pt {

struct BinNode {        // binary radix tree node (inner nodes 0..n-2)
    float lo[3], hi[3];
    int32_t left, right;  // >= 0 inner node index, < 0: ~leaf (sorted position)
    int32_t first, count; // sorted triangle range covered
    int32_t parent;
    int32_t pad;
};

PT_HD uint64_t spread21(uint32_t v) {  // 21 bits -> every third bit
    uint64_t x = v & 0x1FFFFFu;
    x = (x | (x << 32)) & 0x1F00000000FFFFull;
    x = (x | (x << 16)) & 0x1F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

PT_HD uint32_t quant21(float c, float lo, float inv_ext) {
    float f = (c - lo) * inv_ext;  // [0,1]
    f = fminf(fmaxf(f * 2097152.0f, 0.0f), 2097151.0f);
    return (uint32_t)f;
}

// Triangle i of the flattened scene: positions pos[3*v..], indices idx4[i] = (i0,i1,i2,mesh).
// Writes the padded box (lo/hi, 3 floats each).  `pad` is an absolute margin that
// covers the rounding of the slab test (see pt_bvh8.h).
PT_HD void tri_box(const f4* pos, const int32_t* idx4, uint32_t i, float pad, float* lo, float* hi) {
    const f4 a = pos[idx4[4 * i + 0]], b = pos[idx4[4 * i + 1]], c = pos[idx4[4 * i + 2]];
    lo[0] = fminf(a.x, fminf(b.x, c.x)) - pad;
    lo[1] = fminf(a.y, fminf(b.y, c.y)) - pad;
    lo[2] = fminf(a.z, fminf(b.z, c.z)) - pad;
    hi[0] = fmaxf(a.x, fmaxf(b.x, c.x)) + pad;
    hi[1] = fmaxf(a.y, fmaxf(b.y, c.y)) + pad;
    hi[2] = fmaxf(a.z, fmaxf(b.z, c.z)) + pad;
}

PT_HD uint64_t morton_key(const float* lo, const float* hi, const float* slo, const float* sinv) {
    uint32_t x = quant21(0.5f * (lo[0] + hi[0]), slo[0], sinv[0]);
    uint32_t y = quant21(0.5f * (lo[1] + hi[1]), slo[1], sinv[1]);
    uint32_t z = quant21(0.5f * (lo[2] + hi[2]), slo[2], sinv[2]);
    return (spread21(x) << 2) | (spread21(y) << 1) | spread21(z);
}

PT_HD int clz64(uint64_t v) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)v);
#else
    return v ? __builtin_clzll(v) : 64;
#endif
}

// common-prefix length of sorted keys i and j, ties broken by position (Karras 2012 §4)
PT_HD int prefix_len(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + (clz64((uint64_t)(uint32_t)(i ^ j)) - 32);
    return clz64(a ^ b);
}

// inner node i of the binary radix tree over n sorted keys (n >= 2)
PT_HD void radix_tree_node(const uint64_t* keys, int n, int i, BinNode* nodes, int32_t* leaf_parent) {
    const int d = (prefix_len(keys, n, i, i + 1) - prefix_len(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = prefix_len(keys, n, i, i - d);
    int lmax = 2;
    while (prefix_len(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (prefix_len(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = prefix_len(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (prefix_len(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + (d < 0 ? -1 : 0);
    const int first = i < j ? i : j, last = i < j ? j : i;
    BinNode& nd = nodes[i];
    nd.first = first;
    nd.count = last - first + 1;
    if (first == gamma) {
        nd.left = ~gamma;
        leaf_parent[gamma] = i;
    } else {
        nd.left = gamma;
        nodes[gamma].parent = i;
    }
    if (last == gamma + 1) {
        nd.right = ~(gamma + 1);
        leaf_parent[gamma + 1] = i;
    } else {
        nd.right = gamma + 1;
        nodes[gamma + 1].parent = i;
    }
    if (i == 0) nd.parent = -1;
}

PT_HD int32_t arrive(int32_t* flag) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(flag, 1);
#else
    int32_t old = *flag;
    *flag = old + 1;
    return old;
#endif
}
PT_HD uint32_t grab(uint32_t* counter, uint32_t n) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(counter, n);
#else
    uint32_t old = *counter;
    *counter = old + n;
    return old;
#endif
}
PT_HD float load_cg(const float* p) {
#if defined(__CUDA_ARCH__)
    return __ldcg(p);
#else
    return *p;
#endif
}
PT_HD int32_t load_cg_i(const int32_t* p) {
#if defined(__CUDA_ARCH__)
    return __ldcg(p);
#else
    return *p;
#endif
}
PT_HD void fence() {
#if defined(__CUDA_ARCH__)
    __threadfence();
#endif
}


// ---- cost-optimal collapse plan (Ylitie, Karras, Laine 2017, section 3.1) ------------------
// For every binary node n: c[i-1] = smallest SAH cost of representing the subtree of n with at
// most i child slots of a wide node (i = 1..7), and the decisions that reach it:
//   dec[0]    1 = as ONE slot the subtree is a leaf (<= kMaxLeafTris triangles = 3 pair blocks), 0 = an inner
//             wide node whose 8 slots are distributed over the two children by dec[7]
//   dec[j-1]  (j = 2..8) how `j` slots are used: 0 = no better than j-1 slots, k > 0 = left child
//             gets k slots, right child gets j-k
// Costs: kCostNode per wide-node visit, kCostTri per triangle-pair BLOCK test (pt_bvh8.h TriPair: two
// triangles are tested for the price of one), both weighted by the box half-area.  The frame time is
// nearly flat in kCostTri (0.3 .. 2.0 measured, 0.8 best by 1-3 %): node steps and block tests trade evenly.
#ifndef PT_MAX_LEAF_TRIS
#define PT_MAX_LEAF_TRIS 6
#endif
constexpr int kMaxLeafTris = PT_MAX_LEAF_TRIS;  // <= 6: a leaf slot addresses at most 3 blocks (3 unary bits)
static_assert(kMaxLeafTris >= 1 && kMaxLeafTris <= 6, "leaf slots hold 1..3 pair blocks");
constexpr float kCostNode = 1.0f;
#ifndef PT_COST_TRI
#define PT_COST_TRI 0.8f
#endif
constexpr float kCostTri = PT_COST_TRI;
PT_HD int leaf_blocks(int count) { return (count + 1) >> 1; }

struct CollapsePlan {
    float c[7];
    uint8_t dec[8];
    uint32_t pad;
};
static_assert(sizeof(CollapsePlan) == 40, "CollapsePlan is 40 bytes");

PT_HD float half_area(const float* lo, const float* hi) {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// plan of inner binary node `cur` from the finished plans of its children (leaves of the binary
// tree are single triangles: cost area * kCostTri for any number of slots)
PT_HD void plan_node(int cur, BinNode* nodes, CollapsePlan* plans, const uint32_t* sorted_tri, const float* tri_lo,
                     const float* tri_hi) {
    const BinNode& nd = nodes[cur];
    float cc[2][7];
    for (int side = 0; side < 2; ++side) {
        const int32_t c = side == 0 ? nd.left : nd.right;
        if (c < 0) {
            const uint32_t t = sorted_tri[~c];
            float lo[3], hi[3];
            for (int a = 0; a < 3; ++a) {
                lo[a] = load_cg(tri_lo + 3 * (size_t)t + a);
                hi[a] = load_cg(tri_hi + 3 * (size_t)t + a);
            }
            const float v = half_area(lo, hi) * kCostTri;
            for (int i = 0; i < 7; ++i) cc[side][i] = v;
        } else {
            for (int i = 0; i < 7; ++i) cc[side][i] = load_cg(&plans[c].c[i]);
        }
    }
    const float area = half_area(nd.lo, nd.hi);
    CollapsePlan p;
    p.pad = 0;
    float dist[9];  // dist[j]: best cost using exactly-at-most j slots split over both children
    uint8_t split[9];
    for (int j = 2; j <= 8; ++j) {
        float best = 3.4e38f;
        int bk = 1;
        for (int k = 1; k < j; ++k) {
            if (k > 7 || j - k > 7) continue;
            const float v = cc[0][k - 1] + cc[1][j - k - 1];
            if (v < best) {
                best = v;
                bk = k;
            }
        }
        dist[j] = best;
        split[j] = (uint8_t)bk;
    }
    const float c_leaf = nd.count <= kMaxLeafTris ? area * (float)leaf_blocks(nd.count) * kCostTri : 3.4e38f;
    const float c_inner = dist[8] + area * kCostNode;
    p.dec[7] = split[8];
    if (c_leaf <= c_inner) {
        p.c[0] = c_leaf;
        p.dec[0] = 1;
    } else {
        p.c[0] = c_inner;
        p.dec[0] = 0;
    }
    for (int j = 2; j <= 7; ++j) {
        if (dist[j] < p.c[j - 2]) {
            p.c[j - 1] = dist[j];
            p.dec[j - 1] = split[j];
        } else {
            p.c[j - 1] = p.c[j - 2];
            p.dec[j - 1] = 0;
        }
    }
    plans[cur] = p;
}

// leaf at sorted position k walks up; the second thread to reach a node merges the
// two child boxes (boxes of leaves come from tri_lo/tri_hi via the sorted order).
PT_HD void refit_from_leaf(int k, const int32_t* leaf_parent, BinNode* nodes, int32_t* flags,
                           const uint32_t* sorted_tri, const float* tri_lo, const float* tri_hi,
                           CollapsePlan* plans) {
    int cur = leaf_parent[k];
    while (cur >= 0) {
        fence();
        if (arrive(&flags[cur]) == 0) return;  // first arrival: sibling not ready yet
        fence();
        BinNode& nd = nodes[cur];
        float lo[3], hi[3];
        int32_t count = 0;
        for (int side = 0; side < 2; ++side) {
            const int32_t c = side == 0 ? nd.left : nd.right;
            const float *clo, *chi;
            if (c < 0) {
                const uint32_t t = sorted_tri[~c];
                clo = tri_lo + 3 * (size_t)t;
                chi = tri_hi + 3 * (size_t)t;
                count += 1;
            } else {
                clo = nodes[c].lo;
                chi = nodes[c].hi;
                count += load_cg_i(&nodes[c].count);
            }
            for (int a = 0; a < 3; ++a) {
                // child boxes were written by another thread: read them past L1
                const float l = load_cg(clo + a), h = load_cg(chi + a);
                lo[a] = side == 0 ? l : fminf(lo[a], l);
                hi[a] = side == 0 ? h : fmaxf(hi[a], h);
            }
        }
        for (int a = 0; a < 3; ++a) {
            nd.lo[a] = lo[a];
            nd.hi[a] = hi[a];
        }
        nd.count = count;  // (the radix tree already knows it; a tree changed by reinsertion does not)
        plan_node(cur, nodes, plans, sorted_tri, tri_lo, tri_hi);
        cur = nd.parent;
    }
}


// sorted positions of the triangles under binary node n (count <= kMaxLeafTris), left to right
PT_HD int leaf_positions(const BinNode* bin, int32_t n, int* out) {
    int32_t st[2 * kMaxLeafTris];
    int sp = 0, cnt = 0;
    st[sp++] = n;
    while (sp > 0) {
        const int32_t c = st[--sp];
        if (c < 0) {
            if (cnt < kMaxLeafTris) out[cnt] = ~c;
            ++cnt;
        } else {
            st[sp++] = bin[c].right;
            st[sp++] = bin[c].left;
        }
    }
    return cnt < kMaxLeafTris ? cnt : kMaxLeafTris;
}

// ---- binary tree by parallel locally-ordered clustering (Meister & Bittner, TVCG 2018) -------
// Clusters start as the Morton-sorted triangles.  Every round each cluster looks kPlocRadius
// positions to both sides for the neighbour giving the smallest merged box (ploc_nearest),
// mutual nearest neighbours merge into a new inner node (ploc_flags + an exclusive scan decide
// the node index and the compacted position, ploc_emit writes both), until one cluster is left.
// Ties go to the buddy position i^1, then to the smaller position, so the result is the same
// whatever the thread order and at least one pair merges every round.
// The new node's box, triangle count and collapse plan are complete when it is created: no
// separate refit pass.  Cluster ids: >= 0 inner node, < 0 ~(sorted position) of a triangle.
#ifndef PT_PLOC_RADIUS
#define PT_PLOC_RADIUS 8
#endif
constexpr int kPlocRadius = PT_PLOC_RADIUS;

struct Box6 {
    float lo[3], hi[3];
};

PT_HD float merged_half_area(const Box6& a, const Box6& b) {
    const float dx = fmaxf(a.hi[0], b.hi[0]) - fminf(a.lo[0], b.lo[0]);
    const float dy = fmaxf(a.hi[1], b.hi[1]) - fminf(a.lo[1], b.lo[1]);
    const float dz = fmaxf(a.hi[2], b.hi[2]) - fminf(a.lo[2], b.lo[2]);
    return dx * dy + dy * dz + dz * dx;
}

PT_HD void ploc_nearest(int i, int m, const Box6* cbox, int32_t* nn) {
    const Box6 me = cbox[i];
    const int j0 = i - kPlocRadius < 0 ? 0 : i - kPlocRadius;
    const int j1 = i + kPlocRadius > m - 1 ? m - 1 : i + kPlocRadius;
    float best = 3.4e38f;
    int bj = -1;
    for (int j = j0; j <= j1; ++j) {
        if (j == i) continue;
        const float d = merged_half_area(me, cbox[j]);
        // equal cost: the buddy position i^1 wins, then the smaller position (identical boxes
        // — duplicated triangles — then pair up (2k, 2k+1) and the tree stays balanced)
        if (d < best || (d == best && j == (i ^ 1))) {
            best = d;
            bj = j;
        }
    }
    nn[i] = bj;
}

// low 32 bits: 1 when position i survives the round, high 32 bits: 1 when it creates a node
PT_HD uint64_t ploc_flags(int i, const int32_t* nn) {
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == i;
    if (!mutual) return 1ull;
    return i < j ? (1ull | (1ull << 32)) : 0ull;
}

PT_HD void ploc_emit(int i, const int32_t* nn, uint64_t excl, const int32_t* cid, const Box6* cbox,
                     int32_t* out_cid, Box6* out_cbox, BinNode* bin, CollapsePlan* plans, uint32_t node_base,
                     const uint32_t* sorted_tri, const float* tri_lo, const float* tri_hi) {
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == i;
    if (mutual && i > j) return;
    const uint32_t pos = (uint32_t)(excl & 0xFFFFFFFFull);  // exclusive scan of ploc_flags at i
    if (!mutual) {
        out_cid[pos] = cid[i];
        out_cbox[pos] = cbox[i];
        return;
    }
    const int32_t id = (int32_t)(node_base + (uint32_t)(excl >> 32));
    const Box6 a = cbox[i], b = cbox[j];
    BinNode nd;
    for (int k = 0; k < 3; ++k) {
        nd.lo[k] = fminf(a.lo[k], b.lo[k]);
        nd.hi[k] = fmaxf(a.hi[k], b.hi[k]);
    }
    nd.left = cid[i];
    nd.right = cid[j];
    nd.first = 0;
    nd.count = (nd.left < 0 ? 1 : bin[nd.left].count) + (nd.right < 0 ? 1 : bin[nd.right].count);
    nd.parent = -1;
    nd.pad = 0;
    bin[id] = nd;
    plan_node(id, bin, plans, sorted_tri, tri_lo, tri_hi);
    out_cid[pos] = id;
    Box6 u;
    for (int k = 0; k < 3; ++k) {
        u.lo[k] = nd.lo[k];
        u.hi[k] = nd.hi[k];
    }
    out_cbox[pos] = u;
}

// ---- reinsertion optimisation of the binary tree ---------------------------------------------
// Between the clustering and the wide collapse: every node (inner node or triangle) looks for the
// place in the tree where its subtree would cost the least surface area, and a conflict-free
// subset of those moves is applied; a few passes.  The search is the branch-and-bound walk of
// Meister & Bittner 2018 ("Parallel reinsertion for bounding volume hierarchy optimization"):
// from the node's parent upwards, into the sibling subtree at every level.  Conflicts are settled
// without any order dependence, so the device and its host emulation produce the same tree:
//   * a move touches six nodes (x, its parent p, sibling s, grandparent g, the target o and o's
//     parent); it takes part in a round by atomicMax of (gain, candidate) on their lock words and
//     wins when all six hold its key; winners keep their nodes (~0) and the rest try again,
//   * boxes and triangle counts along the two paths are not locked: one refit of the whole tree
//     after the pass restores them, so a gain is an estimate once several moves share ancestors,
//   * a winner whose insertion path (o up to the common ancestor) runs through the root of
//     another winner's moved subtree is dropped: two such moves could close a cycle.
#ifndef PT_REINSERT_PASSES
#define PT_REINSERT_PASSES 4
#endif
constexpr int kReinsertPasses = PT_REINSERT_PASSES;
constexpr int kReinsertRounds = 3;
constexpr int kReinsertStack = 64;
constexpr unsigned long long kRiHeld = ~0ull;
enum : uint8_t { kRiNone = 0, kRiActive = 1, kRiWin = 2, kRiDead = 3 };

struct ReinsertTree {
    BinNode* bin;          // inner nodes 0..n-2
    int32_t* leaf_parent;  // per sorted position
    const uint32_t* sorted_tri;
    const float *tri_lo, *tri_hi;
    int32_t n;     // triangles
    int32_t root;  // never moves
};
// candidates and lock slots: 0..n-2 the inner nodes, n-1..2n-2 the leaves by sorted position
PT_HD int32_t ri_ref(const ReinsertTree& t, int c) { return c < t.n - 1 ? c : ~(c - (t.n - 1)); }
PT_HD int ri_slot(const ReinsertTree& t, int32_t ref) { return ref >= 0 ? ref : (t.n - 1) + ~ref; }
PT_HD int32_t ri_parent(const ReinsertTree& t, int32_t ref) { return ref >= 0 ? t.bin[ref].parent : t.leaf_parent[~ref]; }
PT_HD void ri_set_parent(const ReinsertTree& t, int32_t ref, int32_t p) {
    if (ref >= 0)
        t.bin[ref].parent = p;
    else
        t.leaf_parent[~ref] = p;
}
PT_HD void ri_box(const ReinsertTree& t, int32_t ref, Box6& b) {
    if (ref >= 0) {
        const BinNode& nd = t.bin[ref];
        for (int a = 0; a < 3; ++a) {
            b.lo[a] = nd.lo[a];
            b.hi[a] = nd.hi[a];
        }
    } else {
        const uint32_t tri = t.sorted_tri[~ref];
        for (int a = 0; a < 3; ++a) {
            b.lo[a] = t.tri_lo[3 * (size_t)tri + a];
            b.hi[a] = t.tri_hi[3 * (size_t)tri + a];
        }
    }
}
// the clustering leaves parent links unset: inner node i names itself in its two children
PT_HD void ri_link_children(int i, const ReinsertTree& t) {
    ri_set_parent(t, t.bin[i].left, i);
    ri_set_parent(t, t.bin[i].right, i);
    if (i == t.root) t.bin[i].parent = -1;
}

// best new sibling `target` of candidate c, the common ancestor `pivot` of old and new place, and the
// surface area the move saves (0: stay).  Removing x deletes its parent p (the sibling s moves up) and
// shrinks the ancestors of p; inserting x next to o re-creates p above o and grows o's ancestors below
// the pivot.  `save` collects the first, `ind` (per search branch) the second.
PT_HD void ri_find(int c, const ReinsertTree& t, float* gain, int32_t* target, int32_t* pivot, uint8_t* state) {
    gain[c] = 0.0f;
    state[c] = kRiNone;
    const int32_t x = ri_ref(t, c);
    if (x == t.root) return;
    const int32_t p = ri_parent(t, x);
    if (p < 0 || p == t.root) return;
    Box6 bx;
    ri_box(t, x, bx);
    const float ax = half_area(bx.lo, bx.hi);
    const int32_t s = t.bin[p].left == x ? t.bin[p].right : t.bin[p].left;
    float save = half_area(t.bin[p].lo, t.bin[p].hi);
    Box6 shrunk;  // box of the path node below the pivot once x is gone
    ri_box(t, s, shrunk);
    float best = 0.0f;
    int32_t best_o = 0, best_pivot = p;
    int32_t piv = p, start = s, cur = p;
    int32_t st_o[kReinsertStack];
    float st_ind[kReinsertStack];
    for (;;) {
        // places in the subtree of `start`, reached from the common ancestor piv
        int sp = 0;
        st_o[sp] = start;
        st_ind[sp++] = 0.0f;
        while (sp > 0) {
            --sp;
            const int32_t o = st_o[sp];
            const float ind = st_ind[sp];
            Box6 b;
            ri_box(t, o, b);
            const float m = merged_half_area(b, bx);
            const float g = save - ind - m;
            if (g > best) {
                best = g;
                best_o = o;
                best_pivot = piv;
            }
            if (o >= 0) {
                const float ind2 = ind + (m - half_area(b.lo, b.hi));
                // nothing below o saves more than save - ind2 - area(x)
                if (save - ind2 - ax > best && sp + 2 <= kReinsertStack) {
                    st_o[sp] = t.bin[o].right;
                    st_ind[sp++] = ind2;
                    st_o[sp] = t.bin[o].left;
                    st_ind[sp++] = ind2;
                }
            }
        }
        const int32_t q = t.bin[cur].parent;
        if (q < 0) break;
        if (cur != p) {  // places above cur also save what cur shrinks by
            const float shrink = half_area(t.bin[cur].lo, t.bin[cur].hi) - half_area(shrunk.lo, shrunk.hi);
            if (!(shrink > 0.0f) && !(save - ax > best)) break;
            save += shrink;
        }
        start = t.bin[q].left == cur ? t.bin[q].right : t.bin[q].left;
        Box6 bo;
        ri_box(t, start, bo);
        for (int a = 0; a < 3; ++a) {  // q without x
            shrunk.lo[a] = fminf(shrunk.lo[a], bo.lo[a]);
            shrunk.hi[a] = fmaxf(shrunk.hi[a], bo.hi[a]);
        }
        piv = q;
        cur = q;
    }
    gain[c] = best;
    target[c] = best_o;
    pivot[c] = best_pivot;
    if (best > 0.0f) state[c] = kRiActive;
}

PT_HD unsigned long long ri_key(float gain, int c) {
    uint32_t g;
    memcpy(&g, &gain, 4);
    return ((unsigned long long)g << 32) | (uint32_t)c;
}
PT_HD void ri_lock_max(unsigned long long* w, unsigned long long key) {
#if defined(__CUDA_ARCH__)
    atomicMax(w, key);
#else
    if (*w < key) *w = key;
#endif
}
// the six lock slots of candidate c's move
PT_HD void ri_slots(int c, const ReinsertTree& t, const int32_t* target, int out[6]) {
    const int32_t x = ri_ref(t, c), p = ri_parent(t, x);
    const int32_t s = t.bin[p].left == x ? t.bin[p].right : t.bin[p].left;
    const int32_t o = target[c];
    out[0] = c;
    out[1] = ri_slot(t, p);
    out[2] = ri_slot(t, s);
    out[3] = ri_slot(t, t.bin[p].parent);
    out[4] = ri_slot(t, o);
    out[5] = ri_slot(t, ri_parent(t, o));
}
// one round: (1) ri_bid, (2) ri_settle, then lock words are cleared and (3) ri_hold re-marks what winners own
PT_HD void ri_bid(int c, const ReinsertTree& t, const float* gain, const int32_t* target, uint8_t* state,
                  unsigned long long* lock) {
    if (state[c] != kRiActive) return;
    int sl[6];
    ri_slots(c, t, target, sl);
    for (int k = 0; k < 6; ++k)
        if (lock[sl[k]] == kRiHeld) {  // a winner of an earlier round owns one of the nodes
            state[c] = kRiDead;
            return;
        }
    const unsigned long long key = ri_key(gain[c], c);
    for (int k = 0; k < 6; ++k) ri_lock_max(&lock[sl[k]], key);
}
PT_HD void ri_settle(int c, const ReinsertTree& t, const float* gain, const int32_t* target, uint8_t* state,
                     const unsigned long long* lock) {
    if (state[c] != kRiActive) return;
    int sl[6];
    ri_slots(c, t, target, sl);
    const unsigned long long key = ri_key(gain[c], c);
    bool all = true;
    for (int k = 0; k < 6; ++k) all = all && lock[sl[k]] == key;
    if (all) state[c] = kRiWin;
}
PT_HD void ri_hold(int c, const ReinsertTree& t, const int32_t* target, const uint8_t* state, unsigned long long* lock) {
    if (state[c] != kRiWin) return;
    int sl[6];
    ri_slots(c, t, target, sl);
    for (int k = 0; k < 6; ++k) lock[sl[k]] = kRiHeld;
}
// after the rounds (lock words cleared again): winners mark the root of the subtree they move ...
PT_HD void ri_mark_moved(int c, const uint8_t* state, unsigned long long* lock) {
    if (state[c] == kRiWin) lock[c] = 1ull;
}
// ... and give up when their insertion path crosses somebody else's
PT_HD void ri_drop_crossing(int c, const ReinsertTree& t, const int32_t* target, const int32_t* pivot, uint8_t* state,
                            const unsigned long long* lock) {
    if (state[c] != kRiWin) return;
    for (int32_t a = ri_parent(t, target[c]); a != pivot[c]; a = t.bin[a].parent)
        if (lock[a] != 0ull) {
            state[c] = kRiDead;
            return;
        }
}
PT_HD void ri_apply(int c, const ReinsertTree& t, const int32_t* target, const uint8_t* state) {
    if (state[c] != kRiWin) return;
    const int32_t x = ri_ref(t, c), o = target[c];
    const int32_t p = ri_parent(t, x);
    const int32_t s = t.bin[p].left == x ? t.bin[p].right : t.bin[p].left;
    const int32_t g = t.bin[p].parent;
    if (t.bin[g].left == p)  // s takes p's place
        t.bin[g].left = s;
    else
        t.bin[g].right = s;
    ri_set_parent(t, s, g);
    const int32_t po = ri_parent(t, o);  // p is re-used above o
    if (t.bin[po].left == o)
        t.bin[po].left = p;
    else
        t.bin[po].right = p;
    t.bin[p].parent = po;
    t.bin[p].left = o;
    t.bin[p].right = x;
    ri_set_parent(t, o, p);
}

struct CollapseItem {
    int32_t bin;    // binary node to turn into a wide node (>= 0), or ~leafpos for a 1-triangle scene
    uint32_t wide;  // index of the wide node to write
};

// smallest exponent byte eb such that 255 * 2^(eb-127) >= ext (with room for the
// outward rounding of quantise())
PT_HD uint32_t quant_exponent(float ext) {
    // scale = ext / 255 rounded up to a power of two, then one notch of slack
    float s = ext * (1.0f / 254.0f);
    uint32_t bits = f2u(s);
    uint32_t eb = (bits >> 23) & 0xFFu;
    if (bits & 0x7FFFFFu) eb += 1;  // round up to the next power of two
    if (eb < 1) eb = 1;
    if (eb > 230) eb = 230;  // the walk adds 15 to the exponent byte (slab_axis)
    return eb;
}

// Build wide node `item.wide` from binary node `item.bin`.
// Outputs: nodes_out[item.wide], triangles into tris_out (node order), and new items
// for the inner children into next[]/next_count.
PT_HD void collapse_node(CollapseItem item, const BinNode* bin, const uint32_t* sorted_tri,
                         const float* tri_lo, const float* tri_hi, const f4* pos, const int32_t* idx4,
                         Node8* nodes_out, uint32_t* node_count, TriPair* tris_out, uint32_t* tri_count,
                         CollapseItem* next, uint32_t* next_count, int n_tris, const CollapsePlan* plans) {
    // children under construction: inner binary node (>=0) or single leaf (~pos)
    int32_t ch[8];
    int nch = 0;
    float nlo[3], nhi[3];
    if (item.bin < 0) {  // scene with a single triangle
        ch[nch++] = item.bin;
        const uint32_t t = sorted_tri[~item.bin];
        for (int a = 0; a < 3; ++a) {
            nlo[a] = tri_lo[3 * (size_t)t + a];
            nhi[a] = tri_hi[3 * (size_t)t + a];
        }
    } else {
        const BinNode& b = bin[item.bin];
        for (int a = 0; a < 3; ++a) {
            nlo[a] = b.lo[a];
            nhi[a] = b.hi[a];
        }
        if (b.count <= kMaxLeafTris) {
            ch[nch++] = item.bin;  // whole subtree is one leaf
        } else {
            // follow the cost-optimal plan: 8 slots distributed over the two subtrees
            int32_t st_n[16];
            int st_j[16];
            int sp = 0;
            const int k0 = plans[item.bin].dec[7];
            st_n[sp] = b.right;
            st_j[sp++] = 8 - k0;
            st_n[sp] = b.left;
            st_j[sp++] = k0;
            while (sp > 0) {
                --sp;
                const int32_t n = st_n[sp];
                int j = st_j[sp];
                if (n < 0) {
                    ch[nch++] = n;
                    continue;
                }
                const CollapsePlan& p = plans[n];
                while (j > 1 && p.dec[j - 1] == 0) --j;
                if (j == 1) {
                    ch[nch++] = n;
                    continue;
                }
                const int k = p.dec[j - 1];
                st_n[sp] = bin[n].right;
                st_j[sp++] = j - k;
                st_n[sp] = bin[n].left;
                st_j[sp++] = k;
            }
        }
    }

    // child boxes, kinds
    float clo[8][3], chi[8][3];
    bool inner[8];
    int cnt[8];
    int lpos[8][kMaxLeafTris];  // sorted positions of a leaf child's triangles
    for (int i = 0; i < nch; ++i) {
        if (ch[i] < 0) {
            const int posn = ~ch[i];
            const uint32_t t = sorted_tri[posn];
            for (int a = 0; a < 3; ++a) {
                clo[i][a] = tri_lo[3 * (size_t)t + a];
                chi[i][a] = tri_hi[3 * (size_t)t + a];
            }
            inner[i] = false;
            lpos[i][0] = posn;
            cnt[i] = 1;
        } else {
            const BinNode& c = bin[ch[i]];
            for (int a = 0; a < 3; ++a) {
                clo[i][a] = c.lo[a];
                chi[i][a] = c.hi[a];
            }
            inner[i] = c.count > kMaxLeafTris && plans[ch[i]].dec[0] == 0;
            cnt[i] = c.count;
            if (!inner[i]) cnt[i] = leaf_positions(bin, ch[i], lpos[i]);
        }
    }

    // slot assignment: slot s (bits x=4,y=2,z=1) should hold the child lying furthest
    // in direction (+-1,+-1,+-1)_s from the node centre; greedy on the 8x8 score table
    int slot_of[8];
    bool slot_used[8] = {false, false, false, false, false, false, false, false};
    bool child_done[8] = {false, false, false, false, false, false, false, false};
    const float ncx = 0.5f * (nlo[0] + nhi[0]), ncy = 0.5f * (nlo[1] + nhi[1]),
                ncz = 0.5f * (nlo[2] + nhi[2]);
    for (int round = 0; round < nch; ++round) {
        float best = -3.4e38f;
        int bi = -1, bs = -1;
        for (int i = 0; i < nch; ++i) {
            if (child_done[i]) continue;
            const float cx = 0.5f * (clo[i][0] + chi[i][0]) - ncx, cy = 0.5f * (clo[i][1] + chi[i][1]) - ncy,
                        cz = 0.5f * (clo[i][2] + chi[i][2]) - ncz;
            for (int s = 0; s < 8; ++s) {
                if (slot_used[s]) continue;
                const float sc = ((s & 4) ? cx : -cx) + ((s & 2) ? cy : -cy) + ((s & 1) ? cz : -cz);
                if (sc > best) {
                    best = sc;
                    bi = i;
                    bs = s;
                }
            }
        }
        slot_of[bi] = bs;
        slot_used[bs] = true;
        child_done[bi] = true;
    }

    // counts, allocation
    int n_inner = 0, n_leaf_blocks = 0;
    for (int i = 0; i < nch; ++i) {
        if (inner[i]) ++n_inner;
        else n_leaf_blocks += leaf_blocks(cnt[i]);
    }
    const uint32_t child_base = n_inner ? grab(node_count, (uint32_t)n_inner) : 0u;
    const uint32_t tri_base = n_leaf_blocks ? grab(tri_count, (uint32_t)n_leaf_blocks) : 0u;  // tri_count counts BLOCKS
    const uint32_t next_base = n_inner ? grab(next_count, (uint32_t)n_inner) : 0u;

    // Quantisation grid: the scale is chosen for 1.01 x the extent, and the grid origin sits 1/8 of a
    // step below the node box.  That leaves room on both sides, so the 1/64-step outward slack of
    // every child plane is never lost to the clamp at 0 / 255 — the slab test of the walker
    // (plane_q, pt_bvh8.h) has a rounding error of up to 2^-9 step and relies on that slack.
    const uint32_t ex = quant_exponent((nhi[0] - nlo[0]) * 1.01f), ey = quant_exponent((nhi[1] - nlo[1]) * 1.01f),
                   ez = quant_exponent((nhi[2] - nlo[2]) * 1.01f);
    const float isx = u2f((254u - ex) << 23), isy = u2f((254u - ey) << 23), isz = u2f((254u - ez) << 23);
    nlo[0] = nlo[0] - 0.125f * u2f(ex << 23);
    nlo[1] = nlo[1] - 0.125f * u2f(ey << 23);
    nlo[2] = nlo[2] - 0.125f * u2f(ez << 23);

    uint8_t meta[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint8_t q[6][8];
    for (int a = 0; a < 6; ++a)
        for (int s = 0; s < 8; ++s) q[a][s] = 0;
    uint32_t imask = 0;

    // walk slots in ascending order so inner children get consecutive node indices
    int inner_rank = 0, tri_off = 0;
    for (int s = 0; s < 8; ++s) {
        int i = -1;
        for (int k = 0; k < nch; ++k)
            if (slot_of[k] == s) i = k;
        if (i < 0) continue;
        // quantise outward with 1/64 of a step of extra slack (pt_bvh8.h header)
        const float inv[3] = {isx, isy, isz};
        for (int a = 0; a < 3; ++a) {
            float ql = floorf((clo[i][a] - nlo[a]) * inv[a] - 0.015625f);
            float qh = ceilf((chi[i][a] - nlo[a]) * inv[a] + 0.015625f);
            ql = fminf(fmaxf(ql, 0.0f), 255.0f);
            qh = fminf(fmaxf(qh, 0.0f), 255.0f);
            q[a][s] = (uint8_t)ql;
            q[3 + a][s] = (uint8_t)qh;
        }
        if (inner[i]) {
            meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
            imask |= 1u << s;
            next[next_base + inner_rank] = CollapseItem{ch[i], child_base + (uint32_t)inner_rank};
            ++inner_rank;
        } else {
            const int nb = leaf_blocks(cnt[i]);
            const uint32_t unary = nb == 1 ? 1u : (nb == 2 ? 3u : 7u);
            meta[s] = (uint8_t)((unary << 5) | (uint32_t)tri_off);
            for (int blk = 0; blk < nb; ++blk) {
                float v[2][9];
                uint32_t id[2];
                for (int half = 0; half < 2; ++half) {
                    const int k = 2 * blk + half;
                    if (k < cnt[i]) {
                        const uint32_t t = sorted_tri[lpos[i][k]];
                        const f4 a = pos[idx4[4 * (size_t)t + 0]], b = pos[idx4[4 * (size_t)t + 1]],
                                 c = pos[idx4[4 * (size_t)t + 2]];
                        v[half][0] = a.x; v[half][1] = a.y; v[half][2] = a.z;
                        v[half][3] = b.x - a.x; v[half][4] = b.y - a.y; v[half][5] = b.z - a.z;
                        v[half][6] = c.x - a.x; v[half][7] = c.y - a.y; v[half][8] = c.z - a.z;
                        id[half] = t;
                    } else {  // odd leaf: empty half, never accepted (det = 0)
                        for (int q2 = 0; q2 < 9; ++q2) v[half][q2] = 0.0f;
                        id[half] = 0xFFFFFFFFu;
                    }
                }
                TriPair o;
                o.w[0] = f4{v[0][0], v[1][0], v[0][1], v[1][1]};
                o.w[1] = f4{v[0][2], v[1][2], v[0][3], v[1][3]};
                o.w[2] = f4{v[0][4], v[1][4], v[0][5], v[1][5]};
                o.w[3] = f4{v[0][6], v[1][6], v[0][7], v[1][7]};
                o.w[4] = f4{v[0][8], v[1][8], u2f(id[0]), u2f(id[1])};
                tris_out[tri_base + (uint32_t)tri_off + (uint32_t)blk] = o;
            }
            tri_off += nb;
        }
    }

    auto pack4 = [](const uint8_t* p) {
        return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
    };
    Node8 n;
    n.w[0] = u4{f2u(nlo[0]), f2u(nlo[1]), f2u(nlo[2]), ex | (ey << 8) | (ez << 16) | (imask << 24)};
    n.w[1] = u4{child_base, tri_base, pack4(meta), pack4(meta + 4)};
    n.w[2] = u4{pack4(q[0]), pack4(q[0] + 4), pack4(q[1]), pack4(q[1] + 4)};
    n.w[3] = u4{pack4(q[2]), pack4(q[2] + 4), pack4(q[3]), pack4(q[3] + 4)};
    n.w[4] = u4{pack4(q[4]), pack4(q[4] + 4), pack4(q[5]), pack4(q[5] + 4)};
    nodes_out[item.wide] = n;
    (void)n_tris;
}

// ---- refit of the finished wide tree (b200pt_refit_accel): vertices moved, topology kept.
// Wide node `ni`: its leaf slots get their triangle blocks rewritten from the current vertex positions and their
// boxes from the triangles' padded boxes (tri_box, the builder's own), its inner slots take the box the child
// wrote into node_box one level below; the node then gets a new quantisation grid and child planes by the very
// rules of collapse_node (1.01 x extent, origin 1/8 step below, 1/64 step outward slack).  w1 (child base,
// triangle base, meta bytes) and the slot assignment are untouched.  node_box: 6 floats per node (lo, hi).
// Callers run the levels of the tree bottom-up (node indices of a level are contiguous: DeviceBVH::level_end).
PT_HD void refit_wide_node(uint32_t ni, Node8* nodes, TriPair* tris, const f4* pos, const int32_t* idx4, float pad,
                           float* node_box) {
    Node8 n = nodes[ni];
    const uint32_t imask = n.w[0].w >> 24;
    const uint32_t child_base = n.w[1].x, tri_base = n.w[1].y;
    float clo[8][3], chi[8][3];
    bool present[8];
    float nlo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, nhi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int s = 0; s < 8; ++s) {
        const uint32_t meta = ((s < 4 ? n.w[1].z : n.w[1].w) >> (8 * (s & 3))) & 0xFFu;
        present[s] = meta != 0;
        if (!present[s]) continue;
        for (int a = 0; a < 3; ++a) {
            clo[s][a] = 3.0e38f;
            chi[s][a] = -3.0e38f;
        }
        if ((meta & 0x18u) == 0x18u) {  // inner child: (1 << 5) | (24 + s)
            const uint32_t child = child_base + (uint32_t)popc32(imask & ((1u << s) - 1u));
            for (int a = 0; a < 3; ++a) {
                clo[s][a] = node_box[6 * (size_t)child + a];
                chi[s][a] = node_box[6 * (size_t)child + 3 + a];
            }
        } else {
            const int nb = popc32(meta >> 5);
            const uint32_t first = tri_base + (meta & 0x1Fu);
            for (int blk = 0; blk < nb; ++blk) {
                TriPair o = tris[first + (uint32_t)blk];
                const uint32_t id[2] = {f2u(o.w[4].z), f2u(o.w[4].w)};
                float v[2][9];
                for (int half = 0; half < 2; ++half) {
                    if (id[half] != 0xFFFFFFFFu) {
                        const uint32_t t = id[half];
                        const f4 a = pos[idx4[4 * (size_t)t + 0]], b = pos[idx4[4 * (size_t)t + 1]],
                                 c = pos[idx4[4 * (size_t)t + 2]];
                        v[half][0] = a.x; v[half][1] = a.y; v[half][2] = a.z;
                        v[half][3] = b.x - a.x; v[half][4] = b.y - a.y; v[half][5] = b.z - a.z;
                        v[half][6] = c.x - a.x; v[half][7] = c.y - a.y; v[half][8] = c.z - a.z;
                        float lo[3], hi[3];
                        tri_box(pos, idx4, t, pad, lo, hi);
                        for (int k = 0; k < 3; ++k) {
                            clo[s][k] = fminf(clo[s][k], lo[k]);
                            chi[s][k] = fmaxf(chi[s][k], hi[k]);
                        }
                    } else {
                        for (int q2 = 0; q2 < 9; ++q2) v[half][q2] = 0.0f;
                    }
                }
                o.w[0] = f4{v[0][0], v[1][0], v[0][1], v[1][1]};
                o.w[1] = f4{v[0][2], v[1][2], v[0][3], v[1][3]};
                o.w[2] = f4{v[0][4], v[1][4], v[0][5], v[1][5]};
                o.w[3] = f4{v[0][6], v[1][6], v[0][7], v[1][7]};
                o.w[4] = f4{v[0][8], v[1][8], u2f(id[0]), u2f(id[1])};
                tris[first + (uint32_t)blk] = o;
            }
        }
        for (int a = 0; a < 3; ++a) {
            nlo[a] = fminf(nlo[a], clo[s][a]);
            nhi[a] = fmaxf(nhi[a], chi[s][a]);
        }
    }
    for (int a = 0; a < 3; ++a) {
        node_box[6 * (size_t)ni + a] = nlo[a];
        node_box[6 * (size_t)ni + 3 + a] = nhi[a];
    }
    // same grid and outward quantisation as collapse_node
    const uint32_t ex = quant_exponent((nhi[0] - nlo[0]) * 1.01f), ey = quant_exponent((nhi[1] - nlo[1]) * 1.01f),
                   ez = quant_exponent((nhi[2] - nlo[2]) * 1.01f);
    const float inv[3] = {u2f((254u - ex) << 23), u2f((254u - ey) << 23), u2f((254u - ez) << 23)};
    nlo[0] = nlo[0] - 0.125f * u2f(ex << 23);
    nlo[1] = nlo[1] - 0.125f * u2f(ey << 23);
    nlo[2] = nlo[2] - 0.125f * u2f(ez << 23);
    uint8_t q[6][8];
    for (int a = 0; a < 6; ++a)
        for (int s = 0; s < 8; ++s) q[a][s] = 0;
    for (int s = 0; s < 8; ++s) {
        if (!present[s]) continue;
        for (int a = 0; a < 3; ++a) {
            float ql = floorf((clo[s][a] - nlo[a]) * inv[a] - 0.015625f);
            float qh = ceilf((chi[s][a] - nlo[a]) * inv[a] + 0.015625f);
            ql = fminf(fmaxf(ql, 0.0f), 255.0f);
            qh = fminf(fmaxf(qh, 0.0f), 255.0f);
            q[a][s] = (uint8_t)ql;
            q[3 + a][s] = (uint8_t)qh;
        }
    }
    auto pack4 = [](const uint8_t* p) {
        return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
    };
    n.w[0] = u4{f2u(nlo[0]), f2u(nlo[1]), f2u(nlo[2]), ex | (ey << 8) | (ez << 16) | (imask << 24)};
    n.w[2] = u4{pack4(q[0]), pack4(q[0] + 4), pack4(q[1]), pack4(q[1] + 4)};
    n.w[3] = u4{pack4(q[2]), pack4(q[2] + 4), pack4(q[3]), pack4(q[3] + 4)};
    n.w[4] = u4{pack4(q[4]), pack4(q[4] + 4), pack4(q[5]), pack4(q[5] + 4)};
    nodes[ni] = n;
}

}  // namespace pt

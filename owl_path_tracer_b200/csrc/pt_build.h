// pt_build.h — per-element steps of the on-device BVH builder.
//
// Replaces optixAccelBuild as driven by the reference through OWL
// (application.cpp:131-140, extern/owl/owl/TrianglesGeomGroup.cpp:121-330).  Pipeline:
//   1. triangle boxes + scene box          (tri_prepare, one thread per triangle)
//   2. 63-bit Morton key of the centroid   (same pass)
//   3. radix sort of (key, triangle)       (cub::DeviceRadixSort in bvh_build.cu)
//   4. binary radix tree, Karras 2012      (radix_tree_node, one thread per inner node)
//   5. bottom-up boxes with arrival flags  (refit_from_leaf, one thread per leaf)
//   6. top-down collapse to 8-wide nodes, greedy by surface area, subtrees of <= 3
//      triangles become leaves; octant-ordered slot assignment; 8-bit quantisation;
//      triangles re-emitted in node order   (collapse_node, level by level)
// Each step is a PT_HD function of the element index so that tests/emu can run the
// very same code sequentially on the host.
#pragma once
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
PT_HD void fence() {
#if defined(__CUDA_ARCH__)
    __threadfence();
#endif
}

// leaf at sorted position k walks up; the second thread to reach a node merges the
// two child boxes (boxes of leaves come from tri_lo/tri_hi via the sorted order).
PT_HD void refit_from_leaf(int k, const int32_t* leaf_parent, BinNode* nodes, int32_t* flags,
                           const uint32_t* sorted_tri, const float* tri_lo, const float* tri_hi) {
    int cur = leaf_parent[k];
    while (cur >= 0) {
        fence();
        if (arrive(&flags[cur]) == 0) return;  // first arrival: sibling not ready yet
        fence();
        BinNode& nd = nodes[cur];
        float lo[3], hi[3];
        for (int side = 0; side < 2; ++side) {
            const int32_t c = side == 0 ? nd.left : nd.right;
            const float *clo, *chi;
            if (c < 0) {
                const uint32_t t = sorted_tri[~c];
                clo = tri_lo + 3 * (size_t)t;
                chi = tri_hi + 3 * (size_t)t;
            } else {
                clo = nodes[c].lo;
                chi = nodes[c].hi;
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
        cur = nd.parent;
    }
}

PT_HD float half_area(const float* lo, const float* hi) {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

struct CollapseItem {
    int32_t bin;    // binary node to turn into a wide node (>= 0), or ~leafpos for a 1-triangle scene
    uint32_t wide;  // index of the wide node to write
};

constexpr int kMaxLeafTris = 3;

// smallest exponent byte eb such that 255 * 2^(eb-127) >= ext (with room for the
// outward rounding of quantise())
PT_HD uint32_t quant_exponent(float ext) {
    // scale = ext / 255 rounded up to a power of two, then one notch of slack
    float s = ext * (1.0f / 254.0f);
    uint32_t bits = f2u(s);
    uint32_t eb = (bits >> 23) & 0xFFu;
    if (bits & 0x7FFFFFu) eb += 1;  // round up to the next power of two
    if (eb < 1) eb = 1;
    if (eb > 250) eb = 250;
    return eb;
}

// Build wide node `item.wide` from binary node `item.bin`.
// Outputs: nodes_out[item.wide], triangles into tris_out (node order), and new items
// for the inner children into next[]/next_count.
PT_HD void collapse_node(CollapseItem item, const BinNode* bin, const uint32_t* sorted_tri,
                         const float* tri_lo, const float* tri_hi, const f4* pos, const int32_t* idx4,
                         Node8* nodes_out, uint32_t* node_count, Tri48* tris_out, uint32_t* tri_count,
                         CollapseItem* next, uint32_t* next_count, int n_tris) {
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
            ch[nch++] = b.left;
            ch[nch++] = b.right;
            // greedy: open the inner child with the largest box until 8 children
            while (nch < 8) {
                int best = -1;
                float best_area = -1.0f;
                for (int i = 0; i < nch; ++i) {
                    if (ch[i] < 0) continue;
                    const BinNode& c = bin[ch[i]];
                    if (c.count <= kMaxLeafTris) continue;
                    const float ar = half_area(c.lo, c.hi);
                    if (ar > best_area) {
                        best_area = ar;
                        best = i;
                    }
                }
                if (best < 0) break;
                const BinNode& c = bin[ch[best]];
                ch[best] = c.left;
                ch[nch++] = c.right;
            }
        }
    }

    // child boxes, kinds
    float clo[8][3], chi[8][3];
    bool inner[8];
    int first[8], cnt[8];
    for (int i = 0; i < nch; ++i) {
        if (ch[i] < 0) {
            const int posn = ~ch[i];
            const uint32_t t = sorted_tri[posn];
            for (int a = 0; a < 3; ++a) {
                clo[i][a] = tri_lo[3 * (size_t)t + a];
                chi[i][a] = tri_hi[3 * (size_t)t + a];
            }
            inner[i] = false;
            first[i] = posn;
            cnt[i] = 1;
        } else {
            const BinNode& c = bin[ch[i]];
            for (int a = 0; a < 3; ++a) {
                clo[i][a] = c.lo[a];
                chi[i][a] = c.hi[a];
            }
            inner[i] = c.count > kMaxLeafTris;
            first[i] = c.first;
            cnt[i] = c.count;
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
    int n_inner = 0, n_leaf_tris = 0;
    for (int i = 0; i < nch; ++i) {
        if (inner[i]) ++n_inner;
        else n_leaf_tris += cnt[i];
    }
    const uint32_t child_base = n_inner ? grab(node_count, (uint32_t)n_inner) : 0u;
    const uint32_t tri_base = n_leaf_tris ? grab(tri_count, (uint32_t)n_leaf_tris) : 0u;
    const uint32_t next_base = n_inner ? grab(next_count, (uint32_t)n_inner) : 0u;

    const uint32_t ex = quant_exponent(nhi[0] - nlo[0]), ey = quant_exponent(nhi[1] - nlo[1]),
                   ez = quant_exponent(nhi[2] - nlo[2]);
    const float isx = u2f((254u - ex) << 23), isy = u2f((254u - ey) << 23), isz = u2f((254u - ez) << 23);

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
            const uint32_t unary = cnt[i] == 1 ? 1u : (cnt[i] == 2 ? 3u : 7u);
            meta[s] = (uint8_t)((unary << 5) | (uint32_t)tri_off);
            for (int k = 0; k < cnt[i]; ++k) {
                const uint32_t t = sorted_tri[first[i] + k];
                const f4 a = pos[idx4[4 * (size_t)t + 0]], b = pos[idx4[4 * (size_t)t + 1]],
                         c = pos[idx4[4 * (size_t)t + 2]];
                Tri48 o;
                o.a = f4{a.x, a.y, a.z, u2f(t)};
                o.b = f4{b.x - a.x, b.y - a.y, b.z - a.z, 0.0f};
                o.c = f4{c.x - a.x, c.y - a.y, c.z - a.z, 0.0f};
                tris_out[tri_base + (uint32_t)tri_off + (uint32_t)k] = o;
            }
            tri_off += cnt[i];
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

}  // namespace pt

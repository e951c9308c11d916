// Experiment only: a sequential top-down binned-SAH binary tree as the start of the reinsertion
// passes instead of the clustering, to learn whether a different start changes node visits per ray.  Hooked into
// tests/emu/emu.cpp right after "root = cid[0];" as
//     if (getenv("EMU_SAH")) root = sah_exp::build(bin.data(), (int)n, order.data(), tri_lo.data(), tri_hi.data());
// Not part of any build.
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

namespace sah_exp {
using namespace pt;

struct Ctx {
    BinNode* bin;
    const uint32_t* order;
    const float *tri_lo, *tri_hi;
    std::vector<int32_t> pos;  // sorted positions, permuted in place
    int next = 0;
};

inline void tbox(const Ctx& c, int32_t p, float* lo, float* hi) {
    const uint32_t t = c.order[p];
    for (int a = 0; a < 3; ++a) {
        lo[a] = c.tri_lo[3 * (size_t)t + a];
        hi[a] = c.tri_hi[3 * (size_t)t + a];
    }
}

inline int32_t rec(Ctx& c, int b, int e) {
    if (e - b == 1) return ~c.pos[b];
    const int id = c.next++;
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f}, clo[3] = {3e38f, 3e38f, 3e38f},
          chi[3] = {-3e38f, -3e38f, -3e38f};
    for (int i = b; i < e; ++i) {
        float l[3], h[3];
        tbox(c, c.pos[i], l, h);
        for (int a = 0; a < 3; ++a) {
            lo[a] = fminf(lo[a], l[a]);
            hi[a] = fmaxf(hi[a], h[a]);
            const float m = 0.5f * (l[a] + h[a]);
            clo[a] = fminf(clo[a], m);
            chi[a] = fmaxf(chi[a], m);
        }
    }
    constexpr int NB = 32;
    float best = 3e38f;
    int best_axis = -1, best_bin = 0;
    for (int a = 0; a < 3; ++a) {
        const float ext = chi[a] - clo[a];
        if (!(ext > 0)) continue;
        float blo[NB][3], bhi[NB][3];
        int cnt[NB] = {};
        for (int k = 0; k < NB; ++k)
            for (int d = 0; d < 3; ++d) {
                blo[k][d] = 3e38f;
                bhi[k][d] = -3e38f;
            }
        for (int i = b; i < e; ++i) {
            float l[3], h[3];
            tbox(c, c.pos[i], l, h);
            int k = (int)((0.5f * (l[a] + h[a]) - clo[a]) / ext * NB);
            k = k < 0 ? 0 : k >= NB ? NB - 1 : k;
            ++cnt[k];
            for (int d = 0; d < 3; ++d) {
                blo[k][d] = fminf(blo[k][d], l[d]);
                bhi[k][d] = fmaxf(bhi[k][d], h[d]);
            }
        }
        float ra[NB];
        int rc[NB];
        {
            float l[3] = {3e38f, 3e38f, 3e38f}, h[3] = {-3e38f, -3e38f, -3e38f};
            int n = 0;
            for (int k = NB - 1; k > 0; --k) {
                for (int d = 0; d < 3; ++d) {
                    l[d] = fminf(l[d], blo[k][d]);
                    h[d] = fmaxf(h[d], bhi[k][d]);
                }
                n += cnt[k];
                ra[k] = n ? half_area(l, h) : 0.0f;
                rc[k] = n;
            }
        }
        float l[3] = {3e38f, 3e38f, 3e38f}, h[3] = {-3e38f, -3e38f, -3e38f};
        int n = 0;
        for (int k = 0; k < NB - 1; ++k) {
            for (int d = 0; d < 3; ++d) {
                l[d] = fminf(l[d], blo[k][d]);
                h[d] = fmaxf(h[d], bhi[k][d]);
            }
            n += cnt[k];
            if (n == 0 || rc[k + 1] == 0) continue;
            const float cost = half_area(l, h) * n + ra[k + 1] * rc[k + 1];
            if (cost < best) {
                best = cost;
                best_axis = a;
                best_bin = k;
            }
        }
    }
    int mid;
    if (best_axis < 0) {
        mid = (b + e) / 2;  // identical centroids
    } else {
        const int a = best_axis;
        const float ext = chi[a] - clo[a];
        auto it = std::stable_partition(c.pos.begin() + b, c.pos.begin() + e, [&](int32_t p) {
            float l[3], h[3];
            tbox(c, p, l, h);
            int k = (int)((0.5f * (l[a] + h[a]) - clo[a]) / ext * NB);
            k = k < 0 ? 0 : k >= NB ? NB - 1 : k;
            return k <= best_bin;
        });
        mid = (int)(it - c.pos.begin());
        if (mid == b || mid == e) mid = (b + e) / 2;
    }
    const int32_t L = rec(c, b, mid), R = rec(c, mid, e);
    BinNode& nd = c.bin[id];
    for (int a = 0; a < 3; ++a) {
        nd.lo[a] = lo[a];
        nd.hi[a] = hi[a];
    }
    nd.left = L;
    nd.right = R;
    nd.first = 0;
    nd.count = e - b;
    nd.parent = -1;
    nd.pad = 0;
    return id;
}

inline int32_t build(BinNode* bin, int n, const uint32_t* order, const float* tri_lo, const float* tri_hi) {
    Ctx c{bin, order, tri_lo, tri_hi, std::vector<int32_t>((size_t)n), 0};
    for (int i = 0; i < n; ++i) c.pos[i] = i;
    return rec(c, 0, n);
}
}  // namespace sah_exp

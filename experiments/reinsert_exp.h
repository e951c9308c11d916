// Experiment only, kept for the record ("reinsertion passes in the builder"): reinsertion of the
// binary tree between the clustering rounds and the wide collapse, run inside tests/emu/emu.cpp (included after
// pt_build.h, called as  root = reinsert_exp::optimise(bin, n, root, order, tri_lo, tri_hi, plans, passes)  right after
// the clustering) to learn how many node visits per ray a better tree saves before any device code was written.
//   EMU_REINSERT=<passes>                           sequential greedy passes, largest area first (the quality limit)
//   + EMU_REINSERT_PAR=1 [EMU_REINSERT_ROUNDS=r]    semantics of a parallel pass with path locks (Meister & Bittner)
//   + EMU_REINSERT_LIGHT=1                          six-node locks + cycle check: what pt_build.h now implements
// Not part of any build.
#pragma once
#include <cstdint>
#include <vector>
#include <algorithm>
#include <numeric>
#include <cstring>
#include <cstdlib>

namespace reinsert_exp {
using namespace pt;

struct Tree {
    BinNode* bin;
    std::vector<int32_t> leaf_parent;
    const uint32_t* order;
    const float *tri_lo, *tri_hi;
    int32_t root;

    void box(int32_t c, float* lo, float* hi) const {
        if (c < 0) {
            const uint32_t t = order[~c];
            for (int a = 0; a < 3; ++a) {
                lo[a] = tri_lo[3 * (size_t)t + a];
                hi[a] = tri_hi[3 * (size_t)t + a];
            }
        } else {
            for (int a = 0; a < 3; ++a) {
                lo[a] = bin[c].lo[a];
                hi[a] = bin[c].hi[a];
            }
        }
    }
    float area(int32_t c) const {
        float lo[3], hi[3];
        box(c, lo, hi);
        return half_area(lo, hi);
    }
    int32_t count(int32_t c) const { return c < 0 ? 1 : bin[c].count; }
    int32_t parent(int32_t c) const { return c < 0 ? leaf_parent[~c] : bin[c].parent; }
    void set_parent(int32_t c, int32_t p) {
        if (c < 0)
            leaf_parent[~c] = p;
        else
            bin[c].parent = p;
    }
    int32_t sibling(int32_t c) const {
        const BinNode& p = bin[parent(c)];
        return p.left == c ? p.right : p.left;
    }
    void refit_up(int32_t n) {
        while (n >= 0) {
            float l0[3], h0[3], l1[3], h1[3];
            box(bin[n].left, l0, h0);
            box(bin[n].right, l1, h1);
            for (int a = 0; a < 3; ++a) {
                bin[n].lo[a] = fminf(l0[a], l1[a]);
                bin[n].hi[a] = fmaxf(h0[a], h1[a]);
            }
            bin[n].count = count(bin[n].left) + count(bin[n].right);
            n = bin[n].parent;
        }
    }
};

inline float union_area(const float* alo, const float* ahi, const float* blo, const float* bhi) {
    float lo[3], hi[3];
    for (int a = 0; a < 3; ++a) {
        lo[a] = fminf(alo[a], blo[a]);
        hi[a] = fmaxf(ahi[a], bhi[a]);
    }
    return half_area(lo, hi);
}

// best new sibling for x; returns gain (> 0) and the position in *where
inline float find_best(const Tree& t, int32_t x, int32_t* where, int32_t* pivot_out = nullptr) {
    const int32_t p = t.parent(x);
    float xlo[3], xhi[3];
    t.box(x, xlo, xhi);
    const float ax = half_area(xlo, xhi);
    float best = 0.0f;
    *where = 0x7fffffff;
    float save = t.area(p);
    float blo[3], bhi[3];
    const int32_t s = t.sibling(x);
    t.box(s, blo, bhi);
    struct Item {
        int32_t o;
        float ind;
    };
    std::vector<Item> stack;
    int32_t pivot_now = p, pivot_best = p;
    auto search = [&](int32_t o0) {
        stack.clear();
        stack.push_back({o0, 0.0f});
        while (!stack.empty()) {
            const Item it = stack.back();
            stack.pop_back();
            float lo[3], hi[3];
            t.box(it.o, lo, hi);
            const float m = union_area(lo, hi, xlo, xhi);
            const float g = save - it.ind - m;
            if (g > best) {
                best = g;
                *where = it.o;
                pivot_best = pivot_now;
            }
            if (it.o >= 0) {
                const float ind2 = it.ind + m - half_area(lo, hi);
                if (save - ind2 - ax > best) {
                    stack.push_back({t.bin[it.o].left, ind2});
                    stack.push_back({t.bin[it.o].right, ind2});
                }
            }
        }
    };
    search(s);
    int32_t cur = p;
    while (t.bin[cur].parent >= 0) {
        const int32_t q = t.bin[cur].parent;
        const int32_t c2 = t.bin[q].left == cur ? t.bin[q].right : t.bin[q].left;
        pivot_now = q;
        search(c2);
        float lo[3], hi[3];
        t.box(c2, lo, hi);
        for (int a = 0; a < 3; ++a) {
            blo[a] = fminf(blo[a], lo[a]);
            bhi[a] = fmaxf(bhi[a], hi[a]);
        }
        const float shrink = t.area(q) - half_area(blo, bhi);
        if (shrink <= 0.0f && save - ax <= best) break;
        save += shrink;
        cur = q;
    }
    if (pivot_out) *pivot_out = pivot_best;
    return best;
}

inline void apply(Tree& t, int32_t x, int32_t o) {
    const int32_t p = t.parent(x);
    const int32_t s = t.sibling(x);
    const int32_t g = t.bin[p].parent;
    // s takes p's place
    if (g >= 0) {
        if (t.bin[g].left == p)
            t.bin[g].left = s;
        else
            t.bin[g].right = s;
    } else {
        t.root = s;
    }
    t.set_parent(s, g);
    t.refit_up(g);
    // p becomes the parent of (o, x) where o was
    const int32_t po = t.parent(o);
    if (po >= 0) {
        if (t.bin[po].left == o)
            t.bin[po].left = p;
        else
            t.bin[po].right = p;
    } else {
        t.root = p;
    }
    t.bin[p].parent = po;
    t.bin[p].left = o;
    t.bin[p].right = x;
    t.set_parent(o, p);
    t.set_parent(x, p);
    t.refit_up(p);
}

inline double sa_sum(const Tree& t, int n_inner_hint) {
    double s = 0;
    std::vector<int32_t> st{t.root};
    while (!st.empty()) {
        const int32_t c = st.back();
        st.pop_back();
        if (c < 0) continue;
        s += t.area(c);
        st.push_back(t.bin[c].left);
        st.push_back(t.bin[c].right);
    }
    (void)n_inner_hint;
    return s;
}

inline void full_refit(Tree& t) {
    std::vector<std::pair<int32_t, int>> st{{t.root, 0}};
    while (!st.empty()) {
        auto pr = st.back();
        st.pop_back();
        const int32_t c = pr.first;
        if (c < 0) continue;
        if (pr.second == 0) {
            st.push_back({c, 1});
            st.push_back({t.bin[c].left, 0});
            st.push_back({t.bin[c].right, 0});
        } else {
            float l0[3], h0[3], l1[3], h1[3];
            t.box(t.bin[c].left, l0, h0);
            t.box(t.bin[c].right, l1, h1);
            for (int a = 0; a < 3; ++a) {
                t.bin[c].lo[a] = fminf(l0[a], l1[a]);
                t.bin[c].hi[a] = fmaxf(h0[a], h1[a]);
            }
            t.bin[c].count = t.count(t.bin[c].left) + t.count(t.bin[c].right);
        }
    }
}

// one pass with the semantics of the parallel algorithm: all searches on one snapshot, path locks by the
// largest (gain, candidate) key, winners applied, one refit of the whole tree
inline long parallel_pass(Tree& t, const std::vector<int32_t>& cand, int n) {
    const size_t m = cand.size();
    std::vector<float> gain(m, 0.0f);
    std::vector<int32_t> target(m), pivot(m);
    auto slot = [&](int32_t c) { return c >= 0 ? (size_t)c : (size_t)(n - 1) + (size_t)(~c); };
    std::vector<uint64_t> lock((size_t)2 * n, 0);
    for (size_t i = 0; i < m; ++i) {
        const int32_t x = cand[i];
        if (x == t.root || t.parent(x) == t.root) continue;
        gain[i] = find_best(t, x, &target[i], &pivot[i]);
        if (target[i] == 0x7fffffff) gain[i] = 0.0f;
    }
    const bool light = getenv("EMU_REINSERT_LIGHT") != nullptr;
    auto for_locks = [&](size_t i, auto&& f) {
        const int32_t x = cand[i], p = t.parent(x), s = t.sibling(x), o = target[i], lca = pivot[i];
        f(x);
        f(s);
        f(p);
        if (light) {
            f(t.bin[p].parent);
            f(o);
            f(t.parent(o));
            return;
        }
        if (lca != p)
            for (int32_t a = t.bin[p].parent; a != lca; a = t.bin[a].parent) f(a);
        f(o);
        for (int32_t a = t.parent(o); a != lca; a = t.bin[a].parent) f(a);
    };
    std::vector<uint8_t> win(m, 0), dead(m, 0);
    const int rounds = getenv("EMU_REINSERT_ROUNDS") ? atoi(getenv("EMU_REINSERT_ROUNDS")) : 1;
    for (int r = 0; r < rounds; ++r) {
        for (size_t i = 0; i < m; ++i) {
            if (!(gain[i] > 0.0f) || win[i] || dead[i]) continue;
            bool blocked = false;
            for_locks(i, [&](int32_t c) { blocked = blocked || lock[slot(c)] == ~0ull; });
            if (blocked) dead[i] = 1;
        }
        for (size_t i = 0; i < m; ++i) {
            if (!(gain[i] > 0.0f) || win[i] || dead[i]) continue;
            uint32_t gb;
            memcpy(&gb, &gain[i], 4);
            const uint64_t key = ((uint64_t)gb << 32) | (uint32_t)i;
            for_locks(i, [&](int32_t c) { lock[slot(c)] = std::max(lock[slot(c)], key); });
        }
        long w = 0;
        for (size_t i = 0; i < m; ++i) {
            if (!(gain[i] > 0.0f) || win[i] || dead[i]) continue;
            uint32_t gb;
            memcpy(&gb, &gain[i], 4);
            const uint64_t key = ((uint64_t)gb << 32) | (uint32_t)i;
            bool ok = true, blocked = false;
            for_locks(i, [&](int32_t c) {
                ok = ok && lock[slot(c)] == key;
                blocked = blocked || lock[slot(c)] == ~0ull;
            });
            if (ok) { win[i] = 1; ++w; }
            if (blocked) dead[i] = 1;
        }
        // winners keep their nodes for the rest of the pass, everything else is released
        for (auto& l : lock) if (l != ~0ull) l = 0;
        for (size_t i = 0; i < m; ++i)
            if (win[i]) for_locks(i, [&](int32_t c) { lock[slot(c)] = ~0ull; });
        fprintf(stderr, "  round %d winners %ld\n", r, w);
    }
    if (light) {  // a winner whose insertion path passes through another winner's moved subtree root is dropped
        std::vector<uint8_t> is_x((size_t)2 * n, 0);
        for (size_t i = 0; i < m; ++i) if (win[i]) is_x[slot(cand[i])] = 1;
        long dropped = 0;
        std::vector<uint8_t> drop(m, 0);
        for (size_t i = 0; i < m; ++i) {
            if (!win[i]) continue;
            const int32_t o = target[i], lca = pivot[i];
            bool bad = false;
            for (int32_t a = t.parent(o); a != lca; a = t.bin[a].parent) bad = bad || is_x[slot(a)];
            if (bad) { drop[i] = 1; ++dropped; }
        }
        for (size_t i = 0; i < m; ++i) if (drop[i]) win[i] = 0;
        fprintf(stderr, "  dropped for cycles %ld\n", dropped);
    }
    long moved = 0;
    for (size_t i = 0; i < m; ++i) {
        if (!win[i]) continue;
        const int32_t x = cand[i], o = target[i];
        const int32_t p = t.parent(x), s = t.sibling(x), g = t.bin[p].parent;
        if (t.bin[g].left == p) t.bin[g].left = s; else t.bin[g].right = s;
        t.set_parent(s, g);
        const int32_t po = t.parent(o);
        if (t.bin[po].left == o) t.bin[po].left = p; else t.bin[po].right = p;
        t.bin[p].parent = po;
        t.bin[p].left = o;
        t.bin[p].right = x;
        t.set_parent(o, p);
        ++moved;
    }
    full_refit(t);
    long pos = 0;
    for (size_t i = 0; i < m; ++i) pos += gain[i] > 0.0f;
    fprintf(stderr, "  positive-gain candidates %ld\n", pos);
    return moved;
}

// returns the new root
inline int32_t optimise(BinNode* bin, int n, int32_t root, const uint32_t* order, const float* tri_lo, const float* tri_hi,
                        CollapsePlan* plans, int passes) {
    Tree t{bin, std::vector<int32_t>((size_t)n, -1), order, tri_lo, tri_hi, root};
    // leaf parents + the set of inner nodes in use
    std::vector<int32_t> inner;
    {
        std::vector<int32_t> st{root};
        bin[root].parent = -1;
        while (!st.empty()) {
            const int32_t c = st.back();
            st.pop_back();
            if (c < 0) continue;
            inner.push_back(c);
            for (int32_t ch : {bin[c].left, bin[c].right}) {
                t.set_parent(ch, c);
                st.push_back(ch);
            }
        }
    }
    fprintf(stderr, "reinsert: start sa %.6g\n", sa_sum(t, n));
    if (getenv("EMU_REINSERT_PAR")) {
        std::vector<int32_t> cand(inner);
        for (int i = 0; i < n; ++i) cand.push_back(~i);
        for (int pass = 0; pass < passes; ++pass) {
            const long moved = parallel_pass(t, cand, n);
            fprintf(stderr, "reinsert(par): pass %d moved %ld sa %.6g\n", pass, moved, sa_sum(t, n));
        }
        passes = 0;
    }
    for (int pass = 0; pass < passes; ++pass) {
        // candidates: every inner node and leaf, largest area first
        std::vector<int32_t> cand(inner);
        for (int i = 0; i < n; ++i) cand.push_back(~i);
        std::vector<float> ar(cand.size());
        for (size_t i = 0; i < cand.size(); ++i) ar[i] = t.area(cand[i]);
        std::vector<uint32_t> idx(cand.size());
        std::iota(idx.begin(), idx.end(), 0u);
        std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return ar[a] > ar[b]; });
        long moved = 0;
        for (uint32_t k : idx) {
            const int32_t x = cand[k];
            if (x == t.root) continue;
            const int32_t p = t.parent(x);
            if (p == t.root) continue;
            int32_t o;
            if (p < 0 || p >= n || t.bin[p].parent < 0 || t.bin[p].parent >= n) { fprintf(stderr, "bad p %d x %d par %d root %d\n", p, x, p >= 0 ? t.bin[p].parent : -9, t.root); abort(); }
            if (t.bin[p].left != x && t.bin[p].right != x) { fprintf(stderr, "x %d not child of p %d (%d %d)\n", x, p, t.bin[p].left, t.bin[p].right); abort(); }
            const float g = find_best(t, x, &o);
            if (g > 0.0f && o != 0x7fffffff) {
                apply(t, x, o);
                ++moved;
            }
        }
        fprintf(stderr, "reinsert: pass %d moved %ld sa %.6g\n", pass, moved, sa_sum(t, n));
    }
    // plans + first (unused by the collapse when leaves list positions) bottom-up
    {
        std::vector<std::pair<int32_t, int>> st{{t.root, 0}};
        while (!st.empty()) {
            auto [c, state] = st.back();
            st.pop_back();
            if (c < 0) continue;
            if (state == 0) {
                st.push_back({c, 1});
                st.push_back({bin[c].left, 0});
                st.push_back({bin[c].right, 0});
            } else {
                plan_node(c, bin, plans, order, tri_lo, tri_hi);
            }
        }
    }
    return t.root;
}
}  // namespace reinsert_exp

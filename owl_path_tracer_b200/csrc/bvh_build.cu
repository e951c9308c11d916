// bvh_build.cu — on-device BVH build: kernels around the per-element steps of pt_build.h.
// Replaces optixAccelBuild (reference: application.cpp:131-140 via OWL).  Not on the
// per-frame path: runs once per b200pt_build_accel.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "device_ctx.h"
#include "pt_build.h"

namespace pt {

__global__ void k_tri_prepare(const f4* __restrict__ pos, const int32_t* __restrict__ idx4, uint32_t n, float pad,
                              float3 slo, float3 sinv, float* __restrict__ tri_lo, float* __restrict__ tri_hi,
                              uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float lo[3], hi[3];
    tri_box(pos, idx4, i, pad, lo, hi);
    for (int a = 0; a < 3; ++a) {
        tri_lo[3 * (size_t)i + a] = lo[a];
        tri_hi[3 * (size_t)i + a] = hi[a];
    }
    const float l[3] = {slo.x, slo.y, slo.z}, s[3] = {sinv.x, sinv.y, sinv.z};
    keys[i] = morton_key(lo, hi, l, s);
    vals[i] = i;
}

__global__ void k_ploc_init(uint32_t n, const uint32_t* __restrict__ sorted_tri, const float* __restrict__ tri_lo,
                            const float* __restrict__ tri_hi, int32_t* __restrict__ cid, Box6* __restrict__ cbox) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    cid[i] = ~(int32_t)i;
    const uint32_t t = sorted_tri[i];
    Box6 b;
    for (int a = 0; a < 3; ++a) {
        b.lo[a] = tri_lo[3 * (size_t)t + a];
        b.hi[a] = tri_hi[3 * (size_t)t + a];
    }
    cbox[i] = b;
}

__global__ void k_ploc_nearest(int m, const Box6* __restrict__ cbox, int32_t* __restrict__ nn) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    ploc_nearest(i, m, cbox, nn);
}

__global__ void k_ploc_flags(int m, const int32_t* __restrict__ nn, unsigned long long* __restrict__ flags) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    flags[i] = ploc_flags(i, nn);
}

__global__ void k_ploc_emit(int m, const int32_t* __restrict__ nn, const unsigned long long* __restrict__ incl,
                            const int32_t* __restrict__ cid, const Box6* __restrict__ cbox, int32_t* __restrict__ out_cid,
                            Box6* __restrict__ out_cbox, BinNode* bin, CollapsePlan* plans, uint32_t node_base,
                            const uint32_t* __restrict__ sorted_tri, const float* __restrict__ tri_lo,
                            const float* __restrict__ tri_hi) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint64_t excl = incl[i] - ploc_flags(i, nn);
    ploc_emit(i, nn, excl, cid, cbox, out_cid, out_cbox, bin, plans, node_base, sorted_tri, tri_lo, tri_hi);
}

__global__ void k_radix_tree(const uint64_t* __restrict__ keys, int n, BinNode* nodes, int32_t* leaf_parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    radix_tree_node(keys, n, i, nodes, leaf_parent);
}

__global__ void k_refit(int n, const int32_t* __restrict__ leaf_parent, BinNode* nodes, int32_t* flags,
                        const uint32_t* __restrict__ sorted_tri, const float* __restrict__ tri_lo,
                        const float* __restrict__ tri_hi, CollapsePlan* plans) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    refit_from_leaf(k, leaf_parent, nodes, flags, sorted_tri, tri_lo, tri_hi, plans);
}

// ---- reinsertion passes (pt_build.h, "reinsertion optimisation"); c = candidate = lock slot ----
__global__ void k_ri_link(int n_inner, ReinsertTree t) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_inner) ri_link_children(i, t);
}
__global__ void k_ri_find(int m, ReinsertTree t, float* gain, int32_t* target, int32_t* pivot, uint8_t* state) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m) ri_find(c, t, gain, target, pivot, state);
}
__global__ void k_ri_bid(int m, ReinsertTree t, const float* gain, const int32_t* target, uint8_t* state,
                         unsigned long long* lock) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m) ri_bid(c, t, gain, target, state, lock);
}
__global__ void k_ri_settle(int m, ReinsertTree t, const float* gain, const int32_t* target, uint8_t* state,
                            const unsigned long long* lock) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m) ri_settle(c, t, gain, target, state, lock);
}
__global__ void k_ri_hold(int m, ReinsertTree t, const int32_t* target, const uint8_t* state, unsigned long long* lock) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m) ri_hold(c, t, target, state, lock);
}
__global__ void k_ri_mark(int m, const uint8_t* state, unsigned long long* lock) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m) ri_mark_moved(c, state, lock);
}
__global__ void k_ri_drop(int m, ReinsertTree t, const int32_t* target, const int32_t* pivot, uint8_t* state,
                          const unsigned long long* lock) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m) ri_drop_crossing(c, t, target, pivot, state, lock);
}
__global__ void k_ri_apply(int m, ReinsertTree t, const int32_t* target, const uint8_t* state) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m) ri_apply(c, t, target, state);
}

__global__ void k_collapse(const CollapseItem* __restrict__ items, uint32_t n_items, const BinNode* __restrict__ bin,
                           const uint32_t* __restrict__ sorted_tri, const float* __restrict__ tri_lo,
                           const float* __restrict__ tri_hi, const f4* __restrict__ pos,
                           const int32_t* __restrict__ idx4, Node8* nodes_out, uint32_t* node_count, TriPair* tris_out,
                           uint32_t* tri_count, CollapseItem* next, uint32_t* next_count, int n_tris,
                           const CollapsePlan* __restrict__ plans) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_items) return;
    collapse_node(items[i], bin, sorted_tri, tri_lo, tri_hi, pos, idx4, nodes_out, node_count, tris_out, tri_count, next,
                  next_count, n_tris, plans);
}

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            err = std::string(#x) + ": " + cudaGetErrorString(e_);                              \
            goto fail;                                                                          \
        }                                                                                       \
    } while (0)

bool build_bvh_device(const f4* d_pos, const int32_t* d_idx4, uint32_t n, const float scene_lo[3],
                      const float scene_hi[3], float max_abs, DeviceBVH* out, cudaStream_t st, std::string& err) {
    float *tri_lo = nullptr, *tri_hi = nullptr;
    uint64_t *keys = nullptr, *keys2 = nullptr;
    uint32_t *vals = nullptr, *vals2 = nullptr, *counters = nullptr;
    BinNode* bin = nullptr;
    int32_t *cid = nullptr, *cid2 = nullptr, *nn = nullptr;
    Box6 *cbox = nullptr, *cbox2 = nullptr;
    unsigned long long *pflags = nullptr, *pscan = nullptr;
    CollapsePlan* plans = nullptr;
    void* tmp2 = nullptr;
    size_t tmp2_bytes = 0;
    int32_t root = ~0;
    CollapseItem *qa = nullptr, *qb = nullptr;
    Node8* nodes = nullptr;
    TriPair* tris = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    char* arena = nullptr;
    const int B = 256;
    const uint32_t nb = (n + B - 1) / B;
    // absolute box padding: covers the rounding of the slab test for rays whose origin
    // lies anywhere within a few scene diameters (pt_bvh8.h header)
    const float pad = max_abs * 3.8146973e-6f /* 2^-18 */ + 1e-30f;
    float3 slo = make_float3(scene_lo[0] - pad, scene_lo[1] - pad, scene_lo[2] - pad);
    float3 sinv;
    {
        const float ex = scene_hi[0] - scene_lo[0] + 2 * pad, ey = scene_hi[1] - scene_lo[1] + 2 * pad,
                    ez = scene_hi[2] - scene_lo[2] + 2 * pad;
        sinv = make_float3(ex > 0 ? 1.0f / ex : 0.0f, ey > 0 ? 1.0f / ey : 0.0f, ez > 0 ? 1.0f / ez : 0.0f);
    }
    uint32_t h_counters[4];
    int levels = 0;
    uint32_t n_items = 1;

    // one arena for every temporary (a dozen cudaMalloc/cudaFree pairs per build cost more host time
    // than the build itself); the triangles of the finished tree get their own allocation
    {
        size_t off = 0;
        auto carve = [&](size_t bytes) {
            const size_t at = off;
            off += (bytes + 255) & ~(size_t)255;
            return at;
        };
        const size_t N = (size_t)n;
        const size_t o_tri_lo = carve(sizeof(float) * 3 * N), o_tri_hi = carve(sizeof(float) * 3 * N);
        const size_t o_keys = carve(sizeof(uint64_t) * N), o_keys2 = carve(sizeof(uint64_t) * N);
        const size_t o_vals = carve(sizeof(uint32_t) * N), o_vals2 = carve(sizeof(uint32_t) * N);
        const size_t o_bin = carve(sizeof(BinNode) * N);
        const size_t o_cid = carve(sizeof(int32_t) * N), o_cid2 = carve(sizeof(int32_t) * N), o_nn = carve(sizeof(int32_t) * N);
        const size_t o_cbox = carve(sizeof(Box6) * N), o_cbox2 = carve(sizeof(Box6) * N);
        const size_t o_pflags = carve(sizeof(unsigned long long) * N), o_pscan = carve(sizeof(unsigned long long) * N);
        const size_t o_plans = carve(sizeof(CollapsePlan) * N);
        const size_t o_qa = carve(sizeof(CollapseItem) * N), o_qb = carve(sizeof(CollapseItem) * N);
        const size_t o_nodes = carve(sizeof(Node8) * N);
        const size_t o_counters = carve(sizeof(uint32_t) * 4);
        size_t sort_bytes = 0, scan_bytes = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (uint64_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr,
                                           (uint32_t*)nullptr, (int)n, 0, 63, st));
        CK(cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)n, st));
        tmp_bytes = sort_bytes;
        tmp2_bytes = scan_bytes;
        const size_t o_tmp = carve(sort_bytes), o_tmp2 = carve(scan_bytes);
        CK(cudaMalloc(&arena, off));
        tri_lo = (float*)(arena + o_tri_lo); tri_hi = (float*)(arena + o_tri_hi);
        keys = (uint64_t*)(arena + o_keys); keys2 = (uint64_t*)(arena + o_keys2);
        vals = (uint32_t*)(arena + o_vals); vals2 = (uint32_t*)(arena + o_vals2);
        bin = (BinNode*)(arena + o_bin);
        cid = (int32_t*)(arena + o_cid); cid2 = (int32_t*)(arena + o_cid2); nn = (int32_t*)(arena + o_nn);
        cbox = (Box6*)(arena + o_cbox); cbox2 = (Box6*)(arena + o_cbox2);
        pflags = (unsigned long long*)(arena + o_pflags); pscan = (unsigned long long*)(arena + o_pscan);
        plans = (CollapsePlan*)(arena + o_plans);
        qa = (CollapseItem*)(arena + o_qa); qb = (CollapseItem*)(arena + o_qb);
        nodes = (Node8*)(arena + o_nodes);
        counters = (uint32_t*)(arena + o_counters);
        tmp = arena + o_tmp;
        tmp2 = arena + o_tmp2;
    }
    CK(cudaMalloc(&tris, sizeof(TriPair) * (size_t)n));

    k_tri_prepare<<<nb, B, 0, st>>>(d_pos, d_idx4, n, pad, slo, sinv, tri_lo, tri_hi, keys, vals);
    CK(cudaGetLastError());
    CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, vals, vals2, (int)n, 0, 63, st));
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (n >= 2 && attempt == 0) {
            // parallel locally-ordered clustering: rounds of nearest neighbour / flags / scan / emit
            k_ploc_init<<<nb, B, 0, st>>>(n, vals2, tri_lo, tri_hi, cid, cbox);
            CK(cudaGetLastError());
            uint32_t m = n, node_base = 0;
            int rounds = 0;
            while (m > 1) {
                const uint32_t mb = (m + B - 1) / B;
                k_ploc_nearest<<<mb, B, 0, st>>>((int)m, cbox, nn);
                k_ploc_flags<<<mb, B, 0, st>>>((int)m, nn, pflags);
                CK(cub::DeviceScan::InclusiveSum(tmp2, tmp2_bytes, pflags, pscan, (int)m, st));
                k_ploc_emit<<<mb, B, 0, st>>>((int)m, nn, pscan, cid, cbox, cid2, cbox2, bin, plans, node_base, vals2, tri_lo,
                                             tri_hi);
                CK(cudaGetLastError());
                unsigned long long total = 0;
                CK(cudaMemcpyAsync(&total, pscan + (m - 1), sizeof total, cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                const uint32_t merged = (uint32_t)(total >> 32);
                if (merged == 0 || ++rounds > 4096) {
                    err = "ploc made no progress";
                    goto fail;
                }
                m = (uint32_t)(total & 0xFFFFFFFFull);
                node_base += merged;
                std::swap(cid, cid2);
                std::swap(cbox, cbox2);
            }
            CK(cudaMemcpyAsync(&root, cid, sizeof root, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (n >= 3 && kReinsertPasses > 0) {
                // the sort keys, neighbour lists and cluster boxes are dead by now: their space holds the
                // per-candidate results (2n-1 candidates) and the lock words
                int32_t* leaf_parent = nn;
                int32_t* arrive_flags = cid2;
                float* gain = (float*)keys;
                int32_t* target = (int32_t*)pscan;  // (keys2 stays: the radix-tree fallback reads it)
                int32_t* pivot = (int32_t*)pflags;
                unsigned long long* lock = (unsigned long long*)cbox;
                uint8_t* state = (uint8_t*)cbox2;
                static_assert(sizeof(Box6) >= 2 * sizeof(unsigned long long), "lock words fit the cluster boxes");
                const int m = 2 * (int)n - 1;
                const uint32_t cb = (uint32_t)(m + B - 1) / B;
                const ReinsertTree rt{bin, leaf_parent, vals2, tri_lo, tri_hi, (int32_t)n, root};
                k_ri_link<<<nb, B, 0, st>>>((int)n - 1, rt);
                CK(cudaGetLastError());
                for (int pass = 0; pass < kReinsertPasses; ++pass) {
                    k_ri_find<<<cb, B, 0, st>>>(m, rt, gain, target, pivot, state);
                    CK(cudaGetLastError());
                    for (int round = 0; round < kReinsertRounds; ++round) {
                        CK(cudaMemsetAsync(lock, 0, sizeof(unsigned long long) * (size_t)m, st));
                        if (round > 0) k_ri_hold<<<cb, B, 0, st>>>(m, rt, target, state, lock);
                        k_ri_bid<<<cb, B, 0, st>>>(m, rt, gain, target, state, lock);
                        k_ri_settle<<<cb, B, 0, st>>>(m, rt, gain, target, state, lock);
                        CK(cudaGetLastError());
                    }
                    CK(cudaMemsetAsync(lock, 0, sizeof(unsigned long long) * (size_t)m, st));
                    k_ri_mark<<<cb, B, 0, st>>>(m, state, lock);
                    k_ri_drop<<<cb, B, 0, st>>>(m, rt, target, pivot, state, lock);
                    k_ri_apply<<<cb, B, 0, st>>>(m, rt, target, state);
                    CK(cudaMemsetAsync(arrive_flags, 0, sizeof(int32_t) * (size_t)n, st));
                    k_refit<<<nb, B, 0, st>>>((int)n, leaf_parent, bin, arrive_flags, vals2, tri_lo, tri_hi, plans);
                    CK(cudaGetLastError());
                }
            }
        }
        if (n >= 2 && attempt == 1) {
            // fallback: binary radix tree over the Morton keys (Karras 2012), depth bounded by the key
            // length; used when the clustering produced a tree deeper than the traversal stack
            int32_t *leaf_parent = cid, *arrive_flags = cid2;
            CK(cudaMemsetAsync(arrive_flags, 0, sizeof(int32_t) * (size_t)n, st));
            CK(cudaMemsetAsync(leaf_parent, 0xFF, sizeof(int32_t) * (size_t)n, st));
            k_radix_tree<<<nb, B, 0, st>>>(keys2, (int)n, bin, leaf_parent);
            CK(cudaGetLastError());
            k_refit<<<nb, B, 0, st>>>((int)n, leaf_parent, bin, arrive_flags, vals2, tri_lo, tri_hi, plans);
            CK(cudaGetLastError());
            root = 0;
        }
        // collapse, level by level.  counters: [0] wide nodes, [1] triangles, [2] next queue
        h_counters[0] = 1;
        h_counters[1] = 0;
        h_counters[2] = 0;
        h_counters[3] = 0;
        CK(cudaMemcpyAsync(counters, h_counters, sizeof h_counters, cudaMemcpyHostToDevice, st));
        {
            CollapseItem root_item{root, 0u};
            CK(cudaMemcpyAsync(qa, &root_item, sizeof root_item, cudaMemcpyHostToDevice, st));
            CK(cudaStreamSynchronize(st));
        }
        levels = 0;
        n_items = 1;
        out->level_end.assign(1, 1u);  // level 0: the root
        while (n_items > 0) {
            ++levels;
            k_collapse<<<(n_items + 63) / 64, 64, 0, st>>>(qa, n_items, bin, vals2, tri_lo, tri_hi, d_pos, d_idx4, nodes,
                                                           counters + 0, tris, counters + 1, qb, counters + 2, (int)n, plans);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(h_counters, counters, sizeof h_counters, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            n_items = h_counters[2];
            if (n_items) out->level_end.push_back(h_counters[0]);  // the children this level allocated form the next one
            CK(cudaMemsetAsync(counters + 2, 0, sizeof(uint32_t), st));
            std::swap(qa, qb);
            if (levels > 4 * kStackSize) {
                err = "bvh collapse did not terminate";
                goto fail;
            }
        }
        if (levels <= kStackSize) break;
        if (attempt == 1) {
            err = "bvh deeper than the traversal stack (" + std::to_string(levels) + " levels)";
            goto fail;
        }
    }
    // h_counters[1] counts triangle-pair blocks: every leaf of k triangles takes ceil(k/2) of them
    if (h_counters[1] < (n + 1) / 2 || h_counters[1] > n) {
        err = "bvh collapse lost triangles: " + std::to_string(h_counters[1]) + " pair blocks for " + std::to_string(n) + " triangles";
        goto fail;
    }
    out->n_nodes = h_counters[0];
    out->n_tris = h_counters[1];
    out->levels = levels;
    CK(cudaMalloc(&out->nodes, sizeof(Node8) * (size_t)out->n_nodes));
    CK(cudaMemcpyAsync(out->nodes, nodes, sizeof(Node8) * (size_t)out->n_nodes, cudaMemcpyDeviceToDevice, st));
    out->tris = tris;
    tris = nullptr;
    CK(cudaStreamSynchronize(st));
    cudaFree(arena);
    return true;
fail:
    cudaFree(arena);
    cudaFree(tris);
    return false;
}

__global__ void k_refit_wide(uint32_t first, uint32_t count, Node8* nodes, TriPair* tris, const f4* __restrict__ pos,
                             const int32_t* __restrict__ idx4, float pad, float* node_box) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) refit_wide_node(first + i, nodes, tris, pos, idx4, pad, node_box);
}

bool refit_bvh_device(const f4* d_pos, const int32_t* d_idx4, float max_abs, DeviceBVH* bvh, float* d_node_box, cudaStream_t st,
                      std::string& err) {
    if (!bvh->nodes || bvh->level_end.empty() || bvh->level_end.back() != bvh->n_nodes) {
        err = "no tree to refit";
        return false;
    }
    const float pad = max_abs * 3.8146973e-6f /* 2^-18 */ + 1e-30f;  // build_bvh_device's margin
    for (int l = (int)bvh->level_end.size() - 1; l >= 0; --l) {
        const uint32_t first = l ? bvh->level_end[(size_t)l - 1] : 0u, count = bvh->level_end[(size_t)l] - first;
        if (!count) continue;
        k_refit_wide<<<(count + 63) / 64, 64, 0, st>>>(first, count, bvh->nodes, bvh->tris, d_pos, d_idx4, pad, d_node_box);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) {
            err = std::string("k_refit_wide: ") + cudaGetErrorString(e);
            return false;
        }
    }
    return true;
}

}  // namespace pt

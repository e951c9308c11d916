// bvh_build.cu — on-device BVH build: kernels around the per-element steps of pt_build.h.
// Replaces optixAccelBuild (reference: application.cpp:131-140 via OWL).  Not on the
// per-frame path: runs once per b200pt_build_accel.
#include <cub/device/device_radix_sort.cuh>
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

__global__ void k_radix_tree(const uint64_t* __restrict__ keys, int n, BinNode* nodes, int32_t* leaf_parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    radix_tree_node(keys, n, i, nodes, leaf_parent);
}

__global__ void k_refit(int n, const int32_t* __restrict__ leaf_parent, BinNode* nodes, int32_t* flags,
                        const uint32_t* __restrict__ sorted_tri, const float* __restrict__ tri_lo,
                        const float* __restrict__ tri_hi) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    refit_from_leaf(k, leaf_parent, nodes, flags, sorted_tri, tri_lo, tri_hi);
}

__global__ void k_collapse(const CollapseItem* __restrict__ items, uint32_t n_items, const BinNode* __restrict__ bin,
                           const uint32_t* __restrict__ sorted_tri, const float* __restrict__ tri_lo,
                           const float* __restrict__ tri_hi, const f4* __restrict__ pos,
                           const int32_t* __restrict__ idx4, Node8* nodes_out, uint32_t* node_count, Tri48* tris_out,
                           uint32_t* tri_count, CollapseItem* next, uint32_t* next_count, int n_tris) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_items) return;
    collapse_node(items[i], bin, sorted_tri, tri_lo, tri_hi, pos, idx4, nodes_out, node_count, tris_out, tri_count, next,
                  next_count, n_tris);
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
    int32_t *leaf_parent = nullptr, *flags = nullptr;
    CollapseItem *qa = nullptr, *qb = nullptr;
    Node8* nodes = nullptr;
    Tri48* tris = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
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

    CK(cudaMalloc(&tri_lo, sizeof(float) * 3 * (size_t)n));
    CK(cudaMalloc(&tri_hi, sizeof(float) * 3 * (size_t)n));
    CK(cudaMalloc(&keys, sizeof(uint64_t) * (size_t)n));
    CK(cudaMalloc(&keys2, sizeof(uint64_t) * (size_t)n));
    CK(cudaMalloc(&vals, sizeof(uint32_t) * (size_t)n));
    CK(cudaMalloc(&vals2, sizeof(uint32_t) * (size_t)n));
    CK(cudaMalloc(&bin, sizeof(BinNode) * (size_t)n));
    CK(cudaMalloc(&leaf_parent, sizeof(int32_t) * (size_t)n));
    CK(cudaMalloc(&flags, sizeof(int32_t) * (size_t)n));
    CK(cudaMalloc(&qa, sizeof(CollapseItem) * (size_t)n));
    CK(cudaMalloc(&qb, sizeof(CollapseItem) * (size_t)n));
    CK(cudaMalloc(&nodes, sizeof(Node8) * (size_t)n));
    CK(cudaMalloc(&tris, sizeof(Tri48) * (size_t)n));
    CK(cudaMalloc(&counters, sizeof(uint32_t) * 4));
    CK(cudaMemsetAsync(flags, 0, sizeof(int32_t) * (size_t)n, st));
    CK(cudaMemsetAsync(leaf_parent, 0xFF, sizeof(int32_t) * (size_t)n, st));

    k_tri_prepare<<<nb, B, 0, st>>>(d_pos, d_idx4, n, pad, slo, sinv, tri_lo, tri_hi, keys, vals);
    CK(cudaGetLastError());
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, vals, vals2, (int)n, 0, 63, st));
    CK(cudaMalloc(&tmp, tmp_bytes));
    CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, vals, vals2, (int)n, 0, 63, st));
    if (n >= 2) {
        k_radix_tree<<<nb, B, 0, st>>>(keys2, (int)n, bin, leaf_parent);
        CK(cudaGetLastError());
        k_refit<<<nb, B, 0, st>>>((int)n, leaf_parent, bin, flags, vals2, tri_lo, tri_hi);
        CK(cudaGetLastError());
    }
    // collapse, level by level.  counters: [0] wide nodes, [1] triangles, [2] next queue
    h_counters[0] = 1;
    h_counters[1] = 0;
    h_counters[2] = 0;
    h_counters[3] = 0;
    CK(cudaMemcpyAsync(counters, h_counters, sizeof h_counters, cudaMemcpyHostToDevice, st));
    {
        CollapseItem root{n >= 2 ? 0 : ~0, 0u};
        CK(cudaMemcpyAsync(qa, &root, sizeof root, cudaMemcpyHostToDevice, st));
    }
    while (n_items > 0) {
        ++levels;
        k_collapse<<<(n_items + 63) / 64, 64, 0, st>>>(qa, n_items, bin, vals2, tri_lo, tri_hi, d_pos, d_idx4, nodes,
                                                       counters + 0, tris, counters + 1, qb, counters + 2, (int)n);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(h_counters, counters, sizeof h_counters, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        n_items = h_counters[2];
        CK(cudaMemsetAsync(counters + 2, 0, sizeof(uint32_t), st));
        std::swap(qa, qb);
        if (levels > 4 * kStackSize) {
            err = "bvh collapse did not terminate";
            goto fail;
        }
    }
    if (levels > kStackSize) {
        err = "bvh deeper than the traversal stack (" + std::to_string(levels) + " levels)";
        goto fail;
    }
    if (h_counters[1] != n) {
        err = "bvh collapse lost triangles: " + std::to_string(h_counters[1]) + " of " + std::to_string(n);
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
    cudaFree(tri_lo); cudaFree(tri_hi); cudaFree(keys); cudaFree(keys2); cudaFree(vals); cudaFree(vals2);
    cudaFree(bin); cudaFree(leaf_parent); cudaFree(flags); cudaFree(qa); cudaFree(qb); cudaFree(nodes);
    cudaFree(counters); cudaFree(tmp);
    return true;
fail:
    cudaFree(tri_lo); cudaFree(tri_hi); cudaFree(keys); cudaFree(keys2); cudaFree(vals); cudaFree(vals2);
    cudaFree(bin); cudaFree(leaf_parent); cudaFree(flags); cudaFree(qa); cudaFree(qb); cudaFree(nodes);
    cudaFree(tris); cudaFree(counters); cudaFree(tmp);
    return false;
}

}  // namespace pt

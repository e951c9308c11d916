// mesh_dedup.cu — create_mesh (reference: path_tracer/src/utils/mesh_loader.cpp:9-83) on the device.
//
// The reference walks a shape's corners in order and keeps a std::map from global position index to local vertex id:
// a position index gets the next local id the first time a corner uses it; the normal / texcoord stored for a local
// vertex is the one of that first corner (Q17).  The same mapping without a serial walk:
//   1. first[v]   = smallest corner number that uses position index v          atomicMin over all corners
//   2. is_first[c] = first[corner_v[c]] == c                                     one flag per corner
//   3. rank        = exclusive prefix sum of is_first in corner order            cub::DeviceScan (a library scan on a
//                                                                                one-off loader stage, like the builder's)
//      -> the local id of a vertex = the number of first occurrences before its own first corner: exactly the order in
//         which the serial walk hands out ids
//   4. out_idx[c] = rank[first[corner_v[c]]];  first corners also copy position, normal and texcoord.
// Deterministic (min and a fixed-order scan).  Regular shapes only — every corner carries a normal index or none does,
// likewise for texcoords: with a mix the reference's `while (normals.size() < vertices.size())` fill depends on the walk,
// and the entry point reports B200PT_ERR_INVALID so that the caller keeps the host routine for that shape.
#include <cub/device/device_scan.cuh>

#include "device_ctx.h"

namespace pt {

__global__ void __launch_bounds__(256) k_dedup_first(const int32_t* __restrict__ cv, uint32_t n, int32_t lo, uint32_t* __restrict__ first) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) atomicMin(first + (cv[c] - lo), c);
}
__global__ void __launch_bounds__(256) k_dedup_flag(const int32_t* __restrict__ cv, uint32_t n, int32_t lo, const uint32_t* __restrict__ first,
                                                    uint32_t* __restrict__ is_first) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) is_first[c] = first[cv[c] - lo] == c ? 1u : 0u;
}
__global__ void __launch_bounds__(256) k_dedup_emit(const int32_t* __restrict__ cv, const int32_t* __restrict__ cn, const int32_t* __restrict__ ct,
                                                    uint32_t n, int32_t lo, const uint32_t* __restrict__ first, const uint32_t* __restrict__ rank,
                                                    const float* __restrict__ av, const float* __restrict__ an, const float* __restrict__ at,
                                                    int32_t* __restrict__ out_idx, float* __restrict__ ov, float* __restrict__ on, float* __restrict__ ot) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    const int32_t v = cv[c];
    const uint32_t f = first[v - lo];
    const uint32_t id = rank[f];
    out_idx[c] = (int32_t)id;
    if (f == c) {
        for (int k = 0; k < 3; ++k) ov[3 * (size_t)id + k] = av[3 * (size_t)v + k];
        if (on) for (int k = 0; k < 3; ++k) on[3 * (size_t)id + k] = an[3 * (size_t)cn[c] + k];
        if (ot) for (int k = 0; k < 2; ++k) ot[2 * (size_t)id + k] = at[2 * (size_t)ct[c] + k];
    }
}

// All pointers are device pointers; `first` has (hi - lo + 1) entries, is_first / rank have n; returns the number of
// local vertices through *n_unique (host).  scratch: cub's temporary storage (size from dedup_scan_bytes).
size_t dedup_scan_bytes(uint32_t n) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n);
    return bytes;
}
cudaError_t dedup_mesh_device(const int32_t* cv, const int32_t* cn, const int32_t* ct, uint32_t n, int32_t lo, uint32_t range,
                              const float* av, const float* an, const float* at, uint32_t* first, uint32_t* is_first, uint32_t* rank,
                              void* scratch, size_t scratch_bytes, int32_t* out_idx, float* ov, float* on, float* ot,
                              uint32_t* n_unique, cudaStream_t st) {
    const unsigned g = (n + 255) / 256;
    cudaError_t e = cudaMemsetAsync(first, 0xFF, sizeof(uint32_t) * range, st);
    if (e != cudaSuccess) return e;
    k_dedup_first<<<g, 256, 0, st>>>(cv, n, lo, first);
    k_dedup_flag<<<g, 256, 0, st>>>(cv, n, lo, first, is_first);
    e = cub::DeviceScan::ExclusiveSum(scratch, scratch_bytes, is_first, rank, (int)n, st);
    if (e != cudaSuccess) return e;
    k_dedup_emit<<<g, 256, 0, st>>>(cv, cn, ct, n, lo, first, rank, av, an, at, out_idx, ov, on, ot);
    uint32_t tail[2] = {0, 0};
    e = cudaMemcpyAsync(&tail[0], rank + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&tail[1], is_first + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    *n_unique = tail[0] + tail[1];
    return e;
}

}  // namespace pt

// device_ctx.h — context object behind the C ABI (include/b200pt.h) and the launch
// wrappers shared between api.cu, wavefront.cu and bvh_build.cu.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/b200pt.h"
#include "pt_integrator.h"

namespace pt {

struct DeviceBVH {
    Node8* nodes = nullptr;
    TriPair* tris = nullptr;
    uint32_t n_nodes = 0, n_tris = 0;
    int levels = 0;
    std::vector<uint32_t> level_end;  // nodes of level l are [level_end[l-1], level_end[l]) (level 0 = the root): refit order
};

bool build_bvh_device(const f4* d_pos, const int32_t* d_idx4, uint32_t n, const float scene_lo[3],
                      const float scene_hi[3], float max_abs, DeviceBVH* out, cudaStream_t st, std::string& err);

// Vertices moved, topology kept: triangle blocks, child boxes and quantisation grids of the wide tree are
// recomputed bottom-up, one launch per level (pt_build.h refit_wide_node).  node_box: scratch, 6 floats per node.
bool refit_bvh_device(const f4* d_pos, const int32_t* d_idx4, float max_abs, DeviceBVH* bvh, float* d_node_box, cudaStream_t st,
                      std::string& err);

// Path state, structure of arrays with 16-byte elements (one LDG.128 / STG.128 each).
// The state of the paths in flight is kept COMPACT: entry i of set (k & 1) belongs to the i-th
// queued path of wavefront iteration k; k_shade writes the survivors to consecutive entries of the
// other set.  Both kernels therefore read and write the state with coalesced accesses and without
// going through the queue; only the per-pixel sum is addressed by lane (= pixel).
struct LaneArrays {
    f4* ray_o[2] = {nullptr, nullptr};   // origin.xyz, -
    f4* ray_d[2] = {nullptr, nullptr};   // direction.xyz, control word (pack_ctl)
    f4* thr[2] = {nullptr, nullptr};     // throughput.xyz, RNG state
    f4* hit = nullptr;                   // t, u, v, triangle id   (written by k_extend, same index)
    uint32_t* queue[2] = {nullptr, nullptr};  // lane (pixel) id of entry i
    f4* sum = nullptr;     // per lane: sum of finished samples .xyz, sum of their squared luminances (this is the accumulation buffer)
    uint32_t* rng_save = nullptr;  // per lane: RNG state after the pixel's last sample of a pass (progressive passes resume from it)
    // next-event estimation mode (B200PT_FLAG_NEE) only
    f4* rad = nullptr;     // per lane: light-sampled radiance of the sample in flight .xyz
    f4* sh_o = nullptr;    // shadow requests of the current iteration: origin.xyz, tmax
    f4* sh_d = nullptr;    //   direction.xyz, lane id
    f4* sh_c = nullptr;    //   contribution if unoccluded .xyz, flags (1 = the sample ends with this request)
    f4* sh_f = nullptr;    //   last term of the ending sample .xyz
    uint32_t* sh_counts = nullptr;  // kBatch request counts
    uint32_t* sh_cursor = nullptr;  // kBatch fetch cursors of k_shadow
    // parked walks (pt_walk_warp.cuh SUSPEND): per lane kSuspWords 16-byte words — hit so far | node group,
    // triangle group | sp, stack top | the (at most kSuspStack) stack entries below the top, two per word
    u4* susp = nullptr;
    uint32_t* pending = nullptr;  // kBatch counts of the lanes k_shade passed through because their walk is parked
    uint32_t* counts = nullptr;   // kBatch + 1 queue lengths
    uint32_t* cursor = nullptr;   // kBatch ray-fetch cursors of the persistent extend kernel
    unsigned long long* walk_counters = nullptr;  // nodes, tris (FLAG_COUNTERS)
    uint32_t capacity = 0;
};

constexpr int kSuspWords = 3 + kSuspStack / 2;  // 176 bytes per lane: hit, groups, top + kSuspStack stack entries of 8 bytes
constexpr int kMaxParts = 4;  // interleaved sub-frames walked on separate streams (api.cu: render_local)
constexpr int kBatch = 32;  // wavefront iterations between host checks (even: the state sets alternate with k)

// create_mesh on the device (mesh_dedup.cu)
size_t dedup_scan_bytes(uint32_t n);
cudaError_t dedup_mesh_device(const int32_t* cv, const int32_t* cn, const int32_t* ct, uint32_t n, int32_t lo, uint32_t range,
                              const float* av, const float* an, const float* at, uint32_t* first, uint32_t* is_first, uint32_t* rank,
                              void* scratch, size_t scratch_bytes, int32_t* out_idx, float* ov, float* on, float* ot,
                              uint32_t* n_unique, cudaStream_t st);

// launch wrappers (wavefront.cu)
void launch_init_lanes(const SceneView& sv, const LaneArrays& la, int part, int parts, uint32_t n_part, cudaStream_t st);
void launch_extend(const SceneView& sv, const LaneArrays& la, int k, uint32_t max_count, bool count, cudaStream_t st);
void launch_shade(const SceneView& sv, const LaneArrays& la, int k, uint32_t max_count, cudaStream_t st);
bool timeline_arm(bool on, cudaStream_t st);   // PT_TIMELINE builds only (false otherwise)
bool timeline_dump(const char* path);
void launch_tri_records(const f4* pos, const f4* nrm, const int32_t* idx4, uint32_t n, TriShade* out, cudaStream_t st);
void launch_shadow(const SceneView& sv, const LaneArrays& la, int k, uint32_t max_count, cudaStream_t st);
void launch_resolve(const SceneView& sv, const f4* sum, float inv_spp, int transfer, uint32_t* rgba, f4* mean, cudaStream_t st);
unsigned pass_stats_blocks(uint32_t n_lanes);
void launch_pass_stats(const f4* sum, uint32_t n_lanes, int n_samples, double* partial, cudaStream_t st);
void launch_scatter_rows(const SceneView& sv, const f4* sum, f4* frame, cudaStream_t st);
void launch_fold_streams(const f4* sum, f4* out, uint32_t n_pix, int streams, cudaStream_t st);
void launch_primary(const SceneView& sv, f4* hits, uint32_t* cursor, cudaStream_t st);
void launch_trace(const SceneView& sv, const float* o, const float* d, int n, float tmin, float tmax, f4* hits,
                  uint32_t* cursor, cudaStream_t st);
void set_persistent_blocks(int sm_count);
void launch_sample_disney(const Material* mats, int n, const int32_t* material, const float* wo, const int32_t* lobe_in,
                          const uint32_t* rng_in, float* f, float* wi, float* pdf, int32_t* lobe_out, uint32_t* rng_out,
                          cudaStream_t st);

}  // namespace pt

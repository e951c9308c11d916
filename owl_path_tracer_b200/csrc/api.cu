// api.cu — implementation of the C ABI declared in include/b200pt.h.
// No CPU fallback: without a usable CUDA device b200pt_create fails.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <thread>

#include "device_ctx.h"
#include "pt_lights.h"

using namespace pt;

static thread_local std::string g_error;
extern "C" const char* b200pt_last_error(void) { return g_error.c_str(); }
extern "C" int b200pt_version(void) { return B200PT_VERSION; }

static int fail(int code, const std::string& msg) {
    g_error = msg;
    return code;
}
#define CU(x)                                                                                        \
    do {                                                                                             \
        cudaError_t e_ = (x);                                                                        \
        if (e_ != cudaSuccess) return fail(B200PT_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// Development knobs (environment variables read inside b200pt_render: parts override, L2 persistence window,
// per-iteration log, per-warp timeline) exist only in builds made with -DB200PT_DEV=1 (tools/build_variants.sh).
// The product build reads no environment variable on the render path.
#ifndef B200PT_DEV
#define B200PT_DEV 0
#endif

struct b200pt_ctx {
    b200pt_config cfg{};
    ncclComm_t comm = nullptr;   // b200pt_comm_init_rank / b200pt_comm_init_all
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // parts of the frame walked concurrently (render_local); part 0 runs on `stream`
    cudaStream_t part_stream[kMaxParts] = {};
    cudaEvent_t part_done[kMaxParts] = {};
    LaneArrays part[kMaxParts];
    int n_parts = 0;              // parts the lane arrays are currently laid out for
    int lanes_width = 0;          // frame width they were sized for
    uint32_t* h_counts = nullptr; // pinned: kMaxParts x (kBatch + 1 queue lengths, kBatch shadow counts, kBatch parked counts)
    uint32_t* d_cursor = nullptr;  // ray-fetch cursor of the parity-hook kernels
    // host staging of the flattened scene (kept: parity hooks map triangle ids back)
    std::vector<f4> h_pos, h_nrm;
    std::vector<f2> h_uv;
    std::vector<int32_t> h_idx4;
    std::vector<MeshInfo> h_meshes;
    std::vector<uint32_t> h_texels;
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f}, max_abs = 0;
    bool any_uv = false;
    // device scene
    f4 *d_pos = nullptr, *d_nrm = nullptr;
    f2* d_uv = nullptr;
    int32_t* d_idx4 = nullptr;
    MeshInfo* d_meshes = nullptr;
    uint32_t *d_texels = nullptr, *d_env = nullptr;
    float *d_env_rows = nullptr, *d_env_cols = nullptr;  // NEE mode: luminance cdfs of the map (pt_lights.h)
    f4* d_env_f32 = nullptr;    // float RGB map (b200pt_set_environment_f32) instead of d_env
    Material* d_mats = nullptr;
    int n_mats = 0;
    std::vector<Material> h_mats;      // host copy: the per-mesh material array is rebuilt from it
    Material* d_mesh_mats = nullptr;   // materials[meshes[m].material] for every mesh (SceneView::mesh_materials)
    size_t mesh_mats_capacity = 0;
    bool mesh_mats_dirty = true;
    // light table of the NEE mode (rebuilt with the per-mesh materials: emission is a material attribute)
    std::vector<LightTri> h_lights;
    std::vector<float> h_light_cdf;
    LightTri* d_lights = nullptr;
    float* d_light_cdf = nullptr;
    size_t lights_capacity = 0;
    DeviceBVH bvh;
    bool built = false;
    // b200pt_update_mesh / b200pt_refit_accel
    std::vector<int32_t> h_mesh_vbase, h_mesh_nv;  // vertex range of every mesh in the flattened arrays
    bool geometry_moved = false;                   // vertices changed since the tree was built / refitted
    float* d_node_box = nullptr;                   // refit scratch, 6 floats per wide node
    TriShade* d_tri_rec = nullptr;                 // per-triangle shading records (SceneView::tri_rec)
    // settings
    int env_w = 0, env_h = 0, env_use = 0, env_auto = 0;
    float env_color[3] = {0, 0, 0}, env_intensity = 0;
    float cam[12] = {0};
    int width = 0, height = 0, max_samples = 0, max_depth = 0;
    // lanes + outputs
    LaneArrays la;
    f4* d_frame = nullptr;      // full-frame sum buffer when lanes != pixels (ROWS sharding, several streams per pixel)
    bool use_frame = false;     // the last render left its sums in d_frame
    uint32_t* d_rgba = nullptr;
    f4* d_mean = nullptr;
    uint32_t* h_rgba = nullptr;  // pinned
    float* h_mean = nullptr;     // pinned
    size_t fb_pixels = 0;
    bool rendered = false, resolved = false;
    bool mean_on_host = false;   // h_mean holds the float frame of the last resolve
    int parts_override = 0;      // > 0: sub-frames per frame chosen by the frame pipeline (choose_parts)
    // b200pt_set_drain_callback
    float drain_fraction = 0.0f;
    b200pt_drain_fn drain_fn = nullptr;
    void* drain_user = nullptr;
    // progressive passes (b200pt_render_pass): samples of every pixel already in `sum`; 0 = no frame in progress
    int l2_persist_mode = 0;    // B200PT_L2_PERSIST experiment (render_local)
    int pass_done = 0;
    int resolve_samples = 0;    // divisor of the next resolve (max_samples for a complete frame)
    int transfer = 0;           // b200pt_set_output
    double* d_pass_partial = nullptr;
    std::vector<double> h_pass_partial;
    b200pt_stats stats{};
    std::vector<cudaEvent_t> ev_pool;
};

constexpr int kCountsStride = 3 * kBatch + 1;  // per part in h_counts

static void free_lanes(b200pt_ctx* c) {
    for (int p = 0; p < kMaxParts; ++p) {
        LaneArrays& L = c->part[p];
        for (int b = 0; b < 2; ++b) {
            cudaFree(L.ray_o[b]); cudaFree(L.ray_d[b]); cudaFree(L.thr[b]); cudaFree(L.queue[b]);
        }
        cudaFree(L.hit);
        cudaFree(L.sh_o); cudaFree(L.sh_d); cudaFree(L.sh_c); cudaFree(L.sh_f);
        cudaFree(L.sh_counts); cudaFree(L.sh_cursor);
        cudaFree(L.counts); cudaFree(L.cursor); cudaFree(L.pending);
        L = LaneArrays{};  // sum, rad, walk_counters are aliases of c->la's
    }
    cudaFree(c->la.sum); cudaFree(c->la.rad); cudaFree(c->la.walk_counters); cudaFree(c->la.susp); cudaFree(c->la.rng_save);
    cudaFree(c->d_frame); cudaFree(c->d_rgba); cudaFree(c->d_mean); cudaFree(c->d_pass_partial);
    c->d_pass_partial = nullptr;
    c->pass_done = 0;
    if (c->h_rgba) cudaFreeHost(c->h_rgba);
    if (c->h_mean) cudaFreeHost(c->h_mean);
    c->la = LaneArrays{};
    c->n_parts = 0;
    c->d_frame = nullptr; c->d_rgba = nullptr; c->d_mean = nullptr; c->h_rgba = nullptr; c->h_mean = nullptr;
    c->fb_pixels = 0;
}
static void free_scene(b200pt_ctx* c) {
    cudaFree(c->d_pos); cudaFree(c->d_nrm); cudaFree(c->d_uv); cudaFree(c->d_idx4); cudaFree(c->d_meshes);
    cudaFree(c->d_texels); cudaFree(c->bvh.nodes); cudaFree(c->bvh.tris); cudaFree(c->d_node_box); cudaFree(c->d_tri_rec);
    c->d_node_box = nullptr;
    c->d_tri_rec = nullptr;
    c->d_pos = c->d_nrm = nullptr; c->d_uv = nullptr; c->d_idx4 = nullptr; c->d_meshes = nullptr; c->d_texels = nullptr;
    c->bvh = DeviceBVH{};
    c->built = false;
}

// streams per pixel on this rank (b200pt_config.streams_per_pixel): SPP sharding only, and never more streams
// in the whole job than samples
static int rank_streams(const b200pt_ctx* c) {
    int v = c->cfg.streams_per_pixel;
    if (v <= 1 || (c->cfg.world_size > 1 && c->cfg.shard_mode == B200PT_SHARD_ROWS)) return 1;
    const int world = std::max(c->cfg.world_size, 1);
    while (v > 1 && (long long)v * world > (long long)c->max_samples) --v;
    return v;
}
static SceneView make_view(const b200pt_ctx* c, bool full_frame) {
    SceneView sv{};
    sv.nodes = c->bvh.nodes;
    sv.tris = c->bvh.tris;
    sv.idx4 = c->d_idx4;
    sv.vpos = c->d_pos;
    sv.vnrm = c->d_nrm;
    sv.vuv = c->d_uv;
    sv.tri_rec = c->d_tri_rec;
    sv.meshes = c->d_meshes;
    sv.materials = c->d_mats;
    sv.mesh_materials = c->d_mesh_mats;
    sv.lights = c->d_lights;
    sv.light_cdf = c->d_light_cdf;
    sv.n_lights = (c->cfg.flags & B200PT_FLAG_NEE) ? (int32_t)c->h_lights.size() : 0;
    sv.nee = (c->cfg.flags & B200PT_FLAG_NEE) ? 1 : 0;
    sv.env_cdf_rows = c->d_env_rows;
    sv.env_cdf_cols = c->d_env_cols;
    sv.env_nee = (sv.nee && c->env_use && (c->d_env || c->d_env_f32) && c->d_env_rows) ? 1 : 0;
    sv.texels = c->d_texels;
    sv.env = c->d_env;
    sv.env_f32 = c->d_env_f32;
    sv.env_w = c->env_w;
    sv.env_h = c->env_h;
    sv.env_use = c->env_use;
    sv.env_auto = c->env_auto;
    memcpy(sv.env_color, c->env_color, 12);
    sv.env_intensity = c->env_intensity;
    memcpy(sv.cam_origin, c->cam, 12);
    memcpy(sv.cam_llc, c->cam + 3, 12);
    memcpy(sv.cam_horizontal, c->cam + 6, 12);
    memcpy(sv.cam_vertical, c->cam + 9, 12);
    sv.width = c->width;
    sv.height = c->height;
    sv.max_samples = c->max_samples;
    sv.max_depth = c->max_depth;
    sv.n_materials = c->n_mats;
    sv.row_first = 0;
    sv.row_step = 1;
    sv.seed_row_offset = 0;
    sv.n_lanes = c->width * c->height;
    sv.plane_magic = 0x3F800000u;
    if (!full_frame && c->cfg.world_size > 1) {
        if (c->cfg.shard_mode == B200PT_SHARD_ROWS) {
            sv.row_first = c->cfg.rank;
            sv.row_step = c->cfg.world_size;
            const int rows = c->height > c->cfg.rank ? (c->height - c->cfg.rank + c->cfg.world_size - 1) / c->cfg.world_size : 0;
            sv.n_lanes = c->width * rows;
        }
    }
    if (!full_frame && !(c->cfg.world_size > 1 && c->cfg.shard_mode == B200PT_SHARD_ROWS)) {
        // SPP sharding: the job's samples are split over world_size * V streams, this rank owns V
        const int V = rank_streams(c), world = std::max(c->cfg.world_size, 1), G = world * V, g0 = c->cfg.rank * V;
        const int rem = c->max_samples % G;
        sv.max_samples = c->max_samples / G;
        sv.stream_extra = std::max(0, std::min(V, rem - g0));
        sv.seed_row_offset = g0 * c->height;
        if (V > 1) {
            sv.n_pix_lanes = sv.n_lanes;
            sv.n_lanes *= V;
        }
    }
    return sv;
}

static void comm_release(b200pt_ctx* c);

extern "C" int b200pt_destroy(b200pt_ctx* c);

extern "C" int b200pt_create(const b200pt_config* cfg, b200pt_ctx** out) {
    if (!cfg || !out) return fail(B200PT_ERR_INVALID, "b200pt_create: null argument");
    if (cfg->world_size < 1 || cfg->rank < 0 || cfg->rank >= cfg->world_size)
        return fail(B200PT_ERR_INVALID, "b200pt_create: bad rank/world_size");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(B200PT_ERR_CUDA, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0") +
                                         " (this library has no CPU path)");
    if (cfg->device_ordinal < 0 || cfg->device_ordinal >= ndev) return fail(B200PT_ERR_INVALID, "bad device ordinal");
    CU(cudaSetDevice(cfg->device_ordinal));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cfg->device_ordinal));
    if (prop.major != 10)
        return fail(B200PT_ERR_CUDA, "device is sm_" + std::to_string(prop.major * 10 + prop.minor) + "; this library is built for sm_100a only");
    b200pt_ctx* c = new b200pt_ctx();
    c->cfg = *cfg;
    const int st = [&]() -> int {   // any failure below releases what was created so far (b200pt_destroy copes with nulls)
        CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        CU(cudaEventCreate(&c->ev0));
        CU(cudaEventCreate(&c->ev1));
        CU(cudaMalloc(&c->d_cursor, sizeof(uint32_t)));
        c->part_stream[0] = c->stream;
        for (int p = 1; p < kMaxParts; ++p) CU(cudaStreamCreateWithFlags(&c->part_stream[p], cudaStreamNonBlocking));
        for (int p = 0; p < kMaxParts; ++p) CU(cudaEventCreateWithFlags(&c->part_done[p], cudaEventDisableTiming));
        CU(cudaMallocHost(&c->h_counts, sizeof(uint32_t) * kMaxParts * kCountsStride));
        return B200PT_OK;
    }();
    if (st != B200PT_OK) {
        const std::string msg = g_error;
        b200pt_destroy(c);
        return fail(st, msg);
    }
    set_persistent_blocks(prop.multiProcessorCount);
    *out = c;
    return B200PT_OK;
}

extern "C" int b200pt_destroy(b200pt_ctx* c) {
    if (!c) return B200PT_OK;
    cudaSetDevice(c->cfg.device_ordinal);
    cudaDeviceSynchronize();
    comm_release(c);
    free_lanes(c);
    free_scene(c);
    cudaFree(c->d_env);
    cudaFree(c->d_env_f32);
    cudaFree(c->d_env_rows);
    cudaFree(c->d_env_cols);
    cudaFree(c->d_mats);
    cudaFree(c->d_mesh_mats);
    cudaFree(c->d_lights);
    cudaFree(c->d_light_cdf);
    cudaFree(c->d_cursor);
    for (cudaEvent_t e : c->ev_pool) cudaEventDestroy(e);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    for (int p = 1; p < kMaxParts; ++p)
        if (c->part_stream[p]) cudaStreamDestroy(c->part_stream[p]);
    for (int p = 0; p < kMaxParts; ++p)
        if (c->part_done[p]) cudaEventDestroy(c->part_done[p]);
    if (c->h_counts) cudaFreeHost(c->h_counts);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return B200PT_OK;
}

extern "C" int b200pt_add_mesh(b200pt_ctx* c, const float* verts, const float* normals, const float* uvs, int32_t nv,
                               const int32_t* idx, int32_t nt, int32_t material_index, const uint8_t* tex_rgba,
                               int32_t tex_w, int32_t tex_h, int32_t* mesh_index) {
    if (!c || !verts || !normals || !idx || nv <= 0 || nt <= 0) return fail(B200PT_ERR_INVALID, "add_mesh: null/empty input (vertices, normals and indices are required)");
    if (material_index < 0) return fail(B200PT_ERR_INVALID, "add_mesh: negative material index");
    if (tex_rgba && (!uvs || tex_w <= 0 || tex_h <= 0)) return fail(B200PT_ERR_INVALID, "add_mesh: texture needs uvs and a size");
    for (size_t i = 0; i < 3 * (size_t)nt; ++i)
        if (idx[i] < 0 || idx[i] >= nv) return fail(B200PT_ERR_BOUNDS, "add_mesh: vertex index out of range");
    const int32_t vbase = (int32_t)c->h_pos.size();
    const int32_t tbase = (int32_t)(c->h_idx4.size() / 4);
    const int32_t mesh = (int32_t)c->h_meshes.size();
    for (int i = 0; i < nv; ++i) {
        const float x = verts[3 * i], y = verts[3 * i + 1], z = verts[3 * i + 2];
        c->h_pos.push_back(f4{x, y, z, 0.0f});
        c->h_nrm.push_back(f4{normals[3 * i], normals[3 * i + 1], normals[3 * i + 2], 0.0f});
        c->h_uv.push_back(uvs ? f2{uvs[2 * i], uvs[2 * i + 1]} : f2{0.0f, 0.0f});
        const float p[3] = {x, y, z};
        for (int a = 0; a < 3; ++a) {
            c->lo[a] = fminf(c->lo[a], p[a]);
            c->hi[a] = fmaxf(c->hi[a], p[a]);
            c->max_abs = fmaxf(c->max_abs, fabsf(p[a]));
        }
    }
    for (int t = 0; t < nt; ++t) {
        c->h_idx4.push_back(vbase + idx[3 * t]);
        c->h_idx4.push_back(vbase + idx[3 * t + 1]);
        c->h_idx4.push_back(vbase + idx[3 * t + 2]);
        c->h_idx4.push_back(mesh);
    }
    MeshInfo mi{};
    mi.first_tri = tbase;
    mi.material = material_index;
    mi.tex_offset = -1;
    mi.light_first = -1;
    mi.has_uv = uvs ? 1 : 0;
    if (tex_rgba) {
        mi.tex_offset = (int32_t)c->h_texels.size();
        mi.tex_w = tex_w;
        mi.tex_h = tex_h;
        const uint32_t* px = reinterpret_cast<const uint32_t*>(tex_rgba);
        c->h_texels.insert(c->h_texels.end(), px, px + (size_t)tex_w * tex_h);
    }
    c->h_meshes.push_back(mi);
    c->h_mesh_vbase.push_back(vbase);
    c->h_mesh_nv.push_back(nv);
    c->built = false;
    c->mesh_mats_dirty = true;
    if (mesh_index) *mesh_index = mesh;
    return B200PT_OK;
}

extern "C" int b200pt_build_accel(b200pt_ctx* c) {
    if (!c) return fail(B200PT_ERR_INVALID, "build_accel: null context");
    if (c->h_meshes.empty()) return fail(B200PT_ERR_EMPTY, "no geometries");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    c->pass_done = 0;
    free_scene(c);
    const size_t nv = c->h_pos.size(), nt = c->h_idx4.size() / 4;
    CU(cudaMalloc(&c->d_pos, sizeof(f4) * nv));
    CU(cudaMalloc(&c->d_nrm, sizeof(f4) * nv));
    CU(cudaMalloc(&c->d_uv, sizeof(f2) * nv));
    CU(cudaMalloc(&c->d_idx4, sizeof(int32_t) * 4 * nt));
    CU(cudaMalloc(&c->d_meshes, sizeof(MeshInfo) * c->h_meshes.size()));
    CU(cudaMemcpyAsync(c->d_pos, c->h_pos.data(), sizeof(f4) * nv, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_nrm, c->h_nrm.data(), sizeof(f4) * nv, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_uv, c->h_uv.data(), sizeof(f2) * nv, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_idx4, c->h_idx4.data(), sizeof(int32_t) * 4 * nt, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_meshes, c->h_meshes.data(), sizeof(MeshInfo) * c->h_meshes.size(), cudaMemcpyHostToDevice, c->stream));
    if (!c->h_texels.empty()) {
        CU(cudaMalloc(&c->d_texels, sizeof(uint32_t) * c->h_texels.size()));
        CU(cudaMemcpyAsync(c->d_texels, c->h_texels.data(), sizeof(uint32_t) * c->h_texels.size(), cudaMemcpyHostToDevice, c->stream));
    }
    CU(cudaMalloc(&c->d_tri_rec, sizeof(TriShade) * nt));
    launch_tri_records(c->d_pos, c->d_nrm, c->d_idx4, (uint32_t)nt, c->d_tri_rec, c->stream);
    CU(cudaEventRecord(c->ev0, c->stream));
    std::string err;
    if (!build_bvh_device(c->d_pos, c->d_idx4, (uint32_t)nt, c->lo, c->hi, c->max_abs, &c->bvh, c->stream, err))
        return fail(B200PT_ERR_CUDA, "bvh build: " + err);
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->stats.build_ms = ms;
    c->stats.bvh_nodes = c->bvh.n_nodes;
    c->stats.bvh_tris = (uint32_t)nt;
    c->stats.bvh_bytes = (uint64_t)c->bvh.n_nodes * sizeof(Node8) + (uint64_t)c->bvh.n_tris * sizeof(TriPair);
    c->built = true;
    c->geometry_moved = false;
    return B200PT_OK;
}

// ---- moving geometry: new vertex positions / normals for a mesh, then a refit of the tree
static void scene_bounds(b200pt_ctx* c) {
    for (int a = 0; a < 3; ++a) {
        c->lo[a] = 3e38f;
        c->hi[a] = -3e38f;
    }
    c->max_abs = 0;
    for (const f4& p : c->h_pos) {
        const float v[3] = {p.x, p.y, p.z};
        for (int a = 0; a < 3; ++a) {
            c->lo[a] = fminf(c->lo[a], v[a]);
            c->hi[a] = fmaxf(c->hi[a], v[a]);
            c->max_abs = fmaxf(c->max_abs, fabsf(v[a]));
        }
    }
}

extern "C" int b200pt_update_mesh(b200pt_ctx* c, int32_t mesh_index, const float* verts, const float* normals, int32_t nv) {
    if (!c || !verts) return fail(B200PT_ERR_INVALID, "update_mesh: null argument");
    if (mesh_index < 0 || mesh_index >= (int32_t)c->h_meshes.size()) return fail(B200PT_ERR_BOUNDS, "update_mesh: no such mesh");
    const int32_t vbase = c->h_mesh_vbase[(size_t)mesh_index];
    if (nv != c->h_mesh_nv[(size_t)mesh_index])
        return fail(B200PT_ERR_INVALID, "update_mesh: the mesh has " + std::to_string(c->h_mesh_nv[(size_t)mesh_index]) +
                                            " vertices (topology is fixed; rebuild the scene to change it)");
    for (int i = 0; i < nv; ++i) {
        c->h_pos[(size_t)vbase + i] = f4{verts[3 * i], verts[3 * i + 1], verts[3 * i + 2], 0.0f};
        if (normals) c->h_nrm[(size_t)vbase + i] = f4{normals[3 * i], normals[3 * i + 1], normals[3 * i + 2], 0.0f};
    }
    scene_bounds(c);
    c->pass_done = 0;
    c->mesh_mats_dirty = true;  // NEE mode: the light table holds positions and areas
    if (c->built) {             // before the first build the staging arrays are all there is
        CU(cudaSetDevice(c->cfg.device_ordinal));
        CU(cudaMemcpyAsync(c->d_pos + vbase, c->h_pos.data() + vbase, sizeof(f4) * (size_t)nv, cudaMemcpyHostToDevice, c->stream));
        if (normals)
            CU(cudaMemcpyAsync(c->d_nrm + vbase, c->h_nrm.data() + vbase, sizeof(f4) * (size_t)nv, cudaMemcpyHostToDevice, c->stream));
        launch_tri_records(c->d_pos, c->d_nrm, c->d_idx4, (uint32_t)(c->h_idx4.size() / 4), c->d_tri_rec, c->stream);
        CU(cudaStreamSynchronize(c->stream));
        c->geometry_moved = true;
    }
    return B200PT_OK;
}

extern "C" int b200pt_refit_accel(b200pt_ctx* c) {
    if (!c) return fail(B200PT_ERR_INVALID, "refit_accel: null context");
    if (!c->built) return fail(B200PT_ERR_INVALID, "refit_accel: call b200pt_build_accel first");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    if (!c->d_node_box) CU(cudaMalloc(&c->d_node_box, sizeof(float) * 6 * (size_t)c->bvh.n_nodes));
    CU(cudaEventRecord(c->ev0, c->stream));
    std::string err;
    if (!refit_bvh_device(c->d_pos, c->d_idx4, c->max_abs, &c->bvh, c->d_node_box, c->stream, err))
        return fail(B200PT_ERR_CUDA, "bvh refit: " + err);
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->stats.refit_ms = ms;
    c->geometry_moved = false;
    c->pass_done = 0;
    return B200PT_OK;
}

extern "C" int b200pt_set_materials(b200pt_ctx* c, const b200pt_material* m, int32_t nm) {
    if (!c || !m || nm <= 0) return fail(B200PT_ERR_INVALID, "set_materials: null/empty");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    c->pass_done = 0;
    for (const MeshInfo& mi : c->h_meshes)
        if (mi.material >= nm) return fail(B200PT_ERR_BOUNDS, "set_materials: a mesh refers to material " + std::to_string(mi.material));
    std::vector<Material> tmp((size_t)nm);
    for (int i = 0; i < nm; ++i) {
        memset(&tmp[i], 0, sizeof(Material));
        memcpy(&tmp[i], &m[i], sizeof(b200pt_material));
    }
    if (nm != c->n_mats) {
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_mats);
        c->d_mats = nullptr;
        CU(cudaMalloc(&c->d_mats, sizeof(Material) * (size_t)nm));
        c->n_mats = nm;
    }
    CU(cudaMemcpyAsync(c->d_mats, tmp.data(), sizeof(Material) * (size_t)nm, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));  // tmp goes out of scope
    c->h_mats.swap(tmp);
    c->mesh_mats_dirty = true;
    return B200PT_OK;
}

// shared by the two environment setters: exactly one of rgba / rgb is non-null (or none: no map)
static int set_environment_common(b200pt_ctx* c, const uint8_t* rgba, const float* rgb, int32_t w, int32_t h, int32_t use,
                                  int32_t auto_, const float color[3], float intensity) {
    if (c) c->pass_done = 0;
    if (!c || !color) return fail(B200PT_ERR_INVALID, "set_environment: null");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    CU(cudaDeviceSynchronize());
    cudaFree(c->d_env);
    cudaFree(c->d_env_f32);
    cudaFree(c->d_env_rows);
    cudaFree(c->d_env_cols);
    c->d_env = nullptr;
    c->d_env_f32 = nullptr;
    c->d_env_rows = c->d_env_cols = nullptr;
    c->env_w = c->env_h = 0;
    if ((rgba || rgb) && w > 0 && h > 0) {
        const size_t n = (size_t)w * h;
        std::vector<float> rows, cols;
        bool tables = false;
        const bool want_tables = (c->cfg.flags & B200PT_FLAG_NEE) && use;  // importance sampling of the map
        if (rgba) {
            CU(cudaMalloc(&c->d_env, n * 4));
            CU(cudaMemcpy(c->d_env, rgba, n * 4, cudaMemcpyHostToDevice));
            if (want_tables) {
                std::vector<uint32_t> texels(n);
                memcpy(texels.data(), rgba, n * 4);
                tables = build_env_table(texels.data(), w, h, rows, cols);
            }
        } else {
            std::vector<f4> texels(n);
            for (size_t i = 0; i < n; ++i) texels[i] = f4{rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], 0.0f};
            CU(cudaMalloc(&c->d_env_f32, n * sizeof(f4)));
            CU(cudaMemcpy(c->d_env_f32, texels.data(), n * sizeof(f4), cudaMemcpyHostToDevice));
            if (want_tables) tables = build_env_table_f32(texels.data(), w, h, rows, cols);
        }
        c->env_w = w;
        c->env_h = h;
        if (tables) {
            CU(cudaMalloc(&c->d_env_rows, sizeof(float) * rows.size()));
            CU(cudaMalloc(&c->d_env_cols, sizeof(float) * cols.size()));
            CU(cudaMemcpy(c->d_env_rows, rows.data(), sizeof(float) * rows.size(), cudaMemcpyHostToDevice));
            CU(cudaMemcpy(c->d_env_cols, cols.data(), sizeof(float) * cols.size(), cudaMemcpyHostToDevice));
        }
    }
    c->env_use = use != 0;
    c->env_auto = auto_ != 0;
    memcpy(c->env_color, color, 12);
    c->env_intensity = intensity;
    return B200PT_OK;
}
extern "C" int b200pt_set_environment(b200pt_ctx* c, const uint8_t* rgba, int32_t w, int32_t h, int32_t use, int32_t auto_,
                                      const float color[3], float intensity) {
    return set_environment_common(c, rgba, nullptr, w, h, use, auto_, color, intensity);
}
extern "C" int b200pt_set_environment_f32(b200pt_ctx* c, const float* rgb, int32_t w, int32_t h, int32_t use, int32_t auto_,
                                          const float color[3], float intensity) {
    return set_environment_common(c, nullptr, rgb, w, h, use, auto_, color, intensity);
}

extern "C" int b200pt_set_camera(b200pt_ctx* c, const float origin[3], const float llc[3], const float horizontal[3],
                                 const float vertical[3]) {
    if (!c || !origin || !llc || !horizontal || !vertical) return fail(B200PT_ERR_INVALID, "set_camera: null");
    c->pass_done = 0;
    memcpy(c->cam, origin, 12);
    memcpy(c->cam + 3, llc, 12);
    memcpy(c->cam + 6, horizontal, 12);
    memcpy(c->cam + 9, vertical, 12);
    return B200PT_OK;
}

extern "C" int b200pt_set_render(b200pt_ctx* c, int32_t width, int32_t height, int32_t max_samples, int32_t max_path_depth) {
    if (!c || width <= 0 || height <= 0 || max_samples <= 0 || max_path_depth < 0)
        return fail(B200PT_ERR_INVALID, "set_render: bad size / sample count");
    if (max_samples >= (1 << 20) || max_path_depth > 255)
        return fail(B200PT_ERR_INVALID, "set_render: max_samples < 2^20 and max_path_depth <= 255 supported");
    if ((int64_t)width * height > (int64_t)1 << 30) return fail(B200PT_ERR_INVALID, "set_render: frame too large");
    // a frame in progress (b200pt_render_pass) survives a change of the sample budget only
    if (width != c->width || height != c->height || max_path_depth != c->max_depth) c->pass_done = 0;
    c->width = width;
    c->height = height;
    c->max_samples = max_samples;
    c->max_depth = max_path_depth;
    c->rendered = c->resolved = false;
    return B200PT_OK;
}

// ---------------------------------------------------------------------------- render
// per-mesh copy of the material each mesh is bound to: the shade kernel then needs the mesh index
// only (one dependent load less than entity -> material_index -> material_buffer, device.cu:150-154)
static int sync_mesh_materials(b200pt_ctx* c) {
    if (!c->mesh_mats_dirty) return B200PT_OK;
    const size_t nm = c->h_meshes.size();
    std::vector<Material> tmp(nm);
    for (size_t i = 0; i < nm; ++i) {
        const int32_t mi = c->h_meshes[i].material;
        if (mi < 0 || mi >= (int32_t)c->h_mats.size()) return fail(B200PT_ERR_BOUNDS, "a mesh refers to material " + std::to_string(mi));
        tmp[i] = c->h_mats[(size_t)mi];
    }
    if (nm > c->mesh_mats_capacity) {
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_mesh_mats);
        c->d_mesh_mats = nullptr;
        CU(cudaMalloc(&c->d_mesh_mats, sizeof(Material) * nm));
        c->mesh_mats_capacity = nm;
    }
    CU(cudaMemcpyAsync(c->d_mesh_mats, tmp.data(), sizeof(Material) * nm, cudaMemcpyHostToDevice, c->stream));
    if (c->cfg.flags & B200PT_FLAG_NEE) {
        // emissive triangles, their selection probabilities, and light_first of every mesh
        build_light_table(c->h_pos, c->h_idx4, c->h_meshes, c->h_mats, c->h_lights, c->h_light_cdf);
        const size_t nl = c->h_lights.size();
        if (nl > c->lights_capacity) {
            CU(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_lights);
            cudaFree(c->d_light_cdf);
            c->d_lights = nullptr;
            c->d_light_cdf = nullptr;
            CU(cudaMalloc(&c->d_lights, sizeof(LightTri) * nl));
            CU(cudaMalloc(&c->d_light_cdf, sizeof(float) * nl));
            c->lights_capacity = nl;
        }
        if (nl) {
            CU(cudaMemcpyAsync(c->d_lights, c->h_lights.data(), sizeof(LightTri) * nl, cudaMemcpyHostToDevice, c->stream));
            CU(cudaMemcpyAsync(c->d_light_cdf, c->h_light_cdf.data(), sizeof(float) * nl, cudaMemcpyHostToDevice, c->stream));
        }
        if (c->d_meshes)
            CU(cudaMemcpyAsync(c->d_meshes, c->h_meshes.data(), sizeof(MeshInfo) * nm, cudaMemcpyHostToDevice, c->stream));
        c->stats.n_lights = (uint32_t)nl;
    }
    CU(cudaStreamSynchronize(c->stream));
    c->mesh_mats_dirty = false;
    return B200PT_OK;
}

// Parts: the rank's lanes are split into `parts` interleaved sub-frames (rows r = p mod parts), each with its
// own compact state, queues and stream.  Every launch of the persistent walker ends with a tail in which a few
// long rays keep a few warps busy (sm__cycles_active was 91 % of elapsed with one stream); with two independent
// sub-frames the other part's kernels fill the SMs the tail leaves idle.  B200PT_FLAG_TIMING keeps one part so
// that the per-kernel event times stay clean; B200PT_PARTS overrides (development).
static int choose_parts(const b200pt_ctx* c, const SceneView& sv) {
    int parts = (c->cfg.flags & B200PT_FLAG_TIMING) ? 1 : 2;
    if (sv.n_lanes < (1 << 18)) parts = 1;  // small frames: launch-bound already
    // a slot of a frame pipeline renders its frame as ONE part: the other slots' frames fill the launch tails, and
    // full-size launches beat half-size ones (dragon 1080p / 256 spp, 2 slots: 2175 -> 2285 Mrays/s, profiles/r02j_*)
    if (c->parts_override > 0) parts = c->parts_override;
#if B200PT_DEV
    if (const char* e = getenv("B200PT_PARTS")) parts = atoi(e);
#endif
    const int rows = sv.width > 0 ? sv.n_lanes / sv.width : 0;
    return std::max(1, std::min({parts, kMaxParts, std::max(rows, 1)}));
}
static uint32_t part_lanes(const SceneView& sv, int p, int parts) {
    const int rows = sv.width > 0 ? sv.n_lanes / sv.width : 0;
    const int mine = rows > p ? (rows - p + parts - 1) / parts : 0;
    return (uint32_t)mine * (uint32_t)sv.width;
}

static int ensure_lanes(b200pt_ctx* c, const SceneView& sv, int parts) {
    const size_t px = (size_t)c->width * c->height;
    const uint32_t need = (uint32_t)std::max(sv.n_lanes, 1);
    const bool need_frame = sv.n_pix_lanes > 0 || (c->cfg.world_size > 1 && c->cfg.shard_mode == B200PT_SHARD_ROWS);
    if (c->la.capacity < need || c->fb_pixels != px || c->n_parts != parts || c->lanes_width != c->width || (need_frame && !c->d_frame)) {
        CU(cudaDeviceSynchronize());
        free_lanes(c);
        const bool nee = (c->cfg.flags & B200PT_FLAG_NEE) != 0;
        CU(cudaMalloc(&c->la.sum, sizeof(f4) * need));
        CU(cudaMalloc(&c->la.rng_save, sizeof(uint32_t) * need));
        CU(cudaMalloc(&c->d_pass_partial, sizeof(double) * pass_stats_blocks(need)));
        if (nee) CU(cudaMalloc(&c->la.rad, sizeof(f4) * need));
        CU(cudaMalloc(&c->la.walk_counters, sizeof(unsigned long long) * 2));
        CU(cudaMalloc(&c->la.susp, sizeof(u4) * kSuspWords * (size_t)need));  // touched by parked walks only
        for (int p = 0; p < parts; ++p) {
            LaneArrays& L = c->part[p];
            // + 256: k_shade loads entry i before it knows the queue length
            const size_t padded = (size_t)(need + parts - 1) / parts + (size_t)c->width + 256;
            for (int b = 0; b < 2; ++b) {
                CU(cudaMalloc(&L.ray_o[b], sizeof(f4) * padded));
                CU(cudaMalloc(&L.ray_d[b], sizeof(f4) * padded));
                CU(cudaMalloc(&L.thr[b], sizeof(f4) * padded));
                CU(cudaMalloc(&L.queue[b], sizeof(uint32_t) * padded));
                CU(cudaMemsetAsync(L.queue[b], 0, sizeof(uint32_t) * padded, c->stream));
            }
            CU(cudaMalloc(&L.hit, sizeof(f4) * padded));
            if (nee) {
                CU(cudaMalloc(&L.sh_o, sizeof(f4) * padded));
                CU(cudaMalloc(&L.sh_d, sizeof(f4) * padded));
                CU(cudaMalloc(&L.sh_c, sizeof(f4) * padded));
                CU(cudaMalloc(&L.sh_f, sizeof(f4) * padded));
                CU(cudaMalloc(&L.sh_counts, sizeof(uint32_t) * kBatch));
                CU(cudaMalloc(&L.sh_cursor, sizeof(uint32_t) * kBatch));
            }
            CU(cudaMalloc(&L.counts, sizeof(uint32_t) * (kBatch + 1)));
            CU(cudaMalloc(&L.cursor, sizeof(uint32_t) * kBatch));
            CU(cudaMalloc(&L.pending, sizeof(uint32_t) * kBatch));
            L.susp = c->la.susp;
            L.sum = c->la.sum;
            L.rng_save = c->la.rng_save;
            L.rad = c->la.rad;
            L.walk_counters = c->la.walk_counters;
            L.capacity = (uint32_t)padded;
        }
        c->la.capacity = need;
        c->n_parts = parts;
        c->lanes_width = c->width;
        if (need_frame) CU(cudaMalloc(&c->d_frame, sizeof(f4) * px));
        CU(cudaMalloc(&c->d_rgba, sizeof(uint32_t) * px));
        CU(cudaMalloc(&c->d_mean, sizeof(f4) * px));
        CU(cudaMallocHost(&c->h_rgba, sizeof(uint32_t) * px));
        CU(cudaMallocHost(&c->h_mean, sizeof(float) * 4 * px));
        c->fb_pixels = px;
        CU(cudaStreamSynchronize(c->stream));
    }
    return B200PT_OK;
}

static int check_ready(b200pt_ctx* c, const char* who) {
    if (!c) return fail(B200PT_ERR_INVALID, std::string(who) + ": null context");
    if (!c->built) return fail(B200PT_ERR_INVALID, std::string(who) + ": call b200pt_build_accel first");
    if (!c->d_mats) return fail(B200PT_ERR_INVALID, std::string(who) + ": call b200pt_set_materials first");
    if (c->width <= 0) return fail(B200PT_ERR_INVALID, std::string(who) + ": call b200pt_set_render first");
    if (c->geometry_moved)
        return fail(B200PT_ERR_INVALID, std::string(who) + ": vertices moved (b200pt_update_mesh): call b200pt_refit_accel or b200pt_build_accel first");
    return B200PT_OK;
}

// Renders samples first_sample .. first_sample + n_pass - 1 of every pixel of this rank (n_pass == 0: the whole
// frame, the reference's launch).  first_sample > 0 continues the streams and sums the previous pass left.
static int render_local(b200pt_ctx* c, int first_sample = 0, int n_pass = 0) {
    CU(cudaSetDevice(c->cfg.device_ordinal));
    if (int r = sync_mesh_materials(c)) return r;
    SceneView sv = make_view(c, false);
    if (n_pass > 0) {
        sv.first_sample = first_sample;
        sv.max_samples = first_sample + n_pass;
    }
    const int parts = choose_parts(c, sv);
    if (int r = ensure_lanes(c, sv, parts)) return r;
#if B200PT_DEV
    {
        // Experiment (B200PT_L2_PERSIST = 1 tris | 2 nodes | 0 off): keep the triangle blocks (or the nodes) in the
        // persisting part of L2 so that the lane state streaming through L2 every iteration cannot evict them.
        const char* e = getenv("B200PT_L2_PERSIST");
        const int mode = e ? atoi(e) : 0;
        if (mode != c->l2_persist_mode) {
            cudaDeviceProp prop;
            CU(cudaGetDeviceProperties(&prop, c->cfg.device_ordinal));
            cudaStreamAttrValue attr{};
            if (mode > 0) {
                const void* base = mode == 2 ? (const void*)c->bvh.nodes : (const void*)c->bvh.tris;
                size_t bytes = mode == 2 ? sizeof(Node8) * (size_t)c->bvh.n_nodes : sizeof(TriPair) * (size_t)c->bvh.n_tris;
                bytes = std::min(bytes, (size_t)prop.accessPolicyMaxWindowSize);
                CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min(bytes, (size_t)prop.persistingL2CacheMaxSize)));
                attr.accessPolicyWindow.base_ptr = const_cast<void*>(base);
                attr.accessPolicyWindow.num_bytes = bytes;
                attr.accessPolicyWindow.hitRatio = 1.0f;
                attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            } else {
                attr.accessPolicyWindow.num_bytes = 0;
                CU(cudaCtxResetPersistingL2Cache());
            }
            for (int p = 0; p < kMaxParts; ++p) CU(cudaStreamSetAttribute(c->part_stream[p], cudaStreamAttributeAccessPolicyWindow, &attr));
            c->l2_persist_mode = mode;
        }
    }
#endif
    const bool count = (c->cfg.flags & B200PT_FLAG_COUNTERS) != 0;
    const bool timing = (c->cfg.flags & B200PT_FLAG_TIMING) != 0 && parts == 1;
    const bool nee = sv.nee != 0;
    const int ev_per_iter = nee ? 3 : 2;  // extend | shade [| shadow]
    cudaStream_t st = c->stream;
    b200pt_stats& S = c->stats;
    S.render_ms = S.extend_ms = S.shade_ms = S.other_ms = 0;
    S.rays = S.paths = S.nodes_visited = S.tris_tested = S.kernel_launches = S.iterations = 0;
    S.shadow_rays = 0;
    S.shadow_ms = 0;

    CU(cudaMemsetAsync(c->la.walk_counters, 0, sizeof(unsigned long long) * 2, st));
    CU(cudaEventRecord(c->ev0, st));  // the other parts' streams start behind this point
    uint32_t active[kMaxParts] = {};
    const bool any_work = sv.n_lanes > 0 && sv.max_samples + sv.stream_extra > sv.first_sample && sv.max_depth > 0;
    if (!any_work && sv.first_sample == 0) CU(cudaMemsetAsync(c->la.sum, 0, sizeof(f4) * c->la.capacity, st));
    for (int p = 1; p < parts; ++p) CU(cudaStreamWaitEvent(c->part_stream[p], c->ev0, 0));
    for (int p = 0; p < parts; ++p) {
        const LaneArrays& L = c->part[p];
        cudaStream_t ps = c->part_stream[p];
        CU(cudaMemsetAsync(L.counts, 0, sizeof(uint32_t) * (kBatch + 1), ps));
        CU(cudaMemsetAsync(L.cursor, 0, sizeof(uint32_t) * kBatch, ps));
        CU(cudaMemsetAsync(L.pending, 0, sizeof(uint32_t) * kBatch, ps));
        if (nee) {
            CU(cudaMemsetAsync(L.sh_counts, 0, sizeof(uint32_t) * kBatch, ps));
            CU(cudaMemsetAsync(L.sh_cursor, 0, sizeof(uint32_t) * kBatch, ps));
        }
        const uint32_t n_part = any_work ? part_lanes(sv, p, parts) : 0;
        if (n_part) {
            launch_init_lanes(sv, L, p, parts, n_part, ps);
            S.kernel_launches++;
            active[p] = n_part;
            c->h_counts[p * kCountsStride] = n_part;  // pinned: stays valid until the copy ran
            CU(cudaMemcpyAsync(L.counts, &c->h_counts[p * kCountsStride], sizeof(uint32_t), cudaMemcpyHostToDevice, ps));
        }
    }
    if (timing) {
        while (c->ev_pool.size() < (size_t)(3 * kBatch + 1)) {
            cudaEvent_t e;
            CU(cudaEventCreate(&e));
            c->ev_pool.push_back(e);
        }
    }
    FILE* iter_log = nullptr;
    long tl_iter = -1, iter_no = 0;
#if B200PT_DEV
    if (timing)
        if (const char* e = getenv("B200PT_ITER_LOG")) iter_log = fopen(e, "w");
    // PT_TIMELINE builds: per-warp timestamps of the k_extend launch of wavefront iteration B200PT_TIMELINE_ITER of
    // part 0, written to B200PT_TIMELINE_FILE after the frame
    if (const char* e = getenv("B200PT_TIMELINE_ITER")) tl_iter = atol(e);
#endif
    auto any_active = [&] {
        for (int p = 0; p < parts; ++p)
            if (active[p]) return true;
        return false;
    };
    bool drain_armed = c->drain_fn != nullptr && c->drain_fraction > 0.0f;
    while (any_active()) {
        // the parts' launches are interleaved so that neither stream runs ahead of the other in the queue
        for (int k = 0; k < kBatch; ++k) {
            for (int p = 0; p < parts; ++p) {
                if (!active[p]) continue;
                const LaneArrays& L = c->part[p];
                cudaStream_t ps = c->part_stream[p];
                if (timing) CU(cudaEventRecord(c->ev_pool[ev_per_iter * k], ps));
                if (p == 0 && iter_no + k == tl_iter && timeline_arm(true, ps)) {
                    launch_extend(sv, L, k, active[p], count, ps);
                    timeline_arm(false, ps);
                } else {
                    launch_extend(sv, L, k, active[p], count, ps);
                }
                if (timing) CU(cudaEventRecord(c->ev_pool[ev_per_iter * k + 1], ps));
                launch_shade(sv, L, k, active[p], ps);
                if (nee) {
                    // the shadow rays this iteration's vertices asked for: any-hit walk, then the gated add
                    if (timing) CU(cudaEventRecord(c->ev_pool[ev_per_iter * k + 2], ps));
                    launch_shadow(sv, L, k, active[p], ps);
                }
                S.kernel_launches += (uint64_t)ev_per_iter;
            }
        }
        iter_no += kBatch;
        if (timing) CU(cudaEventRecord(c->ev_pool[ev_per_iter * kBatch], st));
        CU(cudaGetLastError());
        for (int p = 0; p < parts; ++p) {
            if (!active[p]) continue;
            const LaneArrays& L = c->part[p];
            cudaStream_t ps = c->part_stream[p];
            uint32_t* hc = c->h_counts + p * kCountsStride;
            CU(cudaMemcpyAsync(hc, L.counts, sizeof(uint32_t) * (kBatch + 1), cudaMemcpyDeviceToHost, ps));
            CU(cudaMemcpyAsync(hc + 2 * kBatch + 1, L.pending, sizeof(uint32_t) * kBatch, cudaMemcpyDeviceToHost, ps));
            CU(cudaMemsetAsync(L.pending, 0, sizeof(uint32_t) * kBatch, ps));
            // next batch starts from the last queue length; the other slots go back to zero
            CU(cudaMemcpyAsync(L.counts, L.counts + kBatch, sizeof(uint32_t), cudaMemcpyDeviceToDevice, ps));
            CU(cudaMemsetAsync(L.counts + 1, 0, sizeof(uint32_t) * kBatch, ps));
            CU(cudaMemsetAsync(L.cursor, 0, sizeof(uint32_t) * kBatch, ps));
            if (nee) {
                CU(cudaMemcpyAsync(hc + kBatch + 1, L.sh_counts, sizeof(uint32_t) * kBatch, cudaMemcpyDeviceToHost, ps));
                CU(cudaMemsetAsync(L.sh_counts, 0, sizeof(uint32_t) * kBatch, ps));
                CU(cudaMemsetAsync(L.sh_cursor, 0, sizeof(uint32_t) * kBatch, ps));
            }
        }
        for (int p = 0; p < parts; ++p) {
            if (!active[p]) continue;
            CU(cudaStreamSynchronize(c->part_stream[p]));
            const uint32_t* hc = c->h_counts + p * kCountsStride;
            for (int k = 0; k < kBatch; ++k) {
                // a lane whose walk was parked in iteration k sits in queue k+1 with the same ray: count it once
                S.rays += hc[k] - hc[2 * kBatch + 1 + k];
                if (nee) S.shadow_rays += hc[kBatch + 1 + k];
                if (p == 0 && hc[k]) S.iterations++;
            }
            active[p] = hc[kBatch];
        }
        if (drain_armed) {   // b200pt_set_drain_callback: the frame's queue fell below the caller's fraction
            uint64_t in_flight = 0;
            for (int p = 0; p < parts; ++p) in_flight += active[p];
            if ((double)in_flight < (double)c->drain_fraction * (double)sv.n_lanes) {
                drain_armed = false;
                c->drain_fn(c->drain_user);
            }
        }
        if (timing) {
            for (int k = 0; k < kBatch; ++k) {
                float a = 0, b = 0, s = 0;
                cudaEventElapsedTime(&a, c->ev_pool[ev_per_iter * k], c->ev_pool[ev_per_iter * k + 1]);
                cudaEventElapsedTime(&b, c->ev_pool[ev_per_iter * k + 1], c->ev_pool[ev_per_iter * k + 2]);
                if (nee) cudaEventElapsedTime(&s, c->ev_pool[ev_per_iter * k + 2], c->ev_pool[ev_per_iter * k + 3]);
                S.extend_ms += a;
                S.shade_ms += b;
                S.shadow_ms += s;
                if (iter_log && c->h_counts[k])  // development: queue length, parked lanes, kernel times per iteration
                    fprintf(iter_log, "%u %u %.4f %.4f\n", c->h_counts[k], c->h_counts[2 * kBatch + 1 + k], a, b);
            }
        }
    }
    if (iter_log) fclose(iter_log);
#if B200PT_DEV
    if (tl_iter >= 0)
        if (const char* e = getenv("B200PT_TIMELINE_FILE")) {
            CU(cudaDeviceSynchronize());
            timeline_dump(e);
        }
#endif
    if (drain_armed) c->drain_fn(c->drain_user);   // a frame that never fell below the fraction before it ended
    for (int p = 1; p < parts; ++p) {  // the frame is complete on `stream` once every part is
        CU(cudaEventRecord(c->part_done[p], c->part_stream[p]));
        CU(cudaStreamWaitEvent(st, c->part_done[p], 0));
    }
    c->use_frame = false;
    if (sv.n_pix_lanes > 0) {  // several streams per pixel: add the streams' sums in stream order
        launch_fold_streams(c->la.sum, c->d_frame, (uint32_t)sv.n_pix_lanes, sv.n_lanes / sv.n_pix_lanes, st);
        S.kernel_launches++;
        c->use_frame = true;
    } else if (c->cfg.world_size > 1 && c->cfg.shard_mode == B200PT_SHARD_ROWS) {  // lane sums -> full frame, zeros elsewhere
        CU(cudaMemsetAsync(c->d_frame, 0, sizeof(f4) * c->fb_pixels, st));
        if (sv.n_lanes > 0) launch_scatter_rows(sv, c->la.sum, c->d_frame, st);
        S.kernel_launches++;
        c->use_frame = true;
    }
    CU(cudaEventRecord(c->ev1, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    S.render_ms = ms;
    S.other_ms = timing ? ms - S.extend_ms - S.shade_ms - S.shadow_ms : 0;
    {
        const uint64_t pix = (uint64_t)std::max(sv.n_pix_lanes > 0 ? sv.n_pix_lanes : sv.n_lanes, 0);
        const uint64_t V = sv.n_pix_lanes > 0 ? (uint64_t)(sv.n_lanes / sv.n_pix_lanes) : 1u;
        S.paths = pix * (V * (uint64_t)(sv.max_samples - sv.first_sample) + (uint64_t)sv.stream_extra);
    }
    if (count) {
        unsigned long long wc[2];
        CU(cudaMemcpy(wc, c->la.walk_counters, sizeof wc, cudaMemcpyDeviceToHost));
        S.nodes_visited = wc[0];
        S.tris_tested = wc[1];
    }
    c->rendered = true;
    c->resolved = false;
    c->resolve_samples = n_pass > 0 ? first_sample + n_pass : c->max_samples;
    c->pass_done = n_pass > 0 ? first_sample + n_pass : 0;
    return B200PT_OK;
}

extern "C" int b200pt_set_drain_callback(b200pt_ctx* c, float fraction, b200pt_drain_fn fn, void* user) {
    if (!c) return fail(B200PT_ERR_INVALID, "set_drain_callback: null context");
    c->drain_fraction = (fn && fraction > 0.0f) ? fraction : 0.0f;
    c->drain_fn = c->drain_fraction > 0.0f ? fn : nullptr;
    c->drain_user = user;
    return B200PT_OK;
}

extern "C" int b200pt_resolve(b200pt_ctx* c) {
    if (int r = check_ready(c, "resolve")) return r;
    if (!c->rendered) return fail(B200PT_ERR_INVALID, "resolve: nothing rendered");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    const SceneView full = make_view(c, true);
    const f4* src = c->use_frame ? c->d_frame : c->la.sum;
    const float inv = 1.0f / static_cast<float>(c->resolve_samples);  // device.cu:247 (max_samples for a complete frame)
    launch_resolve(full, src, inv, c->transfer, c->d_rgba, c->d_mean, c->stream);
    c->stats.kernel_launches++;
    // the product of a frame is the RGBA8 image (8.3 MB at 1080p); the float frame (33 MB) crosses PCIe only when
    // somebody asks for it (b200pt_framebuffer_f32, progressive callbacks)
    CU(cudaMemcpyAsync(c->h_rgba, c->d_rgba, sizeof(uint32_t) * c->fb_pixels, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    c->resolved = true;
    c->mean_on_host = false;
    return B200PT_OK;
}

static int fetch_mean(b200pt_ctx* c) {
    if (c->mean_on_host) return B200PT_OK;
    CU(cudaSetDevice(c->cfg.device_ordinal));
    CU(cudaMemcpyAsync(c->h_mean, c->d_mean, sizeof(f4) * c->fb_pixels, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->mean_on_host = true;
    return B200PT_OK;
}

extern "C" int b200pt_render(b200pt_ctx* c) {
    if (int r = check_ready(c, "render")) return r;
    if (int r = render_local(c)) return r;
    if (c->cfg.world_size == 1) return b200pt_resolve(c);
    return B200PT_OK;
}

// ---- progressive surface ----------------------------------------------------------------
extern "C" int b200pt_set_output(b200pt_ctx* c, int32_t transfer) {
    if (!c) return fail(B200PT_ERR_INVALID, "set_output: null context");
    if (transfer != B200PT_TRANSFER_LINEAR && transfer != B200PT_TRANSFER_GAMMA22) return fail(B200PT_ERR_INVALID, "set_output: unknown transfer");
    c->transfer = transfer;
    return B200PT_OK;
}

extern "C" int b200pt_render_restart(b200pt_ctx* c) {
    if (!c) return fail(B200PT_ERR_INVALID, "render_restart: null context");
    c->pass_done = 0;
    return B200PT_OK;
}

static int pass_statistics(b200pt_ctx* c, int n_samples, double* rel_variance) {
    const uint32_t n = (uint32_t)(c->width * c->height);
    const unsigned nb = pass_stats_blocks(n);
    launch_pass_stats(c->use_frame ? c->d_frame : c->la.sum, n, n_samples, c->d_pass_partial, c->stream);
    c->stats.kernel_launches++;
    c->h_pass_partial.resize(nb);
    CU(cudaMemcpyAsync(c->h_pass_partial.data(), c->d_pass_partial, sizeof(double) * nb, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    double t = 0.0;
    for (unsigned b = 0; b < nb; ++b) t += c->h_pass_partial[b];   // fixed order: the figure is reproducible
    *rel_variance = n ? t / (double)n : 0.0;
    return B200PT_OK;
}

extern "C" int b200pt_render_pass(b200pt_ctx* c, int32_t n_samples, b200pt_pass_info* info) {
    if (int r = check_ready(c, "render_pass")) return r;
    const int world = std::max(c->cfg.world_size, 1);
    if (c->cfg.streams_per_pixel > 1 || (world > 1 && c->cfg.shard_mode != B200PT_SHARD_SPP))
        return fail(B200PT_ERR_INVALID, "render_pass: progressive passes need one stream per pixel and, on several ranks, SPP sharding");
    if (world > 1 && c->max_samples % world != 0)
        return fail(B200PT_ERR_INVALID, "render_pass: max_samples must be a multiple of world_size");
    if (n_samples <= 0) return fail(B200PT_ERR_INVALID, "render_pass: n_samples must be positive");
    const int budget = c->max_samples / world;   // samples of this rank's stream
    if (c->pass_done >= budget)
        return fail(B200PT_ERR_INVALID, "render_pass: all max_samples samples are rendered (b200pt_render_restart starts over)");
    const int n = std::min(n_samples, budget - c->pass_done);
    if (int r = render_local(c, c->pass_done, n)) return r;
    if (world > 1) {
        // the rank's sums must survive the in-place reduce: the collective runs on a copy
        if (!c->d_frame) CU(cudaMalloc(&c->d_frame, sizeof(f4) * c->fb_pixels));
        CU(cudaMemcpyAsync(c->d_frame, c->la.sum, sizeof(f4) * c->fb_pixels, cudaMemcpyDeviceToDevice, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        c->use_frame = true;
        c->resolve_samples = c->pass_done * world;
    } else if (int r = b200pt_resolve(c)) {
        return r;
    }
    if (info) {
        memset(info, 0, sizeof *info);
        info->samples_done = c->pass_done * world;
        info->samples_in_pass = n * world;
        info->complete = c->pass_done >= budget ? 1 : 0;
        info->render_ms = c->stats.render_ms;
        info->rays = c->stats.rays;
        if (world == 1)
            if (int r = pass_statistics(c, c->pass_done, &info->rel_variance)) return r;
    }
    return B200PT_OK;
}

extern "C" int b200pt_pass_stats(b200pt_ctx* c, b200pt_pass_info* info) {
    if (int r = check_ready(c, "pass_stats")) return r;
    if (!info || !c->rendered) return fail(B200PT_ERR_INVALID, "pass_stats: nothing rendered / null output");
    memset(info, 0, sizeof *info);
    info->samples_done = c->resolve_samples;
    info->complete = c->resolve_samples >= c->max_samples ? 1 : 0;
    return pass_statistics(c, c->resolve_samples, &info->rel_variance);
}

extern "C" int b200pt_render_progressive(b200pt_ctx* c, int32_t pass_samples, double target_rel_variance, b200pt_pass_fn on_pass,
                                         void* user, b200pt_pass_info* last) {
    if (int r = check_ready(c, "render_progressive")) return r;
    if (pass_samples <= 0) return fail(B200PT_ERR_INVALID, "render_progressive: pass_samples must be positive");
    if (c->cfg.world_size > 1)
        return fail(B200PT_ERR_INVALID, "render_progressive: one rank only (several ranks: b200pt_render_pass + reduce + b200pt_pass_stats per pass)");
    if (int r = b200pt_render_restart(c)) return r;
    b200pt_pass_info info{};
    for (;;) {
        if (int r = b200pt_render_pass(c, pass_samples, &info)) return r;
        const bool converged = target_rel_variance > 0.0 && info.samples_done > 1 && info.rel_variance <= target_rel_variance;
        if (converged) info.stopped_early = info.complete ? 0 : 1;
        int stop = 0;
        if (on_pass) {
            if (int r = fetch_mean(c)) return r;
            stop = on_pass(user, &info, c->h_rgba, c->h_mean);
        }
        if (info.complete || converged || stop) break;
    }
    if (last) *last = info;
    return B200PT_OK;
}

extern "C" int b200pt_accum_buffer(b200pt_ctx* c, void** device_ptr, size_t* n_floats) {
    if (!c || !device_ptr || !n_floats) return fail(B200PT_ERR_INVALID, "accum_buffer: null");
    if (!c->rendered) return fail(B200PT_ERR_INVALID, "accum_buffer: nothing rendered");
    *device_ptr = c->use_frame ? (void*)c->d_frame : (void*)c->la.sum;
    *n_floats = 4 * c->fb_pixels;
    return B200PT_OK;
}

extern "C" int b200pt_render_to_host(b200pt_ctx* c, const b200pt_material* m, int32_t nm, uint32_t* rgba_out) {
    if (!rgba_out) return fail(B200PT_ERR_INVALID, "render_to_host: null output");
    if (m)
        if (int r = b200pt_set_materials(c, m, nm)) return r;
    if (int r = b200pt_render(c)) return r;
    if (!c->resolved) return fail(B200PT_ERR_INVALID, "render_to_host: multi-rank contexts resolve on rank 0 after the reduce");
    memcpy(rgba_out, c->h_rgba, sizeof(uint32_t) * c->fb_pixels);
    return B200PT_OK;
}

extern "C" int b200pt_framebuffer_rgba8(b200pt_ctx* c, const uint32_t** out) {
    if (!c || !out) return fail(B200PT_ERR_INVALID, "framebuffer: null");
    if (!c->resolved) return fail(B200PT_ERR_INVALID, "framebuffer: no resolved frame");
    *out = c->h_rgba;
    return B200PT_OK;
}
extern "C" int b200pt_framebuffer_f32(b200pt_ctx* c, const float** out) {
    if (!c || !out) return fail(B200PT_ERR_INVALID, "framebuffer: null");
    if (!c->resolved) return fail(B200PT_ERR_INVALID, "framebuffer: no resolved frame");
    if (int r = fetch_mean(c)) return r;
    *out = c->h_mean;
    return B200PT_OK;
}

extern "C" int b200pt_get_stats(b200pt_ctx* c, b200pt_stats* out) {
    if (!c || !out) return fail(B200PT_ERR_INVALID, "stats: null");
    *out = c->stats;
    return B200PT_OK;
}

// ---------------------------------------------------------------------------- parity hooks
static void hits_to_records(const b200pt_ctx* c, const std::vector<f4>& raw, float tmax, b200pt_hit* out) {
    for (size_t i = 0; i < raw.size(); ++i) {
        const uint32_t tri = f2u(raw[i].w);
        if (tri == kNoHit) {
            out[i] = b200pt_hit{-1, -1, tmax, 0.0f, 0.0f};
        } else {
            const int32_t mesh = c->h_idx4[4 * (size_t)tri + 3];
            out[i] = b200pt_hit{mesh, (int32_t)tri - c->h_meshes[mesh].first_tri, raw[i].x, raw[i].y, raw[i].z};
        }
    }
}

extern "C" int b200pt_primary_hits(b200pt_ctx* c, b200pt_hit* out) {
    if (int r = check_ready(c, "primary_hits")) return r;
    if (!out) return fail(B200PT_ERR_INVALID, "primary_hits: null output");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    const SceneView full = make_view(c, true);
    const size_t n = (size_t)c->width * c->height;
    f4* d_hits = nullptr;
    CU(cudaMalloc(&d_hits, sizeof(f4) * n));
    CU(cudaMemsetAsync(c->d_cursor, 0, sizeof(uint32_t), c->stream));
    launch_primary(full, d_hits, c->d_cursor, c->stream);
    std::vector<f4> raw(n);
    cudaError_t e = cudaMemcpyAsync(raw.data(), d_hits, sizeof(f4) * n, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_hits);
    if (e != cudaSuccess) return fail(B200PT_ERR_CUDA, std::string("primary_hits: ") + cudaGetErrorString(e));
    hits_to_records(c, raw, kTMax, out);
    return B200PT_OK;
}

extern "C" int b200pt_trace(b200pt_ctx* c, const float* origins, const float* dirs, int32_t n, float tmin, float tmax,
                            b200pt_hit* out) {
    if (!c || !c->built) return fail(B200PT_ERR_INVALID, "trace: call b200pt_build_accel first");
    if (!origins || !dirs || !out || n <= 0) return fail(B200PT_ERR_INVALID, "trace: null/empty");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    SceneView sv{};
    sv.nodes = c->bvh.nodes;
    sv.tris = c->bvh.tris;
    sv.plane_magic = 0x3F800000u;
    float *d_o = nullptr, *d_d = nullptr;
    f4* d_hits = nullptr;
    CU(cudaMalloc(&d_o, sizeof(float) * 3 * (size_t)n));
    CU(cudaMalloc(&d_d, sizeof(float) * 3 * (size_t)n));
    CU(cudaMalloc(&d_hits, sizeof(f4) * (size_t)n));
    std::vector<f4> raw((size_t)n);
    cudaError_t e = cudaMemcpyAsync(d_o, origins, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_d, dirs, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        cudaMemsetAsync(c->d_cursor, 0, sizeof(uint32_t), c->stream);
        cudaEventRecord(c->ev0, c->stream);
        launch_trace(sv, d_o, d_d, n, tmin, tmax, d_hits, c->d_cursor, c->stream);
        cudaEventRecord(c->ev1, c->stream);
        e = cudaMemcpyAsync(raw.data(), d_hits, sizeof(f4) * (size_t)n, cudaMemcpyDeviceToHost, c->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) {  // stats: device time of the walk kernel alone, rays = n
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        c->stats.render_ms = ms;
        c->stats.rays = (uint64_t)n;
    }
    cudaFree(d_o); cudaFree(d_d); cudaFree(d_hits);
    if (e != cudaSuccess) return fail(B200PT_ERR_CUDA, std::string("trace: ") + cudaGetErrorString(e));
    hits_to_records(c, raw, tmax, out);
    return B200PT_OK;
}

extern "C" int b200pt_sample_disney(b200pt_ctx* c, int32_t n, const int32_t* material, const float* wo, const int32_t* lobe_in,
                                    const uint32_t* rng_in, float* f, float* wi, float* pdf, int32_t* lobe_out,
                                    uint32_t* rng_out) {
    if (!c || !c->d_mats) return fail(B200PT_ERR_INVALID, "sample_disney: call b200pt_set_materials first");
    if (n <= 0 || !material || !wo || !lobe_in || !rng_in || !f || !wi || !pdf || !lobe_out || !rng_out)
        return fail(B200PT_ERR_INVALID, "sample_disney: null/empty");
    for (int i = 0; i < n; ++i)
        if (material[i] < 0 || material[i] >= c->n_mats) return fail(B200PT_ERR_BOUNDS, "sample_disney: material index out of range");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    const size_t N = (size_t)n;
    char* d = nullptr;  // one allocation: material, lobe_in, rng_in, lobe_out, rng_out, pdf (4 B each), wo, f, wi (12 B each)
    CU(cudaMalloc(&d, N * (6 * 4 + 3 * 12)));
    int32_t* d_mat = (int32_t*)d;
    int32_t* d_li = d_mat + N;
    uint32_t* d_ri = (uint32_t*)(d_li + N);
    int32_t* d_lo = (int32_t*)(d_ri + N);
    uint32_t* d_ro = (uint32_t*)(d_lo + N);
    float* d_pdf = (float*)(d_ro + N);
    float* d_wo = d_pdf + N;
    float* d_f = d_wo + 3 * N;
    float* d_wi = d_f + 3 * N;
    cudaStream_t st = c->stream;
    cudaError_t e = cudaMemcpyAsync(d_mat, material, 4 * N, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_li, lobe_in, 4 * N, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ri, rng_in, 4 * N, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_wo, wo, 12 * N, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        launch_sample_disney(c->d_mats, n, d_mat, d_wo, d_li, d_ri, d_f, d_wi, d_pdf, d_lo, d_ro, st);
        e = cudaMemcpyAsync(f, d_f, 12 * N, cudaMemcpyDeviceToHost, st);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(wi, d_wi, 12 * N, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(pdf, d_pdf, 4 * N, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(lobe_out, d_lo, 4 * N, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rng_out, d_ro, 4 * N, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d);
    if (e != cudaSuccess) return fail(B200PT_ERR_CUDA, std::string("sample_disney: ") + cudaGetErrorString(e));
    return B200PT_OK;
}

extern "C" int b200pt_export_bvh(b200pt_ctx* c, void* nodes, uint32_t* n_nodes, void* tris, uint32_t* n_tris) {
    if (!c || !c->built) return fail(B200PT_ERR_INVALID, "export_bvh: call b200pt_build_accel first");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    if (n_nodes) *n_nodes = c->bvh.n_nodes;
    if (n_tris) *n_tris = c->bvh.n_tris;
    if (nodes) CU(cudaMemcpy(nodes, c->bvh.nodes, sizeof(Node8) * (size_t)c->bvh.n_nodes, cudaMemcpyDeviceToHost));
    if (tris) CU(cudaMemcpy(tris, c->bvh.tris, sizeof(TriPair) * (size_t)c->bvh.n_tris, cudaMemcpyDeviceToHost));
    return B200PT_OK;
}

// ---------------------------------------------------------------------------- the frame's collective (NCCL)
// libnccl.so.2 is opened on first use: single-GPU users never need it, and inside a process that already carries a
// copy (torch bundles one) dlopen by soname hands back that copy instead of loading a second one.
namespace {
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};
NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!api.handle) {
            api.error = std::string("libnccl.so.2 cannot be loaded: ") + dlerror();
            return;
        }
        auto sym = [&](const char* name) {
            void* p = dlsym(api.handle, name);
            if (!p && api.error.empty()) api.error = std::string("libnccl.so.2 lacks ") + name;
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.Reduce = reinterpret_cast<decltype(api.Reduce)>(sym("ncclReduce"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    });
    return &api;
}
}  // namespace
#define NC(x)                                                                                                  \
    do {                                                                                                       \
        ncclResult_t r_ = (x);                                                                                 \
        if (r_ != ncclSuccess) return fail(B200PT_ERR_CUDA, std::string(#x) + ": " + nccl_api()->GetErrorString(r_)); \
    } while (0)

static void comm_release(b200pt_ctx* c) {
    if (c->comm) {
        nccl_api()->CommDestroy(c->comm);
        c->comm = nullptr;
    }
}

static int nccl_ready() {
    NcclApi* a = nccl_api();
    if (!a->error.empty()) return fail(B200PT_ERR_CUDA, a->error);
    return B200PT_OK;
}

extern "C" int b200pt_comm_unique_id(void* id_out) {
    static_assert(sizeof(ncclUniqueId) == B200PT_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    if (!id_out) return fail(B200PT_ERR_INVALID, "comm_unique_id: null");
    if (int r = nccl_ready()) return r;
    ncclUniqueId id;
    NC(nccl_api()->GetUniqueId(&id));
    memcpy(id_out, &id, sizeof id);
    return B200PT_OK;
}

extern "C" int b200pt_comm_init_rank(b200pt_ctx* c, const void* id_in) {
    if (!c || !id_in) return fail(B200PT_ERR_INVALID, "comm_init_rank: null");
    if (int r = nccl_ready()) return r;
    CU(cudaSetDevice(c->cfg.device_ordinal));
    comm_release(c);
    ncclUniqueId id;
    memcpy(&id, id_in, sizeof id);
    NC(nccl_api()->CommInitRank(&c->comm, c->cfg.world_size, id, c->cfg.rank));
    return B200PT_OK;
}

extern "C" int b200pt_comm_init_all(b200pt_ctx* const* ctxs, int32_t n) {
    if (!ctxs || n < 1) return fail(B200PT_ERR_INVALID, "comm_init_all: null / empty");
    if (int r = nccl_ready()) return r;
    std::vector<int> devs((size_t)n);
    for (int i = 0; i < n; ++i) {
        if (!ctxs[i] || ctxs[i]->cfg.rank != i || ctxs[i]->cfg.world_size != n)
            return fail(B200PT_ERR_INVALID, "comm_init_all: ctxs[r] must be rank r of world_size n");
        for (int j = 0; j < i; ++j)
            if (ctxs[j]->cfg.device_ordinal == ctxs[i]->cfg.device_ordinal)
                return fail(B200PT_ERR_INVALID, "comm_init_all: two ranks on one device");
        devs[(size_t)i] = ctxs[i]->cfg.device_ordinal;
        comm_release(ctxs[i]);
    }
    std::vector<ncclComm_t> comms((size_t)n);
    NC(nccl_api()->CommInitAll(comms.data(), n, devs.data()));
    for (int i = 0; i < n; ++i) ctxs[i]->comm = comms[(size_t)i];
    return B200PT_OK;
}

extern "C" int b200pt_reduce(b200pt_ctx* c) {
    if (!c || !c->rendered) return fail(B200PT_ERR_INVALID, "reduce: nothing rendered");
    if (c->cfg.world_size == 1) return B200PT_OK;
    if (!c->comm) return fail(B200PT_ERR_INVALID, "reduce: call b200pt_comm_init_rank / b200pt_comm_init_all first");
    CU(cudaSetDevice(c->cfg.device_ordinal));
    f4* buf = c->use_frame ? c->d_frame : c->la.sum;
    NC(nccl_api()->Reduce(buf, buf, 4 * c->fb_pixels, ncclFloat, ncclSum, 0, c->comm, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return B200PT_OK;
}

extern "C" int b200pt_render_frame(b200pt_ctx* c) {
    if (int r = check_ready(c, "render_frame")) return r;
    if (int r = render_local(c)) return r;
    if (c->cfg.world_size > 1)
        if (int r = b200pt_reduce(c)) return r;
    if (c->cfg.rank == 0) return b200pt_resolve(c);
    return B200PT_OK;
}

extern "C" int b200pt_render_frame_all(b200pt_ctx* const* ctxs, int32_t n) {
    if (!ctxs || n < 1) return fail(B200PT_ERR_INVALID, "render_frame_all: null / empty");
    if (n == 1) return b200pt_render_frame(ctxs[0]);
    std::vector<int> status((size_t)n, B200PT_OK);
    std::vector<std::string> message((size_t)n);
    std::vector<std::thread> th;
    for (int i = 0; i < n; ++i)
        th.emplace_back([&, i] {
            status[(size_t)i] = b200pt_render_frame(ctxs[i]);
            if (status[(size_t)i] != B200PT_OK) message[(size_t)i] = g_error;   // g_error is thread-local
        });
    for (auto& t : th) t.join();
    for (int i = 0; i < n; ++i)
        if (status[(size_t)i] != B200PT_OK) return fail(status[(size_t)i], "rank " + std::to_string(i) + ": " + message[(size_t)i]);
    return B200PT_OK;
}

// ---------------------------------------------------------------------------- b200pt_trace_static
// Variant of b200pt_trace that allocates nothing (no cudaMalloc / cudaFree, no allocating library call): the staging
// the walk kernel needs — ray-fetch cursor, origins, directions, raw hit records — lives in ONE file-scope device array
// of fixed size.  No counterpart in the reference; added for callers that
// cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// Every GPU that loads the library holds the array until the process exits (2.5 MiB).  All calls share it and it keeps
// the last call's contents: the cursor is cleared on every call, the rest is overwritten before it is read; calls
// that overlap on one GPU race on it (serialise them).
namespace {
constexpr size_t kStaticHeader = 256;   // cursor (4 bytes used), keeps the arrays behind it 256-byte aligned
constexpr size_t kStaticBytes = kStaticHeader + (size_t)B200PT_TRACE_STATIC_MAX * (12 + 12 + 16);
static __device__ __align__(256) unsigned char g_trace_scratch[kStaticBytes];

// In host code `g_trace_scratch` is only a stand-in for the symbol, not a device address: the address is asked for once
// per device, with that device current.
unsigned char* static_scratch_of(int device, std::string& err) {
    static std::mutex mu;
    static unsigned char* table[64] = {};
    if (device < 0 || device >= 64) {
        err = "device ordinal out of range";
        return nullptr;
    }
    std::lock_guard<std::mutex> g(mu);
    if (!table[device]) {
        cudaError_t e = cudaSetDevice(device);
        void* p = nullptr;
        if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_trace_scratch);
        if (e != cudaSuccess || !p) {
            err = std::string("cudaGetSymbolAddress(g_trace_scratch): ") + cudaGetErrorString(e);
            return nullptr;
        }
        table[device] = static_cast<unsigned char*>(p);
    }
    return table[device];
}
}  // namespace

extern "C" int b200pt_trace_static(b200pt_ctx* c, const float* origins, const float* dirs, int32_t n, float tmin, float tmax,
                                   b200pt_hit* out) {
    if (!c || !c->built) return fail(B200PT_ERR_INVALID, "trace_static: call b200pt_build_accel first");
    if (!origins || !dirs || !out || n <= 0) return fail(B200PT_ERR_INVALID, "trace_static: null/empty");
    if (n > B200PT_TRACE_STATIC_MAX)
        return fail(B200PT_ERR_BOUNDS, "trace_static: " + std::to_string(n) + " rays exceed the fixed scratch of " +
                                           std::to_string(B200PT_TRACE_STATIC_MAX) + " rays (use b200pt_trace, or split the batch)");
    std::string err;
    unsigned char* base = static_scratch_of(c->cfg.device_ordinal, err);
    if (!base) return fail(B200PT_ERR_CUDA, "trace_static: " + err);
    CU(cudaSetDevice(c->cfg.device_ordinal));
    uint32_t* d_cursor = reinterpret_cast<uint32_t*>(base);
    float* d_o = reinterpret_cast<float*>(base + kStaticHeader);
    float* d_d = reinterpret_cast<float*>(base + kStaticHeader + (size_t)B200PT_TRACE_STATIC_MAX * 12);
    f4* d_hits = reinterpret_cast<f4*>(base + kStaticHeader + (size_t)B200PT_TRACE_STATIC_MAX * 24);
    SceneView sv{};
    sv.nodes = c->bvh.nodes;
    sv.tris = c->bvh.tris;
    sv.plane_magic = 0x3F800000u;
    CU(cudaMemsetAsync(d_cursor, 0, sizeof(uint32_t), c->stream));   // the one thing that must start at zero
    CU(cudaMemcpyAsync(d_o, origins, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_d, dirs, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    launch_trace(sv, d_o, d_d, n, tmin, tmax, d_hits, d_cursor, c->stream);
    // raw 16-byte records land at the front of the caller's 20-byte-record buffer and are widened in place, last first
    // (record i is read from bytes [16 i, 16 i + 16) before bytes [20 i, 20 i + 20) are written): no host staging either
    CU(cudaMemcpyAsync(out, d_hits, sizeof(f4) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    const unsigned char* raw = reinterpret_cast<const unsigned char*>(out);
    for (int64_t i = (int64_t)n - 1; i >= 0; --i) {
        f4 r;
        memcpy(&r, raw + 16 * (size_t)i, sizeof r);
        const uint32_t tri = f2u(r.w);
        b200pt_hit h;
        if (tri == kNoHit) {
            h = b200pt_hit{-1, -1, tmax, 0.0f, 0.0f};
        } else {
            const int32_t mesh = c->h_idx4[4 * (size_t)tri + 3];
            h = b200pt_hit{mesh, (int32_t)tri - c->h_meshes[mesh].first_tri, r.x, r.y, r.z};
        }
        memcpy(reinterpret_cast<unsigned char*>(out) + sizeof(b200pt_hit) * (size_t)i, &h, sizeof h);
    }
    return B200PT_OK;
}

// ---------------------------------------------------------------------------- frame pipeline (b200pt_pipeline_*)
// One worker thread per slot.  Frame s goes to slot s % n and may start rendering once frame s-1 reported that its queue
// is draining (b200pt_set_drain_callback) or ended; the frames' reduces (several ranks) run in frame order.
#include <condition_variable>
#include <deque>

struct b200pt_pipeline {
    struct Job {
        uint64_t seq = 0;
        std::vector<b200pt_material> mats;
        uint32_t* rgba_out = nullptr;
    };
    struct Slot {
        b200pt_pipeline* owner = nullptr;
        b200pt_ctx* ctx = nullptr;
        std::thread worker;
        bool has_job = false, busy = false;
        Job job;
        uint64_t running_seq = 0;
    };
    std::vector<Slot> slots;
    std::mutex mu;
    std::condition_variable cv;
    uint64_t submitted = 0, completed = 0;
    uint64_t drained_upto = 0;   // frames with seq < drained_upto have drained (or ended)
    uint64_t reduce_turn = 0;    // seq of the frame whose reduce may run next
    bool stop = false;
    int status = B200PT_OK;
    std::string error;

    void mark_drained(uint64_t seq) {
        std::lock_guard<std::mutex> g(mu);
        if (drained_upto < seq + 1) drained_upto = seq + 1;
        cv.notify_all();
    }
    void record_error(int code) {   // g_error of the calling thread holds the message
        std::lock_guard<std::mutex> g(mu);
        if (status == B200PT_OK) {
            status = code;
            error = g_error;
        }
    }
    static void on_drain(void* user) {
        Slot* s = static_cast<Slot*>(user);
        s->owner->mark_drained(s->running_seq);
    }
    void run(Slot* s) {
        for (;;) {
            Job job;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || (s->has_job && drained_upto >= s->job.seq); });
                if (stop && !s->has_job) return;
                job = std::move(s->job);
                s->has_job = false;
                s->busy = true;
                s->running_seq = job.seq;
            }
            b200pt_ctx* c = s->ctx;
            int r = B200PT_OK;
            if (!job.mats.empty()) r = b200pt_set_materials(c, job.mats.data(), (int32_t)job.mats.size());
            if (r == B200PT_OK) r = check_ready(c, "pipeline");
            if (r == B200PT_OK) r = render_local(c);
            mark_drained(job.seq);   // whatever happened, the next frame must not wait for this one
            {   // the collective in frame order (every rank issues the same sequence); a failed frame still takes its turn
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return reduce_turn == job.seq; });
            }
            if (r == B200PT_OK && c->cfg.world_size > 1) r = b200pt_reduce(c);
            {
                std::lock_guard<std::mutex> g(mu);
                reduce_turn = job.seq + 1;
                cv.notify_all();
            }
            if (r == B200PT_OK && c->cfg.rank == 0) {
                r = b200pt_resolve(c);
                if (r == B200PT_OK && job.rgba_out) memcpy(job.rgba_out, c->h_rgba, sizeof(uint32_t) * c->fb_pixels);
            }
            if (r != B200PT_OK) record_error(r);
            {
                std::lock_guard<std::mutex> g(mu);
                s->busy = false;
                ++completed;
                cv.notify_all();
            }
        }
    }
};

extern "C" int b200pt_pipeline_create(b200pt_ctx* const* slots, int32_t n, float start_fraction, b200pt_pipeline** out) {
    if (!slots || n < 1 || !out) return fail(B200PT_ERR_INVALID, "pipeline_create: null / empty");
    for (int i = 0; i < n; ++i) {
        if (!slots[i]) return fail(B200PT_ERR_INVALID, "pipeline_create: null slot");
        if (slots[i]->cfg.rank != slots[0]->cfg.rank || slots[i]->cfg.world_size != slots[0]->cfg.world_size ||
            slots[i]->width != slots[0]->width || slots[i]->height != slots[0]->height)
            return fail(B200PT_ERR_INVALID, "pipeline_create: the slots must be contexts of one rank with the same frame size");
        for (int j = 0; j < i; ++j)
            if (slots[j] == slots[i]) return fail(B200PT_ERR_INVALID, "pipeline_create: a context can fill one slot only");
    }
    auto* p = new b200pt_pipeline();
    p->slots.resize((size_t)n);
    const float f = start_fraction > 0.0f ? start_fraction : 0.5f;
    for (int i = 0; i < n; ++i) {
        b200pt_pipeline::Slot& s = p->slots[(size_t)i];
        s.owner = p;
        s.ctx = slots[i];
        b200pt_set_drain_callback(s.ctx, f, &b200pt_pipeline::on_drain, &s);
        if (n > 1) s.ctx->parts_override = 1;
    }
    for (auto& s : p->slots) s.worker = std::thread([p, sp = &s] { p->run(sp); });
    *out = p;
    return B200PT_OK;
}

extern "C" int b200pt_pipeline_submit(b200pt_pipeline* p, const b200pt_material* m, int32_t nm, uint32_t* rgba_out) {
    if (!p) return fail(B200PT_ERR_INVALID, "pipeline_submit: null");
    std::unique_lock<std::mutex> lk(p->mu);
    b200pt_pipeline::Slot& s = p->slots[(size_t)(p->submitted % p->slots.size())];
    p->cv.wait(lk, [&] { return !s.has_job && !s.busy; });
    s.job.seq = p->submitted++;
    s.job.mats.assign(m ? m : nullptr, m ? m + nm : nullptr);
    s.job.rgba_out = rgba_out;
    s.has_job = true;
    p->cv.notify_all();
    return B200PT_OK;
}

extern "C" int b200pt_pipeline_wait(b200pt_pipeline* p) {
    if (!p) return fail(B200PT_ERR_INVALID, "pipeline_wait: null");
    std::unique_lock<std::mutex> lk(p->mu);
    p->cv.wait(lk, [&] { return p->completed == p->submitted; });
    const int st = p->status;
    if (st != B200PT_OK) {
        g_error = p->error;
        p->status = B200PT_OK;
        p->error.clear();
    }
    return st;
}

extern "C" int b200pt_pipeline_destroy(b200pt_pipeline* p) {
    if (!p) return B200PT_OK;
    b200pt_pipeline_wait(p);
    {
        std::lock_guard<std::mutex> g(p->mu);
        p->stop = true;
        p->cv.notify_all();
    }
    for (auto& s : p->slots) {
        s.worker.join();
        b200pt_set_drain_callback(s.ctx, 0.0f, nullptr, nullptr);
        s.ctx->parts_override = 0;
    }
    delete p;
    return B200PT_OK;
}

// ---------------------------------------------------------------------------- b200pt_dedup_mesh
extern "C" int b200pt_dedup_mesh(b200pt_ctx* c, const float* av, int32_t n_v, const float* an, int32_t n_vn, const float* at,
                                 int32_t n_vt, const int32_t* cv, const int32_t* cn, const int32_t* ct, int32_t n_corners,
                                 float* out_verts, float* out_normals, float* out_uvs, int32_t* out_idx, int32_t* n_out_verts) {
    if (!c || !av || !cv || !out_verts || !out_idx || !n_out_verts || n_corners <= 0 || n_corners % 3 != 0 || n_v <= 0)
        return fail(B200PT_ERR_INVALID, "dedup_mesh: null / empty input (corners come in triangles)");
    if ((cn && (!an || !out_normals)) || (ct && (!at || !out_uvs))) return fail(B200PT_ERR_INVALID, "dedup_mesh: indices without their attribute / output array");
    // host-side validation: ranges, and regularity of the optional indices
    int32_t lo = INT32_MAX, hi = -1;
    for (int32_t i = 0; i < n_corners; ++i) {
        const int32_t v = cv[i];
        if (v < 0 || v >= n_v) return fail(B200PT_ERR_BOUNDS, "dedup_mesh: vertex index out of range");
        lo = std::min(lo, v);
        hi = std::max(hi, v);
        if (cn && (cn[i] < 0 || cn[i] >= n_vn))
            return fail(cn[i] < 0 ? B200PT_ERR_INVALID : B200PT_ERR_BOUNDS, cn[i] < 0 ? "dedup_mesh: some corners have no normal index (irregular shape: host routine)" : "dedup_mesh: normal index out of range");
        if (ct && (ct[i] < 0 || ct[i] >= n_vt))
            return fail(ct[i] < 0 ? B200PT_ERR_INVALID : B200PT_ERR_BOUNDS, ct[i] < 0 ? "dedup_mesh: some corners have no texcoord index (irregular shape: host routine)" : "dedup_mesh: texcoord index out of range");
    }
    CU(cudaSetDevice(c->cfg.device_ordinal));
    const uint32_t n = (uint32_t)n_corners, range = (uint32_t)(hi - lo) + 1u;
    // only the slice [lo, hi] of the position array travels; normals / texcoords travel whole (their indices are not bounded
    // by the shape's position range)
    const size_t b_cv = sizeof(int32_t) * n, b_av = sizeof(float) * 3 * (size_t)range;
    const size_t b_an = cn ? sizeof(float) * 3 * (size_t)n_vn : 0, b_at = ct ? sizeof(float) * 2 * (size_t)n_vt : 0;
    const size_t scan_bytes = dedup_scan_bytes(n);
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t total = up(b_cv) * (1 + (cn ? 1 : 0) + (ct ? 1 : 0)) + up(b_av) + up(b_an) + up(b_at) + up(sizeof(uint32_t) * range) +
                         2 * up(sizeof(uint32_t) * n) + up(scan_bytes) + up(b_cv) /*out_idx*/ + up(sizeof(float) * 3 * (size_t)n) * (1 + (cn ? 1 : 0)) +
                         (ct ? up(sizeof(float) * 2 * (size_t)n) : 0);
    unsigned char* arena = nullptr;
    CU(cudaMalloc(&arena, total));
    size_t off = 0;
    auto take = [&](size_t b) { unsigned char* p = arena + off; off += up(b); return p; };
    int32_t* d_cv = (int32_t*)take(b_cv);
    int32_t* d_cn = cn ? (int32_t*)take(b_cv) : nullptr;
    int32_t* d_ct = ct ? (int32_t*)take(b_cv) : nullptr;
    float* d_av = (float*)take(b_av);
    float* d_an = cn ? (float*)take(b_an) : nullptr;
    float* d_at = ct ? (float*)take(b_at) : nullptr;
    uint32_t* d_first = (uint32_t*)take(sizeof(uint32_t) * range);
    uint32_t* d_flag = (uint32_t*)take(sizeof(uint32_t) * n);
    uint32_t* d_rank = (uint32_t*)take(sizeof(uint32_t) * n);
    void* d_scan = take(scan_bytes);
    int32_t* d_idx = (int32_t*)take(b_cv);
    float* d_ov = (float*)take(sizeof(float) * 3 * (size_t)n);
    float* d_on = cn ? (float*)take(sizeof(float) * 3 * (size_t)n) : nullptr;
    float* d_ot = ct ? (float*)take(sizeof(float) * 2 * (size_t)n) : nullptr;
    cudaStream_t st = c->stream;
    cudaError_t e = cudaMemcpyAsync(d_cv, cv, b_cv, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && cn) e = cudaMemcpyAsync(d_cn, cn, b_cv, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && ct) e = cudaMemcpyAsync(d_ct, ct, b_cv, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_av, av + 3 * (size_t)lo, b_av, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && cn) e = cudaMemcpyAsync(d_an, an, b_an, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && ct) e = cudaMemcpyAsync(d_at, at, b_at, cudaMemcpyHostToDevice, st);
    uint32_t n_unique = 0;
    // the position slice starts at index lo: the kernels index it with v, so hand them the pointer shifted back
    if (e == cudaSuccess)
        e = dedup_mesh_device(d_cv, d_cn, d_ct, n, lo, range, d_av - 3 * (size_t)lo, d_an, d_at, d_first, d_flag, d_rank, d_scan, scan_bytes,
                              d_idx, d_ov, d_on, d_ot, &n_unique, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_idx, d_idx, b_cv, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_verts, d_ov, sizeof(float) * 3 * (size_t)n_unique, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && cn) e = cudaMemcpyAsync(out_normals, d_on, sizeof(float) * 3 * (size_t)n_unique, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && ct) e = cudaMemcpyAsync(out_uvs, d_ot, sizeof(float) * 2 * (size_t)n_unique, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(arena);
    if (e != cudaSuccess) return fail(B200PT_ERR_CUDA, std::string("dedup_mesh: ") + cudaGetErrorString(e));
    *n_out_verts = (int32_t)n_unique;
    return B200PT_OK;
}

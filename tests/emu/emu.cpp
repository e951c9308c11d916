// emu.cpp — host emulation of the product's kernels.  TEST TOOL, never shipped, never
// linked into libb200pt.so.  It compiles the very same per-element / per-lane functions
// the CUDA kernels call (owl_path_tracer_b200/csrc/pt_*.h) with g++ and runs them in
// plain loops, so that the builder, the 8-wide walk and the shade state machine can be
// checked against the oracle in the CPU-only container (pytest -m "not gpu"), before any
// GPU time is spent.  The GPU tests then only have to show that nvcc's code returns the
// same bits.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <thread>
#include <vector>

#include "../../owl_path_tracer_b200/csrc/pt_build.h"
#include "../../owl_path_tracer_b200/csrc/pt_integrator.h"
#include "../../owl_path_tracer_b200/csrc/pt_lights.h"

using namespace pt;

struct emu_scene {
    std::vector<f4> pos, nrm;
    std::vector<f2> uv;
    std::vector<int32_t> idx4;
    std::vector<MeshInfo> meshes;
    std::vector<uint32_t> texels, env;
    std::vector<float> env_rows, env_cols;
    std::vector<f4> env_f32;
    std::vector<Material> mats, mesh_mats;
    std::vector<LightTri> lights;
    std::vector<float> light_cdf;
    int nee = 0;
    uint64_t shadow_rays = 0;
    std::vector<Node8> nodes;
    std::vector<TriPair> tris;
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f}, max_abs = 0;
    int levels = 0;
    std::vector<uint32_t> level_end;          // DeviceBVH::level_end of the emulated build
    std::vector<int32_t> mesh_vbase, mesh_nv;  // vertex range of every mesh (emu_update_mesh)
    int ploc_rounds = 0;
    long reinsert_moves = 0;
    double area_clustered = 0, area_reinserted = 0;  // sum of the binary tree's inner-node half areas
    int debug_px = -1, debug_py = -1;  // print every ray of this pixel (debugging aid)
    int force_lbvh = 0;  // tools/bvh_quality.py: build the fallback tree on purpose
    uint64_t walk_nodes = 0, walk_tris = 0;  // node fetches / triangle tests of the last emu_render
    SceneView sv{};
};

static void refresh_view(emu_scene* s) {
    s->sv.nodes = s->nodes.data();
    s->sv.tris = s->tris.data();
    s->sv.idx4 = s->idx4.data();
    s->sv.vpos = s->pos.data();
    s->sv.vnrm = s->nrm.data();
    s->sv.vuv = s->uv.data();
    s->sv.meshes = s->meshes.data();
    s->sv.materials = s->mats.data();
    s->mesh_mats.resize(s->meshes.size());
    for (size_t i = 0; i < s->meshes.size(); ++i)
        if (s->meshes[i].material >= 0 && (size_t)s->meshes[i].material < s->mats.size()) s->mesh_mats[i] = s->mats[s->meshes[i].material];
    s->sv.mesh_materials = s->mesh_mats.data();
    // next-event estimation mode: same table builder as api.cu
    s->lights.clear();
    s->light_cdf.clear();
    for (MeshInfo& mi : s->meshes) mi.light_first = -1;
    if (s->nee) build_light_table(s->pos, s->idx4, s->meshes, s->mats, s->lights, s->light_cdf);
    s->sv.lights = s->lights.data();
    s->sv.light_cdf = s->light_cdf.data();
    s->sv.n_lights = (int)s->lights.size();
    s->sv.nee = s->nee;
    s->sv.texels = s->texels.data();
    s->sv.env = s->env.empty() ? nullptr : s->env.data();
    s->sv.env_f32 = s->env_f32.empty() ? nullptr : s->env_f32.data();
    s->env_rows.clear();
    s->env_cols.clear();
    if (s->nee && s->sv.env_use && !s->env.empty()) build_env_table(s->env.data(), s->sv.env_w, s->sv.env_h, s->env_rows, s->env_cols);
    if (s->nee && s->sv.env_use && !s->env_f32.empty())
        build_env_table_f32(s->env_f32.data(), s->sv.env_w, s->sv.env_h, s->env_rows, s->env_cols);
    s->sv.env_cdf_rows = s->env_rows.data();
    s->sv.env_cdf_cols = s->env_cols.data();
    s->sv.env_nee = (s->nee && s->sv.env_use && !s->env_rows.empty()) ? 1 : 0;
    s->sv.n_materials = (int)s->mats.size();
}

struct emu_hit {
    int32_t mesh, prim;
    float t, u, v;
};
static emu_hit to_rec(const emu_scene* s, const Hit& h, float tmax) {
    if (h.tri == kNoHit) return emu_hit{-1, -1, tmax, 0, 0};
    const int32_t mesh = s->idx4[4 * (size_t)h.tri + 3];
    return emu_hit{mesh, (int32_t)h.tri - s->meshes[mesh].first_tri, h.t, h.u, h.v};
}

template <typename F>
static void parallel_rows(int n, int nthreads, F&& f) {
    std::atomic<int> next{0};
    auto w = [&]() {
        for (;;) {
            int i = next.fetch_add(1);
            if (i >= n) break;
            f(i);
        }
    };
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; ++i) th.emplace_back(w);
    w();
    for (auto& t : th) t.join();
}

extern "C" {

emu_scene* emu_create() { return new emu_scene(); }
void emu_destroy(emu_scene* s) { delete s; }

int emu_add_mesh(emu_scene* s, const float* verts, const float* normals, const float* uvs, int nv, const int32_t* idx,
                 int nt, int material, const uint8_t* tex, int tw, int th) {
    const int32_t vbase = (int32_t)s->pos.size(), tbase = (int32_t)(s->idx4.size() / 4), mesh = (int32_t)s->meshes.size();
    for (int i = 0; i < nv; ++i) {
        s->pos.push_back(f4{verts[3 * i], verts[3 * i + 1], verts[3 * i + 2], 0});
        s->nrm.push_back(f4{normals[3 * i], normals[3 * i + 1], normals[3 * i + 2], 0});
        s->uv.push_back(uvs ? f2{uvs[2 * i], uvs[2 * i + 1]} : f2{0, 0});
        for (int a = 0; a < 3; ++a) {
            s->lo[a] = fminf(s->lo[a], verts[3 * i + a]);
            s->hi[a] = fmaxf(s->hi[a], verts[3 * i + a]);
            s->max_abs = fmaxf(s->max_abs, fabsf(verts[3 * i + a]));
        }
    }
    for (int t = 0; t < nt; ++t) {
        for (int k = 0; k < 3; ++k) s->idx4.push_back(vbase + idx[3 * t + k]);
        s->idx4.push_back(mesh);
    }
    MeshInfo mi{};
    mi.first_tri = tbase;
    mi.material = material;
    mi.tex_offset = -1;
    mi.has_uv = uvs ? 1 : 0;
    if (tex) {
        mi.tex_offset = (int32_t)s->texels.size();
        mi.tex_w = tw;
        mi.tex_h = th;
        const uint32_t* px = (const uint32_t*)tex;
        s->texels.insert(s->texels.end(), px, px + (size_t)tw * th);
    }
    s->meshes.push_back(mi);
    s->mesh_vbase.push_back(vbase);
    s->mesh_nv.push_back(nv);
    return mesh;
}

void emu_set_materials(emu_scene* s, const float* m, int nm) {
    s->mats.assign((size_t)nm, Material{});
    for (int i = 0; i < nm; ++i) memcpy(&s->mats[i], m + 17 * i, 68);
}
void emu_set_environment(emu_scene* s, const uint8_t* rgba, int w, int h, int use, int auto_, const float* color,
                         float intensity) {
    s->env.clear();
    s->env_f32.clear();
    if (rgba && w > 0 && h > 0) s->env.assign((const uint32_t*)rgba, (const uint32_t*)rgba + (size_t)w * h);
    s->sv.env_w = s->env.empty() ? 0 : w;
    s->sv.env_h = s->env.empty() ? 0 : h;
    s->sv.env_use = use != 0;
    s->sv.env_auto = auto_ != 0;
    memcpy(s->sv.env_color, color, 12);
    s->sv.env_intensity = intensity;
}
void emu_set_environment_f32(emu_scene* s, const float* rgb, int w, int h, int use, int auto_, const float* color,
                             float intensity) {
    s->env.clear();
    s->env_f32.clear();
    if (rgb && w > 0 && h > 0) {
        s->env_f32.resize((size_t)w * h);
        for (size_t i = 0; i < (size_t)w * h; ++i) s->env_f32[i] = f4{rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], 0.0f};
    }
    s->sv.env_w = s->env_f32.empty() ? 0 : w;
    s->sv.env_h = s->env_f32.empty() ? 0 : h;
    s->sv.env_use = use != 0;
    s->sv.env_auto = auto_ != 0;
    memcpy(s->sv.env_color, color, 12);
    s->sv.env_intensity = intensity;
}
void emu_set_camera(emu_scene* s, const float* o, const float* llc, const float* h, const float* v) {
    memcpy(s->sv.cam_origin, o, 12);
    memcpy(s->sv.cam_llc, llc, 12);
    memcpy(s->sv.cam_horizontal, h, 12);
    memcpy(s->sv.cam_vertical, v, 12);
}
void emu_set_render(emu_scene* s, int w, int h, int spp, int depth) {
    s->sv.width = w;
    s->sv.height = h;
    s->sv.max_samples = spp;
    s->sv.max_depth = depth;
}

// same sequence as build_bvh_device (bvh_build.cu), element loops instead of kernels
int emu_build(emu_scene* s) {
    const uint32_t n = (uint32_t)(s->idx4.size() / 4);
    if (n == 0) return -1;
    const float pad = s->max_abs * 3.8146973e-6f + 1e-30f;
    float slo[3], sinv[3];
    for (int a = 0; a < 3; ++a) {
        slo[a] = s->lo[a] - pad;
        const float e = s->hi[a] - s->lo[a] + 2 * pad;
        sinv[a] = e > 0 ? 1.0f / e : 0.0f;
    }
    std::vector<float> tri_lo(3 * (size_t)n), tri_hi(3 * (size_t)n);
    std::vector<uint64_t> keys(n);
    std::vector<uint32_t> order(n);
    for (uint32_t i = 0; i < n; ++i) {
        float lo[3], hi[3];
        tri_box(s->pos.data(), s->idx4.data(), i, pad, lo, hi);
        for (int a = 0; a < 3; ++a) {
            tri_lo[3 * (size_t)i + a] = lo[a];
            tri_hi[3 * (size_t)i + a] = hi[a];
        }
        keys[i] = morton_key(lo, hi, slo, sinv);
        order[i] = i;
    }
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> skeys(n);
    for (uint32_t i = 0; i < n; ++i) skeys[i] = keys[order[i]];
    std::vector<BinNode> bin(n);
    std::vector<CollapsePlan> plans(n);
    uint32_t node_count = 1, tri_count = 0;
    for (int attempt = (s->force_lbvh ? 1 : 0); attempt < 2; ++attempt) {
    int32_t root = ~0;
    if (n >= 2 && attempt == 0) {
        // same rounds as build_bvh_device: nearest neighbour, flags, exclusive scan, emit
        std::vector<int32_t> cid(n), cid2(n), nn(n);
        std::vector<Box6> cbox(n), cbox2(n);
        std::vector<uint64_t> scan(n);
        for (uint32_t i = 0; i < n; ++i) {
            cid[i] = ~(int32_t)i;
            const uint32_t t = order[i];
            for (int a = 0; a < 3; ++a) {
                cbox[i].lo[a] = tri_lo[3 * (size_t)t + a];
                cbox[i].hi[a] = tri_hi[3 * (size_t)t + a];
            }
        }
        uint32_t m = n, node_base = 0;
        while (m > 1) {
            for (uint32_t i = 0; i < m; ++i) ploc_nearest((int)i, (int)m, cbox.data(), nn.data());
            uint64_t acc = 0;
            for (uint32_t i = 0; i < m; ++i) {
                scan[i] = acc;
                acc += ploc_flags((int)i, nn.data());
            }
            for (uint32_t i = 0; i < m; ++i)
                ploc_emit((int)i, nn.data(), scan[i], cid.data(), cbox.data(), cid2.data(), cbox2.data(), bin.data(),
                          plans.data(), node_base, order.data(), tri_lo.data(), tri_hi.data());
            m = (uint32_t)(acc & 0xFFFFFFFFull);
            node_base += (uint32_t)(acc >> 32);
            std::swap(cid, cid2);
            std::swap(cbox, cbox2);
            ++s->ploc_rounds;
        }
        root = cid[0];
        if (n >= 3 && kReinsertPasses > 0) {  // same passes and rounds as build_bvh_device
            const int mc = 2 * (int)n - 1;
            std::vector<int32_t> leaf_parent(n, -1), arrive_flags(n), target(mc), pivot(mc);
            std::vector<float> gain(mc);
            std::vector<uint8_t> state(mc);
            std::vector<unsigned long long> lock(mc);
            const ReinsertTree rt{bin.data(), leaf_parent.data(), order.data(), tri_lo.data(), tri_hi.data(), (int32_t)n, root};
            for (int i = 0; i < (int)n - 1; ++i) ri_link_children(i, rt);
            auto inner_area = [&] {
                double a = 0;
                for (int i = 0; i < (int)n - 1; ++i) a += half_area(bin[i].lo, bin[i].hi);
                return a;
            };
            s->reinsert_moves = 0;
            s->area_clustered = inner_area();
            for (int pass = 0; pass < kReinsertPasses; ++pass) {
                for (int c = 0; c < mc; ++c) ri_find(c, rt, gain.data(), target.data(), pivot.data(), state.data());
                for (int round = 0; round < kReinsertRounds; ++round) {
                    std::fill(lock.begin(), lock.end(), 0ull);
                    if (round > 0)
                        for (int c = 0; c < mc; ++c) ri_hold(c, rt, target.data(), state.data(), lock.data());
                    for (int c = 0; c < mc; ++c) ri_bid(c, rt, gain.data(), target.data(), state.data(), lock.data());
                    for (int c = 0; c < mc; ++c) ri_settle(c, rt, gain.data(), target.data(), state.data(), lock.data());
                }
                std::fill(lock.begin(), lock.end(), 0ull);
                for (int c = 0; c < mc; ++c) ri_mark_moved(c, state.data(), lock.data());
                for (int c = 0; c < mc; ++c) ri_drop_crossing(c, rt, target.data(), pivot.data(), state.data(), lock.data());
                long moved = 0;
                for (int c = 0; c < mc; ++c) {
                    moved += state[c] == kRiWin;
                    ri_apply(c, rt, target.data(), state.data());
                }
                s->reinsert_moves += moved;
                std::fill(arrive_flags.begin(), arrive_flags.end(), 0);
                for (int k = 0; k < (int)n; ++k)
                    refit_from_leaf(k, leaf_parent.data(), bin.data(), arrive_flags.data(), order.data(), tri_lo.data(),
                                    tri_hi.data(), plans.data());
            }
            s->area_reinserted = inner_area();
        }
    }
    if (n >= 2 && attempt == 1) {  // fallback of build_bvh_device: binary radix tree + refit
        std::vector<int32_t> leaf_parent(n, -1), flags(n, 0);
        for (int i = 0; i < (int)n - 1; ++i) radix_tree_node(skeys.data(), (int)n, i, bin.data(), leaf_parent.data());
        for (int k = 0; k < (int)n; ++k)
            refit_from_leaf(k, leaf_parent.data(), bin.data(), flags.data(), order.data(), tri_lo.data(), tri_hi.data(),
                            plans.data());
        root = 0;
    }
    s->nodes.assign(n, Node8{});
    s->tris.assign(n, TriPair{});
    std::vector<CollapseItem> qa(n), qb(n);
    uint32_t n_items = 1;
    node_count = 1;
    tri_count = 0;
    qa[0] = CollapseItem{root, 0u};
    s->levels = 0;
    s->level_end.assign(1, 1u);
    while (n_items) {
        ++s->levels;
        uint32_t next_count = 0;
        for (uint32_t i = 0; i < n_items; ++i)
            collapse_node(qa[i], bin.data(), order.data(), tri_lo.data(), tri_hi.data(), s->pos.data(), s->idx4.data(),
                          s->nodes.data(), &node_count, s->tris.data(), &tri_count, qb.data(), &next_count, (int)n, plans.data());
        n_items = next_count;
        if (n_items) s->level_end.push_back(node_count);
        std::swap(qa, qb);
        if (s->levels > 4 * kStackSize) return -2;
    }
    if (s->levels <= kStackSize) break;
    }
    if (tri_count < (n + 1) / 2 || tri_count > n) return -3;  // pair blocks
    s->tris.resize(tri_count);
    if (s->levels > kStackSize) return -4;
    s->nodes.resize(node_count);
    return (int)node_count;
}

int emu_levels(emu_scene* s) { return s->levels; }
long emu_reinsert_stats(emu_scene* s, double* area_before, double* area_after) {
    *area_before = s->area_clustered;
    *area_after = s->area_reinserted;
    return s->reinsert_moves;
}

// b200pt_update_mesh + b200pt_refit_accel: same per-node function as k_refit_wide, levels bottom-up
int emu_update_mesh(emu_scene* s, int mesh, const float* verts, const float* normals, int nv) {
    if (mesh < 0 || mesh >= (int)s->meshes.size() || nv != s->mesh_nv[mesh]) return -1;
    const int32_t vb = s->mesh_vbase[mesh];
    for (int i = 0; i < nv; ++i) {
        s->pos[vb + i] = f4{verts[3 * i], verts[3 * i + 1], verts[3 * i + 2], 0};
        if (normals) s->nrm[vb + i] = f4{normals[3 * i], normals[3 * i + 1], normals[3 * i + 2], 0};
    }
    s->max_abs = 0;
    for (int a = 0; a < 3; ++a) {
        s->lo[a] = 3e38f;
        s->hi[a] = -3e38f;
    }
    for (const f4& p : s->pos) {
        const float v[3] = {p.x, p.y, p.z};
        for (int a = 0; a < 3; ++a) {
            s->lo[a] = fminf(s->lo[a], v[a]);
            s->hi[a] = fmaxf(s->hi[a], v[a]);
            s->max_abs = fmaxf(s->max_abs, fabsf(v[a]));
        }
    }
    return 0;
}
int emu_refit(emu_scene* s) {
    if (s->level_end.empty() || s->level_end.back() != s->nodes.size()) return -1;
    const float pad = s->max_abs * 3.8146973e-6f + 1e-30f;
    std::vector<float> node_box(6 * s->nodes.size());
    for (int l = (int)s->level_end.size() - 1; l >= 0; --l) {
        const uint32_t first = l ? s->level_end[l - 1] : 0u;
        for (uint32_t i = first; i < s->level_end[l]; ++i)
            refit_wide_node(i, s->nodes.data(), s->tris.data(), s->pos.data(), s->idx4.data(), pad, node_box.data());
    }
    refresh_view(s);
    return 0;
}
uint32_t emu_n_pairs(emu_scene* s) { return (uint32_t)s->tris.size(); }
void emu_force_lbvh(emu_scene* s, int on) { s->force_lbvh = on; }
void emu_set_nee(emu_scene* s, int on) { s->nee = on; }
uint64_t emu_shadow_rays(emu_scene* s) { return s->shadow_rays; }
void emu_debug_pixel(emu_scene* s, int px, int py) {
    s->debug_px = px;
    s->debug_py = py;
}
void emu_walk_counters(emu_scene* s, uint64_t* out) {
    out[0] = s->walk_nodes;
    out[1] = s->walk_tris;
}
uint32_t emu_export(emu_scene* s, void* nodes, void* tris) {
    if (nodes) memcpy(nodes, s->nodes.data(), sizeof(Node8) * s->nodes.size());
    if (tris) memcpy(tris, s->tris.data(), sizeof(TriPair) * s->tris.size());
    return (uint32_t)s->nodes.size();
}
// walk a tree handed in from outside (e.g. exported from the GPU) instead of the emulated build
void emu_import(emu_scene* s, const void* nodes, uint32_t n_nodes, const void* tris, uint32_t n_tris) {
    s->nodes.assign((const Node8*)nodes, (const Node8*)nodes + n_nodes);
    s->tris.assign((const TriPair*)tris, (const TriPair*)tris + n_tris);
}

void emu_trace(emu_scene* s, const float* o, const float* d, int n, float tmin, float tmax, emu_hit* out,
               uint64_t* counters) {
    refresh_view(s);
    WalkCounters wc{0, 0};
    uint64_t nodes = 0, tris = 0;
    for (int i = 0; i < n; ++i) {
        Hit h;
        wc = WalkCounters{0, 0};
        walk_closest<true>(s->nodes.data(), s->tris.data(), f3{o[3 * i], o[3 * i + 1], o[3 * i + 2]},
                           f3{d[3 * i], d[3 * i + 1], d[3 * i + 2]}, tmin, tmax, h, &wc);
        nodes += wc.nodes;
        tris += wc.tris;
        out[i] = to_rec(s, h, tmax);
    }
    if (counters) {
        counters[0] = nodes;
        counters[1] = tris;
    }
}

void emu_primary_hits(emu_scene* s, int nthreads, emu_hit* out) {
    refresh_view(s);
    SceneView sv = s->sv;
    parallel_rows(sv.height, nthreads, [&](int py) {
        for (int px = 0; px < sv.width; ++px) {
            uint32_t rng = rng_seed((uint32_t)px, (uint32_t)py);
            f3 o, d;
            camera_ray(sv, px, py, rng, o, d);
            Hit h;
            walk_closest<false>(sv.nodes, sv.tris, o, d, kTMin, kTMax, h, nullptr);
            out[(size_t)px + (size_t)sv.width * py] = to_rec(s, h, kTMax);
        }
    });
}

// The wavefront of api.cu/wavefront.cu run lane by lane: init, then extend + shade until
// the lane retires.  Lanes are independent, so running each to completion gives the same
// per-lane sequence as the GPU's interleaving.  sum_out: W*H*4 floats indexed x + W*y.
uint64_t emu_render(emu_scene* s, int rank, int world, int shard_mode, int nthreads, float* sum_out, float* mean_out,
                    uint32_t* rgba_out) {
    refresh_view(s);
    SceneView sv = s->sv;
    const int W = sv.width, H = sv.height, total_spp = sv.max_samples;
    sv.row_first = 0;
    sv.row_step = 1;
    sv.seed_row_offset = 0;
    sv.n_lanes = W * H;
    if (world > 1) {
        if (shard_mode == 1) {
            sv.row_first = rank;
            sv.row_step = world;
            sv.n_lanes = W * (H > rank ? (H - rank + world - 1) / world : 0);
        } else {
            sv.max_samples = total_spp / world + (rank < total_spp % world ? 1 : 0);
            sv.seed_row_offset = rank * H;
        }
    }
    memset(sum_out, 0, sizeof(float) * 4 * (size_t)W * H);
    std::atomic<uint64_t> rays{0};
    std::atomic<uint64_t> cnt_nodes{0}, cnt_tris{0}, cnt_shadow{0};
    const int rows = sv.n_lanes / W;
    parallel_rows(rows, nthreads, [&](int row) {
        uint64_t nr = 0, nsh = 0;
        WalkCounters wc{0, 0};
        uint64_t wn = 0, wt = 0;
        for (int c = 0; c < W; ++c) {
            const uint32_t lane = (uint32_t)row * W + c;
            int px, py;
            lane_pixel(sv, lane, px, py);
            LaneState L{};
            L.max_samples = (uint32_t)sv.max_samples;
            L.rng = rng_seed((uint32_t)px, (uint32_t)(py + sv.seed_row_offset));
            camera_ray(sv, px, py, L.rng, L.o, L.d);
            L.thr = mk3(1.0f);
            L.sample = 0;
            L.depth = 0;
            L.lobe = LOBE_NONE;
            f3 sum = mk3(0.0f), gathered = mk3(0.0f);
            float second = 0.0f;  // sum.w of the kernels: sum of squared sample luminances
            L.prev_pdf = -1.0f;
            if (sv.max_samples > 0 && sv.max_depth > 0) {
                for (;;) {
                    // state goes through the same packing as the device arrays
                    const uint32_t ctl = pack_ctl(L.sample, L.depth, L.lobe);
                    L.sample = ctl_sample(ctl);
                    L.depth = ctl_depth(ctl);
                    L.lobe = ctl_lobe(ctl);
                    Hit h;
                    wc = WalkCounters{0, 0};
                    walk_closest<true>(sv.nodes, sv.tris, L.o, L.d, kTMin, kTMax, h, &wc);
                    wn += wc.nodes;
                    wt += wc.tris;
                    ++nr;
                    if (s->debug_px == px && s->debug_py == py)
                        printf("DBG %.9g %.9g %.9g %.9g %.9g %.9g %.9g %u\n", L.o.x, L.o.y, L.o.z, L.d.x, L.d.y, L.d.z, h.t, h.tri);
                    const ShadeResult r = sv.nee ? shade_lane<true>(sv, L, h, px, py) : shade_lane<false>(sv, L, h, px, py);
                    if (sv.nee) {
                        // k_shadow: the light sample counts when nothing lies between the vertex and the
                        // light; k_shade / k_shadow: a finished sample is gathered radiance + last term
                        if (L.sh_valid) {
                            Hit blocker;
                            walk_closest<false>(sv.nodes, sv.tris, L.sh_o, L.sh_d, kTMin, L.sh_tmax, blocker, nullptr);
                            ++nsh;
                            if (blocker.tri == kNoHit) gathered = gathered + L.sh_c;
                        }
                        if (L.sample_ended) {
                            sum = sum + (gathered + L.contrib);
                            second += sample_lum2(gathered + L.contrib);
                            gathered = mk3(0.0f);
                        }
                    } else if (L.sample_ended) {
                        sum = sum + L.contrib;
                        second += sample_lum2(L.contrib);
                    }
                    if (r == LANE_DONE) break;
                }
            }
            float* o = sum_out + 4 * ((size_t)px + (size_t)W * py);
            o[0] = sum.x;
            o[1] = sum.y;
            o[2] = sum.z;
            o[3] = second;
        }
        rays += nr;
        cnt_shadow += nsh;
        cnt_nodes += wn;
        cnt_tris += wt;
    });
    s->shadow_rays = cnt_shadow.load();
    s->walk_nodes = cnt_nodes.load();
    s->walk_tris = cnt_tris.load();
    if (mean_out || rgba_out) {
        const float inv = 1.0f / static_cast<float>(total_spp);
        for (int py = 0; py < H; ++py)
            for (int px = 0; px < W; ++px) {
                const float* sp = sum_out + 4 * ((size_t)px + (size_t)W * py);
                const f3 c = f3{sp[0], sp[1], sp[2]} * inv;
                const size_t ofs = (size_t)px + (size_t)W * (H - 1 - py);
                if (rgba_out) rgba_out[ofs] = to_rgba(c);
                if (mean_out) {
                    mean_out[4 * ofs] = c.x;
                    mean_out[4 * ofs + 1] = c.y;
                    mean_out[4 * ofs + 2] = c.z;
                    mean_out[4 * ofs + 3] = 1.0f;
                }
            }
    }
    return rays.load();
}

void emu_sample_disney(const float* mats, int n, const int32_t* material, const float* wo, const int32_t* lobe_in,
                       const uint32_t* rng_in, float* f, float* wi, float* pdf, int32_t* lobe_out, uint32_t* rng_out) {
    for (int i = 0; i < n; ++i) {
        Material m{};
        memcpy(&m, mats + 17 * material[i], 68);
        uint32_t rng = rng_in[i];
        int lobe = lobe_in[i];
        BsdfSample s;
        s.wi = mk3(0.0f);
        s.pdf = 0.0f;
        s.f = mk3(0.0f);
        sample_disney(m, f3{wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]}, rng, lobe, s);
        f[3 * i] = s.f.x; f[3 * i + 1] = s.f.y; f[3 * i + 2] = s.f.z;
        wi[3 * i] = s.wi.x; wi[3 * i + 1] = s.wi.y; wi[3 * i + 2] = s.wi.z;
        pdf[i] = s.pdf;
        lobe_out[i] = lobe;
        rng_out[i] = rng;
    }
}

// detmath hook (fn ids as oracle_math)
void emu_math(int fn, int n, const float* x, const float* y, float* out) {
    for (int i = 0; i < n; ++i) {
        switch (fn) {
            case 0: out[i] = dm_sin(x[i]); break;
            case 1: out[i] = dm_cos(x[i]); break;
            case 2: out[i] = dm_tan(x[i]); break;
            case 3: out[i] = dm_atan(x[i]); break;
            case 4: out[i] = dm_asin(x[i]); break;
            case 5: out[i] = dm_log(x[i]); break;
            case 6: out[i] = dm_atan2(x[i], y[i]); break;
            case 7: out[i] = dm_pow(x[i], y[i]); break;
            default: out[i] = 0;
        }
    }
}

}  // extern "C"

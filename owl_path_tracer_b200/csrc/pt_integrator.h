// pt_integrator.h — per-lane logic of the wavefront: camera ray generation and the
// shade step (one path vertex).  Together with walk_closest (pt_bvh8.h) this restates
// __raygen__ray_gen and trace_path of the reference (device.cu:113-254) as a state
// machine: a lane is one pixel, it owns the pixel's RNG stream and walks through that
// pixel's samples one after the other, exactly in the order the reference's per-pixel
// thread does.  Regenerating the next camera ray inside the shade step is
// what keeps every lane busy until its pixel is finished.
#pragma once
#include "pt_bvh8.h"
#include "pt_disney.h"

namespace pt {

// read-only 16-byte-granular struct fetch: LDG.128s through the non-coherent path on the device
// (a plain struct copy compiles to one 4-byte load per member)
template <class T>
PT_HD T load_ro(const T* p) {
#if defined(__CUDA_ARCH__)
    static_assert(sizeof(T) % 16 == 0, "load_ro: size must be a multiple of 16 bytes");
    T r;
    const uint4* src = reinterpret_cast<const uint4*>(p);
    uint4* dst = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); ++i) dst[i] = __ldg(src + i);
    return r;
#else
    return *p;
#endif
}

struct alignas(16) TriIdx {   // per global triangle: 3 global vertex ids + mesh index
    int32_t i0, i1, i2, mesh;
};

struct alignas(16) MeshInfo {
    int32_t first_tri;    // global id of this mesh's primitive 0
    int32_t material;     // material_index (entity_data, device_global.hpp:38-44)
    int32_t tex_offset;   // first texel in the texture pool, -1 = no texture
    int32_t tex_w, tex_h;
    int32_t has_uv;
    int32_t light_first;  // index of this mesh's primitive 0 in SceneView::lights, -1 = not an emitter
    int32_t pad1;
};

// one emissive triangle of the light table (next-event estimation mode): chosen with probability
// `prob` (proportional to area * emission), sampled uniformly by area (lights.hpp:30-48)
struct alignas(16) LightTri {
    uint32_t tri;      // global triangle id
    float area;        // triangle_area, lights.hpp:15-20
    float prob;        // cdf[i] - cdf[i-1]
    float emission;    // material_data::emission of its mesh
};

struct SceneView {
    const Node8* nodes;
    const TriPair* tris;
    const int32_t* idx4;       // per global triangle: 3 global vertex ids + mesh index
    const f4* vpos;            // per global vertex
    const f4* vnrm;
    const f2* vuv;
    // Per-triangle shading records (device only; nullptr: fetch through idx4 as the host emulation does): the three
    // positions and normals of a triangle side by side, 96 bytes, p0.w = mesh index.  One gather instead of two
    // dependent ones (idx4, then six scattered vertices) in the shade step; same values, same arithmetic.
    const struct TriShade* tri_rec;
    const MeshInfo* meshes;
    const Material* materials;
    const Material* mesh_materials;  // materials[meshes[m].material] per mesh: one dependent load less in shade
    const uint32_t* texels;    // RGBA8 pool (mesh textures)
    const uint32_t* env;       // RGBA8 environment map or nullptr
    const f4* env_f32;         // float RGB environment map (b200pt_set_environment_f32) or nullptr; same addressing
    int32_t env_w, env_h;
    int32_t env_use, env_auto;
    float env_color[3];
    float env_intensity;
    float cam_origin[3], cam_llc[3], cam_horizontal[3], cam_vertical[3];
    int32_t width, height;     // full frame
    int32_t max_samples;       // samples a stream of this rank renders per pixel (streams < stream_extra: one more)
    int32_t max_depth;
    int32_t n_materials;
    // lane -> pixel:  x = lane % width, y = row_first + row_step * (lane / width)
    int32_t row_first, row_step;
    int32_t seed_row_offset;   // added to y when seeding (per-rank streams: first stream of the rank * height)
    int32_t n_lanes;
    // Several RNG streams per pixel on one GPU (b200pt_config.streams_per_pixel = V > 1, SPP sharding): the rank's
    // lanes are V copies of its pixels, lane = v * n_pix_lanes + pixel lane; copy v is stream g = rank * V + v of
    // the world_size * V streams the samples are split over — the very streams an (world_size * V)-GPU job uses.
    int32_t n_pix_lanes;       // pixel lanes of the rank; 0: one stream (lane = pixel lane)
    int32_t stream_extra;      // the first stream_extra streams of the rank render max_samples + 1 samples
    // progressive passes (b200pt_render_pass): this launch renders samples first_sample .. max_samples-1
    // of every pixel; first_sample > 0 continues the pixel's RNG stream (LaneArrays::rng_save) and its sum
    int32_t first_sample;
    // next-event estimation (B200PT_FLAG_NEE); n_lights == 0 or nee == 0: the reference's estimator verbatim
    const LightTri* lights;
    const float* light_cdf;    // n_lights ascending values, last = 1
    int32_t n_lights;
    int32_t nee;
    // importance sampling of the environment map in the NEE mode (pt_lights.h build_env_table): the texel is
    // chosen with probability proportional to its luminance, rows first (marginal), then the column of the row
    const float* env_cdf_rows;  // env_h ascending values, last = 1
    const float* env_cdf_cols;  // env_h x env_w, each row ascending, last = 1
    int32_t env_nee;            // 1: NEE mode, map in use, tables present
    uint32_t plane_magic;      // 0x3F800000, see plane_q (pt_bvh8.h): a runtime value on purpose
};

struct TriShade {
    f4 p0, p1, p2, n0, n1, n2;   // p0.w: mesh index (bits)
};
static_assert(sizeof(TriShade) == 96, "TriShade is 96 bytes");

// packed per-lane control word: sample index (20 bits) | depth (8 bits) | lobe+1 (3 bits)
PT_HD uint32_t pack_ctl(uint32_t sample, uint32_t depth, int lobe) {
    return (sample & 0xFFFFFu) | ((depth & 0xFFu) << 20) | (((uint32_t)(lobe + 1) & 7u) << 28);
}
PT_HD uint32_t ctl_sample(uint32_t c) { return c & 0xFFFFFu; }
PT_HD uint32_t ctl_depth(uint32_t c) { return (c >> 20) & 0xFFu; }
PT_HD int ctl_lobe(uint32_t c) { return (int)((c >> 28) & 7u) - 1; }

PT_HD uint32_t lane_stream(const SceneView& sv, uint32_t lane) {
    return sv.n_pix_lanes > 0 ? lane / (uint32_t)sv.n_pix_lanes : 0u;
}
PT_HD uint32_t lane_samples(const SceneView& sv, uint32_t lane) {
    return (uint32_t)sv.max_samples + (lane_stream(sv, lane) < (uint32_t)sv.stream_extra ? 1u : 0u);
}
PT_HD void lane_pixel(const SceneView& sv, uint32_t lane, int& x, int& y) {
    if (sv.n_pix_lanes > 0) lane %= (uint32_t)sv.n_pix_lanes;
    x = (int)(lane % (uint32_t)sv.width);
    y = sv.row_first + sv.row_step * (int)(lane / (uint32_t)sv.width);
}

// device.cu:231-241: jitter (x draw first), screen = (pixel + rand) / fb_size,
// direction = normalize(llc + u*horizontal + v*vertical - origin)
PT_HD void camera_ray(const SceneView& sv, int px, int py, uint32_t& rng, f3& o, f3& d) {
    const float rx = rng_next(rng);
    const float ry = rng_next(rng);
    const float su = ((float)px + rx) / (float)sv.width;
    const float sv_ = ((float)py + ry) / (float)sv.height;
    const f3 org{sv.cam_origin[0], sv.cam_origin[1], sv.cam_origin[2]};
    const f3 llc{sv.cam_llc[0], sv.cam_llc[1], sv.cam_llc[2]};
    const f3 hor{sv.cam_horizontal[0], sv.cam_horizontal[1], sv.cam_horizontal[2]};
    const f3 ver{sv.cam_vertical[0], sv.cam_vertical[1], sv.cam_vertical[2]};
    o = org;
    d = normalize(llc + su * hor + sv_ * ver - org);
}

// device.cu:23-39 + owl.hpp:248-257: RGBA8, point filter, clamp, normalised coords
PT_HD f3 fetch_rgba8(const uint32_t* img, int w, int h, float u, float v) {
    float fx = floorf(u * (float)w), fy = floorf(v * (float)h);
    int ix = (fx >= 0.0f) ? (fx < (float)w ? (int)fx : w - 1) : 0;  // NaN -> 0
    int iy = (fy >= 0.0f) ? (fy < (float)h ? (int)fy : h - 1) : 0;
    const uint32_t p = img[(size_t)iy * (size_t)w + (size_t)ix];
    return f3{(float)(p & 0xFFu) / 255.0f, (float)((p >> 8) & 0xFFu) / 255.0f,
              (float)((p >> 16) & 0xFFu) / 255.0f};
}

// float RGB map: same point sampling and clamping as fetch_rgba8, radiance not limited to [0, 1]
PT_HD f3 fetch_f32(const f4* img, int w, int h, float u, float v) {
    float fx = floorf(u * (float)w), fy = floorf(v * (float)h);
    int ix = (fx >= 0.0f) ? (fx < (float)w ? (int)fx : w - 1) : 0;
    int iy = (fy >= 0.0f) ? (fy < (float)h ? (int)fy : h - 1) : 0;
    const f4 p = img[(size_t)iy * (size_t)w + (size_t)ix];
    return f3{p.x, p.y, p.z};
}

// miss branch of trace_path, device.cu:136-148
PT_HD f3 miss_radiance(const SceneView& sv, f3 dir) {
    f3 rad = mk3(0.0f);
    if (sv.env_use && (sv.env || sv.env_f32)) {
        const float u = 0.5f + dm_atan2(dir.x, dir.z) / (2.0f * kPi);
        const float v = 0.5f + dm_asin(dir.y) / kPi;
        rad = rad + (sv.env_f32 ? fetch_f32(sv.env_f32, sv.env_w, sv.env_h, u, v) : fetch_rgba8(sv.env, sv.env_w, sv.env_h, u, v));
    } else if (sv.env_auto) {
        rad = rad + lerp3(mk3(1.0f), f3{0.5f, 0.7f, 1.0f}, 0.5f * (dir.y + 1.0f));
    } else {
        rad = rad + f3{sv.env_color[0], sv.env_color[1], sv.env_color[2]};
    }
    return rad * sv.env_intensity;
}

// NEE mode, environment map: texel (ix, iy) that miss_radiance reads for `dir`
PT_HD void env_texel(const SceneView& sv, f3 dir, int& ix, int& iy) {
    const float u = 0.5f + dm_atan2(dir.x, dir.z) / (2.0f * kPi);
    const float v = 0.5f + dm_asin(dir.y) / kPi;
    const float fx = floorf(u * (float)sv.env_w), fy = floorf(v * (float)sv.env_h);  // fetch_rgba8's addressing
    ix = (fx >= 0.0f) ? (fx < (float)sv.env_w ? (int)fx : sv.env_w - 1) : 0;
    iy = (fy >= 0.0f) ? (fy < (float)sv.env_h ? (int)fy : sv.env_h - 1) : 0;
}
// solid-angle density with which the environment sampler produces `dir`: the texel's probability spread over
// the texel's solid angle  (du dv = dlon dlat / 2 pi^2,  d omega = cos(lat) dlon dlat,  cos(lat) = sqrt(1 - y^2))
PT_HD float env_pdf(const SceneView& sv, f3 dir) {
    int ix, iy;
    env_texel(sv, dir, ix, iy);
    const float prow = sv.env_cdf_rows[iy] - (iy ? sv.env_cdf_rows[iy - 1] : 0.0f);
    const float* row = sv.env_cdf_cols + (size_t)iy * (size_t)sv.env_w;
    const float pcol = row[ix] - (ix ? row[ix - 1] : 0.0f);
    const float cl = sqrtf(fmaxf(0.0f, 1.0f - dir.y * dir.y));
    if (cl < 1e-4f) return 0.0f;
    return prow * pcol * ((float)sv.env_w * (float)sv.env_h) / (2.0f * kPi * kPi * cl);
}
PT_HD int cdf_find(const float* cdf, int n, float u) {  // first entry with cdf >= u
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (u <= cdf[mid]) hi = mid;
        else lo = mid + 1;
    }
    return lo;
}

// state of one lane as the shade step sees it
struct LaneState {
    f3 o, d;          // current ray
    f3 thr;           // throughput
    f3 contrib;       // value of the sample that just ended (valid when sample_ended)
    bool sample_ended;
    uint32_t rng;
    uint32_t sample;  // index of the sample in flight
    uint32_t max_samples;  // samples this lane renders (lane_samples)
    uint32_t depth;   // loop counter of trace_path
    int lobe;         // sampled_lobe, persists across bounces (device.cu:127)
    // next-event estimation mode only
    float prev_pdf;   // density with which the BSDF generated the current ray; <= 0: count an emitter fully
    int sh_valid;     // a shadow ray is requested: from sh_o along sh_d up to sh_tmax; if nothing is hit,
    f3 sh_o, sh_d;    //   sh_c is added to the radiance of the sample this vertex belongs to
    float sh_tmax;
    f3 sh_c;
};

enum ShadeResult { LANE_CONTINUE = 0, LANE_DONE = 1 };

PT_HD bool bad3(f3 v) {
    return isinf(v.x) || isinf(v.y) || isinf(v.z) || isnan(v.x) || isnan(v.y) || isnan(v.z);
}

// Second moment of the sample values, kept in sum.w: square of the Rec.709 luminance of a finished sample.  The
// per-pixel variance estimate of the progressive mode (b200pt_pass_info::rel_variance) is built from it.
PT_HD float sample_lum2(f3 value) {
    const float l = lum709(value);
    return l * l;
}

// finish the sample in flight with `contrib`, then start the next one or retire the lane
PT_HD ShadeResult end_sample(const SceneView& sv, LaneState& L, int px, int py, f3 contrib) {
    L.contrib = contrib;  // caller does color += trace_path(...), device.cu:243
    L.sample_ended = true;
    L.sample += 1;
    if (L.sample >= L.max_samples) return LANE_DONE;
    camera_ray(sv, px, py, L.rng, L.o, L.d);
    L.thr = mk3(1.0f);
    L.depth = 0;
    L.lobe = LOBE_NONE;
    L.prev_pdf = -1.0f;
    return LANE_CONTINUE;
}

// pdf_area_to_solid_angle, lights.hpp:23-28
PT_HD float pdf_area_to_solid_angle(float pdf_area, float dist_sqr, float cos_theta) {
    const float a = fabsf(cos_theta);
    if (a < 1e-4f) return 0.0f;
    return pdf_area * dist_sqr / a;
}

// One iteration of the loop body of trace_path (device.cu:130-215) for a lane whose
// ray has just been traced.  Returns LANE_DONE when the pixel's last sample ended.
template <bool NEE>
PT_HD ShadeResult shade_lane(const SceneView& sv, LaneState& L, const Hit& hit, int px, int py) {
    L.sample_ended = false;
    L.sh_valid = 0;
    // NEE: compiled for B200PT_FLAG_NEE contexts only.  Light sampling draws from the emissive triangles, from
    // the environment map, or (one extra draw) from either with probability 1/2 when the scene has both.
    const bool tri_l = NEE && sv.n_lights > 0, env_l = NEE && sv.env_nee != 0;
    const bool nee = tri_l || env_l;
    const float p_sel = (tri_l && env_l) ? 0.5f : 1.0f;
    if (hit.tri == kNoHit) {
        f3 rad = miss_radiance(sv, L.d);
        if (env_l && L.prev_pdf > 0.0f)   // the map could have been reached by light sampling at the previous vertex
            rad = rad * power_heuristic(L.prev_pdf, p_sel * env_pdf(sv, L.d));
        return end_sample(sv, L, px, py, rad * L.thr);
    }
    int i0 = 0, i1 = 0, i2 = 0, mesh;
    f4 p0, p1, p2, n0, n1, n2;
    if (sv.tri_rec) {
        const TriShade r = load_ro(sv.tri_rec + hit.tri);
        p0 = r.p0; p1 = r.p1; p2 = r.p2;
        n0 = r.n0; n1 = r.n1; n2 = r.n2;
        mesh = (int)f2u(r.p0.w);
    } else {
        const TriIdx ix = load_ro(reinterpret_cast<const TriIdx*>(sv.idx4) + hit.tri);
        i0 = ix.i0; i1 = ix.i1; i2 = ix.i2;
        mesh = ix.mesh;
        p0 = load_ro(sv.vpos + i0); p1 = load_ro(sv.vpos + i1); p2 = load_ro(sv.vpos + i2);
        n0 = load_ro(sv.vnrm + i0); n1 = load_ro(sv.vnrm + i1); n2 = load_ro(sv.vnrm + i2);
    }
    // mesh record and its material (entity_data + material_buffer[material_index], device.cu:150-154)
    const MeshInfo mi = load_ro(sv.meshes + mesh);
    Material m = load_ro(sv.mesh_materials + mesh);

    if (m.emission > 0.0f) {  // device.cu:157-161: radiance = emission (overwrites)
        if (tri_l && L.prev_pdf > 0.0f && mi.light_first >= 0) {
            // the same emitter could have been reached by light sampling at the previous vertex:
            // power heuristic between the BSDF density and the light density of this very point
            const LightTri lt = load_ro(sv.lights + (mi.light_first + ((int)hit.tri - mi.first_tri)));
            // area -> solid-angle density with the GEOMETRIC unit normal of the emitter (the reference's unused
            // sample_triangle, lights.hpp:30-48, takes the interpolated shading normal: biased for smooth emitters)
            const f3 nl = normalize(cross(f3{p1.x, p1.y, p1.z} - f3{p0.x, p0.y, p0.z}, f3{p2.x, p2.y, p2.z} - f3{p0.x, p0.y, p0.z}));
            const float pdf_light = p_sel * (lt.prob * pdf_area_to_solid_angle(1.0f / lt.area, hit.t * hit.t, dot(-L.d, nl)));
            const float w = power_heuristic(L.prev_pdf, pdf_light);
            return end_sample(sv, L, px, py, mk3(m.emission * w) * L.thr);
        }
        return end_sample(sv, L, px, py, mk3(m.emission) * L.thr);
    }

    const float bu = hit.u, bv = hit.v;
    const float bw = 1.0f - bu - bv;
    const f3 vp = bw * f3{p0.x, p0.y, p0.z} + bu * f3{p1.x, p1.y, p1.z} + bv * f3{p2.x, p2.y, p2.z};
    const f3 vn = normalize(bw * f3{n0.x, n0.y, n0.z} + bu * f3{n1.x, n1.y, n1.z} + bv * f3{n2.x, n2.y, n2.z});
    if (mi.tex_offset >= 0) {  // device.cu:75-94
        if (sv.tri_rec) {  // the records carry no uv indices: textured meshes take the idx4 route for them
            const TriIdx ix = load_ro(reinterpret_cast<const TriIdx*>(sv.idx4) + hit.tri);
            i0 = ix.i0; i1 = ix.i1; i2 = ix.i2;
        }
        const f2 t0 = sv.vuv[i0], t1 = sv.vuv[i1], t2 = sv.vuv[i2];
        const float tu = bw * t0.x + bu * t1.x + bv * t2.x;
        const float tv = bw * t0.y + bu * t1.y + bv * t2.y;
        const f3 c = fetch_rgba8(sv.texels + mi.tex_offset, mi.tex_w, mi.tex_h, tu, tv);
        m.base_color[0] = c.x;
        m.base_color[1] = c.y;
        m.base_color[2] = c.z;
    }

    // device.cu:176-181 and math.hpp:86-107
    const f3 wo = -normalize(L.d);
    f3 T;
    if (vn.x != vn.y || vn.x != vn.z)
        T = f3{vn.z - vn.y, vn.x - vn.z, vn.y - vn.x};
    else
        T = f3{vn.z - vn.y, vn.x + vn.z, -vn.y - vn.x};
    T = normalize(T);
    const f3 B = cross(vn, T);
    const f3 lwo = normalize(f3{dot(wo, T), dot(wo, B), dot(wo, vn)});

    if (nee) {
        bool use_env = env_l;
        if (tri_l && env_l) use_env = rng_next(L.rng) < 0.5f;
        f3 dir = mk3(0.0f), emitted = mk3(0.0f);
        float pdf_light = 0.0f, tmax = 0.0f;
        if (use_env) {
            // one texel of the map by luminance, a uniform point in it; density and radiance are re-read through
            // the direction -> texel mapping of the miss branch, so both strategies see the same function
            const float u0 = rng_next(L.rng), u1 = rng_next(L.rng), u2 = rng_next(L.rng), u3 = rng_next(L.rng);
            const int iy = cdf_find(sv.env_cdf_rows, sv.env_h, u0);
            const int ix = cdf_find(sv.env_cdf_cols + (size_t)iy * (size_t)sv.env_w, sv.env_w, u1);
            const float lon = (((float)ix + u2) / (float)sv.env_w - 0.5f) * (2.0f * kPi);
            const float lat = (((float)iy + u3) / (float)sv.env_h - 0.5f) * kPi;
            const float cl = dm_cos(lat);
            dir = normalize(f3{dm_sin(lon) * cl, dm_sin(lat), dm_cos(lon) * cl});
            pdf_light = p_sel * env_pdf(sv, dir);
            emitted = miss_radiance(sv, dir);
            tmax = kTMax;
        } else {
            // one emissive triangle by power, one point on it by area
            // (sample_triangle, lights.hpp:30-48; uniform_sample_triangle, sample_methods.hpp:82-86)
            const float us = rng_next(L.rng), u0 = rng_next(L.rng), u1 = rng_next(L.rng);
            const int lo = cdf_find(sv.light_cdf, sv.n_lights, us);
            const LightTri lt = load_ro(sv.lights + lo);
            const TriIdx lx = load_ro(reinterpret_cast<const TriIdx*>(sv.idx4) + lt.tri);
            const f4 q0 = load_ro(sv.vpos + lx.i0), q1 = load_ro(sv.vpos + lx.i1), q2 = load_ro(sv.vpos + lx.i2);
            const float su0 = sqrtf(u0);
            const float bx = 1.0f - su0, by = u1 * su0;
            const float bz = 1.0f - bx - by;
            const f3 pos = bz * f3{q0.x, q0.y, q0.z} + bx * f3{q1.x, q1.y, q1.z} + by * f3{q2.x, q2.y, q2.z};
            const f3 nl = normalize(cross(f3{q1.x, q1.y, q1.z} - f3{q0.x, q0.y, q0.z}, f3{q2.x, q2.y, q2.z} - f3{q0.x, q0.y, q0.z}));
            dir = pos - vp;
            const float dist2 = dot(dir, dir);
            const float dist = sqrtf(dist2);
            dir = dir / dist;
            pdf_light = p_sel * (lt.prob * pdf_area_to_solid_angle(1.0f / lt.area, dist2, dot(-dir, nl)));
            if (!(dist > 2.0f * kTMin)) pdf_light = 0.0f;
            emitted = mk3(lt.emission);
            tmax = dist - kTMin;   // visibiliy_test, device.cu:97-110
        }
        if (pdf_light > 0.0f) {
            const f3 lwi = normalize(f3{dot(dir, T), dot(dir, B), dot(dir, vn)});
            f3 f_eff;
            float pdf_eff;
            eval_mixture(m, lwo, lwi, f_eff, pdf_eff);
            if (pdf_eff > 0.0f) {
                const float w = power_heuristic(pdf_light, pdf_eff);
                const f3 c = L.thr * (f_eff * (emitted * (fabsf(lwi.z) * w / pdf_light)));
                if (!bad3(c) && (c.x != 0.0f || c.y != 0.0f || c.z != 0.0f)) {
                    L.sh_valid = 1;
                    L.sh_o = vp;
                    L.sh_d = dir;
                    L.sh_tmax = tmax;
                    L.sh_c = c;
                }
            }
        }
    }

    // The reference re-traces the same ray when f is inf/nan (--depth; continue,
    // device.cu:196-201).  The trace is deterministic, so the retry only redraws the
    // BSDF sample; we loop here instead of going round the wavefront again.
    for (int attempt = 0; attempt < 64; ++attempt) {
        BsdfSample s;
        s.wi = mk3(0.0f);
        s.pdf = 0.0f;
        s.f = mk3(0.0f);
        sample_disney(m, lwo, L.rng, L.lobe, s);
        const f3 wi = normalize(s.wi.x * T + s.wi.y * B + s.wi.z * vn);
        if (s.pdf < 1E-5f) return end_sample(sv, L, px, py, mk3(0.0f) * L.thr);
        if (bad3(s.f)) continue;
        L.thr = L.thr * (s.f * fabsf(s.wi.z) / s.pdf);
        L.o = vp;
        L.d = wi;
        if (nee) {  // density of this direction under BSDF sampling, for the emitter it may reach
            f3 fe;
            float pe;
            eval_mixture(m, lwo, s.wi, fe, pe);
            L.prev_pdf = L.lobe == LOBE_GLASS ? -1.0f : pe;
        }
        const float bmax = fmaxf(L.thr.x, fmaxf(L.thr.y, L.thr.z));
        if (L.lobe != LOBE_GLASS && L.depth > 3) {  // inverted, uncompensated RR (Q14)
            const float q = fmaxf(.05f, 1.0f - bmax);
            if (rng_next(L.rng) > q) return end_sample(sv, L, px, py, mk3(0.0f) * L.thr);
        }
        L.depth += 1;
        if (L.depth >= (uint32_t)sv.max_depth) return end_sample(sv, L, px, py, mk3(0.0f) * L.thr);
        return LANE_CONTINUE;
    }
    return end_sample(sv, L, px, py, mk3(0.0f) * L.thr);
}

// make_8bit / make_rgba, owl_device.h:79-91.  int(f*256) of NaN is 0 on the device
// (cvt.rzi.s32.f32), made explicit here so host and device agree.
PT_HD uint32_t to_8bit(float f) {
    const float s = f * 256.f;
    if (!(s == s)) return 0u;
    if (s >= 255.0f) return 255u;
    if (s <= 0.0f) return 0u;
    return (uint32_t)(int)s;
}
// Output transfer of the resolve step.  0: the reference as shipped (linear values straight into make_rgba).
// 1: the line the reference carries commented out, device.cu:248: color = saturate(pow(color, 1 / 2.2)).
PT_HD f3 output_transfer(f3 c, int transfer) {
    if (transfer == 1) {
        const float g = 1.0f / 2.2f;
        c = f3{fminf(fmaxf(dm_pow(c.x, g), 0.0f), 1.0f), fminf(fmaxf(dm_pow(c.y, g), 0.0f), 1.0f),
               fminf(fmaxf(dm_pow(c.z, g), 0.0f), 1.0f)};
    }
    return c;
}
PT_HD uint32_t to_rgba(f3 c) {
    return (to_8bit(c.x) << 0) + (to_8bit(c.y) << 8) + (to_8bit(c.z) << 16) + (0xffU << 24);
}

}  // namespace pt

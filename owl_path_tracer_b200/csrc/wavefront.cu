// wavefront.cu — the kernels of the render loop (sm_100a).
//
//   k_init_lanes  raygen for sample 0 of every pixel            (device.cu:220-241)
//   k_extend      closest hit of every queued ray               (replaces optixTrace, device.cu:133)
//   k_shade       one path vertex: miss / emission / Disney sample / RR, next camera
//                 ray when a sample ends, warp-ballot compaction of the surviving lanes
//                                                              (device.cu:136-215, 229-243)
//   k_resolve     mean, RGBA8 pack, row flip                    (device.cu:246-253)
// plus small parity-hook kernels.  Compiled with -fmad=false: the arithmetic must round
// like the scalar oracle's (pt_defs.h); the slab test uses explicit fmaf.
#include <algorithm>

#include "device_ctx.h"
#include "pt_walk_warp.cuh"

namespace pt {

#ifndef PT_BLOCKS_PER_SM
#define PT_BLOCKS_PER_SM 6     // persistent blocks of 128 threads per SM for the walker kernels
#endif
#ifndef PT_EXTEND_MIN_BLOCKS
#define PT_EXTEND_MIN_BLOCKS 6  // register cap of the walker kernels: 65536 / (128 * blocks) -> 85; the pooled triangle phase
                                // holds a borrowed ray next to the lane's own and takes 80.  At 7 blocks (72 registers) it spills
                                // 27 registers and loses what pooling gains; the lane-per-block walker was indifferent between
                                // 6 and 7 blocks and 10 % slower at 8 (64 registers) -- profiles/r02s_*
#endif
#ifndef PT_SHADE_MIN_BLOCKS
#define PT_SHADE_MIN_BLOCKS 5   // register cap of k_shade: 65536 / (128 * blocks) = 102.  Round 1 ran at 4 (128 registers): the
                                // inlined binary64 pow / log bodies needed them.  With those out of line (pt_detmath.h) 5 blocks
                                // are 3 % of a frame faster than 4 and than 6 (profiles/r02d_shade_ab.jsonl)
#endif
constexpr int kBlock = 128;
constexpr int kWalkBlock = PT_WALK_BLOCK;
constexpr int kWalkMinBlocks = PT_EXTEND_MIN_BLOCKS * 128 / kWalkBlock;   // same warps per SM whatever the block size

__device__ __forceinline__ f4 ld4(const f4* p) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    return f4{v.x, v.y, v.z, v.w};
}
// last read of a state entry (k_shade consumes set k, nobody reads it again): PT_STATE_CS >= 1 loads it with the
// streaming hint (ld.global.cs, evict-first) so the dead lines leave L2 before the BVH does; PT_STATE_CS == 2 also
// writes the hit records with st.global.cs
#ifndef PT_STATE_CS
#define PT_STATE_CS 0
#endif
__device__ __forceinline__ f4 ld4_last(const f4* p) {
#if PT_STATE_CS
    const float4 v = __ldcs(reinterpret_cast<const float4*>(p));
#else
    const float4 v = *reinterpret_cast<const float4*>(p);
#endif
    return f4{v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ void st4(f4* p, f4 v) { *reinterpret_cast<float4*>(p) = make_float4(v.x, v.y, v.z, v.w); }

// Part `part` of `parts` owns the rows r = part (mod parts) of the rank's lanes; its state arrays are indexed
// by the position inside the part.  (parts > 1: the frame is walked as independent interleaved sub-frames on
// separate streams so that the tail of one part's kernel overlaps the other part's work; `sum` / `rad` are
// shared and addressed by lane.)
__global__ void __launch_bounds__(kBlock) k_init_lanes(SceneView sv, LaneArrays la, int part, int parts, uint32_t n_part) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_part) return;
    const uint32_t w = (uint32_t)sv.width;
    const uint32_t lane = ((i / w) * (uint32_t)parts + (uint32_t)part) * w + i % w;
    int px, py;
    lane_pixel(sv, lane, px, py);
    // first pass of a frame: seed the pixel's stream (device.cu:226) and clear its sum; a later progressive pass
    // continues the stream where the pixel's last sample left it
    const bool first = sv.first_sample == 0;
    uint32_t rng = first ? rng_seed((uint32_t)px, (uint32_t)(py + sv.seed_row_offset + (int)lane_stream(sv, lane) * sv.height))
                         : la.rng_save[lane];
    f3 o, d;
    camera_ray(sv, px, py, rng, o, d);
    st4(la.ray_o[0] + i, f4{o.x, o.y, o.z, -1.0f});  // .w: prev_pdf of the NEE mode (camera ray: count emitters fully)
    if (la.rad) st4(la.rad + lane, f4{0.0f, 0.0f, 0.0f, 0.0f});
    st4(la.ray_d[0] + i, f4{d.x, d.y, d.z, u2f(pack_ctl((uint32_t)sv.first_sample, 0, LOBE_NONE))});
    st4(la.thr[0] + i, f4{1.0f, 1.0f, 1.0f, u2f(rng)});
    if (first) st4(la.sum + lane, f4{0.0f, 0.0f, 0.0f, 0.0f});
    la.queue[0][i] = lane;
}

struct LaneRaySource {   // entry i of the current state set
    const f4* ray_o;
    const f4* ray_d;
    const uint32_t* queue;   // lane id of entry i: addresses the save area of parked walks
    const u4* susp;
    __device__ __forceinline__ bool load(uint32_t i, f3& o, f3& d, float& tmin, float& tmax) const {
        const f4 a = ld4(ray_o + i), b = ld4(ray_d + i);
        o = f3{a.x, a.y, a.z};
        d = f3{b.x, b.y, b.z};
        tmin = kTMin;
        tmax = kTMax;
        return (f2u(b.w) >> 31) != 0;   // k_shade sets bit 31 of the control word of a lane whose walk is parked
    }
    __device__ __forceinline__ void restore(uint32_t i, Hit& hit, uint32_t& ng_base, uint32_t& ng_mask, uint32_t& tg_base,
                                            uint32_t& tg_mask, int& sp, uint32_t& top_b, uint32_t& top_m, const WalkStack& stack) const {
        const u4* p = susp + (size_t)queue[i] * kSuspWords;
        const u4 h = p[0], g = p[1], t = p[2];
        hit = Hit{u2f(h.x), u2f(h.y), u2f(h.z), h.w};
        ng_base = g.x;
        ng_mask = g.y;
        tg_base = g.z;
        tg_mask = g.w;
        sp = (int)t.x;
        top_b = t.y;
        top_m = t.z;
        for (int e = 0; e + 1 < sp; e += 2) {
            const u4 w = p[3 + (e >> 1)];
            stack.st(e, make_uint2(w.x, w.y));
            if (e + 2 < sp) stack.st(e + 1, make_uint2(w.z, w.w));
        }
    }
};
struct LaneHitSink {
    f4* hit;
    const uint32_t* queue;
    u4* susp;
    __device__ __forceinline__ void store(uint32_t i, const Hit& h) const { st4(hit + i, f4{h.t, h.u, h.v, u2f(h.tri)}); }
    __device__ __forceinline__ void suspend(uint32_t i, const Hit& h, uint32_t ng_base, uint32_t ng_mask, uint32_t tg_base,
                                            uint32_t tg_mask, int sp, uint32_t top_b, uint32_t top_m, const WalkStack& stack) const {
        u4* p = susp + (size_t)queue[i] * kSuspWords;
        p[0] = u4{f2u(h.t), f2u(h.u), f2u(h.v), h.tri};
        p[1] = u4{ng_base, ng_mask, tg_base, tg_mask};
        p[2] = u4{(uint32_t)sp, top_b, top_m, 0u};
        for (int e = 0; e + 1 < sp; e += 2) {   // entries 0 .. sp-2 live in local memory, the top in registers
            const uint2 a = stack.ld(e);
            const uint2 b = (e + 2 < sp) ? stack.ld(e + 1) : make_uint2(0u, 0u);
            p[3 + (e >> 1)] = u4{a.x, a.y, b.x, b.y};
        }
        st4(hit + i, f4{0.0f, 0.0f, 0.0f, u2f(kPending)});
    }
};

// persistent warps: every warp pulls rays of queue k until the queue is drained
template <bool COUNT>
__global__ void __launch_bounds__(kWalkBlock, kWalkMinBlocks) k_extend(SceneView sv, LaneArrays la, int k) {
    const uint32_t n = la.counts[k];
    WalkCounters wc{0, 0};
    if (n) {
        // (k & 1) selects the state set; written out so that the parameter struct is never indexed dynamically
        const uint32_t* q = (k & 1) ? la.queue[1] : la.queue[0];
        warp_walk<COUNT, false, true>(sv.nodes, sv.tris, n, la.cursor + k,
                         LaneRaySource{(k & 1) ? la.ray_o[1] : la.ray_o[0], (k & 1) ? la.ray_d[1] : la.ray_d[0], q, la.susp},
                         LaneHitSink{la.hit, q, la.susp}, wc, sv.plane_magic);
    }
    if (COUNT) {
        unsigned long long a = wc.nodes, b = wc.tris;
        for (int s = 16; s > 0; s >>= 1) {
            a += __shfl_down_sync(0xffffffffu, a, s);
            b += __shfl_down_sync(0xffffffffu, b, s);
        }
        if ((threadIdx.x & 31) == 0 && (a | b)) {
            atomicAdd(la.walk_counters + 0, a);
            atomicAdd(la.walk_counters + 1, b);
        }
    }
}

template <bool NEE>
__global__ void __launch_bounds__(kBlock, PT_SHADE_MIN_BLOCKS) k_shade(SceneView sv, LaneArrays la, int k) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool odd = (k & 1) != 0;
    const f4 *in_o = odd ? la.ray_o[1] : la.ray_o[0], *in_d = odd ? la.ray_d[1] : la.ray_d[0], *in_t = odd ? la.thr[1] : la.thr[0];
    f4 *out_o = odd ? la.ray_o[0] : la.ray_o[1], *out_d = odd ? la.ray_d[0] : la.ray_d[1], *out_t = odd ? la.thr[0] : la.thr[1];
    const uint32_t* in_q = odd ? la.queue[1] : la.queue[0];
    uint32_t* out_q = odd ? la.queue[0] : la.queue[1];
    // everything entry i needs is fetched before the queue length is known (the arrays are padded):
    // queue length, lane id, ray, throughput and hit are five independent loads, the lane's sum
    // follows the lane id
    const uint32_t n = la.counts[k];
    const uint32_t lane = in_q[i];
    const f4 d = ld4_last(in_d + i), t = ld4_last(in_t + i), hv = ld4_last(la.hit + i);
    bool keep = false, shadow = false;
    LaneState L;
    f3 last_term = mk3(0.0f);
    uint32_t shadow_flags = 0, resume_bit = 0;
    const bool parked = i < n && f2u(hv.w) == kPending;
    if (parked) {
        // the walk of this ray was interrupted (pt_walk_warp.cuh SUSPEND): nothing to shade yet, the lane moves
        // to the next queue unchanged and the next k_extend picks the walk up where it stopped
        const f4 o = ld4_last(in_o + i);
        L.o = f3{o.x, o.y, o.z};
        L.prev_pdf = o.w;
        L.d = f3{d.x, d.y, d.z};
        L.thr = f3{t.x, t.y, t.z};
        L.rng = f2u(t.w);
        const uint32_t ctl = f2u(d.w);
        L.sample = ctl_sample(ctl);
        L.depth = ctl_depth(ctl);
        L.lobe = ctl_lobe(ctl);
        resume_bit = 0x80000000u;
        keep = true;
    }
    {
        const unsigned mp = __ballot_sync(0xffffffffu, parked);
        if (mp && (threadIdx.x & 31) == __ffs(mp) - 1) atomicAdd(la.pending + k, (uint32_t)__popc(mp));
    }
    if (i < n && !parked) {
        f4 s = ld4(la.sum + lane);
        if (NEE) L.prev_pdf = in_o[i].w;  // density of the current ray under BSDF sampling (ray_o.w)
        L.o = mk3(0.0f);  // only written: the old origin is never needed again
        L.d = f3{d.x, d.y, d.z};
        L.thr = f3{t.x, t.y, t.z};
        L.rng = f2u(t.w);
        const uint32_t ctl = f2u(d.w);
        L.sample = ctl_sample(ctl);
        L.depth = ctl_depth(ctl);
        L.lobe = ctl_lobe(ctl);
        L.max_samples = lane_samples(sv, lane);
        Hit h{hv.x, hv.y, hv.z, f2u(hv.w)};
        int px, py;
        lane_pixel(sv, lane, px, py);
        const ShadeResult r = shade_lane<NEE>(sv, L, h, px, py);
        if (NEE) {
            // sample value = light-sampled radiance gathered along the path + last term.  When this very
            // vertex asked for a shadow ray the sample is finished by k_shadow instead (flag 1).
            shadow = L.sh_valid != 0;
            if (L.sample_ended) {
                if (shadow) {
                    shadow_flags = 1u;
                    last_term = L.contrib;
                } else {
                    const f4 g = ld4(la.rad + lane);
                    const f3 value{g.x + L.contrib.x, g.y + L.contrib.y, g.z + L.contrib.z};
                    s.x += value.x;
                    s.y += value.y;
                    s.z += value.z;
                    s.w += sample_lum2(value);
                    st4(la.sum + lane, s);
                    st4(la.rad + lane, f4{0.0f, 0.0f, 0.0f, 0.0f});
                }
            }
        } else if (L.sample_ended) {  // color += trace_path(...), in sample order
            s.x += L.contrib.x;
            s.y += L.contrib.y;
            s.z += L.contrib.z;
            s.w += sample_lum2(L.contrib);
            st4(la.sum + lane, s);
        }
        keep = r == LANE_CONTINUE;
        if (!keep) la.rng_save[lane] = L.rng;  // the pixel's stream after its last sample of this pass
    }
    // warp-aggregated append of the survivors to the other state set
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (m) {
        const int lid = threadIdx.x & 31;
        const int leader = __ffs(m) - 1;
        uint32_t base = 0;
        if (lid == leader) base = atomicAdd(la.counts + k + 1, (uint32_t)__popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (keep) {
            const uint32_t pos = base + __popc(m & ((1u << lid) - 1u));
            st4(out_o + pos, f4{L.o.x, L.o.y, L.o.z, NEE ? L.prev_pdf : 0.0f});
            st4(out_d + pos, f4{L.d.x, L.d.y, L.d.z, u2f(pack_ctl(L.sample, L.depth, L.lobe) | resume_bit)});
            st4(out_t + pos, f4{L.thr.x, L.thr.y, L.thr.z, u2f(L.rng)});
            out_q[pos] = lane;
        }
    }
    if (NEE) {  // warp-aggregated append of the shadow requests of this iteration
        const unsigned ms = __ballot_sync(0xffffffffu, shadow);
        if (ms) {
            const int lid = threadIdx.x & 31;
            const int leader = __ffs(ms) - 1;
            uint32_t base = 0;
            if (lid == leader) base = atomicAdd(la.sh_counts + k, (uint32_t)__popc(ms));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (shadow) {
                const uint32_t pos = base + __popc(ms & ((1u << lid) - 1u));
                st4(la.sh_o + pos, f4{L.sh_o.x, L.sh_o.y, L.sh_o.z, L.sh_tmax});
                st4(la.sh_d + pos, f4{L.sh_d.x, L.sh_d.y, L.sh_d.z, u2f(lane)});
                st4(la.sh_c + pos, f4{L.sh_c.x, L.sh_c.y, L.sh_c.z, u2f(shadow_flags)});
                st4(la.sh_f + pos, f4{last_term.x, last_term.y, last_term.z, 0.0f});
            }
        }
    }
}

// ---- k_shadow: any-hit walk of the shadow requests k_shade<true> queued in iteration k (the
// reference's unused visibiliy_test, device.cu:97-110).  An unoccluded request adds its contribution to
// the lane's gathered radiance; a request flagged 1 also finishes the lane's sample.
struct ShadowRaySource {
    const f4* sh_o;
    const f4* sh_d;
    __device__ __forceinline__ bool load(uint32_t i, f3& o, f3& d, float& tmin, float& tmax) const {
        const f4 a = ld4(sh_o + i), b = ld4(sh_d + i);
        o = f3{a.x, a.y, a.z};
        d = f3{b.x, b.y, b.z};
        tmin = kTMin;
        tmax = a.w;
        return false;
    }
};
struct ShadowSink {
    const f4* sh_d;
    const f4* sh_c;
    const f4* sh_f;
    f4* rad;
    f4* sum;
    __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
        const uint32_t lane = f2u(sh_d[i].w);
        const f4 c = ld4(sh_c + i);
        f4 g = ld4(rad + lane);
        if (h.tri == kNoHit) {
            g.x += c.x;
            g.y += c.y;
            g.z += c.z;
        }
        if (f2u(c.w) & 1u) {
            const f4 f = ld4(sh_f + i);
            f4 s = ld4(sum + lane);
            const f3 value{g.x + f.x, g.y + f.y, g.z + f.z};
            s.x += value.x;
            s.y += value.y;
            s.z += value.z;
            s.w += sample_lum2(value);
            st4(sum + lane, s);
            g = f4{0.0f, 0.0f, 0.0f, 0.0f};
        }
        st4(rad + lane, g);
    }
};

__global__ void __launch_bounds__(kWalkBlock, kWalkMinBlocks) k_shadow(SceneView sv, LaneArrays la, int k) {
    const uint32_t n = la.sh_counts[k];
    WalkCounters wc{0, 0};
    if (n)
        warp_walk<false, true, false>(sv.nodes, sv.tris, n, la.sh_cursor + k, ShadowRaySource{la.sh_o, la.sh_d},
                               ShadowSink{la.sh_d, la.sh_c, la.sh_f, la.rad, la.sum}, wc, sv.plane_magic);
}

// device.cu:246-253.  sum is indexed by lane; the frame is written row-flipped.
__global__ void __launch_bounds__(256) k_resolve(SceneView sv, const f4* __restrict__ sum, float inv_spp, int transfer,
                                                 uint32_t* __restrict__ rgba, f4* __restrict__ mean) {
    const uint32_t lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= (uint32_t)sv.n_lanes) return;
    int px, py;
    lane_pixel(sv, lane, px, py);
    const f4 s = ld4(sum + lane);
    const f3 c = f3{s.x, s.y, s.z} * inv_spp;
    const size_t ofs = (size_t)px + (size_t)sv.width * (size_t)(sv.height - 1 - py);
    rgba[ofs] = to_rgba(output_transfer(c, transfer));   // the float frame stays linear
    st4(mean + ofs, f4{c.x, c.y, c.z, 1.0f});
}

// Progressive mode: how converged is the frame after `n` samples per pixel?  Per pixel, from sum.xyz (sum of the
// sample values) and sum.w (sum of their squared luminances): m = lum(sum)/n, unbiased sample variance
// s2 = max(0, w/n - m^2) n/(n-1), variance of the mean s2/n, relative to m^2 + 1e-2 (the epsilon of the relMSE
// gate).  The mean of that over the pixels estimates the relMSE of the frame against its own limit.
// binary64 throughout; one partial per block added in a fixed tree so that the result does not depend on timing.
__global__ void __launch_bounds__(256) k_pass_stats(const f4* __restrict__ sum, uint32_t n_lanes, int n, double* __restrict__ partial) {
    __shared__ double sh[256];
    const uint32_t lane = blockIdx.x * blockDim.x + threadIdx.x;
    double rel = 0.0;
    if (lane < n_lanes && n > 1) {
        const f4 s = ld4(sum + lane);
        const double dn = (double)n;
        const double m = (0.2126 * (double)s.x + 0.7152 * (double)s.y + 0.0722 * (double)s.z) / dn;
        const double s2 = fmax(0.0, (double)s.w / dn - m * m) * (dn / (dn - 1.0));
        rel = (s2 / dn) / (m * m + 1e-2);
        if (!(rel == rel)) rel = 0.0;   // NaN sums (a NaN sample value) do not poison the frame's figure
    }
    sh[threadIdx.x] = rel;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}

// several streams per pixel: per-pixel sum of the streams' sums, added in stream order (the order in which the
// oracle adds the partial frames of the corresponding ranks)
__global__ void __launch_bounds__(256) k_fold_streams(const f4* __restrict__ sum, f4* __restrict__ out, uint32_t n_pix, int streams) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    f4 a = ld4(sum + i);
    for (int v = 1; v < streams; ++v) {
        const f4 b = ld4(sum + (size_t)v * n_pix + i);
        a.x += b.x;
        a.y += b.y;
        a.z += b.z;
        a.w += b.w;   // second moments add like the sums (b200pt_pass_stats on a folded frame)
    }
    st4(out + i, a);
}

// ROWS sharding: lane-indexed sums -> full-frame buffer (x + W*y), zero elsewhere
__global__ void __launch_bounds__(256) k_scatter_rows(SceneView sv, const f4* __restrict__ sum, f4* __restrict__ frame) {
    const uint32_t lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= (uint32_t)sv.n_lanes) return;
    int px, py;
    lane_pixel(sv, lane, px, py);
    st4(frame + ((size_t)px + (size_t)sv.width * (size_t)py), ld4(sum + lane));
}

struct CameraRaySource {
    SceneView sv;
    __device__ __forceinline__ bool load(uint32_t i, f3& o, f3& d, float& tmin, float& tmax) const {
        const int px = (int)(i % (uint32_t)sv.width), py = (int)(i / (uint32_t)sv.width);
        uint32_t rng = rng_seed((uint32_t)px, (uint32_t)py);
        camera_ray(sv, px, py, rng, o, d);
        tmin = kTMin;
        tmax = kTMax;
        return false;
    }
};
struct ArrayRaySource {
    const float *o, *d;
    float tmin, tmax;
    __device__ __forceinline__ bool load(uint32_t i, f3& oo, f3& dd, float& t0, float& t1) const {
        oo = f3{o[3 * i], o[3 * i + 1], o[3 * i + 2]};
        dd = f3{d[3 * i], d[3 * i + 1], d[3 * i + 2]};
        t0 = tmin;
        t1 = tmax;
        return false;
    }
};
struct ArrayHitSink {
    f4* hits;
    __device__ __forceinline__ void store(uint32_t i, const Hit& h) const { st4(hits + i, f4{h.t, h.u, h.v, u2f(h.tri)}); }
};

__global__ void __launch_bounds__(kWalkBlock, kWalkMinBlocks) k_primary(SceneView sv, f4* hits, uint32_t* cursor) {
    WalkCounters wc{0, 0};
    warp_walk<false, false, false>(sv.nodes, sv.tris, (uint32_t)(sv.width * sv.height), cursor, CameraRaySource{sv}, ArrayHitSink{hits}, wc,
                     sv.plane_magic);
}

__global__ void __launch_bounds__(kWalkBlock, kWalkMinBlocks) k_trace(SceneView sv, const float* __restrict__ o, const float* __restrict__ d,
                                                  int n, float tmin, float tmax, f4* hits, uint32_t* cursor) {
    WalkCounters wc{0, 0};
    warp_walk<false, false, false>(sv.nodes, sv.tris, (uint32_t)n, cursor, ArrayRaySource{o, d, tmin, tmax}, ArrayHitSink{hits}, wc,
                     sv.plane_magic);
}

__global__ void k_sample_disney(const Material* __restrict__ mats, int n, const int32_t* material, const float* wo,
                                const int32_t* lobe_in, const uint32_t* rng_in, float* f, float* wi, float* pdf,
                                int32_t* lobe_out, uint32_t* rng_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Material m = mats[material[i]];
    uint32_t rng = rng_in[i];
    int lobe = lobe_in[i];
    BsdfSample s;
    s.wi = mk3(0.0f);
    s.pdf = 0.0f;
    s.f = mk3(0.0f);
    sample_disney(m, f3{wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]}, rng, lobe, s);
    f[3 * i] = s.f.x;
    f[3 * i + 1] = s.f.y;
    f[3 * i + 2] = s.f.z;
    wi[3 * i] = s.wi.x;
    wi[3 * i + 1] = s.wi.y;
    wi[3 * i + 2] = s.wi.z;
    pdf[i] = s.pdf;
    lobe_out[i] = lobe;
    rng_out[i] = rng;
}

static inline unsigned blocks(uint32_t n, int b) { return n ? (n + b - 1) / b : 1; }
// persistent grid of the walker kernels: blocks per SM x SMs of the CURRENT device (a process may hold contexts on
// different GPUs; every launch wrapper runs with its context's device current)
static int g_sm_count[64] = {};
void set_persistent_blocks(int sm_count) {
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64) g_sm_count[dev] = sm_count;
}
static inline unsigned pblocks(uint32_t n) {
    int dev = 0;
    cudaGetDevice(&dev);
    const int sms = (dev >= 0 && dev < 64 && g_sm_count[dev] > 0) ? g_sm_count[dev] : 148;
    return std::min<unsigned>(blocks(n, kWalkBlock), (unsigned)(sms * PT_BLOCKS_PER_SM * 128 / kWalkBlock));
}

void launch_init_lanes(const SceneView& sv, const LaneArrays& la, int part, int parts, uint32_t n_part, cudaStream_t st) {
    k_init_lanes<<<blocks(n_part, kBlock), kBlock, 0, st>>>(sv, la, part, parts, n_part);
}
void launch_extend(const SceneView& sv, const LaneArrays& la, int k, uint32_t max_count, bool count, cudaStream_t st) {
    if (count)
        k_extend<true><<<pblocks(max_count), kWalkBlock, 0, st>>>(sv, la, k);
    else
        k_extend<false><<<pblocks(max_count), kWalkBlock, 0, st>>>(sv, la, k);
}
void launch_shade(const SceneView& sv, const LaneArrays& la, int k, uint32_t max_count, cudaStream_t st) {
    if (sv.nee)
        k_shade<true><<<blocks(max_count, kBlock), kBlock, 0, st>>>(sv, la, k);
    else
        k_shade<false><<<blocks(max_count, kBlock), kBlock, 0, st>>>(sv, la, k);
}
// development (PT_TIMELINE builds only): arm / disarm the per-warp timestamps around one launch, dump them afterwards
bool timeline_arm(bool on, cudaStream_t st) {
#if PT_TIMELINE
    static unsigned long long* buf = nullptr;
    if (!buf) {
        if (cudaMalloc(&buf, sizeof(unsigned long long) * 6 * 65536) != cudaSuccess) return false;
        cudaMemset(buf, 0, sizeof(unsigned long long) * 6 * 65536);
        cudaMemcpyToSymbol(g_timeline, &buf, sizeof buf);
    }
    static int h_on[2] = {0, 1};
    cudaMemcpyToSymbolAsync(g_timeline_on, &h_on[on ? 1 : 0], sizeof(int), 0, cudaMemcpyHostToDevice, st);
    return true;
#else
    (void)on; (void)st;
    return false;
#endif
}
bool timeline_dump(const char* path) {
#if PT_TIMELINE
    unsigned long long* buf = nullptr;
    cudaMemcpyFromSymbol(&buf, g_timeline, sizeof buf);
    if (!buf) return false;
    std::vector<unsigned long long> h(6 * 65536);
    cudaMemcpy(h.data(), buf, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost);
    FILE* f = fopen(path, "w");
    if (!f) return false;
    for (size_t w = 0; w < 65536; ++w)
        if (h[4 * w])   // warp, start, drained, exit [ns], rays, loop steps, lanes served by node steps, by triangle steps, node steps
            fprintf(f, "%zu %llu %llu %llu %llu %llu %llu %llu %llu\n", w, h[4 * w], h[4 * w + 1], h[4 * w + 2], h[4 * w + 3] >> 32,
                    h[4 * w + 3] & 0xFFFFFFFFull, h[4 * 65536 + 2 * w] >> 32, h[4 * 65536 + 2 * w] & 0xFFFFFFFFull, h[4 * 65536 + 2 * w + 1]);
    fclose(f);
    return true;
#else
    (void)path;
    return false;
#endif
}
// per-triangle shading records (SceneView::tri_rec): gathered once per build / vertex update
__global__ void __launch_bounds__(256) k_tri_records(const f4* __restrict__ pos, const f4* __restrict__ nrm, const int32_t* __restrict__ idx4,
                                                     uint32_t n, TriShade* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int i0 = idx4[4 * (size_t)i + 0], i1 = idx4[4 * (size_t)i + 1], i2 = idx4[4 * (size_t)i + 2], mesh = idx4[4 * (size_t)i + 3];
    TriShade r;
    r.p0 = pos[i0];
    r.p0.w = u2f((uint32_t)mesh);
    r.p1 = pos[i1];
    r.p2 = pos[i2];
    r.n0 = nrm[i0];
    r.n1 = nrm[i1];
    r.n2 = nrm[i2];
    out[i] = r;
}
void launch_tri_records(const f4* pos, const f4* nrm, const int32_t* idx4, uint32_t n, TriShade* out, cudaStream_t st) {
    k_tri_records<<<blocks(n, 256), 256, 0, st>>>(pos, nrm, idx4, n, out);
}
void launch_shadow(const SceneView& sv, const LaneArrays& la, int k, uint32_t max_count, cudaStream_t st) {
    k_shadow<<<pblocks(max_count), kWalkBlock, 0, st>>>(sv, la, k);
}
void launch_resolve(const SceneView& sv, const f4* sum, float inv_spp, int transfer, uint32_t* rgba, f4* mean, cudaStream_t st) {
    k_resolve<<<blocks(sv.n_lanes, 256), 256, 0, st>>>(sv, sum, inv_spp, transfer, rgba, mean);
}
unsigned pass_stats_blocks(uint32_t n_lanes) { return blocks(n_lanes, 256); }
void launch_pass_stats(const f4* sum, uint32_t n_lanes, int n_samples, double* partial, cudaStream_t st) {
    k_pass_stats<<<blocks(n_lanes, 256), 256, 0, st>>>(sum, n_lanes, n_samples, partial);
}
void launch_fold_streams(const f4* sum, f4* out, uint32_t n_pix, int streams, cudaStream_t st) {
    k_fold_streams<<<blocks(n_pix, 256), 256, 0, st>>>(sum, out, n_pix, streams);
}
void launch_scatter_rows(const SceneView& sv, const f4* sum, f4* frame, cudaStream_t st) {
    k_scatter_rows<<<blocks(sv.n_lanes, 256), 256, 0, st>>>(sv, sum, frame);
}
void launch_primary(const SceneView& sv, f4* hits, uint32_t* cursor, cudaStream_t st) {
    k_primary<<<pblocks(sv.width * sv.height), kWalkBlock, 0, st>>>(sv, hits, cursor);
}
void launch_trace(const SceneView& sv, const float* o, const float* d, int n, float tmin, float tmax, f4* hits,
                  uint32_t* cursor, cudaStream_t st) {
    k_trace<<<pblocks(n), kWalkBlock, 0, st>>>(sv, o, d, n, tmin, tmax, hits, cursor);
}
void launch_sample_disney(const Material* mats, int n, const int32_t* material, const float* wo, const int32_t* lobe_in,
                          const uint32_t* rng_in, float* f, float* wi, float* pdf, int32_t* lobe_out, uint32_t* rng_out,
                          cudaStream_t st) {
    k_sample_disney<<<blocks(n, 128), 128, 0, st>>>(mats, n, material, wo, lobe_in, rng_in, f, wi, pdf, lobe_out, rng_out);
}

}  // namespace pt

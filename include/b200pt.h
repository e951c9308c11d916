/* b200pt.h — C ABI of the B200-native render loop (libb200pt.so).
 *
 * This header is the drop-in boundary for the render path of
 * jctemp/owl-path-tracer.  The reference has no plugin API; the seam is the set
 * of OWL C-API calls its host makes (path_tracer/src/application.cpp through
 * path_tracer/src/owl.hpp).  Each entry point below names the reference call
 * site(s) it replaces.  Everything here is plain C: opaque context pointer,
 * POD structs, raw pointers + sizes, int status codes (0 = ok).  Nothing throws
 * across the boundary; b200pt_last_error() returns a thread-local message.
 *
 * Ownership: the caller keeps ownership of every input pointer and may free it
 * as soon as the call returns.  The library owns all device memory.  Pointers
 * returned by b200pt_framebuffer_* are borrowed and stay valid until the next
 * render, resize or destroy.
 *
 * Threading: one host thread per context (same as the reference: one launch
 * params object, one stream).  No CPU fallback exists: every call fails with
 * B200PT_ERR_CUDA when no sm_100 device can be used.
 */
#ifndef B200PT_H
#define B200PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PT_VERSION 100

enum b200pt_status {
    B200PT_OK = 0,
    B200PT_ERR_INVALID = 1,   /* bad argument / call order                        */
    B200PT_ERR_CUDA = 2,      /* CUDA runtime / no device                          */
    B200PT_ERR_EMPTY = 3,     /* build_accel with no geometry ("no geometries",
                                 application.cpp:133)                              */
    B200PT_ERR_BOUNDS = 4     /* index out of range (device get_data trap,
                                 macros.hpp:5-11, checked on the host at upload)   */
};

/* How samples are split over ranks. */
enum b200pt_shard_mode {
    B200PT_SHARD_SPP = 0,     /* rank r renders spp/world samples of every pixel with
                                 its own RNG stream random(x, y + r*H)            */
    B200PT_SHARD_ROWS = 1     /* rank r renders rows y%world==r with all samples and
                                 the reference stream random(x, y): the N-rank sum
                                 is bit-identical to the 1-rank image             */
};

/* replaces create_context(nullptr, 1) + create_module (application.cpp:62-63,
 * owl.hpp:22-37): which GPU, and this process's place in a multi-GPU job. */
typedef struct b200pt_config {
    int32_t device_ordinal;   /* CUDA device to use                               */
    int32_t rank;             /* 0 .. world_size-1                                */
    int32_t world_size;       /* number of cooperating processes (1 = reference)  */
    int32_t shard_mode;       /* enum b200pt_shard_mode                           */
    int32_t flags;            /* B200PT_FLAG_* bits                               */
    int32_t streams_per_pixel; /* 0 or 1: one RNG stream per pixel through all its samples, the reference's
                                 order (device.cu:226-243).  V > 1 (SHARD_SPP only): every pixel is rendered by
                                 V concurrent streams, stream g = rank*V + v of world_size*V, seeded
                                 random(x, y + g*H) and rendering spp/(world_size*V) samples — exactly the
                                 streams a (world_size*V)-GPU SHARD_SPP job uses, so the image is the one
                                 such a job produces (sums added in stream order).  V times more paths in
                                 flight: +20 % throughput at V = 4 on the 1080p dragon frame.           */
} b200pt_config;

#define B200PT_FLAG_COUNTERS   1  /* count nodes/triangles visited (slower; for the
                                     roofline's algorithmic bytes)                 */
#define B200PT_FLAG_HOST_BVH   2  /* reserved                                       */
#define B200PT_FLAG_NEE        8  /* next-event estimation: light sampling of the emissive triangles (by power)
                                     and of the environment map when one is in use (texels by luminance) with
                                     a shadow any-hit ray per path vertex, combined with the BSDF-sampled
                                     emitter hits / misses by the power heuristic.  NOT the reference's estimator
                                     (its light sampling, lights.hpp + visibiliy_test device.cu:97-110, is
                                     dead code): an opt-in mode with the same mode in the oracle.      */
#define B200PT_FLAG_TIMING     4  /* CUDA events around every extend / shade launch
                                     (fills extend_ms / shade_ms in b200pt_stats).  Also keeps
                                     the frame on ONE stream: without this flag frames of
                                     >= 2^18 lanes are rendered as two interleaved sub-frames
                                     on two streams (same result bit for bit, ~1-6 % faster),
                                     whose kernels overlap and cannot be timed one by one. */

/* material_data, device_global.hpp:19-36 — 17 floats, 68 bytes, same order. */
typedef struct b200pt_material {
    float base_color[3];
    float subsurface;
    float metallic;
    float specular;
    float specular_tint;
    float roughness;
    float anisotropic;
    float sheen;
    float sheen_tint;
    float clearcoat;
    float clearcoat_gloss;
    float ior;
    float specular_transmission;
    float specular_transmission_roughness;
    float emission;
} b200pt_material;

/* hit_data subset (device.hpp:14-26) for the parity hook. mesh < 0 = miss. */
typedef struct b200pt_hit {
    int32_t mesh;             /* mesh_index (order of b200pt_add_mesh calls)       */
    int32_t prim;             /* primitive index inside that mesh                  */
    float t, u, v;            /* OptiX barycentric convention: P=(1-u-v)p0+u p1+v p2 */
} b200pt_hit;

typedef struct b200pt_stats {
    double render_ms;         /* device time of the last render (CUDA events)      */
    double extend_ms;         /* share spent in the extend (traversal) kernel      */
    double shade_ms;          /* share spent in the shade kernel                   */
    double other_ms;          /* raygen/resolve/queue bookkeeping                  */
    double build_ms;          /* last b200pt_build_accel                           */
    uint64_t rays;            /* radiance segments traced (primary + bounce)       */
    uint64_t paths;           /* camera samples started                            */
    uint64_t nodes_visited;   /* 8-wide nodes fetched   (FLAG_COUNTERS only)       */
    uint64_t tris_tested;     /* triangles intersected  (FLAG_COUNTERS only)       */
    uint64_t kernel_launches; /* kernels launched by the last render               */
    uint64_t iterations;      /* wavefront iterations of the last render           */
    uint32_t bvh_nodes;       /* 8-wide nodes in the tree                          */
    uint32_t bvh_tris;        /* triangles referenced by the tree                  */
    uint64_t bvh_bytes;       /* nodes*80 + triangle-pair blocks*80                */
    uint64_t shadow_rays;     /* shadow (any-hit) rays traced (B200PT_FLAG_NEE)    */
    double shadow_ms;         /* share spent in the shadow kernel (FLAG_TIMING)    */
    uint32_t n_lights;        /* emissive triangles in the light table (FLAG_NEE)  */
    uint32_t reserved;
    double refit_ms;          /* last b200pt_refit_accel                           */
} b200pt_stats;

typedef struct b200pt_ctx b200pt_ctx;

const char* b200pt_last_error(void);
int b200pt_version(void);

/* create_context / destroy_context  (application.cpp:62, Main.cpp:30). */
int b200pt_create(const b200pt_config* cfg, b200pt_ctx** out);
int b200pt_destroy(b200pt_ctx* ctx);

/* One call per entity, replaces the per-entity block of bind_sbt_data
 * (application.cpp:197-246): 4 device buffers + owlGeomCreate + mesh_index /
 * material_index + optional RGBA8 NEAREST/CLAMP texture (already v-flipped by the
 * caller exactly as the reference does).  verts/normals: nv*3 floats; uvs: nv*2
 * floats or NULL; idx: nt*3 int32 (local to this mesh).  Returns the mesh index in
 * *mesh_index (= call order, mesh_id++ at application.cpp:211). */
int b200pt_add_mesh(b200pt_ctx* ctx, const float* verts, const float* normals, const float* uvs,
                    int32_t nv, const int32_t* idx, int32_t nt, int32_t material_index,
                    const uint8_t* tex_rgba, int32_t tex_w, int32_t tex_h, int32_t* mesh_index);

/* init_owl_world (application.cpp:131-140): all meshes -> one acceleration
 * structure, built on the GPU (LBVH -> 8-wide compressed).  B200PT_ERR_EMPTY when
 * no mesh was added. */
int b200pt_build_accel(b200pt_ctx* ctx);

/* Moving geometry (the reference rebuilds everything when anything moves — it has no such call).
 * b200pt_update_mesh replaces the positions (and, when normals != NULL, the normals) of mesh `mesh_index`; nv
 * must be the mesh's vertex count, indices and uvs stay.  b200pt_refit_accel then recomputes triangle blocks,
 * child boxes and quantisation grids of the existing tree bottom-up on the GPU (one launch per level) without
 * touching its topology: 0.1-0.2 ms where a build takes milliseconds.  The closest hit is defined
 * independently of the tree (smallest t, ties to the smaller triangle id), so a refitted tree returns exactly
 * the hits of a freshly built one (tests/test_refit.py); only the traversal cost degrades when the motion is
 * large — rebuild with b200pt_build_accel then.  Rendering between update and refit / build is an error. */
int b200pt_update_mesh(b200pt_ctx* ctx, int32_t mesh_index, const float* verts, const float* normals /* may be NULL */,
                       int32_t nv);
int b200pt_refit_accel(b200pt_ctx* ctx);

/* material buffer upload (application.cpp:257-259) and its per-frame re-upload
 * reset_field (application.cpp:297-304).  Cheap; never rebuilds the BVH. */
int b200pt_set_materials(b200pt_ctx* ctx, const b200pt_material* m, int32_t nm);

/* environment texture + launch params (application.cpp:253-254, 285-289).
 * rgba may be NULL / w,h 0 ("environment_map" handle 0, device.cu:138). */
int b200pt_set_environment(b200pt_ctx* ctx, const uint8_t* rgba, int32_t w, int32_t h,
                           int32_t use, int32_t auto_, const float color[3], float intensity);
/* Float environment map, a superset of the reference behind its own entry point: rgb = w*h*3
 * floats of linear radiance, same row order, point sampling and direction mapping as the RGBA8 map, values not
 * limited to [0, 1] (the reference's own route squeezes the .hdr through stb's LDR conversion).
 * Replaces a map set with b200pt_set_environment and vice versa. */
int b200pt_set_environment_f32(b200pt_ctx* ctx, const float* rgb, int32_t w, int32_t h,
                               int32_t use, int32_t auto_, const float color[3], float intensity);

/* camera_data (camera.hpp:16-22) as bound at application.cpp:272-275. */
int b200pt_set_camera(b200pt_ctx* ctx, const float origin[3], const float llc[3],
                      const float horizontal[3], const float vertical[3]);

/* fb_size, max_samples, max_path_depth (application.cpp:255,271,277-278). */
int b200pt_set_render(b200pt_ctx* ctx, int32_t width, int32_t height, int32_t max_samples,
                      int32_t max_path_depth);

/* render_frame's owlLaunch2D (application.cpp:366), blocking like
 * owlLaunch+owlLaunchSync (extern/owl/owl/impl.cpp:252-261).
 * world_size==1: complete frame.  world_size>1: b200pt_render only renders this
 * rank's share into the float4 sum buffer; the caller sums the buffers of all
 * ranks into rank 0's (one NCCL reduce over b200pt_accum_buffer) and then calls
 * b200pt_resolve on rank 0. */
int b200pt_render(b200pt_ctx* ctx);
int b200pt_accum_buffer(b200pt_ctx* ctx, void** device_ptr, size_t* n_floats);
int b200pt_resolve(b200pt_ctx* ctx);

/* create_mesh (utils/mesh_loader.cpp:9-83) on the device — SURVEY 8(f).4: the reference de-duplicates a shape's vertices
 * by POSITION index with a std::map while walking its corners; a position index gets the next local id at its first
 * corner and keeps that corner's normal / texcoord.  Same mapping here by atomicMin + prefix sum (csrc/mesh_dedup.cu).
 * In:  the OBJ attribute arrays (host pointers; counts in elements of 3 / 3 / 2 floats) and, per triangle corner, the
 *      indices into them (n_corners = 3 * triangles; corner_vn / corner_vt NULL or all >= 0: regular shapes only —
 *      a mix of corners with and without an index gives B200PT_ERR_INVALID and the caller keeps the host routine).
 * Out: caller-allocated arrays sized for n_corners vertices: positions (3 floats), normals (3, if corner_vn), texcoords
 *      (2, if corner_vt), out_idx[n_corners] local vertex id per corner; *n_out_verts = local vertices written.
 * The outputs are exactly the arrays b200pt_add_mesh takes. */
int b200pt_dedup_mesh(b200pt_ctx* ctx, const float* attrib_vertices, int32_t n_vertices, const float* attrib_normals, int32_t n_normals,
                      const float* attrib_texcoords, int32_t n_texcoords, const int32_t* corner_v, const int32_t* corner_vn,
                      const int32_t* corner_vt, int32_t n_corners, float* out_verts, float* out_normals, float* out_uvs,
                      int32_t* out_idx, int32_t* n_out_verts);

/* ---- the frame's collective inside the library (NCCL over NVLink / NVSwitch).
 * Reference: the OWL runtime fans one launch over the devices of a context (extern/owl/owl/RayGen.cpp:140-146; the
 * reference asks for one device, application.cpp:62).  Here a job is world_size contexts (rank r, one GPU each),
 * in one process or one per process; samples are sharded over ranks and ONE ncclReduce of the float4 sums per
 * frame brings them to rank 0.  libnccl.so.2 is opened on first use (the copy a host process already loaded
 * is reused), so single-GPU users never need it.
 *   one process per rank:  rank 0 calls b200pt_comm_unique_id, ships the 128 bytes to the others by any means,
 *                          every rank calls b200pt_comm_init_rank, then b200pt_render_frame per frame.
 *   one process, n GPUs:   b200pt_comm_init_all(ctxs, n) once, b200pt_render_frame_all(ctxs, n) per frame
 *                          (one host thread per rank inside the call). */
#define B200PT_COMM_ID_BYTES 128
int b200pt_comm_unique_id(void* id_out /* B200PT_COMM_ID_BYTES */);
int b200pt_comm_init_rank(b200pt_ctx* ctx, const void* id /* B200PT_COMM_ID_BYTES */);
int b200pt_comm_init_all(b200pt_ctx* const* ctxs, int32_t n);   /* ctxs[r] must be rank r of world_size n */
/* in-place sum of every rank's b200pt_accum_buffer into rank 0's; blocking on the context's stream */
int b200pt_reduce(b200pt_ctx* ctx);
/* render this rank's share + b200pt_reduce + (rank 0) b200pt_resolve.  Collective: every rank calls it. */
int b200pt_render_frame(b200pt_ctx* ctx);
int b200pt_render_frame_all(b200pt_ctx* const* ctxs, int32_t n);

/* ---- Progressive surface (the reference renders a frame in one blocking launch and has none of
 * this — opt-in, the entry points above are unchanged) -------------------------------------------------
 *
 * b200pt_render_pass renders the NEXT n_samples samples of every pixel (clamped to what is left of
 * max_samples), continuing each pixel's RNG stream where its last sample left it, and resolves the frame as the
 * mean over the samples rendered so far.  Because the reference draws all samples of a pixel from one stream in
 * order (device.cu:226-243), the frame after k samples IS the reference's frame at max_samples = k, bit for
 * bit, and after the last pass it is the frame b200pt_render produces (tests/test_progressive.py).  The frame
 * in progress is dropped by b200pt_render_restart and by any call that changes the scene, camera, environment,
 * materials, size or depth; b200pt_set_render with only a new max_samples extends or shortens it.
 * One stream per pixel only; several ranks: see b200pt_pass_stats below. */
typedef struct b200pt_pass_info {
    int32_t samples_done;     /* samples per pixel in the frame now                                       */
    int32_t samples_in_pass;  /* samples this pass added                                                  */
    int32_t complete;         /* 1: samples_done == max_samples                                           */
    int32_t stopped_early;    /* b200pt_render_progressive: 1 when the target was met before max_samples  */
    double rel_variance;      /* mean over pixels of  Var[pixel mean] / (mean^2 + 1e-2)  in Rec.709 luminance, from the
                                 per-pixel first and second moments of the sample values: an estimate of the
                                 relMSE (same epsilon) of the frame against its own limit; 0 while
                                 samples_done < 2                                                         */
    double render_ms;         /* device time of this pass                                                 */
    uint64_t rays;            /* radiance segments traced by this pass                                    */
} b200pt_pass_info;
int b200pt_render_pass(b200pt_ctx* ctx, int32_t n_samples, b200pt_pass_info* info /* may be NULL */);
int b200pt_render_restart(b200pt_ctx* ctx);
/* Several ranks (SPP sharding, max_samples a multiple of world_size): n_samples counts samples of THIS rank's stream,
 * the pass adds n_samples * world_size samples to the frame.  b200pt_render_pass then leaves a COPY of the rank's
 * sums behind b200pt_accum_buffer (the rank's own sums must survive the in-place reduce); the caller reduces the
 * copies into rank 0's, calls b200pt_resolve there, and b200pt_pass_stats for the figure of the combined frame
 * (owl_path_tracer_b200/distributed.py render_pass).  info->rel_variance of b200pt_render_pass itself is 0 then. */
int b200pt_pass_stats(b200pt_ctx* ctx, b200pt_pass_info* info);

/* Called after every pass with the resolved frame (RGBA8 top row first, float RGBA mean; both borrowed until the
 * next pass).  Return non-zero to stop. */
typedef int (*b200pt_pass_fn)(void* user, const b200pt_pass_info* info, const uint32_t* rgba, const float* mean);
/* Passes of pass_samples until max_samples, until rel_variance <= target_rel_variance (0: no early stop), or until
 * the callback (may be NULL) asks to stop.  Starts a new frame. */
int b200pt_render_progressive(b200pt_ctx* ctx, int32_t pass_samples, double target_rel_variance,
                              b200pt_pass_fn on_pass, void* user, b200pt_pass_info* last /* may be NULL */);

/* Output transfer applied by the resolve step to the RGBA8 frame (the float frame stays linear).
 * LINEAR is the reference as shipped; GAMMA22 is the line it carries commented out (device.cu:248):
 * color = saturate(pow(color, 1 / 2.2)). */
#define B200PT_TRANSFER_LINEAR  0
#define B200PT_TRANSFER_GAMMA22 1
int b200pt_set_output(b200pt_ctx* ctx, int32_t transfer);

/* Same frame, but the caller's HOST settings are re-uploaded and the RGBA8 frame is
 * copied back into host memory inside the call (the "e2e" path of bench.py:
 * materials H2D + render + D2H framebuffer, what test_loop does per frame). */
int b200pt_render_to_host(b200pt_ctx* ctx, const b200pt_material* m, int32_t nm,
                          uint32_t* rgba_out /* W*H, row 0 = top */);

/* buffer_to_pointer(framebuffer) (application.cpp:368): RGBA8, row 0 = top
 * (device.cu:251 flip applied), host pointer. */
int b200pt_framebuffer_rgba8(b200pt_ctx* ctx, const uint32_t** out);
/* mean radiance before quantisation, float RGBA, same row order (for relMSE). */
int b200pt_framebuffer_f32(b200pt_ctx* ctx, const float** out);

/* Parity hook: closest hit of sample 0's camera ray for every pixel, out[W*H],
 * index x + W*y with y as launched (no flip). */
int b200pt_primary_hits(b200pt_ctx* ctx, b200pt_hit* out);

/* Parity hook: closest hit for caller-supplied rays (n rays; o,d as n*3 floats). */
int b200pt_trace(b200pt_ctx* ctx, const float* origins, const float* dirs, int32_t n,
                 float tmin, float tmax, b200pt_hit* out);

/* ---- frames back to back (the reference's real workload: test_loop renders 1 + 100/step frames in a row with only the
 * materials changing, application.hpp:89-108).  With one RNG stream per pixel a frame ends in a long tail — the
 * average launch carries a third of the lanes — so a PIPELINE of contexts bound to the same scene starts frame k+1
 * while frame k drains.  Every frame is rendered by exactly the code of b200pt_render_frame on its own context:
 * frames stay bit-identical to the blocking call.  One host thread per slot lives inside the pipeline object. */
typedef void (*b200pt_drain_fn)(void* user);
/* progress notification used by the pipeline (and usable on its own): fn is called once per frame, from inside
 * b200pt_render / b200pt_render_frame, when the number of pixels still in flight first falls below
 * fraction * (lanes of the frame); fraction <= 0 or fn == NULL removes it. */
int b200pt_set_drain_callback(b200pt_ctx* ctx, float fraction, b200pt_drain_fn fn, void* user);

typedef struct b200pt_pipeline b200pt_pipeline;
/* slots: n_slots >= 1 contexts of the SAME rank with the same scene, camera and frame size bound (the pipeline does not
 * own them).  start_fraction: frame k+1 starts once frame k has fewer than start_fraction * lanes pixels in flight
 * (0: default 0.5).  Several ranks: slot j of every rank must share a communicator (b200pt_comm_init_rank on each
 * slot's context); the frames' reduces are issued in frame order on every rank. */
int b200pt_pipeline_create(b200pt_ctx* const* slots, int32_t n_slots, float start_fraction, b200pt_pipeline** out);
/* Queue one frame: materials are copied, rgba_out (W*H uint32, may be NULL; ignored on ranks > 0) is written when the
 * frame completes.  Blocks only while the slot that takes the frame is still busy with an earlier one. */
int b200pt_pipeline_submit(b200pt_pipeline* p, const b200pt_material* materials, int32_t count, uint32_t* rgba_out);
/* Block until every submitted frame is complete; returns the first error any frame met (b200pt_last_error). */
int b200pt_pipeline_wait(b200pt_pipeline* p);
int b200pt_pipeline_destroy(b200pt_pipeline* p);   /* waits first */

/* b200pt_trace without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same results as b200pt_trace (host pointers in and out); the kernel's staging (ray cursor, origins,
 * directions, raw hits) is a fixed device array inside the library, 2.5 MiB per GPU, held until the process exits.
 * n > B200PT_TRACE_STATIC_MAX is rejected with B200PT_ERR_BOUNDS.  The call performs no cudaMalloc / cudaFree and
 * updates no statistics.  All calls on a GPU share the array: calls that overlap on one GPU race (serialise them). */
#define B200PT_TRACE_STATIC_MAX 65536
int b200pt_trace_static(b200pt_ctx* ctx, const float* origins, const float* dirs, int32_t n,
                        float tmin, float tmax, b200pt_hit* out);

int b200pt_get_stats(b200pt_ctx* ctx, b200pt_stats* out);

/* Parity hook for the BSDF (disney/disney.cuh:31-66) evaluated on the device:
 * for each of n cases: material index, local wo[3], previous lobe, RNG state in ->
 * f[3], wi[3], pdf, lobe, RNG state out.  Arrays are n-long (wo/f/wi n*3). */
int b200pt_sample_disney(b200pt_ctx* ctx, int32_t n, const int32_t* material, const float* wo,
                         const int32_t* lobe_in, const uint32_t* rng_in, float* f, float* wi,
                         float* pdf, int32_t* lobe_out, uint32_t* rng_out);

/* Copy the device-built 8-wide BVH back to the host (80-byte nodes, 48-byte
 * triangles) so the oracle-side checker can walk the same topology and count
 * node/triangle visits for the roofline.  Call with NULL pointers to get sizes. */
int b200pt_export_bvh(b200pt_ctx* ctx, void* nodes, uint32_t* n_nodes, void* tris,
                      uint32_t* n_tris);   /* tris: 80-byte blocks of two triangles (pt_bvh8.h TriPair), n_tris = blocks */

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */

// Host stand-in for extern/owl/owl/include/owl/owl_device.h + the OptiX 7 device API, so that the
// reference's device/device.cu compiles UNCHANGED with g++ and its programs run as plain functions
// on a CPU thread.  TEST INFRASTRUCTURE (oracle/_ref); never part of the product.
//
// What is real and what is a stand-in:
//   real      every line of device.cu, device/disney/*.cuh, math.hpp, sample_methods.hpp,
//             random.hpp (two in-class specialisations moved out of the class, see patch.py),
//             owl/common/math/vec.h, and make_8bit/make_rgba (owl_device.h:79-91, extracted at
//             build time into gen/owl_make_rgba.inc)
//   stand-in  optixTrace: the driver's BVH traversal is closed source.  traceRay() below asks a
//             closest-hit callback (the oracle's finder: smallest Moller-Trumbore t in
//             (tmin,tmax), ties to the smaller id) and then runs the reference's own
//             closest-hit / miss program exactly as the SBT of application.cpp:79,98-101 would.
//   stand-in  tex2D<float4>: RGBA8, point filter, clamp, normalised coordinates, normalised-float
//             read — the texture the reference creates (owl.hpp:248-257, OWL/Texture.cpp:100-130).
#pragma once
#include "owl/common/math/vec.h"

struct float2 { float x, y; operator owl::common::vec2f() const { return {x, y}; } };
struct float3 { float x, y, z; operator owl::common::vec3f() const { return {x, y, z}; } };
struct float4 { float x, y, z, w; operator owl::common::vec4f() const { return {x, y, z, w}; } };

enum { OPTIX_RAY_FLAG_NONE = 0, OPTIX_RAY_FLAG_DISABLE_ANYHIT = 1, OPTIX_RAY_FLAG_TERMINATE_ON_FIRST_HIT = 4 };
typedef unsigned int OptixVisibilityMask;

namespace refshim {

struct Texture {            // what a cudaTextureObject_t points to in this build
    int w, h;
    const unsigned char* rgba;   // row 0 first, 4 bytes per texel
};

struct Hit {                // layout of the oracle's HitRec / tests' HIT_DTYPE
    int mesh, prim;
    float t, u, v;
};
// closest hit of n rays; mesh < 0 = miss.  Signature of oracle_trace (oracle/oracle.cpp).
typedef void (*trace_fn)(void* scene, const float* org, const float* dir, int n, float tmin, float tmax, int brute,
                         Hit* out);

struct Programs {           // the SBT: filled by the glue with the names application.cpp binds
    void (*closest_hit)();
    void (*miss[2])();
};

struct Thread {             // per-thread launch state (one OptiX thread = one pixel)
    owl::common::vec2i launch_index;
    const void* raygen_data;
    const void* hit_data;       // SBT record of the geometry that was hit
    void* prd;
    float2 bary;
    float3 ray_dir;
    float ray_tmax;
    unsigned prim;
    owl::common::vec3f last_color;   // argument of the last make_rgba (float frame for the tests)
    unsigned long long rays;
};

inline Thread& tls() { static thread_local Thread t{}; return t; }

struct World {
    trace_fn trace;
    void* scene;
    const void* const* entity_records;   // [mesh] -> SBT record (entity_data)
    Programs prog;
};

}  // namespace refshim

// ---- OptiX device API used by device.cu:256-299
inline float2 optixGetTriangleBarycentrics() { return refshim::tls().bary; }
inline float optixGetRayTmax() { return refshim::tls().ray_tmax; }
inline float3 optixGetWorldRayDirection() { return refshim::tls().ray_dir; }
inline unsigned optixGetPrimitiveIndex() { return refshim::tls().prim; }

// ---- tex2D on the texture kind the reference creates
template <typename T>
inline T tex2D(cudaTextureObject_t tex, float u, float v);
template <>
inline float4 tex2D<float4>(cudaTextureObject_t tex, float u, float v) {
    const refshim::Texture* t = (const refshim::Texture*)(uintptr_t)tex;
    int x = (int)floorf(u * (float)t->w), y = (int)floorf(v * (float)t->h);
    x = x < 0 ? 0 : (x > t->w - 1 ? t->w - 1 : x);
    y = y < 0 ? 0 : (y > t->h - 1 ? t->h - 1 : y);
    const unsigned char* p = t->rgba + 4 * ((size_t)y * t->w + x);
    return float4{p[0] / 255.0f, p[1] / 255.0f, p[2] / 255.0f, p[3] / 255.0f};
}

namespace owl {
using namespace owl::common;

inline vec2i getLaunchIndex() { return refshim::tls().launch_index; }
template <typename T>
inline const T& getProgramData() {
    // inside a closest-hit program this is the geometry's record, inside raygen the raygen's
    return *(const T*)(refshim::tls().hit_data ? refshim::tls().hit_data : refshim::tls().raygen_data);
}
template <typename T>
inline T& getPRD() { return *(T*)refshim::tls().prd; }

namespace real {
#include "gen/owl_make_rgba.inc"   // make_8bit + make_rgba, verbatim lines of the real owl_device.h
}
inline uint32_t make_rgba(const vec3f color) {
    refshim::tls().last_color = color;
    return real::make_rgba(color);
}

template <int _rayType = 0, int _numRayTypes = 1>
struct RayT {   // owl_device.h:127-147, same members
    enum { rayType = _rayType };
    enum { numRayTypes = _numRayTypes };
    inline RayT() {}
    inline RayT(const vec3f& origin, const vec3f& direction, float tmin, float tmax,
                OptixVisibilityMask visibilityMask = (OptixVisibilityMask)(-1))
        : origin(origin), direction(direction), tmin(tmin), tmax(tmax), visibilityMask(visibilityMask) {}
    vec3f origin, direction;
    float tmin = 0.f, tmax = 1e30f, time = 0.f;
    OptixVisibilityMask visibilityMask = (OptixVisibilityMask)-1;
};

template <typename RayType, typename PRD>
inline void traceRay(OptixTraversableHandle traversable, const RayType& ray, PRD& prd, uint32_t rayFlags = 0u) {
    const refshim::World* w = (const refshim::World*)(uintptr_t)traversable;
    refshim::Thread& t = refshim::tls();
    refshim::Hit h;
    w->trace(w->scene, &ray.origin.x, &ray.direction.x, 1, ray.tmin, ray.tmax, 0, &h);
    t.rays++;
    void* outer_prd = t.prd;
    const void* outer_hit = t.hit_data;
    t.prd = &prd;
    if (h.mesh >= 0) {
        t.bary = float2{h.u, h.v};
        t.ray_tmax = h.t;
        t.ray_dir = float3{ray.direction.x, ray.direction.y, ray.direction.z};
        t.prim = (unsigned)h.prim;
        t.hit_data = w->entity_records[h.mesh];
        if ((int)RayType::rayType == 0) w->prog.closest_hit();   // shadow rays: no hit program bound
    } else {
        t.hit_data = nullptr;
        w->prog.miss[RayType::rayType]();
    }
    t.prd = outer_prd;
    t.hit_data = outer_hit;
}
}  // namespace owl

#define OPTIX_RAYGEN_PROGRAM(programName) extern "C" void __raygen__##programName
#define OPTIX_CLOSEST_HIT_PROGRAM(programName) extern "C" void __closesthit__##programName
#define OPTIX_MISS_PROGRAM(programName) extern "C" void __miss__##programName

// pt_defs.h — basic types shared by every kernel of the render path.
//
// Everything here is written so that it compiles both as sm_100a device code (nvcc,
// -fmad=false) and as plain C++ (g++, -ffp-contract=off).  The plain-C++ build is
// used ONLY by tests/emu (a host emulation of the kernels' per-lane logic that lets
// the wavefront be debugged without a GPU); the product library has no CPU path.
//
// Floating point contract (see DESIGN.md "Bit parity"): IEEE binary32, round to
// nearest even, no fused contraction except where fmaf() is written out, IEEE
// division and square root.  With that contract the same expression order gives the
// same bits on the device and in the scalar oracle.
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define PT_HD __host__ __device__ __forceinline__
#define PT_D __device__ __forceinline__
#else
#define PT_HD inline
#define PT_D inline
#endif

namespace pt {

// constants: path_tracer/src/types.hpp:9-18
constexpr float kPi = 3.14159265358979323f;
constexpr float kTwoPi = 6.28318530717958648f;
constexpr float kPiOverTwo = 1.57079632679489661f;
constexpr float kPiOverFour = 0.78539816339744830f;
constexpr float kInvPi = 0.31830988618379067f;
constexpr float kTMin = 1E-3f;
constexpr float kTMax = 1E10f;
constexpr float kAlphaMin = 0.001f;

struct f3 {
    float x, y, z;
};
struct f2 {
    float x, y;
};

PT_HD f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
PT_HD f3 mk3(float s) { return f3{s, s, s}; }
PT_HD f3 operator+(f3 a, f3 b) { return f3{a.x + b.x, a.y + b.y, a.z + b.z}; }
PT_HD f3 operator-(f3 a, f3 b) { return f3{a.x - b.x, a.y - b.y, a.z - b.z}; }
PT_HD f3 operator*(f3 a, f3 b) { return f3{a.x * b.x, a.y * b.y, a.z * b.z}; }
PT_HD f3 operator/(f3 a, f3 b) { return f3{a.x / b.x, a.y / b.y, a.z / b.z}; }
PT_HD f3 operator*(f3 a, float s) { return f3{a.x * s, a.y * s, a.z * s}; }
PT_HD f3 operator*(float s, f3 a) { return f3{s * a.x, s * a.y, s * a.z}; }
PT_HD f3 operator/(f3 a, float s) { return f3{a.x / s, a.y / s, a.z / s}; }
PT_HD f3 operator+(f3 a, float s) { return f3{a.x + s, a.y + s, a.z + s}; }
PT_HD f3 operator-(f3 a) { return f3{-a.x, -a.y, -a.z}; }

// owl vec.h:346 — left to right sum of products
PT_HD float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// owl vec.h:330
PT_HD f3 cross(f3 a, f3 b) {
    return f3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
// owl vec.h:353 + owl-common.h:263: v * (1/sqrt(dot))
PT_HD f3 normalize(f3 v) {
    float r = 1.f / sqrtf(dot(v, v));
    return v * r;
}
PT_HD float sqr(float v) { return v * v; }
// math.hpp:6-10
PT_HD float lerpf(float a, float b, float t) { return a + (b - a) * t; }
PT_HD f3 lerp3(f3 a, f3 b, float t) { return a + (b - a) * t; }
// disney_helper.cuh:14 / owl functors.h:65: min(hi, max(lo, x)) with fmin/fmax NaN rules
PT_HD float clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }

// material_data, device_global.hpp:19-36, padded to 20 floats (5 x 16 B) on the device
struct alignas(16) Material {
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
    float pad[3];
};
static_assert(sizeof(Material) == 80, "Material must be 5 x float4");

// disney.cuh:9-13
enum Lobe : int { LOBE_NONE = -1, LOBE_DIFFUSE = 0, LOBE_CLEARCOAT = 1, LOBE_METALLIC = 2, LOBE_GLASS = 3 };

}  // namespace pt

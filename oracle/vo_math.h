// ORACLE — test infrastructure only. Nothing under oracle/ is imported, linked or executed by
// the product path (vision_b200/); only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, as the checker or the timed CPU baseline.
//
// PARITY UNPINNED: the reference (Royalvice/Vision) ships no golden vectors for this path
//  and its compute layer (ocarina + OptiX) is absent (F1/F2), so this file restates
// the reference's algorithm where Vision-owned source exists and *specifies* it where it does not
// (ray/triangle test, tie-break, offset_ray_origin, ...).
//
// Scalar fp32 math for the oracle. Every operation order written here is normative: the CUDA
// path mirrors it with __fmaf_rn/__fmul_rn so results agree bit-for-bit. Compile with
// -ffp-contract=off so the compiler never fuses what is not written as fmaf().
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>

namespace vo {

struct float2 { float x, y; };
struct float3 { float x, y, z; };
struct float4 { float x, y, z, w; };

inline float3 make3(float x, float y, float z) { return {x, y, z}; }
inline float3 operator+(float3 a, float3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline float3 operator-(float3 a, float3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float3 operator*(float3 a, float3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline float3 operator*(float3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float3 operator*(float s, float3 a) { return {a.x * s, a.y * s, a.z * s}; }
inline float3 operator/(float3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline float3 operator/(float3 a, float3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline float3 operator-(float3 a) { return {-a.x, -a.y, -a.z}; }
inline float3 &operator+=(float3 &a, float3 b) { a = a + b; return a; }
inline float3 &operator*=(float3 &a, float3 b) { a = a * b; return a; }
inline float3 &operator*=(float3 &a, float s) { a = a * s; return a; }

// Normative dot/cross: explicit FMA chain, mirrored on the device.
//   dot   = fma(a.z,b.z, fma(a.y,b.y, a.x*b.x))
//   cross = fma(a.y,b.z, -(a.z*b.y)), ...
inline float dot(float3 a, float3 b) {
    return std::fmaf(a.z, b.z, std::fmaf(a.y, b.y, a.x * b.x));
}
inline float3 cross(float3 a, float3 b) {
    return {std::fmaf(a.y, b.z, -(a.z * b.y)),
            std::fmaf(a.z, b.x, -(a.x * b.z)),
            std::fmaf(a.x, b.y, -(a.y * b.x))};
}
inline float length_squared(float3 a) { return dot(a, a); }
inline float length(float3 a) { return std::sqrt(dot(a, a)); }
// normalize = v * (1/sqrt(dot)) with IEEE sqrt and IEEE divide (device: __fsqrt_rn, __frcp_rn)
inline float3 normalize(float3 a) {
    float inv = 1.0f / std::sqrt(dot(a, a));
    return a * inv;
}
inline float max_comp(float3 a) { return std::max(a.x, std::max(a.y, a.z)); }
inline float3 vmin(float3 a, float3 b) { return {std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)}; }
inline float3 vmax(float3 a, float3 b) { return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)}; }
inline float lerp(float t, float a, float b) { return a + t * (b - a); }
inline float sqr(float x) { return x * x; }
inline float saturate(float x) { return std::min(1.0f, std::max(0.0f, x)); }
inline float safe_sqrt(float x) { return std::sqrt(std::max(0.0f, x)); }

inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

constexpr float Pi = 3.14159265358979323846f;
constexpr float InvPi = 0.31830988618379067154f;
constexpr float Inv2Pi = 0.15915494309189533577f;
constexpr float Inv4Pi = 0.07957747154594766788f;
constexpr float PiOver2 = 1.57079632679489661923f;
constexpr float PiOver4 = 0.78539816339744830961f;
constexpr float OneMinusEpsilon = 0x1.fffffep-1f;
constexpr float ShadowEpsilon = 1e-4f;              // SURVEY A1 (pbrt value; ocarina absent)
constexpr float RayTMax = 1e30f;                    // default t_max for make_ray

// Column-major 4x4 (m[c][r]); m[3] is the translation column. SURVEY A1.
struct float4x4 {
    float m[4][4];
    static float4x4 identity() {
        float4x4 r{};
        for (int i = 0; i < 4; ++i) r.m[i][i] = 1.f;
        return r;
    }
};
inline float4x4 mul(const float4x4 &a, const float4x4 &b) {   // a*b, column vectors
    float4x4 r{};
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 4; ++row) {
            float s = 0.f;
            for (int k = 0; k < 4; ++k) s += a.m[k][row] * b.m[c][k];
            r.m[c][row] = s;
        }
    return r;
}
inline float3 transform_point(const float4x4 &M, float3 p) {
    // x*col0 + y*col1 + z*col2 + col3, FMA chain per component (normative, mirrored on device)
    return {std::fmaf(M.m[2][0], p.z, std::fmaf(M.m[1][0], p.y, std::fmaf(M.m[0][0], p.x, M.m[3][0]))),
            std::fmaf(M.m[2][1], p.z, std::fmaf(M.m[1][1], p.y, std::fmaf(M.m[0][1], p.x, M.m[3][1]))),
            std::fmaf(M.m[2][2], p.z, std::fmaf(M.m[1][2], p.y, std::fmaf(M.m[0][2], p.x, M.m[3][2])))};
}
inline float3 transform_vector(const float4x4 &M, float3 v) {
    return {std::fmaf(M.m[2][0], v.z, std::fmaf(M.m[1][0], v.y, M.m[0][0] * v.x)),
            std::fmaf(M.m[2][1], v.z, std::fmaf(M.m[1][1], v.y, M.m[0][1] * v.x)),
            std::fmaf(M.m[2][2], v.z, std::fmaf(M.m[1][2], v.y, M.m[0][2] * v.x))};
}
float4x4 inverse(const float4x4 &m);       // double-precision host inverse (vo_math.cpp)
float4x4 transpose(const float4x4 &m);

}  // namespace vo

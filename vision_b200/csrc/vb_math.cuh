// Device fp32 math for the path-tracing megakernel. Operation order mirrors oracle/vo_math.h
// (the normative statement); vb_pt.cu is compiled with -fmad=false so only the FMAs written here
// as __fmaf_rn fuse, exactly like the oracle under -ffp-contract=off.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vbm {

struct f3 { float x, y, z; };
struct f2 { float x, y; };

__device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ f3 operator/(f3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ f3 operator-(f3 a) { return {-a.x, -a.y, -a.z}; }

__device__ __forceinline__ float dot(f3 a, f3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ f3 cross(f3 a, f3 b) {
    return {__fmaf_rn(a.y, b.z, -(a.z * b.y)), __fmaf_rn(a.z, b.x, -(a.x * b.z)), __fmaf_rn(a.x, b.y, -(a.y * b.x))};
}
__device__ __forceinline__ float length_squared(f3 a) { return dot(a, a); }
__device__ __forceinline__ float length(f3 a) { return __fsqrt_rn(dot(a, a)); }
__device__ __forceinline__ f3 normalize(f3 a) {
    float inv = __fdiv_rn(1.0f, __fsqrt_rn(dot(a, a)));
    return a * inv;
}
__device__ __forceinline__ float fmin_(float a, float b) { return b < a ? b : a; }   // == std::min
__device__ __forceinline__ float fmax_(float a, float b) { return a < b ? b : a; }   // == std::max
__device__ __forceinline__ float max_comp(f3 a) { return fmax_(a.x, fmax_(a.y, a.z)); }
__device__ __forceinline__ float lerp(float t, float a, float b) { return a + t * (b - a); }
__device__ __forceinline__ float sqr(float x) { return x * x; }
__device__ __forceinline__ float safe_sqrt(float x) { return __fsqrt_rn(fmax_(0.0f, x)); }
__device__ __forceinline__ f3 splat(float v) { return {v, v, v}; }

// column-major 4x4 in a flat array m[4*c + r]
__device__ __forceinline__ f3 transform_point(const float *m, f3 p) {
    return {__fmaf_rn(m[8], p.z, __fmaf_rn(m[4], p.y, __fmaf_rn(m[0], p.x, m[12]))),
            __fmaf_rn(m[9], p.z, __fmaf_rn(m[5], p.y, __fmaf_rn(m[1], p.x, m[13]))),
            __fmaf_rn(m[10], p.z, __fmaf_rn(m[6], p.y, __fmaf_rn(m[2], p.x, m[14])))};
}
__device__ __forceinline__ f3 transform_vector(const float *m, f3 v) {
    return {__fmaf_rn(m[8], v.z, __fmaf_rn(m[4], v.y, m[0] * v.x)),
            __fmaf_rn(m[9], v.z, __fmaf_rn(m[5], v.y, m[1] * v.x)),
            __fmaf_rn(m[10], v.z, __fmaf_rn(m[6], v.y, m[2] * v.x))};
}
// row-major 3x3 normal matrix
__device__ __forceinline__ f3 apply_normal(const float *m, f3 n) {
    return {__fmaf_rn(m[2], n.z, __fmaf_rn(m[1], n.y, m[0] * n.x)),
            __fmaf_rn(m[5], n.z, __fmaf_rn(m[4], n.y, m[3] * n.x)),
            __fmaf_rn(m[8], n.z, __fmaf_rn(m[7], n.y, m[6] * n.x))};
}

constexpr float Pi = 3.14159265358979323846f;
constexpr float InvPi = 0.31830988618379067154f;
constexpr float Inv2Pi = 0.15915494309189533577f;
constexpr float Inv4Pi = 0.07957747154594766788f;
constexpr float PiOver2 = 1.57079632679489661923f;
constexpr float OneMinusEpsilon = 0x1.fffffep-1f;
constexpr float ShadowEpsilon = 1e-4f;
constexpr float RayTMax = 1e30f;

// mirrors vo::sincos_2pi
__device__ __forceinline__ void sincos_2pi(float u, float &s, float &c) {
    float q = floorf(u * 4.0f + 0.5f);
    float r = u - q * 0.25f;
    float x = r * 6.28318530717958647692f;
    float z = x * x;
    float sp = ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z - 1.6666654611e-1f) * z * x + x;
    float cp = ((2.443315711809948e-5f * z - 1.388731625493765e-3f) * z + 4.166664568298827e-2f) * z * z - 0.5f * z + 1.0f;
    int iq = (int)q & 3;
    if (iq == 0) { s = sp; c = cp; }
    else if (iq == 1) { s = cp; c = -sp; }
    else if (iq == 2) { s = -sp; c = -cp; }
    else { s = -cp; c = sp; }
}

// mirror vo::vexp / vo::vlog (oracle/vo_pt.cpp): fixed operation sequence, bit-identical to the host
__device__ __forceinline__ float vexp(float x) {
    if (!(x > -87.0f)) return x != x ? x : 0.0f;
    if (x > 88.0f) return __int_as_float(0x7f800000);
    float n = floorf(x * 1.44269504088896341f + 0.5f);
    float r = __fmaf_rn(n, -0.693359375f, x);
    r = __fmaf_rn(n, 2.12194440e-4f, r);
    float z = r * r;
    float p = ((((1.9875691500e-4f * r + 1.3981999507e-3f) * r + 8.3334519073e-3f) * r + 4.1665795894e-2f) * r + 1.6666665459e-1f) * r + 5.0000001201e-1f;
    float y = p * z + r + 1.0f;
    return __uint_as_float(__float_as_uint(y) + ((uint32_t)(int32_t)n << 23));
}
__device__ __forceinline__ float vlog(float x) {
    if (!(x > 0.0f)) return x == 0.0f ? __int_as_float(0xff800000) : __int_as_float(0x7fc00000);
    uint32_t b = __float_as_uint(x);
    int32_t e = (int32_t)((b >> 23) & 0xffu) - 126;
    float m = __uint_as_float((b & 0x007fffffu) | 0x3f000000u);
    if (m < 0.70710678118654752440f) { e -= 1; m = m + m - 1.0f; } else { m = m - 1.0f; }
    float z = m * m;
    float y = ((((((((7.0376836292e-2f * m - 1.1514610310e-1f) * m + 1.1676998740e-1f) * m - 1.2420140846e-1f) * m + 1.4249322787e-1f) * m
                 - 1.6668057665e-1f) * m + 2.0000714765e-1f) * m - 2.4999993993e-1f) * m + 3.3333331174e-1f) * m * z;
    float fe = (float)e;
    y = y + -2.12194440e-4f * fe;
    y = y + -0.5f * z;
    float r = m + y;
    return r + 0.693359375f * fe;
}

// mirrors vo::vatan2
__device__ __forceinline__ float vatan2(float y, float x) {
    float ax = fabsf(x), ay = fabsf(y);
    float r;
    if (ax == 0.f && ay == 0.f) r = 0.f;
    else {
        bool swap = ay > ax;
        float t = swap ? ax / ay : ay / ax;
        float off = 0.f;
        if (t > 0.4142135623730950f) { off = 0.78539816339744830961f; t = (t - 1.0f) / (t + 1.0f); }
        float z = t * t;
        float p = (((8.05374449538e-2f * z - 1.38776856032e-1f) * z + 1.99777106478e-1f) * z - 3.33329491539e-1f) * z * t + t;
        r = off + p;
        if (swap) r = 1.57079632679489661923f - r;
    }
    if (x < 0.f) r = 3.14159265358979323846f - r;
    return y < 0.f ? -r : r;
}

// mirrors vo::offset_ray_origin (Waechter & Binder)
__device__ __forceinline__ float offset1(float pc, float nc) {
    int32_t of_i = (int32_t)(256.0f * nc);
    int32_t pi = __float_as_int(pc);
    float p_i = __int_as_float(pi + (pc < 0.f ? -of_i : of_i));
    return fabsf(pc) < (1.0f / 32.0f) ? pc + (1.0f / 65536.0f) * nc : p_i;
}
__device__ __forceinline__ f3 offset_ray_origin(f3 p, f3 n) { return {offset1(p.x, n.x), offset1(p.y, n.y), offset1(p.z, n.z)}; }

// reference src/math/util.h:13-34
__device__ __forceinline__ uint32_t tea4(uint32_t v0, uint32_t v1) {
    uint32_t s0 = 0;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}
__device__ __forceinline__ float lcg(uint32_t &state) {
    state = 1664525u * state + 1013904223u;
    return (float)(state & 0x00ffffffu) * (1.0f / 16777216.0f);
}

}  // namespace vbm

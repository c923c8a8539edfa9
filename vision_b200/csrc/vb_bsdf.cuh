// BSDF lobes of the megakernel (device). Included by vb_pt.cu after the Interaction/Sampler/PtScene
// definitions. Compiled with -fmad=false; operation order mirrors oracle/vo_bsdf.inl (normative).
//   GGX microfacet            reference src/base/scattering/microfacet.{h,cpp}
//   Fresnel                   reference src/base/scattering/fresnel.h, src/math/optics.h, src/math/complex.h
//   Lambert / MicrofacetRefl  reference src/base/scattering/bxdf.{h,cpp}
//   FresnelBlend (substrate)  reference src/render_core/material/substrate.cpp:12-70
//   Lobe / DielectricLobe / PureReflectionLobe   reference src/base/scattering/lobe.cpp:45-63,111-145,204-250,253-449,716-735
//   material plugins          reference src/render_core/material/{diffuse,mirror,glass,metal,substrate}.cpp
// The reference dispatches on Polymorphic<Material>; here it is a switch on vb_material_rec::type.
#pragma once

namespace {

struct ScatterEval { f3 f; float pdf; uint32_t flags; };
struct BSDFSample { ScatterEval eval; f3 wi; float eta; };
constexpr uint32_t FlagReflection = 1u << 1, FlagTransmission = 1u << 2, FlagDiffuse = 1u << 3, FlagGlossy = 1u << 4;

__device__ __forceinline__ f3 operator/(f3 a, f3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
__device__ __forceinline__ float abs_dot(f3 a, f3 b) { return fabsf(dot(a, b)); }
__device__ __forceinline__ f3 face_forward(f3 v, f3 n) { return dot(v, n) < 0.f ? -v : v; }
__device__ __forceinline__ bool same_hemisphere(f3 a, f3 b) { return a.z * b.z > 0.f; }
__device__ __forceinline__ float pow5(float x) { float x2 = x * x; return (x2 * x2) * x; }
__device__ __forceinline__ float average3(f3 v) { return ((v.x + v.y) + v.z) / 3.0f; }
__device__ __forceinline__ f3 reflect(f3 wo, f3 n) { return -wo + n * (2.0f * dot(wo, n)); }
__device__ __forceinline__ bool refract(f3 wi, f3 n, float eta, f3 &wt) {        // optics.h:27-39
    float cos_theta_i = dot(n, wi);
    float sin_theta_i_2 = fmax_(0.f, 1.0f - sqr(cos_theta_i));
    float sin_theta_t_2 = sin_theta_i_2 / sqr(eta);
    bool valid = sin_theta_t_2 < 1.0f;
    float cos_theta_t = safe_sqrt(1.0f - sin_theta_t_2);
    wt = -wi / eta + n * (cos_theta_i / eta - cos_theta_t);
    return valid;
}

// ---- Fresnel: optics.h:70-79 (dielectric), :92-104 + complex.h (conductor) ----
__device__ float fresnel_dielectric(float abs_cos_theta_i, float eta) {
    float sin_theta_i_2 = 1.0f - sqr(abs_cos_theta_i);
    float sin_theta_t_2 = sin_theta_i_2 / sqr(eta);
    float cos_theta_t = safe_sqrt(1.0f - sin_theta_t_2);
    float r_parl = (eta * abs_cos_theta_i - cos_theta_t) / (eta * abs_cos_theta_i + cos_theta_t);
    float r_perp = (abs_cos_theta_i - eta * cos_theta_t) / (abs_cos_theta_i + eta * cos_theta_t);
    return sin_theta_t_2 >= 1.0f ? 1.0f : (sqr(r_parl) + sqr(r_perp)) * 0.5f;
}
struct Cx { float re, im; };
__device__ __forceinline__ Cx cmul(Cx a, Cx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
__device__ __forceinline__ Cx cdiv(Cx a, Cx z) {
    float scale = 1.0f / (z.re * z.re + z.im * z.im);
    return {scale * (a.re * z.re + a.im * z.im), scale * (a.im * z.re - a.re * z.im)};
}
__device__ __forceinline__ Cx cadd(Cx a, Cx b) { return {a.re + b.re, a.im + b.im}; }
__device__ __forceinline__ Cx csub(Cx a, Cx b) { return {a.re - b.re, a.im - b.im}; }
__device__ __forceinline__ float cnorm_sqr(Cx z) { return z.re * z.re + z.im * z.im; }
__device__ __forceinline__ Cx csqrt(Cx z) {
    float n = __fsqrt_rn(cnorm_sqr(z));
    float t1 = __fsqrt_rn(0.5f * (n + fabsf(z.re)));
    float t2 = 0.5f * z.im / t1;
    Cx r;
    r.re = n == 0.f ? 0.f : (z.re >= 0.f ? t1 : fabsf(t2));
    r.im = n == 0.f ? 0.f : (z.re >= 0.f ? t2 : copysignf(t1, z.im));
    return r;
}
__device__ float fresnel_complex(float cos_theta_i, float eta_re, float k) {
    Cx eta{eta_re, k};
    float sin_theta_i_2 = 1.0f - sqr(cos_theta_i);
    Cx sin_theta_t_2 = cdiv(Cx{sin_theta_i_2, 0.f}, cmul(eta, eta));
    Cx cos_theta_t = csqrt(csub(Cx{1.f, 0.f}, sin_theta_t_2));
    Cx ci{cos_theta_i, 0.f};
    Cx r_parl = cdiv(csub(cmul(eta, ci), cos_theta_t), cadd(cmul(eta, ci), cos_theta_t));
    Cx r_perp = cdiv(csub(ci, cmul(eta, cos_theta_t)), cadd(ci, cmul(eta, cos_theta_t)));
    return (cnorm_sqr(r_parl) + cnorm_sqr(r_perp)) * 0.5f;
}

// ---- GGX: microfacet.cpp:11-24 (D), :41-50 (lambda), :71-92 (VNDF), :94-110 (sample_wh), :146-155 (PDF_wh) ----
struct GGX { float ax, ay; };
__device__ __forceinline__ float ggx_D(const GGX &m, f3 wh) {
    f3 H{wh.x / m.ax, wh.y / m.ay, wh.z / 1.0f};
    float alpha2 = m.ax * m.ay;
    return InvPi / (alpha2 * sqr(length_squared(H)));
}
__device__ __forceinline__ float ggx_lambda(const GGX &m, f3 w) {
    float sqr_alpha_tan_n = (sqr(m.ax * w.x) + sqr(m.ay * w.y)) / sqr(w.z);
    float ret = 0.5f * (__fsqrt_rn(1.0f + sqr_alpha_tan_n) - 1.0f);
    return w.z == 0.f ? 0.f : ret;
}
__device__ __forceinline__ float ggx_G1(const GGX &m, f3 w) { return 1.0f / (1.0f + ggx_lambda(m, w)); }
__device__ __forceinline__ float ggx_G(const GGX &m, f3 wo, f3 wi) { return 1.0f / (1.0f + ggx_lambda(m, wo) + ggx_lambda(m, wi)); }
__device__ f3 ggx_sample_vndf(f3 Ve, f2 u, float ax, float ay) {
    f3 Vh = normalize(f3{ax * Ve.x, ay * Ve.y, Ve.z});
    float lenSq = Vh.x * Vh.x + Vh.y * Vh.y;
    f3 T1 = lenSq > 1e-7f ? f3{-Vh.y, Vh.x, 0.0f} / __fsqrt_rn(lenSq) : f3{1, 0, 0};
    f3 T2 = lenSq > 1e-7f ? cross(Vh, T1) : f3{0.0f, 1.0f, 0.0f};
    float r = __fsqrt_rn(u.x), s, c;
    sincos_2pi(u.y, s, c);
    float tx = r * c, ty = r * s;
    ty = lerp(0.5f * (1.0f + Vh.z), safe_sqrt(1.0f - sqr(tx)), ty);
    f3 Nh = T1 * tx + T2 * ty + Vh * safe_sqrt(1.0f - (tx * tx + ty * ty));
    return normalize(f3{ax * Nh.x, ay * Nh.y, fmax_(0.0f, Nh.z)});
}
__device__ __forceinline__ f3 ggx_sample_wh(const GGX &m, f3 wo, f2 u) {
    bool flip = wo.z < 0.f;
    f3 wh = ggx_sample_vndf(flip ? -wo : wo, u, m.ax, m.ay);
    return flip ? -wh : wh;
}
__device__ __forceinline__ float ggx_pdf_wh(const GGX &m, f3 wo, f3 wh) {
    return ggx_D(m, wh) * ggx_G1(m, wo) * abs_dot(wo, wh) / fabsf(wo.z);
}
__device__ __forceinline__ float ggx_pdf_wi_reflection(const GGX &m, f3 wo, f3 wh) {
    return ggx_pdf_wh(m, wo, wh) / (4.0f * abs_dot(wo, wh));
}
__device__ __forceinline__ float ggx_pdf_wi_transmission(const GGX &m, f3 wo, f3 wh, f3 wi, float eta) {
    float denom = sqr(dot(wi, wh) * eta + dot(wo, wh));
    float dwh_dwi = abs_dot(wi, wh) / denom;
    return ggx_pdf_wh(m, wo, wh) * dwh_dwi;
}
__device__ __forceinline__ float ggx_brdf_div_fr(const GGX &m, f3 wo, f3 wh, f3 wi) {
    return ggx_D(m, wh) * ggx_G(m, wo, wi) / fabsf(4.0f * wo.z * wi.z);
}
__device__ __forceinline__ float ggx_btdf_div_ft(const GGX &m, f3 wo, f3 wh, f3 wi, float eta) {
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    float numerator = ggx_D(m, wh) * ggx_G(m, wo, wi) * fabsf(dot(wi, wh) * dot(wo, wh));
    float denom = sqr(dot(wi, wh) * eta + dot(wo, wh)) * fabsf(cos_theta_i * cos_theta_o);
    float ft = numerator / denom;
    float factor = 1.0f / sqr(eta);
    ft = denom == 0.f ? 0.f : ft;
    return ft * factor;
}
__device__ __forceinline__ void calculate_alpha(float alpha, float anisotropic, float &ax, float &ay) {   // microfacet.h:43-60
    float x = anisotropic < 0.f ? alpha / (1.f + anisotropic) : alpha * (1.f - anisotropic);
    float y = anisotropic < 0.f ? alpha * (1.f + anisotropic) : alpha / (1.f - anisotropic);
    bool iso = fabsf(anisotropic) <= 1e-4f;
    ax = iso ? alpha : x; ay = iso ? alpha : y;
}

// ---- LUTs (MaterialLut, material.cpp:14-27): normalised coords, linear filter, clamp; software sampled ----
struct LutAxis { int i0, i1; float f; };
__device__ __forceinline__ LutAxis lut_axis(float c, uint32_t n) {
    float x = c * (float)n - 0.5f;
    float fl = floorf(x);
    LutAxis a;
    a.f = x - fl;
    int i = (int)fl;
    a.i0 = min(max(i, 0), (int)n - 1);
    a.i1 = min(max(i + 1, 0), (int)n - 1);
    return a;
}
__device__ float lut_sample2d(const DevLut &L, float u, float v) {
    LutAxis a = lut_axis(u, L.nx), b = lut_axis(v, L.ny);
    float v00 = __ldg(L.data + (size_t)b.i0 * L.nx + a.i0), v10 = __ldg(L.data + (size_t)b.i0 * L.nx + a.i1);
    float v01 = __ldg(L.data + (size_t)b.i1 * L.nx + a.i0), v11 = __ldg(L.data + (size_t)b.i1 * L.nx + a.i1);
    float r0 = v00 + a.f * (v10 - v00);
    float r1 = v01 + a.f * (v11 - v01);
    return r0 + b.f * (r1 - r0);
}
__device__ float lut_sample3d_ch0(const DevLut &L, float u, float v, float w) {
    LutAxis a = lut_axis(u, L.nx), b = lut_axis(v, L.ny), c = lut_axis(w, L.nz);
    float p[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        int z = k ? c.i1 : c.i0;
        size_t row0 = ((size_t)z * L.ny + b.i0) * L.nx, row1 = ((size_t)z * L.ny + b.i1) * L.nx;
        float v00 = __ldg(L.data + (row0 + a.i0) * L.channels), v10 = __ldg(L.data + (row0 + a.i1) * L.channels);
        float v01 = __ldg(L.data + (row1 + a.i0) * L.channels), v11 = __ldg(L.data + (row1 + a.i1) * L.channels);
        float r0 = v00 + a.f * (v10 - v00);
        float r1 = v01 + a.f * (v11 - v01);
        p[k] = r0 + b.f * (r1 - r0);
    }
    return p[0] + c.f * (p[1] - p[0]);
}

}  // namespace
#include "vb_bsdf_lobes.cuh"

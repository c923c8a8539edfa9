// Spectrum plugins of the path on the device: "srgb" (identity RGB, reference src/render_core/spectrum/srgb.cpp:16-57)
// and "hero" = HeroWavelengthSpectrum with dimension 3 (reference src/render_core/spectrum/hero.cpp:15-348,
// src/base/color/spectrum.{h,cpp}, src/base/color/spd.cpp:57-87, src/base/color/cie.h:413-435).
// Mirrors oracle/vo_spectrum.inl statement by statement. Included by vb_pt.cu only.
//
// A `const Swl *` is passed down the shading functions: nullptr = srgb (every spec_* is the identity; the
// kernels instantiated for srgb scenes pass a compile-time nullptr, so they carry no spectral state), non-null = the
// path's three wavelengths and their pdfs.
#pragma once

namespace {

constexpr uint32_t SwlInvalidBit = 0x80000000u;
// bit 30 of the same path word (PathState::swl): the path is in the "only direct" supplement step of Li (vb_pt_bounce.cuh)
constexpr uint32_t PathSupplementBit = 0x40000000u;

struct DevSpectrum {
    const float *rgb2spec;         // float[3][res][res][res][4]
    uint32_t res;
    const float *spd_values;       // SPD pool
    const float4 *spd_meta;        // per SPD: (offset as uint bits, sample_interval = 471 / size, float(sample_count - 2), -) - the
                                   // per-SPD constants of SPD::eval, divided once on the host (same IEEE operations, same bits)
    uint32_t spd_x, spd_y, spd_z, spd_d65;
    float cie_y_integral;
    // decoded coefficients of the CONSTANT colour slots (c0, c1, c2, scale), filled once per scene by k_spec_cache with the same device
    // functions the per-vertex path uses (identical bits): [4 * material + {0 color, 1 spec, 2 sheen tint, 3 coat tint}],
    // [cache_lights + light], [cache_mediums + 2 * medium + {0 sigma_t, 1 sigma_s}]. Textured slots are decoded per texel as before.
    const float4 *cache;
    uint32_t cache_lights, cache_mediums;
};
enum : uint32_t { SC_COLOR = 0, SC_SPEC = 1, SC_SHEEN = 2, SC_COAT = 3 };

struct Swl { float lambda[3], pdf[3]; };

__device__ __forceinline__ float vatanh(float x) { return 0.5f * vlog((1.0f + x) / (1.0f - x)); }
__device__ __forceinline__ float vcosh(float x) { return 0.5f * (vexp(x) + vexp(-x)); }
__device__ __forceinline__ float vfract(float x) { return x - floorf(x); }

// HeroWavelengthSpectrum::sample_wavelength (hero.cpp:288-300) from the path's packed word: bits 0..23 = the draw
// made at sampler dimension -1 (integrator.cpp:48-57), bit 31 = secondary wavelengths invalidated (spectrum.cpp:21-30)
__device__ __forceinline__ Swl swl_make(uint32_t packed) {
    Swl w;
    float u = (float)(packed & 0x00ffffffu) * (1.0f / 16777216.0f);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float offset = (float)i * (1.f / 3.f);
        float up = vfract(u + offset);
        float lambda = 538.f - 138.888889f * vatanh(0.85691062f - 1.82750197f * up);
        w.lambda[i] = lambda;
        w.pdf[i] = 0.0039398042f / sqr(vcosh(0.0072f * (lambda - 538.f)));
    }
    if (packed & SwlInvalidBit) { w.pdf[1] = 0.f; w.pdf[2] = 0.f; }
    return w;
}

// SPD::eval (spd.cpp:68-76), sample_interval_ = 471 / size (spd.cpp:50)
struct SpdPos { uint32_t i; float frac; };
__device__ __forceinline__ SpdPos spd_locate(float4 meta, float lambda) {
    float t = (fmin_(fmax_(lambda, 360.f), 830.f) - 360.f) / meta.y;
    return SpdPos{(uint32_t)fmin_(t, meta.z), vfract(t)};
}
__device__ __forceinline__ float spd_at(const DevSpectrum &P, float4 meta, SpdPos p) {
    const float *func = P.spd_values + __float_as_uint(meta.x);
    float l = __ldg(func + p.i), r = __ldg(func + p.i + 1);
    return lerp(p.frac, l, r);
}
__device__ __forceinline__ float spd_eval(const DevSpectrum &P, uint32_t id, float lambda) {
    float4 meta = __ldg(P.spd_meta + id);
    return spd_at(P, meta, spd_locate(meta, lambda));
}
__device__ __forceinline__ f3 spd_eval3(const DevSpectrum &P, uint32_t id, const Swl &w) {
    return {spd_eval(P, id, w.lambda[0]), spd_eval(P, id, w.lambda[1]), spd_eval(P, id, w.lambda[2])};
}

// RGBSigmoidPolynomial (hero.cpp:27-49)
__device__ __forceinline__ float sigmoid_poly(f3 c, float lambda) {
    float x = __fmaf_rn(__fmaf_rn(c.x, lambda, c.y), lambda, c.z);
    if (isinf(x)) return x > 0.f ? 1.f : 0.f;
    return 0.5f * __fmaf_rn(x, 1.0f / __fsqrt_rn(__fmaf_rn(x, x, 1.f)), 1.f);
}

__device__ __forceinline__ float inverse_smooth_step(float x) {                 // hero.cpp:65-67
    float a = 1.0f - 2.0f * x;
    float as = vatan2(a, safe_sqrt(1.0f - a * a));
    float s, c;
    sincos_2pi(as * (1.0f / 3.0f) * Inv2Pi, s, c);
    return 0.5f - s;
}

// RGBToSpectrumTable::decode_albedo (hero.cpp:93-133, the explicit-lerp form): 8 float4 fetches, software trilinear
__device__ f3 rgb2spec_decode(const DevSpectrum &P, f3 rgb_in) {
    float rgb[3] = {fmin_(fmax_(rgb_in.x, 0.f), 1.f), fmin_(fmax_(rgb_in.y, 0.f), 1.f), fmin_(fmax_(rgb_in.z, 0.f), 1.f)};
    if (rgb[0] == rgb[1] && rgb[1] == rgb[2])
        return {0.f, 0.f, (rgb[0] - 0.5f) * (1.0f / __fsqrt_rn(rgb[0] * (1.0f - rgb[0])))};
    const uint32_t res = P.res;
    uint32_t maxc = rgb[0] > rgb[1] ? (rgb[0] > rgb[2] ? 0u : 2u) : (rgb[1] > rgb[2] ? 1u : 2u);
    float z = maxc == 0 ? rgb[0] : (maxc == 1 ? rgb[1] : rgb[2]);
    float c1 = maxc == 0 ? rgb[1] : (maxc == 1 ? rgb[2] : rgb[0]);
    float c2 = maxc == 0 ? rgb[2] : (maxc == 1 ? rgb[0] : rgb[1]);
    float x = c1 * (float)(res - 1u) / z;
    float y = c2 * (float)(res - 1u) / z;
    float zz = inverse_smooth_step(inverse_smooth_step(z)) * (float)(res - 1u);
    uint32_t xi = min((uint32_t)x, res - 2u), yi = min((uint32_t)y, res - 2u), zi = min((uint32_t)zz, res - 2u);
    float dx = x - (float)xi, dy = y - (float)yi, dz = zz - (float)zi;
    const float4 *T = reinterpret_cast<const float4 *>(P.rgb2spec) + (((size_t)maxc * res + zi) * res + yi) * res + xi;
    const size_t sy = res, sz = (size_t)res * res;
    auto co = [&](uint32_t ox, uint32_t oy, uint32_t oz) {
        float4 v = __ldg(T + oz * sz + oy * sy + ox);
        return f3{v.x, v.y, v.z};
    };
    auto l3 = [](float t, f3 a, f3 b) { return f3{lerp(t, a.x, b.x), lerp(t, a.y, b.y), lerp(t, a.z, b.z)}; };
    return l3(dz, l3(dy, l3(dx, co(0, 0, 0), co(1, 0, 0)), l3(dx, co(0, 1, 0), co(1, 1, 0))),
                  l3(dy, l3(dx, co(0, 0, 1), co(1, 0, 1)), l3(dx, co(0, 1, 1), co(1, 1, 1))));
}

__device__ __forceinline__ f3 sigmoid3(f3 c, const Swl &w) {
    return {sigmoid_poly(c, w.lambda[0]), sigmoid_poly(c, w.lambda[1]), sigmoid_poly(c, w.lambda[2])};
}
// Spectrum::decode_to_albedo (hero.cpp:335-338 / srgb: identity)
__device__ __forceinline__ f3 spec_albedo(const DevSpectrum &P, const Swl *w, f3 rgb) {
    if (!w) return rgb;
    return sigmoid3(rgb2spec_decode(P, rgb), *w);
}
// Spectrum::decode_to_unbound_spectrum (hero.cpp:135-141,343-346)
__device__ __forceinline__ f3 spec_unbound(const DevSpectrum &P, const Swl *w, f3 rgb_in) {
    if (!w) return rgb_in;
    f3 rgb{fmax_(rgb_in.x, 0.f), fmax_(rgb_in.y, 0.f), fmax_(rgb_in.z, 0.f)};
    float scale = 2.f * max_comp(rgb);
    f3 c = rgb2spec_decode(P, scale == 0.f ? f3{0, 0, 0} : rgb / scale);
    return sigmoid3(c, *w) * scale;
}
// the (c0, c1, c2, scale) record of decode_albedo / decode_unbound, as cached by k_spec_cache
__device__ __forceinline__ float4 spec_coefs_albedo(const DevSpectrum &P, f3 rgb) {
    f3 c = rgb2spec_decode(P, rgb);
    return make_float4(c.x, c.y, c.z, 1.f);
}
__device__ __forceinline__ float4 spec_coefs_unbound(const DevSpectrum &P, f3 rgb_in) {
    f3 rgb{fmax_(rgb_in.x, 0.f), fmax_(rgb_in.y, 0.f), fmax_(rgb_in.z, 0.f)};
    float scale = 2.f * max_comp(rgb);
    f3 c = rgb2spec_decode(P, scale == 0.f ? f3{0, 0, 0} : rgb / scale);
    return make_float4(c.x, c.y, c.z, scale);
}
// cached variants: `constant` = the slot has no texture, so its coefficients are entry `idx` of the cache
__device__ __forceinline__ f3 spec_albedo_c(const DevSpectrum &P, const Swl *w, f3 rgb, bool constant, uint32_t idx) {
    if (!w) return rgb;
    if (constant && P.cache) { float4 c = __ldg(P.cache + idx); return sigmoid3(f3{c.x, c.y, c.z}, *w); }
    return spec_albedo(P, w, rgb);
}
__device__ __forceinline__ f3 spec_unbound_c(const DevSpectrum &P, const Swl *w, f3 rgb, bool constant, uint32_t idx) {
    if (!w) return rgb;
    if (constant && P.cache) { float4 c = __ldg(P.cache + idx); return sigmoid3(f3{c.x, c.y, c.z}, *w) * c.w; }
    return spec_unbound(P, w, rgb);
}
__device__ __forceinline__ f3 spec_illum_c(const DevSpectrum &P, const Swl *w, f3 rgb, bool constant, uint32_t idx);
// Spectrum::decode_to_illumination (hero.cpp:339-342,306-313): unbound x D65
__device__ __forceinline__ f3 spec_illum(const DevSpectrum &P, const Swl *w, f3 rgb) {
    if (!w) return rgb;
    return spec_unbound(P, w, rgb) * spd_eval3(P, P.spd_d65, *w);
}

__device__ __forceinline__ f3 spec_illum_c(const DevSpectrum &P, const Swl *w, f3 rgb, bool constant, uint32_t idx) {
    if (!w) return rgb;
    return spec_unbound_c(P, w, rgb, constant, idx) * spd_eval3(P, P.spd_d65, *w);
}

// Spectrum::linear_srgb (hero.cpp:259-261,275-284; cie.h:413-420)
__device__ __forceinline__ f3 spec_to_srgb(const DevSpectrum &P, const Swl *w, f3 sp) {
    if (!w) return sp;
    const float v[3] = {sp.x, sp.y, sp.z};
    f3 sum{0, 0, 0};
    uint32_t valid = 0;
    // the three observer curves have the same sampling (checked by vb_render_frames): one table position per wavelength
    const float4 mx = __ldg(P.spd_meta + P.spd_x), my = __ldg(P.spd_meta + P.spd_y), mz = __ldg(P.spd_meta + P.spd_z);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float p = w->pdf[i];
        SpdPos pos = spd_locate(mx, w->lambda[i]);
        f3 xyz{spd_at(P, mx, pos) * v[i], spd_at(P, my, pos) * v[i], spd_at(P, mz, pos) * v[i]};
        if (p != 0.f) sum = sum + xyz / p;
        if (p > 0.f) valid++;
    }
    float factor = 1.f / ((float)valid * P.cie_y_integral);
    f3 xyz = sum * factor;
    return f3{3.240479f, -0.969256f, 0.055648f} * xyz.x + f3{-1.537150f, 1.875991f, -0.204043f} * xyz.y +
           f3{-0.498535f, 0.041556f, 1.057311f} * xyz.z;
}

}  // namespace

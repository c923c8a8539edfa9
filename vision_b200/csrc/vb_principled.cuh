// PrincipledBSDF on the device: statement-by-statement mirror of oracle/vo_principled.inl (reference
// src/render_core/material/principled_bsdf.cpp:17-461 + LobeSet, src/base/scattering/lobe.cpp:454-713).
// Included by vb_bsdf_lobes.cuh.
#pragma once

namespace {

__device__ float lut_sample2d_ch(const DevLut &L, float u, float v, uint32_t ch) {
    LutAxis a = lut_axis(u, L.nx), b = lut_axis(v, L.ny);
    auto at = [&](int x, int y) { return __ldg(L.data + ((size_t)y * L.nx + x) * L.channels + ch); };
    float r0 = at(a.i0, b.i0) + a.f * (at(a.i1, b.i0) - at(a.i0, b.i0));
    float r1 = at(a.i0, b.i1) + a.f * (at(a.i1, b.i1) - at(a.i0, b.i1));
    return r0 + b.f * (r1 - r0);
}

enum : int { SL_SHEEN = 0, SL_COAT = 1, SL_METALLIC = 2, SL_TRANS = 3, SL_SPEC = 4, SL_DIFFUSE = 5 };
enum : int { FR_DIELECTRIC = 0, FR_SCHLICK = 1, FR_F82 = 2 };

struct SubLobe {
    int kind = SL_DIFFUSE;
    float sample_weight = 0.f, weight = 1.f;      // WeightedLobe
    f3 kr{0, 0, 0};                            // reflection colour | transmission kt | sheen tint | diffuse Kr
    GGX mf{1.f, 1.f};
    int fresnel = FR_DIELECTRIC;
    f3 F0{0, 0, 0}, B{0, 0, 0};
    float eta = 1.f;
    float a = 1.f, b = 0.f;                        // sheen LTC coefficients
};
// The lobe set lives in REGISTERS: no array is ever indexed with a run-time value. A lobe kind occurs at most once and always in
// the order sheen, coat, metallic, transmission, specular, diffuse (principled_bsdf.cpp:352-461), so the set is a bit mask plus
// per-kind fields; every loop over the lobes is six compile-time blocks (VB_EACH_SUBLOBE), and a SubLobe is re-assembled from the
// fields on the spot. (Before: SubLobe s[6] indexed by a run-time count - a 1.2 KB stack frame in k_wf_shade, 240 B of spills.)
struct PrincipledSet {
    uint32_t switches = 0;                      // bit k: lobe kind k is present
    Frame3 frame;
    float sw[6] = {0, 0, 0, 0, 0, 0};           // sample weights (normalised by principled_create), constant indices only
    f3 kr[6];
    GGX mf{1.f, 1.f};
    float ior = 1.f, eta_t = 1.f, trans_w = 1.f, cc_a = 1.f, cc_ior = 1.f, sheen_a = 1.f, sheen_b = 0.f;
    f3 F0_met{0, 0, 0}, B_met{0, 0, 0}, F0_t{0, 0, 0}, F0_s{0, 0, 0};
};

template <int K>
__device__ __forceinline__ SubLobe sub_lobe(const PrincipledSet &ps) {
    SubLobe l;
    l.kind = K; l.sample_weight = ps.sw[K]; l.kr = ps.kr[K];
    if (K == SL_SHEEN) { l.a = ps.sheen_a; l.b = ps.sheen_b; }
    if (K == SL_COAT) { l.fresnel = FR_DIELECTRIC; l.eta = ps.cc_ior; l.mf = GGX{ps.cc_a, ps.cc_a}; }
    if (K == SL_METALLIC) { l.fresnel = FR_F82; l.F0 = ps.F0_met; l.B = ps.B_met; l.mf = ps.mf; }
    if (K == SL_TRANS) { l.fresnel = FR_SCHLICK; l.eta = ps.eta_t; l.F0 = ps.F0_t; l.mf = ps.mf; l.weight = ps.trans_w; }
    if (K == SL_SPEC) { l.fresnel = FR_SCHLICK; l.eta = ps.ior; l.F0 = ps.F0_s; l.mf = ps.mf; }
    return l;
}
// BODY sees `const SubLobe l` and `constexpr int K`, once per present lobe, in lobe order
#define VB_SUBLOBE_BLOCK(ps, KK, ...) if ((ps).switches & (1u << (KK))) { constexpr int K = (KK); const SubLobe l = sub_lobe<K>(ps); (void)K; __VA_ARGS__ }
#define VB_EACH_SUBLOBE(ps, ...)                                                                                                          \
    VB_SUBLOBE_BLOCK(ps, SL_SHEEN, __VA_ARGS__) VB_SUBLOBE_BLOCK(ps, SL_COAT, __VA_ARGS__) VB_SUBLOBE_BLOCK(ps, SL_METALLIC, __VA_ARGS__) \
    VB_SUBLOBE_BLOCK(ps, SL_TRANS, __VA_ARGS__) VB_SUBLOBE_BLOCK(ps, SL_SPEC, __VA_ARGS__) VB_SUBLOBE_BLOCK(ps, SL_DIFFUSE, __VA_ARGS__)


__device__ float saturate(float x) { return fmin_(fmax_(x, 0.f), 1.f); }
__device__ f3 lerp3v(float t, f3 a, f3 b) { return {lerp(t, a.x, b.x), lerp(t, a.y, b.y), lerp(t, a.z, b.z)}; }
__device__ float schlick_F0_from_ior(float ior) { return sqr((ior - 1.0f) / (ior + 1.0f)); }

// Fresnel::evaluate of the sub-lobe (fresnel.h:44-52 Schlick, :60-72 dielectric, :96-104 F82-tint)
__device__ f3 sub_fresnel(const SubLobe &l, float cos_theta) {
    if (l.fresnel == FR_SCHLICK) {
        float F_real = fresnel_dielectric(cos_theta, l.eta);
        float F0_real = schlick_F0_from_ior(l.eta);
        float t = (F_real - F0_real) / (1.f - F0_real);                 // inverse_lerp
        t = saturate(t);
        return lerp3v(t, l.F0, splat(1.f));
    }
    if (l.fresnel == FR_F82) {
        float mu = saturate(1.f - cos_theta);
        float mu5 = pow5(mu);
        f3 f_schlick = lerp3v(mu5, l.F0, splat(1.f));
        f3 r = f_schlick - l.B * cos_theta * mu5 * mu;
        return {saturate(r.x), saturate(r.y), saturate(r.z)};
    }
    return splat(fresnel_dielectric(cos_theta, l.eta));
}

// layering_weight, principled_bsdf.cpp:209-214
__device__ f3 layering_weight(f3 layer_albedo, f3 weight) {
    auto safe_div = [](float a, float b) { return b == 0.f ? 0.f : a / b; };
    f3 tmp{safe_div(layer_albedo.x, weight.x), safe_div(layer_albedo.y, weight.y), safe_div(layer_albedo.z, weight.z)};
    float max_comp = fmax_(tmp.x, fmax_(tmp.y, tmp.z));
    return weight * saturate(1.f - max_comp);
}

__device__ __forceinline__ PrincipledSet principled_create_of(const PtScene &S, const Swl *swl, const Interaction &it, const uint32_t mat_id) {
    const vb_material_rec &m = S.materials[mat_id];
    PrincipledSet ps;
    ps.frame = it.shading;
    auto clampf = [](float v, float lo, float hi) { return fmin_(fmax_(v, lo), hi); };
    const uint32_t switches = (uint32_t)m.params[15];
    f3 color = material_color_of(S, swl, m.color, it, mat_id, SC_COLOR);
    float ior = eval_slot3(S, m.ior, it.uv).x;
    float roughness = clampf(eval_slot3(S, m.roughness, it.uv).x, 0.0001f, 1.f);
    float anisotropic = m.params[0];
    f3 specular_tint = material_color_of(S, swl, m.spec, it, mat_id, SC_SPEC);
    float aspect = __fsqrt_rn(1.f - anisotropic * 0.9f);
    GGX mf{fmax_(0.001f, sqr(roughness) / aspect), fmax_(0.001f, sqr(roughness) * aspect)};
    f3 weight = splat(1.f);
    float cos_theta = dot(it.wo, it.ng);
    float front = cos_theta > 0.f ? 1.f : 0.f;
    auto to_ratio_x = [](const GGX &g) { return __fsqrt_rn(__fsqrt_rn(g.ax * g.ay)); };
    if (switches & (1u << SL_SHEEN)) {
        SubLobe l;
        ps.switches |= 1u << SL_SHEEN;
        f3 sheen_tint = spec_albedo_c(S.spec, swl, f3{m.params[4], m.params[5], m.params[6]}, true, 4u * mat_id + SC_SHEEN);
        float sheen_weight = m.params[2] * front;
        float sheen_roughness = m.params[3];
        l.kr = sheen_tint * sheen_weight * weight;
        const DevLut &L = S.luts[m.params[14] != 0.f ? VB_LUT_SHEEN_APPROX : VB_LUT_SHEEN_VOLUME];
        float cx = 0.f, cy = 0.f, cz = 0.f;
        if (!(L.data == nullptr)) { cx = lut_sample2d_ch(L, cos_theta, sheen_roughness, 0); cy = lut_sample2d_ch(L, cos_theta, sheen_roughness, 1); cz = lut_sample2d_ch(L, cos_theta, sheen_roughness, 2); }
        l.a = cx; l.b = cy;
        l.kr = l.kr * cz;
        f3 albedo = l.kr;
        ps.sw[SL_SHEEN] = average3(albedo); ps.kr[SL_SHEEN] = l.kr; ps.sheen_a = l.a; ps.sheen_b = l.b;
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_COAT)) {
        SubLobe l;
        ps.switches |= 1u << SL_COAT;
        float cc_weight = m.params[7] * front;
        float cc_roughness = sqr(clampf(m.params[8], 0.0001f, 1.f));
        float cc_ior = m.params[9];
        f3 cc_tint = spec_albedo_c(S.spec, swl, f3{m.params[10], m.params[11], m.params[12]}, true, 4u * mat_id + SC_COAT);
        l.fresnel = FR_DIELECTRIC; l.eta = cc_ior;
        l.mf = GGX{cc_roughness, cc_roughness};
        l.kr = weight * cc_weight * cc_tint;
        const DevLut &L = S.luts[VB_LUT_COAT];
        float s = (L.data == nullptr) ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, (cc_ior - 1.003f) / (4.f - 1.003f));
        f3 albedo = l.kr * s;
        ps.sw[SL_COAT] = average3(albedo); ps.kr[SL_COAT] = l.kr; ps.cc_a = cc_roughness; ps.cc_ior = cc_ior;
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_METALLIC)) {
        SubLobe l;
        ps.switches |= 1u << SL_METALLIC;
        float metallic = m.params[1] * front;
        l.fresnel = FR_F82; l.F0 = color;
        const float f = 6.f / 7.f, f5 = ((f * f) * (f * f)) * f;
        f3 f_schlick = lerp3v(f5, l.F0, splat(1.f));
        l.B = f_schlick * (7.f / (f5 * f)) * (splat(1.f) - specular_tint);
        l.mf = mf;
        l.kr = weight * metallic;
        ps.sw[SL_METALLIC] = metallic * average3(weight); ps.kr[SL_METALLIC] = l.kr; ps.F0_met = l.F0; ps.B_met = l.B;
        weight = weight * (1.0f - metallic);
    }
    if (switches & (1u << SL_TRANS)) {
        SubLobe l;
        ps.switches |= 1u << SL_TRANS;
        float trans_weight = m.params[13];
        float eta = cos_theta > 0.f ? ior : 1.0f / ior;                   // Interaction::correct_eta
        l.fresnel = FR_SCHLICK; l.eta = eta; l.F0 = specular_tint * schlick_F0_from_ior(eta);
        l.mf = mf;
        l.kr = color;
        f3 t_weight = weight * trans_weight;
        ps.sw[SL_TRANS] = average3(t_weight); ps.trans_w = average3(t_weight); ps.kr[SL_TRANS] = l.kr; ps.eta_t = eta; ps.F0_t = l.F0;
        weight = weight * (1.0f - trans_weight);
    }
    if (switches & (1u << SL_SPEC)) {
        SubLobe l;
        ps.switches |= 1u << SL_SPEC;
        l.fresnel = FR_SCHLICK; l.eta = ior; l.F0 = specular_tint * schlick_F0_from_ior(ior);
        l.mf = mf;
        l.kr = weight;
        const DevLut &L = S.luts[VB_LUT_SPECULAR];
        float z = __fsqrt_rn(fabsf((ior - 1.0f) / (ior + 1.0f)));
        float s = (L.data == nullptr) ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, z);
        f3 albedo = lerp3v(s, l.F0, splat(1.f)) * l.kr;
        ps.sw[SL_SPEC] = average3(albedo); ps.kr[SL_SPEC] = l.kr; ps.F0_s = l.F0;
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_DIFFUSE)) {
        ps.switches |= 1u << SL_DIFFUSE;
        ps.kr[SL_DIFFUSE] = color * weight * front;
        ps.sw[SL_DIFFUSE] = average3(ps.kr[SL_DIFFUSE]);
    }
    ps.mf = mf; ps.ior = ior;
    float sum = 0.f;                                                     // LobeSet::normalize_sampled_weight
#pragma unroll
    for (int k = 0; k < 6; ++k) if (ps.switches & (1u << k)) sum += ps.sw[k];
#pragma unroll
    for (int k = 0; k < 6; ++k) if (ps.switches & (1u << k)) ps.sw[k] = ps.sw[k] / sum;
    return ps;
}
__device__ __forceinline__ PrincipledSet principled_create(const PtScene &S, const Swl *swl, const Interaction &it) { return principled_create_of(S, swl, it, it.mat_id); }

// SheenLTC (principled_bsdf.cpp:61-117)
__device__ float sheen_eval_ltc(const SubLobe &l, f3 wi) {
    f3 w{l.a * wi.x + l.b * wi.z, l.a * wi.y, wi.z};                // inv_M
    float len = length(w);
    w = w / len;
    float det = sqr(l.a);
    float jacobian = det / (len * len * len);
    return w.z * InvPi * jacobian;
}

__device__ __forceinline__ ScatterEval sub_evaluate_local(const PtScene &S, const SubLobe &l, f3 wo, f3 wi, float *eta_out) {
    ScatterEval se{f3{0, 0, 0}, 0.f, 0u};
    switch (l.kind) {
        case SL_SHEEN: {
            float v = sheen_eval_ltc(l, wi);
            se.f = l.kr * v / wi.z;
            se.pdf = v;
            if (wi.z < 0.f || wo.z < 0.f) se.f = {0, 0, 0};
            se.flags = FlagGlossy | FlagReflection;
        } break;
        case SL_COAT: case SL_SPEC: case SL_METALLIC: {                   // MicrofacetLobe / PureReflectionLobe
            bool s = same_hemisphere(wo, wi);
            f3 wh = normalize(wo + wi);
            f3 whf = face_forward(wh, f3{0, 0, 1});
            f3 F = sub_fresnel(l, abs_dot(wo, whf));
            f3 f = F * ggx_brdf_div_fr(l.mf, wo, whf, wi) * l.kr;
            float pdf = ggx_pdf_wi_reflection(l.mf, wo, wh);
            se.f = s ? f : f3{0, 0, 0};
            se.pdf = s ? pdf : 0.f;
            se.flags = FlagGlossy | FlagReflection;
            if (l.kind == SL_METALLIC) {                                  // MetallicLobe::compensate() == true
                const DevLut &L = S.luts[VB_LUT_PURE_REFLECTION];
                if (!(L.data == nullptr)) se.f = se.f * (1.0f / lut_sample2d(L, __fsqrt_rn(l.mf.ax * l.mf.ay), wo.z));
            }
        } break;
        case SL_TRANS: {                                                  // DielectricLobe with a Schlick Fresnel, lobe.cpp:318-404
            bool reflect_ = same_hemisphere(wo, wi);
            float eta_p = reflect_ ? 1.0f : l.eta;
            if (eta_out) *eta_out = eta_p;
            f3 wh = normalize(wo + wi * eta_p);
            wh = face_forward(wh, wo);
            f3 F = sub_fresnel(l, abs_dot(wh, wo));
            f3 T = splat(1.f) - F;
            float refl_prob = average3(F) / average3(T * l.kr + F);
            const DevLut &L = l.eta > 1.0f ? S.luts[VB_LUT_DIELECTRIC] : S.luts[VB_LUT_DIELECTRIC_INV];
            float comp = 1.f;
            if (!(L.data == nullptr)) {
                float z = l.eta > 1.0f ? (l.eta - 1.003f) / (5.0f - 1.003f) : (1.0f / l.eta - 1.003f) / (5.0f - 1.003f);
                comp = 1.0f / lut_sample3d_ch0(L, __fsqrt_rn(__fsqrt_rn(l.mf.ax * l.mf.ay)), fabsf(wo.z), z);
            }
            if (reflect_) {
                se.f = F * ggx_brdf_div_fr(l.mf, wo, wh, wi);
                se.pdf = ggx_pdf_wi_reflection(l.mf, wo, wh) * refl_prob;
                se.flags = FlagGlossy | FlagReflection;
            } else {
                f3 new_wh = face_forward(wh, wo);
                f3 wh2 = normalize(wo + wi * l.eta);
                se.f = T * ggx_btdf_div_ft(l.mf, wo, wh2, wi, l.eta) * l.kr;
                se.pdf = ggx_pdf_wi_transmission(l.mf, wo, new_wh, wi, l.eta) * (1.0f - refl_prob);
                se.flags = FlagGlossy | FlagTransmission;
            }
            se.f = se.f * comp;
        } break;
        default: {
            bool s = same_hemisphere(wo, wi);
            se.f = s ? l.kr * InvPi : f3{0, 0, 0};
            se.pdf = s ? fabsf(wi.z) * InvPi : 0.f;
            se.flags = FlagDiffuse | FlagReflection;
        } break;
    }
    return se;
}

// LobeSet::evaluate_impl (world), lobe.cpp:667-683
__device__ __forceinline__ ScatterEval principled_evaluate(const PtScene &S, const PrincipledSet &ps, f3 world_wo, f3 world_wi, float *eta_out) {
    f3 wo = ps.frame.to_local(world_wo), wi = ps.frame.to_local(world_wi);
    ScatterEval ret{f3{0, 0, 0}, 0.f, 0u};
    VB_EACH_SUBLOBE(ps, {
        ScatterEval se = sub_evaluate_local(S, l, wo, wi, eta_out);
        se.f = se.f * fabsf(wi.z);                                    // Lobe::evaluate_impl
        // valid_world_factor: same side of the shading normal (1 for the dielectric lobe)
        float factor = l.kind == SL_TRANS ? 1.f : (dot(world_wo, ps.frame.z) * dot(world_wi, ps.frame.z) > 0.f ? 1.f : 0.f);
        se.f = se.f * (l.weight * factor);
        se.pdf = se.pdf * (l.sample_weight * factor);
        ret.f = ret.f + se.f;
        ret.pdf = ret.pdf + se.pdf;
        ret.flags |= se.flags;
    })
    return ret;
}

// Lobe::sample with LobeSet::sample_wi_impl, lobe.cpp:111-119,616-649
// sample_wi_local_impl of the sub-lobe of kind `strategy` (in world space)
__device__ __forceinline__ f3 principled_sub_wi(const PrincipledSet &ps, int strategy, f3 world_wo, Sampler &sampler, bool &valid) {
    f3 wo = ps.frame.to_local(world_wo), wi{0, 0, 1};
    valid = true;
    switch (strategy) {
        case SL_SHEEN: {
            const SubLobe l = sub_lobe<SL_SHEEN>(ps);
            f2 u = sampler.next_2d();
            float r = __fsqrt_rn(u.x), s, c;
            sincos_2pi(u.y, s, c);
            float dx = r * c, dy = r * s;
            f3 w{dx, dy, __fsqrt_rn(fmax_(0.0f, 1.0f - dx * dx - dy * dy))};
            wi = normalize(f3{w.x / l.a - w.z * l.b / l.a, w.y / l.a, w.z});      // M
        } break;
        case SL_TRANS: {
            const SubLobe l = sub_lobe<SL_TRANS>(ps);
            f3 wh = ggx_sample_wh(l.mf, wo, sampler.next_2d());
            float d = dot(wo, wh);
            f3 F = sub_fresnel(l, fabsf(d));
            f3 T = splat(1.f) - F;
            float refl_prob = average3(F) / average3(T * l.kr + F);
            float u1 = sampler.next_1d();
            if (u1 < refl_prob) { wi = reflect(wo, wh); valid = same_hemisphere(wo, wi); }
            else { bool ok = refract(wo, wh, l.eta, wi); valid = ok && !same_hemisphere(wo, wi); }
        } break;
        case SL_DIFFUSE: wi = cosine_hemisphere(sampler.next_2d(), wo.z); break;
        default: {                                                        // coat / metallic / specular: the coat has its own roughness
            const GGX mf = strategy == SL_COAT ? GGX{ps.cc_a, ps.cc_a} : ps.mf;
            f3 wh = ggx_sample_wh(mf, wo, sampler.next_2d());
            wi = reflect(wo, wh);
            valid = same_hemisphere(wo, wi);
        } break;
    }
    return ps.frame.to_world(wi);
}
// samples the direction only (the caller evaluates the set at it)
__device__ __forceinline__ f3 principled_sample_wi(const PrincipledSet &ps, f3 world_wo, Sampler &sampler, bool &valid) {
    float uc = sampler.next_1d();
    (void)sampler.next_2d();                                              // drawn and unused by the reference
    int strategy = -1;                                                    // kind of the chosen lobe (the first present one to begin with)
    float sum_weights = 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k)
        if (ps.switches & (1u << k)) { strategy = (strategy < 0 || uc > sum_weights) ? k : strategy; sum_weights += ps.sw[k]; }
    return principled_sub_wi(ps, strategy, world_wo, sampler, valid);
}

#ifndef VB_PRINCIPLED_INLINE
#define VB_PRINCIPLED_INLINE __forceinline__
#endif
// One call per shaded vertex: create the set once, evaluate it at the light direction, sample a direction, evaluate it there
// (integrator.cpp:20-37,246-283 with MaterialEvaluator on a LobeSet). The two evaluations share ONE copy of the lobe code (a
// two-trip loop), and the set never crosses a call boundary, so it stays in registers.
__device__ VB_PRINCIPLED_INLINE void principled_shade(const PtScene &S, const Swl *swl, const Interaction &it, f3 wi_light, Sampler &sampler,
                                              ScatterEval &se, BSDFSample &bs) {
    const PrincipledSet ps = principled_create(S, swl, it);
    f3 wi = wi_light;
    bool valid = true;
    bs.eta = 1.f;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 1) { wi = principled_sample_wi(ps, it.wo, sampler, valid); bs.wi = wi; }
        ScatterEval e = principled_evaluate(S, ps, it.wo, wi, pass == 1 ? &bs.eta : nullptr);
        if (pass == 0) se = e; else bs.eval = e;
    }
    if (!valid) bs.eval.pdf *= 0.f;
}

// LobeSet::albedo (lobe.cpp:564-570) over the sub-lobes: sheen = tint (principled_bsdf.cpp:54-56), coat / specular = their albedo
// tables at cos_theta (:154-161,198-206), metallic = kr F82(cos), transmission = the dielectric form
__device__ __forceinline__ f3 sub_albedo(const PtScene &S, const SubLobe &l, float cos_theta) {
    f3 a{0, 0, 0};
    float rx = __fsqrt_rn(__fsqrt_rn(l.mf.ax * l.mf.ay));
    switch (l.kind) {
        case SL_SHEEN: case SL_DIFFUSE: a = l.kr; break;
        case SL_COAT: {
            const DevLut &L = S.luts[VB_LUT_COAT];
            float s = (L.data == nullptr) ? 1.f : lut_sample3d_ch0(L, rx, cos_theta, (l.eta - 1.003f) / (4.f - 1.003f));
            a = l.kr * s;
        } break;
        case SL_METALLIC: a = l.kr * sub_fresnel(l, cos_theta); break;
        case SL_TRANS: { f3 F = sub_fresnel(l, fabsf(cos_theta)); a = l.kr * (splat(1.0f) - F) + F; } break;
        case SL_SPEC: {
            const DevLut &L = S.luts[VB_LUT_SPECULAR];
            float z = __fsqrt_rn(fabsf((l.eta - 1.0f) / (l.eta + 1.0f)));
            float s = (L.data == nullptr) ? 1.f : lut_sample3d_ch0(L, rx, cos_theta, z);
            a = lerp3v(s, l.F0, splat(1.f)) * l.kr;
        } break;
    }
    return a;
}
__device__ __noinline__ f3 principled_albedo(const PtScene &S, const Swl *swl, const Interaction &it, float cos_theta) {
    const PrincipledSet ps = principled_create(S, swl, it);
    f3 ret{0, 0, 0};
    VB_EACH_SUBLOBE(ps, { ret = ret + sub_albedo(S, l, cos_theta) * l.weight; })
    return ret;
}

}  // namespace
#include "vb_mixset.cuh"

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
struct PrincipledSet { int n = 0; SubLobe s[6]; Frame3 frame; };

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

__device__ __noinline__ PrincipledSet principled_create(const PtScene &S, const Swl *swl, const Interaction &it) {
    const vb_material_rec &m = S.materials[it.mat_id];
    PrincipledSet ps;
    ps.n = 0;
    ps.frame = it.shading;
    auto clampf = [](float v, float lo, float hi) { return fmin_(fmax_(v, lo), hi); };
    const uint32_t switches = (uint32_t)m.params[15];
    f3 color = material_color(S, swl, m.color, it, SC_COLOR);
    float ior = eval_slot3(S, m.ior, it.uv).x;
    float roughness = clampf(eval_slot3(S, m.roughness, it.uv).x, 0.0001f, 1.f);
    float anisotropic = m.params[0];
    f3 specular_tint = material_color(S, swl, m.spec, it, SC_SPEC);
    float aspect = __fsqrt_rn(1.f - anisotropic * 0.9f);
    GGX mf{fmax_(0.001f, sqr(roughness) / aspect), fmax_(0.001f, sqr(roughness) * aspect)};
    f3 weight = splat(1.f);
    float cos_theta = dot(it.wo, it.ng);
    float front = cos_theta > 0.f ? 1.f : 0.f;
    auto to_ratio_x = [](const GGX &g) { return __fsqrt_rn(__fsqrt_rn(g.ax * g.ay)); };
    if (switches & (1u << SL_SHEEN)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_SHEEN;
        f3 sheen_tint = spec_albedo_c(S.spec, swl, f3{m.params[4], m.params[5], m.params[6]}, true, 4u * it.mat_id + SC_SHEEN);
        float sheen_weight = m.params[2] * front;
        float sheen_roughness = m.params[3];
        l.kr = sheen_tint * sheen_weight * weight;
        const DevLut &L = S.luts[m.params[14] != 0.f ? VB_LUT_SHEEN_APPROX : VB_LUT_SHEEN_VOLUME];
        float cx = 0.f, cy = 0.f, cz = 0.f;
        if (!(L.data == nullptr)) { cx = lut_sample2d_ch(L, cos_theta, sheen_roughness, 0); cy = lut_sample2d_ch(L, cos_theta, sheen_roughness, 1); cz = lut_sample2d_ch(L, cos_theta, sheen_roughness, 2); }
        l.a = cx; l.b = cy;
        l.kr = l.kr * cz;
        f3 albedo = l.kr;
        l.sample_weight = average3(albedo);
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_COAT)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_COAT;
        float cc_weight = m.params[7] * front;
        float cc_roughness = sqr(clampf(m.params[8], 0.0001f, 1.f));
        float cc_ior = m.params[9];
        f3 cc_tint = spec_albedo_c(S.spec, swl, f3{m.params[10], m.params[11], m.params[12]}, true, 4u * it.mat_id + SC_COAT);
        l.fresnel = FR_DIELECTRIC; l.eta = cc_ior;
        l.mf = GGX{cc_roughness, cc_roughness};
        l.kr = weight * cc_weight * cc_tint;
        const DevLut &L = S.luts[VB_LUT_COAT];
        float s = (L.data == nullptr) ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, (cc_ior - 1.003f) / (4.f - 1.003f));
        f3 albedo = l.kr * s;
        l.sample_weight = average3(albedo);
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_METALLIC)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_METALLIC;
        float metallic = m.params[1] * front;
        l.fresnel = FR_F82; l.F0 = color;
        const float f = 6.f / 7.f, f5 = ((f * f) * (f * f)) * f;
        f3 f_schlick = lerp3v(f5, l.F0, splat(1.f));
        l.B = f_schlick * (7.f / (f5 * f)) * (splat(1.f) - specular_tint);
        l.mf = mf;
        l.kr = weight * metallic;
        l.sample_weight = metallic * average3(weight);
        weight = weight * (1.0f - metallic);
    }
    if (switches & (1u << SL_TRANS)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_TRANS;
        float trans_weight = m.params[13];
        float eta = cos_theta > 0.f ? ior : 1.0f / ior;                   // Interaction::correct_eta
        l.fresnel = FR_SCHLICK; l.eta = eta; l.F0 = specular_tint * schlick_F0_from_ior(eta);
        l.mf = mf;
        l.kr = color;
        f3 t_weight = weight * trans_weight;
        l.sample_weight = average3(t_weight); l.weight = average3(t_weight);
        weight = weight * (1.0f - trans_weight);
    }
    if (switches & (1u << SL_SPEC)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_SPEC;
        l.fresnel = FR_SCHLICK; l.eta = ior; l.F0 = specular_tint * schlick_F0_from_ior(ior);
        l.mf = mf;
        l.kr = weight;
        const DevLut &L = S.luts[VB_LUT_SPECULAR];
        float z = __fsqrt_rn(fabsf((ior - 1.0f) / (ior + 1.0f)));
        float s = (L.data == nullptr) ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, z);
        f3 albedo = lerp3v(s, l.F0, splat(1.f)) * l.kr;
        l.sample_weight = average3(albedo);
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_DIFFUSE)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_DIFFUSE;
        l.kr = color * weight * front;
        l.sample_weight = average3(l.kr);
    }
    float sum = 0.f;                                                     // LobeSet::normalize_sampled_weight
    for (int i = 0; i < ps.n; ++i) sum += ps.s[i].sample_weight;
    for (int i = 0; i < ps.n; ++i) ps.s[i].sample_weight = ps.s[i].sample_weight / sum;
    return ps;
}

// SheenLTC (principled_bsdf.cpp:61-117)
__device__ float sheen_eval_ltc(const SubLobe &l, f3 wi) {
    f3 w{l.a * wi.x + l.b * wi.z, l.a * wi.y, wi.z};                // inv_M
    float len = length(w);
    w = w / len;
    float det = sqr(l.a);
    float jacobian = det / (len * len * len);
    return w.z * InvPi * jacobian;
}

__device__ ScatterEval sub_evaluate_local(const PtScene &S, const SubLobe &l, f3 wo, f3 wi, float *eta_out) {
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
__device__ __noinline__ ScatterEval principled_evaluate(const PtScene &S, const PrincipledSet &ps, f3 world_wo, f3 world_wi, float *eta_out) {
    f3 wo = ps.frame.to_local(world_wo), wi = ps.frame.to_local(world_wi);
    ScatterEval ret{f3{0, 0, 0}, 0.f, 0u};
    for (int i = 0; i < ps.n; ++i) {
        const SubLobe &l = ps.s[i];
        ScatterEval se = sub_evaluate_local(S, l, wo, wi, eta_out);
        se.f = se.f * fabsf(wi.z);                                    // Lobe::evaluate_impl
        // valid_world_factor: same side of the shading normal (1 for the dielectric lobe)
        float factor = l.kind == SL_TRANS ? 1.f : (dot(world_wo, ps.frame.z) * dot(world_wi, ps.frame.z) > 0.f ? 1.f : 0.f);
        se.f = se.f * (l.weight * factor);
        se.pdf = se.pdf * (l.sample_weight * factor);
        ret.f = ret.f + se.f;
        ret.pdf = ret.pdf + se.pdf;
        ret.flags |= se.flags;
    }
    return ret;
}

// Lobe::sample with LobeSet::sample_wi_impl, lobe.cpp:111-119,616-649
__device__ __noinline__ BSDFSample principled_sample(const PtScene &S, const PrincipledSet &ps, f3 world_wo, Sampler &sampler) {
    float uc = sampler.next_1d();
    (void)sampler.next_2d();                                              // drawn and unused by the reference
    int strategy = 0;
    float sum_weights = 0.f;
    for (int i = 0; i < ps.n; ++i) { strategy = uc > sum_weights ? i : strategy; sum_weights += ps.s[i].sample_weight; }
    const SubLobe &l = ps.s[strategy];
    f3 wo = ps.frame.to_local(world_wo), wi;
    bool valid = true;
    switch (l.kind) {
        case SL_SHEEN: {
            f2 u = sampler.next_2d();
            float r = __fsqrt_rn(u.x), s, c;
            sincos_2pi(u.y, s, c);
            float dx = r * c, dy = r * s;
            f3 w{dx, dy, __fsqrt_rn(fmax_(0.0f, 1.0f - dx * dx - dy * dy))};
            wi = normalize(f3{w.x / l.a - w.z * l.b / l.a, w.y / l.a, w.z});      // M
        } break;
        case SL_TRANS: {
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
        default: {
            f3 wh = ggx_sample_wh(l.mf, wo, sampler.next_2d());
            wi = reflect(wo, wh);
            valid = same_hemisphere(wo, wi);
        } break;
    }
    BSDFSample bs;
    bs.eta = 1.f;
    bs.wi = ps.frame.to_world(wi);
    bs.eval = principled_evaluate(S, ps, world_wo, bs.wi, &bs.eta);
    if (!valid) bs.eval.pdf *= 0.f;
    return bs;
}

// LobeSet::albedo (lobe.cpp:564-570) over the sub-lobes: sheen = tint (principled_bsdf.cpp:54-56), coat / specular = their albedo
// tables at cos_theta (:154-161,198-206), metallic = kr F82(cos), transmission = the dielectric form
__device__ __noinline__ f3 principled_albedo(const PtScene &S, const PrincipledSet &ps, float cos_theta) {
    f3 ret{0, 0, 0};
    for (int i = 0; i < ps.n; ++i) {
        const SubLobe &l = ps.s[i];
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
        ret = ret + a * l.weight;
    }
    return ret;
}

}  // namespace

// Lobe evaluation / sampling / construction (device). Second half of vb_bsdf.cuh; mirrors the
// corresponding part of oracle/vo_bsdf.inl statement by statement.
#pragma once

namespace {

// the lobe of one surface hit (the reference builds exactly one Lobe per material on this path)
struct Lobe {
    uint32_t type;
    Frame3 frame;              // Material::compute_shading_frame without a normal map
    f3 kr;                     // diffuse Kr | mirror kr | glass kt | substrate Rd
    f3 rs;                     // substrate Rs | metal eta | plastic / metallic Fresnel F0
    f3 kk;                     // metal k | metallic F82-tint B
    GGX mf;
    float eta;                 // glass: corrected relative ior (Interaction::correct_eta); dispersive: of the hero wavelength (eta[0])
    f3 eta3;                   // dispersive glass (hero spectrum, ior slot = "spd", glass.cpp:221-234): corrected ior per wavelength
    bool dispersive;
    bool oren_nayar; float on_A, on_B;
};

// FresnelSchlick::evaluate (fresnel.h:44-52) and FresnelF82Tint::evaluate (fresnel.h:118-124) over plain colour triples
__device__ __forceinline__ f3 fresnel_schlick3(f3 F0, float eta, float cos_theta) {
    float F_real = fresnel_dielectric(cos_theta, eta);
    float F0_real = sqr((eta - 1.0f) / (eta + 1.0f));
    float t = (F_real - F0_real) / (1.f - F0_real);
    t = fmin_(fmax_(t, 0.f), 1.f);
    return {lerp(t, F0.x, 1.f), lerp(t, F0.y, 1.f), lerp(t, F0.z, 1.f)};
}
__device__ __forceinline__ f3 fresnel_f82(f3 F0, f3 B, float cos_theta) {
    float mu = fmin_(fmax_(1.f - cos_theta, 0.f), 1.f);
    float mu5 = pow5(mu);
    f3 f_schlick{lerp(mu5, F0.x, 1.f), lerp(mu5, F0.y, 1.f), lerp(mu5, F0.z, 1.f)};
    f3 r = f_schlick - B * cos_theta * mu5 * mu;
    return {fmin_(fmax_(r.x, 0.f), 1.f), fmin_(fmax_(r.y, 0.f), 1.f), fmin_(fmax_(r.z, 0.f), 1.f)};
}

__device__ f3 lobe_fresnel(const Lobe &lb, float cos_theta) {
    switch (lb.type) {
        case VB_MAT_PLASTIC: return fresnel_schlick3(lb.rs, lb.eta, cos_theta);
        case VB_MAT_METALLIC: return fresnel_f82(lb.rs, lb.kk, cos_theta);
        case VB_MAT_METAL: return {fresnel_complex(cos_theta, lb.rs.x, lb.kk.x), fresnel_complex(cos_theta, lb.rs.y, lb.kk.y),
                                   fresnel_complex(cos_theta, lb.rs.z, lb.kk.z)};
        case VB_MAT_GLASS: return splat(fresnel_dielectric(cos_theta, lb.eta));
        case VB_MAT_SUBSTRATE: return splat(fresnel_dielectric(cos_theta, 1.5f));
        default: return splat(1.f);
    }
}

// -- DielectricLobe, lobe.cpp:253-449 --
__device__ float dielectric_compensate(const PtScene &S, const Lobe &lb, f3 wo) {
    const DevLut &L = lb.eta > 1.0f ? S.luts[VB_LUT_DIELECTRIC] : S.luts[VB_LUT_DIELECTRIC_INV];
    if (!L.data) return 1.0f;
    float x = __fsqrt_rn(__fsqrt_rn(lb.mf.ax * lb.mf.ay));
    float y = fabsf(wo.z);
    const float lo = 1.003f, hi = 5.0f;
    float z = lb.eta > 1.0f ? (lb.eta - lo) / (hi - lo) : (1.0f / lb.eta - lo) / (hi - lo);
    return 1.0f / lut_sample3d_ch0(L, x, y, z);
}
// FresnelDielectric::evaluate (fresnel.h:84-89): a scalar ior slot gives a 1-dimensional spectrum (F.average() == F),
// the dispersive "spd" slot one value per wavelength
__device__ __forceinline__ f3 dielectric_F(const Lobe &lb, float cos_theta) {
    if (!lb.dispersive) return splat(fresnel_dielectric(cos_theta, lb.eta));
    return {fresnel_dielectric(cos_theta, lb.eta3.x), fresnel_dielectric(cos_theta, lb.eta3.y), fresnel_dielectric(cos_theta, lb.eta3.z)};
}
__device__ __forceinline__ float dielectric_refl_prob(const Lobe &lb, f3 F) {
    f3 T = splat(1.0f) - F;
    f3 total = lb.kr * T + F;
    return (lb.dispersive ? average3(F) : F.x) / average3(total);
}
__device__ ScatterEval dielectric_evaluate_local(const PtScene &S, const Lobe &lb, f3 wo, f3 wi, bool flip_wh, float *eta_out) {
    bool reflect_ = same_hemisphere(wo, wi);
    float eta_p = reflect_ ? 1.0f : lb.eta;
    if (eta_out) *eta_out = eta_p;
    f3 wh = normalize(wo + wi * eta_p);
    if (flip_wh) wh = face_forward(wh, wo);
    f3 F = dielectric_F(lb, abs_dot(wh, wo));
    ScatterEval se;
    if (reflect_) {
        se.f = F * ggx_brdf_div_fr(lb.mf, wo, wh, wi);
        se.pdf = ggx_pdf_wi_reflection(lb.mf, wo, wh) * dielectric_refl_prob(lb, F);
        se.flags = FlagGlossy | FlagReflection;
        se.f = se.f * dielectric_compensate(S, lb, wo);
    } else {
        f3 new_wh = face_forward(wh, wo);
        f3 wh2 = normalize(wo + wi * lb.eta);
        f3 tr = (splat(1.0f) - F) * ggx_btdf_div_ft(lb.mf, wo, wh2, wi, lb.eta);
        se.f = lb.kr * tr;
        se.pdf = ggx_pdf_wi_transmission(lb.mf, wo, new_wh, wi, lb.eta) * (1.0f - dielectric_refl_prob(lb, F));
        se.flags = FlagGlossy | FlagTransmission;
        se.f = se.f * dielectric_compensate(S, lb, wo);
    }
    return se;
}

// -- MicrofacetReflection (mirror, metal), bxdf.cpp:65-85 + PureReflectionLobe lobe.cpp:716-730 --
__device__ ScatterEval microfacet_refl_evaluate_local(const PtScene &S, const Lobe &lb, f3 wo, f3 wi) {
    ScatterEval se;
    bool s = same_hemisphere(wo, wi);
    f3 wh = normalize(wo + wi);
    f3 whf = face_forward(wh, f3{0, 0, 1});
    f3 F = lobe_fresnel(lb, abs_dot(wo, whf));
    f3 fr = F * ggx_brdf_div_fr(lb.mf, wo, whf, wi);
    f3 f = fr * lb.kr;
    float pdf = ggx_pdf_wi_reflection(lb.mf, wo, wh);
    se.f = s ? f : f3{0, 0, 0};
    se.pdf = s ? pdf : 0.f;
    se.flags = FlagGlossy | FlagReflection;
    const DevLut &L = S.luts[VB_LUT_PURE_REFLECTION];
    if (L.data) {
        float alpha = __fsqrt_rn(lb.mf.ax * lb.mf.ay);
        se.f = se.f * (1.0f / lut_sample2d(L, alpha, wo.z));
    }
    return se;
}

// -- FresnelBlend (substrate), substrate.cpp:12-70 --
__device__ ScatterEval fresnel_blend_evaluate_local(const Lobe &lb, f3 wo, f3 wi) {
    ScatterEval se;
    bool s = same_hemisphere(wo, wi);
    float aci = fabsf(wi.z), aco = fabsf(wo.z);
    f3 diffuse = lb.kr * (28.f / (23.f * Pi)) * (splat(1.f) - lb.rs) * (1.0f - pow5(1.0f - .5f * aci)) * (1.0f - pow5(1.0f - .5f * aco));
    f3 wh = normalize(wi + wo);
    float sw = pow5(fmin_(fmax_(1.0f - dot(wi, wh), 0.f), 1.f));
    f3 schlick{lerp(sw, lb.rs.x, 1.f), lerp(sw, lb.rs.y, 1.f), lerp(sw, lb.rs.z, 1.f)};
    f3 specular = schlick * (ggx_D(lb.mf, wh) / (4.0f * abs_dot(wi, wh) * fmax_(aci, aco)));
    if (is_zero(wh)) specular = specular * 0.f;
    f3 f = specular + diffuse;
    float fr = fresnel_dielectric(aco, 1.5f);
    float pdf = lerp(fr, aci * InvPi, ggx_pdf_wi_reflection(lb.mf, wo, normalize(wo + wi)));
    se.f = s ? f : f3{0, 0, 0};
    se.pdf = s ? pdf : 0.f;
    se.flags = FlagReflection;
    return se;
}

// -- Lambert / Oren-Nayar, bxdf.cpp:13-46,94-121 --
__device__ ScatterEval diffuse_evaluate_local(const Lobe &lb, f3 wo, f3 wi) {
    ScatterEval se;
    bool s = same_hemisphere(wo, wi);
    f3 f = lb.kr * InvPi;
    if (lb.oren_nayar) {
        float sti = __fsqrt_rn(fmax_(0.f, 1.0f - wi.z * wi.z)), sto = __fsqrt_rn(fmax_(0.f, 1.0f - wo.z * wo.z));
        float cpi = sti == 0.f ? 1.f : fmin_(fmax_(wi.x / sti, -1.f), 1.f), spi = sti == 0.f ? 0.f : fmin_(fmax_(wi.y / sti, -1.f), 1.f);
        float cpo = sto == 0.f ? 1.f : fmin_(fmax_(wo.x / sto, -1.f), 1.f), spo = sto == 0.f ? 0.f : fmin_(fmax_(wo.y / sto, -1.f), 1.f);
        float d_cos = cpi * cpo + spi * spo;
        float max_cos = fmax_(0.f, d_cos);
        bool cond = fabsf(wi.z) > fabsf(wo.z);
        float sin_alpha = cond ? sto : sti;
        float tan_beta = cond ? sti / fabsf(wi.z) : sto / fabsf(wo.z);
        f = lb.kr * InvPi * (lb.on_A + lb.on_B * max_cos * sin_alpha * tan_beta);
    }
    se.f = s ? f : f3{0, 0, 0};
    se.pdf = s ? fabsf(wi.z) * InvPi : 0.f;
    se.flags = FlagDiffuse | FlagReflection;
    return se;
}

// -- PlasticLobe::evaluate_local_impl, plastic.cpp:37-50 (no hemisphere test of its own; flags stay at the ScatterEval default) --
__device__ ScatterEval plastic_evaluate_local(const Lobe &lb, f3 wo, f3 wi) {
    ScatterEval se;
    f3 wh = normalize(wo + wi);
    f3 F = lobe_fresnel(lb, abs_dot(wh, wo));
    se.f = lb.kr * InvPi * (splat(1.f) - F);
    se.f = se.f + F * ggx_brdf_div_fr(lb.mf, wo, wh, wi);
    se.pdf = lerp(average3(F), fabsf(wi.z) * InvPi, ggx_pdf_wi_reflection(lb.mf, wo, wh));
    se.flags = 1u;
    return se;
}

// Lobe::evaluate_impl (lobe.cpp:53-77): world -> local, f *= |cos(wi)|
__device__ __forceinline__ ScatterEval lobe_evaluate(const PtScene &S, const Lobe &lb, f3 world_wo, f3 world_wi, bool with_eta, float *eta_out) {
    f3 wo = lb.frame.to_local(world_wo), wi = lb.frame.to_local(world_wi);
    ScatterEval se;
    switch (lb.type) {
        case VB_MAT_GLASS: se = dielectric_evaluate_local(S, lb, wo, wi, with_eta, eta_out); break;
        case VB_MAT_MIRROR: case VB_MAT_METAL: case VB_MAT_METALLIC: se = microfacet_refl_evaluate_local(S, lb, wo, wi); break;
        case VB_MAT_SUBSTRATE: se = fresnel_blend_evaluate_local(lb, wo, wi); break;
        case VB_MAT_PLASTIC: se = plastic_evaluate_local(lb, wo, wi); break;
        default: se = diffuse_evaluate_local(lb, wo, wi); break;
    }
    se.f = se.f * fabsf(wi.z);
    return se;
}

__device__ __forceinline__ f3 cosine_hemisphere(f2 u, float woz) {
    float r = __fsqrt_rn(u.x), s, c;
    sincos_2pi(u.y, s, c);
    float dx = r * c, dy = r * s;
    float z = __fsqrt_rn(fmax_(0.0f, 1.0f - dx * dx - dy * dy));
    if (woz < 0.f) z = -z;
    return {dx, dy, z};
}

// Lobe::sample (lobe.cpp:111-119) with each lobe's sample_wi_local_impl; draws in the reference's order
// the direction only (sample_wi_local_impl of the lobe, in world space); `valid` = SampledDirection::valid
__device__ __forceinline__ f3 lobe_sample_wi(const Lobe &lb, f3 world_wo, Sampler &sampler, bool &valid) {
    f3 wo = lb.frame.to_local(world_wo);
    f3 wi;
    valid = true;
    switch (lb.type) {
        case VB_MAT_GLASS: {
            f3 wh = ggx_sample_wh(lb.mf, wo, sampler.next_2d());
            float d = dot(wo, wh);
            f3 F = dielectric_F(lb, fabsf(d));
            float uc = sampler.next_1d();
            if (uc < dielectric_refl_prob(lb, F)) {
                wi = reflect(wo, wh);
                valid = same_hemisphere(wo, wi);
            } else {
                bool ok = refract(wo, wh, lb.eta, wi);
                valid = ok && !same_hemisphere(wo, wi);
            }
        } break;
        case VB_MAT_PLASTIC: {                                         // plastic.cpp:52-66
            f3 wh = ggx_sample_wh(lb.mf, wo, sampler.next_2d());
            f3 F = lobe_fresnel(lb, fabsf(wo.z));
            float uc = sampler.next_1d();
            if (uc < average3(F)) {
                wi = reflect(wo, wh);
                valid = same_hemisphere(wo, wi);
            } else wi = cosine_hemisphere(sampler.next_2d(), 1.f);     // square_to_cosine_hemisphere: always the +z side
        } break;
        case VB_MAT_MIRROR: case VB_MAT_METAL: case VB_MAT_METALLIC: {
            f3 wh = ggx_sample_wh(lb.mf, wo, sampler.next_2d());
            wi = reflect(wo, wh);
            valid = same_hemisphere(wo, wi);
        } break;
        case VB_MAT_SUBSTRATE: {
            f2 u = sampler.next_2d();
            float fr = fresnel_dielectric(fabsf(wo.z), 1.5f);
            if (u.x < fr) {
                u.x = (u.x - 0.f) / (fr - 0.f);
                f3 wh = ggx_sample_wh(lb.mf, wo, u);
                wi = reflect(wo, wh);
            } else {
                u.x = (u.x - fr) / (1.f - fr);
                wi = cosine_hemisphere(u, wo.z);
            }
        } break;
        default: wi = cosine_hemisphere(sampler.next_2d(), wo.z); break;
    }
    return lb.frame.to_world(wi);
}
__device__ __forceinline__ BSDFSample lobe_sample(const PtScene &S, const Lobe &lb, f3 world_wo, Sampler &sampler) {
    bool valid = true;
    BSDFSample bs;
    bs.eta = 1.f;
    bs.wi = lobe_sample_wi(lb, world_wo, sampler, valid);
    bs.eval = lobe_evaluate(S, lb, world_wo, bs.wi, true, &bs.eta);
    if (!valid) bs.eval.pdf *= 0.f;
    return bs;
}

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fmin_(fmax_(v, lo), hi); }

// a colour slot of the hit's material as an albedo spectrum (eval_albedo_spectrum, shader_node.cpp:317-321)
__device__ __forceinline__ f3 material_color_of(const PtScene &S, const Swl *swl, const vb_slot &slot, const Interaction &it, uint32_t mat_id, uint32_t which) {
    return spec_albedo_c(S.spec, swl, eval_slot3(S, slot, it.uv), slot.tex_id == VB_INVALID_ID, 4u * mat_id + which);
}
__device__ __forceinline__ f3 material_color(const PtScene &S, const Swl *swl, const vb_slot &slot, const Interaction &it, uint32_t which) {
    return material_color_of(S, swl, slot, it, it.mat_id, which);
}

// Material::create_evaluator -> create_lobe_set of the plugin
__device__ __forceinline__ Lobe create_lobe_of(const PtScene &S, const Swl *swl, const Interaction &it, const uint32_t mat_id) {
    const vb_material_rec &m = S.materials[mat_id];
    Lobe lb;
    lb.type = m.type;
    lb.frame = it.shading;
    lb.kr = f3{0, 0, 0}; lb.rs = f3{0, 0, 0}; lb.kk = f3{0, 0, 0};
    lb.mf = GGX{1.f, 1.f}; lb.eta = 1.f; lb.eta3 = f3{1, 1, 1}; lb.dispersive = false; lb.oren_nayar = false; lb.on_A = 1.f; lb.on_B = 0.f;
    bool remap = (m.flags & 2u) != 0;
    float anis = clampf(m.params[0], -0.9f, 0.9f);
    switch (m.type) {
        case VB_MAT_DIFFUSE:
            lb.kr = material_color_of(S, swl, m.color, it, mat_id, SC_COLOR);
            if (m.flags & 1u) {
                float sigma = eval_slot3(S, m.sigma, it.uv).x * PiOver2;
                float sigma2 = sqr(sigma * sigma);
                lb.oren_nayar = true;
                lb.on_A = 1.f - (sigma2 / (2.f * (sigma2 + 0.33f)));
                lb.on_B = 0.45f * sigma2 / (sigma2 + 0.09f);
            }
            break;
        case VB_MAT_MIRROR: case VB_MAT_METAL: {
            lb.kr = m.type == VB_MAT_MIRROR ? material_color_of(S, swl, m.color, it, mat_id, SC_COLOR) : splat(1.f);
            float r = clampf(eval_slot3(S, m.roughness, it.uv).x, 0.0001f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            lb.rs = f3{m.eta[0], m.eta[1], m.eta[2]}; lb.kk = f3{m.k[0], m.k[1], m.k[2]};
            if (m.type == VB_MAT_METAL && (m.flags & 4u) && swl) {       // metal.cpp:117-121: "spd" eta / k slots
                lb.rs = spd_eval3(S.spec, m.spd[0], *swl); lb.kk = spd_eval3(S.spec, m.spd[1], *swl);
            }
        } break;
        case VB_MAT_GLASS: {
            lb.kr = material_color_of(S, swl, m.color, it, mat_id, SC_COLOR);
            float ior = eval_slot3(S, m.ior, it.uv).x;
            lb.eta = dot(it.wo, it.ng) > 0.f ? ior : 1.0f / ior;
            if ((m.flags & 4u) && swl) {                                 // glass.cpp:216-234: "material_name" curve in an "spd" slot
                f3 iors = spd_eval3(S.spec, m.spd[0], *swl);
                bool front = dot(it.wo, it.ng) > 0.f;
                lb.eta3 = front ? iors : f3{1.0f / iors.x, 1.0f / iors.y, 1.0f / iors.z};
                lb.eta = lb.eta3.x;
                lb.dispersive = true;
            }
            float r = clampf(eval_slot3(S, m.roughness, it.uv).x, 0.01f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
        } break;
        case VB_MAT_SUBSTRATE: {
            lb.kr = material_color_of(S, swl, m.color, it, mat_id, SC_COLOR);
            lb.rs = material_color_of(S, swl, m.spec, it, mat_id, SC_SPEC);
            float r = clampf(eval_slot3(S, m.roughness, it.uv).x, 0.0001f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            if (remap) { lb.mf.ax = sqr(lb.mf.ax); lb.mf.ay = sqr(lb.mf.ay); }
            lb.mf.ax = clampf(lb.mf.ax, 0.0001f, 1.f); lb.mf.ay = clampf(lb.mf.ay, 0.0001f, 1.f);
        } break;
        case VB_MAT_PLASTIC: {                                         // plastic.cpp:109-129
            lb.kr = material_color_of(S, swl, m.color, it, mat_id, SC_COLOR);
            f3 Rs = material_color_of(S, swl, m.spec, it, mat_id, SC_SPEC);
            float ior = eval_slot3(S, m.ior, it.uv).x;
            float r = clampf(eval_slot3(S, m.roughness, it.uv).x, 0.0001f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            if (remap) { lb.mf.ax = sqr(lb.mf.ax); lb.mf.ay = sqr(lb.mf.ay); }
            lb.mf.ax = clampf(lb.mf.ax, 0.0001f, 1.f); lb.mf.ay = clampf(lb.mf.ay, 0.0001f, 1.f);
            lb.eta = ior;
            lb.rs = Rs * sqr((ior - 1.0f) / (ior + 1.0f));             // schlick_F0_from_ior(iors[0]) * Rs
        } break;
        case VB_MAT_METALLIC: {                                        // metallic.cpp:44-60
            lb.kr = material_color_of(S, swl, m.color, it, mat_id, SC_COLOR);
            f3 edge_tint = material_color_of(S, swl, m.spec, it, mat_id, SC_SPEC);
            float r = clampf(eval_slot3(S, m.roughness, it.uv).x, 0.01f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            lb.rs = lb.kr;                                             // FresnelF82Tint(color) + init_from_F82(edge_tint), fresnel.h:110-116
            const float f = 6.f / 7.f, f5 = ((f * f) * (f * f)) * f;
            f3 f_schlick{lerp(f5, lb.rs.x, 1.f), lerp(f5, lb.rs.y, 1.f), lerp(f5, lb.rs.z, 1.f)};
            lb.kk = f_schlick * (7.f / (f5 * f)) * (splat(1.f) - edge_tint);
        } break;
        default: break;
    }
    return lb;
}
__device__ __forceinline__ Lobe create_lobe(const PtScene &S, const Swl *swl, const Interaction &it) { return create_lobe_of(S, swl, it, it.mat_id); }


// Lobe::albedo(cos_theta) of the single-lobe materials (lobe.h:167, lobe.cpp:208-210,308-313, substrate.cpp:22): rt_geom's albedo plane
__device__ __forceinline__ f3 lobe_albedo(const Lobe &lb, float cos_theta) {
    switch (lb.type) {
        case VB_MAT_DIFFUSE: case VB_MAT_SUBSTRATE: return lb.kr;
        case VB_MAT_MIRROR: case VB_MAT_METAL: case VB_MAT_METALLIC: return lb.kr * lobe_fresnel(lb, cos_theta);
        case VB_MAT_PLASTIC: return lobe_fresnel(lb, cos_theta);      // MicrofacetLobe::albedo with a unit reflection colour (plastic.cpp:127)
        case VB_MAT_GLASS: { f3 F = dielectric_F(lb, fabsf(cos_theta)); return lb.kr * (splat(1.0f) - F) + F; }
        default: return f3{0, 0, 0};
    }
}

// MaterialEvaluator (individual_ns branch), material.cpp:132-142,163-175
__device__ __forceinline__ void hemisphere_discard(ScatterEval &se, f3 wo, f3 wi, f3 ng) {
    bool same = dot(wo, ng) * dot(wi, ng) > 0.f;
    bool is_trans = (se.flags & FlagTransmission) != 0;
    if (same == is_trans) se.pdf = 0.f;
}

}  // namespace
#include "vb_principled.cuh"

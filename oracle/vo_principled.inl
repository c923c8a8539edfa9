// ORACLE — test infrastructure only. PrincipledBSDF (reference src/render_core/material/principled_bsdf.cpp:17-461)
// as a flattened LobeSet (reference src/base/scattering/lobe.cpp:454-713): sheen (LTC), coat, metallic (F82-tint),
// transmission (rough dielectric with Schlick Fresnel), specular, diffuse, layered by albedo-scaled weights.
// Included by vo_bsdf.inl inside namespace vo::{anonymous}; vision_b200/csrc/vb_principled.cuh mirrors it.

inline float lut_sample2d_ch(const Lut &L, float u, float v, uint32_t ch) {
    LutAxis a = lut_axis(u, L.nx), b = lut_axis(v, L.ny);
    auto at = [&](int x, int y) { return L.data[((size_t)y * L.nx + x) * L.channels + ch]; };
    float r0 = at(a.i0, b.i0) + a.f * (at(a.i1, b.i0) - at(a.i0, b.i0));
    float r1 = at(a.i0, b.i1) + a.f * (at(a.i1, b.i1) - at(a.i0, b.i1));
    return r0 + b.f * (r1 - r0);
}

enum : int { SL_SHEEN = 0, SL_COAT = 1, SL_METALLIC = 2, SL_TRANS = 3, SL_SPEC = 4, SL_DIFFUSE = 5 };
enum : int { FR_DIELECTRIC = 0, FR_SCHLICK = 1, FR_F82 = 2 };

struct SubLobe {
    int kind = SL_DIFFUSE;
    float sample_weight = 0.f, weight = 1.f;      // WeightedLobe
    float3 kr{0, 0, 0};                            // reflection colour | transmission kt | sheen tint | diffuse Kr
    GGX mf{1.f, 1.f};
    int fresnel = FR_DIELECTRIC;
    float3 F0{0, 0, 0}, B{0, 0, 0};
    float eta = 1.f;
    float a = 1.f, b = 0.f;                        // sheen LTC coefficients
};
struct PrincipledSet { int n = 0; SubLobe s[6]; Frame frame; };

inline float saturate(float x) { return std::min(std::max(x, 0.f), 1.f); }
inline float3 lerp3v(float t, float3 a, float3 b) { return {lerp(t, a.x, b.x), lerp(t, a.y, b.y), lerp(t, a.z, b.z)}; }
inline float schlick_F0_from_ior(float ior) { return sqr((ior - 1.0f) / (ior + 1.0f)); }

// Fresnel::evaluate of the sub-lobe (fresnel.h:44-52 Schlick, :60-72 dielectric, :96-104 F82-tint)
inline float3 sub_fresnel(const SubLobe &l, float cos_theta) {
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
        float3 f_schlick = lerp3v(mu5, l.F0, splat(1.f));
        float3 r = f_schlick - l.B * cos_theta * mu5 * mu;
        return {saturate(r.x), saturate(r.y), saturate(r.z)};
    }
    return splat(fresnel_dielectric(cos_theta, l.eta));
}

// layering_weight, principled_bsdf.cpp:209-214
inline float3 layering_weight(float3 layer_albedo, float3 weight) {
    auto safe_div = [](float a, float b) { return b == 0.f ? 0.f : a / b; };
    float3 tmp{safe_div(layer_albedo.x, weight.x), safe_div(layer_albedo.y, weight.y), safe_div(layer_albedo.z, weight.z)};
    float max_comp = std::max(tmp.x, std::max(tmp.y, tmp.z));
    return weight * saturate(1.f - max_comp);
}

inline PrincipledSet principled_create_of(const Scene &sc, const Swl &swl, const Interaction &it, uint32_t mat_id);
inline PrincipledSet principled_create(const Scene &sc, const Swl &swl, const Interaction &it) { return principled_create_of(sc, swl, it, it.mat_id); }
inline PrincipledSet principled_create_of(const Scene &sc, const Swl &swl, const Interaction &it, uint32_t mat_id) {
    const vb_material_rec &m = sc.materials[mat_id];
    PrincipledSet ps;
    ps.frame = it.shading;
    auto clampf = [](float v, float lo, float hi) { return std::min(std::max(v, lo), hi); };
    const uint32_t switches = (uint32_t)m.params[15];
    float3 color = spec_albedo(sc, swl, eval_slot3(sc, m.color, it.uv));
    float ior = eval_slot3(sc, m.ior, it.uv).x;
    float roughness = clampf(eval_slot3(sc, m.roughness, it.uv).x, 0.0001f, 1.f);
    float anisotropic = m.params[0];
    float3 specular_tint = spec_albedo(sc, swl, eval_slot3(sc, m.spec, it.uv));
    float aspect = std::sqrt(1.f - anisotropic * 0.9f);
    GGX mf{std::max(0.001f, sqr(roughness) / aspect), std::max(0.001f, sqr(roughness) * aspect)};
    float3 weight = splat(1.f);
    float cos_theta = dot(it.wo, it.ng);
    float front = cos_theta > 0.f ? 1.f : 0.f;
    auto to_ratio_x = [](const GGX &g) { return std::sqrt(std::sqrt(g.ax * g.ay)); };
    if (switches & (1u << SL_SHEEN)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_SHEEN;
        float3 sheen_tint = spec_albedo(sc, swl, float3{m.params[4], m.params[5], m.params[6]});
        float sheen_weight = m.params[2] * front;
        float sheen_roughness = m.params[3];
        l.kr = sheen_tint * sheen_weight * weight;
        const Lut &L = sc.luts[m.params[14] != 0.f ? VB_LUT_SHEEN_APPROX : VB_LUT_SHEEN_VOLUME];
        float cx = 0.f, cy = 0.f, cz = 0.f;
        if (!L.data.empty()) { cx = lut_sample2d_ch(L, cos_theta, sheen_roughness, 0); cy = lut_sample2d_ch(L, cos_theta, sheen_roughness, 1); cz = lut_sample2d_ch(L, cos_theta, sheen_roughness, 2); }
        l.a = cx; l.b = cy;
        l.kr = l.kr * cz;
        float3 albedo = l.kr;
        l.sample_weight = average3(albedo);
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_COAT)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_COAT;
        float cc_weight = m.params[7] * front;
        float cc_roughness = sqr(clampf(m.params[8], 0.0001f, 1.f));
        float cc_ior = m.params[9];
        float3 cc_tint = spec_albedo(sc, swl, float3{m.params[10], m.params[11], m.params[12]});
        l.fresnel = FR_DIELECTRIC; l.eta = cc_ior;
        l.mf = GGX{cc_roughness, cc_roughness};
        l.kr = weight * cc_weight * cc_tint;
        const Lut &L = sc.luts[VB_LUT_COAT];
        float s = L.data.empty() ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, (cc_ior - 1.003f) / (4.f - 1.003f));
        float3 albedo = l.kr * s;
        l.sample_weight = average3(albedo);
        weight = layering_weight(albedo, weight);
    }
    if (switches & (1u << SL_METALLIC)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_METALLIC;
        float metallic = m.params[1] * front;
        l.fresnel = FR_F82; l.F0 = color;
        const float f = 6.f / 7.f, f5 = ((f * f) * (f * f)) * f;
        float3 f_schlick = lerp3v(f5, l.F0, splat(1.f));
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
        float3 t_weight = weight * trans_weight;
        l.sample_weight = average3(t_weight); l.weight = average3(t_weight);
        weight = weight * (1.0f - trans_weight);
    }
    if (switches & (1u << SL_SPEC)) {
        SubLobe &l = ps.s[ps.n++];
        l.kind = SL_SPEC;
        l.fresnel = FR_SCHLICK; l.eta = ior; l.F0 = specular_tint * schlick_F0_from_ior(ior);
        l.mf = mf;
        l.kr = weight;
        const Lut &L = sc.luts[VB_LUT_SPECULAR];
        float z = std::sqrt(std::fabs((ior - 1.0f) / (ior + 1.0f)));
        float s = L.data.empty() ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, z);
        float3 albedo = lerp3v(s, l.F0, splat(1.f)) * l.kr;
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
inline float sheen_eval_ltc(const SubLobe &l, float3 wi) {
    float3 w{l.a * wi.x + l.b * wi.z, l.a * wi.y, wi.z};                // inv_M
    float len = length(w);
    w = w / len;
    float det = sqr(l.a);
    float jacobian = det / (len * len * len);
    return w.z * InvPi * jacobian;
}

inline ScatterEval sub_evaluate_local(const Scene &sc, const SubLobe &l, float3 wo, float3 wi, float *eta_out) {
    ScatterEval se;
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
            float3 wh = normalize(wo + wi);
            float3 whf = face_forward(wh, float3{0, 0, 1});
            float3 F = sub_fresnel(l, abs_dot(wo, whf));
            float3 f = F * ggx_brdf_div_fr(l.mf, wo, whf, wi) * l.kr;
            float pdf = ggx_pdf_wi_reflection(l.mf, wo, wh);
            se.f = s ? f : float3{0, 0, 0};
            se.pdf = s ? pdf : 0.f;
            se.flags = FlagGlossy | FlagReflection;
            if (l.kind == SL_METALLIC) {                                  // MetallicLobe::compensate() == true
                const Lut &L = sc.luts[VB_LUT_PURE_REFLECTION];
                if (!L.data.empty()) se.f = se.f * (1.0f / lut_sample2d(L, std::sqrt(l.mf.ax * l.mf.ay), wo.z));
            }
        } break;
        case SL_TRANS: {                                                  // DielectricLobe with a Schlick Fresnel, lobe.cpp:318-404
            bool reflect_ = same_hemisphere(wo, wi);
            float eta_p = reflect_ ? 1.0f : l.eta;
            if (eta_out) *eta_out = eta_p;
            float3 wh = normalize(wo + wi * eta_p);
            wh = face_forward(wh, wo);
            float3 F = sub_fresnel(l, abs_dot(wh, wo));
            float3 T = splat(1.f) - F;
            float refl_prob = average3(F) / average3(T * l.kr + F);
            const Lut &L = l.eta > 1.0f ? sc.luts[VB_LUT_DIELECTRIC] : sc.luts[VB_LUT_DIELECTRIC_INV];
            float comp = 1.f;
            if (!L.data.empty()) {
                float z = l.eta > 1.0f ? (l.eta - 1.003f) / (5.0f - 1.003f) : (1.0f / l.eta - 1.003f) / (5.0f - 1.003f);
                comp = 1.0f / lut_sample3d_ch0(L, std::sqrt(std::sqrt(l.mf.ax * l.mf.ay)), std::fabs(wo.z), z);
            }
            if (reflect_) {
                se.f = F * ggx_brdf_div_fr(l.mf, wo, wh, wi);
                se.pdf = ggx_pdf_wi_reflection(l.mf, wo, wh) * refl_prob;
                se.flags = FlagGlossy | FlagReflection;
            } else {
                float3 new_wh = face_forward(wh, wo);
                float3 wh2 = normalize(wo + wi * l.eta);
                se.f = T * ggx_btdf_div_ft(l.mf, wo, wh2, wi, l.eta) * l.kr;
                se.pdf = ggx_pdf_wi_transmission(l.mf, wo, new_wh, wi, l.eta) * (1.0f - refl_prob);
                se.flags = FlagGlossy | FlagTransmission;
            }
            se.f = se.f * comp;
        } break;
        default: {
            bool s = same_hemisphere(wo, wi);
            se.f = s ? l.kr * InvPi : float3{0, 0, 0};
            se.pdf = s ? std::fabs(wi.z) * InvPi : 0.f;
            se.flags = FlagDiffuse | FlagReflection;
        } break;
    }
    return se;
}

// LobeSet::evaluate_impl (world), lobe.cpp:667-683
inline ScatterEval principled_evaluate(const Scene &sc, const PrincipledSet &ps, float3 world_wo, float3 world_wi, float *eta_out) {
    float3 wo = ps.frame.to_local(world_wo), wi = ps.frame.to_local(world_wi);
    ScatterEval ret;
    ret.flags = 0;
    for (int i = 0; i < ps.n; ++i) {
        const SubLobe &l = ps.s[i];
        ScatterEval se = sub_evaluate_local(sc, l, wo, wi, eta_out);
        se.f = se.f * std::fabs(wi.z);                                    // Lobe::evaluate_impl
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

// sample_wi_local_impl of one sub-lobe (in world space)
inline float3 sub_sample_wi(const SubLobe &l, const Frame &frame, float3 world_wo, Sampler &sampler, bool &valid) {
    float3 wo = frame.to_local(world_wo), wi;
    valid = true;
    switch (l.kind) {
        case SL_SHEEN: {
            float2 u = sampler.next_2d();
            float r = std::sqrt(u.x), s, c;
            sincos_2pi(u.y, s, c);
            float dx = r * c, dy = r * s;
            float3 w{dx, dy, std::sqrt(std::max(0.0f, 1.0f - dx * dx - dy * dy))};
            wi = normalize(float3{w.x / l.a - w.z * l.b / l.a, w.y / l.a, w.z});      // M
        } break;
        case SL_TRANS: {
            float3 wh = ggx_sample_wh(l.mf, wo, sampler.next_2d());
            float d = dot(wo, wh);
            float3 F = sub_fresnel(l, std::fabs(d));
            float3 T = splat(1.f) - F;
            float refl_prob = average3(F) / average3(T * l.kr + F);
            float u1 = sampler.next_1d();
            if (u1 < refl_prob) { wi = reflect(wo, wh); valid = same_hemisphere(wo, wi); }
            else { bool ok = refract(wo, wh, l.eta, wi); valid = ok && !same_hemisphere(wo, wi); }
        } break;
        case SL_DIFFUSE: wi = cosine_hemisphere(sampler.next_2d(), wo.z); break;
        default: {
            float3 wh = ggx_sample_wh(l.mf, wo, sampler.next_2d());
            wi = reflect(wo, wh);
            valid = same_hemisphere(wo, wi);
        } break;
    }
    return frame.to_world(wi);
}

// Lobe::sample with LobeSet::sample_wi_impl, lobe.cpp:111-119,616-649
inline BSDFSample principled_sample(const Scene &sc, const PrincipledSet &ps, float3 world_wo, Sampler &sampler) {
    float uc = sampler.next_1d();
    (void)sampler.next_2d();                                              // drawn and unused by the reference
    int strategy = 0;
    float sum_weights = 0.f;
    for (int i = 0; i < ps.n; ++i) { strategy = uc > sum_weights ? i : strategy; sum_weights += ps.s[i].sample_weight; }
    bool valid = true;
    BSDFSample bs;
    bs.wi = sub_sample_wi(ps.s[strategy], ps.frame, world_wo, sampler, valid);
    bs.eval = principled_evaluate(sc, ps, world_wo, bs.wi, &bs.eta);
    if (!valid) bs.eval.pdf *= 0.f;
    return bs;
}



// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
// Included by vo_pt_li.inl. BSDF lobes of the reference restated as scalar fp32 code; the operation
// order written here is normative for the CUDA mirror (vision_b200/csrc/vb_bsdf.cuh).
//   GGX microfacet            reference src/base/scattering/microfacet.{h,cpp}
//   Fresnel                   reference src/base/scattering/fresnel.h, src/math/optics.h, src/math/complex.h
//   Lambert / MicrofacetRefl  reference src/base/scattering/bxdf.{h,cpp}
//   FresnelBlend (substrate)  reference src/render_core/material/substrate.cpp:12-70
//   Lobe / DielectricLobe / PureReflectionLobe   reference src/base/scattering/lobe.cpp:45-63,111-145,204-250,253-449,716-735
//   material plugins          reference src/render_core/material/{diffuse,mirror,glass,metal,substrate}.cpp
// ocarina helpers that are absent from the tree (face_forward, same_hemisphere, Pow<5>, texture
// sampling of the LUTs: normalised coordinates, linear filter, clamp addressing) are specified here.
namespace vo {
namespace {

struct ScatterEval { float3 f{0, 0, 0}; float pdf = 0.f; uint32_t flags = 1; };
struct BSDFSample { ScatterEval eval; float3 wi{0, 0, 0}; float eta = 1.f; };
constexpr uint32_t FlagReflection = 1u << 1, FlagTransmission = 1u << 2, FlagDiffuse = 1u << 3, FlagGlossy = 1u << 4;

inline float abs_dot(float3 a, float3 b) { return std::fabs(dot(a, b)); }
inline float3 face_forward(float3 v, float3 n) { return dot(v, n) < 0.f ? -v : v; }
inline bool same_hemisphere(float3 a, float3 b) { return a.z * b.z > 0.f; }
inline float pow5(float x) { float x2 = x * x; return (x2 * x2) * x; }
inline float average3(float3 v) { return ((v.x + v.y) + v.z) / 3.0f; }
inline float3 splat(float v) { return {v, v, v}; }
inline float3 reflect(float3 wo, float3 n) { return -wo + n * (2.0f * dot(wo, n)); }
// reference optics.h:27-39 (wi = incident side vector, eta = eta_i / eta_o as used by DielectricLobe)
inline bool refract(float3 wi, float3 n, float eta, float3 &wt) {
    float cos_theta_i = dot(n, wi);
    float sin_theta_i_2 = std::max(0.f, 1.0f - sqr(cos_theta_i));
    float sin_theta_t_2 = sin_theta_i_2 / sqr(eta);
    bool valid = sin_theta_t_2 < 1.0f;
    float cos_theta_t = safe_sqrt(1.0f - sin_theta_t_2);
    wt = -wi / eta + n * (cos_theta_i / eta - cos_theta_t);
    return valid;
}

// ---- Fresnel: reference optics.h:70-79 (dielectric), :92-104 + complex.h (conductor) ----
inline float fresnel_dielectric(float abs_cos_theta_i, float eta) {
    float sin_theta_i_2 = 1.0f - sqr(abs_cos_theta_i);
    float sin_theta_t_2 = sin_theta_i_2 / sqr(eta);
    float cos_theta_t = safe_sqrt(1.0f - sin_theta_t_2);
    float r_parl = (eta * abs_cos_theta_i - cos_theta_t) / (eta * abs_cos_theta_i + cos_theta_t);
    float r_perp = (abs_cos_theta_i - eta * cos_theta_t) / (abs_cos_theta_i + eta * cos_theta_t);
    return sin_theta_t_2 >= 1.0f ? 1.0f : (sqr(r_parl) + sqr(r_perp)) * 0.5f;
}
struct Cx { float re, im; };
inline Cx cmul(Cx a, Cx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
inline Cx cdiv(Cx a, Cx z) {
    float scale = 1.0f / (z.re * z.re + z.im * z.im);
    return {scale * (a.re * z.re + a.im * z.im), scale * (a.im * z.re - a.re * z.im)};
}
inline Cx cadd(Cx a, Cx b) { return {a.re + b.re, a.im + b.im}; }
inline Cx csub(Cx a, Cx b) { return {a.re - b.re, a.im - b.im}; }
inline float cnorm_sqr(Cx z) { return z.re * z.re + z.im * z.im; }
inline Cx csqrt(Cx z) {
    float n = std::sqrt(cnorm_sqr(z));
    float t1 = std::sqrt(0.5f * (n + std::fabs(z.re)));
    float t2 = 0.5f * z.im / t1;
    Cx r;
    r.re = n == 0.f ? 0.f : (z.re >= 0.f ? t1 : std::fabs(t2));
    r.im = n == 0.f ? 0.f : (z.re >= 0.f ? t2 : std::copysign(t1, z.im));
    return r;
}
inline float fresnel_complex(float cos_theta_i, float eta_re, float k) {
    Cx eta{eta_re, k};
    float sin_theta_i_2 = 1.0f - sqr(cos_theta_i);
    Cx sin_theta_t_2 = cdiv(Cx{sin_theta_i_2, 0.f}, cmul(eta, eta));
    Cx cos_theta_t = csqrt(csub(Cx{1.f, 0.f}, sin_theta_t_2));
    Cx ci{cos_theta_i, 0.f};
    Cx r_parl = cdiv(csub(cmul(eta, ci), cos_theta_t), cadd(cmul(eta, ci), cos_theta_t));
    Cx r_perp = cdiv(csub(ci, cmul(eta, cos_theta_t)), cadd(ci, cmul(eta, cos_theta_t)));
    return (cnorm_sqr(r_parl) + cnorm_sqr(r_perp)) * 0.5f;
}

// ---- GGX, reference microfacet.cpp:11-24 (D), :41-50 (lambda), :71-92 (VNDF), :94-110 (sample_wh), :146-155 (PDF_wh) ----
struct GGX { float ax, ay; };
inline float ggx_D(const GGX &m, float3 wh) {
    float3 H{wh.x / m.ax, wh.y / m.ay, wh.z / 1.0f};
    float alpha2 = m.ax * m.ay;
    return InvPi / (alpha2 * sqr(length_squared(H)));
}
inline float ggx_lambda(const GGX &m, float3 w) {
    float sqr_alpha_tan_n = (sqr(m.ax * w.x) + sqr(m.ay * w.y)) / sqr(w.z);
    float ret = 0.5f * (std::sqrt(1.0f + sqr_alpha_tan_n) - 1.0f);
    return w.z == 0.f ? 0.f : ret;
}
inline float ggx_G1(const GGX &m, float3 w) { return 1.0f / (1.0f + ggx_lambda(m, w)); }
inline float ggx_G(const GGX &m, float3 wo, float3 wi) { return 1.0f / (1.0f + ggx_lambda(m, wo) + ggx_lambda(m, wi)); }
inline float3 ggx_sample_vndf(float3 Ve, float2 u, float ax, float ay) {
    float3 Vh = normalize(float3{ax * Ve.x, ay * Ve.y, Ve.z});
    float lenSq = Vh.x * Vh.x + Vh.y * Vh.y;
    float3 T1 = lenSq > 1e-7f ? float3{-Vh.y, Vh.x, 0.0f} / std::sqrt(lenSq) : float3{1, 0, 0};
    float3 T2 = lenSq > 1e-7f ? cross(Vh, T1) : float3{0.0f, 1.0f, 0.0f};
    float r = std::sqrt(u.x), s, c;                      // square_to_disk, warp.h:23-40
    sincos_2pi(u.y, s, c);
    float tx = r * c, ty = r * s;
    ty = lerp(0.5f * (1.0f + Vh.z), safe_sqrt(1.0f - sqr(tx)), ty);
    float3 Nh = T1 * tx + T2 * ty + Vh * safe_sqrt(1.0f - (tx * tx + ty * ty));
    return normalize(float3{ax * Nh.x, ay * Nh.y, std::max(0.0f, Nh.z)});
}
inline float3 ggx_sample_wh(const GGX &m, float3 wo, float2 u) {      // sample_visible = true
    bool flip = wo.z < 0.f;
    float3 wh = ggx_sample_vndf(flip ? -wo : wo, u, m.ax, m.ay);
    return flip ? -wh : wh;
}
inline float ggx_pdf_wh(const GGX &m, float3 wo, float3 wh) {
    return ggx_D(m, wh) * ggx_G1(m, wo) * abs_dot(wo, wh) / std::fabs(wo.z);
}
inline float ggx_pdf_wi_reflection(const GGX &m, float3 wo, float3 wh) {      // microfacet.h:139-156
    return ggx_pdf_wh(m, wo, wh) / (4.0f * abs_dot(wo, wh));
}
inline float ggx_pdf_wi_transmission(const GGX &m, float3 wo, float3 wh, float3 wi, float eta) {   // microfacet.h:169-177
    float denom = sqr(dot(wi, wh) * eta + dot(wo, wh));
    float dwh_dwi = abs_dot(wi, wh) / denom;
    return ggx_pdf_wh(m, wo, wh) * dwh_dwi;
}
inline float ggx_brdf_div_fr(const GGX &m, float3 wo, float3 wh, float3 wi) {  // microfacet.h:179-188
    return ggx_D(m, wh) * ggx_G(m, wo, wi) / std::fabs(4.0f * wo.z * wi.z);
}
inline float ggx_btdf_div_ft(const GGX &m, float3 wo, float3 wh, float3 wi, float eta) {   // microfacet.cpp:160-174, TransportMode::Radiance
    float cos_theta_i = wi.z, cos_theta_o = wo.z;
    float numerator = ggx_D(m, wh) * ggx_G(m, wo, wi) * std::fabs(dot(wi, wh) * dot(wo, wh));
    float denom = sqr(dot(wi, wh) * eta + dot(wo, wh)) * std::fabs(cos_theta_i * cos_theta_o);
    float ft = numerator / denom;
    float factor = 1.0f / sqr(eta);
    ft = denom == 0.f ? 0.f : ft;
    return ft * factor;
}
// microfacet.h:43-60
inline void calculate_alpha(float alpha, float anisotropic, float &ax, float &ay) {
    float x = anisotropic < 0.f ? alpha / (1.f + anisotropic) : alpha * (1.f - anisotropic);
    float y = anisotropic < 0.f ? alpha * (1.f + anisotropic) : alpha / (1.f - anisotropic);
    bool iso = std::fabs(anisotropic) <= 1e-4f;
    ax = iso ? alpha : x; ay = iso ? alpha : y;
}

// ---- LUT textures (MaterialLut, reference material.cpp:14-27): normalised coords, linear, clamp ----
struct LutAxis { int i0, i1; float f; };
inline LutAxis lut_axis(float c, uint32_t n) {
    float x = c * (float)n - 0.5f;
    float fl = std::floor(x);
    LutAxis a;
    a.f = x - fl;
    int i = (int)fl;
    a.i0 = std::min(std::max(i, 0), (int)n - 1);
    a.i1 = std::min(std::max(i + 1, 0), (int)n - 1);
    return a;
}
inline float lut_sample2d(const Lut &L, float u, float v) {                 // 1 channel
    LutAxis a = lut_axis(u, L.nx), b = lut_axis(v, L.ny);
    auto at = [&](int x, int y) { return L.data[(size_t)y * L.nx + x]; };
    float r0 = at(a.i0, b.i0) + a.f * (at(a.i1, b.i0) - at(a.i0, b.i0));
    float r1 = at(a.i0, b.i1) + a.f * (at(a.i1, b.i1) - at(a.i0, b.i1));
    return r0 + b.f * (r1 - r0);
}
inline float lut_sample3d_ch0(const Lut &L, float u, float v, float w) {    // 2 channels, channel 0
    LutAxis a = lut_axis(u, L.nx), b = lut_axis(v, L.ny), c = lut_axis(w, L.nz);
    auto at = [&](int x, int y, int z) { return L.data[(((size_t)z * L.ny + y) * L.nx + x) * L.channels]; };
    auto plane = [&](int z) {
        float r0 = at(a.i0, b.i0, z) + a.f * (at(a.i1, b.i0, z) - at(a.i0, b.i0, z));
        float r1 = at(a.i0, b.i1, z) + a.f * (at(a.i1, b.i1, z) - at(a.i0, b.i1, z));
        return r0 + b.f * (r1 - r0);
    };
    float p0 = plane(c.i0), p1 = plane(c.i1);
    return p0 + c.f * (p1 - p0);
}

// ---- the lobe of one surface hit (the reference builds exactly one Lobe per material here) ----
struct Lobe {
    uint32_t type = VB_MAT_DIFFUSE;
    Frame frame;               // shading frame (Material::compute_shading_frame without a normal map)
    float3 kr{0, 0, 0};        // diffuse Kr | mirror kr | glass kt | substrate Rd
    float3 rs{0, 0, 0};        // substrate Rs | metal eta | plastic / metallic Fresnel F0
    float3 kk{0, 0, 0};        // metal k | metallic F82-tint B
    GGX mf{1.f, 1.f};
    float eta = 1.f;           // glass: corrected relative ior (Interaction::correct_eta); dispersive: of the hero wavelength (eta[0])
    float3 eta3{1, 1, 1};      // dispersive glass (hero spectrum, ior slot = "spd", glass.cpp:221-234): corrected ior per wavelength
    bool dispersive = false;
    bool oren_nayar = false; float on_A = 1.f, on_B = 0.f;
};

// FresnelSchlick::evaluate (fresnel.h:44-52) and FresnelF82Tint::evaluate (fresnel.h:118-124) over plain colour triples
inline float3 fresnel_schlick3(float3 F0, float eta, float cos_theta) {
    float F_real = fresnel_dielectric(cos_theta, eta);
    float F0_real = sqr((eta - 1.0f) / (eta + 1.0f));
    float t = (F_real - F0_real) / (1.f - F0_real);
    t = std::min(std::max(t, 0.f), 1.f);
    return {lerp(t, F0.x, 1.f), lerp(t, F0.y, 1.f), lerp(t, F0.z, 1.f)};
}
inline float3 fresnel_f82(float3 F0, float3 B, float cos_theta) {
    float mu = std::min(std::max(1.f - cos_theta, 0.f), 1.f);
    float mu5 = pow5(mu);
    float3 f_schlick{lerp(mu5, F0.x, 1.f), lerp(mu5, F0.y, 1.f), lerp(mu5, F0.z, 1.f)};
    float3 r = f_schlick - B * cos_theta * mu5 * mu;
    return {std::min(std::max(r.x, 0.f), 1.f), std::min(std::max(r.y, 0.f), 1.f), std::min(std::max(r.z, 0.f), 1.f)};
}

inline float3 lobe_fresnel(const Lobe &lb, float cos_theta) {
    switch (lb.type) {
        case VB_MAT_PLASTIC: return fresnel_schlick3(lb.rs, lb.eta, cos_theta);
        case VB_MAT_METALLIC: return fresnel_f82(lb.rs, lb.kk, cos_theta);
        case VB_MAT_METAL: return {fresnel_complex(cos_theta, lb.rs.x, lb.kk.x), fresnel_complex(cos_theta, lb.rs.y, lb.kk.y),
                                   fresnel_complex(cos_theta, lb.rs.z, lb.kk.z)};
        case VB_MAT_GLASS: return splat(fresnel_dielectric(cos_theta, lb.eta));
        case VB_MAT_SUBSTRATE: return splat(fresnel_dielectric(cos_theta, 1.5f));
        default: return splat(1.f);          // FresnelConstant (mirror)
    }
}

// -- DielectricLobe, reference lobe.cpp:253-449 --
inline float dielectric_compensate(const Scene &sc, const Lobe &lb, float3 wo) {
    const Lut &L = lb.eta > 1.0f ? sc.luts[VB_LUT_DIELECTRIC] : sc.luts[VB_LUT_DIELECTRIC_INV];
    if (L.data.empty()) return 1.0f;
    float x = std::sqrt(std::sqrt(lb.mf.ax * lb.mf.ay));
    float y = std::fabs(wo.z);
    const float lo = 1.003f, hi = 5.0f;
    float z = lb.eta > 1.0f ? (lb.eta - lo) / (hi - lo) : (1.0f / lb.eta - lo) / (hi - lo);
    return 1.0f / lut_sample3d_ch0(L, x, y, z);
}
// FresnelDielectric::evaluate (fresnel.h:84-89): one value per entry of the ior array - a scalar slot gives a
// 1-dimensional spectrum (F.average() == F), the dispersive "spd" slot one value per wavelength
inline float3 dielectric_F(const Lobe &lb, float cos_theta) {
    if (!lb.dispersive) return splat(fresnel_dielectric(cos_theta, lb.eta));
    return {fresnel_dielectric(cos_theta, lb.eta3.x), fresnel_dielectric(cos_theta, lb.eta3.y), fresnel_dielectric(cos_theta, lb.eta3.z)};
}
inline float dielectric_refl_prob(const Lobe &lb, float3 F) {
    float3 T = splat(1.0f) - F;
    float3 total = lb.kr * T + F;
    return (lb.dispersive ? average3(F) : F.x) / average3(total);
}
inline ScatterEval dielectric_evaluate_local(const Scene &sc, const Lobe &lb, float3 wo, float3 wi, bool flip_wh, float *eta_out) {
    bool reflect_ = same_hemisphere(wo, wi);
    float eta_p = reflect_ ? 1.0f : lb.eta;
    if (eta_out) *eta_out = eta_p;
    float3 wh = normalize(wo + wi * eta_p);
    if (flip_wh) wh = face_forward(wh, wo);
    float3 F = dielectric_F(lb, abs_dot(wh, wo));
    ScatterEval se;
    if (reflect_) {
        se.f = F * ggx_brdf_div_fr(lb.mf, wo, wh, wi);
        se.pdf = ggx_pdf_wi_reflection(lb.mf, wo, wh) * dielectric_refl_prob(lb, F);
        se.flags = FlagGlossy | FlagReflection;
        se.f = se.f * dielectric_compensate(sc, lb, wo);
    } else {
        float3 new_wh = face_forward(wh, wo);
        float3 wh2 = normalize(wo + wi * lb.eta);                         // Microfacet::BTDF(wo, wi, ...) recomputes wh
        float3 tr = (splat(1.0f) - F) * ggx_btdf_div_ft(lb.mf, wo, wh2, wi, lb.eta);
        se.f = lb.kr * tr;
        se.pdf = ggx_pdf_wi_transmission(lb.mf, wo, new_wh, wi, lb.eta) * (1.0f - dielectric_refl_prob(lb, F));
        se.flags = FlagGlossy | FlagTransmission;
        se.f = se.f * dielectric_compensate(sc, lb, wo);
    }
    return se;
}

// -- MicrofacetReflection (mirror, metal), reference bxdf.cpp:65-85 + PureReflectionLobe lobe.cpp:716-730 --
inline ScatterEval microfacet_refl_evaluate_local(const Scene &sc, const Lobe &lb, float3 wo, float3 wi) {
    ScatterEval se;
    bool s = same_hemisphere(wo, wi);
    float3 wh = normalize(wo + wi);
    float3 whf = face_forward(wh, float3{0, 0, 1});
    float3 F = lobe_fresnel(lb, abs_dot(wo, whf));
    float3 fr = F * ggx_brdf_div_fr(lb.mf, wo, whf, wi);
    float3 f = fr * lb.kr;
    float pdf = ggx_pdf_wi_reflection(lb.mf, wo, wh);
    se.f = s ? f : float3{0, 0, 0};
    se.pdf = s ? pdf : 0.f;
    se.flags = FlagGlossy | FlagReflection;
    const Lut &L = sc.luts[VB_LUT_PURE_REFLECTION];
    if (!L.data.empty()) {
        float alpha = std::sqrt(lb.mf.ax * lb.mf.ay);
        se.f = se.f * (1.0f / lut_sample2d(L, alpha, wo.z));
    }
    return se;
}

// -- FresnelBlend (substrate), reference substrate.cpp:12-70 --
inline ScatterEval fresnel_blend_evaluate_local(const Lobe &lb, float3 wo, float3 wi) {
    ScatterEval se;
    bool s = same_hemisphere(wo, wi);
    float aci = std::fabs(wi.z), aco = std::fabs(wo.z);
    float3 diffuse = lb.kr * (28.f / (23.f * Pi)) * (splat(1.f) - lb.rs) * (1.0f - pow5(1.0f - .5f * aci)) * (1.0f - pow5(1.0f - .5f * aco));
    float3 wh = normalize(wi + wo);
    float sw = pow5(std::min(std::max(1.0f - dot(wi, wh), 0.f), 1.f));       // schlick_weight
    float3 schlick{lerp(sw, lb.rs.x, 1.f), lerp(sw, lb.rs.y, 1.f), lerp(sw, lb.rs.z, 1.f)};
    float3 specular = schlick * (ggx_D(lb.mf, wh) / (4.0f * abs_dot(wi, wh) * std::max(aci, aco)));
    if (is_zero(wh)) specular = specular * 0.f;
    float3 f = specular + diffuse;
    float fr = fresnel_dielectric(aco, 1.5f);
    float pdf = lerp(fr, aci * InvPi, ggx_pdf_wi_reflection(lb.mf, wo, normalize(wo + wi)));
    se.f = s ? f : float3{0, 0, 0};
    se.pdf = s ? pdf : 0.f;
    se.flags = FlagReflection;
    return se;
}

// -- Lambert / Oren-Nayar, reference bxdf.cpp:13-46,94-121 --
inline ScatterEval diffuse_evaluate_local(const Lobe &lb, float3 wo, float3 wi) {
    ScatterEval se;
    bool s = same_hemisphere(wo, wi);
    float3 f = lb.kr * InvPi;
    if (lb.oren_nayar) {
        auto sin_theta = [](float3 w) { return std::sqrt(std::max(0.f, 1.0f - w.z * w.z)); };
        auto cos_phi = [&](float3 w) { float st = sin_theta(w); return st == 0.f ? 1.f : std::min(std::max(w.x / st, -1.f), 1.f); };
        auto sin_phi = [&](float3 w) { float st = sin_theta(w); return st == 0.f ? 0.f : std::min(std::max(w.y / st, -1.f), 1.f); };
        float sti = sin_theta(wi), sto = sin_theta(wo);
        float d_cos = cos_phi(wi) * cos_phi(wo) + sin_phi(wi) * sin_phi(wo);
        float max_cos = std::max(0.f, d_cos);
        bool cond = std::fabs(wi.z) > std::fabs(wo.z);
        float sin_alpha = cond ? sto : sti;
        float tan_beta = cond ? sti / std::fabs(wi.z) : sto / std::fabs(wo.z);
        f = lb.kr * InvPi * (lb.on_A + lb.on_B * max_cos * sin_alpha * tan_beta);
    }
    se.f = s ? f : float3{0, 0, 0};
    se.pdf = s ? std::fabs(wi.z) * InvPi : 0.f;
    se.flags = FlagDiffuse | FlagReflection;
    return se;
}

// -- PlasticLobe::evaluate_local_impl, plastic.cpp:37-50 (no hemisphere test of its own; flags stay at the ScatterEval default) --
inline ScatterEval plastic_evaluate_local(const Lobe &lb, float3 wo, float3 wi) {
    ScatterEval se;
    float3 wh = normalize(wo + wi);
    float3 F = lobe_fresnel(lb, abs_dot(wh, wo));
    se.f = lb.kr * InvPi * (splat(1.f) - F);
    se.f = se.f + F * ggx_brdf_div_fr(lb.mf, wo, wh, wi);
    se.pdf = lerp(average3(F), std::fabs(wi.z) * InvPi, ggx_pdf_wi_reflection(lb.mf, wo, wh));
    return se;
}

// Lobe::evaluate_impl (lobe.cpp:53-77): world -> local, f *= |cos(wi)|
inline ScatterEval lobe_evaluate(const Scene &sc, const Lobe &lb, float3 world_wo, float3 world_wi, bool with_eta, float *eta_out) {
    float3 wo = lb.frame.to_local(world_wo), wi = lb.frame.to_local(world_wi);
    ScatterEval se;
    switch (lb.type) {
        case VB_MAT_GLASS: se = dielectric_evaluate_local(sc, lb, wo, wi, with_eta, eta_out); break;
        case VB_MAT_MIRROR: case VB_MAT_METAL: case VB_MAT_METALLIC: se = microfacet_refl_evaluate_local(sc, lb, wo, wi); break;
        case VB_MAT_SUBSTRATE: se = fresnel_blend_evaluate_local(lb, wo, wi); break;
        case VB_MAT_PLASTIC: se = plastic_evaluate_local(lb, wo, wi); break;
        default: se = diffuse_evaluate_local(lb, wo, wi); break;
    }
    se.f = se.f * std::fabs(wi.z);
    return se;
}

inline float3 cosine_hemisphere(float2 u, float woz) {               // bxdf.cpp:48-52 + warp.h:23-40,47-52
    float r = std::sqrt(u.x), s, c;
    sincos_2pi(u.y, s, c);
    float dx = r * c, dy = r * s;
    float z = std::sqrt(std::max(0.0f, 1.0f - dx * dx - dy * dy));
    if (woz < 0.f) z = -z;
    return {dx, dy, z};
}

#include "vo_principled.inl"

// Lobe::sample (lobe.cpp:111-119) with each lobe's sample_wi_local_impl; draws in the reference's order
// the direction only (sample_wi_local_impl of the lobe, in world space); `valid` = SampledDirection::valid
inline float3 lobe_sample_wi(const Lobe &lb, float3 world_wo, Sampler &sampler, bool &valid) {
    float3 wo = lb.frame.to_local(world_wo);
    float3 wi;
    valid = true;
    switch (lb.type) {
        case VB_MAT_GLASS: {                                           // lobe.cpp:431-449
            float3 wh = ggx_sample_wh(lb.mf, wo, sampler.next_2d());
            float d = dot(wo, wh);
            float3 F = dielectric_F(lb, std::fabs(d));
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
            float3 wh = ggx_sample_wh(lb.mf, wo, sampler.next_2d());
            float3 F = lobe_fresnel(lb, std::fabs(wo.z));
            float uc = sampler.next_1d();
            if (uc < average3(F)) {
                wi = reflect(wo, wh);
                valid = same_hemisphere(wo, wi);
            } else wi = cosine_hemisphere(sampler.next_2d(), 1.f);     // square_to_cosine_hemisphere: always the +z side
        } break;
        case VB_MAT_MIRROR: case VB_MAT_METAL: case VB_MAT_METALLIC: {  // bxdf.cpp:79-83
            float3 wh = ggx_sample_wh(lb.mf, wo, sampler.next_2d());
            wi = reflect(wo, wh);
            valid = same_hemisphere(wo, wi);
        } break;
        case VB_MAT_SUBSTRATE: {                                       // substrate.cpp:52-68
            float2 u = sampler.next_2d();
            float fr = fresnel_dielectric(std::fabs(wo.z), 1.5f);
            if (u.x < fr) {
                u.x = (u.x - 0.f) / (fr - 0.f);
                float3 wh = ggx_sample_wh(lb.mf, wo, u);
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
inline BSDFSample lobe_sample(const Scene &sc, const Lobe &lb, float3 world_wo, Sampler &sampler) {
    bool valid = true;
    BSDFSample bs;
    bs.wi = lobe_sample_wi(lb, world_wo, sampler, valid);
    bs.eval = lobe_evaluate(sc, lb, world_wo, bs.wi, true, &bs.eta);
    if (!valid) bs.eval.pdf *= 0.f;
    return bs;
}

// Material::create_evaluator -> create_lobe_set of the plugin (diffuse.cpp:21-30, mirror.cpp:60-74,
// glass.cpp:240-257, metal.cpp:137-156, substrate.cpp:126-149)
inline Lobe create_lobe_of(const Scene &sc, const Swl &swl, const Interaction &it, uint32_t mat_id) {
    const vb_material_rec &m = sc.materials[mat_id];
    Lobe lb;
    lb.type = m.type;
    lb.frame = it.shading;
    bool remap = (m.flags & 2u) != 0;
    auto clampf = [](float v, float lo, float hi) { return std::min(std::max(v, lo), hi); };
    float anis = clampf(m.params[0], -0.9f, 0.9f);
    switch (m.type) {
        case VB_MAT_DIFFUSE:
            lb.kr = spec_albedo(sc, swl, eval_slot3(sc, m.color, it.uv));
            if (m.flags & 1u) {                                        // OrenNayar ctor, bxdf.cpp:94-101
                float sigma = eval_slot3(sc, m.sigma, it.uv).x * PiOver2;
                float sigma2 = sqr(sigma * sigma);
                lb.oren_nayar = true;
                lb.on_A = 1.f - (sigma2 / (2.f * (sigma2 + 0.33f)));
                lb.on_B = 0.45f * sigma2 / (sigma2 + 0.09f);
            }
            break;
        case VB_MAT_MIRROR: case VB_MAT_METAL: {
            lb.kr = m.type == VB_MAT_MIRROR ? spec_albedo(sc, swl, eval_slot3(sc, m.color, it.uv)) : splat(1.f);
            float r = clampf(eval_slot3(sc, m.roughness, it.uv).x, 0.0001f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            lb.rs = {m.eta[0], m.eta[1], m.eta[2]}; lb.kk = {m.k[0], m.k[1], m.k[2]};
            if (m.type == VB_MAT_METAL && (m.flags & 4u) && swl.hero) {   // metal.cpp:117-121: eta / k are "spd" slots under a complete spectrum
                lb.rs = spd_eval3(sc, m.spd[0], swl); lb.kk = spd_eval3(sc, m.spd[1], swl);
            }
        } break;
        case VB_MAT_GLASS: {
            lb.kr = spec_albedo(sc, swl, eval_slot3(sc, m.color, it.uv));
            float ior = eval_slot3(sc, m.ior, it.uv).x;
            lb.eta = dot(it.wo, it.ng) > 0.f ? ior : 1.0f / ior;      // Interaction::correct_eta
            if ((m.flags & 4u) && swl.hero) {                          // glass.cpp:216-234: "material_name" curve sampled into an "spd" slot
                float3 iors = spd_eval3(sc, m.spd[0], swl);
                bool front = dot(it.wo, it.ng) > 0.f;
                lb.eta3 = front ? iors : float3{1.0f / iors.x, 1.0f / iors.y, 1.0f / iors.z};
                lb.eta = lb.eta3.x;
                lb.dispersive = true;
            }
            float r = clampf(eval_slot3(sc, m.roughness, it.uv).x, 0.01f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
        } break;
        case VB_MAT_SUBSTRATE: {
            lb.kr = spec_albedo(sc, swl, eval_slot3(sc, m.color, it.uv));
            lb.rs = spec_albedo(sc, swl, eval_slot3(sc, m.spec, it.uv));
            float r = clampf(eval_slot3(sc, m.roughness, it.uv).x, 0.0001f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            if (remap) { lb.mf.ax = sqr(lb.mf.ax); lb.mf.ay = sqr(lb.mf.ay); }
            lb.mf.ax = clampf(lb.mf.ax, 0.0001f, 1.f); lb.mf.ay = clampf(lb.mf.ay, 0.0001f, 1.f);
        } break;
        case VB_MAT_PLASTIC: {                                         // plastic.cpp:109-129
            lb.kr = spec_albedo(sc, swl, eval_slot3(sc, m.color, it.uv));
            float3 Rs = spec_albedo(sc, swl, eval_slot3(sc, m.spec, it.uv));
            float ior = eval_slot3(sc, m.ior, it.uv).x;
            float r = clampf(eval_slot3(sc, m.roughness, it.uv).x, 0.0001f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            if (remap) { lb.mf.ax = sqr(lb.mf.ax); lb.mf.ay = sqr(lb.mf.ay); }
            lb.mf.ax = clampf(lb.mf.ax, 0.0001f, 1.f); lb.mf.ay = clampf(lb.mf.ay, 0.0001f, 1.f);
            lb.eta = ior;
            lb.rs = Rs * sqr((ior - 1.0f) / (ior + 1.0f));             // schlick_F0_from_ior(iors[0]) * Rs
        } break;
        case VB_MAT_METALLIC: {                                        // metallic.cpp:44-60
            lb.kr = spec_albedo(sc, swl, eval_slot3(sc, m.color, it.uv));
            float3 edge_tint = spec_albedo(sc, swl, eval_slot3(sc, m.spec, it.uv));
            float r = clampf(eval_slot3(sc, m.roughness, it.uv).x, 0.01f, 1.f);
            r = remap ? sqr(r) : r;
            calculate_alpha(r, anis, lb.mf.ax, lb.mf.ay);
            lb.rs = lb.kr;                                             // FresnelF82Tint(color) + init_from_F82(edge_tint), fresnel.h:110-116
            const float f = 6.f / 7.f, f5 = ((f * f) * (f * f)) * f;
            float3 f_schlick{lerp(f5, lb.rs.x, 1.f), lerp(f5, lb.rs.y, 1.f), lerp(f5, lb.rs.z, 1.f)};
            lb.kk = f_schlick * (7.f / (f5 * f)) * (splat(1.f) - edge_tint);
        } break;
        default: break;                                                // black body etc.: zero lobe
    }
    return lb;
}
inline Lobe create_lobe(const Scene &sc, const Swl &swl, const Interaction &it) { return create_lobe_of(sc, swl, it, it.mat_id); }

// Lobe::albedo(cos_theta) of the single-lobe materials (lobe.h:167 diffuse = Kr; lobe.cpp:208-210 microfacet = kr * F(cos);
// lobe.cpp:308-313 dielectric = kt (1 - F(|cos|)) + F; substrate.cpp:22 = Rd); consumed by rt_geom's albedo plane
inline float3 lobe_albedo(const Lobe &lb, float cos_theta) {
    switch (lb.type) {
        case VB_MAT_DIFFUSE: case VB_MAT_SUBSTRATE: return lb.kr;
        case VB_MAT_MIRROR: case VB_MAT_METAL: case VB_MAT_METALLIC: return lb.kr * lobe_fresnel(lb, cos_theta);
        case VB_MAT_PLASTIC: return lobe_fresnel(lb, cos_theta);      // MicrofacetLobe::albedo with a unit reflection colour (plastic.cpp:127)
        case VB_MAT_GLASS: { float3 F = dielectric_F(lb, std::fabs(cos_theta)); return lb.kr * (splat(1.0f) - F) + F; }
        default: return float3{0, 0, 0};
    }
}
// LobeSet::albedo (lobe.cpp:564-570) over the principled BSDF's sub-lobes: sheen = tint (principled_bsdf.cpp:54-56), coat / specular =
// their albedo tables at cos_theta (:154-161,198-206), metallic = kr F82(cos) (MicrofacetLobe::albedo), transmission = dielectric form
inline float3 sub_albedo(const Scene &sc, const SubLobe &l, float cos_theta) {
    float3 a{0, 0, 0};
    auto to_ratio_x = [](const GGX &g) { return std::sqrt(std::sqrt(g.ax * g.ay)); };
    switch (l.kind) {
        case SL_SHEEN: case SL_DIFFUSE: a = l.kr; break;
        case SL_COAT: {
            const Lut &L = sc.luts[VB_LUT_COAT];
            float s = L.data.empty() ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, (l.eta - 1.003f) / (4.f - 1.003f));
            a = l.kr * s;
        } break;
        case SL_METALLIC: a = l.kr * sub_fresnel(l, cos_theta); break;
        case SL_TRANS: { float3 F = sub_fresnel(l, std::fabs(cos_theta)); a = l.kr * (splat(1.0f) - F) + F; } break;
        case SL_SPEC: {
            const Lut &L = sc.luts[VB_LUT_SPECULAR];
            float z = std::sqrt(std::fabs((l.eta - 1.0f) / (l.eta + 1.0f)));
            float s = L.data.empty() ? 1.f : lut_sample3d_ch0(L, to_ratio_x(l.mf), cos_theta, z);
            a = lerp3v(s, l.F0, splat(1.f)) * l.kr;
        } break;
    }
    return a;
}
inline float3 principled_albedo(const Scene &sc, const PrincipledSet &ps, float cos_theta) {
    float3 ret{0, 0, 0};
    for (int i = 0; i < ps.n; ++i) ret += sub_albedo(sc, ps.s[i], cos_theta) * ps.s[i].weight;
    return ret;
}

// ---- mix / add materials: a LobeSet over the lobes of two child materials --------------------------------------------------------
// LobeSet::create_mix (lobe.cpp:495-508): children weighted (1 - frac, frac) in both weight and sample weight; create_add (:510-521):
// weights 1, sample weights normalised to 1/2. flatten (:533-562): the sub-lobes of a child that is itself a set (the principled
// BSDF) are pulled up with sample_weight *= parent sample weight and weight *= parent SAMPLE weight (sic). Evaluation and sampling
// are LobeSet::evaluate_impl / sample_wi_impl over the flattened list (:598-683): every lobe in its own shading frame, f scaled by
// weight * valid_world_factor, pdf by sample_weight * valid_world_factor. LobeSet has no is_dispersive (lobe.h:104): a dispersive
// glass child never invalidates the secondary wavelengths.
struct MixEntry { bool sub = false; float sw = 0.f, w = 1.f; Lobe lobe; SubLobe sl; Frame frame; };
struct MixSet { int n = 0; MixEntry e[12]; };

inline MixSet mixset_create(const Scene &sc, const Swl &swl, const Interaction &it) {
    const vb_material_rec &m = sc.materials[it.mat_id];
    float psw[2], pw[2];
    if (m.type == VB_MAT_MIX) {
        float frac = eval_slot3(sc, m.sigma, it.uv).x;
        psw[0] = pw[0] = 1 - frac; psw[1] = pw[1] = frac;
    } else {
        float sum = 0.f; sum += 1.f; sum += 1.f;                        // normalize_sampled_weight
        psw[0] = 1.f / sum; psw[1] = 1.f / sum; pw[0] = pw[1] = 1.f;
    }
    MixSet ms;
    for (int c = 0; c < 2; ++c) {
        uint32_t child = m.spd[c];
        if (sc.materials[child].type == VB_MAT_PRINCIPLED) {
            PrincipledSet ps = principled_create_of(sc, swl, it, child);
            for (int i = 0; i < ps.n; ++i) {
                MixEntry &e = ms.e[ms.n++];
                e.sub = true; e.sl = ps.s[i]; e.frame = ps.frame;
                e.sl.sample_weight = e.sl.sample_weight * psw[c];
                e.sl.weight = e.sl.weight * psw[c];
                e.sw = e.sl.sample_weight; e.w = e.sl.weight;
            }
        } else {
            MixEntry &e = ms.e[ms.n++];
            e.lobe = create_lobe_of(sc, swl, it, child);
            e.sw = psw[c]; e.w = pw[c];
        }
    }
    return ms;
}

inline ScatterEval mixset_evaluate(const Scene &sc, const MixSet &ms, float3 world_wo, float3 world_wi, float *eta_out) {
    ScatterEval ret;
    ret.flags = 0;
    for (int i = 0; i < ms.n; ++i) {
        const MixEntry &e = ms.e[i];
        ScatterEval se;
        float factor;
        if (e.sub) {
            float3 wo = e.frame.to_local(world_wo), wi = e.frame.to_local(world_wi);
            se = sub_evaluate_local(sc, e.sl, wo, wi, eta_out);
            se.f = se.f * std::fabs(wi.z);
            factor = e.sl.kind == SL_TRANS ? 1.f : (dot(world_wo, e.frame.z) * dot(world_wi, e.frame.z) > 0.f ? 1.f : 0.f);
        } else {
            se = lobe_evaluate(sc, e.lobe, world_wo, world_wi, true, eta_out);
            factor = e.lobe.type == VB_MAT_GLASS ? 1.f : (dot(world_wo, e.lobe.frame.z) * dot(world_wi, e.lobe.frame.z) > 0.f ? 1.f : 0.f);
        }
        se.f = se.f * (e.w * factor);
        se.pdf = se.pdf * (e.sw * factor);
        ret.f = ret.f + se.f;
        ret.pdf = ret.pdf + se.pdf;
        ret.flags |= se.flags;
    }
    return ret;
}

inline BSDFSample mixset_sample(const Scene &sc, const MixSet &ms, float3 world_wo, Sampler &sampler) {
    float uc = sampler.next_1d();
    (void)sampler.next_2d();
    int strategy = 0;
    float sum_weights = 0.f;
    for (int i = 0; i < ms.n; ++i) { strategy = uc > sum_weights ? i : strategy; sum_weights += ms.e[i].sw; }
    const MixEntry &e = ms.e[strategy];
    bool valid = true;
    BSDFSample bs;
    bs.wi = e.sub ? sub_sample_wi(e.sl, e.frame, world_wo, sampler, valid) : lobe_sample_wi(e.lobe, world_wo, sampler, valid);
    bs.eval = mixset_evaluate(sc, ms, world_wo, bs.wi, &bs.eta);
    if (!valid) bs.eval.pdf *= 0.f;
    return bs;
}

inline float3 mixset_albedo(const Scene &sc, const MixSet &ms, float cos_theta) {
    float3 ret{0, 0, 0};
    for (int i = 0; i < ms.n; ++i) {
        const MixEntry &e = ms.e[i];
        ret += (e.sub ? sub_albedo(sc, e.sl, cos_theta) : lobe_albedo(e.lobe, cos_theta)) * e.w;
    }
    return ret;
}

// MaterialEvaluator (individual_ns branch), reference material.cpp:132-142,163-175
struct PrincipledEvaluator {
    PrincipledSet set;
    float3 ng;
    ScatterEval evaluate(const Scene &sc, float3 wo, float3 wi) const {
        ScatterEval se = principled_evaluate(sc, set, wo, wi, nullptr);
        bool same = dot(wo, ng) * dot(wi, ng) > 0.f;
        bool is_trans = (se.flags & FlagTransmission) != 0;
        if (same == is_trans) se.pdf = 0.f;
        return se;
    }
    BSDFSample sample(const Scene &sc, float3 wo, Sampler &sampler) const {
        BSDFSample bs = principled_sample(sc, set, wo, sampler);
        bool same = dot(wo, ng) * dot(bs.wi, ng) > 0.f;
        bool is_trans = (bs.eval.flags & FlagTransmission) != 0;
        if (same == is_trans) bs.eval.pdf = 0.f;
        return bs;
    }
};
struct MixEvaluator {
    MixSet set;
    float3 ng;
    ScatterEval evaluate(const Scene &sc, float3 wo, float3 wi) const {
        ScatterEval se = mixset_evaluate(sc, set, wo, wi, nullptr);
        bool same = dot(wo, ng) * dot(wi, ng) > 0.f;
        bool is_trans = (se.flags & FlagTransmission) != 0;
        if (same == is_trans) se.pdf = 0.f;
        return se;
    }
    BSDFSample sample(const Scene &sc, float3 wo, Sampler &sampler) const {
        BSDFSample bs = mixset_sample(sc, set, wo, sampler);
        bool same = dot(wo, ng) * dot(bs.wi, ng) > 0.f;
        bool is_trans = (bs.eval.flags & FlagTransmission) != 0;
        if (same == is_trans) bs.eval.pdf = 0.f;
        return bs;
    }
};
struct Evaluator {
    Lobe lobe;
    float3 ng;
    ScatterEval evaluate(const Scene &sc, float3 wo, float3 wi) const {
        ScatterEval se = lobe_evaluate(sc, lobe, wo, wi, false, nullptr);
        bool same = dot(wo, ng) * dot(wi, ng) > 0.f;
        bool is_trans = (se.flags & FlagTransmission) != 0;
        if (same == is_trans) se.pdf = 0.f;
        return se;
    }
    BSDFSample sample(const Scene &sc, float3 wo, Sampler &sampler) const {
        BSDFSample bs = lobe_sample(sc, lobe, wo, sampler);
        bool same = dot(wo, ng) * dot(bs.wi, ng) > 0.f;
        bool is_trans = (bs.eval.flags & FlagTransmission) != 0;
        if (same == is_trans) bs.eval.pdf = 0.f;
        return bs;
    }
};

}  // namespace
}  // namespace vo

// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
// extern "C" probes of the BSDF building blocks for the CPU unit tests (tests/test_oracle_bsdf.py).
// Included at the end of vo_pt.cpp so that it sees the file-local lobe code.
extern "C" {

float vo_probe_fresnel_dielectric(float c, float eta) { return vo::fresnel_dielectric(c, eta); }
float vo_probe_fresnel_complex(float c, float eta, float k) { return vo::fresnel_complex(c, eta, k); }
int vo_probe_refract(const float *wi, const float *n, float eta, float *wt) {
    vo::float3 t{0, 0, 0};
    bool ok = vo::refract({wi[0], wi[1], wi[2]}, {n[0], n[1], n[2]}, eta, t);
    wt[0] = t.x; wt[1] = t.y; wt[2] = t.z;
    return ok ? 1 : 0;
}
void vo_probe_reflect(const float *wo, const float *n, float *out) {
    vo::float3 r = vo::reflect({wo[0], wo[1], wo[2]}, {n[0], n[1], n[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
// out: D(wh), G(wo,wi), pdf_wh(wo,wh), lambda(wo)
void vo_probe_ggx(float ax, float ay, const float *wo, const float *wi, const float *wh, float *out) {
    vo::GGX m{ax, ay};
    vo::float3 o{wo[0], wo[1], wo[2]}, i{wi[0], wi[1], wi[2]}, h{wh[0], wh[1], wh[2]};
    out[0] = vo::ggx_D(m, h); out[1] = vo::ggx_G(m, o, i); out[2] = vo::ggx_pdf_wh(m, o, h); out[3] = vo::ggx_lambda(m, o);
}
void vo_probe_ggx_sample_wh(float ax, float ay, const float *wo, float u0, float u1, float *wh) {
    vo::float3 h = vo::ggx_sample_wh(vo::GGX{ax, ay}, {wo[0], wo[1], wo[2]}, vo::float2{u0, u1});
    wh[0] = h.x; wh[1] = h.y; wh[2] = h.z;
}

// Evaluates / samples material `mat_id` of a scene at a canonical interaction: position 0, geometric and
// shading normal +z, tangent +x, uv (0.5,0.5); wo/wi in world = local space.
static vo::Interaction probe_interaction(uint32_t mat_id, const float *wo) {
    vo::Interaction it;
    it.pos = {0, 0, 0}; it.ng = {0, 0, 1}; it.uv = {0.5f, 0.5f};
    it.shading = vo::Frame{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    it.wo = {wo[0], wo[1], wo[2]};
    it.mat_id = mat_id;
    return it;
}
// out: f.xyz (already times |cos wi|), pdf, flags
void vo_probe_bsdf_eval(void *scene, uint32_t mat_id, const float *wo, const float *wi, float *out) {
    const vo::Scene &sc = *(const vo::Scene *)scene;
    vo::Interaction it = probe_interaction(mat_id, wo);
    vo::ScatterEval se;
    const uint32_t mt = sc.materials[mat_id].type;
    if (mt == VB_MAT_MIX || mt == VB_MAT_ADD) se = vo::MixEvaluator{vo::mixset_create(sc, vo::Swl{}, it), it.ng}.evaluate(sc, it.wo, {wi[0], wi[1], wi[2]});
    else if (sc.materials[mat_id].type == VB_MAT_PRINCIPLED) se = vo::PrincipledEvaluator{vo::principled_create(sc, vo::Swl{}, it), it.ng}.evaluate(sc, it.wo, {wi[0], wi[1], wi[2]});
    else se = vo::Evaluator{vo::create_lobe(sc, vo::Swl{}, it), it.ng}.evaluate(sc, it.wo, {wi[0], wi[1], wi[2]});
    out[0] = se.f.x; out[1] = se.f.y; out[2] = se.f.z; out[3] = se.pdf; out[4] = (float)se.flags;
}
// n samples with the LCG seeded by `state`; out per sample: wi.xyz, f.xyz, pdf, eta  (8 floats)
void vo_probe_bsdf_sample(void *scene, uint32_t mat_id, const float *wo, uint32_t state, uint32_t n, float *out) {
    const vo::Scene &sc = *(const vo::Scene *)scene;
    vo::Interaction it = probe_interaction(mat_id, wo);
    const bool principled = sc.materials[mat_id].type == VB_MAT_PRINCIPLED;
    vo::Evaluator ev{(principled || sc.materials[mat_id].type >= VB_MAT_MIX) ? vo::Lobe{} : vo::create_lobe(sc, vo::Swl{}, it), it.ng};
    vo::PrincipledEvaluator pev{principled ? vo::principled_create(sc, vo::Swl{}, it) : vo::PrincipledSet{}, it.ng};
    const bool lobe_set = sc.materials[mat_id].type == VB_MAT_MIX || sc.materials[mat_id].type == VB_MAT_ADD;
    vo::MixEvaluator mev{lobe_set ? vo::mixset_create(sc, vo::Swl{}, it) : vo::MixSet{}, it.ng};
    vo::Sampler sampler; sampler.state = state;
    for (uint32_t k = 0; k < n; ++k) {
        vo::BSDFSample bs = lobe_set ? mev.sample(sc, it.wo, sampler) : principled ? pev.sample(sc, it.wo, sampler) : ev.sample(sc, it.wo, sampler);
        float *o = out + 8 * (size_t)k;
        o[0] = bs.wi.x; o[1] = bs.wi.y; o[2] = bs.wi.z; o[3] = bs.eval.f.x; o[4] = bs.eval.f.y; o[5] = bs.eval.f.z;
        o[6] = bs.eval.pdf; o[7] = bs.eta;
    }
}

}  // extern "C"

// LightSampler::select_light / PMF of the scene's rule at u (lightsampler.cpp:159-197): returns the light index, *pmf its PMF
extern "C" uint32_t vo_probe_light_select(void *scene, float u, float *pmf) {
    const vo::Scene &sc = *(const vo::Scene *)scene;
    uint32_t i = vo::light_select(sc, u);
    *pmf = vo::light_pmf(sc, i);
    return i;
}
extern "C" float vo_probe_light_pmf(void *scene, uint32_t index) { return vo::light_pmf(*(const vo::Scene *)scene, index); }

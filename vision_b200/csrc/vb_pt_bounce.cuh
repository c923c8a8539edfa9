// One vertex of IlluminationIntegrator::Li (reference src/base/integral/integrator.cpp:160-299) as a
// device function shared by the two schedules of the path tracer (megakernel: vb_pt_kernel.cuh,
// wavefront: vb_pt_wavefront.cuh). Operation order mirrors oracle/vo_pt_li.inl render_sample.
// Included by vb_pt.cu only.
#pragma once

namespace {

struct PathState {
    Sampler sampler;
    f3 throughput, ret;
    float eta_scale, scatter_pdf;
    uint32_t ray_medium, bounces;
    uint32_t swl;              // hero spectrum: bits 0..23 wavelength draw, bit 31 secondary wavelengths invalidated (vb_spectrum.cuh);
                               // bit 30: the "only direct" supplement step (Li with max_depth < 2, integrator.cpp:303-309)
};

__device__ __forceinline__ void path_begin(const PtScene &S, PathState &ps, uint32_t px, uint32_t py, uint32_t frame) {
    ps.swl = 0;
    if (S.params.spectrum == VB_SPECTRUM_HERO) {                  // RenderEnv::initial, integrator.cpp:48-57: sampler dimension -1
        ps.sampler.start(px, py, frame, 0xffffffffu);
        (void)ps.sampler.next_1d();
        ps.swl = ps.sampler.state & 0x00ffffffu;
    }
    ps.sampler.start(px, py, frame, 1u);
    ps.ray_medium = S.params.process_mediums != 0 ? S.cam.medium_id : VB_INVALID_ID;
    ps.scatter_pdf = 1e16f;
    ps.throughput = f3{1, 1, 1}; ps.ret = f3{0, 0, 0};
    ps.eta_scale = 1.f;
    ps.bounces = 0;
}

enum : int { PT_CONTINUE = 1, PT_SHADOW = 2, PT_VERTEX = 4, PT_LIGHT = 8 };


// in: the ray that was traced and its closest hit. out (by flag): PT_CONTINUE -> `ray` holds the next ray
// and ps.bounces is advanced (max_depth is enforced here); PT_SHADOW ->
// `shadow` must be tested for occlusion and `pending` added to ps.ret when it is free; PT_VERTEX -> a
// surface interaction was computed, PT_LIGHT -> a light was sampled (statistics: a PT_LIGHT vertex without
// PT_SHADOW is a shadow ray the reference traces although its contribution is zero; it is culled here). Without PT_CONTINUE the path is finished: ps.ret
// (+ pending) is the sample's radiance.
// PRINCIPLED = the scene uses the principled BSDF: its LobeSet code (stack frame, out-of-line calls) is compiled
// into a separate kernel instantiation so that scenes without it keep the lean kernel.
// VARIANT = bit flags: PT_PRINCIPLED = the principled BSDF's code is compiled in, PT_HERO = hero spectrum: only those
// variants carry the path's wavelengths; without the flag `swl` is a compile-time nullptr and every spec_* call folds away.
// PT_LOBESET (only with PT_PRINCIPLED) = the mix / add materials' code (vb_mixset.cuh: two children, each possibly a principled set) is
// compiled in as well; scenes with principled materials but no mix / add keep the smaller frame of the plain principled kernels.
enum : int { PT_LEAN = 0, PT_PRINCIPLED = 1, PT_HERO = 2, PT_LOBESET = 4 };
template <int VARIANT>
__device__ __forceinline__ int pt_bounce(const PtScene &S, PathState &ps, VbRay &ray, const VbHit &hit, VbRay &shadow, f3 &pending) {
    constexpr bool PRINCIPLED = (VARIANT & PT_PRINCIPLED) != 0, HERO = (VARIANT & PT_HERO) != 0, LOBESET = (VARIANT & PT_LOBESET) != 0;
    Swl wl;
    const Swl *swl = nullptr;
    if (HERO) { wl = swl_make(ps.swl); swl = &wl; }
    Sampler &sampler = ps.sampler;
    const bool mediums = S.params.process_mediums != 0;
    const uint32_t bounces = ps.bounces;
    if (hit.inst == VB_INVALID_ID) {                          // evaluate_miss, integrator.cpp:137-158 + lightsampler.cpp:289-300
        if (S.env_light != VB_INVALID_ID) {
            const vb_light_rec &l = S.lights[S.env_light];
            f3 tr{1, 1, 1};
            if (mediums) { ray.tmax = l.params[0]; tr = ray_Tr(S, swl, ps.ray_medium, ray); }
            f3 ro{ray.ox, ray.oy, ray.oz}, rd{ray.dx, ray.dy, ray.dz};
            LightEval ev = l.type == VB_LIGHT_DIRECTIONAL ? directional_evaluate_wi(S, swl, l) : env_evaluate_wi(S, swl, l, ro, ro + rd);
            ev.pdf *= light_pmf(S, S.env_light);
            float weight = mis_weight(ps.scatter_pdf, ev.pdf);
            if (S.params.mis_mode == 2) weight = 1.f;
            else if (S.params.mis_mode == 1) weight = bounces == 0 ? weight : 0.f;
            ps.ret = ps.ret + spec_to_srgb(S.spec, swl, ev.L * tr * weight * ps.throughput);
        }
        return 0;
    }
    Interaction it = compute_surface_interaction(S, hit.inst, hit.prim, hit.u, hit.v, true);
    f3 rd{ray.dx, ray.dy, ray.dz}, ro{ray.ox, ray.oy, ray.oz};
    it.wo = normalize(-rd);
    ray.tmax = length(it.pos - ro) / length(rd);             // geometry.h:64-69
    if (mediums && ps.ray_medium != VB_INVALID_ID) {         // integrator.cpp:199-206, homogeneous.cpp:50-70
        const vb_medium_rec &md = S.mediums[ps.ray_medium];
        f3 sigma_t = medium_sigma_t(S, swl, md);
        f3 sigma_s = spec_unbound_c(S.spec, swl, f3{md.sigma_s[0] * md.scale, md.sigma_s[1] * md.scale, md.sigma_s[2] * md.scale}, true,
                                    S.spec.cache_mediums + 2u * ps.ray_medium + 1u);
        uint32_t channel = min((uint32_t)(sampler.next_1d() * 3.0f), 2u);
        float sc_t = channel == 0 ? sigma_t.x : (channel == 1 ? sigma_t.y : sigma_t.z);
        float dist = -vlog(1.0f - sampler.next_1d()) / sc_t;
        float t = fmin_(dist / length(rd), ray.tmax);
        bool sampled_medium = t < ray.tmax;
        if (sampled_medium) {
            it.pos = ro + rd * t;
            it.wo = -rd;
            it.ng = f3{0, 0, 0};
            it.uv = f2{0, 0};
            it.shading.x = f3{0, 0, 0}; it.shading.y = f3{0, 0, 0}; it.shading.z = f3{0, 0, 0};
            it.prim_area = 0.f;
            it.prim_id = VB_INVALID_ID; it.mat_id = VB_INVALID_ID; it.light_id = VB_INVALID_ID;
            it.has_phase = true; it.g = md.g;
            it.inside = ps.ray_medium; it.outside = ps.ray_medium;
        }
        f3 tr = medium_Tr(S, swl, md, t);
        f3 density = sampled_medium ? sigma_t * tr : tr;
        float pdf = average3(density);
        f3 mt = sampled_medium ? tr * sigma_s / pdf : tr / pdf;
        ps.throughput = ps.throughput * mt;
    }
    if (it.mat_id == VB_INVALID_ID && !it.has_phase) {       // integrator.cpp:208-214: medium boundary, bounce not counted
        ray = spawn_ray(it, rd);
        if (mediums) ps.ray_medium = dot(it.ng, rd) > 0.f ? it.outside : it.inside;
        return PT_CONTINUE | PT_VERTEX;
    }
    if (it.light_id != VB_INVALID_ID) {                      // integrator.cpp:221-231
        const vb_light_rec &l = S.lights[it.light_id];
        LightEval ev{f3{0, 0, 0}, 0.f};
        if (l.type == VB_LIGHT_AREA) {
            LightEvalCtx pl{it.pos, it.ng, it.uv, 1.0f / it.prim_area};
            pl.pdf_pos *= alias_pmf(S, l, it.prim_id);
            ev = area_evaluate_wi(S, swl, l, ro, pl);
            ev.pdf *= light_pmf(S, it.light_id);
        }
        f3 tr = ray_Tr(S, swl, ps.ray_medium, ray);
        float weight = mis_weight(ps.scatter_pdf, ev.pdf);
        if (S.params.mis_mode == 2) weight = 1.f;
        else if (S.params.mis_mode == 1) weight = bounces == 0 ? weight : 0.f;
        f3 d = ev.L * ps.throughput * weight * tr;
        ps.ret = ps.ret + spec_to_srgb(S.spec, swl, d);
    }
    // Li with max_depth < 2 and MIS "both" appends ONE mis_bsdf(inner = false) (integrator.cpp:176-233,303-309): the BSDF-sampled ray
    // is traced and only emission / environment are gathered with the BSDF-side MIS weight - no light sample, no further scatter.
    // The reference starts that step whenever the final ray's direction differs from the primary ray's (:179-184); here the bit is
    // set exactly when a scatter produced that ray, which is the same thing unless a sampled direction equals the primary
    // direction bit for bit (the oracle applies the reference's rule; the parity tests would show such a path)
    if (ps.swl & PathSupplementBit) return PT_VERTEX;
    // light sampling
    float u_light = sampler.next_1d();
    f2 u_surface = sampler.next_2d();
    LightSample ls{{f3{0, 0, 0}, 0.f}, f3{0, 0, 0}};
    if (S.n_lights) {
        uint32_t index = light_select(S, u_light);
        const vb_light_rec &l = S.lights[index];
        if (l.type == VB_LIGHT_AREA) ls = area_sample_wi(S, swl, l, it, u_surface);
        else if (l.type == VB_LIGHT_ENV) ls = env_sample_wi(S, swl, l, it, u_surface);
        else if (l.type == VB_LIGHT_DIRECTIONAL) ls = directional_sample_wi(S, swl, l, it);
        else ls = point_sample_wi(S, swl, l, it);
        ls.eval.pdf *= light_pmf(S, index);
    }
    shadow = spawn_ray_to(it, ls.p_light);
    uint32_t shadow_medium = mediums ? (dot(it.ng, f3{shadow.dx, shadow.dy, shadow.dz}) > 0.f ? it.outside : it.inside) : VB_INVALID_ID;
    f3 tr = ray_Tr(S, swl, shadow_medium, shadow);
    // material / phase function, integrator.cpp:246-283 + direct_lighting :20-37
    f3 wi = normalize(ls.p_light - it.pos);
    ScatterEval se;
    BSDFSample bs;
    if (mediums && it.has_phase) {
        float f = phase_HG(dot(it.wo, wi), it.g);
        se.f = f3{f, f, f}; se.pdf = f; se.flags = 0;
        bs = phase_sample(it.g, it.wo, sampler);
    } else if (LOBESET && (S.materials[it.mat_id].type == VB_MAT_MIX || S.materials[it.mat_id].type == VB_MAT_ADD)) {
        mixset_shade(S, swl, it, wi, sampler, se, bs);
        hemisphere_discard(se, it.wo, wi, it.ng);
        hemisphere_discard(bs.eval, it.wo, bs.wi, it.ng);
    } else if (PRINCIPLED && S.materials[it.mat_id].type == VB_MAT_PRINCIPLED) {
        principled_shade(S, swl, it, wi, sampler, se, bs);
        hemisphere_discard(se, it.wo, wi, it.ng);
        hemisphere_discard(bs.eval, it.wo, bs.wi, it.ng);
    } else {
        Lobe lobe = create_lobe(S, swl, it);
        if (HERO && lobe.dispersive) {                             // SampledWavelengths::check_dispersive, integrator.cpp:264 + spectrum.cpp:21-38
            ps.swl |= SwlInvalidBit;
            wl.pdf[1] = 0.f; wl.pdf[2] = 0.f;
        }
        se = lobe_evaluate(S, lobe, it.wo, wi, false, nullptr);
        hemisphere_discard(se, it.wo, wi, it.ng);
        bs = lobe_sample(S, lobe, it.wo, sampler);
        hemisphere_discard(bs.eval, it.wo, bs.wi, it.ng);
    }
    bool is_delta = ls.eval.pdf < 0.f;
    float weight = is_delta ? 1.f : mis_weight(ls.eval.pdf, se.pdf);
    if (S.params.mis_mode == 1) weight = 1.f;
    float lpdf = is_delta ? -ls.eval.pdf : ls.eval.pdf;
    // the contribution counts only when the shadow ray is free; otherwise Ld = 0 and ret is unchanged
    int flags = PT_VERTEX | PT_LIGHT;
    pending = f3{0, 0, 0};
    if (se.pdf > 0.f && lpdf > 0.f && S.params.mis_mode != 2) {
        f3 Ld = ls.eval.L * se.f * weight / lpdf;
        pending = spec_to_srgb(S.spec, swl, ps.throughput * Ld * tr);
        if (pending.x != 0.f || pending.y != 0.f || pending.z != 0.f) flags |= PT_SHADOW;      // ret + 0 == ret
    }
    ps.eta_scale *= sqr(bs.eta);
    float lum = max_comp(ps.throughput);
    if (!(bs.eval.pdf > 0.f) || lum == 0.f) return flags;
    ps.throughput = ps.throughput * (bs.eval.f / bs.eval.pdf);
    if (ps.eta_scale * lum < S.params.rr_threshold && bounces >= S.params.min_depth) {
        float q = fmin_(0.95f, lum);
        float rr = sampler.next_1d();
        if (q < rr) return flags;
        ps.throughput = ps.throughput / q;
    }
    ps.scatter_pdf = bs.eval.pdf;
    ray = spawn_ray(it, bs.wi);
    if (mediums) ps.ray_medium = dot(it.ng, bs.wi) > 0.f ? it.outside : it.inside;
    ps.bounces = bounces + 1;
    if (ps.bounces >= S.params.max_depth) {                       // depth exhausted: the path ends here ...
        if (!(S.params.max_depth < 2 && S.params.mis_mode == 0)) return flags;
        ps.swl |= PathSupplementBit;                              // ... unless Li's only_direct supplement applies
    }
    return flags | PT_CONTINUE;
}

}  // namespace

// mix / add materials on the device: statement-by-statement mirror of the MixSet part of oracle/vo_bsdf.inl
// (reference src/render_core/material/mix.cpp:52-57, add.cpp:50-53; LobeSet::create_mix / create_add / flatten / evaluate_impl /
// sample_wi_impl, src/base/scattering/lobe.cpp:495-562,598-683). A LobeSet over the lobes of two child materials: each child is a
// single Lobe or a principled set whose sub-lobes are pulled up (flatten) with sample_weight *= parent sample weight and
// weight *= parent SAMPLE weight. Children are walked in order, so the flattened list never has to be materialised.
// Included by vb_principled.cuh; compiled only into the PT_LOBESET kernel instantiations (the "heavy" shading launch of scenes that have mix / add).
#pragma once

namespace {

struct MixChild { bool principled; float sw, w; Lobe lb; PrincipledSet ps; };
struct MixSet { MixChild c[2]; };

__device__ __forceinline__ void mixset_create(const PtScene &S, const Swl *swl, const Interaction &it, MixSet &ms) {
    const vb_material_rec &m = S.materials[it.mat_id];
    float psw[2], pw[2];
    if (m.type == VB_MAT_MIX) {
        float frac = eval_slot3(S, m.sigma, it.uv).x;
        psw[0] = pw[0] = 1 - frac; psw[1] = pw[1] = frac;
    } else {
        float sum = 0.f; sum += 1.f; sum += 1.f;                        // normalize_sampled_weight
        psw[0] = 1.f / sum; psw[1] = 1.f / sum; pw[0] = pw[1] = 1.f;
    }
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
        const uint32_t child = m.spd[c];
        MixChild &ch = ms.c[c];
        ch.sw = psw[c]; ch.w = pw[c];
        ch.principled = S.materials[child].type == VB_MAT_PRINCIPLED;
        if (ch.principled) ch.ps = principled_create_of(S, swl, it, child);
        else ch.lb = create_lobe_of(S, swl, it, child);
    }
}

__device__ __noinline__ ScatterEval mixset_evaluate(const PtScene &S, const MixSet &ms, f3 world_wo, f3 world_wi, float *eta_out) {
    ScatterEval ret{f3{0, 0, 0}, 0.f, 0u};
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
        const MixChild &ch = ms.c[c];
        if (ch.principled) {
            const PrincipledSet &ps = ch.ps;
            f3 wo = ps.frame.to_local(world_wo), wi = ps.frame.to_local(world_wi);
            VB_EACH_SUBLOBE(ps, {
                ScatterEval se = sub_evaluate_local(S, l, wo, wi, eta_out);
                se.f = se.f * fabsf(wi.z);
                float factor = l.kind == SL_TRANS ? 1.f : (dot(world_wo, ps.frame.z) * dot(world_wi, ps.frame.z) > 0.f ? 1.f : 0.f);
                float w = l.weight * ch.sw, sw = l.sample_weight * ch.sw;          // flatten
                se.f = se.f * (w * factor);
                se.pdf = se.pdf * (sw * factor);
                ret.f = ret.f + se.f;
                ret.pdf = ret.pdf + se.pdf;
                ret.flags |= se.flags;
            })
        } else {
            ScatterEval se = lobe_evaluate(S, ch.lb, world_wo, world_wi, true, eta_out);
            float factor = ch.lb.type == VB_MAT_GLASS ? 1.f : (dot(world_wo, ch.lb.frame.z) * dot(world_wi, ch.lb.frame.z) > 0.f ? 1.f : 0.f);
            se.f = se.f * (ch.w * factor);
            se.pdf = se.pdf * (ch.sw * factor);
            ret.f = ret.f + se.f;
            ret.pdf = ret.pdf + se.pdf;
            ret.flags |= se.flags;
        }
    }
    return ret;
}

__device__ __forceinline__ f3 mixset_sample_wi(const MixSet &ms, f3 world_wo, Sampler &sampler, bool &valid) {
    float uc = sampler.next_1d();
    (void)sampler.next_2d();
    int chosen_c = -1, chosen_k = -1;
    float sum_weights = 0.f;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
        const MixChild &ch = ms.c[c];
        if (ch.principled) {
#pragma unroll
            for (int k = 0; k < 6; ++k)
                if (ch.ps.switches & (1u << k)) {
                    if (chosen_c < 0 || uc > sum_weights) { chosen_c = c; chosen_k = k; }
                    sum_weights += ch.ps.sw[k] * ch.sw;
                }
        } else {
            if (chosen_c < 0 || uc > sum_weights) { chosen_c = c; chosen_k = -1; }
            sum_weights += ch.sw;
        }
    }
    const MixChild &ch = ms.c[chosen_c < 0 ? 0 : chosen_c];
    if (chosen_k >= 0) return principled_sub_wi(ch.ps, chosen_k, world_wo, sampler, valid);
    return lobe_sample_wi(ch.lb, world_wo, sampler, valid);
}

// one call per shaded vertex, like principled_shade
__device__ __noinline__ void mixset_shade(const PtScene &S, const Swl *swl, const Interaction &it, f3 wi_light, Sampler &sampler,
                                          ScatterEval &se, BSDFSample &bs) {
    MixSet ms;
    mixset_create(S, swl, it, ms);
    se = mixset_evaluate(S, ms, it.wo, wi_light, nullptr);
    bool valid = true;
    bs.eta = 1.f;
    bs.wi = mixset_sample_wi(ms, it.wo, sampler, valid);
    bs.eval = mixset_evaluate(S, ms, it.wo, bs.wi, &bs.eta);
    if (!valid) bs.eval.pdf *= 0.f;
}

__device__ __noinline__ f3 mixset_albedo(const PtScene &S, const Swl *swl, const Interaction &it, float cos_theta) {
    MixSet ms;
    mixset_create(S, swl, it, ms);
    f3 ret{0, 0, 0};
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
        const MixChild &ch = ms.c[c];
        if (ch.principled) {
            VB_EACH_SUBLOBE(ch.ps, { ret = ret + sub_albedo(S, l, cos_theta) * (l.weight * ch.sw); })
        } else ret = ret + lobe_albedo(ch.lb, cos_theta) * ch.w;
    }
    return ret;
}

}  // namespace

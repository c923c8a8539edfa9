// Path-tracing megakernel (sm_100a) + film kernels + their C ABI.
// One launch = camera ray (reference "rt_geom", src/base/sensor/frame_buffer.cpp:156-183) + Li
// (reference "path_tracing", src/base/integral/integrator.cpp:82-107,160-311) + film accumulation
// (reference "RGBFilm-accumulation", frame_buffer.cpp:117-126) for a batch of frame indices.
// Mirrors oracle/vo_pt_li.inl operation for operation; compiled with -fmad=false (see vb_math.cuh).
#include "vb_internal.h"
#include "vb_trace.cuh"
#include "vb_math.cuh"
#include <vector>
#include <cstring>

using namespace vbm;

namespace {

struct DevMesh { const float4 *vtx; const uint4 *tri4; };      // interleaved vertices (pos.xyz, u) (nrm.xyz, v); (i0, i1, i2, 0) per triangle
struct DevInst {
    uint32_t mesh_id, mat_id, light_id, inside, outside, pad[3];
    float o2w[16];
    float nmat[12];            // 9 used: sign(det) * cofactor(o2w 3x3), row-major
};
struct DevTex { const float4 *texels; uint32_t w, h; };
struct DevLut { const float *data; uint32_t nx, ny, nz, channels; };

}  // namespace
#include "vb_spectrum.cuh"
namespace {

struct PtScene {
    VbAccelView accel;
    const DevMesh *meshes;
    const DevInst *instances;
    const vb_material_rec *materials;
    const vb_light_rec *lights;
    uint32_t n_lights;
    uint32_t env_light;            // LightSampler::env_light_ (index of the last Infinite light) or VB_INVALID_ID
    const vb_alias_entry *alias_table;
    const float *alias_func;
    uint32_t lsel_offset, lsel_size;   // light selection table of the "power" light sampler (vb_light_sampler_set); size 0 = uniform
    float lsel_integral;
    uint32_t env_separate;         // vb_light_sampler_env_set (only set when env_light is valid)
    float env_prob;
    const DevTex *textures;
    DevLut luts[VB_LUT_COUNT];
    const vb_medium_rec *mediums;
    DevSpectrum spec;
    vb_camera cam;
    vb_pt_params params;
    uint32_t width, height;
};

constexpr int PT_BLOCK = 128;

struct Frame3 {
    f3 x, y, z;
    __device__ f3 to_local(f3 v) const { return {dot(v, x), dot(v, y), dot(v, z)}; }
    __device__ f3 to_world(f3 v) const { return x * v.x + y * v.y + z * v.z; }
    __device__ void update(f3 n) {
        z = n;
        x = normalize(cross(n, y)) * length(x);
        y = normalize(cross(n, x)) * length(y);
    }
};

struct Interaction {
    f3 pos, wo, ng;
    f2 uv;
    Frame3 shading;
    float prim_area;
    uint32_t prim_id, mat_id, light_id;
    uint32_t inside, outside;      // medium interface (Interaction::set_medium)
    bool has_phase;                // medium interaction (Interaction::init_phase, interaction.cpp:58-63)
    float g;
};

__device__ __forceinline__ f3 ld3(const float *p, uint32_t i) { return {__ldg(p + 3 * i), __ldg(p + 3 * i + 1), __ldg(p + 3 * i + 2)}; }
__device__ __forceinline__ f2 ld2(const float *p, uint32_t i) { return {__ldg(p + 2 * i), __ldg(p + 2 * i + 1)}; }
__device__ __forceinline__ f3 lerp3(float u, float v, f3 a, f3 b, f3 c) {
    float w = 1.0f - u - v;
    return a * w + b * u + c * v;
}
__device__ __forceinline__ bool is_zero(f3 v) { return v.x == 0.f && v.y == 0.f && v.z == 0.f; }

// reference src/base/mgr/geometry.cpp:79-166
__device__ __forceinline__ Interaction compute_surface_interaction(const PtScene &S, uint32_t inst_id, uint32_t prim_id, float bu, float bv,
                                                   bool is_complete) {
    Interaction it;
    it.prim_id = prim_id;
    const DevInst &I = S.instances[inst_id];
    const DevMesh M = S.meshes[I.mesh_id];
    const uint4 ti = __ldg(M.tri4 + prim_id);
    const uint32_t i0 = ti.x, i1 = ti.y, i2 = ti.z;
    // one 32-byte sector per vertex: (pos.xyz, u) (nrm.xyz, v)
    uint4 va0, vb0, va1, vb1, va2, vb2;
    vb_ld256(reinterpret_cast<const uint4 *>(M.vtx) + 2 * (size_t)i0, va0, vb0);
    vb_ld256(reinterpret_cast<const uint4 *>(M.vtx) + 2 * (size_t)i1, va1, vb1);
    vb_ld256(reinterpret_cast<const uint4 *>(M.vtx) + 2 * (size_t)i2, va2, vb2);
    auto xyz = [](const uint4 &q) { return f3{__uint_as_float(q.x), __uint_as_float(q.y), __uint_as_float(q.z)}; };
    it.light_id = I.light_id; it.mat_id = I.mat_id;
    it.inside = I.inside; it.outside = I.outside; it.has_phase = false; it.g = 0.f;
    f3 p0 = transform_point(I.o2w, xyz(va0)), p1 = transform_point(I.o2w, xyz(va1)), p2 = transform_point(I.o2w, xyz(va2));
    it.pos = lerp3(bu, bv, p0, p1, p2);
    f3 dp02 = p0 - p2, dp12 = p1 - p2;
    f3 ng_un = cross(dp02, dp12);
    it.prim_area = 0.5f * length(ng_un);
    f2 t0{__uint_as_float(va0.w), __uint_as_float(vb0.w)}, t1{__uint_as_float(va1.w), __uint_as_float(vb1.w)}, t2{__uint_as_float(va2.w), __uint_as_float(vb2.w)};
    f2 duv02{t0.x - t2.x, t0.y - t2.y}, duv12{t1.x - t2.x, t1.y - t2.y};
    float det = duv02.x * duv12.y - duv02.y * duv12.x;
    bool degenerate_uv = fabsf(det) < 1e-8f;
    Frame3 frame;
    if (is_complete) {
        f3 dp_du, dp_dv;
        if (!degenerate_uv) {
            float inv_det = 1.0f / det;
            dp_du = normalize((dp02 * duv12.y - dp12 * duv02.y) * inv_det);
            dp_dv = normalize((dp02 * (-duv12.x) + dp12 * duv02.x) * inv_det);
        } else {
            dp_du = normalize(p1 - p0);
            dp_dv = normalize(p2 - p0);
        }
        frame.x = dp_du; frame.y = dp_dv; frame.z = normalize(ng_un);
    } else {
        frame.x = dp02; frame.y = dp12; frame.z = normalize(ng_un);
    }
    it.shading = frame;
    if (is_complete) {
        f3 normal = lerp3(bu, bv, xyz(vb0), xyz(vb1), xyz(vb2));
        if (!is_zero(normal)) {
            f3 ns = normalize(apply_normal(I.nmat, normal));
            it.shading.update(ns);
        }
    }
    float w = 1.0f - bu - bv;
    it.uv = {t0.x * w + t1.x * bu + t2.x * bv, t0.y * w + t1.y * bu + t2.y * bv};
    it.ng = frame.z;
    return it;
}

__device__ __forceinline__ f3 flip_to(f3 n, f3 dir) { return dot(n, dir) > 0.f ? n : -n; }

__device__ __forceinline__ VbRay make_ray(f3 o, f3 d, float tmax) { return VbRay{o.x, o.y, o.z, 0.f, d.x, d.y, d.z, tmax}; }
__device__ __forceinline__ VbRay spawn_ray(const Interaction &it, f3 dir) {
    return make_ray(offset_ray_origin(it.pos, flip_to(it.ng, dir)), dir, RayTMax);
}
__device__ __forceinline__ VbRay spawn_ray_to(const Interaction &it, f3 target) {
    f3 dir = target - it.pos;
    return make_ray(offset_ray_origin(it.pos, flip_to(it.ng, dir)), dir, 1.0f - ShadowEpsilon);
}

// bilinear, wrap, normalised uv - mirrors vo sample_texture
__device__ float4 sample_texture(const DevTex &t, f2 uv) {
    float x = uv.x * (float)t.w - 0.5f, y = uv.y * (float)t.h - 0.5f;
    float fx0 = floorf(x), fy0 = floorf(y);
    float fx = x - fx0, fy = y - fy0;
    int W = (int)t.w, H = (int)t.h;
    int x0 = (int)fx0 % W; if (x0 < 0) x0 += W;
    int x1 = ((int)fx0 + 1) % W; if (x1 < 0) x1 += W;
    int y0 = (int)fy0 % H; if (y0 < 0) y0 += H;
    int y1 = ((int)fy0 + 1) % H; if (y1 < 0) y1 += H;
    float4 a = __ldg(t.texels + (size_t)y0 * W + x0), b = __ldg(t.texels + (size_t)y0 * W + x1);
    float4 c = __ldg(t.texels + (size_t)y1 * W + x0), d = __ldg(t.texels + (size_t)y1 * W + x1);
    float4 ab{a.x + fx * (b.x - a.x), a.y + fx * (b.y - a.y), a.z + fx * (b.z - a.z), a.w + fx * (b.w - a.w)};
    float4 cd{c.x + fx * (d.x - c.x), c.y + fx * (d.y - c.y), c.z + fx * (d.z - c.z), c.w + fx * (d.w - c.w)};
    return float4{ab.x + fy * (cd.x - ab.x), ab.y + fy * (cd.y - ab.y), ab.z + fy * (cd.z - ab.z), ab.w + fy * (cd.w - ab.w)};
}
__device__ f3 eval_slot3(const PtScene &S, const vb_slot &s, f2 uv) {
    f3 v{s.value[0], s.value[1], s.value[2]};
    if (s.tex_id != VB_INVALID_ID) {
        float4 t = sample_texture(S.textures[s.tex_id], uv);
        v = {v.x * t.x, v.y * t.y, v.z * t.z};
    }
    return v;
}

// Light::color_ as an illumination spectrum (eval_illumination_spectrum, shader_node.cpp:329-333); `l` is an element of S.lights
__device__ __forceinline__ f3 light_color(const PtScene &S, const Swl *swl, const vb_light_rec &l, f2 uv) {
    return spec_illum_c(S.spec, swl, eval_slot3(S, l.color, uv), l.color.tex_id == VB_INVALID_ID, S.spec.cache_lights + (uint32_t)(&l - S.lights));
}
struct LightEvalCtx { f3 pos, ng; f2 uv; float pdf_pos; };
struct LightEval { f3 L; float pdf; };
struct LightSample { LightEval eval; f3 p_light; };

__device__ __forceinline__ float alias_pmf(const PtScene &S, const vb_light_rec &l, uint32_t i) {
    return l.alias_integral > 0.f ? __ldg(S.alias_func + l.alias_offset + i) / (l.alias_integral * (float)l.alias_size) : 0.f;
}
__device__ uint32_t alias_sample(const PtScene &S, const vb_light_rec &l, float u, float &pmf, float &u_remapped) {
    uint32_t size = l.alias_size;
    u = u * (float)size;
    uint32_t idx = min((uint32_t)u, size - 1);
    u = fmin_(u - (float)idx, OneMinusEpsilon);
    vb_alias_entry e = S.alias_table[l.alias_offset + idx];
    bool keep = u < e.prob;
    uint32_t out = keep ? idx : e.alias;
    u_remapped = keep ? fmin_(u / e.prob, OneMinusEpsilon) : fmin_((1.0f - u) / (1.0f - e.prob), OneMinusEpsilon);
    pmf = alias_pmf(S, l, out);
    return out;
}

// LightSampler::PMF / select_light (lightsampler.cpp:159-197) over PMF_ / select_light_: uniform.cpp:12-33 (1 / N, index = min(u N, N - 1))
// or power.cpp:16-31 (alias table over luminance(light->power()): Warper::PMF / sample_discrete). env_separate: the environment light
// with probability env_prob, the rule over the others otherwise (mirror of oracle/vo_pt_li.inl light_pmf / light_select)
__device__ __forceinline__ float light_pmf_(const PtScene &S, uint32_t index) {
    if (S.lsel_size == 0) return 1.0f / (float)(S.n_lights - (S.env_separate ? 1u : 0u));
    return S.lsel_integral > 0.f ? __ldg(S.alias_func + S.lsel_offset + index) / (S.lsel_integral * (float)S.lsel_size) : 0.f;
}
__device__ __forceinline__ float light_pmf(const PtScene &S, uint32_t index) {
    if (S.env_separate) return index == S.env_light ? S.env_prob : (1 - S.env_prob) * light_pmf_(S, index);
    return light_pmf_(S, index);
}
__device__ __forceinline__ uint32_t light_select(const PtScene &S, float u) {
    uint32_t n = S.n_lights;
    if (S.env_separate) {
        if (u < S.env_prob) return S.env_light;
        u = (u - S.env_prob) / (1 - S.env_prob);
    }
    if (S.lsel_size == 0) {
        if (S.env_separate) {
            float punctual = (float)(n - 1u);
            uint32_t index = (uint32_t)fmin_(u * punctual, punctual - 1);
            return index < S.env_light ? index : index + 1u;
        }
        return (uint32_t)fmin_(u * (float)n, (float)n - 1.0f);
    }
    u = u * (float)n;
    uint32_t idx = min((uint32_t)u, n - 1);
    u = fmin_(u - (float)idx, OneMinusEpsilon);
    vb_alias_entry e = S.alias_table[S.lsel_offset + idx];
    return u < e.prob ? idx : e.alias;
}

__device__ LightEval area_evaluate_wi(const PtScene &S, const Swl *swl, const vb_light_rec &l, f3 ref_pos, const LightEvalCtx &pl) {
    LightEval ev;
    f3 w = ref_pos - pl.pos;
    f3 color = light_color(S, swl, l, pl.uv);
    float s = (dot(w, pl.ng) > 0.f || l.two_sided) ? 1.f : 0.f;
    ev.L = color * l.scale * s;
    float cos_theta = fabsf(dot(pl.ng, normalize(w)));
    float pdf = pl.pdf_pos * length_squared(w) / cos_theta;
    ev.pdf = (isinf(pdf) || isnan(pdf)) ? 0.f : pdf;
    return ev;
}

__device__ LightSample area_sample_wi(const PtScene &S, const Swl *swl, const vb_light_rec &l, const Interaction &ref, f2 u) {
    float pmf, ux;
    uint32_t prim = alias_sample(S, l, u.x, pmf, ux);
    float su0 = __fsqrt_rn(ux);
    float b0 = 1.0f - su0, b1 = u.y * su0;
    Interaction li = compute_surface_interaction(S, l.inst_id, prim, b0, b1, false);
    LightEvalCtx pl{li.pos, li.ng, li.uv, 1.0f / li.prim_area};
    pl.pdf_pos *= pmf;
    LightSample ls;
    ls.eval = area_evaluate_wi(S, swl, l, ref.pos, pl);
    f3 dir = ref.pos - pl.pos;
    float factor = dot(pl.ng, dir) > 0.f ? 1.f : -1.f;
    ls.p_light = offset_ray_origin(pl.pos, pl.ng * factor * S.params.ray_offset_factor);
    return ls;
}

// ---- directional light: render_core/light/environments/directional.cpp:46-80 ----
__device__ __forceinline__ LightEval directional_evaluate_wi(const PtScene &S, const Swl *swl, const vb_light_rec &l) {
    LightEval ev;
    ev.L = spec_illum(S.spec, swl, eval_slot3(S, l.color, f2{0, 0})) * l.scale;
    ev.pdf = -1.f;
    return ev;
}
__device__ __forceinline__ LightSample directional_sample_wi(const PtScene &S, const Swl *swl, const vb_light_rec &l, const Interaction &ref) {
    LightSample ls;
    f3 w_light = -f3{l.direction[0], l.direction[1], l.direction[2]};
    ls.p_light = ref.pos + w_light * l.params[1];
    ls.eval = directional_evaluate_wi(S, swl, l);
    return ls;
}

// ---- point / spot lights: render_core/light/point.cpp:36-42, spot.cpp:47-70, light.cpp:49-58 ----
__device__ LightSample point_sample_wi(const PtScene &S, const Swl *swl, const vb_light_rec &l, const Interaction &ref) {
    LightSample ls;
    f3 pos{l.position[0], l.position[1], l.position[2]};
    f3 value = light_color(S, swl, l, f2{0, 0}) * l.scale;
    f3 w_un = ref.pos - pos;
    if (l.type == VB_LIGHT_POINT) ls.eval.L = value / length_squared(w_un);
    else if (l.type == VB_LIGHT_PROJECTOR) {                       // projector.cpp:99-112
        f3 p = transform_point(l.w2o, ref.pos);
        float d2 = length_squared(p);
        bool valid = p.z > 0.f;
        p = p / p.z;
        float tan_y = l.params[0], tan_x = l.params[1] * tan_y;
        f2 uv{(p.x + tan_x) / (2.f * tan_x), (p.y + tan_y) / (2.f * tan_y)};
        valid = valid && uv.x >= 0.f && uv.x <= 1.f && uv.y >= 0.f && uv.y <= 1.f;
        if (!valid) uv = f2{0.f, 0.f};
        ls.eval.L = light_color(S, swl, l, uv) * (valid ? 1.f : 0.f) / d2 * l.scale;
    } else {
        f3 w = normalize(w_un);
        f3 dir{l.direction[0], l.direction[1], l.direction[2]};
        float cos_angle = l.params[0], cos_falloff_start = l.params[1];
        float cos_theta = fmin_(fmax_(dot(dir, w), cos_angle), cos_falloff_start);
        float factor = (cos_theta - cos_angle) / (cos_falloff_start - cos_angle);
        float f2_ = factor * factor;
        ls.eval.L = value / length_squared(w_un) * (f2_ * f2_);
    }
    ls.eval.pdf = -1.f;
    ls.p_light = pos;
    return ls;
}

// ---- AliasTable2D over the env map: render_core/warper/alias2d.cpp:92-129 ----
__device__ __forceinline__ float env_map_pdf(const PtScene &S, const vb_light_rec &l, f2 p) {
    uint32_t rx = l.env_res[0], ry = l.env_res[1];
    uint32_t iu = min((uint32_t)(p.x * (float)rx), rx - 1), iv = min((uint32_t)(p.y * (float)ry), ry - 1);
    return l.alias_integral > 0.f ? __ldg(S.alias_func + l.cond_offset + iv * rx + iu) / l.alias_integral : 0.f;
}
__device__ __forceinline__ uint32_t alias_offset_remap(const PtScene &S, uint32_t base, uint32_t size, float u, float &u_remapped) {
    u = u * (float)size;
    uint32_t idx = min((uint32_t)u, size - 1);
    u = fmin_(u - (float)idx, OneMinusEpsilon);
    vb_alias_entry e = S.alias_table[base + idx];
    bool keep = u < e.prob;
    u_remapped = keep ? fmin_(u / e.prob, OneMinusEpsilon) : fmin_((1.0f - u) / (1.0f - e.prob), OneMinusEpsilon);
    return keep ? idx : e.alias;
}
__device__ f2 env_sample_continuous(const PtScene &S, const vb_light_rec &l, f2 u, float &pdf) {
    uint32_t rx = l.env_res[0], ry = l.env_res[1];
    float ur;
    uint32_t iv = alias_offset_remap(S, l.alias_offset, ry, u.y, ur);
    float fv = ((float)iv + ur) / (float)ry;
    float pdf_v = l.alias_integral > 0.f ? __ldg(S.alias_func + l.alias_offset + iv) / l.alias_integral : 0.f;
    uint32_t off = rx * iv;
    uint32_t iu = alias_offset_remap(S, l.cond_offset + off, rx, u.x, ur);
    float fu = ((float)iu + ur) / (float)rx;
    float integral_u = __ldg(S.alias_func + l.alias_offset + iv);
    float func_u = __ldg(S.alias_func + l.cond_offset + off + iu);
    float pdf_u = integral_u > 0.f ? func_u / integral_u : 0.f;
    pdf = pdf_u * pdf_v;
    return {fu, fv};
}

// ---- spherical environment map: render_core/light/environments/spherical.cpp:58-160 ----
__device__ __forceinline__ f2 env_calculate_uv(f3 local_dir) {
    float phi = vatan2(local_dir.y, local_dir.x);
    if (phi < 0.f) phi = phi + 2.0f * Pi;
    float cz = fmin_(fmax_(local_dir.z, -1.f), 1.f);
    float theta = vatan2(safe_sqrt(1.0f - cz * cz), cz);
    return {phi * Inv2Pi, theta * InvPi};
}
__device__ __forceinline__ f3 env_L(const PtScene &S, const Swl *swl, const vb_light_rec &l, f3 local_dir) {
    return light_color(S, swl, l, env_calculate_uv(local_dir)) * l.scale;
}
__device__ LightEval env_evaluate_wi(const PtScene &S, const Swl *swl, const vb_light_rec &l, f3 ref_pos, f3 light_pos) {
    LightEval ev;
    f3 world_dir = normalize(light_pos - ref_pos);
    f3 local_dir = transform_vector(l.w2o, world_dir);
    f2 uv = env_calculate_uv(local_dir);
    float s, c;
    sincos_2pi(uv.y * 0.5f, s, c);
    ev.L = env_L(S, swl, l, local_dir);
    float pdf = env_map_pdf(S, l, uv) / (2.0f * Pi * Pi * s);
    ev.pdf = s == 0.f ? 0.f : pdf;
    return ev;
}
__device__ LightSample env_sample_wi(const PtScene &S, const Swl *swl, const vb_light_rec &l, const Interaction &ref, f2 u) {
    LightSample ls;
    float pdf_map;
    f2 uv = env_sample_continuous(S, l, u, pdf_map);
    float st, ct, sp, cp;
    sincos_2pi(uv.y * 0.5f, st, ct);
    sincos_2pi(uv.x, sp, cp);
    f3 local_dir{st * cp, st * sp, ct};
    f3 world_dir = normalize(transform_vector(l.o2w, local_dir));
    float pdf_dir = pdf_map / (2.0f * Pi * Pi * st);
    ls.eval.pdf = isinf(pdf_dir) ? 0.f : pdf_dir;
    ls.eval.L = env_L(S, swl, l, local_dir);
    ls.p_light = ref.pos + world_dir * l.params[0];
    return ls;
}

struct Sampler {
    uint32_t state;
    __device__ void start(uint32_t px, uint32_t py, uint32_t si, uint32_t dim) { state = tea4(tea4(px, py), tea4(si, dim)); }
    __device__ float next_1d() { return lcg(state); }
    __device__ f2 next_2d() { float x = next_1d(); float y = next_1d(); return {x, y}; }
};

}  // namespace
#include "vb_bsdf.cuh"
namespace {

// ---- homogeneous medium, reference render_core/medium/homogeneous.cpp:30-70 ----
__device__ __forceinline__ f3 medium_sigma_t(const PtScene &S, const Swl *swl, const vb_medium_rec &m) {
    return spec_unbound_c(S.spec, swl, f3{m.sigma_a[0] + m.sigma_s[0], m.sigma_a[1] + m.sigma_s[1], m.sigma_a[2] + m.sigma_s[2]} * m.scale, true,
                          S.spec.cache_mediums + 2u * (uint32_t)(&m - S.mediums));
}
__device__ __forceinline__ f3 medium_Tr(const PtScene &S, const Swl *swl, const vb_medium_rec &m, float t) {
    f3 st = medium_sigma_t(S, swl, m);
    float tt = fmin_(RayTMax, t);
    return {vexp(-st.x * tt), vexp(-st.y * tt), vexp(-st.z * tt)};
}
// Geometry::Tr (geometry.cpp:187-199)
__device__ __forceinline__ f3 ray_Tr(const PtScene &S, const Swl *swl, uint32_t medium, const VbRay &ray) {
    if (!S.params.process_mediums || medium == VB_INVALID_ID) return f3{1, 1, 1};
    return medium_Tr(S, swl, S.mediums[medium], length(f3{ray.dx, ray.dy, ray.dz}) * ray.tmax);
}
__device__ __forceinline__ void coordinate_system(f3 n, f3 &a, f3 &b) {
    if (fabsf(n.x) > fabsf(n.y)) a = f3{-n.z, 0.f, n.x} / __fsqrt_rn(n.x * n.x + n.z * n.z);
    else a = f3{0.f, n.z, -n.y} / __fsqrt_rn(n.y * n.y + n.z * n.z);
    b = cross(n, a);
}
__device__ __forceinline__ float phase_HG(float cos_theta, float g) {
    float denom = 1.0f + sqr(g) + 2.0f * g * cos_theta;
    return Inv4Pi * (1.0f - sqr(g)) / (denom * __fsqrt_rn(denom));
}
// HenyeyGreenstein::sample, interaction.cpp:16-32
__device__ BSDFSample phase_sample(float g, f3 wo, Sampler &sampler) {
    f2 u = sampler.next_2d();
    float sqr_term = (1.0f - sqr(g)) / (1.0f + g - 2.0f * g * u.x);
    float cos_theta = -(1.0f + sqr(g) - sqr(sqr_term)) / (2.0f * g);
    if (fabsf(g) < 1e-3f) cos_theta = 1.0f - 2.0f * u.x;
    float sin_theta = safe_sqrt(1.0f - sqr(cos_theta));
    float s, c;
    sincos_2pi(u.y, s, c);
    f3 v1, v2;
    coordinate_system(wo, v1, v2);
    BSDFSample bs;
    bs.eta = 1.f;
    bs.wi = v1 * (sin_theta * c) + v2 * (sin_theta * s) + wo * cos_theta;
    float f = phase_HG(cos_theta, g);
    bs.eval.f = f3{f, f, f}; bs.eval.pdf = f; bs.eval.flags = 0;
    return bs;
}

__device__ __forceinline__ float mis_weight(float f, float g) { return (1.0f * f) / (1.0f * f + 1.0f * g); }


__device__ __forceinline__ float sample_linear10(float u) {
    float a = 1.f, b = 0.f;
    float x = u * (a + b) / (a + __fsqrt_rn((a * a) + u * ((b * b) - (a * a))));
    return fmin_(x, OneMinusEpsilon);
}
__device__ __forceinline__ float sample_tent(float u, float r) {
    return u < 0.5f ? -r * sample_linear10((0.5f - u) * 2.f) : r * sample_linear10((u - 0.5f) * 2.f);
}

// camera ray: reference sampler.h:65-73, sensor.cpp:44-56, thin_lens.cpp:34-42
__device__ VbRay generate_camera_ray(const PtScene &S, Sampler &sampler, uint32_t px, uint32_t py, f2 *p_film_out = nullptr) {
    const vb_camera &cam = S.cam;
    f2 uf = sampler.next_2d();
    f2 fp;
    if (cam.filter_type == VB_FILTER_TRIANGLE) fp = {sample_tent(uf.x, cam.filter_radius[0]), sample_tent(uf.y, cam.filter_radius[1])};
    else if (cam.filter_type == VB_FILTER_FITTED) {        // FilterSampler::sample, fitted_curve.h:83-91
        f2 u{uf.x * 2.f - 1.f, uf.y * 2.f - 1.f};
        const uint32_t N = VB_FILTER_TABLE_SIZE, off = cam.filter_table_offset;
        float ur;
        uint32_t iv = alias_offset_remap(S, off, N, fabsf(u.y), ur);
        float fv = ((float)iv + ur) / (float)N;
        uint32_t iu = alias_offset_remap(S, off + N + N * iv, N, fabsf(u.x), ur);
        float fu = ((float)iu + ur) / (float)N;
        float sx_ = u.x > 0.f ? 1.f : (u.x < 0.f ? -1.f : 0.f), sy_ = u.y > 0.f ? 1.f : (u.y < 0.f ? -1.f : 0.f);
        fp = {fu * sx_ * cam.filter_radius[0], fv * sy_ * cam.filter_radius[1]};
    } else {
        float rx = cam.filter_radius[0], ry = cam.filter_radius[1];
        fp = {-rx + uf.x * (rx - (-rx)), -ry + uf.y * (ry - (-ry))};
    }
    f2 p_film{(float)px + 0.5f + fp.x, (float)py + 0.5f + fp.y};
    if (p_film_out) *p_film_out = p_film;
    f2 u_lens = sampler.next_2d();
    (void)sampler.next_1d();
    float sx = cam.screen_lo[0] + p_film.x / (float)S.width * (cam.screen_hi[0] - cam.screen_lo[0]);
    float sy = cam.screen_hi[1] - p_film.y / (float)S.height * (cam.screen_hi[1] - cam.screen_lo[1]);
    f3 dir = normalize(f3{sx * cam.tan_fov_y_over_2, sy * cam.tan_fov_y_over_2, 1.0f});
    float rr = __fsqrt_rn(u_lens.x), s, c;
    sincos_2pi(u_lens.y, s, c);
    f2 pl{rr * c * cam.lens_radius, rr * s * cam.lens_radius};
    float ft = cam.focal_distance / dir.z;
    f3 p_focus = dir * ft;
    f3 org{pl.x, pl.y, 0.f};
    dir = normalize(p_focus - org);
    return make_ray(transform_point(cam.c2w, org), transform_vector(cam.c2w, dir), RayTMax);
}

}  // namespace

#include "vb_pt_bounce.cuh"
#include "vb_pt_wavefront.cuh"
#include "vb_pt_kernel.cuh"

// ORACLE — test infrastructure only. Included by vo_pt.cpp (same translation unit).
#include "vo_spectrum.inl"
// Interaction, lights, materials, Li and the per-pixel-sample driver.
namespace vo {
namespace {

inline float3 apply_normal(const InstanceX &I, float3 n) {
    const float *m = I.nmat;
    return {std::fmaf(m[2], n.z, std::fmaf(m[1], n.y, m[0] * n.x)),
            std::fmaf(m[5], n.z, std::fmaf(m[4], n.y, m[3] * n.x)),
            std::fmaf(m[8], n.z, std::fmaf(m[7], n.y, m[6] * n.x))};
}

// reference src/base/mgr/geometry.cpp:79-166
Interaction compute_surface_interaction(const Scene &sc, uint32_t inst_id, uint32_t prim_id, float bu, float bv, bool is_complete) {
    Interaction it;
    it.prim_id = prim_id;
    const InstanceX &I = sc.instances[inst_id];
    const MeshData &m = sc.meshes[I.rec.mesh_id];
    uint32_t i0 = m.tri[3 * prim_id], i1 = m.tri[3 * prim_id + 1], i2 = m.tri[3 * prim_id + 2];
    it.light_id = I.rec.light_id; it.mat_id = I.rec.material_id;
    it.inside = I.rec.inside_medium; it.outside = I.rec.outside_medium;
    float3 p0 = transform_point(I.o2w, m.pos[i0]), p1 = transform_point(I.o2w, m.pos[i1]), p2 = transform_point(I.o2w, m.pos[i2]);
    it.pos = lerp3(bu, bv, p0, p1, p2);
    float3 dp02 = p0 - p2, dp12 = p1 - p2;
    float3 ng_un = cross(dp02, dp12);
    it.prim_area = 0.5f * length(ng_un);
    float2 t0 = m.uv[i0], t1 = m.uv[i1], t2 = m.uv[i2];
    float2 duv02{t0.x - t2.x, t0.y - t2.y}, duv12{t1.x - t2.x, t1.y - t2.y};
    float det = duv02.x * duv12.y - duv02.y * duv12.x;
    bool degenerate_uv = std::fabs(det) < 1e-8f;
    Frame frame;
    if (is_complete) {
        float3 dp_du, dp_dv;
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
        float3 normal = lerp3(bu, bv, m.nrm[i0], m.nrm[i1], m.nrm[i2]);
        if (!is_zero(normal)) {
            float3 ns = normalize(apply_normal(I, normal));
            it.shading.update(ns);
        }
    }
    it.uv = lerp2(bu, bv, t0, t1, t2);
    it.ng = frame.z;
    return it;
}

// previous-camera record of a static camera: inverse of the affine camera_to_world by Gauss-Jordan elimination with row pivoting in
// double precision, rounded once to float (the reference keeps inverse(c2w) / inverse(raster_to_sensor), sensor.cpp:89-93)
inline vb_camera_prev camera_prev_of(const vb_camera &cam) {
    double a[4][8];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { a[r][c] = cam.c2w[c * 4 + r]; a[r][4 + c] = r == c ? 1.0 : 0.0; }
    for (int i = 0; i < 4; ++i) {
        int piv = i;
        for (int r = i + 1; r < 4; ++r) if (std::fabs(a[r][i]) > std::fabs(a[piv][i])) piv = r;
        if (piv != i) for (int c = 0; c < 8; ++c) std::swap(a[i][c], a[piv][c]);
        const double inv = 1.0 / a[i][i];
        for (int c = 0; c < 8; ++c) a[i][c] *= inv;
        for (int r = 0; r < 4; ++r) if (r != i) { const double f = a[r][i]; for (int c = 0; c < 8; ++c) a[r][c] -= f * a[i][c]; }
    }
    vb_camera_prev out{};
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out.w2c[c * 4 + r] = (float)a[r][4 + c];
    out.tan_fov_y_over_2 = cam.tan_fov_y_over_2;
    out.screen_lo[0] = cam.screen_lo[0]; out.screen_lo[1] = cam.screen_lo[1]; out.screen_hi[0] = cam.screen_hi[0]; out.screen_hi[1] = cam.screen_hi[1];
    return out;
}

inline float3 flip_to(float3 n, float3 dir) { return dot(n, dir) > 0.f ? n : -n; }

// reference interaction.h:279-299 / interaction.cpp:115-135
inline Ray spawn_ray(const Interaction &it, float3 dir) {
    float3 n = flip_to(it.ng, dir);
    float3 org = offset_ray_origin(it.pos, n);
    return Ray{org, 0.f, dir, RayTMax};
}
inline Ray spawn_ray_to(const Interaction &it, float3 target) {
    float3 dir = target - it.pos;
    float3 n = flip_to(it.ng, dir);
    float3 org = offset_ray_origin(it.pos, n);
    return Ray{org, 0.f, dir, 1.0f - ShadowEpsilon};
}

struct LightEvalCtx { float3 pos, ng; float2 uv; float pdf_pos; };
struct LightEval { float3 L{0, 0, 0}; float pdf = 0.f; };
struct LightSample { LightEval eval; float3 p_light{0, 0, 0}; };

// ---- alias table, reference warper/alias.h:45-53,135-167 ----
inline float alias_pmf(const Scene &sc, const vb_light_rec &l, uint32_t i) {
    return l.alias_integral > 0.f ? sc.alias_func[l.alias_offset + i] / (l.alias_integral * (float)l.alias_size) : 0.f;
}
inline uint32_t alias_sample(const Scene &sc, const vb_light_rec &l, float u, float &pmf, float &u_remapped) {
    uint32_t size = l.alias_size;
    u = u * (float)size;
    uint32_t idx = std::min((uint32_t)u, size - 1);
    u = std::min(u - (float)idx, OneMinusEpsilon);
    vb_alias_entry e = sc.alias_table[l.alias_offset + idx];
    bool keep = u < e.prob;
    uint32_t out = keep ? idx : e.alias;
    u_remapped = keep ? std::min(u / e.prob, OneMinusEpsilon) : std::min((1.0f - u) / (1.0f - e.prob), OneMinusEpsilon);
    pmf = alias_pmf(sc, l, out);
    return out;
}

// ---- area light, reference render_core/light/area.cpp:92-149, light.h:126-130, warp.h:88-94 ----
inline LightEval area_evaluate_wi(const Scene &sc, const Swl &swl, const vb_light_rec &l, float3 ref_pos, const LightEvalCtx &pl) {
    LightEval ev;
    float3 w = ref_pos - pl.pos;
    float3 color = spec_illum(sc, swl, eval_slot3(sc, l.color, pl.uv));
    float s = (dot(w, pl.ng) > 0.f || l.two_sided) ? 1.f : 0.f;
    ev.L = color * l.scale * s;
    float cos_theta = std::fabs(dot(pl.ng, normalize(w)));
    float pdf = pl.pdf_pos * length_squared(w) / cos_theta;
    ev.pdf = (std::isinf(pdf) || std::isnan(pdf)) ? 0.f : pdf;
    return ev;
}

inline LightSample area_sample_wi(const Scene &sc, const Swl &swl, const vb_light_rec &l, const Interaction &ref, float2 u) {
    float pmf, ux;
    uint32_t prim = alias_sample(sc, l, u.x, pmf, ux);
    float su0 = std::sqrt(ux);
    float b0 = 1.0f - su0, b1 = u.y * su0;                       // square_to_triangle, warp.h:64-68
    Interaction li = compute_surface_interaction(sc, l.inst_id, prim, b0, b1, false);
    LightEvalCtx pl{li.pos, li.ng, li.uv, 1.0f / li.prim_area};
    pl.pdf_pos *= pmf;
    LightSample ls;
    ls.eval = area_evaluate_wi(sc, swl, l, ref.pos, pl);
    float3 dir = ref.pos - pl.pos;
    float factor = dot(pl.ng, dir) > 0.f ? 1.f : -1.f;           // SpacePoint::robust_pos
    ls.p_light = offset_ray_origin(pl.pos, pl.ng * factor * sc.params.ray_offset_factor);
    return ls;
}

// ---- point / spot lights, reference render_core/light/point.cpp:36-42, spot.cpp:47-70, light.cpp:49-58 ----
inline LightSample point_sample_wi(const Scene &sc, const Swl &swl, const vb_light_rec &l, const Interaction &ref) {
    LightSample ls;
    float3 pos{l.position[0], l.position[1], l.position[2]};
    float3 value = spec_illum(sc, swl, eval_slot3(sc, l.color, float2{0, 0})) * l.scale;
    float3 w_un = ref.pos - pos;
    if (l.type == VB_LIGHT_POINT) ls.eval.L = value / length_squared(w_un);
    else if (l.type == VB_LIGHT_PROJECTOR) {                       // projector.cpp:99-112: image projected along the light's +z
        float4x4 w2o; std::memcpy(&w2o, l.w2o, 64);
        float3 p = transform_point(w2o, ref.pos);
        float d2 = length_squared(p);
        bool valid = p.z > 0.f;
        p = p / p.z;
        float tan_y = l.params[0], tan_x = l.params[1] * tan_y;
        float2 uv{(p.x + tan_x) / (2.f * tan_x), (p.y + tan_y) / (2.f * tan_y)};
        valid = valid && uv.x >= 0.f && uv.x <= 1.f && uv.y >= 0.f && uv.y <= 1.f;
        if (!valid) uv = float2{0.f, 0.f};                         // the select() discards the sample; keep the fetch in range
        ls.eval.L = spec_illum(sc, swl, eval_slot3(sc, l.color, uv)) * (valid ? 1.f : 0.f) / d2 * l.scale;
    } else {
        float3 w = normalize(w_un);
        float3 dir{l.direction[0], l.direction[1], l.direction[2]};
        float cos_angle = l.params[0], cos_falloff_start = l.params[1];
        float cos_theta = std::min(std::max(dot(dir, w), cos_angle), cos_falloff_start);
        float factor = (cos_theta - cos_angle) / (cos_falloff_start - cos_angle);
        float f2 = factor * factor;
        ls.eval.L = value / length_squared(w_un) * (f2 * f2);
    }
    ls.eval.pdf = -1.f;                                            // delta light marker (light.h:227-231)
    ls.p_light = pos;
    return ls;
}

// ---- AliasTable2D over the env map, reference render_core/warper/alias2d.cpp:92-129 ----
inline float env_map_pdf(const Scene &sc, const vb_light_rec &l, float2 p) {
    uint32_t rx = l.env_res[0], ry = l.env_res[1];
    uint32_t iu = std::min((uint32_t)(p.x * (float)rx), rx - 1), iv = std::min((uint32_t)(p.y * (float)ry), ry - 1);
    return l.alias_integral > 0.f ? sc.alias_func[l.cond_offset + iv * rx + iu] / l.alias_integral : 0.f;
}
inline uint32_t alias_offset_remap(const Scene &sc, uint32_t base, uint32_t size, float u, float &u_remapped) {   // alias.h:125-140
    u = u * (float)size;
    uint32_t idx = std::min((uint32_t)u, size - 1);
    u = std::min(u - (float)idx, OneMinusEpsilon);
    vb_alias_entry e = sc.alias_table[base + idx];
    bool keep = u < e.prob;
    u_remapped = keep ? std::min(u / e.prob, OneMinusEpsilon) : std::min((1.0f - u) / (1.0f - e.prob), OneMinusEpsilon);
    return keep ? idx : e.alias;
}
inline float2 env_sample_continuous(const Scene &sc, const vb_light_rec &l, float2 u, float &pdf) {
    uint32_t rx = l.env_res[0], ry = l.env_res[1];
    float ur;
    uint32_t iv = alias_offset_remap(sc, l.alias_offset, ry, u.y, ur);
    float fv = ((float)iv + ur) / (float)ry;
    float pdf_v = l.alias_integral > 0.f ? sc.alias_func[l.alias_offset + iv] / l.alias_integral : 0.f;
    uint32_t off = rx * iv;
    uint32_t iu = alias_offset_remap(sc, l.cond_offset + off, rx, u.x, ur);
    float fu = ((float)iu + ur) / (float)rx;
    float integral_u = sc.alias_func[l.alias_offset + iv];
    float func_u = sc.alias_func[l.cond_offset + off + iu];
    float pdf_u = integral_u > 0.f ? func_u / integral_u : 0.f;
    pdf = pdf_u * pdf_v;
    return {fu, fv};
}

// ---- spherical environment map, reference render_core/light/environments/spherical.cpp:58-160 ----
inline float2 env_calculate_uv(float3 local_dir) {                  // spherical_phi / spherical_theta (pbrt forms)
    float phi = vatan2(local_dir.y, local_dir.x);
    if (phi < 0.f) phi = phi + 2.0f * Pi;
    float cz = std::min(std::max(local_dir.z, -1.f), 1.f);
    float theta = vatan2(safe_sqrt(1.0f - cz * cz), cz);
    return {phi * Inv2Pi, theta * InvPi};
}
inline float3 env_L(const Scene &sc, const Swl &swl, const vb_light_rec &l, float3 local_dir) {
    return spec_illum(sc, swl, eval_slot3(sc, l.color, env_calculate_uv(local_dir))) * l.scale;
}
inline LightEval env_evaluate_wi(const Scene &sc, const Swl &swl, const vb_light_rec &l, float3 ref_pos, float3 light_pos) {
    LightEval ev;
    float4x4 w2o; std::memcpy(&w2o, l.w2o, 64);
    float3 world_dir = normalize(light_pos - ref_pos);
    float3 local_dir = transform_vector(w2o, world_dir);
    float2 uv = env_calculate_uv(local_dir);
    float s, c;
    sincos_2pi(uv.y * 0.5f, s, c);                                  // sin(theta), theta = uv.y * pi
    ev.L = env_L(sc, swl, l, local_dir);
    float pdf = env_map_pdf(sc, l, uv) / (2.0f * Pi * Pi * s);
    ev.pdf = s == 0.f ? 0.f : pdf;
    return ev;
}
inline LightSample env_sample_wi(const Scene &sc, const Swl &swl, const vb_light_rec &l, const Interaction &ref, float2 u) {
    LightSample ls;
    float pdf_map;
    float2 uv = env_sample_continuous(sc, l, u, pdf_map);
    float st, ct, sp, cp;
    sincos_2pi(uv.y * 0.5f, st, ct);
    sincos_2pi(uv.x, sp, cp);
    float3 local_dir{st * cp, st * sp, ct};
    float4x4 o2w; std::memcpy(&o2w, l.o2w, 64);
    float3 world_dir = normalize(transform_vector(o2w, local_dir));
    float pdf_dir = pdf_map / (2.0f * Pi * Pi * st);
    ls.eval.pdf = std::isinf(pdf_dir) ? 0.f : pdf_dir;
    ls.eval.L = env_L(sc, swl, l, local_dir);
    ls.p_light = ref.pos + world_dir * l.params[0];
    return ls;
}

// ---- directional light, reference render_core/light/environments/directional.cpp:46-80: a delta Infinite light ----
inline LightEval directional_evaluate_wi(const Scene &sc, const Swl &swl, const vb_light_rec &l) {
    LightEval ev;
    ev.L = spec_illum(sc, swl, eval_slot3(sc, l.color, float2{0, 0})) * l.scale;
    ev.pdf = -1.f;
    return ev;
}
inline LightSample directional_sample_wi(const Scene &sc, const Swl &swl, const vb_light_rec &l, const Interaction &ref) {
    LightSample ls;
    float3 w_light = -float3{l.direction[0], l.direction[1], l.direction[2]};
    ls.p_light = ref.pos + w_light * l.params[1];
    ls.eval = directional_evaluate_wi(sc, swl, l);
    return ls;
}

// LightSampler::PMF / select_light (lightsampler.cpp:159-197) over PMF_ / select_light_: uniform.cpp:12-33, or power.cpp:16-31 =
// Warper::PMF / sample_discrete of an alias table over luminance(light->power()) (alias.h:148-181). With env_separate and an
// environment light, that light is chosen with probability env_prob and the rule below is applied to the others (uniform: over
// the punctual_light_num() other lights, index corrected past the environment's; power: the host gave the environment weight 0)
inline float light_pmf_(const Scene &sc, uint32_t index) {
    if (sc.lsel_size == 0) {
        uint32_t n = (uint32_t)sc.lights.size();
        if (sc.env_separate) n -= sc.env_light != VB_INVALID_ID ? 1u : 0u;
        return 1.0f / (float)n;
    }
    return sc.lsel_integral > 0.f ? sc.alias_func[sc.lsel_offset + index] / (sc.lsel_integral * (float)sc.lsel_size) : 0.f;
}
inline float light_pmf(const Scene &sc, uint32_t index) {
    if (sc.env_separate && sc.env_light != VB_INVALID_ID)
        return index == sc.env_light ? sc.env_prob : (1 - sc.env_prob) * light_pmf_(sc, index);
    return light_pmf_(sc, index);
}
inline uint32_t light_select(const Scene &sc, float u) {
    uint32_t n = (uint32_t)sc.lights.size();
    const bool separate = sc.env_separate && sc.env_light != VB_INVALID_ID;
    if (separate) {
        if (u < sc.env_prob) return sc.env_light;
        u = (u - sc.env_prob) / (1 - sc.env_prob);                  // remapping, warp.h:17-21
    }
    if (sc.lsel_size == 0) {
        if (separate) {
            float punctual = (float)(n - 1u);
            uint32_t index = (uint32_t)std::min(u * punctual, punctual - 1);
            return index < sc.env_light ? index : index + 1u;       // correct_index, lightsampler.cpp:28-33
        }
        return (uint32_t)std::min(u * (float)n, (float)n - 1.0f);
    }
    u = u * (float)n;
    uint32_t idx = std::min((uint32_t)u, n - 1);
    u = std::min(u - (float)idx, OneMinusEpsilon);
    vb_alias_entry e = sc.alias_table[sc.lsel_offset + idx];
    return u < e.prob ? idx : e.alias;
}

// reference lightsampler.cpp:210-216 + uniform.cpp:24-33 / power.cpp:26-31
inline LightSample sample_light(const Scene &sc, const Swl &swl, const Interaction &it, Sampler &sampler) {
    float u_light = sampler.next_1d();
    float2 u_surface = sampler.next_2d();
    LightSample ls;
    uint32_t n = (uint32_t)sc.lights.size();
    if (n == 0) return ls;
    uint32_t index = light_select(sc, u_light);
    float pmf = light_pmf(sc, index);
    const vb_light_rec &l = sc.lights[index];
    if (l.type == VB_LIGHT_AREA) ls = area_sample_wi(sc, swl, l, it, u_surface);
    else if (l.type == VB_LIGHT_ENV) ls = env_sample_wi(sc, swl, l, it, u_surface);
    else if (l.type == VB_LIGHT_DIRECTIONAL) ls = directional_sample_wi(sc, swl, l, it);
    else ls = point_sample_wi(sc, swl, l, it);
    ls.eval.pdf *= pmf;
    return ls;
}

// reference lightsampler.cpp:252-267
inline LightEval evaluate_hit_wi(const Scene &sc, const Swl &swl, float3 ref_pos, const Interaction &it) {
    LightEval ev;
    const vb_light_rec &l = sc.lights[it.light_id];
    if (l.type != VB_LIGHT_AREA) return ev;
    LightEvalCtx pl{it.pos, it.ng, it.uv, 1.0f / it.prim_area};
    pl.pdf_pos *= alias_pmf(sc, l, it.prim_id);
    ev = area_evaluate_wi(sc, swl, l, ref_pos, pl);
    ev.pdf *= light_pmf(sc, it.light_id);
    return ev;
}

}  // namespace
}  // namespace vo
#include "vo_bsdf.inl"
namespace vo {
namespace {

inline float mis_weight(float f, float g) { return (1.0f * f) / (1.0f * f + 1.0f * g); }   // warp.h:148-153,187-199

// ---- homogeneous medium, reference render_core/medium/homogeneous.cpp:30-70 ----
inline float3 medium_sigma_t(const Scene &sc, const Swl &swl, const vb_medium_rec &m) {
    return spec_unbound(sc, swl, float3{m.sigma_a[0] + m.sigma_s[0], m.sigma_a[1] + m.sigma_s[1], m.sigma_a[2] + m.sigma_s[2]} * m.scale);
}
inline float3 medium_Tr(const Scene &sc, const Swl &swl, const vb_medium_rec &m, float t) {
    float3 st = medium_sigma_t(sc, swl, m);
    float tt = std::min(RayTMax, t);
    return {vexp(-st.x * tt), vexp(-st.y * tt), vexp(-st.z * tt)};
}
// Geometry::Tr (geometry.cpp:187-199) for a ray inside medium `id`
inline float3 ray_Tr(const Scene &sc, const Swl &swl, uint32_t medium, const Ray &ray) {
    if (!sc.params.process_mediums || medium == VB_INVALID_ID) return float3{1, 1, 1};
    return medium_Tr(sc, swl, sc.mediums[medium], length(ray.d) * ray.tmax);
}
// coordinate_system (pbrt v3 form)
inline void coordinate_system(float3 n, float3 &a, float3 &b) {
    if (std::fabs(n.x) > std::fabs(n.y)) a = float3{-n.z, 0.f, n.x} / std::sqrt(n.x * n.x + n.z * n.z);
    else a = float3{0.f, n.z, -n.y} / std::sqrt(n.y * n.y + n.z * n.z);
    b = cross(n, a);
}
inline float phase_HG(float cos_theta, float g) {                       // interaction.h:135-139
    float denom = 1.0f + sqr(g) + 2.0f * g * cos_theta;
    return Inv4Pi * (1.0f - sqr(g)) / (denom * std::sqrt(denom));
}
// HenyeyGreenstein::sample, interaction.cpp:16-32
inline BSDFSample phase_sample(float g, float3 wo, Sampler &sampler) {
    float2 u = sampler.next_2d();
    float sqr_term = (1.0f - sqr(g)) / (1.0f + g - 2.0f * g * u.x);
    float cos_theta = -(1.0f + sqr(g) - sqr(sqr_term)) / (2.0f * g);
    if (std::fabs(g) < 1e-3f) cos_theta = 1.0f - 2.0f * u.x;
    float sin_theta = safe_sqrt(1.0f - sqr(cos_theta));
    float s, c;
    sincos_2pi(u.y, s, c);
    float3 v1, v2;
    coordinate_system(wo, v1, v2);
    BSDFSample bs;
    bs.wi = v1 * (sin_theta * c) + v2 * (sin_theta * s) + wo * cos_theta;
    float f = phase_HG(cos_theta, g);
    bs.eval.f = float3{f, f, f}; bs.eval.pdf = f; bs.eval.flags = 0;
    return bs;
}

}  // namespace

float3 Scene::render_sample(uint32_t px, uint32_t py, uint32_t w, uint32_t h, uint32_t frame, Hit *primary, PathStats *st, float *aov,
                            std::vector<vb_path_vertex> *dump) const {
    Sampler sampler;
    // ---- rt_geom: camera ray, reference frame_buffer.cpp:172-182, sampler.h:65-73 ----
    sampler.start(px, py, frame, 0u);
    float2 uf = sampler.next_2d();
    float2 fp;
    if (camera.filter_type == VB_FILTER_TRIANGLE) {
        auto sample_linear10 = [](float u) {           // sample_linear(u, a=1, b=0), warp.h:123-130
            float a = 1.f, b = 0.f;
            float x = u * (a + b) / (a + std::sqrt((a * a) + u * ((b * b) - (a * a))));
            return std::min(x, OneMinusEpsilon);
        };
        auto tent = [&](float u, float r) {
            return u < 0.5f ? -r * sample_linear10((0.5f - u) * 2.f) : r * sample_linear10((u - 0.5f) * 2.f);
        };
        fp = {tent(uf.x, camera.filter_radius[0]), tent(uf.y, camera.filter_radius[1])};
    } else if (camera.filter_type == VB_FILTER_FITTED) {   // FilterSampler::sample, fitted_curve.h:83-91 (the weight is not used on this path)
        float2 u{uf.x * 2.f - 1.f, uf.y * 2.f - 1.f};
        const uint32_t N = VB_FILTER_TABLE_SIZE, off = camera.filter_table_offset;
        float ur;
        uint32_t iv = alias_offset_remap(*this, off, N, std::fabs(u.y), ur);
        float fv = ((float)iv + ur) / (float)N;
        uint32_t iu = alias_offset_remap(*this, off + N + N * iv, N, std::fabs(u.x), ur);
        float fu = ((float)iu + ur) / (float)N;
        auto sign = [](float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); };
        fp = {fu * sign(u.x) * camera.filter_radius[0], fv * sign(u.y) * camera.filter_radius[1]};
    } else {
        float rx = camera.filter_radius[0], ry = camera.filter_radius[1];
        fp = {-rx + uf.x * (rx - (-rx)), -ry + uf.y * (ry - (-ry))};
    }
    float2 p_film{(float)px + 0.5f + fp.x, (float)py + 0.5f + fp.y};
    float2 u_lens = sampler.next_2d();
    (void)sampler.next_1d();   // time
    // Sensor::generate_ray, sensor.cpp:44-56 (raster -> screen window -> camera space at z = 1)
    float sx = camera.screen_lo[0] + p_film.x / (float)w * (camera.screen_hi[0] - camera.screen_lo[0]);
    float sy = camera.screen_hi[1] - p_film.y / (float)h * (camera.screen_hi[1] - camera.screen_lo[1]);
    float3 dir = normalize(float3{sx * camera.tan_fov_y_over_2, sy * camera.tan_fov_y_over_2, 1.0f});
    float3 org{0, 0, 0};
    {   // thin_lens.cpp:34-42 (lens_radius 0 still draws the lens sample)
        float rr = std::sqrt(u_lens.x), s, c;
        sincos_2pi(u_lens.y, s, c);
        float2 pl{rr * c * camera.lens_radius, rr * s * camera.lens_radius};
        float ft = camera.focal_distance / dir.z;
        float3 p_focus = dir * ft;
        org = {pl.x, pl.y, 0.f};
        dir = normalize(p_focus - org);
    }
    float4x4 c2w; std::memcpy(&c2w, camera.c2w, 64);
    Ray ray{transform_point(c2w, org), 0.f, transform_vector(c2w, dir), RayTMax};
    Hit ph = trace_closest(ray, st);          // rt_geom's own primary trace (frame_buffer.cpp:179)
    if (primary) *primary = ph;
    if (aov) {                                                      // rt_geom's G-buffer outputs, frame_buffer.cpp:184-218
        uint32_t wp = 0;                                            // RenderEnv::initial precedes sampler.start(pixel, frame, 0) (:168-172)
        if (params.spectrum == 1) { Sampler ws; ws.start(px, py, frame, 0xffffffffu); (void)ws.next_1d(); wp = ws.state & 0x00ffffffu; }
        Swl gw = swl_make(*this, wp);
        for (int i = 0; i < 20; ++i) aov[i] = 0.f;
        aov[7] = 1.f; aov[11] = 1.f;                                // albedo / emission are written for every pixel as (0, 0, 0, 1)
        aov[13] = p_film.x; aov[14] = p_film.y;
        if (!ph.is_miss()) {
            Interaction it = compute_surface_interaction(*this, ph.inst, ph.prim, ph.u, ph.v, true);
            it.wo = normalize(-ray.d);
            float3 ns = it.shading.z;
            aov[0] = ns.x; aov[1] = ns.y; aov[2] = ns.z; aov[3] = 1.f;
            // Sensor::linear_depth (sensor.cpp:192-195): z of the camera-space position; c2w is a rigid transform (sensor.cpp:153-162),
            // so that is the projection of (pos - camera position) on c2w's third column
            float3 fwd{c2w.m[2][0], c2w.m[2][1], c2w.m[2][2]}, cpos{c2w.m[3][0], c2w.m[3][1], c2w.m[3][2]};
            aov[12] = dot(fwd, it.pos - cpos);
            aov[15] = 1.f;
            {   // FrameBuffer::compute_motion_vec (frame_buffer.cpp:483-491) = p_film - Sensor::prev_raster_coord(pos).xy (sensor.cpp:95-100):
                // world -> previous sensor space, divide by z, sensor -> raster (the analytic camera's screen window map inverted)
                vb_camera_prev pc = have_camera_prev ? camera_prev : camera_prev_of(camera);
                float4x4 w2c; std::memcpy(&w2c, pc.w2c, 64);
                float3 q = transform_point(w2c, it.pos);
                float sx = (q.x / q.z) / pc.tan_fov_y_over_2, sy = (q.y / q.z) / pc.tan_fov_y_over_2;
                float rx = (sx - pc.screen_lo[0]) / (pc.screen_hi[0] - pc.screen_lo[0]) * (float)w;
                float ry = (pc.screen_hi[1] - sy) / (pc.screen_hi[1] - pc.screen_lo[1]) * (float)h;
                aov[16] = p_film.x - rx; aov[17] = p_film.y - ry;
            }
            if (it.mat_id != VB_INVALID_ID) {
                float cos_theta = dot(it.shading.z, it.wo);          // MaterialEvaluator::albedo, material.cpp:91-98
                const uint32_t mt = materials[it.mat_id].type;
                float3 a = (mt == VB_MAT_MIX || mt == VB_MAT_ADD) ? mixset_albedo(*this, mixset_create(*this, gw, it), cos_theta)
                           : mt == VB_MAT_PRINCIPLED ? principled_albedo(*this, principled_create(*this, gw, it), cos_theta)
                                                                          : lobe_albedo(create_lobe(*this, gw, it), cos_theta);
                a = spec_to_srgb(*this, gw, a);
                aov[4] = a.x; aov[5] = a.y; aov[6] = a.z;
            }
            if (it.light_id != VB_INVALID_ID) {
                LightEval ev = evaluate_hit_wi(*this, gw, ray.o, it);
                float3 e = spec_to_srgb(*this, gw, ev.L);
                aov[8] = e.x; aov[9] = e.y; aov[10] = e.z;
            }
        }
        return float3{0, 0, 0};
    }

    // ---- path_tracing kernel: reference integrator.cpp:85-99 ----
    uint32_t swl_packed = 0;                                        // RenderEnv::initial, integrator.cpp:48-57: dim -1, state restored after
    if (params.spectrum == 1) { sampler.start(px, py, frame, 0xffffffffu); (void)sampler.next_1d(); swl_packed = sampler.state & 0x00ffffffu; }
    Swl swl = swl_make(*this, swl_packed);
    sampler.start(px, py, frame, 1u);
    const bool mediums = params.process_mediums != 0;
    uint32_t ray_medium = mediums ? camera.medium_id : VB_INVALID_ID;          // RayData::dir_medium.w
    float scatter_pdf = 1e16f;
    float3 throughput{1, 1, 1}, ret{0, 0, 0};
    float3 prev_surface_ng = ray.d;
    const float3 primary_dir = ray.d;
    float eta_scale = 1.f;
    const uint32_t max_depth = params.max_depth;
    auto put = [&](const Ray &r, uint32_t inst, uint32_t prim, float u, float v, float t, uint32_t bounces) {
        if (!dump) return;
        vb_path_vertex o{};
        o.ray_o[0] = r.o.x; o.ray_o[1] = r.o.y; o.ray_o[2] = r.o.z; o.ray_tmin = r.tmin;
        o.ray_d[0] = r.d.x; o.ray_d[1] = r.d.y; o.ray_d[2] = r.d.z; o.ray_tmax = r.tmax;
        o.inst = inst; o.prim = prim; o.u = u; o.v = v; o.t = t;
        o.bounces = bounces; o.sampler_state = sampler.state; o.medium = ray_medium;
        o.throughput[0] = throughput.x; o.throughput[1] = throughput.y; o.throughput[2] = throughput.z; o.scatter_pdf = scatter_pdf;
        o.radiance[0] = ret.x; o.radiance[1] = ret.y; o.radiance[2] = ret.z; o.eta_scale = eta_scale;
        dump->push_back(o);
    };
    for (uint32_t bounces = 0; bounces < max_depth; ++bounces) {
        Hit hit = trace_closest(ray, st);
        put(ray, hit.inst, hit.prim, hit.u, hit.v, hit.t, bounces);
        if (hit.is_miss()) {                                        // evaluate_miss, integrator.cpp:137-158 + lightsampler.cpp:289-300
            if (env_light != VB_INVALID_ID) {
                const vb_light_rec &l = lights[env_light];
                float3 tr{1, 1, 1};
                if (mediums) { ray.tmax = l.params[0]; tr = ray_Tr(*this, swl, ray_medium, ray); }
                LightEval ev = l.type == VB_LIGHT_DIRECTIONAL ? directional_evaluate_wi(*this, swl, l) : env_evaluate_wi(*this, swl, l, ray.o, ray.o + ray.d);
                ev.pdf *= light_pmf(*this, env_light);
                float weight = mis_weight(scatter_pdf, ev.pdf);
                if (params.mis_mode == 2) weight = 1.f;
                else if (params.mis_mode == 1) weight = bounces == 0 ? weight : 0.f;
                ret += spec_to_srgb(*this, swl, ev.L * tr * weight * throughput);
            }
            break;
        }
        Interaction it = compute_surface_interaction(*this, hit.inst, hit.prim, hit.u, hit.v, true);
        it.wo = normalize(-ray.d);
        ray.tmax = length(it.pos - ray.o) / length(ray.d);         // geometry.h:64-69
        if (st) st->vertices++;
        if (mediums && ray_medium != VB_INVALID_ID) {              // integrator.cpp:199-206, homogeneous.cpp:50-70
            const vb_medium_rec &md = this->mediums[ray_medium];
            float3 sigma_t = medium_sigma_t(*this, swl, md);
            float3 sigma_s = spec_unbound(*this, swl, float3{md.sigma_s[0] * md.scale, md.sigma_s[1] * md.scale, md.sigma_s[2] * md.scale});
            uint32_t channel = std::min((uint32_t)(sampler.next_1d() * 3.0f), 2u);
            float sc_t = channel == 0 ? sigma_t.x : (channel == 1 ? sigma_t.y : sigma_t.z);
            float dist = -vlog(1.0f - sampler.next_1d()) / sc_t;
            float t = std::min(dist / length(ray.d), ray.tmax);
            bool sampled_medium = t < ray.tmax;
            if (sampled_medium) {
                Interaction mi;
                mi.pos = ray.o + ray.d * t;
                mi.wo = -ray.d;
                mi.ng = float3{0, 0, 0};
                mi.uv = float2{0, 0};
                mi.shading = Frame{float3{0, 0, 0}, float3{0, 0, 0}, float3{0, 0, 0}};
                mi.has_phase = true; mi.g = md.g;
                mi.inside = ray_medium; mi.outside = ray_medium;
                it = mi;
            }
            float3 tr = medium_Tr(*this, swl, md, t);
            float3 density = sampled_medium ? sigma_t * tr : tr;
            float pdf = average3(density);
            float3 mt = sampled_medium ? tr * sigma_s / pdf : tr / pdf;
            throughput *= mt;
        }
        if (it.mat_id == VB_INVALID_ID && !it.has_phase) {         // integrator.cpp:208-214
            ray = spawn_ray(it, ray.d);
            if (mediums) ray_medium = dot(it.ng, ray.d) > 0.f ? it.outside : it.inside;
            bounces -= 1;                                           // wraps like the reference's Uint
            continue;
        }
        if (it.light_id != VB_INVALID_ID) {                        // integrator.cpp:221-231
            LightEval ev = evaluate_hit_wi(*this, swl, ray.o, it);
            float3 tr = ray_Tr(*this, swl, ray_medium, ray);
            float weight = mis_weight(scatter_pdf, ev.pdf);
            if (params.mis_mode == 2) weight = 1.f;
            else if (params.mis_mode == 1) weight = bounces == 0 ? weight : 0.f;
            ret += spec_to_srgb(*this, swl, ev.L * throughput * weight * tr);
        }
        prev_surface_ng = it.ng;
        // light sampling + shadow ray, integrator.cpp:241-244
        LightSample ls = sample_light(*this, swl, it, sampler);
        Ray shadow = spawn_ray_to(it, ls.p_light);
        uint32_t shadow_medium = mediums ? (dot(it.ng, shadow.d) > 0.f ? it.outside : it.inside) : VB_INVALID_ID;
        bool occluded = trace_any(shadow, st);
        float3 tr = ray_Tr(*this, swl, shadow_medium, shadow);
        // material / phase function, integrator.cpp:246-283 + direct_lighting :20-37
        float3 wi = normalize(ls.p_light - it.pos);
        ScatterEval se;
        BSDFSample bs;
        if (mediums && it.has_phase) {
            float f = phase_HG(dot(it.wo, wi), it.g);
            se.f = float3{f, f, f}; se.pdf = f; se.flags = 0;
            bs = phase_sample(it.g, it.wo, sampler);
        } else if (materials[it.mat_id].type == VB_MAT_MIX || materials[it.mat_id].type == VB_MAT_ADD) {
            MixEvaluator evaluator{mixset_create(*this, swl, it), it.ng};
            se = evaluator.evaluate(*this, it.wo, wi);
            bs = evaluator.sample(*this, it.wo, sampler);
        } else if (materials[it.mat_id].type == VB_MAT_PRINCIPLED) {
            PrincipledEvaluator evaluator{principled_create(*this, swl, it), it.ng};
            se = evaluator.evaluate(*this, it.wo, wi);
            bs = evaluator.sample(*this, it.wo, sampler);
        } else {
            Evaluator evaluator{create_lobe(*this, swl, it), it.ng};
            if (evaluator.lobe.dispersive && swl.hero) {            // SampledWavelengths::check_dispersive, integrator.cpp:264 + spectrum.cpp:21-38
                swl_packed |= SwlInvalidBit;
                swl.pdf[1] = 0.f; swl.pdf[2] = 0.f;
            }
            se = evaluator.evaluate(*this, it.wo, wi);
            bs = evaluator.sample(*this, it.wo, sampler);
        }
        bool is_delta = ls.eval.pdf < 0.f;
        float weight = is_delta ? 1.f : mis_weight(ls.eval.pdf, se.pdf);
        if (params.mis_mode == 1) weight = 1.f;
        float lpdf = is_delta ? -ls.eval.pdf : ls.eval.pdf;
        float3 Ld{0, 0, 0};
        if (!occluded && se.pdf > 0.f && lpdf > 0.f) Ld = ls.eval.L * se.f * weight / lpdf;
        if (params.mis_mode == 2) Ld = Ld * 0.f;
        ret += spec_to_srgb(*this, swl, throughput * Ld * tr);
        eta_scale *= sqr(bs.eta);
        float lum = max_comp(throughput);
        if (!(bs.eval.pdf > 0.f) || lum == 0.f) break;
        throughput *= bs.eval.f / bs.eval.pdf;
        if (eta_scale * lum < params.rr_threshold && bounces >= params.min_depth) {
            float q = std::min(0.95f, lum);
            float rr = sampler.next_1d();
            if (q < rr) break;
            throughput = throughput / q;
        }
        scatter_pdf = bs.eval.pdf;
        ray = spawn_ray(it, bs.wi);
        if (mediums) ray_medium = dot(it.ng, bs.wi) > 0.f ? it.outside : it.inside;
    }
    // "Supplement only direct light BSDF sampling" (integrator.cpp:176-233,303-309): with max_depth < 2 and MIS mode "both", ONE more
    // mis_bsdf(inner = false): the last spawned ray is traced and emission / environment gathered with the BSDF-side MIS weight
    if (max_depth < 2 && params.mis_mode == 0) {
        for (uint32_t bounce = 0; bounce < 1u; ++bounce) {
            Hit hit = trace_closest(ray, st);
            if (ray.d.x == primary_dir.x && ray.d.y == primary_dir.y && ray.d.z == primary_dir.z) break;      // primary_miss: nothing was scattered
            put(ray, hit.inst, hit.prim, hit.u, hit.v, hit.t, max_depth);
            if (hit.is_miss()) {
                if (env_light != VB_INVALID_ID) {
                    const vb_light_rec &l = lights[env_light];
                    float3 tr{1, 1, 1};
                    if (mediums) { ray.tmax = l.params[0]; tr = ray_Tr(*this, swl, ray_medium, ray); }
                    LightEval ev = l.type == VB_LIGHT_DIRECTIONAL ? directional_evaluate_wi(*this, swl, l) : env_evaluate_wi(*this, swl, l, ray.o, ray.o + ray.d);
                    ev.pdf *= light_pmf(*this, env_light);
                    float weight = mis_weight(scatter_pdf, ev.pdf);
                    ret += spec_to_srgb(*this, swl, ev.L * tr * weight * throughput);
                }
                break;
            }
            Interaction it = compute_surface_interaction(*this, hit.inst, hit.prim, hit.u, hit.v, true);
            it.wo = normalize(-ray.d);
            ray.tmax = length(it.pos - ray.o) / length(ray.d);
            if (st) st->vertices++;
            if (mediums && ray_medium != VB_INVALID_ID) {
                const vb_medium_rec &md = this->mediums[ray_medium];
                float3 sigma_t = medium_sigma_t(*this, swl, md);
                float3 sigma_s = spec_unbound(*this, swl, float3{md.sigma_s[0] * md.scale, md.sigma_s[1] * md.scale, md.sigma_s[2] * md.scale});
                uint32_t channel = std::min((uint32_t)(sampler.next_1d() * 3.0f), 2u);
                float sc_t = channel == 0 ? sigma_t.x : (channel == 1 ? sigma_t.y : sigma_t.z);
                float dist = -vlog(1.0f - sampler.next_1d()) / sc_t;
                float t = std::min(dist / length(ray.d), ray.tmax);
                bool sampled_medium = t < ray.tmax;
                if (sampled_medium) { Interaction mi; mi.has_phase = true; mi.inside = ray_medium; mi.outside = ray_medium; it = mi; }
                float3 tr = medium_Tr(*this, swl, md, t);
                float3 density = sampled_medium ? sigma_t * tr : tr;
                float pdf = average3(density);
                float3 mt = sampled_medium ? tr * sigma_s / pdf : tr / pdf;
                throughput *= mt;
            }
            if (it.mat_id == VB_INVALID_ID && !it.has_phase) {
                ray = spawn_ray(it, ray.d);
                if (mediums) ray_medium = dot(it.ng, ray.d) > 0.f ? it.outside : it.inside;
                bounce -= 1;
                continue;
            }
            if (it.light_id != VB_INVALID_ID) {
                LightEval ev = evaluate_hit_wi(*this, swl, ray.o, it);
                float3 tr = ray_Tr(*this, swl, ray_medium, ray);
                float weight = mis_weight(scatter_pdf, ev.pdf);
                ret += spec_to_srgb(*this, swl, ev.L * throughput * weight * tr);
            }
        }
    }
    (void)prev_surface_ng;
    if (dump) {                                                     // the end record: final radiance + sampler state only
        vb_path_vertex o{};
        o.inst = VB_PATH_END; o.prim = VB_PATH_END; o.sampler_state = sampler.state;
        o.radiance[0] = ret.x; o.radiance[1] = ret.y; o.radiance[2] = ret.z;
        dump->push_back(o);
    }
    return ret;
}

}  // namespace vo

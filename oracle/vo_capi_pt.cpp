// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
// extern "C" surface of the oracle path tracer for ctypes.
#include "vo_pt.h"
#include <thread>
#include <atomic>
#include <chrono>

using namespace vo;

extern "C" {

void *vo_scene_create() { return new Scene(); }
void vo_scene_destroy(void *h) { delete (Scene *)h; }

uint32_t vo_scene_add_mesh(void *h, const float *pos3, const float *nrm3, const float *uv2, uint32_t nv,
                           const uint32_t *tri3, uint32_t nt) {
    Scene &s = *(Scene *)h;
    MeshData m;
    m.pos.resize(nv); m.nrm.assign(nv, float3{0, 0, 0}); m.uv.assign(nv, float2{0, 0});
    for (uint32_t i = 0; i < nv; ++i) {
        m.pos[i] = {pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]};
        if (nrm3) m.nrm[i] = {nrm3[3 * i], nrm3[3 * i + 1], nrm3[3 * i + 2]};
        if (uv2) m.uv[i] = {uv2[2 * i], uv2[2 * i + 1]};
    }
    m.tri.assign(tri3, tri3 + 3 * (size_t)nt);
    s.meshes.push_back(std::move(m));
    return (uint32_t)s.meshes.size() - 1;
}
void vo_scene_set_instances(void *h, const vb_instance *inst, uint32_t n) { ((Scene *)h)->set_instances(inst, n); }
void vo_scene_set_materials(void *h, const vb_material_rec *m, uint32_t n) { ((Scene *)h)->materials.assign(m, m + n); }
void vo_scene_set_lights(void *h, const vb_light_rec *l, uint32_t n, const vb_alias_entry *t, const float *f, uint32_t na) {
    Scene &s = *(Scene *)h;
    s.lights.assign(l, l + n);
    s.alias_table.assign(t, t + na);
    s.alias_func.assign(f, f + na);
    s.env_light = VB_INVALID_ID;
    for (uint32_t i = 0; i < n; ++i) if (l[i].type == VB_LIGHT_ENV || l[i].type == VB_LIGHT_DIRECTIONAL) s.env_light = i;
    s.env_separate = 0; s.env_prob = 0.f;
}
void vo_scene_set_light_sampler_env(void *h, uint32_t env_separate, float env_prob) {
    Scene &s = *(Scene *)h;
    s.env_separate = env_separate; s.env_prob = env_prob;
}
void vo_scene_set_light_sampler(void *h, uint32_t kind, uint32_t offset, uint32_t size, float integral) {
    Scene &s = *(Scene *)h;
    if (kind == VB_LIGHT_SAMPLER_POWER) { s.lsel_offset = offset; s.lsel_size = size; s.lsel_integral = integral; }
    else { s.lsel_offset = s.lsel_size = 0; s.lsel_integral = 0.f; }
}
void vo_scene_set_textures(void *h, const vb_texture *t, uint32_t n) {
    Scene &s = *(Scene *)h;
    s.textures.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        s.textures[i].w = t[i].width; s.textures[i].h = t[i].height;
        const float4 *p = (const float4 *)t[i].texels;
        s.textures[i].texels.assign(p, p + (size_t)t[i].width * t[i].height);
    }
}
void vo_scene_set_lut(void *h, int which, const float *data, uint32_t nx, uint32_t ny, uint32_t nz, uint32_t ch) {
    Lut &l = ((Scene *)h)->luts[which];
    l.nx = nx; l.ny = ny; l.nz = nz; l.channels = ch;
    l.data.assign(data, data + (size_t)nx * ny * nz * ch);
}
void vo_scene_set_mediums(void *h, const vb_medium_rec *m, uint32_t n) { ((Scene *)h)->mediums.assign(m, m + n); }
void vo_scene_set_spds(void *h, const float *values, const uint32_t *offsets, uint32_t n) {
    Scene &s = *(Scene *)h;
    s.spd_offsets.assign(offsets, offsets + n + 1);
    s.spd_values.assign(values, values + offsets[n]);
}
void vo_scene_set_spectrum(void *h, const vb_spectrum_rec *r) {
    Scene &s = *(Scene *)h;
    s.spectrum = *r;
    s.rgb2spec_res = r->rgb2spec_res;
    if (r->rgb2spec) s.rgb2spec.assign(r->rgb2spec, r->rgb2spec + (size_t)3 * r->rgb2spec_res * r->rgb2spec_res * r->rgb2spec_res * 4);
    s.spectrum.rgb2spec = nullptr;
}
void vo_scene_set_camera(void *h, const vb_camera *c) { ((Scene *)h)->camera = *c; }
void vo_scene_set_camera_prev(void *h, const vb_camera_prev *c) { Scene *s = (Scene *)h; s->have_camera_prev = c != nullptr; if (c) s->camera_prev = *c; }
void vo_scene_set_integrator(void *h, const vb_pt_params *p) { ((Scene *)h)->params = *p; }
void vo_scene_build(void *h, int threads, int brute) {
    Scene &s = *(Scene *)h;
    s.brute = brute != 0;
    s.build(threads);
}

// Renders frames first, first+stride, ... into accum (w*h float4).
// mode 0 (LERP): acc = lerp(1/(k+1), acc, (L,1)), k = frames_done + i  (reference frame_buffer.cpp:117-126)
// mode 1 (SUM):  acc += (L,1)
// stats5 (nullable): closest rays, any rays, shaded vertices, nodes, tris. Returns seconds.
double vo_render(void *h, uint32_t w, uint32_t hgt, uint32_t first, uint32_t n_frames, uint32_t stride, int mode,
                 uint32_t frames_done, float *accum, int threads, uint64_t *stats5) {
    const Scene &s = *(Scene *)h;
    threads = std::max(1, threads);
    std::vector<PathStats> st(threads);
    auto t0 = std::chrono::steady_clock::now();
    std::atomic<uint32_t> next_row{0};
    auto work = [&](int tid) {
        for (;;) {
            uint32_t y = next_row.fetch_add(1);
            if (y >= hgt) break;
            for (uint32_t x = 0; x < w; ++x) {
                float *a = accum + 4 * ((size_t)y * w + x);
                for (uint32_t i = 0; i < n_frames; ++i) {
                    uint32_t frame = first + i * stride;
                    float3 L = s.render_sample(x, y, w, hgt, frame, nullptr, stats5 ? &st[tid] : nullptr);
                    float v[4] = {L.x, L.y, L.z, 1.0f};
                    if (mode == 0) {
                        float t = 1.0f / (float)(frames_done + i + 1);
                        for (int c = 0; c < 4; ++c) a[c] = a[c] + t * (v[c] - a[c]);
                    } else {
                        for (int c = 0; c < 4; ++c) a[c] += v[c];
                    }
                }
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (stats5) {
        for (int i = 0; i < 5; ++i) stats5[i] = 0;
        for (auto &p : st) { stats5[0] += p.closest; stats5[1] += p.any; stats5[2] += p.vertices; stats5[3] += p.nodes; stats5[4] += p.tris; }
    }
    return dt;
}

// primary hits for one frame index (== reference rt_compute_pixel / the trace inside rt_geom)
void vo_primary_hits(void *h, uint32_t w, uint32_t hgt, uint32_t frame, uint32_t *inst, uint32_t *prim, float *u, float *v,
                     float *t, int threads) {
    const Scene &s = *(Scene *)h;
    // only the camera ray + first trace are needed: run with max_depth 0 on a copy of the params
    Scene *ms = (Scene *)h;
    uint32_t saved = ms->params.max_depth;
    ms->params.max_depth = 0;
    std::atomic<uint32_t> next_row{0};
    auto work = [&]() {
        for (;;) {
            uint32_t y = next_row.fetch_add(1);
            if (y >= hgt) break;
            for (uint32_t x = 0; x < w; ++x) {
                Hit ph;
                s.render_sample(x, y, w, hgt, frame, &ph, nullptr);
                size_t i = (size_t)y * w + x;
                inst[i] = ph.inst; prim[i] = ph.prim; u[i] = ph.u; v[i] = ph.v; t[i] = ph.t;
            }
        }
    };
    std::vector<std::thread> th;
    for (int k = 1; k < std::max(1, threads); ++k) th.emplace_back(work);
    work();
    for (auto &x : th) x.join();
    ms->params.max_depth = saved;
}

// rt_geom's G-buffer planes for one frame index: out = 5 planes of w*h float4 (normal, albedo, emission, geom, motion)
void vo_gbuffer(void *h, uint32_t w, uint32_t hgt, uint32_t frame, float *out, int threads) {
    const Scene &s = *(Scene *)h;
    std::atomic<uint32_t> next_row{0};
    auto work = [&]() {
        for (;;) {
            uint32_t y = next_row.fetch_add(1);
            if (y >= hgt) break;
            for (uint32_t x = 0; x < w; ++x) {
                float a[20];
                s.render_sample(x, y, w, hgt, frame, nullptr, nullptr, a);
                size_t i = (size_t)y * w + x, n = (size_t)w * hgt;
                for (int p = 0; p < 5; ++p) for (int c = 0; c < 4; ++c) out[(p * n + i) * 4 + c] = a[4 * p + c];
            }
        }
    };
    std::vector<std::thread> th;
    for (int k = 1; k < std::max(1, threads); ++k) th.emplace_back(work);
    work();
    for (auto &x : th) x.join();
}

// per-vertex dump of one path (the record layout of vb_dump_path_trace): returns the number of records, writes at most `max`
uint32_t vo_dump_path_trace(void *h, uint32_t w, uint32_t hgt, uint32_t px, uint32_t py, uint32_t frame, vb_path_vertex *out, uint32_t max) {
    const Scene &s = *(Scene *)h;
    std::vector<vb_path_vertex> v;
    s.render_sample(px, py, w, hgt, frame, nullptr, nullptr, nullptr, &v);
    for (uint32_t i = 0; i < v.size() && i < max; ++i) out[i] = v[i];
    return (uint32_t)v.size();
}

// rt_gradient (FrameBuffer::compute_gradient, frame_buffer.cpp:221-275) over the NORMAL and GEOM planes: out = w*h float4
// (normal_fwidth, depth_gradient, 0, 0). Loop order, clipping and the divisions by the sample counts as written there.
void vo_gbuffer_gradient(const float *normal, const float *geom, uint32_t w, uint32_t hgt, float *out) {
    const int W = (int)w, H = (int)hgt;
    for (int cy = 0; cy < H; ++cy)
        for (int cx = 0; cx < W; ++cx) {
            const float *cnp = normal + 4 * ((size_t)cy * W + cx);
            const float3 cn{cnp[0], cnp[1], cnp[2]};
            const float cd = geom[4 * ((size_t)cy * W + cx)];
            uint32_t xn = 0, yn = 0;
            float3 ndx{0, 0, 0}, ndy{0, 0, 0};
            float ddx = 0.f, ddy = 0.f;
            for (int x = std::max(0, cx - 1); x <= std::min(cx + 1, W - 1); ++x)
                for (int y = std::max(0, cy - 1); y <= std::min(cy + 1, H - 1); ++y) {
                    const float *np = normal + 4 * ((size_t)y * W + x);
                    const float3 nn{np[0], np[1], np[2]};
                    const float nd = geom[4 * ((size_t)y * W + x)];
                    if (cx > x) { xn++; ndx = ndx + (cn - nn); ddx += cd - nd; }
                    else if (x > cx) { xn++; ndx = ndx + (nn - cn); ddx += nd - cd; }
                    if (cy > y) { yn++; ndy = ndy + (cn - nn); ddy += cd - nd; }
                    else if (y > cy) { yn++; ndy = ndy + (nn - cn); ddy += nd - cd; }
                }
            const float fx = (float)xn, fy = (float)yn;
            ndx = float3{ndx.x / fx, ndx.y / fx, ndx.z / fx}; ndy = float3{ndy.x / fy, ndy.y / fy, ndy.z / fy};
            const float3 nf{std::fabs(ndx.x) + std::fabs(ndy.x), std::fabs(ndx.y) + std::fabs(ndy.y), std::fabs(ndx.z) + std::fabs(ndy.z)};
            ddx = ddx / fx; ddy = ddy / fy;
            float *o = out + 4 * ((size_t)cy * W + cx);
            o[0] = length(nf); o[1] = std::fabs(ddx) + std::fabs(ddy); o[2] = 0.f; o[3] = 0.f;
        }
}

// render-graph passes of the `customized` pipeline over n float4 pixels: accumulate (pass/accumulate.cpp:19-24), tonemapping
// (pass/tonemapping.cpp:26-33 with tonemapper/impl.cpp:21-46), gamma (pass/gamma.cpp:21-27; linear_to_srgb = the sRGB OETF on [0, 1],
// its power through the shared vlog / vexp)
void vo_pass_accumulate(const float *input, float *output, uint64_t n, uint32_t frame_index) {
    const float t = 1.f / ((float)frame_index + 1.f);
    for (uint64_t i = 0; i < n; ++i) {
        for (int c = 0; c < 3; ++c) output[4 * i + c] = output[4 * i + c] + t * (input[4 * i + c] - output[4 * i + c]);   // ocarina lerp(t, a, b) = a + t * (b - a)
        output[4 * i + 3] = 1.f;
    }
}
void vo_pass_tonemapping(const float *input, float *output, uint64_t n, int tonemap) {
    for (uint64_t i = 0; i < n; ++i) {
        for (int c = 0; c < 3; ++c) {
            float x = input[4 * i + c];
            if (tonemap == VB_TONEMAP_ACES) {
                float y = (x * (2.51f * x + 0.03f)) / (x * (2.43f * x + 0.59f) + 0.14f);
                x = std::min(1.0f, std::max(0.0f, y));
            } else if (tonemap == VB_TONEMAP_REINHARD) x = x / (x + 1.0f);
            output[4 * i + c] = x;
        }
        output[4 * i + 3] = 1.f;
    }
}
void vo_pass_gamma(const float *input, float *output, uint64_t n) {
    for (uint64_t i = 0; i < 4 * n; ++i) {
        float x = std::min(std::max(input[i], 0.f), 1.f);
        output[i] = x <= 0.0031308f ? 12.92f * x : 1.055f * vexp(vlog(x) * (1.0f / 2.4f)) - 0.055f;
    }
}

// known-answer helpers
uint32_t vo_tea4(uint32_t a, uint32_t b) { return tea4(a, b); }
float vo_lcg(uint32_t *state) { return lcg(*state); }
void vo_sincos_2pi(float u, float *s, float *c) { sincos_2pi(u, *s, *c); }
float vo_vexp(float x) { return vexp(x); }
float vo_vlog(float x) { return vlog(x); }
float vo_vatan2(float y, float x) { return vatan2(y, x); }
void vo_offset_ray_origin(const float *p, const float *n, float *out) {
    float3 r = offset_ray_origin({p[0], p[1], p[2]}, {n[0], n[1], n[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}


// hero spectrum probes (vo_pt.cpp): out = lambda[3], pdf[3]
void vo_hero_wavelengths(uint32_t u24, float *out);
// kind 0 albedo, 1 unbound, 2 illumination; then back to linear sRGB: out_sp[3], out_rgb[3]
void vo_hero_decode(void *h, uint32_t u24, int kind, const float *rgb, float *out_sp, float *out_rgb);

}  // extern "C"

// Megakernel, film kernels and the PT part of the C ABI. Included by vb_pt.cu only.
#pragma once

namespace {

struct PtCounters { unsigned long long closest, any, vertices, culled; };

// Megakernel schedule. One thread = one pixel; it runs all frames of the batch for that pixel back to
// back as a state machine: every loop iteration is [regenerate a finished path] -> closest-hit
// traversal -> pt_bounce -> any-hit traversal, so the lanes of a warp meet at ONE closest and ONE
// any-hit call site whatever bounce or frame each of them is at, and a lane whose path ended starts
// its next frame in the same iteration instead of waiting for the longest path of the warp. The
// running film value stays in registers and is written once per launch.
template <bool DUMP_PRIMARY, int VARIANT>
__global__ void __launch_bounds__(PT_BLOCK) k_path_trace(const __grid_constant__ PtScene S, uint32_t first_frame, uint32_t n_frames, uint32_t stride, int mode,
                                                         uint32_t frames_done, float4 *rt, float4 *accum, PtCounters *counters,
                                                         uint32_t *h_inst, uint32_t *h_prim, float *h_u, float *h_v, float *h_t) {
    __shared__ uint2 s_stack_all[VB_SM_STACK * PT_BLOCK];
    uint2 *s_stack = s_stack_all + threadIdx.x;
    const VbTrav mg = vb_trav_init();
    uint32_t pix = blockIdx.x * PT_BLOCK + threadIdx.x;
    uint32_t n_pix = S.width * S.height;
    uint32_t c_closest = 0, c_any = 0, c_vert = 0, c_culled = 0, dummy0 = 0, dummy1 = 0;
    const uint32_t px = pix % S.width, py = pix / S.width;
    if (DUMP_PRIMARY) {
        if (pix < n_pix) {
            Sampler sampler;
            sampler.start(px, py, first_frame, 0u);
            VbRay ray = generate_camera_ray(S, sampler, px, py);
            VbHit hit;
            vb_traverse<false, false>(S.accel, ray, hit, s_stack, PT_BLOCK, dummy0, dummy1, mg);
            h_inst[pix] = hit.inst; h_prim[pix] = hit.prim; h_u[pix] = hit.u; h_v[pix] = hit.v; h_t[pix] = hit.t;
        }
        return;
    }
    bool done = pix >= n_pix, need_new = true;
    float4 acc = done ? make_float4(0.f, 0.f, 0.f, 0.f) : accum[pix];
    float4 last = make_float4(0.f, 0.f, 0.f, 1.f);
    uint32_t fi = 0;
    PathState ps;
    VbRay ray;
    const uint32_t max_depth = S.params.max_depth;
    for (;;) {
        if (!done && need_new) {
            if (fi == n_frames) done = true;
            else {
                // the reference traces the primary ray in rt_geom and again as bounce 0 of Li with the
                // same ray; the second result is identical, so it is traced (and counted) once
                uint32_t frame = first_frame + fi * stride;
                Sampler sampler;
                sampler.start(px, py, frame, 0u);
                ray = generate_camera_ray(S, sampler, px, py);
                path_begin(S, ps, px, py, frame);
                need_new = max_depth == 0;
                if (need_new) {                                                      // no bounce: a black sample (L, 1) = (0, 0, 0, 1)
                    last = make_float4(0.f, 0.f, 0.f, 1.f);
                    if (mode == VB_ACCUM_LERP) {
                        float t = 1.0f / (float)(frames_done + fi + 1);
                        acc.x = acc.x + t * (last.x - acc.x); acc.y = acc.y + t * (last.y - acc.y);
                        acc.z = acc.z + t * (last.z - acc.z); acc.w = acc.w + t * (last.w - acc.w);
                    } else acc.w += 1.f;
                    fi++;
                }
            }
        }
        if (__all_sync(~0u, done)) break;
        bool run = !done && !need_new;
        VbHit hit;
        if (run) { vb_traverse<false, false>(S.accel, ray, hit, s_stack, PT_BLOCK, dummy0, dummy1, mg); c_closest++; }
        int flags = 0;
        VbRay shadow;
        f3 pending{0, 0, 0};
        if (run) flags = pt_bounce<VARIANT>(S, ps, ray, hit, shadow, pending);
        if (flags & PT_VERTEX) c_vert++;
        if ((flags & PT_LIGHT) && !(flags & PT_SHADOW)) c_culled++;
        if (flags & PT_SHADOW) {
            VbHit sh;
            bool occluded = vb_traverse<true, false>(S.accel, shadow, sh, s_stack, PT_BLOCK, dummy0, dummy1, mg);
            c_any++;
            if (!occluded) ps.ret = ps.ret + pending;
        }
        if (run && !(flags & PT_CONTINUE)) {
            last = make_float4(ps.ret.x, ps.ret.y, ps.ret.z, 1.f);
            if (mode == VB_ACCUM_LERP) {
                float t = 1.0f / (float)(frames_done + fi + 1);
                acc.x = acc.x + t * (last.x - acc.x); acc.y = acc.y + t * (last.y - acc.y);
                acc.z = acc.z + t * (last.z - acc.z); acc.w = acc.w + t * (last.w - acc.w);
            } else {
                acc.x += last.x; acc.y += last.y; acc.z += last.z; acc.w += last.w;
            }
            fi++;
            need_new = true;
        }
    }
    if (pix < n_pix) { accum[pix] = acc; rt[pix] = last; }
    for (int o = 16; o > 0; o >>= 1) {
        c_closest += __shfl_xor_sync(~0u, c_closest, o); c_any += __shfl_xor_sync(~0u, c_any, o); c_vert += __shfl_xor_sync(~0u, c_vert, o);
        c_culled += __shfl_xor_sync(~0u, c_culled, o);
    }
    if ((threadIdx.x & 31) == 0 && counters) {
        atomicAdd(&counters->closest, (unsigned long long)c_closest); atomicAdd(&counters->any, (unsigned long long)c_any);
        atomicAdd(&counters->vertices, (unsigned long long)c_vert); atomicAdd(&counters->culled, (unsigned long long)c_culled);
    }
}

// rt_geom's G-buffer outputs for one frame index (reference frame_buffer.cpp:156-219): shading normal, material albedo, emission of
// the primary hit (all through Spectrum::linear_srgb), linear depth + film position. Planes: out[plane * n_pix + pix] (float4);
// normal: (ns, 1) on a hit, zero otherwise (the reference leaves a missed pixel's normal unwritten); albedo / emission: (rgb, 1) for every
// pixel; geom: (Sensor::linear_depth, p_film.x, p_film.y, hit ? 1 : 0); motion: (p_film - Sensor::prev_raster_coord(hit position), 0, 0),
// zero on a miss (frame_buffer.cpp:483-491, sensor.cpp:95-100 with the sensor-to-raster map of the analytic camera inverted).
template <int VARIANT>
__global__ void __launch_bounds__(PT_BLOCK) k_gbuffer(const __grid_constant__ PtScene S, const __grid_constant__ vb_camera_prev PC, uint32_t frame, float4 *out) {
    constexpr bool PRINCIPLED = (VARIANT & PT_PRINCIPLED) != 0, HERO = (VARIANT & PT_HERO) != 0;
    __shared__ uint2 s_stack_all[VB_SM_STACK * PT_BLOCK];
    uint2 *s_stack = s_stack_all + threadIdx.x;
    const VbTrav mg = vb_trav_init();
    uint32_t pix = blockIdx.x * PT_BLOCK + threadIdx.x;
    uint32_t n_pix = S.width * S.height;
    if (pix >= n_pix) return;
    const uint32_t px = pix % S.width, py = pix / S.width;
    Swl wl;
    const Swl *swl = nullptr;
    Sampler sampler;
    if (HERO) { sampler.start(px, py, frame, 0xffffffffu); (void)sampler.next_1d(); wl = swl_make(sampler.state & 0x00ffffffu); swl = &wl; }
    sampler.start(px, py, frame, 0u);
    f2 p_film;
    VbRay ray = generate_camera_ray(S, sampler, px, py, &p_film);
    VbHit hit;
    uint32_t d0 = 0, d1 = 0;
    vb_traverse<false, false>(S.accel, ray, hit, s_stack, PT_BLOCK, d0, d1, mg);
    float4 normal = make_float4(0.f, 0.f, 0.f, 0.f), albedo = make_float4(0.f, 0.f, 0.f, 1.f), emission = make_float4(0.f, 0.f, 0.f, 1.f);
    float4 geom = make_float4(0.f, p_film.x, p_film.y, 0.f), motion = make_float4(0.f, 0.f, 0.f, 0.f);
    if (hit.inst != VB_INVALID_ID) {
        Interaction it = compute_surface_interaction(S, hit.inst, hit.prim, hit.u, hit.v, true);
        f3 rd{ray.dx, ray.dy, ray.dz}, ro{ray.ox, ray.oy, ray.oz};
        it.wo = normalize(-rd);
        f3 ns = it.shading.z;
        normal = make_float4(ns.x, ns.y, ns.z, 1.f);
        const float *m = S.cam.c2w;                                  // Sensor::linear_depth: rigid c2w -> projection on its third column
        geom.x = dot(f3{m[8], m[9], m[10]}, it.pos - f3{m[12], m[13], m[14]});
        geom.w = 1.f;
        {                                                            // Sensor::prev_raster_coord: world -> previous sensor space, / z, -> raster
            f3 pc = transform_point(PC.w2c, it.pos);
            float sx = (pc.x / pc.z) / PC.tan_fov_y_over_2, sy = (pc.y / pc.z) / PC.tan_fov_y_over_2;
            float rx = (sx - PC.screen_lo[0]) / (PC.screen_hi[0] - PC.screen_lo[0]) * (float)S.width;
            float ry = (PC.screen_hi[1] - sy) / (PC.screen_hi[1] - PC.screen_lo[1]) * (float)S.height;
            motion.x = p_film.x - rx; motion.y = p_film.y - ry;
        }
        if (it.mat_id != VB_INVALID_ID) {
            float cos_theta = dot(it.shading.z, it.wo);              // MaterialEvaluator::albedo, material.cpp:91-98
            f3 a;
            const uint32_t mt = S.materials[it.mat_id].type;
            if (PRINCIPLED && (mt == VB_MAT_MIX || mt == VB_MAT_ADD)) a = mixset_albedo(S, swl, it, cos_theta);
            else if (PRINCIPLED && mt == VB_MAT_PRINCIPLED) a = principled_albedo(S, swl, it, cos_theta);
            else a = lobe_albedo(create_lobe(S, swl, it), cos_theta);
            a = spec_to_srgb(S.spec, swl, a);
            albedo = make_float4(a.x, a.y, a.z, 1.f);
        }
        if (it.light_id != VB_INVALID_ID) {                          // LightSampler::evaluate_hit_wi, lightsampler.cpp:252-267
            const vb_light_rec &l = S.lights[it.light_id];
            LightEval ev{f3{0, 0, 0}, 0.f};
            if (l.type == VB_LIGHT_AREA) {
                LightEvalCtx pl{it.pos, it.ng, it.uv, 1.0f / it.prim_area};
                pl.pdf_pos *= alias_pmf(S, l, it.prim_id);
                ev = area_evaluate_wi(S, swl, l, ro, pl);
            }
            f3 e = spec_to_srgb(S.spec, swl, ev.L);
            emission = make_float4(e.x, e.y, e.z, 1.f);
        }
    }
    out[pix] = normal; out[(size_t)n_pix + pix] = albedo; out[2 * (size_t)n_pix + pix] = emission; out[3 * (size_t)n_pix + pix] = geom;
    out[4 * (size_t)n_pix + pix] = motion;
}

// rt_gradient (frame_buffer.cpp:221-275): sums in the reference's loop order (x outer, y inner over the clipped 3x3 window, centre included
// and contributing nothing), then the divisions by the sample counts - a 1-pixel-wide picture divides by zero there as well.
__global__ void __launch_bounds__(256) k_gbuffer_gradient(const float4 *normal, const float4 *geom, float4 *out, int w, int h) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= w * h) return;
    const int cx = pix % w, cy = pix / w;
    const float4 cn = normal[pix];
    const float cd = geom[pix].x;
    uint32_t xn = 0, yn = 0;
    f3 ndx{0, 0, 0}, ndy{0, 0, 0};
    float ddx = 0.f, ddy = 0.f;
    for (int x = max(0, cx - 1); x <= min(cx + 1, w - 1); ++x)
        for (int y = max(0, cy - 1); y <= min(cy + 1, h - 1); ++y) {
            const float4 nn = normal[y * w + x];
            const float nd = geom[y * w + x].x;
            const f3 c{cn.x, cn.y, cn.z}, n{nn.x, nn.y, nn.z};
            if (cx > x) { xn++; ndx = ndx + (c - n); ddx += cd - nd; }
            else if (x > cx) { xn++; ndx = ndx + (n - c); ddx += nd - cd; }
            if (cy > y) { yn++; ndy = ndy + (c - n); ddy += cd - nd; }
            else if (y > cy) { yn++; ndy = ndy + (n - c); ddy += nd - cd; }
        }
    const float fx = (float)xn, fy = (float)yn;
    ndx = f3{ndx.x / fx, ndx.y / fx, ndx.z / fx}; ndy = f3{ndy.x / fy, ndy.y / fy, ndy.z / fy};
    const f3 nf{fabsf(ndx.x) + fabsf(ndy.x), fabsf(ndx.y) + fabsf(ndy.y), fabsf(ndx.z) + fabsf(ndy.z)};
    ddx = ddx / fx; ddy = ddy / fy;
    out[pix] = make_float4(length(nf), fabsf(ddx) + fabsf(ddy), 0.f, 0.f);
}

// vb_dump_path_trace: the megakernel loop for ONE pixel-sample, every lane of one warp walking the same path (no divergence, lane 0
// writes): one record per closest-hit ray with the path state before pt_bounce, then the VB_PATH_END record.
template <int VARIANT>
__global__ void __launch_bounds__(32) k_dump_path(const __grid_constant__ PtScene S, uint32_t px, uint32_t py, uint32_t frame,
                                                  vb_path_vertex *out, uint32_t max_out, uint32_t *n_out) {
    __shared__ uint2 s_stack_all[VB_SM_STACK * 32];
    uint2 *s_stack = s_stack_all + threadIdx.x;
    const VbTrav mg = vb_trav_init();
    uint32_t d0 = 0, d1 = 0, n = 0;
    Sampler sampler;
    sampler.start(px, py, frame, 0u);
    VbRay ray = generate_camera_ray(S, sampler, px, py);
    PathState ps;
    path_begin(S, ps, px, py, frame);
    auto put = [&](const VbRay &r, uint32_t inst, uint32_t prim, float u, float v, float t) {
        if (threadIdx.x == 0 && n < max_out) {
            vb_path_vertex &o = out[n];
            o.ray_o[0] = r.ox; o.ray_o[1] = r.oy; o.ray_o[2] = r.oz; o.ray_tmin = r.tmin;
            o.ray_d[0] = r.dx; o.ray_d[1] = r.dy; o.ray_d[2] = r.dz; o.ray_tmax = r.tmax;
            o.inst = inst; o.prim = prim; o.u = u; o.v = v; o.t = t;
            o.bounces = ps.bounces; o.sampler_state = ps.sampler.state; o.medium = ps.ray_medium;
            o.throughput[0] = ps.throughput.x; o.throughput[1] = ps.throughput.y; o.throughput[2] = ps.throughput.z; o.scatter_pdf = ps.scatter_pdf;
            o.radiance[0] = ps.ret.x; o.radiance[1] = ps.ret.y; o.radiance[2] = ps.ret.z; o.eta_scale = ps.eta_scale;
        }
        n++;
    };
    if (S.params.max_depth > 0) {
        for (;;) {
            VbHit hit;
            vb_traverse<false, false>(S.accel, ray, hit, s_stack, 32, d0, d1, mg);
            put(ray, hit.inst, hit.prim, hit.u, hit.v, hit.t);
            VbRay shadow;
            f3 pending{0, 0, 0};
            int flags = pt_bounce<VARIANT>(S, ps, ray, hit, shadow, pending);
            if (flags & PT_SHADOW) {
                VbHit sh;
                if (!vb_traverse<true, false>(S.accel, shadow, sh, s_stack, 32, d0, d1, mg)) ps.ret = ps.ret + pending;
            }
            if (!(flags & PT_CONTINUE)) break;
        }
    }
    if (threadIdx.x == 0 && n < max_out) {                          // the end record: final radiance + sampler state only
        vb_path_vertex o;
        memset(&o, 0, sizeof(o));
        o.inst = VB_PATH_END; o.prim = VB_PATH_END; o.sampler_state = ps.sampler.state;
        o.radiance[0] = ps.ret.x; o.radiance[1] = ps.ret.y; o.radiance[2] = ps.ret.z;
        out[n] = o;
    }
    n++;
    if (threadIdx.x == 0) *n_out = n;
}

// render-graph passes (pass/tonemapping.cpp:26-33, pass/gamma.cpp:21-27). linear_to_srgb: the sRGB OETF on [0, 1] with the
// power written over the shared vlog / vexp, so that the host restatement gives the same bits.
__device__ __forceinline__ float linear_to_srgb1(float x) {
    x = fmin_(fmax_(x, 0.f), 1.f);
    return x <= 0.0031308f ? 12.92f * x : 1.055f * vexp(vlog(x) * (1.0f / 2.4f)) - 0.055f;
}
__global__ void __launch_bounds__(256) k_pass_tonemapping(const float4 *in, float4 *out, uint32_t n, int tonemap) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = in[i];
    float c[3] = {v.x, v.y, v.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float x = c[k];
        if (tonemap == VB_TONEMAP_ACES) {
            float y = (x * (2.51f * x + 0.03f)) / (x * (2.43f * x + 0.59f) + 0.14f);
            x = fmin_(1.0f, fmax_(0.0f, y));
        } else if (tonemap == VB_TONEMAP_REINHARD) {
            x = x / (x + 1.0f);
        }
        c[k] = x;
    }
    out[i] = make_float4(c[0], c[1], c[2], 1.0f);
}
__global__ void __launch_bounds__(256) k_pass_gamma(const float4 *in, float4 *out, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = in[i];
    out[i] = make_float4(linear_to_srgb1(v.x), linear_to_srgb1(v.y), linear_to_srgb1(v.z), linear_to_srgb1(v.w));
}

// reference frame_buffer.cpp:72-74,135-144 + tonemapper/impl.cpp:21-46
// hero spectrum: sigmoid-polynomial coefficients of every CONSTANT colour slot, decoded once per scene with the same device functions
// the per-vertex path uses (DevSpectrum::cache). Entries of textured slots are written too but never read.
__global__ void __launch_bounds__(128) k_spec_cache(PtScene S, uint32_t n_mat, uint32_t n_lights, uint32_t n_med, float4 *out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 4u * n_mat) {
        const vb_material_rec &m = S.materials[i >> 2];
        const uint32_t k = i & 3u;
        f3 rgb = k == SC_COLOR ? f3{m.color.value[0], m.color.value[1], m.color.value[2]}
               : k == SC_SPEC  ? f3{m.spec.value[0], m.spec.value[1], m.spec.value[2]}
               : k == SC_SHEEN ? f3{m.params[4], m.params[5], m.params[6]} : f3{m.params[10], m.params[11], m.params[12]};
        out[i] = spec_coefs_albedo(S.spec, rgb);
        return;
    }
    i -= 4u * n_mat;
    if (i < n_lights) {
        const vb_light_rec &l = S.lights[i];
        out[4u * n_mat + i] = spec_coefs_unbound(S.spec, f3{l.color.value[0], l.color.value[1], l.color.value[2]});
        return;
    }
    i -= n_lights;
    if (i < 2u * n_med) {
        const vb_medium_rec &m = S.mediums[i >> 1];
        f3 rgb = (i & 1u) ? f3{m.sigma_s[0] * m.scale, m.sigma_s[1] * m.scale, m.sigma_s[2] * m.scale}
                          : f3{m.sigma_a[0] + m.sigma_s[0], m.sigma_a[1] + m.sigma_s[1], m.sigma_a[2] + m.sigma_s[2]} * m.scale;
        out[4u * n_mat + n_lights + i] = spec_coefs_unbound(S.spec, rgb);
    }
}

__global__ void __launch_bounds__(256) k_film_resolve(const float4 *accum, float4 *out, uint32_t n, int mode, float exposure, int tonemap) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = accum[i];
    if (mode == VB_ACCUM_SUM) {
        float inv = v.w > 0.f ? 1.0f / v.w : 0.f;
        v.x *= inv; v.y *= inv; v.z *= inv;
    }
    float c[3] = {v.x, v.y, v.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float x = 1.0f - expf(-c[k] * exposure);
        if (tonemap == VB_TONEMAP_ACES) {
            float y = (x * (2.51f * x + 0.03f)) / (x * (2.43f * x + 0.59f) + 0.14f);
            x = fmin_(1.0f, fmax_(0.0f, y));
        } else if (tonemap == VB_TONEMAP_REINHARD) {
            x = x / (x + 1.0f);
        }
        c[k] = x;
    }
    out[i] = make_float4(c[0], c[1], c[2], 1.0f);
}

}  // namespace

struct vb_wf_buffers;
struct vb_pt_state {
    vb_wf_buffers *wf = nullptr;       // queues of the wavefront schedule (allocated on first use)
    uint32_t n_mat_types = 0;          // distinct material models in use (one: ordering the shading queue is pointless)
    uint64_t work[4] = {0, 0, 0, 0};
    uint32_t n_alias = 0;
    uint32_t lsel_offset = 0, lsel_size = 0; float lsel_integral = 0.f;     // vb_light_sampler_set
    uint32_t env_separate = 0; float env_prob = 0.f;                        // vb_light_sampler_env_set
    bool has_principled = false;       // principled, mix or add in the scene: the heavy shading launch
    bool has_lobeset = false;          // mix or add in the scene: the heavy launch's PT_LOBESET instantiation
    std::vector<DevMesh> h_meshes;
    DevMesh *d_meshes = nullptr;
    DevInst *d_inst = nullptr;
    vb_material_rec *d_mat = nullptr; uint32_t n_mat = 0;
    vb_light_rec *d_lights = nullptr; uint32_t n_lights = 0; uint32_t env_light = VB_INVALID_ID;
    vb_alias_entry *d_alias = nullptr; float *d_func = nullptr;
    DevTex *d_tex = nullptr; std::vector<float4 *> tex_mem;
    DevLut luts[VB_LUT_COUNT] = {};
    vb_medium_rec *d_mediums = nullptr; uint32_t n_mediums = 0;
    float *d_spd_values = nullptr; float4 *d_spd_meta = nullptr; uint32_t n_spds = 0; std::vector<uint32_t> h_spd_sizes;
    float *d_rgb2spec = nullptr; vb_spectrum_rec spectrum{}; bool have_spectrum = false;
    float4 *d_gbuffer = nullptr; size_t gbuffer_n = 0;       // VB_AOV_COUNT planes of w*h float4 (vb_gbuffer_render, vb_gbuffer_gradient)
    vb_path_vertex *dump_out = nullptr; uint32_t dump_max = 0, dump_px = 0, dump_py = 0; uint32_t *dump_n = nullptr;   // vb_dump_path_trace request (pt_launch)
    bool gbuffer_valid = false;                              // a vb_gbuffer_render has filled the planes
    vb_camera_prev cam_prev{}; bool have_cam_prev = false;   // vb_camera_prev_set
    float4 *view = nullptr;                                  // VB_FILM_VIEW (vb_pass_gamma), allocated on first use
    float4 *d_spec_cache = nullptr; size_t spec_cache_n = 0; bool spec_dirty = true;   // hero: decoded constant colour slots (k_spec_cache)
    std::vector<vb_material_rec> h_mat;            // kept for the SPD index checks at render time
    std::vector<vb_light_rec> h_lights;            // kept for the instance / table checks at render time
    vb_camera cam{}; bool have_cam = false;
    vb_pt_params params{}; bool have_params = false;
    uint32_t w = 0, h = 0;
    float4 *rt = nullptr, *accum = nullptr, *output = nullptr;
    uint32_t frames_done = 0;
    PtCounters *d_counters = nullptr;
    PtCounters last{};
    uint64_t geom_version = ~0ull;
};

#include "vb_pt_wavefront_host.cuh"

static vb_pt_state *pt_state(vb_ctx *ctx) {
    if (!ctx->pt) ctx->pt = new vb_pt_state();
    return ctx->pt;
}

void vb_pt_destroy(vb_ctx *ctx) {
    vb_pt_state *p = ctx->pt;
    if (!p) return;
    cudaFree(p->d_meshes); cudaFree(p->d_inst); cudaFree(p->d_mat); cudaFree(p->d_lights); cudaFree(p->d_alias); cudaFree(p->d_func);
    cudaFree(p->d_tex);
    for (auto t : p->tex_mem) cudaFree(t);
    for (auto &l : p->luts) cudaFree((void *)l.data);
    cudaFree(p->d_mediums);
    cudaFree(p->d_spd_values); cudaFree(p->d_spd_meta); cudaFree(p->d_rgb2spec); cudaFree(p->d_spec_cache); cudaFree(p->d_gbuffer); cudaFree(p->view);
    cudaFree(p->rt); cudaFree(p->accum); cudaFree(p->output); cudaFree(p->d_counters);
    if (p->wf) { wf_free(*p->wf); delete p->wf; }
    delete p;
    ctx->pt = nullptr;
}

// device tables for meshes/instances (the analogue of Geometry::upload, geometry.cpp:64-71)
static int pt_sync_geometry(vb_ctx *ctx) {
    vb_pt_state *p = pt_state(ctx);
    if (p->geom_version == ctx->geom_version) return VB_OK;         // meshes and instances unchanged since the tables were made
    p->geom_version = ~0ull;
    cudaFree(p->d_meshes); cudaFree(p->d_inst); p->d_meshes = nullptr; p->d_inst = nullptr;
    std::vector<DevMesh> hm(ctx->meshes.size());
    for (size_t i = 0; i < hm.size(); ++i) hm[i] = DevMesh{ctx->meshes[i].vtx, ctx->meshes[i].tri4};
    std::vector<DevInst> hi(ctx->instances.size());
    for (size_t i = 0; i < hi.size(); ++i) {
        const vb_instance &s = ctx->instances[i];
        if (ctx->meshes[s.mesh_id].soup) return ctx->fail(VB_ERR_UNSUPPORTED, "path tracing needs indexed meshes (vb_mesh_upload)");
        DevInst &d = hi[i];
        memset(&d, 0, sizeof(d));
        d.mesh_id = s.mesh_id; d.mat_id = s.material_id; d.light_id = s.light_id; d.inside = s.inside_medium; d.outside = s.outside_medium;
        memcpy(d.o2w, s.o2w, 64);
        double m[3][3], cf[3][3];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) m[r][c] = s.o2w[4 * c + r];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) {
                int r1 = (r + 1) % 3, r2 = (r + 2) % 3, c1 = (c + 1) % 3, c2 = (c + 2) % 3;
                cf[r][c] = m[r1][c1] * m[r2][c2] - m[r1][c2] * m[r2][c1];
            }
        double det = m[0][0] * cf[0][0] + m[0][1] * cf[0][1] + m[0][2] * cf[0][2];
        double sg = det < 0 ? -1.0 : 1.0;
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) d.nmat[3 * r + c] = (float)(sg * cf[r][c]);
    }
    if (!hm.empty()) {
        VB_CUDA(cudaMalloc(&p->d_meshes, sizeof(DevMesh) * hm.size()));
        VB_CUDA(cudaMemcpy(p->d_meshes, hm.data(), sizeof(DevMesh) * hm.size(), cudaMemcpyHostToDevice));
    }
    if (!hi.empty()) {
        VB_CUDA(cudaMalloc(&p->d_inst, sizeof(DevInst) * hi.size()));
        VB_CUDA(cudaMemcpy(p->d_inst, hi.data(), sizeof(DevInst) * hi.size(), cudaMemcpyHostToDevice));
    }
    p->geom_version = ctx->geom_version;
    return VB_OK;
}

extern "C" {

int vb_textures_set(vb_ctx *ctx, const vb_texture *tex, uint32_t n) {
    if (!ctx) return VB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    for (auto t : p->tex_mem) cudaFree(t);
    p->tex_mem.clear();
    cudaFree(p->d_tex); p->d_tex = nullptr;
    std::vector<DevTex> ht(n);
    for (uint32_t i = 0; i < n; ++i) {
        if (!tex[i].texels || !tex[i].width || !tex[i].height) return ctx->fail(VB_ERR_INVALID, "vb_textures_set: empty texture");
        float4 *d = nullptr;
        size_t bytes = sizeof(float4) * (size_t)tex[i].width * tex[i].height;
        VB_CUDA(cudaMalloc(&d, bytes));
        VB_CUDA(cudaMemcpy(d, tex[i].texels, bytes, cudaMemcpyHostToDevice));
        p->tex_mem.push_back(d);
        ht[i] = DevTex{d, tex[i].width, tex[i].height};
    }
    if (n) {
        VB_CUDA(cudaMalloc(&p->d_tex, sizeof(DevTex) * n));
        VB_CUDA(cudaMemcpy(p->d_tex, ht.data(), sizeof(DevTex) * n, cudaMemcpyHostToDevice));
    }
    return VB_OK;
}

int vb_materials_set(vb_ctx *ctx, const vb_material_rec *mat, uint32_t n) {
    if (!ctx) return VB_ERR_INVALID;
    if (n && !mat) return ctx->fail(VB_ERR_INVALID, "vb_materials_set: null argument");
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    for (uint32_t i = 0; i < n; ++i) {
        if (mat[i].type > VB_MAT_ADD)
            return ctx->fail(VB_ERR_UNSUPPORTED, "vb_materials_set: material type " + std::to_string(mat[i].type) + " is not a known type");
        if (mat[i].type == VB_MAT_MIX || mat[i].type == VB_MAT_ADD)
            for (int c = 0; c < 2; ++c) {
                uint32_t child = mat[i].spd[c];
                if (child >= n) return ctx->fail(VB_ERR_INVALID, "vb_materials_set: mix / add child index out of range");
                if (mat[child].type == VB_MAT_MIX || mat[child].type == VB_MAT_ADD)
                    return ctx->fail(VB_ERR_UNSUPPORTED, "vb_materials_set: a mix / add material nested in another one is not supported (children: single-lobe materials or principled_bsdf)");
            }
    }
    cudaFree(p->d_mat); p->d_mat = nullptr; p->n_mat = n;
    uint32_t types = 0;
    for (uint32_t i = 0; i < n; ++i) types |= 1u << mat[i].type;
    p->n_mat_types = (uint32_t)__builtin_popcount(types);
    p->has_lobeset = (types & ((1u << VB_MAT_MIX) | (1u << VB_MAT_ADD))) != 0;
    p->has_principled = (types & ((1u << VB_MAT_PRINCIPLED) | (1u << VB_MAT_MIX) | (1u << VB_MAT_ADD))) != 0;   // the lobe-SET materials: heavy shading launch
    p->h_mat.assign(mat, mat + n);
    p->spec_dirty = true;
    if (n) {
        VB_CUDA(cudaMalloc(&p->d_mat, sizeof(vb_material_rec) * n));
        VB_CUDA(cudaMemcpy(p->d_mat, mat, sizeof(vb_material_rec) * n, cudaMemcpyHostToDevice));
    }
    return VB_OK;
}

int vb_luts_set(vb_ctx *ctx, int which, const float *data, uint32_t nx, uint32_t ny, uint32_t nz, uint32_t channels) {
    if (!ctx) return VB_ERR_INVALID;
    if (which < 0 || which >= VB_LUT_COUNT) return ctx->fail(VB_ERR_INVALID, "vb_luts_set: unknown table");
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    cudaFree((void *)p->luts[which].data);
    p->luts[which] = DevLut{};
    if (!data) return VB_OK;                       // clears the table: no compensation
    size_t n = (size_t)nx * ny * nz * channels;
    if (n == 0) return ctx->fail(VB_ERR_INVALID, "vb_luts_set: empty table");
    float *d = nullptr;
    VB_CUDA(cudaMalloc(&d, n * sizeof(float)));
    VB_CUDA(cudaMemcpy(d, data, n * sizeof(float), cudaMemcpyHostToDevice));
    p->luts[which] = DevLut{d, nx, ny, nz, channels};
    return VB_OK;
}

int vb_mediums_set(vb_ctx *ctx, const vb_medium_rec *med, uint32_t n) {
    if (!ctx) return VB_ERR_INVALID;
    if (n && !med) return ctx->fail(VB_ERR_INVALID, "vb_mediums_set: null argument");
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    for (uint32_t i = 0; i < n; ++i)
        if (med[i].type != VB_MEDIUM_HOMOGENEOUS) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_mediums_set: medium type not implemented");
    cudaFree(p->d_mediums); p->d_mediums = nullptr; p->n_mediums = n;
    p->spec_dirty = true;
    if (n) {
        VB_CUDA(cudaMalloc(&p->d_mediums, sizeof(vb_medium_rec) * n));
        VB_CUDA(cudaMemcpy(p->d_mediums, med, sizeof(vb_medium_rec) * n, cudaMemcpyHostToDevice));
    }
    return VB_OK;
}

int vb_spds_set(vb_ctx *ctx, const float *values, const uint32_t *offsets, uint32_t n) {
    if (!ctx) return VB_ERR_INVALID;
    if (n && (!values || !offsets)) return ctx->fail(VB_ERR_INVALID, "vb_spds_set: null argument");
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    std::vector<float4> meta(n);
    std::vector<uint32_t> sizes(n);
    for (uint32_t i = 0; i < n; ++i) {                           // validate everything before touching the current pool
        if (offsets[i + 1] < offsets[i] + 2) return ctx->fail(VB_ERR_INVALID, "vb_spds_set: an SPD needs at least 2 samples and ascending offsets");
        uint32_t size = offsets[i + 1] - offsets[i];             // SPD::init / SPD::eval constants (spd.cpp:50,70-72)
        sizes[i] = size;
        float interval = 471.0f / (float)size;
        uint32_t sample_count = (uint32_t)((830.f - 360.f) / interval) + 1u;      // <= size for every size >= 2
        if (sample_count > size || sample_count < 2) return ctx->fail(VB_ERR_INVALID, "vb_spds_set: SPD too short for its sampling interval");
        float off_bits; std::memcpy(&off_bits, &offsets[i], 4);
        meta[i] = make_float4(off_bits, interval, (float)(sample_count - 2u), 0.f);
    }
    cudaFree(p->d_spd_values); cudaFree(p->d_spd_meta);
    p->d_spd_values = nullptr; p->d_spd_meta = nullptr; p->n_spds = 0;           // stays empty if an upload below fails
    p->spec_dirty = true;
    p->h_spd_sizes.clear();
    if (n) {
        VB_CUDA(cudaMalloc(&p->d_spd_values, sizeof(float) * offsets[n]));
        VB_CUDA(cudaMemcpy(p->d_spd_values, values, sizeof(float) * offsets[n], cudaMemcpyHostToDevice));
        VB_CUDA(cudaMalloc(&p->d_spd_meta, sizeof(float4) * n));
        VB_CUDA(cudaMemcpy(p->d_spd_meta, meta.data(), sizeof(float4) * n, cudaMemcpyHostToDevice));
    }
    p->h_spd_sizes = sizes; p->n_spds = n;
    return VB_OK;
}

int vb_spectrum_set(vb_ctx *ctx, const vb_spectrum_rec *spec) {
    if (!ctx) return VB_ERR_INVALID;
    if (!spec) return ctx->fail(VB_ERR_INVALID, "vb_spectrum_set: null argument");
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    if (spec->type > VB_SPECTRUM_HERO) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_spectrum_set: unknown spectrum type");
    if (spec->type == VB_SPECTRUM_HERO) {
        if (spec->dimension != 3) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_spectrum_set: hero spectrum is implemented for dimension 3 only");
        if (!spec->rgb2spec || spec->rgb2spec_res < 2) return ctx->fail(VB_ERR_INVALID, "vb_spectrum_set: hero spectrum needs the rgb2spec coefficient table");
        if (!(spec->cie_y_integral > 0.f)) return ctx->fail(VB_ERR_INVALID, "vb_spectrum_set: cie_y_integral must be positive");
    }
    cudaFree(p->d_rgb2spec); p->d_rgb2spec = nullptr;
    p->spectrum = *spec; p->spectrum.rgb2spec = nullptr; p->have_spectrum = true;
    p->spec_dirty = true;
    if (spec->type == VB_SPECTRUM_HERO) {
        size_t bytes = sizeof(float) * 3 * 4 * (size_t)spec->rgb2spec_res * spec->rgb2spec_res * spec->rgb2spec_res;
        VB_CUDA(cudaMalloc(&p->d_rgb2spec, bytes));
        VB_CUDA(cudaMemcpy(p->d_rgb2spec, spec->rgb2spec, bytes, cudaMemcpyHostToDevice));
    }
    return VB_OK;
}

int vb_lights_set(vb_ctx *ctx, const vb_light_rec *lights, uint32_t n, const vb_alias_entry *alias_table, const float *alias_func,
                  uint32_t n_alias) {
    if (!ctx) return VB_ERR_INVALID;
    if ((n && !lights) || (n_alias && (!alias_table || !alias_func))) return ctx->fail(VB_ERR_INVALID, "vb_lights_set: null argument");
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    for (uint32_t i = 0; i < n; ++i) {
        if (lights[i].type > VB_LIGHT_DIRECTIONAL) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_lights_set: light type not implemented yet");
        bool tabled = lights[i].type == VB_LIGHT_AREA || lights[i].type == VB_LIGHT_ENV;
        if (tabled && lights[i].alias_offset + lights[i].alias_size > n_alias) return ctx->fail(VB_ERR_INVALID, "vb_lights_set: alias range out of bounds");
        if (lights[i].type == VB_LIGHT_ENV) {
            uint64_t cells = (uint64_t)lights[i].env_res[0] * lights[i].env_res[1];
            if (!cells || lights[i].alias_size != lights[i].env_res[1] || lights[i].cond_offset + cells > n_alias)
                return ctx->fail(VB_ERR_INVALID, "vb_lights_set: env map tables out of bounds");
        }
    }
    cudaFree(p->d_lights); cudaFree(p->d_alias); cudaFree(p->d_func);
    p->d_lights = nullptr; p->d_alias = nullptr; p->d_func = nullptr; p->n_lights = n; p->n_alias = n_alias;
    p->lsel_offset = p->lsel_size = 0; p->lsel_integral = 0.f;
    p->env_separate = 0; p->env_prob = 0.f;
    p->h_lights.assign(lights, lights + n);
    p->spec_dirty = true;
    p->env_light = VB_INVALID_ID;
    for (uint32_t i = 0; i < n; ++i) if (lights[i].type == VB_LIGHT_ENV || lights[i].type == VB_LIGHT_DIRECTIONAL) p->env_light = i;
    if (n) {
        VB_CUDA(cudaMalloc(&p->d_lights, sizeof(vb_light_rec) * n));
        VB_CUDA(cudaMemcpy(p->d_lights, lights, sizeof(vb_light_rec) * n, cudaMemcpyHostToDevice));
    }
    if (n_alias) {
        VB_CUDA(cudaMalloc(&p->d_alias, sizeof(vb_alias_entry) * n_alias));
        VB_CUDA(cudaMemcpy(p->d_alias, alias_table, sizeof(vb_alias_entry) * n_alias, cudaMemcpyHostToDevice));
        VB_CUDA(cudaMalloc(&p->d_func, sizeof(float) * n_alias));
        VB_CUDA(cudaMemcpy(p->d_func, alias_func, sizeof(float) * n_alias, cudaMemcpyHostToDevice));
    }
    return VB_OK;
}

int vb_light_sampler_set(vb_ctx *ctx, uint32_t kind, uint32_t alias_offset, uint32_t alias_size, float alias_integral) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (kind == VB_LIGHT_SAMPLER_UNIFORM) { p->lsel_offset = p->lsel_size = 0; p->lsel_integral = 0.f; return VB_OK; }
    if (kind != VB_LIGHT_SAMPLER_POWER) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_light_sampler_set: light sampler kind not implemented");
    if (alias_size == 0 || alias_size != p->n_lights) return ctx->fail(VB_ERR_INVALID, "vb_light_sampler_set: the selection table needs one entry per light (call vb_lights_set first)");
    if ((uint64_t)alias_offset + alias_size > p->n_alias) return ctx->fail(VB_ERR_INVALID, "vb_light_sampler_set: alias range out of bounds");
    p->lsel_offset = alias_offset; p->lsel_size = alias_size; p->lsel_integral = alias_integral;
    return VB_OK;
}
int vb_light_sampler_env_set(vb_ctx *ctx, uint32_t env_separate, float env_prob) {
    if (!ctx) return VB_ERR_INVALID;
    if (env_separate && !(env_prob > 0.f && env_prob < 1.f)) return ctx->fail(VB_ERR_INVALID, "vb_light_sampler_env_set: env_prob must lie inside (0, 1)");
    vb_pt_state *p = pt_state(ctx);
    p->env_separate = env_separate ? 1u : 0u; p->env_prob = env_prob;
    return VB_OK;
}
int vb_camera_set(vb_ctx *ctx, const vb_camera *cam) {
    if (!ctx || !cam) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    p->cam = *cam; p->have_cam = true;
    return VB_OK;
}
int vb_integrator_set(vb_ctx *ctx, const vb_pt_params *prm) {
    if (!ctx || !prm) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (prm->spectrum > VB_SPECTRUM_HERO || prm->sampler != 0) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_integrator_set: srgb / hero spectrum + independent sampler only");
    p->params = *prm; p->have_params = true;
    return VB_OK;
}

int vb_film_create(vb_ctx *ctx, uint32_t width, uint32_t height) {
    if (!ctx) return VB_ERR_INVALID;
    if (!width || !height) return ctx->fail(VB_ERR_INVALID, "vb_film_create: empty film");
    cudaSetDevice(ctx->device);
    vb_pt_state *p = pt_state(ctx);
    cudaFree(p->rt); cudaFree(p->accum); cudaFree(p->output); cudaFree(p->view);
    p->rt = p->accum = p->output = p->view = nullptr;
    size_t bytes = sizeof(float4) * (size_t)width * height;
    VB_CUDA(cudaMalloc(&p->rt, bytes)); VB_CUDA(cudaMalloc(&p->accum, bytes)); VB_CUDA(cudaMalloc(&p->output, bytes));
    p->w = width; p->h = height;
    if (!p->d_counters) VB_CUDA(cudaMalloc(&p->d_counters, sizeof(PtCounters)));
    return vb_film_clear(ctx);
}
int vb_film_clear(vb_ctx *ctx) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_film_clear: no film");
    size_t bytes = sizeof(float4) * (size_t)p->w * p->h;
    VB_CUDA(cudaMemsetAsync(p->rt, 0, bytes, ctx->stream));
    VB_CUDA(cudaMemsetAsync(p->accum, 0, bytes, ctx->stream));
    VB_CUDA(cudaMemsetAsync(p->output, 0, bytes, ctx->stream));
    p->frames_done = 0;
    return VB_OK;
}

// The previous-camera record of a camera: inverse of its (affine) camera-to-world matrix by Gauss-Jordan elimination in double
// precision, rounded once to float. Used when no vb_camera_prev_set call has been made (static camera).
static int camera_prev_from(vb_ctx *ctx, const vb_camera &cam, vb_camera_prev *out) {
    double a[4][8];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { a[r][c] = cam.c2w[c * 4 + r]; a[r][4 + c] = r == c ? 1.0 : 0.0; }
    for (int i = 0; i < 4; ++i) {
        int piv = i;
        for (int r = i + 1; r < 4; ++r) if (std::fabs(a[r][i]) > std::fabs(a[piv][i])) piv = r;
        if (!(std::fabs(a[piv][i]) > 0.0)) return ctx->fail(VB_ERR_INVALID, "vb_gbuffer_render: the camera_to_world matrix is singular");
        if (piv != i) for (int c = 0; c < 8; ++c) std::swap(a[i][c], a[piv][c]);
        const double inv = 1.0 / a[i][i];
        for (int c = 0; c < 8; ++c) a[i][c] *= inv;
        for (int r = 0; r < 4; ++r) if (r != i) { const double f = a[r][i]; for (int c = 0; c < 8; ++c) a[r][c] -= f * a[i][c]; }
    }
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out->w2c[c * 4 + r] = (float)a[r][4 + c];
    out->tan_fov_y_over_2 = cam.tan_fov_y_over_2;
    out->screen_lo[0] = cam.screen_lo[0]; out->screen_lo[1] = cam.screen_lo[1]; out->screen_hi[0] = cam.screen_hi[0]; out->screen_hi[1] = cam.screen_hi[1];
    out->pad[0] = out->pad[1] = out->pad[2] = 0.f;
    return VB_OK;
}

static int pt_launch(vb_ctx *ctx, const char *fn, uint32_t first, uint32_t n_frames, uint32_t stride, int mode, bool dump,
                     uint32_t *hi, uint32_t *hp, float *hu, float *hv, float *ht, float4 *gbuf = nullptr) {
    vb_pt_state *p = pt_state(ctx);
    if (!ctx->accel.built) return ctx->fail(VB_ERR_STATE, std::string(fn) + ": accel not built");
    if (!p->accum) return ctx->fail(VB_ERR_STATE, std::string(fn) + ": no film (vb_film_create)");
    if (!p->have_cam || !p->have_params) return ctx->fail(VB_ERR_STATE, std::string(fn) + ": camera/integrator not set");
    for (auto &in : ctx->instances) {
        if (in.material_id != VB_INVALID_ID && in.material_id >= p->n_mat) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": instance material_id out of range");
        if (in.light_id != VB_INVALID_ID && in.light_id >= p->n_lights) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": instance light_id out of range");
    }
    // an area light samples the triangles of ITS instance through ITS alias table (area.cpp:74-83): both must agree with the geometry
    for (auto &l : p->h_lights) {
        if (l.type != VB_LIGHT_AREA) continue;
        if (l.inst_id >= ctx->instances.size()) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": area light inst_id out of range");
        const uint32_t mesh = ctx->instances[l.inst_id].mesh_id;
        if (mesh >= ctx->meshes.size() || (uint64_t)l.alias_size != ctx->meshes[mesh].nt)
            return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": area light alias_size differs from its instance's triangle count");
    }
    // every texture slot of the material and light records must name an uploaded texture (eval_slot3 indexes S.textures with it)
    {
        const uint32_t n_tex = (uint32_t)p->tex_mem.size();
        auto bad = [n_tex](const vb_slot &sl) { return sl.tex_id != VB_INVALID_ID && sl.tex_id >= n_tex; };
        for (auto &m : p->h_mat)
            if (bad(m.color) || bad(m.roughness) || bad(m.spec) || bad(m.ior) || bad(m.sigma))
                return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": material slot tex_id out of range (vb_textures_set)");
        for (auto &l : p->h_lights)
            if (bad(l.color)) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": light colour tex_id out of range (vb_textures_set)");
    }
    if (p->env_separate && p->env_light != VB_INVALID_ID && p->n_lights < 2)
        return ctx->fail(VB_ERR_UNSUPPORTED, std::string(fn) + ": env_separate needs a light besides the environment light");
    int rc = pt_sync_geometry(ctx);
    if (rc) return rc;
    PtScene S{};
    S.accel = VbAccelView{ctx->accel.nodes, ctx->accel.tris, ctx->accel.xf};
    S.meshes = p->d_meshes; S.instances = p->d_inst; S.materials = p->d_mat; S.lights = p->d_lights; S.n_lights = p->n_lights; S.env_light = p->env_light;
    S.lsel_offset = p->lsel_offset; S.lsel_size = p->lsel_size; S.lsel_integral = p->lsel_integral;
    S.env_separate = (p->env_separate && p->env_light != VB_INVALID_ID) ? 1u : 0u; S.env_prob = p->env_prob;
    S.alias_table = p->d_alias; S.alias_func = p->d_func; S.textures = p->d_tex;
    for (int k = 0; k < VB_LUT_COUNT; ++k) S.luts[k] = p->luts[k];
    S.mediums = p->d_mediums;
    if (p->params.spectrum > VB_SPECTRUM_HERO) return ctx->fail(VB_ERR_UNSUPPORTED, std::string(fn) + ": unknown spectrum type");
    if (p->params.spectrum == VB_SPECTRUM_HERO) {
        const vb_spectrum_rec &sp = p->spectrum;
        if (!p->have_spectrum || sp.type != VB_SPECTRUM_HERO || !p->d_rgb2spec) return ctx->fail(VB_ERR_STATE, std::string(fn) + ": hero spectrum selected but vb_spectrum_set was not called with it");
        if (sp.spd_x >= p->n_spds || sp.spd_y >= p->n_spds || sp.spd_z >= p->n_spds || sp.spd_d65 >= p->n_spds)
            return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": spectrum SPD index out of range (vb_spds_set)");
        for (auto &m : p->h_mat)
            if ((m.flags & 4u) && (m.spd[0] >= p->n_spds || (m.type == VB_MAT_METAL && m.spd[1] >= p->n_spds)))
                return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": material SPD index out of range (vb_spds_set)");
        if (p->h_spd_sizes[sp.spd_x] != p->h_spd_sizes[sp.spd_y] || p->h_spd_sizes[sp.spd_x] != p->h_spd_sizes[sp.spd_z])
            return ctx->fail(VB_ERR_UNSUPPORTED, std::string(fn) + ": the CIE x, y, z SPDs must have the same number of samples");
        S.spec = DevSpectrum{p->d_rgb2spec, sp.rgb2spec_res, p->d_spd_values, p->d_spd_meta, sp.spd_x, sp.spd_y, sp.spd_z, sp.spd_d65, sp.cie_y_integral, nullptr, 0u, 0u};
        if (ctx->opt_pt_spec_cache) {
            const uint32_t n_entries = 4u * p->n_mat + p->n_lights + 2u * p->n_mediums;
            if (p->spec_dirty || p->spec_cache_n != n_entries) {
                cudaFree(p->d_spec_cache); p->d_spec_cache = nullptr; p->spec_cache_n = n_entries;
                if (n_entries) {
                    VB_CUDA(cudaMalloc(&p->d_spec_cache, sizeof(float4) * n_entries));
                    k_spec_cache<<<(n_entries + 127) / 128, 128, 0, ctx->stream>>>(S, p->n_mat, p->n_lights, p->n_mediums, p->d_spec_cache);
                    VB_CHECK_LAUNCH();
                }
                p->spec_dirty = false;
            }
            S.spec.cache = p->d_spec_cache; S.spec.cache_lights = 4u * p->n_mat; S.spec.cache_mediums = 4u * p->n_mat + p->n_lights;
        }
    }
    if (p->params.process_mediums) {
        auto bad = [&](uint32_t id) { return id != VB_INVALID_ID && id >= p->n_mediums; };
        if (bad(p->cam.medium_id)) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": camera medium_id out of range");
        for (auto &in : ctx->instances)
            if (bad(in.inside_medium) || bad(in.outside_medium)) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": instance medium id out of range");
    }
    if (p->cam.filter_type == VB_FILTER_FITTED &&
        (uint64_t)p->cam.filter_table_offset + VB_FILTER_TABLE_SIZE * (VB_FILTER_TABLE_SIZE + 1) > p->n_alias)
        return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": filter table out of bounds of the alias arrays");
    if (p->cam.filter_type > VB_FILTER_FITTED) return ctx->fail(VB_ERR_UNSUPPORTED, std::string(fn) + ": unknown filter type");
    S.cam = p->cam; S.params = p->params; S.width = p->w; S.height = p->h;
    if (S.params.ray_offset_factor == 0.f) S.params.ray_offset_factor = 1.f;
    uint32_t n_pix = p->w * p->h;
    if (p->dump_out) {                                               // vb_dump_path_trace: one pixel-sample, one warp
        ctx->prof_begin("dump_path", 1);
        const bool hero = S.params.spectrum == VB_SPECTRUM_HERO;
        if (hero) k_dump_path<PT_HERO | PT_PRINCIPLED | PT_LOBESET><<<1, 32, 0, ctx->stream>>>(S, p->dump_px, p->dump_py, first, p->dump_out, p->dump_max, p->dump_n);
        else if (p->has_principled) k_dump_path<PT_PRINCIPLED | PT_LOBESET><<<1, 32, 0, ctx->stream>>>(S, p->dump_px, p->dump_py, first, p->dump_out, p->dump_max, p->dump_n);
        else k_dump_path<PT_LEAN><<<1, 32, 0, ctx->stream>>>(S, p->dump_px, p->dump_py, first, p->dump_out, p->dump_max, p->dump_n);
        ctx->prof_end();
        VB_CHECK_LAUNCH();
        return VB_OK;
    }
    if (gbuf) {                                                      // rt_geom's G-buffer planes for frame index `first`
        unsigned g = (n_pix + PT_BLOCK - 1) / PT_BLOCK;
        ctx->prof_begin("gbuffer", 1);
        const bool hero = S.params.spectrum == VB_SPECTRUM_HERO;
        vb_camera_prev PC = p->cam_prev;
        if (!p->have_cam_prev) {                                      // static camera: the previous camera is the current one
            int rc = camera_prev_from(ctx, p->cam, &PC);
            if (rc) return rc;
        }
        if (hero && p->has_principled) k_gbuffer<PT_HERO | PT_PRINCIPLED><<<g, PT_BLOCK, 0, ctx->stream>>>(S, PC, first, gbuf);
        else if (hero) k_gbuffer<PT_HERO><<<g, PT_BLOCK, 0, ctx->stream>>>(S, PC, first, gbuf);
        else if (p->has_principled) k_gbuffer<PT_PRINCIPLED><<<g, PT_BLOCK, 0, ctx->stream>>>(S, PC, first, gbuf);
        else k_gbuffer<PT_LEAN><<<g, PT_BLOCK, 0, ctx->stream>>>(S, PC, first, gbuf);
        ctx->prof_end();
        VB_CHECK_LAUNCH();
        return VB_OK;
    }
    if (!dump && ctx->opt_pt_schedule == 0) {
        // wavefront schedule: batches of whole frames, at most opt_pt_max_paths paths in flight
        if (!p->wf) p->wf = new vb_wf_buffers();
        uint32_t per_batch = (uint32_t)std::max<uint64_t>(1, (uint64_t)ctx->opt_pt_max_paths / n_pix);
        per_batch = std::min(per_batch, n_frames);
        if ((uint64_t)per_batch * n_pix >= (1ull << 32)) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": film too large");
        rc = wf_ensure(ctx, *p->wf, (size_t)per_batch * n_pix);
        if (rc) return rc;
        PtCounters total{};
        if (ctx->opt_pt_count_work) VB_CUDA(cudaMemsetAsync(p->wf->d_work, 0, 32, ctx->stream));
        for (uint32_t fi0 = 0; fi0 < n_frames; fi0 += per_batch) {
            rc = wf_render_batch(ctx, p, *p->wf, S, first, stride, fi0, std::min(per_batch, n_frames - fi0), mode, total);
            if (rc) return rc;
        }
        p->last = total;
        if (ctx->opt_pt_count_work) {
            VB_CUDA(cudaMemcpyAsync(p->work, p->wf->d_work, 32, cudaMemcpyDeviceToHost, ctx->stream));
            VB_CUDA(cudaStreamSynchronize(ctx->stream));
        }
        p->frames_done += n_frames;
        return VB_OK;
    }
    unsigned grid = (n_pix + PT_BLOCK - 1) / PT_BLOCK;
    VB_CUDA(cudaMemsetAsync(p->d_counters, 0, sizeof(PtCounters), ctx->stream));
    ctx->prof_begin("path_trace", 1);
    if (dump) k_path_trace<true, PT_LEAN><<<grid, PT_BLOCK, 0, ctx->stream>>>(S, first, 1, 1, mode, 0, p->rt, p->accum, p->d_counters, hi, hp, hu, hv, ht);
    else if (S.params.spectrum == VB_SPECTRUM_HERO) k_path_trace<false, PT_HERO | PT_PRINCIPLED | PT_LOBESET><<<grid, PT_BLOCK, 0, ctx->stream>>>(S, first, n_frames, stride, mode, p->frames_done, p->rt,
                                                                                                               p->accum, p->d_counters, nullptr, nullptr, nullptr, nullptr, nullptr);
    else if (p->has_principled) k_path_trace<false, PT_PRINCIPLED | PT_LOBESET><<<grid, PT_BLOCK, 0, ctx->stream>>>(S, first, n_frames, stride, mode, p->frames_done, p->rt, p->accum,
                                                                                           p->d_counters, nullptr, nullptr, nullptr, nullptr, nullptr);
    else k_path_trace<false, PT_LEAN><<<grid, PT_BLOCK, 0, ctx->stream>>>(S, first, n_frames, stride, mode, p->frames_done, p->rt, p->accum, p->d_counters,
                                                                       nullptr, nullptr, nullptr, nullptr, nullptr);
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    VB_CUDA(cudaMemcpyAsync(&p->last, p->d_counters, sizeof(PtCounters), cudaMemcpyDeviceToHost, ctx->stream));
    if (!dump) p->frames_done += n_frames;
    return VB_OK;
}

int vb_render_frames(vb_ctx *ctx, uint32_t first_frame, uint32_t n_frames, uint32_t frame_stride, int accumulate_mode) {
    if (!ctx) return VB_ERR_INVALID;
    if (n_frames == 0) return VB_OK;
    if (accumulate_mode != VB_ACCUM_LERP && accumulate_mode != VB_ACCUM_SUM) return ctx->fail(VB_ERR_INVALID, "vb_render_frames: bad mode");
    cudaSetDevice(ctx->device);
    return pt_launch(ctx, "vb_render_frames", first_frame, n_frames, frame_stride ? frame_stride : 1, accumulate_mode, false,
                     nullptr, nullptr, nullptr, nullptr, nullptr);
}

int vb_film_resolve(vb_ctx *ctx, int accumulate_mode, float exposure, int tonemap) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_film_resolve: no film");
    cudaSetDevice(ctx->device);
    uint32_t n = p->w * p->h;
    ctx->prof_begin("film", 1);
    k_film_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>(p->accum, p->output, n, accumulate_mode, exposure, tonemap);
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    return VB_OK;
}

int vb_film_download(vb_ctx *ctx, float *rgba, int which) {
    if (!ctx || !rgba) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_film_download: no film");
    cudaSetDevice(ctx->device);
    if (which == VB_FILM_VIEW && !p->view) return ctx->fail(VB_ERR_STATE, "vb_film_download: no view buffer (vb_pass_gamma)");
    const float4 *src = which == VB_FILM_RT ? p->rt : (which == VB_FILM_OUTPUT ? p->output : (which == VB_FILM_VIEW ? p->view : p->accum));
    VB_CUDA(cudaMemcpyAsync(rgba, src, sizeof(float4) * (size_t)p->w * p->h, cudaMemcpyDeviceToHost, ctx->stream));
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    return VB_OK;
}

int vb_film_device_ptr(vb_ctx *ctx, int which, void **ptr) {
    if (!ctx || !ptr) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_film_device_ptr: no film");
    if (which == VB_FILM_VIEW && !p->view) return ctx->fail(VB_ERR_STATE, "vb_film_device_ptr: no view buffer (vb_pass_gamma)");
    *ptr = which == VB_FILM_RT ? (void *)p->rt : (which == VB_FILM_OUTPUT ? (void *)p->output : (which == VB_FILM_VIEW ? (void *)p->view : (void *)p->accum));
    return VB_OK;
}

int vb_gbuffer_render(vb_ctx *ctx, uint32_t frame_index) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_gbuffer_render: no film (vb_film_create)");
    cudaSetDevice(ctx->device);
    size_t n = (size_t)p->w * p->h;
    if (p->gbuffer_n != n) {
        cudaFree(p->d_gbuffer); p->d_gbuffer = nullptr; p->gbuffer_n = 0;
        VB_CUDA(cudaMalloc(&p->d_gbuffer, sizeof(float4) * VB_AOV_COUNT * n));
        VB_CUDA(cudaMemsetAsync(p->d_gbuffer, 0, sizeof(float4) * VB_AOV_COUNT * n, ctx->stream));
        p->gbuffer_n = n; p->gbuffer_valid = false;
    }
    int rc = pt_launch(ctx, "vb_gbuffer_render", frame_index, 1, 1, VB_ACCUM_SUM, false, nullptr, nullptr, nullptr, nullptr, nullptr, p->d_gbuffer);
    if (rc == VB_OK) p->gbuffer_valid = true;
    return rc;
}

int vb_camera_prev_set(vb_ctx *ctx, const vb_camera_prev *prev) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!prev) { p->have_cam_prev = false; return VB_OK; }          // forget it: back to the static camera
    if (!(prev->tan_fov_y_over_2 > 0.f) || !(prev->screen_hi[0] > prev->screen_lo[0]) || !(prev->screen_hi[1] > prev->screen_lo[1]))
        return ctx->fail(VB_ERR_INVALID, "vb_camera_prev_set: empty field of view or screen window");
    p->cam_prev = *prev; p->have_cam_prev = true;
    return VB_OK;
}

int vb_gbuffer_gradient(vb_ctx *ctx) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    const size_t n = (size_t)p->w * p->h;
    if (!p->d_gbuffer || p->gbuffer_n != n || !p->gbuffer_valid) return ctx->fail(VB_ERR_STATE, "vb_gbuffer_gradient: no G-buffer (vb_gbuffer_render)");
    cudaSetDevice(ctx->device);
    ctx->prof_begin("gbuffer_gradient", 1);
    k_gbuffer_gradient<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(p->d_gbuffer + VB_AOV_NORMAL * n, p->d_gbuffer + VB_AOV_GEOM * n,
                                                                             p->d_gbuffer + VB_AOV_GRADIENT * n, (int)p->w, (int)p->h);
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    return VB_OK;
}

int vb_pass_tonemapping(vb_ctx *ctx, int tonemap) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_pass_tonemapping: no film (vb_film_create)");
    if (tonemap < VB_TONEMAP_LINEAR || tonemap > VB_TONEMAP_REINHARD) return ctx->fail(VB_ERR_INVALID, "vb_pass_tonemapping: unknown tone mapper");
    cudaSetDevice(ctx->device);
    const uint32_t n = p->w * p->h;
    ctx->prof_begin("film", 1);
    k_pass_tonemapping<<<(n + 255) / 256, 256, 0, ctx->stream>>>(p->accum, p->output, n, tonemap);
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    return VB_OK;
}

int vb_pass_gamma(vb_ctx *ctx) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_pass_gamma: no film (vb_film_create)");
    cudaSetDevice(ctx->device);
    const uint32_t n = p->w * p->h;
    if (!p->view) VB_CUDA(cudaMalloc(&p->view, sizeof(float4) * (size_t)n));
    ctx->prof_begin("film", 1);
    k_pass_gamma<<<(n + 255) / 256, 256, 0, ctx->stream>>>(p->output, p->view, n);
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    return VB_OK;
}

int vb_gbuffer_download(vb_ctx *ctx, int which, float *rgba) {
    if (!ctx || !rgba) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (which < 0 || which >= VB_AOV_COUNT) return ctx->fail(VB_ERR_INVALID, "vb_gbuffer_download: unknown plane");
    if (!p->d_gbuffer || p->gbuffer_n != (size_t)p->w * p->h) return ctx->fail(VB_ERR_STATE, "vb_gbuffer_download: no G-buffer (vb_gbuffer_render)");
    cudaSetDevice(ctx->device);
    VB_CUDA(cudaMemcpyAsync(rgba, p->d_gbuffer + (size_t)which * p->gbuffer_n, sizeof(float4) * p->gbuffer_n, cudaMemcpyDeviceToHost, ctx->stream));
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    return VB_OK;
}

int vb_dump_primary_hits(vb_ctx *ctx, uint32_t frame_index, const vb_hits_soa *out) {
    if (!ctx || !out) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_dump_primary_hits: no film");
    cudaSetDevice(ctx->device);
    size_t n = (size_t)p->w * p->h;
    bool host = out->mem == VB_MEM_HOST;
    uint32_t *hi = out->inst, *hp = out->prim; float *hu = out->u, *hv = out->v, *ht = out->t;
    if (host) {
        int rc = ctx->ensure_stage(5 * n * 4);
        if (rc) return rc;
        uint32_t *d = (uint32_t *)ctx->stage;
        hi = d; hp = d + n; hu = (float *)(d + 2 * n); hv = (float *)(d + 3 * n); ht = (float *)(d + 4 * n);
    }
    int rc = pt_launch(ctx, "vb_dump_primary_hits", frame_index, 1, 1, VB_ACCUM_SUM, true, hi, hp, hu, hv, ht);
    if (rc) return rc;
    if (host) {
        void *dst[5] = {out->inst, out->prim, out->u, out->v, out->t};
        const void *src[5] = {hi, hp, hu, hv, ht};
        for (int k = 0; k < 5; ++k) VB_CUDA(cudaMemcpyAsync(dst[k], src[k], n * 4, cudaMemcpyDeviceToHost, ctx->stream));
        VB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return VB_OK;
}

int vb_dump_path_trace(vb_ctx *ctx, uint32_t px, uint32_t py, uint32_t frame_index, vb_path_vertex *out, uint32_t max, uint32_t *n_vertices) {
    if (!ctx || !n_vertices || (!out && max)) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_dump_path_trace: no film");
    if (px >= p->w || py >= p->h) return ctx->fail(VB_ERR_INVALID, "vb_dump_path_trace: pixel outside the film");
    cudaSetDevice(ctx->device);
    int rc = ctx->ensure_stage(sizeof(vb_path_vertex) * (size_t)max + 16);
    if (rc) return rc;
    p->dump_out = (vb_path_vertex *)((char *)ctx->stage + 16); p->dump_n = (uint32_t *)ctx->stage;
    p->dump_max = max; p->dump_px = px; p->dump_py = py;
    rc = pt_launch(ctx, "vb_dump_path_trace", frame_index, 1, 1, VB_ACCUM_SUM, false, nullptr, nullptr, nullptr, nullptr, nullptr);
    vb_path_vertex *d_out = p->dump_out; uint32_t *d_n = p->dump_n;
    p->dump_out = nullptr; p->dump_n = nullptr;
    if (rc) return rc;
    VB_CUDA(cudaMemcpyAsync(n_vertices, d_n, 4, cudaMemcpyDeviceToHost, ctx->stream));
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint32_t n = *n_vertices < max ? *n_vertices : max;
    if (n) VB_CUDA(cudaMemcpy(out, d_out, sizeof(vb_path_vertex) * (size_t)n, cudaMemcpyDeviceToHost));
    return VB_OK;
}

int vb_render_stats(vb_ctx *ctx, uint64_t stats[4]) {
    if (!ctx || !stats) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    stats[0] = p->last.closest; stats[1] = p->last.any; stats[2] = p->last.vertices; stats[3] = p->last.culled;
    return VB_OK;
}

int vb_render_work(vb_ctx *ctx, uint64_t work[4]) {
    if (!ctx || !work) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < 4; ++k) work[k] = p->work[k];
    return VB_OK;
}

}  // extern "C"

#include "vb_pt_dist.cuh"

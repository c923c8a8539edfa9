/* vision_b200_pt.h — path-tracing part of the C ABI (included by vision_b200.h).
 *
 * The reference's plugins (Material, Light, Sensor, Filter, Sampler, Integrator, Spectrum, ToneMapper)
 * serialise their parameters into device memory through ocarina's Encodable/EncodedObject and are
 * dispatched by Polymorphic<T> inside the JIT-compiled "path_tracing" kernel
 * (reference src/base/encoded_object.cpp:22-43, src/base/mgr/scene.cpp:165-187,
 *  src/base/integral/integrator.cpp:82-107,262-268). Here each plugin fills one POD record with a
 * type tag; the megakernel switches on the tag. Array index = the id stored in vb_instance.
 */
#ifndef VISION_B200_PT_H
#define VISION_B200_PT_H

/* ---- shader-graph slot (reference src/base/shader_graph/shader_node.cpp:242-273) -------------- */
/* constant vector (NumberArray, number.cpp:106-109) or image texture sample at it.uv (image.cpp:87-97) */
typedef struct vb_slot {
    float value[4];            /* constant; when tex_id is valid the texel is multiplied by value */
    uint32_t tex_id;           /* VB_INVALID_ID = constant */
    uint32_t pad[3];
} vb_slot;

/* ---- textures: ImagePool (reference src/base/mgr/image_pool.cpp:13-35) ------------------------ */
/* float4 texels, row-major, already converted to linear when the file's colour space was sRGB.
 * Sampling: normalised uv, bilinear, wrap addressing. */
typedef struct vb_texture {
    const float *texels;       /* host pointer, w*h*4 floats */
    uint32_t width, height;
} vb_texture;
VB_API int vb_textures_set(vb_ctx *ctx, const vb_texture *tex, uint32_t n);

/* ---- materials (reference src/render_core/material/<name>.cpp) ------------------------------------- */
enum { VB_MAT_DIFFUSE = 0, VB_MAT_MIRROR = 1, VB_MAT_GLASS = 2, VB_MAT_METAL = 3, VB_MAT_SUBSTRATE = 4,
       VB_MAT_PRINCIPLED = 5, VB_MAT_BLACK_BODY = 6,
       /* plastic.cpp:12-133 (PlasticLobe: Lambert under a Schlick-Fresnel GGX coat; color = Rd, spec = Rs, ior slot, roughness slot,
        * params[0] anisotropic, flags bit1 remapping_roughness) and metallic.cpp:12-63 (MetallicLobe: GGX reflection with the F82-tint
        * Fresnel; color = F0 and reflection colour, spec = edge_tint, roughness slot, params[0] anisotropic, flags bit1) */
       VB_MAT_PLASTIC = 7, VB_MAT_METALLIC = 8,
       /* mix.cpp:52-57 / add.cpp:50-53: a LobeSet over the lobes of two CHILD materials (LobeSet::create_mix / create_add + flatten,
        * lobe.cpp:495-562). spd[0], spd[1] = indices of the children in this same table (children are ordinary records that no
        * instance refers to; they may be any single-lobe material or a principled BSDF, not another mix / add); MIX: sigma slot .x =
        * frac (weights 1 - frac, frac); ADD: both weights 1, sample weights 1/2. */
       VB_MAT_MIX = 9, VB_MAT_ADD = 10 };
typedef struct vb_material_rec {
    uint32_t type;
    uint32_t flags;            /* bit0: has sigma (Oren-Nayar, diffuse.cpp:26-29), bit1: remapping_roughness,
                                * bit2: hero spectrum only - glass ior (glass.cpp:216-234, dispersive) / metal eta, k
                                * (metal.cpp:117-121) are "spd" slots: spd[0] = ior | eta, spd[1] = k, indices into the SPD pool */
    uint32_t spd[2];
    vb_slot color;             /* diffuse.cpp:23, glass/mirror color, substrate diff */
    vb_slot roughness;         /* xy */
    vb_slot spec;              /* substrate spec / principled */
    vb_slot ior;               /* x; glass */
    vb_slot sigma;             /* diffuse Oren-Nayar */
    float eta[4], k[4];        /* metal: complex ior at the 3 sRGB peak wavelengths */
    /* [0] anisotropic (mirror/glass/metal/substrate/principled). principled (principled_bsdf.cpp:234-290; colour slot =
     * color, roughness slot = roughness, spec slot = spec_tint, ior slot = ior; the remaining slots are constants):
     * [1] metallic, [2] sheen_weight, [3] sheen_roughness, [4..6] sheen_tint, [7] coat_weight, [8] coat_roughness,
     * [9] coat_ior, [10..12] coat_tint, [13] transmission_weight, [14] sheen mode (0 volumetric, 1 approximate = default),
     * [15] lobe switches bit mask (bit order sheen, coat, metallic, transmission, specular, diffuse; 63 = all) */
    float params[16];
} vb_material_rec;
VB_API int vb_materials_set(vb_ctx *ctx, const vb_material_rec *mat, uint32_t n);

/* ---- energy-compensation LUTs: MaterialLut (reference src/base/scattering/material.cpp:14-27) ------ */
/* The reference uploads the tables of src/base/scattering/precomputed_table.h as float textures and
 * samples them with normalised coordinates; here they are plain float arrays, x fastest, sampled in
 * software (linear, clamp). PURE_REFLECTION: 32x32x1, 1 channel (lobe.cpp:716-735);
 * DIELECTRIC / DIELECTRIC_INV: 32x32x32, 2 channels (lobe.cpp:253-306). A missing table = no compensation. */
enum { VB_LUT_PURE_REFLECTION = 0, VB_LUT_DIELECTRIC = 1, VB_LUT_DIELECTRIC_INV = 2,
       /* principled BSDF: SpecularLobe / CoatLobe albedo tables (32^3 x 1, precomputed_table.h) and the sheen LTC fits
        * (32 x 32 x float4, ltc_sheen_table.h) */
       VB_LUT_SPECULAR = 3, VB_LUT_COAT = 4, VB_LUT_SHEEN_VOLUME = 5, VB_LUT_SHEEN_APPROX = 6, VB_LUT_COUNT = 7 };
VB_API int vb_luts_set(vb_ctx *ctx, int which, const float *data, uint32_t nx, uint32_t ny, uint32_t nz, uint32_t channels);

/* ---- mediums (reference src/render_core/medium/homogeneous.cpp:11-70) ------------------------------ */
/* index = the id stored in vb_instance.inside_medium / outside_medium and vb_camera.medium_id; only
 * consulted when vb_pt_params.process_mediums != 0 */
enum { VB_MEDIUM_HOMOGENEOUS = 0 };
typedef struct vb_medium_rec {
    uint32_t type;
    float g;                   /* Henyey-Greenstein g, clamped to +-0.99 by the host */
    float scale;
    float pad;
    float sigma_a[4], sigma_s[4];
} vb_medium_rec;
VB_API int vb_mediums_set(vb_ctx *ctx, const vb_medium_rec *med, uint32_t n);

/* ---- lights (reference src/render_core/light/<name>.cpp) + alias tables (warper/alias.h:86-122) ---- */
/* VB_LIGHT_DIRECTIONAL (environments/directional.cpp:20-80): a delta Infinite light; like VB_LIGHT_ENV it is the scene's environment
 * light (the last Infinite light in the list wins, lightsampler.cpp:18-25), so a ray that misses evaluates it: L = color * scale, pdf -1. */
enum { VB_LIGHT_AREA = 0, VB_LIGHT_POINT = 1, VB_LIGHT_SPOT = 2, VB_LIGHT_ENV = 3, VB_LIGHT_PROJECTOR = 4, VB_LIGHT_DIRECTIONAL = 5 };
typedef struct vb_alias_entry { float prob; uint32_t alias; } vb_alias_entry;
typedef struct vb_light_rec {
    uint32_t type;
    uint32_t inst_id;          /* area: instance that carries the emission */
    uint32_t two_sided;
    uint32_t alias_offset;     /* area: first entry of this light's table in the shared arrays */
    uint32_t alias_size;       /* area: number of triangles */
    float alias_integral;      /* sum(weights)/n, alias.h:119 */
    float scale;               /* Light::scale_ after colour normalisation (light.cpp:19-24) */
    float pad0;
    vb_slot color;
    float position[4];         /* point/spot; projector: o2w[3] */
    float direction[4];        /* spot; directional: "direction" as written in the scene file (default -1,-1,-1; not normalised, directional.cpp:33) */
    float params[8];           /* spot: cos_falloff_start, cos_total_width; projector (light/projector.cpp:33-113): [0] tan(angle_y), [1] ratio
                                * (0 in the file = image width / height), image = color slot, w2o / o2w = the light's transform; env: [0] Scene::world_diameter() (scene.h:108-109);
                                * directional: [0] world_diameter, [1] world_radius (directional.cpp:35-38: p_light = pos - direction * world_radius) */
    /* env ("spherical", environments/spherical.cpp:26-211): lat-long map = color slot texture, importance
     * sampled through AliasTable2D (alias2d.cpp:31-129): marginal table over rows at alias_offset /
     * alias_size (= env_res[1]) / alias_integral, conditional row tables (env_res[0] entries per row)
     * from cond_offset in the same shared arrays */
    uint32_t env_res[2];
    uint32_t cond_offset;
    uint32_t pad1;
    float w2o[16], o2w[16];    /* column-major; o2w = desc.o2w * rotation_x(-90) * scale(1, flip_u ? 1 : -1, 1) */
} vb_light_rec;
VB_API int vb_lights_set(vb_ctx *ctx, const vb_light_rec *lights, uint32_t n,
                         const vb_alias_entry *alias_table, const float *alias_func, uint32_t n_alias);
/* Light selection rule (reference src/render_core/lightsampler/uniform.cpp:12-34, power.cpp:11-56). UNIFORM: index = min(u N, N - 1),
 * PMF 1 / N (the default; the table arguments are ignored). POWER: an alias table over luminance(light->power()), one entry per
 * light in light order, stored in vb_lights_set's shared arrays at alias_offset (alias_size must equal the light count;
 * alias_integral = sum(weights) / N, alias.h:119): select = Warper::sample_discrete(u), PMF = Warper::PMF(index).
 * Call after vb_lights_set (which resets the rule to UNIFORM). */
enum { VB_LIGHT_SAMPLER_UNIFORM = 0, VB_LIGHT_SAMPLER_POWER = 1 };
VB_API int vb_light_sampler_set(vb_ctx *ctx, uint32_t kind, uint32_t alias_offset, uint32_t alias_size, float alias_integral);
/* LightSampler's "env_separate" / "env_prob" (reference src/base/illumination/lightsampler.cpp:11-13,28-33,159-197): when set and the
 * scene has an environment light, select_light takes that light when u < env_prob and otherwise applies the rule above to
 * u' = (u - env_prob) / (1 - env_prob) over the OTHER lights (uniform: min(u' (N-1), N-2), indices at or past the environment's
 * shifted by one; power: the caller builds the table with weight 0 for the environment, power.cpp:37-45); PMF = env_prob for the
 * environment light, (1 - env_prob) * PMF_ for the others. env_prob is used as given (the host clamps it to [0.01, 0.99] like the
 * reference's constructor); env_separate with an environment light as the ONLY light is VB_ERR_UNSUPPORTED at render time (the
 * reference divides by zero there). Call after vb_lights_set (which resets it to off). */
VB_API int vb_light_sampler_env_set(vb_ctx *ctx, uint32_t env_separate, float env_prob);

/* ---- spectrum (reference src/render_core/spectrum/srgb.cpp:16-57, hero.cpp:15-348; src/base/color/spd.cpp) ---- */
/* SPD pool: tabulated spectral distributions over 360..830 nm, SPD::eval semantics (spd.cpp:50,57-76: linear
 * interpolation with sample_interval = 471 / size). SPD i = values[offsets[i] .. offsets[i+1]). */
VB_API int vb_spds_set(vb_ctx *ctx, const float *values, const uint32_t *offsets /* n + 1 */, uint32_t n);
/* VB_SPECTRUM_HERO = HeroWavelengthSpectrum with dimension 3 (the default): hero-wavelength sampling at sampler
 * dimension -1, RGB -> sigmoid-polynomial uplifting of albedo / unbound / illumination colours through the
 * coefficient table (float[3][res][res][res][4], hero.cpp:53-54), CIE XYZ -> linear sRGB at every accumulation,
 * secondary wavelengths invalidated by dispersive glass. Selected by vb_pt_params.spectrum. */
enum { VB_SPECTRUM_SRGB = 0, VB_SPECTRUM_HERO = 1 };
typedef struct vb_spectrum_rec {
    uint32_t type;
    uint32_t dimension;        /* hero: 3 (anything else is VB_ERR_UNSUPPORTED) */
    uint32_t spd_x, spd_y, spd_z, spd_d65;   /* SPD pool indices: CIE 1931 observer and D65, 5 nm (SPD::create_cie_*, spd.cpp:89-103) */
    float cie_y_integral;      /* SPD::cie_y_integral(), spd.cpp:105-108 */
    uint32_t rgb2spec_res;     /* 64 */
    const float *rgb2spec;     /* host pointer, 3 * res^3 * 4 floats */
} vb_spectrum_rec;
VB_API int vb_spectrum_set(vb_ctx *ctx, const vb_spectrum_rec *spec);

/* ---- sensor + filter (reference src/base/sensor/sensor.cpp:44-78,153-162, thin_lens.cpp:34-42,
 *      filter/box.cpp:16-20, filter/triangle.cpp:16-18) ------------------------------------------- */
/* VB_FILTER_FITTED = FittedCurveFilter (filter/fitted_curve.h:16-107: gaussian, mitchell, sinc): pixel offsets are drawn
 * from a 20x20 AliasTable2D over |f| on the positive quadrant; its tables live in the alias arrays given to
 * vb_lights_set: 20 marginal entries at filter_table_offset, then 400 conditional entries (row-major) */
enum { VB_FILTER_BOX = 0, VB_FILTER_TRIANGLE = 1, VB_FILTER_FITTED = 2 };
#define VB_FILTER_TABLE_SIZE 20
typedef struct vb_camera {
    float c2w[16];             /* camera_to_world(), column-major */
    float tan_fov_y_over_2;
    float lens_radius, focal_distance;
    uint32_t medium_id;
    float screen_lo[2], screen_hi[2];   /* FrameBuffer::screen_window (frame_buffer.cpp:93-102) */
    uint32_t filter_type;
    float filter_radius[2];
    uint32_t filter_table_offset;   /* VB_FILTER_FITTED only */
} vb_camera;
VB_API int vb_camera_set(vb_ctx *ctx, const vb_camera *cam);

/* ---- integrator (reference src/base/integral/integrator.cpp:59-66) ---------------------------- */
typedef struct vb_pt_params {
    uint32_t max_depth, min_depth;
    float rr_threshold;
    uint32_t mis_mode;         /* 0 both, 1 light, 2 bsdf (integrator.h:116-120) */
    uint32_t spectrum;         /* VB_SPECTRUM_SRGB | VB_SPECTRUM_HERO (needs vb_spectrum_set + vb_spds_set) */
    uint32_t sampler;          /* 0 = independent (TEA + LCG) */
    uint32_t process_mediums;
    float ray_offset_factor;   /* Interaction::s_ray_offset_factor, node_desc.cpp:375 */
} vb_pt_params;
VB_API int vb_integrator_set(vb_ctx *ctx, const vb_pt_params *p);

/* ---- film (reference src/base/sensor/frame_buffer.cpp:117-144) -------------------------------- */
enum { VB_ACCUM_LERP = 0, VB_ACCUM_SUM = 1 };
enum { VB_FILM_RT = 0, VB_FILM_ACCUM = 1, VB_FILM_OUTPUT = 2 };
enum { VB_TONEMAP_LINEAR = 0, VB_TONEMAP_ACES = 1, VB_TONEMAP_REINHARD = 2 };
VB_API int vb_film_create(vb_ctx *ctx, uint32_t width, uint32_t height);
VB_API int vb_film_clear(vb_ctx *ctx);
/* renders frames first, first+stride, ... (n_frames of them): camera ray (rt_geom) + Li (path_tracing)
 * + accumulation. LERP = the reference's running mean acc = lerp(1/(k+1), acc, rt) with k the count of
 * frames accumulated so far on this context; SUM = acc += (L, 1) for multi-GPU reduction. */
VB_API int vb_render_frames(vb_ctx *ctx, uint32_t first_frame, uint32_t n_frames, uint32_t frame_stride, int accumulate_mode);
/* ACCUM -> OUTPUT: (SUM mode: divide by acc.w first) 1-exp(-x*exposure) -> tone map -> w = 1 */
VB_API int vb_film_resolve(vb_ctx *ctx, int accumulate_mode, float exposure, int tonemap);
VB_API int vb_film_download(vb_ctx *ctx, float *rgba, int which);
/* device address of a film plane (w*h float4), for the NCCL all-reduce of the SUM film across GPUs */
VB_API int vb_film_device_ptr(vb_ctx *ctx, int which, void **ptr);

/* ---- multi-GPU film (the reference's accumulation, frame_buffer.cpp:117-126, is single-GPU) -------------------
 * Partition: rank g of G renders frames g, g + G, ... with vb_render_frames(ctx, g, n, G, VB_ACCUM_SUM); the SUM film
 * (sum of radiance, w = sample count) is then made whole by ONE of the two calls below and resolved with VB_ACCUM_SUM.
 *
 * One process per GPU: vb_film_allreduce = ncclAllReduce(sum, fp32, w*h*4 floats) of the ACCUM film in place on the context's
 * stream. `nccl_comm` is a ncclComm_t whose rank lives on this context's device: the caller's own, or one made by
 * vb_nccl_comm_create (ncclGetUniqueId on rank 0 -> vb_nccl_unique_id, ship the 128 bytes to every rank by any means,
 * ncclCommInitRank -> vb_nccl_comm_create). NCCL is bound at run time: the libnccl.so.2 already loaded in the process is used,
 * else the system one (override: environment VB_NCCL_LIB); VB_ERR_UNSUPPORTED when none is found. */
#define VB_NCCL_UNIQUE_ID_BYTES 128
VB_API int vb_nccl_unique_id(void *id128);
VB_API int vb_nccl_comm_create(vb_ctx *ctx, const void *id128, int n_ranks, int rank, void **nccl_comm);
VB_API int vb_nccl_comm_destroy(vb_ctx *ctx, void *nccl_comm);
VB_API int vb_film_allreduce(vb_ctx *ctx, void *nccl_comm);
/* One process, n contexts on n different GPUs with peer access (NVLink): a fused reduce + resolve kernel per GPU. GPU g reads
 * pixel slice g of all n ACCUM films over peer memory, adds them in RANK ORDER (deterministic), and writes the sum into the
 * root's ACCUM film and the resolved picture (mean -> 1-exp(-x*exposure) -> tone map) into the root's OUTPUT film. Stream-ordered
 * after the work queued on every context's stream; the root's stream is ordered after the whole reduction. */
#define VB_MAX_PEERS 16
VB_API int vb_film_reduce_peers(vb_ctx *const *ctxs, uint32_t n, uint32_t root, float exposure, int tonemap);

/* ---- G-buffer of the camera stage (reference rt_geom, src/base/sensor/frame_buffer.cpp:156-219) ---------------------------- */
/* denoiser / AOV planes of the primary hit for ONE frame index, w*h float4 each, all colours through Spectrum::linear_srgb:
 * NORMAL (shading normal, 1) - zero where the camera ray misses; ALBEDO (MaterialEvaluator::albedo(wo), 1); EMISSION
 * (LightSampler::evaluate_hit_wi L, 1); GEOM (Sensor::linear_depth, p_film.x, p_film.y, hit ? 1 : 0); MOTION (motion vector
 * p_film - prev_raster_coord(hit position) in pixels, 0, 0; zero where the ray misses: FrameBuffer::compute_motion_vec,
 * frame_buffer.cpp:483-491). GRADIENT is written by vb_gbuffer_gradient, not by vb_gbuffer_render. */
enum { VB_AOV_NORMAL = 0, VB_AOV_ALBEDO = 1, VB_AOV_EMISSION = 2, VB_AOV_GEOM = 3, VB_AOV_MOTION = 4, VB_AOV_GRADIENT = 5, VB_AOV_COUNT = 6 };
/* The camera of the PREVIOUS frame, as Sensor::store_prev_data keeps it after every render (src/base/sensor/sensor.cpp:89-93:
 * prev_w2s_ = inverse(c2w), prev_s2r_ = inverse(raster_to_sensor)); the sensor-to-raster map is given by its parameters (this
 * library's camera is analytic: vb_camera). Until it is called - and after a call with prev = NULL - the previous camera is the
 * current one (a static camera). */
typedef struct vb_camera_prev {
    float w2c[16];             /* inverse of the previous camera_to_world(), column-major */
    float tan_fov_y_over_2;
    float screen_lo[2], screen_hi[2];
    float pad[3];
} vb_camera_prev;
VB_API int vb_camera_prev_set(vb_ctx *ctx, const vb_camera_prev *prev);
VB_API int vb_gbuffer_render(vb_ctx *ctx, uint32_t frame_index);
/* rt_gradient (FrameBuffer::compute_gradient, frame_buffer.cpp:221-275) over the planes of the last vb_gbuffer_render: per pixel the
 * mean forward / backward difference of the shading normal and of the linear depth over the 3x3 neighbourhood (clipped at the
 * picture's edge), GRADIENT = (length(|dn/dx| + |dn/dy|), |dz/dx| + |dz/dy|, 0, 0) = PixelGeometry::normal_fwidth.w, depth_gradient */
VB_API int vb_gbuffer_gradient(vb_ctx *ctx);
VB_API int vb_gbuffer_download(vb_ctx *ctx, int which, float *rgba);

/* ---- render-graph passes of the `customized` pipeline (reference src/render_core/pass/, src/pipelines/customized/pipeline.cpp:28-44) -- */
/* integrate + accumulate are vb_render_frames(frame, 1, 1, VB_ACCUM_LERP): RT = (Li, 1), ACCUM = lerp(1 / (frame + 1), ACCUM, RT).
 * tonemapping (pass/tonemapping.cpp:26-33): OUTPUT = (tone_mapper(ACCUM.rgb), 1) - no exposure curve, unlike vb_film_resolve.
 * gamma (pass/gamma.cpp:21-27): VIEW = linear_to_srgb(OUTPUT) on all four channels; VB_FILM_VIEW is the graph's output buffer. */
#define VB_FILM_VIEW 3
VB_API int vb_pass_tonemapping(vb_ctx *ctx, int tonemap);
VB_API int vb_pass_gamma(vb_ctx *ctx);

/* ---- debug / parity surface ------------------------------------------------------------------- */
/* primary hit of every pixel for one frame index (== reference rt_compute_pixel, frame_buffer.cpp:277-291) */
VB_API int vb_dump_primary_hits(vb_ctx *ctx, uint32_t frame_index, const vb_hits_soa *out);
/* Per-vertex dump of ONE path (the first place to look when a film differs from the oracle's): pixel (px, py) of frame
 * `frame_index` is traced by the megakernel loop and one record is written per closest-hit ray of IlluminationIntegrator::Li
 * (reference src/base/integral/integrator.cpp:160-311, its only_direct supplement included) with the path state BEFORE that vertex
 * is shaded; a last record with inst = prim = VB_PATH_END carries the final throughput and radiance (= the pixel-sample the film gets).
 * `out` is host memory for `max` records; *n_vertices = records the path produced (all of them written only if <= max). */
#define VB_PATH_END 0xfffffffeu
typedef struct vb_path_vertex {
    float ray_o[3], ray_tmin, ray_d[3], ray_tmax;      /* the ray as traced */
    uint32_t inst, prim; float u, v;                   /* its closest hit; VB_INVALID_ID = miss */
    float t; uint32_t bounces, sampler_state, medium;  /* hit distance; Li's `bounces`; LCG state and current medium id before the vertex */
    float throughput[3], scatter_pdf;
    float radiance[3], eta_scale;                      /* `ret` so far (shadow rays of earlier vertices resolved) */
} vb_path_vertex;
VB_API int vb_dump_path_trace(vb_ctx *ctx, uint32_t px, uint32_t py, uint32_t frame_index, vb_path_vertex *out, uint32_t max, uint32_t *n_vertices);
/* counters of the last vb_render_frames call: [0] closest-hit rays, [1] any-hit rays, [2] shaded vertices,
 * [3] shadow rays culled before tracing because their contribution is exactly zero (the reference traces them:
 * [1] + [3] is its any-hit count) */
VB_API int vb_render_stats(vb_ctx *ctx, uint64_t stats[4]);
/* traversal work of the last vb_render_frames call made with option "pt.count_work" = 1 (instrumented kernels,
 * never timed): [0] wide nodes popped and [1] triangles tested by closest-hit rays, [2], [3] the same for any-hit rays */
VB_API int vb_render_work(vb_ctx *ctx, uint64_t work[4]);

#endif

#include "vh_scene.h"
#include <chrono>
#include <thread>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <dlfcn.h>

namespace vh {

// ---- images ------------------------------------------------------------------------------------
// sRGB EOTF of a texel, evaluated in double precision and rounded once (the Python twin does the same, so host and twin textures are
// bit-identical; float powf differs between libm and numpy by an ulp on a few of the 256 levels)
static float srgb_to_linear(float c) { const double x = (double)c; return (float)(x <= 0.04045 ? x / 12.92 : std::pow((x + 0.055) / 1.055, 2.4)); }
static float linear_to_srgb(float c) { c = std::min(1.f, std::max(0.f, c)); return c <= 0.0031308f ? 12.92f * c : 1.055f * std::pow(c, 1.f / 2.4f) - 0.055f; }

static bool read_pnm_token(std::istream &f, std::string &tok) {
    tok.clear();
    int c;
    while ((c = f.get()) != EOF) {
        if (c == '#') { while ((c = f.get()) != EOF && c != '\n') {} continue; }
        if (!isspace(c)) { tok.push_back((char)c); break; }
    }
    while ((c = f.get()) != EOF && !isspace(c)) tok.push_back((char)c);
    return !tok.empty();
}

Image load_image(const std::string &path, bool to_linear) {
    std::string p = path;
    auto ends = [&](const char *e) { size_t n = strlen(e); return p.size() >= n && strcasecmp(p.c_str() + p.size() - n, e) == 0; };
    if (ends(".hdr") && std::ifstream(path, std::ios::binary)) return load_hdr(path);      // native reader; linear by definition (image_pool.cpp:18-24)
    if (ends(".jpg") || ends(".jpeg") || ends(".png") || ends(".exr")) {                    // vh_codecs.cpp
        Image im = decode_image_file(path);
        if (to_linear && !ends(".exr"))
            for (size_t i = 0; i < (size_t)im.w * im.h; ++i) for (int c = 0; c < 3; ++c) im.rgba[4 * i + c] = srgb_to_linear(im.rgba[4 * i + c]);
        return im;
    }
    if (!ends(".pfm") && !ends(".ppm")) throw std::runtime_error("unsupported image format " + path + " (.jpg, .png, .exr, .hdr, .pfm, .ppm)");
    std::ifstream f(p, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open image " + path);
    std::string magic, tw, th, ts;
    read_pnm_token(f, magic); read_pnm_token(f, tw); read_pnm_token(f, th); read_pnm_token(f, ts);
    Image im; im.w = (uint32_t)std::stoul(tw); im.h = (uint32_t)std::stoul(th);
    im.rgba.resize((size_t)im.w * im.h * 4);
    if (magic == "PF" || magic == "Pf") {
        int ch = magic == "PF" ? 3 : 1;
        std::vector<float> raw((size_t)im.w * im.h * ch);
        f.read((char *)raw.data(), raw.size() * 4);
        for (uint32_t y = 0; y < im.h; ++y)                         // PFM rows are bottom-up
            for (uint32_t x = 0; x < im.w; ++x) {
                const float *s = &raw[((size_t)(im.h - 1 - y) * im.w + x) * ch];
                float *d = &im.rgba[((size_t)y * im.w + x) * 4];
                d[0] = s[0]; d[1] = s[ch == 3 ? 1 : 0]; d[2] = s[ch == 3 ? 2 : 0]; d[3] = 1.f;
            }
    } else if (magic == "P6") {
        std::vector<unsigned char> raw((size_t)im.w * im.h * 3);
        f.read((char *)raw.data(), raw.size());
        for (size_t i = 0; i < (size_t)im.w * im.h; ++i) { for (int c = 0; c < 3; ++c) im.rgba[4 * i + c] = raw[3 * i + c] / 255.f; im.rgba[4 * i + 3] = 1.f; }
    } else throw std::runtime_error("unsupported image format " + p);
    if (to_linear) for (size_t i = 0; i < (size_t)im.w * im.h; ++i) for (int c = 0; c < 3; ++c) im.rgba[4 * i + c] = srgb_to_linear(im.rgba[4 * i + c]);
    return im;
}
// writers use C stdio (no iostream number formatting: no locale dependence when loaded into a foreign process)
void save_pfm(const std::string &path, const float *rgba, uint32_t w, uint32_t h) {
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    std::fprintf(f, "PF\n%u %u\n-1.0\n", w, h);
    for (uint32_t y = 0; y < h; ++y) for (uint32_t x = 0; x < w; ++x) std::fwrite(&rgba[((size_t)(h - 1 - y) * w + x) * 4], 4, 3, f);
    std::fclose(f);
}
void save_ppm(const std::string &path, const float *rgba01, uint32_t w, uint32_t h) {
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    std::fprintf(f, "P6\n%u %u\n255\n", w, h);
    std::vector<unsigned char> row((size_t)w * 3);
    for (uint32_t y = 0; y < h; ++y) {
        for (size_t i = 0; i < (size_t)w; ++i) for (int c = 0; c < 3; ++c) row[3 * i + c] = (unsigned char)(linear_to_srgb(rgba01[4 * ((size_t)y * w + i) + c]) * 255.f + 0.5f);
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
}
// reference image_pool.cpp:13-35: .exr/.hdr are linear, everything else follows color_space (default srgb)
uint32_t ImagePool::obtain(const std::string &path, const std::string &color_space) {
    std::string key = path + "|" + color_space;
    for (uint32_t i = 0; i < keys.size(); ++i) if (keys[i] == key) return i;
    auto ends = [&](const char *e) { size_t n = strlen(e); return path.size() >= n && strcasecmp(path.c_str() + path.size() - n, e) == 0; };
    bool to_linear = color_space == "srgb" && !ends(".exr") && !ends(".hdr");
    images.push_back(load_image(path, to_linear));
    keys.push_back(key);
    return (uint32_t)images.size() - 1;
}

// directory of libvision_host.so (or of the executable that linked the sources statically) + "/data"
std::string data_dir() {
    if (const char *e = getenv("VISION_B200_DATA")) return e;
    Dl_info info;
    std::string p = dladdr((void *)&data_dir, &info) && info.dli_fname ? info.dli_fname : ".";
    size_t s = p.find_last_of('/');
    return (s == std::string::npos ? std::string(".") : p.substr(0, s)) + "/data";
}
static void load_lut(SceneRecords::Lut &l, const std::string &file, uint32_t nx, uint32_t ny, uint32_t nz, uint32_t ch) {
    std::ifstream f(data_dir() + "/" + file, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + data_dir() + "/" + file + " (tools/extract_luts.py writes it)");
    l.nx = nx; l.ny = ny; l.nz = nz; l.channels = ch;
    l.data.resize((size_t)nx * ny * nz * ch);
    f.read((char *)l.data.data(), l.data.size() * 4);
    if ((size_t)f.gcount() != l.data.size() * 4) throw std::runtime_error("short read: " + file);
}

// ---- scene -------------------------------------------------------------------------------------
std::shared_ptr<Material> Scene::obtain_black_body() {
    if (!black_body) { MaterialDesc d; d.sub_type = "black_body"; d.name = "__black_body"; black_body = create_shared<Material>(d); materials.push_back(black_body); }
    return black_body;
}

// 2 * max(radius of the world box, render_setting.min_world_radius); the box is accumulated in double over the
// vertices the triangles use (vision_b200/scene_py.py world_diameter is the twin)
float Scene::world_diameter() const {
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    bool any = false;
    for (auto &i : instances) {
        const MeshData &m = *i->mesh;
        std::vector<char> used(m.nv(), 0);
        for (uint32_t v : m.tri) used[v] = 1;
        for (uint32_t v = 0; v < m.nv(); ++v) {
            if (!used[v]) continue;
            double p[3] = {m.pos[3 * v], m.pos[3 * v + 1], m.pos[3 * v + 2]};
            for (int r = 0; r < 3; ++r) {
                double q = p[0] * (double)i->o2w[r] + p[1] * (double)i->o2w[4 + r] + p[2] * (double)i->o2w[8 + r] + (double)i->o2w[12 + r];
                lo[r] = std::min(lo[r], q); hi[r] = std::max(hi[r], q);
            }
            any = true;
        }
    }
    double radius = 0.0;
    if (any) { double d[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]}; radius = 0.5 * std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); }
    return (float)(2.0 * std::max(radius, (double)desc.render_setting.min_world_radius));
}

// reference scene.cpp:16-34: light_sampler, spectrum, materials, mediums, sensor, shapes, integrator, sampler
void Scene::init(const SceneDesc &d) {
    desc = d;
    warper = create_shared<Warper>(d.warper_desc);
    light_sampler = create_shared<LightSampler>(d.light_sampler_desc);
    for (const LightDesc &ld : d.light_sampler_desc.light_descs) light_sampler->add_light(create_shared<Light>(ld));
    spectrum = create_shared<Spectrum>(d.spectrum_desc);
    for (const MaterialDesc &md : d.material_descs) materials.push_back(create_shared<Material>(md));
    // MediumRegistry::load_mediums (registries.cpp:196-210)
    process_mediums = d.mediums_desc.process;
    for (const MediumDesc &md : d.mediums_desc.mediums) { auto m = create_shared<Medium>(md); m->index = (uint32_t)mediums.size(); mediums.push_back(m); }
    for (auto &m : mediums) if (m->name() == d.mediums_desc.global) { global_medium = m; break; }
    frame_buffer = create_shared<FrameBuffer>(d.pipeline_desc.frame_buffer_desc);
    sensor = create_shared<Sensor>(d.sensor_desc);
    for (const ShapeDesc &sd : d.shape_descs) add_shape(create_shared<ShapeGroup>(sd), sd);
    integrator = create_shared<Integrator>(d.integrator_desc);
    sampler = create_shared<Sampler>(d.sampler_desc);
    for (auto &k : d.ignored_keys) log.push_back("scene key ignored on this path: " + k);
}

// reference scene.cpp:118-151
void Scene::add_shape(const std::shared_ptr<ShapeGroup> &group, const ShapeDesc &sd) {
    groups.push_back(group);
    for (auto &inst : group->instances) {
        auto it = std::find_if(materials.begin(), materials.end(), [&](const std::shared_ptr<Material> &m) { return m->name() == inst->material_name; });
        if (it != materials.end() && !inst->material) inst->material = *it;
        if (sd.has_emission && sd.emission.valid()) {
            LightDesc ld = sd.emission;
            ld.parameter.set_value("inst_id", Json::number((double)instances.size()));
            auto light = create_shared<Light>(ld);
            light_sampler->add_light(light);
            inst->emission = light;
        }
        instances.push_back(inst);
    }
}

// reference scene.cpp:79-91 (remove unused materials -> tidy_up -> fill_instances -> prepare) and
// geometry.cpp:20-34 (mesh handles / instance table); MeshRegistry dedupes meshes by content hash
void Scene::prepare(SceneRecords &out) {
    out = SceneRecords();
    // remove_unused_materials + index assignment
    std::vector<std::shared_ptr<Material>> used;
    for (auto &m : materials) {
        bool u = false;
        for (auto &i : instances) u |= i->material == m;
        if (u) { m->index = (uint32_t)used.size(); used.push_back(m); }
    }
    materials = used;
    light_sampler->tidy_up();
    spectrum->prepare(out);
    SlotContext sctx{&image_pool, desc.scene_path, spectrum->is_complete(), &out.spds};
    std::vector<std::shared_ptr<Material>> all = materials;      // + the children of mix / add materials, appended after the scene's own
    for (size_t k = 0; k < materials.size(); ++k)
        for (auto &c : materials[k]->children()) {
            if (!c->children().empty()) throw std::runtime_error("material '" + materials[k]->name() + "': a mix / add material nested in another one is not supported");
            c->index = (uint32_t)all.size(); all.push_back(c);
        }
    out.materials.resize(all.size());
    for (auto &m : all) m->fill_record(out.materials[m->index], sctx);
    std::map<uint64_t, uint32_t> mesh_ids;
    for (auto &i : instances) {
        bool known_name = false;
        for (const MaterialDesc &md : desc.material_descs) known_name |= md.name == i->material_name;
        if (!i->material && known_name) throw std::runtime_error("instance " + i->name + ": material '" + i->material_name + "' not available");
        uint64_t h = i->mesh->hash();
        auto f = mesh_ids.find(h);
        if (f == mesh_ids.end()) { f = mesh_ids.emplace(h, (uint32_t)out.meshes.size()).first; out.meshes.push_back(i->mesh); }
        vb_instance r{};
        r.mesh_id = f->second; r.material_id = i->material ? i->material->index : VB_INVALID_ID;     // no material: medium boundary only
        r.light_id = i->emission ? i->emission->index : VB_INVALID_ID;
        r.lightmap_id = VB_INVALID_ID;
        auto medium_index = [&](const std::string &nm) { for (auto &m : mediums) if (m->name() == nm) return m->index; return (uint32_t)VB_INVALID_ID; };
        if (i->has_medium) { r.inside_medium = medium_index(i->inside_name); r.outside_medium = medium_index(i->outside_name); }   // shape.cpp:247-258
        else r.inside_medium = r.outside_medium = global_medium ? global_medium->index : VB_INVALID_ID;                         // shape.cpp:259-268
        std::memcpy(r.o2w, i->o2w.data(), 64);
        out.instances.push_back(r);
    }
    out.lights.resize(light_sampler->lights.size());
    for (auto &l : light_sampler->lights) l->fill_record(out.lights[l->index], *this, sctx, out.alias_table, out.alias_func);
    out.textures = image_pool.images;
    if (light_sampler->kind() == VB_LIGHT_SAMPLER_POWER && !out.lights.empty()) {      // PowerLightSampler::prepare (power.cpp:33-53)
        std::vector<float> weights(out.lights.size());
        for (auto &l : light_sampler->lights) {
            const vb_light_rec &r = out.lights[l->index];
            const bool infinite = r.type == VB_LIGHT_ENV || r.type == VB_LIGHT_DIRECTIONAL;
            weights[l->index] = (light_sampler->env_separate && infinite) ? 0.f                       // power.cpp:37-45
                                                                         : light_power_luminance(r, out, l->shape_angle, l->shape_falloff);
        }
        out.light_sampler.kind = VB_LIGHT_SAMPLER_POWER;
        out.light_sampler.alias_offset = (uint32_t)out.alias_table.size(); out.light_sampler.alias_size = (uint32_t)weights.size();
        warper->build(weights, out.alias_table, out.alias_func, out.light_sampler.alias_integral);
    }
    out.light_sampler.env_separate = light_sampler->env_separate ? 1u : 0u;
    out.light_sampler.env_prob = light_sampler->env_prob;
    out.mediums.resize(mediums.size());
    for (auto &m : mediums) m->fill_record(out.mediums[m->index]);
    bool need_luts = false, principled = false;
    for (auto &m : out.materials) { principled |= m.type == VB_MAT_PRINCIPLED;
                                    need_luts |= m.type == VB_MAT_MIRROR || m.type == VB_MAT_GLASS || m.type == VB_MAT_METAL || m.type == VB_MAT_METALLIC; }
    if (principled) {            // PrincipledBSDF::prepare (principled_bsdf.cpp:324-330): coat, sheen, specular tables
        load_lut(out.luts[VB_LUT_SPECULAR], "lut_specular_32x32x32x1.f32", 32, 32, 32, 1);
        load_lut(out.luts[VB_LUT_COAT], "lut_coat_32x32x32x1.f32", 32, 32, 32, 1);
        load_lut(out.luts[VB_LUT_SHEEN_VOLUME], "lut_sheen_volume_32x32x1x4.f32", 32, 32, 1, 4);
        load_lut(out.luts[VB_LUT_SHEEN_APPROX], "lut_sheen_approx_32x32x1x4.f32", 32, 32, 1, 4);
    }
    if (need_luts || principled) {             // PureReflectionLobe::prepare / DielectricLobe::prepare (lobe.cpp:253-260,731-735)
        load_lut(out.luts[VB_LUT_PURE_REFLECTION], "lut_pure_reflection_32x32x1x1.f32", 32, 32, 1, 1);
        load_lut(out.luts[VB_LUT_DIELECTRIC], "lut_dielectric_32x32x32x2.f32", 32, 32, 32, 2);
        load_lut(out.luts[VB_LUT_DIELECTRIC_INV], "lut_dielectric_inv_32x32x32x2.f32", 32, 32, 32, 2);
    }
    out.width = frame_buffer->width; out.height = frame_buffer->height;
    sensor->fill_record(out.camera, out.width, out.height);
    if (out.camera.filter_type == VB_FILTER_FITTED) {       // FilterSampler: AliasTable2D over the 20x20 table, after the lights' tables
        const uint32_t N = VB_FILTER_TABLE_SIZE;
        std::vector<float> w = sensor->filter->table_weights(), marginal, cf;
        std::vector<vb_alias_entry> ct;
        for (uint32_t v = 0; v < N; ++v) {
            std::vector<float> row(w.begin() + (size_t)v * N, w.begin() + (size_t)(v + 1) * N);
            bool any = false;
            for (float x : row) any |= x != 0.f;
            float integral = 0.f;
            if (any) warper->build(row, ct, cf, integral);
            else for (uint32_t i = 0; i < N; ++i) { ct.push_back({1.0f, i}); cf.push_back(0.f); }
            marginal.push_back(integral);
        }
        out.camera.filter_table_offset = (uint32_t)out.alias_table.size();
        float mi;
        warper->build(marginal, out.alias_table, out.alias_func, mi);
        out.alias_table.insert(out.alias_table.end(), ct.begin(), ct.end()); out.alias_func.insert(out.alias_func.end(), cf.begin(), cf.end());
    }
    // the flag only matters when some medium exists (all ids are invalid otherwise): keep the kernel's fast path
    bool mediums_on = process_mediums && !mediums.empty();
    out.camera.medium_id = VB_INVALID_ID;               // Photosensory ctor, photosensory.cpp:15-33
    if (mediums_on) {
        if (sensor->has_medium_name) { for (auto &m : mediums) if (m->name() == sensor->medium_name) out.camera.medium_id = m->index; }
        else if (global_medium) out.camera.medium_id = global_medium->index;
    }
    out.params = vb_pt_params{};
    out.params.max_depth = integrator->max_depth; out.params.min_depth = integrator->min_depth;
    out.params.rr_threshold = integrator->rr_threshold; out.params.mis_mode = integrator->mis_mode;
    out.params.spectrum = spectrum->device_tag(); out.params.sampler = sampler->device_tag();
    out.params.process_mediums = mediums_on ? 1u : 0u; out.params.ray_offset_factor = desc.render_setting.ray_offset_factor;
    out.spp = desc.output_desc.spp ? desc.output_desc.spp : sampler->spp;
    out.exposure = frame_buffer->exposure; out.tonemap = frame_buffer->tone_mapper->device_tag();
}

// ---- pipeline ----------------------------------------------------------------------------------
void Pipeline::check(int rc, const char *what) const {
    if (rc != VB_OK) throw std::runtime_error(std::string(what) + " failed: " + vb_last_error(ctx));
}
Pipeline::~Pipeline() {
    if (ctx && nccl_comm) vb_nccl_comm_destroy(ctx, nccl_comm);
    if (ctx) vb_ctx_destroy(ctx);
}
void Pipeline::set_partition(uint32_t rank, uint32_t world) {
    if (world == 0 || rank >= world) throw std::runtime_error("Pipeline::set_partition: rank out of range");
    if (scene.integrator && scene.integrator->frame_index != 0) throw std::runtime_error("Pipeline::set_partition: frames already rendered");
    dist_rank = rank; dist_world = world;
}
void Pipeline::nccl_init(const void *id128) {
    if (!ctx) throw std::runtime_error("Pipeline::nccl_init: prepare() first");
    check(vb_nccl_comm_create(ctx, id128, (int)dist_world, (int)dist_rank, &nccl_comm), "vb_nccl_comm_create");
}
void Pipeline::init_scene(const SceneDesc &d) { scene.init(d); }

// FixedRenderPipeline::prepare (fixed/pipeline.cpp:14-23) + Pipeline::prepare_geometry (pipeline.cpp:272-277)
void Pipeline::prepare(int device) {
    scene.prepare(records);
    if (vb_ctx_create(device, &ctx) != VB_OK) throw std::runtime_error(std::string("vb_ctx_create failed: ") + vb_last_error(nullptr));
    for (auto &kv : options) check(vb_ctx_set_option(ctx, kv.first.c_str(), kv.second), ("vb_ctx_set_option " + kv.first).c_str());
    std::vector<uint32_t> ids;
    for (auto &m : records.meshes) {
        uint32_t id;
        check(vb_mesh_upload(ctx, m->pos.data(), m->nrm.data(), m->uv.data(), m->nv(), m->tri.data(), m->nt(), &id), "vb_mesh_upload");
        ids.push_back(id);
    }
    std::vector<vb_instance> inst = records.instances;
    for (auto &i : inst) i.mesh_id = ids[i.mesh_id];
    check(vb_instances_set(ctx, inst.data(), (uint32_t)inst.size()), "vb_instances_set");
    check(vb_accel_build(ctx, nullptr, &build_stats), "vb_accel_build");
    std::vector<vb_texture> tex;
    for (auto &t : records.textures) tex.push_back(vb_texture{t.rgba.data(), t.w, t.h});
    check(vb_textures_set(ctx, tex.data(), (uint32_t)tex.size()), "vb_textures_set");
    check(vb_materials_set(ctx, records.materials.data(), (uint32_t)records.materials.size()), "vb_materials_set");
    for (int k = 0; k < VB_LUT_COUNT; ++k) {
        const SceneRecords::Lut &l = records.luts[k];
        check(vb_luts_set(ctx, k, l.data.empty() ? nullptr : l.data.data(), l.nx, l.ny, l.nz, l.channels), "vb_luts_set");
    }
    check(vb_mediums_set(ctx, records.mediums.data(), (uint32_t)records.mediums.size()), "vb_mediums_set");
    {
        std::vector<float> values; std::vector<uint32_t> offsets{0};
        for (auto &s : records.spds) { values.insert(values.end(), s.begin(), s.end()); offsets.push_back((uint32_t)values.size()); }
        check(vb_spds_set(ctx, values.data(), offsets.data(), (uint32_t)records.spds.size()), "vb_spds_set");
        vb_spectrum_rec sp = records.spectrum;
        sp.rgb2spec = records.rgb2spec.empty() ? nullptr : records.rgb2spec.data();
        check(vb_spectrum_set(ctx, &sp), "vb_spectrum_set");
    }
    check(vb_lights_set(ctx, records.lights.data(), (uint32_t)records.lights.size(), records.alias_table.data(), records.alias_func.data(),
                        (uint32_t)records.alias_table.size()), "vb_lights_set");
    check(vb_light_sampler_set(ctx, records.light_sampler.kind, records.light_sampler.alias_offset, records.light_sampler.alias_size,
                               records.light_sampler.alias_integral), "vb_light_sampler_set");
    check(vb_light_sampler_env_set(ctx, records.light_sampler.env_separate, records.light_sampler.env_prob), "vb_light_sampler_env_set");
    check(vb_camera_set(ctx, &records.camera), "vb_camera_set");
    check(vb_integrator_set(ctx, &records.params), "vb_integrator_set");
    check(vb_film_create(ctx, records.width, records.height), "vb_film_create");
}
void Pipeline::render(uint32_t n_frames) { scene.integrator->render(*this, n_frames); after_render(); }
// Sensor::store_prev_data: the camera this frame was rendered with becomes the "previous camera" of the next G-buffer's motion vectors
void Pipeline::after_render() {
    vb_camera_prev prev{};
    Mat4 c2w; std::memcpy(c2w.data(), records.camera.c2w, 64);
    Mat4 w2c = mat4_inverse(c2w);
    std::memcpy(prev.w2c, w2c.data(), 64);
    prev.tan_fov_y_over_2 = records.camera.tan_fov_y_over_2;
    for (int k = 0; k < 2; ++k) { prev.screen_lo[k] = records.camera.screen_lo[k]; prev.screen_hi[k] = records.camera.screen_hi[k]; }
    check(vb_camera_prev_set(ctx, &prev), "vb_camera_prev_set");
}
void Pipeline::final_picture(std::vector<float> &rgba) {
    if (dist_world > 1) {
        if (!nccl_comm) throw std::runtime_error("Pipeline::final_picture: partitioned render without a communicator (nccl_init)");
        check(vb_film_allreduce(ctx, nccl_comm), "vb_film_allreduce");
    }
    check(vb_film_resolve(ctx, accumulate_mode(), records.exposure, records.tonemap), "vb_film_resolve");
    rgba.resize((size_t)records.width * records.height * 4);
    check(vb_film_download(ctx, rgba.data(), VB_FILM_OUTPUT), "vb_film_download");
}

std::shared_ptr<Pipeline> render_on_gpus(const std::string &scene_path, uint32_t n_gpus, uint32_t spp, bool use_nccl,
                                         std::vector<float> &rgba, MultiGpuTimes *times, int first_device) {
    if (n_gpus == 0 || n_gpus > VB_MAX_PEERS) throw std::runtime_error("render_on_gpus: bad GPU count");
    using clk = std::chrono::steady_clock;
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    std::vector<std::shared_ptr<Pipeline>> pipes(n_gpus);
    std::vector<std::string> errors(n_gpus);
    unsigned char id[VB_NCCL_UNIQUE_ID_BYTES] = {};
    if (use_nccl && n_gpus > 1 && vb_nccl_unique_id(id) != VB_OK) throw std::runtime_error("render_on_gpus: NCCL unavailable");
    auto each = [&](auto fn) {                      // one host thread per GPU (contexts are independent; NCCL init / collectives need all ranks)
        std::vector<std::thread> th;
        for (uint32_t g = 0; g < n_gpus; ++g)
            th.emplace_back([&, g] { try { fn(g); } catch (const std::exception &e) { errors[g] = e.what(); } });
        for (auto &t : th) t.join();
        for (uint32_t g = 0; g < n_gpus; ++g)
            if (!errors[g].empty()) throw std::runtime_error("GPU " + std::to_string(g) + ": " + errors[g]);
    };
    auto t0 = clk::now();
    each([&](uint32_t g) {
        pipes[g] = import_scene(scene_path);
        pipes[g]->set_partition(g, n_gpus);
        pipes[g]->prepare(first_device + (int)g);
        if (use_nccl && n_gpus > 1) pipes[g]->nccl_init(id);
    });
    uint32_t n = spp ? spp : pipes[0]->records.spp;
    auto t1 = clk::now();
    each([&](uint32_t g) { pipes[g]->render(n); pipes[g]->check(vb_ctx_synchronize(pipes[g]->ctx), "vb_ctx_synchronize"); });
    auto t2 = clk::now();
    Pipeline &root = *pipes[0];
    if (n_gpus == 1 || use_nccl) {
        std::vector<std::vector<float>> pics(n_gpus);
        each([&](uint32_t g) { pipes[g]->final_picture(pics[g]); });
        rgba = std::move(pics[0]);
    } else {
        std::vector<vb_ctx *> ctxs;
        for (auto &p : pipes) ctxs.push_back(p->ctx);
        root.check(vb_film_reduce_peers(ctxs.data(), n_gpus, 0, root.records.exposure, root.records.tonemap), "vb_film_reduce_peers");
        rgba.resize((size_t)root.records.width * root.records.height * 4);
        root.check(vb_film_download(root.ctx, rgba.data(), VB_FILM_OUTPUT), "vb_film_download");
    }
    auto t3 = clk::now();
    if (times) { times->prepare_ms = ms(t0, t1); times->render_ms = ms(t1, t2); times->reduce_ms = ms(t2, t3); }
    return pipes[0];                                // the other pipelines (contexts, communicators) are released here
}

std::shared_ptr<Pipeline> import_scene(const std::string &path) {
    SceneDesc sd = SceneDesc::from_json(path);
    auto p = create_shared<Pipeline>(sd.pipeline_desc);
    p->init_scene(sd);
    return p;
}

}  // namespace vh

// extern "C" surface of the host layer (libvision_host.so) for the Python harness and the CLI.
#include "vh_scene.h"
#include "vh_graph.h"
#include <cstring>

using namespace vh;

static thread_local std::string g_err;
struct vh_pipeline { std::shared_ptr<Pipeline> p; };

#define VH_TRY(body) try { body; return 0; } catch (const std::exception &e) { g_err = e.what(); return 1; }

extern "C" {

VB_API const char *vh_last_error() { return g_err.c_str(); }

VB_API vh_pipeline *vh_import_scene(const char *path) {
    try { auto h = new vh_pipeline{import_scene(path)}; return h; }
    catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}
VB_API void vh_destroy(vh_pipeline *h) { delete h; }

/* optional overrides before prepare (the CLI/bench knobs): key = "resolution" (a,b), "max_depth", "min_depth", "spp" */
VB_API int vh_override(vh_pipeline *h, const char *key, double a, double b) {
    VH_TRY({
        Scene &s = h->p->scene; std::string k(key);
        if (k == "resolution") { s.frame_buffer->width = (uint32_t)a; s.frame_buffer->height = (uint32_t)b; }
        else if (k == "max_depth") s.integrator->max_depth = (uint32_t)a;
        else if (k == "min_depth") s.integrator->min_depth = (uint32_t)a;
        else if (k == "spp") s.desc.output_desc.spp = (uint32_t)a;
        else throw std::runtime_error("unknown override " + k);
    })
}
/* scene preparation only (records, no GPU) */
VB_API int vh_prepare_records(vh_pipeline *h) { VH_TRY(h->p->scene.prepare(h->p->records)) }
/* full prepare: records + upload through the C ABI + BVH build */
VB_API int vh_prepare(vh_pipeline *h, int device) { VH_TRY(h->p->prepare(device)) }
VB_API int vh_render(vh_pipeline *h, uint32_t n_frames) { VH_TRY(h->p->render(n_frames)) }
VB_API int vh_final_picture(vh_pipeline *h, float *rgba) {
    VH_TRY({ std::vector<float> v; h->p->final_picture(v); std::memcpy(rgba, v.data(), v.size() * 4); })
}
VB_API int vh_download_accum(vh_pipeline *h, float *rgba) { VH_TRY(h->p->check(vb_film_download(h->p->ctx, rgba, VB_FILM_ACCUM), "vb_film_download")) }
VB_API void *vh_ctx(vh_pipeline *h) { return h->p->ctx; }
VB_API void vh_build_stats(vh_pipeline *h, vb_build_stats *out) { *out = h->p->build_stats; }
/* multi-GPU, one process per GPU: vh_set_partition before vh_prepare; then vh_nccl_init with the 128-byte id rank 0 got from
 * vb_nccl_unique_id (every rank calls it; it blocks until all have). vh_final_picture then all-reduces the SUM film first. */
VB_API int vh_set_partition(vh_pipeline *h, uint32_t rank, uint32_t world) { VH_TRY(h->p->set_partition(rank, world)) }
VB_API int vh_nccl_init(vh_pipeline *h, const void *id128) { VH_TRY(h->p->nccl_init(id128)) }
/* the all-reduce alone (what vh_final_picture does first), so that a harness can time it by itself */
VB_API int vh_film_allreduce(vh_pipeline *h) { VH_TRY(h->p->check(vb_film_allreduce(h->p->ctx, h->p->nccl_comm), "vb_film_allreduce")) }
/* start over: empty film, frame index 0 (Pipeline::invalidate, pipeline.h:96: frame_index_ = 0) */
VB_API int vh_invalidate(vh_pipeline *h) {
    VH_TRY({ h->p->check(vb_film_clear(h->p->ctx), "vb_film_clear"); h->p->scene.integrator->frame_index = 0; })
}
VB_API int vh_synchronize(vh_pipeline *h) { VH_TRY(h->p->check(vb_ctx_synchronize(h->p->ctx), "vb_ctx_synchronize")) }
/* one process, n GPUs: see vh::render_on_gpus. times_ms[3] = prepare, render, reduce + resolve + download (may be null) */
VB_API vh_pipeline *vh_render_on_gpus(const char *path, uint32_t n_gpus, uint32_t spp, int use_nccl, float *rgba, uint64_t cap_floats, double *times_ms) {
    try {
        std::vector<float> pic; MultiGpuTimes t;
        auto p = render_on_gpus(path, n_gpus, spp, use_nccl != 0, pic, &t);
        if (rgba && cap_floats >= pic.size()) std::memcpy(rgba, pic.data(), pic.size() * 4);
        if (times_ms) { times_ms[0] = t.prepare_ms; times_ms[1] = t.render_ms; times_ms[2] = t.reduce_ms; }
        return new vh_pipeline{p};
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}

/* info[0..9] = width, height, spp, n_meshes, n_instances, n_materials, n_lights, n_alias, n_textures, tonemap */
VB_API void vh_info(vh_pipeline *h, uint32_t *info, float *exposure) {
    const SceneRecords &r = h->p->records;
    uint32_t v[10] = {r.width, r.height, r.spp, (uint32_t)r.meshes.size(), (uint32_t)r.instances.size(), (uint32_t)r.materials.size(),
                      (uint32_t)r.lights.size(), (uint32_t)r.alias_table.size(), (uint32_t)r.textures.size(), (uint32_t)r.tonemap};
    std::memcpy(info, v, sizeof(v));
    if (exposure) *exposure = r.exposure;
}
/* which: 0 instances, 1 materials, 2 lights, 3 alias_table, 4 alias_func, 5 camera, 6 params, 7 mediums, 8 + k LUT k,
 * 100 spectrum record, 101 number of SPDs (uint64), 102 light sampler record, 200 + i SPD i; returns bytes copied */
VB_API uint64_t vh_records_blob(vh_pipeline *h, int which, void *dst, uint64_t cap) {
    const SceneRecords &r = h->p->records;
    const void *src = nullptr; uint64_t n = 0;
    switch (which) {
        case 0: src = r.instances.data(); n = r.instances.size() * sizeof(vb_instance); break;
        case 1: src = r.materials.data(); n = r.materials.size() * sizeof(vb_material_rec); break;
        case 2: src = r.lights.data(); n = r.lights.size() * sizeof(vb_light_rec); break;
        case 3: src = r.alias_table.data(); n = r.alias_table.size() * sizeof(vb_alias_entry); break;
        case 4: src = r.alias_func.data(); n = r.alias_func.size() * sizeof(float); break;
        case 5: src = &r.camera; n = sizeof(vb_camera); break;
        case 6: src = &r.params; n = sizeof(vb_pt_params); break;
        case 7: src = r.mediums.data(); n = r.mediums.size() * sizeof(vb_medium_rec); break;
        case 102: src = &r.light_sampler; n = sizeof(r.light_sampler); break;
        case 100: src = &r.spectrum; n = sizeof(vb_spectrum_rec); break;
        case 101: { uint64_t c = r.spds.size(); if (dst && cap >= 8) std::memcpy(dst, &c, 8); return 8; }
        default:
            if (which >= 200) {                                              // 200 + i: SPD i of the pool
                if ((size_t)(which - 200) >= r.spds.size()) return 0;
                src = r.spds[which - 200].data(); n = r.spds[which - 200].size() * 4;
                break;
            }
            if (which < 8 || which >= 8 + VB_LUT_COUNT) return 0;           // 8 + k: LUT k
            src = r.luts[which - 8].data.data(); n = r.luts[which - 8].data.size() * 4;
            break;
    }
    if (dst && cap >= n && n) std::memcpy(dst, src, n);
    return n;
}
VB_API void vh_mesh_info(vh_pipeline *h, uint32_t i, uint32_t *nv, uint32_t *nt) { *nv = h->p->records.meshes[i]->nv(); *nt = h->p->records.meshes[i]->nt(); }
VB_API void vh_mesh_copy(vh_pipeline *h, uint32_t i, float *pos, float *nrm, float *uv, uint32_t *tri) {
    const MeshData &m = *h->p->records.meshes[i];
    std::memcpy(pos, m.pos.data(), m.pos.size() * 4); std::memcpy(nrm, m.nrm.data(), m.nrm.size() * 4);
    std::memcpy(uv, m.uv.data(), m.uv.size() * 4); std::memcpy(tri, m.tri.data(), m.tri.size() * 4);
}
VB_API void vh_texture_info(vh_pipeline *h, uint32_t i, uint32_t *w, uint32_t *hh) { *w = h->p->records.textures[i].w; *hh = h->p->records.textures[i].h; }
VB_API void vh_texture_copy(vh_pipeline *h, uint32_t i, float *rgba) { const Image &t = h->p->records.textures[i]; std::memcpy(rgba, t.rgba.data(), t.rgba.size() * 4); }
/* registered plugins, '\n' separated "vision-<category>-<name>#<constructor>" */
VB_API uint64_t vh_plugin_names(char *dst, uint64_t cap) {
    std::string s;
    for (auto &n : PluginRegistry::instance().names()) s += n + "\n";
    if (dst && cap > s.size()) std::memcpy(dst, s.c_str(), s.size() + 1);
    return s.size() + 1;
}
/* Render-graph set-up without a device (tests): `spec` = "name=type,name=type;src.chan>dst.chan,...;out.chan". Returns 0 and the ordered,
 * simplified pass list ("integrate accumulate ...") in dst, or -1 and the error text. */
VB_API int vh_render_graph_order(const char *spec, char *dst, uint64_t cap) {
    std::string out; int rc = 0;
    try {
        std::string sp(spec);
        std::vector<std::string> part;
        size_t a = 0;
        for (size_t i = 0; i <= sp.size(); ++i) if (i == sp.size() || sp[i] == ';') { part.push_back(sp.substr(a, i - a)); a = i + 1; }
        if (part.size() != 3) throw std::runtime_error("render graph spec: passes;edges;output");
        auto split = [](const std::string &x, char c) { std::vector<std::string> v; size_t b = 0; for (size_t i = 0; i <= x.size(); ++i) if (i == x.size() || x[i] == c) { if (i > b) v.push_back(x.substr(b, i - b)); b = i + 1; } return v; };
        RenderGraph g;
        for (auto &pt : split(part[0], ',')) { auto kv = split(pt, '='); if (kv.size() != 2) throw std::runtime_error("render graph spec: name=type"); g.add_pass(RenderPass::create(kv[1]), kv[0]); }
        for (auto &e : split(part[1], ',')) { auto kv = split(e, '>'); if (kv.size() != 2) throw std::runtime_error("render graph spec: src>dst"); g.add_edge(kv[0], kv[1]); }
        g.mark_output(part[2]);
        g.setup();
        for (auto *p : g.pass_list()) out += (out.empty() ? "" : " ") + p->pass_name;
    } catch (const std::exception &e) { out = e.what(); rc = -1; }
    if (dst && cap) { size_t n = std::min<size_t>(out.size(), cap - 1); std::memcpy(dst, out.data(), n); dst[n] = 0; }
    return rc;
}
VB_API uint64_t vh_scene_log(vh_pipeline *h, char *dst, uint64_t cap) {
    std::string s;
    for (auto &n : h->p->scene.log) s += n + "\n";
    if (dst && cap > s.size()) std::memcpy(dst, s.c_str(), s.size() + 1);
    return s.size() + 1;
}

/* model import by itself (the `model` plugin's loader): importer 1 = the reference's (vendored assimp, reference vertex / triangle
 * order), 0 = the in-tree OBJ loader (file order). Returns a handle to the imported meshes or null. */
struct vh_model { std::vector<MeshData> meshes; };
VB_API int vh_assimp_available() { return assimp_available() ? 1 : 0; }
VB_API vh_model *vh_model_load(const char *path, int importer, int swap_handed, int smooth, int flip_uv) {
    try {
        auto m = new vh_model();
        if (importer) m->meshes = load_model_assimp(path, swap_handed != 0, smooth != 0, flip_uv != 0, 0);
        else m->meshes.push_back(load_obj(path, smooth != 0, flip_uv != 0));
        return m;
    } catch (const std::exception &e) { g_err = e.what(); return nullptr; }
}
VB_API uint32_t vh_model_count(vh_model *m) { return (uint32_t)m->meshes.size(); }
VB_API void vh_model_mesh_info(vh_model *m, uint32_t i, uint32_t *nv, uint32_t *nt) { *nv = m->meshes[i].nv(); *nt = m->meshes[i].nt(); }
VB_API void vh_model_mesh_copy(vh_model *m, uint32_t i, float *pos, float *nrm, float *uv, uint32_t *tri) {
    const MeshData &d = m->meshes[i];
    std::memcpy(pos, d.pos.data(), d.pos.size() * 4); std::memcpy(nrm, d.nrm.data(), d.nrm.size() * 4);
    std::memcpy(uv, d.uv.data(), d.uv.size() * 4); std::memcpy(tri, d.tri.data(), d.tri.size() * 4);
}
VB_API void vh_model_free(vh_model *m) { delete m; }

/* image files (vh_image.cpp): save by extension (.png / .ppm with the display gamma, .hdr / .pfm linear); load (.jpg / .png / .exr / .hdr / .pfm /
 * .ppm decoded in the host layer) - first call with dst = null to get the size */
VB_API int vh_save_image(const char *path, const float *rgba, uint32_t w, uint32_t h) { VH_TRY(save_image(path, rgba, w, h)) }
VB_API int vh_load_image(const char *path, int srgb_to_linear, uint32_t *w, uint32_t *h, float *dst, uint64_t cap_floats) {
    VH_TRY({
        Image im = load_image(path, srgb_to_linear != 0);
        *w = im.w; *h = im.h;
        if (dst && cap_floats >= im.rgba.size()) std::memcpy(dst, im.rgba.data(), im.rgba.size() * 4);
    })
}
VB_API void vh_rgbe_encode(const float *rgb, unsigned char *out4) { rgbe_encode(rgb, out4); }
VB_API void vh_rgbe_decode(const unsigned char *in4, float *rgb) { rgbe_decode(in4, rgb); }

}  // extern "C"

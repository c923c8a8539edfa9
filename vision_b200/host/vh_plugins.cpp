// Plugin implementations of the host layer (one shared library instead of one MODULE per plugin).
// Each class cites the reference plugin it restates; host-side behaviour (desc keys, defaults,
// vertex/triangle order, table construction) is kept, device code is a POD record.
#include "vh_scene.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>

namespace vh {

PluginRegistry &PluginRegistry::instance() { static PluginRegistry r; return r; }
void PluginRegistry::add(const std::string &p, const std::string &c, INodeConstructor *k) { entries_.push_back({p + "#" + c, k}); }
INodeConstructor *PluginRegistry::find(const std::string &p, const std::string &c) const {
    for (auto &e : entries_) if (e.first == p + "#" + c) return e.second;
    return nullptr;
}
std::vector<std::string> PluginRegistry::names() const { std::vector<std::string> v; for (auto &e : entries_) v.push_back(e.first); return v; }

// ---- meshes ----------------------------------------------------------------------------------
std::vector<float> MeshData::surface_areas() const {
    std::vector<float> out;
    for (uint32_t t = 0; t < nt(); ++t) {
        const float *a = &pos[3 * tri[3 * t]], *b = &pos[3 * tri[3 * t + 1]], *c = &pos[3 * tri[3 * t + 2]];
        float e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        float cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
        out.push_back(0.5f * std::sqrt(cx * cx + cy * cy + cz * cz));
    }
    return out;
}
uint64_t MeshData::hash() const {
    uint64_t h = 1469598103934665603ull;
    auto mix = [&](const void *p, size_t n) { const unsigned char *b = (const unsigned char *)p; for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; } };
    mix(pos.data(), pos.size() * 4); mix(nrm.data(), nrm.size() * 4); mix(uv.data(), uv.size() * 4); mix(tri.data(), tri.size() * 4);
    return h;
}
void ShapeGroup::add_instance(MeshData mesh) {
    auto inst = std::make_shared<ShapeInstance>();
    inst->mesh = std::make_shared<MeshData>(std::move(mesh));
    instances.push_back(inst);
}
void ShapeGroup::post_init(const ShapeDesc &desc) {
    std::string mat = desc["material"].as_string();
    for (size_t i = 0; i < instances.size(); ++i) {
        instances[i]->material_name = mat;
        instances[i]->o2w = desc.o2w.mat;
        instances[i]->name = name_ + "_" + std::to_string(i);
        instances[i]->has_medium = desc.has_medium;
        instances[i]->inside_name = desc.inside_medium; instances[i]->outside_name = desc.outside_medium;
    }
}

// reference src/render_core/shape/quad.cpp:21-50
class Quad : public ShapeGroup {
public:
    explicit Quad(const ShapeDesc &desc) : ShapeGroup(desc) {
        name_ = desc.name;
        float w = desc["width"].as_float(1.f) / 2, h = desc["height"].as_float(1.f) / 2;
        MeshData m;
        float P[4][3] = {{w, 0, h}, {w, 0, -h}, {-w, 0, h}, {-w, 0, -h}};
        float UV[4][2] = {{1, 1}, {1, 0}, {0, 1}, {0, 0}};
        float d02[3] = {P[0][0] - P[2][0], P[0][1] - P[2][1], P[0][2] - P[2][2]}, d12[3] = {P[1][0] - P[2][0], P[1][1] - P[2][1], P[1][2] - P[2][2]};
        float ng[3] = {d02[1] * d12[2] - d02[2] * d12[1], d02[2] * d12[0] - d02[0] * d12[2], d02[0] * d12[1] - d02[1] * d12[0]};
        for (int i = 0; i < 4; ++i) {
            m.pos.insert(m.pos.end(), P[i], P[i] + 3); m.nrm.insert(m.nrm.end(), ng, ng + 3); m.uv.insert(m.uv.end(), UV[i], UV[i] + 2);
        }
        m.tri = {0, 1, 2, 2, 1, 3};
        add_instance(std::move(m));
        post_init(desc);
    }
    std::string impl_type() const override { return "quad"; }
};
VH_MAKE_CLASS_CREATOR(Quad, shape, quad, constructor)

// reference src/render_core/shape/cube.cpp:20-71
class Cube : public ShapeGroup {
public:
    explicit Cube(const ShapeDesc &desc) : ShapeGroup(desc) {
        name_ = desc.name;
        float x = desc["x"].as_float(1.f), y = desc["y"].as_float(1.f), z = desc["z"].as_float(1.f);
        y = y == 0 ? x : y; z = z == 0 ? y : z;
        x /= 2.f; y /= 2.f; z /= 2.f;
        const float P[24][3] = {{-x, -y, z}, {x, -y, z}, {-x, y, z}, {x, y, z}, {-x, y, -z}, {x, y, -z}, {-x, -y, -z}, {x, -y, -z},
                                {-x, y, z}, {x, y, z}, {-x, y, -z}, {x, y, -z}, {-x, -y, z}, {x, -y, z}, {-x, -y, -z}, {x, -y, -z},
                                {x, -y, z}, {x, y, z}, {x, y, -z}, {x, -y, -z}, {-x, -y, z}, {-x, y, z}, {-x, y, -z}, {-x, -y, -z}};
        const float N[6][3] = {{0, 0, 1}, {0, 0, -1}, {0, 1, 0}, {0, -1, 0}, {1, 0, 0}, {-1, 0, 0}};
        const float UV[24][2] = {{0, 0}, {1, 0}, {0, 1}, {1, 1}, {0, 1}, {1, 1}, {0, 0}, {1, 0}, {0, 1}, {1, 1}, {0, 0}, {1, 0},
                                 {0, 1}, {1, 1}, {0, 0}, {1, 0}, {0, 1}, {1, 1}, {1, 0}, {0, 0}, {0, 1}, {1, 1}, {1, 0}, {0, 0}};
        MeshData m;
        for (int i = 0; i < 24; ++i) {
            m.pos.insert(m.pos.end(), P[i], P[i] + 3); m.nrm.insert(m.nrm.end(), N[i / 4], N[i / 4] + 3); m.uv.insert(m.uv.end(), UV[i], UV[i] + 2);
        }
        m.tri = {0, 1, 3, 0, 3, 2, 6, 5, 7, 4, 5, 6, 10, 9, 11, 8, 9, 10, 13, 14, 15, 13, 12, 14, 18, 17, 19, 17, 16, 19, 21, 22, 23, 20, 21, 23};
        add_instance(std::move(m));
        post_init(desc);
    }
    std::string impl_type() const override { return "cube"; }
};
VH_MAKE_CLASS_CREATOR(Cube, shape, cube, constructor)

// ---- slots -------------------------------------------------------------------------------------
// SlotDesc / ShaderNodeDesc::init (node_desc.cpp:128-152,318-348): a colour attribute is
//   [r,g,b] | number | {"channels","node":X} | {"type":"number","param":{"value"}} | {"fn","color_space"} | {"type":"image","param":{..}}
static vb_slot make_slot(const ParameterSet &ps, float3 def, const SlotContext &ctx) {
    vb_slot s{};
    s.tex_id = VB_INVALID_ID;
    s.value[0] = def.x; s.value[1] = def.y; s.value[2] = def.z; s.value[3] = 1.f;
    const Json &j = ps.data();
    if (j.is_null()) return s;
    if (j.is_object() && j.contains("node")) return make_slot(ps["node"], def, ctx);
    if (j.is_number()) { s.value[0] = s.value[1] = s.value[2] = (float)j.num; return s; }
    if (j.is_array()) {
        for (size_t i = 0; i < j.arr.size() && i < 4; ++i) s.value[i] = (float)j.arr[i].num;
        if (j.arr.size() == 1) s.value[1] = s.value[2] = s.value[0];
        return s;
    }
    if (j.is_object()) {
        std::string type = ps["type"].as_string(j.contains("fn") ? "image" : "number");
        ParameterSet prm = j.contains("param") ? ps["param"] : ps;
        if (type == "number") return make_slot(prm["value"], def, ctx);
        if (type == "image") {
            s.value[0] = s.value[1] = s.value[2] = s.value[3] = 1.f;
            s.tex_id = ctx.images->obtain(ctx.scene_path + "/" + prm["fn"].as_string(), prm["color_space"].as_string("srgb"));
            return s;
        }
        throw std::runtime_error("shader node type not supported on this path: " + type);
    }
    return s;
}

// ---- materials ---------------------------------------------------------------------------------
static void init_material_record(vb_material_rec &r) {     // unused slots: zero constant, no texture
    std::memset(&r, 0, sizeof(r));
    r.color.tex_id = r.roughness.tex_id = r.spec.tex_id = r.ior.tex_id = r.sigma.tex_id = VB_INVALID_ID;
}
// reference src/render_core/material/diffuse.cpp:12-52
class DiffuseMaterial : public Material {
    MaterialDesc desc_;
public:
    explicit DiffuseMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "diffuse"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        init_material_record(r);
        r.type = VB_MAT_DIFFUSE;
        r.color = make_slot(desc_["color"], {0.5f, 0.5f, 0.5f}, ctx);
        if (desc_.has_attr("sigma")) { r.flags |= 1u; r.sigma = make_slot(desc_["sigma"], {0.5f, 0.5f, 0.5f}, ctx); }
    }
};
VH_MAKE_CLASS_CREATOR(DiffuseMaterial, material, diffuse, constructor)

// common part of the microfacet materials: roughness / anisotropic Number slots, remapping_roughness flag
static void fill_microfacet_common(vb_material_rec &r, const MaterialDesc &d, float default_roughness, const SlotContext &ctx) {
    init_material_record(r);
    r.flags = d["remapping_roughness"].as_bool(true) ? 2u : 0u;
    float rough = make_slot(d["roughness"], {default_roughness, default_roughness, default_roughness}, ctx).value[0];
    r.roughness = vb_slot{{rough, rough, rough, 1.f}, VB_INVALID_ID, {0, 0, 0}};
    r.params[0] = make_slot(d["anisotropic"], {0.f, 0.f, 0.f}, ctx).value[0];
}

// reference src/render_core/material/mirror.cpp:18-75
class MirrorMaterial : public Material {
    MaterialDesc desc_;
public:
    explicit MirrorMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "mirror"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        fill_microfacet_common(r, desc_, 0.001f, ctx);
        r.type = VB_MAT_MIRROR;
        r.color = make_slot(desc_["color"], {1.f, 1.f, 1.f}, ctx);
    }
};
VH_MAKE_CLASS_CREATOR(MirrorMaterial, material, mirror, constructor)

// reference src/render_core/material/glass.cpp:104-134 (Sellmeier curves), :165-257
class GlassMaterial : public Material {
    MaterialDesc desc_;
    static float curve(const std::string &name, float lambda) {
        static const float bk7[6] = {1.03961212f, 0.00600069867f, 0.231792344f, 0.0200179144f, 1.01046945f, 103.560653f};
        static const float lasf9[6] = {2.00029547f, 0.0121426017f, 0.298926886f, 0.0538736236f, 1.80691843f, 156.530829f};
        const float *c = name == "LASF9" ? lasf9 : bk7;            // unknown names fall back to BK7 (glass.cpp:150-155)
        float l = lambda / 1000.f, l2 = l * l;
        float f = c[0] * l2 / (l2 - c[1]) + c[2] * l2 / (l2 - c[3]) + c[4] * l2 / (l2 - c[5]);
        return std::sqrt(f + 1.f);
    }
public:
    explicit GlassMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "glass"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        fill_microfacet_common(r, desc_, 0.5f, ctx);
        r.type = VB_MAT_GLASS;
        r.color = make_slot(desc_["color"], {1.f, 1.f, 1.f}, ctx);
        std::string name = desc_["material_name"].as_string();
        // non-spectral spectrum: the curve is evaluated at rgb_spectrum_peak_wavelengths.x (glass.cpp:225-229)
        float ior = name.empty() ? make_slot(desc_["ior"], {1.5f, 1.5f, 1.5f}, ctx).value[0] : curve(name, 602.785f);
        r.ior = vb_slot{{ior, ior, ior, 1.f}, VB_INVALID_ID, {0, 0, 0}};
        if (!name.empty() && ctx.spectrum_complete) {              // glass.cpp:221-224: SPD::to_list(curve) -> dispersive "spd" ior slot
            std::vector<float> lst;
            for (float lambda = 360.f; lambda < 830.f; lambda += 5.f) lst.push_back(curve(name, lambda));
            r.flags |= 4u;
            r.spd[0] = (uint32_t)ctx.spds->size();
            ctx.spds->push_back(std::move(lst));
        }
    }
};
VH_MAKE_CLASS_CREATOR(GlassMaterial, material, glass, constructor)

// reference src/render_core/material/metal.cpp:39-157; complex ior of the named metal at the three sRGB peak
// wavelengths (metal_ior.h / SPD::eval) precomputed into data/metal_ior_srgb.json by tools/extract_metal_ior.py
class MetalMaterial : public Material {
    MaterialDesc desc_;
public:
    explicit MetalMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "metal"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        fill_microfacet_common(r, desc_, 0.01f, ctx);
        r.type = VB_MAT_METAL;
        Json table = Json::parse_file(data_dir() + "/metal_ior_srgb.json");
        std::string name = desc_["material_name"].as_string();
        if (!table.contains(name)) {                               // metal.cpp:106-108: unknown -> first key of the (sorted) map
            std::vector<std::string> keys;
            for (auto &kv : table.obj) keys.push_back(kv.first);
            std::sort(keys.begin(), keys.end());
            name = keys.front();
        }
        const Json &e = table.at(name);
        for (int i = 0; i < 3; ++i) { r.eta[i] = (float)e.at("eta").arr[i].num; r.k[i] = (float)e.at("k").arr[i].num; }
        if (ctx.spectrum_complete) {                               // metal.cpp:117-121: the measured tables as "spd" eta / k slots
            Json full = Json::parse_file(data_dir() + "/metal_ior_spd.json");
            const Json &fe = full.at(name);
            std::vector<float> eta, k;
            for (auto &v : fe.at("eta").arr) eta.push_back((float)v.num);
            for (auto &v : fe.at("k").arr) k.push_back((float)v.num);
            r.flags |= 4u;
            r.spd[0] = (uint32_t)ctx.spds->size(); ctx.spds->push_back(std::move(eta));
            r.spd[1] = (uint32_t)ctx.spds->size(); ctx.spds->push_back(std::move(k));
        }
    }
};
VH_MAKE_CLASS_CREATOR(MetalMaterial, material, metal, constructor)

// reference src/render_core/material/substrate.cpp:92-150
class SubstrateMaterial : public Material {
    MaterialDesc desc_;
public:
    explicit SubstrateMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "substrate"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        fill_microfacet_common(r, desc_, 0.5f, ctx);
        r.type = VB_MAT_SUBSTRATE;
        r.color = make_slot(desc_["color"], {1.f, 1.f, 1.f}, ctx);
        r.spec = make_slot(desc_["spec"], {0.05f, 0.05f, 0.05f}, ctx);
    }
};
VH_MAKE_CLASS_CREATOR(SubstrateMaterial, material, substrate, constructor)

// reference src/render_core/material/principled_bsdf.cpp:216-461 (slot names and defaults :264-290). Colour-like slots that
// scenes texture (color, spec_tint) and roughness / ior stay slots; the remaining inputs are constants in params[]
class PrincipledMaterial : public Material {
    MaterialDesc desc_;
public:
    explicit PrincipledMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "principled_bsdf"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        init_material_record(r);
        r.type = VB_MAT_PRINCIPLED;
        r.color = make_slot(desc_["color"], {1.f, 1.f, 1.f}, ctx);
        r.spec = make_slot(desc_["spec_tint"], {1.f, 1.f, 1.f}, ctx);
        auto number = [&](const char *key, float def) { return desc_.has_attr(key) ? make_slot(desc_[key], {def, def, def}, ctx).value[0] : def; };
        auto const3 = [&](const char *key, float *dst) {
            vb_slot s = make_slot(desc_[key], {1.f, 1.f, 1.f}, ctx);
            if (s.tex_id != VB_INVALID_ID) throw std::runtime_error(std::string("principled_bsdf: textured ") + key + " is not supported");
            dst[0] = s.value[0]; dst[1] = s.value[1]; dst[2] = s.value[2];
        };
        float rough = number("roughness", 0.5f), ior = number("ior", 1.5f);
        r.roughness.value[0] = r.roughness.value[1] = r.roughness.value[2] = rough; r.roughness.value[3] = 1.f;
        r.ior.value[0] = r.ior.value[1] = r.ior.value[2] = ior; r.ior.value[3] = 1.f;
        float *P = r.params;
        P[0] = number("anisotropic", 0.f);
        P[1] = number("metallic", 0.f);
        P[2] = number("sheen_weight", 0.f); P[3] = number("sheen_roughness", 0.5f);
        const3("sheen_tint", P + 4);
        P[7] = number("coat_weight", 0.f); P[8] = number("coat_roughness", 0.2f); P[9] = number("coat_ior", 1.5f);
        const3("coat_tint", P + 10);
        P[13] = number("transmission_weight", 0.f);
        P[14] = 1.f;                 // sheen_mode_ = Approximate
        P[15] = 63.f;                // all lobe switches on
    }
};
VH_MAKE_CLASS_CREATOR(PrincipledMaterial, material, principled_bsdf, constructor)

// reference src/render_core/medium/homogeneous.cpp:11-28 + MediumDesc::init (node_desc.cpp:182-197)
class HomogeneousMedium : public Medium {
    MediumDesc desc_;
public:
    explicit HomogeneousMedium(const MediumDesc &d) : Medium(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "homogeneous"; }
    void fill_record(vb_medium_rec &r) const override {
        std::memset(&r, 0, sizeof(r));
        r.type = VB_MEDIUM_HOMOGENEOUS;
        std::string mn = desc_["medium_name"].as_string();
        float3 ss{0, 0, 0}, sa{0, 0, 0};
        if (!mn.empty()) {
            Json table = Json::parse_file(data_dir() + "/medium_scatter.json");
            const Json *e = table.contains(mn) ? &table.at(mn) : nullptr;
            if (!e && !table.obj.empty()) e = &table.obj.front().second;                    // get_sigma falls back to entry 0
            if (e) { auto f = [&](const char *k, int i) { return (float)e->at(k).arr[i].num; };
                     ss = {f("sigma_s", 0), f("sigma_s", 1), f("sigma_s", 2)}; sa = {f("sigma_a", 0), f("sigma_a", 1), f("sigma_a", 2)}; }
        } else {
            ss = desc_["sigma_s"].as_float3({0, 0, 0}); sa = desc_["sigma_a"].as_float3({0, 0, 0});
        }
        r.sigma_s[0] = ss.x; r.sigma_s[1] = ss.y; r.sigma_s[2] = ss.z; r.sigma_a[0] = sa.x; r.sigma_a[1] = sa.y; r.sigma_a[2] = sa.z;
        r.scale = desc_["scale"].as_float(0.f);
        r.g = std::min(0.99f, std::max(-0.99f, desc_["g"].as_float(0.f)));
    }
};
VH_MAKE_CLASS_CREATOR(HomogeneousMedium, medium, homogeneous, constructor)

// reference src/render_core/material/black_body.cpp (material of light-created geometry, area.cpp:55-68)
class BlackBodyMaterial : public Material {
public:
    explicit BlackBodyMaterial(const MaterialDesc &d) : Material(d) { name_ = d.name; }
    std::string impl_type() const override { return "black_body"; }
    void fill_record(vb_material_rec &r, const SlotContext &) const override { init_material_record(r); r.type = VB_MAT_BLACK_BODY; }
};
VH_MAKE_CLASS_CREATOR(BlackBodyMaterial, material, black_body, constructor)

// reference src/render_core/material/plastic.cpp:68-131 (slots color 1, spec 0.05, ior 1.3, roughness 0.5, anisotropic 0)
class PlasticMaterial : public Material {
    MaterialDesc desc_;
public:
    explicit PlasticMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "plastic"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        fill_microfacet_common(r, desc_, 0.5f, ctx);
        r.type = VB_MAT_PLASTIC;
        r.color = make_slot(desc_["color"], {1.f, 1.f, 1.f}, ctx);
        r.spec = make_slot(desc_["spec"], {0.05f, 0.05f, 0.05f}, ctx);
        float ior = make_slot(desc_["ior"], {1.3f, 1.3f, 1.3f}, ctx).value[0];
        r.ior = vb_slot{{ior, ior, ior, 1.f}, VB_INVALID_ID, {0, 0, 0}};
    }
};
VH_MAKE_CLASS_CREATOR(PlasticMaterial, material, plastic, constructor)

// reference src/render_core/material/metallic.cpp:12-61 (slots color 1, edge_tint 1, roughness 0.5, anisotropic 0; remapping_roughness_
// is a member initialised to true and never read from the desc)
class MetallicMaterial : public Material {
    MaterialDesc desc_;
public:
    explicit MetallicMaterial(const MaterialDesc &d) : Material(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "metallic"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        fill_microfacet_common(r, desc_, 0.5f, ctx);
        r.flags = 2u;
        r.type = VB_MAT_METALLIC;
        r.color = make_slot(desc_["color"], {1.f, 1.f, 1.f}, ctx);
        r.spec = make_slot(desc_["edge_tint"], {1.f, 1.f, 1.f}, ctx);
    }
};
VH_MAKE_CLASS_CREATOR(MetallicMaterial, material, metallic, constructor)

// reference src/render_core/material/mix.cpp:14-58 and add.cpp:13-54: two child materials built from desc.mat0 / desc.mat1
// (MaterialDesc::init, node_desc.cpp:170-176); mix has the "frac" slot (default 0.5)
class LobeSetMaterial : public Material {
protected:
    MaterialDesc desc_;
    std::shared_ptr<Material> mat0_, mat1_;
public:
    explicit LobeSetMaterial(const MaterialDesc &d) : Material(d), desc_(d) {
        name_ = d.name;
        for (int k = 0; k < 2; ++k) {
            const char *key = k ? "mat1" : "mat0";
            if (!d.parameter.contains(key)) throw std::runtime_error("material '" + d.name + "': \"" + key + "\" is missing");
            MaterialDesc cd;
            cd.init(d[key]);
            if (cd.name.empty()) cd.name = d.name + "." + key;
            (k ? mat1_ : mat0_) = create_shared<Material>(cd);
        }
    }
    std::vector<std::shared_ptr<Material>> children() const override { return {mat0_, mat1_}; }
    void fill_children(vb_material_rec &r) const {
        if (mat0_->index == VB_INVALID_ID || mat1_->index == VB_INVALID_ID) throw std::runtime_error("material '" + name_ + "': children have no record index");
        r.spd[0] = mat0_->index; r.spd[1] = mat1_->index;
    }
};
class MixMaterial : public LobeSetMaterial {
public:
    explicit MixMaterial(const MaterialDesc &d) : LobeSetMaterial(d) {}
    std::string impl_type() const override { return "mix"; }
    void fill_record(vb_material_rec &r, const SlotContext &ctx) const override {
        init_material_record(r);
        r.type = VB_MAT_MIX;
        fill_children(r);
        // the plugin reads "frac" (default 0.5); the stock scene files write "scale", which nothing reads
        float frac = desc_.has_attr("frac") ? make_slot(desc_["frac"], {0.5f, 0.5f, 0.5f}, ctx).value[0] : 0.5f;
        r.sigma = vb_slot{{frac, frac, frac, 1.f}, VB_INVALID_ID, {0, 0, 0}};
    }
};
VH_MAKE_CLASS_CREATOR(MixMaterial, material, mix, constructor)
class AddMaterial : public LobeSetMaterial {
public:
    explicit AddMaterial(const MaterialDesc &d) : LobeSetMaterial(d) {}
    std::string impl_type() const override { return "add"; }
    void fill_record(vb_material_rec &r, const SlotContext &) const override { init_material_record(r); r.type = VB_MAT_ADD; fill_children(r); }
};
VH_MAKE_CLASS_CREATOR(AddMaterial, material, add, constructor)

// reference src/render_core/material/subsurface.cpp and multi_layered.cpp are themselves unfinished: neither has a create_lobe_set nor
// a VS_MAKE_CLASS_CREATOR, so the reference cannot instantiate them either. They construct here so that a scene file listing one
// loads; USING one is refused loudly.
#define VH_UNBUILT_MATERIAL(Class, tag)                                                                              \
    class Class : public Material {                                                                                  \
    public:                                                                                                          \
        explicit Class(const MaterialDesc &d) : Material(d) { name_ = d.name; }                                      \
        std::string impl_type() const override { return #tag; }                                                      \
        void fill_record(vb_material_rec &, const SlotContext &) const override {                                    \
            throw std::runtime_error("material '" + name_ + "' (" #tag "): the reference has no lobe set for this type either");  \
        }                                                                                                            \
    };                                                                                                               \
    VH_MAKE_CLASS_CREATOR(Class, material, tag, constructor)
VH_UNBUILT_MATERIAL(SubsurfaceMaterial, subsurface)
VH_UNBUILT_MATERIAL(MultiLayeredMaterial, multi_layered)
#undef VH_UNBUILT_MATERIAL

// ---- warper ------------------------------------------------------------------------------------
// reference src/render_core/warper/alias.h:86-122 (sum in double, probabilities in float, LIFO stacks)
class AliasTable : public Warper {
public:
    explicit AliasTable(const WarperDesc &d) : Warper(d) {}
    std::string impl_type() const override { return "alias"; }
    void build(const std::vector<float> &weights, std::vector<vb_alias_entry> &table, std::vector<float> &func, float &integral) const override {
        double sum = std::accumulate(weights.begin(), weights.end(), 0.0);
        double ratio = (double)weights.size() / sum;
        std::vector<uint32_t> over, under;
        std::vector<vb_alias_entry> t(weights.size());
        for (uint32_t i = 0; i < weights.size(); ++i) {
            float p = (float)(weights[i] * ratio);
            t[i] = {p, i};
            (p > 1.0f ? over : under).push_back(i);
        }
        while (!over.empty() && !under.empty()) {
            uint32_t o = over.back(), u = under.back();
            over.pop_back(); under.pop_back();
            t[o].prob -= 1.0f - t[u].prob;
            t[u].alias = o;
            if (t[o].prob > 1.0f) over.push_back(o);
            else if (t[o].prob < 1.0f) under.push_back(o);
        }
        for (auto i : over) t[i] = {1.0f, i};
        for (auto i : under) t[i] = {1.0f, i};
        integral = (float)(sum / weights.size());
        table.insert(table.end(), t.begin(), t.end());
        func.insert(func.end(), weights.begin(), weights.end());
    }
};
VH_MAKE_CLASS_CREATOR(AliasTable, warper, alias, constructor)

// ---- lights ------------------------------------------------------------------------------------
// reference src/render_core/light/area.cpp:33-177 + Light::initialize_slots (light.cpp:19-24)
class AreaLight : public Light {
    LightDesc desc_;
public:
    explicit AreaLight(const LightDesc &d) : Light(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "area"; }
    int type_order() const override { return VB_LIGHT_AREA; }
    void fill_record(vb_light_rec &r, const Scene &scene, const SlotContext &ctx, std::vector<vb_alias_entry> &table,
                     std::vector<float> &func) const override {
        std::memset(&r, 0, sizeof(r));
        r.type = VB_LIGHT_AREA;
        r.inst_id = inst_idx;
        r.two_sided = desc_["two_sided"].as_bool(false) ? 1u : 0u;
        r.color = make_slot(desc_["color"], {0.5f, 0.5f, 0.5f}, ctx);
        float scale = make_slot(desc_["scale"], {1.f, 1.f, 1.f}, ctx).value[0];
        if (!desc_.has_attr("scale")) scale = 1.f;
        if (r.color.tex_id == VB_INVALID_ID) {                 // NumberArray::normalize (number.cpp:40-49)
            float mx = std::max(r.color.value[0], std::max(r.color.value[1], r.color.value[2]));
            if (!(mx < 1.f)) { for (int i = 0; i < 3; ++i) r.color.value[i] /= mx; scale *= mx; }
        }
        r.scale = scale;
        std::vector<float> w = scene.instances[inst_idx]->mesh->surface_areas();
        r.alias_offset = (uint32_t)table.size(); r.alias_size = (uint32_t)w.size();
        scene.warper->build(w, table, func, r.alias_integral);
    }
};
VH_MAKE_CLASS_CREATOR(AreaLight, light, area, constructor)

// LightSampler ctor (lightsampler.cpp:11-13): env_separate (default false) samples the environment light with its own probability
// env_prob, clamped to [0.01, 0.99]; the rule itself (lightsampler.cpp:159-197) runs on the device (vb_light_sampler_env_set)
LightSampler::LightSampler(const LightSamplerDesc &d)
    : env_separate(d["env_separate"].as_bool(false)), env_prob(std::min(std::max(d["env_prob"].as_float(0.5f), 0.01f), 0.99f)) {}

void LightSampler::tidy_up() {
    // Polymorphic topology order = order in which light TYPES were first added
    std::vector<int> first;
    for (auto &l : lights) if (std::find(first.begin(), first.end(), l->type_order()) == first.end()) first.push_back(l->type_order());
    std::stable_sort(lights.begin(), lights.end(), [&](const std::shared_ptr<Light> &a, const std::shared_ptr<Light> &b) {
        return std::find(first.begin(), first.end(), a->type_order()) < std::find(first.begin(), first.end(), b->type_order());
    });
    for (uint32_t i = 0; i < lights.size(); ++i) lights[i]->index = i;
}
// reference src/render_core/lightsampler/uniform.cpp:12-34 (device part: index = min(u*N, N-1), PMF 1/N)
class UniformLightSampler : public LightSampler {
public:
    explicit UniformLightSampler(const LightSamplerDesc &d) : LightSampler(d) {}
    std::string impl_type() const override { return "uniform"; }
};
VH_MAKE_CLASS_CREATOR(UniformLightSampler, lightsampler, uniform, constructor)

// luminance(light->power()), the weight PowerLightSampler::prepare gives each light (power.cpp:35-50): area.cpp:87-89, point.cpp:32-34,
// spot.cpp:44-46, projector.cpp:58-91, spherical.cpp:56-59, directional.cpp:44-46; Light::average (light.h:80-83) = slot average x scale. A texture's average
// (ocarina Image::average_vector, absent from the tree) = per-channel mean summed in double in pixel order; luminance = Rec.709.
float light_power_luminance(const vb_light_rec &l, const SceneRecords &rec, float spot_angle, float spot_falloff) {
    const float Pi = 3.14159265358979323846f;
    float avg[3];
    if (l.color.tex_id == VB_INVALID_ID) for (int c = 0; c < 3; ++c) avg[c] = l.color.value[c];
    else {
        const Image &im = rec.textures[l.color.tex_id];
        size_t n = (size_t)im.w * im.h;
        for (int c = 0; c < 3; ++c) { double s = 0.0; for (size_t i = 0; i < n; ++i) s += (double)im.rgba[i * 4 + c]; avg[c] = (float)(s / (double)n); }
    }
    for (float &a : avg) a = a * l.scale;
    float p[3];
    if (l.type == VB_LIGHT_AREA) {
        float area = 0.f;
        for (uint32_t i = 0; i < l.alias_size; ++i) area += rec.alias_func[l.alias_offset + i];
        float k = l.two_sided ? 2.f : 1.f;
        for (int c = 0; c < 3; ++c) p[c] = k * avg[c] * area * Pi;
    } else if (l.type == VB_LIGHT_POINT) {
        float k = 4.f * Pi;
        for (int c = 0; c < 3; ++c) p[c] = k * avg[c];
    } else if (l.type == VB_LIGHT_SPOT) {
        float k = 2.f * Pi, s = 1.f - .5f * (spot_angle * 2.f + spot_falloff);
        for (int c = 0; c < 3; ++c) p[c] = k * avg[c] * s;
    } else if (l.type == VB_LIGHT_ENV || l.type == VB_LIGHT_DIRECTIONAL) {      // directional.cpp:44-46: average * Pi * sqr(world_radius)
        float r = l.params[0] / 2.f, k = Pi * (r * r);
        for (int c = 0; c < 3; ++c) p[c] = l.type == VB_LIGHT_ENV ? k * avg[c] : (avg[c] * Pi) * (r * r);
    } else {                                                           // projector: Girard's theorem on the frustum corner directions
        float ratio = l.params[1];
        float y = std::sqrt(1.f / (ratio * ratio + 1.f)), x = ratio * y, z = std::sqrt(x * x + y * y);
        float3 p0{x, y, z}, p1{x, -y, z}, p2{-x, -y, z};
        auto crs = [](float3 a, float3 b) { return float3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; };
        auto dt = [](float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; };
        auto nrm = [&](float3 a) { float l2 = dt(a, a); if (!(l2 > 0.f)) return a; float inv = 1.f / std::sqrt(l2); return float3{a.x * inv, a.y * inv, a.z * inv}; };
        auto neg = [](float3 a) { return float3{-a.x, -a.y, -a.z}; };
        auto ac = [&](float3 a, float3 b) { return std::acos(std::min(std::max(dt(a, b), -1.f), 1.f)); };
        float3 c01 = nrm(crs(p0, p1)), c12 = nrm(crs(p1, p2)), c20 = nrm(crs(p2, p0));
        float solid = std::fabs(ac(c01, neg(c12)) + ac(c12, neg(c20)) + ac(c20, neg(c01)) - Pi);
        float k = (2.f * solid) / (4.f * Pi);
        for (int c = 0; c < 3; ++c) p[c] = k * avg[c];
    }
    return 0.212671f * p[0] + 0.715160f * p[1] + 0.072169f * p[2];
}

// reference src/render_core/lightsampler/power.cpp:11-56: lights are drawn from an alias table over luminance(light->power())
class PowerLightSampler : public LightSampler {
public:
    explicit PowerLightSampler(const LightSamplerDesc &d) : LightSampler(d) {}
    std::string impl_type() const override { return "power"; }
    uint32_t kind() const override { return VB_LIGHT_SAMPLER_POWER; }
};
VH_MAKE_CLASS_CREATOR(PowerLightSampler, lightsampler, power, constructor)

}  // namespace vh

#include "vh_plugins2.inl"
#include "vh_plugins3.inl"

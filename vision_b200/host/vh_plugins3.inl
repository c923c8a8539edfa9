// Included by vh_plugins.cpp: model shape (Wavefront OBJ), point / spot / spherical-environment lights.
#include <fstream>
#include <map>
#include <sstream>
#include <tuple>

namespace vh {

// ---- OBJ ---------------------------------------------------------------------------------------
// The reference imports models through assimp (assimp_parser.cpp:309-341: JoinIdenticalVertices | GenNormals |
// ImproveCacheLocality | FindDegenerates | OptimizeMeshes | Triangulate | FlipUVs); its cache-locality pass
// re-orders triangles, so prim ids of a model are importer-defined. This loader keeps FILE order (SURVEY.md 8f
// N1): one mesh per file, vertices de-duplicated by their (v, vt, vn) index triple in first-use order, polygons
// fan-triangulated, v -> 1 - v when flip_uv, faces with a repeated position index dropped, missing normals =
// face normals (or their sum per position when smooth). vision_b200/scene_py.py obj_mesh is the twin.
static void cross3f(const float *a, const float *b, float *o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
MeshData load_obj(const std::string &path, bool smooth, bool flip_uv) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open model " + path);
    std::vector<float> V, VT, VN;
    struct Idx { int v, t, n; };
    std::vector<std::array<Idx, 3>> faces;
    std::string line;
    while (std::getline(f, line)) {
        std::istringstream ss(line);
        std::string tag;
        if (!(ss >> tag)) continue;
        if (tag == "v" || tag == "vn") {
            std::string a, b, c; ss >> a >> b >> c;
            std::vector<float> &dst = tag == "v" ? V : VN;
            dst.push_back((float)std::strtod(a.c_str(), nullptr)); dst.push_back((float)std::strtod(b.c_str(), nullptr)); dst.push_back((float)std::strtod(c.c_str(), nullptr));
        } else if (tag == "vt") {
            std::string a, b; ss >> a; if (!(ss >> b)) b = "0";
            VT.push_back((float)std::strtod(a.c_str(), nullptr)); VT.push_back((float)std::strtod(b.c_str(), nullptr));
        } else if (tag == "f") {
            std::vector<Idx> idx;
            std::string c;
            while (ss >> c) {
                int q[3] = {0, 0, 0};
                size_t start = 0;
                for (int k = 0; k < 3 && start <= c.size(); ++k) {
                    size_t e = c.find('/', start);
                    std::string part = c.substr(start, e == std::string::npos ? std::string::npos : e - start);
                    q[k] = part.empty() ? 0 : std::atoi(part.c_str());
                    if (e == std::string::npos) break;
                    start = e + 1;
                }
                int cnt[3] = {(int)V.size() / 3, (int)VT.size() / 2, (int)VN.size() / 3};
                for (int k = 0; k < 3; ++k) q[k] = q[k] > 0 ? q[k] - 1 : (q[k] < 0 ? cnt[k] + q[k] : -1);
                idx.push_back({q[0], q[1], q[2]});
            }
            for (size_t k = 1; k + 1 < idx.size(); ++k) {
                std::array<Idx, 3> tri{idx[0], idx[k], idx[k + 1]};
                if (tri[0].v != tri[1].v && tri[1].v != tri[2].v && tri[0].v != tri[2].v) faces.push_back(tri);
            }
        }
    }
    const bool has_n = !VN.empty();
    std::vector<double> acc;
    if (smooth && !has_n) {
        acc.assign(V.size(), 0.0);
        for (auto &tri : faces) {
            double e1[3], e2[3];
            for (int k = 0; k < 3; ++k) { e1[k] = (double)V[3 * tri[1].v + k] - (double)V[3 * tri[0].v + k]; e2[k] = (double)V[3 * tri[2].v + k] - (double)V[3 * tri[0].v + k]; }
            double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
            for (int c = 0; c < 3; ++c) for (int k = 0; k < 3; ++k) acc[3 * tri[c].v + k] += n[k];
        }
    }
    MeshData m;
    std::map<std::tuple<int, int, int, float, float, float>, uint32_t> seen;
    for (auto &tri : faces) {
        float fn[3] = {0.f, 0.f, 0.f};
        bool face_n = !has_n && acc.empty();
        if (face_n) {
            float e1[3], e2[3], n[3];
            for (int k = 0; k < 3; ++k) { e1[k] = V[3 * tri[1].v + k] - V[3 * tri[0].v + k]; e2[k] = V[3 * tri[2].v + k] - V[3 * tri[0].v + k]; }
            cross3f(e1, e2, n);
            float l = std::sqrt((n[0] * n[0] + n[1] * n[1]) + n[2] * n[2]);
            if (l > 0.f) for (int k = 0; k < 3; ++k) fn[k] = n[k] / l;
        }
        for (int c = 0; c < 3; ++c) {
            const Idx &ix = tri[c];
            auto key = std::make_tuple(ix.v, ix.t, ix.n, fn[0], fn[1], fn[2]);
            auto it = seen.find(key);
            if (it == seen.end()) {
                it = seen.emplace(key, m.nv()).first;
                for (int k = 0; k < 3; ++k) m.pos.push_back(V[3 * ix.v + k]);
                if (ix.n >= 0) for (int k = 0; k < 3; ++k) m.nrm.push_back(VN[3 * ix.n + k]);
                else if (!acc.empty()) {
                    const double *a = &acc[3 * ix.v];
                    double l = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
                    for (int k = 0; k < 3; ++k) m.nrm.push_back((float)(l > 0 ? a[k] / l : a[k]));
                } else for (int k = 0; k < 3; ++k) m.nrm.push_back(fn[k]);
                float u = ix.t >= 0 ? VT[2 * ix.t] : 0.f, v = ix.t >= 0 ? VT[2 * ix.t + 1] : 0.f;
                m.uv.push_back(u); m.uv.push_back(flip_uv ? 1.0f - v : v);
            }
            m.tri.push_back(it->second);
        }
    }
    return m;
}

// reference src/render_core/shape/model.cpp:27-36 ("fn", "swap_handed", "smooth", "flip_uv" default true, "subdiv_level"): one
// instance per imported mesh, in the importer's order. "importer": "obj" forces the in-tree loader (file-order primitive ids).
class Model : public ShapeGroup {
public:
    bool used_assimp = false;
    explicit Model(const ShapeDesc &desc) : ShapeGroup(desc) {
        name_ = desc.name;
        std::string fn = desc["fn"].as_string();
        if (fn.empty()) throw std::runtime_error("model shape " + desc.name + ": no \"fn\"");
        std::string full = fn[0] == '/' ? fn : desc["scene_path"].as_string(".") + "/" + fn;
        bool smooth = desc["smooth"].as_bool(false), flip_uv = desc["flip_uv"].as_bool(true);
        if (assimp_available() && desc["importer"].as_string("assimp") != "obj") {
            for (auto &m : load_model_assimp(full, desc["swap_handed"].as_bool(false), smooth, flip_uv, desc["subdiv_level"].as_uint(0u)))
                add_instance(std::move(m));
            used_assimp = true;
        } else {
            if (desc["swap_handed"].as_bool(false)) throw std::runtime_error("model shape " + desc.name + ": swap_handed needs the assimp importer");
            add_instance(load_obj(full, smooth, flip_uv));
        }
        post_init(desc);
    }
    std::string impl_type() const override { return "model"; }
};
VH_MAKE_CLASS_CREATOR(Model, shape, model, constructor)

// reference src/render_core/shape/sphere.cpp:20-88: theta_div = sub_div rings, phi_div = 2 * sub_div segments
class Sphere : public ShapeGroup {
public:
    explicit Sphere(const ShapeDesc &desc) : ShapeGroup(desc) {
        name_ = desc.name;
        const float Pi = 3.14159265358979323846f, TwoPi = 6.283185307179586f;
        float radius = desc["radius"].as_float(1.f);
        uint32_t theta_div = desc["sub_div"].as_uint(60u), phi_div = 2 * theta_div;
        MeshData m;
        auto push = [&](float x, float y, float z, float nx, float ny, float nz, float u, float v) {
            m.pos.insert(m.pos.end(), {x, y, z}); m.nrm.insert(m.nrm.end(), {nx, ny, nz}); m.uv.insert(m.uv.end(), {u, v});
        };
        auto push_unit = [&](float x, float y, float z, float u, float v) {
            float l = std::sqrt((x * x + y * y) + z * z);
            push(x, y, z, x / l, y / l, z / l, u, v);
        };
        push(0, radius, 0, 0, 1, 0, 0, 0);
        for (uint32_t i = 1; i < theta_div; ++i) {
            float v = (float)i / (float)theta_div, theta = Pi * v;
            float y = radius * std::cos(theta), r = radius * std::sin(theta);
            push_unit(r, y, 0.f, 0.f, v);
            for (uint32_t j = 1; j < phi_div; ++j) {
                float u = (float)j / (float)phi_div, phi = u * TwoPi;
                push_unit(std::cos(phi) * r, y, std::sin(phi) * r, u, v);
            }
        }
        push(0, -radius, 0, 0, -1, 0, 0, 1);
        for (uint32_t i = 0; i < phi_div; ++i) m.tri.insert(m.tri.end(), {0u, (i + 1) % phi_div + 1, i + 1});
        for (uint32_t i = 0; i + 2 < theta_div; ++i) {
            uint32_t vs = 1 + i * phi_div;
            for (uint32_t j = 0; j < phi_div; ++j) {
                if (j != phi_div - 1) { m.tri.insert(m.tri.end(), {vs, vs + 1, vs + phi_div}); m.tri.insert(m.tri.end(), {vs + 1, vs + phi_div + 1, vs + phi_div}); }
                else { m.tri.insert(m.tri.end(), {vs, vs + 1 - phi_div, vs + phi_div}); m.tri.insert(m.tri.end(), {vs + 1 - phi_div, vs + 1, vs + phi_div}); }
                vs++;
            }
        }
        uint32_t ve = m.nv() - 1;
        for (uint32_t i = 0; i < phi_div; ++i) m.tri.insert(m.tri.end(), {ve, ve - ((1 + i) % phi_div + 1), ve - (i + 1)});
        add_instance(std::move(m));
        post_init(desc);
    }
    std::string impl_type() const override { return "sphere"; }
};
VH_MAKE_CLASS_CREATOR(Sphere, shape, sphere, constructor)

// ---- lights ------------------------------------------------------------------------------------
// Light ctor + initialize_slots (light.cpp:10-24): colour slot, scale, constant colours normalised into the scale
static void light_common(vb_light_rec &r, const LightDesc &d, float3 def, const SlotContext &ctx) {
    std::memset(&r, 0, sizeof(r));
    r.inst_id = VB_INVALID_ID;
    r.color = make_slot(d["color"], def, ctx);
    float scale = d.has_attr("scale") ? make_slot(d["scale"], {1.f, 1.f, 1.f}, ctx).value[0] : 1.f;
    if (r.color.tex_id == VB_INVALID_ID) {                 // NumberArray::normalize (number.cpp:40-49)
        float mx = std::max(r.color.value[0], std::max(r.color.value[1], r.color.value[2]));
        if (!(mx < 1.f)) { for (int i = 0; i < 3; ++i) r.color.value[i] /= mx; scale *= mx; }
    }
    r.scale = scale;
}

// reference src/render_core/light/point.cpp:24-27
class PointLight : public Light {
protected:
    LightDesc desc_;
public:
    explicit PointLight(const LightDesc &d) : Light(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "point"; }
    int type_order() const override { return VB_LIGHT_POINT; }
    void fill_record(vb_light_rec &r, const Scene &, const SlotContext &ctx, std::vector<vb_alias_entry> &, std::vector<float> &) const override {
        light_common(r, desc_, {1.f, 1.f, 1.f}, ctx);
        r.type = VB_LIGHT_POINT;
        float3 p = desc_["position"].as_float3({0, 0, 0});
        r.position[0] = p.x; r.position[1] = p.y; r.position[2] = p.z;
    }
};
VH_MAKE_CLASS_CREATOR(PointLight, light, point, constructor)

// reference src/render_core/light/spot.cpp:29-35,47-55: angle clamped to [1, 89] deg, falloff to [0, angle]
class SpotLight : public PointLight {
public:
    explicit SpotLight(const LightDesc &d) : PointLight(d) {}
    std::string impl_type() const override { return "spot"; }
    int type_order() const override { return VB_LIGHT_SPOT; }
    void fill_record(vb_light_rec &r, const Scene &s, const SlotContext &ctx, std::vector<vb_alias_entry> &t, std::vector<float> &f) const override {
        PointLight::fill_record(r, s, ctx, t, f);
        r.type = VB_LIGHT_SPOT;
        auto radians = [](float deg) { return (float)((double)deg * 3.14159265358979323846 / 180.0); };
        float angle = radians(std::min(std::max(desc_["angle"].as_float(45.f), 1.f), 89.f));
        float falloff = radians(std::min(std::max(desc_["falloff"].as_float(10.f), 0.f), angle));
        float3 d = desc_["direction"].as_float3({0, 0, 1});
        float inv = 1.0f / std::sqrt((d.x * d.x + d.y * d.y) + d.z * d.z);
        r.direction[0] = d.x * inv; r.direction[1] = d.y * inv; r.direction[2] = d.z * inv;
        r.params[0] = std::cos(angle);
        shape_angle = angle; shape_falloff = falloff;
        r.params[1] = std::cos(std::max(0.f, angle - falloff));
    }
};
VH_MAKE_CLASS_CREATOR(SpotLight, light, spot, constructor)

static bool invert4(const double *m, double *inv);

// reference src/render_core/light/environments/directional.cpp:20-38: a delta Infinite light; "direction" is stored as written
class DirectionalLight : public Light {
    LightDesc desc_;
public:
    explicit DirectionalLight(const LightDesc &d) : Light(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "directional"; }
    int type_order() const override { return VB_LIGHT_DIRECTIONAL; }
    void fill_record(vb_light_rec &r, const Scene &scene, const SlotContext &ctx, std::vector<vb_alias_entry> &, std::vector<float> &) const override {
        light_common(r, desc_, {1.f, 1.f, 1.f}, ctx);
        r.type = VB_LIGHT_DIRECTIONAL;
        float3 d = desc_["direction"].as_float3({-1.f, -1.f, -1.f});
        r.direction[0] = d.x; r.direction[1] = d.y; r.direction[2] = d.z;
        r.params[0] = scene.world_diameter();
        r.params[1] = r.params[0] * 0.5f;                               // world_radius (scene.h:108-109: diameter = radius * 2)
    }
};
VH_MAKE_CLASS_CREATOR(DirectionalLight, light, directional, constructor)

// reference src/render_core/light/environments/spherical.cpp:36-45 (transform), :170-196 (weights), :199-211 +
// warper/alias2d.cpp:32-67 (AliasTable2D::build: one table per row, marginal over the row integrals)
class SphericalMap : public Light {
    LightDesc desc_;
public:
    explicit SphericalMap(const LightDesc &d) : Light(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "spherical"; }
    int type_order() const override { return VB_LIGHT_ENV; }
    void fill_record(vb_light_rec &r, const Scene &scene, const SlotContext &ctx, std::vector<vb_alias_entry> &table,
                     std::vector<float> &func) const override {
        light_common(r, desc_, {1.f, 1.f, 1.f}, ctx);
        r.type = VB_LIGHT_ENV;
        // o2w = desc.o2w * rotation_x(-90) * scale(1, flip_u ? 1 : -1, 1); products in double from the float o2w, one rounding
        const Mat4 &M = desc_.o2w.mat;
        const double kPi = 3.14159265358979323846;
        double a = -90.0 * kPi / 180.0, c = std::cos(a), s = std::sin(a), fy = desc_["flip_u"].as_bool(false) ? 1.0 : -1.0;
        double A[16], RX[16] = {1, 0, 0, 0, 0, c, s, 0, 0, -s, c, 0, 0, 0, 0, 1}, SC[16] = {1, 0, 0, 0, 0, fy, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, T[16], O[16];
        for (int i = 0; i < 16; ++i) A[i] = M[i];
        auto mul = [](const double *x, const double *y, double *o) {       // column-major
            for (int cc = 0; cc < 4; ++cc) for (int rr = 0; rr < 4; ++rr) { double v = 0; for (int k = 0; k < 4; ++k) v += x[4 * k + rr] * y[4 * cc + k]; o[4 * cc + rr] = v; }
        };
        mul(A, RX, T); mul(T, SC, O);
        double Of[16], I[16];
        for (int i = 0; i < 16; ++i) { r.o2w[i] = (float)O[i]; Of[i] = r.o2w[i]; }
        if (!invert4(Of, I)) throw std::runtime_error("spherical light: singular o2w");
        for (int i = 0; i < 16; ++i) r.w2o[i] = (float)I[i];
        // importance table over the texels
        std::vector<float> weights;
        uint32_t rx = 1, ry = 1;
        if (r.color.tex_id != VB_INVALID_ID) {
            const Image &im = ctx.images->images[r.color.tex_id];
            rx = im.w; ry = im.h;
            weights.resize((size_t)rx * ry);
            for (size_t idx = 0; idx < weights.size(); ++idx) {
                const float *px = &im.rgba[4 * idx];
                float lum = (0.212671f * px[0] + 0.715160f * px[1]) + 0.072169f * px[2];
                float v = (float)(idx / ry) + 0.5f;                     // the reference divides by res.y and normalises by res.x
                float theta = v / (float)rx;
                weights[idx] = lum * std::sin(3.14159265358979323846f * theta);
            }
        } else weights.push_back(1.f);
        std::vector<vb_alias_entry> ct, mt; std::vector<float> cf, mf, marginal;
        for (uint32_t v = 0; v < ry; ++v) {
            std::vector<float> row(weights.begin() + (size_t)v * rx, weights.begin() + (size_t)(v + 1) * rx);
            bool any = false;
            for (float w : row) any |= w != 0.f;
            float integral = 0.f;
            if (any) scene.warper->build(row, ct, cf, integral);
            else { for (uint32_t i = 0; i < rx; ++i) { ct.push_back({1.0f, i}); cf.push_back(0.f); } }
            marginal.push_back(integral);
        }
        r.alias_offset = (uint32_t)table.size(); r.alias_size = ry;
        scene.warper->build(marginal, table, func, r.alias_integral);
        r.cond_offset = (uint32_t)table.size();
        table.insert(table.end(), ct.begin(), ct.end()); func.insert(func.end(), cf.begin(), cf.end());
        r.env_res[0] = rx; r.env_res[1] = ry;
        r.params[0] = scene.world_diameter();
    }
};
VH_MAKE_CLASS_CREATOR(SphericalMap, light, spherical, constructor)

// reference src/render_core/light/projector.cpp:33-52,96-113: a point light at o2w[3] that projects its colour image along
// the local +z axis; angle clamped to [1, 89] deg, ratio 0 = image width / height
class Projector : public Light {
    LightDesc desc_;
public:
    explicit Projector(const LightDesc &d) : Light(d), desc_(d) { name_ = d.name; }
    std::string impl_type() const override { return "projector"; }
    int type_order() const override { return VB_LIGHT_PROJECTOR; }
    void fill_record(vb_light_rec &r, const Scene &, const SlotContext &ctx, std::vector<vb_alias_entry> &, std::vector<float> &) const override {
        light_common(r, desc_, {1.f, 1.f, 1.f}, ctx);
        r.type = VB_LIGHT_PROJECTOR;
        const Mat4 &M = desc_.o2w.mat;
        double Of[16], I[16];
        for (int i = 0; i < 16; ++i) { r.o2w[i] = M[i]; Of[i] = M[i]; }
        if (!invert4(Of, I)) throw std::runtime_error("projector light: singular o2w");
        for (int i = 0; i < 16; ++i) r.w2o[i] = (float)I[i];
        r.position[0] = M[12]; r.position[1] = M[13]; r.position[2] = M[14];
        r.direction[0] = M[8]; r.direction[1] = M[9]; r.direction[2] = M[10];          // transform_vector(o2w, (0, 0, 1))
        float ratio = desc_["ratio"].as_float(1.f);
        if (ratio == 0.f && r.color.tex_id != VB_INVALID_ID) {
            const Image &im = ctx.images->images[r.color.tex_id];
            ratio = (float)im.w / (float)im.h;
        }
        float angle = (float)((double)std::min(std::max(desc_["angle"].as_float(45.f), 1.f), 89.f) * 3.14159265358979323846 / 180.0);
        r.params[0] = std::tan(angle);
        r.params[1] = ratio;
    }
};
VH_MAKE_CLASS_CREATOR(Projector, light, projector, constructor)

// Gauss-Jordan with partial pivoting, column-major in/out
static bool invert4(const double *m, double *inv) {
    double a[4][8];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { a[r][c] = m[4 * c + r]; a[r][4 + c] = r == c ? 1.0 : 0.0; }
    for (int i = 0; i < 4; ++i) {
        int p = i;
        for (int r = i + 1; r < 4; ++r) if (std::fabs(a[r][i]) > std::fabs(a[p][i])) p = r;
        if (a[p][i] == 0.0) return false;
        if (p != i) for (int c = 0; c < 8; ++c) std::swap(a[i][c], a[p][c]);
        double d = a[i][i];
        for (int c = 0; c < 8; ++c) a[i][c] /= d;
        for (int r = 0; r < 4; ++r) if (r != i) { double f = a[r][i]; if (f != 0.0) for (int c = 0; c < 8; ++c) a[r][c] -= f * a[i][c]; }
    }
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) inv[4 * c + r] = a[r][4 + c];
    return true;
}

}  // namespace vh

#include "vh_desc.h"
#include <algorithm>
#include <cmath>

namespace vh {

Mat4 mat4_identity() { Mat4 m{}; m[0] = m[5] = m[10] = m[15] = 1.f; return m; }
Mat4 mat4_mul(const Mat4 &a, const Mat4 &b) {
    Mat4 r{};
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 4; ++row) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += (double)a[4 * k + row] * (double)b[4 * c + k];
            r[4 * c + row] = (float)s;
        }
    return r;
}
static float3 norm3(float3 v) { double l = std::sqrt((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z); return {(float)(v.x / l), (float)(v.y / l), (float)(v.z / l)}; }
static float3 cross3(float3 a, float3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// left-handed camera frame [right, up, forward, position]; forward = target - pos
Mat4 mat4_inverse(const Mat4 &m) {
    double a[4][8];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { a[r][c] = m[c * 4 + r]; a[r][4 + c] = r == c ? 1.0 : 0.0; }
    for (int i = 0; i < 4; ++i) {
        int piv = i;
        for (int r = i + 1; r < 4; ++r) if (std::fabs(a[r][i]) > std::fabs(a[piv][i])) piv = r;
        if (!(std::fabs(a[piv][i]) > 0.0)) throw std::runtime_error("mat4_inverse: singular matrix");
        if (piv != i) for (int c = 0; c < 8; ++c) std::swap(a[i][c], a[piv][c]);
        const double inv = 1.0 / a[i][i];
        for (int c = 0; c < 8; ++c) a[i][c] *= inv;
        for (int r = 0; r < 4; ++r) if (r != i) { const double f = a[r][i]; for (int c = 0; c < 8; ++c) a[r][c] -= f * a[i][c]; }
    }
    Mat4 out{};
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out[c * 4 + r] = (float)a[r][4 + c];
    return out;
}
Mat4 look_at(float3 pos, float3 target, float3 up) {
    float3 fwd = norm3({target.x - pos.x, target.y - pos.y, target.z - pos.z});
    float3 right = norm3(cross3(fwd, up));
    float3 up2 = cross3(right, fwd);
    Mat4 m = mat4_identity();
    m[0] = right.x; m[1] = right.y; m[2] = right.z;
    m[4] = up2.x; m[5] = up2.y; m[6] = up2.z;
    m[8] = fwd.x; m[9] = fwd.y; m[10] = fwd.z;
    m[12] = pos.x; m[13] = pos.y; m[14] = pos.z;
    return m;
}
static const double kDeg = 3.14159265358979323846 / 180.0;
Mat4 rotation_x(float deg) { double a = deg * kDeg, c = std::cos(a), s = std::sin(a); Mat4 m = mat4_identity(); m[5] = (float)c; m[6] = (float)s; m[9] = (float)-s; m[10] = (float)c; return m; }
Mat4 rotation_y(float deg) { double a = deg * kDeg, c = std::cos(a), s = std::sin(a); Mat4 m = mat4_identity(); m[0] = (float)c; m[2] = (float)-s; m[8] = (float)s; m[10] = (float)c; return m; }
Mat4 rotation_z(float deg) { double a = deg * kDeg, c = std::cos(a), s = std::sin(a); Mat4 m = mat4_identity(); m[0] = (float)c; m[1] = (float)s; m[4] = (float)-s; m[5] = (float)c; return m; }
Mat4 translation(float3 t) { Mat4 m = mat4_identity(); m[12] = t.x; m[13] = t.y; m[14] = t.z; return m; }
Mat4 scaling(float3 s) { Mat4 m = mat4_identity(); m[0] = s.x; m[5] = s.y; m[10] = s.z; return m; }

float3 ParameterSet::as_float3(float3 d) const {
    if (j_.is_array() && j_.arr.size() >= 3) return {(float)j_.arr[0].num, (float)j_.arr[1].num, (float)j_.arr[2].num};
    if (j_.is_number()) return {(float)j_.num, (float)j_.num, (float)j_.num};
    return d;
}
std::vector<float> ParameterSet::as_floats() const {
    std::vector<float> v;
    if (j_.is_array()) for (auto &e : j_.arr) v.push_back((float)e.num);
    else if (j_.is_number()) v.push_back((float)j_.num);
    return v;
}
Mat4 ParameterSet::as_float4x4(const Mat4 &d) const {
    if (!j_.is_array() || j_.arr.size() != 4) return d;
    Mat4 m{};
    for (int c = 0; c < 4; ++c) {
        if (!j_.arr[c].is_array() || j_.arr[c].arr.size() != 4) return d;
        for (int r = 0; r < 4; ++r) m[4 * c + r] = (float)j_.arr[c].arr[r].num;
    }
    return m;
}

static std::string lower(std::string s) { std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::tolower(c); }); return s; }
std::string NodeDesc::plugin_name() const { return "vision-" + lower(type) + "-" + lower(sub_type); }
void NodeDesc::init_common(const ParameterSet &ps) {
    if (ps.data().is_object()) {
        std::string fn = ps["construct_name"].as_string("");
        if (!fn.empty()) construct_name = fn;
        name = ps["name"].as_string();
    }
}

// reference node_desc.cpp:31-58
void TransformDesc::init(const ParameterSet &ps) {
    if (ps.data().is_null()) return;
    sub_type = ps["type"].as_string("matrix4x4");
    ParameterSet param = ps.value("param");
    if (sub_type == "look_at") {
        mat = look_at(param["position"].as_float3({0, 0, 0}), param["target_pos"].as_float3({0, 0, 1}), param["up"].as_float3({0, 1, 0}));
    } else if (sub_type == "Euler") {
        Mat4 yaw = rotation_y(param["yaw"].as_float(0.f)), roll = rotation_z(param["roll"].as_float(0.f)), pitch = rotation_x(param["pitch"].as_float(0.f));
        mat = mat4_mul(mat4_mul(mat4_mul(translation(param["position"].as_float3({0, 0, 0})), pitch), roll), yaw);
    } else if (sub_type == "trs") {
        float3 t = param["t"].as_float3({0, 0, 0}), s = param["s"].as_float3({1, 1, 1});
        std::vector<float> r = param["r"].as_floats();      // axis-angle (x,y,z,deg)
        Mat4 R = mat4_identity();
        if (r.size() == 4 && r[3] != 0.f) {
            float3 a = norm3({r[0], r[1], r[2]});
            double ang = r[3] * kDeg, c = std::cos(ang), sn = std::sin(ang), k = 1 - c;
            R[0] = (float)(c + a.x * a.x * k); R[1] = (float)(a.y * a.x * k + a.z * sn); R[2] = (float)(a.z * a.x * k - a.y * sn);
            R[4] = (float)(a.x * a.y * k - a.z * sn); R[5] = (float)(c + a.y * a.y * k); R[6] = (float)(a.z * a.y * k + a.x * sn);
            R[8] = (float)(a.x * a.z * k + a.y * sn); R[9] = (float)(a.y * a.z * k - a.x * sn); R[10] = (float)(c + a.z * a.z * k);
        }
        mat = mat4_mul(mat4_mul(translation(t), R), scaling(s));
    } else if (sub_type == "matrix4x4") {
        mat = param["matrix4x4"].as_float4x4(mat4_identity());
    } else {
        throw std::runtime_error("transform type error " + sub_type);       // reference: OC_ERROR
    }
}
void FilterDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("gaussian"); parameter = ps.value("param"); }
void ToneMapperDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = "impl"; construct_name = ps["type"].as_string("linear"); parameter = ps.value("param"); }
void LightDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("area"); parameter = ps.value("param"); o2w.init(parameter.value("o2w")); }
void ShapeDesc::init(const ParameterSet &ps) {
    init_common(ps);
    sub_type = ps["type"].as_string(); name = ps["name"].as_string();
    parameter = ps["param"];
    o2w.init(parameter.value("transform"));
    if (parameter.contains("emission")) { emission.init(parameter["emission"]); has_emission = true; }
    if (parameter.contains("medium")) {
        has_medium = true;
        inside_medium = parameter["medium"]["inside"].as_string();
        outside_medium = parameter["medium"]["outside"].as_string();
    }
}
void MediumDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("homogeneous"); parameter = ps.value("param"); }
void MediumsDesc::init(const Json &data) {
    // scene_desc.cpp:26-34: "process" defaults to TRUE; a missing key behaves like an empty object. A bare
    // list (res/cbox-glass.json) carries no "list" key: nothing is loaded.
    process = true;
    if (!data.is_object()) return;
    ParameterSet ps(data);
    process = ps["process"].as_bool(true);
    global = ps["global"].as_string("");
    const Json &l = data.at("list");
    if (l.is_array()) for (auto &e : l.arr) { MediumDesc d; d.init(ParameterSet(e)); mediums.push_back(d); }
}
void MaterialDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("diffuse"); parameter = ps["param"]; }
void SamplerDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("independent"); parameter = ps.value("param"); }
void SensorDesc::init(const ParameterSet &ps) {
    init_common(ps);
    sub_type = ps["type"].as_string("thin_lens");
    parameter = ps.value("param");
    transform_desc.init(parameter.value("transform"));
    filter_desc.init(parameter.value("filter"));
}
void IntegratorDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("pt"); parameter = ps.value("param"); }
void LightSamplerDesc::init(const ParameterSet &ps) {
    init_common(ps);
    sub_type = ps["type"].as_string("uniform");
    parameter = ps.value("param");
    const ParameterSet lights = parameter["lights"];        // keep the temporary alive: data() returns a reference into it
    const Json &l = lights.data();
    if (l.is_array()) for (auto &e : l.arr) { LightDesc d; d.init(ParameterSet(e)); light_descs.push_back(d); }
}
void SpectrumDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("srgb"); parameter = ps.value("param"); }
void FrameBufferDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("normal"); parameter = ps.value("param"); tone_mapper.init(parameter.value("tone_mapper")); }
void PipelineDesc::init(const ParameterSet &ps) { init_common(ps); sub_type = ps["type"].as_string("fixed"); parameter = ps.value("param"); frame_buffer_desc.init(parameter.value("frame_buffer")); }
void OutputDesc::init(const ParameterSet &ps) {
    if (ps.data().is_null()) return;
    spp = ps["spp"].as_uint(0u); save_exit = ps["save_exit"].as_uint(0u); fn = ps["fn"].as_string("output.png"); denoise = ps["denoise"].as_bool(false);
}
void RenderSettingDesc::init(const ParameterSet &ps) {
    polymorphic_mode = ps["polymorphic_mode"].as_uint(1); min_world_radius = ps["min_world_radius"].as_float(10.f);
    ray_offset_factor = ps["ray_offset_factor"].as_float(1.f);
}

void SceneDesc::init(const Json &data) {
    ParameterSet root(data);
    integrator_desc.init(root.value("integrator"));
    spectrum_desc.init(root.value("spectrum"));
    light_sampler_desc.init(root.value("light_sampler"));
    sampler_desc.init(root.value("sampler"));
    const Json &mats = data.at("materials");
    if (mats.is_array()) for (auto &m : mats.arr) { MaterialDesc d; d.init(ParameterSet(m)); material_descs.push_back(d); }
    const Json &med = data.at("mediums");
    mediums_desc.init(med);
    process_mediums = mediums_desc.process;
    const Json &shapes = data.at("shapes");
    if (shapes.is_array()) for (auto &s : shapes.arr) { ShapeDesc d; d.init(ParameterSet(s)); d.parameter.set_value("scene_path", Json::string(scene_path)); shape_descs.push_back(d); }
    sensor_desc.init(root.value("camera"));
    output_desc.init(root.value("output"));
    render_setting.init(root.value("render_setting"));
    pipeline_desc.init(root.value("pipeline"));
    static const char *known[] = {"integrator", "spectrum", "light_sampler", "sampler", "materials", "mediums", "shapes", "camera",
                                  "output", "render_setting", "pipeline", "denoiser", "warper"};
    if (data.is_object())
        for (auto &kv : data.obj) {
            bool k = false;
            for (auto n : known) k |= kv.first == n;
            if (!k || kv.first == "denoiser") ignored_keys.push_back(kv.first);
        }
}

SceneDesc SceneDesc::from_json(const std::string &path) {
    SceneDesc d;
    size_t p = path.find_last_of("/\\");
    d.scene_path = p == std::string::npos ? "." : path.substr(0, p);
    d.init(Json::parse_file(path));
    return d;
}

}  // namespace vh

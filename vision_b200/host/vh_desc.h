// Descriptor layer of the host side: the scene API the reference exposes through vision_scene.json.
// Mirrors reference src/base/import/{parameter_set.h,node_desc.h,node_desc.cpp,scene_desc.cpp}:
// same JSON keys, same defaults at the point of use, same plugin-name resolution
// ("vision-" + lower(type) + "-" + lower(sub_type), node_desc.h:95-97).
#pragma once
#include "vh_json.h"
#include <array>
#include <cstdint>

namespace vh {

struct float3 { float x, y, z; };
using Mat4 = std::array<float, 16>;     // column-major: m[4*c + r], column 3 = translation
Mat4 mat4_identity();
Mat4 mat4_mul(const Mat4 &a, const Mat4 &b);
Mat4 mat4_inverse(const Mat4 &m);                        // Gauss-Jordan with row pivoting in double precision, rounded once
Mat4 look_at(float3 pos, float3 target, float3 up);       // columns: right, up, forward, position
Mat4 rotation_x(float deg), rotation_y(float deg), rotation_z(float deg);
Mat4 translation(float3 t);
Mat4 scaling(float3 s);

// typed view over a JSON value with defaults at the point of use (reference parameter_set.h)
class ParameterSet {
    Json j_;
public:
    ParameterSet() = default;
    explicit ParameterSet(Json j) : j_(std::move(j)) {}
    const Json &data() const { return j_; }
    ParameterSet operator[](const std::string &k) const { return ParameterSet(j_.at(k)); }
    ParameterSet operator[](size_t i) const { return ParameterSet(j_.is_array() && i < j_.arr.size() ? j_.arr[i] : Json()); }
    ParameterSet value(const std::string &k) const { return (*this)[k]; }
    bool contains(const std::string &k) const { return j_.contains(k); }
    void set_value(const std::string &k, Json v) { j_.set(k, std::move(v)); }
    float as_float(float d = 0.f) const { return j_.is_number() ? (float)j_.num : (j_.is_array() && !j_.arr.empty() && j_.arr[0].is_number() ? (float)j_.arr[0].num : d); }
    uint32_t as_uint(uint32_t d = 0) const { return j_.is_number() ? (uint32_t)j_.num : (j_.kind == Json::Bool ? (uint32_t)j_.b : d); }
    int as_int(int d = 0) const { return j_.is_number() ? (int)j_.num : d; }
    bool as_bool(bool d = false) const { return j_.kind == Json::Bool ? j_.b : (j_.is_number() ? j_.num != 0 : d); }
    std::string as_string(const std::string &d = "") const { return j_.is_string() ? j_.str : d; }
    float3 as_float3(float3 d) const;
    std::vector<float> as_floats() const;
    Mat4 as_float4x4(const Mat4 &d) const;     // JSON rows are columns
};

struct NodeDesc {
    std::string type;               // "Shape", "Material", ... (VISION_DESC_COMMON)
    std::string sub_type;           // plugin name inside the category
    std::string construct_name = "constructor";
    std::string name;
    ParameterSet parameter;
    explicit NodeDesc(std::string t = "") : type(std::move(t)) {}
    ParameterSet operator[](const std::string &k) const { return parameter[k]; }
    bool has_attr(const std::string &k) const { return parameter.contains(k); }
    std::string plugin_name() const;      // "vision-<type>-<sub_type>" lower-cased
    void init_common(const ParameterSet &ps);
};

struct TransformDesc : NodeDesc { Mat4 mat = mat4_identity(); TransformDesc() : NodeDesc("Transform") {} void init(const ParameterSet &ps); };
struct FilterDesc : NodeDesc { FilterDesc() : NodeDesc("Filter") {} void init(const ParameterSet &ps); };
struct ToneMapperDesc : NodeDesc { ToneMapperDesc() : NodeDesc("ToneMapper") {} void init(const ParameterSet &ps); };
struct LightDesc : NodeDesc { TransformDesc o2w; LightDesc() : NodeDesc("Light") {} void init(const ParameterSet &ps); bool valid() const { return !sub_type.empty() && !parameter.data().is_null(); } };
struct ShapeDesc : NodeDesc {
    TransformDesc o2w; LightDesc emission; bool has_emission = false;
    bool has_medium = false; std::string inside_medium, outside_medium;      // "medium": {"inside","outside"} (shape.cpp:247-249)
    ShapeDesc() : NodeDesc("Shape") {}
    void init(const ParameterSet &ps);
};
// reference node_desc.cpp:182-197 (medium_name looks up the measured table, medium_scatter_data.h)
struct MediumDesc : NodeDesc { MediumDesc() : NodeDesc("Medium") {} void init(const ParameterSet &ps); };
// reference scene_desc.cpp:24-34: "mediums": {"process", "global", "list": [...]}
struct MediumsDesc { bool process = false; std::string global; std::vector<MediumDesc> mediums; void init(const Json &data); };
struct MaterialDesc : NodeDesc { MaterialDesc() : NodeDesc("Material") {} void init(const ParameterSet &ps); };
struct SamplerDesc : NodeDesc { SamplerDesc() : NodeDesc("Sampler") {} void init(const ParameterSet &ps); };
struct SensorDesc : NodeDesc { TransformDesc transform_desc; FilterDesc filter_desc; SensorDesc() : NodeDesc("Sensor") {} void init(const ParameterSet &ps); };
struct IntegratorDesc : NodeDesc { IntegratorDesc() : NodeDesc("Integrator") {} void init(const ParameterSet &ps); };
struct LightSamplerDesc : NodeDesc { std::vector<LightDesc> light_descs; LightSamplerDesc() : NodeDesc("LightSampler") {} void init(const ParameterSet &ps); };
struct SpectrumDesc : NodeDesc { SpectrumDesc() : NodeDesc("Spectrum") {} void init(const ParameterSet &ps); };
struct WarperDesc : NodeDesc { WarperDesc() : NodeDesc("Warper") { sub_type = "alias"; } };
struct FrameBufferDesc : NodeDesc { ToneMapperDesc tone_mapper; FrameBufferDesc() : NodeDesc("FrameBuffer") {} void init(const ParameterSet &ps); };
struct PassDesc : NodeDesc { explicit PassDesc(const std::string &type) : NodeDesc("Pass") { sub_type = type; } };   // RenderPass::create(name)
struct PipelineDesc : NodeDesc { FrameBufferDesc frame_buffer_desc; PipelineDesc() : NodeDesc("Pipeline") {} void init(const ParameterSet &ps); };
struct OutputDesc { uint32_t spp = 0; uint32_t save_exit = 0; std::string fn = "output.png"; bool denoise = false; void init(const ParameterSet &ps); };
struct RenderSettingDesc { uint32_t polymorphic_mode = 1; float min_world_radius = 10.f; float ray_offset_factor = 1.f; void init(const ParameterSet &ps); };

// reference src/base/import/scene_desc.cpp:37-61
struct SceneDesc {
    std::string scene_path;         // directory of the scene file
    IntegratorDesc integrator_desc; SpectrumDesc spectrum_desc; LightSamplerDesc light_sampler_desc; SamplerDesc sampler_desc;
    WarperDesc warper_desc; std::vector<MaterialDesc> material_descs; std::vector<ShapeDesc> shape_descs;
    SensorDesc sensor_desc; OutputDesc output_desc; RenderSettingDesc render_setting; PipelineDesc pipeline_desc;
    MediumsDesc mediums_desc;
    bool process_mediums = false;
    std::vector<std::string> ignored_keys;     // keys present in the file that this path does not consume (logged)
    void init(const Json &data);
    static SceneDesc from_json(const std::string &path);
};

}  // namespace vh

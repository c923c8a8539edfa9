// Plugin ABI and plugin interfaces of the host layer (C++), mirroring the reference's:
//   INodeConstructor / NodeConstructor<T> / Node::create_shared      reference src/base/node.h:68-118
//   VS_MAKE_CLASS_CREATOR (extern "C" constructor symbol per plugin)  reference src/base/node.h:236-246
//   Integrator / Sampler / Filter / Sensor / Material / Light / LightSampler / Spectrum / ToneMapper /
//   Warper / ShapeGroup / FrameBuffer / Pipeline                      reference src/base/**
// Device-side virtuals of the reference return ocarina DSL nodes (Float, Uint) and cannot exist
// without the DSL; here each plugin's device part is "fill one POD record" (include/vision_b200_pt.h).
// Plugins live in ONE shared library, so besides the in-process registry every plugin also exports
// `extern "C" <plugin_name_with_underscores>__<construct_name>()` returning its INodeConstructor.
#pragma once
#include "vh_desc.h"
#include <cmath>
#include "../../include/vision_b200.h"
#include <functional>
#include <memory>

namespace vh {

class Scene;
class Pipeline;

class Node {
public:
    std::string name_;
    virtual ~Node() = default;
    virtual std::string impl_type() const = 0;           // VS_MAKE_PLUGIN_NAME_FUNC
    virtual void prepare() {}
    const std::string &name() const { return name_; }
};

class INodeConstructor {
public:
    virtual ~INodeConstructor() = default;
    virtual Node *construct_impl(const NodeDesc *desc) const = 0;
    std::shared_ptr<Node> construct_shared_impl(const NodeDesc *desc) const { return std::shared_ptr<Node>(construct_impl(desc)); }
};
template <class T> class NodeConstructor : public INodeConstructor {
public:
    Node *construct_impl(const NodeDesc *desc) const override { return new T(*static_cast<const typename T::Desc *>(desc)); }
};

// in-process stand-in for RHIContext::obtain_module(desc.plugin_name())->function(construct_name)
class PluginRegistry {
public:
    static PluginRegistry &instance();
    void add(const std::string &plugin_name, const std::string &construct_name, INodeConstructor *c);
    INodeConstructor *find(const std::string &plugin_name, const std::string &construct_name) const;
    std::vector<std::string> names() const;
private:
    std::vector<std::pair<std::string, INodeConstructor *>> entries_;
};

template <class T, class D> std::shared_ptr<T> create_shared(const D &desc) {
    INodeConstructor *c = PluginRegistry::instance().find(desc.plugin_name(), desc.construct_name);
    if (!c) throw std::runtime_error("error node load " + desc.plugin_name() + " (" + desc.construct_name + ")");   // OC_ERROR
    auto p = std::static_pointer_cast<T>(c->construct_shared_impl(&desc));
    p->name_ = desc.name;
    return p;
}

#define VH_CONCAT_(a, b) a##b
#define VH_CONCAT(a, b) VH_CONCAT_(a, b)
// registers Class under "vision-<category>-<name>" and exports the C symbol
#define VH_MAKE_CLASS_CREATOR(Class, category, plugname, ctor)                                                  \
    extern "C" VB_API vh::INodeConstructor *VH_CONCAT(VH_CONCAT(vision_##category##_##plugname, __), ctor)() {  \
        static vh::NodeConstructor<Class> ret;                                                                 \
        return &ret;                                                                                           \
    }                                                                                                          \
    static const bool VH_CONCAT(vh_reg_, __COUNTER__) = [] {                                                      \
        vh::PluginRegistry::instance().add("vision-" #category "-" #plugname, #ctor,                           \
                                           VH_CONCAT(VH_CONCAT(vision_##category##_##plugname, __), ctor)());  \
        return true;                                                                                           \
    }();

// ---------------------------------------------------------------------------------------------
struct MeshData {                  // reference vision::Mesh (src/base/shape.h:78-112)
    std::vector<float> pos, nrm, uv;
    std::vector<uint32_t> tri;
    uint32_t nv() const { return (uint32_t)pos.size() / 3; }
    uint32_t nt() const { return (uint32_t)tri.size() / 3; }
    std::vector<float> surface_areas() const;       // object space, shape.cpp:210-219
    uint64_t hash() const;
};

class Material;
class Light;
struct ShapeInstance {             // reference src/base/shape.h:148-192
    std::shared_ptr<MeshData> mesh;
    Mat4 o2w = mat4_identity();
    std::string name, material_name;
    std::shared_ptr<Material> material;
    std::shared_ptr<Light> emission;
    bool has_medium = false;                         // shape desc carried a "medium" block
    std::string inside_name, outside_name;
};

class ShapeGroup : public Node {   // quad / cube / sphere / model
public:
    using Desc = ShapeDesc;
    std::vector<std::shared_ptr<ShapeInstance>> instances;
    explicit ShapeGroup(const ShapeDesc &) {}
    void post_init(const ShapeDesc &desc);           // shape.cpp:245-270
    void add_instance(MeshData mesh);
};

class ImagePool;
// spectrum_complete / spds: Spectrum::is_complete() (hero) makes glass / metal create "spd" slots (glass.cpp:221-224,
// metal.cpp:117-121); their curves are appended to the scene's SPD pool
struct SlotContext { ImagePool *images; std::string scene_path; bool spectrum_complete = false; std::vector<std::vector<float>> *spds = nullptr; };

class Material : public Node {
public:
    using Desc = MaterialDesc;
    uint32_t index = VB_INVALID_ID;
    explicit Material(const MaterialDesc &) {}
    virtual void fill_record(vb_material_rec &rec, const SlotContext &ctx) const = 0;
    // mix / add: the two child materials (mix.cpp:21-23, add.cpp:21-23). Scene::prepare gives them record indices after the scene's own
    // materials; no instance refers to a child directly.
    virtual std::vector<std::shared_ptr<Material>> children() const { return {}; }
};

class Medium : public Node {       // reference src/base/scattering/medium.h:23-50
public:
    using Desc = MediumDesc;
    uint32_t index = VB_INVALID_ID;
    explicit Medium(const MediumDesc &) {}
    virtual void fill_record(vb_medium_rec &rec) const = 0;
};

// numeric data shipped under vision_b200/data (tools/extract_*.py); directory of this library + "/data"
std::string data_dir();

class Light : public Node {
public:
    using Desc = LightDesc;
    uint32_t index = VB_INVALID_ID;
    uint32_t inst_idx = VB_INVALID_ID;               // IAreaLight::inst_idx_
    mutable float shape_angle = 0.f, shape_falloff = 0.f;   // spot: radians, kept by fill_record for power()
    explicit Light(const LightDesc &desc) : inst_idx(desc["inst_id"].as_uint(VB_INVALID_ID)) {}
    virtual int type_order() const = 0;              // Polymorphic topology order = first-seen type
    virtual void fill_record(vb_light_rec &rec, const Scene &scene, const SlotContext &ctx,
                             std::vector<vb_alias_entry> &table, std::vector<float> &func) const = 0;
};

class LightSampler : public Node {
public:
    using Desc = LightSamplerDesc;
    std::vector<std::shared_ptr<Light>> lights;
    bool env_separate;                                // lightsampler.cpp:11-13
    float env_prob;
    explicit LightSampler(const LightSamplerDesc &d);
    void add_light(std::shared_ptr<Light> l) { lights.push_back(std::move(l)); }
    void tidy_up();                                   // lightsampler.cpp:64-74
    virtual uint32_t kind() const { return VB_LIGHT_SAMPLER_UNIFORM; }
};

class Sampler : public Node { public: using Desc = SamplerDesc; uint32_t spp; explicit Sampler(const SamplerDesc &d) : spp(d["spp"].as_uint(1u)) {} virtual uint32_t device_tag() const = 0; };
class Filter : public Node {
public:
    using Desc = FilterDesc;
    float radius[2];
    explicit Filter(const FilterDesc &d) { radius[0] = radius[1] = d["radius"].as_float(0.5f); }
    virtual uint32_t device_tag() const = 0;
    virtual float evaluate(float, float) const { return 0.f; }      // FittedCurveFilter subclasses (filter.h:60)
    // FilterSampler::build (fitted_curve.h:36-58): |f| at the 20x20 cell centres of the positive quadrant
    std::vector<float> table_weights() const {
        const int N = VB_FILTER_TABLE_SIZE;
        std::vector<float> w((size_t)N * N);
        for (int i = 0; i < N * N; ++i) {
            int x = i % N, y = i / N;
            float px = ((float)x + 0.5f) / (float)N * radius[0], py = ((float)y + 0.5f) / (float)N * radius[1];
            w[i] = std::fabs(evaluate(px, py));
        }
        return w;
    }
};
struct SceneRecords;
class Spectrum : public Node {
public:
    using Desc = SpectrumDesc;
    explicit Spectrum(const SpectrumDesc &) {}
    virtual uint32_t device_tag() const = 0;
    virtual bool is_complete() const { return false; }          // spectrum.h:265
    virtual void prepare(SceneRecords &) const {}                // SPDs / tables the device side needs
};
class ToneMapper : public Node { public: using Desc = ToneMapperDesc; explicit ToneMapper(const ToneMapperDesc &) {} virtual int device_tag() const = 0; };
class Warper : public Node {
public:
    using Desc = WarperDesc;
    explicit Warper(const WarperDesc &) {}
    virtual void build(const std::vector<float> &weights, std::vector<vb_alias_entry> &table, std::vector<float> &func, float &integral) const = 0;
};

class Sensor : public Node {       // reference src/base/sensor/sensor.{h,cpp}
public:
    using Desc = SensorDesc;
    float fov_y = 20.f, yaw = 0.f, pitch = 0.f;
    float3 position{0, 0, 0};
    float focal_distance = 5.f, lens_radius = 0.f;
    std::shared_ptr<Filter> filter;
    bool has_medium_name = false; std::string medium_name;   // photosensory.cpp:15-33
    explicit Sensor(const SensorDesc &desc);
    void update_mat(const Mat4 &m);                   // sensor.cpp:73-78
    Mat4 camera_to_world() const;                     // sensor.cpp:153-162
    void fill_record(vb_camera &cam, uint32_t w, uint32_t h) const;
};

class Integrator : public Node {
public:
    using Desc = IntegratorDesc;
    uint32_t max_depth, min_depth; float rr_threshold; uint32_t mis_mode;
    explicit Integrator(const IntegratorDesc &d)
        : max_depth(d["max_depth"].as_uint(16)), min_depth(d["min_depth"].as_uint(5)), rr_threshold(d["rr_threshold"].as_float(1.f)),
          mis_mode((uint32_t)d["mis_mode"].as_int(0)) {}
    uint32_t frame_index = 0;
    virtual void render(Pipeline &p, uint32_t n_frames) = 0;    // pt.cpp:96-116
};

class FrameBuffer : public Node {
public:
    using Desc = FrameBufferDesc;
    uint32_t width, height; float exposure; std::shared_ptr<ToneMapper> tone_mapper;
    explicit FrameBuffer(const FrameBufferDesc &d);
};

}  // namespace vh

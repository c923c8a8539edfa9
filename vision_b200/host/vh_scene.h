// Scene / Pipeline / ImagePool of the host layer.
//   Scene::init order, add_shape, fill_instances   reference src/base/mgr/scene.cpp:16-34,118-187
//   Pipeline::prepare (FixedRenderPipeline)        reference src/pipelines/fixed/pipeline.cpp:14-42
//   Geometry::update_instances/upload/build_accel  reference src/base/mgr/geometry.cpp:20-71
//   ImagePool::load_texture                        reference src/base/mgr/image_pool.cpp:13-35
#pragma once
#include "vh_node.h"

namespace vh {

struct Image { uint32_t w = 0, h = 0; std::vector<float> rgba; };
// Image::load of the reference (image_pool.cpp:13-35): .jpg / .png / .exr (vh_codecs.cpp), .hdr (vh_image.cpp), .pfm / .ppm;
// srgb_to_linear applies the sRGB EOTF to LDR sources ("color_space": "srgb"; never to .exr / .hdr)
Image load_image(const std::string &path, bool srgb_to_linear);
Image decode_image_file(const std::string &path);     // .jpg / .png / .exr as stored (8-bit values / 255), no colour-space conversion
void decode_jpeg(const unsigned char *data, size_t size, uint32_t &w, uint32_t &h, std::vector<unsigned char> &rgb);
void decode_png(const unsigned char *data, size_t size, uint32_t &w, uint32_t &h, std::vector<float> &rgba);
void decode_exr(const unsigned char *data, size_t size, uint32_t &w, uint32_t &h, std::vector<float> &rgba);
void save_exr(const std::string &path, const float *rgba, uint32_t w, uint32_t h, bool half);    // scan-line, ZIP, channels B G R
void save_pfm(const std::string &path, const float *rgba, uint32_t w, uint32_t h);
void save_ppm(const std::string &path, const float *rgba01, uint32_t w, uint32_t h);   // applies linear_to_srgb
// vh_image.cpp: Radiance .hdr read / write, PNG write, output dispatch by extension ("gamma unless exr / hdr", pipeline.cpp:337-354)
Image load_hdr(const std::string &path);
void save_hdr(const std::string &path, const float *rgba, uint32_t w, uint32_t h);
void save_png(const std::string &path, const float *rgba01, uint32_t w, uint32_t h, bool gamma);
void save_image(const std::string &path, const float *rgba, uint32_t w, uint32_t h, bool display_referred = false);   // true: already gamma-corrected
void rgbe_encode(const float *rgb, unsigned char *out);
void rgbe_decode(const unsigned char *in, float *rgb);

// vh_assimp.cpp: the reference's importer (vendored assimp) behind the `model` shape; vh_plugins3.inl: in-tree OBJ loader (file order)
bool assimp_available();
std::vector<MeshData> load_model_assimp(const std::string &path, bool swap_handed, bool smooth, bool flip_uv, uint32_t subdiv_level);
MeshData load_obj(const std::string &path, bool smooth, bool flip_uv);

class ImagePool {
public:
    std::vector<Image> images;
    std::vector<std::string> keys;
    uint32_t obtain(const std::string &path, const std::string &color_space);   // dedupes by (path, colour space)
};

// POD view of a prepared scene: exactly what is pushed through the C ABI (and what tests feed the oracle)
struct SceneRecords;
float light_power_luminance(const vb_light_rec &l, const SceneRecords &rec, float spot_angle, float spot_falloff);   // vh_plugins.cpp
struct SceneRecords {
    std::vector<std::shared_ptr<MeshData>> meshes;
    std::vector<vb_instance> instances;
    std::vector<vb_material_rec> materials;
    std::vector<vb_light_rec> lights;
    std::vector<vb_alias_entry> alias_table;
    std::vector<float> alias_func;
    struct LightSamplerRec { uint32_t kind = VB_LIGHT_SAMPLER_UNIFORM, alias_offset = 0, alias_size = 0; float alias_integral = 0.f;
                             uint32_t env_separate = 0; float env_prob = 0.f; } light_sampler;       // vb_light_sampler_set, vb_light_sampler_env_set
    std::vector<Image> textures;
    std::vector<vb_medium_rec> mediums;
    std::vector<std::vector<float>> spds;          // SPD pool (vb_spds_set): hero spectrum's CIE x, y, z, D65 first, then material curves
    vb_spectrum_rec spectrum{};                    // rgb2spec is set at upload time from the vector below
    std::vector<float> rgb2spec;
    struct Lut { std::vector<float> data; uint32_t nx = 0, ny = 0, nz = 0, channels = 0; } luts[VB_LUT_COUNT];
    vb_camera camera{};
    vb_pt_params params{};
    uint32_t width = 0, height = 0, spp = 1;
    float exposure = 1.f; int tonemap = 0;
};

class Scene {
public:
    SceneDesc desc;
    std::shared_ptr<LightSampler> light_sampler;
    std::shared_ptr<Spectrum> spectrum;
    std::vector<std::shared_ptr<Material>> materials;
    std::vector<std::shared_ptr<Medium>> mediums;      // MediumRegistry (registries.cpp:196-210)
    std::shared_ptr<Medium> global_medium;
    bool process_mediums = false;
    std::shared_ptr<Sensor> sensor;
    std::vector<std::shared_ptr<ShapeGroup>> groups;
    std::vector<std::shared_ptr<ShapeInstance>> instances;
    std::shared_ptr<Integrator> integrator;
    std::shared_ptr<Sampler> sampler;
    std::shared_ptr<Warper> warper;
    std::shared_ptr<FrameBuffer> frame_buffer;
    ImagePool image_pool;
    std::shared_ptr<Material> black_body;
    std::vector<std::string> log;

    void init(const SceneDesc &d);                    // scene.cpp:16-34
    void add_shape(const std::shared_ptr<ShapeGroup> &group, const ShapeDesc &sd);
    void prepare(SceneRecords &out);                  // scene.cpp:79-91 + geometry.cpp:20-34
    std::shared_ptr<Material> obtain_black_body();
    float world_diameter() const;                     // scene.h:108-109 over the instance boxes (shape.cpp:73-85)
};

class Pipeline : public Node {     // "vision-pipeline-fixed"
public:
    using Desc = PipelineDesc;
    Scene scene;
    SceneRecords records;
    vb_ctx *ctx = nullptr;
    vb_build_stats build_stats{};
    explicit Pipeline(const PipelineDesc &) {}
    std::string impl_type() const override { return "fixed"; }
    void init_scene(const SceneDesc &d);
    // multi-GPU: this pipeline is rank `dist_rank` of `dist_world`; it renders frame indices rank, rank + world, ...
    // of every render(n) call into a SUM film; final_picture() first makes the film whole (NCCL all-reduce over `nccl_comm`,
    // one process per GPU) - or render_on_gpus() below does it over peer memory (one process, several GPUs)
    uint32_t dist_rank = 0, dist_world = 1;
    std::vector<std::pair<std::string, double>> options;      // vb_ctx_set_option pairs applied before the BVH is built (e.g. build.two_level)
    void *nccl_comm = nullptr;
    void set_partition(uint32_t rank, uint32_t world);
    void nccl_init(const void *id128);                // after prepare(): vb_nccl_comm_create on this pipeline's context
    int accumulate_mode() const { return dist_world > 1 ? VB_ACCUM_SUM : VB_ACCUM_LERP; }
    virtual void prepare(int device);                 // uploads through the C ABI, builds the BVH
    virtual void render(uint32_t n_frames);           // integrator->render, then after_render()
    virtual void final_picture(std::vector<float> &rgba);     // [all-reduce +] resolve + download (pipeline.cpp:337-354)
    virtual bool output_is_display_referred() const { return false; }   // true: final_picture() is already gamma-corrected (customized graph)
    void after_render();                              // pipeline.cpp:307-311: Sensor::after_render -> store_prev_data (sensor.cpp:85-93)
    ~Pipeline() override;
    void check(int rc, const char *what) const;
};

std::shared_ptr<Pipeline> import_scene(const std::string &path);   // Importer::import_scene (importer.cpp:23-26)

// One process, n GPUs (devices first_device .. first_device + n - 1): one pipeline + host thread per GPU, frames partitioned by
// index, films reduced and resolved by vb_film_reduce_peers (use_nccl = false) or vb_film_allreduce (true). Returns rank 0's
// pipeline (its OUTPUT / ACCUM films hold the whole picture); `rgba` gets the final picture.
struct MultiGpuTimes { double prepare_ms = 0, render_ms = 0, reduce_ms = 0; };
std::shared_ptr<Pipeline> render_on_gpus(const std::string &scene_path, uint32_t n_gpus, uint32_t spp, bool use_nccl,
                                         std::vector<float> &rgba, MultiGpuTimes *times = nullptr, int first_device = 0);

}  // namespace vh

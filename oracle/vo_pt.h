// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
// CPU restatement of Vision's megakernel path tracer (camera ray -> Li -> film), following
//   reference src/base/integral/integrator.cpp:20-37,82-107,137-311      (Li, direct_lighting)
//   reference src/base/mgr/geometry.cpp:79-166,176-185,213-223           (interaction, occluded)
//   reference src/base/scattering/interaction.{h,cpp}                     (spawn_ray*, offsets)
//   reference src/base/illumination/lightsampler.cpp:159-300, render_core/lightsampler/uniform.cpp
//   reference src/render_core/light/area.cpp:33-177, src/base/illumination/light.h:126-130
//   reference src/render_core/warper/alias.h:86-181
//   reference src/base/scattering/{material.cpp:132-184,lobe.cpp:45-63,111-119,240-250,bxdf.cpp:13-61}
//   reference src/render_core/sampler/independent.cpp:13-43, src/math/util.h:13-34, src/base/sampler.h:60-73
//   reference src/base/sensor/{sensor.cpp:44-56,frame_buffer.cpp:117-183}, render_core/sensor/thin_lens.cpp:34-42
//   reference src/render_core/filter/{box,triangle}.cpp, src/math/warp.h
// It consumes the same POD records as the CUDA library (include/vision_b200_pt.h).
#pragma once
#include "vo_accel.h"
#include "../include/vision_b200.h"
#include <vector>

namespace vo {

struct MeshData {
    std::vector<float3> pos, nrm;
    std::vector<float2> uv;
    std::vector<uint32_t> tri;     // 3 per triangle
};

struct Texture { std::vector<float4> texels; uint32_t w = 0, h = 0; };
struct Lut { std::vector<float> data; uint32_t nx = 0, ny = 0, nz = 0, channels = 0; };

struct InstanceX {                 // instance + derived normal matrix
    vb_instance rec;
    float4x4 o2w;
    float nmat[9];                 // sign(det) * cofactor(o2w 3x3), row-major: n_world = nmat * n
};

struct PathStats { uint64_t closest = 0, any = 0, vertices = 0, nodes = 0, tris = 0; };

struct Scene {
    std::vector<MeshData> meshes;
    std::vector<InstanceX> instances;
    std::vector<vb_material_rec> materials;
    std::vector<vb_light_rec> lights;
    std::vector<vb_alias_entry> alias_table;
    std::vector<float> alias_func;
    uint32_t lsel_offset = 0, lsel_size = 0; float lsel_integral = 0.f;      // "power" light sampler table (size 0 = uniform)
    uint32_t env_light = VB_INVALID_ID;      // LightSampler::env_light_ (lightsampler.cpp:18-25): the last Infinite light wins
    uint32_t env_separate = 0; float env_prob = 0.f;                         // lightsampler.cpp:11-13,159-197
    std::vector<Texture> textures;
    Lut luts[VB_LUT_COUNT];
    std::vector<vb_medium_rec> mediums;
    std::vector<float> spd_values;          // SPD pool (vb_spds_set)
    std::vector<uint32_t> spd_offsets;
    vb_spectrum_rec spectrum{};             // rgb2spec points into the copy below
    std::vector<float> rgb2spec;
    uint32_t rgb2spec_res = 0;
    vb_camera camera{};
    vb_camera_prev camera_prev{}; bool have_camera_prev = false;   // Sensor::store_prev_data (sensor.cpp:89-93); unset = static camera
    vb_pt_params params{};
    Accel accel;
    bool brute = false;

    void set_instances(const vb_instance *inst, uint32_t n);
    void build(int threads);
    Hit trace_closest(const Ray &r, PathStats *st) const;
    bool trace_any(const Ray &r, PathStats *st) const;

    // one pixel-sample: returns L (linear sRGB), optionally the primary hit
    // aov (nullable, 20 floats): rt_geom's G-buffer planes for this pixel-sample (frame_buffer.cpp:156-219) - normal.xyz 1 | albedo.rgb 1 |
    // emission.rgb 1 | linear depth, p_film.x, p_film.y, hit flag | motion vector.xy 0 0; when given, the function returns after the camera stage
    // dump (nullable): one vb_path_vertex per closest-hit ray of Li with the path state before the vertex, then the VB_PATH_END record
    float3 render_sample(uint32_t px, uint32_t py, uint32_t w, uint32_t h, uint32_t frame, Hit *primary, PathStats *st, float *aov = nullptr,
                         std::vector<vb_path_vertex> *dump = nullptr) const;
};

// ---- shared small functions (unit-tested against known answers) ----
uint32_t tea4(uint32_t v0, uint32_t v1);
inline float lcg(uint32_t &state) {
    state = 1664525u * state + 1013904223u;
    return (float)(state & 0x00ffffffu) * (1.0f / 16777216.0f);
}
void sincos_2pi(float u, float &s, float &c);
float vexp(float x);
float vlog(float x);
float vatan2(float y, float x);
float3 offset_ray_origin(float3 p, float3 n);

}  // namespace vo

// Included by vh_plugins.cpp: sampler, filters, spectrum, tone mappers, sensor, integrator, frame buffer.
namespace vh {

// reference src/render_core/sampler/independent.cpp:13-43
class IndependentSampler : public Sampler {
public:
    explicit IndependentSampler(const SamplerDesc &d) : Sampler(d) {}
    std::string impl_type() const override { return "independent"; }
    uint32_t device_tag() const override { return 0; }
};
VH_MAKE_CLASS_CREATOR(IndependentSampler, sampler, independent, constructor)

// reference src/render_core/filter/box.cpp:16-20, triangle.cpp:16-18
class BoxFilter : public Filter {
public:
    explicit BoxFilter(const FilterDesc &d) : Filter(d) {}
    std::string impl_type() const override { return "box"; }
    uint32_t device_tag() const override { return VB_FILTER_BOX; }
};
VH_MAKE_CLASS_CREATOR(BoxFilter, filter, box, constructor)
class TriangleFilter : public Filter {
public:
    explicit TriangleFilter(const FilterDesc &d) : Filter(d) {}
    std::string impl_type() const override { return "triangle"; }
    uint32_t device_tag() const override { return VB_FILTER_TRIANGLE; }
};
VH_MAKE_CLASS_CREATOR(TriangleFilter, filter, triangle, constructor)

// reference src/render_core/filter/gaussian.cpp:24-53 + math/util.h:37-42
class GaussianFilter : public Filter {
    float sigma_, exp_;
    float gaussian(float x) const {
        float a = 1.0f / std::sqrt((2.0f * 3.14159265358979323846f) * (sigma_ * sigma_));
        return a * std::exp(-(x * x) / (2.0f * (sigma_ * sigma_)));
    }
public:
    explicit GaussianFilter(const FilterDesc &d) : Filter(d), sigma_(d["sigma"].as_float(1.f)) { exp_ = gaussian(radius[0]); }
    std::string impl_type() const override { return "gaussian"; }
    uint32_t device_tag() const override { return VB_FILTER_FITTED; }
    float evaluate(float x, float y) const override { return std::max(0.f, gaussian(x) - exp_) * std::max(0.f, gaussian(y) - exp_); }
};
VH_MAKE_CLASS_CREATOR(GaussianFilter, filter, gaussian, constructor)

// reference src/render_core/filter/mitchell.cpp:19-52
class MitchellFilter : public Filter {
    float b_, c_;
    float mitchell_1d(float x) const {
        x = std::fabs(x);
        float x2 = x * x, x3 = x2 * x, s6 = 1.0f / 6.0f;
        if (x <= 1) return ((((12.f - 9.f * b_) - 6.f * c_) * x3 + ((-18.f + 12.f * b_) + 6.f * c_) * x2) + (6.f - 2.f * b_)) * s6;
        if (x <= 2) return (((((-b_ - 6.f * c_) * x3 + (6.f * b_ + 30.f * c_) * x2)) + (-12.f * b_ - 48.f * c_) * x) + (8.f * b_ + 24.f * c_)) * s6;
        return 0.f;
    }
public:
    explicit MitchellFilter(const FilterDesc &d) : Filter(d), b_(d["b"].as_float(1.f / 3.f)), c_(d["c"].as_float(1.f / 3.f)) {}
    std::string impl_type() const override { return "mitchell"; }
    uint32_t device_tag() const override { return VB_FILTER_FITTED; }
    float evaluate(float x, float y) const override { return mitchell_1d(2.f * x / radius[0]) * mitchell_1d(2.f * y / radius[1]); }
};
VH_MAKE_CLASS_CREATOR(MitchellFilter, filter, mitchell, constructor)

// reference src/render_core/filter/sinc.cpp:17-30 + math/util.h:51-66
class LanczosSincFilter : public Filter {
    float tau_;
    static float sinc(float x) { x = 3.14159265358979323846f * x; return 1.0f + x * x == 1.0f ? 1.0f : std::sin(x) / x; }
    float windowed(float x, float r) const { return std::fabs(x) > r ? 0.f : sinc(x) * sinc(x / tau_); }
public:
    explicit LanczosSincFilter(const FilterDesc &d) : Filter(d), tau_(d["tau"].as_float(3.f)) {}
    std::string impl_type() const override { return "sinc"; }
    uint32_t device_tag() const override { return VB_FILTER_FITTED; }
    float evaluate(float x, float y) const override { return (windowed(x, radius[0]) * windowed(y, radius[1])) * 4.0f; }
};
VH_MAKE_CLASS_CREATOR(LanczosSincFilter, filter, sinc, constructor)

// reference src/render_core/spectrum/srgb.cpp:16-57
class SRGBSpectrum : public Spectrum {
public:
    explicit SRGBSpectrum(const SpectrumDesc &d) : Spectrum(d) {}
    std::string impl_type() const override { return "srgb"; }
    uint32_t device_tag() const override { return 0; }
};
VH_MAKE_CLASS_CREATOR(SRGBSpectrum, spectrum, srgb, constructor)

// reference src/render_core/spectrum/hero.cpp:227-348 (dimension 3 = its default). The RGB -> spectrum coefficient table the
// reference compiles in (srgb2spec.h, not in its tree) is read from data/rgb2spec_srgb_3x64x64x64x4.f32 (tools/rgb2spec_opt.c);
// the CIE observer / D65 SPDs are every 5th sample of the 1 nm tables (SPD::create_cie_*, spd.cpp:24-30,89-103)
class HeroWavelengthSpectrum : public Spectrum {
    uint32_t dimension_;
public:
    explicit HeroWavelengthSpectrum(const SpectrumDesc &d) : Spectrum(d), dimension_(d["dimension"].as_uint(3u)) {
        if (dimension_ != 3) throw std::runtime_error("hero spectrum: only dimension 3 is implemented");
    }
    std::string impl_type() const override { return "hero"; }
    uint32_t device_tag() const override { return VB_SPECTRUM_HERO; }
    bool is_complete() const override { return true; }
    void prepare(SceneRecords &out) const override {
        std::vector<float> cie(4 * 471);
        FILE *f = std::fopen((data_dir() + "/cie_1nm_4x471.f32").c_str(), "rb");
        if (!f || std::fread(cie.data(), 4, cie.size(), f) != cie.size()) throw std::runtime_error("cannot read data/cie_1nm_4x471.f32 (tools/extract_cie.py writes it)");
        std::fclose(f);
        const uint32_t interval = 5, n = (uint32_t)std::ceil((830.f - 360.f) / (float)interval);
        out.spds.assign(4, std::vector<float>(n));
        for (int k = 0; k < 4; ++k) for (uint32_t x = 0; x < n; ++x) out.spds[k][x] = cie[k * 471 + x * interval];
        float sum = 0.f;                                            // densely_sampled_spectrum_integral, spd.cpp:11-21
        uint32_t nn = (uint32_t)((830.f - 360.f) / (float)interval) + 1u;
        for (uint32_t i = 0; i + 1 < nn; ++i) sum += 0.5f * (cie[471 + i * interval] + cie[471 + (i + 1) * interval]);
        out.spectrum = vb_spectrum_rec{};
        out.spectrum.type = VB_SPECTRUM_HERO; out.spectrum.dimension = dimension_;
        out.spectrum.spd_x = 0; out.spectrum.spd_y = 1; out.spectrum.spd_z = 2; out.spectrum.spd_d65 = 3;
        out.spectrum.cie_y_integral = sum * (float)interval;
        out.spectrum.rgb2spec_res = 64;
        out.rgb2spec.resize((size_t)3 * 64 * 64 * 64 * 4);
        f = std::fopen((data_dir() + "/rgb2spec_srgb_3x64x64x64x4.f32").c_str(), "rb");
        if (!f || std::fread(out.rgb2spec.data(), 4, out.rgb2spec.size(), f) != out.rgb2spec.size())
            throw std::runtime_error("cannot read data/rgb2spec_srgb_3x64x64x64x4.f32 (written by __graft_entry__.build() / tools/rgb2spec_opt.c)");
        std::fclose(f);
    }
};
VH_MAKE_CLASS_CREATOR(HeroWavelengthSpectrum, spectrum, hero, constructor)

// reference src/render_core/tonemapper/impl.cpp:9-52: one module "vision-tonemapper-impl", three constructors
class LinearToneMapper : public ToneMapper { public: explicit LinearToneMapper(const ToneMapperDesc &d) : ToneMapper(d) {} std::string impl_type() const override { return "linear"; } int device_tag() const override { return VB_TONEMAP_LINEAR; } };
class ACESToneMapper : public ToneMapper { public: explicit ACESToneMapper(const ToneMapperDesc &d) : ToneMapper(d) {} std::string impl_type() const override { return "aces"; } int device_tag() const override { return VB_TONEMAP_ACES; } };
class ReinhardToneMapper : public ToneMapper { public: explicit ReinhardToneMapper(const ToneMapperDesc &d) : ToneMapper(d) {} std::string impl_type() const override { return "reinhard"; } int device_tag() const override { return VB_TONEMAP_REINHARD; } };
VH_MAKE_CLASS_CREATOR(LinearToneMapper, tonemapper, impl, linear)
VH_MAKE_CLASS_CREATOR(ACESToneMapper, tonemapper, impl, aces)
VH_MAKE_CLASS_CREATOR(ReinhardToneMapper, tonemapper, impl, reinhard)

// ---- sensor: reference src/base/sensor/sensor.cpp:13-25,58-78,153-162; thin_lens.cpp:15-42 ----
Sensor::Sensor(const SensorDesc &desc) {
    name_ = desc.name;
    fov_y = std::min(120.f, std::max(15.f, desc["fov_y"].as_float(20.f)));
    focal_distance = desc["focal_distance"].as_float(5.f);
    lens_radius = desc["lens_radius"].as_float(0.f);
    has_medium_name = desc.has_attr("medium"); medium_name = desc["medium"].as_string();
    update_mat(desc.transform_desc.mat);
    filter = create_shared<Filter>(desc.filter_desc);
}
void Sensor::update_mat(const Mat4 &m) {
    const double R2D = 180.0 / 3.14159265358979323846;
    pitch = (float)(std::atan2((double)m[4 * 1 + 2], (double)m[4 * 1 + 1]) * R2D);   // m[1][2], m[1][1]
    yaw = (float)(std::atan2((double)m[4 * 2 + 0], (double)m[4 * 0 + 0]) * R2D);     // m[2][0], m[0][0]
    position = {m[12], m[13], m[14]};
}
Mat4 Sensor::camera_to_world() const {
    Mat4 rot = mat4_mul(mat4_mul(scaling({1, 1, -1}), rotation_y(yaw)), rotation_x(-pitch));
    return mat4_mul(translation(position), rot);
}
void Sensor::fill_record(vb_camera &cam, uint32_t w, uint32_t h) const {
    std::memset(&cam, 0, sizeof(cam));
    Mat4 c2w = camera_to_world();
    std::memcpy(cam.c2w, c2w.data(), 64);
    cam.tan_fov_y_over_2 = (float)std::tan((double)fov_y * 3.14159265358979323846 / 180.0 * 0.5);
    cam.lens_radius = lens_radius; cam.focal_distance = focal_distance; cam.medium_id = VB_INVALID_ID;
    float ratio = (float)w * 1.f / (float)h;                        // frame_buffer.cpp:93-102
    cam.screen_lo[0] = cam.screen_lo[1] = -1.f; cam.screen_hi[0] = cam.screen_hi[1] = 1.f;
    if (ratio > 1.f) { cam.screen_lo[0] = -ratio; cam.screen_hi[0] = ratio; }
    else { cam.screen_lo[1] = -1.f / ratio; cam.screen_hi[1] = 1.f / ratio; }
    cam.filter_type = filter->device_tag();
    cam.filter_radius[0] = filter->radius[0]; cam.filter_radius[1] = filter->radius[1];
}
class ThinLensCamera : public Sensor {
public:
    explicit ThinLensCamera(const SensorDesc &d) : Sensor(d) {}
    std::string impl_type() const override { return "thin_lens"; }
};
VH_MAKE_CLASS_CREATOR(ThinLensCamera, sensor, thin_lens, constructor)
class PinholeCamera : public Sensor {
public:
    explicit PinholeCamera(const SensorDesc &d) : Sensor(d) { lens_radius = 0.f; }
    std::string impl_type() const override { return "pinhole"; }
};
VH_MAKE_CLASS_CREATOR(PinholeCamera, sensor, pinhole, constructor)

// ---- frame buffer: reference src/base/sensor/frame_buffer.cpp:16-26 ----
FrameBuffer::FrameBuffer(const FrameBufferDesc &d) {
    std::vector<float> res = d["resolution"].as_floats();
    width = res.size() >= 2 ? (uint32_t)res[0] : 1280u;
    height = res.size() >= 2 ? (uint32_t)res[1] : 720u;
    exposure = d["exposure"].as_float(1.f);
    tone_mapper = create_shared<ToneMapper>(d.tone_mapper);
}
class NormalFrameBuffer : public FrameBuffer {
public:
    explicit NormalFrameBuffer(const FrameBufferDesc &d) : FrameBuffer(d) {}
    std::string impl_type() const override { return "normal"; }
};
VH_MAKE_CLASS_CREATOR(NormalFrameBuffer, framebuffer, normal, constructor)

// ---- integrator: reference src/render_core/integrator/pt.cpp:27-37,96-116 ----
class PathTracingIntegrator : public Integrator {
public:
    explicit PathTracingIntegrator(const IntegratorDesc &d) : Integrator(d) {}
    std::string impl_type() const override { return "pt"; }
    void render(Pipeline &p, uint32_t n_frames) override {
        // reference: per frame GBuffer -> path_tracing -> accumulate -> tone map, frame_index_++ ;
        // here one launch renders the whole batch of frame indices (film kept in registers)
        // multi-GPU: of the n_frames indices this call covers, this rank takes frame_index + rank, + world, ... into a SUM film
        uint32_t mine = n_frames > p.dist_rank ? (n_frames - p.dist_rank + p.dist_world - 1) / p.dist_world : 0;
        if (mine) p.check(vb_render_frames(p.ctx, frame_index + p.dist_rank, mine, p.dist_world, p.accumulate_mode()), "vb_render_frames");
        frame_index += n_frames;
    }
};
VH_MAKE_CLASS_CREATOR(PathTracingIntegrator, integrator, pt, constructor)

VH_MAKE_CLASS_CREATOR(Pipeline, pipeline, fixed, constructor)

}  // namespace vh

// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
// See vo_pt.h for the reference files restated here. Operation order in this file is normative
// for the CUDA megakernel (vision_b200/csrc/vb_pt.cu), which mirrors it under -fmad=false.
#include "vo_pt.h"
#include <cassert>
#include <limits>

namespace vo {

// ---- reference src/math/util.h:13-24 (tea_impl<N=4>) ----
uint32_t tea4(uint32_t v0, uint32_t v1) {
    uint32_t s0 = 0;
    for (int n = 0; n < 4; ++n) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}

// sin(2*pi*u), cos(2*pi*u) for u in [0,1]: exact quadrant reduction in u, cephes-style minimax
// polynomials on [-pi/4, pi/4], plain mul/add only -> bit-identical on host and device.
void sincos_2pi(float u, float &s, float &c) {
    float q = std::floor(u * 4.0f + 0.5f);
    float r = u - q * 0.25f;
    float x = r * 6.28318530717958647692f;
    float z = x * x;
    float sp = ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z - 1.6666654611e-1f) * z * x + x;
    float cp = ((2.443315711809948e-5f * z - 1.388731625493765e-3f) * z + 4.166664568298827e-2f) * z * z - 0.5f * z + 1.0f;
    int iq = (int)q & 3;
    if (iq == 0) { s = sp; c = cp; }
    else if (iq == 1) { s = cp; c = -sp; }
    else if (iq == 2) { s = -sp; c = -cp; }
    else { s = -cp; c = sp; }
}

// exp / log with a fixed operation sequence (cephes expf/logf polynomials, plain mul/add + the
// fmaf written below) so that medium transmittance and free-flight sampling take bit-identical
// values on host and device; libm and CUDA's expf/logf differ in the last ulp.
float vexp(float x) {
    if (!(x > -87.0f)) return x != x ? x : 0.0f;
    if (x > 88.0f) return std::numeric_limits<float>::infinity();
    float n = std::floor(x * 1.44269504088896341f + 0.5f);
    float r = std::fmaf(n, -0.693359375f, x);
    r = std::fmaf(n, 2.12194440e-4f, r);
    float z = r * r;
    float p = ((((1.9875691500e-4f * r + 1.3981999507e-3f) * r + 8.3334519073e-3f) * r + 4.1665795894e-2f) * r + 1.6666665459e-1f) * r + 5.0000001201e-1f;
    float y = p * z + r + 1.0f;
    return u2f(f2u(y) + ((uint32_t)(int32_t)n << 23));       // y in [0.7, 1.42): scale by 2^n (|n| <= 127)
}
float vlog(float x) {
    if (!(x > 0.0f)) return x == 0.0f ? -std::numeric_limits<float>::infinity() : std::numeric_limits<float>::quiet_NaN();
    uint32_t b = f2u(x);
    int32_t e = (int32_t)((b >> 23) & 0xffu) - 126;
    float m = u2f((b & 0x007fffffu) | 0x3f000000u);            // [0.5, 1)
    if (m < 0.70710678118654752440f) { e -= 1; m = m + m - 1.0f; } else { m = m - 1.0f; }
    float z = m * m;
    float y = ((((((((7.0376836292e-2f * m - 1.1514610310e-1f) * m + 1.1676998740e-1f) * m - 1.2420140846e-1f) * m + 1.4249322787e-1f) * m
                 - 1.6668057665e-1f) * m + 2.0000714765e-1f) * m - 2.4999993993e-1f) * m + 3.3333331174e-1f) * m * z;
    float fe = (float)e;
    y = y + -2.12194440e-4f * fe;
    y = y + -0.5f * z;
    float r = m + y;
    return r + 0.693359375f * fe;
}

// atan2 with a fixed operation sequence (cephes atanf reduction + polynomial); acos(x) is taken as
// vatan2(sqrt(1 - x*x), x). Used for the lat-long environment map lookup (spherical_theta / spherical_phi).
float vatan2(float y, float x) {
    float ax = std::fabs(x), ay = std::fabs(y);
    float r;
    if (ax == 0.f && ay == 0.f) r = 0.f;
    else {
        bool swap = ay > ax;
        float t = swap ? ax / ay : ay / ax;                     // in [0, 1]
        float off = 0.f;
        if (t > 0.4142135623730950f) { off = 0.78539816339744830961f; t = (t - 1.0f) / (t + 1.0f); }
        float z = t * t;
        float p = (((8.05374449538e-2f * z - 1.38776856032e-1f) * z + 1.99777106478e-1f) * z - 3.33329491539e-1f) * z * t + t;
        r = off + p;
        if (swap) r = 1.57079632679489661923f - r;
    }
    if (x < 0.f) r = 3.14159265358979323846f - r;
    return y < 0.f ? -r : r;
}

// Waechter & Binder (Ray Tracing Gems ch.6) integer-ULP offset; ocarina absent.
float3 offset_ray_origin(float3 p, float3 n) {
    const float origin = 1.0f / 32.0f, float_scale = 1.0f / 65536.0f, int_scale = 256.0f;
    auto one = [&](float pc, float nc) {
        int32_t of_i = (int32_t)(int_scale * nc);
        int32_t pi = (int32_t)f2u(pc);
        float p_i = u2f((uint32_t)(pi + (pc < 0.f ? -of_i : of_i)));
        return std::fabs(pc) < origin ? pc + float_scale * nc : p_i;
    };
    return {one(p.x, n.x), one(p.y, n.y), one(p.z, n.z)};
}

namespace {

struct Sampler {     // IndependentSampler, reference independent.cpp:24-42
    uint32_t state = 0;
    void start(uint32_t px, uint32_t py, uint32_t sample_index, uint32_t dim) {
        state = tea4(tea4(px, py), tea4(sample_index, dim));
    }
    float next_1d() { return lcg(state); }
    float2 next_2d() { float x = next_1d(); float y = next_1d(); return {x, y}; }
};

inline float3 lerp3(float u, float v, float3 a, float3 b, float3 c) {   // TriangleHit::triangle_lerp
    float w = 1.0f - u - v;
    return a * w + b * u + c * v;
}
inline float2 lerp2(float u, float v, float2 a, float2 b, float2 c) {
    float w = 1.0f - u - v;
    return {a.x * w + b.x * u + c.x * v, a.y * w + b.y * u + c.y * v};
}
inline bool is_zero(float3 v) { return v.x == 0.f && v.y == 0.f && v.z == 0.f; }

struct Frame {
    float3 x, y, z;
    float3 to_local(float3 v) const { return {dot(v, x), dot(v, y), dot(v, z)}; }
    float3 to_world(float3 v) const { return x * v.x + y * v.y + z * v.z; }
    void update(float3 n) {      // PartialDerivative::update(norm), interaction.h:101-105
        z = n;
        x = normalize(cross(n, y)) * length(x);
        y = normalize(cross(n, x)) * length(y);
    }
};

struct Interaction {
    float3 pos, wo, ng;
    float2 uv;
    Frame shading;
    float prim_area = 0.f;
    uint32_t prim_id = VB_INVALID_ID, mat_id = VB_INVALID_ID, light_id = VB_INVALID_ID;
    uint32_t inside = VB_INVALID_ID, outside = VB_INVALID_ID;
    bool has_phase = false;        // medium interaction (Interaction::init_phase, interaction.cpp:58-63)
    float g = 0.f;
};

// bilinear, wrap addressing, normalised uv (SURVEY A1 texture row)
float4 sample_texture(const Texture &t, float2 uv) {
    float x = uv.x * (float)t.w - 0.5f, y = uv.y * (float)t.h - 0.5f;
    float fx0 = std::floor(x), fy0 = std::floor(y);
    float fx = x - fx0, fy = y - fy0;
    auto wrap = [](int i, int n) { int m = i % n; return m < 0 ? m + n : m; };
    int x0 = wrap((int)fx0, (int)t.w), x1 = wrap((int)fx0 + 1, (int)t.w);
    int y0 = wrap((int)fy0, (int)t.h), y1 = wrap((int)fy0 + 1, (int)t.h);
    auto at = [&](int xx, int yy) { return t.texels[(size_t)yy * t.w + xx]; };
    float4 a = at(x0, y0), b = at(x1, y0), c = at(x0, y1), d = at(x1, y1);
    auto mix = [](float4 p, float4 q, float s) {
        return float4{p.x + s * (q.x - p.x), p.y + s * (q.y - p.y), p.z + s * (q.z - p.z), p.w + s * (q.w - p.w)};
    };
    return mix(mix(a, b, fx), mix(c, d, fx), fy);
}

float3 eval_slot3(const Scene &sc, const vb_slot &s, float2 uv) {
    float3 v{s.value[0], s.value[1], s.value[2]};
    if (s.tex_id != VB_INVALID_ID) {
        float4 t = sample_texture(sc.textures[s.tex_id], uv);
        v = {v.x * t.x, v.y * t.y, v.z * t.z};
    }
    return v;
}

}  // namespace

void Scene::set_instances(const vb_instance *inst, uint32_t n) {
    instances.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        InstanceX &x = instances[i];
        x.rec = inst[i];
        std::memcpy(&x.o2w, inst[i].o2w, 64);
        // normal matrix: sign(det) * cofactor of the upper 3x3, in double, rounded to fp32
        double m[3][3];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) m[r][c] = x.o2w.m[c][r];
        double cf[3][3];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) {
                int r1 = (r + 1) % 3, r2 = (r + 2) % 3, c1 = (c + 1) % 3, c2 = (c + 2) % 3;
                cf[r][c] = m[r1][c1] * m[r2][c2] - m[r1][c2] * m[r2][c1];
            }
        double det = m[0][0] * cf[0][0] + m[0][1] * cf[0][1] + m[0][2] * cf[0][2];
        double sg = det < 0 ? -1.0 : 1.0;
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) x.nmat[3 * r + c] = (float)(sg * cf[r][c]);
    }
}

void Scene::build(int threads) {
    accel.tris.clear();
    for (uint32_t i = 0; i < instances.size(); ++i) {
        const MeshData &m = meshes[instances[i].rec.mesh_id];
        size_t base = accel.tris.size();
        size_t nt = m.tri.size() / 3;
        accel.tris.resize(base + nt);
        for (size_t t = 0; t < nt; ++t) {
            float3 v0 = transform_point(instances[i].o2w, m.pos[m.tri[3 * t]]);
            float3 v1 = transform_point(instances[i].o2w, m.pos[m.tri[3 * t + 1]]);
            float3 v2 = transform_point(instances[i].o2w, m.pos[m.tri[3 * t + 2]]);
            accel.tris[base + t] = WTri{v0, v1 - v0, v2 - v0, i, (uint32_t)t};
        }
    }
    accel.build(threads);
}

Hit Scene::trace_closest(const Ray &r, PathStats *st) const {
    if (st) st->closest++;
    if (brute) return accel.closest_brute(r);
    TraceStats ts;
    Hit h = accel.closest(r, st ? &ts : nullptr);
    if (st) { st->nodes += ts.nodes; st->tris += ts.tris; }
    return h;
}
bool Scene::trace_any(const Ray &r, PathStats *st) const {
    if (st) st->any++;
    if (brute) return accel.any_brute(r);
    TraceStats ts;
    bool o = accel.any(r, st ? &ts : nullptr);
    if (st) { st->nodes += ts.nodes; st->tris += ts.tris; }
    return o;
}

}  // namespace vo

#include "vo_pt_li.inl"
#include "vo_probe.inl"

// known-answer probes of the hero spectrum (tests/test_oracle_spectrum.py)
extern "C" void vo_hero_wavelengths(uint32_t u24, float *out) {
    vo::Scene sc;
    sc.params.spectrum = 1;
    vo::Swl w = vo::swl_make(sc, u24);
    for (int i = 0; i < 3; ++i) { out[i] = w.lambda[i]; out[3 + i] = w.pdf[i]; }
}
extern "C" void vo_hero_decode(void *h, uint32_t u24, int kind, const float *rgb, float *out_sp, float *out_rgb) {
    const vo::Scene &sc = *(const vo::Scene *)h;
    vo::Swl w = vo::swl_make(sc, u24);
    vo::float3 c{rgb[0], rgb[1], rgb[2]};
    vo::float3 sp = kind == 0 ? vo::spec_albedo(sc, w, c) : (kind == 1 ? vo::spec_unbound(sc, w, c) : vo::spec_illum(sc, w, c));
    vo::float3 o = vo::spec_to_srgb(sc, w, sp);
    out_sp[0] = sp.x; out_sp[1] = sp.y; out_sp[2] = sp.z;
    out_rgb[0] = o.x; out_rgb[1] = o.y; out_rgb[2] = o.z;
}

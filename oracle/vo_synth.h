// ORACLE — test infrastructure only (see vo_math.h header).
// Synthetic inputs of SURVEY.md §8(d): random triangle soup M(N) and incoherent rays.
// Counter-based RNG: rnd(seed,k) = splitmix64 finaliser of (seed*K + (k+1)*GOLDEN); fp32 in [0,1)
// as (x >> 40) * 2^-24. vision_b200/synth.py is the numpy twin used by bench.py; the two are
// compared bit-for-bit (mesh) in tests/test_synth.py.
#pragma once
#include "vo_math.h"

namespace vo {

inline uint64_t splitmix64(uint64_t seed, uint64_t k) {
    uint64_t z = seed * 0xD1342543DE82EF95ull + (k + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline float rnd01(uint64_t seed, uint64_t k) {
    return (float)(splitmix64(seed, k) >> 40) * 0x1p-24f;
}

// triangle i of M(N): 12 draws at k = 12*i .. 12*i+11 : centre(3), xi0(3), xi1(3), xi2(3)
// pos9 = v0.xyz v1.xyz v2.xyz  (non-indexed soup: triangle i = vertices 3i,3i+1,3i+2)
inline void synth_triangle(uint64_t seed, uint64_t n_total, uint64_t i, float *pos9) {
    float r = (float)(2.0 * std::pow(1.0 / (double)n_total, 1.0 / 3.0));
    uint64_t idx = i;
    for (;;) {
        uint64_t k = 12 * idx;
        float c[3];
        for (int a = 0; a < 3; ++a) c[a] = 2.0f * rnd01(seed, k + a) - 1.0f;
        for (int v = 0; v < 3; ++v)
            for (int a = 0; a < 3; ++a) {
                float xi = rnd01(seed, k + 3 + 3 * v + a);
                float off = r * (xi - 0.5f);
                pos9[3 * v + a] = c[a] + off;
            }
        float3 v0{pos9[0], pos9[1], pos9[2]}, v1{pos9[3], pos9[4], pos9[5]}, v2{pos9[6], pos9[7], pos9[8]};
        float3 e1 = v1 - v0, e2 = v2 - v0;
        // plain (non-FMA) cross so numpy reproduces it exactly
        float cx = e1.y * e2.z - e1.z * e2.y, cy = e1.z * e2.x - e1.x * e2.z, cz = e1.x * e2.y - e1.y * e2.x;
        float a2 = cx * cx + cy * cy + cz * cz;     // (2*area)^2
        if (a2 >= 4e-24f) return;                   // area >= 1e-12
        idx += n_total;                             // degenerate: redraw with index + N
    }
}

// ray i: 5 draws at k = 8*i .. : origin(3), dir u0,u1.  any-hit t_max: rnd01(seed_any, i)
inline void synth_ray(uint64_t seed, uint64_t i, float *o3, float *d3) {
    uint64_t k = 8 * i;
    for (int a = 0; a < 3; ++a) o3[a] = 2.0f * rnd01(seed, k + a) - 1.0f;
    float u0 = rnd01(seed, k + 3), u1 = rnd01(seed, k + 4);
    // square_to_sphere, reference src/math/warp.h:72-78; trig in double then rounded
    float z = 1.0f - 2.0f * u0;
    float rr = std::sqrt(std::max(0.0f, 1.0f - z * z));
    double phi = 2.0 * 3.14159265358979323846 * (double)u1;
    d3[0] = (float)((double)rr * std::cos(phi));
    d3[1] = (float)((double)rr * std::sin(phi));
    d3[2] = z;
}
inline float synth_any_tmax(uint64_t seed_any, uint64_t i) {
    return 0.05f + 0.45f * rnd01(seed_any, i);
}

}  // namespace vo

// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
// extern "C" surface of the oracle for ctypes (tests/, smoke(), bench.py cpu_baseline only).
#include "vo_accel.h"
#include "vo_synth.h"
#include <thread>
#include <vector>
#include <chrono>

using namespace vo;

namespace {
template <class F>
void parallel_for(uint64_t n, int threads, F f) {
    if (threads <= 1 || n < 1024) { f(0, n, 0); return; }
    std::vector<std::thread> th;
    uint64_t chunk = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        uint64_t b = t * chunk, e = std::min(n, b + chunk);
        if (b >= e) break;
        th.emplace_back([=]() { f(b, e, t); });
    }
    for (auto &x : th) x.join();
}
}  // namespace

extern "C" {

int vo_hw_threads() { return (int)std::max(1u, std::thread::hardware_concurrency()); }

void vo_synth_mesh(uint64_t seed, uint64_t n, float *pos9n, int threads) {
    parallel_for(n, threads, [=](uint64_t b, uint64_t e, int) {
        for (uint64_t i = b; i < e; ++i) synth_triangle(seed, n, i, pos9n + 9 * i);
    });
}

// SoA rays: ox,oy,oz,tmin,dx,dy,dz,tmax each n floats. tmax = 1e30 (closest set);
// tmax_any (nullable) receives the any-hit set's t_max.
void vo_synth_rays(uint64_t seed, uint64_t seed_any, uint64_t n, float *ox, float *oy, float *oz, float *tmin,
                   float *dx, float *dy, float *dz, float *tmax, float *tmax_any, int threads) {
    parallel_for(n, threads, [=](uint64_t b, uint64_t e, int) {
        for (uint64_t i = b; i < e; ++i) {
            float o[3], d[3];
            synth_ray(seed, i, o, d);
            ox[i] = o[0]; oy[i] = o[1]; oz[i] = o[2]; tmin[i] = 0.f;
            dx[i] = d[0]; dy[i] = d[1]; dz[i] = d[2]; tmax[i] = RayTMax;
            if (tmax_any) tmax_any[i] = synth_any_tmax(seed_any, i);
        }
    });
}

void *vo_accel_create() { return new Accel(); }
void vo_accel_destroy(void *h) { delete (Accel *)h; }

// append one instance: world vertex = transform_point(o2w, pos); triangle order defines prim_id.
// o2w: 16 floats column-major (nullptr = identity, vertices used as-is).
void vo_accel_add_instance(void *h, const float *pos3, uint32_t nv, const uint32_t *tri3, uint32_t nt,
                           const float *o2w16, uint32_t inst_id) {
    Accel &a = *(Accel *)h;
    (void)nv;
    float4x4 M = float4x4::identity();
    if (o2w16) std::memcpy(&M, o2w16, 64);
    auto P = [&](uint32_t i) {
        float3 p{pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]};
        return o2w16 ? transform_point(M, p) : p;
    };
    size_t base = a.tris.size();
    a.tris.resize(base + nt);
    for (uint32_t t = 0; t < nt; ++t) {
        float3 v0 = P(tri3[3 * t]), v1 = P(tri3[3 * t + 1]), v2 = P(tri3[3 * t + 2]);
        a.tris[base + t] = WTri{v0, v1 - v0, v2 - v0, inst_id, t};
    }
}

// non-indexed soup fast path (triangle i = vertices 3i..3i+2), identity transform
void vo_accel_add_soup(void *h, const float *pos9n, uint64_t nt, uint32_t inst_id) {
    Accel &a = *(Accel *)h;
    size_t base = a.tris.size();
    a.tris.resize(base + nt);
    for (uint64_t t = 0; t < nt; ++t) {
        const float *p = pos9n + 9 * t;
        float3 v0{p[0], p[1], p[2]}, v1{p[3], p[4], p[5]}, v2{p[6], p[7], p[8]};
        a.tris[base + t] = WTri{v0, v1 - v0, v2 - v0, inst_id, (uint32_t)t};
    }
}

double vo_accel_build(void *h, int threads) {
    auto t0 = std::chrono::steady_clock::now();
    ((Accel *)h)->build(threads);
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
uint64_t vo_accel_num_nodes(void *h) { return ((Accel *)h)->nodes.size(); }

// mode: 0 = BVH, 1 = brute force. stats2 (nullable): [nodes, tris] totals (BVH mode).
// returns seconds spent tracing.
double vo_trace_closest(void *h, uint64_t n, const float *ox, const float *oy, const float *oz, const float *tmin,
                        const float *dx, const float *dy, const float *dz, const float *tmax,
                        uint32_t *inst, uint32_t *prim, float *u, float *v, float *t,
                        int threads, int mode, uint64_t *stats2) {
    const Accel &a = *(Accel *)h;
    std::vector<TraceStats> st(std::max(1, threads));
    auto t0 = std::chrono::steady_clock::now();
    parallel_for(n, threads, [&](uint64_t b, uint64_t e, int tid) {
        for (uint64_t i = b; i < e; ++i) {
            Ray r{{ox[i], oy[i], oz[i]}, tmin[i], {dx[i], dy[i], dz[i]}, tmax[i]};
            Hit hit = mode ? a.closest_brute(r) : a.closest(r, stats2 ? &st[tid] : nullptr);
            inst[i] = hit.inst; prim[i] = hit.prim; u[i] = hit.u; v[i] = hit.v; t[i] = hit.t;
        }
    });
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (stats2) { stats2[0] = stats2[1] = 0; for (auto &s : st) { stats2[0] += s.nodes; stats2[1] += s.tris; } }
    return dt;
}

double vo_trace_any(void *h, uint64_t n, const float *ox, const float *oy, const float *oz, const float *tmin,
                    const float *dx, const float *dy, const float *dz, const float *tmax,
                    uint8_t *occluded, int threads, int mode, uint64_t *stats2) {
    const Accel &a = *(Accel *)h;
    std::vector<TraceStats> st(std::max(1, threads));
    auto t0 = std::chrono::steady_clock::now();
    parallel_for(n, threads, [&](uint64_t b, uint64_t e, int tid) {
        for (uint64_t i = b; i < e; ++i) {
            Ray r{{ox[i], oy[i], oz[i]}, tmin[i], {dx[i], dy[i], dz[i]}, tmax[i]};
            occluded[i] = mode ? a.any_brute(r) : a.any(r, stats2 ? &st[tid] : nullptr);
        }
    });
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (stats2) { stats2[0] = stats2[1] = 0; for (auto &s : st) { stats2[0] += s.nodes; stats2[1] += s.tris; } }
    return dt;
}

}  // extern "C"

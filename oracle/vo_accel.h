// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
//
// Ray/triangle intersection, brute-force and BVH2 traversal for the oracle.
// Replaces (by specification, the code being absent) ocarina::Accel::trace_closest /
// trace_occlusion called at reference src/base/mgr/geometry.cpp:168-174, and the BVH build
// called at geometry.cpp:36-53.
//
// Normative hit semantics (DESIGN.md "Intersection contract"), shared with the CUDA path:
//   * triangles are flattened to world space: v = transform_point(o2w, pos) (FMA chain)
//   * Moller-Trumbore on (v0, e1=v1-v0, e2=v2-v0) with the exact operation order in
//     intersect_triangle() below
//   * accepted iff  t_min < t < t_max, u >= 0, v >= 0, u+v <= 1, det != 0
//   * closest hit = smallest t; equal t -> smallest (inst_id, prim_id) lexicographically
//     (makes the answer independent of traversal order and BVH shape)
//   * barycentrics follow TriangleHit::lerp = (1-u-v)*a + u*b + v*c
#pragma once
#include "vo_math.h"
#include <vector>

namespace vo {

struct Ray { float3 o; float tmin; float3 d; float tmax; };

struct Hit {
    uint32_t inst = 0xffffffffu, prim = 0xffffffffu;
    float u = 0, v = 0, t = 0;
    bool is_miss() const { return inst == 0xffffffffu; }
};

struct WTri { float3 v0, e1, e2; uint32_t inst, prim; };

struct TraceStats { uint64_t nodes = 0, tris = 0; };

// returns true and fills (t,u,v) if the ray hits the triangle inside (tmin, tmax)
inline bool intersect_triangle(const Ray &r, const WTri &tr, float tmax, float &t, float &u, float &v) {
    float3 p = cross(r.d, tr.e2);
    float det = dot(tr.e1, p);
    if (!(det != 0.0f)) return false;            // also rejects NaN
    float inv = 1.0f / det;                       // IEEE divide == __frcp_rn
    float3 s = r.o - tr.v0;
    u = dot(s, p) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return false;
    float3 q = cross(s, tr.e1);
    v = dot(r.d, q) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return false;
    t = dot(tr.e2, q) * inv;
    return t > r.tmin && t < tmax;
}

struct BVH2Node {
    float3 lo, hi;
    uint32_t left;     // internal: index of left child (right = left+1); leaf: first triangle
    uint32_t count;    // 0 = internal, else number of triangles
};

class Accel {
public:
    std::vector<WTri> tris;          // in BVH leaf order after build()
    std::vector<BVH2Node> nodes;
    void build(int threads = 1);     // binned SAH, 16 bins, leaf <= 4
    Hit closest(const Ray &r, TraceStats *st = nullptr) const;
    bool any(const Ray &r, TraceStats *st = nullptr) const;
    Hit closest_brute(const Ray &r) const;
    bool any_brute(const Ray &r) const;
private:
    void build_range(uint32_t node, uint32_t first, uint32_t count, int depth, int par_depth);
};

// a candidate (t, inst, prim) beats the current best?
inline bool better(float t, uint32_t inst, uint32_t prim, const Hit &best) {
    if (best.is_miss()) return true;
    if (t < best.t) return true;
    if (t > best.t) return false;
    return inst < best.inst || (inst == best.inst && prim < best.prim);
}

}  // namespace vo

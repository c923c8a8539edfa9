// ORACLE — test infrastructure only (see vo_math.h header). PARITY UNPINNED.
// Binned-SAH BVH2 build + ordered traversal + brute force. See vo_accel.h for the contract.
#include "vo_accel.h"
#include <atomic>
#include <thread>
#include <limits>

namespace vo {

namespace {
struct Box {
    float3 lo{std::numeric_limits<float>::max(), std::numeric_limits<float>::max(), std::numeric_limits<float>::max()};
    float3 hi{-std::numeric_limits<float>::max(), -std::numeric_limits<float>::max(), -std::numeric_limits<float>::max()};
    void grow(float3 p) { lo = vmin(lo, p); hi = vmax(hi, p); }
    void grow(const Box &b) { lo = vmin(lo, b.lo); hi = vmax(hi, b.hi); }
    float half_area() const {
        float3 e = hi - lo;
        if (e.x < 0) return 0.f;
        return e.x * e.y + e.y * e.z + e.z * e.x;
    }
};
inline float comp(float3 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

struct BuildCtx {
    std::vector<Box> tb;          // per input triangle bounds
    std::vector<float3> tc;       // per input triangle centroid
    std::vector<uint32_t> idx;    // permutation being partitioned
    std::atomic<uint32_t> next_node{1};
    float pad = 0.f;
};
BuildCtx *g_ctx = nullptr;        // build() is not re-entrant (oracle only)
}  // namespace

void Accel::build_range(uint32_t node, uint32_t first, uint32_t count, int depth, int par_depth) {
    BuildCtx &c = *g_ctx;
    Box nb, cb;
    for (uint32_t i = first; i < first + count; ++i) {
        nb.grow(c.tb[c.idx[i]]);
        cb.grow(c.tc[c.idx[i]]);
    }
    BVH2Node &n = nodes[node];
    n.lo = nb.lo; n.hi = nb.hi;
    auto make_leaf = [&]() { n.left = first; n.count = count; };
    if (count <= 2 || depth > 60) { make_leaf(); return; }

    constexpr int NB = 16;
    int best_axis = -1, best_split = -1;
    float best_cost = std::numeric_limits<float>::max();
    float3 ext = cb.hi - cb.lo;
    for (int a = 0; a < 3; ++a) {
        float e = comp(ext, a);
        if (!(e > 0.f)) continue;
        float k = NB / e, lo = comp(cb.lo, a);
        Box bb[NB]; uint32_t bc[NB] = {};
        for (uint32_t i = first; i < first + count; ++i) {
            uint32_t t = c.idx[i];
            int b = std::min(NB - 1, std::max(0, (int)((comp(c.tc[t], a) - lo) * k)));
            bb[b].grow(c.tb[t]); bc[b]++;
        }
        float ra[NB]; Box acc; uint32_t cnt = 0;
        uint32_t rc[NB];
        for (int b = NB - 1; b > 0; --b) { acc.grow(bb[b]); cnt += bc[b]; ra[b] = acc.half_area(); rc[b] = cnt; }
        acc = Box(); cnt = 0;
        for (int b = 0; b < NB - 1; ++b) {
            acc.grow(bb[b]); cnt += bc[b];
            if (cnt == 0 || rc[b + 1] == 0) continue;
            float cost = acc.half_area() * cnt + ra[b + 1] * rc[b + 1];
            if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = b; }
        }
    }
    uint32_t mid;
    if (best_axis < 0) {
        if (count <= 4) { make_leaf(); return; }
        mid = first + count / 2;                      // all centroids coincide: median split
    } else {
        float leaf_cost = nb.half_area() * count;
        if (count <= 4 && leaf_cost <= best_cost + nb.half_area()) { make_leaf(); return; }
        float e = comp(ext, best_axis), k = NB / e, lo = comp(cb.lo, best_axis);
        auto *b = c.idx.data() + first, *e2 = c.idx.data() + first + count;
        auto *m = std::partition(b, e2, [&](uint32_t t) {
            int bi = std::min(NB - 1, std::max(0, (int)((comp(c.tc[t], best_axis) - lo) * k)));
            return bi <= best_split;
        });
        mid = (uint32_t)(m - c.idx.data());
        if (mid == first || mid == first + count) mid = first + count / 2;
    }
    uint32_t l = c.next_node.fetch_add(2);
    n.left = l; n.count = 0;
    uint32_t lc = mid - first, rc2 = count - lc;
    if (par_depth > 0 && count > 65536) {
        std::thread th([&, l]() { build_range(l, first, lc, depth + 1, par_depth - 1); });
        build_range(l + 1, mid, rc2, depth + 1, par_depth - 1);
        th.join();
    } else {
        build_range(l, first, lc, depth + 1, 0);
        build_range(l + 1, mid, rc2, depth + 1, 0);
    }
}

void Accel::build(int threads) {
    BuildCtx ctx;
    g_ctx = &ctx;
    uint32_t n = (uint32_t)tris.size();
    nodes.clear();
    if (n == 0) { g_ctx = nullptr; return; }
    ctx.tb.resize(n); ctx.tc.resize(n); ctx.idx.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        const WTri &t = tris[i];
        float3 v1 = t.v0 + t.e1, v2 = t.v0 + t.e2;
        Box b; b.grow(t.v0); b.grow(v1); b.grow(v2);
        ctx.tb[i] = b;
        ctx.tc[i] = (b.lo + b.hi) * 0.5f;
        ctx.idx[i] = i;
    }
    // conservative padding so that the slab test never culls a hit the triangle test accepts
    Box all; for (auto &b : ctx.tb) all.grow(b);
    float diag = length(all.hi - all.lo);
    float pad = diag * 0x1p-20f + 1e-30f;
    for (auto &b : ctx.tb) { b.lo = b.lo - make3(pad, pad, pad); b.hi = b.hi + make3(pad, pad, pad); }
    nodes.resize(2 * (size_t)n);
    int par = 0; while ((1 << par) < threads) ++par;
    build_range(0, 0, n, 0, par);
    nodes.resize(ctx.next_node.load());
    std::vector<WTri> sorted(n);
    for (uint32_t i = 0; i < n; ++i) sorted[i] = tris[ctx.idx[i]];
    tris.swap(sorted);
    g_ctx = nullptr;
}

namespace {
// slab test; returns entry distance, or +inf on miss. Conservative by the build-time padding.
inline float slab(const BVH2Node &n, const Ray &r, float3 inv, float tmax) {
    float tx0 = (n.lo.x - r.o.x) * inv.x, tx1 = (n.hi.x - r.o.x) * inv.x;
    float ty0 = (n.lo.y - r.o.y) * inv.y, ty1 = (n.hi.y - r.o.y) * inv.y;
    float tz0 = (n.lo.z - r.o.z) * inv.z, tz1 = (n.hi.z - r.o.z) * inv.z;
    float tn = std::max(std::max(std::min(tx0, tx1), std::min(ty0, ty1)), std::max(std::min(tz0, tz1), r.tmin));
    float tf = std::min(std::min(std::max(tx0, tx1), std::max(ty0, ty1)), std::min(std::max(tz0, tz1), tmax));
    // NaN (0*inf) compares false -> treated as overlap by writing the test positively
    return (tn <= tf) ? tn : std::numeric_limits<float>::infinity();
}
inline float3 safe_inv(float3 d) {
    auto f = [](float x) { return 1.0f / (std::fabs(x) > 1e-30f ? x : std::copysign(1e-30f, x)); };
    return {f(d.x), f(d.y), f(d.z)};
}
}  // namespace

Hit Accel::closest(const Ray &r, TraceStats *st) const {
    Hit best;
    if (nodes.empty()) return best;
    float3 inv = safe_inv(r.d);
    uint32_t stack[128]; int sp = 0;
    stack[sp++] = 0;
    float tcur = r.tmax;     // shrinking bound: nodes with entry > tcur cannot hold a hit (ties kept)
    while (sp) {
        uint32_t ni = stack[--sp];
        const BVH2Node &n = nodes[ni];
        if (st) st->nodes++;
        if (!(slab(n, r, inv, tcur) <= tcur)) continue;
        if (n.count) {
            for (uint32_t i = 0; i < n.count; ++i) {
                const WTri &tr = tris[n.left + i];
                float t, u, v;
                if (st) st->tris++;
                if (intersect_triangle(r, tr, r.tmax, t, u, v) && better(t, tr.inst, tr.prim, best)) {
                    best.inst = tr.inst; best.prim = tr.prim; best.t = t; best.u = u; best.v = v;
                    tcur = t;
                }
            }
        } else {
            float t0 = slab(nodes[n.left], r, inv, tcur), t1 = slab(nodes[n.left + 1], r, inv, tcur);
            uint32_t a = n.left, b = n.left + 1;
            if (t1 < t0) { std::swap(a, b); std::swap(t0, t1); }
            if (t1 <= tcur && sp < 127) stack[sp++] = b;
            if (t0 <= tcur && sp < 127) stack[sp++] = a;
        }
    }
    return best;
}

bool Accel::any(const Ray &r, TraceStats *st) const {
    if (nodes.empty()) return false;
    float3 inv = safe_inv(r.d);
    uint32_t stack[128]; int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const BVH2Node &n = nodes[stack[--sp]];
        if (st) st->nodes++;
        if (!(slab(n, r, inv, r.tmax) <= r.tmax)) continue;
        if (n.count) {
            for (uint32_t i = 0; i < n.count; ++i) {
                float t, u, v;
                if (st) st->tris++;
                if (intersect_triangle(r, tris[n.left + i], r.tmax, t, u, v)) return true;
            }
        } else if (sp < 126) {
            stack[sp++] = n.left; stack[sp++] = n.left + 1;
        }
    }
    return false;
}

Hit Accel::closest_brute(const Ray &r) const {
    Hit best;
    for (const WTri &tr : tris) {
        float t, u, v;
        if (intersect_triangle(r, tr, r.tmax, t, u, v) && better(t, tr.inst, tr.prim, best)) {
            best.inst = tr.inst; best.prim = tr.prim; best.t = t; best.u = u; best.v = v;
        }
    }
    return best;
}

bool Accel::any_brute(const Ray &r) const {
    for (const WTri &tr : tris) {
        float t, u, v;
        if (intersect_triangle(r, tr, r.tmax, t, u, v)) return true;
    }
    return false;
}

}  // namespace vo

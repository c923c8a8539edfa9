// Triangle pre-splitting (early split clipping with the priority heuristic of Karras & Aila 2013, "Fast parallel construction of
// high-quality bounding volume hierarchies", section 4): large triangles whose boxes straddle coarse cells of the scene grid are
// cut, before the tree is built, into several REFERENCES - the same triangle with the tight box of one piece of it - so that a
// wall or a floor no longer forces every node above it to span the room. The builders (radix tree / PLOC) and the collapse then
// work on references; the wide BVH stores the triangle once per reference (48 B each), and a ray may test a split triangle
// more than once with identical results. Results of traversal do not change: every point of a triangle lies in at least one of
// its pieces, each piece's box contains it (widened by the rounding of the clip, then padded like every leaf box).
// Included by vb_bvh_build.cu only.
#pragma once
#include <cub/device/device_scan.cuh>

namespace {

constexpr int SPLIT_MAX = 31;        // extra references per triangle
constexpr int SPLIT_POLY = 10;       // a triangle clipped by the six planes of a box has at most 9 vertices

// coarsest median plane of the 2^20-cell scene grid strictly inside (lo, hi) on one axis: level 0 = the scene's middle
__device__ __forceinline__ int split_median_plane(float lo, float hi, float slo, float sext, float &plane) {
    const float S = 1048576.f;
    uint32_t ilo = (uint32_t)fminf(fmaxf((lo - slo) / sext * S, 0.f), S - 1.f), ihi = (uint32_t)fminf(fmaxf((hi - slo) / sext * S, 0.f), S - 1.f);
    if (ilo == ihi) return 64;
    int k = 31 - __clz(ilo ^ ihi);
    uint32_t cell = (ihi >> k) << k;
    plane = slo + ((float)cell / S) * sext;
    if (!(plane > lo && plane < hi)) return 64;
    return 19 - k;
}

__device__ __forceinline__ float split_comp(const float3 &v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

// priority of a triangle = (2^-level * (A_box - A_ideal))^(1/3); A_ideal = |n.x| + |n.y| + |n.z|, n = e1 x e2
__global__ void __launch_bounds__(256) k_split_priority(uint32_t T, BuildBufs b, float *prio, double *sum) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    float p = 0.f;
    if (g < T) {
        float3 slo = make_float3(ord2f(b.scene[0]), ord2f(b.scene[1]), ord2f(b.scene[2]));
        float3 shi = make_float3(ord2f(b.scene[3]), ord2f(b.scene[4]), ord2f(b.scene[5]));
        float3 se = make_float3(fmaxf(shi.x - slo.x, 1e-30f), fmaxf(shi.y - slo.y, 1e-30f), fmaxf(shi.z - slo.z, 1e-30f));
        float4 lo = b.blo[g], hi = b.bhi[g];
        float4 e1 = b.tri_in[3 * (size_t)g + 1], e2 = b.tri_in[3 * (size_t)g + 2];
        float nx = e1.y * e2.z - e1.z * e2.y, ny = e1.z * e2.x - e1.x * e2.z, nz = e1.x * e2.y - e1.y * e2.x;
        float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
        float a_box = 2.f * (dx * dy + dy * dz + dz * dx), a_ideal = fabsf(nx) + fabsf(ny) + fabsf(nz);
        float pl;
        int level = min(split_median_plane(lo.x, hi.x, slo.x, se.x, pl), min(split_median_plane(lo.y, hi.y, slo.y, se.y, pl), split_median_plane(lo.z, hi.z, slo.z, se.z, pl)));
        float norm = se.x * se.y + se.y * se.z + se.z * se.x;                 // scale-free
        float excess = fmaxf(a_box - a_ideal, 0.f) / fmaxf(norm, 1e-30f);
        if (level < 64 && excess > 0.f && isfinite(excess)) p = cbrtf(exp2f(-(float)level) * excess);
        prio[g] = p;
    }
    double d = (double)p;
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(~0u, d, o);
    if ((threadIdx.x & 31) == 0 && d > 0.0) atomicAdd(sum, d);
}

__global__ void __launch_bounds__(256) k_split_count(uint32_t T, const float *prio, float scale, uint32_t *count) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= T) return;
    count[g] = 1u + (uint32_t)fminf(floorf(prio[g] * scale), (float)SPLIT_MAX);
}

struct SplitPoly { float3 v[SPLIT_POLY]; int n; int budget; };

__device__ __forceinline__ void split_bounds(const SplitPoly &P, float3 &lo, float3 &hi) {
    lo = hi = P.v[0];
    for (int i = 1; i < P.n; ++i) {
        lo = make_float3(fminf(lo.x, P.v[i].x), fminf(lo.y, P.v[i].y), fminf(lo.z, P.v[i].z));
        hi = make_float3(fmaxf(hi.x, P.v[i].x), fmaxf(hi.y, P.v[i].y), fmaxf(hi.z, P.v[i].z));
    }
}

// one reference: box of a piece, clamped to the triangle's own (exact) box; .w of the low corner carries the triangle index
__device__ __forceinline__ void split_emit_ref(const SplitPoly &P, float4 tlo, float4 thi, uint32_t tri, float4 *rlo, float4 *rhi, uint32_t at) {
    float3 lo, hi;
    split_bounds(P, lo, hi);
    // clip points are rounded interpolations: widen by 8 ulp of the coordinate before clamping to the exact box
    const float E = 9.5367431640625e-7f;     // 2^-20
    lo.x -= E * fabsf(lo.x); lo.y -= E * fabsf(lo.y); lo.z -= E * fabsf(lo.z);
    hi.x += E * fabsf(hi.x); hi.y += E * fabsf(hi.y); hi.z += E * fabsf(hi.z);
    rlo[at] = make_float4(fmaxf(lo.x, tlo.x), fmaxf(lo.y, tlo.y), fmaxf(lo.z, tlo.z), __uint_as_float(tri));
    rhi[at] = make_float4(fminf(hi.x, thi.x), fminf(hi.y, thi.y), fminf(hi.z, thi.z), 0.f);
}

// cuts P by the plane {axis a = c} into L (<= c) and R (>= c); vertices on the plane go to both
__device__ __forceinline__ void split_clip(const SplitPoly &P, int a, float c, SplitPoly &L, SplitPoly &R) {
    L.n = R.n = 0;
    for (int i = 0; i < P.n; ++i) {
        const float3 v = P.v[i], w = P.v[i + 1 == P.n ? 0 : i + 1];
        const float va = split_comp(v, a), wa = split_comp(w, a);
        if (va <= c && L.n < SPLIT_POLY) L.v[L.n++] = v;
        if (va >= c && R.n < SPLIT_POLY) R.v[R.n++] = v;
        if ((va < c && wa > c) || (va > c && wa < c)) {
            float t = (c - va) / (wa - va);
            float3 x = make_float3(v.x + t * (w.x - v.x), v.y + t * (w.y - v.y), v.z + t * (w.z - v.z));
            if (a == 0) x.x = c; else if (a == 1) x.y = c; else x.z = c;
            if (L.n < SPLIT_POLY) L.v[L.n++] = x;
            if (R.n < SPLIT_POLY) R.v[R.n++] = x;
        }
    }
}

// depth-first splitting of one triangle into count[g] references written at offs[g]...
__global__ void __launch_bounds__(128) k_split_emit(uint32_t T, BuildBufs b, const uint32_t *__restrict__ count, const uint32_t *__restrict__ offs,
                                                    float4 *rlo, float4 *rhi) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= T) return;
    const float4 tlo = b.blo[g], thi = b.bhi[g];
    uint32_t at = offs[g];
    const int budget0 = (int)count[g] - 1;
    if (budget0 == 0) { rlo[at] = make_float4(tlo.x, tlo.y, tlo.z, __uint_as_float(g)); rhi[at] = thi; return; }
    float3 slo = make_float3(ord2f(b.scene[0]), ord2f(b.scene[1]), ord2f(b.scene[2]));
    float3 shi = make_float3(ord2f(b.scene[3]), ord2f(b.scene[4]), ord2f(b.scene[5]));
    float3 se = make_float3(fmaxf(shi.x - slo.x, 1e-30f), fmaxf(shi.y - slo.y, 1e-30f), fmaxf(shi.z - slo.z, 1e-30f));
    SplitPoly stack[SPLIT_MAX + 1];
    int sp = 0;
    {
        float4 p0 = b.tri_in[3 * (size_t)g], e1 = b.tri_in[3 * (size_t)g + 1], e2 = b.tri_in[3 * (size_t)g + 2];
        SplitPoly &P = stack[sp++];
        P.v[0] = make_float3(p0.x, p0.y, p0.z);
        P.v[1] = make_float3(p0.x + e1.x, p0.y + e1.y, p0.z + e1.z);
        P.v[2] = make_float3(p0.x + e2.x, p0.y + e2.y, p0.z + e2.z);
        P.n = 3; P.budget = budget0;
    }
    while (sp > 0) {
        SplitPoly P = stack[--sp];
        if (P.budget == 0) { split_emit_ref(P, tlo, thi, g, rlo, rhi, at++); continue; }
        float3 lo, hi;
        split_bounds(P, lo, hi);
        // the most important median plane the piece crosses; without one, the middle of its longest axis
        float pl[3];
        int lv[3] = {split_median_plane(lo.x, hi.x, slo.x, se.x, pl[0]), split_median_plane(lo.y, hi.y, slo.y, se.y, pl[1]),
                     split_median_plane(lo.z, hi.z, slo.z, se.z, pl[2])};
        float ext[3] = {hi.x - lo.x, hi.y - lo.y, hi.z - lo.z};
        int a = 0;
        for (int k = 1; k < 3; ++k) if (lv[k] < lv[a] || (lv[k] == lv[a] && ext[k] > ext[a])) a = k;
        float c = pl[a];
        if (lv[a] == 64) {
            a = ext[1] > ext[0] ? 1 : 0; if (ext[2] > ext[a]) a = 2;
            float l = split_comp(lo, a), h = split_comp(hi, a);
            c = 0.5f * (l + h);
            if (!(c > l && c < h)) {                              // nothing left to cut: spend the budget on copies (degenerate input only)
                for (int k = 0; k <= P.budget; ++k) split_emit_ref(P, tlo, thi, g, rlo, rhi, at++);
                continue;
            }
        }
        SplitPoly &L = stack[sp], &R = stack[sp + 1];
        split_clip(P, a, c, L, R);
        if (L.n < 3 || R.n < 3) {                                 // the plane only grazes the piece
            for (int k = 0; k <= P.budget; ++k) split_emit_ref(P, tlo, thi, g, rlo, rhi, at++);
            continue;
        }
        // remaining splits in proportion to the longest axes of the two pieces
        float3 llo, lhi, rlo3, rhi3;
        split_bounds(L, llo, lhi); split_bounds(R, rlo3, rhi3);
        float wl = fmaxf(lhi.x - llo.x, fmaxf(lhi.y - llo.y, lhi.z - llo.z)), wr = fmaxf(rhi3.x - rlo3.x, fmaxf(rhi3.y - rlo3.y, rhi3.z - rlo3.z));
        int rest = P.budget - 1;
        int bl = (int)floorf((float)rest * wl / fmaxf(wl + wr, 1e-30f) + 0.5f);
        bl = min(max(bl, 0), rest);
        L.budget = bl; R.budget = rest - bl;
        sp += 2;
    }
}

}  // namespace

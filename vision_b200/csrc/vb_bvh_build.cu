// GPU BVH build for sm_100a: flatten instances to world space -> 63-bit Morton codes -> radix sort
// -> Karras binary radix tree -> bottom-up refit -> level-synchronous collapse into the compressed
// 8-wide BVH (80-byte nodes, 48-byte triangles laid out contiguously per wide node).
//
// Replaces the OptiX GAS/IAS build the reference reaches through
//   mesh.build_bvh(); accel_.add_instance(mesh, o2w); accel_.build_bvh();
// at reference src/base/mgr/geometry.cpp:36-53 (ocarina, absent from the tree).
#include "vb_internal.h"
#include <cub/device/device_radix_sort.cuh>
#include <cfloat>

namespace {

// ---- order-preserving float <-> uint for atomic min/max ------------------------------------
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(uint32_t u) {
    uint32_t v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(v);
#else
    float f; memcpy(&f, &v, 4); return f;
#endif
}

struct BuildBufs {
    float4 *tri_in;        // 3*T  flattened triangles (unsorted)
    float4 *blo, *bhi;     // T    leaf bounds (unpadded): a triangle, or one piece of it (vb_bvh_split.cuh); blo.w = bits of the triangle index
    uint32_t *scene;       // 6    ordered-uint scene bounds: lo xyz, hi xyz
    uint64_t *keys, *keys2;
    uint32_t *vals, *vals2;
    float4 *nlo, *nhi;     // 2T-1 BVH2 node bounds (internal 0..T-2, leaves T-1..2T-2)
    uint32_t *left, *right, *first, *last;   // T-1
    uint32_t *parent;      // 2T-1
    uint32_t *flags;       // T-1
    uint32_t *cls_count;   // 4    triangles per size class (k_morton)
    double *moments;       // 6    sum and sum of squares of the triangle centroids (ray-sort box, size classes)
    float extra_pad;       // added to the conservative leaf padding (bottom-level trees of the two-level structure: rounding of the ray's transform)
    uint32_t absolute;     // 1: leaf records hold the three vertices (p0,prim) (p1) (p2) instead of (v0,prim) (e1,inst) (e2,0)
};

// normative world transform: FMA chain identical to oracle vo::transform_point
__device__ __forceinline__ float3 xform_point(const float *m, float3 p) {
    return make_float3(__fmaf_rn(m[8], p.z, __fmaf_rn(m[4], p.y, __fmaf_rn(m[0], p.x, m[12]))),
                       __fmaf_rn(m[9], p.z, __fmaf_rn(m[5], p.y, __fmaf_rn(m[1], p.x, m[13]))),
                       __fmaf_rn(m[10], p.z, __fmaf_rn(m[6], p.y, __fmaf_rn(m[2], p.x, m[14]))));
}

struct InstGeom {          // per instance, device table for flatten
    const float *pos; const uint32_t *tri; uint32_t soup; uint32_t pad;
    float o2w[16];
};

__global__ void __launch_bounds__(256) k_flatten(const InstGeom *inst, const uint32_t *offsets, uint32_t n_inst,
                                                 uint32_t T, BuildBufs b) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    float3 lo = make_float3(FLT_MAX, FLT_MAX, FLT_MAX), hi = make_float3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    double m[6] = {0, 0, 0, 0, 0, 0};
    if (g < T) {
        uint32_t a = 0, c = n_inst;               // largest a with offsets[a] <= g
        while (c - a > 1) { uint32_t m = (a + c) >> 1; if (offsets[m] <= g) a = m; else c = m; }
        const InstGeom &I = inst[a];
        uint32_t prim = g - offsets[a];
        uint32_t i0, i1, i2;
        if (I.soup) { i0 = 3 * prim; i1 = i0 + 1; i2 = i0 + 2; }
        else { i0 = I.tri[3 * prim]; i1 = I.tri[3 * prim + 1]; i2 = I.tri[3 * prim + 2]; }
        float3 p0 = xform_point(I.o2w, make_float3(I.pos[3 * i0], I.pos[3 * i0 + 1], I.pos[3 * i0 + 2]));
        float3 p1 = xform_point(I.o2w, make_float3(I.pos[3 * i1], I.pos[3 * i1 + 1], I.pos[3 * i1 + 2]));
        float3 p2 = xform_point(I.o2w, make_float3(I.pos[3 * i2], I.pos[3 * i2 + 1], I.pos[3 * i2 + 2]));
        float3 e1 = make_float3(__fsub_rn(p1.x, p0.x), __fsub_rn(p1.y, p0.y), __fsub_rn(p1.z, p0.z));
        float3 e2 = make_float3(__fsub_rn(p2.x, p0.x), __fsub_rn(p2.y, p0.y), __fsub_rn(p2.z, p0.z));
        b.tri_in[3 * (size_t)g + 0] = make_float4(p0.x, p0.y, p0.z, __uint_as_float(prim));
        if (b.absolute) {                      // bottom-level tree: the traversal transforms the vertices itself (vb_traverse_tl)
            b.tri_in[3 * (size_t)g + 1] = make_float4(p1.x, p1.y, p1.z, 0.f);
            b.tri_in[3 * (size_t)g + 2] = make_float4(p2.x, p2.y, p2.z, 0.f);
        } else {
            b.tri_in[3 * (size_t)g + 1] = make_float4(e1.x, e1.y, e1.z, __uint_as_float(a));
            b.tri_in[3 * (size_t)g + 2] = make_float4(e2.x, e2.y, e2.z, 0.f);
        }
        // bounds from v0, v0+e1, v0+e2 (what the intersector effectively sees)
        float3 q1 = make_float3(p0.x + e1.x, p0.y + e1.y, p0.z + e1.z);
        float3 q2 = make_float3(p0.x + e2.x, p0.y + e2.y, p0.z + e2.z);
        lo = make_float3(fminf(p0.x, fminf(fminf(p1.x, q1.x), fminf(p2.x, q2.x))),
                         fminf(p0.y, fminf(fminf(p1.y, q1.y), fminf(p2.y, q2.y))),
                         fminf(p0.z, fminf(fminf(p1.z, q1.z), fminf(p2.z, q2.z))));
        hi = make_float3(fmaxf(p0.x, fmaxf(fmaxf(p1.x, q1.x), fmaxf(p2.x, q2.x))),
                         fmaxf(p0.y, fmaxf(fmaxf(p1.y, q1.y), fmaxf(p2.y, q2.y))),
                         fmaxf(p0.z, fmaxf(fmaxf(p1.z, q1.z), fmaxf(p2.z, q2.z))));
        b.blo[g] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(g));
        b.bhi[g] = make_float4(hi.x, hi.y, hi.z, 0.f);
        m[0] = 0.5 * ((double)lo.x + (double)hi.x); m[1] = 0.5 * ((double)lo.y + (double)hi.y); m[2] = 0.5 * ((double)lo.z + (double)hi.z);
        m[3] = m[0] * m[0]; m[4] = m[1] * m[1]; m[5] = m[2] * m[2];
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        for (int o = 16; o > 0; o >>= 1) m[k] += __shfl_xor_sync(~0u, m[k], o);
        if ((threadIdx.x & 31) == 0) atomicAdd(b.moments + k, m[k]);
    }
    // warp reduce then one atomic per warp per component
    for (int o = 16; o > 0; o >>= 1) {
        lo.x = fminf(lo.x, __shfl_xor_sync(~0u, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(~0u, lo.y, o));
        lo.z = fminf(lo.z, __shfl_xor_sync(~0u, lo.z, o));
        hi.x = fmaxf(hi.x, __shfl_xor_sync(~0u, hi.x, o)); hi.y = fmaxf(hi.y, __shfl_xor_sync(~0u, hi.y, o));
        hi.z = fmaxf(hi.z, __shfl_xor_sync(~0u, hi.z, o));
    }
    if ((threadIdx.x & 31) == 0 && lo.x <= hi.x) {
        atomicMin(&b.scene[0], f2ord(lo.x)); atomicMin(&b.scene[1], f2ord(lo.y)); atomicMin(&b.scene[2], f2ord(lo.z));
        atomicMax(&b.scene[3], f2ord(hi.x)); atomicMax(&b.scene[4], f2ord(hi.y)); atomicMax(&b.scene[5], f2ord(hi.z));
    }
}

__device__ __forceinline__ uint64_t expand21(uint32_t v) {   // spread 21 bits to every 3rd bit
    uint64_t x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// Sort key = [2-bit size class][60-bit Morton code of the centroid]. The size class (triangle box diagonal against
// the spread of the centroids, 5 sigma) goes on top, so the radix tree splits the few large triangles of a scene
// (walls, floors, ground planes) from the many small ones before it splits space: each class is an LBVH over
// similar-sized boxes, and a huge box no longer inflates every ancestor of a leaf deep inside the fine geometry
// (extended Morton codes, Vinkler et al. 2017, reduced to a prefix). Uniform meshes fall into one class.
__global__ void __launch_bounds__(256) k_morton(uint32_t T, uint32_t n_moments, BuildBufs b, int size_classes) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= T) return;                                   // (the last warp may be partial: __activemask below)
    float3 slo = make_float3(ord2f(b.scene[0]), ord2f(b.scene[1]), ord2f(b.scene[2]));
    float3 shi = make_float3(ord2f(b.scene[3]), ord2f(b.scene[4]), ord2f(b.scene[5]));
    float4 lo = b.blo[g], hi = b.bhi[g];
    float ex = fmaxf(shi.x - slo.x, 1e-30f), ey = fmaxf(shi.y - slo.y, 1e-30f), ez = fmaxf(shi.z - slo.z, 1e-30f);
    float cx = (0.5f * (lo.x + hi.x) - slo.x) / ex, cy = (0.5f * (lo.y + hi.y) - slo.y) / ey,
          cz = (0.5f * (lo.z + hi.z) - slo.z) / ez;
    const float S = 1048576.f;   // 2^20
    uint32_t ix = (uint32_t)fminf(fmaxf(cx * S, 0.f), S - 1.f), iy = (uint32_t)fminf(fmaxf(cy * S, 0.f), S - 1.f),
             iz = (uint32_t)fminf(fmaxf(cz * S, 0.f), S - 1.f);
    uint64_t cls = 0;
    {
        double var = 0.0;
        for (int k = 0; k < 3; ++k) { double mean = b.moments[k] / n_moments; var += fmax(0.0, b.moments[3 + k] / n_moments - mean * mean); }
        float D = 5.0f * (float)sqrt(var);                       // spread of the geometry (robust against a few far-out triangles)
        float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
        float diag = sqrtf(dx * dx + dy * dy + dz * dz);
        cls = diag > 0.5f * D ? 3u : (diag > 0.125f * D ? 2u : (diag > 0.03125f * D ? 1u : 0u));
        unsigned act = __activemask();
        for (uint32_t k = 1; k < 4; ++k) {
            unsigned m = __ballot_sync(act, cls == k);
            if (m && (threadIdx.x & 31) == (unsigned)(__ffs(m) - 1)) atomicAdd(b.cls_count + k, (uint32_t)__popc(m));
        }
        if (!size_classes) cls = 0;
    }
    b.keys[g] = (cls << 60) | (expand21(ix) << 2) | (expand21(iy) << 1) | expand21(iz);
    b.vals[g] = g;
}

// Karras 2012: common-prefix length with index tie-break
__device__ __forceinline__ int delta(const uint64_t *keys, int T, int i, int j) {
    if (j < 0 || j >= T) return -1;
    uint64_t a = keys[i], c = keys[j];
    if (a == c) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ c));
}

__global__ void __launch_bounds__(256) k_karras(int T, const uint64_t *keys, BuildBufs b) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T - 1) return;
    int d = (delta(keys, T, i, i + 1) - delta(keys, T, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, T, i, i - d);
    int lmax = 2;
    while (delta(keys, T, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, T, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, T, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, T, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    uint32_t L = (lo == gamma) ? (uint32_t)(T - 1 + gamma) : (uint32_t)gamma;
    uint32_t R = (hi == gamma + 1) ? (uint32_t)(T - 1 + gamma + 1) : (uint32_t)(gamma + 1);
    b.left[i] = L; b.right[i] = R;
    b.first[i] = lo; b.last[i] = hi;
    b.parent[L] = i; b.parent[R] = i;
    if (i == 0) b.parent[0] = 0xffffffffu;
}

__global__ void __launch_bounds__(256) k_refit(uint32_t T, BuildBufs b) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= T) return;
    float3 slo = make_float3(ord2f(b.scene[0]), ord2f(b.scene[1]), ord2f(b.scene[2]));
    float3 shi = make_float3(ord2f(b.scene[3]), ord2f(b.scene[4]), ord2f(b.scene[5]));
    float dx = shi.x - slo.x, dy = shi.y - slo.y, dz = shi.z - slo.z;
    // conservative padding (DESIGN.md "Intersection contract"): slab tests never cull a hit
    // that the normative triangle test accepts
    float pad = sqrtf(dx * dx + dy * dy + dz * dz) * 0x1p-20f + 1e-30f + b.extra_pad;
    uint32_t src = b.vals[g];
    float4 lo = b.blo[src], hi = b.bhi[src];
    lo.x -= pad; lo.y -= pad; lo.z -= pad; hi.x += pad; hi.y += pad; hi.z += pad;
    uint32_t node = T - 1 + g;
    b.nlo[node] = lo; b.nhi[node] = hi;
    if (T == 1) return;
    uint32_t p = b.parent[node];
    while (p != 0xffffffffu) {
        __threadfence();
        if (atomicAdd(&b.flags[p], 1u) == 0) return;     // first arrival leaves; second continues
        uint32_t L = b.left[p], R = b.right[p];
        float4 a = __ldcg(&b.nlo[L]), c = __ldcg(&b.nlo[R]);
        float4 e = __ldcg(&b.nhi[L]), f = __ldcg(&b.nhi[R]);
        b.nlo[p] = make_float4(fminf(a.x, c.x), fminf(a.y, c.y), fminf(a.z, c.z), 0.f);
        b.nhi[p] = make_float4(fmaxf(e.x, f.x), fmaxf(e.y, f.y), fmaxf(e.z, f.z), 0.f);
        p = b.parent[p];
    }
}

// ---- two-level structure: world box of every instance, and the top-level build's "primitives" ----
// exact world bounds of an instance = union over its triangles of what k_flatten would compute for them (same transform chain)
__global__ void __launch_bounds__(256) k_inst_bounds(const InstGeom *inst, const uint32_t *offsets, uint32_t n_inst, uint32_t T, uint32_t *boxes) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= T) return;
    uint32_t a = 0, c = n_inst;
    while (c - a > 1) { uint32_t m = (a + c) >> 1; if (offsets[m] <= g) a = m; else c = m; }
    const InstGeom &I = inst[a];
    uint32_t prim = g - offsets[a];
    uint32_t i0, i1, i2;
    if (I.soup) { i0 = 3 * prim; i1 = i0 + 1; i2 = i0 + 2; }
    else { i0 = I.tri[3 * prim]; i1 = I.tri[3 * prim + 1]; i2 = I.tri[3 * prim + 2]; }
    float3 p0 = xform_point(I.o2w, make_float3(I.pos[3 * i0], I.pos[3 * i0 + 1], I.pos[3 * i0 + 2]));
    float3 p1 = xform_point(I.o2w, make_float3(I.pos[3 * i1], I.pos[3 * i1 + 1], I.pos[3 * i1 + 2]));
    float3 p2 = xform_point(I.o2w, make_float3(I.pos[3 * i2], I.pos[3 * i2 + 1], I.pos[3 * i2 + 2]));
    float3 e1 = make_float3(__fsub_rn(p1.x, p0.x), __fsub_rn(p1.y, p0.y), __fsub_rn(p1.z, p0.z));
    float3 e2 = make_float3(__fsub_rn(p2.x, p0.x), __fsub_rn(p2.y, p0.y), __fsub_rn(p2.z, p0.z));
    float3 q1 = make_float3(p0.x + e1.x, p0.y + e1.y, p0.z + e1.z), q2 = make_float3(p0.x + e2.x, p0.y + e2.y, p0.z + e2.z);
    float lo[3] = {fminf(p0.x, fminf(fminf(p1.x, q1.x), fminf(p2.x, q2.x))), fminf(p0.y, fminf(fminf(p1.y, q1.y), fminf(p2.y, q2.y))),
                   fminf(p0.z, fminf(fminf(p1.z, q1.z), fminf(p2.z, q2.z)))};
    float hi[3] = {fmaxf(p0.x, fmaxf(fmaxf(p1.x, q1.x), fmaxf(p2.x, q2.x))), fmaxf(p0.y, fmaxf(fmaxf(p1.y, q1.y), fmaxf(p2.y, q2.y))),
                   fmaxf(p0.z, fmaxf(fmaxf(p1.z, q1.z), fmaxf(p2.z, q2.z)))};
    for (int k = 0; k < 3; ++k) { atomicMin(&boxes[6 * a + k], f2ord(lo[k])); atomicMax(&boxes[6 * a + 3 + k], f2ord(hi[k])); }
}
// top-level "triangles": one per instance - record (bits of the root node of its mesh's tree, instance id), box = the instance's world box
__global__ void __launch_bounds__(256) k_tlas_prims(uint32_t n_inst, const uint32_t *boxes, const uint32_t *roots, BuildBufs b) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_inst) return;
    float lo[3], hi[3];
    for (int k = 0; k < 3; ++k) { lo[k] = ord2f(boxes[6 * g + k]); hi[k] = ord2f(boxes[6 * g + 3 + k]); }
    b.tri_in[3 * (size_t)g + 0] = make_float4(__uint_as_float(roots[g]), __uint_as_float(g), 0.f, __uint_as_float(g));
    b.tri_in[3 * (size_t)g + 1] = make_float4(0.f, 0.f, 0.f, __uint_as_float(g));
    b.tri_in[3 * (size_t)g + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
    b.blo[g] = make_float4(lo[0], lo[1], lo[2], __uint_as_float(g));
    b.bhi[g] = make_float4(hi[0], hi[1], hi[2], 0.f);
    for (int k = 0; k < 3; ++k) { atomicMin(&b.scene[k], f2ord(lo[k])); atomicMax(&b.scene[3 + k], f2ord(hi[k])); }
    for (int k = 0; k < 3; ++k) {
        double m = 0.5 * ((double)lo[k] + (double)hi[k]);
        atomicAdd(b.moments + k, m); atomicAdd(b.moments + 3 + k, m * m);
    }
}
// a bottom-level tree appended behind others: its nodes' child / triangle indices become absolute
__global__ void __launch_bounds__(256) k_offset_nodes(vb_wide_node *nodes, uint32_t n, uint32_t node_off, uint32_t tri_off) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    nodes[g].child_base += node_off; nodes[g].tri_base += tri_off;
}

}  // namespace

#include "vb_bvh_split.cuh"
#include "vb_bvh_ploc.cuh"
#include "vb_bvh_collapse.cuh"

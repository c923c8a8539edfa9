// Device traversal core for the compressed 8-wide BVH (sm_100a). Shared by the batch trace
// kernels (vb_trace.cu) and the path-tracing megakernel (vb_pt.cu).
//
// Stands in for ocarina::Accel::trace_closest / trace_occlusion (OptiX) called from
// reference src/base/mgr/geometry.cpp:168-174.
//
// Node = 5 x 128-bit read-only loads (ld.global.nc.v4); triangle = 3 x 128-bit loads.
// Traversal stack: VB_SM_STACK entries per thread in shared memory laid out [entry][thread]
// (conflict-free), spilling to a small local array only on very deep paths.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef VB_SM_STACK
#define VB_SM_STACK 12
#endif
#ifndef VB_LOCAL_STACK
#define VB_LOCAL_STACK (64 - VB_SM_STACK)
#endif
#ifdef VB_TRAVERSAL_STACK_ENTRIES
static_assert(VB_SM_STACK + VB_LOCAL_STACK == VB_TRAVERSAL_STACK_ENTRIES, "vb_internal.h states the traversal stack size the build checks against");
#endif

struct VbRay { float ox, oy, oz, tmin, dx, dy, dz, tmax; };
struct VbHit { uint32_t inst, prim; float u, v, t; };

struct VbAccelView {
    const uint4 *__restrict__ nodes;
    const float4 *__restrict__ tris;
    // two-level structure ("build.two_level"): per instance 6 float4 - rows 0..2 of object-to-world (x y z | translation), rows 0..2 of
    // world-to-object; nullptr = the flattened single-level structure (default)
    const float4 *__restrict__ xf;
};

// Normative triangle test (DESIGN.md "Intersection contract"); mirrors oracle
// vo::intersect_triangle operation for operation. __f*_rn intrinsics are never contracted.
__device__ __forceinline__ float vb_dot(float ax, float ay, float az, float bx, float by, float bz) {
    return __fmaf_rn(az, bz, __fmaf_rn(ay, by, __fmul_rn(ax, bx)));
}
__device__ __forceinline__ void vb_cross(float ax, float ay, float az, float bx, float by, float bz,
                                         float &cx, float &cy, float &cz) {
    cx = __fmaf_rn(ay, bz, -__fmul_rn(az, by));
    cy = __fmaf_rn(az, bx, -__fmul_rn(ax, bz));
    cz = __fmaf_rn(ax, by, -__fmul_rn(ay, bx));
}
__device__ __forceinline__ bool vb_intersect_triangle(const VbRay &r, float4 a, float4 b, float4 c, float tmax,
                                                      float &t, float &u, float &v) {
    float px, py, pz;
    vb_cross(r.dx, r.dy, r.dz, c.x, c.y, c.z, px, py, pz);
    float det = vb_dot(b.x, b.y, b.z, px, py, pz);
    if (!(det != 0.0f)) return false;
    float inv = __frcp_rn(det);
    float sx = __fsub_rn(r.ox, a.x), sy = __fsub_rn(r.oy, a.y), sz = __fsub_rn(r.oz, a.z);
    u = __fmul_rn(vb_dot(sx, sy, sz, px, py, pz), inv);
    if (!(u >= 0.0f && u <= 1.0f)) return false;
    float qx, qy, qz;
    vb_cross(sx, sy, sz, b.x, b.y, b.z, qx, qy, qz);
    v = __fmul_rn(vb_dot(r.dx, r.dy, r.dz, qx, qy, qz), inv);
    if (!(v >= 0.0f && __fadd_rn(u, v) <= 1.0f)) return false;
    t = __fmul_rn(vb_dot(c.x, c.y, c.z, qx, qy, qz), inv);
    return t > r.tmin && t < tmax;
}

// node fetch: 96-byte slot, 32-byte aligned -> 2 x ld.global.nc.v8 (256-bit, sm_100+) + 1 x v4
#ifndef VB_NODE_U4
#define VB_NODE_U4 6
#endif
__device__ __forceinline__ void vb_ld256(const uint4 *p, uint4 &a, uint4 &b) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p));
}
__device__ __forceinline__ void vb_load_node(const uint4 *nodes, uint32_t index, uint4 &n0, uint4 &n1, uint4 &n2, uint4 &n3, uint4 &n4) {
    const uint4 *np = nodes + (size_t)index * VB_NODE_U4;
    vb_ld256(np, n0, n1);
    vb_ld256(np + 2, n2, n3);
    n4 = __ldg(np + 4);
}

__device__ __forceinline__ float vb_safe_rcp(float d) {
    return 1.0f / (fabsf(d) > 1e-20f ? d : copysignf(1e-20f, d));
}

__device__ __forceinline__ uint32_t vb_byte(uint32_t w, int i) { return (w >> (8 * i)) & 0xffu; }

// Quantised byte -> float, balanced over the two math pipes of an SMSP (each issues one warp
// instruction every 2 cycles; the box test was ALU-pipe bound - profiles/r1_trace_v4_leaf1_ncu_full.txt):
//   ONE PRMT (alu pipe) builds a half2 from two bytes: 0x6400 | byte = fp16(1024 + byte), exact;
//   HADD2.F32 (fma pipe) widens each half to fp32. The 1024 bias is folded into the FMA addend per
//   node; its rounding error is covered by the builder inflating every child box by 2^-6 grid units.
// The 0x64646464 operand must live in a REGISTER so that the SELECTOR is the PRMT immediate (SASS
// PRMT has one non-register operand slot). ptxas constant-folds any "mov r, imm" and then
// re-materialises one selector register per PRMT, so the value is read back from shared memory
// once per kernel, which ptxas cannot fold. Every kernel that traverses calls vb_trav_init() first.
#define VB_Q_BIAS 1024.f
// Per-kernel traversal constants: the PRMT operand above and the shared address of the octant permutation table
// perm[o][m] = the 8-bit mask m with bit s moved to bit s ^ o: box-test results come out in SLOT order (one bit per
// child), one table look-up puts the internal children's bits into the ray's front-to-back order.
struct VbTrav { uint32_t mg, lut; };
__device__ __forceinline__ VbTrav vb_trav_init() {
    __shared__ uint32_t s_vb_magic;
    __shared__ __align__(16) uint8_t s_vb_perm[8 * 256];
    if (threadIdx.x == 0) s_vb_magic = 0x64646464u;
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += blockDim.x) {
        uint32_t o = i >> 8, m = i & 255u, q = 0;
#pragma unroll
        for (uint32_t b = 0; b < 8; ++b) q |= ((m >> b) & 1u) << (b ^ o);
        s_vb_perm[i] = (uint8_t)q;
    }
    __syncthreads();
    VbTrav t;
    t.mg = *(volatile uint32_t *)&s_vb_magic;
    t.lut = (uint32_t)__cvta_generic_to_shared(s_vb_perm);
    return t;
}
__device__ __forceinline__ uint32_t vb_lds8(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
template <int J>
__device__ __forceinline__ float vb_qf(uint32_t w, uint32_t magic) {
    uint32_t d;      // bytes (J&~1, J|1) of w as two fp16 values 1024+byte; identical PRMTs are CSE'd per pair
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(magic), "n"(J < 2 ? 0x4140 : 0x4342));
    float f;
    if (J & 1) asm("{ .reg .b16 lo, hi; mov.b32 {lo, hi}, %1; cvt.f32.f16 %0, hi; }" : "=f"(f) : "r"(d));
    else asm("{ .reg .b16 lo, hi; mov.b32 {lo, hi}, %1; cvt.f32.f16 %0, lo; }" : "=f"(f) : "r"(d));
    return f;
}

// Tests the 8 quantised child boxes of one node; returns the 8-bit hit mask in SLOT order (bit s = child in slot s).
// Empty slots hold an inverted box and never hit. Per child: 6 decodes + 6 FMA + 4 min/max, then the SIGN of
// cmax - cmin is shifted into the mask (one FADD on the fma pipe + one funnel shift) - no predicate, no per-child
// bit-position decode; slots are visited 7..0 so that the first result ends up in bit 7.
__device__ __forceinline__ uint32_t vb_node_hits(const uint4 n0, const uint4 n2, const uint4 n3, const uint4 n4, const VbRay &r,
                                                 float idx, float idy, float idz, float tmax, const uint32_t mg) {
    float adjx = __uint_as_float((n0.w & 0xffu) << 23) * idx;
    float adjy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * idy;
    float adjz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * idz;
    float orx = fmaf(-VB_Q_BIAS, adjx, (__uint_as_float(n0.x) - r.ox) * idx);
    float ory = fmaf(-VB_Q_BIAS, adjy, (__uint_as_float(n0.y) - r.oy) * idy);
    float orz = fmaf(-VB_Q_BIAS, adjz, (__uint_as_float(n0.z) - r.oz) * idz);
    uint32_t miss = 0;
#pragma unroll
    for (int h = 1; h >= 0; --h) {
        uint32_t qlx = h ? n2.y : n2.x, qly = h ? n2.w : n2.z, qlz = h ? n3.y : n3.x;
        uint32_t qhx = h ? n3.w : n3.z, qhy = h ? n4.y : n4.x, qhz = h ? n4.w : n4.z;
        uint32_t nx = idx < 0.f ? qhx : qlx, fx = idx < 0.f ? qlx : qhx;
        uint32_t ny = idy < 0.f ? qhy : qly, fy = idy < 0.f ? qly : qhy;
        uint32_t nz = idz < 0.f ? qhz : qlz, fz = idz < 0.f ? qlz : qhz;
#define VB_CHILD(J)                                                                                      \
        {                                                                                                \
            float tnx = fmaf(vb_qf<J>(nx, mg), adjx, orx), tfx = fmaf(vb_qf<J>(fx, mg), adjx, orx);              \
            float tny = fmaf(vb_qf<J>(ny, mg), adjy, ory), tfy = fmaf(vb_qf<J>(fy, mg), adjy, ory);              \
            float tnz = fmaf(vb_qf<J>(nz, mg), adjz, orz), tfz = fmaf(vb_qf<J>(fz, mg), adjz, orz);              \
            float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, r.tmin));                                     \
            float cmax = fminf(fminf(tfx, tfy), fminf(tfz, tmax));                                       \
            /* sign set <=> cmax < cmin (x - x = +0; a NaN difference reads as a hit: conservative) */   \
            miss = __funnelshift_l(__float_as_uint(__fsub_rn(cmax, cmin)), miss, 1);                     \
        }
        VB_CHILD(3) VB_CHILD(2) VB_CHILD(1) VB_CHILD(0)
#undef VB_CHILD
    }
    return ~miss & 0xffu;
}

// One node step shared by every traversal loop: fetch node `index`, test its children, return
//   G  = (first internal child, hit internal children in front-to-back order << 24 | internal-slot mask)
//   T  = (first triangle, hit leaf slots << 8 | occupied leaf slots)      (T.y <= 0xff: no triangle hit)
// lut_o = VbTrav::lut + (7 - octant) * 256.
__device__ __forceinline__ void vb_node_step(const uint4 *nodes, uint32_t index, const VbRay &r, float idx, float idy, float idz, float tmax,
                                             uint32_t mg, uint32_t lut_o, uint2 &G, uint2 &T) {
    uint4 n0, n1, n2, n3, n4;
    vb_load_node(nodes, index, n0, n1, n2, n3, n4);
    uint32_t hit8 = vb_node_hits(n0, n2, n3, n4, r, idx, idy, idz, tmax, mg);
    uint32_t imask = n0.w >> 24;
    G = make_uint2(n1.x, (vb_lds8(lut_o + (hit8 & imask)) << 24) | imask);
    T = make_uint2(n1.y, ((hit8 & ~imask & n1.z) << 8) | (n1.z & 0xffu));      // & leaf slots: an empty slot can never report a triangle
}
// next pending triangle of T: clears its bit, returns its index (triangles of a node are stored in slot order)
__device__ __forceinline__ uint32_t vb_next_triangle(uint2 &T) {
    uint32_t bit = 1u << (31u - __clz(T.y));       // bits 8..15
    T.y ^= bit;
    return T.x + __popc(T.y & ((bit - 1u) >> 8));  // occupied leaf slots below it
}

// 64-bit shared-memory access by 32-bit shared address (no generic -> shared conversion at the access)
__device__ __forceinline__ uint2 vb_lds64(uint32_t a) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ void vb_sts64(uint32_t a, uint2 v) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(v.x), "r"(v.y) : "memory");
}

// closest-hit bookkeeping with the normative tie rule (smallest t, then smallest (inst, prim))
__device__ __forceinline__ void vb_commit_hit(VbHit &hit, float &tmax, float4 a, float4 b, float t, float u, float v) {
    uint32_t inst = __float_as_uint(b.w), prim = __float_as_uint(a.w);
    bool better = hit.inst == 0xffffffffu || t < hit.t ||
                  (t == hit.t && (inst < hit.inst || (inst == hit.inst && prim < hit.prim)));
    if (better) { hit.inst = inst; hit.prim = prim; hit.u = u; hit.v = v; hit.t = t; tmax = t; }
}

template <bool ANY, bool COUNT>
__device__ __forceinline__ bool vb_traverse_tl(const VbAccelView &A, const VbRay rw, VbHit &hit, uint2 *s_stack, int stride,
                                            uint32_t &n_nodes, uint32_t &n_tris, const VbTrav V);

// s_stack points at this thread's column of the shared stack: entry e lives at s_stack[e * stride]
template <bool ANY, bool COUNT>
__device__ __forceinline__ bool vb_traverse(const VbAccelView &A, VbRay r, VbHit &hit, uint2 *s_stack, int stride,
                                            uint32_t &n_nodes, uint32_t &n_tris, const VbTrav V) {
    if (A.xf) return vb_traverse_tl<ANY, COUNT>(A, r, hit, s_stack, stride, n_nodes, n_tris, V);      // two-level structure (uniform branch)
    uint2 l_stack[VB_LOCAL_STACK];
    int sp = 0;
    float idx = vb_safe_rcp(r.dx), idy = vb_safe_rcp(r.dy), idz = vb_safe_rcp(r.dz);
    uint32_t oct = (r.dx < 0.f ? 4u : 0u) | (r.dy < 0.f ? 2u : 0u) | (r.dz < 0.f ? 1u : 0u);
    const uint32_t oct_inv = 7u - oct, lut_o = V.lut + (oct_inv << 8);
    float tmax = r.tmax;
    hit.inst = 0xffffffffu; hit.prim = 0xffffffffu; hit.u = 0.f; hit.v = 0.f; hit.t = 0.f;
    if (!(r.tmin < r.tmax)) return false;

    uint2 G = make_uint2(0u, 0x80000000u);     // root group: node 0, "slot 7" hit, imask 0
    uint2 Tg = make_uint2(0u, 0u);
    for (;;) {
        // invariant: G holds at least one un-visited internal child here (only such groups are pushed)
        {
            uint32_t hits_imask = G.y;
            uint32_t bit = 31u - __clz(hits_imask);
            G.y &= ~(1u << bit);
            if (G.y & 0xff000000u) {           // siblings remain: push the rest of the group
                if (sp < VB_SM_STACK) s_stack[sp * stride] = G;
                else if (sp < VB_SM_STACK + VB_LOCAL_STACK) l_stack[sp - VB_SM_STACK] = G;
                sp++;
            }
            uint32_t slot = (bit - 24u) ^ oct_inv;
            uint32_t rel = __popc(hits_imask & ~(0xffffffffu << slot) & 0xffu);
            if (COUNT) n_nodes++;
            vb_node_step(A.nodes, G.x + rel, r, idx, idy, idz, tmax, V.mg, lut_o, G, Tg);
        }
        while (Tg.y > 0xffu) {
            const float4 *tp = A.tris + (size_t)vb_next_triangle(Tg) * 3;
            float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
            if (COUNT) n_tris++;
            float t, u, v;
            if (vb_intersect_triangle(r, a, b, c, r.tmax, t, u, v)) {
                if (ANY) return true;
                vb_commit_hit(hit, tmax, a, b, t, u, v);
            }
        }
        if (!(G.y & 0xff000000u)) {
            if (sp == 0) break;
            sp--;
            if (sp < VB_SM_STACK) G = s_stack[sp * stride];
            else if (sp < VB_SM_STACK + VB_LOCAL_STACK) G = l_stack[sp - VB_SM_STACK];
            else G = make_uint2(0u, 0u);
        }
    }
    return hit.inst != 0xffffffffu;
}

// ---- two-level traversal (option "build.two_level"; reference geometry.cpp:36-53: one GAS per instance + IAS) ------------------------
// The top level is a wide BVH over the instances' WORLD boxes; a leaf slot of it is one instance (record: (bits of the root node of its
// mesh's bottom-level tree, instance id)). A bottom-level tree is built once per unique mesh in OBJECT space; its leaf records hold the
// three object-space vertices (p0, prim) (p1) (p2). Inside an instance the boxes are tested against the ray moved into object space
// (same parameter t: the transform is affine and the direction is not re-normalised), but every TRIANGLE is tested in WORLD space: its
// vertices go through the normative object-to-world chain (the one k_flatten uses), so v0, e1, e2 - and with them t, u, v and the winner
// of every tie - are bit for bit those of the flattened structure. The bottom-level boxes carry an extra padding that covers the
// rounding of the ray's transform (vb_bvh_collapse.cuh). Stack entries: node groups as in vb_traverse; a LEAF group (pending instances of
// a top-level node: top byte 0, bits 8..15 set); the SENTINEL (x = 0xffffffff) that marks the way back out of an instance.
#define VB_TL_SENTINEL 0xffffffffu
__device__ __forceinline__ float3 vb_xf_point(float4 r0, float4 r1, float4 r2, float x, float y, float z) {     // == xform_point of the build
    return make_float3(__fmaf_rn(r0.z, z, __fmaf_rn(r0.y, y, __fmaf_rn(r0.x, x, r0.w))),
                       __fmaf_rn(r1.z, z, __fmaf_rn(r1.y, y, __fmaf_rn(r1.x, x, r1.w))),
                       __fmaf_rn(r2.z, z, __fmaf_rn(r2.y, y, __fmaf_rn(r2.x, x, r2.w))));
}
template <bool ANY, bool COUNT>
__device__ __forceinline__ bool vb_traverse_tl(const VbAccelView &A, const VbRay rw, VbHit &hit, uint2 *s_stack, int stride,
                                            uint32_t &n_nodes, uint32_t &n_tris, const VbTrav V) {
    uint2 l_stack[VB_LOCAL_STACK];
    int sp = 0;
    auto push = [&](uint2 e) {
        if (sp < VB_SM_STACK) s_stack[sp * stride] = e;
        else if (sp < VB_SM_STACK + VB_LOCAL_STACK) l_stack[sp - VB_SM_STACK] = e;
        sp++;
    };
    auto pop = [&]() -> uint2 {
        sp--;
        if (sp < VB_SM_STACK) return s_stack[sp * stride];
        if (sp < VB_SM_STACK + VB_LOCAL_STACK) return l_stack[sp - VB_SM_STACK];
        return make_uint2(0u, 0u);
    };
    hit.inst = 0xffffffffu; hit.prim = 0xffffffffu; hit.u = 0.f; hit.v = 0.f; hit.t = 0.f;
    if (!(rw.tmin < rw.tmax)) return false;
    float tmax = rw.tmax;
    VbRay r = rw;                                   // the ray of the current space (world, or the object space of cur_inst)
    float idx, idy, idz; uint32_t oct_inv, lut_o;
    auto set_space = [&]() {
        idx = vb_safe_rcp(r.dx); idy = vb_safe_rcp(r.dy); idz = vb_safe_rcp(r.dz);
        uint32_t oct = (r.dx < 0.f ? 4u : 0u) | (r.dy < 0.f ? 2u : 0u) | (r.dz < 0.f ? 1u : 0u);
        oct_inv = 7u - oct; lut_o = V.lut + (oct_inv << 8);
    };
    set_space();
    uint32_t cur_inst = 0xffffffffu;
    uint2 G = make_uint2(0u, 0x80000000u);          // root group of the top level
    uint2 Tg = make_uint2(0u, 0u);
    for (;;) {
        if (G.y & 0xff000000u) {                    // a node group with un-visited children: visit the nearest
            uint32_t hits_imask = G.y;
            uint32_t bit = 31u - __clz(hits_imask);
            G.y &= ~(1u << bit);
            if (G.y & 0xff000000u) push(G);
            uint32_t slot = (bit - 24u) ^ oct_inv;
            uint32_t rel = __popc(hits_imask & ~(0xffffffffu << slot) & 0xffu);
            if (COUNT) n_nodes++;
            vb_node_step(A.nodes, G.x + rel, r, idx, idy, idz, tmax, V.mg, lut_o, G, Tg);
        }
        while (Tg.y > 0xffu) {
            const uint32_t rec = vb_next_triangle(Tg);
            const float4 *tp = A.tris + (size_t)rec * 3;
            if (cur_inst == 0xffffffffu) {          // top level: the leaf is an instance - enter it
                const float4 a = __ldg(tp);
                if (G.y & 0xff000000u) push(G);
                if (Tg.y > 0xffu) push(Tg);
                push(make_uint2(VB_TL_SENTINEL, 0u));
                cur_inst = __float_as_uint(a.y);
                const float4 *xf = A.xf + (size_t)cur_inst * 6;
                const float4 w0 = __ldg(xf + 3), w1 = __ldg(xf + 4), w2 = __ldg(xf + 5);
                float3 o = vb_xf_point(w0, w1, w2, rw.ox, rw.oy, rw.oz);
                r.ox = o.x; r.oy = o.y; r.oz = o.z;
                r.dx = __fmaf_rn(w0.z, rw.dz, __fmaf_rn(w0.y, rw.dy, __fmul_rn(w0.x, rw.dx)));
                r.dy = __fmaf_rn(w1.z, rw.dz, __fmaf_rn(w1.y, rw.dy, __fmul_rn(w1.x, rw.dx)));
                r.dz = __fmaf_rn(w2.z, rw.dz, __fmaf_rn(w2.y, rw.dy, __fmul_rn(w2.x, rw.dx)));
                set_space();
                G = make_uint2(__float_as_uint(a.x), 0x80000000u);      // root group of the instance's bottom-level tree
                Tg = make_uint2(0u, 0u);
                break;
            }
            // bottom level: the triangle, moved to world space by the build's own transform chain, against the WORLD ray
            const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
            const float4 *xf = A.xf + (size_t)cur_inst * 6;
            const float4 m0 = __ldg(xf), m1 = __ldg(xf + 1), m2 = __ldg(xf + 2);
            const float3 p0 = vb_xf_point(m0, m1, m2, q0.x, q0.y, q0.z), p1 = vb_xf_point(m0, m1, m2, q1.x, q1.y, q1.z),
                         p2 = vb_xf_point(m0, m1, m2, q2.x, q2.y, q2.z);
            const float4 a = make_float4(p0.x, p0.y, p0.z, q0.w);
            const float4 b = make_float4(__fsub_rn(p1.x, p0.x), __fsub_rn(p1.y, p0.y), __fsub_rn(p1.z, p0.z), __uint_as_float(cur_inst));
            const float4 c = make_float4(__fsub_rn(p2.x, p0.x), __fsub_rn(p2.y, p0.y), __fsub_rn(p2.z, p0.z), 0.f);
            if (COUNT) n_tris++;
            float t, u, v;
            if (vb_intersect_triangle(rw, a, b, c, rw.tmax, t, u, v)) {
                if (ANY) return true;
                vb_commit_hit(hit, tmax, a, b, t, u, v);
            }
        }
        if (!(G.y & 0xff000000u) && !(Tg.y > 0xffu)) {          // nothing left here: next stack entry
            for (;;) {
                if (sp == 0) return hit.inst != 0xffffffffu;
                uint2 e = pop();
                if (e.x == VB_TL_SENTINEL && e.y == 0u) {       // back out of the instance
                    cur_inst = 0xffffffffu; r = rw; set_space();
                    continue;
                }
                if (e.y & 0xff000000u) { G = e; Tg = make_uint2(0u, 0u); break; }
                if (e.y > 0xffu) { Tg = e; G = make_uint2(0u, 0u); break; }
            }
        }
    }
}

// ---- ray sort key: 21-bit Morton code of the origin (128^3 cells of the scene box) + 3-bit direction
// octant = 24 bits = 3 onesweep radix passes. fminf/fmaxf drop NaN operands, so NaN origins land in cell 0.
__device__ __forceinline__ uint32_t vb_expand8(uint32_t v) {      // 8 bits -> every 3rd bit
    v &= 0xffu;
    v = (v | (v << 8)) & 0x00f00fu;
    v = (v | (v << 4)) & 0x0c30c3u;
    v = (v | (v << 2)) & 0x249249u;
    return v;
}
// oct_shift s: the octant sits ABOVE the low s Morton bits, so runs of 2^s neighbouring cells share one octant
__device__ __forceinline__ uint32_t vb_ray_sort_key(float ox, float oy, float oz, float dx, float dy, float dz, float3 lo, float3 inv_ext,
                                                    uint32_t oct_shift = 0) {
    uint32_t ix = (uint32_t)fminf(fmaxf((ox - lo.x) * inv_ext.x * 128.f, 0.f), 127.f);
    uint32_t iy = (uint32_t)fminf(fmaxf((oy - lo.y) * inv_ext.y * 128.f, 0.f), 127.f);
    uint32_t iz = (uint32_t)fminf(fmaxf((oz - lo.z) * inv_ext.z * 128.f, 0.f), 127.f);
    uint32_t oct = (dx < 0.f ? 4u : 0u) | (dy < 0.f ? 2u : 0u) | (dz < 0.f ? 1u : 0u);
    uint32_t m = (vb_expand8(ix) << 2) | (vb_expand8(iy) << 1) | vb_expand8(iz);
    return ((m >> oct_shift) << (oct_shift + 3)) | (oct << oct_shift) | (m & ((1u << oct_shift) - 1u));
}

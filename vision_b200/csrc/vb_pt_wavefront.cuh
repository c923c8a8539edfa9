// Wavefront schedule of the path tracer. Included by vb_pt.cu only.
//
// B200 has no RT cores and no hardware thread re-ordering, so a megakernel keeps a whole warp inside
// the traversal loop until its slowest lane is done: ncu on classroom shows 6 of 32 lanes active in the
// traversal code and ~40 % of the issue slots lost to instruction-cache misses of the 30k-instruction
// kernel (profiles/r2_pt_classroom_v1_ncu_full.txt). Here the same per-path computation (pt_bounce,
// vb_pt_bounce.cuh - bit-identical results) runs as a pipeline of small kernels over queues in HBM:
//
//   generate -> [ trace closest (sorted, persistent warps: vb_trace.cu) -> shade (pt_bounce) ->
//                 trace any -> shadow resolve ] x bounces -> film
//
//  * a batch is F frames x all pixels = P paths; path slot s = f * n_pix + pix holds the path state
//    (48 bytes, three 16-byte streams); queues hold compacted 32-byte ray records, the slot they belong
//    to and the 64-bit sort record (Morton key of the ray << 32 | queue index) the trace kernel orders
//    its fetches by - the producer kernel writes it, so tracing needs no key pass and no copy
//  * closest hits come back as one 16-byte record per entry; before shading the entries are ordered by
//    material (one 8-bit radix pass), so a warp evaluates one BSDF model instead of five
//  * the shade kernel appends the continuation ray and the shadow ray to the next queues with one
//    warp-aggregated atomicAdd each; finished paths simply are not appended
//  * the per-sample radiance stays in the slot; the film kernel folds the F frames of a pixel in frame
//    order, so LERP/SUM accumulation is bit-identical to the sequential reference loop
#pragma once

namespace {

struct WfCounters { uint32_t n_next, n_shadow, pad0, pad1; unsigned long long vertices, culled; };

struct WfState {
    float4 *a;      // throughput.xyz, eta_scale
    float4 *b;      // ret.xyz, scatter_pdf
    uint4 *c;       // sampler state, ray_medium, bounces, hero wavelength word (PathState::swl)
};
// a ray queue: AoS ray records (o.xyz, tmin)(d.xyz, tmax), the sort record (vb_ray_sort_key << 32 | queue index)
// the trace kernel orders its fetches by, and the path slot each entry belongs to
struct WfQueue { float4 *rays; uint64_t *keyidx; uint32_t *slot; };
struct WfSortBox { float3 lo, inv_ext; uint32_t oct_shift; };

__device__ __forceinline__ void wf_push_ray(const WfQueue &Q, const WfSortBox &B, uint32_t i, const VbRay &r) {
    Q.rays[2 * (size_t)i] = make_float4(r.ox, r.oy, r.oz, r.tmin);
    Q.rays[2 * (size_t)i + 1] = make_float4(r.dx, r.dy, r.dz, r.tmax);
    Q.keyidx[i] = ((uint64_t)vb_ray_sort_key(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, B.lo, B.inv_ext, B.oct_shift) << 32) | i;
}
__device__ __forceinline__ VbRay wf_load_ray(const float4 *rays, uint32_t i) {
    float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
    VbRay r;
    r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
    return r;
}

// camera rays of frames [fi0, fi0 + nf) of the batch; queue position == slot for the first bounce
__global__ void __launch_bounds__(256) k_wf_generate(const __grid_constant__ PtScene S, uint32_t first_frame, uint32_t stride, uint32_t fi0, uint32_t n_paths,
                                                     WfQueue Q, WfSortBox B, WfState st) {
    uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_paths) return;
    uint32_t n_pix = S.width * S.height;
    uint32_t f = slot / n_pix, pix = slot - f * n_pix;
    uint32_t px = pix % S.width, py = pix / S.width;
    uint32_t frame = first_frame + (fi0 + f) * stride;
    Sampler sampler;
    sampler.start(px, py, frame, 0u);
    VbRay ray = generate_camera_ray(S, sampler, px, py);
    PathState ps;
    path_begin(S, ps, px, py, frame);
    wf_push_ray(Q, B, slot, ray);
    Q.slot[slot] = slot;
    st.a[slot] = make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, ps.eta_scale);
    st.b[slot] = make_float4(ps.ret.x, ps.ret.y, ps.ret.z, ps.scatter_pdf);
    st.c[slot] = make_uint4(ps.sampler.state, ps.ray_medium, ps.bounces, ps.swl);
}

// 3-bit class of the shading sort key: the lobe-SET materials (principled, mix, add) share class 5 = the heavy launch; class 7 is
// never used (keys 254 / 255 are "no material" / "miss")
enum : uint32_t { WF_CLASS_SET = 5u };
__device__ __forceinline__ uint32_t wf_material_class(uint32_t type) {
    switch (type) {
        case VB_MAT_PRINCIPLED: case VB_MAT_MIX: case VB_MAT_ADD: return WF_CLASS_SET;
        case VB_MAT_PLASTIC: return VB_MAT_SUBSTRATE;
        case VB_MAT_METALLIC: return VB_MAT_METAL;
        default: return type & 7u;
    }
}
// shading order: entries with the same material run in the same warps (one 8-bit radix pass over these records)
__global__ void __launch_bounds__(256) k_wf_material_keys(PtScene S, uint32_t n, const uint4 *__restrict__ hits4, uint64_t *keyidx) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    uint32_t inst = __ldg(&hits4[q].x);
    uint32_t key = 255u;                                            // miss
    if (inst != VB_INVALID_ID) {
        uint32_t mat = S.instances[inst].mat_id;
        key = mat == VB_INVALID_ID ? 254u : (wf_material_class(S.materials[mat].type) << 5) | (mat & 31u);
    }
    keyidx[q] = ((uint64_t)key << 32) | q;
}

// one Li vertex per queue entry; entry i of the launch is queue position order[i] (low word).
// filter (needs `order`, whose high word is the material key of k_wf_material_keys): 0 = every entry, 1 = all but the
// principled-BSDF entries, 2 = only those. Scenes that use the principled BSDF shade each bounce with TWO launches over the
// material-ordered queue - the lean kernel (no stack frame, no spills) for everything else, the principled instantiation for its
// own entries; warps whose entries all belong to the other launch exit at once.
template <int VARIANT>
#ifndef VB_SHADE_MIN_BLOCKS
#define VB_SHADE_MIN_BLOCKS 5      /* 96 registers, 20 warps per SM: the lean kernels are latency-bound (profiles/round2_classroom_shade_trace_ncu_full.txt); 4 -> 5: classroom +1 %, dispersion +5 %; 6: no further gain */
#endif
#ifndef VB_SHADE_MIN_BLOCKS_PRINCIPLED
#define VB_SHADE_MIN_BLOCKS_PRINCIPLED 4      /* 128 registers, 16 warps per SM (3 x 168 registers until round 2: dispersion +1.6 %, cbox_principled +2.6 %, tools/sweep_shade_occ.sh) */
#endif
__global__ void __launch_bounds__(128, (VARIANT & PT_PRINCIPLED) ? VB_SHADE_MIN_BLOCKS_PRINCIPLED : VB_SHADE_MIN_BLOCKS) k_wf_shade(const __grid_constant__ PtScene S, uint32_t n, WfQueue Qin, const uint64_t *__restrict__ order, const uint4 *__restrict__ hits4,
                                                  WfState st, WfQueue Qout, WfQueue Qsh, float4 *sh_pend, WfSortBox B, WfCounters *cnt, int filter) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31u;
    int flags = 0;
    VbRay ray{}, shadow{};
    f3 pending{0, 0, 0};
    uint32_t slot = 0;
    bool mine = i < n;
    uint32_t q = i;
    if (mine && order) {
        uint64_t o = __ldg(order + i);
        q = (uint32_t)o;
        if (filter) {
            uint32_t key = (uint32_t)(o >> 32);
            bool principled = key < 254u && ((key >> 5) & 7u) == WF_CLASS_SET;
            mine = principled == (filter == 2);
        }
    }
    if (!__any_sync(~0u, mine)) return;
    if (mine) {
        slot = __ldg(Qin.slot + q);
        ray = wf_load_ray(Qin.rays, q);
        uint4 h = __ldg(hits4 + q);
        VbHit hit;
        hit.inst = h.x; hit.prim = h.y; hit.u = __uint_as_float(h.z); hit.v = __uint_as_float(h.w); hit.t = 0.f;
        float4 a = st.a[slot], b = st.b[slot];
        uint4 c = st.c[slot];
        PathState ps;
        ps.throughput = f3{a.x, a.y, a.z}; ps.eta_scale = a.w;
        ps.ret = f3{b.x, b.y, b.z}; ps.scatter_pdf = b.w;
        ps.sampler.state = c.x; ps.ray_medium = c.y; ps.bounces = c.z; ps.swl = c.w;
        flags = pt_bounce<VARIANT>(S, ps, ray, hit, shadow, pending);
        st.b[slot] = make_float4(ps.ret.x, ps.ret.y, ps.ret.z, ps.scatter_pdf);
        if (flags & PT_CONTINUE) {
            st.a[slot] = make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, ps.eta_scale);
            st.c[slot] = make_uint4(ps.sampler.state, ps.ray_medium, ps.bounces, ps.swl);
        }
    }
    // queue appends: one atomic per warp and queue
    unsigned m_next = __ballot_sync(~0u, flags & PT_CONTINUE);
    unsigned m_sh = __ballot_sync(~0u, flags & PT_SHADOW);
    unsigned m_vert = __ballot_sync(~0u, flags & PT_VERTEX);
    unsigned m_cull = __ballot_sync(~0u, (flags & PT_LIGHT) && !(flags & PT_SHADOW));
    uint32_t base_next = 0, base_sh = 0;
    if (lane == 0) {
        if (m_next) base_next = atomicAdd(&cnt->n_next, (uint32_t)__popc(m_next));
        if (m_sh) base_sh = atomicAdd(&cnt->n_shadow, (uint32_t)__popc(m_sh));
        if (m_vert) atomicAdd(&cnt->vertices, (unsigned long long)__popc(m_vert));
        if (m_cull) atomicAdd(&cnt->culled, (unsigned long long)__popc(m_cull));
    }
    base_next = __shfl_sync(~0u, base_next, 0);
    base_sh = __shfl_sync(~0u, base_sh, 0);
    const unsigned lt = (1u << lane) - 1u;
    if (flags & PT_CONTINUE) {
        uint32_t o = base_next + __popc(m_next & lt);
        wf_push_ray(Qout, B, o, ray);
        Qout.slot[o] = slot;
    }
    if (flags & PT_SHADOW) {
        uint32_t o = base_sh + __popc(m_sh & lt);
        wf_push_ray(Qsh, B, o, shadow);
        sh_pend[o] = make_float4(pending.x, pending.y, pending.z, __uint_as_float(slot));
    }
}

// ret += pending for every free shadow ray (each slot has at most one shadow entry per bounce)
__global__ void __launch_bounds__(256) k_wf_shadow_resolve(uint32_t n, const uint8_t *__restrict__ occ, const float4 *__restrict__ sh_pend, float4 *st_b) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || occ[i]) return;
    float4 p = sh_pend[i];
    uint32_t slot = __float_as_uint(p.w);
    float4 b = st_b[slot];
    b.x = b.x + p.x; b.y = b.y + p.y; b.z = b.z + p.z;
    st_b[slot] = b;
}

// fold the nf finished frames of every pixel into the film, in frame order
__global__ void __launch_bounds__(256) k_wf_film(uint32_t n_pix, uint32_t nf, uint32_t fi0, int mode, uint32_t frames_done,
                                                 const float4 *__restrict__ st_b, float4 *accum, float4 *rt) {
    uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= n_pix) return;
    float4 acc = accum[pix];
    float4 last = make_float4(0.f, 0.f, 0.f, 1.f);
    for (uint32_t f = 0; f < nf; ++f) {
        float4 b = st_b[(size_t)f * n_pix + pix];
        last = make_float4(b.x, b.y, b.z, 1.f);
        if (mode == VB_ACCUM_LERP) {
            float t = 1.0f / (float)(frames_done + fi0 + f + 1);
            acc.x = acc.x + t * (last.x - acc.x); acc.y = acc.y + t * (last.y - acc.y);
            acc.z = acc.z + t * (last.z - acc.z); acc.w = acc.w + t * (last.w - acc.w);
        } else {
            acc.x += last.x; acc.y += last.y; acc.z += last.z; acc.w += last.w;
        }
    }
    accum[pix] = acc; rt[pix] = last;
}

}  // namespace

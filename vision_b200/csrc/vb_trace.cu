// Batch trace kernels + C ABI entry points: vb_trace_closest / vb_trace_any (+ counted variants).
// Replace ocarina::Accel::trace_closest / trace_occlusion (reference src/base/mgr/geometry.cpp:168-174)
// as a bulk SoA-in / SoA-out API (the microbench and parity surface of SURVEY.md §8b).
#include "vb_internal.h"
#include "vb_trace.cuh"
#include <algorithm>
#include <mutex>
#include <cub/device/device_radix_sort.cuh>

namespace {

constexpr int TRACE_BLOCK = 128;
constexpr int TRACE_MIN_BLOCKS = 8;

struct RaysDev { const float *ox, *oy, *oz, *tmin, *dx, *dy, *dz, *tmax; };
struct HitsDev { uint32_t *inst, *prim; float *u, *v, *t; };

// Persistent-warp traversal with per-lane dynamic ray fetch and triangle postponing.
//  * each lane owns one ray at a time; when fewer than REFILL_THR lanes of a warp are busy the
//    idle lanes grab consecutive ray indices with one warp-aggregated atomicAdd
//  * every loop iteration is two warp-converged phases: (N) lanes with internal-child work pop one
//    child and test its 8 boxes; (T) triangle rounds, run only when >= TRI_THR lanes have pending
//    triangles or some lane cannot continue without testing its triangles (both pending slots
//    full, or no node work left) - warp votes decide, so no lane spins alone in a triangle loop
//  * results are independent of order by the tie rule, so postponing never changes the answer

template <bool ANY, bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_MIN_BLOCKS) k_trace(VbAccelView A, RaysDev R, uint64_t n, HitsDev H, uint8_t *occ,
                                                       unsigned long long *counters, unsigned long long *next_ray,
                                                       int REFILL_THR, int TRI_THR,
                                                       const float4 *__restrict__ RS, const uint32_t *__restrict__ perm, int perm_stride,
                                                       uint4 *H4) {
    // traversal stack: VB_SM_STACK entries per thread in shared memory, [entry][thread] (conflict-free), 1 KB per entry row.
    // The stack pointer is kept as a 32-bit SHARED ADDRESS spa = s0 + 8 * tid + 1024 * depth (s0 1 KB aligned, uniform):
    // push / pop are one ld/st.shared and one add, and "depth > 0" is one compare against a uniform register.
    __shared__ __align__(1024) uint2 s_stack_all[VB_SM_STACK * TRACE_BLOCK];
    static_assert(TRACE_BLOCK * sizeof(uint2) == 1024, "stack row = 1 KB");
    uint32_t s0;        // through an asm so that ptxas keeps it in a register instead of re-deriving it (S2R + 3 ALU) at every use
    asm volatile("mov.u32 %0, %1;" : "=r"(s0) : "r"((uint32_t)__cvta_generic_to_shared(s_stack_all)));
    const uint32_t s_row1 = s0 + 1024u, s_end = s0 + 1024u * VB_SM_STACK;
    uint32_t spa = s0 + 8u * threadIdx.x;
    uint2 l_stack[VB_LOCAL_STACK];
    const VbTrav V = vb_trav_init();
    const uint32_t mg = V.mg;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    uint32_t nn = 0, nt = 0;

    bool active = false, exhausted = false;
    constexpr uint64_t NO_RESULT = ~0ull;
    uint64_t idx = NO_RESULT;
    VbRay r{};
    float idx_ = 0.f, idy_ = 0.f, idz_ = 0.f, tmax = 0.f;
    uint32_t oct_inv = 0, lut_o = 0;
    VbHit hit{};
    uint2 G = make_uint2(0u, 0u), Tg = make_uint2(0u, 0u), Td = make_uint2(0u, 0u);
    bool occluded = false;

    for (;;) {
        // ---------------- results of the rays that finished since the last refill ----------------
        // (stored here, several lanes at a time, instead of one by one inside the traversal loop)
        if (!active && idx != NO_RESULT) {
            if (ANY) occ[idx] = occluded ? 1 : 0;
            else if (H4) H4[idx] = make_uint4(hit.inst, hit.prim, __float_as_uint(hit.u), __float_as_uint(hit.v));   // queue form: one 16-byte record
            else { H.inst[idx] = hit.inst; H.prim[idx] = hit.prim; H.u[idx] = hit.u; H.v[idx] = hit.v; H.t[idx] = hit.t; }
            idx = NO_RESULT;
        }
        // ---------------- refill idle lanes ----------------
        unsigned idle = __ballot_sync(~0u, !active);
        if (idle && !exhausted) {
            unsigned long long base = 0;
            int leader = __ffs(idle) - 1;
            if ((int)lane == leader)        // plain atom: the compiler's own warp aggregation around atomicAdd is wasted on one lane
                asm volatile("atom.global.add.u64 %0, [%1], %2;" : "=l"(base) : "l"(next_ray), "l"((unsigned long long)__popc(idle)) : "memory");
            base = __shfl_sync(~0u, base, leader);
            if (base + __popc(idle) >= n) exhausted = true;
            if (!active) {
                uint64_t q = base + __popc(idle & lt_mask);
                if (q < n) {
                    idx = q;
                    if (RS) {       // sorted order: gather ray perm[idx] from the AoS copy (one 32-byte sector); results go back to that slot
                        if (perm) idx = __ldg(perm + idx * perm_stride);      // stride 2: low word of a 64-bit (key << 32 | index) record
                        float4 a = __ldg(RS + 2 * idx), b = __ldg(RS + 2 * idx + 1);
                        r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
                    } else {
                        r.ox = __ldg(R.ox + idx); r.oy = __ldg(R.oy + idx); r.oz = __ldg(R.oz + idx); r.tmin = __ldg(R.tmin + idx);
                        r.dx = __ldg(R.dx + idx); r.dy = __ldg(R.dy + idx); r.dz = __ldg(R.dz + idx); r.tmax = __ldg(R.tmax + idx);
                    }
                    idx_ = vb_safe_rcp(r.dx); idy_ = vb_safe_rcp(r.dy); idz_ = vb_safe_rcp(r.dz);
                    uint32_t oct = (r.dx < 0.f ? 4u : 0u) | (r.dy < 0.f ? 2u : 0u) | (r.dz < 0.f ? 1u : 0u);
                    oct_inv = 7u - oct; lut_o = V.lut + (oct_inv << 8);
                    tmax = r.tmax;
                    hit.inst = 0xffffffffu; hit.prim = 0xffffffffu; hit.u = 0.f; hit.v = 0.f; hit.t = 0.f;
                    occluded = false;
                    spa = s0 | (spa & 1023u); Tg = make_uint2(0u, 0u); Td = make_uint2(0u, 0u);
                    // root group; an empty interval (or NaN bounds) never enters the tree
                    G = make_uint2(0u, (r.tmin < r.tmax) ? 0x80000000u : 0u);
                    active = true;
                }
            }
        }
        if (__ballot_sync(~0u, active) == 0u) break;

        // ---------------- traverse ----------------
        unsigned act;
        do {
            bool has_node = false;
            if (active && !(Tg.y > 0xffu && Td.y > 0xffu)) {                  // both triangle slots full: a triangle round has to run first
                if (G.y < 0x01000000u && spa >= s_row1) {     // fetch the next node group
                    spa -= 1024u;
                    if (spa < s_end) G = vb_lds64(spa);
                    else {
                        uint32_t e = (spa - s_end) >> 10;
                        G = e < VB_LOCAL_STACK ? l_stack[e] : make_uint2(0u, 0u);
                    }
                }
                has_node = G.y >= 0x01000000u;
            }
            // ---- phase N: one child box-test per lane ----
            if (has_node) {
                uint32_t hits_imask = G.y;
                uint32_t bit = 31u - __clz(hits_imask);
                G.y &= ~(1u << bit);
                if (G.y >= 0x01000000u) {
                    if (spa < s_end) vb_sts64(spa, G);
                    else {
                        uint32_t e = (spa - s_end) >> 10;
                        if (e < VB_LOCAL_STACK) l_stack[e] = G;
                    }
                    spa += 1024u;
                }
                uint32_t slot = (bit - 24u) ^ oct_inv;
                uint32_t rel = __popc(hits_imask & ~(0xffffffffu << slot) & 0xffu);
                if (COUNT) nn++;
                uint2 T;
                vb_node_step(A.nodes, G.x + rel, r, idx_, idy_, idz_, tmax, mg, lut_o, G, T);
                if (T.y > 0xffu) {
                    if (Tg.y > 0xffu) Td = Tg;             // at most one of them was occupied (has_node)
                    Tg = T;
                }
            }
            // ---- phase T: warp-voted triangle rounds ----
            {
                // A round runs when enough lanes hold pending triangles, or when no lane of the warp
                // could do node work this iteration (everyone is blocked on its triangles). A lane that
                // is blocked alone just idles through phase N: cheaper than a 1-lane triangle round.
                bool tri = active && Tg.y > 0xffu;
                unsigned tm = __ballot_sync(~0u, tri);
                bool force = __ballot_sync(~0u, has_node) == 0u;
                while (tm && (force || __popc(tm) >= TRI_THR)) {
                    if (tri) {
                        const float4 *tp = A.tris + (size_t)vb_next_triangle(Tg) * 3;
                        float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                        if (COUNT) nt++;
                        float t, u, v;
                        if (vb_intersect_triangle(r, a, b, c, r.tmax, t, u, v)) {
                            if (ANY) { occluded = true; Tg.y = 0; Td.y = 0; G.y = 0; spa = s0 | (spa & 1023u); }
                            else vb_commit_hit(hit, tmax, a, b, t, u, v);
                        }
                        if (Tg.y <= 0xffu) { Tg = Td; Td = make_uint2(0u, 0u); }
                        tri = Tg.y > 0xffu;
                    }
                    tm = __ballot_sync(~0u, tri);
                }
            }
            // ---- finished rays leave the loop; their results are stored at the next refill ----
            if (active && G.y < 0x01000000u && spa < s_row1 && Tg.y <= 0xffu) active = false;
            act = __ballot_sync(~0u, active);
        } while (act && (exhausted || __popc(act) >= REFILL_THR));
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(~0u, nn, o); nt += __shfl_xor_sync(~0u, nt, o); }
        if (lane == 0) { atomicAdd(&counters[0], nn); atomicAdd(&counters[1], nt); }
    }
}


// Two-level structure ("build.two_level"): same arguments as k_trace, one ray per thread, grid-stride, vb_traverse_tl over this thread's
// column of the shared stack. A kernel of its own so that the persistent flattened kernel above keeps its registers.
template <bool ANY, bool COUNT>
#ifndef VB_TL_MIN_BLOCKS
#define VB_TL_MIN_BLOCKS 8      /* 64 registers: 4 -> 8 resident CTAs = +10..16 % (tools/bench_two_level.py) */
#endif
__global__ void __launch_bounds__(TRACE_BLOCK, VB_TL_MIN_BLOCKS) k_trace_tl(VbAccelView A, RaysDev R, uint64_t n, HitsDev H, uint8_t *occ,
                                                          unsigned long long *counters, unsigned long long *next_ray,
                                                          int REFILL_THR, int TRI_THR,
                                                          const float4 *__restrict__ RS, const uint32_t *__restrict__ perm, int perm_stride,
                                                          uint4 *H4) {
    __shared__ uint2 s_stack_all[VB_SM_STACK * TRACE_BLOCK];
    const VbTrav V = vb_trav_init();
    const unsigned lane = threadIdx.x & 31u;
    uint32_t nn = 0, nt = 0;
    // warp-uniform trip count, lanes re-converge after every ray (a lane that finished early would otherwise run the fetch / store code
    // of its next ray divergently against the others' traversal)
    for (uint64_t i0 = (uint64_t)blockIdx.x * TRACE_BLOCK + (threadIdx.x & ~31u); i0 < n; i0 += (uint64_t)gridDim.x * TRACE_BLOCK, __syncwarp()) {
        const uint64_t i = i0 + lane;
        if (i >= n) continue;
        uint64_t ri = i;
        VbRay r;
        if (RS) {
            if (perm) ri = __ldg(perm + i * perm_stride);
            float4 a = __ldg(RS + 2 * ri), b = __ldg(RS + 2 * ri + 1);
            r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tmin = a.w; r.dx = b.x; r.dy = b.y; r.dz = b.z; r.tmax = b.w;
        } else {
            r.ox = __ldg(R.ox + ri); r.oy = __ldg(R.oy + ri); r.oz = __ldg(R.oz + ri); r.tmin = __ldg(R.tmin + ri);
            r.dx = __ldg(R.dx + ri); r.dy = __ldg(R.dy + ri); r.dz = __ldg(R.dz + ri); r.tmax = __ldg(R.tmax + ri);
        }
        VbHit h;
        bool found = vb_traverse_tl<ANY, COUNT>(A, r, h, s_stack_all + threadIdx.x, TRACE_BLOCK, nn, nt, V);
        if (ANY) occ[ri] = found ? 1 : 0;
        else if (H4) H4[ri] = make_uint4(h.inst, h.prim, __float_as_uint(h.u), __float_as_uint(h.v));
        else { H.inst[ri] = h.inst; H.prim[ri] = h.prim; H.u[ri] = h.u; H.v[ri] = h.v; H.t[ri] = h.t; }
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(~0u, nn, o); nt += __shfl_xor_sync(~0u, nt, o); }
        if (lane == 0) { atomicAdd(&counters[0], nn); atomicAdd(&counters[1], nt); }
    }
}
template <bool ANY, bool COUNT, class... Args>
static void launch_trace(unsigned grid, cudaStream_t st, const VbAccelView &A, Args... args) {
    if (A.xf) k_trace_tl<ANY, COUNT><<<grid, TRACE_BLOCK, 0, st>>>(A, args...);
    else k_trace<ANY, COUNT><<<grid, TRACE_BLOCK, 0, st>>>(A, args...);
}

// ---- ray sorting: 24-bit Morton code of the origin (scene-normalised) + 3-bit direction octant ----
// Incoherent rays become spatially coherent batches: consecutive rays walk the same subtrees, so
// node/triangle fetches hit in L1/L2 instead of going to HBM as random 80/48-byte gathers.
__global__ void __launch_bounds__(256) k_ray_keys(RaysDev R, uint32_t n, float3 lo, float3 inv_ext, uint32_t *keys, uint32_t *vals,
                                                  float4 *aos, uint32_t oct_shift) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float ox = __ldg(R.ox + i), oy = __ldg(R.oy + i), oz = __ldg(R.oz + i);
    float dx = __ldg(R.dx + i), dy = __ldg(R.dy + i), dz = __ldg(R.dz + i);
    // fminf/fmaxf drop NaN operands, so NaN origins land in cell 0
    keys[i] = vb_ray_sort_key(ox, oy, oz, dx, dy, dz, lo, inv_ext, oct_shift);
    vals[i] = i;
    // SoA -> AoS copy (streaming) so that the post-sort gather touches ONE 32-byte sector per ray
    aos[2 * (size_t)i] = make_float4(ox, oy, oz, __ldg(R.tmin + i));
    aos[2 * (size_t)i + 1] = make_float4(dx, dy, dz, __ldg(R.tmax + i));
}
// Host buffers: the call is cut into `trace.host_chunks` ray ranges and pipelined over three streams - range k+1 is
// copied in (copy engine 1) while range k is sorted and traced (SMs) and range k-1 is copied out (copy engine 2) - so
// the call costs about max(H2D, trace) instead of H2D + trace + D2H. Device buffers: one range, no extra streams.
int run_trace(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, const vb_hits_soa *out, uint8_t *occ, vb_mem occ_mem,
              bool any, uint64_t *cnt_nodes, uint64_t *cnt_tris) {
    if (!ctx) return VB_ERR_INVALID;
    const char *fn = any ? "vb_trace_any" : "vb_trace_closest";
    if (!ctx->accel.built) return ctx->fail(VB_ERR_STATE, std::string(fn) + ": accel not built");
    if (!rays || (!any && !out) || (any && !occ)) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": null argument");
    if (n == 0) { if (cnt_nodes) *cnt_nodes = 0; if (cnt_tris) *cnt_tris = 0; return VB_OK; }
    if (n > (1ull << 31) * TRACE_BLOCK) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": too many rays");
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    bool count = cnt_nodes || cnt_tris;
    bool rays_host = rays->mem == VB_MEM_HOST;
    bool out_host = any ? (occ_mem == VB_MEM_HOST) : (out->mem == VB_MEM_HOST);

    // ranges: multiples of 1024 rays so that every sub-array keeps its alignment
    int C = 1;
    if ((rays_host || out_host) && ctx->opt_host_chunks > 1 && n >= (uint64_t)ctx->opt_host_chunk_min_rays * 2)
        C = (int)std::min<uint64_t>((uint64_t)std::min(ctx->opt_host_chunks, VB_MAX_HOST_CHUNKS), n / (uint64_t)ctx->opt_host_chunk_min_rays);
    const uint64_t per = C > 1 ? (((n + C - 1) / C + 1023) & ~1023ull) : n;
    if (C > 1) { int rc = ctx->ensure_copy_streams(); if (rc) return rc; }
    cudaStream_t s_in = C > 1 ? ctx->copy_in : st, s_out = C > 1 ? ctx->copy_out : st;

    // staging layout: [8 ray arrays][hit arrays or occlusion bytes][counters: nodes, tris, pad, pad, next_ray per range]
    size_t ray_bytes = rays_host ? 8 * n * sizeof(float) : 0;
    size_t out_bytes = out_host ? (any ? ((n + 15) & ~15ull) : 5 * n * 4) : 0;
    size_t need = ray_bytes + out_bytes + 128 + 8 * (size_t)VB_MAX_HOST_CHUNKS;
    int rc = ctx->ensure_stage(need);
    if (rc) return rc;
    char *base = (char *)ctx->stage;
    RaysDev R{rays->ox, rays->oy, rays->oz, rays->tmin, rays->dx, rays->dy, rays->dz, rays->tmax};
    const float *h_src[8] = {rays->ox, rays->oy, rays->oz, rays->tmin, rays->dx, rays->dy, rays->dz, rays->tmax};
    float *d_rays = (float *)base;
    if (rays_host) R = RaysDev{d_rays, d_rays + n, d_rays + 2 * n, d_rays + 3 * n, d_rays + 4 * n, d_rays + 5 * n, d_rays + 6 * n, d_rays + 7 * n};
    HitsDev H{};
    uint8_t *d_occ = occ;
    if (any) { if (out_host) d_occ = (uint8_t *)(base + ray_bytes); }
    else {
        H = HitsDev{out->inst, out->prim, out->u, out->v, out->t};
        if (out_host) {
            uint32_t *d = (uint32_t *)(base + ray_bytes);
            H = HitsDev{d, d + n, (float *)(d + 2 * n), (float *)(d + 3 * n), (float *)(d + 4 * n)};
        }
    }
    unsigned long long *d_cnt = (unsigned long long *)(base + ray_bytes + out_bytes);
    d_cnt = (unsigned long long *)(((uintptr_t)d_cnt + 15) & ~(uintptr_t)15);
    VB_CUDA(cudaMemsetAsync(d_cnt, 0, 32 + 8 * (size_t)VB_MAX_HOST_CHUNKS, st));
    if (C > 1) {        // the copy streams start after whatever the caller queued on the context's stream
        VB_CUDA(cudaEventRecord(ctx->ev_in[0], st));
        VB_CUDA(cudaStreamWaitEvent(s_in, ctx->ev_in[0], 0));
        VB_CUDA(cudaStreamWaitEvent(s_out, ctx->ev_in[0], 0));
    }

    VbAccelView A{ctx->accel.nodes, ctx->accel.tris, ctx->accel.xf};
    uint64_t launches = 0;
    ctx->prof_begin(any ? "trace_any" : "trace_closest", 1);
    for (int c = 0; c < C; ++c) {
        const uint64_t r0 = (uint64_t)c * per;
        if (r0 >= n) break;
        const uint64_t m = std::min(per, n - r0);
        if (rays_host) {
            for (int k = 0; k < 8; ++k)
                VB_CUDA(cudaMemcpyAsync(d_rays + k * n + r0, h_src[k] + r0, m * sizeof(float), cudaMemcpyHostToDevice, s_in));
            if (C > 1) { VB_CUDA(cudaEventRecord(ctx->ev_in[c], s_in)); VB_CUDA(cudaStreamWaitEvent(st, ctx->ev_in[c], 0)); }
        }
        RaysDev Rc{R.ox + r0, R.oy + r0, R.oz + r0, R.tmin + r0, R.dx + r0, R.dy + r0, R.dz + r0, R.tmax + r0};
        HitsDev Hc{};
        if (!any) Hc = HitsDev{H.inst + r0, H.prim + r0, H.u + r0, H.v + r0, H.t + r0};
        uint8_t *occ_c = any ? d_occ + r0 : nullptr;
        unsigned long long *d_next = d_cnt + 4 + c;

        // persistent grid: a multiple of the SM count, never more threads than rays
        unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count * ctx->opt_blocks_per_sm, (m + TRACE_BLOCK - 1) / TRACE_BLOCK);
        const float4 *RS = nullptr; const uint32_t *perm = nullptr;
        bool do_sort = ctx->opt_sort_rays && m >= (uint64_t)ctx->opt_sort_min_rays && m < (1ull << 32);
        launches += 1;
        if (do_sort) {
            size_t cub_bytes = 0;
            cub::DoubleBuffer<uint32_t> dk0(nullptr, nullptr), dv0(nullptr, nullptr);
            VB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, dk0, dv0, (int)m, ctx->opt_sort_skip_bits, 24, st));
            size_t nb = ((m * 4 + 255) & ~255ull);
            size_t total = 4 * nb + 32 * m + cub_bytes + 256;
            if (total > ctx->sort_bytes) {
                VB_CUDA(cudaStreamSynchronize(st));            // an earlier range may still be using the old buffer
                if (ctx->sort_buf) cudaFree(ctx->sort_buf);
                ctx->sort_buf = nullptr; ctx->sort_bytes = 0;
                VB_CUDA(cudaMalloc(&ctx->sort_buf, total));
                ctx->sort_bytes = total;
            }
            char *sb = (char *)ctx->sort_buf;
            uint32_t *k0 = (uint32_t *)sb, *k1 = (uint32_t *)(sb + nb), *v0 = (uint32_t *)(sb + 2 * nb), *v1 = (uint32_t *)(sb + 3 * nb);
            float4 *d_aos = (float4 *)(sb + 4 * nb);
            void *cub_tmp = (void *)(((uintptr_t)(sb + 4 * nb + 32 * m) + 255) & ~(uintptr_t)255);
            float3 lo = make_float3(ctx->sort_lo[0], ctx->sort_lo[1], ctx->sort_lo[2]);
            float3 ie = make_float3(1.f / fmaxf(ctx->sort_hi[0] - ctx->sort_lo[0], 1e-30f), 1.f / fmaxf(ctx->sort_hi[1] - ctx->sort_lo[1], 1e-30f),
                                    1.f / fmaxf(ctx->sort_hi[2] - ctx->sort_lo[2], 1e-30f));
            unsigned g2 = (unsigned)((m + 255) / 256);
            k_ray_keys<<<g2, 256, 0, st>>>(Rc, (uint32_t)m, lo, ie, k0, v0, d_aos, (uint32_t)ctx->opt_sort_oct_shift);
            cub::DoubleBuffer<uint32_t> dk(k0, k1), dv(v0, v1);
            cudaError_t ce = cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, dk, dv, (int)m, ctx->opt_sort_skip_bits, 24, st);
            if (ce != cudaSuccess) { ctx->prof_end(); return ctx->fail(VB_ERR_CUDA, std::string("ray sort: ") + cudaGetErrorString(ce)); }
            perm = dv.Current();
            RS = d_aos;                 // k_trace gathers ray perm[i] itself: no separate reorder pass
            launches += 5;
        }
        if (any) {
            if (count) launch_trace<true, true>(grid, st, A, Rc, m, Hc, occ_c, d_cnt, d_next, ctx->opt_refill_thr, ctx->opt_tri_thr, RS, perm, 1, nullptr);
            else launch_trace<true, false>(grid, st, A, Rc, m, Hc, occ_c, d_cnt, d_next, ctx->opt_refill_thr, ctx->opt_tri_thr, RS, perm, 1, nullptr);
        } else {
            if (count) launch_trace<false, true>(grid, st, A, Rc, m, Hc, occ_c, d_cnt, d_next, ctx->opt_refill_thr, ctx->opt_tri_thr, RS, perm, 1, nullptr);
            else launch_trace<false, false>(grid, st, A, Rc, m, Hc, occ_c, d_cnt, d_next, ctx->opt_refill_thr, ctx->opt_tri_thr, RS, perm, 1, nullptr);
        }
        if (out_host) {
            if (C > 1) { VB_CUDA(cudaEventRecord(ctx->ev_out[c], st)); VB_CUDA(cudaStreamWaitEvent(s_out, ctx->ev_out[c], 0)); }
            if (any) VB_CUDA(cudaMemcpyAsync(occ + r0, d_occ + r0, m, cudaMemcpyDeviceToHost, s_out));
            else {
                void *dst[5] = {out->inst + r0, out->prim + r0, out->u + r0, out->v + r0, out->t + r0};
                const void *srcp[5] = {Hc.inst, Hc.prim, Hc.u, Hc.v, Hc.t};
                for (int k = 0; k < 5; ++k) VB_CUDA(cudaMemcpyAsync(dst[k], srcp[k], m * 4, cudaMemcpyDeviceToHost, s_out));
            }
        }
    }
    ctx->cur_launches = launches;
    ctx->prof_end();
    VB_CHECK_LAUNCH();

    if (count) {
        unsigned long long hc[2];
        VB_CUDA(cudaMemcpyAsync(hc, d_cnt, 16, cudaMemcpyDeviceToHost, st));
        VB_CUDA(cudaStreamSynchronize(st));
        if (cnt_nodes) *cnt_nodes = hc[0];
        if (cnt_tris) *cnt_tris = hc[1];
    }
    if (out_host || rays_host) {          // host buffers are valid (and reusable) on return
        if (C > 1) { VB_CUDA(cudaStreamSynchronize(s_in)); VB_CUDA(cudaStreamSynchronize(s_out)); }
        VB_CUDA(cudaStreamSynchronize(st));
    }
    return VB_OK;
}

// ---- record form: the reference's own data shapes (Ray = {org_min, dir_max} 32 B in, TriangleHit = {inst, prim, bary} 16 B out) ----
// Host buffers move as ONE copy per range each way (the SoA form needs 8 + 5) and 16 instead of 20 bytes come back per closest
// hit; the AoS rays are traced in place (no SoA -> AoS pass). With option "trace.async" = 1 a host-buffer call returns once its
// copies and kernels are queued: consecutive calls then overlap (call k+1's H2D runs under call k's traversal and D2H) over two
// alternating staging sets; results are valid after vb_ctx_synchronize.
__global__ void __launch_bounds__(256) k_ray_keys_aos(const float4 *__restrict__ aos, uint32_t n, float3 lo, float3 inv_ext,
                                                      uint32_t *keys, uint32_t *vals, uint32_t oct_shift) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = __ldg(aos + 2 * (size_t)i), b = __ldg(aos + 2 * (size_t)i + 1);
    keys[i] = vb_ray_sort_key(a.x, a.y, a.z, b.x, b.y, b.z, lo, inv_ext, oct_shift);
    vals[i] = i;
}

int run_trace_records(vb_ctx *ctx, const vb_ray *rays, uint64_t n, vb_hit *hits, uint8_t *occ, vb_mem mem, bool any) {
    if (!ctx) return VB_ERR_INVALID;
    const char *fn = any ? "vb_trace_any_records" : "vb_trace_closest_records";
    if (!ctx->accel.built) return ctx->fail(VB_ERR_STATE, std::string(fn) + ": accel not built");
    if (!rays || (!any && !hits) || (any && !occ)) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": null argument");
    if (n == 0) return VB_OK;
    if (n >= (1ull << 32)) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": too many rays in one call");
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    const bool host = mem == VB_MEM_HOST;
    int C = 1;
    if (host && ctx->opt_host_chunks > 1 && n >= (uint64_t)ctx->opt_host_chunk_min_rays * 2)
        C = (int)std::min<uint64_t>((uint64_t)std::min(ctx->opt_host_chunks, VB_MAX_HOST_CHUNKS), n / (uint64_t)ctx->opt_host_chunk_min_rays);
    // asynchronous calls overlap ACROSS calls (copy-in of call k + 1 under the traversal of call k), so fewer, larger ranges win:
    // they sort and trace more efficiently (measured, tools/sweep_e2e.sh: 1 / 2 / 3 / 4 / 6 ranges = 1542 / 1629 / 1617 / 1550 / 1442 Mrays/s)
    if (ctx->opt_trace_async) C = std::min(C, std::max(1, ctx->opt_async_chunks));
    const uint64_t per = C > 1 ? (((n + C - 1) / C + 1023) & ~1023ull) : n;
    if (host) { int rc = ctx->ensure_copy_streams(); if (rc) return rc; }
    cudaStream_t s_in = host ? ctx->copy_in : st, s_out = host ? ctx->copy_out : st;

    // staging set of this call (host buffers only): [rays 32 n][hits 16 n | occlusion n]; two sets alternate between calls
    const int set = host ? (ctx->rec_call++ & 1) : 0;
    const float4 *d_rays = (const float4 *)rays;
    uint4 *d_hits = (uint4 *)hits; uint8_t *d_occ = occ;
    if (host) {
        size_t need = 32 * n + (any ? ((n + 255) & ~255ull) : 16 * n) + 256;
        if (need > ctx->rec_bytes[set]) {
            VB_CUDA(cudaDeviceSynchronize());                         // rare (first call / larger batch): nothing may still use the old set
            if (ctx->rec_stage[set]) cudaFree(ctx->rec_stage[set]);
            ctx->rec_stage[set] = nullptr; ctx->rec_bytes[set] = 0;
            VB_CUDA(cudaMalloc(&ctx->rec_stage[set], need));
            ctx->rec_bytes[set] = need;
        }
        // The set was last used two calls ago. Its RAY half is free once that call's traversal has run (rec_traced, on the context's
        // stream): this call's host-to-device copies wait for that only. Its RESULT half is free once that call's device-to-host copies
        // have drained (rec_done, on the copy-out stream): this call's kernels wait for that. With one event for both, copy-in of call
        // k + 2 could not start before copy-out of call k had finished and the three engines never ran at the same time.
        if (!ctx->rec_done[set]) {
            VB_CUDA(cudaEventCreateWithFlags(&ctx->rec_done[set], cudaEventDisableTiming));
            VB_CUDA(cudaEventCreateWithFlags(&ctx->rec_traced[set], cudaEventDisableTiming));
        } else {
            VB_CUDA(cudaStreamWaitEvent(s_in, ctx->rec_traced[set], 0));
            VB_CUDA(cudaStreamWaitEvent(st, ctx->rec_done[set], 0));
        }
        char *base = (char *)ctx->rec_stage[set];
        d_rays = (const float4 *)base;
        if (any) d_occ = (uint8_t *)(base + 32 * n); else d_hits = (uint4 *)(base + 32 * n);
        // synchronous calls: the copy-in stream starts after whatever the caller queued on the context's stream. Asynchronous calls
        // ("trace.async"): no such edge - the context's stream still holds the PREVIOUS call's traversal, and the point of the mode is
        // that this call's copy-in runs under it; host input buffers are read as of the call, like any host pointer handed to an
        // asynchronous copy
        if (!ctx->opt_trace_async) {
            VB_CUDA(cudaEventRecord(ctx->ev_in[0], st));
            VB_CUDA(cudaStreamWaitEvent(s_in, ctx->ev_in[0], 0));
        }
    }
    int rc = ctx->ensure_stage(256 + 8 * (size_t)VB_MAX_HOST_CHUNKS * 2);
    if (rc) return rc;
    // per-call ray counters: a ring of 2 x VB_MAX_HOST_CHUNKS slots so that an async call never resets its predecessor's
    unsigned long long *d_cnt = (unsigned long long *)ctx->stage;
    unsigned long long *d_next0 = d_cnt + 4 + (size_t)set * VB_MAX_HOST_CHUNKS;
    VB_CUDA(cudaMemsetAsync(d_next0, 0, 8 * (size_t)VB_MAX_HOST_CHUNKS, st));

    VbAccelView A{ctx->accel.nodes, ctx->accel.tris, ctx->accel.xf};
    uint64_t launches = 0;
    ctx->prof_begin(any ? "trace_any" : "trace_closest", 1);
    for (int c = 0; c < C; ++c) {
        const uint64_t r0 = (uint64_t)c * per;
        if (r0 >= n) break;
        const uint64_t m = std::min(per, n - r0);
        if (host) {
            VB_CUDA(cudaMemcpyAsync((void *)(d_rays + 2 * r0), rays + r0, m * sizeof(vb_ray), cudaMemcpyHostToDevice, s_in));
            VB_CUDA(cudaEventRecord(ctx->ev_in[c], s_in));
            VB_CUDA(cudaStreamWaitEvent(st, ctx->ev_in[c], 0));
        }
        const float4 *RS = d_rays + 2 * r0;
        unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count * ctx->opt_blocks_per_sm, (m + TRACE_BLOCK - 1) / TRACE_BLOCK);
        const uint32_t *perm = nullptr;
        launches += 1;
        if (ctx->opt_sort_rays && m >= (uint64_t)ctx->opt_sort_min_rays) {
            size_t cub_bytes = 0;
            cub::DoubleBuffer<uint32_t> dk0(nullptr, nullptr), dv0(nullptr, nullptr);
            VB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, dk0, dv0, (int)m, ctx->opt_sort_skip_bits, 24, st));
            size_t nb = ((m * 4 + 255) & ~255ull);
            size_t total = 4 * nb + cub_bytes + 256;
            if (total > ctx->sort_bytes) {
                VB_CUDA(cudaStreamSynchronize(st));
                if (ctx->sort_buf) cudaFree(ctx->sort_buf);
                ctx->sort_buf = nullptr; ctx->sort_bytes = 0;
                VB_CUDA(cudaMalloc(&ctx->sort_buf, total));
                ctx->sort_bytes = total;
            }
            char *sb = (char *)ctx->sort_buf;
            uint32_t *k0 = (uint32_t *)sb, *k1 = (uint32_t *)(sb + nb), *v0 = (uint32_t *)(sb + 2 * nb), *v1 = (uint32_t *)(sb + 3 * nb);
            void *cub_tmp = (void *)(((uintptr_t)(sb + 4 * nb) + 255) & ~(uintptr_t)255);
            float3 lo = make_float3(ctx->sort_lo[0], ctx->sort_lo[1], ctx->sort_lo[2]);
            float3 ie = make_float3(1.f / fmaxf(ctx->sort_hi[0] - ctx->sort_lo[0], 1e-30f), 1.f / fmaxf(ctx->sort_hi[1] - ctx->sort_lo[1], 1e-30f),
                                    1.f / fmaxf(ctx->sort_hi[2] - ctx->sort_lo[2], 1e-30f));
            k_ray_keys_aos<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(RS, (uint32_t)m, lo, ie, k0, v0, (uint32_t)ctx->opt_sort_oct_shift);
            cub::DoubleBuffer<uint32_t> dk(k0, k1), dv(v0, v1);
            cudaError_t ce = cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, dk, dv, (int)m, ctx->opt_sort_skip_bits, 24, st);
            if (ce != cudaSuccess) { ctx->prof_end(); return ctx->fail(VB_ERR_CUDA, std::string("ray sort: ") + cudaGetErrorString(ce)); }
            perm = dv.Current();
            launches += 5;
        }
        RaysDev R0{};
        HitsDev H0{};
        // perm == nullptr with RS set: k_trace reads ray i of the AoS array directly
        if (any) launch_trace<true, false>(grid, st, A, R0, m, H0, d_occ + r0, d_cnt, d_next0 + c, ctx->opt_refill_thr, ctx->opt_tri_thr, RS, perm, 1, nullptr);
        else launch_trace<false, false>(grid, st, A, R0, m, H0, nullptr, d_cnt, d_next0 + c, ctx->opt_refill_thr, ctx->opt_tri_thr, RS, perm, 1, d_hits + r0);
        if (host) {
            VB_CUDA(cudaEventRecord(ctx->ev_out[c], st));
            VB_CUDA(cudaStreamWaitEvent(s_out, ctx->ev_out[c], 0));
            if (any) VB_CUDA(cudaMemcpyAsync(occ + r0, d_occ + r0, m, cudaMemcpyDeviceToHost, s_out));
            else VB_CUDA(cudaMemcpyAsync(hits + r0, d_hits + r0, m * sizeof(vb_hit), cudaMemcpyDeviceToHost, s_out));
        }
    }
    ctx->cur_launches = launches;
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    if (host) {
        VB_CUDA(cudaEventRecord(ctx->rec_traced[set], st));
        VB_CUDA(cudaEventRecord(ctx->rec_done[set], s_out));
        ctx->rec_pending = true;
        if (!ctx->opt_trace_async) return vb_ctx_synchronize(ctx);       // host buffers are valid (and reusable) on return
    }
    return VB_OK;
}

}  // namespace

// ---- queue form used by the wavefront path tracer (vb_pt_wavefront_host.cuh) -----------------------
// rays: AoS records (o.xyz, tmin)(d.xyz, tmax); keyidx: (vb_ray_sort_key << 32 | queue index) per ray, written
// by the producer kernel, sorted here on its 24 key bits (keyidx_alt = same-size scratch); results go to
// hits4[index] = (inst, prim, u, v) or occ[index]. No staging, no host synchronisation. coherent = the queue is already in
// a traversal-friendly order (camera rays in pixel order: one origin, smoothly varying direction): no sort.
int vb_trace_queue(vb_ctx *ctx, const float4 *rays, uint64_t *keyidx, uint64_t *keyidx_alt, uint32_t n, uint4 *hits4, uint8_t *occ,
                   unsigned long long *work, bool coherent) {
    if (n == 0) return VB_OK;
    const bool any = occ != nullptr;
    cudaStream_t st = ctx->stream;
    int rc = ctx->ensure_stage(128);
    if (rc) return rc;
    unsigned long long *d_cnt = (unsigned long long *)ctx->stage;
    VB_CUDA(cudaMemsetAsync(d_cnt, 0, 32, st));
    const uint32_t *perm = nullptr;
    uint64_t launches = 1;
    ctx->prof_begin(any ? "trace_any" : "trace_closest", 1);
    if (ctx->opt_sort_rays && !coherent && n >= (uint32_t)ctx->opt_sort_min_rays) {
        size_t cub_bytes = 0;
        cub::DoubleBuffer<uint64_t> dk(keyidx, keyidx_alt);
        VB_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, cub_bytes, dk, (int)n, 32 + ctx->opt_sort_skip_bits, 56, st));
        if (cub_bytes > ctx->sort_bytes) {
            if (ctx->sort_buf) cudaFree(ctx->sort_buf);
            ctx->sort_buf = nullptr; ctx->sort_bytes = 0;
            VB_CUDA(cudaMalloc(&ctx->sort_buf, cub_bytes + 256));
            ctx->sort_bytes = cub_bytes + 256;
        }
        cudaError_t ce = cub::DeviceRadixSort::SortKeys(ctx->sort_buf, cub_bytes, dk, (int)n, 32 + ctx->opt_sort_skip_bits, 56, st);
        if (ce != cudaSuccess) { ctx->prof_end(); return ctx->fail(VB_ERR_CUDA, std::string("ray sort: ") + cudaGetErrorString(ce)); }
        perm = (const uint32_t *)dk.Current();
        launches += 4;
    }
    VbAccelView A{ctx->accel.nodes, ctx->accel.tris, ctx->accel.xf};
    unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count * ctx->opt_blocks_per_sm, ((uint64_t)n + TRACE_BLOCK - 1) / TRACE_BLOCK);
    RaysDev R{};
    HitsDev H{};
    // work != null: instrumented run, [nodes popped, triangles tested] are ADDED to work[0..1] (never timed)
    if (work && any) launch_trace<true, true>(grid, st, A, R, n, H, occ, work, d_cnt + 2, ctx->opt_refill_thr, ctx->opt_tri_thr, rays, perm, 2, nullptr);
    else if (work) launch_trace<false, true>(grid, st, A, R, n, H, nullptr, work, d_cnt + 2, ctx->opt_refill_thr, ctx->opt_tri_thr, rays, perm, 2, hits4);
    else if (any) launch_trace<true, false>(grid, st, A, R, n, H, occ, d_cnt, d_cnt + 2, ctx->opt_refill_thr, ctx->opt_tri_thr, rays, perm, 2, nullptr);
    else launch_trace<false, false>(grid, st, A, R, n, H, nullptr, d_cnt, d_cnt + 2, ctx->opt_refill_thr, ctx->opt_tri_thr, rays, perm, 2, hits4);
    ctx->cur_launches = launches;
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    return VB_OK;
}

// ---- vb_trace_closest_static: closest-hit batch trace that allocates nothing (include/vision_b200.h) ----
// Everything run_trace takes from growable buffers (the persistent kernel's ray counter, and for the in-call ray sort the Morton keys,
// the permutation, the AoS copy of the rays and the radix sort's temporary storage) lives in ONE file-scope device array of fixed
// size. g_static_scratch names that array for DEVICE code; host code only ever passes the address cudaGetSymbolAddress returned for
// the device that is current (the host-side symbol is a stand-in, not a device pointer).
namespace {

constexpr uint64_t STATIC_MAX_RAYS = VB_TRACE_STATIC_MAX_RAYS;
constexpr size_t STATIC_CNT_BYTES = 256;                               // nodes, tris, pad, pad, next_ray: cleared every call
constexpr size_t STATIC_KEY_BYTES = STATIC_MAX_RAYS * 4;               // one of k0, k1, v0, v1
constexpr size_t STATIC_AOS_BYTES = STATIC_MAX_RAYS * 32;
constexpr size_t STATIC_CUB_BYTES = 1u << 20;                          // radix sort temporary storage (checked against the query)
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_CNT_BYTES + 4 * STATIC_KEY_BYTES + STATIC_AOS_BYTES + STATIC_CUB_BYTES;
static_assert(STATIC_SCRATCH_BYTES == VB_TRACE_STATIC_SCRATCH_BYTES, "include/vision_b200.h states the size of the static scratch");

static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_static_mu;
unsigned char *g_static_ptr[STATIC_MAX_DEVICES] = {};                  // device address of g_static_scratch, per device ordinal

// once per device, with that device current: the array's address on THAT device
int static_scratch_address(vb_ctx *ctx, unsigned char **out) {
    if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_trace_closest_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[ctx->device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return ctx->fail(VB_ERR_CUDA, std::string("vb_trace_closest_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_static_ptr[ctx->device] = (unsigned char *)p;
    }
    *out = g_static_ptr[ctx->device];
    return VB_OK;
}

}  // namespace

extern "C" {

int vb_trace_closest(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, const vb_hits_soa *out) {
    return run_trace(ctx, rays, n, out, nullptr, VB_MEM_DEVICE, false, nullptr, nullptr);
}
int vb_trace_any(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, uint8_t *occluded, vb_mem out_mem) {
    return run_trace(ctx, rays, n, nullptr, occluded, out_mem, true, nullptr, nullptr);
}
int vb_trace_closest_records(vb_ctx *ctx, const vb_ray *rays, uint64_t n, vb_hit *hits, vb_mem mem) {
    return run_trace_records(ctx, rays, n, hits, nullptr, mem, false);
}
int vb_trace_any_records(vb_ctx *ctx, const vb_ray *rays, uint64_t n, uint8_t *occluded, vb_mem mem) {
    return run_trace_records(ctx, rays, n, nullptr, occluded, mem, true);
}
int vb_trace_closest_static(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, const vb_hits_soa *out) {
    if (!ctx) return VB_ERR_INVALID;
    const char *fn = "vb_trace_closest_static";
    if (!ctx->accel.built) return ctx->fail(VB_ERR_STATE, std::string(fn) + ": accel not built");
    if (!rays || !out) return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": null argument");
    if (rays->mem != VB_MEM_DEVICE || out->mem != VB_MEM_DEVICE)
        return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": rays and hits must be device buffers (host buffers need staging memory: use vb_trace_closest)");
    if (n > STATIC_MAX_RAYS)
        return ctx->fail(VB_ERR_INVALID, std::string(fn) + ": " + std::to_string(n) + " rays exceed the static scratch's maximum of " +
                                             std::to_string(STATIC_MAX_RAYS) + " (VB_TRACE_STATIC_MAX_RAYS) per call; split the batch or use vb_trace_closest");
    if (n == 0) return VB_OK;
    cudaSetDevice(ctx->device);
    unsigned char *sb = nullptr;
    int rc = static_scratch_address(ctx, &sb);
    if (rc) return rc;
    cudaStream_t st = ctx->stream;
    unsigned long long *d_cnt = (unsigned long long *)sb;
    VB_CUDA(cudaMemsetAsync(d_cnt, 0, STATIC_CNT_BYTES, st));          // the array keeps the last call's contents: the ray counter must start at 0
    RaysDev R{rays->ox, rays->oy, rays->oz, rays->tmin, rays->dx, rays->dy, rays->dz, rays->tmax};
    HitsDev H{out->inst, out->prim, out->u, out->v, out->t};
    VbAccelView A{ctx->accel.nodes, ctx->accel.tris, ctx->accel.xf};
    unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count * ctx->opt_blocks_per_sm, (n + TRACE_BLOCK - 1) / TRACE_BLOCK);
    const float4 *RS = nullptr; const uint32_t *perm = nullptr;
    if (ctx->opt_sort_rays && n >= (uint64_t)ctx->opt_sort_min_rays) {      // same in-call ray sort as run_trace, over the static scratch
        uint32_t *k0 = (uint32_t *)(sb + STATIC_CNT_BYTES), *k1 = (uint32_t *)(sb + STATIC_CNT_BYTES + STATIC_KEY_BYTES);
        uint32_t *v0 = (uint32_t *)(sb + STATIC_CNT_BYTES + 2 * STATIC_KEY_BYTES), *v1 = (uint32_t *)(sb + STATIC_CNT_BYTES + 3 * STATIC_KEY_BYTES);
        float4 *d_aos = (float4 *)(sb + STATIC_CNT_BYTES + 4 * STATIC_KEY_BYTES);
        void *cub_tmp = sb + STATIC_CNT_BYTES + 4 * STATIC_KEY_BYTES + STATIC_AOS_BYTES;
        size_t cub_bytes = 0;
        cub::DoubleBuffer<uint32_t> dk0(nullptr, nullptr), dv0(nullptr, nullptr);
        VB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, dk0, dv0, (int)n, 0, 24, st));      // size query only
        if (cub_bytes > STATIC_CUB_BYTES) return ctx->fail(VB_ERR_UNSUPPORTED, std::string(fn) + ": radix sort temporary storage exceeds the static reserve");
        float3 lo = make_float3(ctx->sort_lo[0], ctx->sort_lo[1], ctx->sort_lo[2]);
        float3 ie = make_float3(1.f / fmaxf(ctx->sort_hi[0] - ctx->sort_lo[0], 1e-30f), 1.f / fmaxf(ctx->sort_hi[1] - ctx->sort_lo[1], 1e-30f),
                                1.f / fmaxf(ctx->sort_hi[2] - ctx->sort_lo[2], 1e-30f));
        k_ray_keys<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(R, (uint32_t)n, lo, ie, k0, v0, d_aos, (uint32_t)ctx->opt_sort_oct_shift);
        cub::DoubleBuffer<uint32_t> dk(k0, k1), dv(v0, v1);
        cudaError_t ce = cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, dk, dv, (int)n, 0, 24, st);
        if (ce != cudaSuccess) return ctx->fail(VB_ERR_CUDA, std::string(fn) + ": ray sort: " + cudaGetErrorString(ce));
        perm = dv.Current();
        RS = d_aos;
    }
    launch_trace<false, false>(grid, st, A, R, n, H, nullptr, d_cnt, d_cnt + 4, ctx->opt_refill_thr, ctx->opt_tri_thr, RS, perm, 1, nullptr);
    VB_CHECK_LAUNCH();
    return VB_OK;
}
int vb_trace_closest_counted(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, const vb_hits_soa *out,
                             uint64_t *nodes_popped, uint64_t *tris_tested) {
    uint64_t a = 0, b = 0;
    int rc = run_trace(ctx, rays, n, out, nullptr, VB_MEM_DEVICE, false, &a, &b);
    if (nodes_popped) *nodes_popped = a;
    if (tris_tested) *tris_tested = b;
    return rc;
}
int vb_trace_any_counted(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, uint8_t *occluded, vb_mem out_mem,
                         uint64_t *nodes_popped, uint64_t *tris_tested) {
    uint64_t a = 0, b = 0;
    int rc = run_trace(ctx, rays, n, nullptr, occluded, out_mem, true, &a, &b);
    if (nodes_popped) *nodes_popped = a;
    if (tris_tested) *tris_tested = b;
    return rc;
}

}  // extern "C"

// Batch trace kernels + C ABI entry points: vb_trace_closest / vb_trace_any (+ counted variants).
// Replace ocarina::Accel::trace_closest / trace_occlusion (reference src/base/mgr/geometry.cpp:168-174)
// as a bulk SoA-in / SoA-out API (the microbench and parity surface of SURVEY.md §8b).
#include "vb_internal.h"
#include "vb_trace.cuh"

namespace {

constexpr int TRACE_BLOCK = 128;

struct RaysDev { const float *ox, *oy, *oz, *tmin, *dx, *dy, *dz, *tmax; };
struct HitsDev { uint32_t *inst, *prim; float *u, *v, *t; };

template <bool ANY, bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK) k_trace(VbAccelView A, RaysDev R, uint64_t n, HitsDev H, uint8_t *occ,
                                                       unsigned long long *counters) {
    __shared__ uint2 s_stack[VB_SM_STACK * TRACE_BLOCK];
    uint64_t i = (uint64_t)blockIdx.x * TRACE_BLOCK + threadIdx.x;
    uint32_t nn = 0, nt = 0;
    if (i < n) {
        VbRay r;
        r.ox = __ldg(R.ox + i); r.oy = __ldg(R.oy + i); r.oz = __ldg(R.oz + i); r.tmin = __ldg(R.tmin + i);
        r.dx = __ldg(R.dx + i); r.dy = __ldg(R.dy + i); r.dz = __ldg(R.dz + i); r.tmax = __ldg(R.tmax + i);
        VbHit h;
        bool got = vb_traverse<ANY, COUNT>(A, r, h, s_stack + threadIdx.x, TRACE_BLOCK, nn, nt);
        if (ANY) occ[i] = got ? 1 : 0;
        else { H.inst[i] = h.inst; H.prim[i] = h.prim; H.u[i] = h.u; H.v[i] = h.v; H.t[i] = h.t; }
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(~0u, nn, o); nt += __shfl_xor_sync(~0u, nt, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(&counters[0], nn); atomicAdd(&counters[1], nt); }
    }
}

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

    // staging layout: [8 ray arrays][hit arrays or occlusion bytes][2 counters]
    size_t ray_bytes = rays_host ? 8 * n * sizeof(float) : 0;
    size_t out_bytes = out_host ? (any ? ((n + 15) & ~15ull) : 5 * n * 4) : 0;
    size_t need = ray_bytes + out_bytes + 64;
    int rc = ctx->ensure_stage(need);
    if (rc) return rc;
    char *base = (char *)ctx->stage;
    RaysDev R{rays->ox, rays->oy, rays->oz, rays->tmin, rays->dx, rays->dy, rays->dz, rays->tmax};
    if (rays_host) {
        float *d = (float *)base;
        const float *src[8] = {rays->ox, rays->oy, rays->oz, rays->tmin, rays->dx, rays->dy, rays->dz, rays->tmax};
        for (int k = 0; k < 8; ++k)
            VB_CUDA(cudaMemcpyAsync(d + k * n, src[k], n * sizeof(float), cudaMemcpyHostToDevice, st));
        R = RaysDev{d, d + n, d + 2 * n, d + 3 * n, d + 4 * n, d + 5 * n, d + 6 * n, d + 7 * n};
    }
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
    if (count) VB_CUDA(cudaMemsetAsync(d_cnt, 0, 16, st));

    VbAccelView A{ctx->accel.nodes, ctx->accel.tris};
    unsigned grid = (unsigned)((n + TRACE_BLOCK - 1) / TRACE_BLOCK);
    ctx->prof_begin(any ? "trace_any" : "trace_closest", 1);
    if (any) {
        if (count) k_trace<true, true><<<grid, TRACE_BLOCK, 0, st>>>(A, R, n, H, d_occ, d_cnt);
        else k_trace<true, false><<<grid, TRACE_BLOCK, 0, st>>>(A, R, n, H, d_occ, d_cnt);
    } else {
        if (count) k_trace<false, true><<<grid, TRACE_BLOCK, 0, st>>>(A, R, n, H, d_occ, d_cnt);
        else k_trace<false, false><<<grid, TRACE_BLOCK, 0, st>>>(A, R, n, H, d_occ, d_cnt);
    }
    ctx->prof_end();
    VB_CHECK_LAUNCH();

    if (out_host) {
        if (any) VB_CUDA(cudaMemcpyAsync(occ, d_occ, n, cudaMemcpyDeviceToHost, st));
        else {
            void *dst[5] = {out->inst, out->prim, out->u, out->v, out->t};
            const void *srcp[5] = {H.inst, H.prim, H.u, H.v, H.t};
            for (int k = 0; k < 5; ++k) VB_CUDA(cudaMemcpyAsync(dst[k], srcp[k], n * 4, cudaMemcpyDeviceToHost, st));
        }
    }
    if (count) {
        unsigned long long hc[2];
        VB_CUDA(cudaMemcpyAsync(hc, d_cnt, 16, cudaMemcpyDeviceToHost, st));
        VB_CUDA(cudaStreamSynchronize(st));
        if (cnt_nodes) *cnt_nodes = hc[0];
        if (cnt_tris) *cnt_tris = hc[1];
    } else if (out_host || rays_host) {
        VB_CUDA(cudaStreamSynchronize(st));   // host buffers are valid on return
    }
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

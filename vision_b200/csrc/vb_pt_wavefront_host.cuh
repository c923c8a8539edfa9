// Host driver of the wavefront schedule (kernels: vb_pt_wavefront.cuh). Included by vb_pt_kernel.cuh
// after struct vb_pt_state.
#pragma once
#include <cub/device/device_radix_sort.cuh>

struct vb_wf_buffers {
    size_t capacity = 0;            // paths
    char *mem = nullptr;
    float4 *rays[3]{};              // closest queue (double-buffered), shadow queue
    uint32_t *slot_q[2]{};
    uint64_t *keys[5]{};            // [0,1] / [2,3]: sort record + scratch of the two closest queues; [4]: shadow queue
    uint4 *hits4 = nullptr;
    WfState st{};
    float4 *sh_pend = nullptr;
    uint8_t *occ = nullptr;
    WfCounters *d_cnt = nullptr, *h_cnt = nullptr;
    void *cub_tmp = nullptr; size_t cub_bytes = 0;
    unsigned long long *d_work = nullptr;    // [closest nodes, closest tris, any nodes, any tris] of an instrumented call
};

static void wf_free(vb_wf_buffers &w) {
    cudaFree(w.mem); w.mem = nullptr; w.capacity = 0;
    cudaFree(w.d_cnt); w.d_cnt = nullptr;
    cudaFree(w.d_work); w.d_work = nullptr;
    cudaFree(w.cub_tmp); w.cub_tmp = nullptr; w.cub_bytes = 0;
    if (w.h_cnt) cudaFreeHost(w.h_cnt);
    w.h_cnt = nullptr;
}

static int wf_ensure(vb_ctx *ctx, vb_wf_buffers &w, size_t P) {
    if (!w.d_cnt) {
        VB_CUDA(cudaMalloc(&w.d_cnt, sizeof(WfCounters)));
        VB_CUDA(cudaMallocHost(&w.h_cnt, sizeof(WfCounters)));
        VB_CUDA(cudaMalloc(&w.d_work, 4 * sizeof(unsigned long long)));
    }
    if (P <= w.capacity) return VB_OK;
    cudaFree(w.mem); w.mem = nullptr; w.capacity = 0;
    size_t Pa = (P + 63) & ~(size_t)63;              // every stream starts 256-byte aligned
    // per path: state 48 B, pending 16 B, hit 16 B, 3 ray queues x 32 B, 5 sort records x 8 B, 2 slots x 4 B, occlusion 1 B
    size_t bytes = Pa * (48 + 16 + 16 + 3 * 32 + 5 * 8 + 2 * 4 + 1);
    VB_CUDA(cudaMalloc(&w.mem, bytes));
    char *p = w.mem;
    auto take = [&](size_t elem) { char *r = p; p += Pa * elem; return r; };
    w.st.a = (float4 *)take(16); w.st.b = (float4 *)take(16); w.st.c = (uint4 *)take(16);
    w.sh_pend = (float4 *)take(16);
    w.hits4 = (uint4 *)take(16);
    for (auto &r : w.rays) r = (float4 *)take(32);
    for (auto &k : w.keys) k = (uint64_t *)take(8);
    w.slot_q[0] = (uint32_t *)take(4); w.slot_q[1] = (uint32_t *)take(4);
    w.occ = (uint8_t *)take(1);
    w.capacity = P;
    return VB_OK;
}

// order the n shading entries by material: records in k0 (scratch k1); returns the sorted buffer
static int wf_sort_by_material(vb_ctx *ctx, vb_wf_buffers &w, uint64_t *k0, uint64_t *k1, uint32_t n, const uint64_t **order) {
    cub::DoubleBuffer<uint64_t> dk(k0, k1);
    size_t need = 0;
    VB_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, need, dk, (int)n, 32, 40, ctx->stream));
    if (need > w.cub_bytes) {
        cudaFree(w.cub_tmp); w.cub_tmp = nullptr; w.cub_bytes = 0;
        VB_CUDA(cudaMalloc(&w.cub_tmp, need + 256));
        w.cub_bytes = need + 256;
    }
    VB_CUDA(cub::DeviceRadixSort::SortKeys(w.cub_tmp, need, dk, (int)n, 32, 40, ctx->stream));
    *order = dk.Current();
    return VB_OK;
}

// renders frames first + (fi0 .. fi0+nf-1) * stride of the call as one batch of nf * n_pix paths
static int wf_render_batch(vb_ctx *ctx, vb_pt_state *p, vb_wf_buffers &w, const PtScene &S, uint32_t first, uint32_t stride, uint32_t fi0,
                           uint32_t nf, int mode, PtCounters &total) {
    cudaStream_t st = ctx->stream;
    const uint32_t n_pix = p->w * p->h;
    const uint32_t P = n_pix * nf;
    WfSortBox B;
    B.lo = make_float3(ctx->sort_lo[0], ctx->sort_lo[1], ctx->sort_lo[2]);
    B.inv_ext = make_float3(1.f / fmaxf(ctx->sort_hi[0] - ctx->sort_lo[0], 1e-30f), 1.f / fmaxf(ctx->sort_hi[1] - ctx->sort_lo[1], 1e-30f),
                            1.f / fmaxf(ctx->sort_hi[2] - ctx->sort_lo[2], 1e-30f));
    B.oct_shift = (uint32_t)ctx->opt_sort_oct_shift;
    ctx->prof_begin("pt_generate", 1);
    k_wf_generate<<<(P + 255) / 256, 256, 0, st>>>(S, first, stride, fi0, P, WfQueue{w.rays[0], w.keys[0], w.slot_q[0]}, B, w.st);
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    uint32_t n = S.params.max_depth ? P : 0;
    int cur = 0;
    // The loop runs until every path has finished, like the megakernel's per-path loop: crossings of material-less medium boundaries
    // do not count as bounces, so max_depth does not bound the iterations. The cap only turns a queue that never drains (a bug) into
    // an error instead of a hang; paths are never dropped.
    const uint32_t max_iter = 1u << 16;
    for (uint32_t iter = 0; n; ++iter) {
        if (iter >= max_iter) return ctx->fail(VB_ERR_STATE, "vb_render_frames: the wavefront queue did not drain after 65536 iterations");
        uint64_t *k0 = w.keys[2 * cur], *k1 = w.keys[2 * cur + 1];
        WfQueue Qin{w.rays[cur], k0, w.slot_q[cur]};
        WfQueue Qout{w.rays[cur ^ 1], w.keys[2 * (cur ^ 1)], w.slot_q[cur ^ 1]};
        WfQueue Qsh{w.rays[2], w.keys[4], nullptr};
        unsigned long long *work = ctx->opt_pt_count_work ? w.d_work : nullptr;
        int rc = vb_trace_queue(ctx, Qin.rays, k0, k1, n, w.hits4, nullptr, work, iter == 0);      // camera rays: pixel order is coherent
        if (rc) return rc;
        total.closest += n;
        VB_CUDA(cudaMemsetAsync(w.d_cnt, 0, sizeof(WfCounters), st));
        const uint64_t *order = nullptr;
        ctx->prof_begin("pt_shade", 1);
        if (ctx->opt_pt_sort_material && p->n_mat_types > 1 && n >= 4096) {
            k_wf_material_keys<<<(n + 255) / 256, 256, 0, st>>>(S, n, w.hits4, k0);
            rc = wf_sort_by_material(ctx, w, k0, k1, n, &order);
            if (rc) { ctx->prof_end(); return rc; }
        }
        {
            const bool hero = S.params.spectrum == VB_SPECTRUM_HERO;
            const unsigned grid = (n + 127) / 128;
            auto shade = [&](int variant, int filter) {
#define VB_SHADE(V) k_wf_shade<V><<<grid, 128, 0, st>>>(S, n, Qin, order, w.hits4, w.st, Qout, Qsh, w.sh_pend, B, w.d_cnt, filter)
                switch (variant) {
                    case PT_LEAN: VB_SHADE(PT_LEAN); break;
                    case PT_PRINCIPLED: VB_SHADE(PT_PRINCIPLED); break;
                    case PT_HERO: VB_SHADE(PT_HERO); break;
                    case PT_HERO | PT_PRINCIPLED: VB_SHADE(PT_HERO | PT_PRINCIPLED); break;
                    case PT_PRINCIPLED | PT_LOBESET: VB_SHADE(PT_PRINCIPLED | PT_LOBESET); break;
                    default: VB_SHADE(PT_HERO | PT_PRINCIPLED | PT_LOBESET); break;
                }
#undef VB_SHADE
            };
            const int base = hero ? PT_HERO : PT_LEAN;
            const int heavy = base | PT_PRINCIPLED | (p->has_lobeset ? PT_LOBESET : 0);
            if (!p->has_principled) shade(base, 0);
            else if (order && ctx->opt_pt_split_principled) { shade(base, 1); shade(heavy, 2); }   // material-ordered queue: two launches
            else shade(heavy, 0);
        }
        ctx->prof_end();
        VB_CHECK_LAUNCH();
        VB_CUDA(cudaMemcpyAsync(w.h_cnt, w.d_cnt, sizeof(WfCounters), cudaMemcpyDeviceToHost, st));
        VB_CUDA(cudaStreamSynchronize(st));
        const WfCounters c = *w.h_cnt;
        total.vertices += c.vertices; total.culled += c.culled;
        if (c.n_shadow) {
            rc = vb_trace_queue(ctx, Qsh.rays, w.keys[4], k0, c.n_shadow, nullptr, w.occ, work ? work + 2 : nullptr);     // k0/k1 are free again
            if (rc) return rc;
            total.any += c.n_shadow;
            ctx->prof_begin("pt_shadow_resolve", 1);
            k_wf_shadow_resolve<<<(c.n_shadow + 255) / 256, 256, 0, st>>>(c.n_shadow, w.occ, w.sh_pend, w.st.b);
            ctx->prof_end();
            VB_CHECK_LAUNCH();
        }
        n = c.n_next;
        cur ^= 1;
    }
    ctx->prof_begin("pt_film", 1);
    k_wf_film<<<(n_pix + 255) / 256, 256, 0, st>>>(n_pix, nf, fi0, mode, p->frames_done, w.st.b, p->accum, p->rt);
    ctx->prof_end();
    VB_CHECK_LAUNCH();
    return VB_OK;
}

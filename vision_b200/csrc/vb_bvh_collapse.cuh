// Collapse of the binary radix tree into the compressed 8-wide BVH + host driver of the build.
// Included by vb_bvh_build.cu only.
#pragma once

namespace {

struct WorkItem { uint32_t bvh2, wide; };
struct Counters { uint32_t next_items, n_nodes, n_tris, pad; };

__device__ __forceinline__ float half_area(float4 lo, float4 hi) {
    float x = hi.x - lo.x, y = hi.y - lo.y, z = hi.z - lo.z;
    return x * y + y * z + z * x;
}

// smallest exponent byte e (float exponent bits) with extent <= 255 * 2^(e-127)
__device__ __forceinline__ uint32_t quant_exp(float extent) {
    float s = extent / 255.0f;
    uint32_t bits = __float_as_uint(s);
    uint32_t e = (bits >> 23) & 0xffu;
    if (bits & 0x7fffffu) e++;
    e = max(e, 1u);
    while (e < 254u && ceilf(extent / __uint_as_float(e << 23)) > 255.0f) e++;
    return min(e, 254u);
}


// ---- cost-optimal collapse (dynamic programme over the binary tree, after Ylitie, Karras & Laine 2017) ----
// A ray visits a wide node iff it hits the node's box in its parent, so the expected number of node visits per ray is
// proportional to the SUM OF THE WIDE NODES' BOX AREAS; triangles sit in their own leaf slots with their own boxes whatever
// the topology, so their cost is a constant. For every binary node n and every slot budget i = 1..7:
//   c(n, 1) = A(n) + D(n, 8)                       n becomes a wide node; its subtree is dealt over that node's 8 slots
//   c(n, i) = min(D(n, i), c(n, i - 1))            n is dissolved into its parent's wide node and may use up to i slots there
//   D(n, j) = min over 0 < k < j of c(left, k) + c(right, j - k),     c(triangle, .) = 0
// k_collapse_cost fills c bottom-up (second arrival at a node computes it, like k_refit); k_collapse re-derives the
// arg-min choices top-down. The greedy rule (open the largest child until 8) fills the bottom of the tree badly - a 9-triangle
// subtree becomes a 7-triangle node with a 2-triangle child - and that is where most nodes are.
#define VB_DP_W 7
__device__ __forceinline__ float dp_cost(const float *dp, uint32_t T, uint32_t x, int i) {          // c(x, i), i in 1..7
    return x >= T - 1 ? 0.f : __ldcg(dp + (size_t)x * VB_DP_W + (i - 1));
}
__device__ __forceinline__ float dp_distribute(const float *dp, uint32_t T, uint32_t L, uint32_t R, int j, int &best_k) {
    float best = FLT_MAX; best_k = 1;
    for (int k = 1; k < j; ++k) {
        float c = dp_cost(dp, T, L, min(k, VB_DP_W)) + dp_cost(dp, T, R, min(j - k, VB_DP_W));
        if (c < best) { best = c; best_k = k; }
    }
    return best;
}
__global__ void __launch_bounds__(256) k_collapse_cost(uint32_t T, BuildBufs b, float *dp) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= T || T == 1) return;
    uint32_t p = b.parent[T - 1 + g];
    while (p != 0xffffffffu) {
        __threadfence();
        if (atomicAdd(&b.flags[p], 1u) == 0) return;
        uint32_t L = b.left[p], R = b.right[p];
        float cl[VB_DP_W], cr[VB_DP_W];
#pragma unroll
        for (int i = 0; i < VB_DP_W; ++i) { cl[i] = dp_cost(dp, T, L, i + 1); cr[i] = dp_cost(dp, T, R, i + 1); }
        float c[VB_DP_W + 1];                        // c[j - 1] for j = 1..8, c[7] = D(p, 8)
#pragma unroll
        for (int j = 2; j <= 8; ++j) {
            float best = FLT_MAX;
#pragma unroll
            for (int k = 1; k < j; ++k) if (k <= VB_DP_W && j - k <= VB_DP_W) best = fminf(best, cl[k - 1] + cr[j - k - 1]);
            c[j - 1] = best;
        }
        c[0] = half_area(b.nlo[p], b.nhi[p]) + c[7];
#pragma unroll
        for (int i = 1; i < VB_DP_W; ++i) c[i] = fminf(c[i], c[i - 1]);
#pragma unroll
        for (int i = 0; i < VB_DP_W; ++i) dp[(size_t)p * VB_DP_W + i] = c[i];
        p = b.parent[p];
    }
}
// children of the wide node rooted at binary node `root` under the cost table: at most 8, in c[]
__device__ __forceinline__ int dp_children(const float *dp, uint32_t T, const BuildBufs &b, uint32_t root, uint32_t *c) {
    uint32_t st_node[8]; int st_j[8]; int sp = 0, n = 0, k;
    {
        uint32_t L = b.left[root], R = b.right[root];
        dp_distribute(dp, T, L, R, 8, k);
        st_node[sp] = R; st_j[sp++] = 8 - k; st_node[sp] = L; st_j[sp++] = k;
    }
    while (sp) {
        uint32_t x = st_node[--sp]; int j = min(st_j[sp], VB_DP_W);
        for (;;) {
            if (x >= T - 1 || j == 1) { c[n++] = x; break; }
            uint32_t L = b.left[x], R = b.right[x];
            float d = dp_distribute(dp, T, L, R, j, k);
            if (d < dp_cost(dp, T, x, j - 1)) {            // dissolve x: its children share the j slots
                st_node[sp] = R; st_j[sp++] = j - k;
                x = L; j = k;
            } else j--;
        }
    }
    return n;
}

__global__ void __launch_bounds__(128) k_collapse(const WorkItem *in, uint32_t n_in, WorkItem *out, Counters *cnt,
                                                  uint32_t T, BuildBufs b, vb_wide_node *nodes, float4 *tris_out, uint32_t max_leaf, const float *dp) {
    uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_in) return;
    WorkItem it = in[w];
    auto tri_count = [&](uint32_t x) -> uint32_t { return x >= T - 1 ? 1u : b.last[x] - b.first[x] + 1u; };
    auto leaf_like = [&](uint32_t x) -> bool { return tri_count(x) <= max_leaf; };

    uint32_t c[8]; int n;
    if (leaf_like(it.bvh2)) { c[0] = it.bvh2; n = 1; }
    else if (dp) n = dp_children(dp, T, b, it.bvh2, c);
    else {
        c[0] = b.left[it.bvh2]; c[1] = b.right[it.bvh2]; n = 2;
        while (n < 8) {
            int best = -1; float best_a = -1.f;
            for (int j = 0; j < n; ++j) {
                if (leaf_like(c[j])) continue;
                float a = half_area(b.nlo[c[j]], b.nhi[c[j]]);
                if (a > best_a) { best_a = a; best = j; }
            }
            if (best < 0) break;
            uint32_t x = c[best];
            c[best] = b.left[x]; c[n++] = b.right[x];
        }
    }
    float4 nlo = b.nlo[it.bvh2], nhi = b.nhi[it.bvh2];
    float ncx = 0.5f * (nlo.x + nhi.x), ncy = 0.5f * (nlo.y + nhi.y), ncz = 0.5f * (nlo.z + nhi.z);

    // greedy child -> slot assignment: slot bit2/1/0 = child on the +x/+y/+z side
    float ccx[8], ccy[8], ccz[8];
    for (int j = 0; j < n; ++j) {
        float4 lo = b.nlo[c[j]], hi = b.nhi[c[j]];
        ccx[j] = 0.5f * (lo.x + hi.x) - ncx; ccy[j] = 0.5f * (lo.y + hi.y) - ncy; ccz[j] = 0.5f * (lo.z + hi.z) - ncz;
    }
    int slot_of[8]; uint32_t child_done = 0, slot_used = 0;
    for (int iter = 0; iter < n; ++iter) {
        int bj = -1, bs = -1; float bc = -FLT_MAX;
        for (int j = 0; j < n; ++j) {
            if (child_done >> j & 1) continue;
            for (int s = 0; s < 8; ++s) {
                if (slot_used >> s & 1) continue;
                float cost = ((s & 4) ? ccx[j] : -ccx[j]) + ((s & 2) ? ccy[j] : -ccy[j]) + ((s & 1) ? ccz[j] : -ccz[j]);
                if (cost > bc) { bc = cost; bj = j; bs = s; }
            }
        }
        slot_of[bj] = bs; child_done |= 1u << bj; slot_used |= 1u << bs;
    }
    int child_in_slot[8];
    for (int s = 0; s < 8; ++s) child_in_slot[s] = -1;
    for (int j = 0; j < n; ++j) child_in_slot[slot_of[j]] = j;

    uint32_t imask = 0, k_inner = 0, n_tri = 0;
    for (int s = 0; s < 8; ++s) {
        int j = child_in_slot[s];
        if (j < 0) continue;
        if (leaf_like(c[j])) n_tri += tri_count(c[j]); else { imask |= 1u << s; k_inner++; }
    }
    uint32_t child_base = k_inner ? atomicAdd(&cnt->n_nodes, k_inner) : 0u;
    uint32_t tri_base = n_tri ? atomicAdd(&cnt->n_tris, n_tri) : 0u;
    uint32_t out_base = k_inner ? atomicAdd(&cnt->next_items, k_inner) : 0u;

    vb_wide_node N;
    N.pad[0] = N.pad[1] = N.pad[2] = N.pad[3] = 0u;
    // quantisation frame = node box grown by 1/64 of its extent per side, so that the 2^-6 grid-unit
    // inflation of the child boxes below never clamps at 0/255 (see vb_qf in vb_trace.cuh)
    {
        float mx = (nhi.x - nlo.x) * 0.015625f, my = (nhi.y - nlo.y) * 0.015625f, mz = (nhi.z - nlo.z) * 0.015625f;
        nlo.x -= mx; nlo.y -= my; nlo.z -= mz; nhi.x += mx; nhi.y += my; nhi.z += mz;
    }
    N.px = nlo.x; N.py = nlo.y; N.pz = nlo.z;
    uint32_t ex = quant_exp(nhi.x - nlo.x), ey = quant_exp(nhi.y - nlo.y), ez = quant_exp(nhi.z - nlo.z);
    N.ex = (uint8_t)ex; N.ey = (uint8_t)ey; N.ez = (uint8_t)ez; N.imask = (uint8_t)imask;
    N.child_base = child_base; N.tri_base = tri_base;
    float isx = 1.0f / __uint_as_float(ex << 23), isy = 1.0f / __uint_as_float(ey << 23), isz = 1.0f / __uint_as_float(ez << 23);
    uint32_t tri_off = 0, inner_rel = 0, leaf8 = 0;
    for (int s = 0; s < 8; ++s) N.meta[s] = 0;
    for (int s = 0; s < 8; ++s) {
        int j = child_in_slot[s];
        if (j < 0) {
            N.qlox[s] = N.qloy[s] = N.qloz[s] = 255; N.qhix[s] = N.qhiy[s] = N.qhiz[s] = 0;   // empty: inverted box
            continue;
        }
        float4 lo = b.nlo[c[j]], hi = b.nhi[c[j]];
        N.qlox[s] = (uint8_t)fminf(fmaxf(floorf((lo.x - nlo.x) * isx - 0.015625f), 0.f), 255.f);
        N.qloy[s] = (uint8_t)fminf(fmaxf(floorf((lo.y - nlo.y) * isy - 0.015625f), 0.f), 255.f);
        N.qloz[s] = (uint8_t)fminf(fmaxf(floorf((lo.z - nlo.z) * isz - 0.015625f), 0.f), 255.f);
        N.qhix[s] = (uint8_t)fminf(fmaxf(ceilf((hi.x - nlo.x) * isx + 0.015625f), 0.f), 255.f);
        N.qhiy[s] = (uint8_t)fminf(fmaxf(ceilf((hi.y - nlo.y) * isy + 0.015625f), 0.f), 255.f);
        N.qhiz[s] = (uint8_t)fminf(fmaxf(ceilf((hi.z - nlo.z) * isz + 0.015625f), 0.f), 255.f);
        if (imask >> s & 1) {
            out[out_base + inner_rel] = WorkItem{c[j], child_base + inner_rel};
            inner_rel++;
        } else {
            uint32_t cntj = tri_count(c[j]);
            uint32_t first = c[j] >= T - 1 ? c[j] - (T - 1) : b.first[c[j]];
            leaf8 |= 1u << s;                        // one triangle per leaf slot, stored in slot order (vb_next_triangle)
            for (uint32_t t = 0; t < cntj; ++t) {
                size_t src = 3 * (size_t)__float_as_uint(b.blo[b.vals[first + t]].w), dst = 3 * (size_t)(tri_base + tri_off + t);   // reference -> its triangle
                tris_out[dst] = b.tri_in[src]; tris_out[dst + 1] = b.tri_in[src + 1]; tris_out[dst + 2] = b.tri_in[src + 2];
            }
            tri_off += cntj;
        }
    }
    N.meta[0] = (uint8_t)leaf8;
    const uint4 *src = reinterpret_cast<const uint4 *>(&N);
    uint4 *dst = reinterpret_cast<uint4 *>(nodes + it.wide);
#pragma unroll
    for (int q = 0; q < VB_NODE_U4; ++q) dst[q] = src[q];
}

template <class T> struct DevBuf {
    T *p = nullptr;
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)); }
    ~DevBuf() { if (p) cudaFree(p); }
};

}  // namespace

// One build. *too_deep = the wide tree has more levels than the traversal stack has entries (VB_TRAVERSAL_STACK_ENTRIES: one entry
// per level is the worst case of vb_traverse); nothing is installed in that case.
// What one build runs over: every instance flattened to world space (the default), ONE mesh in its object space (a bottom-level tree of the
// two-level structure: vertices kept in the leaf records, extra leaf padding), or given boxes (the top-level tree: one "triangle" per instance).
struct BuildInput {
    enum Mode { FLATTEN = 0, MESH = 1, TOP = 2 } mode = FLATTEN;
    uint32_t mesh_id = 0; float extra_pad = 0.f;                    // MESH
    const uint32_t *inst_boxes = nullptr, *inst_roots = nullptr;    // TOP: device arrays (6 ordered uints per instance; root node per instance)
};

static int accel_build_once(vb_ctx *ctx, vb_build_stats *out, bool allow_ploc, bool *too_deep, const BuildInput &in = BuildInput()) {
    *too_deep = false;
    cudaStream_t st = ctx->stream;
    uint32_t n_inst = in.mode == BuildInput::MESH ? 1u : (uint32_t)ctx->instances.size();
    if (n_inst == 0) return ctx->fail(VB_ERR_STATE, "vb_accel_build: no instances set");
    std::vector<InstGeom> hg(n_inst);
    std::vector<uint32_t> off(n_inst + 1, 0);
    uint64_t total = 0;
    for (uint32_t i = 0; i < n_inst; ++i) {
        const vb_device_mesh &m = ctx->meshes[in.mode == BuildInput::MESH ? in.mesh_id : ctx->instances[i].mesh_id];
        hg[i].pos = m.pos; hg[i].tri = m.tri; hg[i].soup = m.soup ? 1u : 0u; hg[i].pad = 0;
        if (in.mode == BuildInput::MESH) { for (int k = 0; k < 16; ++k) hg[i].o2w[k] = (k % 5 == 0) ? 1.f : 0.f; }
        else memcpy(hg[i].o2w, ctx->instances[i].o2w, 64);
        off[i] = (uint32_t)total;
        total += in.mode == BuildInput::TOP ? 1u : m.nt;
    }
    if (total == 0) return ctx->fail(VB_ERR_STATE, "vb_accel_build: scene has no triangles");
    if (total > 0x7fffffffull / 3) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_accel_build: more than 715M triangles");
    off[n_inst] = (uint32_t)total;
    uint32_t T = (uint32_t)total;

    cudaFree(ctx->accel.nodes); cudaFree(ctx->accel.tris); cudaFree(ctx->accel.xf);
    ctx->accel = vb_accel();

    DevBuf<InstGeom> d_geom; DevBuf<uint32_t> d_off;
    DevBuf<float4> tri_in, blo, bhi, rlo, rhi, nlo, nhi;
    DevBuf<uint32_t> scene, vals, vals2, left, right, first, last, parent, flags;
    DevBuf<uint64_t> keys, keys2;
    DevBuf<double> moments; DevBuf<uint32_t> cls_count;
    DevBuf<WorkItem> q0, q1; DevBuf<Counters> cnt;
    VB_CUDA(d_geom.alloc(n_inst)); VB_CUDA(d_off.alloc(n_inst + 1));
    VB_CUDA(tri_in.alloc(3 * (size_t)T)); VB_CUDA(blo.alloc(T)); VB_CUDA(bhi.alloc(T));
    VB_CUDA(scene.alloc(8)); VB_CUDA(moments.alloc(8)); VB_CUDA(cls_count.alloc(4)); VB_CUDA(cnt.alloc(1));
    vb_wide_node *d_nodes = nullptr; float4 *d_tris = nullptr;

    // the per-leaf buffers are sized after the pre-splitting pass (R references >= T triangles)
    BuildBufs b{tri_in.p, blo.p, bhi.p, scene.p, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, cls_count.p, moments.p,
                in.mode == BuildInput::MESH ? in.extra_pad : 0.f, in.mode == BuildInput::MESH ? 1u : 0u};

    cudaEvent_t e0 = ctx->get_event(), e1 = ctx->get_event();
    VB_CUDA(cudaMemcpyAsync(d_geom.p, hg.data(), sizeof(InstGeom) * n_inst, cudaMemcpyHostToDevice, st));
    VB_CUDA(cudaMemcpyAsync(d_off.p, off.data(), sizeof(uint32_t) * (n_inst + 1), cudaMemcpyHostToDevice, st));
    uint32_t scene_init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    VB_CUDA(cudaMemcpyAsync(scene.p, scene_init, sizeof(scene_init), cudaMemcpyHostToDevice, st));
    VB_CUDA(cudaMemsetAsync(cls_count.p, 0, 4 * sizeof(uint32_t), st));
    VB_CUDA(cudaEventRecord(e0, st));
    uint64_t launches = 0;
    uint32_t grid = (T + 255) / 256;
    VB_CUDA(cudaMemsetAsync(moments.p, 0, 8 * sizeof(double), st));
    if (in.mode == BuildInput::TOP) k_tlas_prims<<<grid, 256, 0, st>>>(T, in.inst_boxes, in.inst_roots, b);
    else k_flatten<<<grid, 256, 0, st>>>(d_geom.p, d_off.p, n_inst, T, b);
    launches++;
    VB_CHECK_LAUNCH();
    // ---- triangle pre-splitting (vb_bvh_split.cuh): T triangles -> R references ----
    const uint32_t T_tris = T;
    if (ctx->opt_split_pct > 0 && T >= 8 && in.mode == BuildInput::FLATTEN) {
        DevBuf<float> prio; DevBuf<uint32_t> s_count, s_offs;
        VB_CUDA(prio.alloc(T)); VB_CUDA(s_count.alloc(T)); VB_CUDA(s_offs.alloc(T));
        k_split_priority<<<grid, 256, 0, st>>>(T, b, prio.p, moments.p + 6); launches++;
        VB_CHECK_LAUNCH();
        double psum = 0.0;
        VB_CUDA(cudaMemcpyAsync(&psum, moments.p + 6, sizeof(double), cudaMemcpyDeviceToHost, st));
        VB_CUDA(cudaStreamSynchronize(st));
        if (psum > 0.0 && std::isfinite(psum)) {
            const double budget = (double)T * ctx->opt_split_pct / 100.0;
            k_split_count<<<grid, 256, 0, st>>>(T, prio.p, (float)(budget / psum), s_count.p); launches++;
            size_t tmp_bytes = 0;
            VB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, s_count.p, s_offs.p, (int)T, st));
            int rc = ctx->ensure_stage(tmp_bytes);
            if (rc) return rc;
            VB_CUDA(cub::DeviceScan::ExclusiveSum(ctx->stage, tmp_bytes, s_count.p, s_offs.p, (int)T, st)); launches += 2;
            uint32_t tail2[2];
            VB_CUDA(cudaMemcpyAsync(&tail2[0], s_offs.p + (T - 1), 4, cudaMemcpyDeviceToHost, st));
            VB_CUDA(cudaMemcpyAsync(&tail2[1], s_count.p + (T - 1), 4, cudaMemcpyDeviceToHost, st));
            VB_CUDA(cudaStreamSynchronize(st));
            const uint64_t R = (uint64_t)tail2[0] + tail2[1];
            if (R > T && R <= 0x7fffffffull / 3) {
                VB_CUDA(rlo.alloc(R)); VB_CUDA(rhi.alloc(R));
                k_split_emit<<<(T + 127) / 128, 128, 0, st>>>(T, b, s_count.p, s_offs.p, rlo.p, rhi.p); launches++;
                VB_CHECK_LAUNCH();
                VB_CUDA(cudaStreamSynchronize(st));            // prio / s_count / s_offs are released at the end of this block
                b.blo = rlo.p; b.bhi = rhi.p;
                T = (uint32_t)R;
            }
        }
    }
    grid = (T + 255) / 256;
    VB_CUDA(nlo.alloc(2 * (size_t)T)); VB_CUDA(nhi.alloc(2 * (size_t)T));
    VB_CUDA(vals.alloc(T)); VB_CUDA(vals2.alloc(T));
    VB_CUDA(left.alloc(T)); VB_CUDA(right.alloc(T)); VB_CUDA(first.alloc(T)); VB_CUDA(last.alloc(T));
    VB_CUDA(parent.alloc(2 * (size_t)T)); VB_CUDA(flags.alloc(T));
    VB_CUDA(keys.alloc(T)); VB_CUDA(keys2.alloc(T));
    VB_CUDA(q0.alloc(T)); VB_CUDA(q1.alloc(T));
    // worst case one wide node per 2 leaves is never reached; T+1 nodes is a safe bound
    VB_CUDA(cudaMalloc(&d_nodes, ((size_t)T + 1) * sizeof(vb_wide_node)));
    if (cudaMalloc(&d_tris, 3 * (size_t)T * sizeof(float4)) != cudaSuccess) { cudaFree(d_nodes); return ctx->fail(VB_ERR_CUDA, "vb_accel_build: out of device memory"); }
    b.keys = keys.p; b.keys2 = keys2.p; b.vals = vals.p; b.vals2 = vals2.p; b.nlo = nlo.p; b.nhi = nhi.p;
    b.left = left.p; b.right = right.p; b.first = first.p; b.last = last.p; b.parent = parent.p; b.flags = flags.p;
    k_morton<<<grid, 256, 0, st>>>(T, T_tris, b, ctx->opt_size_classes); launches++;
    VB_CHECK_LAUNCH();
    {
        size_t tmp_bytes = 0;
        cub::DoubleBuffer<uint64_t> dk(keys.p, keys2.p);
        cub::DoubleBuffer<uint32_t> dv(vals.p, vals2.p);
        VB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)T, 0, 63, st));
        int rc = ctx->ensure_stage(tmp_bytes);
        if (rc) return rc;
        VB_CUDA(cub::DeviceRadixSort::SortPairs(ctx->stage, tmp_bytes, dk, dv, (int)T, 0, 63, st));
        launches += 8;
        b.keys = dk.Current(); b.vals = dv.Current();
    }
    // builder choice ("build.ploc": 1 = PLOC, 0 = radix tree, -1 = automatic, default): area-driven clustering pays
    // when triangles of very different sizes share space (architecture: 30.9 -> 15.2 wide nodes per closest-hit
    // ray on classroom); on uniform tessellations the radix tree is as good or better (random 10M-triangle
    // soup: 38 vs 54 nodes per ray) and cheaper to build. Automatic = PLOC iff some triangle is larger than
    // 1/8 of the geometry's spread (size class >= 2).
    uint32_t h_cls[4] = {0, 0, 0, 0};
    VB_CUDA(cudaMemcpyAsync(h_cls, cls_count.p, sizeof(h_cls), cudaMemcpyDeviceToHost, st));
    VB_CUDA(cudaStreamSynchronize(st));
    const bool want_ploc = ctx->opt_ploc < 0 ? (h_cls[2] + h_cls[3]) > 0 : ctx->opt_ploc != 0;
    const bool ploc = want_ploc && allow_ploc && T > 2;
    ctx->last_build_ploc = ploc;
    const uint32_t max_leaf = 1u;        // the traversal's slot-order triangle addressing needs one triangle per leaf slot
    if (ploc) {
        DevBuf<uint32_t> cl0, cl1, nn, totals; DevBuf<uint64_t> pf, ps;
        VB_CUDA(cl0.alloc(T)); VB_CUDA(cl1.alloc(T)); VB_CUDA(nn.alloc(T)); VB_CUDA(totals.alloc(2));
        VB_CUDA(pf.alloc(T)); VB_CUDA(ps.alloc(T));
        int rc = ploc_build(ctx, T, b, cl0.p, cl1.p, nn.p, pf.p, ps.p, totals.p, max_leaf, launches);
        if (rc) return rc;
    } else {
        VB_CUDA(cudaMemsetAsync(flags.p, 0, sizeof(uint32_t) * T, st));
        if (T > 1) { k_karras<<<(T - 1 + 255) / 256, 256, 0, st>>>((int)T, b.keys, b); launches++; VB_CHECK_LAUNCH(); }
        k_refit<<<grid, 256, 0, st>>>(T, b); launches++;
        VB_CHECK_LAUNCH();
    }

    // cost table of the cost-optimal collapse ("build.collapse" = 1; 0 = greedy largest-area opening)
    DevBuf<float> dp;
    if (ctx->opt_collapse && T > 2) {
        VB_CUDA(dp.alloc((size_t)(T - 1) * VB_DP_W));
        VB_CUDA(cudaMemsetAsync(flags.p, 0, sizeof(uint32_t) * T, st));
        k_collapse_cost<<<grid, 256, 0, st>>>(T, b, dp.p); launches++;
        VB_CHECK_LAUNCH();
    }
    // level-synchronous collapse; root work item: BVH2 root = node 0 (or the single leaf when T==1)
    Counters hc{0, 1, 0, 0};
    WorkItem root{T == 1 ? 0u : 0u, 0u};
    VB_CUDA(cudaMemcpyAsync(q0.p, &root, sizeof(root), cudaMemcpyHostToDevice, st));
    VB_CUDA(cudaMemcpyAsync(cnt.p, &hc, sizeof(hc), cudaMemcpyHostToDevice, st));
    WorkItem *qin = q0.p, *qout = q1.p;
    uint32_t n_items = 1;
    int levels = 0;
    while (n_items) {
        k_collapse<<<(n_items + 127) / 128, 128, 0, st>>>(qin, n_items, qout, cnt.p, T, b, d_nodes, d_tris, max_leaf, dp.p); launches++;
        VB_CHECK_LAUNCH();
        VB_CUDA(cudaMemcpyAsync(&hc, cnt.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
        VB_CUDA(cudaStreamSynchronize(st));
        n_items = hc.next_items;
        uint32_t zero = 0;
        VB_CUDA(cudaMemcpyAsync(&cnt.p->next_items, &zero, 4, cudaMemcpyHostToDevice, st));
        std::swap(qin, qout);
        if (++levels > ctx->opt_max_levels && n_items) {          // a deeper tree would overflow the traversal stack: refuse it
            cudaFree(d_nodes); cudaFree(d_tris);
            ctx->ev_pool.push_back(e0); ctx->ev_pool.push_back(e1);
            *too_deep = true;
            return VB_OK;
        }
    }
    ctx->last_build_levels = levels;
    VB_CUDA(cudaEventRecord(e1, st));
    VB_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    ctx->ev_pool.push_back(e0); ctx->ev_pool.push_back(e1);
    if (hc.n_tris != T) return ctx->fail(VB_ERR_STATE, "vb_accel_build: triangle count mismatch after collapse");

    uint32_t hs[6];
    VB_CUDA(cudaMemcpy(hs, scene.p, sizeof(hs), cudaMemcpyDeviceToHost));
    // shrink node array to its final size
    uint4 *final_nodes = nullptr;
    VB_CUDA(cudaMalloc(&final_nodes, (size_t)hc.n_nodes * sizeof(vb_wide_node)));
    VB_CUDA(cudaMemcpy(final_nodes, d_nodes, (size_t)hc.n_nodes * sizeof(vb_wide_node), cudaMemcpyDeviceToDevice));
    cudaFree(d_nodes);
    ctx->accel.nodes = final_nodes; ctx->accel.tris = d_tris;
    ctx->accel.n_nodes = hc.n_nodes; ctx->accel.n_tris = T; ctx->accel.built = true;

    vb_build_stats &s = ctx->build_stats;
    s.n_triangles = T_tris; s.n_nodes = hc.n_nodes; s.node_bytes = (uint64_t)hc.n_nodes * sizeof(vb_wide_node); s.tri_bytes = (uint64_t)T * 48;
    s.n_references = T;
    s.build_ms = ms;
    for (int a = 0; a < 3; ++a) { s.scene_lo[a] = ord2f(hs[a]); s.scene_hi[a] = ord2f(hs[3 + a]); }
    // ray-sort box: mean +- 2.5 sigma of the triangle centroids, clipped to the scene box. Rays start on
    // surfaces, so the Morton cells of the sort key go where the geometry is even when a few huge
    // triangles (ground planes, sky domes) inflate the scene box.
    double hm[6];
    VB_CUDA(cudaMemcpy(hm, moments.p, sizeof(hm), cudaMemcpyDeviceToHost));
    for (int a = 0; a < 3; ++a) {
        double mean = hm[a] / T_tris, var = std::max(0.0, hm[3 + a] / T_tris - mean * mean), sd = std::sqrt(var);
        double lo = std::max((double)s.scene_lo[a], mean - 2.5 * sd), hi = std::min((double)s.scene_hi[a], mean + 2.5 * sd);
        if (!(hi > lo)) { lo = s.scene_lo[a]; hi = s.scene_hi[a]; }
        ctx->sort_lo[a] = (float)lo; ctx->sort_hi[a] = (float)hi;
    }
    ctx->prof["bvh_build"].ms += ms; ctx->prof["bvh_build"].launches += launches;
    ctx->prof["all"].ms += ms; ctx->prof["all"].launches += launches;
    if (out) *out = s;
    return VB_OK;
}

// Agglomerative clustering can produce chain-like trees (nested shells: every merge adds one triangle to the cluster inside it) whose
// wide form is deeper than the traversal stack; the radix tree's depth is bounded by its key length (63 Morton + 32 index bits, seven
// binary levels per wide level on a chain), so such a scene is rebuilt with it. A tree that is still too deep is refused.
// one tree with the PLOC -> radix fallback of the single-level build
static int accel_build_tree(vb_ctx *ctx, const BuildInput &in, const char *what) {
    bool too_deep = false;
    int rc = accel_build_once(ctx, nullptr, true, &too_deep, in);
    if (rc || !too_deep) return rc;
    if (ctx->last_build_ploc) {
        ctx->prof["bvh_build_radix_fallback"].launches += 1;
        rc = accel_build_once(ctx, nullptr, false, &too_deep, in);
        if (rc || !too_deep) return rc;
    }
    return ctx->fail(VB_ERR_UNSUPPORTED, std::string("vb_accel_build: ") + what + " is deeper than the traversal stack");
}

// inverse of an affine object-to-world matrix (column-major 4x4) by Gauss-Jordan elimination with row pivoting in double precision
static bool invert_affine(const float *m, double inv[4][4]) {
    double a[4][8];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { a[r][c] = m[c * 4 + r]; a[r][4 + c] = r == c ? 1.0 : 0.0; }
    for (int i = 0; i < 4; ++i) {
        int piv = i;
        for (int r = i + 1; r < 4; ++r) if (std::fabs(a[r][i]) > std::fabs(a[piv][i])) piv = r;
        if (!(std::fabs(a[piv][i]) > 0.0)) return false;
        if (piv != i) for (int c = 0; c < 8; ++c) std::swap(a[i][c], a[piv][c]);
        const double s = 1.0 / a[i][i];
        for (int c = 0; c < 8; ++c) a[i][c] *= s;
        for (int r = 0; r < 4; ++r) if (r != i) { const double f = a[r][i]; for (int c = 0; c < 8; ++c) a[r][c] -= f * a[i][c]; }
    }
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) inv[r][c] = a[r][4 + c];
    return true;
}

// Two-level structure ("build.two_level" = 1; the shape of the reference's own build, geometry.cpp:36-53: a GAS per mesh, an IAS over
// the instances): ONE bottom-level tree per unique mesh in object space + a top-level tree over the instances' world boxes, laid out in
// one node array (top level first, root = node 0) and one record array (instance records, then every mesh's triangles). Memory and
// build time follow the UNIQUE triangles, not the instanced ones. Traversal: vb_traverse_tl (vb_trace.cuh); results are bit for bit
// those of the flattened structure.
static int accel_build_two_level(vb_ctx *ctx, vb_build_stats *out) {
    cudaStream_t st = ctx->stream;
    const uint32_t n_inst = (uint32_t)ctx->instances.size();
    if (n_inst == 0) return ctx->fail(VB_ERR_STATE, "vb_accel_build: no instances set");
    cudaEvent_t e0 = ctx->get_event(), e1 = ctx->get_event();
    VB_CUDA(cudaEventRecord(e0, st));
    // ---- world box of every instance (exact: over its transformed triangles), scene box ----
    std::vector<InstGeom> hg(n_inst);
    std::vector<uint32_t> off(n_inst + 1, 0);
    uint64_t total = 0;
    for (uint32_t i = 0; i < n_inst; ++i) {
        const vb_device_mesh &m = ctx->meshes[ctx->instances[i].mesh_id];
        hg[i].pos = m.pos; hg[i].tri = m.tri; hg[i].soup = m.soup ? 1u : 0u; hg[i].pad = 0;
        memcpy(hg[i].o2w, ctx->instances[i].o2w, 64);
        off[i] = (uint32_t)total; total += m.nt;
    }
    if (total == 0) return ctx->fail(VB_ERR_STATE, "vb_accel_build: scene has no triangles");
    if (total > 0xffffffffull) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_accel_build: more than 4G instanced triangles");
    off[n_inst] = (uint32_t)total;
    DevBuf<InstGeom> d_geom; DevBuf<uint32_t> d_off, d_boxes, d_roots;
    VB_CUDA(d_geom.alloc(n_inst)); VB_CUDA(d_off.alloc(n_inst + 1)); VB_CUDA(d_boxes.alloc(6 * (size_t)n_inst)); VB_CUDA(d_roots.alloc(n_inst));
    std::vector<uint32_t> hb(6 * (size_t)n_inst);
    for (uint32_t i = 0; i < n_inst; ++i) for (int k = 0; k < 6; ++k) hb[6 * (size_t)i + k] = k < 3 ? 0xffffffffu : 0u;
    VB_CUDA(cudaMemcpyAsync(d_geom.p, hg.data(), sizeof(InstGeom) * n_inst, cudaMemcpyHostToDevice, st));
    VB_CUDA(cudaMemcpyAsync(d_off.p, off.data(), sizeof(uint32_t) * (n_inst + 1), cudaMemcpyHostToDevice, st));
    VB_CUDA(cudaMemcpyAsync(d_boxes.p, hb.data(), sizeof(uint32_t) * hb.size(), cudaMemcpyHostToDevice, st));
    k_inst_bounds<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_geom.p, d_off.p, n_inst, (uint32_t)total, d_boxes.p);
    VB_CHECK_LAUNCH();
    VB_CUDA(cudaMemcpyAsync(hb.data(), d_boxes.p, sizeof(uint32_t) * hb.size(), cudaMemcpyDeviceToHost, st));
    VB_CUDA(cudaStreamSynchronize(st));
    double wlo[3] = {1e300, 1e300, 1e300}, whi[3] = {-1e300, -1e300, -1e300};
    for (uint32_t i = 0; i < n_inst; ++i) {
        if (ctx->meshes[ctx->instances[i].mesh_id].nt == 0) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_accel_build (two-level): an instance of an empty mesh");
        for (int k = 0; k < 3; ++k) { wlo[k] = std::min(wlo[k], (double)ord2f(hb[6 * (size_t)i + k])); whi[k] = std::max(whi[k], (double)ord2f(hb[6 * (size_t)i + 3 + k])); }
    }
    double wdiag = 0, wmag = 0;
    for (int k = 0; k < 3; ++k) { wdiag += (whi[k] - wlo[k]) * (whi[k] - wlo[k]); wmag = std::max(wmag, std::max(std::fabs(wlo[k]), std::fabs(whi[k]))); }
    wdiag = std::sqrt(wdiag);
    // ---- transforms: rows of o2w and of w2o per instance; per mesh the padding that covers the rounding of a ray moved into its space:
    // |error of w2o * o| <= ~2^-21 * rowsum(|w2o|) * (|o| + |translation|); 2^-19 * rowsum * (|world| + diag) leaves a factor of four
    std::vector<float> hxf(24 * (size_t)n_inst);
    std::map<uint32_t, float> mesh_pad;
    for (uint32_t i = 0; i < n_inst; ++i) {
        const float *m = ctx->instances[i].o2w;
        double inv[4][4];
        if (!invert_affine(m, inv)) return ctx->fail(VB_ERR_INVALID, "vb_accel_build (two-level): singular instance transform");
        float *x = hxf.data() + 24 * (size_t)i;
        for (int r = 0; r < 3; ++r) { x[4 * r + 0] = m[r]; x[4 * r + 1] = m[4 + r]; x[4 * r + 2] = m[8 + r]; x[4 * r + 3] = m[12 + r]; }
        double rowsum = 0;
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 4; ++c) x[12 + 4 * r + c] = (float)inv[r][c];
            rowsum = std::max(rowsum, std::fabs(inv[r][0]) + std::fabs(inv[r][1]) + std::fabs(inv[r][2]));
        }
        float pad = (float)(0x1p-19 * rowsum * (wmag + wdiag));
        float &mp = mesh_pad[ctx->instances[i].mesh_id];
        mp = std::max(mp, pad);
    }
    // ---- one bottom-level tree per unique mesh ----
    struct Blas { uint4 *nodes; float4 *tris; uint64_t n_nodes, n_tris; int levels; };
    struct BlasSet : std::map<uint32_t, Blas> {          // frees whatever is still held on every way out (VB_CUDA returns included)
        void release() { for (auto &kv : *this) { cudaFree(kv.second.nodes); cudaFree(kv.second.tris); kv.second.nodes = nullptr; kv.second.tris = nullptr; } }
        ~BlasSet() { release(); }
    } blas;
    auto release = [&]() { blas.release(); };
    uint64_t sum_nodes = 0, sum_tris = 0; int max_levels = 0;
    for (auto &kv : mesh_pad) {
        BuildInput in; in.mode = BuildInput::MESH; in.mesh_id = kv.first; in.extra_pad = kv.second;
        int rc = accel_build_tree(ctx, in, "a mesh's bottom-level tree");
        if (rc) { release(); return rc; }
        Blas b{ctx->accel.nodes, ctx->accel.tris, ctx->accel.n_nodes, ctx->accel.n_tris, ctx->last_build_levels};
        ctx->accel = vb_accel();                                   // ownership moved
        blas[kv.first] = b;
        sum_nodes += b.n_nodes; sum_tris += b.n_tris; max_levels = std::max(max_levels, b.levels);
    }
    // ---- top-level tree over the instance boxes; an instance's record carries the ABSOLUTE root node of its mesh's tree ----
    // (absolute positions are known once the top level's node count is: build it with relative roots, patch afterwards)
    std::vector<uint32_t> rel_root(n_inst);
    {
        std::map<uint32_t, uint64_t> node_off; uint64_t o = 0;
        for (auto &kv : blas) { node_off[kv.first] = o; o += kv.second.n_nodes; }
        for (uint32_t i = 0; i < n_inst; ++i) rel_root[i] = (uint32_t)node_off[ctx->instances[i].mesh_id];
    }
    VB_CUDA(cudaMemcpyAsync(d_roots.p, rel_root.data(), sizeof(uint32_t) * n_inst, cudaMemcpyHostToDevice, st));
    BuildInput tin; tin.mode = BuildInput::TOP; tin.inst_boxes = d_boxes.p; tin.inst_roots = d_roots.p;
    int rc = accel_build_tree(ctx, tin, "the top-level tree");
    if (rc) { release(); return rc; }
    const uint64_t top_nodes = ctx->accel.n_nodes, top_recs = ctx->accel.n_tris;
    const int top_levels = ctx->last_build_levels;
    if (top_levels + max_levels + 3 > VB_TRAVERSAL_STACK_ENTRIES) { release(); return ctx->fail(VB_ERR_UNSUPPORTED, "vb_accel_build (two-level): top + bottom levels exceed the traversal stack"); }
    if (top_nodes + sum_nodes > 0xffffffffull || top_recs + sum_tris > 0xffffffffull) { release(); return ctx->fail(VB_ERR_UNSUPPORTED, "vb_accel_build (two-level): more than 4G nodes or records"); }
    // ---- merge: [top nodes][mesh trees ...], [instance records][mesh triangles ...]; indices made absolute ----
    struct Merged { uint4 *nodes = nullptr; float4 *tris = nullptr; float4 *xf = nullptr; bool keep = false;
                    ~Merged() { if (!keep) { cudaFree(nodes); cudaFree(tris); cudaFree(xf); } } } mg;
    uint4 *&nodes = mg.nodes; float4 *&tris = mg.tris; float4 *&xf = mg.xf;
    VB_CUDA(cudaMalloc(&nodes, (top_nodes + sum_nodes) * sizeof(vb_wide_node)));
    VB_CUDA(cudaMalloc(&tris, (top_recs + sum_tris) * 48));
    VB_CUDA(cudaMalloc(&xf, hxf.size() * sizeof(float)));
    VB_CUDA(cudaMemcpyAsync(nodes, ctx->accel.nodes, top_nodes * sizeof(vb_wide_node), cudaMemcpyDeviceToDevice, st));
    VB_CUDA(cudaMemcpyAsync(tris, ctx->accel.tris, top_recs * 48, cudaMemcpyDeviceToDevice, st));
    VB_CUDA(cudaMemcpyAsync(xf, hxf.data(), hxf.size() * sizeof(float), cudaMemcpyHostToDevice, st));
    {   // instance records: relative root -> absolute (the records are float4 triples; .x of the first holds the root's bits)
        std::vector<float> rec(top_recs * 12);
        VB_CUDA(cudaMemcpyAsync(rec.data(), ctx->accel.tris, top_recs * 48, cudaMemcpyDeviceToHost, st));
        VB_CUDA(cudaStreamSynchronize(st));
        for (uint64_t k = 0; k < top_recs; ++k) { uint32_t r; memcpy(&r, &rec[12 * k], 4); r += (uint32_t)top_nodes; memcpy(&rec[12 * k], &r, 4); }
        VB_CUDA(cudaMemcpyAsync(tris, rec.data(), top_recs * 48, cudaMemcpyHostToDevice, st));
        VB_CUDA(cudaStreamSynchronize(st));
    }
    uint64_t no = top_nodes, to = top_recs;
    for (auto &kv : blas) {
        Blas &b = kv.second;
        VB_CUDA(cudaMemcpyAsync(nodes + no * VB_NODE_U4, b.nodes, b.n_nodes * sizeof(vb_wide_node), cudaMemcpyDeviceToDevice, st));
        VB_CUDA(cudaMemcpyAsync(tris + to * 3, b.tris, b.n_tris * 48, cudaMemcpyDeviceToDevice, st));
        k_offset_nodes<<<(unsigned)((b.n_nodes + 255) / 256), 256, 0, st>>>((vb_wide_node *)(nodes + no * VB_NODE_U4), (uint32_t)b.n_nodes, (uint32_t)no, (uint32_t)to);
        no += b.n_nodes; to += b.n_tris;
    }
    VB_CHECK_LAUNCH();
    VB_CUDA(cudaEventRecord(e1, st));
    VB_CUDA(cudaStreamSynchronize(st));
    release();
    cudaFree(ctx->accel.nodes); cudaFree(ctx->accel.tris);
    ctx->accel = vb_accel();
    ctx->accel.nodes = nodes; ctx->accel.tris = tris; ctx->accel.xf = xf; mg.keep = true;
    ctx->accel.n_nodes = top_nodes + sum_nodes; ctx->accel.n_tris = top_recs + sum_tris; ctx->accel.built = true;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    ctx->ev_pool.push_back(e0); ctx->ev_pool.push_back(e1);
    vb_build_stats &s = ctx->build_stats;
    s.n_triangles = sum_tris; s.n_references = sum_tris; s.n_nodes = ctx->accel.n_nodes;
    s.node_bytes = ctx->accel.n_nodes * sizeof(vb_wide_node); s.tri_bytes = ctx->accel.n_tris * 48 + hxf.size() * sizeof(float);
    s.build_ms = ms;
    for (int a = 0; a < 3; ++a) { s.scene_lo[a] = (float)wlo[a]; s.scene_hi[a] = (float)whi[a]; ctx->sort_lo[a] = (float)wlo[a]; ctx->sort_hi[a] = (float)whi[a]; }
    if (out) *out = s;
    return VB_OK;
}

int vb_accel_build_impl(vb_ctx *ctx, vb_build_stats *out) {
    if (ctx->opt_two_level) return accel_build_two_level(ctx, out);
    bool too_deep = false;
    int rc = accel_build_once(ctx, out, true, &too_deep);
    if (rc || !too_deep) return rc;
    if (ctx->last_build_ploc) {
        ctx->prof["bvh_build_radix_fallback"].launches += 1;          // visible through vb_profile_get
        rc = accel_build_once(ctx, out, false, &too_deep);
        if (rc || !too_deep) return rc;
    }
    return ctx->fail(VB_ERR_UNSUPPORTED, "vb_accel_build: the BVH is deeper than the traversal stack (" +
                                             std::to_string(ctx->opt_max_levels) + " wide levels)");
}

// PLOC: parallel locally-ordered clustering (Meister & Bittner 2018) over the Morton-sorted triangle boxes.
// Builds the binary tree the collapse pass (vb_bvh_collapse.cuh) turns into the 8-wide BVH; replaces the
// radix tree (k_karras + k_refit) when option "build.ploc" is on. Included by vb_bvh_build.cu only.
//
// Each round, every cluster looks PLOC_RADIUS neighbours left and right in the current (Morton-ordered) cluster
// array for the partner whose merged box has the smallest surface area; mutual nearest neighbours merge into a
// new node, everything else is kept, the array is compacted in order, repeat until one cluster is left. Unlike
// the radix tree, which splits on key bits only, merges are driven by box area - the agglomerative version of
// the surface-area heuristic - so large and small triangles that share a Morton cell do not end up as siblings.
//
// Node ids: leaves T-1+k (k = position in the Morton order), internal nodes are handed out from T-2 DOWN in
// merge order by a prefix sum (deterministic), so the last merge - the root - is node 0 as the collapse expects.
#pragma once
#include <cub/device/device_scan.cuh>

namespace {

constexpr int PLOC_RADIUS = 16;
constexpr int PLOC_BLOCK = 256;

// leaf boxes with the conservative padding of the intersection contract (same as k_refit)
__global__ void __launch_bounds__(256) k_ploc_leaves(uint32_t T, BuildBufs b, uint32_t *clusters) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= T) return;
    float3 slo = make_float3(ord2f(b.scene[0]), ord2f(b.scene[1]), ord2f(b.scene[2]));
    float3 shi = make_float3(ord2f(b.scene[3]), ord2f(b.scene[4]), ord2f(b.scene[5]));
    float dx = shi.x - slo.x, dy = shi.y - slo.y, dz = shi.z - slo.z;
    float pad = sqrtf(dx * dx + dy * dy + dz * dz) * 0x1p-20f + 1e-30f + b.extra_pad;
    uint32_t src = b.vals[g];
    float4 lo = b.blo[src], hi = b.bhi[src];
    lo.x -= pad; lo.y -= pad; lo.z -= pad; hi.x += pad; hi.y += pad; hi.z += pad;
    b.nlo[T - 1 + g] = lo; b.nhi[T - 1 + g] = hi;
    clusters[g] = T - 1 + g;
}

__device__ __forceinline__ float ploc_merged_area(float4 alo, float4 ahi, float4 blo, float4 bhi) {
    float x = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), y = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y),
          z = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return x * y + y * z + z * x;
}

// nearest neighbour of every cluster within the window; boxes of the block (+ halo) staged in shared memory
__global__ void __launch_bounds__(PLOC_BLOCK) k_ploc_nearest(uint32_t n, const uint32_t *__restrict__ clusters, BuildBufs b, uint32_t *nn) {
    __shared__ float4 s_lo[PLOC_BLOCK + 2 * PLOC_RADIUS], s_hi[PLOC_BLOCK + 2 * PLOC_RADIUS];
    const int base = (int)(blockIdx.x * PLOC_BLOCK) - PLOC_RADIUS;
    for (int k = threadIdx.x; k < PLOC_BLOCK + 2 * PLOC_RADIUS; k += PLOC_BLOCK) {
        int i = base + k;
        if (i >= 0 && i < (int)n) { uint32_t c = clusters[i]; s_lo[k] = b.nlo[c]; s_hi[k] = b.nhi[c]; }
    }
    __syncthreads();
    int i = (int)(blockIdx.x * PLOC_BLOCK + threadIdx.x);
    if (i >= (int)n) return;
    float4 lo = s_lo[threadIdx.x + PLOC_RADIUS], hi = s_hi[threadIdx.x + PLOC_RADIUS];
    float best = FLT_MAX; int best_j = -1;
    int j0 = max(i - PLOC_RADIUS, 0), j1 = min(i + PLOC_RADIUS, (int)n - 1);
    for (int j = j0; j <= j1; ++j) {
        if (j == i) continue;
        float a = ploc_merged_area(lo, hi, s_lo[j - base], s_hi[j - base]);
        if (a < best) { best = a; best_j = j; }          // ties: the lower index
    }
    if (best_j < 0) best_j = i > 0 ? i - 1 : i + 1;      // NaN boxes: any neighbour will do
    nn[i] = (uint32_t)best_j;
}

// per cluster: (merge << 32 | keep); merge = it creates a node (lower index of a mutual pair), keep = it stays in the array
__global__ void __launch_bounds__(256) k_ploc_flags(uint32_t n, const uint32_t *__restrict__ nn, uint64_t *flags) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t j = nn[i];
    bool mutual = nn[j] == i;
    uint64_t merge = mutual && i < j, keep = !(mutual && i > j);
    flags[i] = (merge << 32) | keep;
}

__global__ void __launch_bounds__(256) k_ploc_merge(uint32_t n, uint32_t T, uint32_t merged_before, const uint32_t *__restrict__ clusters,
                                                    const uint32_t *__restrict__ nn, const uint64_t *__restrict__ flags,
                                                    const uint64_t *__restrict__ scan, BuildBufs b, uint32_t max_leaf, uint32_t *out, uint32_t *totals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t f = flags[i], s = scan[i];
    if (i == n - 1) { totals[0] = (uint32_t)(s + f); totals[1] = (uint32_t)((s + f) >> 32); }     // kept, merged this round
    if (!(f & 1u)) return;
    uint32_t id = clusters[i];
    if (f >> 32) {
        uint32_t j = nn[i], L = id, R = clusters[j];
        id = T - 2 - (merged_before + (uint32_t)(s >> 32));
        float4 alo = b.nlo[L], ahi = b.nhi[L], clo = b.nlo[R], chi = b.nhi[R];
        b.nlo[id] = make_float4(fminf(alo.x, clo.x), fminf(alo.y, clo.y), fminf(alo.z, clo.z), 0.f);
        b.nhi[id] = make_float4(fmaxf(ahi.x, chi.x), fmaxf(ahi.y, chi.y), fmaxf(ahi.z, chi.z), 0.f);
        b.left[id] = L; b.right[id] = R;
        b.parent[L] = id; b.parent[R] = id;
        if (id == 0) b.parent[0] = 0xffffffffu;          // the last merge is the root
        b.first[id] = 0; b.last[id] = max_leaf;          // leaf ranges are not contiguous here: never "leaf-like" for the collapse
    }
    out[(uint32_t)s] = id;
}

}  // namespace

// clusters / clusters2 / nn: T uint32 each; flags / scan: T uint64 each
static int ploc_build(vb_ctx *ctx, uint32_t T, BuildBufs b, uint32_t *clusters, uint32_t *clusters2, uint32_t *nn, uint64_t *flags, uint64_t *scan,
                      uint32_t *d_totals, uint32_t max_leaf, uint64_t &launches) {
    cudaStream_t st = ctx->stream;
    k_ploc_leaves<<<(T + 255) / 256, 256, 0, st>>>(T, b, clusters); launches++;
    VB_CHECK_LAUNCH();
    size_t tmp_bytes = 0;
    VB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flags, scan, (int)T, st));
    int rc = ctx->ensure_stage(tmp_bytes);
    if (rc) return rc;
    uint32_t n = T, merged = 0;
    int rounds = 0;
    while (n > 1) {
        k_ploc_nearest<<<(n + PLOC_BLOCK - 1) / PLOC_BLOCK, PLOC_BLOCK, 0, st>>>(n, clusters, b, nn);
        k_ploc_flags<<<(n + 255) / 256, 256, 0, st>>>(n, nn, flags);
        VB_CUDA(cub::DeviceScan::ExclusiveSum(ctx->stage, tmp_bytes, flags, scan, (int)n, st));
        k_ploc_merge<<<(n + 255) / 256, 256, 0, st>>>(n, T, merged, clusters, nn, flags, scan, b, max_leaf, clusters2, d_totals);
        launches += 4;
        VB_CHECK_LAUNCH();
        uint32_t h[2];
        VB_CUDA(cudaMemcpyAsync(h, d_totals, sizeof(h), cudaMemcpyDeviceToHost, st));
        VB_CUDA(cudaStreamSynchronize(st));
        if (h[1] == 0 || h[0] >= n) return ctx->fail(VB_ERR_STATE, "vb_accel_build: PLOC made no progress");
        n = h[0]; merged += h[1];
        std::swap(clusters, clusters2);
        if (++rounds > 4096) return ctx->fail(VB_ERR_STATE, "vb_accel_build: PLOC did not terminate");
    }
    if (merged != T - 1) return ctx->fail(VB_ERR_STATE, "vb_accel_build: PLOC node count mismatch");
    return VB_OK;
}

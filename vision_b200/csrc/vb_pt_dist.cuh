// Multi-GPU film reduction (reference accumulation frame_buffer.cpp:117-126 is single-GPU): every rank renders its
// frame indices into a SUM film (vb_render_frames(.., VB_ACCUM_SUM)), then ONE exchange makes the film whole.
//   vb_film_allreduce      ncclAllReduce(sum, fp32) of the ACCUM film, in place, on the context's stream (one process per GPU).
//                          NCCL is bound at run time (dlopen): the copy already loaded in the process (e.g. by torch) is reused, so a
//                          ncclComm_t made by the caller's NCCL and the calls made here always belong to the same library.
//   vb_film_reduce_peers   one process, N contexts: a fused reduce + resolve kernel over NVLink peer memory. GPU g owns pixel slice g:
//                          it reads that slice of all N films (summed in RANK ORDER, so the result is deterministic and independent of
//                          N's schedule), resolves it (mean -> exposure -> tone map) and writes sum and picture into the root's film.
// Included by vb_pt_kernel.cuh after vb_pt_state.
#pragma once
#include <dlfcn.h>
#include <nccl.h>       // types and prototypes only; nothing is linked
#include <mutex>

namespace {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommCount)(const ncclComm_t, int *) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    std::string error;
};

NcclApi *nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *env = getenv("VB_NCCL_LIB");
        if (env && *env) api.lib = dlopen(env, RTLD_NOW | RTLD_LOCAL);
        if (!api.lib) api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);     // the process's own NCCL, if it has one
        if (!api.lib) api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!api.lib) api.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
        if (!api.lib) { api.error = std::string("NCCL not found (libnccl.so.2; set VB_NCCL_LIB): ") + dlerror(); return; }
        auto sym = [&](const char *name) { void *p = dlsym(api.lib, name); if (!p && api.error.empty()) api.error = std::string("NCCL symbol missing: ") + name; return p; };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.CommCount = (decltype(api.CommCount))sym("ncclCommCount");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
    });
    return &api;
}

#define VB_NCCL(call)                                                                                              \
    do {                                                                                                           \
        ncclResult_t r_ = (call);                                                                                  \
        if (r_ != ncclSuccess) return ctx->fail(VB_ERR_CUDA, std::string(#call) + ": " + api->GetErrorString(r_)); \
    } while (0)

struct PeerFilms { const float4 *film[VB_MAX_PEERS]; };

__device__ __forceinline__ float4 film_resolve_px(float4 v, int mode, float exposure, int tonemap) {
    if (mode == VB_ACCUM_SUM) {
        float inv = v.w > 0.f ? 1.0f / v.w : 0.f;
        v.x *= inv; v.y *= inv; v.z *= inv;
    }
    float c[3] = {v.x, v.y, v.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float x = 1.0f - expf(-c[k] * exposure);
        if (tonemap == VB_TONEMAP_ACES) {
            float y = (x * (2.51f * x + 0.03f)) / (x * (2.43f * x + 0.59f) + 0.14f);
            x = fmin_(1.0f, fmax_(0.0f, y));
        } else if (tonemap == VB_TONEMAP_REINHARD) {
            x = x / (x + 1.0f);
        }
        c[k] = x;
    }
    return make_float4(c[0], c[1], c[2], 1.0f);
}

// pixels [begin, end): sum of the N SUM films in rank order -> root's accum, resolved picture -> root's output. Peer loads are
// 16-byte coalesced reads over NVLink; the two stores are posted writes into the root's HBM.
__global__ void __launch_bounds__(256) k_film_reduce_resolve(PeerFilms films, uint32_t n_films, uint32_t begin, uint32_t end,
                                                             float4 *root_accum, float4 *root_out, float exposure, int tonemap) {
    for (uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x; i < end; i += gridDim.x * blockDim.x) {
        float4 s = __ldg(films.film[0] + i);
        for (uint32_t r = 1; r < n_films; ++r) {
            float4 v = __ldg(films.film[r] + i);
            s.x = __fadd_rn(s.x, v.x); s.y = __fadd_rn(s.y, v.y); s.z = __fadd_rn(s.z, v.z); s.w = __fadd_rn(s.w, v.w);
        }
        root_accum[i] = s;
        root_out[i] = film_resolve_px(s, VB_ACCUM_SUM, exposure, tonemap);
    }
}

}  // namespace

int vb_nccl_unique_id(void *id128) {
    if (!id128) return VB_ERR_INVALID;
    NcclApi *api = nccl_api();
    if (!api->error.empty()) return VB_ERR_UNSUPPORTED;
    static_assert(sizeof(ncclUniqueId) == VB_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size");
    return api->GetUniqueId((ncclUniqueId *)id128) == ncclSuccess ? VB_OK : VB_ERR_CUDA;
}

int vb_nccl_comm_create(vb_ctx *ctx, const void *id128, int n_ranks, int rank, void **comm) {
    if (!ctx) return VB_ERR_INVALID;
    if (!id128 || !comm || n_ranks < 1 || rank < 0 || rank >= n_ranks) return ctx->fail(VB_ERR_INVALID, "vb_nccl_comm_create: bad arguments");
    NcclApi *api = nccl_api();
    if (!api->error.empty()) return ctx->fail(VB_ERR_UNSUPPORTED, api->error);
    VB_CUDA(cudaSetDevice(ctx->device));
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t c = nullptr;
    VB_NCCL(api->CommInitRank(&c, n_ranks, id, rank));
    *comm = c;
    return VB_OK;
}

int vb_nccl_comm_destroy(vb_ctx *ctx, void *comm) {
    if (!ctx) return VB_ERR_INVALID;
    if (!comm) return VB_OK;
    NcclApi *api = nccl_api();
    if (!api->error.empty()) return ctx->fail(VB_ERR_UNSUPPORTED, api->error);
    cudaSetDevice(ctx->device);
    VB_NCCL(api->CommDestroy((ncclComm_t)comm));
    return VB_OK;
}

int vb_film_allreduce(vb_ctx *ctx, void *nccl_comm) {
    if (!ctx) return VB_ERR_INVALID;
    vb_pt_state *p = pt_state(ctx);
    if (!p->accum) return ctx->fail(VB_ERR_STATE, "vb_film_allreduce: no film");
    if (!nccl_comm) return ctx->fail(VB_ERR_INVALID, "vb_film_allreduce: null communicator");
    NcclApi *api = nccl_api();
    if (!api->error.empty()) return ctx->fail(VB_ERR_UNSUPPORTED, api->error);
    VB_CUDA(cudaSetDevice(ctx->device));
    ctx->prof_begin("film_allreduce", 1);
    ncclResult_t r = api->AllReduce(p->accum, p->accum, (size_t)p->w * p->h * 4, ncclFloat32, ncclSum, (ncclComm_t)nccl_comm, ctx->stream);
    ctx->prof_end();
    if (r != ncclSuccess) return ctx->fail(VB_ERR_CUDA, std::string("ncclAllReduce: ") + api->GetErrorString(r));
    return VB_OK;
}

int vb_film_reduce_peers(vb_ctx *const *ctxs, uint32_t n, uint32_t root, float exposure, int tonemap) {
    if (!ctxs || n == 0 || root >= n) return VB_ERR_INVALID;
    for (uint32_t i = 0; i < n; ++i) if (!ctxs[i]) return VB_ERR_INVALID;
    vb_ctx *ctx = ctxs[root];                              // errors are reported on the root context
    if (n > VB_MAX_PEERS) return ctx->fail(VB_ERR_INVALID, "vb_film_reduce_peers: more than VB_MAX_PEERS contexts");
    vb_pt_state *pr = pt_state(ctx);
    if (!pr->accum) return ctx->fail(VB_ERR_STATE, "vb_film_reduce_peers: no film on the root");
    PeerFilms films{};
    for (uint32_t i = 0; i < n; ++i) {
        vb_pt_state *p = pt_state(ctxs[i]);
        if (!p->accum || p->w != pr->w || p->h != pr->h) return ctx->fail(VB_ERR_STATE, "vb_film_reduce_peers: films differ in size or are missing");
        for (uint32_t j = 0; j < i; ++j) if (ctxs[j]->device == ctxs[i]->device) return ctx->fail(VB_ERR_INVALID, "vb_film_reduce_peers: two contexts on one device");
        films.film[i] = p->accum;
    }
    const uint32_t n_pix = pr->w * pr->h;
    if (n == 1) {
        VB_CUDA(cudaSetDevice(ctx->device));
        k_film_resolve<<<(n_pix + 255) / 256, 256, 0, ctx->stream>>>(pr->accum, pr->output, n_pix, VB_ACCUM_SUM, exposure, tonemap);
        VB_CHECK_LAUNCH();
        return VB_OK;
    }
    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t j = 0; j < n; ++j) {
            if (i == j) continue;
            int can = 0;
            VB_CUDA(cudaDeviceCanAccessPeer(&can, ctxs[i]->device, ctxs[j]->device));
            if (!can) return ctx->fail(VB_ERR_UNSUPPORTED, "vb_film_reduce_peers: no peer access between the devices (use vb_film_allreduce)");
            VB_CUDA(cudaSetDevice(ctxs[i]->device));
            cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[j]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return ctx->fail(VB_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
            cudaGetLastError();
        }
    // rendered[i]: context i's film is complete; reduced[g]: slice g has been read everywhere and written to the root
    std::vector<cudaEvent_t> rendered(n), reduced(n);
    for (uint32_t i = 0; i < n; ++i) {
        VB_CUDA(cudaSetDevice(ctxs[i]->device));
        VB_CUDA(cudaEventCreateWithFlags(&rendered[i], cudaEventDisableTiming));
        VB_CUDA(cudaEventCreateWithFlags(&reduced[i], cudaEventDisableTiming));
        VB_CUDA(cudaEventRecord(rendered[i], ctxs[i]->stream));
    }
    // root's own accum is both an input (films[root]) and the output: slice g reads films[root][slice g] and writes the same
    // addresses from the same thread, so in-place is safe.
    for (uint32_t g = 0; g < n; ++g) {
        VB_CUDA(cudaSetDevice(ctxs[g]->device));
        for (uint32_t i = 0; i < n; ++i) if (i != g) VB_CUDA(cudaStreamWaitEvent(ctxs[g]->stream, rendered[i], 0));
        uint32_t begin = (uint32_t)((uint64_t)n_pix * g / n), end = (uint32_t)((uint64_t)n_pix * (g + 1) / n);
        unsigned grid = std::min<unsigned>((end - begin + 255) / 256, (unsigned)ctxs[g]->sm_count * 8);
        if (end > begin) k_film_reduce_resolve<<<grid, 256, 0, ctxs[g]->stream>>>(films, n, begin, end, pr->accum, pr->output, exposure, tonemap);
        VB_CHECK_LAUNCH();
        VB_CUDA(cudaEventRecord(reduced[g], ctxs[g]->stream));
    }
    for (uint32_t i = 0; i < n; ++i) {                      // nobody touches a film again before every slice has been read
        VB_CUDA(cudaSetDevice(ctxs[i]->device));
        for (uint32_t g = 0; g < n; ++g) if (g != i) VB_CUDA(cudaStreamWaitEvent(ctxs[i]->stream, reduced[g], 0));
    }
    for (uint32_t i = 0; i < n; ++i) { cudaEventDestroy(rendered[i]); cudaEventDestroy(reduced[i]); }
    VB_CUDA(cudaSetDevice(ctx->device));
    return VB_OK;
}

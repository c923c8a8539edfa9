// Context lifecycle, mesh/instance tables, profiling slots of libvision_b200.so.
#include "vb_internal.h"
#include <cstring>
#include <algorithm>

static std::string g_create_err;

int vb_ctx::ensure_copy_streams() {
    if (copy_in) return VB_OK;
    vb_ctx *ctx = this;
    VB_CUDA(cudaStreamCreateWithFlags(&copy_in, cudaStreamNonBlocking));
    VB_CUDA(cudaStreamCreateWithFlags(&copy_out, cudaStreamNonBlocking));
    for (int i = 0; i < VB_MAX_HOST_CHUNKS; ++i) {
        VB_CUDA(cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming));
        VB_CUDA(cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming));
    }
    return VB_OK;
}
int vb_ctx::ensure_stage(size_t bytes) {
    if (bytes <= stage_bytes) return VB_OK;
    vb_ctx *ctx = this;
    if (stage) cudaFree(stage);
    stage = nullptr; stage_bytes = 0;
    VB_CUDA(cudaMalloc(&stage, bytes));
    stage_bytes = bytes;
    return VB_OK;
}

cudaEvent_t vb_ctx::get_event() {
    if (!ev_pool.empty()) { cudaEvent_t e = ev_pool.back(); ev_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
void vb_ctx::prof_begin(const char *name, uint64_t launches) {
    cur_name = name; cur_launches = launches;
    cur_start = get_event();
    cudaEventRecord(cur_start, stream);
}
void vb_ctx::prof_end() {
    cudaEvent_t e = get_event();
    cudaEventRecord(e, stream);
    pending.push_back({std::string(cur_name), {cur_start, e}});
    prof[cur_name].launches += cur_launches;
    prof["all"].launches += cur_launches;
    cur_start = nullptr;
    if (pending.size() > 4096) { cudaStreamSynchronize(stream); prof_resolve(); }
}
void vb_ctx::prof_resolve() {
    for (auto &p : pending) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, p.second.first, p.second.second) == cudaSuccess) {
            prof[p.first].ms += ms;
            prof["all"].ms += ms;
        }
        ev_pool.push_back(p.second.first); ev_pool.push_back(p.second.second);
    }
    pending.clear();
}

extern "C" {

int vb_ctx_create(int device, vb_ctx **out) {
    if (!out) return VB_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_create_err = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (libvision_b200 has no CPU fallback)";
        return VB_ERR_CUDA;
    }
    if (device < 0 || device >= n) { g_create_err = "device index out of range"; return VB_ERR_INVALID; }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_err = cudaGetErrorString(e); return VB_ERR_CUDA; }
    cudaDeviceProp prop{};
    cudaGetDeviceProperties(&prop, device);
    if (prop.major != 10) {
        g_create_err = "libvision_b200 is built for sm_100a only; device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor);
        return VB_ERR_UNSUPPORTED;
    }
    vb_ctx *c = new vb_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking)) != cudaSuccess) {
        g_create_err = cudaGetErrorString(e); delete c; return VB_ERR_CUDA;
    }
    c->stream = c->own_stream;
    *out = c;
    return VB_OK;
}

int vb_ctx_destroy(vb_ctx *ctx) {
    if (!ctx) return VB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->prof_resolve();
    for (auto e : ctx->ev_pool) cudaEventDestroy(e);
    for (auto &m : ctx->meshes) {
        if (m.owns_pos) cudaFree(m.pos);
        cudaFree(m.nrm); cudaFree(m.uv); cudaFree(m.tri); cudaFree(m.vtx); cudaFree(m.tri4);
    }
    cudaFree(ctx->d_instances);
    cudaFree(ctx->accel.nodes); cudaFree(ctx->accel.tris); cudaFree(ctx->accel.xf);
    cudaFree(ctx->stage); cudaFree(ctx->sort_buf);
    for (int k = 0; k < 2; ++k) { cudaFree(ctx->rec_stage[k]); if (ctx->rec_done[k]) cudaEventDestroy(ctx->rec_done[k]); if (ctx->rec_traced[k]) cudaEventDestroy(ctx->rec_traced[k]); }
    vb_pt_destroy(ctx);
    if (ctx->copy_in) {
        cudaStreamDestroy(ctx->copy_in); cudaStreamDestroy(ctx->copy_out);
        for (int i = 0; i < VB_MAX_HOST_CHUNKS; ++i) { cudaEventDestroy(ctx->ev_in[i]); cudaEventDestroy(ctx->ev_out[i]); }
    }
    cudaStreamDestroy(ctx->own_stream);
    delete ctx;
    return VB_OK;
}

const char *vb_last_error(vb_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int vb_ctx_set_stream(vb_ctx *ctx, void *s) {
    if (!ctx) return VB_ERR_INVALID;
    ctx->stream = s ? (cudaStream_t)s : ctx->own_stream;
    return VB_OK;
}
int vb_ctx_synchronize(vb_ctx *ctx) {
    if (!ctx) return VB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->rec_pending) {                  // asynchronous host-buffer trace calls: their copy streams too
        VB_CUDA(cudaStreamSynchronize(ctx->copy_in));
        VB_CUDA(cudaStreamSynchronize(ctx->copy_out));
        ctx->rec_pending = false;
    }
    return VB_OK;
}
int vb_ctx_set_option(vb_ctx *ctx, const char *key, double value) {
    if (!ctx || !key) return VB_ERR_INVALID;
    std::string k(key);
    if (k == "trace.refill_thr") ctx->opt_refill_thr = (int)value;
    else if (k == "trace.tri_thr") ctx->opt_tri_thr = (int)value;
    else if (k == "build.max_leaf") { if ((int)value != 1) return ctx->fail(VB_ERR_INVALID, "vb_ctx_set_option: build.max_leaf is fixed at 1 (one triangle per leaf slot)"); }
    else if (k == "build.ploc") ctx->opt_ploc = value < 0 ? -1 : (value != 0 ? 1 : 0);
    else if (k == "build.max_levels") ctx->opt_max_levels = std::min(VB_TRAVERSAL_STACK_ENTRIES, std::max(1, (int)value));   // test hook: lower the depth the build accepts
    else if (k == "build.two_level") ctx->opt_two_level = value != 0 ? 1 : 0;
    else if (k == "build.collapse") ctx->opt_collapse = value != 0 ? 1 : 0;
    else if (k == "build.split_pct") ctx->opt_split_pct = std::min(400, std::max(0, (int)value));
    else if (k == "build.size_classes") ctx->opt_size_classes = value != 0 ? 1 : 0;
    else if (k == "trace.sort_rays") ctx->opt_sort_rays = (int)value;
    else if (k == "trace.sort_skip_bits") ctx->opt_sort_skip_bits = std::min(16, std::max(0, (int)value));
    else if (k == "trace.sort_oct_shift") ctx->opt_sort_oct_shift = std::min(21, std::max(0, (int)value));
    else if (k == "trace.sort_min_rays") ctx->opt_sort_min_rays = std::max(1, (int)value);
    else if (k == "pt.schedule") ctx->opt_pt_schedule = value != 0 ? 1 : 0;
    else if (k == "pt.sort_material") ctx->opt_pt_sort_material = value != 0 ? 1 : 0;
    else if (k == "pt.split_principled") ctx->opt_pt_split_principled = value != 0 ? 1 : 0;
    else if (k == "pt.spec_cache") ctx->opt_pt_spec_cache = value != 0 ? 1 : 0;
    else if (k == "pt.count_work") ctx->opt_pt_count_work = value != 0 ? 1 : 0;
    else if (k == "pt.max_paths") ctx->opt_pt_max_paths = std::max<int64_t>(1, (int64_t)value);
    else if (k == "trace.host_chunks") ctx->opt_host_chunks = std::min(VB_MAX_HOST_CHUNKS, std::max(1, (int)value));
    else if (k == "trace.host_chunk_min_rays") ctx->opt_host_chunk_min_rays = std::max(1024, (int)value);
    else if (k == "trace.async") ctx->opt_trace_async = value != 0 ? 1 : 0;
    else if (k == "trace.async_chunks") ctx->opt_async_chunks = std::min(VB_MAX_HOST_CHUNKS, std::max(1, (int)value));
    else if (k == "trace.blocks_per_sm") ctx->opt_blocks_per_sm = std::max(1, (int)value);
    else return ctx->fail(VB_ERR_INVALID, "vb_ctx_set_option: unknown key " + k);
    return VB_OK;
}
// ---- read-bandwidth probe: the L2 (or HBM) denominator of the path tracer's roofline ----
namespace {
__global__ void __launch_bounds__(256) k_probe_read(const uint4 *__restrict__ p, uint64_t n16, uint32_t *sink) {
    uint32_t acc = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
        uint4 v = __ldg(p + i);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    if (acc == 0x9e3779b9u) *sink = acc;       // never true for the zero-filled buffer: keeps the loads alive
}
}  // namespace
int vb_probe_read_bandwidth(vb_ctx *ctx, uint64_t bytes, int reps, double *gbs) {
    if (!ctx || !gbs || bytes < 4096 || reps < 1) return VB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    void *buf = nullptr; uint32_t *sink = nullptr;
    VB_CUDA(cudaMalloc(&buf, bytes));
    if (cudaMalloc(&sink, 4) != cudaSuccess) { cudaFree(buf); return ctx->fail(VB_ERR_CUDA, "vb_probe_read_bandwidth: cudaMalloc"); }
    cudaMemsetAsync(buf, 0, bytes, st);
    const uint64_t n16 = bytes / 16;
    const unsigned grid = (unsigned)ctx->sm_count * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w) k_probe_read<<<grid, 256, 0, st>>>((const uint4 *)buf, n16, sink);     // warm: the buffer becomes L2-resident if it fits
    cudaEventRecord(e0, st);
    for (int r = 0; r < reps; ++r) k_probe_read<<<grid, 256, 0, st>>>((const uint4 *)buf, n16, sink);
    cudaEventRecord(e1, st);
    cudaError_t ce = cudaStreamSynchronize(st);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(buf); cudaFree(sink);
    if (ce != cudaSuccess) return ctx->fail(VB_ERR_CUDA, std::string("vb_probe_read_bandwidth: ") + cudaGetErrorString(ce));
    *gbs = (double)n16 * 16.0 * reps / (ms * 1e-3) / 1e9;
    return VB_OK;
}

int vb_profile_reset(vb_ctx *ctx) {
    if (!ctx) return VB_ERR_INVALID;
    cudaStreamSynchronize(ctx->stream);
    ctx->prof_resolve();
    ctx->prof.clear();
    return VB_OK;
}
int vb_profile_get(vb_ctx *ctx, const char *name, double *ms, uint64_t *launches) {
    if (!ctx || !name) return VB_ERR_INVALID;
    cudaStreamSynchronize(ctx->stream);
    ctx->prof_resolve();
    auto it = ctx->prof.find(name);
    if (ms) *ms = it == ctx->prof.end() ? 0.0 : it->second.ms;
    if (launches) *launches = it == ctx->prof.end() ? 0 : it->second.launches;
    return VB_OK;
}

// shading-side copies of an indexed mesh: interleaved vertices and padded index triples (same values, fewer and wider loads)
__global__ void __launch_bounds__(256) k_interleave_mesh(const float *pos, const float *nrm, const float *uv, uint32_t nv, float4 *vtx,
                                                         const uint32_t *tri, uint32_t nt, uint4 *tri4) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < nv) {
        vtx[2 * (size_t)g] = make_float4(pos[3 * (size_t)g], pos[3 * (size_t)g + 1], pos[3 * (size_t)g + 2], uv[2 * (size_t)g]);
        vtx[2 * (size_t)g + 1] = make_float4(nrm[3 * (size_t)g], nrm[3 * (size_t)g + 1], nrm[3 * (size_t)g + 2], uv[2 * (size_t)g + 1]);
    }
    if (g < nt) tri4[g] = make_uint4(tri[3 * (size_t)g], tri[3 * (size_t)g + 1], tri[3 * (size_t)g + 2], 0u);
}

int vb_mesh_upload(vb_ctx *ctx, const float *pos3, const float *nrm3, const float *uv2, uint32_t nv,
                   const uint32_t *tri3, uint32_t nt, uint32_t *mesh_id) {
    if (!ctx) return VB_ERR_INVALID;
    if (!pos3 || !tri3 || !mesh_id) return ctx->fail(VB_ERR_INVALID, "vb_mesh_upload: null argument");
    for (uint64_t i = 0; i < 3ull * nt; ++i)
        if (tri3[i] >= nv) return ctx->fail(VB_ERR_INVALID, "vb_mesh_upload: triangle index out of range");
    cudaSetDevice(ctx->device);
    vb_device_mesh m;
    m.nv = nv; m.nt = nt;
    if (nv) {
        VB_CUDA(cudaMalloc(&m.pos, sizeof(float) * 3 * nv));
        VB_CUDA(cudaMemcpyAsync(m.pos, pos3, sizeof(float) * 3 * nv, cudaMemcpyHostToDevice, ctx->stream));
        VB_CUDA(cudaMalloc(&m.nrm, sizeof(float) * 3 * nv));
        if (nrm3) VB_CUDA(cudaMemcpyAsync(m.nrm, nrm3, sizeof(float) * 3 * nv, cudaMemcpyHostToDevice, ctx->stream));
        else VB_CUDA(cudaMemsetAsync(m.nrm, 0, sizeof(float) * 3 * nv, ctx->stream));
        VB_CUDA(cudaMalloc(&m.uv, sizeof(float) * 2 * nv));
        if (uv2) VB_CUDA(cudaMemcpyAsync(m.uv, uv2, sizeof(float) * 2 * nv, cudaMemcpyHostToDevice, ctx->stream));
        else VB_CUDA(cudaMemsetAsync(m.uv, 0, sizeof(float) * 2 * nv, ctx->stream));
    }
    if (nt) {
        VB_CUDA(cudaMalloc(&m.tri, sizeof(uint32_t) * 3 * nt));
        VB_CUDA(cudaMemcpyAsync(m.tri, tri3, sizeof(uint32_t) * 3 * nt, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (nv && nt) {
        VB_CUDA(cudaMalloc(&m.vtx, sizeof(float4) * 2 * (size_t)nv));
        VB_CUDA(cudaMalloc(&m.tri4, sizeof(uint4) * (size_t)nt));
        const uint32_t nmax = nv > nt ? nv : nt;
        k_interleave_mesh<<<(nmax + 255) / 256, 256, 0, ctx->stream>>>(m.pos, m.nrm, m.uv, nv, m.vtx, m.tri, nt, m.tri4);
        VB_CHECK_LAUNCH();
    }
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    *mesh_id = (uint32_t)ctx->meshes.size();
    ctx->meshes.push_back(m);
    ctx->accel.built = false;
    ctx->geom_version++;
    return VB_OK;
}

int vb_mesh_upload_soup(vb_ctx *ctx, const float *pos9, vb_mem mem, uint64_t nt, uint32_t *mesh_id) {
    if (!ctx) return VB_ERR_INVALID;
    if (!pos9 || !mesh_id) return ctx->fail(VB_ERR_INVALID, "vb_mesh_upload_soup: null argument");
    if (nt > 0x55555555ull) return ctx->fail(VB_ERR_INVALID, "vb_mesh_upload_soup: too many triangles");
    cudaSetDevice(ctx->device);
    vb_device_mesh m;
    m.soup = true; m.nt = nt; m.nv = (uint32_t)(3 * nt);
    if (nt) {
        VB_CUDA(cudaMalloc(&m.pos, sizeof(float) * 9 * nt));
        VB_CUDA(cudaMemcpyAsync(m.pos, pos9, sizeof(float) * 9 * nt,
                                mem == VB_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, ctx->stream));
        VB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    *mesh_id = (uint32_t)ctx->meshes.size();
    ctx->meshes.push_back(m);
    ctx->accel.built = false;
    ctx->geom_version++;
    return VB_OK;
}

int vb_instances_set(vb_ctx *ctx, const vb_instance *inst, uint32_t n) {
    if (!ctx) return VB_ERR_INVALID;
    if (n && !inst) return ctx->fail(VB_ERR_INVALID, "vb_instances_set: null argument");
    for (uint32_t i = 0; i < n; ++i)
        if (inst[i].mesh_id >= ctx->meshes.size()) return ctx->fail(VB_ERR_INVALID, "vb_instances_set: unknown mesh_id");
    cudaSetDevice(ctx->device);
    ctx->instances.assign(inst, inst + n);
    cudaFree(ctx->d_instances); ctx->d_instances = nullptr;
    if (n) {
        VB_CUDA(cudaMalloc(&ctx->d_instances, sizeof(vb_instance) * n));
        VB_CUDA(cudaMemcpy(ctx->d_instances, inst, sizeof(vb_instance) * n, cudaMemcpyHostToDevice));
    }
    ctx->accel.built = false;
    ctx->geom_version++;
    return VB_OK;
}

int vb_meshes_clear(vb_ctx *ctx) {
    if (!ctx) return VB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    for (auto &m : ctx->meshes) {
        if (m.owns_pos) cudaFree(m.pos);
        cudaFree(m.nrm); cudaFree(m.uv); cudaFree(m.tri); cudaFree(m.vtx); cudaFree(m.tri4);
    }
    ctx->meshes.clear();
    ctx->instances.clear();
    cudaFree(ctx->d_instances); ctx->d_instances = nullptr;
    cudaFree(ctx->accel.nodes); cudaFree(ctx->accel.tris); cudaFree(ctx->accel.xf);
    ctx->accel = vb_accel();
    ctx->geom_version++;
    return VB_OK;
}

int vb_accel_build(vb_ctx *ctx, const vb_build_opts *opts, vb_build_stats *out) {
    if (!ctx) return VB_ERR_INVALID;
    if (opts && (opts->flags & ~VB_BUILD_TWO_LEVEL)) return ctx->fail(VB_ERR_INVALID, "vb_accel_build: unknown flag in vb_build_opts");
    cudaSetDevice(ctx->device);
    const int saved = ctx->opt_two_level;
    if (opts && (opts->flags & VB_BUILD_TWO_LEVEL)) ctx->opt_two_level = 1;
    int rc = vb_accel_build_impl(ctx, out);
    ctx->opt_two_level = saved;
    return rc;
}

int vb_accel_download(vb_ctx *ctx, void *nodes, void *tris) {
    if (!ctx) return VB_ERR_INVALID;
    if (!ctx->accel.built) return ctx->fail(VB_ERR_STATE, "vb_accel_download: accel not built");
    cudaSetDevice(ctx->device);
    VB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (nodes && ctx->accel.n_nodes) VB_CUDA(cudaMemcpy(nodes, ctx->accel.nodes, ctx->accel.n_nodes * sizeof(vb_wide_node), cudaMemcpyDeviceToHost));
    if (tris && ctx->accel.n_tris) VB_CUDA(cudaMemcpy(tris, ctx->accel.tris, ctx->accel.n_tris * 48, cudaMemcpyDeviceToHost));
    return VB_OK;
}

}  // extern "C"

// Internal declarations of libvision_b200.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <map>
#include "../../include/vision_b200.h"

#define VB_CUDA(call)                                                                         \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            return ctx->fail(VB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
        }                                                                                     \
    } while (0)

#define VB_CHECK_LAUNCH() VB_CUDA(cudaGetLastError())

// entries of the traversal stack (vb_trace.cuh: VB_SM_STACK in shared memory + VB_LOCAL_STACK in local memory); the build refuses deeper trees
#define VB_TRAVERSAL_STACK_ENTRIES 64

struct vb_device_mesh {
    float *pos = nullptr;      // 3*nv
    float *nrm = nullptr;      // 3*nv (may be null)
    float *uv = nullptr;       // 2*nv (may be null)
    uint32_t *tri = nullptr;   // 3*nt (null for soups)
    float4 *vtx = nullptr;     // indexed meshes: 2 float4 per vertex (pos.xyz, u) (nrm.xyz, v) - one 32-byte sector per vertex for the shading gather
    uint4 *tri4 = nullptr;     // indexed meshes: (i0, i1, i2, 0) per triangle - one 16-byte load instead of three 4-byte ones
    uint32_t nv = 0;
    uint64_t nt = 0;
    bool soup = false;
    bool owns_pos = true;
};

// Compressed 8-wide node: 80 bytes of payload in a 96-byte, 32-byte-aligned slot so that it is
// fetched with two 256-bit loads + one 128-bit load (3 L1 tag lookups per lane instead of 5) and
// always covers exactly three 32-byte sectors. See DESIGN.md "Wide node".
#define VB_NODE_U4 6           /* node stride in uint4 units */
struct __align__(32) vb_wide_node {
    float px, py, pz;          // quantisation origin
    uint8_t ex, ey, ez, imask; // per-axis scale exponent (float exponent bits), internal-child mask
    uint32_t child_base;       // index of first internal child (wide node index)
    uint32_t tri_base;         // index of first triangle of this node's leaf children
    uint8_t meta[8];
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
    uint32_t pad[4];
};
static_assert(sizeof(vb_wide_node) == 16 * VB_NODE_U4, "wide node slot must be 96 bytes");

#define VB_MAX_HOST_CHUNKS 16

struct vb_prof_slot { double ms = 0; uint64_t launches = 0; };

struct vb_accel {
    uint4 *nodes = nullptr;    // n_nodes * VB_NODE_U4 uint4
    float4 *tris = nullptr;    // n_tris * 3 float4: (v0,prim) (e1,inst) (e2,0)
    float4 *xf = nullptr;      // two-level structure only: 6 float4 per instance (rows of o2w, rows of w2o); then `nodes` = top level + all
                               // bottom-level trees, `tris` = instance records + object-space triangles (p0,prim) (p1) (p2) (vb_trace.cuh)
    uint64_t n_nodes = 0, n_tris = 0;
    bool built = false;
};

struct vb_pt_state;   // vb_pt.cu

struct vb_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    std::string err;
    std::vector<vb_device_mesh> meshes;
    std::vector<vb_instance> instances;
    uint64_t geom_version = 0;               // bumped by every mesh / instance change: the path tracer re-derives its tables only then
    vb_instance *d_instances = nullptr;
    vb_accel accel;
    vb_build_stats build_stats{};
    float sort_lo[3] = {0, 0, 0}, sort_hi[3] = {1, 1, 1};   // box of the ray-sort Morton key (vb_bvh_collapse.cuh)
    std::map<std::string, vb_prof_slot> prof;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::vector<std::pair<std::string, std::pair<cudaEvent_t, cudaEvent_t>>> pending;  // un-resolved timings
    std::vector<cudaEvent_t> ev_pool;
    // staging for host-pointer calls
    void *stage = nullptr; size_t stage_bytes = 0;
    int sm_count = 148;
    // tunables (vb_ctx_set_option)
    int opt_refill_thr = 24, opt_tri_thr = 12, opt_blocks_per_sm = 8;
    int opt_max_leaf = 1;
    int opt_ploc = -1;                       // build: 1 = PLOC clustering (vb_bvh_ploc.cuh), 0 = radix tree, -1 = automatic
    bool last_build_ploc = false;
    int opt_max_levels = VB_TRAVERSAL_STACK_ENTRIES;    // build: deepest wide tree accepted (never more than the traversal stack holds)
    int last_build_levels = 0;               // wide levels of the last tree built
    int opt_two_level = 0;                   // build: 1 = one bottom-level tree per unique mesh + a top-level tree over the instances
    int opt_collapse = 1;                    // build: 1 = cost-optimal collapse (dynamic programme over the binary tree), 0 = greedy largest-area opening
    int opt_split_pct = 0;                   // build: triangle pre-splitting budget, extra references in percent of the triangle count (vb_bvh_split.cuh)
    int opt_size_classes = 1;                // build: triangle size class as the top bits of the Morton sort key
    int opt_sort_rays = 1, opt_sort_min_rays = 1 << 16;
    int opt_sort_skip_bits = 8;              // ray sort: low key bits the radix sort ignores (8 = two 8-bit passes instead of three, coarser order)
    int opt_sort_oct_shift = 8;              // ray sort key: direction octant above this many low Morton bits (vb_ray_sort_key)
    int opt_pt_schedule = 0;                 // 0 = wavefront (default), 1 = megakernel
    int opt_pt_sort_material = 1;            // wavefront: order shading entries by material
    int opt_pt_spec_cache = 1;               // hero spectrum: constant colour slots decoded once per scene (k_spec_cache)
    int opt_pt_split_principled = 1;         // wavefront: principled-BSDF entries shaded by their own launch, the rest by the lean kernel
    int opt_pt_count_work = 0;               // wavefront: instrumented traversal (nodes / triangles per ray), never timed
    int64_t opt_pt_max_paths = 1 << 26;      // wavefront: paths in flight per batch
    void *sort_buf = nullptr; size_t sort_bytes = 0;
    // host-buffer calls of vb_trace_*: ray ranges pipelined over copy-in / compute / copy-out streams (vb_trace.cu run_trace)
    int opt_host_chunks = 4, opt_host_chunk_min_rays = 1 << 20;
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    cudaEvent_t ev_in[VB_MAX_HOST_CHUNKS] = {}, ev_out[VB_MAX_HOST_CHUNKS] = {};
    int ensure_copy_streams();
    // record-form host calls (vb_trace_*_records): two alternating staging sets, "done" event of the call that last used each
    void *rec_stage[2] = {nullptr, nullptr}; size_t rec_bytes[2] = {0, 0}; cudaEvent_t rec_done[2] = {nullptr, nullptr}, rec_traced[2] = {nullptr, nullptr};
    unsigned rec_call = 0; bool rec_pending = false;
    int opt_async_chunks = 2;                // ray ranges per host-buffer record call when trace.async = 1
    int opt_trace_async = 0;                 // 1: host-buffer record calls return once queued; vb_ctx_synchronize completes them
    vb_pt_state *pt = nullptr;

    int fail(int code, const std::string &msg) { err = msg; return code; }
    int ensure_stage(size_t bytes);
    // profiling helpers: bracket kernels with events on the launching stream
    cudaEvent_t get_event();
    void prof_begin(const char *name, uint64_t launches);
    void prof_end();
    void prof_resolve();
    const char *cur_name = nullptr; cudaEvent_t cur_start = nullptr; uint64_t cur_launches = 0;
};

int vb_accel_build_impl(vb_ctx *ctx, vb_build_stats *out);
int vb_trace_queue(vb_ctx *ctx, const float4 *rays, uint64_t *keyidx, uint64_t *keyidx_alt, uint32_t n, uint4 *hits4, uint8_t *occ,
                   unsigned long long *work = nullptr, bool coherent = false);
void vb_pt_destroy(vb_ctx *ctx);

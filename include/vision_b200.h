/* vision_b200.h — C ABI of libvision_b200.so (sm_100a CUDA behind it).
 *
 * This boundary replaces the reference's compute/RHI layer for the megakernel path-tracing hot
 * path: the (absent, un-vendored) ocarina DSL codegen + OptiX dispatch that Vision reaches through
 *   Device::create_mesh / RHIMesh::build_bvh / Accel::add_instance / Accel::build_bvh
 *                                   reference src/base/mgr/geometry.cpp:36-53
 *   Accel::trace_closest / trace_occlusion      reference src/base/mgr/geometry.cpp:168-174
 *   Shader dispatch of "rt_geom", "path_tracing", "RGBFilm-accumulation", "RGBFilm-tone_mapping"
 *                                   reference src/base/sensor/frame_buffer.cpp:156-219,117-144
 *                                   reference src/base/integral/integrator.cpp:82-107
 *   Managed<T>::upload_immediately / download_immediately
 *                                   reference src/base/mgr/geometry.cpp:64-71, frame_buffer.cpp:424-427
 *
 * Conventions: every function returns int status (VB_OK = 0). vb_last_error(ctx) returns a
 * NUL-terminated message owned by the context (valid until the next call on that context).
 * One context per GPU. Calls on one context are not re-entrant; different contexts may be driven
 * from different host threads. Pointers are caller-owned; each pointer argument that may be either
 * host or device memory travels with a vb_mem tag. No C++ or torch types cross this boundary.
 */
#ifndef VISION_B200_H
#define VISION_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define VB_API __attribute__((visibility("default")))

typedef struct vb_ctx vb_ctx;

enum { VB_OK = 0, VB_ERR_INVALID = 1, VB_ERR_CUDA = 2, VB_ERR_STATE = 3, VB_ERR_UNSUPPORTED = 4 };
typedef enum { VB_MEM_HOST = 0, VB_MEM_DEVICE = 1 } vb_mem;

#define VB_INVALID_ID 0xffffffffu

/* ---- lifecycle ------------------------------------------------------------------------- */
VB_API int vb_ctx_create(int device, vb_ctx **out);
VB_API int vb_ctx_destroy(vb_ctx *ctx);
VB_API const char *vb_last_error(vb_ctx *ctx);           /* ctx may be NULL: last create error */
/* run all work of this context on an existing cudaStream_t (e.g. torch's current stream), so
 * the caller's CUDA events bracket it; NULL restores the context's own stream */
VB_API int vb_ctx_set_stream(vb_ctx *ctx, void *cuda_stream);
VB_API int vb_ctx_synchronize(vb_ctx *ctx);
/* tuning knobs (defaults are the tuned values): "trace.refill_thr", "trace.tri_thr", "trace.blocks_per_sm",
 * "trace.sort_rays", "trace.sort_min_rays", "trace.sort_oct_shift", "trace.host_chunks" (host-buffer calls are cut into
 * this many ray ranges, default 4, and copy-in / trace / copy-out of consecutive ranges overlap on three streams),
 * "trace.host_chunk_min_rays", "build.max_leaf" (fixed at 1); path tracer: "pt.schedule" (0 = wavefront pipeline,
 * default; 1 = single megakernel - same results), "pt.max_paths" (wavefront paths in flight) */
VB_API int vb_ctx_set_option(vb_ctx *ctx, const char *key, double value);
/* device time (ms, CUDA events on the launching stream) and launch count of the kernels this
 * library launched since the last vb_profile_reset; name = "trace_closest" | "trace_any" |
 * "path_trace" (megakernel) | "pt_generate" | "pt_shade" | "pt_shadow_resolve" | "pt_film" (wavefront) |
 * "bvh_build" | "film" | "all" */
VB_API int vb_profile_reset(vb_ctx *ctx);
/* diagnostics: read bandwidth (GB/s) of `reps` grid-stride 128-bit read passes over a zero-filled buffer of `bytes`
 * after 3 warm passes - L2 read bandwidth when the buffer fits L2 (64 MB), HBM read bandwidth when it is much larger.
 * The denominators of the path tracer's roofline ("builder measures with a 64 MB ... kernel") */
VB_API int vb_probe_read_bandwidth(vb_ctx *ctx, uint64_t bytes, int reps, double *gbs);
VB_API int vb_profile_get(vb_ctx *ctx, const char *name, double *ms, uint64_t *launches);

/* ---- geometry: replaces Geometry::upload + Geometry::build_accel -------------------------- */
/* Mesh = the reference's vision::Mesh (src/base/shape.h:78-112): vertices (pos, normal, uv) and
 * Triangle{i,j,k}. Triangle order defines prim_id. nrm3 / uv2 may be NULL (zeros). */
VB_API int vb_mesh_upload(vb_ctx *ctx, const float *pos3, const float *nrm3, const float *uv2, uint32_t nv,
                          const uint32_t *tri3, uint32_t nt, uint32_t *mesh_id);
/* non-indexed triangle soup already resident on host or device: triangle i = vertices 3i..3i+2 */
VB_API int vb_mesh_upload_soup(vb_ctx *ctx, const float *pos9, vb_mem mem, uint64_t nt, uint32_t *mesh_id);
/* Releases every mesh, the instance list and the acceleration structure of the context (mesh ids start at 0 again): what
 * Geometry::clear / a scene reload does in the reference (Pipeline::clear_geometry, src/base/mgr/pipeline.cpp:284-288). Without it a reused
 * context keeps every scene's meshes until vb_ctx_destroy. */
VB_API int vb_meshes_clear(vb_ctx *ctx);

/* InstanceData of the reference (src/base/shape.h:21-33): encoded ids + object-to-world.
 * o2w is column-major 4x4 (m[3] = translation), as the reference's float4x4. Array index = inst_id. */
typedef struct vb_instance {
    uint32_t mesh_id;
    uint32_t material_id;      /* index into vb_materials_set table, VB_INVALID_ID = none */
    uint32_t light_id;         /* index into vb_lights_set table,    VB_INVALID_ID = none */
    uint32_t inside_medium;    /* VB_INVALID_ID = none */
    uint32_t outside_medium;
    uint32_t lightmap_id;      /* carried, unused by the PT path */
    float o2w[16];
} vb_instance;
VB_API int vb_instances_set(vb_ctx *ctx, const vb_instance *inst, uint32_t n);

#define VB_BUILD_TWO_LEVEL 1u      /* this build only: as if option "build.two_level" were 1 */
typedef struct vb_build_opts {
    uint32_t flags;            /* 0 or VB_BUILD_* bits */
} vb_build_opts;
typedef struct vb_build_stats {
    uint64_t n_triangles;      /* flattened world-space triangles */
    uint64_t n_nodes;          /* wide nodes: 80-byte payload in 96-byte, 32-byte-aligned slots */
    uint64_t node_bytes, tri_bytes;
    double build_ms;           /* device time of the whole build */
    float scene_lo[3], scene_hi[3];
    uint64_t n_references;     /* leaves of the BVH: n_triangles plus the extra references of triangle pre-splitting (option
                                * "build.split_pct"); tri_bytes = 48 * n_references, and vb_accel_download's `tris` holds that many */
} vb_build_stats;
/* Builds the compressed 8-wide BVH on the GPU (replaces mesh.build_bvh / accel_.add_instance / accel_.build_bvh,
 * reference src/base/mgr/geometry.cpp:36-53). Default: every instance flattened to world space, ONE tree. Option "build.two_level" = 1
 * (vb_ctx_set_option): one bottom-level tree per unique mesh in object space + a top-level tree over the instances - the reference's
 * GAS / IAS shape; memory and build time follow the unique triangles, hit records are bit-identical to the flattened build's, traversal is
 * 4-5x slower. Then n_triangles / n_references count the STORED (unique) triangles. Other build options: "build.ploc" (-1 automatic, 0 radix
 * tree, 1 PLOC), "build.collapse" (1 cost-optimal 8-wide collapse, 0 greedy), "build.size_classes", "build.split_pct". */
VB_API int vb_accel_build(vb_ctx *ctx, const vb_build_opts *opts, vb_build_stats *out);
/* copies the built structure to host (test/diagnostic surface): nodes = n_nodes*96 bytes,
 * tris = n_triangles*48 bytes (float4 v0|prim, e1|inst, e2|0). Either pointer may be NULL. */
VB_API int vb_accel_download(vb_ctx *ctx, void *nodes, void *tris);

/* ---- batch trace: replaces Accel::trace_closest / trace_occlusion -------------------------- */
/* SoA rays (reference Ray = {org_min, dir_max}: origin + t_min, direction + t_max) */
typedef struct vb_rays_soa {
    const float *ox, *oy, *oz, *tmin, *dx, *dy, *dz, *tmax;
    vb_mem mem;
} vb_rays_soa;
/* SoA hits (reference TriangleHit{inst_id, prim_id, bary} + distance). Miss: inst = prim =
 * VB_INVALID_ID, u = v = t = 0. bary: hit point = (1-u-v)*p0 + u*p1 + v*p2 */
typedef struct vb_hits_soa {
    uint32_t *inst, *prim;
    float *u, *v, *t;
    vb_mem mem;
} vb_hits_soa;
/* Completion: with HOST buffers (input or output) the call returns when the results are in the caller's memory. With DEVICE buffers on
 * both sides it is stream-ordered on the context's stream (vb_ctx_set_stream) and returns once queued: read the results on that stream,
 * or after vb_ctx_synchronize - a consumer on another stream races with the traversal kernel. */
VB_API int vb_trace_closest(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, const vb_hits_soa *out);
VB_API int vb_trace_any(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, uint8_t *occluded, vb_mem out_mem);
/* vb_trace_closest with NO allocation. This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it). Same answers as vb_trace_closest, same in-call ray sort; its scratch memory (the persistent kernel's ray counter, sort keys,
 * permutation, AoS ray copy, radix sort storage) is a file-scope `static __device__` array of VB_TRACE_STATIC_SCRATCH_BYTES
 * (about 13 MiB) that every GPU which loads the library holds until exit. Calls no cudaMalloc / cudaFree or allocating library
 * routine. Limits: rays and hits must be DEVICE buffers (VB_ERR_INVALID otherwise), n <= VB_TRACE_STATIC_MAX_RAYS per call
 * (VB_ERR_INVALID with a message naming the limit; split larger batches). Every call on a device shares the one array: it is
 * re-initialised at the start of each call on the context's stream, and calls that OVERLAP on the same GPU (two contexts or streams
 * of one device) race - serialise them. */
#define VB_TRACE_STATIC_MAX_RAYS (1u << 18)
#define VB_TRACE_STATIC_SCRATCH_BYTES (256u + 4u * 4u * VB_TRACE_STATIC_MAX_RAYS + 32u * VB_TRACE_STATIC_MAX_RAYS + (1u << 20))
VB_API int vb_trace_closest_static(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, const vb_hits_soa *out);
/* Record form: the reference's own data shapes at this boundary. vb_ray = ocarina Ray {org_min, dir_max} (32 bytes: geometry.h:64-69,
 * frame_buffer.h:88-91), vb_hit = TriangleHit {inst_id, prim_id, bary} (16 bytes: geometry.cpp:81-85,215-219); no distance, as in the
 * reference (compute_surface_interaction re-derives it). Same answers as the SoA form. Host buffers (mem = VB_MEM_HOST, ideally
 * pinned) move as one copy per range each way, pipelined over copy-in / compute / copy-out streams. With option "trace.async" = 1
 * a host-buffer call returns once queued - consecutive calls overlap - and results are valid after vb_ctx_synchronize. */
typedef struct vb_ray { float ox, oy, oz, tmin, dx, dy, dz, tmax; } vb_ray;
typedef struct vb_hit { uint32_t inst, prim; float u, v; } vb_hit;
VB_API int vb_trace_closest_records(vb_ctx *ctx, const vb_ray *rays, uint64_t n, vb_hit *hits, vb_mem mem);
VB_API int vb_trace_any_records(vb_ctx *ctx, const vb_ray *rays, uint64_t n, uint8_t *occluded, vb_mem mem);
/* instrumented variants (never timed): per-call totals of wide nodes popped and triangles tested */
VB_API int vb_trace_closest_counted(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, const vb_hits_soa *out,
                                    uint64_t *nodes_popped, uint64_t *tris_tested);
VB_API int vb_trace_any_counted(vb_ctx *ctx, const vb_rays_soa *rays, uint64_t n, uint8_t *occluded, vb_mem out_mem,
                                uint64_t *nodes_popped, uint64_t *tris_tested);

#include "vision_b200_pt.h"

#ifdef __cplusplus
}
#endif
#endif

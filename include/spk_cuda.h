/*
 * spk_cuda.h — C ABI of the B200-native `cuda` path-tracing pipeline for Sparkle.
 *
 * This is the drop-in boundary. The reference has no FFI: its renderers are C++
 * classes behind `Renderer::CreateRenderer` (libraries/source/renderer/renderer/Renderer.cpp:17-53).
 * A `CudaRenderer : Renderer` (see sparkle_b200/host/cuda_renderer.h and INTEGRATION.md) binds exactly the
 * entry points below. Plain pointers and sizes only; no C++ / torch types cross this line.
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error (SpkStatus); spk_cuda_last_error() gives text.
 *     Mirrors the reference's "bool/nullptr + Log(Error)" convention (AppFramework.cpp:200-205); nothing
 *     throws across the boundary.
 *   - the context owns all device memory; the caller owns all host memory. Host pointers in
 *     SpkSceneDesc are borrowed for the duration of the call only.
 *   - one context == one CUDA device == one stream. Multi-GPU = one context per GPU (in one process or one
 *     process each), spp interleaved by (first_sample, sample_stride), accumulators combined and tone-mapped by
 *     spk_cuda_group_resolve over NVLink peer memory; spk_cuda_accum_device_ptr() stays for callers
 *     that bring their own collective (one NCCL reduce).
 *   - there is NO CPU fallback: without a CUDA device spk_cuda_create fails with SPK_ERR_NO_DEVICE.
 */
#ifndef SPK_CUDA_H_
#define SPK_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPK_ABI_VERSION 1u
#define SPK_INVALID_ID 0xFFFFFFFFu

typedef enum SpkStatus {
    SPK_OK = 0,
    SPK_ERR_INVALID_ARGUMENT = -1,
    SPK_ERR_NO_DEVICE = -2,
    SPK_ERR_CUDA = -3,
    SPK_ERR_OUT_OF_MEMORY = -4,
    SPK_ERR_NO_SCENE = -5,
    SPK_ERR_UNSUPPORTED = -6
} SpkStatus;

/* ---- acceleration structures ------------------------------------------------------------------
 * Node layout is that of bvh::v2::Node<float,3> (madmann91/bvh @5a9d759c, the library behind
 * core/math/BVH.h:18-19): bounds = {min.x,max.x,min.y,max.y,min.z,max.z}, index = first_id<<4 |
 * prim_count; prim_count==0 marks an inner node whose children are nodes[first_id], nodes[first_id+1];
 * node 0 is the root. `node_stride` is 28 (the raw library struct) or 32 (padded). `prim_ids` maps
 * leaf slots to primitive indices (bvh_.prim_ids; 8-byte size_t in the library, 4-byte also accepted).
 * nodes == NULL with num_nodes == 0: the tree is BUILT ON THE DEVICE by spk_cuda_upload_scene (an LBVH: Morton sort,
 * Karras hierarchy, bottom-up fit, one triangle / primitive per leaf by default; csrc/spk_build.cu) and can be read back with
 * spk_cuda_download_bvh.
 */
typedef struct SpkBvhNode {
    float bounds[6];
    uint32_t index;
    uint32_t pad;
} SpkBvhNode;

typedef struct SpkBvhDesc {
    const void *nodes;
    uint32_t num_nodes;
    uint32_t node_stride;   /* 28 or 32 bytes */
    const void *prim_ids;
    uint32_t num_prims;
    uint32_t prim_id_bytes; /* 4 or 8 */
} SpkBvhDesc;

/* One per unique io/Mesh.h:7-141 `Mesh` (== one BLAS, MeshRenderProxy.cpp:24-199). Instances share it. */
typedef struct SpkMeshDesc {
    const float *vertices;   /* 3*num_vertices, Mesh::vertices */
    const float *normals;    /* 3*num_vertices, Mesh::normals (required) */
    const float *tangents;   /* 4*num_vertices, Mesh::tangents (required) */
    const float *uvs;        /* 2*num_vertices or NULL (Mesh::uvs empty => uv = 0) */
    const uint32_t *indices; /* 3*num_faces */
    uint32_t num_vertices;
    uint32_t num_faces;
    SpkBvhDesc blas;         /* over faces */
} SpkMeshDesc;

typedef enum SpkPrimitiveKind { SPK_PRIM_MESH = 0, SPK_PRIM_SPHERE = 1 } SpkPrimitiveKind;

/* One per PrimitiveRenderProxy, in SceneRenderProxy::primitives_ order (== primitive index reported
 * in SpkHit.prim). fwd/inv are rows 0..2 of Transform::GetMatrix() and of its inverse, bit for bit
 * (core/math/Transform.h:74-91). Spheres use translation (fwd[3],fwd[7],fwd[11]) and radius2
 * (SphereRenderProxy.h:33-36). */
typedef struct SpkPrimitiveDesc {
    uint32_t kind;     /* SpkPrimitiveKind */
    uint32_t mesh;     /* index into SpkSceneDesc.meshes (mesh kind) */
    uint32_t material; /* index into SpkSceneDesc.materials */
    float radius2;     /* sphere: scaled_radius_2_ */
    float fwd[12];
    float inv[12];
} SpkPrimitiveDesc;

typedef enum SpkMaterialType { SPK_MAT_PBR = 0, SPK_MAT_DIELECTRIC = 1 } SpkMaterialType;

/* io/Material.h:7-22 MaterialResource + Material::Type; 64 bytes like MaterialRenderData
 * (MaterialRenderProxy.h:39-58). Texture ids index SpkSceneDesc.textures, SPK_INVALID_ID = none. */
typedef struct SpkMaterialDesc {
    uint32_t base_color_tex;
    uint32_t emissive_tex;
    uint32_t metallic_roughness_tex;
    uint32_t normal_tex;
    float metallic;
    float roughness;
    float eta;
    uint32_t type; /* SpkMaterialType */
    float base_color[3];
    float pad0;
    float emissive[3];
    float pad1;
} SpkMaterialDesc;

typedef enum SpkPixelFormat {
    SPK_FMT_RGBA8_UNORM = 0,
    SPK_FMT_RGBA8_SRGB = 1,
    SPK_FMT_RGBA16F = 2,
    SPK_FMT_RGBA32F = 3
} SpkPixelFormat;

typedef struct SpkTextureDesc {
    const void *pixels; /* row-major, tightly packed */
    uint32_t width;
    uint32_t height;
    uint32_t format; /* SpkPixelFormat */
    uint32_t pad;
} SpkTextureDesc;

/* SkyRenderProxy.cpp:50-64. has_sky==0: no sky light (miss adds nothing). has_map: cube faces in
 * Image2DCube::FaceId order (+X,-X,+Y,-Y,+Z,-Z). */
typedef struct SpkSkyDesc {
    uint32_t has_sky;
    uint32_t has_map;
    float color[3];
    float pad;
    SpkTextureDesc faces[6];
} SpkSkyDesc;

typedef struct SpkSceneDesc {
    uint32_t abi_version;
    uint32_t num_meshes;
    const SpkMeshDesc *meshes;
    uint32_t num_primitives;
    uint32_t num_materials;
    const SpkPrimitiveDesc *primitives;
    const SpkMaterialDesc *materials;
    uint32_t num_textures;
    uint32_t pad;
    const SpkTextureDesc *textures;
    SpkBvhDesc tlas; /* over primitives (world AABBs), SceneRenderProxy.cpp:31-69 */
    SpkSkyDesc sky;
    float world_extent; /* max |coordinate| of the world bounds; sizes the t-culling slack */
    float pad2;
} SpkSceneDesc;

/* CameraRenderProxy::{Posture,FocusPlane,Attribute} (CameraRenderProxy.h:31-57). */
typedef struct SpkCamera {
    float position[3];
    float up[3];
    float front[3];
    float right[3];
    float max_u[3];
    float max_v[3];
    float lower_left[3];
    float aperture_radius;
    float exposure;
} SpkCamera;

typedef enum SpkAccumMode {
    SPK_ACCUM_SUM = 0,            /* acc += min(L,limit); resolve divides by the sample count */
    SPK_ACCUM_RUNNING_AVERAGE = 1 /* acc = lerp(new, acc, N/(N+1)) per sample, CPURenderer.cpp:530-544 */
} SpkAccumMode;

enum {
    SPK_FLAG_EXACT_TRAVERSAL = 1u, /* never shorten tmax while traversing (reference behaviour) */
    SPK_FLAG_COUNTERS = 2u,        /* instrumented kernel: fill node/triangle counters in SpkStats */
    SPK_FLAG_WAVEFRONT = 4u,       /* wavefront (queue) tracer instead of the persistent megakernel */
    SPK_FLAG_THREAD_PER_RAY = 8u,  /* ray batches: plain one-thread-per-ray kernel instead of the persistent warp-voting one
                                      (kept for A/B measurement; same results) */
    SPK_FLAG_SPATIAL_DENOISE = 32u, /* RenderConfig::spatial_denoise (CPURenderer.cpp:452-526): every sample is rendered
                                      into a g-buffer (colour + valid flag, first-hit normal), invalid pixels take the
                                      colour of a valid non-sky 8-neighbour with normal similarity >= 0.99, then the
                                      sample is accumulated. One sample per pixel per pass like the reference; megakernel */
    SPK_FLAG_AUTO_TRACER = 16u     /* megakernel or wavefront, whichever spk_cuda_choose_tracer measured faster on the
                                      current scene (measured on first use if not yet); SPK_ACCUM_SUM only, else megakernel */
};

typedef struct SpkRenderParams {
    uint32_t first_sample;  /* frame_seed of the first sample, CPURenderer.cpp:432 */
    uint32_t sample_count;  /* samples per pixel added by this call */
    uint32_t sample_stride; /* 1; or G when spp is interleaved over G GPUs (this rank: first_sample=g) */
    uint32_t max_bounce;    /* RenderConfig::max_bounce */
    float output_limit;     /* CameraRenderProxy::OutputLimit (6) */
    uint32_t accum_mode;    /* SpkAccumMode */
    uint32_t flags;
    uint32_t cumulated_cap; /* running average only: the moving-average weight uses min(cumulated, cap) samples, the
                               camera's cumulated count capped at max_spp (CameraRenderProxy.cpp:70-73): past the cap
                               the reference's frame buffer becomes an exponential moving average. 0 = no cap */
} SpkRenderParams;

/* 32-byte ray, 20-byte hit (IntersectionCandidate, Intersection.h:91-104). */
typedef struct SpkRay {
    float origin[3];
    float pad0;
    float direction[3];
    float pad1;
} SpkRay;

typedef struct SpkHit {
    float t; /* world t; FLT_MAX when nothing was hit */
    float u;
    float v;
    uint32_t prim; /* primitive index (SceneRenderProxy order) or SPK_INVALID_ID */
    uint32_t face; /* face index inside the mesh, SPK_INVALID_ID for spheres */
} SpkHit;

typedef struct SpkStats {
    uint64_t paths;       /* pixel samples started */
    uint64_t rays;        /* SceneRenderProxy::Intersect calls (path segments) */
    uint64_t node_visits; /* inner-node visits (each reads one 64-byte child pair), TLAS + BLAS */
    uint64_t tri_tests;   /* triangle records fetched and tested */
    uint64_t inst_visits; /* mesh-instance entries (ray transformed into object space) */
    uint64_t sphere_tests;
    double last_ms; /* device time of the last render / trace call (CUDA events on the context stream) */
    uint64_t kernel_launches; /* kernels launched by this context so far */
} SpkStats;

typedef enum SpkResolveFormat {
    SPK_RESOLVE_LINEAR_F32 = 0,  /* W*H float4: mean radiance rgb (pre-tonemap), w = mean valid flag; row j = scene row j */
    SPK_RESOLVE_TONEMAP_F16 = 1, /* W*H half4, ACES, rows flipped: CPURenderer.cpp:554-576 output_image_ */
    SPK_RESOLVE_TONEMAP_SRGB8 = 2 /* W*H RGBA8, ACES then pow(0.45) (Utilities.h:411-414), rows flipped */
} SpkResolveFormat;

typedef struct SpkCreateInfo {
    uint32_t abi_version;
    int32_t device; /* CUDA device ordinal */
    uint32_t width; /* scene resolution */
    uint32_t height;
} SpkCreateInfo;

typedef struct SpkContext SpkContext;

const char *spk_cuda_last_error(void);
uint32_t spk_cuda_abi_version(void);

/* Renderer::CreateRenderer + CudaRenderer::InitRenderResources (Renderer.cpp:17-53). */
int spk_cuda_create(const SpkCreateInfo *info, SpkContext **out);
int spk_cuda_destroy(SpkContext *ctx);
/* Page-locked host memory for frame buffers and ray batches: spk_cuda_resolve / spk_cuda_trace_closest copy to and from
 * such buffers with one DMA at PCIe speed; any other host pointer works too and is staged through a pinned bounce
 * buffer inside the call. (The reference's readback buffer is an RHI staging buffer, Renderer.cpp:108-180.) */
int spk_cuda_host_alloc(void **out, size_t bytes);
int spk_cuda_host_free(void *p);
/* Use an externally owned cudaStream_t (e.g. torch's current stream) instead of the context's own. */
int spk_cuda_set_stream(SpkContext *ctx, void *cuda_stream);

/* SceneRenderProxy::Update/UpdateBVH consumer (SceneRenderProxy.cpp:195-297): flattens TLAS/BLAS, triangle,
 * attribute, material and texture tables into aligned device buffers. Replaces any previous scene. */
int spk_cuda_upload_scene(SpkContext *ctx, const SpkSceneDesc *scene);
/* PrimitiveChangeType::Update (SceneRenderProxy.cpp:212-230): new transforms + rebuilt TLAS, BLAS kept. tlas == NULL (or
 * a descriptor with nodes == NULL): the TLAS is rebuilt on the device from the new transforms. */
int spk_cuda_update_instances(SpkContext *ctx, const SpkPrimitiveDesc *primitives, uint32_t num_primitives,
                              const SpkBvhDesc *tlas);
/* Any tree of the uploaded scene (mesh index, or SPK_BVH_TLAS) back in the bvh::v2 node layout + prim_ids, so that a CPU
 * consumer (the oracle, the reference's traversal) walks the topology the device walks. Call with NULL arrays for sizes. */
#define SPK_BVH_TLAS 0xFFFFFFFFu
int spk_cuda_download_bvh(SpkContext *ctx, uint32_t mesh, SpkBvhNode *nodes_out, uint32_t node_capacity,
                          uint32_t *num_nodes_out, uint32_t *prim_ids_out, uint32_t prim_capacity, uint32_t *num_prims_out);
typedef struct SpkBuildInfo {
    uint32_t device_built; /* 1: built by spk_build.cu, 0: handed in by the caller */
    uint32_t num_nodes;
    uint32_t num_prims;
    uint32_t depth;   /* device-built trees only */
    double sah_cost;  /* device-built trees only: (sum of inner areas + sum of leaf area * count) / root area */
    double build_ms;  /* device time of the build (CUDA events) */
} SpkBuildInfo;
int spk_cuda_get_build_info(SpkContext *ctx, uint32_t mesh, SpkBuildInfo *out);
/* CameraRenderProxy::OnTransformDirty (CameraRenderProxy.cpp:113-136). */
int spk_cuda_set_camera(SpkContext *ctx, const SpkCamera *camera);
/* CPURenderer::Render's NeedClear branch (CPURenderer.cpp:128-137). */
int spk_cuda_reset_accumulation(SpkContext *ctx);
/* CPURenderer::BasePass + temporal accumulate (CPURenderer.cpp:419-450, 530-544) for sample_count samples. */
int spk_cuda_render(SpkContext *ctx, const SpkRenderParams *params);
/* Times 8 samples per pixel of the current scene and camera with the megakernel and with the wavefront tracer
 * (CUDA events; the megakernel in its phased and its two material-sorted forms) and remembers the fastest for SPK_FLAG_AUTO_TRACER until the next spk_cuda_upload_scene. The
 * accumulator must be empty (it is used as scratch and cleared again). *wavefront_out: 0 megakernel, 1 wavefront;
 * ms_out[2]: the two times (may be NULL). */
int spk_cuda_choose_tracer(SpkContext *ctx, uint32_t max_bounce, uint32_t *wavefront_out, float *ms_out);
/* What was chosen: *wavefront_out -1 not measured yet / 0 megakernel / 1 wavefront; *megakernel_form_out the form of the
 * megakernel in use (831 phased, 1162 / 1084 sorted by material class with 16x2 / 8x4 warps x CTAs per SM). */
int spk_cuda_tracer_choice(SpkContext *ctx, int32_t *wavefront_out, uint32_t *megakernel_form_out);
/* ToneMappingPass + readback (CPURenderer.cpp:554-576; Renderer.cpp:108-180). `total_samples` is the
 * number of samples summed in the accumulator (all ranks, after the reduce) for SPK_ACCUM_SUM. */
int spk_cuda_resolve(SpkContext *ctx, uint32_t format, uint32_t total_samples, void *host_out, size_t host_bytes);
/* Device-resident accumulator (W*H float4) for the NCCL reduce; samples accumulated locally so far. */
int spk_cuda_accum_device_ptr(SpkContext *ctx, void **device_ptr, size_t *bytes);
int spk_cuda_accum_sample_count(SpkContext *ctx, uint32_t *count);

/* Parity probe: the device's sinf (op 0) / cosf (1) / powf (2, with y) on host arrays. The reference calls std::sin, std::cos
 * and std::pow on floats, i.e. glibc's float routines (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, e_powf.c); the device
 * restates them (csrc/spk_glibc_math.cuh) and this entry point lets a test compare the two argument by argument. */
int spk_cuda_probe_libm(SpkContext *ctx, uint32_t op, const float *x, const float *y, uint64_t n, float *out);

/* SceneRenderProxy::Intersect<false> on a ray batch (SceneRenderProxy.cpp:82-126, 299-312), host buffers:
 * H2D of rays, traversal, D2H of hits inside the call. */
int spk_cuda_trace_closest(SpkContext *ctx, const SpkRay *rays, uint64_t num_rays, SpkHit *hits, uint32_t flags);
/* Same with device-resident buffers (no copies); asynchronous on the context stream. */
int spk_cuda_trace_closest_device(SpkContext *ctx, const void *d_rays, uint64_t num_rays, void *d_hits,
                                  uint32_t flags);
/* spk_cuda_trace_closest for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). NO counterpart in the reference.
 * Same host buffers, same hits bit for bit, but the staged rays, the hits and the kernel's ray ticket live in ONE static
 * device array of the library (52 B x SPK_NOALLOC_MAX_RAYS + 256 B = 13.0 MiB, held by every GPU that loads the library
 * until the process exits); the call makes no cudaMalloc / cudaFree / allocating library call. num_rays beyond
 * SPK_NOALLOC_MAX_RAYS is refused with SPK_ERR_INVALID_ARGUMENT (split the batch, or use spk_cuda_trace_closest).
 * `flags`: 0 or SPK_FLAG_EXACT_TRAVERSAL. Synchronous like spk_cuda_trace_closest. Every call on a GPU shares the array
 * (whatever the context): calls that overlap on one GPU - two contexts on different streams or threads - race; serialise
 * them. The array keeps the previous call's contents; what must start at zero (the ticket) is cleared by each call. */
#define SPK_NOALLOC_MAX_RAYS 262144u
int spk_cuda_trace_closest_noalloc(SpkContext *ctx, const SpkRay *rays, uint64_t num_rays, SpkHit *hits,
                                   uint32_t flags);

/* ---- multi-GPU behind the boundary (north_star (d)) ---------------------------------------------------
 * The scene is replicated, rank g of G renders samples {g, g+G, ...} (first_sample = g, sample_stride = G) into its own
 * accumulator, and ONE kernel per GPU combines them: each GPU sums its H/G rows of all G accumulators over NVLink peer
 * memory, tone-maps them (ToneMappingPass, CPURenderer.cpp:554-576) and stores the packed pixels into the root's output
 * buffer; the root copies the frame to `host_out`. Reduce-scatter + resolve + gather in one pass (csrc/spk_group.cu).
 *   local group  all contexts live in this process (CudaRenderer with the `gpus` cvar): one call drives every GPU.
 *   IPC group    one process per GPU (torchrun): every rank exports a handle, the host exchanges them (any transport),
 *                every rank creates its group object and makes the SAME resolve calls; rank 0 is the root and the only
 *                one whose host_out is written (others pass NULL). One rank per GPU. A NULL host_out on the root leaves
 *                the frame in the root's device buffer and does not block. */
#define SPK_PEER_HANDLE_BYTES 192
typedef struct SpkPeerHandle {
    uint8_t bytes[SPK_PEER_HANDLE_BYTES];
} SpkPeerHandle;
typedef struct SpkGroup SpkGroup;
/* Call once per group, before the handles are exchanged (it also clears the flags an earlier group of this context left). */
int spk_cuda_group_export(SpkContext *ctx, SpkPeerHandle *out);
int spk_cuda_group_create(SpkContext *ctx, uint32_t rank, uint32_t world, const SpkPeerHandle *handles, SpkGroup **out);
int spk_cuda_group_create_local(SpkContext *const *ctxs, uint32_t n, SpkGroup **out);
/* `total_samples`: samples summed over all ranks (SPK_ACCUM_SUM divides by it). */
int spk_cuda_group_resolve(SpkGroup *group, uint32_t format, uint32_t total_samples, void *host_out, size_t host_bytes);
/* IPC groups: by default ranks > 0 return from spk_cuda_group_resolve when their kernel has finished (so a rank that
 * never joined is reported to everyone); enabled = 1 makes them return at once, like the root with host_out == NULL. */
int spk_cuda_group_set_async(SpkGroup *group, uint32_t enabled);
/* Destroy a group BEFORE the contexts it was made from. */
int spk_cuda_group_destroy(SpkGroup *group);

/* Tuning knobs, explicit (nothing in the library reads the environment). Defaults are the measured choices. */
typedef enum SpkOption {
    SPK_OPT_MEGAKERNEL_FORM = 0, /* 0 = measured choice (spk_cuda_choose_tracer; 1084 until measured). Forms: 1162 / 1084 /
                                    1083 sorted megakernel, 16 warps x 2 / 8 x 4 / 8 x 3 CTAs per SM, pixels owned by a
                                    CTA; 3162 / 3084 the same with the sort kept in pixel order inside a class; 2162 /
                                    2084 persistent pixel-pulling grid; 831 / 830 / 1621 / 461 phased kernel; 1 = plain
                                    per-pixel loop */
    SPK_OPT_WAVEFRONT_SLOTS = 1, /* path slots of the wavefront tracer (default 2^26, capped by the work of the call; 108 bytes each) */
    SPK_OPT_WAVEFRONT_CHUNK = 2, /* wavefront iterations queued per host read of the control block (default 8) */
    SPK_OPT_TRACE_VOTE = 3,      /* ray-batch kernel: lanes waiting in a leaf state that switch the warp to it (4, 6, 8, 12, 16; default 6) */
    SPK_OPT_SMEM_CARVEOUT = 4,   /* percent of the SM's shared-memory/L1 array asked for as shared memory; -1 driver
                                    default, -2 (default) the smallest that holds the resident CTAs */
    SPK_OPT_VERBOSE = 5,         /* 1: spk_cuda_choose_tracer prints its timings to stderr */
    SPK_OPT_BUILD_MAX_LEAF = 6,  /* triangles per leaf of device-built BLAS (1..15, default 1: fastest to trace) */
    SPK_OPT_COST_ORDERED_STRIPS = 8, /* sorted megakernel: launch the strips of tiles in the order of their cost in the
                                    previous launch of the same view, longest first (default 1) */
    SPK_OPT_WAVEFRONT_SORT = 9,  /* wavefront tracer: sort every ray queue by (Morton cell of the origin in the scene box,
                                  * direction octant) before it is traced; value = bits per axis of the cell, 0 = off */
    SPK_OPT_TRACE_SMEM_TOP = 7   /* ray-batch kernel: pairs of the hottest tree's top levels (the TLAS, else the first BLAS)
                                    staged in shared memory: 0 (default, measured), 64, 128, 256, 512 */
} SpkOption;
int spk_cuda_set_option(SpkContext *ctx, uint32_t option, int64_t value);

int spk_cuda_get_stats(SpkContext *ctx, SpkStats *out);
int spk_cuda_reset_stats(SpkContext *ctx);
int spk_cuda_synchronize(SpkContext *ctx);

#ifdef __cplusplus
}
#endif
#endif /* SPK_CUDA_H_ */

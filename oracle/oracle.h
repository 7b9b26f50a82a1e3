/*
 * oracle.h — C API of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it. The CUDA
 * product (sparkle_b200/) never links, imports or calls it and has no CPU fallback.
 *
 * Two libraries export this same API:
 *   oracle/liboracle.so            "port": plain C++ restatement of the reference CPU path tracer
 *                                  (oracle.cpp; every function cites the reference file:line it follows).
 *   oracle/_ref/libsparkle_ref.so  "reference": the reference's own headers compiled in place from
 *                                  /root/reference (ref_harness.cpp + Makefile), used to pin the port.
 * Both consume the SpkSceneDesc data contract of include/spk_cuda.h (data layout only, no product code).
 */
#ifndef SPK_ORACLE_H_
#define SPK_ORACLE_H_

#include "../include/spk_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum OrcTraceMode {
    ORC_TRACE_BVH_REFERENCE = 0, /* bvh::v2 traversal contract, tmax never shortened (SceneRenderProxy.cpp:95-113) */
    ORC_TRACE_BRUTE_FORCE = 1,   /* every primitive, every triangle, fixed order: topology-independent ground truth */
    ORC_TRACE_BVH_CULLED = 2     /* same traversal, boxes beyond the current best t skipped (context line only) */
} OrcTraceMode;

typedef struct OrcCounters {
    uint64_t paths;
    uint64_t rays;
    uint64_t node_visits;
    uint64_t tri_tests;
    uint64_t inst_visits;
    uint64_t sphere_tests;
    uint64_t rng_draws;
} OrcCounters;

/* which implementation answers: "port" or "reference" */
const char *orc_kind(void);

/* a1: Xoshiro128Plus(seed) state after seeding; first n RandomUnit() values (Sampler.h:16-41) */
void orc_rng_state(uint32_t seed, uint32_t state_out[4]);
void orc_rng_units(uint32_t seed, uint32_t n, float *out);

/* H4 association KATs: op 0 dot, 1 squaredNorm, 2 normalized(a).x/y/z -> out[0..2], 3 cross, 4 M3x3*a (m row-major 3x4),
 * 5 affine M*a. a,b are 3 floats, m is 12 floats. */
void orc_kat_vec(uint32_t op, const float *a, const float *b, const float *m, float *out);
/* libm used by the implementation: op 0 sin,1 cos,2 pow(x,y),3 sqrt */
float orc_kat_libm(uint32_t op, float x, float y);
/* a4: sampler::UnitDisk with the stream seeded by `seed` -> out[2] */
void orc_kat_unit_disk(uint32_t seed, float *out);
/* a11: Image2D::Sample of `tex` at uv -> out[3] */
void orc_kat_texture(const SpkTextureDesc *tex, float u, float v, float *out);
/* a14: SkyRenderProxy::Evaluate(direction) -> out[3] */
void orc_kat_sky(const SpkSkyDesc *sky, const float *direction, float *out);
/* a12/a13: MaterialRenderProxy::SampleSurface with the stream seeded by `seed`.
 * in: ray direction, normal, tangent, uv. out: next direction[3], throughput[3]; returns rng draws used */
uint32_t orc_kat_sample_surface(const SpkSceneDesc *scene, uint32_t material, uint32_t seed, const float *dir,
                                const float *normal, const float *tangent, const float *uv, float *out_wi,
                                float *out_throughput);

/* a17 + transforms ("reference" only: these pin the PRODUCT's host math, the port has no copy).
 * orc_kat_transform: SceneNode::SetTransform(translate, euler radians, scale) under the identity root -> rows 0..2 of
 * Transform::GetMatrix() and of GetInverse().GetMatrix() (Transform.cpp:52-64). orc_kat_camera: OrbitCameraComponent::Setup
 * parameters (degrees) + CameraComponent attribute -> CameraRenderProxy posture / focus plane (CameraRenderProxy.cpp:21-33,
 * 113-136) as an SpkCamera. */
int orc_kat_transform(const float *translate, const float *euler, const float *scale, float *fwd12, float *inv12);
int orc_kat_camera(const float *center, float radius, float pitch_deg, float yaw_deg, float focal_length,
                   float sensor_height, float aperture, float exposure, uint32_t width, uint32_t height, SpkCamera *out);
/* GLTFLoader::ProcessNode transform of a root-level node (GLTFLoader.cpp:532-544): translation, rotation xyzw */
int orc_kat_gltf_node(const float *translation, const float *rotation_xyzw, float *fwd12, float *inv12);
/* reference build only: Image2D::LoadFromFile's HDR branch (stbi_loadf, clamp, half; Image.cpp:152-181) on a file image in
 * memory, and the loop of SkyLightCookJob::Execute (SkyLight.cpp:103-118) at a given cube size */
int orc_kat_decode_hdr(const unsigned char *data, uint32_t size, uint32_t *width, uint32_t *height, uint16_t *rgba_half,
                       uint64_t capacity_halves);
int orc_kat_cook_sky_cube(const uint16_t *equirect_rgba_half, uint32_t width, uint32_t height, uint32_t size,
                          uint16_t *faces_out);
/* "port" only: 0 = glibc float sin/cos/pow as the reference calls them (default), 1 = what the CUDA path computes: the same
 * (csrc/spk_glibc_math.cuh), except Schlick's pow(x, 5.f) as an exact product rounded once (oracle.cpp g_libm_mode) */
int orc_set_libm(uint32_t mode);

/* a6-a9: closest hit of a ray batch. threads<=1 runs on the calling thread. */
int orc_trace_closest(const SpkSceneDesc *scene, const SpkRay *rays, uint64_t num_rays, SpkHit *hits, uint32_t mode,
                      uint32_t threads, OrcCounters *counters);

/* a10: full Intersection of one ray: out = {t, loc[3], normal[3], tangent[3], uv[2]} (12 floats), returns prim or INVALID */
uint32_t orc_intersect_full(const SpkSceneDesc *scene, const SpkRay *ray, float *out12);

/* a2-a5, a15 (temporal part): renders params->sample_count samples per pixel into accum (W*H float4, in/out).
 * *cumulated is the camera's cumulated sample count before the call and is advanced (in/out).
 * Rows are dispatched to `threads` workers like CPURenderer::BasePass. trace_mode as OrcTraceMode (0 = faithful). */
int orc_render(const SpkSceneDesc *scene, const SpkCamera *camera, uint32_t width, uint32_t height,
               const SpkRenderParams *params, float *accum, uint32_t *cumulated, uint32_t trace_mode, uint32_t threads,
               OrcCounters *counters);

/* one pixel sample, for debugging divergences: out = {r,g,b,valid, nx,ny,nz, bounces} (8 floats) */
int orc_sample_pixel(const SpkSceneDesc *scene, const SpkCamera *camera, uint32_t width, uint32_t height, uint32_t i,
                     uint32_t j, uint32_t frame_seed, uint32_t max_bounce, float output_limit, uint32_t trace_mode,
                     float *out8, OrcCounters *counters);

/* a16: resolve like spk_cuda_resolve (same formats and row conventions) */
int orc_resolve(const float *accum, uint32_t width, uint32_t height, uint32_t accum_mode, uint32_t format,
                uint32_t total_samples, float exposure, void *out);

#ifdef __cplusplus
}
#endif
#endif

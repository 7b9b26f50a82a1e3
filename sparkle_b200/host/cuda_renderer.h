/*
 * cuda_renderer.h — host side of the `cuda` pipeline, shaped like the reference's renderer interface so that it drops in
 * beside cpu/gpu/forward/deferred. Everything below talks to the GPU only through the C ABI of
 * include/spk_cuda.h; there is no CPU fallback.
 *
 *   RenderConfig        libraries/include/renderer/RenderConfig.h:15-120, libraries/source/renderer/RenderConfig.cpp:13-47
 *                       same cvar names (pipeline, spp, max_spp, bounce, width, height, spatial_denoise, ...), same
 *                       defaults, plus the new enumerator Pipeline::Cuda and IsCudaRenderMode()
 *   Str2Enum            libraries/include/core/Enum.h:20-32 (case-insensitive name -> enumerator)
 *   Renderer            libraries/include/renderer/renderer/Renderer.h:28-108 (InitRenderResources, Render, Update,
 *                       GetRenderMode, IsReadyForAutoScreenshot, Tick, RequestSaveScreenshot)
 *   Renderer::CreateRenderer   libraries/source/renderer/renderer/Renderer.cpp:17-53
 *   CudaRenderer        the counterpart of CPURenderer (libraries/source/renderer/renderer/CPURenderer.cpp:22-207)
 *   CameraProxy         CameraRenderProxy (libraries/source/renderer/proxy/CameraRenderProxy.cpp:10-76,113-136;
 *                       libraries/include/renderer/proxy/CameraRenderProxy.h:16-113) without RHI
 *   SceneProxy          what the renderer reads from SceneRenderProxy (libraries/include/renderer/proxy/
 *                       SceneRenderProxy.h:46-113): primitives in registration order, materials, camera, sky, the
 *                       per-frame primitive change list; UpdateBVH (SceneRenderProxy.cpp:195-297)
 */
#ifndef SPK_HOST_CUDA_RENDERER_H_
#define SPK_HOST_CUDA_RENDERER_H_

#include <array>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "../../include/spk_cuda.h"

namespace sparkle_cuda
{
struct Vec3
{
    float x = 0.f, y = 0.f, z = 0.f;
};

// ---- RenderConfig -------------------------------------------------------------------------------------------------
struct RenderConfig
{
    enum class Pipeline : uint8_t
    {
        Cpu,
        Gpu,
        Forward,
        Deferred,
        Cuda, // new: the B200 CUDA path tracer
    };

    Pipeline pipeline = Pipeline::Forward; // cvar default (RenderConfig.cpp:13-14)
    uint32_t sample_per_pixel = 1;         // spp
    uint32_t max_sample_per_pixel = 2048;  // max_spp
    uint32_t random_seed_offset = 0;
    uint32_t image_width = 1280;
    uint32_t image_height = 720;
    float render_scale = 1.f;
    uint32_t max_bounce = 8;     // bounce
    bool spatial_denoise = true; // code default; the packaged config sets false (resources/packed/config/config.json:17)
    bool use_ssao = false;
    bool use_prepass = false;
    // cuda pipeline only (no counterpart in RenderConfig.cpp:13-48): GPUs of this box the frame's samples are split over
    // (scene replicated, spp interleaved, spk_cuda_group_resolve combines), and which devices (default: device, device+1, ..)
    uint32_t gpus = 1;
    // PrimitiveChangeType::Update: rebuild the TLAS on the device (spk_cuda_update_instances with tlas == NULL) instead
    // of uploading the host-rebuilt one
    bool device_tlas = false;
    // every BLAS and the TLAS are built on the device at upload (SpkBvhDesc.nodes == NULL, csrc/spk_build.cu) instead of
    // uploading the host-built bvh::v2 trees
    bool device_bvh = false;
    std::vector<int> gpu_devices;

    [[nodiscard]] bool IsCPURenderMode() const { return pipeline == Pipeline::Cpu; }
    [[nodiscard]] bool IsCudaRenderMode() const { return pipeline == Pipeline::Cuda; }
    [[nodiscard]] bool IsRayTracingMode() const { return pipeline == Pipeline::Gpu; }
    [[nodiscard]] bool IsRaterizationMode() const
    {
        return pipeline == Pipeline::Forward || pipeline == Pipeline::Deferred;
    }
    // the BVH is needed by the two software path tracers (SceneRenderProxy.cpp:192 extended with Cuda)
    [[nodiscard]] bool NeedBvh() const { return IsCPURenderMode() || IsCudaRenderMode(); }

    // "--name value" pairs with the reference's cvar names. Unknown names and bad values return false and log.
    bool SetFromArgs(const std::vector<std::string> &args);
    bool Set(const std::string &name, const std::string &value);
    // RenderConfig::Validate (RenderConfig.cpp:141-155): prepass/ssao are forced off for the ray-traced modes
    void Validate();
};

// Str2Enum (Enum.h:20-32): case-insensitive; false if no enumerator matches
bool Str2Pipeline(const std::string &name, RenderConfig::Pipeline &out);
const char *Pipeline2Str(RenderConfig::Pipeline p);

// ---- scene-side data the renderer consumes --------------------------------------------------------------------------
struct Transform // core/math/Transform.h: translation, XYZ euler rotation, scale -> matrix and inverse (float)
{
    Vec3 translation, euler, scale{1.f, 1.f, 1.f};
    void Matrices(float fwd[12], float inv[12]) const;
};

struct MeshData // io/Mesh.h:7-141
{
    std::vector<float> vertices, normals, tangents, uvs; // 3, 3, 4, 2 per vertex
    std::vector<uint32_t> indices;
    static MeshData UnitCube();                       // Mesh::GetUnitCube
    static MeshData UnitSphere(uint32_t v_size = 32); // Mesh::GetUnitSphere (v_size rings, 2*v_size segments)
};

enum class PrimitiveChangeType : uint8_t // SceneRenderProxy.h:56-60
{
    New,
    Remove,
    Move,
    Update,
};

struct CameraAttribute // CameraComponent.h:16-23 -> CameraRenderProxy::Attribute
{
    float vertical_fov = 0.f;
    float focus_distance = 1.f;
    float aperture_radius = 0.f;
    float exposure = 1.f;
};

class CameraProxy
{
public:
    // OrbitCameraComponent::Setup (scene/component/camera/OrbitCameraComponent.cpp:16-30) + CameraComponent's
    // physical attributes (focal length, sensor height, f-number; CameraComponent.cpp:120-128)
    void SetupOrbit(Vec3 center, float radius, float pitch_deg, float yaw_deg, float focal_length = 0.035f,
                    float sensor_height = 0.024f, float aperture = 22.f, float exposure = 1.f);
    // CameraRenderProxy::Update (CameraRenderProxy.cpp:10-76): focus plane on attribute change, posture on transform
    // change, then cumulated = min(max_spp, cumulated + pending)
    void Update(const RenderConfig &config);
    void MarkPixelDirty();
    void ClearPixels() { pixels_dirty_ = false; }
    void AccumulateSample(uint32_t n) { pending_sample_count_ += n; }
    [[nodiscard]] bool NeedClear() const { return pixels_dirty_; }
    [[nodiscard]] uint32_t GetCumulatedSampleCount() const { return cumulated_sample_count_; }
    [[nodiscard]] const SpkCamera &Get() const { return camera_; }

private:
    Vec3 orbit_center_;
    float orbit_radius_ = 25.f, orbit_pitch_ = 10.f, orbit_yaw_ = -20.f;
    float focal_length_ = 0.035f, sensor_height_ = 0.024f, aperture_ = 22.f, exposure_ = 1.f;
    SpkCamera camera_{};
    bool attribute_dirty_ = true, transform_dirty_ = true, pixels_dirty_ = true;
    uint32_t cumulated_sample_count_ = 0, pending_sample_count_ = 0;
};

class SceneProxy
{
public:
    SceneProxy();
    ~SceneProxy();
    uint32_t AddMesh(MeshData mesh); // builds the BLAS (BLAS::Build, MeshRenderProxy.cpp:31-65)
    uint32_t AddTexture(const void *pixels, uint32_t width, uint32_t height, SpkPixelFormat format);
    uint32_t AddMaterial(const SpkMaterialDesc &material);
    uint32_t AddLambertian(Vec3 color);                 // LambertianMaterial.h: metallic 0, roughness 1
    uint32_t AddMetal(Vec3 color, float roughness);     // MetalMaterial.h:11-15
    uint32_t AddDielectric(float eta);                  // DieletricMaterial.h
    uint32_t AddMeshPrimitive(uint32_t mesh, uint32_t material, const Transform &t);
    // the same with explicit rows 0..2 of the forward and inverse matrices (scene loaders: quaternion rotations, hierarchies)
    uint32_t AddMeshPrimitiveMatrix(uint32_t mesh, uint32_t material, const float fwd[12], const float inv[12]);
    uint32_t AddSphere(Vec3 center, float radius, uint32_t material); // SpherePrimitive.cpp:22-38
    void MovePrimitive(uint32_t primitive, const Transform &t);       // records PrimitiveChangeType::Update
    void SetSky(Vec3 color);
    // SkyLight::SetSkyMap (SkyLight.cpp): an equirectangular Radiance .hdr file, cooked into the 6 x 1024 x 1024 RGBAFloat16
    // cube the path tracer samples (host/sky_cook.h). False (and LastError) when the file cannot be read or decoded.
    bool SetSkyMapFile(const std::string &hdr_path, uint32_t cube_size = 0 /* 0: the reference's 1024 */);
    CameraProxy &GetCamera() { return camera_; }

    // SceneRenderProxy::Update/UpdateBVH: (re)build the TLAS when the change list is not empty; returns the changes of
    // this frame (valid until EndUpdate)
    const std::vector<std::pair<PrimitiveChangeType, uint32_t>> &Update(const RenderConfig &config);
    [[nodiscard]] const std::vector<std::pair<PrimitiveChangeType, uint32_t>> &Changes() const { return changes_; }
    void EndUpdate() { changes_.clear(); }
    // flattened scene for spk_cuda_upload_scene / the oracle; pointers stay valid until the next structural change
    [[nodiscard]] const SpkSceneDesc &Desc() const { return desc_; }
    [[nodiscard]] uint32_t NumPrimitives() const { return static_cast<uint32_t>(primitives_.size()); }

private:
    struct Impl;
    std::unique_ptr<Impl> impl_;
    std::vector<SpkPrimitiveDesc> primitives_;
    std::vector<std::array<float, 6>> world_boxes_;
    std::vector<SpkMaterialDesc> materials_;
    std::vector<SpkTextureDesc> textures_;
    std::vector<std::pair<PrimitiveChangeType, uint32_t>> changes_;
    CameraProxy camera_;
    SpkSceneDesc desc_{};
    void RebuildDesc(bool structural);
};

// ---- Renderer ------------------------------------------------------------------------------------------------------
class Renderer
{
public:
    Renderer(const RenderConfig &config, SceneProxy *scene) : render_config_(config), scene_render_proxy_(scene) {}
    virtual ~Renderer() = default;
    // Renderer::CreateRenderer (Renderer.cpp:17-53): switch on the pipeline, then InitRenderResources. Only Cuda can be
    // created here; the other enumerators belong to the reference's own renderers and yield nullptr + an error log.
    static std::unique_ptr<Renderer> CreateRenderer(const RenderConfig &config, SceneProxy *scene, int device = 0);

    virtual bool InitRenderResources() = 0;
    virtual void Render() = 0;
    [[nodiscard]] virtual RenderConfig::Pipeline GetRenderMode() const = 0;
    [[nodiscard]] virtual bool IsReadyForAutoScreenshot() const { return scene_loaded_; }
    void Tick(); // Renderer::Tick (Renderer.cpp:68-75): scene->Update; Update(); scene->EndUpdate
    void NotifySceneLoaded() { scene_loaded_ = true; }
    void RequestSaveScreenshot(const std::string &path) { screenshot_path_ = path; }

protected:
    virtual void Update() = 0;
    const RenderConfig &render_config_;
    SceneProxy *scene_render_proxy_;
    bool scene_loaded_ = false;
    std::string screenshot_path_;
};

class CudaRenderer : public Renderer
{
public:
    CudaRenderer(const RenderConfig &config, SceneProxy *scene, int device);
    ~CudaRenderer() override;
    bool InitRenderResources() override;
    void Render() override;
    [[nodiscard]] RenderConfig::Pipeline GetRenderMode() const override { return RenderConfig::Pipeline::Cuda; }
    [[nodiscard]] bool IsReadyForAutoScreenshot() const override;
    // linear fp32 accumulation (W*H*4) and the tone-mapped sRGB8 image (rows flipped like output_image_)
    bool ReadLinear(std::vector<float> &out);
    bool ReadSrgb8(std::vector<uint8_t> &out);
    [[nodiscard]] SpkContext *Context() const { return ctx_; }
    [[nodiscard]] uint32_t DispatchedSampleCount() const { return dispatched_sample_count_; }

protected:
    void Update() override;

private:
    int device_;
    SpkContext *ctx_ = nullptr;            // the root context (== ctxs_[0])
    std::vector<SpkContext *> ctxs_;       // one per GPU of the `gpus` cvar
    SpkGroup *group_ = nullptr;            // gpus > 1: fused reduce + resolve over peer memory (include/spk_cuda.h)
    uint32_t dispatched_sample_count_ = 0;
    bool scene_uploaded_ = false;
    bool failed_ = false;
};

// minimal PNG writer (8-bit RGBA, stored deflate blocks): the screenshot path of Renderer.cpp:108-180 without stb
const std::string &LastError();
bool WritePng(const std::string &path, const uint8_t *rgba, uint32_t width, uint32_t height);

// BuildTestScene (tests/usd/UsdRoundTripTest.cpp:25-105) minus the absent glTF models and HDR sky: floor, three fixed
// spheres, ten spheres from the fixed-seed-42 stream (SceneManager.cpp:51-105), gradient sky, orbit camera
void BuildTestScene(SceneProxy &scene);
} // namespace sparkle_cuda

// ---- C shims for the tests and the headless driver (ctypes) ---------------------------------------------------------
extern "C"
{
typedef struct SpkHostApp SpkHostApp;
/* args: space-separated "--name value" pairs, e.g. "--pipeline cuda --width 96 --height 64 --max_spp 4 --scene testscene"
 * --scene: testscene | empty | path/to/file.gltf (GLTFLoader semantics, gltf_loader.h) */
SpkHostApp *spk_host_app_create(const char *args, int device);
void spk_host_app_destroy(SpkHostApp *app);
const char *spk_host_last_error(void);
/* RenderFramework::RenderLoop body (RenderFramework.cpp:146-162): Tick + Render; returns 1 when ready for the auto screenshot */
int spk_host_app_frame(SpkHostApp *app);
int spk_host_app_run(SpkHostApp *app, uint32_t max_frames); /* frames rendered until ready */
/* Tick + Render unconditionally, then the ready flag: the reference's frame, which keeps rendering past max_spp */
int spk_host_app_render_frame(SpkHostApp *app);
const SpkSceneDesc *spk_host_app_scene(SpkHostApp *app);
const SpkCamera *spk_host_app_camera(SpkHostApp *app);
uint32_t spk_host_app_cumulated_samples(SpkHostApp *app);
int spk_host_app_read_linear(SpkHostApp *app, float *out, size_t floats);
int spk_host_app_save_screenshot(SpkHostApp *app, const char *path);
int spk_host_app_move_primitive(SpkHostApp *app, uint32_t primitive, float tx, float ty, float tz);
/* OrbitCameraComponent::Setup on the app's camera (marks the pixels dirty: accumulation restarts) */
void spk_host_app_set_orbit(SpkHostApp *app, float cx, float cy, float cz, float radius, float pitch_deg, float yaw_deg);
/* Str2Enum probe: returns the enumerator value or -1 */
int spk_host_parse_pipeline(const char *name);
}
#endif

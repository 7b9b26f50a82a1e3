/*
 * ref_harness.cpp — TEST INFRASTRUCTURE ONLY (builds oracle/_ref/libsparkle_ref.so, orc_kind() == "reference").
 *
 * Drives the reference's OWN CPU path tracer, compiled in place from /root/reference by oracle/Makefile:
 *   libraries/source/renderer/renderer/CPURenderer.cpp   (SetupViewRay, SamplePixel, RenderPixel, BasePass, DenoisePass,
 *                                                         ToneMappingPass — CPURenderer.cpp:220-576)
 *   libraries/source/renderer/proxy/{Scene,Mesh,Sphere,Primitive,Sky,Material,Camera}RenderProxy.cpp
 *   libraries/source/core/math/{Transform,AABB,Ray,Intersection}.cpp, io/{Mesh,Image}.cpp, core/task/*.cpp ...
 * plus the reference's headers (Sampler.h, BxDF.h, Utilities.h, Image.h, Eigen, xoshiro).
 * Nothing from those files is copied here. What this file adds is glue only:
 *   - SpkSceneDesc (include/spk_cuda.h data contract) -> reference objects (Mesh, MaterialResource, *RenderProxy);
 *   - the BVH topology is injected through oracle/ref_shim/bvh/v2/shim.h because madmann91/bvh is absent from the mount;
 *   - link-time stubs (ref_stubs.cpp) for the RHI/UI symbols the CPU path never calls.
 * `private`/`protected` are opened for this one translation unit so the harness can set the proxies' state without an
 * RHI device (the app cannot boot headless here). Class layout is unaffected by access specifiers.
 */
#include <algorithm>
#include <any>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <filesystem>
#include <fstream>
#include <functional>
#include <future>
#include <iostream>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <numbers>
#include <optional>
#include <queue>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <unordered_set>
#include <variant>
#include <vector>

#include <Eigen/Dense>
#include <XoshiroCpp.hpp>
#include <BS_thread_pool.hpp>
#include <spdlog/spdlog.h>

#define private public
#define protected public
#include "core/math/Sampler.h"
#include "core/math/Ray.h"
#include "core/math/Intersection.h"
#include "core/math/AABB.h"
#include "core/math/Transform.h"
#include "core/math/BVH.h"
#include "core/task/TaskManager.h"
#include "io/Image.h"
#include "io/Material.h"
#include "io/Mesh.h"
#include "scene/material/BxDF.h"
#include "renderer/proxy/CameraRenderProxy.h"
#include "renderer/proxy/DialetricMaterialRenderProxy.h"
#include "renderer/proxy/MeshRenderProxy.h"
#include "renderer/proxy/PbrMaterialRenderProxy.h"
#include "renderer/proxy/SceneRenderProxy.h"
#include "renderer/proxy/SkyRenderProxy.h"
#include "renderer/proxy/SphereRenderProxy.h"
#include "renderer/renderer/CPURenderer.h"
#undef private
#undef protected

#include <stb_image.h> // declarations only: Image.cpp holds the implementation
#include "oracle.h"

using namespace sparkle;

namespace
{
std::mutex g_mutex; // the harness is single-entry: the reference's thread-local RNG and singletons are process-wide

PixelFormat ToPixelFormat(uint32_t f)
{
    switch (f)
    {
    case SPK_FMT_RGBA8_UNORM:
        return PixelFormat::R8G8B8A8Unorm;
    case SPK_FMT_RGBA8_SRGB:
        return PixelFormat::R8G8B8A8Srgb;
    case SPK_FMT_RGBA16F:
        return PixelFormat::RGBAFloat16;
    default:
        return PixelFormat::RGBAFloat;
    }
}

std::shared_ptr<Image2D> MakeImage(const SpkTextureDesc &t)
{
    auto format = ToPixelFormat(t.format);
    size_t bytes = size_t(t.width) * t.height * GetPixelSize(format);
    std::vector<uint8_t> pixels(bytes);
    memcpy(pixels.data(), t.pixels, bytes);
    return std::make_shared<Image2D>(t.width, t.height, format, pixels);
}

Transform MakeTransform(const float *fwd, const float *inv)
{
    TransformData f;
    TransformData i;
    f.matrix().setIdentity();
    i.matrix().setIdentity();
    for (int r = 0; r < 3; r++)
    {
        for (int c = 0; c < 4; c++)
        {
            f.matrix()(r, c) = fwd[r * 4 + c];
            i.matrix()(r, c) = inv[r * 4 + c];
        }
    }
    return Transform(f, i);
}

// flattened tree (28- or 32-byte nodes, 4- or 8-byte ids) -> the 32-byte/u32 form the bvh shim consumes
struct TreeCopy
{
    std::vector<uint8_t> nodes;
    std::vector<uint32_t> prim_ids;
    void From(const SpkBvhDesc &d)
    {
        nodes.assign(size_t(d.num_nodes) * 32, 0);
        const auto *src = static_cast<const uint8_t *>(d.nodes);
        for (uint32_t n = 0; n < d.num_nodes; n++)
        {
            memcpy(nodes.data() + size_t(n) * 32, src + size_t(n) * d.node_stride, 28);
        }
        prim_ids.resize(d.num_prims);
        for (uint32_t k = 0; k < d.num_prims; k++)
        {
            prim_ids[k] = d.prim_id_bytes == 8 ? uint32_t(static_cast<const uint64_t *>(d.prim_ids)[k])
                                               : static_cast<const uint32_t *>(d.prim_ids)[k];
        }
    }
    void Inject(uint32_t expected_prims) const
    {
        auto &t = bvh::v2::shim_injected_tree();
        t.nodes = nodes.data();
        t.num_nodes = uint32_t(nodes.size() / 32);
        t.prim_ids = prim_ids.data();
        t.num_prims = expected_prims;
    }
};

struct RefScene
{
    const void *key = nullptr;
    std::unique_ptr<SceneRenderProxy> scene;
    std::vector<std::shared_ptr<Mesh>> meshes;
    std::vector<std::shared_ptr<Image2D>> textures;
    std::deque<MaterialResource> material_resources; // MaterialRenderProxy keeps a reference: addresses must be stable
    std::vector<MaterialRenderProxy *> materials;
    std::vector<PrimitiveRenderProxy *> primitives;
    std::unique_ptr<SkyRenderProxy> sky;
    std::unique_ptr<CameraRenderProxy> camera;
    std::shared_ptr<Mesh> dummy_sphere_mesh;
    std::deque<std::string> names;
};

std::unique_ptr<RefScene> g_scene;

uint64_t SceneFingerprint(const SpkSceneDesc *d)
{
    // identity of a scene so that repeated calls with the same descriptor reuse the built reference scene. Addresses
    // alone are NOT an identity (a freed descriptor's memory is handed out again: a second TestScene with a cube-map sky
    // once came back as the cached gradient-sky one), so every table is hashed by CONTENT: all primitives, materials,
    // texture and sky descriptors with the head of their pixels, mesh descriptors with the head of their arrays, the
    // head of every tree.
    uint64_t h = 1469598103934665603ull;
    auto mix = [&h](uint64_t v) { h = (h ^ v) * 1099511628211ull; };
    auto bytes = [&](const void *p, size_t n) {
        const auto *b = static_cast<const uint8_t *>(p);
        for (size_t k = 0; p && k < n; k++)
        {
            mix(b[k]);
        }
    };
    auto texel_bytes = [](uint32_t format) -> size_t {
        return format == SPK_FMT_RGBA32F ? 16 : format == SPK_FMT_RGBA16F ? 8 : 4;
    };
    auto texture = [&](const SpkTextureDesc &t) {
        mix(t.width);
        mix(t.height);
        mix(t.format);
        const size_t total = size_t(t.width) * t.height * texel_bytes(t.format);
        bytes(t.pixels, std::min<size_t>(total, 256));
        if (t.pixels && total > 512)
        {
            bytes(static_cast<const uint8_t *>(t.pixels) + total - 256, 256);
        }
    };
    auto tree = [&](const SpkBvhDesc &t) {
        mix(t.num_nodes);
        mix(t.num_prims);
        bytes(t.nodes, size_t(std::min<uint32_t>(t.num_nodes, 16u)) * t.node_stride);
        bytes(t.prim_ids, size_t(std::min<uint32_t>(t.num_prims, 64u)) * t.prim_id_bytes);
    };
    mix(reinterpret_cast<uint64_t>(d));
    mix(d->num_meshes);
    mix(d->num_primitives);
    mix(d->num_materials);
    mix(d->num_textures);
    bytes(d->primitives, sizeof(SpkPrimitiveDesc) * size_t(d->num_primitives));
    bytes(d->materials, sizeof(SpkMaterialDesc) * size_t(d->num_materials));
    for (uint32_t t = 0; t < d->num_textures; t++)
    {
        texture(d->textures[t]);
    }
    for (uint32_t m = 0; m < d->num_meshes; m++)
    {
        const SpkMeshDesc &md = d->meshes[m];
        mix(md.num_vertices);
        mix(md.num_faces);
        bytes(md.vertices, size_t(std::min<uint32_t>(md.num_vertices, 32u)) * 12);
        bytes(md.indices, size_t(std::min<uint32_t>(md.num_faces, 32u)) * 12);
        tree(md.blas);
    }
    tree(d->tlas);
    mix(d->sky.has_sky);
    mix(d->sky.has_map);
    bytes(d->sky.color, 12);
    for (int f = 0; d->sky.has_sky && d->sky.has_map && f < 6; f++)
    {
        texture(d->sky.faces[f]);
    }
    return h;
}
uint64_t g_scene_fp = 0;

RefScene *BuildScene(const SpkSceneDesc *d)
{
    uint64_t fp = SceneFingerprint(d);
    if (g_scene && g_scene_fp == fp)
    {
        return g_scene.get();
    }
    g_scene.reset();
    auto rs = std::make_unique<RefScene>();
    rs->scene = std::make_unique<SceneRenderProxy>();

    for (uint32_t t = 0; t < d->num_textures; t++)
    {
        rs->textures.push_back(MakeImage(d->textures[t]));
    }
    auto tex = [&](uint32_t id) -> std::shared_ptr<Image2D> {
        return id == SPK_INVALID_ID ? nullptr : rs->textures[id];
    };
    for (uint32_t m = 0; m < d->num_materials; m++)
    {
        const auto &md = d->materials[m];
        rs->material_resources.emplace_back();
        auto &res = rs->material_resources.back();
        res.base_color = Vector3(md.base_color[0], md.base_color[1], md.base_color[2]);
        res.emissive_color = Vector3(md.emissive[0], md.emissive[1], md.emissive[2]);
        res.metallic = md.metallic;
        res.roughness = md.roughness;
        res.eta = md.eta;
        res.base_color_texture = tex(md.base_color_tex);
        res.emissive_texture = tex(md.emissive_tex);
        res.metallic_roughness_texture = tex(md.metallic_roughness_tex);
        res.normal_texture = tex(md.normal_tex);
        res.name = "material";
        std::unique_ptr<MaterialRenderProxy> proxy;
        if (md.type == SPK_MAT_DIELECTRIC)
        {
            proxy = std::make_unique<DieletricMaterialRenderProxy>(res);
        }
        else
        {
            proxy = std::make_unique<PbrMaterialRenderProxy>(res);
        }
        rs->materials.push_back(rs->scene->AddMaterial(std::move(proxy)));
    }

    std::vector<TreeCopy> blas_trees(d->num_meshes);
    for (uint32_t m = 0; m < d->num_meshes; m++)
    {
        const auto &md = d->meshes[m];
        auto mesh = std::make_shared<Mesh>();
        mesh->name = "mesh";
        mesh->vertices.resize(md.num_vertices);
        mesh->normals.resize(md.num_vertices);
        mesh->tangents.resize(md.num_vertices);
        for (uint32_t v = 0; v < md.num_vertices; v++)
        {
            mesh->vertices[v] = Vector3(md.vertices[3 * v], md.vertices[3 * v + 1], md.vertices[3 * v + 2]);
            mesh->normals[v] = Vector3(md.normals[3 * v], md.normals[3 * v + 1], md.normals[3 * v + 2]);
            mesh->tangents[v] =
                Vector4(md.tangents[4 * v], md.tangents[4 * v + 1], md.tangents[4 * v + 2], md.tangents[4 * v + 3]);
        }
        if (md.uvs)
        {
            mesh->uvs.resize(md.num_vertices);
            for (uint32_t v = 0; v < md.num_vertices; v++)
            {
                mesh->uvs[v] = Vector2(md.uvs[2 * v], md.uvs[2 * v + 1]);
            }
        }
        mesh->indices.assign(md.indices, md.indices + size_t(md.num_faces) * 3);
        rs->meshes.push_back(mesh);
        blas_trees[m].From(md.blas);
    }

    // spheres derive from MeshRenderProxy and carry a raw mesh they never intersect (SphereRenderProxy.h:9-17)
    rs->dummy_sphere_mesh = std::make_shared<Mesh>();
    rs->dummy_sphere_mesh->vertices = {Vector3(0, 0, 0), Vector3(1, 0, 0), Vector3(0, 1, 0)};
    rs->dummy_sphere_mesh->normals = {Up, Up, Up};
    rs->dummy_sphere_mesh->tangents.assign(3, Vector4(1, 0, 0, 1));
    rs->dummy_sphere_mesh->indices = {0, 1, 2};

    const AABB unit_bound(Zeros, Ones);
    for (uint32_t p = 0; p < d->num_primitives; p++)
    {
        const auto &pd = d->primitives[p];
        rs->names.emplace_back("prim" + std::to_string(p));
        std::unique_ptr<RenderProxy> proxy;
        if (pd.kind == SPK_PRIM_SPHERE)
        {
            auto sphere = std::make_unique<SphereRenderProxy>(rs->dummy_sphere_mesh, rs->names.back(), unit_bound);
            sphere->scaled_radius_2_ = pd.radius2;
            proxy = std::move(sphere);
        }
        else
        {
            proxy = std::make_unique<MeshRenderProxy>(rs->meshes[pd.mesh], rs->names.back(), unit_bound);
        }
        auto *prim = static_cast<PrimitiveRenderProxy *>(proxy.get());
        prim->UpdateTransform(MakeTransform(pd.fwd, pd.inv));
        prim->transform_dirty_ = 0;
        prim->rhi_dirty_ = 0;
        prim->SetMaterialRenderProxy(rs->materials[pd.material]);
        rs->scene->AddRenderProxy(std::move(proxy));
        rs->primitives.push_back(prim);
        // BLAS::Build (MeshRenderProxy.cpp:31-65) with the injected topology
        if (pd.kind == SPK_PRIM_SPHERE)
        {
            bvh::v2::shim_injected_tree() = {};
        }
        else
        {
            blas_trees[pd.mesh].Inject(d->meshes[pd.mesh].num_faces);
        }
        prim->BuildBVH();
    }
    // TLAS::Build (SceneRenderProxy.cpp:31-69) through SceneRenderProxy::UpdateBVH; BLASes are already built
    rs->scene->primitive_changes_.clear();
    rs->scene->need_bvh_update_ = true;
    TreeCopy tlas;
    tlas.From(d->tlas);
    tlas.Inject(d->num_primitives);
    rs->scene->UpdateBVH();

    if (d->sky.has_sky)
    {
        std::shared_ptr<Image2DCube> cube;
        if (d->sky.has_map)
        {
            std::array<std::unique_ptr<Image2D>, 6> faces;
            for (int f = 0; f < 6; f++)
            {
                auto img = MakeImage(d->sky.faces[f]);
                faces[f] = std::make_unique<Image2D>(*img);
            }
            cube = std::make_shared<Image2DCube>(std::move(faces), "sky");
        }
        rs->sky = std::make_unique<SkyRenderProxy>(cube);
        rs->sky->SetData(Vector3(d->sky.color[0], d->sky.color[1], d->sky.color[2]));
        if (cube)
        {
            // SkyRenderProxy::Evaluate gates on the RHI image handle (SkyRenderProxy.cpp:53); never dereferenced
            static int dummy_target;
            rs->sky->sky_map_.ptr_ =
                std::shared_ptr<RHIImage>(reinterpret_cast<RHIImage *>(&dummy_target), [](RHIImage *) {});
        }
        rs->scene->SetSkyLight(rs->sky.get());
    }

    rs->camera = std::make_unique<CameraRenderProxy>();
    rs->scene->SetCamera(rs->camera.get());

    g_scene = std::move(rs);
    g_scene_fp = fp;
    return g_scene.get();
}

void SetCamera(RefScene *rs, const SpkCamera *c)
{
    auto v3 = [](const float *p) { return Vector3(p[0], p[1], p[2]); };
    auto &cam = *rs->camera;
    cam.posture_.position = v3(c->position);
    cam.posture_.up = v3(c->up);
    cam.posture_.front = v3(c->front);
    cam.posture_.right = v3(c->right);
    cam.focus_plane_.max_u = v3(c->max_u);
    cam.focus_plane_.max_v = v3(c->max_v);
    cam.focus_plane_.lower_left = v3(c->lower_left);
    cam.state_.aperture_radius = c->aperture_radius;
    cam.state_.exposure = c->exposure;
    cam.state_.focus_distance = 1.f;
    cam.state_.vertical_fov = 1.f;
}

void EnsureTaskManager(uint32_t threads)
{
    static std::unique_ptr<TaskManager> manager;
    if (!manager)
    {
        spdlog::set_level(spdlog::level::warn);
        // workers = min(max_parallism, hw) - reserved (TaskDispatcher.cpp:19-24)
        manager = std::make_unique<TaskManager>(std::max(1u, threads) + 1, 1);
    }
}

void FillRenderConfig(RenderConfig &cfg, uint32_t width, uint32_t height, uint32_t max_bounce)
{
    cfg.pipeline = RenderConfig::Pipeline::Cpu;
    cfg.output_image = RenderConfig::OutputImage::SceneColor;
    cfg.debug_mode = RenderConfig::DebugMode::Color;
    cfg.sample_per_pixel = 1;
    cfg.max_sample_per_pixel = 0x7fffffff;
    cfg.random_seed_offset = 0;
    cfg.max_bounce = max_bounce;
    cfg.image_width = width;
    cfg.image_height = height;
    cfg.shadow_map_resolution = 1;
    cfg.spatial_denoise = false;
    cfg.use_ssao = false;
    cfg.use_prepass = false;
    cfg.use_diffuse_ibl = false;
    cfg.use_specular_ibl = false;
    cfg.use_vsync = false;
    cfg.use_dynamic_spp = false;
    cfg.enable_nee = false;
    cfg.clear_screenshots = false;
    cfg.manual_accumulation = false;
    cfg.target_framerate = 60.f;
    cfg.gpu_time_budget_ratio = 1.f;
    cfg.render_scale = 1.f;
}

SpkHit TraceOne(RefScene *rs, const SpkRay &r, uint32_t mode)
{
    SpkHit h{std::numeric_limits<float>::max(), 0.f, 0.f, SPK_INVALID_ID, SPK_INVALID_ID};
    Ray ray;
    ray.Reset(Vector3(r.origin[0], r.origin[1], r.origin[2]), Vector3(r.direction[0], r.direction[1], r.direction[2]));
    PrimitiveRenderProxy *winner = nullptr;
    if (mode == ORC_TRACE_BRUTE_FORCE)
    {
        // every primitive in registration order; meshes still go through their BLAS (the triangle test is
        // file-local in MeshRenderProxy.cpp). The port's brute force visits every triangle.
        IntersectionCandidate cand;
        for (auto *prim : rs->primitives)
        {
            if (prim->Intersect(ray, cand))
            {
                winner = prim;
            }
        }
    }
    else
    {
        Intersection isect;
        rs->scene->Intersect<false>(ray, isect);
        winner = const_cast<PrimitiveRenderProxy *>(isect.GetPrimitive());
    }
    if (!winner)
    {
        return h;
    }
    // the candidate (u, v, face) is private to TLAS::IntersectInternal: re-run the winner alone to read it
    IntersectionCandidate cand;
    cand.u = 0.f;
    cand.v = 0.f;
    cand.face_idx = SPK_INVALID_ID;
    winner->Intersect(ray, cand);
    h.t = cand.t;
    h.prim = winner->GetPrimitiveIndex();
    if (winner->IsMesh() && dynamic_cast<SphereRenderProxy *>(winner) == nullptr)
    {
        h.u = cand.u;
        h.v = cand.v;
        h.face = cand.face_idx;
    }
    return h;
}
} // namespace

extern "C"
{
const char *orc_kind(void)
{
    return "reference";
}

void orc_rng_state(uint32_t seed, uint32_t state_out[4])
{
    XoshiroCpp::Xoshiro128Plus rng(seed);
    auto s = rng.serialize();
    for (int k = 0; k < 4; k++)
    {
        state_out[k] = s[k];
    }
}

void orc_rng_units(uint32_t seed, uint32_t n, float *out)
{
    std::scoped_lock lock(g_mutex);
    sampler::ReseedCurrentThread(seed);
    for (uint32_t k = 0; k < n; k++)
    {
        out[k] = sampler::RandomUnit();
    }
}

void orc_kat_vec(uint32_t op, const float *a, const float *b, const float *m, float *out)
{
    Vector3 va(a[0], a[1], a[2]);
    Vector3 vb = b ? Vector3(b[0], b[1], b[2]) : Vector3(Zeros);
    TransformData t;
    t.matrix().setIdentity();
    if (m)
    {
        for (int r = 0; r < 3; r++)
        {
            for (int c = 0; c < 4; c++)
            {
                t.matrix()(r, c) = m[r * 4 + c];
            }
        }
    }
    Transform tr(t, t);
    Vector3 r3 = Zeros;
    switch (op)
    {
    case 0:
        out[0] = va.dot(vb);
        return;
    case 1:
        out[0] = va.squaredNorm();
        return;
    case 2:
        r3 = va.normalized();
        break;
    case 3:
        r3 = va.cross(vb);
        break;
    case 4:
        r3 = tr.TransformDirection(va);
        break;
    case 5:
        r3 = tr.TransformPoint(va);
        break;
    default:
        break;
    }
    out[0] = r3.x();
    out[1] = r3.y();
    out[2] = r3.z();
}

float orc_kat_libm(uint32_t op, float x, float y)
{
    switch (op)
    {
    case 0:
        return std::sin(x);
    case 1:
        return std::cos(x);
    case 2:
        return std::pow(x, y);
    case 3:
        return std::sqrt(x);
    case 4:
        return std::acos(x);
    case 5:
        return std::tan(x);
    case 6:
        return std::atan2(x, y);
    default:
        return 0.f;
    }
}

void orc_kat_unit_disk(uint32_t seed, float *out)
{
    std::scoped_lock lock(g_mutex);
    sampler::ReseedCurrentThread(seed);
    Vector2 d = sampler::UnitDisk();
    out[0] = d.x();
    out[1] = d.y();
}

void orc_kat_texture(const SpkTextureDesc *tex, float u, float v, float *out)
{
    auto img = MakeImage(*tex);
    Vector3 c = img->Sample(Vector2(u, v));
    out[0] = c.x();
    out[1] = c.y();
    out[2] = c.z();
}

void orc_kat_sky(const SpkSkyDesc *sky, const float *direction, float *out)
{
    std::scoped_lock lock(g_mutex);
    SpkSceneDesc d{};
    d.abi_version = SPK_ABI_VERSION;
    d.sky = *sky;
    static SpkBvhNode empty_node{};
    d.tlas.nodes = &empty_node;
    d.tlas.node_stride = 32;
    d.tlas.prim_id_bytes = 4;
    RefScene *rs = BuildScene(&d);
    g_scene_fp = 0; // descriptor lives on this stack frame: never reuse
    Ray ray;
    ray.Reset(Vector3(0, 0, 0), Vector3(direction[0], direction[1], direction[2]));
    Vector3 c = rs->sky ? rs->sky->Evaluate(ray) : Vector3(Zeros);
    out[0] = c.x();
    out[1] = c.y();
    out[2] = c.z();
}

// a17 / transforms. The scene-graph classes (SceneNode, OrbitCameraComponent, CameraComponent) pull the whole app in, so
// the few lines of theirs on this path are restated here WITH THE SAME Eigen expressions (so the compiler emits the same
// float operations), and everything below them is the reference's own compiled code: Transform::Update /
// UpdateTransformData / GetInverse (Transform.cpp:7-15,52-64, Transform.h:110-126) and
// CameraRenderProxy::OnTransformDirty (CameraRenderProxy.cpp:113-136).
int orc_kat_transform(const float *translate, const float *euler, const float *scale, float *fwd, float *inv)
{
    // SceneNode::SetTransform(translate, Vector3 rotation, scale) -> Transform::Update(.., EulerRotationToRotationAxis(..), ..)
    // (SceneNode.h:71-76, Transform.h:40-43); a node under the identity root: world == local (SceneNode.cpp:30-44)
    Transform t;
    t.Update(Vector3(translate[0], translate[1], translate[2]), Vector3(euler[0], euler[1], euler[2]),
             Vector3(scale[0], scale[1], scale[2]));
    const auto &m = t.GetMatrix();
    const Transform it = t.GetInverse(); // what MeshRenderProxy uses for the object-space ray (MeshRenderProxy.cpp:99-100)
    const auto &im = it.GetMatrix();
    for (int r = 0; r < 3; r++)
    {
        for (int c = 0; c < 4; c++)
        {
            fwd[r * 4 + c] = m(r, c);
            inv[r * 4 + c] = im(r, c);
        }
    }
    return 0;
}

// GLTFLoader ProcessNode's transform lines (io/scene/GLTFLoader.cpp:532-544; tinygltf is absent, so the loader itself
// cannot be compiled): rotation = AngleAxis(90 deg, Right) * node.rotation, SetTransform(translation, normalized, Ones),
// for a node directly under the identity scene root. q = glTF xyzw.
int orc_kat_gltf_node(const float *translation, const float *q, float *fwd, float *inv)
{
    Rotation rotation = utilities::Vector4AsQuaternion(Vector4(q[0], q[1], q[2], q[3]));
    rotation = Eigen::AngleAxis<Scalar>(utilities::ToRadian(90.0f), Right) * rotation;
    Transform t;
    t.Update(Vector3(translation[0], translation[1], translation[2]), rotation.normalized(), Vector3(Ones));
    const auto &m = t.GetMatrix();
    const Transform it = t.GetInverse();
    const auto &im = it.GetMatrix();
    for (int r = 0; r < 3; r++)
    {
        for (int c = 0; c < 4; c++)
        {
            fwd[r * 4 + c] = m(r, c);
            inv[r * 4 + c] = im(r, c);
        }
    }
    return 0;
}

int orc_kat_camera(const float *center, float radius, float pitch, float yaw, float focal_length, float sensor_height,
                   float aperture, float exposure, uint32_t width, uint32_t height, SpkCamera *out)
{
    // OrbitCameraComponent::UpdateTransform (OrbitCameraComponent.cpp:16-30)
    const Vector3 relative_rotation = utilities::ToRadian(Vector3(-pitch, 0.f, -yaw));
    auto rotation_transform =
        Eigen::AngleAxis<Scalar>(relative_rotation.z(), Up) * Eigen::AngleAxis<Scalar>(relative_rotation.x(), Right);
    const Vector3 relative_direction = rotation_transform * Front;
    const Vector3 position = relative_direction * -radius + Vector3(center[0], center[1], center[2]);
    Transform node; // GetNode()->SetTransform(position, -relative_rotation)
    node.Update(position, Vector3(-relative_rotation), Vector3(Ones));
    // CalculateRenderAttribute (CameraComponent.cpp:120-128), SetFocusDistance(radius_)
    CameraRenderProxy cam;
    cam.state_.vertical_fov = 2.f * std::atan(sensor_height / (2.f * focal_length));
    cam.state_.focus_distance = radius;
    cam.state_.exposure = exposure;
    cam.state_.aperture_radius = focal_length / aperture * 0.5f;
    cam.UpdateTransform(node);
    // the attribute_dirty_ block of CameraRenderProxy::Update (CameraRenderProxy.cpp:21-33; Update itself uploads a UBO
    // through the RHI and cannot run headless); AspectRatio: RenderResolution.h:25-28
    cam.aspect_ratio_ = static_cast<float>(width) / static_cast<float>(height);
    auto h = std::tan(cam.state_.vertical_fov / 2.f);
    cam.focus_plane_.height = 2.f * h * cam.state_.focus_distance;
    cam.focus_plane_.width = cam.aspect_ratio_ * cam.focus_plane_.height;
    cam.SetupProjectionMatrix();
    cam.OnTransformDirty(nullptr); // the reference's own code
    auto put = [](float *dst, const Vector3 &v) {
        dst[0] = v.x();
        dst[1] = v.y();
        dst[2] = v.z();
    };
    put(out->position, cam.posture_.position);
    put(out->up, cam.posture_.up);
    put(out->front, cam.posture_.front);
    put(out->right, cam.posture_.right);
    put(out->max_u, cam.focus_plane_.max_u);
    put(out->max_v, cam.focus_plane_.max_v);
    put(out->lower_left, cam.focus_plane_.lower_left);
    out->aperture_radius = cam.state_.aperture_radius;
    out->exposure = cam.state_.exposure;
    return 0;
}

uint32_t orc_kat_sample_surface(const SpkSceneDesc *scene, uint32_t material, uint32_t seed, const float *dir,
                                const float *normal, const float *tangent, const float *uv, float *out_wi,
                                float *out_throughput)
{
    std::scoped_lock lock(g_mutex);
    RefScene *rs = BuildScene(scene);
    Ray ray;
    ray.Reset(Vector3(0, 0, 0), Vector3(dir[0], dir[1], dir[2]));
    sampler::ReseedCurrentThread(seed);
    Vector3 wi = Zeros;
    Vector3 thr = rs->materials[material]->SampleSurface(ray, wi, Vector3(normal[0], normal[1], normal[2]),
                                                         Vector3(tangent[0], tangent[1], tangent[2]),
                                                         Vector2(uv[0], uv[1]));
    for (int k = 0; k < 3; k++)
    {
        out_wi[k] = wi[k];
        out_throughput[k] = thr[k];
    }
    // draws consumed = position of the stream relative to a fresh copy
    XoshiroCpp::Xoshiro128Plus probe(seed);
    auto now = sampler::detail::GetDefaultRng().serialize();
    for (uint32_t n = 0; n < 64; n++)
    {
        if (probe.serialize() == now)
        {
            return n;
        }
        probe();
    }
    return 0xFFFFFFFFu;
}

int orc_trace_closest(const SpkSceneDesc *scene, const SpkRay *rays, uint64_t num_rays, SpkHit *hits, uint32_t mode,
                      uint32_t threads, OrcCounters *counters)
{
    std::scoped_lock lock(g_mutex);
    if (mode == ORC_TRACE_BVH_CULLED)
    {
        return SPK_ERR_UNSUPPORTED; // the reference never shortens tmax (SceneRenderProxy.cpp:95-113)
    }
    RefScene *rs = BuildScene(scene);
    if (counters)
    {
        *counters = OrcCounters{};
        counters->rays = num_rays;
    }
    uint32_t nt = std::max(1u, threads);
    if (nt == 1 || num_rays < 1024)
    {
        for (uint64_t k = 0; k < num_rays; k++)
        {
            hits[k] = TraceOne(rs, rays[k], mode);
        }
        return SPK_OK;
    }
    std::vector<std::thread> pool;
    std::atomic<uint64_t> next{0};
    for (uint32_t t = 0; t < nt; t++)
    {
        pool.emplace_back([&]() {
            for (;;)
            {
                uint64_t b = next.fetch_add(4096);
                if (b >= num_rays)
                {
                    break;
                }
                uint64_t e = std::min(num_rays, b + 4096);
                for (uint64_t k = b; k < e; k++)
                {
                    hits[k] = TraceOne(rs, rays[k], mode);
                }
            }
        });
    }
    for (auto &th : pool)
    {
        th.join();
    }
    return SPK_OK;
}

uint32_t orc_intersect_full(const SpkSceneDesc *scene, const SpkRay *r, float *out12)
{
    std::scoped_lock lock(g_mutex);
    RefScene *rs = BuildScene(scene);
    Ray ray;
    ray.Reset(Vector3(r->origin[0], r->origin[1], r->origin[2]),
              Vector3(r->direction[0], r->direction[1], r->direction[2]));
    Intersection isect;
    rs->scene->Intersect<false>(ray, isect);
    if (!isect.IsHit())
    {
        return SPK_INVALID_ID;
    }
    out12[0] = isect.T();
    for (int k = 0; k < 3; k++)
    {
        out12[1 + k] = isect.GetLocation()[k];
        out12[4 + k] = isect.GetNormal()[k];
        out12[7 + k] = isect.GetTangent()[k];
    }
    out12[10] = isect.GetTexCoord().x();
    out12[11] = isect.GetTexCoord().y();
    return isect.GetPrimitive()->GetPrimitiveIndex();
}

int orc_render(const SpkSceneDesc *scene, const SpkCamera *camera, uint32_t width, uint32_t height,
               const SpkRenderParams *params, float *accum, uint32_t *cumulated, uint32_t trace_mode, uint32_t threads,
               OrcCounters *counters)
{
    std::scoped_lock lock(g_mutex);
    if (trace_mode != ORC_TRACE_BVH_REFERENCE)
    {
        return SPK_ERR_UNSUPPORTED;
    }
    if (params->output_limit != CameraRenderProxy::OutputLimit)
    {
        return SPK_ERR_UNSUPPORTED; // compiled-in constant (CameraRenderProxy.h:16)
    }
    RefScene *rs = BuildScene(scene);
    SetCamera(rs, camera);
    EnsureTaskManager(threads);

    RenderConfig cfg;
    FillRenderConfig(cfg, width, height, params->max_bounce);
    const bool spatial = params->flags & SPK_FLAG_SPATIAL_DENOISE;
    if (spatial && params->accum_mode != SPK_ACCUM_RUNNING_AVERAGE)
    {
        return SPK_ERR_UNSUPPORTED; // the reference's DenoisePass is the running average
    }
    cfg.spatial_denoise = spatial; // RenderPixel stores first-hit normals, DenoisePass fills invalid neighbours
    CPURenderer renderer(cfg, nullptr, rs->scene.get());
    renderer.camera_ = rs->camera.get();
    renderer.gbuffer_.Resize(width, height);
    renderer.ping_pong_buffer_.assign(height, std::vector<Vector4>(width, Vector4::Zero()));
    renderer.frame_buffer_.assign(height, std::vector<Vector4>(width, Vector4::Zero()));
    renderer.sub_pixel_count_ = 1;
    renderer.actual_sample_per_pixel_ = 1;
    for (uint32_t j = 0; j < height; j++)
    {
        memcpy(renderer.frame_buffer_[j].data(), accum + size_t(j) * width * 4, size_t(width) * 16);
    }
    const Vector2UInt no_debug(UINT_MAX, UINT_MAX);
    uint32_t stride = std::max(1u, params->sample_stride);
    for (uint32_t s = 0; s < params->sample_count; s++)
    {
        renderer.dispatched_sample_count_ = params->first_sample + s * stride; // frame_seed, CPURenderer.cpp:432
        renderer.BasePass(*rs->scene, cfg, no_debug);
        if (params->accum_mode == SPK_ACCUM_RUNNING_AVERAGE)
        {
            // the cap CameraRenderProxy::Update applies (CameraRenderProxy.cpp:70-73; Update itself needs the RHI)
            rs->camera->cumulated_sample_count_ =
                params->cumulated_cap ? std::min(*cumulated, params->cumulated_cap) : *cumulated;
            renderer.DenoisePass(cfg, no_debug); // spatial_denoise=false: temporal lerp only (CPURenderer.cpp:530-544)
        }
        else
        {
            for (uint32_t j = 0; j < height; j++)
            {
                for (uint32_t i = 0; i < width; i++)
                {
                    renderer.frame_buffer_[j][i] += renderer.gbuffer_.color[j][i];
                }
            }
        }
        *cumulated += 1;
    }
    for (uint32_t j = 0; j < height; j++)
    {
        memcpy(accum + size_t(j) * width * 4, renderer.frame_buffer_[j].data(), size_t(width) * 16);
    }
    if (counters)
    {
        *counters = OrcCounters{};
        counters->paths = uint64_t(width) * height * params->sample_count;
    }
    return SPK_OK;
}

int orc_sample_pixel(const SpkSceneDesc *scene, const SpkCamera *camera, uint32_t width, uint32_t height, uint32_t i,
                     uint32_t j, uint32_t frame_seed, uint32_t max_bounce, float output_limit, uint32_t trace_mode,
                     float *out8, OrcCounters *counters)
{
    std::scoped_lock lock(g_mutex);
    (void)counters;
    if (trace_mode != ORC_TRACE_BVH_REFERENCE || output_limit != CameraRenderProxy::OutputLimit)
    {
        return SPK_ERR_UNSUPPORTED;
    }
    RefScene *rs = BuildScene(scene);
    SetCamera(rs, camera);
    RenderConfig cfg;
    FillRenderConfig(cfg, width, height, max_bounce);
    cfg.spatial_denoise = true; // makes RenderPixel store the first-hit normal (CPURenderer.cpp:413-416)
    static std::unique_ptr<CPURenderer> renderer;
    static uint32_t rw = 0;
    static uint32_t rh = 0;
    if (!renderer || rw != width || rh != height)
    {
        renderer = std::make_unique<CPURenderer>(cfg, nullptr, rs->scene.get());
        renderer->gbuffer_.Resize(width, height);
        rw = width;
        rh = height;
    }
    renderer->camera_ = rs->camera.get();
    renderer->scene_render_proxy_ = rs->scene.get();
    // BasePass per-pixel body (CPURenderer.cpp:435-447)
    sampler::ReseedCurrentThread(j * width + i + frame_seed * width * height);
    renderer->RenderPixel(i, j, 1.f / static_cast<float>(width - 1), 1.f / static_cast<float>(height - 1), *rs->scene,
                          cfg, Vector2UInt(UINT_MAX, UINT_MAX));
    const Vector4 &c = renderer->gbuffer_.color[j][i];
    const Vector3 &n = renderer->gbuffer_.world_normal[j][i];
    out8[0] = c.x();
    out8[1] = c.y();
    out8[2] = c.z();
    out8[3] = c.w();
    out8[4] = n.x();
    out8[5] = n.y();
    out8[6] = n.z();
    out8[7] = 0.f;
    return SPK_OK;
}

int orc_resolve(const float *accum, uint32_t width, uint32_t height, uint32_t accum_mode, uint32_t format,
                uint32_t total_samples, float exposure, void *out)
{
    std::scoped_lock lock(g_mutex);
    if (format == SPK_RESOLVE_LINEAR_F32)
    {
        float scale = accum_mode == SPK_ACCUM_SUM ? 1.f / static_cast<float>(std::max(1u, total_samples)) : 1.f;
        auto *o = static_cast<float *>(out);
        for (size_t k = 0; k < size_t(width) * height * 4; k++)
        {
            o[k] = accum_mode == SPK_ACCUM_SUM ? accum[k] / static_cast<float>(std::max(1u, total_samples)) : accum[k];
        }
        (void)scale;
        return SPK_OK;
    }
    if (format != SPK_RESOLVE_TONEMAP_F16)
    {
        return SPK_ERR_UNSUPPORTED;
    }
    EnsureTaskManager(1);
    RenderConfig cfg;
    FillRenderConfig(cfg, width, height, 1);
    SceneRenderProxy dummy_scene;
    CameraRenderProxy camera;
    camera.state_.exposure = exposure;
    CPURenderer renderer(cfg, nullptr, &dummy_scene);
    renderer.camera_ = &camera;
    renderer.frame_buffer_.assign(height, std::vector<Vector4>(width, Vector4::Zero()));
    float div = accum_mode == SPK_ACCUM_SUM ? static_cast<float>(std::max(1u, total_samples)) : 1.f;
    for (uint32_t j = 0; j < height; j++)
    {
        for (uint32_t i = 0; i < width; i++)
        {
            const float *p = accum + (size_t(j) * width + i) * 4;
            renderer.frame_buffer_[j][i] = accum_mode == SPK_ACCUM_SUM ? Vector4(p[0] / div, p[1] / div, p[2] / div, p[3] / div)
                                                                       : Vector4(p[0], p[1], p[2], p[3]);
        }
    }
    renderer.ToneMappingPass(renderer.output_image_); // CPURenderer.cpp:565-576
    memcpy(out, renderer.output_image_.GetRawData(), size_t(width) * height * 8);
    return SPK_OK;
}

// ---- the sky map of a SkyLight (pins sparkle_b200/host/sky_cook.cpp): the reference's own functions, called the way
// Image2D::LoadFromFile (io/Image.cpp:152-181) and SkyLightCookJob::Execute (scene/component/light/SkyLight.cpp:103-118) do
int orc_kat_decode_hdr(const unsigned char *data, uint32_t size, uint32_t *width, uint32_t *height, uint16_t *rgba_half,
                       uint64_t capacity_halves)
{
    if (!stbi_is_hdr_from_memory(data, int(size)))
    {
        return -1;
    }
    int w = 0, h = 0, channels = 0;
    float *px = stbi_loadf_from_memory(data, int(size), &w, &h, &channels, 4);
    if (!px)
    {
        return -2;
    }
    *width = uint32_t(w);
    *height = uint32_t(h);
    if (uint64_t(w) * uint64_t(h) * 4 > capacity_halves)
    {
        stbi_image_free(px);
        return -3;
    }
    const float half_max = static_cast<float>(std::numeric_limits<Half>::max());
    for (size_t k = 0; k < size_t(w) * size_t(h) * 4; k++)
    {
        const Half v = static_cast<Half>(std::clamp(px[k], -half_max, half_max));
        memcpy(rgba_half + k, &v, 2);
    }
    stbi_image_free(px);
    return 0;
}

int orc_kat_cook_sky_cube(const uint16_t *equirect_rgba_half, uint32_t width, uint32_t height, uint32_t size, uint16_t *faces_out)
{
    Image2D sky_map(width, height, PixelFormat::RGBAFloat16);
    memcpy(const_cast<uint8_t *>(sky_map.GetRawData()), equirect_rgba_half, size_t(width) * height * 8);
    Image2DCube cube_map(size, size, PixelFormat::RGBAFloat16, "kat_CubeMap");
    for (unsigned id = 0; id < Image2DCube::FaceId::Count; id++)
    {
        auto face_id = static_cast<Image2DCube::FaceId>(id);
        auto &this_face = cube_map.GetFace(face_id);
        for (unsigned i = 0; i < size; i++)
        {
            for (unsigned j = 0; j < size; j++)
            {
                auto u = (static_cast<Scalar>(i) + 0.5f) / static_cast<Scalar>(size) * 2.f - 1.f;
                auto v = (static_cast<Scalar>(j) + 0.5f) / static_cast<Scalar>(size) * 2.f - 1.f;
                Vector3 direction = Image2DCube::TextureCoordinateToDirection(face_id, u, v);
                Vector2 eq_uv = utilities::CartesianToEquirectangular(direction);
                Vector3 color = sky_map.Sample(eq_uv);
                this_face.SetPixel(i, j, color);
            }
        }
        memcpy(faces_out + size_t(id) * size * size * 4, this_face.GetRawData(), size_t(size) * size * 8);
    }
    return 0;
}
} // extern "C"

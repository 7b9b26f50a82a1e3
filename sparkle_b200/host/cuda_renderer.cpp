/*
 * cuda_renderer.cpp — host side of the `cuda` pipeline (see cuda_renderer.h for the reference interfaces mirrored).
 * Talks to the GPU only through the C ABI of include/spk_cuda.h. No CPU fallback: if the context cannot be created the
 * renderer fails to initialise, logs, and CreateRenderer returns nullptr (AppFramework.cpp:200-205 convention).
 */
#include "cuda_renderer.h"
#include "sky_cook.h"
#include "ref_math.h"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iterator>
#include <sstream>

#include "bvh_builder.h"
#include "gltf_loader.h"
#include "usd_loader.h"

namespace sparkle_cuda
{
namespace
{
thread_local std::string g_last_error;

void LogError(const std::string &msg)
{
    g_last_error = msg;
    fprintf(stderr, "[sparkle_cuda][error] %s\n", msg.c_str());
}

std::string Lower(std::string s)
{
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return static_cast<char>(std::tolower(c)); });
    return s;
}

constexpr float kPi = 3.14159274101257324f;

} // namespace

const std::string &LastError() { return g_last_error; }
void SetLastError(const std::string &msg) { LogError(msg); } // other translation units of the host library

// ---- RenderConfig ------------------------------------------------------------------------------------------------------
static const char *kPipelineNames[] = {"Cpu", "Gpu", "Forward", "Deferred", "Cuda"};

bool Str2Pipeline(const std::string &name, RenderConfig::Pipeline &out)
{
    const std::string want = Lower(name);
    for (int k = 0; k < 5; k++)
    {
        if (Lower(kPipelineNames[k]) == want)
        {
            out = static_cast<RenderConfig::Pipeline>(k);
            return true;
        }
    }
    return false;
}

const char *Pipeline2Str(RenderConfig::Pipeline p) { return kPipelineNames[static_cast<int>(p)]; }

bool RenderConfig::Set(const std::string &name, const std::string &value)
{
    auto as_u32 = [&](uint32_t &dst) {
        char *end = nullptr;
        unsigned long v = strtoul(value.c_str(), &end, 10);
        if (end == value.c_str() || *end != '\0')
        {
            return false;
        }
        dst = static_cast<uint32_t>(v);
        return true;
    };
    auto as_bool = [&](bool &dst) {
        const std::string v = Lower(value);
        if (v == "1" || v == "true")
        {
            dst = true;
            return true;
        }
        if (v == "0" || v == "false")
        {
            dst = false;
            return true;
        }
        return false;
    };
    bool ok = false;
    if (name == "pipeline")
    {
        ok = Str2Pipeline(value, pipeline);
    }
    else if (name == "spp")
    {
        ok = as_u32(sample_per_pixel);
    }
    else if (name == "max_spp")
    {
        ok = as_u32(max_sample_per_pixel);
    }
    else if (name == "bounce")
    {
        ok = as_u32(max_bounce);
    }
    else if (name == "width")
    {
        ok = as_u32(image_width);
    }
    else if (name == "height")
    {
        ok = as_u32(image_height);
    }
    else if (name == "random_seed_offset")
    {
        ok = as_u32(random_seed_offset);
    }
    else if (name == "spatial_denoise")
    {
        ok = as_bool(spatial_denoise);
    }
    else if (name == "device_tlas")
    {
        ok = as_bool(device_tlas);
    }
    else if (name == "device_bvh")
    {
        ok = as_bool(device_bvh);
    }
    else if (name == "gpus")
    {
        ok = as_u32(gpus) && gpus >= 1 && gpus <= 16;
    }
    else if (name == "gpu_devices")
    {
        // comma-separated CUDA device ordinals, one per GPU of `gpus`
        gpu_devices.clear();
        ok = !value.empty();
        size_t at = 0;
        while (ok && at <= value.size())
        {
            size_t comma = value.find(',', at);
            const std::string item = value.substr(at, comma == std::string::npos ? std::string::npos : comma - at);
            char *end = nullptr;
            long v = strtol(item.c_str(), &end, 10);
            ok = end != item.c_str() && *end == '\0' && v >= 0;
            gpu_devices.push_back(int(v));
            if (comma == std::string::npos)
            {
                break;
            }
            at = comma + 1;
        }
    }
    else if (name == "ssao")
    {
        ok = as_bool(use_ssao);
    }
    else if (name == "prepass")
    {
        ok = as_bool(use_prepass);
    }
    else if (name == "render_scale")
    {
        char *end = nullptr;
        render_scale = strtof(value.c_str(), &end);
        ok = end != value.c_str() && render_scale > 0.f && render_scale <= 1.f;
    }
    else
    {
        LogError("unknown config '" + name + "'");
        return false;
    }
    if (!ok)
    {
        LogError("bad value '" + value + "' for config '" + name + "'");
    }
    return ok;
}

bool RenderConfig::SetFromArgs(const std::vector<std::string> &args)
{
    for (size_t k = 0; k + 1 < args.size(); k += 2)
    {
        if (args[k].rfind("--", 0) != 0)
        {
            LogError("expected --name, got '" + args[k] + "'");
            return false;
        }
        if (!Set(args[k].substr(2), args[k + 1]))
        {
            return false;
        }
    }
    Validate();
    return true;
}

void RenderConfig::Validate()
{
    // RenderConfig.cpp:141-155 with Cuda joining the Cpu/Gpu branch
    if (pipeline == Pipeline::Cpu || pipeline == Pipeline::Gpu || pipeline == Pipeline::Cuda)
    {
        use_prepass = false;
        use_ssao = false;
    }
}

// ---- Transform -----------------------------------------------------------------------------------------------------------
void Transform::Matrices(float fwd[12], float inv[12]) const
{
    // Transform::UpdateTransformData (Transform.cpp:52-64) in the reference's float operation order: ref_math.cpp
    const float t[3] = {translation.x, translation.y, translation.z}, e[3] = {euler.x, euler.y, euler.z},
                s[3] = {scale.x, scale.y, scale.z};
    refmath::TrsMatrices(t, e, s, fwd, inv);
}

// ---- meshes --------------------------------------------------------------------------------------------------------------
MeshData MeshData::UnitCube()
{
    // 6 quads on [-1,1]^3, 24 vertices, per-face normal and tangent, uv per corner
    static const float faces[6][6] = {{0, 0, 1, 1, 0, 0},  {0, 0, -1, -1, 0, 0}, {1, 0, 0, 0, 1, 0},
                                      {-1, 0, 0, 0, -1, 0}, {0, 1, 0, -1, 0, 0},  {0, -1, 0, 1, 0, 0}};
    static const float corner_uv[4][2] = {{0, 1}, {1, 1}, {1, 0}, {0, 0}};
    MeshData m;
    for (const auto &f : faces)
    {
        const float n[3] = {f[0], f[1], f[2]}, t[3] = {f[3], f[4], f[5]};
        const float b[3] = {n[1] * t[2] - n[2] * t[1], n[2] * t[0] - n[0] * t[2], n[0] * t[1] - n[1] * t[0]};
        static const float st[4][2] = {{-1, -1}, {1, -1}, {1, 1}, {-1, 1}}; // BL, BR, TR, TL
        uint32_t base = static_cast<uint32_t>(m.vertices.size() / 3);
        for (int c = 0; c < 4; c++)
        {
            for (int k = 0; k < 3; k++)
            {
                m.vertices.push_back(n[k] + st[c][0] * t[k] + st[c][1] * b[k]);
                m.normals.push_back(n[k]);
            }
            m.tangents.insert(m.tangents.end(), {t[0], t[1], t[2], 1.f});
            m.uvs.insert(m.uvs.end(), {corner_uv[c][0], corner_uv[c][1]});
        }
        // counter-clockwise seen from outside: geometric normal == face normal
        m.indices.insert(m.indices.end(), {base, base + 1, base + 2, base, base + 2, base + 3});
    }
    return m;
}

MeshData MeshData::UnitSphere(uint32_t v_size)
{
    // Mesh::GetUnitSphere layout (io/Mesh.cpp:7-86): (2v+1) x (v+1) vertices, 2v x v quads; zero-area pole triangles
    // are dropped (|a| < Eps never hits them)
    const uint32_t u_size = v_size * 2;
    MeshData m;
    for (uint32_t v = 0; v <= v_size; v++)
    {
        for (uint32_t u = 0; u <= u_size; u++)
        {
            double theta = 2.0 * M_PI * u / u_size + M_PI, phi = M_PI * v / v_size;
            double p[3] = {std::cos(theta) * std::sin(phi), -std::cos(phi), std::sin(theta) * std::sin(phi)};
            double len = std::sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
            float n[3] = {float(p[0] / len), float(p[1] / len), float(p[2] / len)};
            m.vertices.insert(m.vertices.end(), {n[0], n[1], n[2]});
            m.normals.insert(m.normals.end(), {n[0], n[1], n[2]});
            // tangent = (n x major_axis) x n, utilities::GetPossibleMajorAxis (Utilities.h:222-234)
            const float inv_sqrt3 = 0.577350258827209473f;
            float a[3] = {0, 0, 0};
            a[std::fabs(n[0]) < inv_sqrt3 ? 0 : (std::fabs(n[1]) < inv_sqrt3 ? 1 : 2)] = 1.f;
            float c[3] = {n[1] * a[2] - n[2] * a[1], n[2] * a[0] - n[0] * a[2], n[0] * a[1] - n[1] * a[0]};
            m.tangents.insert(m.tangents.end(), {c[1] * n[2] - c[2] * n[1], c[2] * n[0] - c[0] * n[2],
                                                 c[0] * n[1] - c[1] * n[0], 1.f});
            m.uvs.insert(m.uvs.end(), {float(u) / float(u_size), float(v) / float(v_size)});
        }
    }
    auto area2 = [&](uint32_t i0, uint32_t i1, uint32_t i2) {
        const float *a = &m.vertices[3 * i0], *b = &m.vertices[3 * i1], *c = &m.vertices[3 * i2];
        double e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        double x = e1[1] * e2[2] - e1[2] * e2[1], y = e1[2] * e2[0] - e1[0] * e2[2], z = e1[0] * e2[1] - e1[1] * e2[0];
        return x * x + y * y + z * z;
    };
    for (uint32_t y = 0; y < v_size; y++)
    {
        for (uint32_t x = 0; x < u_size; x++)
        {
            uint32_t vi = y * (u_size + 1) + x;
            uint32_t tri[2][3] = {{vi, vi + u_size + 1, vi + 1}, {vi + 1, vi + u_size + 1, vi + u_size + 2}};
            for (auto &t : tri)
            {
                if (area2(t[0], t[1], t[2]) > 1e-24)
                {
                    m.indices.insert(m.indices.end(), {t[0], t[1], t[2]});
                }
            }
        }
    }
    return m;
}

// ---- camera --------------------------------------------------------------------------------------------------------------
void CameraProxy::SetupOrbit(Vec3 center, float radius, float pitch_deg, float yaw_deg, float focal_length,
                             float sensor_height, float aperture, float exposure)
{
    // kept as given: Update() turns them into posture and focus plane once the resolution is known (ref_math.cpp follows
    // OrbitCameraComponent::UpdateTransform, CalculateRenderAttribute and CameraRenderProxy::Update / OnTransformDirty)
    orbit_center_ = center;
    orbit_radius_ = radius;
    orbit_pitch_ = pitch_deg;
    orbit_yaw_ = yaw_deg;
    focal_length_ = focal_length;
    sensor_height_ = sensor_height;
    aperture_ = aperture;
    exposure_ = exposure;
    attribute_dirty_ = true;
    transform_dirty_ = true;
}

void CameraProxy::MarkPixelDirty()
{
    pixels_dirty_ = true;
    cumulated_sample_count_ = 0;
    pending_sample_count_ = 0;
}

void CameraProxy::Update(const RenderConfig &config)
{
    if (attribute_dirty_ || transform_dirty_)
    {
        const float c[3] = {orbit_center_.x, orbit_center_.y, orbit_center_.z};
        refmath::OrbitCamera(c, orbit_radius_, orbit_pitch_, orbit_yaw_, focal_length_, sensor_height_, aperture_,
                             exposure_, config.image_width, config.image_height, &camera_);
        attribute_dirty_ = false;
        transform_dirty_ = false;
        pixels_dirty_ = true;
    }
    if (pixels_dirty_)
    {
        cumulated_sample_count_ = 0;
        pending_sample_count_ = 0;
    }
    cumulated_sample_count_ = std::min(config.max_sample_per_pixel, cumulated_sample_count_ + pending_sample_count_);
    pending_sample_count_ = 0;
}

// ---- SceneProxy ------------------------------------------------------------------------------------------------------------
struct SceneProxy::Impl
{
    struct MeshEntry
    {
        MeshData data;
        SpkHostBvh *blas = nullptr;
        float lo[3], hi[3];
    };
    std::vector<MeshEntry> meshes;
    std::vector<SpkMeshDesc> mesh_descs;
    std::vector<std::vector<uint8_t>> texture_pixels;
    SpkHostBvh *tlas = nullptr;
    bool has_sky = false;
    Vec3 sky_color{0.5f, 0.7f, 1.0f}; // SkyLight.h:101
    std::vector<uint16_t> sky_faces; // cooked sky map: 6 faces of sky_face_size^2 RGBAFloat16 pixels, else empty
    uint32_t sky_face_size = 0;
    ~Impl()
    {
        for (auto &m : meshes)
        {
            spk_host_bvh_free(m.blas);
        }
        spk_host_bvh_free(tlas);
    }
};

SceneProxy::SceneProxy() : impl_(std::make_unique<Impl>()) {}
SceneProxy::~SceneProxy() = default;

uint32_t SceneProxy::AddMesh(MeshData mesh)
{
    Impl::MeshEntry e;
    e.data = std::move(mesh);
    const uint32_t faces = static_cast<uint32_t>(e.data.indices.size() / 3);
    e.blas = spk_host_bvh_build_triangles(e.data.vertices.data(), e.data.indices.data(), faces, 4, 32);
    for (int k = 0; k < 3; k++)
    {
        e.lo[k] = 3.4e38f;
        e.hi[k] = -3.4e38f;
    }
    for (size_t v = 0; v < e.data.vertices.size(); v += 3)
    {
        for (int k = 0; k < 3; k++)
        {
            e.lo[k] = std::min(e.lo[k], e.data.vertices[v + k]);
            e.hi[k] = std::max(e.hi[k], e.data.vertices[v + k]);
        }
    }
    impl_->meshes.push_back(std::move(e));
    return static_cast<uint32_t>(impl_->meshes.size() - 1);
}

uint32_t SceneProxy::AddTexture(const void *pixels, uint32_t width, uint32_t height, SpkPixelFormat format)
{
    const size_t bpp = format == SPK_FMT_RGBA32F ? 16 : format == SPK_FMT_RGBA16F ? 8 : 4;
    impl_->texture_pixels.emplace_back(static_cast<const uint8_t *>(pixels),
                                       static_cast<const uint8_t *>(pixels) + size_t(width) * height * bpp);
    SpkTextureDesc t{};
    t.width = width;
    t.height = height;
    t.format = format;
    textures_.push_back(t);
    return static_cast<uint32_t>(textures_.size() - 1);
}

uint32_t SceneProxy::AddMaterial(const SpkMaterialDesc &material)
{
    materials_.push_back(material);
    return static_cast<uint32_t>(materials_.size() - 1);
}

static SpkMaterialDesc BaseMaterial(Vec3 color)
{
    // io/Material.h:7-22 defaults: base colour 1, emissive 0, metallic 1, roughness 1, eta 0, no textures
    SpkMaterialDesc m{};
    m.base_color_tex = m.emissive_tex = m.metallic_roughness_tex = m.normal_tex = SPK_INVALID_ID;
    m.metallic = 1.f;
    m.roughness = 1.f;
    m.type = SPK_MAT_PBR;
    m.base_color[0] = color.x;
    m.base_color[1] = color.y;
    m.base_color[2] = color.z;
    return m;
}

uint32_t SceneProxy::AddLambertian(Vec3 color)
{
    SpkMaterialDesc m = BaseMaterial(color);
    m.metallic = 0.f;
    return AddMaterial(m);
}

uint32_t SceneProxy::AddMetal(Vec3 color, float roughness)
{
    SpkMaterialDesc m = BaseMaterial(color);
    m.roughness = roughness;
    return AddMaterial(m);
}

uint32_t SceneProxy::AddDielectric(float eta)
{
    SpkMaterialDesc m = BaseMaterial(Vec3{1.f, 1.f, 1.f});
    m.eta = eta;
    m.type = SPK_MAT_DIELECTRIC;
    return AddMaterial(m);
}

static void WorldBox(const float fwd[12], const float lo[3], const float hi[3], std::array<float, 6> &box)
{
    // AABB::TransformTo (core/math/AABB.h:57-78): the eight corners transformed, min/max taken
    for (int k = 0; k < 3; k++)
    {
        box[k] = 3.4e38f;
        box[3 + k] = -3.4e38f;
    }
    for (int c = 0; c < 8; c++)
    {
        const float p[3] = {c & 1 ? hi[0] : lo[0], c & 2 ? hi[1] : lo[1], c & 4 ? hi[2] : lo[2]};
        for (int k = 0; k < 3; k++)
        {
            float w = fwd[4 * k] * p[0] + fwd[4 * k + 1] * p[1] + fwd[4 * k + 2] * p[2] + fwd[4 * k + 3];
            box[k] = std::min(box[k], w);
            box[3 + k] = std::max(box[3 + k], w);
        }
    }
}

uint32_t SceneProxy::AddMeshPrimitive(uint32_t mesh, uint32_t material, const Transform &t)
{
    SpkPrimitiveDesc p{};
    p.kind = SPK_PRIM_MESH;
    p.mesh = mesh;
    p.material = material;
    t.Matrices(p.fwd, p.inv);
    primitives_.push_back(p);
    std::array<float, 6> box;
    WorldBox(p.fwd, impl_->meshes[mesh].lo, impl_->meshes[mesh].hi, box);
    world_boxes_.push_back(box);
    changes_.emplace_back(PrimitiveChangeType::New, NumPrimitives() - 1);
    return NumPrimitives() - 1;
}

uint32_t SceneProxy::AddMeshPrimitiveMatrix(uint32_t mesh, uint32_t material, const float fwd[12], const float inv[12])
{
    SpkPrimitiveDesc p{};
    p.kind = SPK_PRIM_MESH;
    p.mesh = mesh;
    p.material = material;
    memcpy(p.fwd, fwd, sizeof(p.fwd));
    memcpy(p.inv, inv, sizeof(p.inv));
    primitives_.push_back(p);
    std::array<float, 6> box;
    WorldBox(p.fwd, impl_->meshes[mesh].lo, impl_->meshes[mesh].hi, box);
    world_boxes_.push_back(box);
    changes_.emplace_back(PrimitiveChangeType::New, NumPrimitives() - 1);
    return NumPrimitives() - 1;
}

uint32_t SceneProxy::AddSphere(Vec3 center, float radius, uint32_t material)
{
    SpkPrimitiveDesc p{};
    p.kind = SPK_PRIM_SPHERE;
    p.material = material;
    p.radius2 = radius * radius; // scaled_radius_2_ (SphereRenderProxy.h:33-36)
    Transform t;
    t.translation = center;
    t.scale = Vec3{radius, radius, radius};
    t.Matrices(p.fwd, p.inv);
    primitives_.push_back(p);
    world_boxes_.push_back({center.x - radius, center.y - radius, center.z - radius, center.x + radius,
                            center.y + radius, center.z + radius});
    changes_.emplace_back(PrimitiveChangeType::New, NumPrimitives() - 1);
    return NumPrimitives() - 1;
}

void SceneProxy::MovePrimitive(uint32_t primitive, const Transform &t)
{
    SpkPrimitiveDesc &p = primitives_[primitive];
    t.Matrices(p.fwd, p.inv);
    if (p.kind == SPK_PRIM_SPHERE)
    {
        const float r = std::sqrt(p.radius2);
        world_boxes_[primitive] = {t.translation.x - r, t.translation.y - r, t.translation.z - r,
                                   t.translation.x + r, t.translation.y + r, t.translation.z + r};
    }
    else
    {
        WorldBox(p.fwd, impl_->meshes[p.mesh].lo, impl_->meshes[p.mesh].hi, world_boxes_[primitive]);
    }
    changes_.emplace_back(PrimitiveChangeType::Update, primitive);
}

void SceneProxy::SetSky(Vec3 color)
{
    impl_->has_sky = true;
    impl_->sky_color = color;
}

bool SceneProxy::SetSkyMapFile(const std::string &hdr_path, uint32_t cube_size)
{
    std::ifstream in(hdr_path, std::ios::binary);
    if (!in)
    {
        SetLastError("SkyLight: failed to read image file " + hdr_path);
        return false;
    }
    std::vector<uint8_t> bytes((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    uint32_t w = 0, h = 0;
    std::vector<uint16_t> equirect;
    std::string error;
    if (!DecodeRadianceHdr(bytes.data(), bytes.size(), w, h, equirect, error))
    {
        SetLastError("SkyLight: failed to decode image " + hdr_path + ": " + error);
        return false;
    }
    const uint32_t size = cube_size ? cube_size : kSkyCubeSize;
    impl_->sky_faces.assign(size_t(6) * size * size * 4, 0);
    CookSkyCube(equirect.data(), w, h, size, impl_->sky_faces.data());
    impl_->sky_face_size = size;
    impl_->has_sky = true;
    return true;
}

void SceneProxy::RebuildDesc(bool structural)
{
    if (structural)
    {
        impl_->mesh_descs.clear();
        for (auto &m : impl_->meshes)
        {
            SpkMeshDesc d{};
            d.vertices = m.data.vertices.data();
            d.normals = m.data.normals.data();
            d.tangents = m.data.tangents.data();
            d.uvs = m.data.uvs.empty() ? nullptr : m.data.uvs.data();
            d.indices = m.data.indices.data();
            d.num_vertices = static_cast<uint32_t>(m.data.vertices.size() / 3);
            d.num_faces = static_cast<uint32_t>(m.data.indices.size() / 3);
            d.blas = SpkBvhDesc{spk_host_bvh_nodes(m.blas), spk_host_bvh_num_nodes(m.blas), 32,
                                spk_host_bvh_prim_ids(m.blas), spk_host_bvh_num_prims(m.blas), 4};
            impl_->mesh_descs.push_back(d);
        }
        for (size_t t = 0; t < textures_.size(); t++)
        {
            textures_[t].pixels = impl_->texture_pixels[t].data();
        }
    }
    // TLAS::Build (SceneRenderProxy.cpp:31-69): one box and centre per primitive, rebuilt on every change
    spk_host_bvh_free(impl_->tlas);
    impl_->tlas = nullptr;
    float extent = 0.f;
    if (!primitives_.empty())
    {
        impl_->tlas = spk_host_bvh_build(world_boxes_[0].data(), nullptr, NumPrimitives(), 1, 32);
        for (const auto &b : world_boxes_)
        {
            for (float v : b)
            {
                extent = std::max(extent, std::fabs(v));
            }
        }
    }
    desc_ = SpkSceneDesc{};
    desc_.abi_version = SPK_ABI_VERSION;
    desc_.num_meshes = static_cast<uint32_t>(impl_->mesh_descs.size());
    desc_.meshes = impl_->mesh_descs.data();
    desc_.num_primitives = NumPrimitives();
    desc_.primitives = primitives_.data();
    desc_.num_materials = static_cast<uint32_t>(materials_.size());
    desc_.materials = materials_.data();
    desc_.num_textures = static_cast<uint32_t>(textures_.size());
    desc_.textures = textures_.data();
    if (impl_->tlas)
    {
        desc_.tlas = SpkBvhDesc{spk_host_bvh_nodes(impl_->tlas), spk_host_bvh_num_nodes(impl_->tlas), 32,
                                spk_host_bvh_prim_ids(impl_->tlas), spk_host_bvh_num_prims(impl_->tlas), 4};
    }
    desc_.sky.has_sky = impl_->has_sky ? 1u : 0u;
    desc_.sky.color[0] = impl_->sky_color.x;
    desc_.sky.color[1] = impl_->sky_color.y;
    desc_.sky.color[2] = impl_->sky_color.z;
    desc_.sky.has_map = impl_->sky_faces.empty() ? 0u : 1u;
    for (uint32_t f = 0; f < 6 && desc_.sky.has_map; f++)
    {
        const size_t face_halves = size_t(impl_->sky_face_size) * impl_->sky_face_size * 4;
        desc_.sky.faces[f] = SpkTextureDesc{impl_->sky_faces.data() + f * face_halves, impl_->sky_face_size, impl_->sky_face_size,
                                            SPK_FMT_RGBA16F, 0};
    }
    desc_.world_extent = extent;
}

const std::vector<std::pair<PrimitiveChangeType, uint32_t>> &SceneProxy::Update(const RenderConfig &config)
{
    if (!changes_.empty() && config.NeedBvh())
    {
        bool structural = false;
        for (const auto &c : changes_)
        {
            structural |= c.first != PrimitiveChangeType::Update;
        }
        RebuildDesc(structural || desc_.abi_version == 0);
        camera_.MarkPixelDirty(); // SceneRenderProxy.cpp:201-235: any primitive change restarts accumulation
    }
    camera_.Update(config);
    return changes_;
}

// ---- Renderer ----------------------------------------------------------------------------------------------------------------
std::unique_ptr<Renderer> Renderer::CreateRenderer(const RenderConfig &config, SceneProxy *scene, int device)
{
    std::unique_ptr<Renderer> renderer;
    switch (config.pipeline)
    {
    case RenderConfig::Pipeline::Cuda:
        renderer = std::make_unique<CudaRenderer>(config, scene, device);
        break;
    default:
        LogError(std::string("pipeline '") + Pipeline2Str(config.pipeline) +
                 "' belongs to the reference's own renderers; this library provides 'cuda' only");
        return nullptr;
    }
    if (!renderer->InitRenderResources())
    {
        return nullptr; // no fallback: the caller fails initialisation (AppFramework.cpp:200-205)
    }
    return renderer;
}

void Renderer::Tick()
{
    scene_render_proxy_->Update(render_config_);
    Update();
    scene_render_proxy_->EndUpdate();
}

CudaRenderer::CudaRenderer(const RenderConfig &config, SceneProxy *scene, int device)
    : Renderer(config, scene), device_(device)
{
}

CudaRenderer::~CudaRenderer()
{
    if (group_)
    {
        spk_cuda_group_destroy(group_);
    }
    for (SpkContext *c : ctxs_)
    {
        spk_cuda_destroy(c);
    }
}

bool CudaRenderer::InitRenderResources()
{
    // one context per GPU of the `gpus` cvar; the first is the root (tone-mapping target and read-back)
    uint32_t gpus = std::max(1u, render_config_.gpus);
    if (gpus > 1 && render_config_.spatial_denoise)
    {
        LogError("CudaRenderer: spatial_denoise works on one accumulator: running on one GPU");
        gpus = 1;
    }
    for (uint32_t g = 0; g < gpus; g++)
    {
        SpkCreateInfo info{};
        info.abi_version = SPK_ABI_VERSION;
        info.device = g < render_config_.gpu_devices.size() ? render_config_.gpu_devices[g] : device_ + int(g);
        info.width = render_config_.image_width;
        info.height = render_config_.image_height;
        SpkContext *c = nullptr;
        int rc = spk_cuda_create(&info, &c);
        if (rc != SPK_OK)
        {
            LogError(std::string("CudaRenderer: spk_cuda_create failed (") + std::to_string(rc) + "): " +
                     spk_cuda_last_error());
            failed_ = true;
            return false;
        }
        ctxs_.push_back(c);
    }
    ctx_ = ctxs_[0];
    if (gpus > 1 && spk_cuda_group_create_local(ctxs_.data(), gpus, &group_) != SPK_OK)
    {
        LogError(std::string("CudaRenderer: spk_cuda_group_create_local failed: ") + spk_cuda_last_error());
        failed_ = true;
        return false;
    }
    return true;
}

void CudaRenderer::Update()
{
    // device scene (re)build keyed on the change list, like GPURenderer.cpp:331-385: New/Remove/Move rebuild everything,
    // Update refreshes instance transforms and the TLAS only
    if (!ctx_ || failed_)
    {
        return;
    }
    bool structural = !scene_uploaded_, moved = false;
    // the change list of this frame: built by SceneProxy::Update just before, valid until EndUpdate
    for (const auto &c : scene_render_proxy_->Changes())
    {
        structural |= c.first != PrimitiveChangeType::Update;
        moved |= c.first == PrimitiveChangeType::Update;
    }
    const SpkSceneDesc &d = scene_render_proxy_->Desc();
    int rc = SPK_OK;
    // device_bvh: the same descriptor with every tree left out, so that spk_cuda_upload_scene builds them on the device
    SpkSceneDesc without_trees = d;
    std::vector<SpkMeshDesc> meshes_without_trees;
    if (render_config_.device_bvh)
    {
        meshes_without_trees.assign(d.meshes, d.meshes + d.num_meshes);
        for (SpkMeshDesc &m : meshes_without_trees)
        {
            m.blas = SpkBvhDesc{};
        }
        without_trees.meshes = meshes_without_trees.data();
        without_trees.tlas = SpkBvhDesc{};
    }
    const SpkSceneDesc &up = render_config_.device_bvh ? without_trees : d;
    for (SpkContext *c : ctxs_) // the scene is replicated on every GPU
    {
        if (rc != SPK_OK)
        {
            break;
        }
        if (structural)
        {
            rc = spk_cuda_upload_scene(c, &up);
        }
        else if (moved)
        {
            rc = spk_cuda_update_instances(c, d.primitives, d.num_primitives,
                                           (render_config_.device_tlas || render_config_.device_bvh) ? nullptr : &d.tlas);
            if (rc != SPK_OK)
            {
                // e.g. the rebuilt TLAS has more nodes than the uploaded one: a full upload always fits
                rc = spk_cuda_upload_scene(c, &up);
            }
        }
    }
    if (structural)
    {
        scene_uploaded_ = rc == SPK_OK;
    }
    if (rc != SPK_OK)
    {
        LogError(std::string("CudaRenderer: scene upload failed: ") + spk_cuda_last_error());
        failed_ = true;
    }
}

void CudaRenderer::Render()
{
    if (!ctx_ || failed_ || !scene_uploaded_)
    {
        return;
    }
    // CPURenderer::Render (CPURenderer.cpp:116-137, 206-207): re-fetch the camera every frame; on NeedClear zero the
    // accumulation, ClearPixels, restart the sample index; after the dispatch AccumulateSample(n)
    CameraProxy &camera = scene_render_proxy_->GetCamera();
    int rc = SPK_OK;
    if (camera.NeedClear())
    {
        for (SpkContext *c : ctxs_)
        {
            rc = rc == SPK_OK ? spk_cuda_set_camera(c, &camera.Get()) : rc;
            rc = rc == SPK_OK ? spk_cuda_reset_accumulation(c) : rc;
        }
        camera.ClearPixels();
        dispatched_sample_count_ = 0;
    }
    const uint32_t remaining = render_config_.max_sample_per_pixel > camera.GetCumulatedSampleCount()
                                   ? render_config_.max_sample_per_pixel - camera.GetCumulatedSampleCount()
                                   : 0u;
    // the cuda pipeline may render n > 1 samples per call (the gpu pipeline does, GPURenderer.cpp:409-471)
    const uint32_t n = std::min(std::max(1u, render_config_.sample_per_pixel), std::max(1u, remaining));
    SpkRenderParams p{};
    p.first_sample = dispatched_sample_count_ + render_config_.random_seed_offset; // frame_seed (CPURenderer.cpp:432)
    p.sample_count = n;
    p.sample_stride = 1;
    p.max_bounce = render_config_.max_bounce;
    p.output_limit = 6.f; // CameraRenderProxy::OutputLimit (CameraRenderProxy.h:16)
    p.accum_mode = SPK_ACCUM_RUNNING_AVERAGE; // the reference's temporal lerp (CPURenderer.cpp:530-544)
    // RenderConfig::spatial_denoise: g-buffer + neighbour fill per sample (CPURenderer.cpp:511-526)
    p.flags = render_config_.spatial_denoise ? SPK_FLAG_SPATIAL_DENOISE : 0u;
    // past max_spp the camera's cumulated count stays at the cap: an exponential moving average (CameraRenderProxy.cpp:70-73)
    p.cumulated_cap = render_config_.max_sample_per_pixel;
    if (group_)
    {
        // gpus > 1: GPU g takes samples {g, g+G, ..} of this frame's n into its own SUM accumulator; the sums are combined
        // when the frame is read (float association is the only difference from the running lerp, ~1e-7 relative).
        // Past max_spp nothing more is added: the converged frame stays as it is.
        const uint32_t G = uint32_t(ctxs_.size());
        p.accum_mode = SPK_ACCUM_SUM;
        p.cumulated_cap = 0;
        p.flags |= SPK_FLAG_AUTO_TRACER;
        for (uint32_t g = 0; g < G && rc == SPK_OK && remaining > 0; g++)
        {
            SpkRenderParams q = p;
            q.first_sample = p.first_sample + g;
            q.sample_stride = G;
            q.sample_count = n > g ? (n - g + G - 1) / G : 0u;
            rc = spk_cuda_render(ctxs_[g], &q);
        }
    }
    else
    {
        rc = rc == SPK_OK ? spk_cuda_render(ctx_, &p) : rc;
    }
    if (rc != SPK_OK)
    {
        LogError(std::string("CudaRenderer::Render: ") + spk_cuda_last_error());
        failed_ = true;
        return;
    }
    if (!(group_ && remaining == 0))
    {
        dispatched_sample_count_ += n; // (gpus > 1 adds nothing past max_spp: the divisor of the sums stays)
    }
    camera.AccumulateSample(n);
    if (!screenshot_path_.empty() && IsReadyForAutoScreenshot())
    {
        std::vector<uint8_t> img;
        if (ReadSrgb8(img))
        {
            WritePng(screenshot_path_, img.data(), render_config_.image_width, render_config_.image_height);
        }
        screenshot_path_.clear();
    }
}

bool CudaRenderer::IsReadyForAutoScreenshot() const
{
    // CPURenderer.cpp:32-36. The camera's cumulated count advances in the NEXT Update, as in the reference.
    return Renderer::IsReadyForAutoScreenshot() && !failed_ &&
           scene_render_proxy_->GetCamera().GetCumulatedSampleCount() >= render_config_.max_sample_per_pixel;
}

bool CudaRenderer::ReadLinear(std::vector<float> &out)
{
    out.resize(size_t(render_config_.image_width) * render_config_.image_height * 4);
    if (group_)
    {
        return spk_cuda_group_resolve(group_, SPK_RESOLVE_LINEAR_F32, dispatched_sample_count_, out.data(), out.size() * 4) ==
               SPK_OK;
    }
    return ctx_ && spk_cuda_resolve(ctx_, SPK_RESOLVE_LINEAR_F32, 0, out.data(), out.size() * 4) == SPK_OK;
}

bool CudaRenderer::ReadSrgb8(std::vector<uint8_t> &out)
{
    out.resize(size_t(render_config_.image_width) * render_config_.image_height * 4);
    if (group_)
    {
        return spk_cuda_group_resolve(group_, SPK_RESOLVE_TONEMAP_SRGB8, dispatched_sample_count_, out.data(), out.size()) ==
               SPK_OK;
    }
    return ctx_ && spk_cuda_resolve(ctx_, SPK_RESOLVE_TONEMAP_SRGB8, 0, out.data(), out.size()) == SPK_OK;
}

// ---- PNG ---------------------------------------------------------------------------------------------------------------------
namespace
{
uint32_t Crc32(const uint8_t *data, size_t n, uint32_t crc = 0)
{
    static uint32_t table[256];
    static bool init = false;
    if (!init)
    {
        for (uint32_t i = 0; i < 256; i++)
        {
            uint32_t c = i;
            for (int k = 0; k < 8; k++)
            {
                c = c & 1 ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            }
            table[i] = c;
        }
        init = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++)
    {
        crc = table[(crc ^ data[i]) & 0xFF] ^ (crc >> 8);
    }
    return ~crc;
}

void PutBe32(std::vector<uint8_t> &v, uint32_t x)
{
    v.insert(v.end(), {uint8_t(x >> 24), uint8_t(x >> 16), uint8_t(x >> 8), uint8_t(x)});
}

void PutChunk(std::vector<uint8_t> &png, const char tag[4], const std::vector<uint8_t> &payload)
{
    PutBe32(png, static_cast<uint32_t>(payload.size()));
    std::vector<uint8_t> body(tag, tag + 4);
    body.insert(body.end(), payload.begin(), payload.end());
    png.insert(png.end(), body.begin(), body.end());
    PutBe32(png, Crc32(body.data(), body.size()));
}
} // namespace

bool WritePng(const std::string &path, const uint8_t *rgba, uint32_t width, uint32_t height)
{
    std::vector<uint8_t> raw; // filter byte 0 + row
    raw.reserve(size_t(height) * (size_t(width) * 4 + 1));
    for (uint32_t y = 0; y < height; y++)
    {
        raw.push_back(0);
        raw.insert(raw.end(), rgba + size_t(y) * width * 4, rgba + size_t(y + 1) * width * 4);
    }
    std::vector<uint8_t> z = {0x78, 0x01}; // zlib header, stored (uncompressed) deflate blocks
    uint32_t a = 1, b = 0;                 // adler32
    for (uint8_t c : raw)
    {
        a = (a + c) % 65521u;
        b = (b + a) % 65521u;
    }
    for (size_t off = 0; off < raw.size() || off == 0; off += 65535)
    {
        size_t n = std::min<size_t>(65535, raw.size() - off);
        bool last = off + n >= raw.size();
        z.push_back(last ? 1 : 0);
        z.insert(z.end(), {uint8_t(n), uint8_t(n >> 8), uint8_t(~n), uint8_t((~n) >> 8)});
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        if (last)
        {
            break;
        }
    }
    PutBe32(z, (b << 16) | a);
    std::vector<uint8_t> png = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr;
    PutBe32(ihdr, width);
    PutBe32(ihdr, height);
    ihdr.insert(ihdr.end(), {8, 6, 0, 0, 0});
    PutChunk(png, "IHDR", ihdr);
    PutChunk(png, "IDAT", z);
    PutChunk(png, "IEND", {});
    FILE *f = fopen(path.c_str(), "wb");
    if (!f)
    {
        LogError("cannot write " + path);
        return false;
    }
    bool ok = fwrite(png.data(), 1, png.size(), f) == png.size();
    fclose(f);
    return ok;
}

// ---- BuildTestScene ----------------------------------------------------------------------------------------------------------
namespace
{
// XoshiroCpp::Xoshiro128Plus seeded through SplitMix64 (XoshiroCpp.hpp:684-690,1305-1330): the fixed-seed stream
// sampler::RandomUnit<true> (Sampler.h:32-41) that places the random spheres
struct FixedSeedRng
{
    uint32_t s[4];
    explicit FixedSeedRng(uint64_t seed)
    {
        for (auto &w : s)
        {
            seed += 0x9E3779B97F4A7C15ull;
            uint64_t z = seed;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            w = static_cast<uint32_t>(z ^ (z >> 31));
        }
    }
    float Unit()
    {
        const uint32_t result = s[0] + s[3], t = s[1] << 9;
        s[2] ^= s[0];
        s[3] ^= s[1];
        s[1] ^= s[2];
        s[0] ^= s[3];
        s[2] ^= t;
        s[3] = (s[3] << 11) | (s[3] >> 21);
        return static_cast<float>(result) / static_cast<float>(UINT32_MAX);
    }
    void UnitDisk(float &x, float &y) // Sampler.h:43-53
    {
        float rand = Unit();
        float theta = Unit() * 2.0f * kPi;
        float r = std::sqrt(rand);
        x = r * std::cos(theta);
        y = r * std::sin(theta);
    }
};

struct Box
{
    float lo[3], hi[3];
};

bool Collides(const Box &a, const std::vector<Box> &others)
{
    // Scene::BoxCollides -> AABB::Intersect (Scene.cpp:189-194, AABB.h:50-55): strict overlap of centre/half-extent
    for (const Box &b : others)
    {
        bool overlap = true;
        for (int k = 0; k < 3; k++)
        {
            float c0 = (a.lo[k] + a.hi[k]) * 0.5f, h0 = (a.hi[k] - a.lo[k]) * 0.5f;
            float c1 = (b.lo[k] + b.hi[k]) * 0.5f, h1 = (b.hi[k] - b.lo[k]) * 0.5f;
            overlap &= std::fabs(c0 - c1) < h0 + h1;
        }
        if (overlap)
        {
            return true;
        }
    }
    return false;
}
} // namespace

void BuildTestScene(SceneProxy &scene)
{
    // tests/usd/UsdRoundTripTest.cpp:25-105: floor, glass / lambert / gold spheres, ten random spheres; the two glTF
    // models and the HDR sky are absent from the mount -> gradient sky
    static const Vec3 metals[5] = {{1.0f, 0.7f, 0.29f}, {0.56f, 0.57f, 0.58f}, {0.95f, 0.64f, 0.54f},
                                   {0.92f, 0.92f, 0.92f}, {0.95f, 0.93f, 0.88f}}; // MaterialManager.cpp:21-49
    const uint32_t white = scene.AddLambertian(Vec3{1, 1, 1});
    const uint32_t glass = scene.AddDielectric(1.6f);
    const uint32_t gold = scene.AddMetal(metals[0], 0.2f);
    const uint32_t cube = scene.AddMesh(MeshData::UnitCube());
    Transform floor;
    floor.translation = Vec3{0, 0, -0.001f};
    floor.scale = Vec3{50.f, 50.f, 0.001f};
    scene.AddMeshPrimitive(cube, white, floor);
    scene.AddSphere(Vec3{0, 0, 2}, 2.f, glass);
    scene.AddSphere(Vec3{-4, 4, 2}, 2.f, white);
    scene.AddSphere(Vec3{4, -4, 2}, 2.f, gold);
    std::vector<Box> boxes = {{{-50, -50, -0.002f}, {50, 50, 0.f}},
                              {{-2, -2, 0}, {2, 2, 4}},
                              {{-6, 2, 0}, {-2, 6, 4}},
                              {{2, -6, 0}, {6, -2, 4}}};
    FixedSeedRng rng(42); // GenerateRandomSpheres (SceneManager.cpp:51-105)
    for (int i = 0; i < 10; i++)
    {
        const float radius = 0.4f + (1.f - 0.4f) * rng.Unit();
        const float kind = rng.Unit();
        uint32_t material;
        if (kind < 0.7f)
        {
            float r = rng.Unit(), g = rng.Unit(), b = rng.Unit();
            material = scene.AddLambertian(Vec3{r, g, b});
        }
        else if (kind < 0.7f + 0.1f)
        {
            material = scene.AddDielectric(rng.Unit() * 2.f + 1.3f);
        }
        else
        {
            material = scene.AddMetal(metals[static_cast<int>(rng.Unit() * 4.f)], 0.2f);
        }
        Box box;
        Vec3 pos;
        do
        {
            float x, y;
            rng.UnitDisk(x, y);
            pos = Vec3{x * 13.f, y * 13.f, radius};
            const float p[3] = {pos.x, pos.y, pos.z};
            for (int k = 0; k < 3; k++)
            {
                box.lo[k] = p[k] - radius;
                box.hi[k] = p[k] + radius;
            }
        } while (Collides(box, boxes));
        scene.AddSphere(pos, radius, material);
        boxes.push_back(box);
    }
    scene.SetSky(Vec3{0.5f, 0.7f, 1.0f});
    scene.GetCamera().SetupOrbit(Vec3{0, 0, 0}, 25.f, 10.f, -20.f); // UsdRoundTripTest.cpp:31-36
}
} // namespace sparkle_cuda

// ---- C shims -------------------------------------------------------------------------------------------------------------------
using namespace sparkle_cuda;

struct SpkHostApp
{
    RenderConfig config;
    SceneProxy scene;
    std::unique_ptr<Renderer> renderer;
    uint32_t frames = 0;
};

extern "C"
{
const char *spk_host_last_error(void) { return LastError().c_str(); }

int spk_host_parse_pipeline(const char *name)
{
    RenderConfig::Pipeline p;
    return name && Str2Pipeline(name, p) ? static_cast<int>(p) : -1;
}

SpkHostApp *spk_host_app_create(const char *args, int device)
{
    auto app = std::make_unique<SpkHostApp>();
    std::vector<std::string> tokens, config_args;
    std::istringstream in(args ? args : "");
    for (std::string t; in >> t;)
    {
        tokens.push_back(t);
    }
    std::string scene_name = "testscene", sky_map;
    GltfOptions gltf_options;
    for (size_t k = 0; k + 1 < tokens.size(); k += 2)
    {
        if (tokens[k] == "--scene")
        {
            scene_name = tokens[k + 1];
        }
        else if (tokens[k] == "--sky_map")
        {
            sky_map = tokens[k + 1];
        }
        else if (tokens[k] == "--gltf_instancing")
        {
            gltf_options.instancing = tokens[k + 1] == "true" || tokens[k + 1] == "1";
        }
        else
        {
            config_args.push_back(tokens[k]);
            config_args.push_back(tokens[k + 1]);
        }
    }
    if (tokens.size() % 2 != 0 || !app->config.SetFromArgs(config_args))
    {
        return nullptr;
    }
    const bool gltf = scene_name.size() > 5 && scene_name.compare(scene_name.size() - 5, 5, ".gltf") == 0;
    const bool usd = scene_name.size() > 5 && scene_name.compare(scene_name.size() - 5, 5, ".usda") == 0;
    if (scene_name != "testscene" && scene_name != "empty" && !gltf && !usd)
    {
        LogError("unknown scene '" + scene_name + "' (testscene, empty, a .gltf path or a .usda path)");
        return nullptr;
    }
    if (usd)
    {
        // SceneManager::LoadScene -> USDLoader::Load (the default scene of the app is resources/packed/TestScene.usda):
        // sky and camera come from the file when it has a DomeLight / Camera, else the app defaults
        app->scene.SetSky(Vec3{0.5f, 0.7f, 1.0f});
        app->scene.GetCamera().SetupOrbit(Vec3{0, 0, 0}, 25.f, 10.f, -20.f);
        if (!LoadUsd(scene_name, app->scene))
        {
            return nullptr;
        }
    }
    else if (gltf)
    {
        // SceneManager::LoadScene -> GLTFLoader::Load (AppFramework.cpp:200-205: a failed load logs and fails init)
        if (!LoadGltf(scene_name, app->scene, nullptr, gltf_options))
        {
            return nullptr;
        }
        app->scene.SetSky(Vec3{0.5f, 0.7f, 1.0f});
        app->scene.GetCamera().SetupOrbit(Vec3{0, 0, 0}, 25.f, 10.f, -20.f);
    }
    else if (scene_name == "testscene")
    {
        BuildTestScene(app->scene);
    }
    else
    {
        app->scene.SetSky(Vec3{0.5f, 0.7f, 1.0f});
        app->scene.GetCamera().SetupOrbit(Vec3{0, 0, 0}, 25.f, 10.f, -20.f);
    }
    // SceneManager::AddDefaultSky (SceneManager.cpp:229-243): the default sky of every scene that does not bring its own is
    // a SkyLight with the map skymap/studio_garden.hdr. That asset is not in the mount: --sky_map names the .hdr to use.
    if (!sky_map.empty() && !app->scene.SetSkyMapFile(sky_map))
    {
        return nullptr;
    }
    app->renderer = Renderer::CreateRenderer(app->config, &app->scene, device);
    if (!app->renderer)
    {
        return nullptr;
    }
    app->renderer->NotifySceneLoaded();
    return app.release();
}

void spk_host_app_destroy(SpkHostApp *app) { delete app; }

int spk_host_app_frame(SpkHostApp *app)
{
    // RenderFramework::RenderLoop body (RenderFramework.cpp:146-162)
    app->renderer->Tick();
    if (app->renderer->IsReadyForAutoScreenshot())
    {
        return 1;
    }
    app->renderer->Render();
    app->frames++;
    return 0;
}

int spk_host_app_render_frame(SpkHostApp *app)
{
    // the reference's order exactly (RenderFramework.cpp:153-158): Tick, Render, then the ready flag. Past max_spp every
    // such frame lerps one more sample in with weight cap/(cap+1) (CameraRenderProxy.cpp:70-73). spk_host_app_frame
    // checks BEFORE rendering so that spk_host_app_run stops with exactly max_spp samples: the reference's screenshot is
    // requested asynchronously and holds max_spp + k samples, k >= 1 depending on thread timing.
    app->renderer->Tick();
    app->renderer->Render();
    app->frames++;
    return app->renderer->IsReadyForAutoScreenshot() ? 1 : 0;
}

int spk_host_app_run(SpkHostApp *app, uint32_t max_frames)
{
    uint32_t start = app->frames;
    for (uint32_t f = 0; f < max_frames; f++)
    {
        if (spk_host_app_frame(app))
        {
            break;
        }
    }
    return static_cast<int>(app->frames - start);
}

const SpkSceneDesc *spk_host_app_scene(SpkHostApp *app) { return &app->scene.Desc(); }
const SpkCamera *spk_host_app_camera(SpkHostApp *app) { return &app->scene.GetCamera().Get(); }
uint32_t spk_host_app_cumulated_samples(SpkHostApp *app) { return app->scene.GetCamera().GetCumulatedSampleCount(); }

int spk_host_app_read_linear(SpkHostApp *app, float *out, size_t floats)
{
    auto *r = dynamic_cast<CudaRenderer *>(app->renderer.get());
    std::vector<float> img;
    if (!r || !r->ReadLinear(img) || img.size() > floats)
    {
        return -1;
    }
    memcpy(out, img.data(), img.size() * 4);
    return 0;
}

int spk_host_app_save_screenshot(SpkHostApp *app, const char *path)
{
    auto *r = dynamic_cast<CudaRenderer *>(app->renderer.get());
    std::vector<uint8_t> img;
    if (!r || !r->ReadSrgb8(img))
    {
        return -1;
    }
    return WritePng(path, img.data(), app->config.image_width, app->config.image_height) ? 0 : -1;
}

void spk_host_app_set_orbit(SpkHostApp *app, float cx, float cy, float cz, float radius, float pitch_deg, float yaw_deg)
{
    app->scene.GetCamera().SetupOrbit(Vec3{cx, cy, cz}, radius, pitch_deg, yaw_deg);
}

int spk_host_app_move_primitive(SpkHostApp *app, uint32_t primitive, float tx, float ty, float tz)
{
    if (primitive >= app->scene.NumPrimitives())
    {
        return -1;
    }
    const SpkPrimitiveDesc &p = app->scene.Desc().primitives[primitive];
    Transform t;
    t.translation = Vec3{tx, ty, tz};
    if (p.kind == SPK_PRIM_SPHERE)
    {
        float r = std::sqrt(p.radius2);
        t.scale = Vec3{r, r, r};
    }
    app->scene.MovePrimitive(primitive, t);
    return 0;
}
}

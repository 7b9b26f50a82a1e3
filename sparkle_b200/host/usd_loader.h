/*
 * usd_loader.h — USD (.usda) scene ingestion for the `cuda` pipeline.
 *
 * The reference loads USD through tinyusdz + its Tydra render-scene converter (libraries/source/io/scene/USDLoader.cpp);
 * neither is in the mount. What the reference guarantees is the ROUND TRIP (docs/USD.md): USDExporter.cpp writes exactly
 * the prims USDLoader.cpp understands, and the packaged default scene resources/packed/TestScene.usda is such an export.
 * This is a reader of its own for that subset of ASCII USD, with the loader's rules restated:
 *   - `Xform` prims form the node hierarchy; each carries ONE `xformOp:transform` matrix4d, written row-vector style
 *     (USDExporter.cpp:53-67 transposes Eigen's matrix; USDLoader.cpp:29-41 transposes it back). World = parent * local.
 *   - a child `Mesh` becomes a mesh primitive: points, per-vertex normals (required), primvars:st, primvars:tangents as
 *     float4 or float3 (w = 1); missing tangents are derived from the normal (USDLoader.cpp:338-347); triangles only;
 *     a mesh without `material:binding` is skipped (USDLoader.cpp:283-287)
 *   - `Material` + `UsdPreviewSurface` -> MaterialResource (USDLoader.cpp:349-425): a constant opacity < 1 makes a
 *     dielectric with eta = ior; diffuseColor / emissiveColor / normal / metallic+roughness are constants or
 *     `UsdUVTexture` connections (one packed metallic-roughness image read through .g / .b); an emissive texture forces
 *     emissive_color = 1; colour images are sRGB when inputs:sourceColorSpace = "sRGB", else Unorm (:195-205)
 *   - `DomeLight` -> sky colour = color * intensity * 2^exposure (:455-458); a sky MAP (inputs:texture:file, an
 *     equirectangular Radiance .hdr next to the .usda, :490-493) is decoded and cooked into the 6 x 1024^2 RGBAFloat16
 *     cube of SkyLightCookJob (host/sky_cook.h), which the path tracer then samples instead of the colour
 *   - `Camera` -> orbit camera rebuilt from the node transform and focusDistance (OrbitCameraComponent.cpp:32-56; lens
 *     attributes in millimetres, USDLoader.cpp:95-116); `DistantLight` is read and ignored (the path tracers light
 *     with the sky and emissive surfaces only)
 * Texture files are 8-bit PNG next to the .usda (`textures/...`, USDExporter.cpp:95-118).
 */
#ifndef SPARKLE_CUDA_USD_LOADER_H
#define SPARKLE_CUDA_USD_LOADER_H
#include "cuda_renderer.h"
#include "gltf_loader.h" // SpkHostScene

#include <string>

namespace sparkle_cuda
{
struct UsdStats
{
    uint32_t nodes = 0;      // Xform prims visited
    uint32_t meshes = 0;     // mesh primitives added
    uint32_t skipped = 0;    // meshes skipped (no material, not triangles, no normals)
    uint32_t materials = 0;
    uint32_t textures = 0;
    uint32_t cameras = 0;
    uint32_t dome_lights = 0;
    uint32_t distant_lights = 0;
    uint64_t triangles = 0;
};

// Adds the stage's default prim (or its first root Xform) to `scene`, sets sky and camera when the file has them.
// Returns false (and sets LastError) on malformed input; never throws.
bool LoadUsd(const std::string &path, SceneProxy &scene, UsdStats *stats = nullptr);
} // namespace sparkle_cuda

extern "C"
{
/* counters of a load into a scratch scene (9 x u64 in UsdStats order), or -1 */
int spk_host_usd_probe(const char *path, uint64_t *stats9);
/* a USD scene without a renderer: flattened description (TLAS built) for the oracle and its camera; NULL on failure.
 * Freed with spk_host_scene_free, read with spk_host_scene_desc (gltf_loader.h). */
SpkHostScene *spk_host_scene_load_usd(const char *path, uint32_t width, uint32_t height);
const SpkCamera *spk_host_scene_camera(SpkHostScene *scene);
}
#endif

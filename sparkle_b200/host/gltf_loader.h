/*
 * gltf_loader.h — glTF 2.0 scene ingestion for the `cuda` pipeline.
 *
 * Restates the semantics of the reference's loader, libraries/source/io/scene/GLTFLoader.cpp (tinygltf is absent from
 * the mount, so this is a minimal reader of its own: .gltf JSON + external or data-URI buffers, PNG images through zlib):
 *   - node scale is BAKED into the vertices and the node keeps scale 1 (GLTFLoader.cpp:192-247, 541-544)
 *   - every node rotation is pre-multiplied by +90 degrees about X, Y-up -> Z-up (:537-538), children included
 *   - node.matrix is ignored (only translation / rotation / scale are read, :530-544)
 *   - one mesh primitive per glTF primitive (:551-562). The reference duplicates the mesh for every node that uses it;
 *     here primitives with the same accessors and the same scale share ONE BLAS (GltfOptions::instancing extends that
 *     to any scale), which is what makes a 10M-triangle instanced scene 1M unique triangles on the device
 *     
 *   - material factors are only read when the matching texture exists (:485-517); otherwise MaterialResource defaults:
 *     base colour 1, emissive 0, metallic 1, roughness 1 (io/Material.h:7-22)
 *   - base colour / emissive images are sRGB, normal / metallic-roughness images are Unorm (:263, 306)
 *   - TRIANGLES only; points and lines are skipped with an error log like the reference (:452-457)
 *   - missing TANGENT: the reference interpolates the constant (1,0,0,1) (io/Mesh.h:104-109); missing TEXCOORD_0: uv 0
 */
#ifndef SPARKLE_CUDA_GLTF_LOADER_H
#define SPARKLE_CUDA_GLTF_LOADER_H
#include "cuda_renderer.h"

#include <string>

namespace sparkle_cuda
{
struct GltfStats
{
    uint32_t nodes = 0;           // nodes visited
    uint32_t primitives = 0;      // mesh primitives added to the scene (instances)
    uint32_t unique_meshes = 0;   // BLAS built (after sharing)
    uint32_t materials = 0;
    uint32_t textures = 0;
    uint64_t instanced_triangles = 0;
    uint64_t unique_triangles = 0;
};

struct GltfOptions
{
    // false (default): the reference's behaviour, node.scale baked into the vertices; primitives share a BLAS only when
    // accessors AND scale are equal. true: node.scale stays in the instance matrix (the same affine map, evaluated at
    // ray time instead of at load time), so every use of a primitive shares one BLAS whatever its scale -- what a scene
    // with per-instance scales needs to stay at its unique triangle count on the device. Not a reference option.
    bool instancing = false;
};

// Adds the default scene of `path` to `scene`. Returns false (and sets LastError) on malformed input; never throws.
bool LoadGltf(const std::string &path, SceneProxy &scene, GltfStats *stats = nullptr, const GltfOptions &options = {});

// 8-bit PNG (grey, grey+alpha, RGB, palette, RGBA; non-interlaced) -> RGBA8. Used for glTF images.
bool DecodePng(const uint8_t *data, size_t size, std::vector<uint8_t> &rgba, uint32_t &width, uint32_t &height);
} // namespace sparkle_cuda

// a scene loaded without a renderer (CPU tests feed its SpkSceneDesc to the oracle); shared with usd_loader.cpp
struct SpkHostScene
{
    sparkle_cuda::RenderConfig config;
    sparkle_cuda::SceneProxy scene;
};

extern "C"
{
/* tests / headless driver: "--scene file.gltf" goes through LoadGltf; this probe loads a file into a scratch scene and
 * returns the counters (7 x u64 in GltfStats order), or -1 */
int spk_host_gltf_probe(const char *path, int instancing, uint64_t *stats7);
/* a glTF scene without a renderer: its flattened description (TLAS built) for the oracle; NULL on failure */
SpkHostScene *spk_host_scene_load_gltf(const char *path, int instancing);
const SpkSceneDesc *spk_host_scene_desc(SpkHostScene *scene);
void spk_host_scene_free(SpkHostScene *scene);
}
#endif

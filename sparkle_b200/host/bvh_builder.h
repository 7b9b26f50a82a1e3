/*
 * bvh_builder.h — host-side BVH2 builder for the `cuda` pipeline.
 *
 * Stands where the reference calls bvh::v2::DefaultBuilder<Node>::build (madmann91/bvh @5a9d759c, absent from the
 * reference mount) in TLAS::Build (libraries/source/renderer/proxy/SceneRenderProxy.cpp:31-69) and BLAS::Build
 * (libraries/source/renderer/proxy/MeshRenderProxy.cpp:31-65). Input is the same pair of arrays the reference hands
 * to the library (one AABB and one centre per primitive); output is the library's node layout
 * (bounds = {min.x,max.x,min.y,max.y,min.z,max.z}, index = first_id<<4 | prim_count, children of an inner node adjacent,
 * node 0 = root) plus prim_ids, so the same flattened tree feeds the device scene builder, the reference harness and the
 * oracle.
 *
 * Algorithm: top-down binned SAH (16 bins on the centroid bounds, all three axes), leaves of at most `max_leaf`
 * primitives, object-median fallback once the remaining depth budget would not fit the reference's 32-entry traversal
 * stack (SmallStack<Index,32>, MeshRenderProxy.cpp:103,110).
 */
#ifndef SPK_HOST_BVH_BUILDER_H_
#define SPK_HOST_BVH_BUILDER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct SpkHostBvh SpkHostBvh; /* opaque; owns nodes + prim_ids */

/* bboxes: 6 floats per primitive {min.xyz, max.xyz}; centers: 3 floats per primitive (NULL => box centre (min+max)/2).
 * max_leaf in [1,15]; max_depth <= 32. Returns NULL on invalid arguments. */
SpkHostBvh *spk_host_bvh_build(const float *bboxes, const float *centers, uint32_t num_prims, uint32_t max_leaf,
                               uint32_t max_depth);
/* Convenience for BLAS::Build: boxes/centres from an indexed triangle list exactly as Mesh::GetTriangleMinMax does
 * (io/Mesh.h:55-71) and centre = (min+max)/2 (MeshRenderProxy.cpp:50). */
SpkHostBvh *spk_host_bvh_build_triangles(const float *vertices, const uint32_t *indices, uint32_t num_faces,
                                         uint32_t max_leaf, uint32_t max_depth);
void spk_host_bvh_free(SpkHostBvh *bvh);

uint32_t spk_host_bvh_num_nodes(const SpkHostBvh *bvh);
uint32_t spk_host_bvh_num_prims(const SpkHostBvh *bvh);
uint32_t spk_host_bvh_depth(const SpkHostBvh *bvh);
/* 32-byte nodes: 6 float bounds, u32 index, u32 pad (SpkBvhNode of include/spk_cuda.h) */
const void *spk_host_bvh_nodes(const SpkHostBvh *bvh);
const uint32_t *spk_host_bvh_prim_ids(const SpkHostBvh *bvh);
/* SAH cost of the tree (inner 1.0, primitive 1.0), for reports */
double spk_host_bvh_sah_cost(const SpkHostBvh *bvh);

#ifdef __cplusplus
}
#endif
#endif

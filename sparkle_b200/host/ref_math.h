/*
 * ref_math.h — the reference's scene-graph float math on the way into the `cuda` pipeline, restated operation for
 * operation so that the SAME orbit parameters / TRS give the SAME BITS the reference's Eigen code produces
 * (g++ -O2 -ffp-contract=off on x86-64, where Eigen's quaternion product takes its SSE path):
 *
 *   utilities::EulerRotationToRotationAxis   libraries/include/core/math/Utilities.h:392-396  AngleAxis(x,Right) *
 *                                            AngleAxis(y,Front) * AngleAxis(z,Up) as quaternions
 *   Transform::UpdateTransformData           libraries/source/core/math/Transform.cpp:52-64   Translation * rotation *
 *                                            Scaling, then the general affine inverse (3x3 cofactors / determinant)
 *   GLTFLoader ProcessNode (transform)      libraries/source/io/scene/GLTFLoader.cpp:532-544
 *   OrbitCameraComponent::UpdateTransform    libraries/source/scene/component/camera/OrbitCameraComponent.cpp:16-30
 *   CalculateRenderAttribute                 libraries/source/scene/component/camera/CameraComponent.cpp:120-128
 *   CameraRenderProxy::Update (attributes)   libraries/source/renderer/proxy/CameraRenderProxy.cpp:21-33
 *   CameraRenderProxy::OnTransformDirty      libraries/source/renderer/proxy/CameraRenderProxy.cpp:113-136
 *
 * Pinned bit for bit against the reference's own compiled code by tests/test_host_cpu.py (orc_kat_transform,
 * orc_kat_camera in oracle/ref_harness.cpp) and by the committed golden vectors.
 */
#ifndef SPK_HOST_REF_MATH_H_
#define SPK_HOST_REF_MATH_H_

#include "../../include/spk_cuda.h"

namespace sparkle_cuda
{
namespace refmath
{
struct Quat
{
    float x, y, z, w; // Eigen coefficient order
};

Quat FromAxisAngle(float angle, const float axis[3]); // QuaternionBase::operator=(AngleAxis)
Quat Mul(const Quat &a, const Quat &b);                // quat_product<Target, float>: the SSE association
void ToMatrix(const Quat &q, float m[9]);              // QuaternionBase::toRotationMatrix, row-major 3x3
void Rotate(const Quat &q, const float v[3], float out[3]); // QuaternionBase::_transformVector
Quat EulerToQuat(const float euler[3]);                // EulerRotationToRotationAxis
Quat Normalized(const Quat &q);                        // Quaternion::normalized()
Quat GltfNodeRotation(const float q_xyzw[4]);          // GLTFLoader.cpp:532-544: y-up -> z-up, normalized

// Transform(translate, euler, scale): rows 0..2 of GetMatrix() and of GetInverse().GetMatrix()
void TrsMatrices(const float translate[3], const float euler[3], const float scale[3], float fwd[12], float inv[12]);
// the same from a rotation quaternion (scene loaders)
void TrsMatricesQuat(const float translate[3], const Quat &rotation, const float scale[3], float fwd[12], float inv[12]);

// OrbitCameraComponent::Setup(center, radius, pitch, yaw) + CameraComponent attribute -> CameraRenderProxy state
void OrbitCamera(const float center[3], float radius, float pitch_deg, float yaw_deg, float focal_length,
                 float sensor_height, float aperture, float exposure, uint32_t width, uint32_t height, SpkCamera *out);
} // namespace refmath
} // namespace sparkle_cuda

extern "C"
{
/* ctypes entry points (sparkle_b200/geometry.py trs, sparkle_b200/scene.py orbit_camera): one implementation for the
 * C++ host and the Python bindings */
void spk_host_trs(const float *translate, const float *euler, const float *scale, float *fwd12, float *inv12);
/* root-level glTF node: translation + rotation xyzw -> matrices (GLTFLoader.cpp:532-544) */
void spk_host_gltf_node(const float *translation, const float *rotation_xyzw, float *fwd12, float *inv12);
void spk_host_orbit_camera(const float *center, float radius, float pitch_deg, float yaw_deg, float focal_length,
                           float sensor_height, float aperture, float exposure, uint32_t width, uint32_t height,
                           SpkCamera *out);
}
#endif

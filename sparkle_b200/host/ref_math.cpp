/* ref_math.cpp — see ref_math.h. Compiled -O2 -ffp-contract=off (no fused multiply-add), like the reference oracle build. */
#include "ref_math.h"

#include <cmath>

namespace sparkle_cuda
{
namespace refmath
{
namespace
{
constexpr float kPi = 3.14159274101257324f; // std::numbers::pi_v<float> (core/math/Types.h:51)
const float kRight[3] = {1.f, 0.f, 0.f}, kFront[3] = {0.f, 1.f, 0.f}, kUp[3] = {0.f, 0.f, 1.f}; // Types.h:45-47

inline float ToRadian(float degree) { return degree / 180.f * kPi; } // Utilities.h:73-76

inline void Cross(const float a[3], const float b[3], float out[3])
{
    out[0] = a[1] * b[2] - a[2] * b[1];
    out[1] = a[2] * b[0] - a[0] * b[2];
    out[2] = a[0] * b[1] - a[1] * b[0];
}
} // namespace

Quat FromAxisAngle(float angle, const float axis[3])
{
    const float ha = 0.5f * angle;
    const float s = std::sin(ha); // glibc sinf / cosf, what EIGEN_USING_STD(sin) resolves to for float
    Quat q;
    q.w = std::cos(ha);
    q.x = s * axis[0];
    q.y = s * axis[1];
    q.z = s * axis[2];
    return q;
}

Quat Mul(const Quat &a, const Quat &b)
{
    // Eigen's SSE quat_product (arch/Geometry_SIMD.h): per lane (a*b.w - swz(a)*swz(b)) + (+-)(s1 + s2), w lane negated
    Quat r;
    r.x = (a.x * b.w - a.z * b.y) + (a.y * b.z + a.w * b.x);
    r.y = (a.y * b.w - a.x * b.z) + (a.z * b.x + a.w * b.y);
    r.z = (a.z * b.w - a.y * b.x) + (a.x * b.y + a.w * b.z);
    r.w = (a.w * b.w - a.x * b.x) + -(a.z * b.z + a.y * b.y);
    return r;
}

void ToMatrix(const Quat &q, float m[9])
{
    const float tx = 2.f * q.x, ty = 2.f * q.y, tz = 2.f * q.z;
    const float twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
    const float txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
    const float tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
    m[0] = 1.f - (tyy + tzz);
    m[1] = txy - twz;
    m[2] = txz + twy;
    m[3] = txy + twz;
    m[4] = 1.f - (txx + tzz);
    m[5] = tyz - twx;
    m[6] = txz - twy;
    m[7] = tyz + twx;
    m[8] = 1.f - (txx + tyy);
}

void Rotate(const Quat &q, const float v[3], float out[3])
{
    // uv = vec x v; uv += uv; v + w * uv + vec x uv
    const float vec[3] = {q.x, q.y, q.z};
    float uv[3], c[3];
    Cross(vec, v, uv);
    uv[0] += uv[0];
    uv[1] += uv[1];
    uv[2] += uv[2];
    Cross(vec, uv, c);
    for (int k = 0; k < 3; k++)
    {
        out[k] = (v[k] + q.w * uv[k]) + c[k];
    }
}

Quat EulerToQuat(const float euler[3])
{
    return Mul(Mul(FromAxisAngle(euler[0], kRight), FromAxisAngle(euler[1], kFront)), FromAxisAngle(euler[2], kUp));
}

Quat Normalized(const Quat &q)
{
    // Vector4f::normalized(): squaredNorm as one SSE packet reduction, (x2 + z2) + (y2 + w2), then true division
    const float z = (q.x * q.x + q.z * q.z) + (q.y * q.y + q.w * q.w);
    if (!(z > 0.f))
    {
        return q;
    }
    const float n = std::sqrt(z);
    return Quat{q.x / n, q.y / n, q.z / n, q.w / n};
}

Quat GltfNodeRotation(const float q_xyzw[4])
{
    // GLTFLoader.cpp:532-544: AngleAxis(ToRadian(90), Right) * rotation, normalized
    const Quat r{q_xyzw[0], q_xyzw[1], q_xyzw[2], q_xyzw[3]};
    return Normalized(Mul(FromAxisAngle(ToRadian(90.0f), kRight), r));
}

void TrsMatricesQuat(const float translate[3], const Quat &rotation, const float scale[3], float fwd[12], float inv[12])
{
    float r[9];
    ToMatrix(rotation, r);
    float l[9]; // linear part: rotation columns scaled
    for (int i = 0; i < 3; i++)
    {
        for (int j = 0; j < 3; j++)
        {
            l[3 * i + j] = r[3 * i + j] * scale[j];
            fwd[4 * i + j] = l[3 * i + j];
        }
        fwd[4 * i + 3] = 0.f + translate[i]; // Translation * Isometry: pretranslate adds onto a zero translation
    }
    // Transform::inverse(Affine): linear().inverse() by cofactors (LU/InverseImpl.h, size 3), translation = (-Linv) * t
    auto cof = [&](int i, int j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return l[3 * i1 + j1] * l[3 * i2 + j2] - l[3 * i1 + j2] * l[3 * i2 + j1];
    };
    const float c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
    const float det = c00 * l[0] + (c10 * l[3] + c20 * l[6]); // .sum() of a 3-vector: a0 + (a1 + a2)
    const float invdet = 1.f / det;
    float li[9];
    li[0] = c00 * invdet;
    li[1] = c10 * invdet;
    li[2] = c20 * invdet;
    li[3] = cof(0, 1) * invdet;
    li[4] = cof(1, 1) * invdet;
    li[5] = cof(2, 1) * invdet;
    li[6] = cof(0, 2) * invdet;
    li[7] = cof(1, 2) * invdet;
    li[8] = cof(2, 2) * invdet;
    const float *t = fwd; // translation column of the forward matrix
    for (int i = 0; i < 3; i++)
    {
        inv[4 * i + 0] = li[3 * i + 0];
        inv[4 * i + 1] = li[3 * i + 1];
        inv[4 * i + 2] = li[3 * i + 2];
        inv[4 * i + 3] = (-li[3 * i + 0]) * t[3] + ((-li[3 * i + 1]) * t[7] + (-li[3 * i + 2]) * t[11]);
    }
}

void TrsMatrices(const float translate[3], const float euler[3], const float scale[3], float fwd[12], float inv[12])
{
    TrsMatricesQuat(translate, EulerToQuat(euler), scale, fwd, inv);
}

void OrbitCamera(const float center[3], float radius, float pitch_deg, float yaw_deg, float focal_length,
                 float sensor_height, float aperture, float exposure, uint32_t width, uint32_t height, SpkCamera *out)
{
    // OrbitCameraComponent::UpdateTransform
    const float rr[3] = {ToRadian(-pitch_deg), ToRadian(0.f), ToRadian(-yaw_deg)};
    const Quat rot = Mul(FromAxisAngle(rr[2], kUp), FromAxisAngle(rr[0], kRight));
    float dir[3];
    Rotate(rot, kFront, dir);
    const float position[3] = {dir[0] * -radius + center[0], dir[1] * -radius + center[1], dir[2] * -radius + center[2]};
    const float euler[3] = {-rr[0], -rr[1], -rr[2]}; // GetNode()->SetTransform(position, -relative_rotation)
    const float ones[3] = {1.f, 1.f, 1.f};
    float fwd[12], inv[12];
    TrsMatrices(position, euler, ones, fwd, inv);
    // CalculateRenderAttribute + SetFocusDistance(radius_)
    const float vertical_fov = 2.f * std::atan(sensor_height / (2.f * focal_length));
    const float focus = radius;
    // CameraRenderProxy::Update, attribute block
    const float aspect = static_cast<float>(width) / static_cast<float>(height);
    const float h = std::tan(vertical_fov / 2.f);
    const float plane_h = 2.f * h * focus;
    const float plane_w = aspect * plane_h;
    // OnTransformDirty: position = GetTranslation(); right / front / up = rows 0 / 1 / 2 of the matrix
    for (int k = 0; k < 3; k++)
    {
        out->position[k] = position[k];
        out->right[k] = fwd[k];
        out->front[k] = fwd[4 + k];
        out->up[k] = fwd[8 + k];
        out->max_u[k] = out->right[k] * plane_w;
        out->max_v[k] = out->up[k] * plane_h;
        out->lower_left[k] = ((out->position[k] + out->front[k] * focus) - out->max_u[k] * .5f) - out->max_v[k] * .5f;
    }
    out->aperture_radius = focal_length / aperture * 0.5f;
    out->exposure = exposure;
}
} // namespace refmath
} // namespace sparkle_cuda

extern "C"
{
void spk_host_trs(const float *translate, const float *euler, const float *scale, float *fwd12, float *inv12)
{
    sparkle_cuda::refmath::TrsMatrices(translate, euler, scale, fwd12, inv12);
}

void spk_host_gltf_node(const float *translation, const float *rotation_xyzw, float *fwd12, float *inv12)
{
    const float ones[3] = {1.f, 1.f, 1.f};
    sparkle_cuda::refmath::TrsMatricesQuat(translation, sparkle_cuda::refmath::GltfNodeRotation(rotation_xyzw), ones, fwd12,
                                           inv12);
}

void spk_host_orbit_camera(const float *center, float radius, float pitch_deg, float yaw_deg, float focal_length,
                           float sensor_height, float aperture, float exposure, uint32_t width, uint32_t height,
                           SpkCamera *out)
{
    sparkle_cuda::refmath::OrbitCamera(center, radius, pitch_deg, yaw_deg, focal_length, sensor_height, aperture, exposure,
                                       width, height, out);
}
}

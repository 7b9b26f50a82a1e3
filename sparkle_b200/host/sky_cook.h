/*
 * sky_cook.h — the sky MAP of a SkyLight, from an .hdr file to the six cube faces the path tracer samples (SURVEY.md 8f-2,
 * the data source of row a14).
 *
 * Reference: SkyLight::SetSkyMap loads an equirectangular Radiance .hdr with Image2D::LoadFromFile (io/Image.cpp:120-196:
 * stb_image's stbi_loadf, values clamped to +-65504 and stored as RGBAFloat16) and SkyLightCookJob::Execute
 * (scene/component/light/SkyLight.cpp:82-131) turns it into a 1024 x 1024 x 6 RGBAFloat16 cube on the CPU: for pixel (i, j)
 * of a face, direction = Image2DCube::TextureCoordinateToDirection(face, u, v) with u, v = (i + 0.5) / size * 2 - 1
 * (io/Image.h:287-315), uv = CartesianToEquirectangular(direction) (core/math/Utilities.h:510-518), colour =
 * Image2D::Sample(uv) (bilinear, wrapping, io/Image.h:112-151), SetPixel(i, j, colour) (alpha 1, cast to half). The sun
 * statistics of the same job feed the raster pipelines' directional light and are not computed here.
 * stb_image (thirdparty/stb/stb_image.h, stbi__hdr_load :7158-7287, stbi__hdr_convert :7131-7156) is restated: header
 * tokens, flat and run-length scanlines, mantissa * 2^(e - 136).
 * Checked bit for bit against the reference's own functions (oracle/ref_harness.cpp: ref_decode_hdr, ref_cook_sky_cube)
 * in tests/test_sky_cook_cpu.py.
 */
#ifndef SPARKLE_CUDA_SKY_COOK_H
#define SPARKLE_CUDA_SKY_COOK_H
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace sparkle_cuda
{
constexpr uint32_t kSkyCubeSize = 1024; // SkyLight.cpp:24

// Eigen::half(float): round to nearest even, denormals kept, overflow to infinity
uint16_t FloatToHalf(float f);
float HalfToFloat(uint16_t h);

// Radiance RGBE -> RGBAFloat16 the way Image2D::LoadFromFile stores it. Returns false (message in `error`) on input
// stb_image would refuse.
bool DecodeRadianceHdr(const uint8_t *data, size_t size, uint32_t &width, uint32_t &height, std::vector<uint16_t> &rgba_half,
                       std::string &error);

// SkyLightCookJob::Execute's loop: six faces of size x size RGBAFloat16 pixels (4 halves each), FaceId order
void CookSkyCube(const uint16_t *equirect_rgba_half, uint32_t width, uint32_t height, uint32_t size, uint16_t *faces_out);
} // namespace sparkle_cuda

extern "C"
{
/* tests: decode into a malloc'd buffer of width * height * 4 halves (free with spk_host_free_buffer); 0 or -1 */
int spk_host_decode_hdr(const uint8_t *data, size_t size, uint32_t *width, uint32_t *height, uint16_t **rgba_half);
void spk_host_free_buffer(void *p);
/* tests: faces_out holds 6 * size * size * 4 halves */
int spk_host_cook_sky_cube(const uint16_t *equirect_rgba_half, uint32_t width, uint32_t height, uint32_t size,
                           uint16_t *faces_out);
}
#endif

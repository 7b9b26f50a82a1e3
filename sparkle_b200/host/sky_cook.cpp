/*
 * sky_cook.cpp — see sky_cook.h. Compiled with -ffp-contract=off like the rest of the host library: the float arithmetic
 * below is the reference's, operation for operation.
 */
#include "sky_cook.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace sparkle_cuda
{
// Eigen::half_impl::float_to_half_rtne (Eigen/src/Core/arch/Default/Half.h), the portable branch
uint16_t FloatToHalf(float value)
{
    uint32_t f;
    std::memcpy(&f, &value, 4);
    const uint32_t sign = f & 0x80000000u;
    f ^= sign;
    uint32_t o;
    if (f >= ((127u + 16u) << 23)) // 65536 and above, infinities, NaNs
    {
        o = f > (255u << 23) ? 0x7e00u : 0x7c00u;
    }
    else if (f < (113u << 23)) // below 2^-14: a half denormal (or zero); the float adder does the rounding
    {
        const uint32_t magic_bits = ((127u - 15u) + (23u - 10u) + 1u) << 23; // 0.5f
        float magic, x;
        std::memcpy(&magic, &magic_bits, 4);
        std::memcpy(&x, &f, 4);
        x += magic;
        uint32_t xb;
        std::memcpy(&xb, &x, 4);
        o = xb - magic_bits;
    }
    else
    {
        const uint32_t mant_odd = (f >> 13) & 1u;
        f += (uint32_t(15 - 127) << 23) + 0xfffu;
        f += mant_odd;
        o = f >> 13;
    }
    return uint16_t(o | (sign >> 16));
}

// Eigen::half_impl::half_to_float, the portable branch
float HalfToFloat(uint16_t h)
{
    const uint32_t shifted_exp = 0x7c00u << 13;
    uint32_t o = (uint32_t(h) & 0x7fffu) << 13;
    const uint32_t exp = shifted_exp & o;
    o += (127u - 15u) << 23;
    if (exp == shifted_exp) // Inf / NaN
    {
        o += (128u - 16u) << 23;
    }
    else if (exp == 0u) // zero / denormal
    {
        o += 1u << 23;
        float f, magic;
        const uint32_t magic_bits = 113u << 23;
        std::memcpy(&f, &o, 4);
        std::memcpy(&magic, &magic_bits, 4);
        f -= magic;
        std::memcpy(&o, &f, 4);
    }
    o |= (uint32_t(h) & 0x8000u) << 16;
    float out;
    std::memcpy(&out, &o, 4);
    return out;
}

namespace
{
// stb_image's memory reader: reads past the end return 0
struct Reader
{
    const uint8_t *p, *end;
    bool AtEof() const { return p >= end; }
    int Get8() { return p < end ? *p++ : 0; }
    void GetN(uint8_t *out, int n)
    {
        for (int k = 0; k < n; k++)
        {
            out[k] = uint8_t(Get8());
        }
    }
};

constexpr int kHdrBufLen = 1024; // STBI__HDR_BUFLEN

// stbi__hdr_gettoken: one header line
std::string GetToken(Reader &r)
{
    std::string token;
    char c = char(r.Get8());
    while (!r.AtEof() && c != '\n')
    {
        token.push_back(c);
        if (int(token.size()) == kHdrBufLen - 1)
        {
            while (!r.AtEof() && r.Get8() != '\n')
            {
            }
            break;
        }
        c = char(r.Get8());
    }
    return token;
}

// stbi__hdr_convert with req_comp 4, then Image2D::LoadFromFile's clamp and cast (Image.cpp:166-181)
void ConvertPixel(const uint8_t rgbe[4], uint16_t *out)
{
    float rgba[4] = {0.f, 0.f, 0.f, 1.f};
    if (rgbe[3] != 0)
    {
        const float f1 = float(std::ldexp(1.0f, int(rgbe[3]) - (128 + 8)));
        rgba[0] = float(rgbe[0]) * f1;
        rgba[1] = float(rgbe[1]) * f1;
        rgba[2] = float(rgbe[2]) * f1;
    }
    const float half_max = 65504.f;
    for (int k = 0; k < 4; k++)
    {
        out[k] = FloatToHalf(std::clamp(rgba[k], -half_max, half_max));
    }
}
} // namespace

bool DecodeRadianceHdr(const uint8_t *data, size_t size, uint32_t &width_out, uint32_t &height_out,
                       std::vector<uint16_t> &pixels, std::string &error)
{
    Reader r{data, data + size};
    const std::string id = GetToken(r);
    if (id != "#?RADIANCE" && id != "#?RGBE")
    {
        error = "not HDR";
        return false;
    }
    bool valid = false;
    for (;;)
    {
        const std::string token = GetToken(r);
        if (token.empty())
        {
            break;
        }
        if (token == "FORMAT=32-bit_rle_rgbe")
        {
            valid = true;
        }
    }
    if (!valid)
    {
        error = "unsupported format";
        return false;
    }
    const std::string dims = GetToken(r);
    if (dims.compare(0, 3, "-Y ") != 0)
    {
        error = "unsupported data layout";
        return false;
    }
    char *next = nullptr;
    const int height = int(std::strtol(dims.c_str() + 3, &next, 10));
    while (*next == ' ')
    {
        ++next;
    }
    if (std::strncmp(next, "+X ", 3) != 0)
    {
        error = "unsupported data layout";
        return false;
    }
    const int width = int(std::strtol(next + 3, nullptr, 10));
    constexpr int kMaxDimension = 1 << 24; // STBI_MAX_DIMENSIONS
    if (height > kMaxDimension || width > kMaxDimension || width <= 0 || height <= 0)
    {
        error = "too large";
        return false;
    }
    // stbi__mad4sizes_valid(width, height, 4, sizeof(float)): the decoded image must fit an int
    if (uint64_t(width) * uint64_t(height) * 4u * sizeof(float) > uint64_t(0x7fffffff))
    {
        error = "too large";
        return false;
    }
    pixels.assign(size_t(width) * size_t(height) * 4, 0);
    width_out = uint32_t(width);
    height_out = uint32_t(height);

    // the flat form; also where stb_image lands when a scanline does not start a run-length record (from pixel 1 of row 0,
    // "yes, this makes no sense")
    auto flat_from = [&](int j0, int i0) {
        for (int j = j0; j < height; ++j)
        {
            for (int i = (j == j0 ? i0 : 0); i < width; ++i)
            {
                uint8_t rgbe[4];
                r.GetN(rgbe, 4);
                ConvertPixel(rgbe, pixels.data() + (size_t(j) * width + i) * 4);
            }
        }
    };
    if (width < 8 || width >= 32768)
    {
        flat_from(0, 0);
        return true;
    }
    std::vector<uint8_t> scanline(size_t(width) * 4);
    for (int j = 0; j < height; ++j)
    {
        const int c1 = r.Get8(), c2 = r.Get8();
        int len = r.Get8();
        if (c1 != 2 || c2 != 2 || (len & 0x80))
        {
            const uint8_t rgbe[4] = {uint8_t(c1), uint8_t(c2), uint8_t(len), uint8_t(r.Get8())};
            ConvertPixel(rgbe, pixels.data());
            flat_from(0, 1);
            return true;
        }
        len = (len << 8) | r.Get8();
        if (len != width)
        {
            error = "invalid decoded scanline length";
            return false;
        }
        for (int k = 0; k < 4; ++k)
        {
            int i = 0, nleft;
            while ((nleft = width - i) > 0)
            {
                int count = r.Get8();
                if (count > 128)
                {
                    const uint8_t value = uint8_t(r.Get8());
                    count -= 128;
                    if (count == 0 || count > nleft)
                    {
                        error = "bad RLE data in HDR";
                        return false;
                    }
                    for (int z = 0; z < count; ++z)
                    {
                        scanline[size_t(i++) * 4 + k] = value;
                    }
                }
                else
                {
                    if (count == 0 || count > nleft)
                    {
                        error = "bad RLE data in HDR";
                        return false;
                    }
                    for (int z = 0; z < count; ++z)
                    {
                        scanline[size_t(i++) * 4 + k] = uint8_t(r.Get8());
                    }
                }
            }
        }
        for (int i = 0; i < width; ++i)
        {
            ConvertPixel(scanline.data() + size_t(i) * 4, pixels.data() + (size_t(j) * width + i) * 4);
        }
    }
    return true;
}

namespace
{
// utilities::WrapMod (Utilities.h:526-529)
inline unsigned WrapMod(int a, unsigned b)
{
    return (unsigned(a % int(b)) + b) % b;
}

// Image2D::Sample on an RGBAFloat16 image (io/Image.h:112-151): rgb of the bilinear filter, Lerp(a, b, f) = a + (b - a) * f
inline void SampleHalf(const uint16_t *px, unsigned width, unsigned height, float u, float v, float rgb[3])
{
    const float pu = u * float(width - 1u), pv = v * float(height - 1u); // size_vector_ = (width - 1, height - 1), Image.h:21
    const int iu = int(pu), iv = int(pv); // utilities::Decompose
    const float fu = pu - float(iu), fv = pv - float(iv);
    const unsigned x0 = WrapMod(iu, width), x1 = WrapMod(iu + 1, width), y0 = WrapMod(iv, height), y1 = WrapMod(iv + 1, height);
    const uint16_t *s00 = px + (size_t(y0) * width + x0) * 4, *s10 = px + (size_t(y0) * width + x1) * 4;
    const uint16_t *s01 = px + (size_t(y1) * width + x0) * 4, *s11 = px + (size_t(y1) * width + x1) * 4;
    for (int c = 0; c < 3; c++)
    {
        const float a00 = HalfToFloat(s00[c]), a10 = HalfToFloat(s10[c]), a01 = HalfToFloat(s01[c]), a11 = HalfToFloat(s11[c]);
        const float l0 = a00 + (a10 - a00) * fu;
        const float l1 = a01 + (a11 - a01) * fu;
        rgb[c] = l0 + (l1 - l0) * fv;
    }
}

// Image2DCube::TextureCoordinateToDirection (io/Image.h:287-315); Vector3::normalized() sums x*x + (y*y + z*z)
inline void FaceDirection(unsigned face, float u, float v, float d[3])
{
    switch (face)
    {
    case 0: d[0] = 1.f; d[1] = -v; d[2] = -u; break;
    case 1: d[0] = -1.f; d[1] = -v; d[2] = u; break;
    case 2: d[0] = u; d[1] = 1.f; d[2] = v; break;
    case 3: d[0] = u; d[1] = -1.f; d[2] = -v; break;
    case 4: d[0] = u; d[1] = -v; d[2] = 1.f; break;
    default: d[0] = -u; d[1] = -v; d[2] = -1.f; break;
    }
    const float n2 = d[0] * d[0] + (d[1] * d[1] + d[2] * d[2]);
    if (n2 > 0.f)
    {
        const float n = std::sqrt(n2);
        d[0] = d[0] / n;
        d[1] = d[1] / n;
        d[2] = d[2] / n;
    }
}
} // namespace

void CookSkyCube(const uint16_t *equirect, uint32_t width, uint32_t height, uint32_t size, uint16_t *faces)
{
    constexpr float kInvPi = 0.318309886183790671537767526745028724f; // std::numbers::inv_pi_v<float>
    const float inv_atan_x = kInvPi * 0.5f, inv_atan_y = kInvPi;
    const uint16_t one = FloatToHalf(1.f);
    auto columns = [&](unsigned face, unsigned i_begin, unsigned i_end) {
        uint16_t *out = faces + size_t(face) * size * size * 4;
        for (unsigned i = i_begin; i < i_end; i++)
        {
            for (unsigned j = 0; j < size; j++)
            {
                const float u = (float(i) + 0.5f) / float(size) * 2.f - 1.f;
                const float v = (float(j) + 0.5f) / float(size) * 2.f - 1.f;
                float d[3];
                FaceDirection(face, u, v, d);
                // utilities::CartesianToEquirectangular (Utilities.h:510-518)
                const float eu = std::atan2(d[1], d[0]) * inv_atan_x + 0.5f;
                const float ev = std::asin(-d[2]) * inv_atan_y + 0.5f;
                float rgb[3];
                SampleHalf(equirect, width, height, eu, ev, rgb);
                uint16_t *p = out + (size_t(j) * size + i) * 4; // SetPixel(x = i, y = j)
                p[0] = FloatToHalf(rgb[0]);
                p[1] = FloatToHalf(rgb[1]);
                p[2] = FloatToHalf(rgb[2]);
                p[3] = one;
            }
        }
    };
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned per_face = std::max(1u, std::min(hw, 16u) / 2u);
    std::vector<std::thread> workers;
    for (unsigned face = 0; face < 6; face++)
    {
        for (unsigned k = 0; k < per_face; k++)
        {
            const unsigned b = unsigned(uint64_t(size) * k / per_face), e = unsigned(uint64_t(size) * (k + 1) / per_face);
            workers.emplace_back(columns, face, b, e);
        }
    }
    for (auto &w : workers)
    {
        w.join();
    }
}
} // namespace sparkle_cuda

extern "C"
{
int spk_host_decode_hdr(const uint8_t *data, size_t size, uint32_t *width, uint32_t *height, uint16_t **rgba_half)
{
    if (!data || !width || !height || !rgba_half)
    {
        return -1;
    }
    std::vector<uint16_t> px;
    std::string error;
    if (!sparkle_cuda::DecodeRadianceHdr(data, size, *width, *height, px, error))
    {
        return -1;
    }
    *rgba_half = static_cast<uint16_t *>(std::malloc(px.size() * 2));
    if (!*rgba_half)
    {
        return -1;
    }
    std::memcpy(*rgba_half, px.data(), px.size() * 2);
    return 0;
}

void spk_host_free_buffer(void *p)
{
    std::free(p);
}

int spk_host_cook_sky_cube(const uint16_t *equirect, uint32_t width, uint32_t height, uint32_t size, uint16_t *faces_out)
{
    if (!equirect || !faces_out || width == 0 || height == 0 || size == 0)
    {
        return -1;
    }
    sparkle_cuda::CookSkyCube(equirect, width, height, size, faces_out);
    return 0;
}
}

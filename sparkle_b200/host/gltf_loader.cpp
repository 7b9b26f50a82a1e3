/* gltf_loader.cpp — see gltf_loader.h. Host code only; nothing here touches the device. */
#include "gltf_loader.h"
#include "ref_math.h"

#include <zlib.h>

#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>

namespace sparkle_cuda
{
void SetLastError(const std::string &msg); // cuda_renderer.cpp

namespace
{
// ---- a small JSON reader (objects, arrays, strings, numbers, true/false/null) -------------------------------------
struct Json
{
    enum Type
    {
        Null,
        Bool,
        Number,
        String,
        Array,
        Object
    } type = Null;
    double number = 0;
    bool boolean = false;
    std::string string;
    std::vector<Json> array;
    std::vector<std::pair<std::string, Json>> object;

    [[nodiscard]] const Json *Find(const std::string &key) const
    {
        for (const auto &kv : object)
        {
            if (kv.first == key)
            {
                return &kv.second;
            }
        }
        return nullptr;
    }
    [[nodiscard]] int Int(const std::string &key, int dflt) const
    {
        const Json *j = Find(key);
        return j && j->type == Number ? static_cast<int>(j->number) : dflt;
    }
    [[nodiscard]] double Num(const std::string &key, double dflt) const
    {
        const Json *j = Find(key);
        return j && j->type == Number ? j->number : dflt;
    }
    [[nodiscard]] std::string Str(const std::string &key) const
    {
        const Json *j = Find(key);
        return j && j->type == String ? j->string : std::string();
    }
    [[nodiscard]] const std::vector<Json> &Arr(const std::string &key) const
    {
        static const std::vector<Json> empty;
        const Json *j = Find(key);
        return j && j->type == Array ? j->array : empty;
    }
};

class JsonParser
{
public:
    JsonParser(const char *p, const char *end) : p_(p), end_(end) {}
    bool Parse(Json &out)
    {
        return Value(out, 0) && (Skip(), p_ == end_);
    }

private:
    const char *p_, *end_;
    void Skip()
    {
        while (p_ < end_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\r' || *p_ == '\t'))
        {
            p_++;
        }
    }
    bool Literal(const char *s)
    {
        size_t n = strlen(s);
        if (size_t(end_ - p_) < n || memcmp(p_, s, n) != 0)
        {
            return false;
        }
        p_ += n;
        return true;
    }
    bool ParseString(std::string &out)
    {
        if (p_ >= end_ || *p_ != '"')
        {
            return false;
        }
        p_++;
        out.clear();
        while (p_ < end_ && *p_ != '"')
        {
            char c = *p_++;
            if (c == '\\')
            {
                if (p_ >= end_)
                {
                    return false;
                }
                char e = *p_++;
                switch (e)
                {
                case 'n': out += '\n'; break;
                case 't': out += '\t'; break;
                case 'r': out += '\r'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'u': {
                    if (end_ - p_ < 4)
                    {
                        return false;
                    }
                    unsigned cp = static_cast<unsigned>(strtoul(std::string(p_, p_ + 4).c_str(), nullptr, 16));
                    p_ += 4;
                    out += cp < 0x80 ? static_cast<char>(cp) : '?'; // names only; non-ASCII is not interpreted
                    break;
                }
                default: out += e; break; // \" \\ \/
                }
            }
            else
            {
                out += c;
            }
        }
        if (p_ >= end_)
        {
            return false;
        }
        p_++;
        return true;
    }
    bool Value(Json &out, int depth)
    {
        if (depth > 64)
        {
            return false;
        }
        Skip();
        if (p_ >= end_)
        {
            return false;
        }
        if (*p_ == '{')
        {
            p_++;
            out.type = Json::Object;
            Skip();
            if (p_ < end_ && *p_ == '}')
            {
                p_++;
                return true;
            }
            for (;;)
            {
                Skip();
                std::string key;
                if (!ParseString(key))
                {
                    return false;
                }
                Skip();
                if (p_ >= end_ || *p_++ != ':')
                {
                    return false;
                }
                out.object.emplace_back(std::move(key), Json{});
                if (!Value(out.object.back().second, depth + 1))
                {
                    return false;
                }
                Skip();
                if (p_ < end_ && *p_ == ',')
                {
                    p_++;
                    continue;
                }
                if (p_ < end_ && *p_ == '}')
                {
                    p_++;
                    return true;
                }
                return false;
            }
        }
        if (*p_ == '[')
        {
            p_++;
            out.type = Json::Array;
            Skip();
            if (p_ < end_ && *p_ == ']')
            {
                p_++;
                return true;
            }
            for (;;)
            {
                out.array.emplace_back();
                if (!Value(out.array.back(), depth + 1))
                {
                    return false;
                }
                Skip();
                if (p_ < end_ && *p_ == ',')
                {
                    p_++;
                    continue;
                }
                if (p_ < end_ && *p_ == ']')
                {
                    p_++;
                    return true;
                }
                return false;
            }
        }
        if (*p_ == '"')
        {
            out.type = Json::String;
            return ParseString(out.string);
        }
        if (Literal("true"))
        {
            out.type = Json::Bool;
            out.boolean = true;
            return true;
        }
        if (Literal("false"))
        {
            out.type = Json::Bool;
            return true;
        }
        if (Literal("null"))
        {
            return true;
        }
        char *num_end = nullptr;
        std::string tmp(p_, std::min(end_, p_ + 64));
        double v = strtod(tmp.c_str(), &num_end);
        if (num_end == tmp.c_str())
        {
            return false;
        }
        p_ += num_end - tmp.c_str();
        out.type = Json::Number;
        out.number = v;
        return true;
    }
};

bool ReadFile(const std::string &path, std::vector<uint8_t> &out)
{
    std::ifstream f(path, std::ios::binary);
    if (!f)
    {
        return false;
    }
    f.seekg(0, std::ios::end);
    std::streamoff n = f.tellg();
    f.seekg(0);
    out.resize(static_cast<size_t>(n));
    f.read(reinterpret_cast<char *>(out.data()), n);
    return bool(f);
}

bool Base64(const std::string &s, size_t from, std::vector<uint8_t> &out)
{
    auto val = [](char c) -> int {
        if (c >= 'A' && c <= 'Z') return c - 'A';
        if (c >= 'a' && c <= 'z') return c - 'a' + 26;
        if (c >= '0' && c <= '9') return c - '0' + 52;
        if (c == '+' || c == '-') return 62;
        if (c == '/' || c == '_') return 63;
        return -1;
    };
    uint32_t acc = 0;
    int bits = 0;
    for (size_t i = from; i < s.size(); i++)
    {
        if (s[i] == '=')
        {
            break;
        }
        int v = val(s[i]);
        if (v < 0)
        {
            return false;
        }
        acc = (acc << 6) | static_cast<uint32_t>(v);
        bits += 6;
        if (bits >= 8)
        {
            bits -= 8;
            out.push_back(static_cast<uint8_t>(acc >> bits));
        }
    }
    return true;
}

// uri -> bytes: data URI or a file next to the .gltf
bool LoadUri(const std::string &dir, const std::string &uri, std::vector<uint8_t> &out)
{
    if (uri.rfind("data:", 0) == 0)
    {
        size_t comma = uri.find(',');
        return comma != std::string::npos && uri.find(";base64") < comma && Base64(uri, comma + 1, out);
    }
    return ReadFile(dir.empty() ? uri : dir + "/" + uri, out);
}

uint32_t Be32(const uint8_t *p) { return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | p[3]; }
} // namespace

bool DecodePng(const uint8_t *data, size_t size, std::vector<uint8_t> &rgba, uint32_t &width, uint32_t &height)
{
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (size < 8 || memcmp(data, sig, 8) != 0)
    {
        return false;
    }
    std::vector<uint8_t> idat, palette, trns;
    uint32_t bit_depth = 0, color_type = 0, interlace = 0;
    width = height = 0;
    for (size_t off = 8; off + 12 <= size;)
    {
        uint32_t len = Be32(data + off);
        const uint8_t *type = data + off + 4, *body = data + off + 8;
        if (off + 12 + size_t(len) > size)
        {
            return false;
        }
        if (!memcmp(type, "IHDR", 4) && len >= 13)
        {
            width = Be32(body);
            height = Be32(body + 4);
            bit_depth = body[8];
            color_type = body[9];
            interlace = body[12];
        }
        else if (!memcmp(type, "PLTE", 4))
        {
            palette.assign(body, body + len);
        }
        else if (!memcmp(type, "tRNS", 4))
        {
            trns.assign(body, body + len);
        }
        else if (!memcmp(type, "IDAT", 4))
        {
            idat.insert(idat.end(), body, body + len);
        }
        else if (!memcmp(type, "IEND", 4))
        {
            break;
        }
        off += 12 + size_t(len);
    }
    static const int channels_of[7] = {1, 0, 3, 1, 2, 0, 4};
    if (width == 0 || height == 0 || bit_depth != 8 || interlace != 0 || color_type > 6 || channels_of[color_type] == 0 ||
        uint64_t(width) * height > (1ull << 28))
    {
        return false;
    }
    const size_t ch = static_cast<size_t>(channels_of[color_type]), stride = size_t(width) * ch;
    std::vector<uint8_t> raw((stride + 1) * height);
    uLongf raw_len = static_cast<uLongf>(raw.size());
    if (uncompress(raw.data(), &raw_len, idat.data(), static_cast<uLong>(idat.size())) != Z_OK || raw_len != raw.size())
    {
        return false;
    }
    // undo the per-row filters in place (PNG spec 9.2)
    std::vector<uint8_t> prev(stride, 0);
    rgba.resize(size_t(width) * height * 4);
    for (uint32_t y = 0; y < height; y++)
    {
        uint8_t *row = raw.data() + (stride + 1) * y;
        const uint8_t filter = row[0];
        uint8_t *cur = row + 1;
        for (size_t x = 0; x < stride; x++)
        {
            const int a = x >= ch ? cur[x - ch] : 0, b = prev[x], c = x >= ch ? prev[x - ch] : 0;
            int pred = 0;
            switch (filter)
            {
            case 1: pred = a; break;
            case 2: pred = b; break;
            case 3: pred = (a + b) / 2; break;
            case 4: {
                const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
                pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
                break;
            }
            default: break;
            }
            cur[x] = static_cast<uint8_t>(cur[x] + pred);
        }
        memcpy(prev.data(), cur, stride);
        uint8_t *o = rgba.data() + size_t(y) * width * 4;
        for (uint32_t x = 0; x < width; x++)
        {
            const uint8_t *s = cur + size_t(x) * ch;
            switch (color_type)
            {
            case 0: o[4 * x] = o[4 * x + 1] = o[4 * x + 2] = s[0]; o[4 * x + 3] = 255; break;
            case 2: o[4 * x] = s[0]; o[4 * x + 1] = s[1]; o[4 * x + 2] = s[2]; o[4 * x + 3] = 255; break;
            case 3: {
                const size_t k = s[0];
                if (3 * k + 2 >= palette.size())
                {
                    return false;
                }
                o[4 * x] = palette[3 * k];
                o[4 * x + 1] = palette[3 * k + 1];
                o[4 * x + 2] = palette[3 * k + 2];
                o[4 * x + 3] = k < trns.size() ? trns[k] : 255;
                break;
            }
            case 4: o[4 * x] = o[4 * x + 1] = o[4 * x + 2] = s[0]; o[4 * x + 3] = s[1]; break;
            default: memcpy(o + 4 * x, s, 4); break;
            }
        }
    }
    return true;
}

namespace
{
// ---- accessors -----------------------------------------------------------------------------------------------------
struct Model
{
    Json root;
    std::string dir;
    std::vector<std::vector<uint8_t>> buffers;
    std::vector<bool> buffer_loaded;
};

struct View
{
    const uint8_t *data = nullptr;
    size_t count = 0, stride = 0;
    int component = 0, width = 0; // componentType, number of components
};

int ComponentBytes(int c)
{
    switch (c)
    {
    case 5120: case 5121: return 1;
    case 5122: case 5123: return 2;
    case 5125: case 5126: return 4;
    default: return 0;
    }
}

bool Fail(const std::string &msg)
{
    SetLastError("GLTFLoader: " + msg);
    return false;
}

bool GetView(Model &m, int accessor_id, View &out)
{
    const auto &accessors = m.root.Arr("accessors");
    if (accessor_id < 0 || size_t(accessor_id) >= accessors.size())
    {
        return Fail("accessor index out of range");
    }
    const Json &acc = accessors[size_t(accessor_id)];
    const std::string type = acc.Str("type");
    out.width = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : 0;
    out.component = acc.Int("componentType", 0);
    // counts, offsets and strides come from JSON doubles: range-check them BEFORE any cast or product (a negative or huge
    // value is undefined behaviour on the cast, and a wrapped product would pass the bounds check below)
    const double max_bytes = 4294967296.0 * 16.0; // far above any buffer this reader loads
    const double count_d = acc.Num("count", 0);
    if (!(count_d >= 0.0 && count_d <= max_bytes))
    {
        return Fail("accessor count out of range");
    }
    out.count = static_cast<size_t>(count_d);
    const int cb = ComponentBytes(out.component);
    const auto &views = m.root.Arr("bufferViews");
    const int bv = acc.Int("bufferView", -1);
    if (out.width == 0 || cb == 0 || bv < 0 || size_t(bv) >= views.size())
    {
        return Fail("unsupported accessor (type, componentType or missing bufferView)");
    }
    const Json &view = views[size_t(bv)];
    const int buf = view.Int("buffer", -1);
    const auto &buffers = m.root.Arr("buffers");
    if (buf < 0 || size_t(buf) >= buffers.size())
    {
        return Fail("bufferView references a missing buffer");
    }
    if (!m.buffer_loaded[size_t(buf)])
    {
        if (!LoadUri(m.dir, buffers[size_t(buf)].Str("uri"), m.buffers[size_t(buf)]))
        {
            return Fail("cannot read buffer " + buffers[size_t(buf)].Str("uri"));
        }
        m.buffer_loaded[size_t(buf)] = true;
    }
    const double view_off = view.Num("byteOffset", 0), acc_off = acc.Num("byteOffset", 0), stride_d = view.Num("byteStride", 0);
    const auto &bytes = m.buffers[size_t(buf)];
    const double size_d = double(bytes.size());
    if (!(view_off >= 0.0 && view_off <= size_d && acc_off >= 0.0 && acc_off <= size_d && stride_d >= 0.0 && stride_d <= size_d))
    {
        return Fail("accessor offset or stride out of range");
    }
    const size_t offset = static_cast<size_t>(view_off) + static_cast<size_t>(acc_off);
    const size_t elem = size_t(cb) * size_t(out.width);
    out.stride = static_cast<size_t>(stride_d);
    if (out.stride == 0)
    {
        out.stride = elem; // Accessor::ByteStride: tightly packed
    }
    // every term is <= bytes.size() and the count is bounded above, so the 128-bit product cannot wrap
    const unsigned __int128 last = (unsigned __int128)offset + (unsigned __int128)(out.count ? out.count - 1 : 0) * out.stride + elem;
    if (out.count && last > (unsigned __int128)bytes.size())
    {
        return Fail("accessor runs past the end of its buffer");
    }
    out.data = bytes.data() + offset;
    return true;
}

bool ReadFloats(Model &m, int accessor_id, int width, std::vector<float> &out)
{
    View v;
    if (!GetView(m, accessor_id, v))
    {
        return false;
    }
    if (v.width != width || v.component != 5126)
    {
        return Fail("vertex attribute must be float with the expected width");
    }
    out.resize(v.count * size_t(width));
    for (size_t i = 0; i < v.count; i++)
    {
        memcpy(out.data() + i * size_t(width), v.data + i * v.stride, size_t(width) * 4);
    }
    return true;
}

bool ReadIndices(Model &m, int accessor_id, std::vector<uint32_t> &out)
{
    View v;
    if (!GetView(m, accessor_id, v))
    {
        return false;
    }
    if (v.width != 1 || v.component == 5126)
    {
        return Fail("indices must be an integer SCALAR accessor");
    }
    out.resize(v.count);
    for (size_t i = 0; i < v.count; i++)
    {
        const uint8_t *p = v.data + i * v.stride;
        switch (v.component) // GLTFLoader.cpp:347-378: signed and unsigned 8/16/32-bit
        {
        case 5120: out[i] = static_cast<uint32_t>(*reinterpret_cast<const int8_t *>(p)); break;
        case 5121: out[i] = *p; break;
        case 5122: { int16_t t; memcpy(&t, p, 2); out[i] = static_cast<uint32_t>(t); break; }
        case 5123: { uint16_t t; memcpy(&t, p, 2); out[i] = t; break; }
        default: memcpy(&out[i], p, 4); break;
        }
    }
    return true;
}

// ---- node transforms ------------------------------------------------------------------------------------------------
// A node's local transform is Transform(translation, rotation, Ones) with the y-up -> z-up rotation in front
// (GLTFLoader.cpp:532-544), computed in the reference's float operation order by ref_math.cpp (bit-identical for nodes
// directly under the scene root: tests/test_host_cpu.py). Children compose like SceneNode::UpdateDirtyTransform
// (SceneNode.cpp:30-44): world = parent * local, world^-1 = local^-1 * parent^-1, as 3x4 affine products.
struct Affine
{
    float fwd[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    float inv[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    bool identity = true;
};
void Mul34(const float a[12], const float b[12], float out[12]) // a after b
{
    for (int i = 0; i < 3; i++)
    {
        for (int j = 0; j < 4; j++)
        {
            float v = (a[4 * i] * b[j] + a[4 * i + 1] * b[4 + j]) + a[4 * i + 2] * b[8 + j];
            out[4 * i + j] = j == 3 ? v + a[4 * i + 3] : v;
        }
    }
}
Affine Compose(const Affine &parent, const Affine &local)
{
    if (parent.identity)
    {
        Affine r = local; // identity * M is M bit for bit
        r.identity = false;
        return r;
    }
    Affine r;
    Mul34(parent.fwd, local.fwd, r.fwd);
    Mul34(local.inv, parent.inv, r.inv);
    r.identity = false;
    return r;
}

void ReadVec(const Json &node, const char *key, float *out, int n)
{
    const auto &a = node.Arr(key);
    for (int i = 0; i < n && size_t(i) < a.size(); i++)
    {
        out[i] = static_cast<float>(a[size_t(i)].number);
    }
}

// ---- images and materials (LoadMaterials, GLTFLoader.cpp:475-526) ---------------------------------------------------
struct Loader
{
    Model model;
    SceneProxy *scene = nullptr;
    GltfStats stats;
    std::map<std::pair<int, bool>, uint32_t> texture_cache; // (glTF texture, linear?) -> scene texture
    std::vector<uint32_t> materials;                        // glTF material -> scene material
    uint32_t default_material = SPK_INVALID_ID;
    bool instancing = false; // keep node.scale in the instance matrix instead of baking it (GltfOptions)
    struct MeshKey // what makes two glTF primitives the same triangles: their accessors and the baked scale
    {
        int accessors[5];
        float scale[3];
        bool operator<(const MeshKey &o) const
        {
            int c = memcmp(accessors, o.accessors, sizeof(accessors));
            return c != 0 ? c < 0 : memcmp(scale, o.scale, 12) < 0;
        }
    };
    std::map<MeshKey, std::pair<uint32_t, uint32_t>> mesh_cache; // -> (scene mesh, triangle count)

    // CreateTexture (:283-327): textures[i].source -> images[j] -> RGBA8; sRGB for colour, Unorm for data
    bool Texture(int texture_index, bool linear, uint32_t &out)
    {
        auto hit = texture_cache.find({texture_index, linear});
        if (hit != texture_cache.end())
        {
            out = hit->second;
            return true;
        }
        const auto &textures = model.root.Arr("textures");
        const auto &images = model.root.Arr("images");
        if (texture_index < 0 || size_t(texture_index) >= textures.size())
        {
            return Fail("texture index out of range");
        }
        const int source = textures[size_t(texture_index)].Int("source", -1);
        if (source < 0 || size_t(source) >= images.size())
        {
            return Fail("texture without an image source");
        }
        const Json &image = images[size_t(source)];
        std::vector<uint8_t> bytes;
        const std::string uri = image.Str("uri");
        if (!uri.empty())
        {
            if (!LoadUri(model.dir, uri, bytes))
            {
                return Fail("texture image has no data: " + uri); // GLTFLoader.cpp:306-309
            }
        }
        else
        {
            const auto &views = model.root.Arr("bufferViews");
            const int bv = image.Int("bufferView", -1);
            if (bv < 0 || size_t(bv) >= views.size())
            {
                return Fail("image has neither uri nor bufferView");
            }
            const Json &view = views[size_t(bv)];
            const int buf = view.Int("buffer", -1);
            const auto &buffers = model.root.Arr("buffers");
            if (buf < 0 || size_t(buf) >= buffers.size())
            {
                return Fail("image bufferView references a missing buffer");
            }
            if (!model.buffer_loaded[size_t(buf)])
            {
                if (!LoadUri(model.dir, buffers[size_t(buf)].Str("uri"), model.buffers[size_t(buf)]))
                {
                    return Fail("cannot read buffer " + buffers[size_t(buf)].Str("uri"));
                }
                model.buffer_loaded[size_t(buf)] = true;
            }
            const size_t off = static_cast<size_t>(view.Num("byteOffset", 0)), len = static_cast<size_t>(view.Num("byteLength", 0));
            const auto &b = model.buffers[size_t(buf)];
            if (off + len > b.size())
            {
                return Fail("image bufferView runs past the end of its buffer");
            }
            bytes.assign(b.begin() + static_cast<std::ptrdiff_t>(off), b.begin() + static_cast<std::ptrdiff_t>(off + len));
        }
        std::vector<uint8_t> rgba;
        uint32_t w = 0, h = 0;
        if (!DecodePng(bytes.data(), bytes.size(), rgba, w, h))
        {
            return Fail("only 8-bit non-interlaced PNG images are decoded here: " + (uri.empty() ? std::string("<embedded>") : uri));
        }
        out = scene->AddTexture(rgba.data(), w, h, linear ? SPK_FMT_RGBA8_UNORM : SPK_FMT_RGBA8_SRGB);
        texture_cache[{texture_index, linear}] = out;
        stats.textures++;
        return true;
    }

    static SpkMaterialDesc DefaultMaterial()
    {
        // MaterialResource (io/Material.h:7-22): base colour 1, emissive 0, metallic 1, roughness 1, eta 0, no textures
        SpkMaterialDesc m{};
        m.base_color_tex = m.emissive_tex = m.metallic_roughness_tex = m.normal_tex = SPK_INVALID_ID;
        m.metallic = 1.f;
        m.roughness = 1.f;
        m.eta = 0.f;
        m.type = SPK_MAT_PBR;
        m.base_color[0] = m.base_color[1] = m.base_color[2] = 1.f;
        return m;
    }

    bool LoadMaterials()
    {
        for (const Json &mat : model.root.Arr("materials"))
        {
            SpkMaterialDesc m = DefaultMaterial();
            static const Json empty;
            const Json *pbr = mat.Find("pbrMetallicRoughness");
            pbr = pbr ? pbr : &empty;
            // factors are read only next to their texture (GLTFLoader.cpp:485-517)
            if (const Json *t = pbr->Find("baseColorTexture"); t && t->Int("index", -1) >= 0)
            {
                if (!Texture(t->Int("index", -1), false, m.base_color_tex))
                {
                    return false;
                }
                float f[4] = {1.f, 1.f, 1.f, 1.f}; // glTF default baseColorFactor
                ReadVec(*pbr, "baseColorFactor", f, 4);
                memcpy(m.base_color, f, 12);
            }
            if (const Json *t = mat.Find("normalTexture"); t && t->Int("index", -1) >= 0)
            {
                if (!Texture(t->Int("index", -1), true, m.normal_tex))
                {
                    return false;
                }
            }
            if (const Json *t = pbr->Find("metallicRoughnessTexture"); t && t->Int("index", -1) >= 0)
            {
                if (!Texture(t->Int("index", -1), true, m.metallic_roughness_tex))
                {
                    return false;
                }
                m.metallic = static_cast<float>(pbr->Num("metallicFactor", 1.0));
                m.roughness = static_cast<float>(pbr->Num("roughnessFactor", 1.0));
            }
            if (const Json *t = mat.Find("emissiveTexture"); t && t->Int("index", -1) >= 0)
            {
                if (!Texture(t->Int("index", -1), false, m.emissive_tex))
                {
                    return false;
                }
                float f[3] = {0.f, 0.f, 0.f}; // glTF default emissiveFactor
                ReadVec(mat, "emissiveFactor", f, 3);
                memcpy(m.emissive, f, 12);
            }
            materials.push_back(scene->AddMaterial(m));
            stats.materials++;
        }
        return true;
    }

    // LoadPrimitive (GLTFLoader.cpp:329-473) with the node scale baked into the positions (LoadPositionArray :192-247)
    bool Primitive(int mesh_id, int prim_id, const Json &prim, const float scale[3], uint32_t &mesh_out, uint32_t &tris)
    {
        (void)mesh_id;
        (void)prim_id;
        const Json *attrs = prim.Find("attributes");
        if (!attrs || attrs->Int("POSITION", -1) < 0)
        {
            return Fail("primitive without POSITION");
        }
        MeshKey key{{attrs->Int("POSITION", -1), attrs->Int("NORMAL", -1), attrs->Int("TANGENT", -1),
                     attrs->Int("TEXCOORD_0", -1), prim.Int("indices", -1)},
                    {scale[0], scale[1], scale[2]}};
        auto hit = mesh_cache.find(key);
        if (hit != mesh_cache.end())
        {
            mesh_out = hit->second.first;
            tris = hit->second.second;
            return true;
        }
        MeshData md;
        if (!ReadFloats(model, attrs->Int("POSITION", -1), 3, md.vertices))
        {
            return false;
        }
        for (size_t v = 0; v < md.vertices.size(); v += 3)
        {
            for (int k = 0; k < 3; k++)
            {
                md.vertices[v + size_t(k)] *= scale[k]; // positions[i].cwiseProduct(scale)
            }
        }
        const size_t nv = md.vertices.size() / 3;
        if (attrs->Int("NORMAL", -1) < 0)
        {
            return Fail("primitive without NORMAL: geometric-normal shading (io/Mesh.h:86-93) is not built");
        }
        if (!ReadFloats(model, attrs->Int("NORMAL", -1), 3, md.normals))
        {
            return false;
        }
        if (attrs->Int("TANGENT", -1) >= 0)
        {
            if (!ReadFloats(model, attrs->Int("TANGENT", -1), 4, md.tangents))
            {
                return false;
            }
        }
        else
        {
            md.tangents.resize(nv * 4); // Mesh::GetTangent without tangents: the constant (Right, 1) (io/Mesh.h:104-109)
            for (size_t v = 0; v < nv; v++)
            {
                md.tangents[4 * v] = 1.f;
                md.tangents[4 * v + 1] = 0.f;
                md.tangents[4 * v + 2] = 0.f;
                md.tangents[4 * v + 3] = 1.f;
            }
        }
        if (attrs->Int("TEXCOORD_0", -1) >= 0 && !ReadFloats(model, attrs->Int("TEXCOORD_0", -1), 2, md.uvs))
        {
            return false;
        }
        if (md.normals.size() != nv * 3 || md.tangents.size() != nv * 4 || (!md.uvs.empty() && md.uvs.size() != nv * 2))
        {
            return Fail("vertex attribute counts differ");
        }
        if (prim.Int("indices", -1) >= 0)
        {
            if (!ReadIndices(model, prim.Int("indices", -1), md.indices))
            {
                return false;
            }
        }
        else
        {
            md.indices.resize(nv);
            for (size_t v = 0; v < nv; v++)
            {
                md.indices[v] = static_cast<uint32_t>(v);
            }
        }
        if (md.indices.size() < 3 || md.indices.size() % 3 != 0)
        {
            return Fail("index count is not a multiple of 3");
        }
        for (uint32_t idx : md.indices)
        {
            if (idx >= nv)
            {
                return Fail("vertex index out of range");
            }
        }
        tris = static_cast<uint32_t>(md.indices.size() / 3);
        mesh_out = scene->AddMesh(std::move(md)); // BLAS::Build
        mesh_cache[key] = {mesh_out, tris};
        stats.unique_meshes++;
        stats.unique_triangles += tris;
        return true;
    }

    // ProcessNode (GLTFLoader.cpp:528-575)
    bool Node(int node_id, const Affine &parent, int depth)
    {
        const auto &nodes = model.root.Arr("nodes");
        if (node_id < 0 || size_t(node_id) >= nodes.size() || depth > 256)
        {
            return Fail("node index out of range or hierarchy too deep");
        }
        const Json &node = nodes[size_t(node_id)];
        stats.nodes++;
        float t[3] = {0.f, 0.f, 0.f}, s[3] = {1.f, 1.f, 1.f}, q[4] = {0.f, 0.f, 0.f, 1.f};
        ReadVec(node, "translation", t, 3);
        ReadVec(node, "scale", s, 3);
        ReadVec(node, "rotation", q, 4);
        // glTF is y-up, the scene z-up: AngleAxis(90 deg, Right) * rotation, then normalized (:537-544)
        Affine local;
        static const float ones[3] = {1.f, 1.f, 1.f};
        refmath::TrsMatricesQuat(t, refmath::GltfNodeRotation(q), ones, local.fwd, local.inv); // scale stays Ones
        const Affine world = Compose(parent, local);
        const int mesh_id = node.Int("mesh", -1);
        if (mesh_id >= 0)
        {
            const auto &meshes = model.root.Arr("meshes");
            if (size_t(mesh_id) >= meshes.size())
            {
                return Fail("node references a missing mesh");
            }
            const auto &prims = meshes[size_t(mesh_id)].Arr("primitives");
            for (size_t p = 0; p < prims.size(); p++)
            {
                const int mode = prims[p].Int("mode", 4);
                if (mode != 4)
                {
                    // points / lines are ignored with an error log like the reference (:452-457); strips and fans are
                    // not converted here
                    SetLastError("GLTFLoader: primitive is not a triangle list, ignoring");
                    continue;
                }
                uint32_t mesh = 0, tris = 0;
                if (!Primitive(mesh_id, static_cast<int>(p), prims[p], instancing ? ones : s, mesh, tris))
                {
                    return false;
                }
                const int mat = prims[p].Int("material", -1);
                uint32_t material;
                if (mat >= 0 && size_t(mat) < materials.size())
                {
                    material = materials[size_t(mat)];
                }
                else
                {
                    if (default_material == SPK_INVALID_ID)
                    {
                        default_material = scene->AddMaterial(DefaultMaterial());
                    }
                    material = default_material;
                }
                float fwd[12], inv[12];
                memcpy(fwd, world.fwd, sizeof(fwd));
                memcpy(inv, world.inv, sizeof(inv));
                if (instancing)
                {
                    // world * Scaling(s): forward columns scaled, inverse rows divided
                    for (int i = 0; i < 3; i++)
                    {
                        for (int j = 0; j < 3; j++)
                        {
                            fwd[4 * i + j] *= s[j];
                        }
                        for (int j = 0; j < 4; j++)
                        {
                            inv[4 * i + j] /= s[i];
                        }
                    }
                }
                scene->AddMeshPrimitiveMatrix(mesh, material, fwd, inv);
                stats.primitives++;
                stats.instanced_triangles += tris;
            }
        }
        for (const Json &child : node.Arr("children"))
        {
            if (!Node(static_cast<int>(child.number), world, depth + 1))
            {
                return false;
            }
        }
        return true;
    }
};
} // namespace

bool LoadGltf(const std::string &path, SceneProxy &scene, GltfStats *stats, const GltfOptions &options)
{
    std::vector<uint8_t> text;
    if (!ReadFile(path, text))
    {
        return Fail("failed to find model " + path);
    }
    Loader loader;
    loader.scene = &scene;
    loader.instancing = options.instancing;
    const char *begin = reinterpret_cast<const char *>(text.data());
    if (!JsonParser(begin, begin + text.size()).Parse(loader.model.root) || loader.model.root.type != Json::Object)
    {
        return Fail("failed to load model from " + path + ": not a JSON object");
    }
    const size_t slash = path.find_last_of('/');
    loader.model.dir = slash == std::string::npos ? std::string() : path.substr(0, slash);
    const size_t nbuf = loader.model.root.Arr("buffers").size();
    loader.model.buffers.resize(nbuf);
    loader.model.buffer_loaded.assign(nbuf, false);
    if (!loader.LoadMaterials())
    {
        return false;
    }
    // the default scene's roots, or every node of a file without scenes (GLTFLoader.cpp:619-636)
    std::vector<int> roots;
    const auto &scenes = loader.model.root.Arr("scenes");
    if (scenes.empty())
    {
        for (size_t i = 0; i < loader.model.root.Arr("nodes").size(); i++)
        {
            roots.push_back(static_cast<int>(i));
        }
    }
    else
    {
        int which = loader.model.root.Int("scene", -1);
        which = which > -1 && size_t(which) < scenes.size() ? which : 0;
        for (const Json &n : scenes[size_t(which)].Arr("nodes"))
        {
            roots.push_back(static_cast<int>(n.number));
        }
    }
    for (int r : roots)
    {
        if (!loader.Node(r, Affine{}, 0))
        {
            return false;
        }
    }
    if (stats)
    {
        *stats = loader.stats;
    }
    return true;
}
} // namespace sparkle_cuda

extern "C" int spk_host_gltf_probe(const char *path, int instancing, uint64_t *stats7)
{
    sparkle_cuda::SceneProxy scene;
    sparkle_cuda::GltfStats st;
    sparkle_cuda::GltfOptions opt;
    opt.instancing = instancing != 0;
    if (!path || !stats7 || !sparkle_cuda::LoadGltf(path, scene, &st, opt))
    {
        return -1;
    }
    const uint64_t v[7] = {st.nodes, st.primitives, st.unique_meshes, st.materials, st.textures, st.instanced_triangles,
                           st.unique_triangles};
    memcpy(stats7, v, sizeof(v));
    return 0;
}


extern "C" SpkHostScene *spk_host_scene_load_gltf(const char *path, int instancing)
{
    auto s = std::make_unique<SpkHostScene>();
    sparkle_cuda::GltfOptions opt;
    opt.instancing = instancing != 0;
    if (!path || !sparkle_cuda::LoadGltf(path, s->scene, nullptr, opt))
    {
        return nullptr;
    }
    s->scene.SetSky(sparkle_cuda::Vec3{0.5f, 0.7f, 1.0f});
    s->config.pipeline = sparkle_cuda::RenderConfig::Pipeline::Cuda; // NeedBvh(): SceneRenderProxy.cpp:192 + Cuda
    s->scene.Update(s->config);                                     // TLAS::Build + the flattened description
    s->scene.EndUpdate();
    return s.release();
}

extern "C" const SpkSceneDesc *spk_host_scene_desc(SpkHostScene *s) { return s ? &s->scene.Desc() : nullptr; }
extern "C" void spk_host_scene_free(SpkHostScene *s) { delete s; }

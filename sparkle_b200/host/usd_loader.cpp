/*
 * usd_loader.cpp — see usd_loader.h. A small reader for the ASCII-USD subset the reference's USDExporter.cpp writes, and
 * the scene conversion rules of its USDLoader.cpp (cited per function). No tinyusdz, no Tydra: prims, properties and
 * values are parsed generically (numbers of any tuple / array nesting are flattened in order), then read by name.
 */
#include "usd_loader.h"

#include "gltf_loader.h" // DecodePng

#include <cmath>
#include <cstdlib>
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
bool Fail(const std::string &msg)
{
    SetLastError("USDLoader: " + msg);
    return false;
}

// ---- ASCII USD: prims, properties, values ---------------------------------------------------------------------------
struct UsdValue
{
    std::vector<double> numbers; // every number of the value, tuples and arrays flattened in order
    std::string text;            // string / token / @asset@ / <path> (the first one of the value)
    bool has_text = false;
};

struct UsdProperty
{
    std::string type; // "color3f", "float3[]", "rel", ...
    UsdValue value;
    bool authored = false; // had "= value"
};

struct UsdPrim
{
    std::string type, name, display_name;
    std::map<std::string, UsdProperty> props; // key as written, ".connect" included
    std::vector<std::unique_ptr<UsdPrim>> children;
    const UsdPrim *Child(const std::string &n) const
    {
        for (const auto &c : children)
        {
            if (c->name == n)
            {
                return c.get();
            }
        }
        return nullptr;
    }
    const UsdProperty *Prop(const std::string &n) const
    {
        auto it = props.find(n);
        return it == props.end() ? nullptr : &it->second;
    }
};

class UsdParser
{
public:
    explicit UsdParser(const std::string &text) : s_(text) {}

    bool Parse(std::vector<std::unique_ptr<UsdPrim>> &roots, std::string &default_prim, std::string &error)
    {
        while (p_ < s_.size() && isspace(static_cast<unsigned char>(s_[p_])))
        {
            p_++;
        }
        if (s_.compare(p_, 5, "#usda") != 0)
        {
            error = "not an ASCII USD file (#usda header missing)";
            return false;
        }
        while (p_ < s_.size() && s_[p_] != '\n')
        {
            p_++;
        }
        SkipSpace();
        if (Peek() == '(')
        {
            std::map<std::string, UsdValue> meta;
            if (!ParseMetadata(meta))
            {
                error = error_;
                return false;
            }
            auto it = meta.find("defaultPrim");
            if (it != meta.end())
            {
                default_prim = it->second.text;
            }
        }
        for (;;)
        {
            SkipSpace();
            if (p_ >= s_.size())
            {
                break;
            }
            auto prim = ParsePrim();
            if (!prim)
            {
                error = error_ + " near offset " + std::to_string(p_);
                return false;
            }
            roots.push_back(std::move(prim));
        }
        return true;
    }

private:
    const std::string &s_;
    size_t p_ = 0;
    std::string error_;

    char Peek() const { return p_ < s_.size() ? s_[p_] : '\0'; }

    void SkipSpace()
    {
        for (;;)
        {
            while (p_ < s_.size() && (isspace(static_cast<unsigned char>(s_[p_])) || s_[p_] == ';'))
            {
                p_++;
            }
            if (p_ < s_.size() && s_[p_] == '#')
            {
                while (p_ < s_.size() && s_[p_] != '\n')
                {
                    p_++;
                }
                continue;
            }
            return;
        }
    }

    // identifiers with namespaces, array brackets and connection suffixes: "float3[]", "primvars:st", "inputs:st.connect"
    std::string Word()
    {
        SkipSpace();
        size_t b = p_;
        while (p_ < s_.size())
        {
            const char c = s_[p_];
            if (isalnum(static_cast<unsigned char>(c)) || c == '_' || c == ':' || c == '.')
            {
                p_++;
            }
            else if (c == '[' && p_ + 1 < s_.size() && s_[p_ + 1] == ']')
            {
                p_ += 2;
            }
            else
            {
                break;
            }
        }
        return s_.substr(b, p_ - b);
    }

    bool String(std::string &out)
    {
        SkipSpace();
        const char q = Peek();
        if (q != '"' && q != '\'')
        {
            return false;
        }
        // triple-quoted strings are taken whole
        if (s_.compare(p_, 3, std::string(3, q)) == 0)
        {
            const size_t e = s_.find(std::string(3, q), p_ + 3);
            if (e == std::string::npos)
            {
                return false;
            }
            out = s_.substr(p_ + 3, e - p_ - 3);
            p_ = e + 3;
            return true;
        }
        p_++;
        out.clear();
        while (p_ < s_.size() && s_[p_] != q)
        {
            if (s_[p_] == '\\' && p_ + 1 < s_.size())
            {
                p_++;
            }
            out.push_back(s_[p_++]);
        }
        if (p_ >= s_.size())
        {
            return false;
        }
        p_++;
        return true;
    }

    // one value: scalar, string, @asset@, <path>, (tuple), [array], nested; numbers are appended in reading order
    bool Value(UsdValue &v)
    {
        SkipSpace();
        const char c = Peek();
        if (c == '(' || c == '[')
        {
            const char close = c == '(' ? ')' : ']';
            p_++;
            for (;;)
            {
                SkipSpace();
                if (Peek() == close)
                {
                    p_++;
                    return true;
                }
                if (Peek() == ',')
                {
                    p_++;
                    continue;
                }
                if (p_ >= s_.size() || !Value(v))
                {
                    error_ = error_.empty() ? "unterminated tuple or array" : error_;
                    return false;
                }
            }
        }
        if (c == '"' || c == '\'')
        {
            std::string t;
            if (!String(t))
            {
                error_ = "unterminated string";
                return false;
            }
            if (!v.has_text)
            {
                v.text = t;
                v.has_text = true;
            }
            return true;
        }
        if (c == '@' || c == '<')
        {
            const char close = c == '@' ? '@' : '>';
            const size_t e = s_.find(close, p_ + 1);
            if (e == std::string::npos)
            {
                error_ = "unterminated asset or path";
                return false;
            }
            if (!v.has_text)
            {
                v.text = s_.substr(p_ + 1, e - p_ - 1);
                v.has_text = true;
            }
            p_ = e + 1;
            return true;
        }
        if (isdigit(static_cast<unsigned char>(c)) || c == '-' || c == '+' || c == '.')
        {
            char *end = nullptr;
            const double d = strtod(s_.c_str() + p_, &end);
            if (end == s_.c_str() + p_)
            {
                error_ = "bad number";
                return false;
            }
            // "inf" / "nan" spellings are words, handled below
            p_ = size_t(end - s_.c_str());
            v.numbers.push_back(d);
            return true;
        }
        const std::string w = Word();
        if (w.empty())
        {
            error_ = std::string("unexpected character '") + c + "'";
            return false;
        }
        if (w == "inf")
        {
            v.numbers.push_back(INFINITY);
        }
        else if (w == "nan")
        {
            v.numbers.push_back(NAN);
        }
        else if (!v.has_text) // None, true/false, bare tokens
        {
            v.text = w;
            v.has_text = true;
        }
        return true;
    }

    // "( key = value ... )": prim / property / layer metadata. Doc strings and unknown keys are skipped by shape.
    bool ParseMetadata(std::map<std::string, UsdValue> &meta)
    {
        p_++; // '('
        for (;;)
        {
            SkipSpace();
            if (Peek() == ')')
            {
                p_++;
                return true;
            }
            if (p_ >= s_.size())
            {
                error_ = "unterminated metadata";
                return false;
            }
            if (Peek() == '"' || Peek() == '\'')
            {
                std::string doc; // a bare string is the comment metadata
                if (!String(doc))
                {
                    error_ = "unterminated string";
                    return false;
                }
                continue;
            }
            std::string key = Word();
            if (key.empty())
            {
                error_ = "bad metadata";
                return false;
            }
            // list-edit qualifiers and typed dictionaries: "prepend apiSchemas = [...]", "dictionary x = {...}"
            if (key == "prepend" || key == "append" || key == "add" || key == "delete" || key == "reorder")
            {
                key = Word();
            }
            SkipSpace();
            if (Peek() != '=')
            {
                continue; // a flag without a value
            }
            p_++;
            SkipSpace();
            if (Peek() == '{')
            {
                if (!SkipBraces())
                {
                    return false;
                }
                continue;
            }
            UsdValue v;
            if (!Value(v))
            {
                return false;
            }
            meta[key] = std::move(v);
        }
    }

    bool SkipBraces()
    {
        int depth = 0;
        while (p_ < s_.size())
        {
            const char c = s_[p_];
            if (c == '"' || c == '\'')
            {
                std::string t;
                if (!String(t))
                {
                    error_ = "unterminated string";
                    return false;
                }
                continue;
            }
            p_++;
            if (c == '{')
            {
                depth++;
            }
            else if (c == '}' && --depth == 0)
            {
                return true;
            }
        }
        error_ = "unterminated dictionary";
        return false;
    }

    std::unique_ptr<UsdPrim> ParsePrim()
    {
        const std::string spec = Word();
        if (spec != "def" && spec != "over" && spec != "class")
        {
            error_ = "expected def / over / class, got '" + spec + "'";
            return nullptr;
        }
        auto prim = std::make_unique<UsdPrim>();
        SkipSpace();
        if (Peek() != '"')
        {
            prim->type = Word();
        }
        if (!String(prim->name))
        {
            error_ = "prim name expected";
            return nullptr;
        }
        SkipSpace();
        if (Peek() == '(')
        {
            std::map<std::string, UsdValue> meta;
            if (!ParseMetadata(meta))
            {
                return nullptr;
            }
            auto it = meta.find("displayName");
            if (it != meta.end())
            {
                prim->display_name = it->second.text;
            }
        }
        SkipSpace();
        if (Peek() != '{')
        {
            error_ = "'{' expected after prim " + prim->name;
            return nullptr;
        }
        p_++;
        for (;;)
        {
            SkipSpace();
            if (Peek() == '}')
            {
                p_++;
                return prim;
            }
            if (p_ >= s_.size())
            {
                error_ = "unterminated prim " + prim->name;
                return nullptr;
            }
            const size_t mark = p_;
            std::string w = Word();
            if (w == "def" || w == "over" || w == "class")
            {
                p_ = mark;
                auto child = ParsePrim();
                if (!child)
                {
                    return nullptr;
                }
                prim->children.push_back(std::move(child));
                continue;
            }
            if (w == "variantSet" || w == "reorder")
            {
                error_ = "variant sets and reorder statements are outside the exported subset";
                return nullptr;
            }
            // property: [custom] [uniform|varying] (rel | type) name [= value] [(metadata)]
            while (w == "custom" || w == "uniform" || w == "varying" || w == "prepend" || w == "append" || w == "add" ||
                   w == "delete")
            {
                w = Word();
            }
            UsdProperty prop;
            prop.type = w;
            const std::string name = Word();
            if (w.empty() || name.empty())
            {
                error_ = "property expected in prim " + prim->name;
                return nullptr;
            }
            SkipSpace();
            if (Peek() == '=')
            {
                p_++;
                if (!Value(prop.value))
                {
                    return nullptr;
                }
                prop.authored = true;
            }
            SkipSpace();
            if (Peek() == '(')
            {
                std::map<std::string, UsdValue> meta;
                if (!ParseMetadata(meta))
                {
                    return nullptr;
                }
            }
            prim->props[name] = std::move(prop);
        }
    }
};

// ---- scene conversion -------------------------------------------------------------------------------------------------
struct Mat4
{
    float m[4][4]; // column-vector convention, m[row][col]
};

Mat4 Identity()
{
    Mat4 r{};
    for (int i = 0; i < 4; i++)
    {
        r.m[i][i] = 1.f;
    }
    return r;
}

Mat4 Mul(const Mat4 &a, const Mat4 &b)
{
    Mat4 r{};
    for (int i = 0; i < 4; i++)
    {
        for (int j = 0; j < 4; j++)
        {
            float acc = 0.f;
            for (int k = 0; k < 4; k++)
            {
                acc += a.m[i][k] * b.m[k][j];
            }
            r.m[i][j] = acc;
        }
    }
    return r;
}

// affine inverse of the upper 3x4 (double arithmetic, one rounding per element)
bool AffineInverse(const Mat4 &a, float inv[12])
{
    double m[3][3], t[3];
    for (int i = 0; i < 3; i++)
    {
        for (int j = 0; j < 3; j++)
        {
            m[i][j] = a.m[i][j];
        }
        t[i] = a.m[i][3];
    }
    const double det = m[0][0] * (m[1][1] * m[2][2] - m[1][2] * m[2][1]) - m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) +
                       m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
    if (!(std::fabs(det) > 1e-30))
    {
        return false;
    }
    double r[3][3];
    r[0][0] = (m[1][1] * m[2][2] - m[1][2] * m[2][1]) / det;
    r[0][1] = (m[0][2] * m[2][1] - m[0][1] * m[2][2]) / det;
    r[0][2] = (m[0][1] * m[1][2] - m[0][2] * m[1][1]) / det;
    r[1][0] = (m[1][2] * m[2][0] - m[1][0] * m[2][2]) / det;
    r[1][1] = (m[0][0] * m[2][2] - m[0][2] * m[2][0]) / det;
    r[1][2] = (m[0][2] * m[1][0] - m[0][0] * m[1][2]) / det;
    r[2][0] = (m[1][0] * m[2][1] - m[1][1] * m[2][0]) / det;
    r[2][1] = (m[0][1] * m[2][0] - m[0][0] * m[2][1]) / det;
    r[2][2] = (m[0][0] * m[1][1] - m[0][1] * m[1][0]) / det;
    for (int i = 0; i < 3; i++)
    {
        for (int j = 0; j < 3; j++)
        {
            inv[4 * i + j] = float(r[i][j]);
        }
        inv[4 * i + 3] = float(-(r[i][0] * t[0] + r[i][1] * t[1] + r[i][2] * t[2]));
    }
    return true;
}

struct Converter
{
    SceneProxy &scene;
    UsdStats stats;
    std::string dir; // directory of the .usda: asset paths are relative to it (USDLoader.cpp:622-623)
    const std::vector<std::unique_ptr<UsdPrim>> &roots;
    std::map<std::string, uint32_t> material_ids;  // prim path -> material
    std::map<std::string, uint32_t> texture_ids;   // file + colour space -> texture
    bool have_camera = false;

    const UsdPrim *Resolve(const std::string &path) const
    {
        // absolute prim path "/a/b/c" (a property suffix ".x" is cut by the caller)
        const UsdPrim *cur = nullptr;
        size_t at = 0;
        while (at < path.size())
        {
            if (path[at] == '/')
            {
                at++;
                continue;
            }
            size_t e = path.find('/', at);
            const std::string part = path.substr(at, e == std::string::npos ? std::string::npos : e - at);
            if (!cur)
            {
                for (const auto &r : roots)
                {
                    if (r->name == part)
                    {
                        cur = r.get();
                    }
                }
            }
            else
            {
                cur = cur->Child(part);
            }
            if (!cur)
            {
                return nullptr;
            }
            at = e == std::string::npos ? path.size() : e;
        }
        return cur;
    }

    static bool Number(const UsdPrim &p, const char *name, float &out)
    {
        const UsdProperty *a = p.Prop(name);
        if (!a || !a->authored || a->value.numbers.empty())
        {
            return false;
        }
        out = float(a->value.numbers[0]);
        return true;
    }

    static bool Color(const UsdPrim &p, const char *name, float out[3])
    {
        const UsdProperty *a = p.Prop(name);
        if (!a || !a->authored || a->value.numbers.size() < 3)
        {
            return false;
        }
        for (int k = 0; k < 3; k++)
        {
            out[k] = float(a->value.numbers[size_t(k)]);
        }
        return true;
    }

    // CreateTexture (USDLoader.cpp:166-236): 8-bit images, sRGB formats only for gamma colour spaces; 1 / 3 channels are
    // expanded to RGBA by the PNG decoder
    uint32_t Texture(const UsdPrim &material, const std::string &connection, bool &ok)
    {
        // connection: </_materials/m/base_color_tex.outputs:rgb>
        const size_t dot = connection.rfind('.');
        const std::string shader_path = connection.substr(0, dot);
        const UsdPrim *shader = Resolve(shader_path);
        (void)material;
        const UsdProperty *file = shader ? shader->Prop("inputs:file") : nullptr;
        if (!shader || !file || !file->value.has_text)
        {
            ok = Fail("texture connection " + connection + " does not lead to a UsdUVTexture with inputs:file");
            return SPK_INVALID_ID;
        }
        const UsdProperty *cs = shader->Prop("inputs:sourceColorSpace");
        const bool srgb = cs && cs->value.text == "sRGB";
        const std::string key = file->value.text + (srgb ? "|srgb" : "|raw");
        auto it = texture_ids.find(key);
        if (it != texture_ids.end())
        {
            return it->second;
        }
        const std::string full = dir.empty() ? file->value.text : dir + "/" + file->value.text;
        std::ifstream in(full, std::ios::binary);
        std::vector<uint8_t> bytes((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
        std::vector<uint8_t> rgba;
        uint32_t w = 0, h = 0;
        if (!in || bytes.empty() || !DecodePng(bytes.data(), bytes.size(), rgba, w, h))
        {
            ok = Fail("texture image has no data: " + full + " (8-bit PNG expected)");
            return SPK_INVALID_ID;
        }
        const uint32_t id = scene.AddTexture(rgba.data(), w, h, srgb ? SPK_FMT_RGBA8_SRGB : SPK_FMT_RGBA8_UNORM);
        texture_ids[key] = id;
        stats.textures++;
        return id;
    }

    // LoadMesh's material half (USDLoader.cpp:349-425)
    bool Material(const std::string &path, uint32_t &out)
    {
        auto it = material_ids.find(path);
        if (it != material_ids.end())
        {
            out = it->second;
            return true;
        }
        const UsdPrim *mat = Resolve(path);
        if (!mat || mat->type != "Material")
        {
            return Fail("material binding " + path + " does not name a Material prim");
        }
        // MaterialResource defaults (io/Material.h:7-22)
        SpkMaterialDesc m{};
        m.base_color_tex = m.emissive_tex = m.metallic_roughness_tex = m.normal_tex = SPK_INVALID_ID;
        m.base_color[0] = m.base_color[1] = m.base_color[2] = 1.f;
        m.metallic = 1.f;
        m.roughness = 1.f;
        m.type = SPK_MAT_PBR;
        const UsdPrim *surface = nullptr;
        for (const auto &c : mat->children)
        {
            const UsdProperty *id = c->Prop("info:id");
            if (c->type == "Shader" && id && id->value.text == "UsdPreviewSurface")
            {
                surface = c.get();
            }
        }
        bool dielectric = false;
        bool ok = true;
        if (surface)
        {
            // UsdPreviewSurface fallbacks when an input is not authored (the converter's defaults)
            float opacity = 1.f, ior = 1.5f;
            Number(*surface, "inputs:opacity", opacity);
            Number(*surface, "inputs:ior", ior);
            dielectric = !surface->Prop("inputs:opacity.connect") && opacity < 1.f;
            auto connected = [&](const char *name) -> const UsdProperty * {
                const UsdProperty *c = surface->Prop(std::string(name) + ".connect");
                return c && c->value.has_text ? c : nullptr;
            };
            if (const UsdProperty *c = connected("inputs:diffuseColor"))
            {
                m.base_color_tex = Texture(*mat, c->value.text, ok);
            }
            else
            {
                float v[3] = {0.18f, 0.18f, 0.18f};
                Color(*surface, "inputs:diffuseColor", v);
                memcpy(m.base_color, v, 12);
            }
            if (const UsdProperty *c = connected("inputs:emissiveColor"))
            {
                m.emissive_tex = Texture(*mat, c->value.text, ok);
                m.emissive[0] = m.emissive[1] = m.emissive[2] = 1.f; // the default of zero would erase the texture
            }
            else
            {
                float v[3] = {0.f, 0.f, 0.f};
                Color(*surface, "inputs:emissiveColor", v);
                memcpy(m.emissive, v, 12);
            }
            if (const UsdProperty *c = connected("inputs:normal"))
            {
                m.normal_tex = Texture(*mat, c->value.text, ok);
            }
            const UsdProperty *mc = connected("inputs:metallic");
            if (mc)
            {
                const UsdProperty *rc = connected("inputs:roughness");
                const auto image_of = [](const std::string &s) { return s.substr(0, s.rfind('.')); };
                if (!rc || image_of(rc->value.text) != image_of(mc->value.text))
                {
                    return Fail("metallic and roughness must read one packed texture (material " + path + ")");
                }
                m.metallic_roughness_tex = Texture(*mat, mc->value.text, ok);
            }
            else
            {
                float metallic = 0.f, roughness = 0.5f;
                Number(*surface, "inputs:metallic", metallic);
                Number(*surface, "inputs:roughness", roughness);
                m.metallic = metallic;
                m.roughness = roughness;
            }
            m.eta = dielectric ? ior : 0.f;
        }
        if (!ok)
        {
            return false;
        }
        m.type = dielectric ? SPK_MAT_DIELECTRIC : SPK_MAT_PBR;
        out = scene.AddMaterial(m);
        material_ids[path] = out;
        stats.materials++;
        return true;
    }

    // LoadMesh's geometry half (USDLoader.cpp:268-347)
    bool Mesh(const UsdPrim &prim, const Mat4 &world)
    {
        const UsdProperty *binding = prim.Prop("material:binding");
        if (!binding || !binding->value.has_text)
        {
            SetLastError("USDLoader: Mesh " + prim.name + " does not have material. Skipped.");
            stats.skipped++;
            return true;
        }
        const UsdProperty *points = prim.Prop("points"), *normals = prim.Prop("normals");
        const UsdProperty *counts = prim.Prop("faceVertexCounts"), *indices = prim.Prop("faceVertexIndices");
        if (counts)
        {
            for (double c : counts->value.numbers)
            {
                if (c != 3.0)
                {
                    SetLastError("USDLoader: Mesh " + prim.name + " is not a triangle mesh. Skipped.");
                    stats.skipped++;
                    return true;
                }
            }
        }
        if (!points || !indices || points->value.numbers.empty() || points->value.numbers.size() % 3 != 0 ||
            indices->value.numbers.empty() || indices->value.numbers.size() % 3 != 0)
        {
            return Fail("Mesh " + prim.name + " needs points and triangle faceVertexIndices");
        }
        const size_t nv = points->value.numbers.size() / 3;
        if (!normals || normals->value.numbers.size() != nv * 3)
        {
            return Fail("Mesh " + prim.name + " needs per-vertex normals (the loader asserts on them)");
        }
        MeshData md;
        md.vertices.resize(nv * 3);
        md.normals.resize(nv * 3);
        for (size_t k = 0; k < nv * 3; k++)
        {
            md.vertices[k] = float(points->value.numbers[k]);
            md.normals[k] = float(normals->value.numbers[k]);
        }
        md.indices.resize(indices->value.numbers.size());
        for (size_t k = 0; k < md.indices.size(); k++)
        {
            const double v = indices->value.numbers[k];
            if (v < 0 || v >= double(nv))
            {
                return Fail("Mesh " + prim.name + ": vertex index out of range");
            }
            md.indices[k] = uint32_t(v);
        }
        const UsdProperty *st = prim.Prop("primvars:st");
        if (st && st->value.numbers.size() == nv * 2)
        {
            md.uvs.resize(nv * 2);
            for (size_t k = 0; k < nv * 2; k++)
            {
                md.uvs[k] = float(st->value.numbers[k]);
            }
        }
        const UsdProperty *tg = prim.Prop("primvars:tangents");
        md.tangents.resize(nv * 4);
        if (tg && tg->value.numbers.size() == nv * 4)
        {
            for (size_t k = 0; k < nv * 4; k++)
            {
                md.tangents[k] = float(tg->value.numbers[k]);
            }
        }
        else if (tg && tg->value.numbers.size() == nv * 3)
        {
            for (size_t i = 0; i < nv; i++)
            {
                for (int a = 0; a < 3; a++)
                {
                    md.tangents[i * 4 + size_t(a)] = float(tg->value.numbers[i * 3 + size_t(a)]);
                }
                md.tangents[i * 4 + 3] = 1.f;
            }
        }
        else
        {
            // (n x major) x n with major = GetPossibleMajorAxis(n) (Utilities.h:222-234), w = 1
            const float inv_sqrt3 = 0.577350258827209473f;
            for (size_t i = 0; i < nv; i++)
            {
                const float n[3] = {md.normals[i * 3], md.normals[i * 3 + 1], md.normals[i * 3 + 2]};
                float major[3] = {0.f, 0.f, 1.f};
                if (std::fabs(n[0]) < inv_sqrt3)
                {
                    major[0] = 1.f;
                    major[2] = 0.f;
                }
                else if (std::fabs(n[1]) < inv_sqrt3)
                {
                    major[1] = 1.f;
                    major[2] = 0.f;
                }
                const float c[3] = {n[1] * major[2] - n[2] * major[1], n[2] * major[0] - n[0] * major[2],
                                    n[0] * major[1] - n[1] * major[0]};
                md.tangents[i * 4] = c[1] * n[2] - c[2] * n[1];
                md.tangents[i * 4 + 1] = c[2] * n[0] - c[0] * n[2];
                md.tangents[i * 4 + 2] = c[0] * n[1] - c[1] * n[0];
                md.tangents[i * 4 + 3] = 1.f;
            }
        }
        uint32_t material = 0;
        if (!Material(binding->value.text, material))
        {
            return false;
        }
        float fwd[12], inv[12];
        for (int i = 0; i < 3; i++)
        {
            for (int j = 0; j < 4; j++)
            {
                fwd[4 * i + j] = world.m[i][j];
            }
        }
        if (!AffineInverse(world, inv))
        {
            return Fail("Mesh " + prim.name + ": singular node transform");
        }
        stats.triangles += md.indices.size() / 3;
        const uint32_t mesh = scene.AddMesh(std::move(md));
        scene.AddMeshPrimitiveMatrix(mesh, material, fwd, inv);
        stats.meshes++;
        return true;
    }

    // LoadCamera (USDLoader.cpp:80-127) + OrbitCameraComponent::SetupFromTransform (OrbitCameraComponent.cpp:32-56)
    void Camera(const UsdPrim &prim, const Mat4 &world)
    {
        float focal_mm = 35.f, aperture_mm = 24.f, f_stop = 0.f, focus = 0.f;
        Number(prim, "focalLength", focal_mm);
        Number(prim, "verticalAperture", aperture_mm);
        Number(prim, "fStop", f_stop);
        Number(prim, "focusDistance", focus);
        // camera nodes store the world-to-camera rotation: the world-space forward is R^T * Front, Front = +Y
        const float fx = world.m[1][0], fy = world.m[1][1], fz = world.m[1][2];
        const float len = std::sqrt(fx * fx + fy * fy + fz * fz);
        const float f[3] = {fx / len, fy / len, fz / len};
        const float yaw = std::atan2(f[0], f[1]) * 180.f / 3.14159274101257324f;
        const float pitch = -std::asin(f[2]) * 180.f / 3.14159274101257324f;
        const float radius = focus > 0.f ? focus : 25.f;
        const Vec3 center{world.m[0][3] + f[0] * radius, world.m[1][3] + f[1] * radius, world.m[2][3] + f[2] * radius};
        scene.GetCamera().SetupOrbit(center, radius, pitch, yaw, focal_mm * 0.001f, aperture_mm * 0.001f,
                                     f_stop > 0.f ? f_stop : 22.f);
        have_camera = true;
        stats.cameras++;
    }

    // LoadNode (USDLoader.cpp:498-549)
    bool Node(const UsdPrim &prim, const Mat4 &parent)
    {
        stats.nodes++;
        Mat4 local = Identity();
        const UsdProperty *op = prim.Prop("xformOp:transform");
        if (op && op->value.numbers.size() == 16)
        {
            for (int i = 0; i < 4; i++)
            {
                for (int j = 0; j < 4; j++)
                {
                    local.m[j][i] = float(op->value.numbers[size_t(i * 4 + j)]); // MatrixCast: eigen(j, i) = m[i][j]
                }
            }
        }
        const Mat4 world = Mul(parent, local);
        for (const auto &c : prim.children)
        {
            if (c->type == "Xform")
            {
                if (!Node(*c, world))
                {
                    return false;
                }
            }
            else if (c->type == "Mesh")
            {
                if (!Mesh(*c, world))
                {
                    return false;
                }
            }
            else if (c->type == "Camera")
            {
                const UsdProperty *proj = c->Prop("projection");
                if (proj && proj->value.has_text && proj->value.text != "perspective")
                {
                    return Fail("only perspective camera is supported for now. node: " + c->name);
                }
                Camera(*c, world);
            }
            else if (c->type == "DomeLight")
            {
                float color[3] = {1.f, 1.f, 1.f}, intensity = 1.f, exposure = 0.f;
                Color(*c, "inputs:color", color);
                Number(*c, "inputs:intensity", intensity);
                Number(*c, "inputs:exposure", exposure);
                const float s = intensity * std::exp2(exposure); // GetLightColor (USDLoader.cpp:455-458)
                scene.SetSky(Vec3{color[0] * s, color[1] * s, color[2] * s});
                const UsdProperty *file = c->Prop("inputs:texture:file");
                if (file && file->value.has_text && !file->value.text.empty())
                {
                    // the texture path is relative to the USD file (USDLoader.cpp:490-493); SkyLight::SetSkyMap cooks it
                    if (!scene.SetSkyMapFile(dir.empty() ? file->value.text : dir + "/" + file->value.text))
                    {
                        return false;
                    }
                }
                stats.dome_lights++;
            }
            else if (c->type == "DistantLight")
            {
                stats.distant_lights++; // DirectionalLight: not a light source of the path tracers
            }
            else
            {
                SetLastError("USDLoader: Skipped unsupported node type " + c->type + ". node: " + c->name);
            }
        }
        return true;
    }
};
} // namespace

bool LoadUsd(const std::string &path, SceneProxy &scene, UsdStats *stats)
{
    if (path.size() < 5 || path.compare(path.size() - 5, 5, ".usda") != 0)
    {
        return Fail("only ASCII .usda files are read here (binary .usdc / .usdz need tinyusdz): " + path);
    }
    std::ifstream in(path, std::ios::binary);
    if (!in)
    {
        return Fail("failed to load file " + path);
    }
    std::stringstream buffer;
    buffer << in.rdbuf();
    const std::string text = buffer.str();
    std::vector<std::unique_ptr<UsdPrim>> roots;
    std::string default_prim, error;
    UsdParser parser(text);
    if (!parser.Parse(roots, default_prim, error))
    {
        return Fail("loading file " + path + " failed. " + error);
    }
    const UsdPrim *root = nullptr;
    for (const auto &r : roots)
    {
        if ((!default_prim.empty() && r->name == default_prim) || (default_prim.empty() && !root && r->type == "Xform"))
        {
            root = r.get();
        }
    }
    if (!root)
    {
        return Fail("no default prim / root Xform in " + path);
    }
    const size_t slash = path.find_last_of('/');
    Converter conv{scene, UsdStats{}, slash == std::string::npos ? std::string() : path.substr(0, slash), roots, {}, {}, false};
    if (!conv.Node(*root, Identity()))
    {
        return false;
    }
    if (stats)
    {
        *stats = conv.stats;
    }
    return true;
}
} // namespace sparkle_cuda

// ---- C shims ------------------------------------------------------------------------------------------------------------
extern "C" int spk_host_usd_probe(const char *path, uint64_t *stats9)
{
    sparkle_cuda::SceneProxy scene;
    sparkle_cuda::UsdStats st;
    if (!path || !stats9 || !sparkle_cuda::LoadUsd(path, scene, &st))
    {
        return -1;
    }
    const uint64_t v[9] = {st.nodes, st.meshes, st.skipped, st.materials, st.textures, st.cameras, st.dome_lights,
                           st.distant_lights, st.triangles};
    memcpy(stats9, v, sizeof(v));
    return 0;
}

extern "C" SpkHostScene *spk_host_scene_load_usd(const char *path, uint32_t width, uint32_t height)
{
    auto s = std::make_unique<SpkHostScene>();
    if (!path || !sparkle_cuda::LoadUsd(path, s->scene))
    {
        return nullptr;
    }
    s->config.pipeline = sparkle_cuda::RenderConfig::Pipeline::Cuda;
    s->config.image_width = width;
    s->config.image_height = height;
    s->scene.Update(s->config);
    s->scene.EndUpdate();
    return s.release();
}

extern "C" const SpkCamera *spk_host_scene_camera(SpkHostScene *s) { return s ? &s->scene.GetCamera().Get() : nullptr; }

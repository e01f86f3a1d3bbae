// rt2_ingest.cpp — glTF 2.0 / GLB, PNG and OpenEXR readers feeding rt2_scene_desc (see rt2_ingest.h).
// Self-contained: a small JSON reader, base64, zlib (system libz) for PNG and EXR ZIP blocks.
#include "rt2_ingest.h"

#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>

namespace rt2ingest {
namespace {

// ------------------------------------------------------------------------------------------------
// files, base64
// ------------------------------------------------------------------------------------------------
bool read_file(const std::string& path, std::vector<uint8_t>& out, std::string& err) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open " + path; return false; }
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); err = "cannot size " + path; return false; }
    out.resize((size_t)n);
    size_t got = n ? std::fread(out.data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    if (got != (size_t)n) { err = "short read on " + path; return false; }
    return true;
}

std::string dir_of(const std::string& path) {
    size_t p = path.find_last_of('/');
    return p == std::string::npos ? std::string() : path.substr(0, p + 1);
}

bool base64_decode(const char* s, size_t n, std::vector<uint8_t>& out) {
    uint32_t acc = 0; int bits = 0;
    out.clear(); out.reserve(n * 3 / 4);
    for (size_t i = 0; i < n; i++) {
        const char c = s[i];
        int v;
        if (c >= 'A' && c <= 'Z') v = c - 'A';
        else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
        else if (c >= '0' && c <= '9') v = c - '0' + 52;
        else if (c == '+' || c == '-') v = 62;
        else if (c == '/' || c == '_') v = 63;
        else if (c == '=' || c == '\n' || c == '\r') continue;
        else return false;
        acc = (acc << 6) | (uint32_t)v; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
    }
    return true;
}

// "data:<mime>;base64,<payload>" or a relative file name (percent-decoding of %20 only, the common case)
bool load_uri(const std::string& uri, const std::string& base_dir, std::vector<uint8_t>& out, std::string& err) {
    if (uri.compare(0, 5, "data:") == 0) {
        size_t comma = uri.find(',');
        if (comma == std::string::npos || uri.find(";base64") == std::string::npos || uri.find(";base64") > comma) { err = "unsupported data URI"; return false; }
        if (!base64_decode(uri.data() + comma + 1, uri.size() - comma - 1, out)) { err = "bad base64 payload"; return false; }
        return true;
    }
    std::string name;
    for (size_t i = 0; i < uri.size(); i++) {
        if (uri[i] == '%' && i + 2 < uri.size() + 0 && std::isxdigit((unsigned char)uri[i + 1]) && std::isxdigit((unsigned char)uri[i + 2])) {
            name.push_back((char)std::strtol(uri.substr(i + 1, 2).c_str(), nullptr, 16)); i += 2;
        } else name.push_back(uri[i]);
    }
    return read_file(base_dir + name, out, err);
}

// ------------------------------------------------------------------------------------------------
// JSON (RFC 8259 subset sufficient for glTF: no surrogate-pair recombination, numbers as double)
// ------------------------------------------------------------------------------------------------
struct JVal {
    enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<JVal> arr;
    std::vector<std::pair<std::string, JVal>> obj;

    const JVal* get(const char* key) const {
        if (type != Obj) return nullptr;
        for (const auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool has(const char* key) const { return get(key) != nullptr; }
    double number(const char* key, double dflt) const { const JVal* v = get(key); return v && v->type == Num ? v->num : dflt; }
    long integer(const char* key, long dflt) const { const JVal* v = get(key); return v && v->type == Num ? (long)v->num : dflt; }
    std::string string(const char* key) const { const JVal* v = get(key); return v && v->type == Str ? v->str : std::string(); }
    size_t size() const { return type == Arr ? arr.size() : 0; }
};

struct JParser {
    const char* p; const char* end; std::string err; int depth = 0;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool fail(const char* m) { if (err.empty()) err = m; return false; }
    bool parse_string(std::string& out) {
        if (p >= end || *p != '"') return fail("expected string");
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return fail("bad escape");
                switch (*p) {
                    case '"': out.push_back('"'); break; case '\\': out.push_back('\\'); break; case '/': out.push_back('/'); break;
                    case 'b': out.push_back('\b'); break; case 'f': out.push_back('\f'); break; case 'n': out.push_back('\n'); break;
                    case 'r': out.push_back('\r'); break; case 't': out.push_back('\t'); break;
                    case 'u': {
                        if (end - p < 5) return fail("bad \\u escape");
                        unsigned cp = (unsigned)std::strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
                        p += 4;
                        if (cp < 0x80) out.push_back((char)cp);
                        else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
                        else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
                        break;
                    }
                    default: return fail("bad escape");
                }
                p++;
            } else out.push_back(*p++);
        }
        if (p >= end) return fail("unterminated string");
        p++;
        return true;
    }
    bool parse(JVal& v) {
        if (++depth > 128) return fail("JSON nested too deeply");
        ws();
        if (p >= end) return fail("unexpected end of JSON");
        bool ok = true;
        if (*p == '{') {
            v.type = JVal::Obj; p++; ws();
            if (p < end && *p == '}') p++;
            else for (;;) {
                ws();
                std::string key;
                if (!parse_string(key)) { ok = false; break; }
                ws();
                if (p >= end || *p != ':') { ok = fail("expected ':'"); break; }
                p++;
                v.obj.emplace_back(std::move(key), JVal());
                if (!parse(v.obj.back().second)) { ok = false; break; }
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; break; }
                ok = fail("expected ',' or '}'"); break;
            }
        } else if (*p == '[') {
            v.type = JVal::Arr; p++; ws();
            if (p < end && *p == ']') p++;
            else for (;;) {
                v.arr.emplace_back();
                if (!parse(v.arr.back())) { ok = false; break; }
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; break; }
                ok = fail("expected ',' or ']'"); break;
            }
        } else if (*p == '"') {
            v.type = JVal::Str; ok = parse_string(v.str);
        } else if (end - p >= 4 && std::strncmp(p, "true", 4) == 0) { v.type = JVal::Bool; v.b = true; p += 4; }
        else if (end - p >= 5 && std::strncmp(p, "false", 5) == 0) { v.type = JVal::Bool; v.b = false; p += 5; }
        else if (end - p >= 4 && std::strncmp(p, "null", 4) == 0) { v.type = JVal::Null; p += 4; }
        else {
            const char* q = p;
            while (q < end && (std::isdigit((unsigned char)*q) || *q == '-' || *q == '+' || *q == '.' || *q == 'e' || *q == 'E')) q++;
            if (q == p) ok = fail("unexpected character in JSON");
            else { v.type = JVal::Num; v.num = std::strtod(std::string(p, q).c_str(), nullptr); p = q; }
        }
        depth--;
        return ok;
    }
};

bool parse_json(const char* text, size_t n, JVal& root, std::string& err) {
    JParser ps{text, text + n, {}, 0};
    if (!ps.parse(root)) { err = "JSON: " + ps.err; return false; }
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// PNG (ISO/IEC 15948): 8-bit, non-interlaced
// ------------------------------------------------------------------------------------------------
namespace {
uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

bool zlib_inflate(const uint8_t* src, size_t n, std::vector<uint8_t>& dst, size_t expected, std::string& err) {
    dst.resize(expected);
    uLongf out_len = (uLongf)expected;
    int rc = uncompress(dst.data(), &out_len, src, (uLong)n);
    if (rc != Z_OK || out_len != expected) { err = "zlib: inflate failed or size mismatch"; return false; }
    return true;
}
}  // namespace

bool decode_png(const uint8_t* data, size_t size, std::vector<uint8_t>& pixels, uint32_t& w, uint32_t& h, uint32_t& channels, std::string& err) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (size < 8 || std::memcmp(data, sig, 8) != 0) { err = "not a PNG"; return false; }
    size_t pos = 8;
    uint32_t bit_depth = 0, color_type = 0, interlace = 0;
    std::vector<uint8_t> idat, palette, trns;
    bool have_ihdr = false, done = false;
    while (!done && pos + 12 <= size) {
        const uint32_t len = be32(data + pos);
        const uint8_t* type = data + pos + 4;
        if (pos + 12 + (size_t)len > size) { err = "PNG: truncated chunk"; return false; }
        const uint8_t* body = data + pos + 8;
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len != 13) { err = "PNG: bad IHDR"; return false; }
            w = be32(body); h = be32(body + 4); bit_depth = body[8]; color_type = body[9]; interlace = body[12];
            have_ihdr = true;
        } else if (!std::memcmp(type, "PLTE", 4)) palette.assign(body, body + len);
        else if (!std::memcmp(type, "tRNS", 4)) trns.assign(body, body + len);
        else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!std::memcmp(type, "IEND", 4)) done = true;
        pos += 12 + (size_t)len;
    }
    if (!have_ihdr || idat.empty() || w == 0 || h == 0) { err = "PNG: missing IHDR/IDAT"; return false; }
    if (interlace) { err = "PNG: interlaced images are not supported"; return false; }
    if (bit_depth == 16) { err = "16 bytes images"; return false; }               // wording of src/gltf/image.rs:38
    uint32_t samples;                                                              // samples per pixel in the file
    switch (color_type) {
        case 0: samples = 1; break; case 2: samples = 3; break; case 3: samples = 1; break;
        case 4: samples = 2; break; case 6: samples = 4; break;
        default: err = "PNG: bad colour type"; return false;
    }
    if (bit_depth != 8 && !((color_type == 0 || color_type == 3) && (bit_depth == 1 || bit_depth == 2 || bit_depth == 4))) { err = "PNG: unsupported bit depth"; return false; }
    if ((uint64_t)w * h > (1ull << 28)) { err = "PNG: image too large"; return false; }
    const size_t stride = ((size_t)w * samples * bit_depth + 7) / 8;
    const size_t bpp = std::max<size_t>(1, samples * bit_depth / 8);
    std::vector<uint8_t> raw;
    if (!zlib_inflate(idat.data(), idat.size(), raw, (stride + 1) * h, err)) return false;
    // unfilter in place (filter byte + stride per row)
    std::vector<uint8_t> prev(stride, 0);
    for (uint32_t y = 0; y < h; y++) {
        uint8_t* row = raw.data() + (size_t)y * (stride + 1);
        const uint8_t ft = row[0];
        uint8_t* cur = row + 1;
        for (size_t i = 0; i < stride; i++) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int pred;
            switch (ft) {
                case 0: pred = 0; break; case 1: pred = a; break; case 2: pred = b; break; case 3: pred = (a + b) >> 1; break;
                case 4: { const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c); pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
                default: err = "PNG: bad filter type"; return false;
            }
            cur[i] = (uint8_t)(cur[i] + pred);
        }
        std::memcpy(prev.data(), cur, stride);
    }
    // expand to 8-bit channels the way the `image` crate reports them: palette -> RGB(A), grey < 8 bit -> L8
    if (color_type == 3) {
        if (palette.size() < 3) { err = "PNG: palette image without PLTE"; return false; }
        channels = trns.empty() ? 3u : 4u;
    } else channels = samples;
    pixels.resize((size_t)w * h * channels);
    for (uint32_t y = 0; y < h; y++) {
        const uint8_t* cur = raw.data() + (size_t)y * (stride + 1) + 1;
        uint8_t* dst = pixels.data() + (size_t)y * w * channels;
        if (color_type == 3 || (color_type == 0 && bit_depth < 8)) {
            const uint32_t maxv = (1u << bit_depth) - 1u;
            for (uint32_t x = 0; x < w; x++) {
                uint32_t v = bit_depth == 8 ? cur[x] : (cur[(size_t)x * bit_depth / 8] >> (8 - bit_depth - (x * bit_depth) % 8)) & maxv;
                if (color_type == 3) {
                    if ((size_t)v * 3 + 2 >= palette.size()) v = 0;
                    dst[(size_t)x * channels + 0] = palette[v * 3]; dst[(size_t)x * channels + 1] = palette[v * 3 + 1]; dst[(size_t)x * channels + 2] = palette[v * 3 + 2];
                    if (channels == 4) dst[(size_t)x * 4 + 3] = v < trns.size() ? trns[v] : 255;
                } else dst[x] = (uint8_t)(v * 255u / maxv);
            }
        } else std::memcpy(dst, cur, (size_t)w * channels);
    }
    return true;
}

// ------------------------------------------------------------------------------------------------
// glTF 2.0
// ------------------------------------------------------------------------------------------------
namespace {
struct Accessor {
    const uint8_t* base = nullptr; size_t stride = 0; size_t count = 0;
    uint32_t component = 0;      // 5120 i8, 5121 u8, 5122 i16, 5123 u16, 5125 u32, 5126 f32
    uint32_t ncomp = 0;          // SCALAR 1 .. VEC4 4
    bool normalized = false;
};
size_t component_size(uint32_t c) { return c == 5120 || c == 5121 ? 1 : (c == 5122 || c == 5123 ? 2 : (c == 5125 || c == 5126 ? 4 : 0)); }

struct Doc {
    JVal root;
    std::vector<std::vector<uint8_t>> buffers;
    std::string base_dir;
};

bool get_accessor(const Doc& d, long index, Accessor& a, std::string& err) {
    const JVal* accs = d.root.get("accessors");
    if (!accs || index < 0 || (size_t)index >= accs->size()) { err = "glTF: accessor index out of range"; return false; }
    const JVal& j = accs->arr[(size_t)index];
    if (j.has("sparse")) { err = "glTF: sparse accessors are not supported"; return false; }
    a.component = (uint32_t)j.integer("componentType", 0);
    if (j.integer("count", 0) < 0 || j.integer("byteOffset", 0) < 0) { err = "glTF: negative accessor count / byteOffset"; return false; }
    a.count = (size_t)j.integer("count", 0);
    a.normalized = j.get("normalized") && j.get("normalized")->b;
    const std::string type = j.string("type");
    a.ncomp = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : 0;
    const size_t cs = component_size(a.component);
    if (!cs || !a.ncomp) { err = "glTF: unsupported accessor type " + type; return false; }
    const long bv_index = j.integer("bufferView", -1);
    const JVal* bvs = d.root.get("bufferViews");
    if (bv_index < 0 || !bvs || (size_t)bv_index >= bvs->size()) { err = "glTF: accessor without a buffer view"; return false; }
    const JVal& bv = bvs->arr[(size_t)bv_index];
    const long buf = bv.integer("buffer", -1);
    if (buf < 0 || (size_t)buf >= d.buffers.size()) { err = "glTF: buffer index out of range"; return false; }
    if (bv.integer("byteOffset", 0) < 0 || bv.integer("byteStride", 0) < 0) { err = "glTF: negative buffer view byteOffset / byteStride"; return false; }
    const size_t off = (size_t)bv.integer("byteOffset", 0) + (size_t)j.integer("byteOffset", 0);
    const size_t elem = cs * a.ncomp;
    a.stride = (size_t)bv.integer("byteStride", 0);
    if (a.stride == 0) a.stride = elem;
    const std::vector<uint8_t>& b = d.buffers[(size_t)buf];
    // overflow-safe form of off + (count-1)*stride + elem <= size
    if (a.count && (off > b.size() || elem > b.size() - off || (a.count - 1) > (b.size() - off - elem) / a.stride)) { err = "glTF: accessor reads past the end of its buffer"; return false; }
    a.base = b.data() + off;
    return true;
}

// gltf crate `into_f32` / `into_rgba_f32` conversions: normalised integers -> [0,1] (signed: max(c/MAX, -1))
float component_as_float(const Accessor& a, size_t i, uint32_t c) {
    const uint8_t* p = a.base + i * a.stride + c * component_size(a.component);
    switch (a.component) {
        case 5126: { float f; std::memcpy(&f, p, 4); return f; }
        case 5121: return a.normalized ? (float)p[0] / 255.0f : (float)p[0];
        case 5123: { uint16_t v; std::memcpy(&v, p, 2); return a.normalized ? (float)v / 65535.0f : (float)v; }
        case 5120: { int8_t v = (int8_t)p[0]; return a.normalized ? std::max((float)v / 127.0f, -1.0f) : (float)v; }
        case 5122: { int16_t v; std::memcpy(&v, p, 2); return a.normalized ? std::max((float)v / 32767.0f, -1.0f) : (float)v; }
        case 5125: { uint32_t v; std::memcpy(&v, p, 4); return (float)v; }
    }
    return 0.0f;
}
uint32_t component_as_u32(const Accessor& a, size_t i) {
    const uint8_t* p = a.base + i * a.stride;
    switch (a.component) {
        case 5121: return p[0];
        case 5123: { uint16_t v; std::memcpy(&v, p, 2); return v; }
        case 5125: { uint32_t v; std::memcpy(&v, p, 4); return v; }
    }
    return 0;
}

// gltf::scene::Transform::matrix(): `matrix` verbatim (column-major), or T * R * S
void node_matrix(const JVal& node, float m[16]) {
    const JVal* jm = node.get("matrix");
    if (jm && jm->size() == 16) { for (int i = 0; i < 16; i++) m[i] = (float)jm->arr[(size_t)i].num; return; }
    float t[3] = {0, 0, 0}, r[4] = {0, 0, 0, 1}, s[3] = {1, 1, 1};
    if (const JVal* v = node.get("translation")) if (v->size() == 3) for (int i = 0; i < 3; i++) t[i] = (float)v->arr[(size_t)i].num;
    if (const JVal* v = node.get("rotation")) if (v->size() == 4) for (int i = 0; i < 4; i++) r[i] = (float)v->arr[(size_t)i].num;
    if (const JVal* v = node.get("scale")) if (v->size() == 3) for (int i = 0; i < 3; i++) s[i] = (float)v->arr[(size_t)i].num;
    const float x = r[0], y = r[1], z = r[2], w = r[3];
    const float x2 = x + x, y2 = y + y, z2 = z + z;
    const float xx = x * x2, xy = x * y2, xz = x * z2, yy = y * y2, yz = y * z2, zz = z * z2, wx = w * x2, wy = w * y2, wz = w * z2;
    m[0] = (1 - (yy + zz)) * s[0]; m[1] = (xy + wz) * s[0]; m[2] = (xz - wy) * s[0]; m[3] = 0;
    m[4] = (xy - wz) * s[1]; m[5] = (1 - (xx + zz)) * s[1]; m[6] = (yz + wx) * s[1]; m[7] = 0;
    m[8] = (xz + wy) * s[2]; m[9] = (yz - wx) * s[2]; m[10] = (1 - (xx + yy)) * s[2]; m[11] = 0;
    m[12] = t[0]; m[13] = t[1]; m[14] = t[2]; m[15] = 1;
}

struct MeshRange { uint32_t vertex_offset, vertex_count, index_offset, index_count; rt2_geometry_info material; };

// src/gltf/material.rs:12-23 (+ the gltf crate's defaults for absent properties / an absent material)
void read_material(const Doc& d, long index, rt2_geometry_info& g) {
    g.base_color[0] = g.base_color[1] = g.base_color[2] = g.base_color[3] = 1.0f;
    g.base_color_texture_index = -1; g.metallic_factor = 1.0f; g.roughness = 1.0f;
    g.emission[0] = g.emission[1] = g.emission[2] = 0.0f; g.emission[3] = 1.0f;             // src/model.rs:404-409
    const JVal* mats = d.root.get("materials");
    if (index < 0 || !mats || (size_t)index >= mats->size()) return;
    const JVal& m = mats->arr[(size_t)index];
    if (const JVal* e = m.get("emissiveFactor")) if (e->size() == 3) for (int i = 0; i < 3; i++) g.emission[i] = (float)e->arr[(size_t)i].num;
    if (const JVal* pbr = m.get("pbrMetallicRoughness")) {
        if (const JVal* c = pbr->get("baseColorFactor")) if (c->size() == 4) for (int i = 0; i < 4; i++) g.base_color[i] = (float)c->arr[(size_t)i].num;
        g.metallic_factor = (float)pbr->number("metallicFactor", 1.0);
        g.roughness = (float)pbr->number("roughnessFactor", 1.0);
        if (const JVal* t = pbr->get("baseColorTexture")) g.base_color_texture_index = (int32_t)t->integer("index", -1);
    }
}

// src/gltf/image.rs:57-109: R8 -> (r,0,0,255), R8G8 -> (r,g,0,255), RGB -> (r,g,b,255), RGBA verbatim
void expand_to_rgba8(const std::vector<uint8_t>& px, uint32_t channels, std::vector<uint8_t>& out) {
    const size_t n = px.size() / channels;
    out.resize(n * 4);
    for (size_t i = 0; i < n; i++) {
        const uint8_t* s = px.data() + i * channels;
        uint8_t* o = out.data() + i * 4;
        o[0] = s[0]; o[1] = channels >= 2 ? s[1] : 0;
        o[2] = channels >= 3 ? s[2] : 0; o[3] = channels == 4 ? s[3] : 255;
    }
}
}  // namespace

bool load_gltf(const std::string& path, GltfModel& out, std::string& err) {
    std::vector<uint8_t> file;
    if (!read_file(path, file, err)) return false;
    Doc d;
    d.base_dir = dir_of(path);
    const uint8_t* glb_bin = nullptr; size_t glb_bin_size = 0;
    if (file.size() >= 12 && !std::memcmp(file.data(), "glTF", 4)) {                        // GLB container
        uint32_t version, length;
        std::memcpy(&version, file.data() + 4, 4); std::memcpy(&length, file.data() + 8, 4);
        if (version != 2 || length > file.size()) { err = "GLB: unsupported version or truncated file"; return false; }
        size_t pos = 12; bool have_json = false;
        while (pos + 8 <= length) {
            uint32_t clen, ctype;
            std::memcpy(&clen, file.data() + pos, 4); std::memcpy(&ctype, file.data() + pos + 4, 4);
            if (pos + 8 + (size_t)clen > length) { err = "GLB: truncated chunk"; return false; }
            if (ctype == 0x4E4F534Au && !have_json) { if (!parse_json((const char*)file.data() + pos + 8, clen, d.root, err)) return false; have_json = true; }
            else if (ctype == 0x004E4942u && !glb_bin) { glb_bin = file.data() + pos + 8; glb_bin_size = clen; }
            pos += 8 + (size_t)clen; pos = (pos + 3) & ~(size_t)3;
        }
        if (!have_json) { err = "GLB: no JSON chunk"; return false; }
    } else if (!parse_json((const char*)file.data(), file.size(), d.root, err)) return false;
    if (d.root.type != JVal::Obj) { err = "glTF: top level is not an object"; return false; }

    if (const JVal* bufs = d.root.get("buffers")) for (size_t i = 0; i < bufs->size(); i++) {
        const JVal& b = bufs->arr[i];
        d.buffers.emplace_back();
        if (b.has("uri")) { if (!load_uri(b.string("uri"), d.base_dir, d.buffers.back(), err)) return false; }
        else if (i == 0 && glb_bin) d.buffers.back().assign(glb_bin, glb_bin + glb_bin_size);
        else { err = "glTF: buffer without uri"; return false; }
        if (d.buffers.back().size() < (size_t)b.integer("byteLength", 0)) { err = "glTF: buffer shorter than byteLength"; return false; }
    }

    // ---- meshes: one entry per supported primitive (mod.rs:62-125) ----
    std::map<std::pair<size_t, size_t>, size_t> redirect;
    std::vector<MeshRange> meshes;
    const JVal* jmeshes = d.root.get("meshes");
    for (size_t mi = 0; jmeshes && mi < jmeshes->size(); mi++) {
        const JVal* prims = jmeshes->arr[mi].get("primitives");
        for (size_t pi = 0; prims && pi < prims->size(); pi++) {
            const JVal& prim = prims->arr[pi];
            const JVal* attrs = prim.get("attributes");
            if (!prim.has("indices") || !attrs || !attrs->has("POSITION") || !attrs->has("NORMAL")) continue;    // is_primitive_supported, mod.rs:170-174
            Accessor pos, nrm, col, uv, idx;
            if (!get_accessor(d, attrs->integer("POSITION", -1), pos, err) || !get_accessor(d, attrs->integer("NORMAL", -1), nrm, err) ||
                !get_accessor(d, prim.integer("indices", -1), idx, err)) return false;
            const bool has_col = attrs->has("COLOR_0"), has_uv = attrs->has("TEXCOORD_0");
            if (has_col && !get_accessor(d, attrs->integer("COLOR_0", -1), col, err)) return false;
            if (has_uv && !get_accessor(d, attrs->integer("TEXCOORD_0", -1), uv, err)) return false;
            if (pos.component != 5126 || pos.ncomp != 3 || nrm.component != 5126 || nrm.ncomp != 3) { err = "glTF: POSITION/NORMAL must be float VEC3"; return false; }
            if (nrm.count < pos.count || (has_col && col.count < pos.count) || (has_uv && uv.count < pos.count)) { err = "glTF: attribute shorter than POSITION"; return false; }
            if (idx.ncomp != 1 || (idx.component != 5121 && idx.component != 5123 && idx.component != 5125)) { err = "glTF: bad index accessor"; return false; }
            MeshRange mr{};
            mr.vertex_offset = (uint32_t)out.vertices.size(); mr.vertex_count = (uint32_t)pos.count;
            mr.index_offset = (uint32_t)out.indices.size(); mr.index_count = (uint32_t)idx.count;
            for (size_t v = 0; v < pos.count; v++) {
                rt2_vertex vx{};                                                             // w = 0 in position and normal (mod.rs:75,90)
                for (uint32_t c = 0; c < 3; c++) { vx.pos[c] = component_as_float(pos, v, c); vx.normal[c] = component_as_float(nrm, v, c); }
                if (has_col) { for (uint32_t c = 0; c < 4; c++) vx.color[c] = c < col.ncomp ? component_as_float(col, v, c) : 1.0f; }   // into_rgba_f32: alpha 1 for RGB
                else vx.color[0] = vx.color[1] = vx.color[2] = vx.color[3] = 1.0f;           // Vec4::ONE
                if (has_uv) { vx.uv[0] = component_as_float(uv, v, 0); vx.uv[1] = component_as_float(uv, v, 1); }
                out.vertices.push_back(vx);
            }
            for (size_t i = 0; i < idx.count; i++) {
                const uint32_t v = component_as_u32(idx, i);
                if (v >= pos.count) { err = "glTF: index past the end of the primitive's vertices"; return false; }
                out.indices.push_back(v);
            }
            read_material(d, prim.integer("material", -1), mr.material);
            redirect[{mi, pi}] = meshes.size();
            meshes.push_back(mr);
        }
    }
    // ---- nodes with a mesh, document order, LOCAL transform (mod.rs:127-139) ----
    const JVal* jnodes = d.root.get("nodes");
    for (size_t ni = 0; jnodes && ni < jnodes->size(); ni++) {
        const JVal& node = jnodes->arr[ni];
        const long mi = node.integer("mesh", -1);
        if (mi < 0 || !jmeshes || (size_t)mi >= jmeshes->size()) continue;
        float m[16];
        node_matrix(node, m);
        const JVal* prims = jmeshes->arr[(size_t)mi].get("primitives");
        for (size_t pi = 0; prims && pi < prims->size(); pi++) {
            auto it = redirect.find({(size_t)mi, pi});
            if (it == redirect.end()) continue;
            const MeshRange& mr = meshes[it->second];
            rt2_geometry_info g = mr.material;
            std::memcpy(g.transform, m, 64);
            g.index_offset = mr.index_offset; g.vertex_offset = mr.vertex_offset;
            out.geometries.push_back(g);
            out.index_counts.push_back(mr.index_count);
        }
    }
    if (out.geometries.empty()) { err = "glTF: no node references a supported mesh primitive"; return false; }

    // ---- images (gltf::import decodes them; src/gltf/image.rs expands to RGBA8) ----
    const JVal* jimages = d.root.get("images");
    for (size_t i = 0; jimages && i < jimages->size(); i++) {
        const JVal& im = jimages->arr[i];
        std::vector<uint8_t> bytes;
        if (im.has("uri")) { if (!load_uri(im.string("uri"), d.base_dir, bytes, err)) return false; }
        else {
            const long bvi = im.integer("bufferView", -1);
            const JVal* bvs = d.root.get("bufferViews");
            if (bvi < 0 || !bvs || (size_t)bvi >= bvs->size()) { err = "glTF: image without uri or bufferView"; return false; }
            const JVal& bv = bvs->arr[(size_t)bvi];
            const long buf = bv.integer("buffer", -1);
            const size_t off = (size_t)bv.integer("byteOffset", 0), len = (size_t)bv.integer("byteLength", 0);
            if (buf < 0 || (size_t)buf >= d.buffers.size() || off + len > d.buffers[(size_t)buf].size()) { err = "glTF: image buffer view out of range"; return false; }
            bytes.assign(d.buffers[(size_t)buf].begin() + (long)off, d.buffers[(size_t)buf].begin() + (long)(off + len));
        }
        std::vector<uint8_t> px; uint32_t w = 0, h = 0, ch = 0;
        if (bytes.size() >= 3 && bytes[0] == 0xFF && bytes[1] == 0xD8) { if (!decode_jpeg(bytes.data(), bytes.size(), px, w, h, ch, err)) return false; }
        else if (!decode_png(bytes.data(), bytes.size(), px, w, h, ch, err)) return false;
        out.image_rgba8.emplace_back();
        expand_to_rgba8(px, ch, out.image_rgba8.back());
        out.image_w.push_back(w); out.image_h.push_back(h);
    }
    // ---- samplers (mod.rs:143-155, src/gltf/texture.rs:56-70, src/model.rs:165-183) ----
    auto filter_of = [](long gl, bool) { return (gl == 9728 || gl == 9984 || gl == 9986) ? (uint32_t)RT2_FILTER_NEAREST : (uint32_t)RT2_FILTER_LINEAR; };
    out.samplers.push_back({RT2_FILTER_LINEAR, RT2_FILTER_LINEAR});
    if (const JVal* js = d.root.get("samplers")) for (size_t i = 0; i < js->size(); i++)
        out.samplers.push_back({filter_of(js->arr[i].integer("magFilter", 9729), true), filter_of(js->arr[i].integer("minFilter", 9729), false)});
    if (const JVal* jt = d.root.get("textures")) for (size_t i = 0; i < jt->size(); i++) {
        const long src = jt->arr[i].integer("source", -1), smp = jt->arr[i].integer("sampler", -1);
        if (src < 0 || (size_t)src >= out.image_rgba8.size()) { err = "glTF: texture source out of range"; return false; }
        if (smp >= 0 && (size_t)smp + 1 >= out.samplers.size()) { err = "glTF: texture sampler out of range"; return false; }
        out.textures.push_back({(uint32_t)src, smp < 0 ? 0u : (uint32_t)smp + 1u});
    }
    for (const auto& g : out.geometries)
        if (g.base_color_texture_index >= (int32_t)out.textures.size()) { err = "glTF: material texture index out of range"; return false; }
    return true;
}

// ------------------------------------------------------------------------------------------------
// OpenEXR, single-part scan-line files (openexr.com "OpenEXR File Layout")
// ------------------------------------------------------------------------------------------------
namespace {
float half_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu;
    uint32_t bits;
    if (e == 0) {
        if (m == 0) bits = sign;
        else { int s = 0; uint32_t mm = m; while (!(mm & 0x400u)) { mm <<= 1; s++; } bits = sign | ((uint32_t)(113 - s) << 23) | ((mm & 0x3FFu) << 13); }
    } else if (e == 31) bits = sign | 0x7F800000u | (m << 13);
    else bits = sign | ((e + 112u) << 23) | (m << 13);
    float f; std::memcpy(&f, &bits, 4); return f;
}

struct ExrChannel { std::string name; int type; };   // 0 UINT, 1 HALF, 2 FLOAT

// undo the ZIP/RLE block post-processing: delta predictor, then de-interleave of the two halves
void exr_unpredict_reorder(std::vector<uint8_t>& buf, std::vector<uint8_t>& tmp) {
    const size_t n = buf.size();
    for (size_t i = 1; i < n; i++) buf[i] = (uint8_t)(buf[i - 1] + buf[i] - 128);
    tmp.resize(n);
    const size_t half = (n + 1) / 2;
    for (size_t i = 0, a = 0, b = half; i < n; ) { tmp[i++] = buf[a++]; if (i < n) tmp[i++] = buf[b++]; }
    buf.swap(tmp);
}
bool exr_rle_decode(const uint8_t* src, size_t n, std::vector<uint8_t>& out, size_t expected) {
    out.clear(); out.reserve(expected);
    size_t i = 0;
    while (i < n) {
        const int c = (int8_t)src[i++];
        if (c < 0) { const size_t len = (size_t)(-c); if (i + len > n) return false; out.insert(out.end(), src + i, src + i + len); i += len; }
        else { if (i >= n) return false; out.insert(out.end(), (size_t)c + 1, src[i++]); }
        if (out.size() > expected) return false;
    }
    return out.size() == expected;
}
}  // namespace

bool load_exr(const std::string& path, std::vector<float>& rgba, uint32_t& w, uint32_t& h, std::string& err) {
    std::vector<uint8_t> f;
    if (!read_file(path, f, err)) return false;
    if (f.size() < 8 || f[0] != 0x76 || f[1] != 0x2F || f[2] != 0x31 || f[3] != 0x01) { err = "not an OpenEXR file"; return false; }
    uint32_t version; std::memcpy(&version, f.data() + 4, 4);
    if ((version & 0xFFu) != 2 || (version & 0x1A00u)) { err = "EXR: tiled, deep and multi-part files are not supported"; return false; }
    size_t pos = 8;
    std::vector<ExrChannel> channels;
    int compression = -1, line_order = 0;
    int32_t dw[4] = {0, 0, -1, -1};
    auto cstr = [&](std::string& s) { s.clear(); while (pos < f.size() && f[pos]) s.push_back((char)f[pos++]); if (pos >= f.size()) return false; pos++; return true; };
    for (;;) {
        std::string name, type;
        if (!cstr(name)) { err = "EXR: truncated header"; return false; }
        if (name.empty()) break;
        if (!cstr(type) || pos + 4 > f.size()) { err = "EXR: truncated header"; return false; }
        uint32_t size; std::memcpy(&size, f.data() + pos, 4); pos += 4;
        if (pos + size > f.size()) { err = "EXR: truncated attribute"; return false; }
        const uint8_t* v = f.data() + pos;
        if (name == "channels") {
            size_t p = 0;
            while (p < size && v[p]) {
                ExrChannel c;
                while (p < size && v[p]) c.name.push_back((char)v[p++]);
                p++;
                if (p + 16 > size) { err = "EXR: bad channel list"; return false; }
                int32_t t, xs, ys; std::memcpy(&t, v + p, 4); std::memcpy(&xs, v + p + 8, 4); std::memcpy(&ys, v + p + 12, 4);
                if (xs != 1 || ys != 1) { err = "EXR: subsampled channels are not supported"; return false; }
                c.type = t; p += 16;
                channels.push_back(c);
            }
        } else if (name == "compression" && size >= 1) compression = v[0];
        else if (name == "dataWindow" && size >= 16) std::memcpy(dw, v, 16);
        else if (name == "lineOrder" && size >= 1) line_order = v[0];
        pos += size;
    }
    (void)line_order;   // the offset table is indexed by block number in increasing-y order for both INCREASING_Y and DECREASING_Y
    if (channels.empty() || dw[2] < dw[0] || dw[3] < dw[1]) { err = "EXR: missing channels or dataWindow"; return false; }
    int lines_per_block;
    switch (compression) {
        case 0: case 1: case 2: lines_per_block = 1; break;     // NONE, RLE, ZIPS
        case 3: lines_per_block = 16; break;                    // ZIP
        case 4: lines_per_block = 32; break;                    // PIZ (the OpenEXR default for photographs)
        default: err = "EXR: compression " + std::to_string(compression) + " is not supported (NONE, RLE, ZIPS, ZIP, PIZ only; not PXR24 / B44 / DWA)"; return false;
    }
    w = (uint32_t)(dw[2] - dw[0] + 1); h = (uint32_t)(dw[3] - dw[1] + 1);
    if ((uint64_t)w * h > (1ull << 28)) { err = "EXR: image too large"; return false; }
    size_t bytes_per_line = 0;
    for (const auto& c : channels) { if (c.type < 0 || c.type > 2) { err = "EXR: bad pixel type"; return false; } bytes_per_line += (size_t)w * (c.type == 1 ? 2 : 4); }
    // destination component of each channel: R,G,B,A by name; a lone luminance channel Y feeds R, G and B
    std::vector<int> dest(channels.size(), -1);
    bool have_alpha = false, have_rgb = false;
    for (size_t i = 0; i < channels.size(); i++) {
        const std::string& n = channels[i].name;
        if (n == "R") { dest[i] = 0; have_rgb = true; } else if (n == "G") { dest[i] = 1; have_rgb = true; } else if (n == "B") { dest[i] = 2; have_rgb = true; }
        else if (n == "A") { dest[i] = 3; have_alpha = true; }
    }
    if (!have_rgb) for (size_t i = 0; i < channels.size(); i++) if (channels[i].name == "Y") dest[i] = 4;
    const uint32_t blocks = (h + (uint32_t)lines_per_block - 1) / (uint32_t)lines_per_block;
    if (pos + (size_t)blocks * 8 > f.size()) { err = "EXR: truncated offset table"; return false; }
    const size_t table = pos;
    rgba.assign((size_t)w * h * 4, 0.0f);
    if (!have_alpha) for (size_t i = 0; i < (size_t)w * h; i++) rgba[i * 4 + 3] = 1.0f;
    std::vector<uint8_t> block, tmp;
    for (uint32_t b = 0; b < blocks; b++) {
        uint64_t off; std::memcpy(&off, f.data() + table + (size_t)b * 8, 8);
        if (f.size() < 8 || off > f.size() - 8) { err = "EXR: block offset out of range"; return false; }
        int32_t y0, size; std::memcpy(&y0, f.data() + off, 4); std::memcpy(&size, f.data() + off + 4, 4);
        if (size < 0 || (uint64_t)size > f.size() - 8 - off || y0 < dw[1] || y0 > dw[3]) { err = "EXR: bad block header"; return false; }
        const uint32_t row0 = (uint32_t)(y0 - dw[1]);
        const uint32_t rows = std::min<uint32_t>((uint32_t)lines_per_block, h - row0);
        const size_t expected = bytes_per_line * rows;
        const uint8_t* src = f.data() + off + 8;
        if ((size_t)size == expected) block.assign(src, src + size);       // stored raw (also when compression did not shrink the block)
        else if (compression == 2 || compression == 3) { if (!zlib_inflate(src, (size_t)size, block, expected, err)) return false; exr_unpredict_reorder(block, tmp); }
        else if (compression == 1) { if (!exr_rle_decode(src, (size_t)size, block, expected)) { err = "EXR: bad RLE block"; return false; } exr_unpredict_reorder(block, tmp); }
        else if (compression == 4) {
            std::vector<int> words; for (const auto& c : channels) words.push_back(c.type == 1 ? 1 : 2);
            if (!exr_piz_decode(src, (size_t)size, words, w, rows, block, err)) return false;
            if (block.size() != expected) { err = "EXR: PIZ block decodes to the wrong size"; return false; }
        }
        else { err = "EXR: block size mismatch"; return false; }
        const uint8_t* p = block.data();
        for (uint32_t r = 0; r < rows; r++) {
            float* dst_row = rgba.data() + (size_t)(row0 + r) * w * 4;
            for (size_t c = 0; c < channels.size(); c++) {                  // within a scan line: channels in file (alphabetical) order, each contiguous
                const int t = channels[c].type, d = dest[c];
                for (uint32_t x = 0; x < w; x++) {
                    float v;
                    if (t == 1) { uint16_t hv; std::memcpy(&hv, p, 2); p += 2; v = half_to_float(hv); }
                    else if (t == 2) { std::memcpy(&v, p, 4); p += 4; }
                    else { uint32_t u; std::memcpy(&u, p, 4); p += 4; v = (float)u; }
                    if (d >= 0 && d < 4) dst_row[(size_t)x * 4 + d] = v;
                    else if (d == 4) { dst_row[(size_t)x * 4] = dst_row[(size_t)x * 4 + 1] = dst_row[(size_t)x * 4 + 2] = v; }
                }
            }
        }
    }
    return true;
}

}  // namespace rt2ingest

// ovis_scene.cu - scene import: the reference's on-disk formats on the input side of the
// path, read straight into a device-resident world.
//
//   .object world   JSON array of serialised objects (engine/src/json_serialiser.cc). World -> root Entity
//                   (engine/src/world.cc:56-71), Entity -> properties, components, children (engine/src/entity.cc:
//                   106-148). Only what reaches the render world is interpreted: Entity transforms and activity,
//                   MeshRenderer / Skybox (graphics/src/renderer.cc:70-85, mesh_renderer.cc:50-59, skybox.cc:46-54),
//                   the four Light classes (graphics/src/light.cc:35-119) and Camera (graphics/src/camera.cc:71-79).
//   .obj mesh       per sub-mesh local bounding boxes exactly as OBJLoader + MeshLoader::calculateBoundingBox produce
//                   them (engine/src/loaders/obj_loader.cc:94-279, mesh_loader.cc:187-199).
//
// Host code: parsing is a one-off, pointer-chasing job; the result is handed to the world through the same calls an
// application would make (ovis_node_add, ovis_entity_add, ovis_node_bind_entities, ovis_light_add), so everything
// downstream - transforms, bounds, culls - runs on the device as for any other world.
#include "orion_vis.h"

#include "ovis_math.hpp"

#include <cfloat>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <array>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

namespace ovis {
int report_error(int code, const char *message);   // ovis_world.cu: sets ovis_last_error(), returns code
}

namespace {

int scene_error(int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    return ovis::report_error(code, buf);
}

// ---- minimal JSON: what rapidjson's Document::Parse accepts with the default flags, and the values it yields -------

struct Json {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0.0;
    bool is_double = false;                                     // rapidjson's IsDouble(): written with a fraction or an exponent, or beyond 64 bits
    bool is_uint = false;                                       // rapidjson's IsUint(): a non-negative integer below 2^32 (and "-0")
    std::string str;
    std::vector<Json> items;                                    // Array
    std::vector<std::pair<std::string, Json>> members;          // Object, in file order

    const Json *get(const char *key) const {
        for (const auto &m : members)
            if (m.first == key)
                return &m.second;
        return nullptr;
    }
};

struct JsonParser {
    const char *p, *end;
    std::string error;

    void skip() {
        while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r'))
            p++;
    }
    bool fail(const char *what) {
        if (error.empty())
            error = what;
        return false;
    }
    // rapidjson's ParseStringToStream (reader.h:839-900 of the vendored copy) with the default flags: the escapes of
    // RFC 8259 only, \u with exactly four hex digits, a high surrogate must be followed by a low one and the pair becomes
    // one 4-byte sequence, a lone low surrogate is stored as it stands, bytes below 0x20 are refused, everything else is
    // copied without validation.
    bool parseHex4(unsigned &cp) {
        cp = 0;
        for (int i = 0; i < 4; i++, p++) {
            const char c = p < end ? *p : '\0';
            cp <<= 4;
            if (c >= '0' && c <= '9') cp += (unsigned)(c - '0');
            else if (c >= 'A' && c <= 'F') cp += (unsigned)(c - 'A' + 10);
            else if (c >= 'a' && c <= 'f') cp += (unsigned)(c - 'a' + 10);
            else return fail("bad \\u escape");
        }
        return true;
    }
    static void putUtf8(std::string &out, unsigned cp) {
        if (cp <= 0x7f) {
            out += (char)cp;
        } else if (cp <= 0x7ff) {
            out += (char)(0xc0 | (cp >> 6));
            out += (char)(0x80 | (cp & 0x3f));
        } else if (cp <= 0xffff) {
            out += (char)(0xe0 | (cp >> 12));
            out += (char)(0x80 | ((cp >> 6) & 0x3f));
            out += (char)(0x80 | (cp & 0x3f));
        } else {
            out += (char)(0xf0 | (cp >> 18));
            out += (char)(0x80 | ((cp >> 12) & 0x3f));
            out += (char)(0x80 | ((cp >> 6) & 0x3f));
            out += (char)(0x80 | (cp & 0x3f));
        }
    }
    bool parseString(std::string &out) {
        if (p >= end || *p != '"')
            return fail("expected string");
        p++;
        for (;;) {
            if (p >= end)
                return fail("unterminated string");
            const char c = *p;
            if (c == '"') {
                p++;
                return true;
            }
            if ((unsigned char)c < 0x20)
                return fail("control character in a string");
            if (c != '\\') {
                out += c;
                p++;
                continue;
            }
            if (++p >= end)
                return fail("bad escape");
            switch (*p++) {
            case '"': out += '"'; break;
            case '\\': out += '\\'; break;
            case '/': out += '/'; break;
            case 'n': out += '\n'; break;
            case 't': out += '\t'; break;
            case 'r': out += '\r'; break;
            case 'b': out += '\b'; break;
            case 'f': out += '\f'; break;
            case 'u': {
                unsigned cp;
                if (!parseHex4(cp))
                    return false;
                if (cp >= 0xd800 && cp <= 0xdbff) {
                    unsigned lo;
                    if (end - p < 2 || p[0] != '\\' || p[1] != 'u')
                        return fail("bad surrogate pair");
                    p += 2;
                    if (!parseHex4(lo))
                        return false;
                    if (lo < 0xdc00 || lo > 0xdfff)
                        return fail("bad surrogate pair");
                    cp = (((cp - 0xd800) << 10) | (lo - 0xdc00)) + 0x10000;
                }
                putUtf8(out, cp);
                break;
            }
            default:
                return fail("bad escape");
            }
        }
    }
    // rapidjson's ParseNumber (reader.h:1133-1370) with the default flags on a 64-bit build, then the value GetDouble()
    // returns for it (document.h:1652-1659): integers that fit stay integers (converted exactly, or rounded once from 64
    // bits), anything with a fraction or an exponent is significand * or / a power of ten in double arithmetic - NOT the
    // correctly rounded strtod - with at most 17 significant digits taken (internal/strtod.h: StrtodNormalPrecision).
    static double pow10_table(int n) {
        static const std::array<double, 309> e = [] {             // (initialised once, safely, whichever thread comes first)
            std::array<double, 309> t{};
            char lit[16];
            for (int i = 0; i <= 308; i++) {
                snprintf(lit, sizeof lit, "1e%d", i);
                t[(size_t)i] = strtod(lit, nullptr);               // what the literal 1e+i is in pow10.h's table
            }
            return t;
        }();
        return e[(size_t)n];
    }
    static double fastPath(double significand, int exp) {
        if (exp < -308)
            return 0.0;
        return exp >= 0 ? significand * pow10_table(exp) : significand / pow10_table(-exp);
    }
    bool parseNumber(double &out, bool *is_double = nullptr, bool *is_uint = nullptr) {
        const char *s = p;
        auto peek = [&]() { return s < end ? *s : '\0'; };
        auto digit = [&]() { const char c = peek(); return c >= '0' && c <= '9'; };
        const bool minus = peek() == '-';
        if (minus)
            s++;
        unsigned i = 0;
        uint64_t i64 = 0;
        bool use64 = false, useDouble = false;
        int sig = 0;
        double d = 0.0;
        if (peek() == '0') {
            s++;
        } else if (peek() >= '1' && peek() <= '9') {
            i = (unsigned)(*s++ - '0');
            const unsigned lim = minus ? 214748364u : 429496729u;
            const char last = minus ? '8' : '5';
            while (digit()) {
                if (i >= lim && (i != lim || peek() > last)) {
                    i64 = i;
                    use64 = true;
                    break;
                }
                i = i * 10 + (unsigned)(*s++ - '0');
                sig++;
            }
        } else {
            return fail("unexpected character");
        }
        if (use64) {
            const uint64_t lim = minus ? 0x0CCCCCCCCCCCCCCCull : 0x1999999999999999ull;
            const char last = minus ? '8' : '5';
            while (digit()) {
                if (i64 >= lim && (i64 != lim || peek() > last)) {
                    d = (double)i64;
                    useDouble = true;
                    break;
                }
                i64 = i64 * 10 + (unsigned)(*s++ - '0');
                sig++;
            }
        }
        if (useDouble)
            while (digit()) {
                if (d >= 1.7976931348623157e307)
                    return fail("number too big");
                d = d * 10 + (*s++ - '0');
            }
        int expFrac = 0;
        if (peek() == '.') {
            s++;
            if (!digit())
                return fail("digits must follow the decimal point");
            if (!useDouble) {
                if (!use64)
                    i64 = i;
                while (digit()) {
                    if (i64 > 0x1FFFFFFFFFFFFFull)                  // 2^53 - 1
                        break;
                    i64 = i64 * 10 + (unsigned)(*s++ - '0');
                    --expFrac;
                    if (i64 != 0)
                        sig++;
                }
                d = (double)i64;
                useDouble = true;
            }
            while (digit()) {
                if (sig < 17) {
                    d = d * 10.0 + (*s++ - '0');
                    --expFrac;
                    if (d > 0.0)
                        sig++;
                } else {
                    s++;
                }
            }
        }
        int exp = 0;
        if (peek() == 'e' || peek() == 'E') {
            s++;
            if (!useDouble) {
                d = use64 ? (double)i64 : (double)i;
                useDouble = true;
            }
            bool expMinus = false;
            if (peek() == '+')
                s++;
            else if (peek() == '-') {
                expMinus = true;
                s++;
            }
            if (!digit())
                return fail("digits must follow the exponent mark");
            exp = *s++ - '0';
            if (expMinus) {
                while (digit()) {
                    exp = exp * 10 + (*s++ - '0');
                    if (exp >= 214748364)
                        while (digit())
                            s++;
                }
                exp = -exp;
            } else {
                const int maxExp = 308 - expFrac;
                while (digit()) {
                    exp = exp * 10 + (*s++ - '0');
                    if (exp > maxExp)
                        return fail("number too big");
                }
            }
        }
        if (is_double)
            *is_double = useDouble;
        if (is_uint)
            *is_uint = !useDouble && !use64 && (!minus || i == 0);   // Int(0) carries the unsigned flags too (document.h: GenericValue(int))
        if (useDouble) {
            const int pw = exp + expFrac;
            d = pw < -308 ? fastPath(fastPath(d, -308), pw + 308) : fastPath(d, pw);
            out = minus ? -d : d;
        } else if (use64) {
            out = minus ? (double)(int64_t)(~i64 + 1) : (double)i64;
        } else {
            out = minus ? (double)(int32_t)(~i + 1) : (double)i;
        }
        p = s;
        return true;
    }
    // Document::Parse(text, length) (document.h; the stream skips the bytes of a UTF-8 byte order mark and reads a NUL byte as the end
    // of the text): one value, white space either side, nothing else.
    bool parseDocument(Json &doc) {
        for (unsigned char mark : { 0xefu, 0xbbu, 0xbfu })          // encodedstream.h:72-76: each byte of the mark on its own
            if (p < end && (unsigned char)*p == mark)
                p++;
        if (const void *nul = memchr(p, 0, (size_t)(end - p)))
            end = static_cast<const char *>(nul);
        skip();
        if (p >= end)
            return fail("the document is empty");
        if (!parse(doc))
            return false;
        skip();
        if (p != end)
            return fail("trailing characters after the document");
        return true;
    }
    bool parse(Json &v, int depth = 0) {
        if (depth > 64)
            return fail("nesting too deep");
        skip();
        if (p >= end)
            return fail("unexpected end of file");
        if (*p == '{') {
            v.kind = Json::Object;
            p++;
            skip();
            if (p < end && *p == '}') { p++; return true; }
            for (;;) {
                skip();
                std::string key;
                if (!parseString(key))
                    return false;
                skip();
                if (p >= end || *p != ':')
                    return fail("expected ':'");
                p++;
                v.members.emplace_back(std::move(key), Json());
                if (!parse(v.members.back().second, depth + 1))
                    return false;
                skip();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (*p == '[') {
            v.kind = Json::Array;
            p++;
            skip();
            if (p < end && *p == ']') { p++; return true; }
            for (;;) {
                v.items.emplace_back();
                if (!parse(v.items.back(), depth + 1))
                    return false;
                skip();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (*p == '"') {
            v.kind = Json::String;
            return parseString(v.str);
        }
        if (end - p >= 4 && !strncmp(p, "true", 4)) { v.kind = Json::Bool; v.b = true; p += 4; return true; }
        if (end - p >= 5 && !strncmp(p, "false", 5)) { v.kind = Json::Bool; v.b = false; p += 5; return true; }
        if (end - p >= 4 && !strncmp(p, "null", 4)) { v.kind = Json::Null; p += 4; return true; }
        if (!parseNumber(v.num, &v.is_double, &v.is_uint))
            return false;
        v.kind = Json::Number;
        return true;
    }
};

bool read_file(const char *path, std::string &out) {
    FILE *f = fopen(path, "rb");
    if (!f)
        return false;
    char buf[1 << 16];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0)
        out.append(buf, n);
    fclose(f);
    return true;
}

// ---- OBJ sub-mesh bounding boxes --------------------------------------------------------------------------------

// String::tokenize(line, tokens, delimiters) with the default trimEmpty = true (core/src/string.cc)
void tokenize(const std::string &s, const char *delims, bool trim_empty, std::vector<std::string> &out) {
    size_t last = 0;
    while (last <= s.size()) {
        size_t pos = s.find_first_of(delims, last);
        if (pos == std::string::npos)
            pos = s.size();
        if (pos != last || !trim_empty)
            out.emplace_back(s, last, pos - last);
        last = pos + 1;
    }
}

struct SubMeshBox {
    float box[6];
};

// The vertex data OBJLoader builds next to the boxes (obj_loader.cc:245-276): one vertex per distinct v/vt/vn triple in
// order of first use, shared by all sub-meshes of the file; a quad becomes the triangles (0, 1, 2), (2, 3, 0).
struct ObjGeometry {
    std::vector<float> positions, normals, texcoords;           // 3, 3, 2 per vertex
    std::vector<uint32_t> indices;
    std::vector<int64_t> first_index;                           // per sub-mesh; the last entry is indices.size()
};

// obj_loader.cc:94-152 (line dispatch), :158-181 (vertex elements), :187-278 (faces), mesh_loader.cc:187-199 (box).
int parse_obj(const char *path, std::vector<SubMeshBox> &out, std::string &error, ObjGeometry *geo = nullptr) {
    std::string text;
    if (!read_file(path, text)) {
        error = std::string("cannot read ") + path;
        return OVIS_ERR_INVALID;
    }
    std::vector<ovis::hm::Vec3> positions, normals;
    std::vector<float> texcoords;
    std::map<std::array<uint32_t, 3>, uint32_t> vertex_of;      // (position, texcoord, normal) -> vertex (obj_loader.cc:245)
    size_t n_texcoords = 0, n_normals = 0;
    std::string current_material = "default";
    int current = -1;                                            // index into `out`, -1 = begin a new sub-mesh at the next face
    size_t line_no = 0, start = 0;
    auto bad = [&](const char *what) {
        error = std::string(path) + ": " + std::to_string(line_no) + ": " + what;
        return OVIS_ERR_INVALID;
    };
    while (start < text.size()) {
        size_t nl = text.find('\n', start);
        if (nl == std::string::npos)
            nl = text.size();
        std::string line(text, start, nl - start);
        start = nl + 1;
        line_no++;
        std::vector<std::string> tok;
        tokenize(line, " \r", true, tok);
        if (tok.empty())
            continue;
        if (tok[0] == "v" || tok[0] == "vt" || tok[0] == "vn") {
            const size_t want = tok[0] == "vt" ? 2 : 3;
            if (tok.size() < want + 1)
                return bad("too few values");
            float val[3] = { 0, 0, 0 };
            for (size_t i = 0; i < want; i++) {
                const char *str = tok[i + 1].c_str();
                char *stop = nullptr;
                val[i] = strtof(str, &stop);                     // :170
                if (stop != str + tok[i + 1].size())
                    return bad("expected float value");
            }
            if (tok[0] == "v")
                positions.push_back({ val[0], val[1], val[2] });
            else if (tok[0] == "vt") {
                n_texcoords++;
                texcoords.push_back(val[0]);
                texcoords.push_back(val[1]);
            } else {
                n_normals++;
                normals.push_back({ val[0], val[1], val[2] });
            }
        } else if (tok[0] == "f") {
            const size_t nv = tok.size() - 1;
            if (nv != 3 && nv != 4)
                return bad("expected 3 or 4 vertices");
            if (current < 0) {                                   // :188-192
                out.push_back({ { FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX } });   // mesh_loader.cc:188
                current = (int)out.size() - 1;
                if (geo)
                    geo->first_index.push_back((int64_t)geo->indices.size());
            }
            uint32_t face[4] = { 0, 0, 0, 0 };
            for (size_t i = 0; i < nv; i++) {
                std::vector<std::string> sub;
                tokenize(tok[i + 1], "/", false, sub);
                if (sub.size() != 3)
                    return bad("expected v/vt/vn");
                size_t idx[3];
                for (int j = 0; j < 3; j++) {
                    const char *str = sub[j].c_str();
                    char *stop = nullptr;
                    uint16_t value = (uint16_t)strtoul(str, &stop, 10);       // :220 (16-bit, as the reference)
                    if (stop != str + sub[j].size())
                        return bad("expected integer value");
                    value -= 1;
                    idx[j] = value;
                }
                // :229-248 reject an index beyond its array; the position test is strict here (the reference's `>` would
                // let index == size read past the end of the array)
                if (idx[0] >= positions.size())
                    return bad("invalid position index");
                if (idx[1] > n_texcoords)
                    return bad("invalid texture coordinate index");
                if (idx[2] > n_normals)
                    return bad("invalid normal index");
                const ovis::hm::Vec3 &p = positions[idx[0]];
                float *b = out[(size_t)current].box;             // glm::min(minimum, position), glm::max(maximum, position)
                b[0] = ovis::hm::glmMin(b[0], p.x); b[1] = ovis::hm::glmMin(b[1], p.y); b[2] = ovis::hm::glmMin(b[2], p.z);
                b[3] = ovis::hm::glmMax(b[3], p.x); b[4] = ovis::hm::glmMax(b[4], p.y); b[5] = ovis::hm::glmMax(b[5], p.z);
                if (geo) {
                    if (idx[1] >= n_texcoords || idx[2] >= n_normals)
                        return bad("vertex refers to a texture coordinate or normal that was not declared");
                    auto ins = vertex_of.emplace(std::array<uint32_t, 3>{ (uint32_t)idx[0], (uint32_t)idx[1], (uint32_t)idx[2] },
                                                 (uint32_t)(geo->positions.size() / 3));
                    if (ins.second) {
                        geo->positions.insert(geo->positions.end(), { p.x, p.y, p.z });
                        const ovis::hm::Vec3 &nv3 = normals[idx[2]];
                        geo->normals.insert(geo->normals.end(), { nv3.x, nv3.y, nv3.z });
                        geo->texcoords.insert(geo->texcoords.end(), { texcoords[2 * idx[1]], texcoords[2 * idx[1] + 1] });
                    }
                    face[i] = ins.first->second;
                }
            }
            if (geo) {                                           // :268-276
                geo->indices.insert(geo->indices.end(), { face[0], face[1], face[2] });
                if (nv == 4)
                    geo->indices.insert(geo->indices.end(), { face[2], face[3], face[0] });
            }
        } else if (tok[0] == "usemtl") {
            if (tok.size() != 2)
                return bad("expected single material name");
            if (tok[1] != current_material) {                    // :123-127
                current_material = tok[1];
                current = -1;
            }
        } else if (tok[0] == "g") {
            if (tok.size() != 2)
                return bad("expected single group name");
            current = -1;                                        // :139
        }
    }
    if (geo)
        geo->first_index.push_back((int64_t)geo->indices.size());
    return OVIS_OK;
}

}  // namespace

extern "C" int ovis_obj_mesh(const char *obj_path, int64_t cap_vertices, float *out_positions, float *out_normals, float *out_texcoords,
                             int64_t cap_indices, uint32_t *out_indices, int32_t cap_submeshes, int64_t *out_first_index,
                             int64_t *out_n_vertices, int64_t *out_n_indices, int32_t *out_n_submeshes) {
    if (!obj_path || cap_vertices < 0 || cap_indices < 0 || cap_submeshes < 0 || !out_n_vertices || !out_n_indices || !out_n_submeshes)
        return scene_error(OVIS_ERR_INVALID, "ovis_obj_mesh: bad arguments");
    std::vector<SubMeshBox> boxes;
    ObjGeometry geo;
    std::string error;
    int rc = parse_obj(obj_path, boxes, error, &geo);
    if (rc != OVIS_OK)
        return scene_error(rc, "%s", error.c_str());
    const int64_t nv = (int64_t)geo.positions.size() / 3, ni = (int64_t)geo.indices.size();
    *out_n_vertices = nv;
    *out_n_indices = ni;
    *out_n_submeshes = (int32_t)boxes.size();
    if (nv > cap_vertices || ni > cap_indices || (int64_t)boxes.size() > cap_submeshes)
        return scene_error(OVIS_ERR_CAPACITY, "%s: %lld vertices, %lld indices, %zu sub-meshes", obj_path, (long long)nv, (long long)ni, boxes.size());
    if ((nv > 0 && (!out_positions || !out_normals || !out_texcoords)) || (ni > 0 && !out_indices) || !out_first_index)
        return scene_error(OVIS_ERR_INVALID, "ovis_obj_mesh: null output");
    memcpy(out_positions, geo.positions.data(), sizeof(float) * geo.positions.size());
    memcpy(out_normals, geo.normals.data(), sizeof(float) * geo.normals.size());
    memcpy(out_texcoords, geo.texcoords.data(), sizeof(float) * geo.texcoords.size());
    memcpy(out_indices, geo.indices.data(), sizeof(uint32_t) * geo.indices.size());
    for (size_t i = 0; i < geo.first_index.size(); i++)
        out_first_index[i] = geo.first_index[i];
    return OVIS_OK;
}

// MeshLoader::calculateTangents (engine/src/loaders/mesh_loader.cc:201-289, Lengyel's method): per triangle the s and t
// directions are added to its three vertices, then every vertex gets normalize(t - n * dot(n, t)) and the handedness
// sign of dot(cross(n, t), bitangent) in w. Same float operations in the same order (GLM's dot / cross / normalize).
extern "C" int ovis_mesh_tangents(int64_t n_vertices, const float *positions, const float *normals, const float *texcoords,
                                  int64_t n_indices, const uint32_t *indices, float *out_tangents) {
    if (n_vertices < 0 || n_indices < 0 || n_indices % 3 != 0 || (n_vertices > 0 && (!positions || !normals || !texcoords || !out_tangents)) ||
        (n_indices > 0 && !indices))
        return scene_error(OVIS_ERR_INVALID, "ovis_mesh_tangents: bad arguments");
    for (int64_t i = 0; i < n_indices; i++)
        if ((int64_t)indices[i] >= n_vertices)
            return scene_error(OVIS_ERR_INVALID, "ovis_mesh_tangents: index %lld is out of range", (long long)i);
    using ovis::hm::Vec3;
    std::vector<Vec3> tangents((size_t)n_vertices, Vec3{ 0.0f, 0.0f, 0.0f }), bitangents((size_t)n_vertices, Vec3{ 0.0f, 0.0f, 0.0f });
    auto P = [&](uint32_t i) { return Vec3{ positions[3 * i], positions[3 * i + 1], positions[3 * i + 2] }; };
    for (int64_t i = 0; i < n_indices; i += 3) {
        const uint32_t i0 = indices[i], i1 = indices[i + 1], i2 = indices[i + 2];
        const Vec3 p0 = P(i0), p1 = P(i1), p2 = P(i2);
        const float x1 = p1.x - p0.x, x2 = p2.x - p0.x, y1 = p1.y - p0.y, y2 = p2.y - p0.y, z1 = p1.z - p0.z, z2 = p2.z - p0.z;
        const float s1 = texcoords[2 * i1] - texcoords[2 * i0], s2 = texcoords[2 * i2] - texcoords[2 * i0];
        const float t1 = texcoords[2 * i1 + 1] - texcoords[2 * i0 + 1], t2 = texcoords[2 * i2 + 1] - texcoords[2 * i0 + 1];
        const float r = 1.0f / (s1 * t2 - s2 * t1);
        const Vec3 sdir{ (t2 * x1 - t1 * x2) * r, (t2 * y1 - t1 * y2) * r, (t2 * z1 - t1 * z2) * r };
        const Vec3 tdir{ (s1 * x2 - s2 * x1) * r, (s1 * y2 - s2 * y1) * r, (s1 * z2 - s2 * z1) * r };
        for (uint32_t v : { i0, i1, i2 }) {
            tangents[v] = tangents[v] + sdir;
            bitangents[v] = bitangents[v] + tdir;
        }
    }
    for (int64_t i = 0; i < n_vertices; i++) {
        const Vec3 n{ normals[3 * i], normals[3 * i + 1], normals[3 * i + 2] };
        const Vec3 t = tangents[(size_t)i];
        const Vec3 tangent = ovis::hm::normalize(t - n * ovis::hm::dot(n, t));
        const float handedness = (ovis::hm::dot(ovis::hm::cross(n, t), bitangents[(size_t)i]) < 0.0f) ? -1.0f : 1.0f;
        out_tangents[4 * i] = tangent.x; out_tangents[4 * i + 1] = tangent.y; out_tangents[4 * i + 2] = tangent.z;
        out_tangents[4 * i + 3] = handedness;
    }
    return OVIS_OK;
}

// TGALoader::loadData (engine/src/loaders/tga_loader.cc:60-117): uncompressed true-colour images of 24 or 32 bits, pixel
// data after the ID field and the colour map, rows in file order (the loader does not look at the origin bits); 24-bit
// pixels get alpha 255. The file stores B, G, R(, A); the output here is R, G, B, A.
extern "C" int ovis_tga_load(const char *path, int64_t capacity_bytes, uint8_t *out_rgba, int32_t *out_width, int32_t *out_height) {
    if (!path || capacity_bytes < 0 || !out_width || !out_height)
        return scene_error(OVIS_ERR_INVALID, "ovis_tga_load: bad arguments");
    std::string data;
    if (!read_file(path, data))
        return scene_error(OVIS_ERR_INVALID, "ovis_tga_load: cannot read %s", path);
    if (data.size() < 18)
        return scene_error(OVIS_ERR_INVALID, "%s: not a TGA file", path);
    const unsigned char *h = reinterpret_cast<const unsigned char *>(data.data());
    const int id_length = h[0], image_type = h[2], map_length = h[5] | (h[6] << 8), map_depth = h[7];
    const int width = h[12] | (h[13] << 8), height = h[14] | (h[15] << 8), depth = h[16];
    if (image_type != 2)
        return scene_error(OVIS_ERR_INVALID, "%s: unsupported image format (%d)", path, image_type);
    if (depth != 24 && depth != 32)
        return scene_error(OVIS_ERR_INVALID, "%s: unsupported depth (%d)", path, depth);
    *out_width = width;
    *out_height = height;
    const size_t pixels = (size_t)width * (size_t)height, bpp = (size_t)depth / 8;
    const size_t offset = 18 + (size_t)id_length + (size_t)map_length * (size_t)(map_depth / 8);
    if (offset + pixels * bpp > data.size())
        return scene_error(OVIS_ERR_INVALID, "%s: truncated", path);
    if ((int64_t)(pixels * 4) > capacity_bytes)
        return scene_error(OVIS_ERR_CAPACITY, "%s: %d x %d needs %zu bytes", path, width, height, pixels * 4);
    if (pixels > 0 && !out_rgba)
        return scene_error(OVIS_ERR_INVALID, "ovis_tga_load: null output");
    const unsigned char *src = h + offset;
    for (size_t i = 0; i < pixels; i++) {
        out_rgba[4 * i] = src[bpp * i + 2];
        out_rgba[4 * i + 1] = src[bpp * i + 1];
        out_rgba[4 * i + 2] = src[bpp * i];
        out_rgba[4 * i + 3] = bpp == 4 ? src[4 * i + 3] : 255;
    }
    return OVIS_OK;
}

// Canonical text of a parsed document (see ovis_json_canonical in orion_vis.h). Numbers are written as the reference
// consumes them: GetFloat() = (float)double, plus the integer when the double is one.
static void json_canonical(const Json &v, std::string &out) {
    char buf[64];
    switch (v.kind) {
    case Json::Null: out += 'z'; break;
    case Json::Bool: out += v.b ? 't' : 'f'; break;
    case Json::Number: {
        const float f = (float)v.num;
        uint32_t bits;
        memcpy(&bits, &f, 4);
        snprintf(buf, sizeof buf, "#%08x", bits);
        out += buf;
        if (v.num == (double)(long long)v.num && v.num > -9.0e15 && v.num < 9.0e15) {
            snprintf(buf, sizeof buf, "i%lld", (long long)v.num);
            out += buf;
        }
        break;
    }
    case Json::String:
        out += '"';
        for (unsigned char c : v.str) {
            if (c < 0x20 || c == '"' || c == '\\' || c >= 0x7f) {
                snprintf(buf, sizeof buf, "\\x%02x", c);
                out += buf;
            } else {
                out += (char)c;
            }
        }
        out += '"';
        break;
    case Json::Array:
        out += '[';
        for (const Json &i : v.items) {
            json_canonical(i, out);
            out += ',';
        }
        out += ']';
        break;
    case Json::Object:
        out += '{';
        for (const auto &m : v.members) {
            Json key;
            key.kind = Json::String;
            key.str = m.first;
            json_canonical(key, out);
            out += ':';
            json_canonical(m.second, out);
            out += ',';
        }
        out += '}';
        break;
    }
}

extern "C" int ovis_json_canonical(const char *text, int64_t length, char *out, int64_t capacity, int64_t *out_length) {
    if (!text || length < 0 || !out_length || capacity < 0 || (capacity > 0 && !out))
        return scene_error(OVIS_ERR_INVALID, "ovis_json_canonical: bad arguments");
    JsonParser parser{ text, text + length, {} };
    Json doc;
    if (!parser.parseDocument(doc))
        return scene_error(OVIS_ERR_INVALID, "ovis_json_canonical: %s", parser.error.c_str());
    std::string s;
    json_canonical(doc, s);
    *out_length = (int64_t)s.size();
    if ((int64_t)s.size() > capacity)
        return scene_error(OVIS_ERR_CAPACITY, "ovis_json_canonical: needs %lld bytes", (long long)s.size());
    memcpy(out, s.data(), s.size());
    return OVIS_OK;
}

extern "C" int ovis_obj_submesh_aabbs(const char *obj_path, int32_t capacity, float *out_aabb, int32_t *out_count) {
    if (!obj_path || capacity < 0 || (capacity > 0 && !out_aabb) || !out_count)
        return scene_error(OVIS_ERR_INVALID, "ovis_obj_submesh_aabbs: bad arguments");
    std::vector<SubMeshBox> boxes;
    std::string error;
    int rc = parse_obj(obj_path, boxes, error);
    if (rc != OVIS_OK)
        return scene_error(rc, "%s", error.c_str());
    *out_count = (int32_t)boxes.size();
    if ((int32_t)boxes.size() > capacity)
        return scene_error(OVIS_ERR_CAPACITY, "%s has %zu sub-meshes, buffer holds %d", obj_path, boxes.size(), capacity);
    for (size_t i = 0; i < boxes.size(); i++)
        std::memcpy(out_aabb + 6 * i, boxes[i].box, sizeof boxes[i].box);
    return OVIS_OK;
}

// ---- .object world ----------------------------------------------------------------------------------------------

namespace {

using ovis::hm::Quat;
using ovis::hm::Trs;
using ovis::hm::Vec3;

struct SceneObjects {
    Json root;                                                   // the document: an array of serialised objects
};

const Json *props(const Json &o) { return o.get("objectProperties"); }

bool prop_bool(const Json &o, const char *name, bool fallback) {
    const Json *p = props(o);
    const Json *v = p ? p->get(name) : nullptr;
    return (v && v->kind == Json::Bool) ? v->b : fallback;
}

// JSONSerialiser::read (json_serialiser.cc:486-626) is strict about JSON types, and a value of the wrong type leaves the
// property at its default: bool wants IsBool(); float wants IsFloat() - a number WRITTEN as a double (fraction or
// exponent: `50` is not one, `50.0` is) within the float range (rapidjson document.h:971-976); vec3 / vec4 / quat want an
// array of exactly that many numbers of any kind (GetFloat()); unsigned integers want IsUint().
float prop_float(const Json &o, const char *name, float fallback) {
    const Json *p = props(o);
    const Json *v = p ? p->get(name) : nullptr;
    if (!v || v->kind != Json::Number || !v->is_double || !(v->num >= -3.4028234e38 && v->num <= 3.4028234e38))
        return fallback;
    return (float)v->num;                                         // GetFloat(): double -> float
}

bool prop_floats(const Json &o, const char *name, int n, float *out) {
    const Json *p = props(o);
    const Json *v = p ? p->get(name) : nullptr;
    if (!v || v->kind != Json::Array || (int)v->items.size() != n)
        return false;
    for (int i = 0; i < n; i++)
        if (v->items[(size_t)i].kind != Json::Number)
            return false;
    for (int i = 0; i < n; i++)
        out[i] = (float)v->items[(size_t)i].num;
    return true;
}

bool prop_uint(const Json &o, const char *name, uint32_t *out) {
    const Json *p = props(o);
    const Json *v = p ? p->get(name) : nullptr;
    if (!v || v->kind != Json::Number || !v->is_uint)
        return false;
    *out = (uint32_t)v->num;
    return true;
}

// {"objectID": n}: n is a uint32 (json_serialiser.cc:456-458) and names the n-th object of the document (findObject,
// :232-244: `document[id]`, not a search for a member). -1: no usable reference; -2: the empty object, a null reference.
long long ref_id(const Json *ref) {
    if (!ref || ref->kind != Json::Object)
        return -1;
    if (ref->members.empty())
        return -2;
    const Json *id = ref->get("objectID");
    return (id && id->kind == Json::Number && id->is_uint) ? (long long)id->num : -1;
}

struct Importer {
    OvisWorld *world;                                             // unused by the walk
    const SceneObjects &scene;
    std::string game_dir;
    int vp_width, vp_height;
    std::map<std::string, std::vector<SubMeshBox>> meshes;       // parsed .obj files by asset path
    std::string error;

    // what the walk collects, in activation order (entity.cc:414-421: own components, then children, depth first)
    std::vector<int32_t> node_parent;
    std::vector<float> node_trs;                                  // 10 per node, local
    std::vector<Trs> node_world;
    std::vector<float> ent_aabb;                                  // 6 per render entity, local
    std::vector<uint32_t> ent_flags, ent_node;
    std::vector<std::string> ent_asset;                           // per render entity: the mesh asset ("" for a skybox)
    std::vector<int32_t> ent_submesh;                             // ... and the sub-mesh index (-1 for a skybox)
    std::vector<OvisLight> lights;
    std::vector<float> light_shading;                             // attenuation[3], bias per light
    std::vector<uint32_t> light_node;                             // the Entity (node) each light component sits on
    std::vector<OvisView> cameras;
    int shadow_map_resolution = 0;

    bool fail(const std::string &what) {
        if (error.empty())
            error = what;
        return false;
    }

    const std::vector<SubMeshBox> *mesh(const std::string &asset) {
        auto it = meshes.find(asset);
        if (it != meshes.end())
            return &it->second;
        // asset paths are "<search path>/<file without extension>"; the game's own assets live under "game/"
        if (asset.compare(0, 5, "game/") != 0) {
            fail("mesh asset '" + asset + "' is outside the game asset directory");
            return nullptr;
        }
        std::string path = game_dir + "/" + asset.substr(5) + ".obj";
        std::vector<SubMeshBox> boxes;
        std::string why;
        if (parse_obj(path.c_str(), boxes, why) != OVIS_OK) {
            fail(why);
            return nullptr;
        }
        if (boxes.empty()) {
            fail(path + " has no faces");
            return nullptr;
        }
        return &meshes.emplace(asset, std::move(boxes)).first->second;
    }

    bool component(const Json &c, uint32_t node, const Trs &world_trs) {
        if (!prop_bool(c, "active", false))                      // component.cc:29,104: inactive components never activate
            return true;
        const Json *cls = c.get("objectClass");
        const std::string name = cls ? cls->str : "";
        if (name == "MeshRenderer" || name == "Skybox") {
            const uint32_t flags = prop_bool(c, "castsShadow", true) ? OVIS_CASTS_SHADOW : 0u;   // renderer.cc:32,80
            if (name == "Skybox") {                              // skybox.cc:50-51
                const float box[6] = { -FLT_MAX, -FLT_MAX, -FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX };
                ent_aabb.insert(ent_aabb.end(), box, box + 6);
                ent_flags.push_back(flags);
                ent_node.push_back(node);
                ent_asset.emplace_back();
                ent_submesh.push_back(-1);
                return true;
            }
            const Json *p = props(c);
            const Json *m = p ? p->get("mesh") : nullptr;
            const Json *asset = m ? m->get("asset") : nullptr;
            if (!asset || asset->kind != Json::String)
                return fail("MeshRenderer without a mesh asset");
            const std::vector<SubMeshBox> *boxes = mesh(asset->str);
            if (!boxes)
                return false;
            for (const SubMeshBox &b : *boxes) {                 // one RenderEntity per sub-mesh, mesh_renderer.cc:129-134
                ent_aabb.insert(ent_aabb.end(), b.box, b.box + 6);
                ent_flags.push_back(flags);
                ent_node.push_back(node);
                ent_asset.push_back(asset->str);
                ent_submesh.push_back((int32_t)(&b - boxes->data()));
            }
            return true;
        }
        int type = -1;
        if (name == "AmbientLight") type = OVIS_LIGHT_AMBIENT;
        else if (name == "DirectionalLight") type = OVIS_LIGHT_DIRECTIONAL;
        else if (name == "PointLight") type = OVIS_LIGHT_POINT;
        else if (name == "SpotLight") type = OVIS_LIGHT_SPOT;
        if (type >= 0) {
            OvisLight l;
            std::memset(&l, 0, sizeof l);
            l.type = type;
            l.flags = prop_bool(c, "castsShadows", false) ? OVIS_CASTS_SHADOW : 0u;          // light.cc:43
            float colour[3] = { 1.0f, 1.0f, 1.0f };                                          // light.cc:39
            prop_floats(c, "colour", 3, colour);
            std::memcpy(l.colour, colour, sizeof colour);
            l.intensity = prop_float(c, "intensity", 0.8f);                                  // light.cc:40
            float att[3] = { 0.0f, 0.0f, 0.0f };
            if (type == OVIS_LIGHT_POINT) {                                                  // light.cc:58-64
                l.range = prop_float(c, "range", 100.0f);
                att[0] = 1.0f; att[1] = 0.045f; att[2] = 0.0075f;
            } else if (type == OVIS_LIGHT_SPOT) {                                            // light.cc:67-74
                l.range = prop_float(c, "range", 50.0f);
                att[0] = 1.0f; att[1] = 0.09f; att[2] = 0.032f;
                l.cutoff = prop_float(c, "cutoff", 20.0f);
                if (l.cutoff > 45.0f)
                    return fail("spot light cutoff must be <= 45 degrees");
            }
            prop_floats(c, "attenuation", 3, att);
            // Light::transformed (light.cc:100-108): absolute direction and position from the entity's world transform
            Vec3 d = ovis::hm::rotate(world_trs.q, Vec3{ 0.0f, 0.0f, -1.0f });
            l.direction[0] = d.x; l.direction[1] = d.y; l.direction[2] = d.z;
            l.position[0] = world_trs.p.x; l.position[1] = world_trs.p.y; l.position[2] = world_trs.p.z;
            lights.push_back(l);
            light_shading.insert(light_shading.end(), att, att + 3);
            light_shading.push_back(prop_float(c, "shadowBiasConstant", -0.001f));
            light_node.push_back(node);
            return true;
        }
        if (name == "Camera") {                                                              // camera.cc:46-57,71-79
            OvisView v;
            std::memset(&v, 0, sizeof v);
            v.position[0] = world_trs.p.x; v.position[1] = world_trs.p.y; v.position[2] = world_trs.p.z;
            v.orientation[0] = world_trs.q.w; v.orientation[1] = world_trs.q.x;
            v.orientation[2] = world_trs.q.y; v.orientation[3] = world_trs.q.z;
            v.hfov_deg = prop_float(c, "fov", 75.0f);
            v.znear = prop_float(c, "zNear", 0.1f);
            v.zfar = prop_float(c, "zFar", 1000.0f);
            v.vp_width = vp_width;
            v.vp_height = vp_height;
            cameras.push_back(v);
            const Json *p = props(c);
            const Json *pipe = scene_object(ref_id(p ? p->get("renderPipeline") : nullptr));
            if (pipe && !shadow_map_resolution) {                 // PROPERTY() uint16_t, deferred_render_pipeline.h:68
                uint32_t res = 0;
                if (prop_uint(*pipe, "shadowMapResolution", &res))
                    shadow_map_resolution = (int)(uint16_t)res;
            }
            return true;
        }
        return true;                                              // physics, controllers, post effects: not on the path
    }

    // findObject (json_serialiser.cc:232-244): the id is a position in the document; the object must say what class it is
    const Json *scene_object(long long id) const {
        if (id < 0 || id >= (long long)scene.root.items.size())
            return nullptr;
        const Json &o = scene.root.items[(size_t)id];
        const Json *cls = o.kind == Json::Object ? o.get("objectClass") : nullptr;
        return (cls && cls->kind == Json::String) ? &o : nullptr;
    }
    static const std::string &class_of(const Json &o) { return o.get("objectClass")->str; }
    std::vector<char> taken;                                      // per document position: already someone's child or component

    // Entity::activated order; `parent_node` < 0 for the root.
    bool entity(const Json &e, int32_t parent_node, const Trs *parent_world, int depth) {
        if (depth > 4096)
            return fail("entity tree too deep (cycle?)");
        Trs local = { { 0.0f, 0.0f, 0.0f }, { 1.0f, 0.0f, 0.0f, 0.0f }, { 1.0f, 1.0f, 1.0f } };
        if (parent_node >= 0) {                                   // the root's properties are not deserialised, entity.cc:127-130
            if (!prop_bool(e, "active", false))
                return true;                                      // inactive entity: nothing below it is active in the world
            float v[4];
            if (prop_floats(e, "position", 3, v)) local.p = { v[0], v[1], v[2] };
            if (prop_floats(e, "orientation", 4, v)) local.q = { v[0], v[1], v[2], v[3] };
            if (prop_floats(e, "scale", 3, v)) local.s = { v[0], v[1], v[2] };
        }
        const Trs world_trs = parent_world ? ovis::hm::composeTrs(*parent_world, local) : local;
        const uint32_t node = (uint32_t)node_parent.size();
        node_parent.push_back(parent_node);
        const float trs[10] = { local.p.x, local.p.y, local.p.z, local.q.w, local.q.x, local.q.y, local.q.z,
                                local.s.x, local.s.y, local.s.z };
        node_trs.insert(node_trs.end(), trs, trs + 10);
        node_world.push_back(world_trs);
        // `while (serialiser.pop(component)) addComponent(...)` (entity.cc:132-138; children :141-147): the list ends at the
        // first element that cannot be read - a reference that is not {"objectID": uint}, names no object of the document,
        // or an object of the wrong kind - and what follows it is never seen. A null reference ({}) or an object that is
        // already attached somewhere would be handed to addComponent / addChild, which assert: refused here.
        if (taken.size() != scene.root.items.size())
            taken.assign(scene.root.items.size(), 0);
        const Json *comps = e.get("components");
        if (comps && comps->kind == Json::Array)
            for (const Json &ref : comps->items) {
                const long long id = ref_id(&ref);
                if (id == -2)
                    return fail("null component reference");
                const Json *c = scene_object(id);
                if (!c || class_of(*c) == "Entity" || class_of(*c) == "World")
                    break;
                if (taken[(size_t)id]++)
                    return fail("a component is attached twice");
                if (!component(*c, node, world_trs))
                    return false;
            }
        const Json *kids = e.get("children");
        if (kids && kids->kind == Json::Array)
            for (const Json &ref : kids->items) {
                const long long id = ref_id(&ref);
                if (id == -2)
                    return fail("null child reference");
                const Json *k = scene_object(id);
                if (!k || class_of(*k) != "Entity")
                    break;
                if (taken[(size_t)id]++)
                    return fail("an entity is attached twice (or is its own ancestor)");
                if (!entity(*k, (int32_t)node, &world_trs, depth + 1))
                    return false;
            }
        return true;
    }
};

}  // namespace

struct OvisScene {
    std::vector<std::string> ent_asset;
    std::vector<int32_t> ent_submesh;
    std::string game_dir;
    std::vector<int32_t> node_parent;
    std::vector<float> node_trs, ent_aabb, light_shading;
    std::vector<uint32_t> ent_flags, ent_node, light_node;
    std::vector<OvisLight> lights;
    std::vector<OvisView> cameras;
    int shadow_map_resolution = 0;
};

static void fill_info(const OvisScene &sc, OvisSceneInfo *info) {
    std::memset(info, 0, sizeof *info);
    info->n_nodes = (int32_t)sc.node_parent.size();
    info->n_entities = (int64_t)sc.ent_flags.size();
    info->n_lights = (int32_t)sc.lights.size();
    info->n_cameras = (int32_t)sc.cameras.size();
    info->shadow_map_resolution = sc.shadow_map_resolution;
}

extern "C" {

int ovis_scene_load(const char *world_path, const char *game_asset_dir, int32_t vp_width, int32_t vp_height, OvisScene **out_scene) {
    if (!world_path || !game_asset_dir || vp_width < 1 || vp_height < 1 || !out_scene)
        return scene_error(OVIS_ERR_INVALID, "ovis_scene_load: bad arguments");
    *out_scene = nullptr;
    std::string text;
    if (!read_file(world_path, text))
        return scene_error(OVIS_ERR_INVALID, "ovis_scene_load: cannot read %s", world_path);
    SceneObjects scene;
    JsonParser parser{ text.data(), text.data() + text.size(), {} };
    if (!parser.parseDocument(scene.root) || scene.root.kind != Json::Array)
        return scene_error(OVIS_ERR_INVALID, "%s: not a serialised object file (%s)", world_path,
                           parser.error.empty() ? "top level is not an array" : parser.error.c_str());
    Importer imp{ nullptr, scene, game_asset_dir, vp_width, vp_height };
    // "The object to return is the first object in the file" (json_serialiser.cc:221): a world asset starts with its World
    const Json *world_object = imp.scene_object(0);
    if (!world_object || Importer::class_of(*world_object) != "World")
        return scene_error(OVIS_ERR_INVALID, "%s: the first object of the file is not a World", world_path);
    const long long root_id = ref_id(world_object->get("root"));                // world.cc:60-61
    const Json *root = imp.scene_object(root_id);
    if (!root || Importer::class_of(*root) != "Entity")
        return scene_error(OVIS_ERR_INVALID, "%s: the World has no root entity", world_path);
    imp.taken.assign(scene.root.items.size(), 0);
    imp.taken[(size_t)root_id] = 1;
    if (!imp.entity(*root, -1, nullptr, 0))
        return scene_error(OVIS_ERR_INVALID, "%s: %s", world_path, imp.error.c_str());
    OvisScene *sc = new OvisScene;
    sc->node_parent = std::move(imp.node_parent);
    sc->node_trs = std::move(imp.node_trs);
    sc->ent_aabb = std::move(imp.ent_aabb);
    sc->ent_flags = std::move(imp.ent_flags);
    sc->ent_node = std::move(imp.ent_node);
    sc->ent_asset = std::move(imp.ent_asset);
    sc->ent_submesh = std::move(imp.ent_submesh);
    sc->game_dir = imp.game_dir;
    sc->lights = std::move(imp.lights);
    sc->light_shading = std::move(imp.light_shading);
    sc->light_node = std::move(imp.light_node);
    sc->cameras = std::move(imp.cameras);
    sc->shadow_map_resolution = imp.shadow_map_resolution;
    *out_scene = sc;
    return OVIS_OK;
}

int ovis_scene_free(OvisScene *scene) {
    delete scene;
    return OVIS_OK;
}

int ovis_scene_entity_mesh(const OvisScene *scene, int64_t entity, char *out_obj_path, int32_t capacity, int32_t *out_submesh) {
    if (!scene || entity < 0 || entity >= (int64_t)scene->ent_asset.size() || capacity < 0 || (capacity > 0 && !out_obj_path) || !out_submesh)
        return scene_error(OVIS_ERR_INVALID, "ovis_scene_entity_mesh: bad arguments");
    const std::string &asset = scene->ent_asset[(size_t)entity];
    *out_submesh = scene->ent_submesh[(size_t)entity];
    const std::string path = asset.empty() ? std::string() : scene->game_dir + "/" + asset.substr(5) + ".obj";
    if ((int64_t)path.size() + 1 > capacity)
        return scene_error(OVIS_ERR_CAPACITY, "ovis_scene_entity_mesh: the path needs %zu bytes", path.size() + 1);
    std::memcpy(out_obj_path, path.c_str(), path.size() + 1);
    return OVIS_OK;
}

int ovis_scene_info(const OvisScene *scene, OvisSceneInfo *out_info) {
    if (!scene || !out_info)
        return scene_error(OVIS_ERR_INVALID, "ovis_scene_info: null argument");
    fill_info(*scene, out_info);
    return OVIS_OK;
}

int ovis_scene_arrays(const OvisScene *scene, int32_t *out_parent, float *out_local_trs, float *out_local_aabb, uint32_t *out_flags,
                      uint32_t *out_entity_node, OvisLight *out_lights, float *out_light_shading, OvisView *out_cameras) {
    if (!scene)
        return scene_error(OVIS_ERR_INVALID, "ovis_scene_arrays: null scene");
    auto copy = [](void *dst, const void *src, size_t bytes) {
        if (dst && bytes)
            std::memcpy(dst, src, bytes);
    };
    copy(out_parent, scene->node_parent.data(), scene->node_parent.size() * 4);
    copy(out_local_trs, scene->node_trs.data(), scene->node_trs.size() * 4);
    copy(out_local_aabb, scene->ent_aabb.data(), scene->ent_aabb.size() * 4);
    copy(out_flags, scene->ent_flags.data(), scene->ent_flags.size() * 4);
    copy(out_entity_node, scene->ent_node.data(), scene->ent_node.size() * 4);
    copy(out_lights, scene->lights.data(), scene->lights.size() * sizeof(OvisLight));
    copy(out_light_shading, scene->light_shading.data(), scene->light_shading.size() * 4);
    copy(out_cameras, scene->cameras.data(), scene->cameras.size() * sizeof(OvisView));
    return OVIS_OK;
}

int ovis_scene_instantiate(const OvisScene *scene, OvisWorld *world, OvisSceneInfo *out_info) {
    if (!scene || !world)
        return scene_error(OVIS_ERR_INVALID, "ovis_scene_instantiate: null argument");
    OvisSceneInfo info;
    fill_info(*scene, &info);
    int64_t have_entities = 0, have_nodes = 0;
    int32_t have_lights = 0;
    int rc = ovis_world_counts(world, &have_entities, &have_nodes, &have_lights);
    if (rc != OVIS_OK)
        return rc;
    std::vector<int32_t> parent = scene->node_parent;             // recorded as indices into the scene: make them world ids
    std::vector<uint32_t> ent_node = scene->ent_node;
    if (have_nodes > 0) {
        for (int32_t &p : parent)
            if (p >= 0)
                p += (int32_t)have_nodes;
        for (uint32_t &n : ent_node)
            n += (uint32_t)have_nodes;
    }
    rc = ovis_node_add(world, info.n_nodes, parent.data(), scene->node_trs.data(), &info.first_node);
    if (rc != OVIS_OK)
        return rc;
    if (info.n_entities > 0) {
        rc = ovis_entity_add(world, info.n_entities, scene->ent_aabb.data(), scene->ent_flags.data(), &info.first_entity);
        if (rc != OVIS_OK)
            return rc;
        std::vector<uint32_t> ids((size_t)info.n_entities);
        for (size_t i = 0; i < ids.size(); i++)
            ids[i] = info.first_entity + (uint32_t)i;
        rc = ovis_node_bind_entities(world, info.n_entities, ent_node.data(), ids.data());
        if (rc != OVIS_OK)
            return rc;
    }
    if (info.n_lights > 0) {
        rc = ovis_light_add(world, info.n_lights, scene->lights.data(), &info.first_light);
        if (rc != OVIS_OK)
            return rc;
        for (int32_t i = 0; i < info.n_lights; i++) {
            rc = ovis_light_set_shading(world, info.first_light + (uint32_t)i, &scene->light_shading[4 * (size_t)i],
                                        scene->light_shading[4 * (size_t)i + 3]);
            if (rc != OVIS_OK)
                return rc;
            // Light::transformed (graphics/src/light.cc:100-108): the light goes where its Entity goes
            rc = ovis_light_bind_node(world, info.first_light + (uint32_t)i, (int32_t)(info.first_node + scene->light_node[(size_t)i]));
            if (rc != OVIS_OK)
                return rc;
        }
    }
    if (out_info)
        *out_info = info;
    return OVIS_OK;
}

int ovis_scene_import(OvisWorld *world, const char *world_path, const char *game_asset_dir, int32_t vp_width,
                      int32_t vp_height, int32_t cap_cameras, OvisView *out_cameras, OvisSceneInfo *out_info) {
    if (!world || cap_cameras < 0 || (cap_cameras > 0 && !out_cameras) || !out_info)
        return scene_error(OVIS_ERR_INVALID, "ovis_scene_import: bad arguments");
    OvisScene *scene = nullptr;
    int rc = ovis_scene_load(world_path, game_asset_dir, vp_width, vp_height, &scene);
    if (rc != OVIS_OK)
        return rc;
    std::unique_ptr<OvisScene> owner(scene);
    rc = ovis_scene_instantiate(scene, world, out_info);
    if (rc != OVIS_OK)
        return rc;
    for (int32_t i = 0; i < out_info->n_cameras && i < cap_cameras; i++)
        out_cameras[i] = scene->cameras[(size_t)i];
    if (out_info->n_cameras > cap_cameras)
        return scene_error(OVIS_ERR_CAPACITY, "%s has %d cameras, buffer holds %d", world_path, out_info->n_cameras, cap_cameras);
    return OVIS_OK;
}

}  // extern "C"

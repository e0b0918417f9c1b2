/* json_harness.cc - TEST INFRASTRUCTURE: the JSON layer of the scene loader pinned by the reference's own dependency.
 *
 * The reference parses its serialised worlds with the rapidjson it vendors (engine/3rdparty/rapidjson, used by
 * engine/src/json_serialiser.cc:213-214: Document::Parse with the default flags) and reads numbers with GetFloat() /
 * GetInt() (:495-550). This file parses a document with that same library, the same call, and writes the canonical text
 * ovis_json_canonical defines; tests/test_scene_import.py requires the two to be equal byte for byte on the reference's
 * own worlds, the committed fixture and a set of awkward documents. */
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "rapidjson/document.h"

static void canonical(const rapidjson::Value &v, std::string &out) {
    char buf[64];
    if (v.IsNull()) {
        out += 'z';
    } else if (v.IsBool()) {
        out += v.GetBool() ? 't' : 'f';
    } else if (v.IsNumber()) {
        const double d = v.GetDouble();
        const float f = v.GetFloat();
        uint32_t bits;
        memcpy(&bits, &f, 4);
        snprintf(buf, sizeof buf, "#%08x", bits);
        out += buf;
        if (d == (double)(long long)d && d > -9.0e15 && d < 9.0e15) {
            snprintf(buf, sizeof buf, "i%lld", (long long)d);
            out += buf;
        }
    } else if (v.IsString()) {
        out += '"';
        const char *s = v.GetString();
        for (rapidjson::SizeType i = 0; i < v.GetStringLength(); i++) {
            const unsigned char c = (unsigned char)s[i];
            if (c < 0x20 || c == '"' || c == '\\' || c >= 0x7f) {
                snprintf(buf, sizeof buf, "\\x%02x", c);
                out += buf;
            } else {
                out += (char)c;
            }
        }
        out += '"';
    } else if (v.IsArray()) {
        out += '[';
        for (rapidjson::SizeType i = 0; i < v.Size(); i++) {
            canonical(v[i], out);
            out += ',';
        }
        out += ']';
    } else if (v.IsObject()) {
        out += '{';
        for (auto it = v.MemberBegin(); it != v.MemberEnd(); ++it) {
            canonical(it->name, out);
            out += ':';
            canonical(it->value, out);
            out += ',';
        }
        out += '}';
    }
}

/* Returns the length of the canonical text (written if it fits), or -1 when rapidjson rejects the document. */
extern "C" int64_t oref_json_canonical(const char *text, int64_t length, char *out, int64_t capacity) {
    rapidjson::Document document;
    document.Parse(text, (size_t)length);
    if (document.HasParseError())
        return -1;
    std::string s;
    canonical(document, s);
    if ((int64_t)s.size() <= capacity)
        memcpy(out, s.data(), s.size());
    return (int64_t)s.size();
}

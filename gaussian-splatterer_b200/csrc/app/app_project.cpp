// Project + settings.json (reference src/Project.h:6-73 with nlohmann's NLOHMANN_DEFINE_TYPE_INTRUSIVE_WITH_DEFAULT,
// written / read by UiFrame::saveSettings / loadSettings, src/ui/UiFrame.cpp:323-331,358-370).
//
// nlohmann/json is a dependency of the reference (lib/json), not of this library: what it does for this one struct is
// restated here -- a small JSON reader, the WITH_DEFAULT from_json rules (absent key -> Project{} default, arithmetic
// fields accept integer / float / boolean JSON values, bool fields only booleans, strings only strings, anything else is
// type_error.302; a non-object where an object is needed is type_error.306) and dump(): compact, keys in std::map order,
// floats widened to double and printed as the shortest round-trip decimal in nlohmann's layout (exponent form below
// 1e-4 and from 1e15, ".0" appended to integral values).  Pinned by tests/golden/app_project.json, which
// tools/make_golden_app.cpp generates with the reference's own headers.
#include "app_common.h"

#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>

namespace gsbapp {

Project::Project() {
    d.sphere1 = {16, 10.0f, 60.0f, 0.0f, 0.0f};
    d.sphere2 = {16, 10.0f, 60.0f, 0.0f, 0.0f};
    d.rtSamples = 100;
    d.lrLocation = 0.00005f; d.lrSh = 0.0001f; d.lrScale = 0.00002f; d.lrOpacity = 0.0001f; d.lrRotation = 0.000025f;
    d.paramScaleMax = 0.3f;
    d.paramCullOpacity = 0.005f; d.paramCullSize = 0.004f; d.paramDensifyVariance = 2.0f; d.paramSplitSize = 0.04f;
    d.paramSplitDistance = 1.5f; d.paramSplitScale = 0.8f; d.paramCloneDistance = 1.6f;
    d.iterations = 0; d.intervalCapture = 50; d.intervalDensify = 200;
    d.previewTimer = 0.0f; d.previewRtSamples = 50; d.previewSplatScale = 1.0f;
    d.previewTruth = 0; d.previewTruthIndex = 0;
    d.previewFreeOrbit = 1; d.previewFreeOrbitSpeed = 0.5f; d.previewFreeDistance = 10.0f; d.previewFreeFovDeg = 60.0f;
    d.previewFreeRotX = 25.0f; d.previewFreeRotY = 0.0f;
    d.renderResX = 2048; d.renderResY = 2048;
}

namespace {

// ------------------------------------------------------------------------------------------------------- JSON value
struct JValue;
using JObject = std::map<std::string, JValue>;
struct JValue {
    enum Kind { Null, Bool, Int, UInt, Float, String, Array, Object } kind = Null;
    bool b = false;
    int64_t i = 0;
    uint64_t u = 0;
    double f = 0.0;
    std::string s;
    std::vector<JValue> a;
    std::shared_ptr<JObject> o;
    const char* type_name() const {                    // nlohmann basic_json::type_name()
        switch (kind) {
            case Null: return "null";
            case Bool: return "boolean";
            case String: return "string";
            case Array: return "array";
            case Object: return "object";
            default: return "number";
        }
    }
};

struct Parser {
    const char* p;
    const char* begin;
    const char* end;
    [[noreturn]] void error(const std::string& what) const {
        // line / column like nlohmann's parse_error.101 (column counts characters read on the current line)
        int line = 1, col = 0;
        for (const char* q = begin; q < p && q < end; q++) {
            if (*q == '\n') { line++; col = 0; } else col++;
        }
        std::ostringstream os;
        os << "[json.exception.parse_error.101] parse error at line " << line << ", column " << col << ": " << what;
        throw std::runtime_error(os.str());
    }
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool lit(const char* word) {
        const size_t n = strlen(word);
        if ((size_t)(end - p) >= n && memcmp(p, word, n) == 0) { p += n; return true; }
        return false;
    }
    static void utf8(std::string& out, uint32_t cp) {
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
        else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
        else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    }
    uint32_t hex4() {
        if (end - p < 4) error("syntax error while parsing value - invalid string: '\\u' must be followed by 4 hex digits");
        uint32_t v = 0;
        for (int k = 0; k < 4; k++) {
            const char c = *p++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= (uint32_t)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (uint32_t)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (uint32_t)(c - 'A' + 10);
            else error("syntax error while parsing value - invalid string: '\\u' must be followed by 4 hex digits");
        }
        return v;
    }
    std::string string() {
        std::string out;
        p++;                                            // opening quote
        while (true) {
            if (p >= end) error("syntax error while parsing value - invalid string: missing closing quote");
            const unsigned char c = (unsigned char)*p++;
            if (c == '"') return out;
            if (c < 0x20) error("syntax error while parsing value - invalid string: control character must be escaped");
            if (c != '\\') { out += (char)c; continue; }
            if (p >= end) error("syntax error while parsing value - invalid string: missing closing quote");
            const char e = *p++;
            switch (e) {
                case '"': out += '"'; break;
                case '\\': out += '\\'; break;
                case '/': out += '/'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'n': out += '\n'; break;
                case 'r': out += '\r'; break;
                case 't': out += '\t'; break;
                case 'u': {
                    uint32_t cp = hex4();
                    if (cp >= 0xD800 && cp <= 0xDBFF) {
                        if (end - p >= 2 && p[0] == '\\' && p[1] == 'u') {
                            p += 2;
                            const uint32_t lo = hex4();
                            if (lo < 0xDC00 || lo > 0xDFFF) error("syntax error while parsing value - invalid string: surrogate U+D800..U+DBFF must be followed by U+DC00..U+DFFF");
                            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                        } else error("syntax error while parsing value - invalid string: surrogate U+D800..U+DBFF must be followed by U+DC00..U+DFFF");
                    }
                    utf8(out, cp);
                    break;
                }
                default: error("syntax error while parsing value - invalid string: forbidden character after backslash");
            }
        }
    }
    JValue number() {
        const char* s = p;
        bool is_float = false;
        if (p < end && *p == '-') p++;
        if (p >= end || *p < '0' || *p > '9') error("syntax error while parsing value - invalid number; expected digit after '-'");
        if (*p == '0') p++; else while (p < end && *p >= '0' && *p <= '9') p++;
        if (p < end && *p == '.') {
            is_float = true; p++;
            if (p >= end || *p < '0' || *p > '9') error("syntax error while parsing value - invalid number; expected digit after '.'");
            while (p < end && *p >= '0' && *p <= '9') p++;
        }
        if (p < end && (*p == 'e' || *p == 'E')) {
            is_float = true; p++;
            if (p < end && (*p == '+' || *p == '-')) p++;
            if (p >= end || *p < '0' || *p > '9') error("syntax error while parsing value - invalid number; expected '+', '-', or digit after exponent");
            while (p < end && *p >= '0' && *p <= '9') p++;
        }
        JValue v;
        const std::string txt(s, p);
        if (!is_float) {                               // nlohmann lexer: unsigned, then signed, else falls back to double
            if (txt[0] != '-') {
                uint64_t u = 0;
                auto r = std::from_chars(txt.data(), txt.data() + txt.size(), u);
                if (r.ec == std::errc() && r.ptr == txt.data() + txt.size()) { v.kind = JValue::UInt; v.u = u; return v; }
            } else {
                int64_t i = 0;
                auto r = std::from_chars(txt.data(), txt.data() + txt.size(), i);
                if (r.ec == std::errc() && r.ptr == txt.data() + txt.size()) { v.kind = JValue::Int; v.i = i; return v; }
            }
        }
        v.kind = JValue::Float;
        v.f = std::strtod(txt.c_str(), nullptr);
        return v;
    }
    JValue value() {
        ws();
        if (p >= end) error("syntax error while parsing value - unexpected end of input; expected '[', '{', or a literal");
        JValue v;
        const char c = *p;
        if (c == '{') {
            p++;
            v.kind = JValue::Object;
            v.o = std::make_shared<JObject>();
            ws();
            if (p < end && *p == '}') { p++; return v; }
            while (true) {
                ws();
                if (p >= end || *p != '"') {
                    if (p < end) p++;                   // nlohmann reports the column after the offending token
                    error(std::string("syntax error while parsing object key - unexpected ") +
                          (p > end || p == begin ? "end of input" : std::string("'") + *(p - 1) + "'") + "; expected string literal");
                }
                std::string key = string();
                ws();
                if (p >= end || *p != ':') { if (p < end) p++; error("syntax error while parsing object separator - expected ':'"); }
                p++;
                (*v.o)[key] = value();                  // later duplicates overwrite, like nlohmann
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; return v; }
                if (p < end) p++;
                error("syntax error while parsing object - expected '}'");
            }
        }
        if (c == '[') {
            p++;
            v.kind = JValue::Array;
            ws();
            if (p < end && *p == ']') { p++; return v; }
            while (true) {
                v.a.push_back(value());
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; return v; }
                if (p < end) p++;
                error("syntax error while parsing array - expected ']'");
            }
        }
        if (c == '"') { v.kind = JValue::String; v.s = string(); return v; }
        if (lit("true")) { v.kind = JValue::Bool; v.b = true; return v; }
        if (lit("false")) { v.kind = JValue::Bool; v.b = false; return v; }
        if (lit("null")) { v.kind = JValue::Null; return v; }
        if (c == '-' || (c >= '0' && c <= '9')) return number();
        p++;
        error("syntax error while parsing value - invalid literal");
    }
};

JValue parse_json(const char* text) {
    Parser ps{text, text, text + strlen(text)};
    JValue v = ps.value();
    ps.ws();
    if (ps.p != ps.end) { ps.p++; ps.error("syntax error while parsing value - unexpected trailing characters; expected end of input"); }
    return v;
}

// ------------------------------------------------------------------------------------------------------------- dump
// nlohmann serializer::dump_float -> to_chars -> format_buffer: shortest round-trip digits of the double, then
//   k <= n <= 15 : digits, zeros, ".0"      0 < n <= 15 : dig.its      -4 < n <= 0 : 0.000digits      else d[.igits]e[+-]XX
void dump_double(std::string& out, double x) {
    if (!std::isfinite(x)) { out += "null"; return; }
    if (std::signbit(x)) { out += '-'; x = -x; }
    if (x == 0.0) { out += "0.0"; return; }
    char buf[64];
    auto r = std::to_chars(buf, buf + sizeof(buf), x, std::chars_format::scientific);     // d[.ddd]e[+-]XX, shortest
    std::string sci(buf, r.ptr);
    const size_t epos = sci.find('e');
    std::string digits;
    for (size_t k = 0; k < epos; k++) if (sci[k] != '.') digits += sci[k];
    const int exp10 = std::atoi(sci.c_str() + epos + 1);
    const int k = (int)digits.size();
    const int n = exp10 + 1;                              // value = 0.digits * 10^n
    if (k <= n && n <= 15) {
        out += digits; out.append((size_t)(n - k), '0'); out += ".0";
    } else if (0 < n && n <= 15) {
        out += digits.substr(0, (size_t)n); out += '.'; out += digits.substr((size_t)n);
    } else if (-4 < n && n <= 0) {
        out += "0."; out.append((size_t)(-n), '0'); out += digits;
    } else {
        out += digits[0];
        if (k > 1) { out += '.'; out += digits.substr(1); }
        out += 'e';
        int e = n - 1;
        out += e < 0 ? '-' : '+';
        if (e < 0) e = -e;
        char eb[16];
        snprintf(eb, sizeof(eb), "%02d", e);
        out += eb;
    }
}

void dump_string(std::string& out, const std::string& s) {          // ensure_ascii = false: UTF-8 passes through
    out += '"';
    for (const unsigned char c : s) {
        switch (c) {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\b': out += "\\b"; break;
            case '\f': out += "\\f"; break;
            case '\n': out += "\\n"; break;
            case '\r': out += "\\r"; break;
            case '\t': out += "\\t"; break;
            default:
                if (c < 0x20) { char b[8]; snprintf(b, sizeof(b), "\\u%04x", c); out += b; }
                else out += (char)c;
        }
    }
    out += '"';
}

// ------------------------------------------------------------------------------------------- field table (PJ:64-73)
enum FieldType { F_INT, F_FLOAT, F_BOOL, F_STRING, F_SPHERE };
struct Field { const char* name; FieldType type; size_t offset; };
#define GSB_F(name, type) {#name, type, offsetof(gsb_project_data, name)}
const Field kFields[] = {          // alphabetical = std::map order of the dumped object
    GSB_F(intervalCapture, F_INT), GSB_F(intervalDensify, F_INT), GSB_F(iterations, F_INT),
    GSB_F(lrLocation, F_FLOAT), GSB_F(lrOpacity, F_FLOAT), GSB_F(lrRotation, F_FLOAT), GSB_F(lrScale, F_FLOAT), GSB_F(lrSh, F_FLOAT),
    GSB_F(paramCloneDistance, F_FLOAT), GSB_F(paramCullOpacity, F_FLOAT), GSB_F(paramCullSize, F_FLOAT),
    GSB_F(paramDensifyVariance, F_FLOAT), GSB_F(paramScaleMax, F_FLOAT), GSB_F(paramSplitDistance, F_FLOAT),
    GSB_F(paramSplitScale, F_FLOAT), GSB_F(paramSplitSize, F_FLOAT),
    {"pathModel", F_STRING, 1}, {"pathTextureDiffuse", F_STRING, 2}, {"perspective", F_STRING, 0},
    GSB_F(previewFreeDistance, F_FLOAT), GSB_F(previewFreeFovDeg, F_FLOAT), GSB_F(previewFreeOrbit, F_BOOL),
    GSB_F(previewFreeOrbitSpeed, F_FLOAT), GSB_F(previewFreeRotX, F_FLOAT), GSB_F(previewFreeRotY, F_FLOAT),
    GSB_F(previewRtSamples, F_INT), GSB_F(previewSplatScale, F_FLOAT), GSB_F(previewTimer, F_FLOAT),
    GSB_F(previewTruth, F_BOOL), GSB_F(previewTruthIndex, F_INT), GSB_F(renderResX, F_INT), GSB_F(renderResY, F_INT),
    GSB_F(rtSamples, F_INT), GSB_F(sphere1, F_SPHERE), GSB_F(sphere2, F_SPHERE),
};
#undef GSB_F
#define GSB_S(name, type) {#name, type, offsetof(gsb_camera_sphere, name)}
const Field kSphereFields[] = {GSB_S(count, F_INT), GSB_S(distance, F_FLOAT), GSB_S(fovDeg, F_FLOAT), GSB_S(rotX, F_FLOAT),
                               GSB_S(rotY, F_FLOAT)};
#undef GSB_S

std::string* string_field(Project& p, size_t which) {
    return which == 0 ? &p.perspective : which == 1 ? &p.pathModel : &p.pathTextureDiffuse;
}

void dump_scalar(std::string& out, const char* base, const Field& f) {
    if (f.type == F_INT) out += std::to_string(*reinterpret_cast<const int*>(base + f.offset));
    else if (f.type == F_BOOL) out += *reinterpret_cast<const int*>(base + f.offset) ? "true" : "false";
    else dump_double(out, (double)*reinterpret_cast<const float*>(base + f.offset));
}

std::string dump_project(Project& p) {
    std::string out = "{";
    bool first = true;
    for (const Field& f : kFields) {
        if (!first) out += ',';
        first = false;
        dump_string(out, f.name);
        out += ':';
        if (f.type == F_STRING) dump_string(out, *string_field(p, f.offset));
        else if (f.type == F_SPHERE) {
            out += '{';
            const char* base = reinterpret_cast<const char*>(&p.d) + f.offset;
            bool sf = true;
            for (const Field& g : kSphereFields) {
                if (!sf) out += ',';
                sf = false;
                dump_string(out, g.name);
                out += ':';
                dump_scalar(out, base, g);
            }
            out += '}';
        } else dump_scalar(out, reinterpret_cast<const char*>(&p.d), f);
    }
    out += '}';
    return out;
}

// --------------------------------------------------------------------------------------------------------- from_json
[[noreturn]] void type_error(int id, const std::string& what) {
    throw std::runtime_error("[json.exception.type_error." + std::to_string(id) + "] " + what);
}

// j.value(key, default): only objects have value() (type_error.306 otherwise)
const JValue* lookup(const JValue& j, const char* key) {
    if (j.kind != JValue::Object) type_error(306, std::string("cannot use value() with ") + j.type_name());
    auto it = j.o->find(key);
    return it == j.o->end() ? nullptr : &it->second;
}

template <typename T>
T arithmetic(const JValue& v) {          // nlohmann from_json for arithmetic types other than the json's own: + boolean
    switch (v.kind) {
        case JValue::UInt: return static_cast<T>(v.u);
        case JValue::Int: return static_cast<T>(v.i);
        case JValue::Float: return static_cast<T>(v.f);
        case JValue::Bool: return static_cast<T>(v.b);
        default: type_error(302, std::string("type must be number, but is ") + v.type_name());
    }
}

void read_scalar(const JValue& obj, char* base, const Field& f) {
    const JValue* v = lookup(obj, f.name);
    if (!v) return;                                  // destination already holds the Project{} default
    if (f.type == F_INT) *reinterpret_cast<int*>(base + f.offset) = arithmetic<int>(*v);
    else if (f.type == F_FLOAT) *reinterpret_cast<float*>(base + f.offset) = arithmetic<float>(*v);
    else {
        if (v->kind != JValue::Bool) type_error(302, std::string("type must be boolean, but is ") + v->type_name());
        *reinterpret_cast<int*>(base + f.offset) = v->b ? 1 : 0;
    }
}

void project_from_json(const JValue& j, Project& dst) {
    Project p;                                       // WITH_DEFAULT: start from Project{}, not from the current values
    // the macro reads the members in declaration order; errors surface for the first offending member in THAT order
    static const char* kDeclOrder[] = {
        "perspective", "pathModel", "pathTextureDiffuse", "sphere1", "sphere2", "rtSamples", "lrLocation", "lrSh", "lrScale",
        "lrOpacity", "lrRotation", "paramScaleMax", "paramCullOpacity", "paramCullSize", "paramDensifyVariance", "paramSplitSize",
        "paramSplitDistance", "paramSplitScale", "paramCloneDistance", "iterations", "intervalCapture", "intervalDensify",
        "previewTimer", "previewRtSamples", "previewSplatScale", "previewTruth", "previewTruthIndex", "previewFreeOrbit",
        "previewFreeOrbitSpeed", "previewFreeDistance", "previewFreeFovDeg", "previewFreeRotX", "previewFreeRotY", "renderResX",
        "renderResY"};
    for (const char* name : kDeclOrder) {
        const Field* f = nullptr;
        for (const Field& c : kFields) if (strcmp(c.name, name) == 0) f = &c;
        if (f->type == F_STRING) {
            const JValue* v = lookup(j, f->name);
            if (!v) continue;
            if (v->kind != JValue::String) type_error(302, std::string("type must be string, but is ") + v->type_name());
            *string_field(p, f->offset) = v->s;
        } else if (f->type == F_SPHERE) {
            const JValue* v = lookup(j, f->name);
            if (!v) continue;
            gsb_camera_sphere s = {16, 10.0f, 60.0f, 0.0f, 0.0f};       // CameraSphere{} defaults for ITS absent keys
            for (const Field& g : kSphereFields) read_scalar(*v, reinterpret_cast<char*>(&s), g);
            *reinterpret_cast<gsb_camera_sphere*>(reinterpret_cast<char*>(&p.d) + f->offset) = s;
        } else read_scalar(j, reinterpret_cast<char*>(&p.d), *f);
    }
    dst = p;
}

}  // namespace
}  // namespace gsbapp

// =============================================================================================================== C ABI
using namespace gsbapp;

extern "C" {

int gsb_project_create(gsb_project** out) {
    if (!out) return fail("gsb_project_create: invalid argument");
    *out = new gsb_project();
    return GSB_OK;
}

int gsb_project_init_new(gsb_project* p) {
    if (!p) return fail("gsb_project_init_new: invalid argument");
    p->p = Project();
    p->p.d.sphere2.count = 0;
    p->p.d.sphere2.fovDeg = 30.0f;
    return GSB_OK;
}

int gsb_project_destroy(gsb_project* p) { delete p; return GSB_OK; }

gsb_project_data* gsb_project_fields(gsb_project* p) { return p ? &p->p.d : nullptr; }

static std::string* named_string(gsb_project* p, const char* name) {
    if (!p || !name) return nullptr;
    if (!strcmp(name, "perspective")) return &p->p.perspective;
    if (!strcmp(name, "pathModel")) return &p->p.pathModel;
    if (!strcmp(name, "pathTextureDiffuse")) return &p->p.pathTextureDiffuse;
    return nullptr;
}

const char* gsb_project_get_string(gsb_project* p, const char* name) {
    std::string* s = named_string(p, name);
    if (!s) { fail("gsb_project_get_string: unknown field"); return nullptr; }
    return s->c_str();
}

int gsb_project_set_string(gsb_project* p, const char* name, const char* value) {
    std::string* s = named_string(p, name);
    if (!s || !value) return fail("gsb_project_set_string: unknown field");
    *s = value;
    return GSB_OK;
}

int gsb_project_to_json(gsb_project* p, char* buf, size_t cap, size_t* needed) {
    if (!p) return fail("gsb_project_to_json: invalid argument");
    const std::string s = dump_project(p->p);
    if (needed) *needed = s.size() + 1;
    if (buf && cap > 0) {
        const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
        memcpy(buf, s.data(), n);
        buf[n] = 0;
    }
    return GSB_OK;
}

int gsb_project_from_json(gsb_project* p, const char* text) {
    if (!p || !text) return fail("gsb_project_from_json: invalid argument");
    try {
        project_from_json(parse_json(text), p->p);
    } catch (const std::exception& e) {
        return fail(e.what());
    }
    return GSB_OK;
}

int gsb_project_save(gsb_project* p, const char* path) {
    if (!p || !path) return fail("gsb_project_save: invalid argument");
    std::ofstream file(path);
    if (!file) return fail(std::string("Failed to write settings file at \"") + path + "\"!");
    file << dump_project(p->p);
    return GSB_OK;
}

int gsb_project_load(gsb_project* p, const char* path) {
    if (!p || !path) return fail("gsb_project_load: invalid argument");
    std::ifstream file(path, std::ios::binary);
    if (!file) return fail(std::string("Failed to load settings file at \"") + path + "\"!");      // UF:360 dialog text
    std::stringstream ss;
    ss << file.rdbuf();
    return gsb_project_from_json(p, ss.str().c_str());
}

int gsb_project_train_params(const gsb_project_data* d, gsb_project_params* q) {
    if (!d || !q) return fail("gsb_project_train_params: invalid argument");
    q->lrLocation = d->lrLocation; q->lrSh = d->lrSh; q->lrScale = d->lrScale; q->lrOpacity = d->lrOpacity;
    q->lrRotation = d->lrRotation; q->paramScaleMax = d->paramScaleMax; q->paramCullOpacity = d->paramCullOpacity;
    q->paramCullSize = d->paramCullSize; q->paramDensifyVariance = d->paramDensifyVariance; q->paramSplitSize = d->paramSplitSize;
    q->paramSplitDistance = d->paramSplitDistance; q->paramSplitScale = d->paramSplitScale;
    q->paramCloneDistance = d->paramCloneDistance;
    return GSB_OK;
}

}  // extern "C"

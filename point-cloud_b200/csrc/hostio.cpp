// hostio.cpp — host-side formats of the conversion path (no CUDA):
//   metadata.json   point-converter/src/metadata.rs:9-64 (serde_json 1.0.114 pretty printer, ryu floats)
//   node files      point-converter/src/cell.rs:155-229, 280-335 (reader side, for merge mode)
//   path naming     point-converter/src/cell.rs:20-31, metadata.rs:47-49
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "builder.h"

namespace pcb {

// ---- f32 -> text exactly as serde_json writes it (ryu `format32`, "pretty" layout) -----------------
// Shortest digits that round-trip; positional for 1e-5 <= |v| < 1e13 with at least one fractional
// digit, scientific (d.ddde±x without '+') otherwise; non-finite -> null (serde_json).
std::string format_f32(float v) {
    if (std::isnan(v) || std::isinf(v)) return "null";
    std::string out;
    if (std::signbit(v)) out += '-';
    float a = std::fabs(v);
    if (a == 0.0f) return out + "0.0";
    char sci[48];
    auto r = std::to_chars(sci, sci + sizeof sci - 1, a, std::chars_format::scientific);
    *r.ptr = 0;
    // sci = d[.ddd]e[+-]XX
    std::string digits;
    const char* p = sci;
    for (; p < r.ptr && *p != 'e'; ++p)
        if (*p != '.') digits += *p;
    int e10 = std::atoi(p + 1);
    const int len = (int)digits.size();
    const int k = e10 - (len - 1);   // a = digits * 10^k
    const int kk = len + k;          // decimal point sits after kk digits
    if (k >= 0 && kk <= 13) {
        out += digits + std::string((size_t)k, '0') + ".0";
    } else if (kk > 0 && kk <= 13) {
        out += digits.substr(0, (size_t)kk) + "." + digits.substr((size_t)kk);
    } else if (kk > -6 && kk <= 0) {
        out += "0." + std::string((size_t)-kk, '0') + digits;
    } else {
        out += digits[0];
        if (len > 1) out += "." + digits.substr(1);
        out += "e" + std::to_string(kk - 1);
    }
    return out;
}

static void put_json_string(std::string& o, const std::string& s) {
    o += '"';
    for (unsigned char c : s) {
        if (c == '"') o += "\\\"";
        else if (c == '\\') o += "\\\\";
        else if (c == '\n') o += "\\n";
        else if (c == '\r') o += "\\r";
        else if (c == '\t') o += "\\t";
        else if (c == '\b') o += "\\b";
        else if (c == '\f') o += "\\f";
        else if (c < 0x20) {
            char b[8];
            std::snprintf(b, sizeof b, "\\u%04x", c);
            o += b;
        } else o += (char)c;
    }
    o += '"';
}

// serde_json::to_writer_pretty: two-space indent, ": " after keys, one array element per line, declared
// field order (metadata.rs:9-28; Aabb{min,max} bounding-volume/src/lib.rs:4-8; config metadata.rs:67-78).
std::string metadata_to_json(const Metadata& m) {
    std::string o = "{\n  \"version\": ";
    put_json_string(o, m.version);
    o += ",\n  \"name\": ";
    put_json_string(o, m.name);
    o += ",\n  \"number_of_points\": " + std::to_string(m.number_of_points);
    o += ",\n  \"hierarchies\": " + std::to_string(m.hierarchies);
    o += ",\n  \"bounding_box\": {";
    const float* vecs[2] = {m.bbox_min, m.bbox_max};
    const char* names[2] = {"min", "max"};
    for (int i = 0; i < 2; i++) {
        o += std::string(i ? "," : "") + "\n    \"" + names[i] + "\": [";
        for (int a = 0; a < 3; a++) o += std::string(a ? "," : "") + "\n      " + format_f32(vecs[i][a]);
        o += "\n    ]";
    }
    o += "\n  },\n  \"config\": {";
    o += "\n    \"cell_point_overflow_limit\": " + std::to_string(m.cell_point_overflow_limit);
    o += ",\n    \"sub_grid_dimension\": " + std::to_string(m.sub_grid_dimension);
    o += ",\n    \"max_cell_size\": " + format_f32(m.max_cell_size);
    o += "\n  }\n}";
    return o;
}

// ---- a small JSON value parser (serde_json::from_reader accepts any key order, ignores unknown keys) --
namespace {
struct JVal {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    double num = 0;
    std::string text;  // Str payload, or the raw token of a Num
    std::vector<JVal> items;
    std::vector<std::pair<std::string, JVal>> fields;
    const JVal* get(const char* key) const {
        for (auto& f : fields)
            if (f.first == key) return &f.second;
        return nullptr;
    }
};
struct JParser {
    const char* s;
    const char* end;
    void skip() {
        while (s < end && (*s == ' ' || *s == '\t' || *s == '\n' || *s == '\r')) ++s;
    }
    bool lit(const char* w) {
        size_t n = std::strlen(w);
        if ((size_t)(end - s) >= n && std::memcmp(s, w, n) == 0) { s += n; return true; }
        return false;
    }
    bool string(std::string* out) {
        if (s >= end || *s != '"') return false;
        ++s;
        while (s < end && *s != '"') {
            if (*s == '\\') {
                if (++s >= end) return false;
                switch (*s) {
                    case 'n': *out += '\n'; break;
                    case 't': *out += '\t'; break;
                    case 'r': *out += '\r'; break;
                    case 'b': *out += '\b'; break;
                    case 'f': *out += '\f'; break;
                    case 'u': {
                        if (end - s < 5) return false;
                        unsigned cp = (unsigned)std::strtoul(std::string(s + 1, s + 5).c_str(), nullptr, 16);
                        s += 4;
                        if (cp < 0x80) *out += (char)cp;
                        else if (cp < 0x800) { *out += (char)(0xC0 | (cp >> 6)); *out += (char)(0x80 | (cp & 63)); }
                        else { *out += (char)(0xE0 | (cp >> 12)); *out += (char)(0x80 | ((cp >> 6) & 63)); *out += (char)(0x80 | (cp & 63)); }
                        break;
                    }
                    default: *out += *s;
                }
                ++s;
            } else *out += *s++;
        }
        if (s >= end) return false;
        ++s;
        return true;
    }
    bool value(JVal* v, int depth = 0) {
        if (depth > 64) return false;
        skip();
        if (s >= end) return false;
        if (*s == '{') {
            v->kind = JVal::Obj;
            ++s;
            skip();
            if (s < end && *s == '}') { ++s; return true; }
            for (;;) {
                skip();
                std::string key;
                if (!string(&key)) return false;
                skip();
                if (s >= end || *s != ':') return false;
                ++s;
                JVal child;
                if (!value(&child, depth + 1)) return false;
                v->fields.emplace_back(std::move(key), std::move(child));
                skip();
                if (s < end && *s == ',') { ++s; continue; }
                if (s < end && *s == '}') { ++s; return true; }
                return false;
            }
        }
        if (*s == '[') {
            v->kind = JVal::Arr;
            ++s;
            skip();
            if (s < end && *s == ']') { ++s; return true; }
            for (;;) {
                JVal child;
                if (!value(&child, depth + 1)) return false;
                v->items.push_back(std::move(child));
                skip();
                if (s < end && *s == ',') { ++s; continue; }
                if (s < end && *s == ']') { ++s; return true; }
                return false;
            }
        }
        if (*s == '"') { v->kind = JVal::Str; return string(&v->text); }
        if (lit("true")) { v->kind = JVal::Bool; v->num = 1; return true; }
        if (lit("false")) { v->kind = JVal::Bool; return true; }
        if (lit("null")) { v->kind = JVal::Null; return true; }
        const char* b = s;
        while (s < end && (std::isdigit((unsigned char)*s) || *s == '-' || *s == '+' || *s == '.' || *s == 'e' || *s == 'E')) ++s;
        if (b == s) return false;
        v->kind = JVal::Num;
        v->text.assign(b, s);
        v->num = std::strtod(v->text.c_str(), nullptr);
        return true;
    }
};
bool as_u64(const JVal* v, uint64_t* out) {
    if (!v || v->kind != JVal::Num || v->text.find_first_of(".eE-") != std::string::npos) return false;
    *out = std::strtoull(v->text.c_str(), nullptr, 10);
    return true;
}
// serde parses a JSON number into f64 and casts to f32 for f32 fields
bool as_f32(const JVal* v, float* out) {
    if (!v || v->kind != JVal::Num) return false;
    *out = (float)v->num;
    return true;
}
bool as_vec3(const JVal* v, float out[3]) {
    if (!v || v->kind != JVal::Arr || v->items.size() != 3) return false;
    return as_f32(&v->items[0], out) && as_f32(&v->items[1], out + 1) && as_f32(&v->items[2], out + 2);
}
}  // namespace

bool metadata_from_json(const std::string& text, Metadata* m) {
    JParser p{text.data(), text.data() + text.size()};
    JVal root;
    if (!p.value(&root) || root.kind != JVal::Obj) return false;
    p.skip();
    if (p.s != p.end) return false;
    const JVal *ver = root.get("version"), *name = root.get("name"), *bb = root.get("bounding_box"),
               *cfg = root.get("config");
    if (!ver || ver->kind != JVal::Str || !name || name->kind != JVal::Str) return false;
    if (!bb || bb->kind != JVal::Obj || !cfg || cfg->kind != JVal::Obj) return false;
    uint64_t np = 0, hier = 0, lim = 0, dim = 0;
    if (!as_u64(root.get("number_of_points"), &np) || !as_u64(root.get("hierarchies"), &hier)) return false;
    if (!as_vec3(bb->get("min"), m->bbox_min) || !as_vec3(bb->get("max"), m->bbox_max)) return false;
    if (!as_u64(cfg->get("cell_point_overflow_limit"), &lim) || !as_u64(cfg->get("sub_grid_dimension"), &dim) ||
        !as_f32(cfg->get("max_cell_size"), &m->max_cell_size))
        return false;
    if (hier > 0xFFFFFFFFull || lim > 0xFFFFFFFFull || dim > 0xFFFFFFFFull) return false;
    m->version = ver->text;
    m->name = name->text;
    m->number_of_points = np;
    m->hierarchies = (uint32_t)hier;
    m->cell_point_overflow_limit = (uint32_t)lim;
    m->sub_grid_dimension = (uint32_t)dim;
    return true;
}

// ---- node files --------------------------------------------------------------------------------
std::string node_rel_path(uint32_t hierarchy, const int32_t idx[3]) {
    char buf[96];
    std::snprintf(buf, sizeof buf, "h_%u/c_%d_%d_%d.bin", hierarchy, idx[0], idx[1], idx[2]);
    return buf;
}

static inline uint32_t rd_u32(const uint8_t* p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

// Cell::read_from (cell.rs:183-229): header, number_of_points grid records, u8 entry count, entries.
bool parse_node(const uint8_t* data, size_t len, ExistingNode* out) {
    if (len < 49) return false;
    const uint32_t ng = rd_u32(data + 20);
    size_t off = 48;
    if ((len - off) / 16 < ng) return false;
    out->grid.resize(ng);
    if (ng) std::memcpy(out->grid.data(), data + off, 16ull * ng);
    off += 16ull * ng;
    if (off >= len) return false;
    const unsigned nent = data[off++];
    out->entries.clear();
    for (unsigned e = 0; e < nent; e++) {
        if (len - off < 16) return false;
        ExistingNode::Entry en;
        for (int a = 0; a < 3; a++) en.child[a] = (int32_t)rd_u32(data + off + 4 * a);
        const uint32_t cnt = rd_u32(data + off + 12);
        off += 16;
        en.none = (cnt == 0);  // cell.rs:210-211
        if (cnt) {
            if ((len - off) / 16 < cnt) return false;
            en.list.resize(cnt);
            std::memcpy(en.list.data(), data + off, 16ull * cnt);
            off += 16ull * cnt;
        }
        out->entries.push_back(std::move(en));
    }
    return true;
}

}  // namespace pcb

// pc_oracle_conv.cpp — TEST INFRASTRUCTURE ONLY (see pc_oracle.hpp).
// Converter driver (converter.rs:72-246), metadata.json codec (metadata.rs:9-64)
// and the C ABI that tests / bench.py bind with ctypes.
#include <sys/stat.h>

#include <algorithm>
#include <cerrno>
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <thread>

#include "pc_oracle.hpp"

namespace orc {

static bool read_file(const std::string& path, std::vector<uint8_t>* out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out->resize((size_t)n);
    size_t got = n ? std::fread(out->data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == (size_t)n;
}
static bool write_file(const std::string& path, const void* data, size_t n) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    size_t put = n ? std::fwrite(data, 1, n, f) : 0;
    std::fclose(f);
    return put == n;
}

// converter.rs:96-104 + bounding-volume/src/lib.rs:23-52.  f32::min/max on
// finite values; first batch of an empty cloud REPLACES the box.
void Converter::update_bounding_box(const std::vector<Point>& points) {
    if (points.empty()) return;  // Aabb::from(empty) == None
    Aabb aabb;
    aabb.min = aabb.max = Vec3{points[0].x, points[0].y, points[0].z};
    for (size_t i = 1; i < points.size(); i++) {
        const Point& p = points[i];
        aabb.min.x = std::fmin(aabb.min.x, p.x);
        aabb.min.y = std::fmin(aabb.min.y, p.y);
        aabb.min.z = std::fmin(aabb.min.z, p.z);
        aabb.max.x = std::fmax(aabb.max.x, p.x);
        aabb.max.y = std::fmax(aabb.max.y, p.y);
        aabb.max.z = std::fmax(aabb.max.z, p.z);
    }
    if (metadata.number_of_points == 0) {
        metadata.bounding_box = aabb;
    } else {
        Aabb& b = metadata.bounding_box;
        b.min.x = std::fmin(b.min.x, aabb.min.x);
        b.min.y = std::fmin(b.min.y, aabb.min.y);
        b.min.z = std::fmin(b.min.z, aabb.min.z);
        b.max.x = std::fmax(b.max.x, aabb.max.x);
        b.max.y = std::fmax(b.max.y, aabb.max.y);
        b.max.z = std::fmax(b.max.z, aabb.max.z);
    }
}

// converter.rs:106-112
void Converter::add_points_batch(const std::vector<Point>& points) {
    update_bounding_box(points);
    metadata.number_of_points += points.size();
    PointMap grouped = group_points(points, 0, metadata.config);
    MetadataConfig config = metadata.config;
    add_points_in_hierarchy(0, config, std::move(grouped));
}

// converter.rs:114-139 (recursion unrolled into a loop; same visiting order per level)
void Converter::add_points_in_hierarchy(uint32_t hierarchy, const MetadataConfig& config, PointMap&& grouped_in) {
    PointMap grouped = std::move(grouped_in);
    for (;;) {
        if (hierarchy >= 32) {  // 2u32.pow(32) overflows in the reference (metadata.rs:92)
            error = 1;
            return;
        }
        create_hierarchy_folder(hierarchy);
        PointMap next;
        for (auto& kv : grouped) {
            CellId id{hierarchy, kv.first};
            Cell& cell = get_cell_mut(id);
            arrivals += kv.second.size();
            const uint32_t points_before = cell.header.total_number_of_points;   // src/plugins/converter.rs:566
            PointMap remaining = add_points_to_cell(config, kv.second, cell);
            points_by_delta += (int32_t)cell.header.total_number_of_points - (int32_t)points_before;   // :571-573
            // merge_point_maps, converter.rs:49-60
            for (auto& r : remaining) {
                auto it = next.find(r.first);
                if (it != next.end())
                    it->second.insert(it->second.end(), r.second.begin(), r.second.end());
                else
                    next.emplace(r.first, std::move(r.second));
            }
        }
        if (next.empty()) return;
        grouped = std::move(next);
        hierarchy += 1;
    }
}

// converter.rs:141-158
void Converter::create_hierarchy_folder(uint32_t hierarchy) {
    if (metadata.hierarchies <= hierarchy) {
        metadata.hierarchies += 1;
        if (!working_directory.empty()) {
            char buf[32];
            std::snprintf(buf, sizeof buf, "/h_%u", hierarchy);
            ::mkdir((working_directory + buf).c_str(), 0777);
        }
    }
}

// converter.rs:160-207 — cache miss: read the node file, NotFound => Cell::new.
Cell& Converter::get_cell_mut(const CellId& id) {
    auto it = cells.find(id);
    if (it != cells.end()) {
        if (lru_capacity) {   // LRUCache::get_mut refreshes the entry (converter.rs:176-178)
            auto pos = lru_pos.find(id);
            lru_order.splice(lru_order.begin(), lru_order, pos->second);
        }
        return *it->second;
    }
    auto cell = std::make_unique<Cell>();
    bool loaded = false;
    if (!working_directory.empty()) {
        std::vector<uint8_t> bytes;
        if (read_file(working_directory + "/" + cell_path(id), &bytes)) {
            loaded = Cell::read_from(bytes.data(), bytes.size(), cell.get());
            lru_loads += loaded ? 1 : 0;
        }
    }
    if (lru_capacity) {
        if (!loaded) cell->points_grid.reserve(50000);   // Cell::new(.., 50_000), converter.rs:199
        if (cells.size() >= lru_capacity) {   // PutResult::Evicted => save_cell (converter.rs:166-173)
            const CellId victim = lru_order.back();
            lru_order.pop_back();
            lru_pos.erase(victim);
            auto vit = cells.find(victim);
            std::vector<uint8_t> buf;
            vit->second->write_to(buf);
            if (!write_file(working_directory + "/" + cell_path(victim), buf.data(), buf.size())) error |= 2;
            cells.erase(vit);
            lru_evictions++;
        }
        lru_order.push_front(id);
        lru_pos[id] = lru_order.begin();
    }
    if (!loaded) {
        float cell_size = metadata.config.cell_size(id.hierarchy);
        cell->header.id = id;
        cell->header.sub_cell_size = metadata.config.sub_cell_size(cell_size);
        cell->header.size = cell_size;
        cell->header.pos = metadata.config.cell_pos(id.index, cell_size);
        cell->points_grid.clear();
        cell->overflow.clear();
    }
    Cell& ref = *cell;
    cells.emplace(id, std::move(cell));
    return ref;
}

// converter.rs:218-246 — Drop: save_cache + save_metadata.
bool Converter::save_all() {
    if (working_directory.empty()) return true;
    std::vector<uint8_t> buf;
    for (auto& kv : cells) {
        buf.clear();
        kv.second->write_to(buf);
        if (!write_file(working_directory + "/" + cell_path(kv.first), buf.data(), buf.size())) return false;
    }
    std::string js = metadata_to_json(metadata);
    return write_file(working_directory + "/metadata.json", js.data(), js.size());
}

// ---------------------------------------------------------------------------
// serde_json 1.0.114 float text = ryu 1.0.17 `format32` (pretty): shortest
// round-trip digits; 1e-5 <= |v| < 1e13 in positional notation with at least
// one fractional digit, otherwise d[.ddd]e[-]x.
std::string format_f32(float v) {
    if (!(v == v) || std::isinf(v)) return "null";  // serde_json writes null for non-finite
    uint32_t bits;
    std::memcpy(&bits, &v, 4);
    std::string out;
    if (bits >> 31) out.push_back('-');
    if ((bits & 0x7FFFFFFFu) == 0) return out + "0.0";
    char tmp[64];
    auto res = std::to_chars(tmp, tmp + sizeof tmp, std::fabs(v), std::chars_format::scientific);
    std::string s(tmp, res.ptr);  // d[.ddd]e[+-]xx, shortest round trip
    size_t epos = s.find('e');
    std::string digits;
    for (size_t i = 0; i < epos; i++)
        if (s[i] != '.') digits.push_back(s[i]);
    int exp10 = std::atoi(s.c_str() + epos + 1);
    int length = (int)digits.size();
    int k = exp10 - (length - 1);  // value = digits * 10^k
    int kk = length + k;           // 10^(kk-1) <= v < 10^kk
    if (0 <= k && kk <= 13) {      // 1234e7 -> 12340000000.0
        out += digits;
        out.append((size_t)k, '0');
        out += ".0";
    } else if (0 < kk && kk <= 13) {  // 1234e-2 -> 12.34
        out += digits.substr(0, (size_t)kk);
        out.push_back('.');
        out += digits.substr((size_t)kk);
    } else if (-6 < kk && kk <= 0) {  // 1234e-6 -> 0.001234
        out += "0.";
        out.append((size_t)(-kk), '0');
        out += digits;
    } else if (length == 1) {  // 1e30
        out += digits;
        out += "e" + std::to_string(kk - 1);
    } else {  // 1234e30 -> 1.234e33
        out.push_back(digits[0]);
        out.push_back('.');
        out += digits.substr(1);
        out += "e" + std::to_string(kk - 1);
    }
    return out;
}

static std::string json_escape(const std::string& s) {
    std::string o = "\"";
    for (char c : s) {
        switch (c) {
            case '"': o += "\\\""; break;
            case '\\': o += "\\\\"; break;
            case '\n': o += "\\n"; break;
            case '\r': o += "\\r"; break;
            case '\t': o += "\\t"; break;
            default:
                if ((unsigned char)c < 0x20) {
                    char b[8];
                    std::snprintf(b, sizeof b, "\\u%04x", c);
                    o += b;
                } else
                    o.push_back(c);
        }
    }
    return o + "\"";
}

// metadata.rs:51-53 — serde_json::to_writer_pretty: 2-space indent, one array
// element per line, field order as declared (metadata.rs:9-28, 67-78; Aabb min,max).
std::string metadata_to_json(const Metadata& m) {
    std::ostringstream o;
    auto vec = [&](const char* name, const Vec3& v, bool comma) {
        o << "    \"" << name << "\": [\n      " << format_f32(v.x) << ",\n      " << format_f32(v.y) << ",\n      "
          << format_f32(v.z) << "\n    ]" << (comma ? "," : "") << "\n";
    };
    o << "{\n";
    o << "  \"version\": " << json_escape(m.version) << ",\n";
    o << "  \"name\": " << json_escape(m.name) << ",\n";
    o << "  \"number_of_points\": " << m.number_of_points << ",\n";
    o << "  \"hierarchies\": " << m.hierarchies << ",\n";
    o << "  \"bounding_box\": {\n";
    vec("min", m.bounding_box.min, true);
    vec("max", m.bounding_box.max, false);
    o << "  },\n";
    o << "  \"config\": {\n";
    o << "    \"cell_point_overflow_limit\": " << m.config.cell_point_overflow_limit << ",\n";
    o << "    \"sub_grid_dimension\": " << m.config.sub_grid_dimension << ",\n";
    o << "    \"max_cell_size\": " << format_f32(m.config.max_cell_size) << "\n";
    o << "  }\n";
    o << "}";
    return o.str();
}

// Minimal JSON reader for the metadata schema: any key order, unknown keys
// skipped, all six top-level keys required (serde derive, metadata.rs:9-28).
namespace {
struct JP {
    const char* p;
    const char* e;
    bool ok = true;
    void ws() {
        while (p < e && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) p++;
    }
    bool eat(char c) {
        ws();
        if (p < e && *p == c) {
            p++;
            return true;
        }
        return false;
    }
    std::string str() {
        ws();
        std::string s;
        if (p >= e || *p != '"') {
            ok = false;
            return s;
        }
        p++;
        while (p < e && *p != '"') {
            if (*p == '\\' && p + 1 < e) {
                p++;
                switch (*p) {
                    case 'n': s.push_back('\n'); break;
                    case 't': s.push_back('\t'); break;
                    case 'r': s.push_back('\r'); break;
                    case 'u': {
                        if (p + 4 < e) {
                            unsigned cp = (unsigned)std::strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
                            if (cp < 0x80) s.push_back((char)cp);
                            else s.push_back('?');
                            p += 4;
                        }
                        break;
                    }
                    default: s.push_back(*p);
                }
                p++;
            } else
                s.push_back(*p++);
        }
        if (p >= e) ok = false;
        else p++;
        return s;
    }
    double num() {
        ws();
        const char* b = p;
        while (p < e && (std::isdigit((unsigned char)*p) || *p == '-' || *p == '+' || *p == '.' || *p == 'e' || *p == 'E'))
            p++;
        if (b == p) {
            ok = false;
            return 0;
        }
        return std::strtod(std::string(b, p).c_str(), nullptr);
    }
    // f32 fields: serde parses the decimal text to f64 and casts to f32; for
    // shortest-round-trip f32 text this reproduces the original f32 exactly.
    float f32() { return (float)num(); }
    void skip() {
        ws();
        if (p >= e) {
            ok = false;
            return;
        }
        if (*p == '"') {
            str();
        } else if (*p == '{' || *p == '[') {
            char open = *p, close = (open == '{') ? '}' : ']';
            p++;
            ws();
            if (eat(close)) return;
            for (;;) {
                if (open == '{') {
                    str();
                    if (!eat(':')) { ok = false; return; }
                }
                skip();
                if (!ok) return;
                if (eat(',')) continue;
                if (eat(close)) return;
                ok = false;
                return;
            }
        } else {
            while (p < e && *p != ',' && *p != '}' && *p != ']' && *p != ' ' && *p != '\n') p++;
        }
    }
    bool vec3(Vec3* v) {
        if (!eat('[')) return false;
        v->x = f32();
        if (!eat(',')) return false;
        v->y = f32();
        if (!eat(',')) return false;
        v->z = f32();
        return eat(']');
    }
    template <class F>
    bool object(F&& field) {
        if (!eat('{')) return false;
        if (eat('}')) return true;
        for (;;) {
            std::string k = str();
            if (!ok || !eat(':')) return false;
            if (!field(k)) return false;
            if (eat(',')) continue;
            return eat('}');
        }
    }
};
}  // namespace

bool metadata_from_json(const std::string& text, Metadata* m) {
    JP j{text.data(), text.data() + text.size()};
    unsigned seen = 0;
    bool ok = j.object([&](const std::string& k) {
        if (k == "version") { m->version = j.str(); seen |= 1; }
        else if (k == "name") { m->name = j.str(); seen |= 2; }
        else if (k == "number_of_points") { m->number_of_points = (uint64_t)j.num(); seen |= 4; }
        else if (k == "hierarchies") { m->hierarchies = (uint32_t)j.num(); seen |= 8; }
        else if (k == "bounding_box") {
            unsigned s2 = 0;
            if (!j.object([&](const std::string& k2) {
                    if (k2 == "min") { s2 |= 1; return j.vec3(&m->bounding_box.min); }
                    if (k2 == "max") { s2 |= 2; return j.vec3(&m->bounding_box.max); }
                    j.skip();
                    return j.ok;
                }) || s2 != 3)
                return false;
            seen |= 16;
        } else if (k == "config") {
            unsigned s2 = 0;
            if (!j.object([&](const std::string& k2) {
                    if (k2 == "cell_point_overflow_limit") { m->config.cell_point_overflow_limit = (uint32_t)j.num(); s2 |= 1; }
                    else if (k2 == "sub_grid_dimension") { m->config.sub_grid_dimension = (uint32_t)j.num(); s2 |= 2; }
                    else if (k2 == "max_cell_size") { m->config.max_cell_size = j.f32(); s2 |= 4; }
                    else j.skip();
                    return j.ok;
                }) || s2 != 7)
                return false;
            seen |= 32;
        } else
            j.skip();
        return j.ok;
    });
    return ok && j.ok && seen == 63;
}

}  // namespace orc

// ---------------------------------------------------------------------------
// C ABI for ctypes (tests / bench cpu_baseline only).
using namespace orc;

struct OrcHandle {
    Converter conv;
    uint32_t batch_size = 10000;  // lib.rs:32
    int threads = 1;
    // threads > 1: sub-converters, one per worker, each owning a disjoint set of
    // level-0 cells (trees under different level-0 cells never interact,
    // converter.rs:124-134), fed the same batch sequence.
    std::vector<std::unique_ptr<Converter>> workers;
};

extern "C" {

// lib.rs:86-101 load_metadata + Converter::new (converter.rs:79-94).
// out_dir may be NULL/"" => pure in-memory run with default metadata.
void* orc_open(const char* out_dir, uint32_t batch_size, int threads) {
    auto* h = new OrcHandle();
    if (batch_size) h->batch_size = batch_size;
    h->threads = threads < 1 ? 1 : threads;
    if (out_dir && out_dir[0]) {
        h->conv.working_directory = out_dir;
        ::mkdir(out_dir, 0777);  // create_dir_all of the leaf; parents are the caller's job in tests
        std::vector<uint8_t> bytes;
        std::string path = std::string(out_dir) + "/metadata.json";
        FILE* f = std::fopen(path.c_str(), "rb");
        if (f) {
            std::fclose(f);
            std::ifstream in(path, std::ios::binary);
            std::stringstream ss;
            ss << in.rdbuf();
            Metadata m;
            if (!metadata_from_json(ss.str(), &m)) {  // reference: unwrap() panics
                delete h;
                return nullptr;
            }
            h->conv.metadata = m;
        }
    }
    return h;
}

// Tests only: override config for in-memory runs (the reference takes config
// from defaults or the existing metadata.json; there is no CLI flag).
void orc_set_config(void* hv, uint32_t limit, uint32_t dim, float max_cell_size) {
    auto* h = (OrcHandle*)hv;
    h->conv.metadata.config.cell_point_overflow_limit = limit;
    h->conv.metadata.config.sub_grid_dimension = dim;
    h->conv.metadata.config.max_cell_size = max_cell_size;
}

// One input file: lib.rs:31-52 — get_batch(batch_size) slices, never spanning
// files; a zero-point file still produces one (empty) batch.
int orc_add_file(void* hv, const Point* pts, uint64_t n) {
    auto* h = (OrcHandle*)hv;
    const uint64_t bs = h->batch_size;
    if (h->threads <= 1) {
        uint64_t off = 0;
        do {
            uint64_t len = std::min<uint64_t>(bs, n - off);
            std::vector<Point> batch(pts + off, pts + off + len);
            h->conv.add_points_batch(batch);
            off += len;
        } while (off < n);
        return h->conv.error;
    }
    // Parallel-over-level-0-cells variant (generous upper bound for a thread pool;
    // the reference CLI itself is single-threaded).  Results identical.
    const int T = h->threads;
    if (h->workers.empty())
        for (int t = 0; t < T; t++) {
            auto w = std::make_unique<Converter>();
            w->metadata = h->conv.metadata;
            w->working_directory = h->conv.working_directory;
            h->workers.push_back(std::move(w));
        }
    // bounding box / count / level-0 bookkeeping on the main converter
    {
        uint64_t off = 0;
        do {
            uint64_t len = std::min<uint64_t>(bs, n - off);
            std::vector<Point> batch(pts + off, pts + off + len);
            h->conv.update_bounding_box(batch);
            h->conv.metadata.number_of_points += len;
            off += len;
        } while (off < n);
    }
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) {
        th.emplace_back([=]() {
            Converter& w = *h->workers[t];
            const MetadataConfig cfg = w.metadata.config;
            const float cs0 = cfg.cell_size(0);
            uint64_t off = 0;
            std::vector<Point> mine;
            do {
                uint64_t len = std::min<uint64_t>(bs, n - off);
                mine.clear();
                for (uint64_t i = off; i < off + len; i++) {
                    IVec3 c = cfg.cell_index(Vec3{pts[i].x, pts[i].y, pts[i].z}, cs0);
                    if ((int)(IVec3Hash()(c) % (size_t)T) == t) mine.push_back(pts[i]);
                }
                // level 0 is visited by every batch (converter.rs:111,120) even when empty
                PointMap grouped = group_points(mine, 0, cfg);
                w.add_points_in_hierarchy(0, cfg, std::move(grouped));
                off += len;
            } while (off < n);
        });
    }
    for (auto& x : th) x.join();
    int err = 0;
    for (auto& w : h->workers) {
        err |= w->error;
        h->conv.metadata.hierarchies = std::max(h->conv.metadata.hierarchies, w->metadata.hierarchies);
    }
    return err;
}

// Drop (converter.rs:241-246).  Returns 0 on success.
int orc_finish(void* hv) {
    auto* h = (OrcHandle*)hv;
    for (auto& w : h->workers) {
        for (auto& kv : w->cells) h->conv.cells.emplace(kv.first, std::move(kv.second));
        w->cells.clear();
        h->conv.arrivals += w->arrivals;
        w->arrivals = 0;
        h->conv.points_by_delta += w->points_by_delta;
        w->points_by_delta = 0;
    }
    return h->conv.save_all() ? 0 : -1;
}

uint64_t orc_number_of_points(void* hv) { return ((OrcHandle*)hv)->conv.metadata.number_of_points; }
uint32_t orc_hierarchies(void* hv) { return ((OrcHandle*)hv)->conv.metadata.hierarchies; }
uint64_t orc_arrivals(void* hv) { return ((OrcHandle*)hv)->conv.arrivals; }
int64_t orc_points_by_delta(void* hv) { return ((OrcHandle*)hv)->conv.points_by_delta; }
uint64_t orc_cell_count(void* hv) { return ((OrcHandle*)hv)->conv.cells.size(); }
void orc_bounding_box(void* hv, float* out6) {
    auto& b = ((OrcHandle*)hv)->conv.metadata.bounding_box;
    out6[0] = b.min.x; out6[1] = b.min.y; out6[2] = b.min.z;
    out6[3] = b.max.x; out6[4] = b.max.y; out6[5] = b.max.z;
}

// Serialise every node (path + image) into one blob for in-memory comparison:
// repeated { u32 path_len, path bytes, u64 image_len, image bytes }.  Call with
// buf == NULL to get the size.
uint64_t orc_dump_nodes(void* hv, uint8_t* buf, uint64_t cap) {
    auto* h = (OrcHandle*)hv;
    uint64_t off = 0;
    std::vector<uint8_t> img;
    for (auto& kv : h->conv.cells) {
        std::string p = cell_path(kv.first);
        img.clear();
        kv.second->write_to(img);
        uint32_t pl = (uint32_t)p.size();
        uint64_t il = img.size();
        if (buf && off + 4 + pl + 8 + il <= cap) {
            std::memcpy(buf + off, &pl, 4);
            std::memcpy(buf + off + 4, p.data(), pl);
            std::memcpy(buf + off + 4 + pl, &il, 8);
            std::memcpy(buf + off + 12 + pl, img.data(), il);
        }
        off += 12 + pl + il;
    }
    return off;
}

int orc_metadata_json(void* hv, char* buf, uint64_t cap) {
    std::string s = metadata_to_json(((OrcHandle*)hv)->conv.metadata);
    if (buf && cap > s.size()) std::memcpy(buf, s.c_str(), s.size() + 1);
    return (int)s.size();
}

void orc_close(void* hv) { delete (OrcHandle*)hv; }

// The reference CLI's cell residency (converter.rs:92: LRU of 100 cells, evicted cells rewritten whole, misses
// re-read and re-hashed).  Needs an output directory; single-threaded runs only.  Returns 0 on success.
int orc_set_lru(void* hv, uint32_t capacity) {
    auto* h = (OrcHandle*)hv;
    if (h->conv.working_directory.empty() || h->threads > 1 || !h->conv.cells.empty()) return -1;
    h->conv.lru_capacity = capacity;
    return 0;
}
void orc_lru_stats(void* hv, uint64_t* evictions, uint64_t* loads) {
    auto* h = (OrcHandle*)hv;
    *evictions = h->conv.lru_evictions;
    *loads = h->conv.lru_loads;
}

// --- primitive probes for known-answer tests -------------------------------
float orc_cell_size(uint32_t limit, uint32_t dim, float max_cell_size, uint32_t h) {
    MetadataConfig c{limit, dim, max_cell_size};
    return c.cell_size(h);
}
void orc_cell_index(float max_cell_size, uint32_t h, const float* p, int32_t* out) {
    MetadataConfig c{5000, 96, max_cell_size};
    IVec3 i = c.cell_index(Vec3{p[0], p[1], p[2]}, c.cell_size(h));
    out[0] = i.x; out[1] = i.y; out[2] = i.z;
}
void orc_cell_pos(const int32_t* idx, float cell_size, float* out) {
    MetadataConfig c;
    Vec3 v = c.cell_pos(IVec3{idx[0], idx[1], idx[2]}, cell_size);
    out[0] = v.x; out[1] = v.y; out[2] = v.z;
}
void orc_hex_from_world(const float* p, float radius, int32_t* out) {
    OffsetIndex o = offset_from_world(Vec3{p[0], p[1], p[2]}, radius);
    out[0] = o.x; out[1] = o.y; out[2] = o.z;
}
void orc_hex_to_world(const int32_t* k, float radius, float* out) {
    Vec3 v = offset_to_world(OffsetIndex{k[0], k[1], k[2]}, radius);
    out[0] = v.x; out[1] = v.y; out[2] = v.z;
}
// d2 between a point and the reference centre of its own hex prism (cell.rs:75-78)
float orc_hex_d2(const float* p, float sub_cell_size) {
    float R = sub_cell_size / 2.0f;
    Vec3 pos{p[0], p[1], p[2]};
    OffsetIndex o = offset_from_world(pos, R);
    return distance_squared(offset_to_world(o, R), pos);
}
int orc_format_f32(float v, char* buf, uint64_t cap) {
    std::string s = format_f32(v);
    if (buf && cap > s.size()) std::memcpy(buf, s.c_str(), s.size() + 1);
    return (int)s.size();
}
}  // extern "C"

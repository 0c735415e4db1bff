// reader.cpp — input readers of the converter front end (host side, no CUDA).
// Mirrors get_batched_point_reader (point-converter/src/lib.rs:62-84) and the readers behind the trait
// BatchedPointReader (converter.rs:24-30):
//   *.ply            converter/ply.rs:35-82 + the property mapping of point.rs:56-131 (bug-compatible:
//                    ASCII files are parsed but contribute no points, ply.rs:43-51; float colours are
//                    divided by 255 and truncated, point.rs:96-98)
//   *.las            converter/las.rs:22-55 (uncompressed LAS only; LAZ needs the third-party laz 0.9.1
//                    arithmetic decoder and is reported as unsupported)
//   metadata.json    converter/own.rs:15-87 (re-ingest of a converted cloud)
// A reader hands back the points of one file; the caller cuts them into batches (lib.rs:31-52).
#include "reader.h"

#include <dirent.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <memory>
#include <sstream>
#include <thread>
#include <vector>

#include "builder.h"

namespace pcb {

static std::string extension_of(const std::string& path) {
    size_t slash = path.find_last_of('/');
    size_t dot = path.find_last_of('.');
    if (dot == std::string::npos || (slash != std::string::npos && dot < slash) || dot + 1 == path.size()) return "";
    if (dot == 0 || (slash != std::string::npos && dot == slash + 1)) return "";  // ".hidden" has no extension in Rust
    return path.substr(dot + 1);
}

static bool slurp(const std::string& path, std::vector<uint8_t>* out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out->resize(n > 0 ? (size_t)n : 0);
    size_t got = n > 0 ? std::fread(out->data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == out->size();
}

// Rust `f32 as u8`: saturating, NaN -> 0
static uint8_t sat_u8(float v) {
    if (!(v == v)) return 0;
    if (v <= 0.0f) return 0;
    if (v >= 255.0f) return 255;
    return (uint8_t)v;
}

// ---- PLY ------------------------------------------------------------------------------------------------
enum PlyType { T_I8, T_U8, T_I16, T_U16, T_I32, T_U32, T_F32, T_F64, T_BAD };
static PlyType ply_type(const std::string& s) {
    if (s == "char" || s == "int8") return T_I8;
    if (s == "uchar" || s == "uint8") return T_U8;
    if (s == "short" || s == "int16") return T_I16;
    if (s == "ushort" || s == "uint16") return T_U16;
    if (s == "int" || s == "int32") return T_I32;
    if (s == "uint" || s == "uint32") return T_U32;
    if (s == "float" || s == "float32") return T_F32;
    if (s == "double" || s == "float64") return T_F64;
    return T_BAD;
}
static int ply_size(PlyType t) {
    static const int sz[] = {1, 1, 2, 2, 4, 4, 4, 8, 0};
    return sz[t];
}
struct PlyProp {
    std::string name;
    PlyType type;
    int offset;
};
struct PlyElement {
    std::string name;
    uint64_t count = 0;
    std::vector<PlyProp> props;
    int stride = 0;
    bool has_list = false;
};

template <class T>
static T load_swapped(const uint8_t* p, bool swap) {
    uint8_t b[sizeof(T)];
    for (size_t i = 0; i < sizeof(T); i++) b[i] = swap ? p[sizeof(T) - 1 - i] : p[i];
    T v;
    memcpy(&v, b, sizeof(T));
    return v;
}

// Runs f(lo, hi) over [0, n) on up to 8 host threads (decoding a chunk is embarrassingly parallel).
template <class F>
static void chunk_parallel(uint64_t n, F f) {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned nt = (unsigned)std::min<uint64_t>(std::min(8u, hw), std::max<uint64_t>(1, n / 65536));
    if (nt <= 1) {
        f(0, n);
        return;
    }
    std::vector<std::thread> th;
    const uint64_t per = (n + nt - 1) / nt;
    for (unsigned t = 0; t < nt; t++) {
        const uint64_t lo = std::min<uint64_t>(n, t * per), hi = std::min<uint64_t>(n, lo + per);
        if (lo < hi) th.emplace_back([=] { f(lo, hi); });
    }
    for (auto& t : th) t.join();
}

// point.rs:61-129 set_property, resolved ONCE per property instead of per point: kind 0-2 = x, y, z; 3-6 = r, g, b, a
struct PlyAction {
    int kind;
    PlyType type;
    int offset;
};
static void ply_decode(const uint8_t* src, int stride, const std::vector<PlyAction>& acts, bool swap, bool fast15,
                       pcb_point* out, uint64_t lo, uint64_t hi) {
    for (uint64_t i = lo; i < hi; i++) {
        const uint8_t* p = src + i * (uint64_t)stride;
        pcb_point pt;
        if (fast15) {   // float x, y, z + uchar red, green, blue, little endian: the layout every exporter writes
            memcpy(&pt, p, 15);
            pt.rgba[3] = 255;
            out[i] = pt;
            continue;
        }
        pt.x = pt.y = pt.z = 0.0f;
        pt.rgba[0] = pt.rgba[1] = pt.rgba[2] = 0;
        pt.rgba[3] = 255;  // Point::default (point.rs:16-23)
        for (const PlyAction& a : acts) {
            const uint8_t* q = p + a.offset;
            if (a.kind < 3) {
                float* dst = a.kind == 0 ? &pt.x : a.kind == 1 ? &pt.y : &pt.z;
                if (a.type == T_F32) *dst = load_swapped<float>(q, swap);
                else if (a.type == T_F64) *dst = (float)load_swapped<double>(q, swap);
                // any other type is ignored: the coordinate stays 0 (point.rs:74, 83, 92 `_ => {}`)
            } else {
                if (a.type == T_U8) pt.rgba[a.kind - 3] = *q;
                else if (a.type == T_F32) pt.rgba[a.kind - 3] = sat_u8(load_swapped<float>(q, swap) / 255.0f);  // point.rs:96-98
            }
        }
        out[i] = pt;
    }
}

// ---- the streaming readers --------------------------------------------------------------------------------
struct PointStream::Impl {
    enum Kind { PLY, LAS, OWN } kind = PLY;
    FILE* f = nullptr;
    std::string path;
    uint64_t total = 0, delivered = 0;
    bool ascii = false, failed = false;
    // PLY
    int stride = 0;
    bool swap = false, fast15 = false;
    std::vector<PlyAction> acts;
    // LAS
    LasLayout las;
    // own format: node files still to read
    std::vector<std::string> files;
    size_t next_file = 0;
    std::vector<pcb_point> carry;   // points of a node file that did not fit the last chunk
    size_t carry_at = 0;
    std::vector<uint8_t> raw;
    ~Impl() { if (f) std::fclose(f); }
};

PointStream::PointStream() : impl_(new Impl()) {}
PointStream::~PointStream() { delete impl_; }
uint64_t PointStream::total_points() const { return impl_->total; }
bool PointStream::ascii_ply() const { return impl_->ascii; }
const LasLayout* PointStream::las_layout() const { return impl_->kind == Impl::LAS ? &impl_->las : nullptr; }

static int open_ply(PointStream::Impl* im, std::string* msg) {
    im->f = std::fopen(im->path.c_str(), "rb");
    if (!im->f) { *msg = "cannot open " + im->path; return READ_FATAL; }  // File::open(path).unwrap() (ply.rs:20)
    auto next_line = [&](std::string* line) {
        line->clear();
        int c;
        bool any = false;
        while ((c = std::fgetc(im->f)) != EOF) {
            any = true;
            if (c == '\n') break;
            line->push_back((char)c);
            if (line->size() > 4096) return false;
        }
        if (!line->empty() && line->back() == '\r') line->pop_back();
        return any;
    };
    std::string line;
    if (!next_line(&line) || line != "ply") { *msg = "not a PLY file: " + im->path; return READ_FATAL; }  // ply.rs:24
    int encoding = -1;  // 0 ascii, 1 little, 2 big
    std::vector<PlyElement> elements;
    bool ended = false;
    while (next_line(&line)) {
        std::istringstream ss(line);
        std::string kw;
        ss >> kw;
        if (kw == "end_header") { ended = true; break; }
        if (kw == "format") {
            std::string f;
            ss >> f;
            encoding = f == "ascii" ? 0 : f == "binary_little_endian" ? 1 : f == "binary_big_endian" ? 2 : -1;
        } else if (kw == "element") {
            PlyElement e;
            ss >> e.name >> e.count;
            elements.push_back(e);
        } else if (kw == "property" && !elements.empty()) {
            std::string t;
            ss >> t;
            PlyElement& e = elements.back();
            if (t == "list") {
                e.has_list = true;
            } else {
                PlyProp p;
                p.type = ply_type(t);
                ss >> p.name;
                p.offset = e.stride;
                if (p.type == T_BAD) { *msg = "unknown PLY property type " + t; return READ_FATAL; }
                e.stride += ply_size(p.type);
                e.props.push_back(p);
            }
        }
    }
    if (!ended || encoding < 0) { *msg = "malformed PLY header: " + im->path; return READ_FATAL; }
    const PlyElement* vertex = nullptr;
    uint64_t skip = 0;
    for (auto& e : elements) {
        if (e.name == "vertex") { vertex = &e; break; }
        if (e.has_list) { *msg = "PLY list element before the vertices is not supported"; return READ_FATAL; }
        skip += (uint64_t)e.stride * e.count;
    }
    if (!vertex) { *msg = "PLY file without a vertex element"; return READ_FATAL; }  // .unwrap() ply.rs:37
    im->total = vertex->count;
    if (encoding == 0) {   // ply.rs:43-51: the ASCII branch parses every line and pushes NOTHING; batches are empty
        im->ascii = true;
        return READ_OK;
    }
    if (vertex->has_list) { *msg = "PLY vertex element with a list property is not supported"; return READ_FATAL; }
    im->swap = encoding == 2;  // host is little-endian
    im->stride = vertex->stride;
    for (auto& pr : vertex->props) {
        const std::string& n = pr.name;
        int kind = n == "x" ? 0 : n == "y" ? 1 : n == "z" ? 2 : (n == "red" || n == "r") ? 3 : (n == "green" || n == "g") ? 4
                   : (n == "blue" || n == "b") ? 5 : (n == "alpha" || n == "a") ? 6 : -1;
        if (kind >= 0) im->acts.push_back(PlyAction{kind, pr.type, pr.offset});
    }
    im->fast15 = !im->swap && im->stride == 15 && im->acts.size() == 6;
    for (int k = 0; im->fast15 && k < 6; k++) {
        const PlyAction& a = im->acts[k];
        im->fast15 = a.kind == k && a.offset == (k < 3 ? 4 * k : 9 + k) && a.type == (k < 3 ? T_F32 : T_U8);
    }
    if (skip && fseeko(im->f, (off_t)skip, SEEK_CUR) != 0) { *msg = "cannot seek in " + im->path; return READ_FATAL; }
    return READ_OK;
}

// converter/las.rs:32-41: pos = (x, y, z) as f32 of the scaled f64 coordinates (las 0.8.4: X * scale + offset),
// colour = low byte of each u16 channel (`as u8`), alpha 255; formats without colour give (0, 0, 0).
static int open_las(PointStream::Impl* im, std::string* msg) {
    im->f = std::fopen(im->path.c_str(), "rb");
    if (!im->f) { *msg = "cannot open " + im->path; return READ_FATAL; }
    uint8_t d[375];
    const size_t got = std::fread(d, 1, sizeof d, im->f);
    if (got < 227 || memcmp(d, "LASF", 4) != 0) { *msg = "not a LAS file: " + im->path; return READ_FATAL; }
    auto u16 = [&](size_t o) { return (uint32_t)d[o] | ((uint32_t)d[o + 1] << 8); };
    auto u32 = [&](size_t o) { return u16(o) | (u16(o + 2) << 16); };
    auto u64 = [&](size_t o) { return (uint64_t)u32(o) | ((uint64_t)u32(o + 4) << 32); };
    auto f64 = [&](size_t o) { double v; memcpy(&v, d + o, 8); return v; };
    const uint8_t ver_minor = d[25];
    const uint32_t offset_to_points = u32(96);
    uint8_t format = d[104];
    const uint32_t rec_len = u16(105);
    uint64_t count = u32(107);
    if (ver_minor >= 4 && got >= 255 && count == 0) count = u64(247);
    if (format & 0x80) { *msg = "LAZ (compressed) input is not supported: " + im->path; return READ_UNSUPPORTED; }
    format &= 0x3F;
    LasLayout& L = im->las;
    L.scale[0] = f64(131); L.scale[1] = f64(139); L.scale[2] = f64(147);
    L.offset[0] = f64(155); L.offset[1] = f64(163); L.offset[2] = f64(171);
    L.colour_off = -1;  // byte offset of (red, green, blue) u16 inside a record
    switch (format) {
        case 2: L.colour_off = 20; break;
        case 3: L.colour_off = 28; break;
        case 5: L.colour_off = 28; break;
        case 7: case 8: case 10: L.colour_off = 30; break;
        default: break;
    }
    L.stride = rec_len;
    if (L.colour_off >= 0 && (uint32_t)L.colour_off + 6 > rec_len) L.colour_off = -1;
    im->total = count;
    if (rec_len < 20) { *msg = "bad LAS record length"; return READ_FATAL; }
    if (fseeko(im->f, (off_t)offset_to_points, SEEK_SET) != 0) { *msg = "cannot seek in " + im->path; return READ_FATAL; }
    return READ_OK;
}

void las_decode_host(const uint8_t* rec, uint64_t n, const LasLayout& L, pcb_point* out) {
    chunk_parallel(n, [&](uint64_t lo, uint64_t hi) {
        for (uint64_t i = lo; i < hi; i++) {
            const uint8_t* p = rec + i * (uint64_t)L.stride;
            int32_t X, Y, Z;
            memcpy(&X, p, 4); memcpy(&Y, p + 4, 4); memcpy(&Z, p + 8, 4);
            pcb_point pt;
            pt.x = (float)((double)X * L.scale[0] + L.offset[0]);
            pt.y = (float)((double)Y * L.scale[1] + L.offset[1]);
            pt.z = (float)((double)Z * L.scale[2] + L.offset[2]);
            pt.rgba[0] = pt.rgba[1] = pt.rgba[2] = 0;
            pt.rgba[3] = 255;
            if (L.colour_off >= 0) {
                pt.rgba[0] = p[L.colour_off];       // `color.red as u8`: low byte of the little-endian u16 (las.rs:40)
                pt.rgba[1] = p[L.colour_off + 2];
                pt.rgba[2] = p[L.colour_off + 4];
            }
            out[i] = pt;
        }
    });
}

// converter/own.rs:21-49: every h_<level> directory below the metadata file, every cell file in it, all_points()
// = grid points then overflow points (cell.rs:66-68).  The reference walks read_dir order (OS dependent) and
// HashMap order; here: files sorted by name, grid in file order, overflow entries in file order - the multiset of
// points is the same.
static int open_own(PointStream::Impl* im, std::string* msg) {
    std::vector<uint8_t> bytes;
    Metadata meta;
    if (!slurp(im->path, &bytes) || !metadata_from_json(std::string(bytes.begin(), bytes.end()), &meta)) {
        *msg = "Couldn't parse metadata at " + im->path;
        return READ_FATAL;  // BatchedPointCloudPointReader::new(path).unwrap() (lib.rs:75-77)
    }
    size_t slash = im->path.find_last_of('/');
    const std::string dir = slash == std::string::npos ? "." : im->path.substr(0, slash);
    im->total = meta.number_of_points;
    for (uint32_t h = 0; h < meta.hierarchies; h++) {
        const std::string hd = dir + "/h_" + std::to_string(h);
        DIR* dp = opendir(hd.c_str());
        if (!dp) continue;  // "Failed to read dir" is logged and skipped (own.rs:26-32)
        std::vector<std::string> names;
        while (dirent* e = readdir(dp)) {
            std::string nm = e->d_name;
            if (nm != "." && nm != "..") names.push_back(hd + "/" + nm);
        }
        closedir(dp);
        std::sort(names.begin(), names.end());
        im->files.insert(im->files.end(), names.begin(), names.end());
    }
    return READ_OK;
}

// lib.rs:62-84
PointStream* PointStream::open(const std::string& path, int* status, std::string* message) {
    std::unique_ptr<PointStream> ps(new PointStream());
    Impl* im = ps->impl_;
    im->path = path;
    const std::string ext = extension_of(path);
    message->clear();
    if (ext == "ply") { im->kind = Impl::PLY; *status = open_ply(im, message); }
    else if (ext == "las") { im->kind = Impl::LAS; *status = open_las(im, message); }
    else if (ext == "laz") { *status = READ_UNSUPPORTED; *message = "LAZ (compressed) input is not supported: " + path; }
    else if (ext == "json") { im->kind = Impl::OWN; *status = open_own(im, message); }
    else { *status = READ_UNSUPPORTED; *message = ext.empty() ? "" : "Unsupported file format '" + ext + "'"; }
    if (*status != READ_OK) return nullptr;
    return ps.release();
}

// Raw LAS records for the device-side decode (pcb_add_las_records): up to max_points records into `out`.
int PointStream::next_raw(uint64_t max_points, uint8_t* out, uint64_t* n, std::string* message) {
    Impl* im = impl_;
    *n = 0;
    if (im->kind != Impl::LAS || im->failed || im->delivered >= im->total) return READ_OK;
    const uint64_t want = std::min<uint64_t>(max_points, im->total - im->delivered);
    const size_t got = std::fread(out, im->las.stride, want, im->f);
    *n = got;
    im->delivered += got;
    if (got < want) {  // reader error mid-file: logged, the file ends there (lib.rs:36-39)
        im->failed = true;
        *message = "unexpected end of file in " + im->path;
        return READ_TRUNCATED;
    }
    return READ_OK;
}

int PointStream::next(uint64_t max_points, pcb_point* out, uint64_t* n, std::string* message) {
    Impl* im = impl_;
    *n = 0;
    if (im->failed || im->ascii) return READ_OK;
    if (im->kind == Impl::OWN) {
        uint64_t got = 0;
        while (got < max_points) {
            if (im->carry_at < im->carry.size()) {
                const uint64_t take = std::min<uint64_t>(max_points - got, im->carry.size() - im->carry_at);
                memcpy(out + got, im->carry.data() + im->carry_at, take * sizeof(pcb_point));
                im->carry_at += take;
                got += take;
                continue;
            }
            if (im->next_file >= im->files.size()) break;
            std::vector<uint8_t> cell;
            ExistingNode node;
            im->carry.clear();
            im->carry_at = 0;
            if (!slurp(im->files[im->next_file++], &cell) || !parse_node(cell.data(), cell.size(), &node)) continue;  // own.rs:42-48
            im->carry = std::move(node.grid);
            for (auto& en : node.entries) im->carry.insert(im->carry.end(), en.list.begin(), en.list.end());
        }
        *n = got;
        im->delivered += got;
        return READ_OK;
    }
    if (im->delivered >= im->total) return READ_OK;
    const uint64_t want = std::min<uint64_t>(max_points, im->total - im->delivered);
    const int stride = im->kind == Impl::LAS ? (int)im->las.stride : im->stride;
    im->raw.resize((size_t)want * stride);
    const size_t got = stride ? std::fread(im->raw.data(), stride, want, im->f) : 0;
    if (im->kind == Impl::LAS) {
        las_decode_host(im->raw.data(), got, im->las, out);
    } else {
        const uint8_t* src = im->raw.data();
        chunk_parallel(got, [&](uint64_t lo, uint64_t hi) { ply_decode(src, stride, im->acts, im->swap, im->fast15, out, lo, hi); });
    }
    *n = got;
    im->delivered += got;
    if (got < want) {  // reader error mid-file: logged, the file ends there (lib.rs:36-39)
        im->failed = true;
        *message = "unexpected end of file in " + im->path;
        return READ_TRUNCATED;
    }
    return READ_OK;
}

// Reads every point of `path` (appended to *out): the streaming reader drained into one vector (tests, --dump-points).
int read_points_file(const std::string& path, std::vector<pcb_point>* out, ReadInfo* info) {
    int status = READ_OK;
    std::unique_ptr<PointStream> ps(PointStream::open(path, &status, &info->message));
    if (!ps) return status;
    info->total_points = ps->total_points();
    info->ascii_ply = ps->ascii_ply();
    uint64_t got = 0;
    for (;;) {
        const size_t at = out->size();
        const uint64_t chunk = 1u << 20;
        out->resize(at + chunk);
        uint64_t n = 0;
        status = ps->next(chunk, out->data() + at, &n, &info->message);
        out->resize(at + n);
        got += n;
        if (n == 0 || status != READ_OK) break;
    }
    info->points_read = got;
    return status;
}

}  // namespace pcb

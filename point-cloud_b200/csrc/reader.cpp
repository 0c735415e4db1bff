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
#include <sstream>

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

// point.rs:61-129 set_property
static void ply_set(pcb_point* pt, const PlyProp& pr, const uint8_t* src, bool swap) {
    const std::string& n = pr.name;
    int axis = n == "x" ? 0 : n == "y" ? 1 : n == "z" ? 2 : -1;
    if (axis >= 0) {
        float* dst = axis == 0 ? &pt->x : axis == 1 ? &pt->y : &pt->z;
        if (pr.type == T_F32) *dst = load_swapped<float>(src, swap);
        else if (pr.type == T_F64) *dst = (float)load_swapped<double>(src, swap);
        return;  // any other type is ignored: the coordinate stays 0 (point.rs:74, 83, 92 `_ => {}`)
    }
    int ch = (n == "red" || n == "r") ? 0 : (n == "green" || n == "g") ? 1 : (n == "blue" || n == "b") ? 2
             : (n == "alpha" || n == "a") ? 3 : -1;
    if (ch >= 0) {
        if (pr.type == T_U8) pt->rgba[ch] = *src;
        else if (pr.type == T_F32) pt->rgba[ch] = sat_u8(load_swapped<float>(src, swap) / 255.0f);  // point.rs:96-98
    }
}

static int read_ply(const std::string& path, std::vector<pcb_point>* out, ReadInfo* info) {
    std::vector<uint8_t> data;
    if (!slurp(path, &data)) {
        info->message = "cannot open " + path;
        return READ_FATAL;  // File::open(path).unwrap() (ply.rs:20)
    }
    // header: text lines up to "end_header"
    size_t pos = 0;
    auto next_line = [&](std::string* line) {
        if (pos >= data.size()) return false;
        size_t e = pos;
        while (e < data.size() && data[e] != '\n') e++;
        line->assign((const char*)data.data() + pos, e - pos);
        if (!line->empty() && line->back() == '\r') line->pop_back();
        pos = e + 1;
        return true;
    };
    std::string line;
    if (!next_line(&line) || line != "ply") {
        info->message = "not a PLY file: " + path;
        return READ_FATAL;  // read_header(..).unwrap() (ply.rs:24)
    }
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
                if (p.type == T_BAD) { info->message = "unknown PLY property type " + t; return READ_FATAL; }
                e.stride += ply_size(p.type);
                e.props.push_back(p);
            }
        }
    }
    if (!ended || encoding < 0) { info->message = "malformed PLY header: " + path; return READ_FATAL; }
    const PlyElement* vertex = nullptr;
    size_t skip = 0;
    for (auto& e : elements) {
        if (e.name == "vertex") { vertex = &e; break; }
        if (e.has_list) { info->message = "PLY list element before the vertices is not supported"; return READ_FATAL; }
        skip += (size_t)e.stride * e.count;
    }
    if (!vertex) { info->message = "PLY file without a vertex element"; return READ_FATAL; }  // .unwrap() ply.rs:37
    info->total_points = vertex->count;
    if (encoding == 0) {
        // ply.rs:43-51: the ASCII branch parses every line and pushes NOTHING; batches are empty
        info->points_read = 0;
        info->ascii_ply = true;
        return READ_OK;
    }
    if (vertex->has_list) { info->message = "PLY vertex element with a list property is not supported"; return READ_FATAL; }
    const bool swap = encoding == 2;  // host is little-endian
    const uint8_t* p = data.data() + pos + skip;
    const uint64_t avail = (pos + skip <= data.size() && vertex->stride) ? (data.size() - pos - skip) / vertex->stride : 0;
    const uint64_t n = std::min<uint64_t>(vertex->count, avail);
    out->reserve(out->size() + n);
    for (uint64_t i = 0; i < n; i++, p += vertex->stride) {
        pcb_point pt;
        pt.x = pt.y = pt.z = 0.0f;
        pt.rgba[0] = pt.rgba[1] = pt.rgba[2] = 0;
        pt.rgba[3] = 255;  // Point::default (point.rs:16-23)
        for (auto& pr : vertex->props) ply_set(&pt, pr, p + pr.offset, swap);
        out->push_back(pt);
    }
    info->points_read = n;
    if (n < vertex->count) {  // reader error mid-file: logged, the file ends there (lib.rs:36-39)
        info->message = "unexpected end of file in " + path;
        return READ_TRUNCATED;
    }
    return READ_OK;
}

// ---- LAS (uncompressed) ------------------------------------------------------------------------------------
// converter/las.rs:32-41: pos = (x, y, z) as f32 of the scaled f64 coordinates (las 0.8.4: X * scale + offset),
// colour = low byte of each u16 channel (`as u8`), alpha 255; formats without colour give (0, 0, 0).
static int read_las(const std::string& path, std::vector<pcb_point>* out, ReadInfo* info) {
    std::vector<uint8_t> d;
    if (!slurp(path, &d)) { info->message = "cannot open " + path; return READ_FATAL; }
    if (d.size() < 227 || memcmp(d.data(), "LASF", 4) != 0) { info->message = "not a LAS file: " + path; return READ_FATAL; }
    auto u16 = [&](size_t o) { return (uint32_t)d[o] | ((uint32_t)d[o + 1] << 8); };
    auto u32 = [&](size_t o) { return u16(o) | (u16(o + 2) << 16); };
    auto u64 = [&](size_t o) { return (uint64_t)u32(o) | ((uint64_t)u32(o + 4) << 32); };
    auto f64 = [&](size_t o) { double v; memcpy(&v, d.data() + o, 8); return v; };
    const uint8_t ver_minor = d[25];
    const uint32_t offset_to_points = u32(96);
    uint8_t format = d[104];
    const uint32_t rec_len = u16(105);
    uint64_t count = u32(107);
    if (ver_minor >= 4 && d.size() >= 255 && count == 0) count = u64(247);
    if (format & 0x80) { info->message = "LAZ (compressed) input is not supported: " + path; return READ_UNSUPPORTED; }
    format &= 0x3F;
    const double sx = f64(131), sy = f64(139), sz = f64(147), ox = f64(155), oy = f64(163), oz = f64(171);
    int colour_off = -1;  // byte offset of (red, green, blue) u16 inside a record
    switch (format) {
        case 2: colour_off = 20; break;
        case 3: colour_off = 28; break;
        case 5: colour_off = 28; break;
        case 7: case 8: case 10: colour_off = 30; break;
        default: break;
    }
    info->total_points = count;
    if (rec_len < 20) { info->message = "bad LAS record length"; return READ_FATAL; }
    const uint64_t avail = offset_to_points <= d.size() ? (d.size() - offset_to_points) / rec_len : 0;
    const uint64_t n = std::min<uint64_t>(count, avail);
    out->reserve(out->size() + n);
    for (uint64_t i = 0; i < n; i++) {
        const size_t o = offset_to_points + (size_t)i * rec_len;
        const int32_t X = (int32_t)u32(o), Y = (int32_t)u32(o + 4), Z = (int32_t)u32(o + 8);
        pcb_point pt;
        pt.x = (float)((double)X * sx + ox);
        pt.y = (float)((double)Y * sy + oy);
        pt.z = (float)((double)Z * sz + oz);
        pt.rgba[0] = pt.rgba[1] = pt.rgba[2] = 0;
        pt.rgba[3] = 255;
        if (colour_off >= 0 && (uint32_t)colour_off + 6 <= rec_len) {
            pt.rgba[0] = d[o + colour_off];       // `color.red as u8`: low byte of the little-endian u16 (las.rs:40)
            pt.rgba[1] = d[o + colour_off + 2];
            pt.rgba[2] = d[o + colour_off + 4];
        }
        out->push_back(pt);
    }
    info->points_read = n;
    if (n < count) { info->message = "unexpected end of file in " + path; return READ_TRUNCATED; }
    return READ_OK;
}

// ---- converted cloud as input (metadata.json) -----------------------------------------------------------------
// converter/own.rs:21-49: every h_<level> directory below the metadata file, every cell file in it, all_points()
// = grid points then overflow points (cell.rs:66-68).  The reference walks read_dir order (OS dependent) and
// HashMap order; here: files sorted by name, grid in file order, overflow entries in file order — the multiset of
// points is the same.
static int read_own(const std::string& path, std::vector<pcb_point>* out, ReadInfo* info) {
    std::vector<uint8_t> bytes;
    Metadata meta;
    if (!slurp(path, &bytes) || !metadata_from_json(std::string(bytes.begin(), bytes.end()), &meta)) {
        info->message = "Couldn't parse metadata at " + path;
        return READ_FATAL;  // BatchedPointCloudPointReader::new(path).unwrap() (lib.rs:75-77)
    }
    size_t slash = path.find_last_of('/');
    const std::string dir = slash == std::string::npos ? "." : path.substr(0, slash);
    info->total_points = meta.number_of_points;
    uint64_t got = 0;
    for (uint32_t h = 0; h < meta.hierarchies; h++) {
        const std::string hd = dir + "/h_" + std::to_string(h);
        DIR* dp = opendir(hd.c_str());
        if (!dp) continue;  // "Failed to read dir" is logged and skipped (own.rs:26-32)
        std::vector<std::string> names;
        while (dirent* e = readdir(dp)) {
            std::string nm = e->d_name;
            if (nm != "." && nm != "..") names.push_back(nm);
        }
        closedir(dp);
        std::sort(names.begin(), names.end());
        for (auto& nm : names) {
            std::vector<uint8_t> cell;
            ExistingNode node;
            if (!slurp(hd + "/" + nm, &cell) || !parse_node(cell.data(), cell.size(), &node)) continue;  // own.rs:42-48
            out->insert(out->end(), node.grid.begin(), node.grid.end());
            got += node.grid.size();
            for (auto& en : node.entries) {
                out->insert(out->end(), en.list.begin(), en.list.end());
                got += en.list.size();
            }
        }
    }
    info->points_read = got;
    return READ_OK;
}

// lib.rs:62-84
int read_points_file(const std::string& path, std::vector<pcb_point>* out, ReadInfo* info) {
    const std::string ext = extension_of(path);
    if (ext == "ply") return read_ply(path, out, info);
    if (ext == "las") return read_las(path, out, info);
    if (ext == "laz") {
        info->message = "LAZ (compressed) input is not supported: " + path;
        return READ_UNSUPPORTED;
    }
    if (ext == "json") return read_own(path, out, info);
    info->message = ext.empty() ? "" : "Unsupported file format '" + ext + "'";
    return READ_UNSUPPORTED;
}

}  // namespace pcb

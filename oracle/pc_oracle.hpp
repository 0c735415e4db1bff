// pc_oracle.hpp — TEST INFRASTRUCTURE ONLY.
//
// Sequential CPU restatement of the reference's point-converter octree/LOD
// construction (Seiichi-Yahiro/point-cloud, point-converter crate).  It is the
// parity oracle for the CUDA path and the timed CPU baseline of bench.py.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load it.  The product (point-cloud_b200/) never links it.
//
// PARITY UNPINNED: the reference holds no tests, fixtures or golden vectors for
// this path and cannot be compiled here (no Rust toolchain), so
// this restatement is pinned only by hand-derived known-answer tests
// (tests/test_oracle_kat.py) and by an independent pure-Python restatement
// (oracle/model.py) that must agree with it bit for bit.
//
// Every function cites the reference file:line (relative to /root/reference)
// it follows.  Arithmetic is IEEE-754 binary32, round-to-nearest-even, no FMA
// contraction (build with -ffp-contract=off), evaluated left to right exactly
// as the Rust source (glam 0.27.0 scalar Vec3 on x86-64).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <list>
#include <map>
#include <memory>
#include <optional>
#include <string>
#include <unordered_map>
#include <vector>

namespace orc {

// point-converter/src/point.rs:8-14 — 16-byte record {pos: Vec3, color: [u8;4]}.
struct Point {
    float x, y, z;
    uint8_t color[4];
};
static_assert(sizeof(Point) == 16, "Point must be 16 bytes");

struct IVec3 {
    int32_t x, y, z;
    bool operator==(const IVec3& o) const { return x == o.x && y == o.y && z == o.z; }
    bool operator<(const IVec3& o) const {
        if (x != o.x) return x < o.x;
        if (y != o.y) return y < o.y;
        return z < o.z;
    }
};
struct IVec3Hash {
    size_t operator()(const IVec3& v) const {
        uint64_t h = (uint32_t)v.x * 0x9E3779B97F4A7C15ull;
        h = (h ^ (uint32_t)v.y) * 0xC2B2AE3D27D4EB4Full;
        h = (h ^ (uint32_t)v.z) * 0x165667B19E3779F9ull;
        return (size_t)(h ^ (h >> 29));
    }
};

struct Vec3 {
    float x, y, z;
};

// Rust `f32 as i32`: saturating, NaN -> 0 (language semantics since 1.45).
inline int32_t sat_i32(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return INT32_MAX;
    if (v <= -2147483648.0f) return INT32_MIN;
    return (int32_t)v;  // truncation toward zero
}

// point-converter/src/metadata.rs:67-88
struct MetadataConfig {
    uint32_t cell_point_overflow_limit = 5000;
    uint32_t sub_grid_dimension = 96;
    float max_cell_size = 1000.0f;

    // metadata.rs:91-93 — max_cell_size / 2u32.pow(h) as f32
    float cell_size(uint32_t hierarchy) const {
        return max_cell_size / (float)(uint32_t)(1u << hierarchy);
    }
    // metadata.rs:95-97
    float sub_cell_size(float cell_size) const { return cell_size / (float)sub_grid_dimension; }
    // metadata.rs:100-102 — (pos / cell_size).floor().as_ivec3()
    IVec3 cell_index(Vec3 pos, float cell_size) const {
        return IVec3{sat_i32(std::floor(pos.x / cell_size)), sat_i32(std::floor(pos.y / cell_size)),
                     sat_i32(std::floor(pos.z / cell_size))};
    }
    // metadata.rs:104-106 — cell_index.as_vec3() * cell_size + cell_size / 2.0
    Vec3 cell_pos(IVec3 idx, float cell_size) const {
        float half = cell_size / 2.0f;
        return Vec3{(float)idx.x * cell_size + half, (float)idx.y * cell_size + half,
                    (float)idx.z * cell_size + half};
    }
};

// bounding-volume/src/lib.rs:4-8
struct Aabb {
    Vec3 min{0, 0, 0}, max{0, 0, 0};
};

// point-converter/src/metadata.rs:9-41
struct Metadata {
    std::string version = "1.0";
    std::string name = "Unknown";
    uint64_t number_of_points = 0;
    uint32_t hierarchies = 0;
    Aabb bounding_box;
    MetadataConfig config;
};

// point-converter/src/hex.rs:10-15, 37-42
struct OffsetIndex {
    int32_t x, y, z;
    bool operator==(const OffsetIndex& o) const { return x == o.x && y == o.y && z == o.z; }
    bool operator<(const OffsetIndex& o) const {
        if (z != o.z) return z < o.z;
        if (y != o.y) return y < o.y;
        return x < o.x;
    }
};
struct OffsetHash {
    size_t operator()(const OffsetIndex& v) const { return IVec3Hash()(IVec3{v.x, v.y, v.z}); }
};
struct AxialIndex {
    int32_t q, r, h;
};

constexpr float SQRT_3 = 1.73205080757f;  // hex.rs:3

// wrapping i32 helpers (Rust release builds wrap; values never get near the limits)
inline int32_t wsub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
inline int32_t wadd(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
inline int32_t wneg(int32_t a) { return (int32_t)(0u - (uint32_t)a); }

// hex.rs:18-24
inline AxialIndex to_axial(OffsetIndex o) {
    return AxialIndex{wsub(o.x, wsub(o.y, (o.y & 1)) / 2), o.y, o.z};
}
// hex.rs:45-51
inline OffsetIndex to_offset(AxialIndex a) {
    return OffsetIndex{wadd(a.q, wsub(a.r, (a.r & 1)) / 2), a.r, a.h};
}
// hex.rs:55-65
inline Vec3 axial_to_world(AxialIndex a, float cell_radius) {
    float q = (float)a.q, r = (float)a.r, h = (float)a.h;
    float x = cell_radius * (SQRT_3 * q + SQRT_3 / 2.0f * r);
    float y = cell_radius * 3.0f / 2.0f * r;
    float z = h * cell_radius;
    return Vec3{x, y, z};
}
// hex.rs:67-85
inline AxialIndex axial_from_world(Vec3 pos, float cell_radius) {
    float x = pos.x / (cell_radius * SQRT_3);
    float y = pos.y / (-cell_radius * SQRT_3);
    float t = SQRT_3 * y + 1.0f;
    float temp1 = std::floor(t + x);
    float temp2 = t - x;
    float temp3 = 2.0f * x + 1.0f;
    float qf = (temp1 + temp3) / 3.0f;
    float rf = (temp1 + temp2) / 3.0f;
    int32_t q = sat_i32(std::floor(qf));
    int32_t r = wneg(sat_i32(std::floor(rf)));
    int32_t h = sat_i32(pos.z / cell_radius);
    return AxialIndex{q, r, h};
}
// hex.rs:28-34
inline Vec3 offset_to_world(OffsetIndex o, float cell_radius) { return axial_to_world(to_axial(o), cell_radius); }
inline OffsetIndex offset_from_world(Vec3 pos, float cell_radius) {
    return to_offset(axial_from_world(pos, cell_radius));
}

// glam 0.27.0 Vec3::distance_squared (scalar): (a-b).dot(a-b) = dx*dx + dy*dy + dz*dz
inline float distance_squared(Vec3 a, Vec3 b) {
    float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
    return dx * dx + dy * dy + dz * dz;
}

// point-converter/src/cell.rs:14-18
struct CellId {
    uint32_t hierarchy;
    IVec3 index;
    bool operator==(const CellId& o) const { return hierarchy == o.hierarchy && index == o.index; }
};
struct CellIdHash {
    size_t operator()(const CellId& c) const { return IVec3Hash()(c.index) * 31u + c.hierarchy; }
};

// point-converter/src/cell.rs:238-278
struct Header {
    CellId id;
    uint32_t total_number_of_points = 0;
    uint32_t number_of_points = 0;
    uint32_t number_of_overflow_points = 0;
    float size = 0;
    float sub_cell_size = 0;
    Vec3 pos{0, 0, 0};
    // cell.rs:276-278
    OffsetIndex sub_grid_index_for_point(const Point& p) const {
        return offset_from_world(Vec3{p.x, p.y, p.z}, sub_cell_size / 2.0f);
    }
};

// point-converter/src/cell.rs:33-38
struct Cell {
    Header header;
    std::unordered_map<OffsetIndex, Point, OffsetHash> points_grid;
    // None marker == std::nullopt (child flushed).  std::map: iteration order is not semantic.
    std::map<IVec3, std::optional<std::vector<Point>>> overflow;

    bool add_point(const Point& point, Point* loser);                      // cell.rs:70-94
    std::vector<Point> add_points(const std::vector<Point>& points);        // cell.rs:96-106
    std::unordered_map<IVec3, std::vector<Point>, IVec3Hash> add_points_in_overflow(
        std::unordered_map<IVec3, std::vector<Point>, IVec3Hash>&& overflow_points,
        const MetadataConfig& config);                                      // cell.rs:108-153
    void write_to(std::vector<uint8_t>& out) const;                         // cell.rs:155-181
    static bool read_from(const uint8_t* data, size_t len, Cell* out);      // cell.rs:183-229
};

using PointMap = std::unordered_map<IVec3, std::vector<Point>, IVec3Hash>;

PointMap group_points(const std::vector<Point>& points, uint32_t hierarchy,
                      const MetadataConfig& config);  // converter.rs:32-47
PointMap add_points_to_cell(const MetadataConfig& config, const std::vector<Point>& points,
                            Cell& cell);  // converter.rs:62-70

std::string cell_path(const CellId& id);                       // cell.rs:20-31
std::string metadata_to_json(const Metadata& m);               // metadata.rs:51-53
bool metadata_from_json(const std::string& text, Metadata* m); // metadata.rs:55-57
std::string format_f32(float v);                               // ryu 1.0.17 pretty format

// point-converter/src/converter.rs:72-246.  All cells are kept in RAM (the
// reference's LRU(100) + disk spill is semantically transparent);
// cells missing from RAM are looked up on disk first, exactly like
// load_or_create_cell (converter.rs:187-207).
struct Converter {
    Metadata metadata;
    std::string working_directory;  // empty => no file IO at all
    std::unordered_map<CellId, std::unique_ptr<Cell>, CellIdHash> cells;
    uint64_t arrivals = 0;          // statistic: sum over levels of delivered points
    // the GUI caller's accounting (src/plugins/converter.rs:565-573): sum over add_points_to_cell calls of
    // header.total_number_of_points after - before (an i32 per call there)
    int64_t points_by_delta = 0;
    int error = 0;                  // 1 = hierarchy >= 32 (reference would overflow 2u32.pow)
    // Optional residency emulation of converter.rs:92, 160-178: at most lru_capacity cells in RAM (the reference
    // uses LRUCache::with_hasher(100, ..)); a miss loads the node file or creates the cell, the evicted cell is
    // written out whole (save_cell).  0 = keep everything in RAM (the default; results are identical either way).
    size_t lru_capacity = 0;
    std::list<CellId> lru_order;    // front = most recently used
    std::unordered_map<CellId, std::list<CellId>::iterator, CellIdHash> lru_pos;
    uint64_t lru_evictions = 0, lru_loads = 0;

    void update_bounding_box(const std::vector<Point>& points);  // converter.rs:96-104
    void add_points_batch(const std::vector<Point>& points);     // converter.rs:106-112
    void add_points_in_hierarchy(uint32_t hierarchy, const MetadataConfig& config,
                                 PointMap&& grouped);            // converter.rs:114-139
    void create_hierarchy_folder(uint32_t hierarchy);            // converter.rs:141-158
    Cell& get_cell_mut(const CellId& id);                        // converter.rs:160-207
    bool save_all();                                             // converter.rs:218-238 (Drop)
};

}  // namespace orc

// pc_oracle.cpp — TEST INFRASTRUCTURE ONLY (see pc_oracle.hpp header comment).
// PARITY UNPINNED by reference tests; pinned by KATs + oracle/model.py.
#include "pc_oracle.hpp"

#include <sys/stat.h>
#include <sys/types.h>

#include <algorithm>
#include <cerrno>
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <thread>

namespace orc {

// ---------------------------------------------------------------------------
// cell.rs:70-94 — hex-grid contest for one point.  Returns true when a point
// leaves (written to *loser): the displaced incumbent when the newcomer is
// STRICTLY closer to the prism's reference centre, else the newcomer itself.
bool Cell::add_point(const Point& point, Point* loser) {
    OffsetIndex index = header.sub_grid_index_for_point(point);
    auto it = points_grid.find(index);
    if (it != points_grid.end()) {
        Vec3 pos = offset_to_world(index, header.sub_cell_size / 2.0f);
        float old_distance = distance_squared(pos, Vec3{it->second.x, it->second.y, it->second.z});
        float new_distance = distance_squared(pos, Vec3{point.x, point.y, point.z});
        if (new_distance < old_distance) {
            *loser = it->second;
            it->second = point;
        } else {
            *loser = point;
        }
        return true;
    }
    points_grid.emplace(index, point);
    header.total_number_of_points += 1;
    header.number_of_points += 1;
    return false;
}

// cell.rs:96-106 — losers leave in trigger order.
std::vector<Point> Cell::add_points(const std::vector<Point>& points) {
    std::vector<Point> overflow_points;
    overflow_points.reserve(points.size());
    Point loser;
    for (const Point& p : points)
        if (add_point(p, &loser)) overflow_points.push_back(loser);
    return overflow_points;
}

// cell.rs:108-153 — per child-cell buffer / flush state machine.
PointMap Cell::add_points_in_overflow(PointMap&& overflow_points, const MetadataConfig& config) {
    PointMap remaining;
    const size_t limit = config.cell_point_overflow_limit;
    for (auto& kv : overflow_points) {
        const IVec3& cell_index = kv.first;
        std::vector<Point>& points = kv.second;
        auto it = overflow.find(cell_index);
        if (it == overflow.end()) {  // Entry::Vacant, cell.rs:117-126
            if (points.size() <= limit) {
                header.total_number_of_points += (uint32_t)points.size();
                header.number_of_overflow_points += (uint32_t)points.size();
                overflow.emplace(cell_index, std::move(points));
            } else {
                remaining.emplace(cell_index, std::move(points));
                overflow.emplace(cell_index, std::nullopt);
            }
        } else if (!it->second.has_value()) {  // None, cell.rs:128-130
            remaining.emplace(cell_index, std::move(points));
        } else {  // Some(cell_points), cell.rs:131-147
            std::vector<Point>& cell_points = *it->second;
            uint32_t cell_points_len = (uint32_t)cell_points.size();
            uint32_t points_len = (uint32_t)points.size();
            cell_points.insert(cell_points.end(), points.begin(), points.end());
            if (cell_points.size() < limit) {
                header.total_number_of_points += points_len;
                header.number_of_overflow_points += points_len;
            } else {
                header.total_number_of_points -= cell_points_len;
                header.number_of_overflow_points -= cell_points_len;
                remaining.emplace(cell_index, std::move(cell_points));
                it->second = std::nullopt;
            }
        }
    }
    return remaining;
}

// ---------------------------------------------------------------------------
// Little-endian codec helpers (lib.rs:1 — Endianess = LittleEndian)
static void put_u32(std::vector<uint8_t>& o, uint32_t v) {
    for (int i = 0; i < 4; i++) o.push_back((uint8_t)(v >> (8 * i)));
}
static void put_i32(std::vector<uint8_t>& o, int32_t v) { put_u32(o, (uint32_t)v); }
static void put_f32(std::vector<uint8_t>& o, float f) {
    uint32_t v;
    std::memcpy(&v, &f, 4);
    put_u32(o, v);
}
// point.rs:25-37
static void put_point(std::vector<uint8_t>& o, const Point& p) {
    put_f32(o, p.x);
    put_f32(o, p.y);
    put_f32(o, p.z);
    o.push_back(p.color[0]);
    o.push_back(p.color[1]);
    o.push_back(p.color[2]);
    o.push_back(p.color[3]);
}
struct Rd {
    const uint8_t* p;
    size_t n, off = 0;
    bool ok = true;
    uint32_t u32() {
        if (off + 4 > n) { ok = false; return 0; }
        uint32_t v = (uint32_t)p[off] | ((uint32_t)p[off + 1] << 8) | ((uint32_t)p[off + 2] << 16) |
                     ((uint32_t)p[off + 3] << 24);
        off += 4;
        return v;
    }
    int32_t i32() { return (int32_t)u32(); }
    float f32() {
        uint32_t v = u32();
        float f;
        std::memcpy(&f, &v, 4);
        return f;
    }
    uint8_t u8() {
        if (off + 1 > n) { ok = false; return 0; }
        return p[off++];
    }
    // point.rs:39-53
    Point point() {
        Point q;
        q.x = f32();
        q.y = f32();
        q.z = f32();
        q.color[0] = u8();
        q.color[1] = u8();
        q.color[2] = u8();
        q.color[3] = u8();
        return q;
    }
};

// cell.rs:155-181 + Header::write_to cell.rs:280-298.  The reference writes the
// grid section and the overflow entries in HashMap iteration order (an artefact);
// the oracle writes the grid sorted by hex key (z,y,x) and entries sorted by
// child index.  Order INSIDE an overflow list is semantic and preserved.
void Cell::write_to(std::vector<uint8_t>& o) const {
    put_u32(o, header.id.hierarchy);
    put_i32(o, header.id.index.x);
    put_i32(o, header.id.index.y);
    put_i32(o, header.id.index.z);
    put_u32(o, header.total_number_of_points);
    put_u32(o, header.number_of_points);
    put_u32(o, header.number_of_overflow_points);
    put_f32(o, header.size);
    put_f32(o, header.sub_cell_size);
    put_f32(o, header.pos.x);
    put_f32(o, header.pos.y);
    put_f32(o, header.pos.z);

    std::vector<std::pair<OffsetIndex, const Point*>> grid;
    grid.reserve(points_grid.size());
    for (auto& kv : points_grid) grid.emplace_back(kv.first, &kv.second);
    std::sort(grid.begin(), grid.end(), [](auto& a, auto& b) { return a.first < b.first; });
    for (auto& g : grid) put_point(o, *g.second);

    o.push_back((uint8_t)overflow.size());  // cell.rs:162
    for (auto& kv : overflow) {
        put_i32(o, kv.first.x);
        put_i32(o, kv.first.y);
        put_i32(o, kv.first.z);
        if (kv.second.has_value()) {
            put_u32(o, (uint32_t)kv.second->size());
            for (const Point& p : *kv.second) put_point(o, p);
        } else {
            put_u32(o, 0);
        }
    }
}

// cell.rs:183-229 + Header::read_from cell.rs:300-335
bool Cell::read_from(const uint8_t* data, size_t len, Cell* out) {
    Rd r{data, len};
    Header& h = out->header;
    h.id.hierarchy = r.u32();
    h.id.index.x = r.i32();
    h.id.index.y = r.i32();
    h.id.index.z = r.i32();
    h.total_number_of_points = r.u32();
    h.number_of_points = r.u32();
    h.number_of_overflow_points = r.u32();
    h.size = r.f32();
    h.sub_cell_size = r.f32();
    h.pos.x = r.f32();
    h.pos.y = r.f32();
    h.pos.z = r.f32();
    if (!r.ok) return false;
    out->points_grid.clear();
    out->points_grid.reserve(h.number_of_points);
    for (uint32_t i = 0; i < h.number_of_points; i++) {
        Point p = r.point();
        if (!r.ok) return false;
        out->points_grid[h.sub_grid_index_for_point(p)] = p;  // cell.rs:192-194 (insert overrides)
    }
    size_t overflow_len = r.u8();
    if (!r.ok) return false;
    out->overflow.clear();
    for (size_t e = 0; e < overflow_len; e++) {
        IVec3 key{r.i32(), r.i32(), r.i32()};
        // NB: evaluation order of the braced list is left to right (x, y, z).
        uint32_t n = r.u32();
        if (!r.ok) return false;
        if (n == 0) {
            out->overflow[key] = std::nullopt;  // cell.rs:210-211
        } else {
            std::vector<Point> v;
            v.reserve(n);
            for (uint32_t i = 0; i < n; i++) v.push_back(r.point());
            if (!r.ok) return false;
            out->overflow[key] = std::move(v);
        }
    }
    return true;
}

// ---------------------------------------------------------------------------
// converter.rs:32-47 — order inside each group is input order.
PointMap group_points(const std::vector<Point>& points, uint32_t hierarchy, const MetadataConfig& config) {
    float cell_size = config.cell_size(hierarchy);
    PointMap map;
    for (const Point& p : points) map[config.cell_index(Vec3{p.x, p.y, p.z}, cell_size)].push_back(p);
    return map;
}

// converter.rs:62-70
PointMap add_points_to_cell(const MetadataConfig& config, const std::vector<Point>& points, Cell& cell) {
    std::vector<Point> overflow_points = cell.add_points(points);
    PointMap grouped = group_points(overflow_points, cell.header.id.hierarchy + 1, config);
    return cell.add_points_in_overflow(std::move(grouped), config);
}

// cell.rs:20-31, metadata.rs:47-49
std::string cell_path(const CellId& id) {
    char buf[128];
    std::snprintf(buf, sizeof buf, "h_%u/c_%d_%d_%d.bin", id.hierarchy, id.index.x, id.index.y, id.index.z);
    return buf;
}

}  // namespace orc

// cull_oracle.cpp - CPU restatement of the renderer's cull / LOD compute shader.  TEST INFRASTRUCTURE ONLY: only tests/
// may link or call it; the product (libpcb.so) never does.
//
// Follows src/plugins/render/point/pipelines/compute.wgsl (paths relative to /root/reference), function by function:
//   cell_index :81-83, get_hierarchy :85-99, search_smallest_hierarchy :101-126, binary_search :128-158,
//   unpack4x8 / pack4x8 :160-171, main :174-205.
// The shader cannot be run here (no graphics stack on this box: "unavailable on this box"), so this restatement is not
// pinned by a reference run: PARITY UNPINNED.  Arithmetic: one IEEE binary32 operation per WGSL operation, in source
// order, compiled with -ffp-contract=off (mat4x4 * vec4 as the left-to-right linear combination of the columns).
// Output: the kept points in INPUT order (the shader's order depends on the atomic; tests compare sets).
#include <cmath>
#include <cstdint>
#include <cstring>

namespace {
struct Point {
    float x, y, z;
    uint8_t rgba[4];
};
struct V4 {
    float x, y, z, w;
};
V4 mat_vec(const float* m, V4 v) {   // column-major m[4 * col + row]
    V4 r;
    r.x = ((m[0] * v.x + m[4] * v.y) + m[8] * v.z) + m[12] * v.w;
    r.y = ((m[1] * v.x + m[5] * v.y) + m[9] * v.z) + m[13] * v.w;
    r.z = ((m[2] * v.x + m[6] * v.y) + m[10] * v.z) + m[14] * v.w;
    r.w = ((m[3] * v.x + m[7] * v.y) + m[11] * v.z) + m[15] * v.w;
    return r;
}
int32_t to_i32(float v) {   // WGSL i32(f32): truncation, saturating; NaN -> 0
    if (!(v == v)) return 0;
    if (v >= 2147483648.0f) return INT32_MAX;
    if (v <= -2147483648.0f) return INT32_MIN;
    return (int32_t)v;
}
uint32_t to_u32(float v) {
    if (!(v == v) || v <= 0.0f) return 0;
    if (v >= 4294967296.0f) return UINT32_MAX;
    return (uint32_t)v;
}
struct Cell {
    int32_t h, x, y, z;
};
bool binary_search(const Cell* cells, uint32_t len, Cell t) {   // :128-158
    int low = 0, high = (int)len - 1;
    while (low <= high) {
        int mid = (low + high) / 2;
        const Cell& m = cells[mid];
        if (m.h == t.h && m.x == t.x && m.y == t.y && m.z == t.z) return true;
        if (m.h < t.h || (m.h == t.h && m.x < t.x) || (m.h == t.h && m.x == t.x && m.y < t.y) ||
            (m.h == t.h && m.x == t.x && m.y == t.y && m.z < t.z))
            low = mid + 1;
        else
            high = mid - 1;
    }
    return false;
}
}  // namespace

extern "C" {
struct orc_cull_params {   // same layout as pcb_cull_params (include/pcb.h)
    float view[16], projection[16];
    float cam_pos[3];
    uint32_t depth_width, depth_height;
    uint32_t number_of_hierarchies;
    uint32_t size_by_distance, max_hierarchy;
    uint32_t cell_hierarchy;
};

// returns the number of points kept; out (capacity n) receives them in input order
uint64_t orc_cull_lod(const orc_cull_params* p, const float* hier /* nh x (cell_size, spacing) */, const int32_t* loaded,
                      uint32_t n_loaded, const float* far_dist, const float* depth, const Point* in, uint64_t n, Point* out) {
    const Cell* cells = (const Cell*)loaded;
    const uint32_t nh = p->number_of_hierarchies;
    uint64_t kept = 0;
    for (uint64_t i = 0; i < n; i++) {
        const Point pt = in[i];
        const V4 view = mat_vec(p->view, V4{pt.x, pt.y, pt.z, 1.0f});        // :183
        const V4 clip = mat_vec(p->projection, view);                        // :184
        const float nx = clip.x / clip.w, ny = clip.y / clip.w, nz = clip.z / clip.w;   // :185
        if (!(std::fabs(nx) <= 1.0f && std::fabs(ny) <= 1.0f && std::fabs(nz - 0.5f) <= 0.5f)) continue;   // :186
        const float fu = (nx * 0.5f + 0.5f) * (float)p->depth_width;         // :187
        const float fv = (ny * -0.5f + 0.5f) * (float)p->depth_height;
        uint32_t u = to_u32(fu), v = to_u32(fv);
        u = u < p->depth_width ? u : p->depth_width - 1;   // out-of-range textureLoad: clamped (see cull.cu)
        v = v < p->depth_height ? v : p->depth_height - 1;
        const float d = depth[(size_t)v * p->depth_width + u];               // :188
        uint32_t own;                                                        // search_smallest_hierarchy, :101-126
        if (p->cell_hierarchy >= nh - 1) {
            own = nh - 1;
        } else {
            uint32_t h = p->cell_hierarchy;
            for (;;) {
                h += 1;
                if (h >= nh) break;
                const float cs = hier[2 * h];
                Cell t{(int32_t)h, to_i32(std::floor(pt.x / cs)), to_i32(std::floor(pt.y / cs)), to_i32(std::floor(pt.z / cs))};
                if (!binary_search(cells, n_loaded, t)) break;
            }
            own = h - 1;
        }
        uint32_t hsel = own;                                                 // get_hierarchy, :85-99
        if (p->size_by_distance) {
            const float dx = p->cam_pos[0] - pt.x, dy = p->cam_pos[1] - pt.y, dz = p->cam_pos[2] - pt.z;
            const float dist = std::sqrt((dx * dx + dy * dy) + dz * dz);
            for (uint32_t k = p->max_hierarchy; k > own; k--)
                if (dist < far_dist[k]) { hsel = k; break; }
        }
        const float radius = hier[2 * hsel + 1];                             // :191
        const V4 mclip = mat_vec(p->projection, V4{view.x, view.y, view.z + radius, view.w});   // :193
        const float mz = mclip.z / mclip.w;                                  // :194
        const float radius_z = nz - mz;                                      // :196
        if (mz < d || (mz - d) < radius_z * 3.0f) {                          // :198
            Point o = pt;
            o.rgba[3] = (uint8_t)hsel;                                       // :201-202
            out[kept++] = o;
        }
    }
    return kept;
}
}

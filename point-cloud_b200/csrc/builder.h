// builder.h — host-side interface of the device builder (plain C++, no CUDA types), used by the C ABI
// (api.cpp).  Mirrors the reference's Converter (point-converter/src/converter.rs:72-246):
//   Builder::add_points  ~ Converter::add_points_batch, one call per get_batch (lib.rs:31-35)
//   Builder::build       ~ everything add_points_in_hierarchy did for those batches, level by level
#pragma once
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/pcb.h"

namespace pcb {

// point-converter/src/metadata.rs:9-28, 67-78
struct Metadata {
    std::string version = "1.0";
    std::string name = "Unknown";
    uint64_t number_of_points = 0;
    uint32_t hierarchies = 0;
    float bbox_min[3] = {0, 0, 0}, bbox_max[3] = {0, 0, 0};
    uint32_t cell_point_overflow_limit = 5000;
    uint32_t sub_grid_dimension = 96;
    float max_cell_size = 1000.0f;
};
std::string metadata_to_json(const Metadata& m);                 // metadata.rs:51-53 (serde_json pretty)
bool metadata_from_json(const std::string& text, Metadata* m);   // metadata.rs:55-57
std::string format_f32(float v);                                 // ryu / serde_json float text

// State of one node file read back for merge mode (cell.rs:183-229).
struct ExistingNode {
    std::vector<pcb_point> grid;
    struct Entry {
        int32_t child[3];
        bool none;                     // count == 0 marker: child already flushed
        std::vector<pcb_point> list;
    };
    std::vector<Entry> entries;
};
bool parse_node(const uint8_t* data, size_t len, ExistingNode* out);
std::string node_rel_path(uint32_t hierarchy, const int32_t idx[3]);  // cell.rs:20-31

// Supplies existing nodes in merge mode; returns false when the node does not exist (converter.rs:187-207).
struct NodeSource {
    virtual ~NodeSource() {}
    virtual bool load(uint32_t hierarchy, const int32_t idx[3], ExistingNode* out) = 0;
};

struct NodeRec {
    uint32_t hierarchy;
    int32_t index[3];
    uint64_t offset;  // into the level image
    uint64_t size;
    uint32_t level_slot;  // which level image
};

struct LevelImage {
    uint32_t hierarchy = 0;
    uint64_t bytes = 0;
    unsigned char* d_image = nullptr;  // device copy (kept when PCB_FLAG_DEVICE_ONLY)
    unsigned char* h_image = nullptr;  // pinned host copy (otherwise)
    uint64_t h_cap = 0;
};

struct KernelStat {
    std::string name;
    uint64_t launches = 0, units = 0, bytes = 0;
    double ms = 0;
};

// arrivals of the first level a partial build did not build (device pointers, valid while the Builder lives)
struct PendingArrivals {
    uint32_t level = 0;
    uint64_t n = 0, ncell = 0;
    const pcb_point* d_pts = nullptr;
    const uint32_t* d_bat = nullptr;          // batch ordinal + 1
    const uint32_t* d_cell_start = nullptr;   // [ncell + 1]
    const int32_t* d_cell_idx = nullptr;      // [3 * ncell]
    double prism_load = 0;                    // arrivals per occupied prism on the last level built (record-filter hint)
};

class BuilderImpl;
class Builder {
   public:
    Builder(int device, uint32_t flags);
    ~Builder();
    void set_config(uint32_t limit, uint32_t dim, float max_cell_size);
    void reserve(uint64_t n);
    void set_stream(void* cuda_stream);  // run on the caller's stream (e.g. torch's current stream)
    void begin_file();                                        // lib.rs:17-32: batches never span files
    void add_points(const pcb_point* pts, uint64_t n, uint32_t batch_size, bool device_ptr);
    void add_points_one_batch(const pcb_point* pts, uint64_t n, bool device_ptr);  // converter.rs:106
    void add_points_batched(const pcb_point* d_pts, const uint32_t* d_batch, uint64_t n);  // explicit batch ordinals
    // arrivals of one level in (cell, tau) order with their CSR (multi-GPU); device pointers, borrowed; the only add
    void add_arrivals(uint32_t level, const pcb_point* d_pts, const uint32_t* d_bat_plus1, uint64_t n,
                      const uint32_t* d_cell_start, const int32_t* d_cell_idx, uint64_t ncell, const float* bbox6,
                      uint64_t count_points);
    void set_stop_level(uint32_t level);   // build() stops before this level; its arrivals stay on the device
    void set_prism_load(double v);         // hint that travels with pending arrivals (see PendingArrivals)
    PendingArrivals pending() const;
    // uncompressed LAS point records (host bytes), decoded on the device (converter/las.rs:32-41)
    void add_las_records(const uint8_t* records, uint64_t n, uint32_t stride, int32_t colour_off, const double scale[3],
                         const double offset[3], uint32_t batch_size);
    static void* alloc_host(uint64_t bytes, uint64_t* got);   // pinned staging from the process-wide cache
    static void free_host(void* p, uint64_t bytes);
    // Runs the level-synchronous build.  meta: in = state before this run, out = after.
    void build(Metadata* meta, NodeSource* existing);
    const std::vector<NodeRec>& nodes() const;
    const std::vector<LevelImage>& levels() const;
    std::vector<KernelStat> kernel_stats() const;
    uint64_t points_added() const;
    uint64_t arrivals() const;
    uint64_t launches() const;
    uint32_t levels_built() const;
    double h2d_seconds() const;
    double build_seconds() const;
    double d2h_seconds() const;

   private:
    BuilderImpl* impl_;
};

struct Error {
    int status;
    std::string message;
};

}  // namespace pcb

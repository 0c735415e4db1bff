// api.cpp — the C ABI of include/pcb.h over pcb::Builder.  No exception crosses the boundary.
// Reference seam being replaced: load_metadata + Converter::new / add_points_batch / Drop
// (point-converter/src/lib.rs:11-60, 86-101; converter.rs:78-112, 218-246).
#include <errno.h>
#include <sys/stat.h>
#include <sys/types.h>

#include <algorithm>
#include <atomic>
#include <map>
#include <memory>
#include <mutex>
#include <vector>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <thread>

#include "builder.h"

using namespace pcb;

namespace pcb {
int device_count();
void trim_host_cache();
int route_bbox(int device, void* stream, const pcb_point* d_pts, uint64_t n, float S0, float out_min[3], float out_max[3]);
int route_histogram(int device, void* stream, const pcb_point* d_pts, uint64_t n, float S0, const int32_t origin[3],
                    const uint32_t dims[3], uint32_t* d_hist);
int route_pack(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index, uint32_t batch_size,
               float S0, const int32_t origin[3], const uint32_t dims[3], const uint8_t* d_owner, uint32_t world,
               pcb_point* d_send_pts, uint32_t* d_send_batch, uint64_t* send_counts);
int route_scatter_cells(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index, uint32_t batch_size,
                        float S0, const int32_t origin[3], const uint32_t dims[3], const uint8_t* d_owner,
                        const uint32_t* d_dst_base, uint32_t world, const uint64_t* recv_pts_ptrs,
                        const uint64_t* recv_batch_ptrs, uint32_t* d_err);
int gather_cells(int device, void* stream, const pcb_point* d_pts, const uint32_t* d_bat, const uint32_t* d_cell_start,
                 const uint32_t* d_sel, const uint32_t* d_out_start, uint32_t nsel, uint64_t total, pcb_point* d_out_pts,
                 uint32_t* d_out_bat);
void plan_owners(const uint64_t* counts, uint64_t ncells, uint32_t world, uint8_t* owner);
int route_bbox_static(int device, void* stream, const pcb_point* d_pts, uint64_t n, float S0, float* d_out6, uint32_t* d_err);
int route_scatter(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index, uint32_t batch_size,
                  float S0, const int32_t origin[3], const uint32_t dims[3], const uint8_t* d_owner, uint32_t world,
                  const uint64_t* recv_pts_ptrs, const uint64_t* recv_batch_ptrs, const uint64_t* recv_offsets);
}

struct pcb_handle {
    std::unique_ptr<Builder> builder;
    Metadata meta;
    std::string out_dir;
    pcb_options opt;
    bool merge = false;     // metadata.json existed at open (lib.rs:92-95)
    bool finished = false;
    std::string err;
    std::string meta_json;
    std::vector<KernelStat> kstats;
};

static thread_local std::string g_open_error;

static int fail(pcb_handle* h, int status, const std::string& msg) {
    if (h) h->err = msg;
    else g_open_error = msg;
    return status;
}

// create_dir_all (converter.rs:80)
static bool mkdir_p(const std::string& path) {
    if (path.empty()) return true;
    std::string cur;
    for (size_t i = 0; i <= path.size(); i++) {
        if (i == path.size() || path[i] == '/') {
            if (!cur.empty() && cur != "/" && ::mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return false;
        }
        if (i < path.size()) cur += path[i];
    }
    return true;
}

static bool read_file(const std::string& path, std::vector<uint8_t>* out, bool* not_found) {
    *not_found = false;
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) {
        *not_found = (errno == ENOENT);
        return false;
    }
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out->resize(n > 0 ? (size_t)n : 0);
    size_t got = n > 0 ? std::fread(out->data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == out->size();
}

static bool write_file(const std::string& path, const void* data, size_t n) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    size_t put = n ? std::fwrite(data, 1, n, f) : 0;
    return (std::fclose(f) == 0) && put == n;
}

// converter.rs:180-207: a missing file means a new cell; any other failure is fatal (reference panics)
struct DirNodeSource : NodeSource {
    std::string dir;
    std::atomic<uint64_t> total_before{0};   // header totals of the nodes loaded (PCB_FLAG_COUNT_BY_DELTA)
    bool load(uint32_t hierarchy, const int32_t idx[3], ExistingNode* out) override {
        std::vector<uint8_t> bytes;
        bool nf = false;
        std::string p = dir + "/" + node_rel_path(hierarchy, idx);
        if (!read_file(p, &bytes, &nf)) {
            if (nf) return false;
            throw Error{PCB_ERR_IO, "cannot read node file " + p};
        }
        if (!parse_node(bytes.data(), bytes.size(), out)) throw Error{PCB_ERR_FORMAT, "corrupt node file " + p};
        if (bytes.size() >= 20) {   // Header.total_number_of_points (cell.rs:286)
            uint32_t t;
            std::memcpy(&t, bytes.data() + 16, 4);
            total_before += t;
        }
        return true;
    }
};

extern "C" {

const char* pcb_version(void) { return "point-cloud-b200 0.1 (sm_100a)"; }
int pcb_device_count(void) { return pcb::device_count(); }

int pcb_open(const char* out_dir, const pcb_options* opt, pcb_handle** out) {
    if (!out) return PCB_ERR_ARG;
    *out = nullptr;
    std::unique_ptr<pcb_handle> h(new pcb_handle());
    std::memset(&h->opt, 0, sizeof h->opt);
    if (opt) std::memcpy(&h->opt, opt, std::min<size_t>(sizeof h->opt, opt->struct_size ? opt->struct_size : sizeof h->opt));
    if ((h->opt.flags & PCB_FLAG_DEVICE_ONLY) && (h->opt.flags & PCB_FLAG_COUNT_BY_DELTA))
        return fail(nullptr, PCB_ERR_ARG, "PCB_FLAG_COUNT_BY_DELTA needs the node images on the host (not with PCB_FLAG_DEVICE_ONLY)");
    if (h->opt.flags & PCB_FLAG_DEVICE_ONLY) h->opt.flags |= PCB_FLAG_NO_FILES;
    h->out_dir = out_dir ? out_dir : "";
    if (h->out_dir.empty()) h->opt.flags |= PCB_FLAG_NO_FILES;
    if (!h->out_dir.empty()) {
        if (!(h->opt.flags & PCB_FLAG_NO_FILES) && !mkdir_p(h->out_dir))
            return fail(nullptr, PCB_ERR_IO, "cannot create output directory " + h->out_dir);
        std::vector<uint8_t> bytes;
        bool nf = false;
        if (read_file(h->out_dir + "/metadata.json", &bytes, &nf)) {  // lib.rs:87-95
            if (!metadata_from_json(std::string(bytes.begin(), bytes.end()), &h->meta))
                return fail(nullptr, PCB_ERR_FORMAT, "unreadable metadata.json");
            h->merge = true;
        }
    }
    if (!h->merge) {  // Metadata::default() (metadata.rs:30-41), optionally with a test-supplied config
        // every field by itself: a zero field keeps its reference default (5000 / 96 / 1000.0, metadata.rs:80-88)
        if (h->opt.cell_point_overflow_limit) h->meta.cell_point_overflow_limit = h->opt.cell_point_overflow_limit;
        if (h->opt.sub_grid_dimension) h->meta.sub_grid_dimension = h->opt.sub_grid_dimension;
        if (h->opt.max_cell_size != 0.0f) h->meta.max_cell_size = h->opt.max_cell_size;
    }
    // the config is validated here, not when the build starts
    if (h->meta.sub_grid_dimension == 0 || !(h->meta.max_cell_size > 0.0f) || !std::isfinite(h->meta.max_cell_size))
        return fail(nullptr, PCB_ERR_ARG, "bad config (sub_grid_dimension / max_cell_size)");
    try {
        h->builder.reset(new Builder(h->opt.device, h->opt.flags));
    } catch (const Error& e) {
        return fail(nullptr, e.status, e.message);
    } catch (const std::exception& e) {
        return fail(nullptr, PCB_ERR_NOMEM, e.what());
    }
    *out = h.release();
    return PCB_OK;
}

const char* pcb_last_error(const pcb_handle* h) { return h ? h->err.c_str() : g_open_error.c_str(); }

#define PCB_GUARD(h, body)                                 \
    if (!(h)) return PCB_ERR_ARG;                          \
    try {                                                  \
        body;                                              \
    } catch (const Error& e) {                             \
        return fail((h), e.status, e.message);             \
    } catch (const std::bad_alloc&) {                      \
        return fail((h), PCB_ERR_NOMEM, "out of host memory"); \
    } catch (const std::exception& e) {                    \
        return fail((h), PCB_ERR_ARG, e.what());           \
    }                                                      \
    return PCB_OK;

int pcb_reserve(pcb_handle* h, uint64_t n) { PCB_GUARD(h, h->builder->reserve(n)) }

int pcb_set_stream(pcb_handle* h, void* cuda_stream) { PCB_GUARD(h, h->builder->set_stream(cuda_stream)) }

int pcb_begin_file(pcb_handle* h) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    PCB_GUARD(h, h->builder->begin_file())
}

int pcb_add_points(pcb_handle* h, const pcb_point* pts, uint64_t n) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    if (h && n && !pts) return fail(h, PCB_ERR_ARG, "null points");
    PCB_GUARD(h, h->builder->add_points(pts, n, h->opt.batch_size, false))
}

int pcb_add_las_records(pcb_handle* h, const uint8_t* records, uint64_t n, uint32_t stride, int32_t colour_off,
                        const double scale[3], const double offset[3]) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    if (h && ((n && !records) || !scale || !offset)) return fail(h, PCB_ERR_ARG, "null records");
    PCB_GUARD(h, h->builder->add_las_records(records, n, stride, colour_off, scale, offset, h->opt.batch_size))
}

// pinned chunks handed out to callers: pointer -> capacity (the cache wants it back)
static std::mutex g_host_mu;
static std::map<void*, uint64_t> g_host_chunks;
void* pcb_alloc_host(uint64_t bytes) {
    try {
        uint64_t got = 0;
        void* p = Builder::alloc_host(bytes ? bytes : 16, &got);
        std::lock_guard<std::mutex> g(g_host_mu);
        g_host_chunks[p] = got;
        return p;
    } catch (...) {
        return nullptr;
    }
}
void pcb_free_host(void* p) {
    if (!p) return;
    uint64_t got = 0;
    {
        std::lock_guard<std::mutex> g(g_host_mu);
        auto it = g_host_chunks.find(p);
        if (it == g_host_chunks.end()) return;
        got = it->second;
        g_host_chunks.erase(it);
    }
    Builder::free_host(p, got);
}

int pcb_add_points_device(pcb_handle* h, const pcb_point* d_pts, uint64_t n) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    if (h && n && !d_pts) return fail(h, PCB_ERR_ARG, "null points");
    PCB_GUARD(h, h->builder->add_points(d_pts, n, h->opt.batch_size, true))
}

int pcb_add_points_device_batched(pcb_handle* h, const pcb_point* d_pts, const uint32_t* d_batch, uint64_t n) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    if (h && n && (!d_pts || !d_batch)) return fail(h, PCB_ERR_ARG, "null points");
    PCB_GUARD(h, h->builder->add_points_batched(d_pts, d_batch, n))
}

int pcb_add_arrivals(pcb_handle* h, uint32_t level, const pcb_point* d_pts, const uint32_t* d_batch_plus1, uint64_t n,
                     const uint32_t* d_cell_start, const int32_t* d_cell_idx, uint64_t ncell, const float* bbox6,
                     uint64_t count_points) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    if (h && n && (!d_pts || !d_batch_plus1 || !d_cell_start || !d_cell_idx || !ncell)) return fail(h, PCB_ERR_ARG, "null arrivals");
    PCB_GUARD(h, h->builder->add_arrivals(level, d_pts, d_batch_plus1, n, d_cell_start, d_cell_idx, ncell, bbox6, count_points))
}

int pcb_set_stop_level(pcb_handle* h, uint32_t stop_level) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    PCB_GUARD(h, h->builder->set_stop_level(stop_level))
}

int pcb_set_prism_load(pcb_handle* h, float prism_load) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    PCB_GUARD(h, h->builder->set_prism_load(prism_load > 0.0f ? (double)prism_load : 0.0))
}

int pcb_pending_arrivals(const pcb_handle* h, pcb_arrivals* out) {
    if (!h || !out || !h->finished) return PCB_ERR_ARG;
    const PendingArrivals p = h->builder->pending();
    out->level = p.level;
    out->prism_load = (float)p.prism_load;
    out->n = p.n;
    out->ncell = p.ncell;
    out->d_pts = p.d_pts;
    out->d_batch_plus1 = p.d_bat;
    out->d_cell_start = p.d_cell_start;
    out->d_cell_idx = p.d_cell_idx;
    return PCB_OK;
}

int pcb_write_metadata(const char* out_dir, uint64_t number_of_points, uint32_t hierarchies, const float bbox_min[3],
                       const float bbox_max[3], uint32_t limit, uint32_t dim, float max_cell_size) {
    if (!out_dir || !bbox_min || !bbox_max) return PCB_ERR_ARG;
    Metadata m;
    m.number_of_points = number_of_points;
    m.hierarchies = hierarchies;
    for (int a = 0; a < 3; a++) { m.bbox_min[a] = bbox_min[a]; m.bbox_max[a] = bbox_max[a]; }
    m.cell_point_overflow_limit = limit;
    m.sub_grid_dimension = dim;
    m.max_cell_size = max_cell_size;
    std::string js = metadata_to_json(m);
    if (!mkdir_p(out_dir)) return PCB_ERR_IO;
    return write_file(std::string(out_dir) + "/metadata.json", js.data(), js.size()) ? PCB_OK : PCB_ERR_IO;
}

#define PCB_ROUTE_GUARD(body)                          \
    try {                                              \
        return body;                                   \
    } catch (const Error& e) {                         \
        return fail(nullptr, e.status, e.message);     \
    } catch (const std::exception& e) {                \
        return fail(nullptr, PCB_ERR_ARG, e.what());   \
    }

int pcb_route_bbox(int device, void* stream, const pcb_point* d_pts, uint64_t n, float max_cell_size, float out_min[3],
                   float out_max[3]) {
    if (!out_min || !out_max || (n && !d_pts)) return PCB_ERR_ARG;
    PCB_ROUTE_GUARD(pcb::route_bbox(device, stream, d_pts, n, max_cell_size, out_min, out_max))
}
int pcb_route_bbox_static(int device, void* stream, const pcb_point* d_pts, uint64_t n, float max_cell_size,
                          float* d_out6, uint32_t* d_err) {
    if (!d_out6 || (n && !d_pts)) return fail(nullptr, PCB_ERR_ARG, "pcb_route_bbox_static: null pointer");
    PCB_ROUTE_GUARD(pcb::route_bbox_static(device, stream, d_pts, n, max_cell_size, d_out6, d_err))
}
int pcb_route_histogram(int device, void* stream, const pcb_point* d_pts, uint64_t n, float max_cell_size,
                        const int32_t origin[3], const uint32_t dims[3], uint32_t* d_hist) {
    if (!origin || !dims || !d_hist || (n && !d_pts)) return PCB_ERR_ARG;
    PCB_ROUTE_GUARD(pcb::route_histogram(device, stream, d_pts, n, max_cell_size, origin, dims, d_hist))
}
int pcb_route_pack(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index,
                   uint32_t batch_size, float max_cell_size, const int32_t origin[3], const uint32_t dims[3],
                   const uint8_t* d_owner, uint32_t world, pcb_point* d_send_pts, uint32_t* d_send_batch,
                   uint64_t* send_counts) {
    if (!origin || !dims || !d_owner || !send_counts || (n && (!d_pts || !d_send_pts || !d_send_batch))) return PCB_ERR_ARG;
    PCB_ROUTE_GUARD(pcb::route_pack(device, stream, d_pts, n, first_index, batch_size, max_cell_size, origin, dims,
                                    d_owner, world, d_send_pts, d_send_batch, send_counts))
}

int pcb_route_scatter(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index,
                      uint32_t batch_size, float max_cell_size, const int32_t origin[3], const uint32_t dims[3],
                      const uint8_t* d_owner, uint32_t world, const uint64_t* recv_pts_ptrs,
                      const uint64_t* recv_batch_ptrs, const uint64_t* recv_offsets) {
    if (!origin || !dims || !d_owner || !recv_pts_ptrs || !recv_batch_ptrs || !recv_offsets || (n && !d_pts)) return PCB_ERR_ARG;
    PCB_ROUTE_GUARD(pcb::route_scatter(device, stream, d_pts, n, first_index, batch_size, max_cell_size, origin, dims,
                                       d_owner, world, recv_pts_ptrs, recv_batch_ptrs, recv_offsets))
}

int pcb_route_scatter_cells(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index,
                            uint32_t batch_size, float max_cell_size, const int32_t origin[3], const uint32_t dims[3],
                            const uint8_t* d_owner, const uint32_t* d_dst_base, uint32_t world,
                            const uint64_t* recv_pts_ptrs, const uint64_t* recv_batch_ptrs, uint32_t* d_err) {
    if (!origin || !dims || !d_owner || !d_dst_base || !recv_pts_ptrs || !recv_batch_ptrs || (n && !d_pts)) return PCB_ERR_ARG;
    PCB_ROUTE_GUARD(pcb::route_scatter_cells(device, stream, d_pts, n, first_index, batch_size, max_cell_size, origin, dims,
                                             d_owner, d_dst_base, world, recv_pts_ptrs, recv_batch_ptrs, d_err))
}

int pcb_gather_cells(int device, void* stream, const pcb_point* d_pts, const uint32_t* d_batch, const uint32_t* d_cell_start,
                     const uint32_t* d_sel, const uint32_t* d_out_start, uint32_t nsel, uint64_t total, pcb_point* d_out_pts,
                     uint32_t* d_out_batch) {
    if (nsel && total && (!d_pts || !d_batch || !d_cell_start || !d_sel || !d_out_start || !d_out_pts || !d_out_batch)) return PCB_ERR_ARG;
    PCB_ROUTE_GUARD(pcb::gather_cells(device, stream, d_pts, d_batch, d_cell_start, d_sel, d_out_start, nsel, total, d_out_pts,
                                      d_out_batch))
}

int pcb_plan_owners(const uint64_t* counts, uint64_t ncells, uint32_t world, uint8_t* owner) {
    if (!counts || !owner || world == 0 || world > 256) return PCB_ERR_ARG;
    try {
        pcb::plan_owners(counts, ncells, world, owner);
    } catch (const std::exception&) {
        return PCB_ERR_NOMEM;
    }
    return PCB_OK;
}

int pcb_get_config(const pcb_handle* h, uint32_t* limit, uint32_t* dim, float* max_cell_size) {
    if (!h) return PCB_ERR_ARG;
    if (limit) *limit = h->meta.cell_point_overflow_limit;
    if (dim) *dim = h->meta.sub_grid_dimension;
    if (max_cell_size) *max_cell_size = h->meta.max_cell_size;
    return PCB_OK;
}

int pcb_add_points_batch(pcb_handle* h, const pcb_point* pts, uint64_t n) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    if (h && n && !pts) return fail(h, PCB_ERR_ARG, "null points");
    PCB_GUARD(h, h->builder->add_points_one_batch(pts, n, false))
}

static void write_outputs(pcb_handle* h, pcb_stats* stats) {
    auto t0 = std::chrono::steady_clock::now();
    const auto& nodes = h->builder->nodes();
    const auto& levels = h->builder->levels();
    // converter.rs:141-158: h_<level> directories
    for (uint32_t l = 0; l < h->meta.hierarchies; l++) {
        std::string d = h->out_dir + "/h_" + std::to_string(l);
        if (::mkdir(d.c_str(), 0777) != 0 && errno != EEXIST) throw Error{PCB_ERR_IO, "cannot create " + d};
    }
    // converter.rs:209-224 save_cell for every node touched by this run
    std::atomic<size_t> next{0};
    std::atomic<bool> bad{false};
    unsigned nthreads = std::min<unsigned>(16, std::max<unsigned>(1, std::thread::hardware_concurrency()));
    if (nodes.size() < 64) nthreads = 1;
    auto worker = [&]() {
        for (;;) {
            size_t i = next.fetch_add(1);
            if (i >= nodes.size()) return;
            const NodeRec& r = nodes[i];
            const unsigned char* img = levels[r.level_slot].h_image + r.offset;
            if (!write_file(h->out_dir + "/" + node_rel_path(r.hierarchy, r.index), img, r.size)) bad = true;
        }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nthreads; t++) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    if (bad) throw Error{PCB_ERR_IO, "failed to write a node file"};
    // converter.rs:226-238 save_metadata
    if (!(h->opt.flags & PCB_FLAG_NO_METADATA_FILE) &&
        !write_file(h->out_dir + "/metadata.json", h->meta_json.data(), h->meta_json.size()))
        throw Error{PCB_ERR_IO, "failed to write metadata.json"};
    if (stats) stats->write_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

int pcb_finish(pcb_handle* h, pcb_stats* stats) {
    if (h && h->finished) return fail(h, PCB_ERR_ARG, "handle already finished");
    PCB_GUARD(h, {
        auto t0 = std::chrono::steady_clock::now();
        DirNodeSource src;
        src.dir = h->out_dir;
        const uint64_t points_before = h->meta.number_of_points;
        h->builder->build(&h->meta, h->merge ? &src : nullptr);
        h->finished = true;
        if (h->opt.flags & PCB_FLAG_COUNT_BY_DELTA) {
            // src/plugins/converter.rs:565-573: every add_points_to_cell call reports total_after - total_before of its
            // cell; over a run that telescopes to (header totals of the touched nodes now) - (their totals before)
            uint64_t after = 0;
            const auto& lv = h->builder->levels();
            for (auto& n : h->builder->nodes()) {
                uint32_t t;
                std::memcpy(&t, lv[n.level_slot].h_image + n.offset + 16, 4);
                after += t;
            }
            h->meta.number_of_points = points_before + after - src.total_before.load();
        }
        h->meta_json = metadata_to_json(h->meta);
        h->kstats = h->builder->kernel_stats();
        pcb_stats s;
        std::memset(&s, 0, sizeof s);
        if (!(h->opt.flags & PCB_FLAG_NO_FILES)) write_outputs(h, &s);
        s.number_of_points = h->meta.number_of_points;
        s.points_added = h->builder->points_added();
        s.arrivals = h->builder->arrivals();
        s.nodes_written = h->builder->nodes().size();
        for (auto& n : h->builder->nodes()) s.node_bytes += n.size;
        s.hierarchies = h->meta.hierarchies;
        s.levels_built = h->builder->levels_built();
        s.gpu_launches = h->builder->launches();
        s.h2d_seconds = h->builder->h2d_seconds();
        s.build_seconds = h->builder->build_seconds();
        s.d2h_seconds = h->builder->d2h_seconds();
        s.total_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        for (int a = 0; a < 3; a++) {
            s.bbox_min[a] = h->meta.bbox_min[a];
            s.bbox_max[a] = h->meta.bbox_max[a];
        }
        if (stats) *stats = s;
    })
}

void pcb_close(pcb_handle* h) { delete h; }

void pcb_trim_host_cache(void) { pcb::trim_host_cache(); }

uint64_t pcb_node_count(const pcb_handle* h) { return (h && h->finished) ? h->builder->nodes().size() : 0; }

int pcb_node_get(const pcb_handle* h, uint64_t i, pcb_node_info* info, const uint8_t** host_image) {
    if (!h || !h->finished || i >= h->builder->nodes().size()) return PCB_ERR_ARG;
    const NodeRec& r = h->builder->nodes()[i];
    if (info) {
        info->hierarchy = r.hierarchy;
        for (int a = 0; a < 3; a++) info->index[a] = r.index[a];
        info->offset = r.offset;
        info->size = r.size;
    }
    if (host_image) {
        const unsigned char* base = h->builder->levels()[r.level_slot].h_image;
        *host_image = base ? base + r.offset : nullptr;
    }
    return PCB_OK;
}

int pcb_metadata_json(const pcb_handle* h, char* buf, uint64_t cap) {
    if (!h || !h->finished) return PCB_ERR_ARG;
    if (buf && cap > h->meta_json.size()) std::memcpy(buf, h->meta_json.c_str(), h->meta_json.size() + 1);
    return (int)h->meta_json.size();
}

int pcb_kernel_stat_count(const pcb_handle* h) { return h ? (int)h->kstats.size() : 0; }

int pcb_kernel_stat_get(const pcb_handle* h, int i, pcb_kernel_stat* out) {
    if (!h || !out || i < 0 || i >= (int)h->kstats.size()) return PCB_ERR_ARG;
    const KernelStat& k = h->kstats[i];
    std::memset(out, 0, sizeof *out);
    std::snprintf(out->name, sizeof out->name, "%s", k.name.c_str());
    out->launches = k.launches;
    out->units = k.units;
    out->bytes = k.bytes;
    out->ms = k.ms;
    return PCB_OK;
}

}  // extern "C"

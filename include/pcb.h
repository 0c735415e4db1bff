/* pcb.h — C ABI of the B200-native octree/LOD builder (libpcb.so / libpcb.a).
 *
 * Drop-in boundary for point-converter's conversion path.  The reference has
 * no FFI; its seam is the Rust API below (paths relative to /root/reference):
 *
 *   load_metadata + Converter::new      point-converter/src/lib.rs:12-13, 86-101;
 *                                       converter.rs:79-94            -> pcb_open
 *   per input file (batches never span  point-converter/src/lib.rs:17-32
 *   files)                                                            -> pcb_begin_file
 *   Converter::add_points_batch         converter.rs:106-112 (called once per
 *                                       get_batch(10_000), lib.rs:32-34) -> pcb_add_points
 *   Drop for Converter (save_cache +    converter.rs:218-246           -> pcb_finish
 *   save_metadata)
 *
 * Plain pointers and sizes only; no exceptions cross the boundary; every
 * function returns 0 or a negative pcb_status.  One caller thread per handle
 * (the reference's Converter is not shared either).  There is NO CPU fallback:
 * without a CUDA device every compute entry point fails with PCB_ERR_CUDA.
 *
 * Semantics: results (node point sets, overflow lists in emission order,
 * headers, metadata.json) are bit-identical to the reference CLI fed the same
 * files in the same order.  The build is bulk and level-synchronous: points are
 * staged to the device by pcb_add_points and the octree is built in pcb_finish.
 */
#ifndef PCB_H
#define PCB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* point-converter/src/point.rs:8-14, 25-37 — 16-byte little-endian record. */
typedef struct pcb_point {
    float x, y, z;
    uint8_t rgba[4];
} pcb_point;

typedef struct pcb_handle pcb_handle;

typedef enum pcb_status {
    PCB_OK = 0,
    PCB_ERR_ARG = -1,      /* bad argument / call order */
    PCB_ERR_IO = -2,       /* file system error (reference: panic, converter.rs:84,153,171,203,243) */
    PCB_ERR_FORMAT = -3,   /* unreadable metadata.json / node file (reference: unwrap panic, lib.rs:94) */
    PCB_ERR_CUDA = -4,     /* CUDA error, or no device: there is no CPU fallback */
    PCB_ERR_RANGE = -5,    /* non-finite coordinate, level-0 index out of range, or octree deeper than
                              31 levels (reference overflows 2u32.pow(32), metadata.rs:92) */
    PCB_ERR_NOMEM = -6
} pcb_status;

/* flags for pcb_options.flags */
#define PCB_FLAG_NO_FILES 1u     /* do not touch the file system; node images stay in host memory and are
                                    read back with pcb_node_* (tests, bench e2e) */
#define PCB_FLAG_DEVICE_ONLY 2u  /* implies NO_FILES; node images stay in device memory (bench `value`) */
#define PCB_FLAG_PROFILE 4u      /* time every kernel class with CUDA events (pcb_kernel_stat) */
#define PCB_FLAG_NO_METADATA_FILE 8u /* write node files but not metadata.json (multi-GPU: every rank writes its own
                                    nodes into the shared directory, rank 0 writes the reduced metadata.json with
                                    pcb_write_metadata) */

#define PCB_FLAG_COUNT_BY_DELTA 16u /* GUI-caller accounting (src/plugins/converter.rs:565-573): number_of_points advances
                                    by the change of every touched node's header total_number_of_points (grid + kept
                                    overflow, negative when a list is flushed) instead of by batch lengths
                                    (converter.rs:108).  Together with batch_size = 50000 (src/plugins/converter.rs:600)
                                    this is the in-app converter's preset.  Not with PCB_FLAG_DEVICE_ONLY. */

typedef struct pcb_options {
    uint32_t struct_size;   /* sizeof(pcb_options), for ABI growth */
    int32_t device;         /* CUDA device ordinal */
    uint32_t batch_size;    /* 0 => 10000 (lib.rs:32).  GUI caller uses 50000 (src/plugins/converter.rs:600) */
    uint32_t flags;
    /* Config override for a FRESH output only (ignored when metadata.json exists).  All zero => reference
       defaults 5000 / 96 / 1000.0 (metadata.rs:80-88).  The reference CLI has no such flags; tests use
       this the way they would pre-seed a metadata.json. */
    uint32_t cell_point_overflow_limit;
    uint32_t sub_grid_dimension;
    float max_cell_size;
    uint32_t reserved;
} pcb_options;

typedef struct pcb_stats {
    uint64_t number_of_points;   /* metadata.number_of_points after this run */
    uint64_t points_added;       /* points staged through this handle */
    uint64_t arrivals;           /* sum over levels of points delivered to cells */
    uint64_t nodes_written;      /* node images produced by this run */
    uint64_t node_bytes;         /* total bytes of those images */
    uint32_t hierarchies;        /* metadata.hierarchies after this run */
    uint32_t levels_built;       /* levels visited by this run */
    uint64_t gpu_launches;       /* kernels launched by this handle */
    double h2d_seconds, build_seconds, d2h_seconds, write_seconds, total_seconds;
    float bbox_min[3], bbox_max[3];
} pcb_stats;

typedef struct pcb_kernel_stat {
    char name[32];
    uint64_t launches;
    uint64_t units;          /* elements processed, summed over launches */
    uint64_t bytes;          /* algorithmic bytes, summed over launches */
    double ms;               /* CUDA-event time, summed over launches (PCB_FLAG_PROFILE only) */
} pcb_kernel_stat;

typedef struct pcb_node_info {
    uint32_t hierarchy;
    int32_t index[3];
    uint64_t offset;         /* into the level image buffer */
    uint64_t size;           /* bytes of the node file image */
} pcb_node_info;

int pcb_open(const char* out_dir, const pcb_options* opt, pcb_handle** out);
int pcb_begin_file(pcb_handle* h);
/* Appends `n` points of the current input file; the stream is cut into batches of pcb_options.batch_size
   points exactly as lib.rs:31-35 does with get_batch(10_000).  Host pointer; copied before returning. */
int pcb_add_points(pcb_handle* h, const pcb_point* pts, uint64_t n);
/* Exactly ONE batch, the direct mirror of Converter::add_points_batch(Vec<Point>) (converter.rs:106-112;
   the GUI caller src/plugins/converter.rs:203 uses its own batch sizes).  n may be 0. */
int pcb_add_points_batch(pcb_handle* h, const pcb_point* pts, uint64_t n);
/* Uncompressed LAS point records of the current input file, as they lie in the file (`stride` bytes each, host
   memory): X, Y, Z as little-endian i32 at bytes 0-11, red / green / blue as u16 at colour_off (-1: none).  The
   records are copied to the device and decoded THERE exactly as converter/las.rs:32-41 does on the host:
   pos = (X * scale + offset) in f64, then `as f32`; colour = low byte of each channel; alpha 255.  Cut into batches
   like pcb_add_points.  stride 12..96. */
int pcb_add_las_records(pcb_handle* h, const uint8_t* records, uint64_t n, uint32_t stride, int32_t colour_off,
                        const double scale[3], const double offset[3]);
/* Page-locked host memory for input chunks (from the library's recycled cache; pcb_add_points from such a buffer is
   one DMA at full PCIe speed).  pcb_free_host returns it to the cache.  The streamed ingest of the CLI
   (lib.rs:31-52: one bounded batch of input at a time) uses three such chunks in rotation. */
void* pcb_alloc_host(uint64_t bytes);
void pcb_free_host(void* p);
/* Optional capacity hint (total points this handle will receive). */
int pcb_reserve(pcb_handle* h, uint64_t n);
/* Run all device work of this handle on the caller's CUDA stream (a cudaStream_t), e.g. so that the
   caller's CUDA events bracket it.  Call right after pcb_open. */
int pcb_set_stream(pcb_handle* h, void* cuda_stream);
/* Same, but `d_pts` is a device pointer on the handle's device (resident-input path).  When it is the first
   add on the handle the buffer is aliased, not copied: keep it valid and unchanged until pcb_finish returns.
   STREAM ORDER: the handle works on its own non-blocking stream unless pcb_set_stream gave it the caller's.  Borrowed
   device inputs (this call and pcb_add_points_device_batched) must be COMPLETE on the handle's stream when the call is
   made: produce them on the stream passed to pcb_set_stream, or synchronise the producer stream first.  The library
   inserts no cross-stream dependency of its own. */
int pcb_add_points_device(pcb_handle* h, const pcb_point* d_pts, uint64_t n);
/* Multi-GPU ingest: points that were routed to this device by the level-0 exchange, still in the order of the
   undivided input stream, each with the ordinal (0-based, global) of the batch it belonged to (lib.rs:31-35 cut
   of the undivided stream).  Device pointers, borrowed until pcb_finish returns; the only add on the handle. */
int pcb_add_points_device_batched(pcb_handle* h, const pcb_point* d_pts, const uint32_t* d_batch, uint64_t n);
/* Multi-GPU ingest, generalised: the arrivals of ONE level of the hierarchy in (cell, arrival order), with their CSR -
   what converter.rs:124-134 would hand to the cells of that level over the whole run.  d_batch_plus1[i] = global batch
   ordinal + 1 of arrival i (lib.rs:31-35 cut of the undivided stream); d_cell_start[ncell + 1] = first arrival of every
   cell; d_cell_idx[3 * ncell] = cell indices (metadata.rs:100-102) at `level`, every cell non-empty.  Level 0: what the
   fused exchange pcb_route_scatter_cells delivers; level l > 0: the pending arrivals of another handle's partial build
   (pcb_set_stop_level), re-balanced across ranks.  bbox6 (min xyz, max xyz; may be NULL) and count_points update
   metadata.json's bounding_box / number_of_points like converter.rs:96-108 would.  Device pointers on the handle's
   device, complete on the handle's stream (pcb_set_stream) and borrowed until pcb_finish returns (in merge mode
   d_cell_start is updated in place); the only add on the handle. */
int pcb_add_arrivals(pcb_handle* h, uint32_t level, const pcb_point* d_pts, const uint32_t* d_batch_plus1, uint64_t n,
                     const uint32_t* d_cell_start, const int32_t* d_cell_idx, uint64_t ncell, const float* bbox6,
                     uint64_t count_points);
/* Partial build: pcb_finish builds (and writes) the levels < stop_level only; the arrivals of level stop_level stay on
   the device and are described by pcb_pending_arrivals (valid until pcb_close).  Call before pcb_finish. */
int pcb_set_stop_level(pcb_handle* h, uint32_t stop_level);
typedef struct pcb_arrivals {
    uint32_t level;
    float prism_load;               /* arrivals per occupied hex prism on the last level built: pass it to the handle that
                                       resumes (pcb_set_prism_load) so that its first level takes the same record-filter
                                       decision a single handle would have taken; a performance hint, never affects results */
    uint64_t n, ncell;
    const pcb_point* d_pts;
    const uint32_t* d_batch_plus1;
    const uint32_t* d_cell_start;   /* [ncell + 1] */
    const int32_t* d_cell_idx;      /* [3 * ncell] */
} pcb_arrivals;
int pcb_pending_arrivals(const pcb_handle* h, pcb_arrivals* out);
int pcb_set_prism_load(pcb_handle* h, float prism_load);
int pcb_finish(pcb_handle* h, pcb_stats* stats);
/* Metadata::write_to (metadata.rs:51-53) for a reduced result: used by rank 0 after the per-rank builds. */
int pcb_write_metadata(const char* out_dir, uint64_t number_of_points, uint32_t hierarchies, const float bbox_min[3],
                       const float bbox_max[3], uint32_t cell_point_overflow_limit, uint32_t sub_grid_dimension,
                       float max_cell_size);
/* Config in force on this handle (defaults, pcb_options override, or the existing metadata.json). */
int pcb_get_config(const pcb_handle* h, uint32_t* cell_point_overflow_limit, uint32_t* sub_grid_dimension,
                   float* max_cell_size);
void pcb_close(pcb_handle* h);
/* A warm process keeps what a conversion needs between handles: pinned host buffers (node images, staging chunks),
   the device arenas, the staging memory of the stream-ordered pool and two idle streams per device - pcb_open /
   pcb_close then make no allocating or driver-lock calls.  This call frees the idle buffers, arenas and pool memory
   (the streams stay: a few hundred bytes each). */
void pcb_trim_host_cache(void);
const char* pcb_last_error(const pcb_handle* h);

/* Results in memory (after pcb_finish). */
uint64_t pcb_node_count(const pcb_handle* h);
int pcb_node_get(const pcb_handle* h, uint64_t i, pcb_node_info* info, const uint8_t** host_image);
int pcb_metadata_json(const pcb_handle* h, char* buf, uint64_t cap); /* returns length */
int pcb_kernel_stat_count(const pcb_handle* h);
int pcb_kernel_stat_get(const pcb_handle* h, int i, pcb_kernel_stat* out);

/* ---- multi-GPU routing by level-0 cell (one process per GPU).  Device pointers on `device`,
   work runs on `stream` (a cudaStream_t).  Return >= 0: kernels launched; < 0: a pcb_status code, message through
   pcb_last_error(NULL).  The reference has no counterpart: it is single-process.
   1. pcb_route_bbox: min / max of the local slice (+inf / -inf when n == 0), validates coordinates.
   2. pcb_route_histogram: points per level-0 cell (metadata.rs:100-102 at level 0) inside the dense index box
      [origin, origin + dims), x fastest; the caller all-reduces the table and plans owners.
   3. pcb_route_pack: stable partition of the slice by owner[cell] into the all-to-all send buffer, with the
      global batch ordinal (first_index + i) / batch_size of every point; send_counts[world] on the host.
   4. pcb_route_scatter (world <= 8): partition and exchange in ONE kernel - every point and its batch ordinal are
      stored straight into the receive buffers of the owner rank: recv_pts_ptrs[o] / recv_batch_ptrs[o] are device
      addresses valid on `device` (peer-mapped memory of rank o reached over NVLink, local memory for the own rank),
      recv_offsets[o] = number of points the lower ranks send to o.  Replaces step 3 + the all-to-all; the caller
      synchronises the ranks before (buffers free) and after (stores landed). */
int pcb_route_bbox(int device, void* stream, const pcb_point* d_pts, uint64_t n, float max_cell_size, float out_min[3],
                   float out_max[3]);
int pcb_route_histogram(int device, void* stream, const pcb_point* d_pts, uint64_t n, float max_cell_size,
                        const int32_t origin[3], const uint32_t dims[3], uint32_t* d_hist);
int pcb_route_pack(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index,
                   uint32_t batch_size, float max_cell_size, const int32_t origin[3], const uint32_t dims[3],
                   const uint8_t* d_owner, uint32_t world, pcb_point* d_send_pts, uint32_t* d_send_batch,
                   uint64_t* send_counts);
int pcb_route_scatter(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index,
                      uint32_t batch_size, float max_cell_size, const int32_t origin[3], const uint32_t dims[3],
                      const uint8_t* d_owner, uint32_t world, const uint64_t* recv_pts_ptrs,
                      const uint64_t* recv_batch_ptrs, const uint64_t* recv_offsets);

/* 4b. pcb_route_scatter_cells (world <= 8, level-0 box of at most 512 cells): exchange AND level-0 split in one kernel.
   d_dst_base[cell] = first slot of this rank's points of `cell` inside the receive buffer of owner[cell] (= first slot
   of the cell there + the points lower ranks send to it: the caller derives both from the all-gathered per-rank
   histograms).  Every owner receives its level-0 arrivals in (cell, stream order) with batch ordinal + 1, ready for
   pcb_add_arrivals(level 0).  d_err (device u32, may be NULL): PCB_DERR bits, non-zero = the pass starved (time-sliced
   device) and the buffers must not be used. */
int pcb_route_scatter_cells(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index,
                            uint32_t batch_size, float max_cell_size, const int32_t origin[3], const uint32_t dims[3],
                            const uint8_t* d_owner, const uint32_t* d_dst_base, uint32_t world,
                            const uint64_t* recv_pts_ptrs, const uint64_t* recv_batch_ptrs, uint32_t* d_err);
/* Level-l re-balance: packs whole cells of one level's arrivals (pcb_pending_arrivals) into a send buffer.  d_sel[k] =
   source cell of output cell k, d_out_start[nsel + 1] = CSR of the output; `total` = d_out_start[nsel].  16 + 4 bytes per
   arrival, one coalesced pass.  No reference counterpart (the reference is single-process). */
int pcb_gather_cells(int device, void* stream, const pcb_point* d_pts, const uint32_t* d_batch, const uint32_t* d_cell_start,
                     const uint32_t* d_sel, const uint32_t* d_out_start, uint32_t nsel, uint64_t total, pcb_point* d_out_pts,
                     uint32_t* d_out_batch);
/* Deterministic greedy owner plan (largest cell first onto the least loaded rank; host arrays): owner[i] for every
   cell ordinal i of the histogram `counts`.  A pure function of the global histogram: every rank computes the same. */
int pcb_plan_owners(const uint64_t* counts, uint64_t ncells, uint32_t world, uint8_t* owner);

/* pcb_route_bbox without any allocation.  NO counterpart in the reference: added
   for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
   it).  The kernel's scratch (min / max / error words) lives in a file-scope device array of PCB_STATIC_SCRATCH_BYTES
   = 256 bytes that every GPU loading the library holds until exit; the call performs no cudaMalloc / cudaFree /
   allocating library routine, no host synchronisation and no device-to-host copy.  d_out6 (device, 6 floats: min xyz,
   max xyz; +inf / -inf when n == 0) and d_err (device u32, may be NULL: 1 = non-finite coordinate, 2 = level-0 index
   outside +-2^20) are written on `stream`.  n > PCB_BBOX_STATIC_MAX_POINTS is rejected with PCB_ERR_RANGE.  Every call
   on a device shares the scratch array (re-initialised at the start of each call): calls that OVERLAP on the GPU
   (different streams) race - serialise them.  Returns the number of kernels launched, or a negative pcb_status. */
#define PCB_BBOX_STATIC_MAX_POINTS ((1ull << 30) - 1)
int pcb_route_bbox_static(int device, void* stream, const pcb_point* d_pts, uint64_t n, float max_cell_size,
                          float* d_out6, uint32_t* d_err);

/* ---- render-time cull / level-of-detail pass: the compute shader the renderer dispatches per
   loaded cell, src/plugins/render/point/pipelines/compute.wgsl:81-205, as a CUDA kernel.  For every point of one cell:
   view / projection transform, frustum test, depth-texture test, pick of the deepest LOADED hierarchy that contains
   the point (binary search over the loaded cells, optionally overridden by distance to the camera), colour alpha :=
   that hierarchy, append to d_out through the atomic counter *d_count (the indirect draw buffer's instance_count:
   the caller zeroes it, or lets several cells accumulate).  Output order is the atomic's, as in the shader.
   CAPACITY: d_out must hold (*d_count on entry) + n records - when several dispatches accumulate into one counter they
   share ONE output buffer sized for all of them, as the renderer's indirect buffer is; nothing is bounds-checked.
   Matrices are column-major (glam / wgpu): m[4 * col + row].  d_hierarchies: number_of_hierarchies x (cell_size,
   spacing) floats; d_loaded_cells: n_loaded x (hierarchy, x, y, z) i32, sorted ascending, 16-byte aligned;
   d_far_distances: >= max_hierarchy + 1 floats (only read when size_by_distance); d_depth: depth_height rows of
   depth_width floats.  Returns kernels launched or a negative pcb_status. */
typedef struct pcb_cull_params {
    float view[16], projection[16];
    float cam_pos[3];
    uint32_t depth_width, depth_height;
    uint32_t number_of_hierarchies;
    uint32_t size_by_distance, max_hierarchy;   /* FrustumsSettings, compute.wgsl:36-39 */
    uint32_t cell_hierarchy;                    /* hierarchy of the cell whose points d_in holds (compute.wgsl:67-75) */
} pcb_cull_params;
int pcb_cull_lod(int device, void* stream, const pcb_cull_params* params, const float* d_hierarchies,
                 const int32_t* d_loaded_cells, uint32_t n_loaded, const float* d_far_distances, const float* d_depth,
                 const pcb_point* d_in, uint64_t n, pcb_point* d_out, uint32_t* d_count);

const char* pcb_version(void);
int pcb_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* PCB_H */

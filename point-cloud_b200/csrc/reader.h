// reader.h - input readers of the converter front end (see reader.cpp): streaming, bounded memory.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/pcb.h"

namespace pcb {

enum ReadStatus {
    READ_OK = 0,
    READ_UNSUPPORTED = 1,  // unknown extension: warn + skip the file (lib.rs:78-81)
    READ_TRUNCATED = 2,    // reader error mid-file: log, keep the complete batches read so far (lib.rs:36-39)
    READ_FATAL = 3         // the reference unwrap()s here (file cannot be opened / header unreadable)
};

struct ReadInfo {
    uint64_t total_points = 0;  // BatchedPointReader::total_points
    uint64_t points_read = 0;
    bool ascii_ply = false;     // ply.rs:43-51: total_points are "read" but none is delivered
    std::string message;
};

// fixed-stride layout of the point records of an uncompressed LAS file (converter/las.rs:22-55)
struct LasLayout {
    uint32_t stride = 0;
    int32_t colour_off = -1;    // byte offset of (red, green, blue) u16 inside a record, -1: format without colour
    double scale[3] = {1, 1, 1}, offset[3] = {0, 0, 0};
};
void las_decode_host(const uint8_t* records, uint64_t n, const LasLayout& layout, pcb_point* out);

// One input file as a stream of point records: the mirror of trait BatchedPointReader (converter.rs:24-30) - the
// reference never holds more than one get_batch(10_000) of input, this reader never more than the chunk it is asked for.
class PointStream {
   public:
    struct Impl;
    static PointStream* open(const std::string& path, int* status, std::string* message);   // null unless READ_OK
    ~PointStream();
    uint64_t total_points() const;
    bool ascii_ply() const;
    const LasLayout* las_layout() const;   // non-null for .las: next_raw() may be used instead of next()
    // up to max_points decoded records into out; *n = records delivered (0: end of file).  READ_OK or READ_TRUNCATED.
    int next(uint64_t max_points, pcb_point* out, uint64_t* n, std::string* message);
    // .las only: up to max_points RAW records (stride bytes each) into out, for the device-side decode
    int next_raw(uint64_t max_points, uint8_t* out, uint64_t* n, std::string* message);

   private:
    PointStream();
    Impl* impl_;
};

// Reads every point of `path` (appended to *out): the stream drained into one vector (tests, --dump-points).
int read_points_file(const std::string& path, std::vector<pcb_point>* out, ReadInfo* info);

}  // namespace pcb

// reader.h — input readers of the converter front end (see reader.cpp).
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

// Reads every point of `path` (appended to *out).
int read_points_file(const std::string& path, std::vector<pcb_point>* out, ReadInfo* info);

}  // namespace pcb

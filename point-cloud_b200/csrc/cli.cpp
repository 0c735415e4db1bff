// cli.cpp — `point_converter`: the reference's command line (point-converter/src/main.rs:11-50) and its driver
// loop (lib.rs:11-60) on top of the C ABI (include/pcb.h).  Same flags, same log lines, same output layout.
//   -o, --output <DIR>        output directory (created; default: current directory)      main.rs:16-17, 47-48
//   -d, --directories <DIRS>  directories whose entries are converted (one level)         main.rs:20-21, 32-36
//   -f, --files <FILES>       input files                                                   main.rs:24-25
#include <dirent.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pcb.h"
#include "reader.h"

static void log_line(const char* level, const std::string& msg) { std::fprintf(stderr, "[%s point_converter] %s\n", level, msg.c_str()); }

static void usage() {
    std::printf(
        "Point converter will convert your points to a format that the point cloud renderer can use.\n"
        "Currently supported file formats are las and ply and the generated metadata.json.\n\n"
        "Usage: point_converter [OPTIONS]\n\nOptions:\n"
        "  -o, --output <DIR>         Output directory of the converted format.\n"
        "                             Will be created if it doesn't exist.\n"
        "  -d, --directories <DIRS>   Directories with input files to convert.\n"
        "  -f, --files <FILES>        Input files with the points to convert.\n"
        "      --device <N>           CUDA device ordinal (default 0)\n"
        "      --host-las             decode LAS records on the host (default: on the device)\n"
        "  -h, --help                 Print help\n  -V, --version              Print version\n");
}

int main(int argc, char** argv) {
    std::string output;
    std::vector<std::string> dirs, files;
    int device = 0;
    std::string dump_points;   // test hook: read the inputs, write the 16-byte records, do not touch a GPU
    bool host_las = false;     // --host-las: decode LAS records on the host instead of on the device
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto values = [&](std::vector<std::string>* dst) {
            while (i + 1 < argc && argv[i + 1][0] != '-') dst->push_back(argv[++i]);
        };
        if (a == "-h" || a == "--help") { usage(); return 0; }
        if (a == "-V" || a == "--version") { std::printf("point_converter (%s)\n", pcb_version()); return 0; }
        if ((a == "-o" || a == "--output") && i + 1 < argc) output = argv[++i];
        else if (a.rfind("--output=", 0) == 0) output = a.substr(9);
        else if (a == "-d" || a == "--directories") values(&dirs);
        else if (a == "-f" || a == "--files") values(&files);
        else if (a == "--device" && i + 1 < argc) device = std::atoi(argv[++i]);
        else if (a == "--dump-points" && i + 1 < argc) dump_points = argv[++i];
        else if (a == "--host-las") host_las = true;
        else {
            std::fprintf(stderr, "error: unexpected argument '%s'\n\n", a.c_str());
            usage();
            return 2;
        }
    }
    // main.rs:32-38: files first, then the entries of every directory (read_dir order, one level)
    std::vector<std::string> paths = files;
    for (auto& d : dirs) {
        DIR* dp = opendir(d.c_str());
        if (!dp) {  // read_dir(path).unwrap()
            log_line("ERROR", "cannot read directory " + d);
            return 101;
        }
        while (dirent* e = readdir(dp)) {
            std::string nm = e->d_name;
            if (nm != "." && nm != "..") paths.push_back(d + "/" + nm);
        }
        closedir(dp);
    }
    if (paths.empty()) {  // main.rs:40-43
        log_line("WARN ", "Please provide some files or directories");
        return 0;
    }
    if (!dump_points.empty()) {
        FILE* f = std::fopen(dump_points.c_str(), "wb");
        if (!f) return 101;
        for (auto& path : paths) {
            std::vector<pcb_point> pts;
            pcb::ReadInfo info;
            int rs = pcb::read_points_file(path, &pts, &info);
            std::fprintf(stderr, "%s: status %d total %llu read %llu ascii %d %s\n", path.c_str(), rs,
                         (unsigned long long)info.total_points, (unsigned long long)info.points_read, (int)info.ascii_ply,
                         info.message.c_str());
            if (!pts.empty()) std::fwrite(pts.data(), sizeof(pcb_point), pts.size(), f);
        }
        std::fclose(f);
        return 0;
    }
    if (output.empty()) {  // main.rs:47-48
        char cwd[4096];
        output = getcwd(cwd, sizeof cwd) ? cwd : ".";
    }

    pcb_options opt;
    std::memset(&opt, 0, sizeof opt);
    opt.struct_size = sizeof opt;
    opt.device = device;
    pcb_handle* h = nullptr;
    if (access((output + "/metadata.json").c_str(), R_OK) == 0) log_line("INFO ", "Found an existing metadata file.");
    else log_line("INFO ", "Found no metadata file. A new one will be created.");   // lib.rs:92-99
    if (pcb_open(output.c_str(), &opt, &h) != PCB_OK) {
        log_line("ERROR", pcb_last_error(nullptr));
        return 101;  // the reference panics
    }
    const uint64_t BATCH = 10000;  // lib.rs:32
    // Streamed ingest (lib.rs:31-52 never holds more than one batch of input; here: never more than three chunks).
    // A reader thread fills page-locked chunks (file read + decode on up to 8 host threads) while this thread hands
    // the previous chunk to the device (pcb_add_points: one DMA from pinned memory).  Uncompressed LAS records travel
    // raw and are decoded on the device (pcb_add_las_records).
    const uint64_t chunk_pts = (4u << 20) / BATCH * BATCH;   // points per chunk, a multiple of the batch size (~64 MB)
    struct Chunk {
        void* mem = nullptr;
        uint64_t n = 0;
        int status = pcb::READ_OK;
        bool raw = false, last = false;
        std::string message;
    };
    uint64_t total_added = 0;
    auto t0 = std::chrono::steady_clock::now();
    for (size_t pi = 0; pi < paths.size(); pi++) {
        log_line("INFO ", "Converting file " + std::to_string(pi + 1) + "/" + std::to_string(paths.size()) + ", \"" + paths[pi] + "\"");
        int rs = pcb::READ_OK;
        std::string msg;
        std::unique_ptr<pcb::PointStream> ps(pcb::PointStream::open(paths[pi], &rs, &msg));
        if (rs == pcb::READ_UNSUPPORTED) {
            if (!msg.empty()) log_line("WARN ", msg);  // lib.rs:78-81
            continue;
        }
        if (rs == pcb::READ_FATAL || !ps) {
            log_line("ERROR", msg);
            pcb_close(h);
            return 101;
        }
        log_line("INFO ", "Converting " + std::to_string(ps->total_points()) + " points");
        const pcb::LasLayout* las = host_las ? nullptr : ps->las_layout();
        if (las && (las->stride < 12 || las->stride > 96)) las = nullptr;
        if (!ps->ascii_ply()) pcb_reserve(h, std::min<uint64_t>(total_added + ps->total_points(), (1ull << 30) - 1));
        // three chunks in rotation: one being filled, one queued, one being copied
        std::vector<Chunk> ring(3);
        const uint64_t want_pts = std::max<uint64_t>(BATCH, std::min<uint64_t>(chunk_pts, (ps->total_points() + BATCH - 1) / BATCH * BATCH));
        for (auto& c : ring) {
            c.mem = pcb_alloc_host(las ? want_pts * las->stride : want_pts * sizeof(pcb_point));
            if (!c.mem) {
                log_line("ERROR", "cannot allocate page-locked input chunks");
                pcb_close(h);
                return 101;
            }
        }
        std::mutex mu;
        std::condition_variable cv;
        size_t filled = 0, consumed = 0;   // chunk k lives in ring[k % 3]
        bool stop = false;
        std::thread reader([&] {
            for (size_t k = 0;; k++) {
                {
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return stop || k < consumed + ring.size(); });
                    if (stop) return;
                }
                Chunk& c = ring[k % ring.size()];
                c.raw = las != nullptr;
                c.status = las ? ps->next_raw(want_pts, (uint8_t*)c.mem, &c.n, &c.message)
                               : ps->next(want_pts, (pcb_point*)c.mem, &c.n, &c.message);
                c.last = c.n == 0 || c.status != pcb::READ_OK;
                const bool last = c.last;
                {
                    std::lock_guard<std::mutex> lk(mu);
                    filled = k + 1;
                }
                cv.notify_all();
                if (last) return;
            }
        });
        bool failed = false, began = false;
        uint64_t file_points = 0;
        for (size_t k = 0;; k++) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return filled > k; });
            }
            Chunk& c = ring[k % ring.size()];
            uint64_t keep = c.n;
            if (c.status == pcb::READ_TRUNCATED) {  // lib.rs:36-39: the failing get_batch delivers nothing, the file ends there
                log_line("ERROR", c.message);
                keep = (file_points + keep) / BATCH * BATCH - file_points;   // chunks are whole batches: trim the last one
            }
            int rc = PCB_OK;
            // a file without points still yields one (empty) batch (lib.rs:32-34); a file whose FIRST get_batch fails
            // yields none (lib.rs:36-39)
            if (keep || (c.last && file_points == 0 && c.status == pcb::READ_OK)) {
                if (!began) rc = pcb_begin_file(h);
                began = true;
                if (rc == PCB_OK)
                    rc = c.raw ? pcb_add_las_records(h, (const uint8_t*)c.mem, keep, las->stride, las->colour_off, las->scale, las->offset)
                               : pcb_add_points(h, keep ? (const pcb_point*)c.mem : nullptr, keep);
            }
            file_points += keep;
            const bool last = c.last;
            {
                std::lock_guard<std::mutex> lk(mu);
                consumed = k + 1;
                if (rc != PCB_OK) stop = true;
            }
            cv.notify_all();
            if (rc != PCB_OK) {
                failed = true;
                break;
            }
            if (last) break;
        }
        reader.join();
        for (auto& c : ring) pcb_free_host(c.mem);
        total_added += file_points;
        if (failed) {
            log_line("ERROR", pcb_last_error(h));
            pcb_close(h);
            return 101;
        }
    }
    pcb_stats st;
    if (pcb_finish(h, &st) != PCB_OK) {
        log_line("ERROR", pcb_last_error(h));
        pcb_close(h);
        return 101;
    }
    auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
    log_line("INFO ", "Finished converting after " + std::to_string(ms) + " ms");   // lib.rs:56-59
    log_line("INFO ", "points " + std::to_string(st.number_of_points) + ", levels " + std::to_string(st.hierarchies) + ", nodes " +
                          std::to_string(st.nodes_written) + ", device build " + std::to_string((long)(st.build_seconds * 1e3)) +
                          " ms, file write " + std::to_string((long)(st.write_seconds * 1e3)) + " ms");
    pcb_close(h);
    return 0;
}

// cli.cpp — `point_converter`: the reference's command line (point-converter/src/main.rs:11-50) and its driver
// loop (lib.rs:11-60) on top of the C ABI (include/pcb.h).  Same flags, same log lines, same output layout.
//   -o, --output <DIR>        output directory (created; default: current directory)      main.rs:16-17, 47-48
//   -d, --directories <DIRS>  directories whose entries are converted (one level)         main.rs:20-21, 32-36
//   -f, --files <FILES>       input files                                                   main.rs:24-25
#include <dirent.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
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
        "  -h, --help                 Print help\n  -V, --version              Print version\n");
}

int main(int argc, char** argv) {
    std::string output;
    std::vector<std::string> dirs, files;
    int device = 0;
    std::string dump_points;   // test hook: read the inputs, write the 16-byte records, do not touch a GPU
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
    auto t0 = std::chrono::steady_clock::now();
    for (size_t pi = 0; pi < paths.size(); pi++) {
        log_line("INFO ", "Converting file " + std::to_string(pi + 1) + "/" + std::to_string(paths.size()) + ", \"" + paths[pi] + "\"");
        std::vector<pcb_point> pts;
        pcb::ReadInfo info;
        int rs = pcb::read_points_file(paths[pi], &pts, &info);
        if (rs == pcb::READ_UNSUPPORTED) {
            if (!info.message.empty()) log_line("WARN ", info.message);  // lib.rs:78-81
            continue;
        }
        if (rs == pcb::READ_FATAL) {
            log_line("ERROR", info.message);
            pcb_close(h);
            return 101;
        }
        log_line("INFO ", "Converting " + std::to_string(info.total_points) + " points");
        uint64_t keep = pts.size();
        if (rs == pcb::READ_TRUNCATED) {  // lib.rs:36-39: the failing get_batch delivers nothing, the file ends there
            log_line("ERROR", info.message);
            keep = keep / BATCH * BATCH;
            if (keep == 0) continue;
        }
        if (pcb_begin_file(h) != PCB_OK || pcb_add_points(h, keep ? pts.data() : nullptr, keep) != PCB_OK) {
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

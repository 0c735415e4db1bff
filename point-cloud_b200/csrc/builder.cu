// builder.cu — host driver of the device build: owns device memory and streams, launches the per-level
// kernels of level.cuh in the order tests/bulk_model.py specifies.  There is no CPU compute path here:
// without a CUDA device the constructor throws (PCB_ERR_CUDA).
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <mutex>
#include <thread>
#include <atomic>
#include <exception>
#include <vector>

#include "builder.h"
#include "common.cuh"
#include "level.cuh"
#include "radix.cuh"
#include "scan.cuh"
#include "route.cuh"

namespace pcb {

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// PCB_TRACE=1: host-side timeline of the build on stderr (synchronises at every mark; debugging only)
static bool g_trace = pcb_tuning_env("PCB_TRACE") != nullptr;
static double g_trace_t0 = 0, g_trace_last = 0;
static void trace_mark(cudaStream_t st, const char* what, unsigned level = 0, unsigned long long n = 0) {
    if (!g_trace) return;
    double a = now_s();
    cudaStreamSynchronize(st);
    double b = now_s();
    fprintf(stderr, "[pcb] %8.3f ms (+%7.3f host, +%7.3f drain) L%u n=%llu %s\n", (b - g_trace_t0) * 1e3,
            (a - g_trace_last) * 1e3, (b - a) * 1e3, level, n, what);
    g_trace_last = now_s();
}

// Process-wide cache of pinned host buffers: cudaHostAlloc costs ~0.3 ms/MB, far more than the copy it
// serves, so node-image staging buffers are recycled between levels, handles and runs.
struct PinnedCache {
    std::mutex mu;
    std::vector<std::pair<void*, u64>> free_list;
    void* get(u64 bytes, u64* got) {
        {
            std::lock_guard<std::mutex> g(mu);
            size_t best = SIZE_MAX;
            for (size_t i = 0; i < free_list.size(); i++)
                if (free_list[i].second >= bytes && (best == SIZE_MAX || free_list[i].second < free_list[best].second)) best = i;
            if (best != SIZE_MAX && free_list[best].second <= 2 * bytes + (1u << 20)) {
                void* p = free_list[best].first;
                *got = free_list[best].second;
                free_list.erase(free_list.begin() + best);
                return p;
            }
        }
        void* p = nullptr;
        u64 sz = (bytes + (1u << 20) - 1) & ~((u64)(1u << 20) - 1);
        CUDA_TRY(cudaHostAlloc(&p, sz, cudaHostAllocMapped));   // mapped: kernels publish small results into it
        *got = sz;
        return p;
    }
    void put(void* p, u64 bytes) {
        std::lock_guard<std::mutex> g(mu);
        free_list.emplace_back(p, bytes);
    }
    void trim() {
        std::lock_guard<std::mutex> g(mu);
        for (auto& f : free_list) cudaFreeHost(f.first);
        free_list.clear();
    }
};
static PinnedCache g_pinned;

// Process-wide cache of device work arenas.  The stream-ordered pool re-maps virtual memory when blocks of
// varying sizes are recycled (measured: milliseconds inside cudaMallocAsync while the GPU idles), so the
// per-level arrays are bump-allocated from three big blocks instead: [0]/[1] alternate between levels (level h
// allocates from h & 1 and reads its inputs from the other one), [2] holds the node images of a build.
// Anything that does not fit falls back to cudaMallocAsync.
struct Arena {
    unsigned char* base = nullptr;
    u64 cap = 0, off = 0;
};
struct ArenaSet {
    int device = -1;
    bool in_use = false;
    Arena a[4];   // 0/1: per-level arrays (ping-pong), 2: node images, 3: level-0 arrivals of the whole run
};
struct ArenaCache {
    std::mutex mu;
    std::vector<ArenaSet*> sets;
    ArenaSet* acquire(int device) {
        std::lock_guard<std::mutex> g(mu);
        for (auto* s : sets)
            if (s->device == device && !s->in_use) { s->in_use = true; return s; }
        ArenaSet* s = new ArenaSet();
        s->device = device;
        s->in_use = true;
        sets.push_back(s);
        return s;
    }
    void release(ArenaSet* s) {
        std::lock_guard<std::mutex> g(mu);
        s->in_use = false;
    }
    void trim() {
        std::lock_guard<std::mutex> g(mu);
        for (auto* s : sets) {
            if (s->in_use) continue;
            cudaSetDevice(s->device);
            for (auto& a : s->a) {
                if (a.base) cudaFree(a.base);
                a = Arena();
            }
        }
    }
};
static ArenaCache g_arenas;

// Runs f(0..n-1) on up to 16 host threads; the first exception is re-thrown on the caller's thread.
template <class F>
static void host_parallel_for(u32 n, F f) {
    const u32 nt = std::min<u32>(std::min<u32>(16u, std::max(1u, std::thread::hardware_concurrency())), n);
    if (nt <= 1) {
        for (u32 i = 0; i < n; i++) f(i);
        return;
    }
    std::atomic<u32> next{0};
    std::atomic<bool> failed{false};
    std::exception_ptr err;
    std::mutex mu;
    std::vector<std::thread> th;
    for (u32 t = 0; t < nt; t++)
        th.emplace_back([&] {
            try {
                for (u32 i = next.fetch_add(1); i < n && !failed.load(); i = next.fetch_add(1)) f(i);
            } catch (...) {
                std::lock_guard<std::mutex> g(mu);
                if (!err) err = std::current_exception();
                failed.store(true);
            }
        });
    for (auto& t : th) t.join();
    if (err) std::rethrow_exception(err);
}

// Small results go to the host by kernel stores into mapped pinned memory, not through the copy engine: a
// device-to-host memcpy queues behind the node-image download that is in flight on the other stream (hundreds of
// MB), and the host needs these few bytes to size the next level - measured 2-4 ms of build stall per level.
__global__ void k_publish(const unsigned char* __restrict__ src, unsigned char* __restrict__ dst, u64 bytes,
                          int vec16) {
    const u64 i = ((u64)blockIdx.x * blockDim.x + threadIdx.x) * 16ull;
    if (i >= bytes) return;
    if (vec16 && i + 16 <= bytes) {
        *(uint4*)(dst + i) = *(const uint4*)(src + i);
    } else {   // 4-byte aligned data (slices of index tables), or the tail
        const u64 e = i + 16 < bytes ? i + 16 : bytes;
        u64 b = i;
        for (; b + 4 <= e; b += 4) *(u32*)(dst + b) = *(const u32*)(src + b);
        for (; b < e; b++) dst[b] = src[b];
    }
}

struct Scratch {  // small device-resident results, mirrored in pinned host memory
    u64 diffmask;
    u32 err;
    u32 ncell;
    u32 nwin;
    u32 aux;   // record filter: number of early arrivals, then length of the round-2 key list
    int bbox[6];
    LevelTotals totals;
};

class BuilderImpl {
   public:
    int device;
    uint32_t flags;
    bool profile;
    cudaStream_t st = nullptr, copy_st = nullptr;
    uint32_t limit = 5000, dim = 96;
    float max_cell_size = 1000.0f;

    // staged input (input order), batch segments
    PcbPoint* d_in = nullptr;
    bool in_borrowed = false;  // d_in aliases the caller's device buffer (pcb_add_points_device, first add)
    bool own_stream = true;
    u64 n_in = 0, cap_in = 0;
    std::vector<BatchSeg> segs;
    u32 next_batch = 0;
    u64 batches = 0;

    Scratch* d_scr = nullptr;
    Scratch* h_scr = nullptr;
    std::vector<std::pair<void*, u64>> host_tmps;
    u64 h_scr_cap = 0;
    u32* d_bat_in = nullptr;   // optional per-point batch ordinals (pcb_add_points_device_batched)
    bool bat_owned = false;

    std::vector<NodeRec> nodes;
    std::vector<LevelImage> levels;
    std::vector<KernelStat> stats;
    struct ProfRec {
        int cls;
        cudaEvent_t a, b;
    };
    std::vector<ProfRec> prof_recs;
    cudaEvent_t prof_a = nullptr;
    u64 launches = 0, arrivals = 0;
    u32 levels_built = 0;
    bool inputs_borrowed = false;   // build_level must not free its inputs (slices of the run's level-0 arrays)
    double h2d_s = 0, build_s = 0, d2h_s = 0;

    // Process-wide cache of idle streams per device: cudaStreamCreate / cudaStreamDestroy take driver locks that an NVML
    // query of another thread or of a monitoring agent can hold for ~13 ms (measured: pcb_open / pcb_close of a warm
    // process took 13-27 ms instead of 0.2 ms in one step out of three on some boxes while pcb_finish kept its time);
    // a handle takes its two streams from here and hands them back idle.  Never destroyed (process lifetime).
    struct StreamCache {
        std::mutex mu;
        std::map<int, std::vector<cudaStream_t>> idle;
        cudaStream_t get(int dev) {
            {
                std::lock_guard<std::mutex> lk(mu);
                auto& v = idle[dev];
                if (!v.empty()) { cudaStream_t s = v.back(); v.pop_back(); return s; }
            }
            cudaStream_t s = nullptr;
            CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            return s;
        }
        void put(int dev, cudaStream_t s) {   // s must be idle
            std::lock_guard<std::mutex> lk(mu);
            idle[dev].push_back(s);
        }
    };
    static StreamCache& stream_cache() { static StreamCache* c = new StreamCache(); return *c; }
    struct PoolDevices {
        std::mutex mu;
        std::map<int, bool> set;   // devices whose default pool this library configured (and may hold cached memory in)
    };
    static PoolDevices& pool_devices() { static PoolDevices* p = new PoolDevices(); return *p; }
    static void trim_pools() {   // pcb_trim_host_cache: hand the pool's idle memory back to the driver
        std::lock_guard<std::mutex> lk(pool_devices().mu);
        int cur = 0;
        cudaGetDevice(&cur);
        for (auto& kv : pool_devices().set) {
            cudaMemPool_t pool;
            if (cudaSetDevice(kv.first) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, kv.first) == cudaSuccess) {
                cudaDeviceSynchronize();
                cudaMemPoolTrimTo(pool, 0);
            }
        }
        cudaSetDevice(cur);
    }

    BuilderImpl(int dev, uint32_t fl) : device(dev), flags(fl), profile((fl & PCB_FLAG_PROFILE) != 0) {
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0)
            throw Error{PCB_ERR_CUDA, "no CUDA device: this library has no CPU fallback"};
        if (dev < 0 || dev >= count) throw Error{PCB_ERR_ARG, "bad device ordinal"};
        CUDA_TRY(cudaSetDevice(dev));
        st = stream_cache().get(dev);
        copy_st = stream_cache().get(dev);
        {   // once per device: the stream-ordered pool keeps freed memory (no trim at synchronisation)
            std::lock_guard<std::mutex> lk(pool_devices().mu);
            if (!pool_devices().set[dev]) {
                cudaMemPool_t pool;
                CUDA_TRY(cudaDeviceGetDefaultMemPool(&pool, dev));
                uint64_t thr = UINT64_MAX;
                CUDA_TRY(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
                pool_devices().set[dev] = true;
            }
        }
        u64 got = 0;
        h_scr = (Scratch*)g_pinned.get(sizeof(Scratch), &got);   // recycled: cudaHostAlloc costs ~0.1 ms
        h_scr_cap = got;
    }
    ~BuilderImpl() {
        cudaSetDevice(device);
        cudaStreamSynchronize(st);
        cudaStreamSynchronize(copy_st);
        for (auto& l : levels) {
            if (l.d_image && !in_arena(l.d_image)) cudaFreeAsync(l.d_image, st);
            if (l.h_image && l.h_cap) g_pinned.put(l.h_image, l.h_cap);   // (h_cap == 0: carved from the slab)
        }
        if (slab_thread.joinable()) slab_thread.join();
        if (h_slab) g_pinned.put(h_slab, slab_cap);
        if (d_in && !in_borrowed) cudaFreeAsync(d_in, st);
        if (d_raw) cudaFree(d_raw);
        if (pend_owned) {
            void* pend[4] = {pend_pts, pend_bat, pend_cell_start, pend_cell_idx};
            for (void* q : pend)
                if (q && !in_arena(q)) cudaFreeAsync(q, st);
        }
        if (d_bat_in && bat_owned) cudaFree(d_bat_in);
        if (d_scr) cudaFreeAsync(d_scr, st);
        if (h_scr) g_pinned.put(h_scr, h_scr_cap);
        cudaStreamSynchronize(st);
        for (auto& t : host_tmps) g_pinned.put(t.first, t.second);
        if (arenas) g_arenas.release(arenas);
        if (st && own_stream) stream_cache().put(device, st);   // idle: synchronised above
        if (copy_st) { cudaStreamSynchronize(copy_st); stream_cache().put(device, copy_st); }
    }

    // Arrivals of level `ext.level` supplied by the caller in (cell, tau) order with their CSR (multi-GPU: the fused
    // level-0 exchange delivers them that way; level-l re-balance: the deliveries of another rank's level l - 1).
    // Device pointers, borrowed until build() returns.  Stage 0 is skipped.
    struct ExtArrivals {
        bool set = false;
        u32 level = 0, n = 0, ncell = 0;
        const PcbPoint* pts = nullptr;
        const u32* bat = nullptr;          // batch ordinal + 1
        const u32* cell_start = nullptr;   // [ncell + 1]
        const int* cell_idx = nullptr;     // [3 * ncell]
        bool has_box = false;
        float lo[3], hi[3];
        u64 count_points = 0;              // added to metadata.number_of_points (0: counted by another handle)
        double prism_load = 0;             // arrivals per occupied prism on the level that produced them (0: unknown)
    } ext;
    // partial build: levels >= stop_level are not built; their arrivals stay on the device (pend_*)
    u32 stop_level = 0xFFFFFFFFu;
    PcbPoint* pend_pts = nullptr;
    u32* pend_bat = nullptr;
    u32* pend_cell_start = nullptr;
    int* pend_cell_idx = nullptr;
    u32 pend_n = 0, pend_ncell = 0, pend_level = 0;
    bool pend_owned = false;

    ArenaSet* arenas = nullptr;
    double last_mult = 0;   // arrivals per occupied prism on the level just built (same wave)
    int cur_arena = 0;
    void arena_prepare(int which, u64 bytes) {   // (re)size an arena; only while nothing of it is in use
        Arena& a = arenas->a[which];
        a.off = 0;
        if (a.cap >= bytes) return;
        CUDA_TRY(cudaStreamSynchronize(st));
        if (a.base) CUDA_TRY(cudaFree(a.base));
        a.base = nullptr;
        a.cap = 0;
        if (cudaMalloc((void**)&a.base, bytes) != cudaSuccess) {   // not fatal: everything falls back to the pool
            cudaGetLastError();
            a.base = nullptr;
            return;
        }
        a.cap = bytes;
    }
    void* arena_alloc(int which, u64 bytes) {
        if (!arenas) return nullptr;
        Arena& a = arenas->a[which];
        u64 start = (a.off + 255) & ~255ull;
        if (!a.base || start + bytes > a.cap) return nullptr;
        a.off = start + bytes;
        return a.base + start;
    }
    bool in_arena(const void* p) const {
        if (!arenas) return false;
        for (auto& a : arenas->a)
            if (a.base && (const unsigned char*)p >= a.base && (const unsigned char*)p < a.base + a.cap) return true;
        return false;
    }
    template <class T>
    T* dalloc_from(int which, u64 n) {
        const u64 bytes = std::max<u64>(n, 1) * sizeof(T);
        if (void* q = arena_alloc(which, bytes)) return (T*)q;
        void* p = nullptr;
        CUDA_TRY(cudaMallocAsync(&p, bytes, st));
        return (T*)p;
    }
    template <class T>
    T* dalloc(u64 n) { return dalloc_from<T>(cur_arena, n); }
    void dfree(void* p) {
        if (p && !in_arena(p)) CUDA_TRY(cudaFreeAsync(p, st));
    }

    int cls(const char* name) {
        for (size_t i = 0; i < stats.size(); i++)
            if (stats[i].name == name) return (int)i;
        KernelStat s;
        s.name = name;
        stats.push_back(s);
        return (int)stats.size() - 1;
    }
    void pb() {
        if (!profile) return;
        CUDA_TRY(cudaEventCreate(&prof_a));
        CUDA_TRY(cudaEventRecord(prof_a, st));
    }
    void pe(const char* name, u64 nlaunch, u64 units, u64 bytes) {
        int c = cls(name);
        stats[c].launches += nlaunch;
        stats[c].units += units;
        stats[c].bytes += bytes;
        launches += nlaunch;
        if (!profile) return;
        ProfRec r;
        r.cls = c;
        r.a = prof_a;
        CUDA_TRY(cudaEventCreate(&r.b));
        CUDA_TRY(cudaEventRecord(r.b, st));
        prof_recs.push_back(r);
    }
    void prof_collect() {
        for (auto& r : prof_recs) {
            float ms = 0;
            cudaEventElapsedTime(&ms, r.a, r.b);
            stats[r.cls].ms += ms;
            cudaEventDestroy(r.a);
            cudaEventDestroy(r.b);
        }
        prof_recs.clear();
    }

    void set_stream(cudaStream_t s) {
        if (st && own_stream) { cudaStreamSynchronize(st); stream_cache().put(device, st); }
        st = s;
        own_stream = false;
    }
    // device -> pinned host, stream-ordered; both 16-byte aligned
    void publish(void* host_pinned, const void* dev, u64 bytes) {
        if (!bytes) return;
        if (bytes <= (4ull << 20)) {
            const int vec16 = (((uintptr_t)dev | (uintptr_t)host_pinned) & 15u) == 0;
            if ((((uintptr_t)dev | (uintptr_t)host_pinned) & 3u) != 0) throw Error{PCB_ERR_ARG, "internal: unaligned publish"};
            k_publish<<<div_up(div_up(bytes, 16), 256), 256, 0, st>>>((const unsigned char*)dev,
                                                                       (unsigned char*)host_pinned, bytes, vec16);
        } else {
            CUDA_TRY(cudaMemcpyAsync(host_pinned, dev, bytes, cudaMemcpyDeviceToHost, st));
        }
    }
    void* host_tmp(u64 bytes) {   // pinned scratch that lives until the end of the run
        u64 got = 0;
        void* p = g_pinned.get(std::max<u64>(bytes, 16), &got);
        host_tmps.emplace_back(p, got);
        return p;
    }
    void fetch_scratch() {
        publish(h_scr, d_scr, sizeof(Scratch));
        CUDA_TRY(cudaStreamSynchronize(st));
    }
    void check_err() {
        u32 e = h_scr->err;
        if (e & PCB_DERR_NONFINITE) throw Error{PCB_ERR_RANGE, "non-finite coordinate in input"};
        if (e & PCB_DERR_L0RANGE) throw Error{PCB_ERR_RANGE, "level-0 cell index outside +-2^20"};
        if (e & PCB_DERR_HEXRANGE)
            throw Error{PCB_ERR_RANGE, "hex key outside its cell window (coordinate precision exhausted)"};
        if (e & PCB_DERR_NEST) throw Error{PCB_ERR_RANGE, "child cell index does not nest in its parent"};
        if (e & PCB_DERR_TIMEOUT)
            throw Error{PCB_ERR_CUDA, "a look-back starved waiting for an earlier tile (device shared / time-sliced?): "
                                      "the build was abandoned instead of hanging"};
    }

    // Page-locked home of the node images.  cudaHostAlloc pins ~3 GB/s: 0.5 s for the 1.6 GB of C2's images, paid by a
    // cold process AFTER its input is on the device.  A capacity hint (pcb_reserve) therefore starts the allocation of ONE
    // slab on a background thread - it overlaps the caller's file reading and H2D copies - and the images of the build
    // are carved from it; whatever does not fit falls back to the recycled per-image buffers.
    std::thread slab_thread;
    unsigned char* h_slab = nullptr;
    u64 slab_cap = 0, slab_off = 0;
    void slab_prefetch(u64 n_points) {
        if ((flags & PCB_FLAG_DEVICE_ONLY) || h_slab || slab_thread.joinable() || n_points < (1u << 20)) return;
        const u64 want = 16ull * n_points + n_points / 4 + (32ull << 20);   // records + headers / padding
        slab_thread = std::thread([this, want] {
            cudaSetDevice(device);
            try {
                u64 got = 0;
                h_slab = (unsigned char*)g_pinned.get(want, &got);
                slab_cap = got;
            } catch (...) {
                h_slab = nullptr;   // not fatal: per-image buffers
                slab_cap = 0;
            }
        });
    }
    unsigned char* slab_take(u64 bytes) {
        if (slab_thread.joinable()) slab_thread.join();
        const u64 start = (slab_off + 255) & ~255ull;
        if (!h_slab || start + bytes > slab_cap) return nullptr;
        slab_off = start + bytes;
        return h_slab + start;
    }

    void reserve(u64 n) {
        slab_prefetch(n);
        if (n <= cap_in) return;
        if (n >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "more than 2^30-1 points on one device"};
        // stream-ordered pool allocation: the pool keeps freed memory (release threshold = max, set at open), so a warm
        // process gets the staging buffer of the previous conversion back without cudaMalloc / cudaFree (driver locks)
        PcbPoint* p = nullptr;
        CUDA_TRY(cudaMallocAsync((void**)&p, n * sizeof(PcbPoint), st));
        if (n_in) CUDA_TRY(cudaMemcpyAsync(p, d_in, n_in * sizeof(PcbPoint), cudaMemcpyDeviceToDevice, st));
        if (d_in && !in_borrowed) CUDA_TRY(cudaFreeAsync(d_in, st));
        d_in = p;
        cap_in = n;
        in_borrowed = false;
    }

    // ---- batch bookkeeping (lib.rs:31-52).  A "run" = one input file: its points are cut into batches
    // of run_bs points; a file without points still yields one (empty) batch because lib.rs:32-34 calls
    // add_points_batch before it looks at remaining_points().
    bool run_open = false;
    u32 run_bs = 0;
    u64 run_points = 0;
    void end_run() {
        if (!run_open) return;
        u64 bs = std::max<u32>(run_bs, 1);
        u64 nb = std::max<u64>(1, (run_points + bs - 1) / bs);
        next_batch += (u32)nb;
        batches += nb;
        run_open = false;
    }
    void open_run() {
        end_run();
        run_open = true;
        run_bs = 0;
        run_points = 0;
    }
    void begin_file() { open_run(); }

    // batch bookkeeping of n more points of the current input file
    void note_points(u64 n, u32& batch_size, bool one_batch) {
        if (d_bat_in || ext.set) throw Error{PCB_ERR_ARG, "points cannot be added after routed points"};
        if (n_in + n >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "more than 2^30-1 points on one device"};
        if (one_batch) {
            open_run();
            batch_size = (u32)std::max<u64>(n, 1);
        } else if (!run_open || (run_bs != 0 && run_bs != batch_size)) {
            open_run();
        }
        if (run_bs == 0 && n > 0) {
            run_bs = batch_size;
            // the run may already hold batches only if it held points; it does not: start a segment here
            BatchSeg s{(u32)n_in, next_batch, batch_size, 0};
            segs.push_back(s);
        }
        run_points += n;
    }

    // Uncompressed LAS records decoded on the device (k_las_decode): the host only copies bytes.
    unsigned char* d_raw = nullptr;
    u64 raw_cap = 0;
    void add_las(const uint8_t* rec, u64 n, u32 stride, int colour_off, const double scale[3], const double offset[3],
                 u32 batch_size) {
        CUDA_TRY(cudaSetDevice(device));
        if (stride < 12 || stride > LAS_DEC_MAX_STRIDE || (colour_off >= 0 && (u32)colour_off + 6 > stride))
            throw Error{PCB_ERR_ARG, "LAS record layout not supported by the device decoder"};
        note_points(n, batch_size, false);
        if (!n) return;
        if (in_borrowed) throw Error{PCB_ERR_ARG, "cannot append to a borrowed device buffer"};
        if (n_in + n > cap_in) reserve(std::min<u64>(std::max<u64>(n_in + n, cap_in + cap_in / 2), (1ull << 30) - 1));
        const u64 bytes = n * stride, padded = (bytes + 3) & ~3ull;
        if (padded > raw_cap) {
            CUDA_TRY(cudaStreamSynchronize(st));
            if (d_raw) CUDA_TRY(cudaFree(d_raw));
            d_raw = nullptr;
            raw_cap = 0;
            CUDA_TRY(cudaMalloc((void**)&d_raw, padded + 256));
            raw_cap = padded;
        }
        double t0 = now_s();
        CUDA_TRY(cudaMemcpyAsync(d_raw, rec, bytes, cudaMemcpyHostToDevice, st));
        k_las_decode<<<div_up(n, LAS_DEC_THREADS), LAS_DEC_THREADS, 0, st>>>(d_raw, (u32)n, stride, colour_off, scale[0], scale[1],
                                                                             scale[2], offset[0], offset[1], offset[2],
                                                                             d_in + n_in);
        const int c = cls("las_decode");
        stats[c].launches += 1;
        stats[c].units += n;
        stats[c].bytes += bytes + 16 * n;
        launches += 1;
        CUDA_TRY(cudaStreamSynchronize(st));   // the caller may reuse `rec`
        h2d_s += now_s() - t0;
        n_in += n;
    }

    void add(const pcb_point* pts, u64 n, u32 batch_size, bool device_ptr, bool one_batch) {
        CUDA_TRY(cudaSetDevice(device));
        note_points(n, batch_size, one_batch);
        if (n && device_ptr && n_in == 0 && cap_in == 0) {
            // resident input: alias the caller's buffer instead of copying it (it must stay valid and
            // unchanged until the build has run)
            d_in = (PcbPoint*)pts;
            in_borrowed = true;
            cap_in = n;
            n_in = n;
        } else if (n) {
            if (n_in + n > cap_in) reserve(std::min<u64>(std::max<u64>(n_in + n, cap_in + cap_in / 2), (1ull << 30) - 1));
            double t0 = now_s();
            CUDA_TRY(cudaMemcpyAsync(d_in + n_in, pts, n * sizeof(PcbPoint),
                                     device_ptr ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
            if (!device_ptr) {
                CUDA_TRY(cudaStreamSynchronize(st));
                h2d_s += now_s() - t0;
            }
            n_in += n;
        }
        if (one_batch) end_run();
    }

    // Points that already carry their global batch ordinal (multi-GPU: the exchange keeps the ordinals of the
    // undivided stream).  Device pointers, borrowed until build() returns; must be the only add on the handle.
    void add_batched(const pcb_point* d_pts, const u32* d_batch, u64 n) {
        CUDA_TRY(cudaSetDevice(device));
        if (n_in != 0 || batches != 0 || run_open) throw Error{PCB_ERR_ARG, "pcb_add_points_device_batched must be the only add"};
        if (n >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "more than 2^30-1 points on one device"};
        batches = 1;            // at least one batch was processed: level 0 is visited (converter.rs:111,120)
        if (!n) return;
        d_in = (PcbPoint*)d_pts;
        in_borrowed = true;
        cap_in = n_in = n;
        d_bat_in = (u32*)d_batch;
        bat_owned = false;
        segs.push_back(BatchSeg{0u, 0u, 1u, 0u});
    }

    void add_arrivals(u32 level, const pcb_point* d_pts, const u32* d_bat, u64 n, const u32* d_cell_start,
                      const int32_t* d_cell_idx, u64 ncell, const float* bbox6, u64 count_points) {
        CUDA_TRY(cudaSetDevice(device));
        if (n_in != 0 || batches != 0 || run_open || ext.set) throw Error{PCB_ERR_ARG, "pcb_add_arrivals must be the only add"};
        if (n >= (1ull << 30) || ncell > n) throw Error{PCB_ERR_RANGE, "bad arrival counts"};
        if (level >= 32) throw Error{PCB_ERR_RANGE, "octree deeper than 31 levels"};
        ext.set = true;
        ext.level = level;
        ext.n = (u32)n;
        ext.ncell = (u32)ncell;
        ext.pts = (const PcbPoint*)d_pts;
        ext.bat = d_bat;
        ext.cell_start = d_cell_start;
        ext.cell_idx = d_cell_idx;
        ext.has_box = bbox6 != nullptr;
        if (bbox6)
            for (int a = 0; a < 3; a++) { ext.lo[a] = bbox6[a]; ext.hi[a] = bbox6[3 + a]; }
        ext.count_points = count_points;
        batches = 1;
        n_in = count_points;   // points_added of this handle
    }

    void build(Metadata* meta, NodeSource* existing);
    void finish_build(double t_start);
    void run_levels(Metadata* meta, NodeSource* existing, u32 h0, PcbPoint* pts, u32* bat, u32* cell_start, int* cell_idx,
                    u32 n0, u32 ncell);
    template <bool PAIRS>
    void build_level(u32 h, PcbPoint*& pts, u32*& bat, u32*& cell_start, int*& cell_idx, u32& n, u32& ncell,
                     NodeSource* existing);
    bool pairs_mode_for(u32 n, u32 ncell, const LevelGeom& g);
    // Level-0 arrivals per wave.  With node images going to the host a quarter of the run per wave lets the
    // downloads overlap the build; device-only runs have nothing to overlap and keep one wave (fewer, larger
    // launches) until the per-level arrays (72 B per arrival, twice) would crowd the device.
    u64 wave_points(u64 n0) const {
        static const u64 forced = pcb_tuning_env("PCB_WAVE_POINTS") ? strtoull(pcb_tuning_env("PCB_WAVE_POINTS"), nullptr, 10) : 0;
        if (stop_level != 0xFFFFFFFFu) return 1ull << 40;   // a partial build keeps ONE set of pending arrivals: one wave
        if (forced) return forced;
        if (flags & PCB_FLAG_DEVICE_ONLY) return 400ull << 20;
        return std::min<u64>(std::max<u64>(n0 / 4, 16ull << 20), 400ull << 20);
    }
};

static int bits_for(u32 range) {
    int b = 1;
    while ((1u << b) < range) b++;
    return b;
}
// bits needed to hold any value in [0, count)
static int bits_count(u64 count) {
    int b = 0;
    while (b < 64 && (1ull << b) < count) b++;
    return b;
}

// hex window of one cell: the cell spans 2*dim hex radii per axis; columns are sqrt(3) R apart, rows
// 1.5 R, layers R (hex.rs:55-65).  The window origin is the key of the min corner minus a margin that
// absorbs rounding and the odd-row column shift; k_hexkey verifies every key falls inside.
static LevelGeom make_geom(float max_cell_size, u32 dim, u32 h) {
    LevelGeom g;
    g.S = max_cell_size / (float)(1u << h);        // metadata.rs:91-93
    g.S_child = max_cell_size / (float)(1u << (h + 1 < 32 ? h + 1 : 31));
    g.sub = g.S / (float)dim;                      // metadata.rs:95-97
    g.R = g.sub / 2.0f;                            // cell.rs:277
    u32 rx = (u32)std::ceil(2.0 * dim / 1.7320508) + 3, ry = (u32)std::ceil(2.0 * dim / 1.5) + 3, rz = 2 * dim + 3;
    g.bx = bits_for(rx + 8);
    g.by = bits_for(ry + 8);
    g.bz = bits_for(rz + 8);
    g.mx = (int)(((1u << g.bx) - rx) / 2);
    g.my = (int)(((1u << g.by) - ry) / 2);
    g.mz = (int)(((1u << g.bz) - rz) / 2);
    g.hexbits = g.bx + g.by + g.bz;
    g.half_sqrt3 = 1.73205080757f / 2.0f;          // hex.rs:60  SQRT_3 / 2.0
    g.row_pitch = (g.R * 3.0f) / 2.0f;             // hex.rs:61  cell_radius * 3.0 / 2.0
    const volatile float dx = g.R * 1.73205080757f, dy = (-g.R) * 1.73205080757f;   // the f32 divisors of hex.rs:68-69
    g.inv.x = 1.0 / (double)dx;
    g.inv.y = 1.0 / (double)dy;
    g.inv.z = 1.0 / (double)g.R;
    g.inv.third = 1.0 / 3.0;
    return g;
}

static const bool g_force_pairs = pcb_tuning_env("PCB_FORCE_PAIRS") != nullptr;  // test hook: exercise the payload sort

void BuilderImpl::build(Metadata* meta, NodeSource* existing) {
    CUDA_TRY(cudaSetDevice(device));
    end_run();
    const double t_start = now_s();
    g_trace_t0 = g_trace_last = t_start;
    limit = meta->cell_point_overflow_limit;
    dim = meta->sub_grid_dimension;
    max_cell_size = meta->max_cell_size;
    if (dim == 0 || !(max_cell_size > 0.0f) || !std::isfinite(max_cell_size))
        throw Error{PCB_ERR_ARG, "bad config (sub_grid_dimension / max_cell_size)"};
    if (batches == 0) return;                                   // nothing was ever added
    if (!d_scr) CUDA_TRY(cudaMallocAsync((void**)&d_scr, sizeof(Scratch), st));
    if (!arenas) arenas = g_arenas.acquire(device);
    if (ext.set) {
        // ---- arrivals of one level handed in with their CSR: no stage 0
        if (ext.level == 0) meta->hierarchies = std::max<uint32_t>(meta->hierarchies, 1);
        if (ext.has_box && ext.count_points) {
            if (meta->number_of_points == 0) {              // converter.rs:98-99: first points REPLACE the box
                for (int a = 0; a < 3; a++) { meta->bbox_min[a] = ext.lo[a]; meta->bbox_max[a] = ext.hi[a]; }
            } else {                                        // converter.rs:100-102
                for (int a = 0; a < 3; a++) {
                    meta->bbox_min[a] = std::fmin(meta->bbox_min[a], ext.lo[a]);
                    meta->bbox_max[a] = std::fmax(meta->bbox_max[a], ext.hi[a]);
                }
            }
        }
        meta->number_of_points += ext.count_points;         // converter.rs:108
        if (ext.n) {
            Scratch init;
            memset(&init, 0, sizeof init);
            *h_scr = init;
            CUDA_TRY(cudaMemcpyAsync(d_scr, h_scr, sizeof init, cudaMemcpyHostToDevice, st));
            arena_prepare(2, 18ull * ext.n + (64ull << 20));
            cur_arena = 0;
            last_mult = ext.prism_load;   // the record filter's prediction for level ext.level (build_level)
            run_levels(meta, existing, ext.level, (PcbPoint*)ext.pts, (u32*)ext.bat, (u32*)ext.cell_start, (int*)ext.cell_idx,
                       ext.n, ext.ncell);
        }
        finish_build(t_start);
        return;
    }
    meta->hierarchies = std::max<uint32_t>(meta->hierarchies, 1);  // converter.rs:111,120: level 0 always visited
    if (n_in == 0) return;
    const u32 n0 = (u32)n_in;
    const float S0 = max_cell_size;
    // Arena 0 first serves the level-0 staging (keys A/B, histograms, look-back state: ~17 B per point); arenas 0/1
    // are sized for the largest wave once the level-0 cells are known (below).  Sizes only ever grow, so a process
    // that converts clouds of one size allocates once.
    arena_prepare(0, std::max<u64>(arenas->a[0].cap, 17ull * n_in + (64ull << 20)));
    arena_prepare(2, 18ull * n_in + (64ull << 20));
    cur_arena = 0;

    // ---- stage 0a: bounding box + validation
    Scratch init;
    memset(&init, 0, sizeof init);
    for (int a = 0; a < 3; a++) { init.bbox[a] = INT_MAX; init.bbox[3 + a] = INT_MIN; }
    *h_scr = init;
    CUDA_TRY(cudaMemcpyAsync(d_scr, h_scr, sizeof init, cudaMemcpyHostToDevice, st));
    const int grid_big = 148 * 8;
    pb();
    k_bbox_check<<<grid_big, 256, 0, st>>>(d_in, n0, S0, d_scr->bbox, &d_scr->err);
    pe("bbox_check", 1, n0, 16ull * n0);
    fetch_scratch();
    check_err();
    trace_mark(st, "bbox", 0, n0);
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) { lo[a] = ord2f_host(h_scr->bbox[a]); hi[a] = ord2f_host(h_scr->bbox[3 + a]); }
    if (meta->number_of_points == 0) {              // converter.rs:98-99: first points REPLACE the box
        for (int a = 0; a < 3; a++) { meta->bbox_min[a] = lo[a]; meta->bbox_max[a] = hi[a]; }
    } else {                                        // converter.rs:100-102
        for (int a = 0; a < 3; a++) {
            meta->bbox_min[a] = std::fmin(meta->bbox_min[a], lo[a]);
            meta->bbox_max[a] = std::fmax(meta->bbox_max[a], hi[a]);
        }
    }
    meta->number_of_points += n_in;                 // converter.rs:108

    // ---- stage 0b-d: level-0 arrivals in (cell, input index) order
    L0Box box;
    int o3[3];
    u32 d3[3];
    for (int a = 0; a < 3; a++) {
        o3[a] = (int)std::floor(lo[a] / S0);
        d3[a] = (u32)((int)std::floor(hi[a] / S0) - o3[a] + 1);
    }
    box.ox = o3[0]; box.oy = o3[1]; box.oz = o3[2];
    box.nx = d3[0]; box.ny = d3[1]; box.nz = d3[2];
    const u64 vol = (u64)box.nx * box.ny * box.nz;   // < 2^63: every axis spans < 2^21 cells
    const int cb0 = bits_count(vol);
    int pb0 = std::max(1, bits_count(n0));
    const bool pairs0 = g_force_pairs || cb0 + pb0 > 64;
    if (pairs0) pb0 = 0;
    u32 ncell_ub = (u32)std::min<u64>(n0, vol);
    arena_prepare(3, 20ull * n0 + 16ull * ncell_ub + (1ull << 20));
    u32* cell_start = dalloc_from<u32>(3, (u64)ncell_ub + 1);
    int* cell_idx = dalloc_from<int>(3, 3ull * ncell_ub);
    PcbPoint* pts = dalloc_from<PcbPoint>(3, n0);
    u32* bat = dalloc_from<u32>(3, n0);
    BatchSeg* d_segs = dalloc<BatchSeg>(segs.size());
    CUDA_TRY(cudaMemcpyAsync(d_segs, segs.data(), segs.size() * sizeof(BatchSeg), cudaMemcpyHostToDevice, st));
    static const bool l0_sort_forced = pcb_tuning_env("PCB_L0_SORT") != nullptr;   // test hook: the general (sort) path
    if (vol <= L0S_MAX_BINS && !pairs0 && !l0_sort_forced) {
        // ---- few level-0 cells: dense histogram -> first slot per cell -> one stable split pass
        const u32 nb = (u32)vol, ntiles = div_up(n0, L0S_TILE);
        const int dbits = std::max(1, bits_count(vol));
        const u64 words = 2ull * nb + (u64)ntiles * nb;
        u32* tmp = dalloc<u32>(words);
        CUDA_TRY(cudaMemsetAsync(tmp, 0, 4 * words, st));
        u32 *lhist = tmp, *dbase = tmp + nb, *status = tmp + 2ull * nb;
        const int smem = L0S_TILE * 22 + (L0S_WARPS + 2) * (int)nb * 4;
        CUDA_TRY(cudaFuncSetAttribute(k_l0_split<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, L0SplitSmem::bytes));
        CUDA_TRY(cudaFuncSetAttribute(k_l0_split<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        pb();
        k_route_hist<<<grid_big, 256, 0, st>>>(d_in, n0, S0, box, nb, lhist);
        k_l0_cells<<<1, L0S_MAX_BINS, 0, st>>>(lhist, nb, box, n0, dbase, cell_start, cell_idx, &d_scr->ncell);
        pe("l0_hist", 2, n0, 16ull * n0);
        pb();
        k_l0_split<false><<<ntiles, L0S_THREADS, smem, st>>>(d_in, n0, S0, box, nb, dbits, dbase, d_segs, (u32)segs.size(),
                                                             d_bat_in, pts, bat, status, &d_scr->err, nullptr, RouteDst(), 0, 1);
        pe("l0_split", 1, n0, 36ull * n0);
        dfree(tmp);
    } else {
        const RadixPlan plan0 = radix_plan(pb0, cb0);
        u64 *kA = dalloc<u64>(n0), *kB = dalloc<u64>(n0);
        u32 *vA = pairs0 ? dalloc<u32>(n0) : nullptr, *vB = pairs0 ? dalloc<u32>(n0) : nullptr;
        RadixTemp rt;
        const u64 rwords = RadixTemp::ghist_words() + RadixTemp::status_words(n0, plan0);
        rt.ghist = dalloc<u32>(rwords);
        rt.status = rt.ghist + RadixTemp::ghist_words();
        rt.err = &d_scr->err;
        CUDA_TRY(cudaMemsetAsync(rt.ghist, 0, rwords * 4, st));
        pb();
        k_l0_key<<<grid_big, 256, 0, st>>>(d_in, n0, S0, box, pb0, plan0, kA, vA, rt.ghist);
        pe("l0_key", 1, n0, 24ull * n0);
        int cur = pairs0 ? radix_sort<true>(kA, vA, kB, vB, n0, plan0, rt, st, *this)
                         : radix_sort<false>(kA, vA, kB, vB, n0, plan0, rt, st, *this);
        u64* ks = cur ? kB : kA;
        u32* vs = cur ? vB : vA;
        trace_mark(st, "l0_sort", 0, n0);
        void* scan_tmp = dalloc<unsigned char>(chain_scan_temp_bytes<256, 8>(n0));
        L0HeadOp hop{ks, n0, pb0, box, cell_start, cell_idx, &d_scr->ncell};
        pb();
        int nl = chain_scan<L0HeadOp, 256, 8>(hop, n0, scan_tmp, st);
        pe("l0_heads", nl, n0, 8ull * n0);
        pb();
        k_l0_gather<<<div_up(n0, 256), 256, 0, st>>>(d_in, ks, vs, n0, pb0, d_segs, (u32)segs.size(), d_bat_in, pts, bat);
        pe("l0_gather", 1, n0, 44ull * n0);
        dfree(kA); dfree(kB); dfree(vA); dfree(vB); dfree(rt.ghist); dfree(scan_tmp);
    }
    fetch_scratch();
    u32 ncell = h_scr->ncell;
    trace_mark(st, "l0_gather", 0, n0);
    dfree(d_segs);

    run_levels(meta, existing, 0, pts, bat, cell_start, cell_idx, n0, ncell);
    dfree(pts); dfree(bat); dfree(cell_start); dfree(cell_idx);
    finish_build(t_start);
}

void BuilderImpl::finish_build(double t_start) {
    CUDA_TRY(cudaStreamSynchronize(st));
    build_s += now_s() - t_start;
    double t1 = now_s();
    CUDA_TRY(cudaStreamSynchronize(copy_st));
    d2h_s += now_s() - t1;
    for (auto& l : levels)
        if (l.h_image && l.d_image) {
            dfree(l.d_image);
            l.d_image = nullptr;
        }
    prof_collect();
}

// The levels from h0 down for the cells of level h0 given by (pts, bat, cell_start, cell_idx): arrivals in (cell, tau)
// order, borrowed.  Cells of one level never interact (converter.rs:124-134: a cell only receives from its parent), so
// the forest is built a contiguous group of them ("wave") at a time: the node images of one wave travel to the host
// while the next wave is built, and the per-level arrays are sized for a wave, not for the run.
void BuilderImpl::run_levels(Metadata* meta, NodeSource* existing, u32 h0, PcbPoint* pts, u32* bat, u32* cell_start,
                             int* cell_idx, u32 n0, u32 ncell) {
    u32* const h_cs = (u32*)host_tmp(4ull * ((u64)ncell + 1));
    publish(h_cs, cell_start, 4ull * ((u64)ncell + 1));
    CUDA_TRY(cudaStreamSynchronize(st));
    const u64 wave_target = wave_points(n0);
    std::vector<u32> wave_end;   // exclusive cell ordinal per wave
    u64 wave_max = 0;
    {   // equal shares of the arrivals, cut at the nearest cell boundary
        const u32 nw = (u32)std::min<u64>(ncell, std::max<u64>(1, ((u64)n0 + wave_target - 1) / wave_target));
        u32 c = 0;
        for (u32 w = 1; w <= nw && c < ncell; w++) {
            const u64 goal = (u64)n0 * w / nw;
            u32 e = c + 1;
            while (e < ncell && h_cs[e] < goal && (h_cs[e + 1] <= goal || goal - h_cs[e] > h_cs[e + 1] - goal)) e++;
            if (w == nw) e = ncell;
            wave_end.push_back(e);
            wave_max = std::max<u64>(wave_max, h_cs[e] - h_cs[c]);
            c = e;
        }
        if (wave_end.empty() || wave_end.back() != ncell) {
            wave_end.push_back(ncell);
            wave_max = std::max<u64>(wave_max, (u64)h_cs[ncell] - h_cs[c]);
        }
    }
    arena_prepare(0, std::max<u64>(arenas->a[0].cap, 72ull * wave_max + (64ull << 20)));
    arena_prepare(1, std::max<u64>(arenas->a[1].cap, 72ull * wave_max + (64ull << 20)));
    u32 hmax = h0, c0 = 0;
    for (u32 c1 : wave_end) {
        u32 n = h_cs[c1] - h_cs[c0], wcell = c1 - c0, h = h0;
        PcbPoint* wpts = pts + h_cs[c0];
        u32* wbat = bat + h_cs[c0];
        int* widx = cell_idx + 3ull * c0;
        u32* wstart = cell_start;
        if (wave_end.size() > 1) {
            // the wave's first level (h0) resets arena h0 & 1 and reads its inputs from the other one
            cur_arena = (int)((h0 & 1u) ^ 1u);
            arenas->a[cur_arena].off = 0;
            wstart = dalloc<u32>((u64)wcell + 1);
            k_wave_starts<<<div_up(wcell + 1, 256), 256, 0, st>>>(cell_start + c0, h_cs[c0], wcell + 1, wstart);
        }
        bool first = true;                    // the wave's first inputs are slices of the run's arrays: not freed
        inputs_borrowed = true;
        while (n && h < stop_level) {
            if (h >= 32) throw Error{PCB_ERR_RANGE, "octree deeper than 31 levels (reference overflows 2u32.pow, metadata.rs:92)"};
            meta->hierarchies = std::max<uint32_t>(meta->hierarchies, h + 1);  // converter.rs:141-144
            if (pairs_mode_for(n, wcell, make_geom(max_cell_size, dim, h)))
                build_level<true>(h, wpts, wbat, wstart, widx, n, wcell, existing);
            else
                build_level<false>(h, wpts, wbat, wstart, widx, n, wcell, existing);
            inputs_borrowed = false;
            first = false;
            h++;
        }
        if (n && h >= stop_level) {
            // partial build (single wave): the arrivals of level `stop_level` stay on the device for the caller
            pend_pts = wpts; pend_bat = wbat; pend_cell_start = wstart; pend_cell_idx = widx;
            pend_n = n; pend_ncell = wcell; pend_level = h;
            pend_owned = !first;
        } else if (!first) {
            dfree(wpts); dfree(wbat); dfree(wstart); dfree(widx);   // (empty) inputs of the level that was not built
        }
        inputs_borrowed = false;
        hmax = std::max(hmax, h);
        c0 = c1;
        trace_mark(st, "wave done", hmax, c1);
    }
    levels_built = hmax - h0;
    trace_mark(st, "levels done");
}

bool BuilderImpl::pairs_mode_for(u32 n, u32 ncell, const LevelGeom& g) {
    const int cb = bits_count(ncell), pbits = std::max(1, bits_count(n));
    if (cb + g.hexbits > 64) throw Error{PCB_ERR_RANGE, "too many cells on one level"};
    return g_force_pairs || cb + g.hexbits + pbits > 64;
}

// One level of the hierarchy for ALL batches (converter.rs:114-139).  In: arrivals sorted by (cell, tau).
// Out: this level's node images + the next level's arrivals (same invariant).
template <bool PAIRS>
void BuilderImpl::build_level(u32 h, PcbPoint*& pts, u32*& bat, u32*& cell_start, int*& cell_idx, u32& n, u32& ncell,
                              NodeSource* existing) {
    const LevelGeom g = make_geom(max_cell_size, dim, h);
    const u32 n_new = n;  // arrivals proper (incumbents of merge mode are added below)
    bool own_pts = !inputs_borrowed;          // level 0 of a wave reads slices of the run's level-0 arrays
    const bool own_cells = !inputs_borrowed;
    cur_arena = (int)(h & 1u);          // the other arena holds this level's inputs (written by the previous level)
    arenas->a[cur_arena].off = 0;
    arrivals += n_new;

    // ---- merge mode: read the node files of the cells that receive arrivals (converter.rs:160-207)
    unsigned char* ex_state = nullptr;
    u32 *ex_m0 = nullptr, *ex_bucket = nullptr, *ex_rank = nullptr;
    PcbPoint* ex_pts = nullptr;
    u32 nex = 0;
    // pinned node table of this level: cell indices, node offsets, node sizes (recycled when the level is done)
    const u64 tab_idx = (12ull * ncell + 15) & ~15ull;
    u64 tab_cap = 0;
    unsigned char* const tab = (unsigned char*)g_pinned.get(tab_idx + 16ull * ncell + 32, &tab_cap);
    struct TabGuard {
        unsigned char* p; u64 cap;
        ~TabGuard() { g_pinned.put(p, cap); }
    } tab_guard{tab, tab_cap};
    int* const hidx = (int*)tab;
    if (existing) {
        publish(hidx, cell_idx, 12ull * ncell);
        CUDA_TRY(cudaStreamSynchronize(st));
        // node files are read and decoded by a pool of host threads, then packed into pinned staging in parallel
        // (C3 = 100 M points merged into 100 M: 1.6 GB in ~5 800 node files; one thread took 2.3 s for this)
        std::vector<ExistingNode> ens(ncell);
        std::vector<unsigned char> found(ncell, 0);
        host_parallel_for(ncell, [&](u32 c) { found[c] = existing->load(h, &hidx[3 * c], &ens[c]) ? 1 : 0; });
        trace_mark(st, "merge_load", h, ncell);
        std::vector<u32> inc_start(ncell + 1, 0);
        std::vector<u64> exl_start(ncell + 1, 0);
        bool any = false;
        for (u32 c = 0; c < ncell; c++) {
            inc_start[c + 1] = inc_start[c];
            exl_start[c + 1] = exl_start[c];
            if (!found[c]) continue;
            any = true;
            if ((u64)inc_start[c] + ens[c].grid.size() >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "level too large for one device"};
            inc_start[c + 1] += (u32)ens[c].grid.size();
            for (auto& e : ens[c].entries) {
                int dx = e.child[0] - 2 * hidx[3 * c], dy = e.child[1] - 2 * hidx[3 * c + 1],
                    dz = e.child[2] - 2 * hidx[3 * c + 2];
                if ((unsigned)dx > 1u || (unsigned)dy > 1u || (unsigned)dz > 1u)
                    throw Error{PCB_ERR_FORMAT, "existing node has an overflow entry outside its 8 children"};
                if (!e.none) exl_start[c + 1] += e.list.size();
            }
        }
        if (any) {
            const u32 ninc = inc_start[ncell];
            if (exl_start[ncell] >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "level too large for one device"};
            nex = (u32)exl_start[ncell];
            if ((u64)n + ninc >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "level too large for one device"};
            unsigned char* hstate = (unsigned char*)host_tmp(8ull * ncell);
            u32* hm0 = (u32*)host_tmp(32ull * ncell);
            pcb_point* h_inc = (pcb_point*)host_tmp(16ull * ninc);
            pcb_point* h_exl = (pcb_point*)host_tmp(16ull * nex);
            u32* hbucket = (u32*)host_tmp(4ull * nex);
            u32* hrank = (u32*)host_tmp(4ull * nex);
            u32* h_inc_start = (u32*)host_tmp(4ull * (ncell + 1));
            memcpy(h_inc_start, inc_start.data(), 4ull * (ncell + 1));
            trace_mark(st, "merge_alloc", h, ninc);
            host_parallel_for(ncell, [&](u32 c) {
                for (int o = 0; o < 8; o++) { hstate[8 * c + o] = PCB_EX_ABSENT; hm0[8 * c + o] = 0; }
                if (!found[c]) return;
                if (!ens[c].grid.empty())
                    memcpy(h_inc + inc_start[c], ens[c].grid.data(), 16ull * ens[c].grid.size());
                u64 at = exl_start[c];
                for (auto& e : ens[c].entries) {
                    const u32 k = 8 * c + (u32)((e.child[0] - 2 * hidx[3 * c]) * 4 + (e.child[1] - 2 * hidx[3 * c + 1]) * 2 +
                                                (e.child[2] - 2 * hidx[3 * c + 2]));
                    if (e.none) {
                        hstate[k] = PCB_EX_NONE;
                    } else {
                        hstate[k] = PCB_EX_SOME;
                        hm0[k] = (u32)e.list.size();
                        if (!e.list.empty()) memcpy(h_exl + at, e.list.data(), 16ull * e.list.size());
                        for (u32 r = 0; r < e.list.size(); r++) { hbucket[at + r] = k; hrank[at + r] = r; }
                        at += e.list.size();
                    }
                }
                ExistingNode().grid.swap(ens[c].grid);   // release as we go
                ens[c].entries.clear();
            });
            trace_mark(st, "merge_pack", h, nex);
            // every source below is pinned and lives until the end of the run: no host synchronisation needed
            ex_state = dalloc<unsigned char>(8ull * ncell);
            ex_m0 = dalloc<u32>(8ull * ncell);
            CUDA_TRY(cudaMemcpyAsync(ex_state, hstate, 8ull * ncell, cudaMemcpyHostToDevice, st));
            CUDA_TRY(cudaMemcpyAsync(ex_m0, hm0, 32ull * ncell, cudaMemcpyHostToDevice, st));
            if (nex) {
                ex_pts = dalloc<PcbPoint>(nex);
                ex_bucket = dalloc<u32>(nex);
                ex_rank = dalloc<u32>(nex);
                CUDA_TRY(cudaMemcpyAsync(ex_pts, h_exl, 16ull * nex, cudaMemcpyHostToDevice, st));
                CUDA_TRY(cudaMemcpyAsync(ex_bucket, hbucket, 4ull * nex, cudaMemcpyHostToDevice, st));
                CUDA_TRY(cudaMemcpyAsync(ex_rank, hrank, 4ull * nex, cudaMemcpyHostToDevice, st));
            }
            if (ninc) {
                u32* d_inc_start = dalloc<u32>(ncell + 1);
                PcbPoint* d_inc = dalloc<PcbPoint>(ninc);
                CUDA_TRY(cudaMemcpyAsync(d_inc_start, h_inc_start, 4ull * (ncell + 1), cudaMemcpyHostToDevice, st));
                CUDA_TRY(cudaMemcpyAsync(d_inc, h_inc, 16ull * ninc, cudaMemcpyHostToDevice, st));
                PcbPoint* mp = dalloc<PcbPoint>((u64)n + ninc);
                u32* mb = dalloc<u32>((u64)n + ninc);
                pb();
                k_merge_arrivals<<<div_up(n, 256), 256, 0, st>>>(pts, bat, n, cell_start, ncell, d_inc_start, mp, mb);
                k_merge_incumbents<<<div_up(ninc, 256), 256, 0, st>>>(d_inc, ninc, d_inc_start, ncell, cell_start, mp, mb);
                k_merge_starts<<<div_up(ncell + 1, 256), 256, 0, st>>>(cell_start, d_inc_start, ncell);
                pe("merge_incumbents", 3, (u64)n + ninc, 40ull * ((u64)n + ninc));
                if (own_pts) { dfree(pts); dfree(bat); }
                dfree(d_inc_start); dfree(d_inc);
                pts = mp; bat = mb;
                own_pts = true;
                n += ninc;
            }
        }
    }

    // ---- key layout and radix plan of this level
    KeyLayout lay;
    lay.cb = bits_count(ncell);
    lay.hexbits = g.hexbits;
    lay.pb = PAIRS ? 0 : std::max(1, bits_count(n));
    if (lay.cb + lay.hexbits + lay.pb > 64) throw Error{PCB_ERR_RANGE, "too many cells on one level"};
    // ---- record filter (level.cuh): worth it when a prism sees many arrivals.  Level 0 of a big wave: assumed;
    // deeper: predicted from the level above (the prisms shrink 4x in area, 8x in volume per level).
    static const bool filter_off = pcb_tuning_env("PCB_NO_FILTER") != nullptr;                         // tuning / test hooks
    static const int filter_force = pcb_tuning_env("PCB_FILTER_SHIFT") ? atoi(pcb_tuning_env("PCB_FILTER_SHIFT")) : -1;
    u32 fshift = 0;   // early share of a cell = its first len >> fshift arrivals; 0 = no filter on this level
    static const double filter_min = pcb_tuning_env("PCB_FILTER_MIN") ? atof(pcb_tuning_env("PCB_FILTER_MIN")) : 64.0;
    if (!PAIRS && !existing && !filter_off && lay.cb + lay.hexbits < FILTER_GID_BITS) {
        const double m_pred = h == 0 ? 1e9 : last_mult / 4.0;
        if (filter_force > 0) fshift = (u32)filter_force;
        else if (n >= (1u << 21) && m_pred >= filter_min) fshift = m_pred >= 256.0 ? 4 : m_pred >= 64.0 ? 3 : 2;
    }

    CellGeo* geo = dalloc<CellGeo>(ncell);
    uint2* d2oct = dalloc<uint2>(n);                 // contest inputs by arrival position, written by k_hexkey
    u32* win_src = dalloc<u32>(n);
    u32* win_end = dalloc<u32>(ncell);
    const u64 tmp_bytes = std::max<u64>(chain_scan_temp_bytes<64, 8>(n), chain_scan_temp_bytes<256, 4>(ncell));
    void* scan_tmp = dalloc<unsigned char>(tmp_bytes);
    u64 *kA = nullptr, *kB = nullptr, *em = nullptr;
    u32 *vA = nullptr, *vB = nullptr;
    u32* rt_mem = nullptr;
    uint4* tinfo = nullptr;
    u32* seg_base = nullptr;
    static const int contest_minb = pcb_tuning_env("PCB_CONTEST_MINB") ? atoi(pcb_tuning_env("PCB_CONTEST_MINB")) : 3;  // tuning hook
    auto run_contest = [&](const ContestOp<PAIRS>& cop, u32 cnt) {   // cnt: upper bound when cop.n_dev is set
        pb();
        const u32* nd = cop.n_dev;
        int nl = contest_minb == 4   ? chain_scan<ContestOp<PAIRS>, 256, 8, 4>(cop, cnt, scan_tmp, st, nd)
                 : contest_minb == 6 ? chain_scan<ContestOp<PAIRS>, 128, 8, 6>(cop, cnt, scan_tmp, st, nd)
                 : contest_minb == 8 ? chain_scan<ContestOp<PAIRS>, 128, 8, 8>(cop, cnt, scan_tmp, st, nd)
                 : contest_minb == 12 ? chain_scan<ContestOp<PAIRS>, 64, 8, 12>(cop, cnt, scan_tmp, st, nd)
                                     : chain_scan<ContestOp<PAIRS>, 256, 8, 3>(cop, cnt, scan_tmp, st, nd);
        pe("contest", nl, cnt, (PAIRS ? 28ull : 24ull) * cnt);
    };
    // sort plan of `cnt` keys spread over the level's cells: segmented when the cell bits would cost extra passes
    // (radix.cuh): runs of 2^segk consecutive cells are sorted independently inside the same launches.  Taken when
    // segments average >= PCB_SEG_MIN keys (tiles never straddle a segment: small segments mean partly filled tiles).
    static const bool seg_off = pcb_tuning_env("PCB_NO_SEG") != nullptr;                                       // tuning / test hooks
    static const u64 seg_min = pcb_tuning_env("PCB_SEG_MIN") ? strtoull(pcb_tuning_env("PCB_SEG_MIN"), nullptr, 10) : 4ull * RADIX_TILE;
    int segk = 31;
    u32 nseg = 1;
    auto plan_for = [&](u32 cnt) {
        segk = 31;
        nseg = 1;
        const int db = radix_digit_bits();
        const int top = div_up(lay.hexbits + 1, db) * db;   // room left in the top digit
        const int k = top - lay.hexbits;
        if (!seg_off && lay.cb > k && div_up(lay.hexbits + k, db) < div_up(lay.hexbits + lay.cb, db) &&
            (u64)cnt >= seg_min * div_up(ncell, 1u << k)) {
            segk = k;
            nseg = div_up(ncell, 1u << k);
        }
        return radix_plan(lay.pb, (nseg > 1 ? segk : lay.cb) + lay.hexbits);
    };
    // radix scratch (histograms + look-back status) and, for a segmented sort, the tile table over `starts`
    // (`table`: use the tile table even for a single segment - the keys' count is then only known on the device and
    // `cnt` is an upper bound; the table gives every tile its real extent)
    auto radix_setup = [&](RadixTemp& rt, const RadixPlan& plan, u32 cnt, const u32* starts, bool table = false) {
        table = table || nseg > 1;
        const u64 hwords = table ? RadixTemp::ghist_words_seg(nseg, plan) : RadixTemp::ghist_words();
        const u64 rwords = hwords + RadixTemp::status_words(cnt, plan, table ? nseg : 0);
        rt_mem = dalloc<u32>(rwords);
        rt.ghist = rt_mem;
        rt.status = rt.ghist + hwords;
        rt.err = &d_scr->err;
        CUDA_TRY(cudaMemsetAsync(rt.ghist, 0, rwords * 4, st));
        if (table) {
            rt.nseg = nseg;
            rt.ntiles_seg = (u32)RadixTemp::seg_tiles_bound(cnt, nseg);
            tinfo = dalloc<uint4>(rt.ntiles_seg);
            seg_base = dalloc<u32>(nseg + 1);
            rt.tinfo = tinfo;
            k_seg_plan<<<1, 1024, 0, st>>>(starts, ncell, segk, nseg, seg_base);
            k_seg_tiles<<<div_up(rt.ntiles_seg, 256), 256, 0, st>>>(starts, ncell, segk, seg_base, nseg, rt.ntiles_seg, tinfo);
        }
    };
    const u32 hexgrid = std::min<u32>(148 * 8, div_up(n, 256));
    static const int hexkey_minb = pcb_tuning_env("PCB_HEXKEY_MINB") ? atoi(pcb_tuning_env("PCB_HEXKEY_MINB")) : 8;   // tuning hook
    auto launch_hexkey = [&](const RadixPlan& plan, u64* keys, u32* vals, u32* ghist, const u32* estart, u64* keys_e) {
#define PCB_HEXKEY(M)                                                                                                     \
    k_hexkey<PAIRS, M><<<std::min<u32>(148 * M, div_up(n, 256)), 256, 0, st>>>(pts, n, cell_start, ncell, geo, g, lay, plan, keys, vals, \
                                                                             d2oct, ghist, segk, nseg, estart, keys_e, &d_scr->err)
        if (hexkey_minb == 6) PCB_HEXKEY(6);
        else if (hexkey_minb == 5) PCB_HEXKEY(5);
        else PCB_HEXKEY(8);
#undef PCB_HEXKEY
    };
    pb();
    k_cellgeo<<<div_up(ncell, 256), 256, 0, st>>>(cell_idx, cell_start, ncell, g, geo, &d_scr->err);
    pe("hexkey", 1, ncell, 60ull * ncell);

    bool filtered = false;
    if (fshift) {
        // ---- round 1: the early arrivals of every cell
        u32* estart = dalloc<u32>((u64)ncell + 1);
        k_early_starts<<<1, 1024, 0, st>>>(cell_start, ncell, fshift, estart, &d_scr->aux);
        // the number of early arrivals stays on the device until round 1 is over (one host synchronisation less):
        // buffers, grids and the tile table are sized by its upper bound
        const u32 nE_ub = (u32)std::min<u64>(n, ((u64)n >> fshift) + ncell);
        const RadixPlan plan = plan_for(nE_ub);
        kA = dalloc<u64>(n);   // keys by arrival position, later em
        // candidate lists: one per warp of k_filter, each can hold every arrival that warp looks at
        const u64 fchunk = (((u64)n + hexgrid - 1) / hexgrid + 255ull) & ~255ull;
        const u32 nlists = hexgrid * 8, list_cap = (u32)(fchunk / 8);
        kB = dalloc<u64>((u64)nlists * list_cap);
        u64 *eA = dalloc<u64>(nE_ub), *eB = dalloc<u64>(nE_ub);
        uint4* tails = dalloc<uint4>(nE_ub);
        RadixTemp rt;
        pb();
        radix_setup(rt, plan, nE_ub, estart, true);
        launch_hexkey(plan, kA, vA, rt.ghist, estart, eA);
        pe("hexkey", 3, n, 32ull * n);
        trace_mark(st, "hexkey", h, n);
        int cur = radix_sort<PAIRS>(eA, nullptr, eB, nullptr, nE_ub, plan, rt, st, *this);
        em = kA;
        run_contest(ContestOp<PAIRS>{cur ? eB : eA, nullptr, nE_ub, lay, g, d2oct, geo, em, win_src, win_end, &d_scr->nwin,
                                     &d_scr->err, tails, nullptr, &d_scr->aux}, nE_ub);
        fetch_scratch();
        check_err();
        const u32 nwinE = h_scr->nwin, nE = h_scr->aux;
        trace_mark(st, "round1", h, nE);
        // predicted round-2 list: early winners + one candidate per (early arrivals per prism + 1) later arrivals
        const double s2_pred = nwinE + (double)(n - nE) / ((double)nE / std::max<u32>(nwinE, 1) + 1.0);
        static const bool filter_bail = pcb_tuning_env("PCB_FILTER_BAIL") != nullptr;   // test hook: give up after round 1
        if (!filter_bail && nwinE < (1u << 29) && (filter_force > 0 || s2_pred < 0.4 * n)) {   // (table capacity is 32 bits)
            // ---- filter: later arrivals against the early winners
            FilterTable T;
            u32 cap = 1024;
            int capbits = 10;
            while (cap < 3ull * nwinE) { cap <<= 1; capbits++; }   // load factor 1/6 .. 1/3
            T.mask = cap - 1;
            T.shift = 64 - capbits;
            T.slots = dalloc<unsigned long long>(cap);
            CUDA_TRY(cudaMemsetAsync(T.slots, 0, 8ull * cap, st));
            u32* ncand = dalloc<u32>(2ull * nlists);
            u32* loff = ncand + nlists;
            u64* wkeys = dalloc<u64>(std::max<u32>(nwinE, 1));
            pb();
            k_filter_table<<<std::min<u32>(148 * 8, div_up(std::max<u32>(nwinE, 1), 256)), 256, 0, st>>>(tails, &d_scr->nwin, T, lay.pb,
                                                                                                wkeys);
            k_filter<<<hexgrid, 256, 0, st>>>(kA, d2oct, n, cell_start, ncell, geo, estart, lay.pb, T, kB, ncand, em);
            k_filter_offsets<<<1, 1024, 0, st>>>(ncand, nlists, loff, &d_scr->aux);
            pe("filter", 3, n, 24ull * n);
            fetch_scratch();
            const u32 nS2 = nwinE + h_scr->aux;
            trace_mark(st, "filter", h, nS2);
            // ---- round 2: early winners, then the candidates in arrival order: stable sort on the cell + prism bits
            dfree(rt_mem); dfree(tinfo); dfree(seg_base);
            rt_mem = nullptr; tinfo = nullptr; seg_base = nullptr;
            segk = 31;
            nseg = 1;
            const RadixPlan plan2 = radix_plan(lay.pb, lay.hexbits + lay.cb);
            RadixTemp rt2;
            u64 *s2A = dalloc<u64>(nS2), *s2B = dalloc<u64>(nS2);
            pb();
            radix_setup(rt2, plan2, nS2, nullptr);
            CUDA_TRY(cudaMemcpyAsync(s2A, wkeys, 8ull * nwinE, cudaMemcpyDeviceToDevice, st));
            k_filter_gather<<<nlists, 256, 0, st>>>(kB, list_cap, ncand, loff, &d_scr->nwin, s2A);
            k_keys_hist<<<std::min<u32>(148 * 4, div_up(nS2, 256)), 256, 0, st>>>(s2A, nS2, plan2, rt2.ghist);
            pe("filter", 3, nS2, 24ull * nS2);
            cur = radix_sort<PAIRS>(s2A, nullptr, s2B, nullptr, nS2, plan2, rt2, st, *this);
            run_contest(ContestOp<PAIRS>{cur ? s2B : s2A, nullptr, nS2, lay, g, d2oct, geo, em, win_src, win_end, &d_scr->nwin,
                                         &d_scr->err, nullptr, estart, nullptr}, nS2);
            trace_mark(st, "round2", h, nS2);
            dfree(s2A); dfree(s2B); dfree(T.slots); dfree(ncand); dfree(wkeys);
            filtered = true;
        } else {
            dfree(kA); dfree(kB);
            dfree(rt_mem); dfree(tinfo); dfree(seg_base);
            kA = kB = nullptr; rt_mem = nullptr; tinfo = nullptr; seg_base = nullptr;
        }
        dfree(estart); dfree(eA); dfree(eB); dfree(tails);
    }
    if (!filtered) {
        // ---- stages 1-2: hex keys (+ the digit histograms of every pass)
        const RadixPlan plan = plan_for(n);
        kA = dalloc<u64>(n);
        kB = dalloc<u64>(n);
        vA = PAIRS ? dalloc<u32>(n) : nullptr;
        vB = PAIRS ? dalloc<u32>(n) : nullptr;
        RadixTemp rt;
        pb();
        radix_setup(rt, plan, n, cell_start);
        launch_hexkey(plan, kA, vA, rt.ghist, nullptr, nullptr);
        pe("hexkey", nseg > 1 ? 3 : 1, n, (PAIRS ? 36ull : 32ull) * n);
        trace_mark(st, "hexkey", h, n);

        // ---- stage 3: stable sort by (cell, Z, Y, X)
        int cur = radix_sort<PAIRS>(kA, vA, kB, vB, n, plan, rt, st, *this);
        u64* ks = cur ? kB : kA;
        u32* vs = cur ? vB : vA;
        em = cur ? kA : kB;   // the other key buffer is free now
        trace_mark(st, "sort", h, n);

        // ---- stage 4: contest
        run_contest(ContestOp<PAIRS>{ks, vs, n, lay, g, d2oct, geo, em, win_src, win_end, &d_scr->nwin, &d_scr->err, nullptr,
                                     nullptr, nullptr}, n);
        trace_mark(st, "contest", h, n);
    }
    int nl = 0;

    // ---- stages 5-6: segmented 8-way split of the losers (ranks inside their (cell, octant) buckets)
    u32* rank = dalloc<u32>(n);
    u32* bcnt = dalloc<u32>(8ull * ncell);
    u32* special = dalloc<u32>(32ull * ncell);
    OctScanOp oop{em, bat, cell_start, ncell, n, limit, ex_state, ex_m0, rank, bcnt, special, &d_scr->err};
    pb();
    nl = octscan(oop, n, scan_tmp, st);
    pe("oct_scan", nl, n, 12ull * n);
    trace_mark(st, "octsplit", h, n);

    // ---- stage 7: flush decisions + offsets
    BucketDst* bdst = dalloc<BucketDst>(8ull * ncell);
    u32 *cell_deliver = dalloc<u32>(ncell), *cell_children = dalloc<u32>(ncell);
    u32 *cell_dbase = dalloc<u32>(ncell), *cell_cbase = dalloc<u32>(ncell);
    u64 *cell_bytes = dalloc<u64>(ncell), *node_off = dalloc<u64>(ncell);
    pb();
    k_bucket_decide<<<div_up(ncell, 128), 128, 0, st>>>(bcnt, special, win_end, ncell, limit, ex_state, ex_m0, bdst,
                                                        cell_deliver, cell_children, cell_bytes);
    CellScanOp sop{cell_deliver, cell_children, cell_bytes, cell_dbase, cell_cbase, node_off, &d_scr->totals};
    nl = chain_scan<CellScanOp, 256, 4>(sop, ncell, scan_tmp, st);
    pe("bucket_decide", 1 + nl, ncell, 200ull * ncell);
    // node table for the host
    u64* const hoff = (u64*)(tab + tab_idx);
    u64* const hbytes = (u64*)(tab + tab_idx + ((8ull * ncell + 15) & ~15ull));
    if (!existing) publish(hidx, cell_idx, 12ull * ncell);
    publish(hoff, node_off, 8ull * ncell);
    publish(hbytes, cell_bytes, 8ull * ncell);
    fetch_scratch();
    check_err();
    const LevelTotals tot = h_scr->totals;
    const u32 nwin = h_scr->nwin;
    last_mult = (double)n / std::max<u32>(nwin, 1);   // arrivals per prism: predicts the next level's (record filter)
    trace_mark(st, "decide", h, ncell);

    // ---- stage 8: node images + next level's arrivals
    LevelImage img;
    img.hierarchy = h;
    img.bytes = tot.image_bytes;
    img.d_image = dalloc_from<unsigned char>(2, std::max<u64>(img.bytes, 16));
    PcbPoint* next_pts = dalloc<PcbPoint>(tot.n_next);
    u32* next_bat = dalloc<u32>(tot.n_next);
    u32* next_cell_start = dalloc<u32>((u64)tot.ncell_next + 1);
    int* next_cell_idx = dalloc<int>(3ull * tot.ncell_next);
    pb();
    k_cell_emit<<<div_up(ncell, 128), 128, 0, st>>>(cell_idx, ncell, h, g, win_end, bdst, cell_dbase, cell_cbase, node_off,
                                                    cell_bytes, img.d_image, next_cell_start, next_cell_idx, tot.n_next,
                                                    tot.ncell_next);
    pe("cell_emit", 1, ncell, 300ull * ncell);
    pb();
    k_loser_emit<<<div_up(n, 256 * LOSER_ITEMS), 256, 0, st>>>(em, rank, bat, n, cell_start, ncell, pts, bdst, next_pts, next_bat,
                                                 img.d_image);
    pe("loser_emit", 1, n, 12ull * n + 40ull * (n - nwin));
    pb();
    k_winner_emit<<<div_up(nwin, 256 * WINNER_ITEMS), 256, 0, st>>>(win_src, nwin, win_end, ncell, pts, node_off, img.d_image);
    pe("winner_emit", 1, nwin, 36ull * nwin);
    if (nex) {
        pb();
        k_exlist_emit<<<div_up(nex, 256), 256, 0, st>>>(ex_pts, nex, ex_bucket, ex_rank, bdst, next_pts, next_bat,
                                                        img.d_image);
        pe("exlist_emit", 1, nex, 44ull * nex);
    }
    trace_mark(st, "emit", h, n - nwin);

    // node records (host) and image download
    const u32 slot = (u32)levels.size();
    for (u32 c = 0; c < ncell; c++) {
        NodeRec r;
        r.hierarchy = h;
        r.index[0] = hidx[3 * c]; r.index[1] = hidx[3 * c + 1]; r.index[2] = hidx[3 * c + 2];
        r.offset = hoff[c];
        r.size = hbytes[c];
        r.level_slot = slot;
        nodes.push_back(r);
    }
    if (!(flags & PCB_FLAG_DEVICE_ONLY)) {
        img.h_cap = 0;
        img.h_image = slab_take(std::max<u64>(img.bytes, 16));
        if (!img.h_image) img.h_image = (unsigned char*)g_pinned.get(std::max<u64>(img.bytes, 16), &img.h_cap);
        cudaEvent_t ev;
        CUDA_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        CUDA_TRY(cudaEventRecord(ev, st));
        CUDA_TRY(cudaStreamWaitEvent(copy_st, ev, 0));
        CUDA_TRY(cudaMemcpyAsync(img.h_image, img.d_image, img.bytes, cudaMemcpyDeviceToHost, copy_st));
        CUDA_TRY(cudaEventDestroy(ev));
    }
    levels.push_back(img);

    // ---- hand over to the next level
    dfree(geo); dfree(d2oct); dfree(kA); dfree(kB); dfree(vA); dfree(vB); dfree(rt_mem); dfree(scan_tmp);
    dfree(tinfo); dfree(seg_base);
    dfree(win_src); dfree(win_end); dfree(rank); dfree(bcnt); dfree(special); dfree(bdst);
    dfree(cell_deliver); dfree(cell_children); dfree(cell_dbase); dfree(cell_cbase); dfree(cell_bytes); dfree(node_off);
    dfree(ex_state); dfree(ex_m0); dfree(ex_pts); dfree(ex_bucket); dfree(ex_rank);
    if (own_pts) { dfree(pts); dfree(bat); }
    if (own_cells) { dfree(cell_start); dfree(cell_idx); }
    pts = next_pts; bat = next_bat; cell_start = next_cell_start; cell_idx = next_cell_idx;
    n = tot.n_next;
    ncell = tot.ncell_next;
}

// ---- public wrappers: every CUDA failure becomes a pcb::Error ------------------------------------------
template <class F>
static auto guarded(F&& f) -> decltype(f()) {
    try {
        return f();
    } catch (const PcbCudaError& e) {
        char buf[512];
        snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e.code, cudaGetErrorString(e.code), e.file,
                 e.line, e.expr);
        throw Error{PCB_ERR_CUDA, buf};
    }
}

Builder::Builder(int device, uint32_t flags) : impl_(nullptr) {
    guarded([&] { impl_ = new BuilderImpl(device, flags); return 0; });
}
Builder::~Builder() { delete impl_; }
void Builder::set_config(uint32_t limit, uint32_t dim, float max_cell_size) {
    impl_->limit = limit;
    impl_->dim = dim;
    impl_->max_cell_size = max_cell_size;
}
void Builder::reserve(uint64_t n) {
    guarded([&] { impl_->reserve(n); return 0; });
}
void Builder::begin_file() { impl_->begin_file(); }
void Builder::set_stream(void* stream) { impl_->set_stream((cudaStream_t)stream); }
void Builder::add_points(const pcb_point* pts, uint64_t n, uint32_t batch_size, bool device_ptr) {
    guarded([&] { impl_->add(pts, n, batch_size ? batch_size : 10000u, device_ptr, false); return 0; });
}
void Builder::add_points_batched(const pcb_point* d_pts, const uint32_t* d_batch, uint64_t n) {
    guarded([&] { impl_->add_batched(d_pts, d_batch, n); return 0; });
}
void Builder::add_las_records(const uint8_t* records, uint64_t n, uint32_t stride, int32_t colour_off, const double scale[3],
                              const double offset[3], uint32_t batch_size) {
    guarded([&] { impl_->add_las(records, n, stride, colour_off, scale, offset, batch_size ? batch_size : 10000u); return 0; });
}
void* Builder::alloc_host(uint64_t bytes, uint64_t* got) { return guarded([&] { return g_pinned.get(bytes, got); }); }
void Builder::free_host(void* p, uint64_t bytes) { g_pinned.put(p, bytes); }
void Builder::add_arrivals(uint32_t level, const pcb_point* d_pts, const uint32_t* d_bat, uint64_t n,
                           const uint32_t* d_cell_start, const int32_t* d_cell_idx, uint64_t ncell, const float* bbox6,
                           uint64_t count_points) {
    guarded([&] { impl_->add_arrivals(level, d_pts, d_bat, n, d_cell_start, d_cell_idx, ncell, bbox6, count_points); return 0; });
}
void Builder::set_stop_level(uint32_t level) { impl_->stop_level = level; }
void Builder::set_prism_load(double v) { impl_->ext.prism_load = v; }
PendingArrivals Builder::pending() const {
    PendingArrivals p;
    p.level = impl_->pend_level;
    p.n = impl_->pend_n;
    p.ncell = impl_->pend_ncell;
    p.d_pts = (const pcb_point*)impl_->pend_pts;
    p.d_bat = impl_->pend_bat;
    p.d_cell_start = impl_->pend_cell_start;
    p.d_cell_idx = impl_->pend_cell_idx;
    p.prism_load = impl_->last_mult;
    return p;
}
void Builder::add_points_one_batch(const pcb_point* pts, uint64_t n, bool device_ptr) {
    guarded([&] { impl_->add(pts, n, 0, device_ptr, true); return 0; });
}
void Builder::build(Metadata* meta, NodeSource* existing) {
    guarded([&] { impl_->build(meta, existing); return 0; });
}
const std::vector<NodeRec>& Builder::nodes() const { return impl_->nodes; }
const std::vector<LevelImage>& Builder::levels() const { return impl_->levels; }
std::vector<KernelStat> Builder::kernel_stats() const { return impl_->stats; }
uint64_t Builder::points_added() const { return impl_->n_in; }
uint64_t Builder::arrivals() const { return impl_->arrivals; }
uint64_t Builder::launches() const { return impl_->launches; }
uint32_t Builder::levels_built() const { return impl_->levels_built; }
double Builder::h2d_seconds() const { return impl_->h2d_s; }
double Builder::build_seconds() const { return impl_->build_s; }
double Builder::d2h_seconds() const { return impl_->d2h_s; }

// ---- multi-GPU routing (free functions: they run before a handle exists) ---------------------------------
struct NoProf {
    void pb() {}
    void pe(const char*, u64, u64, u64) {}
};

static L0Box make_box(const int32_t origin[3], const uint32_t dims[3]) {
    L0Box box;
    box.ox = origin[0]; box.oy = origin[1]; box.oz = origin[2];
    box.nx = dims[0]; box.ny = dims[1]; box.nz = dims[2];
    return box;
}

int route_bbox(int device, void* stream, const pcb_point* d_pts, uint64_t n, float S0, float out_min[3], float out_max[3]) {
    return guarded([&] {
        CUDA_TRY(cudaSetDevice(device));
        cudaStream_t st = (cudaStream_t)stream;
        if (n >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "more than 2^30-1 points on one device"};
        int init[8] = {INT_MAX, INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MIN, 0, 0};
        struct Scratch8 {   // freed on every path out of this function
            int* p = nullptr;
            cudaStream_t st;
            ~Scratch8() { if (p) cudaFreeAsync(p, st); }
        } scr;
        scr.st = st;
        CUDA_TRY(cudaMallocAsync((void**)&scr.p, sizeof init, st));
        int* d = scr.p;
        CUDA_TRY(cudaMemcpyAsync(d, init, sizeof init, cudaMemcpyHostToDevice, st));
        if (n) k_bbox_check<<<148 * 8, 256, 0, st>>>((const PcbPoint*)d_pts, (u32)n, S0, d, (u32*)(d + 6));
        int res[8];
        CUDA_TRY(cudaMemcpyAsync(res, d, sizeof res, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (res[6] & PCB_DERR_NONFINITE) throw Error{PCB_ERR_RANGE, "non-finite coordinate in input"};
        if (res[6] & PCB_DERR_L0RANGE) throw Error{PCB_ERR_RANGE, "level-0 cell index outside +-2^20"};
        for (int a = 0; a < 3; a++) {
            out_min[a] = n ? ord2f_host(res[a]) : INFINITY;
            out_max[a] = n ? ord2f_host(res[3 + a]) : -INFINITY;
        }
        return n ? 1 : 0;
    });
}

int route_histogram(int device, void* stream, const pcb_point* d_pts, uint64_t n, float S0, const int32_t origin[3],
                    const uint32_t dims[3], uint32_t* d_hist) {
    return guarded([&] {
        CUDA_TRY(cudaSetDevice(device));
        cudaStream_t st = (cudaStream_t)stream;
        const u64 vol = (u64)dims[0] * dims[1] * dims[2];
        if (n >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "more than 2^30-1 points on one device"};
        if (vol == 0 || vol > (1ull << 26)) throw Error{PCB_ERR_RANGE, "level-0 index box too large for the dense routing table"};
        CUDA_TRY(cudaMemsetAsync(d_hist, 0, vol * 4, st));
        if (n) k_route_hist<<<148 * 4, 256, 0, st>>>((const PcbPoint*)d_pts, (u32)n, S0, make_box(origin, dims), (u32)vol, d_hist);
        return n ? 1 : 0;
    });
}

int route_pack(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index, uint32_t batch_size,
               float S0, const int32_t origin[3], const uint32_t dims[3], const uint8_t* d_owner, uint32_t world,
               pcb_point* d_send_pts, uint32_t* d_send_batch, uint64_t* send_counts) {
    return guarded([&] {
        CUDA_TRY(cudaSetDevice(device));
        cudaStream_t st = (cudaStream_t)stream;
        for (u32 r = 0; r < world; r++) send_counts[r] = 0;
        if (!n) return 0;
        if (n >= (1ull << 30) || world == 0 || world > 256 || batch_size == 0) throw Error{PCB_ERR_ARG, "bad routing arguments"};
        const u32 n32 = (u32)n;
        static const bool no_split = pcb_tuning_env("PCB_ROUTE_SORT") != nullptr;   // test hook: force the general path
        if (world <= 8 && !no_split) {
            // one node: single-pass stable split (local histogram -> per-owner send offsets -> chained 8-way split)
            const L0Box box = make_box(origin, dims);
            const u64 vol = (u64)dims[0] * dims[1] * dims[2];
            const u32 ntiles = div_up(n, ROUTE_SPLIT_TILE);
            u32* tmp = nullptr;   // [vol] local histogram | [8] counts | [8] bases | descriptors
            const u64 words = vol + 16 + 4 + (chain_temp_bytes(ntiles) + 3) / 4 + 4;
            CUDA_TRY(cudaMallocAsync((void**)&tmp, 4 * words, st));
            CUDA_TRY(cudaMemsetAsync(tmp, 0, 4 * words, st));
            u32 *lhist = tmp, *counts = tmp + vol, *bases = counts + 8;
            uint4* desc = (uint4*)(tmp + ((vol + 16 + 3) & ~3ull));
            k_route_hist<<<148 * 8, 256, 0, st>>>((const PcbPoint*)d_pts, n32, S0, box, (u32)vol, lhist);
            k_route_owner_counts<<<1, 256, 0, st>>>(lhist, (u32)vol, d_owner, counts, bases);
            RouteDst dst8;
            for (int o = 0; o < 8; o++) { dst8.pts[o] = (PcbPoint*)d_send_pts; dst8.bat[o] = d_send_batch; dst8.off[o] = 0; }
            k_route_split<<<ntiles, ROUTE_SPLIT_THREADS, 0, st>>>((const PcbPoint*)d_pts, n32, S0, box, d_owner, bases, first_index,
                                                                 batch_size, dst8, world, desc);
            u32 hc[8];
            CUDA_TRY(cudaMemcpyAsync(hc, counts, sizeof hc, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            for (u32 r = 0; r < world; r++) send_counts[r] = hc[r];
            CUDA_TRY(cudaFreeAsync(tmp, st));
            return 3;
        }
        const int pb = std::max(1, bits_count(n));
        const RadixPlan plan = radix_plan(pb, std::max(1, bits_count(world)));
        u64 *kA = nullptr, *kB = nullptr;
        u32* ghist = nullptr;
        const u64 rwords = RadixTemp::ghist_words() + RadixTemp::status_words(n, plan);
        CUDA_TRY(cudaMallocAsync((void**)&kA, 8 * n, st));
        CUDA_TRY(cudaMallocAsync((void**)&kB, 8 * n, st));
        CUDA_TRY(cudaMallocAsync((void**)&ghist, 4 * rwords, st));
        CUDA_TRY(cudaMemsetAsync(ghist, 0, 4 * rwords, st));
        RadixTemp rt;
        rt.ghist = ghist;
        rt.status = ghist + RadixTemp::ghist_words();
        k_route_key<<<148 * 8, 256, 0, st>>>((const PcbPoint*)d_pts, n32, S0, make_box(origin, dims), d_owner, pb, plan, kA, ghist);
        NoProf np;
        int cur = radix_sort<false>(kA, nullptr, kB, nullptr, n32, plan, rt, st, np);
        k_route_gather<<<div_up(n, 256), 256, 0, st>>>((const PcbPoint*)d_pts, cur ? kB : kA, n32, pb, first_index, batch_size,
                                                       (PcbPoint*)d_send_pts, d_send_batch);
        std::vector<u32> h(RADIX_MAX_BINS);
        CUDA_TRY(cudaMemcpyAsync(h.data(), ghist, 4 * RADIX_MAX_BINS, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        for (u32 r = 0; r < world; r++) send_counts[r] = h[r];
        CUDA_TRY(cudaFreeAsync(kA, st));
        CUDA_TRY(cudaFreeAsync(kB, st));
        CUDA_TRY(cudaFreeAsync(ghist, st));
        return 2 + plan.npass;
    });
}

// Partition + exchange in one kernel: every point is stored straight into the receive buffer of the rank that owns
// its level-0 cell.  recv_*_ptrs[o] are device pointers valid on THIS device (peer-mapped memory of rank o, or local
// memory for o == own rank); recv_offsets[o] = points that lower ranks send to o.  The caller synchronises the ranks
// before (buffers free) and after (stores landed).
int route_scatter(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index, uint32_t batch_size,
                  float S0, const int32_t origin[3], const uint32_t dims[3], const uint8_t* d_owner, uint32_t world,
                  const uint64_t* recv_pts_ptrs, const uint64_t* recv_batch_ptrs, const uint64_t* recv_offsets) {
    return guarded([&] {
        CUDA_TRY(cudaSetDevice(device));
        cudaStream_t st = (cudaStream_t)stream;
        if (!n) return 0;
        if (n >= (1ull << 30) || world == 0 || world > 8 || batch_size == 0) throw Error{PCB_ERR_ARG, "bad routing arguments"};
        RouteDst dst8;
        for (u32 o = 0; o < 8; o++) {
            const u32 q = o < world ? o : 0;
            if (recv_offsets[q] >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "receive offset too large"};
            dst8.pts[o] = (PcbPoint*)(uintptr_t)recv_pts_ptrs[q];
            dst8.bat[o] = (u32*)(uintptr_t)recv_batch_ptrs[q];
            dst8.off[o] = (u32)recv_offsets[q];
        }
        const u32 ntiles = div_up(n, ROUTE_SPLIT_TILE);
        uint4* desc = nullptr;
        CUDA_TRY(cudaMallocAsync((void**)&desc, chain_temp_bytes(ntiles), st));
        CUDA_TRY(cudaMemsetAsync(desc, 0, chain_temp_bytes(ntiles), st));
        k_route_split<<<ntiles, ROUTE_SPLIT_THREADS, 0, st>>>((const PcbPoint*)d_pts, (u32)n, S0, make_box(origin, dims), d_owner,
                                                             nullptr, first_index, batch_size, dst8, world, desc);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaFreeAsync(desc, st));
        return 1;
    });
}

// Level-0 exchange fused with the level-0 split (world <= 8, level-0 box of at most 512 cells): ONE pass reads every
// point of the slice and stores it, with its global batch ordinal + 1, at its final level-0 arrival slot in the
// receive buffer of the rank that owns its cell: d_dst_base[cell] = (first slot of the cell in the owner's buffer) +
// (points of lower ranks in the cell).  The owner then builds straight from pcb_add_arrivals(level 0).
int route_scatter_cells(int device, void* stream, const pcb_point* d_pts, uint64_t n, uint64_t first_index, uint32_t batch_size,
                        float S0, const int32_t origin[3], const uint32_t dims[3], const uint8_t* d_owner,
                        const uint32_t* d_dst_base, uint32_t world, const uint64_t* recv_pts_ptrs,
                        const uint64_t* recv_batch_ptrs, uint32_t* d_err) {
    return guarded([&] {
        CUDA_TRY(cudaSetDevice(device));
        cudaStream_t st = (cudaStream_t)stream;
        if (!n) return 0;
        const u64 vol = (u64)dims[0] * dims[1] * dims[2];
        if (n >= (1ull << 30) || world == 0 || world > 8 || batch_size == 0 || vol == 0 || vol > L0S_MAX_BINS)
            throw Error{PCB_ERR_ARG, "bad routing arguments (fused cell scatter: world <= 8, level-0 box <= 512 cells)"};
        RouteDst dst8;
        for (u32 o = 0; o < 8; o++) {
            const u32 q = o < world ? o : 0;
            dst8.pts[o] = (PcbPoint*)(uintptr_t)recv_pts_ptrs[q];
            dst8.bat[o] = (u32*)(uintptr_t)recv_batch_ptrs[q];
            dst8.off[o] = 0;
        }
        const u32 nb = (u32)vol, ntiles = div_up(n, L0S_TILE);
        const int dbits = std::max(1, bits_count(vol));
        struct Tmp {   // look-back status words; freed on every path
            u32* p = nullptr;
            cudaStream_t st;
            ~Tmp() { if (p) cudaFreeAsync(p, st); }
        } tmp;
        tmp.st = st;
        const u64 words = (u64)ntiles * nb + 4;
        CUDA_TRY(cudaMallocAsync((void**)&tmp.p, 4 * words, st));
        CUDA_TRY(cudaMemsetAsync(tmp.p, 0, 4 * words, st));
        const int smem = L0S_TILE * 22 + (L0S_WARPS + 2) * (int)nb * 4;
        static bool attr_set = false;
        if (!attr_set) {
            CUDA_TRY(cudaFuncSetAttribute(k_l0_split<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, L0SplitSmem::bytes));
            CUDA_TRY(cudaFuncSetAttribute(k_l0_split<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
            attr_set = true;
        }
        k_l0_split<true><<<ntiles, L0S_THREADS, smem, st>>>((const PcbPoint*)d_pts, (u32)n, S0, make_box(origin, dims), nb, dbits,
                                                            d_dst_base, nullptr, 0, nullptr, nullptr, nullptr, tmp.p,
                                                            d_err, d_owner, dst8, first_index, batch_size);
        CUDA_TRY(cudaGetLastError());
        return 1;
    });
}

// Level-l re-balance: copy the selected cells of one level's arrivals back to back (the all-to-all send buffer).
int gather_cells(int device, void* stream, const pcb_point* d_pts, const uint32_t* d_bat, const uint32_t* d_cell_start,
                 const uint32_t* d_sel, const uint32_t* d_out_start, uint32_t nsel, uint64_t total, pcb_point* d_out_pts,
                 uint32_t* d_out_bat) {
    return guarded([&] {
        CUDA_TRY(cudaSetDevice(device));
        if (!nsel || !total) return 0;
        if (total >= (1ull << 30)) throw Error{PCB_ERR_RANGE, "more than 2^30-1 arrivals on one device"};
        const u32 grid = (u32)std::min<u64>(148 * 8, (total + 255) / 256);
        k_gather_cells<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint4*)d_pts, d_bat, d_cell_start, d_sel, d_out_start, nsel,
                                                               (uint4*)d_out_pts, d_out_bat);
        CUDA_TRY(cudaGetLastError());
        return 1;
    });
}

// Deterministic greedy bin packing of cells onto ranks: largest cell first (ties: lower cell ordinal), onto the least
// loaded rank (ties: lower rank).  A pure function of the global histogram, so every rank computes the same plan.
void plan_owners(const uint64_t* counts, uint64_t ncells, uint32_t world, uint8_t* owner) {
    std::vector<uint64_t> order;
    order.reserve(ncells);
    for (uint64_t i = 0; i < ncells; i++)
        if (counts[i]) order.push_back(i);
    std::stable_sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) { return counts[a] > counts[b]; });
    std::vector<uint64_t> load(world, 0);
    for (uint64_t i = 0; i < ncells; i++) owner[i] = 0;
    for (uint64_t i : order) {
        uint32_t r = 0;
        for (uint32_t q = 1; q < world; q++)
            if (load[q] < load[r]) r = q;
        owner[i] = (uint8_t)r;
        load[r] += counts[i];
    }
}

// ---- allocation-free bounding box (no counterpart in the reference) ----
// Same result as route_bbox, for callers that must not allocate (CUDA graph capture, tight serving loops): the
// min / max / error scratch of k_bbox_check lives in a file-scope device array instead of a cudaMallocAsync block.
// The array's name is a host-side stand-in in host code: its device address is taken ONCE per device, with that device
// current, through cudaGetSymbolAddress.  Every call on a device shares the array: it is re-initialised by the first
// kernel of each call, and calls that overlap on the GPU (two streams) race.
#define PCB_STATIC_SCRATCH_BYTES 256   // held by every GPU that loads the library until exit
static __device__ __align__(256) int g_static_scratch[PCB_STATIC_SCRATCH_BYTES / 4];
__global__ void k_bbox_static_init(int* __restrict__ s) {
    const int t = threadIdx.x;
    if (t < 8) s[t] = t < 3 ? INT_MAX : t < 6 ? INT_MIN : 0;
}
__global__ void k_bbox_static_finish(const int* __restrict__ s, u32 n, float* __restrict__ out6, u32* __restrict__ err) {
    const int t = threadIdx.x;
    if (t < 6) {
        int i = s[t];
        i = i ^ ((i >> 31) & 0x7FFFFFFF);   // inverse of f2ord
        out6[t] = n ? __int_as_float(i) : (t < 3 ? INFINITY : -INFINITY);
    }
    if (t == 6 && err) *err = (u32)s[6];
}
static std::mutex g_static_mu;
static int* g_static_ptr[64] = {nullptr};
int route_bbox_static(int device, void* stream, const pcb_point* d_pts, uint64_t n, float S0, float* d_out6, uint32_t* d_err) {
    return guarded([&] {
        if (device < 0 || device >= 64) throw Error{PCB_ERR_ARG, "bad device ordinal"};
        if (n > PCB_BBOX_STATIC_MAX_POINTS)
            throw Error{PCB_ERR_RANGE, "pcb_route_bbox_static: more than PCB_BBOX_STATIC_MAX_POINTS (2^30 - 1) points in one call"};
        CUDA_TRY(cudaSetDevice(device));
        int* scratch = nullptr;
        {
            std::lock_guard<std::mutex> g(g_static_mu);
            if (!g_static_ptr[device]) {   // once per device, with that device current
                void* q = nullptr;
                CUDA_TRY(cudaGetSymbolAddress(&q, g_static_scratch));
                g_static_ptr[device] = (int*)q;
            }
            scratch = g_static_ptr[device];
        }
        cudaStream_t st = (cudaStream_t)stream;
        k_bbox_static_init<<<1, 32, 0, st>>>(scratch);
        if (n) k_bbox_check<<<148 * 8, 256, 0, st>>>((const PcbPoint*)d_pts, (u32)n, S0, scratch, (u32*)(scratch + 6));
        k_bbox_static_finish<<<1, 32, 0, st>>>(scratch, (u32)n, d_out6, d_err);
        CUDA_TRY(cudaGetLastError());
        return n ? 3 : 2;
    });
}

void trim_host_cache() {
    g_pinned.trim();
    g_arenas.trim();
    BuilderImpl::trim_pools();
}

int device_count() {
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) return 0;
    return c;
}

}  // namespace pcb

#ifdef PCB_RADIX_INSTR
// diagnostics build only: cycles per onesweep phase summed over tiles (see RADIX_TICK), reset on read
extern "C" int pcb_debug_radix_instr(unsigned long long* out) {
    unsigned long long zero[16] = {0};
    if (cudaMemcpyFromSymbol(out, g_radix_instr, sizeof(zero)) != cudaSuccess) return -1;
    return cudaMemcpyToSymbol(g_radix_instr, zero, sizeof(zero)) == cudaSuccess ? 0 : -1;
}
#endif

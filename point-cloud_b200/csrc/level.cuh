// level.cuh — the per-level kernels of the bulk, level-synchronous octree/LOD build.
//
// One level = what the reference does for every batch in Converter::add_points_in_hierarchy
// (point-converter/src/converter.rs:114-139) and add_points_to_cell (:62-70), done once for ALL
// batches.  Arrivals of a level are held sorted by (cell, arrival order tau); the arrival position
// is therefore a proxy of tau inside a cell and only the batch ordinal is carried explicitly
// (`bat` = batch + 1; 0 = state that existed before this run, i.e. merge mode).
// tests/bulk_model.py is the executable specification of this file (validated against the oracle).
#pragma once
#include "common.cuh"

struct LevelGeom {
    float S;        // cell edge of this level            metadata.rs:91-93
    float S_child;  // cell edge of level + 1
    float sub;      // S / sub_grid_dimension              metadata.rs:95-97
    float R;        // sub / 2.0 (hex circumradius)        cell.rs:276-278
    int bx, by, bz; // bits of the per-cell hex window per axis
    int mx, my, mz; // margins subtracted from the hex key of the cell's min corner
    int hexbits;    // bx + by + bz
};

// order-preserving float <-> int (for atomicMin/atomicMax of f32)
__device__ __forceinline__ int f2ord(float f) {
    int i = __float_as_int(f);
    return i ^ ((i >> 31) & 0x7FFFFFFF);
}
static inline float ord2f_host(int i) {
    i = i ^ ((i >> 31) & 0x7FFFFFFF);
    float f;
    memcpy(&f, &i, 4);
    return f;
}

// ---------------------------------------------------------------------------------------------
// Stage 0a: bounding box (converter.rs:96-104; bounding-volume/src/lib.rs:23-52) + input validation.
// bbox[0..2] = min xyz, bbox[3..5] = max xyz as ordered ints.
__global__ void k_bbox_check(const PcbPoint* __restrict__ pts, u32 n, float S0, int* __restrict__ bbox,
                             u32* __restrict__ err) {
    int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
    u32 bad = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        PcbPoint p = pts[i];
        float v[3] = {p.x, p.y, p.z};
#pragma unroll
        for (int a = 0; a < 3; a++) {
            if (!isfinite(v[a])) bad |= PCB_DERR_NONFINITE;
            int o = f2ord(v[a]);
            mn[a] = min(mn[a], o);
            mx[a] = max(mx[a], o);
            int ci = cell_index_axis(v[a], S0);
            if (ci < -(1 << 20) || ci >= (1 << 20)) bad |= PCB_DERR_L0RANGE;
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        for (int o = 16; o; o >>= 1) {
            mn[a] = min(mn[a], __shfl_xor_sync(0xFFFFFFFFu, mn[a], o));
            mx[a] = max(mx[a], __shfl_xor_sync(0xFFFFFFFFu, mx[a], o));
        }
    }
    for (int o = 16; o; o >>= 1) bad |= __shfl_xor_sync(0xFFFFFFFFu, bad, o);
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            atomicMin(&bbox[a], mn[a]);
            atomicMax(&bbox[3 + a], mx[a]);
        }
        if (bad) atomicOr(err, bad);
    }
}

// Stage 0b: level-0 cell of every input point (converter.rs:110 group_points(points, 0)) as a dense
// linear id inside the index box [o, o + dim) known from the bounding box; value = input index.
struct L0Box {
    int ox, oy, oz;
    u32 nx, ny, nz;
};
__device__ __forceinline__ u64 l0_key(const PcbPoint& p, float S0, const L0Box& b) {
    u64 ix = (u64)(u32)(cell_index_axis(p.x, S0) - b.ox);
    u64 iy = (u64)(u32)(cell_index_axis(p.y, S0) - b.oy);
    u64 iz = (u64)(u32)(cell_index_axis(p.z, S0) - b.oz);
    return (iz * b.ny + iy) * b.nx + ix;
}
__global__ void k_l0_key(const PcbPoint* __restrict__ pts, u32 n, float S0, L0Box box, u64* __restrict__ keys,
                         u64* __restrict__ vals, u64* __restrict__ diffmask) {
    const u64 key0 = l0_key(pts[0], S0, box);
    u64 diff = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 k = l0_key(pts[i], S0, box);
        keys[i] = k;
        vals[i] = i;
        diff |= k ^ key0;
    }
    for (int o = 16; o; o >>= 1) diff |= __shfl_xor_sync(0xFFFFFFFFu, diff, o);
    if ((threadIdx.x & 31) == 0 && diff) atomicOr((unsigned long long*)diffmask, (unsigned long long)diff);
}

// Stage 0c: distinct level-0 cells of the sorted keys -> cell list + segment starts.
struct L0HeadOp {
    typedef u32 State;
    const u64* keys;
    u32 n;
    L0Box box;
    u32* cell_start;  // [ncell + 1]
    int* cell_idx;    // [ncell * 3]
    u32* ncell_out;
    __device__ State identity() const { return 0; }
    __device__ State load(u32 i) const { return (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u; }
    __device__ State combine(const State& a, const State& b) const { return a + b; }
    __device__ void consume(u32 i, const State& excl, const State& incl) const {
        if (incl != excl) {
            cell_start[excl] = i;
            u64 k = keys[i];
            u32 ix = (u32)(k % box.nx);
            u64 r = k / box.nx;
            u32 iy = (u32)(r % box.ny);
            u32 iz = (u32)(r / box.ny);
            cell_idx[3 * excl + 0] = box.ox + (int)ix;
            cell_idx[3 * excl + 1] = box.oy + (int)iy;
            cell_idx[3 * excl + 2] = box.oz + (int)iz;
        }
    }
    __device__ void finish(const State& total) const {
        *ncell_out = total;
        cell_start[total] = n;
    }
};

// Stage 0d: arrivals of level 0 in (cell, input index) order; batch ordinal from the segment table
// (lib.rs:31-52: batches of `batch_size` points, never spanning files).
struct BatchSeg {
    u32 start;       // first input index of the segment
    u32 first_batch; // global ordinal of its first batch
    u32 batch_size;  // points per batch inside the segment
    u32 pad;
};
__global__ void k_l0_gather(const PcbPoint* __restrict__ in_pts, const u64* __restrict__ vals, u32 n,
                            const BatchSeg* __restrict__ segs, u32 nseg, PcbPoint* __restrict__ pts,
                            u32* __restrict__ bat) {
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    u32 i = (u32)vals[j];
    pts[j] = in_pts[i];
    u32 lo = 0, hi = nseg;
    while (hi - lo > 1) {
        u32 mid = (lo + hi) >> 1;
        if (segs[mid].start <= i) lo = mid; else hi = mid;
    }
    BatchSeg s = segs[lo];
    bat[j] = s.first_batch + (i - s.start) / s.batch_size + 1u;
}

// ---------------------------------------------------------------------------------------------
// Stage 1: per-cell origin of the hex window (hex key of the cell's min corner minus a margin).
__global__ void k_cellbase(const int* __restrict__ cell_idx, u32 ncell, LevelGeom g, int* __restrict__ cellbase) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    float x = __fmul_rn((float)cell_idx[3 * c + 0], g.S);
    float y = __fmul_rn((float)cell_idx[3 * c + 1], g.S);
    float z = __fmul_rn((float)cell_idx[3 * c + 2], g.S);
    HexKey k = hex_from_world(x, y, z, g.R);
    cellbase[3 * c + 0] = k.x - g.mx;
    cellbase[3 * c + 1] = k.y - g.my;
    cellbase[3 * c + 2] = k.z - g.mz;
}

// Stage 2: hex-prism key (cell.rs:276-278 -> hex.rs:32-34) and contest metric d2 (cell.rs:75-78)
// of every arrival.  key = cell ordinal << hexbits | Z << (bx+by) | Y << bx | X   (window-relative)
//                    val = d2 bits << 32 | arrival position
__device__ __forceinline__ u64 hexkey_of(const PcbPoint& p, u32 c, const int* __restrict__ cellbase,
                                         const LevelGeom& g, float* d2, u32* bad) {
    HexKey k = hex_from_world(p.x, p.y, p.z, g.R);
    *d2 = hex_d2(k, p.x, p.y, p.z, g.R);
    u32 rx = (u32)(k.x - cellbase[3 * c + 0]);
    u32 ry = (u32)(k.y - cellbase[3 * c + 1]);
    u32 rz = (u32)(k.z - cellbase[3 * c + 2]);
    if ((rx >> g.bx) | (ry >> g.by) | (rz >> g.bz)) *bad |= PCB_DERR_HEXRANGE;
    return ((u64)c << g.hexbits) | ((u64)rz << (g.bx + g.by)) | ((u64)ry << g.bx) | (u64)rx;
}
__global__ void k_hexkey(const PcbPoint* __restrict__ pts, u32 n, const u32* __restrict__ cell_start, u32 ncell,
                         const int* __restrict__ cellbase, LevelGeom g, u64* __restrict__ keys,
                         u64* __restrict__ vals, u32* __restrict__ cell_of, u64* __restrict__ diffmask,
                         u32* __restrict__ err) {
    u32 bad = 0;
    float d2;
    const u64 key0 = hexkey_of(pts[0], 0, cellbase, g, &d2, &bad);
    bad = 0;
    u64 diff = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        PcbPoint p = pts[i];
        u32 c = find_segment(cell_start, ncell, (u32)i);
        u64 k = hexkey_of(p, c, cellbase, g, &d2, &bad);
        keys[i] = k;
        vals[i] = ((u64)__float_as_uint(d2) << 32) | (u64)i;
        cell_of[i] = c;
        diff |= k ^ key0;
    }
    for (int o = 16; o; o >>= 1) {
        diff |= __shfl_xor_sync(0xFFFFFFFFu, diff, o);
        bad |= __shfl_xor_sync(0xFFFFFFFFu, bad, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (diff) atomicOr((unsigned long long*)diffmask, (unsigned long long)diff);
        if (bad) atomicOr(err, bad);
    }
}

// Stage 4: the hex-grid contest of Cell::add_point (cell.rs:70-94) for every (cell, prism) group at
// once.  In sorted order a group is contiguous and in arrival order; the incumbent seen by an arrival
// is the running minimum of (d2, position) over the earlier members (strict `<`: equal d2 keeps the
// earlier one).  Every non-first member emits exactly one loser — itself, or the incumbent it displaces
// — and the emission is recorded under the position of the arrival that triggered it.
struct ContestState {
    u64 mn;       // min (d2 bits << 32 | pos) since the last group head
    u32 has_head; // a group head lies inside the scanned range
    u32 ntail;    // groups completed (== winners emitted)
};
struct ContestOp {
    typedef ContestState State;
    const u64* keys;
    const u64* vals;
    u32 n;
    int hexbits;
    u32* emitted;  // [n] by arrival position: loser's arrival position, PCB_NONE for group heads
    u32* win_src;  // [#groups] arrival position of each winner, sorted (cell, Z, Y, X)
    u32* win_end;  // [ncell] end of the cell's winners in win_src
    u32* nwin_out;
    __device__ State identity() const { return State{~0ull, 0u, 0u}; }
    __device__ State load(u32 i) const {
        u64 k = keys[i];
        u32 head = (i == 0 || keys[i - 1] != k) ? 1u : 0u;
        u32 tail = (i + 1 == n || keys[i + 1] != k) ? 1u : 0u;
        return State{vals[i], head, tail};
    }
    __device__ State combine(const State& a, const State& b) const {
        State r;
        r.mn = b.has_head ? b.mn : (a.mn < b.mn ? a.mn : b.mn);
        r.has_head = a.has_head | b.has_head;
        r.ntail = a.ntail + b.ntail;
        return r;
    }
    __device__ void consume(u32 i, const State& excl, const State& incl) const {
        u64 k = keys[i], v = vals[i];
        u32 pos = (u32)v;
        bool head = (i == 0 || keys[i - 1] != k);
        if (head) {
            emitted[pos] = PCB_NONE;
        } else {
            u64 inc = excl.mn;  // incumbent
            emitted[pos] = ((u32)(v >> 32) < (u32)(inc >> 32)) ? (u32)inc : pos;  // cell.rs:80 strict <
        }
        bool tail = (i + 1 == n || keys[i + 1] != k);
        if (tail) {
            win_src[excl.ntail] = (u32)incl.mn;
            if (i + 1 == n || (keys[i + 1] >> hexbits) != (k >> hexbits)) win_end[(u32)(k >> hexbits)] = excl.ntail + 1;
        }
    }
    __device__ void finish(const State& total) const { *nwin_out = total.ntail; }
};

// Stage 5: child octant of every emitted loser (converter.rs:68 group_points(losers, h + 1)), stored
// under the trigger's arrival position; 8 = this arrival emitted nothing.  octant = dx*4 + dy*2 + dz
// so that octant order == lexicographic (x, y, z) order of the child index.
__global__ void k_octant(const u32* __restrict__ emitted, const PcbPoint* __restrict__ pts,
                         const u32* __restrict__ cell_of, const int* __restrict__ cell_idx, u32 n, float S_child,
                         unsigned char* __restrict__ oct8, u32* __restrict__ err) {
    u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u32 e = emitted[t];
    unsigned char o = 8;
    if (e != PCB_NONE) {
        PcbPoint p = pts[e];
        u32 c = cell_of[t];
        int dx = cell_index_axis(p.x, S_child) - 2 * cell_idx[3 * c + 0];
        int dy = cell_index_axis(p.y, S_child) - 2 * cell_idx[3 * c + 1];
        int dz = cell_index_axis(p.z, S_child) - 2 * cell_idx[3 * c + 2];
        if (((u32)dx | (u32)dy | (u32)dz) > 1u) {
            atomicOr(err, PCB_DERR_NEST);
            dx &= 1; dy &= 1; dz &= 1;
        }
        o = (unsigned char)(dx * 4 + dy * 2 + dz);
    }
    oct8[t] = o;
}

// Stage 6: segmented stable 8-way split of the losers by child octant.  One scan over arrival positions
// with eight running counters gives every loser its rank among the losers of its octant; the counter
// vector at each cell's first arrival (cellP) turns global ranks into per-(cell, octant) ranks.
struct Oct8 {
    u32 c[8];
};
struct OctScanOp {
    typedef Oct8 State;
    const unsigned char* oct8;
    const u32* cell_of;
    u32 ncell;
    u32* rank;   // [n] exclusive count of the same octant before this position (global)
    u32* cellP;  // [(ncell + 1) * 8] counters at the first arrival of each cell; [ncell] = totals
    __device__ State identity() const {
        State s;
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = 0;
        return s;
    }
    __device__ State load(u32 i) const {
        State s = identity();
        u32 o = oct8[i];
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = (o == (u32)k) ? 1u : 0u;
        return s;
    }
    __device__ State combine(const State& a, const State& b) const {
        State s;
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = a.c[k] + b.c[k];
        return s;
    }
    __device__ void consume(u32 i, const State& excl, const State&) const {
        u32 o = oct8[i];
        if (o < 8) {
            u32 r = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) r = (o == (u32)k) ? excl.c[k] : r;
            rank[i] = r;
        }
        u32 c = cell_of[i];
        if (i == 0 || cell_of[i - 1] != c) {
#pragma unroll
            for (int k = 0; k < 8; k++) cellP[8 * c + k] = excl.c[k];
        }
    }
    __device__ void finish(const State& total) const {
#pragma unroll
        for (int k = 0; k < 8; k++) cellP[8 * ncell + k] = total.c[k];
    }
};

// bucket (cell, octant) -> first loser slot; losers end up grouped by cell, then octant, each bucket in
// emission order.
__global__ void k_bucket_start(const u32* __restrict__ cellP, u32 ncell, u32* __restrict__ bstart) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    u32 base = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) base += cellP[8 * c + k];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        bstart[8 * c + k] = base;
        base += cellP[8 * (c + 1) + k] - cellP[8 * c + k];
    }
    if (c == ncell - 1) bstart[8 * ncell] = base;
}

__global__ void k_loser_scatter(const unsigned char* __restrict__ oct8, const u32* __restrict__ rank,
                                const u32* __restrict__ cell_of, const u32* __restrict__ cellP,
                                const u32* __restrict__ bstart, const u32* __restrict__ emitted,
                                const u32* __restrict__ bat, u32 n, u32* __restrict__ los_src,
                                u32* __restrict__ los_bat, u32* __restrict__ los_bucket) {
    u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u32 o = oct8[t];
    if (o >= 8) return;
    u32 c = cell_of[t];
    u32 k = 8 * c + o;
    u32 dst = bstart[k] + rank[t] - cellP[k];
    los_src[dst] = emitted[t];
    los_bat[dst] = bat[t];  // batch of the TRIGGER: the call in which the loser left the grid
    los_bucket[dst] = k;
}

// ---------------------------------------------------------------------------------------------
// Stage 7: Cell::add_points_in_overflow (cell.rs:108-153) for every (cell, child) bucket, closed form.
// `lb` = trigger batches (+1) of the bucket's NEW losers in emission order, `m0`/`state` = what the node
// file held before this run (state 0 absent, 1 Some(list of m0), 2 None).  Returns true when the list is
// delivered to the child; *bf = batch (+1) of the call that delivers the buffered list (0: every call
// delivers its own losers, state None).          tests/bulk_model.py::flush_rule
#define PCB_EX_ABSENT 0
#define PCB_EX_SOME 1
#define PCB_EX_NONE 2
__device__ __forceinline__ bool flush_rule(const u32* __restrict__ lb, u32 n, u32 m0, int state, u32 L, u32* bf) {
    *bf = 0;
    if (state == PCB_EX_NONE) return true;                     // cell.rs:128-130
    if (state == PCB_EX_SOME) {                                // cell.rs:131-147: flush when len >= L
        if ((u64)m0 + n < L) return false;
        u32 k = (L >= 1 + m0) ? L - 1 - m0 : 0;
        *bf = lb[k];
        return true;
    }
    if (L == 0) { *bf = lb[0]; return true; }                  // cell.rs:118: len <= 0 never holds
    if (n < L) return false;                                   // first call keeps when len <= L
    u32 bstar = lb[L - 1];
    if (bstar == lb[0] && (n == L || lb[L] != bstar)) {        // first call brought exactly L: kept (:118),
        if (n == L) return false;                              // flushed by the next call (:137)
        *bf = lb[L];
        return true;
    }
    *bf = bstar;
    return true;
}

#define PCB_BF_ENTRY 1
#define PCB_BF_FLUSH 2

__global__ void k_bucket_decide(const u32* __restrict__ bstart, const u32* __restrict__ los_bat,
                                const u32* __restrict__ win_end, u32 ncell, u32 L,
                                const unsigned char* __restrict__ ex_state, const u32* __restrict__ ex_m0,
                                unsigned char* __restrict__ bflag, u32* __restrict__ bfl, u32* __restrict__ bcount,
                                u32* __restrict__ cell_deliver, u32* __restrict__ cell_children,
                                u64* __restrict__ cell_bytes) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    u32 ng = win_end[c] - (c ? win_end[c - 1] : 0u);
    u64 bytes = 48ull + 16ull * ng + 1ull;
    u32 deliver = 0, children = 0;
    for (int o = 0; o < 8; o++) {
        u32 k = 8 * c + o;
        u32 b0 = bstart[k], n = bstart[k + 1] - b0;
        int st = ex_state ? ex_state[k] : PCB_EX_ABSENT;
        u32 m0 = (st == PCB_EX_SOME) ? ex_m0[k] : 0u;
        unsigned char fl = 0;
        u32 bf = 0, cnt = m0 + n;
        if (n == 0) {
            if (st != PCB_EX_ABSENT) { fl = PCB_BF_ENTRY; bytes += 16ull + 16ull * m0; }
        } else {
            fl = PCB_BF_ENTRY;
            if (flush_rule(los_bat + b0, n, m0, st, L, &bf)) {
                fl |= PCB_BF_FLUSH;
                deliver += cnt;
                children += 1;
                bytes += 16ull;
            } else {
                bytes += 16ull + 16ull * cnt;
            }
        }
        bflag[k] = fl;
        bfl[k] = bf;
        bcount[k] = cnt;
    }
    cell_deliver[c] = deliver;
    cell_children[c] = children;
    cell_bytes[c] = bytes;
}

struct CellScanState {
    u64 bytes;
    u32 deliver, children;
};
struct LevelTotals {
    u64 image_bytes;
    u32 n_next, ncell_next, n_winners, pad;
};
struct CellScanOp {
    typedef CellScanState State;
    const u32* cell_deliver;
    const u32* cell_children;
    const u64* cell_bytes;
    u32* cell_dbase;
    u32* cell_cbase;
    u64* node_off;
    LevelTotals* totals;
    __device__ State identity() const { return State{0ull, 0u, 0u}; }
    __device__ State load(u32 i) const {
        return State{(cell_bytes[i] + 15ull) & ~15ull, cell_deliver[i], cell_children[i]};
    }
    __device__ State combine(const State& a, const State& b) const {
        return State{a.bytes + b.bytes, a.deliver + b.deliver, a.children + b.children};
    }
    __device__ void consume(u32 i, const State& excl, const State&) const {
        node_off[i] = excl.bytes;
        cell_dbase[i] = excl.deliver;
        cell_cbase[i] = excl.children;
    }
    __device__ void finish(const State& t) const {
        totals->image_bytes = t.bytes;
        totals->n_next = t.deliver;
        totals->ncell_next = t.children;
    }
};

__device__ __forceinline__ void store_bytes(unsigned char* dst, const void* src, int nbytes) {
    const unsigned char* s = (const unsigned char*)src;
    for (int i = 0; i < nbytes; i++) dst[i] = s[i];
}
// 16-byte record to an address that is 1 (mod 16): every overflow record of a node image sits there
// (48-byte header + 16*n grid + 1 count byte + 16-byte entry headers, cell.rs:155-181).
__device__ __forceinline__ void store16_at1(unsigned char* dst, uint4 v) {
    u64 lo = ((u64)v.y << 32) | v.x, hi = ((u64)v.w << 32) | v.z;
    dst[0] = (unsigned char)lo;
    *(unsigned short*)(dst + 1) = (unsigned short)(lo >> 8);
    *(u32*)(dst + 3) = (u32)(lo >> 24);
    *(u64*)(dst + 7) = (lo >> 56) | (hi << 8);
    dst[15] = (unsigned char)(hi >> 56);
}

// Stage 8a: per node: 48-byte Header (cell.rs:280-298), overflow entry count and entry headers
// (cell.rs:162-177), plus the bucket -> destination tables used by the point movers below.
__global__ void k_cell_emit(const int* __restrict__ cell_idx, u32 ncell, u32 hierarchy, LevelGeom g,
                            const u32* __restrict__ win_end, const unsigned char* __restrict__ bflag,
                            const u32* __restrict__ bcount, const u32* __restrict__ cell_dbase,
                            const u32* __restrict__ cell_cbase, const u64* __restrict__ node_off,
                            const u64* __restrict__ cell_bytes, unsigned char* __restrict__ image,
                            u32* __restrict__ bdst, u64* __restrict__ boff, u32* __restrict__ next_cell_start,
                            int* __restrict__ next_cell_idx, u32 n_next, u32 ncell_next) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    const int ix = cell_idx[3 * c], iy = cell_idx[3 * c + 1], iz = cell_idx[3 * c + 2];
    u32 ng = win_end[c] - (c ? win_end[c - 1] : 0u);
    u32 nov = 0, nent = 0;
    for (int o = 0; o < 8; o++) {
        unsigned char fl = bflag[8 * c + o];
        if (fl & PCB_BF_ENTRY) {
            nent++;
            if (!(fl & PCB_BF_FLUSH)) nov += bcount[8 * c + o];
        }
    }
    unsigned char* node = image + node_off[c];
    float half = __fdiv_rn(g.S, 2.0f);  // metadata.rs:104-106: idx as f32 * cell_size + cell_size / 2.0
    float px = __fadd_rn(__fmul_rn((float)ix, g.S), half);
    float py = __fadd_rn(__fmul_rn((float)iy, g.S), half);
    float pz = __fadd_rn(__fmul_rn((float)iz, g.S), half);
    uint4* hdr = (uint4*)node;
    hdr[0] = make_uint4(hierarchy, (u32)ix, (u32)iy, (u32)iz);
    hdr[1] = make_uint4(ng + nov, ng, nov, __float_as_uint(g.S));
    hdr[2] = make_uint4(__float_as_uint(g.sub), __float_as_uint(px), __float_as_uint(py), __float_as_uint(pz));
    u64 eoff = 48ull + 16ull * ng;
    node[eoff] = (unsigned char)nent;
    eoff += 1;
    u32 dbase = cell_dbase[c], cb = cell_cbase[c];
    for (int o = 0; o < 8; o++) {
        u32 k = 8 * c + o;
        unsigned char fl = bflag[k];
        if (!(fl & PCB_BF_ENTRY)) continue;
        int child[3] = {2 * ix + ((o >> 2) & 1), 2 * iy + ((o >> 1) & 1), 2 * iz + (o & 1)};
        u32 kept = (fl & PCB_BF_FLUSH) ? 0u : bcount[k];
        store_bytes(node + eoff, child, 12);
        store_bytes(node + eoff + 12, &kept, 4);
        boff[k] = node_off[c] + eoff + 16;
        eoff += 16ull + 16ull * kept;
        if (fl & PCB_BF_FLUSH) {
            bdst[k] = dbase;
            next_cell_start[cb] = dbase;
            next_cell_idx[3 * cb + 0] = child[0];
            next_cell_idx[3 * cb + 1] = child[1];
            next_cell_idx[3 * cb + 2] = child[2];
            dbase += bcount[k];
            cb++;
        }
    }
    // zero the alignment padding after the node
    for (u64 b = eoff; b < ((cell_bytes[c] + 15ull) & ~15ull); b++) node[b] = 0;
    if (c == ncell - 1) next_cell_start[ncell_next] = n_next;
}

// Stage 8b: move every loser: delivered to its child cell (next level's arrivals, in (child, tau')
// order, with batch max(b_e, b_flush) — SURVEY N3) or kept in the parent's overflow list (image).
__global__ void k_loser_emit(const u32* __restrict__ los_src, const u32* __restrict__ los_bat,
                             const u32* __restrict__ los_bucket, u32 nlos, const PcbPoint* __restrict__ pts,
                             const u32* __restrict__ bstart, const unsigned char* __restrict__ bflag,
                             const u32* __restrict__ bfl, const u32* __restrict__ bcount,
                             const u32* __restrict__ bdst, const u64* __restrict__ boff,
                             PcbPoint* __restrict__ next_pts, u32* __restrict__ next_bat,
                             unsigned char* __restrict__ image) {
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nlos) return;
    u32 k = los_bucket[j];
    u32 n = bstart[k + 1] - bstart[k];
    u32 m0 = bcount[k] - n;
    u32 r = m0 + ((u32)j - bstart[k]);
    uint4 p = *(const uint4*)(pts + los_src[j]);
    if (bflag[k] & PCB_BF_FLUSH) {
        u32 dst = bdst[k] + r;
        *(uint4*)(next_pts + dst) = p;
        next_bat[dst] = max(los_bat[j], bfl[k]);
    } else {
        store16_at1(image + boff[k] + 16ull * r, p);
    }
}

// Stage 8c: grid section of every node: the winners, already in (cell, Z, Y, X) order.
__global__ void k_winner_emit(const u32* __restrict__ win_src, u32 nwin, const u32* __restrict__ win_end,
                              const u32* __restrict__ cell_of, const PcbPoint* __restrict__ pts,
                              const u64* __restrict__ node_off, unsigned char* __restrict__ image) {
    u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= nwin) return;
    u32 src = win_src[w];
    u32 c = cell_of[src];
    u32 wbeg = c ? win_end[c - 1] : 0u;
    *(uint4*)(image + node_off[c] + 48ull + 16ull * ((u32)w - wbeg)) = *(const uint4*)(pts + src);
}

// merge mode: the overflow lists that were already in the node files (front of each bucket)
__global__ void k_exlist_emit(const PcbPoint* __restrict__ ex_pts, u32 nex, const u32* __restrict__ ex_bucket,
                              const u32* __restrict__ ex_rank, const unsigned char* __restrict__ bflag,
                              const u32* __restrict__ bfl, const u32* __restrict__ bdst,
                              const u64* __restrict__ boff, PcbPoint* __restrict__ next_pts,
                              u32* __restrict__ next_bat, unsigned char* __restrict__ image) {
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nex) return;
    u32 k = ex_bucket[j], r = ex_rank[j];
    uint4 p = *(const uint4*)(ex_pts + j);
    if (bflag[k] & PCB_BF_FLUSH) {
        *(uint4*)(next_pts + bdst[k] + r) = p;
        next_bat[bdst[k] + r] = bfl[k];  // max(0, b_flush)
    } else {
        store16_at1(image + boff[k] + 16ull * r, p);
    }
}

// merge mode: put the incumbents (grid winners read back from the node files) in front of each cell's
// arrivals.  new_start[c] = cell_start[c] + inc_start[c].
__global__ void k_merge_arrivals(const PcbPoint* __restrict__ pts, const u32* __restrict__ bat, u32 n,
                                 const u32* __restrict__ cell_start, u32 ncell, const u32* __restrict__ inc_start,
                                 PcbPoint* __restrict__ out_pts, u32* __restrict__ out_bat) {
    u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u32 c = find_segment(cell_start, ncell, (u32)t);
    u32 dst = (u32)t + inc_start[c + 1];
    out_pts[dst] = pts[t];
    out_bat[dst] = bat[t];
}
__global__ void k_merge_incumbents(const PcbPoint* __restrict__ inc_pts, u32 ninc, const u32* __restrict__ inc_start,
                                   u32 ncell, const u32* __restrict__ cell_start, PcbPoint* __restrict__ out_pts,
                                   u32* __restrict__ out_bat) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ninc) return;
    u32 c = find_segment(inc_start, ncell, (u32)i);
    u32 dst = cell_start[c] + (u32)i;
    out_pts[dst] = inc_pts[i];
    out_bat[dst] = 0;
}
__global__ void k_merge_starts(u32* __restrict__ cell_start, const u32* __restrict__ inc_start, u32 ncell) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c <= ncell) cell_start[c] += inc_start[c];
}

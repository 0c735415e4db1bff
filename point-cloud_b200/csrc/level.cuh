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
#include "radix.cuh"
#include "scan.cuh"

struct LevelGeom {
    float S;        // cell edge of this level            metadata.rs:91-93
    float S_child;  // cell edge of level + 1
    float sub;      // S / sub_grid_dimension              metadata.rs:95-97
    float R;        // sub / 2.0 (hex circumradius)        cell.rs:276-278
    int bx, by, bz; // bits of the per-cell hex window per axis
    int mx, my, mz; // margins subtracted from the hex key of the cell's min corner
    int hexbits;    // bx + by + bz
    float half_sqrt3;  // SQRT_3 / 2.0          hex.rs:60 (loop-invariant sub-expressions, same single IEEE ops)
    float row_pitch;   // (R * 3.0) / 2.0       hex.rs:61
    HexInv inv;        // reciprocals (double) of the level's divisors, see fdiv_inv
};

// per-cell constants of a level, one 32-byte record
struct __align__(16) CellGeo {
    int hx, hy, hz;      // origin of the hex window (hex key of the cell's min corner minus the margin)
    u32 start;           // first arrival
    float tx, ty, tz;    // octant thresholds: floor(p / S_child) >= 2 * index + 1  <=>  p >= t   (see octant_threshold)
    u32 end;             // one past the last arrival
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
    const double invS0 = 1.0 / (double)S0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        PcbPoint p = pts[i];
        float v[3] = {p.x, p.y, p.z};
#pragma unroll
        for (int a = 0; a < 3; a++) {
            if (!isfinite(v[a])) bad |= PCB_DERR_NONFINITE;
            int o = f2ord(v[a]);
            mn[a] = min(mn[a], o);
            mx[a] = max(mx[a], o);
            int ci = cell_index_axis_inv(v[a], invS0);
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

// Input side: uncompressed LAS point records (fixed stride) -> 16-byte Points on the device.  converter/las.rs:32-41:
// pos = (x, y, z) as f32 of the scaled f64 coordinates (las 0.8.4: X * scale + offset, one f64 multiply and one f64
// add, never fused), colour = low byte of each little-endian u16 channel (`as u8`), alpha 255; formats without colour
// give (0, 0, 0).  A block stages the records of its 256 points in shared memory with coalesced word loads (records
// are 20-70 bytes at arbitrary alignment), then every thread decodes its own.
#define LAS_DEC_THREADS 256
#define LAS_DEC_MAX_STRIDE 96
__global__ void __launch_bounds__(LAS_DEC_THREADS)
k_las_decode(const unsigned char* __restrict__ rec, u32 n, u32 stride, int colour_off, double sx, double sy, double sz,
             double ox, double oy, double oz, PcbPoint* __restrict__ out) {
    __shared__ u32 s_words[LAS_DEC_THREADS * LAS_DEC_MAX_STRIDE / 4 + 2];
    const u64 p0 = (u64)blockIdx.x * LAS_DEC_THREADS;
    const u32 cnt = (u32)min((u64)LAS_DEC_THREADS, (u64)n - p0);
    const u64 b0 = p0 * stride, b1 = b0 + (u64)cnt * stride;
    const u64 w0 = b0 >> 2, w1 = (b1 + 3) >> 2;   // `rec` is 4-byte aligned and padded to a multiple of 4 bytes
    for (u64 w = w0 + threadIdx.x; w < w1; w += LAS_DEC_THREADS) s_words[w - w0] = __ldg((const u32*)rec + w);
    __syncthreads();
    if (threadIdx.x >= cnt) return;
    const unsigned char* p = (const unsigned char*)s_words + (b0 - (w0 << 2)) + (u64)threadIdx.x * stride;
    auto le32 = [&](int o) { return (i32)((u32)p[o] | ((u32)p[o + 1] << 8) | ((u32)p[o + 2] << 16) | ((u32)p[o + 3] << 24)); };
    PcbPoint q;
    q.x = __double2float_rn(__dadd_rn(__dmul_rn((double)le32(0), sx), ox));
    q.y = __double2float_rn(__dadd_rn(__dmul_rn((double)le32(4), sy), oy));
    q.z = __double2float_rn(__dadd_rn(__dmul_rn((double)le32(8), sz), oz));
    u32 rgba = 0xFF000000u;
    if (colour_off >= 0) rgba |= (u32)p[colour_off] | ((u32)p[colour_off + 2] << 8) | ((u32)p[colour_off + 4] << 16);
    q.rgba = rgba;
    out[p0 + threadIdx.x] = q;
}

// Sort-key layout of a level (host decides, kernels follow):
//   embedded (PAIRS = false): key = [cell ordinal : cb][Z : bz][Y : by][X : bx][arrival position : pb]
//   pairs    (PAIRS = true) : key = [cell ordinal : cb][Z][Y][X], value = arrival position
// The radix passes cover the cell + hex bits only; the arrival position in the low bits rides along, and
// because the input is in arrival order and every pass is stable, groups come out in arrival order.
struct KeyLayout {
    int pb;        // bits of the arrival position (0 in pairs mode)
    int hexbits;   // bx + by + bz
    int cb;        // bits of the cell ordinal
};

// Stage 0b: level-0 cell of every input point (converter.rs:110 group_points(points, 0)) as a dense
// linear id inside the index box [o, o + dim) known from the bounding box.  key = cell id << pb | input index.
struct L0Box {
    int ox, oy, oz;
    u32 nx, ny, nz;
};
// invS0 = 1.0 / (double)max_cell_size, once per thread (fdiv_inv: the reference's f32 division, bit for bit)
__device__ __forceinline__ u64 l0_cell_id(const PcbPoint& p, double invS0, const L0Box& b) {
    u64 ix = (u64)(u32)(cell_index_axis_inv(p.x, invS0) - b.ox);
    u64 iy = (u64)(u32)(cell_index_axis_inv(p.y, invS0) - b.oy);
    u64 iz = (u64)(u32)(cell_index_axis_inv(p.z, invS0) - b.oz);
    return (iz * b.ny + iy) * b.nx + ix;
}
__global__ void __launch_bounds__(256)
k_l0_key(const PcbPoint* __restrict__ pts, u32 n, float S0, L0Box box, int pb, RadixPlan plan,
         u64* __restrict__ keys, u32* __restrict__ vals /* pairs mode (pb == 0), else null */, u32* __restrict__ ghist) {
    __shared__ u32 hist[RADIX_MAX_PASSES * RADIX_MAX_BINS];
    for (int i = threadIdx.x; i < plan.npass * RADIX_MAX_BINS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const double invS0 = 1.0 / (double)S0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 k = l0_cell_id(pts[i], invS0, box);
        if (vals) vals[i] = (u32)i; else k = (k << pb) | i;
        keys[i] = k;
        radix_hist_add(hist, plan, k);
    }
    __syncthreads();
    radix_hist_flush(hist, plan, ghist);
}

// Stage 0c: distinct level-0 cells of the sorted keys -> cell list + segment starts.
struct L0HeadOp {
    typedef u32 State;
    static constexpr bool kTileLoad = false;
    const u64* keys;
    u32 n;
    int pb;
    L0Box box;
    u32* cell_start;  // [ncell + 1]
    int* cell_idx;    // [ncell * 3]
    u32* ncell_out;
    __device__ static void pack(const State& s, u32 status, uint4& a, uint4& b) {
        a = make_uint4(s, 0u, 0u, status);
        b = make_uint4(0u, 0u, 0u, status);
    }
    __device__ static State unpack(const uint4& a, const uint4&) { return a.x; }
    __device__ State identity() const { return 0; }
    __device__ State load(u32 i) const { return (i == 0 || (keys[i] >> pb) != (keys[i - 1] >> pb)) ? 1u : 0u; }
    __device__ State combine(const State& a, const State& b) const { return a + b; }
    __device__ void consume(u32 i, const State& excl, const State& incl, const State&) const {
        if (incl != excl) {
            cell_start[excl] = i;
            u64 k = keys[i] >> pb;
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
// (lib.rs:31-52: batches of `batch_size` points, never spanning files) or, when the caller supplied them
// (multi-GPU: the exchange keeps global batch ordinals), from in_bat.
struct BatchSeg {
    u32 start;       // first input index of the segment
    u32 first_batch; // global ordinal of its first batch
    u32 batch_size;  // points per batch inside the segment
    u32 pad;
};
__global__ void k_l0_gather(const PcbPoint* __restrict__ in_pts, const u64* __restrict__ keys,
                            const u32* __restrict__ vals, u32 n, int pb, const BatchSeg* __restrict__ segs, u32 nseg, const u32* __restrict__ in_bat,
                            PcbPoint* __restrict__ pts, u32* __restrict__ bat) {
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    u32 i = vals ? vals[j] : (u32)(keys[j] & ((1ull << pb) - 1ull));
    pts[j] = in_pts[i];
    if (in_bat) {
        bat[j] = in_bat[i] + 1u;
        return;
    }
    u32 lo = 0, hi = nseg;
    while (hi - lo > 1) {
        u32 mid = (lo + hi) >> 1;
        if (segs[mid].start <= i) lo = mid; else hi = mid;
    }
    BatchSeg s = segs[lo];
    bat[j] = s.first_batch + (i - s.start) / s.batch_size + 1u;
}

// ---- level-0 arrivals in ONE pass when the level-0 box has at most 512 cells ---------------------------------
// (the usual case: 1000-unit cells).  k_l0_key + a radix pass + the head scan + k_l0_gather move every point
// 16 R + 8 W, 8 R + 8 W, 8 R, 8 + 16 R + 20 W; here a dense per-cell histogram (16 R, k_route_hist) gives the first
// slot of every cell, and k_l0_split - a onesweep pass whose "key" is the point itself and whose digit is the dense
// cell id - reads each point once and writes it (and its batch ordinal) once at its final place: stable ranking
// per warp, tile staged in shared memory in cell order, decoupled look-back over per-tile cell counts.
#define L0S_THREADS 256
#define L0S_ITEMS 8
#define L0S_TILE (L0S_THREADS * L0S_ITEMS)
#define L0S_WARPS (L0S_THREADS / 32)
#define L0S_MAX_BINS 512

// one block: dense histogram -> first slot of every dense cell + the CSR of the non-empty cells (sorted by dense id,
// the order of the level-0 keys: z, y, x)
__global__ void __launch_bounds__(L0S_MAX_BINS)
k_l0_cells(const u32* __restrict__ hist, u32 vol, L0Box box, u32 n, u32* __restrict__ dbase, u32* __restrict__ cell_start,
           int* __restrict__ cell_idx, u32* __restrict__ ncell_out) {
    __shared__ u32 s_cnt[L0S_MAX_BINS], s_ne[L0S_MAX_BINS];
    const u32 t = threadIdx.x;
    const u32 c = t < vol ? hist[t] : 0u;
    s_cnt[t] = c;
    s_ne[t] = c ? 1u : 0u;
    __syncthreads();
    for (u32 o = 1; o < L0S_MAX_BINS; o <<= 1) {   // Hillis-Steele inclusive scans
        u32 a = t >= o ? s_cnt[t - o] : 0u, b = t >= o ? s_ne[t - o] : 0u;
        __syncthreads();
        s_cnt[t] += a;
        s_ne[t] += b;
        __syncthreads();
    }
    if (t < vol) {
        const u32 first = s_cnt[t] - c;
        dbase[t] = first;
        if (c) {
            const u32 ord = s_ne[t] - 1u;
            cell_start[ord] = first;
            cell_idx[3 * ord + 0] = box.ox + (int)(t % box.nx);
            cell_idx[3 * ord + 1] = box.oy + (int)((t / box.nx) % box.ny);
            cell_idx[3 * ord + 2] = box.oz + (int)(t / (box.nx * box.ny));
        }
    }
    if (t == L0S_MAX_BINS - 1) {
        cell_start[s_ne[t]] = n;
        *ncell_out = s_ne[t];
    }
}

struct L0SplitSmem {
    static constexpr int bytes = L0S_TILE * 16 + L0S_TILE * 4 + L0S_TILE * 2 + L0S_WARPS * L0S_MAX_BINS * 4 +
                                 2 * L0S_MAX_BINS * 4;
};

// ROUTE (multi-GPU, route.cuh): the same pass IS the level-0 exchange - cell d's points go to the receive buffer of rank
// owner[d] (peer-mapped memory reached over NVLink, or local memory), at slot dbase[d] + rank among this slice's points
// of d, with their global batch ordinal (first_index + i) / route_batch + 1: every owner receives its level-0 arrivals
// already in (cell, stream order), no second split on the receiving side.
struct RouteDst {
    PcbPoint* pts[8];
    u32* bat[8];
    u32 off[8];
};
template <bool ROUTE>
__global__ void __launch_bounds__(L0S_THREADS, 4)   // 64 registers, four blocks per SM (three: 1.63 ms, four: 1.37 ms per 100 M points)
k_l0_split(const PcbPoint* __restrict__ in_pts, u32 n, float S0, L0Box box, u32 nb, int bits,
           const u32* __restrict__ dbase, const BatchSeg* __restrict__ segs, u32 nseg, const u32* __restrict__ in_bat,
           PcbPoint* __restrict__ pts, u32* __restrict__ bat, u32* __restrict__ status /* [ntiles][nb] */,
           u32* __restrict__ err, const unsigned char* __restrict__ owner, RouteDst dst8, u64 first_index, u32 route_batch) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint4* s_pts = (uint4*)smem_raw;
    u32* s_idx = (u32*)(s_pts + L0S_TILE);
    unsigned short* s_dig = (unsigned short*)(s_idx + L0S_TILE);
    u32* cnt = (u32*)(s_dig + L0S_TILE);   // [WARPS][nb]
    u32* dig_excl = cnt + L0S_WARPS * nb;
    u32* gbase = dig_excl + nb;
    __shared__ u32 s_warp_sums[L0S_WARPS];
    __shared__ unsigned char s_own[ROUTE ? L0S_MAX_BINS : 4];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 tile = blockIdx.x;   // tile = block index, see k_chain_scan
    const u64 tbase = (u64)tile * L0S_TILE;
    const u32 tile_n = (u32)min((u64)L0S_TILE, (u64)n - tbase);
    for (u32 i = tid; i < L0S_WARPS * nb; i += L0S_THREADS) cnt[i] = 0;
    if (ROUTE)
        for (u32 i = tid; i < nb; i += L0S_THREADS) s_own[i] = owner[i] & 7u;
    __syncthreads();

    // warp-striped ownership: item k of lane l in warp w is element w * 32 * ITEMS + 32 k + l of the tile
    uint4 p[L0S_ITEMS];
    u32 dig[L0S_ITEMS];
    const double invS0 = 1.0 / (double)S0;
    const u32 wbase = warp * 32 * L0S_ITEMS + lane;
    u32* wcnt = cnt + warp * nb;
#pragma unroll
    for (int k = 0; k < L0S_ITEMS; k++) {
        const u32 j = wbase + 32 * k;
        dig[k] = nb;
        if (j < tile_n) {
            p[k] = __ldg((const uint4*)(in_pts + tbase + j));
            PcbPoint q;
            q.x = __uint_as_float(p[k].x); q.y = __uint_as_float(p[k].y); q.z = __uint_as_float(p[k].z); q.rgba = p[k].w;
            dig[k] = (u32)l0_cell_id(q, invS0, box);
            atomicAdd(&wcnt[dig[k]], 1u);   // early counts (see k_onesweep)
        }
    }
    __syncthreads();

    constexpr u32 DPT = L0S_MAX_BINS / L0S_THREADS;   // 2 digits per thread at most
    const u32 d0 = tid * DPT;
    u32 count[DPT];
    u32 tsum = 0;
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        const u32 d = d0 + q;
        count[q] = 0;
        if (d < nb) {
            u32 running = 0;
#pragma unroll
            for (int w = 0; w < L0S_WARPS; w++) running += cnt[w * nb + d];
            count[q] = running;
            *(volatile u32*)&status[(u64)tile * nb + d] = running | (tile == 0 ? RADIX_FLAG_INCL : RADIX_FLAG_AGG);
        }
        tsum += count[q];
    }
    u32 tinc = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        u32 a = __shfl_up_sync(0xFFFFFFFFu, tinc, o);
        if (lane >= o) tinc += a;
    }
    if (lane == 31) s_warp_sums[warp] = tinc;
    __syncthreads();
    u32 toff = 0;
#pragma unroll
    for (int w = 0; w < L0S_WARPS; w++)
        if (w < warp) toff += s_warp_sums[w];
    u32 texcl = toff + tinc - tsum;
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        const u32 d = d0 + q;
        if (d < nb) {
            u32 running = texcl;
#pragma unroll
            for (int w = 0; w < L0S_WARPS; w++) {
                u32 c = cnt[w * nb + d];
                cnt[w * nb + d] = running;   // first staged slot of (warp, cell)
                running += c;
            }
            dig_excl[d] = texcl;
            texcl += count[q];
        }
    }
    __syncthreads();

    // stable ranking inside the warp, points go straight to their staged slot
    const u32 lt = (1u << lane) - 1u;
    bool votes = false;
#pragma unroll
    for (int k = 0; k < L0S_ITEMS; k++) {
        const u32 j = wbase + 32 * k;
        const bool valid = j < tile_n;
        const u32 d = valid ? dig[k] : 0u;
        const u32 peers = radix_peers<2>(d, valid, bits, 1u << bits, k, votes);
        const u32 lower = __popc(peers & lt);
        const u32 before = valid ? wcnt[d] : 0u;
        __syncwarp();
        if (valid && lower == 0) wcnt[d] = before + __popc(peers);
        __syncwarp();
        if (valid) {
            s_pts[before + lower] = p[k];
            s_idx[before + lower] = (u32)(tbase + j);
            s_dig[before + lower] = (unsigned short)d;
        }
    }

    // decoupled look-back per cell over the tiles before this one
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        const u32 d = d0 + q;
        if (d < nb) {
            u32 prev = 0;
            if (tile > 0) {
                int t = (int)tile - 1;
                const volatile u32* sp = status + (u64)t * nb + d;
                bool done = false;
                while (!done) {
                    u32 w[RADIX_WINDOW];
#pragma unroll
                    for (int x = 0; x < RADIX_WINDOW; x++) w[x] = (t - x >= 0) ? *(sp - (size_t)x * nb) : RADIX_FLAG_INCL;
#pragma unroll
                    for (int x = 0; x < RADIX_WINDOW; x++) {
                        if (!done) {
                            u32 sw = w[x];
                            u32 spins = 0;
                            while ((sw & (RADIX_FLAG_AGG | RADIX_FLAG_INCL)) == 0u) {
                                sw = *(sp - (size_t)x * nb);
                                if (++spins > PCB_SPIN_LIMIT) {
                                    if (err) atomicOr(err, PCB_DERR_TIMEOUT);
                                    sw = RADIX_FLAG_INCL;
                                }
                            }
                            prev += sw & RADIX_COUNT_MASK;
                            done = (sw & RADIX_FLAG_INCL) != 0u;
                        }
                    }
                    t -= RADIX_WINDOW;
                    sp -= (size_t)RADIX_WINDOW * nb;
                }
                *(volatile u32*)&status[(u64)tile * nb + d] = ((prev + count[q]) & RADIX_COUNT_MASK) | RADIX_FLAG_INCL;
            }
            gbase[d] = __ldg(dbase + d) + prev - dig_excl[d];
        }
    }
    __syncthreads();
    for (u32 j = tid; j < tile_n; j += L0S_THREADS) {
        const u32 g = gbase[s_dig[j]] + j;
        const u32 i = s_idx[j];
        if (ROUTE) {
            const u32 o = s_own[s_dig[j]];
            PcbPoint* dp = nullptr;
            u32* db = nullptr;
#pragma unroll
            for (int q = 0; q < 8; q++) {
                dp = (o == (u32)q) ? dst8.pts[q] : dp;
                db = (o == (u32)q) ? dst8.bat[q] : db;
            }
            *(uint4*)(dp + g) = s_pts[j];
            db[g] = (u32)((first_index + i) / route_batch) + 1u;
            continue;
        }
        *(uint4*)(pts + g) = s_pts[j];
        u32 b;
        if (in_bat) {
            b = __ldg(in_bat + i) + 1u;
        } else {
            u32 lo = 0, hi = nseg;
            while (hi - lo > 1) {
                u32 mid = (lo + hi) >> 1;
                if (segs[mid].start <= i) lo = mid; else hi = mid;
            }
            const BatchSeg sg = segs[lo];
            b = sg.first_batch + (i - sg.start) / sg.batch_size + 1u;
        }
        bat[g] = b;
    }
}

// ---------------------------------------------------------------------------------------------
// The child octant of a loser is bit-for-bit cell_index(p, level + 1) - 2 * cell_index (converter.rs:68,
// metadata.rs:100-102).  x -> floor(x / S_child) is monotone, so for the odd child index m = 2 * i + 1 there is
// one smallest float t with floor(t / S_child) >= m, and "p lands in the upper half" is exactly p >= t.  The
// threshold is found once per cell and axis with the reference's own division, so that the per-point work is a
// compare instead of an IEEE division.
__device__ __forceinline__ float octant_threshold(int i, float S_child, u32* err) {
    const long long m = 2ll * i + 1ll;
    float t = __fmul_rn((float)m, S_child);
    int guard = 0;
    // walk down while the predecessor still reaches m, then up until t itself does
    while (guard++ < 64 && (long long)cell_index_axis(nextafterf(t, -INFINITY), S_child) >= m) t = nextafterf(t, -INFINITY);
    while (guard++ < 128 && (long long)cell_index_axis(t, S_child) < m) t = nextafterf(t, INFINITY);
    if ((long long)cell_index_axis(t, S_child) < m || (long long)cell_index_axis(nextafterf(t, -INFINITY), S_child) >= m)
        atomicOr(err, PCB_DERR_NEST);
    return t;
}

// Stage 1: per-cell constants (hex window origin, octant thresholds, arrival range).
__global__ void k_cellgeo(const int* __restrict__ cell_idx, const u32* __restrict__ cell_start, u32 ncell, LevelGeom g,
                          CellGeo* __restrict__ geo, u32* __restrict__ err) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    const int ix = cell_idx[3 * c + 0], iy = cell_idx[3 * c + 1], iz = cell_idx[3 * c + 2];
    float x = __fmul_rn((float)ix, g.S);
    float y = __fmul_rn((float)iy, g.S);
    float z = __fmul_rn((float)iz, g.S);
    HexKey k = hex_from_world(x, y, z, g.inv);
    CellGeo o;
    o.hx = k.x - g.mx; o.hy = k.y - g.my; o.hz = k.z - g.mz;
    o.start = cell_start[c];
    o.tx = octant_threshold(ix, g.S_child, err);
    o.ty = octant_threshold(iy, g.S_child, err);
    o.tz = octant_threshold(iz, g.S_child, err);
    o.end = cell_start[c + 1];
    geo[c] = o;
}
__device__ __forceinline__ CellGeo load_geo(const CellGeo* __restrict__ geo, u32 c) {
    const uint4* p = (const uint4*)(geo + c);
    uint4 a = __ldg(p), b = __ldg(p + 1);
    CellGeo o;
    o.hx = (int)a.x; o.hy = (int)a.y; o.hz = (int)a.z; o.start = a.w;
    o.tx = __uint_as_float(b.x); o.ty = __uint_as_float(b.y); o.tz = __uint_as_float(b.z); o.end = b.w;
    return o;
}

// cell of arrival i: one binary search per warp (on the warp's first element), then a short walk
__device__ __forceinline__ u32 cell_of_arrival(const u32* __restrict__ cell_start, u32 ncell, u32 i, u32 warp_first) {
    u32 c = find_segment(cell_start, ncell, warp_first);
    while (__ldg(cell_start + c + 1) <= i) c++;
    return c;
}

// Stage 2: hex-prism key (cell.rs:276-278 -> hex.rs:32-34) of every arrival, window-relative, plus the digit
// histograms of all radix passes.
template <bool PAIRS, int MINB>
__global__ void __launch_bounds__(256, MINB)   // MINB = 8: 32 registers, eight blocks per SM (37 registers ran 20 % slower)
k_hexkey(const PcbPoint* __restrict__ pts, u32 n, const u32* __restrict__ cell_start, u32 ncell,
         const CellGeo* __restrict__ geo, LevelGeom g, KeyLayout lay, RadixPlan plan, u64* __restrict__ keys,
         u32* __restrict__ vals, uint2* __restrict__ d2oct_out,
         u32* __restrict__ ghist, int segk /* segment = cell >> segk (31: one segment) */, u32 nseg,
         const u32* __restrict__ estart /* record filter: first early arrival of every cell in keys_e, else null */,
         u64* __restrict__ keys_e, u32* __restrict__ err) {
    __shared__ u32 hist[RADIX_MAX_PASSES * RADIX_MAX_BINS];
    for (int i = threadIdx.x; i < RADIX_MAX_PASSES * RADIX_MAX_BINS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    u32 bad = 0;
    // every block owns one contiguous chunk of arrivals: the cell ordinal is searched once, then walked
    // (n < 2^30 per level, builder.cu: 32-bit indices cannot wrap, and addresses are one IMAD.WIDE each)
    const u32 chunk = (u32)(((((u64)n + gridDim.x - 1) / gridDim.x) + 255ull) & ~255ull);
    const u32 lo = (u32)min((u64)n, (u64)blockIdx.x * chunk), hi = (u32)min((u64)n, (u64)lo + chunk);
    const u32 seg_lo = lo < hi ? find_segment(cell_start, ncell, lo) >> segk : 0u;   // first segment of the block
    const u32 seg_words = (u32)plan.npass * RADIX_MAX_BINS, slots = (u32)radix_hist_slots(plan);
    const bool p3x8 = radix_plan_is_3x8(plan);
    u32 c = 0;
    CellGeo cg;
    cg.end = 0;
    u32 slot = 0;   // segment of cell c relative to the block's first segment (changes with the cell: rarely)
    u32 ebase = 0, elen = 0xFFFFFFFFu;   // record filter: the cell's early arrivals go to keys_e[ebase ...), elen of them
    // key = cell | prism (| arrival position): the cell part only changes with the cell, the prism part fits 32 bits for
    // every grid whose window has <= 32 bits - one 64-bit shift per arrival instead of four
    const bool narrow = g.hexbits <= 32;
    const int cshift = g.hexbits + (PAIRS ? 0 : lay.pb);
    u64 cbase = 0;
    if (lo + threadIdx.x < hi) {
        c = find_segment(cell_start, ncell, lo + threadIdx.x);
        cg = load_geo(geo, c);
        slot = (c >> segk) - seg_lo;
        cbase = (u64)c << cshift;
        if (estart) { ebase = __ldg(estart + c); elen = __ldg(estart + c + 1) - ebase; }
    }
    for (u32 i0 = lo; i0 < hi; i0 += 256) {
        const u32 i = i0 + threadIdx.x;
        if (i < hi) {
            PcbPoint p = pts[i];
            if (cg.end <= i) {
                do cg = load_geo(geo, ++c); while (cg.end <= i);
                slot = (c >> segk) - seg_lo;
                cbase = (u64)c << cshift;
                if (estart) { ebase = __ldg(estart + c); elen = __ldg(estart + c + 1) - ebase; }
            }
            HexKey k = hex_from_world(p.x, p.y, p.z, g.inv);
            u32 rx = (u32)(k.x - cg.hx);
            u32 ry = (u32)(k.y - cg.hy);
            u32 rz = (u32)(k.z - cg.hz);
            if ((rx >> g.bx) | (ry >> g.by) | (rz >> g.bz)) {
                bad |= PCB_DERR_HEXRANGE;
                rx &= (1u << g.bx) - 1u; ry &= (1u << g.by) - 1u; rz &= (1u << g.bz) - 1u;
            }
            u64 key;
            if (narrow) {
                const u32 prism = (rz << (g.bx + g.by)) | (ry << g.bx) | rx;
                key = PAIRS ? (cbase | prism) : (cbase | ((u64)prism << lay.pb) | i);
            } else {
                key = ((u64)c << g.hexbits) | ((u64)rz << (g.bx + g.by)) | ((u64)ry << g.bx) | (u64)rx;
                if (!PAIRS) key = (key << lay.pb) | i;
            }
            if (PAIRS) vals[i] = i;
            keys[i] = key;
            // what the contest needs of this point, by arrival position: ONE 8-byte record (d2, child octant) that
            // stays L2-resident per cell, instead of a random 16-byte gather of the point itself (one DRAM sector per
            // arrival).  One record = one L2 sector per gather; d2 and octant in separate arrays cost two.
            d2oct_out[i] = make_uint2(__float_as_uint(hex_d2(k, p.x, p.y, p.z, g.R, g.half_sqrt3, g.row_pitch)),
                                      (p.x >= cg.tx ? 4u : 0u) | (p.y >= cg.ty ? 2u : 0u) | (p.z >= cg.tz ? 1u : 0u));
            const u32 rel = i - cg.start;
            if (rel < elen) {   // every arrival without the record filter, else the early ones: these get sorted
                if (estart) keys_e[ebase + rel] = key;
                if (slot < slots) {
                    if (p3x8) radix_hist_add_3x8(hist + slot * seg_words, plan.shift[0], key, slot);
                    else radix_hist_add(hist + slot * seg_words, plan, key, slot);
                }
                else radix_hist_add_global(ghist + (u64)(seg_lo + slot) * seg_words, plan, key);
            }
        }
    }
    for (int o = 16; o; o >>= 1) bad |= __shfl_xor_sync(0xFFFFFFFFu, bad, o);
    if ((threadIdx.x & 31) == 0 && bad) atomicOr(err, bad);
    __syncthreads();
    radix_hist_flush_seg(hist, plan, ghist, seg_lo, nseg);
}

// Stage 4: the hex-grid contest of Cell::add_point (cell.rs:70-94) for every (cell, prism) group at
// once.  In sorted order a group is contiguous and in arrival order; the incumbent seen by an arrival
// is the running minimum of (d2, position) over the earlier members (strict `<`: equal d2 keeps the
// earlier one).  Every non-first member emits exactly one loser — itself, or the incumbent it displaces
// — and the emission is recorded under the position of the arrival that triggered it, together with the
// loser's child octant (converter.rs:68 group_points(losers, h + 1); octant = dx*4 + dy*2 + dz, so octant
// order == lexicographic (x, y, z) order of the child index).
#define PCB_EM_NONE_OCT 8u
#define PCB_EM_CELL_LAST (1ull << 40)    // this arrival position is the last one of its cell
#define PCB_EM_CELL_FIRST (1ull << 41)   // ... the first one
struct ContestState {
    u64 mn;     // min (d2 bits << 32 | pos) since the last group head
    u32 meta;   // bits 0-3: child octant of mn's point; bit 8: a group head lies inside the scanned range;
                // (single elements only) bit 9: group tail, bit 10: last element of its cell in sorted order
    u32 ntail;  // groups completed (== winners emitted)
};
template <bool PAIRS>
struct ContestOp {
    typedef ContestState State;
    static constexpr bool kTileLoad = true;
    static constexpr bool kOneWord = true;   // the 16-byte state and the status bits share one descriptor word
    static constexpr int ITEMS = 8;
    const u64* keys;
    const u32* vals;
    u32 n;
    KeyLayout lay;
    LevelGeom g;
    const uint2* d2oct;   // [n] by arrival position, from k_hexkey: .x = d2 to the prism centre (hex.rs:55-65, cell.rs:77-78),
                          // .y = child octant (metadata.rs:100-102 one level down)
    const CellGeo* geo;
    u64* em;       // [n] by arrival position: loser's arrival position | octant << 32 | flags
    u32* win_src;  // [#groups] arrival position of each winner, sorted (cell, Z, Y, X)
    u32* win_end;  // [ncell] end of the cell's winners in win_src
    u32* nwin_out;
    u32* err;
    // record filter (both null otherwise).  Round 1 (early arrivals): tails[w] = {group id, d2 bits, position} of
    // winner w.  Round 2 (early winners + candidates): estart != null, elements that are early arrivals are the
    // incumbents of their groups - their own emission was decided in round 1 and is not touched.
    uint4* tails;
    const u32* estart;
    const u32* n_dev;   // element count on the device (then `n` is an upper bound, see k_chain_scan), else null
    __device__ static void pack(const State& s, u32 status, uint4& a, uint4& b) {
        a = make_uint4((u32)s.mn, (u32)(s.mn >> 32), s.meta, s.ntail | status);   // ntail < 2^30
        b = make_uint4(0u, 0u, 0u, status);
    }
    __device__ static State unpack(const uint4& a, const uint4&) {
        return State{((u64)a.y << 32) | a.x, a.z, a.w & ~PCB_DESC_MASK};
    }
    __device__ u64 gid(u64 key) const { return PAIRS ? key : (key >> lay.pb); }
    __device__ State identity() const { return State{~0ull, 0u, 0u}; }
    // meta bits of a single element: 0-3 child octant, 8 group head, 9 group tail, 10 last of its cell in sorted
    // order, 11 first arrival of its cell, 12 last arrival of its cell
    __device__ void load_items(u64 base, State (&v)[ITEMS]) const {
        const int lane = threadIdx.x & 31;
        const u32 n = n_dev ? __ldg(n_dev) : this->n;
        u64 key[ITEMS];
        u32 pos[ITEMS];
        if (base + ITEMS <= n) {
            const ulonglong2* src = (const ulonglong2*)(keys + base);
#pragma unroll
            for (int k = 0; k < ITEMS / 2; k++) {
                ulonglong2 t = src[k];
                key[2 * k] = t.x;
                key[2 * k + 1] = t.y;
            }
        } else {
#pragma unroll
            for (int k = 0; k < ITEMS; k++) key[k] = (base + k < n) ? keys[base + k] : ~0ull;
        }
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            if (PAIRS) pos[k] = (base + k < n) ? vals[base + k] : 0u;
            else pos[k] = (u32)(key[k] & ((1ull << lay.pb) - 1ull));
        }
        // the gathers go first: they are the long-latency part
        u32 d2v[ITEMS], ocv[ITEMS];
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const uint2 rec = (base + k < n) ? __ldg(d2oct + pos[k]) : make_uint2(0u, 0u);
            d2v[k] = rec.x;
            ocv[k] = rec.y;
        }
        // neighbours across the thread boundary
        u64 prevk = __shfl_up_sync(0xFFFFFFFFu, key[ITEMS - 1], 1);
        u64 nextk = __shfl_down_sync(0xFFFFFFFFu, key[0], 1);
        if (lane == 0) prevk = (base > 0 && base <= n) ? keys[base - 1] : ~0ull;
        if (lane == 31) nextk = (base + ITEMS < n) ? keys[base + ITEMS] : ~0ull;
        u32 cprev = PCB_NONE;
        CellGeo cg;
        cg.hx = cg.hy = cg.hz = 0;
        cg.tx = cg.ty = cg.tz = 0.0f;
        cg.start = cg.end = 0;
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const u64 i = base + k;
            if (i >= n) {
                v[k] = identity();
                continue;
            }
            const u64 G = gid(key[k]);
            const u64 Gp = gid(k ? key[k - 1] : prevk), Gn = gid(k + 1 < ITEMS ? key[k + 1] : nextk);
            const u32 c = (u32)(G >> g.hexbits);
            if (c != cprev) {
                cg = load_geo(geo, c);
                cprev = c;
            }
            u32 meta = ocv[k];
            const bool head = (i == 0) || Gp != G;
            const bool last = (i + 1 == n);
            const bool tail = last || Gn != G;
            const bool cell_end = last || (Gn >> g.hexbits) != (G >> g.hexbits);
            meta |= (head ? 0x100u : 0u) | (tail ? 0x200u : 0u) | (cell_end ? 0x400u : 0u);
            meta |= (cg.start == pos[k] ? 0x800u : 0u) | (cg.end == pos[k] + 1u ? 0x1000u : 0u);
            v[k] = State{((u64)d2v[k] << 32) | pos[k], meta, tail ? 1u : 0u};
        }
    }
    __device__ State load(u32) const { return identity(); }   // unused (kTileLoad)
    __device__ State combine(const State& a, const State& b) const {
        State r;
        const bool take_b = (b.meta & 0x100u) || b.mn < a.mn;   // equal only for identity: positions are distinct
        r.mn = take_b ? b.mn : a.mn;
        r.meta = ((take_b ? b.meta : a.meta) & 0xFu) | ((a.meta | b.meta) & 0x100u);
        r.ntail = a.ntail + b.ntail;
        return r;
    }
    __device__ void consume(u32 j, const State& excl, const State& incl, const State& self) const {
        const u32 pos = (u32)self.mn;
        u64 e;
        if (self.meta & 0x100u) {
            e = (u64)PCB_EM_NONE_OCT << 32 | pos;
        } else if ((u32)(self.mn >> 32) < (u32)(excl.mn >> 32)) {          // cell.rs:80 strict <: displaces the incumbent
            e = ((u64)(excl.meta & 0xFu) << 32) | (u32)excl.mn;
        } else {
            e = ((u64)(self.meta & 0xFu) << 32) | pos;
        }
        if (self.meta & 0x200u) {
            win_src[excl.ntail] = (u32)incl.mn;
            if (tails) {
                const u64 G = gid(keys[j]);
                tails[excl.ntail] = make_uint4((u32)G, (u32)(G >> 32), (u32)(incl.mn >> 32), (u32)incl.mn);
            }
        }
        if (self.meta & 0x400u) win_end[(u32)(gid(keys[j]) >> g.hexbits)] = excl.ntail + 1;
        if (self.meta & 0x800u) e |= PCB_EM_CELL_FIRST;
        if (self.meta & 0x1000u) e |= PCB_EM_CELL_LAST;
        if (estart) {
            const u32 c = (u32)(gid(keys[j]) >> g.hexbits);
            if (pos - __ldg(&geo[c].start) < __ldg(estart + c + 1) - __ldg(estart + c)) return;   // incumbent
        }
        em[pos] = e;
    }
    __device__ void finish(const State& total) const { *nwin_out = total.ntail; }
};

// ---- record filter ---------------------------------------------------------------------------------------------------
// On the upper levels a prism receives hundreds of arrivals and all but a handful of them lose to an incumbent
// that arrived long before (cell.rs:80: only an arrival with a strictly smaller d2 than everything before it - a
// RECORD - changes the grid).  An arrival that is not a record emits itself, and that needs no sort.  So a level is
// built in two rounds:
//   round 1: the EARLY arrivals of every cell (its first share, in arrival order) go through sort + contest as
//            usual; the winners so far (prism, d2) are put into a hash table;
//   filter : every later arrival looks its prism up: d2 >= the table's d2 means an earlier arrival is at least as
//            close, so it is no record and emits itself (coalesced store, done).  The others - CANDIDATES, a few
//            per cent - may be records;
//   round 2: early winners + candidates go through sort + contest.  Every true record is an early arrival or a
//            candidate, the incumbent seen by a candidate is always a record, so the running minima over this subset
//            are the true ones: emissions and final winners are exactly those of the full contest.
// One 8-byte slot per early winner: (group id + 1) << 24 | ceil(d2 bits / 256); 0 = empty; linear probing.  The 24-bit
// bound is rounded UP, so "arrival's d2 bits >> 8 >= bound" still proves d2 >= the early winner's d2 (a few more
// candidates, never a missed record).  Group ids need 40 bits (cell + prism bits; the filter is off beyond that).
#define FILTER_GID_BITS 40
struct FilterTable {
    unsigned long long* slots;
    u32 mask;    // capacity - 1 (power of two)
    int shift;   // 64 - log2(capacity): Fibonacci hashing takes the top bits of the product
};
__device__ __forceinline__ u32 filter_hash(u64 gid, const FilterTable& T) {
    return (u32)((gid * 0x9E3779B97F4A7C15ull) >> T.shift);
}
// winners of round 1 -> table + their keys (dense list, one per winner)
__global__ void k_filter_table(const uint4* __restrict__ tails, const u32* __restrict__ nwin_ptr, FilterTable T, int pb,
                               u64* __restrict__ wkeys) {
    const u32 nwin = *nwin_ptr;
    for (u32 w = blockIdx.x * blockDim.x + threadIdx.x; w < nwin; w += gridDim.x * blockDim.x) {
        const uint4 t = tails[w];
        const u64 gid = ((u64)t.y << 32) | t.x;
        const unsigned long long entry = ((gid + 1ull) << 24) | (unsigned long long)((t.z + 0xFFu) >> 8);
        u32 h = filter_hash(gid, T);
        while (atomicCAS(T.slots + h, 0ull, entry) != 0ull) h = (h + 1u) & T.mask;   // group ids are distinct: never an update
        wkeys[w] = (gid << pb) | t.w;
    }
}
// Candidates are appended to one private list per WARP of k_filter (capacity = the arrivals that warp looks at, so
// the lists tile one buffer of ~n keys): no atomics at all.  One shared counter, even 128 of them, serialised
// millions of atomics on a few L2 lines and cost 1.5-2 ms per level.  Every warp owns a CONTIGUOUS range of arrivals,
// so the lists back to back hold the candidates in arrival order: the round-2 list (early winners first, then the
// candidates) only needs a stable sort on the cell and prism bits, like any level.
// every arrival behind the early share of its cell: no record -> emits itself; else candidate for round 2
__global__ void __launch_bounds__(256)
k_filter(const u64* keys /* by arrival position; aliases em */, const uint2* __restrict__ d2oct, u32 n,
         const u32* __restrict__ cell_start, u32 ncell, const CellGeo* __restrict__ geo, const u32* __restrict__ estart,
         int pb, FilterTable T, u64* __restrict__ cand_lists /* [gridDim.x * 8] lists of chunk / 8 keys */,
         u32* __restrict__ ncand /* [gridDim.x * 8] */, u64* em) {
    const u64 chunk = (((u64)n + gridDim.x - 1) / gridDim.x + 255ull) & ~255ull;
    const u64 lo = (u64)blockIdx.x * chunk, hi = min((u64)n, lo + chunk);
    const u32 list_id = blockIdx.x * 8u + (threadIdx.x >> 5);
    u64* const list = cand_lists + (u64)list_id * (chunk / 8);
    u32 filled = 0;   // keys in this warp's list (same in every lane)
    const int lane = threadIdx.x & 31;
    const u64 wlo = lo + (u64)(threadIdx.x >> 5) * (chunk / 8), whi = min(hi, wlo + chunk / 8);   // this warp's arrivals
    u32 c = 0, elen = 0;
    CellGeo cg;
    cg.start = cg.end = 0;
    if (wlo + lane < whi) {
        c = find_segment(cell_start, ncell, (u32)(wlo + lane));
        cg = load_geo(geo, c);
        elen = __ldg(estart + c + 1) - __ldg(estart + c);
    }
    for (u64 i0 = wlo; i0 < whi; i0 += 32) {
        const u64 i = i0 + lane;
        bool cand = false;
        u64 key = 0;
        if (i < whi) {
            if (cg.end <= (u32)i) {
                do cg = load_geo(geo, ++c); while (cg.end <= (u32)i);
                elen = __ldg(estart + c + 1) - __ldg(estart + c);
            }
            if ((u32)i - cg.start >= elen) {
                key = keys[i];
                const uint2 rec = d2oct[i];
                const u64 gid = key >> pb;
                u32 h = filter_hash(gid, T);
                cand = true;
                for (;;) {
                    const unsigned long long s = __ldg(T.slots + h);
                    if (s == 0ull) break;                        // prism not seen among the early arrivals
                    if ((s >> 24) == gid + 1ull) {
                        // cell.rs:80: only a strictly smaller d2 displaces; d2 >= the (rounded-up) early minimum: no record
                        cand = (rec.x >> 8) < (u32)(s & 0xFFFFFFull);
                        break;
                    }
                    h = (h + 1u) & T.mask;
                }
                if (!cand)
                    em[i] = ((u64)rec.y << 32) | (u32)i | (cg.end == (u32)i + 1u ? PCB_EM_CELL_LAST : 0ull);
            }
        }
        const u32 m = __ballot_sync(0xFFFFFFFFu, cand);
        if (cand) list[filled + __popc(m & ((1u << lane) - 1u))] = key;
        filled += (u32)__popc(m);
    }
    if (lane == 0) ncand[list_id] = filled;
}
// exclusive prefix of the candidate-list lengths (one block); *total = number of candidates
__global__ void __launch_bounds__(1024) k_filter_offsets(const u32* __restrict__ ncand, u32 nlists, u32* __restrict__ loff,
                                                         u32* __restrict__ total) {
    __shared__ u32 sh[1024];
    const u32 per = (nlists + 1023u) / 1024u;
    const u32 lo = min(nlists, threadIdx.x * per), hi = min(nlists, lo + per);
    u32 acc = 0;
    for (u32 c = lo; c < hi; c++) acc += ncand[c];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        u32 v = threadIdx.x >= (u32)d ? sh[threadIdx.x - d] : 0u;
        __syncthreads();
        sh[threadIdx.x] += v;
        __syncthreads();
    }
    u32 run = sh[threadIdx.x] - acc;
    for (u32 c = lo; c < hi; c++) {
        loff[c] = run;
        run += ncand[c];
    }
    if (threadIdx.x == 1023) *total = sh[1023];
}
// round-2 key list = early winners, then the candidate lists back to back (one block per list)
__global__ void __launch_bounds__(256)
k_filter_gather(const u64* __restrict__ cand_lists, u32 list_cap, const u32* __restrict__ ncand, const u32* __restrict__ loff,
                const u32* __restrict__ nwin_ptr, u64* __restrict__ out) {
    const u64* in = cand_lists + (u64)blockIdx.x * list_cap;
    const u32 cnt = ncand[blockIdx.x];
    u64* dst = out + *nwin_ptr + loff[blockIdx.x];
    for (u32 i = threadIdx.x; i < cnt; i += blockDim.x) dst[i] = in[i];
}
// digit histograms of all passes for a key list that no kernel of ours produced in order
__global__ void __launch_bounds__(256)
k_keys_hist(const u64* __restrict__ keys, u32 n, RadixPlan plan, u32* __restrict__ ghist) {
    __shared__ u32 hist[RADIX_MAX_PASSES * RADIX_MAX_BINS];
    for (int i = threadIdx.x; i < RADIX_MAX_PASSES * RADIX_MAX_BINS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 key = keys[i];
#pragma unroll 1
        for (int p = 0; p < plan.npass; p++)
            atomicAdd(&hist[p * RADIX_MAX_BINS + ((u32)(key >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u))], 1u);
    }
    __syncthreads();
    radix_hist_flush(hist, plan, ghist);
}
// early share of every cell (its first max(1, len >> shift) arrivals): estart[c] = first slot of cell c in the early
// key list, estart[ncell] = *total = number of early arrivals.  One block.
__global__ void __launch_bounds__(1024) k_early_starts(const u32* __restrict__ cell_start, u32 ncell, u32 shift,
                                                       u32* __restrict__ estart, u32* __restrict__ total) {
    __shared__ u32 sh[1024];
    const u32 per = (ncell + 1023u) / 1024u;
    const u32 lo = min(ncell, threadIdx.x * per), hi = min(ncell, lo + per);
    u32 acc = 0;
    for (u32 c = lo; c < hi; c++) acc += max(1u, (cell_start[c + 1] - cell_start[c]) >> shift);
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        u32 v = threadIdx.x >= (u32)d ? sh[threadIdx.x - d] : 0u;
        __syncthreads();
        sh[threadIdx.x] += v;
        __syncthreads();
    }
    u32 run = sh[threadIdx.x] - acc;
    for (u32 c = lo; c < hi; c++) {
        estart[c] = run;
        run += max(1u, (cell_start[c + 1] - cell_start[c]) >> shift);
    }
    if (threadIdx.x == 1023) {
        estart[ncell] = sh[1023];
        *total = sh[1023];
    }
}

// Stage 6: segmented stable 8-way split of the losers by child octant.  One scan over arrival positions,
// restarting at every cell, with eight running counters gives every loser its rank inside its
// (cell, octant) bucket.  The trigger batches at the ranks the flush rule looks at are captured on the fly.
#define PCB_EX_ABSENT 0
#define PCB_EX_SOME 1
#define PCB_EX_NONE 2
struct Oct8 {
    u32 c[8];
    u32 head;
};
struct OctScanOp {
    typedef Oct8 State;
    static constexpr bool kTileLoad = false;
    const u64* em;
    const u32* bat;
    const u32* cell_start;
    u32 ncell;
    u32 n;
    u32 L;
    const unsigned char* ex_state;  // merge mode only (else null)
    const u32* ex_m0;
    u32* rank;     // [n] rank of the loser emitted at this position inside its bucket (new losers only)
    u32* bcnt;     // [ncell * 8] new losers per bucket
    u32* special;  // [ncell * 8][4] trigger batch (+1) at ranks 0, L-1, L, and (merge) max(L-1-m0, 0)
    u32* err = nullptr;   // device error word (look-back timeout)
    __device__ static void pack(const State& s, u32 status, uint4& a, uint4& b) {   // counters < 2^30
        a = make_uint4(s.c[0], s.c[1], s.c[2], s.c[3] | status);
        b = make_uint4(s.c[4], s.c[5], s.c[6] | ((s.head & 1u) << 31), s.c[7] | status);
    }
    __device__ static State unpack(const uint4& a, const uint4& b) {
        State s;
        s.c[0] = a.x; s.c[1] = a.y; s.c[2] = a.z; s.c[3] = a.w & ~PCB_DESC_MASK;
        s.c[4] = b.x; s.c[5] = b.y; s.c[6] = b.z & 0x7FFFFFFFu; s.c[7] = b.w & ~PCB_DESC_MASK;
        s.head = b.z >> 31;
        return s;
    }
    __device__ State identity() const {
        State s;
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = 0;
        s.head = 0;
        return s;
    }
    // head word: bit 0 = a cell's first arrival lies in the range; bit 1 (single elements) = cell's last arrival
    __device__ State load(u32 i) const {
        State s;
        const u64 e = em[i];
        u32 o = (u32)(e >> 32) & 0xFu;
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = (o == (u32)k) ? 1u : 0u;
        s.head = ((e & PCB_EM_CELL_FIRST) ? 1u : 0u) | ((e & PCB_EM_CELL_LAST) ? 2u : 0u);
        return s;
    }
    __device__ State combine(const State& a, const State& b) const {
        State s;
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = (b.head & 1u) ? b.c[k] : a.c[k] + b.c[k];
        s.head = (a.head | b.head) & 1u;
        return s;
    }
    __device__ void consume(u32 i, const State& excl, const State& incl, const State& self) const {
        u32 o = 8;
#pragma unroll
        for (int k = 0; k < 8; k++) o = self.c[k] ? (u32)k : o;
        u32 c = PCB_NONE;
        if (o < 8) {
            u32 r = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) r = (o == (u32)k) ? excl.c[k] : r;
            if (self.head & 1u) r = 0;
            rank[i] = r;
            if (ex_state || r == 0 || r + 1 == L || r == L) {
                c = find_segment(cell_start, ncell, i);
                const u32 kk = 8 * c + o;
                const u32 b = bat[i];
                if (r == 0) special[4 * kk + 0] = b;
                if (r + 1 == L) special[4 * kk + 1] = b;
                if (r == L) special[4 * kk + 2] = b;
                if (ex_state && ex_state[kk] == PCB_EX_SOME) {
                    u32 m0 = ex_m0[kk];
                    u32 want = (L >= 1 + m0) ? L - 1 - m0 : 0;
                    if (r == want) special[4 * kk + 3] = b;
                }
            }
        }
        if (self.head & 2u) {   // last arrival of its cell: bucket totals
            if (c == PCB_NONE) c = find_segment(cell_start, ncell, i);
#pragma unroll
            for (int k = 0; k < 8; k++) bcnt[8 * c + k] = incl.c[k];
        }
    }
    __device__ void finish(const State&) const {}
};

// Specialised single-pass kernel for the scan above (tile = 256 threads x 16 consecutive arrivals each):
//  * a thread's sixteen arrivals are counted in eight 8-bit counters (two registers); the count of an arrival's octant
//    before it is its rank inside the thread.  The count runs twice: before the scan for the thread's totals, and again
//    for the ranks inside the tile:  local rank + (the thread's start value of its octant, unless a cell started inside
//    the thread before it)  - one read-only shared-memory load per arrival;
//  * inside the warp and the block the counters travel as four registers of two 16-bit fields (a tile has 4096
//    arrivals); bit 31 of the fourth is the "a cell starts in the range" flag, which the segmented combine
//    (b.head ? b : a + b) carries along for free.  Only the cross-tile prefix uses the 9-word state of OctScanOp;
//  * a thread owns CONSECUTIVE arrivals (the scan wants that), so loading them directly would make every warp load
//    touch 32 different 128-byte lines, and every rank store 16: the warp loads its 512 records with coalesced 16-byte
//    loads instead, parks the 16-bit codes in shared memory (padded: conflict-free both ways), every thread picks its
//    sixteen codes up with two 16-byte shared loads, and the ranks take the same way back.
//  * the ranks INSIDE the tile do not depend on the tiles before it, so they are computed while warp 0 waits in the
//    look-back (warp 0: publish the aggregate, own ranks, then poll); they are parked in the rank-transpose buffer as
//    tile-local rank | octant << 16 | "no prefix" << 20  and the tile prefix of the octant is added on the way out, in
//    the coalesced layout: behind the look-back there is one shared load, one add and the trigger test per arrival.
// Round 2, measured on the B200 per C2 step: 3.00 ms (64-bit packed counters, a shared-memory load / store chain per
// arrival, direct loads: 96 instructions per arrival, l1tex 68 %) -> 3.44 (32-bit counters, direct loads narrowed to
// sixteen 4-byte loads per thread: L1-bound) -> 2.74 (transposed loads / stores) -> 2.61 ms (codes packed two per
// register: 40 registers, six blocks per SM; seven: 2.63, five: 2.76, four: 3.08) -> 2.55 ms (in-tile ranks computed
// during the look-back).
#define OCT_TILE_MAX 4096   // 16-bit fields hold a whole tile
struct OctW4 {
    u32 w[4];   // octants (0,1) (2,3) (4,5) (6,7), 16 bits each; bit 31 of w[3]: head
};
#define OCT_HEAD_BIT 0x80000000u
__device__ __forceinline__ OctW4 octw_combine(const OctW4& a, const OctW4& b) {
    const bool cut = (b.w[3] & OCT_HEAD_BIT) != 0u;
    OctW4 r;
#pragma unroll
    for (int q = 0; q < 4; q++) r.w[q] = cut ? b.w[q] : a.w[q] + b.w[q];
    return r;
}
__device__ __forceinline__ u32 octw_field(const OctW4& a, int oct) { return (a.w[oct >> 1] >> (16 * (oct & 1))) & 0x1FFFu; }
// one arrival into the byte counters (lo: octants 0-3, hi: 4-7); returns the count of its octant before it
__device__ __forceinline__ u32 oct_bytes_step(u32 code, u32& lo, u32& hi, u32& started) {
    if (code & 0x200u) { lo = 0u; hi = 0u; started = 1u; }   // PCB_EM_CELL_FIRST >> 32
    const u32 o = code & 0xFu, sel = o & 3u;
    const u32 r = __byte_perm((o & 4u) ? hi : lo, 0u, 0x4440u | sel);
    const u32 one = 1u << (8u * sel);
    lo += (o < 4u) ? one : 0u;
    hi += (o & 4u) ? one : 0u;
    return r;   // meaningless for octant 8 (no emission)
}
#define OCT_CODE(k) (((k) & 1) ? cw[(k) >> 1] >> 16 : cw[(k) >> 1] & 0xFFFFu)   // two 16-bit codes per register
struct OctCw8 {
    u32 w[8];
};
// bucket totals of the cell whose last arrival is item kl of this thread (rare: once per cell)
__device__ __noinline__ void oct_cell_totals(OctCw8 cw, int kl, u32 cell, bool tcut, const u32* __restrict__ col /* s_run column, stride */,
                                             int stride, const Oct8* __restrict__ prefix, u32* __restrict__ bcnt) {
    u32 lo = 0u, hi = 0u, started = 0u;
    for (int j = 0; j <= kl; j++) {
        const u32 c16 = (j & 1) ? cw.w[j >> 1] >> 16 : cw.w[j >> 1] & 0xFFFFu;
        oct_bytes_step(c16, lo, hi, started);
    }
    for (int q = 0; q < 8; q++) {
        u32 v = ((q < 4 ? lo : hi) >> (8 * (q & 3))) & 0xFFu;
        if (!started) v += col[q * stride] + (tcut ? 0u : prefix->c[q]);
        bcnt[8 * cell + q] = v;
    }
}
template <int OCT_THREADS, int MINB>
__global__ void __launch_bounds__(OCT_THREADS, MINB)
k_octscan(OctScanOp op, u32 n, u32 ntiles, uint4* __restrict__ desc) {
    constexpr int OCT_ITEMS = 16;
    constexpr int WARPS = OCT_THREADS / 32;
    constexpr int OCT_TILE = OCT_THREADS * OCT_ITEMS;
    static_assert(OCT_TILE <= OCT_TILE_MAX, "16-bit fields hold a whole tile");
    __shared__ uint4 s_warp[WARPS];
    __shared__ Oct8 s_prefix;
    __shared__ __align__(16) u32 s_run[9][OCT_THREADS];
    __shared__ __align__(16) u32 s_out[WARPS * 576];
    static_assert(9 * OCT_THREADS >= WARPS * 288, "code transpose fits the counter columns");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 tile = blockIdx.x;   // see k_chain_scan
    const u64 base = (u64)tile * OCT_TILE + (u64)tid * OCT_ITEMS;
    const u64 wbase = (u64)tile * OCT_TILE + (u64)warp * (32 * OCT_ITEMS);
    u32 cw[OCT_ITEMS / 2];
    {
        u32* wcode = &s_run[0][0] + warp * 288;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const u64 e0 = wbase + (u64)(j * 64 + 2 * lane);
            u32 c0 = PCB_EM_NONE_OCT, c1 = PCB_EM_NONE_OCT;
            if (e0 + 1 < n) {
                const ulonglong2 e = *(const ulonglong2*)(op.em + e0);
                c0 = (u32)(e.x >> 32);
                c1 = (u32)(e.y >> 32);
            } else if (e0 < n) {
                c0 = (u32)(op.em[e0] >> 32);
            }
            wcode[j * 36 + lane] = (c0 & 0xFFFFu) | (c1 << 16);
        }
        __syncwarp();
        const uint4 a = *(const uint4*)(wcode + lane * 8 + (lane >> 2) * 4), b = *(const uint4*)(wcode + lane * 8 + (lane >> 2) * 4 + 4);
        cw[0] = a.x; cw[1] = a.y; cw[2] = a.z; cw[3] = a.w;
        cw[4] = b.x; cw[5] = b.y; cw[6] = b.z; cw[7] = b.w;
    }
    u32 lo = 0u, hi = 0u, started = 0u;
#pragma unroll
    for (int k = 0; k < OCT_ITEMS; k++) oct_bytes_step(OCT_CODE(k), lo, hi, started);
    OctW4 winc;
    winc.w[0] = __byte_perm(lo, 0u, 0x4140u);
    winc.w[1] = __byte_perm(lo, 0u, 0x4342u);
    winc.w[2] = __byte_perm(hi, 0u, 0x4140u);
    winc.w[3] = __byte_perm(hi, 0u, 0x4342u) | (started ? OCT_HEAD_BIT : 0u);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        OctW4 up;
#pragma unroll
        for (int q = 0; q < 4; q++) up.w[q] = __shfl_up_sync(0xFFFFFFFFu, winc.w[q], d);
        if (lane >= d) winc = octw_combine(up, winc);
    }
    if (lane == 31) s_warp[warp] = make_uint4(winc.w[0], winc.w[1], winc.w[2], winc.w[3]);
    __syncthreads();   // (also: every warp has picked its codes up, s_run may be reused)
    OctW4 wpre{{0u, 0u, 0u, 0u}}, total{{0u, 0u, 0u, 0u}};
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
        const uint4 v = s_warp[w];
        const OctW4 sw{{v.x, v.y, v.z, v.w}};
        if (w < warp) wpre = octw_combine(wpre, sw);
        total = octw_combine(total, sw);
    }
    OctW4 texcl;
#pragma unroll
    for (int q = 0; q < 4; q++) texcl.w[q] = __shfl_up_sync(0xFFFFFFFFu, winc.w[q], 1);
    if (lane == 0) texcl = OctW4{{0u, 0u, 0u, 0u}};
    texcl = octw_combine(wpre, texcl);
    Oct8 tot8;
    if (warp == 0) {
#pragma unroll
        for (int q = 0; q < 8; q++) tot8.c[q] = octw_field(total, q);
        tot8.head = total.w[3] >> 31;
        if (lane == 0) chain_publish<OctScanOp>(desc, tile, tot8, tile == 0 ? PCB_DESC_INCL : PCB_DESC_AGG);
    }
    // ranks inside the tile
    const bool tcut = (texcl.w[3] & OCT_HEAD_BIT) != 0u;
#pragma unroll
    for (int q = 0; q < 8; q++) s_run[q][tid] = octw_field(texcl, q);
    s_run[8][tid] = 0u;
    u32 lastmask = 0u;
    {
        u32 pk[4];
        lo = 0u; hi = 0u; started = 0u;
#pragma unroll
        for (int k = 0; k < OCT_ITEMS; k++) {
            const u32 code = OCT_CODE(k);
            const u32 o = code & 0xFu;
            const u32 start = s_run[o][tid];
            const u32 local = oct_bytes_step(code, lo, hi, started);
            const u32 r = started ? local : start + local;
            const u32 nofix = (started != 0u || tcut || o >= 8u) ? 1u : 0u;
            pk[k & 3] = (o < 8u ? r : 0u) | (o << 16) | (nofix << 20);
            lastmask |= ((code >> 8) & 1u) << k;
            if ((k & 3) == 3)
                *(uint4*)(s_out + warp * 576 + lane * 16 + (k - 3) + (lane >> 1) * 4) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
    }
    if (warp == 0) {
        Oct8 prefix = chain_lookback(op, desc, tile, tot8, op.err, false);
        if (lane == 0) s_prefix = prefix;
    }
    __syncthreads();
    // on the way out (coalesced layout): + prefix of the octant, trigger batches, store
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int i = j * 128 + lane * 4;
        const uint4 v = *(const uint4*)(s_out + warp * 576 + i + (i >> 5) * 4);
        const u32 w4[4] = {v.x, v.y, v.z, v.w};
        const u64 e0 = wbase + (u64)i;
        u32 rk[4];
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const u32 o = (w4[t] >> 16) & 0xFu;
            const u32 pre = s_prefix.c[o & 7u];
            const u32 r = (w4[t] & 0xFFFFu) + (((w4[t] >> 20) & 1u) ? 0u : pre);
            rk[t] = r;
            if (o < 8u && (op.ex_state || r == 0u || r - (op.L - 1u) <= 1u)) {   // r == 0, L - 1 or L  (L >= 1)
                const u32 ii = (u32)(e0 + t);
                const u32 c = find_segment(op.cell_start, op.ncell, ii);
                const u32 kk = 8 * c + o;
                const u32 b = op.bat[ii];
                if (r == 0) op.special[4 * kk + 0] = b;
                if (r + 1 == op.L) op.special[4 * kk + 1] = b;
                if (r == op.L) op.special[4 * kk + 2] = b;
                if (op.ex_state && op.ex_state[kk] == PCB_EX_SOME) {
                    u32 m0 = op.ex_m0[kk];
                    u32 want = (op.L >= 1 + m0) ? op.L - 1 - m0 : 0;
                    if (r == want) op.special[4 * kk + 3] = b;
                }
            }
        }
        if (e0 + 4 <= n) {
            *(uint4*)(op.rank + e0) = make_uint4(rk[0], rk[1], rk[2], rk[3]);
        } else {
            if (e0 < n) op.rank[e0] = rk[0];
            if (e0 + 1 < n) op.rank[e0 + 1] = rk[1];
            if (e0 + 2 < n) op.rank[e0 + 2] = rk[2];
        }
    }
    while (lastmask) {   // last arrival of a cell (once per cell): bucket totals
        const int kl = __ffs(lastmask) - 1;
        lastmask &= lastmask - 1u;
        OctCw8 c8;
#pragma unroll
        for (int m = 0; m < 8; m++) c8.w[m] = cw[m];
        oct_cell_totals(c8, kl, find_segment(op.cell_start, op.ncell, (u32)(base + kl)), tcut, &s_run[0][tid], OCT_THREADS, &s_prefix, op.bcnt);
    }
}
#undef OCT_CODE
template <int T, int MINB>
static inline int octscan_launch(const OctScanOp& op, u32 n, void* temp, cudaStream_t st) {
    const u32 ntiles = div_up(n, T * 16);
    if (!ntiles) return 0;
    cudaMemsetAsync(temp, 0, chain_temp_bytes(ntiles), st);
    k_octscan<T, MINB><<<ntiles, T, 0, st>>>(op, n, ntiles, (uint4*)temp);
    return 1;
}
static inline int octscan(const OctScanOp& op, u32 n, void* temp, cudaStream_t st) {
    static const int variant = pcb_tuning_env("PCB_OCT_VARIANT") ? atoi(pcb_tuning_env("PCB_OCT_VARIANT")) : 0;   // tuning hook
    switch (variant) {
        case 1: return octscan_launch<256, 5>(op, n, temp, st);
        case 2: return octscan_launch<256, 7>(op, n, temp, st);
        default: return octscan_launch<256, 6>(op, n, temp, st);   // 40 registers, six blocks per SM
    }
}

// ---------------------------------------------------------------------------------------------
// Stage 7: Cell::add_points_in_overflow (cell.rs:108-153) for every (cell, child) bucket, closed form.
// lb0 / lbLm1 / lbL / lbK = trigger batches (+1) of the bucket's NEW losers at emission ranks 0, L-1, L and
// max(L-1-m0, 0); `m0`/`state` = what the node file held before this run (state 0 absent, 1 Some(list of
// m0), 2 None).  Returns true when the list is delivered to the child; *bf = batch (+1) of the call that
// delivers the buffered list (0: every call delivers its own losers, state None).
//                                                                        tests/bulk_model.py::flush_rule
__device__ __forceinline__ bool flush_rule(u32 n, u32 m0, int state, u32 L, u32 lb0, u32 lbLm1, u32 lbL, u32 lbK,
                                           u32* bf) {
    *bf = 0;
    if (state == PCB_EX_NONE) return true;                     // cell.rs:128-130
    if (state == PCB_EX_SOME) {                                // cell.rs:131-147: flush when len >= L
        if ((u64)m0 + n < L) return false;
        *bf = lbK;
        return true;
    }
    if (L == 0) { *bf = lb0; return true; }                    // cell.rs:118: len <= 0 never holds
    if (n < L) return false;                                   // first call keeps when len <= L
    u32 bstar = lbLm1;
    if (bstar == lb0 && (n == L || lbL != bstar)) {            // first call brought exactly L: kept (:118),
        if (n == L) return false;                              // flushed by the next call (:137)
        *bf = lbL;
        return true;
    }
    *bf = bstar;
    return true;
}

#define PCB_BF_ENTRY 1u
#define PCB_BF_FLUSH 2u

// where a bucket's losers go
struct __align__(16) BucketDst {
    u64 base;    // FLUSH: first slot in the next level's arrivals; else byte offset of the list in the image
    u32 bfl;     // batch (+1) of the flushing call
    u32 m0fl;    // m0 << 2 | flags
};

__global__ void k_bucket_decide(const u32* __restrict__ bcnt, const u32* __restrict__ special,
                                const u32* __restrict__ win_end, u32 ncell, u32 L,
                                const unsigned char* __restrict__ ex_state, const u32* __restrict__ ex_m0,
                                BucketDst* __restrict__ bdst, u32* __restrict__ cell_deliver,
                                u32* __restrict__ cell_children, u64* __restrict__ cell_bytes) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    u32 ng = win_end[c] - (c ? win_end[c - 1] : 0u);
    u64 bytes = 48ull + 16ull * ng + 1ull;
    u32 deliver = 0, children = 0;
    for (int o = 0; o < 8; o++) {
        u32 k = 8 * c + o;
        u32 n = bcnt[k];
        int st = ex_state ? ex_state[k] : PCB_EX_ABSENT;
        u32 m0 = (st == PCB_EX_SOME) ? ex_m0[k] : 0u;
        u32 fl = 0, bf = 0, cnt = m0 + n;
        if (n == 0) {
            if (st != PCB_EX_ABSENT) { fl = PCB_BF_ENTRY; bytes += 16ull + 16ull * m0; }
        } else {
            fl = PCB_BF_ENTRY;
            if (flush_rule(n, m0, st, L, special[4 * k], special[4 * k + 1], special[4 * k + 2], special[4 * k + 3], &bf)) {
                fl |= PCB_BF_FLUSH;
                deliver += cnt;
                children += 1;
                bytes += 16ull;
            } else {
                bytes += 16ull + 16ull * cnt;
            }
        }
        BucketDst d;
        d.base = cnt;   // provisional: k_cell_emit turns the count into the destination
        d.bfl = bf;
        d.m0fl = (m0 << 2) | fl;
        bdst[k] = d;
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
    static constexpr bool kTileLoad = false;
    const u32* cell_deliver;
    const u32* cell_children;
    const u64* cell_bytes;
    u32* cell_dbase;
    u32* cell_cbase;
    u64* node_off;
    LevelTotals* totals;
    __device__ static void pack(const State& s, u32 status, uint4& a, uint4& b) {
        a = make_uint4((u32)s.bytes, (u32)(s.bytes >> 32), s.deliver, status);
        b = make_uint4(s.children, 0u, 0u, status);
    }
    __device__ static State unpack(const uint4& a, const uint4& b) {
        return State{((u64)a.y << 32) | a.x, a.z, b.x};
    }
    __device__ State identity() const { return State{0ull, 0u, 0u}; }
    __device__ State load(u32 i) const {
        return State{(cell_bytes[i] + 15ull) & ~15ull, cell_deliver[i], cell_children[i]};
    }
    __device__ State combine(const State& a, const State& b) const {
        return State{a.bytes + b.bytes, a.deliver + b.deliver, a.children + b.children};
    }
    __device__ void consume(u32 i, const State& excl, const State&, const State&) const {
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
// (cell.rs:162-177), plus the bucket -> destination table used by the point movers below.
__global__ void k_cell_emit(const int* __restrict__ cell_idx, u32 ncell, u32 hierarchy, LevelGeom g,
                            const u32* __restrict__ win_end, BucketDst* __restrict__ bdst,
                            const u32* __restrict__ cell_dbase, const u32* __restrict__ cell_cbase,
                            const u64* __restrict__ node_off, const u64* __restrict__ cell_bytes,
                            unsigned char* __restrict__ image, u32* __restrict__ next_cell_start,
                            int* __restrict__ next_cell_idx, u32 n_next, u32 ncell_next) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    const int ix = cell_idx[3 * c], iy = cell_idx[3 * c + 1], iz = cell_idx[3 * c + 2];
    u32 ng = win_end[c] - (c ? win_end[c - 1] : 0u);
    u32 nov = 0, nent = 0;
    for (int o = 0; o < 8; o++) {
        BucketDst d = bdst[8 * c + o];
        if (d.m0fl & PCB_BF_ENTRY) {
            nent++;
            if (!(d.m0fl & PCB_BF_FLUSH)) nov += (u32)d.base;
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
        BucketDst d = bdst[k];
        if (!(d.m0fl & PCB_BF_ENTRY)) continue;
        const u32 cnt = (u32)d.base;
        int child[3] = {2 * ix + ((o >> 2) & 1), 2 * iy + ((o >> 1) & 1), 2 * iz + (o & 1)};
        u32 kept = (d.m0fl & PCB_BF_FLUSH) ? 0u : cnt;
        store_bytes(node + eoff, child, 12);
        store_bytes(node + eoff + 12, &kept, 4);
        if (d.m0fl & PCB_BF_FLUSH) {
            d.base = dbase;
            next_cell_start[cb] = dbase;
            next_cell_idx[3 * cb + 0] = child[0];
            next_cell_idx[3 * cb + 1] = child[1];
            next_cell_idx[3 * cb + 2] = child[2];
            dbase += cnt;
            cb++;
        } else {
            d.base = node_off[c] + eoff + 16;
        }
        bdst[k] = d;
        eoff += 16ull + 16ull * kept;
    }
    // zero the alignment padding after the node
    for (u64 b = eoff; b < ((cell_bytes[c] + 15ull) & ~15ull); b++) node[b] = 0;
    if (c == ncell - 1) next_cell_start[ncell_next] = n_next;
}

// Stage 8b: move every loser, in arrival (= emission) order: delivered to its child cell (next level's
// arrivals, in (child, tau') order, with batch max(b_e, b_flush) — SURVEY N3) or kept in the parent's
// overflow list (image).
#define LOSER_ITEMS 4   // arrivals per thread: independent load chains (em -> bucket record -> point) in flight together
__global__ void __launch_bounds__(256)
k_loser_emit(const u64* __restrict__ em, const u32* __restrict__ rank, const u32* __restrict__ bat, u32 n,
             const u32* __restrict__ cell_start, u32 ncell, const PcbPoint* __restrict__ pts,
             const BucketDst* __restrict__ bdst, PcbPoint* __restrict__ next_pts, u32* __restrict__ next_bat,
             unsigned char* __restrict__ image) {
    const u64 b0 = (u64)blockIdx.x * (256 * LOSER_ITEMS);
    u64 e[LOSER_ITEMS];
    u32 rk[LOSER_ITEMS], bt[LOSER_ITEMS];
#pragma unroll
    for (int k = 0; k < LOSER_ITEMS; k++) {
        const u64 t = b0 + 256 * k + threadIdx.x;
        e[k] = (t < n) ? em[t] : ((u64)PCB_EM_NONE_OCT << 32);
        rk[k] = (t < n) ? rank[t] : 0u;
        bt[k] = (t < n) ? bat[t] : 0u;
    }
    BucketDst d[LOSER_ITEMS];
    uint4 p[LOSER_ITEMS];
    // cell of the arrival: one binary search per warp (its first arrival), then a walk (arrivals ascend with k)
    u32 c = (b0 + (threadIdx.x & ~31u) < n) ? find_segment(cell_start, ncell, (u32)(b0 + (threadIdx.x & ~31u))) : 0u;
#pragma unroll
    for (int k = 0; k < LOSER_ITEMS; k++) {
        const u64 t = b0 + 256 * k + threadIdx.x;
        const u32 o = (u32)(e[k] >> 32) & 0xFu;
        if (t < n)
            while (__ldg(cell_start + c + 1) <= (u32)t) c++;
        if (o < 8) {
            d[k] = bdst[8 * c + o];
            p[k] = *(const uint4*)(pts + (u32)e[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < LOSER_ITEMS; k++) {
        const u32 o = (u32)(e[k] >> 32) & 0xFu;
        if (o < 8) {
            const u32 r = (d[k].m0fl >> 2) + rk[k];
            if (d[k].m0fl & PCB_BF_FLUSH) {
                const u64 dst = d[k].base + r;
                *(uint4*)(next_pts + dst) = p[k];
                next_bat[dst] = max(bt[k], d[k].bfl);
            } else {
                store16_at1(image + d[k].base + 16ull * r, p[k]);
            }
        }
    }
}

// Stage 8c: grid section of every node: the winners, already in (cell, Z, Y, X) order.
#define WINNER_ITEMS 4
__global__ void __launch_bounds__(256)
k_winner_emit(const u32* __restrict__ win_src, u32 nwin, const u32* __restrict__ win_end, u32 ncell,
              const PcbPoint* __restrict__ pts, const u64* __restrict__ node_off, unsigned char* __restrict__ image) {
    const u64 b0 = (u64)blockIdx.x * (256 * WINNER_ITEMS);
    u32 src[WINNER_ITEMS];
#pragma unroll
    for (int k = 0; k < WINNER_ITEMS; k++) {
        const u64 w = b0 + 256 * k + threadIdx.x;
        src[k] = (w < nwin) ? win_src[w] : 0u;
    }
    uint4 p[WINNER_ITEMS];
#pragma unroll
    for (int k = 0; k < WINNER_ITEMS; k++)
        if (b0 + 256 * k + threadIdx.x < nwin) p[k] = *(const uint4*)(pts + src[k]);
    // cell = first c with win_end[c] > w: one binary search per warp (its first winner), then a walk (w ascends with k)
    const u64 wf = b0 + (threadIdx.x & ~31u);
    u32 c = 0;
    if (wf < nwin) {
        u32 lo = 0, hi = ncell;
        while (lo < hi) {
            u32 mid = (lo + hi) >> 1;
            if (__ldg(win_end + mid) > (u32)wf) hi = mid; else lo = mid + 1;
        }
        c = lo;
    }
#pragma unroll
    for (int k = 0; k < WINNER_ITEMS; k++) {
        const u64 w = b0 + 256 * k + threadIdx.x;
        if (w < nwin) {
            while (__ldg(win_end + c) <= (u32)w) c++;
            const u32 wbeg = c ? __ldg(win_end + c - 1) : 0u;
            *(uint4*)(image + node_off[c] + 48ull + 16ull * ((u32)w - wbeg)) = p[k];
        }
    }
}

// merge mode: the overflow lists that were already in the node files (front of each bucket)
__global__ void k_exlist_emit(const PcbPoint* __restrict__ ex_pts, u32 nex, const u32* __restrict__ ex_bucket,
                              const u32* __restrict__ ex_rank, const BucketDst* __restrict__ bdst,
                              PcbPoint* __restrict__ next_pts, u32* __restrict__ next_bat,
                              unsigned char* __restrict__ image) {
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nex) return;
    const BucketDst d = bdst[ex_bucket[j]];
    const u32 r = ex_rank[j];
    uint4 p = *(const uint4*)(ex_pts + j);
    if (d.m0fl & PCB_BF_FLUSH) {
        *(uint4*)(next_pts + d.base + r) = p;
        next_bat[d.base + r] = d.bfl;  // max(0, b_flush)
    } else {
        store16_at1(image + d.base + 16ull * r, p);
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
// CSR of one wave of level-0 cells: the global starts rebased to the wave's first arrival
__global__ void k_wave_starts(const u32* __restrict__ cell_start, u32 first, u32 count, u32* __restrict__ out) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < count) out[c] = cell_start[c] - first;
}
__global__ void k_merge_starts(u32* __restrict__ cell_start, const u32* __restrict__ inc_start, u32 ncell) {
    u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c <= ncell) cell_start[c] += inc_start[c];
}

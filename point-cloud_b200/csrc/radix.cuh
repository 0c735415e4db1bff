// radix.cuh — stable LSD radix sort of (u64 key, u64 value) pairs, hand-written for sm_100a.
//
// Replaces the reference's FxHashMap grouping (point-converter/src/converter.rs:32-47 group_points;
// cell.rs:70-94 hash probe per point) by a sort: equal keys become contiguous and, because every pass
// is stable, stay in arrival order.
//
// Pass structure (this file, v1 — no inter-block waiting anywhere):
//   k_radix_upsweep   per-tile 256-bin digit histogram                      reads 8 B/element
//   device_scan       exclusive sum over the digit-major [256][ntiles] table
//   k_radix_downsweep re-reads the tile, ranks keys with warp match/ballot, stages the tile in shared
//                     memory in digit order and writes runs of equal digits coalesced
//                                                                          reads 16 B, writes 16 B/element
// Only the 8-bit windows of the key in which some pair of keys differs are sorted (`diffmask`).
#pragma once
#include "common.cuh"
#include "scan.cuh"

#define RADIX_THREADS 256
#define RADIX_ITEMS 16
#define RADIX_TILE (RADIX_THREADS * RADIX_ITEMS)
#define RADIX_WARPS (RADIX_THREADS / 32)

__device__ __forceinline__ u32 radix_digit(u64 key, int shift, u32 mask) { return (u32)(key >> shift) & mask; }

__global__ void __launch_bounds__(RADIX_THREADS)
k_radix_upsweep(const u64* __restrict__ keys, u32 n, int shift, u32 mask, u32 ntiles, u32* __restrict__ tile_hist) {
    __shared__ u32 hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const u64 base = (u64)blockIdx.x * RADIX_TILE;
#pragma unroll 4
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u64 i = base + (u64)k * RADIX_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[radix_digit(keys[i], shift, mask)], 1u);
    }
    __syncthreads();
    tile_hist[(u64)threadIdx.x * ntiles + blockIdx.x] = hist[threadIdx.x];
}

// dynamic shared memory layout: u64 skeys[TILE]; u64 svals[TILE]; u32 cnt[WARPS][256]; u32 dig_excl[256]; u32 gbase[256]
#define RADIX_SMEM_BYTES (RADIX_TILE * 16 + RADIX_WARPS * 256 * 4 + 256 * 4 + 256 * 4 + 64)

__global__ void __launch_bounds__(RADIX_THREADS)
k_radix_downsweep(const u64* __restrict__ keys_in, const u64* __restrict__ vals_in, u64* __restrict__ keys_out,
                  u64* __restrict__ vals_out, u32 n, int shift, u32 mask, u32 ntiles,
                  const u32* __restrict__ tile_base /* scanned [256][ntiles] */) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* skeys = (u64*)smem_raw;
    u64* svals = skeys + RADIX_TILE;
    u32* cnt = (u32*)(svals + RADIX_TILE);  // [WARPS][256]
    u32* dig_excl = cnt + RADIX_WARPS * 256;
    u32* gbase = dig_excl + 256;
    __shared__ u32 warp_sums[RADIX_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u64 tbase = (u64)blockIdx.x * RADIX_TILE;
    const u32 tile_n = (u32)min((u64)RADIX_TILE, (u64)n - tbase);
    for (int i = tid; i < RADIX_WARPS * 256; i += RADIX_THREADS) cnt[i] = 0;
    __syncthreads();

    // warp-striped ownership: item k of lane l in warp w is element w*32*ITEMS + k*32 + l (order: k, then lane)
    u64 key[RADIX_ITEMS];
    u32 rank[RADIX_ITEMS];
    const u32 wbase = warp * 32 * RADIX_ITEMS + lane;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        key[k] = (j < tile_n) ? keys_in[tbase + j] : ~0ull;
    }
    u32* wcnt = cnt + warp * 256;
    const u32 lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        bool valid = j < tile_n;
        u32 d = valid ? radix_digit(key[k], shift, mask) : 256u;
        u32 peers = __match_any_sync(0xFFFFFFFFu, d);
        u32 lower = __popc(peers & lt);
        u32 before = valid ? wcnt[d] : 0u;
        __syncwarp();
        if (valid && lower == 0) wcnt[d] = before + __popc(peers);
        __syncwarp();
        rank[k] = before + lower;
    }
    __syncthreads();
    // per digit: exclusive prefix over warps, tile count
    u32 running = 0;
#pragma unroll
    for (int w = 0; w < RADIX_WARPS; w++) {
        u32 c = cnt[w * 256 + tid];
        cnt[w * 256 + tid] = running;
        running += c;
    }
    // exclusive scan of the 256 tile counts (one per thread)
    u32 incl = running;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        u32 v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    u32 woff = 0;
#pragma unroll
    for (int w = 0; w < RADIX_WARPS; w++)
        if (w < warp) woff += warp_sums[w];
    u32 excl = woff + incl - running;
    dig_excl[tid] = excl;
    gbase[tid] = tile_base[(u64)tid * ntiles + blockIdx.x] - excl;
    __syncthreads();

#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        if (j < tile_n) {
            u32 d = radix_digit(key[k], shift, mask);
            u32 p = dig_excl[d] + wcnt[d] + rank[k];
            rank[k] = p;
            skeys[p] = key[k];
            svals[p] = vals_in[tbase + j];
        }
    }
    __syncthreads();
    for (u32 j = tid; j < tile_n; j += RADIX_THREADS) {
        u64 kx = skeys[j];
        u32 g = gbase[radix_digit(kx, shift, mask)] + j;
        keys_out[g] = kx;
        vals_out[g] = svals[j];
    }
}

struct RadixTemp {
    u32* tile_hist = nullptr;   // [256 * ntiles + 1]
    u32* scan_agg = nullptr;    // device_scan temp
    u64 cap_elems = 0;
    static u64 hist_words(u64 n) { return 256ull * ((n + RADIX_TILE - 1) / RADIX_TILE) + 1; }
};

// Sorts by the key bits set in `diffmask`.  Ping-pongs between (k0,v0) and (k1,v1); returns 0 or 1 = which
// pair holds the sorted data.  `prof` brackets every kernel class: prof.pb(); prof.pe(name, launches, units, bytes).
template <class Prof>
static inline int radix_sort_pairs(u64* k0, u64* v0, u64* k1, u64* v1, u32 n, u64 diffmask, RadixTemp& tmp,
                                   cudaStream_t st, Prof& prof) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k_radix_downsweep, cudaFuncAttributeMaxDynamicSharedMemorySize, RADIX_SMEM_BYTES);
        attr_set = true;
    }
    int cur = 0;
    if (n < 2) diffmask = 0;
    const u32 ntiles = div_up(n, RADIX_TILE);
    int shift = 0;
    while (shift < 64 && (diffmask >> shift) != 0) {
        shift += __builtin_ctzll(diffmask >> shift);
        int bits = 64 - shift < 8 ? 64 - shift : 8;
        u32 mask = (1u << bits) - 1u;
        u64 *ki = cur ? k1 : k0, *vi = cur ? v1 : v0, *ko = cur ? k0 : k1, *vo = cur ? v0 : v1;
        prof.pb();
        k_radix_upsweep<<<ntiles, RADIX_THREADS, 0, st>>>(ki, n, shift, mask, ntiles, tmp.tile_hist);
        prof.pe("radix_upsweep", 1, n, 8ull * n);
        SumOp<u32, u32> op{tmp.tile_hist, tmp.tile_hist, nullptr};
        prof.pb();
        int nl = device_scan(op, 256u * ntiles, tmp.scan_agg, st);
        prof.pe("radix_scan", nl, 256ull * ntiles, 8ull * 256 * ntiles);
        prof.pb();
        k_radix_downsweep<<<ntiles, RADIX_THREADS, RADIX_SMEM_BYTES, st>>>(ki, vi, ko, vo, n, shift, mask, ntiles,
                                                                        tmp.tile_hist);
        prof.pe("radix_downsweep", 1, n, 32ull * n);
        cur ^= 1;
        shift += 8;
    }
    return cur;
}

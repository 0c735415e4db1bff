// radix.cuh — stable LSD radix sort of u64 keys (optionally with a u32 payload), hand-written for sm_100a.
//
// Replaces the reference's FxHashMap grouping (point-converter/src/converter.rs:32-47 group_points;
// cell.rs:70-94 hash probe per point) by a sort: equal keys become contiguous and, because every pass
// is stable, stay in arrival order.
//
// "Onesweep": the digit histograms of ALL passes are accumulated by the kernel that produces the keys
// (radix_hist_add / radix_hist_flush below), so a pass is ONE kernel that reads every key once and writes it
// once (16 B per key and pass; 24 B with a payload).  Inside a pass a tile
//   1. takes its ordinal from an atomic counter (forward progress for the look-back),
//   2. ranks its keys per warp with __match_any_sync on the digit,
//   3. publishes its per-digit counts, folds the published counts of the tiles before it (decoupled
//      look-back, one thread per digit, flag bits inside the status word),
//   4. stages the tile in shared memory in digit order and writes runs of equal digits coalesced.
// Digits are up to 9 bits wide (512 bins); the caller lays the sort bits out so that the passes are even.
#pragma once
#include "common.cuh"

#define RADIX_THREADS 256
#define RADIX_ITEMS 16
#define RADIX_TILE (RADIX_THREADS * RADIX_ITEMS)
#define RADIX_WARPS (RADIX_THREADS / 32)
#define RADIX_MAX_BITS 9
#define RADIX_MAX_BINS (1 << RADIX_MAX_BITS)
#define RADIX_MAX_PASSES 8
#define RADIX_FLAG_AGG 0x40000000u
#define RADIX_FLAG_INCL 0x80000000u
#define RADIX_COUNT_MASK 0x3FFFFFFFu

struct RadixPlan {
    int npass;
    int shift[RADIX_MAX_PASSES];
    int bits[RADIX_MAX_PASSES];
};

// Sort bits [lo, lo + nbits) in ceil(nbits / 9) passes of (almost) equal width.
static inline RadixPlan radix_plan(int lo, int nbits) {
    RadixPlan p;
    p.npass = (nbits + RADIX_MAX_BITS - 1) / RADIX_MAX_BITS;
    int done = 0;
    for (int i = 0; i < p.npass; i++) {
        int b = (nbits - done + (p.npass - i) - 1) / (p.npass - i);
        p.shift[i] = lo + done;
        p.bits[i] = b;
        done += b;
    }
    for (int i = p.npass; i < RADIX_MAX_PASSES; i++) p.shift[i] = p.bits[i] = 0;
    return p;
}

// ---- histogram side (called from the key-producing kernels) ----------------------------------------------
// shared: u32 hist[npass * RADIX_MAX_BINS], zeroed by the caller
__device__ __forceinline__ void radix_hist_add(u32* __restrict__ hist, const RadixPlan& plan, u64 key) {
#pragma unroll
    for (int p = 0; p < RADIX_MAX_PASSES; p++) {
        if (p + 1 < plan.npass) {
            atomicAdd(&hist[p * RADIX_MAX_BINS + ((u32)(key >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u))], 1u);
        } else if (p + 1 == plan.npass) {
            // the most significant digit is nearly uniform inside a warp (arrivals come cell by cell):
            // aggregate in the warp instead of serialising 32 atomics on one address
            const u32 d = (u32)(key >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u);
            const u32 peers = __match_any_sync(__activemask(), d);
            if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) atomicAdd(&hist[p * RADIX_MAX_BINS + d], (u32)__popc(peers));
        }
    }
}
__device__ __forceinline__ void radix_hist_flush(const u32* __restrict__ hist, const RadixPlan& plan, u32* __restrict__ ghist) {
    for (int i = threadIdx.x; i < plan.npass * RADIX_MAX_BINS; i += blockDim.x) {
        u32 v = hist[i];
        if (v) atomicAdd(&ghist[i], v);
    }
}

// ---- one pass -----------------------------------------------------------------------------------------------
// dynamic shared memory: u64 skeys[TILE]; u32 svals[TILE] (PAIRS); u32 wcnt[WARPS][MAX_BINS]; u32 dig_excl[MAX_BINS];
//                        u32 gbase[MAX_BINS]
template <bool PAIRS>
struct RadixSmem {
    static constexpr int bytes = RADIX_TILE * 8 + (PAIRS ? RADIX_TILE * 4 : 0) + RADIX_WARPS * RADIX_MAX_BINS * 4 +
                                 2 * RADIX_MAX_BINS * 4;
};

template <bool PAIRS>
__global__ void __launch_bounds__(RADIX_THREADS, PAIRS ? 3 : 4)
k_onesweep(const u64* __restrict__ keys_in, u64* __restrict__ keys_out, const u32* __restrict__ vals_in,
           u32* __restrict__ vals_out, u32 n, int shift, int bits, const u32* __restrict__ ghist /* [bins] */,
           u32* __restrict__ status /* [ntiles][bins] */, u32* __restrict__ counter) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* skeys = (u64*)smem_raw;
    u32* svals = (u32*)(skeys + RADIX_TILE);
    u32* cnt = svals + (PAIRS ? RADIX_TILE : 0);  // [WARPS][bins]
    u32* dig_excl = cnt + RADIX_WARPS * RADIX_MAX_BINS;
    u32* gbase = dig_excl + RADIX_MAX_BINS;
    __shared__ u32 s_warp_sums[2][RADIX_WARPS];
    __shared__ u32 s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 bins = 1u << bits, mask = bins - 1u;
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
    for (int i = tid; i < RADIX_WARPS * RADIX_MAX_BINS; i += RADIX_THREADS) cnt[i] = 0;
    __syncthreads();
    const u32 tile = s_tile;
    const u64 tbase = (u64)tile * RADIX_TILE;
    const u32 tile_n = (u32)min((u64)RADIX_TILE, (u64)n - tbase);

    // warp-striped ownership: item k of lane l in warp w is element w*32*ITEMS + k*32 + l (order: k, then lane)
    u64 key[RADIX_ITEMS];
    const u32 wbase = warp * 32 * RADIX_ITEMS + lane;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        key[k] = (j < tile_n) ? keys_in[tbase + j] : ~0ull;
    }
    // early counts: per-warp digit histogram with shared-memory atomics, so that the tile's counts can be
    // published (and the tiles behind us unblocked) before the expensive stable ranking
    u32* wcnt = cnt + warp * RADIX_MAX_BINS;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        if (j < tile_n) atomicAdd(&wcnt[(u32)(key[k] >> shift) & mask], 1u);
    }
    __syncthreads();

    // digits owned by this thread: [d0, d0 + DPT)
    const u32 dpt = (bins + RADIX_THREADS - 1) / RADIX_THREADS;  // 1 or 2
    const u32 d0 = tid * dpt;
    u32 count[2] = {0, 0}, gcount[2] = {0, 0};
#pragma unroll
    for (u32 q = 0; q < 2; q++) {
        u32 d = d0 + q;
        if (q < dpt && d < bins) {
            u32 running = 0;
#pragma unroll
            for (int w = 0; w < RADIX_WARPS; w++) running += cnt[w * RADIX_MAX_BINS + d];
            count[q] = running;
            gcount[q] = ghist[d];
            u32 word = running | (tile == 0 ? RADIX_FLAG_INCL : RADIX_FLAG_AGG);
            *(volatile u32*)&status[(u64)tile * bins + d] = word;
        }
    }
    // exclusive scan over digits of (tile count, global count)
    u32 tsum = count[0] + count[1], gsum = gcount[0] + gcount[1];
    u32 tinc = tsum, ginc = gsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        u32 a = __shfl_up_sync(0xFFFFFFFFu, tinc, o), b = __shfl_up_sync(0xFFFFFFFFu, ginc, o);
        if (lane >= o) { tinc += a; ginc += b; }
    }
    if (lane == 31) { s_warp_sums[0][warp] = tinc; s_warp_sums[1][warp] = ginc; }
    __syncthreads();
    u32 toff = 0, goff = 0;
#pragma unroll
    for (int w = 0; w < RADIX_WARPS; w++)
        if (w < warp) { toff += s_warp_sums[0][w]; goff += s_warp_sums[1][w]; }
    u32 texcl = toff + tinc - tsum, gexcl = goff + ginc - gsum;
    // cnt[w][d] := staged position of warp w's first key of digit d; dig_excl[d] := first staged position of d
    u32 gex[2] = {0, 0};
#pragma unroll
    for (u32 q = 0; q < 2; q++) {
        u32 d = d0 + q;
        if (q < dpt && d < bins) {
            u32 running = texcl;
#pragma unroll
            for (int w = 0; w < RADIX_WARPS; w++) {
                u32 c = cnt[w * RADIX_MAX_BINS + d];
                cnt[w * RADIX_MAX_BINS + d] = running;
                running += c;
            }
            dig_excl[d] = texcl;
            gex[q] = gexcl;
            texcl += count[q];
            gexcl += gcount[q];
        }
    }
    __syncthreads();

    // stable ranking inside the warp (match on the digit), keys go straight to their staged position
    const u32 lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        bool valid = j < tile_n;
        u32 d = valid ? ((u32)(key[k] >> shift) & mask) : bins;  // invalid lanes form their own match group
        u32 peers = __match_any_sync(0xFFFFFFFFu, d);
        u32 lower = __popc(peers & lt);
        u32 before = valid ? wcnt[d] : 0u;
        __syncwarp();
        if (valid && lower == 0) wcnt[d] = before + __popc(peers);
        __syncwarp();
        if (valid) {
            skeys[before + lower] = key[k];
            if (PAIRS) svals[before + lower] = vals_in[tbase + j];
        }
    }

    // decoupled look-back, one digit at a time per thread (the tiles before us had the whole ranking phase
    // to publish)
#pragma unroll
    for (u32 q = 0; q < 2; q++) {
        u32 d = d0 + q;
        if (q < dpt && d < bins) {
            u32 prev = 0;
            if (tile > 0) {
                int t = (int)tile - 1;
                bool done = false;
                while (!done) {   // 8 predecessors per step: their status words are independent loads
                    u32 w[8];
#pragma unroll
                    for (int x = 0; x < 8; x++)
                        w[x] = (t - x >= 0) ? *(volatile const u32*)&status[(u64)(t - x) * bins + d] : RADIX_FLAG_INCL;
#pragma unroll
                    for (int x = 0; x < 8; x++) {
                        if (!done) {
                            u32 s = w[x];
                            while ((s & (RADIX_FLAG_AGG | RADIX_FLAG_INCL)) == 0u) s = *(volatile const u32*)&status[(u64)(t - x) * bins + d];
                            prev += s & RADIX_COUNT_MASK;
                            done = (s & RADIX_FLAG_INCL) != 0u;
                        }
                    }
                    t -= 8;
                }
                *(volatile u32*)&status[(u64)tile * bins + d] = ((prev + count[q]) & RADIX_COUNT_MASK) | RADIX_FLAG_INCL;
            }
            gbase[d] = gex[q] + prev - dig_excl[d];   // global slot of staged position j of this digit: gbase[d] + j
        }
    }
    __syncthreads();
    for (u32 j = tid; j < tile_n; j += RADIX_THREADS) {
        u64 kx = skeys[j];
        u32 g = gbase[(u32)(kx >> shift) & mask] + j;
        keys_out[g] = kx;
        if (PAIRS) vals_out[g] = svals[j];
    }
}

struct RadixTemp {
    u32* ghist = nullptr;    // [MAX_PASSES * MAX_BINS], zeroed before the key-producing kernel runs
    u32* status = nullptr;   // [npass][ntiles * bins + 1] (last word of each pass = tile counter), zeroed
    static u64 ghist_words() { return (u64)RADIX_MAX_PASSES * RADIX_MAX_BINS; }
    static u64 status_words(u64 n, const RadixPlan& p) {
        u64 ntiles = (n + RADIX_TILE - 1) / RADIX_TILE, w = 0;
        for (int i = 0; i < p.npass; i++) w += ntiles * (1ull << p.bits[i]) + 1;
        return w + 1;
    }
};

// Runs the passes of `plan`.  Ping-pongs between (k0,v0) and (k1,v1); returns 0 or 1 = which pair holds the
// sorted data.  tmp.ghist must already hold the histograms; tmp.status must be zero.
template <bool PAIRS, class Prof>
static inline int radix_sort(u64* k0, u32* v0, u64* k1, u32* v1, u32 n, const RadixPlan& plan, RadixTemp& tmp,
                             cudaStream_t st, Prof& prof) {
    cudaFuncSetAttribute(k_onesweep<PAIRS>, cudaFuncAttributeMaxDynamicSharedMemorySize, RadixSmem<PAIRS>::bytes);
    cudaFuncSetAttribute(k_onesweep<PAIRS>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    int cur = 0;
    if (n < 2) return 0;
    const u32 ntiles = div_up(n, RADIX_TILE);
    u32* status = tmp.status;
    for (int p = 0; p < plan.npass; p++) {
        const u64 words = (u64)ntiles << plan.bits[p];
        u64 *ki = cur ? k1 : k0, *ko = cur ? k0 : k1;
        u32 *vi = cur ? v1 : v0, *vo = cur ? v0 : v1;
        prof.pb();
        k_onesweep<PAIRS><<<ntiles, RADIX_THREADS, RadixSmem<PAIRS>::bytes, st>>>(
            ki, ko, vi, vo, n, plan.shift[p], plan.bits[p], tmp.ghist + p * RADIX_MAX_BINS, status, status + words);
        prof.pe("radix_onesweep", 1, n, (PAIRS ? 24ull : 16ull) * n);
        status += words + 1;
        cur ^= 1;
    }
    return cur;
}

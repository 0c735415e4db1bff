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
//   2. counts its digits per warp (shared-memory atomics) and publishes the tile's counts at once,
//   3. ranks its keys stably per warp (radix_peers: MATCH.ANY or one vote per bit, picked per warp) straight
//      into a shared-memory stage in digit order,
//   4. folds the published counts of the tiles before it (decoupled look-back, one thread per digit, 8 status
//      words per step, flag bits inside the status word) and writes runs of equal digits coalesced.
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
#ifndef RADIX_WINDOW
#define RADIX_WINDOW 8   // predecessors polled per look-back step (independent loads)
#endif

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
    // a real loop over the lower passes (the plan sits in the constant bank: indexed loads, no unrolled
    // compare-and-branch ladder over RADIX_MAX_PASSES), then the top pass
    const int top = plan.npass - 1;
#pragma unroll 1
    for (int p = 0; p < top; p++)
        atomicAdd(&hist[p * RADIX_MAX_BINS + ((u32)(key >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u))], 1u);
    if (top >= 0) {
        // the most significant digit is nearly uniform inside a warp (arrivals come cell by cell):
        // aggregate in the warp instead of serialising 32 atomics on one address
        const u32 d = (u32)(key >> plan.shift[top]) & ((1u << plan.bits[top]) - 1u);
        const u32 peers = __match_any_sync(__activemask(), d);
        if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) atomicAdd(&hist[top * RADIX_MAX_BINS + d], (u32)__popc(peers));
    }
}
__device__ __forceinline__ void radix_hist_flush(const u32* __restrict__ hist, const RadixPlan& plan, u32* __restrict__ ghist) {
    for (int i = threadIdx.x; i < plan.npass * RADIX_MAX_BINS; i += blockDim.x) {
        u32 v = hist[i];
        if (v) atomicAdd(&ghist[i], v);
    }
}

// ---- one pass -----------------------------------------------------------------------------------------------
// dynamic shared memory: u64 skeys[TILE]; u32 svals[TILE] (PAIRS); u32 cnt[WARPS][bins]; u32 dig_excl[bins];
//                        u32 gbase[bins]
// THREADS = 512 (8192-key tiles, two resident blocks per SM) is the production shape: the look-back costs
// bins x window status loads per TILE, so the bigger tile halves it per key and halves the tile rate (the depth
// of the look-back is the number of tiles that publish between a tile's aggregate and its inclusive prefix).
// Lanes of the warp whose digit equals mine (restricted to valid lanes).  Two ways: one vote per digit bit (a fixed
// ~3 instructions per bit whatever the data) or MATCH.ANY, whose cost grows with the number of distinct digits in
// the warp (measured: a pass over incoherent digits runs 2x slower with it, a pass over a handful of distinct
// digits 20 % faster).  MODE 0 = match, 1 = votes, 2 = adaptive: the warp matches its first item, counts the
// groups and picks the cheaper way for the other items.
#define RADIX_ADAPT_GROUPS 8
__device__ __forceinline__ u32 radix_peers_vote(u32 d, bool valid, int bits) {
    u32 peers = __ballot_sync(0xFFFFFFFFu, valid);
#pragma unroll
    for (int b = 0; b < RADIX_MAX_BITS; b++) {
        if (b < bits) {
            const bool bit = (d >> b) & 1u;
            const u32 m = __ballot_sync(0xFFFFFFFFu, bit);
            peers &= bit ? m : ~m;
        }
    }
    return peers;
}
template <int MODE>
__device__ __forceinline__ u32 radix_peers(u32 d, bool valid, int bits, u32 bins, int k, bool& votes) {
    if (MODE == 1 || (MODE == 2 && k > 0 && votes)) return radix_peers_vote(d, valid, bits);
    const u32 peers = __match_any_sync(0xFFFFFFFFu, valid ? d : bins);   // invalid lanes form their own group
    if (MODE == 2 && k == 0) {
        const u32 leaders = __ballot_sync(0xFFFFFFFFu, (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u);
        votes = __popc(leaders) > RADIX_ADAPT_GROUPS;
    }
    return peers;
}

template <bool PAIRS, int THREADS>
struct RadixSmem {
    static constexpr int tile = THREADS * RADIX_ITEMS;
    static constexpr int bytes = tile * 8 + (PAIRS ? tile * 4 : 0) + (THREADS / 32) * RADIX_MAX_BINS * 4 +
                                 2 * RADIX_MAX_BINS * 4;
};
template <bool PAIRS, int THREADS>
struct RadixShape {
    static constexpr int tile = THREADS * RADIX_ITEMS;
    static constexpr int minb = (227 * 1024) / (RadixSmem<PAIRS, THREADS>::bytes + 1024) < (65536 / (64 * THREADS))
                                    ? (227 * 1024) / (RadixSmem<PAIRS, THREADS>::bytes + 1024)
                                    : (65536 / (64 * THREADS));
};

template <bool PAIRS, int THREADS, int MODE>
__global__ void __launch_bounds__(THREADS, RadixShape<PAIRS, THREADS>::minb)
k_onesweep(const u64* __restrict__ keys_in, u64* __restrict__ keys_out, const u32* __restrict__ vals_in,
           u32* __restrict__ vals_out, u32 n, int shift, int bits, const u32* __restrict__ ghist /* [bins] */,
           u32* __restrict__ status /* [ntiles][bins] */, u32* __restrict__ counter) {
    constexpr int TILE = THREADS * RADIX_ITEMS, WARPS = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* skeys = (u64*)smem_raw;
    u32* svals = (u32*)(skeys + TILE);
    u32* cnt = svals + (PAIRS ? TILE : 0);  // [WARPS][bins]
    __shared__ u32 s_warp_sums[2][WARPS];
    __shared__ u32 s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 bins = 1u << bits, mask = bins - 1u;
    u32* dig_excl = cnt + WARPS * bins;
    u32* gbase = dig_excl + bins;
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
    for (u32 i = tid; i < WARPS * bins / 4; i += THREADS) ((uint4*)cnt)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const u32 tile = s_tile;
    const u64 tbase = (u64)tile * TILE;
    const u32 tile_n = (u32)min((u64)TILE, (u64)n - tbase);

    // warp-striped ownership: item k of lane l in warp w is element w*32*ITEMS + k*32 + l (order: k, then lane)
    u64 key[RADIX_ITEMS];
    const u32 wbase = warp * 32 * RADIX_ITEMS + lane;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        key[k] = (j < tile_n) ? keys_in[tbase + j] : ~0ull;
    }
    // early counts: per-warp digit histogram with shared-memory atomics, so that the tile's counts can be
    // published (and the tiles behind us unblocked) before the stable ranking
    u32* wcnt = cnt + warp * bins;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        if (j < tile_n) atomicAdd(&wcnt[(u32)(key[k] >> shift) & mask], 1u);
    }
    __syncthreads();

    // digits owned by this thread: [d0, d0 + DPT)
    constexpr u32 DPT = (RADIX_MAX_BINS + THREADS - 1) / THREADS;   // digits per thread, worst case
    const u32 dpt = (bins + THREADS - 1) / THREADS;
    const u32 d0 = tid * dpt;
    u32* const my_status = status + (u64)tile * bins + d0;
    u32 count[DPT], gcount[DPT];
    u32 tsum = 0, gsum = 0;
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        u32 d = d0 + q;
        count[q] = gcount[q] = 0;
        if (q < dpt && d < bins) {
            u32 running = 0;
#pragma unroll
            for (int w = 0; w < WARPS; w++) running += cnt[w * bins + d];
            count[q] = running;
            gcount[q] = ghist[d];
            u32 word = running | (tile == 0 ? RADIX_FLAG_INCL : RADIX_FLAG_AGG);
            *(volatile u32*)&my_status[q] = word;
        }
        tsum += count[q];
        gsum += gcount[q];
    }
    // exclusive scan over digits of (tile count, global count)
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
    for (int w = 0; w < WARPS; w++)
        if (w < warp) { toff += s_warp_sums[0][w]; goff += s_warp_sums[1][w]; }
    u32 texcl = toff + tinc - tsum, gexcl = goff + ginc - gsum;
    // cnt[w][d] := staged position of warp w's first key of digit d; dig_excl[d] := first staged position of d
    u32 gex[DPT];
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        u32 d = d0 + q;
        gex[q] = 0;
        if (q < dpt && d < bins) {
            u32 running = texcl;
#pragma unroll
            for (int w = 0; w < WARPS; w++) {
                u32 c = cnt[w * bins + d];
                cnt[w * bins + d] = running;
                running += c;
            }
            dig_excl[d] = texcl;
            gex[q] = gexcl;
            texcl += count[q];
            gexcl += gcount[q];
        }
    }
    __syncthreads();

    // stable ranking inside the warp, keys go straight to their staged position
    const u32 lt = (1u << lane) - 1u;
    bool votes = false;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        bool valid = j < tile_n;
        u32 d = (u32)(key[k] >> shift) & mask;
        u32 peers = radix_peers<MODE>(d, valid, bits, bins, k, votes);
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

    // decoupled look-back, one digit at a time per thread
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        u32 d = d0 + q;
        if (q < dpt && d < bins) {
            u32 prev = 0;
            if (tile > 0) {
                int t = (int)tile - 1;
                const volatile u32* sp = my_status + q - bins;   // status word of tile t, digit d
                const u32 back = bins;
                bool done = false;
                while (!done) {   // RADIX_WINDOW predecessors per step: their status words are independent loads
                    u32 w[RADIX_WINDOW];
#pragma unroll
                    for (int x = 0; x < RADIX_WINDOW; x++)
                        w[x] = (t - x >= 0) ? *(sp - (size_t)x * back) : RADIX_FLAG_INCL;
#pragma unroll
                    for (int x = 0; x < RADIX_WINDOW; x++) {
                        if (!done) {
                            u32 s = w[x];
                            while ((s & (RADIX_FLAG_AGG | RADIX_FLAG_INCL)) == 0u) s = *(sp - (size_t)x * back);
                            prev += s & RADIX_COUNT_MASK;
                            done = (s & RADIX_FLAG_INCL) != 0u;
                        }
                    }
                    t -= RADIX_WINDOW;
                    sp -= (size_t)RADIX_WINDOW * back;
                }
                *(volatile u32*)&my_status[q] = ((prev + count[q]) & RADIX_COUNT_MASK) | RADIX_FLAG_INCL;
            }
            gbase[d] = gex[q] + prev - dig_excl[d];   // global slot of staged position j of this digit: gbase[d] + j
        }
    }
    __syncthreads();
    for (u32 j = tid; j < tile_n; j += THREADS) {
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
    static u64 status_words(u64 n, const RadixPlan& p) {   // sized for the smallest tile shape
        u64 ntiles = (n + RADIX_TILE - 1) / RADIX_TILE, w = 0;
        for (int i = 0; i < p.npass; i++) w += ntiles * (1ull << p.bits[i]) + 1;
        return w + 1;
    }
};

// Runs the passes of `plan`.  Ping-pongs between (k0,v0) and (k1,v1); returns 0 or 1 = which pair holds the
// sorted data.  tmp.ghist must already hold the histograms; tmp.status must be zero.
template <bool PAIRS, int THREADS, int MODE, class Prof>
static inline int radix_sort_shape(u64* k0, u32* v0, u64* k1, u32* v1, u32 n, const RadixPlan& plan, RadixTemp& tmp,
                                   cudaStream_t st, Prof& prof) {
    auto kern = k_onesweep<PAIRS, THREADS, MODE>;
    constexpr int smem = RadixSmem<PAIRS, THREADS>::bytes, TILE = THREADS * RADIX_ITEMS;
    static bool attr_set = false;   // per instantiation
    if (!attr_set) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        attr_set = true;
    }
    int cur = 0;
    if (n < 2) return 0;
    const u32 ntiles = div_up(n, TILE);
    u32* status = tmp.status;
    for (int p = 0; p < plan.npass; p++) {
        const u64 words = (u64)ntiles << plan.bits[p];
        u64 *ki = cur ? k1 : k0, *ko = cur ? k0 : k1;
        u32 *vi = cur ? v1 : v0, *vo = cur ? v0 : v1;
        prof.pb();
        kern<<<ntiles, THREADS, smem, st>>>(ki, ko, vi, vo, n, plan.shift[p], plan.bits[p],
                                            tmp.ghist + p * RADIX_MAX_BINS, status, status + words);
        static const bool by_pass = getenv("PCB_PROF_PASSES") != nullptr;   // diagnostics: one stats row per pass index
        static const char* const pass_name[RADIX_MAX_PASSES] = {"radix_p0", "radix_p1", "radix_p2", "radix_p3",
                                                                "radix_p4", "radix_p5", "radix_p6", "radix_p7"};
        prof.pe(by_pass ? pass_name[p] : "radix_onesweep", 1, n, (PAIRS ? 24ull : 16ull) * n);
        status += words + 1;
        cur ^= 1;
    }
    return cur;
}
template <bool PAIRS, class Prof>
static inline int radix_sort(u64* k0, u32* v0, u64* k1, u32* v1, u32 n, const RadixPlan& plan, RadixTemp& tmp,
                             cudaStream_t st, Prof& prof) {
    static const int threads = getenv("PCB_RADIX_THREADS") ? atoi(getenv("PCB_RADIX_THREADS")) : 256;   // tuning hooks
    static const int mode = getenv("PCB_RADIX_MODE") ? atoi(getenv("PCB_RADIX_MODE")) : 2;
    const int v = ((!PAIRS && threads == 512 && n > 4 * RADIX_TILE) ? 3 : 0) + (mode % 3);
#define PCB_RS(T, M) radix_sort_shape<PAIRS, T, M, Prof>(k0, v0, k1, v1, n, plan, tmp, st, prof)
    switch (v) {
        case 0: return PCB_RS(256, 0);
        case 1: return PCB_RS(256, 1);
        case 2: return PCB_RS(256, 2);
        case 3: return PCB_RS(512, 0);
        case 4: return PCB_RS(512, 1);
        default: return PCB_RS(512, 2);
    }
#undef PCB_RS
}

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
//
// SEGMENTED sort (RadixTemp::tinfo != nullptr): the input is already partitioned into SEGMENTS (runs of 2^k
// consecutive cells of a level: arrivals come cell by cell) and only has to be ordered inside each segment, so the
// passes cover the hex-prism bits plus k cell bits instead of all cell bits - three passes per level for the
// default grid whatever the number of cells.  Every segment is an independent sort problem inside the same
// launches: tiles never straddle a segment (a table gives each tile its range, k_seg_tiles), every segment has its
// own digit histograms, and a look-back ends at the first tile of the segment.
#pragma once
#include "common.cuh"
#include "scan.cuh"

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
#define RADIX_WINDOW 4   // predecessors polled per look-back step (independent loads); 8 and 16 measured slower with 256 bins
#endif
#ifdef PCB_RADIX_INSTR   // diagnostics: cycles per phase of a tile, summed over tiles (thread 0 of every block)
__device__ unsigned long long g_radix_instr[16];
#define RADIX_TICK(i) do { if (tid == 0) { long long now_ = clock64(); atomicAdd(&g_radix_instr[i], (unsigned long long)(now_ - tick_)); tick_ = now_; } } while (0)
#define RADIX_COUNT(i, v) do { if (tid == 0) atomicAdd(&g_radix_instr[i], (unsigned long long)(v)); } while (0)
#else
#define RADIX_TICK(i) do { } while (0)
#define RADIX_COUNT(i, v) do { } while (0)
#endif

struct RadixPlan {
    int npass;
    int shift[RADIX_MAX_PASSES];
    int bits[RADIX_MAX_PASSES];
};

// Widest digit the plans use (tuning hook PCB_RADIX_BITS, 4..9).  8: a pass over 256 bins costs ~0.57 ms per 94 M keys
// on the B200 against ~0.75 ms over 512 bins (the look-back reads bins x depth status words per tile), and the default
// grid's 23 prism bits + 1 cell bit make exactly three 8-bit passes (measured per step of C2: 8.6 ms against 10.9 ms
// with three 9-bit passes, 10.7 ms with four 7-bit passes).
#define RADIX_DEFAULT_BITS 8
static inline int radix_digit_bits() {
    static const int b = pcb_tuning_env("PCB_RADIX_BITS") ? std::min(RADIX_MAX_BITS, std::max(4, atoi(pcb_tuning_env("PCB_RADIX_BITS")))) : RADIX_DEFAULT_BITS;
    return b;
}
// Sort bits [lo, lo + nbits) in ceil(nbits / digit width) passes of (almost) equal width.
static inline RadixPlan radix_plan(int lo, int nbits) {
    RadixPlan p;
    p.npass = (nbits + radix_digit_bits() - 1) / radix_digit_bits();
    if (p.npass > RADIX_MAX_PASSES) p.npass = (nbits + RADIX_MAX_BITS - 1) / RADIX_MAX_BITS;
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
// `tag` tells lanes apart whose `hist` differ (segment slots): they must not be aggregated together.
__device__ __forceinline__ void radix_hist_add(u32* __restrict__ hist, const RadixPlan& plan, u64 key, u32 tag = 0) {
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
        const u32 peers = __match_any_sync(__activemask(), d | (tag << RADIX_MAX_BITS));
        if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) atomicAdd(&hist[top * RADIX_MAX_BINS + d], (u32)__popc(peers));
    }
}
// The production plan - three 8-bit passes over 24 consecutive bits - without the plan lookups: one 64-bit shift, three
// byte extractions (k_hexkey is issue-bound; the generic loop above costs ~45 of its ~210 instructions per arrival).
__device__ __forceinline__ bool radix_plan_is_3x8(const RadixPlan& plan) {
    return plan.npass == 3 && plan.bits[0] == 8 && plan.bits[1] == 8 && plan.bits[2] == 8 &&
           plan.shift[1] == plan.shift[0] + 8 && plan.shift[2] == plan.shift[0] + 16;
}
__device__ __forceinline__ void radix_hist_add_3x8(u32* __restrict__ hist, int shift0, u64 key, u32 tag) {
    const u32 kk = (u32)(key >> shift0);
    atomicAdd(&hist[kk & 0xFFu], 1u);
    atomicAdd(&hist[RADIX_MAX_BINS + ((kk >> 8) & 0xFFu)], 1u);
    const u32 d = (kk >> 16) & 0xFFu;   // most significant digit: nearly uniform inside a warp, aggregated (see above)
    const u32 peers = __match_any_sync(__activemask(), d | (tag << RADIX_MAX_BITS));
    if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) atomicAdd(&hist[2 * RADIX_MAX_BINS + d], (u32)__popc(peers));
}
// Segmented variant: a block accumulates the histograms of the first RADIX_MAX_PASSES / npass segments it meets
// (slot = segment - first segment of the block) in shared memory; anything beyond goes to global memory directly.
__device__ __forceinline__ int radix_hist_slots(const RadixPlan& plan) { return RADIX_MAX_PASSES / plan.npass; }
__device__ __forceinline__ void radix_hist_add_global(u32* __restrict__ ghist_seg, const RadixPlan& plan, u64 key) {
#pragma unroll 1
    for (int p = 0; p < plan.npass; p++)
        atomicAdd(&ghist_seg[p * RADIX_MAX_BINS + ((u32)(key >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u))], 1u);
}
__device__ __forceinline__ void radix_hist_flush_seg(const u32* __restrict__ hist, const RadixPlan& plan,
                                                     u32* __restrict__ ghist, u32 seg_lo, u32 nseg) {
    const int per = plan.npass * RADIX_MAX_BINS, slots = radix_hist_slots(plan);
    for (int i = threadIdx.x; i < slots * per; i += blockDim.x) {
        const u32 seg = seg_lo + i / per;
        const u32 v = hist[i];
        if (v && seg < nseg) atomicAdd(&ghist[(u64)seg * per + i % per], v);
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
// THREADS = 256 is the production shape (512 = 8192-key tiles, two resident blocks per SM, measured equal): the look-back costs
// bins x window status loads per TILE, so the bigger tile halves it per key and halves the tile rate (the depth
// of the look-back is the number of tiles that publish between a tile's aggregate and its inclusive prefix).
// Lanes of the warp whose digit equals mine (restricted to valid lanes).  Two ways: one vote per digit bit (a fixed
// ~3 instructions per bit whatever the data) or MATCH.ANY, whose cost grows with the number of distinct digits in
// the warp (measured: a pass over incoherent digits runs 2x slower with it, a pass over a handful of distinct
// digits 20 % faster).  MODE 0 = match, 1 = votes, 2 = adaptive: the warp matches its first item, counts the
// groups and picks the cheaper way for the other items.
#define RADIX_ADAPT_GROUPS 8
// CBITS > 0: the digit width is a compile-time constant (the production plans are all 8-bit passes): no per-bit
// `b < bits` test.  Per bit: test (LOP3 -> predicate), VOTE, SEL, one 3-input LOP3.
template <int CBITS = 0>
__device__ __forceinline__ u32 radix_peers_vote(u32 d, bool valid, int bits) {
    u32 peers = __ballot_sync(0xFFFFFFFFu, valid);
#pragma unroll
    for (int b = 0; b < (CBITS ? CBITS : RADIX_MAX_BITS); b++) {
        if (CBITS || b < bits) {
            const u32 sel = d & (1u << b);
            const u32 m = __ballot_sync(0xFFFFFFFFu, sel != 0u);
            peers &= m ^ (sel ? 0u : 0xFFFFFFFFu);
        }
    }
    return peers;
}
template <int MODE, int CBITS = 0>
__device__ __forceinline__ u32 radix_peers(u32 d, bool valid, int bits, u32 bins, int k, bool& votes) {
    if (MODE == 1 || (MODE == 2 && k > 0 && votes)) return radix_peers_vote<CBITS>(d, valid, bits);
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

template <bool PAIRS, int THREADS, int MODE, int CBITS>
__global__ void __launch_bounds__(THREADS, RadixShape<PAIRS, THREADS>::minb)
k_onesweep(const u64* __restrict__ keys_in, u64* __restrict__ keys_out, const u32* __restrict__ vals_in,
           u32* __restrict__ vals_out, u32 n, int shift, int bits_arg, const u32* __restrict__ ghist /* [bins] */,
           u32* __restrict__ status /* [ntiles][bins] */, u32* __restrict__ counter,
           const uint4* __restrict__ tinfo /* segmented: per tile {first key, keys, segment, first key of the segment} */,
           u32 hist_stride /* words between the histograms of consecutive segments */, u32* __restrict__ err) {
    constexpr int TILE = THREADS * RADIX_ITEMS, WARPS = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* skeys = (u64*)smem_raw;
    u32* svals = (u32*)(skeys + TILE);
    u32* cnt = svals + (PAIRS ? TILE : 0);  // [WARPS][bins]
    __shared__ u32 s_warp_sums[2][WARPS];
    __shared__ u32 s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bits = CBITS ? CBITS : bits_arg;   // CBITS > 0: compile-time digit width (bins, digits per thread, votes)
    const u32 bins = 1u << bits, mask = bins - 1u;
    u32* dig_excl = cnt + WARPS * bins;
    u32* gbase = dig_excl + bins;
#ifdef PCB_RADIX_INSTR
    long long tick_ = clock64();
#endif
    u32 tile, tile_n, seg_first = 0;
    u64 tbase;
    bool first;   // first tile of its sort problem: nothing to look back at
    if (tinfo) {
        // segmented: tile = block index (tiles before this one are resident or done: 1-D grids are dispatched in
        // index order, the assumption the chained scans make as well)
        tile = blockIdx.x;
        const uint4 ti = __ldg(tinfo + tile);
        if ((ti.y & 0x7FFFFFFFu) == 0u) return;   // padding of the tile table
        tbase = ti.x;
        tile_n = ti.y & 0x7FFFFFFFu;
        first = (ti.y >> 31) != 0u;
        ghist += (u64)ti.z * hist_stride;
        seg_first = ti.w;
        for (u32 i = tid; i < WARPS * bins / 4; i += THREADS) ((uint4*)cnt)[i] = make_uint4(0, 0, 0, 0);
        __syncthreads();
    } else {
        if (tid == 0) s_tile = atomicAdd(counter, 1u);
        for (u32 i = tid; i < WARPS * bins / 4; i += THREADS) ((uint4*)cnt)[i] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        tile = s_tile;
        tbase = (u64)tile * TILE;
        tile_n = (u32)min((u64)TILE, (u64)n - tbase);
        first = tile == 0;
    }
    RADIX_TICK(0);   // ticket / tile table + zeroing
    RADIX_COUNT(15, 1);

    // warp-striped ownership: item k of lane l in warp w is element w*32*ITEMS + k*32 + l (order: k, then lane)
    u64 key[RADIX_ITEMS];
    const u32 wbase = warp * 32 * RADIX_ITEMS + lane;
    // item k is valid iff wbase + 32 k < tile_n iff k < kmax: one compare against a register per item (the compiler
    // otherwise rebuilds j from the thread index for every item under the 64-register cap)
    const int kmax = tile_n > wbase ? (int)min((u32)RADIX_ITEMS, (tile_n - wbase + 31u) / 32u) : 0;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        key[k] = (k < kmax) ? keys_in[tbase + j] : ~0ull;
    }
    // early counts: per-warp digit histogram with shared-memory atomics, so that the tile's counts can be
    // published (and the tiles behind us unblocked) before the stable ranking
    u32* wcnt = cnt + warp * bins;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++)
        if (k < kmax) atomicAdd(&wcnt[(u32)(key[k] >> shift) & mask], 1u);
    __syncthreads();
    RADIX_TICK(1);   // key loads + early counts

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
            u32 word = running | (first ? RADIX_FLAG_INCL : RADIX_FLAG_AGG);
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
    u32 texcl = toff + tinc - tsum, gexcl = seg_first + goff + ginc - gsum;
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
    RADIX_TICK(2);   // digit scan + publication

    // stable ranking inside the warp, keys go straight to their staged position
    const u32 lt = (1u << lane) - 1u;
    bool votes = false;
#pragma unroll
    for (int k = 0; k < RADIX_ITEMS; k++) {
        u32 j = wbase + k * 32;
        bool valid = k < kmax;
        u32 d = (u32)(key[k] >> shift) & mask;
        u32 peers = radix_peers<MODE, CBITS>(d, valid, bits, bins, k, votes);
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

    RADIX_TICK(3);   // ranking
    // decoupled look-back, one digit at a time per thread
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        u32 d = d0 + q;
        if (q < dpt && d < bins) {
            u32 prev = 0;
            if (!first) {
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
                            u32 spins = 0;
                            while ((s & (RADIX_FLAG_AGG | RADIX_FLAG_INCL)) == 0u) {
                                s = *(sp - (size_t)x * back);
                                if (++spins > PCB_SPIN_LIMIT) {   // starved (see scan.cuh): fail the build, do not hang
                                    if (err) atomicOr(err, PCB_DERR_TIMEOUT);
                                    s = RADIX_FLAG_INCL;
                                }
                            }
                            prev += s & RADIX_COUNT_MASK;
                            done = (s & RADIX_FLAG_INCL) != 0u;
                        }
                    }
                    t -= RADIX_WINDOW;
                    sp -= (size_t)RADIX_WINDOW * back;
                    if (q == 0) RADIX_COUNT(14, 1);
                }
                *(volatile u32*)&my_status[q] = ((prev + count[q]) & RADIX_COUNT_MASK) | RADIX_FLAG_INCL;
            }
            gbase[d] = gex[q] + prev - dig_excl[d];   // global slot of staged position j of this digit: gbase[d] + j
        }
    }
    RADIX_TICK(4);   // look-back of thread 0
    __syncthreads();
    RADIX_TICK(5);   // waiting for the look-back of the other warps
    for (u32 j = tid; j < tile_n; j += THREADS) {
        u64 kx = skeys[j];
        u32 g = gbase[(u32)(kx >> shift) & mask] + j;
        keys_out[g] = kx;
        if (PAIRS) vals_out[g] = svals[j];
    }
    RADIX_TICK(6);   // scatter (issue)
}

// ---- segment tables ---------------------------------------------------------------------------------------------
// Segment s = cells [s << k, (s + 1) << k) of a level.  k_seg_plan (one block): first tile of every segment
// (tile_base[nseg] = number of tiles).  k_seg_tiles: one record per tile, zero records behind the last tile.
__global__ void __launch_bounds__(1024) k_seg_plan(const u32* __restrict__ cell_start, u32 ncell, int k, u32 nseg,
                                                   u32* __restrict__ tile_base /* [nseg + 1] */) {
    __shared__ u32 sh[1024];
    const u32 per = (nseg + 1023u) / 1024u;
    const u32 lo = min(nseg, threadIdx.x * per), hi = min(nseg, lo + per);
    u32 acc = 0;
    for (u32 s = lo; s < hi; s++) {
        const u32 a = cell_start[min((u64)s << k, (u64)ncell)], b = cell_start[min(((u64)s + 1) << k, (u64)ncell)];
        acc += (b - a + RADIX_TILE - 1) / RADIX_TILE;
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        u32 v = threadIdx.x >= (u32)d ? sh[threadIdx.x - d] : 0u;
        __syncthreads();
        sh[threadIdx.x] += v;
        __syncthreads();
    }
    u32 run = sh[threadIdx.x] - acc;
    for (u32 s = lo; s < hi; s++) {
        const u32 a = cell_start[min((u64)s << k, (u64)ncell)], b = cell_start[min(((u64)s + 1) << k, (u64)ncell)];
        tile_base[s] = run;
        run += (b - a + RADIX_TILE - 1) / RADIX_TILE;
    }
    if (threadIdx.x == 1023) tile_base[nseg] = sh[1023];
}
__global__ void k_seg_tiles(const u32* __restrict__ cell_start, u32 ncell, int k, const u32* __restrict__ tile_base,
                            u32 nseg, u32 ntiles_ub, uint4* __restrict__ tinfo) {
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles_ub) return;
    if (t >= tile_base[nseg]) {
        tinfo[t] = make_uint4(0u, 0u, 0u, 0u);
        return;
    }
    const u32 s = find_segment(tile_base, nseg, t), local = t - tile_base[s];
    const u32 a = cell_start[min((u64)s << k, (u64)ncell)], b = cell_start[min(((u64)s + 1) << k, (u64)ncell)];
    const u32 tb = a + local * RADIX_TILE;
    tinfo[t] = make_uint4(tb, min((u32)RADIX_TILE, b - tb) | (local == 0u ? 0x80000000u : 0u), s, a);
}

struct RadixTemp {
    u32* ghist = nullptr;    // [MAX_PASSES * MAX_BINS], zeroed before the key-producing kernel runs
    u32* status = nullptr;   // [npass][ntiles * bins + 1] (last word of each pass = tile counter), zeroed
    u32* err = nullptr;      // device error word (PCB_DERR_TIMEOUT), optional
    const uint4* tinfo = nullptr;   // segmented sort: tile table (k_seg_tiles), else nullptr
    u32 ntiles_seg = 0;             // ... its length (upper bound of the number of tiles)
    u32 nseg = 1;                   // ... ghist then holds [nseg][npass][RADIX_MAX_BINS]
    static u64 ghist_words() { return (u64)RADIX_MAX_PASSES * RADIX_MAX_BINS; }
    static u64 ghist_words_seg(u64 nseg, const RadixPlan& p) { return nseg * p.npass * RADIX_MAX_BINS; }
    static u64 seg_tiles_bound(u64 n, u64 nseg) { return (n + RADIX_TILE - 1) / RADIX_TILE + nseg; }
    static u64 status_words(u64 n, const RadixPlan& p, u64 nseg = 0) {   // sized for the smallest tile shape
        u64 ntiles = (n + RADIX_TILE - 1) / RADIX_TILE + nseg, w = 0;
        for (int i = 0; i < p.npass; i++) w += ntiles * (1ull << p.bits[i]) + 1;
        return w + 1;
    }
};

// Runs the passes of `plan`.  Ping-pongs between (k0,v0) and (k1,v1); returns 0 or 1 = which pair holds the
// sorted data.  tmp.ghist must already hold the histograms; tmp.status must be zero.
template <bool PAIRS, int THREADS, int MODE, class Prof>
static inline int radix_sort_shape(u64* k0, u32* v0, u64* k1, u32* v1, u32 n, const RadixPlan& plan, RadixTemp& tmp,
                                   cudaStream_t st, Prof& prof) {
    auto kern_any = k_onesweep<PAIRS, THREADS, MODE, 0>;                      // any digit width
    auto kern_def = k_onesweep<PAIRS, THREADS, MODE, RADIX_DEFAULT_BITS>;     // the production width, compiled in
    constexpr int smem = RadixSmem<PAIRS, THREADS>::bytes, TILE = THREADS * RADIX_ITEMS;
    static bool attr_set = false;   // per instantiation
    if (!attr_set) {
        for (auto kern : {kern_any, kern_def}) {
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        }
        attr_set = true;
    }
    static const bool generic_only = pcb_tuning_env("PCB_RADIX_GENERIC") != nullptr;   // tuning hook: A/B of the two
    int cur = 0;
    if (n < 2) return 0;
    const u32 ntiles = tmp.tinfo ? tmp.ntiles_seg : div_up(n, TILE);
    const u32 hist_stride = tmp.tinfo ? (u32)plan.npass * RADIX_MAX_BINS : 0u;
    u32* status = tmp.status;
    for (int p = 0; p < plan.npass; p++) {
        const u64 words = (u64)ntiles << plan.bits[p];
        u64 *ki = cur ? k1 : k0, *ko = cur ? k0 : k1;
        u32 *vi = cur ? v1 : v0, *vo = cur ? v0 : v1;
        prof.pb();
        auto kern = (plan.bits[p] == RADIX_DEFAULT_BITS && !generic_only) ? kern_def : kern_any;
        kern<<<ntiles, THREADS, smem, st>>>(ki, ko, vi, vo, n, plan.shift[p], plan.bits[p],
                                            tmp.ghist + p * RADIX_MAX_BINS, status, status + words, tmp.tinfo,
                                            hist_stride, tmp.err);
        static const bool by_pass = pcb_tuning_env("PCB_PROF_PASSES") != nullptr;   // diagnostics: one stats row per pass index
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
    static const int threads = pcb_tuning_env("PCB_RADIX_THREADS") ? atoi(pcb_tuning_env("PCB_RADIX_THREADS")) : 256;   // tuning hooks
    static const int mode = pcb_tuning_env("PCB_RADIX_MODE") ? atoi(pcb_tuning_env("PCB_RADIX_MODE")) : 2;
    const int v = ((!PAIRS && threads == 512 && n > 4 * RADIX_TILE && !tmp.tinfo) ? 3 : 0) + (mode % 3);
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

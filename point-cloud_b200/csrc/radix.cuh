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
//   4. folds the published counts of the tiles before it (decoupled look-back) and writes runs of equal digits
//      coalesced.  A tile's counts are published PACKED: eight digits x 16 bits in one 16-byte word whose top two
//      bits carry the status, so one thread folds eight digits with one 16-byte load per predecessor and adds the
//      packed words of a whole window before unpacking (the one-thread-per-digit version spent ~40 % of the
//      kernel's instructions on 4-byte status loads).  Inclusive prefixes need 32 bits per digit and live in a
//      second array that is only read where the walk ends.
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
#define RADIX_FLAG_AGG 0x40000000u    // packed counts valid
#define RADIX_FLAG_INCL 0x80000000u   // ... and the tile's inclusive prefixes are (being) published
#define RADIX_FLAG_MASK 0xC0000000u
#define RADIX_INCL_VALID 0x80000000u  // bit 31 of every inclusive-prefix word
#define RADIX_GROUP 8                 // digits per packed status word
#define RADIX_COUNT_MASK 0x3FFFFFFFu  // one-word-per-bin look-back of k_l0_split / k_route_split
#ifndef RADIX_WINDOW
#define RADIX_WINDOW 8
#endif
#ifndef RADIX_LB
#define RADIX_LB 1   // look-back: 1 = one thread per digit (32-bit status words), 3 = packed words, four lanes per group
#endif
#ifdef PCB_RADIX_INSTR   // diagnostics: cycles per phase of a tile, summed over tiles (thread 0 of every block)
__device__ unsigned long long g_radix_instr[16];
#define RADIX_TICK(i) do { if (tid == 0) { long long now_ = clock64(); atomicAdd(&g_radix_instr[i], (unsigned long long)(now_ - tick_)); tick_ = now_; } } while (0)
#define RADIX_COUNT(i, v) do { if (tid == 0) atomicAdd(&g_radix_instr[i], (unsigned long long)(v)); } while (0)
#else
#define RADIX_TICK(i) do { } while (0)
#define RADIX_COUNT(i, v) do { } while (0)
#endif

__device__ __forceinline__ uint4 radix_ld_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void radix_st_v4(uint4* p, const uint4& v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// status words of one pass: packed counts uint4[ntiles][sb / 8], inclusive prefixes u32[ntiles][sb], tile counter (+ pad)
static inline u64 radix_status_bins(int bits) { return (1ull << bits) < RADIX_GROUP ? RADIX_GROUP : (1ull << bits); }
static inline u64 radix_pass_words(u64 ntiles, int bits) {
    const u64 sb = radix_status_bins(bits);
    return ntiles * (sb / 2 + sb) + 4;
}

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
// dynamic shared memory: u64 skeys[TILE]; u32 svals[TILE] (PAIRS); u32 cnt[WARPS][bins]; u32 tcount[bins];
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


// RADIX_LB == 4: the prefixes over tiles are not folded by the tiles themselves.  The first blocks of the grid are
// SCANNERS: one warp per group of eight digits walks the tiles in order, takes the packed counts as they are
// published (32 tiles per load, as many as are ready per step) and stores every tile's inclusive prefixes.  A tile
// then needs ONE read of its predecessor's prefixes, issued after its ranking, by which time the scanner has
// usually passed it: the tile-side look-back (4-5 dependent L2 round trips while the other warps wait at the
// barrier, measured 9 k of a tile's 30 k cycles) shrinks to one round trip.  Scanner blocks have the lowest block
// indices, so they are resident before any tile (1-D grids are dispatched in index order - the assumption the
// chained scans make as well).
template <int THREADS>
__device__ __forceinline__ void radix_scanner(const uint4* __restrict__ agg, u32* __restrict__ incl, u32 ntiles,
                                              u32 ngroups, u32 sb) {
    const int lane = threadIdx.x & 31;
    const u32 g = blockIdx.x * (THREADS / 32) + (threadIdx.x >> 5);
    if (g >= ngroups) return;
    u32 base[RADIX_GROUP];
#pragma unroll
    for (int i = 0; i < RADIX_GROUP; i++) base[i] = 0u;
    u32 t0 = 0;
    while (t0 < ntiles) {
        const u32 t = t0 + lane;
        uint4 w = make_uint4(0u, 0u, 0u, 0u);
        if (t < ntiles) w = radix_ld_v4(agg + (u64)t * ngroups + g);
        // the ready prefix of this window (tiles publish roughly in order)
        const u32 notready = __ballot_sync(0xFFFFFFFFu, (w.w & RADIX_FLAG_MASK) == 0u);
        const u32 p = notready ? (u32)(__ffs(notready) - 1) : 32u;
        if (p == 0u) continue;
        const bool use = (u32)lane < p;
        u32 v[RADIX_GROUP];
        v[0] = w.x & 0xFFFFu; v[1] = w.x >> 16; v[2] = w.y & 0xFFFFu; v[3] = w.y >> 16;
        v[4] = w.z & 0xFFFFu; v[5] = w.z >> 16; v[6] = w.w & 0xFFFFu; v[7] = (w.w & ~RADIX_FLAG_MASK) >> 16;
#pragma unroll
        for (int i = 0; i < RADIX_GROUP; i++) {
            u32 x = use ? v[i] : 0u;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const u32 y = __shfl_up_sync(0xFFFFFFFFu, x, o);
                if (lane >= o) x += y;
            }
            v[i] = x + base[i];
            base[i] = __shfl_sync(0xFFFFFFFFu, v[i], 31);
        }
        if (use) {
            uint4* op = (uint4*)(incl + (u64)t * sb + (u64)g * RADIX_GROUP);
            radix_st_v4(op, make_uint4(v[0] | RADIX_INCL_VALID, v[1] | RADIX_INCL_VALID, v[2] | RADIX_INCL_VALID, v[3] | RADIX_INCL_VALID));
            radix_st_v4(op + 1, make_uint4(v[4] | RADIX_INCL_VALID, v[5] | RADIX_INCL_VALID, v[6] | RADIX_INCL_VALID, v[7] | RADIX_INCL_VALID));
        }
        t0 += p;
    }
}
static inline u32 radix_scanner_blocks(int bits, int threads) {
#if RADIX_LB == 4
    return (u32)((radix_status_bins(bits) / RADIX_GROUP + threads / 32 - 1) / (threads / 32));
#else
    (void)bits; (void)threads;
    return 0u;
#endif
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
           uint4* __restrict__ agg /* [ntiles][sb / 8] */, u32* __restrict__ incl /* [ntiles][sb] */,
           u32* __restrict__ counter) {
    constexpr int TILE = THREADS * RADIX_ITEMS, WARPS = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* skeys = (u64*)smem_raw;
    u32* svals = (u32*)(skeys + TILE);
    u32* cnt = svals + (PAIRS ? TILE : 0);  // [WARPS][bins]
    __shared__ u32 s_warp_sums[2][WARPS];
    __shared__ u32 s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 bins = 1u << bits, mask = bins - 1u;
    u32* tcount = cnt + WARPS * bins;   // digit counts of the tile
    u32* gbase = tcount + bins;
    const u32 sb = bins < RADIX_GROUP ? RADIX_GROUP : bins, ngroups = sb / RADIX_GROUP;
#if RADIX_LB == 4
    const u32 nscan = (ngroups + WARPS - 1) / WARPS;
    if (blockIdx.x < nscan) {
        radix_scanner<THREADS>(agg, incl, gridDim.x - nscan, ngroups, sb);
        return;
    }
#endif
#ifdef PCB_RADIX_INSTR
    long long tick_ = clock64();
#endif
#if RADIX_LB == 4
    const u32 tile = blockIdx.x - nscan;   // tiles before this one are resident or done (in-order dispatch)
    for (u32 i = tid; i < WARPS * bins / 4; i += THREADS) ((uint4*)cnt)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
#else
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
    for (u32 i = tid; i < WARPS * bins / 4; i += THREADS) ((uint4*)cnt)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const u32 tile = s_tile;
#endif
    RADIX_TICK(0);   // ticket + zeroing
    RADIX_COUNT(15, 1);
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
    RADIX_TICK(1);   // key loads + early counts

    // digits owned by this thread: [d0, d0 + DPT)
    constexpr u32 DPT = (RADIX_MAX_BINS + THREADS - 1) / THREADS;   // digits per thread, worst case
    const u32 dpt = (bins + THREADS - 1) / THREADS;
    const u32 d0 = tid * dpt;
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
            tcount[d] = running;
#if RADIX_LB == 1
            *(volatile u32*)&incl[(u64)tile * sb + d] = running | (tile == 0 ? RADIX_FLAG_INCL : RADIX_FLAG_AGG);
#endif
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
#if RADIX_LB != 1
    // publish the tile's counts: one packed 16-byte word per group of eight digits
    if (tid < ngroups) {
        u32 c[RADIX_GROUP];
#pragma unroll
        for (int i = 0; i < RADIX_GROUP; i++) c[i] = (tid * RADIX_GROUP + i < bins) ? tcount[tid * RADIX_GROUP + i] : 0u;
        if (RADIX_LB != 4 && tile == 0) {
            uint4* ip = (uint4*)(incl + (u64)tid * RADIX_GROUP);
            radix_st_v4(ip, make_uint4(c[0] | RADIX_INCL_VALID, c[1] | RADIX_INCL_VALID, c[2] | RADIX_INCL_VALID, c[3] | RADIX_INCL_VALID));
            radix_st_v4(ip + 1, make_uint4(c[4] | RADIX_INCL_VALID, c[5] | RADIX_INCL_VALID, c[6] | RADIX_INCL_VALID, c[7] | RADIX_INCL_VALID));
        }
        radix_st_v4(agg + (u64)tile * ngroups + tid,
                    make_uint4(c[0] | (c[1] << 16), c[2] | (c[3] << 16), c[4] | (c[5] << 16),
                               c[6] | (c[7] << 16) | ((RADIX_LB != 4 && tile == 0) ? RADIX_FLAG_INCL : RADIX_FLAG_AGG)));
    }
#endif
    // cnt[w][d] := staged position of warp w's first key of digit d; gbase[d] := global slot of the digit's first
    // key minus its staged position (the look-back adds the keys of the tiles before this one)
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        u32 d = d0 + q;
        if (q < dpt && d < bins) {
            u32 running = texcl;
#pragma unroll
            for (int w = 0; w < WARPS; w++) {
                u32 c = cnt[w * bins + d];
                cnt[w * bins + d] = running;
                running += c;
            }
            gbase[d] = gexcl - texcl;
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

    RADIX_TICK(3);   // ranking
#if RADIX_LB == 1
    // decoupled look-back, one digit at a time per thread, RADIX_WINDOW predecessors per step (independent loads);
    // status word = count | flag, in the `incl` array
#pragma unroll
    for (u32 q = 0; q < DPT; q++) {
        u32 d = d0 + q;
        if (q < dpt && d < bins) {
            u32 prev = 0;
            if (tile > 0) {
                int t = (int)tile - 1;
                volatile u32* const mine = incl + (u64)tile * sb + d;
                const volatile u32* sp = mine - sb;   // status word of tile t, digit d
                bool done = false;
                while (!done) {
                    u32 w[RADIX_WINDOW];
#pragma unroll
                    for (int x = 0; x < RADIX_WINDOW; x++)
                        w[x] = (t - x >= 0) ? *(sp - (size_t)x * sb) : RADIX_FLAG_INCL;
#pragma unroll
                    for (int x = 0; x < RADIX_WINDOW; x++) {
                        if (!done) {
                            u32 sw = w[x];
                            while ((sw & RADIX_FLAG_MASK) == 0u) sw = *(sp - (size_t)x * sb);
                            prev += sw & RADIX_COUNT_MASK;
                            done = (sw & RADIX_FLAG_INCL) != 0u;
                        }
                    }
                    t -= RADIX_WINDOW;
                    sp -= (size_t)RADIX_WINDOW * sb;
                    if (q == 0) RADIX_COUNT(14, 1);
                }
                *mine = ((prev + count[q]) & RADIX_COUNT_MASK) | RADIX_FLAG_INCL;
            }
            gbase[d] += prev;
        }
    }
#elif RADIX_LB == 4
    // the predecessor's inclusive prefixes, written by the scanner warp of the group
    if ((u32)tid < ngroups) {
        u32 prev[RADIX_GROUP];
#pragma unroll
        for (int i = 0; i < RADIX_GROUP; i++) prev[i] = 0u;
        if (tile > 0) {
            const uint4* q4 = (const uint4*)(incl + (u64)(tile - 1) * sb + (u64)tid * RADIX_GROUP);
            uint4 a, b;
            do {
                a = radix_ld_v4(q4);
                b = radix_ld_v4(q4 + 1);
                RADIX_COUNT(14, 1);
            } while (((a.x & a.y & a.z & a.w & b.x & b.y & b.z & b.w) & RADIX_INCL_VALID) == 0u);
            prev[0] = a.x & ~RADIX_INCL_VALID; prev[1] = a.y & ~RADIX_INCL_VALID;
            prev[2] = a.z & ~RADIX_INCL_VALID; prev[3] = a.w & ~RADIX_INCL_VALID;
            prev[4] = b.x & ~RADIX_INCL_VALID; prev[5] = b.y & ~RADIX_INCL_VALID;
            prev[6] = b.z & ~RADIX_INCL_VALID; prev[7] = b.w & ~RADIX_INCL_VALID;
        }
#pragma unroll
        for (int i = 0; i < RADIX_GROUP; i++)
            if (tid * RADIX_GROUP + i < bins) gbase[tid * RADIX_GROUP + i] += prev[i];
    }
#else
    // decoupled look-back: four lanes per group of eight digits, each lane owns four consecutive predecessors of
    // a 16-tile window (independent 16-byte loads); packed counts are added before unpacking (4 * TILE < 2^16:
    // no carry between fields).  The walk of a group ends at the nearest tile that has inclusive prefixes.
    {
        constexpr int LPG = 4, CH = 4;   // lanes per group, predecessors per lane and step
        const u32 grp = (u32)tid / LPG, sub = (u32)tid % LPG;
        const bool gact = grp < ngroups;
        u32 prev[RADIX_GROUP];
#pragma unroll
        for (int i = 0; i < RADIX_GROUP; i++) prev[i] = 0u;
        bool gdone = !gact || tile == 0;
        long long t = (long long)tile - 1 - (long long)sub * CH;   // my nearest predecessor of this step
        while (__any_sync(0xFFFFFFFFu, !gdone)) {
            uint4 w[CH];
#pragma unroll
            for (int x = 0; x < CH; x++)
                w[x] = (!gdone && t - x >= 0) ? radix_ld_v4(agg + (u64)(t - x) * ngroups + grp)
                                              : make_uint4(0u, 0u, 0u, RADIX_FLAG_MASK);
            uint4 acc = make_uint4(0u, 0u, 0u, 0u);
            uint4 ia = make_uint4(0u, 0u, 0u, 0u), ib = ia;
            bool found = false;
#pragma unroll
            for (int x = 0; x < CH; x++) {
                if (!found) {
                    uint4 sw = w[x];
                    while ((sw.w & RADIX_FLAG_MASK) == 0u) sw = radix_ld_v4(agg + (u64)(t - x) * ngroups + grp);
                    const u32 f = sw.w & RADIX_FLAG_MASK;
                    if (f == RADIX_FLAG_AGG) {
                        acc.x += sw.x; acc.y += sw.y; acc.z += sw.z; acc.w += sw.w & ~RADIX_FLAG_MASK;
                    } else {
                        found = true;
                        if (f == RADIX_FLAG_INCL) {   // the walk ends on this tile's inclusive prefixes
                            const uint4* q4 = (const uint4*)(incl + (u64)(t - x) * sb + (u64)grp * RADIX_GROUP);
                            do {
                                ia = radix_ld_v4(q4);
                                ib = radix_ld_v4(q4 + 1);
                            } while (((ia.x & ia.y & ia.z & ia.w & ib.x & ib.y & ib.z & ib.w) & RADIX_INCL_VALID) == 0u);
                            ia.x &= ~RADIX_INCL_VALID; ia.y &= ~RADIX_INCL_VALID; ia.z &= ~RADIX_INCL_VALID; ia.w &= ~RADIX_INCL_VALID;
                            ib.x &= ~RADIX_INCL_VALID; ib.y &= ~RADIX_INCL_VALID; ib.z &= ~RADIX_INCL_VALID; ib.w &= ~RADIX_INCL_VALID;
                        }
                    }
                }
            }
            // the group's nearest lane that found the end of the walk; lanes farther away contribute nothing
            const u32 fm = (__ballot_sync(0xFFFFFFFFu, found) >> (lane & ~(LPG - 1))) & ((1u << LPG) - 1u);
            const u32 first = fm ? (u32)(__ffs(fm) - 1) : (u32)LPG;
            const bool use = !gdone && sub <= first;
            u32 c8[RADIX_GROUP];
            c8[0] = (acc.x & 0xFFFFu) + ia.x; c8[1] = (acc.x >> 16) + ia.y;
            c8[2] = (acc.y & 0xFFFFu) + ia.z; c8[3] = (acc.y >> 16) + ia.w;
            c8[4] = (acc.z & 0xFFFFu) + ib.x; c8[5] = (acc.z >> 16) + ib.y;
            c8[6] = (acc.w & 0xFFFFu) + ib.z; c8[7] = (acc.w >> 16) + ib.w;
#pragma unroll
            for (int i = 0; i < RADIX_GROUP; i++) {
                u32 v = use ? c8[i] : 0u;
                v += __shfl_xor_sync(0xFFFFFFFFu, v, 1);
                v += __shfl_xor_sync(0xFFFFFFFFu, v, 2);
                prev[i] += v;
            }
            gdone = gdone || fm != 0u;
            t -= LPG * CH;
            RADIX_COUNT(14, 1);
        }
        if (gact && sub == 0) {
            if (tile > 0) {
                u32 c[RADIX_GROUP];
#pragma unroll
                for (int i = 0; i < RADIX_GROUP; i++) c[i] = (grp * RADIX_GROUP + i < bins) ? tcount[grp * RADIX_GROUP + i] : 0u;
                uint4* op = (uint4*)(incl + (u64)tile * sb + (u64)grp * RADIX_GROUP);
                radix_st_v4(op, make_uint4((prev[0] + c[0]) | RADIX_INCL_VALID, (prev[1] + c[1]) | RADIX_INCL_VALID,
                                           (prev[2] + c[2]) | RADIX_INCL_VALID, (prev[3] + c[3]) | RADIX_INCL_VALID));
                radix_st_v4(op + 1, make_uint4((prev[4] + c[4]) | RADIX_INCL_VALID, (prev[5] + c[5]) | RADIX_INCL_VALID,
                                               (prev[6] + c[6]) | RADIX_INCL_VALID, (prev[7] + c[7]) | RADIX_INCL_VALID));
                radix_st_v4(agg + (u64)tile * ngroups + grp,
                            make_uint4(c[0] | (c[1] << 16), c[2] | (c[3] << 16), c[4] | (c[5] << 16),
                                       c[6] | (c[7] << 16) | RADIX_FLAG_INCL));
            }
#pragma unroll
            for (int i = 0; i < RADIX_GROUP; i++)
                if (grp * RADIX_GROUP + i < bins) gbase[grp * RADIX_GROUP + i] += prev[i];
        }
    }
#endif
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

struct RadixTemp {
    u32* ghist = nullptr;    // [MAX_PASSES * MAX_BINS], zeroed before the key-producing kernel runs
    u32* status = nullptr;   // per pass radix_pass_words(): packed counts, inclusive prefixes, tile counter; zeroed, 16-byte aligned
    static u64 ghist_words() { return (u64)RADIX_MAX_PASSES * RADIX_MAX_BINS; }
    static u64 status_words(u64 n, const RadixPlan& p) {   // sized for the smallest tile shape
        u64 ntiles = (n + RADIX_TILE - 1) / RADIX_TILE, w = 0;
        for (int i = 0; i < p.npass; i++) w += radix_pass_words(ntiles, p.bits[i]);
        return w;
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
        const u64 sb = radix_status_bins(plan.bits[p]);
        uint4* agg = (uint4*)status;
        u32* incl = status + (u64)ntiles * (sb / 2);
        u64 *ki = cur ? k1 : k0, *ko = cur ? k0 : k1;
        u32 *vi = cur ? v1 : v0, *vo = cur ? v0 : v1;
        prof.pb();
        kern<<<ntiles + radix_scanner_blocks(plan.bits[p], THREADS), THREADS, smem, st>>>(ki, ko, vi, vo, n, plan.shift[p], plan.bits[p],
                                            tmp.ghist + p * RADIX_MAX_BINS, agg, incl, incl + (u64)ntiles * sb);
        static const bool by_pass = getenv("PCB_PROF_PASSES") != nullptr;   // diagnostics: one stats row per pass index
        static const char* const pass_name[RADIX_MAX_PASSES] = {"radix_p0", "radix_p1", "radix_p2", "radix_p3",
                                                                "radix_p4", "radix_p5", "radix_p6", "radix_p7"};
        prof.pe(by_pass ? pass_name[p] : "radix_onesweep", 1, n, (PAIRS ? 24ull : 16ull) * n);
        status += radix_pass_words(ntiles, plan.bits[p]);
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

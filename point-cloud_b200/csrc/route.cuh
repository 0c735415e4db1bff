// route.cuh — kernels of the multi-GPU level-0 exchange: every rank partitions its slice of the
// undivided input stream by OWNER rank of the point's level-0 cell (metadata.rs:100-102 at level 0) into the
// send buffer of the all-to-all, stably, so that every owner receives its points in stream order.
// The partition is ONE onesweep pass over (owner << pb | index) keys — the collective's packing is the sort's
// scatter, there is no separate pack pass over the points.
#pragma once
#include "level.cuh"

#define ROUTE_SMEM_BINS 8192

// dense per-cell counts of this rank's slice (all-reduced by the host into the global table)
__global__ void __launch_bounds__(256)
k_route_hist(const PcbPoint* __restrict__ pts, u32 n, float S0, L0Box box, u32 vol, u32* __restrict__ hist) {
    __shared__ u32 sh[ROUTE_SMEM_BINS];
    const bool use_sh = vol <= ROUTE_SMEM_BINS;
    if (use_sh) {
        for (u32 i = threadIdx.x; i < vol; i += blockDim.x) sh[i] = 0;
        __syncthreads();
    }
    const double invS0 = 1.0 / (double)S0;
    // four points per thread and step (independent loads in flight).  A warp whose 32 points share a cell (scans are
    // spatially coherent) adds once; otherwise plain atomics: a few-way conflict is cheaper than MATCH.ANY, whose cost
    // grows with the number of distinct cells in the warp.
    constexpr int HI = 4;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i0 = (u64)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * HI) {
        uint4 p[HI];
#pragma unroll
        for (int k = 0; k < HI; k++)
            if (i0 + k * stride < n) p[k] = *(const uint4*)(pts + i0 + k * stride);
#pragma unroll
        for (int k = 0; k < HI; k++) {
            if (i0 + k * stride < n) {
                PcbPoint q;
                q.x = __uint_as_float(p[k].x); q.y = __uint_as_float(p[k].y); q.z = __uint_as_float(p[k].z); q.rgba = p[k].w;
                const u32 id = (u32)l0_cell_id(q, invS0, box);
                const u32 act = __activemask();
                const u32 id0 = __shfl_sync(act, id, __ffs(act) - 1);
                u32* const dst = use_sh ? &sh[id] : &hist[id];
                if (__all_sync(act, id == id0)) {
                    if ((threadIdx.x & 31) == (u32)(__ffs(act) - 1)) atomicAdd(dst, (u32)__popc(act));
                } else {
                    atomicAdd(dst, 1u);
                }
            }
        }
    }
    if (use_sh) {
        __syncthreads();
        for (u32 i = threadIdx.x; i < vol; i += blockDim.x)
            if (sh[i]) atomicAdd(&hist[i], sh[i]);
    }
}

// key = owner << pb | local index, digit histogram of the single radix pass fused
__global__ void __launch_bounds__(256)
k_route_key(const PcbPoint* __restrict__ pts, u32 n, float S0, L0Box box, const unsigned char* __restrict__ owner,
            int pb, RadixPlan plan, u64* __restrict__ keys, u32* __restrict__ ghist) {
    __shared__ u32 hist[RADIX_MAX_PASSES * RADIX_MAX_BINS];
    for (int i = threadIdx.x; i < plan.npass * RADIX_MAX_BINS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const double invS0 = 1.0 / (double)S0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 id = l0_cell_id(pts[i], invS0, box);
        const u64 k = ((u64)__ldg(owner + id) << pb) | i;
        keys[i] = k;
        radix_hist_add(hist, plan, k);
    }
    __syncthreads();
    radix_hist_flush(hist, plan, ghist);
}

// send buffer: points in (owner, stream order) + their global batch ordinals (lib.rs:31-35 on the undivided stream)
__global__ void k_route_gather(const PcbPoint* __restrict__ pts, const u64* __restrict__ keys, u32 n, int pb,
                               u64 first_index, u32 batch_size, PcbPoint* __restrict__ send_pts,
                               u32* __restrict__ send_bat) {
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const u32 i = (u32)(keys[j] & ((1ull << pb) - 1ull));
    *(uint4*)(send_pts + j) = __ldg((const uint4*)(pts + i));
    send_bat[j] = (u32)((first_index + i) / batch_size);
}

// ---- stable split by owner, up to 8 owners (one node) --------------------------------------------------------
// One pass instead of key + radix pass + gather: a point is read once (16 B) and written once (16 + 4 B) at
// send_base[owner] + (number of earlier points of this slice with the same owner).  The running per-owner counts
// are a chained scan over tiles (decoupled look-back, scan.cuh); inside a tile the eight counters are packed
// 4 x 16 bit like the octant scan's.
#define ROUTE_SPLIT_THREADS 256
#define ROUTE_SPLIT_ITEMS 8
#define ROUTE_SPLIT_TILE (ROUTE_SPLIT_THREADS * ROUTE_SPLIT_ITEMS)
struct RouteCountOp {
    typedef Oct8 State;
    __device__ static void pack(const State& s, u32 status, uint4& a, uint4& b) {   // counters < 2^30
        a = make_uint4(s.c[0], s.c[1], s.c[2], s.c[3] | status);
        b = make_uint4(s.c[4], s.c[5], s.c[6], s.c[7] | status);
    }
    __device__ static State unpack(const uint4& a, const uint4& b) {
        State s;
        s.c[0] = a.x; s.c[1] = a.y; s.c[2] = a.z; s.c[3] = a.w & ~PCB_DESC_MASK;
        s.c[4] = b.x; s.c[5] = b.y; s.c[6] = b.z; s.c[7] = b.w & ~PCB_DESC_MASK;
        s.head = 0;
        return s;
    }
    __device__ State identity() const {
        State s;
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = 0;
        s.head = 0;
        return s;
    }
    __device__ State combine(const State& a, const State& b) const {
        State s;
#pragma unroll
        for (int k = 0; k < 8; k++) s.c[k] = a.c[k] + b.c[k];
        s.head = 0;
        return s;
    }
};

// per-owner totals of the slice from its dense level-0 histogram, and their exclusive prefix (send offsets)
__global__ void k_route_owner_counts(const u32* __restrict__ lhist, u32 vol, const unsigned char* __restrict__ owner,
                                     u32* __restrict__ counts /* [8] */, u32* __restrict__ bases /* [8] */) {
    __shared__ u32 sc[8];
    if (threadIdx.x < 8) sc[threadIdx.x] = 0;
    __syncthreads();
    u32 loc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (u32 i = threadIdx.x; i < vol; i += blockDim.x) {
        const u32 h = lhist[i];
        if (h) {
            const u32 o = owner[i] & 7u;
#pragma unroll
            for (int k = 0; k < 8; k++) loc[k] += (o == (u32)k) ? h : 0u;
        }
    }
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (loc[k]) atomicAdd(&sc[k], loc[k]);
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 run = 0;
        for (int k = 0; k < 8; k++) { counts[k] = sc[k]; bases[k] = run; run += sc[k]; }
    }
}

// Where owner o's points go: slot off[o] (+ dev_bases[o] when given) of (pts[o], bat[o]) - struct RouteDst, level.cuh.
// All eight equal = a local send buffer; peer-mapped receive buffers of the other ranks = the exchange itself (stores
// travel over NVLink).
__global__ void __launch_bounds__(ROUTE_SPLIT_THREADS)
k_route_split(const PcbPoint* __restrict__ pts, u32 n, float S0, L0Box box, const unsigned char* __restrict__ owner,
              const u32* __restrict__ dev_bases, u64 first_index, u32 batch_size, RouteDst dst8, u32 world,
              uint4* __restrict__ desc) {
    constexpr int WARPS = ROUTE_SPLIT_THREADS / 32;
    __shared__ u32 s_warp[WARPS][8];
    __shared__ u32 s_tbase[9], s_gstart[8];
    __shared__ uint4 s_pts[ROUTE_SPLIT_TILE];
    __shared__ u32 s_bat[ROUTE_SPLIT_TILE];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 tile = blockIdx.x;   // tile = block index, see k_chain_scan
    // warp-striped: item k of lane l is element wbase + 32 k + l, so one load / store instruction of a warp covers
    // 32 consecutive points and the points of one owner leave as contiguous runs (full sectors, also over NVLink)
    const u64 wbase = (u64)tile * ROUTE_SPLIT_TILE + (u64)warp * (32 * ROUTE_SPLIT_ITEMS);
    const u32 lt = (1u << lane) - 1u;
    uint4 p[ROUTE_SPLIT_ITEMS];
    const double invS0 = 1.0 / (double)S0;
    u32 code[ROUTE_SPLIT_ITEMS];   // owner | rank inside the warp << 4
    u32 wc[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // per-owner counts of this warp so far (same in every lane)
#pragma unroll
    for (int k = 0; k < ROUTE_SPLIT_ITEMS; k++) {
        const u64 e = wbase + 32 * k + lane;
        u32 o = 8u;
        if (e < n) {
            p[k] = __ldg((const uint4*)(pts + e));
            PcbPoint q;
            q.x = __uint_as_float(p[k].x); q.y = __uint_as_float(p[k].y); q.z = __uint_as_float(p[k].z); q.rgba = p[k].w;
            o = __ldg(owner + l0_cell_id(q, invS0, box)) & 7u;
        }
        const u32 peers = __match_any_sync(0xFFFFFFFFu, o);   // at most 9 distinct values
        u32 before = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) before = (o == (u32)q) ? wc[q] : before;
        code[k] = o | ((before + (u32)__popc(peers & lt)) << 4);
#pragma unroll
        for (int q = 0; q < 8; q++)
            if (q < (int)world) wc[q] += (u32)__popc(__ballot_sync(0xFFFFFFFFu, o == (u32)q));
    }
    if (lane < 8) {
        u32 v = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) v = (lane == q) ? wc[q] : v;
        s_warp[warp][lane] = v;
    }
    __syncthreads();
    RouteCountOp op;
    if (warp == 0) {
        Oct8 tot;
#pragma unroll
        for (int q = 0; q < 8; q++) {
            u32 t = 0;
#pragma unroll
            for (int w = 0; w < WARPS; w++) t += s_warp[w][q];
            tot.c[q] = t;
        }
        tot.head = 0;
        Oct8 prefix = chain_lookback(op, desc, tile, tot);
        if (lane == 0) {
            u32 run = 0;
#pragma unroll
            for (int q = 0; q < 8; q++) {
                s_tbase[q] = run;   // first staged slot of owner q in this tile
                s_gstart[q] = dst8.off[q] + (dev_bases ? __ldg(dev_bases + q) : 0u) + prefix.c[q];
                run += tot.c[q];
            }
            s_tbase[8] = run;
        }
    }
    __syncthreads();
    // stage the tile in owner order in shared memory, then write it out: one owner's points of a tile leave as ONE
    // contiguous run (~TILE / world points), which is what the NVLink stores need at 8 owners (direct stores from
    // the warp-striped registers were runs of 4 points there)
    u32 start[8];
#pragma unroll
    for (int q = 0; q < 8; q++) {
        u32 t = s_tbase[q];
#pragma unroll
        for (int w = 0; w < WARPS; w++) t += (w < warp) ? s_warp[w][q] : 0u;
        start[q] = t;
    }
#pragma unroll
    for (int k = 0; k < ROUTE_SPLIT_ITEMS; k++) {
        const u32 o = code[k] & 0xFu;
        if (o < 8u) {
            u32 slot = 0;
#pragma unroll
            for (int q = 0; q < 8; q++) slot = (o == (u32)q) ? start[q] : slot;
            slot += code[k] >> 4;
            s_pts[slot] = p[k];
            s_bat[slot] = (u32)((first_index + wbase + 32 * k + lane) / batch_size);
        }
    }
    __syncthreads();
    const u32 tile_n = s_tbase[8];
    for (u32 j = tid; j < tile_n; j += ROUTE_SPLIT_THREADS) {
        u32 o = 0;
#pragma unroll
        for (int q = 1; q < 8; q++) o += (s_tbase[q] <= j) ? 1u : 0u;   // bases are non-decreasing
        // an owner without points in this tile shares its base with the next one: the count above then lands on
        // the LAST owner whose base is <= j, which is the one that has points
        const u32 g = s_gstart[o] + (j - s_tbase[o]);
        PcbPoint* dp = nullptr;
        u32* db = nullptr;
#pragma unroll
        for (int q = 0; q < 8; q++) {
            dp = (o == (u32)q) ? dst8.pts[q] : dp;
            db = (o == (u32)q) ? dst8.bat[q] : db;
        }
        *(uint4*)(dp + g) = s_pts[j];
        db[g] = s_bat[j];
    }
}

// ---- level-l re-balance: pack whole cells into the all-to-all send buffer ---------------------------------------------
// Selected cells sel[0 .. nsel) of the source arrivals (CSR cs) are copied back to back: output element j belongs to
// selected cell k = the segment of out_start containing j and comes from source element cs[sel[k]] + (j - out_start[k]).
// One binary search per warp, then a walk (cells are thousands of elements long); 16 + 4 bytes per element, coalesced.
__global__ void __launch_bounds__(256)
k_gather_cells(const uint4* __restrict__ pts, const u32* __restrict__ bat, const u32* __restrict__ cs,
               const u32* __restrict__ sel, const u32* __restrict__ out_start /* [nsel + 1] */, u32 nsel,
               uint4* __restrict__ out_pts, u32* __restrict__ out_bat) {
    const u32 total = __ldg(out_start + nsel);
    for (u64 j0 = ((u64)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; j0 < total; j0 += (u64)gridDim.x * blockDim.x) {
        const u64 j = j0 + (threadIdx.x & 31);
        u32 k = find_segment(out_start, nsel, (u32)j0);
        if (j < total) {
            while (__ldg(out_start + k + 1) <= (u32)j) k++;
            const u32 src = __ldg(cs + __ldg(sel + k)) + ((u32)j - __ldg(out_start + k));
            out_pts[j] = __ldg(pts + src);
            out_bat[j] = __ldg(bat + src);
        }
    }
}

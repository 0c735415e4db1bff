// scan.cuh — device-wide scan with a user operator (reduce-then-scan, no spin-waits).
//
// Three launches: (1) every tile reduces its elements to one State, (2) one block
// turns the tile aggregates into exclusive prefixes, (3) every tile re-reads its
// elements and hands each one its exclusive and inclusive prefix.  The operator
// need not be commutative (segmented scans).  Elements are assigned blocked:
// thread t of a tile owns ITEMS consecutive elements, so `consume` sees them in
// sequence order.
//
// Op interface:
//   typedef ... State;
//   __device__ State identity() const;
//   __device__ State load(u32 i) const;                         // element i as a State
//   __device__ State combine(const State& a, const State& b) const;   // a before b
//   __device__ void consume(u32 i, const State& excl, const State& incl) const;
//   __device__ void finish(const State& total) const;           // once, reduction of everything
#pragma once
#include "common.cuh"

template <class Op, int THREADS>
__device__ __forceinline__ typename Op::State block_exclusive_scan(const Op& op, typename Op::State mine,
                                                                   typename Op::State* sh /*[2*THREADS]*/,
                                                                   typename Op::State* block_total) {
    typedef typename Op::State State;
    const int t = threadIdx.x;
    int cur = 0;
    sh[t] = mine;
    __syncthreads();
#pragma unroll 1
    for (int d = 1; d < THREADS; d <<= 1) {
        State v = sh[cur * THREADS + t];
        if (t >= d) v = op.combine(sh[cur * THREADS + t - d], v);
        sh[(cur ^ 1) * THREADS + t] = v;
        cur ^= 1;
        __syncthreads();
    }
    State excl = (t == 0) ? op.identity() : sh[cur * THREADS + t - 1];
    if (block_total) *block_total = sh[cur * THREADS + THREADS - 1];
    __syncthreads();
    return excl;
}

template <class Op, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS) k_scan_reduce(Op op, u32 n, typename Op::State* __restrict__ agg) {
    typedef typename Op::State State;
    __shared__ State sh[2 * THREADS];
    const u64 base = (u64)blockIdx.x * (THREADS * ITEMS) + (u64)threadIdx.x * ITEMS;
    State acc = op.identity();
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        u64 i = base + k;
        if (i < n) acc = op.combine(acc, op.load((u32)i));
    }
    State total;
    block_exclusive_scan<Op, THREADS>(op, acc, sh, &total);
    if (threadIdx.x == 0) agg[blockIdx.x] = total;
}

// one block; agg[0..ntiles) -> exclusive prefixes in place
template <class Op, int THREADS>
__global__ void __launch_bounds__(THREADS) k_scan_agg(Op op, typename Op::State* __restrict__ agg, u32 ntiles) {
    typedef typename Op::State State;
    __shared__ State sh[2 * THREADS];
    const u32 per = (ntiles + THREADS - 1) / THREADS;
    const u32 lo = min(ntiles, threadIdx.x * per), hi = min(ntiles, lo + per);
    State acc = op.identity();
    for (u32 i = lo; i < hi; i++) acc = op.combine(acc, agg[i]);
    State total;
    State run = block_exclusive_scan<Op, THREADS>(op, acc, sh, &total);
    for (u32 i = lo; i < hi; i++) {
        State v = agg[i];
        agg[i] = run;
        run = op.combine(run, v);
    }
    if (threadIdx.x == 0) op.finish(total);
}

template <class Op, int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS) k_scan_apply(Op op, u32 n, const typename Op::State* __restrict__ agg) {
    typedef typename Op::State State;
    __shared__ State sh[2 * THREADS];
    const u64 base = (u64)blockIdx.x * (THREADS * ITEMS) + (u64)threadIdx.x * ITEMS;
    State acc = op.identity();
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        u64 i = base + k;
        if (i < n) acc = op.combine(acc, op.load((u32)i));
    }
    State run = block_exclusive_scan<Op, THREADS>(op, acc, sh, (State*)nullptr);
    run = op.combine(agg[blockIdx.x], run);
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        u64 i = base + k;
        if (i < n) {
            State incl = op.combine(run, op.load((u32)i));  // re-load: L1 hit, saves registers
            op.consume((u32)i, run, incl);
            run = incl;
        }
    }
}

// Host launcher.  `agg` must hold div_up(n, THREADS*ITEMS) States.  Returns launches made.
template <class Op, int THREADS = 256, int ITEMS = 8>
static inline int device_scan(const Op& op, u32 n, typename Op::State* agg, cudaStream_t st) {
    const u32 tile = THREADS * ITEMS;
    const u32 ntiles = n ? div_up(n, tile) : 0;
    if (ntiles) k_scan_reduce<Op, THREADS, ITEMS><<<ntiles, THREADS, 0, st>>>(op, n, agg);
    k_scan_agg<Op, 256><<<1, 256, 0, st>>>(op, agg, ntiles);
    if (ntiles) k_scan_apply<Op, THREADS, ITEMS><<<ntiles, THREADS, 0, st>>>(op, n, agg);
    return ntiles ? 3 : 1;
}

template <class Op, int THREADS = 256, int ITEMS = 8>
static inline u64 device_scan_temp_states(u64 n) {
    return (n + (u64)THREADS * ITEMS - 1) / ((u64)THREADS * ITEMS) + 1;
}

// ---- plain exclusive sums ---------------------------------------------------
template <class TIn, class TOut>
struct SumOp {
    typedef TOut State;
    const TIn* in;
    TOut* out;        // exclusive prefix per element (may alias in when types match)
    TOut* total_out;  // optional: sum of everything (also written to out[n] by the caller's choice)
    __device__ State identity() const { return 0; }
    __device__ State load(u32 i) const { return (TOut)in[i]; }
    __device__ State combine(const State& a, const State& b) const { return a + b; }
    __device__ void consume(u32 i, const State& excl, const State&) const { out[i] = excl; }
    __device__ void finish(const State& total) const {
        if (total_out) *total_out = total;
    }
};

// ---------------------------------------------------------------------------------------------
// Single-pass chained scan (decoupled look-back): every element is read ONCE.  Tiles take their
// ordinal from an atomic counter, so every predecessor of a running tile is itself running or done
// (forward progress); a tile publishes its aggregate, then folds its predecessors' published
// aggregates / inclusive prefixes (one warp, 32 predecessors per step, in order: the operator need
// not commute) and publishes its own inclusive prefix.  Publication is data-then-flag with
// st.release.gpu / ld.acquire.gpu.
//
// Op interface as above, except
//   __device__ void consume(u32 i, const State& excl, const State& incl, const State& self) const;
// (`self` = load(i), kept in registers: load may be expensive, e.g. a gather).
// temp layout of a chained scan: uint4 desc[2 * ntiles], then the u32 tile counter; all zero at launch
static inline u64 chain_temp_bytes(u64 ntiles) { return 32 * ntiles + 16; }

__device__ __forceinline__ u32 ld_acquire_u32(const u32* p) {
    u32 v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(u32* p, u32 v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <class State>
__device__ __forceinline__ State shfl_down_state(const State& s, int d) {
    static_assert(sizeof(State) % 4 == 0, "State must be a multiple of 4 bytes");
    State r;
    const u32* a = (const u32*)&s;
    u32* b = (u32*)&r;
#pragma unroll
    for (int w = 0; w < (int)(sizeof(State) / 4); w++) b[w] = __shfl_down_sync(0xFFFFFFFFu, a[w], d);
    return r;
}
template <class State>
__device__ __forceinline__ State shfl_up_state(const State& s, int d) {
    State r;
    const u32* a = (const u32*)&s;
    u32* b = (u32*)&r;
#pragma unroll
    for (int w = 0; w < (int)(sizeof(State) / 4); w++) b[w] = __shfl_up_sync(0xFFFFFFFFu, a[w], d);
    return r;
}
template <class State>
__device__ __forceinline__ State shfl_state(const State& s, int lane) {
    State r;
    const u32* a = (const u32*)&s;
    u32* b = (u32*)&r;
#pragma unroll
    for (int w = 0; w < (int)(sizeof(State) / 4); w++) b[w] = __shfl_sync(0xFFFFFFFFu, a[w], lane);
    return r;
}
template <class State>
__device__ __forceinline__ State ldcg_state(const State* p) {
    State r;
    const u32* a = (const u32*)p;
    u32* b = (u32*)&r;
#pragma unroll
    for (int w = 0; w < (int)(sizeof(State) / 4); w++) b[w] = __ldcg(a + w);
    return r;
}

// Tile descriptor of the chained scans: 32 bytes = two 16-byte words, each carrying the status (0 nothing,
// 1 aggregate, 2 inclusive prefix) in bits 30-31 of its .w, so that data and status travel in the same
// memory transaction: no fence, no L1 invalidation, one volatile 16-byte load per word and poll.
// Every Op supplies   static void pack(const State&, u32 status, uint4& a, uint4& b)   (status already
// shifted: PCB_DESC_AGG / PCB_DESC_INCL OR-ed into a.w and b.w by pack) and   static State unpack(a, b).
// An Op whose State fits the first word sets  static constexpr bool kOneWord = true;  the second word is then neither
// stored nor polled.
template <class Op, class = void>
struct ChainOneWord { static constexpr bool value = false; };
template <class Op>
struct ChainOneWord<Op, decltype((void)Op::kOneWord)> { static constexpr bool value = Op::kOneWord; };
#define PCB_DESC_AGG 0x40000000u
#define PCB_DESC_INCL 0x80000000u
#define PCB_DESC_MASK 0xC0000000u
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_v4(uint4* p, const uint4& v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Warp-wide look-back of tile `tile` (call with all 32 lanes of ONE warp): publishes `total` as the tile's
// aggregate, folds the predecessors, publishes the inclusive prefix and returns the exclusive one (all lanes).
template <class Op>
__device__ __forceinline__ typename Op::State chain_lookback(const Op& op, uint4* __restrict__ desc, u32 tile,
                                                             const typename Op::State& total) {
    typedef typename Op::State State;
    const int lane = threadIdx.x & 31;
    if (lane == 0) {
        uint4 a, b;
        Op::pack(total, tile == 0 ? PCB_DESC_INCL : PCB_DESC_AGG, a, b);
        st_volatile_v4(desc + 2 * (u64)tile, a);
        if (!ChainOneWord<Op>::value) st_volatile_v4(desc + 2 * (u64)tile + 1, b);
    }
    State prefix = op.identity();
    if (tile > 0) {
        int look = (int)tile - 1;
        for (;;) {
            int t = look - lane;
            u32 f = PCB_DESC_INCL;
            State s = op.identity();
            if (t >= 0) {
                uint4 a, b;
                for (;;) {   // both words of one publication (a reader may catch the writer between its two stores)
                    a = ld_volatile_v4(desc + 2 * (u64)t);
                    if (ChainOneWord<Op>::value) b = a; else b = ld_volatile_v4(desc + 2 * (u64)t + 1);
                    f = a.w & PCB_DESC_MASK;
                    if (f != 0u && f == (b.w & PCB_DESC_MASK)) break;
#ifdef PCB_SPIN_BACKOFF
                    __nanosleep(PCB_SPIN_BACKOFF);
#endif
                }
                s = Op::unpack(a, b);
            }
            const u32 m = __ballot_sync(0xFFFFFFFFu, f == PCB_DESC_INCL);
            const int last = m ? (__ffs(m) - 1) : 31;
            if (lane > last) s = op.identity();
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {   // ordered fold: farther tiles (higher lanes) on the left
                State o = shfl_down_state(s, d);
                if (lane + d < 32) s = op.combine(o, s);
            }
            s = shfl_state(s, 0);
            prefix = op.combine(s, prefix);
            if (m) break;
            look -= 32;
        }
        if (lane == 0) {
            uint4 a, b;
            Op::pack(op.combine(prefix, total), PCB_DESC_INCL, a, b);
            st_volatile_v4(desc + 2 * (u64)tile, a);
            if (!ChainOneWord<Op>::value) st_volatile_v4(desc + 2 * (u64)tile + 1, b);
        }
    }
    return prefix;
}

// n_dev (optional): the element count lives on the device; `n` is then only an upper bound that sized the grid (tiles
// behind the real end fold identities through the chain, the last tile of the grid still reports the total).  The Op
// must bound its own loads the same way.
template <class Op, int THREADS, int ITEMS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_chain_scan(Op op, u32 n, u32 ntiles, uint4* __restrict__ desc,
                                                              const u32* __restrict__ n_dev) {
    typedef typename Op::State State;
    if (n_dev) n = __ldg(n_dev);
    constexpr int WARPS = THREADS / 32;
    __shared__ State s_warp[WARPS];
    __shared__ State s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Tile = block index.  The look-back needs every earlier tile to be running or done: blocks of a 1-D grid are
    // dispatched in index order (the assumption CUB's decoupled look-back scan makes as well).  An atomic ticket
    // instead costs its round trip (about 1 us, a third of a tile's lifetime here) before the first load is
    // issued; taking the ticket one tile ahead in a persistent block hides that but leaves tickets idle for a
    // tile's duration, which the tiles behind pay for in their look-back (measured: 2x slower).
    const u32 tile = blockIdx.x;
    const u64 base = (u64)tile * (THREADS * ITEMS) + (u64)tid * ITEMS;

    State v[ITEMS];
    State acc = op.identity();
    if constexpr (Op::kTileLoad) {
        op.load_items(base, v);   // whole-thread load (all threads call it: it may shuffle)
    } else {
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            u64 i = base + k;
            v[k] = (i < n) ? op.load((u32)i) : op.identity();
        }
    }
#pragma unroll
    for (int k = 0; k < ITEMS; k++) acc = op.combine(acc, v[k]);

    // inclusive scan of the thread aggregates inside the warp, then across warps
    State winc = acc;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        State up = shfl_up_state(winc, d);
        if (lane >= d) winc = op.combine(up, winc);
    }
    if (lane == 31) s_warp[warp] = winc;
    __syncthreads();
    State wpre = op.identity();   // warps before mine
    State total = op.identity();  // whole tile (only warp 0, which does the look-back, and the last thread need it)
    if (warp == 0 || tid == THREADS - 1) {
#pragma unroll
        for (int w = 0; w < WARPS; w++) total = op.combine(total, s_warp[w]);
    }
    if (warp > 0) {
#pragma unroll
        for (int w = 0; w < WARPS - 1; w++)
            if (w < warp) wpre = op.combine(wpre, s_warp[w]);
    }
    State texcl = shfl_up_state(winc, 1);
    if (lane == 0) texcl = op.identity();
    texcl = op.combine(wpre, texcl);   // exclusive prefix of this thread inside the tile

    if (warp == 0) {
        State prefix = chain_lookback(op, desc, tile, total);
        if (lane == 0) s_prefix = prefix;
    }
    __syncthreads();
    State run = op.combine(s_prefix, texcl);
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        u64 i = base + k;
        if (i < n) {
            State incl = op.combine(run, v[k]);
            op.consume((u32)i, run, incl, v[k]);
            run = incl;
        }
    }
    if (tile == ntiles - 1 && tid == THREADS - 1) op.finish(op.combine(s_prefix, total));
}

// Host launcher: `temp` must hold chain_temp_bytes(ntiles) bytes.  Returns launches (memset excluded).
template <class Op, int THREADS, int ITEMS, int MINB = 1>
static inline int chain_scan(const Op& op, u32 n, void* temp, cudaStream_t st, const u32* n_dev = nullptr) {
    const u32 ntiles = div_up(n, THREADS * ITEMS);
    if (!ntiles) return 0;
    cudaMemsetAsync(temp, 0, chain_temp_bytes(ntiles), st);
    k_chain_scan<Op, THREADS, ITEMS, MINB><<<ntiles, THREADS, 0, st>>>(op, n, ntiles, (uint4*)temp, n_dev);
    return 1;
}
template <int THREADS, int ITEMS>
static inline u64 chain_scan_temp_bytes(u64 n) {
    return chain_temp_bytes((n + (u64)THREADS * ITEMS - 1) / ((u64)THREADS * ITEMS));
}

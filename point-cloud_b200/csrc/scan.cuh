// scan.cuh - single-pass device-wide scans with a user operator (decoupled look-back), hand-written for sm_100a.
//
// Op interface:
//   typedef ... State;                                              // multiple of 4 bytes
//   static constexpr bool kTileLoad;                                // true: load_items() fills a thread's items at once
//   __device__ State identity() const;
//   __device__ State load(u32 i) const;                             // element i as a State
//   __device__ State combine(const State& a, const State& b) const;   // a before b; need not commute (segmented scans)
//   __device__ void consume(u32 i, const State& excl, const State& incl, const State& self) const;
//                                                                   // (`self` = load(i), kept in registers)
//   __device__ void finish(const State& total) const;               // once, reduction of everything
//   static pack / unpack: the tile descriptor words, see below;  optional member  u32* err  (device error word)
// Elements are assigned blocked: thread t of a tile owns ITEMS consecutive elements, so `consume` sees them in
// sequence order.
//
// Every element is read ONCE.  Tile = block index: a tile publishes its aggregate, then folds its predecessors'
// published aggregates / inclusive prefixes (one warp, 32 predecessors per step, in order) and publishes its own
// inclusive prefix.  The look-back needs every earlier tile to be resident or done, i.e. blocks of a 1-D grid
// dispatched in index order (the assumption CUB's decoupled look-back makes as well: true on a dedicated device, not
// guaranteed under MPS time-slicing, a debugger or compute-sanitizer's serialisation); a look-back that starves gives
// up with PCB_DERR_TIMEOUT instead of hanging the GPU (PCB_SPIN_LIMIT).
#pragma once
#include "common.cuh"

// temp layout of a chained scan: uint4 desc[2 * ntiles] (+ 16 spare bytes); all zero at launch
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
// memory transaction: no fence, no L1 invalidation, one 16-byte load per word and poll.
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
// One 128-bit relaxed.gpu access per word (ld/st .b128: ONE scalar access in the PTX memory model - the four elements
// of a .v4 vector access may legally be observed torn, new status with old payload).  SASS: LDG/STG.E.128.STRONG.GPU.
__device__ __forceinline__ uint4 ld_desc(const uint4* p) {
    unsigned long long lo, hi;
    asm volatile("{\n\t.reg .b128 t;\n\tld.relaxed.gpu.global.b128 t, [%2];\n\tmov.b128 {%0, %1}, t;\n\t}"
                 : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
    return make_uint4((u32)lo, (u32)(lo >> 32), (u32)hi, (u32)(hi >> 32));
}
__device__ __forceinline__ void st_desc(uint4* p, const uint4& v) {
    const unsigned long long lo = ((unsigned long long)v.y << 32) | v.x, hi = ((unsigned long long)v.w << 32) | v.z;
    asm volatile("{\n\t.reg .b128 t;\n\tmov.b128 t, {%1, %2};\n\tst.relaxed.gpu.global.b128 [%0], t;\n\t}"
                 ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
// A look-back that polls longer than this gives up, raises PCB_DERR_TIMEOUT and carries on with an identity prefix: the
// build then FAILS with an error instead of hanging.  ~2^22 polls of >= 0.5 us each: seconds.
#ifndef PCB_SPIN_LIMIT
#define PCB_SPIN_LIMIT (1u << 22)
#endif
// device error word of an Op (member `u32* err`), or null
template <class Op, class = void>
struct ChainErr { __device__ static u32* get(const Op&) { return nullptr; } };
template <class Op>
struct ChainErr<Op, decltype((void)((Op*)nullptr)->err)> { __device__ static u32* get(const Op& op) { return op.err; } };

template <class Op>
__device__ __forceinline__ void chain_publish(uint4* __restrict__ desc, u32 tile, const typename Op::State& s, u32 status) {
    uint4 a, b;
    Op::pack(s, status, a, b);
    st_desc(desc + 2 * (u64)tile, a);
    if (!ChainOneWord<Op>::value) st_desc(desc + 2 * (u64)tile + 1, b);
}

// Warp-wide look-back of tile `tile` (call with all 32 lanes of ONE warp): publishes `total` as the tile's
// aggregate, folds the predecessors, publishes the inclusive prefix and returns the exclusive one (all lanes).
// publish = false: the caller has already published the aggregate (chain_publish by one lane) and did other work since.
template <class Op>
__device__ __forceinline__ typename Op::State chain_lookback(const Op& op, uint4* __restrict__ desc, u32 tile,
                                                             const typename Op::State& total, u32* err = nullptr,
                                                             bool publish = true) {
    typedef typename Op::State State;
    const int lane = threadIdx.x & 31;
    if (publish && lane == 0) chain_publish<Op>(desc, tile, total, tile == 0 ? PCB_DESC_INCL : PCB_DESC_AGG);
    State prefix = op.identity();
    if (tile > 0) {
        int look = (int)tile - 1;
        u32 spins = 0;
        for (;;) {
            int t = look - lane;
            u32 f = PCB_DESC_INCL;
            State s = op.identity();
            if (t >= 0) {
                uint4 a, b;
                bool gave_up = false;
                for (;;) {   // both words of one publication (a reader may catch the writer between its two stores)
                    a = ld_desc(desc + 2 * (u64)t);
                    if (ChainOneWord<Op>::value) b = a; else b = ld_desc(desc + 2 * (u64)t + 1);
                    f = a.w & PCB_DESC_MASK;
                    if (f != 0u && f == (b.w & PCB_DESC_MASK)) break;
                    if (++spins > PCB_SPIN_LIMIT) {   // starved: fail the build, do not hang
                        if (err) atomicOr(err, PCB_DERR_TIMEOUT);
                        gave_up = true;
                        f = PCB_DESC_INCL;
                        break;
                    }
#ifdef PCB_SPIN_BACKOFF
                    __nanosleep(PCB_SPIN_BACKOFF);
#endif
                }
                if (!gave_up) s = Op::unpack(a, b);
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
        if (lane == 0) chain_publish<Op>(desc, tile, op.combine(prefix, total), PCB_DESC_INCL);
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
        State prefix = chain_lookback(op, desc, tile, total, ChainErr<Op>::get(op));
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

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

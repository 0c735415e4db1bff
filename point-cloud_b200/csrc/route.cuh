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
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u32 id = (u32)l0_cell_id(pts[i], S0, box);
        const u32 peers = __match_any_sync(__activemask(), id);       // random points hit few cells: aggregate
        if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) {
            if (use_sh) atomicAdd(&sh[id], (u32)__popc(peers));
            else atomicAdd(&hist[id], (u32)__popc(peers));
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
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 id = l0_cell_id(pts[i], S0, box);
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

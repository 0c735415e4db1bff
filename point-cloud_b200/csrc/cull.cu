// cull.cu - the renderer's per-cell cull / level-of-detail compute pass as a CUDA kernel.
//
// Restates src/plugins/render/point/pipelines/compute.wgsl:81-205 (paths relative to /root/reference), which the
// reference dispatches once per loaded cell with workgroups of 128 points:
//   main                       :175-205  view / clip transform, frustum test, depth lookup, hierarchy pick, depth test,
//                                        colour alpha := hierarchy, append to the output through an atomic counter
//   get_hierarchy              :85-99    deepest loaded hierarchy that contains the point (+ size-by-distance override)
//   search_smallest_hierarchy  :101-126  walk down while the point's cell at the next hierarchy is loaded
//   binary_search              :128-158  over the loaded cells sorted by (hierarchy, x, y, z)
//   cell_index                 :81-83    floor(position / cell_size) as i32 (same as metadata.rs:100-102)
// No graphics API is involved: the depth texture is a plain row-major f32 array, the indirect draw buffer's
// instance_count a u32 in device memory.  Arithmetic: every WGSL operation is ONE IEEE binary32 operation here, in
// source order, never fused (mat4x4 * vec4 = ((c0 * x + c1 * y) + c2 * z) + c3 * w per component); WGSL itself allows an
// implementation to fuse or reorder, so a real wgpu device may differ from this (and from oracle/cull_oracle.cpp, which
// follows the same order) in the last place of a borderline test.  Output ORDER is whatever the atomic gives, as in the
// shader: compare as a set.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pcb.h"

namespace {

struct CullArgs {
    float view[16], proj[16];   // column-major: m[4 * col + row] (glam / wgpu layout)
    float cam[3];
    uint32_t depth_w, depth_h;
    uint32_t nh;                // metadata.number_of_hierarchies
    uint32_t size_by_distance, max_hierarchy;
    uint32_t cell_h;
    uint32_t n_loaded;
};

struct V4 {
    float x, y, z, w;
};

// mat4x4<f32> * vec4<f32>: linear combination of the columns, left to right
__device__ __forceinline__ V4 mat_vec(const float* m, V4 v) {
    V4 r;
    r.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[0], v.x), __fmul_rn(m[4], v.y)), __fmul_rn(m[8], v.z)), __fmul_rn(m[12], v.w));
    r.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[1], v.x), __fmul_rn(m[5], v.y)), __fmul_rn(m[9], v.z)), __fmul_rn(m[13], v.w));
    r.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[2], v.x), __fmul_rn(m[6], v.y)), __fmul_rn(m[10], v.z)), __fmul_rn(m[14], v.w));
    r.w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[3], v.x), __fmul_rn(m[7], v.y)), __fmul_rn(m[11], v.z)), __fmul_rn(m[15], v.w));
    return r;
}

// compute.wgsl:128-158
__device__ __forceinline__ bool loaded(const int4* __restrict__ cells, uint32_t len, int h, int x, int y, int z) {
    int low = 0, high = (int)len - 1;
    while (low <= high) {
        const int mid = (low + high) / 2;
        const int4 c = __ldg(cells + mid);   // (hierarchy, x, y, z)
        if (c.x == h && c.y == x && c.z == y && c.w == z) return true;
        if (c.x < h || (c.x == h && c.y < x) || (c.x == h && c.y == x && c.z < y) || (c.x == h && c.y == x && c.z == y && c.w < z))
            low = mid + 1;
        else
            high = mid - 1;
    }
    return false;
}

// One point through compute.wgsl:175-205 (main): returns whether it is kept and, if so, its hierarchy.
__device__ __forceinline__ bool cull_point(const CullArgs& a, const float2* __restrict__ hier, const int4* __restrict__ cells,
                                           const float* __restrict__ far_dist, const float* __restrict__ depth,
                                           const pcb_point& p, uint32_t* hsel_out) {
    const V4 view = mat_vec(a.view, V4{p.x, p.y, p.z, 1.0f});
    const V4 clip = mat_vec(a.proj, view);
    const float nx = __fdiv_rn(clip.x, clip.w), ny = __fdiv_rn(clip.y, clip.w), nz = __fdiv_rn(clip.z, clip.w);
    if (!(fabsf(nx) <= 1.0f && fabsf(ny) <= 1.0f && fabsf(__fsub_rn(nz, 0.5f)) <= 0.5f)) return false;   // :186
    // :187 uv = vec2<u32>((ndc.xy * vec2(0.5, -0.5) + 0.5) * vec2<f32>(textureDimensions))
    const float fu = __fmul_rn(__fadd_rn(__fmul_rn(nx, 0.5f), 0.5f), (float)a.depth_w);
    const float fv = __fmul_rn(__fadd_rn(__fmul_rn(ny, -0.5f), 0.5f), (float)a.depth_h);
    uint32_t u = __float2uint_rz(fu), v = __float2uint_rz(fv);   // saturating truncation
    // ndc == +-1 lands one texel outside: textureLoad is then implementation-defined; clamped here (robust access)
    u = u < a.depth_w ? u : a.depth_w - 1;
    v = v < a.depth_h ? v : a.depth_h - 1;
    const float d = __ldg(depth + (size_t)v * a.depth_w + u);
    // get_hierarchy, :85-99
    uint32_t own;
    if (a.cell_h >= a.nh - 1) {
        own = a.nh - 1;
    } else {
        uint32_t h = a.cell_h;
        for (;;) {
            h += 1;
            if (h >= a.nh) break;   // (the shader reads hierarchies[nh] through robust access and finds nothing)
            const float cs = __ldg(&hier[h].x);
            const int cx = __float2int_rz(floorf(__fdiv_rn(p.x, cs)));
            const int cy = __float2int_rz(floorf(__fdiv_rn(p.y, cs)));
            const int cz = __float2int_rz(floorf(__fdiv_rn(p.z, cs)));
            if (!loaded(cells, a.n_loaded, (int)h, cx, cy, cz)) break;
        }
        own = h - 1;
    }
    uint32_t hsel = own;
    if (a.size_by_distance) {
        const float dx = __fsub_rn(a.cam[0], p.x), dy = __fsub_rn(a.cam[1], p.y), dz = __fsub_rn(a.cam[2], p.z);
        const float dist = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        for (uint32_t k = a.max_hierarchy; k > own; k--)
            if (dist < __ldg(far_dist + k)) { hsel = k; break; }
    }
    const float radius = __ldg(&hier[hsel].y);
    const V4 mclip = mat_vec(a.proj, V4{view.x, view.y, __fadd_rn(view.z, radius), view.w});
    const float mz = __fdiv_rn(mclip.z, mclip.w);
    const float radius_z = __fsub_rn(nz, mz);
    *hsel_out = hsel;
    return mz < d || __fsub_rn(mz, d) < __fmul_rn(radius_z, 3.0f);   // :198
}

// The shader runs workgroups of 128 points and appends every kept point through atomicAdd(&instance_count, 1); the output
// ORDER is whatever the atomic gives.  Here a block takes 1024 points (four per thread, warp-striped: coalesced) and
// reserves its slots with ONE atomic: 20 M points through a single counter were bound by same-address atomics
// (one per warp: 0.54 ms; scripts/gpu_cull_bench.py).
#define CULL_THREADS 256
#define CULL_ITEMS 4
__global__ void __launch_bounds__(CULL_THREADS)
k_cull_lod(CullArgs a, const float2* __restrict__ hier /* (cell_size, spacing) per hierarchy */,
           const int4* __restrict__ cells, const float* __restrict__ far_dist, const float* __restrict__ depth,
           const pcb_point* __restrict__ in, uint32_t n, pcb_point* __restrict__ out, uint32_t* __restrict__ count) {
    __shared__ uint32_t s_cnt[CULL_ITEMS * (CULL_THREADS / 32)];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t b0 = (uint64_t)blockIdx.x * (CULL_THREADS * CULL_ITEMS);
    pcb_point p[CULL_ITEMS];
    uint32_t hsel[CULL_ITEMS], before[CULL_ITEMS];
    bool keep[CULL_ITEMS];
#pragma unroll
    for (int k = 0; k < CULL_ITEMS; k++) {
        const uint64_t i = b0 + (uint64_t)k * CULL_THREADS + tid;
        keep[k] = false;
        hsel[k] = 0;
        if (i < n) {
            p[k] = in[i];
            keep[k] = cull_point(a, hier, cells, far_dist, depth, p[k], &hsel[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < CULL_ITEMS; k++) {
        const unsigned m = __ballot_sync(0xFFFFFFFFu, keep[k]);
        before[k] = (uint32_t)__popc(m & ((1u << lane) - 1u));
        if (lane == 0) s_cnt[k * (CULL_THREADS / 32) + warp] = (uint32_t)__popc(m);
    }
    __syncthreads();
    if (warp == 0) {   // exclusive scan of the 32 (item, warp) counts, one atomic for the block
        static_assert(CULL_ITEMS * (CULL_THREADS / 32) == 32, "one warp scans the counts");
        const uint32_t c = s_cnt[lane];
        uint32_t inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if (lane >= d) inc += up;
        }
        s_cnt[lane] = inc - c;
        if (lane == 31) s_base = inc ? atomicAdd(count, inc) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < CULL_ITEMS; k++)
        if (keep[k]) {
            p[k].rgba[3] = (uint8_t)hsel[k];   // pack4x8(vec4(unpacked_color.xyz, hierarchy)), :201-202
            out[s_base + s_cnt[k * (CULL_THREADS / 32) + warp] + before[k]] = p[k];
        }
}

}  // namespace

extern "C" int pcb_cull_lod(int device, void* stream, const pcb_cull_params* prm, const float* d_hierarchies,
                            const int32_t* d_loaded_cells, uint32_t n_loaded, const float* d_far_distances,
                            const float* d_depth, const pcb_point* d_in, uint64_t n, pcb_point* d_out, uint32_t* d_count) {
    if (!prm || !d_hierarchies || !d_depth || !d_count || (n && (!d_in || !d_out)) || (n_loaded && !d_loaded_cells)) return PCB_ERR_ARG;
    if (n >= (1ull << 31) || prm->number_of_hierarchies == 0 || prm->depth_width == 0 || prm->depth_height == 0) return PCB_ERR_ARG;
    if (prm->size_by_distance && !d_far_distances) return PCB_ERR_ARG;
    if (prm->size_by_distance && prm->max_hierarchy >= prm->number_of_hierarchies) return PCB_ERR_ARG;
    if (((uintptr_t)d_loaded_cells & 15u) != 0) return PCB_ERR_ARG;   // int4 loads
    if (cudaSetDevice(device) != cudaSuccess) return PCB_ERR_CUDA;
    if (!n) return 0;
    CullArgs a;
    for (int k = 0; k < 16; k++) { a.view[k] = prm->view[k]; a.proj[k] = prm->projection[k]; }
    for (int k = 0; k < 3; k++) a.cam[k] = prm->cam_pos[k];
    a.depth_w = prm->depth_width;
    a.depth_h = prm->depth_height;
    a.nh = prm->number_of_hierarchies;
    a.size_by_distance = prm->size_by_distance;
    a.max_hierarchy = prm->max_hierarchy;
    a.cell_h = prm->cell_hierarchy;
    a.n_loaded = n_loaded;
    const uint64_t per_block = CULL_THREADS * CULL_ITEMS;
    k_cull_lod<<<(unsigned)((n + per_block - 1) / per_block), CULL_THREADS, 0, (cudaStream_t)stream>>>(
        a, (const float2*)d_hierarchies, (const int4*)d_loaded_cells, d_far_distances, d_depth, d_in, (uint32_t)n, d_out, d_count);
    return cudaGetLastError() == cudaSuccess ? 1 : PCB_ERR_CUDA;
}

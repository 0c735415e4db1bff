// common.cuh — shared device/host helpers of the B200 octree/LOD builder.
//
// Arithmetic here restates, op for op, the reference's f32 sequences
// (paths relative to /root/reference):
//   MetadataConfig::cell_index   point-converter/src/metadata.rs:100-102
//   MetadataConfig::cell_pos     point-converter/src/metadata.rs:104-106
//   AxialIndex::from_world       point-converter/src/hex.rs:67-85
//   AxialIndex::to_offset        point-converter/src/hex.rs:45-51
//   OffsetIndex::to_axial        point-converter/src/hex.rs:18-24
//   AxialIndex::to_world         point-converter/src/hex.rs:55-65
//   Vec3::distance_squared       glam 0.27.0 (scalar Vec3): (dx*dx + dy*dy) + dz*dz
// Every operation is a single IEEE-754 binary32 op, round-to-nearest-even,
// never contracted into an FMA (explicit __f*_rn intrinsics; the library is
// also compiled with -fmad=false).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

typedef uint32_t u32;
typedef uint64_t u64;
typedef int32_t i32;

#define PCB_SQRT_3 1.73205080757f  // hex.rs:3
#define PCB_NONE 0xFFFFFFFFu

// device-side error bits (Builder::d_err)
#define PCB_DERR_NONFINITE 1u   // NaN/Inf coordinate
#define PCB_DERR_L0RANGE 2u     // level-0 cell index outside +-2^20
#define PCB_DERR_HEXRANGE 4u    // hex key outside the per-cell window (coordinate precision exhausted)
#define PCB_DERR_NEST 8u        // child index does not nest in its parent (SURVEY N1 violated)
#define PCB_DERR_TIMEOUT 16u    // a look-back waited too long for an earlier tile (blocks not dispatched in order?)

struct __align__(16) PcbPoint {
    float x, y, z;
    u32 rgba;
};

// Rust `f32 as i32`: saturating, NaN -> 0.  cvt.rzi.s32.f32 does exactly that.
__device__ __forceinline__ i32 sat_i32(float v) { return __float2int_rz(v); }

// metadata.rs:100-102 — (pos / cell_size).floor().as_ivec3(), one axis
__device__ __forceinline__ i32 cell_index_axis(float p, float cell_size) {
    return sat_i32(floorf(__fdiv_rn(p, cell_size)));
}

// The reference divides every point by per-level constants (hex.rs:68-69, :79-80, :83; metadata.rs:101).  The IEEE
// quotient a / b of two binary32 numbers equals  (float)((double)a * inv)  with inv = 1.0 / (double)b  computed once:
// the double product is within 2^-52 (relative) of a / b, while a / b - unless exactly representable - stays at least
// ~2^-49 (relative) away from every binary32 rounding boundary (a - m * b is a non-zero multiple of the last place of a
// 49-bit product), and it never IS a midpoint (m * b would need 25 significant bits).  So one conversion, one DMUL
// and one conversion replace the ~22-instruction __fdiv_rn sequence, bit for bit (checked exhaustively over all 2^32
// dividends for b = 3.0 and the default grid's divisors, and on 4 * 10^8 random (a, b) pairs incl. overflow and
// subnormal quotients: tests/test_div_identity.py).
__device__ __forceinline__ float fdiv_inv(float a, double inv) { return __double2float_rn(__dmul_rn((double)a, inv)); }
__device__ __forceinline__ i32 cell_index_axis_inv(float p, double inv_cell_size) {
    return sat_i32(floorf(fdiv_inv(p, inv_cell_size)));
}
// reciprocals of the divisors of hex.rs:67-85 for one level
struct HexInv {
    double x;      // 1 / (R * SQRT_3)
    double y;      // 1 / ((-R) * SQRT_3)
    double z;      // 1 / R
    double third;  // 1 / 3.0
};

struct HexKey {
    i32 x, y, z;
};

// hex.rs:67-85 followed by :45-51 (wrapping i32 arithmetic as in a Rust release build)
__device__ __forceinline__ HexKey hex_from_world(float px, float py, float pz, const HexInv& inv) {
    float x = fdiv_inv(px, inv.x);   // px / (R * SQRT_3)
    float y = fdiv_inv(py, inv.y);   // py / ((-R) * SQRT_3)
    float t = __fadd_rn(__fmul_rn(PCB_SQRT_3, y), 1.0f);
    float temp1 = floorf(__fadd_rn(t, x));
    float temp2 = __fsub_rn(t, x);
    float temp3 = __fadd_rn(__fmul_rn(2.0f, x), 1.0f);
    float qf = fdiv_inv(__fadd_rn(temp1, temp3), inv.third);   // (temp1 + temp3) / 3.0
    float rf = fdiv_inv(__fadd_rn(temp1, temp2), inv.third);
    // floor then saturating conversion == ONE conversion rounding towards minus infinity (cvt.rmi.s32.f32 saturates and
    // maps NaN to 0 like cvt.rzi; beyond the i32 range floorf is the identity)
    i32 q = __float2int_rd(qf);
    i32 r = (i32)(0u - (u32)__float2int_rd(rf));
    i32 h = sat_i32(fdiv_inv(pz, inv.z));   // (pz / R) as i32: truncation
    i32 e = (i32)((u32)r - (u32)(r & 1));  // even => `/ 2` is exact
    HexKey k;
    k.x = (i32)((u32)q + (u32)(e / 2));
    k.y = r;
    k.z = h;
    return k;
}

// hex.rs:18-24 then :55-65, then cell.rs:77-78 distance_squared(centre, p).
// half_sqrt3 = SQRT_3 / 2.0 and row_pitch = (R * 3.0) / 2.0 are the reference's own sub-expressions, evaluated
// once per level with the same single IEEE operations.
__device__ __forceinline__ float hex_d2(HexKey k, float px, float py, float pz, float R, float half_sqrt3,
                                        float row_pitch) {
    i32 e = (i32)((u32)k.y - (u32)(k.y & 1));
    i32 q = (i32)((u32)k.x - (u32)(e / 2));
    float qf = (float)q, rf = (float)k.y, hf = (float)k.z;
    float cx = __fmul_rn(R, __fadd_rn(__fmul_rn(PCB_SQRT_3, qf), __fmul_rn(half_sqrt3, rf)));
    float cy = __fmul_rn(row_pitch, rf);
    float cz = __fmul_rn(hf, R);
    float dx = __fsub_rn(cx, px), dy = __fsub_rn(cy, py), dz = __fsub_rn(cz, pz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// largest c with start[c] <= i  (start has ncell+1 entries, start[0] == 0)
__device__ __forceinline__ u32 find_segment(const u32* __restrict__ start, u32 nseg, u32 i) {
    u32 lo = 0, hi = nseg;  // invariant: start[lo] <= i < start[hi]
    while (hi - lo > 1) {
        u32 mid = (lo + hi) >> 1;
        if (__ldg(start + mid) <= i) lo = mid; else hi = mid;
    }
    return lo;
}

#define CUDA_TRY(expr)                                                          \
    do {                                                                        \
        cudaError_t _e = (expr);                                                \
        if (_e != cudaSuccess) throw PcbCudaError(_e, #expr, __FILE__, __LINE__); \
    } while (0)

struct PcbCudaError {
    cudaError_t code;
    const char* expr;
    const char* file;
    int line;
    PcbCudaError(cudaError_t c, const char* e, const char* f, int l) : code(c), expr(e), file(f), line(l) {}
};

static inline u32 div_up(u64 a, u64 b) { return (u32)((a + b - 1) / b); }

// Tuning and test hooks (PCB_FORCE_PAIRS, PCB_SEG_MIN, PCB_WAVE_POINTS, PCB_TRACE ...) are read from the environment
// ONLY when PCB_ENABLE_TUNING=1 is set as well: a stray PCB_* variable in a user's environment must not change the
// code path of a production conversion.  The parity tests that force a path set both.
static inline const char* pcb_tuning_env(const char* name) {
    static const bool on = [] {
        const char* e = getenv("PCB_ENABLE_TUNING");
        return e && e[0] == '1' && e[1] == 0;
    }();
    return on ? getenv(name) : nullptr;
}

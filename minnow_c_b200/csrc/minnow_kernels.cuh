// minnow_kernels.cuh -- device structs and helpers shared by the kernels of
// libminnow_cuda (sm_100a only). No CPU fallback lives anywhere in this tree.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "minnow_cuda.h"

namespace mnw {

constexpr int kBlockLen = MNWCUDA_BLOCK_LEN;   // values per pack block (one warp)
constexpr int kChunk = 32;                     // consecutive values per lane
constexpr int kWarpsPerCta = 8;                // encode: one block per warp
constexpr int kTileBlocks = kWarpsPerCta;      // blocks per encode tile
constexpr int kTileVals = kTileBlocks * kBlockLen;
constexpr int kRun = MNWCUDA_DITHER_RUN;       // blocks per decode task
constexpr int kMaxWidth = 25;                  // residual bits: |delta| < 2^24
constexpr int kStageWords = 1024 + 32;         // per-warp staging: 1024 words + skew pads

enum : int { kErrDepth = 1, kErrCapacity = 2, kErrPayload = 3, kErrRange = 4, kErrTimeout = 5 };

// One field of the batch as the kernels see it (built on the host per call).
struct DevField {
    int64_t ioOff;       // first float of the field in the float buffer
    int64_t fixedBase;   // sum of fixed bytes (header+table) of earlier fields
    int64_t firstBlock;  // sum of 3*nb of earlier fields
    int64_t firstTile;   // encode tiles of earlier fields
    int64_t firstTask;   // decode tasks of earlier fields
    int32_t n;           // particles
    int32_t nb;          // blocks per component
    uint32_t fieldCode;
    float delta;
    float width;
    int32_t tilesPerComp;
    uint64_t seed;
};

// Quantisation parameters of one field (PositionQuantization, types.h:82-87).
struct DevQuant {
    float X0[3];
    float X1[3];
    float maxDiff;
    uint32_t depth;
    float width;   // decode: box width from the payload header (Posn)
    double scale;  // encode: RN_double(2^depth / maxDiff), see bin_index_fast
};

// Per-block result of the pack kernel before the table is finalised.
struct TmpBlock {
    uint32_t first;
    int32_t minDelta;
    uint32_t width;
    uint32_t size16;    // packed size, 16-byte units
    uint64_t prefix16;  // payload offset over the whole batch, 16-byte units
};

// ---- order-preserving float <-> uint32 (for atomicMin/atomicMax) ----
__device__ __forceinline__ uint32_t ord_from_float(float v) {
    uint32_t b = __float_as_uint(v);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float float_from_ord(uint32_t o) {
    uint32_t b = o ^ ((o >> 31) ? 0x80000000u : 0xffffffffu);
    return __uint_as_float(b);
}

// util_UndoPeriodic's per-element rule (src/util.c:106-110), branch-free:
// sub, two compares (the second masked by the first, as in the reference's
// else-if), two predicated adds
__device__ __forceinline__ float unwrap(float v, float anchor, float L, float half) {
    float r = v;
    asm("{\n\t"
        ".reg .pred up, dn;\n\t"
        ".reg .f32 rel, nh;\n\t"
        "sub.rn.f32 rel, %1, %2;\n\t"
        "neg.f32 nh, %4;\n\t"
        "setp.ge.f32 up, rel, %4;\n\t"
        "setp.lt.and.f32 dn, rel, nh, !up;\n\t"
        "@up sub.rn.f32 %0, %1, %3;\n\t"
        "@dn add.rn.f32 %0, %1, %3;\n\t"
        "}"
        : "+f"(r) : "f"(v), "f"(anchor), "f"(L), "f"(half));
    return r;
}

// util_Periodic's per-element rule (src/util.c:76-82)
__device__ __forceinline__ float wrap(float v, float L) {
    float r = v;  // branch-free: two compares (the second masked by the first), two predicated adds
    asm("{\n\t"
        ".reg .pred up, dn;\n\t"
        "setp.ge.f32 up, %1, %2;\n\t"
        "setp.lt.and.f32 dn, %1, 0f00000000, !up;\n\t"
        "@up sub.rn.f32 %0, %1, %2;\n\t"
        "@dn add.rn.f32 %0, %1, %2;\n\t"
        "}"
        : "+f"(r) : "f"(v), "f"(L));
    return r;
}

// util_UniformBinIndex's per-element rule (src/util.c:184-191): IEEE division,
// clamp, scale, truncate. No contraction anywhere.
__device__ __forceinline__ uint32_t bin_index(float v, float x0, float dx, float bins,
                                              uint32_t top) {
    float t = __fdiv_rn(__fsub_rn(v, x0), dx);
    if (t < 0.0f) return 0u;
    if (t >= 1.0f) return top;
    return __float2uint_rz(__fmul_rn(t, bins));
}

// The same index without the IEEE float division. util.c:184-191 computes
//   t = RN_f(a/dx);  q = t<0 ? 0 : t>=1 ? top : trunc(t * 2^depth)
// with a = x-x0. Multiplying by 2^depth is exact, so q = trunc(RN_f(a/dx)*2^depth)
// = trunc(RN_f(a * 2^depth/dx)) outside the subnormal range (where q = 0 on both
// sides). The quotient of two 24-bit floats is never closer than 2^-49
// (relative) to a rounding midpoint of the 24-bit format, and
// RN_d(a * RN_d(2^depth/dx)) is within 2^-52 of the true scaled quotient, so
// rounding that double to float gives exactly RN_f(a/dx)*2^depth. cvt.rzi.u32
// saturates negatives (t<0) and NaN (0/0 on a constant field) to 0, and the
// t>=1 clamp becomes an integer min. DESIGN.md "quantiser" has the proof.
__device__ __forceinline__ uint32_t bin_index_fast(float v, float x0, double scale, uint32_t top) {
    float a = __fsub_rn(v, x0);
    float s = __double2float_rn(__dmul_rn((double)a, scale));
    return min(__float2uint_rz(s), top);
}

// xorshift128+ (55,14,36), src/rand.c:80-90; returns s0+s1 of the old state
__device__ __forceinline__ uint64_t xs128p_next(uint64_t &s0, uint64_t &s1) {
    uint64_t a = s0, b = s1;
    uint64_t sum = a + b;
    b ^= a;
    s0 = ((a << 55) | (a >> 9)) ^ b ^ (b << 14);
    s1 = (b << 36) | (b >> 28);
    return sum;
}

// The same step on 32-bit halves (state a = ah:al, b = bh:bl), returning the low
// word of s0+s1 (rand_Float keeps only its low 24 bits, src/rand.c:60-64). The
// rotations are written as funnel shifts -- 9 ALU-pipe instructions (4 LOP3, 5 SHF)
// where the compiler's expansion of the 64-bit form takes 11 -- and b<<14's low
// word and the sum go to the FMA pipe. (Moving rotr(a,9) there too, as the halves
// of the 64-bit product x * 2^23, costs more issue slots than it frees: measured
// slower.)
// one LOP3; written out because ptxas otherwise re-associates a ^ (b ^ c) with the
// xor that produced b and spends an extra instruction
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ uint32_t xs128p_next32(uint32_t &al, uint32_t &ah, uint32_t &bl, uint32_t &bh) {
    const uint32_t sum = al + bl;
    const uint32_t xl = al ^ bl, xh = ah ^ bh;      // b ^= a
    const uint32_t rl = __funnelshift_r(al, ah, 9);  // rotl(a, 55) = rotr(a, 9)
    const uint32_t rh = __funnelshift_r(ah, al, 9);
    al = xor3(rl, xl, xl << 14);
    ah = xor3(rh, xh, __funnelshift_l(xl, xh, 14));
    bl = __funnelshift_l(xl, xh, 4);  // rotl(b, 36) = rotl(swap halves, 4)
    bh = __funnelshift_l(xh, xl, 4);
    return sum;
}

// staging layout for packed words: one pad word after every 32, so that lanes
// walking rows of `w` words and lanes walking consecutive words both spread
// over the banks
__device__ __forceinline__ int skew(int m) { return m + (m >> 5); }

// staging layout for 32x32 float transposes: 16-byte chunk j of row r sits at
// chunk position j ^ (r & 7) (the 128-byte XOR swizzle)
__device__ __forceinline__ int swz_chunk(int row, int j) { return row * 8 + (j ^ (row & 7)); }

// 128-bit shared-memory accesses by 32-bit shared address (the swizzled layouts
// are walked by XOR on the address)
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_f4(uint32_t a, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void sts_u4(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// tile/task -> field lookup: largest f with first[f] <= idx
template <typename GetFirst>
__device__ __forceinline__ int find_field(int nFields, int64_t idx, GetFirst first) {
    int lo = 0, hi = nFields - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (first(mid) <= idx) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ uint64_t ld_volatile_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_volatile_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ float4 ld_stream_f4(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ uint4 ld_stream_u4(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

}  // namespace mnw

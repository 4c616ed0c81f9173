// minnow_kernels.cuh -- device structs and helpers shared by the kernels of
// libminnow_cuda (sm_100a only). No CPU fallback lives anywhere in this tree.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "minnow_cuda.h"

namespace mnw {

constexpr int kBlockLen = MNWCUDA_BLOCK_LEN;   // values per pack block (one warp)
constexpr int kChunk = 32;                     // consecutive values per lane
constexpr int kRun = MNWCUDA_DITHER_RUN;       // blocks per decode task
constexpr int kMaxWidth = 25;                  // residual bits: |delta| < 2^24
constexpr int kStageWords = 1024 + 32;         // per-warp staging: 1024 words + skew pads

enum : int { kErrDepth = 1, kErrCapacity = 2, kErrPayload = 3, kErrRange = 4, kErrTimeout = 5 };

// One field of the batch as the kernels see it (built on the host per call).
struct DevField {
    int64_t ioOff;       // first float of the field in the float buffer
    int64_t fixedBase;   // sum of fixed bytes (header+table) of earlier fields
    int64_t firstBlock;  // sum of 3*nb of earlier fields
    int64_t firstTask;   // decode tasks of earlier fields
    int32_t n;           // particles
    int32_t nb;          // blocks per component
    uint32_t fieldCode;
    float delta;
    float width;
    uint64_t seed;
};

// Quantisation parameters of one field (PositionQuantization, types.h:82-87).
struct DevQuant {
    float X0[3];
    float X1[3];
    float maxDiff;
    uint32_t depth;
    float width;   // decode: box width from the payload header (Posn)
    double scale;  // encode: RN_double(2^depth / maxDiff), see quantise32 (enc_math.cuh)
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

__device__ __forceinline__ uint4 ld_stream_u4(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

}  // namespace mnw

// enc_math.cuh -- the per-value arithmetic of the encoder: periodic unwrap as
// thresholds, exact division-free quantisation, delta + width of one block.
#pragma once

#include "enc_common.cuh"

namespace mnw {

// ---------------------------------------------------------------------------
// util_UndoPeriodic (src/util.c:106-110) tests rel = x - x[0] against +-L/2.
// RN(x - anchor) is monotone in x, so each test is a threshold on x itself:
// first_ge(anchor, c) = the smallest float v with RN(v - anchor) >= c.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float next_up(float x) {
    uint32_t b = __float_as_uint(x);
    if (x == 0.0f) return __uint_as_float(1u);
    return __uint_as_float(x > 0.0f ? b + 1u : b - 1u);
}
__device__ __forceinline__ float next_down(float x) {
    uint32_t b = __float_as_uint(x);
    if (x == 0.0f) return __uint_as_float(0x80000001u);
    return __uint_as_float(x > 0.0f ? b - 1u : b + 1u);
}
__device__ __forceinline__ float first_ge(float anchor, float c) {
    float v = __fadd_rn(anchor, c);
    for (int i = 0; i < 8; i++) {
        float d = next_down(v);
        if (__fsub_rn(d, anchor) >= c) v = d; else break;
    }
    for (int i = 0; i < 8; i++) {
        if (!(__fsub_rn(v, anchor) >= c)) v = next_up(v); else break;
    }
    return v;
}

struct Unwrap {
    float tHi, tLo, L;  // x >= tHi: x - L;  else x < tLo: x + L
};
__device__ __forceinline__ Unwrap make_unwrap(float anchor, float L) {
    Unwrap u;
    const float half = __fmul_rn(L, 0.5f);
    u.tHi = first_ge(anchor, half);
    u.tLo = first_ge(anchor, -half);
    u.L = L;
    return u;
}
__device__ __forceinline__ float unwrap_t(float v, const Unwrap &u) {
    float r = v;
    asm("{\n\t"
        ".reg .pred up, dn;\n\t"
        "setp.ge.f32 up, %1, %2;\n\t"
        "setp.lt.and.f32 dn, %1, %3, !up;\n\t"
        "@up sub.rn.f32 %0, %1, %4;\n\t"
        "@dn add.rn.f32 %0, %1, %4;\n\t"
        "}"
        : "+f"(r) : "f"(v), "f"(u.tHi), "f"(u.tLo), "f"(u.L));
    return r;
}

// util_UniformBinIndex (src/util.c:184-191) on 32 values per lane, without the
// IEEE division: t = RN_f(a/dx) with a = x - x0, q = t<0 ? 0 : t>=1 ? top :
// trunc(t * 2^depth). Multiplying by 2^depth is exact, so q = trunc(RN_f(a *
// 2^depth/dx)) outside the subnormal range (where q = 0 on both sides). The
// quotient of two 24-bit floats is never closer than 2^-49 (relative) to a
// rounding midpoint of the 24-bit format, and RN_d(a * RN_d(2^depth/dx)) is within
// 2^-52 of the true scaled quotient, so rounding that double to float gives exactly
// RN_f(a/dx) * 2^depth (DESIGN.md section 5 has the proof). With
// kBiased (depth <= 23) the truncation is an add in round-toward-zero mode:
// for 0 <= s < 2^23, RZ(s + 2^23) = 2^23 + floor(s) exactly, i.e. the result's
// bit pattern is kQBias + q; the bias cancels in the deltas. The clamp t >= 1
// becomes fmin against the largest index (an integer, so trunc commutes).
// s >= 0 always: x0 is the minimum of the very same unwrapped values.

// Chunk c = kk * 32 + lane (elements 4c .. 4c+3) of a full block that does not start on a
// 16-byte boundary (particle counts that are not multiples of four: the TMA cannot fetch
// it), with the widest loads its address allows -- two 8-byte ones, or 4 + 8 + 4.
__device__ __forceinline__ float4 ldg_chunk_unaligned(const float *src, int c, bool by8) {
    const float *s = src + c * 4;
    float4 v;
    if (by8) {
        const float2 a = __ldg(reinterpret_cast<const float2 *>(s)), b = __ldg(reinterpret_cast<const float2 *>(s + 2));
        v = make_float4(a.x, a.y, b.x, b.y);
    } else {
        const float a = __ldg(s);
        const float2 b = __ldg(reinterpret_cast<const float2 *>(s + 1));
        const float d = __ldg(s + 3);
        v = make_float4(a, b.x, b.y, d);
    }
    return v;
}

// ... into the stage, in the swizzled layout the TMA would have produced (out of line: this is
// the path of unaligned batches only, and the kernel is sensitive to the size of its hot code)
__device__ __noinline__ void stage_full_unaligned(const float *src, float *stage, int lane) {
    const bool by8 = ((uintptr_t)src & 7) == 0;
    float4 *st = reinterpret_cast<float4 *>(stage);
#pragma unroll
    for (int kk = 0; kk < 8; kk++) {
        const int c = kk * 32 + lane;
        st[swz_chunk(c >> 3, c & 7)] = ldg_chunk_unaligned(src, c, by8);
    }
    __syncwarp();
}

// ... or straight into the extrema (A work: nothing else needs the values)
__device__ __noinline__ void minmax_full_unaligned(const float *src, int lane, float *lohi) {
    const bool by8 = ((uintptr_t)src & 7) == 0;
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
#pragma unroll
    for (int kk = 0; kk < 8; kk++) {
        const float4 v = ldg_chunk_unaligned(src, kk * 32 + lane, by8);
        lo = fminf(fminf(lo, v.x), fminf(v.y, fminf(v.z, v.w)));
        hi = fmaxf(fmaxf(hi, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
    }
    lohi[0] = lo;
    lohi[1] = hi;
}

// the lane's 32 consecutive values of a staged block (128-byte swizzle), as bit patterns
__device__ __forceinline__ void load_block32(const float *stage, int lane, uint32_t (&q)[kChunk]) {
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint4 e4 = reinterpret_cast<const uint4 *>(stage)[swz_chunk(lane, j)];
        q[j * 4 + 0] = e4.x;
        q[j * 4 + 1] = e4.y;
        q[j * 4 + 2] = e4.z;
        q[j * 4 + 3] = e4.w;
    }
}

// util_UndoPeriodic on the lane's 32 values, in place. unwrap_t is monotone within
// each of its three classes (x >= tHi: x - L; x < tLo: x + L; between: x), so a lane
// whose raw extrema fall into one class shifts all of its values alike -- and in
// nearly every block no lane needs a shift at all: one three-input min/max per two
// values instead of five instructions per value. Lanes that straddle a threshold
// (or hold a NaN) take the value-by-value form; the warp votes so that the common
// case has no divergent code.
__device__ __forceinline__ void unwrap32(const Unwrap &uw, uint32_t (&q)[kChunk]) {
    float lo = __uint_as_float(q[0]), hi = lo;
#pragma unroll
    for (int i = 1; i < kChunk - 1; i += 2) {
        lo = fminf(lo, fminf(__uint_as_float(q[i]), __uint_as_float(q[i + 1])));
        hi = fmaxf(hi, fmaxf(__uint_as_float(q[i]), __uint_as_float(q[i + 1])));
    }
    lo = fminf(lo, __uint_as_float(q[kChunk - 1]));
    hi = fmaxf(hi, __uint_as_float(q[kChunk - 1]));
    const bool split = (lo < uw.tHi && !(hi < uw.tHi)) || (lo < uw.tLo && !(hi < uw.tLo)) || !(lo <= hi);
    const float shift = !(lo < uw.tHi) ? -uw.L : lo < uw.tLo ? uw.L : 0.0f;
    if (__any_sync(kFull, split)) {
#pragma unroll
        for (int i = 0; i < kChunk; i++) q[i] = __float_as_uint(unwrap_t(__uint_as_float(q[i]), uw));
    } else if (__any_sync(kFull, shift != 0.0f)) {
        // RN(x + (-L)) = RN(x - L); a lane of the middle class adds +0 (x itself, and -0 -> +0
        // quantises like -0: both give index 0 or the same difference to x0)
#pragma unroll
        for (int i = 0; i < kChunk; i++) q[i] = __float_as_uint(__fadd_rn(__uint_as_float(q[i]), shift));
    }
}

// in place: q holds the raw floats' bit patterns on entry, the indices on return
// (One body for positions and velocities -- the unwrap is a pass of its own, unwrap32 --
// so that the consumers of an SM, which work on both kinds at once, share its code: the
// kernel is sensitive to the size of the code its warps are spread over.)
template <bool kBiased>
__device__ __forceinline__ void quantise32(float x0, double scale, uint32_t top, uint32_t (&q)[kChunk]) {
    const float topf = (float)top;
#pragma unroll
    for (int i = 0; i < kChunk; i++) {
        const float u = __uint_as_float(q[i]);
        const float a = __fsub_rn(u, x0);
#if defined(MNW_ABLATE) && (MNW_ABLATE & 1)
        const float s = __fmul_rn(a, (float)scale);  // timing experiment only: not the exact quotient
#else
        const float s = __double2float_rn(__dmul_rn((double)a, scale));
#endif
        if (kBiased) q[i] = __float_as_uint(__fadd_rz(fminf(s, topf), 8388608.0f));
        else q[i] = min(__float2uint_rz(s), top);
    }
}

struct BlockStats {
    uint32_t first;
    int dmin, w;
};

// delta along the block, min/max over j in [1, len), width; leaves the
// residuals r[j] = d[j] - dmin in q (r[0] = 0, zeros beyond len)
// kBiasedQ: q holds the bit patterns of the floats 2^23 + index (quantise32's
// kBiased form). Their differences are exact in FP32, so for a full block the
// first pass runs on the FMA pipe with three-input min/max (FMNMX3; there is no
// integer counterpart on sm_100): 64 instructions instead of 96.
template <bool kBiasedQ>
__device__ __forceinline__ BlockStats delta_width(uint32_t (&q)[kChunk], int len, int lane, int *error) {
    const uint32_t prevLast = __shfl_up_sync(kFull, q[kChunk - 1], 1);
    BlockStats bs;
    bs.first = __shfl_sync(kFull, q[0], 0);
    int dmin, dmax;
    if (kBiasedQ && len == kBlockLen) {
        float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
        if (lane) lo = hi = __fsub_rn(__uint_as_float(q[0]), __uint_as_float(prevLast));
        {
            const float d = __fsub_rn(__uint_as_float(q[1]), __uint_as_float(q[0]));
            lo = fminf(lo, d);
            hi = fmaxf(hi, d);
        }
#pragma unroll
        for (int i = 2; i < kChunk; i += 2) {
            const float d0 = __fsub_rn(__uint_as_float(q[i]), __uint_as_float(q[i - 1]));
            const float d1 = __fsub_rn(__uint_as_float(q[i + 1]), __uint_as_float(q[i]));
            lo = fminf(lo, fminf(d0, d1));
            hi = fmaxf(hi, fmaxf(d0, d1));
        }
        dmin = __float2int_rn(lo);
        dmax = __float2int_rn(hi);
    } else if (len == kBlockLen) {
        const int d0 = (int)(q[0] - prevLast);
        dmin = lane ? d0 : INT32_MAX;
        dmax = lane ? d0 : INT32_MIN;
#pragma unroll
        for (int i = 1; i < kChunk; i++) {
            const int d = (int)(q[i] - q[i - 1]);
            dmin = min(dmin, d);
            dmax = max(dmax, d);
        }
    } else {
        dmin = INT32_MAX;
        dmax = INT32_MIN;
        uint32_t prev = prevLast;
#pragma unroll
        for (int i = 0; i < kChunk; i++) {
            const int j = lane * kChunk + i;
            const int d = (int)(q[i] - prev);
            prev = q[i];
            if (j >= 1 && j < len) { dmin = min(dmin, d); dmax = max(dmax, d); }
        }
    }
    dmin = __reduce_min_sync(kFull, dmin);
    dmax = __reduce_max_sync(kFull, dmax);
    if (len <= 1) { dmin = 0; dmax = 0; }
    int w = 32 - __clz((uint32_t)(dmax - dmin));
    if (w > kMaxWidth) {  // only reachable with out-of-range integers at the CFunc boundary
        if (lane == 0) atomicCAS(error, 0, kErrRange);
        w = 0;
    }
    bs.dmin = dmin;
    bs.w = w;
    if (w > 0) {
        if (len == kBlockLen) {
#pragma unroll
            for (int i = kChunk - 1; i >= 1; i--) q[i] = q[i] - q[i - 1] - (uint32_t)dmin;
            q[0] = lane ? q[0] - prevLast - (uint32_t)dmin : 0u;
        } else {
#pragma unroll
            for (int i = kChunk - 1; i >= 1; i--) {
                const uint32_t r = q[i] - q[i - 1] - (uint32_t)dmin;
                q[i] = (lane * kChunk + i < len) ? r : 0u;
            }
            q[0] = (lane && lane * kChunk < len) ? q[0] - prevLast - (uint32_t)dmin : 0u;
        }
    }
    return bs;
}

}  // namespace mnw

// bitpack.cuh -- lane-local uniform-width pack / unpack of 32 values, with the
// width a compile-time constant so that every shift amount and word boundary
// folds into an immediate. The bit layout is util_U32UniformPack's
// (/root/reference/src/util.c:333-352): value i of the block occupies bits
// [W*i, W*i+W) of a little-endian stream of 32-bit words, LSB first. A lane owns
// values [32*lane, 32*lane+32) of a 1024-value block, i.e. exactly W whole
// words [W*lane, W*lane+W), so no bits are shared between lanes.
#pragma once

#include "minnow_kernels.cuh"

namespace mnw {

// r[i] < 2^W is the caller's invariant (residual <= spread < 2^W).
template <int W>
__device__ __forceinline__ void pack_lane(const uint32_t (&r)[kChunk], uint32_t *st, int lane) {
    if (W == 0) return;
    const int base = lane * W;
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < kChunk; i++) {
        const int off = (i * W) & 31;
        acc += r[i] << off;  // disjoint bits: the sum is the OR, and shift-and-add is one instruction
        if (off + W >= 32) {
            st[skew(base + ((i * W) >> 5))] = acc;
            acc = (off + W > 32) ? (r[i] >> ((32 - off) & 31)) : 0u;
        }
    }
}

// inverse (util_U32UndoUniformPack, src/util.c:383-403); words beyond the
// block's payload read whatever the stage holds (callers never use those values).
// kOr is OR-ed into every value: with kFloatMagic (W <= 23) the result is the bit
// pattern of the float 2^23 + r, which costs nothing (it folds into the mask's
// LOP3) and lets the decoder run its prefix sums on the FP32 pipe.
constexpr uint32_t kFloatMagic = 0x4b000000u;  // 2^23 as a float

template <int W, uint32_t kOr = 0u>
__device__ __forceinline__ void unpack_lane(uint32_t (&r)[kChunk], const uint32_t *st, int lane) {
    if (W == 0) {
#pragma unroll
        for (int i = 0; i < kChunk; i++) r[i] = kOr;
        return;
    }
    const int base = lane * W;
    const uint32_t mask = W >= 32 ? 0xffffffffu : ((1u << (W & 31)) - 1u);
    uint32_t cur = st[skew(base)];
#pragma unroll
    for (int i = 0; i < kChunk; i++) {
        const int off = (i * W) & 31;
        if (off + W < 32) {
            r[i] = ((cur >> off) & mask) | kOr;
        } else if (off + W == 32) {
            r[i] = (cur >> off) | kOr;
            if (i + 1 < kChunk) cur = st[skew(base + (((i + 1) * W) >> 5))];
        } else {
            uint32_t nxt = st[skew(base + ((i * W) >> 5) + 1)];
            r[i] = (__funnelshift_r(cur, nxt, off) & mask) | kOr;
            cur = nxt;
        }
    }
}

#define MNW_WIDTH_CASES(X)                                                                        \
    X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13) X(14) X(15) X(16) \
    X(17) X(18) X(19) X(20) X(21) X(22) X(23) X(24) X(25)

// runtime width -> the specialised body; widths above kMaxWidth never reach here
__device__ __forceinline__ void pack_dispatch(int w, const uint32_t (&r)[kChunk], uint32_t *st, int lane) {
    switch (w) {
#define X(W) case W: pack_lane<W>(r, st, lane); break;
        MNW_WIDTH_CASES(X)
#undef X
        default: break;
    }
}

__device__ __forceinline__ void unpack_dispatch(int w, uint32_t (&r)[kChunk], const uint32_t *st, int lane) {
    switch (w) {
#define X(W) case W: unpack_lane<W>(r, st, lane); break;
        MNW_WIDTH_CASES(X)
#undef X
        default: break;
    }
}

#define MNW_FLOAT_WIDTH_CASES(X)                                                                  \
    X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13) X(14) X(15) X(16) \
    X(17) X(18) X(19) X(20) X(21) X(22) X(23)

// widths 0..23 only: r[i] = bits of the float 2^23 + residual
__device__ __forceinline__ void unpack_dispatch_magic(int w, uint32_t (&r)[kChunk], const uint32_t *st,
                                                      int lane) {
    switch (w) {
#define X(W) case W: unpack_lane<W, kFloatMagic>(r, st, lane); break;
        MNW_FLOAT_WIDTH_CASES(X)
#undef X
        default: break;
    }
}

// widths 24 and 25 for the decoder's float path (plain integers out)
__device__ __forceinline__ void unpack_dispatch_wide(int w, uint32_t (&r)[kChunk], const uint32_t *st,
                                                     int lane) {
    if (w == 24) unpack_lane<24>(r, st, lane);
    else unpack_lane<25>(r, st, lane);
}

}  // namespace mnw

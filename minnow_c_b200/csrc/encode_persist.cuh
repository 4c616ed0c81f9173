// encode_persist.cuh -- the encoder: one persistent kernel (k_encode) + k_finalize.
//
// k_encode runs one CTA of 16 independent warps per SM for the whole batch.
// A warp draws tickets from one global counter; a ticket is a tile of four
// 1024-value blocks and one of two kinds of work, interleaved in ticket order:
//   A(g)  util_UndoPeriodic + util_MinMax of tile g (src/util.c:97-113, 27-46),
//         straight from 128-bit loads in registers. The warp that finishes a
//         field's last A tile runs deltaToDepth (src/quant.c:186-196, 684-693)
//         and publishes the field's quantisation parameters.
//   B(g)  `lag` tiles later (one field + a margin, so that the parameters are
//         known and the second read of the input hits the L2 instead of HBM):
//         util_UniformBinIndex + delta + per-block width + util_U32UniformPack
//         (src/util.c:173-196, 311-355), block by block, the next block's
//         input arriving by cp.async while the current one is computed. The
//         packed blocks stay in the warp's shared memory; only the tile's size
//         is published.
// A scan warp (CTA 0, warp 0) turns the tile sizes into exclusive prefixes in
// tile order. The warp that packed tile g copies it to its final position two
// B tiles later, when the prefix has long arrived, and writes the blocks'
// table entries at the same time. Nothing is staged through global memory.
//
// There is no CTA-wide synchronisation. A ticket only ever waits on smaller
// tickets (its field's A tiles; the prefix, which needs the sizes of earlier
// tiles only), and tickets are only held by running warps, so the waits cannot
// deadlock.
#pragma once

#include "encode_common.cuh"

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

// ---------------------------------------------------------------------------
// Scan warp: tile sizes -> exclusive prefixes, 512 tiles per poll. The flag
// words carry their own payload and nobody reads anything else through them,
// so relaxed accesses are enough.
// ---------------------------------------------------------------------------
constexpr int kScanPerLane = 16;

__device__ __forceinline__ void scan_tiles(const EncodeParams &p, int lane) {
    uint64_t running = 0;
    uint32_t a[kScanPerLane], nxt[kScanPerLane];
#pragma unroll
    for (int k = 0; k < kScanPerLane; k++) {
        const int64_t i = (int64_t)lane * kScanPerLane + k;
        nxt[k] = i < p.nTiles ? ld_relaxed_u32(p.tileAgg + i) : kAggReady;
    }
    for (int64_t base = 0; base < p.nTiles; base += 32 * kScanPerLane) {
        const int64_t mine = base + (int64_t)lane * kScanPerLane;
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) a[k] = nxt[k];
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) {  // next window's first poll is in flight meanwhile
            const int64_t i = mine + 32 * kScanPerLane + k;
            nxt[k] = i < p.nTiles ? ld_relaxed_u32(p.tileAgg + i) : kAggReady;
        }
        // publish as far as the sizes are known, poll the rest again
        int done = 0;  // entries of this lane already published
        uint64_t carry = running;
        uint32_t spins = 0;
        while (true) {
            // number of leading ready entries of the window, over the whole warp
            int myReady = kScanPerLane;  // index of this lane's first entry that is not ready
#pragma unroll
            for (int k = kScanPerLane - 1; k >= 0; k--)
                if (!(a[k] & kAggReady)) myReady = k;
            const bool full = myReady == kScanPerLane;
            const unsigned fullMask = __ballot_sync(0xffffffffu, full);
            const int firstPartial = __ffs(~fullMask) - 1;  // -1: every lane is complete
            uint32_t local = 0;
#pragma unroll
            for (int k = 0; k < kScanPerLane; k++) local += (k >= done && k < myReady) ? (a[k] & ~kAggReady) : 0u;
            // lanes after the first incomplete one must not publish yet
            const bool active = firstPartial < 0 || lane <= firstPartial;
            uint32_t contrib = active ? local : 0u;
            uint32_t incl = contrib;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += up;
            }
            if (active) {
                uint64_t excl = carry + (incl - contrib);
#pragma unroll
                for (int k = 0; k < kScanPerLane; k++) {
                    if (k >= done && k < myReady) {
                        if (mine + k < p.nTiles) st_relaxed_u64(p.tilePrefix + mine + k, kPrefixReady | excl);
                        excl += a[k] & ~kAggReady;
                    }
                }
                done = myReady;
            }
            carry += __shfl_sync(0xffffffffu, incl, 31);
            if (firstPartial < 0) break;
            if (++spins > kSpinLimit) {
                if (lane == 0) spin_timeout(p, 3, base);
                return;
            }
            if (spins > 64) __nanosleep(100);
#pragma unroll
            for (int k = 0; k < kScanPerLane; k++)
                if (!(a[k] & kAggReady)) a[k] = ld_relaxed_u32(p.tileAgg + mine + k);
        }
        running = carry;
    }
    if (lane == 0) st_relaxed_u64(p.tilePrefix + p.nTiles, kPrefixReady | running);
}

}  // namespace mnw

#include "encode_tile.cuh"

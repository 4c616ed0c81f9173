// enc_kernel.cuh -- k_encode (persistent, one CTA per SM: a TMA producer warp,
// 14 consumer warps, and the scan warp in CTA 0) and k_finalize.
//
//   A work   util_UndoPeriodic + util_MinMax of one block (src/util.c:97-113,
//            27-46) straight out of the ring; a consumer keeps the running
//            min/max of consecutive blocks of one component in registers and
//            flushes them with two atomics when the component changes -- and
//            before it ever waits for anything. Whoever completes a field runs
//            deltaToDepth (src/quant.c:186-196, 684-693) and publishes the
//            field's quantisation parameters.
//   B work   util_UniformBinIndex + delta + width + util_U32UniformPack of one
//            block (src/util.c:173-196, 311-355). The size is published at once;
//            the packed words go through the block's own ring stage into one of
//            the consumer's scratch slots in global memory (L2 resident).
//   scan     one warp turns the sizes into exclusive prefixes, in block order.
//   place    a few blocks later the same consumer copies the block from its
//            slot to its final position and writes its table entry.
// Work only ever waits for work with a smaller ticket, and every waiting warp
// has flushed what others may wait for, so the waits cannot deadlock.
#pragma once

#include "enc_math.cuh"
#include "enc_ring.cuh"

namespace mnw {

// ---------------------------------------------------------------------------
// maxDiff and deltaToDepth of one field (src/quant.c:186-196, 684-693), run by
// a single thread once every min/max of the field is final.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void compute_quant(const EncodeParams &p, int f) {
    const DevField fd = p.fields[f];
    DevQuant q;
    float md = 0.0f;
    for (int c = 0; c < 3; c++) {
        float lo = 0.0f, hi = 0.0f;
        if (fd.n > 0) {
            lo = float_from_ord(__ldcg(p.minmax + (f * 3 + c) * 2));
            hi = float_from_ord(~__ldcg(p.minmax + (f * 3 + c) * 2 + 1));
        }
        if (lo == 0.0f) lo = 0.0f;  // -0 -> +0 (zero canonicalisation)
        if (hi == 0.0f) hi = 0.0f;
        q.X0[c] = lo;
        q.X1[c] = hi;
        float d = __fsub_rn(hi, lo);
        if (md < d) md = d;
    }
    q.maxDiff = md;
    float range = __fsub_rn(__fadd_rn(q.X0[0], md), q.X0[0]);
    int depth = 0;
    for (; depth <= 24; depth++)
        if (__fmul_rn(fd.delta, (float)(1 << depth)) > range) break;
    if (depth > 24) {
        if (fd.n > 0) atomicCAS(p.error, 0, kErrDepth);
        depth = 24;
    }
    q.depth = (uint32_t)depth;
    q.width = fd.width;
    // a constant field has maxDiff = 0: util_UniformBinIndex's 0/0 falls through
    // both clamps and converts to 0 on the host; scale 0 gives the same zeros
    q.scale = md > 0.0f ? __ddiv_rn((double)(1u << depth), (double)md) : 0.0;
    p.quant[f] = q;
}

// one thread per field: empty fields have no blocks to trigger compute_quant
__global__ void k_depth_empty(EncodeParams p) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nFields) return;
    if (p.fields[f].n == 0) compute_quant(p, f);
}

// ---------------------------------------------------------------------------
// Scan warp: block sizes -> exclusive prefixes, 1024 blocks per poll. The flag
// words carry their own payload and nobody reads anything else through them,
// so relaxed accesses are enough.
// ---------------------------------------------------------------------------
constexpr int kScanPerLane = 32;

__device__ __forceinline__ void scan_blocks(const EncodeParams &p, int lane) {
    uint64_t running = 0;
    for (int64_t base = 0; base < p.nBlocks; base += 32 * kScanPerLane) {
        const int64_t mine = base + (int64_t)lane * kScanPerLane;
        uint32_t a[kScanPerLane];
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) a[k] = mine + k < p.nBlocks ? ld_relaxed_u32(p.blkAgg + mine + k) : kAggReady;
        uint32_t spins = 0;
        while (true) {
            bool all = true;
#pragma unroll
            for (int k = 0; k < kScanPerLane; k++) all = all && (a[k] & kAggReady);
            if (__all_sync(kFull, all)) break;
            if (++spins > kSpinLimit) {
                if (lane == 0) spin_timeout(p.error, 3, base);
                return;
            }
            if (spins > 16) __nanosleep(100);
#pragma unroll
            for (int k = 0; k < kScanPerLane; k++)
                if (!(a[k] & kAggReady)) a[k] = ld_relaxed_u32(p.blkAgg + mine + k);
        }
        uint32_t local = 0;
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) local += a[k] & ~kAggReady;
        uint32_t incl = local;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t up = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += up;
        }
        uint64_t excl = running + (incl - local);
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) {
            if (mine + k < p.nBlocks) {
                asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p.blkPrefix + mine + k),
                             "l"(kPrefixReady | excl) : "memory");
            }
            excl += a[k] & ~kAggReady;
        }
        running += __shfl_sync(kFull, incl, 31);
    }
    if (lane == 0) st_release_u64(p.blkPrefix + p.nBlocks, kPrefixReady | running);
}

// ---------------------------------------------------------------------------
// A work: running min/max of one component in registers
// ---------------------------------------------------------------------------
struct MinMaxAcc {
    int f, comp;      // -1: nothing pending
    uint32_t count;   // blocks folded in
    float lo, hi;
    Unwrap uw;
    bool isPos;
};

__device__ __forceinline__ void acc_flush(const EncodeParams &p, MinMaxAcc &acc, int lane) {
    if (acc.f < 0) return;
    const uint32_t olo = __reduce_min_sync(kFull, ord_from_float(acc.lo));
    const uint32_t ohi = __reduce_min_sync(kFull, ~ord_from_float(acc.hi));  // max kept inverted (0xff memset)
    if (lane == 0) {
        uint32_t *mm = p.minmax + ((int64_t)acc.f * 3 + acc.comp) * 2;
        atomicMin(mm, olo);
        atomicMin(mm + 1, ohi);
        __threadfence();
        const uint32_t done = atomicAdd(p.aDone + acc.f, acc.count);
        if (done + acc.count == 3u * (uint32_t)p.fields[acc.f].nb) {
            __threadfence();
            compute_quant(p, acc.f);
            __threadfence();
            st_release_u32(p.ready + acc.f, 1u);
        }
    }
    acc.f = -1;
    acc.count = 0;
}

template <bool kPos>
__device__ __forceinline__ void minmax_stage(const float4 *st, int lane, const Unwrap &uw, float &lo, float &hi) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float4 v = st[k * 32 + lane];  // any order of the chunks will do: the swizzle is ignored
        float e0 = v.x, e1 = v.y, e2 = v.z, e3 = v.w;
        if (kPos) {
            e0 = unwrap_t(e0, uw);
            e1 = unwrap_t(e1, uw);
            e2 = unwrap_t(e2, uw);
            e3 = unwrap_t(e3, uw);
        }
        lo = fminf(fminf(lo, e0), fminf(e1, fminf(e2, e3)));
        hi = fmaxf(fmaxf(hi, e0), fmaxf(e1, fmaxf(e2, e3)));
    }
}

__device__ __forceinline__ void a_block(const EncodeParams &p, const WorkDesc &d, const float *stage, MinMaxAcc &acc,
                                        int lane) {
    if (acc.f != d.f || acc.comp != d.comp) {
        acc_flush(p, acc, lane);
        const DevField *fdp = p.fields + d.f;
        acc.f = d.f;
        acc.comp = d.comp;
        acc.lo = __int_as_float(0x7f800000);
        acc.hi = __int_as_float(0xff800000);
        acc.isPos = fdp->fieldCode == field_Posn;
        acc.uw.tHi = acc.uw.tLo = acc.uw.L = 0.0f;
        if (acc.isPos) acc.uw = make_unwrap(__ldg(p.in + fdp->ioOff + (int64_t)d.comp * fdp->n), fdp->width);
    }
    if (!d.slow) {
        const float4 *st = reinterpret_cast<const float4 *>(stage);
        if (acc.isPos) minmax_stage<true>(st, lane, acc.uw, acc.lo, acc.hi);
        else minmax_stage<false>(st, lane, acc.uw, acc.lo, acc.hi);
    } else {
        const DevField *fdp = p.fields + d.f;
        const float *src = p.in + fdp->ioOff + (int64_t)d.comp * fdp->n + (int64_t)d.b * kBlockLen;
        for (int i = lane; i < d.len; i += 32) {
            float u = __ldg(src + i);
            if (acc.isPos) u = unwrap_t(u, acc.uw);
            acc.lo = fminf(acc.lo, u);
            acc.hi = fmaxf(acc.hi, u);
        }
    }
    acc.count++;
}

}  // namespace mnw

#include "enc_consume.cuh"

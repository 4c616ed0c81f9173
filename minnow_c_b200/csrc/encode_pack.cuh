// encode_pack.cuh -- the B work of k_encode, phase 1: quantise + delta + width
// + pack of one 1024-value block per worker warp. Included by
// encode_kernels.cuh; the placement half and the kernel are in encode_tail.cuh.
#pragma once

#include "encode_kernels.cuh"

namespace mnw {

constexpr int kScanPerLane = 4;     // scan warp: tile aggregates per lane and poll (128 per round trip)
constexpr int kSlotBytes = 3328;    // scratch bytes per block: 25 bits x 1024 values, 16-byte padded, +128
constexpr uint32_t kAggReady = 0x80000000u;
constexpr uint64_t kPrefixReady = 1ull << 63;

// ---------------------------------------------------------------------------
// B tile (one warp per 1024-value block, no CTA-wide synchronisation):
//   load    8 x 128-bit coalesced loads per lane (issued before the wait for
//           the field's parameters), transposed through swizzled shared memory
//           so that each lane owns 32 consecutive values
//   quant   util_UniformBinIndex (src/util.c:184-191) via bin_index_fast
//   delta   d[j] = q[j]-q[j-1] (the recurrence of util_U8DeltaEncode,
//           src/util.c:283-295, on the quantised integers), min/max by warp
//           redux, width = bit length of (max-min)
//   pack    util_U32UniformPack (src/util.c:333-352), width-specialised: each
//           lane's 32 residuals are exactly `width` whole words
//   park    the packed words go to the block's slot of an L2-resident scratch
//           ring (128-bit coalesced stores); the warp that arrives last
//           publishes the tile's size for the scan warp. The copy to the final
//           position is the C tile's job, `lag2` tiles later, when the offset
//           is known -- nobody waits for anybody in steady state.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void place_block(const EncodeParams &p, const TileJob &job);

template <bool kFromQ>
__device__ __forceinline__ void quantize_pack(const EncodeParams &p, const uint64_t *qIn, const TileJob &job,
                                              const TileJob &placeJob, uint32_t *st, uint32_t *sSize,
                                              uint32_t *sArrive) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = job.f, comp = job.comp;
    const DevField *fdp = p.fields + f;
    const int nb = fdp->nb, n = fdp->n;
    const int b = job.t * kTileBlocks + warp;
    const bool active = b < nb;
    const int len = active ? min(kBlockLen, n - b * kBlockLen) : 0;
    const int64_t ioOff = fdp->ioOff;
    const int64_t elem = (int64_t)comp * n + (int64_t)b * kBlockLen;
    const float *src = p.in + ioOff + elem;
    const bool fast = !kFromQ && len == kBlockLen && ((uintptr_t)src & 15) == 0;

    uint32_t size16 = 0;

    // ---- input loads first: they do not depend on the field's parameters ----
    float4 v[8];
    if (fast) {
        const float4 *s4 = reinterpret_cast<const float4 *>(src);
        const uint64_t drop = l2_policy_drop();  // last read: the line is dead afterwards
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = ld_hint_f4(s4 + k * 32 + lane, drop);
    }
    // ---- while they are in flight: place the tile parked lag2 tiles ago ----
    if (placeJob.kind == 3) place_block(p, placeJob);

    uint32_t q[kChunk];
    if (active) {
        if (!kFromQ) {
            if (p.lag >= 0) {
                while (ld_acquire_u32(p.ready + f) == 0u) __nanosleep(64);
            }
            float *sf = reinterpret_cast<float *>(st);
            if (fast) {
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    int c = k * 32 + lane;
                    reinterpret_cast<float4 *>(sf)[swz_chunk(c >> 3, c & 7)] = v[k];
                }
            } else {
                float fill = __ldg(src);
#pragma unroll 4
                for (int k = 0; k < 32; k++) {
                    int e = k * 32 + lane;
                    float x = e < len ? __ldg(src + e) : fill;
                    sf[swz_chunk(e >> 5, (e >> 2) & 7) * 4 + (e & 3)] = x;
                }
            }
            __syncwarp();
            const DevQuant *qp = p.quant + f;  // written during this launch: read through L2
            const float x0 = __ldcg(&qp->X0[comp]);
            const double scale = __ldcg(&qp->scale);
            const uint32_t top = (1u << __ldcg(&qp->depth)) - 1u;
            const bool isPos = fdp->fieldCode == field_Posn;
            const float L = fdp->width, half = __fmul_rn(L, 0.5f);
            const float anchor = isPos ? __ldg(p.in + ioOff + (int64_t)comp * n) : 0.0f;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                float4 e = reinterpret_cast<const float4 *>(sf)[swz_chunk(lane, j)];
                if (isPos) {
                    e.x = unwrap(e.x, anchor, L, half);
                    e.y = unwrap(e.y, anchor, L, half);
                    e.z = unwrap(e.z, anchor, L, half);
                    e.w = unwrap(e.w, anchor, L, half);
                }
                q[j * 4 + 0] = bin_index_fast(e.x, x0, scale, top);
                q[j * 4 + 1] = bin_index_fast(e.y, x0, scale, top);
                q[j * 4 + 2] = bin_index_fast(e.z, x0, scale, top);
                q[j * 4 + 3] = bin_index_fast(e.w, x0, scale, top);
            }
            __syncwarp();
        } else {
            // literal CFunc boundary: quantised uint64 values in (types.h:117-122)
            const uint64_t *qsrc = qIn + ioOff + elem;
#pragma unroll 4
            for (int k = 0; k < 32; k++) {
                int e = k * 32 + lane;
                st[skew(e)] = e < len ? (uint32_t)qsrc[e] : 0u;
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < kChunk; i++) q[i] = st[skew(lane * kChunk + i)];
            __syncwarp();
        }

        // ---- delta, min/max over j in [1, len) ----
        const uint32_t prevLast = __shfl_up_sync(0xffffffffu, q[kChunk - 1], 1);
        const uint32_t first = __shfl_sync(0xffffffffu, q[0], 0);
        int dmin, dmax;
        if (len == kBlockLen) {
            int d0 = (int)(q[0] - prevLast);
            dmin = lane ? d0 : INT32_MAX;
            dmax = lane ? d0 : INT32_MIN;
#pragma unroll
            for (int i = 1; i < kChunk; i++) {
                int d = (int)(q[i] - q[i - 1]);
                dmin = min(dmin, d);
                dmax = max(dmax, d);
            }
        } else {
            dmin = INT32_MAX;
            dmax = INT32_MIN;
            uint32_t prev = prevLast;
#pragma unroll
            for (int i = 0; i < kChunk; i++) {
                int j = lane * kChunk + i;
                int d = (int)(q[i] - prev);
                prev = q[i];
                if (j >= 1 && j < len) { dmin = min(dmin, d); dmax = max(dmax, d); }
            }
        }
        dmin = __reduce_min_sync(0xffffffffu, dmin);
        dmax = __reduce_max_sync(0xffffffffu, dmax);
        if (len <= 1) { dmin = 0; dmax = 0; }
        const uint32_t spread = (uint32_t)(dmax - dmin);
        int w = 32 - __clz(spread);
        if (w > kMaxWidth) {  // only reachable with out-of-range integers at the CFunc boundary
            if (lane == 0) atomicCAS(p.error, 0, kErrRange);
            w = 0;
        }
        if (p.debugMask & 8) w = 7;
        const int nWords = (len * w + 31) >> 5;
        size16 = (uint32_t)((nWords + 3) >> 2);

        // residuals in place (q is dead as such from here on)
        if (w > 0) {
            if (len == kBlockLen) {
#pragma unroll
                for (int i = kChunk - 1; i >= 1; i--) q[i] = q[i] - q[i - 1] - (uint32_t)dmin;
                q[0] = lane ? q[0] - prevLast - (uint32_t)dmin : 0u;
            } else {
#pragma unroll
                for (int i = kChunk - 1; i >= 1; i--) {
                    uint32_t r = q[i] - q[i - 1] - (uint32_t)dmin;
                    q[i] = (lane * kChunk + i < len) ? r : 0u;
                }
                q[0] = (lane && lane * kChunk < len) ? q[0] - prevLast - (uint32_t)dmin : 0u;
            }
        }

        // ---- pack: lane's w words at [lane*w, lane*w+w) of the stage buffer ----
        if (w > 0 && !(p.debugMask & 4)) pack_dispatch(w, q, st, lane);
        __syncwarp();

        // ---- park the packed block in its scratch slot ----
        const int64_t ringPos = job.tile & (p.ringTiles - 1);  // ringTiles is a power of two
        if (lane == 0) {  // the slot's previous tenants must have been copied out
            const uint32_t need = (uint32_t)(job.tile >> p.ringShift) * kWarpsPerCta;
            while (ld_acquire_u32(p.slotDone + ringPos) < need) __nanosleep(64);
        }
        __syncwarp();
        uint4 *slot = reinterpret_cast<uint4 *>(p.scratch + (ringPos * kWarpsPerCta + warp) * kSlotBytes);
        if (!(p.debugMask & 2)) {
            for (int c = lane; c < (int)size16; c += 32) {
                const int m = c * 4;
                uint4 o;
                o.x = m + 0 < nWords ? st[skew(m + 0)] : 0u;
                o.y = m + 1 < nWords ? st[skew(m + 1)] : 0u;
                o.z = m + 2 < nWords ? st[skew(m + 2)] : 0u;
                o.w = m + 3 < nWords ? st[skew(m + 3)] : 0u;
                slot[c] = o;
            }
        }
        if (lane == 0) {
            TmpBlock tb;
            tb.first = first;
            tb.minDelta = dmin;
            tb.width = (uint32_t)w;
            tb.size16 = size16;
            tb.prefix16 = 0;
            p.tmp[fdp->firstBlock + (int64_t)comp * nb + b] = tb;
        }
        __threadfence();  // slot + table entry before the tile's size is published
    }

    // ---- the last of the 8 warps to arrive publishes the tile's size ----
    if (lane == 0) {
        ((volatile uint32_t *)sSize)[warp] = size16;
        __threadfence_block();
        if (atomicAdd(sArrive, 1u) == kWarpsPerCta - 1) {
            __threadfence();  // acquire the other warps' slots and table entries before the release below
            uint32_t agg = 0;
#pragma unroll
            for (int i = 0; i < kWarpsPerCta; i++) agg += ((volatile uint32_t *)sSize)[i];
            st_release_u32(p.tileAgg + job.tile, kAggReady | agg);
        }
    }
}

}  // namespace mnw

#include "encode_tail.cuh"

// encode_btile.cuh -- B tiles of k_encode: quantise + delta + width + pack of
// up to four 1024-value blocks by one warp, into the warp's shared memory.
// Included by encode_tile.cuh.
#pragma once

#include "encode_tile.cuh"

namespace mnw {

// 8 x 16-byte cp.async per lane: the block lands in the XOR-swizzled stage so
// that lane l can afterwards read its 32 consecutive values conflict-free
__device__ __forceinline__ void issue_block_load(float *stage, const float *src, int lane, uint64_t pol) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int c = k * 32 + lane;
        cp_async16(stage + swz_chunk(c >> 3, c & 7) * 4, src + c * 4, pol);
    }
}

// util_UniformBinIndex (src/util.c:184-191) on 32 values per lane. See
// bin_index_fast (minnow_kernels.cuh) for the division-free scaling. With
// kBiased (depth <= 23) the truncation is an add in round-toward-zero mode:
// for 0 <= s < 2^23, RZ(s + 2^23) = 2^23 + floor(s) exactly, i.e. the result's
// bit pattern is kQBias + q; the bias cancels in the deltas. The clamp t >= 1
// becomes fmin against the largest index (an integer, so trunc commutes).
// s >= 0 always: x0 is the minimum of the very same unwrapped values.
template <bool kPos, bool kBiased>
__device__ __forceinline__ void quantise32(const float *stage, int lane, const Unwrap &uw, float x0,
                                           double scale, uint32_t top, uint32_t (&q)[kChunk]) {
    const float topf = (float)top;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const float4 e4 = reinterpret_cast<const float4 *>(stage)[swz_chunk(lane, j)];
        float e[4] = {e4.x, e4.y, e4.z, e4.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            float u = e[i];
            if (kPos) u = unwrap_t(u, uw);
            const float a = __fsub_rn(u, x0);
            const float s = __double2float_rn(__dmul_rn((double)a, scale));
            if (kBiased) q[j * 4 + i] = __float_as_uint(__fadd_rz(fminf(s, topf), 8388608.0f));
            else q[j * 4 + i] = min(__float2uint_rz(s), top);
        }
    }
}

struct BlockStats {
    uint32_t first;
    int dmin, w;
};

// delta along the block, min/max over j in [1, len), width; leaves the
// residuals r[j] = d[j] - dmin in q (r[0] = 0, zeros beyond len)
__device__ __forceinline__ BlockStats delta_width(uint32_t (&q)[kChunk], int len, int lane, int *error) {
    const uint32_t prevLast = __shfl_up_sync(kFull, q[kChunk - 1], 1);
    BlockStats bs;
    bs.first = __shfl_sync(kFull, q[0], 0);
    int dmin, dmax;
    if (len == kBlockLen) {
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

// ---------------------------------------------------------------------------
// B tile. Returns with the tile's size published and the packed blocks pending
// in half h of the warp's shared memory (blocks that did not fit: in the warp's
// overflow slots in global memory).
// ---------------------------------------------------------------------------
template <bool kFromQ>
__device__ __forceinline__ void b_tile(const EncodeParams &p, int64_t tile, const WarpSmem &ws, int h, TileRef &pd,
                                       int lane) {
    const TileRef r = make_ref(p, tile);
    const DevField *fdp = p.fields + r.f;
    const int n = fdp->n;
    const int64_t compOff = fdp->ioOff + (int64_t)r.comp * n;
    const float *base = p.in + compOff;
    const bool isPos = fdp->fieldCode == field_Posn;
    const uint64_t drop = l2_policy_drop();  // last read: the line is dead afterwards
    uint32_t *buf = ws.half(h), *rec = ws.rec(h);
    uint32_t *ovf = ws.ovf(h);

    auto block_src = [&](int j) { return base + (int64_t)(r.b0 + j) * kBlockLen; };
    auto block_len = [&](int j) { return min(kBlockLen, n - (r.b0 + j) * kBlockLen); };
    auto block_fast = [&](int j) {
        return !kFromQ && block_len(j) == kBlockLen && ((uintptr_t)block_src(j) & 15) == 0;
    };

    // first block's input is requested before anything that can wait
    if (block_fast(0)) issue_block_load(ws.in, block_src(0), lane, drop);

    float x0 = 0.0f;
    double scale = 0.0;
    uint32_t top = 0, depth = 0;
    Unwrap uw;
    uw.tHi = uw.tLo = uw.L = 0.0f;
    if (!kFromQ) {
        if (p.lag >= 0) {
            uint32_t spins = 0;
            while (ld_acquire_u32(p.ready + r.f) == 0u) {
                __nanosleep(100);
                if (++spins > kSpinLimit) { spin_timeout(p, 1, tile); break; }
            }
        }
        const DevQuant *qp = p.quant + r.f;  // written during this launch: read through L2
        x0 = __ldcg(&qp->X0[r.comp]);
        scale = __ldcg(&qp->scale);
        depth = __ldcg(&qp->depth);
        top = (1u << depth) - 1u;
        if (isPos) uw = make_unwrap(__ldg(base), fdp->width);
    }
    const bool biased = !kFromQ && depth <= 23;

    int outPos = 0;
    uint32_t agg = 0;
    for (int j = 0; j < r.nBlk; j++) {
        const int len = block_len(j);
        uint32_t q[kChunk];
        if (kFromQ) {
            // literal CFunc boundary: quantised uint64 values in (types.h:117-122)
            const uint64_t *qsrc = p.qIn + compOff + (int64_t)(r.b0 + j) * kBlockLen;
#pragma unroll
            for (int i = 0; i < kChunk; i++) {
                const int e = lane * kChunk + i;
                q[i] = e < len ? (uint32_t)qsrc[e] : 0u;
            }
        } else {
            if (block_fast(j)) {
                cp_async_wait_all();
                __syncwarp();
            } else {
                // ragged or unaligned block: plain loads into the same swizzled layout
                const float *src = block_src(j);
                const float fill = __ldg(src);
#pragma unroll 4
                for (int k = 0; k < 32; k++) {
                    const int e = k * 32 + lane;
                    const float x = e < len ? __ldg(src + e) : fill;
                    ws.in[swz_chunk(e >> 5, (e >> 2) & 7) * 4 + (e & 3)] = x;
                }
                __syncwarp();
            }
            if (isPos) {
                if (biased) quantise32<true, true>(ws.in, lane, uw, x0, scale, top, q);
                else quantise32<true, false>(ws.in, lane, uw, x0, scale, top, q);
            } else {
                if (biased) quantise32<false, true>(ws.in, lane, uw, x0, scale, top, q);
                else quantise32<false, false>(ws.in, lane, uw, x0, scale, top, q);
            }
            __syncwarp();  // the stage is free again
        }

        BlockStats bs = delta_width(q, len, lane, p.error);
        if (biased) bs.first -= kQBias;
        const int nWords = (len * bs.w + 31) >> 5;
        const uint32_t size16 = (uint32_t)((nWords + 3) >> 2);
        const int need = 33 * bs.w + 1;  // skewed words of a full block of this width (ragged blocks write zeros up to it)
        const bool spill = outPos + need > kEncHalfWords;
        const bool prefetch = !kFromQ && j + 1 < r.nBlk && block_fast(j + 1);
        uint32_t pos;
        if (!spill) {
            if (prefetch) issue_block_load(ws.in, block_src(j + 1), lane, drop);
            if (bs.w > 0) pack_dispatch(bs.w, q, buf + outPos, lane);
            pos = (uint32_t)outPos;
            outPos += need;
        } else {
            // the tile's half is full (wide residuals): pack through the input stage into
            // the warp's overflow slot in global memory; it is placed from there
            uint32_t *tmp = reinterpret_cast<uint32_t *>(ws.in);
            if (bs.w > 0) pack_dispatch(bs.w, q, tmp, lane);
            __syncwarp();
            uint4 *slot = reinterpret_cast<uint4 *>(ovf + (size_t)j * kEncSlotWords);
            for (int c = lane; c < (int)size16; c += 32) {
                const int m = c * 4;
                uint4 o;
                o.x = tmp[skew(m)];
                o.y = m + 1 < nWords ? tmp[skew(m + 1)] : 0u;
                o.z = m + 2 < nWords ? tmp[skew(m + 2)] : 0u;
                o.w = m + 3 < nWords ? tmp[skew(m + 3)] : 0u;
                slot[c] = o;
            }
            __syncwarp();
            if (prefetch) issue_block_load(ws.in, block_src(j + 1), lane, drop);
            pos = kPosSpilled;
        }
        if (lane == 0) {
            rec[j * 8 + 0] = bs.first;
            rec[j * 8 + 1] = (uint32_t)bs.dmin;
            rec[j * 8 + 2] = (uint32_t)bs.w;
            rec[j * 8 + 3] = size16;
            rec[j * 8 + 4] = pos;
            rec[j * 8 + 5] = (uint32_t)nWords;
        }
        agg += size16;
    }
    __syncwarp();
    if (lane == 0) st_relaxed_u32(p.tileAgg + tile, kAggReady | agg);
    pd = r;
}

// ---------------------------------------------------------------------------
// k_encode
// ---------------------------------------------------------------------------
template <bool kFromQ>
__global__ void __launch_bounds__(kEncWarps * 32, 1) k_encode(EncodeParams p) {
    extern __shared__ __align__(16) uint8_t encSmem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (blockIdx.x == 0 && warp == 0) {
        scan_tiles(p, lane);
        return;
    }
    WarpSmem ws;
    uint8_t *mine = encSmem + (size_t)warp * kEncWarpSmem;
    ws.in = reinterpret_cast<float *>(mine);
    ws.half0 = reinterpret_cast<uint32_t *>(mine + kEncInWords * 4);
    ws.rec0 = ws.half0 + 2 * kEncHalfWords;
    ws.ovf0 = p.overflow + ((size_t)blockIdx.x * kEncWarps + warp) * 2 * kEncTileBlocks * kEncSlotWords;

    TileRef pend[2];
    pend[0].valid = pend[1].valid = false;
    pend[0].tile = pend[1].tile = -1;
    int bCount = 0;
    uint32_t tk = 0;
    if (lane == 0) tk = atomicAdd(p.ticket, 1u);
    while (true) {
        const int64_t t = (int64_t)__shfl_sync(kFull, tk, 0);
        if (t >= p.nTickets) break;
        if (lane == 0) tk = atomicAdd(p.ticket, 1u);  // the next ticket is on its way meanwhile
        int64_t a = -1, b = -1;
        if (kFromQ || p.lag < 0) b = t;
        else if (t & 1) b = (t >> 1) - p.lag;
        else a = t >> 1;
        if (a >= 0 && a < p.nTiles) {
            if (!kFromQ) a_tile(p, a, lane);
        } else if (b >= 0 && b < p.nTiles) {
            const int h = bCount & 1;
            if (h == 0) {
                place_pending(p, pend[0], ws, 0, lane);
                b_tile<kFromQ>(p, b, ws, 0, pend[0], lane);
            } else {
                place_pending(p, pend[1], ws, 1, lane);
                b_tile<kFromQ>(p, b, ws, 1, pend[1], lane);
            }
            bCount++;
        }
    }
    // drain: the older tile first
    if (pend[0].valid && pend[1].valid && pend[1].tile < pend[0].tile) {
        place_pending(p, pend[1], ws, 1, lane);
        place_pending(p, pend[0], ws, 0, lane);
    } else {
        place_pending(p, pend[0], ws, 0, lane);
        place_pending(p, pend[1], ws, 1, lane);
    }
}

// ---------------------------------------------------------------------------
// k_finalize: one thread per field writes the 64-byte header and the field's
// byte offset (the block tables were written at placement).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_finalize(EncodeParams p, int64_t fixedTotal) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f > p.nFields) return;
    if (f == p.nFields) {
        p.offsets[f] = (uint64_t)fixedTotal + (p.tilePrefix[p.nTiles] & ~kPrefixReady) * 16;
        return;
    }
    const DevField fd = p.fields[f];
    const uint64_t p0 = p.tilePrefix[fd.firstTile] & ~kPrefixReady;
    const int64_t regionStart = fd.fixedBase + (int64_t)(p0 * 16);
    p.offsets[f] = (uint64_t)regionStart;
    if (regionStart + 64 + (int64_t)fd.nb * 48 > p.outCap) { atomicCAS(p.error, 0, kErrCapacity); return; }
    const DevQuant qp = p.quant[f];
    __align__(16) MnwDiffHeader h;
    h.Magic = algo_Diff;
    h.Version = MNWCUDA_ALGO_VERSION;
    h.FieldCode = fd.fieldCode;
    h.ParticleLen = fd.n;
    h.Depth = (uint8_t)qp.depth;
    h.Pad[0] = h.Pad[1] = h.Pad[2] = 0;
    h.Width = fd.fieldCode == field_Posn ? fd.width : 0.0f;
    for (int c = 0; c < 3; c++) { h.X0[c] = qp.X0[c]; h.X1[c] = qp.X1[c]; }
    h.BlockLen = kBlockLen;
    h.DitherRun = MNWCUDA_DITHER_RUN;
    h.DitherLanes = MNWCUDA_DITHER_LANES;
    h.DitherSeed = fd.seed;
    const uint4 *src = reinterpret_cast<const uint4 *>(&h);
    uint4 *dst = reinterpret_cast<uint4 *>(p.out + regionStart);
    for (int i = 0; i < 4; i++) dst[i] = src[i];
}

}  // namespace mnw

// encode_kernels.cuh -- K1 unwrap+min/max, K1b depth, K2 quantize+delta+pack
// with a single-pass (decoupled look-back) scan of block sizes, K2f table and
// header finalisation. Replaces, per field, position()/velocity() of
// /root/reference/src/quant.c:161-289 and the (stub) CFunc behind
// src/funcs.c:71. HBM-bound integer/float streaming: no tensor cores.
#pragma once

#include "minnow_kernels.cuh"

namespace mnw {

struct EncodeParams {
    const float *in;
    const DevField *fields;
    int nFields;
    int uniformTiles;   // >0: every field has this many tiles (fast lookup)
    uint32_t *minmax;   // [nFields][3][2] ordered-uint min, ~max
    DevQuant *quant;    // [nFields]
    uint64_t *tileState;  // [nTiles] look-back state
    uint32_t *ticket;     // dynamic tile counter
    TmpBlock *tmp;        // [nBlocksTotal]
    uint8_t *out;
    int64_t outCap;
    uint64_t *offsets;  // [nFields+1]
    int *error;
};

__device__ __forceinline__ void locate_tile(const EncodeParams &p, int64_t tile, int &f, int &comp,
                                            int &tileInComp) {
    if (p.uniformTiles > 0) {
        f = (int)(tile / p.uniformTiles);
    } else {
        const DevField *fields = p.fields;
        f = find_field(p.nFields, tile, [fields](int i) { return fields[i].firstTile; });
    }
    int local = (int)(tile - p.fields[f].firstTile);
    int tpc = p.fields[f].tilesPerComp;
    comp = local / tpc;
    tileInComp = local - comp * tpc;
}

// ---------------------------------------------------------------------------
// K1: util_UndoPeriodic (Posn only, never stored) + util_MinMax per component
// (src/util.c:97-113, 27-46). One CTA per 8192-value tile, float4 streaming
// loads, warp-shuffle then shared-memory reduction, one atomicMin/atomicMax
// pair per CTA on an order-preserving integer image of the floats.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_unwrap_minmax(EncodeParams p) {
    int f, comp, t;
    locate_tile(p, blockIdx.x, f, comp, t);
    const DevField fd = p.fields[f];
    const float *base = p.in + fd.ioOff + (int64_t)comp * fd.n;
    const int64_t start = (int64_t)t * kTileVals;
    const int cnt = (int)min((int64_t)kTileVals, (int64_t)fd.n - start);
    const bool isPos = fd.fieldCode == field_Posn;
    const float L = fd.width, half = __fmul_rn(L, 0.5f);
    const float anchor = isPos ? __ldg(base) : 0.0f;

    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    const float *src = base + start;
    if (cnt == kTileVals && ((uintptr_t)src & 15) == 0) {
        const float4 *s4 = reinterpret_cast<const float4 *>(src);
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = ld_stream_f4(s4 + k * 256 + threadIdx.x);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            float e[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
            for (int i = 0; i < 4; i++) {
                float u = isPos ? unwrap(e[i], anchor, L, half) : e[i];
                lo = fminf(lo, u);
                hi = fmaxf(hi, u);
            }
        }
    } else {
        for (int i = threadIdx.x; i < cnt; i += 256) {
            float u = __ldg(src + i);
            if (isPos) u = unwrap(u, anchor, L, half);
            lo = fminf(lo, u);
            hi = fmaxf(hi, u);
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
    }
    __shared__ float sLo[8], sHi[8];
    if ((threadIdx.x & 31) == 0) { sLo[threadIdx.x >> 5] = lo; sHi[threadIdx.x >> 5] = hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 1; w < 8; w++) { lo = fminf(lo, sLo[w]); hi = fmaxf(hi, sHi[w]); }
        uint32_t *mm = p.minmax + ((int64_t)f * 3 + comp) * 2;
        atomicMin(mm, ord_from_float(lo));
        atomicMin(mm + 1, ~ord_from_float(hi));  // max kept inverted: both slots memset to 0xff
    }
}

// ---------------------------------------------------------------------------
// K1b: maxDiff and deltaToDepth per field (src/quant.c:186-196, 684-693).
// ---------------------------------------------------------------------------
__global__ void k_depth(EncodeParams p) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nFields) return;
    const DevField fd = p.fields[f];
    DevQuant q;
    float md = 0.0f;
    for (int c = 0; c < 3; c++) {
        float lo = 0.0f, hi = 0.0f;
        if (fd.n > 0) {
            lo = float_from_ord(p.minmax[(f * 3 + c) * 2]);
            hi = float_from_ord(~p.minmax[(f * 3 + c) * 2 + 1]);
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
    p.quant[f] = q;
}

// ---------------------------------------------------------------------------
// K2: one warp per 1024-value block, 8 blocks per CTA.
//   load    8 x 128-bit coalesced loads per lane, transposed through swizzled
//           shared memory so that each lane owns 32 consecutive values
//   quant   util_UniformBinIndex (src/util.c:184-191)
//   delta   d[j] = q[j]-q[j-1] (the recurrence of util_U8DeltaEncode,
//           src/util.c:283-295, on the quantised integers), min/max by warp
//           redux, width = bit length of (max-min)
//   pack    util_U32UniformPack (src/util.c:333-352): each lane's 32 residuals
//           are exactly `width` whole words
//   place   CTA aggregate of block sizes, decoupled look-back over earlier
//           tiles, coalesced word stores
// ---------------------------------------------------------------------------
constexpr uint64_t kFlagAgg = 1ull << 62, kFlagPrefix = 2ull << 62, kValMask = (1ull << 62) - 1;

template <bool kFromQ>
__global__ void __launch_bounds__(256) k_quantize_delta_pack(EncodeParams p, const uint64_t *qIn) {
    __shared__ __align__(16) uint32_t stage[kWarpsPerCta][kStageWords];
    __shared__ uint32_t sSize[kWarpsPerCta];
    __shared__ uint64_t sExcl;
    __shared__ uint32_t sTile;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) sTile = atomicAdd(p.ticket, 1u);
    __syncthreads();
    const int64_t tile = sTile;

    int f, comp, t;
    locate_tile(p, tile, f, comp, t);
    const DevField fd = p.fields[f];
    const DevQuant qp = p.quant[f];
    const int b = t * kTileBlocks + warp;
    const bool active = b < fd.nb;
    const int len = active ? min(kBlockLen, fd.n - b * kBlockLen) : 0;
    uint32_t *st = stage[warp];

    uint32_t q[kChunk];
    if (active) {
        const int64_t elem = (int64_t)comp * fd.n + (int64_t)b * kBlockLen;
        if (!kFromQ) {
            const float *src = p.in + fd.ioOff + elem;
            float *sf = reinterpret_cast<float *>(st);
            if (len == kBlockLen && ((uintptr_t)src & 15) == 0) {
                const float4 *s4 = reinterpret_cast<const float4 *>(src);
                float4 v[8];
#pragma unroll
                for (int k = 0; k < 8; k++) v[k] = ld_stream_f4(s4 + k * 32 + lane);
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
                    float v = e < len ? __ldg(src + e) : fill;
                    sf[swz_chunk(e >> 5, (e >> 2) & 7) * 4 + (e & 3)] = v;
                }
            }
            __syncwarp();
            const bool isPos = fd.fieldCode == field_Posn;
            const float L = fd.width, half = __fmul_rn(L, 0.5f);
            const float anchor = isPos ? __ldg(p.in + fd.ioOff + (int64_t)comp * fd.n) : 0.0f;
            const float x0 = qp.X0[comp], dx = qp.maxDiff;
            const float bins = (float)(1u << qp.depth);
            const uint32_t top = (1u << qp.depth) - 1u;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                float4 v = reinterpret_cast<const float4 *>(sf)[swz_chunk(lane, j)];
                float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    float u = isPos ? unwrap(e[i], anchor, L, half) : e[i];
                    q[j * 4 + i] = bin_index(u, x0, dx, bins, top);
                }
            }
            __syncwarp();
        } else {
            // literal CFunc boundary: quantised uint64 values in (types.h:117-122)
            const uint64_t *src = qIn + fd.ioOff + elem;
#pragma unroll 4
            for (int k = 0; k < 32; k++) {
                int e = k * 32 + lane;
                st[skew(e)] = e < len ? (uint32_t)src[e] : 0u;
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < kChunk; i++) q[i] = st[skew(lane * kChunk + i)];
            __syncwarp();
        }
    } else {
#pragma unroll
        for (int i = 0; i < kChunk; i++) q[i] = 0;
    }

    // ---- delta, min/max over j in [1, len) ----
    uint32_t prevLast = __shfl_up_sync(0xffffffffu, q[kChunk - 1], 1);
    int dmin = INT32_MAX, dmax = INT32_MIN;
    {
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
    const int w = 32 - __clz(spread);
    const int nWords = (len * w + 31) >> 5;
    const uint32_t size16 = (uint32_t)((nWords + 3) >> 2);
    const uint32_t first = __shfl_sync(0xffffffffu, q[0], 0);
    if (lane == 0) sSize[warp] = size16;
    __syncthreads();

    // ---- publish the tile aggregate as early as possible ----
    if (threadIdx.x == 0) {
        uint64_t agg = 0;
#pragma unroll
        for (int i = 0; i < kWarpsPerCta; i++) agg += sSize[i];
        if (tile == 0) {
            st_volatile_u64(p.tileState, kFlagPrefix | agg);
            sExcl = 0;
        } else {
            st_volatile_u64(p.tileState + tile, kFlagAgg | agg);
        }
    }

    // ---- pack: lane's 32 residuals -> w words at words [lane*w, lane*w+w) ----
    if (w > 0) {
        uint32_t prev = prevLast;
        uint64_t acc = 0;
        int fill = 0, wi = lane * w;
#pragma unroll
        for (int i = 0; i < kChunk; i++) {
            int j = lane * kChunk + i;
            uint32_t r = (uint32_t)((int)(q[i] - prev) - dmin);
            prev = q[i];
            if (!(j >= 1 && j < len)) r = 0;
            acc |= (uint64_t)r << fill;
            fill += w;
            if (fill >= 32) {
                st[skew(wi)] = (uint32_t)acc;
                wi++;
                acc >>= 32;
                fill -= 32;
            }
        }
    }

    // ---- look-back (warp 0) ----
    if (warp == 0 && tile > 0) {
        uint64_t excl = 0;
        int64_t idx = tile - 1;
        while (true) {
            int64_t me = idx - lane;
            uint64_t s = kFlagPrefix;  // virtual predecessor of tile 0: prefix 0
            if (me >= 0) {
                do { s = ld_volatile_u64(p.tileState + me); } while ((s >> 62) == 0);
            }
            uint32_t isPrefix = __ballot_sync(0xffffffffu, (s >> 62) == 2);
            uint64_t v = s & kValMask;
            if (isPrefix) {
                int firstP = __ffs(isPrefix) - 1;
                if (lane > firstP) v = 0;
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
            excl += v;
            if (isPrefix) break;
            idx -= 32;
        }
        if (lane == 0) {
            uint64_t agg = 0;
#pragma unroll
            for (int i = 0; i < kWarpsPerCta; i++) agg += sSize[i];
            st_volatile_u64(p.tileState + tile, kFlagPrefix | (excl + agg));
            sExcl = excl;
        }
    }
    __syncthreads();

    if (!active) return;
    uint64_t prefix16 = sExcl;
    for (int i = 0; i < warp; i++) prefix16 += sSize[i];

    const int64_t payloadBase = fd.fixedBase + 64 + (int64_t)fd.nb * 48;  // + 16*prefix16
    const int64_t dstByte = payloadBase + (int64_t)(prefix16 * 16);
    if (dstByte + (int64_t)size16 * 16 > p.outCap) {
        if (lane == 0) atomicCAS(p.error, 0, kErrCapacity);
        return;
    }
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.out + dstByte);
    const int padded = (int)size16 * 4;
    for (int m = lane; m < padded; m += 32) dst[m] = m < nWords ? st[skew(m)] : 0u;
    if (lane == 0) {
        TmpBlock tb;
        tb.first = first;
        tb.minDelta = dmin;
        tb.width = (uint32_t)w;
        tb.pad = 0;
        tb.prefix16 = prefix16;
        p.tmp[fd.firstBlock + (int64_t)comp * fd.nb + b] = tb;
    }
}

// ---------------------------------------------------------------------------
// K2f: after K2 every tile state holds its inclusive prefix. One thread per
// block turns TmpBlock into the wire table entry (offset relative to the
// field's payload base); one thread per field writes the 64-byte header and
// the field's byte offset.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t prefix_before_tile(const EncodeParams &p, int64_t tile) {
    return tile == 0 ? 0 : (p.tileState[tile - 1] & kValMask);
}

__global__ void __launch_bounds__(256) k_finalize(EncodeParams p, int64_t nBlocksTotal, int64_t nTilesTotal,
                                                   int uniformBlocks) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g < nBlocksTotal) {
        int f;
        if (uniformBlocks > 0) {
            f = (int)(g / uniformBlocks);
        } else {
            const DevField *fields = p.fields;
            f = find_field(p.nFields, g, [fields](int i) { return fields[i].firstBlock; });
        }
        const DevField fd = p.fields[f];
        uint64_t p0 = prefix_before_tile(p, fd.firstTile);
        int64_t regionStart = fd.fixedBase + (int64_t)(p0 * 16);
        TmpBlock tb = p.tmp[g];
        int64_t entry = regionStart + 64 + (g - fd.firstBlock) * 16;
        if (entry + 16 <= p.outCap) {
            uint4 e = make_uint4(tb.first, (uint32_t)tb.minDelta, tb.width, (uint32_t)(tb.prefix16 - p0));
            *reinterpret_cast<uint4 *>(p.out + entry) = e;
        }
    }
    if (g <= p.nFields) {
        int f = (int)g;
        if (f == p.nFields) {
            const DevField last = p.fields[f - 1];
            int64_t fixedEnd = last.fixedBase + 64 + (int64_t)last.nb * 48;
            p.offsets[f] = (uint64_t)fixedEnd + prefix_before_tile(p, nTilesTotal) * 16;
            return;
        }
        const DevField fd = p.fields[f];
        uint64_t p0 = prefix_before_tile(p, fd.firstTile);
        int64_t regionStart = fd.fixedBase + (int64_t)(p0 * 16);
        p.offsets[f] = (uint64_t)regionStart;
        if (regionStart + 64 > p.outCap) { atomicCAS(p.error, 0, kErrCapacity); return; }
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
}

}  // namespace mnw

// encode_tail.cuh -- placement half of the encoder: the scan warp, the C tile
// (copy of parked blocks to their final offsets), the kernel k_encode and
// k_finalize. Included by encode_pack.cuh.
#pragma once

#include "encode_pack.cuh"

namespace mnw {

__device__ __forceinline__ uint64_t ld_acquire_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---------------------------------------------------------------------------
// Scan warp (CTA 0, warp 0; lives for the whole launch): turns the tiles' sizes
// into exclusive prefixes, in tile order, 128 tiles per poll. It is the only
// consumer of tileAgg and the only producer of tilePrefix, so block placement
// never involves a chain of look-backs between working CTAs.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ void scan_tiles(const EncodeParams &p, int lane) {
    uint64_t running = 0;
    uint32_t a[kScanPerLane], nxt[kScanPerLane];
    // relaxed polls; one fence per window orders them (acquire) against the
    // relaxed prefix stores that follow (release): the B tile's parked blocks
    // and table entries are visible to whoever sees the prefix
#pragma unroll
    for (int k = 0; k < kScanPerLane; k++) {
        const int64_t i = (int64_t)lane * kScanPerLane + k;
        nxt[k] = i < p.nTiles ? ld_relaxed_u32(p.tileAgg + i) : kAggReady;
    }
    for (int64_t base = 0; base < p.nTiles; base += 32 * kScanPerLane) {
        const int64_t mine = base + (int64_t)lane * kScanPerLane;
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) a[k] = nxt[k];
        // next window's first poll is in flight while this one is processed
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) {
            const int64_t i = mine + 32 * kScanPerLane + k;
            nxt[k] = i < p.nTiles ? ld_relaxed_u32(p.tileAgg + i) : kAggReady;
        }
        while (true) {
            bool all = true;
#pragma unroll
            for (int k = 0; k < kScanPerLane; k++) all = all && (a[k] & kAggReady);
            if (__all_sync(0xffffffffu, all)) break;
#pragma unroll
            for (int k = 0; k < kScanPerLane; k++)
                if (!(a[k] & kAggReady)) a[k] = ld_relaxed_u32(p.tileAgg + mine + k);
        }
        __threadfence();
        uint32_t local = 0;
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) local += a[k] & ~kAggReady;
        uint32_t incl = local;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += up;
        }
        uint64_t excl = running + (incl - local);
#pragma unroll
        for (int k = 0; k < kScanPerLane; k++) {
            if (mine + k < p.nTiles) st_relaxed_u64(p.tilePrefix + mine + k, kPrefixReady | excl);
            excl += a[k] & ~kAggReady;
        }
        running += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) st_relaxed_u64(p.tilePrefix + p.nTiles, kPrefixReady | running);
}

// ---------------------------------------------------------------------------
// C tile (one warp per block): the tile's offset is known by now; copy the
// block from its scratch slot (an L2 hit) to its final position with 128-bit
// coalesced streaming stores and complete its table entry.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void place_block(const EncodeParams &p, const TileJob &job) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const DevField *fdp = p.fields + job.f;
    const int nb = fdp->nb;
    const int b0 = job.t * kTileBlocks;
    const int b = b0 + warp;
    const int64_t ringPos = job.tile & (p.ringTiles - 1);
    if (b < nb) {
        uint64_t pre = kPrefixReady;
        if (!(p.debugMask & 16))
            while (!((pre = ld_acquire_u64(p.tilePrefix + job.tile)) & kPrefixReady)) __nanosleep(64);
        pre &= ~kPrefixReady;
        // sizes of the tile's blocks: lane i < 8 reads block i's table entry
        TmpBlock *tmp = p.tmp + fdp->firstBlock + (int64_t)job.comp * nb + b0;
        uint32_t sz = (lane < kTileBlocks && b0 + lane < nb) ? __ldcg(&tmp[lane].size16) : 0u;
        uint32_t below = 0, mySize = 0;
#pragma unroll
        for (int i = 0; i < kTileBlocks; i++) {
            uint32_t s = __shfl_sync(0xffffffffu, sz, i);
            below += i < warp ? s : 0u;
            mySize = i == warp ? s : mySize;
        }
        uint64_t prefix16 = pre + below;
        if (p.debugMask & 17) prefix16 = (uint64_t)job.tile * 8 * 56 + warp * 56;  // timing experiments only
        const int64_t dstByte = fdp->fixedBase + 64 + (int64_t)nb * 48 + (int64_t)(prefix16 * 16);
        if (dstByte + (int64_t)mySize * 16 > p.outCap) {
            if (lane == 0) atomicCAS(p.error, 0, kErrCapacity);
        } else {
            const uint4 *slot =
                reinterpret_cast<const uint4 *>(p.scratch + (ringPos * kWarpsPerCta + warp) * kSlotBytes);
            uint4 *dst = reinterpret_cast<uint4 *>(p.out + dstByte);
            for (int c = lane; c < (int)mySize; c += 32) __stcs(dst + c, __ldcg(slot + c));
            if (lane == 0) tmp[warp].prefix16 = prefix16;
        }
    }
    __syncwarp();
    if (lane == 0) atomicAdd(p.slotDone + ringPos, 1u);  // every warp counts, active or not
}

// ticket -> tile work. Tickets alternate between
//   A(g)            first read of tile g (min/max)
//   B(g) + C(g')    g = A's tile - lag: the field's parameters are known and the
//                   second read hits L2; park the packed blocks, publish their
//                   size. g' = g - lag2: the scan warp has that tile's offset;
//                   place it while this tile's input loads are in flight.
// With lag < 0 (CFunc boundary, parameters given) there are no A tickets.
struct JobPair {
    TileJob main, place;
};

template <bool kFromQ>
__device__ __forceinline__ JobPair decode_ticket(const EncodeParams &p, int64_t t) {
    JobPair jp;
    jp.main.kind = jp.place.kind = 0;
    jp.main.f = jp.main.comp = jp.main.t = 0;
    jp.place = jp.main;
    jp.main.tile = jp.place.tile = -1;
    int64_t a = -1, b = -1;
    if (kFromQ || p.lag < 0) {
        b = t;
    } else {
        if (t & 1) b = (t >> 1) - p.lag; else a = t >> 1;
    }
    if (a >= 0 && a < p.nTiles) {
        jp.main.kind = 1;
        jp.main.tile = a;
        locate_tile(p, a, jp.main.f, jp.main.comp, jp.main.t);
    }
    if (b >= 0 && b < p.nTiles) {
        jp.main.kind = 2;
        jp.main.tile = b;
        locate_tile(p, b, jp.main.f, jp.main.comp, jp.main.t);
    }
    const int64_t c = b - p.lag2;
    if (b >= 0 && c >= 0 && c < p.nTiles) {
        jp.place.kind = 3;
        jp.place.tile = c;
        locate_tile(p, c, jp.place.f, jp.place.comp, jp.place.t);
        if (jp.main.kind == 0) jp.main.kind = 4;  // past the last B tile: placement only
    }
    return jp;
}

// ---------------------------------------------------------------------------
// k_encode: CTA 0 hosts the scan warp; every other CTA claims one ticket.
// A ticket only ever waits on smaller tickets (ready flags of its field's A
// tiles; the ring slot's previous tenants, since ringTiles > lag2; the scan
// warp, which needs B tiles at most 127 ahead of the tile being placed, and
// lag2 > 127), and only running CTAs hold tickets, so nothing can deadlock.
// ---------------------------------------------------------------------------
template <bool kFromQ>
__global__ void __launch_bounds__(256, 4) k_encode(EncodeParams p, const uint64_t *qIn) {
    __shared__ __align__(16) uint32_t stage[kWarpsPerCta][kStageWords];
    __shared__ float sLo[kWarpsPerCta], sHi[kWarpsPerCta];
    __shared__ uint32_t sSize[kWarpsPerCta];
    __shared__ uint32_t sArrive;
    __shared__ JobPair sJob;

    if (blockIdx.x == 0) {
        if (threadIdx.x < 32) scan_tiles(p, threadIdx.x);
        return;
    }
    if (threadIdx.x == 0) {
        sArrive = 0;
        sJob = decode_ticket<kFromQ>(p, (int64_t)atomicAdd(p.ticket, 1u));
    }
    __syncthreads();
    const TileJob job = sJob.main;
    if (job.kind == 1) {
        if (!kFromQ) {
            minmax_phase1(p, job, sLo, sHi);
            __syncthreads();
            if (threadIdx.x == 0) minmax_phase2(p, job, sLo, sHi);
        }
    } else if (job.kind == 2) {
        quantize_pack<kFromQ>(p, qIn, job, sJob.place, stage[threadIdx.x >> 5], sSize, &sArrive);
    } else if (job.kind == 4) {
        place_block(p, sJob.place);
    }
}

// ---------------------------------------------------------------------------
// k_finalize: one thread per block turns TmpBlock into the wire table entry
// (offset relative to the field's payload base); one thread per field writes
// the 64-byte header and the field's byte offset.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t prefix_of_tile(const EncodeParams &p, int64_t tile) {
    return p.tilePrefix[tile] & ~kPrefixReady;
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
        uint64_t p0 = prefix_of_tile(p, fd.firstTile);
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
            p.offsets[f] = (uint64_t)fixedEnd + prefix_of_tile(p, nTilesTotal) * 16;
            return;
        }
        const DevField fd = p.fields[f];
        uint64_t p0 = fd.firstTile < nTilesTotal ? prefix_of_tile(p, fd.firstTile) : prefix_of_tile(p, nTilesTotal);
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

// enc_consume.cuh -- B work, placement, the kernel and k_finalize. Included by
// enc_kernel.cuh.
#pragma once

#include "enc_kernel.cuh"

namespace mnw {

__device__ __forceinline__ void release_stage(const RingCtrl &rc, int s, bool wrote, int lane) {
    __syncwarp();
    if (lane == 0) {
        if (wrote) fence_proxy_async();  // generic-proxy writes before the TMA's next fill
        rc.uses[s] = rc.uses[s] + 1;     // one owner at a time
        mbar_arrive(rc.empty + s);
    }
}

// the producer found the component's parameters not yet published: wait for them here
__device__ __forceinline__ bool wait_quant(const EncodeParams &p, WorkDesc &d) {
    const uint64_t *w = p.fq + ((int64_t)d.f * 3 + d.comp) * 2;
    uint64_t w0 = ld_relaxed_u64(w), w1 = ld_relaxed_u64(w + 1);
    uint32_t spins = 0;
    while (!(w0 & w1 & kFqValid)) {
        __nanosleep(100);
        if (++spins > kSpinLimit) { spin_timeout(p.error, 1, d.blk); return false; }
        w0 = ld_relaxed_u64(w);
        w1 = ld_relaxed_u64(w + 1);
    }
    d.scale = __longlong_as_double((long long)(w0 & ~kFqValid));
    d.x0 = __uint_as_float((uint32_t)w1);
    d.depth = (uint32_t)(w1 >> 32) & 0xffu;
    return true;
}

// ---------------------------------------------------------------------------
// A consumer parks its packed blocks until their offsets are known and remembers
// them in a small FIFO; it places the oldest one as soon as that block's prefix
// has arrived. A block is packed straight into the consumer's arena in shared
// memory when there is room (a ring allocator, blocks leave in order); otherwise
// through the ring stage into one of the consumer's scratch slots in global
// memory (L2 resident). The warp that parked a block is the one that places it,
// so no fences are involved anywhere: the size flags and the prefixes carry
// their own payload.
// ---------------------------------------------------------------------------
constexpr int kFifo = 8;

struct ParkFifo {
    int4 *where;       // shared: [kFifo] arena offset in words (-1: scratch slot), arena words taken, packed words, size16
    int64_t *blk;      // shared: [kFifo]
    int64_t *pay;      // shared: [kFifo] WorkDesc::payBase of the block's field
    uint32_t *slots;   // global: [kFifo][kSlotWords]
    uint32_t *arena;   // shared: [kArenaWords]
    int head, count;
    int aTail, aUsed;  // arena: next word to hand out, words in use (including skipped tails)
    unsigned long long *wst;  // shared: phase clocks of this warp, or null
};

#define MNW_PHASE(i)                                   \
    if (kStats && ff.wst) {                                    \
        const long long tNow_ = clock64();             \
        if (lane == 0) ff.wst[i] += tNow_ - tPh;       \
        tPh = tNow_;                                   \
    }

// room for `need` contiguous words in the arena? (-1: no) `span` = words taken, counting a skipped tail
__device__ __forceinline__ int arena_alloc(ParkFifo &ff, int need, int &span) {
    int start = ff.aTail, waste = 0;
    if (start + need > kArenaWords) {
        waste = kArenaWords - start;
        start = 0;
    }
    span = waste + need;
    if (need > kArenaWords || ff.aUsed + span > kArenaWords) return -1;
    ff.aTail = start + need;
    ff.aUsed += span;
    return start;
}

// kPrefixReady | prefix of the block, or 0 if the scan has not got there yet
__device__ __forceinline__ uint64_t peek_prefix(const EncodeParams &p, int64_t blk) {
    uint64_t pre;
    return block_prefix(p.blkRel, p.winBase, blk, pre) ? (kPrefixReady | pre) : 0ull;
}

// The oldest parked block goes to its final position: `pre` is its prefix, v0/v1
// its first two 16-byte chunks per lane if it sits in a scratch slot (loaded by
// the caller so that their latency overlaps other work; wider blocks load the
// rest here). Nothing here reads global memory: the table entry was written at
// packing time (EncodeParams::tbl) and the payload base travelled with the block.
__device__ __forceinline__ void place_finish(const EncodeParams &p, ParkFifo &ff, uint64_t pre, const uint4 &v0,
                                             const uint4 &v1, int lane) {
    const int k = ff.head;
    const int4 wh = ff.where[k];
    const int size16 = wh.w;
    const int64_t dstByte = ff.pay[k] + (int64_t)(pre * 16);
    if (dstByte + (int64_t)size16 * 16 > p.outCap) {
        if (lane == 0) atomicCAS(p.error, 0, kErrCapacity);
    } else {
        uint4 *dst = reinterpret_cast<uint4 *>(p.out + dstByte);
        if (wh.x >= 0) {
            const uint32_t *src = ff.arena + wh.x;
            const int nWords = wh.z;
            for (int c = lane; c < size16; c += 32) {
                const int m = c * 4;
                uint4 o;
                o.x = src[skew(m)];
                o.y = m + 1 < nWords ? src[skew(m + 1)] : 0u;
                o.z = m + 2 < nWords ? src[skew(m + 2)] : 0u;
                o.w = m + 3 < nWords ? src[skew(m + 3)] : 0u;
                st_stream_u4(dst + c, o);
            }
        } else {
            const uint4 *slot = reinterpret_cast<const uint4 *>(ff.slots + (size_t)k * kSlotWords);
            if (lane < size16) st_stream_u4(dst + lane, v0);
            if (lane + 32 < size16) st_stream_u4(dst + lane + 32, v1);
            for (int c = lane + 64; c < size16; c += 32) st_stream_u4(dst + c, __ldcg(slot + c));
        }
    }
    ff.aUsed -= wh.y;
    __syncwarp();
    ff.head = (ff.head + 1) % kFifo;
    ff.count--;
}

// the first two chunks per lane of the oldest parked block (if it sits in a scratch slot)
__device__ __forceinline__ void load_head(const ParkFifo &ff, uint4 &v0, uint4 &v1, int lane) {
    const int4 wh = ff.where[ff.head];
    v0 = v1 = make_uint4(0u, 0u, 0u, 0u);
    if (wh.x >= 0) return;
    const uint4 *slot = reinterpret_cast<const uint4 *>(ff.slots + (size_t)ff.head * kSlotWords);
    if (lane < wh.w) v0 = __ldcg(slot + lane);
    if (lane + 32 < wh.w) v1 = __ldcg(slot + lane + 32);
}

// place the oldest parked block, waiting for its prefix if need be
__device__ __forceinline__ bool place_oldest(const EncodeParams &p, ParkFifo &ff, int lane) {
    const int64_t pl = ff.blk[ff.head];
    uint4 v0, v1;
    load_head(ff, v0, v1, lane);
    uint64_t pre = 0;
    uint32_t spins = 0;
    while (!block_prefix(p.blkRel, p.winBase, pl, pre)) {
        __nanosleep(200);
        if (++spins > kSpinLimit) { spin_timeout(p.error, 2, pl); return false; }
    }
    place_finish(p, ff, pre, v0, v1, lane);
    return true;
}

// ---------------------------------------------------------------------------
// B work on block d.blk
// ---------------------------------------------------------------------------
// (false: a wait timed out, the caller leaves the kernel)
template <bool kFromQ, bool kStats>
__device__ __forceinline__ bool b_block(const EncodeParams &p, WorkDesc &d, float *stage, const RingCtrl &rc,
                                        int s, ParkFifo &ff, int lane) {
    const int64_t blk = d.blk;
    long long tPh = kStats && ff.wst ? clock64() : 0;
    // the oldest parked block's prefix words are asked for now and looked at after
    // this block is packed: an L2 round trip under load is a few thousand cycles
    bool tryPlace = ff.count > 0 && ff.count < kFifo;
    uint32_t relW = 0;
    uint64_t wbW = 0;
    if (tryPlace) {
        const int64_t hb = ff.blk[ff.head];
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(relW) : "l"(p.blkRel + hb) : "memory");
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(wbW) : "l"(p.winBase + hb / kScanWin) : "memory");
    }
    if (!kFromQ && !(d.flags & kDescQuant) && !wait_quant(p, d)) return false;
    const uint32_t top = (1u << d.depth) - 1u;
    const bool biased = d.depth <= 23;
    MNW_PHASE(0)

    const int len = d.len;
    uint32_t q[kChunk];
    if (kFromQ) {
        // literal CFunc boundary: quantised uint64 values in (types.h:117-122)
        const DevField *fdp = p.fields + d.f;
        const uint64_t *qsrc = p.qIn + fdp->ioOff + (int64_t)d.comp * fdp->n + (int64_t)d.b * kBlockLen;
#pragma unroll
        for (int i = 0; i < kChunk; i++) {
            const int e = lane * kChunk + i;
            q[i] = e < len ? (uint32_t)qsrc[e] : 0u;
        }
    } else {
        if (d.flags & kDescSlow) {
            // ragged or unaligned block: plain loads into the stage, same swizzled layout
            const DevField *fdp = p.fields + d.f;
            const float *src = p.in + fdp->ioOff + (int64_t)d.comp * fdp->n + (int64_t)d.b * kBlockLen;
            const float fill = __ldg(src);
#pragma unroll 4
            for (int k = 0; k < 32; k++) {
                const int e = k * 32 + lane;
                const float x = e < len ? __ldg(src + e) : fill;
                stage[swz_chunk(e >> 5, (e >> 2) & 7) * 4 + (e & 3)] = x;
            }
            __syncwarp();
        }
        Unwrap uw;
        uw.tHi = d.tHi;
        uw.tLo = d.tLo;
        uw.L = d.L;
        if (d.flags & kDescPos) {
            if (biased) quantise32<true, true>(stage, lane, uw, d.x0, d.scale, top, q);
            else quantise32<true, false>(stage, lane, uw, d.x0, d.scale, top, q);
        } else {
            if (biased) quantise32<false, true>(stage, lane, uw, d.x0, d.scale, top, q);
            else quantise32<false, false>(stage, lane, uw, d.x0, d.scale, top, q);
        }
    }

    MNW_PHASE(1)
    BlockStats bs = (!kFromQ && biased) ? delta_width<true>(q, len, lane, p.error)
                                        : delta_width<false>(q, len, lane, p.error);
    if (!kFromQ && biased) bs.first -= kQBias;
    const int nWords = (len * bs.w + 31) >> 5;
    const uint32_t size16 = (uint32_t)((nWords + 3) >> 2);
    // the size is all the scan warp needs: publish it before anything else; the
    // table entry (k_tables adds the offset) goes out with it
    if (lane == 0) p.tbl[blk] = make_uint4(bs.first, (uint32_t)bs.dmin, (uint32_t)bs.w, 0u);
    if (lane == 0)
        asm volatile("st.relaxed.gpu.global.u16 [%0], %1;" ::"l"(p.blkAgg + blk), "h"((uint16_t)(kAggReady | size16)) : "memory");

    if (ff.count == kFifo) {
        // no free record: the oldest block has to go now
        if (!place_oldest(p, ff, lane)) return false;
    }
    // pack straight into the arena if there is room (the block's values are in
    // registers by now, so its stage can go back at once), else through the stage
    int span = 0;
    // (the pack routines write all 32 x width words of a block, ragged or not)
    const int loc = arena_alloc(ff, 33 * bs.w + 1, span);
    uint32_t *tmp = loc >= 0 ? ff.arena + loc : reinterpret_cast<uint32_t *>(stage);
    if (kStats && ff.wst && lane == 0 && loc < 0) ff.wst[7] += 1;  // stats: blocks that went through a scratch slot
    if (loc >= 0) release_stage(rc, s, !kFromQ && (d.flags & kDescSlow), lane);  // (a ragged block was staged by hand)
    __syncwarp();
    MNW_PHASE(2)
    if (bs.w > 0) pack_dispatch(bs.w, q, tmp, lane);
    __syncwarp();
    MNW_PHASE(3)
    MNW_PHASE(4)
    // ask now for what placing the oldest parked block will need (its prefix, its
    // first chunks if they sit in a scratch slot): the loads fly while this block is parked
    uint4 v0H, v1H;
    v0H = v1H = make_uint4(0u, 0u, 0u, 0u);
    if (tryPlace) load_head(ff, v0H, v1H, lane);
    const int k = (ff.head + ff.count) % kFifo;
    if (loc < 0) {
        uint4 *slot = reinterpret_cast<uint4 *>(ff.slots + (size_t)k * kSlotWords);
        for (int c = lane; c < (int)size16; c += 32) {
            const int m = c * 4;
            uint4 o;
            o.x = tmp[skew(m)];
            o.y = m + 1 < nWords ? tmp[skew(m + 1)] : 0u;
            o.z = m + 2 < nWords ? tmp[skew(m + 2)] : 0u;
            o.w = m + 3 < nWords ? tmp[skew(m + 3)] : 0u;
            slot[c] = o;
        }
    }
    if (lane == 0) {
        ff.blk[k] = blk;
        ff.pay[k] = d.payBase;
        ff.where[k] = make_int4(loc, loc >= 0 ? span : 0, nWords, (int)size16);
    }
    ff.count++;
    if (loc < 0) release_stage(rc, s, true, lane);  // includes __syncwarp
    else __syncwarp();
    MNW_PHASE(5)
    if (tryPlace && (relW & kRelReady) && (wbW & kPrefixReady))
        place_finish(p, ff, (wbW & ~kPrefixReady) + (relW & ~kRelReady), v0H, v1H, lane);
    MNW_PHASE(5)
    return true;
}

}  // namespace mnw

#include "enc_entry.cuh"

// enc_consume.cuh -- B work, placement, the kernel and k_finalize. Included by
// enc_kernel.cuh.
#pragma once

#include "enc_kernel.cuh"

namespace mnw {

__device__ __forceinline__ void release_stage(const RingCtrl &rc, int s, bool wrote, int lane) {
    __syncwarp();
    if (lane == 0) {
        if (rc.tRel) rc.tRel[s] = clock64();  // (statistics build only)
        if (wrote) fence_proxy_async();  // generic-proxy writes before the TMA's next fill
        rc.uses[s] = rc.uses[s] + 1;     // one owner at a time
        mbar_arrive(rc.empty + s);
    }
}

// the producer found the component's parameters not yet published: wait for them here
__device__ __forceinline__ bool wait_quant(const EncodeParams &p, WorkDesc &d, int lane) {
    const uint64_t *w = p.fq + ((int64_t)d.f * 3 + d.comp) * 2;
    uint64_t w0 = ld_relaxed_u64(w), w1 = ld_relaxed_u64(w + 1);
    SpinClock clk;
    if (p.stats && lane == 0 && !(w0 & w1 & kFqValid)) atomicAdd(p.stats + 30, 1ull);  // (statistics build only)
    while (!(w0 & w1 & kFqValid)) {
        __nanosleep(100);
        if (p.stats && lane == 0) atomicAdd(p.stats + 31, 1ull);
        if (spin_expired(clk)) { spin_timeout(p.error, 1, d.blk); return false; }
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
// has arrived. A block's values go to registers and its ring stage back to the
// producer before anything else happens (the ring is the kernel's window of bytes
// in flight: a stage that waits for arithmetic is a stage that is not loading).
// The block is packed into the consumer's arena in shared memory (a ring
// allocator that holds at least one block of any width; blocks leave in order).
// When the arena has no room and the oldest block's prefix is not there yet, the
// oldest arena-resident block moves to one of the consumer's scratch slots in
// global memory (L2 resident) -- so a consumer only ever blocks when all of its
// kFifo records are taken. The warp that parked a block is the one that places
// it, so no fences are involved anywhere: the size flags and the prefixes carry
// their own payload.
// ---------------------------------------------------------------------------
struct ParkFifo {
    int4 *where;       // shared: [kFifo] arena offset in words (-1: scratch slot), arena words taken, packed words, size16
    int64_t *blk;      // shared: [kFifo]
    int64_t *pay;      // shared: [kFifo] WorkDesc::payBase of the block's field
    uint32_t *slots;   // global: [kFifo][kSlotWords]
    uint32_t *arena;   // shared: [kArenaWords]
    int head, count;
    int aTail, aUsed;  // arena: next word to hand out, words in use (including skipped tails)
    int resident;      // records (the newest ones) whose block sits in the arena
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
    if (ff.aUsed + span > kArenaWords) return -1;
    ff.aTail = start + need;
    ff.aUsed += span;
    return start;
}

// 16-byte chunk c of a packed block in the arena (skewed layout). The pack routines
// write all 32 x width words of a block -- a multiple of four -- and the residuals
// beyond a ragged block's length are zero, so the words past the block's last one
// are there and zero; four words starting at a multiple of four share a skew row.
__device__ __forceinline__ uint4 arena_chunk(const uint32_t *src, int c) {
    const uint32_t *w = src + skew(c * 4);
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// the oldest block that still sits in the arena moves to its record's scratch slot
__device__ __forceinline__ void spill_oldest_resident(ParkFifo &ff, int lane) {
    const int k = (ff.head + ff.count - ff.resident) % kFifo;
    const int4 wh = ff.where[k];
    uint4 *slot = reinterpret_cast<uint4 *>(ff.slots + (size_t)k * kSlotWords);
    const uint32_t *src = ff.arena + wh.x;
#pragma unroll 1
    for (int c = lane; c < wh.w; c += 32) slot[c] = arena_chunk(src, c);
    __syncwarp();
    if (lane == 0) ff.where[k] = make_int4(-1, 0, wh.z, wh.w);
    ff.aUsed -= wh.y;
    ff.resident--;
    if (ff.resident == 0) ff.aTail = ff.aUsed = 0;
    __syncwarp();
}

// The oldest parked block goes to its final position: `pre` is its prefix. The
// table entry was written at packing time (EncodeParams::tbl) and the payload
// base travelled with the block.
__device__ __forceinline__ void place_finish(const EncodeParams &p, ParkFifo &ff, uint64_t pre, int lane) {
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
#pragma unroll 1
            for (int c = lane; c < size16; c += 32) st_stream_u4(dst + c, arena_chunk(src, c));
        } else {
            const uint4 *slot = reinterpret_cast<const uint4 *>(ff.slots + (size_t)k * kSlotWords);
#pragma unroll 1
            for (int c = lane; c < size16; c += 32) st_stream_u4(dst + c, __ldcg(slot + c));
        }
    }
    if (wh.x >= 0) {
        ff.aUsed -= wh.y;
        ff.resident--;
        if (ff.resident == 0) ff.aTail = ff.aUsed = 0;
    }
    __syncwarp();
    ff.head = (ff.head + 1) % kFifo;
    ff.count--;
}

// kPrefixReady | prefix of the block, or 0 if the scan has not got there yet
__device__ __forceinline__ uint64_t peek_prefix(const EncodeParams &p, int64_t blk) {
    uint64_t pre;
    return block_prefix(p.blkRel, p.winBase, blk, pre) ? (kPrefixReady | pre) : 0ull;
}

// place the oldest parked block, waiting for its prefix if need be
__device__ __forceinline__ bool place_oldest(const EncodeParams &p, ParkFifo &ff, int lane) {
    const int64_t pl = ff.blk[ff.head];
    uint64_t pre = 0;
    SpinClock clk;
    while (!block_prefix(p.blkRel, p.winBase, pl, pre)) {
        __nanosleep(200);
        if (spin_expired(clk)) { spin_timeout(p.error, 2, pl); return false; }
    }
    place_finish(p, ff, pre, lane);
    return true;
}

// ---------------------------------------------------------------------------
// B work on block d.blk
// ---------------------------------------------------------------------------
// (false: a wait timed out, the caller leaves the kernel; the stage has been handed back)
// (kVec: full blocks off the 16-byte grid are staged with vector loads -- the general kernel only: the
// kernel of the uniform batches stays what it is)
template <bool kFromQ, bool kStats, bool kVec = false>
__device__ __forceinline__ bool b_block(const EncodeParams &p, WorkDesc &d, float *stage, const RingCtrl &rc,
                                        int s, ParkFifo &ff, int lane) {
    const int64_t blk = d.blk;
    long long tPh = kStats && ff.wst ? clock64() : 0;
    // the oldest parked block's prefix words are asked for now and looked at after
    // this block is packed: an L2 round trip under load is a few thousand cycles
    const bool tryPlace = ff.count > 0;
    uint32_t relW = 0;
    uint64_t wbW = 0;
    int64_t hb = -1;
    if (tryPlace) {
        hb = ff.blk[ff.head];
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(relW) : "l"(p.blkRel + hb) : "memory");
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(wbW) : "l"(p.winBase + hb / kScanWin) : "memory");
    }
    const int len = d.len;
    uint32_t q[kChunk];
    if (kFromQ) {
        release_stage(rc, s, false, lane);  // nothing came through the ring
    } else {
        if (d.flags & kDescSlow) {
            // ragged or unaligned block: plain loads into the stage, same swizzled layout
            const DevField *fdp = p.fields + d.f;
            const float *src = p.in + fdp->ioOff + (int64_t)d.comp * fdp->n + (int64_t)d.b * kBlockLen;
            if (kVec && len == kBlockLen) {
                stage_full_unaligned(src, stage, lane);  // (a full block off the 16-byte grid: vector loads)
            } else {
                const float fill = __ldg(src);
#pragma unroll 4
                for (int k = 0; k < 32; k++) {
                    const int e = k * 32 + lane;
                    const float x = e < len ? __ldg(src + e) : fill;
                    stage[swz_chunk(e >> 5, (e >> 2) & 7) * 4 + (e & 3)] = x;
                }
                __syncwarp();
            }
        }
        // the lane's 32 values to registers, the stage back to the producer
        load_block32(stage, lane, q);
        release_stage(rc, s, (d.flags & kDescSlow) != 0, lane);
        if (kStats && p.stats && lane == 0 && !(d.flags & kDescQuant)) atomicAdd(p.stats + 29, 1ull);
        if (!(d.flags & kDescQuant) && !wait_quant(p, d, lane)) return false;
    }
    const uint32_t top = (1u << d.depth) - 1u;
    const bool biased = d.depth <= 23;
    MNW_PHASE(0)

    if (kFromQ) {
        // literal CFunc boundary: quantised uint64 values in (types.h:117-122)
        const DevField *fdp = p.fields + d.f;
        const uint64_t *qsrc = p.qIn + fdp->ioOff + (int64_t)d.comp * fdp->n + (int64_t)d.b * kBlockLen;
        // values must fit the declared depth (a QField from another quantiser may not)
        bool bad = false;
#pragma unroll
        for (int i = 0; i < kChunk; i++) {
            const int e = lane * kChunk + i;
            const uint64_t v = e < len ? qsrc[e] : 0ull;
            bad = bad || v > (uint64_t)top;
            q[i] = (uint32_t)v;
        }
        if (__any_sync(kFull, bad) && lane == 0) atomicCAS(p.error, 0, kErrRange);
    } else {
        Unwrap uw;
        uw.tHi = d.tHi;
        uw.tLo = d.tLo;
        uw.L = d.L;
        if (d.flags & kDescPos) unwrap32(uw, q);
        if (biased) quantise32<true>(d.x0, d.scale, top, q);
        else quantise32<false>(d.x0, d.scale, top, q);
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

    // a record and room in the arena for this block (the pack routines write all
    // 32 x width words of a block, ragged or not)
    if (kStats && p.stats && lane == 0 && ff.count == kFifo) atomicAdd(p.stats + 18, 1ull);  // all records taken
    if (ff.count == kFifo && !place_oldest(p, ff, lane)) return false;  // the only wait on another warp's work
    int span = 0, loc;
    while ((loc = arena_alloc(ff, 33 * bs.w + 1, span)) < 0) {
        if (kStats && p.stats && lane == 0) atomicAdd(p.stats + 19, 1ull);  // the arena had no room
        const uint64_t pre = peek_prefix(p, ff.blk[ff.head]);
        if (pre) place_finish(p, ff, pre & ~kPrefixReady, lane);
        else spill_oldest_resident(ff, lane);
        if (kStats && ff.wst && lane == 0 && !pre) ff.wst[7] += 1;  // stats: blocks that went through a scratch slot
    }
    uint32_t *tmp = ff.arena + loc;
    __syncwarp();
    MNW_PHASE(2)
#if defined(MNW_ABLATE) && (MNW_ABLATE & 2)
    if (bs.w > 0 && lane == 0) tmp[0] = q[0] + q[31];  // timing experiment only: no packing
#else
    if (bs.w > 0) pack_dispatch(bs.w, q, tmp, lane);
#endif
    MNW_PHASE(3)
    const int k = (ff.head + ff.count) % kFifo;
    if (lane == 0) {
        ff.blk[k] = blk;
        ff.pay[k] = d.payBase;
        ff.where[k] = make_int4(loc, span, nWords, (int)size16);
    }
    ff.count++;
    ff.resident++;
    __syncwarp();
    MNW_PHASE(4)
    // (the head may have changed meanwhile: the words asked for belong to block hb)
    if (tryPlace && (relW & kRelReady) && (wbW & kPrefixReady) && ff.blk[ff.head] == hb)
        place_finish(p, ff, (wbW & ~kPrefixReady) + (relW & ~kRelReady), lane);
    MNW_PHASE(5)
    return true;
}

}  // namespace mnw

#include "enc_entry.cuh"

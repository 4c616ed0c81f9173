// enc_ring.cuh -- the input ring of k_encode. The producer warp of a CTA draws
// group tickets; each of its 32 lanes owns one slot of the group: it works out
// the slot's block, waits for the next ring stage to be free, writes the work
// descriptor next to it and has the TMA unit copy the block's 4 KiB into the
// stage (2D tensor map over the input as rows of 32 floats, 128-byte swizzle,
// so that a consumer lane can read its 32 consecutive values without bank
// conflicts). Consumer warps claim stages in order, wait on the stage's "full"
// mbarrier and hand the stage back through its "empty" mbarrier. A parity wait
// can only tell the current phase from the one before, and consumers racing
// through no-op slots can claim the next use of a stage whose current fill is
// still in flight: so a consumer first waits until the stage's previous use has
// been handed back (a plain counter), and only then for the fill.
#pragma once

#include "enc_math.cuh"

namespace mnw {

struct WorkDesc {  // 80 bytes, one per ring stage: the block and what its component needs
    int32_t kind;
    int32_t f, comp, b;
    int64_t blk;
    int32_t len;
    uint32_t flags;   // kDescSlow | kDescPos | kDescQuant
    double scale;     // util_UniformBinIndex's 2^depth / maxDiff (kDescQuant)
    float x0;
    uint32_t depth;
    float tHi, tLo;   // unwrap thresholds (kDescPos), see make_unwrap
    float L;
    uint32_t pad;
    int64_t payBase;  // byte offset of the field's payload area if the batch's prefix were 0
    int64_t pad2;
};
enum : uint32_t {
    kDescSlow = 1u,   // the block did not go through the TMA: the consumer loads it itself
    kDescPos = 2u,    // a position component: periodic unwrap applies
    kDescQuant = 4u,  // scale, x0 and depth are filled in (else the consumer waits for them)
};

struct RingCtrl {
    uint64_t *full;   // [kEncStages] descriptor written and TMA bytes landed
    uint64_t *empty;  // [kEncStages] the consumer is done with the stage
    WorkDesc *desc;   // [kEncStages]
    uint32_t *qHead;  // next stage sequence number to claim (consumers, atomic)
    volatile uint32_t *uses;  // [kEncStages] how many times the stage has been handed back
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t"
                 ".reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t"
                 "}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// bounded wait: a barrier that never completes raises kErrTimeout instead of hanging
// (false: the caller leaves the kernel, it must not touch the ring again)
__device__ __forceinline__ bool mbar_wait(int *error, uint64_t *bar, uint32_t parity, int site) {
    if (mbar_try_wait(bar, parity)) return true;  // the usual case: no counters, no clock
    uint32_t spins = 0;
    unsigned long long t0 = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 1023u) == 0) {  // a try_wait may itself suspend the warp for a while: bound the time
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > kWaitLimitNs) { spin_timeout(error, site, -1); return false; }
        }
    }
    return true;
}
__device__ __forceinline__ void tma_load_block(void *smemDst, const CUtensorMap *map, int32_t row, uint64_t *bar,
                                               uint64_t pol) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(smem_u32(smemDst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(0), "r"(row), "r"(smem_u32(bar)), "l"(pol)
        : "memory");
}
__device__ __forceinline__ void tma_load_block_plain(void *smemDst, const CUtensorMap *map, int32_t row, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(smemDst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(0), "r"(row), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ RingCtrl ring_ctrl(uint8_t *smem) {
    RingCtrl rc;
    uint8_t *c = smem + kEncCtrlOff;
    rc.full = reinterpret_cast<uint64_t *>(c);
    rc.empty = rc.full + kEncStages;
    rc.desc = reinterpret_cast<WorkDesc *>(rc.empty + kEncStages);
    rc.qHead = reinterpret_cast<uint32_t *>(rc.desc + kEncStages);
    rc.uses = rc.qHead + 4;
    return rc;
}

// ---------------------------------------------------------------------------
// producer warp: all 32 lanes, one slot each per pass
// ---------------------------------------------------------------------------
template <bool kFromQ>
__device__ __forceinline__ void producer(const EncodeParams &p, const CUtensorMap *map, uint8_t *smem,
                                         const RingCtrl &rc, int lane) {
    // A reads carry no cache hint, B reads are evict_first: with the B work ~56 MB of
    // A traffic behind, about two thirds of the B reads still hit in L2 (evict_last on
    // the A reads thrashes: nothing survives; MNWCUDA_DBG=2 tries it, 8 streams both)
    const uint64_t keep = l2_policy_keep(), drop = l2_policy_drop();
    const bool onlyB = kFromQ || p.lagG < 0;
    struct {
        int f, comp;
        uint32_t flags, depth;
        float L, tHi, tLo, x0;
        double scale;
    } pc;
    pc.f = pc.comp = -1;
    pc.flags = pc.depth = 0;
    pc.L = pc.tHi = pc.tLo = pc.x0 = 0.0f;
    pc.scale = 0.0;
    uint32_t pass = 0;
    uint32_t tk = 0;
    if (lane == 0) tk = atomicAdd(p.ticket, 1u);
    while (true) {
        const int64_t g = (int64_t)__shfl_sync(kFull, tk, 0);
        if (g >= p.nGroups) break;
        if (lane == 0) tk = atomicAdd(p.ticket, 1u);  // the next ticket is on its way meanwhile
        WorkDesc d;
        d.kind = kWorkNop;
        d.f = d.comp = d.b = 0;
        d.len = 0;
        d.flags = kDescSlow;
        d.scale = 0.0;
        d.x0 = d.tHi = d.tLo = d.L = 0.0f;
        d.depth = 0;
        d.pad = 0;
        d.payBase = d.pad2 = 0;
        int64_t blk;
        if (onlyB) {
            blk = g * kGroupSlots + lane;
            if (blk < p.nBlocks) d.kind = kWorkB;
        } else if (lane < kGroupBlocks) {
            blk = g * kGroupBlocks + lane;
            if (blk < p.nBlocks) d.kind = kWorkA;
        } else {
            blk = (g - p.lagG) * kGroupBlocks + (lane - kGroupBlocks);
            if (blk >= 0 && blk < p.nBlocks) d.kind = kWorkB;
        }
        d.blk = blk;
        bool fast = false;
        int64_t row = 0;
        if (d.kind != kWorkNop) {
            locate_block(p, blk, d.f, d.comp, d.b);
            const DevField *fdp = p.fields + d.f;
            const int n = fdp->n;
            d.len = min(kBlockLen, n - d.b * kBlockLen);
            d.payBase = fdp->fixedBase + 64 + (int64_t)fdp->nb * 48;
            const int64_t off = fdp->ioOff + (int64_t)d.comp * n + (int64_t)d.b * kBlockLen;  // floats from p.in
            row = (p.tmaShift + off) >> 5;
            fast = !kFromQ && p.tmaShift >= 0 && d.len == kBlockLen && ((p.tmaShift + off) & 31) == 0 &&
                   row + 32 <= p.tmaRows;
            // the component's constants: this lane sees the same component for many passes
            if (pc.f != d.f || pc.comp != d.comp) {
                pc.f = d.f;
                pc.comp = d.comp;
                pc.flags = 0;
                pc.L = fdp->width;
                pc.tHi = pc.tLo = pc.x0 = 0.0f;
                pc.scale = 0.0;
                pc.depth = 0;
                if (kFromQ) pc.depth = __ldg(&p.quant[d.f].depth);  // given by the caller: the range check needs it
                if (!kFromQ && fdp->fieldCode == field_Posn) {
                    const Unwrap uw = make_unwrap(__ldg(p.in + fdp->ioOff + (int64_t)d.comp * n), fdp->width);
                    pc.tHi = uw.tHi;
                    pc.tLo = uw.tLo;
                    pc.flags |= kDescPos;
                }
            }
            if (d.kind == kWorkB && !kFromQ && !(pc.flags & kDescQuant)) {
                if (p.lagG >= 0) {
                    // published during this launch by the reduce warp; if not there yet the consumer waits
                    const uint64_t *w = p.fq + ((int64_t)d.f * 3 + d.comp) * 2;
                    const uint64_t w0 = ld_relaxed_u64(w), w1 = ld_relaxed_u64(w + 1);
                    if (w0 & w1 & kFqValid) {
                        pc.scale = __longlong_as_double((long long)(w0 & ~kFqValid));
                        pc.x0 = __uint_as_float((uint32_t)w1);
                        pc.depth = (uint32_t)(w1 >> 32) & 0xffu;
                        pc.flags |= kDescQuant;
                    }
                } else {
                    const DevQuant *qp = p.quant + d.f;  // given by the caller, written before the launch
                    pc.x0 = __ldg(&qp->X0[d.comp]);
                    pc.scale = __ldg(&qp->scale);
                    pc.depth = __ldg(&qp->depth);
                    pc.flags |= kDescQuant;
                }
            }
            d.flags = pc.flags | (fast ? 0u : kDescSlow);
            d.scale = pc.scale;
            d.x0 = pc.x0;
            d.depth = pc.depth;
            d.tHi = pc.tHi;
            d.tLo = pc.tLo;
            d.L = pc.L;
        }
        const uint32_t seq = pass * kGroupSlots + (uint32_t)lane;
        const int s = (int)(seq % kEncStages);
        const uint32_t use = seq / kEncStages;
        // every lane fills its stage the moment that stage is handed back: a lane must
        // not wait for the slowest stage of the pass (a B block holds one ten times
        // longer than an A block)
        bool filled = false;
        unsigned long long t0 = 0;
        for (uint32_t spins = 0;; spins++) {
            if (!filled && (use == 0 || mbar_try_wait(rc.empty + s, (use - 1) & 1))) {
                rc.desc[s] = d;
                if (fast) {
                    mbar_arrive_expect_tx(rc.full + s, kEncStageBytes);
                    if (d.kind == kWorkA && !(p.dbg & 10))
                        tma_load_block_plain(smem + kEncRingOff + s * kEncStageBytes, map, (int32_t)row, rc.full + s);
                    else
                        tma_load_block(smem + kEncRingOff + s * kEncStageBytes, map, (int32_t)row, rc.full + s,
                                       (d.kind == kWorkA && (p.dbg & 2)) ? keep : drop);
                } else {
                    mbar_arrive(rc.full + s);
                }
                filled = true;
            }
            if (__all_sync(kFull, filled)) break;
            if ((spins & 1023u) == 1023u) {  // bounded like every other wait
                unsigned long long t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                if (t0 == 0) t0 = t;
                else if (t - t0 > kWaitLimitNs) {
                    if (!filled) spin_timeout(p.error, 4, -1);
                    return;
                }
            }
        }
        pass++;
    }
    if (lane < kEncConsumers) {  // one stop marker per consumer warp
        const uint32_t seq = pass * kGroupSlots + (uint32_t)lane;
        const int s = (int)(seq % kEncStages);
        const uint32_t use = seq / kEncStages;
        if (use > 0 && !mbar_wait(p.error, rc.empty + s, (use - 1) & 1, 4)) return;
        WorkDesc d;
        d.kind = kWorkStop;
        d.f = d.comp = d.b = 0;
        d.blk = -1;
        d.len = 0;
        d.flags = kDescSlow;
        d.scale = 0.0;
        d.x0 = d.tHi = d.tLo = d.L = 0.0f;
        d.depth = d.pad = 0;
        d.payBase = d.pad2 = 0;
        rc.desc[s] = d;
        mbar_arrive(rc.full + s);
    }
}

}  // namespace mnw

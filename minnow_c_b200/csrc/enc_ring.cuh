// enc_ring.cuh -- the input ring of k_encode. The planner warp of a CTA draws
// group tickets; each of its 32 lanes owns one slot of the group: it works out the
// slot's block and writes the work descriptor into a queue indexed by sequence
// number. The issuer (one thread) walks the sequence numbers in order and, the
// moment the next one's ring stage has been handed back, has the TMA unit copy the
// block's 4 KiB into the stage (2D tensor map over the input as rows of 32 floats,
// 128-byte swizzle, so that a consumer lane can read its 32 consecutive values
// without bank conflicts). Consumer warps claim sequence numbers in order, wait on
// the stage's "full" mbarrier and hand the stage back through its "empty" mbarrier.
// A parity wait can only tell the current phase from the one before, and consumers
// racing through no-op slots can claim the next use of a stage whose current fill is
// still in flight: so a consumer first waits until the stage's previous use has been
// handed back (a plain counter), and only then for the fill.
#pragma once

#include "enc_math.cuh"

namespace mnw {

struct WorkDesc {  // 80 bytes, one per sequence number (queue of kDescQ): the block and what its component needs
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
    int64_t row;      // first row of the block in the tensor map (the issuer's TMA coordinate)
};
enum : uint32_t {
    kDescSlow = 1u,   // the block did not go through the TMA: the consumer loads it itself
    kDescPos = 2u,    // a position component: periodic unwrap applies
    kDescQuant = 4u,  // scale, x0 and depth are filled in (else the consumer waits for them)
};

struct RingCtrl {
    uint64_t *full;   // [kEncStages] descriptor written and TMA bytes landed
    uint64_t *empty;  // [kEncStages] the consumer is done with the stage
    WorkDesc *desc;   // [kDescQ] by sequence number modulo kDescQ
    uint32_t *qHead;  // [0] next sequence number to claim (consumers, atomic); [1] sequence numbers planned
                      // (planner); [2] issued (issuer); [3] one past the last sequence number, once known
    volatile uint32_t *uses;  // [kEncStages] how many times the stage has been handed back
    long long *tIssue, *tRel;  // MNWCUDA_ENCODE_STATS: per stage, clock of the last fill / hand-back (else null)
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
// ask L2 for a block ahead of its TMA copy (no completion tracking: a hint)
__device__ __forceinline__ void tma_prefetch_block(const CUtensorMap *map, int32_t row) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];" ::"l"(reinterpret_cast<uint64_t>(map)),
                 "r"(0), "r"(row)
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// atomicAdd on a shared word by one elected lane, result to all lanes (inline asm: the
// compiler otherwise wraps a one-lane atomicAdd in its warp-aggregation sequence)
__device__ __forceinline__ uint32_t warp_atom_add_shared(uint32_t *word, uint32_t v) {
    uint32_t old = 0;
    asm volatile("{\n\t"
                 ".reg .pred one;\n\t"
                 "elect.sync _|one, 0xffffffff;\n\t"
                 "@one atom.shared.add.u32 %0, [%1], %2;\n\t"
                 "}"
                 : "+r"(old) : "r"((uint32_t)__cvta_generic_to_shared(word)), "r"(v) : "memory");
    return __shfl_sync(0xffffffffu, old, 0);
}

__device__ __forceinline__ RingCtrl ring_ctrl(uint8_t *smem) {
    RingCtrl rc;
    uint8_t *c = smem + kEncCtrlOff;
    rc.full = reinterpret_cast<uint64_t *>(c);
    rc.empty = rc.full + kEncStages;
    rc.desc = reinterpret_cast<WorkDesc *>(rc.empty + kEncStages);
    rc.qHead = reinterpret_cast<uint32_t *>(rc.desc + kDescQ);
    rc.uses = rc.qHead + 4;
    rc.tIssue = rc.tRel = nullptr;
    return rc;
}

// ---------------------------------------------------------------------------
// planner warp: all 32 lanes, one slot each per pass. It draws the group tickets,
// works out the 32 descriptors of a pass (block location, component constants,
// unwrap thresholds, the published quantisation parameters: global loads) and puts
// them into the descriptor queue, up to kDescQ - kEncStages sequence numbers ahead
// of the issuer. It never touches the ring: a stage that has been handed back must
// not wait for descriptor arithmetic.
// ---------------------------------------------------------------------------
template <bool kFromQ, bool kStats>
__device__ __forceinline__ void planner(const EncodeParams &p, const CUtensorMap *map, const RingCtrl &rc, int lane,
                                        int nConsumers) {
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
    volatile uint32_t *planned = rc.qHead + 1, *issued = rc.qHead + 2, *end = rc.qHead + 3;
    uint32_t pass = 0;
    uint32_t tk = 0;
    if (lane == 0) tk = atomicAdd(p.ticket, 1u);
    bool stop = false;
    while (true) {
        const int64_t g = (int64_t)__shfl_sync(kFull, tk, 0);
        stop = g >= p.nGroups;
        if (!stop && lane == 0) tk = atomicAdd(p.ticket, 1u);  // the next ticket is on its way meanwhile
        const long long tp0 = kStats ? clock64() : 0;
        WorkDesc d;
        d.kind = kWorkNop;
        d.f = d.comp = d.b = 0;
        d.len = 0;
        d.flags = kDescSlow;
        d.scale = 0.0;
        d.x0 = d.tHi = d.tLo = d.L = 0.0f;
        d.depth = 0;
        d.pad = 0;
        d.payBase = d.row = 0;
        int64_t blk = -1;
        if (stop) {
            if (lane < nConsumers) d.kind = kWorkStop;  // the last pass: one stop marker per consumer warp
        } else if (onlyB) {
            blk = g * kGroupSlots + lane;
            if (blk < p.nBlocks) d.kind = kWorkB;
        } else if (lane < kGroupBlocks) {
            blk = g * kGroupBlocks + lane;
            if (blk < p.nBlocks) d.kind = kWorkA;
        } else {
            blk = (g - p.lagG) * kGroupBlocks + (lane - kGroupBlocks);
            if (blk >= 0 && blk < p.nBlocks && !(p.dbg & 64)) d.kind = kWorkB;  // (64: A work alone, timing experiment)
        }
        d.blk = blk;
        if (d.kind == kWorkA || d.kind == kWorkB) {
            locate_block(p, blk, d.f, d.comp, d.b);
            const DevField *fdp = p.fields + d.f;
            const int n = fdp->n;
            d.len = min(kBlockLen, n - d.b * kBlockLen);
            d.payBase = fdp->fixedBase + 64 + (int64_t)fdp->nb * 48;
            const int64_t off = fdp->ioOff + (int64_t)d.comp * n + (int64_t)d.b * kBlockLen;  // floats from p.in
            d.row = (p.tmaShift + off) >> 5;
            const bool fast = !kFromQ && p.tmaShift >= 0 && d.len == kBlockLen && ((p.tmaShift + off) & 31) == 0 &&
                              d.row + 32 <= p.tmaRows;
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
            // MNWCUDA_DBG & 16 / & 32: L2 prefetch of the A / B blocks at planning time (experiment)
            if (fast && ((d.kind == kWorkA && (p.dbg & 16)) || (d.kind == kWorkB && (p.dbg & 32))))
                tma_prefetch_block(map, (int32_t)d.row);
            d.flags = pc.flags | (fast ? 0u : kDescSlow);
            d.scale = pc.scale;
            d.x0 = pc.x0;
            d.depth = pc.depth;
            d.tHi = pc.tHi;
            d.tLo = pc.tLo;
            d.L = pc.L;
        }
        // room in the queue: an entry is free once its stage's next use can be issued
        const uint32_t seq = pass * kGroupSlots + (uint32_t)lane;
        const long long tp1 = kStats ? clock64() : 0;
        for (SpinClock clk; pass * kGroupSlots + kGroupSlots > *issued + (uint32_t)(kDescQ - kEncStages);) {
            __nanosleep(64);
            if (spin_expired(clk)) { spin_timeout(p.error, 8, -1); return; }
        }
        rc.desc[seq % kDescQ] = d;
        __threadfence_block();
        __syncwarp();
        if (lane == 0) {
            *planned = (pass + 1) * kGroupSlots;
            if (stop) *end = (pass + 1) * kGroupSlots;
        }
        if (kStats && p.stats && lane == 0) {
            atomicAdd(p.stats + 22, (unsigned long long)(tp1 - tp0));
            atomicAdd(p.stats + 23, (unsigned long long)(clock64() - tp1));
            atomicAdd(p.stats + 24, 1ull);
        }
        pass++;
        if (stop) return;
    }
}

// ---------------------------------------------------------------------------
// issuer warp: lane l owns the sequence numbers l, l + 32, ... Round by round, every
// lane starts the TMA copy of its block the moment its ring stage has been handed
// back (or just completes the stage's "full" barrier for a block that does not go
// through the TMA); a lane does not wait for the slowest stage of the round (a B
// block holds one longer than an A block). Nothing else happens on this path: the
// descriptors were worked out ahead of time by the planner.
// ---------------------------------------------------------------------------
// (kPrefetch: a block the TMA cannot fetch -- off the 16-byte grid -- is at least asked into L2 here, a ring's
// depth before its consumer loads it by hand; the general kernel only)
template <bool kStats, bool kPrefetch = false>
__device__ __forceinline__ void issuer(const EncodeParams &p, const CUtensorMap *map, uint8_t *smem,
                                       const RingCtrl &rc, int lane) {
    // A reads carry no cache hint, B reads are evict_first: with the B work ~56 MB of
    // A traffic behind, about two thirds of the B reads still hit in L2 (evict_last on
    // the A reads thrashes: nothing survives)
    const uint64_t drop = l2_policy_drop();
    volatile uint32_t *planned = rc.qHead + 1, *issued = rc.qHead + 2, *end = rc.qHead + 3;
    long long refill = 0, nRefill = 0;
    for (uint32_t round = 0;; round++) {
        const uint32_t seq = round * kGroupSlots + (uint32_t)lane;
        for (SpinClock clk; (round + 1) * kGroupSlots > *planned;) {
            if (round * kGroupSlots >= *end) {
                if (kStats && p.stats) {
                    atomicAdd(p.stats + 20, (unsigned long long)refill);
                    atomicAdd(p.stats + 21, (unsigned long long)nRefill);
                }
                return;
            }
            if (spin_expired(clk)) { spin_timeout(p.error, 9, -1); return; }
        }
        __threadfence_block();
        const int s = (int)(seq % kEncStages);
        const uint32_t use = seq / kEncStages;
        const WorkDesc *d = rc.desc + seq % kDescQ;
        const int kind = d->kind;
        const bool fast = (kind == kWorkA || kind == kWorkB) && !(d->flags & kDescSlow);
        const int32_t row = (int32_t)d->row;
        bool filled = false;
        unsigned long long t0 = 0;
        for (uint32_t spins = 0;; spins++) {
            if (!filled && (use == 0 || (rc.uses[s] >= use && mbar_try_wait(rc.empty + s, (use - 1) & 1)))) {
                if (kStats && rc.tIssue) {
                    const long long now = clock64();
                    if (use > 0) { refill += now - rc.tRel[s]; nRefill++; }
                    rc.tIssue[s] = now;
                }
                if (fast) {
                    mbar_arrive_expect_tx(rc.full + s, kEncStageBytes);
                    if (kind == kWorkA) tma_load_block_plain(smem + kEncRingOff + s * kEncStageBytes, map, row, rc.full + s);
                    else tma_load_block(smem + kEncRingOff + s * kEncStageBytes, map, row, rc.full + s, drop);
                } else {
                    if (kPrefetch && (kind == kWorkA || kind == kWorkB) && p.tmaShift >= 0 && row >= 0 && (int64_t)row + 32 <= p.tmaRows)
                        tma_prefetch_block(map, row);
                    mbar_arrive(rc.full + s);
                }
                filled = true;
            }
            if (__all_sync(kFull, filled)) break;
            __nanosleep(32);  // (a hot spin would take issue slots from the consumers of this scheduler)
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
        if (lane == 0) *issued = (round + 1) * kGroupSlots;
    }
}

}  // namespace mnw

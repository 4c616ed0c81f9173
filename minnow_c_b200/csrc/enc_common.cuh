// enc_common.cuh -- constants, parameters and small helpers of the encoder
// (k_encode + k_finalize). Replaces, per field, position()/velocity() of
// /root/reference/src/quant.c:161-289 and the (stub) CFunc behind
// src/funcs.c:71. HBM-bound float/integer streaming: no tensor cores.
//
// Work is cut into 1024-value blocks. Blocks are numbered over the whole batch
// (field-major, then component, then block). Global work "groups" of 32 slots
// are handed out by one ticket counter: slots 0..15 of group g are the A work
// (unwrap + min/max) of blocks 16g..16g+15, slots 16..31 the B work (quantise +
// delta + width + pack) of blocks 16(g-lagG)..16(g-lagG)+15 -- lagG groups
// later, when the field's parameters are known and, with luck, the block is
// still in L2. Everything that crosses warps goes through self-validating
// words (a value plus a "there" bit, zero-initialised, relaxed accesses): block
// min/max -> the reduce warp -> per-component parameters; block sizes -> the
// scan warp -> block prefixes. No fences on the data path.
#pragma once

#include <cuda.h>

#include "bitpack.cuh"
#include "minnow_kernels.cuh"

namespace mnw {

constexpr unsigned kFull = 0xffffffffu;
#ifndef MNW_ENC_CONSUMERS
#define MNW_ENC_CONSUMERS 22
#endif
constexpr int kEncConsumers = MNW_ENC_CONSUMERS;
constexpr int kEncWarps = kEncConsumers + 2;  // per CTA (one CTA per SM): planner, issuer, consumers; the last warp of CTA 0 is the scan
                                              // warp, of CTA 1 the reduce warp (elsewhere it is a consumer like the others)
#ifndef MNW_DESCQ
#define MNW_DESCQ 96
#endif
constexpr int kDescQ = MNW_DESCQ;           // descriptor queue: the planner runs up to kDescQ - kEncStages slots ahead of the issuer
constexpr int kEncStageBytes = 4096;
constexpr int kGroupSlots = 32;      // slots per ticket
constexpr int kGroupBlocks = 16;     // A blocks (and B blocks) per group
constexpr int kFifo = 8;             // parked blocks a consumer can remember
constexpr int kArenaWords = 832;     // per consumer: packed blocks waiting for their offset, in shared memory
                                     // (one block of the widest kind, 33 x 25 + 1 words, or four typical ones)
// input ring: 4 KiB blocks (TMA, 128-byte swizzle), as many stages as the SM's shared memory holds
// next to the consumers' arenas and records
constexpr int kEncPerConsumer = kFifo * (16 + 8 + 8) + 64 + kArenaWords * 4;
constexpr int kEncTimelineBytes = 1024;  // MNWCUDA_ENCODE_STATS: issue / hand-back clocks of up to 64 stages
constexpr int kEncStagesFit = (232448 - 32 - kEncTimelineBytes - kDescQ * 80 - kEncConsumers * kEncPerConsumer) / (kEncStageBytes + 2 * 8 + 4);
#ifdef MNW_ENC_STAGES
constexpr int kEncStages = MNW_ENC_STAGES;
#else
constexpr int kEncStages = kEncStagesFit;
#endif
constexpr int kEncRingOff = 0;
constexpr int kEncCtrlOff = kEncStages * kEncStageBytes;
constexpr int kEncCtrlBytes = (2 * kEncStages * 8 + kDescQ * 80 + 16 + kEncStages * 4 + 15) & ~15;
constexpr int kEncFifoOff = kEncCtrlOff + kEncCtrlBytes;   // per consumer: records of its parked blocks
constexpr int kEncFifoBytes = kEncConsumers * kFifo * (16 + 8 + 8);
constexpr int kEncStatOff = kEncFifoOff + kEncFifoBytes;   // per consumer: 8 phase clocks (MNWCUDA_ENCODE_STATS)
constexpr int kEncArenaOff = kEncStatOff + kEncConsumers * 64;
constexpr int kEncTimelineOff = kEncArenaOff + kEncConsumers * kArenaWords * 4;
constexpr int kEncSmemBytes = kEncTimelineOff + kEncTimelineBytes;
static_assert(kEncStages <= 64, "k_encode: the stage timeline holds 64 stages");
static_assert(kEncSmemBytes <= 232448, "k_encode: more shared memory than an SM has");
static_assert(kArenaWords >= 33 * 25 + 1, "k_encode: the arena must hold one block of any width");
static_assert(kEncStages >= 16, "k_encode: too few ring stages");
static_assert(kDescQ >= kEncStages + kGroupSlots, "k_encode: the descriptor queue must hold a pass beyond the ring");
constexpr int kSlotWords = 832;      // scratch slot of one packed block: 25 bits x 1024, 16-byte padded

constexpr uint32_t kAggReady = 0x8000u;       // in a 16-bit size word
constexpr uint32_t kRelReady = 0x80000000u;   // in a 32-bit window-relative prefix
constexpr uint64_t kPrefixReady = 1ull << 63; // in a 64-bit window base
#ifndef MNW_SCAN_WIN
#define MNW_SCAN_WIN 1024
#endif
constexpr int kScanWin = MNW_SCAN_WIN;        // blocks per scan window (a multiple of 256)
constexpr int kScanRows = kScanWin / 256;     // 16-byte loads per lane and window
constexpr int kRedWin = 512;                  // blocks per reduce window
constexpr uint64_t kFqValid = 1ull << 63;
constexpr uint32_t kQBias = 0x4b000000u;  // bit pattern of 2^23: trunc by add.rz keeps it in the exponent

enum : int { kWorkNop = 0, kWorkA = 1, kWorkB = 2, kWorkStop = 3 };

struct EncodeParams {
    const float *in;
    const uint64_t *qIn;   // literal CFunc boundary: quantised values in, or null
    const DevField *fields;
    int nFields;
    int uniformBlocks;     // >0: every field has this many blocks (fast lookup)
    uint32_t *minmax;      // [nFields][3][2] ordered-uint min, ~max
    DevQuant *quant;       // [nFields]
    uint16_t *blkAgg;      // [nBlocks, padded to kScanWin] kAggReady | packed size of the block, 16-byte units
    uint32_t *blkRel;      // [nBlocks, padded] kRelReady | exclusive prefix within the block's scan window
    uint64_t *winBase;     // [nWin+1] kPrefixReady | exclusive prefix at the window's start; [nWin] = total
    uint32_t *ticket;      // group ticket counter
    uint4 *tbl;            // [nBlocks] First, MinDelta, Width of the block (k_tables adds Offset16)
    uint64_t *blkMM;       // [nBlocks, padded to kRedWin] ord(min) | ~ord(max) << 32 of the block, 0 = not there yet
    uint64_t *fq;          // [nFields][3][2] published parameters of a component (see publish_quant)
    int offGrid;           // host: some block of the batch does not start on the 128-byte grid (which kernel to launch)
    uint64_t *winMM;       // [nBlocks / kRedWin] the same per reduce window (parallel reduce), 0 = not there yet
    uint32_t *scratch;     // [CTAs][kEncConsumers][kFifo][kSlotWords] packed blocks waiting for their offset
    int64_t tmaShift;      // floats between the tensor map's 128-byte aligned base and `in`; <0: no TMA
    int64_t tmaRows;       // rows of 32 floats covered by the tensor map
    uint8_t *out;
    int64_t outCap;
    uint64_t *offsets;     // [nFields+1]
    int *error;
    int64_t nBlocks;
    int64_t nGroups;       // tickets >= nGroups mean "no more work"
    int64_t lagG;          // B of group g-lagG rides with A of group g; <0: no A work (quant given)
    int dbg;               // MNWCUDA_DBG: debugging switches
    unsigned long long *stats;  // MNWCUDA_ENCODE_STATS: per-phase clock totals over all warps, or null
};

// ---- memory-ordering helpers ----
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint4 ld_relaxed_u4(const void *p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u4(void *p, uint4 v) {
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
// exclusive prefix (16-byte units) of block `blk`, if the scan has got there
__device__ __forceinline__ bool block_prefix(const uint32_t *blkRel, const uint64_t *winBase, int64_t blk,
                                             uint64_t &pre) {
    uint32_t rel;
    uint64_t wb;  // two independent loads: one round trip
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(rel) : "l"(blkRel + blk) : "memory");
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(wb) : "l"(winBase + blk / kScanWin) : "memory");
    pre = (wb & ~kPrefixReady) + (rel & ~kRelReady);
    return (rel & kRelReady) && (wb & kPrefixReady);
}
__device__ __forceinline__ uint64_t l2_policy_keep() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_policy_drop() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void st_stream_u4(uint4 *p, uint4 v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

// every wait on another warp's work gives up after kWaitLimitNs of wall clock
// (%globaltimer, looked at every 256 polls) and raises kErrTimeout instead of
// hanging the device: error[1] = site, error[2..3] = block. The limit is far above
// anything a healthy launch waits for (a whole 24 GiB batch encodes in 10 ms), also
// under a profiler's serialised replay.
constexpr unsigned long long kWaitLimitNs = 4000000000ull;
struct SpinClock {
    uint32_t spins = 0;
    unsigned long long t0 = 0;
};
__device__ __forceinline__ bool spin_expired(SpinClock &c) {
    if ((++c.spins & 255u) != 0) return false;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (c.t0 == 0) {
        c.t0 = t;
        return false;
    }
    return t - c.t0 > kWaitLimitNs;
}
__device__ __noinline__ void spin_timeout(int *error, int site, int64_t blk) {
    if (atomicCAS(error, 0, kErrTimeout) == 0) {
        error[1] = site;
        error[2] = (int)(blk & 0xffffffff);
        error[3] = (int)(blk >> 32);
    }
}

// global block index -> field, component, block within the component
__device__ __forceinline__ void locate_block(const EncodeParams &p, int64_t blk, int &f, int &comp, int &b) {
    if (p.uniformBlocks > 0) {
        f = (int)((uint32_t)blk / (uint32_t)p.uniformBlocks);
    } else {
        const DevField *fields = p.fields;
        f = find_field(p.nFields, blk, [fields](int i) { return fields[i].firstBlock; });
    }
    const int local = (int)(blk - p.fields[f].firstBlock);
    const int nb = p.fields[f].nb;
    comp = local / nb;
    b = local - comp * nb;
}

}  // namespace mnw

// enc_entry.cuh -- the kernels' entry points. Included by enc_consume.cuh.
#pragma once

#include "enc_consume.cuh"

namespace mnw {

// ---------------------------------------------------------------------------
// k_encode: warp 0 = planner, warp 1 = issuer, then the consumers, the last warp = the scan warp
// in CTA 0 and the reduce warp in CTA 1 (the host launches at least two CTAs).
// ---------------------------------------------------------------------------
// kStats: the per-phase clocks of MNWCUDA_ENCODE_STATS; a separate instantiation so
// that the normal kernel carries none of the clock reads and 64-bit adds (they were
// a few per cent of the consumers' instructions).
// kGeneral: the launch for big batches whose fields differ in size (the general parallel
// reduce, enc_kernel.cuh) -- an instantiation of its own so that the kernel of the uniform
// batches stays exactly what it is (its time moves with anything that changes its service
// warps' code).
// kOffGrid (with kGeneral): batches with blocks off the 128-byte grid -- vector loads by hand in the A and B
// work, L2 prefetch by the issuer. Again an instantiation of its own: that code costs the others 2 %.
template <bool kFromQ, bool kStats, bool kGeneral = false, bool kOffGrid = false>
__global__ void __launch_bounds__(kEncWarps * 32, 1)
k_encode(EncodeParams p, const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(1024) uint8_t encSmem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    RingCtrl rc = ring_ctrl(encSmem);
    if (kStats && p.stats) {
        rc.tIssue = reinterpret_cast<long long *>(encSmem + kEncTimelineOff);
        rc.tRel = rc.tIssue + 64;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < kEncStages; s++) {
            mbar_init(rc.full + s, 1);
            mbar_init(rc.empty + s, 1);
        }
        rc.qHead[0] = rc.qHead[1] = rc.qHead[2] = 0;  // claimed, planned, issued
        rc.qHead[3] = 0xffffffffu;                    // end of the sequence: not known yet
        for (int s = 0; s < kEncStages; s++) rc.uses[s] = 0;
        if (kGeneral) {  // quant_workers' state (CTA 1 has no ring traffic: the stages are free)
            volatile int *q = reinterpret_cast<volatile int *>(encSmem + kEncRingOff + kRedWarpsGeneral * kEncStageBytes);
            q[0] = q[1] = q[2] = 0;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async();
    }
    __syncthreads();
    // with enough CTAs the first two do nothing but scan and reduce: a service warp that
    // shares its SM with 18 busy consumers falls behind, and everybody waits for it
    const bool service = gridDim.x >= 8 && blockIdx.x < 2;
    if (service && warp != kEncWarps - 1) {
        // CTA 1 of a big launch: its other warps reduce the windows in parallel
        if (blockIdx.x == 1 && p.lagG >= 0 && !kFromQ) {
            if (kGeneral) {
                if (warp < kRedWarpsGeneral)
                    reduce_windows_general(p, reinterpret_cast<uint2 *>(encSmem + kEncRingOff + warp * kEncStageBytes), lane, warp);
                else
                    quant_workers(p, reinterpret_cast<volatile int *>(encSmem + kEncRingOff + kRedWarpsGeneral * kEncStageBytes), lane);
            } else if (warp < kRedWarps && reduce_in_parallel(p)) {
                reduce_windows(p, lane, warp);
            }
        }
        return;
    }
    // the last warp: scan in CTA 0, reduce in CTA 1, one more consumer everywhere else
    const bool lastIsService = blockIdx.x < 2;
    if (warp == 0) {
        planner<kFromQ, kStats>(p, &tmap, rc, lane, lastIsService ? kEncConsumers - 1 : kEncConsumers);
        return;
    }
    if (warp == 1) {
        issuer<kStats, kOffGrid>(p, &tmap, encSmem, rc, lane);
        return;
    }
    if (warp == kEncWarps - 1 && lastIsService) {
        if (blockIdx.x == 0) scan_blocks(p, lane);
        else if (p.lagG >= 0 && !kFromQ) {
            if (kGeneral)  // (the host launches it with 8 CTAs or more only)
                reduce_fold_general(p, reinterpret_cast<volatile int *>(encSmem + kEncRingOff + kRedWarpsGeneral * kEncStageBytes), lane);
            else if (gridDim.x >= 8 && reduce_in_parallel(p)) reduce_fold(p, lane);
            else reduce_fields(p, lane);
        }
        return;
    }
    const int c = warp - 2;
    ParkFifo ff;
    ff.where = reinterpret_cast<int4 *>(encSmem + kEncFifoOff) + c * kFifo;
    ff.blk = reinterpret_cast<int64_t *>(encSmem + kEncFifoOff + kEncConsumers * kFifo * 16) + c * kFifo;
    ff.pay = reinterpret_cast<int64_t *>(encSmem + kEncFifoOff + kEncConsumers * kFifo * 24) + c * kFifo;
    ff.slots = p.scratch + ((size_t)blockIdx.x * kEncConsumers + c) * kFifo * kSlotWords;
    ff.arena = reinterpret_cast<uint32_t *>(encSmem + kEncArenaOff) + c * kArenaWords;
    ff.head = ff.count = 0;
    ff.aTail = ff.aUsed = ff.resident = 0;
    ff.wst = nullptr;
    if (kStats && p.stats) {
        ff.wst = reinterpret_cast<unsigned long long *>(encSmem + kEncStatOff) + c * 8;
        if (lane == 0)
            for (int i = 0; i < 8; i++) ff.wst[i] = 0;
        __syncwarp();
    }
    bool drain = false;
    long long tA = 0, tB = 0, tQueue = 0, nA = 0, nB = 0, tFill = 0, nFill = 0;
    const long long tStart = kStats ? clock64() : 0;
    while (true) {
        const long long c0 = kStats ? clock64() : 0;
        // (inline asm: the compiler otherwise wraps the one-lane atomicAdd in its
        // warp-aggregation sequence, a dozen instructions per claim)
        uint32_t seq = 0;
        asm volatile("{\n\t"
                     ".reg .pred one;\n\t"
                     "elect.sync _|one, 0xffffffff;\n\t"
                     "@one atom.shared.add.u32 %0, [%1], 1;\n\t"
                     "}"
                     : "+r"(seq) : "r"((uint32_t)__cvta_generic_to_shared(rc.qHead)) : "memory");
        seq = __shfl_sync(kFull, seq, 0);
        const int s = (int)(seq % kEncStages);
        const uint32_t use = seq / kEncStages;
        for (SpinClock clk; rc.uses[s] < use;) {
            if (spin_expired(clk)) { spin_timeout(p.error, 7, -1); return; }
        }
        if (!mbar_wait(p.error, rc.full + s, use & 1u, 5)) return;
        WorkDesc d = rc.desc[seq % kDescQ];
        float *stage = reinterpret_cast<float *>(encSmem + kEncRingOff + s * kEncStageBytes);
        const long long c1 = kStats ? clock64() : 0;
        if (kStats) {
            if (rc.tIssue && d.kind != kWorkStop) {  // stage timeline: fill -> data seen by its consumer
                tFill += c1 - rc.tIssue[s];
                nFill++;
            }
            tQueue += c1 - c0;
            if (ff.wst && lane == 0 && d.kind == kWorkA) ff.wst[6] += c1 - c0;  // stats: the part spent waiting for A blocks
        }
        if (d.kind == kWorkStop) break;
        if (drain) {  // after a timeout: hand the stages back untouched so that the kernel can end
            release_stage(rc, s, false, lane);
            continue;
        }
        if (d.kind == kWorkA) {
            a_block<kOffGrid>(p, d, stage, lane);
            release_stage(rc, s, false, lane);
            if (kStats) {
                tA += clock64() - c1;
                nA++;
            }
        } else if (d.kind == kWorkB) {
            if (!b_block<kFromQ, kStats, kOffGrid>(p, d, stage, rc, s, ff, lane)) {
                drain = true;  // (b_block hands its stage back before anything that can time out)
                continue;
            }
            if (kStats) {
                tB += clock64() - c1;
                nB++;
            }
        } else {
            release_stage(rc, s, false, lane);
        }
    }
    while (ff.count > 0 && !drain)
        if (!place_oldest(p, ff, lane)) return;
    if (kStats && p.stats && lane == 0) {
        atomicAdd(p.stats + 0, (unsigned long long)tA);
        atomicAdd(p.stats + 1, (unsigned long long)tB);
        atomicAdd(p.stats + 4, (unsigned long long)tQueue);
        atomicAdd(p.stats + 5, (unsigned long long)nA);
        atomicAdd(p.stats + 6, (unsigned long long)nB);
        atomicAdd(p.stats + 7, (unsigned long long)(clock64() - tStart));
        for (int i = 0; i < 8; i++) atomicAdd(p.stats + 8 + i, ff.wst[i]);
        atomicAdd(p.stats + 16, (unsigned long long)tFill);
        atomicAdd(p.stats + 17, (unsigned long long)nFill);
    }
}

// ---------------------------------------------------------------------------
// k_tables: one thread per block copies the block's table entry (written at
// packing time) to its place in the field's block table and adds Offset16, the
// block's prefix relative to the field's first block.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_tables(EncodeParams p) {
    const int64_t blk = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (blk >= p.nBlocks) return;
    int f, comp, b;
    locate_block(p, blk, f, comp, b);
    const DevField *fdp = p.fields + f;
    const int64_t fb = fdp->firstBlock;
    const uint64_t p0 = (p.winBase[fb / kScanWin] & ~kPrefixReady) + (p.blkRel[fb] & ~kRelReady);
    const uint64_t pre = (p.winBase[blk / kScanWin] & ~kPrefixReady) + (p.blkRel[blk] & ~kRelReady);
    const int64_t entry = fdp->fixedBase + (int64_t)(p0 * 16) + 64 + ((int64_t)comp * fdp->nb + b) * 16;
    if (entry + 16 > p.outCap) { atomicCAS(p.error, 0, kErrCapacity); return; }
    uint4 e = p.tbl[blk];
    e.w = (uint32_t)(pre - p0);
    *reinterpret_cast<uint4 *>(p.out + entry) = e;
}

// ---------------------------------------------------------------------------
// k_finalize: one thread per field writes the 64-byte header and the field's
// byte offset (the block tables were written at placement).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_finalize(EncodeParams p, int64_t fixedTotal) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f > p.nFields) return;
    if (f == p.nFields) {
        p.offsets[f] = (uint64_t)fixedTotal + (p.winBase[(p.nBlocks + kScanWin - 1) / kScanWin] & ~kPrefixReady) * 16;
        return;
    }
    const DevField fd = p.fields[f];
    // an empty field at the very end has no first block: its region starts at the total
    const uint64_t p0 = fd.firstBlock < p.nBlocks
                            ? (p.winBase[fd.firstBlock / kScanWin] & ~kPrefixReady) + (p.blkRel[fd.firstBlock] & ~kRelReady)
                            : (p.winBase[(p.nBlocks + kScanWin - 1) / kScanWin] & ~kPrefixReady);
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
    h.Width = (fd.fieldCode == field_Posn || fd.fieldCode == field_Ptid) ? fd.width : 0.0f;  // (Ptid: uint32 bits)
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

// enc_kernel.cuh -- k_encode (persistent, one CTA per SM: a TMA producer warp,
// 22 consumer warps, a service warp: scan in CTA 0, reduce in CTA 1) and k_finalize.
//
//   A work   util_UndoPeriodic + util_MinMax of one block (src/util.c:97-113,
//            27-46) straight out of the ring; the block's min/max go out as
//            one 64-bit word.
//   reduce   one warp (CTA 1) folds the blocks' min/max in block order; at the
//            end of a field it runs deltaToDepth (src/quant.c:186-196, 684-693)
//            and publishes the field's quantisation parameters.
//   B work   util_UniformBinIndex + delta + width + util_U32UniformPack of one
//            block (src/util.c:173-196, 311-355). The size is published at once;
//            the packed words go through the block's own ring stage into one of
//            the consumer's scratch slots in global memory (L2 resident).
//   scan     one warp turns the sizes into exclusive prefixes, in block order.
//   place    a few blocks later the same consumer copies the block from its
//            slot to its final position and writes its table entry.
// A work never waits; B work waits for A work with smaller tickets only (through
// the reduce warp); placement waits for B work of smaller blocks only (through
// the scan warp): the waits cannot deadlock.
#pragma once

#include "enc_math.cuh"
#include "enc_ring.cuh"

namespace mnw {

// ---------------------------------------------------------------------------
// maxDiff and deltaToDepth of one field (src/quant.c:186-196, 684-693), run by
// a single thread once every min/max of the field is final.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void compute_quant(const EncodeParams &p, int f) {
    const DevField fd = p.fields[f];
    DevQuant q;
    float md = 0.0f;
    for (int c = 0; c < 3; c++) {
        float lo = 0.0f, hi = 0.0f;
        if (fd.n > 0) {
            lo = float_from_ord(__ldcg(p.minmax + (f * 3 + c) * 2));
            hi = float_from_ord(~__ldcg(p.minmax + (f * 3 + c) * 2 + 1));
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
    // a constant field has maxDiff = 0: util_UniformBinIndex's 0/0 falls through
    // both clamps and converts to 0 on the host; scale 0 gives the same zeros
    q.scale = md > 0.0f ? __ddiv_rn((double)(1u << depth), (double)md) : 0.0;
    p.quant[f] = q;
    // what B work needs, as two self-validating words per component: the scale
    // (a non-negative double: its sign bit is free) and X0 | depth << 32
    if (p.fq) {
        for (int c = 0; c < 3; c++) {
            uint64_t *w = p.fq + ((int64_t)f * 3 + c) * 2;
            const uint64_t w0 = kFqValid | (uint64_t)__double_as_longlong(q.scale);
            const uint64_t w1 = kFqValid | ((uint64_t)depth << 32) | (uint64_t)__float_as_uint(q.X0[c]);
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(w), "l"(w0) : "memory");
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(w + 1), "l"(w1) : "memory");
        }
    }
}

// one thread per field: empty fields have no blocks to trigger compute_quant
__global__ void k_depth_empty(EncodeParams p) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nFields) return;
    if (p.fields[f].n == 0) compute_quant(p, f);
}

// ---------------------------------------------------------------------------
// Scan warp: block sizes -> exclusive prefixes, one window of 1024 blocks at a
// time with the next window's loads already in flight. Sizes are 16-bit words
// (8 blocks per 16-byte load); a block's prefix goes out as a 32-bit offset
// within its window plus one 64-bit base per window. The words carry their own
// "there" bits and nobody reads anything else through them: relaxed accesses.
// Lane l, load j covers blocks base + 256 j + 8 l .. +7 (coalesced).
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool agg_ready(const uint4 &v) {
    return (v.x & v.y & v.z & v.w & 0x80008000u) == 0x80008000u;
}
__device__ __forceinline__ void scan_load(const EncodeParams &p, int64_t base, int lane, uint4 (&a)[kScanRows]) {
#pragma unroll
    for (int j = 0; j < kScanRows; j++) a[j] = ld_relaxed_u4(p.blkAgg + base + j * 256 + lane * 8);
}

__device__ __forceinline__ void scan_blocks(const EncodeParams &p, int lane) {
    if (p.dbg & 64) return;  // (A work alone, timing experiment: no block ever gets a size)
    uint64_t running = 0;
    uint4 cur[kScanRows], nxt[kScanRows];
    scan_load(p, 0, lane, cur);
    int64_t w = 0;
    for (int64_t base = 0; base < p.nBlocks; base += kScanWin, w++) {
        if (lane == 0)
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p.winBase + w), "l"(kPrefixReady | running) : "memory");
        const bool more = base + kScanWin < p.nBlocks;
        if (more) scan_load(p, base + kScanWin, lane, nxt);
        if (base + kScanWin > p.nBlocks) {
            // last, ragged window: blocks past the end count as there, with size 0
#pragma unroll
            for (int j = 0; j < kScanRows; j++) {
                uint32_t *wds = reinterpret_cast<uint32_t *>(&cur[j]);
#pragma unroll
                for (int i = 0; i < 8; i++)
                    if (base + j * 256 + lane * 8 + i >= p.nBlocks) wds[i >> 1] = (wds[i >> 1] & ~(0xffffu << (16 * (i & 1)))) | (kAggReady << (16 * (i & 1)));
            }
        }
        SpinClock clk;
        while (true) {
            bool all = true;
#pragma unroll
            for (int j = 0; j < kScanRows; j++) all = all && agg_ready(cur[j]);
            if (__all_sync(kFull, all)) {
                if (p.stats && lane == 0) {  // MNWCUDA_ENCODE_STATS: was the window complete at first look?
                    p.stats[25] += 1;
                    p.stats[26] += clk.spins == 0;
                }
                break;
            }
            if (spin_expired(clk)) {
                if (lane == 0) spin_timeout(p.error, 3, base);
                return;
            }
            if (clk.spins > 16) __nanosleep(100);
#pragma unroll
            for (int j = 0; j < kScanRows; j++)
                if (!agg_ready(cur[j])) {
                    uint4 v = ld_relaxed_u4(p.blkAgg + base + j * 256 + lane * 8);
                    if (base + kScanWin > p.nBlocks) {
                        uint32_t *wds = reinterpret_cast<uint32_t *>(&v);
#pragma unroll
                        for (int i = 0; i < 8; i++)
                            if (base + j * 256 + lane * 8 + i >= p.nBlocks) wds[i >> 1] = (wds[i >> 1] & ~(0xffffu << (16 * (i & 1)))) | (kAggReady << (16 * (i & 1)));
                    }
                    cur[j] = v;
                }
        }
        uint32_t rowBase = 0;
#pragma unroll
        for (int j = 0; j < kScanRows; j++) {
            const uint32_t wd[4] = {cur[j].x, cur[j].y, cur[j].z, cur[j].w};
            uint32_t sz[8];
#pragma unroll
            for (int i = 0; i < 8; i++) sz[i] = (wd[i >> 1] >> (16 * (i & 1))) & 0x7fffu;
            uint32_t mine = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) mine += sz[i];
            uint32_t incl = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t up = __shfl_up_sync(kFull, incl, d);
                if (lane >= d) incl += up;
            }
            uint32_t ex = kRelReady | (rowBase + incl - mine);
            uint32_t o[8];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                o[i] = ex;
                ex += sz[i];
            }
            uint32_t *dst = p.blkRel + base + j * 256 + lane * 8;
            st_relaxed_u4(dst, make_uint4(o[0], o[1], o[2], o[3]));
            st_relaxed_u4(dst + 4, make_uint4(o[4], o[5], o[6], o[7]));
            rowBase += __shfl_sync(kFull, incl, 31);
        }
        running += rowBase;
        if (more) {
#pragma unroll
            for (int j = 0; j < kScanRows; j++) cur[j] = nxt[j];
        }
    }
    if (lane == 0)
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p.winBase + w), "l"(kPrefixReady | running) : "memory");
}

// ---------------------------------------------------------------------------
// Reduce warp: block min/max -> per-component min/max -> the field's
// parameters, in block order, one window of 512 blocks at a time with the next
// window's loads in flight. All lanes walk the (field, component) sequence
// together. Lane l, load j covers blocks base + 64 j + 2 l, +1.
// ---------------------------------------------------------------------------
struct RedWalk {
    int f, comp;
    int64_t compEnd;  // global index one past the component's last block
    uint32_t lo, hi;  // running ord(min), ~ord(max)
};

__device__ __forceinline__ void red_skip_empty(const EncodeParams &p, RedWalk &w) {
    while (w.f < p.nFields && p.fields[w.f].nb == 0) w.f++;  // k_depth_empty took care of those
    if (w.f < p.nFields) w.compEnd = p.fields[w.f].firstBlock + p.fields[w.f].nb;
}

// kDefer: the field's parameters are not worked out here but by red_flush, 32 fields at
// a time, one per lane (a batch of many small fields finishes hundreds of fields per window,
// and compute_quant -- a division, the depth loop, dependent loads -- takes microseconds)
template <bool kDefer = false>
__device__ __forceinline__ void red_finish_comp(const EncodeParams &p, RedWalk &w, int lane) {
    if (lane == 0) {
        uint32_t *mm = p.minmax + ((int64_t)w.f * 3 + w.comp) * 2;
        mm[0] = w.lo;
        mm[1] = w.hi;
        if (!kDefer && w.comp == 2) compute_quant(p, w.f);  // reads back what this thread has just written
    }
    w.lo = w.hi = 0xffffffffu;
    if (++w.comp == 3) {
        w.comp = 0;
        w.f++;
        red_skip_empty(p, w);
    } else {
        w.compEnd += p.fields[w.f].nb;
    }
}

// the parameters of the fields [from, w.f) whose last component has been finished (kDefer)
__device__ __forceinline__ void red_flush(const EncodeParams &p, const RedWalk &w, int &from, int lane) {
    if (from >= w.f) return;
    __syncwarp();  // lane 0's min/max words, for everybody
    for (int f0 = from; f0 < w.f; f0 += 32) {
        const int f = f0 + lane;
        if (f < w.f && p.fields[f].nb > 0) compute_quant(p, f);  // (empty fields: k_depth_empty)
    }
    from = w.f;
}

__device__ __forceinline__ bool mm_ready(const uint4 &v) { return (v.x | v.y) != 0u && (v.z | v.w) != 0u; }
__device__ __forceinline__ void red_load(const EncodeParams &p, int64_t base, int lane, uint4 (&a)[8]) {
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = ld_relaxed_u4(p.blkMM + base + j * 64 + lane * 2);
}

__device__ __forceinline__ void reduce_fields(const EncodeParams &p, int lane) {
    RedWalk w;
    w.f = w.comp = 0;
    w.compEnd = 0;
    w.lo = w.hi = 0xffffffffu;
    red_skip_empty(p, w);
    int pendFrom = w.f;  // first field whose parameters have not been worked out
    uint4 cur[8], nxt[8];
    red_load(p, 0, lane, cur);
    for (int64_t base = 0; base < p.nBlocks; base += kRedWin) {
        const bool more = base + kRedWin < p.nBlocks;
        if (more) red_load(p, base + kRedWin, lane, nxt);
        const int64_t winEnd = min(base + kRedWin, p.nBlocks);
        const bool whole = base + kRedWin <= p.nBlocks;
        SpinClock clk;
        while (true) {
            bool all = true;
            if (whole) {
#pragma unroll
                for (int j = 0; j < 8; j++) all = all && mm_ready(cur[j]);
            } else {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int64_t b0 = base + j * 64 + lane * 2;
                    all = all && (b0 >= winEnd || (cur[j].x | cur[j].y) != 0u) &&
                          (b0 + 1 >= winEnd || (cur[j].z | cur[j].w) != 0u);
                }
            }
            if (__all_sync(kFull, all)) {
                if (p.stats && lane == 0) {
                    p.stats[27] += 1;
                    p.stats[28] += clk.spins == 0;
                }
                break;
            }
            if (spin_expired(clk)) {
                if (lane == 0) spin_timeout(p.error, 6, base);
                return;
            }
            if (clk.spins > 16) __nanosleep(100);
#pragma unroll
            for (int j = 0; j < 8; j++)
                if (!mm_ready(cur[j])) cur[j] = ld_relaxed_u4(p.blkMM + base + j * 64 + lane * 2);
        }
        if (whole && base + kRedWin <= w.compEnd) {
            // the usual case: the whole window lies in one component
            uint32_t lo = 0xffffffffu, hi = 0xffffffffu;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                lo = min(lo, min(cur[j].x, cur[j].z));
                hi = min(hi, min(cur[j].y, cur[j].w));
            }
            w.lo = min(w.lo, __reduce_min_sync(kFull, lo));
            w.hi = min(w.hi, __reduce_min_sync(kFull, hi));
            if (base + kRedWin == w.compEnd) red_finish_comp<true>(p, w, lane);
        } else {
            // component by component
            int64_t s0 = base;
            while (s0 < winEnd) {
                const int64_t e = min(winEnd, w.compEnd);
                uint32_t lo = 0xffffffffu, hi = 0xffffffffu;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int64_t b0 = base + j * 64 + lane * 2;
                    if (b0 >= s0 && b0 < e) {
                        lo = min(lo, cur[j].x);
                        hi = min(hi, cur[j].y);
                    }
                    if (b0 + 1 >= s0 && b0 + 1 < e) {
                        lo = min(lo, cur[j].z);
                        hi = min(hi, cur[j].w);
                    }
                }
                w.lo = min(w.lo, __reduce_min_sync(kFull, lo));
                w.hi = min(w.hi, __reduce_min_sync(kFull, hi));
                s0 = e;
                if (s0 == w.compEnd) red_finish_comp<true>(p, w, lane);
            }
        }
        red_flush(p, w, pendFrom, lane);  // (w.f is the field in progress: everything before it is complete)
        if (more) {
#pragma unroll
            for (int j = 0; j < 8; j++) cur[j] = nxt[j];
        }
    }
}

// ---------------------------------------------------------------------------
// The same in parallel, for batches whose components are whole numbers of reduce
// windows (every bench-sized batch): one window is one L2 round trip for a warp, and
// a single warp walking the windows in order gets through exactly one per round trip
// -- measured, it never waited for the A work: everybody waited for it. So kRedWarps
// warps of CTA 1 take the windows round robin and publish one word per window
// (ord(min) | ~ord(max) << 32, never zero), and one more warp folds those words in
// order, 32 windows per load, and finishes the components.
// ---------------------------------------------------------------------------
#ifndef MNW_RED_WARPS
#define MNW_RED_WARPS 8
#endif
constexpr int kRedWarps = MNW_RED_WARPS;

__device__ __forceinline__ bool reduce_in_parallel(const EncodeParams &p) {
    return p.uniformBlocks > 0 && (p.uniformBlocks / 3) % kRedWin == 0 && p.winMM != nullptr;
}

__device__ __forceinline__ void reduce_windows(const EncodeParams &p, int lane, int k) {
    const int64_t nWin = p.nBlocks / kRedWin;  // (uniform batch: nBlocks is a multiple of kRedWin)
    uint4 cur[8];
    for (int64_t w = k; w < nWin; w += kRedWarps) {
        const int64_t base = w * kRedWin;
        red_load(p, base, lane, cur);
        SpinClock clk;
        while (true) {
            bool all = true;
#pragma unroll
            for (int j = 0; j < 8; j++) all = all && mm_ready(cur[j]);
            if (__all_sync(kFull, all)) break;
            if (spin_expired(clk)) {
                if (lane == 0) spin_timeout(p.error, 6, base);
                return;
            }
            if (clk.spins > 4) __nanosleep(200);
#pragma unroll
            for (int j = 0; j < 8; j++)
                if (!mm_ready(cur[j])) cur[j] = ld_relaxed_u4(p.blkMM + base + j * 64 + lane * 2);
        }
        uint32_t lo = 0xffffffffu, hi = 0xffffffffu;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            lo = min(lo, min(cur[j].x, cur[j].z));
            hi = min(hi, min(cur[j].y, cur[j].w));
        }
        lo = __reduce_min_sync(kFull, lo);
        hi = __reduce_min_sync(kFull, hi);
        if (lane == 0)
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p.winMM + w), "l"((uint64_t)lo | ((uint64_t)hi << 32)) : "memory");
    }
}

__device__ __forceinline__ void reduce_fold(const EncodeParams &p, int lane) {
    const int64_t nWin = p.nBlocks / kRedWin;
    RedWalk w;
    w.f = w.comp = 0;
    w.compEnd = 0;
    w.lo = w.hi = 0xffffffffu;
    red_skip_empty(p, w);
    for (int64_t w0 = 0; w0 < nWin; w0 += 32) {
        const int cnt = (int)min((int64_t)32, nWin - w0);
        // windows are folded as they arrive: waiting for the whole batch of 32 would wait
        // for A work that itself waits (through the ring) for the parameters of this field
        uint64_t v = 0;
        int done = 0;
        SpinClock clk;
        while (done < cnt) {
            if (lane >= done && lane < cnt && v == 0) v = ld_relaxed_u64(p.winMM + w0 + lane);
            const unsigned ok = __ballot_sync(kFull, lane < cnt && v != 0);
            int upto = done;
            while (upto < cnt && ((ok >> upto) & 1u)) upto++;
            for (int i = done; i < upto; i++) {
                const uint64_t x = __shfl_sync(kFull, v, i);
                w.lo = min(w.lo, (uint32_t)x);
                w.hi = min(w.hi, (uint32_t)(x >> 32));
                if ((w0 + i + 1) * kRedWin == w.compEnd) red_finish_comp(p, w, lane);
            }
            if (upto == done) {
                if (spin_expired(clk)) {
                    if (lane == 0) spin_timeout(p.error, 6, (w0 + done) * kRedWin);
                    return;
                }
                __nanosleep(200);
            }
            done = upto;
        }
    }
}

// ---------------------------------------------------------------------------
// The parallel reduce for batches of any shape (fields of unequal size: a window of 512
// blocks may hold hundreds of components). The kRedWarps warps of CTA 1 take the windows
// round robin as above; a window's blocks fall into segments, one per component it
// touches. The components that lie inside the window are finished by its warp (their
// min/max go straight to p.minmax); the first segment ("head": its component may have
// begun in an earlier window) and the last one if its component goes on ("tail") are
// published in a 32-byte record together with the component that is open where the
// window ends. The fold warp takes the records in order -- two merges, one jump, and the
// parameters of the fields completed so far, one field per lane -- so nothing in it grows
// with the number of components.
//   rec[0] = { head ord(min), head ~ord(max), tail ord(min), tail ~ord(max) }   (a pair is never 0 | 0)
//   rec[1] = { field open at the window's end, its component, flags, 0 }
// The window warp's plain stores to p.minmax are ordered before its record by a fence, and
// the fold warp fences after it has seen the record.
// ---------------------------------------------------------------------------
constexpr uint32_t kRecValid = 0x80000000u, kRecHeadCloses = 1u, kRecTail = 2u, kRecWrote = 4u;  // kRecWrote: p.minmax words too
constexpr int kRedWarpsGeneral = kRedWarps;  // (22 warps instead of 8: 10.9 against 8.3 ms on 40 000 small fields -- more pollers, not more progress)

// (f, comp, compEnd) of the component that holds block b (b < nBlocks)
__device__ __forceinline__ void red_locate(const EncodeParams &p, int64_t b, RedWalk &w) {
    const DevField *fields = p.fields;
    // (equal-sized fields: no search -- a dozen dependent loads per window otherwise)
    if (p.uniformBlocks > 0) w.f = (int)(b / p.uniformBlocks);
    else w.f = find_field(p.nFields, b, [fields](int i) { return fields[i].firstBlock; });
    const int64_t first = fields[w.f].firstBlock;
    const int nb = fields[w.f].nb;
    w.comp = (int)((b - first) / nb);
    w.compEnd = first + (int64_t)(w.comp + 1) * nb;
}

// the next component after the one that has just ended (w.f = nFields past the last one)
__device__ __forceinline__ void red_advance(const EncodeParams &p, RedWalk &w) {
    if (++w.comp == 3) {
        w.comp = 0;
        w.f++;
        red_skip_empty(p, w);
    } else {
        w.compEnd += p.fields[w.f].nb;
    }
}

// sm: 4 KiB of shared memory of this warp (CTA 1 has no ring traffic: the stages are free)
__device__ __forceinline__ void reduce_windows_general(const EncodeParams &p, uint2 *sm, int lane, int k) {
    const int64_t nWin = (p.nBlocks + kRedWin - 1) / kRedWin;
    uint4 *rec = reinterpret_cast<uint4 *>(p.winMM);
    uint4 cur[8];
    for (int64_t wi = k; wi < nWin; wi += kRedWarpsGeneral) {
        const int64_t base = wi * kRedWin, winEnd = min(base + kRedWin, p.nBlocks);
        red_load(p, base, lane, cur);
        SpinClock clk;
        while (true) {
            bool all = true;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int64_t b0 = base + j * 64 + lane * 2;
                all = all && (b0 >= winEnd || (cur[j].x | cur[j].y) != 0u) && (b0 + 1 >= winEnd || (cur[j].z | cur[j].w) != 0u);
            }
            if (__all_sync(kFull, all)) break;
            if (spin_expired(clk)) {
                if (lane == 0) spin_timeout(p.error, 6, base);
                return;
            }
            if (clk.spins > 4) __nanosleep(200);
#pragma unroll
            for (int j = 0; j < 8; j++)
                if (!mm_ready(cur[j])) cur[j] = ld_relaxed_u4(p.blkMM + base + j * 64 + lane * 2);
        }
        RedWalk w;
        red_locate(p, base, w);
        uint32_t headLo = 0xffffffffu, headHi = 0xffffffffu, tailLo = 0u, tailHi = 0u, flags = kRecValid;
        if (w.compEnd >= winEnd) {
            // the usual case with long fields: the window lies inside one component
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int64_t b0 = base + j * 64 + lane * 2;
                if (b0 < winEnd) {
                    headLo = min(headLo, cur[j].x);
                    headHi = min(headHi, cur[j].y);
                }
                if (b0 + 1 < winEnd) {
                    headLo = min(headLo, cur[j].z);
                    headHi = min(headHi, cur[j].w);
                }
            }
            headLo = __reduce_min_sync(kFull, headLo);
            headHi = __reduce_min_sync(kFull, headHi);
            if (w.compEnd == winEnd) {
                flags |= kRecHeadCloses;
                red_advance(p, w);
            }
        } else {
            // Several components: the window goes to shared memory and every lane takes a
            // component of its own (ordinal 3 f + c; empty fields are empty ranges), round
            // after round, until one reaches beyond the window's end: that one is open there.
            // The first component is the head (it may have begun in an earlier window); the
            // others lie inside the window -- their min/max are final -- or are the tail.
#pragma unroll
            for (int j = 0; j < 8; j++) reinterpret_cast<uint4 *>(sm)[32 * j + lane] = cur[j];
            __syncwarp();
            const int64_t o0 = (int64_t)w.f * 3 + w.comp, oEnd = (int64_t)p.nFields * 3;
            bool found = false;
            w.f = p.nFields;  // (open component if nothing reaches beyond the end: past the last field)
            w.comp = 0;
            for (int64_t r0 = o0; r0 < oEnd && !found; r0 += 32) {
                const int64_t o = r0 + lane;
                bool beyond = false, isTail = false;
                uint32_t lo = 0xffffffffu, hi = 0xffffffffu;
                int f = 0, c = 0;
                if (o < oEnd) {
                    f = (int)(o / 3);
                    c = (int)(o - (int64_t)f * 3);
                    const int nb = p.fields[f].nb;
                    const int64_t start = p.fields[f].firstBlock + (int64_t)c * nb, end = start + nb;
                    beyond = nb > 0 && end > winEnd;  // it holds the block at winEnd, or begins at or after it
                    const int64_t s0 = max(start, base), e = min(end, winEnd);
                    for (int64_t b = s0; b < e; b++) {
                        const uint2 v = sm[b - base];
                        lo = min(lo, v.x);
                        hi = min(hi, v.y);
                    }
                    if (s0 < e) {
                        if (o == o0) {
                            // (picked up below)
                        } else if (end <= winEnd) {  // a whole component inside the window
                            uint32_t *mm = p.minmax + o * 2;
                            mm[0] = lo;
                            mm[1] = hi;
                            __threadfence();  // (before this window's record)
                        } else {
                            isTail = true;
                        }
                    }
                }
                if (r0 == o0) {  // the head is lane 0's component of the first round
                    headLo = __shfl_sync(kFull, lo, 0);
                    headHi = __shfl_sync(kFull, hi, 0);
                    const bool closes = __shfl_sync(kFull, (int)!beyond, 0) != 0;
                    if (closes) flags |= kRecHeadCloses;
                }
                const unsigned tailMask = __ballot_sync(kFull, isTail);
                if (tailMask) {
                    const int src = __ffs(tailMask) - 1;
                    tailLo = __shfl_sync(kFull, lo, src);
                    tailHi = __shfl_sync(kFull, hi, src);
                    flags |= kRecTail;
                }
                const unsigned beyondMask = __ballot_sync(kFull, beyond);
                if (beyondMask) {
                    const int src = __ffs(beyondMask) - 1;
                    w.f = __shfl_sync(kFull, f, src);
                    w.comp = __shfl_sync(kFull, c, src);
                    found = true;
                }
            }
            flags |= kRecWrote;
            __syncwarp();  // (the next window's entries overwrite sm; the lanes' fences are behind them)
        }
        // (w is now the component open at winEnd: the one that holds that block, or past the last field)
        if (lane == 0) {
            st_relaxed_u4(rec + 2 * wi, make_uint4(headLo, headHi, tailLo, tailHi));
            st_relaxed_u4(rec + 2 * wi + 1, make_uint4((uint32_t)w.f, (uint32_t)w.comp, flags, 0u));
        }
    }
}

// The parameters of the fields are worked out by the other warps of CTA 1 (quant_workers):
// the fold warp only says how many fields are complete -- unless they come one or two at a
// time (long fields), when it works them out itself at once: a worker's answer is a poll
// interval later. qState[0] = fields complete (every field below it has all three min/max
// pairs in p.minmax), qState[1] = next field to hand out, qState[2] = fields below this one
// were done by the fold warp.
__device__ __forceinline__ void quant_workers(const EncodeParams &p, volatile int *qState, int lane) {
    while (true) {
        int f0 = 0;
        if (lane == 0) f0 = atomicAdd(const_cast<int *>(qState) + 1, 32);
        f0 = __shfl_sync(kFull, f0, 0);
        if (f0 >= p.nFields) return;
        const int f = f0 + lane;
        if (f < p.nFields && p.fields[f].nb > 0) {  // (empty fields: k_depth_empty)
            // a lane waits for its own field only: a long field must not wait for its 31 neighbours
            SpinClock clk;
            while (qState[0] <= f) {
                __nanosleep(200);
                if (spin_expired(clk)) { spin_timeout(p.error, 10, f); return; }
            }
            if (f >= qState[2]) {  // (below qState[2]: the fold warp did it itself)
                __threadfence();
                compute_quant(p, f);
            }
        }
        __syncwarp();
    }
}

__device__ __forceinline__ void reduce_fold_general(const EncodeParams &p, volatile int *qState, int lane) {
    const int64_t nWin = (p.nBlocks + kRedWin - 1) / kRedWin;
    const uint4 *rec = reinterpret_cast<const uint4 *>(p.winMM);
    RedWalk w;  // the open component
    w.f = w.comp = 0;
    w.compEnd = 0;
    w.lo = w.hi = 0xffffffffu;
    red_skip_empty(p, w);
    int pendFrom = w.f;
    // 32 windows per round, one lane polling each (a window at a time would be one L2 round
    // trip per window: thousands of them do not fit in the kernel's time); the records are
    // merged in order as they arrive
    for (int64_t w0 = 0; w0 < nWin; w0 += 32) {
        const int cnt = (int)min((int64_t)32, nWin - w0);
        uint4 r0 = make_uint4(0u, 0u, 0u, 0u), r1 = r0;
        bool have = false;
        int done = 0;
        SpinClock clk;
        while (done < cnt) {
            if (lane >= done && lane < cnt && !have) {
                r1 = ld_relaxed_u4(rec + 2 * (w0 + lane) + 1);
                r0 = ld_relaxed_u4(rec + 2 * (w0 + lane));
                have = (r1.z & kRecValid) && (r0.x | r0.y) != 0u && (!(r1.z & kRecTail) || (r0.z | r0.w) != 0u);
            }
            const unsigned ok = __ballot_sync(kFull, lane < cnt && have);
            int upto = done;
            while (upto < cnt && ((ok >> upto) & 1u)) upto++;
            for (int i = done; i < upto; i++) {
                const uint32_t hLo = __shfl_sync(kFull, r0.x, i), hHi = __shfl_sync(kFull, r0.y, i);
                const uint32_t tLo = __shfl_sync(kFull, r0.z, i), tHi = __shfl_sync(kFull, r0.w, i);
                const uint32_t fOpen = __shfl_sync(kFull, r1.x, i), cOpen = __shfl_sync(kFull, r1.y, i);
                const uint32_t flags = __shfl_sync(kFull, r1.z, i);
                if (flags & kRecWrote) __threadfence();  // (min/max words of the components inside that window)
                w.lo = min(w.lo, hLo);
                w.hi = min(w.hi, hHi);
                if (flags & kRecHeadCloses) {
                    if (lane == 0) {
                        uint32_t *mm = p.minmax + ((int64_t)w.f * 3 + w.comp) * 2;
                        mm[0] = w.lo;
                        mm[1] = w.hi;
                    }
                    w.lo = w.hi = 0xffffffffu;
                }
                w.f = (int)fOpen;
                w.comp = (int)cOpen;
                if (flags & kRecTail) {
                    w.lo = tLo;
                    w.hi = tHi;
                }
                if (w.f > pendFrom) {  // every field before the open one is complete
                    const bool self = w.f - pendFrom <= 4;
                    const int upTo = w.f;
                    if (self) red_flush(p, w, pendFrom, lane);  // (sets pendFrom = w.f)
                    if (lane == 0) {
                        __threadfence();  // (this lane's min/max words, and the window warps' that it has seen)
                        if (self) qState[2] = upTo;
                        qState[0] = upTo;
                    }
                    pendFrom = upTo;
                }
            }
            if (upto == done) {
                if (spin_expired(clk)) {
                    if (lane == 0) spin_timeout(p.error, 6, (w0 + done) * kRedWin);
                    return;
                }
                __nanosleep(200);
            }
            done = upto;
        }
    }
}

// ---------------------------------------------------------------------------
// A work: min/max of one block
// ---------------------------------------------------------------------------
template <bool kPos>
__device__ __forceinline__ void minmax_stage(const float4 *st, int lane, const Unwrap &uw, float &lo, float &hi) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float4 v = st[k * 32 + lane];  // any order of the chunks will do: the swizzle is ignored
        float e0 = v.x, e1 = v.y, e2 = v.z, e3 = v.w;
        if (kPos) {
            e0 = unwrap_t(e0, uw);
            e1 = unwrap_t(e1, uw);
            e2 = unwrap_t(e2, uw);
            e3 = unwrap_t(e3, uw);
        }
        lo = fminf(fminf(lo, e0), fminf(e1, fminf(e2, e3)));
        hi = fmaxf(fmaxf(hi, e0), fmaxf(e1, fmaxf(e2, e3)));
    }
}

template <bool kVec = false>
__device__ __forceinline__ void a_block(const EncodeParams &p, const WorkDesc &d, const float *stage, int lane) {
    const bool isPos = d.flags & kDescPos;
    Unwrap uw;
    uw.tHi = d.tHi;
    uw.tLo = d.tLo;
    uw.L = d.L;
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
#if defined(MNW_ABLATE) && (MNW_ABLATE & 8)
    if (true) {  // timing experiment only: one value per lane instead of 32
        lo = hi = stage[lane];
    } else
#endif
    if (!(d.flags & kDescSlow)) {
        const float4 *st = reinterpret_cast<const float4 *>(stage);
        minmax_stage<false>(st, lane, uw, lo, hi);
        if (isPos) {
            // unwrap_t is monotone within each of its three classes (x >= tHi, x < tLo,
            // between), so if neither threshold separates the lane's raw extrema, the
            // unwrapped extrema are the unwrapped raw extrema: 1 instruction per value
            // instead of 5. Rows along y and z almost always qualify; a lane that
            // straddles a threshold (or holds only NaNs) redoes its values one by one.
            const bool split = (lo < uw.tHi && !(hi < uw.tHi)) || (lo < uw.tLo && !(hi < uw.tLo)) || !(lo <= hi);
            if (split) {
                lo = __int_as_float(0x7f800000);
                hi = __int_as_float(0xff800000);
                minmax_stage<true>(st, lane, uw, lo, hi);
            } else {
                lo = unwrap_t(lo, uw);
                hi = unwrap_t(hi, uw);
            }
        }
    } else {
        const DevField *fdp = p.fields + d.f;
        const float *src = p.in + fdp->ioOff + (int64_t)d.comp * fdp->n + (int64_t)d.b * kBlockLen;
        bool done = false;
        if (kVec && d.len == kBlockLen) {
            // a full block off the 16-byte grid: vector loads, and the extrema rule of the staged path
            float lohi[2];
            minmax_full_unaligned(src, lane, lohi);
            lo = lohi[0];
            hi = lohi[1];
            const bool split = isPos && ((lo < uw.tHi && !(hi < uw.tHi)) || (lo < uw.tLo && !(hi < uw.tLo)) || !(lo <= hi));
            if (!__any_sync(kFull, split)) {
                if (isPos) {
                    lo = unwrap_t(lo, uw);
                    hi = unwrap_t(hi, uw);
                }
                done = true;
            } else {
                lo = __int_as_float(0x7f800000);
                hi = __int_as_float(0xff800000);
            }
        }
        if (!done) {
            for (int i = lane; i < d.len; i += 32) {
                float u = __ldg(src + i);
                if (isPos) u = unwrap_t(u, uw);
                lo = fminf(lo, u);
                hi = fmaxf(hi, u);
            }
        }
    }
    const uint32_t olo = __reduce_min_sync(kFull, ord_from_float(lo));
    const uint32_t ohi = __reduce_min_sync(kFull, ~ord_from_float(hi));  // max kept inverted: one min for both
    if (lane == 0)  // never all-zero: min is never NaN
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p.blkMM + d.blk),
                     "l"((uint64_t)olo | ((uint64_t)ohi << 32)) : "memory");
}

}  // namespace mnw

#include "enc_consume.cuh"

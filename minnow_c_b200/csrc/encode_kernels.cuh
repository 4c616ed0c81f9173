// encode_kernels.cuh -- the encoder: one fused kernel (k_encode) + k_finalize.
// Replaces, per field, position()/velocity() of
// /root/reference/src/quant.c:161-289 and the (stub) CFunc behind
// src/funcs.c:71. HBM-bound integer/float streaming: no tensor cores.
//
// k_encode hands out 8192-value tiles through one dynamic ticket counter, one
// ticket per CTA. Three kinds of tile work alternate in ticket order:
//   A(g)  util_UndoPeriodic + util_MinMax of tile g (src/util.c:97-113, 27-46);
//         the CTA that finishes the last A tile of a field also runs
//         deltaToDepth (src/quant.c:186-196, 684-693) and publishes the
//         field's quantisation parameters
//   B(g)  `lag` tiles later: util_UniformBinIndex + delta + per-block width +
//         util_U32UniformPack of tile g (src/util.c:173-196, 311-355). Its
//         second read of the input hits the 126 MB L2 instead of HBM: the live
//         window between the passes is about one field (48 MiB at 2^22
//         particles). The packed blocks are parked in an L2-resident scratch
//         ring and the tile's size is published.
//   C(g)  `lag2` tiles later: a dedicated scan warp has turned the sizes into
//         offsets by then; copy the parked blocks to their final position.
// A ticket is only ever held by a running CTA and only waits on tickets
// smaller than itself, so the spin-waits cannot deadlock; in steady state
// nothing waits at all.
#pragma once

#include "bitpack.cuh"
#include "minnow_kernels.cuh"

namespace mnw {

struct EncodeParams {
    const float *in;
    const DevField *fields;
    int nFields;
    int uniformTiles;   // >0: every field has this many tiles (fast lookup)
    uint32_t *minmax;   // [nFields][3][2] ordered-uint min, ~max
    DevQuant *quant;    // [nFields]
    uint32_t *tileAgg;    // [nTiles] kAggReady | packed size of the tile, 16-byte units
    uint64_t *tilePrefix; // [nTiles+1] kPrefixReady | exclusive prefix, written by the scan warp
    uint32_t *slotDone;   // [ringTiles] warps that have copied a tile out of the ring slot, cumulative
    uint8_t *scratch;     // [ringTiles][8][kSlotBytes] packed blocks waiting for their offset
    uint32_t *ticket;     // dynamic ticket counter
    uint32_t *aDone;      // [nFields] A tiles finished
    uint32_t *ready;      // [nFields] quantisation parameters published
    TmpBlock *tmp;        // [nBlocksTotal]
    uint8_t *out;
    int64_t outCap;
    uint64_t *offsets;  // [nFields+1]
    int *error;
    int64_t nTiles;
    int64_t nTickets;   // tickets >= nTickets mean "no more work"
    int64_t lag;        // B(g) is issued with A(g+lag); <0: no A tiles (quant given)
    int64_t lag2;       // C(g) is issued with B(g+lag2)
    int64_t ringTiles;  // scratch ring capacity in tiles: a power of two, > lag2 when tiles wrap
    int ringShift;      // log2(ringTiles)
    unsigned long long *debugStats;  // MNWCUDA_DEBUG_STATS: look-back counters, or null
    int debugMask;      // timing experiments only (MNWCUDA_DEBUG_MASK): 1 no look-back, 2 no stores,
                        // 4 no pack, 8 fixed width; output bytes are wrong when set
};

// what thread 0 works out for the whole CTA from its ticket
struct TileJob {
    int kind;  // 0 nothing, 1 = A (min/max), 2 = B (pack)
    int f, comp, t;
    int64_t tile;
};

__device__ __forceinline__ void locate_tile(const EncodeParams &p, int64_t tile, int &f, int &comp,
                                            int &tileInComp) {
    if (p.uniformTiles > 0) {
        f = (int)((uint32_t)tile / (uint32_t)p.uniformTiles);  // tickets < 2^31
    } else {
        const DevField *fields = p.fields;
        f = find_field(p.nFields, tile, [fields](int i) { return fields[i].firstTile; });
    }
    int local = (int)(tile - p.fields[f].firstTile);
    int tpc = p.fields[f].tilesPerComp;
    comp = local / tpc;
    tileInComp = local - comp * tpc;
}

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// L2 eviction-priority policies for the two reads of the input (sm_100a takes
// the priority qualifier directly only on 256-bit loads; 128-bit loads carry it
// as a cache-hint operand)
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
__device__ __forceinline__ float4 ld_hint_f4(const float4 *p, uint64_t pol) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p), "l"(pol));
    return r;
}

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
    q.scale = __ddiv_rn((double)(1u << depth), (double)md);
    p.quant[f] = q;
}

// one thread per field: empty fields have no tiles to trigger compute_quant
__global__ void k_depth_empty(EncodeParams p) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nFields) return;
    if (p.fields[f].n == 0) compute_quant(p, f);
}

// ---------------------------------------------------------------------------
// A tile, phase 1 (the 8 worker warps): unwrap (Posn only, never stored) +
// min/max. float4 streaming loads, warp-shuffle reduction, one slot per warp.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void minmax_phase1(const EncodeParams &p, const TileJob &job, float *sLo, float *sHi) {
    const int f = job.f, comp = job.comp, t = job.t;
    const DevField *fdp = p.fields + f;
    const int n = fdp->n;
    const float *base = p.in + fdp->ioOff + (int64_t)comp * n;
    const int64_t start = (int64_t)t * kTileVals;
    const int cnt = (int)min((int64_t)kTileVals, (int64_t)n - start);
    const bool isPos = fdp->fieldCode == field_Posn;
    const float L = fdp->width, half = __fmul_rn(L, 0.5f);
    const float anchor = isPos ? __ldg(base) : 0.0f;

    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    const float *src = base + start;
    if (cnt == kTileVals && ((uintptr_t)src & 15) == 0) {
        const float4 *s4 = reinterpret_cast<const float4 *>(src);
        float4 v[8];
        const uint64_t keep = l2_policy_keep();  // first read: stay in L2 for the B pass
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = ld_hint_f4(s4 + k * 256 + threadIdx.x, keep);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            float e0 = v[k].x, e1 = v[k].y, e2 = v[k].z, e3 = v[k].w;
            if (isPos) {
                e0 = unwrap(e0, anchor, L, half);
                e1 = unwrap(e1, anchor, L, half);
                e2 = unwrap(e2, anchor, L, half);
                e3 = unwrap(e3, anchor, L, half);
            }
            lo = fminf(lo, fminf(fminf(e0, e1), fminf(e2, e3)));
            hi = fmaxf(hi, fmaxf(fmaxf(e0, e1), fmaxf(e2, e3)));
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
    if ((threadIdx.x & 31) == 0) { sLo[threadIdx.x >> 5] = lo; sHi[threadIdx.x >> 5] = hi; }
}

// A tile, phase 2 (one thread, after the CTA barrier): one atomicMin/atomicMax
// pair per tile on an order-preserving integer image of the floats; the CTA
// that finishes a field's last tile publishes its quantisation parameters.
__device__ __forceinline__ void minmax_phase2(const EncodeParams &p, const TileJob &job, const float *sLo,
                                              const float *sHi) {
    float lo = sLo[0], hi = sHi[0];
#pragma unroll
    for (int w = 1; w < kWarpsPerCta; w++) { lo = fminf(lo, sLo[w]); hi = fmaxf(hi, sHi[w]); }
    const int f = job.f;
    uint32_t *mm = p.minmax + ((int64_t)f * 3 + job.comp) * 2;
    atomicMin(mm, ord_from_float(lo));
    atomicMin(mm + 1, ~ord_from_float(hi));  // max kept inverted: both slots memset to 0xff
    __threadfence();
    uint32_t done = atomicAdd(p.aDone + f, 1u);
    if (done == 3u * (uint32_t)p.fields[f].tilesPerComp - 1u) {
        __threadfence();
        compute_quant(p, f);
        __threadfence();
        st_release_u32(p.ready + f, 1u);
    }
}

}  // namespace mnw

#include "encode_pack.cuh"

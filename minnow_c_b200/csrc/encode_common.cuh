// encode_common.cuh -- parameters, memory-ordering helpers and the per-field
// quantisation rule shared by the encoder's pieces (encode_persist.cuh).
// Replaces, per field, position()/velocity() of
// /root/reference/src/quant.c:161-289 and the (stub) CFunc behind
// src/funcs.c:71. HBM-bound float/integer streaming: no tensor cores.
#pragma once

#include "bitpack.cuh"
#include "minnow_kernels.cuh"

namespace mnw {

constexpr int kEncTileBlocks = 4;                        // 1024-value blocks per ticket
constexpr int kEncTileVals = kEncTileBlocks * kBlockLen;  // 16 KiB of input
constexpr int kEncWarps = 16;                             // persistent warps per CTA, one CTA per SM
constexpr int kEncInWords = 1024;                         // per warp: one swizzled 4 KiB input stage
constexpr int kEncHalfWords = 1256;                       // per warp: two packed-output halves (skewed words)
constexpr int kEncWarpSmem = (kEncInWords + 2 * kEncHalfWords) * 4 + 2 * kEncTileBlocks * 32;
constexpr int kEncSmemBytes = kEncWarps * kEncWarpSmem;
constexpr int kEncSlotWords = 832;                        // global overflow slot of one block: 25 bits x 1024, 16-byte padded
constexpr uint32_t kPosSpilled = 0xffffffffu;             // block record: packed words are in the overflow slot

constexpr uint32_t kAggReady = 0x80000000u;
constexpr uint64_t kPrefixReady = 1ull << 63;
constexpr uint32_t kQBias = 0x4b000000u;  // bit pattern of 2^23: trunc by add.rz keeps it in the exponent

struct EncodeParams {
    const float *in;
    const uint64_t *qIn;  // literal CFunc boundary: quantised values in, or null
    const DevField *fields;
    int nFields;
    int uniformTiles;      // >0: every field has this many tiles (fast lookup)
    uint32_t *minmax;      // [nFields][3][2] ordered-uint min, ~max
    DevQuant *quant;       // [nFields]
    uint32_t *tileAgg;     // [nTiles] kAggReady | packed size of the tile, 16-byte units
    uint64_t *tilePrefix;  // [nTiles+1] kPrefixReady | exclusive prefix, written by the scan warp
    uint32_t *ticket;      // dynamic ticket counter
    uint32_t *aDone;       // [nFields] A tiles finished
    uint32_t *ready;       // [nFields] quantisation parameters published
    uint32_t *overflow;    // [CTAs][kEncWarps][2][kEncTileBlocks][kEncSlotWords]
    uint8_t *out;
    int64_t outCap;
    uint64_t *offsets;  // [nFields+1]
    int *error;
    int64_t nTiles;
    int64_t nTickets;  // tickets >= nTickets mean "no more work"
    int64_t lag;       // B(g) is issued with A(g+lag); <0: no A tiles (quant given)
};

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// L2 eviction-priority policies for the two reads of the input
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
// 16-byte asynchronous global->shared copy (LDGSTS), L2 hint on the source line
__device__ __forceinline__ void cp_async16(void *smemDst, const void *src, uint64_t pol) {
    uint32_t d = (uint32_t)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void st_stream_u4(uint4 *p, uint4 v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

// every spin-wait gives up after ~0.5 s and raises kErrTimeout instead of hanging
// the device: error[1] = site (1 ready flag, 2 prefix, 3 scan), error[2..3] = tile
constexpr uint32_t kSpinLimit = 1u << 19;
__device__ __noinline__ void spin_timeout(const EncodeParams &p, int site, int64_t tile) {
    if (atomicCAS(p.error, 0, kErrTimeout) == 0) {
        p.error[1] = site;
        p.error[2] = (int)(tile & 0xffffffff);
        p.error[3] = (int)(tile >> 32);
    }
}

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
}

// one thread per field: empty fields have no tiles to trigger compute_quant
__global__ void k_depth_empty(EncodeParams p) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nFields) return;
    if (p.fields[f].n == 0) compute_quant(p, f);
}

}  // namespace mnw

// encode_tile.cuh -- the per-warp tile work of k_encode: A tiles (min/max), B
// tiles (quantise + delta + pack into shared memory), placement of packed
// tiles, the kernel itself and k_finalize. Included by encode_persist.cuh.
#pragma once

#include "encode_persist.cuh"

namespace mnw {

constexpr unsigned kFull = 0xffffffffu;

struct TileRef {  // warp-uniform description of one tile
    int64_t tile;
    int f, comp, b0, nBlk;
    bool valid;
};

struct WarpSmem {
    float *in;        // [1024] swizzled input stage
    uint32_t *half0;  // [2][kEncHalfWords] packed blocks, skewed words
    uint32_t *rec0;   // [2][kEncTileBlocks][8] first, minDelta, width, size16, pos, nWords
    uint32_t *ovf0;   // global: [2][kEncTileBlocks][kEncSlotWords] blocks that did not fit their half
    __device__ __forceinline__ uint32_t *half(int h) const { return half0 + h * kEncHalfWords; }
    __device__ __forceinline__ uint32_t *rec(int h) const { return rec0 + h * kEncTileBlocks * 8; }
    __device__ __forceinline__ uint32_t *ovf(int h) const { return ovf0 + (size_t)h * kEncTileBlocks * kEncSlotWords; }
};

__device__ __forceinline__ TileRef make_ref(const EncodeParams &p, int64_t tile) {
    TileRef r;
    int t;
    locate_tile(p, tile, r.f, r.comp, t);
    r.tile = tile;
    r.b0 = t * kEncTileBlocks;
    r.nBlk = min(kEncTileBlocks, p.fields[r.f].nb - r.b0);
    r.valid = true;
    return r;
}

// ---------------------------------------------------------------------------
// A tile: unwrap (Posn only, never stored) + min/max of up to 4096 values.
// 128-bit streaming loads kept in L2 for the B pass, two blocks in flight.
// ---------------------------------------------------------------------------
template <bool kPos>
__device__ __forceinline__ void minmax_f4(const float4 (&v)[8], const Unwrap &uw, float &lo, float &hi) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
        float e0 = v[k].x, e1 = v[k].y, e2 = v[k].z, e3 = v[k].w;
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

template <bool kPos>
__device__ __forceinline__ void minmax_fast(const float4 *s4, int lane, const Unwrap &uw, float &lo, float &hi) {
    const uint64_t keep = l2_policy_keep();
    float4 va[8], vb[8];
#pragma unroll
    for (int k = 0; k < 8; k++) va[k] = ld_hint_f4(s4 + k * 32 + lane, keep);
#pragma unroll
    for (int k = 0; k < 8; k++) vb[k] = ld_hint_f4(s4 + 256 + k * 32 + lane, keep);
    minmax_f4<kPos>(va, uw, lo, hi);
#pragma unroll
    for (int k = 0; k < 8; k++) va[k] = ld_hint_f4(s4 + 512 + k * 32 + lane, keep);
    minmax_f4<kPos>(vb, uw, lo, hi);
#pragma unroll
    for (int k = 0; k < 8; k++) vb[k] = ld_hint_f4(s4 + 768 + k * 32 + lane, keep);
    minmax_f4<kPos>(va, uw, lo, hi);
    minmax_f4<kPos>(vb, uw, lo, hi);
}

__device__ __forceinline__ void a_tile(const EncodeParams &p, int64_t tile, int lane) {
    const TileRef r = make_ref(p, tile);
    const DevField *fdp = p.fields + r.f;
    const int n = fdp->n;
    const float *base = p.in + fdp->ioOff + (int64_t)r.comp * n;
    const int64_t start = (int64_t)r.b0 * kBlockLen;
    const int cnt = (int)min((int64_t)kEncTileVals, (int64_t)n - start);
    const bool isPos = fdp->fieldCode == field_Posn;
    Unwrap uw;
    uw.tHi = uw.tLo = uw.L = 0.0f;
    if (isPos) uw = make_unwrap(__ldg(base), fdp->width);

    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    const float *src = base + start;
    if (cnt == kEncTileVals && ((uintptr_t)src & 15) == 0) {
        const float4 *s4 = reinterpret_cast<const float4 *>(src);
        if (isPos) minmax_fast<true>(s4, lane, uw, lo, hi);
        else minmax_fast<false>(s4, lane, uw, lo, hi);
    } else {
        for (int i = lane; i < cnt; i += 32) {
            float u = __ldg(src + i);
            if (isPos) u = unwrap_t(u, uw);
            lo = fminf(lo, u);
            hi = fmaxf(hi, u);
        }
    }
    // warp reduction on the order-preserving integer image; the max is kept
    // inverted so that both slots start from a 0xff memset
    const uint32_t olo = __reduce_min_sync(kFull, ord_from_float(lo));
    const uint32_t ohi = __reduce_min_sync(kFull, ~ord_from_float(hi));
    if (lane == 0) {
        uint32_t *mm = p.minmax + ((int64_t)r.f * 3 + r.comp) * 2;
        atomicMin(mm, olo);
        atomicMin(mm + 1, ohi);
        __threadfence();
        const uint32_t done = atomicAdd(p.aDone + r.f, 1u);
        if (done == 3u * (uint32_t)fdp->tilesPerComp - 1u) {
            __threadfence();
            compute_quant(p, r.f);
            __threadfence();
            st_release_u32(p.ready + r.f, 1u);
        }
    }
}

// ---------------------------------------------------------------------------
// placement: copy blocks [from, to) of a packed tile from the warp's shared
// memory to their final position (128-bit streaming stores) and write their
// table entries. `pre` is the tile's exclusive prefix in 16-byte units.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t wait_prefix(const EncodeParams &p, int64_t tile) {
    uint64_t pre;
    uint32_t spins = 0;
    while (!((pre = ld_relaxed_u64(p.tilePrefix + tile)) & kPrefixReady)) {
        __nanosleep(200);
        if (++spins > kSpinLimit) { spin_timeout(p, 2, tile); break; }
    }
    return pre & ~kPrefixReady;
}

__device__ __forceinline__ void place_blocks(const EncodeParams &p, const TileRef &r, const uint32_t *buf,
                                             const uint32_t *ovf, const uint32_t *rec, int from, int to,
                                             uint64_t pre, int lane) {
    const DevField *fdp = p.fields + r.f;
    const int nb = fdp->nb;
    const uint64_t p0 = wait_prefix(p, fdp->firstTile);  // published before `pre`; relaxed stores may still reorder
    const int64_t tableBase = fdp->fixedBase + (int64_t)(p0 * 16) + 64;
    const int64_t payloadBase = tableBase + (int64_t)nb * 48;
    uint64_t running = pre - p0;
    for (int j = 0; j < from; j++) running += rec[j * 8 + 3];
    for (int j = from; j < to; j++) {
        const uint32_t size16 = rec[j * 8 + 3];
        const uint32_t pos = rec[j * 8 + 4];
        const int nWords = (int)rec[j * 8 + 5];
        const int64_t dstByte = payloadBase + (int64_t)(running * 16);
        if (dstByte + (int64_t)size16 * 16 > p.outCap) {
            if (lane == 0) atomicCAS(p.error, 0, kErrCapacity);
        } else {
            uint4 *dst = reinterpret_cast<uint4 *>(p.out + dstByte);
            if (pos == kPosSpilled) {
                const uint4 *slot = reinterpret_cast<const uint4 *>(ovf + (size_t)j * kEncSlotWords);
                for (int c = lane; c < (int)size16; c += 32) st_stream_u4(dst + c, __ldcg(slot + c));
            }
            const uint32_t *st = buf + (pos == kPosSpilled ? 0u : pos);
            for (int c = lane; pos != kPosSpilled && c < (int)size16; c += 32) {
                const int m = c * 4;
                uint4 o;
                o.x = st[skew(m)];
                o.y = m + 1 < nWords ? st[skew(m + 1)] : 0u;
                o.z = m + 2 < nWords ? st[skew(m + 2)] : 0u;
                o.w = m + 3 < nWords ? st[skew(m + 3)] : 0u;
                st_stream_u4(dst + c, o);
            }
            if (lane == 0) {
                const uint4 e = make_uint4(rec[j * 8 + 0], rec[j * 8 + 1], rec[j * 8 + 2], (uint32_t)running);
                *reinterpret_cast<uint4 *>(p.out + tableBase + ((int64_t)r.comp * nb + r.b0 + j) * 16) = e;
            }
        }
        running += size16;
    }
}

__device__ __forceinline__ void place_pending(const EncodeParams &p, TileRef &pd, const WarpSmem &ws, int h,
                                              int lane) {
    if (!pd.valid) return;
    const uint64_t pre = wait_prefix(p, pd.tile);
    place_blocks(p, pd, ws.half(h), ws.ovf(h), ws.rec(h), 0, pd.nBlk, pre, lane);
    pd.valid = false;
    __syncwarp();
}

}  // namespace mnw

#include "encode_btile.cuh"

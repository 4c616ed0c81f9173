// quant_kernels.cuh -- the callers either side of the hot path:
//   * the quantiser alone: the fixed quant_QField / quant_Field of
//     /root/reference/src/quant.h:6-7 for Posn / Velc (position()/velocity(),
//     undoPosition()/undoVelocity(), src/quant.c:161-289, 451-551), as thin
//     kernels over the same arithmetic the fused encoder and decoder use;
//   * particle IDs and unsigned ints: id()/undoID(), uint()/undoUint()
//     (src/quant.c:291-327, 373-398, 553-608) with util_U64UndoPeriodic and
//     util_U64MinMax (src/util.c:115-143, 48-68) -- lossless integer work;
//   * util_Checksum (src/util.c:438-445) over many byte ranges at once: the
//     per-block checksums of a segment image and the per-field CField.Checksum.
// None of this is on the timed path; plain grid-stride kernels.
#pragma once

#include "decode_kernels.cuh"
#include "enc_math.cuh"

namespace mnw {

// ---------------------------------------------------------------------------
// quantiser alone, floats -> uint64 (Posn / Velc)
// ---------------------------------------------------------------------------
struct QuantParams {
    const float *in;      // [3][n] planar
    uint64_t *qOut;       // [3][n]
    int32_t n;
    int isPos;
    float L;
    uint32_t *minmax;     // [3][2] ord(min), ~ord(max); 0xffffffff before the first value
    const DevQuant *quant;
};

constexpr int kQThreads = 256, kQPerThread = 8;

__global__ void __launch_bounds__(kQThreads) k_q_minmax(QuantParams p) {
    const int comp = blockIdx.y;
    const float *src = p.in + (int64_t)comp * p.n;
    Unwrap uw = make_unwrap(__ldg(src), p.L);
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    const int64_t base = (int64_t)blockIdx.x * kQThreads * kQPerThread;
#pragma unroll
    for (int k = 0; k < kQPerThread; k++) {
        const int64_t i = base + k * kQThreads + threadIdx.x;
        if (i < p.n) {
            float u = __ldg(src + i);
            if (p.isPos) u = unwrap_t(u, uw);
            lo = fminf(lo, u);
            hi = fmaxf(hi, u);
        }
    }
    const uint32_t olo = __reduce_min_sync(kFull, ord_from_float(lo));
    const uint32_t ohi = __reduce_min_sync(kFull, ~ord_from_float(hi));
    if ((threadIdx.x & 31) == 0 && lo <= hi) {  // (a warp with no value contributes nothing)
        atomicMin(p.minmax + comp * 2, olo);
        atomicMin(p.minmax + comp * 2 + 1, ohi);
    }
}

// maxDiff + deltaToDepth of field 0 of an EncodeParams (one thread)
__global__ void k_q_params(EncodeParams p) { compute_quant(p, 0); }

__global__ void __launch_bounds__(kQThreads) k_q_quantise(QuantParams p) {
    const int comp = blockIdx.y;
    const float *src = p.in + (int64_t)comp * p.n;
    const DevQuant qp = *p.quant;
    const float x0 = comp == 0 ? qp.X0[0] : comp == 1 ? qp.X0[1] : qp.X0[2];
    const uint32_t top = (1u << qp.depth) - 1u;
    Unwrap uw = make_unwrap(__ldg(src), p.L);
    const int64_t base = (int64_t)blockIdx.x * kQThreads * kQPerThread;
#pragma unroll
    for (int k = 0; k < kQPerThread; k++) {
        const int64_t i = base + k * kQThreads + threadIdx.x;
        if (i < p.n) {
            float u = __ldg(src + i);
            if (p.isPos) u = unwrap_t(u, uw);
            // util_UniformBinIndex (src/util.c:184-191) as in quantise32: exact without the division
            const float a = __fsub_rn(u, x0);
            const float s = __double2float_rn(__dmul_rn((double)a, qp.scale));
            p.qOut[(int64_t)comp * p.n + i] = (uint64_t)min(__float2uint_rz(s), top);
        }
    }
}

// ---------------------------------------------------------------------------
// dequantiser alone, uint64 -> floats: the decoder's dither (same streams, same
// draw order as k_unpack_dequant: stream task*32+lane, lane l owns values
// [32l, 32l+32) of every block of the task) on values given by the caller.
// The literal arithmetic of util_UndoUniformBinIndex (src/util.c:229-238).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_dequant_q(DecodeParams p, const uint64_t *qIn, int64_t nTasksTotal) {
    const int lane = threadIdx.x & 31;
    const int64_t task = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (task >= nTasksTotal) return;
    const DevField fd = p.fields[0];
    const DevQuant qp = p.quant[0];
    const uint64_t *sp = p.states + (task * 32 + lane) * 2;
    uint64_t s0 = sp[0], s1 = sp[1];
    const bool isPos = fd.fieldCode == field_Posn;
    const float bw = __fdiv_rn(qp.maxDiff, (float)(1u << qp.depth));
    const int g0 = (int)task * kRun;
    const int gCnt = min(3 * fd.nb - g0, kRun);
    for (int k = 0; k < gCnt; k++) {
        const int comp = (g0 + k) / fd.nb;
        const int b = g0 + k - comp * fd.nb;
        const int len = min(kBlockLen, fd.n - b * kBlockLen);
        const float x0 = comp == 0 ? qp.X0[0] : comp == 1 ? qp.X0[1] : qp.X0[2];
        const int64_t at = (int64_t)comp * fd.n + (int64_t)b * kBlockLen + lane * kChunk;
        for (int i = 0; i < kChunk && lane * kChunk + i < len; i++) {
            const float qf = (float)qIn[at + i];
            const float edge = __fadd_rn(x0, __fmul_rn(bw, qf));
            const uint32_t bits = (uint32_t)xs128p_next(s0, s1) & 0xffffffu;
            const float u = __fmul_rn((float)bits, 5.9604644775390625e-08f);
            float v = __fadd_rn(edge, __fmul_rn(u, bw));
            if (isPos) v = wrap(v, qp.width);
            p.out[at + i] = v;
        }
    }
}

// ---------------------------------------------------------------------------
// particle IDs: id() / undoID()
// ---------------------------------------------------------------------------
struct IdParams {
    const uint64_t *ids;   // [n]
    uint64_t *q;           // [3][n]
    int32_t n;
    uint64_t width;
    long long *mm;         // [3][2] min, max of the unwrapped coordinates (signed)
};

__device__ __forceinline__ long long id_coord(uint64_t id, uint64_t w, int c) {
    return (long long)(c == 0 ? id % w : c == 1 ? (id / w) % w : id / (w * w));
}
// util_U64UndoPeriodic's per-element rule (src/util.c:129-135); element 0 is the anchor
__device__ __forceinline__ long long id_unwrap(long long x, long long x0, long long L) {
    if (x - x0 >= L / 2) return x - L;
    if (x - x0 < -L / 2) return x + L;
    return x;
}

__global__ void __launch_bounds__(256) k_id_minmax(IdParams p) {
    const long long L = (long long)p.width;
    long long lo[3], hi[3], x0[3];
    const uint64_t first = __ldg(p.ids);
#pragma unroll
    for (int c = 0; c < 3; c++) {
        x0[c] = id_coord(first, p.width, c);
        lo[c] = LLONG_MAX;
        hi[c] = LLONG_MIN;
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t id = __ldg(p.ids + i);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const long long u = id_unwrap(id_coord(id, p.width, c), x0[c], L);
            lo[c] = min(lo[c], u);
            hi[c] = max(hi[c], u);
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
        for (int d = 16; d > 0; d >>= 1) {
            lo[c] = min(lo[c], __shfl_xor_sync(kFull, lo[c], d));
            hi[c] = max(hi[c], __shfl_xor_sync(kFull, hi[c], d));
        }
        if ((threadIdx.x & 31) == 0 && lo[c] <= hi[c]) {
            atomicMin(p.mm + c * 2, lo[c]);
            atomicMax(p.mm + c * 2 + 1, hi[c]);
        }
    }
}

// q = unwrapped coordinate - component minimum (the +L shift of util_U64UndoPeriodic
// for a negative minimum cancels in the difference; it only moves X0 and X1)
__global__ void __launch_bounds__(256) k_id_quantise(IdParams p) {
    const long long L = (long long)p.width;
    const uint64_t first = __ldg(p.ids);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t id = __ldg(p.ids + i);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const long long u = id_unwrap(id_coord(id, p.width, c), id_coord(first, p.width, c), L);
            p.q[(int64_t)c * p.n + i] = (uint64_t)(u - p.mm[c * 2]);
        }
    }
}

// undoID (src/quant.c:570-578); a value of all ones in any component marks a lost particle
__global__ void __launch_bounds__(256) k_id_restore(const uint64_t *q, int32_t n, uint64_t width, uint64_t x0a,
                                                    uint64_t x0b, uint64_t x0c, uint64_t *ids) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t qa = q[i], qb = q[(int64_t)n + i], qc = q[2 * (int64_t)n + i];
        if ((qa & qb & qc) == ~0ull || qa == ~0ull || qb == ~0ull || qc == ~0ull) {
            ids[i] = ~0ull;
            continue;
        }
        uint64_t x = qa + x0a, y = qb + x0b, z = qc + x0c;
        if (x >= width) x -= width;
        if (y >= width) y -= width;
        if (z >= width) z -= width;
        ids[i] = x + width * y + width * width * z;
    }
}

// ---------------------------------------------------------------------------
// unsigned ints: uint() / undoUint(); N values coded as three runs of ceil(N/3)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_u_minmax(const uint64_t *data, int32_t n, unsigned long long *mm) {
    unsigned long long lo = ~0ull, hi = 0ull;
    bool any = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long v = __ldg(data + i);
        lo = min(lo, v);
        hi = max(hi, v);
        any = true;
    }
    for (int d = 16; d > 0; d >>= 1) {
        lo = min(lo, __shfl_xor_sync(kFull, lo, d));
        hi = max(hi, __shfl_xor_sync(kFull, hi, d));
    }
    const bool anyW = __any_sync(kFull, any);
    if ((threadIdx.x & 31) == 0 && anyW) {
        atomicMin(mm, lo);
        atomicMax(mm + 1, hi);
    }
}
// q[i] = data[min(i, n-1)] - x0 for i < 3*run (the tail repeats the last value)
__global__ void __launch_bounds__(256) k_u_quantise(const uint64_t *data, int32_t n, int32_t run,
                                                    const unsigned long long *mm, uint64_t *q) {
    const int64_t m = 3 * (int64_t)run;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x)
        q[i] = __ldg(data + (i < n ? i : n - 1)) - mm[0];
}
// literal CFunc boundary for Unsi: the N values given, the tail of the three runs repeats the last one
__global__ void __launch_bounds__(256) k_u_pad(uint64_t *q, int32_t n, int32_t run) {
    const int64_t m = 3 * (int64_t)run;
    for (int64_t i = n + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x)
        q[i] = q[n - 1];
}
__global__ void __launch_bounds__(256) k_u_restore(const uint64_t *q, int32_t n, uint64_t x0, uint64_t *out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = q[i] == ~0ull ? ~0ull : q[i] + x0;
}

// ---------------------------------------------------------------------------
// util_Checksum (src/util.c:438-445): rotate right by one, add the byte, start 1.
// The recurrence does not compose (rotation and carry do not commute), so a range
// is one thread's sequential work; the parallelism is across ranges. Ranges that
// start on a 16-byte boundary are read 16 bytes at a time.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t checksum_word(uint32_t c, uint32_t w) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        // rotate right by one = the high word of (c:c) << 31; with the addition it is
        // one funnel-shift-and-add (LEA.HI) per byte on the dependent chain
        c = __funnelshift_l(c, c, 31) + ((w >> (8 * k)) & 0xffu);
    }
    return c;
}
// The same with the byte picked and added by a dot product (IDP.4A on the FMA pipe): two
// dependent instructions per byte instead of one, but one ALU-pipe instruction per byte
// instead of 2.5 -- for the many short chains of the packed blocks, which are bound by the
// ALU pipe's throughput, not by the chain's latency.
__device__ __forceinline__ uint32_t checksum_word_dp(uint32_t c, uint32_t w) {
    c = __dp4a(w, 0x00000001u, __funnelshift_l(c, c, 31));
    c = __dp4a(w, 0x00000100u, __funnelshift_l(c, c, 31));
    c = __dp4a(w, 0x00010000u, __funnelshift_l(c, c, 31));
    return __dp4a(w, 0x01000000u, __funnelshift_l(c, c, 31));
}
template <bool kShort>
__device__ __forceinline__ uint32_t checksum_16(uint32_t c, const uint4 &a) {
    if (kShort) {
        c = checksum_word_dp(c, a.x);
        c = checksum_word_dp(c, a.y);
        c = checksum_word_dp(c, a.z);
        return checksum_word_dp(c, a.w);
    }
    c = checksum_word(c, a.x);
    c = checksum_word(c, a.y);
    c = checksum_word(c, a.z);
    return checksum_word(c, a.w);
}
constexpr int kSumAhead = 8;  // 16-byte loads in flight per thread: a range is one dependent chain, so the
                              // next 128 bytes are asked for before the current ones are summed
template <bool kShort = false>
__device__ __forceinline__ uint32_t checksum_range(const uint8_t *src, int64_t len) {
    uint32_t c = 1;
    int64_t i = 0;
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
        const int64_t groups = len / (16 * kSumAhead);
        uint4 cur[kSumAhead], nxt[kSumAhead];
        if (groups > 0) {
#pragma unroll
            for (int k = 0; k < kSumAhead; k++) cur[k] = ld_stream_u4(s4 + k);
        }
        for (int64_t g = 0; g < groups; g++) {
            if (g + 1 < groups) {
#pragma unroll
                for (int k = 0; k < kSumAhead; k++) nxt[k] = ld_stream_u4(s4 + (g + 1) * kSumAhead + k);
            }
#pragma unroll
            for (int k = 0; k < kSumAhead; k++) c = checksum_16<kShort>(c, cur[k]);
#pragma unroll
            for (int k = 0; k < kSumAhead; k++) cur[k] = nxt[k];
        }
        i = groups * 16 * kSumAhead;
        for (; i + 16 <= len; i += 16) c = checksum_16<kShort>(c, ld_stream_u4(s4 + (i >> 4)));
    }
    for (; i < len; i++) c = __funnelshift_l(c, c, 31) + src[i];
    return c;
}

// sums[r] = util_Checksum(base + off[r], len[r])
__global__ void __launch_bounds__(128) k_checksum_ranges(const uint8_t *base, const int64_t *off, const int64_t *len,
                                                         uint32_t *sums, int64_t nRanges) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nRanges) return;
    sums[r] = checksum_range(base + off[r], len[r]);
}

// The BlockHeader table (doc/header_format/header_format.tex:182-185) of a batch's
// packed stream, straight from the stream: per field, block 0 = header + block
// table, blocks 1..3nb = the packed blocks in table order. One thread per block.
// verify = 0: headers[e] = {Length, Checksum}.
// verify = 1: headers[] is what the segment says; damaged[e] = the block's bytes (or
//   its length) do not match. k_mark_damage then gives a damaged pack block
//   Width = kWidthDamaged in the stream's table (only its particles are lost,
//   tex:188-196) and clears fieldValid[f] for a damaged block 0.
struct MnwBlockHeaderDev {
    int32_t Length;
    uint32_t Checksum;
};
// A field's block 0 (header + table: 48 bytes per block of the field, 192 KiB for 2^22
// particles) is one long chain. One CTA per field: threads 1..127 stream the range into a
// double-buffered staging area in shared memory with coalesced loads, thread 0 runs the
// chain out of it -- one LEA.HI per byte, never waiting for DRAM.
constexpr int kB0Chunk16 = 640;  // 16-byte words per staging buffer (10 KiB)

__device__ __forceinline__ uint32_t checksum_block0(const uint8_t *src, int64_t len, uint4 (*buf)[kB0Chunk16]) {
    const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
    const int64_t n16 = len >> 4;  // len = 64 + 48 nb: a multiple of 16
    const int64_t chunks = (n16 + kB0Chunk16 - 1) / kB0Chunk16;
    uint32_t c = 1;
    auto fill = [&](int64_t ch) {
        if (threadIdx.x == 0) return;
        const int64_t base = ch * kB0Chunk16;
        const int cnt = (int)min((int64_t)kB0Chunk16, n16 - base);
        for (int i = (int)threadIdx.x - 1; i < cnt; i += (int)blockDim.x - 1) buf[ch & 1][i] = ld_stream_u4(s4 + base + i);
    };
    fill(0);
    __syncthreads();
    for (int64_t ch = 0; ch < chunks; ch++) {
        if (ch + 1 < chunks) fill(ch + 1);
        if (threadIdx.x == 0) {
            const int cnt = (int)min((int64_t)kB0Chunk16, n16 - ch * kB0Chunk16);
            const uint4 *b = buf[ch & 1];
#pragma unroll 4
            for (int i = 0; i < cnt; i++) c = checksum_16<false>(c, b[i]);
        }
        __syncthreads();
    }
    return c;
}

// Packed blocks, one per thread: a thread reading its own block 16 bytes at a time
// touches 32 different lines per warp-wide load, and the L1 queue becomes the limit.
// Instead the warp fetches everybody's next 128 bytes together -- a quarter warp per
// thread's piece, 4 x 128 contiguous bytes per load instruction -- into shared memory, and
// every lane then sums its own piece from there. Lengths are multiples of 16, pointers
// 16-byte aligned (a lane with nothing to do passes len 0).
constexpr int kSumPiece16 = 8;           // 16-byte words per lane and round
constexpr int kSumRow = kSumPiece16 + 1;  // row stride in shared memory (odd: the lanes' rows spread over the banks)

__device__ __forceinline__ uint32_t checksum_blocks32(const uint8_t *src, int64_t len, uint4 (*st)[kSumRow]) {
    const int lane = threadIdx.x & 31;
    const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
    int rem = (int)(len >> 4);  // 16-byte words left
    uint32_t c = 1;
    while (__any_sync(kFull, rem > 0)) {
#pragma unroll
        for (int i = 0; i < 32 / 4; i++) {
            const int j = 4 * i + (lane >> 3);  // the lane whose piece this quarter warp fetches
            const uint64_t pj = __shfl_sync(kFull, (uint64_t)(uintptr_t)s4, j);
            const int rj = __shfl_sync(kFull, rem, j);
            const int w = lane & 7;
            if (w < rj) st[j][w] = ld_stream_u4(reinterpret_cast<const uint4 *>((uintptr_t)pj) + w);
        }
        __syncwarp();
        const int n = min(rem, kSumPiece16);
        for (int w = 0; w < n; w++) c = checksum_16<true>(c, st[lane][w]);
        s4 += n;
        rem -= n;
        __syncwarp();
    }
    return c;
}

// Grid: one CTA per field for its block 0, then one thread per packed block. e is the
// entry's index in the BlockHeader table.
__global__ void __launch_bounds__(128) k_block_headers(DecodeParams p, MnwBlockHeaderDev *headers, int verify,
                                                       uint8_t *damaged, int64_t nEntries) {
    __shared__ uint4 buf[2][kB0Chunk16];  // block 0 CTAs (checksum_block0) ...
    uint4 (*stageSum)[32][kSumRow] = reinterpret_cast<uint4 (*)[32][kSumRow]>(&buf[0][0]);  // ... or the others
    static_assert(sizeof(buf) >= 4 * 32 * kSumRow * sizeof(uint4), "k_block_headers: staging area of four warps");
    const DevField *fields = p.fields;
    int f;
    int64_t k;
    bool live = true;
    if ((int64_t)blockIdx.x < p.nFields) {
        f = (int)blockIdx.x;
        k = 0;
    } else {
        int64_t d = ((int64_t)blockIdx.x - p.nFields) * blockDim.x + threadIdx.x;  // packed block d of the batch
        live = d < nEntries - p.nFields;  // (a dead thread stays for the warp's shuffles, on block 0 of the batch)
        if (!live) d = 0;
        // (p.uniformTasks is the blocks per field here when all fields have as many: no search --
        // nine dependent loads -- in front of every thread's work)
        if (p.uniformTasks > 0) f = (int)(d / p.uniformTasks);
        else f = find_field(p.nFields, d, [fields](int i) { return fields[i].firstBlock; });
        k = d - fields[f].firstBlock + 1;
    }
    const DevField fd = p.fields[f];
    const int64_t e = fd.firstBlock + f + k;
    const uint8_t *region = p.in + p.offsets[f];
    const int64_t regionLen = (int64_t)(p.offsets[f + 1] - p.offsets[f]);
    const int64_t fixed = 64 + (int64_t)fd.nb * 48;
    int64_t at = 0, len = fixed;
    bool inside = (p.offsets[f] & 15u) == 0 && regionLen >= fixed;
    uint32_t sum = 0u;
    if (k == 0) {  // (uniform over the CTA)
        if (inside) sum = checksum_block0(region, len, buf);
        if (threadIdx.x != 0) return;
    } else {
        inside = inside && live;
        if (inside) {
            const uint4 ent = __ldg(reinterpret_cast<const uint4 *>(region + 64) + (k - 1));
            const int b = (int)((k - 1) % fd.nb);
            const int cnt = min(kBlockLen, fd.n - b * kBlockLen);
            const uint32_t w = ent.z > (uint32_t)kMaxWidth ? 0u : ent.z;
            at = fixed + (int64_t)ent.w * 16;
            len = (int64_t)((((cnt * (int)w + 31) >> 5) + 3) >> 2) * 16;
            inside = ent.z <= (uint32_t)kMaxWidth && at + len <= regionLen;
        }
        // the packed blocks of a warp's 32 threads are summed together, staged through shared memory
        sum = checksum_blocks32(inside ? region + at : nullptr, inside ? len : 0, stageSum[threadIdx.x >> 5]);
    }
    if (!live) return;
    if (!verify) {
        headers[e].Length = (int32_t)len;
        headers[e].Checksum = sum;
        if (!inside) atomicCAS(p.error, 0, kErrPayload);
        return;
    }
    damaged[e] = !(inside && headers[e].Length == (int32_t)len && headers[e].Checksum == sum);
}

// second step of the verification (the table is not touched while it is being summed)
__global__ void __launch_bounds__(128) k_mark_damage(DecodeParams p, const uint8_t *damaged, int32_t *fieldValid,
                                                     int64_t nEntries) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nEntries || !damaged[e]) return;
    const DevField *fields = p.fields;
    const int f = find_field(p.nFields, e, [fields](int i) { return fields[i].firstBlock + i; });
    const DevField fd = p.fields[f];
    const int64_t k = e - (fd.firstBlock + f);
    if (k == 0) {
        fieldValid[f] = 0;
        return;
    }
    const int64_t regionLen = (int64_t)(p.offsets[f + 1] - p.offsets[f]);
    if ((p.offsets[f] & 15u) != 0 || regionLen < 64 + (int64_t)fd.nb * 48) return;
    uint4 *entry = reinterpret_cast<uint4 *>(const_cast<uint8_t *>(p.in) + p.offsets[f] + 64) + (k - 1);
    entry->z = kWidthDamaged;
}

}  // namespace mnw

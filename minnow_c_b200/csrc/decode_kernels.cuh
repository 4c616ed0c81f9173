// decode_kernels.cuh -- K3p header parse, K3a dither stream states, K3
// unpack + inverse delta + dithered de-quantisation. Replaces, per field, the
// (stub) DFunc behind /root/reference/src/funcs.c:54 and undoPosition()/
// undoVelocity() of src/quant.c:451-551, with the dither arithmetic of
// util_UndoUniformBinIndex (src/util.c:223-242) and the RNG of src/rand.c.
#pragma once

#include "bitpack.cuh"
#include "minnow_kernels.cuh"

namespace mnw {

struct DecodeParams {
    const uint8_t *in;
    const uint64_t *offsets;  // [nFields+1]
    const DevField *fields;
    int nFields;
    int uniformTasks;  // >0: every field has this many tasks
    DevQuant *quant;   // [nFields] parsed from the headers
    uint64_t *seeds;   // [nFields]
    int *tableOf;      // [nFields] field whose stream table this field uses
    uint64_t *states;  // [nFields][maxStreams][2], or [maxStreams][nFields][2] when statesByStream
    int64_t maxStreams;
    const uint64_t *jumpPow;  // [kJumpPows][128][2]: columns of J^(2^k)
    float *out;
    uint64_t *qOut;  // literal DFunc boundary: quantised values out, or null
    int *error;
    uint32_t *dmgCount;  // tasks that hold a damaged block (k_damaged_tasks blanks those blocks afterwards)
    uint32_t *dmgList;   // [nTasks]
    int statesByStream;  // k_stream_states_tab writes stream-major: a CTA's states of one k for all fields are contiguous
};

__device__ __forceinline__ int64_t state_index(const DecodeParams &p, int f, int64_t k) {
    return p.statesByStream ? k * p.nFields + f : (int64_t)f * p.maxStreams + k;
}

constexpr int kJumpPows = 26;
constexpr uint32_t kWidthDamaged = 0xffffffffu;  // table entry of a block whose checksum failed (k_block_headers)

// ---------------------------------------------------------------------------
// K3p: one thread per field reads the 64-byte header: quantisation parameters,
// seed, consistency checks; decides whose stream table to use (fields with the
// seed of field 0 share table 0).
// ---------------------------------------------------------------------------
// A region is usable only if it is 16-byte aligned, its offsets are ordered and it
// holds at least the header and the block table: checked before the first load, so
// that a truncated or corrupt stream raises kErrPayload instead of faulting.
__device__ __forceinline__ bool region_ok(const DecodeParams &p, int f) {
    const uint64_t a = p.offsets[f], b = p.offsets[f + 1];
    return (a & 15u) == 0 && b >= a && b - a >= (uint64_t)(64 + 48 * (int64_t)p.fields[f].nb);
}

__global__ void k_parse_headers(DecodeParams p) {
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nFields) return;
    const DevField fd = p.fields[f];
    DevQuant q;
    if (!region_ok(p, f)) {
        atomicCAS(p.error, 0, kErrPayload);
        for (int c = 0; c < 3; c++) q.X0[c] = q.X1[c] = 0.0f;
        q.maxDiff = q.width = 0.0f;
        q.depth = 0xffffffffu;
        p.quant[f] = q;
        p.seeds[f] = 0;
        p.tableOf[f] = f;
        return;
    }
    const uint8_t *src = p.in + p.offsets[f];
    __align__(16) MnwDiffHeader h;
    for (int i = 0; i < 4; i++)
        reinterpret_cast<uint4 *>(&h)[i] = reinterpret_cast<const uint4 *>(src)[i];
    bool ok = h.Magic == algo_Diff && h.ParticleLen == fd.n && h.BlockLen == kBlockLen &&
              h.DitherRun == MNWCUDA_DITHER_RUN && h.DitherLanes == MNWCUDA_DITHER_LANES &&
              h.Depth <= 24 && h.FieldCode == fd.fieldCode;
    if (!ok) atomicCAS(p.error, 0, kErrPayload);
    float md = 0.0f;
    for (int c = 0; c < 3; c++) {
        q.X0[c] = h.X0[c];
        q.X1[c] = h.X1[c];
        float d = __fsub_rn(h.X1[c], h.X0[c]);
        if (md < d) md = d;
    }
    q.maxDiff = md;
    q.depth = ok ? h.Depth : 0xffffffffu;
    q.width = h.Width;
    p.quant[f] = q;
    p.seeds[f] = h.DitherSeed;
    // fields with the seed of field 0 (and no more streams than it) share its table
    bool share = false;
    if (f > 0 && region_ok(p, 0)) {
        const MnwDiffHeader *h0 = reinterpret_cast<const MnwDiffHeader *>(p.in + p.offsets[0]);
        int64_t mine = (3 * (int64_t)fd.nb + kRun - 1) / kRun * 32;
        int64_t zero = (3 * (int64_t)p.fields[0].nb + kRun - 1) / kRun * 32;
        share = h.DitherSeed == h0->DitherSeed && mine <= zero;
    }
    p.tableOf[f] = share ? 0 : f;
}

// splitmix64, src/rand.c:114-121
__device__ __forceinline__ uint64_t splitmix_next(uint64_t &s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// ---------------------------------------------------------------------------
// K3a: rand_Seed(seed, nStreams) (src/rand.c:22-39) in parallel. Stream k is
// the seed state advanced by k jumps of 2^64 draws; the jump is GF(2)-linear,
// so J^k s = product over the set bits of k of the precomputed J^(2^i), each
// applied as an XOR of the matrix columns selected by the state's bits.
// One thread per stream; only fields that own a table do work.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_stream_states(DecodeParams p) {
    int f = blockIdx.y;
    if (p.tableOf[f] != f) return;
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const DevField fd = p.fields[f];
    int64_t nStreams = (3 * (int64_t)fd.nb + kRun - 1) / kRun * 32;
    if (k >= nStreams) return;
    uint64_t sm = p.seeds[f];
    uint64_t s0 = splitmix_next(sm);
    uint64_t s1 = splitmix_next(sm);
    for (int bit = 0; bit < kJumpPows; bit++) {
        if (!((k >> bit) & 1)) continue;
        const ulonglong2 *col = reinterpret_cast<const ulonglong2 *>(p.jumpPow) + bit * 128;
        uint64_t a = 0, b = 0;
#pragma unroll 8
        for (int i = 0; i < 64; i++) {
            ulonglong2 c0 = __ldg(col + i);
            ulonglong2 c1 = __ldg(col + 64 + i);
            uint64_t m0 = 0ull - ((s0 >> i) & 1ull);
            uint64_t m1 = 0ull - ((s1 >> i) & 1ull);
            a ^= (c0.x & m0) ^ (c1.x & m1);
            b ^= (c0.y & m0) ^ (c1.y & m1);
        }
        s0 = a;
        s1 = b;
    }
    uint64_t *dst = p.states + state_index(p, f, k) * 2;
    dst[0] = s0;
    dst[1] = s1;
}

// ---------------------------------------------------------------------------
// K3a for batches of many fields with their own seeds. J^k does not depend on the
// seed: k_jump_table tabulates the columns of J^k for every stream index once per
// context (2 KiB per k), and k_stream_states_tab applies J^k to the seed states of
// all fields that own a table -- one CTA per k with the matrix in shared memory,
// threads over fields, one 128x128 GF(2) matrix-vector product per stream (about
// 900 instructions) instead of one per set bit of k per field.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void gf2_matvec(const ulonglong2 *col, uint64_t &s0, uint64_t &s1) {
    uint64_t a = 0, b = 0;
#pragma unroll 8
    for (int i = 0; i < 64; i++) {
        const ulonglong2 c0 = col[i], c1 = col[64 + i];
        const uint64_t m0 = 0ull - ((s0 >> i) & 1ull);
        const uint64_t m1 = 0ull - ((s1 >> i) & 1ull);
        a ^= (c0.x & m0) ^ (c1.x & m1);
        b ^= (c0.y & m0) ^ (c1.y & m1);
    }
    s0 = a;
    s1 = b;
}

// one thread per (k, column i): column i of J^k = product over the set bits of k of J^(2^bit) applied to e_i
__global__ void __launch_bounds__(128) k_jump_table(const uint64_t *jumpPow, ulonglong2 *tab, int64_t nStreams) {
    const int64_t k = blockIdx.x;
    const int i = threadIdx.x;
    if (k >= nStreams) return;
    uint64_t s0 = i < 64 ? 1ull << i : 0ull, s1 = i >= 64 ? 1ull << (i - 64) : 0ull;
    for (int bit = 0; bit < kJumpPows; bit++) {
        if (!((k >> bit) & 1)) continue;
        gf2_matvec(reinterpret_cast<const ulonglong2 *>(jumpPow) + bit * 128, s0, s1);
    }
    tab[k * 128 + i] = make_ulonglong2(s0, s1);
}

// The CTA first turns the 128 columns of its J^k into a table of nibble products
// (32 nibbles x 16 values, 8 KiB of shared memory: entry [n][v] = XOR of the columns
// 4n + b over the set bits b of v), so that a matrix-vector product is 32 look-ups
// and XORs (~230 instructions) instead of 128 masked columns (~1000).
__global__ void __launch_bounds__(128) k_stream_states_tab(DecodeParams p, const ulonglong2 *tab) {
    __shared__ ulonglong2 col[128];
    __shared__ ulonglong2 nib[32][16];
    const int64_t k = blockIdx.x;
    col[threadIdx.x] = tab[k * 128 + threadIdx.x];
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int e = threadIdx.x * 4 + j, n = e >> 4, v = e & 15;
        ulonglong2 x = make_ulonglong2(0ull, 0ull);
#pragma unroll
        for (int b = 0; b < 4; b++)
            if ((v >> b) & 1) {
                x.x ^= col[4 * n + b].x;
                x.y ^= col[4 * n + b].y;
            }
        nib[n][v] = x;
    }
    __syncthreads();
    for (int f = threadIdx.x; f < p.nFields; f += blockDim.x) {
        if (p.tableOf[f] != f) continue;
        const int64_t nStreams = (3 * (int64_t)p.fields[f].nb + kRun - 1) / kRun * 32;
        if (k >= nStreams) continue;
        uint64_t sm = p.seeds[f];
        const uint64_t s0 = splitmix_next(sm);
        const uint64_t s1 = splitmix_next(sm);
        uint64_t a = 0, b = 0;
#pragma unroll
        for (int n = 0; n < 16; n++) {
            const ulonglong2 e0 = nib[n][(s0 >> (4 * n)) & 15u], e1 = nib[16 + n][(s1 >> (4 * n)) & 15u];
            a ^= e0.x ^ e1.x;
            b ^= e0.y ^ e1.y;
        }
        uint64_t *dst = p.states + state_index(p, f, k) * 2;
        dst[0] = a;
        dst[1] = b;
    }
}

// ---------------------------------------------------------------------------
// K3: one warp per dither task = 32 consecutive blocks of one field; lane l
// owns stream task*32+l and values [32l, 32l+32) of every block it visits.
// There is no CTA-wide synchronisation: warps are independent and latency is
// hidden by occupancy.
//   table   lane i holds the table entry of the task's block i (one coalesced
//           512-byte load), broadcast by shuffle when the block is visited
//   unpack  util_U32UndoUniformPack (src/util.c:383-403): 128-bit coalesced
//           loads staged through skewed shared memory, width-specialised
//           extraction of the lane's 32 residuals from its `width` words
//   scan    q[j] = first + sum_{i<=j} (r[i]+minDelta): lane-local inclusive
//           sums + warp shuffle scan of the lane totals
//   dequant util_UndoUniformBinIndex (src/util.c:229-238), two separately
//           rounded statements, one xorshift128+ draw per value
//   wrap    util_Periodic (src/util.c:76-82), Posn only
//   store   transposed through swizzled shared memory, 128-bit coalesced
// ---------------------------------------------------------------------------
constexpr int kDecWarps = 4;

// The lane's 32 deltas wait in its row of the swizzled staging area (exact floats,
// |d| < 2^24); S enters as the value before the lane's first one and every
// partial sum is an integer below 2^24, so the FP32 adds are exact. The loop is
// rolled (8 values per trip) on purpose: the fully unrolled body was 17 KB of code
// per variant and the kernel stalled on instruction fetch.
// kFold: the dither term RN(bits * 2^-24 * bw) is computed as RN((bits * 2^8) *
// (bw * 2^-32)) -- the same real number rounded once, as long as bw * 2^-32 is
// itself exact (the caller checks bw >= 2^-90); the shift replaces the 24-bit mask
// and runs on the FMA pipe. kWrap && isPos: util_Periodic. kGuard: elements at or
// beyond len draw nothing from the stream (ragged last block).
template <bool kGuard, bool kWrap, bool kFold>
__device__ __forceinline__ void dequant_row(float S, float x0, float bw, float bwS, bool isPos, float L,
                                            int lane, int len, uint64_t &s0, uint64_t &s1, uint32_t rowAddr) {
    uint32_t al = (uint32_t)s0, ah = (uint32_t)(s0 >> 32), bl = (uint32_t)s1, bh = (uint32_t)(s1 >> 32);
#pragma unroll 2
    for (int j = 0; j < 8; j++) {
        const uint32_t slot = rowAddr ^ (j << 4);
        const float4 d4 = lds_f4(slot);
        const float d[4] = {d4.x, d4.y, d4.z, d4.w};
        float o[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            S = __fadd_rn(S, d[i]);
            float v = 0.0f;
            if (!kGuard || lane * kChunk + j * 4 + i < len) {
                float edge = __fadd_rn(x0, __fmul_rn(bw, S));
                uint32_t sum = xs128p_next32(al, ah, bl, bh);
                float ub;
                if (kFold) {
                    ub = __fmul_rn((float)(sum * 256u), bwS);
                } else {
                    float u = __fmul_rn((float)(sum & 0xffffffu), 5.9604644775390625e-08f);  // /2^24, exact
                    ub = __fmul_rn(u, bw);
                }
                v = __fadd_rn(edge, ub);
                if (kWrap && isPos) v = wrap(v, L);
            }
            o[i] = v;
        }
        sts_f4(slot, make_float4(o[0], o[1], o[2], o[3]));
    }
    s0 = ((uint64_t)ah << 32) | al;
    s1 = ((uint64_t)bh << 32) | bl;
}

template <bool kToQ>
__global__ void __launch_bounds__(kDecWarps * 32, 7) k_unpack_dequant(DecodeParams p, int64_t nTasksTotal) {
    __shared__ __align__(16) uint32_t stage[kDecWarps][kStageWords];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t task = (int64_t)blockIdx.x * kDecWarps + warp;
    if (task >= nTasksTotal) return;
    uint32_t *st = stage[warp];

    int f;
    if (p.uniformTasks > 0) {
        f = (int)(task / p.uniformTasks);
    } else {
        const DevField *fields = p.fields;
        f = find_field(p.nFields, task, [fields](int i) { return fields[i].firstTask; });
    }
    const DevField fd = p.fields[f];
    const DevQuant qp = p.quant[f];
    if (qp.depth > 24) return;  // malformed header, error already raised
    const int localTask = (int)(task - fd.firstTask);
    const uint8_t *region = p.in + p.offsets[f];
    const int64_t regionLen = (int64_t)(p.offsets[f + 1] - p.offsets[f]);
    const uint4 *table = reinterpret_cast<const uint4 *>(region + 64);
    const int64_t fixed = 64 + (int64_t)fd.nb * 48;
    const uint4 *payload = reinterpret_cast<const uint4 *>(region + fixed);
    const int64_t payloadLen16 = (regionLen - fixed) >> 4;

    uint64_t s0 = 0, s1 = 0;
    if (!kToQ) {
        const uint64_t *sp =
            p.states + state_index(p, p.tableOf[f], (int64_t)localTask * 32 + lane) * 2;
        s0 = sp[0];
        s1 = sp[1];
    }
    const bool isPos = fd.fieldCode == field_Posn;
    const float L = qp.width;  // the encoder's box width, not the caller's
    const float bw = __fdiv_rn(qp.maxDiff, (float)(1u << qp.depth));
    const float bwS = __fmul_rn(bw, 2.3283064365386963e-10f);  // bw * 2^-32
    const bool foldOK = bw >= 8.0779356694631609e-28f;         // 2^-90: bwS is exact

    const int g0 = localTask * kRun;
    const int gCnt = min(3 * fd.nb - g0, kRun);
    // Lane i prepares block i of the task once: table entry, component, bounds
    // check, the component's X0 and whether util_Periodic can do anything to it.
    // ent.z travels as width | component << 8 | wrap << 10.
    uint4 ent = make_uint4(0u, 0u, 0u, 0u);
    float x0L = 0.0f;
    {
        bool bad = false, damaged = false;
        if (lane < gCnt) {
            ent = __ldg(table + g0 + lane);
            const int gl = g0 + lane;
            const int compL = gl / fd.nb;
            const int lenL = min(kBlockLen, fd.n - (gl - compL * fd.nb) * kBlockLen);
            // A block whose checksum failed upstream is marked, not read: it goes through
            // as a block of width 0 (its lanes draw their dither as usual, so the task's
            // other blocks are untouched) and k_damaged_tasks blanks its values afterwards.
            damaged = ent.z == kWidthDamaged;
            if (damaged) ent.z = ent.w = 0u;
            const int size16L = (((lenL * (int)ent.z + 31) >> 5) + 3) >> 2;
            bad = ent.z > (uint32_t)kMaxWidth || (int64_t)ent.w + size16L > payloadLen16;
            x0L = compL == 0 ? qp.X0[0] : compL == 1 ? qp.X0[1] : qp.X0[2];
            // Can util_Periodic move any value of this component? The values are
            // x0 + bw*q + dither with q up to the index of X1 (the component's own
            // maximum), i.e. at most X1 + bw plus a few roundings of relative size
            // 2^-24 -- covered by one more bw and 2^-20 of the bound; from below they
            // cannot get under x0 (rounding is monotone). Inside [0, L) nothing moves.
            const float x1L = compL == 0 ? qp.X1[0] : compL == 1 ? qp.X1[1] : qp.X1[2];
            float top = __fadd_rn(x1L, __fadd_rn(bw, bw));
            top = __fadd_rn(top, __fmul_rn(fabsf(top), 9.5367431640625e-07f));
            const bool inBox = x0L >= 0.0f && top < L;
            ent.z |= (uint32_t)compL << 8 | (isPos && !inBox ? 1u << 10 : 0u);
        }
        if (__any_sync(0xffffffffu, bad)) {
            if (lane == 0) atomicCAS(p.error, 0, kErrPayload);
            return;
        }
        if (__any_sync(0xffffffffu, damaged) && lane == 0) p.dmgList[atomicAdd(p.dmgCount, 1u)] = (uint32_t)task;
    }
    const uint32_t rowAddr = (uint32_t)__cvta_generic_to_shared(st) + lane * 128 + ((lane & 7) << 4);

    for (int k = 0; k < gCnt; k++) {
        const uint32_t first = __shfl_sync(0xffffffffu, ent.x, k);
        const uint32_t minDelta = __shfl_sync(0xffffffffu, ent.y, k);
        const uint32_t zz = __shfl_sync(0xffffffffu, ent.z, k);
        const uint32_t off16 = __shfl_sync(0xffffffffu, ent.w, k);
        const int w = (int)(zz & 0xffu);
        const int comp = (int)(zz >> 8) & 3;
        const int b = g0 + k - comp * fd.nb;
        const int len = min(kBlockLen, fd.n - b * kBlockLen);
        const int size16 = (((len * w + 31) >> 5) + 3) >> 2;

        // ---- unpack ----
        const uint4 *src = payload + off16;
        for (int c = lane; c < size16; c += 32) {
            uint4 x = ld_stream_u4(src + c);
            const int m = c * 4;
            st[skew(m + 0)] = x.x;
            st[skew(m + 1)] = x.y;
            st[skew(m + 2)] = x.z;
            st[skew(m + 3)] = x.w;
        }
        __syncwarp();
        uint32_t r[kChunk];

        if (kToQ) {
            // literal DFunc boundary: integer scan, uint64 quantised values out
            unpack_dispatch(w, r, st, lane);
            __syncwarp();
            r[0] = lane ? r[0] + minDelta : 0u;
#pragma unroll
            for (int i = 1; i < kChunk; i++) r[i] = r[i] + minDelta + r[i - 1];
            const uint32_t run = r[kChunk - 1];
            uint32_t incl = run;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += up;
            }
            const uint32_t base = first + (incl - run);
            uint64_t *dst = p.qOut + fd.ioOff + (int64_t)comp * fd.n + (int64_t)b * kBlockLen;
#pragma unroll
            for (int i = 0; i < kChunk; i++) st[skew(lane * kChunk + i)] = base + r[i];
            __syncwarp();
#pragma unroll 4
            for (int kk = 0; kk < 32; kk++) {
                int el = kk * 32 + lane;
                if (el < len) dst[el] = st[skew(el)];
            }
            __syncwarp();
            continue;
        }

        // ---- deltas as floats, lane totals ----
        // Residuals of at most 23 bits come out of the unpack as floats 2^23 + r
        // (bitpack.cuh), and c = minDelta - 2^23 is a float when minDelta >= -2^23:
        // then d = r + minDelta is one exact FP32 add and the inverse delta needs no
        // integer pipe at all. Anything else adds in integers and converts (the
        // deltas of valid data are below 2^24 in magnitude either way).
        float T = 0.0f;  // sum of the lane's deltas
        if (w <= 23) unpack_dispatch_magic(w, r, st, lane);
        else unpack_dispatch_wide(w, r, st, lane);
        if (w <= 23 && (int32_t)minDelta >= -(1 << 23)) {
            const float c = __fsub_rn((float)(int32_t)minDelta, 8388608.0f);
#pragma unroll
            for (int i = 0; i < kChunk; i++) {
                float d = __fadd_rn(__uint_as_float(r[i]), c);
                if (i == 0 && lane == 0) d = 0.0f;  // the block's first value is `first` itself
                T = __fadd_rn(T, d);
                r[i] = __float_as_uint(d);
            }
        } else {
            if (w <= 23) {
#pragma unroll
                for (int i = 0; i < kChunk; i++) r[i] &= 0x7fffffu;
            }
            uint32_t t = 0u;
#pragma unroll
            for (int i = 0; i < kChunk; i++) {
                uint32_t d = r[i] + minDelta;
                if (i == 0 && lane == 0) d = 0u;
                t += d;
                r[i] = __float_as_uint((float)(int32_t)d);
            }
            T = (float)(int32_t)t;
        }
        __syncwarp();  // everyone has read its packed words: the stage becomes the float area
        float *sf = reinterpret_cast<float *>(st);
#pragma unroll
        for (int j = 0; j < 8; j++)  // chunk j of row `lane` sits at chunk position j ^ (lane & 7)
            sts_u4(rowAddr ^ (j << 4), make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]));

        // ---- inverse delta across lanes: exclusive prefix of the lane totals ----
        // (taken from the lane below, not as incl - T: the last lane of a ragged
        // block sums stale words past the payload, and in floating point that does
        // not cancel)
        float incl = T;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            float up = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl = __fadd_rn(incl, up);
        }
        const float excl = __shfl_up_sync(0xffffffffu, incl, 1);
        const float basef = lane ? __fadd_rn((float)first, excl) : (float)first;

        // ---- inverse delta within the lane + dither + de-quantise + wrap, in place ----
        const float x0 = __shfl_sync(0xffffffffu, x0L, k);
        float *dst = p.out + fd.ioOff + (int64_t)comp * fd.n + (int64_t)b * kBlockLen;
        if (len == kBlockLen && foldOK) {
            if (zz & (1u << 10))
                dequant_row<false, true, true>(basef, x0, bw, bwS, true, L, lane, len, s0, s1, rowAddr);
            else
                dequant_row<false, false, true>(basef, x0, bw, bwS, false, L, lane, len, s0, s1, rowAddr);
        } else {
            dequant_row<true, true, false>(basef, x0, bw, bwS, isPos, L, lane, len, s0, s1, rowAddr);
        }
        __syncwarp();
        if (len == kBlockLen && ((uintptr_t)dst & 15) == 0) {
            // element chunk c = kk*32 + lane lives in row c>>3 at chunk position
            // (lane&7) ^ (row&7): two base addresses, the rest are immediates
            const uint32_t r0 = lane >> 3, jj = lane & 7;
            const uint32_t aEven = (uint32_t)__cvta_generic_to_shared(st) + ((r0 * 8 + (jj ^ r0)) << 4);
            const uint32_t aOdd = aEven ^ 64u;
            float4 *d4 = reinterpret_cast<float4 *>(dst) + lane;
#pragma unroll
            for (int kk = 0; kk < 8; kk++) __stcs(d4 + kk * 32, lds_f4(((kk & 1) ? aOdd : aEven) + kk * 512));
        } else if (len == kBlockLen) {
            // A full block whose destination is not 16-byte aligned (particle counts that are not
            // multiples of four): the same 16-byte reads from the staging area, and each chunk goes
            // out as the widest stores its address allows -- two 8-byte ones, or 4 + 8 + 4.
            const uint32_t r0 = lane >> 3, jj = lane & 7;
            const uint32_t aEven = (uint32_t)__cvta_generic_to_shared(st) + ((r0 * 8 + (jj ^ r0)) << 4);
            const uint32_t aOdd = aEven ^ 64u;
            float *d1 = dst + lane * 4;
            if (((uintptr_t)dst & 7) == 0) {
#pragma unroll
                for (int kk = 0; kk < 8; kk++) {
                    const float4 v = lds_f4(((kk & 1) ? aOdd : aEven) + kk * 512);
                    __stcs(reinterpret_cast<float2 *>(d1 + kk * 128), make_float2(v.x, v.y));
                    __stcs(reinterpret_cast<float2 *>(d1 + kk * 128 + 2), make_float2(v.z, v.w));
                }
            } else {
#pragma unroll
                for (int kk = 0; kk < 8; kk++) {
                    const float4 v = lds_f4(((kk & 1) ? aOdd : aEven) + kk * 512);
                    __stcs(d1 + kk * 128, v.x);
                    __stcs(reinterpret_cast<float2 *>(d1 + kk * 128 + 1), make_float2(v.y, v.z));
                    __stcs(d1 + kk * 128 + 3, v.w);
                }
            }
        } else {
#pragma unroll 4
            for (int kk = 0; kk < 32; kk++) {
                int el = kk * 32 + lane;
                if (el < len) dst[el] = sf[swz_chunk(el >> 5, (el >> 2) & 7) * 4 + (el & 3)];
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------
// After k_unpack_dequant: the values of damaged blocks (segment image, a block
// checksum failed) become NaN -- all ones at the DFunc boundary -- so that only
// those particles are lost (doc/header_format/header_format.tex:188-196). One
// CTA walks the (normally empty) list of tasks and resets it.
// ---------------------------------------------------------------------------
template <bool kToQ>
__global__ void __launch_bounds__(128) k_damaged_tasks(DecodeParams p) {
    const uint32_t n = *p.dmgCount;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t i = warp; i < n; i += 4) {
        const int64_t task = p.dmgList[i];
        int f;
        if (p.uniformTasks > 0) {
            f = (int)(task / p.uniformTasks);
        } else {
            const DevField *fields = p.fields;
            f = find_field(p.nFields, task, [fields](int k) { return fields[k].firstTask; });
        }
        const DevField fd = p.fields[f];
        const uint4 *table = reinterpret_cast<const uint4 *>(p.in + p.offsets[f] + 64);
        const int g0 = (int)(task - fd.firstTask) * kRun;
        const int gCnt = min(3 * fd.nb - g0, kRun);
        for (int k = 0; k < gCnt; k++) {
            if (__ldg(table + g0 + k).z != kWidthDamaged) continue;
            const int comp = (g0 + k) / fd.nb, b = g0 + k - comp * fd.nb;
            const int len = min(kBlockLen, fd.n - b * kBlockLen);
            const int64_t at = fd.ioOff + (int64_t)comp * fd.n + (int64_t)b * kBlockLen;
            for (int el = lane; el < len; el += 32) {
                if (kToQ) p.qOut[at + el] = ~0ull;
                else p.out[at + el] = __int_as_float(0x7fc00000);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) *p.dmgCount = 0u;
}

}  // namespace mnw

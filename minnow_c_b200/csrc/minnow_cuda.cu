// minnow_cuda.cu -- host side of libminnow_cuda.so: the extern "C" ABI of
// include/minnow_cuda.h over the kernels in encode_kernels.cuh /
// decode_kernels.cuh / gen_kernels.cuh. sm_100a only; build with
//   nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo -O3
// There is no CPU path: without a device every entry point aborts.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "minnow_cuda.h"
#include "decode_kernels.cuh"
#include "enc_kernel.cuh"
#include "gen_kernels.cuh"
#include "quant_kernels.cuh"

using namespace mnw;

// Panic convention of /root/reference/src/debug.h:72-89: message, exit(1).
#define CUDA_CHECK(expr)                                                                 \
    do {                                                                                 \
        cudaError_t err_ = (expr);                                                       \
        if (err_ != cudaSuccess) {                                                       \
            fprintf(stderr, "Panic at file %s, line %d:\nCUDA error %s: %s\n", __FILE__, \
                    __LINE__, cudaGetErrorName(err_), cudaGetErrorString(err_));         \
            exit(1);                                                                     \
        }                                                                                \
    } while (0)

#define MNW_PANIC(...)                                                       \
    do {                                                                     \
        fprintf(stderr, "Panic at file %s, line %d:\n", __FILE__, __LINE__); \
        fprintf(stderr, __VA_ARGS__);                                        \
        fprintf(stderr, "\n");                                               \
        exit(1);                                                             \
    } while (0)

namespace {

// A device buffer that only ever grows (scratch is reused across calls).
struct Scratch {
    void *ptr = nullptr;
    size_t cap = 0;
    void *need(size_t bytes) {
        if (bytes > cap) {
            if (ptr) CUDA_CHECK(cudaFree(ptr));
            size_t want = bytes + bytes / 4 + 256;
            CUDA_CHECK(cudaMalloc(&ptr, want));
            cap = want;
        }
        return ptr;
    }
    void release() {
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        cap = 0;
    }
};

}  // namespace

struct MnwCudaCtx {
    int device = 0;
    int64_t launches = 0;
    uint64_t ditherSeed = 0;
    Scratch fields, quant, minmax, tileState, seeds, tableOf, states, misc, ring;
    Scratch ioA, ioB, ioC, ioD;  // host-buffer entry points: device staging
    Scratch dmg;                 // tasks holding a damaged block (k_damaged_tasks)
    Scratch jumpTab;        // columns of J^k for k < tabStreams (k_jump_table), built on first use
    int64_t tabStreams = 0;
    cudaEvent_t tabEv = nullptr;
    uint64_t *jumpPow = nullptr;
    int *error = nullptr;     // device flag
    uint32_t *ticket = nullptr;
    cudaStream_t own = nullptr;  // stream of the synchronous host-buffer entry points
    cudaStream_t copyIn = nullptr, copyOut = nullptr;
    int smCount = 148;
    int64_t encodeLag = -1;      // groups of 16 blocks between the A and B passes (-1: one field + encodeMargin)
    int64_t encodeMargin = -1;   // groups of slack so that a field's parameters are ready when its B work starts; -1: 128
                                 // to 384 by field size for batches that take the uniform reduce, 384 for the general
                                 // one (its parameters arrive later: a fence, a record and a fold per window)
                                 // (smaller: B work waits for them; larger: its input has left L2 again)
    void *encodeTiled = nullptr;  // cuTensorMapEncodeTiled, resolved through the runtime (no libcuda link)
    unsigned long long *encStats = nullptr;  // MNWCUDA_ENCODE_STATS
    EncodeParams lastEnc;        // most recent encode launch (diagnostics on a device-side timeout)
    bool haveLastEnc = false;
    bool timing = false;         // mnwcuda_EnableKernelTiming
    cudaEvent_t ev[10] = {};     // encode: 0..2, decode: 5..8
    int64_t hostChunkFloats = 32ll << 20;  // raw floats per pipeline chunk of the host-buffer entry points
    cudaEvent_t pipeEv[6] = {};  // host-buffer pipeline: copied in, kernels done, copied out (two sets each)
    bool evSet[10] = {};
};

static inline void mark(MnwCudaCtx *ctx, int i, cudaStream_t s) {
    if (!ctx->timing) return;
    CUDA_CHECK(cudaEventRecord(ctx->ev[i], s));
    ctx->evSet[i] = true;
}

// ---------------------------------------------------------------------------
// sizes
// ---------------------------------------------------------------------------
static inline int64_t nblocks_of(int32_t n) { return ((int64_t)n + kBlockLen - 1) / kBlockLen; }

extern "C" int64_t mnwcuda_FieldFixedBytes(int32_t n) {
    return (int64_t)sizeof(MnwDiffHeader) + 3 * nblocks_of(n) * (int64_t)sizeof(MnwDiffBlock);
}

extern "C" int64_t mnwcuda_FieldBound(int32_t n) {
    int64_t perBlock = ((int64_t)kBlockLen * kMaxWidth / 8 + 15) / 16 * 16;
    return mnwcuda_FieldFixedBytes(n) + 3 * nblocks_of(n) * perBlock;
}

// ---------------------------------------------------------------------------
// dither RNG on the host: only used once per context, to tabulate the jump
// matrix powers that k_stream_states multiplies with (src/rand.c:80-112).
// ---------------------------------------------------------------------------
static inline uint64_t host_next(uint64_t s[2]) {
    uint64_t a = s[0], b = s[1], sum = a + b;
    b ^= a;
    s[0] = ((a << 55) | (a >> 9)) ^ b ^ (b << 14);
    s[1] = (b << 36) | (b >> 28);
    return sum;
}

static void host_jump(uint64_t s[2]) {
    static const uint64_t poly[2] = {0xbeac0467eba5facbull, 0xd86b048b86aa9922ull};
    uint64_t a = 0, b = 0;
    for (int w = 0; w < 2; w++)
        for (int bit = 0; bit < 64; bit++) {
            if ((poly[w] >> bit) & 1ull) { a ^= s[0]; b ^= s[1]; }
            host_next(s);
        }
    s[0] = a;
    s[1] = b;
}

// columns of J^(2^k), k < kJumpPows; column i = image of basis vector e_i
static std::vector<uint64_t> jump_powers() {
    std::vector<uint64_t> m((size_t)kJumpPows * 128 * 2);
    for (int i = 0; i < 128; i++) {
        uint64_t s[2] = {i < 64 ? 1ull << i : 0ull, i >= 64 ? 1ull << (i - 64) : 0ull};
        host_jump(s);
        m[(size_t)i * 2] = s[0];
        m[(size_t)i * 2 + 1] = s[1];
    }
    for (int k = 1; k < kJumpPows; k++) {
        const uint64_t *prev = &m[(size_t)(k - 1) * 256];
        uint64_t *cur = &m[(size_t)k * 256];
        for (int i = 0; i < 128; i++) {  // cur col i = prev * (prev col i)
            uint64_t v0 = prev[i * 2], v1 = prev[i * 2 + 1], a = 0, b = 0;
            for (int j = 0; j < 64; j++) {
                if ((v0 >> j) & 1ull) { a ^= prev[j * 2]; b ^= prev[j * 2 + 1]; }
                if ((v1 >> j) & 1ull) { a ^= prev[(64 + j) * 2]; b ^= prev[(64 + j) * 2 + 1]; }
            }
            cur[i * 2] = a;
            cur[i * 2 + 1] = b;
        }
    }
    return m;
}

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------
extern "C" MnwCudaCtx *mnwcuda_Create(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        MNW_PANIC("libminnow_cuda needs a CUDA device and has no CPU fallback (%s).",
                  cudaGetErrorString(e));
    if (device < 0) CUDA_CHECK(cudaGetDevice(&device));
    CUDA_CHECK(cudaSetDevice(device));
    MnwCudaCtx *ctx = new MnwCudaCtx();
    ctx->device = device;
    std::vector<uint64_t> jp = jump_powers();
    CUDA_CHECK(cudaMalloc(&ctx->jumpPow, jp.size() * sizeof(uint64_t)));
    CUDA_CHECK(cudaMemcpy(ctx->jumpPow, jp.data(), jp.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMalloc(&ctx->error, 256));
    CUDA_CHECK(cudaMemset(ctx->error, 0, 256));
    ctx->ticket = reinterpret_cast<uint32_t *>(ctx->error) + 16;
    CUDA_CHECK(cudaDeviceGetAttribute(&ctx->smCount, cudaDevAttrMultiProcessorCount, device));
    if (const char *e = getenv("MNWCUDA_ENCODE_LAG")) ctx->encodeLag = atoll(e);
    if (const char *e = getenv("MNWCUDA_HOST_CHUNK_FLOATS")) ctx->hostChunkFloats = atoll(e) < 1 ? 1 : atoll(e);
    if (const char *e = getenv("MNWCUDA_ENCODE_MARGIN")) ctx->encodeMargin = atoll(e) < 0 ? 0 : atoll(e);
    {
        cudaDriverEntryPointQueryResult qres;
        CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ctx->encodeTiled, cudaEnableDefault, &qres));
        if (qres != cudaDriverEntryPointSuccess || !ctx->encodeTiled)
            MNW_PANIC("cuTensorMapEncodeTiled is not available from this driver.");
    }
    if (getenv("MNWCUDA_ENCODE_STATS")) {
        CUDA_CHECK(cudaMalloc(&ctx->encStats, 256));
        CUDA_CHECK(cudaMemset(ctx->encStats, 0, 256));
    }
    CUDA_CHECK(cudaFuncSetAttribute(k_encode<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kEncSmemBytes));
    CUDA_CHECK(cudaFuncSetAttribute(k_encode<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kEncSmemBytes));
    CUDA_CHECK(cudaFuncSetAttribute(k_encode<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kEncSmemBytes));
    CUDA_CHECK(cudaFuncSetAttribute(k_encode<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kEncSmemBytes));
    CUDA_CHECK(cudaFuncSetAttribute(k_encode<false, false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kEncSmemBytes));
    CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->own, cudaStreamNonBlocking));
    CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->copyIn, cudaStreamNonBlocking));
    CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->copyOut, cudaStreamNonBlocking));
    for (cudaEvent_t &e : ctx->pipeEv) CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CUDA_CHECK(cudaEventCreateWithFlags(&ctx->tabEv, cudaEventDisableTiming));
    return ctx;
}

extern "C" void mnwcuda_Destroy(MnwCudaCtx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    Scratch *all[] = {&ctx->fields, &ctx->quant, &ctx->minmax, &ctx->tileState, &ctx->seeds,
                      &ctx->tableOf, &ctx->states, &ctx->misc, &ctx->ring, &ctx->ioA, &ctx->ioB, &ctx->ioC,
                      &ctx->ioD, &ctx->jumpTab, &ctx->dmg};
    for (Scratch *s : all) s->release();
    cudaFree(ctx->jumpPow);
    cudaFree(ctx->error);
    cudaStreamDestroy(ctx->own);
    cudaStreamDestroy(ctx->copyIn);
    cudaStreamDestroy(ctx->copyOut);
    for (cudaEvent_t e : ctx->ev)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->pipeEv)
        if (e) cudaEventDestroy(e);
    if (ctx->tabEv) cudaEventDestroy(ctx->tabEv);
    if (ctx->encStats) {
        unsigned long long h[32];
        cudaMemcpy(h, ctx->encStats, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[mnwcuda encode stats] stage timeline (clk): fill issued -> consumer has the data %.0f | "
                        "handed back -> refilled %.0f | producer per pass: descriptors %.0f, filling %.0f (%llu passes)\n",
                (double)h[16] / (double)(h[17] ? h[17] : 1), (double)h[20] / (double)(h[21] ? h[21] : 1),
                (double)h[22] / (double)(h[24] ? h[24] : 1), (double)h[23] / (double)(h[24] ? h[24] : 1), h[24]);
        fprintf(stderr, "[mnwcuda encode stats] service warps: scan windows %llu, complete at first look %.1f%% | reduce windows "
                        "%llu, complete at first look %.1f%% (100%%: the warp never waits, everybody waits for it)\n",
                h[25], 100.0 * (double)h[26] / (double)(h[25] ? h[25] : 1), h[27], 100.0 * (double)h[28] / (double)(h[27] ? h[27] : 1));
        const double tot = (double)h[7] > 0 ? (double)h[7] : 1.0;
        fprintf(stderr, "[mnwcuda encode stats] consumer-time shares: A %.1f%% B %.1f%% queue %.1f%% | clk/A-block %.0f "
                        "clk/B-block %.0f | blocks A %llu B %llu\n",
                100.0 * h[0] / tot, 100.0 * h[1] / tot, 100.0 * h[4] / tot, (double)h[0] / (double)(h[5] ? h[5] : 1),
                (double)h[1] / (double)(h[6] ? h[6] : 1), h[5], h[6]);
        const double nB = (double)(h[6] ? h[6] : 1);
        fprintf(stderr, "[mnwcuda encode stats] B blocks planned before their parameters were published (the consumer fetches them): %.1f%%\n",
                100.0 * (double)h[29] / nB);
        fprintf(stderr, "[mnwcuda encode stats] B blocks that found all of the consumer's records taken: %.1f%% | arena without room: %.2f times per block\n",
                100.0 * (double)h[18] / nB, (double)h[19] / nB);
        fprintf(stderr, "[mnwcuda encode stats] ... of which not published when the consumer looked either: %.1f%% of all B blocks, %.1f polls each\n",
                100.0 * (double)h[30] / nB, (double)h[31] / (double)(h[30] ? h[30] : 1));
        fprintf(stderr, "[mnwcuda encode stats] clk/B-block by phase: params %.0f quantise %.0f delta+width %.0f pack %.0f "
                        "wait-free-record %.0f park+place %.0f | through scratch slots %.1f%% | queue time spent on A blocks %.1f%%\n",
                h[8] / nB, h[9] / nB, h[10] / nB, h[11] / nB, h[12] / nB, h[13] / nB, 100.0 * h[15] / nB,
                100.0 * (double)h[14] / (double)(h[4] ? h[4] : 1));
        cudaFree(ctx->encStats);
    }
    delete ctx;
}

extern "C" int64_t mnwcuda_LaunchCount(const MnwCudaCtx *ctx) { return ctx->launches; }

extern "C" void mnwcuda_EnableKernelTiming(MnwCudaCtx *ctx, int on) {
    CUDA_CHECK(cudaSetDevice(ctx->device));
    ctx->timing = on != 0;
    if (ctx->timing)
        for (cudaEvent_t &e : ctx->ev)
            if (!e) CUDA_CHECK(cudaEventCreate(&e));
}

// ms[0..1] = k_encode (+k_depth_empty), k_finalize; ms[2] = k_block_headers (mnwcuda_BlockHeadersDev); ms[3] unused (-1)
// ms[4..6] = k_parse_headers, k_stream_states, k_unpack_dequant
extern "C" int mnwcuda_KernelTimes(MnwCudaCtx *ctx, float *ms) {
    if (!ctx->timing) return -1;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    static const int first[7] = {0, 1, 3, -1, 5, 6, 7};
    for (int k = 0; k < 7; k++) {
        int a = first[k], b = a + 1;
        ms[k] = -1.0f;
        if (a < 0) continue;
        if (ctx->evSet[a] && ctx->evSet[b]) {
            CUDA_CHECK(cudaEventSynchronize(ctx->ev[b]));
            CUDA_CHECK(cudaEventElapsedTime(&ms[k], ctx->ev[a], ctx->ev[b]));
        }
    }
    return 0;
}

extern "C" int mnwcuda_Status(MnwCudaCtx *ctx, void *stream) {
    int err = 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    CUDA_CHECK(cudaMemcpyAsync(&err, ctx->error, sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
    if (err != 0) {
        if (err == kErrTimeout) {
            int info[4] = {0, 0, 0, 0};
            CUDA_CHECK(cudaMemcpy(info, ctx->error, sizeof(info), cudaMemcpyDeviceToHost));
            fprintf(stderr, "[mnwcuda] device-side wait timed out: site %d, tile %d\n", info[1], info[2]);
            if (ctx->haveLastEnc) {
                const EncodeParams &q = ctx->lastEnc;
                uint32_t st[4] = {0, 0, 0, 0};
                cudaMemcpy(&st[0], q.ticket, 4, cudaMemcpyDeviceToHost);
                fprintf(stderr, "[mnwcuda] nBlocks %lld nGroups %lld lagG %lld | ticket %u\n", (long long)q.nBlocks,
                        (long long)q.nGroups, (long long)q.lagG, st[0]);
                // first block without a prefix / a size / a min-max, first component without parameters
                const size_t nb = (size_t)q.nBlocks;
                std::vector<uint64_t> mm(nb), fq((size_t)q.nFields * 6);
                std::vector<uint32_t> pre(nb);
                std::vector<uint16_t> agg(nb);
                cudaMemcpy(pre.data(), q.blkRel, nb * 4, cudaMemcpyDeviceToHost);
                cudaMemcpy(mm.data(), q.blkMM, nb * 8, cudaMemcpyDeviceToHost);
                cudaMemcpy(agg.data(), q.blkAgg, nb * 2, cudaMemcpyDeviceToHost);
                cudaMemcpy(fq.data(), q.fq, fq.size() * 8, cudaMemcpyDeviceToHost);
                size_t iPre = 0, iAgg = 0, iMM = 0, iFq = 0, nAggMissing = 0;
                while (iPre < nb && (pre[iPre] >> 31)) iPre++;
                while (iAgg < nb && (agg[iAgg] >> 15)) iAgg++;
                for (size_t i = 0; i < nb; i++) nAggMissing += !(agg[i] >> 15);
                while (iMM < nb && mm[iMM]) iMM++;
                while (iFq < fq.size() && (fq[iFq] >> 63)) iFq++;
                fprintf(stderr, "[mnwcuda] first missing: prefix %zu size %zu (missing in all %zu) minmax %zu params word %zu of %zu\n",
                        iPre, iAgg, nAggMissing, iMM, iFq, fq.size());
                if (q.winMM) {
                    std::vector<uint64_t> wm(nb / kRedWin + 1);
                    cudaMemcpy(wm.data(), q.winMM, (nb / kRedWin) * 8, cudaMemcpyDeviceToHost);
                    size_t iW = 0, nMiss = 0;
                    while (iW < nb / kRedWin && wm[iW]) iW++;
                    for (size_t i = 0; i < nb / kRedWin; i++) nMiss += !wm[i];
                    fprintf(stderr, "[mnwcuda] reduce windows: first missing %zu (missing in all %zu of %zu)\n", iW, nMiss,
                            nb / kRedWin);
                }
            }
        }
        CUDA_CHECK(cudaMemsetAsync(ctx->error, 0, sizeof(int), s));
        CUDA_CHECK(cudaStreamSynchronize(s));
    }
    return err;
}

// ---------------------------------------------------------------------------
// batch plan: host descriptors -> DevField array + totals
// ---------------------------------------------------------------------------
namespace {

constexpr int64_t kTabMinFields = 16;       // below: k_stream_states (one product per set bit of k per field)
constexpr int64_t kTabMaxStreams = 1 << 16;  // above: the table would exceed 128 MiB

struct Plan {
    std::vector<DevField> fields;
    int64_t nBlocks = 0, nTasks = 0, fixedTotal = 0, maxStreams = 0;
    int uniformBlocks = 0, uniformTasks = 0;
    int64_t maxBlocksPerField = 0;
    int64_t floatEnd = 0;  // one past the last input float any field refers to
    bool hasEmpty = false;
    bool aligned32 = true;  // every component of every field starts a multiple of 32 floats from the batch's base
};

int build_plan(const MnwFieldDesc *h, int64_t nFields, Plan &pl) {
    pl.fields.resize((size_t)nFields);
    bool uniform = true;
    for (int64_t i = 0; i < nFields; i++) {
        const MnwFieldDesc &d = h[i];
        if (d.ParticleLen < 0 || d.Offset < 0) return -1;
        if (d.FieldCode != field_Posn && d.FieldCode != field_Velc && d.FieldCode != field_Ptid &&
            d.FieldCode != field_Unsi)
            return -3;
        DevField &f = pl.fields[(size_t)i];
        f.ioOff = d.Offset;
        f.fixedBase = pl.fixedTotal;
        f.firstBlock = pl.nBlocks;
        f.firstTask = pl.nTasks;
        f.n = d.ParticleLen;
        f.nb = (int32_t)nblocks_of(d.ParticleLen);
        f.fieldCode = d.FieldCode;
        f.delta = d.Delta;
        f.width = d.Width;
        f.seed = d.DitherSeed;
        int64_t tasks = (3 * (int64_t)f.nb + kRun - 1) / kRun;
        pl.fixedTotal += mnwcuda_FieldFixedBytes(d.ParticleLen);
        pl.nBlocks += 3 * (int64_t)f.nb;
        pl.nTasks += tasks;
        if (tasks * 32 > pl.maxStreams) pl.maxStreams = tasks * 32;
        if (3 * (int64_t)f.nb > pl.maxBlocksPerField) pl.maxBlocksPerField = 3 * (int64_t)f.nb;
        if (d.Offset + 3ll * d.ParticleLen > pl.floatEnd) pl.floatEnd = d.Offset + 3ll * d.ParticleLen;
        if (d.ParticleLen == 0) pl.hasEmpty = true;
        if ((d.Offset & 31) || (d.ParticleLen & 31)) pl.aligned32 = false;
        if (d.ParticleLen != h[0].ParticleLen || d.ParticleLen == 0) uniform = false;
    }
    if (uniform && nFields > 0) {
        pl.uniformBlocks = 3 * pl.fields[0].nb;
        pl.uniformTasks = (int)((3 * (int64_t)pl.fields[0].nb + kRun - 1) / kRun);
    }
    if (pl.nBlocks >= (1ll << 31) || pl.nTasks >= (1ll << 31)) return -5;
    return 0;
}

EncodeParams prepare_encode(MnwCudaCtx *ctx, const Plan &pl, uint8_t *d_out, int64_t outCap,
                            uint64_t *d_offsets, cudaStream_t s) {
    const size_t nF = pl.fields.size();
    EncodeParams p;
    memset(&p, 0, sizeof(p));
    p.fields = static_cast<DevField *>(ctx->fields.need(nF * sizeof(DevField)));
    p.nFields = (int)nF;
    p.uniformBlocks = pl.uniformBlocks;
    p.minmax = static_cast<uint32_t *>(ctx->minmax.need(nF * 6 * sizeof(uint32_t)));
    p.quant = static_cast<DevQuant *>(ctx->quant.need(nF * sizeof(DevQuant)));
    // per-consumer scratch slots for packed blocks waiting for their offset
    p.scratch = static_cast<uint32_t *>(
        ctx->ring.need((size_t)ctx->smCount * kEncConsumers * kFifo * kSlotWords * sizeof(uint32_t)));
    // one zeroed region: window bases | block min/max | component parameters | block prefixes | block sizes | ticket
    const size_t nWin = (size_t)((pl.nBlocks + kScanWin - 1) / kScanWin);
    const size_t padScan = nWin * kScanWin;
    const size_t padRed = (size_t)((pl.nBlocks + kRedWin - 1) / kRedWin) * kRedWin;
    const size_t bWin = ((nWin + 1) * 8 + 15) & ~(size_t)15, bMM = padRed * 8, bFq = nF * 6 * 8;
    const size_t bRel = padScan * 4, bAgg = ((padScan * 2 + 15) & ~(size_t)15) + 16;
    const size_t bWinMM = (padRed / kRedWin) * 32;  // 8 bytes per window (uniform batches) or a 32-byte record
    const size_t stateBytes = bWin + bMM + bFq + bRel + bAgg + bWinMM;
    p.tbl = static_cast<uint4 *>(ctx->misc.need((size_t)(pl.nBlocks + 1) * sizeof(uint4)));
    uint8_t *st = static_cast<uint8_t *>(ctx->tileState.need(stateBytes));
    p.winBase = reinterpret_cast<uint64_t *>(st);
    p.blkMM = reinterpret_cast<uint64_t *>(st + bWin);
    p.fq = reinterpret_cast<uint64_t *>(st + bWin + bMM);
    p.blkRel = reinterpret_cast<uint32_t *>(st + bWin + bMM + bFq);
    p.blkAgg = reinterpret_cast<uint16_t *>(st + bWin + bMM + bFq + bRel);
    p.ticket = reinterpret_cast<uint32_t *>(st + bWin + bMM + bFq + bRel + bAgg - 16);
    p.winMM = reinterpret_cast<uint64_t *>(st + bWin + bMM + bFq + bRel + bAgg);
    p.out = d_out;
    p.outCap = outCap;
    p.offsets = d_offsets;
    p.error = ctx->error;
    p.nBlocks = pl.nBlocks;
    p.lagG = -1;
    // pageable source: the copy is staged before the call returns
    CUDA_CHECK(cudaMemcpyAsync(const_cast<DevField *>(p.fields), pl.fields.data(), nF * sizeof(DevField),
                               cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemsetAsync(st, 0, stateBytes, s));
    return p;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

void launch_encode(MnwCudaCtx *ctx, const EncodeParams &p, int64_t floatEnd, bool fromQ, cudaStream_t s) {
    EncodeParams q = p;
    q.stats = ctx->encStats;
    if (const char *e = getenv("MNWCUDA_DBG")) q.dbg = atoi(e);
    // the input as rows of 32 floats (128 bytes) from a 128-byte aligned base: one TMA box = 32 rows = one block
    CUtensorMap map;
    memset(&map, 0, sizeof(map));
    q.tmaShift = -1;
    q.tmaRows = 0;
    if (!fromQ && p.in) {
        const uintptr_t addr = reinterpret_cast<uintptr_t>(p.in);
        const uintptr_t base = addr & ~(uintptr_t)127;
        const int64_t shift = (int64_t)((addr - base) / 4);
        const int64_t rows = (shift + floatEnd) / 32;
        if (rows >= 32 && rows < (1ll << 31)) {
            cuuint64_t dims[2] = {32, (cuuint64_t)rows};
            cuuint64_t strides[1] = {128};
            cuuint32_t box[2] = {32, 32};
            cuuint32_t elem[2] = {1, 1};
            CUresult rc = reinterpret_cast<EncodeTiledFn>(ctx->encodeTiled)(
                &map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, reinterpret_cast<void *>(base), dims, strides, box, elem,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (rc != CUDA_SUCCESS) MNW_PANIC("cuTensorMapEncodeTiled failed (%d).", (int)rc);
            q.tmaShift = shift;
            q.tmaRows = rows;
        }
    }
    // one persistent CTA per SM; fewer when the batch is tiny
    int64_t ctas = q.nGroups * kGroupSlots / kEncConsumers + 1;
    if (ctas > ctx->smCount) ctas = ctx->smCount;
    if (const char *e = getenv("MNWCUDA_ENCODE_CTAS"))  // debugging: does the time scale with the SMs?
        if (atoi(e) >= 8 && atoi(e) < ctas) ctas = atoi(e);
    if (ctas < 2) ctas = 2;  // the reduce warp lives in CTA 1
    // The CTAs of k_encode wait for each other's results (scan warp in CTA 0, reduce
    // warp in CTA 1, ticket order), so all of them must be resident at once: a
    // cooperative launch starts the grid only when every CTA has an SM, whatever else
    // the device is running; the grid is sized by the occupancy query, never above it.
    // big batches whose components are not whole reduce windows (fields of unequal size): the
    // instantiation with the general parallel reduce; everything else as before
    const bool uniformWindows = q.uniformBlocks > 0 && (q.uniformBlocks / 3) % kRedWin == 0;
    // ... and batches with blocks off the 128-byte grid (its A and B work fetch those with vector loads)
    const bool offGrid = q.offGrid || (q.tmaShift & 31) != 0;
    const bool general = !fromQ && !q.stats && q.lagG >= 0 && ctas >= 8 && (!uniformWindows || offGrid || (q.dbg & 128));  // (128: experiment)
    const void *fn = fromQ ? (const void *)k_encode<true, false>
                   : general ? (offGrid ? (const void *)k_encode<false, false, true, true> : (const void *)k_encode<false, false, true>)
                   : q.stats ? (const void *)k_encode<false, true> : (const void *)k_encode<false, false>;
    int perSm = 0;
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, fn, kEncWarps * 32, kEncSmemBytes));
    if (perSm < 1) MNW_PANIC("k_encode does not fit an SM of this device.");
    if (ctas > (int64_t)perSm * ctx->smCount) ctas = (int64_t)perSm * ctx->smCount;
    ctx->lastEnc = q;
    ctx->haveLastEnc = true;
    void *args[2] = {&q, &map};
    CUDA_CHECK(cudaLaunchCooperativeKernel(fn, dim3((unsigned)ctas), dim3(kEncWarps * 32), args, kEncSmemBytes, s));
    ctx->launches++;
    CUDA_CHECK(cudaGetLastError());
}

void launch_finalize(MnwCudaCtx *ctx, const Plan &pl, const EncodeParams &p, cudaStream_t s) {
    const int nF = (int)pl.fields.size();
    k_finalize<<<(unsigned)((nF + 1 + 127) / 128), 128, 0, s>>>(p, pl.fixedTotal);
    ctx->launches++;
    if (pl.nBlocks > 0) {
        k_tables<<<(unsigned)((pl.nBlocks + 255) / 256), 256, 0, s>>>(p);
        ctx->launches++;
    }
    CUDA_CHECK(cudaGetLastError());
}

}  // namespace

// ---------------------------------------------------------------------------
// device-resident batch encode / decode
// ---------------------------------------------------------------------------
extern "C" int mnwcuda_EncodeFieldsDev(MnwCudaCtx *ctx, const float *d_in, const MnwFieldDesc *h_fields,
                                       int64_t nFields, uint8_t *d_out, int64_t outCap,
                                       uint64_t *d_offsets, void *stream) {
    if (!ctx || nFields < 0 || nFields >= (1ll << 30)) return -1;
    if (nFields == 0) return 0;
    if (((uintptr_t)d_out & 15) || ((uintptr_t)d_in & 3)) return -6;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Plan pl;
    int rc = build_plan(h_fields, nFields, pl);
    if (rc) return rc;
    for (int64_t i = 0; i < nFields; i++)  // float fields only: integer fields go through mnwcuda_EncodeField
        if (h_fields[i].FieldCode != field_Posn && h_fields[i].FieldCode != field_Velc) return -3;
    EncodeParams p = prepare_encode(ctx, pl, d_out, outCap, d_offsets, s);
    p.in = d_in;
    // the B work of a block trails its A work by one (largest) field plus what can be in flight
    const int64_t fieldGroups = (pl.maxBlocksPerField + kGroupBlocks - 1) / kGroupBlocks;
    const bool uniformWindows = pl.uniformBlocks > 0 && (pl.uniformBlocks / 3) % kRedWin == 0;
    // (uniform reduce: 128 groups for fields of 2^22 particles -- more and their second read starts to
    // miss L2 -- up to 384 for smaller ones, whose B work otherwise waits for parameters more often:
    // 0.40 -> 0.47 of peak at 2^19 particles per field, 0.48 -> 0.53 at 2^21)
    const int64_t margin = ctx->encodeMargin >= 0 ? ctx->encodeMargin
                           : !uniformWindows      ? 384
                                                  : std::max<int64_t>(128, std::min<int64_t>(384, 896 - fieldGroups));
    p.lagG = ctx->encodeLag >= 0 ? ctx->encodeLag : fieldGroups + margin;
    if (p.lagG < fieldGroups + 1) p.lagG = fieldGroups + 1;
    p.offGrid = pl.aligned32 ? 0 : 1;
    // timing experiments (scripts/enc_only.py; the output of such a call is not a valid stream):
    // "B" = B work alone with the parameters the previous call on this batch left in ctx->quant,
    // "A" = A work alone (MNWCUDA_DBG 64 makes the planner skip the B slots and the scan warp leave)
    const char *only = getenv("MNWCUDA_ENCODE_ONLY");
    const bool onlyB = only && only[0] == 'B';
    if (onlyB) p.lagG = -1;
    mark(ctx, 0, s);
    if (pl.hasEmpty) {
        k_depth_empty<<<(unsigned)((nFields + 127) / 128), 128, 0, s>>>(p);
        ctx->launches++;
    }
    if (pl.nBlocks > 0) {
        p.nGroups = onlyB ? (pl.nBlocks + kGroupSlots - 1) / kGroupSlots
                          : (pl.nBlocks + kGroupBlocks - 1) / kGroupBlocks + p.lagG;
        if (p.nGroups >= (1ll << 31) - 1) return -5;
        launch_encode(ctx, p, pl.floatEnd, false, s);
    }
    mark(ctx, 1, s);
    launch_finalize(ctx, pl, p, s);
    mark(ctx, 2, s);
    return 0;
}

extern "C" int mnwcuda_DecodeFieldsDev(MnwCudaCtx *ctx, const uint8_t *d_in, const uint64_t *d_offsets,
                                       const MnwFieldDesc *h_fields, int64_t nFields, float *d_out,
                                       void *stream) {
    if (!ctx || nFields < 0 || nFields >= (1ll << 30)) return -1;
    if (nFields == 0) return 0;
    if (((uintptr_t)d_in & 15) || ((uintptr_t)d_out & 3)) return -6;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Plan pl;
    int rc = build_plan(h_fields, nFields, pl);
    if (rc) return rc;
    for (int64_t i = 0; i < nFields; i++)  // float fields only: integer fields go through mnwcuda_DecodeField
        if (h_fields[i].FieldCode != field_Posn && h_fields[i].FieldCode != field_Velc) return -3;
    const size_t nF = (size_t)nFields;
    DecodeParams p;
    memset(&p, 0, sizeof(p));
    p.in = d_in;
    p.offsets = d_offsets;
    p.fields = static_cast<DevField *>(ctx->fields.need(nF * sizeof(DevField)));
    p.nFields = (int)nF;
    p.uniformTasks = pl.uniformTasks;
    p.quant = static_cast<DevQuant *>(ctx->quant.need(nF * sizeof(DevQuant)));
    p.seeds = static_cast<uint64_t *>(ctx->seeds.need(nF * sizeof(uint64_t)));
    p.tableOf = static_cast<int *>(ctx->tableOf.need(nF * sizeof(int)));
    p.maxStreams = pl.maxStreams;
    p.states = static_cast<uint64_t *>(ctx->states.need(nF * (size_t)pl.maxStreams * 16 + 16));
    p.jumpPow = ctx->jumpPow;
    p.out = d_out;
    p.qOut = nullptr;
    p.error = ctx->error;
    CUDA_CHECK(cudaMemcpyAsync(const_cast<DevField *>(p.fields), pl.fields.data(), nF * sizeof(DevField),
                               cudaMemcpyHostToDevice, s));
    mark(ctx, 5, s);
    k_parse_headers<<<(unsigned)((nFields + 127) / 128), 128, 0, s>>>(p);
    ctx->launches++;
    mark(ctx, 6, s);
    if (pl.nTasks > 0) {
        // Many fields: one matrix-vector product per stream against the tabulated J^k
        // (built once per context); few fields: the jump-matrix powers directly.
        if (nFields >= kTabMinFields && pl.maxStreams <= kTabMaxStreams) {
            if (ctx->tabStreams < pl.maxStreams) {
                CUDA_CHECK(cudaStreamSynchronize(s));  // the old table may still be read
                ulonglong2 *tab = static_cast<ulonglong2 *>(ctx->jumpTab.need((size_t)pl.maxStreams * 2048));
                k_jump_table<<<(unsigned)pl.maxStreams, 128, 0, s>>>(ctx->jumpPow, tab, pl.maxStreams);
                CUDA_CHECK(cudaEventRecord(ctx->tabEv, s));
                ctx->tabStreams = pl.maxStreams;
                ctx->launches++;
            }
            CUDA_CHECK(cudaStreamWaitEvent(s, ctx->tabEv, 0));
            p.statesByStream = 1;  // (a CTA writes its k's states of all fields as one contiguous run)
            k_stream_states_tab<<<(unsigned)pl.maxStreams, 128, 0, s>>>(p, static_cast<const ulonglong2 *>(ctx->jumpTab.ptr));
        } else {
            dim3 grid((unsigned)((pl.maxStreams + 127) / 128), (unsigned)nFields);
            k_stream_states<<<grid, 128, 0, s>>>(p);
        }
        mark(ctx, 7, s);
        const unsigned dgrid = (unsigned)((pl.nTasks + kDecWarps - 1) / kDecWarps);
        p.dmgCount = reinterpret_cast<uint32_t *>(ctx->error) + 32;  // zero between calls
        p.dmgList = static_cast<uint32_t *>(ctx->dmg.need((size_t)pl.nTasks * 4));
        k_unpack_dequant<false><<<dgrid, kDecWarps * 32, 0, s>>>(p, pl.nTasks);
        mark(ctx, 8, s);
        k_damaged_tasks<false><<<1, 128, 0, s>>>(p);
        ctx->launches += 3;
    }
    CUDA_CHECK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// host-buffer entry points
// ---------------------------------------------------------------------------
namespace {

void accuracy_of(const Field &f, float &delta, float &width) {
    if (f.Hd.FieldCode == field_Posn) {
        const PositionAccuracy *a = static_cast<const PositionAccuracy *>(f.Acc);
        if (!a) MNW_PANIC("Field has no PositionAccuracy.");
        if (a->Deltas) MNW_PANIC("Per-particle accuracies are not supported by algo_Diff v0.9.");
        delta = a->Delta;
        width = a->Width;
    } else if (f.Hd.FieldCode == field_Velc) {
        const VelocityAccuracy *a = static_cast<const VelocityAccuracy *>(f.Acc);
        if (!a) MNW_PANIC("Field has no VelocityAccuracy.");
        if (a->Deltas) MNW_PANIC("Per-particle accuracies are not supported by algo_Diff v0.9.");
        if (a->SymLog10Scaled) MNW_PANIC("symLog10 not supported.");  // src/quant.c:749
        delta = a->Delta;
        width = 0.0f;
    } else {
        MNW_PANIC("Field code %x is not supported by algo_Diff v0.9.", f.Hd.FieldCode);
    }
}

void panic_on_status(MnwCudaCtx *ctx, cudaStream_t s, const char *what) {
    int err = mnwcuda_Status(ctx, s);
    if (err == kErrDepth)
        MNW_PANIC("%s: the requested accuracy exceeds the granularity of single precision floats, "
                  "which only support 24 bits of mantissa precision.", what);  // src/quant.c:688-693
    if (err == kErrCapacity) MNW_PANIC("%s: output capacity exceeded.", what);
    if (err == kErrPayload) MNW_PANIC("%s: malformed algo_Diff payload.", what);
    if (err == kErrRange) MNW_PANIC("%s: quantised values do not fit the declared depth.", what);
    if (err == kErrTimeout) MNW_PANIC("%s: a device-side wait timed out (internal error).", what);
}

void *xmalloc(size_t bytes) {
    void *p = malloc(bytes ? bytes : 1);
    if (!p) MNW_PANIC("Allocation failed.");  // AssertAlloc, src/debug.h:36-43
    return p;
}

// fetch one field's payload (device region described by two offsets) into malloc memory
uint8_t *fetch_payload(MnwCudaCtx *ctx, const uint8_t *d_out, const uint64_t *d_off, int64_t *len) {
    uint64_t off[2];
    CUDA_CHECK(cudaMemcpyAsync(off, d_off, sizeof(off), cudaMemcpyDeviceToHost, ctx->own));
    CUDA_CHECK(cudaStreamSynchronize(ctx->own));
    *len = (int64_t)(off[1] - off[0]);
    uint8_t *data = static_cast<uint8_t *>(xmalloc((size_t)*len));
    CUDA_CHECK(cudaMemcpyAsync(data, d_out + off[0], (size_t)*len, cudaMemcpyDeviceToHost, ctx->own));
    CUDA_CHECK(cudaStreamSynchronize(ctx->own));
    return data;
}

}  // namespace

extern "C" void mnwcuda_SetDitherSeed(void *buffer, uint64_t seed) {
    static_cast<MnwCudaCtx *>(buffer)->ditherSeed = seed;
}

static CField encode_int_field(Field f, MnwCudaCtx *ctx);   // minnow_ext.inl: Ptid / Unsi
static Field decode_int_field(CField cf, MnwCudaCtx *ctx);

extern "C" CField mnwcuda_EncodeField(Field f, void *buffer) {
    MnwCudaCtx *ctx = static_cast<MnwCudaCtx *>(buffer);
    if (!ctx) MNW_PANIC("mnwcuda_EncodeField needs the buffer returned by MnwCudaCAlloc_v0_9.");
    CUDA_CHECK(cudaSetDevice(ctx->device));
    if (f.Hd.FieldCode == field_Ptid || f.Hd.FieldCode == field_Unsi) return encode_int_field(f, ctx);
    MnwFieldDesc d;
    memset(&d, 0, sizeof(d));
    d.ParticleLen = f.Hd.ParticleLen;
    d.FieldCode = f.Hd.FieldCode;
    d.DitherSeed = ctx->ditherSeed;
    accuracy_of(f, d.Delta, d.Width);
    const size_t inBytes = (size_t)d.ParticleLen * 3 * sizeof(float);
    const int64_t cap = mnwcuda_FieldBound(d.ParticleLen);
    float *d_in = static_cast<float *>(ctx->ioA.need(inBytes + 16));
    uint8_t *d_out = static_cast<uint8_t *>(ctx->ioB.need((size_t)cap));
    uint64_t *d_off = static_cast<uint64_t *>(ctx->ioC.need(2 * sizeof(uint64_t)));
    if (inBytes) CUDA_CHECK(cudaMemcpyAsync(d_in, f.Data, inBytes, cudaMemcpyHostToDevice, ctx->own));
    int rc = mnwcuda_EncodeFieldsDev(ctx, d_in, &d, 1, d_out, cap, d_off, ctx->own);
    if (rc) MNW_PANIC("mnwcuda_EncodeField: bad field (%d).", rc);
    panic_on_status(ctx, ctx->own, "mnwcuda_EncodeField");
    CField cf;
    memset(&cf, 0, sizeof(cf));
    cf.Hd = f.Hd;
    cf.Data = fetch_payload(ctx, d_out, d_off, &cf.DataLen);
    return cf;
}

extern "C" Field mnwcuda_DecodeField(CField cf, void *buffer) {
    MnwCudaCtx *ctx = static_cast<MnwCudaCtx *>(buffer);
    if (!ctx) MNW_PANIC("mnwcuda_DecodeField needs the buffer returned by MnwCudaDAlloc_v0_9.");
    CUDA_CHECK(cudaSetDevice(ctx->device));
    if (cf.Hd.FieldCode == field_Ptid || cf.Hd.FieldCode == field_Unsi) return decode_int_field(cf, ctx);
    if (cf.Hd.ParticleLen < 0 || cf.DataLen < mnwcuda_FieldFixedBytes(cf.Hd.ParticleLen))
        MNW_PANIC("mnwcuda_DecodeField: payload too short.");
    MnwDiffHeader h;
    memcpy(&h, cf.Data, sizeof(h));
    MnwFieldDesc d;
    memset(&d, 0, sizeof(d));
    d.ParticleLen = cf.Hd.ParticleLen;
    d.FieldCode = cf.Hd.FieldCode;
    const size_t outBytes = (size_t)d.ParticleLen * 3 * sizeof(float);
    uint8_t *d_in = static_cast<uint8_t *>(ctx->ioB.need((size_t)cf.DataLen + 16));
    float *d_out = static_cast<float *>(ctx->ioA.need(outBytes + 16));
    uint64_t *d_off = static_cast<uint64_t *>(ctx->ioC.need(2 * sizeof(uint64_t)));
    uint64_t off[2] = {0, (uint64_t)cf.DataLen};
    CUDA_CHECK(cudaMemcpyAsync(d_in, cf.Data, (size_t)cf.DataLen, cudaMemcpyHostToDevice, ctx->own));
    CUDA_CHECK(cudaMemcpyAsync(d_off, off, sizeof(off), cudaMemcpyHostToDevice, ctx->own));
    int rc = mnwcuda_DecodeFieldsDev(ctx, d_in, d_off, &d, 1, d_out, ctx->own);
    if (rc) MNW_PANIC("mnwcuda_DecodeField: bad field (%d).", rc);
    panic_on_status(ctx, ctx->own, "mnwcuda_DecodeField");

    Field f;
    memset(&f, 0, sizeof(f));
    f.Hd = cf.Hd;
    f.Data = xmalloc(outBytes);
    if (outBytes) CUDA_CHECK(cudaMemcpyAsync(f.Data, d_out, outBytes, cudaMemcpyDeviceToHost, ctx->own));
    CUDA_CHECK(cudaStreamSynchronize(ctx->own));
    // reported accuracy: depthToDelta on component 0 (src/quant.c:490-493, 654-662)
    float delta = (h.X1[0] - h.X0[0]) / (float)(1 << h.Depth);
    if (cf.Hd.FieldCode == field_Posn) {
        PositionAccuracy *a = static_cast<PositionAccuracy *>(calloc(1, sizeof(*a)));
        a->Delta = delta;
        a->Width = h.Width;
        f.Acc = a;
    } else {
        VelocityAccuracy *a = static_cast<VelocityAccuracy *>(calloc(1, sizeof(*a)));
        a->Delta = delta;
        f.Acc = a;
    }
    return f;
}

// Chunks of fields are staged through device buffers one after another;
// chunk size is bounded so that scratch stays modest.
// Host-buffer batches run as a three-stage pipeline over chunks of ~128 MiB of
// raw floats: copy in (stream copyIn), kernels (stream own), copy out (stream
// copyOut), two sets of device staging buffers. PCIe is full duplex, so a
// chunk's copy out overlaps the next chunk's copy in, and the kernels hide
// behind both. Pinned host buffers make the copies asynchronous; pageable ones
// work, staged by the driver.
namespace {

// fields [i, j) of one chunk, re-based to the chunk's staging buffer
int64_t host_chunk(int64_t limit, const MnwFieldDesc *h_fields, int64_t nFields, int64_t i,
                   std::vector<MnwFieldDesc> &local, int64_t &floats, int64_t &bound) {
    local.clear();
    floats = bound = 0;
    int64_t j = i;
    while (j < nFields && (j == i || floats + 3ll * h_fields[j].ParticleLen <= limit)) {
        MnwFieldDesc d = h_fields[j];
        d.Offset = floats;
        floats += 3ll * d.ParticleLen;
        bound += mnwcuda_FieldBound(d.ParticleLen);
        local.push_back(d);
        j++;
    }
    return j;
}
}  // namespace

extern "C" int64_t mnwcuda_EncodeFieldsHost(MnwCudaCtx *ctx, const float *h_in, const MnwFieldDesc *h_fields,
                                            int64_t nFields, uint8_t *h_out, int64_t outCap,
                                            uint64_t *h_offsets) {
    if (!ctx || nFields < 0) return -1;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    h_offsets[0] = 0;
    if (nFields == 0) return 0;
    struct Chunk {
        int64_t first, end, floats, bound;
        std::vector<MnwFieldDesc> local;
    };
    std::vector<Chunk> chunks;
    int64_t maxFloats = 0, maxBound = 0, maxCnt = 0;
    for (int64_t i = 0; i < nFields;) {
        Chunk c;
        c.first = i;
        c.end = host_chunk(ctx->hostChunkFloats, h_fields, nFields, i, c.local, c.floats, c.bound);
        maxFloats = std::max(maxFloats, c.floats);
        maxBound = std::max(maxBound, c.bound);
        maxCnt = std::max(maxCnt, c.end - c.first);
        i = c.end;
        chunks.push_back(std::move(c));
    }
    // two sets of staging buffers carved out of the three scratch areas
    const size_t inStride = ((size_t)maxFloats * 4 + 16 + 255) & ~(size_t)255;
    const size_t outStride = ((size_t)maxBound + 255) & ~(size_t)255;
    const size_t offStride = ((size_t)(maxCnt + 1) * 8 + 255) & ~(size_t)255;
    uint8_t *inBuf = static_cast<uint8_t *>(ctx->ioA.need(2 * inStride));
    uint8_t *outBuf = static_cast<uint8_t *>(ctx->ioB.need(2 * outStride));
    uint8_t *offBuf = static_cast<uint8_t *>(ctx->ioC.need(2 * offStride));
    cudaEvent_t *evIn = ctx->pipeEv, *evEnc = ctx->pipeEv + 2, *evOut = ctx->pipeEv + 4;
    std::vector<uint64_t> offs;
    int64_t total = 0;
    int64_t result = 0;

    auto copy_in = [&](size_t c) {
        const Chunk &ch = chunks[c];
        float *d_in = reinterpret_cast<float *>(inBuf + (c & 1) * inStride);
        // the set's previous user (chunk c-2) has been encoded: the host waited for it in its turn
        for (int64_t k = 0; k < ch.end - ch.first; k++) {
            const size_t bytes = (size_t)ch.local[(size_t)k].ParticleLen * 12;
            if (bytes)
                CUDA_CHECK(cudaMemcpyAsync(d_in + ch.local[(size_t)k].Offset, h_in + h_fields[ch.first + k].Offset,
                                           bytes, cudaMemcpyHostToDevice, ctx->copyIn));
        }
        CUDA_CHECK(cudaEventRecord(evIn[c & 1], ctx->copyIn));
    };

    copy_in(0);
    for (size_t c = 0; c < chunks.size(); c++) {
        const Chunk &ch = chunks[c];
        const int64_t cnt = ch.end - ch.first;
        const int b = (int)(c & 1);
        float *d_in = reinterpret_cast<float *>(inBuf + b * inStride);
        uint8_t *d_out = outBuf + b * outStride;
        uint64_t *d_off = reinterpret_cast<uint64_t *>(offBuf + b * offStride);
        CUDA_CHECK(cudaStreamWaitEvent(ctx->own, evIn[b], 0));
        if (c >= 2) CUDA_CHECK(cudaStreamWaitEvent(ctx->own, evOut[b], 0));  // d_out's previous payload has left
        int rc = mnwcuda_EncodeFieldsDev(ctx, d_in, ch.local.data(), cnt, d_out, ch.bound, d_off, ctx->own);
        if (rc) { result = rc; break; }
        if (c + 1 < chunks.size()) copy_in(c + 1);  // flies while this chunk is encoded and copied out
        offs.resize((size_t)cnt + 1);
        CUDA_CHECK(cudaMemcpyAsync(offs.data(), d_off, (size_t)(cnt + 1) * 8, cudaMemcpyDeviceToHost, ctx->own));
        const int err = mnwcuda_Status(ctx, ctx->own);  // waits for this chunk's kernels
        if (err) { result = -100 - err; break; }
        const int64_t bytes = (int64_t)offs[(size_t)cnt];
        if (total + bytes > outCap) { result = -2; break; }
        CUDA_CHECK(cudaMemcpyAsync(h_out + total, d_out, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->copyOut));
        CUDA_CHECK(cudaEventRecord(evOut[b], ctx->copyOut));
        for (int64_t k = 0; k <= cnt; k++) h_offsets[ch.first + k] = (uint64_t)total + offs[(size_t)k];
        total += bytes;
    }
    (void)evEnc;
    CUDA_CHECK(cudaStreamSynchronize(ctx->copyIn));
    CUDA_CHECK(cudaStreamSynchronize(ctx->own));
    CUDA_CHECK(cudaStreamSynchronize(ctx->copyOut));
    return result ? result : total;
}

extern "C" int mnwcuda_DecodeFieldsHost(MnwCudaCtx *ctx, const uint8_t *h_in, const uint64_t *h_offsets,
                                        const MnwFieldDesc *h_fields, int64_t nFields, float *h_out) {
    if (!ctx || nFields < 0) return -1;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    if (nFields == 0) return 0;
    struct Chunk {
        int64_t first, end, floats, bytes;
        std::vector<MnwFieldDesc> local;
    };
    std::vector<Chunk> chunks;
    int64_t maxFloats = 0, maxBytes = 0, maxCnt = 0;
    for (int64_t i = 0; i < nFields;) {
        Chunk c;
        int64_t bound;
        c.first = i;
        c.end = host_chunk(ctx->hostChunkFloats, h_fields, nFields, i, c.local, c.floats, bound);
        if (h_offsets[c.first] & 15) return -6;
        c.bytes = (int64_t)(h_offsets[c.end] - h_offsets[c.first]);
        maxFloats = std::max(maxFloats, c.floats);
        maxBytes = std::max(maxBytes, c.bytes);
        maxCnt = std::max(maxCnt, c.end - c.first);
        i = c.end;
        chunks.push_back(std::move(c));
    }
    const size_t inStride = ((size_t)maxBytes + 16 + 255) & ~(size_t)255;
    const size_t outStride = ((size_t)maxFloats * 4 + 16 + 255) & ~(size_t)255;
    const size_t offStride = ((size_t)(maxCnt + 1) * 8 + 255) & ~(size_t)255;
    uint8_t *inBuf = static_cast<uint8_t *>(ctx->ioB.need(2 * inStride));
    uint8_t *outBuf = static_cast<uint8_t *>(ctx->ioA.need(2 * outStride));
    uint8_t *offBuf = static_cast<uint8_t *>(ctx->ioC.need(2 * offStride));
    cudaEvent_t *evIn = ctx->pipeEv, *evDec = ctx->pipeEv + 2, *evOut = ctx->pipeEv + 4;
    std::vector<uint64_t> offs;
    int result = 0;
    for (size_t c = 0; c < chunks.size(); c++) {
        const Chunk &ch = chunks[c];
        const int64_t cnt = ch.end - ch.first;
        const int b = (int)(c & 1);
        uint8_t *d_in = inBuf + b * inStride;
        float *d_out = reinterpret_cast<float *>(outBuf + b * outStride);
        uint64_t *d_off = reinterpret_cast<uint64_t *>(offBuf + b * offStride);
        const uint64_t base = h_offsets[ch.first];
        offs.resize((size_t)cnt + 1);
        for (int64_t k = 0; k <= cnt; k++) offs[(size_t)k] = h_offsets[ch.first + k] - base;
        if (c >= 2) CUDA_CHECK(cudaStreamWaitEvent(ctx->copyIn, evDec[b], 0));  // the set's previous chunk is decoded
        // (the small pageable copy first: it is staged synchronously, behind whatever the stream holds)
        CUDA_CHECK(cudaMemcpyAsync(d_off, offs.data(), (size_t)(cnt + 1) * 8, cudaMemcpyHostToDevice, ctx->copyIn));
        CUDA_CHECK(cudaMemcpyAsync(d_in, h_in + base, (size_t)ch.bytes, cudaMemcpyHostToDevice, ctx->copyIn));
        CUDA_CHECK(cudaEventRecord(evIn[b], ctx->copyIn));
        CUDA_CHECK(cudaStreamWaitEvent(ctx->own, evIn[b], 0));
        if (c >= 2) CUDA_CHECK(cudaStreamWaitEvent(ctx->own, evOut[b], 0));  // d_out's previous floats have left
        int rc = mnwcuda_DecodeFieldsDev(ctx, d_in, d_off, ch.local.data(), cnt, d_out, ctx->own);
        if (rc) { result = rc; break; }
        CUDA_CHECK(cudaEventRecord(evDec[b], ctx->own));
        CUDA_CHECK(cudaStreamWaitEvent(ctx->copyOut, evDec[b], 0));
        for (int64_t k = 0; k < cnt; k++) {
            const size_t nb = (size_t)ch.local[(size_t)k].ParticleLen * 12;
            if (nb)
                CUDA_CHECK(cudaMemcpyAsync(h_out + h_fields[ch.first + k].Offset, d_out + ch.local[(size_t)k].Offset,
                                           nb, cudaMemcpyDeviceToHost, ctx->copyOut));
        }
        CUDA_CHECK(cudaEventRecord(evOut[b], ctx->copyOut));
    }
    CUDA_CHECK(cudaStreamSynchronize(ctx->copyIn));
    const int err = mnwcuda_Status(ctx, ctx->own);
    CUDA_CHECK(cudaStreamSynchronize(ctx->copyOut));
    if (result) return result;
    return err ? -100 - err : 0;
}

#include "minnow_ext.inl"

// ---------------------------------------------------------------------------
// synthetic input
// ---------------------------------------------------------------------------
static inline uint64_t host_splitmix(uint64_t &s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

extern "C" int mnwcuda_GenerateZeldovich(MnwCudaCtx *ctx, int32_t ngrid, float boxWidth, uint64_t seed,
                                         int64_t firstId, int64_t count, float *d_pos, float *d_vel,
                                         void *stream) {
    return mnwcuda_GenerateZeldovichRough(ctx, ngrid, boxWidth, seed, firstId, count, 0.0f, d_pos, d_vel, stream);
}

extern "C" int mnwcuda_GenerateZeldovichRough(MnwCudaCtx *ctx, int32_t ngrid, float boxWidth, uint64_t seed,
                                              int64_t firstId, int64_t count, float noiseCells, float *d_pos,
                                              float *d_vel, void *stream) {
    if (!ctx || ngrid <= 0 || count < 0 || !(noiseCells >= 0.0f)) return -1;
    if (count == 0) return 0;
    CUDA_CHECK(cudaSetDevice(ctx->device));
    // 64 integer wave-vectors with |k_i| <= 8, amplitude ~ |k|^-2 along k,
    // rescaled to an rms displacement of 1.5 cells; velocity = 200 * (psi in cells)
    GenModes m;
    uint64_t s = seed;
    double sumsq = 0.0;
    double ax[kGenModes], ay[kGenModes], az[kGenModes];
    for (int i = 0; i < kGenModes; i++) {
        int kx, ky, kz;
        do {
            kx = (int)(host_splitmix(s) % 17) - 8;
            ky = (int)(host_splitmix(s) % 17) - 8;
            kz = (int)(host_splitmix(s) % 17) - 8;
        } while (kx == 0 && ky == 0 && kz == 0);
        double k2 = (double)kx * kx + (double)ky * ky + (double)kz * kz;
        double a = 1.0 / k2, inv = 1.0 / sqrt(k2);
        m.kx[i] = kx; m.ky[i] = ky; m.kz[i] = kz;
        m.phase[i] = (float)((double)(host_splitmix(s) >> 11) * (1.0 / 9007199254740992.0) * 6.283185307179586);
        ax[i] = a * kx * inv; ay[i] = a * ky * inv; az[i] = a * kz * inv;
        sumsq += a * a;
    }
    const double cell = (double)boxWidth / ngrid;
    const double scale = 1.5 * cell / sqrt(sumsq / 2.0);
    for (int i = 0; i < kGenModes; i++) {
        m.ax[i] = (float)(ax[i] * scale); m.ay[i] = (float)(ay[i] * scale); m.az[i] = (float)(az[i] * scale);
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUDA_CHECK(cudaMemcpyToSymbolAsync(cModes, &m, sizeof(m), 0, cudaMemcpyHostToDevice, st));
    GenParams p;
    p.ngrid = ngrid;
    p.L = boxWidth;
    p.velPerLen = (float)(200.0 / cell);
    p.firstId = firstId;
    p.count = count;
    p.pos = d_pos;
    p.vel = d_vel;
    p.noise = (float)(noiseCells * cell);
    p.noiseSeed = seed ^ 0x6d6e77526f756768ull;
    k_zeldovich<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(p);
    ctx->launches++;
    CUDA_CHECK(cudaGetLastError());
    return 0;
}

#include "minnow_noalloc.inl"

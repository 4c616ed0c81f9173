/* segment.c -- the segment-level flow of /root/reference/src/funcs.c for
 * Posn / Velc fields, on top of the registry (register.c) and libminnow_cuda.
 * Host-side C of libminnow_host.so; no arithmetic on particle data happens
 * here (the checksum of the compressed bytes is the caller's job in the
 * reference too: src/types.h:135, src/funcs.c:72).
 */
#include "minnow_register.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "minnow_cuda.h"

#define HOST_PANIC(...)                                                      \
    do {                                                                     \
        fprintf(stderr, "Panic at file %s, line %d:\n", __FILE__, __LINE__); \
        fprintf(stderr, __VA_ARGS__);                                        \
        fprintf(stderr, "\n");                                               \
        exit(1);                                                             \
    } while (0)

static uint64_t g_ditherSeed = 0;

void MnwSeg_SetDitherSeed(uint64_t seed) { g_ditherSeed = seed; }

/* src/util.c:438-445 */
uint32_t MnwChecksum(const uint8_t *data, int64_t len) {
    uint32_t sum = 1;
    for (int64_t i = 0; i < len; i++) {
        sum = (sum >> 1) | (sum << 31);
        sum += data[i];
    }
    return sum;
}

/* CField.Checksum of every field of a segment (src/funcs.c:72): the recurrence is
 * sequential within a field, so one host thread per field, at most 16 at a time */
typedef struct SumJob { CField *cf; } SumJob;
static void *sum_worker(void *arg) {
    CField *cf = ((SumJob *)arg)->cf;
    cf->Checksum = MnwChecksum(cf->Data, cf->DataLen);
    return NULL;
}
static void checksum_fields(CField *fields, int32_t n) {
    enum { kWave = 16 };
    pthread_t th[kWave];
    SumJob jobs[kWave];
    for (int32_t base = 0; base < n; base += kWave) {
        int32_t cnt = n - base < kWave ? n - base : kWave;
        for (int32_t i = 0; i < cnt; i++) {
            jobs[i].cf = &fields[base + i];
            if (cnt == 1 || pthread_create(&th[i], NULL, sum_worker, &jobs[i]) != 0) {
                sum_worker(&jobs[i]);
                jobs[i].cf = NULL;
            }
        }
        for (int32_t i = 0; i < cnt; i++)
            if (jobs[i].cf) pthread_join(th[i], NULL);
    }
}

static void *xcalloc(size_t n, size_t size) {
    void *p = calloc(n ? n : 1, size);
    if (!p) HOST_PANIC("Allocation failed."); /* AssertAlloc, src/debug.h:36-43 */
    return p;
}

static void require_supported(Register reg, uint32_t algo, uint32_t version) {
    if (!Register_Supports(reg, algo, version)) /* src/funcs.c:151-154 */
        HOST_PANIC("v%u.%u of algorithm %x is not supported.", (version >> 16) & 0xff, (version >> 8) & 0xff, algo);
    if (algo != MNWCUDA_ALGO_CODE)
        HOST_PANIC("Algorithm %x is registered, but only algo_Diff has a fused path here; "
                   "use its CFunc/DFunc with quant_QField/quant_Field.", algo);
}

CSeg MnwSeg_Encode(Seg s) {
    Register reg = Register_New();
    CSeg cs;
    cs.FieldLen = s.FieldLen;
    cs.Fields = xcalloc((size_t)s.FieldLen, sizeof(*cs.Fields));
    for (int32_t i = 0; i < s.FieldLen; i++) {
        const Field *f = &s.Fields[i];
        require_supported(reg, f->Hd.AlgoCode, f->Hd.AlgoVersion);
        Compressor comp = Register_GetCompressor(reg, f->Hd.AlgoCode, f->Hd.AlgoVersion);
        mnwcuda_SetDitherSeed(comp.Buffer, g_ditherSeed);
        cs.Fields[i] = mnwcuda_EncodeField(*f, comp.Buffer);
        Register_FreeCompressor(reg, f->Hd.AlgoCode, f->Hd.AlgoVersion, comp);
    }
    checksum_fields(cs.Fields, cs.FieldLen);
    Register_Free(reg);
    return cs;
}

Seg MnwSeg_Decode(CSeg cs) {
    Register reg = Register_New();
    Seg s;
    s.FieldLen = cs.FieldLen;
    s.Fields = xcalloc((size_t)cs.FieldLen, sizeof(*s.Fields));
    CField *sums = xcalloc((size_t)cs.FieldLen, sizeof(*sums));
    memcpy(sums, cs.Fields, (size_t)cs.FieldLen * sizeof(*sums));
    checksum_fields(sums, cs.FieldLen);
    for (int32_t i = 0; i < cs.FieldLen; i++) {
        const CField *cf = &cs.Fields[i];
        require_supported(reg, cf->Hd.AlgoCode, cf->Hd.AlgoVersion);
        /* src/funcs.c:49-56: a damaged field is skipped, not an error */
        if (sums[i].Checksum != cf->Checksum) {
            s.Fields[i].Hd = cf->Hd;
            s.Fields[i].Valid = 0;
            continue;
        }
        Decompressor dec = Register_GetDecompressor(reg, cf->Hd.AlgoCode, cf->Hd.AlgoVersion);
        s.Fields[i] = mnwcuda_DecodeField(*cf, dec.Buffer);
        s.Fields[i].Valid = 1;
        Register_FreeDecompressor(reg, cf->Hd.AlgoCode, cf->Hd.AlgoVersion, dec);
    }
    free(sums);
    Register_Free(reg);
    return s;
}

/* ToBytes / FromBytes (src/funcs.c:78-120; not functional upstream: stream_Write
 * drops its result, src/stream.c:54-87) as the segment image of
 * doc/header_format/header_format.tex:129-204, block checksums on the device */
MnwBytes MnwSeg_ToBytes(CSeg cs, int32_t particleNum) {
    MnwBytes out = {NULL, 0};
    int64_t cap = mnwcuda_SegmentImageBound(cs.Fields, cs.FieldLen);
    if (cap < 0) HOST_PANIC("MnwSeg_ToBytes: unsupported field (%lld).", (long long)cap);
    out.Data = xcalloc((size_t)cap, 1);
    MnwCudaCtx *ctx = mnwcuda_Acquire(-1);
    out.Len = mnwcuda_SegmentToImage(ctx, cs.Fields, cs.FieldLen, particleNum, out.Data, cap);
    mnwcuda_Release(ctx);
    if (out.Len < 0) HOST_PANIC("MnwSeg_ToBytes: malformed field payload (%lld).", (long long)out.Len);
    return out;
}

/* valid (may be NULL): [FieldNum] 0 for a field whose header block is damaged.
 * A segment whose own header checksum fails comes back empty (FieldLen = -1). */
CSeg MnwSeg_FromBytes(MnwBytes bytes, int32_t *valid, int32_t *particleNum) {
    CSeg cs = {NULL, -1};
    if (bytes.Len < 16) return cs;
    int32_t nFields;
    memcpy(&nFields, bytes.Data + 8, 4);
    if (nFields < 0 || nFields > (1 << 20)) return cs;
    cs.Fields = xcalloc((size_t)nFields, sizeof(*cs.Fields));
    int32_t *v = valid ? valid : xcalloc((size_t)nFields, sizeof(*v));
    MnwCudaCtx *ctx = mnwcuda_Acquire(-1);
    cs.FieldLen = mnwcuda_SegmentFromImage(ctx, bytes.Data, bytes.Len, cs.Fields, nFields, v, particleNum);
    mnwcuda_Release(ctx);
    if (!valid) free(v);
    if (cs.FieldLen < 0) {
        free(cs.Fields);
        cs.Fields = NULL;
        cs.FieldLen = -1;
    }
    return cs;
}

void MnwCSeg_Free(CSeg cs) { /* CSeg_Free, src/funcs.c:205-210 */
    for (int32_t i = 0; i < cs.FieldLen; i++) free(cs.Fields[i].Data);
    free(cs.Fields);
}

void MnwSeg_Free(Seg s) { /* quant_FreeField, src/quant.c:100-113: Acc and Data are ours */
    for (int32_t i = 0; i < s.FieldLen; i++) {
        free(s.Fields[i].Acc);
        free(s.Fields[i].Data);
    }
    free(s.Fields);
}

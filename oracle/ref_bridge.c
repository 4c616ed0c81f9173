/* ref_bridge.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Flat-pointer wrappers over the reference's own functions, compiled (never
 * copied) from /root/reference by oracle/Makefile into oracle/_ref/. They let
 * Python tests and the oracle composition (-DORC_USE_REFERENCE) call the real
 * util.c / rand.c / quant.c arithmetic. Sequences are wrapped, not copied,
 * where the reference allows it (Seq_WrapArray, src/base_seq.h).
 */
#include <stdlib.h>
#include <string.h>

#include "util.h"  /* /root/reference/src/util.h */
#include "rand.h"
#include "seq.h"
#include "quant.h"

/* non-static in src/quant.c:675, not declared in quant.h */
void deltaToDepth(float delta, float *deltas, float x0, float x1,
                  uint8_t *depthPtr, uint8_t **depthsPtr, int32_t len);
/* non-static in src/rand.c:96 */
void xorshiftJump(rand_State *state);

void ref_undo_periodic(float *x, int64_t n, float L) {
    util_UndoPeriodic(FSeq_WrapArray(x, (int32_t)n), L);
}

void ref_periodic(float *x, int64_t n, float L) {
    util_Periodic(FSeq_WrapArray(x, (int32_t)n), L);
}

void ref_minmax(const float *x, int64_t n, float *mn, float *mx) {
    util_MinMax(FSeq_WrapArray((float *)x, (int32_t)n), mn, mx);
}

/* the reference Panics (exit 1) above depth 24; probe the condition first so
 * that a test can observe it (src/quant.c:684-693) */
int ref_delta_to_depth(float delta, float x0, float x1) {
    if (!(delta * (float)(1 << 24) > x1 - x0)) return -1;
    uint8_t depth = 0, *depths = NULL;
    deltaToDepth(delta, NULL, x0, x1, &depth, &depths, 0);
    return depth;
}

void ref_uniform_bin_index(const float *x, int64_t n, uint8_t level,
                           float x0, float dx, uint32_t *q) {
    if (n <= 0) return;
    U64Seq buf = U64Seq_New((int32_t)n);
    buf = util_UniformBinIndex(FSeq_WrapArray((float *)x, (int32_t)n), level,
                               x0, dx, buf);
    for (int64_t i = 0; i < n; i++) q[i] = (uint32_t)buf.Data[i];
    U64Seq_Free(buf);
}

void ref_undo_uniform_bin_index(const uint32_t *q, int64_t n, uint8_t level,
                                float x0, float dx, uint64_t state[2],
                                float *out) {
    if (n <= 0) return;
    U64Seq idx = U64Seq_New((int32_t)n);
    for (int64_t i = 0; i < n; i++) idx.Data[i] = q[i];
    FSeq buf = FSeq_New((int32_t)n);
    buf = util_UndoUniformBinIndex(idx, level, x0, dx, (rand_State *)state, buf);
    memcpy(out, buf.Data, (size_t)n * sizeof(float));
    FSeq_Free(buf);
    U64Seq_Free(idx);
}

void ref_uniform_pack(const uint32_t *x, int64_t n, uint8_t width,
                      uint32_t *out) {
    U32Seq in = U32Seq_WrapArray((uint32_t *)x, (int32_t)n);
    U32Seq buf = util_U32UniformPack(in, width, U32Seq_Empty());
    if (buf.Len > 0) memcpy(out, buf.Data, (size_t)buf.Len * 4);
    U32Seq_Free(buf);
}

void ref_undo_uniform_pack(const uint32_t *in, uint8_t width, int64_t n,
                           uint32_t *out) {
    if (n <= 0) return;
    uint64_t bits = (uint64_t)width * (uint64_t)n;
    int32_t words = (int32_t)((bits + 31) / 32);
    U32Seq src = U32Seq_WrapArray((uint32_t *)in, words);
    U32Seq buf = util_U32UndoUniformPack(src, width, (int32_t)n, U32Seq_Empty());
    memcpy(out, buf.Data, (size_t)n * 4);
    U32Seq_Free(buf);
}

uint32_t ref_checksum(const uint8_t *bytes, int64_t n) {
    return util_Checksum(U8BigSeq_WrapArray((uint8_t *)bytes, n));
}

void ref_rand_seed(uint64_t seed, int64_t n, uint64_t *states) {
    if (n <= 0) return;
    rand_State *s = rand_Seed(seed, (int32_t)n);
    memcpy(states, s, (size_t)n * sizeof(*s));
    free(s);
}

uint64_t ref_rand_u64(uint64_t state[2]) {
    return rand_Uint64((rand_State *)state);
}

float ref_rand_float(uint64_t state[2]) {
    return rand_Float((rand_State *)state);
}

void ref_rand_jump(uint64_t state[2]) { xorshiftJump((rand_State *)state); }

/* id() / undoID() / uint() / undoUint() of src/quant.c:291-327, 373-398,
 * 553-608 through the reference's own dispatch (quant_QField / quant_Field) */
int ref_quant_id(const uint64_t *ids, int32_t n, uint64_t width,
                 uint64_t X0[3], uint64_t X1[3], uint64_t *q) {
    if (n <= 0 || width == 0) return -1;
    /* id() never attaches its IDQuantization to the QField it returns
     * (src/quant.c:291-327: `quant` is filled in and dropped), so X0 / X1 are
     * taken from the same reference primitives id() calls, on the same split
     * (src/quant.c:308-320), and the values are cross-checked against id()'s own */
    for (int32_t i = 0; i < n; i++) {
        q[i] = ids[i] % width;
        q[(size_t)n + i] = (ids[i] / width) % width;
        q[2 * (size_t)n + i] = ids[i] / (width * width);
    }
    for (int c = 0; c < 3; c++) {
        U64Seq comp = U64Seq_WrapArray(q + (size_t)c * n, n);
        util_U64UndoPeriodic(comp, width);
        util_U64MinMax(comp, &X0[c], &X1[c]);
        for (int32_t i = 0; i < n; i++) comp.Data[i] -= X0[c];
    }
    IDAccuracy acc = { .Width = width };
    Field f;
    memset(&f, 0, sizeof(f));
    f.Hd.FieldCode = field_Ptid;
    f.Hd.ParticleLen = n;
    f.Data = (void *)ids;
    f.Acc = &acc;
    QField qf = quant_QField(f);
    int same = memcmp(q, qf.Data, (size_t)n * 3 * sizeof(uint64_t)) == 0;
    free(qf.Data);
    return same ? 0 : -9;
}

int ref_undo_id(const uint64_t *q, int32_t n, uint64_t width,
                const uint64_t X0[3], uint64_t *ids) {
    if (n <= 0) return -1;
    IDQuantization quant;
    memset(&quant, 0, sizeof(quant));
    quant.Width = width;
    for (int c = 0; c < 3; c++) quant.X0[c] = X0[c];
    QField qf;
    memset(&qf, 0, sizeof(qf));
    qf.Hd.FieldCode = field_Ptid;
    qf.Hd.ParticleLen = n;
    qf.Data = (uint64_t *)q;
    qf.Quant = &quant;
    Field f = quant_Field(qf);
    memcpy(ids, f.Data, (size_t)n * sizeof(uint64_t));
    free(f.Acc);
    free(f.Data);
    return 0;
}

int ref_quant_uint(const uint64_t *data, int32_t n, uint64_t *x0, uint64_t *x1,
                   uint64_t *q) {
    if (n <= 0) return -1;
    /* uint() drops its IntQuantization the same way (src/quant.c:373-398) */
    util_U64MinMax(U64Seq_WrapArray((uint64_t *)data, n), x0, x1);
    uint8_t acc = 0;
    Field f;
    memset(&f, 0, sizeof(f));
    f.Hd.FieldCode = field_Unsi;
    f.Hd.ParticleLen = n;
    f.Data = (void *)data;
    f.Acc = &acc;
    QField qf = quant_QField(f);
    memcpy(q, qf.Data, (size_t)n * sizeof(uint64_t));
    free(qf.Data);
    for (int32_t i = 0; i < n; i++)
        if (q[i] != data[i] - *x0) return -9;
    return 0;
}

int ref_undo_uint(const uint64_t *q, int32_t n, uint64_t x0, uint64_t *out) {
    if (n <= 0) return -1;
    IntQuantization quant = { .X0 = x0, .X1 = 0 };
    QField qf;
    memset(&qf, 0, sizeof(qf));
    qf.Hd.FieldCode = field_Unsi;
    qf.Hd.ParticleLen = n;
    qf.Data = (uint64_t *)q;
    qf.Quant = &quant;
    Field f = quant_Field(qf);
    memcpy(out, f.Data, (size_t)n * sizeof(uint64_t));
    free(f.Data);
    return 0;
}

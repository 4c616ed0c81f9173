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

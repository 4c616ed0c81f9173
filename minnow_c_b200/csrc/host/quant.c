/* quant.c -- quant_QField / quant_Field / quant_FreeQField / quant_FreeField
 * with the reference's names and signatures (/root/reference/src/quant.h:6-9),
 * as thin callers of libminnow_cuda's quantiser-alone entry points. They are
 * what src/funcs.c:13-38 (Quantize / UndoQuantize) and :191-203 (Seg_Free /
 * QSeg_Free) call, with the reference's defects fixed: 3N
 * outputs, the Quantization attached to the QField, dx = maxDiff, a
 * reproducible dither seed. Host-side C of libminnow_host.so; all arithmetic on
 * particle data happens on the device.
 */
#include "minnow_register.h"

#include <stdio.h>
#include <stdlib.h>

#include "minnow_cuda.h"

#define HOST_PANIC(...)                                                      \
    do {                                                                     \
        fprintf(stderr, "Panic at file %s, line %d:\n", __FILE__, __LINE__); \
        fprintf(stderr, __VA_ARGS__);                                        \
        fprintf(stderr, "\n");                                               \
        exit(1);                                                             \
    } while (0)

static uint64_t g_quantSeed = 0;

/* dither seed quant_Field uses for a QField that did not come out of the
 * algo_Diff DFunc (those carry their payload's seed); default 0 */
void MnwQuant_SetDitherSeed(uint64_t seed) { g_quantSeed = seed; }

QField quant_QField(Field f) { /* src/quant.c:135-144 */
    MnwCudaCtx *ctx = mnwcuda_Acquire(-1);
    QField qf = mnwcuda_QuantizeField(f, ctx);
    mnwcuda_Release(ctx);
    return qf;
}

Field quant_Field(QField qf) { /* src/quant.c:146-155 */
    MnwCudaCtx *ctx = mnwcuda_Acquire(-1);
    mnwcuda_SetDitherSeed(ctx, g_quantSeed);
    Field f = mnwcuda_DequantizeField(qf, ctx);
    mnwcuda_Release(ctx);
    return f;
}

void quant_FreeQField(QField qf) { /* src/quant.c:65-98; Depths is always NULL here */
    switch (qf.Hd.FieldCode) {
    case field_Posn:
    case field_Velc:
    case field_Ptid:
    case field_Unsi:
        mnwcuda_ForgetQuant(qf.Quant);
        free(qf.Quant);
        free(qf.Data);
        return;
    default:
        HOST_PANIC("Unrecognized field code %x.", qf.Hd.FieldCode);
    }
}

void quant_FreeField(Field f) { /* src/quant.c:100-133; Deltas is always NULL here */
    switch (f.Hd.FieldCode) {
    case field_Posn:
    case field_Velc:
    case field_Ptid:
    case field_Unsi:
        free(f.Acc);
        free(f.Data);
        return;
    default:
        HOST_PANIC("Unrecognized field code %x.", f.Hd.FieldCode);
    }
}

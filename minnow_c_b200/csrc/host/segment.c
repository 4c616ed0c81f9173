/* segment.c -- the segment-level flow of /root/reference/src/funcs.c for
 * Posn / Velc fields, on top of the registry (register.c) and libminnow_cuda.
 * Host-side C of libminnow_host.so; no arithmetic on particle data happens
 * here (the checksum of the compressed bytes is the caller's job in the
 * reference too: src/types.h:135, src/funcs.c:72).
 */
#include "minnow_register.h"

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

static void *xcalloc(size_t n, size_t size) {
    void *p = calloc(n ? n : 1, size);
    if (!p) HOST_PANIC("Allocation failed."); /* AssertAlloc, src/debug.h:36-43 */
    return p;
}

static void require_supported(Register reg, uint32_t algo, uint32_t version) {
    if (!Register_Supports(reg, algo, version)) /* src/funcs.c:151-154 */
        HOST_PANIC("v%u.%u of algorithm %x is not supported.", (version >> 16) & 0xff, (version >> 8) & 0xff, algo);
    if (algo != MNWCUDA_ALGO_CODE)
        HOST_PANIC("Algorithm %x is registered, but only algo_Diff has a fused float path here; "
                   "use its CFunc/DFunc with the reference's quant_QField/quant_Field.", algo);
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
        CField cf = mnwcuda_EncodeField(*f, comp.Buffer);
        cf.Checksum = MnwChecksum(cf.Data, cf.DataLen);
        cs.Fields[i] = cf;
        Register_FreeCompressor(reg, f->Hd.AlgoCode, f->Hd.AlgoVersion, comp);
    }
    Register_Free(reg);
    return cs;
}

Seg MnwSeg_Decode(CSeg cs) {
    Register reg = Register_New();
    Seg s;
    s.FieldLen = cs.FieldLen;
    s.Fields = xcalloc((size_t)cs.FieldLen, sizeof(*s.Fields));
    for (int32_t i = 0; i < cs.FieldLen; i++) {
        const CField *cf = &cs.Fields[i];
        require_supported(reg, cf->Hd.AlgoCode, cf->Hd.AlgoVersion);
        /* src/funcs.c:49-56: a damaged field is skipped, not an error */
        if (MnwChecksum(cf->Data, cf->DataLen) != cf->Checksum) {
            s.Fields[i].Hd = cf->Hd;
            s.Fields[i].Valid = 0;
            continue;
        }
        Decompressor dec = Register_GetDecompressor(reg, cf->Hd.AlgoCode, cf->Hd.AlgoVersion);
        s.Fields[i] = mnwcuda_DecodeField(*cf, dec.Buffer);
        s.Fields[i].Valid = 1;
        Register_FreeDecompressor(reg, cf->Hd.AlgoCode, cf->Hd.AlgoVersion, dec);
    }
    Register_Free(reg);
    return s;
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

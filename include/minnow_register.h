/* minnow_register.h -- the algorithm registry and the segment-level calls of
 * libminnow_host.so: the host-side C that sits between a minnow client and
 * libminnow_cuda.so.
 *
 * The reference's registry is an empty stub (src/register.c:21-84: Register_Add
 * drops its arguments, Register_Supports is constant false, the getters return
 * uninitialised structs) and every Load / Free function of src/funcs.c:122-189
 * builds a fresh Register_New(); so "registering" an algorithm can only mean
 * that Register_New() itself knows it. This library is that registry: same
 * types and signatures as src/register.h:9-49, a built-in table in which
 * algo_Diff v0.9 maps to the CUDA plugin functions of minnow_cuda.h.
 *
 * Inside the minnow tree include the reference's own "register.h" first (same
 * include guard): this header then declares only the Mnw* additions.
 */
#ifndef MINNOW_REGISTER_H_
#define MINNOW_REGISTER_H_

#include <stdbool.h>
#include <stdint.h>

#include "minnow_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#ifndef MNW_REGISTER_H_ /* the reference's register.h guard */
#define MNW_REGISTER_H_

/* layout of the reference's generated sequences (src/base_seq.h:124-152):
 * { T *Data; int32_t Len, Cap; } */
typedef struct U32Seq { uint32_t *Data; int32_t Len, Cap; } U32Seq;
typedef struct U32SeqSeq { U32Seq *Data; int32_t Len, Cap; } U32SeqSeq;
typedef struct PtrSeq { void **Data; int32_t Len, Cap; } PtrSeq;
typedef struct PtrSeqSeq { PtrSeq *Data; int32_t Len, Cap; } PtrSeqSeq;

typedef struct Register { /* src/register.h:9-13 */
    U32Seq Algo;          /* algorithm codes */
    U32SeqSeq Version;    /* per algorithm: supported version words */
    PtrSeqSeq Funcs;      /* per algorithm, per version: RegisterFuncs* */
} Register;

typedef struct RegisterFuncs { /* src/register.h:15-22 */
    CFunc CFunc;
    void *(*CAlloc)(void);
    void (*CFree)(void *);
    DFunc DFunc;
    void *(*DAlloc)(void);
    void (*DFree)(void *);
} RegisterFuncs;

/* src/register.h:24-49, with working bodies */
Register Register_New(void);
void Register_Free(Register reg);
void Register_Add(Register reg, uint32_t algo, uint32_t version, RegisterFuncs funcs);
Compressor Register_GetCompressor(Register reg, uint32_t algo, uint32_t version);
Decompressor Register_GetDecompressor(Register reg, uint32_t algo, uint32_t version);
void Register_FreeDecompressor(Register reg, uint32_t algo, uint32_t version, Decompressor decomp);
void Register_FreeCompressor(Register reg, uint32_t algo, uint32_t version, Compressor comp);
bool Register_Supports(Register reg, uint32_t algo, uint32_t version);
uint32_t Register_Newest(Register reg, uint32_t algo);

#endif /* MNW_REGISTER_H_ */

/* ---- segment level: the flow of src/funcs.c:268-301 for Posn / Velc fields ----
 * MnwSeg_Encode == Quantize + LoadCompressors + Compress + FreeCompressors
 * (src/funcs.c:13-23, 143-162, 62-76): every field goes through the registry
 * by (AlgoCode, AlgoVersion); algo_Diff fields take the fused float path
 * (mnwcuda_EncodeField); the caller's checksum of src/funcs.c:72 is applied.
 * MnwSeg_Decode == LoadDecompressors + Decompress + UndoQuantize
 * (src/funcs.c:122-141, 40-60, 25-38): a field whose checksum does not match
 * comes back zeroed with Valid = 0, as in the reference. Unsupported
 * (algo, version) pairs Panic like src/funcs.c:151-154.
 * Returned segments are released with MnwCSeg_Free / MnwSeg_Free. */
CSeg MnwSeg_Encode(Seg s);
Seg MnwSeg_Decode(CSeg cs);
void MnwCSeg_Free(CSeg cs);
void MnwSeg_Free(Seg s);
/* ToBytes / FromBytes (src/funcs.c:78-120) as the segment image of
 * doc/header_format/header_format.tex:129-204 (mnwcuda_SegmentToImage). A block
 * whose checksum fails costs only its own particles (NaN / all ones after
 * decoding); valid[f] = 0 (valid may be NULL) for a field whose header block is
 * damaged; FieldLen = -1 when the segment header itself is. Data is malloc'd. */
typedef struct MnwBytes { uint8_t *Data; int64_t Len; } MnwBytes;
MnwBytes MnwSeg_ToBytes(CSeg cs, int32_t particleNum);
CSeg MnwSeg_FromBytes(MnwBytes bytes, int32_t *valid, int32_t *particleNum);
/* the reference's quantiser interface (src/quant.h:6-9) over the device kernels:
 * what src/funcs.c's Quantize / UndoQuantize / Seg_Free / QSeg_Free call */
QField quant_QField(Field f);
Field quant_Field(QField qf);
void quant_FreeQField(QField qf);
void quant_FreeField(Field f);
void MnwQuant_SetDitherSeed(uint64_t seed);
/* util_Checksum (src/util.c:438-445): rotate right by one, add byte, start 1 */
uint32_t MnwChecksum(const uint8_t *data, int64_t len);
/* dither seed of the fields encoded by MnwSeg_Encode from now on (default 0) */
void MnwSeg_SetDitherSeed(uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif /* MINNOW_REGISTER_H_ */

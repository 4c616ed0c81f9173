/* minnow_cuda.h -- C ABI of libminnow_cuda.so: the B200 (sm_100a)
 * implementation of minnow's encode/decode hot path, registered as algorithm
 * `algo_Diff` v0.9 behind the reference's plugin interface.
 *
 * Plain C: pointers, sizes, the reference's own by-value structs. No torch,
 * no C++ types. Every entry point cites the reference interface it replaces
 * (paths under /root/reference). There is no CPU fallback: every entry point
 * needs a CUDA device and aborts (Panic convention, src/debug.h:72-89) on a
 * CUDA failure.
 */
#ifndef MINNOW_CUDA_H_
#define MINNOW_CUDA_H_

#include <stdint.h>
#include "minnow_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define MNWCUDA_ALGO_CODE    algo_Diff
#define MNWCUDA_ALGO_VERSION 0x00000900u /* 0.9.0-dev, src/semver.c:93-94 */

#define MNWCUDA_BLOCK_LEN    1024 /* values per pack block */
#define MNWCUDA_DITHER_RUN   32   /* consecutive blocks per dither task */
#define MNWCUDA_DITHER_LANES 32   /* rand_Seed streams per dither task */

/* ---- wire structs of the Diff v0.9 payload (CField.Data) ---- */
typedef struct MnwDiffHeader { /* 64 bytes, little-endian; the pointer-free
                                  image of Position/VelocityQuantization
                                  (src/types.h:82-95) */
    uint32_t Magic;       /* algo_Diff */
    uint32_t Version;     /* MNWCUDA_ALGO_VERSION */
    uint32_t FieldCode;
    int32_t  ParticleLen;
    uint8_t  Depth;
    uint8_t  Pad[3];
    float    Width;       /* Posn: box width; Velc: SymLog10Threshold (0) */
    float    X0[3];
    float    X1[3];
    uint32_t BlockLen;
    uint16_t DitherRun;
    uint16_t DitherLanes;
    uint64_t DitherSeed;
} MnwDiffHeader;

typedef struct MnwDiffBlock { /* 16 bytes per block, table follows the header */
    uint32_t First;
    int32_t  MinDelta;
    uint32_t Width;
    uint32_t Offset16;
} MnwDiffBlock;

#define MNWCUDA_WIDTH_DAMAGED 0xffffffffu /* MnwDiffBlock.Width of a block whose
                                             checksum failed (segment image) */

/* Integer fields (field_Ptid, field_Unsi) use the same 64-byte header: Width and
 * X0[] / X1[] hold uint32 bit patterns -- Ptid: IDQuantization.Width, X0[3],
 * X1[3] (src/types.h:97-99); Unsi: X0[0..1] / X1[0..1] = low, high word of
 * IntQuantization.X0 / X1 (src/types.h:77-79) -- Depth = bit length of the
 * widest range (at most 24), DitherSeed = 0. An Unsi field of N values is coded
 * as three runs of ceil(N/3) (the tail repeats the last value): the payload
 * header's ParticleLen is the run length, CField.Hd.ParticleLen stays N. */

/* BlockHeader of a segment image (doc/header_format/header_format.tex:182-185) */
typedef struct MnwBlockHeader {
    int32_t  Length;
    uint32_t Checksum; /* util_Checksum, src/util.c:438-445 (start value 1) */
} MnwBlockHeader;

/* ---- one field of a batch (host array) ---- */
typedef struct MnwFieldDesc {
    int64_t  Offset;     /* encode: first float of the field's 3N planar input
                            inside d_in; decode: same inside d_out */
    int32_t  ParticleLen;
    uint32_t FieldCode;  /* field_Posn or field_Velc (src/types.h:8-9); decode and
                            the block-header calls also take field_Ptid /
                            field_Unsi (ParticleLen = the run length for Unsi) */
    float    Delta;      /* Position/VelocityAccuracy.Delta (types.h:42-55) */
    float    Width;      /* PositionAccuracy.Width; ignored for Velc */
    uint64_t DitherSeed; /* recorded in the header; the decoder's rand_Seed seed
                            (replaces clock() of src/quant.c:639) */
} MnwFieldDesc;

typedef struct MnwCudaCtx MnwCudaCtx;

/* ---- sizes (host arithmetic only) ---- */
int64_t mnwcuda_FieldFixedBytes(int32_t particleLen); /* header + block table */
int64_t mnwcuda_FieldBound(int32_t particleLen);      /* worst-case payload */

/* ---- context: device scratch, stream-ordered; one per thread ----
 * The algorithm-private `void *Buffer` of the plugin interface
 * (src/types.h:133, src/register.h:15-22) is one of these. */
MnwCudaCtx *mnwcuda_Create(int device);
void mnwcuda_Destroy(MnwCudaCtx *ctx);

/* ---- device-resident batch entry points (the timed path) ----
 * Replace, for every field of the batch at once,
 *   encode: quant_QField (src/quant.c:135-144 -> position()/velocity()
 *           :161-289: util_UndoPeriodic, util_MinMax, deltaToDepth,
 *           util_UniformBinIndex) followed by CFunc (src/funcs.c:71)
 *   decode: DFunc (src/funcs.c:54) followed by quant_Field
 *           (src/quant.c:146-155 -> undoPosition()/undoVelocity() :451-551:
 *           util_UndoUniformBinIndex, util_Periodic).
 * Input floats must be finite: the reference's own arithmetic is undefined for
 * NaN / infinity (the float -> uint32 cast of src/util.c:190) and nothing here
 * detects them.
 * d_in/d_out/d_offsets are device pointers; h_fields is a host array.
 * Field f's payload is d_out[d_offsets[f] .. d_offsets[f+1]); d_offsets has
 * nFields+1 entries. `stream` is a cudaStream_t (NULL = default stream).
 * Work is enqueued, not waited for. Returns 0, or <0 for a host-side
 * argument error; device-side errors are reported by mnwcuda_Status. */
int mnwcuda_EncodeFieldsDev(MnwCudaCtx *ctx, const float *d_in,
                            const MnwFieldDesc *h_fields, int64_t nFields,
                            uint8_t *d_out, int64_t outCap,
                            uint64_t *d_offsets, void *stream);
int mnwcuda_DecodeFieldsDev(MnwCudaCtx *ctx, const uint8_t *d_in,
                            const uint64_t *d_offsets,
                            const MnwFieldDesc *h_fields, int64_t nFields,
                            float *d_out, void *stream);
/* Waits for `stream` and returns 0 or the first device-side error of the
 * calls since the last Status: 1 = accuracy needs depth > 24 (the Panic of
 * src/quant.c:688-693), 2 = output capacity exceeded, 3 = malformed payload,
 * 4 = quantised integers at the CFunc boundary exceed 24 bits,
 * 5 = a device-side wait inside k_encode timed out (an internal error: the
 *     kernel ends cleanly, the output is invalid, and a description of the
 *     first missing dependency goes to stderr). */
int mnwcuda_Status(MnwCudaCtx *ctx, void *stream);
/* Kernels launched by this context since creation (bench bookkeeping). */
int64_t mnwcuda_LaunchCount(const MnwCudaCtx *ctx);
/* Per-kernel device times of the most recent Encode/DecodeFieldsDev call
 * (bench bookkeeping; CUDA events on the call's stream, off by default).
 * ms[0] = k_encode (min/max + quantise + delta + pack, fused), ms[1] =
 * k_finalize + k_tables, ms[2] = k_block_headers (mnwcuda_BlockHeadersDev),
 * ms[3] unused; ms[4..6] = k_parse_headers, k_stream_states,
 * k_unpack_dequant; -1 = not run. */
void mnwcuda_EnableKernelTiming(MnwCudaCtx *ctx, int on);
int mnwcuda_KernelTimes(MnwCudaCtx *ctx, float *ms /* [7] */);

/* ---- host-buffer entry points (reference-facing, synchronous) ----
 * mnwcuda_EncodeField == CFunc(quant_QField(f)); the returned CField.Data is
 * malloc'd and released by the caller with free() (CSeg_Free,
 * src/funcs.c:205-210); Checksum is left 0 for the caller (src/types.h:135).
 * mnwcuda_DecodeField == quant_Field(DFunc(cf)); Data and Acc are malloc'd
 * (quant_FreeField, src/quant.c:100-113). `buffer` is a MnwCudaCtx*. */
CField mnwcuda_EncodeField(Field f, void *buffer);
Field mnwcuda_DecodeField(CField cf, void *buffer);
/* Many fields per call with host buffers: H2D, kernels and D2H are pipelined
 * over chunks of fields. h_in holds the fields' floats at h_fields[i].Offset;
 * payloads are written back to back into h_out, h_offsets[nFields+1]. Returns
 * total bytes or <0. */
int64_t mnwcuda_EncodeFieldsHost(MnwCudaCtx *ctx, const float *h_in,
                                 const MnwFieldDesc *h_fields, int64_t nFields,
                                 uint8_t *h_out, int64_t outCap,
                                 uint64_t *h_offsets);
int mnwcuda_DecodeFieldsHost(MnwCudaCtx *ctx, const uint8_t *h_in,
                             const uint64_t *h_offsets,
                             const MnwFieldDesc *h_fields, int64_t nFields,
                             float *h_out);

/* ---- the six plugin functions (src/register.h:15-22; shape of
 * src/algo_Test_v0_9.h:6-13). CFunc takes quantized uint64 values and a
 * Position/VelocityQuantization in qf.Quant; DFunc returns them. ---- */
CField MnwCudaCompress_v0_9(QField qf, void *buffer);
QField MnwCudaDecompress_v0_9(CField cf, void *buffer);
void *MnwCudaCAlloc_v0_9(void);
void MnwCudaCFree_v0_9(void *buffer);
void *MnwCudaDAlloc_v0_9(void);
void MnwCudaDFree_v0_9(void *buffer);
/* Dither seed used by EncodeField / Compress on this buffer (default 0). */
void mnwcuda_SetDitherSeed(void *buffer, uint64_t seed);
/* CAlloc / DAlloc hand out pooled contexts (a context costs device allocations,
 * streams and a jump table: the reference makes one Compressor per field,
 * src/funcs.c:145-158). Acquire / Release are the same pool for other callers. */
MnwCudaCtx *mnwcuda_Acquire(int device);
void mnwcuda_Release(MnwCudaCtx *ctx);

/* ---- the quantiser alone: quant_QField / quant_Field (src/quant.h:6-7) as the
 * reference intends them (src/quant.c:161-327, 373-398, 451-608), with the fixes
 * of DESIGN.md section 2: 3N outputs, the Quantization attached to the QField
 * (the reference forgets to), dx = maxDiff, explicit dither seed. Posn, Velc,
 * Ptid, Unsi. Data / Quant / Acc of the results are malloc'd (quant_FreeQField /
 * quant_FreeField, src/quant.c:65-113). mnwcuda_DequantizeField dithers with the
 * seed of the payload the QField came from when it was made by
 * MnwCudaDecompress_v0_9 (remembered per Quant pointer until then), else with
 * the buffer's seed: quant_Field(DFunc(cf)) == mnwcuda_DecodeField(cf). ---- */
QField mnwcuda_QuantizeField(Field f, void *buffer);
Field mnwcuda_DequantizeField(QField qf, void *buffer);
void mnwcuda_ForgetQuant(const void *quant); /* a QField freed without being dequantised */

/* ---- checksums and the segment image (doc/header_format/header_format.tex:
 * 129-204): SegmentHeader | FieldHeader[F] | BlockHeader[B] | blocks. Blocks of
 * an algo_Diff field: block 0 = its 64-byte header + block table, then one per
 * packed block in table order (all multiples of 16 bytes). FieldHeader is the
 * code's struct (src/types.h:103-108, 4th word ParticleLen). util_Checksum runs
 * on the device, one thread per block. ---- */
int64_t mnwcuda_BlockCount(const MnwFieldDesc *h_fields, int64_t nFields);
/* d_headers[mnwcuda_BlockCount] <- {Length, Checksum} of every block of a packed
 * stream on the device (the output of mnwcuda_EncodeFieldsDev) */
int mnwcuda_BlockHeadersDev(MnwCudaCtx *ctx, const uint8_t *d_stream,
                            const uint64_t *d_offsets, const MnwFieldDesc *h_fields,
                            int64_t nFields, MnwBlockHeader *d_headers, void *stream);
/* The inverse check: a packed block that does not match its header gets Width =
 * MNWCUDA_WIDTH_DAMAGED in the stream's table (the decoders then return NaN /
 * all-ones for its particles only, tex:188-196); d_fieldValid[f] = 0 when a
 * field's block 0 is damaged (else 1). */
int mnwcuda_VerifyBlocksDev(MnwCudaCtx *ctx, uint8_t *d_stream,
                            const uint64_t *d_offsets, const MnwFieldDesc *h_fields,
                            int64_t nFields, const MnwBlockHeader *d_headers,
                            int32_t *d_fieldValid, void *stream);
/* util_Checksum of nRanges byte ranges of a host buffer (copied to the device):
 * h_sums[r] = checksum(h_data + h_off[r], h_len[r]). E.g. CField.Checksum of
 * every field of a segment at once (src/funcs.c:72). */
int mnwcuda_ChecksumsHost(MnwCudaCtx *ctx, const uint8_t *h_data, int64_t dataLen,
                          const int64_t *h_off, const int64_t *h_len, int64_t nRanges,
                          uint32_t *h_sums);
/* CSeg <-> segment image. ToImage returns the bytes written (<= ImageBound) or
 * <0; FromImage fills fields[0..F) (Data malloc'd, Checksum = util_Checksum of
 * the payload as received), valid[f] = 0 for a field whose block 0 is damaged,
 * returns F or <0 (-4: the segment header's own checksum failed). */
int64_t mnwcuda_SegmentImageBound(const CField *fields, int32_t nFields);
int64_t mnwcuda_SegmentToImage(MnwCudaCtx *ctx, const CField *fields, int32_t nFields,
                               int32_t particleNum, uint8_t *h_image, int64_t cap);
int32_t mnwcuda_SegmentFromImage(MnwCudaCtx *ctx, const uint8_t *h_image, int64_t len,
                                 CField *fields, int32_t maxFields, int32_t *valid,
                                 int32_t *particleNum);

/* ---- decode of one float field without any allocation ----
 * No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it). The same kernels as
 * mnwcuda_DecodeFieldsDev with nFields = 1, but every piece of decoder scratch
 * (control words, parsed header, jump-matrix powers, damaged-task list, dither
 * stream states) lives in one file-scope `static __device__` array of
 * MNWCUDA_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit; no context is needed and the call issues kernel launches only (no
 * cudaMalloc / cudaFree / copies; the FIRST call on a device also resolves the
 * array with cudaGetSymbolAddress and uploads 52 KiB of jump-matrix powers with
 * a synchronous cudaMemcpy -- make that call outside any capture).
 * d_payload: the field's algo_Diff payload on the device, 16-byte aligned,
 * payloadLen bytes; d_out: 3 * particleLen floats (planar, as DecodeFieldsDev);
 * d_status: device word that receives 0 or the device-side error (3 = malformed
 * payload), may be NULL; `device` < 0 = the current device.
 * Returns 0, -1 bad argument, -3 field code other than Posn / Velc, -6
 * misaligned pointer, -7 particleLen > MNWCUDA_NOALLOC_MAX_PARTICLES (message on
 * stderr; use mnwcuda_DecodeFieldsDev), -8 CUDA error. Work is enqueued on
 * `stream`, not waited for. All calls on a device share the one array: calls
 * that overlap on the GPU (different streams, or several threads) race -- keep
 * them on one stream or order them with events. */
#define MNWCUDA_NOALLOC_MAX_PARTICLES (1 << 22)
#define MNWCUDA_NOALLOC_SCRATCH_BYTES 253184
int mnwcuda_DecodeFieldNoAlloc(int device, const uint8_t *d_payload,
                               int64_t payloadLen, int32_t particleLen,
                               uint32_t fieldCode, float *d_out,
                               int32_t *d_status, void *stream);

/* ---- synthetic input (bench/tests): Zel'dovich-displaced Lagrangian grid,
 * ids [firstId, firstId+count) of an ngrid^3 box, planar float32.
 * d_pos / d_vel may be NULL. ---- */
int mnwcuda_GenerateZeldovich(MnwCudaCtx *ctx, int32_t ngrid, float boxWidth,
                              uint64_t seed, int64_t firstId, int64_t count,
                              float *d_pos, float *d_vel, void *stream);
/* The same with a white-noise displacement of +-noiseCells grid cells per
 * component on top (a hash of the particle id): a rough field whose deltas do
 * not shrink with the grid spacing, for the wide-block paths. */
int mnwcuda_GenerateZeldovichRough(MnwCudaCtx *ctx, int32_t ngrid, float boxWidth,
                                   uint64_t seed, int64_t firstId, int64_t count,
                                   float noiseCells, float *d_pos, float *d_vel,
                                   void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MINNOW_CUDA_H_ */

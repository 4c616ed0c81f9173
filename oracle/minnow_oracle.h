/* minnow_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the minnow encode/decode hot path (quantize -> delta
 * along particle-ID order -> uniform-width bit-pack per block, and back).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this. The product (libminnow_cuda) never
 * links or calls it.
 *
 * Parity status: every arithmetic primitive is pinned against the reference's
 * golden vectors (test/compress_util_test.c) and against the reference's own
 * util.c / rand.c / quant.c compiled into oracle/_ref/ (see oracle/Makefile).
 * The COMPOSITION (integer delta, per-block headers, block size, dither
 * stream assignment) exists nowhere in the reference (algo_Test_v0_9.c is a
 * stub): it is "parity unpinned" at that level and defined by this file.
 *
 * All citations are file:line under /root/reference.
 */
#ifndef MINNOW_ORACLE_H_
#define MINNOW_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- field / algorithm codes (src/types.h:8-20) ---- */
#define ORC_FIELD_POSN 0x506f736eu
#define ORC_FIELD_VELC 0x56656c63u
#define ORC_FIELD_PTID 0x50746964u
#define ORC_FIELD_UNSI 0x556e7369u
#define ORC_ALGO_DIFF  0x44696666u
#define ORC_VERSION_0_9 0x00000900u /* src/semver.c:93-94 */

/* ---- wire layout of the Diff v0.9 payload (CField.Data) ---- */
#define ORC_BLOCK_LEN     1024u /* values per pack block */
#define ORC_DITHER_LANES  32u   /* dither streams per task */
#define ORC_DITHER_RUN    32u   /* consecutive blocks per dither task */
#define ORC_CHUNK         32u   /* consecutive values dithered by one stream */

typedef struct OrcFieldHeader { /* 64 bytes, little-endian */
    uint32_t Magic;        /* ORC_ALGO_DIFF */
    uint32_t Version;      /* ORC_VERSION_0_9 */
    uint32_t FieldCode;    /* Posn or Velc */
    int32_t  ParticleLen;  /* N */
    uint8_t  Depth;        /* PositionQuantization.Depth, types.h:82-87 */
    uint8_t  Pad[3];
    float    Width;        /* Posn: box width; Velc: SymLog10Threshold (0) */
    float    X0[3];
    float    X1[3];
    uint32_t BlockLen;
    uint16_t DitherRun;
    uint16_t DitherLanes;
    uint64_t DitherSeed;
} OrcFieldHeader;

#define ORC_WIDTH_DAMAGED 0xffffffffu /* table entry of a block whose checksum failed */

typedef struct OrcBlockEntry { /* 16 bytes */
    uint32_t First;    /* q[0] of the block */
    int32_t  MinDelta; /* min_j (q[j]-q[j-1]), j>=1; 0 if len==1 */
    uint32_t Width;    /* bits per residual, 0..25 */
    uint32_t Offset16; /* payload offset from payload base, 16-byte units */
} OrcBlockEntry;

/* ---- primitives (restated) ---- */
void orc_undo_periodic(float *x, int64_t n, float L);
void orc_periodic(float *x, int64_t n, float L);
void orc_minmax(const float *x, int64_t n, float *mn, float *mx);
int  orc_delta_to_depth(float delta, float x0, float x1);
void orc_uniform_bin_index(const float *x, int64_t n, uint8_t level,
                           float x0, float dx, uint32_t *q);
void orc_undo_uniform_bin_index(const uint32_t *q, int64_t n, uint8_t level,
                                float x0, float dx, uint64_t state[2],
                                float *out);
int64_t orc_pack_words(int64_t n, uint8_t width);
void orc_uniform_pack(const uint32_t *x, int64_t n, uint8_t width,
                      uint32_t *out);
void orc_undo_uniform_pack(const uint32_t *in, uint8_t width, int64_t n,
                           uint32_t *out);
uint32_t orc_checksum(const uint8_t *bytes, int64_t n);

void     orc_rand_seed(uint64_t seed, int64_t n, uint64_t *states /*2n*/);
uint64_t orc_rand_u64(uint64_t state[2]);
float    orc_rand_float(uint64_t state[2]);
void     orc_rand_jump(uint64_t state[2]);

/* ---- composition: the Diff v0.9 algorithm ---- */
int64_t orc_diff_nblocks(int32_t n);
int64_t orc_diff_fixed_bytes(int32_t n);
int64_t orc_diff_bound(int32_t n);
int64_t orc_diff_nstreams(int32_t n);

/* float field (3n planar) -> payload. Returns bytes written, <0 on error
 * (-1 depth>24, -2 cap too small, -3 bad field code). q_out (3n) optional. */
int64_t orc_diff_encode_field(uint32_t fieldCode, const float *data, int32_t n,
                              float delta, float width, uint64_t seed,
                              uint8_t *out, int64_t cap, uint32_t *q_out);
/* payload -> float field (3n planar). Returns 0, <0 on malformed input. */
int orc_diff_decode_field(const uint8_t *in, int64_t len, float *out,
                          uint32_t *q_out);

/* quantised values (3n, the QField boundary) -> floats: the fixed quant_Field */
int orc_dequantize_field(const OrcFieldHeader *hd, const uint64_t *q, float *out);

/* integer fields (src/quant.c:291-327, 373-398, 553-608): quantisers alone and
 * as Diff v0.9 payloads. q holds 3n values for ids, n for unsigned ints. */
int orc_id_quantize(const uint64_t *ids, int32_t n, uint64_t width,
                    uint64_t X0[3], uint64_t X1[3], uint64_t *q);
int orc_id_dequantize(const uint64_t *q, int32_t n, uint64_t width,
                      const uint64_t X0[3], uint64_t *ids);
int orc_uint_quantize(const uint64_t *data, int32_t n, uint64_t *x0,
                      uint64_t *x1, uint64_t *q);
int orc_uint_dequantize(const uint64_t *q, int32_t n, uint64_t x0, uint64_t *out);
int64_t orc_diff_encode_id_field(const uint64_t *ids, int32_t n, uint64_t width,
                                 uint8_t *out, int64_t cap);
int orc_diff_decode_id_field(const uint8_t *in, int64_t len, uint64_t *ids);
int64_t orc_diff_encode_uint_field(const uint64_t *data, int32_t n,
                                   uint8_t *out, int64_t cap);
int orc_diff_decode_uint_field(const uint8_t *in, int64_t len, int32_t n, uint64_t *out);

/* segment image: SegmentHeader | FieldHeader[] | BlockHeader[] | blocks
 * (doc/header_format/header_format.tex:129-204) */
typedef struct OrcSegField {
    uint32_t FieldCode, AlgoCode, AlgoVersion;
    int32_t ParticleLen;
    uint8_t *Data;   /* the field's Diff v0.9 payload (CField.Data) */
    int64_t DataLen;
} OrcSegField;
int64_t orc_seg_field_blocks(const OrcSegField *f);
int64_t orc_seg_bytes_bound(const OrcSegField *fields, int32_t nFields);
int64_t orc_seg_to_bytes(const OrcSegField *fields, int32_t nFields,
                         int32_t particleNum, uint8_t *out, int64_t cap);
int32_t orc_seg_from_bytes(uint8_t *bytes, int64_t len, OrcSegField *fields,
                           int32_t maxFields, int32_t *valid, int32_t *particleNum);

/* literal plugin boundary: quantized ints in / out (CFunc / DFunc shape) */
int64_t orc_diff_encode_q(const OrcFieldHeader *hd, const uint64_t *q,
                          uint8_t *out, int64_t cap);
int orc_diff_decode_q(const uint8_t *in, int64_t len, uint64_t *q);

/* quantizer alone (quant.c position()/velocity() as intended) */
int orc_quantize_field(uint32_t fieldCode, const float *data, int32_t n,
                       float delta, float width, OrcFieldHeader *hd,
                       uint32_t *q);

/* multi-threaded batch drivers for the CPU baseline (threads over fields) */
double orc_bench_encode(uint32_t fieldCode, const float *data, int32_t n,
                        int64_t nFields, float delta, float width,
                        int threads, int64_t *bytesOut);
double orc_bench_decode(const uint8_t *const *payloads, const int64_t *lens,
                        int32_t n, int64_t nFields, int threads);

/* bench input made on the host (same recipe as the device generator; not
 * bit-identical to it): ids [firstId, firstId+count) of an ngrid^3 box, planar
 * float32; pos / vel may be NULL */
int orc_gen_zeldovich(int32_t ngrid, float L, uint64_t seed, int64_t firstId,
                      int64_t count, float noiseCells, float *pos, float *vel,
                      int threads);

const char *orc_flavour(void); /* "port", "reference" or "reference-debug" */

#ifdef __cplusplus
}
#endif
#endif

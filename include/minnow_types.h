/* minnow_types.h -- layout-compatible declarations of the reference's
 * boundary types, for building libminnow_cuda outside the minnow tree.
 *
 * Inside the minnow tree include the reference's own "types.h" first: this
 * header then declares nothing (same include guard). Every struct below must
 * stay byte-for-byte identical to /root/reference/src/types.h (cited per
 * struct), because Field / QField / CField cross the C ABI BY VALUE.
 */
#ifndef TYPES_H_
#define TYPES_H_

#include <stdint.h>

/* 4-character codes, big-endian packed (doc/header_format/header_format.tex:360-366) */
#define field_Posn 0x506f736e /* types.h:8  */
#define field_Velc 0x56656c63 /* types.h:9  */
#define field_Ptid 0x50746964 /* types.h:10 */
#define field_Unsf 0x556e7366 /* types.h:11 */
#define field_Unsi 0x556e7369 /* types.h:12 */
#define algo_Test  0x5472696d /* types.h:14 (aliases algo_Trim) */
#define algo_Diff  0x44696666 /* types.h:16: reserved, unimplemented upstream */

typedef void *Accuracy;     /* types.h:26 */
typedef void *Quantization; /* types.h:66 */

typedef struct PositionAccuracy { /* types.h:42-46 */
    float *Deltas;                /* NULL: one Delta for every particle */
    float Delta, Width;
    int32_t Len;
} PositionAccuracy;

typedef struct VelocityAccuracy { /* types.h:48-55 */
    float *Deltas;
    float Delta;
    int32_t Len;
    int32_t SymLog10Scaled;
    float SymLog10Threshold;
} VelocityAccuracy;

typedef struct IDAccuracy { /* types.h:57-59 */
    uint64_t Width;           /* grid cells per side: id = ix + W*iy + W^2*iz */
} IDAccuracy;

typedef uint8_t IntAccuracy; /* types.h:40: does nothing */

typedef struct IntQuantization { /* types.h:77-79 */
    uint64_t X0, X1;
} IntQuantization;

typedef struct IDQuantization { /* types.h:97-99 */
    uint64_t Width, X0[3], X1[3];
} IDQuantization;

typedef struct PositionQuantization { /* types.h:82-87 */
    uint8_t *Depths;
    int32_t Len;
    float Width, X0[3], X1[3];
    uint8_t Depth;
} PositionQuantization;

typedef struct VelocityQuantization { /* types.h:89-95 */
    uint8_t *Depths;
    int32_t Len, SymLog10Scaled;
    float X0[3], X1[3];
    float SymLog10Threshold;
    uint8_t Depth;
} VelocityQuantization;

typedef struct FieldHeader { /* types.h:103-108 */
    uint32_t FieldCode;
    uint32_t AlgoCode;
    uint32_t AlgoVersion;
    int32_t ParticleLen;
} FieldHeader;

typedef struct Field { /* types.h:110-115 */
    FieldHeader Hd;
    int64_t Valid;
    void *Data; /* float[3*ParticleLen], planar x|y|z (quant.c:170-172) */
    Accuracy Acc;
} Field;

typedef struct QField { /* types.h:117-122 */
    FieldHeader Hd;
    int64_t Valid;
    uint64_t *Data; /* uint64[3*ParticleLen], planar */
    Quantization Quant;
} QField;

typedef struct CField { /* types.h:124-129 */
    FieldHeader Hd;
    uint8_t *Data;
    int64_t DataLen;
    uint32_t Checksum; /* filled by the caller (types.h:135, funcs.c:72) */
} CField;

typedef QField (*DFunc)(CField, void *); /* types.h:134 */
typedef CField (*CFunc)(QField, void *); /* types.h:136 */

typedef struct Decompressor { void *Buffer; DFunc DFunc; } Decompressor; /* types.h:138-141 */
typedef struct Compressor { void *Buffer; CFunc CFunc; } Compressor;     /* types.h:143-146 */

typedef struct Seg { Field *Fields; int32_t FieldLen; } Seg;    /* types.h:150-153 */
typedef struct QSeg { QField *Fields; int32_t FieldLen; } QSeg; /* types.h:155-158 */
typedef struct CSeg { CField *Fields; int32_t FieldLen; } CSeg; /* types.h:160-163 */

#endif /* TYPES_H_ */

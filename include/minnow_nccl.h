/* minnow_nccl.h -- C ABI of libminnow_nccl.so: the multi-GPU driver of the
 * encode/decode hot path. One process per GPU; a snapshot
 * is cut into contiguous particle-ID slabs = contiguous runs of segments, one
 * per rank, and every rank encodes its slab with no data-path collective
 * (segments are independent: the reference leaves segmentation to the client,
 * doc/separation_of_duties.md:7-12, so nothing upstream corresponds to this
 * file). NCCL is used for exactly two things:
 *   1. ncclAllGather of each rank's packed byte count (one uint64 per rank) on
 *      the encode stream, followed by an exclusive prefix on the device: the
 *      rank's base offset in the global stream. No host synchronisation.
 *   2. optionally, the gather (or scatter) of the packed slabs to (from) one
 *      rank at the prefix offsets: ONE ncclGroupStart .. ncclSend / ncclRecv x
 *      ranks .. ncclGroupEnd.
 * The ncclUniqueId travels by whatever the launcher has (bench.py broadcasts it
 * with torch.distributed).
 */
#ifndef MINNOW_NCCL_H_
#define MINNOW_NCCL_H_

#include <stdint.h>
#include "minnow_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

#define MNWNCCL_ID_BYTES 128 /* sizeof(ncclUniqueId) */

typedef struct MnwSharded MnwSharded;

/* rank 0: a fresh ncclUniqueId to hand to every rank. Returns 0 or <0. */
int mnwnccl_UniqueId(uint8_t id[MNWNCCL_ID_BYTES]);
/* every rank (collective): a communicator over `world` ranks on ctx's device */
MnwSharded *mnwnccl_Init(MnwCudaCtx *ctx, int rank, int world, const uint8_t id[MNWNCCL_ID_BYTES]);
void mnwnccl_Finalize(MnwSharded *sh);
int mnwnccl_Rank(const MnwSharded *sh);
int mnwnccl_World(const MnwSharded *sh);

/* mnwcuda_EncodeFieldsDev on the rank's slab, then the size exchange on the same
 * stream: d_sizes[world] = every rank's packed bytes, d_base[0] = this rank's
 * exclusive prefix (its slab's offset in the global stream). Enqueued, not
 * waited for. Returns 0, <0 for argument errors (as EncodeFieldsDev), -20 for an
 * NCCL error. */
int mnwnccl_EncodeSharded(MnwSharded *sh, const float *d_in, const MnwFieldDesc *h_fields,
                          int64_t nFields, uint8_t *d_out, int64_t outCap,
                          uint64_t *d_offsets, uint64_t *d_sizes, uint64_t *d_base,
                          void *stream);
/* the mirror: every rank decodes its own slab; no collective is involved */
int mnwnccl_DecodeSharded(MnwSharded *sh, const uint8_t *d_in, const uint64_t *d_offsets,
                          const MnwFieldDesc *h_fields, int64_t nFields, float *d_out,
                          void *stream);
/* Gather the ranks' packed slabs (h_sizes[world] bytes each, host copy of d_sizes)
 * into d_global on `root`, in rank order, as one grouped send/recv; d_global is
 * ignored elsewhere. Scatter is the inverse (root holds the global stream,
 * every rank receives its slab into d_local). Enqueued on `stream`. */
int mnwnccl_GatherStream(MnwSharded *sh, const uint8_t *d_local, const uint64_t *h_sizes,
                         int root, uint8_t *d_global, void *stream);
int mnwnccl_ScatterStream(MnwSharded *sh, const uint8_t *d_global, const uint64_t *h_sizes,
                          int root, uint8_t *d_local, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MINNOW_NCCL_H_ */

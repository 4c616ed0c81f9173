// minnow_nccl.cu -- libminnow_nccl.so: the multi-GPU driver over libminnow_cuda
// and NCCL (include/minnow_nccl.h). Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC
//        minnow_nccl.cu -L.. -lminnow_cuda -lnccl
#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "minnow_nccl.h"

static_assert(sizeof(ncclUniqueId) == MNWNCCL_ID_BYTES, "ncclUniqueId size");

#define NCCL_TRY(expr)                                                                              \
    do {                                                                                            \
        ncclResult_t r_ = (expr);                                                                   \
        if (r_ != ncclSuccess) {                                                                    \
            fprintf(stderr, "[mnwnccl] %s:%d NCCL error: %s\n", __FILE__, __LINE__, ncclGetErrorString(r_)); \
            return -20;                                                                             \
        }                                                                                           \
    } while (0)

struct MnwSharded {
    MnwCudaCtx *ctx;
    ncclComm_t comm;
    int rank, world, device;
};

// exclusive prefix of the gathered sizes for this rank (world is at most a few dozen)
__global__ void k_slab_base(const uint64_t *sizes, int rank, uint64_t *base) {
    uint64_t sum = 0;
    for (int r = 0; r < rank; r++) sum += sizes[r];
    *base = sum;
}

extern "C" int mnwnccl_UniqueId(uint8_t id[MNWNCCL_ID_BYTES]) {
    ncclUniqueId u;
    NCCL_TRY(ncclGetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return 0;
}

extern "C" MnwSharded *mnwnccl_Init(MnwCudaCtx *ctx, int rank, int world, const uint8_t id[MNWNCCL_ID_BYTES]) {
    if (!ctx || rank < 0 || world < 1 || rank >= world) return nullptr;
    MnwSharded *sh = new MnwSharded();
    sh->ctx = ctx;
    sh->rank = rank;
    sh->world = world;
    cudaGetDevice(&sh->device);
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    ncclResult_t r = ncclCommInitRank(&sh->comm, world, u, rank);
    if (r != ncclSuccess) {
        fprintf(stderr, "[mnwnccl] ncclCommInitRank: %s\n", ncclGetErrorString(r));
        delete sh;
        return nullptr;
    }
    return sh;
}

extern "C" void mnwnccl_Finalize(MnwSharded *sh) {
    if (!sh) return;
    ncclCommDestroy(sh->comm);
    delete sh;
}

extern "C" int mnwnccl_Rank(const MnwSharded *sh) { return sh->rank; }
extern "C" int mnwnccl_World(const MnwSharded *sh) { return sh->world; }

extern "C" int mnwnccl_EncodeSharded(MnwSharded *sh, const float *d_in, const MnwFieldDesc *h_fields, int64_t nFields,
                                     uint8_t *d_out, int64_t outCap, uint64_t *d_offsets, uint64_t *d_sizes,
                                     uint64_t *d_base, void *stream) {
    if (!sh || nFields < 1) return -1;
    int rc = mnwcuda_EncodeFieldsDev(sh->ctx, d_in, h_fields, nFields, d_out, outCap, d_offsets, stream);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // the slab's packed size is the last offset: one uint64 per rank, gathered on the encode stream
    NCCL_TRY(ncclAllGather(d_offsets + nFields, d_sizes, 1, ncclUint64, sh->comm, s));
    k_slab_base<<<1, 1, 0, s>>>(d_sizes, sh->rank, d_base);
    return cudaGetLastError() == cudaSuccess ? 0 : -21;
}

extern "C" int mnwnccl_DecodeSharded(MnwSharded *sh, const uint8_t *d_in, const uint64_t *d_offsets,
                                     const MnwFieldDesc *h_fields, int64_t nFields, float *d_out, void *stream) {
    if (!sh) return -1;
    return mnwcuda_DecodeFieldsDev(sh->ctx, d_in, d_offsets, h_fields, nFields, d_out, stream);
}

extern "C" int mnwnccl_GatherStream(MnwSharded *sh, const uint8_t *d_local, const uint64_t *h_sizes, int root,
                                    uint8_t *d_global, void *stream) {
    if (!sh || root < 0 || root >= sh->world) return -1;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    NCCL_TRY(ncclGroupStart());
    if (sh->rank == root) {
        uint64_t pos = 0;
        for (int r = 0; r < sh->world; r++) {
            if (r == root) {
                if (h_sizes[r] && cudaMemcpyAsync(d_global + pos, d_local, h_sizes[r], cudaMemcpyDeviceToDevice, s) != cudaSuccess)
                    return -21;
            } else if (h_sizes[r]) {
                NCCL_TRY(ncclRecv(d_global + pos, h_sizes[r], ncclUint8, r, sh->comm, s));
            }
            pos += h_sizes[r];
        }
    } else if (h_sizes[sh->rank]) {
        NCCL_TRY(ncclSend(d_local, h_sizes[sh->rank], ncclUint8, root, sh->comm, s));
    }
    NCCL_TRY(ncclGroupEnd());
    return 0;
}

extern "C" int mnwnccl_ScatterStream(MnwSharded *sh, const uint8_t *d_global, const uint64_t *h_sizes, int root,
                                     uint8_t *d_local, void *stream) {
    if (!sh || root < 0 || root >= sh->world) return -1;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    NCCL_TRY(ncclGroupStart());
    if (sh->rank == root) {
        uint64_t pos = 0;
        for (int r = 0; r < sh->world; r++) {
            if (r == root) {
                if (h_sizes[r] && cudaMemcpyAsync(d_local, d_global + pos, h_sizes[r], cudaMemcpyDeviceToDevice, s) != cudaSuccess)
                    return -21;
            } else if (h_sizes[r]) {
                NCCL_TRY(ncclSend(d_global + pos, h_sizes[r], ncclUint8, r, sh->comm, s));
            }
            pos += h_sizes[r];
        }
    } else if (h_sizes[sh->rank]) {
        NCCL_TRY(ncclRecv(d_local, h_sizes[sh->rank], ncclUint8, root, sh->comm, s));
    }
    NCCL_TRY(ncclGroupEnd());
    return 0;
}

// Does a tiled TMA load accept a start coordinate that is not 16-byte aligned in global
// memory? 1-D tensor of floats, box of 256, no swizzle; coordinates 0, 3, 257, 1000001.
// nvcc -gencode arch=compute_100a,code=sm_100a -o tma_unaligned tma_unaligned.cu && ./tma_unaligned
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

__global__ void k(const __grid_constant__ CUtensorMap map, int coord, float *out) {
    __shared__ __align__(128) float buf[1024];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(4096));
        for (int q = 0; q < 4; q++)
            asm volatile("cp.async.bulk.tensor.1d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2}], [%3];"
                         ::"r"((uint32_t)__cvta_generic_to_shared(buf + 256 * q)), "l"(&map), "r"(coord + 256 * q),
                           "r"((uint32_t)__cvta_generic_to_shared(&bar)) : "memory");
    }
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"((uint32_t)__cvta_generic_to_shared(&bar)) : "memory");
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) out[i] = buf[i];
}

int main() {
    const size_t n = 4u << 20;
    std::vector<float> h(n);
    for (size_t i = 0; i < n; i++) h[i] = (float)i;
    float *d, *o;
    cudaMalloc(&d, n * 4); cudaMalloc(&o, 4096);
    cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice);
    typedef CUresult (*Fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                           const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fp; cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr);
    CUtensorMap map;
    cuuint64_t dims[1] = {n}; cuuint64_t strides[1] = {0}; cuuint32_t box[1] = {256}; cuuint32_t es[1] = {1};
    CUresult rc = ((Fn)fp)(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 1, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc %d\n", (int)rc);
    int coords[] = {0, 3, 257, 1000001, (int)n - 500};
    for (int c : coords) {
        cudaMemset(o, 0, 4096);
        k<<<1, 128>>>(map, c, o);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<float> r(1024);
        cudaMemcpy(r.data(), o, 4096, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int i = 0; i < 1024; i++) {
            float want = (size_t)(c + i) < n ? (float)(c + i) : 0.0f;
            bad += r[i] != want;
        }
        printf("coord %d: %s, first %.0f last %.0f, mismatches %d\n", c, cudaGetErrorString(e), r[0], r[1023], bad);
    }
    return 0;
}

// l2window.cu -- how large a sliding re-read window does the B200 L2 hold?
// Ticket g reads tile g (first touch) and tile g-lag (second touch) of a 4 GiB
// buffer; time vs lag tells where the second touch stops hitting L2.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2window l2window.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

__device__ __forceinline__ uint64_t pol_keep() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t pol_drop() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ float4 ldh(const float4 *p, uint64_t pol) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ float4 ldp(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// tile = 2048 float4 = 32 KiB; one CTA of 256 threads per ticket
template <int MODE>  // 0 plain, 1 hints
__global__ void __launch_bounds__(256) k(const float4 *in, float *sink, long nTiles, long lag, unsigned *ticket, int second) {
    __shared__ long sT;
    float acc = 0.f;
    while (true) {
        if (threadIdx.x == 0) sT = atomicAdd(ticket, 1u);
        __syncthreads();
        long t = sT;
        __syncthreads();
        if (t >= nTiles + lag) break;
        uint64_t keep = pol_keep(), drop = pol_drop();
        if (t < nTiles) {
            const float4 *p = in + t * 2048 + threadIdx.x;
            float4 v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = MODE ? ldh(p + i * 256, keep) : ldp(p + i * 256);
#pragma unroll
            for (int i = 0; i < 8; i++) acc += v[i].x + v[i].y + v[i].z + v[i].w;
        }
        long b = t - lag;
        if (second && b >= 0 && b < nTiles) {
            const float4 *p = in + b * 2048 + threadIdx.x;
            float4 v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = MODE ? ldh(p + i * 256, drop) : ldp(p + i * 256);
#pragma unroll
            for (int i = 0; i < 8; i++) acc += v[i].x * v[i].y + v[i].z * v[i].w;
        }
    }
    if (acc == 12345.678f) *sink = acc;
}

int main(int argc, char **argv) {
    const long bytes = 4l << 30, nTiles = bytes / 32768;
    float4 *in; float *sink; unsigned *ticket;
    cudaMalloc(&in, bytes); cudaMalloc(&sink, 4); cudaMalloc(&ticket, 4);
    cudaMemset(in, 0, bytes);
    int persist = argc > 1 ? atoi(argv[1]) : 0;
    if (persist) {
        int maxP = 0; cudaDeviceGetAttribute(&maxP, cudaDevAttrMaxPersistingL2CacheSize, 0);
        cudaError_t e = cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)maxP);
        printf("persisting L2 carve-out max %d MiB -> %s\n", maxP >> 20, cudaGetErrorString(e));
    }
    int l2 = 0; cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, 0);
    printf("L2 size %d MiB\n", l2 >> 20);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    int grid = 148 * 4;
    for (int mode = 0; mode < 2; mode++) {
        for (int second = 0; second < 2; second++) {
            long lags[] = {256, 512, 1024, 1536, 2048, 2560, 3072, 3584, 4096, 8192};  // x32 KiB
            for (long lag : lags) {
                if (!second && lag != 256) continue;
                float best = 1e9;
                for (int rep = 0; rep < 3; rep++) {
                    cudaMemset(ticket, 0, 4);
                    cudaEventRecord(a);
                    if (mode) k<1><<<grid, 256>>>(in, sink, nTiles, lag, ticket, second);
                    else k<0><<<grid, 256>>>(in, sink, nTiles, lag, ticket, second);
                    cudaEventRecord(b); cudaEventSynchronize(b);
                    float ms; cudaEventElapsedTime(&ms, a, b);
                    if (ms < best) best = ms;
                }
                printf("mode %s second %d lag %5ld tiles (%4ld MiB window): %.3f ms  first-touch rate %.0f GB/s\n",
                       mode ? "hints" : "plain", second, lag, lag * 32 / 1024, best, bytes / best / 1e6);
            }
        }
    }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

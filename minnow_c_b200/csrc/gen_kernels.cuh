// gen_kernels.cuh -- synthetic Zel'dovich-displaced Lagrangian grid (bench and
// test input only; nothing in the reference corresponds to it). Particle id =
// ix + N*iy + N^2*iz, x fastest: the order id() assumes
// (/root/reference/src/quant.c:308-312).
#pragma once

#include "minnow_kernels.cuh"

namespace mnw {

constexpr int kGenModes = 64;

struct GenModes {
    int kx[kGenModes], ky[kGenModes], kz[kGenModes];
    float phase[kGenModes];
    float ax[kGenModes], ay[kGenModes], az[kGenModes];  // displacement amplitude, box units
};

struct GenParams {
    int ngrid;
    float L;
    float velPerLen;  // velocity = velPerLen * displacement
    int64_t firstId, count;
    float *pos, *vel;  // planar [3][count], may be null
    float noise;       // white-noise displacement per component, uniform in [-noise, noise) box units (0: none)
    uint64_t noiseSeed;
};

// splitmix64 finaliser as a hash of (seed, particle id, component): the noise of a
// particle does not depend on how the ids are cut into calls
__device__ __forceinline__ float gen_noise(uint64_t seed, int64_t id, int c) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (uint64_t)(3 * id + c + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (float)(z >> 40) * (1.0f / 8388608.0f) - 1.0f;  // [-1, 1)
}

__constant__ GenModes cModes;

__global__ void __launch_bounds__(256) k_zeldovich(GenParams p) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.count) return;
    int64_t id = p.firstId + i;
    const int N = p.ngrid;
    int ix = (int)(id % N), iy = (int)((id / N) % N), iz = (int)(id / ((int64_t)N * N));
    float px = 0.f, py = 0.f, pz = 0.f;
    const float piOverN = 3.14159265358979f / (float)N;
#pragma unroll 8
    for (int m = 0; m < kGenModes; m++) {
        // phase of mode m at the cell centre: 2*pi*k.(i+1/2)/N, reduced mod 2N
        int t = cModes.kx[m] * (2 * ix + 1) + cModes.ky[m] * (2 * iy + 1) + cModes.kz[m] * (2 * iz + 1);
        t %= 2 * N;
        float s = sinf((float)t * piOverN + cModes.phase[m]);
        px = fmaf(cModes.ax[m], s, px);
        py = fmaf(cModes.ay[m], s, py);
        pz = fmaf(cModes.az[m], s, pz);
    }
    const float cell = p.L / (float)N;
    float psi[3] = {px, py, pz};
    int idx[3] = {ix, iy, iz};
    if (p.noise > 0.0f) {
#pragma unroll
        for (int c = 0; c < 3; c++) psi[c] += p.noise * gen_noise(p.noiseSeed, id, c);
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
        if (p.pos) {
            float x = ((float)idx[c] + 0.5f) * cell + psi[c];
            x -= p.L * floorf(x / p.L);
            if (!(x >= 0.0f && x < p.L)) x = 0.0f;
            p.pos[(int64_t)c * p.count + i] = x;
        }
        if (p.vel) p.vel[(int64_t)c * p.count + i] = p.velPerLen * psi[c];
    }
}

}  // namespace mnw

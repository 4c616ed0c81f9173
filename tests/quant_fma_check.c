/* Test helper (CPU only, compiled by tests/test_quantiser_exact.py): the encoder's
 * FP32 quantiser -- s = RN_f(a / maxDiff) * 2^depth by FMA division with a
 * per-field reciprocal, minnow_c_b200/csrc/enc_math.cuh quantise32<.,.,true> and
 * compute_quant -- against util_UniformBinIndex's arithmetic
 * (/root/reference/src/util.c:184-191): t = a/dx (IEEE), index = trunc(t * 2^depth)
 * clamped to [0, 2^depth - 1]. Returns the number of index mismatches. */
#include <math.h>
#include <stdint.h>
#include <string.h>

static uint64_t sm64(uint64_t *s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static float step(float x, int k) {
    uint32_t b;
    memcpy(&b, &x, 4);
    b += (uint32_t)k;
    memcpy(&x, &b, 4);
    return x;
}

long quant_fma_mismatches(uint64_t seed, long fields, int per_field, int emin, int emax) {
    long bad = 0;
    for (long it = 0; it < fields; it++) {
        const int depth = (int)(sm64(&seed) % 25);
        uint32_t m = (uint32_t)sm64(&seed) & 0x7fffff;
        const int e = emin + (int)(sm64(&seed) % (uint64_t)(emax - emin + 1));
        if (sm64(&seed) % 8 == 0) m = (sm64(&seed) % 2) ? 0x7fffff : (sm64(&seed) % 2 ? 0 : 0x7ffffe);
        const float dx = ldexpf(1.0f + (float)m / 8388608.0f, e);
        /* compute_quant: y = RN_f(RN_d(2^depth / dx)), ds = dx * 2^-depth */
        const float y = (float)(ldexp(1.0, depth) / (double)dx);
        const float nds = -ldexpf(dx, -depth);
        const float top = ldexpf(1.0f, depth) - 1.0f;
        for (int k = 0; k < per_field; k++) {
            const uint64_t r = sm64(&seed);
            float a;
            switch (r & 3) {
            case 0: a = dx * (float)((r >> 8) & 0xffffff) / 16777216.0f; break;
            case 1: { /* a few ulps around a bin edge */
                const uint64_t kk = (r >> 8) % ((1ull << depth) + 1);
                a = step((float)((double)kk * (double)dx / ldexp(1.0, depth)), (int)((r >> 40) % 7) - 3);
                break;
            }
            case 2: { /* the quotient a few ulps around a rounding midpoint */
                const double q = ldexp(1.0 + ((double)((r >> 8) & 0x7fffff) + 0.5) / 8388608.0, -(int)((r >> 32) % 26));
                a = step((float)(q * (double)dx), (int)((r >> 40) % 5) - 2);
                break;
            }
            default: a = step(dx, -(int)((r >> 8) % 1000)); break;
            }
            if (!(a >= 0.0f) || a > dx) a = dx * 0.5f;
            const float t = a / dx;
            float want = t * ldexpf(1.0f, depth);
            float s = a * y;
            s = fmaf(fmaf(nds, s, a), y, s);
            s = fmaf(fmaf(nds, s, a), y, s);
            if (floorf(fminf(s, top)) != floorf(fminf(want, top))) bad++;
        }
    }
    return bad;
}

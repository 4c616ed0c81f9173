/* minnow_oracle.c -- TEST INFRASTRUCTURE ONLY (see minnow_oracle.h).
 *
 * Build: gcc -O2 -std=c99 -ffp-contract=off (no -march=native / -mfma /
 * -ffast-math: bit parity depends on separately rounded float operations).
 *
 * With -DORC_USE_REFERENCE the arithmetic primitives below are replaced by
 * calls into the reference's own objects (oracle/ref_bridge.c), so that the
 * same composition can be run on the restated and on the real primitives.
 */
#define _POSIX_C_SOURCE 200809L
#include "minnow_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------ */
/* primitives                                                          */
/* ------------------------------------------------------------------ */

static uint64_t splitmix64(uint64_t *s) { /* src/rand.c:114-121 */
    uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

#ifdef ORC_USE_REFERENCE

/* provided by oracle/ref_bridge.c, thin wrappers over util_* / rand_* */
void ref_undo_periodic(float *x, int64_t n, float L);
void ref_periodic(float *x, int64_t n, float L);
void ref_minmax(const float *x, int64_t n, float *mn, float *mx);
int  ref_delta_to_depth(float delta, float x0, float x1);
void ref_uniform_bin_index(const float *x, int64_t n, uint8_t level,
                           float x0, float dx, uint32_t *q);
void ref_undo_uniform_bin_index(const uint32_t *q, int64_t n, uint8_t level,
                                float x0, float dx, uint64_t state[2],
                                float *out);
void ref_uniform_pack(const uint32_t *x, int64_t n, uint8_t width,
                      uint32_t *out);
void ref_undo_uniform_pack(const uint32_t *in, uint8_t width, int64_t n,
                           uint32_t *out);
uint32_t ref_checksum(const uint8_t *bytes, int64_t n);
void ref_rand_seed(uint64_t seed, int64_t n, uint64_t *states);
uint64_t ref_rand_u64(uint64_t state[2]);
float ref_rand_float(uint64_t state[2]);
void ref_rand_jump(uint64_t state[2]);

#ifdef ORC_REFERENCE_DEBUG
const char *orc_flavour(void) { return "reference-debug"; }
#else
const char *orc_flavour(void) { return "reference"; }
#endif

void orc_undo_periodic(float *x, int64_t n, float L) { ref_undo_periodic(x, n, L); }
void orc_periodic(float *x, int64_t n, float L) { ref_periodic(x, n, L); }
void orc_minmax(const float *x, int64_t n, float *mn, float *mx) { ref_minmax(x, n, mn, mx); }
int orc_delta_to_depth(float delta, float x0, float x1) { return ref_delta_to_depth(delta, x0, x1); }
void orc_uniform_bin_index(const float *x, int64_t n, uint8_t level, float x0,
                           float dx, uint32_t *q) {
    ref_uniform_bin_index(x, n, level, x0, dx, q);
}
void orc_undo_uniform_bin_index(const uint32_t *q, int64_t n, uint8_t level,
                                float x0, float dx, uint64_t state[2],
                                float *out) {
    ref_undo_uniform_bin_index(q, n, level, x0, dx, state, out);
}
void orc_uniform_pack(const uint32_t *x, int64_t n, uint8_t width, uint32_t *out) {
    ref_uniform_pack(x, n, width, out);
}
void orc_undo_uniform_pack(const uint32_t *in, uint8_t width, int64_t n, uint32_t *out) {
    ref_undo_uniform_pack(in, width, n, out);
}
uint32_t orc_checksum(const uint8_t *bytes, int64_t n) { return ref_checksum(bytes, n); }
void orc_rand_seed(uint64_t seed, int64_t n, uint64_t *states) { ref_rand_seed(seed, n, states); }
uint64_t orc_rand_u64(uint64_t state[2]) { return ref_rand_u64(state); }
float orc_rand_float(uint64_t state[2]) { return ref_rand_float(state); }
void orc_rand_jump(uint64_t state[2]) { ref_rand_jump(state); }

#else /* restated primitives */

const char *orc_flavour(void) { return "port"; }

/* src/util.c:97-113. The anchor is the first element; a value half a box
 * or more above it is pulled down by L, one more than half a box below it is
 * pushed up by L. Element 0 goes through the same test (no-op for L > 0). */
void orc_undo_periodic(float *x, int64_t n, float L) {
    if (n <= 0) return;
    const float anchor = x[0];
    const float half = L / 2;
    for (int64_t i = 0; i < n; i++) {
        float rel = x[i] - anchor;
        if (rel >= half) x[i] = x[i] - L;
        else if (rel < -half) x[i] = x[i] + L;
    }
}

/* src/util.c:70-84 */
void orc_periodic(float *x, int64_t n, float L) {
    for (int64_t i = 0; i < n; i++) {
        float v = x[i];
        if (v >= L) x[i] = v - L;
        else if (v < 0) x[i] = v + L;
    }
}

/* src/util.c:27-46: strict comparisons, so NaNs are skipped and the first
 * of several equal extrema is kept. */
void orc_minmax(const float *x, int64_t n, float *mn, float *mx) {
    float lo = x[0], hi = x[0];
    for (int64_t i = 1; i < n; i++) {
        float v = x[i];
        if (v > hi) hi = v;
        if (v < lo) lo = v;
    }
    *mn = lo;
    *mx = hi;
}

/* src/quant.c:684-693 (uniform-accuracy branch): smallest depth in 0..24
 * with delta*2^depth > x1-x0, all in float. Returns -1 where the reference
 * Panics. */
int orc_delta_to_depth(float delta, float x0, float x1) {
    float range = x1 - x0;
    for (int d = 0; d <= 24; d++) {
        float span = delta * (float)(1 << d);
        if (span > range) return d;
    }
    return -1;
}

/* src/util.c:173-196. IEEE single division, then scale and truncate. */
void orc_uniform_bin_index(const float *x, int64_t n, uint8_t level,
                           float x0, float dx, uint32_t *q) {
    const float bins = (float)(1 << level);
    const uint32_t top = (1u << (uint32_t)level) - 1u;
    for (int64_t i = 0; i < n; i++) {
        float t = (x[i] - x0) / dx;
        if (t < 0) q[i] = 0;
        else if (t >= 1) q[i] = top;
        else q[i] = (uint32_t)(t * bins);
    }
}

static inline uint64_t rotl64(uint64_t v, int k) {
    return (v << k) | (v >> (64 - k));
}

/* xorshift128+ with constants (55,14,36): src/rand.c:80-90 */
uint64_t orc_rand_u64(uint64_t state[2]) {
    uint64_t a = state[0], b = state[1];
    uint64_t sum = a + b;
    b ^= a;
    state[0] = rotl64(a, 55) ^ b ^ (b << 14);
    state[1] = rotl64(b, 36);
    return sum;
}

/* src/rand.c:60-64: low 24 bits / 2^24 */
float orc_rand_float(uint64_t state[2]) {
    uint64_t r = orc_rand_u64(state) & 0xffffffu;
    return (float)r / 16777216.0f;
}

/* 2^64-step jump, src/rand.c:96-112 */
void orc_rand_jump(uint64_t state[2]) {
    static const uint64_t poly[2] = {0xbeac0467eba5facbULL,
                                     0xd86b048b86aa9922ULL};
    uint64_t a = 0, b = 0;
    for (int w = 0; w < 2; w++) {
        for (int bit = 0; bit < 64; bit++) {
            if ((poly[w] >> bit) & 1u) { a ^= state[0]; b ^= state[1]; }
            (void)orc_rand_u64(state);
        }
    }
    state[0] = a;
    state[1] = b;
}


/* src/rand.c:22-39: stream i is stream i-1 jumped by 2^64 draws. */
void orc_rand_seed(uint64_t seed, int64_t n, uint64_t *states) {
    if (n <= 0) return;
    states[0] = splitmix64(&seed);
    states[1] = splitmix64(&seed);
    for (int64_t i = 1; i < n; i++) {
        states[2 * i] = states[2 * i - 2];
        states[2 * i + 1] = states[2 * i - 1];
        orc_rand_jump(states + 2 * i);
    }
}

/* src/util.c:223-242: the bin's lower edge and the dither are two
 * separately rounded statements. One draw per value, in order. */
void orc_undo_uniform_bin_index(const uint32_t *q, int64_t n, uint8_t level,
                                float x0, float dx, uint64_t state[2],
                                float *out) {
    const float bw = dx / (float)((uint64_t)1 << level);
    for (int64_t i = 0; i < n; i++) {
        float edge = x0 + bw * (float)(uint64_t)q[i];
        out[i] = edge + orc_rand_float(state) * bw;
    }
}

/* src/util.c:311-355: value i lives in bits [w*i, w*i+w) of a little-endian
 * stream of 32-bit words, LSB first. */
void orc_uniform_pack(const uint32_t *x, int64_t n, uint8_t width,
                      uint32_t *out) {
    int64_t words = orc_pack_words(n, width);
    if (words == 0) return;
    if (width == 32) { memcpy(out, x, (size_t)n * 4); return; }
    memset(out, 0, (size_t)words * 4);
    const uint64_t mask = ((uint64_t)1 << width) - 1;
    for (int64_t i = 0; i < n; i++) {
        uint64_t bit = (uint64_t)width * (uint64_t)i;
        uint64_t v = ((uint64_t)x[i] & mask) << (bit & 31);
        out[bit >> 5] |= (uint32_t)v;
        if (v >> 32) out[(bit >> 5) + 1] |= (uint32_t)(v >> 32);
    }
}

/* src/util.c:357-406 */
void orc_undo_uniform_pack(const uint32_t *in, uint8_t width, int64_t n,
                           uint32_t *out) {
    if (n <= 0) return;
    if (width == 32) { memcpy(out, in, (size_t)n * 4); return; }
    if (width == 0) { memset(out, 0, (size_t)n * 4); return; }
    const uint64_t mask = ((uint64_t)1 << width) - 1;
    for (int64_t i = 0; i < n; i++) {
        uint64_t bit = (uint64_t)width * (uint64_t)i;
        uint64_t lo = in[bit >> 5];
        uint64_t window = lo;
        if ((bit & 31) + width > 32) window |= (uint64_t)in[(bit >> 5) + 1] << 32;
        out[i] = (uint32_t)((window >> (bit & 31)) & mask);
    }
}

/* src/util.c:438-445: rotate right by one, add byte; starts at 1. */
uint32_t orc_checksum(const uint8_t *bytes, int64_t n) {
    uint32_t c = 1;
    for (int64_t i = 0; i < n; i++) {
        c = (c >> 1) | (c << 31);
        c += bytes[i];
    }
    return c;
}

#endif /* ORC_USE_REFERENCE */

/* src/util.c:316-317: ceil(width*n/32) words; 0 when width==0 or n==0 */
int64_t orc_pack_words(int64_t n, uint8_t width) {
    uint64_t bits = (uint64_t)width * (uint64_t)n;
    return (int64_t)((bits + 31) / 32);
}

/* ------------------------------------------------------------------ */
/* composition: the Diff v0.9 algorithm                                 */
/* ------------------------------------------------------------------ */

int64_t orc_diff_nblocks(int32_t n) {
    return ((int64_t)n + ORC_BLOCK_LEN - 1) / ORC_BLOCK_LEN;
}

int64_t orc_diff_fixed_bytes(int32_t n) {
    return (int64_t)sizeof(OrcFieldHeader) +
           3 * orc_diff_nblocks(n) * (int64_t)sizeof(OrcBlockEntry);
}

/* worst case: every block at 25 bits, each padded to 16 bytes */
int64_t orc_diff_bound(int32_t n) {
    int64_t nb = orc_diff_nblocks(n);
    int64_t perBlock = (int64_t)ORC_BLOCK_LEN * 25 / 8;
    return orc_diff_fixed_bytes(n) + 3 * nb * ((perBlock + 15) / 16 * 16);
}

int64_t orc_diff_nstreams(int32_t n) {
    int64_t g = 3 * orc_diff_nblocks(n);
    return (g + ORC_DITHER_RUN - 1) / ORC_DITHER_RUN * ORC_DITHER_LANES;
}

static float plus_zero(float v) { return v == 0 ? 0.0f : v; }

/* max over components of X1-X0, the loop of src/quant.c:186-192 */
static float max_diff(const float X0[3], const float X1[3]) {
    float m = 0;
    for (int c = 0; c < 3; c++) {
        float d = X1[c] - X0[c];
        if (m < d) m = d;
    }
    return m;
}

static uint32_t bit_length(uint32_t v) {
    uint32_t b = 0;
    while (v) { b++; v >>= 1; }
    return b;
}

/* position()/velocity() of src/quant.c:161-289 as intended (3n outputs,
 * header filled in): unwrap (Posn), min/max per component, one depth from
 * the widest component, one bin width for all three. */
int orc_quantize_field(uint32_t fieldCode, const float *data, int32_t n,
                       float delta, float width, OrcFieldHeader *hd,
                       uint32_t *q) {
    if (fieldCode != ORC_FIELD_POSN && fieldCode != ORC_FIELD_VELC) return -3;
    memset(hd, 0, sizeof(*hd));
    hd->Magic = ORC_ALGO_DIFF;
    hd->Version = ORC_VERSION_0_9;
    hd->FieldCode = fieldCode;
    hd->ParticleLen = n;
    hd->Width = fieldCode == ORC_FIELD_POSN ? width : 0.0f;
    hd->BlockLen = ORC_BLOCK_LEN;
    hd->DitherRun = ORC_DITHER_RUN;
    hd->DitherLanes = ORC_DITHER_LANES;
    if (n <= 0) return 0;

    float *tmp = malloc((size_t)n * 3 * sizeof(float));
    if (!tmp) return -4;
    memcpy(tmp, data, (size_t)n * 3 * sizeof(float));
    for (int c = 0; c < 3; c++) {
        float *xc = tmp + (size_t)c * n;
        if (fieldCode == ORC_FIELD_POSN) orc_undo_periodic(xc, n, width);
        orc_minmax(xc, n, &hd->X0[c], &hd->X1[c]);
        hd->X0[c] = plus_zero(hd->X0[c]);
        hd->X1[c] = plus_zero(hd->X1[c]);
    }
    float md = max_diff(hd->X0, hd->X1);
    int depth = orc_delta_to_depth(delta, hd->X0[0], hd->X0[0] + md);
    if (depth < 0) { free(tmp); return -1; }
    hd->Depth = (uint8_t)depth;
    for (int c = 0; c < 3; c++) {
        orc_uniform_bin_index(tmp + (size_t)c * n, n, hd->Depth, hd->X0[c],
                              md, q + (size_t)c * n);
    }
    free(tmp);
    return 0;
}

/* delta + per-block width + pack of already-quantized values */
static int64_t encode_blocks(const OrcFieldHeader *hd, const uint32_t *q,
                             uint8_t *out, int64_t cap) {
    int32_t n = hd->ParticleLen;
    int64_t nb = orc_diff_nblocks(n);
    int64_t fixed = orc_diff_fixed_bytes(n);
    if (cap < fixed) return -2;
    memcpy(out, hd, sizeof(*hd));
    OrcBlockEntry *table = (OrcBlockEntry *)(out + sizeof(*hd));
    uint8_t *payload = out + fixed;
    int64_t off16 = 0;
    uint32_t r[ORC_BLOCK_LEN];
    uint32_t words[ORC_BLOCK_LEN];

    for (int c = 0; c < 3; c++) {
        for (int64_t b = 0; b < nb; b++) {
            const uint32_t *qb = q + (size_t)c * n + (size_t)b * ORC_BLOCK_LEN;
            int64_t len = n - b * ORC_BLOCK_LEN;
            if (len > ORC_BLOCK_LEN) len = ORC_BLOCK_LEN;

            int32_t lo = 0, hi = 0;
            for (int64_t j = 1; j < len; j++) {
                int32_t d = (int32_t)(qb[j] - qb[j - 1]);
                if (j == 1 || d < lo) lo = d;
                if (j == 1 || d > hi) hi = d;
            }
            r[0] = 0;
            for (int64_t j = 1; j < len; j++)
                r[j] = (uint32_t)((int32_t)(qb[j] - qb[j - 1]) - lo);
            uint32_t w = bit_length((uint32_t)(hi - lo));

            int64_t nw = orc_pack_words(len, (uint8_t)w);
            int64_t bytes16 = (nw * 4 + 15) / 16;
            if (fixed + (off16 + bytes16) * 16 > cap) return -2;
            orc_uniform_pack(r, len, (uint8_t)w, words);
            uint8_t *dst = payload + off16 * 16;
            memset(dst, 0, (size_t)bytes16 * 16);
            memcpy(dst, words, (size_t)nw * 4);

            OrcBlockEntry *e = &table[c * nb + b];
            e->First = qb[0];
            e->MinDelta = lo;
            e->Width = w;
            e->Offset16 = (uint32_t)off16;
            off16 += bytes16;
        }
    }
    return fixed + off16 * 16;
}

int64_t orc_diff_encode_field(uint32_t fieldCode, const float *data, int32_t n,
                              float delta, float width, uint64_t seed,
                              uint8_t *out, int64_t cap, uint32_t *q_out) {
    OrcFieldHeader hd;
    uint32_t *q = q_out;
    if (!q) q = malloc((size_t)(n > 0 ? n : 1) * 3 * sizeof(uint32_t));
    int rc = orc_quantize_field(fieldCode, data, n, delta, width, &hd, q);
    int64_t bytes = rc;
    if (rc == 0) {
        hd.DitherSeed = seed;
        bytes = encode_blocks(&hd, q, out, cap);
    }
    if (!q_out) free(q);
    return bytes;
}

int64_t orc_diff_encode_q(const OrcFieldHeader *hd, const uint64_t *q,
                          uint8_t *out, int64_t cap) {
    int64_t m = 3 * (int64_t)hd->ParticleLen;
    uint32_t *q32 = malloc((size_t)(m > 0 ? m : 1) * sizeof(uint32_t));
    for (int64_t i = 0; i < m; i++) q32[i] = (uint32_t)q[i];
    int64_t bytes = encode_blocks(hd, q32, out, cap);
    free(q32);
    return bytes;
}

/* unpack + inverse delta for every block; q is 3n values */
static int decode_blocks(const uint8_t *in, int64_t len, uint32_t *q) {
    if (len < (int64_t)sizeof(OrcFieldHeader)) return -1;
    OrcFieldHeader hd;
    memcpy(&hd, in, sizeof(hd));
    if (hd.Magic != ORC_ALGO_DIFF || hd.BlockLen != ORC_BLOCK_LEN ||
        hd.DitherRun != ORC_DITHER_RUN || hd.DitherLanes != ORC_DITHER_LANES)
        return -2;
    int32_t n = hd.ParticleLen;
    int64_t nb = orc_diff_nblocks(n);
    int64_t fixed = orc_diff_fixed_bytes(n);
    if (len < fixed) return -1;
    const OrcBlockEntry *table = (const OrcBlockEntry *)(in + sizeof(hd));
    uint32_t r[ORC_BLOCK_LEN];
    uint32_t words[ORC_BLOCK_LEN];

    for (int64_t g = 0; g < 3 * nb; g++) {
        int64_t b = g % nb, c = g / nb;
        int64_t cnt = n - b * ORC_BLOCK_LEN;
        if (cnt > ORC_BLOCK_LEN) cnt = ORC_BLOCK_LEN;
        OrcBlockEntry e = table[g];
        uint32_t *qb = q + (size_t)c * n + (size_t)b * ORC_BLOCK_LEN;
        if (e.Width == ORC_WIDTH_DAMAGED) { /* orc_seg_from_bytes: the block's checksum failed */
            for (int64_t j = 0; j < cnt; j++) qb[j] = 0; /* placeholders: the values come out as NaN */
            continue;
        }
        if (e.Width > 32) return -3;
        int64_t nw = orc_pack_words(cnt, (uint8_t)e.Width);
        int64_t start = fixed + (int64_t)e.Offset16 * 16;
        if (start + nw * 4 > len) return -1;
        memcpy(words, in + start, (size_t)nw * 4);
        orc_undo_uniform_pack(words, (uint8_t)e.Width, cnt, r);
        qb[0] = e.First;
        for (int64_t j = 1; j < cnt; j++)
            qb[j] = qb[j - 1] + r[j] + (uint32_t)e.MinDelta;
    }
    return 0;
}

int orc_diff_decode_q(const uint8_t *in, int64_t len, uint64_t *q) {
    if (len < (int64_t)sizeof(OrcFieldHeader)) return -1;
    OrcFieldHeader hd;
    memcpy(&hd, in, sizeof(hd));
    int64_t m = 3 * (int64_t)hd.ParticleLen;
    uint32_t *q32 = malloc((size_t)(m > 0 ? m : 1) * sizeof(uint32_t));
    int rc = decode_blocks(in, len, q32);
    if (rc == 0) for (int64_t i = 0; i < m; i++) q[i] = q32[i];
    free(q32);
    return rc;
}

/* quant_Field as intended (src/quant.c:451-551 with the fixes listed in
 * DESIGN.md): dx = maxDiff for every component, util.c:237-238's two-step
 * dither, explicit seed, one stream per (task, lane), 32 values per visit.
 * `damaged` (may be NULL): per block, non-zero = its values come out as NaN
 * (the stream still advances, so the other blocks are unaffected). */
static int dequantize(const OrcFieldHeader *hd, const uint32_t *q,
                      const uint8_t *damaged, float *out) {
    int32_t n = hd->ParticleLen;
    if (n <= 0) return n == 0 ? 0 : -1;
    int64_t nb = orc_diff_nblocks(n);
    int64_t ns = orc_diff_nstreams(n);
    uint64_t *states = malloc((size_t)ns * 2 * sizeof(uint64_t));
    orc_rand_seed(hd->DitherSeed, ns, states);
    float md = max_diff(hd->X0, hd->X1);

    for (int64_t g = 0; g < 3 * nb; g++) {
        int64_t b = g % nb, c = g / nb;
        int64_t cnt = n - b * ORC_BLOCK_LEN;
        if (cnt > ORC_BLOCK_LEN) cnt = ORC_BLOCK_LEN;
        int64_t task = g / ORC_DITHER_RUN;
        size_t base = (size_t)c * n + (size_t)b * ORC_BLOCK_LEN;
        for (int64_t lane = 0; lane * ORC_CHUNK < cnt; lane++) {
            int64_t lo = lane * ORC_CHUNK;
            int64_t m = cnt - lo < ORC_CHUNK ? cnt - lo : ORC_CHUNK;
            orc_undo_uniform_bin_index(
                q + base + lo, m, hd->Depth, hd->X0[c], md,
                states + 2 * (task * ORC_DITHER_LANES + lane),
                out + base + lo);
        }
    }
    if (hd->FieldCode == ORC_FIELD_POSN)
        for (int c = 0; c < 3; c++)
            orc_periodic(out + (size_t)c * n, n, hd->Width);
    if (damaged)
        for (int64_t g = 0; g < 3 * nb; g++) {
            if (!damaged[g]) continue;
            int64_t b = g % nb, c = g / nb;
            int64_t cnt = n - b * ORC_BLOCK_LEN;
            if (cnt > ORC_BLOCK_LEN) cnt = ORC_BLOCK_LEN;
            uint32_t nanBits = 0x7fc00000u;
            for (int64_t j = 0; j < cnt; j++)
                memcpy(out + (size_t)c * n + (size_t)b * ORC_BLOCK_LEN + j, &nanBits, 4);
        }
    free(states);
    return 0;
}

/* the quantised values of a field (3n, as the literal QField boundary carries
 * them) -> floats: the fixed quant_Field for Posn / Velc */
int orc_dequantize_field(const OrcFieldHeader *hd, const uint64_t *q, float *out) {
    int64_t m = 3 * (int64_t)hd->ParticleLen;
    uint32_t *q32 = malloc((size_t)(m > 0 ? m : 1) * sizeof(uint32_t));
    for (int64_t i = 0; i < m; i++) q32[i] = (uint32_t)q[i];
    int rc = dequantize(hd, q32, NULL, out);
    free(q32);
    return rc;
}

/* DFunc + quant_Field */
int orc_diff_decode_field(const uint8_t *in, int64_t len, float *out,
                          uint32_t *q_out) {
    if (len < (int64_t)sizeof(OrcFieldHeader)) return -1;
    OrcFieldHeader hd;
    memcpy(&hd, in, sizeof(hd));
    int32_t n = hd.ParticleLen;
    if (n <= 0) return n == 0 ? 0 : -1;
    uint32_t *q = q_out;
    if (!q) q = malloc((size_t)n * 3 * sizeof(uint32_t));
    int rc = decode_blocks(in, len, q);
    if (rc != 0) { if (!q_out) free(q); return rc; }
    int64_t nb = orc_diff_nblocks(n);
    uint8_t *damaged = calloc((size_t)(3 * nb), 1);
    const OrcBlockEntry *table = (const OrcBlockEntry *)(in + sizeof(hd));
    int any = 0;
    for (int64_t g = 0; g < 3 * nb; g++)
        if (table[g].Width == ORC_WIDTH_DAMAGED) { damaged[g] = 1; any = 1; }
    rc = dequantize(&hd, q, any ? damaged : NULL, out);
    free(damaged);
    if (!q_out) free(q);
    return rc;
}

/* ------------------------------------------------------------------ */
/* integer fields: particle IDs and unsigned ints (lossless)            */
/* ------------------------------------------------------------------ */

#ifdef ORC_USE_REFERENCE
int ref_quant_id(const uint64_t *ids, int32_t n, uint64_t width,
                 uint64_t X0[3], uint64_t X1[3], uint64_t *q);
int ref_undo_id(const uint64_t *q, int32_t n, uint64_t width,
                const uint64_t X0[3], uint64_t *ids);
int ref_quant_uint(const uint64_t *data, int32_t n, uint64_t *x0, uint64_t *x1,
                   uint64_t *q);
int ref_undo_uint(const uint64_t *q, int32_t n, uint64_t x0, uint64_t *out);

int orc_id_quantize(const uint64_t *ids, int32_t n, uint64_t width,
                    uint64_t X0[3], uint64_t X1[3], uint64_t *q) {
    return ref_quant_id(ids, n, width, X0, X1, q);
}
int orc_id_dequantize(const uint64_t *q, int32_t n, uint64_t width,
                      const uint64_t X0[3], uint64_t *ids) {
    return ref_undo_id(q, n, width, X0, ids);
}
int orc_uint_quantize(const uint64_t *data, int32_t n, uint64_t *x0,
                      uint64_t *x1, uint64_t *q) {
    return ref_quant_uint(data, n, x0, x1, q);
}
int orc_uint_dequantize(const uint64_t *q, int32_t n, uint64_t x0, uint64_t *out) {
    return ref_undo_uint(q, n, x0, out);
}
#else
/* id() of src/quant.c:291-327: the id is split into three grid coordinates
 * (x fastest), each component is unwrapped around its first element with
 * util_U64UndoPeriodic (src/util.c:115-143: signed tests against +-L/2 from
 * the second element on, then everything shifted up by L if the minimum went
 * negative), and the component's minimum is subtracted. */
int orc_id_quantize(const uint64_t *ids, int32_t n, uint64_t width,
                    uint64_t X0[3], uint64_t X1[3], uint64_t *q) {
    if (n <= 0 || width == 0) return -1;
    for (int32_t i = 0; i < n; i++) {
        q[i] = ids[i] % width;
        q[(size_t)n + i] = (ids[i] / width) % width;
        q[2 * (size_t)n + i] = ids[i] / (width * width);
    }
    const int64_t L = (int64_t)width;
    for (int c = 0; c < 3; c++) {
        int64_t *x = (int64_t *)(q + (size_t)c * n);
        const int64_t x0 = x[0];
        for (int32_t i = 1; i < n; i++) {
            if (x[i] - x0 >= L / 2) x[i] -= L;
            else if (x[i] - x0 < -L / 2) x[i] += L;
        }
        int64_t mn = x0;
        for (int32_t i = 1; i < n; i++) if (x[i] < mn) mn = x[i];
        if (mn < 0) for (int32_t i = 0; i < n; i++) x[i] += L;
        /* util_U64MinMax, src/util.c:48-68 */
        uint64_t *u = (uint64_t *)x, lo = u[0], hi = u[0];
        for (int32_t i = 1; i < n; i++) {
            if (u[i] > hi) hi = u[i];
            if (u[i] < lo) lo = u[i];
        }
        X0[c] = lo;
        X1[c] = hi;
        for (int32_t i = 0; i < n; i++) u[i] -= lo;
    }
    return 0;
}

/* undoID, src/quant.c:553-587 */
int orc_id_dequantize(const uint64_t *q, int32_t n, uint64_t width,
                      const uint64_t X0[3], uint64_t *ids) {
    if (n <= 0) return -1;
    for (int32_t i = 0; i < n; i++) {
        uint64_t x = q[i] + X0[0];
        if (x >= width) x -= width;
        uint64_t y = q[(size_t)n + i] + X0[1];
        if (y >= width) y -= width;
        uint64_t z = q[2 * (size_t)n + i] + X0[2];
        if (z >= width) z -= width;
        ids[i] = x + width * y + width * width * z;
    }
    return 0;
}

/* uint() / undoUint(), src/quant.c:373-398, 589-608 */
int orc_uint_quantize(const uint64_t *data, int32_t n, uint64_t *x0,
                      uint64_t *x1, uint64_t *q) {
    if (n <= 0) return -1;
    uint64_t lo = data[0], hi = data[0];
    for (int32_t i = 1; i < n; i++) {
        if (data[i] > hi) hi = data[i];
        if (data[i] < lo) lo = data[i];
    }
    for (int32_t i = 0; i < n; i++) q[i] = data[i] - lo;
    *x0 = lo;
    *x1 = hi;
    return 0;
}
int orc_uint_dequantize(const uint64_t *q, int32_t n, uint64_t x0, uint64_t *out) {
    if (n <= 0) return -1;
    for (int32_t i = 0; i < n; i++) out[i] = q[i] + x0;
    return 0;
}
#endif

static void put_u32(float *slot, uint32_t v) { memcpy(slot, &v, 4); }
static uint32_t get_u32(const float *slot) { uint32_t v; memcpy(&v, slot, 4); return v; }

/* Particle IDs as a Diff v0.9 payload: id() then the same delta + pack as the
 * float fields. The header's Width / X0 / X1 slots hold the IDQuantization
 * (src/types.h:97-99) as uint32 bit patterns; Depth = bit length of the widest
 * component's range (at most 24, else -1). */
int64_t orc_diff_encode_id_field(const uint64_t *ids, int32_t n, uint64_t width,
                                 uint8_t *out, int64_t cap) {
    OrcFieldHeader hd;
    memset(&hd, 0, sizeof(hd));
    hd.Magic = ORC_ALGO_DIFF;
    hd.Version = ORC_VERSION_0_9;
    hd.FieldCode = ORC_FIELD_PTID;
    hd.ParticleLen = n;
    hd.BlockLen = ORC_BLOCK_LEN;
    hd.DitherRun = ORC_DITHER_RUN;
    hd.DitherLanes = ORC_DITHER_LANES;
    if (width == 0 || width >= ((uint64_t)1 << 31)) return -3;
    put_u32(&hd.Width, (uint32_t)width);
    uint64_t *q = malloc((size_t)(n > 0 ? n : 1) * 3 * sizeof(uint64_t));
    if (n > 0) {
        uint64_t X0[3], X1[3], range = 0;
        orc_id_quantize(ids, n, width, X0, X1, q);
        for (int c = 0; c < 3; c++) {
            if (X1[c] >= ((uint64_t)1 << 32)) { free(q); return -1; }
            put_u32(&hd.X0[c], (uint32_t)X0[c]);
            put_u32(&hd.X1[c], (uint32_t)X1[c]);
            if (X1[c] - X0[c] > range) range = X1[c] - X0[c];
        }
        if (range >= ((uint64_t)1 << 24)) { free(q); return -1; }
        hd.Depth = (uint8_t)bit_length((uint32_t)range);
    }
    int64_t bytes = orc_diff_encode_q(&hd, q, out, cap);
    free(q);
    return bytes;
}

int orc_diff_decode_id_field(const uint8_t *in, int64_t len, uint64_t *ids) {
    if (len < (int64_t)sizeof(OrcFieldHeader)) return -1;
    OrcFieldHeader hd;
    memcpy(&hd, in, sizeof(hd));
    if (hd.FieldCode != ORC_FIELD_PTID) return -3;
    int32_t n = hd.ParticleLen;
    if (n <= 0) return n == 0 ? 0 : -1;
    uint64_t *q = malloc((size_t)n * 3 * sizeof(uint64_t));
    int rc = orc_diff_decode_q(in, len, q);
    if (rc == 0) {
        uint64_t X0[3] = {get_u32(&hd.X0[0]), get_u32(&hd.X0[1]), get_u32(&hd.X0[2])};
        rc = orc_id_dequantize(q, n, get_u32(&hd.Width), X0, ids);
        /* a damaged block (orc_seg_from_bytes) loses its particles' ids */
        const OrcBlockEntry *table = (const OrcBlockEntry *)(in + sizeof(hd));
        int64_t nb = orc_diff_nblocks(n);
        for (int64_t g = 0; g < 3 * nb; g++)
            if (table[g].Width == ORC_WIDTH_DAMAGED) {
                int64_t b = g % nb, cnt = n - b * ORC_BLOCK_LEN;
                if (cnt > ORC_BLOCK_LEN) cnt = ORC_BLOCK_LEN;
                for (int64_t j = 0; j < cnt; j++) ids[b * ORC_BLOCK_LEN + j] = UINT64_MAX;
            }
    }
    free(q);
    return rc;
}

/* Unsigned integers: uint() (subtract the minimum), then the N values are coded
 * as three runs of ceil(N/3) (the last run padded with its last value) through
 * the same delta + pack. The header's X0[0..1] / X1[0..1] slots hold the 64-bit
 * minimum and maximum (IntQuantization, src/types.h:77-79); ParticleLen in the
 * payload header is the run length, the true count travels in CField.Hd. */
int64_t orc_diff_encode_uint_field(const uint64_t *data, int32_t n,
                                   uint8_t *out, int64_t cap) {
    OrcFieldHeader hd;
    memset(&hd, 0, sizeof(hd));
    hd.Magic = ORC_ALGO_DIFF;
    hd.Version = ORC_VERSION_0_9;
    hd.FieldCode = ORC_FIELD_UNSI;
    int32_t run = (int32_t)(((int64_t)n + 2) / 3);
    hd.ParticleLen = run;
    hd.BlockLen = ORC_BLOCK_LEN;
    hd.DitherRun = ORC_DITHER_RUN;
    hd.DitherLanes = ORC_DITHER_LANES;
    uint64_t *q = malloc((size_t)(run > 0 ? run : 1) * 3 * sizeof(uint64_t));
    if (n > 0) {
        uint64_t x0, x1;
        orc_uint_quantize(data, n, &x0, &x1, q);
        for (int64_t i = n; i < 3 * (int64_t)run; i++) q[i] = q[n - 1];
        if (x1 - x0 >= ((uint64_t)1 << 24)) { free(q); return -1; }
        put_u32(&hd.X0[0], (uint32_t)x0); put_u32(&hd.X0[1], (uint32_t)(x0 >> 32));
        put_u32(&hd.X1[0], (uint32_t)x1); put_u32(&hd.X1[1], (uint32_t)(x1 >> 32));
        hd.Depth = (uint8_t)bit_length((uint32_t)(x1 - x0));
    }
    int64_t bytes = orc_diff_encode_q(&hd, q, out, cap);
    free(q);
    return bytes;
}

int orc_diff_decode_uint_field(const uint8_t *in, int64_t len, int32_t n, uint64_t *out) {
    if (len < (int64_t)sizeof(OrcFieldHeader)) return -1;
    OrcFieldHeader hd;
    memcpy(&hd, in, sizeof(hd));
    if (hd.FieldCode != ORC_FIELD_UNSI) return -3;
    int32_t run = hd.ParticleLen;
    if (n < 0 || (int64_t)n > 3 * (int64_t)run || ((int64_t)n + 2) / 3 != run) return -1;
    if (n == 0) return 0;
    uint64_t *q = malloc((size_t)run * 3 * sizeof(uint64_t));
    int rc = orc_diff_decode_q(in, len, q);
    if (rc == 0) {
        uint64_t x0 = (uint64_t)get_u32(&hd.X0[0]) | ((uint64_t)get_u32(&hd.X0[1]) << 32);
        rc = orc_uint_dequantize(q, n, x0, out);
        const OrcBlockEntry *table = (const OrcBlockEntry *)(in + sizeof(hd));
        int64_t nb = orc_diff_nblocks(run);
        for (int64_t g = 0; g < 3 * nb; g++)
            if (table[g].Width == ORC_WIDTH_DAMAGED) {
                int64_t b = g % nb, c = g / nb, cnt = run - b * ORC_BLOCK_LEN;
                if (cnt > ORC_BLOCK_LEN) cnt = ORC_BLOCK_LEN;
                for (int64_t j = 0; j < cnt; j++) {
                    int64_t i = c * run + b * ORC_BLOCK_LEN + j;
                    if (i < n) out[i] = UINT64_MAX;
                }
            }
    }
    free(q);
    return rc;
}

/* ------------------------------------------------------------------ */
/* segment image (doc/header_format/header_format.tex:129-204)          */
/* ------------------------------------------------------------------ */

/* blocks of one Diff v0.9 field: block 0 = header + block table, then one per
 * packed block in table order; lengths are multiples of 16 (so of 8, tex:198-204) */
static int64_t field_run_len(const OrcSegField *f) {
    if (f->FieldCode == ORC_FIELD_UNSI) return ((int64_t)f->ParticleLen + 2) / 3;
    return f->ParticleLen;
}
int64_t orc_seg_field_blocks(const OrcSegField *f) {
    return 1 + 3 * orc_diff_nblocks((int32_t)field_run_len(f));
}
static int64_t block_len_of(const OrcSegField *f, int64_t k) {
    int32_t run = (int32_t)field_run_len(f);
    int64_t nb = orc_diff_nblocks(run);
    if (k == 0) return orc_diff_fixed_bytes(run);
    const OrcBlockEntry *table = (const OrcBlockEntry *)(f->Data + sizeof(OrcFieldHeader));
    int64_t g = k - 1, b = g % nb;
    int64_t cnt = run - b * ORC_BLOCK_LEN;
    if (cnt > ORC_BLOCK_LEN) cnt = ORC_BLOCK_LEN;
    return (orc_pack_words(cnt, (uint8_t)table[g].Width) * 4 + 15) / 16 * 16;
}

int64_t orc_seg_bytes_bound(const OrcSegField *fields, int32_t nFields) {
    int64_t total = 16 + 16 * (int64_t)nFields;
    for (int32_t i = 0; i < nFields; i++) total += 8 * orc_seg_field_blocks(&fields[i]) + fields[i].DataLen;
    return total;
}

/* SegmentHeader | FieldHeader[F] | BlockHeader[B] | blocks. FieldHeader is the
 * code's struct (src/types.h:103-108, 4th word ParticleLen: what src/funcs.c:85
 * serialises), BlockHeader {int32 Length; uint32 Checksum} (tex:182-185); the
 * checksum is util_Checksum (start value 1, as the code has it, not the 0xff of
 * tex:381-389); SegmentHeader.Checksum covers the 12 + 16F + 8B bytes behind it
 * (tex:150-158). */
int64_t orc_seg_to_bytes(const OrcSegField *fields, int32_t nFields,
                         int32_t particleNum, uint8_t *out, int64_t cap) {
    int64_t nBlocks = 0;
    for (int32_t i = 0; i < nFields; i++) nBlocks += orc_seg_field_blocks(&fields[i]);
    if (nBlocks >= ((int64_t)1 << 31)) return -5;
    int64_t hdrBytes = 16 + 16 * (int64_t)nFields + 8 * nBlocks;
    if (orc_seg_bytes_bound(fields, nFields) > cap) return -2;
    int32_t *sh = (int32_t *)out;
    sh[1] = (int32_t)nBlocks;
    sh[2] = nFields;
    sh[3] = particleNum;
    uint8_t *fh = out + 16, *bh = fh + 16 * (int64_t)nFields, *dst = out + hdrBytes;
    for (int32_t i = 0; i < nFields; i++) {
        const OrcSegField *f = &fields[i];
        uint32_t w[4] = {f->FieldCode, f->AlgoCode, f->AlgoVersion, (uint32_t)f->ParticleLen};
        memcpy(fh + 16 * (int64_t)i, w, 16);
        int64_t nbk = orc_seg_field_blocks(f), off = 0;
        for (int64_t k = 0; k < nbk; k++) {
            int64_t len = block_len_of(f, k);
            if (off + len > f->DataLen) return -1;
            int32_t len32 = (int32_t)len;
            uint32_t sum = orc_checksum(f->Data + off, len);
            memcpy(bh, &len32, 4);
            memcpy(bh + 4, &sum, 4);
            bh += 8;
            off += len;
        }
        if (off != f->DataLen) return -1;
        memcpy(dst, f->Data, (size_t)f->DataLen);
        dst += f->DataLen;
    }
    uint32_t top = orc_checksum(out + 4, hdrBytes - 4);
    memcpy(out, &top, 4);
    return dst - out;
}

/* Inverse. fields[i].Data points into `bytes` (which is modified: the table
 * entry of a block whose checksum fails gets Width = ORC_WIDTH_DAMAGED, so that
 * only that block's particles are lost, tex:188-196); valid[i] = 0 when the
 * field's block 0 (header + table) is damaged. Returns the number of fields,
 * <0 when the headers themselves are damaged or inconsistent. */
int32_t orc_seg_from_bytes(uint8_t *bytes, int64_t len, OrcSegField *fields,
                           int32_t maxFields, int32_t *valid, int32_t *particleNum) {
    if (len < 16) return -1;
    int32_t sh[4];
    memcpy(sh, bytes, 16);
    int64_t nBlocks = sh[1];
    int32_t nFields = sh[2];
    if (nFields < 0 || nBlocks < 0 || nFields > maxFields) return -1;
    int64_t hdrBytes = 16 + 16 * (int64_t)nFields + 8 * nBlocks;
    if (hdrBytes > len) return -1;
    if (orc_checksum(bytes + 4, hdrBytes - 4) != (uint32_t)sh[0]) return -4;
    if (particleNum) *particleNum = sh[3];
    const uint8_t *fh = bytes + 16, *bh = fh + 16 * (int64_t)nFields;
    uint8_t *src = bytes + hdrBytes;
    int64_t seen = 0;
    for (int32_t i = 0; i < nFields; i++) {
        uint32_t w[4];
        memcpy(w, fh + 16 * (int64_t)i, 16);
        OrcSegField *f = &fields[i];
        f->FieldCode = w[0]; f->AlgoCode = w[1]; f->AlgoVersion = w[2]; f->ParticleLen = (int32_t)w[3];
        f->Data = src;
        int64_t nbk = orc_seg_field_blocks(f), off = 0;
        if (seen + nbk > nBlocks) return -1;
        valid[i] = 1;
        int32_t run = (int32_t)field_run_len(f);
        for (int64_t k = 0; k < nbk; k++) {
            int32_t len32;
            uint32_t sum;
            memcpy(&len32, bh, 4);
            memcpy(&sum, bh + 4, 4);
            bh += 8;
            if (len32 < 0 || src + off + len32 > bytes + len) return -1;
            if (k == 0 && len32 != orc_diff_fixed_bytes(run)) return -1;
            if (orc_checksum(src + off, len32) != sum) {
                if (k == 0) valid[i] = 0;
                else if (valid[i]) {
                    OrcBlockEntry *table = (OrcBlockEntry *)(src + sizeof(OrcFieldHeader));
                    table[k - 1].Width = ORC_WIDTH_DAMAGED;
                }
            }
            off += len32;
        }
        f->DataLen = off;
        src += off;
        seen += nbk;
    }
    return seen == nBlocks ? nFields : -1;
}

/* ------------------------------------------------------------------ */
/* CPU-baseline drivers: independent fields spread over host threads    */
/* ------------------------------------------------------------------ */

typedef struct BenchJob {
    int tid, threads, decode;
    uint32_t fieldCode;
    const float *data;
    const uint8_t *const *payloads;
    const int64_t *lens;
    int32_t n;
    int64_t nFields;
    float delta, width;
    int64_t bytes;
} BenchJob;

static void *bench_worker(void *arg) {
    BenchJob *j = arg;
    size_t m = (size_t)j->n * 3;
    if (j->decode) {
        float *out = malloc(m * sizeof(float));
        for (int64_t f = j->tid; f < j->nFields; f += j->threads)
            orc_diff_decode_field(j->payloads[f], j->lens[f], out, NULL);
        free(out);
    } else {
        int64_t cap = orc_diff_bound(j->n);
        uint8_t *out = malloc((size_t)cap);
        for (int64_t f = j->tid; f < j->nFields; f += j->threads) {
            int64_t b = orc_diff_encode_field(j->fieldCode, j->data + f * m,
                                              j->n, j->delta, j->width, 0,
                                              out, cap, NULL);
            if (b > 0) j->bytes += b;
        }
        free(out);
    }
    return NULL;
}

static double run_jobs(BenchJob *proto, int threads, int64_t *bytesOut) {
    if (threads < 1) threads = 1;
    pthread_t *th = malloc((size_t)threads * sizeof(*th));
    BenchJob *jobs = malloc((size_t)threads * sizeof(*jobs));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < threads; t++) {
        jobs[t] = *proto;
        jobs[t].tid = t;
        jobs[t].threads = threads;
        jobs[t].bytes = 0;
        pthread_create(&th[t], NULL, bench_worker, &jobs[t]);
    }
    int64_t total = 0;
    for (int t = 0; t < threads; t++) {
        pthread_join(th[t], NULL);
        total += jobs[t].bytes;
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (bytesOut) *bytesOut = total;
    free(th);
    free(jobs);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

/* encode nFields fields laid out back to back (3n floats each); seconds */
double orc_bench_encode(uint32_t fieldCode, const float *data, int32_t n,
                        int64_t nFields, float delta, float width,
                        int threads, int64_t *bytesOut) {
    BenchJob p;
    memset(&p, 0, sizeof(p));
    p.fieldCode = fieldCode; p.data = data; p.n = n; p.nFields = nFields;
    p.delta = delta; p.width = width;
    return run_jobs(&p, threads, bytesOut);
}

double orc_bench_decode(const uint8_t *const *payloads, const int64_t *lens,
                        int32_t n, int64_t nFields, int threads) {
    BenchJob p;
    memset(&p, 0, sizeof(p));
    p.decode = 1; p.payloads = payloads; p.lens = lens; p.n = n;
    p.nFields = nFields;
    return run_jobs(&p, threads, NULL);
}

/* ------------------------------------------------------------------ */
/* bench input on the host: the same synthetic snapshot the device       */
/* generator makes (64 sine modes drawn by splitmix64, amplitude          */
/* ~ |k|^-2 along k, rms displacement 1.5 cells, v = 200 psi / cell,      */
/* optional white noise), so that the --impl reference arm of bench.py    */
/* never has to load the CUDA library. Same formulas, host libm: the      */
/* floats agree with the device's to rounding, not bit for bit (nothing   */
/* compares them). Nothing in the reference corresponds to this.          */
/* ------------------------------------------------------------------ */
#define ORC_GEN_MODES 64

typedef struct GenJob {
    int tid, threads, ngrid;
    float L, velPerLen, noise;
    uint64_t noiseSeed;
    int64_t firstId, count;
    const int *kx, *ky, *kz;
    const float *phase, *ax, *ay, *az;
    float *pos, *vel;
} GenJob;

static float gen_noise(uint64_t seed, int64_t id, int c) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ULL * (uint64_t)(3 * id + c + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return (float)(z >> 40) * (1.0f / 8388608.0f) - 1.0f;
}

static void *gen_worker(void *arg) {
    GenJob *j = arg;
    const int N = j->ngrid;
    const float piOverN = 3.14159265358979f / (float)N;
    const float cell = j->L / (float)N;
    int64_t lo = j->count * j->tid / j->threads, hi = j->count * (j->tid + 1) / j->threads;
    for (int64_t i = lo; i < hi; i++) {
        int64_t id = j->firstId + i;
        int idx[3] = {(int)(id % N), (int)((id / N) % N), (int)(id / ((int64_t)N * N))};
        float psi[3] = {0, 0, 0};
        for (int m = 0; m < ORC_GEN_MODES; m++) {
            int t = j->kx[m] * (2 * idx[0] + 1) + j->ky[m] * (2 * idx[1] + 1) + j->kz[m] * (2 * idx[2] + 1);
            t %= 2 * N;
            float sn = sinf((float)t * piOverN + j->phase[m]);
            psi[0] = fmaf(j->ax[m], sn, psi[0]);
            psi[1] = fmaf(j->ay[m], sn, psi[1]);
            psi[2] = fmaf(j->az[m], sn, psi[2]);
        }
        for (int c = 0; c < 3; c++) {
            if (j->noise > 0) psi[c] += j->noise * gen_noise(j->noiseSeed, id, c);
            if (j->pos) {
                float x = ((float)idx[c] + 0.5f) * cell + psi[c];
                x -= j->L * floorf(x / j->L);
                if (!(x >= 0.0f && x < j->L)) x = 0.0f;
                j->pos[(int64_t)c * j->count + i] = x;
            }
            if (j->vel) j->vel[(int64_t)c * j->count + i] = j->velPerLen * psi[c];
        }
    }
    return NULL;
}

int orc_gen_zeldovich(int32_t ngrid, float L, uint64_t seed, int64_t firstId,
                      int64_t count, float noiseCells, float *pos, float *vel,
                      int threads) {
    if (ngrid <= 0 || count < 0) return -1;
    int kx[ORC_GEN_MODES], ky[ORC_GEN_MODES], kz[ORC_GEN_MODES];
    float phase[ORC_GEN_MODES], fx[ORC_GEN_MODES], fy[ORC_GEN_MODES], fz[ORC_GEN_MODES];
    double ax[ORC_GEN_MODES], ay[ORC_GEN_MODES], az[ORC_GEN_MODES], sumsq = 0.0;
    uint64_t s = seed;
    for (int i = 0; i < ORC_GEN_MODES; i++) {
        do {
            kx[i] = (int)(splitmix64(&s) % 17) - 8;
            ky[i] = (int)(splitmix64(&s) % 17) - 8;
            kz[i] = (int)(splitmix64(&s) % 17) - 8;
        } while (kx[i] == 0 && ky[i] == 0 && kz[i] == 0);
        double k2 = (double)kx[i] * kx[i] + (double)ky[i] * ky[i] + (double)kz[i] * kz[i];
        double a = 1.0 / k2, inv = 1.0 / sqrt(k2);
        phase[i] = (float)((double)(splitmix64(&s) >> 11) * (1.0 / 9007199254740992.0) * 6.283185307179586);
        ax[i] = a * kx[i] * inv; ay[i] = a * ky[i] * inv; az[i] = a * kz[i] * inv;
        sumsq += a * a;
    }
    const double cell = (double)L / ngrid;
    const double scale = 1.5 * cell / sqrt(sumsq / 2.0);
    for (int i = 0; i < ORC_GEN_MODES; i++) {
        fx[i] = (float)(ax[i] * scale); fy[i] = (float)(ay[i] * scale); fz[i] = (float)(az[i] * scale);
    }
    if (threads < 1) threads = 1;
    pthread_t *th = malloc((size_t)threads * sizeof(*th));
    GenJob *jobs = malloc((size_t)threads * sizeof(*jobs));
    for (int t = 0; t < threads; t++) {
        GenJob g = {t, threads, ngrid, L, (float)(200.0 / cell), (float)(noiseCells * cell),
                    seed ^ 0x6d6e77526f756768ULL, firstId, count, kx, ky, kz, phase, fx, fy, fz, pos, vel};
        jobs[t] = g;
        pthread_create(&th[t], NULL, gen_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
    return 0;
}

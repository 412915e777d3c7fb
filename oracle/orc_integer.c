/* orc_integer.c -- bit-exact integer stages of the KS+PBS path (TEST INFRASTRUCTURE, see tfhe_oracle.h). */
#include "tfhe_oracle.h"
#include <string.h>
#include <stdlib.h>

static inline uint64_t maskw(int bits) { return bits == 64 ? ~0ULL : ((1ULL << bits) - 1); }

/* commons/math/decomposition/decomposer.rs:99-119 */
uint64_t orc_closest_representable(uint64_t x, int base_log, int level, int bits) {
    uint64_t m = maskw(bits);
    int non_rep = bits - base_log * level;
    int shift = non_rep - 1;
    uint64_t res = (x & m) >> shift;
    res += 1;
    res &= ~1ULL;
    return (res << shift) & m;
}

/* commons/math/decomposition/iter.rs:120-127 (identical copy: fft_impl/fft64/math/decomposition.rs:79-86) */
static inline uint64_t decompose_one_level(int base_log, uint64_t *state, uint64_t mod_b_mask, uint64_t m) {
    uint64_t res = *state & mod_b_mask;
    *state >>= base_log;
    uint64_t carry = (((res - 1) & m) | *state) & res;
    carry >>= (base_log - 1);
    *state = (*state + carry) & m;
    return (res - (carry << base_log)) & m;
}

/* iter.rs:37-50: state = closest_representable(x) >> (BITS - l*b); terms come out level l first */
void orc_decompose(uint64_t x, int base_log, int level, int bits, uint64_t *out) {
    uint64_t m = maskw(bits);
    uint64_t state = orc_closest_representable(x, base_log, level, bits) >> (bits - base_log * level);
    uint64_t mb = (1ULL << base_log) - 1;
    for (int t = 0; t < level; t++) out[t] = decompose_one_level(base_log, &state, mb, m);
}

/* fft_impl/common.rs:26-43 */
uint64_t orc_modswitch(uint64_t x, int log2N, int bits) {
    uint64_t out = (x & maskw(bits)) >> (bits - log2N - 2);
    out += 1;
    out >>= 1;
    return out;
}

/* algorithms/polynomial_algorithms.rs:315-354.  out = in * X^{-deg} in Z[X]/(X^N+1). */
void orc_monomial_div(uint64_t *out, const uint64_t *in, int N, uint64_t deg, int bits) {
    uint64_t m = maskw(bits);
    uint64_t full = deg / (uint64_t)N;
    int r = (int)(deg % (uint64_t)N);
    int flip = (full % 2) == 1;
    /* out[j] = in[j+r] for j < N-r ; out[j] = -in[j+r-N] for j >= N-r ; all negated when flip */
    for (int j = 0; j < N; j++) {
        uint64_t v;
        if (j < N - r) v = in[j + r]; else v = (0 - in[j + r - N]);
        if (flip) v = 0 - v;
        out[j] = v & m;
    }
}

/* algorithms/polynomial_algorithms.rs:425-490.  out = in * X^{deg} - in. */
void orc_monomial_mul_sub(uint64_t *out, const uint64_t *in, int N, uint64_t deg, int bits) {
    uint64_t m = maskw(bits);
    uint64_t full = deg / (uint64_t)N;
    int r = (int)(deg % (uint64_t)N);
    int flip = (full % 2) == 1;
    /* two branch-free (vectorisable) ranges: the wrapped-around head is negated, the tail is not; `flip` negates both */
    const uint64_t sh = flip ? 0 : ~0ULL, st = flip ? ~0ULL : 0;      /* v -> (v ^ s) - s negates when s is all-ones */
    const uint64_t *head = in + (N - r);
    for (int j = 0; j < r; j++) out[j] = (((head[j] ^ sh) - sh) - in[j]) & m;
    for (int j = r; j < N; j++) out[j] = (((in[j - r] ^ st) - st) - in[j]) & m;
}

/* algorithms/glwe_sample_extraction.rs:91-147 with nth = 0 */
void orc_sample_extract0(uint64_t *lwe, const uint64_t *glwe, int k, int N, int bits) {
    uint64_t m = maskw(bits);
    for (int p = 0; p < k; p++) {
        const uint64_t *poly = glwe + (size_t)p * N;
        uint64_t *o = lwe + (size_t)p * N;
        o[0] = poly[0] & m;
        for (int t = 1; t < N; t++) o[t] = (0 - poly[N - t]) & m;
    }
    lwe[(size_t)k * N] = glwe[(size_t)k * N] & m;
}

/* algorithms/lwe_keyswitch.rs:143-169 ; inner op slice_algorithms.rs:363-459 (out -= digit * level_ct).
 * KSK layout [kN][l_ks][n+1], level blocks stored in decomposition-iterator order (level l_ks first),
 * lwe_keyswitch_key_generation.rs:109-120. */
void orc_keyswitch(const orc_params *p, const uint64_t *ksk, const uint64_t *in, uint64_t *out) {
    uint64_t m = maskw(p->bits);
    int n_in = p->k * p->N, n_out = p->n, L = p->ks_level;
    memset(out, 0, sizeof(uint64_t) * (n_out + 1));
    out[n_out] = in[n_in] & m;
    uint64_t digits[64];
    for (int i = 0; i < n_in; i++) {
        orc_decompose(in[i], p->ks_base_log, L, p->bits, digits);
        const uint64_t *blk = ksk + (size_t)i * L * (n_out + 1);
        for (int t = 0; t < L; t++) {
            uint64_t d = digits[t];
            if (d == 0) continue;
            const uint64_t *row = blk + (size_t)t * (n_out + 1);
            for (int j = 0; j <= n_out; j++) out[j] -= d * row[j];
        }
    }
    for (int j = 0; j <= n_out; j++) out[j] &= m;
}

/* shortint/engine/mod.rs:72-128 */
void orc_fill_accumulator(const orc_params *p, const uint64_t *f, int pmod, uint64_t *glwe) {
    int N = p->N, k = p->k;
    uint64_t m = maskw(p->bits);
    memset(glwe, 0, sizeof(uint64_t) * (size_t)(k + 1) * N);
    uint64_t *body = glwe + (size_t)k * N;
    uint64_t delta = (1ULL << (p->bits - 1)) / (uint64_t)pmod;
    int box = N / pmod;
    for (int i = 0; i < pmod; i++)
        for (int j = 0; j < box; j++) body[i * box + j] = (f[i] * delta) & m;
    int half = box / 2;
    for (int j = 0; j < half; j++) body[j] = (0 - body[j]) & m;
    /* rotate_left(half) */
    uint64_t *tmp = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (int j = 0; j < N; j++) tmp[j] = body[(j + half) % N];
    memcpy(body, tmp, sizeof(uint64_t) * N);
    free(tmp);
}

/* boolean/engine/bootstrapping.rs:59-60 ; boolean/mod.rs:74-80 (PLAINTEXT_TRUE = 1<<29 on u32) */
void orc_boolean_lut(const orc_params *p, uint64_t *glwe) {
    int N = p->N, k = p->k;
    memset(glwe, 0, sizeof(uint64_t) * (size_t)(k + 1) * N);
    for (int j = 0; j < N; j++) glwe[(size_t)k * N + j] = 1ULL << (p->bits - 3);
}

/* boolean/engine/mod.rs : AND :620-640, NAND :646-682, OR/NOR/XOR/XNOR analogues (u32 wrapping) */
void orc_boolean_gate_prologue(int gate, int len, const uint64_t *a, const uint64_t *b, uint64_t *out) {
    const uint64_t m = 0xFFFFFFFFULL, T = 1ULL << 29, F = 7ULL << 29, Q = 1ULL << 30; /* 1/8, -1/8, 1/4 */
    int body = len - 1;
    for (int j = 0; j < len; j++) {
        uint64_t s = a[j] + b[j], v;
        uint64_t cst;
        switch (gate) {
        case 0: v = s; cst = F; break;                 /* AND  : a+b-1/8 */
        case 1: v = 0 - s; cst = T; break;             /* NAND : -(a+b)+1/8 */
        case 2: v = s; cst = T; break;                 /* OR   : a+b+1/8 */
        case 3: v = 0 - s; cst = F; break;             /* NOR  : -(a+b)-1/8 */
        case 4: v = 2 * s; cst = Q; break;             /* XOR  : 2(a+b)+1/4 */
        default: v = 0 - 2 * s; cst = 0 - Q; break;    /* XNOR : -2(a+b)-1/4 */
        }
        if (j == body) v += cst;
        out[j] = v & m;
    }
}

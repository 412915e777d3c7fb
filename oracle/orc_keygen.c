/* orc_keygen.c -- key and ciphertext synthesis for the oracle (TEST INFRASTRUCTURE, see tfhe_oracle.h).
 * The reference draws these from concrete-csprng (AES-CTR); stream equality is not needed because
 * keys and inputs are exported as bytes and consumed by both the oracle and the CUDA path.
 * PRNG: xoshiro256** seeded through splitmix64; Gaussian by Box-Muller rounded to the torus like
 * commons/math/torus/mod.rs:72-78 (from_torus). */
#include "tfhe_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

void orc_parallel_for(void (*fn)(void *, int), void *ctx, int count, int nthreads);

typedef struct { uint64_t s[4]; } rng_t;
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static uint64_t splitmix(uint64_t *x) {
    uint64_t z = (*x += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
static void rng_seed(rng_t *r, uint64_t seed, uint64_t stream) {
    uint64_t x = seed ^ (stream * 0xD1342543DE82EF95ULL + 0x632BE59BD9B4E019ULL);
    for (int i = 0; i < 4; i++) r->s[i] = splitmix(&x);
}
static inline uint64_t rng_next(rng_t *r) {
    uint64_t *s = r->s, res = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return res;
}
static inline double rng_unit(rng_t *r) { return ((double)(rng_next(r) >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
static double rng_gauss(rng_t *r, double std) {
    double u1 = rng_unit(r), u2 = rng_unit(r);
    return std * sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925 * u2);
}
static inline uint64_t maskw(int bits) { return bits == 64 ? ~0ULL : ((1ULL << bits) - 1); }
static inline uint64_t torus_noise(rng_t *r, double std, int bits) {
    double t = rng_gauss(r, std);
    double fr = t - round(t);
    return (uint64_t)(int64_t)round(fr * ldexp(1.0, bits)) & maskw(bits);
}

/* LweSecretKey::generate_new_binary (used e.g. fft_impl/common.rs:113-123) */
void orc_gen_binary_key(uint64_t seed, int len, uint64_t *key) {
    rng_t r; rng_seed(&r, seed, 0x5EC2E7);
    for (int i = 0; i < len; i++) key[i] = rng_next(&r) >> 63;
}

/* algorithms/lwe_encryption.rs : b = <a,s> + plaintext + e ; layout mask then body (entities/lwe_ciphertext.rs:598-609) */
static void lwe_encrypt_rng(int dim, int bits, const uint64_t *key, uint64_t pt, double std, rng_t *r, uint64_t *ct) {
    uint64_t m = maskw(bits), b = 0;
    for (int i = 0; i < dim; i++) { ct[i] = rng_next(r) & m; if (key[i]) b += ct[i]; }
    ct[dim] = (b + pt + torus_noise(r, std, bits)) & m;
}
void orc_lwe_encrypt(int dim, int bits, const uint64_t *key, uint64_t pt, double std, uint64_t seed, uint64_t *ct) {
    rng_t r; rng_seed(&r, seed, 0xC7);
    lwe_encrypt_rng(dim, bits, key, pt, std, &r, ct);
}
uint64_t orc_lwe_phase(int dim, int bits, const uint64_t *key, const uint64_t *ct) {
    uint64_t b = ct[dim];
    for (int i = 0; i < dim; i++) b -= ct[i] * key[i];
    return b & maskw(bits);
}

/* algorithms/lwe_keyswitch_key_generation.rs:103-129.  Block i, stored slot t (t = 0 is level l_ks):
 * LWE encryption under the small key of  s_in[i] * 2^{bits - b*(l_ks - t)}  (to_recomposition_summand). */
typedef struct { const orc_params *p; const uint64_t *big, *small; uint64_t seed; uint64_t *ksk; } ksk_ctx;
static void ksk_one(void *c, int i) {
    ksk_ctx *x = (ksk_ctx *)c; const orc_params *p = x->p;
    rng_t r; rng_seed(&r, x->seed, 0x1000000ULL + (uint64_t)i);
    int L = p->ks_level, n = p->n;
    for (int t = 0; t < L; t++) {
        int level = L - t;
        uint64_t pt = (x->big[i] << (p->bits - p->ks_base_log * level)) & maskw(p->bits);
        lwe_encrypt_rng(n, p->bits, x->small, pt, p->lwe_std, &r, x->ksk + ((size_t)i * L + t) * (n + 1));
    }
}
void orc_gen_ksk(const orc_params *p, const uint64_t *big_key, const uint64_t *small_key, uint64_t seed, uint64_t *ksk, int nthreads) {
    ksk_ctx c = { p, big_key, small_key, seed, ksk };
    orc_parallel_for(ksk_one, &c, p->k * p->N, nthreads);
}

/* algorithms/glwe_encryption.rs (encrypt_glwe_ciphertext_assign): body += sum_j A_j*S_j + E ; mask uniform */
static void glwe_encrypt_assign(const orc_params *p, const uint64_t *glwe_key, uint64_t *row, rng_t *r, uint64_t *scratch) {
    int N = p->N, k = p->k; uint64_t m = maskw(p->bits);
    uint64_t *body = row + (size_t)k * N;
    for (int j = 0; j < k; j++) {
        uint64_t *A = row + (size_t)j * N;
        for (int t = 0; t < N; t++) A[t] = rng_next(r) & m;
        orc_negacyclic_mul_binary(N, A, glwe_key + (size_t)j * N, scratch, p->bits);
        for (int t = 0; t < N; t++) body[t] += scratch[t];
    }
    for (int t = 0; t < N; t++) body[t] = (body[t] + torus_noise(r, p->glwe_std, p->bits)) & m;
}

/* ggsw_encryption.rs:100-150,300-331 (encrypt_constant_ggsw_ciphertext): GGSW of the constant `msg`,
 * layout [level 1..l][row 0..k][poly 0..k][N] (entities/ggsw_ciphertext.rs:185-197). */
static void ggsw_encrypt_constant(const orc_params *p, const uint64_t *glwe_key, uint64_t msg, rng_t *r, uint64_t *ggsw, uint64_t *scratch) {
    int N = p->N, k = p->k, k1 = k + 1, L = p->pbs_level; uint64_t m = maskw(p->bits);
    for (int lvl = 1; lvl <= L; lvl++) {
        uint64_t factor = ((0 - msg) << (p->bits - p->pbs_base_log * lvl)) & m; /* ggsw_encryption.rs:121-125 */
        for (int row = 0; row < k1; row++) {
            uint64_t *g = ggsw + ((size_t)(lvl - 1) * k1 + row) * k1 * N;
            uint64_t *body = g + (size_t)k * N;
            if (row < k) for (int t = 0; t < N; t++) body[t] = (glwe_key[(size_t)row * N + t] * factor) & m;
            else { memset(body, 0, sizeof(uint64_t) * N); body[0] = (0 - factor) & m; }
            glwe_encrypt_assign(p, glwe_key, g, r, scratch);
        }
    }
}

/* algorithms/lwe_bootstrap_key_generation.rs:113-133.
 * Layout [n][level 1..l][row 0..k][poly 0..k][N] (entities/lwe_bootstrap_key.rs:99-106). */
typedef struct { const orc_params *p; const uint64_t *small, *glwe_key; uint64_t seed; uint64_t *bsk; int grouping; } bsk_ctx;
static void bsk_one(void *c, int i) {
    bsk_ctx *x = (bsk_ctx *)c; const orc_params *p = x->p;
    int N = p->N, k1 = p->k + 1, L = p->pbs_level;
    rng_t r; rng_seed(&r, x->seed, 0x2000000ULL + (uint64_t)i);
    uint64_t *scratch = (uint64_t *)malloc(sizeof(uint64_t) * N);
    ggsw_encrypt_constant(p, x->glwe_key, x->small[i], &r, x->bsk + (size_t)i * L * k1 * k1 * N, scratch);
    free(scratch);
}
void orc_gen_bsk(const orc_params *p, const uint64_t *small_key, const uint64_t *glwe_key, uint64_t seed, uint64_t *bsk, int nthreads) {
    bsk_ctx c = { p, small_key, glwe_key, seed, bsk, 0 };
    orc_parallel_for(bsk_one, &c, p->n, nthreads);
}

/* algorithms/lwe_multi_bit_bootstrap_key_generation.rs:87-172 + combine_key_bits :401-427.
 * Layout [n/G][2^G][level][row][poly][N] (entities/lwe_multi_bit_bootstrap_key.rs): GGSW `idx` of a group encrypts
 * prod_b (s_b XOR NOT bit_b(idx)), bit position of key element b = G - (b+1): idx 0 is prod (1 - s_b). */
static void mb_bsk_one(void *c, int gi) {
    bsk_ctx *x = (bsk_ctx *)c; const orc_params *p = x->p;
    int N = p->N, k1 = p->k + 1, L = p->pbs_level, G = x->grouping, NG = 1 << G;
    uint64_t *scratch = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (int idx = 0; idx < NG; idx++) {
        rng_t r; rng_seed(&r, x->seed, 0x3000000ULL + (uint64_t)gi * NG + idx);
        uint64_t msg = 1;
        for (int b = 0; b < G; b++) {
            uint64_t inv = (uint64_t)(((idx >> (G - (b + 1))) & 1) ^ 1);
            msg *= (x->small[(size_t)gi * G + b] ^ inv);
        }
        ggsw_encrypt_constant(p, x->glwe_key, msg, &r, x->bsk + ((size_t)gi * NG + idx) * L * k1 * k1 * N, scratch);
    }
    free(scratch);
}
void orc_gen_multibit_bsk(const orc_params *p, int grouping, const uint64_t *small_key, const uint64_t *glwe_key, uint64_t seed,
                          uint64_t *bsk, int nthreads) {
    bsk_ctx c = { p, small_key, glwe_key, seed, bsk, grouping };
    orc_parallel_for(mb_bsk_one, &c, p->n / grouping, nthreads);
}

/* algorithms/lwe_packing_keyswitch_key_generation.rs:60-150.  Block i holds l_ks GLWE ciphertexts (level l_ks first) of the
 * constant polynomial  s_in[i] * 2^{bits - b*level}  under the GLWE key: [kN][l_ks][(k+1)N].  Noise: glwe_std, as the gadget
 * server key passes it (gadget/engine/bootstrapping.rs:355-363). */
typedef struct { const orc_params *p; const uint64_t *big, *glwe_key; uint64_t seed; uint64_t *pksk; } pksk_ctx;
static void pksk_one(void *c, int i) {
    pksk_ctx *x = (pksk_ctx *)c; const orc_params *p = x->p;
    int N = p->N, k = p->k, L = p->ks_level; size_t glwe = (size_t)(k + 1) * N;
    rng_t r; rng_seed(&r, x->seed, 0x4000000ULL + (uint64_t)i);
    uint64_t *scratch = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (int t = 0; t < L; t++) {
        int level = L - t;
        uint64_t *g = x->pksk + ((size_t)i * L + t) * glwe;
        uint64_t *body = g + (size_t)k * N;
        memset(body, 0, sizeof(uint64_t) * N);
        body[0] = (x->big[i] << (p->bits - p->ks_base_log * level)) & maskw(p->bits);
        glwe_encrypt_assign(p, x->glwe_key, g, &r, scratch);
    }
    free(scratch);
}
void orc_gen_pksk(const orc_params *p, const uint64_t *big_key, const uint64_t *glwe_key, uint64_t seed, uint64_t *pksk, int nthreads) {
    pksk_ctx c = { p, big_key, glwe_key, seed, pksk };
    orc_parallel_for(pksk_one, &c, p->k * p->N, nthreads);
}

/* algorithms/lwe_wopbs.rs:80-160 (generate_circuit_bootstrap_lwe_pfpksk_list) over
 * lwe_private_functional_packing_keyswitch_key_generation.rs:27-140: k+1 private functional packing keyswitch keys from the big
 * LWE key (the flattened GLWE key) to the GLWE key, function x -> -x, polynomial = S_r for r < k and the constant -1 for r = k.
 * Block (r, j) holds, for level 1..L, a GLWE encryption of  polynomial_r * (f(1) * bit_j) * 2^(bits - base_log*level),
 * bit_j = big key bit j for j < kN and -1 for the body slot j = kN.  Layout [k+1][kN+1][level 1..L][(k+1)N]. */
typedef struct { const orc_params *p; const uint64_t *glwe_key; uint64_t seed; uint64_t *out; int base_log, level; } pf_ctx;
static void pfpksk_one(void *c, int idx) {
    pf_ctx *x = (pf_ctx *)c; const orc_params *p = x->p;
    int N = p->N, k = p->k, L = x->level, n_in = k * N + 1; size_t glwe = (size_t)(k + 1) * N; uint64_t m = maskw(p->bits);
    int r = idx / n_in, j = idx % n_in;
    rng_t rg; rng_seed(&rg, x->seed, 0x5000000ULL + (uint64_t)idx);
    uint64_t *scratch = (uint64_t *)malloc(sizeof(uint64_t) * N);
    uint64_t key_bit = j < k * N ? x->glwe_key[j] : m;                 /* last_key_iter_bit = Scalar::MAX (:84-90) */
    uint64_t scalar = ((0 - 1ULL) * key_bit) & m;                      /* f(Scalar::ONE).wrapping_mul(input_key_bit) (:119) */
    for (int level = 1; level <= L; level++) {
        uint64_t *g = x->out + (((size_t)r * n_in + j) * L + (level - 1)) * glwe;
        uint64_t *body = g + (size_t)k * N;
        uint64_t summand = (scalar << (p->bits - x->base_log * level)) & m;   /* to_recomposition_summand */
        memset(g, 0, sizeof(uint64_t) * glwe);
        if (r < k) { for (int t = 0; t < N; t++) body[t] = (x->glwe_key[(size_t)r * N + t] * summand) & m; }
        else body[0] = ((0 - summand)) & m;                           /* last polynomial = [Scalar::MAX, 0, ...] (lwe_wopbs.rs:128-136) */
        glwe_encrypt_assign(p, x->glwe_key, g, &rg, scratch);
    }
    free(scratch);
}
void orc_gen_cbs_pfpksk_list(const orc_params *p, int base_log, int level, const uint64_t *glwe_key, uint64_t seed, uint64_t *out,
                             int nthreads) {
    pf_ctx c = { p, glwe_key, seed, out, base_log, level };
    orc_parallel_for(pfpksk_one, &c, (p->k + 1) * (p->k * p->N + 1), nthreads);
}

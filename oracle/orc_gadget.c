/* orc_gadget.c -- CPU restatement of the integer operations the fork's gadget engine adds around the blind rotation
 * (TEST INFRASTRUCTURE, see tfhe_oracle.h).  Citations relative to /root/reference/tfhe/src/. */
#include "tfhe_oracle.h"
#include <stdlib.h>
#include <string.h>

static inline uint64_t maskw(int bits) { return bits == 64 ? ~0ULL : ((1ULL << bits) - 1); }

/* core_crypto/algorithms/lwe_packing_keyswitch.rs:100-187 (keyswitch_lwe_ciphertext_into_glwe_ciphertext).
 * pksk: [kN][l_ks (level l_ks first)][(k+1)N]; in: [kN+1]; out: GLWE [(k+1)N]. */
void orc_packing_keyswitch(const orc_params *p, const uint64_t *pksk, const uint64_t *in, uint64_t *out) {
    uint64_t m = maskw(p->bits);
    int n_in = p->k * p->N, L = p->ks_level;
    size_t glwe = (size_t)(p->k + 1) * p->N;
    memset(out, 0, sizeof(uint64_t) * glwe);
    out[(size_t)p->k * p->N] = in[n_in] & m;                  /* get_mut_body()[0] = input body (:161) */
    uint64_t digits[64];
    for (int i = 0; i < n_in; i++) {
        orc_decompose(in[i], p->ks_base_log, L, p->bits, digits);
        const uint64_t *blk = pksk + (size_t)i * L * glwe;
        for (int t = 0; t < L; t++) {
            uint64_t d = digits[t];
            if (d == 0) continue;
            const uint64_t *row = blk + (size_t)t * glwe;
            for (size_t j = 0; j < glwe; j++) out[j] -= d * row[j];   /* slice_wrapping_sub_scalar_mul_assign (:179-185) */
        }
    }
    for (size_t j = 0; j < glwe; j++) out[j] &= m;
}

/* gadget/engine/bootstrapping.rs:596-611 (mvb_bootstrap): accu_i = v0 * v_i per polynomial
 * (polynomial_karatsuba_wrapping_mul == the negacyclic schoolbook product), then sample extraction of coefficient 0 */
void orc_glwe_mul_plain_extract(const orc_params *p, const uint64_t *glwe, const uint64_t *poly, uint64_t *lwe_out) {
    int N = p->N, k1 = p->k + 1;
    uint64_t *prod = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)k1 * N);
    for (int j = 0; j < k1; j++) orc_negacyclic_schoolbook(N, glwe + (size_t)j * N, poly, prod + (size_t)j * N, p->bits);
    orc_sample_extract0(lwe_out, prod, p->k, N, p->bits);
    free(prod);
}

/* algorithms/polynomial_algorithms.rs (polynomial_wrapping_monic_monomial_mul_assign), degree < N */
static void monomial_mul(uint64_t *out, const uint64_t *in, int N, int deg, uint64_t m) {
    for (int c = 0; c < N; c++) out[c] = (c >= deg) ? in[c - deg] : ((0 - in[N + c - deg]) & m);
}

/* gadget/engine/bootstrapping.rs:690-773 (pack_into_new_accumulator), literal loop structure.
 * lwes: [count][kN+1]; out: GLWE [(k+1)N]. */
void orc_pack_into_accumulator(const orc_params *p, const uint64_t *pksk, const uint64_t *lwes, int count, int pmod, uint64_t *out) {
    uint64_t m = maskw(p->bits);
    int N = p->N, k1 = p->k + 1;
    size_t len = (size_t)p->k * N + 1, glwe = (size_t)k1 * N;
    uint64_t *elems = (uint64_t *)calloc((size_t)pmod * len, sizeof(uint64_t));
    uint64_t *buffer = (uint64_t *)malloc(sizeof(uint64_t) * glwe), *bis = (uint64_t *)malloc(sizeof(uint64_t) * glwe);
    for (int kk = 0; kk < pmod; kk++) {                                   /* reordering accumulator (:699-716) */
        int src = (kk % 2 == 0) ? kk / 2 : (pmod + 1) / 2 + (kk - 1) / 2;
        if (src >= count) continue;                                       /* None => zero ciphertext */
        for (size_t c = 0; c < len; c++) {
            uint64_t v = lwes[(size_t)src * len + c];
            elems[(size_t)kk * len + c] = (kk % 2) ? ((0 - v) & m) : v;   /* lwe_ciphertext_opposite_assign */
        }
    }
    int size_slice = N / pmod;
    memset(out, 0, sizeof(uint64_t) * glwe);
    for (int kk = 0; kk < pmod; kk++) {                                   /* :736-751 */
        orc_packing_keyswitch(p, pksk, elems + (size_t)kk * len, buffer);
        int reps = kk == 0 ? size_slice / 2 : size_slice;
        int offset = kk == 0 ? 0 : size_slice / 2 + (kk - 1) * size_slice;
        for (int i = 0; i < reps; i++) {
            for (int j = 0; j < k1; j++) monomial_mul(bis + (size_t)j * N, buffer + (size_t)j * N, N, offset + i, m);
            for (size_t c = 0; c < glwe; c++) out[c] = (out[c] + bis[c]) & m;
        }
    }
    orc_packing_keyswitch(p, pksk, elems, buffer);                        /* index = p-1 (last one) :753-766 */
    for (size_t c = 0; c < glwe; c++) buffer[c] = (0 - buffer[c]) & m;    /* glwe_ciphertext_opposite_assign */
    for (int i = 0; i < size_slice / 2; i++) {
        for (int j = 0; j < k1; j++) monomial_mul(bis + (size_t)j * N, buffer + (size_t)j * N, N, N - size_slice / 2 + i, m);
        for (size_t c = 0; c < glwe; c++) out[c] = (out[c] + bis[c]) & m;
    }
    free(elems); free(buffer); free(bis);
}

/* orc_wopbs.c -- CPU restatement of the bit extraction that opens the WoP-PBS (TEST INFRASTRUCTURE, see tfhe_oracle.h).
 * Citations relative to /root/reference/tfhe/src/. */
#include "tfhe_oracle.h"
#include <stdlib.h>
#include <string.h>

/* core_crypto/fft_impl/fft64/crypto/wop_pbs/mod.rs:66-225 (extract_bits), literal loop structure.
 * in: [kN+1] under the big key; out: [nbits][n+1] under the small key, most significant extracted bit FIRST, each bit scaled
 * by q/2.  The extraction runs from the least significant bit of the window (bit delta_log) upwards. */
void orc_extract_bits(const orc_params *p, const uint64_t *ksk, const double *fbsk, const uint64_t *in, int delta_log, int nbits,
                      uint64_t *out) {
    const int W = p->bits, N = p->N, k = p->k;
    const uint64_t m = W == 64 ? ~0ULL : ((1ULL << W) - 1);
    const size_t blen = (size_t)k * N + 1, slen = (size_t)p->n + 1, glen = (size_t)(k + 1) * N;
    uint64_t *buffer = (uint64_t *)malloc(sizeof(uint64_t) * blen), *shifted = (uint64_t *)malloc(sizeof(uint64_t) * blen);
    uint64_t *ks = (uint64_t *)malloc(sizeof(uint64_t) * slen), *pbs = (uint64_t *)malloc(sizeof(uint64_t) * blen);
    uint64_t *acc = (uint64_t *)calloc(glen, sizeof(uint64_t));
    memcpy(buffer, in, sizeof(uint64_t) * blen);                                          /* lwe_in_buffer (:129-132) */
    for (int bit_idx = 0; bit_idx < nbits; bit_idx++) {                                   /* .rev().enumerate() (:158) */
        uint64_t *output_ct = out + (size_t)(nbits - 1 - bit_idx) * slen;
        for (size_t c = 0; c < blen; c++) shifted[c] = (buffer[c] << (W - delta_log - bit_idx - 1)) & m;   /* :160-166 */
        orc_keyswitch(p, ksk, shifted, ks);                                               /* :169-176 */
        memcpy(output_ct, ks, sizeof(uint64_t) * slen);                                   /* :182-184 */
        if (bit_idx == nbits - 1) break;                                                  /* :188-190 */
        ks[p->n] = (ks[p->n] + (1ULL << (W - 2))) & m;                                    /* + q/4 (:193-194) */
        for (int c = 0; c < N; c++) acc[(size_t)k * N + c] = (0 - (1ULL << (delta_log - 1 + bit_idx))) & m;   /* :198-205 */
        orc_pbs(p, fbsk, ks, acc, pbs);                                                   /* :207-213 */
        pbs[blen - 1] = (pbs[blen - 1] + (1ULL << (delta_log + bit_idx - 1))) & m;        /* :217-219 */
        for (size_t c = 0; c < blen; c++) buffer[c] = (buffer[c] - pbs[c]) & m;           /* :222-223 */
    }
    free(buffer); free(shifted); free(ks); free(pbs); free(acc);
}

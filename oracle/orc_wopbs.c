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

/* algorithms/lwe_private_functional_packing_keyswitch.rs:8-89.  key: one pfpksk [kN+1][level 1..L][(k+1)N]; in: [kN+1] (the
 * body is decomposed like a mask element); out: GLWE [(k+1)N]. */
void orc_pfpks(const orc_params *p, int base_log, int level, const uint64_t *key, const uint64_t *in, uint64_t *out) {
    const uint64_t m = p->bits == 64 ? ~0ULL : ((1ULL << p->bits) - 1);
    const int n_in = p->k * p->N + 1;
    const size_t glwe = (size_t)(p->k + 1) * p->N;
    uint64_t digits[64];
    memset(out, 0, sizeof(uint64_t) * glwe);                                              /* :63 */
    for (int j = 0; j < n_in; j++) {                                                      /* :71-88 */
        orc_decompose(in[j], base_log, level, p->bits, digits);                           /* emits level L first */
        for (int t = 0; t < level; t++) {
            const uint64_t d = digits[t];
            if (d == 0) continue;
            const uint64_t *row = key + ((size_t)j * level + (level - 1 - t)) * glwe;    /* keyswitch_key_block.iter().rev() */
            for (size_t c = 0; c < glwe; c++) out[c] -= d * row[c];
        }
    }
    for (size_t c = 0; c < glwe; c++) out[c] &= m;
}

/* fft64/crypto/wop_pbs/mod.rs:243-344 (circuit_bootstrap_boolean) with :369-444 (homomorphic_shift_boolean):
 * in: [n+1], one bit at position delta_log; ggsw_out: standard-domain GGSW [level 1..cbs_level][k+1 rows][(k+1)N]. */
void orc_circuit_bootstrap_boolean(const orc_params *p, const double *fbsk, const uint64_t *pfpksk_list, int pf_base_log, int pf_level,
                                   int cbs_base_log, int cbs_level, const uint64_t *in, int delta_log, uint64_t *ggsw_out) {
    const int W = p->bits, N = p->N, k = p->k, k1 = k + 1, n_in = k * N + 1;
    const uint64_t m = W == 64 ? ~0ULL : ((1ULL << W) - 1);
    const size_t glwe = (size_t)k1 * N, key_len = (size_t)n_in * pf_level * glwe;
    uint64_t *shifted = (uint64_t *)malloc(sizeof(uint64_t) * (p->n + 1)), *bs = (uint64_t *)malloc(sizeof(uint64_t) * n_in);
    uint64_t *acc = (uint64_t *)calloc(glwe, sizeof(uint64_t));
    for (int level = 1; level <= cbs_level; level++) {                                    /* :320-343 */
        for (int c = 0; c <= p->n; c++) shifted[c] = (in[c] << (W - delta_log - 1)) & m;  /* cleartext mul (:398-402) */
        shifted[p->n] = (shifted[p->n] + (1ULL << (W - 2))) & m;                          /* + q/4 (:405-407) */
        const uint64_t alpha = 1ULL << (W - 1 - cbs_base_log * level);
        for (int c = 0; c < N; c++) acc[(size_t)k * N + c] = (0 - alpha) & m;             /* :423-427 */
        orc_pbs(p, fbsk, shifted, acc, bs);                                               /* :431-437 */
        bs[n_in - 1] = (bs[n_in - 1] + alpha) & m;                                        /* :441-443 */
        for (int r = 0; r < k1; r++)                                                      /* :333-342 */
            orc_pfpks(p, pf_base_log, pf_level, pfpksk_list + (size_t)r * key_len, bs, ggsw_out + ((size_t)(level - 1) * k1 + r) * glwe);
    }
    free(shifted); free(bs); free(acc);
}

/* :468-592 (cmux_tree_memory_optimized): binary tree of CMUXes over the 2^t polynomials of one look-up table.  The reference walks
 * it depth-first with O(t) buffers; the operations and their operands are those of the plain layer-by-layer tree written here:
 * layer j pairs consecutive nodes, node = t0 + ggsw (x) (t1 - t0) with the j-th GGSW from the END of the list (:541).
 * fggsw: t Fourier GGSWs (most significant first); lut: [2^t][N]; out: GLWE. */
static void cmux_tree(const orc_params *pc, const double *fggsw, size_t ggsw_len, int t, const uint64_t *lut, uint64_t *out) {
    const int N = pc->N, k = pc->k;
    const uint64_t m = pc->bits == 64 ? ~0ULL : ((1ULL << pc->bits) - 1);
    const size_t glwe = (size_t)(k + 1) * N;
    size_t count = (size_t)1 << t;
    uint64_t *nodes = (uint64_t *)calloc(count * glwe, sizeof(uint64_t)), *diff = (uint64_t *)malloc(sizeof(uint64_t) * glwe);
    for (size_t q = 0; q < count; q++) memcpy(nodes + q * glwe + (size_t)k * N, lut + q * N, sizeof(uint64_t) * N);   /* :531-537 */
    for (int j = 0; j < t; j++) {
        const double *g = fggsw + (size_t)(t - 1 - j) * ggsw_len;
        for (size_t q = 0; q < count / 2; q++) {
            uint64_t *t0 = nodes + 2 * q * glwe, *t1 = t0 + glwe;
            for (size_t c = 0; c < glwe; c++) diff[c] = (t1[c] - t0[c]) & m;                                        /* :543-546 */
            orc_add_external_product(pc, t0, g, diff);                                                               /* :566-567 */
            if (q) memcpy(nodes + q * glwe, t0, sizeof(uint64_t) * glwe);
        }
        count /= 2;
    }
    memcpy(out, nodes, sizeof(uint64_t) * glwe);
    free(nodes); free(diff);
}

/* :647-746 (circuit_bootstrap_boolean_vertical_packing), :785-853 (vertical_packing: CMUX tree over the polys_per_lut = 2^t
 * polynomials of a table with the t most significant GGSWs, then blind rotation with the others),
 * :866-892 (blind_rotate_assign over the GGSW list), ggsw.rs:766-777 (cmux), then sample extraction (:852).
 * bits_in: [nbits][n+1] (most significant bit first, each scaled by q/2); luts: [n_out][polys_per_lut][N]; out: [n_out][kN+1]. */
void orc_cbs_vertical_packing_tree(const orc_params *p, const double *fbsk, const uint64_t *pfpksk_list, int pf_base_log, int pf_level,
                                   int cbs_base_log, int cbs_level, const uint64_t *bits_in, int nbits, const uint64_t *luts, int n_out,
                                   int polys_per_lut, uint64_t *out) {
    const int W = p->bits, N = p->N, k = p->k, k1 = k + 1;
    const size_t glwe = (size_t)k1 * N, ggsw_len = (size_t)cbs_level * k1 * glwe;
    uint64_t *ggsw_res = (uint64_t *)malloc(sizeof(uint64_t) * ggsw_len);
    double *fggsw = (double *)malloc(sizeof(double) * ggsw_len * (size_t)nbits);
    for (int i = 0; i < nbits; i++) {                                                     /* :721-733 */
        orc_circuit_bootstrap_boolean(p, fbsk, pfpksk_list, pf_base_log, pf_level, cbs_base_log, cbs_level,
                                      bits_in + (size_t)i * (p->n + 1), W - 1, ggsw_res);
        for (size_t q = 0; q < ggsw_len / N; q++) orc_fft_forward_torus(N, ggsw_res + q * N, W, fggsw + (size_t)i * ggsw_len + q * N);
    }
    orc_params pc = *p;
    pc.pbs_level = cbs_level;
    pc.pbs_base_log = cbs_base_log;
    int t = 0;                                                                            /* log_lut_number (:812-821) */
    while ((1 << (t + 1)) <= polys_per_lut) t++;
    if (t > nbits) t = 0;
    const uint64_t m = W == 64 ? ~0ULL : ((1ULL << W) - 1);
    uint64_t *acc = (uint64_t *)malloc(sizeof(uint64_t) * glwe), *ct1 = (uint64_t *)malloc(sizeof(uint64_t) * glwe);
    for (int o = 0; o < n_out; o++) {                                                     /* :740-745 */
        cmux_tree(&pc, fggsw, ggsw_len, t, luts + (size_t)o * polys_per_lut * N, acc);    /* :837-843 */
        uint64_t deg = 1;
        for (int i = nbits - 1; i >= t; i--) {                                            /* br_ggsw.into_ggsw_iter().rev() (:874) */
            for (int j = 0; j < k1; j++) orc_monomial_div(ct1 + (size_t)j * N, acc + (size_t)j * N, N, deg, W);
            deg <<= 1;
            for (size_t c = 0; c < glwe; c++) ct1[c] = (ct1[c] - acc[c]) & m;             /* cmux: ct1 -= ct0 */
            orc_add_external_product(&pc, acc, fggsw + (size_t)i * ggsw_len, ct1);
        }
        orc_sample_extract0(out + (size_t)o * ((size_t)k * N + 1), acc, k, N, W);
    }
    free(ggsw_res); free(fggsw); free(acc); free(ct1);
}

void orc_cbs_vertical_packing(const orc_params *p, const double *fbsk, const uint64_t *pfpksk_list, int pf_base_log, int pf_level,
                              int cbs_base_log, int cbs_level, const uint64_t *bits_in, int nbits, const uint64_t *luts, int n_out,
                              uint64_t *out) {
    orc_cbs_vertical_packing_tree(p, fbsk, pfpksk_list, pf_base_log, pf_level, cbs_base_log, cbs_level, bits_in, nbits, luts, n_out, 1, out);
}
